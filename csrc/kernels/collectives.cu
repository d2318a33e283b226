// Cross-GPU collectives over peer-mapped memory (NVLink 5 / NVSwitch), sm_100a.
//
// These kernels replace the reference's MPI calls on CuArrays, which as written stage every
// message through the host (D2H copy -> MPI -> H2D copy into a fresh array,
// /root/reference/src/mpi_extensions.jl:97-106 allreduce!, :119-128 bcast!, :141-150 reduce!;
// SURVEY §2.3) and its explicit host round trip in allreduce_gradients
// (/root/reference/src/optimizer.jl:46-47,62).
//
// Algorithms
//   oneshot_push   small messages: every rank stores its vector straight into all peers'
//                  staging slots (posted NVLink writes, no read round trip), one flag barrier,
//                  then reduces its W local slots in rank order -> latency ~ one NVLink hop.
//   twoshot_symm   in place on a symmetric window: rank r loads slice r from all peers, reduces
//                  in rank order and stores the result into every peer's window
//                  (dest_root >= 0: into root's only = reduce!).
//   twoshot_stage  arbitrary user buffers: scatter-push slices into the owners' staging,
//                  owner reduces locally, pushes the result to all peers' staging, local copy out.
//                  All NVLink traffic is posted writes.
//   bcast_*        root pushes (direct for small, scatter + all-gather for large so the root's
//                  link carries N*(W-1)/W instead of N*(W-1)).
// One owner per slice + fixed rank order 0..W-1 => every rank ends with bit-identical data.
#include <cstdarg>
#include <cstdio>
#include <string>

#include "common.cuh"

namespace fmpi_host {
static thread_local std::string g_err;
int set_error(const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return 1;
}
int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  return set_error("%s: %s", what, cudaGetErrorString(e));
}
}  // namespace fmpi_host

extern "C" const char* fmpi_cuda_last_error(void) { return fmpi_host::g_err.c_str(); }
extern "C" size_t fmpi_sigpad_bytes(void) { return fmpi::kSigPadBytes; }

namespace fmpi {

struct PtrTable {
  void* p[FMPI_MAX_RANKS];
};

// Reduce one 16-byte vector of W buffers in rank order. `addr(r)` yields rank r's address;
// with the loops fully unrolled r is a compile-time constant, so peer base pointers come
// straight from the kernel-parameter bank (no pointer arrays on the stack).
template <typename T, bool kPeerWrittenLocal, typename AddrFn>
__device__ __forceinline__ uint4 reduce_vec(AddrFn addr, int world, int op) {
  using V = VecOps<T>;
  uint4 raw[FMPI_MAX_RANKS];
#pragma unroll
  for (int r = 0; r < FMPI_MAX_RANKS; ++r)
    if (r < world) raw[r] = kPeerWrittenLocal ? ld_volatile(addr(r)) : ld_stream(addr(r));
  typename V::Acc acc[V::kN], b[V::kN];
  V::unpack(raw[0], acc);
#pragma unroll
  for (int r = 1; r < FMPI_MAX_RANKS; ++r)
    if (r < world) {
      V::unpack(raw[r], b);
#pragma unroll
      for (int k = 0; k < V::kN; ++k) acc[k] = op_apply(acc[k], b[k], op);
    }
  return V::pack(acc);
}

__host__ __device__ inline uint64_t ceil_div(uint64_t a, uint64_t b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------------------- barrier
__global__ void barrier_kernel(fmpi_devcomm c) {
  Ctx x;
  ctx_init(x, c);
  block_barrier(x, c);
  ctx_finish(x);
}

// ---------------------------------------------------------------------------- one-shot push
// Staging layout at stage_off (per rank): [parity 0|1][src rank 0..W-1][slot_bytes].
// dest_root < 0: allreduce; otherwise only root receives and reduces (reduce!).
template <typename T>
__global__ void __launch_bounds__(512)
oneshot_push_kernel(fmpi_devcomm c, PtrTable in, PtrTable out, uint64_t nbytes, int op,
                    uint64_t stage_off, uint64_t slot_bytes, int dest_root) {
  Ctx x;
  ctx_init(x, c);
  const int ai = arg_index(c);
  const uint32_t seq = seq_read(x);
  const uint64_t parity = seq & 1u;
  const uint64_t nvec = nbytes / 16, tail = nbytes % 16;
  const uint8_t* src = static_cast<const uint8_t*>(in.p[ai]);
  uint8_t* dst = static_cast<uint8_t*>(out.p[ai]);
  const uint64_t my_slot = stage_off + (parity * c.world + x.rank) * slot_bytes;
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t t0 = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;

  // push
  for (uint64_t i = t0; i < nvec; i += stride) {
    const uint4 v = ld_stream(src + i * 16);
#pragma unroll
    for (int r = 0; r < FMPI_MAX_RANKS; ++r)
      if (r < c.world && (dest_root < 0 || r == dest_root))
        st_stream(byte_ptr(c.data[r], my_slot) + i * 16, v);
  }
  if (tail && t0 == 0) {  // last partial vector: zero-padded
    uint4 v = make_uint4(0, 0, 0, 0);
    uint8_t* vb = reinterpret_cast<uint8_t*>(&v);
    for (uint64_t k = 0; k < tail; ++k) vb[k] = src[nvec * 16 + k];
#pragma unroll
    for (int r = 0; r < FMPI_MAX_RANKS; ++r)
      if (r < c.world && (dest_root < 0 || r == dest_root))
        st_stream(byte_ptr(c.data[r], my_slot) + nvec * 16, v);
  }
  block_barrier(x, c);

  // reduce my W slots (local memory that peers wrote)
  if (dest_root < 0 || x.rank == dest_root) {
    const uint8_t* base = byte_ptr(c.data[x.rank], stage_off + parity * c.world * slot_bytes);
    for (uint64_t i = t0; i < nvec; i += stride)
      *reinterpret_cast<uint4*>(dst + i * 16) = reduce_vec<T, true>(
          [&](int r) { return base + uint64_t(r) * slot_bytes + i * 16; }, c.world, op);
    if (tail && t0 == 0) {
      const uint4 v = reduce_vec<T, true>(
          [&](int r) { return base + uint64_t(r) * slot_bytes + nvec * 16; }, c.world, op);
      const uint8_t* vb = reinterpret_cast<const uint8_t*>(&v);
      for (uint64_t k = 0; k < tail; ++k) dst[nvec * 16 + k] = vb[k];
    }
  }
  seq_advance(x, seq);
  ctx_finish(x);
}

// ---------------------------------------------------------------------------- two-shot, symmetric
// In place on the window region [off, off + nvec*16). Slices are vector-granular.
template <typename T>
__global__ void __launch_bounds__(512)
twoshot_symm_kernel(fmpi_devcomm c, uint64_t off, uint64_t nvec, int op, int dest_root) {
  Ctx x;
  ctx_init(x, c);
  block_barrier(x, c);  // every rank's producer kernels have finished
  const uint64_t per = ceil_div(nvec, c.world);
  const uint64_t lo = min(per * x.rank, nvec), hi = min(lo + per, nvec);
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t i = lo + uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < hi; i += stride) {
    const uint4 v = reduce_vec<T, false>(
        [&](int r) { return byte_ptr(c.data[r], off) + i * 16; }, c.world, op);
#pragma unroll
    for (int r = 0; r < FMPI_MAX_RANKS; ++r)
      if (r < c.world && (dest_root < 0 || r == dest_root))
        st_stream(byte_ptr(c.data[r], off) + i * 16, v);
  }
  block_barrier(x, c);  // all results have landed everywhere
  ctx_finish(x);
}

// ---------------------------------------------------------------------------- two-shot, staged
// Staging at stage_off (per rank): scatter[W][per*16] then result[W*per*16].
template <typename T>
__global__ void __launch_bounds__(512)
twoshot_stage_kernel(fmpi_devcomm c, PtrTable in, PtrTable out, uint64_t nvec, int op,
                     uint64_t stage_off, int dest_root) {
  Ctx x;
  ctx_init(x, c);
  const int ai = arg_index(c);
  const uint8_t* src = static_cast<const uint8_t*>(in.p[ai]);
  uint8_t* dst = static_cast<uint8_t*>(out.p[ai]);
  const uint64_t per = ceil_div(nvec, c.world);
  const uint64_t scatter_bytes = per * 16;
  const uint64_t result_off = stage_off + uint64_t(c.world) * scatter_bytes;
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t t0 = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;

  // phase A: scatter-push slice q of my input into owner q's staging slot [my rank]
  for (int q = 0; q < c.world; ++q) {
    const uint64_t lo = min(per * q, nvec), hi = min(lo + per, nvec);
    uint8_t* slot = byte_ptr(c.data[q], stage_off + uint64_t(x.rank) * scatter_bytes);
    for (uint64_t i = lo + t0; i < hi; i += stride)
      st_stream(slot + (i - lo) * 16, ld_stream(src + i * 16));
  }
  block_barrier(x, c);

  // phase B: reduce my slice from local slots, push the result to every rank's result area
  {
    const uint64_t lo = min(per * x.rank, nvec), hi = min(lo + per, nvec);
    const uint8_t* base = byte_ptr(c.data[x.rank], stage_off);
    for (uint64_t i = lo + t0; i < hi; i += stride) {
      const uint4 v = reduce_vec<T, true>(
          [&](int r) { return base + uint64_t(r) * scatter_bytes + (i - lo) * 16; }, c.world, op);
#pragma unroll
      for (int r = 0; r < FMPI_MAX_RANKS; ++r)
        if (r < c.world && (dest_root < 0 || r == dest_root))
          st_stream(byte_ptr(c.data[r], result_off) + i * 16, v);
    }
  }
  block_barrier(x, c);

  // phase C: local copy out (same block -> element mapping as the writers used)
  if (dest_root < 0 || x.rank == dest_root) {
    const uint8_t* res = byte_ptr(c.data[x.rank], result_off);
    for (int q = 0; q < c.world; ++q) {
      const uint64_t lo = min(per * q, nvec), hi = min(lo + per, nvec);
      for (uint64_t i = lo + t0; i < hi; i += stride)
        *reinterpret_cast<uint4*>(dst + i * 16) = ld_volatile(res + i * 16);
    }
  }
  ctx_finish(x);
}

// ---------------------------------------------------------------------------- broadcast
// Symmetric, in place: direct (root stores to all) or scatter + all-gather.
__global__ void __launch_bounds__(512)
bcast_symm_kernel(fmpi_devcomm c, uint64_t off, uint64_t nvec, int root, int sag) {
  Ctx x;
  ctx_init(x, c);
  block_barrier(x, c);  // peers are done with whatever used this region before
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t t0 = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (!sag) {
    if (x.rank == root) {
      const uint8_t* src = byte_ptr(c.data[root], off);
      for (uint64_t i = t0; i < nvec; i += stride) {
        const uint4 v = ld_stream(src + i * 16);
#pragma unroll
        for (int r = 0; r < FMPI_MAX_RANKS; ++r)
          if (r < c.world && r != root) st_stream(byte_ptr(c.data[r], off) + i * 16, v);
      }
    }
    block_barrier(x, c);
  } else {
    const uint64_t per = ceil_div(nvec, c.world);
    if (x.rank == root) {  // scatter: slice q -> rank q
      const uint8_t* src = byte_ptr(c.data[root], off);
      for (int q = 0; q < c.world; ++q) {
        if (q == root) continue;
        const uint64_t lo = min(per * q, nvec), hi = min(lo + per, nvec);
        uint8_t* d = byte_ptr(c.data[q], off);
        for (uint64_t i = lo + t0; i < hi; i += stride) st_stream(d + i * 16, ld_stream(src + i * 16));
      }
    }
    block_barrier(x, c);
    if (x.rank != root) {  // all-gather: my slice -> everyone except root and me
      const uint64_t lo = min(per * x.rank, nvec), hi = min(lo + per, nvec);
      const uint8_t* src = byte_ptr(c.data[x.rank], off);
      for (uint64_t i = lo + t0; i < hi; i += stride) {
        const uint4 v = ld_volatile(src + i * 16);
#pragma unroll
        for (int r = 0; r < FMPI_MAX_RANKS; ++r)
          if (r < c.world && r != root && r != x.rank) st_stream(byte_ptr(c.data[r], off) + i * 16, v);
      }
    }
    block_barrier(x, c);
  }
  ctx_finish(x);
}

// Arbitrary buffers through double-buffered staging: [parity][nbytes rounded to 16].
__global__ void __launch_bounds__(512)
bcast_stage_kernel(fmpi_devcomm c, PtrTable buf, uint64_t nbytes, int root, uint64_t stage_off,
                   uint64_t half_bytes, int sag) {
  Ctx x;
  ctx_init(x, c);
  const int ai = arg_index(c);
  const uint32_t seq = seq_read(x);
  const uint64_t base = stage_off + uint64_t(seq & 1u) * half_bytes;
  const uint64_t nvec = nbytes / 16, tail = nbytes % 16;
  const uint64_t nv = nvec + (tail ? 1 : 0);
  uint8_t* user = static_cast<uint8_t*>(buf.p[ai]);
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t t0 = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const uint64_t per = ceil_div(nv, c.world);

  auto load_user = [&](uint64_t i) -> uint4 {
    if (i < nvec) return ld_stream(user + i * 16);
    uint4 v = make_uint4(0, 0, 0, 0);
    uint8_t* vb = reinterpret_cast<uint8_t*>(&v);
    for (uint64_t k = 0; k < tail; ++k) vb[k] = user[nvec * 16 + k];
    return v;
  };

  if (x.rank == root) {
    if (!sag) {
      for (uint64_t i = t0; i < nv; i += stride) {
        const uint4 v = load_user(i);
#pragma unroll
        for (int r = 0; r < FMPI_MAX_RANKS; ++r)
          if (r < c.world && r != root) st_stream(byte_ptr(c.data[r], base) + i * 16, v);
      }
    } else {
      for (int q = 0; q < c.world; ++q) {
        if (q == root) continue;
        const uint64_t lo = min(per * q, nv), hi = min(lo + per, nv);
        for (uint64_t i = lo + t0; i < hi; i += stride)
          st_stream(byte_ptr(c.data[q], base) + i * 16, load_user(i));
      }
      // root's own slice goes to everyone directly (root is its owner)
      const uint64_t lo = min(per * root, nv), hi = min(lo + per, nv);
      for (uint64_t i = lo + t0; i < hi; i += stride) {
        const uint4 v = load_user(i);
#pragma unroll
        for (int r = 0; r < FMPI_MAX_RANKS; ++r)
          if (r < c.world && r != root) st_stream(byte_ptr(c.data[r], base) + i * 16, v);
      }
    }
  }
  block_barrier(x, c);
  if (sag) {
    if (x.rank != root) {
      const uint64_t lo = min(per * x.rank, nv), hi = min(lo + per, nv);
      const uint8_t* mine = byte_ptr(c.data[x.rank], base);
      for (uint64_t i = lo + t0; i < hi; i += stride) {
        const uint4 v = ld_volatile(mine + i * 16);
#pragma unroll
        for (int r = 0; r < FMPI_MAX_RANKS; ++r)
          if (r < c.world && r != root && r != x.rank) st_stream(byte_ptr(c.data[r], base) + i * 16, v);
      }
    }
    block_barrier(x, c);
  }
  if (x.rank != root) {
    const uint8_t* mine = byte_ptr(c.data[x.rank], base);
    // same block->element mapping as the writers: iterate slice by slice in sag mode
    if (!sag) {
      for (uint64_t i = t0; i < nv; i += stride) {
        const uint4 v = ld_volatile(mine + i * 16);
        if (i < nvec) {
          *reinterpret_cast<uint4*>(user + i * 16) = v;
        } else {
          const uint8_t* vb = reinterpret_cast<const uint8_t*>(&v);
          for (uint64_t k = 0; k < tail; ++k) user[nvec * 16 + k] = vb[k];
        }
      }
    } else {
      for (int q = 0; q < c.world; ++q) {
        const uint64_t lo = min(per * q, nv), hi = min(lo + per, nv);
        for (uint64_t i = lo + t0; i < hi; i += stride) {
          const uint4 v = ld_volatile(mine + i * 16);
          if (i < nvec) {
            *reinterpret_cast<uint4*>(user + i * 16) = v;
          } else {
            const uint8_t* vb = reinterpret_cast<const uint8_t*>(&v);
            for (uint64_t k = 0; k < tail; ++k) user[nvec * 16 + k] = vb[k];
          }
        }
      }
    }
  }
  seq_advance(x, seq);
  ctx_finish(x);
}

}  // namespace fmpi

// ================================================================================ host side
using namespace fmpi;
using fmpi_host::check_cuda;
using fmpi_host::launch_ranked;
using fmpi_host::set_error;

namespace {

int validate(const fmpi_devcomm* c) {
  if (!c) return set_error("null communicator");
  if (c->world < 1 || c->world > FMPI_MAX_RANKS) return set_error("world %d out of range", c->world);
  if (c->channel < 0 || c->channel >= FMPI_NUM_CHANNELS) return set_error("bad channel %d", c->channel);
  if (!c->emulate && (c->rank < 0 || c->rank >= c->world)) return set_error("bad rank %d", c->rank);
  for (int r = 0; r < c->world; ++r)
    if (!c->data[r] || !c->sig[r]) return set_error("peer %d window not mapped", r);
  return 0;
}

struct Shape { int blocks, threads; };

// Defaults: latency-bound messages want few CTAs (one barrier each), bandwidth-bound ones want
// enough 16-byte requests in flight to cover the NVLink round trip (~1 MB per direction).
Shape pick_shape(const fmpi_devcomm* c, const fmpi_launch* l, uint64_t nbytes) {
  Shape s;
  s.threads = (l && l->threads > 0 && l->threads <= 512) ? l->threads : 512;
  if (l && l->blocks > 0) {
    s.blocks = l->blocks;
  } else {
    const uint64_t per_block = uint64_t(s.threads) * 16 * 4;
    uint64_t b = (nbytes + per_block - 1) / per_block;
    const uint64_t cap = nbytes >= (uint64_t(8) << 20) ? 64 : nbytes >= (uint64_t(1) << 20) ? 32 : 16;
    s.blocks = int(b < 1 ? 1 : b > cap ? cap : b);
  }
  int cap = FMPI_MAX_BLOCKS;
  if (c->emulate) cap = 128 / c->world;  // cooperative launch must be co-resident
  if (s.blocks > cap) s.blocks = cap;
  if (s.blocks < 1) s.blocks = 1;
  return s;
}

PtrTable table(const void* const* p, const fmpi_devcomm* c) {
  PtrTable t{};
  const int n = c->emulate ? c->world : 1;
  for (int i = 0; i < n; ++i) t.p[i] = const_cast<void*>(p[i]);
  return t;
}

#define FMPI_DISPATCH_DTYPE(dt, EXPR)                                   \
  switch (dt) {                                                         \
    case FMPI_F32: { using T = float; EXPR; } break;                    \
    case FMPI_F64: { using T = double; EXPR; } break;                   \
    case FMPI_F16: { using T = __half; EXPR; } break;                   \
    case FMPI_BF16: { using T = __nv_bfloat16; EXPR; } break;           \
    case FMPI_I32: { using T = int32_t; EXPR; } break;                  \
    case FMPI_I64: { using T = int64_t; EXPR; } break;                  \
    case FMPI_U8: { using T = uint8_t; EXPR; } break;                   \
    case FMPI_I8: { using T = int8_t; EXPR; } break;                    \
    case FMPI_I16: { using T = int16_t; EXPR; } break;                  \
    default: rc = set_error("unsupported dtype %d", dt);                \
  }

int reduce_symm_impl(const fmpi_devcomm* c, uint64_t off, uint64_t count, int dtype, int op,
                     int dest_root, const fmpi_launch* l, cudaStream_t st) {
  if (validate(c)) return 1;
  const size_t es = dtype_size(dtype);
  if (!es) return set_error("unsupported dtype %d", dtype);
  if (op < FMPI_SUM || op > FMPI_MAX) return set_error("unsupported op %d", op);
  if (off % 16) return set_error("window offset must be 16-byte aligned");
  const uint64_t nbytes = count * es, nvec = (nbytes + 15) / 16;
  if (off + nvec * 16 > c->data_bytes) return set_error("region exceeds the symmetric window");
  if (count == 0) return 0;
  const Shape s = pick_shape(c, l, nbytes / c->world);
  int rc = 0;
  FMPI_DISPATCH_DTYPE(dtype, rc = launch_ranked(twoshot_symm_kernel<T>, *c, s.blocks, s.threads,
                                                st, *c, off, nvec, op, dest_root));
  return rc;
}

int reduce_any_impl(const fmpi_devcomm* c, const void* const* in, void* const* out,
                    uint64_t count, int dtype, int op, int dest_root, uint64_t stage_off,
                    uint64_t stage_bytes, const fmpi_launch* l, cudaStream_t st) {
  if (validate(c)) return 1;
  const size_t es = dtype_size(dtype);
  if (!es) return set_error("unsupported dtype %d", dtype);
  if (op < FMPI_SUM || op > FMPI_MAX) return set_error("unsupported op %d", op);
  if (stage_off % 16) return set_error("staging offset must be 16-byte aligned");
  if (stage_off + stage_bytes > c->data_bytes) return set_error("staging exceeds the window");
  if (count == 0) return 0;
  const uint64_t nbytes = count * es;
  const int algo = l ? l->algo : FMPI_ALGO_AUTO;
  // The staging area is split into two private halves so that the push-without-handshake
  // one-shot kernel can never overwrite bytes a peer is still copying out of a two-shot call:
  //   [one-shot: 2 parities x W slots][two-shot: scatter W x per | result W x per]
  uint64_t os_bytes = 2ull * c->world * ((256ull << 10) + 16);
  if (os_bytes > stage_bytes / 4) os_bytes = stage_bytes / 4;
  os_bytes &= ~uint64_t(15);
  const uint64_t slot = (os_bytes / (2ull * c->world)) & ~uint64_t(15);
  const bool want_oneshot =
      algo == FMPI_ALGO_ONESHOT || (algo == FMPI_ALGO_AUTO && nbytes <= (256u << 10));
  const bool aligned = nbytes % 16 == 0;
  int rc = 0;
  for (int i = 0; i < (c->emulate ? c->world : 1); ++i)
    if ((reinterpret_cast<uintptr_t>(in[i]) % 16) || (reinterpret_cast<uintptr_t>(out[i]) % 16))
      return set_error("collective buffers must be 16-byte aligned");
  if ((want_oneshot || !aligned) && nbytes + 16 <= slot) {
    const Shape s = pick_shape(c, l, nbytes);
    const PtrTable ti = table(in, c), to = table(reinterpret_cast<const void* const*>(out), c);
    FMPI_DISPATCH_DTYPE(dtype, rc = launch_ranked(oneshot_push_kernel<T>, *c, s.blocks, s.threads,
                                                  st, *c, ti, to, nbytes, op, stage_off, slot,
                                                  dest_root));
    return rc;
  }
  if (!aligned)
    return set_error("two-shot path needs a multiple of 16 bytes (got %llu); split off the tail",
                     (unsigned long long)nbytes);
  // two-shot staged, chunked so that scatter (N) + result (N) fit its half of the staging area.
  const uint64_t ts_off = stage_off + os_bytes, ts_bytes = stage_bytes - os_bytes;
  uint64_t max_vec = ts_bytes / 32;                            // 2 areas x 16 B
  max_vec = max_vec / c->world * c->world;
  if (max_vec < 2ull * c->world) return set_error("staging area too small");
  max_vec -= c->world;                                         // room for the per-slice round-up
  const uint64_t nvec = nbytes / 16;
  const int n_ranks = c->emulate ? c->world : 1;
  for (uint64_t done = 0; done < nvec && rc == 0; done += max_vec) {
    const uint64_t n = nvec - done < max_vec ? nvec - done : max_vec;
    PtrTable ti{}, to{};
    for (int i = 0; i < n_ranks; ++i) {
      ti.p[i] = const_cast<uint8_t*>(static_cast<const uint8_t*>(in[i]) + done * 16);
      to.p[i] = static_cast<uint8_t*>(out[i]) + done * 16;
    }
    const Shape s = pick_shape(c, l, n * 16);
    FMPI_DISPATCH_DTYPE(dtype, rc = launch_ranked(twoshot_stage_kernel<T>, *c, s.blocks, s.threads,
                                                  st, *c, ti, to, n, op, ts_off, dest_root));
  }
  return rc;
}

}  // namespace

extern "C" {

int fmpi_barrier_dev(const fmpi_devcomm* c, fmpi_stream st) {
  if (validate(c)) return 1;
  return launch_ranked(barrier_kernel, *c, 1, 32, static_cast<cudaStream_t>(st), *c);
}

int fmpi_allreduce_symm(const fmpi_devcomm* c, uint64_t off, uint64_t count, int dtype, int op,
                        const fmpi_launch* l, fmpi_stream st) {
  return reduce_symm_impl(c, off, count, dtype, op, -1, l, static_cast<cudaStream_t>(st));
}

int fmpi_reduce_symm(const fmpi_devcomm* c, uint64_t off, uint64_t count, int dtype, int op,
                     int root, const fmpi_launch* l, fmpi_stream st) {
  if (c && (root < 0 || root >= c->world)) return set_error("bad root %d", root);
  return reduce_symm_impl(c, off, count, dtype, op, root, l, static_cast<cudaStream_t>(st));
}

int fmpi_allreduce_any(const fmpi_devcomm* c, const void* const* in, void* const* out,
                       uint64_t count, int dtype, int op, uint64_t stage_off,
                       uint64_t stage_bytes, const fmpi_launch* l, fmpi_stream st) {
  return reduce_any_impl(c, in, out, count, dtype, op, -1, stage_off, stage_bytes, l,
                         static_cast<cudaStream_t>(st));
}

int fmpi_reduce_any(const fmpi_devcomm* c, const void* const* in, void* const* out,
                    uint64_t count, int dtype, int op, int root, uint64_t stage_off,
                    uint64_t stage_bytes, const fmpi_launch* l, fmpi_stream st) {
  if (c && (root < 0 || root >= c->world)) return set_error("bad root %d", root);
  return reduce_any_impl(c, in, out, count, dtype, op, root, stage_off, stage_bytes, l,
                         static_cast<cudaStream_t>(st));
}

int fmpi_bcast_symm(const fmpi_devcomm* c, uint64_t off, uint64_t nbytes, int root,
                    const fmpi_launch* l, fmpi_stream st) {
  if (validate(c)) return 1;
  if (root < 0 || root >= c->world) return set_error("bad root %d", root);
  if (off % 16) return set_error("window offset must be 16-byte aligned");
  const uint64_t nvec = (nbytes + 15) / 16;
  if (off + nvec * 16 > c->data_bytes) return set_error("region exceeds the symmetric window");
  if (nbytes == 0) return 0;
  const int algo = l ? l->algo : FMPI_ALGO_AUTO;
  const int sag = (algo == FMPI_ALGO_TWOSHOT || (algo == FMPI_ALGO_AUTO && nbytes >= (1u << 20))) &&
                  c->world > 2;
  const Shape s = pick_shape(c, l, nbytes);
  return launch_ranked(bcast_symm_kernel, *c, s.blocks, s.threads, static_cast<cudaStream_t>(st),
                       *c, off, nvec, root, sag);
}

int fmpi_bcast_any(const fmpi_devcomm* c, void* const* buf, uint64_t nbytes, int root,
                   uint64_t stage_off, uint64_t stage_bytes, const fmpi_launch* l,
                   fmpi_stream st) {
  if (validate(c)) return 1;
  if (root < 0 || root >= c->world) return set_error("bad root %d", root);
  if (stage_off % 16) return set_error("staging offset must be 16-byte aligned");
  if (stage_off + stage_bytes > c->data_bytes) return set_error("staging exceeds the window");
  if (nbytes == 0) return 0;
  const uint64_t half = (stage_bytes / 2) & ~uint64_t(15);
  if (half < 32) return set_error("staging area too small");
  const uint64_t chunk = half - 16;
  const int algo = l ? l->algo : FMPI_ALGO_AUTO;
  const int n_ranks = c->emulate ? c->world : 1;
  int rc = 0;
  for (uint64_t done = 0; done < nbytes && rc == 0; done += chunk) {
    const uint64_t n = nbytes - done < chunk ? nbytes - done : chunk;
    PtrTable t{};
    for (int i = 0; i < n_ranks; ++i) t.p[i] = static_cast<uint8_t*>(buf[i]) + done;
    if (reinterpret_cast<uintptr_t>(t.p[0]) % 16)
      return set_error("bcast buffer must be 16-byte aligned");
    const int sag = (algo == FMPI_ALGO_TWOSHOT || (algo == FMPI_ALGO_AUTO && n >= (1u << 20))) &&
                    c->world > 2;
    const Shape s = pick_shape(c, l, n);
    rc = launch_ranked(bcast_stage_kernel, *c, s.blocks, s.threads, static_cast<cudaStream_t>(st),
                       *c, t, n, root, stage_off, half, sag);
  }
  return rc;
}

__global__ void flush_l2_kernel(uint4* p, uint64_t nvec, uint32_t salt) {
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride)
    p[i] = make_uint4(salt, uint32_t(i), salt ^ uint32_t(i), 0);
}

int fmpi_flush_l2(void* buf, uint64_t nbytes, fmpi_stream st) {
  static uint32_t salt = 0;
  flush_l2_kernel<<<148 * 4, 512, 0, static_cast<cudaStream_t>(st)>>>(static_cast<uint4*>(buf),
                                                                     nbytes / 16, ++salt);
  return check_cuda(cudaGetLastError(), "flush_l2");
}

}  // extern "C"
