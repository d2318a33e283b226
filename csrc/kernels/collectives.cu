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
//                  owner reduces (own slice straight from its input), pushes the result to the
//                  peers' staging and writes its own output directly, peers copy out.
//                  All NVLink traffic is posted writes.
//   nvls_*         in-switch reduction: multimem.ld_reduce pulls ONE reduced vector through the
//                  NVSwitch instead of W peer vectors, multimem.st replicates the result to all
//                  GPUs with one store. Link bytes per direction ~ N instead of 2N(W-1)/W.
//   bcast_*        root pushes (direct / multicast for small, scatter + all-gather for large
//                  P2P so the root's link carries N*(W-1)/W instead of N*(W-1)).
// One owner per slice + fixed rank order 0..W-1 => every rank ends with bit-identical data.
//
// Memory-level parallelism: every loop loads a batch of U vectors before it touches any of
// them (the asm loads are volatile, so the batch is what puts requests in flight).
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
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

// ---- health page: pinned + mapped, one per device of this process
#include <mutex>
namespace {
struct HealthPage { uint32_t* host = nullptr; uint32_t* dev = nullptr; };
HealthPage g_health[64];
std::mutex g_health_mu;
int health_for_current(HealthPage** out) {
  int dev = 0;
  if (fmpi_host::check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return 1;
  if (dev < 0 || dev >= 64) return fmpi_host::set_error("device index %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_health_mu);
  HealthPage& h = g_health[dev];
  if (!h.host) {
    void* p = nullptr;
    if (fmpi_host::check_cuda(cudaHostAlloc(&p, 64, cudaHostAllocMapped | cudaHostAllocPortable), "cudaHostAlloc(health page)")) return 1;
    memset(p, 0, 64);
    void* d = nullptr;
    if (fmpi_host::check_cuda(cudaHostGetDevicePointer(&d, p, 0), "cudaHostGetDevicePointer")) { cudaFreeHost(p); return 1; }
    h.host = static_cast<uint32_t*>(p);
    h.dev = static_cast<uint32_t*>(d);
  }
  *out = &h;
  return 0;
}
}  // namespace
extern "C" int fmpi_health_page(uint32_t** host_ptr, uint32_t** dev_ptr) {
  HealthPage* h;
  if (health_for_current(&h)) return 1;
  if (host_ptr) *host_ptr = h->host;
  if (dev_ptr) *dev_ptr = h->dev;
  return 0;
}
extern "C" int fmpi_health_check(void) {
  HealthPage* h;
  if (health_for_current(&h)) return 1;
  const volatile uint32_t* w = h->host;
  if (w[fmpi::kCtrlError])
    return fmpi_host::set_error("a cross-GPU wait was abandoned after its time budget: a peer rank is lost or "
                                "far behind (results of the collectives in flight are undefined)");
  if (w[fmpi::kCtrlAbort]) return fmpi_host::set_error("the device world was aborted by the host");
  return 0;
}
extern "C" int fmpi_health_abort(void) {
  HealthPage* h;
  if (health_for_current(&h)) return 1;
  reinterpret_cast<volatile uint32_t*>(h->host)[fmpi::kCtrlAbort] = 1;
  return 0;
}
extern "C" int fmpi_health_reset(void) {
  HealthPage* h;
  if (health_for_current(&h)) return 1;
  reinterpret_cast<volatile uint32_t*>(h->host)[fmpi::kCtrlAbort] = 0;
  reinterpret_cast<volatile uint32_t*>(h->host)[fmpi::kCtrlError] = 0;
  return 0;
}

namespace fmpi {

struct PtrTable {
  void* p[FMPI_MAX_RANKS];
};

__host__ __device__ inline uint64_t ceil_div(uint64_t a, uint64_t b) { return (a + b - 1) / b; }

// Vectors j in [lo, hi) with j == lo + t0 (mod stride) belong to this thread; U are in flight.
// The block -> vector mapping does not depend on U, so producer and consumer phases that use
// this helper with the same (lo, t0, stride) meet on the same block (= the same barrier).
template <int U, typename LoadFn, typename StoreFn>
__device__ __forceinline__ void vec_loop(uint64_t lo, uint64_t hi, uint64_t t0, uint64_t stride,
                                         LoadFn ld, StoreFn st) {
  for (uint64_t i = lo + t0; i < hi; i += U * stride) {
    uint4 v[U];
#pragma unroll
    for (int k = 0; k < U; ++k)
      if (i + k * stride < hi) v[k] = ld(i + k * stride);
#pragma unroll
    for (int k = 0; k < U; ++k)
      if (i + k * stride < hi) st(i + k * stride, v[k]);
  }
}

// Same, but each element is the rank-ordered reduction of W vectors; addr(r, j) yields rank r's
// copy of vector j. All U*W loads are issued before any arithmetic.
template <typename T, int U, bool kPeerWritten, typename AddrFn, typename StoreFn>
__device__ __forceinline__ void reduce_loop(uint64_t lo, uint64_t hi, uint64_t t0, uint64_t stride,
                                            int world, int op, AddrFn addr, StoreFn st) {
  using V = VecOps<T>;
  for (uint64_t i = lo + t0; i < hi; i += U * stride) {
    uint4 raw[U][FMPI_MAX_RANKS];
#pragma unroll
    for (int k = 0; k < U; ++k)
      if (i + k * stride < hi) {
#pragma unroll
        for (int r = 0; r < FMPI_MAX_RANKS; ++r)
          if (r < world)
            raw[k][r] = kPeerWritten ? ld_volatile(addr(r, i + k * stride)) : ld_stream(addr(r, i + k * stride));
      }
#pragma unroll
    for (int k = 0; k < U; ++k)
      if (i + k * stride < hi) {
        typename V::Acc acc[V::kN], b[V::kN];
        V::unpack(raw[k][0], acc);
#pragma unroll
        for (int r = 1; r < FMPI_MAX_RANKS; ++r)
          if (r < world) {
            V::unpack(raw[k][r], b);
#pragma unroll
            for (int e = 0; e < V::kN; ++e) acc[e] = op_apply(acc[e], b[e], op);
          }
        st(i + k * stride, V::pack(acc));
      }
  }
}

// The same loop with the register budget of the raw vectors (64 registers) spent according to
// the world size: 8 element groups in flight at W <= 2, 4 at W <= 4, 2 at W <= 8. With few ranks a
// thread has few loads per element group, and a remote load takes microseconds: without the extra
// groups in flight a two-rank reduction is latency-bound at ~300 GB/s (measured, 64 CTAs).
template <typename T, int U, int MAXW, bool kPeerWritten, typename AddrFn, typename StoreFn>
__device__ __forceinline__ void reduce_loop_n(uint64_t lo, uint64_t hi, uint64_t t0, uint64_t stride,
                                              int world, int op, AddrFn addr, StoreFn st) {
  using V = VecOps<T>;
  for (uint64_t i = lo + t0; i < hi; i += U * stride) {
    uint4 raw[U][MAXW];
#pragma unroll
    for (int k = 0; k < U; ++k)
      if (i + k * stride < hi) {
#pragma unroll
        for (int r = 0; r < MAXW; ++r)
          if (r < world)
            raw[k][r] = kPeerWritten ? ld_volatile(addr(r, i + k * stride)) : ld_stream(addr(r, i + k * stride));
      }
#pragma unroll
    for (int k = 0; k < U; ++k)
      if (i + k * stride < hi) {
        typename V::Acc acc[V::kN], b[V::kN];
        V::unpack(raw[k][0], acc);
#pragma unroll
        for (int r = 1; r < MAXW; ++r)
          if (r < world) {
            V::unpack(raw[k][r], b);
#pragma unroll
            for (int e = 0; e < V::kN; ++e) acc[e] = op_apply(acc[e], b[e], op);
          }
        st(i + k * stride, V::pack(acc));
      }
  }
}
template <typename T, bool kPeerWritten, typename AddrFn, typename StoreFn>
__device__ __forceinline__ void reduce_loop_w(uint64_t lo, uint64_t hi, uint64_t t0, uint64_t stride,
                                              int world, int op, AddrFn addr, StoreFn st) {
  if (world <= 2) reduce_loop_n<T, 8, 2, kPeerWritten>(lo, hi, t0, stride, world, op, addr, st);
  else if (world <= 4) reduce_loop_n<T, 4, 4, kPeerWritten>(lo, hi, t0, stride, world, op, addr, st);
  else reduce_loop_n<T, 2, 8, kPeerWritten>(lo, hi, t0, stride, world, op, addr, st);
}

template <typename Fn>
__device__ __forceinline__ void store_to_ranks(const fmpi_devcomm& c, int dest_root, int skip, Fn ptr_of,
                                               uint4 v) {
#pragma unroll
  for (int r = 0; r < FMPI_MAX_RANKS; ++r)
    if (r < c.world && r != skip && (dest_root < 0 || r == dest_root)) st_stream(ptr_of(r), v);
}

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

  vec_loop<2>(0, nvec, t0, stride, [&](uint64_t j) { return ld_stream(src + j * 16); },
              [&](uint64_t j, uint4 v) {
                store_to_ranks(c, dest_root, -1, [&](int r) { return byte_ptr(c.data[r], my_slot) + j * 16; }, v);
              });
  if (tail && t0 == 0) {  // last partial vector: zero-padded
    uint4 v = make_uint4(0, 0, 0, 0);
    uint8_t* vb = reinterpret_cast<uint8_t*>(&v);
    for (uint64_t k = 0; k < tail; ++k) vb[k] = src[nvec * 16 + k];
    store_to_ranks(c, dest_root, -1, [&](int r) { return byte_ptr(c.data[r], my_slot) + nvec * 16; }, v);
  }
  block_barrier(x, c);

  if (dest_root < 0 || x.rank == dest_root) {
    const uint8_t* base = byte_ptr(c.data[x.rank], stage_off + parity * c.world * slot_bytes);
    auto addr = [&](int r, uint64_t j) { return base + uint64_t(r) * slot_bytes + j * 16; };
    reduce_loop<T, 1, true>(0, nvec, t0, stride, c.world, op, addr,
                            [&](uint64_t j, uint4 v) { *reinterpret_cast<uint4*>(dst + j * 16) = v; });
    if (tail && t0 == 0) {
      reduce_loop<T, 1, true>(nvec, nvec + 1, 0, 1, c.world, op, addr, [&](uint64_t, uint4 v) {
        const uint8_t* vb = reinterpret_cast<const uint8_t*>(&v);
        for (uint64_t k = 0; k < tail; ++k) dst[nvec * 16 + k] = vb[k];
      });
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
  const uint64_t t0 = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  reduce_loop_w<T, false>(
      lo, hi, t0, stride, c.world, op, [&](int r, uint64_t j) { return byte_ptr(c.data[r], off) + j * 16; },
      [&](uint64_t j, uint4 v) {
        store_to_ranks(c, dest_root, -1, [&](int r) { return byte_ptr(c.data[r], off) + j * 16; }, v);
      });
  block_barrier(x, c);  // all results have landed everywhere
  ctx_finish(x);
}

// In-switch version (sum of f32 / f16 / bf16): one multimem.ld_reduce + one multimem.st per vector.
template <typename T>
__global__ void __launch_bounds__(512)
nvls_symm_kernel(fmpi_devcomm c, uint64_t off, uint64_t nvec, int dest_root) {
  Ctx x;
  ctx_init(x, c);
  block_barrier(x, c);
  const uint64_t per = ceil_div(nvec, c.world);
  const uint64_t lo = min(per * x.rank, nvec), hi = min(lo + per, nvec);
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t t0 = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  uint8_t* mc = byte_ptr(c.mc_data, off);
  vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return Multimem<T>::ld_reduce(mc + j * 16); },
              [&](uint64_t j, uint4 v) {
                if (dest_root < 0) multimem_st(mc + j * 16, v);
                else st_stream(byte_ptr(c.data[dest_root], off) + j * 16, v);
              });
  block_barrier(x, c);
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
  // (my own slice stays where it is: phase B reads it straight from the input)
  for (int dq = 1; dq < c.world; ++dq) {
    const int q = (x.rank + dq) % c.world;  // staggered so the ranks do not gang up on one peer
    const uint64_t lo = min(per * q, nvec), hi = min(lo + per, nvec);
    uint8_t* slot = byte_ptr(c.data[q], stage_off + uint64_t(x.rank) * scatter_bytes);
    vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_stream(src + j * 16); },
                [&](uint64_t j, uint4 v) { st_stream(slot + (j - lo) * 16, v); });
  }
  block_barrier(x, c);

  // phase B: reduce my slice, push the result to the peers' result areas, write mine directly
  {
    const uint64_t lo = min(per * x.rank, nvec), hi = min(lo + per, nvec);
    const uint8_t* base = byte_ptr(c.data[x.rank], stage_off);
    const int me = x.rank;
    const bool mine_too = dest_root < 0 || dest_root == me;
    reduce_loop_w<T, true>(
        lo, hi, t0, stride, c.world, op,
        [&](int r, uint64_t j) { return r == me ? src + j * 16 : base + uint64_t(r) * scatter_bytes + (j - lo) * 16; },
        [&](uint64_t j, uint4 v) {
          store_to_ranks(c, dest_root, me, [&](int r) { return byte_ptr(c.data[r], result_off) + j * 16; }, v);
          if (mine_too) *reinterpret_cast<uint4*>(dst + j * 16) = v;
        });
  }
  block_barrier(x, c);

  // phase C: copy the other owners' slices out of my result area
  if (dest_root < 0 || x.rank == dest_root) {
    const uint8_t* res = byte_ptr(c.data[x.rank], result_off);
    for (int dq = 1; dq < c.world; ++dq) {
      const int q = (x.rank + dq) % c.world;
      const uint64_t lo = min(per * q, nvec), hi = min(lo + per, nvec);
      vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_volatile(res + j * 16); },
                  [&](uint64_t j, uint4 v) { *reinterpret_cast<uint4*>(dst + j * 16) = v; });
    }
  }
  ctx_finish(x);
}

// In-switch version for arbitrary buffers: copy in -> ld_reduce/st on the multicast alias of the
// staging area (in place: only the owner ever touches a vector) -> copy out.
template <typename T>
__global__ void __launch_bounds__(512)
nvls_stage_kernel(fmpi_devcomm c, PtrTable in, PtrTable out, uint64_t nvec, uint64_t stage_off,
                  int dest_root) {
  Ctx x;
  ctx_init(x, c);
  const int ai = arg_index(c);
  const uint8_t* src = static_cast<const uint8_t*>(in.p[ai]);
  uint8_t* dst = static_cast<uint8_t*>(out.p[ai]);
  const uint64_t per = ceil_div(nvec, c.world);
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t t0 = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  uint8_t* mine = byte_ptr(c.data[x.rank], stage_off);
  uint8_t* mc = byte_ptr(c.mc_data, stage_off);

  for (int q = 0; q < c.world; ++q) {  // slice-wise so every phase maps vectors to the same block
    const uint64_t lo = min(per * q, nvec), hi = min(lo + per, nvec);
    vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_stream(src + j * 16); },
                [&](uint64_t j, uint4 v) { st_stream(mine + j * 16, v); });
  }
  block_barrier(x, c);
  {
    const uint64_t lo = min(per * x.rank, nvec), hi = min(lo + per, nvec);
    vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return Multimem<T>::ld_reduce(mc + j * 16); },
                [&](uint64_t j, uint4 v) {
                  if (dest_root < 0) multimem_st(mc + j * 16, v);
                  else st_stream(byte_ptr(c.data[dest_root], stage_off) + j * 16, v);
                });
  }
  block_barrier(x, c);
  if (dest_root < 0 || x.rank == dest_root) {
    for (int q = 0; q < c.world; ++q) {
      const uint64_t lo = min(per * q, nvec), hi = min(lo + per, nvec);
      vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_volatile(mine + j * 16); },
                  [&](uint64_t j, uint4 v) { *reinterpret_cast<uint4*>(dst + j * 16) = v; });
    }
  }
  ctx_finish(x);
}

// ---------------------------------------------------------------------------- broadcast
// Symmetric, in place. mode 0: root stores to every peer; 1: scatter + all-gather; 2: multicast.
__global__ void __launch_bounds__(512)
bcast_symm_kernel(fmpi_devcomm c, uint64_t off, uint64_t nvec, int root, int mode) {
  Ctx x;
  ctx_init(x, c);
  block_barrier(x, c);  // peers are done with whatever used this region before
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t t0 = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const uint8_t* src = byte_ptr(c.data[x.rank], off);
  if (mode >= 3) {
    // pull + re-emit: rank q loads slice q from the ROOT's window (the root's link carries the
    // message once, spread over W-1 readers) and stores it to every other rank — through the
    // multicast alias when there is one — so all W links send at once. The root's own slice is
    // emitted by the root. No barrier between the two steps: a vector is re-emitted by the
    // thread that fetched it.
    const uint64_t per = ceil_div(nvec, c.world);
    const uint64_t lo = min(per * x.rank, nvec), hi = min(lo + per, nvec);
    const uint8_t* rsrc = byte_ptr(c.data[root], off);
    const int me = x.rank;
    if (c.mc_data && !c.emulate && mode == 3) {
      uint8_t* mc = byte_ptr(c.mc_data, off);
      vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_stream(rsrc + j * 16); },
                  [&](uint64_t j, uint4 v) { multimem_st(mc + j * 16, v); });
    } else {   // mode 4, or no multicast mapping: re-emit with W - 2 peer stores
      vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_stream(rsrc + j * 16); },
                  [&](uint64_t j, uint4 v) {
#pragma unroll
                    for (int r = 0; r < FMPI_MAX_RANKS; ++r)
                      if (r < c.world && r != root) st_stream(byte_ptr(c.data[r], off) + j * 16, v);
                    (void)me;
                  });
    }
    block_barrier(x, c);
  } else if (mode != 1) {
    if (x.rank == root) {
      if (mode == 2) {
        uint8_t* mc = byte_ptr(c.mc_data, off);
        vec_loop<4>(0, nvec, t0, stride, [&](uint64_t j) { return ld_stream(src + j * 16); },
                    [&](uint64_t j, uint4 v) { multimem_st(mc + j * 16, v); });
      } else {
        vec_loop<4>(0, nvec, t0, stride, [&](uint64_t j) { return ld_stream(src + j * 16); },
                    [&](uint64_t j, uint4 v) {
                      store_to_ranks(c, -1, root, [&](int r) { return byte_ptr(c.data[r], off) + j * 16; }, v);
                    });
      }
    }
    block_barrier(x, c);
  } else {
    const uint64_t per = ceil_div(nvec, c.world);
    if (x.rank == root) {
      for (int dq = 1; dq < c.world; ++dq) {  // scatter: slice q -> rank q
        const int q = (root + dq) % c.world;
        const uint64_t lo = min(per * q, nvec), hi = min(lo + per, nvec);
        uint8_t* d = byte_ptr(c.data[q], off);
        vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_stream(src + j * 16); },
                    [&](uint64_t j, uint4 v) { st_stream(d + j * 16, v); });
      }
      // the root owns its slice: it goes to everybody directly
      const uint64_t lo = min(per * root, nvec), hi = min(lo + per, nvec);
      vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_stream(src + j * 16); },
                  [&](uint64_t j, uint4 v) {
                    store_to_ranks(c, -1, root, [&](int r) { return byte_ptr(c.data[r], off) + j * 16; }, v);
                  });
    }
    block_barrier(x, c);
    if (x.rank != root) {  // all-gather: my slice -> everyone except root and me
      const uint64_t lo = min(per * x.rank, nvec), hi = min(lo + per, nvec);
      const int me = x.rank;
      vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_volatile(src + j * 16); },
                  [&](uint64_t j, uint4 v) {
#pragma unroll
                    for (int r = 0; r < FMPI_MAX_RANKS; ++r)
                      if (r < c.world && r != root && r != me) st_stream(byte_ptr(c.data[r], off) + j * 16, v);
                  });
    }
    block_barrier(x, c);
  }
  ctx_finish(x);
}

// Arbitrary buffers through double-buffered staging: [parity][nbytes rounded to 16].
__global__ void __launch_bounds__(512)
bcast_stage_kernel(fmpi_devcomm c, PtrTable buf, uint64_t nbytes, int root, uint64_t stage_off,
                   uint64_t half_bytes, int mode) {
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
  const int me = x.rank;

  auto load_user = [&](uint64_t j) -> uint4 {
    if (j < nvec) return ld_stream(user + j * 16);
    uint4 v = make_uint4(0, 0, 0, 0);
    uint8_t* vb = reinterpret_cast<uint8_t*>(&v);
    for (uint64_t k = 0; k < tail; ++k) vb[k] = user[nvec * 16 + k];
    return v;
  };
  auto store_user = [&](uint64_t j, uint4 v) {
    if (j < nvec) {
      *reinterpret_cast<uint4*>(user + j * 16) = v;
    } else {
      const uint8_t* vb = reinterpret_cast<const uint8_t*>(&v);
      for (uint64_t k = 0; k < tail; ++k) user[nvec * 16 + k] = vb[k];
    }
  };
  auto to_all_but_root = [&](uint64_t j, uint4 v) {
    store_to_ranks(c, -1, root, [&](int r) { return byte_ptr(c.data[r], base) + j * 16; }, v);
  };

  if (me == root) {
    if (mode == 2) {
      uint8_t* mc = byte_ptr(c.mc_data, base);
      vec_loop<4>(0, nv, t0, stride, load_user, [&](uint64_t j, uint4 v) { multimem_st(mc + j * 16, v); });
    } else if (mode == 0) {
      vec_loop<4>(0, nv, t0, stride, load_user, to_all_but_root);
    } else {
      for (int dq = 1; dq < c.world; ++dq) {
        const int q = (root + dq) % c.world;
        const uint64_t lo = min(per * q, nv), hi = min(lo + per, nv);
        uint8_t* d = byte_ptr(c.data[q], base);
        vec_loop<4>(lo, hi, t0, stride, load_user, [&](uint64_t j, uint4 v) { st_stream(d + j * 16, v); });
      }
      const uint64_t lo = min(per * root, nv), hi = min(lo + per, nv);
      vec_loop<4>(lo, hi, t0, stride, load_user, to_all_but_root);
    }
  }
  block_barrier(x, c);
  const uint8_t* mine = byte_ptr(c.data[me], base);
  if (mode == 1) {
    if (me != root) {
      const uint64_t lo = min(per * me, nv), hi = min(lo + per, nv);
      vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_volatile(mine + j * 16); },
                  [&](uint64_t j, uint4 v) {
#pragma unroll
                    for (int r = 0; r < FMPI_MAX_RANKS; ++r)
                      if (r < c.world && r != root && r != me) st_stream(byte_ptr(c.data[r], base) + j * 16, v);
                  });
    }
    block_barrier(x, c);
  }
  if (me != root) {
    if (mode != 1) {
      vec_loop<4>(0, nv, t0, stride, [&](uint64_t j) { return ld_volatile(mine + j * 16); }, store_user);
    } else {  // slice-wise: same block -> vector mapping as the writers
      for (int q = 0; q < c.world; ++q) {
        const uint64_t lo = min(per * q, nv), hi = min(lo + per, nv);
        vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_volatile(mine + j * 16); }, store_user);
      }
    }
  }
  seq_advance(x, seq);
  ctx_finish(x);
}

// ---------------------------------------------------------------------------- chunk pipeline
// Large messages in ordinary (unregistered) buffers. One persistent kernel per rank whose CTAs
// take one of three roles and hand chunks to each other through counters and cross-GPU flags,
// so the staging copies run UNDER the link phase instead of before and after it:
//
//   in    copy chunk k of the user buffer into staging slot k % S       (local HBM traffic)
//         (bcast: the root pushes slice q of the chunk straight into rank q's slot)
//   link  on slot k % S, owner of slice q: multimem.ld_reduce + multimem.st (in-switch sum and
//         multicast) or W peer loads + W peer stores; reduce: result to the root's slot only;
//         bcast: re-emit my slice to everybody
//   out   copy slot k % S into the user buffer                            (local HBM traffic)
//
// Per slot s and use u (= k / S) the hand-over values are U = base[s] + u + 1 (base[s] = uses of
// the slot in all earlier calls, so flags only ever grow):
//   in  -> link : every rank's flag[IN][s]   >= U   (all ranks copied chunk k in)
//   link -> out : every rank's flag[LINK][s] >= U   (all slices of chunk k reduced and landed)
//   out -> in   : flag[LINK][s] >= U on every rank (peers no longer read my slot) and my own
//                 out-counter; bcast: flag[OUT][s] >= U from every receiver
// A role's CTAs count themselves done per slot; the one that completes the count signals.
// All CTAs are co-resident (grid <= SM count), so the waits cannot deadlock.
struct PipeArgs {
  PtrTable in, out;
  uint64_t nvec, chunk_vec, stage_off;
  int nslots, kind, root, op;   // kind: 0 all-reduce, 1 reduce to root, 2 broadcast from root
  int n_link, n_copy;           // CTAs in the link role; the rest are split over in / out
  uint64_t* trace;              // diagnostics (FMPI_PIPE_TRACE=1): [3 roles][256 chunks][4] globaltimer stamps
};
enum { kPipeIn = 0, kPipeLink = 1, kPipeOut = 2 };

template <typename T, bool kNvls>
__global__ void __launch_bounds__(512)
pipe_kernel(fmpi_devcomm c, PipeArgs a) {
  Ctx x;
  ctx_init(x, c);
  const int ai = arg_index(c);
  const int me = x.rank, W = c.world, S = a.nslots;
  const uint8_t* src = static_cast<const uint8_t*>(a.in.p[ai]);
  uint8_t* dst = static_cast<uint8_t*>(a.out.p[ai]);
  uint32_t* pad = x.my_pad;
  __shared__ uint32_t base_s[kPipeSlots];
  if (threadIdx.x < kPipeSlots) base_s[threadIdx.x] = pad[kCtlBase + kCtlPipeBase + threadIdx.x];
  // every rank has entered this call => every rank has left the previous one on this channel
  block_barrier(x, c);

  // ---- roles (the grid is the same on every rank; what a copy CTA does depends on the rank)
  const bool direct = a.kind == 2 && a.n_link == 0;   // broadcast without the re-emit step
  const bool has_in = a.kind != 2 || me == a.root;
  const bool has_out = a.kind == 0 || (a.kind == 1 ? me == a.root : me != a.root);
  const int n_in = has_in ? (has_out ? a.n_copy / 2 : a.n_copy) : 0;
  const int n_out = has_out ? a.n_copy - n_in : 0;
  int role, idx, nrole;
  if (int(blockIdx.x) < a.n_link) { role = kPipeLink; idx = blockIdx.x; nrole = a.n_link; }
  else if (int(blockIdx.x) - a.n_link < n_in) { role = kPipeIn; idx = blockIdx.x - a.n_link; nrole = n_in; }
  else { role = kPipeOut; idx = blockIdx.x - a.n_link - n_in; nrole = n_out; }

  const uint32_t chan_off = uint32_t(c.channel) * kChannelWords;
  auto flag_of = [&](int r, int f, int s, int writer) -> uint32_t* {
    return c.sig[r] + chan_off + kPipeFlagBase + (f * kPipeSlots + s) * FMPI_MAX_RANKS + writer;
  };
  auto counter = [&](int rl, int s) -> uint32_t* { return pad + kCtlBase + kCtlPipeCnt + rl * kPipeSlots + s; };
  // wait until flag f of slot s written by every rank in [0, W) (only_root >= 0: that rank only;
  // skip_root >= 0: everybody but that rank) has reached `value`
  auto wait_peers = [&](int f, int s, uint32_t value, int only, int skip) {
    if (int(threadIdx.x) < W && (only < 0 || int(threadIdx.x) == only) && int(threadIdx.x) != skip && !x.dead)
      x.dead = !wait_flag(flag_of(me, f, s, threadIdx.x), value, c);
    __syncthreads();
  };
  auto wait_local = [&](int rl, int s, uint32_t value) {
    if (threadIdx.x == 0 && !x.dead) x.dead = !wait_flag(counter(rl, s), value, c);
    __syncthreads();
  };
  // this CTA is done with its share of the slot; the CTA that completes the count tells the peers
  // (one sys-scope release fence per CTA and chunk: it orders this CTA's data stores — the
  // __syncthreads before it makes them this thread's predecessors — before the counter; the CTA
  // that completes the count then publishes the flag to every peer with release stores)
  auto done = [&](int rl, int s, uint32_t total, uint32_t value) {
    __syncthreads();
    if (threadIdx.x == 0) {
      fence_acq_rel_sys();
      if (atomicAdd(counter(rl, s), 1u) + 1 == total) {
        fence_acq_rel_sys();   // one fence, then relaxed flag stores: a release store each would
        for (int r = 0; r < W; ++r) st_relaxed_sys(flag_of(r, rl, s, me), value);   // cost W more MEMBAR.SYS
      }
    }
  };

  const uint64_t nchunks = ceil_div(a.nvec, a.chunk_vec);
  const uint64_t t0 = uint64_t(idx) * blockDim.x + threadIdx.x;
  const uint64_t stride = uint64_t(nrole) * blockDim.x;
  const uint64_t slot_bytes = a.chunk_vec * 16;

  if (nrole > 0)
    for (uint64_t k = 0; k < nchunks; ++k) {
      const int s = int(k % S);
      const uint32_t u = uint32_t(k / S);
      const uint32_t U = base_s[s] + u + 1;
      const uint64_t v0 = k * a.chunk_vec;                       // first vector of the chunk
      const uint64_t cv = min(a.chunk_vec, a.nvec - v0);         // vectors in the chunk
      const uint64_t per = ceil_div(cv, uint64_t(W));
      const uint64_t soff = a.stage_off + uint64_t(s) * slot_bytes;

      uint64_t* tr = (a.trace && idx == 0 && threadIdx.x == 0 && k < 256 && arg_index(c) == 0)
                         ? a.trace + (uint64_t(role) * 256 + k) * 4 : nullptr;
      if (tr) tr[0] = globaltimer_ns();
      if (role == kPipeIn) {
        if (u > 0) {
          if (a.kind == 2) wait_peers(kPipeOut, s, U - 1, -1, a.root);  // receivers emptied the slot
          else {
            wait_peers(kPipeLink, s, U - 1, -1, -1);
            if (n_out > 0) wait_local(kPipeOut, s, uint32_t(n_out) * u);
          }
        }
        if (tr) tr[1] = globaltimer_ns();
        if (a.kind != 2) {
          uint8_t* slot = byte_ptr(c.data[me], soff);
          vec_loop<8>(0, cv, t0, stride, [&](uint64_t j) { return ld_stream(src + (v0 + j) * 16); },
                      [&](uint64_t j, uint4 v) { *reinterpret_cast<uint4*>(slot + j * 16) = v; });
        } else if (direct) {                  // the root sends the whole chunk to every receiver
          if (kNvls) {
            uint8_t* mc = byte_ptr(c.mc_data, soff);
            vec_loop<8>(0, cv, t0, stride, [&](uint64_t j) { return ld_stream(src + (v0 + j) * 16); },
                        [&](uint64_t j, uint4 v) { multimem_st(mc + j * 16, v); });
          } else {
            vec_loop<8>(0, cv, t0, stride, [&](uint64_t j) { return ld_stream(src + (v0 + j) * 16); },
                        [&](uint64_t j, uint4 v) {
                          store_to_ranks(c, -1, me, [&](int r) { return byte_ptr(c.data[r], soff) + j * 16; }, v);
                        });
          }
        } else {
          for (int dq = 0; dq < W; ++dq) {    // the root scatters: slice q -> rank q's slot
            const int q = (me + dq) % W;
            const uint64_t lo = min(per * q, cv), hi = min(lo + per, cv);
            uint8_t* slot = byte_ptr(c.data[q], soff);
            vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_stream(src + (v0 + j) * 16); },
                        [&](uint64_t j, uint4 v) { st_stream(slot + j * 16, v); });
          }
        }
        if (tr) tr[2] = globaltimer_ns();
        done(kPipeIn, s, uint32_t(n_in) * (u + 1), U);
        if (tr) tr[3] = globaltimer_ns();
      } else if (role == kPipeLink) {
        wait_peers(kPipeIn, s, U, a.kind == 2 ? a.root : -1, -1);
        if (tr) tr[1] = globaltimer_ns();
        const uint64_t lo = min(per * me, cv), hi = min(lo + per, cv);
        if (a.kind == 2) {
          const uint8_t* mine = byte_ptr(c.data[me], soff);
          if (kNvls) {
            uint8_t* mc = byte_ptr(c.mc_data, soff);
            vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_volatile(mine + j * 16); },
                        [&](uint64_t j, uint4 v) { multimem_st(mc + j * 16, v); });
          } else {
            vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return ld_volatile(mine + j * 16); },
                        [&](uint64_t j, uint4 v) {
#pragma unroll
                          for (int r = 0; r < FMPI_MAX_RANKS; ++r)
                            if (r < W && r != a.root && r != me) st_stream(byte_ptr(c.data[r], soff) + j * 16, v);
                        });
          }
        } else if constexpr (kNvls) {
          uint8_t* mc = byte_ptr(c.mc_data, soff);
          uint8_t* rootslot = byte_ptr(c.data[a.kind == 1 ? a.root : me], soff);
          vec_loop<4>(lo, hi, t0, stride, [&](uint64_t j) { return Multimem<T>::ld_reduce(mc + j * 16); },
                      [&](uint64_t j, uint4 v) {
                        if (a.kind == 0) multimem_st(mc + j * 16, v);
                        else st_stream(rootslot + j * 16, v);
                      });
        } else {
          const int dest_root = a.kind == 1 ? a.root : -1;
          reduce_loop_w<T, false>(
              lo, hi, t0, stride, W, a.op, [&](int r, uint64_t j) { return byte_ptr(c.data[r], soff) + j * 16; },
              [&](uint64_t j, uint4 v) {
                store_to_ranks(c, dest_root, -1, [&](int r) { return byte_ptr(c.data[r], soff) + j * 16; }, v);
              });
        }
        if (tr) tr[2] = globaltimer_ns();
        done(kPipeLink, s, uint32_t(a.n_link) * (u + 1), U);
        if (tr) tr[3] = globaltimer_ns();
      } else {
        if (direct) wait_peers(kPipeIn, s, U, a.root, -1);
        else wait_peers(kPipeLink, s, U, -1, -1);
        if (tr) tr[1] = globaltimer_ns();
        const uint8_t* slot = byte_ptr(c.data[me], soff);
        vec_loop<8>(0, cv, t0, stride, [&](uint64_t j) { return ld_volatile(slot + j * 16); },
                    [&](uint64_t j, uint4 v) { st_stream(dst + (v0 + j) * 16, v); });
        if (tr) tr[2] = globaltimer_ns();
        done(kPipeOut, s, uint32_t(n_out) * (u + 1), U);
        if (tr) tr[3] = globaltimer_ns();
      }
    }

  ctx_finish(x);
  if (last_block_done(x)) {   // every CTA of this rank is through: roll the slot bases forward
    if (int(threadIdx.x) < S) {
      const uint64_t uses = nchunks / S + (uint64_t(threadIdx.x) < nchunks % S ? 1 : 0);
      pad[kCtlBase + kCtlPipeBase + threadIdx.x] = base_s[threadIdx.x] + uint32_t(uses);
    }
    if (threadIdx.x < 3 * kPipeSlots) pad[kCtlBase + kCtlPipeCnt + threadIdx.x] = 0;
  }
}

__global__ void flush_l2_kernel(uint4* p, uint64_t nvec, uint32_t salt) {
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride)
    p[i] = make_uint4(salt, uint32_t(i), salt ^ uint32_t(i), 0);
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
// a CTA per SM so that enough 16-byte requests are in flight to cover HBM and NVLink latency.
Shape pick_shape(const fmpi_devcomm* c, const fmpi_launch* l, uint64_t nbytes) {
  Shape s;
  // the cross-GPU barrier signals with threads 0..world-1: never fewer than a warp, whole warps
  s.threads = (l && l->threads > 0 && l->threads <= 512) ? ((l->threads + 31) / 32 * 32) : 512;
  if (l && l->blocks > 0) {
    s.blocks = l->blocks;
  } else {
    const uint64_t per_block = uint64_t(s.threads) * 16 * 4;
    uint64_t b = (nbytes + per_block - 1) / per_block;
    const uint64_t cap = nbytes >= (uint64_t(4) << 20) ? 148 : nbytes >= (uint64_t(1) << 20) ? 64 : 24;
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

bool nvls_capable(const fmpi_devcomm* c, int dtype, int op) {
  return c->mc_data && !c->emulate && c->world > 1 && op == FMPI_SUM &&
         (dtype == FMPI_F32 || dtype == FMPI_BF16 || dtype == FMPI_F16);
}

// AUTO defaults from the measured grid (profiles/tune_*.json): on a symmetric region the
// in-switch path wins from W = 4 up (665 vs 624 GB/s bus bandwidth at 256 MB, with 16 CTAs);
// through staging the P2P two-shot wins at W <= 4 because the NVLS variant pays a copy-in.
// The Python tuner (parallel/tuner.py) overrides this with the table measured on the box.
bool want_nvls(const fmpi_devcomm* c, int algo, int dtype, int op, uint64_t nbytes, bool symm) {
  if (algo == FMPI_ALGO_NVLS) return true;
  if (algo != FMPI_ALGO_AUTO || !nvls_capable(c, dtype, op)) return false;
  return symm ? (c->world >= 4 && nbytes >= (256u << 10)) : (c->world >= 8 && nbytes >= (4u << 20));
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

#define FMPI_DISPATCH_NVLS(dt, EXPR)                                    \
  switch (dt) {                                                         \
    case FMPI_F32: { using T = float; EXPR; } break;                    \
    case FMPI_F16: { using T = __half; EXPR; } break;                   \
    case FMPI_BF16: { using T = __nv_bfloat16; EXPR; } break;           \
    default: rc = set_error("NVLS supports f32 / f16 / bf16 sums, got dtype %d", dt); \
  }

int pipe_impl(const fmpi_devcomm* c, const void* const* in, void* const* out, uint64_t nvec, int dtype,
              int op, int kind, int root, uint64_t stage_off, uint64_t stage_bytes, const fmpi_launch* l,
              cudaStream_t st);
struct PipeCfg;
const PipeCfg& pipe_cfg();
bool pipe_wanted(const fmpi_devcomm* c, int algo, uint64_t nbytes, int kind);

// The tail of every staging region belongs to the pipelined kernels (at most a quarter of it,
// at most 48 MB): their slots are read and written by peers outside the other families' barriers.
uint64_t pipe_carve(uint64_t stage_bytes) {
  uint64_t p = stage_bytes / 4;
  if (p > (48ull << 20)) p = 48ull << 20;
  return p & ~uint64_t(255);
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
  Shape s = pick_shape(c, l, nbytes / c->world);
  int rc = 0;
  if (want_nvls(c, l ? l->algo : FMPI_ALGO_AUTO, dtype, op, nbytes, true)) {
    if (!nvls_capable(c, dtype, op)) return set_error("NVLS requested but not possible for this call");
    if (!(l && l->blocks > 0) && s.blocks > 16) s.blocks = 16;  // the switch does the work: few CTAs suffice
    FMPI_DISPATCH_NVLS(dtype, rc = launch_ranked(nvls_symm_kernel<T>, *c, s.blocks, s.threads, st, *c,
                                                 off, nvec, dest_root));
    return rc;
  }
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
  const int n_ranks0 = c->emulate ? c->world : 1;
  for (int i = 0; i < n_ranks0; ++i)
    if ((reinterpret_cast<uintptr_t>(in[i]) % 16) || (reinterpret_cast<uintptr_t>(out[i]) % 16))
      return set_error("collective buffers must be 16-byte aligned");
  const uint64_t pipe_bytes = pipe_carve(stage_bytes);
  stage_bytes -= pipe_bytes;
  if (nbytes % 16 == 0 && pipe_wanted(c, algo, nbytes, dest_root < 0 ? 0 : 1)) {
    if (algo == FMPI_ALGO_NVLS && !nvls_capable(c, dtype, op)) return set_error("NVLS requested but not possible for this call");
    return pipe_impl(c, in, out, nbytes / 16, dtype, op, dest_root < 0 ? 0 : 1, dest_root < 0 ? 0 : dest_root,
                     stage_off + stage_bytes, pipe_bytes, l, st);
  }
  // The staging area is split into private parts so that the push-without-handshake one-shot
  // kernel can never overwrite bytes a peer is still copying out of a two-shot call:
  //   [one-shot: 2 parities x W slots][two-shot: scatter W x per | result W x per]
  // (the NVLS kernel uses the two-shot part; both start with a barrier-protected phase).
  uint64_t os_bytes = 2ull * c->world * ((256ull << 10) + 16);
  if (os_bytes > stage_bytes / 4) os_bytes = stage_bytes / 4;
  os_bytes &= ~uint64_t(15);
  const uint64_t slot = (os_bytes / (2ull * c->world)) & ~uint64_t(15);
  const bool want_oneshot =
      algo == FMPI_ALGO_ONESHOT || (algo == FMPI_ALGO_AUTO && nbytes <= (256u << 10));
  const bool aligned = nbytes % 16 == 0;
  const int n_ranks = c->emulate ? c->world : 1;
  int rc = 0;
  for (int i = 0; i < n_ranks; ++i)
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
  const uint64_t ts_off = stage_off + os_bytes, ts_bytes = stage_bytes - os_bytes;
  const bool nvls = want_nvls(c, algo, dtype, op, nbytes, false);
  if (nvls && !nvls_capable(c, dtype, op)) return set_error("NVLS requested but not possible for this call");
  // Chunked so that a chunk fits: P2P uses scatter (N, first half) + result (N, second half).
  // NVLS needs only N and lives in the SECOND half: a P2P call's unsynchronised scatter pushes
  // go to the first half, so they can never hit bytes a peer is still copying out of an
  // earlier NVLS call.
  const uint64_t nvls_off = ts_off + ((ts_bytes / 2) & ~uint64_t(15));
  uint64_t max_vec = ts_bytes / 32;
  max_vec = max_vec / c->world * c->world;
  if (max_vec < 2ull * c->world) return set_error("staging area too small");
  max_vec -= c->world;  // room for the per-slice round-up
  const uint64_t nvec = nbytes / 16;
  for (uint64_t done = 0; done < nvec && rc == 0; done += max_vec) {
    const uint64_t n = nvec - done < max_vec ? nvec - done : max_vec;
    PtrTable ti{}, to{};
    for (int i = 0; i < n_ranks; ++i) {
      ti.p[i] = const_cast<uint8_t*>(static_cast<const uint8_t*>(in[i]) + done * 16);
      to.p[i] = static_cast<uint8_t*>(out[i]) + done * 16;
    }
    const Shape s = pick_shape(c, l, n * 16);
    if (nvls) {
      FMPI_DISPATCH_NVLS(dtype, rc = launch_ranked(nvls_stage_kernel<T>, *c, s.blocks, s.threads, st, *c,
                                                   ti, to, n, nvls_off, dest_root));
    } else {
      FMPI_DISPATCH_DTYPE(dtype, rc = launch_ranked(twoshot_stage_kernel<T>, *c, s.blocks, s.threads,
                                                    st, *c, ti, to, n, op, ts_off, dest_root));
    }
  }
  return rc;
}

// ---- chunk pipeline: launch parameters (env overrides exist for the tuning runs only)
struct PipeCfg { int link_nvls, link_p2p, copy_nvls, copy_p2p, slots, nvls_min_world; uint64_t chunk_bytes, min_ar, min_other; };
const PipeCfg& pipe_cfg() {
  static const PipeCfg cfg = [] {
    auto env = [](const char* k, long d) { const char* v = getenv(k); return v && *v ? atol(v) : d; };
    PipeCfg p;
    p.link_nvls = int(env("FMPI_PIPE_LINK_NVLS", 16));
    p.link_p2p = int(env("FMPI_PIPE_LINK_P2P", 64));
    p.copy_nvls = int(env("FMPI_PIPE_COPY_NVLS", 64));
    p.copy_p2p = int(env("FMPI_PIPE_COPY_P2P", 80));
    p.slots = int(env("FMPI_PIPE_SLOTS", 3));
    if (p.slots < 2) p.slots = 2;
    if (p.slots > kPipeSlots) p.slots = kPipeSlots;
    // the in-switch link phase moves (W+1)/W * N per direction against 2(W-1)/W * N for the
    // peer-to-peer one: a win from four ranks up only
    p.nvls_min_world = int(env("FMPI_PIPE_NVLS_MIN_WORLD", 4));
    p.chunk_bytes = uint64_t(env("FMPI_PIPE_CHUNK_KB", 16384)) << 10;
    p.min_ar = uint64_t(env("FMPI_PIPE_MIN_AR_KB", 48 << 10)) << 10;
    p.min_other = uint64_t(env("FMPI_PIPE_MIN_KB", 8 << 10)) << 10;
    return p;
  }();
  return cfg;
}

bool pipe_wanted(const fmpi_devcomm* c, int algo, uint64_t nbytes, int kind) {
  if (c->world < 2) return false;
  if (algo == FMPI_ALGO_PIPE) return true;
  if (algo != FMPI_ALGO_AUTO) return false;
  // Opt-in only (FMPI_PIPE_AUTO=1 lets AUTO choose it from min_ar / min_other bytes up). Measured
  // on 8 B200s (profiles/runs/r2/pipe_trace_*_8.log): per 16 MB chunk the roles work 25 us (in),
  // 13 us (link, in-switch) and 10 us (out) but each waits 15-50 us for the slowest of the eight
  // ranks at every hand-over, so a 256 MB all-reduce reaches 367-418 GB/s bus bandwidth where the
  // single staged kernel reaches ~500 (and the in-place symmetric kernel 811). At 2 GPUs with the
  // peer-to-peer link phase the three roles share the SMs' load/store bandwidth: 200-300 GB/s
  // against 320-460 staged. The kernel stays as a selectable algorithm (FMPI_ALGO_PIPE).
  const PipeCfg& cfg = pipe_cfg();
  static const bool auto_on = [] { const char* v = getenv("FMPI_PIPE_AUTO"); return v && *v && *v != '0'; }();
  if (!auto_on) return false;
  if (!(c->mc_data && !c->emulate && c->world >= cfg.nvls_min_world)) return false;
  return nbytes >= (kind == 0 ? cfg.min_ar : cfg.min_other);
}

// Diagnostics: with FMPI_PIPE_TRACE=1 the first CTA of every role stamps globaltimer per chunk
// (start, after its wait, after its copy / link work, after its hand-over) into this buffer;
// fmpi_pipe_trace_read copies the stamps of the LAST pipelined call to the host.
constexpr size_t kPipeTraceWords = 3 * 256 * 4;
uint64_t* pipe_trace_buffer(bool clear = true) {
  static uint64_t* buf = [] {
    const char* v = getenv("FMPI_PIPE_TRACE");
    uint64_t* p = nullptr;
    if (v && *v && *v != '0') cudaMalloc(&p, kPipeTraceWords * sizeof(uint64_t));
    return p;
  }();
  if (buf && clear) {
    cudaDeviceSynchronize();
    cudaMemset(buf, 0, kPipeTraceWords * sizeof(uint64_t));
  }
  return buf;
}

// kind: 0 all-reduce, 1 reduce, 2 broadcast. The region [stage_off, stage_off + stage_bytes) is
// private to the pipelined kernels of this channel (their slots are read by peers without a
// handshake with the other kernel families).
int pipe_impl(const fmpi_devcomm* c, const void* const* in, void* const* out, uint64_t nvec, int dtype,
              int op, int kind, int root, uint64_t stage_off, uint64_t stage_bytes, const fmpi_launch* l,
              cudaStream_t st) {
  const PipeCfg& cfg = pipe_cfg();
  const bool forced = l && l->algo == FMPI_ALGO_NVLS;
  const bool nvls = (kind == 2 ? (c->mc_data && !c->emulate && c->world > 1) : nvls_capable(c, dtype, op)) &&
                    (forced || c->world >= cfg.nvls_min_world);
  PipeArgs a{};
  a.in = table(in, c);
  a.out = table(reinterpret_cast<const void* const*>(out), c);
  a.nvec = nvec; a.kind = kind; a.root = root; a.op = op; a.stage_off = stage_off;
  a.nslots = cfg.slots;
  // chunks: large enough that the per-chunk hand-over (a few microseconds) stays small, small
  // enough that the pipeline fills quickly: about 1/12 of the message, 1 MB .. chunk_bytes
  uint64_t chunk = nvec * 16 / 12;
  if (chunk < (1ull << 20)) chunk = 1ull << 20;
  if (chunk > cfg.chunk_bytes) chunk = cfg.chunk_bytes;
  if (chunk * a.nslots > stage_bytes) chunk = stage_bytes / a.nslots;
  a.chunk_vec = chunk / 16 / c->world * c->world;
  if (a.chunk_vec < uint64_t(c->world)) return set_error("pipeline staging area too small");
  int total = (l && l->blocks > 0) ? l->blocks : 0;
  if (total > 0) {
    a.n_link = nvls ? total / 4 : total / 2;
  } else {
    a.n_link = nvls ? cfg.link_nvls : cfg.link_p2p;
    total = a.n_link + (nvls ? cfg.copy_nvls : cfg.copy_p2p);
    if (!nvls && c->world <= 2 && kind != 2) {   // two ranks: the link role carries half the bytes itself
      a.n_link = cfg.link_p2p + cfg.copy_p2p / 2;
      total = cfg.link_p2p + cfg.copy_p2p;
    }
  }
  int cap = 144;                                   // co-resident: one CTA per SM, a few SMs spare
  if (c->emulate) cap = 128 / c->world;
  if (total > cap) { a.n_link = a.n_link * cap / total; total = cap; }
  if (a.n_link < 1) a.n_link = 1;
  if (kind == 2 && c->world <= 2) a.n_link = 0;     // nobody to re-emit to: the root sends directly
  if (total - a.n_link < 2) total = a.n_link + 2;
  a.n_copy = total - a.n_link;
  int threads = (l && l->threads > 0 && l->threads <= 512) ? (l->threads + 31) / 32 * 32 : 512;
  a.trace = pipe_trace_buffer();
  int rc = 0;
  if (nvls) {
    FMPI_DISPATCH_NVLS(kind == 2 ? FMPI_F32 : dtype,
                       rc = launch_ranked(pipe_kernel<T, true>, *c, total, threads, st, *c, a));
  } else {
    FMPI_DISPATCH_DTYPE(kind == 2 ? FMPI_U8 : dtype,
                        rc = launch_ranked(pipe_kernel<T, false>, *c, total, threads, st, *c, a));
  }
  return rc;
}

// 0 direct, 1 scatter + all-gather, 2 multicast from the root, 3 pull + re-emit (symmetric only)
int bcast_mode(const fmpi_devcomm* c, int algo, uint64_t nbytes, bool symm = false) {
  const bool mc = c->mc_data && !c->emulate;
  if (algo == FMPI_ALGO_PIPE) return symm ? 3 : (c->world > 2 ? 1 : 0);
  if (algo == FMPI_ALGO_PULL_P2P) return symm ? 4 : (c->world > 2 ? 1 : 0);
  if (algo == FMPI_ALGO_NVLS) return mc ? 2 : -1;
  if (algo == FMPI_ALGO_ONESHOT) return 0;
  if (algo == FMPI_ALGO_TWOSHOT) return c->world > 2 ? 1 : 0;
  // large symmetric messages: every rank fetches its slice from the root and re-emits it, so
  // all W links send at once (the root's multicast alone tops out near 350-390 GB/s)
  if (symm && c->world > 2 && nbytes >= (8u << 20)) return 3;
  if (mc && c->world > 2) return 2;
  return (nbytes >= (1u << 20) && c->world > 2) ? 1 : 0;
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
  const int mode = bcast_mode(c, l ? l->algo : FMPI_ALGO_AUTO, nbytes, true);
  if (mode < 0) return set_error("multicast broadcast requested but no multicast mapping");
  Shape s = pick_shape(c, l, mode >= 3 ? nbytes / c->world : nbytes);
  // pull + multicast re-emit: the switch replicates, 16 CTAs saturate it and more only add
  // contention (8 GPUs, 1 GB: 513 GB/s with 16 CTAs, 481 with 64, 393 with 148; symm_probe_8.json)
  if (mode == 3 && c->mc_data && !c->emulate && !(l && l->blocks > 0) && s.blocks > 16) s.blocks = 16;
  return launch_ranked(bcast_symm_kernel, *c, s.blocks, s.threads, static_cast<cudaStream_t>(st),
                       *c, off, nvec, root, mode);
}

int fmpi_bcast_any(const fmpi_devcomm* c, void* const* buf, uint64_t nbytes, int root,
                   uint64_t stage_off, uint64_t stage_bytes, const fmpi_launch* l,
                   fmpi_stream st) {
  if (validate(c)) return 1;
  if (root < 0 || root >= c->world) return set_error("bad root %d", root);
  if (stage_off % 16) return set_error("staging offset must be 16-byte aligned");
  if (stage_off + stage_bytes > c->data_bytes) return set_error("staging exceeds the window");
  if (nbytes == 0) return 0;
  const uint64_t pipe_bytes = pipe_carve(stage_bytes);
  stage_bytes -= pipe_bytes;
  if (nbytes % 16 == 0 && pipe_wanted(c, l ? l->algo : FMPI_ALGO_AUTO, nbytes, 2)) {
    const int n = c->emulate ? c->world : 1;
    for (int i = 0; i < n; ++i)
      if (reinterpret_cast<uintptr_t>(buf[i]) % 16) return set_error("bcast buffer must be 16-byte aligned");
    return pipe_impl(c, reinterpret_cast<const void* const*>(buf), buf, nbytes / 16, FMPI_U8, FMPI_SUM, 2, root,
                     stage_off + stage_bytes, pipe_bytes, l, static_cast<cudaStream_t>(st));
  }
  const uint64_t half = (stage_bytes / 2) & ~uint64_t(15);
  if (half < 32) return set_error("staging area too small");
  const uint64_t chunk = half - 16;
  const int algo = l ? l->algo : FMPI_ALGO_AUTO;
  const int n_ranks = c->emulate ? c->world : 1;
  int rc = 0;
  for (uint64_t done = 0; done < nbytes && rc == 0; done += chunk) {
    const uint64_t n = nbytes - done < chunk ? nbytes - done : chunk;
    PtrTable t{};
    for (int i = 0; i < n_ranks; ++i) {
      t.p[i] = static_cast<uint8_t*>(buf[i]) + done;
      if (reinterpret_cast<uintptr_t>(t.p[i]) % 16) return set_error("bcast buffer must be 16-byte aligned");
    }
    const int mode = bcast_mode(c, algo, n);
    if (mode < 0) return set_error("multicast broadcast requested but no multicast mapping");
    const Shape s = pick_shape(c, l, n);
    rc = launch_ranked(bcast_stage_kernel, *c, s.blocks, s.threads, static_cast<cudaStream_t>(st),
                       *c, t, n, root, stage_off, half, mode);
  }
  return rc;
}

int fmpi_pipe_trace_read(uint64_t* host_out, size_t nwords) {
  uint64_t* b = pipe_trace_buffer(false);
  if (!b) return set_error("FMPI_PIPE_TRACE is not set");
  if (nwords > kPipeTraceWords) nwords = kPipeTraceWords;
  if (check_cuda(cudaDeviceSynchronize(), "sync")) return 1;
  return check_cuda(cudaMemcpy(host_out, b, nwords * sizeof(uint64_t), cudaMemcpyDeviceToHost), "trace copy");
}

int fmpi_flush_l2(void* buf, uint64_t nbytes, fmpi_stream st) {
  static uint32_t salt = 0;
  flush_l2_kernel<<<148 * 4, 512, 0, static_cast<cudaStream_t>(st)>>>(static_cast<uint4*>(buf),
                                                                     nbytes / 16, ++salt);
  return check_cuda(cudaGetLastError(), "flush_l2");
}

}  // extern "C"
