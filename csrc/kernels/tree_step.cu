// Tree-level fused kernels: the functional API (`update!(state, params, grads)` with a
// DistributedOptimizer, `allreduce_gradients(tree)`) on leaves that live wherever the autodiff
// put them — Zygote-owned arrays on the Julia side, `.grad` tensors of an arbitrary module here.
//
// Reference path being replaced (/root/reference/src/optimizer.jl:20-23 and :45-65): per leaf a
// host-staged MPI all-reduce, then the rule's elementwise kernels. Round 1 did
// pack -> all-reduce -> unpack -> multi_tensor_step (4 launches, the reduced gradient written
// twice). Here:
//
//   tree_oneshot_step   ONE launch for trees up to the push-slot size: every rank gathers its
//                       gradient leaves through the pointer table and pushes them into the
//                       peers' parity slots, one cross-GPU barrier, then every rank sums the W
//                       copies in rank order and applies the rule straight to its scattered
//                       parameter / moment leaves. The reduced gradient never exists in memory.
//   tree_allreduce      larger trees: gather the leaves into the symmetric arena (local copy), a
//                       grid-wide + cross-GPU barrier, in-place two-shot or in-switch all-reduce
//                       of the arena — one launch; the rule kernel (fmpi_multi_tensor_step) or
//                       the unpack then reads the arena through a pointer table (second launch).
//
// Work distribution (both kernels): leaves are cut into fixed chunks; chunk k of the concatenated
// chunk list belongs to block k mod gridDim.x on every rank and in every phase, so a block only
// ever depends on the same-index block of its peers (= the barrier it takes part in).
#include "common.cuh"

namespace fmpi {

constexpr uint64_t kTreeChunkBytes = 32768;

// device table of one rank: int64 [p[n] | g[n] | m[n] | v[n] | numel[n] | off_bytes[n]]
struct TreeTab {
  const uint64_t* t[FMPI_MAX_RANKS];
};

struct TreeMath {
  int rule;
  float eta, b1, b2, eps, lambda, rho, bc1, bc2, gscale;
  __device__ __forceinline__ float step(float p, float g, float& m, float& v) const {
    g *= gscale;
    switch (rule) {
      case FMPI_RULE_DESCENT: return p - eta * g;
      case FMPI_RULE_MOMENTUM: m = rho * m + eta * g; return p - m;
      case FMPI_RULE_NESTEROV: {
        const float d = -rho * rho * m + (1.f + rho) * eta * g;
        m = rho * m - eta * g;
        return p - d;
      }
      default: {
        m = b1 * m + (1.f - b1) * g;
        v = b2 * v + (1.f - b2) * g * g;
        const float d = (m / bc1) / (sqrtf(v / bc2) + eps) * eta;
        return rule == FMPI_RULE_ADAMW ? p - (d + lambda * p) : p - d;
      }
    }
  }
};

template <typename T> __device__ __forceinline__ float tload(const T* p) { return float(*p); }
template <> __device__ __forceinline__ float tload(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <> __device__ __forceinline__ float tload(const __half* p) { return __half2float(*p); }
template <typename T> __device__ __forceinline__ void tstore(T* p, float v) { *p = T(v); }
template <> __device__ __forceinline__ void tstore(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ void tstore(__half* p, float v) { *p = __float2half_rn(v); }

// Walk the chunks of leaf list that belong to this block: fn(leaf index, first element, count).
template <typename T, typename Fn>
__device__ __forceinline__ void for_my_chunks(int n, const uint64_t* numel, Fn fn) {
  constexpr uint64_t CE = kTreeChunkBytes / sizeof(T);
  uint64_t chunk0 = 0;
  const uint64_t G = gridDim.x;
  for (int t = 0; t < n; ++t) {
    const uint64_t ne = numel[t];
    const uint64_t nch = (ne + CE - 1) / CE;
    const uint64_t first = (blockIdx.x + G - chunk0 % G) % G;
    for (uint64_t c = first; c < nch; c += G) {
      const uint64_t lo = c * CE;
      fn(t, lo, min(CE, ne - lo));
    }
    chunk0 += nch;
  }
}

template <typename T>
__global__ void __launch_bounds__(512)
tree_oneshot_step_kernel(fmpi_devcomm c, int n, TreeTab tabs, uint64_t stage_off, uint64_t slot_bytes, TreeMath sm) {
  using V = VecOps<T>;
  constexpr int N = V::kN;
  Ctx x;
  ctx_init(x, c);
  const int me = x.rank, W = c.world;
  const uint64_t* tab = tabs.t[arg_index(c)];
  const uint64_t* ps = tab;
  const uint64_t* gs = tab + n;
  const uint64_t* ms = tab + 2 * n;
  const uint64_t* vs = tab + 3 * n;
  const uint64_t* numel = tab + 4 * n;
  const uint64_t* offs = tab + 5 * n;
  const uint32_t seq = seq_read(x);
  const uint64_t parity = seq & 1u;
  const uint64_t my_slot = stage_off + (parity * W + me) * slot_bytes;
  const bool has_m = sm.rule != FMPI_RULE_DESCENT;
  const bool has_v = sm.rule == FMPI_RULE_ADAM || sm.rule == FMPI_RULE_ADAMW;

  // ---- phase 1: gather my gradient leaves and push them into every peer's slot [parity][me]
  for_my_chunks<T>(n, numel, [&](int t, uint64_t lo, uint64_t cnt) {
    const T* g = reinterpret_cast<const T*>(gs[t]) + lo;
    const uint64_t foff = offs[t] + lo * sizeof(T);       // 16-byte aligned: chunks are 32 KB
    const bool vec = (reinterpret_cast<uintptr_t>(g) & 15) == 0;
    const uint64_t nv = vec ? cnt / N : 0;
    for (uint64_t i = threadIdx.x; i < nv; i += blockDim.x) {
      const uint4 v = ld_stream(reinterpret_cast<const uint8_t*>(g) + i * 16);
#pragma unroll
      for (int r = 0; r < FMPI_MAX_RANKS; ++r)
        if (r < W && r != me) st_stream(byte_ptr(c.data[r], my_slot + foff) + i * 16, v);
    }
    for (uint64_t i = nv * N + threadIdx.x; i < cnt; i += blockDim.x) {
      const T e = g[i];
#pragma unroll
      for (int r = 0; r < FMPI_MAX_RANKS; ++r)
        if (r < W && r != me) reinterpret_cast<T*>(byte_ptr(c.data[r], my_slot + foff))[i] = e;
    }
  });
  block_barrier(x, c);

  // ---- phase 2: rank-ordered sum of the W copies (my own straight from the leaf), rule applied
  //      to the scattered parameter / moment leaves
  const uint8_t* slots = byte_ptr(c.data[me], stage_off + parity * W * slot_bytes);
  for_my_chunks<T>(n, numel, [&](int t, uint64_t lo, uint64_t cnt) {
    T* p = reinterpret_cast<T*>(ps[t]) + lo;
    const T* g = reinterpret_cast<const T*>(gs[t]) + lo;
    float* m = has_m ? reinterpret_cast<float*>(ms[t]) + lo : nullptr;
    float* v = has_v ? reinterpret_cast<float*>(vs[t]) + lo : nullptr;
    const uint64_t foff = offs[t] + lo * sizeof(T);
    const bool vec = ((reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(p) |
                       (has_m ? reinterpret_cast<uintptr_t>(m) : 0) | (has_v ? reinterpret_cast<uintptr_t>(v) : 0)) & 15) == 0;
    const uint64_t nv = vec ? cnt / N : 0;
    for (uint64_t i = threadIdx.x; i < nv; i += blockDim.x) {
      uint4 raw[FMPI_MAX_RANKS];
#pragma unroll
      for (int r = 0; r < FMPI_MAX_RANKS; ++r)
        if (r < W)
          raw[r] = r == me ? ld_stream(reinterpret_cast<const uint8_t*>(g) + i * 16)
                           : ld_volatile(slots + uint64_t(r) * slot_bytes + foff + i * 16);
      float ge[N], b[N], pe[N], me_[N], ve[N];
      V::unpack(raw[0], ge);
#pragma unroll
      for (int r = 1; r < FMPI_MAX_RANKS; ++r)
        if (r < W) {
          V::unpack(raw[r], b);
#pragma unroll
          for (int k = 0; k < N; ++k) ge[k] += b[k];
        }
      V::unpack(*reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(p) + i * 16), pe);
#pragma unroll
      for (int k = 0; k < N; k += 4) {
        if (has_m) *reinterpret_cast<float4*>(me_ + k) = *reinterpret_cast<const float4*>(m + i * N + k);
        if (has_v) *reinterpret_cast<float4*>(ve + k) = *reinterpret_cast<const float4*>(v + i * N + k);
      }
#pragma unroll
      for (int k = 0; k < N; ++k) {
        float mm = has_m ? me_[k] : 0.f, vv = has_v ? ve[k] : 0.f;
        pe[k] = sm.step(pe[k], ge[k], mm, vv);
        me_[k] = mm;
        ve[k] = vv;
      }
      *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(p) + i * 16) = V::pack(pe);
#pragma unroll
      for (int k = 0; k < N; k += 4) {
        if (has_m) *reinterpret_cast<float4*>(m + i * N + k) = *reinterpret_cast<float4*>(me_ + k);
        if (has_v) *reinterpret_cast<float4*>(v + i * N + k) = *reinterpret_cast<float4*>(ve + k);
      }
    }
    for (uint64_t i = nv * N + threadIdx.x; i < cnt; i += blockDim.x) {
      float acc = 0.f;
      for (int r = 0; r < W; ++r) {
        T raw_e;
        if (r == me) raw_e = g[i];
        else raw_e = __ldcv(reinterpret_cast<const T*>(slots + uint64_t(r) * slot_bytes + foff) + i);  // a peer wrote it: no L1
        const float e = tload(&raw_e);
        acc = r == 0 ? e : acc + e;
      }
      float mm = has_m ? m[i] : 0.f, vv = has_v ? v[i] : 0.f;
      const float pn = sm.step(tload(p + i), acc, mm, vv);
      tstore(p + i, pn);
      if (has_m) m[i] = mm;
      if (has_v) v[i] = vv;
    }
  });
  seq_advance(x, seq);
  ctx_finish(x);
}

// Local grid-wide barrier (all CTAs of this rank's launch), built on one control word that the
// last block of the launch resets. Every CTA must be resident (grids <= FMPI_MAX_BLOCKS).
__device__ __forceinline__ void grid_barrier(Ctx& x, const fmpi_devcomm& c, int which) {
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* w = x.my_pad + kCtlBase + kCtlGrid;
    __threadfence();
    atomicAdd(w, 1u);
    const uint32_t target = uint32_t(which + 1) * gridDim.x;
    if (!x.dead) x.dead = !wait_flag(w, target, c);
    __threadfence();
  }
  __syncthreads();
}

// Gather the leaves into the symmetric arena at `off`, then all-reduce the arena in place
// (two-shot peer-to-peer, or in-switch when kNvls). dir_pack = 0 would skip the gather.
template <typename T, bool kNvls>
__global__ void __launch_bounds__(512)
tree_allreduce_kernel(fmpi_devcomm c, int n, TreeTab tabs, uint64_t off, uint64_t nvec, int op) {
  Ctx x;
  ctx_init(x, c);
  const int me = x.rank;
  const uint64_t* tab = tabs.t[arg_index(c)];
  const uint64_t* gs = tab + n;
  const uint64_t* numel = tab + 4 * n;
  const uint64_t* offs = tab + 5 * n;
  uint8_t* arena = byte_ptr(c.data[me], off);
  constexpr int N = VecOps<T>::kN;

  for_my_chunks<T>(n, numel, [&](int t, uint64_t lo, uint64_t cnt) {
    const T* g = reinterpret_cast<const T*>(gs[t]) + lo;
    uint8_t* d = arena + offs[t] + lo * sizeof(T);
    const bool vec = (reinterpret_cast<uintptr_t>(g) & 15) == 0;
    const uint64_t nv = vec ? cnt / N : 0;
    for (uint64_t i = threadIdx.x; i < nv; i += blockDim.x)
      *reinterpret_cast<uint4*>(d + i * 16) = ld_stream(reinterpret_cast<const uint8_t*>(g) + i * 16);
    for (uint64_t i = nv * N + threadIdx.x; i < cnt; i += blockDim.x) reinterpret_cast<T*>(d)[i] = g[i];
  });
  grid_barrier(x, c, 0);     // all of MY leaves are in my arena ...
  block_barrier(x, c);       // ... and, block by block, so are every peer's

  const uint64_t per = (nvec + c.world - 1) / c.world;
  const uint64_t lo = min(per * me, nvec), hi = min(lo + per, nvec);
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t t0 = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if constexpr (kNvls) {
    uint8_t* mc = byte_ptr(c.mc_data, off);
    for (uint64_t i = lo + t0; i < hi; i += 4 * stride) {
      uint4 v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (i + k * stride < hi) v[k] = Multimem<T>::ld_reduce(mc + (i + k * stride) * 16);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (i + k * stride < hi) multimem_st(mc + (i + k * stride) * 16, v[k]);
    }
  } else {
    using V = VecOps<T>;
    for (uint64_t i = lo + t0; i < hi; i += stride) {
      uint4 raw[FMPI_MAX_RANKS];
#pragma unroll
      for (int r = 0; r < FMPI_MAX_RANKS; ++r)
        if (r < c.world) raw[r] = ld_stream(byte_ptr(c.data[r], off) + i * 16);
      typename V::Acc acc[V::kN], b[V::kN];
      V::unpack(raw[0], acc);
#pragma unroll
      for (int r = 1; r < FMPI_MAX_RANKS; ++r)
        if (r < c.world) {
          V::unpack(raw[r], b);
#pragma unroll
          for (int e = 0; e < V::kN; ++e) acc[e] = op_apply(acc[e], b[e], op);
        }
      const uint4 out = V::pack(acc);
#pragma unroll
      for (int r = 0; r < FMPI_MAX_RANKS; ++r)
        if (r < c.world) st_stream(byte_ptr(c.data[r], off) + i * 16, out);
    }
  }
  block_barrier(x, c);       // every slice has landed everywhere
  ctx_finish(x);
  if (last_block_done(x) && threadIdx.x == 0) x.my_pad[kCtlBase + kCtlGrid] = 0;
}

}  // namespace fmpi

using namespace fmpi;
using fmpi_host::check_cuda;
using fmpi_host::launch_ranked;
using fmpi_host::set_error;

namespace {
int tree_common(const fmpi_devcomm* c, int n, const void* const* tables, int dtype, TreeTab* out) {
  if (!c || !tables) return set_error("null argument");
  if (c->world < 2 || c->world > FMPI_MAX_RANKS) return set_error("tree kernels need 2..%d ranks", FMPI_MAX_RANKS);
  if (n <= 0) return set_error("empty leaf table");
  if (dtype != FMPI_F32 && dtype != FMPI_BF16 && dtype != FMPI_F16)
    return set_error("tree kernels support f32 / bf16 / f16 leaves, got dtype %d", dtype);
  const int nr = c->emulate ? c->world : 1;
  for (int i = 0; i < nr; ++i) {
    if (!tables[i]) return set_error("null leaf table for rank slot %d", i);
    out->t[i] = static_cast<const uint64_t*>(tables[i]);
  }
  return 0;
}
int tree_blocks(const fmpi_devcomm* c, const fmpi_launch* l, uint64_t total_bytes) {
  int blocks = (l && l->blocks > 0) ? l->blocks : int((total_bytes + kTreeChunkBytes - 1) / kTreeChunkBytes);
  if (!(l && l->blocks > 0) && blocks > 64) blocks = 64;
  if (blocks < 1) blocks = 1;
  if (blocks > FMPI_MAX_BLOCKS) blocks = FMPI_MAX_BLOCKS;
  if (c->emulate && blocks > 128 / c->world) blocks = 128 / c->world;
  return blocks;
}
}  // namespace

extern "C" {

int fmpi_tree_step_oneshot(const fmpi_devcomm* c, int n, const void* const* tables, uint64_t total_bytes,
                           int dtype, const fmpi_hyper* h, int64_t step, uint64_t stage_off,
                           uint64_t stage_bytes, const fmpi_launch* l, fmpi_stream st) {
  TreeTab tb{};
  if (tree_common(c, n, tables, dtype, &tb)) return 1;
  if (!h) return set_error("null hyper-parameters");
  if (h->rule < FMPI_RULE_DESCENT || h->rule > FMPI_RULE_ADAMW) return set_error("rule %d has no kernel", h->rule);
  if (stage_off % 16) return set_error("staging offset must be 16-byte aligned");
  if (stage_off + stage_bytes > c->data_bytes) return set_error("staging exceeds the window");
  const uint64_t slot = (stage_bytes / (2ull * c->world)) & ~uint64_t(15);
  if (total_bytes > slot)
    return set_error("tree of %llu bytes exceeds the %llu-byte push slots; use fmpi_tree_allreduce",
                     (unsigned long long)total_bytes, (unsigned long long)slot);
  TreeMath sm;
  sm.rule = h->rule; sm.eta = h->eta; sm.b1 = h->beta1; sm.b2 = h->beta2; sm.eps = h->eps;
  sm.lambda = h->lambda; sm.rho = h->rho; sm.gscale = h->grad_scale;
  sm.bc1 = 1.f - powf(h->beta1, float(step + 1));
  sm.bc2 = 1.f - powf(h->beta2, float(step + 1));
  const int blocks = tree_blocks(c, l, total_bytes);
  cudaStream_t s = static_cast<cudaStream_t>(st);
  switch (dtype) {
    case FMPI_F32: return launch_ranked(tree_oneshot_step_kernel<float>, *c, blocks, 512, s, *c, n, tb, stage_off, slot, sm);
    case FMPI_BF16: return launch_ranked(tree_oneshot_step_kernel<__nv_bfloat16>, *c, blocks, 512, s, *c, n, tb, stage_off, slot, sm);
    default: return launch_ranked(tree_oneshot_step_kernel<__half>, *c, blocks, 512, s, *c, n, tb, stage_off, slot, sm);
  }
}

int fmpi_tree_allreduce(const fmpi_devcomm* c, int n, const void* const* tables, uint64_t off,
                        uint64_t total_bytes, int dtype, int op, const fmpi_launch* l, fmpi_stream st) {
  TreeTab tb{};
  if (tree_common(c, n, tables, dtype, &tb)) return 1;
  if (op < FMPI_SUM || op > FMPI_MAX) return set_error("unsupported op %d", op);
  if (off % 16) return set_error("arena offset must be 16-byte aligned");
  const uint64_t nvec = (total_bytes + 15) / 16;
  if (off + nvec * 16 > c->data_bytes) return set_error("arena exceeds the symmetric window");
  const int algo = l ? l->algo : FMPI_ALGO_AUTO;
  const bool can_nvls = c->mc_data && !c->emulate && op == FMPI_SUM;
  if (algo == FMPI_ALGO_NVLS && !can_nvls) return set_error("NVLS requested but not possible for this call");
  const bool nvls = algo == FMPI_ALGO_NVLS || (algo == FMPI_ALGO_AUTO && can_nvls && c->world >= 4 && total_bytes >= (256u << 10));
  int blocks = (l && l->blocks > 0) ? l->blocks : int((total_bytes / c->world + 32767) / 32768);
  if (!(l && l->blocks > 0) && blocks > 96) blocks = 96;
  if (blocks < 1) blocks = 1;
  if (blocks > FMPI_MAX_BLOCKS) blocks = FMPI_MAX_BLOCKS;
  if (c->emulate && blocks > 128 / c->world) blocks = 128 / c->world;
  cudaStream_t s = static_cast<cudaStream_t>(st);
#define FMPI_TA(T)                                                                                          \
  return nvls ? launch_ranked(tree_allreduce_kernel<T, true>, *c, blocks, 512, s, *c, n, tb, off, nvec, op) \
              : launch_ranked(tree_allreduce_kernel<T, false>, *c, blocks, 512, s, *c, n, tb, off, nvec, op)
  switch (dtype) {
    case FMPI_F32: FMPI_TA(float);
    case FMPI_BF16: FMPI_TA(__nv_bfloat16);
    default: FMPI_TA(__half);
  }
#undef FMPI_TA
}

}  // extern "C"
