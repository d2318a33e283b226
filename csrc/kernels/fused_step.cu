// Fused gradient all-reduce (+) optimiser step — the headline kernel.
//
// Reference path being replaced (/root/reference/src/optimizer.jl:20-23): for every parameter
// leaf, a blocking MPI_Allreduce of the gradient (staged D2H/H2D for CuArrays,
// src/mpi_extensions.jl:97-106) followed by the wrapped rule's several elementwise GPU kernels
// (Optimisers.apply! + subtract!). Here ONE launch per bucket does all
// of it over NVLink:
//
//   start barrier                      (every rank's backward has produced this bucket)
//   rank r, for its 1/W slice:         g = sum over ranks of peer gradients   (rank order, fp32)
//                                      rule update of m, v, (master) p in registers
//                                      store new p into EVERY rank's parameter window
//   end barrier                        (all parameter stores have landed)
//   optional: clear the local gradient bucket for the next accumulation
//
// Link traffic per rank and direction: (W-1)/W * N gradient bytes in + (W-1)/W * N parameter
// bytes in — the same as a two-shot all-reduce — but the optimiser's HBM traffic drops from
// 28 B/element to ~8 + 20/W B/element, the reduced gradient never exists in memory, and the
// moments are only ever touched by their owner (lazily replicated, see parallel/engine.py).
// One owner computes each parameter => replicas stay bit-identical.
//
// kNvls = true uses the NVSwitch's in-fabric reduction instead of W peer loads
// (multimem.ld_reduce) and its multicast store instead of W peer stores (multimem.st).
#include "common.cuh"

namespace fmpi {

struct FusedArgs {
  uint64_t g_off, p_off;  // byte offsets of the flat gradient / parameter buffers in the window
  float* m[FMPI_MAX_RANKS];
  float* v[FMPI_MAX_RANKS];
  float* master[FMPI_MAX_RANKS];
  const int64_t* step[FMPI_MAX_RANKS];
  uint64_t lo, hi;  // element range of this bucket
  fmpi_hyper h;
};

template <int kRule>
struct RuleMath {
  float eta, b1, b2, eps, lambda, rho, bc1, bc2, gscale;
  __device__ __forceinline__ void init(const fmpi_hyper& h, int64_t t) {
    eta = h.eta_dev ? *reinterpret_cast<const float*>(h.eta_dev) : h.eta;
    b1 = h.beta1; b2 = h.beta2; eps = h.eps; lambda = h.lambda; rho = h.rho;
    gscale = h.grad_scale;
    if (kRule == FMPI_RULE_ADAM || kRule == FMPI_RULE_ADAMW) {
      const float n = float(t + 1);
      bc1 = 1.f - powf(b1, n);
      bc2 = 1.f - powf(b2, n);
    } else {
      bc1 = bc2 = 1.f;
    }
  }
  // returns the new parameter; updates m, v in place
  __device__ __forceinline__ float step(float p, float g, float& m, float& v) const {
    g *= gscale;
    if (kRule == FMPI_RULE_DESCENT) {
      return p - eta * g;
    } else if (kRule == FMPI_RULE_MOMENTUM) {
      m = rho * m + eta * g;
      return p - m;
    } else if (kRule == FMPI_RULE_NESTEROV) {
      const float d = -rho * rho * m + (1.f + rho) * eta * g;
      m = rho * m - eta * g;
      return p - d;
    } else {
      m = b1 * m + (1.f - b1) * g;
      v = b2 * v + (1.f - b2) * g * g;
      const float d = (m / bc1) / (sqrtf(v / bc2) + eps) * eta;
      if (kRule == FMPI_RULE_ADAMW) return p - (d + lambda * p);
      return p - d;
    }
  }
};

template <int kRule> struct RuleState { static constexpr bool kM = true, kV = false; };
template <> struct RuleState<FMPI_RULE_DESCENT> { static constexpr bool kM = false, kV = false; };
template <> struct RuleState<FMPI_RULE_ADAM> { static constexpr bool kM = true, kV = true; };
template <> struct RuleState<FMPI_RULE_ADAMW> { static constexpr bool kM = true, kV = true; };

// Local optimiser state of U vectors (p or its fp32 master copy, m, v) as registers.
template <typename T, bool kMaster, int kRule, int U>
struct LocalState {
  using V = VecOps<T>;
  static constexpr int N = V::kN;
  float p[U][N], m[U][N], v[U][N];
  uint4 praw[U];

  // issue the loads (no use of the values: the caller puts the wait where it wants it)
  __device__ __forceinline__ void load(uint64_t i, uint64_t stride, uint64_t vhi, uint64_t elem0,
                                       const uint8_t* pwin, const float* mloc, const float* vloc, const float* mst) {
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t j = i + uint64_t(u) * stride;
      if (j < vhi) {
        const uint64_t e = elem0 + j * N;
        if (kMaster) {
#pragma unroll
          for (int k = 0; k < N; k += 4) *reinterpret_cast<float4*>(&p[u][k]) = __ldcs(reinterpret_cast<const float4*>(mst + e + k));
        } else {
          praw[u] = ld_stream(pwin + j * 16);
        }
        if (RuleState<kRule>::kM) {
#pragma unroll
          for (int k = 0; k < N; k += 4) *reinterpret_cast<float4*>(&m[u][k]) = __ldcs(reinterpret_cast<const float4*>(mloc + e + k));
        }
        if (RuleState<kRule>::kV) {
#pragma unroll
          for (int k = 0; k < N; k += 4) *reinterpret_cast<float4*>(&v[u][k]) = __ldcs(reinterpret_cast<const float4*>(vloc + e + k));
        }
      }
    }
  }
  // apply the rule to vector u with gradient g, write m / v / master back, return the new p
  __device__ __forceinline__ uint4 apply(int u, uint64_t j, uint64_t elem0, const float* g, const RuleMath<kRule>& rule,
                                         float* mloc, float* vloc, float* mst) {
    if (!kMaster) V::unpack(praw[u], p[u]);
#pragma unroll
    for (int k = 0; k < N; ++k) p[u][k] = rule.step(p[u][k], g[k], m[u][k], v[u][k]);
    const uint64_t e = elem0 + j * N;
    if (RuleState<kRule>::kM) {
#pragma unroll
      for (int k = 0; k < N; k += 4) __stcs(reinterpret_cast<float4*>(mloc + e + k), *reinterpret_cast<float4*>(&m[u][k]));
    }
    if (RuleState<kRule>::kV) {
#pragma unroll
      for (int k = 0; k < N; k += 4) __stcs(reinterpret_cast<float4*>(vloc + e + k), *reinterpret_cast<float4*>(&v[u][k]));
    }
    if (kMaster) {
#pragma unroll
      for (int k = 0; k < N; k += 4) __stcs(reinterpret_cast<float4*>(mst + e + k), *reinterpret_cast<float4*>(&p[u][k]));
    }
    return V::pack(p[u]);
  }
};

// Work per thread and iteration: U vectors. The in-switch flavour keeps two 16-byte
// multimem.ld_reduce in flight per thread for fp32 (8 elements either way) and fits two CTAs
// on an SM (<= 64 registers), i.e. 4 reduced vectors in flight per thread slot; the
// peer-to-peer flavour holds W raw vectors per element group instead and stays at U = 1.
// The local state of the FIRST iteration is requested before the start barrier — it does not
// depend on the peers, so its HBM latency hides behind the barrier — and the state of
// iteration k+1 is requested right after the stores of iteration k, together with the
// gradient request of k+1, so the link stream and the HBM stream overlap.
template <typename T, bool kMaster, int kRule, bool kNvls>
__global__ void __launch_bounds__(512, kNvls ? 2 : 1)
fused_step_kernel(fmpi_devcomm c, FusedArgs a) {
  using V = VecOps<T>;
  constexpr int N = V::kN;
  constexpr int U = kNvls ? 8 / N : 1;
  Ctx x;
  ctx_init(x, c);
  const int ai = arg_index(c);
  const uint64_t nvec = (a.hi - a.lo) / N;
  const uint64_t per = (nvec + c.world - 1) / c.world;
  const uint64_t vlo = min(per * x.rank, nvec), vhi = min(vlo + per, nvec);
  const uint64_t gbase = a.g_off + a.lo * sizeof(T), pbase = a.p_off + a.lo * sizeof(T);
  float* __restrict__ mloc = a.m[ai];
  float* __restrict__ vloc = a.v[ai];
  float* __restrict__ mst = a.master[ai];
  const uint8_t* pwin = byte_ptr(c.data[x.rank], pbase);
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  uint64_t i = vlo + uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;

  LocalState<T, kMaster, kRule, U> ls;
  ls.load(i, stride, vhi, a.lo, pwin, mloc, vloc, mst);
  RuleMath<kRule> rule;
  rule.init(a.h, a.step[ai] ? *a.step[ai] : 0);
  block_barrier(x, c);      // every rank's backward has produced this bucket

  while (i < vhi) {
    // ---- reduced gradient, fp32, rank order (or summed inside the switch)
    [[maybe_unused]] uint4 graw[U];
    if constexpr (kNvls) {
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i + uint64_t(u) * stride < vhi)
          graw[u] = Multimem<T>::ld_reduce(byte_ptr(c.mc_data, gbase) + (i + uint64_t(u) * stride) * 16);
    } else {
      uint4 raw[FMPI_MAX_RANKS];
#pragma unroll
      for (int r = 0; r < FMPI_MAX_RANKS; ++r)
        if (r < c.world) raw[r] = ld_stream(byte_ptr(c.data[r], gbase) + i * 16);
      float g[N], b[N];
      V::unpack(raw[0], g);
#pragma unroll
      for (int r = 1; r < FMPI_MAX_RANKS; ++r)
        if (r < c.world) {
          V::unpack(raw[r], b);
#pragma unroll
          for (int k = 0; k < N; ++k) g[k] += b[k];
        }
      const uint4 pv = ls.apply(0, i, a.lo, g, rule, mloc, vloc, mst);
#pragma unroll
      for (int r = 0; r < FMPI_MAX_RANKS; ++r)
        if (r < c.world) st_stream(byte_ptr(c.data[r], pbase) + i * 16, pv);
    }
    if constexpr (kNvls) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint64_t j = i + uint64_t(u) * stride;
        if (j < vhi) {
          float g[N];
          V::unpack(graw[u], g);
          multimem_st(byte_ptr(c.mc_data, pbase) + j * 16, ls.apply(u, j, a.lo, g, rule, mloc, vloc, mst));
        }
      }
    }
    i += uint64_t(U) * stride;
    ls.load(i, stride, vhi, a.lo, pwin, mloc, vloc, mst);
  }

  block_barrier(x, c);      // all parameter stores have landed everywhere
  ctx_finish(x);
  if (a.h.zero_grad) {
    // Clear with the readers' block->vector mapping, slice by slice: vector i of slice q was
    // read by block blockIdx.x of rank q, and that is exactly the block the barrier above
    // synchronised with.
    uint4* gl = reinterpret_cast<uint4*>(byte_ptr(c.data[x.rank], gbase));
    for (int q = 0; q < c.world; ++q) {
      const uint64_t qlo = min(per * q, nvec), qhi = min(qlo + per, nvec);
      for (uint64_t j = qlo + uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; j < qhi; j += stride)
        gl[j] = make_uint4(0, 0, 0, 0);
    }
  }
}

// One-shot flavour for small buckets (a whole Quick-Start-sized model): every rank PUSHES its
// gradient bucket into the peers' staging slots [parity][source rank], ONE barrier, then every
// rank sums the W copies in rank order and applies the rule to ALL elements locally. No slice
// ownership, no second barrier, no parameter traffic: latency ~ one NVLink hop + the barrier.
// The optimiser state is replicated (every rank updates all of m, v) and the replicas stay
// bit-identical because every rank adds the same W values in the same order.
// Replaces the per-leaf allreduce! + apply! of /root/reference/src/optimizer.jl:20-23 for
// latency-bound models (README.md:37-38).
template <typename T, bool kMaster, int kRule>
__global__ void __launch_bounds__(512)
fused_step_oneshot_kernel(fmpi_devcomm c, FusedArgs a, uint64_t stage_off, uint64_t slot_bytes) {
  using V = VecOps<T>;
  constexpr int N = V::kN;
  Ctx x;
  ctx_init(x, c);
  const int ai = arg_index(c);
  const int me = x.rank, W = c.world;
  const uint64_t parity = seq_read(x) & 1u;
  const uint32_t seq = seq_read(x);
  const uint64_t nvec = (a.hi - a.lo) / N;
  const uint64_t gbase = a.g_off + a.lo * sizeof(T), pbase = a.p_off + a.lo * sizeof(T);
  float* __restrict__ mloc = a.m[ai];
  float* __restrict__ vloc = a.v[ai];
  float* __restrict__ mst = a.master[ai];
  uint8_t* gl = byte_ptr(c.data[me], gbase);
  uint8_t* pl = byte_ptr(c.data[me], pbase);
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t t0 = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const uint64_t my_slot = stage_off + (parity * W + me) * slot_bytes;

  for (uint64_t j = t0; j < nvec; j += stride) {
    const uint4 v = ld_stream(gl + j * 16);
#pragma unroll
    for (int r = 0; r < FMPI_MAX_RANKS; ++r)
      if (r < W && r != me) st_stream(byte_ptr(c.data[r], my_slot) + j * 16, v);
  }
  RuleMath<kRule> rule;
  rule.init(a.h, a.step[ai] ? *a.step[ai] : 0);
  LocalState<T, kMaster, kRule, 1> ls;
  ls.load(t0, stride, nvec, a.lo, pl, mloc, vloc, mst);     // local: in flight across the barrier
  block_barrier(x, c);

  const uint8_t* slots = byte_ptr(c.data[me], stage_off + parity * W * slot_bytes);
  const bool zero = a.h.zero_grad != 0;
  for (uint64_t j = t0; j < nvec; j += stride) {
    uint4 raw[FMPI_MAX_RANKS];
#pragma unroll
    for (int r = 0; r < FMPI_MAX_RANKS; ++r)
      if (r < W) raw[r] = r == me ? ld_stream(gl + j * 16) : ld_volatile(slots + uint64_t(r) * slot_bytes + j * 16);
    float g[N], b[N];
    V::unpack(raw[0], g);
#pragma unroll
    for (int r = 1; r < FMPI_MAX_RANKS; ++r)
      if (r < W) {
        V::unpack(raw[r], b);
#pragma unroll
        for (int k = 0; k < N; ++k) g[k] += b[k];
      }
    st_stream(pl + j * 16, ls.apply(0, j, a.lo, g, rule, mloc, vloc, mst));
    if (zero) st_stream(gl + j * 16, make_uint4(0, 0, 0, 0));
    ls.load(j + stride, stride, nvec, a.lo, pl, mloc, vloc, mst);
  }
  seq_advance(x, seq);
  ctx_finish(x);
}

// World of one: no peers, no barriers — a pure HBM stream (read g, p, m, v; write p, m, v and
// the cleared g: 32 B/element for fp32 Adam). Written separately so it does not carry the
// 8-rank load arrays of the multi-GPU kernel (88 registers -> 1 CTA/SM, measured 3.4 TB/s):
// U vectors are loaded back to back before any arithmetic to keep enough bytes in flight.
template <typename T, bool kMaster, int kRule>
__global__ void __launch_bounds__(256)
fused_step_local_kernel(fmpi_devcomm c, FusedArgs a) {
  using V = VecOps<T>;
  constexpr int N = V::kN;
  constexpr int U = N == 4 ? 4 : 2;
  RuleMath<kRule> rule;
  rule.init(a.h, a.step[0] ? *a.step[0] : 0);
  const uint64_t nvec = (a.hi - a.lo) / N;
  uint8_t* gp = byte_ptr(c.data[0], a.g_off + a.lo * sizeof(T));
  uint8_t* pp = byte_ptr(c.data[0], a.p_off + a.lo * sizeof(T));
  float* __restrict__ mloc = a.m[0];
  float* __restrict__ vloc = a.v[0];
  float* __restrict__ mst = a.master[0];
  const bool zero = a.h.zero_grad != 0;
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t i0 = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; i0 < nvec; i0 += U * stride) {
    uint4 graw[U], praw[U];
    float p[U][N], m[U][N], v[U][N];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t i = i0 + u * stride;
      if (i < nvec) {
        const uint64_t e = a.lo + i * N;
        graw[u] = ld_stream(gp + i * 16);
        if (kMaster) {
#pragma unroll
          for (int k = 0; k < N; k += 4) *reinterpret_cast<float4*>(&p[u][k]) = *reinterpret_cast<const float4*>(mst + e + k);
        } else {
          praw[u] = ld_stream(pp + i * 16);
        }
        if (RuleState<kRule>::kM) {
#pragma unroll
          for (int k = 0; k < N; k += 4) *reinterpret_cast<float4*>(&m[u][k]) = *reinterpret_cast<const float4*>(mloc + e + k);
        }
        if (RuleState<kRule>::kV) {
#pragma unroll
          for (int k = 0; k < N; k += 4) *reinterpret_cast<float4*>(&v[u][k]) = *reinterpret_cast<const float4*>(vloc + e + k);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint64_t i = i0 + u * stride;
      if (i < nvec) {
        const uint64_t e = a.lo + i * N;
        float g[N];
        V::unpack(graw[u], g);
        if (!kMaster) V::unpack(praw[u], p[u]);
#pragma unroll
        for (int k = 0; k < N; ++k) p[u][k] = rule.step(p[u][k], g[k], m[u][k], v[u][k]);
        if (RuleState<kRule>::kM) {
#pragma unroll
          for (int k = 0; k < N; k += 4) *reinterpret_cast<float4*>(mloc + e + k) = *reinterpret_cast<float4*>(&m[u][k]);
        }
        if (RuleState<kRule>::kV) {
#pragma unroll
          for (int k = 0; k < N; k += 4) *reinterpret_cast<float4*>(vloc + e + k) = *reinterpret_cast<float4*>(&v[u][k]);
        }
        if (kMaster) {
#pragma unroll
          for (int k = 0; k < N; k += 4) *reinterpret_cast<float4*>(mst + e + k) = *reinterpret_cast<float4*>(&p[u][k]);
        }
        st_stream(pp + i * 16, V::pack(p[u]));
        if (zero) st_stream(gp + i * 16, make_uint4(0, 0, 0, 0));
      }
    }
  }
}

__global__ void step_advance_kernel(int64_t* s) { *s += 1; }

}  // namespace fmpi

using namespace fmpi;
using fmpi_host::check_cuda;
using fmpi_host::launch_ranked;
using fmpi_host::set_error;

namespace {

template <typename T, bool kMaster, bool kNvls>
int launch_rule(const fmpi_devcomm& c, const FusedArgs& a, int blocks, int threads, cudaStream_t st) {
  switch (a.h.rule) {
    case FMPI_RULE_DESCENT:
      return launch_ranked(fused_step_kernel<T, kMaster, FMPI_RULE_DESCENT, kNvls>, c, blocks, threads, st, c, a);
    case FMPI_RULE_MOMENTUM:
      return launch_ranked(fused_step_kernel<T, kMaster, FMPI_RULE_MOMENTUM, kNvls>, c, blocks, threads, st, c, a);
    case FMPI_RULE_NESTEROV:
      return launch_ranked(fused_step_kernel<T, kMaster, FMPI_RULE_NESTEROV, kNvls>, c, blocks, threads, st, c, a);
    case FMPI_RULE_ADAM:
      return launch_ranked(fused_step_kernel<T, kMaster, FMPI_RULE_ADAM, kNvls>, c, blocks, threads, st, c, a);
    case FMPI_RULE_ADAMW:
      return launch_ranked(fused_step_kernel<T, kMaster, FMPI_RULE_ADAMW, kNvls>, c, blocks, threads, st, c, a);
    default:
      return set_error("rule %d has no fused kernel", a.h.rule);
  }
}

template <typename T, bool kNvls>
int launch_master(const fmpi_devcomm& c, const FusedArgs& a, bool master, int blocks, int threads,
                  cudaStream_t st) {
  return master ? launch_rule<T, true, kNvls>(c, a, blocks, threads, st)
                : launch_rule<T, false, kNvls>(c, a, blocks, threads, st);
}

}  // namespace

extern "C" {

int fmpi_fused_step(const fmpi_devcomm* c, uint64_t g_off, int gdt, uint64_t p_off, int pdt,
                    float* const* m, float* const* v, float* const* master, uint64_t lo,
                    uint64_t hi, const fmpi_hyper* h, const int64_t* const* step_dev,
                    const fmpi_launch* l, fmpi_stream stream) {
  if (!c || !h) return set_error("null argument");
  if (c->world < 1 || c->world > FMPI_MAX_RANKS) return set_error("world %d out of range", c->world);
  if (gdt != pdt) return set_error("fused step needs gradients and parameters of one dtype");
  if (gdt != FMPI_F32 && gdt != FMPI_BF16 && gdt != FMPI_F16)
    return set_error("fused step supports f32 / bf16 / f16, got dtype %d", gdt);
  const uint64_t es = dtype_size(gdt), N = 16 / es;
  if (hi < lo) return set_error("bad bucket range");
  if (hi == lo) return 0;
  if ((g_off % 16) || (p_off % 16) || (lo % N) || ((hi - lo) % N))
    return set_error("bucket [%llu, %llu) / offsets must be 16-byte aligned",
                     (unsigned long long)lo, (unsigned long long)hi);
  if (g_off + hi * es > c->data_bytes || p_off + hi * es > c->data_bytes)
    return set_error("bucket exceeds the symmetric window");
  const int n_ranks = c->emulate ? c->world : 1;
  FusedArgs a{};
  a.g_off = g_off; a.p_off = p_off; a.lo = lo; a.hi = hi; a.h = *h;
  const bool need_m = h->rule != FMPI_RULE_DESCENT;
  const bool need_v = h->rule == FMPI_RULE_ADAM || h->rule == FMPI_RULE_ADAMW;
  for (int i = 0; i < n_ranks; ++i) {
    a.m[i] = m ? m[i] : nullptr;
    a.v[i] = v ? v[i] : nullptr;
    a.master[i] = master ? master[i] : nullptr;
    a.step[i] = step_dev ? step_dev[i] : nullptr;
    if ((need_m && !a.m[i]) || (need_v && !a.v[i])) return set_error("rule %d needs moment buffers", h->rule);
  }
  const bool use_master = master && master[0];
  if (gdt == FMPI_F32 && use_master) return set_error("fp32 parameters take no master copy");
  const int algo = l ? l->algo : FMPI_ALGO_AUTO;
  const bool nvls = algo == FMPI_ALGO_NVLS && c->world > 1;
  if (nvls && (!c->mc_data || c->emulate)) return set_error("NVLS requested but no multicast mapping");

  const uint64_t slice_bytes = (hi - lo) * es / c->world;
  int threads = (l && l->threads > 0) ? (l->threads + 31) / 32 * 32 : 512;  // whole warps: the barrier needs >= world threads
  if (threads > 512) threads = 512;
  int blocks;
  if (l && l->blocks > 0) {
    blocks = l->blocks;
  } else {
    const uint64_t per_block = uint64_t(threads) * 16 * 2;
    const uint64_t want = (slice_bytes + per_block - 1) / per_block;
    // multi-GPU: besides the link traffic the owner streams m, v and p through HBM, so the
    // kernel wants more CTAs than a pure all-reduce (measured at 8 GPUs, ResNet-50-sized
    // bucket, NVLS: 16 CTAs 454 us, 48: 301 us, 96: 286 us, 148: 289 us)
    // (round 2 kernel: the in-switch flavour fits two CTAs per SM; 128 of them by default)
    // measured at 2 GPUs (102 MB, peer-to-peer flavour): 48 CTAs 440 us, 128 CTAs 211 us — the
    // owner needs the whole chip's load slots to keep W remote vectors per element in flight
    const uint64_t cap = c->world == 1 ? 148 * 4 : nvls ? 128 : 144;
    blocks = int(want < 1 ? 1 : want > cap ? cap : want);
  }
  if (c->world > 1 && blocks > FMPI_MAX_BLOCKS) blocks = FMPI_MAX_BLOCKS;
  if (c->emulate && blocks > 128 / c->world) blocks = 128 / c->world;
  cudaStream_t st = static_cast<cudaStream_t>(stream);

  if (c->world == 1) {
    // no peers: the local streaming kernel, sized to fill the chip (256-thread CTAs, several
    // resident per SM, a few waves)
    fmpi_devcomm c1 = *c;
    c1.emulate = 0;
    c1.rank = 0;
    const int lthreads = 256;
    const uint64_t nvec1 = (hi - lo) / N;
    uint64_t want1 = (nvec1 + uint64_t(lthreads) * 4 - 1) / (uint64_t(lthreads) * 4);
    const int lblocks = (l && l->blocks > 0) ? l->blocks : int(want1 < 1 ? 1 : want1 > 148 * 8 ? 148 * 8 : want1);
#define FMPI_W1(T, MASTER)                                                                      \
  switch (h->rule) {                                                                            \
    case FMPI_RULE_DESCENT: fused_step_local_kernel<T, MASTER, FMPI_RULE_DESCENT><<<lblocks, lthreads, 0, st>>>(c1, a); break;   \
    case FMPI_RULE_MOMENTUM: fused_step_local_kernel<T, MASTER, FMPI_RULE_MOMENTUM><<<lblocks, lthreads, 0, st>>>(c1, a); break; \
    case FMPI_RULE_NESTEROV: fused_step_local_kernel<T, MASTER, FMPI_RULE_NESTEROV><<<lblocks, lthreads, 0, st>>>(c1, a); break; \
    case FMPI_RULE_ADAM: fused_step_local_kernel<T, MASTER, FMPI_RULE_ADAM><<<lblocks, lthreads, 0, st>>>(c1, a); break;         \
    case FMPI_RULE_ADAMW: fused_step_local_kernel<T, MASTER, FMPI_RULE_ADAMW><<<lblocks, lthreads, 0, st>>>(c1, a); break;       \
    default: return set_error("rule %d has no fused kernel", h->rule);                          \
  }
    if (gdt == FMPI_F32) { FMPI_W1(float, false) }
    else if (gdt == FMPI_BF16) { if (use_master) { FMPI_W1(__nv_bfloat16, true) } else { FMPI_W1(__nv_bfloat16, false) } }
    else { if (use_master) { FMPI_W1(__half, true) } else { FMPI_W1(__half, false) } }
#undef FMPI_W1
    return check_cuda(cudaGetLastError(), "fused_step launch");
  }

  if (gdt == FMPI_F32)
    return nvls ? launch_rule<float, false, true>(*c, a, blocks, threads, st)
                : launch_rule<float, false, false>(*c, a, blocks, threads, st);
  if (gdt == FMPI_BF16)
    return nvls ? launch_master<__nv_bfloat16, true>(*c, a, use_master, blocks, threads, st)
                : launch_master<__nv_bfloat16, false>(*c, a, use_master, blocks, threads, st);
  return nvls ? launch_master<__half, true>(*c, a, use_master, blocks, threads, st)
              : launch_master<__half, false>(*c, a, use_master, blocks, threads, st);
}

int fmpi_fused_step_oneshot(const fmpi_devcomm* c, uint64_t g_off, uint64_t p_off, int dt,
                            float* const* m, float* const* v, float* const* master, uint64_t lo,
                            uint64_t hi, const fmpi_hyper* h, const int64_t* const* step_dev,
                            uint64_t stage_off, uint64_t stage_bytes, const fmpi_launch* l,
                            fmpi_stream stream) {
  if (!c || !h) return set_error("null argument");
  if (c->world < 2 || c->world > FMPI_MAX_RANKS) return set_error("one-shot fused step needs 2..%d ranks", FMPI_MAX_RANKS);
  if (dt != FMPI_F32 && dt != FMPI_BF16 && dt != FMPI_F16)
    return set_error("fused step supports f32 / bf16 / f16, got dtype %d", dt);
  const uint64_t es = dtype_size(dt), N = 16 / es;
  if (hi <= lo) return hi == lo ? 0 : set_error("bad bucket range");
  if ((g_off % 16) || (p_off % 16) || (stage_off % 16) || (lo % N) || ((hi - lo) % N))
    return set_error("bucket / offsets must be 16-byte aligned");
  const uint64_t nbytes = (hi - lo) * es;
  const uint64_t slot = (stage_bytes / (2ull * c->world)) & ~uint64_t(15);
  if (nbytes > slot) return set_error("one-shot fused step: bucket of %llu bytes exceeds the %llu-byte staging slots",
                                      (unsigned long long)nbytes, (unsigned long long)slot);
  if (g_off + hi * es > c->data_bytes || p_off + hi * es > c->data_bytes || stage_off + stage_bytes > c->data_bytes)
    return set_error("bucket or staging exceeds the symmetric window");
  const int n_ranks = c->emulate ? c->world : 1;
  FusedArgs a{};
  a.g_off = g_off; a.p_off = p_off; a.lo = lo; a.hi = hi; a.h = *h;
  const bool need_m = h->rule != FMPI_RULE_DESCENT;
  const bool need_v = h->rule == FMPI_RULE_ADAM || h->rule == FMPI_RULE_ADAMW;
  for (int i = 0; i < n_ranks; ++i) {
    a.m[i] = m ? m[i] : nullptr;
    a.v[i] = v ? v[i] : nullptr;
    a.master[i] = master ? master[i] : nullptr;
    a.step[i] = step_dev ? step_dev[i] : nullptr;
    if ((need_m && !a.m[i]) || (need_v && !a.v[i])) return set_error("rule %d needs moment buffers", h->rule);
  }
  const bool use_master = master && master[0];
  if (dt == FMPI_F32 && use_master) return set_error("fp32 parameters take no master copy");
  int threads = (l && l->threads > 0) ? (l->threads + 31) / 32 * 32 : 512;
  if (threads > 512) threads = 512;
  const uint64_t nvec = nbytes / 16;
  int blocks = (l && l->blocks > 0) ? l->blocks : int((nvec + 2ull * threads - 1) / (2ull * threads));
  if (blocks < 1) blocks = 1;
  if (blocks > 32 && !(l && l->blocks > 0)) blocks = 32;
  if (blocks > FMPI_MAX_BLOCKS) blocks = FMPI_MAX_BLOCKS;
  if (c->emulate && blocks > 128 / c->world) blocks = 128 / c->world;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define FMPI_OS(T, MASTER)                                                                                  \
  switch (h->rule) {                                                                                        \
    case FMPI_RULE_DESCENT: return launch_ranked(fused_step_oneshot_kernel<T, MASTER, FMPI_RULE_DESCENT>, *c, blocks, threads, st, *c, a, stage_off, slot);   \
    case FMPI_RULE_MOMENTUM: return launch_ranked(fused_step_oneshot_kernel<T, MASTER, FMPI_RULE_MOMENTUM>, *c, blocks, threads, st, *c, a, stage_off, slot); \
    case FMPI_RULE_NESTEROV: return launch_ranked(fused_step_oneshot_kernel<T, MASTER, FMPI_RULE_NESTEROV>, *c, blocks, threads, st, *c, a, stage_off, slot); \
    case FMPI_RULE_ADAM: return launch_ranked(fused_step_oneshot_kernel<T, MASTER, FMPI_RULE_ADAM>, *c, blocks, threads, st, *c, a, stage_off, slot);         \
    case FMPI_RULE_ADAMW: return launch_ranked(fused_step_oneshot_kernel<T, MASTER, FMPI_RULE_ADAMW>, *c, blocks, threads, st, *c, a, stage_off, slot);       \
    default: return set_error("rule %d has no fused kernel", h->rule);                                      \
  }
  if (dt == FMPI_F32) { FMPI_OS(float, false) }
  else if (dt == FMPI_BF16) { if (use_master) { FMPI_OS(__nv_bfloat16, true) } else { FMPI_OS(__nv_bfloat16, false) } }
  else { if (use_master) { FMPI_OS(__half, true) } else { FMPI_OS(__half, false) } }
#undef FMPI_OS
}

int fmpi_step_advance(int64_t* step_dev, fmpi_stream st) {
  if (!step_dev) return set_error("null step counter");
  step_advance_kernel<<<1, 1, 0, static_cast<cudaStream_t>(st)>>>(step_dev);
  return check_cuda(cudaGetLastError(), "step_advance");
}

}  // extern "C"
