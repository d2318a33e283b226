// Channels-last batch normalisation (training) fused with the residual add and the ReLU that
// follow it in ResNet bottlenecks — forward and backward, bf16 / f16 / f32 I/O, fp32 math.
//
// Why this is in a data-parallel library: measured with `ncu --metrics gpu__time_duration`
// (profiles/README.md §7), PyTorch's native channels_last batch-norm kernels are 60.7 % of the
// GPU time of a ResNet-50 training step on B200 and the separate ReLU / add kernels another
// 15 %, while the convolutions are 19 %. The headline benchmark (BASELINE.json config 2) is
// therefore bound by memory-bound normalisation code that runs far from the HBM roofline, and
// that is exactly the kind of op to hand-write (the reference has no model code at all; its
// users get whatever Lux/Flux + cuDNN give them).
//
// Layout: x is [M, C] row-major (M = N*H*W rows, C channels; that is NHWC memory). A CTA is a
// (tx channel-vectors) x (ty row-lanes) tile of 256 threads; a thread owns one 16-byte channel
// vector (8 bf16 or 4 f32 channels), keeps that vector's per-channel constants in registers
// and walks down the rows with four 16-byte loads in flight. Column reductions (statistics,
// d_gamma / d_beta) are per-thread fp32 partials -> shared-memory tree over the row-lanes ->
// one partial per CTA in scratch -> a tiny finalize kernel that folds them in double.
//
//   forward : stats -> finalize (mean, invstd, scale, shift, running stats) -> apply
//             y = relu?(x * scale + shift (+ residual))
//   backward: reduce (sum dz, sum dz*xhat; dz = dy * mask when ReLU is fused) -> finalize
//             (d_beta, d_gamma, coefficients) -> apply dx = a * (dz - b1 - xhat * b2), d_res = dz
//
// Statistics are accumulated around a per-channel pivot (the channel's value in row 0):
// sum(x - pivot) and sum((x - pivot)^2) in fp32 per thread, folded in double. With the plain
// E[x^2] - mean^2 form a channel whose |mean| is far above its standard deviation (post-ReLU
// activations, an un-normalised first layer) loses the variance to cancellation.
//
// The ReLU mask of a layer WITHOUT a residual input is recomputed in the backward kernels from
// x (x * scale + shift > 0, the fp32 expression the forward pass rounded) instead of being read
// from the forward output y: two of the seven passes over the activations of such a layer go away.
#include "common.cuh"

namespace fmpi {

struct BNGeom {
  int tx, ty, gy;
  unsigned gx;
  uint64_t rows_per_cta;
};

constexpr int kBNThreads = 256;
constexpr int kBNMaxGx = 1024;

__host__ inline BNGeom bn_geom(uint64_t M, int VC) {
  BNGeom g;
  g.tx = 1;
  while (g.tx < 32 && g.tx * 2 <= VC) g.tx *= 2;
  g.ty = kBNThreads / g.tx;
  g.gy = (VC + g.tx - 1) / g.tx;
  uint64_t want = (148ull * 3 + g.gy - 1) / g.gy;   // ~3 CTAs per SM; fewer partials to fold
  const uint64_t max_gx = (M + g.ty - 1) / g.ty;
  if (want > max_gx) want = max_gx;
  if (want > kBNMaxGx) want = kBNMaxGx;
  if (want < 1) want = 1;
  uint64_t rows = (M + want - 1) / want;
  rows = (rows + g.ty - 1) / g.ty * g.ty;
  g.rows_per_cta = rows;
  g.gx = unsigned((M + rows - 1) / rows);
  return g;
}

struct BNIdx {
  int cx, ry, vc;
  bool act;
  uint64_t r0, r1;
};

template <int N>
__device__ __forceinline__ BNIdx bn_idx(uint64_t M, int C, int tx, uint64_t rows_per_cta) {
  BNIdx i;
  i.cx = threadIdx.x % tx;
  i.ry = threadIdx.x / tx;
  i.vc = blockIdx.y * tx + i.cx;
  i.act = i.vc * N < C;
  i.r0 = uint64_t(blockIdx.x) * rows_per_cta;
  i.r1 = min(i.r0 + rows_per_cta, M);
  return i;
}

template <int N>
__device__ __forceinline__ void load_consts(const float* __restrict__ p, int c0, float* out) {
#pragma unroll
  for (int k = 0; k < N; k += 4) *reinterpret_cast<float4*>(out + k) = *reinterpret_cast<const float4*>(p + c0 + k);
}

// Tree-reduce K floats per thread over the row-lanes (ry) of the CTA; result valid where ry == 0.
template <int K>
__device__ __forceinline__ void lanes_reduce(float* v, float* sm, int tx, int ty, int cx, int ry) {
  float* mine = sm + (size_t(ry) * tx + cx) * K;
#pragma unroll
  for (int k = 0; k < K; ++k) mine[k] = v[k];
  __syncthreads();
  for (int off = ty >> 1; off > 0; off >>= 1) {
    if (ry < off) {
      const float* other = sm + (size_t(ry + off) * tx + cx) * K;
#pragma unroll
      for (int k = 0; k < K; ++k) mine[k] += other[k];
    }
    __syncthreads();
  }
#pragma unroll
  for (int k = 0; k < K; ++k) v[k] = mine[k];
}

// ---------------------------------------------------------------------------- forward
template <typename T>
__global__ void __launch_bounds__(kBNThreads)
bn_stats_kernel(const T* __restrict__ x, uint64_t M, int C, int tx, uint64_t rows_per_cta, float* __restrict__ partial,
                float* __restrict__ pivot_out) {
  using V = VecOps<T>;
  constexpr int N = V::kN;
  extern __shared__ float sm[];
  const int ty = blockDim.x / tx;
  const BNIdx i = bn_idx<N>(M, C, tx, rows_per_cta);
  float acc[2 * N], piv[N];
#pragma unroll
  for (int k = 0; k < 2 * N; ++k) acc[k] = 0.f;
  if (i.act) {
    const uint8_t* base = reinterpret_cast<const uint8_t*>(x) + uint64_t(i.vc) * 16;
    const uint64_t rb = uint64_t(C) * sizeof(T);
    V::unpack(*reinterpret_cast<const uint4*>(base), piv);       // row 0: the same pivot in every CTA
    if (blockIdx.x == 0 && i.ry == 0) {
#pragma unroll
      for (int k = 0; k < N; ++k) pivot_out[i.vc * N + k] = piv[k];
    }
    for (uint64_t r = i.r0 + i.ry; r < i.r1; r += 4ull * ty) {
      uint4 raw[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (r + uint64_t(u) * ty < i.r1) raw[u] = ld_stream(base + (r + uint64_t(u) * ty) * rb);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (r + uint64_t(u) * ty < i.r1) {
          float e[N];
          V::unpack(raw[u], e);
#pragma unroll
          for (int k = 0; k < N; ++k) {
            const float d = e[k] - piv[k];
            acc[k] += d;
            acc[N + k] += d * d;
          }
        }
    }
  }
  lanes_reduce<2 * N>(acc, sm, tx, ty, i.cx, i.ry);
  if (i.ry == 0 && i.act) {
    float* p = partial + uint64_t(blockIdx.x) * 2 * C + i.vc * N;
#pragma unroll
    for (int k = 0; k < N; ++k) {
      p[k] = acc[k];
      p[C + k] = acc[N + k];
    }
  }
}

// Fold the per-CTA partials of 32 channels with a 32 x 32 thread block: lane x = channel,
// lane y strides over the row chunks; doubles in shared memory, tree over y. (A one-thread-
// per-channel loop over ~10^3 partials took > 100 us per layer and was the top kernel of the
// step.) Returns the two sums in (s, q), valid where threadIdx.y == 0.
constexpr int kFoldC = 32, kFoldG = 32;
__device__ __forceinline__ void fold_partials(const float* __restrict__ partial, unsigned gx, int C, int c,
                                              double& s, double& q) {
  __shared__ double sh[2][kFoldG][kFoldC + 1];
  s = 0.0;
  q = 0.0;
  if (c < C)
    for (unsigned g = threadIdx.y; g < gx; g += kFoldG) {
      s += double(partial[uint64_t(g) * 2 * C + c]);
      q += double(partial[uint64_t(g) * 2 * C + C + c]);
    }
  sh[0][threadIdx.y][threadIdx.x] = s;
  sh[1][threadIdx.y][threadIdx.x] = q;
  __syncthreads();
  for (int off = kFoldG / 2; off > 0; off >>= 1) {
    if (int(threadIdx.y) < off) {
      sh[0][threadIdx.y][threadIdx.x] += sh[0][threadIdx.y + off][threadIdx.x];
      sh[1][threadIdx.y][threadIdx.x] += sh[1][threadIdx.y + off][threadIdx.x];
    }
    __syncthreads();
  }
  s = sh[0][0][threadIdx.x];
  q = sh[1][0][threadIdx.x];
}

__global__ void __launch_bounds__(kFoldC * kFoldG)
bn_fwd_finalize_kernel(const float* __restrict__ partial, const float* __restrict__ pivot, unsigned gx, uint64_t M, int C,
                       const float* __restrict__ gamma, const float* __restrict__ beta,
                       float* running_mean, float* running_var, float momentum, float eps,
                       float* save_mean, float* save_invstd, float* scale, float* shift) {
  const int c = blockIdx.x * kFoldC + threadIdx.x;
  double s, q;
  fold_partials(partial, gx, C, c, s, q);
  if (c >= C || threadIdx.y != 0) return;
  const double dm = s / double(M);                   // mean of (x - pivot)
  const double mean = double(pivot[c]) + dm;
  double var = q / double(M) - dm * dm;              // shift-invariant: no cancellation against the mean
  if (var < 0.0) var = 0.0;
  const float invstd = float(1.0 / sqrt(var + double(eps)));
  save_mean[c] = float(mean);
  save_invstd[c] = invstd;
  const float sc = (gamma ? gamma[c] : 1.f) * invstd;
  scale[c] = sc;
  shift[c] = (beta ? beta[c] : 0.f) - float(mean) * sc;
  if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * float(mean);
  if (running_var) {
    const double unbiased = M > 1 ? var * double(M) / double(M - 1) : var;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * float(unbiased);
  }
}

template <typename T, bool kRelu, bool kRes>
__global__ void __launch_bounds__(kBNThreads)
bn_apply_kernel(const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y, uint64_t M, int C, int tx,
                uint64_t rows_per_cta, const float* __restrict__ scale, const float* __restrict__ shift) {
  using V = VecOps<T>;
  constexpr int N = V::kN;
  const int ty = blockDim.x / tx;
  const BNIdx i = bn_idx<N>(M, C, tx, rows_per_cta);
  if (!i.act) return;
  float sc[N], sh[N];
  load_consts<N>(scale, i.vc * N, sc);
  load_consts<N>(shift, i.vc * N, sh);
  const uint64_t rb = uint64_t(C) * sizeof(T);
  const uint64_t col = uint64_t(i.vc) * 16;
  const uint8_t* xb = reinterpret_cast<const uint8_t*>(x) + col;
  const uint8_t* rbp = kRes ? reinterpret_cast<const uint8_t*>(res) + col : nullptr;
  uint8_t* yb = reinterpret_cast<uint8_t*>(y) + col;
  for (uint64_t r = i.r0 + i.ry; r < i.r1; r += 4ull * ty) {
    uint4 xr[4], rr[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (r + uint64_t(u) * ty < i.r1) {
        xr[u] = ld_stream(xb + (r + uint64_t(u) * ty) * rb);
        if (kRes) rr[u] = ld_stream(rbp + (r + uint64_t(u) * ty) * rb);
      }
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (r + uint64_t(u) * ty < i.r1) {
        float e[N], z[N];
        V::unpack(xr[u], e);
        if (kRes) V::unpack(rr[u], z);
#pragma unroll
        for (int k = 0; k < N; ++k) {
          float o = e[k] * sc[k] + sh[k];
          if (kRes) o += z[k];
          if (kRelu) o = fmaxf(o, 0.f);
          e[k] = o;
        }
        *reinterpret_cast<uint4*>(yb + (r + uint64_t(u) * ty) * rb) = V::pack(e);
      }
  }
}

// ---------------------------------------------------------------------------- backward
// kMask: 0 no ReLU, 1 mask from the forward output y, 2 mask recomputed from x (no residual)
template <typename T, int kMask>
__global__ void __launch_bounds__(kBNThreads)
bn_bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ y, const T* __restrict__ x, uint64_t M, int C,
                     int tx, uint64_t rows_per_cta, const float* __restrict__ mean,
                     const float* __restrict__ invstd, const float* __restrict__ gamma,
                     const float* __restrict__ beta, float* __restrict__ partial) {
  constexpr bool kRelu = kMask == 1;
  using V = VecOps<T>;
  constexpr int N = V::kN;
  extern __shared__ float sm[];
  const int ty = blockDim.x / tx;
  const BNIdx i = bn_idx<N>(M, C, tx, rows_per_cta);
  float acc[2 * N];
#pragma unroll
  for (int k = 0; k < 2 * N; ++k) acc[k] = 0.f;
  if (i.act) {
    float mu[N], is[N], sc[N], sh[N];
    load_consts<N>(mean, i.vc * N, mu);
    load_consts<N>(invstd, i.vc * N, is);
    if (kMask == 2) {
#pragma unroll
      for (int k = 0; k < N; ++k) {
        sc[k] = (gamma ? gamma[i.vc * N + k] : 1.f) * is[k];
        sh[k] = (beta ? beta[i.vc * N + k] : 0.f) - mu[k] * sc[k];
      }
    }
    const uint64_t rb = uint64_t(C) * sizeof(T);
    const uint64_t col = uint64_t(i.vc) * 16;
    const uint8_t* db = reinterpret_cast<const uint8_t*>(dy) + col;
    const uint8_t* yb = kRelu ? reinterpret_cast<const uint8_t*>(y) + col : nullptr;
    const uint8_t* xb = reinterpret_cast<const uint8_t*>(x) + col;
    for (uint64_t r = i.r0 + i.ry; r < i.r1; r += 2ull * ty) {
      uint4 dr[2], yr[2], xr[2];
#pragma unroll
      for (int u = 0; u < 2; ++u)
        if (r + uint64_t(u) * ty < i.r1) {
          const uint64_t o = (r + uint64_t(u) * ty) * rb;
          dr[u] = ld_stream(db + o);
          xr[u] = ld_stream(xb + o);
          if (kRelu) yr[u] = ld_stream(yb + o);
        }
#pragma unroll
      for (int u = 0; u < 2; ++u)
        if (r + uint64_t(u) * ty < i.r1) {
          float d[N], xe[N], ye[N];
          V::unpack(dr[u], d);
          V::unpack(xr[u], xe);
          if (kRelu) V::unpack(yr[u], ye);
#pragma unroll
          for (int k = 0; k < N; ++k) {
            const bool off = kMask == 1 ? !(ye[k] > 0.f) : kMask == 2 ? !(xe[k] * sc[k] + sh[k] > 0.f) : false;
            const float dz = off ? 0.f : d[k];
            acc[k] += dz;
            acc[N + k] += dz * (xe[k] - mu[k]) * is[k];
          }
        }
    }
  }
  lanes_reduce<2 * N>(acc, sm, tx, ty, i.cx, i.ry);
  if (i.ry == 0 && i.act) {
    float* p = partial + uint64_t(blockIdx.x) * 2 * C + i.vc * N;
#pragma unroll
    for (int k = 0; k < N; ++k) {
      p[k] = acc[k];
      p[C + k] = acc[N + k];
    }
  }
}

__global__ void __launch_bounds__(kFoldC * kFoldG)
bn_bwd_finalize_kernel(const float* __restrict__ partial, unsigned gx, uint64_t M, int C,
                       const float* __restrict__ gamma, const float* __restrict__ invstd,
                       float* dgamma, float* dbeta, float* a, float* b1, float* b2) {
  const int c = blockIdx.x * kFoldC + threadIdx.x;
  double s1, s2;
  fold_partials(partial, gx, C, c, s1, s2);
  if (c >= C || threadIdx.y != 0) return;
  if (dbeta) dbeta[c] = float(s1);
  if (dgamma) dgamma[c] = float(s2);
  a[c] = (gamma ? gamma[c] : 1.f) * invstd[c];
  b1[c] = float(s1 / double(M));
  b2[c] = float(s2 / double(M));
}

template <typename T, int kMask, bool kRes>
__global__ void __launch_bounds__(kBNThreads)
bn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ y, const T* __restrict__ x, T* __restrict__ dx,
                    T* __restrict__ dres, uint64_t M, int C, int tx, uint64_t rows_per_cta,
                    const float* __restrict__ mean, const float* __restrict__ invstd, const float* __restrict__ a,
                    const float* __restrict__ b1, const float* __restrict__ b2, const float* __restrict__ beta) {
  constexpr bool kRelu = kMask == 1;
  using V = VecOps<T>;
  constexpr int N = V::kN;
  const int ty = blockDim.x / tx;
  const BNIdx i = bn_idx<N>(M, C, tx, rows_per_cta);
  if (!i.act) return;
  float mu[N], is[N], ca[N], c1[N], c2[N];
  load_consts<N>(mean, i.vc * N, mu);
  load_consts<N>(invstd, i.vc * N, is);
  load_consts<N>(a, i.vc * N, ca);
  load_consts<N>(b1, i.vc * N, c1);
  load_consts<N>(b2, i.vc * N, c2);
  float sh[N];
  if (kMask == 2) {     // a = gamma * invstd is the forward scale; shift = beta - mean * scale
#pragma unroll
    for (int k = 0; k < N; ++k) sh[k] = (beta ? beta[i.vc * N + k] : 0.f) - mu[k] * ca[k];
  }
  const uint64_t rb = uint64_t(C) * sizeof(T);
  const uint64_t col = uint64_t(i.vc) * 16;
  const uint8_t* db = reinterpret_cast<const uint8_t*>(dy) + col;
  const uint8_t* yb = kRelu ? reinterpret_cast<const uint8_t*>(y) + col : nullptr;
  const uint8_t* xb = reinterpret_cast<const uint8_t*>(x) + col;
  uint8_t* ob = reinterpret_cast<uint8_t*>(dx) + col;
  uint8_t* rbp = kRes ? reinterpret_cast<uint8_t*>(dres) + col : nullptr;
  for (uint64_t r = i.r0 + i.ry; r < i.r1; r += 2ull * ty) {
    uint4 dr[2], yr[2], xr[2];
#pragma unroll
    for (int u = 0; u < 2; ++u)
      if (r + uint64_t(u) * ty < i.r1) {
        const uint64_t o = (r + uint64_t(u) * ty) * rb;
        dr[u] = ld_stream(db + o);
        xr[u] = ld_stream(xb + o);
        if (kRelu) yr[u] = ld_stream(yb + o);
      }
#pragma unroll
    for (int u = 0; u < 2; ++u)
      if (r + uint64_t(u) * ty < i.r1) {
        const uint64_t o = (r + uint64_t(u) * ty) * rb;
        float d[N], xe[N], ye[N], g[N];
        V::unpack(dr[u], d);
        V::unpack(xr[u], xe);
        if (kRelu) V::unpack(yr[u], ye);
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const bool off = kMask == 1 ? !(ye[k] > 0.f) : kMask == 2 ? !(xe[k] * ca[k] + sh[k] > 0.f) : false;
          const float dz = off ? 0.f : d[k];
          d[k] = dz;
          const float xh = (xe[k] - mu[k]) * is[k];
          g[k] = ca[k] * (dz - c1[k] - xh * c2[k]);
        }
        *reinterpret_cast<uint4*>(ob + o) = V::pack(g);
        if (kRes) *reinterpret_cast<uint4*>(rbp + o) = V::pack(d);
      }
  }
}

}  // namespace fmpi

using namespace fmpi;
using fmpi_host::check_cuda;
using fmpi_host::set_error;

namespace {

// scratch (floats): partial[kBNMaxGx][2][C] | scale | shift | a | b1 | b2 | pivot
size_t bn_scratch_floats(int C) { return size_t(kBNMaxGx) * 2 * C + 6 * size_t(C); }

template <typename T>
int bn_forward_t(const void* x, const void* res, void* y, uint64_t M, int C, const float* gamma, const float* beta,
                 float* rmean, float* rvar, float momentum, float eps, int relu, float* save_mean,
                 float* save_invstd, float* scratch, cudaStream_t st) {
  constexpr int N = VecOps<T>::kN;
  const BNGeom g = bn_geom(M, C / N);
  float* partial = scratch;
  float* scale = scratch + size_t(kBNMaxGx) * 2 * C;
  float* shift = scale + C;
  const dim3 grid(g.gx, g.gy);
  const size_t smem = size_t(kBNThreads) * 2 * N * sizeof(float);
  float* pivot = scale + 5 * size_t(C);
  bn_stats_kernel<T><<<grid, kBNThreads, smem, st>>>(static_cast<const T*>(x), M, C, g.tx, g.rows_per_cta, partial, pivot);
  bn_fwd_finalize_kernel<<<(C + kFoldC - 1) / kFoldC, dim3(kFoldC, kFoldG), 0, st>>>(partial, pivot, g.gx, M, C, gamma, beta, rmean, rvar, momentum,
                                                          eps, save_mean, save_invstd, scale, shift);
  const T* xp = static_cast<const T*>(x);
  const T* rp = static_cast<const T*>(res);
  T* yp = static_cast<T*>(y);
  if (relu) {
    if (res) bn_apply_kernel<T, true, true><<<grid, kBNThreads, 0, st>>>(xp, rp, yp, M, C, g.tx, g.rows_per_cta, scale, shift);
    else bn_apply_kernel<T, true, false><<<grid, kBNThreads, 0, st>>>(xp, rp, yp, M, C, g.tx, g.rows_per_cta, scale, shift);
  } else {
    if (res) bn_apply_kernel<T, false, true><<<grid, kBNThreads, 0, st>>>(xp, rp, yp, M, C, g.tx, g.rows_per_cta, scale, shift);
    else bn_apply_kernel<T, false, false><<<grid, kBNThreads, 0, st>>>(xp, rp, yp, M, C, g.tx, g.rows_per_cta, scale, shift);
  }
  return check_cuda(cudaGetLastError(), "bn_forward launch");
}

template <typename T>
int bn_apply_t(const void* x, const void* res, void* y, uint64_t M, int C, const float* scale, const float* shift,
               int relu, cudaStream_t st) {
  constexpr int N = VecOps<T>::kN;
  const BNGeom g = bn_geom(M, C / N);
  const dim3 grid(g.gx, g.gy);
  const T* xp = static_cast<const T*>(x);
  const T* rp = static_cast<const T*>(res);
  T* yp = static_cast<T*>(y);
  if (relu) {
    if (res) bn_apply_kernel<T, true, true><<<grid, kBNThreads, 0, st>>>(xp, rp, yp, M, C, g.tx, g.rows_per_cta, scale, shift);
    else bn_apply_kernel<T, true, false><<<grid, kBNThreads, 0, st>>>(xp, rp, yp, M, C, g.tx, g.rows_per_cta, scale, shift);
  } else {
    if (res) bn_apply_kernel<T, false, true><<<grid, kBNThreads, 0, st>>>(xp, rp, yp, M, C, g.tx, g.rows_per_cta, scale, shift);
    else bn_apply_kernel<T, false, false><<<grid, kBNThreads, 0, st>>>(xp, rp, yp, M, C, g.tx, g.rows_per_cta, scale, shift);
  }
  return check_cuda(cudaGetLastError(), "bn_apply launch");
}

template <typename T>
int bn_backward_t(const void* dy, const void* y, const void* x, void* dx, void* dres, uint64_t M, int C,
                  const float* gamma, const float* beta, const float* mean, const float* invstd, float* dgamma,
                  float* dbeta, int relu, float* scratch, cudaStream_t st) {
  constexpr int N = VecOps<T>::kN;
  const BNGeom g = bn_geom(M, C / N);
  float* partial = scratch;
  float* a = scratch + size_t(kBNMaxGx) * 2 * C + 2 * size_t(C);
  float* b1 = a + C;
  float* b2 = b1 + C;
  const dim3 grid(g.gx, g.gy);
  const size_t smem = size_t(kBNThreads) * 2 * N * sizeof(float);
  const T* dp = static_cast<const T*>(dy);
  const T* yp = static_cast<const T*>(y);
  const T* xp = static_cast<const T*>(x);
  T* op = static_cast<T*>(dx);
  T* rp = static_cast<T*>(dres);
  // mask: 0 none; 1 from y (a residual was added before the ReLU, or the caller passed y);
  //       2 recomputed from x (ReLU straight after the affine transform, y == NULL)
  const int mask = !relu ? 0 : (yp ? 1 : 2);
#define FMPI_BN_REDUCE(MASK) \
  bn_bwd_reduce_kernel<T, MASK><<<grid, kBNThreads, smem, st>>>(dp, yp, xp, M, C, g.tx, g.rows_per_cta, mean, invstd, gamma, beta, partial)
  if (mask == 0) FMPI_BN_REDUCE(0); else if (mask == 1) FMPI_BN_REDUCE(1); else FMPI_BN_REDUCE(2);
#undef FMPI_BN_REDUCE
  bn_bwd_finalize_kernel<<<(C + kFoldC - 1) / kFoldC, dim3(kFoldC, kFoldG), 0, st>>>(partial, g.gx, M, C, gamma, invstd, dgamma, dbeta, a, b1, b2);
#define FMPI_BN_APPLY(MASK, RES) \
  bn_bwd_apply_kernel<T, MASK, RES><<<grid, kBNThreads, 0, st>>>(dp, yp, xp, op, rp, M, C, g.tx, g.rows_per_cta, mean, invstd, a, b1, b2, beta)
  if (mask == 0) { if (dres) FMPI_BN_APPLY(0, true); else FMPI_BN_APPLY(0, false); }
  else if (mask == 1) { if (dres) FMPI_BN_APPLY(1, true); else FMPI_BN_APPLY(1, false); }
  else { if (dres) FMPI_BN_APPLY(2, true); else FMPI_BN_APPLY(2, false); }
#undef FMPI_BN_APPLY
  return check_cuda(cudaGetLastError(), "bn_backward launch");
}

int bn_check(uint64_t M, int C, int dtype, const void* a, const void* b) {
  if (dtype != FMPI_F32 && dtype != FMPI_BF16 && dtype != FMPI_F16) return set_error("batch norm: f32 / bf16 / f16 only");
  const int N = dtype == FMPI_F32 ? 4 : 8;
  if (M == 0 || C <= 0 || C % N) return set_error("batch norm: C = %d must be a positive multiple of %d", C, N);
  if ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) return set_error("batch norm: 16-byte aligned tensors required");
  return 0;
}

}  // namespace

extern "C" {

size_t fmpi_bn_scratch_bytes(int C) { return bn_scratch_floats(C) * sizeof(float); }

int fmpi_bn_forward(const void* x, const void* res, void* y, uint64_t M, int C, int dtype, const float* gamma,
                    const float* beta, float* running_mean, float* running_var, float momentum, float eps, int relu,
                    float* save_mean, float* save_invstd, void* scratch, fmpi_stream st) {
  if (bn_check(M, C, dtype, x, y)) return 1;
  if (!save_mean || !save_invstd || !scratch) return set_error("batch norm: null output / scratch");
  cudaStream_t s = static_cast<cudaStream_t>(st);
  float* sc = static_cast<float*>(scratch);
  switch (dtype) {
    case FMPI_F32: return bn_forward_t<float>(x, res, y, M, C, gamma, beta, running_mean, running_var, momentum, eps, relu, save_mean, save_invstd, sc, s);
    case FMPI_BF16: return bn_forward_t<__nv_bfloat16>(x, res, y, M, C, gamma, beta, running_mean, running_var, momentum, eps, relu, save_mean, save_invstd, sc, s);
    default: return bn_forward_t<__half>(x, res, y, M, C, gamma, beta, running_mean, running_var, momentum, eps, relu, save_mean, save_invstd, sc, s);
  }
}

int fmpi_bn_apply(const void* x, const void* res, void* y, uint64_t M, int C, int dtype, const float* scale,
                  const float* shift, int relu, fmpi_stream st) {
  if (bn_check(M, C, dtype, x, y)) return 1;
  cudaStream_t s = static_cast<cudaStream_t>(st);
  switch (dtype) {
    case FMPI_F32: return bn_apply_t<float>(x, res, y, M, C, scale, shift, relu, s);
    case FMPI_BF16: return bn_apply_t<__nv_bfloat16>(x, res, y, M, C, scale, shift, relu, s);
    default: return bn_apply_t<__half>(x, res, y, M, C, scale, shift, relu, s);
  }
}

int fmpi_bn_backward(const void* dy, const void* y, const void* x, void* dx, void* dres, uint64_t M, int C, int dtype,
                     const float* gamma, const float* beta, const float* save_mean, const float* save_invstd,
                     float* dgamma, float* dbeta, int relu, void* scratch, fmpi_stream st) {
  if (bn_check(M, C, dtype, dy, dx)) return 1;
  if (relu && !y && dres)
    return set_error("batch norm backward: with a residual input the ReLU mask needs the forward output y");
  if (!scratch) return set_error("batch norm: null scratch");
  cudaStream_t s = static_cast<cudaStream_t>(st);
  float* sc = static_cast<float*>(scratch);
  switch (dtype) {
    case FMPI_F32: return bn_backward_t<float>(dy, y, x, dx, dres, M, C, gamma, beta, save_mean, save_invstd, dgamma, dbeta, relu, sc, s);
    case FMPI_BF16: return bn_backward_t<__nv_bfloat16>(dy, y, x, dx, dres, M, C, gamma, beta, save_mean, save_invstd, dgamma, dbeta, relu, sc, s);
    default: return bn_backward_t<__half>(dy, y, x, dx, dres, M, C, gamma, beta, save_mean, save_invstd, dgamma, dbeta, relu, sc, s);
  }
}

}  // extern "C"
