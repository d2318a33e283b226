// Channels-last max pooling, forward and backward (bf16 / f16 / f32).
//
// After the batch-norm kernels (batchnorm.cu) PyTorch's `max_pool_backward_nhwc` (1.7 ms) and
// `max_pool_forward_nhwc` (0.75 ms) were the next memory-bound items of the ResNet-50 step on
// B200, against ~0.3 ms of HBM time for the bytes involved (profiles/README.md §7). Here:
//   forward : one thread per output pixel x 16-byte channel vector; the K x K window is read
//             with vector loads (neighbouring windows overlap in L1/L2) and the position of the
//             maximum is kept as one byte per channel.
//   backward: GATHER, one thread per input pixel x channel vector: the at most
//             ceil(K/S)^2 windows that contain the pixel are checked against the saved byte
//             and their gradients added — no atomics, every dx vector is written exactly once.
// First maximum in scan order wins, like PyTorch; NaNs propagate (a NaN beats any number).
#include "common.cuh"

namespace fmpi {

struct PoolGeom {
  int N, H, W, C, K, S, P, OH, OW;
};

template <typename T>
__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, uint8_t* __restrict__ idx, PoolGeom g) {
  using V = VecOps<T>;
  constexpr int NV = V::kN;
  const int VC = g.C / NV;
  const uint64_t total = uint64_t(g.N) * g.OH * g.OW * VC;
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t t = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
    const int vc = int(t % VC);
    uint64_t pix = t / VC;
    const int ow = int(pix % g.OW);
    pix /= g.OW;
    const int oh = int(pix % g.OH);
    const int n = int(pix / g.OH);
    float best[NV];
    uint8_t where[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      best[k] = -INFINITY;
      where[k] = 0;
    }
    bool first = true;
    for (int kh = 0; kh < g.K; ++kh) {
      const int ih = oh * g.S - g.P + kh;
      if (ih < 0 || ih >= g.H) continue;
      for (int kw = 0; kw < g.K; ++kw) {
        const int iw = ow * g.S - g.P + kw;
        if (iw < 0 || iw >= g.W) continue;
        const uint64_t off = ((uint64_t(n) * g.H + ih) * g.W + iw) * g.C + uint64_t(vc) * NV;
        float e[NV];
        V::unpack(*reinterpret_cast<const uint4*>(x + off), e);
#pragma unroll
        for (int k = 0; k < NV; ++k)
          if (first || e[k] > best[k] || e[k] != e[k]) {
            best[k] = e[k];
            where[k] = uint8_t(kh * g.K + kw);
          }
        first = false;
      }
    }
    const uint64_t o = ((uint64_t(n) * g.OH + oh) * g.OW + ow) * g.C + uint64_t(vc) * NV;
    *reinterpret_cast<uint4*>(y + o) = V::pack(best);
#pragma unroll
    for (int k = 0; k < NV; ++k) idx[o + k] = where[k];
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ idx, T* __restrict__ dx, PoolGeom g) {
  using V = VecOps<T>;
  constexpr int NV = V::kN;
  const int VC = g.C / NV;
  const uint64_t total = uint64_t(g.N) * g.H * g.W * VC;
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  for (uint64_t t = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; t < total; t += stride) {
    const int vc = int(t % VC);
    uint64_t pix = t / VC;
    const int iw = int(pix % g.W);
    pix /= g.W;
    const int ih = int(pix % g.H);
    const int n = int(pix / g.H);
    float acc[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) acc[k] = 0.f;
    // windows (oh, ow) with oh*S - P <= ih <= oh*S - P + K - 1
    const int oh_lo = max(0, (ih + g.P - g.K + g.S) / g.S);   // ceil((ih + P - K + 1) / S) for the non-negative case
    const int oh_hi = min(g.OH - 1, (ih + g.P) / g.S);
    const int ow_lo = max(0, (iw + g.P - g.K + g.S) / g.S);
    const int ow_hi = min(g.OW - 1, (iw + g.P) / g.S);
    for (int oh = oh_lo; oh <= oh_hi; ++oh) {
      const int kh = ih + g.P - oh * g.S;
      if (kh < 0 || kh >= g.K) continue;
      for (int ow = ow_lo; ow <= ow_hi; ++ow) {
        const int kw = iw + g.P - ow * g.S;
        if (kw < 0 || kw >= g.K) continue;
        const uint8_t code = uint8_t(kh * g.K + kw);
        const uint64_t o = ((uint64_t(n) * g.OH + oh) * g.OW + ow) * g.C + uint64_t(vc) * NV;
        float d[NV];
        V::unpack(*reinterpret_cast<const uint4*>(dy + o), d);
        uint8_t w[NV];
        if (NV == 8) *reinterpret_cast<uint2*>(w) = *reinterpret_cast<const uint2*>(idx + o);
        else *reinterpret_cast<uint32_t*>(w) = *reinterpret_cast<const uint32_t*>(idx + o);
#pragma unroll
        for (int k = 0; k < NV; ++k)
          if (w[k] == code) acc[k] += d[k];
      }
    }
    const uint64_t off = ((uint64_t(n) * g.H + ih) * g.W + iw) * g.C + uint64_t(vc) * NV;
    *reinterpret_cast<uint4*>(dx + off) = V::pack(acc);
  }
}

}  // namespace fmpi

using namespace fmpi;
using fmpi_host::check_cuda;
using fmpi_host::set_error;

namespace {
int pool_check(const PoolGeom& g, int dtype) {
  if (dtype != FMPI_F32 && dtype != FMPI_BF16 && dtype != FMPI_F16) return set_error("max pool: f32 / bf16 / f16 only");
  const int NV = dtype == FMPI_F32 ? 4 : 8;
  if (g.N <= 0 || g.H <= 0 || g.W <= 0 || g.C <= 0 || g.C % NV) return set_error("max pool: C must be a multiple of %d", NV);
  if (g.K < 1 || g.K > 15 || g.S < 1 || g.P < 0 || g.P >= g.K) return set_error("max pool: unsupported window");
  if (g.OH <= 0 || g.OW <= 0) return set_error("max pool: empty output");
  return 0;
}
PoolGeom pool_geom(int N, int H, int W, int C, int K, int S, int P) {
  PoolGeom g{N, H, W, C, K, S, P, 0, 0};
  g.OH = (H + 2 * P - K) / S + 1;
  g.OW = (W + 2 * P - K) / S + 1;
  return g;
}
int pool_blocks(uint64_t total) {
  const uint64_t want = (total + 255) / 256;
  return int(want < 1 ? 1 : want > 148ull * 32 ? 148 * 32 : want);
}
}  // namespace

extern "C" {

int fmpi_maxpool_fwd(const void* x, void* y, uint8_t* idx, int N, int H, int W, int C, int K, int S, int P,
                     int dtype, fmpi_stream st) {
  const PoolGeom g = pool_geom(N, H, W, C, K, S, P);
  if (pool_check(g, dtype)) return 1;
  const int NV = dtype == FMPI_F32 ? 4 : 8;
  const int blocks = pool_blocks(uint64_t(N) * g.OH * g.OW * (C / NV));
  cudaStream_t s = static_cast<cudaStream_t>(st);
  switch (dtype) {
    case FMPI_F32: maxpool_fwd_kernel<float><<<blocks, 256, 0, s>>>(static_cast<const float*>(x), static_cast<float*>(y), idx, g); break;
    case FMPI_BF16: maxpool_fwd_kernel<__nv_bfloat16><<<blocks, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), idx, g); break;
    default: maxpool_fwd_kernel<__half><<<blocks, 256, 0, s>>>(static_cast<const __half*>(x), static_cast<__half*>(y), idx, g); break;
  }
  return check_cuda(cudaGetLastError(), "maxpool_fwd launch");
}

int fmpi_maxpool_bwd(const void* dy, const uint8_t* idx, void* dx, int N, int H, int W, int C, int K, int S, int P,
                     int dtype, fmpi_stream st) {
  const PoolGeom g = pool_geom(N, H, W, C, K, S, P);
  if (pool_check(g, dtype)) return 1;
  const int NV = dtype == FMPI_F32 ? 4 : 8;
  const int blocks = pool_blocks(uint64_t(N) * H * W * (C / NV));
  cudaStream_t s = static_cast<cudaStream_t>(st);
  switch (dtype) {
    case FMPI_F32: maxpool_bwd_kernel<float><<<blocks, 256, 0, s>>>(static_cast<const float*>(dy), idx, static_cast<float*>(dx), g); break;
    case FMPI_BF16: maxpool_bwd_kernel<__nv_bfloat16><<<blocks, 256, 0, s>>>(static_cast<const __nv_bfloat16*>(dy), idx, static_cast<__nv_bfloat16*>(dx), g); break;
    default: maxpool_bwd_kernel<__half><<<blocks, 256, 0, s>>>(static_cast<const __half*>(dy), idx, static_cast<__half*>(dx), g); break;
  }
  return check_cuda(cudaGetLastError(), "maxpool_bwd launch");
}

}  // extern "C"
