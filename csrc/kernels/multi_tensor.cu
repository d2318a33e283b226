// Pointer-table ("multi-tensor") kernels: one launch over many scattered leaves.
//
//  * fmpi_multi_tensor_step — the optimiser rules of fused_step.cu with no communication, for
//    the functional `update(state, params, grads)` API on arbitrary trees (what Optimisers.jl
//    does leaf by leaf with 2-4 broadcast kernels each).
//  * fmpi_pack / fmpi_unpack — gather leaves into / scatter them out of one flat buffer so a
//    whole gradient or parameter tree travels in ONE collective instead of one MPI message per
//    leaf (/root/reference/src/optimizer.jl:49-58, src/synchronize.jl:12). This is the device
//    replacement for the reference's fmap(cpu, ·)/fmap(gpu, ·) staging (optimizer.jl:47,62).
//
// Work distribution: tensors are cut into fixed chunks; chunk k of the concatenated chunk list
// belongs to block k mod gridDim.x. Every block walks the (small) size table itself, so the
// host needs no per-launch prefix sums and the tables can stay resident on the device.
#include "common.cuh"

namespace fmpi {

constexpr uint64_t kChunkElems = 8192;
constexpr uint64_t kChunkBytes = 32768;

template <typename T> __device__ __forceinline__ float ldf(const T* p) { return float(*p); }
template <> __device__ __forceinline__ float ldf(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <> __device__ __forceinline__ float ldf(const __half* p) { return __half2float(*p); }
template <typename T> __device__ __forceinline__ void stf(T* p, float v) { *p = T(v); }
template <> __device__ __forceinline__ void stf(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ void stf(__half* p, float v) { *p = __float2half_rn(v); }

struct StepMath {
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

template <typename T>
__global__ void __launch_bounds__(512)
multi_tensor_step_kernel(int n, T* const* __restrict__ ps, const T* const* __restrict__ gs,
                         float* const* __restrict__ ms, float* const* __restrict__ vs,
                         const uint64_t* __restrict__ numel, StepMath sm) {
  const bool has_m = sm.rule != FMPI_RULE_DESCENT;
  const bool has_v = sm.rule == FMPI_RULE_ADAM || sm.rule == FMPI_RULE_ADAMW;
  uint64_t chunk0 = 0;  // global index of this tensor's first chunk
  for (int t = 0; t < n; ++t) {
    const uint64_t ne = numel[t];
    const uint64_t nch = (ne + kChunkElems - 1) / kChunkElems;
    const uint64_t G = gridDim.x;
    const uint64_t first = (blockIdx.x + G - chunk0 % G) % G;
    T* p = ps[t];
    const T* g = gs[t];
    float* m = has_m ? ms[t] : nullptr;
    float* v = has_v ? vs[t] : nullptr;
    // fp32 leaves whose four arrays are 16-byte aligned go 4 elements at a time with two
    // vectors in flight per thread; everything else takes the scalar loop below
    const bool vec4 = sizeof(T) == 4 &&
        ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) |
          (has_m ? reinterpret_cast<uintptr_t>(m) : 0) | (has_v ? reinterpret_cast<uintptr_t>(v) : 0)) & 15) == 0;
    for (uint64_t c = first; c < nch; c += G) {
      const uint64_t lo = c * kChunkElems;
      uint64_t hi = min(lo + kChunkElems, ne);
      if (vec4) {
        const uint64_t vhi = lo + (hi - lo) / 4 * 4;
        for (uint64_t i = lo + uint64_t(threadIdx.x) * 4; i < vhi; i += uint64_t(blockDim.x) * 8) {
          float4 pv[2], gv[2], mv[2], vv[2];
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const uint64_t j = i + uint64_t(u) * blockDim.x * 4;
            if (j < vhi) {
              pv[u] = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p) + j);
              gv[u] = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(g) + j);
              if (has_m) mv[u] = *reinterpret_cast<const float4*>(m + j);
              if (has_v) vv[u] = *reinterpret_cast<const float4*>(v + j);
            }
          }
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const uint64_t j = i + uint64_t(u) * blockDim.x * 4;
            if (j < vhi) {
              float* pe = reinterpret_cast<float*>(&pv[u]);
              float* ge = reinterpret_cast<float*>(&gv[u]);
              float* me = reinterpret_cast<float*>(&mv[u]);
              float* ve = reinterpret_cast<float*>(&vv[u]);
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                float mm = has_m ? me[k] : 0.f, v2 = has_v ? ve[k] : 0.f;
                pe[k] = sm.step(pe[k], ge[k], mm, v2);
                me[k] = mm;
                ve[k] = v2;
              }
              *reinterpret_cast<float4*>(reinterpret_cast<float*>(p) + j) = pv[u];
              if (has_m) *reinterpret_cast<float4*>(m + j) = mv[u];
              if (has_v) *reinterpret_cast<float4*>(v + j) = vv[u];
            }
          }
        }
        if (vhi == hi) continue;
        // fall through with the < 4 element remainder
        for (uint64_t i = vhi + threadIdx.x; i < hi; i += blockDim.x) {
          float mm = has_m ? m[i] : 0.f, vv2 = has_v ? v[i] : 0.f;
          const float pn = sm.step(ldf(p + i), ldf(g + i), mm, vv2);
          stf(p + i, pn);
          if (has_m) m[i] = mm;
          if (has_v) v[i] = vv2;
        }
        continue;
      }
      for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        float mm = has_m ? m[i] : 0.f, vv = has_v ? v[i] : 0.f;
        const float pn = sm.step(ldf(p + i), ldf(g + i), mm, vv);
        stf(p + i, pn);
        if (has_m) m[i] = mm;
        if (has_v) v[i] = vv;
      }
    }
    chunk0 += nch;
  }
}

// dir = 0: leaf -> flat (pack); dir = 1: flat -> leaf (unpack)
__global__ void __launch_bounds__(512)
pack_kernel(int n, uint8_t* const* __restrict__ leaves, const uint64_t* __restrict__ offs,
            const uint64_t* __restrict__ nbytes, uint8_t* __restrict__ flat, int dir) {
  uint64_t chunk0 = 0;
  for (int t = 0; t < n; ++t) {
    const uint64_t nb = nbytes[t];
    const uint64_t nch = (nb + kChunkBytes - 1) / kChunkBytes;
    const uint64_t G = gridDim.x;
    const uint64_t first = (blockIdx.x + G - chunk0 % G) % G;
    uint8_t* leaf = leaves[t];
    uint8_t* fl = flat + offs[t];
    const bool vec = ((reinterpret_cast<uintptr_t>(leaf) | reinterpret_cast<uintptr_t>(fl)) & 15) == 0;
    for (uint64_t c = first; c < nch; c += G) {
      const uint64_t lo = c * kChunkBytes, hi = min(lo + kChunkBytes, nb);
      uint8_t* d = dir ? leaf : fl;
      const uint8_t* s = dir ? fl : leaf;
      if (vec) {
        const uint64_t v1 = lo + (hi - lo) / 16 * 16;
        for (uint64_t i = lo + uint64_t(threadIdx.x) * 16; i < v1; i += uint64_t(blockDim.x) * 16)
          *reinterpret_cast<uint4*>(d + i) = ld_stream(s + i);
        for (uint64_t i = v1 + threadIdx.x; i < hi; i += blockDim.x) d[i] = s[i];
      } else {
        for (uint64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) d[i] = s[i];
      }
    }
    chunk0 += nch;
  }
}

}  // namespace fmpi

using namespace fmpi;
using fmpi_host::check_cuda;
using fmpi_host::set_error;

extern "C" {

int fmpi_multi_tensor_step(int n, void* const* p, const void* const* g, float* const* m,
                           float* const* v, const uint64_t* numel, int pdt, int gdt,
                           const fmpi_hyper* h, int64_t step, fmpi_stream st) {
  if (n <= 0) return 0;
  if (!p || !g || !numel || !h) return set_error("null argument");
  if (pdt != gdt) return set_error("multi_tensor_step needs parameters and gradients of one dtype");
  if (h->rule < FMPI_RULE_DESCENT || h->rule > FMPI_RULE_ADAMW) return set_error("rule %d has no kernel", h->rule);
  if (h->rule != FMPI_RULE_DESCENT && !m) return set_error("rule %d needs a moment table", h->rule);
  if ((h->rule == FMPI_RULE_ADAM || h->rule == FMPI_RULE_ADAMW) && !v) return set_error("Adam needs a second-moment table");
  StepMath sm;
  sm.rule = h->rule; sm.eta = h->eta; sm.b1 = h->beta1; sm.b2 = h->beta2; sm.eps = h->eps;
  sm.lambda = h->lambda; sm.rho = h->rho; sm.gscale = h->grad_scale;
  sm.bc1 = 1.f - powf(h->beta1, float(step + 1));
  sm.bc2 = 1.f - powf(h->beta2, float(step + 1));
  const int blocks = 148 * 2, threads = 512;
  cudaStream_t s = static_cast<cudaStream_t>(st);
  switch (pdt) {
    case FMPI_F32:
      multi_tensor_step_kernel<float><<<blocks, threads, 0, s>>>(n, reinterpret_cast<float* const*>(p), reinterpret_cast<const float* const*>(g), m, v, numel, sm);
      break;
    case FMPI_BF16:
      multi_tensor_step_kernel<__nv_bfloat16><<<blocks, threads, 0, s>>>(n, reinterpret_cast<__nv_bfloat16* const*>(p), reinterpret_cast<const __nv_bfloat16* const*>(g), m, v, numel, sm);
      break;
    case FMPI_F16:
      multi_tensor_step_kernel<__half><<<blocks, threads, 0, s>>>(n, reinterpret_cast<__half* const*>(p), reinterpret_cast<const __half* const*>(g), m, v, numel, sm);
      break;
    default: return set_error("multi_tensor_step: unsupported dtype %d", pdt);
  }
  return check_cuda(cudaGetLastError(), "multi_tensor_step launch");
}

int fmpi_pack(int n, const void* const* src, const uint64_t* offsets_bytes, const uint64_t* nbytes,
              void* flat, fmpi_stream st) {
  if (n <= 0) return 0;
  pack_kernel<<<148 * 2, 512, 0, static_cast<cudaStream_t>(st)>>>(
      n, reinterpret_cast<uint8_t* const*>(const_cast<void* const*>(src)), offsets_bytes, nbytes,
      static_cast<uint8_t*>(flat), 0);
  return check_cuda(cudaGetLastError(), "pack launch");
}

int fmpi_unpack(int n, void* const* dst, const uint64_t* offsets_bytes, const uint64_t* nbytes,
                const void* flat, fmpi_stream st) {
  if (n <= 0) return 0;
  pack_kernel<<<148 * 2, 512, 0, static_cast<cudaStream_t>(st)>>>(
      n, reinterpret_cast<uint8_t* const*>(dst), offsets_bytes, nbytes,
      static_cast<uint8_t*>(const_cast<void*>(flat)), 1);
  return check_cuda(cudaGetLastError(), "unpack launch");
}

}  // extern "C"
