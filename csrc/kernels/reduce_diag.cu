// Diagnostics on flat buffers: sum of squares (global gradient norm for logging / clipping)
// and a 64-bit checksum (bitwise replica-consistency assertions, SURVEY §5.1 — the reference
// has no such check; its tests compare values rank by rank, test/test_synchronize.jl).
//
// Both are single-launch two-stage reductions: every block writes one partial into a scratch
// buffer, takes a ticket, and the last block to arrive folds the partials in index order, so
// the result is deterministic run to run. The scratch buffer (fmpi_reduce_scratch_bytes(),
// caller-allocated, zero-initialised once) holds the ticket word and the partials; the ticket
// is reset by the last block so the buffer can be reused by the next call on the same stream.
#include <mutex>

#include "common.cuh"

namespace fmpi {

constexpr int kMaxPartials = 1024;
constexpr size_t kScratchBytes = 16 + sizeof(double) * kMaxPartials;

template <typename T> __device__ __forceinline__ double to_double(T x) { return double(x); }
template <> __device__ __forceinline__ double to_double(__nv_bfloat16 x) { return double(__bfloat162float(x)); }
template <> __device__ __forceinline__ double to_double(__half x) { return double(__half2float(x)); }

__device__ __forceinline__ double block_sum(double v, double* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  if (warp == 0) {
    v = lane < (blockDim.x + 31) / 32 ? sh[lane] : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  }
  return v;  // valid in thread 0
}

__device__ __forceinline__ uint64_t block_sum_u64(uint64_t v, uint64_t* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) sh[warp] = v;
  __syncthreads();
  if (warp == 0) {
    v = lane < (blockDim.x + 31) / 32 ? sh[lane] : 0ull;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  }
  return v;
}

// Returns true in thread 0 of the last block to finish.
__device__ __forceinline__ bool take_ticket(unsigned int* ticket) {
  __shared__ bool last;
  if (threadIdx.x == 0) {
    __threadfence();
    last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  return last;
}

template <typename T>
__global__ void __launch_bounds__(512)
sumsq_kernel(const T* __restrict__ x, uint64_t n, double* out, unsigned int* ticket, double* partials) {
  __shared__ double sh[32];
  using V = VecOps<T>;
  constexpr int N = V::kN;
  double acc = 0.0;
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t gtid = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  // scalar head up to the first 16-byte boundary, 16-byte vectors (4 in flight per thread:
  // this is a pure HBM read stream), scalar tail
  const uint64_t mis = reinterpret_cast<uintptr_t>(x) % 16;
  uint64_t head = mis ? (16 - mis) / sizeof(T) : 0;
  if (head > n) head = n;
  const uint64_t nvec = (n - head) / N;
  const uint8_t* xv = reinterpret_cast<const uint8_t*>(x + head);
  for (uint64_t i = gtid; i < nvec; i += 4 * stride) {
    uint4 raw[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * stride < nvec) raw[u] = ld_stream(xv + (i + u * stride) * 16);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * stride < nvec) {
        typename V::Acc e[N];
        V::unpack(raw[u], e);
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const double d = double(e[k]);
          acc += d * d;
        }
      }
  }
  if (gtid == 0) {
    for (uint64_t i = 0; i < head; ++i) { const double d = to_double(x[i]); acc += d * d; }
    for (uint64_t i = head + nvec * N; i < n; ++i) { const double d = to_double(x[i]); acc += d * d; }
  }
  acc = block_sum(acc, sh);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
  if (take_ticket(ticket)) {
    __threadfence();
    double tot = 0.0;
    for (int b = threadIdx.x; b < gridDim.x; b += blockDim.x) tot += partials[b];
    __syncthreads();
    tot = block_sum(tot, sh);
    if (threadIdx.x == 0) {
      *out = tot;
      *ticket = 0;
    }
  }
}

__global__ void __launch_bounds__(512)
checksum_kernel(const uint8_t* __restrict__ x, uint64_t nbytes, uint64_t* out, unsigned int* ticket,
                uint64_t* partials) {
  __shared__ uint64_t sh[32];
  uint64_t acc = 0;
  const uint64_t nwords = nbytes / 4;
  const uint32_t* w = reinterpret_cast<const uint32_t*>(x);
  const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
  const uint64_t gtid = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  // word i contributes (w[i] + C) * (2i + 1): position sensitive, order independent, so the
  // bulk can be read as 16-byte vectors (4 in flight) around a scalar head / tail
  const uint64_t mis = (reinterpret_cast<uintptr_t>(x) % 16) / 4;
  uint64_t head = mis ? 4 - mis : 0;
  if (head > nwords) head = nwords;
  const uint64_t nvec = (nwords - head) / 4;
  const uint8_t* xv = reinterpret_cast<const uint8_t*>(w + head);
  for (uint64_t i = gtid; i < nvec; i += 4 * stride) {
    uint4 raw[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * stride < nvec) raw[u] = ld_stream(xv + (i + u * stride) * 16);
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (i + u * stride < nvec) {
        const uint64_t w0 = head + (i + u * stride) * 4;
        const uint32_t ws[4] = {raw[u].x, raw[u].y, raw[u].z, raw[u].w};
#pragma unroll
        for (int k = 0; k < 4; ++k) acc += (uint64_t(ws[k]) + 0x9E3779B97F4A7C15ull) * (2 * (w0 + k) + 1);
      }
  }
  if (gtid == 0) {
    for (uint64_t i = 0; i < head; ++i) acc += (uint64_t(w[i]) + 0x9E3779B97F4A7C15ull) * (2 * i + 1);
    for (uint64_t i = head + nvec * 4; i < nwords; ++i) acc += (uint64_t(w[i]) + 0x9E3779B97F4A7C15ull) * (2 * i + 1);
    for (uint64_t k = nwords * 4; k < nbytes; ++k) acc += (uint64_t(x[k]) + 0x51ull) * (2 * k + 1);
  }
  acc = block_sum_u64(acc, sh);
  if (threadIdx.x == 0) partials[blockIdx.x] = acc;
  if (take_ticket(ticket)) {
    __threadfence();
    uint64_t tot = 0;
    for (int b = threadIdx.x; b < gridDim.x; b += blockDim.x) tot += partials[b];
    __syncthreads();
    tot = block_sum_u64(tot, sh);
    if (threadIdx.x == 0) {
      *out = tot;
      *ticket = 0;
    }
  }
}

}  // namespace fmpi

using namespace fmpi;
using fmpi_host::check_cuda;
using fmpi_host::set_error;

static int diag_blocks(uint64_t n) {
  const uint64_t want = (n + 512 * 8 - 1) / (512 * 8);
  return int(want < 1 ? 1 : want > 148 * 4 ? 148 * 4 : want);
}

extern "C" {

size_t fmpi_reduce_scratch_bytes(void) { return kScratchBytes; }

int fmpi_sumsq(const void* x, uint64_t count, int dtype, double* out, void* scratch, fmpi_stream st) {
  if (!out || !scratch) return set_error("null output / scratch");
  unsigned int* ticket = static_cast<unsigned int*>(scratch);
  double* partials = reinterpret_cast<double*>(static_cast<uint8_t*>(scratch) + 16);
  const int blocks = diag_blocks(count);
  cudaStream_t s = static_cast<cudaStream_t>(st);
  switch (dtype) {
    case FMPI_F32: sumsq_kernel<float><<<blocks, 512, 0, s>>>(static_cast<const float*>(x), count, out, ticket, partials); break;
    case FMPI_F64: sumsq_kernel<double><<<blocks, 512, 0, s>>>(static_cast<const double*>(x), count, out, ticket, partials); break;
    case FMPI_BF16: sumsq_kernel<__nv_bfloat16><<<blocks, 512, 0, s>>>(static_cast<const __nv_bfloat16*>(x), count, out, ticket, partials); break;
    case FMPI_F16: sumsq_kernel<__half><<<blocks, 512, 0, s>>>(static_cast<const __half*>(x), count, out, ticket, partials); break;
    default: return set_error("sumsq: unsupported dtype %d", dtype);
  }
  return check_cuda(cudaGetLastError(), "sumsq launch");
}

int fmpi_checksum(const void* x, uint64_t nbytes, uint64_t* out, void* scratch, fmpi_stream st) {
  if (!out || !scratch) return set_error("null output / scratch");
  if (reinterpret_cast<uintptr_t>(x) % 4) return set_error("checksum: buffer must be 4-byte aligned");
  unsigned int* ticket = static_cast<unsigned int*>(scratch);
  uint64_t* partials = reinterpret_cast<uint64_t*>(static_cast<uint8_t*>(scratch) + 16);
  checksum_kernel<<<diag_blocks(nbytes / 4), 512, 0, static_cast<cudaStream_t>(st)>>>(
      static_cast<const uint8_t*>(x), nbytes, out, ticket, partials);
  return check_cuda(cudaGetLastError(), "checksum launch");
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// fmpi_sumsq_noalloc — the same sum of squares for callers that must not allocate.
//
// Who that is: code running where a device allocation is either forbidden or would be a bug —
// a Julia/CUDA.jl host that owns the memory pool and does not want a foreign library calling
// the allocator behind its back, a logging hook that fires in the middle of a training step
// whose memory has been planned to the byte, a monitoring thread that may run while another
// thread is in an allocator-sensitive region. fmpi_sumsq asks such a caller to bring a scratch
// buffer; this variant brings its own: a small array that lives in the library's device image.
//
// What to know before using it:
//   * The scratch is ONE array per GPU (every GPU that loads this library holds these 2304
//     bytes until the process exits) and every call on that GPU uses it. Two calls whose
//     kernels can overlap on the device — different streams without an ordering between
//     them — race on it and produce garbage. Calls on one stream are fine.
//   * The array keeps whatever the previous call left behind. The only part that has to start
//     from zero is the arrival ticket; each call clears it on its stream before the kernel.
//   * At most FMPI_NOALLOC_MAX_COUNT elements per call; larger inputs are refused.
//   * The first call on a device looks up the array's device address (once, with that device
//     current). Make that first call outside of any stream capture.
//   * Nothing here calls cudaMalloc, cudaFree or any allocating library routine. The caller
//     still supplies the input and the 8-byte output.
namespace fmpi {
constexpr int kNoallocPartials = 256;
constexpr size_t kNoallocBytes = 2304;  // 16 (ticket) + 8 * 256 partials, rounded up to 256
static_assert(16 + sizeof(double) * kNoallocPartials <= kNoallocBytes, "scratch too small");
}  // namespace fmpi

// In host code this name is only a stand-in for the symbol, not a device address; the address
// is obtained with cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[fmpi::kNoallocBytes];

namespace {
constexpr int kMaxDevices = 64;
void* g_noalloc_ptr[kMaxDevices];
bool g_noalloc_ready[kMaxDevices];
std::mutex g_noalloc_mu;

// Once per device, with that device current.
int noalloc_scratch(void** out) {
  int dev = -1;
  if (check_cuda(cudaGetDevice(&dev), "cudaGetDevice")) return 1;
  if (dev < 0 || dev >= kMaxDevices) return set_error("device %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ready[dev]) {
    void* p = nullptr;
    if (check_cuda(cudaGetSymbolAddress(&p, g_noalloc_scratch), "cudaGetSymbolAddress(noalloc scratch)"))
      return 1;
    g_noalloc_ptr[dev] = p;
    g_noalloc_ready[dev] = true;
  }
  *out = g_noalloc_ptr[dev];
  return 0;
}
}  // namespace

extern "C" int fmpi_sumsq_noalloc(const void* x, uint64_t count, int dtype, double* out,
                                  fmpi_stream st) {
  if (!out) return set_error("null output");
  if (!x && count > 0) return set_error("null input with %llu elements", (unsigned long long)count);
  if (count > FMPI_NOALLOC_MAX_COUNT)
    return set_error("fmpi_sumsq_noalloc: %llu elements exceed the limit of %llu; split the input "
                     "or use fmpi_sumsq with a caller-provided scratch buffer",
                     (unsigned long long)count, (unsigned long long)FMPI_NOALLOC_MAX_COUNT);
  if (dtype != FMPI_F32 && dtype != FMPI_F64 && dtype != FMPI_BF16 && dtype != FMPI_F16)
    return set_error("sumsq: unsupported dtype %d", dtype);
  void* scratch = nullptr;
  if (noalloc_scratch(&scratch)) return 1;
  cudaStream_t s = static_cast<cudaStream_t>(st);
  // the ticket must be zero when the kernel starts, whatever an earlier call left there
  if (check_cuda(cudaMemsetAsync(scratch, 0, 16, s), "cudaMemsetAsync(noalloc ticket)")) return 1;
  unsigned int* ticket = static_cast<unsigned int*>(scratch);
  double* partials = reinterpret_cast<double*>(static_cast<uint8_t*>(scratch) + 16);
  uint64_t want = (count + 512 * 8 - 1) / (512 * 8);
  const int blocks = int(want < 1 ? 1 : want > fmpi::kNoallocPartials ? fmpi::kNoallocPartials : want);
  switch (dtype) {
    case FMPI_F32: sumsq_kernel<float><<<blocks, 512, 0, s>>>(static_cast<const float*>(x), count, out, ticket, partials); break;
    case FMPI_F64: sumsq_kernel<double><<<blocks, 512, 0, s>>>(static_cast<const double*>(x), count, out, ticket, partials); break;
    case FMPI_BF16: sumsq_kernel<__nv_bfloat16><<<blocks, 512, 0, s>>>(static_cast<const __nv_bfloat16*>(x), count, out, ticket, partials); break;
    default: sumsq_kernel<__half><<<blocks, 512, 0, s>>>(static_cast<const __half*>(x), count, out, ticket, partials); break;
  }
  return check_cuda(cudaGetLastError(), "sumsq_noalloc launch");
}
