// Shared device-side building blocks for the cross-GPU kernels (sm_100a).
//
//  * signal-pad layout + the cross-GPU block barrier (replaces MPI_Barrier,
//    /root/reference/src/common.jl:91, and the implicit synchronisation inside every MPI
//    collective of /root/reference/src/mpi_extensions.jl)
//  * 16-byte vector load/store helpers with the cache hints that fit streaming peer traffic
//  * dtype/op-generic element arithmetic (the reference wraps a Julia function in MPI.Op,
//    src/mpi_extensions.jl:52-54; here sum/prod/min/max over nine dtypes, fp32 accumulation
//    for f16/bf16)
//
// Rank identity: real runs take it from fmpi_devcomm::rank; single-GPU emulation (tests on
// one B200, cooperative launch) takes it from blockIdx.y so the same device code — flags,
// barriers, peer pointers — is exercised without a second GPU.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/fmpi.h"

namespace fmpi {

// ---------------------------------------------------------------------------- signal pad
// Per channel (uint32 words):
//   flags[FMPI_MAX_BLOCKS][FMPI_MAX_RANKS]  written by peers: flags[b][src] = src's epoch
//   epoch[FMPI_MAX_BLOCKS]                  local: barriers block b has completed
//   ctl[FMPI_MAX_BLOCKS]                    local control words:
//     ctl[kCtlSeq]     push-mode calls completed on this channel (ONE word for the whole grid:
//                      every block of a launch reads it on entry, the LAST block to finish
//                      advances it, so no block can see the next call's parity)
//     ctl[kCtlTicket]  blocks of the current launch that have finished (reset by the last one)
//     ctl[kCtlPipe..]  state of the chunk-pipelined kernels (collectives.cu)
//   pipe flags [3][kPipeSlots][FMPI_MAX_RANKS]  written by peers (pipelined kernels)
constexpr int kFlagWords = FMPI_MAX_BLOCKS * FMPI_MAX_RANKS;
constexpr int kPipeSlots = 4;
constexpr int kPipeFlagWords = 3 * kPipeSlots * FMPI_MAX_RANKS;
constexpr int kChannelWords = kFlagWords + 2 * FMPI_MAX_BLOCKS + kPipeFlagWords;
constexpr size_t kSigPadBytes = size_t(FMPI_NUM_CHANNELS) * kChannelWords * sizeof(uint32_t);
constexpr int kCtlBase = kFlagWords + FMPI_MAX_BLOCKS;
constexpr int kCtlSeq = 0, kCtlTicket = 1, kCtlGrid = 2;   // kCtlGrid: local grid-wide barrier counter
constexpr int kCtlPipeBase = 8;                      // [kPipeSlots] slot uses in earlier calls
constexpr int kCtlPipeCnt = kCtlPipeBase + kPipeSlots;  // [3 roles][kPipeSlots] local counters
constexpr int kPipeFlagBase = kFlagWords + 2 * FMPI_MAX_BLOCKS;
static_assert(kCtlPipeCnt + 3 * kPipeSlots <= FMPI_MAX_BLOCKS, "control words overflow");

// Process-wide health page (pinned, mapped host memory; fmpi_devcomm::ctrl): the host sets
// ctrl[kCtrlAbort] when it learns that a peer died, a kernel that gave up waiting sets
// ctrl[kCtrlError]; both end every wait in flight and every later one.
constexpr int kCtrlAbort = 0, kCtrlError = 1;

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_acq_rel_sys() {
  asm volatile("fence.acq_rel.sys;" ::: "memory");
}

__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

struct Ctx {
  int rank;
  int world;
  uint32_t flag_off;    // word offset of this block's flags row inside a signal pad
  uint32_t* my_pad;     // this rank's signal pad, at this channel
  uint32_t epoch;
  bool dead;            // a wait of this thread was abandoned (peer lost / abort requested)
};

// Bounded wait for *f to reach `target` (monotonic counters, wrap-safe). Returns false when the
// wait was abandoned: the host asked for an abort, another wait already failed, or the budget
// fmpi_devcomm::timeout_ns ran out (then the error word is set for the host to find). The
// reference has no counterpart: a lost rank leaves its peers inside libmpi for ever.
__device__ __forceinline__ bool wait_flag(const uint32_t* f, uint32_t target, const fmpi_devcomm& c) {
  uint32_t spins = 0;
  uint64_t t0 = 0;
  while (int32_t(ld_acquire_sys(f) - target) < 0) {
    if ((++spins & 0x3ffu) == 0 && c.ctrl) {
      if (ld_relaxed_sys_u32(c.ctrl + kCtrlAbort) | ld_relaxed_sys_u32(c.ctrl + kCtrlError)) return false;
      if (c.timeout_ns) {
        const uint64_t now = globaltimer_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > c.timeout_ns) {
          atomicExch(c.ctrl + kCtrlError, 1u);
          return false;
        }
      }
    }
  }
  return true;
}

__device__ __forceinline__ int rank_of(const fmpi_devcomm& c) {
  return c.emulate ? int(blockIdx.y) : c.rank;
}
// index into the per-emulated-rank argument arrays (real mode uses entry 0)
__device__ __forceinline__ int arg_index(const fmpi_devcomm& c) {
  return c.emulate ? int(blockIdx.y) : 0;
}

__device__ __forceinline__ void ctx_init(Ctx& x, const fmpi_devcomm& c) {
  x.rank = rank_of(c);
  x.world = c.world;
  x.flag_off = uint32_t(c.channel) * kChannelWords + blockIdx.x * FMPI_MAX_RANKS;
  x.my_pad = c.sig[x.rank] + size_t(c.channel) * kChannelWords;
  x.epoch = x.my_pad[kFlagWords + blockIdx.x];  // only this block ever writes it (at kernel end)
  x.dead = false;
}

// Cross-GPU barrier between the same-index blocks of all ranks. Everything this block wrote
// before the barrier (to any GPU) is visible to every peer block after it.
__device__ __forceinline__ void block_barrier(Ctx& x, const fmpi_devcomm& c) {
  __syncthreads();
  const uint32_t target = ++x.epoch;
  if (threadIdx.x < x.world) {   // host entry points guarantee blockDim.x >= 32 >= world
    fence_acq_rel_sys();
    st_release_sys(c.sig[threadIdx.x] + x.flag_off + x.rank, target);
    if (!x.dead) x.dead = !wait_flag(c.sig[x.rank] + x.flag_off + threadIdx.x, target, c);
  }
  __syncthreads();
}

__device__ __forceinline__ void ctx_finish(Ctx& x) {
  if (threadIdx.x == 0) x.my_pad[kFlagWords + blockIdx.x] = x.epoch;
}

// Push-mode call parity: one word per channel, read by every block on entry.
__device__ __forceinline__ uint32_t seq_read(const Ctx& x) {
  return ld_relaxed_sys_u32(x.my_pad + kCtlBase + kCtlSeq);
}
// True in exactly one block of the launch: the last one to get here (so every block of this
// launch has already read whatever the caller is about to change). Resets the ticket.
__device__ __forceinline__ bool last_block_done(Ctx& x) {
  __shared__ uint32_t ticket_s;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    ticket_s = atomicAdd(x.my_pad + kCtlBase + kCtlTicket, 1u);
  }
  __syncthreads();
  const bool last = ticket_s == gridDim.x - 1;
  if (last && threadIdx.x == 0) x.my_pad[kCtlBase + kCtlTicket] = 0;
  return last;
}
__device__ __forceinline__ void seq_advance(Ctx& x, uint32_t cur) {
  if (last_block_done(x) && threadIdx.x == 0) x.my_pad[kCtlBase + kCtlSeq] = cur + 1;
}

// ---------------------------------------------------------------------------- 16-byte vectors
// Streaming loads: peer traffic and one-touch local data do not deserve L1 lines.
__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
// Data another GPU wrote into our memory during this kernel: bypass the (incoherent) L1.
__device__ __forceinline__ uint4 ld_volatile(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_stream(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

__device__ __forceinline__ uint8_t* byte_ptr(void* p, uint64_t off) {
  return static_cast<uint8_t*>(p) + off;
}
__device__ __forceinline__ const uint8_t* byte_ptr(const void* p, uint64_t off) {
  return static_cast<const uint8_t*>(p) + off;
}

// ---------------------------------------------------------------------------- element math
template <typename A>
__device__ __forceinline__ A op_apply(A a, A b, int op) {
  switch (op) {
    case FMPI_SUM: return a + b;
    case FMPI_PROD: return a * b;
    case FMPI_MIN: return b < a ? b : a;
    default: return a < b ? b : a;
  }
}

// VecOps<T>: view a uint4 as kN elements of T, accumulate in Acc.
template <typename T>
struct VecOps;

template <>
struct VecOps<float> {
  using Acc = float;
  static constexpr int kN = 4;
  __device__ static void unpack(const uint4& v, Acc* a) {
    a[0] = __uint_as_float(v.x); a[1] = __uint_as_float(v.y);
    a[2] = __uint_as_float(v.z); a[3] = __uint_as_float(v.w);
  }
  __device__ static uint4 pack(const Acc* a) {
    return make_uint4(__float_as_uint(a[0]), __float_as_uint(a[1]), __float_as_uint(a[2]),
                      __float_as_uint(a[3]));
  }
};

template <>
struct VecOps<double> {
  using Acc = double;
  static constexpr int kN = 2;
  __device__ static void unpack(const uint4& v, Acc* a) {
    a[0] = __longlong_as_double((long long)(uint64_t(v.y) << 32 | v.x));
    a[1] = __longlong_as_double((long long)(uint64_t(v.w) << 32 | v.z));
  }
  __device__ static uint4 pack(const Acc* a) {
    const uint64_t lo = (uint64_t)__double_as_longlong(a[0]);
    const uint64_t hi = (uint64_t)__double_as_longlong(a[1]);
    return make_uint4(uint32_t(lo), uint32_t(lo >> 32), uint32_t(hi), uint32_t(hi >> 32));
  }
};

template <>
struct VecOps<__nv_bfloat16> {
  using Acc = float;
  static constexpr int kN = 8;
  __device__ static void unpack(const uint4& v, Acc* a) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      a[2 * i] = __uint_as_float(w[i] << 16);
      a[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ static uint4 pack(const Acc* a) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(a[2 * i], a[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};

template <>
struct VecOps<__half> {
  using Acc = float;
  static constexpr int kN = 8;
  __device__ static void unpack(const uint4& v, Acc* a) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
      a[2 * i] = f.x;
      a[2 * i + 1] = f.y;
    }
  }
  __device__ static uint4 pack(const Acc* a) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __half2 h = __floats2half2_rn(a[2 * i], a[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};

template <typename I, int N>
struct IntVecOps {
  using Acc = I;
  static constexpr int kN = N;
  __device__ static void unpack(const uint4& v, Acc* a) {
    const I* p = reinterpret_cast<const I*>(&v);
#pragma unroll
    for (int i = 0; i < N; ++i) a[i] = p[i];
  }
  __device__ static uint4 pack(const Acc* a) {
    uint4 v;
    I* p = reinterpret_cast<I*>(&v);
#pragma unroll
    for (int i = 0; i < N; ++i) p[i] = a[i];
    return v;
  }
};
template <> struct VecOps<int32_t> : IntVecOps<int32_t, 4> {};
template <> struct VecOps<int64_t> : IntVecOps<int64_t, 2> {};
template <> struct VecOps<int16_t> : IntVecOps<int16_t, 8> {};
template <> struct VecOps<int8_t> : IntVecOps<int8_t, 16> {};
template <> struct VecOps<uint8_t> : IntVecOps<uint8_t, 16> {};

// ---- in-switch reduction / multicast store (NVLS)
template <typename T> struct Multimem;
template <> struct Multimem<float> {
  __device__ static uint4 ld_reduce(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
    return v;
  }
};
template <> struct Multimem<__nv_bfloat16> {
  __device__ static uint4 ld_reduce(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
    return v;
  }
};
template <> struct Multimem<__half> {
  __device__ static uint4 ld_reduce(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
    return v;
  }
};
__device__ __forceinline__ void multimem_st(void* mc, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

__host__ __device__ inline size_t dtype_size(int dt) {
  switch (dt) {
    case FMPI_F32: case FMPI_I32: return 4;
    case FMPI_F64: case FMPI_I64: return 8;
    case FMPI_F16: case FMPI_BF16: case FMPI_I16: return 2;
    case FMPI_U8: case FMPI_I8: return 1;
    default: return 0;
  }
}

}  // namespace fmpi

// Host-side helpers shared by the .cu translation units -------------------------------------
namespace fmpi_host {
int set_error(const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);

// Launch `kernel` either as a plain grid (real multi-GPU run) or cooperatively with
// gridDim.y = world (single-GPU emulation; co-residency is required because the emulated
// ranks wait on one another).
template <typename... Args>
int launch_ranked(void (*kernel)(Args...), const fmpi_devcomm& c, int blocks, int threads,
                  cudaStream_t st, Args... args) {
  if (blocks < 1) blocks = 1;
  if (blocks > FMPI_MAX_BLOCKS) blocks = FMPI_MAX_BLOCKS;
  if (c.emulate) {
    void* argv[] = {(void*)&args...};
    dim3 grid(blocks, c.world, 1), block(threads, 1, 1);
    return check_cuda(cudaLaunchCooperativeKernel((const void*)kernel, grid, block, argv, 0, st),
                      "cudaLaunchCooperativeKernel");
  }
  kernel<<<blocks, threads, 0, st>>>(args...);
  return check_cuda(cudaGetLastError(), "kernel launch");
}
}  // namespace fmpi_host
