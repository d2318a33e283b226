// POSIX shared-memory collectives for CPU tensors.
//
// This image (and the GPU box) has no MPI, so host tensors — the SPMD test-suite, the
// "Quick-Start MLP on CPU at 3 ranks" config, scalar/metadata exchange — are served by this
// small backend instead. It provides what the reference asks of MPI for `Array`s
// (/root/reference/src/mpi_extensions.jl:108-111 Allreduce!, :130-133 Bcast!, :152-155 Reduce!)
// plus a barrier (src/common.jl:91) and an all-gather for handle exchange.
//
// Design: one shm segment per job. Header (barrier words, attach count) followed by W staging
// slots and one result slot. A message is processed in slot-sized chunks:
//   allreduce: copy-in -> barrier -> rank r reduces slice r over slots 0..W-1 in rank order
//              into the result slot -> barrier -> copy-out.
// One owner per slice + fixed rank order => bitwise-identical results on every rank (the
// replica-consistency rule of SURVEY.md §5.1). Every wait is bounded by a timeout so a dead
// peer surfaces as an error rather than a hang.
#define FMPI_HOST_ONLY
#include "../include/fmpi.h"

#include <atomic>
#include <cerrno>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <future>
#include <string>
#include <vector>

#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

namespace {

thread_local std::string g_err;

int fail(const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return 1;
}

constexpr uint32_t kMagic = 0x464d5049u;  // "FMPI"

struct alignas(64) Header {
  std::atomic<uint32_t> magic;
  uint32_t world;
  uint64_t slot_bytes;
  alignas(64) std::atomic<uint32_t> attached;
  alignas(64) std::atomic<uint32_t> detached;
  alignas(64) std::atomic<uint32_t> bar_count;
  alignas(64) std::atomic<uint32_t> bar_gen;
  alignas(64) std::atomic<uint32_t> aborted;
};

double now_s() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return double(ts.tv_sec) + 1e-9 * double(ts.tv_nsec);
}

// Waiting policy: spin (a peer that is running arrives within microseconds), then yield the
// core (ranks may outnumber cores on CPU test boxes), then sleep in short naps so a job that
// waits for a long time does not burn a core per rank. A nap costs 50+ us of latency, so it is
// kept for waits that are already long.
struct Backoff {
  int n = 0;
  void pause() {
    ++n;
    if (n < 1024) {
#if defined(__x86_64__)
      __builtin_ia32_pause();
#endif
    } else if (n < 1024 + 8192) {
      sched_yield();
    } else {
      timespec ts{0, 50000};
      nanosleep(&ts, nullptr);
    }
  }
};

// bf16 / f16 <-> f32 (host arithmetic happens in float; one rounding at the end).
inline float bf16_to_f(uint16_t h) {
  uint32_t u = uint32_t(h) << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
inline uint16_t f_to_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return uint16_t((u >> 16) | 0x40);  // quiet NaN
  u += 0x7fffu + ((u >> 16) & 1u);                                        // RNE
  return uint16_t(u >> 16);
}
inline float f16_to_f(uint16_t h) {
  uint32_t sign = (uint32_t(h) & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1f;
  uint32_t man = h & 0x3ffu;
  uint32_t u;
  if (exp == 0) {
    if (man == 0) {
      u = sign;
    } else {  // subnormal
      int e = -1;
      do { ++e; man <<= 1; } while ((man & 0x400u) == 0);
      u = sign | uint32_t(127 - 15 - e) << 23 | (man & 0x3ffu) << 13;
    }
  } else if (exp == 31) {
    u = sign | 0x7f800000u | man << 13;
  } else {
    u = sign | (exp + 127 - 15) << 23 | man << 13;
  }
  float f;
  memcpy(&f, &u, 4);
  return f;
}
inline uint16_t f_to_f16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  uint32_t sign = (u >> 16) & 0x8000u;
  int32_t exp = int32_t((u >> 23) & 0xff) - 127 + 15;
  uint32_t man = u & 0x7fffffu;
  if (((u >> 23) & 0xff) == 0xff) return uint16_t(sign | 0x7c00u | (man ? 0x200u : 0));
  if (exp >= 31) return uint16_t(sign | 0x7c00u);
  if (exp <= 0) {
    if (exp < -10) return uint16_t(sign);
    man |= 0x800000u;
    uint32_t shift = uint32_t(14 - exp);
    uint32_t half = man >> shift;
    uint32_t rem = man & ((1u << shift) - 1);
    uint32_t mid = 1u << (shift - 1);
    if (rem > mid || (rem == mid && (half & 1))) ++half;
    return uint16_t(sign | half);
  }
  uint32_t half = uint32_t(exp) << 10 | man >> 13;
  uint32_t rem = man & 0x1fffu;
  if (rem > 0x1000u || (rem == 0x1000u && (half & 1))) ++half;
  return uint16_t(sign | half);
}

template <typename T>
inline T apply_op(T a, T b, int op) {
  switch (op) {
    case FMPI_SUM: return a + b;
    case FMPI_PROD: return a * b;
    case FMPI_MIN: return b < a ? b : a;
    default: return a < b ? b : a;
  }
}

// out[i] = reduce_{r=0..W-1} slot_r[i]   for i in [lo, hi)
template <typename T>
void reduce_native(const uint8_t* slots, size_t slot_stride, int world, uint8_t* out, size_t lo,
                   size_t hi, int op) {
  T* o = reinterpret_cast<T*>(out);
  const T* s0 = reinterpret_cast<const T*>(slots);
  for (size_t i = lo; i < hi; ++i) o[i] = s0[i];
  for (int r = 1; r < world; ++r) {
    const T* s = reinterpret_cast<const T*>(slots + size_t(r) * slot_stride);
    switch (op) {
      case FMPI_SUM: for (size_t i = lo; i < hi; ++i) o[i] = o[i] + s[i]; break;
      case FMPI_PROD: for (size_t i = lo; i < hi; ++i) o[i] = o[i] * s[i]; break;
      case FMPI_MIN: for (size_t i = lo; i < hi; ++i) o[i] = s[i] < o[i] ? s[i] : o[i]; break;
      default: for (size_t i = lo; i < hi; ++i) o[i] = o[i] < s[i] ? s[i] : o[i]; break;
    }
  }
}

template <float (*Load)(uint16_t), uint16_t (*Store)(float)>
void reduce_half(const uint8_t* slots, size_t slot_stride, int world, uint8_t* out, size_t lo,
                 size_t hi, int op) {
  uint16_t* o = reinterpret_cast<uint16_t*>(out);
  for (size_t i = lo; i < hi; ++i) {
    float acc = Load(reinterpret_cast<const uint16_t*>(slots)[i]);
    for (int r = 1; r < world; ++r)
      acc = apply_op<float>(
          acc, Load(reinterpret_cast<const uint16_t*>(slots + size_t(r) * slot_stride)[i]), op);
    o[i] = Store(acc);
  }
}

size_t dtype_size(int dt) {
  switch (dt) {
    case FMPI_F32: case FMPI_I32: return 4;
    case FMPI_F64: case FMPI_I64: return 8;
    case FMPI_F16: case FMPI_BF16: case FMPI_I16: return 2;
    case FMPI_U8: case FMPI_I8: return 1;
    default: return 0;
  }
}

void reduce_range(int dt, const uint8_t* slots, size_t stride, int world, uint8_t* out,
                  size_t lo, size_t hi, int op);

// The fold makes `world` passes over the output range: done in cache-sized blocks so that only
// the first pass of a block misses.
void reduce_dispatch(int dt, const uint8_t* slots, size_t stride, int world, uint8_t* out,
                     size_t lo, size_t hi, int op) {
  const size_t es = dtype_size(dt);
  const size_t blk = es ? (size_t(128) << 10) / es : hi - lo;
  for (size_t b = lo; b < hi; b += blk) reduce_range(dt, slots, stride, world, out, b, b + blk < hi ? b + blk : hi, op);
}

void reduce_range(int dt, const uint8_t* slots, size_t stride, int world, uint8_t* out,
                  size_t lo, size_t hi, int op) {
  switch (dt) {
    case FMPI_F32: reduce_native<float>(slots, stride, world, out, lo, hi, op); break;
    case FMPI_F64: reduce_native<double>(slots, stride, world, out, lo, hi, op); break;
    case FMPI_I32: reduce_native<int32_t>(slots, stride, world, out, lo, hi, op); break;
    case FMPI_I64: reduce_native<int64_t>(slots, stride, world, out, lo, hi, op); break;
    case FMPI_I16: reduce_native<int16_t>(slots, stride, world, out, lo, hi, op); break;
    case FMPI_U8: reduce_native<uint8_t>(slots, stride, world, out, lo, hi, op); break;
    case FMPI_I8: reduce_native<int8_t>(slots, stride, world, out, lo, hi, op); break;
    case FMPI_F16: reduce_half<f16_to_f, f_to_f16>(slots, stride, world, out, lo, hi, op); break;
    case FMPI_BF16: reduce_half<bf16_to_f, f_to_bf16>(slots, stride, world, out, lo, hi, op); break;
  }
}

// acc[i] = acc[i] op src[i] — one step of the rank-ordered fold, same arithmetic as reduce_native
template <typename T>
void accumulate(uint8_t* acc_, const uint8_t* src_, size_t n, int op) {
  T* __restrict__ a = reinterpret_cast<T*>(acc_);
  const T* __restrict__ s = reinterpret_cast<const T*>(src_);
  switch (op) {
    case FMPI_SUM: for (size_t i = 0; i < n; ++i) a[i] = a[i] + s[i]; break;
    case FMPI_PROD: for (size_t i = 0; i < n; ++i) a[i] = a[i] * s[i]; break;
    case FMPI_MIN: for (size_t i = 0; i < n; ++i) a[i] = s[i] < a[i] ? s[i] : a[i]; break;
    default: for (size_t i = 0; i < n; ++i) a[i] = a[i] < s[i] ? s[i] : a[i]; break;
  }
}

void accumulate_dispatch(int dt, uint8_t* acc, const uint8_t* src, size_t n, int op) {
  switch (dt) {
    case FMPI_F32: accumulate<float>(acc, src, n, op); break;
    case FMPI_F64: accumulate<double>(acc, src, n, op); break;
    case FMPI_I32: accumulate<int32_t>(acc, src, n, op); break;
    case FMPI_I64: accumulate<int64_t>(acc, src, n, op); break;
    case FMPI_I16: accumulate<int16_t>(acc, src, n, op); break;
    case FMPI_U8: accumulate<uint8_t>(acc, src, n, op); break;
    case FMPI_I8: accumulate<int8_t>(acc, src, n, op); break;
    default: break;   // 16-bit floats never come here (fmpi_host_allreduce_symm routes them away)
  }
}

}  // namespace

namespace fmpi_host_detail {
int fail(const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return 1;
}
double now_s() { return ::now_s(); }
size_t dtype_size(int dtype) { return ::dtype_size(dtype); }
void reduce_slots(int dtype, const uint8_t* slots, size_t stride, int world, uint8_t* out,
                  size_t lo, size_t hi, int op) {
  reduce_dispatch(dtype, slots, stride, world, out, lo, hi, op);
}
}  // namespace fmpi_host_detail

// `rank` / `world` are the shared-memory participants, i.e. the ranks of THIS node. With a
// leader mesh attached (`net`) the job has nnodes * world ranks and global rank
// node * world + rank; without one the two notions coincide.
struct fmpi_host_comm {
  int rank = 0, world = 1;
  int node = 0, nnodes = 1;
  fmpi_net* net = nullptr;
  bool heap = false;          // base came from aligned_alloc (one rank per node: no shm needed)
  size_t slot_bytes = 0;
  double timeout_s = 300.0;
  std::string name;
  size_t map_bytes = 0;
  uint8_t* base = nullptr;
  Header* hdr = nullptr;
  uint8_t* slots = nullptr;   // world * slot_bytes
  uint8_t* result = nullptr;  // slot_bytes
  uint8_t* result2 = nullptr; // slot_bytes; alternates with `result` in the pipelined path
  unsigned symm_seq = 0;             // symmetric host blocks allocated so far (names them)
  std::vector<uint8_t> symm_acc;     // private accumulator of fmpi_host_allreduce_symm
  uint32_t local_gen = 0;

  int barrier() {
    if (world == 1) return 0;
    const uint32_t gen = local_gen;
    if (hdr->bar_count.fetch_add(1, std::memory_order_acq_rel) + 1 == uint32_t(world)) {
      hdr->bar_count.store(0, std::memory_order_relaxed);
      hdr->bar_gen.store(gen + 1, std::memory_order_release);
    } else {
      Backoff b;
      const double t0 = now_s();
      while (hdr->bar_gen.load(std::memory_order_acquire) == gen) {
        b.pause();
        if ((b.n & 1023) == 0) {
          if (hdr->aborted.load(std::memory_order_relaxed))
            return fail("rank %d: a peer aborted the job", rank);
          if (now_s() - t0 > timeout_s) {
            hdr->aborted.store(1, std::memory_order_relaxed);
            return fail("rank %d: barrier timed out after %.0f s (a peer is dead or out of step)",
                        rank, timeout_s);
          }
        }
      }
    }
    local_gen = gen + 1;
    return 0;
  }

  // Every rank of the node calls this after its inter-node step (rc = what that step returned;
  // 0 for a rank that had nothing to send). A failure poisons the node so the ranks waiting at
  // the barrier give up too; otherwise the node meets and may read what came back.
  int leader_phase(int rc) {
    if (rc) {
      if (hdr) hdr->aborted.store(1, std::memory_order_relaxed);
      return 1;
    }
    return barrier();
  }

  // Slice [lo, hi) of a chunk of n units that this rank owns: it reduces that slice inside the
  // node and it is the one that carries it between nodes — local rank l of every node forms
  // "rail" l with its own sockets, so a node's network traffic is spread over all of its ranks.
  void my_slice(size_t n, size_t* lo, size_t* hi) const {
    const size_t per = (n + size_t(world) - 1) / size_t(world);
    *lo = per * size_t(rank) < n ? per * size_t(rank) : n;
    *hi = *lo + per < n ? *lo + per : n;
  }
};

namespace {
// all-reduces of at least this many bytes in a multi-node job overlap their network step with
// the shared-memory steps of the neighbouring chunks (allreduce_many_pipelined below)
// (measured over loopback TCP, 2 nodes x 2 ranks: 1 MB is faster unpipelined, 8 MB 1.8x faster
// pipelined)
// FMPI_HOST_PIPELINE_MIN_BYTES overrides (tests force the path with tiny messages).
size_t pipeline_min_bytes() {
  static const size_t v = [] {
    const char* e = getenv("FMPI_HOST_PIPELINE_MIN_BYTES");
    const long long x = e ? atoll(e) : 0;
    return x > 0 ? size_t(x) : size_t(4) << 20;
  }();
  return v;
}
int allreduce_many_pipelined(fmpi_host_comm* c, int n, void* const* bufs, const uint64_t* counts,
                             size_t es, int dtype, int op, size_t total);
}  // namespace

extern "C" {

const char* fmpi_last_error(void) { return g_err.c_str(); }

int fmpi_host_init(const char* name, int rank, int world, size_t slot_bytes, double timeout_s,
                   fmpi_host_comm** out) {
  if (!name || !out || world < 1 || rank < 0 || rank >= world)
    return fail("fmpi_host_init: bad arguments (rank %d, world %d)", rank, world);
  slot_bytes = (slot_bytes + 63) & ~size_t(63);
  if (slot_bytes == 0) slot_bytes = size_t(4) << 20;
  auto* c = new fmpi_host_comm;
  c->rank = rank;
  c->world = world;
  c->slot_bytes = slot_bytes;
  c->timeout_s = timeout_s > 0 ? timeout_s : 300.0;
  c->name = name;
  const size_t hdr_bytes = (sizeof(Header) + 4095) & ~size_t(4095);
  c->map_bytes = hdr_bytes + size_t(world + 2) * slot_bytes;

  int fd = -1;
  if (rank == 0) {
    shm_unlink(name);  // stale segment from a crashed job with the same name
    fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) { delete c; return fail("shm_open(create %s): %s", name, strerror(errno)); }
    // allocate the pages now: on a machine whose /dev/shm is too small (64 MB is a common
    // container default) this is an error message here instead of a SIGBUS at the first touch
    const int e = posix_fallocate(fd, 0, off_t(c->map_bytes));
    if (e != 0) {
      close(fd); shm_unlink(name);
      const size_t need = c->map_bytes;
      delete c;
      return fail("cannot allocate %zu bytes of shared memory for the host world: %s "
                  "(is /dev/shm large enough? lower Init(slot_bytes=...) or enlarge it)", need, strerror(e));
    }
  } else {
    Backoff b;
    const double t0 = now_s();
    for (;;) {
      fd = shm_open(name, O_RDWR, 0600);
      if (fd >= 0) {
        struct stat st;
        if (fstat(fd, &st) == 0 && size_t(st.st_size) >= c->map_bytes) break;
        close(fd);
        fd = -1;
      }
      if (now_s() - t0 > c->timeout_s) { delete c; return fail("rank %d: timed out opening %s", rank, name); }
      b.pause();
    }
  }
  void* p = mmap(nullptr, c->map_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (p == MAP_FAILED) { const int e = errno; delete c; return fail("mmap: %s", strerror(e)); }
  c->base = static_cast<uint8_t*>(p);
  c->hdr = reinterpret_cast<Header*>(p);
  c->slots = c->base + hdr_bytes;
  c->result = c->slots + size_t(world) * slot_bytes;
  c->result2 = c->result + slot_bytes;

  if (rank == 0) {
    c->hdr->world = uint32_t(world);
    c->hdr->slot_bytes = slot_bytes;
    c->hdr->attached.store(0);
    c->hdr->detached.store(0);
    c->hdr->bar_count.store(0);
    c->hdr->bar_gen.store(0);
    c->hdr->aborted.store(0);
    c->hdr->magic.store(kMagic, std::memory_order_release);
  } else {
    Backoff b;
    const double t0 = now_s();
    while (c->hdr->magic.load(std::memory_order_acquire) != kMagic) {
      if (now_s() - t0 > c->timeout_s) {
        munmap(p, c->map_bytes); delete c;
        return fail("rank %d: timed out waiting for rank 0 to initialise %s", rank, name);
      }
      b.pause();
    }
    if (c->hdr->world != uint32_t(world) || c->hdr->slot_bytes != slot_bytes) {
      munmap(p, c->map_bytes); delete c;
      return fail("rank %d: world/slot mismatch with rank 0", rank);
    }
  }
  c->hdr->attached.fetch_add(1);
  // Everyone attached => the name is no longer needed; unlink early so nothing leaks on crash.
  {
    Backoff b;
    const double t0 = now_s();
    while (c->hdr->attached.load(std::memory_order_acquire) < uint32_t(world)) {
      if (now_s() - t0 > c->timeout_s) {
        munmap(p, c->map_bytes); delete c;
        return fail("rank %d: timed out waiting for all %d ranks to attach", rank, world);
      }
      b.pause();
    }
  }
  if (rank == 0) shm_unlink(name);
  *out = c;
  return 0;
}

int fmpi_host_finalize(fmpi_host_comm* c) {
  if (!c) return 0;
  if (c->net) fmpi_net_finalize(c->net);
  if (c->heap) {
    free(c->base);
  } else if (c->base) {
    c->hdr->detached.fetch_add(1);
    munmap(c->base, c->map_bytes);
  }
  delete c;
  return 0;
}

int fmpi_host_init_multi(const char* name, int local_rank, int local_world, int node, int nnodes,
                         const char* master_addr, int port, size_t slot_bytes, double timeout_s,
                         fmpi_host_comm** out) {
  if (!out || nnodes < 1 || node < 0 || node >= nnodes || local_world < 1 || local_rank < 0 ||
      local_rank >= local_world)
    return fail("fmpi_host_init_multi: bad arguments (node %d/%d, local rank %d/%d)", node, nnodes,
                local_rank, local_world);
  fmpi_host_comm* c = nullptr;
  if (local_world > 1) {
    if (fmpi_host_init(name, local_rank, local_world, slot_bytes, timeout_s, &c)) return 1;
  } else {
    // a node with a single rank needs no segment: one staging slot + the result slot on the heap
    slot_bytes = (slot_bytes + 63) & ~size_t(63);
    if (slot_bytes == 0) slot_bytes = size_t(4) << 20;
    c = new fmpi_host_comm;
    c->slot_bytes = slot_bytes;
    c->timeout_s = timeout_s > 0 ? timeout_s : 300.0;
    c->name = name ? name : "";
    c->heap = true;
    c->base = static_cast<uint8_t*>(aligned_alloc(4096, 3 * slot_bytes));
    if (!c->base) { delete c; return fail("out of memory for %zu-byte staging", 2 * slot_bytes); }
    c->slots = c->base;
    c->result = c->base + slot_bytes;
    c->result2 = c->result + slot_bytes;
  }
  c->node = node;
  c->nnodes = nnodes;
  if (nnodes > 1) {
    // one mesh per rail: local rank l of every node, rendezvous on port + l of node 0
    if (port + local_world - 1 > 65535) {
      fmpi_host_finalize(c);
      return fail("ports %d..%d are out of range", port, port + local_world - 1);
    }
    if (fmpi_net_init(master_addr, port + local_rank, node, nnodes, c->timeout_s, &c->net)) {
      if (c->hdr) c->hdr->aborted.store(1, std::memory_order_relaxed);
      fmpi_host_finalize(c);
      return 1;
    }
  }
  // nobody leaves before every rail of the node is connected
  if (c->barrier()) { fmpi_host_finalize(c); return 1; }
  *out = c;
  return 0;
}

int fmpi_host_rank(const fmpi_host_comm* c) { return c ? c->node * c->world + c->rank : -1; }
int fmpi_host_world(const fmpi_host_comm* c) { return c ? c->nnodes * c->world : -1; }

int fmpi_host_barrier(fmpi_host_comm* c) {
  if (!c) return fail("null communicator");
  if (c->barrier()) return 1;
  if (c->nnodes == 1) return 0;
  if (c->leader_phase(c->rank == 0 ? fmpi_net_barrier(c->net) : 0)) return 1;   // rail 0 only
  return 0;
}

int fmpi_host_allreduce(fmpi_host_comm* c, const void* send, void* recv, size_t count, int dtype,
                        int op) {
  if (!c) return fail("null communicator");
  const size_t es = dtype_size(dtype);
  if (es == 0) return fail("unsupported dtype %d", dtype);
  if (op < FMPI_SUM || op > FMPI_MAX) return fail("unsupported op %d", op);
  if (c->world == 1 && c->nnodes == 1) {
    if (recv != send) memmove(recv, send, count * es);
    return 0;
  }
  if (c->nnodes > 1 && count * es >= pipeline_min_bytes()) {   // large and over the network: pipelined
    if (recv != send) memmove(recv, send, count * es);
    void* one[1] = {recv};
    const uint64_t cnt[1] = {count};
    return allreduce_many_pipelined(c, 1, one, cnt, es, dtype, op, count);
  }
  const size_t chunk = c->slot_bytes / es;
  const uint8_t* s = static_cast<const uint8_t*>(send);
  uint8_t* d = static_cast<uint8_t*>(recv);
  for (size_t done = 0; done < count; done += chunk) {
    const size_t n = count - done < chunk ? count - done : chunk;
    memcpy(c->slots + size_t(c->rank) * c->slot_bytes, s + done * es, n * es);
    if (c->barrier()) return 1;
    const size_t per = (n + size_t(c->world) - 1) / size_t(c->world);
    const size_t lo = per * size_t(c->rank) < n ? per * size_t(c->rank) : n;
    const size_t hi = lo + per < n ? lo + per : n;
    if (hi > lo) reduce_dispatch(dtype, c->slots, c->slot_bytes, c->world, c->result, lo, hi, op);
    if (c->barrier()) return 1;
    // the node's sum now sits in the result slot; every rank folds the other nodes' into the
    // slice it owns, over its own rail
    if (c->nnodes > 1 &&
        c->leader_phase(hi > lo ? fmpi_net_allreduce(c->net, c->result + lo * es, hi - lo, dtype, op) : 0))
      return 1;
    memcpy(d + done * es, c->result, n * es);
  }
  // The result slot must stay intact until everyone has copied it out.
  return c->barrier();
}

}  // extern "C"

namespace {

// Position in the concatenation of the leaves; only ever moves forward.
struct LeafCursor {
  int leaf = 0;
  size_t leaf_lo = 0;
};

// Copy the elements [start, start + m) of the concatenated leaves into `slot` (to_slot) or back
// out of it.
void leaf_copy(void* const* bufs, const uint64_t* counts, int n, size_t es, size_t start, size_t m,
               LeafCursor& cur, uint8_t* slot, bool to_slot) {
  while (cur.leaf < n && cur.leaf_lo + counts[cur.leaf] <= start) cur.leaf_lo += counts[cur.leaf++];
  int k = cur.leaf;
  size_t klo = cur.leaf_lo;
  while (k < n && klo < start + m) {
    const size_t a = klo > start ? klo : start;
    const size_t kend = klo + counts[k];
    const size_t b = kend < start + m ? kend : start + m;
    if (b > a) {
      uint8_t* user = static_cast<uint8_t*>(bufs[k]) + (a - klo) * es;
      if (to_slot) memcpy(slot + (a - start) * es, user, (b - a) * es);
      else memcpy(user, slot + (a - start) * es, (b - a) * es);
    }
    klo = kend;
    ++k;
  }
}

// Large messages of a job over several nodes: the network step of chunk k runs on a helper
// thread while this thread copies out chunk k-1 and copies in / reduces chunk k+1 (two result
// slots alternate). Per chunk k:   copy-in(k) | barrier | reduce(k) into result[k%2] | barrier |
// join net(k-1), start net(k) | barrier (every rail's net(k-1) is done) | copy-out(k-1).
int allreduce_many_pipelined(fmpi_host_comm* c, int n, void* const* bufs, const uint64_t* counts,
                             size_t es, int dtype, int op, size_t total) {
  size_t chunk_bytes = total * es / 8;
  if (chunk_bytes < (size_t(512) << 10)) chunk_bytes = size_t(512) << 10;
  if (chunk_bytes > c->slot_bytes) chunk_bytes = c->slot_bytes;
  size_t chunk = (chunk_bytes / es) & ~size_t(63);
  if (chunk == 0) chunk = chunk_bytes / es;        // tiny slots (tests)
  uint8_t* mine = c->slots + size_t(c->rank) * c->slot_bytes;
  uint8_t* res[2] = {c->result, c->result2};
  LeafCursor in_cur, out_cur;
  std::future<int> net;             // the network step in flight (chunk k-1)
  size_t prev_start = 0, prev_m = 0;
  int k = 0;
  auto fail_out = [&](int rc) {
    if (net.valid()) net.wait();
    return rc;
  };
  for (size_t done = 0; done < total; done += chunk, ++k) {
    const size_t m = total - done < chunk ? total - done : chunk;
    leaf_copy(bufs, counts, n, es, done, m, in_cur, mine, true);
    if (c->barrier()) return fail_out(1);
    size_t lo, hi;
    c->my_slice(m, &lo, &hi);
    uint8_t* r = res[k & 1];
    if (hi > lo) reduce_dispatch(dtype, c->slots, c->slot_bytes, c->world, r, lo, hi, op);
    if (c->barrier()) return fail_out(1);
    const int rc_prev = net.valid() ? net.get() : 0;
    if (rc_prev) fail("rank %d: the inter-node step of an all-reduce failed (a node is lost or out of step)", c->rank);
    if (rc_prev == 0 && hi > lo) {
      fmpi_net* rail = c->net;
      uint8_t* slice = r + lo * es;
      const size_t cnt = hi - lo;
      net = std::async(std::launch::async, [rail, slice, cnt, dtype, op] {
        return fmpi_net_allreduce(rail, slice, cnt, dtype, op);
      });
    }
    if (c->leader_phase(rc_prev)) return fail_out(1);       // barrier: every rail finished chunk k-1
    if (k > 0) leaf_copy(bufs, counts, n, es, prev_start, prev_m, out_cur, res[(k - 1) & 1], false);
    prev_start = done;
    prev_m = m;
  }
  const int rc_last = net.valid() ? net.get() : 0;
  if (rc_last) fail("rank %d: the inter-node step of an all-reduce failed (a node is lost or out of step)", c->rank);
  if (c->leader_phase(rc_last)) return 1;
  leaf_copy(bufs, counts, n, es, prev_start, prev_m, out_cur, res[(k - 1) & 1], false);
  return c->barrier();
}

}  // namespace

extern "C" {

// The leaves of a gradient tree as ONE message: pieces are copied from the leaves straight
// into the staging slot and from the result slot straight back, so a tree costs the barriers of
// one message and no intermediate flat copy (the reference sends one MPI message per leaf,
// src/optimizer.jl:55-58).
int fmpi_host_allreduce_many(fmpi_host_comm* c, int n, void* const* bufs, const uint64_t* counts,
                             int dtype, int op) {
  if (!c) return fail("null communicator");
  const size_t es = dtype_size(dtype);
  if (es == 0) return fail("unsupported dtype %d", dtype);
  if (op < FMPI_SUM || op > FMPI_MAX) return fail("unsupported op %d", op);
  if (n < 0 || (n > 0 && (!bufs || !counts))) return fail("fmpi_host_allreduce_many: bad arguments");
  if ((c->world == 1 && c->nnodes == 1) || n == 0) return 0;
  size_t total = 0;
  for (int k = 0; k < n; ++k) total += counts[k];
  if (c->nnodes > 1 && total * es >= pipeline_min_bytes())
    return allreduce_many_pipelined(c, n, bufs, counts, es, dtype, op, total);
  const size_t chunk = c->slot_bytes / es;
  uint8_t* mine = c->slots + size_t(c->rank) * c->slot_bytes;
  int leaf = 0;       // first leaf that still has elements at or after `done`
  size_t leaf_lo = 0; // index of that leaf's first element in the concatenation
  for (size_t done = 0; done < total; done += chunk) {
    const size_t m = total - done < chunk ? total - done : chunk;
    while (leaf < n && leaf_lo + counts[leaf] <= done) leaf_lo += counts[leaf++];
    // walk the leaves overlapping [done, done + m) twice: copy in, then copy out
    for (int pass = 0; pass < 2; ++pass) {
      int k = leaf;
      size_t klo = leaf_lo;
      while (k < n && klo < done + m) {
        const size_t a = klo > done ? klo : done;
        const size_t kend = klo + counts[k];
        const size_t b = kend < done + m ? kend : done + m;
        if (b > a) {
          uint8_t* user = static_cast<uint8_t*>(bufs[k]) + (a - klo) * es;
          if (pass == 0) memcpy(mine + (a - done) * es, user, (b - a) * es);
          else memcpy(user, c->result + (a - done) * es, (b - a) * es);
        }
        klo = kend;
        ++k;
      }
      if (pass == 1) break;
      if (c->barrier()) return 1;
      const size_t per = (m + size_t(c->world) - 1) / size_t(c->world);
      const size_t lo = per * size_t(c->rank) < m ? per * size_t(c->rank) : m;
      const size_t hi = lo + per < m ? lo + per : m;
      if (hi > lo) reduce_dispatch(dtype, c->slots, c->slot_bytes, c->world, c->result, lo, hi, op);
      if (c->barrier()) return 1;
      if (c->nnodes > 1 &&
          c->leader_phase(hi > lo ? fmpi_net_allreduce(c->net, c->result + lo * es, hi - lo, dtype, op) : 0))
        return 1;
    }
  }
  return c->barrier();
}

int fmpi_host_reduce(fmpi_host_comm* c, const void* send, void* recv, size_t count, int dtype,
                     int op, int root) {
  if (!c) return fail("null communicator");
  const size_t es = dtype_size(dtype);
  if (es == 0) return fail("unsupported dtype %d", dtype);
  if (op < FMPI_SUM || op > FMPI_MAX) return fail("unsupported op %d", op);
  if (root < 0 || root >= c->nnodes * c->world) return fail("bad root %d", root);
  if (c->world == 1 && c->nnodes == 1) {
    if (recv != send) memmove(recv, send, count * es);
    return 0;
  }
  const int root_node = root / c->world;
  const bool is_root = c->node == root_node && c->rank == root % c->world;
  const size_t chunk = c->slot_bytes / es;
  const uint8_t* s = static_cast<const uint8_t*>(send);
  uint8_t* d = static_cast<uint8_t*>(recv);
  for (size_t done = 0; done < count; done += chunk) {
    const size_t n = count - done < chunk ? count - done : chunk;
    memcpy(c->slots + size_t(c->rank) * c->slot_bytes, s + done * es, n * es);
    if (c->barrier()) return 1;
    const size_t per = (n + size_t(c->world) - 1) / size_t(c->world);
    const size_t lo = per * size_t(c->rank) < n ? per * size_t(c->rank) : n;
    const size_t hi = lo + per < n ? lo + per : n;
    if (hi > lo) reduce_dispatch(dtype, c->slots, c->slot_bytes, c->world, c->result, lo, hi, op);
    if (c->barrier()) return 1;
    if (c->nnodes > 1 &&
        c->leader_phase(hi > lo ? fmpi_net_reduce(c->net, c->result + lo * es, hi - lo, dtype, op, root_node) : 0))
      return 1;
    if (is_root) memcpy(d + done * es, c->result, n * es);
  }
  return c->barrier();
}

// ---- symmetric host memory: the CPU twin of the device side's symmetric segments -------------
// Every rank allocates `nbytes` in a shared-memory object of its own and maps everybody else's.
// An all-reduce on such memory needs no staging copies: rank r reduces slice r over the W
// mapped copies (rank order) into its own copy, and after one barrier everybody copies the
// other slices from their owners. Per rank that is ~3 passes over the message instead of the
// ~5.3 of the slot path (copy in, reduce, copy out).
struct fmpi_host_symm {
  size_t bytes = 0;
  int world = 1, rank = 0;
  std::vector<uint8_t*> peer;   // peer[q] = rank q's block in this process
};

int fmpi_host_symm_alloc(fmpi_host_comm* c, size_t nbytes, fmpi_host_symm** out) {
  if (!c || !out || nbytes == 0) return fail("fmpi_host_symm_alloc: bad arguments");
  const size_t bytes = (nbytes + 4095) & ~size_t(4095);
  auto* s = new fmpi_host_symm;
  s->bytes = bytes;
  s->world = c->world;
  s->rank = c->rank;
  s->peer.assign(size_t(c->world), nullptr);
  const unsigned seq = c->symm_seq++;                 // same on every rank: allocation is collective
  auto name_of = [&](int q) {
    char buf[320];
    snprintf(buf, sizeof(buf), "%.200s_s%u_r%d", c->name.empty() ? "/fmpi_anon" : c->name.c_str(), seq, q);
    return std::string(buf);
  };
  auto map_one = [&](int q, bool create) -> uint8_t* {
    const std::string nm = name_of(q);
    if (create) shm_unlink(nm.c_str());
    const int fd = shm_open(nm.c_str(), create ? (O_CREAT | O_EXCL | O_RDWR) : O_RDWR, 0600);
    if (fd < 0) return nullptr;
    if (create && posix_fallocate(fd, 0, off_t(bytes)) != 0) { close(fd); return nullptr; }  // see fmpi_host_init
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    return p == MAP_FAILED ? nullptr : static_cast<uint8_t*>(p);
  };
  bool ok = (s->peer[size_t(c->rank)] = map_one(c->rank, true)) != nullptr;
  if (ok) memset(s->peer[size_t(c->rank)], 0, bytes);                       // first touch on the owner
  if (c->barrier()) { delete s; return 1; }                                  // every block exists (or failed)
  for (int q = 0; ok && q < c->world; ++q)
    if (q != c->rank) ok = (s->peer[size_t(q)] = map_one(q, false)) != nullptr;
  if (c->barrier()) { delete s; return 1; }                                  // everybody has mapped
  shm_unlink(name_of(c->rank).c_str());                                      // names no longer needed
  // agree on the outcome: one failed rank fails the allocation everywhere
  uint8_t* flag = c->slots + size_t(c->rank) * c->slot_bytes;
  flag[0] = ok ? 1 : 0;
  if (c->barrier()) { delete s; return 1; }
  bool all = true;
  for (int q = 0; q < c->world; ++q) all = all && (c->slots + size_t(q) * c->slot_bytes)[0] == 1;
  if (c->barrier()) { delete s; return 1; }
  if (!all) {
    for (uint8_t* p : s->peer) if (p) munmap(p, bytes);
    delete s;
    return fail("rank %d: could not create or map the symmetric host blocks (%zu bytes each)", c->rank, bytes);
  }
  *out = s;
  return 0;
}

void* fmpi_host_symm_ptr(fmpi_host_symm* s) { return s ? s->peer[size_t(s->rank)] : nullptr; }
size_t fmpi_host_symm_bytes(fmpi_host_symm* s) { return s ? s->bytes : 0; }

int fmpi_host_symm_free(fmpi_host_comm* c, fmpi_host_symm* s) {
  if (!s) return 0;
  const int rc = c ? c->barrier() : 0;        // nobody is still reading a peer's block
  for (uint8_t* p : s->peer) if (p) munmap(p, s->bytes);
  delete s;
  return rc;
}

int fmpi_host_allreduce_symm(fmpi_host_comm* c, fmpi_host_symm* s, size_t offset_bytes, size_t count,
                             int dtype, int op) {
  if (!c || !s) return fail("null communicator or segment");
  const size_t es = dtype_size(dtype);
  if (es == 0) return fail("unsupported dtype %d", dtype);
  if (op < FMPI_SUM || op > FMPI_MAX) return fail("unsupported op %d", op);
  if (offset_bytes % es || offset_bytes + count * es > s->bytes)
    return fail("range [%zu, +%zu elements) is outside the %zu-byte symmetric block", offset_bytes, count, s->bytes);
  if (s->world != c->world) return fail("segment belongs to another world");
  if (count == 0) return 0;
  uint8_t* mine = s->peer[size_t(c->rank)] + offset_bytes;
  // rails need the slots; 16-bit floats accumulate in fp32 across all ranks in the slot path
  // (one rounding), which a pairwise fold here would not reproduce
  if (c->nnodes > 1 || dtype == FMPI_F16 || dtype == FMPI_BF16)
    return fmpi_host_allreduce(c, mine, mine, count, dtype, op);
  if (c->world == 1) return 0;
  if (c->barrier()) return 1;                            // everybody's operand is complete
  size_t lo, hi;
  c->my_slice(count, &lo, &hi);
  if (hi > lo) {
    // slice r of every block is read by rank r only, so the owner may overwrite its own copy of
    // it as soon as it has folded the W copies (rank order, like the slot path) in a private
    // accumulator
    // accumulator, in cache-sized blocks: each block is folded over all W copies and written
    // out before the next one is touched
    const size_t blk = (size_t(256) << 10) / es;
    if (c->symm_acc.size() < blk * es) c->symm_acc.resize(blk * es);
    uint8_t* acc = c->symm_acc.data();
    for (size_t b0 = lo; b0 < hi; b0 += blk) {
      const size_t cnt = hi - b0 < blk ? hi - b0 : blk;
      memcpy(acc, s->peer[0] + offset_bytes + b0 * es, cnt * es);
      for (int q = 1; q < c->world; ++q)
        accumulate_dispatch(dtype, acc, s->peer[size_t(q)] + offset_bytes + b0 * es, cnt, op);
      memcpy(mine + b0 * es, acc, cnt * es);
    }
  }
  if (c->barrier()) return 1;                            // every owner has published its slice
  for (int q = 0; q < c->world; ++q) {
    if (q == c->rank) continue;
    const size_t per = (count + size_t(c->world) - 1) / size_t(c->world);
    const size_t qlo = per * size_t(q) < count ? per * size_t(q) : count;
    const size_t qhi = qlo + per < count ? qlo + per : count;
    if (qhi > qlo) memcpy(mine + qlo * es, s->peer[size_t(q)] + offset_bytes + qlo * es, (qhi - qlo) * es);
  }
  return c->barrier();                                   // nobody is still reading my block
}

// Broadcast on symmetric host memory: everybody copies the range straight out of the root's
// block (one pass, no slots). Over several nodes the slot path carries it.
int fmpi_host_bcast_symm(fmpi_host_comm* c, fmpi_host_symm* s, size_t offset_bytes, size_t nbytes, int root) {
  if (!c || !s) return fail("null communicator or segment");
  if (root < 0 || root >= c->nnodes * c->world) return fail("bad root %d", root);
  if (offset_bytes + nbytes > s->bytes)
    return fail("range [%zu, +%zu) is outside the %zu-byte symmetric block", offset_bytes, nbytes, s->bytes);
  if (s->world != c->world) return fail("segment belongs to another world");
  uint8_t* mine = s->peer[size_t(c->rank)] + offset_bytes;
  if (c->nnodes > 1) return fmpi_host_bcast(c, mine, nbytes, root);
  if (c->world == 1 || nbytes == 0) return 0;
  if (c->barrier()) return 1;                                   // the root's data is complete
  if (c->rank != root) memcpy(mine, s->peer[size_t(root)] + offset_bytes, nbytes);
  return c->barrier();                                          // the root may overwrite it again
}

int fmpi_host_bcast(fmpi_host_comm* c, void* buf, size_t nbytes, int root) {
  if (!c) return fail("null communicator");
  if (root < 0 || root >= c->nnodes * c->world) return fail("bad root %d", root);
  if (c->world == 1 && c->nnodes == 1) return 0;
  const int root_node = root / c->world;
  const bool is_root = c->node == root_node && c->rank == root % c->world;
  uint8_t* b = static_cast<uint8_t*>(buf);
  for (size_t done = 0; done < nbytes; done += c->slot_bytes) {
    const size_t n = nbytes - done < c->slot_bytes ? nbytes - done : c->slot_bytes;
    if (is_root) memcpy(c->result, b + done, n);
    if (c->barrier()) return 1;
    if (c->nnodes > 1) {   // each rail carries its slice of the chunk from the root's node
      size_t lo, hi;
      c->my_slice(n, &lo, &hi);
      if (c->leader_phase(hi > lo ? fmpi_net_bcast(c->net, c->result + lo, hi - lo, root_node) : 0)) return 1;
    }
    if (!is_root) memcpy(b + done, c->result, n);
    if (c->barrier()) return 1;
  }
  return 0;
}

int fmpi_host_allgather(fmpi_host_comm* c, const void* send, void* recv, size_t nbytes_each) {
  if (!c) return fail("null communicator");
  if (nbytes_each > c->slot_bytes)
    return fail("allgather blob of %zu bytes exceeds the %zu-byte slot", nbytes_each, c->slot_bytes);
  if (c->world == 1 && c->nnodes == 1) {
    if (recv != send) memmove(recv, send, nbytes_each);
    return 0;
  }
  uint8_t* d = static_cast<uint8_t*>(recv);
  if (c->nnodes > 1) {
    // node-major result: the leader packs its node's blobs, swaps packs with the other leaders
    // and leaves the whole table in the result slot
    const size_t node_bytes = size_t(c->world) * nbytes_each;
    const size_t all = size_t(c->nnodes) * node_bytes;
    if (all > c->slot_bytes)
      return fail("allgather of %zu bytes in total exceeds the %zu-byte slot", all, c->slot_bytes);
    memcpy(c->result + size_t(c->node) * node_bytes + size_t(c->rank) * nbytes_each, send,
           nbytes_each);
    if (c->barrier()) return 1;
    if (c->leader_phase(c->rank == 0 ? fmpi_net_allgather(c->net, c->result + size_t(c->node) * node_bytes,
                                                    c->result, node_bytes)
                               : 0))
      return 1;
    memcpy(d, c->result, all);
    return c->barrier();
  }
  memcpy(c->slots + size_t(c->rank) * c->slot_bytes, send, nbytes_each);
  if (c->barrier()) return 1;
  for (int r = 0; r < c->world; ++r)
    memcpy(d + size_t(r) * nbytes_each, c->slots + size_t(r) * c->slot_bytes, nbytes_each);
  return c->barrier();
}

}  // extern "C"
