/*
 * fmpi.h — C ABI of the B200-native data-parallel runtime.
 *
 * One ABI, two consumers: the Python package (ctypes) and the Julia binding (ccall).
 * Everything is pointer / size / enum / cudaStream_t based so torch tensors and CUDA.jl
 * CuArrays can be handed over without copies.
 *
 * Two shared objects implement it:
 *   libfmpi_host.so  (csrc/host_backend.cpp, csrc/fdx.cpp)   — no CUDA dependency; POSIX-shm
 *                     collectives for CPU tensors + Unix-socket FD exchange.
 *   libfmpi_cuda.so  (csrc/symm.cpp, csrc/topo.cpp, csrc/kernels/ *.cu) — sm_100a kernels,
 *                     symmetric heap (VMM / multicast / CUDA-IPC), topology discovery.
 *
 * Reference parity (nikopj/FluxMPI.jl): the host collectives stand in for
 * MPI.Allreduce!/Bcast!/Reduce! on Arrays (src/mpi_extensions.jl:108-111,130-133,152-155);
 * the device collectives replace the CuArray methods (src/mpi_extensions.jl:97-106,119-128,
 * 141-150) and the fused kernels replace src/optimizer.jl:20-23.
 *
 * All functions return 0 on success, non-zero on failure; fmpi_last_error() /
 * fmpi_cuda_last_error() return a thread-local message for the last failure.
 */
#ifndef FMPI_H_
#define FMPI_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FMPI_MAX_RANKS 8      /* one NVSwitch domain */
#define FMPI_MAX_BLOCKS 160   /* max CTAs per rank of any cross-GPU kernel */
#define FMPI_NUM_CHANNELS 4   /* independent signal channels (one per concurrent stream) */

typedef enum {
  FMPI_F32 = 0, FMPI_F64 = 1, FMPI_F16 = 2, FMPI_BF16 = 3,
  FMPI_I32 = 4, FMPI_I64 = 5, FMPI_U8 = 6, FMPI_I8 = 7, FMPI_I16 = 8
} fmpi_dtype;

typedef enum { FMPI_SUM = 0, FMPI_PROD = 1, FMPI_MIN = 2, FMPI_MAX = 3 } fmpi_op;

/* Optimiser rules that have a fused kernel. Semantics follow Optimisers.jl
 * (the package the reference plugs into at src/optimizer.jl:20-25). */
typedef enum {
  FMPI_RULE_DESCENT = 0,   /* x -= eta*g */
  FMPI_RULE_MOMENTUM = 1,  /* v = rho*v + eta*g ; x -= v */
  FMPI_RULE_NESTEROV = 2,  /* d = -rho^2*v + (1+rho)*eta*g ; v = rho*v - eta*g ; x -= d */
  FMPI_RULE_ADAM = 3,      /* m,v moments; x -= eta * mhat / (sqrt(vhat)+eps) */
  FMPI_RULE_ADAMW = 4      /* Adam then x -= lambda*x (OptimiserChain(Adam, WeightDecay)) */
} fmpi_rule;

/* Hyper-parameter block of the fused rule kernels (fmpi_fused_step, fmpi_multi_tensor_step).
 * 48 bytes. */
typedef struct {
  int32_t rule;          /* fmpi_rule */
  float eta, beta1, beta2, eps, lambda, rho;
  float grad_scale;      /* multiplies the summed gradient (1 = reference SUM semantics) */
  int32_t zero_grad;     /* != 0: each rank clears its gradient bucket after the end barrier
                            (fused device step only) */
  uint64_t eta_dev;      /* optional device pointer to a float that overrides eta (learning-rate
                            schedules under CUDA-graph replay); 0 = use eta. Device only. */
} fmpi_hyper;

/* ------------------------------------------------------------------ host backend */

typedef struct fmpi_host_comm fmpi_host_comm;

const char* fmpi_last_error(void);

/* Join (rank 0: create) the shared-memory world `name` ("/fmpi_<job>"); slot_bytes is the
 * per-rank staging size (messages larger than that are chunked). timeout_s bounds every wait
 * so a dead peer turns into an error instead of a hang. */
int fmpi_host_init(const char* name, int rank, int world, size_t slot_bytes, double timeout_s,
                   fmpi_host_comm** out);
int fmpi_host_finalize(fmpi_host_comm* c);
int fmpi_host_rank(const fmpi_host_comm* c);
int fmpi_host_world(const fmpi_host_comm* c);
int fmpi_host_barrier(fmpi_host_comm* c);
/* recv may alias send (in-place). Result is bitwise identical on every rank (one owner per
 * slice reduces in rank order 0..W-1). */
int fmpi_host_allreduce(fmpi_host_comm* c, const void* send, void* recv, size_t count,
                        int dtype, int op);
/* In-place all-reduce of n separate arrays of one dtype as a single message (a gradient tree):
 * the barriers of one message, no flattening copy. */
int fmpi_host_allreduce_many(fmpi_host_comm* c, int n, void* const* bufs, const uint64_t* counts,
                             int dtype, int op);
/* Symmetric host memory (the CPU twin of fmpi_symm_*): every rank allocates nbytes in shared
 * memory and maps its node peers' blocks; an all-reduce on a range of it runs in place, without
 * staging copies. alloc / free are collective and must be issued in the same order everywhere.
 * Over several nodes, and for 16-bit floats, the all-reduce takes the slot path on the block. */
typedef struct fmpi_host_symm fmpi_host_symm;
int fmpi_host_symm_alloc(fmpi_host_comm* c, size_t nbytes, fmpi_host_symm** out);
void* fmpi_host_symm_ptr(fmpi_host_symm* s);
size_t fmpi_host_symm_bytes(fmpi_host_symm* s);
int fmpi_host_symm_free(fmpi_host_comm* c, fmpi_host_symm* s);
int fmpi_host_allreduce_symm(fmpi_host_comm* c, fmpi_host_symm* s, size_t offset_bytes, size_t count,
                             int dtype, int op);
int fmpi_host_bcast_symm(fmpi_host_comm* c, fmpi_host_symm* s, size_t offset_bytes, size_t nbytes,
                         int root);
int fmpi_host_bcast(fmpi_host_comm* c, void* buf, size_t nbytes, int root);
/* Only root's recv is written; non-root buffers are left untouched
 * (test/test_mpi_extensions.jl:56-60). */
int fmpi_host_reduce(fmpi_host_comm* c, const void* send, void* recv, size_t count, int dtype,
                     int op, int root);
/* all-gather of equal-sized blobs (handle exchange, print ordering, tests) */
int fmpi_host_allgather(fmpi_host_comm* c, const void* send, void* recv, size_t nbytes_each);

/* Optimiser rule over a table of n host tensors in one call, one pass over memory per leaf
 * (the CPU twin of fmpi_multi_tensor_step; the reference leaves this to Optimisers.jl's
 * per-leaf broadcasts, src/optimizer.jl:20-25). p and g have dtype `dtype` (FMPI_F32 or
 * FMPI_F64), m and v are arrays of the same dtype (entries may be NULL when the rule does not
 * use them). hyper = {eta, beta1, beta2, eps, lambda, rho, grad_scale} in double so fp64
 * tensors lose nothing. `step` = completed steps so far (bias correction uses step + 1).
 * Arithmetic is done in the tensors' own precision, in the operation order of rules.py, so the
 * result matches the per-leaf path to rounding. */
int fmpi_host_multi_tensor_step(int n, void* const* p, const void* const* g, void* const* m,
                                void* const* v, const uint64_t* numel, int dtype, int rule,
                                const double* hyper, int64_t step);

/* Unix-domain-socket file-descriptor exchange (SCM_RIGHTS), used to pass CUDA VMM
 * shareable handles between the ranks of one node. */
int fmpi_fdx_open(const char* job, int rank, int* sock_out);
/* Send `fd` to every peer and receive theirs: fds_out[r] for r != rank (fds_out[rank] = fd).
 * All ranks must have returned from fmpi_fdx_open before anyone calls this. */
int fmpi_fdx_exchange(int sock, const char* job, int rank, int world, int fd, int* fds_out,
                      double timeout_s);
/* Root sends one fd to everybody else (multicast object handle). */
int fmpi_fdx_bcast(int sock, const char* job, int rank, int world, int root, int* fd_inout,
                   double timeout_s);
int fmpi_fdx_close(int sock, const char* job, int rank);
const char* fmpi_fdx_last_error(void);

/* ---- more than one node (csrc/net_backend.cpp) --------------------------------------------
 * The reference spans nodes through whatever `mpiexec` starts. Here the ranks of
 * a node keep talking through shared memory and local rank l of every node joins a full TCP
 * mesh with the other nodes' local rank l ("rail" l), which carries slice l of every chunk.
 * Global rank = node * local_world + local_rank (every node runs the same number of ranks).
 *
 * fmpi_net_*: collectives between the members of one mesh (one per node), in place. Node 0
 * listens on master_addr:port for the rendezvous; the others dial in (retrying until
 * timeout_s). Reductions are done by one owner per slice in node order, so every node ends
 * with the same bits. */
typedef struct fmpi_net fmpi_net;
int fmpi_net_init(const char* master_addr, int port, int node, int nnodes, double timeout_s,
                  fmpi_net** out);
int fmpi_net_finalize(fmpi_net* n);
int fmpi_net_barrier(fmpi_net* n);
int fmpi_net_allreduce(fmpi_net* n, void* buf, size_t count, int dtype, int op);
/* only root_node's buf holds the result afterwards; the others' buffers are scratch */
int fmpi_net_reduce(fmpi_net* n, void* buf, size_t count, int dtype, int op, int root_node);
int fmpi_net_bcast(fmpi_net* n, void* buf, size_t nbytes, int root_node);
int fmpi_net_allgather(fmpi_net* n, const void* send, void* recv, size_t nbytes_each);

/* Two-level host world: shared memory inside the node + one mesh per rail between nodes (local
 * rank l rendezvouses on port + l of node 0). The returned communicator is used with the same
 * fmpi_host_* calls as a single-node one; ranks and roots are global. `name` must differ
 * between nodes that share a machine (tests). */
int fmpi_host_init_multi(const char* name, int local_rank, int local_world, int node, int nnodes,
                         const char* master_addr, int port, size_t slot_bytes, double timeout_s,
                         fmpi_host_comm** out);

/* ------------------------------------------------------------------ device side */
#ifndef FMPI_HOST_ONLY

typedef void* fmpi_stream; /* cudaStream_t */

const char* fmpi_cuda_last_error(void);

/* ---- topology (csrc/topo.cpp); replaces the reference's blind (rank+1)%ndev device pick
 * (src/common.jl:33-34) with discovery. Writes a JSON document into buf. */
int fmpi_topo_json(char* buf, size_t buflen);
int fmpi_device_count(void);

/* ---- symmetric heap (csrc/symm.cpp) */
typedef struct fmpi_symm fmpi_symm;
enum { FMPI_SYMM_VMM = 0, FMPI_SYMM_IPC = 1, FMPI_SYMM_LOCAL = 2 };

/* Allocate `bytes` (rounded up to the allocation / multicast granularity) on `device`.
 * mode VMM: cuMemCreate with a POSIX-fd shareable handle; IPC: cudaMalloc + cudaIpc handle;
 * LOCAL: plain cudaMalloc, no sharing (world 1 and single-GPU emulation). */
int fmpi_symm_create(int device, size_t bytes, int mode, int want_multicast, int world,
                     fmpi_symm** out);
size_t fmpi_symm_size(const fmpi_symm* s);
void* fmpi_symm_local_ptr(const fmpi_symm* s);
int fmpi_symm_mode(const fmpi_symm* s);
/* VMM: the fd to pass to peers. */
int fmpi_symm_export_fd(fmpi_symm* s, int* fd_out);
/* IPC: 64-byte cudaIpcMemHandle_t. */
int fmpi_symm_export_ipc(fmpi_symm* s, void* handle64);
/* Map peer `r`'s allocation; afterwards fmpi_symm_peer_ptr(s, r) is valid. */
int fmpi_symm_import_fd(fmpi_symm* s, int r, int fd);
int fmpi_symm_import_ipc(fmpi_symm* s, int r, const void* handle64);
int fmpi_symm_set_self(fmpi_symm* s, int rank);
void* fmpi_symm_peer_ptr(const fmpi_symm* s, int r);
/* Multicast (NVLS): root creates + exports, everyone imports/adds its device, then (after a
 * host barrier) binds its memory and maps the multicast VA. */
int fmpi_symm_mc_supported(int device);
int fmpi_symm_mc_create(fmpi_symm* s, int* fd_out);          /* root only */
int fmpi_symm_mc_import(fmpi_symm* s, int fd);               /* non-root */
int fmpi_symm_mc_add_device(fmpi_symm* s);                   /* all, then barrier */
int fmpi_symm_mc_bind_map(fmpi_symm* s);                     /* all, then barrier */
void* fmpi_symm_mc_ptr(const fmpi_symm* s);
int fmpi_symm_destroy(fmpi_symm* s);

/* ---- device communicator: peer pointer tables handed to every kernel by value */
typedef struct {
  int32_t rank;           /* this process's rank (ignored when emulate != 0) */
  int32_t world;
  int32_t emulate;        /* != 0: single-GPU emulation, rank = blockIdx.y (cooperative launch) */
  int32_t channel;        /* signal channel in [0, FMPI_NUM_CHANNELS) */
  void* data[FMPI_MAX_RANKS];      /* symmetric data windows (same layout on every rank) */
  uint32_t* sig[FMPI_MAX_RANKS];   /* symmetric signal pads (fmpi_sigpad_bytes() each) */
  void* mc_data;          /* multicast alias of the data window, or NULL */
  uint64_t data_bytes;    /* size of each data window */
  uint32_t* ctrl;         /* device address of the health page from fmpi_health_page(), or NULL:
                             word 0 = abort request (host -> kernels), word 1 = error (kernel ->
                             host). NULL: waits are unbounded (single-GPU emulation) */
  uint64_t timeout_ns;    /* budget of one cross-GPU wait; 0 = no time limit */
} fmpi_devcomm;

/* The health page: 64 bytes of pinned host memory mapped into the device (one per process and
 * device, created on first use with that device current). A cross-GPU wait that has run out of
 * fmpi_devcomm::timeout_ns sets word 1 and returns, and so does every other wait on this GPU
 * from then on, so a lost peer ends in an error on the survivors instead of GPUs that spin
 * inside a kernel for ever; the host can set word 0 to the same effect (it does when the host
 * world reports a dead rank). After an error the data a collective produced is undefined.
 *   fmpi_health_page   host and device address of the page
 *   fmpi_health_check  0 = healthy; 1 = a wait was abandoned (fmpi_cuda_last_error says why) */
int fmpi_health_page(uint32_t** host_ptr, uint32_t** dev_ptr);
int fmpi_health_check(void);
int fmpi_health_abort(void);
int fmpi_health_reset(void);

size_t fmpi_sigpad_bytes(void);

/* Launch-shape hints; 0 = library default chosen from message size. */
typedef struct { int32_t blocks; int32_t threads; int32_t algo; } fmpi_launch;
enum { FMPI_ALGO_AUTO = 0, FMPI_ALGO_ONESHOT = 1, FMPI_ALGO_TWOSHOT = 2, FMPI_ALGO_NVLS = 3,
       FMPI_ALGO_PIPE = 4, /* chunk-pipelined staging (ordinary buffers); symmetric bcast: pull + multicast re-emit */
       FMPI_ALGO_PULL_P2P = 5 /* symmetric bcast: pull the slice from the root, re-emit with peer stores */ };

/* Collectives on buffers that already live in the symmetric window at byte offset `off`
 * (in-place, result in every rank's window for allreduce/bcast, root only for reduce). */
int fmpi_allreduce_symm(const fmpi_devcomm* c, uint64_t off, uint64_t count, int dtype, int op,
                        const fmpi_launch* l, fmpi_stream st);
int fmpi_bcast_symm(const fmpi_devcomm* c, uint64_t off, uint64_t nbytes, int root,
                    const fmpi_launch* l, fmpi_stream st);
int fmpi_reduce_symm(const fmpi_devcomm* c, uint64_t off, uint64_t count, int dtype, int op,
                     int root, const fmpi_launch* l, fmpi_stream st);

/* Collectives on arbitrary device buffers: the kernel itself stages through the symmetric
 * window at `stage_off` (push mode: writes go straight into the peers' slots, so there is no
 * separate copy-in). in/out are per-emulated-rank arrays; real mode uses entry [0].
 * stage_bytes is the size of the staging area available at stage_off. */
int fmpi_allreduce_any(const fmpi_devcomm* c, const void* const* in, void* const* out,
                       uint64_t count, int dtype, int op, uint64_t stage_off,
                       uint64_t stage_bytes, const fmpi_launch* l, fmpi_stream st);
int fmpi_bcast_any(const fmpi_devcomm* c, void* const* buf, uint64_t nbytes, int root,
                   uint64_t stage_off, uint64_t stage_bytes, const fmpi_launch* l,
                   fmpi_stream st);
int fmpi_reduce_any(const fmpi_devcomm* c, const void* const* in, void* const* out,
                    uint64_t count, int dtype, int op, int root, uint64_t stage_off,
                    uint64_t stage_bytes, const fmpi_launch* l, fmpi_stream st);
int fmpi_barrier_dev(const fmpi_devcomm* c, fmpi_stream st);
/* Diagnostics of the chunk-pipelined kernels: with FMPI_PIPE_TRACE=1 in the environment the first
 * CTA of each role (in, link, out) stamps %globaltimer per chunk — start, after its wait, after
 * its work, after its hand-over — and this call copies the [3][256][4] stamps of the last
 * pipelined collective to the host (synchronises the device). */
int fmpi_pipe_trace_read(uint64_t* host_out, size_t nwords);

/* ---- fused gradient allreduce (+) optimiser step (the headline kernel).
 * Gradients (dtype gdt) at g_off and parameters (dtype pdt) at p_off live in the symmetric
 * window; the element range [lo, hi) is one bucket. Each rank owns a 1/W slice of the bucket:
 * it reduces that slice across all ranks (rank-ordered fp32 sum, or multimem.ld_reduce),
 * applies the rule in registers and writes the new parameters into EVERY rank's window, so
 * replicas stay bit-identical. Moments m, v (fp32) and the optional fp32 master copy are
 * process-local arrays indexed like the window (only the owned slice is touched).
 * step_dev points at a device int64 holding the number of completed steps (read only; advance
 * it with fmpi_step_advance so CUDA-graph replays see fresh bias corrections).
 * Hyper-parameters: fmpi_hyper, defined with the rules at the top of this header. */

int fmpi_fused_step(const fmpi_devcomm* c, uint64_t g_off, int gdt, uint64_t p_off, int pdt,
                    float* const* m, float* const* v, float* const* master,
                    uint64_t lo, uint64_t hi, const fmpi_hyper* h, const int64_t* const* step_dev,
                    const fmpi_launch* l, fmpi_stream st);
/* One-shot flavour for small buckets (a Quick-Start-sized model, README.md:37-38 of the
 * reference): every rank pushes its gradient bucket into the peers' staging slots
 * ([2 parities][world][slot] at stage_off, slot = stage_bytes / (2 * world) >= bucket bytes),
 * one cross-GPU barrier, then every rank sums the W copies in rank order and applies the rule
 * to the WHOLE bucket locally. The optimiser state is therefore replicated (m, v are updated
 * everywhere), unlike fmpi_fused_step where only the owner of a slice touches them. */
int fmpi_fused_step_oneshot(const fmpi_devcomm* c, uint64_t g_off, uint64_t p_off, int dtype,
                            float* const* m, float* const* v, float* const* master, uint64_t lo,
                            uint64_t hi, const fmpi_hyper* h, const int64_t* const* step_dev,
                            uint64_t stage_off, uint64_t stage_bytes, const fmpi_launch* l,
                            fmpi_stream st);
int fmpi_step_advance(int64_t* step_dev, fmpi_stream st);

/* Same rule math with no communication, over a table of scattered tensors (functional
 * `update` on arbitrary trees). Pointer tables are device arrays of n entries. */
int fmpi_multi_tensor_step(int n, void* const* p, const void* const* g, float* const* m,
                           float* const* v, const uint64_t* numel, int pdt, int gdt,
                           const fmpi_hyper* h, int64_t step, fmpi_stream st);

/* ---- tree-level fused kernels (csrc/kernels/tree_step.cu): the functional API on leaves that
 * live wherever autodiff put them (replaces /root/reference/src/optimizer.jl:20-23 and :45-65).
 * `tables` is a per-emulated-rank array (real mode: entry [0]) of DEVICE pointers to
 * uint64 [p[n] | g[n] | m[n] | v[n] | numel[n] | flat_offset_bytes[n]] (offsets 16-byte aligned;
 * m / v entries may be 0 when the rule has no such state). Leaves share one dtype
 * (f32 / bf16 / f16); moments are fp32.
 *   fmpi_tree_step_oneshot  ONE launch: push the gradient leaves into the peers' slots
 *       ([2 parities][world][slot] at stage_off, slot = stage_bytes / (2 * world) >= total_bytes),
 *       one barrier, rank-ordered sum + rule applied in place to p / m / v. `step` = completed
 *       steps (bias correction uses step + 1). The reduced gradient never exists in memory.
 *   fmpi_tree_allreduce     gather the gradient leaves into the symmetric window at `off` and
 *       all-reduce that region in place (two-shot peer-to-peer or in-switch), one launch; follow
 *       it with fmpi_multi_tensor_step / fmpi_unpack reading the region through a pointer table. */
int fmpi_tree_step_oneshot(const fmpi_devcomm* c, int n, const void* const* tables, uint64_t total_bytes,
                           int dtype, const fmpi_hyper* h, int64_t step, uint64_t stage_off,
                           uint64_t stage_bytes, const fmpi_launch* l, fmpi_stream st);
int fmpi_tree_allreduce(const fmpi_devcomm* c, int n, const void* const* tables, uint64_t off,
                        uint64_t total_bytes, int dtype, int op, const fmpi_launch* l, fmpi_stream st);

/* ---- pack / unpack scattered leaves into / out of a flat buffer (pointer-table kernels) */
int fmpi_pack(int n, const void* const* src, const uint64_t* offsets_bytes,
              const uint64_t* nbytes, void* flat, fmpi_stream st);
int fmpi_unpack(int n, void* const* dst, const uint64_t* offsets_bytes, const uint64_t* nbytes,
                const void* flat, fmpi_stream st);

/* ---- diagnostics: sum of squares (gradient norm) and a 64-bit order-independent checksum
 * (bitwise replica-consistency checks). scratch: device buffer of fmpi_reduce_scratch_bytes()
 * that the caller allocates once; out: device double / uint64. */
size_t fmpi_reduce_scratch_bytes(void);
int fmpi_sumsq(const void* x, uint64_t count, int dtype, double* out, void* scratch,
               fmpi_stream st);
int fmpi_checksum(const void* x, uint64_t nbytes, uint64_t* out, void* scratch, fmpi_stream st);

/* fmpi_sumsq for callers that must not allocate device memory (no counterpart in the
 * reference, which has no device code of its own): a host runtime that owns the GPU memory
 * pool, a hook running inside a step whose memory is fully planned, a monitoring thread.
 * Same value as fmpi_sumsq up to fp64 summation order (it folds at most 256 partial sums where
 * fmpi_sumsq folds up to 592), but the scratch is a 2304-byte array inside the library's device
 * image instead of a caller buffer, so the call performs no cudaMalloc / cudaFree / allocating
 * library call; the caller passes only the input and the 8-byte device output.
 *   - at most FMPI_NOALLOC_MAX_COUNT elements, more is refused with an error;
 *   - one scratch array per GPU, shared by all calls: calls that can overlap on the device
 *     (unordered streams) race; calls on one stream are safe;
 *   - the first call on a device resolves the array's address (once per device, with that
 *     device current): issue it before starting any stream capture. */
#define FMPI_NOALLOC_MAX_COUNT (1ull << 28)
int fmpi_sumsq_noalloc(const void* x, uint64_t count, int dtype, double* out, fmpi_stream st);

/* ---- channels-last batch normalisation fused with the residual add and ReLU that follow it
 * (csrc/kernels/batchnorm.cu). x, res, y, dy, dx, dres are [M, C] row-major (NHWC memory) in
 * `dtype` (f32 / bf16 / f16; C a multiple of 4 resp. 8, 16-byte aligned); gamma, beta, the
 * running and saved statistics and the gradients of gamma / beta are fp32 [C]. `scratch` is a
 * device buffer of fmpi_bn_scratch_bytes(C) that calls on one stream may share.
 *   forward : y = relu?(batchnorm(x) (+ res)); updates running stats; saves mean / invstd
 *   apply   : y = relu?(x * scale + shift (+ res))           (inference, precomputed affine)
 *   backward: dx, dgamma, dbeta (and dres = dy masked by the ReLU when res was given); `y` is
 *             the forward output: needed only when relu != 0 AND a residual was added — with
 *             y == NULL the ReLU mask is recomputed from x, gamma, beta and the saved statistics
 *             (two fewer passes over the activations).
 * The statistics are accumulated around a per-channel pivot, so channels whose mean is far from
 * zero keep their variance (no E[x^2] - mean^2 cancellation).
 * The reference ships no model code; this exists because these ops, not the
 * convolutions, dominate the headline ResNet-50 step on B200 (profiles/README.md §7). */
size_t fmpi_bn_scratch_bytes(int C);
int fmpi_bn_forward(const void* x, const void* res, void* y, uint64_t M, int C, int dtype,
                    const float* gamma, const float* beta, float* running_mean, float* running_var,
                    float momentum, float eps, int relu, float* save_mean, float* save_invstd,
                    void* scratch, fmpi_stream st);
int fmpi_bn_apply(const void* x, const void* res, void* y, uint64_t M, int C, int dtype,
                  const float* scale, const float* shift, int relu, fmpi_stream st);
int fmpi_bn_backward(const void* dy, const void* y, const void* x, void* dx, void* dres, uint64_t M,
                     int C, int dtype, const float* gamma, const float* beta, const float* save_mean,
                     const float* save_invstd, float* dgamma, float* dbeta, int relu, void* scratch,
                     fmpi_stream st);

/* ---- channels-last max pooling (csrc/kernels/pool.cu): x [N,H,W,C] -> y [N,OH,OW,C] with
 * OH = (H + 2P - K)/S + 1; idx keeps one byte per output element (window position of the
 * maximum) for the atomics-free gather backward. */
int fmpi_maxpool_fwd(const void* x, void* y, uint8_t* idx, int N, int H, int W, int C, int K, int S,
                     int P, int dtype, fmpi_stream st);
int fmpi_maxpool_bwd(const void* dy, const uint8_t* idx, void* dx, int N, int H, int W, int C, int K,
                     int S, int P, int dtype, fmpi_stream st);

/* Write > L2-size bytes (bench hygiene: flush L2 between timed iterations). */
int fmpi_flush_l2(void* buf, uint64_t nbytes, fmpi_stream st);

#endif /* FMPI_HOST_ONLY */

#ifdef __cplusplus
}
#endif
#endif /* FMPI_H_ */
