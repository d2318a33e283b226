/* C driver of the DEVICE half of include/fmpi.h.
 *
 * The Julia binding (julia/FluxMPIB200/src/device.jl, engine.jl) reaches the kernels through
 * plain `ccall`s; no Julia exists on this image, so this program makes the same call sequence
 * from C — no Python, no ctypes in between — and checks every result against a host loop:
 *
 *   fmpi_symm_create (local) x W -> an emulated fmpi_devcomm (W virtual ranks on one GPU)
 *   fmpi_health_page, fmpi_barrier_dev
 *   fmpi_allreduce_any (one-shot and staged two-shot), fmpi_reduce_any, fmpi_bcast_any
 *   fmpi_allreduce_symm, fmpi_bcast_symm
 *   fmpi_fused_step (Adam, two buckets) + fmpi_step_advance, fmpi_fused_step_oneshot
 *   fmpi_tree_step_oneshot, fmpi_tree_allreduce, fmpi_pack / fmpi_unpack
 *   fmpi_multi_tensor_step, fmpi_sumsq, fmpi_sumsq_noalloc, fmpi_checksum
 *
 * Build and run (tests/test_abi_cuda_c.py does this on a GPU box):
 *   gcc -std=c11 -O1 -Iinclude -I/usr/local/cuda/include csrc/abi_test_cuda.c \
 *       -Lfluxmpi_b200/_lib -lfmpi_cuda -L/usr/local/cuda/lib64 -lcudart -lm -o abi_test_cuda
 * Exit code 0 and a last line "ABI_CUDA_OK <n checks>" mean success.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime_api.h>

#include "fmpi.h"

#define W 4
#define MB (1ull << 20)

static int g_checks = 0;

#define CK(call)                                                                                \
  do {                                                                                          \
    int rc_ = (call);                                                                           \
    if (rc_ != 0) {                                                                             \
      fprintf(stderr, "%s:%d: %s failed: %s\n", __FILE__, __LINE__, #call, fmpi_cuda_last_error()); \
      exit(1);                                                                                  \
    }                                                                                           \
  } while (0)
#define CU(call)                                                                                \
  do {                                                                                          \
    cudaError_t e_ = (call);                                                                    \
    if (e_ != cudaSuccess) {                                                                    \
      fprintf(stderr, "%s:%d: %s: %s\n", __FILE__, __LINE__, #call, cudaGetErrorString(e_));    \
      exit(1);                                                                                  \
    }                                                                                           \
  } while (0)
#define EXPECT(cond, ...)                                     \
  do {                                                        \
    ++g_checks;                                               \
    if (!(cond)) {                                            \
      fprintf(stderr, "%s:%d: CHECK FAILED: ", __FILE__, __LINE__); \
      fprintf(stderr, __VA_ARGS__);                           \
      fprintf(stderr, "\n");                                  \
      exit(2);                                                \
    }                                                         \
  } while (0)

static float frand(unsigned* s) { *s = *s * 1664525u + 1013904223u; return (float)((*s >> 8) & 0xffff) / 65536.0f + 0.25f; }

static void* dalloc(size_t n) { void* p; CU(cudaMalloc(&p, n)); CU(cudaMemset(p, 0, n)); return p; }
static void h2d(void* d, const void* h, size_t n) { CU(cudaMemcpy(d, h, n, cudaMemcpyHostToDevice)); }
static void d2h(void* h, const void* d, size_t n) { CU(cudaMemcpy(h, d, n, cudaMemcpyDeviceToHost)); }

static int close_enough(const float* a, const float* b, size_t n, float tol, const char* what) {
  for (size_t i = 0; i < n; ++i) {
    const float d = fabsf(a[i] - b[i]);
    if (!(d <= tol * (1.0f + fabsf(b[i])))) {
      fprintf(stderr, "%s: element %zu: got %.9g want %.9g\n", what, i, a[i], b[i]);
      return 0;
    }
  }
  return 1;
}

int main(void) {
  if (fmpi_device_count() < 1) { printf("no CUDA device\n"); return 77; }
  CU(cudaSetDevice(0));
  const size_t sig = (fmpi_sigpad_bytes() + 65535) / 65536 * 65536;
  const size_t stage = 8 * MB, heap = 16 * MB, window = 2 * stage + heap;
  fmpi_symm* seg[W];
  uint8_t* base[W];
  fmpi_devcomm c;
  memset(&c, 0, sizeof c);
  c.world = W; c.emulate = 1; c.channel = 0; c.data_bytes = window;
  for (int r = 0; r < W; ++r) {
    CK(fmpi_symm_create(0, sig + window, FMPI_SYMM_LOCAL, 0, 1, &seg[r]));
    base[r] = (uint8_t*)fmpi_symm_local_ptr(seg[r]);
    EXPECT(base[r] != NULL && fmpi_symm_size(seg[r]) >= sig + window, "segment %d", r);
    CU(cudaMemset(base[r], 0, sig + window));
    c.sig[r] = (uint32_t*)base[r];
    c.data[r] = base[r] + sig;
  }
  uint32_t *hp = NULL, *dp = NULL;
  CK(fmpi_health_page(&hp, &dp));
  c.ctrl = dp; c.timeout_ns = 20ull * 1000000000ull;
  const uint64_t ar_off = 0, bc_off = stage, heap_off = 2 * stage;
  fmpi_launch la = {0, 0, FMPI_ALGO_AUTO};
  cudaStream_t st; CU(cudaStreamCreate(&st));
  CK(fmpi_barrier_dev(&c, st));
  CK(fmpi_barrier_dev(&c, st));

  /* ---- all-reduce / reduce / bcast on ordinary buffers */
  const size_t sizes[] = {5, 4096, (1u << 20) + 12};
  unsigned seed = 7;
  for (int si = 0; si < 3; ++si) {
    const size_t n = sizes[si];
    float* h[W]; float* want = calloc(n, sizeof(float)); float* got = malloc(n * sizeof(float));
    void* d[W]; const void* in[W]; void* out[W];
    for (int r = 0; r < W; ++r) {
      h[r] = malloc(n * sizeof(float));
      for (size_t i = 0; i < n; ++i) { h[r][i] = frand(&seed); }
      d[r] = dalloc(n * sizeof(float) + 16);
      h2d(d[r], h[r], n * sizeof(float));
      in[r] = d[r]; out[r] = d[r];
    }
    for (size_t i = 0; i < n; ++i) { float s = h[0][i]; for (int r = 1; r < W; ++r) s += h[r][i]; want[i] = s; }
    /* the bulk must be a multiple of 16 bytes on the two-shot path: split like the bindings do */
    const size_t bulk = (n * 4 % 16 == 0 || n * 4 <= 200 * 1024) ? n : (n * 4 / 16 * 16) / 4;
    CK(fmpi_allreduce_any(&c, in, out, bulk, FMPI_F32, FMPI_SUM, ar_off, stage, &la, st));
    if (bulk < n) {
      const void* in2[W]; void* out2[W];
      for (int r = 0; r < W; ++r) { in2[r] = (float*)d[r] + bulk; out2[r] = (float*)d[r] + bulk; }
      CK(fmpi_allreduce_any(&c, in2, out2, n - bulk, FMPI_F32, FMPI_SUM, ar_off, stage, &la, st));
    }
    CU(cudaStreamSynchronize(st));
    for (int r = 0; r < W; ++r) {
      d2h(got, d[r], n * sizeof(float));
      EXPECT(close_enough(got, want, n, 1e-5f, "allreduce_any"), "allreduce_any n=%zu rank %d", n, r);
    }
    /* reduce to root 2 (max): only the root changes */
    for (int r = 0; r < W; ++r) h2d(d[r], h[r], n * sizeof(float));
    if (n * 4 % 16 == 0 || n * 4 <= 200 * 1024) {
      CK(fmpi_reduce_any(&c, in, out, n, FMPI_F32, FMPI_MAX, 2, ar_off, stage, &la, st));
      CU(cudaStreamSynchronize(st));
      for (int r = 0; r < W; ++r) {
        d2h(got, d[r], n * sizeof(float));
        int ok = 1;
        for (size_t i = 0; i < n && ok; ++i) {
          float m = h[0][i]; for (int q = 1; q < W; ++q) m = h[q][i] > m ? h[q][i] : m;
          ok = got[i] == (r == 2 ? m : h[r][i]);
        }
        EXPECT(ok, "reduce_any n=%zu rank %d", n, r);
      }
    }
    /* bcast from root 1 */
    for (int r = 0; r < W; ++r) h2d(d[r], h[r], n * sizeof(float));
    CK(fmpi_bcast_any(&c, out, n * sizeof(float), 1, bc_off, stage, &la, st));
    CU(cudaStreamSynchronize(st));
    for (int r = 0; r < W; ++r) {
      d2h(got, d[r], n * sizeof(float));
      EXPECT(memcmp(got, h[1], n * sizeof(float)) == 0, "bcast_any n=%zu rank %d", n, r);
    }
    for (int r = 0; r < W; ++r) { free(h[r]); CU(cudaFree(d[r])); }
    free(want); free(got);
  }

  /* ---- in place on the symmetric window */
  {
    const size_t n = (1u << 16) + 4;
    float* h[W]; float* want = calloc(n, sizeof(float)); float* got = malloc(n * sizeof(float));
    for (int r = 0; r < W; ++r) {
      h[r] = malloc(n * sizeof(float));
      for (size_t i = 0; i < n; ++i) h[r][i] = frand(&seed);
      h2d((uint8_t*)c.data[r] + heap_off, h[r], n * sizeof(float));
    }
    for (size_t i = 0; i < n; ++i) { float s = h[0][i]; for (int r = 1; r < W; ++r) s += h[r][i]; want[i] = s; }
    CK(fmpi_allreduce_symm(&c, heap_off, n, FMPI_F32, FMPI_SUM, &la, st));
    CU(cudaStreamSynchronize(st));
    for (int r = 0; r < W; ++r) {
      d2h(got, (uint8_t*)c.data[r] + heap_off, n * sizeof(float));
      EXPECT(close_enough(got, want, n, 1e-5f, "allreduce_symm"), "allreduce_symm rank %d", r);
    }
    h2d((uint8_t*)c.data[3] + heap_off, h[3], n * sizeof(float));
    CK(fmpi_bcast_symm(&c, heap_off, n * sizeof(float), 3, &la, st));
    CU(cudaStreamSynchronize(st));
    for (int r = 0; r < W; ++r) {
      d2h(got, (uint8_t*)c.data[r] + heap_off, n * sizeof(float));
      EXPECT(memcmp(got, h[3], n * sizeof(float)) == 0, "bcast_symm rank %d", r);
    }
    for (int r = 0; r < W; ++r) free(h[r]);
    free(want); free(got);
  }

  /* ---- fused all-reduce (+) Adam: two buckets, two steps, against the host rule */
  {
    const size_t n = 64 * 301;
    const uint64_t p_off = heap_off, g_off = heap_off + 4 * MB, os_off = heap_off + 8 * MB;
    float* p = malloc(n * 4); float* m = calloc(n, 4); float* v = calloc(n, 4); float* g = malloc(n * 4);
    float* got = malloc(n * 4);
    float *dm[W], *dv[W]; int64_t* dstep[W]; const int64_t* cstep[W];
    for (size_t i = 0; i < n; ++i) p[i] = frand(&seed) - 0.75f;
    for (int r = 0; r < W; ++r) {
      h2d((uint8_t*)c.data[r] + p_off, p, n * 4);
      dm[r] = dalloc(n * 4); dv[r] = dalloc(n * 4); dstep[r] = dalloc(8); cstep[r] = dstep[r];
    }
    fmpi_hyper hy; memset(&hy, 0, sizeof hy);
    hy.rule = FMPI_RULE_ADAM; hy.eta = 1e-2f; hy.beta1 = 0.9f; hy.beta2 = 0.999f; hy.eps = 1e-8f;
    hy.grad_scale = 1.0f; hy.zero_grad = 1;
    fmpi_devcomm c2 = c; c2.channel = 2;
    for (int step = 0; step < 2; ++step) {
      for (size_t i = 0; i < n; ++i) g[i] = 0.f;
      for (int r = 0; r < W; ++r) {
        float* gr = malloc(n * 4);
        for (size_t i = 0; i < n; ++i) { gr[i] = frand(&seed) - 0.75f; g[i] += gr[i]; }
        h2d((uint8_t*)c.data[r] + g_off, gr, n * 4);
        free(gr);
      }
      /* slice-owner flavour, two buckets */
      CK(fmpi_fused_step(&c2, g_off, FMPI_F32, p_off, FMPI_F32, dm, dv, NULL, 0, 64 * 100, &hy, cstep, &la, st));
      CK(fmpi_fused_step(&c2, g_off, FMPI_F32, p_off, FMPI_F32, dm, dv, NULL, 64 * 100, n, &hy, cstep, &la, st));
      const float bc1 = 1.f - powf(0.9f, (float)(step + 1)), bc2 = 1.f - powf(0.999f, (float)(step + 1));
      for (size_t i = 0; i < n; ++i) {
        m[i] = 0.9f * m[i] + 0.1f * g[i];
        v[i] = 0.999f * v[i] + 0.001f * g[i] * g[i];
        p[i] -= (m[i] / bc1) / (sqrtf(v[i] / bc2) + 1e-8f) * 1e-2f;
      }
      for (int r = 0; r < W; ++r) CK(fmpi_step_advance(dstep[r], st));
      CU(cudaStreamSynchronize(st));
      for (int r = 0; r < W; ++r) {
        d2h(got, (uint8_t*)c.data[r] + p_off, n * 4);
        EXPECT(close_enough(got, p, n, 2e-5f, "fused_step"), "fused_step step %d rank %d", step, r);
        d2h(got, (uint8_t*)c.data[r] + g_off, n * 4);
        int zero = 1; for (size_t i = 0; i < n; ++i) zero &= got[i] == 0.f;
        EXPECT(zero, "gradient bucket cleared, rank %d", r);
      }
    }
    /* one-shot flavour on a fresh problem (replicated state) */
    for (size_t i = 0; i < n; ++i) { p[i] = frand(&seed) - 0.75f; m[i] = v[i] = 0.f; g[i] = 0.f; }
    for (int r = 0; r < W; ++r) {
      h2d((uint8_t*)c.data[r] + p_off, p, n * 4);
      CU(cudaMemset(dm[r], 0, n * 4)); CU(cudaMemset(dv[r], 0, n * 4)); CU(cudaMemset(dstep[r], 0, 8));
      float* gr = malloc(n * 4);
      for (size_t i = 0; i < n; ++i) { gr[i] = frand(&seed) - 0.75f; g[i] += gr[i]; }
      h2d((uint8_t*)c.data[r] + g_off, gr, n * 4);
      free(gr);
    }
    CK(fmpi_fused_step_oneshot(&c2, g_off, p_off, FMPI_F32, dm, dv, NULL, 0, n, &hy, cstep, os_off, 2 * W * n * 4, &la, st));
    CU(cudaStreamSynchronize(st));
    for (size_t i = 0; i < n; ++i) {
      m[i] = 0.1f * g[i]; v[i] = 0.001f * g[i] * g[i];
      p[i] -= (m[i] / 0.1f) / (sqrtf(v[i] / 0.001f) + 1e-8f) * 1e-2f;
    }
    for (int r = 0; r < W; ++r) {
      d2h(got, (uint8_t*)c.data[r] + p_off, n * 4);
      EXPECT(close_enough(got, p, n, 3e-5f, "fused_step_oneshot"), "fused_step_oneshot rank %d", r);
    }
    for (int r = 0; r < W; ++r) { CU(cudaFree(dm[r])); CU(cudaFree(dv[r])); CU(cudaFree(dstep[r])); }
    free(p); free(m); free(v); free(g); free(got);
  }

  /* ---- tree kernels, pack / unpack, multi-tensor step, diagnostics */
  {
    enum { NL = 3 };
    const uint64_t numel[NL] = {7, 5000, 64 * 130};
    uint64_t offs[NL], total = 0;
    for (int t = 0; t < NL; ++t) { offs[t] = total; total += (numel[t] * 4 + 15) / 16 * 16; }
    float *dp_[W][NL], *dg[W][NL], *dmm[W][NL], *dvv[W][NL];
    float* hp0[NL]; float* hg[NL];
    const void* tabs[W];
    for (int t = 0; t < NL; ++t) {
      hp0[t] = malloc(numel[t] * 4); hg[t] = calloc(numel[t], 4);
      for (uint64_t i = 0; i < numel[t]; ++i) hp0[t][i] = frand(&seed);
    }
    for (int r = 0; r < W; ++r) {
      uint64_t row[6 * NL];
      for (int t = 0; t < NL; ++t) {
        dp_[r][t] = dalloc(numel[t] * 4); dg[r][t] = dalloc(numel[t] * 4);
        dmm[r][t] = dalloc(numel[t] * 4); dvv[r][t] = dalloc(numel[t] * 4);
        h2d(dp_[r][t], hp0[t], numel[t] * 4);
        float* gr = malloc(numel[t] * 4);
        for (uint64_t i = 0; i < numel[t]; ++i) { gr[i] = frand(&seed) - 0.75f; hg[t][i] += gr[i]; }
        h2d(dg[r][t], gr, numel[t] * 4);
        free(gr);
        row[t] = (uint64_t)(uintptr_t)dp_[r][t]; row[NL + t] = (uint64_t)(uintptr_t)dg[r][t];
        row[2 * NL + t] = (uint64_t)(uintptr_t)dmm[r][t]; row[3 * NL + t] = (uint64_t)(uintptr_t)dvv[r][t];
        row[4 * NL + t] = numel[t]; row[5 * NL + t] = offs[t];
      }
      void* dt = dalloc(sizeof row);
      h2d(dt, row, sizeof row);
      tabs[r] = dt;
    }
    fmpi_hyper hy; memset(&hy, 0, sizeof hy);
    hy.rule = FMPI_RULE_MOMENTUM; hy.eta = 0.05f; hy.rho = 0.9f; hy.grad_scale = 1.0f;
    CK(fmpi_tree_step_oneshot(&c, NL, tabs, total, FMPI_F32, &hy, 0, heap_off, 8 * MB, &la, st));
    CU(cudaStreamSynchronize(st));
    for (int r = 0; r < W; ++r)
      for (int t = 0; t < NL; ++t) {
        float* got = malloc(numel[t] * 4); float* want = malloc(numel[t] * 4);
        d2h(got, dp_[r][t], numel[t] * 4);
        for (uint64_t i = 0; i < numel[t]; ++i) want[i] = hp0[t][i] - 0.05f * hg[t][i];
        EXPECT(close_enough(got, want, numel[t], 1e-5f, "tree_step_oneshot"), "tree_step_oneshot rank %d leaf %d", r, t);
        free(got); free(want);
      }
    /* gather + all-reduce into the window, then unpack into the parameter leaves of rank 0 */
    CK(fmpi_tree_allreduce(&c, NL, tabs, heap_off + 8 * MB, total, FMPI_F32, FMPI_SUM, &la, st));
    CU(cudaStreamSynchronize(st));
    for (int t = 0; t < NL; ++t) {
      float* got = malloc(numel[t] * 4);
      d2h(got, (uint8_t*)c.data[1] + heap_off + 8 * MB + offs[t], numel[t] * 4);
      EXPECT(close_enough(got, hg[t], numel[t], 1e-5f, "tree_allreduce"), "tree_allreduce leaf %d", t);
      free(got);
    }
    {
      uint64_t tab[3 * NL];
      for (int t = 0; t < NL; ++t) { tab[t] = (uint64_t)(uintptr_t)dp_[0][t]; tab[NL + t] = offs[t]; tab[2 * NL + t] = numel[t] * 4; }
      uint64_t* dtab = dalloc(sizeof tab);
      h2d(dtab, tab, sizeof tab);
      void* flat = dalloc(total);
      CK(fmpi_pack(NL, (const void* const*)dtab, dtab + NL, dtab + 2 * NL, flat, st));
      for (int t = 0; t < NL; ++t) CU(cudaMemsetAsync(dp_[0][t], 0, numel[t] * 4, st));
      CK(fmpi_unpack(NL, (void* const*)dtab, dtab + NL, dtab + 2 * NL, flat, st));
      CU(cudaStreamSynchronize(st));
      float* a = malloc(numel[1] * 4); float* b = malloc(numel[1] * 4);
      d2h(a, dp_[0][1], numel[1] * 4); d2h(b, dp_[1][1], numel[1] * 4);
      EXPECT(memcmp(a, b, numel[1] * 4) == 0, "pack / unpack round trip");
      free(a); free(b);
      /* multi-tensor rule step on rank 2's leaves: Descent */
      uint64_t mt[5 * NL];
      for (int t = 0; t < NL; ++t) { mt[t] = (uint64_t)(uintptr_t)dp_[2][t]; mt[NL + t] = (uint64_t)(uintptr_t)dg[2][t]; mt[4 * NL + t] = numel[t]; }
      uint64_t* dmt = dalloc(sizeof mt);
      h2d(dmt, mt, sizeof mt);
      fmpi_hyper hd; memset(&hd, 0, sizeof hd); hd.rule = FMPI_RULE_DESCENT; hd.eta = 0.5f; hd.grad_scale = 1.0f;
      float* before = malloc(numel[2] * 4); float* gr = malloc(numel[2] * 4); float* after = malloc(numel[2] * 4);
      d2h(before, dp_[2][2], numel[2] * 4); d2h(gr, dg[2][2], numel[2] * 4);
      CK(fmpi_multi_tensor_step(NL, (void* const*)dmt, (const void* const*)(dmt + NL), NULL, NULL, dmt + 4 * NL, FMPI_F32, FMPI_F32, &hd, 0, st));
      CU(cudaStreamSynchronize(st));
      d2h(after, dp_[2][2], numel[2] * 4);
      for (uint64_t i = 0; i < numel[2]; ++i) before[i] -= 0.5f * gr[i];
      EXPECT(close_enough(after, before, numel[2], 1e-6f, "multi_tensor_step"), "multi_tensor_step");
      free(before); free(gr); free(after);
      /* diagnostics: both sum-of-squares entry points and the checksum */
      double* dout = dalloc(16); void* scratch = dalloc(fmpi_reduce_scratch_bytes());
      double s1 = 0, s2 = 0, want = 0;
      float* hx = malloc(numel[2] * 4);
      d2h(hx, dp_[2][2], numel[2] * 4);
      for (uint64_t i = 0; i < numel[2]; ++i) want += (double)hx[i] * hx[i];
      CK(fmpi_sumsq(dp_[2][2], numel[2], FMPI_F32, dout, scratch, st));
      CK(fmpi_sumsq_noalloc(dp_[2][2], numel[2], FMPI_F32, dout + 1, st));
      CU(cudaStreamSynchronize(st));
      d2h(&s1, dout, 8); d2h(&s2, dout + 1, 8);
      EXPECT(fabs(s1 - want) <= 1e-9 * want && fabs(s2 - want) <= 1e-9 * want, "sumsq %.12g / %.12g want %.12g", s1, s2, want);
      uint64_t ck1 = 0, ck2 = 0; uint64_t* dck = dalloc(16);
      CK(fmpi_checksum(dp_[0][1], numel[1] * 4, dck, scratch, st));
      CK(fmpi_checksum(dp_[1][1], numel[1] * 4, dck + 1, scratch, st));
      CU(cudaStreamSynchronize(st));
      d2h(&ck1, dck, 8); d2h(&ck2, dck + 1, 8);
      EXPECT(ck1 == ck2 && ck1 != 0, "checksum of equal buffers");
      free(hx);
    }
  }

  EXPECT(fmpi_health_check() == 0, "health page: %s", fmpi_cuda_last_error());
  for (int r = 0; r < W; ++r) CK(fmpi_symm_destroy(seg[r]));
  printf("ABI_CUDA_OK %d\n", g_checks);
  return 0;
}
