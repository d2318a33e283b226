// Optimiser rules on host tensors: one call per parameter tree, one pass over memory per leaf.
//
// The reference's CPU path is Optimisers.jl's `apply!` per leaf — a chain of broadcast
// expressions, each a pass over the leaf (src/optimizer.jl:20-25 wraps it). The PyTorch
// restatement of that (fluxmpi_b200/rules.py) is ~13 out-of-place tensor ops per Adam leaf; on
// the Quick-Start MLP that is 3.8 ms of a 5.5 ms epoch. Here the same arithmetic, in the same
// order and precision, is a single loop per leaf that the compiler vectorises: read g, m, v, p
// once, write m, v, p once.
#define FMPI_HOST_ONLY
#include "../include/fmpi.h"
#include "host_internal.h"

#include <cmath>

using fmpi_host_detail::fail;

namespace {

// The loops are built for the baseline x86-64 ISA and, where the CPU has them, for AVX2+FMA
// and AVX-512 (picked at load time), without tying the library to the build machine.
#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__)
#define FMPI_CLONES __attribute__((target_clones("default", "avx2", "avx512f")))
#else
#define FMPI_CLONES
#endif

template <typename T>
struct Consts {
  T eta, b1, b2, omb1, omb2, eps, lambda, rho, scale;
  T c1, c2;         // 1 - beta1^(t+1), 1 - beta2^(t+1)
  T nrho2, nest;    // -rho^2, (1 + rho) * eta
};

struct HostHyper {
  int rule;
  double eta, beta1, beta2, eps, lambda, rho, grad_scale;
};

template <typename T>
Consts<T> make_consts(const HostHyper& h, int64_t step) {
  Consts<T> c;
  // Python evaluates the scalar expressions in double and rounds once when they meet the tensor
  const double b1 = h.beta1, b2 = h.beta2, rho = h.rho, eta = h.eta;
  c.eta = T(eta);
  c.b1 = T(b1);
  c.b2 = T(b2);
  c.omb1 = T(1.0 - b1);
  c.omb2 = T(1.0 - b2);
  c.eps = T(h.eps);
  c.lambda = T(h.lambda);
  c.rho = T(rho);
  c.scale = T(h.grad_scale);
  c.c1 = T(1.0 - std::pow(b1, double(step + 1)));
  c.c2 = T(1.0 - std::pow(b2, double(step + 1)));
  c.nrho2 = T(-(rho * rho));
  c.nest = T((1.0 + rho) * eta);
  return c;
}

template <typename T>
FMPI_CLONES void descent(T* __restrict__ p, const T* __restrict__ g, uint64_t n, Consts<T> c, bool scaled) {
  for (uint64_t i = 0; i < n; ++i) {
    const T gi = scaled ? g[i] * c.scale : g[i];
    p[i] -= gi * c.eta;
  }
}

template <typename T>
FMPI_CLONES void momentum(T* __restrict__ p, const T* __restrict__ g, T* __restrict__ m, uint64_t n,
                          Consts<T> c, bool scaled) {
  for (uint64_t i = 0; i < n; ++i) {
    const T gi = scaled ? g[i] * c.scale : g[i];
    const T v = m[i] * c.rho + gi * c.eta;
    m[i] = v;
    p[i] -= v;
  }
}

template <typename T>
FMPI_CLONES void nesterov(T* __restrict__ p, const T* __restrict__ g, T* __restrict__ m, uint64_t n,
                          Consts<T> c, bool scaled) {
  for (uint64_t i = 0; i < n; ++i) {
    const T gi = scaled ? g[i] * c.scale : g[i];
    const T d = m[i] * c.nrho2 + gi * c.nest;
    m[i] = m[i] * c.rho - gi * c.eta;
    p[i] -= d;
  }
}

template <typename T, bool kDecay>
FMPI_CLONES void adam(T* __restrict__ p, const T* __restrict__ g, T* __restrict__ m, T* __restrict__ v,
                      uint64_t n, Consts<T> c, bool scaled) {
  for (uint64_t i = 0; i < n; ++i) {
    const T gi = scaled ? g[i] * c.scale : g[i];
    const T mt = m[i] * c.b1 + gi * c.omb1;
    const T vt = v[i] * c.b2 + gi * gi * c.omb2;
    m[i] = mt;
    v[i] = vt;
    T d = mt / c.c1 / (std::sqrt(vt / c.c2) + c.eps) * c.eta;
    if (kDecay) d += p[i] * c.lambda;
    p[i] -= d;
  }
}

template <typename T>
int run(int n, void* const* p, const void* const* g, void* const* m, void* const* v,
        const uint64_t* numel, const HostHyper& h, int64_t step) {
  const Consts<T> c = make_consts<T>(h, step);
  const bool scaled = h.grad_scale != 1.0;
  const bool need_m = h.rule != FMPI_RULE_DESCENT;
  const bool need_v = h.rule == FMPI_RULE_ADAM || h.rule == FMPI_RULE_ADAMW;
  for (int k = 0; k < n; ++k) {
    if (!p[k] || !g[k]) return fail("leaf %d: null parameter or gradient", k);
    if ((need_m && (!m || !m[k])) || (need_v && (!v || !v[k])))
      return fail("leaf %d: rule %d needs moment arrays", k, h.rule);
    T* pk = static_cast<T*>(p[k]);
    const T* gk = static_cast<const T*>(g[k]);
    T* mk = need_m ? static_cast<T*>(m[k]) : nullptr;
    T* vk = need_v ? static_cast<T*>(v[k]) : nullptr;
    switch (h.rule) {
      case FMPI_RULE_DESCENT: descent<T>(pk, gk, numel[k], c, scaled); break;
      case FMPI_RULE_MOMENTUM: momentum<T>(pk, gk, mk, numel[k], c, scaled); break;
      case FMPI_RULE_NESTEROV: nesterov<T>(pk, gk, mk, numel[k], c, scaled); break;
      case FMPI_RULE_ADAM: adam<T, false>(pk, gk, mk, vk, numel[k], c, scaled); break;
      case FMPI_RULE_ADAMW: adam<T, true>(pk, gk, mk, vk, numel[k], c, scaled); break;
      default: return fail("rule %d has no fused host loop", h.rule);
    }
  }
  return 0;
}

}  // namespace

extern "C" int fmpi_host_multi_tensor_step(int n, void* const* p, const void* const* g,
                                           void* const* m, void* const* v, const uint64_t* numel,
                                           int dtype, int rule, const double* hyper,
                                           int64_t step) {
  if (n < 0 || !hyper || (n > 0 && (!p || !g || !numel))) return fail("fmpi_host_multi_tensor_step: bad arguments");
  if (step < 0) return fail("negative step count");
  const HostHyper h{rule, hyper[0], hyper[1], hyper[2], hyper[3], hyper[4], hyper[5], hyper[6]};
  if (dtype == FMPI_F32) return run<float>(n, p, g, m, v, numel, h, step);
  if (dtype == FMPI_F64) return run<double>(n, p, g, m, v, numel, h, step);
  return fail("host rule loops support f32 / f64, got dtype %d", dtype);
}
