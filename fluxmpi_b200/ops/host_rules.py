"""In-place optimiser step for a whole tree of CPU leaves in one native call
(``csrc/host_rules.cpp``) — the host twin of :func:`ops.fused.try_tree_step_`.

The per-leaf path restates Optimisers.jl's ``apply!`` with out-of-place torch ops (about 13
per Adam leaf, each a pass over the leaf plus an allocation); on the reference's Quick-Start MLP
(README.md:31-70) that is 3.8 ms of a 5.5 ms epoch. The native loop reads g, m, v, p once and
writes m, v, p once, with the same operation order and precision.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native
from .. import rules as R
from .. import tree as T
from . import fused

_CODES = {torch.float32: _native.F32, torch.float64: _native.F64}


def _host_hyper(rule):
    """(rule code, [eta, beta1, beta2, eps, lambda, rho, grad_scale]) in Python floats — the
    device block stores fp32, which would cost fp64 tensors half their digits."""
    r, _ = fused.unwrap_rule(rule)
    eta, b1, b2, eps, lam, rho = 0.0, 0.9, 0.999, 1e-8, 0.0, 0.0
    if type(r) is R.Descent:
        code, eta = _native.RULE_DESCENT, r.eta
    elif type(r) is R.Momentum:
        code, eta, rho = _native.RULE_MOMENTUM, r.eta, r.rho
    elif type(r) is R.Nesterov:
        code, eta, rho = _native.RULE_NESTEROV, r.eta, r.rho
    elif type(r) is R.Adam:
        code, eta, eps, (b1, b2) = _native.RULE_ADAM, r.eta, r.eps, r.beta
    elif (type(r) is R.OptimiserChain and len(r.opts) == 2 and type(r.opts[0]) is R.Adam
          and type(r.opts[1]) is R.WeightDecay):
        a, wd = r.opts
        code, eta, eps, (b1, b2), lam = _native.RULE_ADAMW, a.eta, a.eps, a.beta, wd.gamma
    else:
        return None
    return code, (C.c_double * 7)(eta, b1, b2, eps, lam, rho, 1.0)


def try_tree_step_host_(state_tree, params, grads) -> bool:
    """Returns False — having changed nothing — when any leaf does not qualify (mixed rules or
    dtypes, a rule without a native loop, non-contiguous storage, dtypes other than f32 / f64);
    the caller then walks the tree leaf by leaf."""
    triples, seen = [], set()

    def walk(st, p, g) -> bool:
        if isinstance(st, R.Leaf):
            if g is None or st.frozen or id(p) in seen:
                return True
            seen.add(id(p))
            triples.append((st, p, g))
            return True
        if st is None or not T.is_container(st):
            return True
        cs = [c for _, c in T.children(st)]
        cp = [c for _, c in T.children(p)]
        cg = [c for _, c in T.children(g)] if g is not None and T.is_container(g) else [None] * len(cs)
        if not (len(cs) == len(cp) == len(cg)):
            return False
        return all(walk(a, b, c) for a, b, c in zip(cs, cp, cg))

    if not walk(state_tree, params, grads) or not triples:
        return False
    rule = triples[0][0].rule
    hh = _host_hyper(rule)
    if hh is None:
        return False
    code, hyper = hh
    first = triples[0][1]
    if not isinstance(first, torch.Tensor) or first.dtype not in _CODES:
        return False
    dt = first.dtype
    need_m = code != _native.RULE_DESCENT
    need_v = code in (_native.RULE_ADAM, _native.RULE_ADAMW)
    n = len(triples)
    P, G = (C.c_void_p * n)(), (C.c_void_p * n)()
    M = (C.c_void_p * n)() if need_m else None
    V = (C.c_void_p * n)() if need_v else None
    N = (C.c_uint64 * n)()
    rebuilds, step, keep = [], None, []
    for i, (leaf, p, g) in enumerate(triples):
        if leaf.rule != rule or not (isinstance(p, torch.Tensor) and isinstance(g, torch.Tensor)):
            return False
        pd, gd = (p.detach() if p.requires_grad else p), (g.detach() if g.requires_grad else g)
        if pd.is_cuda or gd.is_cuda or pd.dtype != dt or gd.dtype != dt or pd.shape != gd.shape \
                or not pd.is_contiguous() or not gd.is_contiguous() or pd.numel() == 0:
            return False
        got = fused._rule_state_tensors(rule, leaf.state)
        if got is None:
            return False
        m, v, s, rebuild = got
        for t, need in ((m, need_m), (v, need_v)):
            if need and not (isinstance(t, torch.Tensor) and not t.is_cuda and t.dtype == dt
                             and t.is_contiguous() and t.shape == pd.shape):
                return False
        if step is None:
            step = s
        elif s != step:
            return False
        P[i], G[i], N[i] = pd.data_ptr(), gd.data_ptr(), pd.numel()
        if need_m:
            M[i] = m.data_ptr()
        if need_v:
            V[i] = v.data_ptr()
        keep.append((pd, gd, m, v))
        rebuilds.append(rebuild)
    _native.host_check(
        _native.host().fmpi_host_multi_tensor_step(n, P, G, M, V, N, _CODES[dt], code, hyper, step or 0),
        "host_multi_tensor_step")
    for (leaf, _, _), rebuild in zip(triples, rebuilds):
        leaf.state = rebuild()
    return True
