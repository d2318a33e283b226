"""Optimiser rules with the Optimisers.jl contract the reference plugs into.

The reference is a *wrapper* around any ``Optimisers.AbstractRule``
(/root/reference/src/optimizer.jl:16-25) and reaches into ``Optimisers.Leaf.state``
(/root/reference/src/synchronize.jl:24-27). Optimisers.jl is not part of the reference tree
(SURVEY §2.5), so the contract is restated here for torch tensors:

* ``rule.init(x) -> state``
* ``rule.apply(state, x, dx) -> (state', dx')``; the caller then does ``x' = x - dx'``
* ``setup(rule, tree)`` mirrors the parameter tree with :class:`Leaf` nodes
* ``update(state_tree, params, grads) -> (state_tree', params')`` (non-mutating) and
  ``update_`` (mutating, ``Optimisers.update!``)

Adam keeps ``(mt, vt, (β1^t, β2^t))`` exactly like Optimisers (ϵ outside the square root, so
``torch.optim.Adam`` is a valid numerical oracle).
"""
from __future__ import annotations

from dataclasses import dataclass, replace
from typing import Any

import torch

from . import tree as T


class AbstractRule:
    """Base class of all rules (``Optimisers.AbstractRule``)."""

    def init(self, x: torch.Tensor):
        raise NotImplementedError

    def apply(self, state, x: torch.Tensor, dx: torch.Tensor):
        raise NotImplementedError


@dataclass
class Leaf:
    """Per-parameter optimiser node (``Optimisers.Leaf``): the rule and its state."""
    rule: AbstractRule
    state: Any
    frozen: bool = False

    def __eq__(self, other):
        return (isinstance(other, Leaf) and self.rule == other.rule
                and T.tree_equal(self.state, other.state))


def _zeros(x):
    return torch.zeros_like(x, memory_format=torch.preserve_format)


# ------------------------------------------------------------------------------ rules


@dataclass(frozen=True)
class Descent(AbstractRule):
    eta: float = 0.1

    def init(self, x):
        return None

    def apply(self, state, x, dx):
        return state, dx * self.eta


@dataclass(frozen=True)
class Momentum(AbstractRule):
    eta: float = 0.01
    rho: float = 0.9

    def init(self, x):
        return _zeros(x)

    def apply(self, state, x, dx):
        v = state * self.rho + dx * self.eta
        return v, v


@dataclass(frozen=True)
class Nesterov(AbstractRule):
    eta: float = 0.001
    rho: float = 0.9

    def init(self, x):
        return _zeros(x)

    def apply(self, state, x, dx):
        newdx = state * (-self.rho ** 2) + dx * ((1 + self.rho) * self.eta)
        v = state * self.rho - dx * self.eta
        return v, newdx


@dataclass(frozen=True)
class RMSProp(AbstractRule):
    eta: float = 0.001
    rho: float = 0.9
    eps: float = 1e-8

    def init(self, x):
        return _zeros(x)

    def apply(self, state, x, dx):
        acc = state * self.rho + dx * dx * (1 - self.rho)
        return acc, dx * (self.eta / (acc.sqrt() + self.eps))


@dataclass(frozen=True)
class Adam(AbstractRule):
    eta: float = 0.001
    beta: tuple = (0.9, 0.999)
    eps: float = 1e-8

    def init(self, x):
        return (_zeros(x), _zeros(x), tuple(self.beta))

    def apply(self, state, x, dx):
        mt, vt, bt = state
        b1, b2 = self.beta
        mt = mt * b1 + dx * (1 - b1)
        vt = vt * b2 + dx * dx * (1 - b2)
        dxn = mt / (1 - bt[0]) / ((vt / (1 - bt[1])).sqrt() + self.eps) * self.eta
        return (mt, vt, (bt[0] * b1, bt[1] * b2)), dxn


@dataclass(frozen=True)
class AdaMax(AbstractRule):
    eta: float = 0.001
    beta: tuple = (0.9, 0.999)
    eps: float = 1e-8

    def init(self, x):
        return (_zeros(x), _zeros(x), tuple(self.beta))

    def apply(self, state, x, dx):
        mt, ut, bt = state
        b1, b2 = self.beta
        mt = mt * b1 + dx * (1 - b1)
        ut = torch.maximum(ut * b2, dx.abs())
        dxn = mt / (ut + self.eps) * (self.eta / (1 - bt[0]))
        return (mt, ut, (bt[0] * b1, bt[1] * b2)), dxn


@dataclass(frozen=True)
class AdaGrad(AbstractRule):
    eta: float = 0.1
    eps: float = 1e-8

    def init(self, x):
        return torch.full_like(x, self.eps)

    def apply(self, state, x, dx):
        acc = state + dx * dx
        return acc, dx * (self.eta / (acc.sqrt() + self.eps))


@dataclass(frozen=True)
class AdaDelta(AbstractRule):
    rho: float = 0.9
    eps: float = 1e-8

    def init(self, x):
        return (_zeros(x), _zeros(x))

    def apply(self, state, x, dx):
        acc, dacc = state
        acc = acc * self.rho + dx * dx * (1 - self.rho)
        dxn = dx * ((dacc + self.eps).sqrt() / (acc + self.eps).sqrt())
        dacc = dacc * self.rho + dxn * dxn * (1 - self.rho)
        return (acc, dacc), dxn


@dataclass(frozen=True)
class AMSGrad(AbstractRule):
    eta: float = 0.001
    beta: tuple = (0.9, 0.999)
    eps: float = 1e-8

    def init(self, x):
        return (torch.full_like(x, self.eps), torch.full_like(x, self.eps),
                torch.full_like(x, self.eps))

    def apply(self, state, x, dx):
        mt, vt, vhat = state
        b1, b2 = self.beta
        mt = mt * b1 + dx * (1 - b1)
        vt = vt * b2 + dx * dx * (1 - b2)
        vhat = torch.maximum(vhat, vt)
        return (mt, vt, vhat), mt * self.eta / (vhat.sqrt() + self.eps)


@dataclass(frozen=True)
class NAdam(AbstractRule):
    eta: float = 0.001
    beta: tuple = (0.9, 0.999)
    eps: float = 1e-8

    def init(self, x):
        return (_zeros(x), _zeros(x), tuple(self.beta))

    def apply(self, state, x, dx):
        mt, vt, bt = state
        b1, b2 = self.beta
        mt = mt * b1 + dx * (1 - b1)
        vt = vt * b2 + dx * dx * (1 - b2)
        num = mt * (b1 / (1 - b1 * bt[0])) + dx * ((1 - b1) / (1 - bt[0]))
        dxn = num / ((vt * (b2 / (1 - bt[1]))).sqrt() + self.eps) * self.eta
        return (mt, vt, (bt[0] * b1, bt[1] * b2)), dxn


@dataclass(frozen=True)
class AdaBelief(AbstractRule):
    eta: float = 0.001
    beta: tuple = (0.9, 0.999)
    eps: float = 1e-16

    def init(self, x):
        return (_zeros(x), _zeros(x))

    def apply(self, state, x, dx):
        mt, st = state
        b1, b2 = self.beta
        mt = mt * b1 + dx * (1 - b1)
        st = st * b2 + (dx - mt) ** 2 * (1 - b2) + self.eps
        return (mt, st), mt * self.eta / (st.sqrt() + self.eps)


@dataclass(frozen=True)
class WeightDecay(AbstractRule):
    gamma: float = 5e-4

    def init(self, x):
        return None

    def apply(self, state, x, dx):
        return state, dx + x * self.gamma


@dataclass(frozen=True)
class ClipGrad(AbstractRule):
    delta: float = 10.0

    def init(self, x):
        return None

    def apply(self, state, x, dx):
        return state, dx.clamp(-self.delta, self.delta)


@dataclass(frozen=True)
class ClipNorm(AbstractRule):
    omega: float = 10.0
    p: float = 2.0
    throw: bool = True

    def init(self, x):
        return None

    def apply(self, state, x, dx):
        nrm = torch.linalg.vector_norm(dx.float(), self.p)
        if self.throw and not bool(torch.isfinite(nrm)):
            raise FloatingPointError(f"gradient has {self.p}-norm {float(nrm)}, for array of "
                                     f"size {tuple(dx.shape)}")
        scale = torch.clamp(self.omega / nrm, max=1.0).to(dx.dtype)
        return state, dx * scale


@dataclass(frozen=True)
class OptimiserChain(AbstractRule):
    """Compose rules; each sees the previous one's output as its gradient."""
    opts: tuple = ()

    def __init__(self, *opts):
        object.__setattr__(self, "opts", tuple(opts))

    def init(self, x):
        return tuple(o.init(x) for o in self.opts)

    def apply(self, state, x, dx):
        new = []
        for o, s in zip(self.opts, state):
            s, dx = o.apply(s, x, dx)
            new.append(s)
        return tuple(new), dx


def AdamW(eta: float = 0.001, beta=(0.9, 0.999), gamma: float = 0.0, eps: float = 1e-8):
    """``OptimiserChain(Adam(η, β, ϵ), WeightDecay(γ))`` as in Optimisers.jl: the decay term
    is added to Adam's step un-scaled by η (torch's AdamW corresponds to γ = lr·weight_decay)."""
    return OptimiserChain(Adam(eta, tuple(beta), eps), WeightDecay(gamma))


# ------------------------------------------------------------------------------ setup / update


def _trainable(x) -> bool:
    return isinstance(x, torch.Tensor) and (x.is_floating_point() or x.is_complex())


def setup(rule: AbstractRule, params):
    """Mirror ``params`` with :class:`Leaf` nodes (``Optimisers.setup``). Tied tensors share one
    Leaf; non-trainable leaves map to ``None``."""
    if not isinstance(rule, AbstractRule):
        raise TypeError("setup expects an AbstractRule")
    cache: dict[int, Leaf] = {}
    keep = []

    def mk(x):
        if not _trainable(x):
            return None
        if id(x) not in cache:
            keep.append(x)
            cache[id(x)] = Leaf(rule, rule.init(_data(x)))
        return cache[id(x)]

    return T.fmap(mk, params, _cache={})


def _data(x: torch.Tensor) -> torch.Tensor:
    return x.detach() if x.requires_grad else x


def _is_leaf_node(n) -> bool:
    return isinstance(n, Leaf) or n is None


def _walk_update(st, ps, gs, inplace: bool, seen: dict):
    """Returns (st', ps')."""
    if isinstance(st, Leaf):
        if gs is None or st.frozen:
            return st, ps
        key = id(ps)
        if key in seen:
            return seen[key]
        x = _data(ps)
        g = gs.detach() if isinstance(gs, torch.Tensor) else gs
        apply = getattr(st.rule, "apply_", None) if inplace else None
        if apply is not None:
            new_state, dx = apply(st.state, x, g)
        else:
            new_state, dx = st.rule.apply(st.state, x, g)
        if inplace:
            with torch.no_grad():
                x.sub_(dx.to(x.dtype))
            st.state = new_state
            out = (st, ps)
        else:
            out = (Leaf(st.rule, new_state, st.frozen), x - dx.to(x.dtype))
        seen[key] = out
        return out
    if st is None or not T.is_container(st):
        return st, ps
    cs = [c for _, c in T.children(st)]
    cp = [c for _, c in T.children(ps)]
    cg = [c for _, c in T.children(gs)] if gs is not None and T.is_container(gs) else [None] * len(cs)
    if not (len(cs) == len(cp) == len(cg)):
        raise ValueError("update: optimiser-state, parameter and gradient trees differ in shape")
    res = [_walk_update(a, b, c, inplace, seen) for a, b, c in zip(cs, cp, cg)]
    return T.rebuild(st, [r[0] for r in res]), T.rebuild(ps, [r[1] for r in res])


def update(state_tree, params, grads):
    """Non-mutating step: ``(state', params') = update(state, params, grads)``
    (``Optimisers.update``, used at /root/reference/README.md:66).

    A rule may take over the whole-tree step by defining ``tree_update`` — the
    DistributedOptimizer uses that to reduce every gradient in one fused collective instead of
    one per leaf."""
    hook = getattr(_first_rule(state_tree), "tree_update", None)
    if hook is not None:
        return hook(state_tree, params, grads, False)
    return walk_update(state_tree, params, grads, False)


def update_(state_tree, params, grads):
    """Mutating step (``Optimisers.update!``): parameters and state tensors change in place."""
    hook = getattr(_first_rule(state_tree), "tree_update", None)
    if hook is not None:
        return hook(state_tree, params, grads, True)
    return walk_update(state_tree, params, grads, True)


def walk_update(state_tree, params, grads, inplace: bool):
    """The per-leaf walk (for rules that implement ``tree_update``). An in-place step over a
    tree of CUDA leaves with a fusable rule is ONE pointer-table kernel launch
    (ops/fused.py, csrc/kernels/multi_tensor.cu) instead of 2-4 elementwise kernels per leaf."""
    if inplace and _any_cuda(params):
        from .ops import fused
        if fused.try_tree_step_(state_tree, params, grads):
            return state_tree, params
    elif inplace:
        # CPU leaves: the same idea, one native call and one pass over memory per leaf
        from .ops import host_rules
        if host_rules.try_tree_step_host_(state_tree, params, grads):
            return state_tree, params
    elif not _any_cuda(params) and _first_rule(state_tree) is not None:
        # non-mutating step on CPU leaves: copy parameters and state once, then the native loop
        # on the copies (3 copies + 1 pass instead of ~13 out-of-place passes per Adam leaf)
        from .ops import host_rules
        if host_rules._host_hyper(_first_rule(state_tree)) is not None:
            new_st, new_ps = _copy_for_step(state_tree, params)
            if host_rules.try_tree_step_host_(new_st, new_ps, grads):
                return new_st, new_ps
    return _walk_update(state_tree, params, grads, inplace, {})


def _copy_for_step(state_tree, params):
    """Fresh Leafs with cloned state tensors and detached clones of the trainable parameters;
    ties are kept (a shared Leaf / tensor is copied once), everything else is passed through."""
    leaf_cache: dict[int, Leaf] = {}

    def copy_leaf(n):
        if not isinstance(n, Leaf):
            return n
        if id(n) not in leaf_cache:
            state = T.fmap(lambda t: t.clone() if isinstance(t, torch.Tensor) else t, n.state)
            leaf_cache[id(n)] = Leaf(n.rule, state, n.frozen)
        return leaf_cache[id(n)]

    new_st = T.fmap(copy_leaf, state_tree, is_leaf=_is_leaf_node)
    new_ps = T.fmap(lambda x: _data(x).clone() if _trainable(x) else x, params)
    return new_st, new_ps


def _any_cuda(params) -> bool:
    for x in T.leaves(params):
        if isinstance(x, torch.Tensor):
            return x.is_cuda
    return False


def _first_rule(state_tree):
    for n in T.leaves(state_tree, is_leaf=_is_leaf_node):
        if isinstance(n, Leaf):
            return n.rule
    return None


def adjust(state_tree, eta: float | None = None, **kw):
    """Change hyper-parameters of every Leaf's rule (``Optimisers.adjust``)."""

    def f(n):
        if not isinstance(n, Leaf):
            return n
        return Leaf(_adjust_rule(n.rule, eta, kw), n.state, n.frozen)

    return T.fmap(f, state_tree, is_leaf=_is_leaf_node)


def _adjust_rule(rule, eta, kw):
    if hasattr(rule, "adjusted"):
        return rule.adjusted(eta, kw)
    if isinstance(rule, OptimiserChain):
        return OptimiserChain(*[_adjust_rule(o, eta, kw) for o in rule.opts])
    changes = dict(kw)
    if eta is not None and hasattr(rule, "eta"):
        changes["eta"] = eta
    changes = {k: v for k, v in changes.items() if hasattr(rule, k)}
    return replace(rule, **changes) if changes else rule
