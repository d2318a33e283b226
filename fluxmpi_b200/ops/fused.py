"""Python faces of the non-collective kernels in ``libfmpi_cuda.so``.

* :func:`multi_tensor_step` — one launch that applies an optimiser rule to a whole list of
  scattered (param, grad, m, v) leaves (csrc/kernels/multi_tensor.cu). The reference leaves
  this to Optimisers.jl, which issues 2-4 broadcast kernels per leaf.
* :func:`sumsq` / :func:`checksum` — gradient norm and bitwise replica checks
  (csrc/kernels/reduce_diag.cu).
* :func:`flush_l2` — bench hygiene.

Everything raises if the CUDA library is absent; there is no eager fallback on a GPU box.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from .. import _native
from .. import rules as R
from .._native import Hyper, cuda_check


def _stream_ptr(stream=None) -> int:
    return (torch.cuda.current_stream() if stream is None else stream).cuda_stream


def _dtype_code(dt):
    from ..comm import dtype_code
    return dtype_code(dt)


def unwrap_rule(rule):
    """Strip a DistributedOptimizer wrapper; returns (inner_rule, average)."""
    from ..parallel.optimizer import DistributedOptimizer
    if isinstance(rule, DistributedOptimizer):
        return rule.optimizer, rule.average
    return rule, False


def hyper_for(rule, grad_scale: float = 1.0, zero_grad: bool = False, eta_dev: int = 0) -> Optional[Hyper]:
    """Map an Optimisers-style rule onto the fused kernels' hyper-parameter block, or None when
    the rule has no fused implementation (the caller then reduces with a kernel and applies
    the rule with torch ops)."""
    rule, _ = unwrap_rule(rule)
    h = Hyper()
    h.grad_scale, h.zero_grad, h.eta_dev = grad_scale, int(zero_grad), eta_dev
    h.beta1, h.beta2, h.eps, h.lam, h.rho = 0.9, 0.999, 1e-8, 0.0, 0.0
    if type(rule) is R.Descent:
        h.rule, h.eta = _native.RULE_DESCENT, rule.eta
    elif type(rule) is R.Momentum:
        h.rule, h.eta, h.rho = _native.RULE_MOMENTUM, rule.eta, rule.rho
    elif type(rule) is R.Nesterov:
        h.rule, h.eta, h.rho = _native.RULE_NESTEROV, rule.eta, rule.rho
    elif type(rule) is R.Adam:
        h.rule, h.eta, h.eps = _native.RULE_ADAM, rule.eta, rule.eps
        h.beta1, h.beta2 = rule.beta
    elif (type(rule) is R.OptimiserChain and len(rule.opts) == 2 and type(rule.opts[0]) is R.Adam
          and type(rule.opts[1]) is R.WeightDecay):
        a, wd = rule.opts
        h.rule, h.eta, h.eps, h.lam = _native.RULE_ADAMW, a.eta, a.eps, wd.gamma
        h.beta1, h.beta2 = a.beta
    else:
        return None
    return h


def rule_needs(h: Hyper) -> tuple[bool, bool]:
    """(needs first moment / velocity, needs second moment)"""
    return h.rule != _native.RULE_DESCENT, h.rule in (_native.RULE_ADAM, _native.RULE_ADAMW)


class TensorTable:
    """Device-resident pointer table for a fixed list of leaves (built once, reused per step)."""

    def __init__(self, params: Sequence[torch.Tensor], grads: Sequence[torch.Tensor],
                 m: Optional[Sequence[torch.Tensor]], v: Optional[Sequence[torch.Tensor]]):
        n = len(params)
        dev = params[0].device
        zeros = [0] * n
        rows = [[t.data_ptr() for t in params], [t.data_ptr() for t in grads],
                [t.data_ptr() for t in m] if m is not None else zeros,
                [t.data_ptr() for t in v] if v is not None else zeros,
                [t.numel() for t in params]]
        flat = [x for row in rows for x in row]
        self.n = n
        self.has_m, self.has_v = m is not None, v is not None
        self.dev = torch.tensor(flat, dtype=torch.int64).pin_memory().to(dev, non_blocking=True)
        self.keep = (list(params), list(grads), list(m or []), list(v or []))
        self.key = tuple(rows[0]) + tuple(rows[1])

    def row(self, i: int) -> int:
        return self.dev.data_ptr() + 8 * self.n * i


def multi_tensor_step(table: TensorTable, h: Hyper, step: int, dtype: torch.dtype, stream=None):
    lib = _native.cuda()
    code = _dtype_code(dtype)
    cuda_check(lib.fmpi_multi_tensor_step(
        table.n, table.row(0), table.row(1), table.row(2) if table.has_m else None,
        table.row(3) if table.has_v else None, table.row(4), code, code, C.byref(h), step,
        _stream_ptr(stream)), "multi_tensor_step")
    _native.count_launch()


_scratch: dict = {}


def _scratch_for(device) -> torch.Tensor:
    key = (device.type, device.index, torch.cuda.current_stream(device).cuda_stream)
    s = _scratch.get(key)
    if s is None:
        s = torch.zeros(_native.cuda().fmpi_reduce_scratch_bytes(), dtype=torch.uint8, device=device)
        _scratch[key] = s
    return s


def sumsq(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Sum of squares of ``x`` in float64 (a 0-d device tensor; no host sync)."""
    lib = _native.cuda()
    x = x.detach().contiguous()
    out = torch.empty((), dtype=torch.float64, device=x.device) if out is None else out
    cuda_check(lib.fmpi_sumsq(x.data_ptr(), x.numel(), _dtype_code(x.dtype), out.data_ptr(),
                              _scratch_for(x.device).data_ptr(), _stream_ptr()), "sumsq")
    _native.count_launch()
    return out


def checksum(x: torch.Tensor) -> torch.Tensor:
    """64-bit position-weighted checksum of the bytes of ``x`` (0-d int64 device tensor)."""
    lib = _native.cuda()
    x = x.detach().contiguous()
    out = torch.empty((), dtype=torch.int64, device=x.device)
    cuda_check(lib.fmpi_checksum(x.data_ptr(), x.numel() * x.element_size(), out.data_ptr(),
                                 _scratch_for(x.device).data_ptr(), _stream_ptr()), "checksum")
    _native.count_launch()
    return out


_flush_buf: dict = {}


def flush_l2(device=None, nbytes: int = 256 << 20):
    """Overwrite a buffer larger than the 126 MB L2 so the next timed iteration starts cold."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    buf = _flush_buf.get(device)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        _flush_buf[device] = buf
    cuda_check(_native.cuda().fmpi_flush_l2(buf.data_ptr(), nbytes, _stream_ptr()), "flush_l2")


# ------------------------------------------------------------------------------ functional trees
_table_cache: dict = {}


def _rule_state_tensors(rule, state):
    """(m, v, step, rebuild) for a fusable rule's per-leaf state, or None if the layout is not
    the one ``rule.init`` produces. ``rebuild(step+1)`` returns the state to store back."""
    import math
    r, _ = unwrap_rule(rule)
    if type(r) is R.Descent:
        return None, None, 0, (lambda: state)
    if type(r) in (R.Momentum, R.Nesterov):
        if not isinstance(state, torch.Tensor):
            return None
        return state, None, 0, (lambda: state)
    if type(r) is R.Adam:
        adam, st, wrap = r, state, (lambda s: s)
    elif type(r) is R.OptimiserChain and len(r.opts) == 2 and type(r.opts[0]) is R.Adam:
        adam, st, wrap = r.opts[0], state[0], (lambda s: (s, state[1]))
    else:
        return None
    if not (isinstance(st, tuple) and len(st) == 3 and isinstance(st[0], torch.Tensor)):
        return None
    mt, vt, bt = st
    b1 = adam.beta[0]
    step = int(round(math.log(bt[0]) / math.log(b1))) - 1 if 0 < b1 < 1 else 0
    return mt, vt, max(step, 0), (lambda: wrap((mt, vt, (bt[0] * adam.beta[0], bt[1] * adam.beta[1]))))


def _collect_tree(state_tree, params, grads):
    """The leaves of a whole-tree step as flat lists, or None when any leaf does not qualify for
    the pointer-table kernels (mixed rules or dtypes, CPU tensors, non-contiguous storage, fp64,
    a rule without a fused kernel). Ties are visited once."""
    from .. import tree as T
    triples, seen = [], set()

    def walk(st, p, g) -> bool:
        if isinstance(st, R.Leaf):
            if g is None or st.frozen:
                return True
            if id(p) in seen:
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
        return None
    rule = triples[0][0].rule
    h = hyper_for(rule)
    if h is None:
        return None
    dt = triples[0][1].dtype
    if dt not in (torch.float32, torch.bfloat16, torch.float16):
        return None
    ms, vs, ps, gs, rebuilds, step = [], [], [], [], [], None
    for leaf, p, g in triples:
        if leaf.rule != rule or not (isinstance(p, torch.Tensor) and isinstance(g, torch.Tensor)):
            return None
        pd, gd = (p.detach() if p.requires_grad else p), (g.detach() if g.requires_grad else g)
        if not (pd.is_cuda and gd.is_cuda and pd.dtype == dt and gd.dtype == dt and pd.is_contiguous()
                and gd.is_contiguous() and pd.shape == gd.shape and pd.numel() > 0):
            return None
        got = _rule_state_tensors(rule, leaf.state)
        if got is None:
            return None
        m, v, s, rebuild = got
        for t in (m, v):
            if t is not None and not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()
                                      and t.shape == pd.shape):
                return None
        if step is None:
            step = s
        elif s != step:
            return None
        ps.append(pd); gs.append(gd); ms.append(m); vs.append(v); rebuilds.append(rebuild)
    return rule, h, dt, ps, gs, ms, vs, rebuilds, step or 0, triples


def try_tree_step_(state_tree, params, grads) -> bool:
    """One-kernel in-place optimiser step for a whole tree of CUDA leaves
    (``Optimisers.update!`` semantics): returns False — having changed nothing — when any leaf
    does not qualify; the caller then walks the tree leaf by leaf."""
    got = _collect_tree(state_tree, params, grads)
    if got is None:
        return False
    rule, h, dt, ps, gs, ms, vs, rebuilds, step, triples = got
    need_m, need_v = rule_needs(h)
    key = tuple(t.data_ptr() for t in ps + gs) + tuple(t.data_ptr() for t in ms + vs if t is not None)
    table = _table_cache.get(key)
    if table is None:
        if len(_table_cache) > 16:
            _table_cache.clear()
        table = TensorTable(ps, gs, ms if need_m else None, vs if need_v else None)
        _table_cache[key] = table
    # (averaging, if asked for, has already been applied to the gradients by the caller)
    multi_tensor_step(table, h, step, dt)
    for (leaf, _, _), rebuild in zip(triples, rebuilds):
        leaf.state = rebuild()
    return True


# ------------------------------------------------------------------------------ distributed trees
class TreeTable:
    """Device table [p | g | m | v | numel | flat offset] for the tree kernels
    (csrc/kernels/tree_step.cu); flat offsets are 16-byte aligned."""

    def __init__(self, ps, gs, ms, vs):
        n = len(ps)
        es = ps[0].element_size()
        offs, total = [], 0
        for t in ps:
            offs.append(total)
            total += (t.numel() * es + 15) // 16 * 16
        zeros = [0] * n
        rows = [[t.data_ptr() for t in ps], [t.data_ptr() for t in gs],
                [t.data_ptr() for t in ms] if ms is not None else zeros,
                [t.data_ptr() for t in vs] if vs is not None else zeros,
                [t.numel() for t in ps], offs]
        self.n, self.total, self.offs = n, total, offs
        self.dev = torch.tensor([x for row in rows for x in row], dtype=torch.int64).pin_memory().to(
            ps[0].device, non_blocking=True)
        self.keep = (list(ps), list(gs), list(ms or []), list(vs or []))
        self.ptr = C.c_void_p(self.dev.data_ptr())
        self.ref = C.byref(self.ptr)
        self.arena_table = None          # TensorTable whose gradients point into the pack arena


_tree_cache: dict = {}


def try_tree_step_distributed_(state_tree, params, grads, average: bool) -> bool:
    """``update!`` through a DistributedOptimizer over CUDA leaves on a native multi-GPU world in
    at most TWO launches: small trees in ONE (push + barrier + rank-ordered sum + rule applied to
    the scattered leaves; the reduced gradient never exists in memory), larger trees as
    gather-and-all-reduce into the symmetric arena + the pointer-table rule kernel reading it.
    Returns False, having done nothing, when the tree or the world does not qualify."""
    from .. import runtime
    S = runtime.state()
    dw = S.device_world
    if dw is None or not dw.native or S.world < 2:
        return False
    got = _collect_tree(state_tree, params, grads)
    if got is None:
        return False
    rule, h, dt, ps, gs, ms, vs, rebuilds, step, triples = got
    need_m, need_v = rule_needs(h)
    key = tuple(t.data_ptr() for t in ps + gs) + tuple(t.data_ptr() for t in ms + vs if t is not None)
    tab = _tree_cache.get(key)
    if tab is None:
        if len(_tree_cache) > 16:
            _tree_cache.clear()
        tab = TreeTable(ps, gs, ms if need_m else None, vs if need_v else None)
        _tree_cache[key] = tab
    h.grad_scale = 1.0 / S.world if average else 1.0
    lib = dw._lib
    dw.check_health()
    st = _stream_ptr()
    code = _dtype_code(dt)
    tr_off, tr_bytes = dw.tree_region
    slot = (tr_bytes // (2 * S.world)) & ~15
    if tab.total <= min(slot, int(TREE_ONESHOT_MAX)):
        cuda_check(lib.fmpi_tree_step_oneshot(dw._comm_refs[0], tab.n, tab.ref, tab.total, code, C.byref(h), step,
                                              tr_off, tr_bytes, dw._lp_auto, st), "tree_step_oneshot")
        _native.count_launch()
    else:
        arena, refs, off = dw.pack_arena(tab.total)
        cuda_check(lib.fmpi_tree_allreduce(refs[0], tab.n, tab.ref, off, tab.total, code, _native.SUM,
                                           dw._lp_auto, st), "tree_allreduce")
        _native.count_launch()
        if tab.arena_table is None or tab.arena_table.arena_ptr != arena.data_ptr():
            es = ps[0].element_size()
            views = [arena[o:o + p.numel() * es].view(dt).view(p.shape) for o, p in zip(tab.offs, ps)]
            tab.arena_table = TensorTable(ps, views, ms if need_m else None, vs if need_v else None)
            tab.arena_table.arena_ptr = arena.data_ptr()
        multi_tensor_step(tab.arena_table, h, step, dt)
    for (leaf, _, _), rebuild in zip(triples, rebuilds):
        leaf.state = rebuild()
    return True


TREE_ONESHOT_MAX = 4 << 20     # bytes; above it the two-launch path moves fewer bytes per link
