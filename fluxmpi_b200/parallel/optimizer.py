"""DistributedOptimizer and allreduce_gradients — the tree-level DP algorithms.

Reference (citations into /root/reference/src/optimizer.jl):

* ``DistributedOptimizer(rule)`` :16-25 — an ``AbstractRule`` whose ``apply!`` all-reduces the
  gradient with ``+`` and then delegates to the wrapped rule; ``init`` delegates, so the state
  tree is identical to the unwrapped rule's (test/test_optimizer.jl:13-14). **Sum, not mean**
  (:11-14; README.md:127-128): scale the loss by ``1/total_workers()`` for an average.
* ``allreduce_gradients(gs; on_gpu)`` :45-65 — all leaves in flight at once, then wait.

What is new (SURVEY §2.1 "capability gaps"): the whole gradient tree is reduced in ONE fused
collective (pack → allreduce → unpack by pointer-table kernels on the GPU; one shm pass on the
CPU) instead of one MPI message per leaf, there is no host staging for device tensors, and
``allreduce_gradients`` also accepts a flat vector (:class:`FlatParams` / plain tensor), which
the reference only supports through ``DistributedOptimizer``.
"""
from __future__ import annotations

import threading
from dataclasses import dataclass

import numpy as np
import torch

from .. import comm, runtime
from .. import rules as R
from .. import tree as T
from ..utils import trace
from .flat import FlatParams

_tls = threading.local()


def _already_reduced() -> bool:
    return getattr(_tls, "reduced", False)


class _Reduced:
    def __enter__(self):
        self.prev = getattr(_tls, "reduced", False)
        _tls.reduced = True

    def __exit__(self, *a):
        _tls.reduced = self.prev


@dataclass(frozen=True)
class DistributedOptimizer(R.AbstractRule):
    """Wrap ``optimizer`` so that gradients are summed over all workers before the step."""
    optimizer: R.AbstractRule
    average: bool = False  # SUM stays the default; opt-in mean

    def init(self, x):
        return self.optimizer.init(x)

    def apply(self, state, x, dx):
        if not _already_reduced():
            dx = comm.allreduce_(dx.clone() if dx.requires_grad else dx, "sum")
            if self.average:
                dx = dx / runtime.total_workers()
        return self.optimizer.apply(state, x, dx)

    # whole-tree step: one fused gradient reduction, then the inner rule per leaf
    def tree_update(self, state_tree, params, grads, inplace: bool):
        # CUDA leaves + a rule with a fused kernel on a native multi-GPU world: at most two launches
        # for the whole step (ops/fused.py try_tree_step_distributed_); `update` runs it on copies
        if R._any_cuda(params):
            from ..ops import fused
            if inplace:
                if fused.try_tree_step_distributed_(state_tree, params, grads, self.average):
                    return state_tree, params
            elif fused.hyper_for(self.optimizer) is not None and runtime.total_workers() > 1:
                new_st, new_ps = R._copy_for_step(state_tree, params)
                if fused.try_tree_step_distributed_(new_st, new_ps, grads, self.average):
                    return new_st, new_ps
        grads = allreduce_gradients(grads)
        if self.average:
            w = runtime.total_workers()
            grads = T.fmap(lambda g: g / w if isinstance(g, torch.Tensor) else g, grads)
        with _Reduced():
            return R.walk_update(state_tree, params, grads, inplace)

    def adjusted(self, eta, kw):
        return DistributedOptimizer(R._adjust_rule(self.optimizer, eta, kw), self.average)


def _grad_leaves(gs) -> list[torch.Tensor]:
    out = []
    for leaf in T.leaves(gs):
        if isinstance(leaf, torch.Tensor):
            out.append(leaf)
        elif isinstance(leaf, np.ndarray) and leaf.dtype.kind in "biufc":
            out.append(torch.from_numpy(leaf))
    return out


def allreduce_gradients(gs, on_gpu: bool | None = None):
    """Sum the gradients over all workers; returns the tree (leaves are reduced in place).

    ``gs`` may be any nested container of tensors (dict / tuple / namedtuple / list), a
    :class:`FlatParams`, or a single tensor. Non-array leaves pass through untouched
    (src/optimizer.jl:51). ``on_gpu`` is accepted for API compatibility and ignored: device
    tensors never travel through the host here.
    """
    if not runtime.Initialized():
        raise runtime.FluxMPINotInitializedError()
    if isinstance(gs, FlatParams):
        comm.allreduce_(gs.data, "sum")
        return gs
    if isinstance(gs, torch.Tensor):
        return comm.allreduce_(gs, "sum")
    if runtime.total_workers() == 1:
        return gs
    leaves = [g for g in _grad_leaves(gs) if g.numel() > 0]
    if not leaves:
        return gs
    groups: dict[tuple, list[torch.Tensor]] = {}
    for g in leaves:
        groups.setdefault((g.device, g.dtype), []).append(g)
    reqs = []
    for (dev, dt), ts in groups.items():
        trace.count("allreduce_tree", "cuda" if dev.type == "cuda" else "host",
                    sum(t.numel() * t.element_size() for t in ts))
        if dev.type == "cuda":
            dw = runtime.state().device_world
            if dw is not None and dw.native:
                dw.allreduce_many(ts, comm.op_code("sum"))
                continue
        if dev.type == "cpu" and all(t.is_contiguous() for t in ts):
            # the leaves are copied straight into / out of the shared-memory slots: one message,
            # no flattening copy (csrc/host_backend.cpp fmpi_host_allreduce_many)
            runtime.host_run(comm._host_allreduce_many_now,
                             [t.detach() if t.requires_grad else t for t in ts], comm.op_code("sum"))
            continue
        if len(ts) == 1:
            _, req = comm.Iallreduce(ts[0], "sum")
            reqs.append(req)
            continue
        # one message for the whole group: flatten, reduce, scatter back
        flat = torch.cat([t.detach().reshape(-1) for t in ts])
        _, req = comm.Iallreduce(flat, "sum")
        reqs.append(_Unflatten(req, flat, ts))
    comm.Waitall(reqs)
    return gs


class _Unflatten:
    def __init__(self, req, flat, ts):
        self.req, self.flat, self.ts = req, flat, ts

    def wait(self, sync: bool = False):
        self.req.wait(sync=sync)
        off = 0
        with torch.no_grad():
            for t in self.ts:
                n = t.numel()
                t.detach().copy_(self.flat[off:off + n].view(t.shape))
                off += n
