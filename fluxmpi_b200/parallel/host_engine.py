"""HostDataParallelEngine — the CPU counterpart of :class:`DataParallelEngine`.

Same contract (flat parameter / gradient vectors with the module's tensors as views, gradients
SUMMED over the workers unless ``average``, ``step()`` after ``backward()``, ``gather_state`` /
``load_state``), built from the host pieces: the flat gradient vector lives in symmetric host
memory (``fm.shm_empty``), so the all-reduce runs in place through the peers' mappings, and the
rule is one native loop over the flat vector (``csrc/host_rules.cpp``). There is nothing to
overlap on a CPU (the ranks share the cores), so there is one bucket and no hooks.

The reference's CPU path is the same per-leaf loop as its GPU path (src/optimizer.jl:20-23).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from .. import _native, comm, runtime
from .. import rules as R
from ..ops import fused as F
from ..ops import host_rules
from .flat import Axes, FlatParams

_ELEM_ALIGN = 16


class HostDataParallelEngine:
    def __init__(self, module: torch.nn.Module, rule: R.AbstractRule, *, average: Optional[bool] = None,
                 sync_params: bool = True, root_rank: int = 0):
        if not runtime.Initialized():
            raise runtime.FluxMPINotInitializedError()
        S = runtime.state()
        self.world, self.rank = S.world, S.rank
        self.module = module
        inner, avg = F.unwrap_rule(rule)
        self.rule = inner
        self.average = avg if average is None else average
        self.steps_done = 0

        params, seen = [], set()
        for p in module.parameters():
            if p.requires_grad and id(p) not in seen:
                seen.add(id(p))
                params.append(p)
        if not params:
            raise ValueError("module has no trainable parameters")
        if any(p.is_cuda for p in params):
            raise ValueError("HostDataParallelEngine is for CPU modules; use DataParallelEngine on a GPU")
        dts = {p.dtype for p in params}
        if len(dts) != 1 or next(iter(dts)) not in (torch.float32, torch.float64):
            raise ValueError(f"HostDataParallelEngine needs one parameter dtype, float32 or float64, got {dts}")
        self.dtype = params[0].dtype
        self.params = params
        self.offsets, cur = [], 0
        for p in params:
            self.offsets.append(cur)
            cur += (p.numel() + _ELEM_ALIGN - 1) // _ELEM_ALIGN * _ELEM_ALIGN
        self.numel = cur
        # both flat vectors in symmetric host memory (collective; plain memory at one worker):
        # the gradient all-reduce and the parameter broadcast then run in place
        self.P = comm.shm_empty((cur,), self.dtype)
        self.P.zero_()
        self.G = comm.shm_empty((cur,), self.dtype)
        self.G.zero_()
        with torch.no_grad():
            for p, o in zip(params, self.offsets):
                view = self.P[o:o + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view
                p.grad = self.G[o:o + p.numel()].view(p.shape)

        hh = host_rules._host_hyper(self.rule)
        self.fused = hh is not None
        self.m = self.v = None
        self.state_tree = None
        if self.fused:
            self.code, self.hyper = hh
            if self.code != _native.RULE_DESCENT:
                self.m = torch.zeros(cur, dtype=self.dtype)
            if self.code in (_native.RULE_ADAM, _native.RULE_ADAMW):
                self.v = torch.zeros(cur, dtype=self.dtype)
            self.hyper[6] = 1.0 / self.world if self.average else 1.0      # grad_scale
        if sync_params and self.world > 1:
            self.broadcast_params(root_rank)

    # ------------------------------------------------------------------ views
    def flat_params(self) -> FlatParams:
        ax = Axes()
        names = {id(p): n for n, p in self.module.named_parameters()}
        for p, o in zip(self.params, self.offsets):
            ax.entries[names.get(id(p), f"param{o}")] = (o, tuple(p.shape))
        return FlatParams(self.P, ax)

    def flat_grads(self) -> FlatParams:
        return self.flat_params().like(self.G)

    def broadcast_params(self, root: int = 0):
        comm.bcast_(self.P, root)
        bufs = [b for b in self.module.buffers() if b.numel() > 0]
        if bufs:
            from .sync import synchronize
            synchronize(bufs, root_rank=root)

    # ------------------------------------------------------------------ the step
    def step(self):
        """Sum the flat gradient over the workers in place, apply the rule, clear the gradient."""
        if self.world > 1:
            comm.allreduce_(self.G, "sum")
        if self.fused:
            ptr = lambda t: (C.c_void_p * 1)(t.data_ptr()) if t is not None else None
            n = (C.c_uint64 * 1)(self.numel)
            code = _native.F32 if self.dtype == torch.float32 else _native.F64
            _native.host_check(_native.host().fmpi_host_multi_tensor_step(
                1, ptr(self.P), ptr(self.G), ptr(self.m), ptr(self.v), n, code, self.code, self.hyper,
                self.steps_done), "host_multi_tensor_step")
        else:
            self.materialize_state()
            scale = 1.0 / self.world if self.average else 1.0
            with torch.no_grad():
                for leaf, p in zip(self.state_tree, self.params):
                    g = p.grad if scale == 1.0 else p.grad * scale
                    leaf.state, dx = self.rule.apply(leaf.state, p.data, g)
                    p.data.sub_(dx.to(p.dtype))
        self.G.zero_()
        self.steps_done += 1

    def zero_grad(self):
        self.G.zero_()

    def set_lr(self, eta: float):
        if self.fused:
            self.hyper[0] = float(eta)
        else:
            self.rule = R._adjust_rule(self.rule, eta, {})

    # ------------------------------------------------------------------ state access
    def materialize_state(self):
        """Per-leaf state of an unfused rule is created lazily; make it exist (same tree shape on
        every rank before a checkpoint or a synchronize walks it)."""
        if not self.fused and self.state_tree is None:
            self.state_tree = [R.Leaf(self.rule, self.rule.init(p.data)) for p in self.params]

    def gather_state(self) -> dict:
        out = {"step": self.steps_done}
        if not self.fused:
            self.materialize_state()
            out["leaves"] = [leaf.state for leaf in self.state_tree]
            return out
        for name in ("m", "v"):
            t = getattr(self, name)
            if t is not None:
                out[name] = t.clone()
        return out

    def load_state(self, state: dict):
        self.steps_done = int(state.get("step", 0))
        if not self.fused:
            leaves = state.get("leaves")
            self.state_tree = None if leaves is None else \
                [R.Leaf(self.rule, st.state if isinstance(st, R.Leaf) else st) for st in leaves]
            return
        for name in ("m", "v"):
            if name in state and getattr(self, name) is not None:
                getattr(self, name).copy_(state[name])

    def close(self):
        pass


def make_engine(module: torch.nn.Module, rule: R.AbstractRule, **kw):
    """:class:`DataParallelEngine` for a CUDA module, :class:`HostDataParallelEngine` for a CPU one."""
    if any(p.is_cuda for p in module.parameters()):
        from .engine import DataParallelEngine
        return DataParallelEngine(module, rule, **kw)
    host_kw = {k: v for k, v in kw.items() if k in ("average", "sync_params", "root_rank")}
    return HostDataParallelEngine(module, rule, **host_kw)
