"""``synchronize`` — make every rank's copy of a tree equal to ``root_rank``'s.

Reference: /root/reference/src/synchronize.jl:10-35 (+ ext/FluxMPIComponentArraysExt.jl:6-9,
ext/FluxMPIFluxExt.jl:6-8). Dispatch table reproduced:

=====================================  =====================================================
reference method                       here
=====================================  =====================================================
NamedTuple / Tuple (empty → itself)    dict / list / tuple / namedtuple (empty → itself)
AbstractArray{<:Number} → ``bcast!``   torch.Tensor / numeric ndarray → broadcast in place
AbstractArray (non-numbers)            object ndarray → elementwise
Optimisers.Leaf → sync ``.state``      rules.Leaf → sync ``.state``
Number → bcast 1-element array         int / float / bool / complex / numpy scalar
anything else → identity               None, str, … returned untouched (even if it differs)
ComponentArray → one bcast of data     FlatParams → one bcast of the flat vector
FluxMPIFluxModel → bare model          FluxMPIFluxModel(nn.Module) → bare module
=====================================  =====================================================

What changes is the cost model: the reference issues one blocking ``MPI_Bcast`` per leaf and
one per scalar. Here all array leaves that share a device are packed by a
pointer-table kernel into one staging buffer and travel in **one** broadcast; all scalars
travel in one more small host message.
"""
from __future__ import annotations

import numbers
from typing import Any

import numpy as np
import torch

from .. import comm, runtime
from .. import tree as T
from ..rules import Leaf
from .flat import FlatParams


class FluxMPIFluxModel:
    """Marker wrapper so ``synchronize`` can dispatch on an arbitrary model object
    (/root/reference/src/FluxMPI.jl:84-86). ``synchronize(FluxMPIFluxModel(m))`` broadcasts the
    module's parameters and buffers and returns the **bare** module
    (ext/FluxMPIFluxExt.jl:7)."""

    def __init__(self, model):
        self.model = model


_ALIGN = 16


def _is_scalar(x) -> bool:
    return isinstance(x, (numbers.Number, np.generic)) and not isinstance(x, (torch.Tensor,))


class _Plan:
    """Collects leaves, then synchronises them in as few collectives as possible."""

    def __init__(self, root: int):
        self.root = root
        self.tensors: list[torch.Tensor] = []
        self.seen: set[int] = set()
        self.scalars: list[Any] = []

    def add_tensor(self, t: torch.Tensor, key: int | None = None):
        key = id(t) if key is None else key
        if key not in self.seen and t.numel() > 0:
            self.seen.add(key)
            self.tensors.append(t)

    def add_scalar(self, x) -> int:
        self.scalars.append(x)
        return len(self.scalars) - 1

    def run(self) -> list:
        groups: dict[torch.device, list[torch.Tensor]] = {}
        for t in self.tensors:
            groups.setdefault(t.device, []).append(t)
        for dev, ts in groups.items():
            _bcast_group(ts, self.root)
        return _bcast_scalars(self.scalars, self.root)


def _bcast_group(ts: list[torch.Tensor], root: int):
    if runtime.total_workers() == 1:
        return
    if len(ts) == 1:
        comm.bcast_(_plain(ts[0]), root)
        return
    if ts[0].is_cuda:
        dw = runtime.state().device_world
        if dw is not None and dw.native:
            dw.bcast_many([_plain(t) for t in ts], root)
            return
    # generic path: pack into one byte buffer, one broadcast, unpack
    offs, total = [], 0
    for t in ts:
        offs.append(total)
        total += (t.numel() * t.element_size() + _ALIGN - 1) // _ALIGN * _ALIGN
    flat = torch.empty(total, dtype=torch.uint8, device=ts[0].device)
    if runtime.local_rank() == root:
        for t, o in zip(ts, offs):
            n = t.numel() * t.element_size()
            flat[o:o + n].copy_(_bytes_view(t))
    comm.bcast_(flat, root)
    if runtime.local_rank() != root:
        for t, o in zip(ts, offs):
            n = t.numel() * t.element_size()
            src = flat[o:o + n].view(t.dtype if t.dtype != torch.bool else torch.uint8)
            with torch.no_grad():
                _plain(t).copy_(src.view(t.shape) if t.dtype != torch.bool
                                else src.view(t.shape).bool())


def _plain(t: torch.Tensor) -> torch.Tensor:
    return t.detach() if t.requires_grad else t


def _bytes_view(t: torch.Tensor) -> torch.Tensor:
    c = _plain(t).contiguous()
    if c.dtype == torch.bool:
        c = c.to(torch.uint8)
    return c.view(-1).view(torch.uint8)


def _bcast_scalars(scalars: list, root: int) -> list:
    """All scalars in one message: each is carried as (float64 real, float64 imag, int64)."""
    if not scalars:
        return []
    if runtime.total_workers() == 1:
        return list(scalars)
    n = len(scalars)
    f = torch.zeros(2 * n, dtype=torch.float64)
    i = torch.zeros(n, dtype=torch.int64)
    for k, x in enumerate(scalars):
        if isinstance(x, (bool, np.bool_)) or isinstance(x, (numbers.Integral, np.integer)):
            i[k] = int(x)
        elif isinstance(x, (numbers.Real, np.floating)):
            f[2 * k] = float(x)
        else:
            z = complex(x)
            f[2 * k], f[2 * k + 1] = z.real, z.imag
    buf = torch.cat([f.view(torch.uint8), i.view(torch.uint8)])
    comm.bcast_(buf, root)
    f = buf[: 16 * n].view(torch.float64)
    i = buf[16 * n:].view(torch.int64)
    out = []
    for k, x in enumerate(scalars):
        if isinstance(x, (bool, np.bool_)):
            out.append(type(x)(bool(i[k].item())))
        elif isinstance(x, (numbers.Integral, np.integer)):
            out.append(type(x)(i[k].item()))
        elif isinstance(x, (numbers.Real, np.floating)):
            out.append(type(x)(f[2 * k].item()))
        else:
            out.append(type(x)(complex(f[2 * k].item(), f[2 * k + 1].item())))
    return out


class _ScalarSlot:
    __slots__ = ("idx",)

    def __init__(self, idx):
        self.idx = idx


def _collect(x, plan: _Plan):
    """First pass: register leaves, return a skeleton with scalar placeholders."""
    if isinstance(x, FluxMPIFluxModel):
        return _collect_module(x.model, plan)
    if isinstance(x, torch.nn.Module):
        return _collect_module(x, plan)
    if isinstance(x, FlatParams):
        plan.add_tensor(x.data)
        return x
    if isinstance(x, Leaf):
        return Leaf(x.rule, _collect(x.state, plan), x.frozen)
    if isinstance(x, torch.Tensor):
        plan.add_tensor(x)
        return x
    if isinstance(x, np.ndarray):
        if x.dtype.kind in "biufc":
            if x.size:
                plan.add_tensor(torch.from_numpy(x), key=id(x))
            return x
        out = np.empty_like(x)
        for idx in np.ndindex(x.shape):
            out[idx] = _collect(x[idx], plan)
        return out
    if T.is_container(x):
        kids = list(T.children(x))
        if not kids:
            return x
        return T.rebuild(x, [_collect(c, plan) for _, c in kids])
    if _is_scalar(x):
        return _ScalarSlot(plan.add_scalar(x))
    return x


def _collect_module(m: torch.nn.Module, plan: _Plan):
    for p in m.parameters():
        plan.add_tensor(p)
    for b in m.buffers():
        plan.add_tensor(b)
    return m


def _fill(x, scalars: list):
    if isinstance(x, _ScalarSlot):
        return scalars[x.idx]
    if isinstance(x, Leaf):
        return Leaf(x.rule, _fill(x.state, scalars), x.frozen)
    if isinstance(x, np.ndarray) and x.dtype == object:
        out = np.empty_like(x)
        for idx in np.ndindex(x.shape):
            out[idx] = _fill(x[idx], scalars)
        return out
    if T.is_container(x):
        kids = list(T.children(x))
        if not kids:
            return x
        return T.rebuild(x, [_fill(c, scalars) for _, c in kids])
    return x


def synchronize(x, root_rank: int = 0):
    """Synchronise ``x`` across all processes from ``root_rank`` and return the synced tree.

    Tensors are overwritten in place (so existing references stay valid) *and* the tree is
    returned; scalars are immutable, so callers must use the return value — the same contract
    as the reference (``ps = FluxMPI.synchronize!(ps; root_rank=0)``, README.md:45)."""
    if not runtime.Initialized():
        raise runtime.FluxMPINotInitializedError()
    root = int(root_rank)
    if not (0 <= root < runtime.total_workers()):
        raise ValueError(f"root_rank {root} out of range")
    plan = _Plan(root)
    skel = _collect(x, plan)
    scalars = plan.run()
    return _fill(skel, scalars) if plan.scalars else skel


synchronize_ = synchronize
