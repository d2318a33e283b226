"""FlatParams — a flat vector with named, nested views (the ComponentArrays.jl analogue).

Reference: ``synchronize!(::ComponentArray)`` broadcasts the flat backing vector in one message
and rebuilds the array from its axes (/root/reference/ext/FluxMPIComponentArraysExt.jl:6-9);
test/test_synchronize.jl:56-66 checks that nested property access still works afterwards. A
ComponentArray gradient going through ``DistributedOptimizer`` is one allreduce for the whole
model — the shape the fused kernel wants, so this type is also what the
high-performance engine uses internally (one symmetric buffer, parameters are views).
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Iterator

import torch

from .. import tree as T

_ALIGN_BYTES = 16


class Axes:
    """Nested name → (offset, shape) | Axes description of a flat vector."""

    def __init__(self):
        self.entries: "OrderedDict[Any, tuple[int, tuple] | Axes]" = OrderedDict()
        self.kind: type = dict
        self.nt_type = None

    def __eq__(self, other):
        return isinstance(other, Axes) and list(self.entries.items()) == list(other.entries.items())

    def __repr__(self):
        return f"Axes({dict(self.entries)!r})"


class FlatParams:
    """``FlatParams.from_tree(tree)`` copies every tensor leaf into one contiguous 1-D tensor;
    ``fp.a.b`` / ``fp["a"]["b"]`` return views into it. ``getdata``/``getaxes`` mirror
    ComponentArrays."""

    __slots__ = ("data", "axes")

    def __init__(self, data: torch.Tensor, axes: Axes):
        object.__setattr__(self, "data", data)
        object.__setattr__(self, "axes", axes)

    # -- construction -----------------------------------------------------------------
    @staticmethod
    def from_tree(tree, dtype: torch.dtype | None = None, device=None, align_elems: int = 1,
                  out: torch.Tensor | None = None) -> "FlatParams":
        tensors = [t for t in T.leaves(tree, unique=False) if isinstance(t, torch.Tensor)]
        if not tensors:
            raise ValueError("FlatParams.from_tree: the tree holds no tensors")
        dtype = dtype or tensors[0].dtype
        device = device if device is not None else tensors[0].device
        cursor = [0]

        def build(node) -> Axes | tuple:
            if isinstance(node, torch.Tensor):
                off = (cursor[0] + align_elems - 1) // align_elems * align_elems
                cursor[0] = off + node.numel()
                return (off, tuple(node.shape))
            if T.is_container(node):
                ax = Axes()
                ax.kind = list if isinstance(node, list) else tuple if isinstance(node, tuple) else dict
                if T.is_namedtuple(node):
                    ax.nt_type = type(node)
                for k, c in T.children(node):
                    sub = build(c)
                    if sub is not None:
                        ax.entries[k] = sub
                return ax
            return None

        axes = build(tree)
        if not isinstance(axes, Axes):
            ax = Axes()
            ax.entries["value"] = axes
            axes = ax
            tree = {"value": tree}
        total = cursor[0]
        if out is None:
            data = torch.zeros(total, dtype=dtype, device=device)
        else:
            if out.numel() < total or out.dim() != 1:
                raise ValueError("FlatParams.from_tree: `out` is too small or not 1-D")
            data = out[:total]
        fp = FlatParams(data, axes)

        def fill(node, ax):
            if isinstance(ax, tuple):
                off, shape = ax
                n = 1
                for s in shape:
                    n *= s
                with torch.no_grad():
                    data[off:off + n].view(shape).copy_(node)
                return
            for k, c in T.children(node):
                if k in ax.entries:
                    fill(c, ax.entries[k])

        fill(tree, axes)
        return fp

    # -- ComponentArrays vocabulary -----------------------------------------------------
    def getdata(self) -> torch.Tensor:
        return self.data

    def getaxes(self) -> Axes:
        return self.axes

    def like(self, data: torch.Tensor) -> "FlatParams":
        """Same axes over another flat vector (``ComponentArray(d, getaxes(x))``)."""
        if data.numel() != self.data.numel():
            raise ValueError("like: flat vector has a different length")
        return FlatParams(data, self.axes)

    # -- access -------------------------------------------------------------------------
    def _resolve(self, entry):
        if isinstance(entry, Axes):
            return FlatParams(self.data, entry)
        off, shape = entry
        n = 1
        for s in shape:
            n *= s
        return self.data[off:off + n].view(shape)

    def __getitem__(self, key):
        try:
            return self._resolve(self.axes.entries[key])
        except KeyError:
            raise KeyError(key) from None

    def __getattr__(self, name):
        try:
            entries = object.__getattribute__(self, "axes").entries
        except AttributeError:
            raise AttributeError(name) from None
        if name in entries:
            return self._resolve(entries[name])
        raise AttributeError(name)

    def __setattr__(self, name, value):
        raise AttributeError("FlatParams is immutable; write into the views instead")

    def keys(self):
        return self.axes.entries.keys()

    def __iter__(self) -> Iterator:
        return iter(self.axes.entries.keys())

    def __len__(self) -> int:
        return self.data.numel()

    def to_tree(self):
        """Nested container of views with the original structure."""

        def rec(ax: Axes):
            vals = [rec(e) if isinstance(e, Axes) else self._resolve(e) for e in ax.entries.values()]
            if ax.nt_type is not None:
                return ax.nt_type._make(vals)
            if ax.kind is list:
                return list(vals)
            if ax.kind is tuple:
                return tuple(vals)
            return dict(zip(ax.entries.keys(), vals))

        return rec(self.axes)

    def named_views(self, prefix: str = "") -> Iterator[tuple[str, torch.Tensor]]:
        def rec(ax: Axes, pre: str):
            for k, e in ax.entries.items():
                name = f"{pre}.{k}" if pre else str(k)
                if isinstance(e, Axes):
                    yield from rec(e, name)
                else:
                    yield name, self._resolve(e)

        yield from rec(self.axes, prefix)

    def __repr__(self):
        return f"FlatParams(n={self.data.numel()}, dtype={self.data.dtype}, axes={self.axes!r})"
