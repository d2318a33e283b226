"""Structural map over nested containers — the role Functors.jl's ``fmap`` plays for the
reference (/root/reference/src/synchronize.jl:12, src/optimizer.jl:47,58,62,
ext/FluxMPIFluxExt.jl:7).

Containers: dict (any Mapping subclass, order kept), list, tuple, namedtuple, dataclass
instances marked with ``@functor``, and :class:`fluxmpi_b200.rules.Leaf` is handled by its
callers. Everything else is a leaf. Like Functors, results are cached by object identity so a
tied (shared) tensor is visited once and stays tied in the output.
"""
from __future__ import annotations

import dataclasses
from collections.abc import Mapping
from typing import Any, Callable, Iterator

import numpy as np
import torch

_FUNCTOR_TYPES: set[type] = set()


def functor(cls):
    """Class decorator: treat instances of this dataclass as containers for :func:`fmap`."""
    if not dataclasses.is_dataclass(cls):
        raise TypeError("@functor expects a dataclass")
    _FUNCTOR_TYPES.add(cls)
    return cls


def is_namedtuple(x) -> bool:
    return isinstance(x, tuple) and hasattr(x, "_fields") and hasattr(x, "_make")


def is_container(x) -> bool:
    return isinstance(x, (Mapping, list, tuple)) or type(x) in _FUNCTOR_TYPES


def is_array(x) -> bool:
    return isinstance(x, (torch.Tensor, np.ndarray))


def is_numeric_array(x) -> bool:
    if isinstance(x, torch.Tensor):
        return True
    return isinstance(x, np.ndarray) and x.dtype.kind in "biufc"


def children(x) -> Iterator[tuple[Any, Any]]:
    if isinstance(x, Mapping):
        yield from x.items()
    elif is_namedtuple(x):
        yield from zip(x._fields, x)
    elif isinstance(x, (list, tuple)):
        yield from enumerate(x)
    elif type(x) in _FUNCTOR_TYPES:
        for f in dataclasses.fields(x):
            yield f.name, getattr(x, f.name)


def rebuild(x, new_children: list):
    if isinstance(x, Mapping):
        try:
            out = type(x)()
            for k, v in zip(x.keys(), new_children):
                out[k] = v
            return out
        except TypeError:
            return dict(zip(x.keys(), new_children))
    if is_namedtuple(x):
        return type(x)._make(new_children)
    if isinstance(x, tuple):
        return tuple(new_children)
    if isinstance(x, list):
        return list(new_children)
    names = [f.name for f in dataclasses.fields(x)]
    return dataclasses.replace(x, **dict(zip(names, new_children)))


def fmap(f: Callable[[Any], Any], x, *, is_leaf: Callable[[Any], bool] | None = None,
         _cache: dict | None = None):
    """Apply ``f`` to every leaf of ``x`` and rebuild the same structure."""
    cache = {} if _cache is None else _cache

    def rec(node):
        leaf = (is_leaf is not None and is_leaf(node)) or not is_container(node)
        if leaf:
            cacheable = is_array(node)
            if cacheable and id(node) in cache:
                return cache[id(node)][1]
            out = f(node)
            if cacheable:
                cache[id(node)] = (node, out)  # keep node alive so ids stay unique
            return out
        return rebuild(node, [rec(c) for _, c in children(node)])

    return rec(x)


def fmap2(f: Callable[[Any, Any], Any], x, y, *, is_leaf=None):
    """Map over two trees of the same structure (``y`` may hold ``None`` for a whole subtree)."""

    def rec(a, b):
        leaf = (is_leaf is not None and is_leaf(a)) or not is_container(a)
        if leaf or b is None:
            return f(a, b)
        ca = [c for _, c in children(a)]
        cb = [c for _, c in children(b)]
        if len(ca) != len(cb):
            raise ValueError(f"tree structure mismatch: {type(a).__name__} with {len(ca)} vs "
                             f"{len(cb)} children")
        return rebuild(a, [rec(p, q) for p, q in zip(ca, cb)])

    return rec(x, y)


def leaves(x, *, is_leaf=None, unique: bool = True) -> list:
    """Leaves in traversal order; identical array objects are listed once when ``unique``."""
    out, seen = [], set()

    def rec(node):
        if ((is_leaf is not None and is_leaf(node)) or not is_container(node)):
            if unique and is_array(node):
                if id(node) in seen:
                    return
                seen.add(id(node))
            out.append(node)
            return
        for _, c in children(node):
            rec(c)

    rec(x)
    return out


def leaves_with_path(x, prefix: tuple = ()) -> list[tuple[tuple, Any]]:
    out = []

    def rec(node, path):
        if not is_container(node):
            out.append((path, node))
            return
        for k, c in children(node):
            rec(c, path + (k,))

    rec(x, prefix)
    return out


def tree_equal(a, b) -> bool:
    if is_container(a):
        if not is_container(b):
            return False
        ca, cb = list(children(a)), list(children(b))
        if len(ca) != len(cb):
            return False
        return all(ka == kb and tree_equal(va, vb) for (ka, va), (kb, vb) in zip(ca, cb))
    if isinstance(a, torch.Tensor):
        return isinstance(b, torch.Tensor) and a.shape == b.shape and bool(torch.equal(a, b))
    if isinstance(a, np.ndarray):
        return isinstance(b, np.ndarray) and a.shape == b.shape and bool(np.array_equal(a, b))
    return a == b
