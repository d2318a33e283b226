"""Checkpoint / resume for data-parallel runs.

The reference has no checkpointing; its guide only says to gate side effects on
``local_rank() == 0`` (docs/src/guide.md:19) and ``synchronize!`` from ``root_rank`` is the
building block a user would pair with load-on-root. This module is that pairing:

* ``save(path, tree)`` — rank ``root`` writes (atomically), everyone else waits at a barrier;
* ``load(path, like=tree)`` — rank ``root`` reads, then the tree is synchronised to all ranks,
  so a job can resume with a different number of workers;
* ``save_engine`` / ``load_engine`` — parameters + *full replicated* optimiser state of a
  :class:`DataParallelEngine` (the owner-sharded moments are gathered first).
"""
from __future__ import annotations

import os
import pickle
from pathlib import Path

import torch

from .. import runtime
from .. import tree as T
from ..parallel.sync import synchronize


def _to_cpu(tree):
    return T.fmap(lambda x: x.detach().cpu() if isinstance(x, torch.Tensor) else x, tree)


def save(path, tree, root: int = 0) -> None:
    """Write ``tree`` (any nested container of tensors / scalars) from rank ``root``."""
    if runtime.local_rank() == root:
        p = Path(path)
        p.parent.mkdir(parents=True, exist_ok=True)
        tmp = p.with_suffix(p.suffix + f".tmp{os.getpid()}")
        torch.save(_to_cpu(tree), tmp)
        os.replace(tmp, p)
    runtime.host_barrier()


def load(path, like, root: int = 0):
    """Read on ``root``, broadcast into the structure of ``like`` (same shapes) on every rank.
    Tensors of ``like`` are overwritten in place and the synchronised tree is returned."""
    if runtime.local_rank() == root:
        # tensors, numbers and plain containers only: loading a checkpoint never runs code from it
        data = torch.load(path, map_location="cpu", weights_only=True)

        def put(dst, src):
            if isinstance(dst, torch.Tensor):
                with torch.no_grad():
                    (dst.detach() if dst.requires_grad else dst).copy_(src)
                return dst
            return src

        like = T.fmap2(put, like, data)
    return synchronize(like, root_rank=root)


def _bcast_blob(blob: bytes | None, root: int) -> bytes:
    """Root's byte string on every rank (length first, then the payload), outside the structural
    synchronise: the ranks need not know its size, or even whether there is one, beforehand."""
    from .. import comm
    n = torch.tensor([len(blob) if blob is not None else 0], dtype=torch.int64)
    comm.bcast_(n, root)
    buf = torch.empty(int(n), dtype=torch.uint8)
    if runtime.local_rank() == root and int(n):
        buf.copy_(torch.frombuffer(bytearray(blob), dtype=torch.uint8))
    if int(n):
        comm.bcast_(buf, root)
    return bytes(buf.numpy().tobytes())


def save_engine(path, engine, extra=None, root: int = 0) -> None:
    """Parameters, buffers, the full replicated optimiser state and ``extra`` (any picklable
    object: epoch counters, RNG state, a config dict) of a ``DataParallelEngine``."""
    engine.materialize_state()
    state = engine.gather_state()                      # collective: every rank takes part
    blob = torch.frombuffer(bytearray(pickle.dumps(extra)), dtype=torch.uint8).clone()
    payload = {"params": engine.P, "optimizer": state, "steps_done": engine.steps_done,
               "extra_pickle": blob,
               "buffers": {k: v for k, v in engine.module.named_buffers()}}
    save(path, payload, root)


def load_engine(path, engine, root: int = 0):
    """Restore parameters, buffers and optimiser state into ``engine`` on every rank; returns the
    ``extra`` object that was saved (it is unpickled: load checkpoints you trust)."""
    engine.materialize_state()                         # same tree shape on every rank before the walk
    like = {"params": engine.P, "optimizer": engine.gather_state(), "steps_done": 0,
            "buffers": {k: v for k, v in engine.module.named_buffers()}}
    blob = None
    if runtime.local_rank() == root:
        data = torch.load(path, map_location="cpu", weights_only=True)
        blob = bytes(data.pop("extra_pickle").numpy().tobytes()) if "extra_pickle" in data else None
        if "extra" in data:                            # files written before the blob existed
            blob = pickle.dumps(data.pop("extra"))

        def put(dst, src):
            if isinstance(dst, torch.Tensor):
                with torch.no_grad():
                    (dst.detach() if dst.requires_grad else dst).copy_(src)
                return dst
            return src

        like = T.fmap2(put, like, data)
    got = synchronize(like, root_rank=root)
    raw = _bcast_blob(blob, root)
    engine.load_state(got["optimizer"])
    if getattr(engine, "master", None) is not None:
        engine.master.copy_(got["optimizer"].get("master", engine.P.float()))
    engine.steps_done = int(got["steps_done"])
    return pickle.loads(raw) if raw else None
