"""Data sharding and the host→device input pipeline.

Reference: ``DistributedDataContainer(data)`` (/root/reference/src/data.jl:8-26) keeps the
contiguous block ``partition(1:N, ceil(N/W))[rank+1]`` of an MLUtils-style container
(``length`` / ``getindex``) — test/test_data.jl:15-26. Here the same block layout over any
object with ``__len__`` / ``__getitem__`` (lists, tensors, torch Datasets). One deliberate
difference: when ceil-chunking yields fewer than W blocks (N=9, W=4) the reference
throws a BoundsError; trailing ranks get an empty shard instead.

``DeviceLoader`` is the B200 half of the north star's item (4): batches are assembled in
pinned host buffers and copied to the GPU on a dedicated copy stream, double-buffered so the
copy of batch k+1 overlaps the compute of batch k.
"""
from __future__ import annotations

import math
from typing import Iterator, Sequence

import torch

from . import runtime


class DistributedDataContainer:
    """Shard ``data`` across workers in contiguous rank-ordered blocks."""

    def __init__(self, data, rank: int | None = None, world: int | None = None):
        self.data = data
        total = len(data)
        world = runtime.total_workers() if world is None else world
        rank = runtime.local_rank() if rank is None else rank
        per = int(math.ceil(total / world)) if total else 0
        lo = min(rank * per, total)
        hi = min(lo + per, total)
        self.idxs = range(lo, hi)

    def __len__(self) -> int:
        return len(self.idxs)

    def __getitem__(self, i):
        """``ddc[i]`` = ``data[idxs[i]]`` (src/data.jl:26). Slices and index lists are forwarded
        to the container in one call when it supports batch indexing (tensors, TensorDataset),
        so a mini-batch is one gather instead of ``batch`` Python-level lookups."""
        if isinstance(i, slice):
            r = self.idxs[i]
            try:
                return self.data[slice(r.start, r.stop, r.step)] if len(r) else self.data[0:0]
            except (TypeError, IndexError, KeyError):
                return [self.data[j] for j in r]
        if isinstance(i, torch.Tensor) and i.dim() == 0:
            return self.data[self.idxs[int(i)]]
        if isinstance(i, (list, tuple, torch.Tensor)):
            mapped = [self.idxs[int(j)] for j in i]
            try:
                return self.data[mapped]
            except (TypeError, IndexError, KeyError):
                return [self.data[j] for j in mapped]
        return self.data[self.idxs[i]]

    def __iter__(self):
        for j in self.idxs:
            yield self.data[j]


class TensorDataset:
    """Minimal ``(x[i], y[i], …)`` container over equally long tensors (MLUtils' tuple-of-arrays
    convention), batch-indexable so loaders can gather a whole batch with one index op."""

    def __init__(self, *tensors: torch.Tensor):
        n = tensors[0].shape[0]
        if any(t.shape[0] != n for t in tensors):
            raise ValueError("all tensors must share the first dimension")
        self.tensors = tensors

    def __len__(self):
        return self.tensors[0].shape[0]

    def __getitem__(self, i):
        return tuple(t[i] for t in self.tensors)


def _collate(samples: Sequence):
    first = samples[0]
    if isinstance(first, torch.Tensor):
        return torch.stack(list(samples))
    if isinstance(first, (tuple, list)):
        return tuple(_collate([s[k] for s in samples]) for k in range(len(first)))
    return torch.as_tensor(samples)


class DeviceLoader:
    """Iterate mini-batches of a (sharded) container as device tensors.

    Each batch is collated into one of ``depth`` pinned host buffers, then copied host→device
    with ``non_blocking=True`` on a private copy stream; the consumer's stream waits on the
    copy's event. ``h2d_bytes`` counts the bytes moved (bench.py reports it).
    """

    def __init__(self, data, batch_size: int, device=None, shuffle: bool = False,
                 drop_last: bool = True, depth: int = 2, seed: int = 0):
        self.data = data
        self.batch_size = int(batch_size)
        self.device = torch.device(device) if device is not None else runtime.current_device()
        self.shuffle = shuffle
        self.drop_last = drop_last
        self.depth = max(2, depth)
        self.seed = seed
        self.epoch = 0
        self.h2d_bytes = 0
        self._cuda = self.device.type == "cuda"
        self._stream = torch.cuda.Stream(self.device) if self._cuda else None
        self._pinned: list = [None] * self.depth
        self._dev: list = [None] * self.depth
        self._events: list = [None] * self.depth
        self._consumed: list = [None] * self.depth

    def __len__(self):
        n = len(self.data)
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def _indices(self) -> list[int]:
        n = len(self.data)
        if self.shuffle:
            g = torch.Generator().manual_seed(self.seed + self.epoch)
            return torch.randperm(n, generator=g).tolist()
        return list(range(n))

    def _gather(self, idx: list[int]):
        # a contiguous run (no shuffling) becomes a slice: tensor-backed containers then hand
        # back views and the only host copy left is the one into the pinned slot
        key = slice(idx[0], idx[-1] + 1) if idx and idx[-1] - idx[0] + 1 == len(idx) and \
            all(b - a == 1 for a, b in zip(idx, idx[1:])) else idx
        try:
            batch = self.data[key]       # batch-indexable containers gather in one op
            if isinstance(batch, list):
                batch = _collate(batch)
        except (TypeError, IndexError, KeyError):
            batch = _collate([self.data[i] for i in idx])
        return batch if isinstance(batch, tuple) else (batch,)

    def _stage(self, slot: int, batch: tuple):
        """host batch → pinned slot → device slot (async)."""
        if not self._cuda:
            return batch
        if self._pinned[slot] is None or any(
                p.shape != b.shape or p.dtype != b.dtype for p, b in zip(self._pinned[slot], batch)):
            self._pinned[slot] = tuple(torch.empty(b.shape, dtype=b.dtype, pin_memory=True)
                                       for b in batch)
            self._dev[slot] = tuple(torch.empty(b.shape, dtype=b.dtype, device=self.device)
                                    for b in batch)
        if self._events[slot] is not None:
            self._events[slot].synchronize()     # pinned slot free to overwrite
        # a batch that already sits contiguously in pinned memory (a view of a pinned dataset)
        # is DMA'd from where it is; anything else is first assembled in the pinned slot
        direct = all(b.is_pinned() and b.is_contiguous() for b in batch)
        src = batch if direct else self._pinned[slot]
        if not direct:
            for p, b in zip(self._pinned[slot], batch):
                p.copy_(b)
        with torch.cuda.stream(self._stream):
            if self._consumed[slot] is not None:
                self._stream.wait_event(self._consumed[slot])   # device slot no longer in use
            for d, p in zip(self._dev[slot], src):
                d.copy_(p, non_blocking=True)
                self.h2d_bytes += p.numel() * p.element_size()
            ev = torch.cuda.Event()
            ev.record(self._stream)
        self._events[slot] = ev
        return self._dev[slot]

    def __iter__(self) -> Iterator[tuple]:
        """Keeps ``depth - 1`` batches staged ahead of the one being consumed. When the
        consumer comes back for the next batch an event is recorded on its stream; the device
        slot is only overwritten after that event (so the consumer's queued work is safe)."""
        from collections import deque
        order = self._indices()
        self.epoch += 1
        queue: deque = deque()
        slot = 0
        for b in range(len(self)):
            idx = order[b * self.batch_size:(b + 1) * self.batch_size]
            queue.append((slot, self._stage(slot, self._gather(idx))))
            slot = (slot + 1) % self.depth
            if len(queue) == self.depth:
                yield from self._hand_out(queue.popleft())
        while queue:
            yield from self._hand_out(queue.popleft())

    def _hand_out(self, item):
        slot, staged = item
        if not self._cuda:
            yield staged
            return
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self._events[slot])
        yield staged
        done = torch.cuda.Event()
        done.record(torch.cuda.current_stream(self.device))
        self._consumed[slot] = done
