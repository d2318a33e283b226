"""Lightweight tracing: named regions with NVTX ranges (visible in Nsight) and CUDA-event or
wall-clock timings collected per rank. The reference has no tracing or profiling hooks at all
(its examples hand-time with ``time()``, README.md:59,69).

    with fm.trace.region("backward"):
        loss.backward()
    fm.trace.report()           # rank-ordered table through fluxmpi_println
"""
from __future__ import annotations

import contextlib
import time
from collections import defaultdict

import torch

_enabled = True
_records: dict[str, list[float]] = defaultdict(list)
_pending: list = []


def enable(flag: bool = True):
    global _enabled
    _enabled = flag


@contextlib.contextmanager
def region(name: str, sync: bool = False):
    """Time a region. On CUDA the timing uses events on the current stream and is resolved
    lazily (no host sync unless ``sync=True``)."""
    if not _enabled:
        yield
        return
    cuda = torch.cuda.is_available() and torch.cuda.is_initialized()
    if cuda:
        torch.cuda.nvtx.range_push(name)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.perf_counter()
    try:
        yield
    finally:
        if cuda:
            e1.record()
            torch.cuda.nvtx.range_pop()
            _pending.append((name, e0, e1))
            if sync:
                _resolve()
        else:
            _records[name].append((time.perf_counter() - t0) * 1e3)


def _resolve():
    while _pending:
        name, e0, e1 = _pending.pop(0)
        e1.synchronize()
        _records[name].append(e0.elapsed_time(e1))


def summary() -> dict:
    """{region: {"count", "mean_ms", "min_ms", "max_ms", "total_ms"}} for this rank."""
    _resolve()
    return {k: {"count": len(v), "mean_ms": sum(v) / len(v), "min_ms": min(v), "max_ms": max(v),
                "total_ms": sum(v)} for k, v in _records.items() if v}


def reset():
    _pending.clear()
    _records.clear()
    _counters.clear()


# ---- collective counters: how many calls and bytes went through each public collective, host
# and device side separately (the reference keeps no counters or bandwidth statistics)
_counters: dict = defaultdict(lambda: [0, 0])


def count(kind: str, where: str, nbytes: int) -> None:
    c = _counters[(kind, where)]
    c[0] += 1
    c[1] += int(nbytes)


def counters() -> dict:
    """{"allreduce/host": {"calls": n, "bytes": b}, "bcast/cuda": ...} for this rank."""
    return {f"{k}/{w}": {"calls": c[0], "bytes": c[1]} for (k, w), c in sorted(_counters.items())}


def report():
    from .. import runtime
    for k, s in summary().items():
        runtime.fluxmpi_println(f"{k:24s} n={s['count']:5d} mean={s['mean_ms']:9.3f} ms "
                                f"min={s['min_ms']:9.3f} max={s['max_ms']:9.3f}")
