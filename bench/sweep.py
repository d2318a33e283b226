#!/usr/bin/env python
"""Collective bandwidth sweep — BASELINE config 5: "allreduce/bcast/reduce bandwidth sweep
1 KB-1 GB at 2/4/8 GPUs vs ... NCCL" (the reference's CUDA-aware-MPI build cannot run here).

    python -m fluxmpi_b200.launch -n 8 bench/sweep.py [--max-mb 1024] [--out gpurun_out/sweep_8.json]

Every number is device time (CUDA events on the launching stream, after warm-up), max over
ranks. Bus bandwidth conventions: allreduce busbw = bytes/time * 2(W-1)/W,
bcast / reduce busbw = bytes/time. Between timed iterations the buffers rotate through a pool
larger than the 126 MB L2 for small messages; large messages exceed L2 by themselves.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import fluxmpi_b200 as fm                                    # noqa: E402


def time_op(fn, bufs, iters, warm, graph=False):
    """Average device ms per call of fn(buf) cycling through bufs. With graph=True the calls
    are captured into a CUDA graph first, so the number is device time free of host launch
    overhead (the regime that matters once a training step is itself a graph)."""
    for i in range(warm):
        fn(bufs[i % len(bufs)])
    fm.barrier()
    torch.cuda.synchronize()
    if graph:
        inner = min(iters, 20)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
                for i in range(inner):
                    fn(bufs[i % len(bufs)])
        reps = max(1, iters // inner)
        g.replay()
        torch.cuda.synchronize()
        fm.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / (reps * inner)
        t = torch.tensor([ms], dtype=torch.float64)
        fm.allreduce_(t, "max")
        return float(t)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(bufs[i % len(bufs)])
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    t = torch.tensor([ms], dtype=torch.float64)
    fm.allreduce_(t, "max")
    return float(t)


def nccl_registered_pool(dist, dev):
    """A torch memory pool whose blocks come from ncclMemAlloc and are registered with the
    communicator (NCCL's user-buffer registration: zero-copy / NVLS on the caller's tensor) — the
    like-for-like partner of fm.symm_empty. None when this torch / NCCL build has no such API."""
    try:
        backend = dist.distributed_c10d._get_default_group()._get_backend(dev)
        pool = torch.cuda.MemPool(backend.mem_allocator)
        return backend, pool
    except Exception as e:                                       # noqa: BLE001
        return None, f"{type(e).__name__}: {str(e)[:120]}"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=float, default=1024)
    ap.add_argument("--min-bytes", type=int, default=1024)
    ap.add_argument("--dtype", default="float32")
    ap.add_argument("--out", default="")
    ap.add_argument("--ops", default="allreduce,bcast,reduce")
    ap.add_argument("--no-nccl", action="store_true")
    ap.add_argument("--graph", action="store_true", help="time CUDA-graph replays (no host launch cost)")
    ap.add_argument("--columns", default="any,symm,nccl,nccl_reg",
                    help="any = ours on ordinary tensors, symm = ours on fm.symm_empty tensors, "
                         "nccl = NCCL on ordinary tensors, nccl_reg = NCCL on registered (ncclMemAlloc) tensors")
    a = ap.parse_args()
    cols = a.columns.split(",")

    fm.Init()
    r, w = fm.local_rank(), fm.total_workers()
    dev = fm.current_device()
    dw = fm.runtime.state().device_world
    dt = getattr(torch, a.dtype)
    es = torch.empty((), dtype=dt).element_size()

    dist = None
    reg_backend = reg_pool = None
    reg_note = ""
    if not a.no_nccl and w > 1 and ("nccl" in cols or "nccl_reg" in cols):
        import torch.distributed as dist
        dist.init_process_group("nccl", rank=r, world_size=w, device_id=dev)
        if "nccl_reg" in cols:
            reg_backend, reg_pool = nccl_registered_pool(dist, dev)
            if reg_backend is None:
                reg_note, reg_pool = reg_pool, None

    sizes = []
    s = a.min_bytes
    while s <= a.max_mb * (1 << 20):
        sizes.append(s)
        s *= 4
    rows = []
    pool_bytes = 192 << 20
    symm_cache: dict = {}
    reg_big = []
    if reg_pool is not None:
        # two maximum-size tensors from the registered pool, registered once; smaller messages
        # are slices of them (registration covers the whole allocation)
        try:
            with torch.cuda.use_mem_pool(reg_pool):
                reg_big = [torch.ones(sizes[-1] // es, dtype=dt, device=dev) for _ in range(2)]
            reg_backend.register_mem_pool(reg_pool)
        except Exception as e:                                   # noqa: BLE001
            reg_note, reg_pool, reg_big = f"{type(e).__name__}: {str(e)[:120]}", None, []
    for nbytes in sizes:
        n = nbytes // es
        nbuf = max(1, min(64, pool_bytes // nbytes))
        iters = 200 if nbytes <= (1 << 20) else 40 if nbytes <= (64 << 20) else 10
        warm = max(3, iters // 10)
        sets = {}
        if "any" in cols or "nccl" in cols:
            sets["ordinary"] = [torch.ones(n, dtype=dt, device=dev) for _ in range(nbuf)]
        if "symm" in cols and dw.native:
            k = max(1, min(nbuf, 4))
            if nbytes not in symm_cache:
                symm_cache[nbytes] = [fm.symm_empty(n, dt) for _ in range(k)]
            sets["symm"] = symm_cache[nbytes]
        if reg_big:
            per = max(1, min(nbuf // 2, reg_big[0].numel() // n))
            sets["nccl_reg"] = [big[i * n:(i + 1) * n] for big in reg_big for i in range(per)]
        row = {"bytes": nbytes, "world": w}
        for op in a.ops.split(","):
            if op == "allreduce":
                ours = lambda b: fm.allreduce_(b, "sum")
                nccl = (lambda b: dist.all_reduce(b)) if dist else None
                factor = 2 * (w - 1) / w
            elif op == "bcast":
                ours = lambda b: fm.bcast_(b, 0)
                nccl = (lambda b: dist.broadcast(b, 0)) if dist else None
                factor = 1.0
            else:
                ours = lambda b: fm.reduce_(b, "sum", 0)
                nccl = (lambda b: dist.reduce(b, 0)) if dist else None
                factor = 1.0
            plan = [("any", ours, "ordinary"), ("symm", ours, "symm"), ("nccl", nccl, "ordinary"),
                    ("nccl_reg", nccl, "nccl_reg")]
            for col, fn, key in plan:
                if col not in cols or fn is None or key not in sets:
                    continue
                bufs = sets[key]
                for b in bufs:
                    b.fill_(1)
                ms = time_op(fn, bufs, iters, warm, a.graph)
                row[f"{op}_{col}_us"] = round(ms * 1e3, 2)
                row[f"{op}_{col}_busbw_gbs"] = round(nbytes / (ms * 1e-3) * factor / 1e9, 2)
            best_nccl = min((row[f"{op}_{c}_us"] for c in ("nccl", "nccl_reg") if f"{op}_{c}_us" in row), default=None)
            if best_nccl is not None:
                for col in ("any", "symm"):
                    if f"{op}_{col}_us" in row:
                        row[f"{op}_{col}_speedup"] = round(best_nccl / row[f"{op}_{col}_us"], 3)
        rows.append(row)
        if r == 0:
            print(json.dumps(row), flush=True)
        sets.clear()
        torch.cuda.empty_cache()
    if r == 0 and a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps({
            "world": w, "dtype": a.dtype, "mode": dw.mode, "multicast": dw.has_multicast,
            "nccl_registered": "ncclMemAlloc pool registered with the communicator" if reg_pool is not None
            else f"unavailable ({reg_note})" if reg_note else "not requested",
            "speedup": "NCCL's best column (ordinary or registered) time / ours",
            "timing": "CUDA events, max over ranks" + (", CUDA-graph replay" if a.graph else ", eager launches"),
            "rows": rows}, indent=1))
    if dist:
        dist.destroy_process_group()
    fm.Finalize()


if __name__ == "__main__":
    main()
