#!/usr/bin/env python
"""CPU tier: the host world against gloo, and BASELINE config 1 (Quick-Start MLP, CPU, 3 ranks).

    python -m fluxmpi_b200.launch -n 3 bench/host_bench.py [--out profiles/runs/host_bench_3.json]

Needs no GPU. Two parts:

1. all-reduce / bcast of fp32 host tensors, 4 B .. 64 MB: this library's shared-memory backend
   (csrc/host_backend.cpp) vs ``torch.distributed`` with the gloo backend — the collective
   library a CPU-only PyTorch job would otherwise use (the image has no MPI). Wall time per
   call between barriers, median of the per-iteration MAX over ranks.
2. the reference's Quick Start (README.md:31-70: MLP 1-256-512-256-1, Adam(0.001), 16 samples
   per rank) as three training loops:
     * ``ours``          — ``DistributedOptimizer`` + ``update_``: the whole gradient tree as one
                           shared-memory message, then the native rule loop (csrc/host_rules.cpp);
     * ``ref_shape``     — the reference's shape (src/optimizer.jl:20-23): one blocking
                           all-reduce per leaf (through this library's backend), then the rule as
                           per-leaf array operations like Optimisers.jl's broadcasts;
     * ``ref_shape_gloo``— the same per-leaf loop with gloo doing the all-reduce.
"""
from __future__ import annotations

import argparse
import json
import statistics
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import fluxmpi_b200 as fm                                    # noqa: E402
from fluxmpi_b200 import rules as R                          # noqa: E402
from fluxmpi_b200.models.mlp import functional_forward, quickstart_mlp   # noqa: E402


def timed(fn, iters, warm):
    """Median over iterations of the slowest rank's wall time, in microseconds."""
    for _ in range(warm):
        fn()
    ts = torch.empty(iters, dtype=torch.float64)
    for i in range(iters):
        fm.barrier()
        t0 = time.perf_counter()
        fn()
        ts[i] = time.perf_counter() - t0
    fm.allreduce_(ts, "max")
    return statistics.median(ts.tolist()) * 1e6


def sweep(dist, max_mb):
    rows = []
    n = 1
    while n * 4 <= max_mb * (1 << 20):
        x = torch.ones(n)
        iters = 300 if n * 4 <= (1 << 16) else 60 if n * 4 <= (4 << 20) else 12
        warm = max(3, iters // 10)
        row = {"bytes": n * 4}
        row["allreduce_us"] = round(timed(lambda: fm.allreduce_(x, "sum"), iters, warm), 2)
        if n * 4 >= (1 << 16):          # the same on a tensor in symmetric host memory (fm.shm_empty)
            sx = fm.shm_empty((n,), torch.float32)
            sx.fill_(1)
            row["allreduce_symm_us"] = round(timed(lambda: fm.allreduce_(sx, "sum"), iters, warm), 2)
        x.fill_(1)
        row["bcast_us"] = round(timed(lambda: fm.bcast_(x, 0), iters, warm), 2)
        if dist is not None:
            x.fill_(1)
            row["gloo_allreduce_us"] = round(timed(lambda: dist.all_reduce(x), iters, warm), 2)
            x.fill_(1)
            row["gloo_bcast_us"] = round(timed(lambda: dist.broadcast(x, 0), iters, warm), 2)
            row["allreduce_speedup"] = round(row["gloo_allreduce_us"] / row["allreduce_us"], 2)
            row["bcast_speedup"] = round(row["gloo_bcast_us"] / row["bcast_us"], 2)
        rows.append(row)
        if fm.local_rank() == 0:
            print(json.dumps(row), flush=True)
        n *= 8
    return rows


def quickstart(mode, epochs, dist):
    """Epochs per second of the Quick-Start loop (after 5 warm-up epochs)."""
    r, w = fm.local_rank(), fm.total_workers()
    torch.manual_seed(r)
    model = quickstart_mlp()
    if mode == "engine":        # HostDataParallelEngine: flat vectors, symmetric host gradient buffer
        eng = fm.make_engine(model, fm.DistributedOptimizer(R.Adam(0.001)))
        x = torch.rand(16, 1)
        y = x ** 2

        def estep():
            l = ((model(x) - y) ** 2).sum() / w
            l.backward()
            eng.step()
            return l

        for _ in range(5):
            estep()
        fm.barrier()
        t0 = time.perf_counter()
        for _ in range(epochs):
            l = estep()
        fm.barrier()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64)
        fm.allreduce_(dt, "max")
        return epochs / float(dt), float(l.detach())
    ps = {k: v.detach().clone().requires_grad_(True) for k, v in model.named_parameters()}
    ps = fm.synchronize(ps, root_rank=0)
    x = torch.rand(16, 1)
    y = x ** 2
    adam = R.Adam(0.001)
    if mode == "ours":
        st = R.setup(fm.DistributedOptimizer(adam), ps)
    else:
        st = R.setup(adam, ps)

    def epoch():
        nonlocal st, ps
        l = ((functional_forward(ps, x) - y) ** 2).sum() / w
        gs = dict(zip(ps.keys(), torch.autograd.grad(l, list(ps.values()))))
        if mode == "ours":
            st, ps = R.update_(st, ps, gs)
        else:
            for g in gs.values():                   # one blocking message per leaf
                if mode == "ref_shape":
                    fm.allreduce_(g, "sum")
                else:
                    dist.all_reduce(g)
            st, ps = R._walk_update(st, ps, gs, True, {})    # the rule as per-leaf array ops
        return l

    for _ in range(5):
        epoch()
    fm.barrier()
    t0 = time.perf_counter()
    for _ in range(epochs):
        l = epoch()
    fm.barrier()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64)
    fm.allreduce_(dt, "max")
    return epochs / float(dt), float(l.detach())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=float, default=64)
    ap.add_argument("--epochs", type=int, default=300)
    ap.add_argument("--out", default="")
    ap.add_argument("--no-gloo", action="store_true")
    a = ap.parse_args()
    torch.set_num_threads(1)          # ranks share the machine; keep the compute side comparable
    fm.Init()
    r, w = fm.local_rank(), fm.total_workers()
    dist = None
    if not a.no_gloo and w > 1:
        import torch.distributed as dist
        dist.init_process_group("gloo", rank=r, world_size=w)
    rows = sweep(dist, a.max_mb)
    qs = {}
    for mode in ("ours", "engine", "ref_shape") + (("ref_shape_gloo",) if dist is not None else ()):
        eps, last = quickstart(mode, a.epochs, dist)
        qs[mode] = {"epochs_per_s": round(eps, 1), "last_local_loss": last}
        if r == 0:
            print(json.dumps({"quickstart": mode, **qs[mode]}), flush=True)
    if r == 0 and a.out:
        S = fm.runtime.state()
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps({
            "world": w, "nodes": S.nnodes, "timing": "wall clock between host barriers, max over ranks, median",
            "torch_threads": 1, "sweep_fp32": rows, "quickstart_mlp": qs}, indent=1))
    if dist is not None:
        dist.destroy_process_group()
    fm.Finalize()


if __name__ == "__main__":
    main()
