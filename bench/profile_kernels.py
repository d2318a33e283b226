#!/usr/bin/env python
"""Launch each hot kernel a few times on ONE GPU so `ncu` can capture it (B200_PROFILING.md:
run this plainly first, then under `ncu --set full -k regex:<kernel> ...`).

  fused_step_kernel        world = 1, a ResNet-50-sized flat fp32 vector (25.6 M parameters):
                           the launch bench.py makes once per step at N = 1 — HBM-bound
                           (28 B/element: read g, p, m, v; write p, m, v, and clear g = 32 B).
  fused_step_kernel (W=8)  eight emulated ranks on this GPU (cooperative launch): same code as
                           across NVLink, here the "peer" loads hit local HBM.
  oneshot_push / twoshot_stage / twoshot_symm / bcast_stage   emulated W = 8.
  multi_tensor_step, pack, sumsq, checksum

Also prints CUDA-event timings (the numbers to quote; nothing measured under ncu is).
"""
from __future__ import annotations

import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from fluxmpi_b200 import _native                                   # noqa: E402
from fluxmpi_b200 import rules as R                                # noqa: E402
from fluxmpi_b200.ops import fused as F                            # noqa: E402
from fluxmpi_b200.parallel.device_world import EmulatedWorld       # noqa: E402

DEV = torch.device("cuda", 0)
torch.cuda.set_device(DEV)
N = 25_557_056            # ResNet-50 parameter count rounded to 64 elements


ITERS = int(os.environ.get("PROFILE_ITERS", 10))     # PROFILE_ITERS=1 under ncu: one capture per kernel


def ev_time(fn, iters=None, warm=None, flush=True):
    iters = ITERS if iters is None else iters
    warm = min(3, ITERS) if warm is None else warm
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        if flush:
            F.flush_l2(DEV)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e3    # us


def fused(world, n, dtype=torch.float32, rule=R.Adam(1e-3), blocks=0):
    ew = EmulatedWorld(world, DEV, window_mb=int(2 * n * 4 / (1 << 20)) + 8, stage_mb=1)
    es = torch.empty((), dtype=dtype).element_size()
    p_off, g_off = ew.heap_alloc(n * es), ew.heap_alloc(n * es)
    for r in range(world):
        ew.window(r, p_off, n * es).view(dtype).normal_()
        ew.window(r, g_off, n * es).view(dtype).normal_()
    h = F.hyper_for(rule, zero_grad=True)
    m = [torch.zeros(n, device=DEV) for _ in range(world)]
    v = [torch.zeros(n, device=DEV) for _ in range(world)]
    step = [torch.zeros(1, dtype=torch.int64, device=DEV) for _ in range(world)]
    return lambda: ew.fused_step(g_off, p_off, dtype, m, v, None, 0, n, h, step, blocks=blocks)


def main():
    out = {}
    f1 = fused(1, N)
    us = ev_time(f1)
    out["fused_step W=1 fp32 Adam (ResNet-50 size)"] = (us, 32 * N / us / 1e3)
    f8 = fused(8, 64 * 4096 * 8)
    out["fused_step W=8 emulated (8 x 8 MB)"] = (ev_time(f8, flush=False), None)

    ew = EmulatedWorld(8, DEV, window_mb=16, stage_mb=24)
    small = [torch.ones(4096, device=DEV) for _ in range(8)]
    big = [torch.ones(1 << 20, device=DEV) for _ in range(8)]
    out["oneshot_push W=8 16 KB"] = (ev_time(lambda: ew.allreduce(small, _native.SUM, algo=_native.ALGO_ONESHOT), flush=False), None)
    out["twoshot_stage W=8 4 MB"] = (ev_time(lambda: ew.allreduce(big, _native.SUM, algo=_native.ALGO_TWOSHOT), flush=False), None)
    out["bcast_stage W=8 4 MB"] = (ev_time(lambda: ew.bcast(big, 0, algo=_native.ALGO_TWOSHOT), flush=False), None)
    off = ew.heap_alloc(4 << 20)
    out["twoshot_symm W=8 4 MB"] = (ev_time(lambda: ew.allreduce_symm(off, 1 << 20, torch.float32, _native.SUM), flush=False), None)

    shapes = [(64, 3, 7, 7), (256, 64, 1, 1), (512, 512, 3, 3), (1000, 2048), (2048,)] * 20
    ps = [torch.randn(s, device=DEV) for s in shapes]
    gs = [torch.randn(s, device=DEV) for s in shapes]
    ms = [torch.zeros(s, device=DEV) for s in shapes]
    vs = [torch.zeros(s, device=DEV) for s in shapes]
    tab = F.TensorTable(ps, gs, ms, vs)
    h = F.hyper_for(R.Adam(1e-3))
    nel = sum(p.numel() for p in ps)
    us = ev_time(lambda: F.multi_tensor_step(tab, h, 0, torch.float32))
    out[f"multi_tensor_step Adam {len(ps)} tensors {nel / 1e6:.1f} M elements"] = (us, 28 * nel / us / 1e3)
    x = torch.randn(N, device=DEV)
    us = ev_time(lambda: F.sumsq(x))
    out["sumsq 25.6 M f32"] = (us, 4 * N / us / 1e3)
    us = ev_time(lambda: F.checksum(x))
    out["checksum 102 MB"] = (us, 4 * N / us / 1e3)
    for k, (us, gbs) in out.items():
        print(f"{k:58s} {us:9.1f} us" + (f"  {gbs:8.1f} GB/s" if gbs else ""))


if __name__ == "__main__":
    main()
