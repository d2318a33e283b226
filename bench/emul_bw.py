#!/usr/bin/env python
"""HBM-side efficiency of the cross-GPU kernels' copy / reduce loops, measured on ONE GPU.

In single-GPU emulation every "peer" access lands in local HBM, so a W-rank collective becomes
a pure memory-traffic problem whose ideal time is (bytes moved by all virtual ranks) / (copy
bandwidth). The ratio shows how much of the multi-GPU time can be blamed on the loops rather
than on NVLink, and lets the loops be tuned (and profiled with ncu) without a second GPU.
"""
from __future__ import annotations

import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from fluxmpi_b200 import _native                                   # noqa: E402
from fluxmpi_b200.ops import fused as F                            # noqa: E402
from fluxmpi_b200.parallel.device_world import EmulatedWorld       # noqa: E402

DEV = torch.device("cuda", 0)
torch.cuda.set_device(DEV)
HBM = 6540.8e9


def ev(fn, iters=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        F.flush_l2(DEV)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters * 1e-3


def main():
    for w, mb in ((2, 256), (4, 128), (8, 64)):
        n = mb << 20
        ew = EmulatedWorld(w, DEV, window_mb=mb + 8, stage_mb=2 * mb + 8)
        xs = [torch.ones(n // 4, device=DEV) for _ in range(w)]
        f = (w - 1) / w
        # two-shot staged: A push (read f*N, write f*N), B reduce (read N, write N), C copy out (read f*N, write f*N)
        t = ev(lambda: ew.allreduce(xs, _native.SUM, algo=_native.ALGO_TWOSHOT))
        traffic = w * n * (4 * f + 2)
        print(f"W={w} {mb:4d} MB/rank twoshot_stage {t * 1e6:9.1f} us  traffic {traffic / 1e9:6.2f} GB  "
              f"{traffic / t / 1e12:5.2f} TB/s = {traffic / t / HBM:5.1%} of copy BW")
        # in place on the symmetric window: read W copies of the own slice (N), write W copies (N)
        off = ew.heap_alloc(n)
        t = ev(lambda: ew.allreduce_symm(off, n // 4, torch.float32, _native.SUM))
        traffic = w * n * 2
        print(f"W={w} {mb:4d} MB/rank twoshot_symm  {t * 1e6:9.1f} us  traffic {traffic / 1e9:6.2f} GB  "
              f"{traffic / t / 1e12:5.2f} TB/s = {traffic / t / HBM:5.1%} of copy BW")
        # bcast staged, scatter + all-gather: root reads N, all write; receivers copy out
        t = ev(lambda: ew.bcast(xs, 0, algo=_native.ALGO_TWOSHOT))
        traffic = n * (2 + (w - 1) * 2) + (n * f * (w - 2) * 2 if w > 2 else 0)
        print(f"W={w} {mb:4d} MB/rank bcast_stage   {t * 1e6:9.1f} us  traffic {traffic / 1e9:6.2f} GB  "
              f"{traffic / t / 1e12:5.2f} TB/s = {traffic / t / HBM:5.1%} of copy BW")
        del ew, xs
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
