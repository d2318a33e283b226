#!/usr/bin/env python
"""BatchNorm(+add)(+ReLU) and max-pool kernels on real ResNet-50 layer shapes (batch 256, bf16,
channels_last) — one GPU. Run plainly for CUDA-event timings (L2 flushed between iterations;
algorithmic bytes / time against MEASURED_PEAKS.hbm_gbs), or under
`ncu --set full -k regex:'bn_|maxpool'` with PROFILE_ITERS=1 for one capture per launch.

    python bench/profile_bn.py [--out gpurun_out/bn_kernels.json]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from fluxmpi_b200.ops import fused as F                         # noqa: E402
from fluxmpi_b200.ops.batchnorm import FusedBatchNormAct        # noqa: E402
from fluxmpi_b200.ops.pool import FusedMaxPool2d                # noqa: E402

ITERS = int(os.environ.get("PROFILE_ITERS", 10))
# (H = W, C, relu, residual): the BN layers that carry most of the bytes of a ResNet-50 step
SHAPES = [(112, 64, True, False), (56, 64, True, False), (56, 256, True, True), (56, 256, False, False),
          (28, 128, True, False), (28, 512, True, True), (14, 256, True, False), (14, 1024, True, True),
          (7, 512, True, False), (7, 2048, True, True)]


def peak_gbs() -> float:
    p = Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json"
    try:
        return float(json.loads(p.read_text())["hbm_gbs"])
    except (OSError, ValueError, KeyError):
        return 6650.0


def ev(fn, prep=None):
    for _ in range(min(3, ITERS)):
        if prep:
            prep()
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(ITERS):
        if prep:
            prep()
        F.flush_l2()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--out", default="")
    ap.add_argument("--only", default="", help="hw,C pairs to keep, e.g. 28,512;56,64 (ncu runs: keep it short)")
    ap.add_argument("--no-pool", action="store_true")
    a = ap.parse_args()
    keep = {tuple(int(v) for v in item.split(",")) for item in a.only.split(";") if item}
    dev = torch.device("cuda", 0)
    peak = peak_gbs()
    rows = []
    for hw, C, relu, res in SHAPES:
        if keep and (hw, C) not in keep:
            continue
        shape = (a.batch, C, hw, hw)
        x = torch.randn(shape, device=dev).bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
        r = torch.randn(shape, device=dev).bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True) if res else None
        bn = FusedBatchNormAct(C, relu=relu).to(dev)
        dy = torch.randn(shape, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
        nbytes = x.numel() * 2
        ys = []

        def fwd():
            ys.clear()
            ys.append(bn(x, r) if res else bn(x))

        def bwd():
            ys[0].backward(dy)

        t_f = ev(fwd)
        t_b = ev(bwd, prep=fwd)
        # passes over the activation tensor: fwd = stats read + apply (read x [+res], write y);
        # bwd = reduce (dy, x [, y]) + apply (dy, x [, y] read; dx [, dres] written)
        p_f = 1 + 2 + (1 if res else 0)
        need_y = relu and res               # the ReLU mask is recomputed from x unless a residual went in
        p_b = (2 + (1 if need_y else 0)) * 2 + 1 + (1 if (res and relu) else 0)
        rows.append({"hw": hw, "C": C, "relu": relu, "residual": res, "tensor_mb": round(nbytes / 2 ** 20, 1),
                     "fwd_us": round(t_f, 1), "fwd_passes": p_f, "fwd_gbs": round(p_f * nbytes / t_f / 1e3, 1),
                     "fwd_frac_of_measured_hbm": round(p_f * nbytes / t_f / 1e3 / peak, 3),
                     "bwd_us": round(t_b, 1), "bwd_passes": p_b, "bwd_gbs": round(p_b * nbytes / t_b / 1e3, 1),
                     "bwd_frac_of_measured_hbm": round(p_b * nbytes / t_b / 1e3 / peak, 3)})
        print(json.dumps(rows[-1]), flush=True)
        del x, r, dy, bn, ys
        torch.cuda.empty_cache()
    if a.no_pool:
        return _finish(a, peak, rows)
    # stem max-pool 3x3 / 2 on 112x112x64
    x = torch.randn(a.batch, 64, 112, 112, device=dev).bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    mp = FusedMaxPool2d(3, 2, 1)
    ys = []

    def pf():
        ys.clear()
        ys.append(mp(x))

    t_f = ev(pf)
    dy = torch.randn_like(ys[0])
    t_b = ev(lambda: ys[0].backward(dy), prep=pf)
    nin, nout = x.numel() * 2, ys[0].numel() * 2
    rows.append({"op": "maxpool3x3s2 112x112x64", "fwd_us": round(t_f, 1),
                 "fwd_gbs": round((nin + nout + nout // 2) / t_f / 1e3, 1),
                 "bwd_us": round(t_b, 1), "bwd_gbs": round((nin + nout + nout // 2) / t_b / 1e3, 1)})
    print(json.dumps(rows[-1]), flush=True)
    _finish(a, peak, rows)


def _finish(a, peak, rows):
    if a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps({"hbm_gbs_measured": peak, "batch": a.batch, "dtype": "bf16",
                                           "timing": "CUDA events, median, L2 flushed", "rows": rows}, indent=1))


if __name__ == "__main__":
    main()
