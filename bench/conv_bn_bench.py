#!/usr/bin/env python
"""Per-layer timing of the 1x1-convolution + BatchNorm(+ReLU) pairs of ResNet-50 (batch 256,
bf16, channels_last): cuDNN convolution followed by the fused BatchNorm kernels (what the
headline benchmark runs) against ``ops.conv_bn.conv1x1_bn_act`` (the tcgen05 GEMM that takes the
statistics in its epilogue). One GPU, CUDA events, L2 flushed between iterations.

    python bench/conv_bn_bench.py [--batch 256] [--iters 20] [--out gpurun_out/conv_bn_bench.json]

Prints one JSON line per layer shape and a total weighted by how often the shape occurs in the
network. Decision rule for turning ``fused_conv_bn`` on: the weighted forward+backward total
must be lower, not just the forward.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from fluxmpi_b200.ops import conv_bn                            # noqa: E402
from fluxmpi_b200.ops import fused as F                         # noqa: E402
from fluxmpi_b200.ops.batchnorm import FusedBatchNormAct        # noqa: E402

# (H = W, C_in, C_out, relu, residual, occurrences) of the stride-1 1x1 convolutions of ResNet-50
SHAPES = [
    (56, 64, 64, True, False, 1), (56, 256, 64, True, False, 2), (56, 64, 256, True, True, 3),
    (56, 64, 256, False, False, 1),                                        # layer1 downsample
    (56, 256, 128, True, False, 1), (28, 512, 128, True, False, 3), (28, 128, 512, True, True, 4),
    (28, 512, 256, True, False, 1), (14, 1024, 256, True, False, 5), (14, 256, 1024, True, True, 6),
    (14, 1024, 512, True, False, 1), (7, 2048, 512, True, False, 2), (7, 512, 2048, True, True, 3),
]


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(iters):
        F.flush_l2()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    return ms[len(ms) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--out", default="")
    ap.add_argument("--variant", type=int, default=0, help="0: one SM per tile, 1: CTA pairs (cta_group::2)")
    a = ap.parse_args()
    assert torch.cuda.is_available()
    if conv_bn.gemm_lib() is None:
        raise SystemExit("libfmpi_gemm.so unavailable: " + str(conv_bn.why_unavailable()))
    conv_bn.set_variant(a.variant)
    dev = torch.device("cuda", 0)
    rows, tot = [], {"cudnn_fwd": 0.0, "fused_fwd": 0.0, "cudnn_fwdbwd": 0.0, "fused_fwdbwd": 0.0}
    for hw, cin, cout, relu, res, count in SHAPES:
        torch.manual_seed(0)
        conv = torch.nn.Conv2d(cin, cout, 1, bias=False).to(dev).to(memory_format=torch.channels_last)
        bn = FusedBatchNormAct(cout, relu=relu).to(dev)
        x = torch.randn(a.batch, cin, hw, hw, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
        x.requires_grad_(True)
        r = torch.randn(a.batch, cout, hw, hw, device=dev).bfloat16().contiguous(memory_format=torch.channels_last) \
            if res else None
        g = torch.randn(a.batch, cout, hw, hw, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)

        def cudnn_fwd():
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return bn(conv(x), r)

        def fused_fwd():
            return conv_bn.conv1x1_bn_act(x, conv, bn, r)

        def fwdbwd(f):
            def run():
                out = f()
                torch.autograd.grad(out, [x, conv.weight, bn.weight, bn.bias], g)
            return run

        assert conv_bn.eligible(x, conv, bn, r)
        with torch.no_grad():
            err = float((cudnn_fwd().float() - fused_fwd().float()).abs().max())
        row = {"hw": hw, "cin": cin, "cout": cout, "relu": relu, "residual": res, "count": count,
               "max_abs_diff": round(err, 4),
               "cudnn_fwd_us": round(timed(cudnn_fwd, a.iters) * 1e3, 1),
               "fused_fwd_us": round(timed(fused_fwd, a.iters) * 1e3, 1),
               "cudnn_fwdbwd_us": round(timed(fwdbwd(cudnn_fwd), a.iters) * 1e3, 1),
               "fused_fwdbwd_us": round(timed(fwdbwd(fused_fwd), a.iters) * 1e3, 1)}
        m = a.batch * hw * hw
        row["fused_fwd_gbs"] = round((m * (cin + 2 * cout + (cout if res else 0)) * 2) / (row["fused_fwd_us"] * 1e-6) / 1e9, 1)
        for k in tot:
            tot[k] += row[k + "_us"] * count
        rows.append(row)
        print(json.dumps(row), flush=True)
    summary = {k + "_ms_per_step": round(v / 1e3, 3) for k, v in tot.items()}
    print(json.dumps({"weighted_total": summary}), flush=True)
    if a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps({"batch": a.batch, "timing": "CUDA events, median, L2 flushed",
                                           "rows": rows, "weighted_total": summary}, indent=1))


if __name__ == "__main__":
    main()
