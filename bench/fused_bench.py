#!/usr/bin/env python
"""The headline kernel in isolation: fused gradient all-reduce + Adam step over a model-sized
flat buffer, against the baseline it replaces (NCCL all-reduce + torch's fused Adam).

    python -m fluxmpi_b200.launch -n 8 bench/fused_bench.py --out gpurun_out/fused_8.json

Sizes: ResNet-50 (25.6 M fp32), ViT-B/16 (86.6 M fp32), GPT-2 medium (354.8 M bf16 + fp32
master). Device time, CUDA events, max over ranks; L2 is flushed between iterations
(`fmpi_flush_l2` writes 256 MB). Roofline per BASELINE.json: the slower of the optimiser's HBM
bytes at the measured copy bandwidth and the bytes that must cross NVLink at the measured
770 GB/s per direction:
  P2P two-shot: in = (W-1)/W * N * es (gradient reads) + (W-1)/W * N * es (parameter stores)
  NVLS:         in = N/W * es (reduced gradient) + (W-1)/W * N * es (multicast parameters)
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import fluxmpi_b200 as fm                                          # noqa: E402
from fluxmpi_b200 import rules as R                                # noqa: E402
from fluxmpi_b200.ops import fused as F                            # noqa: E402
from fluxmpi_b200.parallel.engine import DataParallelEngine        # noqa: E402

HBM_GBS, LINK_GBS = 6540.8, 770.0


class Flat(torch.nn.Module):
    def __init__(self, n, dtype):
        super().__init__()
        self.w = torch.nn.Parameter(torch.randn(n, dtype=dtype, device="cuda") * 0.01)


def timeit(fn, prep, iters, dev):
    for _ in range(3):
        prep(); fn()
    fm.barrier()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        prep()
        F.flush_l2(dev)
        fm.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    t = torch.tensor([tot / iters], dtype=torch.float64)
    fm.allreduce_(t, "max")
    return float(t) * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--iters", type=int, default=8)
    ap.add_argument("--blocks", default="16,48")
    ap.add_argument("--algos", default="twoshot,nvls")
    ap.add_argument("--models", default="resnet50,vit_b16,gpt2_medium")
    ap.add_argument("--no-baseline", action="store_true")
    a = ap.parse_args()
    fm.Init()
    r, w = fm.local_rank(), fm.total_workers()
    dev = fm.current_device()
    dw = fm.runtime.state().device_world
    dist = None
    if w > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", rank=r, world_size=w, device_id=dev)
    rows = []
    for name, n, dt in (("resnet50", 25_557_056, torch.float32), ("vit_b16", 86_567_680, torch.float32),
                        ("gpt2_medium", 354_823_168, torch.bfloat16)):
        if name not in a.models.split(","):
            continue
        es = 2 if dt == torch.bfloat16 else 4
        row = {"model": name, "numel": n, "dtype": str(dt).split(".")[-1], "world": w}
        for algo in ([x for x in a.algos.split(",") if x != "nvls" or dw.has_multicast] if w > 1 else ["twoshot"]):
            for blocks in ((0,) if w == 1 else tuple(int(b) for b in a.blocks.split(","))):
                mod = Flat(n, dt)
                eng = DataParallelEngine(mod, R.Adam(1e-3), bucket_mb=1e9, overlap=False, algo=algo,
                                         blocks=blocks, sync_params=False)
                us = timeit(eng.step, lambda: eng.G.normal_(), a.iters, dev)
                key = f"ours_{algo}_b{blocks}_us"
                row[key] = round(us, 1)
                eng.close()
                del eng, mod
                torch.cuda.empty_cache()
        # baseline: NCCL all-reduce + torch fused Adam (fp32 master for bf16, like the engine)
        p = torch.randn(n, dtype=torch.float32, device=dev).requires_grad_(True)
        g = torch.empty(n, dtype=dt, device=dev)
        lp = torch.empty(n, dtype=dt, device=dev) if dt != torch.float32 else None
        opt = torch.optim.Adam([p], lr=1e-3, fused=True)

        def base():
            if dist is not None:
                dist.all_reduce(g)
            p.grad = g if dt == torch.float32 else g.float()
            opt.step()
            if lp is not None:
                lp.copy_(p.detach())

        row["nccl_plus_fused_adam_us"] = None if a.no_baseline else round(timeit(base, lambda: g.normal_(), a.iters, dev), 1)
        hbm_owner = (8 * n * es / 4 + 20 * n / w) if dt == torch.float32 else (n * (2 * es + 2) + 24 * n / w)
        row["roofline_p2p_us"] = round(max(hbm_owner / HBM_GBS / 1e3, 2 * (w - 1) / w * n * es / LINK_GBS / 1e3), 1)
        row["roofline_nvls_us"] = round(max(hbm_owner / HBM_GBS / 1e3, n * es / LINK_GBS / 1e3), 1)
        ours = min(v for k, v in row.items() if k.startswith("ours_"))
        row["best_ours_us"] = ours
        row["speedup_vs_nccl_adam"] = None if a.no_baseline else round(row["nccl_plus_fused_adam_us"] / ours, 2)
        row["fraction_of_roofline"] = round(min(row["roofline_p2p_us"], row["roofline_nvls_us"]) / ours, 3) if w > 1 \
            else round((32 * n / HBM_GBS / 1e3) / ours, 3)
        rows.append(row)
        if r == 0:
            print(json.dumps(row), flush=True)
        del p, g, lp, opt
        torch.cuda.empty_cache()
    if r == 0 and a.out:
        Path(a.out).write_text(json.dumps({"world": w, "rows": rows, "hbm_gbs": HBM_GBS, "link_gbs": LINK_GBS}, indent=1))
    if dist is not None:
        dist.destroy_process_group()
    fm.Finalize()


if __name__ == "__main__":
    main()
