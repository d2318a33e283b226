#!/usr/bin/env python
"""The headline kernel in isolation: fused gradient all-reduce + Adam step over a model-sized
flat buffer, against the baseline it replaces (NCCL all-reduce + torch's fused Adam).

    python -m fluxmpi_b200.launch -n 8 bench/fused_bench.py --out gpurun_out/fused_8.json

Sizes: ResNet-50 (25.6 M fp32), ViT-B/16 (86.6 M fp32), GPT-2 medium (354.8 M bf16 + fp32
master). Device time, CUDA events, max over ranks; L2 is flushed between iterations
(`fmpi_flush_l2` writes 256 MB). Roofline: the slower of the optimiser's HBM bytes at the measured
copy bandwidth (MEASURED_PEAKS.json) and the bytes that must cross NVLink per direction per GPU,
  P2P two-shot: 2 * (W-1)/W * N * es      (gradient slices in, parameter slices in)
  NVLS:         (W+1)/W * N * es          (N/W reduced gradient + N multicast parameters in;
                                           N served + N/W stored out)
divided by the link rate; two fractions are reported: against 900 GB/s (nominal, the figure
BASELINE.json fixes) and against 770 GB/s (the measured peer-copy rate, B200_PROFILING.md).
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

LINK_NOMINAL, LINK_MEASURED = 900.0, 770.0


def hbm_gbs() -> float:
    try:
        return float(json.loads((Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
    except (OSError, ValueError, KeyError):
        return 6650.0


HBM_GBS = hbm_gbs()


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
    ap.add_argument("--blocks", default="48,96,128")
    ap.add_argument("--algos", default="twoshot,nvls")
    ap.add_argument("--models", default="mlp,resnet50,vit_b16,gpt2_medium")
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
    for name, n, dt in (("mlp", 263_744, torch.float32), ("resnet50", 25_557_056, torch.float32),
                        ("vit_b16", 86_567_680, torch.float32), ("gpt2_medium", 354_823_168, torch.bfloat16)):
        if name not in a.models.split(","):
            continue
        es = 2 if dt == torch.bfloat16 else 4
        row = {"model": name, "numel": n, "dtype": str(dt).split(".")[-1], "world": w, "bytes": n * es}
        algos = [x for x in a.algos.split(",") if x != "nvls" or dw.has_multicast] if w > 1 else ["twoshot"]
        if name == "mlp":
            algos = [x for x in algos if x != "oneshot"] + (["oneshot"] if w > 1 else [])
        else:
            algos = [x for x in algos if x != "oneshot"]
        for algo in algos:
            for blocks in ((0,) if (w == 1 or algo == "oneshot") else tuple(int(b) for b in a.blocks.split(","))):
                mod = Flat(n, dt)
                eng = DataParallelEngine(mod, R.Adam(1e-3), bucket_mb=1e9, overlap=False, algo=algo,
                                         blocks=blocks, sync_params=False)
                us = timeit(eng.step, lambda: eng.G.normal_(), a.iters, dev)
                row[f"ours_{algo}_b{blocks}_us"] = round(us, 1)
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
        hbm_us = hbm_owner / HBM_GBS / 1e3
        link = {"p2p": 2 * (w - 1) / w * n * es, "nvls": (w + 1) / w * n * es}
        ours = {k: v for k, v in row.items() if k.startswith("ours_")}
        best_key = min(ours, key=ours.get)
        best = ours[best_key]
        row["best_ours"] = best_key
        row["best_ours_us"] = best
        row["speedup_vs_nccl_adam"] = None if a.no_baseline else round(row["nccl_plus_fused_adam_us"] / best, 2)
        if w > 1:
            kind = "nvls" if "_nvls_" in best_key else "p2p"
            for tag, rate in (("900", LINK_NOMINAL), ("770", LINK_MEASURED)):
                roof = max(hbm_us, link[kind] / rate / 1e3)
                row[f"roofline_{kind}_us_at_{tag}"] = round(roof, 1)
                row[f"fraction_of_roofline_at_{tag}"] = round(roof / best, 3)
            # the bytes-minimal algorithm on this box is the in-switch one: also quote against it
            roof = max(hbm_us, link["nvls" if dw.has_multicast else "p2p"] / LINK_NOMINAL / 1e3)
            row["fraction_of_min_traffic_roofline_at_900"] = round(roof / best, 3)
        else:
            row["fraction_of_hbm_roofline"] = round((32 * n / HBM_GBS / 1e3) / best, 3)
        rows.append(row)
        if r == 0:
            print(json.dumps(row), flush=True)
        del p, g, lp, opt
        torch.cuda.empty_cache()
    if r == 0 and a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps({"world": w, "rows": rows, "hbm_gbs_measured": HBM_GBS,
                                           "link_gbs_nominal": LINK_NOMINAL, "link_gbs_measured": LINK_MEASURED,
                                           "timing": "CUDA events, L2 flushed, max over ranks"}, indent=1))
    if dist is not None:
        dist.destroy_process_group()
    fm.Finalize()


if __name__ == "__main__":
    main()
