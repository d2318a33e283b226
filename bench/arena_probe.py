#!/usr/bin/env python
"""All-reduce of a large ORDINARY tensor: the single staged kernel against the opt-in path through
the symmetric arena (copy in, in-place kernel, copy out; FMPI_ARENA_ALLREDUCE_MB), same process,
correctness checked against the expected sum. Bus bandwidth, CUDA events, max over ranks.

    python -m fluxmpi_b200.launch -n 8 bench/arena_probe.py [--mb 64,256,1024]
"""
import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import fluxmpi_b200 as fm                                    # noqa: E402
from fluxmpi_b200.parallel.device_world import MB             # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mb", default="64,256")
ap.add_argument("--out", default="")
a = ap.parse_args()
fm.Init()
r, w = fm.local_rank(), fm.total_workers()
dev = fm.current_device()
dw = fm.runtime.state().device_world
rows = []
for mb in [float(x) for x in a.mb.split(",")]:
    n = int(mb * MB) // 4
    x = torch.empty(n, device=dev)
    row = {"bytes": n * 4, "world": w}
    for name, thr in (("staged", 0), ("arena", 1)):
        dw.ARENA_MIN_BYTES = thr * MB                        # 0 = off
        x.fill_(float(r + 1))
        fm.allreduce_(x, "sum")
        ok = bool(torch.all(x == w * (w + 1) / 2))
        iters = 6
        fm.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fm.allreduce_(x, "max")
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64)
        fm.allreduce_(ms, "max")
        row[name + "_ok"] = ok
        row[name + "_us"] = round(float(ms) * 1e3, 1)
        row[name + "_busbw_gbs"] = round(n * 4 / (float(ms) * 1e-3) * 2 * (w - 1) / w / 1e9, 1)
    rows.append(row)
    if r == 0:
        print(json.dumps(row), flush=True)
if r == 0 and a.out:
    Path(a.out).parent.mkdir(parents=True, exist_ok=True)
    Path(a.out).write_text(json.dumps({"world": w, "rows": rows}, indent=1))
fm.Finalize()
