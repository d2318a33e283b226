#!/usr/bin/env python
"""In-place broadcast and reduce-to-root on symmetric memory, one algorithm at a time
(bus bandwidth = bytes / time; CUDA events, max over ranks): which of the multicast and the
peer-to-peer forms is faster at this world size.

    python -m fluxmpi_b200.launch -n 8 bench/symm_probe.py [--out gpurun_out/symm_probe_8.json]
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import fluxmpi_b200 as fm                                    # noqa: E402
from fluxmpi_b200 import _native                             # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="")
    ap.add_argument("--sizes-mb", default="16,64,256,1024")
    a = ap.parse_args()
    fm.Init()
    r, w = fm.local_rank(), fm.total_workers()
    dw = fm.runtime.state().device_world
    sizes = [int(float(x) * (1 << 20)) for x in a.sizes_mb.split(",")]
    buf = fm.symm_empty((max(sizes) // 4,), torch.float32)
    buf.fill_(1.0)
    refs, off, has_mc = dw.find_symmetric(buf)
    lib = dw._lib
    st = lambda: torch.cuda.current_stream().cuda_stream
    rows = []
    algos = {"bcast": [("root_multicast", _native.ALGO_NVLS), ("scatter_allgather", _native.ALGO_TWOSHOT),
                       ("pull_multicast", _native.ALGO_PIPE), ("pull_p2p", _native.ALGO_PULL_P2P)],
             "reduce": [("nvls", _native.ALGO_NVLS), ("twoshot", _native.ALGO_TWOSHOT)],
             "allreduce": [("nvls", _native.ALGO_NVLS), ("twoshot", _native.ALGO_TWOSHOT)]}
    for nbytes in sizes:
        n = nbytes // 4
        for op, cands in algos.items():
            for name, algo in cands:
                if not has_mc and algo in (_native.ALGO_NVLS, _native.ALGO_PIPE):
                    continue
                for blocks in ((16, 64, 148) if op != "bcast" or algo != _native.ALGO_NVLS else (16, 64)):
                    lp = C.byref(_native.Launch(blocks, 512, algo))

                    def call():
                        if op == "bcast":
                            rc = lib.fmpi_bcast_symm(refs[0], off, nbytes, 0, lp, st())
                        elif op == "reduce":
                            rc = lib.fmpi_reduce_symm(refs[0], off, n, _native.F32, _native.SUM, 0, lp, st())
                        else:
                            rc = lib.fmpi_allreduce_symm(refs[0], off, n, _native.F32, _native.SUM, lp, st())
                        _native.cuda_check(rc, op)

                    iters = 5 if nbytes >= (256 << 20) else 10
                    for _ in range(2):
                        call()
                    buf.fill_(1.0)
                    fm.barrier()
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    for _ in range(iters):
                        call()
                    e1.record()
                    torch.cuda.synchronize()
                    ms = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64)
                    fm.allreduce_(ms, "max")
                    factor = 2 * (w - 1) / w if op == "allreduce" else 1.0
                    row = {"op": op, "algo": name, "blocks": blocks, "bytes": nbytes, "world": w,
                           "us": round(float(ms) * 1e3, 1), "busbw_gbs": round(nbytes / (float(ms) * 1e-3) * factor / 1e9, 1)}
                    rows.append(row)
                    if r == 0:
                        print(json.dumps(row), flush=True)
                    buf.fill_(1.0)
    if r == 0 and a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps({"world": w, "multicast": has_mc, "timing": "CUDA events, max over ranks",
                                           "rows": rows}, indent=1))
    fm.Finalize()


if __name__ == "__main__":
    main()
