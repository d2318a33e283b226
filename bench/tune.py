#!/usr/bin/env python
"""Tuning grid for the cross-GPU kernels: algorithm x CTA count x threads per message size.

    python -m fluxmpi_b200.launch -n 4 bench/tune.py --out gpurun_out/tune_4.json

For every (op, size) it times each candidate launch shape (CUDA events, max over ranks) on
(a) arbitrary device buffers (the `allreduce_` / `bcast_` / `reduce_` path, staged through the
symmetric window) and (b) buffers that already live in a symmetric segment (the engine's
path: no staging copies, pure link phases). The best row per (op, world, size) is what
`fluxmpi_b200/parallel/tuner.py` loads from `fluxmpi_b200/parallel/tuned.json`.
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
from fluxmpi_b200._native import Launch, cuda_check          # noqa: E402

ALGOS = {"oneshot": _native.ALGO_ONESHOT, "twoshot": _native.ALGO_TWOSHOT, "nvls": _native.ALGO_NVLS}


def timeit(fn, iters, warm=3):
    for _ in range(warm):
        fn()
    fm.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64)
    fm.allreduce_(t, "max")
    return float(t) * 1e3      # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes-mb", default="0.25,1,4,16,64,256")
    ap.add_argument("--blocks", default="8,16,32,64,96,148")
    ap.add_argument("--threads", default="256,512")
    ap.add_argument("--dtype", default="float32")
    ap.add_argument("--out", default="")
    a = ap.parse_args()

    fm.Init()
    r, w = fm.local_rank(), fm.total_workers()
    dev = fm.current_device()
    dw = fm.runtime.state().device_world
    lib = _native.cuda()
    dt = getattr(torch, a.dtype)
    es = torch.empty((), dtype=dt).element_size()
    code = fm.comm.dtype_code(dt)
    st = lambda: torch.cuda.current_stream().cuda_stream
    sizes = [int(float(s) * (1 << 20)) for s in a.sizes_mb.split(",")]
    blocks = [int(b) for b in a.blocks.split(",")]
    threads = [int(t) for t in a.threads.split(",")]
    algos = ["twoshot"] + (["nvls"] if dw.has_multicast else [])

    seg = dw.alloc_segment(max(sizes) + (4 << 20))
    scomm = dw.comm_for(seg, 3)
    rows = []
    for nbytes in sizes:
        n = nbytes // es
        buf = torch.ones(n, dtype=dt, device=dev)
        iters = 50 if nbytes <= (4 << 20) else 15 if nbytes <= (64 << 20) else 6
        for algo in algos + (["oneshot"] if nbytes <= (256 << 10) else []):
            for b in blocks:
                for t in threads:
                    lp = C.byref(Launch(b, t, ALGOS[algo]))
                    row = {"world": w, "bytes": nbytes, "algo": algo, "blocks": b, "threads": t}
                    try:
                        us = timeit(lambda: dw._reduce_native(buf, buf, _native.SUM, -1, 0, launch=lp), iters)
                        row["allreduce_any_us"] = round(us, 2)
                        row["allreduce_any_busbw"] = round(nbytes / us / 1e3 * 2 * (w - 1) / w, 1)
                        us = timeit(lambda: dw._reduce_native(buf, buf, _native.SUM, 0, 0, launch=lp), iters)
                        row["reduce_any_us"] = round(us, 2)
                        row["reduce_any_busbw"] = round(nbytes / us / 1e3, 1)
                        p = C.c_void_p(buf.data_ptr())
                        us = timeit(lambda: cuda_check(lib.fmpi_bcast_any(
                            C.byref(dw.comms[0]), C.byref(p), nbytes, 0, dw.bc_off, dw.bc_bytes, lp, st())), iters)
                        row["bcast_any_us"] = round(us, 2)
                        row["bcast_any_busbw"] = round(nbytes / us / 1e3, 1)
                        if algo != "oneshot":
                            us = timeit(lambda: cuda_check(lib.fmpi_allreduce_symm(
                                C.byref(scomm), 0, n, code, _native.SUM, lp, st())), iters)
                            row["allreduce_symm_us"] = round(us, 2)
                            row["allreduce_symm_busbw"] = round(nbytes / us / 1e3 * 2 * (w - 1) / w, 1)
                            us = timeit(lambda: cuda_check(lib.fmpi_bcast_symm(
                                C.byref(scomm), 0, nbytes, 0, lp, st())), iters)
                            row["bcast_symm_us"] = round(us, 2)
                            row["bcast_symm_busbw"] = round(nbytes / us / 1e3, 1)
                    except _native.NativeError as e:
                        row["error"] = str(e)[:120]
                    rows.append(row)
                    if r == 0:
                        print(json.dumps(row), flush=True)
        del buf
    if r == 0 and a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        Path(a.out).write_text(json.dumps({"world": w, "dtype": a.dtype, "multicast": dw.has_multicast,
                                           "rows": rows}, indent=1))
    fm.Finalize()


if __name__ == "__main__":
    main()
