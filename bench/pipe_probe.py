#!/usr/bin/env python
"""Probe of the chunk-pipelined staged collectives (ALGO_PIPE): bus bandwidth of one message size
for the launch parameters in the environment (FMPI_PIPE_*), and with FMPI_PIPE_TRACE=1 a
per-role, per-chunk timeline of the last call (where each role waits and how long it works).

    FMPI_PIPE_TRACE=1 python -m fluxmpi_b200.launch -n 2 bench/pipe_probe.py --mb 256 --op allreduce
"""
from __future__ import annotations

import argparse
import ctypes as C
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import fluxmpi_b200 as fm                                    # noqa: E402
from fluxmpi_b200 import _native                             # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=float, default=256)
    ap.add_argument("--op", default="allreduce", choices=["allreduce", "bcast", "reduce"])
    ap.add_argument("--iters", type=int, default=8)
    a = ap.parse_args()
    fm.Init()
    r, w = fm.local_rank(), fm.total_workers()
    dev = fm.current_device()
    dw = fm.runtime.state().device_world
    n = int(a.mb * (1 << 20)) // 4
    x = torch.ones(n, device=dev)
    lp = C.byref(_native.Launch(0, 0, _native.ALGO_PIPE))

    def call():
        if a.op == "allreduce":
            dw._reduce_native(x, x, _native.SUM, -1, 0, launch=lp)
        elif a.op == "reduce":
            dw._reduce_native(x, x, _native.SUM, 0, 0, launch=lp)
        else:
            dw._in1.value = x.data_ptr()
            _native.cuda_check(dw._lib.fmpi_bcast_any(dw._comm_refs[0], dw._in1_ref, n * 4, 0, dw.bc_off, dw.bc_bytes, lp,
                                                      torch.cuda.current_stream().cuda_stream), "bcast")

    for _ in range(3):
        call()
        x.fill_(1.0)
    fm.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        call()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / a.iters], dtype=torch.float64)
    fm.allreduce_(ms, "max")
    factor = 2 * (w - 1) / w if a.op == "allreduce" else 1.0
    cfg = {k: v for k, v in os.environ.items() if k.startswith("FMPI_PIPE")}
    if r == 0:
        print(f"{a.op} {a.mb} MB W={w}: {float(ms) * 1e3:.1f} us, busbw {n * 4 / (float(ms) * 1e-3) * factor / 1e9:.1f} GB/s  {cfg}", flush=True)
    if os.environ.get("FMPI_PIPE_TRACE", "0") not in ("0", ""):
        x.fill_(1.0)
        torch.cuda.synchronize()
        fm.barrier()
        call()
        buf = (C.c_uint64 * (3 * 256 * 4))()
        _native.cuda_check(dw._lib.fmpi_pipe_trace_read(buf, len(buf)), "trace")
        if r == 0:
            t0 = min(v for v in buf if v)
            for role, name in enumerate(("in", "link", "out")):
                rows = []
                for k in range(256):
                    s = [buf[(role * 256 + k) * 4 + i] for i in range(4)]
                    if not s[0]:
                        break
                    rows.append(f"k{k}: start {(s[0] - t0) / 1e3:7.1f} wait {(s[1] - s[0]) / 1e3:6.1f} work {(s[2] - s[1]) / 1e3:6.1f} handover {(s[3] - s[2]) / 1e3:5.1f}")
                print(f"--- role {name} (us) ---")
                print("\n".join(rows[:14]))
    fm.Finalize()


if __name__ == "__main__":
    main()
