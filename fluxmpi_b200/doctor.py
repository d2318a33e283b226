"""``python -m fluxmpi_b200.doctor`` — what will this machine do for a job?

Prints one line per check (``ok`` / ``warn`` / ``FAIL``) and exits non-zero if anything that a
job needs is missing: the toolchain and the built libraries, the host world (a two-rank
self-test through the launcher), the multi-node path (two pretend nodes over loopback TCP), and
on a GPU machine the topology, peer access and multicast capability and a two-GPU device
self-test. The reference has no such tool; `Init(verbose=true)` prints the device it picked
(src/common.jl:38) and that is all.
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent

SELF_TEST = r"""
import torch, fluxmpi_b200 as fm
fm.Init()
r, w = fm.local_rank(), fm.total_workers()
dev = fm.current_device()
t = torch.full((1000,), float(r + 1), device=dev)
fm.allreduce_(t, "sum")
assert float(t[0]) == w * (w + 1) / 2, float(t[0])
b = torch.full((7,), float(r), device=dev)
fm.bcast_(b, w - 1)
assert float(b[0]) == w - 1
h = torch.full((3,), float(r))
fm.allreduce_(h, "max")
assert float(h[0]) == w - 1
dw = fm.runtime.state().device_world
mode = "cpu" if dw is None else (f"{dw.mode}{'+multicast' if dw.has_multicast else ''}" if dw.native else "nccl-fallback")
if r == 0:
    print("SELFTEST_OK", mode, flush=True)
fm.Finalize()
"""


class Report:
    def __init__(self):
        self.failed = False
        self.rows = []

    def add(self, status: str, what: str, detail: str = ""):
        self.failed |= status == "FAIL"
        self.rows.append({"status": status, "check": what, "detail": detail})
        print(f"[{status:>4}] {what}" + (f": {detail}" if detail else ""), flush=True)


def _self_test(rep: Report, what: str, nprocs: int, cpu_only: bool = False, **kw):
    from .launch import launch
    with tempfile.NamedTemporaryFile("w", suffix=".py", delete=False) as f:
        f.write(SELF_TEST)
        script = f.name
    try:
        env = {"PYTHONPATH": str(ROOT) + os.pathsep + os.environ.get("PYTHONPATH", ""), "FMPI_TIMEOUT_S": "60"}
        if cpu_only:
            env["CUDA_VISIBLE_DEVICES"] = ""          # the host paths are checked without touching a GPU
        code, outs = launch(nprocs, [script], timeout=180, env_extra=env, capture=True, **kw)
        line = next((ln for ln in outs[0].splitlines() if ln.startswith("SELFTEST_OK")), "")
        if code == 0 and line:
            rep.add("ok", what, line.split(" ", 1)[1])
        else:
            rep.add("FAIL", what, f"exit code {code}: " + outs[0].strip().splitlines()[-1][:200] if outs[0].strip() else f"exit code {code}")
    finally:
        os.unlink(script)


def _device_selftest(rep: Report, nprocs: int):
    """fluxmpi_b200.selftest over all GPUs of the box: every kernel family through the live world
    against float64 host references (what bench.py runs before timing)."""
    from .launch import launch
    env = {"PYTHONPATH": str(ROOT) + os.pathsep + os.environ.get("PYTHONPATH", ""), "FMPI_TIMEOUT_S": "60"}
    code, outs = launch(nprocs, ["fluxmpi_b200.selftest"], module=True, timeout=300, env_extra=env, capture=True)
    line = next((ln for ln in reversed(outs[0].splitlines()) if ln.startswith("{")), "")
    try:
        got = json.loads(line) if line else {}
    except ValueError:
        got = {}
    what = f"device self-test, {nprocs} GPUs"
    if code == 0 and got.get("ok"):
        rep.add("ok", what, f"{got.get('cases')} cases, mode {got.get('mode')}{'+multicast' if got.get('multicast') else ''}")
    else:
        rep.add("FAIL", what, f"exit code {code}; failed: {got.get('failed')}")


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="python -m fluxmpi_b200.doctor")
    ap.add_argument("--json", action="store_true", help="print the report as one JSON object at the end")
    ap.add_argument("--no-selftest", action="store_true", help="skip the launcher self-tests")
    a = ap.parse_args(argv)
    rep = Report()
    import torch
    from . import _build
    rep.add("ok", "python / torch", f"{sys.version.split()[0]} / {torch.__version__}")
    rep.add("ok" if shutil.which("g++") else "FAIL", "g++ (host library)", shutil.which("g++") or "not found")
    nvcc = _build._nvcc()
    rep.add("ok" if nvcc else "warn", "nvcc (CUDA library)", nvcc or "not found: prebuilt libraries only")
    try:
        rep.add("ok", "libfmpi_host.so", str(_build.build_host()))
    except Exception as e:                                       # noqa: BLE001
        rep.add("FAIL", "libfmpi_host.so", str(e)[:200])
    try:
        lib = _build.build_cuda()
        rep.add("ok" if lib and Path(lib).exists() else "warn", "libfmpi_cuda.so", str(lib) if lib else "not built (no nvcc)")
    except Exception as e:                                       # noqa: BLE001
        rep.add("FAIL", "libfmpi_cuda.so", str(e)[:200])
    try:
        import torch.distributed as dist
        rep.add("ok", "torch.distributed backends", ", ".join(b for b in ("nccl", "gloo", "mpi") if getattr(dist, f"is_{b}_available")()))
    except Exception as e:                                       # noqa: BLE001
        rep.add("warn", "torch.distributed", str(e)[:200])

    ngpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    rep.add("ok" if ngpu else "warn", "GPUs", f"{ngpu} x {torch.cuda.get_device_name(0)}" if ngpu else "none: CPU tier only")
    if ngpu:
        cap = torch.cuda.get_device_capability(0)
        rep.add("ok" if cap[0] == 10 else "warn", "compute capability", f"{cap[0]}.{cap[1]} (kernels are built for sm_100a)")
        try:
            import ctypes as C
            from . import _native
            buf = C.create_string_buffer(1 << 16)
            _native.cuda_check(_native.cuda().fmpi_topo_json(buf, len(buf)), "topology")
            topo = json.loads(buf.value.decode())
            devs = topo.get("devices", [])
            mc = all(d.get("multicast") for d in devs) if devs else False
            links = [d.get("nvlink_active", d.get("nvlinks")) for d in devs]
            rep.add("ok", "topology", f"{len(devs)} devices, NVLink links per GPU {links[:1]}, multicast {'yes' if mc else 'no'}")
        except Exception as e:                                   # noqa: BLE001
            rep.add("warn", "topology", str(e)[:200])

    if not a.no_selftest:
        _self_test(rep, "host world, 2 ranks (shared memory)", 2, cpu_only=True)
        _self_test(rep, "two pretend nodes x 1 rank (TCP mesh)", 2, cpu_only=True, sim_nodes=2)
        if ngpu >= 2:
            os.environ.pop("FMPI_FORCE_FALLBACK", None)
            _self_test(rep, "device world, 2 GPUs", 2)
            _device_selftest(rep, min(ngpu, 8))
    print("doctor:", "problems found" if rep.failed else "all required checks passed")
    if a.json:
        print(json.dumps({"failed": rep.failed, "checks": rep.rows}))
    return 1 if rep.failed else 0


if __name__ == "__main__":
    sys.exit(main())
