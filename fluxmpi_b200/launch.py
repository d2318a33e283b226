"""SPMD launcher: ``python -m fluxmpi_b200.launch -n N script.py [args…]``.

Plays the role of ``mpiexecjl -n <np> julia --project=. file.jl`` in the reference
(/root/reference/README.md:72, docs/src/guide.md:21): start N copies of one script, one per
GPU, each told who it is through ``RANK`` / ``WORLD_SIZE`` / ``LOCAL_RANK`` / ``MASTER_ADDR`` /
``MASTER_PORT`` (the same variables ``torchrun`` sets, so scripts run under either).

Failure handling the reference leaves to libmpi: if any rank exits non-zero the launcher
terminates the others (no orphaned ranks spinning in a barrier) and returns that exit code;
``--timeout`` bounds the whole job.
"""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import time
import uuid


def _free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch(nprocs: int, argv: list[str], timeout: float | None = None, module: bool = False,
           env_extra: dict | None = None, capture: bool = False):
    """Run ``argv`` N-way. Returns (exit_code, [stdout per rank] or None)."""
    port = _free_port()
    job = uuid.uuid4().hex[:12]
    procs = []
    for r in range(nprocs):
        env = dict(os.environ)
        env.update({
            "RANK": str(r), "WORLD_SIZE": str(nprocs), "LOCAL_RANK": str(r),
            "LOCAL_WORLD_SIZE": str(nprocs), "MASTER_ADDR": "127.0.0.1",
            "MASTER_PORT": str(port), "FMPI_JOB_ID": job,
        })
        env.setdefault("OMP_NUM_THREADS", str(max(1, (os.cpu_count() or 1) // nprocs)))
        if env_extra:
            env.update({k: str(v) for k, v in env_extra.items()})
        cmd = [sys.executable, "-u"] + (["-m"] if module else []) + list(argv)
        out = subprocess.PIPE if capture else None
        procs.append(subprocess.Popen(cmd, env=env, stdout=out,
                                      stderr=subprocess.STDOUT if capture else None,
                                      start_new_session=True))
    deadline = time.time() + timeout if timeout else None
    code = 0
    outputs = [b""] * nprocs if capture else None
    alive = set(range(nprocs))
    try:
        if capture:
            # drain pipes without deadlocking: communicate() per process in turn is fine because
            # ranks only block on each other for bounded time and pipes are drained in order.
            import threading

            def drain(i):
                outputs[i] = procs[i].stdout.read()

            threads = [threading.Thread(target=drain, args=(i,), daemon=True) for i in range(nprocs)]
            for t in threads:
                t.start()
        while alive:
            for r in list(alive):
                rc = procs[r].poll()
                if rc is not None:
                    alive.discard(r)
                    if rc != 0 and code == 0:
                        code = rc
            if code != 0 and alive:
                _terminate(procs, alive)
                break
            if deadline and time.time() > deadline and alive:
                code = 124
                _terminate(procs, alive)
                break
            time.sleep(0.02)
        if capture:
            for t in threads:
                t.join(timeout=5)
    finally:
        _terminate(procs, {r for r in range(nprocs) if procs[r].poll() is None})
    return code, ([o.decode(errors="replace") for o in outputs] if capture else None)


def _terminate(procs, which):
    for r in which:
        try:
            os.killpg(procs[r].pid, signal.SIGTERM)
        except (ProcessLookupError, PermissionError):
            pass
    t0 = time.time()
    while time.time() - t0 < 3 and any(procs[r].poll() is None for r in which):
        time.sleep(0.05)
    for r in which:
        if procs[r].poll() is None:
            try:
                os.killpg(procs[r].pid, signal.SIGKILL)
            except (ProcessLookupError, PermissionError):
                pass


def main(argv=None):
    ap = argparse.ArgumentParser(prog="python -m fluxmpi_b200.launch",
                                 description="start N SPMD ranks of a script")
    ap.add_argument("-n", "--nprocs", type=int, required=True)
    ap.add_argument("--timeout", type=float, default=None, help="seconds for the whole job")
    ap.add_argument("-m", dest="module", action="store_true", help="run a module (python -m)")
    ap.add_argument("script")
    ap.add_argument("args", nargs=argparse.REMAINDER)
    ns = ap.parse_args(argv)
    code, _ = launch(ns.nprocs, [ns.script] + ns.args, timeout=ns.timeout, module=ns.module)
    sys.exit(code)


if __name__ == "__main__":
    main()
