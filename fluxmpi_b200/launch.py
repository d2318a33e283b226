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
import shutil
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


def _free_port_range(k: int) -> int:
    """First port of k consecutive free ones (the rails of a multi-node host world use
    base .. base + ranks_per_node - 1)."""
    for _ in range(200):
        base = _free_port()
        if base + k > 65535:
            continue
        socks = []
        try:
            for p in range(base, base + k):
                s = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
                socks.append(s)
                s.bind(("0.0.0.0", p))
            return base
        except OSError:
            continue
        finally:
            for s in socks:
                s.close()
    raise RuntimeError(f"no {k} consecutive free ports found")


def _is_command(argv: list[str], module: bool) -> bool:
    """``julia --project script.jl`` and friends are started as they are; anything else is a
    Python script (or module) for this interpreter."""
    if module or not argv:
        return False
    head = argv[0]
    return not head.endswith(".py") and not os.path.isfile(head) and shutil.which(head) is not None


def launch(nprocs: int, argv: list[str], timeout: float | None = None, module: bool = False,
           env_extra: dict | None = None, capture: bool = False, nnodes: int = 1,
           node_rank: int = 0, master_addr: str = "127.0.0.1", master_port: int | None = None,
           job_id: str | None = None, sim_nodes: int = 1):
    """Run ``argv`` on ``nprocs`` ranks of this node. Returns (exit_code, [stdout per rank] or
    None).

    A job over several nodes starts one launcher per node with the same ``nnodes``,
    ``master_addr`` (node 0's address) and ``master_port`` and its own ``node_rank``; global
    ranks are node-major. ``sim_nodes=K`` splits the ranks of ONE machine into K pretend nodes
    (shared memory inside each, TCP between them) so the multi-node paths can be tested anywhere.
    """
    if nnodes > 1 and (master_port is None or sim_nodes > 1):
        raise ValueError("a multi-node job needs an explicit --master-port (the same on every node) "
                         "and cannot be combined with --sim-nodes")
    if sim_nodes < 1 or nprocs % sim_nodes:
        raise ValueError(f"--sim-nodes {sim_nodes} does not divide -n {nprocs}")
    port = master_port if master_port is not None else _free_port()
    job = job_id or (f"mn{port}" if nnodes > 1 else uuid.uuid4().hex[:12])
    lworld = nprocs // sim_nodes
    world = nprocs * nnodes
    as_command = _is_command(list(argv), module)
    # real jobs use MASTER_PORT + 1 .. MASTER_PORT + ranks_per_node (one rail per local rank)
    net_port = _free_port_range(nprocs // sim_nodes) if sim_nodes > 1 else None
    procs = []
    for r in range(nprocs):
        env = dict(os.environ)
        grank = node_rank * nprocs + r
        env.update({
            "RANK": str(grank), "WORLD_SIZE": str(world), "LOCAL_RANK": str(r % lworld),
            "LOCAL_WORLD_SIZE": str(lworld), "GROUP_RANK": str(grank // lworld),
            "MASTER_ADDR": master_addr, "MASTER_PORT": str(port), "FMPI_JOB_ID": job,
        })
        if sim_nodes > 1:
            env["FMPI_SIM_NODES"] = str(sim_nodes)
            env["FMPI_NET_PORT"] = str(net_port)
        env.setdefault("OMP_NUM_THREADS", str(max(1, (os.cpu_count() or 1) // nprocs)))
        if env_extra:
            env.update({k: str(v) for k, v in env_extra.items()})
        cmd = list(argv) if as_command else \
            [sys.executable, "-u"] + (["-m"] if module else []) + list(argv)
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
    ap.add_argument("-n", "--nprocs", type=int, required=True, help="ranks on this node (one per GPU)")
    ap.add_argument("--timeout", type=float, default=None, help="seconds for the whole job")
    ap.add_argument("-m", dest="module", action="store_true", help="run a module (python -m)")
    ap.add_argument("--nnodes", type=int, default=1)
    ap.add_argument("--node-rank", type=int, default=0)
    ap.add_argument("--master-addr", default="127.0.0.1", help="address of node 0")
    ap.add_argument("--master-port", type=int, default=None, help="same free port on every node")
    ap.add_argument("--job-id", default=None, help="same on every node (default: from the port)")
    ap.add_argument("--sim-nodes", type=int, default=1,
                    help="testing: treat the ranks of this machine as this many nodes")
    ap.add_argument("--max-restarts", type=int, default=0,
                    help="single-node jobs: start all ranks again (up to this many times) when one "
                         "fails; the script sees FMPI_RESTART_COUNT and resumes from its checkpoint")
    ap.add_argument("script", help="a .py file, a module with -m, or a command such as julia")
    ap.add_argument("args", nargs=argparse.REMAINDER)
    ns = ap.parse_args(argv)
    if ns.max_restarts and ns.nnodes > 1:
        ap.error("--max-restarts needs all ranks under one launcher (single node)")
    code = run_with_restarts(ns.nprocs, [ns.script] + ns.args, ns.max_restarts, timeout=ns.timeout,
                             module=ns.module, nnodes=ns.nnodes, node_rank=ns.node_rank,
                             master_addr=ns.master_addr, master_port=ns.master_port, job_id=ns.job_id,
                             sim_nodes=ns.sim_nodes)
    sys.exit(code)


def run_with_restarts(nprocs: int, argv: list[str], max_restarts: int = 0, **kw) -> int:
    """Failure handling one step beyond detection (the reference has neither): when
    a rank dies the job is torn down — that part is ``launch`` — and, if restarts are allowed,
    started again as a whole with ``FMPI_RESTART_COUNT`` = attempt number, so a script that
    checkpoints (``fm.checkpoint``) continues from its last save. A timeout (exit 124) is not
    retried."""
    extra = dict(kw.pop("env_extra", None) or {})
    code = 0
    for attempt in range(max_restarts + 1):
        extra["FMPI_RESTART_COUNT"] = attempt
        code, _ = launch(nprocs, argv, env_extra=extra, **kw)
        if code in (0, 124):
            break
        if attempt < max_restarts:
            print(f"[fluxmpi_b200.launch] job failed with exit code {code}; restart "
                  f"{attempt + 1} of {max_restarts}", file=sys.stderr, flush=True)
    return code


if __name__ == "__main__":
    main()
