"""A job over several nodes: shared memory inside a node, the TCP leader mesh between nodes
(csrc/net_backend.cpp). The reference gets this from ``mpiexec`` and never tests
it; here the ranks of one machine are split into pretend nodes (``launch --sim-nodes``), or two
launchers are started the way two real nodes would start them."""
import functools
import os
import subprocess
import sys

import pytest

from conftest import ROOT, run_spmd

SCRIPTS = ["common.py", "data.py", "mpi_extensions.py", "optimizer.py", "synchronize.py"]


@functools.lru_cache(maxsize=None)
def suite(world: int, nodes: int):
    extra = ("quickstart.py:10",) if (world, nodes) == (4, 2) else ()
    return run_spmd("_suite.py", world, timeout=600, args=(*SCRIPTS, *extra), sim_nodes=nodes)


# 2 nodes x 2 ranks (both levels), 3 nodes x 1 rank (no shared memory at all, odd node count)
@pytest.mark.parametrize("world,nodes", [(4, 2), (3, 3)])
@pytest.mark.parametrize("script", SCRIPTS)
def test_spmd_over_nodes(script, world, nodes):
    code, outs = suite(world, nodes)
    for r, o in enumerate(outs):
        if f"SCRIPT_OK {script} rank {r}" not in o:
            joined = "\n".join(f"--- rank {i} ---\n{t}" for i, t in enumerate(outs))
            pytest.fail(f"{script} failed at W={world} over {nodes} nodes (exit {code})\n{joined[-6000:]}")


def test_quickstart_over_two_nodes():
    code, outs = suite(4, 2)
    assert all(f"SCRIPT_OK quickstart.py rank {r}" in o for r, o in enumerate(outs)), outs[0][-3000:]


def _free_port():
    """MASTER_PORT with the ports behind it free too (rails use MASTER_PORT + 1 ...)."""
    from fluxmpi_b200.launch import _free_port_range
    return _free_port_range(8)


def _two_launchers(script, per_node, args=(), timeout=300):
    """One launcher per node, as on a real cluster (same port and job id, own --node-rank)."""
    port = _free_port()
    env = dict(os.environ)
    env["PYTHONPATH"] = str(ROOT) + os.pathsep + env.get("PYTHONPATH", "")
    env["FMPI_TIMEOUT_S"] = "60"
    procs = [subprocess.Popen(
        [sys.executable, "-m", "fluxmpi_b200.launch", "-n", str(per_node), "--nnodes", "2",
         "--node-rank", str(k), "--master-addr", "127.0.0.1", "--master-port", str(port),
         "--timeout", str(timeout), str(ROOT / "tests" / "spmd" / script), *args],
        env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for k in range(2)]
    outs = [p.communicate(timeout=timeout + 30)[0] for p in procs]
    return [p.returncode for p in procs], outs


def test_two_launchers_form_one_job():
    codes, outs = _two_launchers("_suite.py", 2, args=("common.py", "mpi_extensions.py"))
    assert codes == [0, 0], "\n".join(outs)[-6000:]
    assert "SCRIPT_OK mpi_extensions.py rank 0" in outs[0] and "SCRIPT_OK mpi_extensions.py rank 3" in outs[1]


def test_lost_node_is_an_error_not_a_hang():
    codes, outs = _two_launchers("die_node.py", 2, timeout=120)
    assert codes == [0, 0], "\n".join(outs)[-4000:]
    assert outs[0].count("PEER_LOSS_DETECTED") == 2, outs[0][-4000:]


@pytest.mark.parametrize("world,nodes", [(4, 2), (6, 3)])
def test_randomised_collectives_over_nodes(world, nodes):
    """tests/spmd/stress.py: a seeded sequence of collectives with sizes around slot / slice /
    rail boundaries (some with fewer elements than ranks), each checked against an all-gather."""
    code, outs = run_spmd("stress.py", world, timeout=600, args=("120",), sim_nodes=nodes)
    assert code == 0 and all(f"RANK_OK {r}" in o for r, o in enumerate(outs)), "\n".join(outs)[-4000:]


def test_pipelined_all_reduce_over_nodes():
    """Large all-reduces of a multi-node job overlap their network step with the shared-memory
    steps of the neighbouring chunks (helper thread, two result slots). Forced here for every
    message of the randomised test and of the tree / dtype matrix (tiny slots = many chunks)."""
    env = {"FMPI_HOST_PIPELINE_MIN_BYTES": "1"}
    code, outs = run_spmd("stress.py", 4, timeout=600, args=("120",), sim_nodes=2, env=env)
    assert code == 0 and all(f"RANK_OK {r}" in o for r, o in enumerate(outs)), "\n".join(outs)[-4000:]
    code, outs = run_spmd("mpi_extensions.py", 6, timeout=600, sim_nodes=3, env=env)
    assert code == 0 and all(f"RANK_OK {r}" in o for r, o in enumerate(outs)), "\n".join(outs)[-4000:]


def test_host_engine_over_nodes():
    code, outs = run_spmd("host_engine.py", 4, timeout=300, sim_nodes=2)
    assert code == 0 and all(f"RANK_OK {r}" in o for r, o in enumerate(outs)), "\n".join(outs)[-4000:]
