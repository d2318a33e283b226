"""N-way SPMD runs of tests/spmd/*.py on the host (POSIX-shm) backend, at the world sizes the
reference's harness uses (2..4; /root/reference/test/runtests.jl:3-4) including the odd one.

One launch per world size runs all scripts back to back (tests/spmd/_suite.py); the per-script
tests below read that launch's transcript."""
import functools

import pytest

from conftest import run_spmd

SCRIPTS = ["common.py", "data.py", "mpi_extensions.py", "optimizer.py", "synchronize.py"]


@functools.lru_cache(maxsize=None)
def suite(world: int):
    extra = ("quickstart.py:20",) if world == 3 else ()
    code, outs = run_spmd("_suite.py", world, timeout=600, args=(*SCRIPTS, *extra))
    return code, outs


def _check(world, script):
    code, outs = suite(world)
    for r, o in enumerate(outs):
        if f"SCRIPT_OK {script} rank {r}" not in o:
            joined = "\n".join(f"--- rank {i} ---\n{t}" for i, t in enumerate(outs))
            pytest.fail(f"{script} failed at W={world} (suite exit {code})\n{joined[-6000:]}")


@pytest.mark.parametrize("world", [2, 3, 4])
@pytest.mark.parametrize("script", SCRIPTS)
def test_spmd(script, world):
    _check(world, script)


def test_single_worker():
    code, outs = run_spmd("_suite.py", 1, args=tuple(SCRIPTS))
    assert code == 0, outs[0][-4000:]


def test_quickstart_mlp_cpu_3_ranks():
    """BASELINE config 1: Quick-Start MLP + DistributedOptimizer(Adam), CPU, 3 ranks."""
    _check(3, "quickstart.py")
    _, outs = suite(3)
    # rank-ordered printing: rank 0's stream shows its own "[0 / 3]" lines only
    assert "[0 / 3] Epoch 1: Loss" in outs[0] and "[1 / 3] Epoch" not in outs[0]


def test_println_rank_order(spmd, tmp_path):
    """Output of fluxmpi_println is serialised in rank order (src/common.jl:86-92)."""
    log = tmp_path / "order.log"
    spmd("print_order.py", 4, args=(str(log),))
    ranks = [int(line.split()[0]) for line in log.read_text().split("\n") if line.strip()]
    assert ranks == [0, 1, 2, 3] * 3, ranks


def test_checkpoint_roundtrip_and_trace(spmd, tmp_path):
    """Aux subsystems the reference lacks: checkpoint / resume, tracing."""
    outs = spmd("checkpoint.py", 3, args=(str(tmp_path / "ckpt.pt"),))
    assert all(f"RANK_OK {r}" in o for r, o in enumerate(outs))


def test_job_restarts_and_resumes_from_checkpoint(tmp_path):
    """launch --max-restarts: the job that lost a rank is started again as a whole and the
    script continues from its last checkpoint (tests/spmd/restart.py asserts where it resumed)."""
    import os
    from conftest import ROOT
    from fluxmpi_b200.launch import run_with_restarts
    env = {"PYTHONPATH": str(ROOT) + os.pathsep + os.environ.get("PYTHONPATH", ""), "FMPI_TIMEOUT_S": "60"}
    script = str(ROOT / "tests" / "spmd" / "restart.py")
    assert run_with_restarts(3, [script, str(tmp_path / "a.pt")], 0, timeout=120, env_extra=env) == 7
    assert run_with_restarts(3, [script, str(tmp_path / "b.pt")], 1, timeout=240, env_extra=env) == 0


def test_dead_rank_is_detected(spmd):
    """A rank that dies must not leave the others hanging (the reference has no such guard,
    SURVEY §5): the launcher kills the job and reports the failing exit code."""
    code, outs = run_spmd("die.py", 3, timeout=120)
    assert code == 3, (code, outs)


def test_randomised_collectives(spmd):
    """tests/spmd/stress.py on one node at the odd world size."""
    outs = spmd("stress.py", 3, timeout=600, args=("120",))
    assert all(f"RANK_OK {r}" in o for r, o in enumerate(outs))


def test_host_engine_matches_torch_optim(spmd):
    """HostDataParallelEngine: flat vectors, gradient in symmetric host memory, native rule loop."""
    outs = spmd("host_engine.py", 3, timeout=300)
    assert all(f"RANK_OK {r}" in o for r, o in enumerate(outs))


def test_examples_run_on_cpu(tmp_path):
    """examples/ stay runnable: the any-device engine example (with checkpoint + resume) and the
    functional quick start, two ranks each."""
    import os
    from conftest import ROOT
    from fluxmpi_b200.launch import launch
    env = {"PYTHONPATH": str(ROOT) + os.pathsep + os.environ.get("PYTHONPATH", ""), "FMPI_TIMEOUT_S": "60",
           "OMP_NUM_THREADS": "1"}
    ck = str(tmp_path / "ck.pt")
    ex = str(ROOT / "examples" / "engine_any_device.py")
    code, outs = launch(2, [ex, "--epochs", "100", "--checkpoint", ck], timeout=240, env_extra=env, capture=True)
    assert code == 0 and "HostDataParallelEngine" in outs[0], outs[0][-3000:]
    code, outs = launch(2, [ex, "--epochs", "150", "--checkpoint", ck], timeout=240, env_extra=env, capture=True)
    assert code == 0 and "resumed at epoch 100" in outs[0], outs[0][-3000:]
    for name, n in (("quickstart_mlp.py", 2), ("flux_style_module.py", 2), ("rank_groups.py", 4)):
        code, outs = launch(n, [str(ROOT / "examples" / name)], timeout=240, env_extra=env, capture=True)
        assert code == 0, (name, outs[0][-3000:])


def test_doctor_passes_on_this_machine(capsys):
    """python -m fluxmpi_b200.doctor: toolchain, libraries, a 2-rank host self-test and a
    2-pretend-node self-test; required checks must pass on the CPU box."""
    import json
    from fluxmpi_b200 import doctor
    assert doctor.main(["--json"]) == 0
    report = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    checks = {c["check"]: c["status"] for c in report["checks"]}
    assert checks["libfmpi_host.so"] == "ok" and checks["host world, 2 ranks (shared memory)"] == "ok"
    assert checks["two pretend nodes x 1 rank (TCP mesh)"] == "ok" and not report["failed"]


def test_rank_groups_run_different_collectives_concurrently(spmd):
    """fm.Group: {0,1} and {2,3} at the same time, a 3-of-4 group, then the world (the `comm`
    argument of /root/reference/src/mpi_extensions.jl:97,119,141)."""
    outs = spmd("groups.py", 4)
    assert all(f"RANK_OK {r}" in o for r, o in enumerate(outs))


@pytest.mark.parametrize("w", [1, 3])
def test_group_argument_checks_and_selftest_skip(spmd, w):
    outs = spmd("extras.py", w)
    assert all(f"RANK_OK {r}" in o for r, o in enumerate(outs))


def test_randomised_groups_and_world_interleaved(spmd):
    """tests/spmd/stress_groups.py: three overlapping groups and the world taking turns."""
    outs = spmd("stress_groups.py", 4, timeout=600, args=("150",))
    assert all(f"RANK_OK {r}" in o for r, o in enumerate(outs))
