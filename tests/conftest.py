"""Test harness.

The reference's tests are SPMD scripts launched N-way by ``mpiexec``
(/root/reference/test/runtests.jl:11-16); the assertions run inside every rank. Same design
here: ``tests/spmd/*.py`` are rank scripts, ``run_spmd`` launches them with our launcher and
asserts exit code 0. GPU tests carry ``@pytest.mark.gpu``.
"""
import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200)")
    from fluxmpi_b200 import _build
    _build.build_host()


def run_spmd(script: str, nprocs: int, timeout: float = 180.0, env=None, args=(), sim_nodes: int = 1):
    from fluxmpi_b200.launch import launch
    env_extra = {"PYTHONPATH": str(ROOT) + os.pathsep + os.environ.get("PYTHONPATH", ""),
                 "FMPI_TIMEOUT_S": "60"}
    if env:
        env_extra.update(env)
    code, outs = launch(nprocs, [str(ROOT / "tests" / "spmd" / script), *args], timeout=timeout,
                        env_extra=env_extra, capture=True, sim_nodes=sim_nodes)
    return code, outs


@pytest.fixture
def spmd():
    def _run(script, nprocs, **kw):
        code, outs = run_spmd(script, nprocs, **kw)
        if code != 0:
            joined = "\n".join(f"--- rank {i} ---\n{o}" for i, o in enumerate(outs))
            pytest.fail(f"{script} at W={nprocs} exited {code}\n{joined}")
        return outs
    return _run


@pytest.fixture(autouse=True)
def _device_waits_stay_healthy(request):
    """After every GPU test: no cross-rank wait of a kernel may have been abandoned (a protocol
    bug would otherwise pass as a slow test with wrong numbers, or hang without the bound)."""
    yield
    if request.node.get_closest_marker("gpu") is None:
        return
    import torch
    if not torch.cuda.is_available():
        return
    from fluxmpi_b200 import _native
    if _native.cuda_loaded():
        torch.cuda.synchronize()
        lib = _native.cuda()
        rc = lib.fmpi_health_check()
        if rc:
            msg = lib.fmpi_cuda_last_error().decode()
            lib.fmpi_health_reset()
            pytest.fail("device health page: " + msg)
