"""Real multi-GPU runs (one process per GPU) of the device collectives and the engine.
Skipped unless the box has at least two GPUs (`gpurun --gpus N`)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
    pytest.skip("needs at least two CUDA devices", allow_module_level=True)

from conftest import run_spmd

NGPU = torch.cuda.device_count()
# the smallest and the largest world the box allows (2 and 8 on a full node; 3 is covered in
# single-GPU emulation and on the CPU backend) — keeps the tier's wall time bounded
WORLDS = sorted({2, NGPU} & set(range(2, NGPU + 1)))


def _run(script, w, **kw):
    code, outs = run_spmd(script, w, timeout=600, **kw)
    joined = "\n".join(f"--- rank {i} ---\n{o[-3000:]}" for i, o in enumerate(outs))
    assert code == 0, joined
    for r, o in enumerate(outs):
        assert f"RANK_OK {r}" in o, joined


@pytest.mark.parametrize("w", WORLDS)
def test_device_collectives(w):
    _run("gpu_collectives.py", w)


@pytest.mark.parametrize("w", WORLDS)
def test_engine_replicas_and_reference(w):
    _run("gpu_engine.py", w)


def test_ipc_mode_and_no_multicast():
    _run("gpu_collectives.py", 2, env={"FMPI_SYMM_MODE": "ipc"}, args=("quick",))
    _run("gpu_collectives.py", 2, env={"FMPI_DISABLE_MULTICAST": "1"}, args=("quick",))


def test_nccl_fallback_when_p2p_disabled():
    _run("gpu_collectives.py", 2, env={"FMPI_DISABLE_P2P": "1"}, args=("quick",))


def test_engine_without_peer_mappings_over_nccl():
    """profiles/runs/fallback_engine_2gpu.log"""
    _run("gpu_engine_fallback.py", 2, env={"FMPI_FORCE_FALLBACK": "1"})


def test_pretend_nodes_device_tensors_take_nccl():
    """2 pretend nodes x 1 GPU: host world over the TCP mesh, device collectives over NCCL."""
    _run("gpu_collectives.py", 2, args=("quick",), sim_nodes=2)


def test_lost_rank_ends_in_an_error_not_a_spinning_gpu():
    """Device-side failure detection: bounded waits + the health page (csrc/kernels/common.cuh)."""
    import time
    t0 = time.time()
    code, outs = run_spmd("gpu_die.py", 2, timeout=120, env={"FMPI_DEVICE_TIMEOUT_S": "3", "FMPI_TIMEOUT_S": "20"})
    joined = "\n".join(outs)
    assert code == 7, joined
    assert "SURVIVOR_ERROR rank 0" in joined and time.time() - t0 < 90, joined


@pytest.mark.skipif(NGPU < 4, reason="rank groups on device tensors need four GPUs")
def test_rank_groups_on_device_tensors():
    _run("groups.py", 4, args=("cuda",))


def test_selftest_through_the_live_world():
    """fluxmpi_b200.selftest (what bench.py runs before timing at N >= 2) at the largest world."""
    from fluxmpi_b200.launch import launch
    import json, os, sys
    from conftest import ROOT
    env = {"PYTHONPATH": str(ROOT) + os.pathsep + os.environ.get("PYTHONPATH", ""), "FMPI_TIMEOUT_S": "60"}
    code, outs = launch(NGPU, ["fluxmpi_b200.selftest"], module=True, timeout=600, env_extra=env, capture=True)
    assert code == 0, "\n".join(o[-2000:] for o in outs)
    rep = json.loads([ln for ln in outs[0].splitlines() if ln.startswith("{")][-1])
    assert rep["ok"] and rep["cases"] >= 20 and rep["world"] == NGPU, rep
