"""The engine without peer mappings (what a job over several nodes, or one with P2P switched
off, gets): flat buffers in ordinary memory, NCCL per bucket, local fused rule kernel. One GPU is
enough to check the local half; test_gpu_multi.py runs the same script over NCCL."""
import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

from conftest import run_spmd


def test_engine_without_peer_mappings_one_gpu():
    code, outs = run_spmd("gpu_engine_fallback.py", 1, timeout=600, env={"FMPI_FORCE_FALLBACK": "1"})
    assert code == 0 and "RANK_OK 0" in outs[0], outs[0][-4000:]
