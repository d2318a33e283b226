"""The device half of include/fmpi.h driven from plain C (csrc/abi_test_cuda.c): the call sequence
the Julia binding makes (julia/FluxMPIB200/src/device.jl, engine.jl), with no Python or ctypes in
between, every result checked against a host loop inside the C program."""
import os
import subprocess
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

ROOT = Path(__file__).resolve().parent.parent


def test_c_driver_of_the_device_abi(tmp_path):
    from fluxmpi_b200 import _build
    lib = _build.build_cuda()
    cuda_home = Path(os.environ.get("CUDA_HOME", "/usr/local/cuda"))
    exe = tmp_path / "abi_test_cuda"
    cmd = ["gcc", "-std=c11", "-O1", "-Wall", f"-I{ROOT / 'include'}", f"-I{cuda_home / 'include'}",
           str(ROOT / "csrc" / "abi_test_cuda.c"), f"-L{Path(lib).parent}", "-lfmpi_cuda",
           f"-L{cuda_home / 'lib64'}", "-lcudart", "-lm", "-o", str(exe)]
    subprocess.run(cmd, check=True)
    env = dict(os.environ)
    env["LD_LIBRARY_PATH"] = os.pathsep.join([str(Path(lib).parent), str(cuda_home / "lib64"),
                                              env.get("LD_LIBRARY_PATH", "")])
    out = subprocess.run([str(exe)], env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    last = out.stdout.strip().splitlines()[-1]
    assert last.startswith("ABI_CUDA_OK") and int(last.split()[1]) >= 40, out.stdout[-500:]
