"""Host half of the tcgen05 GEMM library (csrc/kernels/gemm_bnstats.cu) on a machine without a
GPU: the library is loaded against the toolkit's link-time stub of libcuda, which is enough for
everything CUTLASS decides on the host — argument plumbing, alignment and shape checks
(``can_implement``), workspace size. Launching needs a real driver; that part is
tests/test_gpu_conv_bn.py."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
PROBE = r"""
import ctypes as C, json, sys
lib = C.CDLL(sys.argv[1])
lib.fmpi_conv1x1_bnstats_workspace.restype = C.c_size_t
lib.fmpi_conv1x1_bnstats_workspace.argtypes = [C.c_int] * 3
lib.fmpi_gemm_last_error.restype = C.c_char_p
lib.fmpi_conv1x1_bnstats.argtypes = [C.c_void_p] * 5 + [C.c_int] * 3 + [C.c_void_p, C.c_size_t, C.c_void_p]
lib.fmpi_conv1x1_bnstats_variant.argtypes = [C.c_int]
assert lib.fmpi_conv1x1_bnstats_variant(int(sys.argv[3])) == 0
assert lib.fmpi_conv1x1_bnstats_variant(7) == 1            # unknown variant refused, selection kept
out = []
for M, N, K in json.loads(sys.argv[2]):
    ws = lib.fmpi_conv1x1_bnstats_workspace(M, N, K)
    rc = lib.fmpi_conv1x1_bnstats(4096, 8192, 12288, 16384, 20480, M, N, K, None, 0, None)
    out.append([ws, rc, lib.fmpi_gemm_last_error().decode()])
print(json.dumps(out))
"""


@pytest.mark.parametrize("variant", [0, 1], ids=["1sm", "2sm"])
def test_host_side_checks_with_a_stub_driver(tmp_path, variant):
    import json
    from fluxmpi_b200 import _build
    lib = _build.build_gemm()
    if lib is None or not lib.exists():
        pytest.skip("libfmpi_gemm.so is not built here (no nvcc or no CUTLASS headers)")
    nvcc = _build._nvcc()
    stub = Path(nvcc).resolve().parent.parent / "lib64" / "stubs" / "libcuda.so" if nvcc else None
    if stub is None or not stub.exists():
        pytest.skip("no libcuda stub in this toolkit")
    (tmp_path / "libcuda.so.1").symlink_to(stub)
    # every stride-1 1x1 convolution of ResNet-50 at batch 256, a ragged M, a K below the tile,
    # and two shapes the entry point must refuse before CUTLASS sees them
    shapes = [(256 * hw * hw, cout, cin) for hw, cin, cout in
              [(56, 64, 64), (56, 256, 64), (56, 64, 256), (56, 256, 128), (28, 512, 128), (28, 128, 512),
               (28, 512, 256), (14, 1024, 256), (14, 256, 1024), (14, 1024, 512), (7, 2048, 512), (7, 512, 2048)]]
    shapes += [(147, 64, 64), (128, 64, 16), (100, 60, 64), (100, 64, 12)]
    env = dict(os.environ, LD_LIBRARY_PATH=f"{tmp_path}:{os.environ.get('LD_LIBRARY_PATH', '')}")
    r = subprocess.run([sys.executable, "-c", PROBE, str(lib), json.dumps(shapes), str(variant)], env=env,
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    res = json.loads(r.stdout.strip().splitlines()[-1])
    for (M, N, K), (ws, rc, msg) in zip(shapes[:-2], res[:-2]):
        assert ws == 0, (M, N, K, ws)                       # atomics: no reduction workspace
        # accepted by can_implement; the failure must be the launch preparation, which needs a driver
        assert rc == 1 and "initialize" in msg and "can_implement" not in msg, (M, N, K, msg)
    for (M, N, K), (ws, rc, msg) in zip(shapes[-2:], res[-2:]):
        assert rc == 1 and "multiples of 8" in msg, (M, N, K, msg)
