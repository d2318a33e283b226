"""The C ABI driven from plain C (csrc/abi_test.c) — the contract the Julia binding's ccalls
rely on, checked without Julia. Also checks that include/fmpi.h is valid C11."""
import os
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "fluxmpi_b200" / "_lib"


def _free_port():
    import socket
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


# (world, nodes): nodes > 1 goes through fmpi_host_init_multi (shared memory + TCP leader mesh)
@pytest.mark.parametrize("world,nodes", [(1, 1), (3, 1), (4, 1), (4, 2), (6, 3), (3, 3)])
def test_host_abi_from_c(tmp_path, world, nodes):
    from fluxmpi_b200 import _build
    _build.build_host()
    exe = tmp_path / "abi_test"
    subprocess.run(["gcc", "-std=c11", "-O2", "-Wall", "-Werror", f"-I{ROOT / 'include'}",
                    str(ROOT / "csrc" / "abi_test.c"), f"-L{LIB}", "-lfmpi_host", "-o", str(exe)], check=True)
    env = dict(os.environ, LD_LIBRARY_PATH=f"{LIB}:{os.environ.get('LD_LIBRARY_PATH', '')}")
    out = subprocess.run([str(exe), str(world), str(nodes), str(_free_port())], env=env,
                         capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    assert f"ABI_OK world={world}" in out.stdout


def test_header_is_valid_c_with_device_half(tmp_path):
    src = tmp_path / "hdr.c"
    src.write_text('#include "fmpi.h"\nint main(void){ fmpi_devcomm c; fmpi_hyper h; fmpi_launch l; '
                   '(void)c; (void)h; (void)l; return sizeof(fmpi_hyper) == 48 ? 0 : 1; }\n')
    exe = tmp_path / "hdr"
    subprocess.run(["gcc", "-std=c11", "-Wall", "-Werror", f"-I{ROOT / 'include'}", str(src), "-o", str(exe)], check=True)
    assert subprocess.run([str(exe)]).returncode == 0
