"""The C ABI driven from plain C (csrc/abi_test.c) — the contract the Julia binding's ccalls
rely on, checked without Julia. Also checks that include/fmpi.h is valid C11."""
import os
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "fluxmpi_b200" / "_lib"


def _free_port():
    """Base of a run of free ports: a multi-node host world uses one port per local rank."""
    from fluxmpi_b200.launch import _free_port_range
    return _free_port_range(8)


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


def test_host_library_under_address_and_ub_sanitizers(tmp_path):
    """The reference has no sanitizer runs. The whole host library (shared-memory
    world, leader mesh, rule loops) is rebuilt with ASan + UBSan and driven by the C test over
    2 pretend nodes x 2 ranks; any report fails the run."""
    csrc = ROOT / "csrc"
    lib = tmp_path / "libfmpi_host.so"
    flags = ["-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=undefined",
             "-fno-omit-frame-pointer"]
    r = subprocess.run(["g++", *flags, "-ffp-contract=off", "-fno-math-errno", "-std=c++17", "-fPIC", "-shared",
                        "-pthread", f"-I{ROOT / 'include'}", *[str(csrc / s) for s in
                        ("host_backend.cpp", "fdx.cpp", "net_backend.cpp", "host_rules.cpp")],
                        "-o", str(lib), "-lrt"], capture_output=True, text=True)
    if r.returncode != 0:
        pytest.skip("this toolchain cannot build with -fsanitize=address,undefined: " + r.stderr[-300:])
    exe = tmp_path / "abi_test"
    subprocess.run(["gcc", "-std=c11", *flags, f"-I{ROOT / 'include'}", str(csrc / "abi_test.c"),
                    f"-L{tmp_path}", "-lfmpi_host", "-o", str(exe)], check=True)
    env = dict(os.environ, LD_LIBRARY_PATH=str(tmp_path), ASAN_OPTIONS="detect_leaks=1:abort_on_error=0")
    for world, nodes in ((3, 1), (4, 2)):
        out = subprocess.run([str(exe), str(world), str(nodes), str(_free_port())], env=env,
                             capture_output=True, text=True, timeout=180)
        assert out.returncode == 0 and f"ABI_OK world={world}" in out.stdout, out.stdout + out.stderr
        assert "ERROR: AddressSanitizer" not in out.stderr and "runtime error" not in out.stderr, out.stderr


def test_header_is_valid_c_with_device_half(tmp_path):
    src = tmp_path / "hdr.c"
    src.write_text('#include "fmpi.h"\nint main(void){ fmpi_devcomm c; fmpi_hyper h; fmpi_launch l; '
                   '(void)c; (void)h; (void)l; return sizeof(fmpi_hyper) == 48 ? 0 : 1; }\n')
    exe = tmp_path / "hdr"
    subprocess.run(["gcc", "-std=c11", "-Wall", "-Werror", f"-I{ROOT / 'include'}", str(src), "-o", str(exe)], check=True)
    assert subprocess.run([str(exe)]).returncode == 0
