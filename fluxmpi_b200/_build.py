"""Build the two native libraries in-tree (``fluxmpi_b200/_lib/*.so``).

* ``libfmpi_host.so`` — g++ only, no CUDA: POSIX-shm host collectives + fd exchange.
* ``libfmpi_cuda.so`` — nvcc, ``-gencode arch=compute_100a,code=sm_100a -lineinfo``: every
  CUDA kernel plus the symmetric-heap / topology runtime. cudart is linked statically and the
  driver API is reached through ``cudaGetDriverEntryPoint``, so the library loads on a box
  without a GPU (the CPU build check) and nvcc cross-compiles it there.

The built objects are git-ignored but travel to the GPU box with the gpurun snapshot.
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
INC = ROOT / "include"
LIB = Path(__file__).resolve().parent / "_lib"

HOST_SOURCES = ["host_backend.cpp", "fdx.cpp"]
CUDA_SOURCES = [
    "symm.cpp",
    "topo.cpp",
    "kernels/collectives.cu",
    "kernels/fused_step.cu",
    "kernels/multi_tensor.cu",
    "kernels/reduce_diag.cu",
]

NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _nvcc() -> str | None:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    return None


def _digest(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(str(x) for x in paths):
        h.update(p.encode())
        h.update(Path(p).read_bytes())
    return h.hexdigest()


def _stale(out: Path, srcs, extra=()) -> bool:
    stamp = out.with_suffix(out.suffix + ".stamp")
    if not out.exists() or not stamp.exists():
        return True
    return stamp.read_text().strip() != _digest(list(srcs) + list(extra))


def _mark(out: Path, srcs, extra=()):
    out.with_suffix(out.suffix + ".stamp").write_text(_digest(list(srcs) + list(extra)))


def _headers():
    return sorted(INC.glob("*.h")) + sorted((CSRC / "kernels").glob("*.cuh")) + sorted(CSRC.glob("*.h"))


def build_host(force: bool = False, verbose: bool = False) -> Path:
    LIB.mkdir(exist_ok=True)
    out = LIB / "libfmpi_host.so"
    srcs = [CSRC / s for s in HOST_SOURCES]
    if not force and not _stale(out, srcs, _headers()):
        return out
    cmd = ["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-Wall", "-pthread",
           f"-I{INC}", *map(str, srcs), "-o", str(out), "-lrt"]
    if verbose:
        print(" ".join(cmd), file=sys.stderr)
    subprocess.run(cmd, check=True)
    _mark(out, srcs, _headers())
    return out


def build_cuda(force: bool = False, verbose: bool = False) -> Path | None:
    """Returns the library path, or None when there is no nvcc on this machine."""
    LIB.mkdir(exist_ok=True)
    out = LIB / "libfmpi_cuda.so"
    srcs = [CSRC / s for s in CUDA_SOURCES if (CSRC / s).exists()]
    if not force and not _stale(out, srcs, _headers()):
        return out
    nvcc = _nvcc()
    if nvcc is None:
        return out if out.exists() else None
    objs = []
    objdir = LIB / "obj"
    objdir.mkdir(exist_ok=True)
    common = [nvcc, "-O3", "-std=c++17", *NVCC_ARCH, "-lineinfo", "-Xcompiler", "-fPIC",
              f"-I{INC}", f"-I{CSRC}", "--expt-relaxed-constexpr"]
    procs = []
    for s in srcs:
        o = objdir / (s.name + ".o")
        cmd = [*common, "-x", "cu", "-c", str(s), "-o", str(o)]
        if os.environ.get("FMPI_PTXAS_V"):
            cmd += ["-Xptxas", "-v"]
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        procs.append((subprocess.Popen(cmd), cmd))
        objs.append(o)
    for p, cmd in procs:
        if p.wait() != 0:
            raise subprocess.CalledProcessError(p.returncode, cmd)
    link = [nvcc, "-shared", *NVCC_ARCH, "-o", str(out), *map(str, objs), "-cudart", "static",
            "-lpthread", "-ldl", "-lrt"]
    if verbose:
        print(" ".join(link), file=sys.stderr)
    subprocess.run(link, check=True)
    _mark(out, srcs, _headers())
    return out


def build_all(force: bool = False, verbose: bool = False):
    return build_host(force, verbose), build_cuda(force, verbose)


if __name__ == "__main__":
    h, c = build_all(force="--force" in sys.argv, verbose=True)
    print("host:", h)
    print("cuda:", c)
