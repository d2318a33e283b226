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

HOST_SOURCES = ["host_backend.cpp", "fdx.cpp", "net_backend.cpp", "host_rules.cpp"]
CUDA_SOURCES = [
    "symm.cpp",
    "topo.cpp",
    "kernels/collectives.cu",
    "kernels/fused_step.cu",
    "kernels/multi_tensor.cu",
    "kernels/tree_step.cu",
    "kernels/reduce_diag.cu",
    "kernels/batchnorm.cu",
    "kernels/pool.cu",
]

NVCC_ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _nvcc() -> str | None:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and Path(cand).exists():
            return cand
    return None


def _digest(paths) -> str:
    # repo-relative names: the tree is copied to a different absolute path on the GPU box and
    # the libraries built here must count as fresh there
    h = hashlib.sha256()
    for p in sorted(str(Path(x).resolve().relative_to(ROOT)) for x in paths):
        h.update(p.encode())
        h.update((ROOT / p).read_bytes())
    return h.hexdigest()


class _BuildLock:
    """One builder at a time (N ranks import the package concurrently)."""

    def __enter__(self):
        import fcntl
        LIB.mkdir(exist_ok=True)
        self.f = open(LIB / ".build.lock", "w")
        fcntl.flock(self.f, fcntl.LOCK_EX)
        return self

    def __exit__(self, *a):
        import fcntl
        fcntl.flock(self.f, fcntl.LOCK_UN)
        self.f.close()


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
    with _BuildLock():
        if not force and not _stale(out, srcs, _headers()):
            return out
        tmp = out.with_suffix(f".so.tmp{os.getpid()}")
        # -O3 for the loop vectoriser; no FMA contraction so every ISA clone of the rule loops
        # (csrc/host_rules.cpp) rounds the same way and replicas on different CPUs stay equal
        # (-fno-math-errno: sqrt without the errno branch, or the Adam loop stays scalar)
        cmd = ["g++", "-O3", "-ffp-contract=off", "-fno-math-errno", "-std=c++17", "-fPIC", "-shared", "-Wall", "-pthread",
               f"-I{INC}", *map(str, srcs), "-o", str(tmp), "-lrt"]
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        subprocess.run(cmd, check=True)
        os.replace(tmp, out)
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
    with _BuildLock():
        if not force and not _stale(out, srcs, _headers()):
            return out
        return _build_cuda_locked(nvcc, out, srcs, verbose)


def _build_cuda_locked(nvcc: str, out: Path, srcs, verbose: bool) -> Path:
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
    tmp = out.with_suffix(f".so.tmp{os.getpid()}")
    link = [nvcc, "-shared", *NVCC_ARCH, "-o", str(tmp), *map(str, objs), "-cudart", "static",
            "-lpthread", "-ldl", "-lrt"]
    if verbose:
        print(" ".join(link), file=sys.stderr)
    subprocess.run(link, check=True)
    os.replace(tmp, out)
    _mark(out, srcs, _headers())
    return out


def build_all(force: bool = False, verbose: bool = False):
    return build_host(force, verbose), build_cuda(force, verbose)


if __name__ == "__main__":
    h, c = build_all(force="--force" in sys.argv, verbose=True)
    print("host:", h)
    print("cuda:", c)
