"""The Julia binding cannot be executed on this image (no Julia). What can be checked without
it: every ``ccall`` in julia/FluxMPIB200 names a function that include/fmpi.h declares, passes
as many arguments as the prototype has, and asks the library that actually exports the symbol
(``hsym`` = libfmpi_host.so, ``csym`` = libfmpi_cuda.so). The struct mirrors are checked by
size against the C compiler."""
import re
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def _prototypes():
    hdr = re.sub(r"/\*.*?\*/", "", (ROOT / "include" / "fmpi.h").read_text(), flags=re.S)
    out = {}
    for m in re.finditer(r"\b(fmpi_\w+)\s*\(([^;{}]*?)\)\s*;", hdr, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else len(args.split(","))
    return out


def _split_top_level(s: str):
    items, depth, tok = [], 0, ""
    for ch in s:
        depth += ch in "({"
        depth -= ch in ")}"
        if ch == "," and depth == 0:
            items.append(tok.strip())
            tok = ""
        else:
            tok += ch
    if tok.strip():
        items.append(tok.strip())
    return items


def _ccalls():
    calls = []
    for f in sorted((ROOT / "julia" / "FluxMPIB200").rglob("*.jl")):
        src = f.read_text()
        for m in re.finditer(r"ccall\(\s*([hc])sym\(:(\w+)\)\s*,\s*[\w{}\.]+\s*,\s*\(", src):
            i = j = m.end()
            depth = 1
            while depth:
                depth += src[j] == "("
                depth -= src[j] == ")"
                j += 1
            calls.append((f.name, m.group(1), m.group(2), len(_split_top_level(src[i:j - 1]))))
    return calls


def _exports(lib: Path):
    out = subprocess.run(["nm", "-D", "--defined-only", str(lib)], capture_output=True, text=True, check=True).stdout
    return {line.split()[-1] for line in out.splitlines() if line.strip()}


def test_every_ccall_matches_the_header_and_the_right_library():
    from fluxmpi_b200 import _build
    protos, calls = _prototypes(), _ccalls()
    assert len(calls) > 40 and len(protos) > 60
    host = _exports(_build.build_host())
    cuda_lib = _build.build_cuda()
    cuda = _exports(cuda_lib) if cuda_lib is not None and Path(cuda_lib).exists() else None
    problems = []
    for fname, lib, name, n in calls:
        if name not in protos:
            problems.append(f"{fname}: {name} is not declared in fmpi.h")
        elif protos[name] != n:
            problems.append(f"{fname}: {name} called with {n} arguments, prototype has {protos[name]}")
        if lib == "h" and name not in host:
            problems.append(f"{fname}: hsym(:{name}) but libfmpi_host.so does not export it")
        if lib == "c" and cuda is not None and name not in cuda:
            problems.append(f"{fname}: csym(:{name}) but libfmpi_cuda.so does not export it")
    assert not problems, "\n".join(problems)


def test_julia_struct_mirrors_have_the_c_sizes(tmp_path):
    """DevComm / Launch / Hyper in abi.jl and engine.jl against sizeof() from the header."""
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "fmpi.h"\nint main(void){ printf("%zu %zu %zu\\n", '
                   'sizeof(fmpi_devcomm), sizeof(fmpi_launch), sizeof(fmpi_hyper)); return 0; }\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-std=c11", f"-I{ROOT / 'include'}", str(src), "-o", str(exe)], check=True)
    devcomm, launch, hyper = map(int, subprocess.run([str(exe)], capture_output=True, text=True).stdout.split())
    sizes = {"Int32": 4, "UInt32": 4, "Float32": 4, "UInt64": 8, "Int64": 8, "Ptr{Cvoid}": 8}

    def julia_sizeof(path: str, struct: str) -> int:
        body = re.search(rf"struct {struct}\n(.*?)\nend", (ROOT / "julia" / "FluxMPIB200" / "src" / path).read_text(),
                         flags=re.S).group(1)
        total = 0
        for line in body.splitlines():
            ty = line.split("::")[1].strip()
            m = re.match(r"NTuple\{(\w+), (.*)\}", ty)
            if m:
                n = 8 if m.group(1) == "FMPI_MAX_RANKS" else int(m.group(1))
                total += n * sizes[m.group(2)]
            else:
                total += sizes[ty]
        return total

    assert julia_sizeof("abi.jl", "DevComm") == devcomm
    assert julia_sizeof("abi.jl", "Launch") == launch
    assert julia_sizeof("engine.jl", "Hyper") == hyper == 48


def test_ccall_argument_types_match_the_prototypes():
    """Class of every argument (32-bit int, 64-bit unsigned / signed, float, double, pointer) in
    the Julia type tuple against the C parameter in the same position."""
    hdr = re.sub(r"/\*.*?\*/", "", (ROOT / "include" / "fmpi.h").read_text(), flags=re.S)
    cmap = {"int": "i32", "int32_t": "i32", "size_t": "u64", "uint64_t": "u64", "int64_t": "i64",
            "double": "f64", "float": "f32"}
    jmap = {"Cint": "i32", "Int32": "i32", "Csize_t": "u64", "UInt64": "u64", "Int64": "i64",
            "Cdouble": "f64", "Cfloat": "f32", "Float32": "f32"}

    def cclass(t: str) -> str:
        t = t.strip()
        if "*" in t or t.startswith("fmpi_stream"):
            return "ptr"
        ty = t.replace("const", "").strip()
        ty = re.sub(r"\s+\w+$", "", ty).strip() if " " in ty else ty
        return cmap.get(ty, "?" + ty)

    def jclass(t: str) -> str:
        t = t.strip()
        if t.startswith(("Ptr", "Ref", "CuPtr")) or t == "Cstring":
            return "ptr"
        return jmap.get(t, "?" + t)

    protos = {}
    for m in re.finditer(r"\b(fmpi_\w+)\s*\(([^;{}]*?)\)\s*;", hdr, flags=re.S):
        args = m.group(2).strip()
        protos[m.group(1)] = [] if args in ("", "void") else [cclass(a) for a in args.split(",")]
    problems = []
    for f in sorted((ROOT / "julia" / "FluxMPIB200").rglob("*.jl")):
        src = f.read_text()
        for m in re.finditer(r"ccall\(\s*[hc]sym\(:(\w+)\)\s*,\s*[\w{}\.]+\s*,\s*\(", src):
            i = j = m.end()
            depth = 1
            while depth:
                depth += src[j] == "("
                depth -= src[j] == ")"
                j += 1
            jt = [jclass(x) for x in _split_top_level(src[i:j - 1])]
            if jt != protos.get(m.group(1)):
                problems.append(f"{f.name}: {m.group(1)}: julia {jt} vs C {protos.get(m.group(1))}")
    assert not problems, "\n".join(problems)
