"""ctypes ``argtypes`` in fluxmpi_b200/_native.py against the prototypes of include/fmpi.h:
same number of arguments and the same class (32-bit int, 64-bit unsigned / signed, float, double,
pointer) in every position. A mismatch here is silent memory corruption at run time, so it is
checked where no GPU is needed (both libraries load without a driver)."""
import ctypes as C
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CMAP = {"int": "i32", "int32_t": "i32", "size_t": "u64", "uint64_t": "u64", "int64_t": "i64",
        "double": "f64", "float": "f32"}


def _cclass(t: str) -> str:
    t = t.strip()
    if "*" in t or t.startswith("fmpi_stream"):
        return "ptr"
    ty = t.replace("const", "").strip()
    ty = re.sub(r"\s+\w+$", "", ty).strip() if " " in ty else ty
    return CMAP.get(ty, "?" + ty)


def _pclass(t) -> str:
    if t in (C.c_int, C.c_int32):
        return "i32"
    if t in (C.c_size_t, C.c_uint64):
        return "u64"
    if t is C.c_int64:
        return "i64"
    if t is C.c_double:
        return "f64"
    if t is C.c_float:
        return "f32"
    if t in (C.c_void_p, C.c_char_p) or isinstance(getattr(t, "_type_", None), type):
        return "ptr"
    return "?" + repr(t)


def test_argtypes_match_the_header():
    from fluxmpi_b200 import _build, _native
    hdr = re.sub(r"/\*.*?\*/", "", (ROOT / "include" / "fmpi.h").read_text(), flags=re.S)
    protos = {}
    for m in re.finditer(r"\b(fmpi_\w+)\s*\(([^;{}]*?)\)\s*;", hdr, flags=re.S):
        args = m.group(2).strip()
        protos[m.group(1)] = [] if args in ("", "void") else [_cclass(a) for a in args.split(",")]
    libs = [_native.host()]
    if _build.build_cuda() is not None:
        libs.append(_native.cuda())
    checked, problems = 0, []
    for lib in libs:
        for name, ct in protos.items():
            try:
                fn = getattr(lib, name)
            except AttributeError:
                continue                                   # lives in the other library
            if fn.argtypes is None:
                continue                                   # never called from Python with arguments
            checked += 1
            pt = [_pclass(t) for t in fn.argtypes]
            if pt != ct:
                problems.append(f"{name}: python {pt} vs C {ct}")
    assert checked >= 20 and not problems, "\n".join(problems)
