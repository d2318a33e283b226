"""ctypes bindings to the C ABI in ``include/fmpi.h``.

The ABI is the single source of truth for both this package and the Julia binding
(``julia/FluxMPIB200``). Nothing here falls back silently: if a CUDA device is present and
``libfmpi_cuda.so`` cannot be loaded, importing the device layer raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

from . import _build

MAX_RANKS = 8
MAX_BLOCKS = 160
NUM_CHANNELS = 4

# dtype / op enums (mirror include/fmpi.h)
F32, F64, F16, BF16, I32, I64, U8, I8, I16 = range(9)
SUM, PROD, MIN, MAX = range(4)
RULE_DESCENT, RULE_MOMENTUM, RULE_NESTEROV, RULE_ADAM, RULE_ADAMW = range(5)
ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_NVLS, ALGO_PIPE, ALGO_PULL_P2P = range(6)
SYMM_VMM, SYMM_IPC, SYMM_LOCAL = range(3)

OPS = {"sum": SUM, "+": SUM, "add": SUM, "prod": PROD, "*": PROD, "mul": PROD,
       "min": MIN, "max": MAX}


def op_code(op) -> int:
    """Accepts "sum"/"+"/"prod"/"*"/"min"/"max", the enum ints, or the python/torch callables
    ``operator.add``/``operator.mul``/``min``/``max``/``torch.add``/… (the reference takes a
    Julia function and wraps it in ``MPI.Op``, src/mpi_extensions.jl:52-54)."""
    if isinstance(op, int):
        if 0 <= op <= 3:
            return op
        raise ValueError(f"unknown reduction op code {op}")
    if isinstance(op, str):
        try:
            return OPS[op.lower()]
        except KeyError:
            raise ValueError(f"unknown reduction op {op!r}") from None
    name = getattr(op, "__name__", "")
    table = {"add": SUM, "iadd": SUM, "sum": SUM, "mul": PROD, "imul": PROD, "prod": PROD,
             "min": MIN, "minimum": MIN, "max": MAX, "maximum": MAX}
    if name in table:
        return table[name]
    raise ValueError(f"unsupported reduction op {op!r}; use sum/prod/min/max")


class NativeError(RuntimeError):
    pass


_host = None
_cuda = None


def _load(path: Path):
    return C.CDLL(str(path), mode=getattr(os, "RTLD_NOW", 2) | getattr(os, "RTLD_LOCAL", 0))


def host():
    global _host
    if _host is None:
        lib = _load(_build.build_host())
        lib.fmpi_last_error.restype = C.c_char_p
        lib.fmpi_fdx_last_error.restype = C.c_char_p
        lib.fmpi_host_init.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_size_t, C.c_double,
                                       C.POINTER(C.c_void_p)]
        lib.fmpi_host_finalize.argtypes = [C.c_void_p]
        lib.fmpi_host_barrier.argtypes = [C.c_void_p]
        lib.fmpi_host_allreduce.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                            C.c_int, C.c_int]
        lib.fmpi_host_bcast.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        lib.fmpi_host_reduce.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                         C.c_int, C.c_int, C.c_int]
        lib.fmpi_host_allgather.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        lib.fmpi_fdx_open.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int)]
        lib.fmpi_fdx_exchange.argtypes = [C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int,
                                          C.POINTER(C.c_int), C.c_double]
        lib.fmpi_fdx_bcast.argtypes = [C.c_int, C.c_char_p, C.c_int, C.c_int, C.c_int,
                                       C.POINTER(C.c_int), C.c_double]
        lib.fmpi_fdx_close.argtypes = [C.c_int, C.c_char_p, C.c_int]
        lib.fmpi_host_init_multi.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                             C.c_char_p, C.c_int, C.c_size_t, C.c_double,
                                             C.POINTER(C.c_void_p)]
        lib.fmpi_net_init.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_double,
                                      C.POINTER(C.c_void_p)]
        lib.fmpi_net_finalize.argtypes = [C.c_void_p]
        lib.fmpi_net_barrier.argtypes = [C.c_void_p]
        lib.fmpi_net_allreduce.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int]
        lib.fmpi_net_reduce.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int,
                                        C.c_int]
        lib.fmpi_net_bcast.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]
        lib.fmpi_net_allgather.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        lib.fmpi_host_allreduce_many.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                                 C.c_int, C.c_int]
        lib.fmpi_host_symm_alloc.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
        lib.fmpi_host_symm_ptr.argtypes = [C.c_void_p]
        lib.fmpi_host_symm_ptr.restype = C.c_void_p
        lib.fmpi_host_symm_bytes.argtypes = [C.c_void_p]
        lib.fmpi_host_symm_bytes.restype = C.c_size_t
        lib.fmpi_host_symm_free.argtypes = [C.c_void_p, C.c_void_p]
        lib.fmpi_host_allreduce_symm.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t,
                                                 C.c_int, C.c_int]
        lib.fmpi_host_bcast_symm.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_int]
        lib.fmpi_host_multi_tensor_step.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                                                    C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                                    C.c_void_p, C.c_int64]
        _host = lib
    return _host


def host_check(rc: int, what: str = ""):
    if rc != 0:
        msg = host().fmpi_last_error().decode(errors="replace")
        raise NativeError(f"{what}: {msg}" if what else msg)


def fdx_check(rc: int, what: str = ""):
    if rc != 0:
        msg = host().fmpi_fdx_last_error().decode(errors="replace")
        raise NativeError(f"{what}: {msg}" if what else msg)


class DevComm(C.Structure):
    _fields_ = [
        ("rank", C.c_int32),
        ("world", C.c_int32),
        ("emulate", C.c_int32),
        ("channel", C.c_int32),
        ("data", C.c_void_p * MAX_RANKS),
        ("sig", C.c_void_p * MAX_RANKS),
        ("mc_data", C.c_void_p),
        ("data_bytes", C.c_uint64),
        ("ctrl", C.c_void_p),
        ("timeout_ns", C.c_uint64),
    ]


class Launch(C.Structure):
    _fields_ = [("blocks", C.c_int32), ("threads", C.c_int32), ("algo", C.c_int32)]


class Hyper(C.Structure):
    _fields_ = [
        ("rule", C.c_int32),
        ("eta", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
        ("lam", C.c_float), ("rho", C.c_float),
        ("grad_scale", C.c_float),
        ("zero_grad", C.c_int32),
        ("eta_dev", C.c_uint64),
    ]


PtrArray = C.c_void_p * MAX_RANKS


def ptr_array(ptrs) -> "PtrArray":
    a = PtrArray()
    for i, p in enumerate(ptrs):
        a[i] = p
    return a


def cuda_lib_path() -> Path:
    return _build.LIB / "libfmpi_cuda.so"


def cuda():
    """Load libfmpi_cuda.so (building it first if stale and nvcc is present)."""
    global _cuda
    if _cuda is None:
        path = _build.build_cuda()
        if path is None or not Path(path).exists():
            raise NativeError(
                "libfmpi_cuda.so is missing and nvcc is not available to build it; "
                "run `python -c 'import __graft_entry__ as g; g.build()'` on a box with nvcc")
        lib = _load(path)
        vp, u64, i32 = C.c_void_p, C.c_uint64, C.c_int
        lib.fmpi_cuda_last_error.restype = C.c_char_p
        lib.fmpi_topo_json.argtypes = [C.c_char_p, C.c_size_t]
        lib.fmpi_symm_create.argtypes = [i32, C.c_size_t, i32, i32, i32, C.POINTER(vp)]
        lib.fmpi_symm_size.argtypes = [vp]
        lib.fmpi_symm_size.restype = C.c_size_t
        lib.fmpi_symm_local_ptr.argtypes = [vp]
        lib.fmpi_symm_local_ptr.restype = vp
        lib.fmpi_symm_mode.argtypes = [vp]
        lib.fmpi_symm_export_fd.argtypes = [vp, C.POINTER(i32)]
        lib.fmpi_symm_export_ipc.argtypes = [vp, vp]
        lib.fmpi_symm_import_fd.argtypes = [vp, i32, i32]
        lib.fmpi_symm_import_ipc.argtypes = [vp, i32, vp]
        lib.fmpi_symm_set_self.argtypes = [vp, i32]
        lib.fmpi_symm_peer_ptr.argtypes = [vp, i32]
        lib.fmpi_symm_peer_ptr.restype = vp
        lib.fmpi_symm_mc_supported.argtypes = [i32]
        lib.fmpi_symm_mc_create.argtypes = [vp, C.POINTER(i32)]
        lib.fmpi_symm_mc_import.argtypes = [vp, i32]
        lib.fmpi_symm_mc_add_device.argtypes = [vp]
        lib.fmpi_symm_mc_bind_map.argtypes = [vp]
        lib.fmpi_symm_mc_ptr.argtypes = [vp]
        lib.fmpi_symm_mc_ptr.restype = vp
        lib.fmpi_symm_destroy.argtypes = [vp]
        lib.fmpi_sigpad_bytes.restype = C.c_size_t
        lib.fmpi_health_page.argtypes = [C.POINTER(vp), C.POINTER(vp)]
        dc, lp = C.POINTER(DevComm), C.POINTER(Launch)
        lib.fmpi_allreduce_symm.argtypes = [dc, u64, u64, i32, i32, lp, vp]
        lib.fmpi_bcast_symm.argtypes = [dc, u64, u64, i32, lp, vp]
        lib.fmpi_reduce_symm.argtypes = [dc, u64, u64, i32, i32, i32, lp, vp]
        lib.fmpi_allreduce_any.argtypes = [dc, vp, vp, u64, i32, i32, u64, u64, lp, vp]
        lib.fmpi_bcast_any.argtypes = [dc, vp, u64, i32, u64, u64, lp, vp]
        lib.fmpi_reduce_any.argtypes = [dc, vp, vp, u64, i32, i32, i32, u64, u64, lp, vp]
        lib.fmpi_barrier_dev.argtypes = [dc, vp]
        lib.fmpi_pipe_trace_read.argtypes = [vp, C.c_size_t]
        lib.fmpi_fused_step.argtypes = [dc, u64, i32, u64, i32, vp, vp, vp, u64, u64,
                                        C.POINTER(Hyper), vp, lp, vp]
        lib.fmpi_fused_step_oneshot.argtypes = [dc, u64, u64, i32, vp, vp, vp, u64, u64,
                                                C.POINTER(Hyper), vp, u64, u64, lp, vp]
        lib.fmpi_step_advance.argtypes = [vp, vp]
        lib.fmpi_multi_tensor_step.argtypes = [i32, vp, vp, vp, vp, vp, i32, i32,
                                               C.POINTER(Hyper), C.c_int64, vp]
        lib.fmpi_tree_step_oneshot.argtypes = [dc, i32, vp, u64, i32, C.POINTER(Hyper), C.c_int64, u64, u64, lp, vp]
        lib.fmpi_tree_allreduce.argtypes = [dc, i32, vp, u64, u64, i32, i32, lp, vp]
        lib.fmpi_pack.argtypes = [i32, vp, vp, vp, vp, vp]
        lib.fmpi_unpack.argtypes = [i32, vp, vp, vp, vp, vp]
        lib.fmpi_reduce_scratch_bytes.restype = C.c_size_t
        lib.fmpi_sumsq.argtypes = [vp, u64, i32, vp, vp, vp]
        lib.fmpi_checksum.argtypes = [vp, u64, vp, vp, vp]
        lib.fmpi_sumsq_noalloc.argtypes = [vp, u64, i32, vp, vp]
        f32 = C.c_float
        lib.fmpi_bn_scratch_bytes.argtypes = [i32]
        lib.fmpi_bn_scratch_bytes.restype = C.c_size_t
        lib.fmpi_bn_forward.argtypes = [vp, vp, vp, u64, i32, i32, vp, vp, vp, vp, f32, f32, i32, vp, vp, vp, vp]
        lib.fmpi_bn_apply.argtypes = [vp, vp, vp, u64, i32, i32, vp, vp, i32, vp]
        lib.fmpi_bn_backward.argtypes = [vp, vp, vp, vp, vp, u64, i32, i32, vp, vp, vp, vp, vp, vp, i32, vp, vp]
        lib.fmpi_maxpool_fwd.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp]
        lib.fmpi_maxpool_bwd.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp]
        lib.fmpi_flush_l2.argtypes = [vp, u64, vp]
        _cuda = lib
    return _cuda


def cuda_check(rc: int, what: str = ""):
    if rc != 0:
        msg = cuda().fmpi_cuda_last_error().decode(errors="replace")
        raise NativeError(f"{what}: {msg}" if what else msg)


def cuda_loaded() -> bool:
    return _cuda is not None


# launch counter: every kernel launch of ours goes through one of the fmpi_* entry points; the
# wrappers in ops/ bump this so bench.py can report `gpu_launches`.
launch_count = 0


def count_launch(n: int = 1):
    global launch_count
    launch_count += n
