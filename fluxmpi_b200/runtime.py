"""Process runtime: Init, rank/size, rank-ordered printing, configuration.

Reference behaviour being reproduced (citations into /root/reference):

* ``Init(; gpu_devices, verbose)`` — src/common.jl:16-45. Idempotent, joins the job, binds a
  device. Ours binds GPU ``LOCAL_RANK`` by default (the reference's ``(rank+1) % ndev`` puts
  rank 0 on GPU 1) and keeps the ``gpu_devices`` override.
* ``Initialized()`` — src/common.jl:6.
* ``local_rank()`` / ``total_workers()`` — src/common.jl:52-69; raise
  ``FluxMPINotInitializedError`` before Init; plain ints, hence constants under autograd
  (the reference marks them ``@non_differentiable``).
* ``fluxmpi_print`` / ``fluxmpi_println`` — src/common.jl:72-98: timestamp only before Init,
  bare print at one worker, else ``"<ts> [rank / size] "`` serialised in rank order.
* ``disable_cudampi_support`` — src/FluxMPI.jl:51-56 → ``disable_p2p_support``.

Process model: SPMD, one process per GPU. Ranks come from ``RANK`` / ``WORLD_SIZE`` /
``LOCAL_RANK`` (set by ``python -m fluxmpi_b200.launch`` or by ``torchrun``). The host world is
our own POSIX-shm backend (no MPI on this image); the device world is a symmetric heap mapped
into every peer (parallel/device_world.py). ``torch.distributed`` is only brought up on demand
for the NCCL fallback / baseline.
"""
from __future__ import annotations

import atexit
import ctypes as C
import datetime as _dt
import logging
import os
import sys
import threading
import warnings
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass, field
from typing import Optional, Sequence

import torch

from . import _native
from .utils import prefs


class FluxMPINotInitializedError(RuntimeError):
    """Raised by local_rank()/total_workers() before Init (src/FluxMPI.jl:59-63)."""

    def __init__(self):
        super().__init__("Please call fluxmpi_b200.Init(...) before using FluxMPI functionalities!")


@dataclass
class _State:
    initialized: bool = False
    rank: int = 0
    world: int = 1
    local_rank: int = 0
    local_world: int = 1
    node: int = 0
    nnodes: int = 1
    device: Optional[torch.device] = None
    job: str = ""
    host_comm: Optional[int] = None           # fmpi_host_comm*
    host_pool: Optional[ThreadPoolExecutor] = None
    device_world: object = None               # parallel.device_world.DeviceWorld
    p2p_disabled: bool = False
    topology: dict = field(default_factory=dict)
    timeout_s: float = 300.0
    slot_bytes: int = 8 << 20                 # staging slot of the host world
    host_symm: list = field(default_factory=list)   # (lo, hi, handle, buffer) of comm.shm_empty blocks
    lock: threading.Lock = field(default_factory=threading.Lock)
    host_exec: threading.Lock = field(default_factory=threading.Lock)   # one host collective at a time
    host_count: threading.Lock = field(default_factory=threading.Lock)  # guards host_pending
    host_pending: int = 0                     # non-blocking host collectives not yet finished


_S = _State()
_log = logging.getLogger("fluxmpi_b200")

# ---------------------------------------------------------------------------- configuration

if "FLUXMPI_DISABLE_CUDAMPI_SUPPORT" in os.environ:
    # src/FluxMPI.jl:17-19 — the legacy env var only warns.
    warnings.warn(
        "FLUXMPI_DISABLE_CUDAMPI_SUPPORT environment variable has been removed and has no "
        "effect. Please use `fluxmpi_b200.disable_p2p_support()` instead.", stacklevel=2)

_S.p2p_disabled = bool(prefs.load_preference(prefs.KEY_DISABLE_P2P, False)) or \
    os.environ.get("FMPI_DISABLE_P2P", "0") not in ("0", "", "false", "False")


def disable_p2p_support(disable: bool = True):
    """Persist the "no NVLink P2P / multicast kernels" switch. Takes effect in the next
    process (same contract as the reference's preference, src/FluxMPI.jl:46-56)."""
    p = prefs.set_preference(prefs.KEY_DISABLE_P2P, bool(disable))
    print(f"[fluxmpi_b200] NVLink P2P support {'disabled' if disable else 'enabled'} in {p}. "
          "Restart the process for this change to take effect!", file=sys.stderr)


# Name kept for scripts written against the reference API.
disable_cudampi_support = disable_p2p_support


def p2p_disabled() -> bool:
    return _S.p2p_disabled


# ---------------------------------------------------------------------------- init


def Initialized() -> bool:
    return _S.initialized


def _env_int(name: str, default: int) -> int:
    v = os.environ.get(name)
    return int(v) if v not in (None, "") else default


def _job_name() -> str:
    job = os.environ.get("FMPI_JOB_ID")
    if job:
        return job
    port = os.environ.get("MASTER_PORT", "0")
    run = os.environ.get("TORCHELASTIC_RUN_ID")
    if run is not None:
        # under torchrun every worker is a child of the same agent process: its pid makes the
        # name unique per launch even when the port and the (default "none") run id repeat, so
        # a segment left behind by a crashed earlier job can never be mistaken for this one
        return f"{port}-{run}-{os.getuid()}-{os.getppid()}"
    return f"{port}-none-{os.getuid()}"


def Init(gpu_devices: Optional[Sequence[int]] = None, verbose: bool = False,
         slot_bytes: int = 8 << 20, timeout_s: Optional[float] = None,
         device_window_bytes: Optional[int] = None) -> None:
    """Set up the data-parallel world. Safe to call more than once."""
    if _S.initialized:
        if verbose:
            fluxmpi_println("FluxMPI already initialized; Skipping...")
        return
    with _S.lock:
        rank = _env_int("RANK", 0)
        world = _env_int("WORLD_SIZE", 1)
        lrank = _env_int("LOCAL_RANK", rank)
        lworld = _env_int("LOCAL_WORLD_SIZE", world)
        if not (0 <= rank < world):
            raise ValueError(f"bad RANK={rank} for WORLD_SIZE={world}")
        _S.rank, _S.world, _S.local_rank, _S.local_world = rank, world, lrank, lworld
        _S.timeout_s = float(timeout_s if timeout_s is not None
                             else os.environ.get("FMPI_TIMEOUT_S", 300.0))
        _S.slot_bytes = (int(slot_bytes) + 63) & ~63 or (4 << 20)
        _S.job = _job_name()

        if world % lworld or not (0 <= lrank < lworld):
            raise ValueError(f"WORLD_SIZE={world} / LOCAL_WORLD_SIZE={lworld} / LOCAL_RANK={lrank}: "
                             "every node must run the same number of ranks")
        nnodes = world // lworld
        node = _env_int("GROUP_RANK", rank // lworld)
        if nnodes > 1 and rank != node * lworld + lrank:
            raise ValueError(f"RANK={rank} is not node-major (node {node}, local rank {lrank} of {lworld})")
        _S.node, _S.nnodes = node, nnodes

        if world > 1:
            lib = _native.host()
            h = C.c_void_p()
            name = ("/fmpi_" + "".join(ch if ch.isalnum() else "_" for ch in _S.job))[:200]
            if nnodes == 1:
                _native.host_check(
                    lib.fmpi_host_init(name.encode(), rank, world, slot_bytes, _S.timeout_s,
                                       C.byref(h)), "fmpi_host_init")
            else:
                # shared memory inside the node, a TCP mesh between the node leaders
                # (csrc/net_backend.cpp). The mesh port sits next to the rendezvous port the
                # launcher handed out unless FMPI_NET_PORT says otherwise.
                addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
                port = _env_int("FMPI_NET_PORT", _env_int("MASTER_PORT", 29500) + 1)
                _native.host_check(
                    lib.fmpi_host_init_multi(f"{name}_n{node}".encode(), lrank, lworld, node, nnodes,
                                             addr.encode(), port, slot_bytes, _S.timeout_s,
                                             C.byref(h)), "fmpi_host_init_multi")
            _S.host_comm = h.value
        # One worker => host collectives (blocking and non-blocking) execute in issue order.
        _S.host_pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix="fmpi-host")
        _S.host_pending = 0
        _S.initialized = True

    if verbose and world == 1:
        warnings.warn("Using FluxMPI with only 1 worker. It might be faster to run the code "
                      "without MPI", stacklevel=2)

    if torch.cuda.is_available():
        ndev = torch.cuda.device_count()
        if gpu_devices is None:
            # pretend nodes (launch --sim-nodes) share one machine's GPUs: spread by global rank
            dev = (rank if os.environ.get("FMPI_SIM_NODES") else lrank) % ndev
        else:
            dev = int(gpu_devices[rank])
        if verbose:
            fluxmpi_println(f"Using GPU {dev}")
        torch.cuda.set_device(dev)
        _S.device = torch.device("cuda", dev)
        from .parallel import device_world
        _S.device_world = device_world.create(
            rank=rank, world=world, device=dev, job=_S.job, disabled=_S.p2p_disabled,
            window_bytes=device_window_bytes, verbose=verbose, nnodes=_S.nnodes)
        _S.topology = _S.device_world.topology
        # one capability message per process, like the reference's load-time probe
        # (src/FluxMPI.jl:25-31: "CUDA-aware MPI support disabled" / "not CUDA Aware" / "is CUDA Aware")
        dw = _S.device_world
        if _S.p2p_disabled and world > 1:
            _log.info("NVLink P2P support disabled using LocalPreferences.toml; device collectives use NCCL.")
        elif world > 1 and not dw.native:
            _log.warning("NVLink peer-to-peer memory is not available (%s); device collectives use NCCL.",
                         dw.why_fallback)
        else:
            _log.info("NVLink P2P kernels active (mode=%s, multicast=%s).", dw.mode, dw.has_multicast)
    else:
        _S.device = torch.device("cpu")
        if verbose:
            fluxmpi_println("Using CPU")
    atexit.register(Finalize)


def Finalize() -> None:
    """Tear the world down (the reference relies on ``MPI.Finalize``, test/test_common.jl:15)."""
    if not _S.initialized:
        return
    with _S.lock:
        from . import group as _group
        _group.free_all()
        if _S.device_world is not None:
            try:
                _S.device_world.close()
            except Exception:
                pass
            _S.device_world = None
        if _S.host_pool is not None:
            _S.host_pool.shutdown(wait=True)
            _S.host_pool = None
        for _lo, _hi, handle, _buf in _S.host_symm:      # unmap without a barrier: peers may be gone
            _native.host().fmpi_host_symm_free(None, handle)
        _S.host_symm = []
        if _S.host_comm is not None:
            _native.host().fmpi_host_finalize(_S.host_comm)
            _S.host_comm = None
        _S.initialized = False


def local_rank() -> int:
    if not _S.initialized:
        raise FluxMPINotInitializedError()
    return _S.rank


def total_workers() -> int:
    if not _S.initialized:
        raise FluxMPINotInitializedError()
    return _S.world


def node_local_rank() -> int:
    """Rank within the node (``LOCAL_RANK``); the reference has no such notion
    (src/common.jl:33-34 uses the global rank)."""
    if not _S.initialized:
        raise FluxMPINotInitializedError()
    return _S.local_rank


def current_device() -> torch.device:
    if not _S.initialized:
        raise FluxMPINotInitializedError()
    return _S.device


def topology() -> dict:
    return dict(_S.topology)


def state() -> _State:
    return _S


# ---------------------------------------------------------------------------- host helpers


def _host_done(_fut):
    with _S.host_count:               # the callback runs on the worker thread
        _S.host_pending -= 1


def host_submit(fn, *args):
    """Queue ``fn`` on the single host-communication worker (non-blocking collectives; keeps
    issue order). Returns the future."""
    if not _S.initialized:
        raise FluxMPINotInitializedError()

    def run():
        with _S.host_exec:
            return fn(*args)

    with _S.host_count:
        _S.host_pending += 1
    fut = _S.host_pool.submit(run)
    fut.add_done_callback(_host_done)
    return fut


def host_run(fn, *args):
    """Run a blocking host collective. With nothing queued it runs on the calling thread (a
    hand-over to the worker and back costs more than a small shared-memory all-reduce);
    otherwise it takes its place behind the non-blocking calls already issued."""
    if not _S.initialized:
        raise FluxMPINotInitializedError()
    try:
        if _S.host_pending > 0:
            return host_submit(fn, *args).result()
        with _S.host_exec:
            return fn(*args)
    except _native.NativeError:
        # the host world gave up on a peer (timeout / abort flag): kernels that are waiting for
        # that peer's flags on this GPU must give up too, or they would spin until their own budget
        dw = _S.device_world
        if dw is not None:
            dw.abort()
        raise


def _host_barrier_now():
    if _S.world > 1:
        _native.host_check(_native.host().fmpi_host_barrier(_S.host_comm), "barrier")


def host_barrier():
    """CPU-side barrier across all ranks (``MPI.Barrier``, src/common.jl:91). Never touches a
    GPU stream."""
    if not _S.initialized:
        raise FluxMPINotInitializedError()
    if _S.world > 1:
        host_run(_host_barrier_now)


def host_allgather_bytes(blob: bytes, size: Optional[int] = None) -> list[bytes]:
    """All-gather one equal-sized blob per rank (handle exchange)."""
    n = len(blob) if size is None else size
    blob = blob.ljust(n, b"\0")
    if _S.world == 1:
        return [blob]

    def run():
        send = C.create_string_buffer(blob, n)
        recv = C.create_string_buffer(n * _S.world)
        _native.host_check(_native.host().fmpi_host_allgather(_S.host_comm, send, recv, n),
                           "allgather")
        raw = recv.raw
        return [raw[i * n:(i + 1) * n] for i in range(_S.world)]

    return host_run(run)


# ---------------------------------------------------------------------------- printing


def _now() -> str:
    return _dt.datetime.now().isoformat(timespec="milliseconds")


def _emit(fn_end: str, prefix: str, args, kwargs):
    file = kwargs.pop("file", sys.stdout)
    end = kwargs.pop("end", fn_end)
    sep = kwargs.pop("sep", "")
    # Julia's print concatenates its arguments without a separator; keep that.
    text = sep.join(str(a) for a in args)
    file.write(prefix + text + end)
    file.flush()


def _fluxmpi_print(end: str, args, kwargs):
    if not _S.initialized:
        _emit(end, f"{_now()} ", args, kwargs)
        return
    if _S.world == 1:
        _emit(end, "", args, kwargs)
        return
    for r in range(_S.world):
        if r == _S.rank:
            _emit(end, f"{_now()} [{_S.rank} / {_S.world}] ", args, dict(kwargs))
        host_barrier()


def fluxmpi_println(*args, **kwargs) -> None:
    """Add rank and size information to the printed statement (src/common.jl:100-105)."""
    _fluxmpi_print("\n", args, kwargs)


def fluxmpi_print(*args, **kwargs) -> None:
    """Like :func:`fluxmpi_println` without the trailing newline (src/common.jl:107-112)."""
    _fluxmpi_print("", args, kwargs)
