"""Communication primitives — the reference's "MPIExtensions" layer.

Reference (citations into /root/reference/src/mpi_extensions.jl):

* ``allreduce!(v, op, comm)``  :97-111   → :func:`allreduce_`
* ``bcast!(v, root, comm)``    :119-133  → :func:`bcast_`
* ``reduce!(v, op, root, comm)`` :141-155 → :func:`reduce_` (non-root buffers untouched,
  test/test_mpi_extensions.jl:56-60)
* ``Iallreduce!(send, recv, op, comm)`` / ``Iallreduce!(buf, op, comm)`` :26-60 → :func:`Iallreduce`
* ``Ibcast!(buf, root, comm)`` :70-88 → :func:`Ibcast`
* ``cpu`` / ``gpu`` movers :5-8 → :func:`cpu`, :func:`gpu`

Differences by design: device tensors are never staged through the host — they go
to the NVLink peer-to-peer kernels of ``libfmpi_cuda.so`` (or NCCL when P2P support has been
disabled with :func:`fluxmpi_b200.disable_p2p_support`); all calls are truly in place *and*
return the tensor, so both calling styles of the reference work. ``comm`` is ``COMM_WORLD`` or a
:class:`fluxmpi_b200.Group` (a rank subset with its own host world and device signal pads);
roots are ranks within ``comm``, as in MPI.
"""
from __future__ import annotations

import ctypes as C
from concurrent.futures import Future
from typing import Optional

import numpy as np
import torch

from . import _native, runtime
from . import group as _group
from ._native import op_code
from .utils import trace as _trace

COMM_WORLD = "COMM_WORLD"

_TORCH_DTYPES = {
    torch.float32: _native.F32, torch.float64: _native.F64, torch.float16: _native.F16,
    torch.bfloat16: _native.BF16, torch.int32: _native.I32, torch.int64: _native.I64,
    torch.uint8: _native.U8, torch.int8: _native.I8, torch.int16: _native.I16,
    torch.bool: _native.U8,
}


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _TORCH_DTYPES[dt]
    except KeyError:
        raise TypeError(f"dtype {dt} is not supported by the collectives") from None


# ------------------------------------------------------------------------------ movers


def cpu(x):
    """Move arrays to host memory; anything that is not an array passes through
    (src/mpi_extensions.jl:5,7)."""
    return x.cpu() if isinstance(x, torch.Tensor) else x


def gpu(x):
    """Move arrays to this rank's GPU; non-arrays pass through (src/mpi_extensions.jl:6,8)."""
    if isinstance(x, torch.Tensor):
        return x.to(runtime.current_device()) if torch.cuda.is_available() else x
    if isinstance(x, np.ndarray) and torch.cuda.is_available():
        return torch.from_numpy(x).to(runtime.current_device())
    return x


# ------------------------------------------------------------------------------ requests


class Request:
    """Handle of a non-blocking collective (``MPI.Request``). ``wait()`` ≙ ``MPI.Wait!``.

    Host tensors: completes when the worker thread has finished the collective.
    Device tensors: wraps a CUDA event recorded on the communication stream; ``wait()`` makes
    the *current stream* wait for it (no host block) unless ``sync=True``.
    """

    def __init__(self, future: Optional[Future] = None, event: "torch.cuda.Event | None" = None,
                 buffer=None):
        self._future = future
        self._event = event
        self.buffer = buffer  # keeps the buffer alive, like req.buffer at mpi_extensions.jl:47

    def wait(self, sync: bool = False):
        if self._future is not None:
            self._future.result()
            self._future = None
        if self._event is not None:
            if sync:
                self._event.synchronize()
            else:
                torch.cuda.current_stream().wait_event(self._event)
            self._event = None
        return self.buffer

    def test(self) -> bool:
        if self._future is not None and not self._future.done():
            return False
        if self._event is not None and not self._event.query():
            return False
        return True


def Waitall(requests, sync: bool = False) -> None:
    """``MPI.Waitall!`` (src/optimizer.jl:59)."""
    for r in requests:
        r.wait(sync=sync)


# ------------------------------------------------------------------------------ helpers


def _as_tensor(v):
    if isinstance(v, torch.Tensor):
        return v
    if isinstance(v, np.ndarray):
        return torch.from_numpy(v)
    raise TypeError(f"collectives need a torch.Tensor or numpy array, got {type(v).__name__}")


def _host_view(t: torch.Tensor):
    """(contiguous work tensor, needs_copy_back)"""
    if t.is_contiguous():
        return t, False
    return t.contiguous(), True


def shm_empty(shape, dtype=torch.float32) -> torch.Tensor:
    """Collective: a CPU tensor in symmetric host memory (every rank must call it with the same
    shape, in the same order) — the host twin of :func:`symm_empty`. ``allreduce_`` on it (or on
    a contiguous slice) runs in place through the peers' mappings, with no staging copies
    (csrc/host_backend.cpp ``fmpi_host_allreduce_symm``). The memory lives until ``Finalize``;
    with one worker this is a plain ``torch.empty``."""
    if not runtime.Initialized():
        raise runtime.FluxMPINotInitializedError()
    S = runtime.state()
    numel = 1
    for d in (shape if isinstance(shape, (tuple, list, torch.Size)) else (shape,)):
        numel *= int(d)
    if S.world == 1 or S.host_comm is None or numel == 0:
        return torch.empty(shape, dtype=dtype)
    nbytes = numel * torch.empty((), dtype=dtype).element_size()
    lib = _native.host()

    def alloc():
        h = C.c_void_p()
        _native.host_check(lib.fmpi_host_symm_alloc(S.host_comm, nbytes, C.byref(h)), "host_symm_alloc")
        return h

    h = runtime.host_run(alloc)
    ptr = lib.fmpi_host_symm_ptr(h)
    buf = (C.c_char * nbytes).from_address(ptr)
    S.host_symm.append((ptr, ptr + nbytes, h, buf))
    return torch.frombuffer(buf, dtype=dtype, count=numel).view(shape)


def shm_free(t: torch.Tensor) -> None:
    """Collective: give back the block a tensor from :func:`shm_empty` lives in (call it on every
    rank, in the same order; the tensor and its views must not be used afterwards). Blocks that
    are never freed are released by ``Finalize``."""
    S = runtime.state()
    p = t.data_ptr()
    for i, (lo, hi, h, _buf) in enumerate(S.host_symm):
        if lo <= p < hi:
            del S.host_symm[i]
            runtime.host_run(lambda: _native.host_check(
                _native.host().fmpi_host_symm_free(S.host_comm, h), "host_symm_free"))
            return


def _host_symm_hit(t: torch.Tensor):
    """(segment handle, byte offset) when ``t`` lies inside a block from :func:`shm_empty`."""
    regs = runtime.state().host_symm
    if not regs or not t.is_contiguous():
        return None
    p, n = t.data_ptr(), t.numel() * t.element_size()
    for lo, hi, h, _ in regs:
        if lo <= p and p + n <= hi:
            return h, p - lo
    return None


class _HC:
    """host communicator view: (handle, rank, world) of the world or of a Group"""
    __slots__ = ("host_comm", "rank", "world", "is_world")

    def __init__(self, grp=None):
        S = runtime.state()
        if grp is None:
            self.host_comm, self.rank, self.world, self.is_world = S.host_comm, S.rank, S.world, True
        else:
            self.host_comm, self.rank, self.world, self.is_world = grp.host_comm, grp.rank, grp.size, False


def _host_allreduce_now(send: torch.Tensor, recv: torch.Tensor, op: int, grp=None):
    S = _HC(grp)
    if S.world == 1:
        if recv.data_ptr() != send.data_ptr():
            recv.copy_(send)
        return
    if S.is_world and send.data_ptr() == recv.data_ptr() and send.dtype != torch.bool:
        hit = _host_symm_hit(recv)
        if hit is not None:
            _native.host_check(
                _native.host().fmpi_host_allreduce_symm(S.host_comm, hit[0], hit[1], recv.numel(),
                                                        dtype_code(recv.dtype), op), "allreduce (symmetric)")
            return
    s, _ = _host_view(send)
    r, back = _host_view(recv)
    _native.host_check(
        _native.host().fmpi_host_allreduce(S.host_comm, s.data_ptr(), r.data_ptr(), s.numel(),
                                           dtype_code(s.dtype), op), "allreduce")
    if back:
        recv.copy_(r)


def _host_allreduce_many_now(ts, op: int):
    """All contiguous CPU tensors of one dtype, reduced in place as one message."""
    S = runtime.state()
    if S.world == 1 or not ts:
        return
    n = len(ts)
    ptrs, counts = (C.c_void_p * n)(), (C.c_uint64 * n)()
    for i, t in enumerate(ts):
        ptrs[i], counts[i] = t.data_ptr(), t.numel()
    _native.host_check(
        _native.host().fmpi_host_allreduce_many(S.host_comm, n, ptrs, counts,
                                                dtype_code(ts[0].dtype), op), "allreduce_many")


def _host_bcast_now(buf: torch.Tensor, root: int, grp=None):
    S = _HC(grp)
    if S.world == 1:
        return
    hit = _host_symm_hit(buf) if S.is_world else None
    if hit is not None:            # symmetric host memory: copied straight out of the root's block
        _native.host_check(
            _native.host().fmpi_host_bcast_symm(S.host_comm, hit[0], hit[1], buf.numel() * buf.element_size(),
                                                root), "bcast (symmetric)")
        return
    b, back = _host_view(buf)
    _native.host_check(
        _native.host().fmpi_host_bcast(S.host_comm, b.data_ptr(), b.numel() * b.element_size(),
                                       root), "bcast")
    if back:
        buf.copy_(b)


def _host_reduce_now(send: torch.Tensor, recv: torch.Tensor, op: int, root: int, grp=None):
    S = _HC(grp)
    if S.world == 1:
        if recv.data_ptr() != send.data_ptr():
            recv.copy_(send)
        return
    s, _ = _host_view(send)
    r, back = _host_view(recv)
    _native.host_check(
        _native.host().fmpi_host_reduce(S.host_comm, s.data_ptr(), r.data_ptr(), s.numel(),
                                        dtype_code(s.dtype), op, root), "reduce")
    if back and S.rank == root:
        recv.copy_(r)


def _check_pair(send: torch.Tensor, recv: torch.Tensor):
    if send.shape != recv.shape or send.dtype != recv.dtype or send.device != recv.device:
        raise ValueError("send and recv buffers must agree in shape, dtype and device")


def _device_world():
    dw = runtime.state().device_world
    if dw is None:
        raise RuntimeError("device collectives need Init() on a CUDA machine")
    return dw


# ------------------------------------------------------------------------------ blocking API


def _resolve_op(op):
    """(native op code, None) for sum / prod / min / max in any spelling, else (None, callable):
    the reference accepts any Julia function and wraps it in ``MPI.Op``
    (src/mpi_extensions.jl:52-54); an arbitrary Python callable ``f(a, b) -> tensor`` takes the
    gather-and-fold path below."""
    try:
        return op_code(op), None
    except ValueError:
        if callable(op):
            return None, op
        raise


def _host_custom_now(send: torch.Tensor, recv: torch.Tensor, fn, root: int, grp=None):
    """Reduction with a user function: all-gather the operands, fold them in rank order on
    every rank (same order everywhere, hence the same bits). ``root < 0`` = all ranks keep the
    result. W copies of the message live on each rank while this runs: meant for the odd custom
    reduction, not for gradients."""
    S = _HC(grp)
    slot_bytes = runtime.state().slot_bytes
    s, _ = _host_view(send)
    if S.world == 1:
        if recv.data_ptr() != s.data_ptr():
            recv.copy_(s)
        return
    raw = s.reshape(-1).view(torch.uint8) if s.dtype != torch.bool else s.reshape(-1).to(torch.uint8)
    n, W = raw.numel(), S.world
    parts = torch.empty(W, n, dtype=torch.uint8)
    chunk = max(1, slot_bytes // W)
    tmp = torch.empty(W * chunk, dtype=torch.uint8)
    lib = _native.host()
    for off in range(0, n, chunk):
        m = min(chunk, n - off)
        _native.host_check(lib.fmpi_host_allgather(S.host_comm, raw.data_ptr() + off, tmp.data_ptr(), m),
                           "allgather")
        parts[:, off:off + m] = tmp[:W * m].view(W, m)
    if root >= 0 and S.rank != root:
        return
    vals = [p.view(s.dtype).view(s.shape) if s.dtype != torch.bool else p.bool().view(s.shape) for p in parts]
    acc = vals[0]
    for x in vals[1:]:
        acc = fn(acc, x)
    with torch.no_grad():
        recv.copy_(torch.as_tensor(acc).reshape(recv.shape))


def _custom(send: torch.Tensor, recv: torch.Tensor, fn, root: int = -1, grp=None):
    if recv.is_cuda:                       # staged through the host, as the reference's MPI would
        h = send.detach().cpu()
        out = h.clone()
        runtime.host_run(_host_custom_now, h, out, fn, root, grp)
        if root < 0 or _HC(grp).rank == root:
            with torch.no_grad():
                recv.copy_(out)
    else:
        runtime.host_run(_host_custom_now, send.detach(), recv.detach(), fn, root, grp)


def _count(kind: str, t: torch.Tensor) -> None:
    _trace.count(kind, "cuda" if t.is_cuda else "host", t.numel() * t.element_size())


def allreduce_(v, op="sum", comm=COMM_WORLD):
    """In-place all-reduce of ``v`` with ``op`` over ``comm``; returns ``v``.

    CPU tensors / numpy arrays: POSIX-shm backend. CUDA tensors: NVLink P2P kernel enqueued on
    the current stream (stream-ordered, like every torch op). ``op`` may also be any callable
    ``f(a, b)`` on tensors (gathered and folded in rank order)."""
    if not runtime.Initialized():
        raise runtime.FluxMPINotInitializedError()
    grp = _group.resolve(comm)
    t = _as_tensor(v)
    code, fn = _resolve_op(op)
    _count("allreduce", t)
    if fn is not None:
        _custom(t, t, fn, grp=grp)
        return v
    dtype_code(t.dtype)
    if t.is_cuda:
        _device_world().allreduce(t, code, group=grp)
    else:
        runtime.host_run(_host_allreduce_now, t, t, code, grp)
    return v


def bcast_(v, root: int = 0, comm=COMM_WORLD):
    """In-place broadcast of ``v`` from ``root`` (a rank within ``comm``); returns ``v``."""
    if not runtime.Initialized():
        raise runtime.FluxMPINotInitializedError()
    grp = _group.resolve(comm)
    t = _as_tensor(v)
    _check_root(root, grp)
    _count("bcast", t)
    if t.is_cuda:
        _device_world().bcast(t, root, group=grp)
    else:
        runtime.host_run(_host_bcast_now, t, root, grp)
    return v


def reduce_(v, op="sum", root: int = 0, comm=COMM_WORLD):
    """Reduce to ``root`` (a rank within ``comm``) in place; non-root buffers are left untouched.
    Returns ``v``."""
    if not runtime.Initialized():
        raise runtime.FluxMPINotInitializedError()
    grp = _group.resolve(comm)
    t = _as_tensor(v)
    code, fn = _resolve_op(op)
    _check_root(root, grp)
    _count("reduce", t)
    if fn is not None:
        _custom(t, t, fn, root, grp)
        return v
    dtype_code(t.dtype)
    if t.is_cuda:
        _device_world().reduce(t, code, root, group=grp)
    else:
        runtime.host_run(_host_reduce_now, t, t, code, root, grp)
    return v


def _group_barrier_now(grp):
    if grp.size > 1:
        _native.host_check(_native.host().fmpi_host_barrier(grp.host_comm), "group barrier")


def barrier(comm=COMM_WORLD):
    """Host barrier (``MPI.Barrier``) over ``comm``."""
    grp = _group.resolve(comm)
    if grp is None:
        runtime.host_barrier()
    else:
        runtime.host_run(_group_barrier_now, grp)


def _check_root(root: int, grp=None):
    n = runtime.total_workers() if grp is None else grp.size
    if not (0 <= int(root) < n):
        raise ValueError(f"root {root} out of range for {n} workers")


# ------------------------------------------------------------------------------ non-blocking API


def Iallreduce(*args, comm=COMM_WORLD):
    """``Iallreduce(send, recv, op)`` or in-place ``Iallreduce(buf, op)`` → ``(recv, request)``.
    The values in ``recv`` are valid after ``request.wait()``."""
    if not runtime.Initialized():
        raise runtime.FluxMPINotInitializedError()
    if len(args) == 3:
        send, recv, op = args
    elif len(args) == 2:
        (recv, op), send = args, args[0]
    else:
        raise TypeError("Iallreduce(send, recv, op) or Iallreduce(buf, op)")
    grp = _group.resolve(comm)
    s, r = _as_tensor(send), _as_tensor(recv)
    _check_pair(s, r)
    code, fn = _resolve_op(op)
    _count("iallreduce", r)
    if fn is not None:                      # custom function: host path, queued like the others
        if r.is_cuda:
            _custom(s, r, fn, grp=grp)
            return recv, Request(buffer=recv)
        fut = runtime.host_submit(_host_custom_now, s.detach(), r.detach(), fn, -1, grp)
        return recv, Request(future=fut, buffer=recv)
    dtype_code(s.dtype)
    if r.is_cuda:
        ev = _device_world().allreduce_async(s, r, code, group=grp)
        return recv, Request(event=ev, buffer=recv)
    fut = runtime.host_submit(_host_allreduce_now, s, r, code, grp)
    return recv, Request(future=fut, buffer=recv)


def Ibcast(buf, root: int = 0, comm=COMM_WORLD):
    """Non-blocking broadcast → ``(buf, request)``."""
    if not runtime.Initialized():
        raise runtime.FluxMPINotInitializedError()
    grp = _group.resolve(comm)
    b = _as_tensor(buf)
    _check_root(root, grp)
    _count("ibcast", b)
    if b.is_cuda:
        ev = _device_world().bcast_async(b, root, group=grp)
        return buf, Request(event=ev, buffer=buf)
    fut = runtime.host_submit(_host_bcast_now, b, root, grp)
    return buf, Request(future=fut, buffer=buf)


# ------------------------------------------------------------------------------ symmetric tensors


def symm_empty(shape, dtype=torch.float32):
    """Collective allocation of a tensor in symmetric memory (every rank must call it with the
    same shape, in the same order). Collectives on such tensors run in place over NVLink with
    no staging copies (the analogue of MPI/NCCL registered buffers; the reference has none).
    On a CPU-only world or in NCCL-fallback mode this is a plain ``torch.empty``."""
    if not runtime.Initialized():
        raise runtime.FluxMPINotInitializedError()
    dw = runtime.state().device_world
    if dw is None or not dw.native:
        return torch.empty(shape, dtype=dtype, device=runtime.current_device())
    return dw.symm_empty(shape, dtype)


def symm_zeros(shape, dtype=torch.float32):
    t = symm_empty(shape, dtype)
    t.zero_()
    return t


def is_symmetric(t) -> bool:
    dw = runtime.state().device_world
    return bool(dw is not None and dw.native and isinstance(t, torch.Tensor) and t.is_cuda
                and dw.find_symmetric(t) is not None)


# Julia-style aliases so reference scripts translate 1:1 (`allreduce!` → `allreduce_b`).
allreduce = allreduce_
bcast = bcast_
reduce = reduce_
