"""The device world: symmetric memory mapped into every peer + the NVLink kernels on top.

This is what replaces ``MPI.COMM_WORLD`` for device tensors. The reference hands a CuArray to
MPI and (as written) stages it through the host both ways
(/root/reference/src/mpi_extensions.jl:97-106,119-128,141-150). Here every rank
owns one *symmetric segment* — ``[signal pads | all-reduce staging | broadcast staging]`` —
that all peers of the NVSwitch domain map (``csrc/symm.cpp``), and the collectives are single
kernels that read / write peer memory directly (``csrc/kernels/collectives.cu``).

Bring-up order (collective, inside ``Init``):
  1. topology JSON (``csrc/topo.cpp``)
  2. allocate the segment: VMM (+ multicast) → fall back to CUDA-IPC if any rank cannot
  3. exchange handles: fds over Unix sockets (VMM) or 64-byte IPC handles over the host world
  4. host barrier, build the ``fmpi_devcomm`` tables

If peer access is impossible (or ``disable_p2p_support()`` was persisted) the world runs in
*fallback* mode: device collectives go to NCCL through ``torch.distributed`` — the baseline
path, kept for A/B runs exactly like the reference's CUDA-aware toggle (src/FluxMPI.jl:46-56).

``EmulatedWorld`` runs W virtual ranks on ONE GPU (cooperative launch, rank = blockIdx.y) so
the very same device code — flags, barriers, peer pointers — is testable on a single B200.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import warnings
from typing import Optional, Sequence

import torch

from .. import _native
from .._native import DevComm, Launch, cuda_check
from ..comm import _TORCH_DTYPES as _DT

MB = 1 << 20


def _align(x: int, a: int) -> int:
    return (x + a - 1) // a * a


class _RawCudaMemory:
    """Expose a raw device pointer through ``__cuda_array_interface__`` so torch can alias it."""

    def __init__(self, ptr: int, nbytes: int, owner=None):
        self.owner = owner
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2, "strides": None}


def tensor_from_ptr(ptr: int, nbytes: int, device, owner=None) -> torch.Tensor:
    return torch.as_tensor(_RawCudaMemory(ptr, nbytes, owner), device=device)


def _dtype_code(dt):
    from ..comm import dtype_code
    return dtype_code(dt)


def _stream_ptr(stream=None) -> int:
    s = torch.cuda.current_stream() if stream is None else stream
    return s.cuda_stream


class Segment:
    """One symmetric allocation (local pointer + peers' mappings + optional multicast alias)."""

    def __init__(self, handle, local_ptr, size, peer_ptrs, mc_ptr, mode):
        self.handle = handle
        self.local_ptr = local_ptr
        self.size = size
        self.peer_ptrs = peer_ptrs
        self.mc_ptr = mc_ptr
        self.mode = mode


class _Base:
    """Shared launch helpers for the real and the emulated world."""

    world: int
    device: torch.device
    sig_bytes: int

    def _launch(self, blocks=0, threads=0, algo=_native.ALGO_AUTO):
        return C.byref(Launch(blocks, threads, algo))

    def _mk_comm(self, seg_data_ptrs, sig_ptrs, data_bytes, channel, rank, emulate, mc_ptr=0):
        c = DevComm()
        c.rank, c.world, c.emulate, c.channel = rank, self.world, emulate, channel
        for r in range(self.world):
            c.data[r] = seg_data_ptrs[r]
            c.sig[r] = sig_ptrs[r]
        c.mc_data = mc_ptr or None
        c.data_bytes = data_bytes
        c.ctrl = getattr(self, "_ctrl_dev", 0) or None      # health page (real worlds only)
        c.timeout_ns = int(getattr(self, "_timeout_ns", 0))
        return c


class DeviceWorld(_Base):
    # staging per signal channel: 0 = blocking collectives, 1 = non-blocking ones (their kernels
    # may overlap, so they must not share slots); the engine's channel 2 needs no staging
    AR_STAGE_MB = (int(os.environ.get("FMPI_AR_STAGE_MB", 160)), int(os.environ.get("FMPI_AR_ASYNC_STAGE_MB", 64)))
    BCAST_STAGE_MB = (int(os.environ.get("FMPI_BCAST_STAGE_MB", 96)), int(os.environ.get("FMPI_BCAST_ASYNC_STAGE_MB", 32)))
    TREE_SLOT_MB = 4
    GROUP_AR_MB, GROUP_BC_MB = 16, 8
    MAX_COMMS = 8            # signal-pad sets: 0 = the world, 1.. = rank groups (fm.Group)

    def __init__(self, rank: int, world: int, device: int, job: str):
        self.rank, self.world = rank, world
        self.device = torch.device("cuda", device)
        self.dev_index = device
        self.job = job
        self.native = False
        self.mode = "fallback"
        self.has_multicast = False
        self.topology: dict = {}
        self.segments: list[Segment] = []
        self.base: Optional[Segment] = None
        self.comm_stream: Optional[torch.cuda.Stream] = None
        self._pg = None
        self._tables: dict = {}
        self._closed = False
        self.why_fallback = ""
        self._heap = None                 # [segment, bump offset] for symm_empty
        self._symm_regs: list = []        # registered symmetric segments (user-visible tensors)
        self.tuner = None
        self._pack_arena = None           # symmetric target of tree-level pack kernels
        self._ctrl_host = None            # health page (host view) of this process's device
        self._ctrl_dev = 0
        self._timeout_ns = 0
        self._groups: dict = {}           # comm id -> Group state

    # ------------------------------------------------------------------ bring-up
    def _topology(self):
        lib = _native.cuda()
        buf = C.create_string_buffer(1 << 16)
        cuda_check(lib.fmpi_topo_json(buf, len(buf)), "topology")
        self.topology = json.loads(buf.value.decode())

    def _all_agree(self, ok: bool) -> bool:
        from .. import runtime
        flags = runtime.host_allgather_bytes(b"\1" if ok else b"\0", 1)
        return all(f == b"\1" for f in flags)

    def alloc_segment(self, nbytes: int, want_multicast: bool = True) -> Segment:
        """Collective: every rank allocates ``nbytes`` and maps everybody else's."""
        from .. import runtime
        lib = _native.cuda()
        host = _native.host()
        W, r = self.world, self.rank
        if W == 1:
            h = C.c_void_p()
            cuda_check(lib.fmpi_symm_create(self.dev_index, nbytes, _native.SYMM_LOCAL, 0, 1, C.byref(h)),
                       "symm_create")
            lib.fmpi_symm_set_self(h, 0)
            seg = Segment(h, lib.fmpi_symm_local_ptr(h), lib.fmpi_symm_size(h),
                          [lib.fmpi_symm_local_ptr(h)], 0, "local")
            self.segments.append(seg)
            return seg

        seg = None
        forced = os.environ.get("FMPI_SYMM_MODE", "").lower()
        if forced in ("", "vmm"):
            seg = self._alloc_vmm(nbytes, want_multicast)
        if seg is None and forced in ("", "ipc"):
            seg = self._alloc_ipc(nbytes)
        if seg is None:
            raise _native.NativeError("could not establish symmetric memory (VMM and IPC failed): "
                                      + self.why_fallback)
        self.segments.append(seg)
        return seg

    def _alloc_vmm(self, nbytes: int, want_multicast: bool) -> Optional[Segment]:
        from .. import runtime
        lib, host = _native.cuda(), _native.host()
        W, r = self.world, self.rank
        mc_want = bool(want_multicast and lib.fmpi_symm_mc_supported(self.dev_index)
                       and os.environ.get("FMPI_DISABLE_MULTICAST", "0") in ("0", ""))
        mc_want = self._all_agree(mc_want)
        h = C.c_void_p()
        ok = lib.fmpi_symm_create(self.dev_index, nbytes, _native.SYMM_VMM, int(mc_want), W, C.byref(h)) == 0
        fd = C.c_int(-1)
        if ok:
            ok = lib.fmpi_symm_export_fd(h, C.byref(fd)) == 0
        err = "" if ok else lib.fmpi_cuda_last_error().decode()
        if not self._all_agree(ok):
            self.why_fallback = f"VMM allocation/export failed on some rank ({err})"
            if h.value:
                lib.fmpi_symm_destroy(h)
            return None
        lib.fmpi_symm_set_self(h, r)
        tag = f"{self.job}-{len(self.segments)}".encode()
        sock = C.c_int(-1)
        ok = host.fmpi_fdx_open(tag, r, C.byref(sock)) == 0
        if not self._all_agree(ok):
            self.why_fallback = "fd-exchange socket could not be opened"
            lib.fmpi_symm_destroy(h)
            return None
        fds = (C.c_int * W)()
        ok = host.fmpi_fdx_exchange(sock, tag, r, W, fd.value, fds, 60.0) == 0
        err = "" if ok else host.fmpi_fdx_last_error().decode()
        if ok:
            for q in range(W):
                if q != r and lib.fmpi_symm_import_fd(h, q, fds[q]) != 0:
                    ok = False
                    err = lib.fmpi_cuda_last_error().decode()
                    break
        os.close(fd.value)
        if not self._all_agree(ok):
            self.why_fallback = f"VMM handle exchange/import failed ({err})"
            host.fmpi_fdx_close(sock, tag, r)
            lib.fmpi_symm_destroy(h)
            return None
        mc_ptr = 0
        if mc_want:
            mc_ptr = self._setup_multicast(h, sock, tag)
        host.fmpi_fdx_close(sock, tag, r)
        runtime.host_barrier()
        peers = [lib.fmpi_symm_peer_ptr(h, q) for q in range(W)]
        return Segment(h, lib.fmpi_symm_local_ptr(h), lib.fmpi_symm_size(h), peers, mc_ptr, "vmm")

    def _setup_multicast(self, h, sock, tag) -> int:
        """Returns the multicast VA or 0. Every step is agreed on by all ranks so nobody waits
        for a peer that gave up."""
        from .. import runtime
        lib, host = _native.cuda(), _native.host()
        W, r = self.world, self.rank
        fd = C.c_int(-1)
        ok = True
        if r == 0:
            ok = lib.fmpi_symm_mc_create(h, C.byref(fd)) == 0
        if not self._all_agree(ok):
            self.why_no_multicast = lib.fmpi_cuda_last_error().decode() if r == 0 else "root failed"
            return 0
        ok = host.fmpi_fdx_bcast(sock, tag, r, W, 0, C.byref(fd), 60.0) == 0
        if ok and r != 0:
            ok = lib.fmpi_symm_mc_import(h, fd.value) == 0
        elif ok and r == 0:
            os.close(fd.value)
        if not self._all_agree(ok):
            return 0
        ok = lib.fmpi_symm_mc_add_device(h) == 0
        if not self._all_agree(ok):
            return 0
        ok = lib.fmpi_symm_mc_bind_map(h) == 0
        if not ok:
            self.why_no_multicast = lib.fmpi_cuda_last_error().decode()
        if not self._all_agree(ok):
            return 0
        return lib.fmpi_symm_mc_ptr(h) or 0

    def _alloc_ipc(self, nbytes: int) -> Optional[Segment]:
        from .. import runtime
        lib = _native.cuda()
        W, r = self.world, self.rank
        h = C.c_void_p()
        ok = lib.fmpi_symm_create(self.dev_index, nbytes, _native.SYMM_IPC, 0, W, C.byref(h)) == 0
        blob = C.create_string_buffer(64)
        if ok:
            ok = lib.fmpi_symm_export_ipc(h, blob) == 0
        err = "" if ok else lib.fmpi_cuda_last_error().decode()
        if not self._all_agree(ok):
            self.why_fallback += f"; IPC allocation/export failed ({err})"
            if h.value:
                lib.fmpi_symm_destroy(h)
            return None
        lib.fmpi_symm_set_self(h, r)
        blobs = runtime.host_allgather_bytes(blob.raw, 64)
        for q in range(W):
            if q != r and lib.fmpi_symm_import_ipc(h, q, blobs[q]) != 0:
                ok = False
                err = lib.fmpi_cuda_last_error().decode()
                break
        if not self._all_agree(ok):
            self.why_fallback += f"; IPC import failed ({err})"
            lib.fmpi_symm_destroy(h)
            return None
        runtime.host_barrier()
        peers = [lib.fmpi_symm_peer_ptr(h, q) for q in range(W)]
        return Segment(h, lib.fmpi_symm_local_ptr(h), lib.fmpi_symm_size(h), peers, 0, "ipc")

    def bring_up(self, disabled: bool, verbose: bool):
        lib = _native.cuda()      # raises if the library is missing: no silent fallback
        self._topology()
        self.pad_bytes = _align(lib.fmpi_sigpad_bytes(), 1 << 16)      # one set of signal pads
        self.sig_bytes = self.pad_bytes * self.MAX_COMMS                # all sets, then the data window
        off = 0
        self.ar_regions, self.bc_regions = [], []
        for mb in self.AR_STAGE_MB:
            self.ar_regions.append((off, mb * MB))
            off += mb * MB
        for mb in self.BCAST_STAGE_MB:
            self.bc_regions.append((off, mb * MB))
            off += mb * MB
        # push slots of the one-launch tree step (ops/fused.py): [2 parities][W][4 MB], private to
        # that kernel family because peers write into them without a handshake
        self.tree_region = (off, 2 * self.world * self.TREE_SLOT_MB * MB)
        off += self.tree_region[1]
        # every rank group gets staging of its own: a non-member may be in a world collective while
        # the members are in a group collective, and neither may push into the other's slots
        self.group_regions = {}
        for gid in range(1, self.MAX_COMMS):
            self.group_regions[gid] = ((off, self.GROUP_AR_MB * MB), (off + self.GROUP_AR_MB * MB, self.GROUP_BC_MB * MB))
            off += (self.GROUP_AR_MB + self.GROUP_BC_MB) * MB
        self._window_bytes = off
        self.ar_off, self.ar_bytes = self.ar_regions[0]
        self.bc_off, self.bc_bytes = self.bc_regions[0]
        # health page: lets a kernel give up on a lost peer and lets the host stop the kernels
        hp, dp = C.c_void_p(), C.c_void_p()
        cuda_check(lib.fmpi_health_page(C.byref(hp), C.byref(dp)), "health page")
        self._ctrl_host = (C.c_uint32 * 16).from_address(hp.value)
        self._ctrl_dev = dp.value
        from .. import runtime as _rt
        self._timeout_ns = int(float(os.environ.get("FMPI_DEVICE_TIMEOUT_S", _rt.state().timeout_s)) * 1e9)
        self.comm_stream = torch.cuda.Stream(self.device, priority=-1)
        if disabled and self.world > 1:
            self.why_fallback = "P2P support disabled by preference"
            return
        if os.environ.get("FMPI_FORCE_FALLBACK", "0") not in ("0", ""):
            self.why_fallback = "FMPI_FORCE_FALLBACK is set"     # A/B runs and tests, any world size
            return
        if getattr(self, "nnodes", 1) > 1:
            # peer mappings and multicast objects exist inside one NVSwitch domain only; a job
            # that spans nodes sends its device tensors through NCCL (which is itself
            # NVLink-inside / network-between) and its host tensors through the leader mesh
            self.why_fallback = f"the job spans {self.nnodes} nodes and the NVLink kernels are node-local"
            return
        try:
            need = self.sig_bytes + self._window_bytes
            if self.world == 1:
                need = self.sig_bytes + 4 * MB
                self.ar_regions = [(0, MB), (MB, MB)]
                self.bc_regions = [(2 * MB, MB), (3 * MB, MB)]
                self.tree_region = (0, 0)
                self.ar_off, self.ar_bytes = self.ar_regions[0]
                self.bc_off, self.bc_bytes = self.bc_regions[0]
            self.base = self.alloc_segment(need)
        except _native.NativeError as e:
            if self.world > 1 and os.environ.get("FMPI_REQUIRE_P2P", "0") not in ("0", ""):
                raise
            self.why_fallback = str(e)
            warnings.warn(f"[fluxmpi_b200] NVLink P2P unavailable, falling back to NCCL: {e}")
            return
        self.native = True
        self.mode = self.base.mode
        self.has_multicast = bool(self.base.mc_ptr)
        self.comms = [self.comm_for(self.base, ch, data_offset=self.sig_bytes,
                                    data_bytes=self.base.size - self.sig_bytes)
                      for ch in range(_native.NUM_CHANNELS)]
        from .tuner import Tuner
        self.tuner = Tuner(self.world, self.has_multicast)
        # call-path caches: the eager collectives are latency-sensitive on the host side too
        self._lib = lib
        self._comm_refs = [C.byref(c) for c in self.comms]
        self._lp_auto_obj = Launch(0, 0, _native.ALGO_AUTO)
        self._lp_auto = C.byref(self._lp_auto_obj)
        self._in1, self._out1 = C.c_void_p(), C.c_void_p()
        self._in1_ref, self._out1_ref = C.byref(self._in1), C.byref(self._out1)

    def comm_for(self, seg: Segment, channel: int, data_offset: int = 0,
                 data_bytes: Optional[int] = None, comm_id: int = 0, members=None) -> DevComm:
        """Device communicator over ``seg`` (signal pads always come from the base segment; set
        ``comm_id`` of them). ``members`` (world ranks, ascending) restricts it to a rank group:
        the tables then hold only the members' pointers and ranks are group ranks."""
        base = self.base if self.base is not None else seg
        members = list(range(self.world)) if members is None else list(members)
        data = [seg.peer_ptrs[r] + data_offset for r in members]
        sig = [base.peer_ptrs[r] + comm_id * self.pad_bytes for r in members]
        whole = len(members) == self.world
        mc = seg.mc_ptr + data_offset if (seg.mc_ptr and whole) else 0   # the multicast object spans the world
        nbytes = (seg.size - data_offset) if data_bytes is None else data_bytes
        c = DevComm()
        c.rank, c.world, c.emulate, c.channel = members.index(self.rank), len(members), 0, channel
        for i in range(len(members)):
            c.data[i] = data[i]
            c.sig[i] = sig[i]
        c.mc_data = mc or None
        c.data_bytes = nbytes
        c.ctrl = self._ctrl_dev or None
        c.timeout_ns = self._timeout_ns
        return c

    def check_health(self):
        """Raise if a kernel gave up waiting for a peer (or the world was aborted)."""
        h = self._ctrl_host
        if h is not None and (h[0] or h[1]):
            rc = self._lib.fmpi_health_check() if hasattr(self, "_lib") else _native.cuda().fmpi_health_check()
            cuda_check(rc or 1, "device world")

    def abort(self):
        """Make every cross-GPU wait on this GPU return (called when the host world reports a
        dead peer): the kernels in flight finish at once with undefined results."""
        if self._ctrl_host is not None:
            self._ctrl_host[0] = 1

    # ------------------------------------------------------------------ NCCL fallback
    def _nccl(self):
        import torch.distributed as dist
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            dist.init_process_group("nccl", rank=self.rank, world_size=self.world,
                                    device_id=self.device)
        return dist

    _DIST_OPS = None

    def _dist_op(self, code):
        import torch.distributed as dist
        return {_native.SUM: dist.ReduceOp.SUM, _native.PROD: dist.ReduceOp.PRODUCT,
                _native.MIN: dist.ReduceOp.MIN, _native.MAX: dist.ReduceOp.MAX}[code]

    # ------------------------------------------------------------------ helpers
    @staticmethod
    def _work_tensor(t: torch.Tensor):
        """(contiguous 16-byte-aligned alias-or-copy viewed without bool, needs_copy_back)"""
        w = t.detach() if t.requires_grad else t
        copy = False
        if not w.is_contiguous():
            w, copy = w.contiguous(), True
        elif w.data_ptr() % 16:
            w, copy = w.clone(), True     # fresh allocations are 512-byte aligned
        if w.dtype == torch.bool:
            w = w.view(torch.uint8)
        return w, copy

    def _reduce_native(self, src: torch.Tensor, dst: torch.Tensor, op: int, root: int, channel: int,
                       stream=None, launch=None):
        lib = self._lib
        comm = self._comm_refs[channel]
        st = (torch.cuda.current_stream() if stream is None else stream).cuda_stream
        code = _DT.get(src.dtype)
        if code is None:
            code = _dtype_code(src.dtype)
        es = src.element_size()
        n = src.numel()
        nbytes = n * es
        bulk = n if nbytes % 16 == 0 or nbytes <= (200 << 10) else (nbytes // 16 * 16) // es
        if launch is not None:
            lp = launch
        else:
            tl = self.tuner.launch("allreduce_any" if root < 0 else "reduce_any", nbytes)
            if tl and tl[1].algo == _native.ALGO_NVLS and not (
                    op == _native.SUM and src.dtype in (torch.float32, torch.bfloat16, torch.float16)):
                tl = None
            lp = tl[0] if tl else self._lp_auto
        i, o = self._in1, self._out1     # reusable one-entry pointer tables (read at launch)
        sp, dp = src.data_ptr(), dst.data_ptr()
        ar_off, ar_bytes = self.ar_regions[1 if channel == 1 else 0]
        off = 0
        for cnt in ((bulk, n - bulk) if bulk < n else (bulk,)):
            if cnt:
                i.value, o.value = sp + off * es, dp + off * es
                if root < 0:
                    rc = lib.fmpi_allreduce_any(comm, self._in1_ref, self._out1_ref, cnt, code, op,
                                                ar_off, ar_bytes, lp, st)
                else:
                    rc = lib.fmpi_reduce_any(comm, self._in1_ref, self._out1_ref, cnt, code, op, root,
                                             ar_off, ar_bytes, lp, st)
                if rc:
                    cuda_check(rc, "allreduce" if root < 0 else "reduce")
                _native.launch_count += 1
            off += cnt

    # ------------------------------------------------------------------ collectives (stream-ordered)
    # ------------------------------------------------------------------ symmetric tensors
    def symm_empty(self, shape, dtype=torch.float32) -> torch.Tensor:
        """Collective: a tensor in symmetric memory (same offset on every rank). Collectives on
        such tensors run in place over NVLink with no staging copies — the analogue of NCCL's
        registered user buffers. Every rank must make the same sequence of symm_empty calls."""
        numel = 1
        for s in (shape if isinstance(shape, (tuple, list, torch.Size)) else (shape,)):
            numel *= int(s)
        nbytes = _align(max(numel, 1) * torch.empty((), dtype=dtype).element_size(), 256)
        heap = self._heap
        if heap is None or heap[1] + nbytes > heap[0].size:
            size = max(nbytes, int(os.environ.get("FMPI_SYMM_HEAP_MB", 256)) * MB)
            seg = self.alloc_segment(size)
            self._register(seg)
            heap = self._heap = [seg, 0]
        seg, off = heap
        heap[1] = off + nbytes
        raw = self.symm_tensor(seg, off, numel * torch.empty((), dtype=dtype).element_size())
        return raw.view(dtype).view(shape)

    def _register(self, seg: Segment):
        comms = [self.comm_for(seg, ch) for ch in range(_native.NUM_CHANNELS)]
        self._symm_regs.append((seg.local_ptr, seg.local_ptr + seg.size, seg, comms,
                                [C.byref(c) for c in comms]))

    def find_symmetric(self, t: torch.Tensor):
        """(comm refs, byte offset, has multicast) if ``t`` lives in a registered symmetric
        segment."""
        p = t.data_ptr()
        for lo, hi, seg, comms, refs in self._symm_regs:
            if lo <= p < hi and p + t.numel() * t.element_size() <= hi:
                return refs, p - lo, bool(comms[0].mc_data)
        return None

    def _symm_collective(self, kind: str, t: torch.Tensor, hit, op: int, root: int, channel: int, stream):
        refs, off, has_mc = hit
        lib = self._lib
        nbytes = t.numel() * t.element_size()
        st = (torch.cuda.current_stream() if stream is None else stream).cuda_stream
        tl = self.tuner.launch("bcast_symm" if kind == "bcast" else "allreduce_symm", nbytes)
        if tl and tl[1].algo == _native.ALGO_NVLS and not has_mc:
            tl = None       # no multicast alias on this segment: peer-to-peer kernels
        if tl and kind != "bcast" and tl[1].algo == _native.ALGO_NVLS and not (
                op == _native.SUM and t.dtype in (torch.float32, torch.bfloat16, torch.float16)):
            tl = None       # the switch only adds f32/f16/bf16: other ops/dtypes take the P2P kernel
        lp = tl[0] if tl else self._lp_auto
        if kind == "bcast":
            rc = lib.fmpi_bcast_symm(refs[channel], off, nbytes, root, lp, st)
        elif root < 0:
            rc = lib.fmpi_allreduce_symm(refs[channel], off, t.numel(), _DT[t.dtype], op, lp, st)
        else:
            rc = lib.fmpi_reduce_symm(refs[channel], off, t.numel(), _DT[t.dtype], op, root, lp, st)
        if rc:
            cuda_check(rc, kind)
        _native.launch_count += 1

    def _symm_hit(self, t: torch.Tensor):
        if not self._symm_regs or not t.is_contiguous() or t.dtype == torch.bool:
            return None
        hit = self.find_symmetric(t)
        if hit is None or hit[1] % 16 or (t.numel() * t.element_size()) % 16:
            return None
        return hit

    # ------------------------------------------------------------------ rank groups
    def _group_call(self, kind: str, t: torch.Tensor, op: int, root: int, group, channel: int = 0, stream=None):
        """A collective over a rank subset (fm.Group). Native worlds: the staged kernels on the
        group's own communicator — its signal pads, the members' windows only, and a staging region
        of its own, so it is independent of whatever the world or another group is doing on the same
        GPUs. Otherwise: through the group's host world."""
        if group.size == 1 or t.numel() == 0:
            return t
        if not self.native:
            h = t.detach().cpu()
            from .. import comm as _comm
            {"allreduce": lambda: _comm._host_allreduce_now(h, h, op, group),
             "reduce": lambda: _comm._host_reduce_now(h, h, op, root, group),
             "bcast": lambda: _comm._host_bcast_now(h, root, group)}[kind]()
            if kind == "allreduce" or (kind == "reduce") == (group.rank == root):
                self._copy_back(t, h.to(t.device))
            return t
        self.check_health()
        lib = self._lib
        ref = group.device_refs()[channel]
        st = (torch.cuda.current_stream() if stream is None else stream).cuda_stream
        w, copy = self._work_tensor(t)
        nbytes = w.numel() * w.element_size()
        (g_ar_off, g_ar_bytes), (g_bc_off, g_bc_bytes) = self.group_regions[group.id]
        half = 1 if channel == 1 else 0                 # blocking / non-blocking calls: separate halves
        if kind == "bcast":
            bc_off, bc_bytes = g_bc_off + half * (g_bc_bytes // 2), g_bc_bytes // 2
            self._in1.value = w.data_ptr()
            rc = lib.fmpi_bcast_any(ref, self._in1_ref, nbytes, root, bc_off, bc_bytes, self._lp_auto, st)
            cuda_check(rc, "group bcast")
        else:
            ar_off, ar_bytes = g_ar_off + half * (g_ar_bytes // 2), g_ar_bytes // 2
            es, n = w.element_size(), w.numel()
            code = _DT[w.dtype]
            bulk = n if nbytes % 16 == 0 or nbytes <= (200 << 10) else (nbytes // 16 * 16) // es
            off = 0
            for cnt in ((bulk, n - bulk) if bulk < n else (bulk,)):
                if cnt:
                    self._in1.value = self._out1.value = w.data_ptr() + off * es
                    if kind == "allreduce":
                        rc = lib.fmpi_allreduce_any(ref, self._in1_ref, self._out1_ref, cnt, code, op, ar_off,
                                                    ar_bytes, self._lp_auto, st)
                    else:
                        rc = lib.fmpi_reduce_any(ref, self._in1_ref, self._out1_ref, cnt, code, op, root, ar_off,
                                                 ar_bytes, self._lp_auto, st)
                    cuda_check(rc, "group " + kind)
                off += cnt
        _native.launch_count += 1
        if copy and (kind == "allreduce" or (kind == "reduce") == (group.rank == root)):
            self._copy_back(t, w)
        return t

    def allreduce(self, t: torch.Tensor, op: int, launch=None, group=None):
        if group is not None:
            return self._group_call("allreduce", t, op, -1, group)
        if self.world == 1 or t.numel() == 0:
            return t
        if self.native:
            self.check_health()
        if self.native and launch is None:
            hit = self._symm_hit(t)
            if hit is not None:
                self._symm_collective("allreduce", t.detach() if t.requires_grad else t, hit, op, -1, 0, None)
                return t
        w, copy = self._work_tensor(t)
        if self.native:
            if launch is None and self._arena_ok(w):
                self._allreduce_through_arena(w, op)
            else:
                self._reduce_native(w, w, op, -1, 0, launch=launch)
        else:
            self._nccl().all_reduce(w, op=self._dist_op(op))
        if copy:
            self._copy_back(t, w)
        return t

    # Opt-in (FMPI_ARENA_ALLREDUCE_MB=<threshold>): a large ordinary tensor is copied into the
    # symmetric arena, all-reduced there in place by the in-switch / two-shot kernel and copied
    # back — three stream-ordered launches over pieces that are each measured: at 8 GPUs the
    # in-place kernel moves 256 MB in 578 us against ~940 us for the single staged kernel, and the
    # two local copies cost ~85 us each. Pieces larger than the arena are done one after another.
    ARENA_MIN_BYTES = int(float(os.environ.get("FMPI_ARENA_ALLREDUCE_MB", 0)) * MB)
    ARENA_MAX_BYTES = int(float(os.environ.get("FMPI_ARENA_MAX_MB", 512)) * MB)

    def _arena_ok(self, w: torch.Tensor) -> bool:
        n = w.numel() * w.element_size()
        return (self.ARENA_MIN_BYTES > 0 and self.world > 1 and n >= self.ARENA_MIN_BYTES and n % 16 == 0
                and w.dtype != torch.bool)

    def _allreduce_through_arena(self, w: torch.Tensor, op: int):
        flat = w.view(-1)
        es = w.element_size()
        piece = self.ARENA_MAX_BYTES // es // 16 * 16
        arena, _refs, _off = self.pack_arena(min(flat.numel() * es, self.ARENA_MAX_BYTES))
        for lo in range(0, flat.numel(), piece):
            part = flat[lo:lo + piece]
            view = arena[:part.numel() * es].view(w.dtype)
            view.copy_(part)
            self._symm_collective("allreduce", view, self.find_symmetric(view), op, -1, 0, None)
            part.copy_(view)

    def reduce(self, t: torch.Tensor, op: int, root: int, group=None):
        if group is not None:
            return self._group_call("reduce", t, op, root, group)
        if self.world == 1 or t.numel() == 0:
            return t
        if self.native:
            hit = self._symm_hit(t)
            if hit is not None:
                self._symm_collective("reduce", t.detach() if t.requires_grad else t, hit, op, root, 0, None)
                return t
        w, copy = self._work_tensor(t)
        if self.native:
            self._reduce_native(w, w, op, root, 0)
        else:
            tmp = w.clone()
            self._nccl().reduce(tmp, dst=root, op=self._dist_op(op))
            if self.rank == root:
                w.copy_(tmp)
        if copy and self.rank == root:
            self._copy_back(t, w)
        return t

    def bcast(self, t: torch.Tensor, root: int, channel: int = 0, stream=None, group=None):
        if group is not None:
            return self._group_call("bcast", t, 0, root, group, channel, stream)
        if self.world == 1 or t.numel() == 0:
            return t
        if self.native:
            hit = self._symm_hit(t)
            if hit is not None:
                self._symm_collective("bcast", t.detach() if t.requires_grad else t, hit, 0, root, channel, stream)
                return t
        w, copy = self._work_tensor(t)
        if self.native:
            lib = self._lib
            nbytes = w.numel() * w.element_size()
            tl = self.tuner.launch("bcast_any", nbytes)
            self._in1.value = w.data_ptr()
            bc_off, bc_bytes = self.bc_regions[1 if channel == 1 else 0]
            rc = lib.fmpi_bcast_any(self._comm_refs[channel], self._in1_ref, nbytes, root, bc_off,
                                    bc_bytes, tl[0] if tl else self._lp_auto,
                                    (torch.cuda.current_stream() if stream is None else stream).cuda_stream)
            if rc:
                cuda_check(rc, "bcast")
            _native.launch_count += 1
        else:
            self._nccl().broadcast(w, src=root)
        if copy and self.rank != root:
            self._copy_back(t, w)
        return t

    @staticmethod
    def _copy_back(t, w):
        with torch.no_grad():
            tt = t.detach() if t.requires_grad else t
            tt.copy_(w.view(torch.bool) if t.dtype == torch.bool else w)

    # ------------------------------------------------------------------ non-blocking
    def allreduce_async(self, send: torch.Tensor, recv: torch.Tensor, op: int, group=None) -> torch.cuda.Event:
        ev = torch.cuda.Event()
        if group is not None:
            cs = self.comm_stream
            cs.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(cs):
                if recv.data_ptr() != send.data_ptr():
                    recv.copy_(send)
                self._group_call("allreduce", recv, op, -1, group, 1, cs)
                for x in (send, recv):
                    x.record_stream(cs)
                ev.record(cs)
            return ev
        if self.world == 1 or send.numel() == 0:
            if recv.data_ptr() != send.data_ptr():
                recv.copy_(send)
            ev.record()
            return ev
        cs = self.comm_stream
        cs.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(cs):
            s, _ = self._work_tensor(send)
            r, copy = self._work_tensor(recv)
            if self.native:
                self._reduce_native(s, r, op, -1, 1, stream=cs)
            else:
                if r.data_ptr() != s.data_ptr():
                    r.copy_(s)
                self._nccl().all_reduce(r, op=self._dist_op(op))
            if copy:
                self._copy_back(recv, r)
            for x in (send, recv):
                x.record_stream(cs)
            ev.record(cs)
        return ev

    def bcast_async(self, buf: torch.Tensor, root: int, group=None) -> torch.cuda.Event:
        ev = torch.cuda.Event()
        if (self.world == 1 and group is None) or buf.numel() == 0:
            ev.record()
            return ev
        cs = self.comm_stream
        cs.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(cs):
            self.bcast(buf, root, channel=1, stream=cs, group=group)
            buf.record_stream(cs)
            ev.record(cs)
        return ev

    # ------------------------------------------------------------------ whole-tree helpers
    def _table(self, ts: Sequence[torch.Tensor]):
        """Device pointer table for a list of leaves (cached: training loops reuse the leaves)."""
        key = tuple((t.data_ptr(), t.numel() * t.element_size()) for t in ts)
        hit = self._tables.get(key)
        if hit is not None:
            return hit
        offs, total = [], 0
        for t in ts:
            offs.append(total)
            total += _align(t.numel() * t.element_size(), 16)
        n = len(ts)
        host = torch.tensor([t.data_ptr() for t in ts] + offs + [k[1] for k in key],
                            dtype=torch.int64).pin_memory()
        dev = host.to(self.device, non_blocking=True)
        need = _align(max(total, 16), 256)
        if self.native and self.world > 1:
            # pack straight into symmetric memory: the collective then runs in place over NVLink
            # (every rank packs the same tree, so the collective allocation below stays in step)
            if self._pack_arena is None or self._pack_arena.numel() < need:
                self._pack_arena = self.symm_empty((max(need, 64 * MB),), torch.uint8)
            flat = self._pack_arena[:need]
        else:
            flat = torch.empty(need, dtype=torch.uint8, device=self.device)
        entry = (dev, flat, total, n)
        if len(self._tables) > 64:
            self._tables.clear()
        self._tables[key] = entry
        return entry

    def pack_arena(self, nbytes: int):
        """(uint8 tensor of >= nbytes in symmetric memory, comm refs of its segment, byte offset of
        the tensor in that segment) — the target of the tree-level gather kernels. Collective the
        first time (and whenever it has to grow): every rank walks the same tree."""
        need = _align(max(nbytes, 16), 256)
        if self._pack_arena is None or self._pack_arena.numel() < need:
            self._pack_arena = self.symm_empty((max(need, 64 * MB),), torch.uint8)
        refs, off, _ = self.find_symmetric(self._pack_arena)
        return self._pack_arena, refs, off

    def _pack(self, ts, unpack: bool):
        lib = _native.cuda()
        dev, flat, total, n = self._table(ts)
        base = dev.data_ptr()
        fn = lib.fmpi_unpack if unpack else lib.fmpi_pack
        cuda_check(fn(n, base, base + 8 * n, base + 16 * n, flat.data_ptr(), _stream_ptr()), "pack")
        _native.count_launch()
        return flat, total

    def allreduce_many(self, ts: Sequence[torch.Tensor], op: int):
        """One collective for a list of same-dtype leaves: pack → all-reduce → unpack."""
        ts = [t.detach() if t.requires_grad else t for t in ts]
        if any(not t.is_contiguous() for t in ts):
            for t in ts:
                self.allreduce(t, op)
            return
        if (self.native and self.world > 1 and ts[0].dtype in (torch.float32, torch.bfloat16, torch.float16)
                and all(t.dtype == ts[0].dtype for t in ts)):
            # two launches: gather the leaves into the symmetric arena and all-reduce it in place
            # (csrc/kernels/tree_step.cu), then scatter the result back into the leaves
            from ..ops import fused as _F
            self.check_health()
            key = ("tree",) + tuple((t.data_ptr(), t.numel()) for t in ts)
            tab = self._tables.get(key)
            if tab is None:
                if len(self._tables) > 64:
                    self._tables.clear()
                tab = _F.TreeTable(ts, ts, None, None)
                self._tables[key] = tab
            arena, refs, off = self.pack_arena(tab.total)
            cuda_check(self._lib.fmpi_tree_allreduce(refs[0], tab.n, tab.ref, off, tab.total, _DT[ts[0].dtype], op,
                                                     self._lp_auto, _stream_ptr()), "tree_allreduce")
            _native.count_launch()
            if getattr(tab, "unpack_tab", None) is None or tab.unpack_arena != arena.data_ptr():
                es = ts[0].element_size()
                tab.unpack_tab = torch.tensor([t.data_ptr() for t in ts] + tab.offs + [t.numel() * es for t in ts],
                                              dtype=torch.int64).pin_memory().to(self.device, non_blocking=True)
                tab.unpack_arena = arena.data_ptr()
            b, n = tab.unpack_tab.data_ptr(), tab.n
            cuda_check(self._lib.fmpi_unpack(n, b, b + 8 * n, b + 16 * n, arena.data_ptr(), _stream_ptr()), "unpack")
            _native.count_launch()
            return
        flat, total = self._pack(ts, unpack=False)
        view = flat[:total].view(ts[0].dtype if ts[0].dtype != torch.bool else torch.uint8)
        self.allreduce(view, op)
        self._pack(ts, unpack=True)

    def bcast_many(self, ts: Sequence[torch.Tensor], root: int):
        ts = [t.detach() if t.requires_grad else t for t in ts]
        if any(not t.is_contiguous() for t in ts):
            for t in ts:
                self.bcast(t, root)
            return
        if self.rank == root:
            flat, total = self._pack(ts, unpack=False)
        else:
            _, flat, total, _ = self._table(ts)
        self.bcast(flat[:max(total, 1)], root)
        if self.rank != root:
            self._pack(ts, unpack=True)

    def barrier(self, stream=None):
        if self.world > 1 and self.native:
            self.check_health()
            cuda_check(_native.cuda().fmpi_barrier_dev(C.byref(self.comms[0]), _stream_ptr(stream)), "barrier")
            _native.count_launch()

    def symm_tensor(self, seg: Segment, offset: int, nbytes: int) -> torch.Tensor:
        """uint8 tensor aliasing ``[offset, offset+nbytes)`` of this rank's part of ``seg``."""
        return tensor_from_ptr(seg.local_ptr + offset, nbytes, self.device, owner=seg)

    def close(self):
        if self._closed:
            return
        self._closed = True
        try:
            torch.cuda.synchronize(self.device)
        except Exception:
            pass
        if _native.cuda_loaded():
            lib = _native.cuda()
            for seg in reversed(self.segments):
                lib.fmpi_symm_destroy(seg.handle)
        self.segments.clear()


def create(rank: int, world: int, device: int, job: str, disabled: bool = False,
           window_bytes: Optional[int] = None, verbose: bool = False, nnodes: int = 1) -> DeviceWorld:
    dw = DeviceWorld(rank, world, device, job)
    dw.nnodes = nnodes
    dw.bring_up(disabled, verbose)
    if verbose:
        from .. import runtime
        runtime.fluxmpi_println(
            f"device world: mode={dw.mode} multicast={dw.has_multicast} "
            f"{'(' + dw.why_fallback + ')' if dw.why_fallback else ''}")
    return dw


class EmulatedWorld(_Base):
    """W virtual ranks on one GPU. Every "rank" gets its own segment (signal pad + data window)
    in local memory; kernels are launched cooperatively with gridDim.y = W."""

    def __init__(self, world: int, device=None, window_mb: int = 64, stage_mb: int = 16):
        lib = _native.cuda()
        self.world = world
        self.device = torch.device(device if device is not None else torch.cuda.current_device())
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.sig_bytes = _align(lib.fmpi_sigpad_bytes(), 1 << 16)
        # bounded waits here too: a protocol bug ends in an error, not in a kernel that spins
        hp, dp = C.c_void_p(), C.c_void_p()
        with torch.cuda.device(self.device):
            cuda_check(lib.fmpi_health_page(C.byref(hp), C.byref(dp)), "health page")
        self._ctrl_host = (C.c_uint32 * 16).from_address(hp.value)
        self._ctrl_dev = dp.value
        self._timeout_ns = int(float(os.environ.get("FMPI_DEVICE_TIMEOUT_S", 20)) * 1e9)
        self.ar_off, self.ar_bytes = 0, stage_mb * MB
        self.bc_off, self.bc_bytes = self.ar_bytes, stage_mb * MB
        self.heap_off = self.bc_off + self.bc_bytes
        self.window_bytes = self.heap_off + window_mb * MB
        total = self.sig_bytes + self.window_bytes
        self.mem = [torch.zeros(total, dtype=torch.uint8, device=self.device) for _ in range(world)]
        base = [m.data_ptr() for m in self.mem]
        assert all(b % 256 == 0 for b in base)
        self.comms = [self._mk_comm([b + self.sig_bytes for b in base], base, self.window_bytes, ch, 0, 1)
                      for ch in range(_native.NUM_CHANNELS)]
        self._heap_top = self.heap_off

    def check_health(self):
        if self._ctrl_host[0] or self._ctrl_host[1]:
            cuda_check(_native.cuda().fmpi_health_check() or 1, "emulated world")

    # window helpers ------------------------------------------------------------------
    def window(self, rank: int, offset: int, nbytes: int) -> torch.Tensor:
        s = self.sig_bytes + offset
        return self.mem[rank][s:s + nbytes]

    def heap_alloc(self, nbytes: int) -> int:
        off = _align(self._heap_top, 256)
        if off + nbytes > self.window_bytes:
            raise MemoryError("emulated window exhausted")
        self._heap_top = off + nbytes
        return off

    def _ptrs(self, ts):
        return _native.ptr_array([t.data_ptr() for t in ts])

    # collectives: one tensor per virtual rank -------------------------------------------
    def allreduce(self, ts: Sequence[torch.Tensor], op: int, outs=None, algo=_native.ALGO_AUTO,
                  blocks=0, root: int = -1):
        lib = _native.cuda()
        outs = ts if outs is None else outs
        t0 = ts[0]
        code = _dtype_code(t0.dtype)
        i, o = self._ptrs(ts), self._ptrs(outs)
        lp = self._launch(blocks, 0, algo)
        if root < 0:
            rc = lib.fmpi_allreduce_any(C.byref(self.comms[0]), i, o, t0.numel(), code, op,
                                        self.ar_off, self.ar_bytes, lp, _stream_ptr())
        else:
            rc = lib.fmpi_reduce_any(C.byref(self.comms[0]), i, o, t0.numel(), code, op, root,
                                     self.ar_off, self.ar_bytes, lp, _stream_ptr())
        cuda_check(rc, "emulated allreduce")
        _native.count_launch()

    def bcast(self, ts: Sequence[torch.Tensor], root: int, algo=_native.ALGO_AUTO, blocks=0):
        lib = _native.cuda()
        t0 = ts[0]
        cuda_check(lib.fmpi_bcast_any(C.byref(self.comms[0]), self._ptrs(ts),
                                      t0.numel() * t0.element_size(), root, self.bc_off, self.bc_bytes,
                                      self._launch(blocks, 0, algo), _stream_ptr()), "emulated bcast")
        _native.count_launch()

    def allreduce_symm(self, offset: int, count: int, dtype: torch.dtype, op: int, root: int = -1,
                       blocks=0):
        lib = _native.cuda()
        code = _dtype_code(dtype)
        lp = self._launch(blocks)
        if root < 0:
            rc = lib.fmpi_allreduce_symm(C.byref(self.comms[0]), offset, count, code, op, lp, _stream_ptr())
        else:
            rc = lib.fmpi_reduce_symm(C.byref(self.comms[0]), offset, count, code, op, root, lp, _stream_ptr())
        cuda_check(rc, "emulated symm allreduce")
        _native.count_launch()

    def bcast_symm(self, offset: int, nbytes: int, root: int, algo=_native.ALGO_AUTO, blocks=0):
        lib = _native.cuda()
        cuda_check(lib.fmpi_bcast_symm(C.byref(self.comms[0]), offset, nbytes, root,
                                       self._launch(blocks, 0, algo), _stream_ptr()), "emulated symm bcast")
        _native.count_launch()

    def barrier(self):
        cuda_check(_native.cuda().fmpi_barrier_dev(C.byref(self.comms[0]), _stream_ptr()), "barrier")
        _native.count_launch()

    def fused_step(self, g_off: int, p_off: int, dtype: torch.dtype, m, v, master, lo: int, hi: int,
                   hyper, step_dev, blocks: int = 0, channel: int = 2):
        """Fused all-reduce + optimiser step over the emulated ranks. ``m``/``v``/``master``/
        ``step_dev`` are per-rank lists (or None)."""
        lib = _native.cuda()
        code = _dtype_code(dtype)
        arr = lambda ts: self._ptrs(ts) if ts is not None else None
        cuda_check(lib.fmpi_fused_step(C.byref(self.comms[channel]), g_off, code, p_off, code, arr(m),
                                       arr(v), arr(master), lo, hi, C.byref(hyper), arr(step_dev),
                                       self._launch(blocks), _stream_ptr()), "emulated fused_step")
        _native.count_launch()

    def fused_step_oneshot(self, g_off: int, p_off: int, dtype: torch.dtype, m, v, master, lo: int, hi: int,
                           hyper, step_dev, stage_off: int, stage_bytes: int, blocks: int = 0, channel: int = 2):
        """One-shot flavour (push to every peer, one barrier, every rank applies the whole rule)."""
        lib = _native.cuda()
        code = _dtype_code(dtype)
        arr = lambda ts: self._ptrs(ts) if ts is not None else None
        cuda_check(lib.fmpi_fused_step_oneshot(C.byref(self.comms[channel]), g_off, p_off, code, arr(m), arr(v),
                                               arr(master), lo, hi, C.byref(hyper), arr(step_dev), stage_off,
                                               stage_bytes, self._launch(blocks), _stream_ptr()),
                   "emulated fused_step_oneshot")
        _native.count_launch()

    def tree_step_oneshot(self, tabs, n: int, total: int, dtype: torch.dtype, hyper, step: int,
                          stage_off: int, stage_bytes: int, blocks: int = 0, channel: int = 0):
        """``tabs``: one device table per virtual rank ([p | g | m | v | numel | off] int64)."""
        lib = _native.cuda()
        cuda_check(lib.fmpi_tree_step_oneshot(C.byref(self.comms[channel]), n, self._ptrs(tabs), total,
                                              _dtype_code(dtype), C.byref(hyper), step, stage_off, stage_bytes,
                                              self._launch(blocks), _stream_ptr()), "emulated tree_step_oneshot")
        _native.count_launch()

    def tree_allreduce(self, tabs, n: int, off: int, total: int, dtype: torch.dtype, op: int, blocks: int = 0,
                       channel: int = 0):
        lib = _native.cuda()
        cuda_check(lib.fmpi_tree_allreduce(C.byref(self.comms[channel]), n, self._ptrs(tabs), off, total,
                                           _dtype_code(dtype), op, self._launch(blocks, 0, _native.ALGO_TWOSHOT),
                                           _stream_ptr()), "emulated tree_allreduce")
        _native.count_launch()
