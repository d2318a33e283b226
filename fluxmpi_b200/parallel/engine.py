"""DataParallelEngine — the B200 training path.

What the reference does per optimiser step (/root/reference/src/optimizer.jl:20-23 inside
``Optimisers.update``, README.md:63-66): after the *whole* backward pass, for every parameter
leaf, a blocking allreduce (host-staged for GPU arrays) and then the rule's elementwise
kernels — strictly serial, one message per leaf.

What happens here:

* every parameter of the module becomes a view into ONE flat buffer ``P`` and every ``.grad`` a
  view into ONE flat buffer ``G``; both live in a symmetric segment that all peers map
  (the ComponentArray-shaped layout of SURVEY §2.1, :class:`FlatParams` gives named access);
* the flat range is cut into buckets in backward order; a post-accumulate-grad hook launches a
  bucket's **fused all-reduce + optimiser kernel** (csrc/kernels/fused_step.cu) on a
  high-priority communication stream as soon as its gradients exist, so NVLink traffic
  overlaps the rest of backward;
* the optimiser moments are written only by the owner of a slice (the kernel pushes the new
  parameters to every replica), so optimiser HBM traffic is ~1/W of the replicated scheme;
  :meth:`gather_state` materialises full replicated moments on demand (checkpointing,
  ``synchronize``), which keeps the *observable* state tree identical to the unwrapped rule's
  (test/test_optimizer.jl:13-14);
* semantics stay the reference's: gradients are SUMMED (``average=True`` opts in to the mean).

Rules without a fused kernel still work: the bucket is all-reduced in place by the two-shot
kernel and the rule is applied with torch ops.

Without peer mappings (``disable_p2p_support()``, ``FMPI_FORCE_FALLBACK=1``, a job over several
nodes) the layout and the bucket schedule stay the same; a bucket is all-reduced by NCCL from
the communication stream and the rule is applied by the local fused kernel, with the optimiser
state replicated (tests/spmd/gpu_engine_fallback.py).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Callable, Optional

import torch

from .. import _native, runtime
from .. import rules as R
from .._native import Hyper, Launch, cuda_check
from ..ops import fused as F
from .flat import Axes, FlatParams

_ELEM_ALIGN = 64          # every parameter starts on a 64-element boundary (>= 128 B)


@dataclass
class _Bucket:
    lo: int
    hi: int
    params: list
    pending: int = 0
    ready: bool = False
    launched: bool = False


class DataParallelEngine:
    def __init__(self, module: torch.nn.Module, rule: R.AbstractRule, *, bucket_mb: float = 32.0,
                 overlap: bool = True, average: Optional[bool] = None, master_weights: bool = True,
                 sync_params: bool = True, root_rank: int = 0, algo: str = "auto",
                 blocks: int = 0, channel: int = 2):
        if not runtime.Initialized():
            raise runtime.FluxMPINotInitializedError()
        S = runtime.state()
        if S.device is None or S.device.type != "cuda":
            raise RuntimeError("DataParallelEngine needs a CUDA device (use rules.update on CPU)")
        self.dw = S.device_world
        self.world, self.rank = S.world, S.rank
        self.device = S.device
        self.module = module
        inner, avg = F.unwrap_rule(rule)
        self.rule = inner
        self.average = avg if average is None else average
        self.overlap = overlap and self.world > 1
        self.channel = channel
        self.blocks = blocks
        self.steps_done = 0
        self.native = self.dw is not None and self.dw.native

        params = [p for p in module.parameters() if p.requires_grad]
        seen, uniq = set(), []
        for p in params:
            if id(p) not in seen:
                seen.add(id(p))
                uniq.append(p)
        if not uniq:
            raise ValueError("module has no trainable parameters")
        dts = {p.dtype for p in uniq}
        if len(dts) != 1:
            raise ValueError(f"DataParallelEngine needs one parameter dtype, got {dts}")
        self.dtype = uniq[0].dtype
        if self.dtype not in (torch.float32, torch.bfloat16, torch.float16):
            raise ValueError(f"unsupported parameter dtype {self.dtype}")
        self.params = uniq
        es = uniq[0].element_size()

        # ---- flat layout
        offs, cur = [], 0
        for p in uniq:
            offs.append(cur)
            cur += (p.numel() + _ELEM_ALIGN - 1) // _ELEM_ALIGN * _ELEM_ALIGN
        self.numel = cur
        self.offsets = offs
        buf_bytes = (cur * es + (1 << 21) - 1) >> 21 << 21

        # ---- symmetric segment  [P | G] (+ [2 parities][W] push slots for the one-shot flavour)
        self.p_off, self.g_off = 0, buf_bytes
        # a model this small is latency-bound: every rank pushes its whole gradient to the peers,
        # one barrier, everybody applies the whole rule (csrc/kernels/fused_step.cu, one-shot)
        oneshot_max = int(float(os.environ.get("FMPI_FUSED_ONESHOT_KB", 2048)) * 1024)
        self.oneshot = bool(self.native and self.world > 1 and F.hyper_for(self.rule) is not None
                            and (algo == "oneshot" or (algo == "auto" and cur * es <= oneshot_max)))
        self.stage_off = 2 * buf_bytes
        self.stage_bytes = 2 * self.world * ((cur * es + 255) // 256 * 256) if self.oneshot else 0
        if self.native:
            self.seg = self.dw.alloc_segment(2 * buf_bytes + self.stage_bytes)
            self.comm = self.dw.comm_for(self.seg, channel)
            rawP = self.dw.symm_tensor(self.seg, self.p_off, cur * es)
            rawG = self.dw.symm_tensor(self.seg, self.g_off, cur * es)
        else:
            # No peer mappings (P2P switched off, or the job spans nodes): the same flat layout
            # in ordinary memory. A bucket is all-reduced by NCCL on the communication stream
            # and the rule is then applied by the local fused kernel, still overlapping the rest
            # of backward; the optimiser state is replicated, as in the reference.
            if self.dw is None:
                raise RuntimeError("DataParallelEngine needs Init() on a CUDA machine")
            self.seg = None
            self._flat = torch.empty(2 * buf_bytes, dtype=torch.uint8, device=self.device)
            self.comm = _native.DevComm()           # a world of one over the flat buffer
            self.comm.rank, self.comm.world, self.comm.emulate, self.comm.channel = 0, 1, 0, channel
            self.comm.data[0] = self._flat.data_ptr()
            self.comm.data_bytes = 2 * buf_bytes
            rawP = self._flat[:cur * es]
            rawG = self._flat[buf_bytes:buf_bytes + cur * es]
        self.P = rawP.view(self.dtype)
        self.G = rawG.view(self.dtype)
        self.G.zero_()
        with torch.no_grad():
            for p, o in zip(uniq, offs):
                # keep the parameter's memory format (channels_last conv weights stay NHWC, so
                # cuDNN never re-lays them out); autograd wants grads with the same strides
                dense = p.is_contiguous() or p.is_contiguous(memory_format=torch.channels_last) \
                    if p.dim() == 4 else p.is_contiguous()
                stride = p.stride() if dense else None
                mk = (lambda buf: buf[o:o + p.numel()].as_strided(p.shape, stride)) if stride \
                    else (lambda buf: buf[o:o + p.numel()].view(p.shape))
                view = mk(self.P)
                view.copy_(p.data)
                p.data = view
                p.grad = mk(self.G)

        # ---- optimiser state (process-local; only the owned slices are ever written)
        self.hyper: Optional[Hyper] = F.hyper_for(self.rule)
        self.fused = self.hyper is not None
        self.eta_dev = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=self.device)
        self.m = self.v = self.master = None
        if self.fused:
            need_m, need_v = F.rule_needs(self.hyper)
            if need_m:
                self.m = torch.zeros(cur, dtype=torch.float32, device=self.device)
            if need_v:
                self.v = torch.zeros(cur, dtype=torch.float32, device=self.device)
            if self.dtype != torch.float32 and master_weights:
                self.master = self.P.float()
            self.eta_dev.fill_(self.hyper.eta)
            self.hyper.eta_dev = self.eta_dev.data_ptr()
            self.hyper.grad_scale = 1.0 / self.world if self.average else 1.0
            self.hyper.zero_grad = 1
        else:
            self.state_tree = None   # built lazily by the unfused path

        # in-switch reduction moves ~N bytes per link instead of 2N(W-1)/W: a win from W = 4 up
        has_mc = self.native and bool(self.seg.mc_ptr)
        use_nvls = (algo == "nvls") or (algo == "auto" and has_mc and self.world >= 4)
        if algo == "nvls" and not has_mc:
            raise RuntimeError("algo='nvls' requested but this world has no multicast mapping")
        # measured choice for in-place all-reduce of a bucket-sized symmetric region, if any
        tuned = self.dw.tuner.choice("allreduce_symm", int(min(bucket_mb * (1 << 20), cur * es))) \
            if (self.native and self.world > 1 and self.dw.tuner is not None) else None
        if tuned is not None:
            if algo == "auto":
                use_nvls = tuned[0] == "nvls" and has_mc
        # CTA count: unlike a pure all-reduce (8-16 CTAs saturate the in-switch path) the fused
        # kernel also streams m, v and p through HBM; measured at 8 GPUs (bench/fused_bench.py)
        # 48 CTAs are 1.5-1.9x faster than 16 for both variants, so default well above that
        if blocks == 0 and self.world > 1 and self.native:
            self.blocks = int(os.environ.get("FMPI_FUSED_BLOCKS", 0))     # 0: the library's default
        self.algo = _native.ALGO_NVLS if use_nvls else _native.ALGO_TWOSHOT
        if self.oneshot:
            self.algo = _native.ALGO_ONESHOT
            bucket_mb = float("inf")            # one launch for the whole model
            self.blocks = blocks

        # ---- buckets in backward (reverse registration) order
        bucket_elems = cur if bucket_mb == float("inf") else max(int(bucket_mb * (1 << 20) / es), 1)
        if self.world == 1:
            bucket_elems = cur          # nothing to overlap with: one launch for the whole model
        self.buckets: list[_Bucket] = []
        hi = cur
        acc: list = []
        for p, o in reversed(list(zip(uniq, offs))):
            acc.append(p)
            if hi - o >= bucket_elems or o == 0:
                self.buckets.append(_Bucket(o, hi, acc))
                hi, acc = o, []
        self._bucket_of = {}
        for i, b in enumerate(self.buckets):
            for p in b.params:
                self._bucket_of[id(p)] = i
        self._next = 0
        self._hooks = []
        if self.overlap:
            for p in uniq:
                self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))
        self.comm_stream = self.dw.comm_stream
        self._armed = False
        self._reset_buckets()

        if sync_params and self.world > 1:
            self.broadcast_params(root_rank)

    # ------------------------------------------------------------------ views
    def flat_params(self) -> FlatParams:
        """Named views over the flat parameter vector (ComponentArray-style)."""
        ax = Axes()
        names = {id(p): n for n, p in self.module.named_parameters()}
        for p, o in zip(self.params, self.offsets):
            ax.entries[names.get(id(p), f"param{o}")] = (o, tuple(p.shape))
        return FlatParams(self.P, ax)

    def flat_grads(self) -> FlatParams:
        return self.flat_params().like(self.G)

    # ------------------------------------------------------------------ collectives on the segment
    def broadcast_params(self, root: int = 0):
        """``synchronize!`` of the parameters in ONE kernel (ext/FluxMPIComponentArraysExt.jl:6-9
        does the same with one MPI_Bcast of the flat vector) + buffers through the tree path."""
        lib = _native.cuda()
        nbytes = self.numel * self.P.element_size()
        if self.native:
            cuda_check(lib.fmpi_bcast_symm(C.byref(self.comm), self.p_off, nbytes, root,
                                           C.byref(Launch(0, 0, 0)), torch.cuda.current_stream().cuda_stream),
                       "bcast_symm")
            _native.count_launch()
        elif self.world > 1:
            self.dw._nccl().broadcast(self.P, root)
        if self.master is not None:
            self.master.copy_(self.P)
        bufs = [b for b in self.module.buffers() if b.numel() > 0]
        if bufs:
            from .sync import synchronize
            synchronize(bufs, root_rank=root)

    # ------------------------------------------------------------------ bucket machinery
    def _reset_buckets(self):
        for b in self.buckets:
            b.pending = len(b.params)
            b.ready = b.launched = False
        self._next = 0
        self._armed = False

    def accumulate(self):
        """Context manager for gradient accumulation: backward passes inside it only add into
        the flat gradient buffer (no bucket is launched); the next backward outside it, or
        :meth:`step`, reduces and applies the accumulated sum."""
        eng = self

        class _Ctx:
            def __enter__(self):
                eng._accumulating = True

            def __exit__(self, *a):
                eng._accumulating = False
                eng._reset_buckets()

        return _Ctx()

    def _on_grad(self, p):
        if getattr(self, "_accumulating", False):
            return
        i = self._bucket_of.get(id(p))
        if i is None:
            return
        b = self.buckets[i]
        b.pending -= 1
        if b.pending == 0:
            b.ready = True
            self._launch_ready()

    def _launch_ready(self):
        while self._next < len(self.buckets) and self.buckets[self._next].ready:
            self._launch_bucket(self._next, self.comm_stream)
            self._next += 1

    def _launch_bucket(self, i: int, stream: torch.cuda.Stream):
        b = self.buckets[i]
        if b.launched:
            return
        cur = torch.cuda.current_stream()
        if stream is not cur and stream.cuda_stream != cur.cuda_stream:
            stream.wait_stream(cur)      # the bucket's gradients are complete on `cur`
            self._armed = True
        lib = _native.cuda()
        if not self.native and self.world > 1:
            # NCCL runs on its own stream; issuing from `stream` makes it start after the
            # bucket's gradients and makes `stream` (not the host) wait for the result
            with torch.cuda.stream(stream):
                self.dw._nccl().all_reduce(self.G[b.lo:b.hi])
        if self.fused:
            one = lambda t: C.byref(C.c_void_p(t.data_ptr())) if t is not None else None
            step_ptr = C.byref(C.c_void_p(self.step_dev.data_ptr()))
            code = F._dtype_code(self.dtype)
            if self.oneshot:
                cuda_check(lib.fmpi_fused_step_oneshot(
                    C.byref(self.comm), self.g_off, self.p_off, code, one(self.m), one(self.v),
                    one(self.master), b.lo, b.hi, C.byref(self.hyper), step_ptr, self.stage_off,
                    self.stage_bytes, C.byref(Launch(self.blocks, 0, self.algo)), stream.cuda_stream),
                    "fused_step_oneshot")
            else:
                cuda_check(lib.fmpi_fused_step(
                    C.byref(self.comm), self.g_off, code, self.p_off, code, one(self.m), one(self.v),
                    one(self.master), b.lo, b.hi, C.byref(self.hyper), step_ptr,
                    C.byref(Launch(self.blocks, 0, self.algo)), stream.cuda_stream), "fused_step")
        else:
            code = F._dtype_code(self.dtype)
            es = self.G.element_size()
            if self.world > 1 and self.native:
                cuda_check(lib.fmpi_allreduce_symm(
                    C.byref(self.comm), self.g_off + b.lo * es, b.hi - b.lo, code, _native.SUM,
                    C.byref(Launch(self.blocks, 0, 0)), stream.cuda_stream), "allreduce_symm")
        _native.count_launch()
        b.launched = True

    # ------------------------------------------------------------------ the step
    def step(self):
        """Finish the optimiser step: launch whatever buckets backward did not (no-overlap
        mode, unused parameters), join the communication stream, advance the step counter."""
        cur = torch.cuda.current_stream()
        use_side = self.overlap
        for i in range(self._next, len(self.buckets)):
            self._launch_bucket(i, self.comm_stream if use_side else cur)
        self._next = len(self.buckets)
        if use_side or self._armed:
            cur.wait_stream(self.comm_stream)
        if self.fused:
            cuda_check(_native.cuda().fmpi_step_advance(self.step_dev.data_ptr(), cur.cuda_stream), "step_advance")
            _native.count_launch()
        else:
            self._unfused_apply()
        self.steps_done += 1
        self._reset_buckets()
        if self.native and self.world > 1:
            self.dw.check_health()       # two host-memory reads: a lost peer surfaces here

    def materialize_state(self):
        """Build the lazily created per-leaf state of an unfused rule (no-op otherwise), so that
        every rank walks a state tree of the same shape (checkpointing, synchronize)."""
        if not self.fused and self.state_tree is None:
            self.state_tree = [R.Leaf(self.rule, self.rule.init(p.data)) for p in self.params]

    def _unfused_apply(self):
        """Arbitrary rule: gradients were all-reduced in place by the kernel; apply with torch."""
        self.materialize_state()
        scale = 1.0 / self.world if self.average else 1.0
        with torch.no_grad():
            for leaf, p in zip(self.state_tree, self.params):
                g = p.grad if scale == 1.0 else p.grad * scale
                leaf.state, dx = self.rule.apply(leaf.state, p.data, g)
                p.data.sub_(dx.to(p.dtype))
            self.G.zero_()

    def zero_grad(self):
        self.G.zero_()

    def set_lr(self, eta: float):
        """Learning-rate change that CUDA-graph replays pick up (device scalar)."""
        if self.fused:
            self.eta_dev.fill_(eta)
            self.hyper.eta = eta
        else:
            self.rule = R._adjust_rule(self.rule, eta, {})

    # ------------------------------------------------------------------ state access
    def owned_ranges(self) -> list[tuple[int, int]]:
        """Element ranges whose optimiser state this rank owns (one slice per bucket)."""
        if not self.native or self.oneshot:  # replicated state: every rank holds all of it
            return [(b.lo, b.hi) for b in self.buckets]
        es = self.P.element_size()
        n_per_vec = 16 // es
        out = []
        for b in self.buckets:
            nvec = (b.hi - b.lo) // n_per_vec
            per = (nvec + self.world - 1) // self.world
            lo = min(per * self.rank, nvec)
            hi = min(lo + per, nvec)
            out.append((b.lo + lo * n_per_vec, b.lo + hi * n_per_vec))
        return out

    def gather_state(self) -> dict:
        """Full, replicated optimiser state (what the unwrapped rule would hold on every rank).
        Owned slices are summed into zeros elsewhere — one all-reduce per moment."""
        out = {"step": int(self.step_dev.item()) if self.fused else self.steps_done}
        if not self.fused:
            self.materialize_state()
            out["leaves"] = [leaf.state for leaf in self.state_tree]   # plain tensors / numbers / tuples
            return out
        for name in ("m", "v", "master"):
            t = getattr(self, name)
            if t is None:
                continue
            if not self.native or self.oneshot:
                out[name] = t.clone()
                continue
            full = torch.zeros_like(t)
            for lo, hi in self.owned_ranges():
                full[lo:hi] = t[lo:hi]
            if self.world > 1:
                self.dw.allreduce(full, _native.SUM)
            out[name] = full
        return out

    def load_state(self, state: dict):
        if not self.fused:
            leaves = state.get("leaves")
            self.state_tree = None if leaves is None else \
                [R.Leaf(self.rule, st.state if isinstance(st, R.Leaf) else st) for st in leaves]
            self.steps_done = state.get("step", 0)
            return
        self.step_dev.fill_(int(state.get("step", 0)))
        for name in ("m", "v", "master"):
            if name in state and getattr(self, name) is not None:
                getattr(self, name).copy_(state[name])

    def state_tree_like_rule(self):
        """Optimisers-shaped view of the state: ``{param_name: Leaf(rule, state)}`` with the
        same layout ``rules.setup(rule, params)`` would produce."""
        st = self.gather_state()
        names = {id(p): n for n, p in self.module.named_parameters()}
        tree = {}
        t = st["step"]
        for p, o in zip(self.params, self.offsets):
            sl = slice(o, o + p.numel())
            if isinstance(self.rule, R.Adam) or (isinstance(self.rule, R.OptimiserChain)):
                a = self.rule if isinstance(self.rule, R.Adam) else self.rule.opts[0]
                state = (st["m"][sl].view(p.shape), st["v"][sl].view(p.shape),
                         (a.beta[0] ** (t + 1), a.beta[1] ** (t + 1)))
                if isinstance(self.rule, R.OptimiserChain):
                    state = (state, None)
            elif "m" in st:
                state = st["m"][sl].view(p.shape)
            else:
                state = None
            tree[names.get(id(p), str(o))] = R.Leaf(self.rule, state)
        return tree

    def close(self):
        for h in self._hooks:
            h.remove()
        self._hooks.clear()


class GraphedStep:
    """Capture ``forward + backward + fused buckets + step`` into ONE CUDA graph.

    ResNet-50 on a B200 is launch-bound at moderate batch sizes (hundreds of short cuDNN
    kernels); replaying a graph removes the CPU from the loop. The bucket kernels are ordinary
    launches on the communication stream, which joins the capture through stream events, and
    all per-step scalars they need (step count, learning rate) live in device memory.
    """

    def __init__(self, engine: DataParallelEngine, fn: Callable, example_inputs: tuple,
                 warmup: int = 3):
        self.engine = engine
        self.static_in = tuple(torch.empty_like(x) for x in example_inputs)
        for s, x in zip(self.static_in, example_inputs):
            s.copy_(x)
        # warm-up and capture run on side streams on purpose; autograd's "AccumulateGrad stream
        # mismatch" notice is about exactly that and is not actionable here
        quiet = getattr(torch.autograd.graph, "set_warn_on_accumulate_grad_stream_mismatch", None)
        if quiet is not None:
            quiet(False)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                loss = fn(*self.static_in)
                engine.step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        if engine.world > 1:
            runtime.host_barrier()
        self.graph = torch.cuda.CUDAGraph()
        before = _native.launch_count
        with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):
            self.static_loss = fn(*self.static_in)
            engine.step()
        self.launches_per_replay = _native.launch_count - before   # our kernels inside the graph
        _native.launch_count = before                              # captured, not executed
        engine.steps_done -= 1
        self.warmup_steps = warmup

    def __call__(self, *inputs):
        for s, x in zip(self.static_in, inputs):
            if s.data_ptr() != x.data_ptr():
                s.copy_(x, non_blocking=True)
        self.graph.replay()
        self.engine.steps_done += 1
        _native.count_launch(self.launches_per_replay)
        return self.static_loss
