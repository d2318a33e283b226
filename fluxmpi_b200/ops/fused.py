"""Python faces of the non-collective kernels in ``libfmpi_cuda.so``.

* :func:`multi_tensor_step` — one launch that applies an optimiser rule to a whole list of
  scattered (param, grad, m, v) leaves (csrc/kernels/multi_tensor.cu). The reference leaves
  this to Optimisers.jl, which issues 2-4 broadcast kernels per leaf.
* :func:`sumsq` / :func:`checksum` — gradient norm and bitwise replica checks
  (csrc/kernels/reduce_diag.cu).
* :func:`flush_l2` — bench hygiene.

Everything raises if the CUDA library is absent; there is no eager fallback on a GPU box.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence

import torch

from .. import _native
from .. import rules as R
from .._native import Hyper, cuda_check


def _stream_ptr(stream=None) -> int:
    return (torch.cuda.current_stream() if stream is None else stream).cuda_stream


def _dtype_code(dt):
    from ..comm import dtype_code
    return dtype_code(dt)


def unwrap_rule(rule):
    """Strip a DistributedOptimizer wrapper; returns (inner_rule, average)."""
    from ..parallel.optimizer import DistributedOptimizer
    if isinstance(rule, DistributedOptimizer):
        return rule.optimizer, rule.average
    return rule, False


def hyper_for(rule, grad_scale: float = 1.0, zero_grad: bool = False, eta_dev: int = 0) -> Optional[Hyper]:
    """Map an Optimisers-style rule onto the fused kernels' hyper-parameter block, or None when
    the rule has no fused implementation (the caller then reduces with a kernel and applies
    the rule with torch ops)."""
    rule, _ = unwrap_rule(rule)
    h = Hyper()
    h.grad_scale, h.zero_grad, h.eta_dev = grad_scale, int(zero_grad), eta_dev
    h.beta1, h.beta2, h.eps, h.lam, h.rho = 0.9, 0.999, 1e-8, 0.0, 0.0
    if type(rule) is R.Descent:
        h.rule, h.eta = _native.RULE_DESCENT, rule.eta
    elif type(rule) is R.Momentum:
        h.rule, h.eta, h.rho = _native.RULE_MOMENTUM, rule.eta, rule.rho
    elif type(rule) is R.Nesterov:
        h.rule, h.eta, h.rho = _native.RULE_NESTEROV, rule.eta, rule.rho
    elif type(rule) is R.Adam:
        h.rule, h.eta, h.eps = _native.RULE_ADAM, rule.eta, rule.eps
        h.beta1, h.beta2 = rule.beta
    elif (type(rule) is R.OptimiserChain and len(rule.opts) == 2 and type(rule.opts[0]) is R.Adam
          and type(rule.opts[1]) is R.WeightDecay):
        a, wd = rule.opts
        h.rule, h.eta, h.eps, h.lam = _native.RULE_ADAMW, a.eta, a.eps, wd.gamma
        h.beta1, h.beta2 = a.beta
    else:
        return None
    return h


def rule_needs(h: Hyper) -> tuple[bool, bool]:
    """(needs first moment / velocity, needs second moment)"""
    return h.rule != _native.RULE_DESCENT, h.rule in (_native.RULE_ADAM, _native.RULE_ADAMW)


class TensorTable:
    """Device-resident pointer table for a fixed list of leaves (built once, reused per step)."""

    def __init__(self, params: Sequence[torch.Tensor], grads: Sequence[torch.Tensor],
                 m: Optional[Sequence[torch.Tensor]], v: Optional[Sequence[torch.Tensor]]):
        n = len(params)
        dev = params[0].device
        zeros = [0] * n
        rows = [[t.data_ptr() for t in params], [t.data_ptr() for t in grads],
                [t.data_ptr() for t in m] if m is not None else zeros,
                [t.data_ptr() for t in v] if v is not None else zeros,
                [t.numel() for t in params]]
        flat = [x for row in rows for x in row]
        self.n = n
        self.has_m, self.has_v = m is not None, v is not None
        self.dev = torch.tensor(flat, dtype=torch.int64).pin_memory().to(dev, non_blocking=True)
        self.keep = (list(params), list(grads), list(m or []), list(v or []))
        self.key = tuple(rows[0]) + tuple(rows[1])

    def row(self, i: int) -> int:
        return self.dev.data_ptr() + 8 * self.n * i


def multi_tensor_step(table: TensorTable, h: Hyper, step: int, dtype: torch.dtype, stream=None):
    lib = _native.cuda()
    code = _dtype_code(dtype)
    cuda_check(lib.fmpi_multi_tensor_step(
        table.n, table.row(0), table.row(1), table.row(2) if table.has_m else None,
        table.row(3) if table.has_v else None, table.row(4), code, code, C.byref(h), step,
        _stream_ptr(stream)), "multi_tensor_step")
    _native.count_launch()


_scratch: dict = {}


def _scratch_for(device) -> torch.Tensor:
    key = (device.type, device.index, torch.cuda.current_stream(device).cuda_stream)
    s = _scratch.get(key)
    if s is None:
        s = torch.zeros(_native.cuda().fmpi_reduce_scratch_bytes(), dtype=torch.uint8, device=device)
        _scratch[key] = s
    return s


def sumsq(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Sum of squares of ``x`` in float64 (a 0-d device tensor; no host sync)."""
    lib = _native.cuda()
    x = x.detach().contiguous()
    out = torch.empty((), dtype=torch.float64, device=x.device) if out is None else out
    cuda_check(lib.fmpi_sumsq(x.data_ptr(), x.numel(), _dtype_code(x.dtype), out.data_ptr(),
                              _scratch_for(x.device).data_ptr(), _stream_ptr()), "sumsq")
    _native.count_launch()
    return out


def checksum(x: torch.Tensor) -> torch.Tensor:
    """64-bit position-weighted checksum of the bytes of ``x`` (0-d int64 device tensor)."""
    lib = _native.cuda()
    x = x.detach().contiguous()
    out = torch.empty((), dtype=torch.int64, device=x.device)
    cuda_check(lib.fmpi_checksum(x.data_ptr(), x.numel() * x.element_size(), out.data_ptr(),
                                 _scratch_for(x.device).data_ptr(), _stream_ptr()), "checksum")
    _native.count_launch()
    return out


_flush_buf: dict = {}


def flush_l2(device=None, nbytes: int = 256 << 20):
    """Overwrite a buffer larger than the 126 MB L2 so the next timed iteration starts cold."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    buf = _flush_buf.get(device)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
        _flush_buf[device] = buf
    cuda_check(_native.cuda().fmpi_flush_l2(buf.data_ptr(), nbytes, _stream_ptr()), "flush_l2")
