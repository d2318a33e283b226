"""FusedBatchNormAct — channels-last BatchNorm2d fused with the residual add and the ReLU that
follow it, forward and backward, on the hand-written kernels of csrc/kernels/batchnorm.cu.

``FusedBatchNormAct`` subclasses ``nn.BatchNorm2d`` (same parameters, buffers and state-dict
keys), so a model can switch between the stock layer and this one without touching its
checkpoints:

    y = bn(x)                      # relu=False:  batchnorm(x)
    y = bn(x)                      # relu=True:   relu(batchnorm(x))
    y = bn(x, residual=identity)   #              relu?(batchnorm(x) + identity)

The kernels need NHWC-dense (``channels_last``) CUDA tensors of f32 / bf16 / f16 with a channel
count that is a multiple of 4 (f32) or 8 (16-bit); anything else — CPU tensors, NCHW layout,
odd channel counts — takes the stock PyTorch ops, so the module is safe everywhere.

Why it exists: PyTorch's native channels_last batch-norm kernels take 60.7 % of a ResNet-50
training step on B200 and the separate ReLU / add kernels 15 % (profiles/README.md §7); the
reference ships no model code at all.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F
from torch import nn

from .. import _native

_DT = {torch.float32: _native.F32, torch.bfloat16: _native.BF16, torch.float16: _native.F16}
_scratch: dict = {}


def _scratch_for(device: torch.device, C: int) -> torch.Tensor:
    key = (device.index, C)
    s = _scratch.get(key)
    if s is None:
        s = torch.empty(_native.cuda().fmpi_bn_scratch_bytes(C), dtype=torch.uint8, device=device)
        _scratch[key] = s
    return s


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def kernel_ok(x: torch.Tensor, residual: Optional[torch.Tensor] = None) -> bool:
    if not (x.is_cuda and x.dim() == 4 and x.dtype in _DT and x.numel() > 0):
        return False
    C = x.shape[1]
    if C % (4 if x.dtype == torch.float32 else 8):
        return False
    if not x.is_contiguous(memory_format=torch.channels_last):
        return False
    if residual is not None and not (residual.shape == x.shape and residual.dtype == x.dtype
                                     and residual.is_contiguous(memory_format=torch.channels_last)):
        return False
    return True


class _BNActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, weight, bias, running_mean, running_var, momentum, eps, relu):
        lib = _native.cuda()
        N, C, H, W = x.shape
        M = N * H * W
        y = torch.empty_like(x)                                   # preserves channels_last
        mean = torch.empty(C, dtype=torch.float32, device=x.device)
        invstd = torch.empty(C, dtype=torch.float32, device=x.device)
        w = weight.detach().float() if weight is not None else None
        b = bias.detach().float() if bias is not None else None
        _native.cuda_check(lib.fmpi_bn_forward(
            x.data_ptr(), residual.data_ptr() if residual is not None else None, y.data_ptr(), M, C,
            _DT[x.dtype], w.data_ptr() if w is not None else None, b.data_ptr() if b is not None else None,
            running_mean.data_ptr() if running_mean is not None else None,
            running_var.data_ptr() if running_var is not None else None,
            float(momentum), float(eps), int(relu), mean.data_ptr(), invstd.data_ptr(),
            _scratch_for(x.device, C).data_ptr(), _stream()), "bn_forward")
        _native.count_launch(3)
        ctx.relu = bool(relu)
        ctx.has_res = residual is not None
        ctx.affine = (weight is not None, bias is not None)
        ctx.wdtype = weight.dtype if weight is not None else None
        # the ReLU mask is recomputed from x in backward unless a residual went into the ReLU
        ctx.save_for_backward(x, y if (relu and residual is not None) else None, w, b, mean, invstd)
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _native.cuda()
        x, y, w, b, mean, invstd = ctx.saved_tensors
        N, C, H, W = x.shape
        M = N * H * W
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty_like(x)
        dres = torch.empty_like(x) if (ctx.has_res and ctx.relu) else None
        dgamma = torch.empty(C, dtype=torch.float32, device=x.device)
        dbeta = torch.empty(C, dtype=torch.float32, device=x.device)
        _native.cuda_check(lib.fmpi_bn_backward(
            dy.data_ptr(), y.data_ptr() if y is not None else None, x.data_ptr(), dx.data_ptr(),
            dres.data_ptr() if dres is not None else None, M, C, _DT[x.dtype],
            w.data_ptr() if w is not None else None, b.data_ptr() if b is not None else None,
            mean.data_ptr(), invstd.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), int(ctx.relu), _scratch_for(x.device, C).data_ptr(),
            _stream()), "bn_backward")
        _native.count_launch(3)
        if ctx.has_res and not ctx.relu:
            dres = dy                                   # d(x + r)/dr = 1: the same gradient, no copy
        has_w, has_b = ctx.affine
        return (dx, dres if ctx.has_res else None,
                dgamma.to(ctx.wdtype) if has_w else None, dbeta.to(ctx.wdtype) if has_b else None,
                None, None, None, None, None)


def fused_bn_act(x, weight, bias, running_mean, running_var, training: bool, momentum: float, eps: float,
                 relu: bool = False, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Functional form. Falls back to the stock ops whenever the kernels do not apply."""
    if not kernel_ok(x, residual):
        y = F.batch_norm(x, running_mean, running_var, weight, bias, training, momentum, eps)
        if residual is not None:
            y = y + residual
        return F.relu(y) if relu else y
    if training or running_mean is None:
        return _BNActFn.apply(x, residual, weight, bias, running_mean, running_var, momentum, eps, relu)
    # inference: precomputed affine, one elementwise kernel
    lib = _native.cuda()
    N, C, H, W = x.shape
    invstd = torch.rsqrt(running_var.float() + eps)
    scale = invstd * (weight.detach().float() if weight is not None else 1.0)
    shift = (bias.detach().float() if bias is not None else 0.0) - running_mean.float() * scale
    y = torch.empty_like(x)
    _native.cuda_check(lib.fmpi_bn_apply(x.data_ptr(), residual.data_ptr() if residual is not None else None,
                                         y.data_ptr(), N * H * W, C, _DT[x.dtype], scale.contiguous().data_ptr(),
                                         shift.contiguous().data_ptr(), int(relu), _stream()), "bn_apply")
    _native.count_launch()
    return y


class FusedBatchNormAct(nn.BatchNorm2d):
    """``nn.BatchNorm2d`` + optional residual add + optional ReLU in one pass over memory."""

    def __init__(self, num_features: int, relu: bool = False, eps: float = 1e-5, momentum: float = 0.1,
                 affine: bool = True, track_running_stats: bool = True):
        super().__init__(num_features, eps=eps, momentum=momentum, affine=affine,
                         track_running_stats=track_running_stats)
        self.relu = relu

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        training = self.training or not self.track_running_stats
        momentum = 0.0 if self.momentum is None else self.momentum
        if self.training and self.track_running_stats and self.num_batches_tracked is not None:
            self.num_batches_tracked.add_(1)
            if self.momentum is None:
                momentum = 1.0 / float(self.num_batches_tracked)
        return fused_bn_act(x, self.weight, self.bias,
                            self.running_mean if self.track_running_stats else None,
                            self.running_var if self.track_running_stats else None,
                            training, momentum, self.eps, self.relu, residual)

    def extra_repr(self) -> str:
        return super().extra_repr() + f", relu={self.relu}"
