"""FusedMaxPool2d — channels-last max pooling on the kernels of csrc/kernels/pool.cu
(vector-wide forward that records the arg-max as one byte per element, atomics-free gather
backward). Drop-in for ``nn.MaxPool2d`` (floor mode, no dilation); anything the kernels do not
cover — CPU tensors, NCHW layout, odd channel counts, ``ceil_mode`` — uses the stock op.

PyTorch's own NHWC max-pool kernels cost 2.5 ms of the ResNet-50 step on B200 for ~0.3 ms worth
of HBM traffic (profiles/README.md §7).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F
from torch import nn

from .. import _native

_DT = {torch.float32: _native.F32, torch.bfloat16: _native.BF16, torch.float16: _native.F16}


def kernel_ok(x: torch.Tensor) -> bool:
    return (x.is_cuda and x.dim() == 4 and x.dtype in _DT and x.numel() > 0
            and x.shape[1] % (4 if x.dtype == torch.float32 else 8) == 0
            and x.is_contiguous(memory_format=torch.channels_last))


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, s, p):
        lib = _native.cuda()
        N, C, H, W = x.shape
        OH, OW = (H + 2 * p - k) // s + 1, (W + 2 * p - k) // s + 1
        y = torch.empty((N, C, OH, OW), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        idx = torch.empty((N, OH, OW, C), dtype=torch.uint8, device=x.device)
        _native.cuda_check(lib.fmpi_maxpool_fwd(x.data_ptr(), y.data_ptr(), idx.data_ptr(), N, H, W, C, k, s, p,
                                                _DT[x.dtype], torch.cuda.current_stream().cuda_stream), "maxpool_fwd")
        _native.count_launch()
        ctx.save_for_backward(idx)
        ctx.geom = (N, C, H, W, k, s, p)
        ctx.dtype = x.dtype
        return y

    @staticmethod
    def backward(ctx, dy):
        lib = _native.cuda()
        (idx,) = ctx.saved_tensors
        N, C, H, W, k, s, p = ctx.geom
        if not dy.is_contiguous(memory_format=torch.channels_last):
            dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty((N, C, H, W), dtype=ctx.dtype, device=dy.device, memory_format=torch.channels_last)
        _native.cuda_check(lib.fmpi_maxpool_bwd(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), N, H, W, C, k, s, p,
                                                _DT[ctx.dtype], torch.cuda.current_stream().cuda_stream), "maxpool_bwd")
        _native.count_launch()
        return dx, None, None, None


def max_pool2d_nhwc(x: torch.Tensor, kernel_size: int, stride: int | None = None, padding: int = 0) -> torch.Tensor:
    stride = kernel_size if stride is None else stride
    if not kernel_ok(x) or padding >= kernel_size or kernel_size > 15:
        return F.max_pool2d(x, kernel_size, stride, padding)
    return _MaxPoolFn.apply(x, int(kernel_size), int(stride), int(padding))


class FusedMaxPool2d(nn.Module):
    def __init__(self, kernel_size: int, stride: int | None = None, padding: int = 0):
        super().__init__()
        self.kernel_size, self.stride, self.padding = kernel_size, stride or kernel_size, padding

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return max_pool2d_nhwc(x, self.kernel_size, self.stride, self.padding)

    def extra_repr(self) -> str:
        return f"kernel_size={self.kernel_size}, stride={self.stride}, padding={self.padding}"
