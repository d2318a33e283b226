"""1x1 convolution + BatchNorm (+ residual add) (+ ReLU) with the BatchNorm statistics taken in
the epilogue of the convolution's GEMM (csrc/kernels/gemm_bnstats.cu, tcgen05 + TMA through
CUTLASS's sm_100 collectives) instead of by a separate pass over the convolution output.

EXPERIMENTAL — written at the end of round 1 when no GPU time was left: the kernel compiles for
sm_100a and its SASS has been inspected, but it has never run. Nothing enables it by default
(``resnet50(fused_conv_bn=True)`` / ``FMPI_CONV_BN=1`` opt in) and the GPU test that checks it
against ``F.conv2d`` + ``F.batch_norm`` is skipped unless ``FMPI_TEST_UNVERIFIED=1``.

    z = conv1x1_bn_act(x, conv, bn, residual=None)    # conv: nn.Conv2d(k=1), bn: FusedBatchNormAct

Forward: one GEMM (Y = X W^T, per-channel sum / sum of squares of Y as a by-product), the
finalize kernel, the fused apply kernel. Backward: the fused BatchNorm backward kernels, then
dX = dY W and dW = dY^T X as two library GEMMs. Whatever does not qualify (not CUDA bf16
channels_last, stride / padding / groups / bias, eval mode, missing library) takes
``bn(conv(x), residual)``.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from .. import _build, _native
from . import batchnorm as B

_lib = None
_lib_failed: Optional[str] = None
_workspace: dict = {}


def gemm_lib():
    """libfmpi_gemm.so, or None (with the reason in ``why_unavailable()``)."""
    global _lib, _lib_failed
    if _lib is not None or _lib_failed is not None:
        return _lib
    try:
        path = _build.build_gemm()
        if path is None or not path.exists():
            raise OSError("libfmpi_gemm.so is not built (no nvcc or no CUTLASS headers)")
        lib = C.CDLL(str(path))
        lib.fmpi_gemm_last_error.restype = C.c_char_p
        lib.fmpi_conv1x1_bnstats_workspace.restype = C.c_size_t
        lib.fmpi_conv1x1_bnstats_workspace.argtypes = [C.c_int, C.c_int, C.c_int]
        lib.fmpi_conv1x1_bnstats.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                             C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
        lib.fmpi_conv1x1_bnstats_variant.argtypes = [C.c_int]
        # FMPI_CONV_BN_VARIANT=1 selects the CTA-pair (cta_group::2) kernel; see set_variant
        if lib.fmpi_conv1x1_bnstats_variant(int(os.environ.get("FMPI_CONV_BN_VARIANT", "0") or 0)):
            raise OSError(lib.fmpi_gemm_last_error().decode(errors="replace"))
        _lib = lib
    except OSError as e:
        _lib_failed = str(e)
    return _lib


def set_variant(variant: int) -> None:
    """0: one SM per 128x64 tile (default); 1: CTA pairs sharing a 256x64 MMA tile."""
    lib = gemm_lib()
    if lib is None:
        raise _native.NativeError("libfmpi_gemm.so unavailable: " + str(_lib_failed))
    if lib.fmpi_conv1x1_bnstats_variant(int(variant)):
        raise ValueError(lib.fmpi_gemm_last_error().decode(errors="replace"))


def why_unavailable() -> Optional[str]:
    gemm_lib()
    return _lib_failed


def _workspace_for(device: torch.device, nbytes: int) -> torch.Tensor:
    w = _workspace.get(device.index)
    if w is None or w.numel() < nbytes:
        w = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _workspace[device.index] = w
    return w


def eligible(x: torch.Tensor, conv: torch.nn.Conv2d, bn: torch.nn.Module, residual) -> bool:
    if not (isinstance(conv, torch.nn.Conv2d) and conv.kernel_size == (1, 1) and conv.stride == (1, 1)
            and conv.padding == (0, 0) and conv.dilation == (1, 1) and conv.groups == 1 and conv.bias is None):
        return False
    if not (isinstance(bn, B.FusedBatchNormAct) and bn.training and bn.num_features == conv.out_channels):
        return False
    if not (x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 4 and x.numel() > 0
            and x.shape[1] == conv.in_channels):
        return False
    if conv.in_channels % 8 or conv.out_channels % 8:
        return False
    if not x.is_contiguous(memory_format=torch.channels_last):
        return False
    if residual is not None:
        shape = (x.shape[0], conv.out_channels, x.shape[2], x.shape[3])
        if not (tuple(residual.shape) == shape and residual.dtype == x.dtype
                and residual.is_contiguous(memory_format=torch.channels_last)):
            return False
    return gemm_lib() is not None


# The three native steps, each behind a function of tensors so that a CPU test can stand in
# plain-PyTorch versions and check the autograd plumbing around them (tests/test_unit_cpu.py).
def _k_gemm_stats(x, w2, y, scratch, M, N, K):
    """y = x w2^T; column sums of the accumulators into scratch[0:N], of their squares into [N:2N]."""
    gl = gemm_lib()
    ws = _workspace_for(x.device, gl.fmpi_conv1x1_bnstats_workspace(M, N, K))
    if gl.fmpi_conv1x1_bnstats(x.data_ptr(), w2.data_ptr(), y.data_ptr(), scratch.data_ptr(),
                               scratch.data_ptr() + 4 * N, M, N, K, ws.data_ptr(), ws.numel(), B._stream()):
        raise _native.NativeError("conv1x1_bnstats: " + gl.fmpi_gemm_last_error().decode(errors="replace"))


def _k_bn_fwd(y, residual, z, M, N, g, b, running_mean, running_var, momentum, eps, relu, mean, invstd, scratch):
    ptr = lambda t: t.data_ptr() if t is not None else None
    _native.cuda_check(_native.cuda().fmpi_bn_forward_presummed(
        y.data_ptr(), ptr(residual), z.data_ptr(), M, N, B._DT[y.dtype], ptr(g), ptr(b), ptr(running_mean),
        ptr(running_var), float(momentum), float(eps), int(relu), mean.data_ptr(), invstd.data_ptr(),
        scratch.data_ptr(), B._stream()), "bn_forward_presummed")


def _k_bn_bwd(dz, z, y, dy, dres, M, N, g, mean, invstd, dgamma, dbeta, relu):
    ptr = lambda t: t.data_ptr() if t is not None else None
    _native.cuda_check(_native.cuda().fmpi_bn_backward(
        dz.data_ptr(), ptr(z), y.data_ptr(), dy.data_ptr(), ptr(dres), M, N, B._DT[y.dtype], ptr(g),
        mean.data_ptr(), invstd.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), int(relu),
        B._scratch_for(y.device, N).data_ptr(), B._stream()), "bn_backward")


def _scratch(device, N):
    return B._scratch_for(device, N)


class _ConvBNActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, residual, gamma, beta, running_mean, running_var, momentum, eps, relu):
        Bn, K, H, W = x.shape
        N = weight.shape[0]
        M = Bn * H * W
        dev = x.device
        w2 = weight.detach().reshape(N, K).to(x.dtype).contiguous()
        y = torch.empty((Bn, N, H, W), dtype=x.dtype, device=dev, memory_format=torch.channels_last)
        scratch = _scratch(dev, N)                             # sums land in its first 2*N floats
        _k_gemm_stats(x, w2, y, scratch, M, N, K)
        z = torch.empty_like(y)
        mean = torch.empty(N, dtype=torch.float32, device=dev)
        invstd = torch.empty(N, dtype=torch.float32, device=dev)
        g = gamma.detach().float() if gamma is not None else None
        b = beta.detach().float() if beta is not None else None
        _k_bn_fwd(y, residual, z, M, N, g, b, running_mean, running_var, momentum, eps, relu, mean, invstd, scratch)
        _native.count_launch(3)
        ctx.relu, ctx.has_res = bool(relu), residual is not None
        ctx.affine = (gamma is not None, beta is not None)
        ctx.gdtype = gamma.dtype if gamma is not None else None
        ctx.wshape, ctx.wdtype = weight.shape, weight.dtype
        ctx.save_for_backward(x, w2, y, z if relu else None, g, mean, invstd)
        return z

    @staticmethod
    def backward(ctx, dz):
        x, w2, y, z, g, mean, invstd = ctx.saved_tensors
        Bn, K, H, W = x.shape
        N = w2.shape[0]
        M = Bn * H * W
        if not dz.is_contiguous(memory_format=torch.channels_last):
            dz = dz.contiguous(memory_format=torch.channels_last)
        dy = torch.empty_like(y)
        dres = torch.empty_like(y) if (ctx.has_res and ctx.relu) else None
        dgamma = torch.empty(N, dtype=torch.float32, device=x.device)
        dbeta = torch.empty(N, dtype=torch.float32, device=x.device)
        _k_bn_bwd(dz, z, y, dy, dres, M, N, g, mean, invstd, dgamma, dbeta, ctx.relu)
        _native.count_launch(3)
        if ctx.has_res and not ctx.relu:
            dres = dz                                          # d(bn + r)/dr = 1: the same gradient
        dy2 = dy.permute(0, 2, 3, 1).reshape(M, N)             # views of NHWC memory
        x2 = x.permute(0, 2, 3, 1).reshape(M, K)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = torch.mm(dy2, w2).view(Bn, H, W, K).permute(0, 3, 1, 2)      # channels_last view
        if ctx.needs_input_grad[1]:
            dw = torch.mm(dy2.t(), x2).to(ctx.wdtype).view(ctx.wshape)
        has_g, has_b = ctx.affine
        return (dx, dw, dres if ctx.has_res else None,
                dgamma.to(ctx.gdtype) if has_g else None, dbeta.to(ctx.gdtype) if has_b else None,
                None, None, None, None, None)


def enabled_by_env() -> bool:
    return os.environ.get("FMPI_CONV_BN", "0") not in ("0", "")


def conv1x1_bn_act(x: torch.Tensor, conv: torch.nn.Conv2d, bn: B.FusedBatchNormAct,
                   residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``bn(conv(x), residual)`` with the statistics fused into the convolution when possible."""
    if not eligible(x, conv, bn, residual):
        return bn(conv(x), residual)
    momentum = 0.0 if bn.momentum is None else bn.momentum
    if bn.track_running_stats and bn.num_batches_tracked is not None:
        bn.num_batches_tracked.add_(1)
        if bn.momentum is None:
            momentum = 1.0 / float(bn.num_batches_tracked)
    return _ConvBNActFn.apply(x, conv.weight, residual, bn.weight, bn.bias,
                              bn.running_mean if bn.track_running_stats else None,
                              bn.running_var if bn.track_running_stats else None,
                              momentum, bn.eps, bn.relu)
