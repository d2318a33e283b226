"""1x1 convolution as a tcgen05 GEMM with the BatchNorm statistics in its epilogue
(csrc/kernels/gemm_bnstats.cu, ops/conv_bn.py) against plain PyTorch ops.

The kernel was written after round 1's GPU time was used up: it compiles for sm_100a but has not
run yet, so these tests are skipped unless FMPI_TEST_UNVERIFIED=1. Once they pass on a B200,
drop the gate (and only then consider turning ``fused_conv_bn`` on anywhere)."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)
if os.environ.get("FMPI_TEST_UNVERIFIED", "0") in ("0", ""):
    pytest.skip("conv + BN-statistics GEMM not yet run on a GPU (set FMPI_TEST_UNVERIFIED=1)",
                allow_module_level=True)

import torch.nn.functional as F

from fluxmpi_b200.ops import conv_bn
from fluxmpi_b200.ops.batchnorm import FusedBatchNormAct


def test_library_loads():
    assert conv_bn.gemm_lib() is not None, conv_bn.why_unavailable()


@pytest.fixture(autouse=True)
def _default_variant():
    yield
    if conv_bn.gemm_lib() is not None:
        conv_bn.set_variant(0)


@pytest.mark.parametrize("variant", [0, 1], ids=["1sm", "2sm"])
@pytest.mark.parametrize("M,N,K", [(128, 64, 64), (147, 64, 64), (4096, 256, 64), (3000, 128, 512), (25088, 64, 256)])
def test_gemm_and_column_sums(M, N, K, variant):
    torch.manual_seed(0)
    lib = conv_bn.gemm_lib()
    conv_bn.set_variant(variant)
    x = torch.randn(M, K, device="cuda").bfloat16()
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).bfloat16()
    y = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    stats = torch.full((2, N), float("nan"), device="cuda")
    ws = torch.empty(max(lib.fmpi_conv1x1_bnstats_workspace(M, N, K), 16), dtype=torch.uint8, device="cuda")
    for _ in range(2):                                      # the call clears its own sums
        rc = lib.fmpi_conv1x1_bnstats(x.data_ptr(), w.data_ptr(), y.data_ptr(), stats[0].data_ptr(),
                                      stats[1].data_ptr(), M, N, K, ws.data_ptr(), ws.numel(),
                                      torch.cuda.current_stream().cuda_stream)
        assert rc == 0, lib.fmpi_gemm_last_error()
    torch.cuda.synchronize()
    ref = x.float() @ w.float().t()
    assert torch.allclose(y.float(), ref, rtol=2e-2, atol=2e-2)
    assert torch.allclose(stats[0], ref.sum(0), rtol=1e-3, atol=1e-2 * M ** 0.5)
    assert torch.allclose(stats[1], (ref * ref).sum(0), rtol=1e-3, atol=1e-2 * M ** 0.5)


@pytest.mark.parametrize("shape", [(8, 64, 56, 56, 256), (3, 256, 7, 7, 64), (4, 128, 28, 28, 128)])
@pytest.mark.parametrize("relu,res", [(True, False), (True, True), (False, False)])
def test_conv_bn_act_matches_reference(shape, relu, res):
    Bn, K, H, W, N = shape
    torch.manual_seed(1)
    conv = torch.nn.Conv2d(K, N, 1, bias=False).cuda().to(memory_format=torch.channels_last)
    bn = FusedBatchNormAct(N, relu=relu).cuda()
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
    ref_bn = torch.nn.BatchNorm2d(N).cuda()
    ref_bn.load_state_dict(bn.state_dict())
    x = torch.randn(Bn, K, H, W, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    r = torch.randn(Bn, N, H, W, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last) if res else None
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    ra = r.clone().requires_grad_(True) if res else None
    rb = r.clone().requires_grad_(True) if res else None
    assert conv_bn.eligible(xa, conv, bn, ra), conv_bn.why_unavailable()

    z = conv_bn.conv1x1_bn_act(xa, conv, bn, ra)
    wa_grad = torch.autograd.grad(z.float().square().sum(), [xa, conv.weight, bn.weight, bn.bias] + ([ra] if res else []))

    y = F.conv2d(xb.float(), conv.weight)                   # fp32 reference of the same op
    t = ref_bn(y)
    if res:
        t = t + rb.float()
    t = torch.relu(t) if relu else t
    wb_grad = torch.autograd.grad(t.square().sum(), [xb, conv.weight, ref_bn.weight, ref_bn.bias] + ([rb] if res else []))

    assert z.is_contiguous(memory_format=torch.channels_last) and z.dtype == torch.bfloat16
    assert torch.allclose(z.float(), t, rtol=3e-2, atol=3e-2)
    assert torch.allclose(bn.running_mean, ref_bn.running_mean, rtol=1e-2, atol=1e-3)
    assert torch.allclose(bn.running_var, ref_bn.running_var, rtol=2e-2, atol=1e-3)
    for a, b in zip(wa_grad, wb_grad):
        scale = float(b.float().abs().mean()) + 1e-6
        assert float((a.float() - b.float()).abs().mean()) < 5e-2 * scale, (a.shape, scale)


def test_tiny_resnet_with_fused_conv_bn_trains():
    from fluxmpi_b200.models import resnet_tiny
    torch.manual_seed(0)
    a = resnet_tiny(fused_bn=True, fused_conv_bn=True).cuda().to(memory_format=torch.channels_last)
    b = resnet_tiny(fused_bn=True, fused_conv_bn=False).cuda().to(memory_format=torch.channels_last)
    b.load_state_dict(a.state_dict())
    x = torch.randn(16, 3, 64, 64, device="cuda").contiguous(memory_format=torch.channels_last)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        ya, yb = a(x), b(x)
    assert torch.allclose(ya.float(), yb.float(), rtol=5e-2, atol=5e-2)
    ya.float().square().mean().backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in a.parameters())
