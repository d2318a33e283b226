"""Fused channels-last BatchNorm (+ residual add + ReLU) kernels against PyTorch's own ops
evaluated in fp32 — forward output, running statistics, and every gradient."""

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

from fluxmpi_b200 import _native
from fluxmpi_b200.models import resnet_tiny
from fluxmpi_b200.ops.batchnorm import FusedBatchNormAct, fused_bn_act, kernel_ok

SHAPES = [(4, 64, 12, 12), (2, 8, 5, 7), (3, 256, 7, 7), (2, 2048, 3, 3), (5, 24, 9, 4), (2, 128, 56, 56)]


def reference(x, res, w, b, rm, rv, relu, momentum=0.1, eps=1e-5):
    """PyTorch ops in fp32 on the (possibly 16-bit) inputs."""
    xf = x.float().detach().requires_grad_(True)
    rf = res.float().detach().requires_grad_(True) if res is not None else None
    wf, bf = w.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True)
    y = F.batch_norm(xf, rm, rv, wf, bf, True, momentum, eps)
    if rf is not None:
        y = y + rf
    if relu:
        y = F.relu(y)
    return xf, rf, wf, bf, y


@pytest.mark.parametrize("shape", SHAPES)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("relu,use_res", [(False, False), (True, False), (True, True), (False, True)])
def test_forward_backward_match_pytorch(shape, dtype, relu, use_res):
    g = torch.Generator().manual_seed(hash((shape, relu, use_res)) % 1000)
    N, C, H, W = shape
    cl = lambda t: t.cuda().to(dtype).contiguous(memory_format=torch.channels_last)
    x = cl(torch.randn(shape, generator=g) * 1.7 + 0.3)
    res = cl(torch.randn(shape, generator=g)) if use_res else None
    w = (torch.rand(C, generator=g) + 0.5).cuda()
    b = (torch.randn(C, generator=g) * 0.2).cuda()
    rm0, rv0 = torch.zeros(C).cuda(), torch.ones(C).cuda()
    assert kernel_ok(x, res)

    xr, rr, wr, br, yr = reference(x, res, w, b, rm0.clone(), rv0.clone(), relu)
    rm_ref, rv_ref = rm0.clone(), rv0.clone()
    F.batch_norm(x.float(), rm_ref, rv_ref, w, b, True, 0.1, 1e-5)

    xk = x.clone().requires_grad_(True)
    rk = res.clone().requires_grad_(True) if res is not None else None
    wk, bk = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    rm, rv = rm0.clone(), rv0.clone()
    before = _native.launch_count
    y = fused_bn_act(xk, wk, bk, rm, rv, True, 0.1, 1e-5, relu=relu, residual=rk)
    assert _native.launch_count - before == 3, "the kernel path must be the one that runs"
    assert y.dtype == dtype and y.is_contiguous(memory_format=torch.channels_last)

    lo = dtype != torch.float32
    tol = dict(rtol=2e-2, atol=3e-2) if lo else dict(rtol=1e-4, atol=1e-4)
    assert torch.allclose(y.float(), yr, **tol), (y.float() - yr).abs().max()
    assert torch.allclose(rm, rm_ref, rtol=1e-4, atol=1e-5) and torch.allclose(rv, rv_ref, rtol=1e-4, atol=1e-5)

    dy = cl(torch.randn(shape, generator=g))
    yr.backward(dy.float())
    y.backward(dy)
    gt = dict(rtol=3e-2, atol=3e-2) if lo else dict(rtol=2e-3, atol=2e-4)
    # the ReLU mask comes from the rounded output; elements within rounding of zero may flip, so
    # allow a handful of outliers in 16-bit and compare the rest tightly
    def close(a, b, frac=0.0):
        bad = ~torch.isclose(a.float(), b.float(), **gt)
        return float(bad.float().mean()) <= frac, float((a.float() - b.float()).abs().max())
    ok, worst = close(xk.grad, xr.grad, 2e-3 if lo else 0.0)
    assert ok, worst
    scale = max(1.0, float(wr.grad.abs().max()))
    assert torch.allclose(wk.grad, wr.grad, rtol=3e-2 if lo else 1e-3, atol=(3e-2 if lo else 1e-3) * scale)
    scale = max(1.0, float(br.grad.abs().max()))
    assert torch.allclose(bk.grad, br.grad, rtol=3e-2 if lo else 1e-3, atol=(3e-2 if lo else 1e-3) * scale)
    if use_res:
        ok, worst = close(rk.grad, rr.grad, 2e-3 if lo else 0.0)
        assert ok, worst


@pytest.mark.parametrize("ratio", [10.0, 100.0, 1000.0])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("relu", [False, True])
def test_statistics_survive_a_mean_far_above_the_deviation(ratio, dtype, relu):
    """|mean| / sigma in {10, 100, 1000} per channel (post-ReLU activations, an un-normalised
    first layer): E[x^2] - mean^2 in fp32 cancels there; the pivoted sums do not. Compared with a
    float64 restatement on the values the kernel actually sees (after the 16-bit rounding)."""
    g = torch.Generator().manual_seed(int(ratio))
    N, C, H, W = 6, 64, 14, 14
    sigma = torch.rand(C, generator=g) * 0.5 + 0.25
    mean = sigma * ratio * torch.where(torch.rand(C, generator=g) < 0.5, -1.0, 1.0)
    x = torch.randn(N, C, H, W, generator=g) * sigma.view(1, C, 1, 1) + mean.view(1, C, 1, 1)
    x = x.cuda().to(dtype).contiguous(memory_format=torch.channels_last)
    w = (torch.rand(C, generator=g) + 0.5).cuda()
    b = (torch.randn(C, generator=g) * 0.2).cuda()
    rm, rv = torch.zeros(C).cuda(), torch.ones(C).cuda()
    xd = x.double().detach().requires_grad_(True)
    m64 = xd.mean(dim=(0, 2, 3))
    v64 = xd.var(dim=(0, 2, 3), unbiased=False)
    y64 = (xd - m64.view(1, C, 1, 1)) / torch.sqrt(v64.view(1, C, 1, 1) + 1e-5) * w.double().view(1, C, 1, 1) + b.double().view(1, C, 1, 1)
    if relu:
        y64 = torch.relu(y64)
    xk = x.clone().requires_grad_(True)
    y = fused_bn_act(xk, w, b, rm, rv, True, 1.0, 1e-5, relu=relu)       # momentum 1: running = batch statistics
    # forward: the normalised output is O(1); a lost variance shows up as O(1) errors at once
    tol = 5e-2 if dtype == torch.bfloat16 else 2e-3 * max(1.0, ratio / 100)
    assert torch.allclose(y.double(), y64, rtol=tol, atol=tol), float((y.double() - y64).abs().max())
    assert torch.allclose(rm.double(), m64, rtol=1e-5, atol=1e-5), float((rm.double() - m64).abs().max())
    n = N * H * W
    assert torch.allclose(rv.double(), v64 * n / (n - 1), rtol=2e-3), float((rv.double() / (v64 * n / (n - 1)) - 1).abs().max())
    dy = torch.randn(N, C, H, W, generator=g).cuda().to(dtype).contiguous(memory_format=torch.channels_last)
    y64.backward(dy.double())
    y.backward(dy)
    gtol = 6e-2 if dtype == torch.bfloat16 else 5e-3 * max(1.0, ratio / 100)
    scale = float(xd.grad.abs().max())
    bad = (xk.grad.double() - xd.grad).abs() > gtol * scale
    assert float(bad.double().mean()) <= (2e-3 if relu else 0.0), float((xk.grad.double() - xd.grad).abs().max() / scale)


def test_relu_mask_is_recomputed_from_x_without_a_residual():
    """BN + ReLU without a residual saves no forward output for backward (the mask is
    x * scale + shift > 0); with a residual it keeps y."""
    x = torch.randn(4, 32, 8, 8, device="cuda").contiguous(memory_format=torch.channels_last).requires_grad_(True)
    r = torch.randn(4, 32, 8, 8, device="cuda").contiguous(memory_format=torch.channels_last)
    bn = FusedBatchNormAct(32, relu=True).cuda()
    y = bn(x)
    saved = y.grad_fn.saved_tensors
    assert not any(t is not None and t.data_ptr() == y.data_ptr() for t in saved)
    y2 = bn(x, r)
    assert any(t is not None and t.data_ptr() == y2.data_ptr() for t in y2.grad_fn.saved_tensors)
    # and the gradient equals the one PyTorch computes from the mask of y
    ref = torch.nn.BatchNorm2d(32).cuda()
    ref.load_state_dict(bn.state_dict())
    xr = x.detach().clone().requires_grad_(True)
    dy = torch.randn_like(y)
    bn2 = FusedBatchNormAct(32, relu=True).cuda()
    bn2.load_state_dict(ref.state_dict())
    bn2(x).backward(dy)
    F.relu(ref(xr)).backward(dy)
    assert torch.allclose(x.grad, xr.grad, rtol=2e-3, atol=2e-4), float((x.grad - xr.grad).abs().max())


def test_module_matches_batchnorm2d_train_and_eval():
    torch.manual_seed(0)
    ref = torch.nn.BatchNorm2d(32).cuda()
    ours = FusedBatchNormAct(32, relu=True).cuda()
    ours.load_state_dict(ref.state_dict())                  # same keys
    x = torch.randn(6, 32, 10, 10, device="cuda").contiguous(memory_format=torch.channels_last)
    for _ in range(3):
        a = F.relu(ref(x))
        b = ours(x)
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-4)
    assert int(ours.num_batches_tracked) == 3
    assert torch.allclose(ours.running_mean, ref.running_mean, rtol=1e-4, atol=1e-5)
    assert torch.allclose(ours.running_var, ref.running_var, rtol=1e-4, atol=1e-5)
    ref.eval(); ours.eval()
    before = _native.launch_count
    assert torch.allclose(F.relu(ref(x)), ours(x), rtol=1e-4, atol=1e-4)
    assert _native.launch_count - before == 1               # inference: one elementwise kernel
    # layouts / devices the kernels do not cover fall back to the stock ops
    xn = torch.randn(6, 32, 10, 10, device="cuda")          # NCHW
    assert not kernel_ok(xn)
    assert torch.allclose(F.relu(ref(xn)), ours(xn), rtol=1e-4, atol=1e-4)
    cpu = FusedBatchNormAct(6, relu=False)
    assert cpu(torch.randn(2, 6, 3, 3)).shape == (2, 6, 3, 3)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape,k,s,p", [((4, 64, 32, 32), 3, 2, 1), ((2, 8, 7, 9), 3, 2, 1), ((3, 16, 10, 10), 2, 2, 0),
                                         ((2, 24, 11, 6), 3, 1, 1), ((1, 8, 5, 5), 5, 3, 2)])
def test_maxpool_nhwc_matches_pytorch(dtype, shape, k, s, p):
    from fluxmpi_b200.ops.pool import FusedMaxPool2d, kernel_ok as pool_ok
    g = torch.Generator().manual_seed(7)
    x = torch.randn(shape, generator=g).cuda().to(dtype).contiguous(memory_format=torch.channels_last)
    assert pool_ok(x)
    xa = x.clone().requires_grad_(True)
    xb = x.clone().requires_grad_(True)
    before = _native.launch_count
    ya = FusedMaxPool2d(k, s, p)(xa)
    assert _native.launch_count - before == 1
    yb = F.max_pool2d(xb, k, s, p)
    assert ya.shape == yb.shape and torch.equal(ya, yb)
    assert ya.is_contiguous(memory_format=torch.channels_last)
    dy = torch.randn(yb.shape, generator=g).cuda().to(dtype).contiguous(memory_format=torch.channels_last)
    ya.backward(dy)
    yb.backward(dy)
    # overlapping windows add several gradients into one input: fp32 accumulation here, so allow
    # a rounding step in 16-bit
    tol = dict(rtol=1e-2, atol=1e-2) if dtype != torch.float32 else dict(rtol=1e-6, atol=1e-6)
    assert torch.allclose(xa.grad.float(), xb.grad.float(), **tol)
    # ties: the first maximum in scan order takes the gradient, like PyTorch
    t = torch.ones(1, 8, 4, 4, device="cuda", dtype=dtype).contiguous(memory_format=torch.channels_last)
    ta, tb = t.clone().requires_grad_(True), t.clone().requires_grad_(True)
    FusedMaxPool2d(2, 2, 0)(ta).sum().backward()
    F.max_pool2d(tb, 2, 2, 0).sum().backward()
    assert torch.equal(ta.grad, tb.grad)


def test_resnet_fused_equals_unfused_training_steps():
    """Whole-model check under the bench configuration (channels_last, bf16 autocast)."""
    torch.manual_seed(3)
    a = resnet_tiny(fused_bn=False).cuda().to(memory_format=torch.channels_last)
    b = resnet_tiny(fused_bn=True).cuda().to(memory_format=torch.channels_last)
    b.load_state_dict(a.state_dict())
    oa = torch.optim.SGD(a.parameters(), lr=0.05)
    ob = torch.optim.SGD(b.parameters(), lr=0.05)
    g = torch.Generator().manual_seed(1)
    for step in range(4):
        x = torch.randn(16, 3, 64, 64, generator=g).cuda().contiguous(memory_format=torch.channels_last)
        y = torch.randint(0, 10, (16,), generator=g).cuda()
        losses = []
        for m, o in ((a, oa), (b, ob)):
            o.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = F.cross_entropy(m(x), y)
            loss.backward()
            o.step()
            losses.append(float(loss.detach()))
        assert abs(losses[0] - losses[1]) < 0.05 * max(1.0, abs(losses[0])), (step, losses)
    for (n, pa), (_, pb) in zip(a.named_parameters(), b.named_parameters()):
        d = (pa - pb).abs()
        assert float(d.mean()) < 5e-3, (n, float(d.mean()))
    for (n, ba), (_, bb) in zip(a.named_buffers(), b.named_buffers()):
        assert torch.allclose(ba.float(), bb.float(), rtol=5e-2, atol=5e-2), n
