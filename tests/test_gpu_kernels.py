"""GPU tests of the sm_100a kernels on ONE B200.

Multi-rank kernels run in *emulation*: W virtual ranks = gridDim.y of one cooperative launch,
each with its own signal pad and data window in local memory (parallel/device_world.py,
EmulatedWorld) — the device code (flags, barriers, peer-pointer tables, slice arithmetic) is
the code that runs across real GPUs. Every numerical check is against a plain PyTorch
reference of the same op computed in float64 / float32.
"""

import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():  # collected on the CPU box, skipped there
    pytest.skip("needs a CUDA device", allow_module_level=True)

from fluxmpi_b200 import _native
from fluxmpi_b200 import rules as R
from fluxmpi_b200.ops import fused as F
from fluxmpi_b200.parallel.device_world import EmulatedWorld

DEV = torch.device("cuda", 0)
OPS = {"sum": _native.SUM, "prod": _native.PROD, "min": _native.MIN, "max": _native.MAX}


def ref_reduce(stack: torch.Tensor, op: str) -> torch.Tensor:
    s = stack.double()
    return {"sum": s.sum(0), "prod": s.prod(0), "min": s.amin(0), "max": s.amax(0)}[op]


def make(dtype, n, w, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    if dtype.is_floating_point:
        x = torch.rand(w, n, generator=g, dtype=torch.float64) + 0.5
    else:
        x = torch.randint(1, 3, (w, n), generator=g).double()
    return [x[r].to(dtype).to(DEV) for r in range(w)]


TOL = {torch.float16: 2e-2, torch.bfloat16: 1e-1, torch.float32: 1e-5, torch.float64: 1e-12}


@pytest.fixture(scope="module")
def worlds():
    cache = {}

    def get(w):
        if w not in cache:
            cache[w] = EmulatedWorld(w, DEV, window_mb=48, stage_mb=24)
        return cache[w]

    return get


def test_native_library_is_loaded():
    lib = _native.cuda()
    assert lib.fmpi_sigpad_bytes() > 0
    maps = open("/proc/self/maps").read()
    assert "libfmpi_cuda.so" in maps


@pytest.mark.parametrize("w", [2, 3, 8])
def test_barrier(worlds, w):
    ew = worlds(w)
    for _ in range(5):
        ew.barrier()
    torch.cuda.synchronize()


@pytest.mark.parametrize("w", [2, 3, 4, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16, torch.float64,
                                   torch.int32, torch.int64, torch.uint8, torch.int8, torch.int16])
def test_allreduce_oneshot_dtypes_ops(worlds, w, dtype):
    ew = worlds(w)
    for n in (1, 5, 1000, 4099):
        for op in ("sum", "prod", "min", "max"):
            if op == "prod" and n > 1000:
                continue
            xs = make(dtype, n, w, seed=n)
            want = ref_reduce(torch.stack(xs), op)
            if not dtype.is_floating_point:
                want = want.to(torch.int64).to(dtype).double()
            ew.allreduce(xs, OPS[op], algo=_native.ALGO_ONESHOT)
            torch.cuda.synchronize()
            tol = TOL.get(dtype, 0)
            for r in range(w):
                assert torch.allclose(xs[r].double(), want, rtol=tol, atol=tol), (dtype, n, op, r)
                assert torch.equal(xs[r], xs[0]), "replicas must be bit-identical"


@pytest.mark.parametrize("w", [2, 3, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float64, torch.int32])
def test_allreduce_twoshot_staged(worlds, w, dtype):
    ew = worlds(w)
    es = torch.empty((), dtype=dtype).element_size()
    for nbytes in (16, 4096, 1 << 20, (3 << 20) + 16 * 7):
        n = nbytes // es
        xs = make(dtype, n, w, seed=nbytes % 1000)
        outs = [torch.empty_like(x) for x in xs]
        want = ref_reduce(torch.stack(xs), "sum")
        keep = [x.clone() for x in xs]
        ew.allreduce(xs, _native.SUM, outs=outs, algo=_native.ALGO_TWOSHOT)
        torch.cuda.synchronize()
        tol = TOL.get(dtype, 0)
        for r in range(w):
            assert torch.allclose(outs[r].double(), want, rtol=tol, atol=tol)
            assert torch.equal(outs[r], outs[0])
            assert torch.equal(xs[r], keep[r]), "out-of-place call must not touch its input"


def test_allreduce_twoshot_chunks_through_small_staging():
    ew = EmulatedWorld(4, DEV, window_mb=1, stage_mb=2)     # 2 MB staging, 8 MB message
    xs = make(torch.float32, 2 << 20, 4, seed=7)
    want = ref_reduce(torch.stack(xs), "sum")
    ew.allreduce(xs, _native.SUM, algo=_native.ALGO_TWOSHOT)
    torch.cuda.synchronize()
    for r in range(4):
        assert torch.allclose(xs[r].double(), want, rtol=1e-5, atol=1e-5) and torch.equal(xs[r], xs[0])


@pytest.mark.parametrize("w", [2, 3, 8])
def test_mixed_call_sequence_reuses_staging_safely(worlds, w):
    """small / large / small / bcast back to back on one stream: the one-shot and two-shot
    areas are private, parity double-buffering protects consecutive pushes."""
    ew = worlds(w)
    for rep in range(6):
        a = make(torch.float32, 257, w, seed=rep)
        b = make(torch.float32, 1 << 18, w, seed=100 + rep)
        c = make(torch.float32, 33, w, seed=200 + rep)
        wa, wb, wc = (ref_reduce(torch.stack(t), "sum") for t in (a, b, c))
        d = [torch.full((1000,), float(r), device=DEV) for r in range(w)]
        ew.allreduce(a, _native.SUM)
        ew.allreduce(b, _native.SUM)
        ew.allreduce(c, _native.SUM)
        ew.bcast(d, rep % w)
        ew.allreduce(a, _native.MAX)
        torch.cuda.synchronize()
        for r in range(w):
            assert torch.allclose(a[r].double(), wa, rtol=1e-5)
            assert torch.allclose(b[r].double(), wb, rtol=1e-5)
            assert torch.allclose(c[r].double(), wc, rtol=1e-5)
            assert torch.equal(d[r], torch.full((1000,), float(rep % w), device=DEV))


@pytest.mark.parametrize("w", [2, 3, 8])
@pytest.mark.parametrize("algo", [_native.ALGO_ONESHOT, _native.ALGO_TWOSHOT])
def test_reduce_to_root_leaves_others_untouched(worlds, w, algo):
    """test/test_mpi_extensions.jl:52-61 on device memory."""
    ew = worlds(w)
    n = 1 << 16
    for root in (0, w - 1):
        xs = make(torch.float32, n, w, seed=root)
        keep = [x.clone() for x in xs]
        want = ref_reduce(torch.stack(xs), "sum")
        ew.allreduce(xs, _native.SUM, algo=algo, root=root)
        torch.cuda.synchronize()
        for r in range(w):
            if r == root:
                assert torch.allclose(xs[r].double(), want, rtol=1e-5)
            else:
                assert torch.equal(xs[r], keep[r])


@pytest.mark.parametrize("w", [2, 3, 8])
@pytest.mark.parametrize("algo", [_native.ALGO_ONESHOT, _native.ALGO_TWOSHOT])
def test_bcast_any(worlds, w, algo):
    ew = worlds(w)
    for nbytes in (1, 15, 16, 17, 6000, (2 << 20) + 5):
        for root in (0, w - 1):
            src = torch.randint(0, 255, (nbytes,), dtype=torch.uint8, device=DEV)
            buf = torch.zeros(w, (nbytes + 64 + 15) // 16 * 16, dtype=torch.uint8, device=DEV)
            ts = [buf[r, :nbytes] for r in range(w)]
            ts[root].copy_(src)
            ew.bcast(ts, root, algo=algo)
            torch.cuda.synchronize()
            for r in range(w):
                assert torch.equal(ts[r], src), (nbytes, root, r)
                assert int(buf[r, nbytes:].sum()) == 0, "wrote past the end of the buffer"


@pytest.mark.parametrize("w", [2, 3, 8])
def test_symmetric_window_collectives(worlds, w):
    ew = worlds(w)
    n = (1 << 18) + 4
    off = ew.heap_alloc(n * 4 + 64)
    # allreduce in place
    xs = make(torch.float32, n, w, seed=3)
    for r in range(w):
        ew.window(r, off, n * 4).view(torch.float32).copy_(xs[r])
    want = ref_reduce(torch.stack(xs), "sum")
    ew.allreduce_symm(off, n, torch.float32, _native.SUM)
    torch.cuda.synchronize()
    got = [ew.window(r, off, n * 4).view(torch.float32) for r in range(w)]
    for r in range(w):
        assert torch.allclose(got[r].double(), want, rtol=1e-5) and torch.equal(got[r], got[0])
    # reduce to root 1: others keep their data
    for r in range(w):
        got[r].copy_(xs[r])
    ew.allreduce_symm(off, n, torch.float32, _native.MAX, root=1)
    torch.cuda.synchronize()
    for r in range(w):
        if r == 1:
            assert torch.equal(got[r].double(), ref_reduce(torch.stack(xs), "max"))
        else:
            assert torch.equal(got[r], xs[r])
    # bcast: direct and scatter+all-gather
    for algo in (_native.ALGO_ONESHOT, _native.ALGO_TWOSHOT):
        for r in range(w):
            got[r].fill_(float(r))
        ew.bcast_symm(off, n * 4, w - 1, algo=algo)
        torch.cuda.synchronize()
        for r in range(w):
            assert torch.equal(got[r], torch.full((n,), float(w - 1), device=DEV))


# ------------------------------------------------------------------ chunk-pipelined kernels
@pytest.mark.parametrize("w", [2, 3, 8])
def test_pipelined_collectives_in_emulation(w):
    """ALGO_PIPE (copy-in | link | copy-out roles handing chunks over through counters and
    cross-rank flags; the peer-to-peer link flavour runs in emulation): all-reduce, reduce to a
    root and broadcast over several chunks and slot reuses, back to back without host syncs,
    with sizes that leave partial chunks and slices."""
    ew = EmulatedWorld(w, DEV, window_mb=1, stage_mb=4)     # pipe region 1 MB => ~341 KB chunks
    for rep, nbytes in enumerate((16 * w, 4096, 600_000 // 16 * 16, (3 << 20) + 16 * 7, 1 << 20)):
        n = nbytes // 4
        xs = make(torch.float32, n, w, seed=rep)
        keep = [x.clone() for x in xs]
        want = ref_reduce(torch.stack(xs), "sum")
        outs = [torch.empty_like(x) for x in xs]
        ew.allreduce(xs, _native.SUM, outs=outs, algo=_native.ALGO_PIPE)
        root = rep % w
        red = [x.clone() for x in xs]
        ew.allreduce(red, _native.MAX, algo=_native.ALGO_PIPE, root=root)
        bc = [torch.full((nbytes,), r + 1, dtype=torch.uint8, device=DEV) for r in range(w)]
        ew.bcast(bc, root, algo=_native.ALGO_PIPE)
        ew.allreduce(xs, _native.SUM, algo=_native.ALGO_PIPE)            # in place, right behind
        torch.cuda.synchronize()
        wmax = ref_reduce(torch.stack(keep), "max")
        for r in range(w):
            assert torch.allclose(outs[r].double(), want, rtol=1e-5, atol=1e-5), (w, nbytes, r)
            assert torch.equal(outs[r], outs[0]) and torch.equal(xs[r], outs[r])
            assert torch.equal(red[r].double(), wmax) if r == root else torch.equal(red[r], keep[r]), (w, nbytes, r)
            assert torch.equal(bc[r], torch.full((nbytes,), root + 1, dtype=torch.uint8, device=DEV)), (w, nbytes, r)


@pytest.mark.parametrize("w", [2, 3, 8])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float64, torch.int32])
def test_pipelined_allreduce_dtypes_and_mixing_with_other_paths(w, dtype):
    ew = EmulatedWorld(w, DEV, window_mb=1, stage_mb=4)
    es = torch.empty((), dtype=dtype).element_size()
    n = (1 << 20) // es + 16 * 3
    small = make(torch.float32, 100, w, seed=5)
    wsmall = ref_reduce(torch.stack(small), "sum")
    for rep in range(3):
        xs = make(dtype, n, w, seed=40 + rep)
        want = ref_reduce(torch.stack(xs), "sum")
        if not dtype.is_floating_point:
            want = want.to(torch.int64).to(dtype).double()
        ew.allreduce(xs, _native.SUM, algo=_native.ALGO_PIPE)
        ys = [t.clone() for t in small]
        ew.allreduce(ys, _native.SUM)                                    # one-shot in between
        zs = make(torch.float32, 1 << 17, w, seed=60 + rep)
        wz = ref_reduce(torch.stack(zs), "sum")
        ew.allreduce(zs, _native.SUM, algo=_native.ALGO_TWOSHOT)         # staged two-shot in between
        torch.cuda.synchronize()
        tol = TOL.get(dtype, 0)
        for r in range(w):
            assert torch.allclose(xs[r].double(), want, rtol=tol, atol=tol) and torch.equal(xs[r], xs[0])
            assert torch.allclose(ys[r].double(), wsmall, rtol=1e-5)
            assert torch.allclose(zs[r].double(), wz, rtol=1e-5)


@pytest.mark.parametrize("w", [2, 3, 8])
def test_symmetric_bcast_pull_and_reemit(worlds, w):
    """Mode 3 of the in-place broadcast: every rank loads its slice from the root's window and
    stores it to all others (ALGO_PIPE selects it; AUTO does from 4 MB up)."""
    ew = worlds(w)
    for n in (4, 1000, (1 << 18) + 12):
        off = ew.heap_alloc(n * 4 + 64)
        got = [ew.window(r, off, n * 4).view(torch.float32) for r in range(w)]
        for root in (0, w - 1):
            src = torch.randn(n, device=DEV)
            for r in range(w):
                got[r].fill_(float(r))
            got[root].copy_(src)
            for algo in (_native.ALGO_PIPE, _native.ALGO_PULL_P2P):
                for r in range(w):
                    if r != root:
                        got[r].fill_(float(r))
                ew.bcast_symm(off, n * 4, root, algo=algo)
                torch.cuda.synchronize()
                for r in range(w):
                    assert torch.equal(got[r], src), (w, n, root, r, algo)


# ------------------------------------------------------------------ fused allreduce + optimiser
def torch_reference(rule, p0, grads_per_step, dtype):
    """fp32 PyTorch restatement through the package's own rule objects (themselves checked
    against torch.optim in tests/test_unit_cpu.py)."""
    p = p0.float().clone()
    st = rule.init(p)
    for g in grads_per_step:
        st, dx = rule.apply(st, p, g.float())
        p = p - dx
    return p


RULES = [R.Descent(0.1), R.Momentum(0.05, 0.9), R.Nesterov(0.05, 0.9), R.Adam(1e-2, (0.9, 0.999), 1e-8),
         R.AdamW(1e-2, (0.9, 0.99), 1e-3)]


@pytest.mark.parametrize("w", [1, 2, 3, 8])
@pytest.mark.parametrize("rule", RULES, ids=lambda r: type(r).__name__ if not isinstance(r, R.OptimiserChain) else "AdamW")
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_fused_step_matches_reference(worlds, w, rule, dtype):
    ew = worlds(w) if w > 1 else EmulatedWorld(1, DEV, window_mb=16, stage_mb=1)
    n = 64 * 1031            # odd number of 64-element groups: slices are uneven
    es = 4 if dtype == torch.float32 else 2
    p_off, g_off = ew.heap_alloc(n * es), ew.heap_alloc(n * es)
    gen = torch.Generator().manual_seed(11)
    p0 = torch.randn(n, generator=gen).to(dtype)
    P = [ew.window(r, p_off, n * es).view(dtype) for r in range(w)]
    G = [ew.window(r, g_off, n * es).view(dtype) for r in range(w)]
    for r in range(w):
        P[r].copy_(p0)
    h = F.hyper_for(rule, zero_grad=True)
    need_m, need_v = F.rule_needs(h)
    m = [torch.zeros(n, device=DEV) for _ in range(w)] if need_m else None
    v = [torch.zeros(n, device=DEV) for _ in range(w)] if need_v else None
    master = [p0.float().to(DEV) for _ in range(w)] if dtype != torch.float32 else None
    step = [torch.zeros(1, dtype=torch.int64, device=DEV) for _ in range(w)]
    summed = []
    buckets = [(0, 64 * 500), (64 * 500, n)]
    for s in range(4):
        gs = [torch.randn(n, generator=gen).to(dtype) for _ in range(w)]
        for r in range(w):
            G[r].copy_(gs[r])
        summed.append(torch.stack([g.float() for g in gs]).sum(0))
        for lo, hi in buckets:
            ew.fused_step(g_off, p_off, dtype, m, v, master, lo, hi, h, step)
        for r in range(w):
            _native.cuda_check(_native.cuda().fmpi_step_advance(step[r].data_ptr(), torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        for r in range(w):
            assert int(G[r].float().abs().sum()) == 0, "gradient bucket must be cleared"
    want = torch_reference(rule, p0, summed, dtype).to(DEV)
    tol = 2e-5 if dtype == torch.float32 else 2e-2
    for r in range(w):
        got = P[r].float()
        assert torch.allclose(got, want, rtol=tol, atol=tol), (w, rule, (got - want).abs().max())
        assert torch.equal(P[r], P[0]), "replicas must be bit-identical"
    if master is not None:
        # the fp32 master copy is only maintained by the owner of a slice (lazily replicated
        # state): stitched together from the owners it must match the fp32 reference tightly
        stitched = torch.empty(n, device=DEV)
        for lo, hi in buckets:
            nvec = (hi - lo) // 8
            per = (nvec + w - 1) // w
            for r in range(w):
                a, b = lo + min(per * r, nvec) * 8, lo + min(per * r + per, nvec) * 8
                stitched[a:b] = master[r][a:b]
        assert torch.allclose(stitched, want, rtol=1e-4, atol=1e-5), (stitched - want).abs().max()
        assert torch.equal(P[0], stitched.to(dtype)), "bf16 parameters are the rounded master"
    assert int(step[0]) == 4


@pytest.mark.parametrize("w", [2, 3, 8])
@pytest.mark.parametrize("rule", RULES, ids=lambda r: type(r).__name__ if not isinstance(r, R.OptimiserChain) else "AdamW")
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_fused_step_oneshot_matches_reference(w, rule, dtype):
    """One-shot flavour (push to all peers, one barrier, every rank applies the whole rule):
    same numbers as the fp32 rule fed the summed gradient, replicated state on every rank,
    parity slots reused over back-to-back steps."""
    ew = EmulatedWorld(w, DEV, window_mb=8, stage_mb=1)
    n = 64 * 257
    es = 4 if dtype == torch.float32 else 2
    p_off, g_off = ew.heap_alloc(n * es), ew.heap_alloc(n * es)
    stage_bytes = 2 * w * (n * es)
    stage_off = ew.heap_alloc(stage_bytes)
    gen = torch.Generator().manual_seed(13)
    p0 = torch.randn(n, generator=gen).to(dtype)
    P = [ew.window(r, p_off, n * es).view(dtype) for r in range(w)]
    G = [ew.window(r, g_off, n * es).view(dtype) for r in range(w)]
    for r in range(w):
        P[r].copy_(p0)
    h = F.hyper_for(rule, zero_grad=True)
    need_m, need_v = F.rule_needs(h)
    m = [torch.zeros(n, device=DEV) for _ in range(w)] if need_m else None
    v = [torch.zeros(n, device=DEV) for _ in range(w)] if need_v else None
    master = [p0.float().to(DEV) for _ in range(w)] if dtype != torch.float32 else None
    step = [torch.zeros(1, dtype=torch.int64, device=DEV) for _ in range(w)]
    summed = []
    for s in range(5):                                   # no host sync between steps 1..3
        gs = [torch.randn(n, generator=gen).to(dtype) for _ in range(w)]
        for r in range(w):
            G[r].copy_(gs[r])
        summed.append(torch.stack([g.float() for g in gs]).sum(0))
        ew.fused_step_oneshot(g_off, p_off, dtype, m, v, master, 0, n, h, step, stage_off, stage_bytes,
                              blocks=(0, 1, 3, 0, 2)[s])
        for r in range(w):
            _native.cuda_check(_native.cuda().fmpi_step_advance(step[r].data_ptr(), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    want = torch_reference(rule, p0, summed, dtype).to(DEV)
    tol = 2e-5 if dtype == torch.float32 else 2e-2
    for r in range(w):
        assert int(G[r].float().abs().sum()) == 0, "gradient bucket must be cleared"
        assert torch.allclose(P[r].float(), want, rtol=tol, atol=tol), (w, rule, (P[r].float() - want).abs().max())
        assert torch.equal(P[r], P[0]), "replicas must be bit-identical"
        if m is not None:
            assert torch.equal(m[r], m[0]), "replicated first moment"
        if v is not None:
            assert torch.equal(v[r], v[0]), "replicated second moment"
    if master is not None:
        assert torch.allclose(master[0], want, rtol=1e-4, atol=1e-5)


def test_fused_step_grad_scale_and_device_lr():
    ew = EmulatedWorld(2, DEV, window_mb=8, stage_mb=1)
    n = 64 * 10
    p_off, g_off = ew.heap_alloc(n * 4), ew.heap_alloc(n * 4)
    P = [ew.window(r, p_off, n * 4).view(torch.float32) for r in range(2)]
    G = [ew.window(r, g_off, n * 4).view(torch.float32) for r in range(2)]
    for r in range(2):
        P[r].fill_(1.0)
        G[r].fill_(float(r + 1))
    lr = torch.tensor([0.5], device=DEV)
    h = F.hyper_for(R.Descent(123.0), grad_scale=0.5, eta_dev=lr.data_ptr())   # mean of (1, 2) = 1.5
    ew.fused_step(g_off, p_off, torch.float32, None, None, None, 0, n, h, None)
    torch.cuda.synchronize()
    for r in range(2):
        assert torch.allclose(P[r], torch.full((n,), 1.0 - 0.5 * 1.5, device=DEV))
        assert torch.equal(G[r], torch.full((n,), float(r + 1), device=DEV))   # zero_grad off


# ------------------------------------------------------------------ pointer-table kernels
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rule", RULES, ids=lambda r: type(r).__name__ if not isinstance(r, R.OptimiserChain) else "AdamW")
def test_multi_tensor_step(rule, dtype):
    gen = torch.Generator().manual_seed(5)
    shapes = [(3,), (17, 5), (8192,), (1, 1), (20000,), (64, 33)]
    ps = [torch.randn(s, generator=gen).to(dtype).to(DEV) for s in shapes]
    p0 = [p.clone() for p in ps]
    h = F.hyper_for(rule)
    need_m, need_v = F.rule_needs(h)
    ms = [torch.zeros(s, device=DEV) for s in shapes] if need_m else None
    vs = [torch.zeros(s, device=DEV) for s in shapes] if need_v else None
    gs = [torch.empty(s, dtype=dtype, device=DEV) for s in shapes]
    table = F.TensorTable(ps, gs, ms, vs)
    hist = [[] for _ in shapes]
    for step in range(3):
        for i, s in enumerate(shapes):
            g = torch.randn(s, generator=gen).to(dtype)
            gs[i].copy_(g)
            hist[i].append(g)
        F.multi_tensor_step(table, h, step, dtype)
    torch.cuda.synchronize()
    tol = 2e-5 if dtype == torch.float32 else 3e-2
    for i in range(len(shapes)):
        want = torch_reference(rule, p0[i].cpu(), hist[i], dtype)
        if dtype != torch.float32:
            # the kernel rounds the parameter to bf16 every step; restate that
            p = p0[i].cpu().float()
            st = rule.init(p)
            for g in hist[i]:
                st, dx = rule.apply(st, p, g.float())
                p = (p - dx).to(dtype).float()
            want = p
        assert torch.allclose(ps[i].float().cpu(), want, rtol=tol, atol=tol), (i, rule)


def _tree_tables(ew, w, shapes, dtype, rule, gen, misalign=False):
    """Per virtual rank: parameter / gradient / moment leaves and the device table of
    csrc/kernels/tree_step.cu ([p | g | m | v | numel | flat offset])."""
    h = F.hyper_for(rule)
    need_m, need_v = F.rule_needs(h)
    es = 4 if dtype == torch.float32 else 2
    p0 = [torch.randn(s, generator=gen).to(dtype) for s in shapes]
    ranks = []
    for r in range(w):
        def leaf(src, dt=dtype):
            if not misalign:
                return src.to(DEV).to(dt).clone()
            buf = torch.zeros(src.numel() + 1, dtype=dt, device=DEV)     # element offset 1: not 16-byte aligned
            buf[1:].copy_(src.reshape(-1))
            return buf[1:].view(src.shape)
        ps = [leaf(p) for p in p0]
        gs = [torch.empty_like(p) for p in ps]
        ms = [torch.zeros(s, device=DEV) for s in shapes] if need_m else None
        vs = [torch.zeros(s, device=DEV) for s in shapes] if need_v else None
        ranks.append((ps, gs, ms, vs))
    offs, total = [], 0
    for s_ in p0:
        offs.append(total)
        total += (s_.numel() * es + 15) // 16 * 16
    n = len(shapes)
    tabs = []
    for ps, gs, ms, vs in ranks:
        z = [0] * n
        rows = [[t.data_ptr() for t in ps], [t.data_ptr() for t in gs], [t.data_ptr() for t in ms] if ms else z,
                [t.data_ptr() for t in vs] if vs else z, [t.numel() for t in ps], offs]
        tabs.append(torch.tensor([x for row in rows for x in row], dtype=torch.int64, device=DEV))
    return h, p0, ranks, tabs, offs, total


@pytest.mark.parametrize("w", [2, 3, 8])
@pytest.mark.parametrize("rule", [R.Descent(0.1), R.Nesterov(0.05, 0.9), R.Adam(1e-2), R.AdamW(1e-2, (0.9, 0.99), 1e-3)],
                         ids=["Descent", "Nesterov", "Adam", "AdamW"])
@pytest.mark.parametrize("dtype,misalign", [(torch.float32, False), (torch.bfloat16, False), (torch.float32, True)])
def test_tree_step_oneshot_matches_reference(w, rule, dtype, misalign):
    """One launch for `update!` through a DistributedOptimizer on scattered leaves: push, one
    barrier, rank-ordered sum, rule applied in place — against the fp32 rule fed the summed
    gradient; every replica ends with the same bits."""
    ew = EmulatedWorld(w, DEV, window_mb=1, stage_mb=8)
    gen = torch.Generator().manual_seed(21)
    shapes = [(3,), (17, 5), (8192 + 5,), (1, 1), (20000,), (64, 33)]
    h, p0, ranks, tabs, offs, total = _tree_tables(ew, w, shapes, dtype, rule, gen, misalign)
    hist = [[] for _ in shapes]
    for step in range(4):
        for i, s_ in enumerate(shapes):
            gsum = None
            for r in range(w):
                g = torch.randn(s_, generator=gen).to(dtype)
                ranks[r][1][i].copy_(g)
                gsum = g.float() if gsum is None else gsum + g.float()
            hist[i].append(gsum)
        ew.tree_step_oneshot(tabs, len(shapes), total, dtype, h, step, ew.ar_off, ew.ar_bytes, blocks=(0, 1, 3, 2)[step])
    torch.cuda.synchronize()
    tol = 2e-5 if dtype == torch.float32 else 3e-2
    for i in range(len(shapes)):
        p = p0[i].float()
        st = rule.init(p)
        for g in hist[i]:
            st, dx = rule.apply(st, p, g)
            p = (p - dx).to(dtype).float()
        for r in range(w):
            got = ranks[r][0][i]
            assert torch.allclose(got.float().cpu(), p, rtol=tol, atol=tol), (w, i, r, (got.float().cpu() - p).abs().max())
            assert torch.equal(got, ranks[0][0][i]), "replicas must be bit-identical"


@pytest.mark.parametrize("w", [2, 3, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_tree_allreduce_gathers_and_reduces_in_one_launch(w, dtype):
    ew = EmulatedWorld(w, DEV, window_mb=4, stage_mb=1)
    gen = torch.Generator().manual_seed(22)
    shapes = [(5,), (300, 7), (70001,), (2,), (4096, 16)]
    h, p0, ranks, tabs, offs, total = _tree_tables(ew, w, shapes, dtype, R.Descent(0.1), gen)
    off = ew.heap_alloc(total + 64)
    es = 4 if dtype == torch.float32 else 2
    for rep in range(3):
        want = []
        for i, s_ in enumerate(shapes):
            gsum = None
            for r in range(w):
                g = torch.randn(s_, generator=gen).to(dtype)
                ranks[r][1][i].copy_(g)
                gsum = g.double() if gsum is None else gsum + g.double()
            want.append(gsum)
        ew.tree_allreduce(tabs, len(shapes), off, total, dtype, _native.SUM, blocks=(0, 2, 5)[rep])
        torch.cuda.synchronize()
        tol = 1e-5 if dtype == torch.float32 else 6e-2
        for r in range(w):
            arena = ew.window(r, off, total)
            for i, s_ in enumerate(shapes):
                n = p0[i].numel()
                got = arena[offs[i]:offs[i] + n * es].view(dtype).view(s_)
                assert torch.allclose(got.double().cpu(), want[i], rtol=tol, atol=tol), (w, rep, r, i)
                assert torch.equal(got, ew.window(0, off, total)[offs[i]:offs[i] + n * es].view(dtype).view(s_))


def test_pack_unpack_roundtrip():
    lib = _native.cuda()
    gen = torch.Generator().manual_seed(9)
    leaves = [torch.randn(n, generator=gen).to(DEV) for n in (1, 7, 4096, 100003, 16)]
    odd = torch.randn(101, generator=gen).to(DEV)[1:]          # misaligned leaf (4-byte offset)
    leaves.append(odd)
    offs, total = [], 0
    for t in leaves:
        offs.append(total)
        total += (t.numel() * 4 + 15) // 16 * 16
    n = len(leaves)
    tab = torch.tensor([t.data_ptr() for t in leaves] + offs + [t.numel() * 4 for t in leaves],
                       dtype=torch.int64).to(DEV)
    flat = torch.zeros(total, dtype=torch.uint8, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    b = tab.data_ptr()
    _native.cuda_check(lib.fmpi_pack(n, b, b + 8 * n, b + 16 * n, flat.data_ptr(), st))
    torch.cuda.synchronize()
    for t, o in zip(leaves, offs):
        assert torch.equal(flat[o:o + t.numel() * 4].view(torch.float32), t)
    keep = [t.clone() for t in leaves]
    for t in leaves:
        t.zero_()
    _native.cuda_check(lib.fmpi_unpack(n, b, b + 8 * n, b + 16 * n, flat.data_ptr(), st))
    torch.cuda.synchronize()
    for t, k in zip(leaves, keep):
        assert torch.equal(t, k)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16, torch.float64])
def test_sumsq_and_checksum(dtype):
    for n in (1, 1000, 1 << 20, (1 << 22) + 3):
        x = torch.randn(n, device=DEV).to(dtype)
        got = F.sumsq(x)
        want = (x.double() ** 2).sum()
        assert torch.allclose(got, want, rtol=1e-9), (dtype, n)
        assert torch.equal(F.sumsq(x), got), "deterministic run to run"
    a = torch.randn(100001, device=DEV)
    b = a.clone()
    assert int(F.checksum(a)) == int(F.checksum(b))
    b[777] += 1e-3
    assert int(F.checksum(a)) != int(F.checksum(b))
    c = a.clone()
    c[[1, 2]] = c[[2, 1]]
    assert int(F.checksum(a)) != int(F.checksum(c)), "checksum must be position sensitive"
