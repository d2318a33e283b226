# Device collectives across real GPUs: the reference's test_mpi_extensions.jl cases on CUDA
# tensors (which the reference never tested) plus dtype/op/size sweeps, bitwise
# replica equality, tree-level synchronize / allreduce_gradients and the non-blocking API.
import sys
import torch
import fluxmpi_b200 as fm
from fluxmpi_b200 import rules as R

quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
fm.Init(verbose=True)
r, w = fm.local_rank(), fm.total_workers()
dev = fm.current_device()
dw = fm.runtime.state().device_world
fm.fluxmpi_println(f"native={dw.native} mode={dw.mode} multicast={dw.has_multicast} {dw.why_fallback}")


def same_everywhere(t):
    lo, hi = t.clone(), t.clone()
    fm.allreduce_(lo, "min"); fm.allreduce_(hi, "max")
    return torch.equal(lo.view(torch.uint8) if False else lo, hi)


# reference cases (test_mpi_extensions.jl:9-62) on the GPU
x = torch.ones(4, dtype=torch.float64, device=dev)
assert torch.equal(fm.allreduce_(x.clone(), "+"), x * w)
assert torch.equal(fm.allreduce_(x.clone(), "*"), x)
b = torch.ones(2, 3, device=dev) if r == 0 else torch.zeros(2, 3, device=dev)
assert torch.equal(fm.bcast_(b, 0), torch.ones(2, 3, device=dev))
y = fm.reduce_(x.clone(), "+", 0)
assert torch.equal(y, x * w if r == 0 else x)
y = torch.zeros(4, dtype=torch.float64, device=dev)
_, req = fm.Iallreduce(x, y, "+"); req.wait()
assert torch.equal(y, x * w)
b = torch.ones(2, 3, device=dev) if r == 0 else torch.zeros(2, 3, device=dev)
_, req = fm.Ibcast(b, 0); req.wait(sync=True)
assert torch.equal(b, torch.ones(2, 3, device=dev))

sizes = (1, 7, 4096, 70001) if quick else (1, 7, 1000, 4099, 70001, 1 << 20, (5 << 20) + 3, 48 << 20)
dtypes = (torch.float32, torch.bfloat16) if quick else (torch.float32, torch.float64, torch.float16,
                                                        torch.bfloat16, torch.int32, torch.int64, torch.uint8)
gen = torch.Generator().manual_seed(99)
for dt in dtypes:
    for n in sizes:
        if n > (1 << 20) and dt not in (torch.float32, torch.bfloat16):
            continue
        full = (torch.rand(w, n, generator=gen, dtype=torch.float64) + 0.5) if dt.is_floating_point \
            else torch.randint(1, 3, (w, n), generator=gen).double()
        mine = full[r].to(dt).to(dev)
        ref_in = full.to(dt).double()
        for op, ref in (("sum", ref_in.sum(0)), ("max", ref_in.amax(0)), ("min", ref_in.amin(0))):
            out = fm.allreduce_(mine.clone(), op)
            want = ref if dt.is_floating_point else ref.to(torch.int64).to(dt).double()
            tol = {torch.float16: 2e-2, torch.bfloat16: 1e-1, torch.float32: 1e-5}.get(dt, 1e-12)
            assert torch.allclose(out.double().cpu(), want, rtol=tol, atol=tol), (dt, n, op)
            assert same_everywhere(out), (dt, n, op, "replica drift")
        for root in {0, w - 1}:
            t = mine.clone()
            fm.bcast_(t, root)
            assert torch.equal(t.cpu(), full[root].to(dt)), (dt, n, root)
            t = mine.clone()
            fm.reduce_(t, "sum", root)
            if r == root:
                want = ref_in.sum(0) if dt.is_floating_point else ref_in.sum(0).to(torch.int64).to(dt).double()
                tol = {torch.float16: 2e-2, torch.bfloat16: 1e-1, torch.float32: 1e-5}.get(dt, 1e-12)
                assert torch.allclose(t.double().cpu(), want, rtol=tol, atol=tol)
            else:
                assert torch.equal(t, mine)

# back-to-back mixed sizes on one stream, no host syncs in between
ts = [torch.full((n,), float(r + 1), device=dev) for n in (3, 1 << 19, 17, 1 << 21, 5)]
for _ in range(5):
    for t in ts:
        fm.allreduce_(t, "max")
torch.cuda.synchronize()
assert all(torch.equal(t, torch.full_like(t, float(w))) for t in ts)

# a non-blocking all-reduce still in flight on the communication stream while a blocking one runs
# on the current stream (legal MPI usage): the two channels have private staging
big_a = torch.full((3 << 18,), float(r + 1), device=dev)
big_b = torch.full((3 << 18,), float(2 * r + 1), device=dev)
sm_a = torch.full((1000,), float(r + 1), device=dev)
sm_b = torch.full((1000,), float(2 * r + 1), device=dev)
for _ in range(4):
    _, q1 = fm.Iallreduce(big_a, "sum")
    fm.allreduce_(big_b, "sum")
    _, q2 = fm.Iallreduce(sm_a, "max")
    fm.allreduce_(sm_b, "max")
    _, q3 = fm.Ibcast(sm_a, w - 1)
    fm.bcast_(sm_b, 0)
    fm.Waitall([q1, q2, q3])
torch.cuda.synchronize()
tri = w * (w + 1) / 2
assert torch.equal(big_a, torch.full_like(big_a, tri * w ** 3)), big_a[:3]
assert torch.equal(big_b, torch.full_like(big_b, (w * w) * w ** 3)), big_b[:3]
assert torch.equal(sm_a, torch.full_like(sm_a, float(w))) and torch.equal(sm_b, torch.full_like(sm_b, 2.0 * w - 1))

# push-mode kernels (one-shot all-reduce, staged broadcast) on the communication stream while the
# SMs are busy with somebody else's work: blocks of one launch may start far apart in time and
# must still agree on the call's parity
if not quick:
    busy = torch.randn(4096, 4096, device=dev)
    xs = [torch.full((30000,), float(r + 1 + i), device=dev) for i in range(8)]
    bs = [torch.full((300000,), float(r + i), device=dev) for i in range(8)]
    reqs = []
    for i in range(8):
        for _ in range(6):
            busy = (busy @ busy).clamp_(-1, 1)
        reqs.append(fm.Iallreduce(xs[i], "sum")[1])
        reqs.append(fm.Ibcast(bs[i], i % w)[1])
    fm.Waitall(reqs, sync=True)
    for i in range(8):
        assert torch.equal(xs[i], torch.full_like(xs[i], tri + w * i)), (i, xs[i][:3])
        assert torch.equal(bs[i], torch.full_like(bs[i], float(i % w + i))), (i, bs[i][:3])

# non-contiguous / misaligned
m = torch.arange(12, dtype=torch.float32, device=dev).reshape(3, 4).t()
want = m.clone() * w
fm.allreduce_(m, "+")
assert torch.equal(m, want)
off = torch.ones(1001, device=dev)[1:]
fm.allreduce_(off, "+")
assert torch.equal(off, torch.full((1000,), float(w), device=dev))

# tree level: synchronize (params + Adam state + scalars) and allreduce_gradients in one collective each
torch.manual_seed(r)
net = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.BatchNorm1d(7), torch.nn.Linear(7, 3)).to(dev)
ps = {k: v.detach() for k, v in net.named_parameters()}
st = R.setup(fm.DistributedOptimizer(R.Adam(1e-3)), ps)
if r == 0:
    for leaf in st.values():
        leaf.state[0].fill_(1.0)
ps = fm.synchronize(ps); st = fm.synchronize(st); fm.synchronize(fm.FluxMPIFluxModel(net))
for v in ps.values():
    assert same_everywhere(v)
for leaf in st.values():
    assert torch.equal(leaf.state[0], torch.ones_like(leaf.state[0]))
gs = {k: torch.full_like(v, float(r + 1)) for k, v in ps.items()}
gs = fm.allreduce_gradients(gs)
for v in gs.values():
    assert torch.equal(v, torch.full_like(v, w * (w + 1) / 2))
# DistributedOptimizer on device tensors == plain rule fed the summed gradient
_, ps2 = R.update(st, ps, {k: torch.ones_like(v) for k, v in ps.items()})
st_ref = R.setup(R.Adam(1e-3), ps)
for k in st_ref:
    st_ref[k].state = st[k].state
_, ps3 = R.update(st_ref, ps, {k: torch.full_like(v, float(w)) for k, v in ps.items()})
for k in ps2:
    assert torch.allclose(ps2[k], ps3[k], atol=1e-6)
fp = fm.FlatParams.from_tree({"a": torch.full((1000,), float(r), device=dev), "b": torch.ones(3, 3, device=dev)})
fm.allreduce_gradients(fp)
assert torch.equal(fp.b, torch.full((3, 3), float(w), device=dev))

# tensors in symmetric memory: same calls, in-place kernels without staging
if dw.native:
    for n in (4, 1000, 1 << 20):
        s = fm.symm_empty((n,), torch.float32)
        assert fm.is_symmetric(s) and not fm.is_symmetric(torch.ones(4, device=dev))
        s.fill_(float(r + 1))
        fm.allreduce_(s, "sum")
        assert torch.equal(s, torch.full((n,), w * (w + 1) / 2, device=dev)), n
        s.fill_(float(r))
        fm.bcast_(s, w - 1)
        assert torch.equal(s, torch.full((n,), float(w - 1), device=dev))
        s.fill_(float(r + 1))
        fm.reduce_(s, "max", 0)
        assert torch.equal(s, torch.full((n,), float(w if r == 0 else r + 1), device=dev))
    hb = fm.symm_empty((3, 1024), torch.bfloat16)
    hb.fill_(1.0)
    fm.allreduce_(hb, "+")
    assert hb.shape == (3, 1024) and torch.equal(hb, torch.full((3, 1024), float(w), dtype=torch.bfloat16, device=dev))

torch.cuda.synchronize()
fm.Finalize()
print("RANK_OK", r)
