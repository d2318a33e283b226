# mirrors /root/reference/test/test_optimizer.jl:1-38
from collections import namedtuple
import torch
import fluxmpi_b200 as fm
from fluxmpi_b200 import rules as R
from fluxmpi_b200.tree import tree_equal, fmap

fm.Init(verbose=True)
r, w = fm.local_rank(), fm.total_workers()

# --- DistributedOptimizer (test_optimizer.jl:5-27)
opt = R.Adam(0.001)
ps = {"a": torch.zeros(4), "b": torch.zeros(4)}
st_opt = R.setup(opt, ps)
dopt = fm.DistributedOptimizer(opt)
st_dopt = R.setup(dopt, ps)

assert tree_equal(st_dopt["a"].state, st_opt["a"].state)      # :13
assert tree_equal(st_dopt["b"].state, st_opt["b"].state)      # :14
st_dopt = fm.synchronize(st_dopt)                              # :16 (no warnings)

gs = {"a": torch.ones(4), "b": torch.ones(4)}
_, ps_dopt = R.update(st_dopt, ps, gs)                         # :20
gs2 = {"a": torch.full((4,), float(w)), "b": torch.full((4,), float(w))}
_, ps_opt = R.update(st_opt, ps, gs2)                          # :23
assert torch.allclose(ps_dopt["a"], ps_opt["a"], atol=1e-5, rtol=1e-5)   # :25
assert torch.allclose(ps_dopt["b"], ps_opt["b"], atol=1e-5, rtol=1e-5)   # :26
assert torch.equal(ps["a"], torch.zeros(4))                   # update is non-mutating

# the per-leaf `apply` entry (what Optimisers.update calls in the reference, optimizer.jl:20-23)
state, dx = dopt.apply(dopt.init(ps["a"]), ps["a"], torch.ones(4))
state2, dx2 = opt.apply(opt.init(ps["a"]), ps["a"], torch.full((4,), float(w)))
assert torch.allclose(dx, dx2)

# mutating flavour, several steps, rank-dependent gradients: replicas stay identical
torch.manual_seed(0)
ps_m = {"w": torch.randn(3, 5), "b": (torch.randn(5), torch.randn(1))}
st_m = R.setup(fm.DistributedOptimizer(R.Momentum(0.1, 0.9)), ps_m)
ref_p = fmap(lambda t: t.clone(), ps_m)
ref_st = R.setup(R.Momentum(0.1, 0.9), ref_p)
for step in range(3):
    g_of = lambda q: fmap(lambda t: torch.full_like(t, float((q + 1) * (step + 1))), ps_m)
    st_m, ps_m = R.update_(st_m, ps_m, g_of(r))
    gsum = fmap(lambda t: torch.full_like(t, float(sum(q + 1 for q in range(w)) * (step + 1))), ps_m)
    ref_st, ref_p = R.update(ref_st, ref_p, gsum)
assert torch.allclose(ps_m["w"], ref_p["w"], atol=1e-6)
assert torch.allclose(ps_m["b"][0], ref_p["b"][0], atol=1e-6)

# average=True opt-in
st_avg = R.setup(fm.DistributedOptimizer(R.Descent(1.0), average=True), {"a": torch.zeros(2)})
_, p_avg = R.update(st_avg, {"a": torch.zeros(2)}, {"a": torch.full((2,), float(r))})
assert torch.allclose(p_avg["a"], torch.full((2,), -(w - 1) / 2))

# --- allreduce_gradients (test_optimizer.jl:29-36)
NT = namedtuple("NT", ["a", "b"])
gs = NT(a=torch.ones(4), b=torch.ones(4))
gs_ = fm.allreduce_gradients(gs, on_gpu=False)
assert isinstance(gs_, NT)
assert torch.equal(gs_.a, torch.full((4,), float(w)))
assert torch.equal(gs_.b, torch.full((4,), float(w)))

# nested tree, mixed dtypes, non-array leaves pass through, tied leaf reduced once
tied = torch.ones(3)
gs = {"l1": {"w": torch.ones(2, 2, dtype=torch.float64), "b": tied}, "l2": (tied, None, "tag", 3),
      "l3": [torch.ones(5, dtype=torch.float32)]}
out = fm.allreduce_gradients(gs)
assert torch.equal(out["l1"]["w"], torch.full((2, 2), float(w), dtype=torch.float64))
assert torch.equal(out["l1"]["b"], torch.full((3,), float(w))) and out["l2"][0] is out["l1"]["b"]
assert out["l2"][1] is None and out["l2"][2] == "tag" and out["l2"][3] == 3
assert torch.equal(out["l3"][0], torch.full((5,), float(w)))

# flat-vector overloads (the ComponentArray gap)
fp = fm.FlatParams.from_tree({"a": {"b": torch.ones(2, 3), "c": torch.ones(4)}, "d": torch.ones(1)})
fm.allreduce_gradients(fp)
assert torch.equal(fp.a.b, torch.full((2, 3), float(w))) and torch.equal(fp.d, torch.full((1,), float(w)))
v = fm.allreduce_gradients(torch.ones(7))
assert torch.equal(v, torch.full((7,), float(w)))

fm.Finalize()
print("RANK_OK", r)
