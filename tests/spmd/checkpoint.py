# checkpoint / resume through rank-0 save + synchronise-on-load; trace regions.
import sys
import torch
import fluxmpi_b200 as fm
from fluxmpi_b200.utils import checkpoint, trace

fm.Init()
r, w = fm.local_rank(), fm.total_workers()
path = sys.argv[1]

tree = {"w": torch.full((3, 2), float(r + 1)), "step": 7 + r, "nested": (torch.arange(4.0) * (r + 1), "tag")}
checkpoint.save(path, tree, root=0)                       # rank 0's values land on disk
mine = {"w": torch.zeros(3, 2), "step": 0, "nested": (torch.zeros(4), "tag")}
with trace.region("load"):
    got = checkpoint.load(path, mine, root=0)
assert torch.equal(got["w"], torch.ones(3, 2)) and got["w"] is mine["w"]
assert got["step"] == 7 and torch.equal(got["nested"][0], torch.arange(4.0)) and got["nested"][1] == "tag"
s = trace.summary()
assert s["load"]["count"] == 1 and s["load"]["total_ms"] >= 0
trace.report()
# collective counters: calls and bytes per public collective, host and device side
trace.reset()
fm.allreduce_(torch.ones(10), "sum")
fm.allreduce_(torch.ones(6, dtype=torch.float64), "max")
fm.bcast_(torch.ones(4), 0)
fm.allreduce_gradients({"a": torch.ones(3), "b": torch.ones(5)})
c = trace.counters()
assert c["allreduce/host"] == {"calls": 2, "bytes": 40 + 48}, c
assert c["bcast/host"] == {"calls": 1, "bytes": 16}
assert c["allreduce_tree/host"] == ({"calls": 1, "bytes": 32} if w > 1 else c["allreduce_tree/host"])

# engine round trip with a non-trivial `extra` (a dict with tensors and numbers that only the root
# has at load time) for a rule with a native loop and for one whose state is created lazily
from fluxmpi_b200 import rules as R
for k, rule in enumerate((R.Adam(1e-2), R.RMSProp(1e-2))):
    torch.manual_seed(5)
    net = torch.nn.Sequential(torch.nn.Linear(4, 8), torch.nn.Tanh(), torch.nn.Linear(8, 1))
    eng = fm.make_engine(net, fm.DistributedOptimizer(rule))
    for _ in range(2):
        net(torch.randn(6, 4)).square().mean().backward()
        eng.step()
    extra = {"epoch": 3, "rng": torch.arange(5), "note": "hello", "lr": [0.1, 0.01]}
    epath = path + f".engine{k}"
    checkpoint.save_engine(epath, eng, extra=extra if r == 0 else None)
    p_saved = eng.P.clone()
    torch.manual_seed(99 + r)
    net2 = torch.nn.Sequential(torch.nn.Linear(4, 8), torch.nn.Tanh(), torch.nn.Linear(8, 1))
    eng2 = fm.make_engine(net2, fm.DistributedOptimizer(rule))     # fresh: lazy state not built yet
    back = checkpoint.load_engine(epath, eng2)
    assert back["epoch"] == 3 and back["note"] == "hello" and torch.equal(back["rng"], torch.arange(5)) and back["lr"] == [0.1, 0.01]
    assert torch.equal(eng2.P, p_saved) and eng2.steps_done == 2
    x = torch.randn(6, 4, generator=torch.Generator().manual_seed(1))
    for e, n in ((eng, net), (eng2, net2)):                        # both continue identically
        n(x).square().mean().backward()
        e.step()
    assert torch.allclose(eng.P, eng2.P, rtol=1e-6, atol=1e-7), (eng.P - eng2.P).abs().max()
fm.Finalize()
print("RANK_OK", r)
