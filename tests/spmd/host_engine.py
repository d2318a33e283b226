# HostDataParallelEngine (CPU): flat parameter / gradient vectors, gradient in symmetric host
# memory reduced in place, native rule loop — against torch.optim on a replica whose gradients
# are summed with the functional API. Any world size, also over pretend nodes.
import copy

import torch

import fluxmpi_b200 as fm
from fluxmpi_b200 import rules as R
from fluxmpi_b200.models.mlp import quickstart_mlp
from fluxmpi_b200.parallel.host_engine import HostDataParallelEngine, make_engine

fm.Init()
r, w = fm.local_rank(), fm.total_workers()
torch.set_num_threads(1)


def batches(n, seed):
    g = torch.Generator().manual_seed(seed)
    xs = [torch.rand(16, 1, generator=g) for _ in range(n)]
    return [(x, x ** 2) for x in xs]


for rule, topt in [(R.Adam(1e-2), lambda ps: torch.optim.Adam(ps, lr=1e-2, eps=1e-8)),
                   (R.Momentum(1e-3, 0.9), lambda ps: torch.optim.SGD(ps, lr=1e-3, momentum=0.9)),
                   (R.AdamW(1e-2, (0.9, 0.999), 1e-3), lambda ps: torch.optim.AdamW(ps, lr=1e-2, weight_decay=0.1)),
                   (R.RMSProp(1e-3), None)]:                    # no native loop: per-leaf path
    torch.manual_seed(1 + r)                                     # replicas start different ...
    ours = quickstart_mlp()
    eng = make_engine(ours, fm.DistributedOptimizer(rule))       # ... and are synchronised from rank 0
    assert isinstance(eng, HostDataParallelEngine)
    ref = copy.deepcopy(ours)
    lo, hi = eng.P.double().sum().reshape(1), eng.P.double().sum().reshape(1)
    fm.allreduce_(lo, "min"); fm.allreduce_(hi, "max")
    assert float(lo) == float(hi), "parameters not synchronised at construction"
    name, p = next(iter(ours.named_parameters()))
    assert eng.flat_params()[name].data_ptr() == p.data_ptr() and p.grad.data_ptr() == eng.flat_grads()[name].data_ptr()
    data = batches(5, seed=50 + r)
    for x, y in data:
        loss = ((ours(x) - y) ** 2).mean()
        loss.backward()
        eng.step()
    assert float(eng.G.abs().sum()) == 0.0 and eng.steps_done == len(data)
    if topt is not None:
        assert eng.fused
        opt = topt(ref.parameters())
        for x, y in data:
            opt.zero_grad(set_to_none=True)
            ((ref(x) - y) ** 2).mean().backward()
            for q in ref.parameters():
                fm.allreduce_(q.grad, "sum")
            opt.step()
        for (n, a), (_, b) in zip(ours.named_parameters(), ref.named_parameters()):
            assert torch.allclose(a, b, rtol=2e-4, atol=2e-6), (type(rule).__name__, n, float((a.detach() - b.detach()).abs().max()))
    chk = eng.P.double().sum().reshape(1)
    lo, hi = chk.clone(), chk.clone()
    fm.allreduce_(lo, "min"); fm.allreduce_(hi, "max")
    assert float(lo) == float(hi), "replicas drifted"
    st = eng.gather_state()
    assert st["step"] == len(data)
    eng2 = HostDataParallelEngine(copy.deepcopy(ours), fm.DistributedOptimizer(rule), sync_params=False)
    eng2.load_state(st)
    assert eng2.steps_done == len(data)
    if eng.fused and eng.m is not None:
        assert torch.equal(eng2.m, eng.m)

# average=True divides the summed gradient by the number of workers
torch.manual_seed(0)
a = quickstart_mlp()
e = HostDataParallelEngine(a, fm.DistributedOptimizer(R.Descent(0.1), average=True))
before = e.P.clone()
x, y = batches(1, seed=9)[0]                                    # same batch on every rank
((a(x) - y) ** 2).sum().backward()
g = e.G.clone()
e.step()
assert torch.allclose(e.P, before - 0.1 * g, rtol=1e-6, atol=1e-7)

fm.fluxmpi_println("host engine ok")
print(f"RANK_OK {r}", flush=True)
fm.Finalize()
