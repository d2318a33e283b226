# DataParallelEngine without peer mappings (FMPI_FORCE_FALLBACK=1 here; in production: P2P
# switched off, or a job over several nodes): flat buffers in ordinary memory, NCCL all-reduce
# per bucket on the communication stream, local fused rule kernel. Checked against torch.optim
# on a replica whose gradients are summed with the functional API. Runs at any world size.
import copy
import os

import torch

import fluxmpi_b200 as fm
from fluxmpi_b200 import rules as R
from fluxmpi_b200.models import resnet_tiny
from fluxmpi_b200.parallel.engine import DataParallelEngine, GraphedStep

assert os.environ.get("FMPI_FORCE_FALLBACK") == "1"
fm.Init()
r, w = fm.local_rank(), fm.total_workers()
dw = fm.runtime.state().device_world
assert dw is not None and not dw.native, "the fallback was not taken"
dev = fm.current_device()
loss_fn = torch.nn.functional.cross_entropy


def batches(n, seed):
    g = torch.Generator().manual_seed(seed)
    return [(torch.randn((8, 3, 32, 32), generator=g).to(dev), torch.randint(0, 10, (8,), generator=g).to(dev))
            for _ in range(n)]


for rule, topt in [(R.Adam(1e-3), lambda ps: torch.optim.Adam(ps, lr=1e-3, eps=1e-8)),
                   (R.Momentum(1e-2, 0.9), lambda ps: torch.optim.SGD(ps, lr=1e-2, momentum=0.9)),
                   (R.RMSProp(1e-3), None)]:                    # no fused kernel: torch-op path
    torch.manual_seed(0)                                         # same initial weights on all ranks
    ref = resnet_tiny().to(dev)
    ours = copy.deepcopy(ref)
    data = batches(4, seed=100 + r)                              # every rank its own shard
    eng = DataParallelEngine(ours, fm.DistributedOptimizer(rule), bucket_mb=0.05)
    assert not eng.native and (w == 1 or len(eng.buckets) > 1)
    losses = []
    for x, y in data:
        loss = loss_fn(ours(x), y)
        loss.backward()
        eng.step()
        losses.append(float(loss.detach()))
    assert float(eng.G.abs().sum()) == 0.0
    if topt is not None:
        assert eng.fused
        opt = topt(ref.parameters())
        ref_losses = []
        for x, y in data:
            opt.zero_grad(set_to_none=True)
            loss = loss_fn(ref(x), y)
            loss.backward()
            for p in ref.parameters():
                fm.allreduce_(p.grad, "sum")                     # the reference sums gradients
            opt.step()
            ref_losses.append(float(loss.detach()))
        assert torch.allclose(torch.tensor(losses), torch.tensor(ref_losses), rtol=5e-3, atol=5e-3), (losses, ref_losses)
        for (n, a), (_, b) in zip(ours.named_parameters(), ref.named_parameters()):
            d = (a.detach() - b.detach()).abs()
            assert float(d.mean()) < 3e-4 and float(d.max()) < 3e-2, (n, float(d.mean()), float(d.max()))
    # replicas stay identical
    chk = eng.P.float().sum().reshape(1).double()
    lo, hi = chk.clone(), chk.clone()
    fm.allreduce_(lo, "min"); fm.allreduce_(hi, "max")
    assert float(lo) == float(hi), (float(lo), float(hi))
    st = eng.gather_state()
    assert st["step"] == len(data)
    eng.close()

# the whole step as one CUDA graph also works on this path at a world of one (NCCL inside a
# capture is left to the user's NCCL/torch versions, so it is only tried here when alone)
if w == 1:
    torch.manual_seed(0)
    net = resnet_tiny().to(dev)
    eng = DataParallelEngine(net, fm.DistributedOptimizer(R.Adam(1e-3)))
    x, y = batches(1, seed=7)[0]

    def fwd_bwd(x, y):
        loss = loss_fn(net(x), y)
        loss.backward()
        return loss

    g = GraphedStep(eng, fwd_bwd, (x, y))
    l0 = float(g(x, y))
    for _ in range(5):
        l1 = float(g(x, y))
    assert l1 < l0, (l0, l1)
    eng.close()

fm.fluxmpi_println("fallback engine ok")
print(f"RANK_OK {r}", flush=True)
fm.Finalize()
