# The reference's Quick Start (/root/reference/README.md:31-70) on CPU — BASELINE config 1:
# MLP Dense(1=>256=>512=>256=>1, tanh) + DistributedOptimizer(Adam(0.001)), run at 3 ranks.
import sys
import torch
import fluxmpi_b200 as fm
from fluxmpi_b200 import rules as R
from fluxmpi_b200.models.mlp import quickstart_mlp, functional_forward

fm.Init()
r, w = fm.local_rank(), fm.total_workers()
epochs = int(sys.argv[1]) if len(sys.argv) > 1 else 20

torch.manual_seed(r)                                    # per-rank seeds differ (README.md:40)
model = quickstart_mlp()
ps = {k: v.detach().clone().requires_grad_(True) for k, v in model.named_parameters()}
st = dict(model.named_buffers())

ps = fm.synchronize(ps, root_rank=0)                    # README.md:45
st = fm.synchronize(st, root_rank=0)                    # README.md:46

x = torch.rand(16, 1)                                   # different data on every rank (README.md:48)
y = x ** 2

opt = fm.DistributedOptimizer(R.Adam(0.001))            # README.md:51
st_opt = R.setup(opt, ps)
st_opt = fm.synchronize(st_opt, root_rank=0)            # README.md:56


def loss(p):
    return ((functional_forward(p, x) - y) ** 2).sum() / fm.total_workers()


def checksum(tree):
    c = torch.stack([t.detach().double().sum() for t in tree.values()]).sum().reshape(1)
    lo, hi = c.clone(), c.clone()
    fm.allreduce_(lo, "min"); fm.allreduce_(hi, "max")
    return float(lo), float(hi)


lo, hi = checksum(ps)
assert lo == hi, "replicas differ after synchronize"

first = None
for epoch in range(1, epochs + 1):
    l = loss(ps)
    gs = dict(zip(ps.keys(), torch.autograd.grad(l, list(ps.values()))))
    gl = l.detach().clone().reshape(1)
    fm.allreduce_(gl, "sum")                            # global loss, identical on all ranks
    if epoch == 1:
        first = float(gl)
    if epoch % 10 == 0 or epoch == 1:
        fm.fluxmpi_println(f"Epoch {epoch}: Loss {float(gl):.6f}")
    st_opt, ps = R.update_(st_opt, ps, gs)              # README.md:66
    lo, hi = checksum(ps)
    assert lo == hi, f"replicas drifted at epoch {epoch}"

assert float(gl) < first, (first, float(gl))
fm.Finalize()
print("RANK_OK", r)
