"""Quick start — the FluxMPI.jl README example (/root/reference/README.md:31-70) in this library.

    python -m fluxmpi_b200.launch -n 3 examples/quickstart_mlp.py           # CPU or one GPU per rank

Functional style, like Lux + Optimisers: parameters are a tree of tensors, the optimiser state is
a tree of Leaf nodes, `update_` steps them; `DistributedOptimizer` sums the gradients first.
"""
import time

import torch

import fluxmpi_b200 as fm
from fluxmpi_b200.models.mlp import functional_forward, quickstart_mlp

fm.Init()
dev = fm.current_device()

torch.manual_seed(fm.local_rank())                              # every rank starts differently ...
model = quickstart_mlp()
ps = {k: v.detach().to(dev).requires_grad_(True) for k, v in model.named_parameters()}
ps = fm.synchronize(ps, root_rank=0)                            # ... until rank 0's values win

x = torch.rand(16, 1, device=dev)
y = x ** 2

opt = fm.DistributedOptimizer(fm.Adam(0.001))
st_opt = fm.synchronize(fm.setup(opt, ps), root_rank=0)


def loss(p):                                                    # SUM semantics: scale by 1 / workers
    return ((functional_forward(p, x) - y) ** 2).sum() / fm.total_workers()


t1 = time.time()
for epoch in range(1, 101):
    l = loss(ps)
    gs = dict(zip(ps, torch.autograd.grad(l, list(ps.values()))))
    if epoch % 20 == 0:
        fm.fluxmpi_println(f"Epoch {epoch}: Loss {float(l):.6f}")
    st_opt, ps = fm.update_(st_opt, ps, gs)
fm.fluxmpi_println(f"{time.time() - t1:.3f} s")
fm.Finalize()
