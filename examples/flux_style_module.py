"""Flux-style usage (/root/reference/docs/src/examples/flux.md): a model *object* is wrapped in
FluxMPIFluxModel so `synchronize` can walk its parameters and buffers; gradients are reduced
with `allreduce_gradients` and a plain (unwrapped) optimiser rule does the step.

    python -m fluxmpi_b200.launch -n 2 examples/flux_style_module.py
"""
import torch

import fluxmpi_b200 as fm

fm.Init()
dev = fm.current_device()
torch.manual_seed(fm.local_rank())
model = torch.nn.Sequential(torch.nn.Linear(4, 32), torch.nn.Tanh(), torch.nn.Linear(32, 1)).to(dev)
model = fm.synchronize(fm.FluxMPIFluxModel(model), root_rank=0)         # returns the bare module

ps = dict(model.named_parameters())
st = fm.setup(fm.Adam(1e-2), ps)                                        # note: NOT DistributedOptimizer
data = fm.DistributedDataContainer(fm.TensorDataset(torch.randn(256, 4), torch.randn(256, 1)))
for epoch in range(5):
    for xb, yb in fm.DeviceLoader(data, 16, device=dev):
        loss = ((model(xb) - yb) ** 2).mean() / fm.total_workers()
        gs = dict(zip(ps, torch.autograd.grad(loss, list(ps.values()))))
        gs = fm.allreduce_gradients(gs)                                 # one fused collective for the tree
        st, _ = fm.update_(st, ps, gs)
    fm.fluxmpi_println(f"epoch {epoch}: loss {float(loss) * fm.total_workers():.5f}")
fm.Finalize()
