"""Sub-communicators: the `comm` argument of the reference's primitives
(/root/reference/src/mpi_extensions.jl:97,119,141 take any MPI communicator). Two "islands" train
the same small model on different data and average within the island every step; every few steps
the islands exchange parameters through the world. Runs on CPU tensors or, with a GPU per rank, on
device tensors (the group collectives then run on the group's own signal pads and staging).

    python -m fluxmpi_b200.launch -n 4 examples/rank_groups.py
"""
import torch

import fluxmpi_b200 as fm

fm.Init()
r, w = fm.local_rank(), fm.total_workers()
dev = fm.current_device()
half = max(w // 2, 1)
islands = [fm.Group(list(range(0, half))), fm.Group(list(range(half, w)))] if w >= 2 else [fm.Group([0])]
mine = next(g for g in islands if g.member)            # creation was collective; use the one I am in

torch.manual_seed(0)
model = torch.nn.Linear(8, 1).to(dev)
ps = {k: p.detach() for k, p in model.named_parameters()}
st = fm.setup(fm.Descent(0.05), ps)
gen = torch.Generator().manual_seed(100 + r)
for step in range(12):
    x = torch.randn(32, 8, generator=gen).to(dev)
    y = x.sum(1, keepdim=True)
    loss = ((model(x) - y) ** 2).mean() / mine.size    # SUM semantics: scale by the island size
    gs = dict(zip(ps, torch.autograd.grad(loss, list(model.parameters()))))
    for g in gs.values():
        fm.allreduce_(g, "sum", comm=mine)             # only my island takes part
    st, _ = fm.update_(st, ps, gs)
    if step % 4 == 3:                                  # islands meet: average the parameters over the world
        for p in ps.values():
            fm.allreduce_(p, "sum")
            p.div_(w)
    fm.fluxmpi_println(f"step {step:2d} island {islands.index(mine)} loss {float(loss) * mine.size:.4f}")
for g in islands:
    g.free()
fm.Finalize()
