"""The engine path on whatever the ranks have: `fm.make_engine` returns `DataParallelEngine` for
a CUDA module (flat symmetric buffers, fused NVLink all-reduce + optimiser kernel, overlap with
backward) and `HostDataParallelEngine` for a CPU one (gradient vector in symmetric host memory
reduced in place, native rule loop). Same training loop either way.

    python -m fluxmpi_b200.launch -n 3 examples/engine_any_device.py          # CPU box or GPUs
    python -m fluxmpi_b200.launch -n 4 --sim-nodes 2 examples/engine_any_device.py
    python -m fluxmpi_b200.launch -n 3 --max-restarts 1 examples/engine_any_device.py --checkpoint /tmp/ck.pt
"""
import argparse
import os
import time

import torch

import fluxmpi_b200 as fm
from fluxmpi_b200.models.mlp import quickstart_mlp

ap = argparse.ArgumentParser()
ap.add_argument("--epochs", type=int, default=200)
ap.add_argument("--checkpoint", default="")
a = ap.parse_args()

fm.Init()
r, w = fm.local_rank(), fm.total_workers()
dev = fm.current_device()

torch.manual_seed(r)                                         # replicas start different; the engine
model = quickstart_mlp().to(dev)                             # synchronises them from rank 0
engine = fm.make_engine(model, fm.DistributedOptimizer(fm.Adam(1e-3), average=True))

start = 0
if a.checkpoint and os.path.exists(a.checkpoint):           # resume (also after --max-restarts)
    state = fm.checkpoint.load(a.checkpoint, {"P": engine.P, "opt": engine.gather_state(), "epoch": 0})
    engine.load_state(state["opt"])
    start = int(state["epoch"])
    fm.fluxmpi_println(f"resumed at epoch {start}")

# every rank owns a different shard of the data
xs = torch.rand(64 * w, 1, generator=torch.Generator().manual_seed(0))
data = fm.DistributedDataContainer(fm.TensorDataset(xs, xs ** 2))
x, y = (t.to(dev) for t in data[list(range(len(data)))])

t0 = time.time()
for epoch in range(start + 1, a.epochs + 1):
    loss = ((model(x) - y) ** 2).mean()
    loss.backward()
    engine.step()                                            # all-reduce + rule; gradients cleared
    if epoch % 50 == 0:
        fm.fluxmpi_println(f"epoch {epoch}: local loss {float(loss):.6f}")
        if a.checkpoint:
            fm.checkpoint.save(a.checkpoint, {"P": engine.P, "opt": engine.gather_state(), "epoch": epoch})
if dev.type == "cuda":
    torch.cuda.synchronize()
fm.fluxmpi_println(f"{type(engine).__name__}: {(a.epochs - start) / (time.time() - t0):.0f} epochs/s")
fm.Finalize()
