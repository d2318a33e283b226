"""Data-parallel ResNet-50 training on B200s with the fast path: DistributedDataContainer ->
DeviceLoader (pinned, double-buffered) -> DataParallelEngine (flat symmetric buffers, fused
all-reduce + Adam per bucket, overlapped with backward) -> one CUDA graph per step, plus
checkpoint / resume.

    python -m fluxmpi_b200.launch -n 8 examples/train_resnet.py --steps 200
    torchrun --standalone --nproc-per-node 8 examples/train_resnet.py
"""
import argparse
import os

import torch
import torch.nn.functional as F

import fluxmpi_b200 as fm
from fluxmpi_b200.models import resnet50
from fluxmpi_b200.parallel.engine import DataParallelEngine, GraphedStep

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=100)
ap.add_argument("--batch", type=int, default=256, help="per GPU")
ap.add_argument("--lr", type=float, default=1e-3)
ap.add_argument("--ckpt", default="")
ap.add_argument("--samples", type=int, default=4096, help="synthetic dataset size (global)")
a = ap.parse_args()

fm.Init(verbose=True)
dev = fm.current_device()
torch.backends.cudnn.benchmark = True

# a synthetic ImageNet-shaped dataset, sharded into contiguous rank-ordered blocks
g = torch.Generator().manual_seed(0)
images = torch.randint(0, 256, (a.samples, 224, 224, 3), dtype=torch.uint8, generator=g).pin_memory()
labels = torch.randint(0, 1000, (a.samples,), generator=g).pin_memory()
shard = fm.DistributedDataContainer(fm.TensorDataset(images, labels))
loader = fm.DeviceLoader(shard, a.batch, shuffle=True, depth=3)

model = resnet50(fused_bn=True).to(dev).to(memory_format=torch.channels_last)
engine = DataParallelEngine(model, fm.DistributedOptimizer(fm.Adam(a.lr)))      # syncs from rank 0
if a.ckpt and os.path.exists(a.ckpt):
    fm.checkpoint.load_engine(a.ckpt, engine)
    fm.fluxmpi_println(f"resumed at step {engine.steps_done}")


def fwd_bwd(xb, yb):
    x = xb.permute(0, 3, 1, 2).float().sub_(127.5).mul_(1 / 58.0)       # NHWC bytes -> channels_last floats
    with torch.autocast("cuda", dtype=torch.bfloat16):
        loss = F.cross_entropy(model(x), yb) / fm.total_workers()       # gradients are SUMMED
    loss.backward()
    return loss


step = GraphedStep(engine, fwd_bwd, next(iter(loader)))
done = 0
while done < a.steps:
    for xb, yb in loader:
        with fm.trace.region("step"):
            loss = step(xb, yb)
        done += 1
        if done % 20 == 0:
            fm.fluxmpi_println(f"step {done}: loss {float(loss) * fm.total_workers():.4f}")
        if done >= a.steps:
            break
fm.trace.report()
if a.ckpt:
    fm.checkpoint.save_engine(a.ckpt, engine)
fm.Finalize()
