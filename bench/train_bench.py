#!/usr/bin/env python
"""Training-step benchmarks for the other BASELINE.json configs (device-timed, both arms):

  --model vit_b16      config 3: flat-vector gradients, all-reduce fused into the Adam update (fp32 params,
                       bf16 autocast), samples/s
  --model gpt2_medium  config 4: bf16 parameters + fp32 master weights, large gradient volume, bucketed
                       fused all-reduce overlapped with backward, tokens/s
  --model mlp          config 1's model on the GPU (latency regime: 1 MB of gradients), steps/s

    python -m torch.distributed.run --nproc-per-node 8 ... bench/train_bench.py --model gpt2_medium --impl ours|ddp
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

import torch
import torch.nn.functional as F

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench as headline                                           # noqa: E402  (shared helpers)


def build(name, dev):
    from fluxmpi_b200 import models
    if name == "vit_b16":
        m = models.vit_b16().to(dev)
        mk = lambda b, g: (torch.randn(b, 3, 224, 224, generator=g).to(dev), torch.randint(0, 1000, (b,), generator=g).to(dev))
        loss = lambda net, x, y: F.cross_entropy(net(x), y)
        return m, mk, loss, 64, "samples/s", 1, torch.float32
    if name == "gpt2_medium":
        m = models.gpt2_medium().to(dev).to(torch.bfloat16)
        mk = lambda b, g: (torch.randint(0, 50257, (b, 1024), generator=g).to(dev), torch.randint(0, 50257, (b, 1024), generator=g).to(dev))
        loss = lambda net, x, y: F.cross_entropy(net(x).float().view(-1, 50257), y.view(-1))
        return m, mk, loss, 8, "tokens/s", 1024, torch.bfloat16
    m = models.quickstart_mlp().to(dev)
    mk = lambda b, g: (torch.rand(b, 1, generator=g).to(dev), None)
    loss = lambda net, x, y: ((net(x) - x ** 2) ** 2).sum()
    return m, mk, loss, 16, "steps/s", 0, torch.float32


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="vit_b16", choices=["vit_b16", "gpt2_medium", "mlp"])
    ap.add_argument("--impl", default="ours", choices=["ours", "functional", "ddp"],
                    help="ours = DataParallelEngine; functional = the reference's loop shape, "
                         "update!(state, params, grads) through a DistributedOptimizer on the module's own leaves")
    ap.add_argument("--steps", type=int, default=15)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--bucket-mb", type=float, default=64.0)
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    lrank = int(os.environ.get("LOCAL_RANK", rank))
    torch.backends.cudnn.benchmark = True
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.allow_tf32 = True

    if a.impl == "ours":
        import fluxmpi_b200 as fm
        from fluxmpi_b200 import _native
        from fluxmpi_b200 import rules as R
        from fluxmpi_b200.parallel.engine import DataParallelEngine, GraphedStep
        fm.Init()
        dev = fm.current_device()
        torch.manual_seed(rank)
        model, mk, loss_fn, b0, unit, per_sample, pdt = build(a.model, dev)
        batch = a.batch or b0
        eng = DataParallelEngine(model, fm.DistributedOptimizer(R.Adam(1e-4)), bucket_mb=a.bucket_mb)
        x, y = mk(batch, torch.Generator().manual_seed(rank))
        use_ac = pdt == torch.float32 and a.model != "mlp"

        def fn(x, y=None):
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=use_ac):
                l = loss_fn(model, x, y) / world
            l.backward()
            return l

        ins = (x,) if y is None else (x, y)
        if a.no_graph:
            def step(*i):
                l = fn(*i)
                eng.step()
                return l
            for _ in range(a.warmup):
                step(*ins)
        else:
            step = GraphedStep(eng, fn, ins, warmup=max(a.warmup, 3))
        before = _native.launch_count
        ms = headline.timed(lambda k: [step(*ins) for _ in range(k)], a.steps, fm.barrier, "ours")
        launches = _native.launch_count - before
        extra = {"buckets": len(eng.buckets), "comm": fm.runtime.state().device_world.mode +
                 ("+nvls" if eng.algo == _native.ALGO_NVLS else ""), "cuda_graph": not a.no_graph,
                 "master_weights": eng.master is not None, "params": eng.numel}
        fin = fm.Finalize
    elif a.impl == "functional":
        # /root/reference/README.md:61-67: pullback, then Optimisers.update through the
        # DistributedOptimizer. Leaves stay where the module put them; on CUDA the whole step after
        # backward is one launch for trees up to 4 MB (csrc/kernels/tree_step.cu), two beyond.
        import fluxmpi_b200 as fm
        from fluxmpi_b200 import _native
        from fluxmpi_b200 import rules as R
        fm.Init()
        dev = fm.current_device()
        torch.manual_seed(rank)
        model, mk, loss_fn, b0, unit, per_sample, pdt = build(a.model, dev)
        batch = a.batch or b0
        model = fm.synchronize(fm.FluxMPIFluxModel(model), root_rank=0)
        ps = {k: p.detach() for k, p in model.named_parameters()}
        st = R.setup(fm.DistributedOptimizer(R.Adam(1e-4)), ps)
        x, y = mk(batch, torch.Generator().manual_seed(rank))
        use_ac = pdt == torch.float32 and a.model != "mlp"
        params = list(model.parameters())

        def step(*i):
            nonlocal st
            if params[0].grad is not None:
                torch._foreach_zero_([p.grad for p in params])      # same gradient storage every step
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=use_ac):
                l = (loss_fn(model, *i) if len(i) == 2 else loss_fn(model, i[0], None)) / world
            l.backward()
            st, _ = R.update_(st, ps, {k: p.grad for k, p in model.named_parameters()})
            return l

        ins = (x,) if y is None else (x, y)
        for _ in range(max(a.warmup, 3)):
            step(*ins)
        before = _native.launch_count
        ms = headline.timed(lambda k: [step(*ins) for _ in range(k)], a.steps, fm.barrier, "ours")
        launches = _native.launch_count - before
        extra = {"comm": fm.runtime.state().device_world.mode, "cuda_graph": False,
                 "our_launches_per_step": launches / max(a.steps, 1), "params": sum(p.numel() for p in params)}
        fin = fm.Finalize
    else:
        import torch.distributed as dist
        from torch.nn.parallel import DistributedDataParallel as DDP
        torch.cuda.set_device(lrank)
        dev = torch.device("cuda", lrank)
        if world > 1:
            dist.init_process_group("nccl", device_id=dev)
        torch.manual_seed(0)
        model, mk, loss_fn, b0, unit, per_sample, pdt = build(a.model, dev)
        batch = a.batch or b0
        net = DDP(model, device_ids=[lrank], gradient_as_bucket_view=True) if world > 1 else model
        if pdt == torch.bfloat16:
            # the usual mixed-precision recipe for bf16 weights: fp32 master copy in the optimiser
            master = [p.detach().float().clone() for p in model.parameters()]
            opt = torch.optim.Adam(master, lr=1e-4, fused=True)
        else:
            master = None
            opt = torch.optim.Adam(model.parameters(), lr=1e-4, fused=True)
        x, y = mk(batch, torch.Generator().manual_seed(rank))
        use_ac = pdt == torch.float32 and a.model != "mlp"

        def step(*i):
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=use_ac):
                l = loss_fn(net, *i) if len(i) == 2 else loss_fn(net, i[0], None)
            for p in model.parameters():
                p.grad = None
            l.backward()
            if master is not None:
                for m, p in zip(master, model.parameters()):
                    m.grad = p.grad.float()
                opt.step()
                with torch.no_grad():
                    torch._foreach_copy_(list(model.parameters()), master)
            else:
                opt.step()
            return l

        ins = (x,) if y is None else (x, y)
        for _ in range(max(a.warmup, 3)):
            step(*ins)
        barrier = (lambda: dist.barrier()) if world > 1 else (lambda: None)
        ms = headline.timed(lambda k: [step(*ins) for _ in range(k)], a.steps, barrier, "ddp")
        launches, extra = 0, {"comm": "NCCL (DDP)", "optimizer": "torch.optim.Adam(fused=True)"}
        fin = (lambda: dist.destroy_process_group()) if world > 1 else (lambda: None)

    per = ms / a.steps
    units = (batch * world * max(per_sample, 1)) if unit != "steps/s" else 1
    if rank == 0:
        print(json.dumps({"metric": f"{a.model} train {unit}", "impl": a.impl, "value": round(units / (per / 1e3), 1),
                          "unit": unit, "n_gpus": world, "ms_per_step": round(per, 3), "steps": a.steps,
                          "warmup": a.warmup, "per_gpu_batch": batch, "gpu_launches": launches, "config": extra}))
    fin()


if __name__ == "__main__":
    main()
