# Randomised differential test of the host world: a seeded sequence of collectives with sizes
# around the slot, slice and rail boundaries (including messages with fewer elements than
# ranks), every result compared with one computed from an all-gather of the inputs. Small
# slots so that chunking, empty slices and multi-chunk trees all occur. Runs on one node and
# over pretend nodes.
import random
import sys

import torch

import fluxmpi_b200 as fm

fm.Init(slot_bytes=4096)
r, w = fm.local_rank(), fm.total_workers()
rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 150
plan = random.Random(2024)                     # same decisions on every rank
data = torch.Generator().manual_seed(77 + r)   # different values on every rank

DTYPES = [torch.float32, torch.float64, torch.int32, torch.int64, torch.bfloat16, torch.uint8]
SIZES = [1, 2, 3, w - 1, w, w + 1, 7, 511, 512, 513, 1023, 1024, 1025, 2047, 2048, 2049, 4096, 5000, 12289]


def sample(n, dt):
    if dt.is_floating_point:
        return (torch.rand(n, generator=data, dtype=torch.float64) + 0.5).to(dt)
    return torch.randint(1, 4, (n,), generator=data).to(dt)


def gather(t):
    """[W, n] tensor of every rank's t (through a sum of one-hot rows in a wide-enough dtype)."""
    wide = torch.float64 if t.dtype.is_floating_point else torch.int64
    g = torch.zeros(w, t.numel(), dtype=wide)
    g[r] = t.to(wide)
    fm.allreduce_(g, "sum")
    return g


def expect(g, op, dt):
    red = {"sum": g.sum(0), "prod": g.prod(0), "min": g.amin(0), "max": g.amax(0)}[op]
    if dt == torch.uint8:
        red = red % 256
    return red


def close(out, want, dt):
    if dt == torch.bfloat16:
        return torch.allclose(out.double(), want.double(), rtol=0.1, atol=0.1)
    if dt.is_floating_point:
        return torch.allclose(out.double(), want.double(), rtol=1e-5 if dt == torch.float32 else 1e-12)
    return torch.equal(out.to(torch.int64), want.to(torch.int64))


for it in range(rounds):
    kind = plan.choice(["allreduce", "allreduce", "bcast", "reduce", "tree", "iallreduce"])
    dt = plan.choice(DTYPES)
    n = max(1, plan.choice(SIZES))
    op = plan.choice(["sum", "min", "max"] if n > 64 else ["sum", "prod", "min", "max"])
    root = plan.randrange(w)
    x = sample(n, dt)
    if kind == "allreduce":
        g = gather(x)
        out = fm.allreduce_(x.clone(), op)
        assert close(out, expect(g, op, dt), dt), (it, kind, dt, n, op)
    elif kind == "iallreduce":
        g = gather(x)
        recv = torch.zeros_like(x)
        _, req = fm.Iallreduce(x, recv, op)
        small = torch.tensor([float(r)])
        fm.allreduce_(small, "max")                    # a blocking call queued behind it
        req.wait()
        assert close(recv, expect(g, op, dt), dt) and float(small) == w - 1, (it, kind, dt, n, op)
    elif kind == "bcast":
        g = gather(x)
        out = fm.bcast_(x.clone(), root)
        assert close(out, g[root], dt), (it, kind, dt, n, root)
    elif kind == "reduce":
        g = gather(x)
        out = fm.reduce_(x.clone(), op, root)
        want = expect(g, op, dt) if r == root else x
        assert close(out, want, dt), (it, kind, dt, n, op, root)
    else:  # a gradient tree as one message
        if not dt.is_floating_point or dt == torch.bfloat16:
            dt = torch.float32
        leaves = [sample(max(1, plan.choice(SIZES)), dt) for _ in range(plan.randrange(1, 6))]
        gs = [gather(t) for t in leaves]
        tree = {"a": leaves[0].clone(), "rest": [t.clone() for t in leaves[1:]], "tag": it}
        out = fm.allreduce_gradients(tree)
        got = [out["a"], *out["rest"]]
        assert all(close(o, g.sum(0), dt) for o, g in zip(got, gs)) and out["tag"] == it, (it, kind, dt)

fm.barrier()
fm.Finalize()
print(f"RANK_OK {r}", flush=True)
