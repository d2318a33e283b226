# Randomised differential test of rank groups on the host backend: a seeded sequence in which the
# world and three overlapping groups take turns — a rank may be inside a group collective while a
# non-member is already in the next world collective — every result compared with one computed
# from a world all-gather of the inputs. Small slots so that chunking occurs inside the groups.
import random
import sys

import torch

import fluxmpi_b200 as fm

fm.Init(slot_bytes=4096)
r, w = fm.local_rank(), fm.total_workers()
assert w >= 4
rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 120
plan = random.Random(99)                       # same decisions on every rank
data = torch.Generator().manual_seed(500 + r)

groups = [fm.Group([0, 1]), fm.Group([1, 2, 3]), fm.Group(list(range(w))[::2])]
SIZES = [1, 2, 3, 7, 511, 1024, 1025, 2049, 5000]


def gather(t):
    g = torch.zeros(w, t.numel(), dtype=torch.float64)
    g[r] = t.double()
    fm.allreduce_(g, "sum")
    return g


for it in range(rounds):
    which = plan.randrange(len(groups) + 1)           # len(groups) = the world
    kind = plan.choice(["allreduce", "bcast", "reduce", "iallreduce"])
    n = plan.choice(SIZES)
    op = plan.choice(["sum", "min", "max"])
    x = (torch.rand(n, generator=data, dtype=torch.float64) + 0.5).float()
    g_all = gather(x)                                 # world collective: everyone takes part
    if which == len(groups):
        members, comm, me = list(range(w)), fm.COMM_WORLD, r
    else:
        grp = groups[which]
        members, comm, me = grp.ranks, grp, grp.rank
        if not grp.member:
            continue                                  # non-members go straight to the next round
    root = plan.randrange(len(members)) if False else it % len(members)
    sub = g_all[members]
    red = {"sum": sub.sum(0), "min": sub.amin(0), "max": sub.amax(0)}[op]
    if kind == "allreduce":
        out = fm.allreduce_(x.clone(), op, comm=comm)
        assert torch.allclose(out.double(), red, rtol=1e-5), (it, which, kind, n, op)
    elif kind == "iallreduce":
        recv = torch.zeros_like(x)
        _, req = fm.Iallreduce(x, recv, op, comm=comm)
        req.wait()
        assert torch.allclose(recv.double(), red, rtol=1e-5), (it, which, kind, n, op)
    elif kind == "bcast":
        out = fm.bcast_(x.clone(), root, comm=comm)
        assert torch.equal(out.double(), sub[root]), (it, which, kind, n, root)
    else:
        out = fm.reduce_(x.clone(), op, root, comm=comm)
        want = red if me == root else x.double()
        assert torch.allclose(out.double(), want, rtol=1e-5), (it, which, kind, n, op, root)

fm.barrier()
for g in groups:
    g.free()
fm.Finalize()
print(f"RANK_OK {r}", flush=True)
