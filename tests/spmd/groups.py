# Rank groups: the `comm` argument of the reference's primitives
# (/root/reference/src/mpi_extensions.jl:97,119,141 take any MPI communicator). Run at W = 4:
# two disjoint pairs run DIFFERENT collectives concurrently, then a 3-of-4 group, then the world.
# With a GPU per rank ("cuda" argument) the same on device tensors.
import sys
import torch
import fluxmpi_b200 as fm

use_cuda = len(sys.argv) > 1 and sys.argv[1] == "cuda"
fm.Init()
r, w = fm.local_rank(), fm.total_workers()
assert w >= 4
dev = fm.current_device() if use_cuda else torch.device("cpu")

lo, hi = fm.Group([0, 1]), fm.Group([2, 3])              # collective over the world
three = fm.Group([0, 2, 3])
assert lo.id != hi.id and lo.size == 2 and three.ranks == [0, 2, 3]
mine = lo if r < 2 else hi if r < 4 else None
assert (mine is None) or (mine.member and mine.rank == r % 2)
if r < 4:
    other = hi if r < 2 else lo
    assert not other.member
    try:
        fm.allreduce_(torch.ones(2, device=dev), "sum", comm=other)
        raise SystemExit("a non-member call must be refused")
    except ValueError:
        pass

for n in (1, 1000, 70001, 1 << 20):
    x = torch.full((n,), float(r + 1), device=dev)
    if r < 2:                                             # pair {0,1}: sum, then bcast from its rank 1
        fm.allreduce_(x, "sum", comm=lo)
        assert torch.equal(x, torch.full((n,), 3.0, device=dev)), (n, x[:3])
        y = torch.full((n,), float(r), device=dev)
        fm.bcast_(y, 1, comm=lo)
        assert torch.equal(y, torch.ones(n, device=dev))
    elif r < 4:                                           # pair {2,3}: max, then reduce to its rank 0
        fm.allreduce_(x, "max", comm=hi)
        assert torch.equal(x, torch.full((n,), 4.0, device=dev)), (n, x[:3])
        y = torch.full((n,), float(r), device=dev)
        fm.reduce_(y, "sum", 0, comm=hi)
        assert torch.equal(y, torch.full((n,), 5.0 if r == 2 else 3.0, device=dev))
    # straight into a collective of another communicator, no host sync in between
    z = torch.full((n,), float(r + 1), device=dev)
    if three.member:
        fm.allreduce_(z, "sum", comm=three)               # ranks 0, 2, 3: 1 + 3 + 4
        assert torch.equal(z, torch.full((n,), 8.0, device=dev)), (n, z[:3])
        _, req = fm.Ibcast(z, 2, comm=three)              # group rank 2 = world rank 3
        req.wait()
        q = torch.full((n,), float(r), device=dev)
        _, rq = fm.Iallreduce(q, "min", comm=three)
        rq.wait()
        if use_cuda:
            torch.cuda.synchronize()
        assert torch.equal(q, torch.zeros(n, device=dev))
    wv = torch.full((n,), float(r + 1), device=dev)
    fm.allreduce_(wv, "sum")                              # and the world right behind
    assert torch.equal(wv, torch.full((n,), w * (w + 1) / 2, device=dev)), (n, wv[:3])

fm.barrier(comm=mine) if mine is not None else None
fm.barrier()
if use_cuda:
    torch.cuda.synchronize()
lo.free(); hi.free(); three.free()
fm.Finalize()
print("RANK_OK", r)
