# Round-2 odds and ends on the host backend: argument checks of fm.Group, a group that equals the
# world, a one-member group, the self-test's answer on a world without a native device side.
import torch
import fluxmpi_b200 as fm

fm.Init()
r, w = fm.local_rank(), fm.total_workers()

for bad in ([], [0, 0], [w], [-1]):
    try:
        fm.Group(bad)
        raise SystemExit(f"Group({bad}) must be refused")
    except ValueError:
        pass
try:
    fm.allreduce_(torch.ones(2), "sum", comm=object())
    raise SystemExit("a comm that is neither the world nor a Group must be refused")
except TypeError:
    pass

whole = fm.Group(list(range(w)))                 # same members as the world, its own host segment
solo = fm.Group([w - 1])                         # one member: collectives are no-ops for it
x = torch.full((5,), float(r + 1))
fm.allreduce_(x, "sum", comm=whole)
assert torch.equal(x, torch.full((5,), w * (w + 1) / 2))
y = torch.arange(3.0) * (r + 1)
fm.bcast_(y, w - 1, comm=whole)
assert torch.equal(y, torch.arange(3.0) * w)
z, req = fm.Iallreduce(torch.full((4,), float(r)), "max", comm=whole)
req.wait()
assert torch.equal(z, torch.full((4,), float(w - 1)))
fm.barrier(comm=whole)
if solo.member:
    s = torch.full((3,), 7.0)
    fm.allreduce_(s, "sum", comm=solo)
    fm.bcast_(s, 0, comm=solo)
    fm.reduce_(s, "max", 0, comm=solo)
    assert torch.equal(s, torch.full((3,), 7.0)) and solo.rank == 0 and solo.size == 1
    try:
        fm.bcast_(s, 1, comm=solo)               # roots are ranks WITHIN the group
        raise SystemExit("root 1 does not exist in a one-member group")
    except ValueError:
        pass
else:
    try:
        fm.barrier(comm=solo)
        raise SystemExit("a non-member must be refused")
    except ValueError:
        pass
whole.free(); solo.free()
try:
    fm.allreduce_(x, "sum", comm=whole)
    raise SystemExit("a freed group must be refused")
except ValueError:
    pass

rep = fm.selftest.run()                          # CPU world: nothing native to test, says so
assert rep["ok"] and rep["cases"] == 0 and "skipped" in rep, rep
fm.Finalize()
print("RANK_OK", r)
