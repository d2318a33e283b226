# mirrors /root/reference/test/test_synchronize.jl:1-99
import warnings
from collections import namedtuple
import numpy as np
import torch
import fluxmpi_b200 as fm
from fluxmpi_b200 import rules as R

fm.Init(verbose=True)
r, w = fm.local_rank(), fm.total_workers()
root_rank = 0


def mk(shape):
    return torch.ones(shape) if r == root_rank else torch.zeros(shape)


# --- nested NamedTuple (test_synchronize.jl:16-25)
Inner = namedtuple("Inner", ["b", "c"])
Outer = namedtuple("Outer", ["a", "d"])
gs = Outer(a=Inner(b=mk((2, 3)), c=mk((2, 3))), d=mk((2, 3)))
gs_ = fm.synchronize(gs, root_rank=root_rank)
assert isinstance(gs_, Outer) and isinstance(gs_.a, Inner)
for t in (gs_.a.b, gs_.a.c, gs_.d):
    assert torch.equal(t, torch.ones(2, 3))
assert gs_.a.b is gs.a.b                     # in place as well as returned

# --- optimiser state (test_synchronize.jl:27-54)
opt = R.Adam(0.001)
st_opt = R.setup(opt, gs)
if r == root_rank:
    st_opt.a.b.state[0].fill_(1); st_opt.a.b.state[1].fill_(1)
    st_opt.a.c.state[0].fill_(1); st_opt.a.c.state[1].fill_(1)
    st_opt.d.state[0].fill_(1); st_opt.d.state[1].fill_(1)
    # scalars inside the state (Adam's β-powers) must come from root too
    st_opt.d.state = (st_opt.d.state[0], st_opt.d.state[1], (0.5, 0.25))
st_opt = fm.synchronize(st_opt, root_rank=root_rank)
for leaf in (st_opt.a.b, st_opt.a.c, st_opt.d):
    assert isinstance(leaf, R.Leaf)
    assert torch.equal(leaf.state[0], torch.ones(2, 3)) and torch.equal(leaf.state[1], torch.ones(2, 3))
assert st_opt.d.state[2] == (0.5, 0.25)
assert st_opt.a.b.state[2] == (0.9, 0.999)

# stateless rule: state is None (test_synchronize.jl:49-53)
st_opt = R.setup(R.Descent(0.001), gs)
with warnings.catch_warnings():
    warnings.simplefilter("error")
    st_opt = fm.synchronize(st_opt, root_rank=root_rank)
assert st_opt.d.state is None

# --- ComponentArray analogue (test_synchronize.jl:56-66)
gs = {"a": {"b": mk((2, 3)), "c": mk((2, 3))}, "d": mk((2, 3))}
cgs = fm.FlatParams.from_tree(gs)
cgs_ = fm.synchronize(cgs, root_rank=root_rank)
assert torch.equal(cgs_.a.b, torch.ones(2, 3))
assert torch.equal(cgs_.a.c, torch.ones(2, 3))
assert torch.equal(cgs_.d, torch.ones(2, 3))
assert torch.equal(cgs_["a"]["b"], torch.ones(2, 3))
assert cgs_.getdata().numel() == 18 and cgs_.getaxes() == cgs.getaxes()

# --- nested plain tuple (test_synchronize.jl:69-79)
gs = ((mk((2, 3)), mk((2, 3))), mk((2, 3)))
gs = fm.synchronize(gs, root_rank=root_rank)
assert torch.equal(gs[0][0], torch.ones(2, 3)) and torch.equal(gs[0][1], torch.ones(2, 3))
assert torch.equal(gs[1], torch.ones(2, 3))

# --- misc (test_synchronize.jl:81-97)
assert fm.synchronize(None, root_rank=root_rank) is None
x = "random" if r == root_rank else "no-random"
assert fm.synchronize(x, root_rank=root_rank) == x            # symbols are left alone
assert fm.synchronize(fm.local_rank(), root_rank=root_rank) == root_rank
assert type(fm.synchronize(fm.local_rank(), root_rank=root_rank)) is int

# --- beyond the reference
assert fm.synchronize((), root_rank=0) == () and fm.synchronize({}, root_rank=0) == {}   # :11
# other roots, floats / bools / complex / numpy scalars
last = w - 1
got = fm.synchronize({"f": 1.5 * r, "b": r == last, "c": complex(r, -r), "n": np.float32(r)}, root_rank=last)
assert got == {"f": 1.5 * last, "b": True, "c": complex(last, -last), "n": np.float32(last)}
assert isinstance(got["n"], np.float32) and isinstance(got["b"], bool)
# arrays of non-numbers are synchronised elementwise (src/synchronize.jl:20-22)
arr = np.empty(2, dtype=object)
arr[0], arr[1] = mk((2,)), float(r)
arr = fm.synchronize(arr, root_rank=0)
assert torch.equal(arr[0], torch.ones(2)) and arr[1] == 0.0
# numpy numeric arrays, int tensors, bool tensors, mixed dtypes in one tree, tied tensors
tied = mk((3,))
tree = {"np": np.full(3, float(r == 0)), "i": torch.full((2,), 7 if r == 0 else 0),
        "m": torch.tensor([r == 0, False]), "h": mk((5,)).to(torch.bfloat16), "t1": tied, "t2": [tied]}
tree = fm.synchronize(tree)
assert np.array_equal(tree["np"], np.ones(3)) and torch.equal(tree["i"], torch.full((2,), 7))
assert torch.equal(tree["m"], torch.tensor([True, False]))
assert torch.equal(tree["h"], torch.ones(5, dtype=torch.bfloat16))
assert tree["t1"] is tree["t2"][0] and torch.equal(tied, torch.ones(3))
# nn.Module through FluxMPIFluxModel returns the bare module (ext/FluxMPIFluxExt.jl:7)
torch.manual_seed(r)
net = torch.nn.Sequential(torch.nn.Linear(3, 4), torch.nn.BatchNorm1d(4))
out = fm.synchronize(fm.FluxMPIFluxModel(net), root_rank=0)
assert out is net
torch.manual_seed(0)
ref = torch.nn.Sequential(torch.nn.Linear(3, 4), torch.nn.BatchNorm1d(4))
for a, b in zip(net.state_dict().values(), ref.state_dict().values()):
    assert torch.equal(a, b)

fm.Finalize()
print("RANK_OK", r)
