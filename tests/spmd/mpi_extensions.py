# mirrors /root/reference/test/test_mpi_extensions.jl:1-64 (+ dtype/op/size matrix)
import numpy as np
import torch
import fluxmpi_b200 as fm

fm.Init(verbose=True, slot_bytes=1 << 16)   # small slots so big messages exercise chunking
r, w = fm.local_rank(), fm.total_workers()

# --- Iallreduce (test_mpi_extensions.jl:9-23)
x = torch.ones(4, dtype=torch.float64)
y = torch.zeros(4, dtype=torch.float64)
y_, req = fm.Iallreduce(x, y, "+")
req.wait()
assert y_ is y and torch.equal(y, x * w)
y2 = torch.zeros(4, dtype=torch.float64)
_, req = fm.Iallreduce(x, y2, "*")
req.wait()
assert torch.equal(y2, x)
# in-place form
z = torch.full((5,), float(r + 1))
_, req = fm.Iallreduce(z, "sum")
fm.Waitall([req])
assert torch.equal(z, torch.full((5,), w * (w + 1) / 2))

# --- Ibcast (test_mpi_extensions.jl:25-32)
b = torch.ones(2, 3) if r == 0 else torch.zeros(2, 3)
b_, req = fm.Ibcast(b, 0)
req.wait()
assert b_ is b and torch.equal(b, torch.ones(2, 3))

# --- blocking wrappers (test_mpi_extensions.jl:34-62)
x = torch.ones(4, dtype=torch.float64)
assert torch.equal(fm.allreduce_(x.clone(), "+"), x * w)
assert torch.equal(fm.allreduce_(x.clone(), "*"), x)
b = torch.ones(2, 3) if r == 0 else torch.zeros(2, 3)
assert torch.equal(fm.bcast_(b, 0), torch.ones(2, 3))
x = torch.ones(4, dtype=torch.float64)
y = fm.reduce_(x.clone(), "+", 0)
if r == 0:
    assert torch.equal(y, x * w)
else:
    assert torch.equal(y, x)           # non-root unchanged (test_mpi_extensions.jl:56-60)

# --- beyond the reference: every dtype x op, odd sizes, chunking, bitwise replica equality
import operator
gen = torch.Generator().manual_seed(1234 + r)
for dt in (torch.float32, torch.float64, torch.float16, torch.bfloat16, torch.int32,
           torch.int64, torch.uint8, torch.int8, torch.int16):
    for n in (1, 7, 1000, 70001):
        if dt.is_floating_point:
            mine = (torch.rand(n, generator=gen, dtype=torch.float64) + 0.5).to(dt)
        else:
            mine = torch.randint(1, 3, (n,), generator=gen).to(dt)
        gathered = [None] * w
        allv = torch.zeros(w, n, dtype=dt)
        allv[r] = mine
        fm.allreduce_(allv, "sum")       # all-gather via sum of one-hot rows
        for op, ref in (("sum", lambda a: a.double().sum(0)), ("prod", lambda a: a.double().prod(0)),
                        ("min", lambda a: a.double().amin(0)), ("max", lambda a: a.double().amax(0))):
            if op == "prod" and n > 1000:
                continue
            out = fm.allreduce_(mine.clone(), op)
            want = ref(allv)
            if dt in (torch.uint8, torch.int8, torch.int16):
                want = want.to(torch.int64).to(dt).double()   # wrap-around like the C types
            tol = {torch.float16: 2e-2, torch.bfloat16: 1e-1, torch.float32: 1e-5}.get(dt, 1e-12)
            assert torch.allclose(out.double(), want, rtol=tol, atol=tol), (dt, n, op)
            # replicas must agree bit for bit
            chk = torch.zeros(w, n, dtype=dt)
            chk[r] = out
            fm.allreduce_(chk, "max")
            assert all(torch.equal(chk[q], out) for q in range(w)), (dt, n, op, "replica drift")

# any function as the reduction (the reference wraps a Julia function in MPI.Op,
# src/mpi_extensions.jl:52-54): folded in rank order, so non-commutative ones are well defined
t = torch.full((70001,), float(r + 1), dtype=torch.float64)         # larger than a 64 KB slot
fm.allreduce_(t, lambda a, b: 2 * a + b)
want = 1.0
for q in range(1, w):
    want = 2 * want + (q + 1)
assert torch.equal(t, torch.full((70001,), want, dtype=torch.float64)), (float(t[0]), want)
t = torch.tensor([r, -r, 7], dtype=torch.int64)
fm.reduce_(t, torch.maximum, w - 1)                                  # a torch function, to a root
assert torch.equal(t, torch.tensor([w - 1, 0, 7]) if r == w - 1 else torch.tensor([r, -r, 7]))
a, b = torch.full((5,), float(r)), torch.zeros(5)
_, req = fm.Iallreduce(a, b, lambda x, y: x - y)
req.wait()
want = 0.0
for q in range(1, w):
    want = want - q
assert torch.equal(b, torch.full((5,), want)) and torch.equal(a, torch.full((5,), float(r)))
try:
    fm.allreduce_(torch.ones(2), "median")
    raise AssertionError("unknown op name accepted")
except ValueError:
    pass

# symmetric host memory: in-place all-reduce through the peers' mappings, bit for bit equal to
# the slot path, on whole tensors, slices at odd offsets, every native dtype and op
for dt in (torch.float32, torch.float64, torch.int32, torch.int64, torch.uint8, torch.bfloat16):
    n = 70001
    sym = fm.shm_empty((n,), dt)
    assert sym.shape == (n,) and sym.dtype == dt and sym.is_contiguous()
    for op in ("sum", "max", "min"):
        src = (torch.rand(n, generator=gen, dtype=torch.float64) * 3 + 1).to(dt)
        ref = fm.allreduce_(src.clone(), op)
        sym.copy_(src)
        out = fm.allreduce_(sym, op)
        assert out is sym and torch.equal(sym, ref), (dt, op)
        sym.copy_(src)
        part = sym[1234:1234 + 4097]                        # a view inside the block
        fm.allreduce_(part, op)
        assert torch.equal(part, ref[1234:1234 + 4097]) and torch.equal(sym[:1234], src[:1234]), (dt, op)
        assert torch.equal(sym[1234 + 4097:], src[1234 + 4097:]), (dt, op)
for root in range(w):                                       # broadcast straight out of the root's block
    sym = fm.shm_empty((20011,), torch.float64)
    sym.fill_(float(r))
    fm.bcast_(sym[7:20000], root)
    assert torch.equal(sym[7:20000], torch.full((19993,), float(root), dtype=torch.float64))
    assert float(sym[0]) == r and float(sym[-1]) == r
two = fm.shm_empty((3, 5), torch.float32)                   # a second block, different shape
two.fill_(float(r + 1))
fm.allreduce_(two, "sum")
assert torch.equal(two, torch.full((3, 5), w * (w + 1) / 2))
fm.shm_free(two)                                            # collective; later blocks still work
three = fm.shm_empty((9,), torch.int64)
three.fill_(r)
fm.allreduce_(three, "max")
assert torch.equal(three, torch.full((9,), w - 1))

# a gradient tree as ONE message: leaves of odd sizes that straddle the 64 KB slots, compared
# bit for bit with one all-reduce per leaf
for dt in (torch.float32, torch.float64):
    tree = {"a": torch.rand(3, generator=gen).to(dt), "b": [torch.rand(20011, generator=gen).to(dt),
            torch.rand(1, generator=gen).to(dt)], "c": {"w": torch.rand(130, 257, generator=gen).to(dt)},
            "note": "not an array"}
    ref = [fm.allreduce_(t.clone(), "sum") for t in (tree["a"], *tree["b"], tree["c"]["w"])]
    out = fm.allreduce_gradients(tree)
    got = (out["a"], *out["b"], out["c"]["w"])
    assert all(torch.equal(g, x) for g, x in zip(got, ref)), dt
    assert out["note"] == "not an array"

# blocking calls issued while non-blocking ones are still in flight keep their place in line on
# every rank (ranks are skewed on purpose so some find the queue empty and others do not)
import random
import time
rnd = random.Random(r)
for it in range(60):
    big = torch.full((50000,), float(r + it))
    small = torch.full((3,), float(r))
    _, req = fm.Iallreduce(big, big, "sum")
    if rnd.random() < 0.5:
        time.sleep(rnd.random() * 2e-3)
    fm.allreduce_(small, "max")                       # must not overtake the queued one
    _, req2 = fm.Ibcast(small, it % w)
    fm.barrier()
    req.wait(); req2.wait()
    assert torch.equal(big, torch.full((50000,), float(sum(range(w)) + w * it))), it
    assert torch.equal(small, torch.full((3,), float(w - 1))), it

# bcast from every root, reduce to every root, big payload
for root in range(w):
    t = torch.full((33333,), float(r))
    fm.bcast_(t, root)
    assert torch.equal(t, torch.full((33333,), float(root)))
    t = torch.full((12345,), 1.0 + r)
    fm.reduce_(t, "max", root)
    assert torch.equal(t, torch.full((12345,), float(w) if r == root else 1.0 + r))

# non-contiguous tensors and numpy arrays go through the same calls
m = torch.arange(12, dtype=torch.float32).reshape(3, 4).t()
assert not m.is_contiguous()
want = m.clone() * w
fm.allreduce_(m, operator.add)
assert torch.equal(m, want)
a = np.full(6, r + 1.0)
fm.allreduce_(a, "+")
assert np.array_equal(a, np.full(6, w * (w + 1) / 2))

# errors
for bad in ("xor", 17):
    try:
        fm.allreduce_(torch.ones(1), bad)
    except ValueError:
        pass
    else:
        raise AssertionError("bad op accepted")
try:
    fm.bcast_(torch.ones(1), w)
except ValueError:
    pass
else:
    raise AssertionError("bad root accepted")

fm.Finalize()
print("RANK_OK", r)
