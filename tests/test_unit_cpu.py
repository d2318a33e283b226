"""Single-process unit tests (no launch): tree walk, rules vs torch.optim oracles, FlatParams,
preferences, data sharding arithmetic."""
import math
import os
from collections import OrderedDict, namedtuple

import numpy as np
import pytest
import torch

import fluxmpi_b200 as fm
from fluxmpi_b200 import rules as R
from fluxmpi_b200 import tree as T
from fluxmpi_b200.utils import prefs


# ------------------------------------------------------------------------------ tree
def test_fmap_structure_and_identity_cache():
    NT = namedtuple("NT", ["x", "y"])
    shared = torch.ones(2)
    tr = {"a": NT(x=shared, y=[torch.zeros(1), None, "s"]), "b": (shared, 3), "c": OrderedDict(z=1.5)}
    calls = []

    def f(v):
        calls.append(v)
        return v * 2 if isinstance(v, torch.Tensor) else v

    out = T.fmap(f, tr)
    assert isinstance(out["a"], NT) and isinstance(out["c"], OrderedDict)
    assert out["a"].x is out["b"][0]                      # tied stays tied
    assert sum(1 for c in calls if c is shared) == 1      # visited once (Functors cache)
    assert out["a"].y[1] is None and out["a"].y[2] == "s" and out["b"][1] == 3
    assert torch.equal(out["a"].x, torch.full((2,), 2.0))
    assert T.fmap(f, ()) == () and T.fmap(f, {}) == {}


def test_leaves_and_tree_equal():
    a = {"p": torch.ones(2), "q": (np.arange(3), 2)}
    b = {"p": torch.ones(2), "q": (np.arange(3), 2)}
    assert T.tree_equal(a, b)
    b["q"] = (np.arange(3), 3)
    assert not T.tree_equal(a, b)
    assert len(T.leaves(a)) == 3
    assert [p for p, _ in T.leaves_with_path(a)] == [("p",), ("q", 0), ("q", 1)]


# ------------------------------------------------------------------------------ rules
def _run_rule(rule, steps, x0, grads):
    ps = {"w": x0.clone()}
    st = R.setup(rule, ps)
    for g in grads[:steps]:
        st, ps = R.update(st, ps, {"w": g})
    return ps["w"]


def _run_torch(opt_cls, kwargs, steps, x0, grads):
    p = torch.nn.Parameter(x0.clone())
    opt = opt_cls([p], **kwargs)
    for g in grads[:steps]:
        p.grad = g.clone()
        opt.step()
    return p.detach()


@pytest.fixture
def problem():
    g = torch.Generator().manual_seed(0)
    x0 = torch.randn(64, generator=g, dtype=torch.float64)
    grads = [torch.randn(64, generator=g, dtype=torch.float64) for _ in range(6)]
    return x0, grads


def test_adam_matches_torch(problem):
    x0, grads = problem
    ours = _run_rule(R.Adam(1e-2, (0.9, 0.999), 1e-8), 6, x0, grads)
    ref = _run_torch(torch.optim.Adam, dict(lr=1e-2, betas=(0.9, 0.999), eps=1e-8), 6, x0, grads)
    assert torch.allclose(ours, ref, rtol=1e-10, atol=1e-12)


def test_adamw_matches_torch_with_gamma_lr_wd(problem):
    x0, grads = problem
    lr, wd = 1e-2, 0.1
    ours = _run_rule(R.AdamW(lr, (0.9, 0.999), lr * wd, 1e-8), 5, x0, grads)
    ref = _run_torch(torch.optim.AdamW, dict(lr=lr, weight_decay=wd, eps=1e-8), 5, x0, grads)
    assert torch.allclose(ours, ref, rtol=1e-9, atol=1e-11)


def test_descent_and_momentum_match_torch(problem):
    x0, grads = problem
    assert torch.allclose(_run_rule(R.Descent(0.1), 4, x0, grads),
                          _run_torch(torch.optim.SGD, dict(lr=0.1), 4, x0, grads))
    # Optimisers' Momentum folds eta into the velocity: v = rho*v + eta*g ; x -= v.
    # torch: buf = rho*buf + g ; x -= lr*buf  — identical trajectories for constant lr.
    assert torch.allclose(_run_rule(R.Momentum(0.05, 0.9), 5, x0, grads),
                          _run_torch(torch.optim.SGD, dict(lr=0.05, momentum=0.9), 5, x0, grads),
                          rtol=1e-10, atol=1e-12)
    assert torch.allclose(_run_rule(R.Nesterov(0.05, 0.9), 5, x0, grads),
                          _run_torch(torch.optim.SGD, dict(lr=0.05, momentum=0.9, nesterov=True),
                                     5, x0, grads), rtol=1e-10, atol=1e-12)


def test_other_rules_match_torch(problem):
    x0, grads = problem
    assert torch.allclose(_run_rule(R.AdaMax(2e-3), 5, x0, grads),
                          _run_torch(torch.optim.Adamax, dict(lr=2e-3, eps=1e-8), 5, x0, grads),
                          rtol=1e-6, atol=1e-8)
    # RMSProp: same recurrence as torch (alpha=rho, eps outside sqrt)
    assert torch.allclose(_run_rule(R.RMSProp(1e-3, 0.9, 1e-8), 5, x0, grads),
                          _run_torch(torch.optim.RMSprop, dict(lr=1e-3, alpha=0.9, eps=1e-8),
                                     5, x0, grads), rtol=1e-9, atol=1e-11)


def test_every_rule_steps_and_keeps_shapes(problem):
    x0, grads = problem
    for rule in (R.AdaGrad(), R.AdaDelta(), R.AMSGrad(), R.NAdam(), R.AdaBelief(), R.WeightDecay(),
                 R.ClipGrad(0.5), R.ClipNorm(1.0), R.OptimiserChain(R.ClipNorm(1.0), R.Adam(1e-3))):
        out = _run_rule(rule, 3, x0, grads)
        assert out.shape == x0.shape and torch.isfinite(out).all() and not torch.equal(out, x0)


def test_setup_tied_and_nontrainable_leaves():
    w = torch.ones(3)
    ps = {"a": w, "b": {"tied": w, "idx": torch.arange(3), "n": 5, "none": None}}
    st = R.setup(R.Adam(), ps)
    assert st["a"] is st["b"]["tied"]
    assert st["b"]["idx"] is None and st["b"]["n"] is None and st["b"]["none"] is None
    st2, ps2 = R.update(st, ps, {"a": torch.ones(3), "b": {"tied": torch.ones(3), "idx": None,
                                                           "n": None, "none": None}})
    assert ps2["a"] is ps2["b"]["tied"]               # stepped once, still tied
    assert ps2["b"]["n"] == 5


def test_update_inplace_mutates_and_adjust():
    ps = {"w": torch.ones(4)}
    st = R.setup(R.Descent(0.5), ps)
    st, ps2 = R.update_(st, ps, {"w": torch.ones(4)})
    assert ps2["w"] is ps["w"] and torch.equal(ps["w"], torch.full((4,), 0.5))
    st = R.adjust(st, 0.25)
    assert st["w"].rule.eta == 0.25
    st = R.adjust(R.setup(fm.DistributedOptimizer(R.Adam(1e-3)), ps), 1e-2)
    assert st["w"].rule.optimizer.eta == 1e-2


@pytest.mark.parametrize("dt", [torch.float32, torch.float64])
@pytest.mark.parametrize("rule", [R.Descent(0.1), R.Momentum(0.01, 0.9), R.Nesterov(0.01, 0.9),
                                  R.Adam(1e-3), R.AdamW(1e-3, (0.9, 0.99), 1e-3)],
                         ids=["descent", "momentum", "nesterov", "adam", "adamw"])
def test_native_host_rule_loop_matches_per_leaf_ops(rule, dt):
    """``update_`` on CPU leaves is one native call (csrc/host_rules.cpp); same numbers as the
    per-leaf array operations to the last bit or two, tied leaves stepped once, state in place."""
    import copy
    from fluxmpi_b200.ops import host_rules
    torch.manual_seed(0)
    ps = {"a": torch.randn(257, 33, dtype=dt), "b": {"w": torch.randn(1000, dtype=dt), "k": 3}}
    ps["b"]["tied"] = ps["b"]["w"]
    ps2 = copy.deepcopy(ps)
    ps2["b"]["tied"] = ps2["b"]["w"]
    st, st2 = R.setup(rule, ps), R.setup(rule, ps2)
    m0 = st["a"].state[0] if isinstance(st["a"].state, tuple) and isinstance(st["a"].state[0], torch.Tensor) else None
    for _ in range(6):
        gs = {"a": torch.randn(257, 33, dtype=dt), "b": {"w": torch.randn(1000, dtype=dt), "k": None}}
        gs["b"]["tied"] = gs["b"]["w"]
        assert host_rules.try_tree_step_host_(st, ps, gs)            # the native path qualifies
        st2, ps2 = R._walk_update(st2, ps2, gs, True, {})             # per-leaf torch ops
    eps = 4 * torch.finfo(dt).eps
    for a, b in ((ps["a"], ps2["a"]), (ps["b"]["w"], ps2["b"]["w"])):
        assert torch.allclose(a, b, rtol=eps, atol=eps), float((a - b).abs().max())
    assert ps["b"]["tied"] is ps["b"]["w"] and ps["b"]["k"] == 3
    assert T.tree_equal(st["a"].state[2] if m0 is not None else None,
                        st2["a"].state[2] if m0 is not None else None)   # same bias-correction powers
    if m0 is not None:
        assert st["a"].state[0] is m0                                   # moments updated in place
    # a rule without a native loop, or a non-contiguous leaf, declines and changes nothing
    before = ps["a"].clone()
    assert not host_rules.try_tree_step_host_(R.setup(R.RMSProp(), ps), ps, gs)
    pt = {"w": torch.randn(8, 6, dtype=dt).t()}
    assert not host_rules.try_tree_step_host_(R.setup(rule, pt), pt, {"w": torch.ones(6, 8, dtype=dt)})
    assert torch.equal(ps["a"], before)


def test_distributed_optimizer_state_equals_inner_rule():
    """test/test_optimizer.jl:13-14 — without any world."""
    ps = {"a": torch.zeros(4), "b": torch.zeros(4)}
    s1 = R.setup(R.Adam(0.001), ps)
    s2 = R.setup(fm.DistributedOptimizer(R.Adam(0.001)), ps)
    assert T.tree_equal(s1["a"].state, s2["a"].state) and T.tree_equal(s1["b"].state, s2["b"].state)


# ------------------------------------------------------------------------------ FlatParams
def test_flatparams_views_and_roundtrip():
    NT = namedtuple("NT", ["u", "v"])
    tree = {"a": NT(u=torch.arange(6.).reshape(2, 3), v=torch.ones(2)), "b": [torch.zeros(1)], "meta": "x"}
    fp = fm.FlatParams.from_tree(tree)
    assert len(fp) == 9 and fp.getdata().shape == (9,)
    assert torch.equal(fp.a.u, tree["a"].u) and fp.a.u.data_ptr() == fp.data.data_ptr()
    fp.a.v.fill_(7)
    assert torch.equal(fp.data[6:8], torch.full((2,), 7.0))
    back = fp.to_tree()
    assert isinstance(back["a"], NT) and isinstance(back["b"], list)
    assert [n for n, _ in fp.named_views()] == ["a.u", "a.v", "b.0"]
    other = fp.like(torch.zeros(9))
    assert other.getaxes() == fp.getaxes() and torch.equal(other.a.u, torch.zeros(2, 3))
    with pytest.raises(AttributeError):
        fp.nope
    aligned = fm.FlatParams.from_tree(tree, align_elems=4)
    assert aligned.axes.entries["a"].entries["v"][0] == 8


# ------------------------------------------------------------------------------ prefs / data
def test_preferences_roundtrip(tmp_path, monkeypatch):
    f = tmp_path / "LocalPreferences.toml"
    f.write_text('[Other]\nkeep = 1\n')
    monkeypatch.setenv("FMPI_PREFS_FILE", str(f))
    assert prefs.load_preference(prefs.KEY_DISABLE_P2P, False) is False
    fm.disable_p2p_support()
    assert prefs.load_preference(prefs.KEY_DISABLE_P2P, False) is True
    assert prefs.load_all()["Other"]["keep"] == 1
    fm.disable_cudampi_support(disable=False)             # reference name kept as alias
    assert prefs.load_preference(prefs.KEY_DISABLE_P2P, True) is False


@pytest.mark.parametrize("n,w", [(10, 2), (10, 3), (10, 4), (9, 4), (1, 3), (0, 2), (16, 8)])
def test_ddc_block_partition(n, w):
    data = list(range(n))
    shards = [fm.DistributedDataContainer(data, rank=r, world=w) for r in range(w)]
    per = math.ceil(n / w) if n else 0
    assert [x for s in shards for x in s] == data                  # disjoint cover, rank order
    for r, s in enumerate(shards):
        assert len(s) == max(0, min(per, n - r * per))
        assert all(s[i] == data[s.idxs[i]] for i in range(len(s)))


def test_not_initialized_error():
    if fm.Initialized():
        pytest.skip("world already up in this process")
    with pytest.raises(fm.FluxMPINotInitializedError):
        fm.local_rank()
    with pytest.raises(fm.FluxMPINotInitializedError):
        fm.allreduce_(torch.ones(1), "sum")


def test_launcher_propagates_exit_code(tmp_path):
    from fluxmpi_b200.launch import launch
    s = tmp_path / "s.py"
    s.write_text("import os, sys, time\nif os.environ['RANK']=='1': sys.exit(7)\ntime.sleep(60)\n")
    code, _ = launch(3, [str(s)], timeout=30, capture=True)
    assert code == 7


# ------------------------------------------------------------------------------ tuner
def test_tuner_lookup(tmp_path, monkeypatch):
    import json
    from fluxmpi_b200.parallel.tuner import Tuner
    f = tmp_path / "tuned.json"
    f.write_text(json.dumps({"4": {"allreduce_any": [
        {"bytes": 1 << 20, "algo": "twoshot", "blocks": 32, "threads": 256},
        {"bytes": 64 << 20, "algo": "nvls", "blocks": 148, "threads": 512}]}}))
    t = Tuner(4, True, str(f))
    assert t.choice("allreduce_any", 2 << 20) == ("twoshot", 32, 256)       # nearest on a log scale
    assert t.choice("allreduce_any", 100 << 20) == ("nvls", 148, 512)
    # large in-place broadcasts are left to the library (pull + re-emit) unless the table was re-measured
    f2 = tmp_path / "tuned2.json"
    f2.write_text(json.dumps({"4": {"bcast_symm": [
        {"bytes": 4 << 20, "algo": "nvls", "blocks": 8, "threads": 512},
        {"bytes": 64 << 20, "algo": "nvls", "blocks": 16, "threads": 512}]}}))
    assert Tuner(4, True, str(f2)).choice("bcast_symm", 6 << 20) == ("nvls", 8, 512)
    assert Tuner(4, True, str(f2)).choice("bcast_symm", 64 << 20) is None
    f2.write_text(json.dumps({"4": {"bcast_symm": [
        {"bytes": 64 << 20, "algo": "pipe", "blocks": 64, "threads": 512}]}}))
    assert Tuner(4, True, str(f2)).choice("bcast_symm", 64 << 20) == ("pipe", 64, 512)
    assert t.choice("allreduce_any", 1000) is None                          # below the measured range
    assert t.choice("bcast_any", 1 << 20) is None                           # no table for this op
    ref, obj = t.launch("allreduce_any", 3 << 20)
    assert (obj.blocks, obj.threads, obj.algo) == (32, 256, 2)
    assert t.launch("allreduce_any", 3 << 20)[1] is obj                     # cached per size bucket
    no_mc = Tuner(4, False, str(f))
    assert no_mc.choice("allreduce_any", 100 << 20) == ("twoshot", 32, 256)  # nvls rows dropped
    assert Tuner(8, True, str(f)).choice("allreduce_any", 1 << 20) is None
    monkeypatch.setenv("FMPI_NO_TUNE", "1")
    assert Tuner(4, True, str(f)).choice("allreduce_any", 2 << 20) is None


def test_shipped_tuned_table_is_well_formed():
    import json
    from pathlib import Path
    import fluxmpi_b200.parallel.tuner as tuner_mod
    data = json.loads(Path(tuner_mod.__file__).with_name("tuned.json").read_text())
    for world in ("2", "4", "8"):
        for op in ("allreduce_any", "reduce_any", "bcast_any", "allreduce_symm", "bcast_symm"):
            rows = data[world][op]
            assert rows and all(r["algo"] in ("oneshot", "twoshot", "nvls") and 1 <= r["blocks"] <= 160 for r in rows)


# ------------------------------------------------------------------------------ model ops (CPU fallbacks)
def test_fused_ops_fall_back_to_stock_ops_off_gpu():
    from fluxmpi_b200.models import resnet_tiny
    from fluxmpi_b200.ops.batchnorm import FusedBatchNormAct
    from fluxmpi_b200.ops.pool import FusedMaxPool2d
    torch.manual_seed(0)
    ref, ours = torch.nn.BatchNorm2d(6), FusedBatchNormAct(6, relu=True)
    ours.load_state_dict(ref.state_dict())                        # identical keys
    x, r = torch.randn(4, 6, 5, 5), torch.randn(4, 6, 5, 5)
    assert torch.allclose(ours(x, r), torch.relu(ref(x) + r), atol=1e-6)
    assert torch.allclose(ours.running_var, ref.running_var) and int(ours.num_batches_tracked) == 1
    ours.eval(); ref.eval()
    assert torch.allclose(ours(x), torch.relu(ref(x)), atol=1e-6)
    assert torch.equal(FusedMaxPool2d(3, 2, 1)(x), torch.nn.functional.max_pool2d(x, 3, 2, 1))
    a, b = resnet_tiny(fused_bn=True), resnet_tiny(fused_bn=False)
    b.load_state_dict(a.state_dict())
    img = torch.randn(2, 3, 32, 32)
    assert torch.allclose(a(img), b(img), atol=1e-5)
    assert sum(p.numel() for p in a.parameters()) == sum(p.numel() for p in b.parameters())


def test_device_loader_cpu_shuffle_and_remainder():
    xs, ys = torch.arange(23.0).reshape(23, 1), torch.arange(23)
    data = fm.DistributedDataContainer(fm.TensorDataset(xs, ys), rank=0, world=1)
    loader = fm.DeviceLoader(data, 5, device="cpu", shuffle=True, drop_last=False, seed=3)
    seen = [int(v) for _, yb in loader for v in yb]
    assert sorted(seen) == list(range(23)) and seen != list(range(23))
    assert len(loader) == 5
    again = [int(v) for _, yb in loader for v in yb]
    assert sorted(again) == list(range(23)) and again != seen      # a new permutation every epoch
    assert len(fm.DeviceLoader(data, 5, device="cpu", drop_last=True)) == 4


def test_symmetric_lookup_and_dispatch_with_a_mock_library(monkeypatch):
    """DeviceWorld's symmetric-buffer lookup and kernel dispatch, driven on CPU with a recording
    stand-in for the CUDA library: offsets, bounds, and the rule that a buffer without a
    multicast alias (a registered user buffer) never gets the in-switch algorithm."""
    import ctypes as C
    from fluxmpi_b200 import _native
    from fluxmpi_b200.parallel import device_world as DW
    from fluxmpi_b200.parallel.tuner import Tuner

    class FakeTensor:
        def __init__(self, ptr, n, dtype=torch.float32):
            self.ptr, self.n, self.dtype, self.requires_grad = ptr, n, dtype, False

        def data_ptr(self): return self.ptr
        def numel(self): return self.n
        def element_size(self): return 4
        def is_contiguous(self): return True

    class FakeLib:
        def __init__(self): self.calls = []
        def fmpi_allreduce_symm(self, ref, off, count, dt, op, lp, st):
            self.calls.append(("allreduce", off, count, C.cast(lp, C.POINTER(_native.Launch)).contents.algo)); return 0
        def fmpi_bcast_symm(self, ref, off, nbytes, root, lp, st):
            self.calls.append(("bcast", off, nbytes, root)); return 0
        def fmpi_reduce_symm(self, ref, off, count, dt, op, root, lp, st):
            self.calls.append(("reduce", off, count, root)); return 0

    class FakeStream:
        cuda_stream = 0

    monkeypatch.setattr(torch.cuda, "current_stream", lambda *a, **k: FakeStream())
    dw = DW.DeviceWorld(0, 8, 0, "job")
    dw._lib = FakeLib()
    dw._lp_auto_obj = _native.Launch(0, 0, _native.ALGO_AUTO)
    dw._lp_auto = C.byref(dw._lp_auto_obj)
    dw.tuner = Tuner(8, True)

    def comm(mc):
        c = _native.DevComm()
        c.mc_data = mc
        return c

    with_mc = [comm(0x7000) for _ in range(_native.NUM_CHANNELS)]
    no_mc = [comm(None) for _ in range(_native.NUM_CHANNELS)]
    dw._symm_regs.append((0x10000, 0x10000 + (256 << 20), "seg", with_mc, [C.byref(c) for c in with_mc]))
    dw._symm_regs.append((0x9000_0000, 0x9000_0000 + (64 << 20), ("ipc", [], None), no_mc, [C.byref(c) for c in no_mc]))

    n = (64 << 20) // 4
    hit = dw._symm_hit(FakeTensor(0x10000 + 4096, n))
    assert hit is not None and hit[1] == 4096 and hit[2] is True
    dw._symm_collective("allreduce", FakeTensor(0x10000 + 4096, n), hit, _native.SUM, -1, 0, None)
    dw._symm_collective("allreduce", FakeTensor(0x10000 + 4096, n), hit, _native.MAX, -1, 0, None)
    hit2 = dw._symm_hit(FakeTensor(0x9000_0000, n))
    assert hit2 is not None and hit2[1] == 0 and hit2[2] is False
    dw._symm_collective("allreduce", FakeTensor(0x9000_0000, n), hit2, _native.SUM, -1, 0, None)
    dw._symm_collective("bcast", FakeTensor(0x9000_0000, n), hit2, 0, 3, 0, None)
    dw._symm_collective("reduce", FakeTensor(0x9000_0000, n), hit2, _native.SUM, 2, 0, None)
    calls = dw._lib.calls
    assert calls[0][:3] == ("allreduce", 4096, n)
    tuned_nvls = Tuner(8, True).choice("allreduce_symm", 64 << 20)
    if tuned_nvls is not None and tuned_nvls[0] == "nvls":
        assert calls[0][3] == _native.ALGO_NVLS                  # multicast + sum: in-switch
    assert calls[1][3] != _native.ALGO_NVLS                      # max is not a switch operation
    assert calls[2][3] != _native.ALGO_NVLS                      # no multicast alias on this buffer
    assert calls[3] == ("bcast", 0, 64 << 20, 3) and calls[4] == ("reduce", 0, n, 2)
    # a tensor that sticks out of the registered range, or is misaligned, is not a hit
    assert dw._symm_hit(FakeTensor(0x9000_0000 + (64 << 20) - 64, 1024)) is None
    assert dw._symm_hit(FakeTensor(0x10000 + 8, 1024)) is None
    assert dw.find_symmetric(FakeTensor(0x5000, 16)) is None


def test_launcher_helpers():
    from fluxmpi_b200 import launch as L
    assert not L._is_command(["train.py", "--x"], False) and not L._is_command(["pkg.mod"], True)
    assert L._is_command(["sh", "-c", "true"], False)            # an executable on PATH runs as it is
    assert not L._is_command(["definitely-not-a-command-xyz"], False)
    base = L._free_port_range(4)
    assert 1024 < base < 65535 - 4
    with pytest.raises(ValueError):
        L.launch(3, ["x.py"], sim_nodes=2)                        # 3 ranks do not split into 2 nodes
    with pytest.raises(ValueError):
        L.launch(2, ["x.py"], nnodes=2)                           # multi-node needs an explicit port
    # a command (not a python script) is started as it is, with the rank environment
    code, outs = L.launch(2, ["sh", "-c", "echo rank=$RANK/$WORLD_SIZE local=$LOCAL_RANK node=$GROUP_RANK"],
                          capture=True, timeout=60, sim_nodes=2)
    assert code == 0 and outs[0].strip() == "rank=0/2 local=0 node=0" and outs[1].strip() == "rank=1/2 local=0 node=1"


@pytest.mark.parametrize("rule", [R.Momentum(0.01, 0.9), R.Adam(1e-3), R.AdamW(1e-3, (0.9, 0.99), 1e-3), R.RMSProp(1e-3)],
                         ids=["momentum", "adam", "adamw", "rmsprop"])
def test_non_mutating_update_on_cpu_uses_copies_and_matches_per_leaf_ops(rule):
    """``update`` (Optimisers.update, the call in the reference's README) on CPU leaves: the
    native loop runs on copies; inputs stay untouched, ties stay tied, numbers equal the
    per-leaf path; a rule without a native loop simply takes the per-leaf path."""
    import copy
    torch.manual_seed(0)
    ps = {"a": torch.randn(65, 9), "b": {"w": torch.randn(100), "k": 3}}
    ps["b"]["tied"] = ps["b"]["w"]
    ps0 = copy.deepcopy(ps)
    st, st2 = R.setup(rule, ps), R.setup(rule, ps)
    st_first = st
    a, b = ps, ps
    for _ in range(4):
        gs = {"a": torch.randn(65, 9), "b": {"w": torch.randn(100), "k": None}}
        gs["b"]["tied"] = gs["b"]["w"]
        st, a = R.update(st, a, gs)
        st2, b = R._walk_update(st2, b, gs, False, {})
    assert torch.equal(ps["a"], ps0["a"]) and torch.equal(ps["b"]["w"], ps0["b"]["w"])      # inputs untouched
    assert T.tree_equal(st_first["a"].state, R.setup(rule, ps)["a"].state)                   # first state untouched
    assert a["b"]["tied"] is a["b"]["w"] and a["b"]["k"] == 3 and st["b"]["tied"] is st["b"]["w"]
    assert torch.allclose(a["a"], b["a"], rtol=1e-6, atol=1e-7) and torch.allclose(a["b"]["w"], b["b"]["w"], rtol=1e-6, atol=1e-7)
    assert T.tree_equal(st["a"].state, st2["a"].state) or torch.allclose(st["a"].state[0], st2["a"].state[0], rtol=1e-6, atol=1e-7)
