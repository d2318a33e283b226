"""Self-test of the device world across the real GPUs of a job.

The reference's tests run their assertions inside every rank of a real MPI job
(/root/reference/test/runtests.jl:11-16). This is the same idea for the NVLink kernels, small
enough to run at the start of a benchmark or from ``python -m fluxmpi_b200.doctor``: every
algorithm family is driven through the live world (VMM / multicast mappings, signal pads) and
compared with a float64 host restatement, and replicas are compared bit for bit.

    import fluxmpi_b200 as fm
    fm.Init()
    report = fm.selftest.run()        # collective; {"ok": bool, "cases": n, "failed": [...], ...}
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _native, runtime
from . import rules as R
from .ops import fused as F


def _all_true(flag: bool) -> bool:
    from . import comm
    t = torch.tensor([1.0 if flag else 0.0], dtype=torch.float64)
    comm.allreduce_(t, "min")
    return bool(t.item() == 1.0)


def _host_sum(x: torch.Tensor, op: str = "sum") -> torch.Tensor:
    """float64 reference through the HOST world (an independent path: shared memory, no GPU)."""
    from . import comm
    h = x.detach().double().cpu().contiguous()
    comm.allreduce_(h, op)
    return h


def _same_bits(t: torch.Tensor) -> bool:
    """every rank holds the same bytes (checksum kernel + host min/max)"""
    from . import comm
    c = F.checksum(t).double().cpu().reshape(1)
    lo, hi = c.clone(), c.clone()
    comm.allreduce_(lo, "min")
    comm.allreduce_(hi, "max")
    return bool(lo.item() == hi.item())


def run(verbose: bool = False) -> dict:
    """Collective over all ranks. Never raises for a failed comparison: the report says which."""
    if not runtime.Initialized():
        raise runtime.FluxMPINotInitializedError()
    from . import comm
    S = runtime.state()
    dw = S.device_world
    r, w = S.rank, S.world
    rep = {"ok": True, "cases": 0, "failed": [], "world": w,
           "mode": getattr(dw, "mode", "none"), "multicast": bool(getattr(dw, "has_multicast", False))}
    if dw is None or w == 1 or not dw.native:
        rep["skipped"] = "no native device world (one worker, CPU, or NCCL fallback)"
        return rep
    dev = S.device
    g = torch.Generator().manual_seed(1234 + r)

    def case(name: str, ok: bool):
        ok = _all_true(bool(ok))
        rep["cases"] += 1
        if not ok:
            rep["ok"] = False
            rep["failed"].append(name)
        if verbose and r == 0:
            print(f"[selftest] {'ok  ' if ok else 'FAIL'} {name}", flush=True)

    def close(got: torch.Tensor, want: torch.Tensor, tol: float) -> bool:
        return bool(torch.allclose(got.double().cpu(), want, rtol=tol, atol=tol))

    lp = lambda algo, blocks=0: C.byref(_native.Launch(blocks, 0, algo))
    tols = {torch.float32: 1e-5, torch.bfloat16: 6e-2}

    # ---- all-reduce on ordinary tensors: one-shot push, staged two-shot, staged NVLS, pipeline
    algos = [("oneshot", _native.ALGO_ONESHOT, 4096), ("twoshot_stage", _native.ALGO_TWOSHOT, 1 << 20),
             ("pipe", _native.ALGO_PIPE, (24 << 20) + 16 * 5)]
    if dw.has_multicast:
        algos.insert(2, ("nvls_stage", _native.ALGO_NVLS, 1 << 20))
    for name, algo, n in algos:
        for dt in (torch.float32, torch.bfloat16):
            x = (torch.rand(n, generator=g) + 0.5).to(dt).to(dev)
            want = _host_sum(x)
            dw.allreduce(x, _native.SUM, launch=lp(algo))
            case(f"allreduce {name} {str(dt)[6:]} n={n}", close(x, want, tols[dt]) and _same_bits(x))
    # ---- in place on symmetric memory: two-shot P2P and in-switch
    for name, algo in (("twoshot_symm", _native.ALGO_TWOSHOT),) + ((("nvls_symm", _native.ALGO_NVLS),) if dw.has_multicast else ()):
        s = dw.symm_empty(((1 << 20) + 4,), torch.float32)
        s.copy_(torch.rand(s.numel(), generator=g) + 0.5)
        want = _host_sum(s)
        refs, off, _ = dw.find_symmetric(s)
        _native.cuda_check(dw._lib.fmpi_allreduce_symm(refs[0], off, s.numel(), _native.F32, _native.SUM, lp(algo),
                                                       torch.cuda.current_stream().cuda_stream), name)
        case(f"allreduce {name} f32", close(s, want, 1e-5) and _same_bits(s))
        # reduce to the last rank: only the root changes
        keep = (torch.rand(s.numel(), generator=g) + 0.5).to(dev)
        s.copy_(keep)
        want = _host_sum(s)
        _native.cuda_check(dw._lib.fmpi_reduce_symm(refs[0], off, s.numel(), _native.F32, _native.SUM, w - 1, lp(algo),
                                                    torch.cuda.current_stream().cuda_stream), name)
        case(f"reduce {name} f32", close(s, want, 1e-5) if r == w - 1 else torch.equal(s, keep))
    # ---- reduce on ordinary tensors (staged + pipelined), non-root untouched
    for name, algo, n in (("staged", _native.ALGO_AUTO, 1 << 18), ("pipe", _native.ALGO_PIPE, (12 << 20) + 16)):
        x = (torch.rand(n, generator=g) + 0.5).to(dev)
        keep, want = x.clone(), _host_sum(x)
        dw._reduce_native(x, x, _native.SUM, 0, 0, launch=lp(algo))
        case(f"reduce {name} f32 n={n}", close(x, want, 1e-5) if r == 0 else torch.equal(x, keep))
    # ---- broadcast: direct / multicast / scatter + all-gather / pipeline; symmetric: pull + re-emit
    for root in (0, w - 1):
        for name, algo, n in (("direct", _native.ALGO_ONESHOT, 5000), ("scatter", _native.ALGO_TWOSHOT, 3 << 20),
                              ("pipe", _native.ALGO_PIPE, (20 << 20) + 32)) + \
                ((("multicast", _native.ALGO_NVLS, 3 << 20),) if dw.has_multicast else ()):
            src = torch.arange(n, dtype=torch.float32) * (root + 1)
            x = src.to(dev) if r == root else torch.full((n,), -1.0, device=dev)
            dw._in1.value = x.data_ptr()
            _native.cuda_check(dw._lib.fmpi_bcast_any(dw._comm_refs[0], dw._in1_ref, n * 4, root, dw.bc_off, dw.bc_bytes,
                                                      lp(algo), torch.cuda.current_stream().cuda_stream), name)
            case(f"bcast {name} root={root} n={n}", torch.equal(x.cpu(), src))
        s = dw.symm_empty((3 << 20,), torch.float32)
        src = torch.arange(s.numel(), dtype=torch.float32) + root
        s.copy_(src if r == root else torch.zeros_like(src))
        refs, off, _ = dw.find_symmetric(s)
        _native.cuda_check(dw._lib.fmpi_bcast_symm(refs[0], off, s.numel() * 4, root, lp(_native.ALGO_PIPE),
                                                   torch.cuda.current_stream().cuda_stream), "bcast_symm pull")
        case(f"bcast symmetric pull+re-emit root={root}", torch.equal(s.cpu(), src))
    # ---- fused all-reduce (+) Adam through the engine: NVLS / two-shot / one-shot flavours against
    #      the fp32 rule fed the host-summed gradient; replicas bit-identical
    from .parallel.engine import DataParallelEngine
    flavours = [("twoshot", 64 * 4099), ("oneshot", 64 * 515)] + ([("nvls", 64 * 4099)] if dw.has_multicast else [])
    for algo, n in flavours:
        mod = torch.nn.Linear(n // 64, 64, bias=False).to(dev)
        with torch.no_grad():
            mod.weight.copy_(torch.randn(64, n // 64, generator=torch.Generator().manual_seed(7)))
        rule = R.Adam(1e-2)
        eng = DataParallelEngine(mod, rule, overlap=False, algo=algo, sync_params=False)
        p = mod.weight.detach().float().cpu().reshape(-1).clone()
        st = rule.init(p)
        ok = True
        for _ in range(3):
            gl = torch.randn(n, generator=g)
            eng.G[:n].copy_(gl)
            want_g = _host_sum(gl.to(dev)).float()
            eng.step()
            st, dx = rule.apply(st, p, want_g)
            p = p - dx
            ok = ok and close(eng.P[:n], p.double(), 2e-5)
        case(f"fused step {algo} f32 Adam", ok and _same_bits(eng.P))
        eng.close()
    torch.cuda.synchronize()
    try:
        dw.check_health()
    except _native.NativeError:
        rep["ok"] = False
        rep["failed"].append("device health page: a cross-GPU wait was abandoned")
    return rep


if __name__ == "__main__":
    import json
    runtime.Init()
    out = run(verbose=True)
    if runtime.local_rank() == 0:
        print(json.dumps(out))
    runtime.Finalize()
