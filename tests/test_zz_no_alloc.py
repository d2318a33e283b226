"""fmpi_sumsq_noalloc through the C ABI (collected last on purpose: a `-x` run has already
executed every other test by the time this one can fail).

The entry point exists for callers that cannot let a library allocate device memory — a host
runtime that owns the memory pool (the Julia binding under CUDA.jl), a hook firing inside a
step whose memory is planned, a monitoring thread. It keeps its scratch in a small array inside
the library's device image, so the only buffers involved are the ones the caller hands in. The
test drives it exactly like such a caller would: raw pointers through ctypes, nothing from the
Python wrappers. It must agree with fmpi_sumsq (caller-provided scratch) to the tolerance the
other sumsq tests use, on two calls in a row — the second call sees whatever the first left in
the shared scratch — and it must refuse inputs beyond its documented limit.

Works in a fresh process (no Init, no other kernels run first) and after the whole suite.
"""

import pytest
import torch

pytestmark = pytest.mark.gpu

if not torch.cuda.is_available():
    pytest.skip("needs a CUDA device", allow_module_level=True)

from fluxmpi_b200 import _native

F32, F64, F16, BF16 = _native.F32, _native.F64, _native.F16, _native.BF16
CODES = {torch.float32: F32, torch.float64: F64, torch.float16: F16, torch.bfloat16: BF16}
MAX_COUNT = 1 << 28      # FMPI_NOALLOC_MAX_COUNT in include/fmpi.h


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _sumsq_with_caller_scratch(lib, x):
    scratch = torch.zeros(lib.fmpi_reduce_scratch_bytes(), dtype=torch.uint8, device=x.device)
    out = torch.zeros((), dtype=torch.float64, device=x.device)
    rc = lib.fmpi_sumsq(x.data_ptr(), x.numel(), CODES[x.dtype], out.data_ptr(), scratch.data_ptr(), _stream())
    assert rc == 0, lib.fmpi_cuda_last_error().decode()
    return out


def _sumsq_noalloc(lib, x):
    out = torch.full((), -1.0, dtype=torch.float64, device=x.device)
    rc = lib.fmpi_sumsq_noalloc(x.data_ptr(), x.numel(), CODES[x.dtype], out.data_ptr(), _stream())
    assert rc == 0, lib.fmpi_cuda_last_error().decode()
    return out


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16, torch.float64])
def test_noalloc_matches_existing_entry_point_twice_in_a_row(dtype):
    lib = _native.cuda()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    g = torch.Generator(device="cpu").manual_seed(1234)
    # different sizes back to back: the second call runs on the scratch contents (partials,
    # ticket) that the first one left behind, with a different number of blocks
    first = torch.randn((1 << 21) + 7, generator=g).to(dtype).to(dev)
    second = torch.randn(4099, generator=g).to(dtype).to(dev)
    for x in (first, second):
        want = _sumsq_with_caller_scratch(lib, x)
        got = _sumsq_noalloc(lib, x)
        torch.cuda.synchronize()
        assert torch.allclose(got, want, rtol=1e-9), (dtype, x.numel(), float(got), float(want))
        assert torch.allclose(got, (x.double() ** 2).sum(), rtol=1e-9)
    # same input twice: identical bits (deterministic fold), and no drift from the shared scratch
    a = _sumsq_noalloc(lib, first)
    b = _sumsq_noalloc(lib, first)
    torch.cuda.synchronize()
    assert torch.equal(a, b)


def test_noalloc_allocates_nothing_and_rejects_oversize():
    lib = _native.cuda()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    x = torch.randn(1 << 16, device=dev)
    _sumsq_noalloc(lib, x)                        # first call on this device does the set-up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info(dev)
    out = torch.zeros((), dtype=torch.float64, device=dev)
    free_before, _ = torch.cuda.mem_get_info(dev)
    for _ in range(50):
        rc = lib.fmpi_sumsq_noalloc(x.data_ptr(), x.numel(), F32, out.data_ptr(), _stream())
        assert rc == 0
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info(dev)
    assert free_after == free_before, "the no-alloc entry point changed the amount of free device memory"
    # beyond the documented maximum: refused before anything is launched, with a clear message
    out.fill_(-3.0)
    rc = lib.fmpi_sumsq_noalloc(x.data_ptr(), MAX_COUNT + 1, F32, out.data_ptr(), _stream())
    assert rc != 0
    msg = lib.fmpi_cuda_last_error().decode()
    assert "exceed" in msg and str(MAX_COUNT) in msg, msg
    torch.cuda.synchronize()
    assert float(out) == -3.0, "a refused call must not touch the output"
    # and the entry point still works afterwards
    got = _sumsq_noalloc(lib, x)
    torch.cuda.synchronize()
    assert torch.allclose(got, (x.double() ** 2).sum(), rtol=1e-9)
