# SPMD rank script — mirrors /root/reference/test/test_common.jl:1-17 (+ checklist 2-4).
import io, re, warnings
import fluxmpi_b200 as fm

# before Init: rank queries raise, printing shows a timestamp only (src/common.jl:53,65,76-79)
for fn in (fm.local_rank, fm.total_workers):
    try:
        fn()
    except fm.FluxMPINotInitializedError as e:
        assert "Init" in str(e)
    else:
        raise AssertionError("expected FluxMPINotInitializedError")
buf = io.StringIO()
fm.fluxmpi_println("pre-init", file=buf)
assert re.match(r"^\d{4}-\d\d-\d\dT[\d:.]+ pre-init\n$", buf.getvalue()), buf.getvalue()

fm.Init(verbose=True)
assert fm.Initialized()
fm.Init(verbose=False)          # idempotent (src/common.jl:17-20)
assert fm.local_rank() < fm.total_workers()

with warnings.catch_warnings():
    warnings.simplefilter("error")
    fm.fluxmpi_println("Hello from rank ", fm.local_rank())
    fm.fluxmpi_print("no newline")
    fm.fluxmpi_println()

# the prefix format "[rank / size]" (src/common.jl:88)
buf = io.StringIO()
fm.fluxmpi_println("x", file=buf)
r, w = fm.local_rank(), fm.total_workers()
if w > 1:
    assert re.search(rf" \[{r} / {w}\] x\n$", buf.getvalue()), buf.getvalue()
else:
    assert buf.getvalue() == "x\n"

# ranks are plain ints => constants under autograd (CRC.@non_differentiable, src/common.jl:57,69)
import torch
p = torch.ones(3, requires_grad=True)
loss = (p * (fm.local_rank() + 1)).sum() / fm.total_workers()
loss.backward()
assert torch.allclose(p.grad, torch.full((3,), (r + 1) / w))

fm.barrier()
fm.Finalize()
assert not fm.Initialized()
print("RANK_OK", r)
