# User-buffer registration (fm.register_buffer): an ordinary torch tensor mapped into the peers
# through CUDA IPC, then reduced / broadcast in place by the peer-to-peer kernels.
# NOT YET RUN when written (no GPU time left in round 1); launched by tests/test_gpu_multi.py
# only with FMPI_TEST_UNVERIFIED=1.
import torch

import fluxmpi_b200 as fm

fm.Init(verbose=True)
r, w = fm.local_rank(), fm.total_workers()
dev = fm.current_device()
dw = fm.runtime.state().device_world
assert dw.native and w > 1

n = (8 << 20) // 4 + 4                                       # 8 MB and a bit, fp32
pad = torch.empty(r * 1000 + 16, device=dev)                 # different offsets inside the blocks
t = torch.full((n,), float(r + 1), device=dev)
assert not fm.is_symmetric(t)
reg = fm.register_buffer(t)
assert reg is not None, "CUDA-IPC registration failed"
assert fm.is_symmetric(t)

tri = w * (w + 1) / 2
fm.allreduce_(t, "sum")
assert torch.equal(t, torch.full((n,), tri, device=dev)), float(t[0])
fm.allreduce_(t, "max")                                      # no multicast on this buffer: P2P kernel
assert torch.equal(t, torch.full((n,), tri, device=dev))
t.fill_(float(r))
fm.bcast_(t, w - 1)
assert torch.equal(t, torch.full((n,), float(w - 1), device=dev))
t.fill_(1.0)
fm.reduce_(t, "sum", 0)
assert torch.equal(t, torch.full((n,), float(w) if r == 0 else 1.0, device=dev))
# a slice at the same relative offset on every rank is still in place
t.fill_(float(r + 1))
s = t[1024:1024 + (1 << 20)]
assert fm.is_symmetric(s)
fm.allreduce_(s, "sum")
assert torch.equal(s, torch.full_like(s, tri)) and float(t[0]) == r + 1 and float(t[-1]) == r + 1
# small messages take the same path (one-shot inside the kernel's own choice)
u = t[:4]
fm.allreduce_(u, "sum")
torch.cuda.synchronize()

fm.deregister_buffer(reg)
assert not fm.is_symmetric(t)
t.fill_(float(r + 1))
fm.allreduce_(t, "sum")                                      # staged path again, same answer
assert torch.equal(t, torch.full((n,), tri, device=dev))
del pad
fm.fluxmpi_println("register_buffer ok")
print(f"RANK_OK {r}", flush=True)
fm.Finalize()
