# A rank that leaves without a word must end in an error on the survivors, not in GPUs that spin
# inside a kernel for ever: the cross-GPU waits are bounded (fmpi_devcomm::timeout_ns) and the
# host finds the error word at its next health check. The reference has no counterpart: a lost
# rank leaves the others inside libmpi.
import os
import sys
import time

import torch
import fluxmpi_b200 as fm
from fluxmpi_b200._native import NativeError

fm.Init()
r, w = fm.local_rank(), fm.total_workers()
dev = fm.current_device()
dw = fm.runtime.state().device_world
assert dw.native, dw.why_fallback
x = torch.ones(1 << 16, device=dev)
fm.allreduce_(x, "sum")
torch.cuda.synchronize()
assert float(x[0]) == w
fm.barrier()
if r == w - 1:
    sys.stdout.flush()
    os._exit(0)                       # exit code 0: the launcher does not stop the others
t0 = time.time()
try:
    fm.allreduce_(x, "sum")           # the kernel waits for the lost rank's flag ...
    torch.cuda.synchronize()          # ... and returns once its budget is spent
    dw.check_health()
except NativeError as e:
    print(f"SURVIVOR_ERROR rank {r} after {time.time() - t0:.1f} s: {e}", flush=True)
    os._exit(7)
print("no error was raised", flush=True)
os._exit(0)
