# A whole node disappears (its launcher cannot tell the other node's launcher): the survivors
# must get an error from the collective instead of waiting forever.
import os
import sys

import torch

import fluxmpi_b200 as fm

fm.Init()
S = fm.runtime.state()
fm.allreduce_(torch.ones(4), "sum")          # the mesh works
if S.node == S.nnodes - 1:
    sys.stdout.flush()
    os._exit(0)                              # gone without Finalize; sockets close with the process
try:
    for _ in range(3):                       # the first call may still find the peer's bytes in flight
        fm.allreduce_(torch.ones(1 << 16), "sum")
except Exception as e:                       # noqa: BLE001
    print(f"PEER_LOSS_DETECTED rank {fm.local_rank()}: {e}", flush=True)
    os._exit(0)
print("NOT_DETECTED", flush=True)
os._exit(1)
