import sys
import torch
import fluxmpi_b200 as fm

fm.Init()
if fm.local_rank() == 1:
    sys.stdout.flush()
    import os
    os._exit(3)
fm.allreduce_(torch.ones(4), "sum")   # would hang forever without failure handling
