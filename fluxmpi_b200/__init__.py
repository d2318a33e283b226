"""fluxmpi_b200 — B200-native data-parallel training with FluxMPI.jl's capabilities.

Public API (reference export list /root/reference/src/FluxMPI.jl:88-96 and the documented
unexported functions docs/src/api.md:17-58):

=================================  ==========================================================
reference                          here
=================================  ==========================================================
FluxMPI.Init / Initialized         Init / Initialized
local_rank / total_workers         local_rank / total_workers
fluxmpi_print / fluxmpi_println    fluxmpi_print / fluxmpi_println
FluxMPI.synchronize!               synchronize (alias synchronize_)
DistributedOptimizer               DistributedOptimizer
allreduce_gradients                allreduce_gradients
DistributedDataContainer           DistributedDataContainer
FluxMPI.allreduce!/bcast!/reduce!  allreduce_ / bcast_ / reduce_
FluxMPI.Iallreduce!/Ibcast!        Iallreduce / Ibcast (+ Request.wait, Waitall)
FluxMPI.disable_cudampi_support    disable_p2p_support (old name kept as alias)
FluxMPIFluxModel                   FluxMPIFluxModel
ComponentArray extension           FlatParams
=================================  ==========================================================

Beyond the reference: ``DataParallelEngine`` (parallel/engine.py) — gradients live in a
persistent symmetric fusion buffer and one hand-written sm_100a kernel reduces them over
NVLink and applies the optimiser step in registers, bucket by bucket, overlapped with backward.
"""
from .runtime import (Init, Initialized, Finalize, local_rank, total_workers, node_local_rank,
                      current_device, topology, fluxmpi_print, fluxmpi_println,
                      disable_p2p_support, disable_cudampi_support, p2p_disabled,
                      FluxMPINotInitializedError, host_barrier)
from .comm import (allreduce_, bcast_, reduce_, Iallreduce, Ibcast, Request, Waitall, barrier,
                   cpu, gpu, COMM_WORLD, symm_empty, symm_zeros, is_symmetric,
                   shm_empty, shm_free)
from .group import Group
from .tree import fmap, functor
from . import rules
from .rules import (AbstractRule, Leaf, setup, update, update_, adjust, Descent, Momentum,
                    Nesterov, RMSProp, Adam, AdamW, AdaMax, AdaGrad, AdaDelta, AMSGrad, NAdam,
                    AdaBelief, WeightDecay, ClipGrad, ClipNorm, OptimiserChain)
from .parallel.flat import FlatParams
from .parallel.sync import synchronize, synchronize_, FluxMPIFluxModel
from .parallel.optimizer import DistributedOptimizer, allreduce_gradients
from .data import DistributedDataContainer, DeviceLoader, TensorDataset

from .utils import checkpoint, trace

__version__ = "0.1.0"


def __getattr__(name):
    # heavy / CUDA-touching pieces are imported on demand
    if name == "DataParallelEngine":
        from .parallel.engine import DataParallelEngine
        return DataParallelEngine
    if name == "selftest":
        import importlib
        return importlib.import_module(".selftest", __name__)
    if name in ("HostDataParallelEngine", "make_engine"):
        from .parallel import host_engine
        return getattr(host_engine, name)
    raise AttributeError(name)
