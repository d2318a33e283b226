"""
    FluxMPIB200

Thin Julia binding over the C ABI of `include/fmpi.h` (the same two shared libraries the Python
package `fluxmpi_b200` loads). It keeps FluxMPI.jl's names and calling conventions —
`Init`, `local_rank`, `total_workers`, `synchronize!`, `DistributedOptimizer`,
`allreduce_gradients`, `DistributedDataContainer`, `fluxmpi_print(ln)`, `allreduce!`, `bcast!`,
`reduce!`, `Iallreduce!`, `Ibcast!` — so an existing Lux/Optimisers/Zygote script switches by
replacing `using FluxMPI` with `using FluxMPIB200`. No MPI is involved: `Array`s go to the
POSIX-shm host backend, `CuArray`s are handed by pointer to the sm_100a NVLink kernels.

Launch with `python -m fluxmpi_b200.launch -n N julia --project script.jl` (or anything that
sets `RANK`, `WORLD_SIZE`, `LOCAL_RANK`). Libraries are found through `ENV["FLUXMPI_B200_LIB"]`
(directory holding `libfmpi_host.so` / `libfmpi_cuda.so`) or `../../fluxmpi_b200/_lib`.

This image has no Julia, so the binding is verified only through the ABI it calls:
`csrc/abi_test.c` drives the same host entry points from C and the Python tests drive the
device ones.
"""
module FluxMPIB200

using CUDA, Dates, Functors, Libdl, Optimisers, Preferences, Setfield
import ChainRulesCore as CRC

export local_rank, total_workers, DistributedOptimizer, fluxmpi_print, fluxmpi_println,
       DistributedDataContainer, allreduce_gradients, FluxMPIFluxModel

include("abi.jl")        # library handles, enums, struct mirrors, error helpers
include("host.jl")       # host world: init, barrier, collectives on Arrays
include("device.jl")     # device world: symmetric segment bring-up, collectives on CuArrays
include("runtime.jl")    # Init / ranks / printing
include("sync_tree.jl")
include("dist_optimizer.jl")
include("sharding.jl")
include("engine.jl")     # FlatEngine: fused all-reduce + optimiser step on a flat vector

export FlatEngine, step!, params, grads

"""
    FluxMPIFluxModel(model)

Marker wrapper so `synchronize!` can dispatch on an arbitrary Flux model
(reference: src/FluxMPI.jl:84-86); the method lives in `ext/FluxMPIB200FluxExt.jl`.
"""
struct FluxMPIFluxModel{M}
    model::M
end

end
