# Init / ranks / printing (reference: src/common.jl).

struct FluxMPINotInitializedError <: Exception end
Base.showerror(io::IO, ::FluxMPINotInitializedError) =
    print(io, "Please call FluxMPIB200.Init(...) before using FluxMPI functionalities!")

"""    Initialized()   (reference: src/common.jl:6)"""
Initialized() = WORLD[] !== nothing

"""
    Init(; gpu_devices=nothing, verbose=false)

Join the job and bind a GPU. Idempotent. Default device is `LOCAL_RANK % ndev` (the
reference's `(rank+1) % ndev` puts rank 0 on GPU 1, src/common.jl:33-34); `gpu_devices[rank+1]`
overrides it.
"""
function Init(; gpu_devices::Union{Nothing, Vector{Int}}=nothing, verbose::Bool=false)
    if Initialized()
        verbose && fluxmpi_println("FluxMPI already initialized; Skipping...")
        return
    end
    w = _host_init()
    if verbose && w.size == 1
        @warn "Using FluxMPI with only 1 worker. It might be faster to run the code without MPI" maxlog=1
    end
    if CUDA.functional()
        dev = gpu_devices === nothing ? w.local_rank % length(CUDA.devices()) : gpu_devices[w.rank + 1]
        verbose && fluxmpi_println("Using GPU $dev")
        CUDA.device!(dev)
        if !@load_preference("FluxMPIDisableP2PSupport", false)
            w.device = _device_bringup(w, dev)
        end
    else
        verbose && fluxmpi_println("Using CPU")
    end
    return
end

"""    disable_p2p_support(; disable=true) — persisted; needs a restart (reference: src/FluxMPI.jl:51-56)"""
function disable_p2p_support(; disable=true)
    @set_preferences!("FluxMPIDisableP2PSupport" => disable)
    @info "NVLink P2P support " * (disable ? "disabled" : "enabled") * ". Restart Julia for this change to take effect!" maxlog=1
end
const disable_cudampi_support = disable_p2p_support

"""    local_rank()   (reference: src/common.jl:52-57)"""
@inline function local_rank()
    !Initialized() && throw(FluxMPINotInitializedError())
    return WORLD[].rank
end
CRC.@non_differentiable local_rank()

"""    total_workers()   (reference: src/common.jl:64-69)"""
@inline function total_workers()
    !Initialized() && throw(FluxMPINotInitializedError())
    return WORLD[].size
end
CRC.@non_differentiable total_workers()

for print_fn in (:println, :print)
    function_name = Symbol("fluxmpi_" * string(print_fn))
    @eval begin
        function $(function_name)(args...; kwargs...)
            if !Initialized()
                $(print_fn)("$(Dates.now()) ", args...; kwargs...)
                return
            end
            rank, size = local_rank(), total_workers()
            if size == 1
                $(print_fn)(args...; kwargs...)
                return
            end
            for r in 0:(size - 1)
                if r == rank
                    $(print_fn)("$(Dates.now()) [$(rank) / $(size)] ", args...; kwargs...)
                    flush(stdout)
                end
                host_barrier()          # host-side only: never touches a GPU stream
            end
            return
        end
        CRC.@non_differentiable $(function_name)(::Any...)
    end
end
