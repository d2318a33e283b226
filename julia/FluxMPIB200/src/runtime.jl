# Process runtime for the Julia side: Init, rank queries, rank-ordered printing.
# Behavioural contract: /root/reference/src/common.jl (Init :16-45, rank/size :52-69,
# printing :72-98). Implementation is independent: ranks are cached in the World object,
# printing serialises over the shm host barrier, and the device world is brought up here.

struct FluxMPINotInitializedError <: Exception end
Base.showerror(io::IO, ::FluxMPINotInitializedError) =
    print(io, "Please call FluxMPIB200.Init(...) before using FluxMPI functionalities!")

Initialized() = WORLD[] !== nothing

_require_world() = (w = WORLD[]; w === nothing ? throw(FluxMPINotInitializedError()) : w)

local_rank() = _require_world().rank
total_workers() = _require_world().size
CRC.@non_differentiable local_rank()
CRC.@non_differentiable total_workers()

function _pick_device(w::World, gpu_devices)
    gpu_devices === nothing || return gpu_devices[w.rank + 1]
    return w.local_rank % length(CUDA.devices())     # rank r -> GPU r (SURVEY §10)
end

"""
    Init(; gpu_devices=nothing, verbose=false)

Join the job (environment `RANK` / `WORLD_SIZE` / `LOCAL_RANK`), bind a GPU and map the peers'
symmetric memory. Calling it again is a no-op.
"""
function Init(; gpu_devices::Union{Nothing, Vector{Int}}=nothing, verbose::Bool=false)
    if Initialized()
        verbose && fluxmpi_println("FluxMPI already initialized; Skipping...")
        return nothing
    end
    w = _host_init()
    verbose && w.size == 1 &&
        @warn "Using FluxMPI with only 1 worker. It might be faster to run the code without MPI" maxlog=1
    if !CUDA.functional()
        verbose && fluxmpi_println("Using CPU")
        return nothing
    end
    dev = _pick_device(w, gpu_devices)
    verbose && fluxmpi_println("Using GPU $dev")
    CUDA.device!(dev)
    # peer mappings exist inside one NVSwitch domain; over several nodes CuArrays are staged
    # through the host world (device.jl), like the reference without CUDA-aware MPI
    one_node = parse(Int, get(ENV, "LOCAL_WORLD_SIZE", string(w.size))) == w.size
    if one_node && !@load_preference("FluxMPIDisableP2PSupport", false)
        w.device = _device_bringup(w, dev)
    end
    return nothing
end

"""
    disable_p2p_support(; disable=true)

Persist the "no NVLink peer-to-peer kernels" preference (takes effect after a restart).
`disable_cudampi_support` is kept as an alias of the reference's name.
"""
function disable_p2p_support(; disable=true)
    @set_preferences!("FluxMPIDisableP2PSupport" => disable)
    state = disable ? "disabled" : "enabled"
    @info "NVLink P2P support $state. Restart Julia for this change to take effect!" maxlog=1
end
const disable_cudampi_support = disable_p2p_support

# One emitter for both print flavours: before Init a timestamp only, at one worker the bare
# text, otherwise "<time> [rank / size] text" with the ranks taking turns around a host barrier.
function _ordered_emit(emit::Function, args, kwargs)
    w = WORLD[]
    if w === nothing
        emit(string(Dates.now(), " "), args...; kwargs...)
    elseif w.size == 1
        emit(args...; kwargs...)
    else
        for turn in 0:(w.size - 1)
            if turn == w.rank
                emit(string(Dates.now(), " [", w.rank, " / ", w.size, "] "), args...; kwargs...)
                flush(stdout)
            end
            host_barrier()
        end
    end
    return nothing
end

"""    fluxmpi_println(args...; kwargs...) — `println` with rank and size, in rank order"""
fluxmpi_println(args...; kwargs...) = _ordered_emit(println, args, kwargs)
"""    fluxmpi_print(args...; kwargs...) — `print` with rank and size, in rank order"""
fluxmpi_print(args...; kwargs...) = _ordered_emit(print, args, kwargs)
CRC.@non_differentiable fluxmpi_println(::Any...)
CRC.@non_differentiable fluxmpi_print(::Any...)
