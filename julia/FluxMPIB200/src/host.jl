# Host world (POSIX shm): what MPI.Init / MPI.Allreduce! / MPI.Bcast! / MPI.Reduce! / MPI.Barrier
# are to the reference (src/common.jl:22,91; src/mpi_extensions.jl:108-111,130-133,152-155).

mutable struct World
    rank::Int
    size::Int
    local_rank::Int
    host::Ptr{Cvoid}          # fmpi_host_comm*
    device::Any               # DeviceWorld or nothing
    job::String
end

const WORLD = Ref{Union{Nothing, World}}(nothing)

function _job_name()
    haskey(ENV, "FMPI_JOB_ID") && return ENV["FMPI_JOB_ID"]
    return string(get(ENV, "MASTER_PORT", "0"), "-", get(ENV, "TORCHELASTIC_RUN_ID", "none"), "-jl")
end

function _host_init()
    rank = parse(Int, get(ENV, "RANK", "0"))
    size = parse(Int, get(ENV, "WORLD_SIZE", "1"))
    lrank = parse(Int, get(ENV, "LOCAL_RANK", string(rank)))
    job = _job_name()
    h = Ref{Ptr{Cvoid}}(C_NULL)
    lsize = parse(Int, get(ENV, "LOCAL_WORLD_SIZE", string(size)))
    nnodes = size ÷ lsize
    if size > 1
        name = "/fmpi_" * replace(job, r"[^A-Za-z0-9]" => "_")
        timeout = parse(Float64, get(ENV, "FMPI_TIMEOUT_S", "300"))
        if nnodes == 1
            hcheck(ccall(hsym(:fmpi_host_init), Cint, (Cstring, Cint, Cint, Csize_t, Cdouble, Ptr{Ptr{Cvoid}}),
                         name, rank, size, 8 << 20, timeout, h), "fmpi_host_init")
        else
            # several nodes: shared memory inside each, a TCP mesh between the node leaders
            node = parse(Int, get(ENV, "GROUP_RANK", string(rank ÷ lsize)))
            port = parse(Int, get(ENV, "FMPI_NET_PORT", string(parse(Int, get(ENV, "MASTER_PORT", "29500")) + 1)))
            hcheck(ccall(hsym(:fmpi_host_init_multi), Cint,
                         (Cstring, Cint, Cint, Cint, Cint, Cstring, Cint, Csize_t, Cdouble, Ptr{Ptr{Cvoid}}),
                         string(name, "_n", node), lrank, lsize, node, nnodes,
                         get(ENV, "MASTER_ADDR", "127.0.0.1"), port, 8 << 20, timeout, h), "fmpi_host_init_multi")
        end
    end
    WORLD[] = World(rank, size, lrank, h[], nothing, job)
    atexit(_finalize)
    return WORLD[]
end

function _finalize()
    w = WORLD[]
    w === nothing && return
    w.device === nothing || _device_close(w.device)
    w.host == C_NULL || ccall(hsym(:fmpi_host_finalize), Cint, (Ptr{Cvoid},), w.host)
    WORLD[] = nothing
    return
end

function host_barrier()
    w = WORLD[]
    (w === nothing || w.size == 1) && return
    hcheck(ccall(hsym(:fmpi_host_barrier), Cint, (Ptr{Cvoid},), w.host), "barrier")
end

function host_allgather(blob::Vector{UInt8})
    w = WORLD[]
    w.size == 1 && return [blob]
    out = Vector{UInt8}(undef, length(blob) * w.size)
    hcheck(ccall(hsym(:fmpi_host_allgather), Cint, (Ptr{Cvoid}, Ptr{UInt8}, Ptr{UInt8}, Csize_t),
                 w.host, blob, out, length(blob)), "allgather")
    return [out[(i * length(blob) + 1):((i + 1) * length(blob))] for i in 0:(w.size - 1)]
end

const NativeOp = Union{typeof(+), typeof(*), typeof(min), typeof(max)}

# Any other function (the reference wraps it in MPI.Op, src/mpi_extensions.jl:52-54): gather
# every rank's array in slot-sized pieces and fold them in rank order — the same order on every
# rank, so the result is identical everywhere even for non-commutative functions.
function _host_fold!(v::Array{T}, f, root::Integer=-1) where {T}
    w = WORLD[]
    w.size == 1 && return v
    isbitstype(T) || throw(FluxMPIError("custom reductions need a bits eltype, got $T"))
    mine = collect(reinterpret(UInt8, vec(v)))
    n = length(mine)
    parts = [Vector{UInt8}(undef, n) for _ in 1:w.size]
    chunk = max(1, (8 << 20) ÷ w.size)
    for lo in 1:chunk:n
        hi = min(lo + chunk - 1, n)
        got = host_allgather(mine[lo:hi])
        for q in 1:w.size
            parts[q][lo:hi] = got[q]
        end
    end
    (root >= 0 && w.rank != root) && return v
    vals = [reshape(collect(reinterpret(T, p)), size(v)) for p in parts]
    copyto!(v, foldl((a, b) -> f.(a, b), vals))
    return v
end

_host_allreduce!(v::Array, f) = _host_fold!(v, f)

function _host_allreduce!(v::Array{T}, op::NativeOp) where {T}
    w = WORLD[]
    w.size == 1 && return v
    hcheck(ccall(hsym(:fmpi_host_allreduce), Cint, (Ptr{Cvoid}, Ptr{T}, Ptr{T}, Csize_t, Cint, Cint),
                 w.host, v, v, length(v), dtype_code(T), op_code(op)), "allreduce!")
    return v
end

# A whole gradient tree as one shared-memory message: pieces go from the leaves straight into
# the staging slot and back (no flattening copy, the barriers of a single message).
function _host_allreduce_many!(leaves::Vector{<:Array{T}}, op) where {T}
    w = WORLD[]
    (w.size == 1 || isempty(leaves)) && return leaves
    ptrs = Ptr{Cvoid}[Base.unsafe_convert(Ptr{Cvoid}, pointer(l)) for l in leaves]
    counts = UInt64[length(l) for l in leaves]
    GC.@preserve leaves begin
        hcheck(ccall(hsym(:fmpi_host_allreduce_many), Cint,
                     (Ptr{Cvoid}, Cint, Ptr{Ptr{Cvoid}}, Ptr{UInt64}, Cint, Cint),
                     w.host, length(leaves), ptrs, counts, dtype_code(T), op_code(op)), "allreduce_many")
    end
    return leaves
end

function _host_bcast!(v::Array{T}, root::Integer) where {T}
    w = WORLD[]
    w.size == 1 && return v
    hcheck(ccall(hsym(:fmpi_host_bcast), Cint, (Ptr{Cvoid}, Ptr{T}, Csize_t, Cint),
                 w.host, v, sizeof(v), root), "bcast!")
    return v
end

_host_reduce!(v::Array, f, root::Integer) = _host_fold!(v, f, root)

function _host_reduce!(v::Array{T}, op::NativeOp, root::Integer) where {T}
    w = WORLD[]
    w.size == 1 && return v
    hcheck(ccall(hsym(:fmpi_host_reduce), Cint, (Ptr{Cvoid}, Ptr{T}, Ptr{T}, Csize_t, Cint, Cint, Cint),
                 w.host, v, v, length(v), dtype_code(T), op_code(op), root), "reduce!")
    return v
end
