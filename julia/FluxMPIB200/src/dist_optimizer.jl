# DistributedOptimizer and allreduce_gradients.
# Contract: /root/reference/src/optimizer.jl:16-25 (wrap any rule, sum gradients, delegate state)
# and :45-65 (all leaves in flight, then wait).

"""
    DistributedOptimizer(rule)

An `Optimisers.AbstractRule` that sums the gradient over all workers and then applies `rule`.
SUM, not mean: divide the loss by `total_workers()` for an average. Its state is `rule`'s state.
"""
struct DistributedOptimizer{R} <: Optimisers.AbstractRule
    optimizer::R
end

Optimisers.init(d::DistributedOptimizer, x::AbstractArray) = Optimisers.init(d.optimizer, x)

function Optimisers.apply!(d::DistributedOptimizer, state, x, dx)
    summed = allreduce!(dx, +)
    return Optimisers.apply!(d.optimizer, state, x, summed)
end

"""
    allreduce_gradients(gs; on_gpu=CUDA.functional())

Sum every numeric array leaf of the tree `gs` over the workers and return the tree. All
reductions are posted before any is waited for. `on_gpu` is accepted for compatibility and
ignored (device arrays are reduced in device memory). A bare array — e.g. a ComponentArray's
flat vector — is accepted too.
"""
# Device trees in two launches (csrc/kernels/tree_step.cu): the leaves are gathered through a
# pointer table into a symmetric arena and all-reduced there in place by ONE kernel; a second one
# scatters the sums back into the leaves — wherever Zygote put them. Falls back to the per-leaf
# path when the leaves are not all CuArrays of one 32- or 16-bit float type.
const _TreeFloat = Union{Float32, Float16}
const _tree_arena = Ref{Any}(nothing)          # (segment handle, DevComm, bytes)

function _arena(w::World, nbytes::Integer)
    a = _tree_arena[]
    if a === nothing || a[3] < nbytes
        dev = CUDA.deviceid(CUDA.device())
        size = max(cld(nbytes, 1 << 21) << 21, 64 << 20)
        no = _segment_counter[]
        _segment_counter[] += 1
        seg, mcptr = _alloc_segment(w, dev, size, no)
        base = (w.device::DeviceWorld).comms[1]
        peers = [ccall(csym(:fmpi_symm_peer_ptr), Ptr{Cvoid}, (Ptr{Cvoid}, Cint), seg, q) for q in 0:(w.size - 1)]
        pad(v) = ntuple(i -> i <= length(v) ? v[i] : C_NULL, FMPI_MAX_RANKS)
        comm = DevComm(w.rank, w.size, 0, 0, pad(peers), base.sig, mcptr, UInt64(size), base.ctrl, base.timeout_ns)
        a = (seg, comm, size)
        _tree_arena[] = a
    end
    return a
end

function _tree_allreduce!(leaves::Vector{<:CuArray{T}}) where {T <: _TreeFloat}
    w = WORLD[]
    n = length(leaves)
    offs = Vector{UInt64}(undef, n)
    total = UInt64(0)
    for (i, l) in enumerate(leaves)
        offs[i] = total
        total += UInt64(cld(sizeof(l), 16) * 16)
    end
    seg, comm, _ = _arena(w, total)
    ptrs = UInt64[UInt64(UInt(pointer(l))) for l in leaves]
    zero64 = zeros(UInt64, n)
    table = CuVector{UInt64}(vcat(ptrs, ptrs, zero64, zero64, UInt64[length(l) for l in leaves], offs))
    tabref = Ref{Ptr{Cvoid}}(Base.unsafe_convert(Ptr{Cvoid}, pointer(table)))
    ccheck(ccall(csym(:fmpi_tree_allreduce), Cint,
                 (Ptr{DevComm}, Cint, Ptr{Ptr{Cvoid}}, UInt64, UInt64, Cint, Cint, Ptr{Launch}, Ptr{Cvoid}),
                 Ref(comm), n, tabref, 0, total, dtype_code(T), op_code(+), Ref(AUTO), _stream()), "tree_allreduce")
    unp = CuVector{UInt64}(vcat(ptrs, offs, UInt64[sizeof(l) for l in leaves]))
    base = Base.unsafe_convert(Ptr{Cvoid}, pointer(unp))
    arena = ccall(csym(:fmpi_symm_local_ptr), Ptr{Cvoid}, (Ptr{Cvoid},), seg)
    ccheck(ccall(csym(:fmpi_unpack), Cint, (Cint, Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}, Ptr{Cvoid}),
                 n, base, base + 8n, base + 16n, arena, _stream()), "unpack")
    CUDA.synchronize()        # `table` / `unp` must outlive the kernels that read them
    return nothing
end

function allreduce_gradients(gs::Union{NamedTuple, Tuple}; on_gpu::Bool=CUDA.functional())
    w = WORLD[]
    if w !== nothing && w.size > 1 && w.device !== nothing
        dev_leaves = CuArray[]
        uniform = Ref(true)
        fmap(gs) do leaf
            if leaf isa CuArray{<:_TreeFloat}
                push!(dev_leaves, leaf)
            elseif leaf isa AbstractArray{<:Number}
                uniform[] = false
            end
            leaf
        end
        if uniform[] && !isempty(dev_leaves) && all(l -> eltype(l) === eltype(dev_leaves[1]), dev_leaves)
            _tree_allreduce!(Vector{CuArray{eltype(dev_leaves[1])}}(dev_leaves))
            return gs
        end
    end
    pending = Request[]
    host = Dict{DataType, Vector{Any}}()      # host leaves by eltype: one message per eltype
    post(leaf) = leaf
    function post(leaf::Array{T}) where {T <: Union{Float32, Float64, Float16, Int32, Int64, Int16, Int8, UInt8}}
        push!(get!(host, T, Any[]), leaf)
        return leaf
    end
    function post(leaf::AbstractArray{<:Number})
        out, req = Iallreduce!(leaf, +)
        push!(pending, req)
        return out
    end
    reduced = fmap(post, gs)
    for (T, leaves) in host
        _host_allreduce_many!(Vector{Array{T}}(leaves), +)
    end
    waitall(pending)
    return reduced
end

allreduce_gradients(g::AbstractArray{<:Number}; on_gpu::Bool=CUDA.functional()) = allreduce!(g, +)
