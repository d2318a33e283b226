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
function allreduce_gradients(gs::Union{NamedTuple, Tuple}; on_gpu::Bool=CUDA.functional())
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
