# DistributedOptimizer / allreduce_gradients (reference: src/optimizer.jl).

"""
    DistributedOptimizer(optimizer)

Wrap any Optimisers.jl rule; gradients are SUMMED over the workers before the wrapped rule
runs (scale the loss by `1 / total_workers()` for an average — reference: src/optimizer.jl:11-14).
The state tree is the wrapped rule's (`Optimisers.init` delegates).
"""
struct DistributedOptimizer{O} <: Optimisers.AbstractRule
    optimizer::O
end

function Optimisers.apply!(o::DistributedOptimizer, state, x, y)
    y_ = allreduce!(y, +)
    return Optimisers.apply!(o.optimizer, state, x, y_)
end

Optimisers.init(o::DistributedOptimizer, x::AbstractArray) = Optimisers.init(o.optimizer, x)

"""
    allreduce_gradients(gs; on_gpu=CUDA.functional())

Sum every array leaf of `gs` over the workers, all in flight at once, and return the tree.
`on_gpu` is accepted for compatibility and ignored: device arrays are reduced where they are
(the reference copies the whole tree to the host first, src/optimizer.jl:46-47). Also accepts a
flat vector (ComponentArray), which the reference's NamedTuple-only method does not.
"""
function allreduce_gradients(gs::Union{NamedTuple, Tuple}; on_gpu::Bool=CUDA.functional())
    requests = Request[]
    nonblocking(g) = g
    function nonblocking(g::AbstractArray{<:Number})
        g, req = Iallreduce!(g, +)
        push!(requests, req)
        return g
    end
    gs = fmap(nonblocking, gs)
    waitall(requests)
    return gs
end

allreduce_gradients(g::AbstractArray{<:Number}; on_gpu::Bool=CUDA.functional()) = allreduce!(g, +)
