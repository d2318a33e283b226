# synchronize! (reference: src/synchronize.jl:10-35). Same dispatch table; the array method is
# truly in place for CuArrays because nothing is staged through the host.

"""
    synchronize!(x; root_rank=0)

Synchronize `x` across all processes from `root_rank`; returns the synchronised tree.
"""
function synchronize!(ps::Union{NamedTuple, Tuple}; root_rank::Integer=0)
    length(ps) == 0 && return ps
    return fmap(x -> synchronize!(x; root_rank), ps)
end

synchronize!(x::AbstractArray{T}; root_rank::Integer=0) where {T <: Number} = bcast!(x, root_rank)

synchronize!(x::AbstractArray; root_rank::Integer=0) = synchronize!.(x; root_rank)

function synchronize!(l::Optimisers.Leaf; root_rank::Integer=0)
    @set! l.state = synchronize!(l.state; root_rank)
    return l
end

synchronize!(x::Number; root_rank::Integer=0) = bcast!([x], root_rank)[1]

# Symbols, Nothing, Missing, Val, ...: left alone
synchronize!(x; root_rank::Integer=0) = x
