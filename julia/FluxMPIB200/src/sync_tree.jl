# synchronize!: make every rank's copy of a parameter / state / optimiser-state tree equal to
# root_rank's. Contract: /root/reference/src/synchronize.jl:10-35 (what each kind of leaf does).
# One recursive walker with explicit leaf classification instead of a method per type, so the
# ComponentArrays / Flux extensions only add a classification rule.

@enum LeafKind NumericArray ObjectArray OptLeaf Scalar Container Opaque

_kind(::AbstractArray{<:Number}) = NumericArray
_kind(::AbstractArray) = ObjectArray
_kind(::Optimisers.Leaf) = OptLeaf
_kind(::Number) = Scalar
_kind(::Union{NamedTuple, Tuple}) = Container
_kind(::Any) = Opaque            # Symbol, Nothing, Missing, Val, ... are left as they are

function _sync(x, root::Integer)
    k = _kind(x)
    if k == NumericArray
        return bcast!(x, root)
    elseif k == ObjectArray
        return map(e -> _sync(e, root), x)
    elseif k == OptLeaf
        return Setfield.@set x.state = _sync(x.state, root)
    elseif k == Scalar
        return bcast!([x], root)[1]
    elseif k == Container
        isempty(x) && return x
        return fmap(e -> _sync(e, root), x)
    end
    return x
end

"""
    synchronize!(x; root_rank=0)

Returns the synchronised tree (arrays are also updated in place — device arrays included,
since nothing is staged through the host).
"""
synchronize!(x; root_rank::Integer=0) = _sync(x, root_rank)
