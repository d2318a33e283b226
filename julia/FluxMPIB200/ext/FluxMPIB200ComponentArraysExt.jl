module FluxMPIB200ComponentArraysExt

using ComponentArrays, FluxMPIB200

# A ComponentArray is one flat vector plus axes: broadcast / reduce the vector once and put the
# axes back (behaviour of /root/reference/ext/FluxMPIComponentArraysExt.jl:6-9; the
# allreduce_gradients method closes the gap named in SURVEY §2.1).
_rewrap(x::ComponentArray, flat) = ComponentArray(flat, getaxes(x))

FluxMPIB200.synchronize!(x::ComponentArray; root_rank::Integer=0) =
    _rewrap(x, FluxMPIB200.bcast!(getdata(x), root_rank))

FluxMPIB200.allreduce_gradients(x::ComponentArray; on_gpu::Bool=true) =
    _rewrap(x, FluxMPIB200.allreduce!(getdata(x), +))

end
