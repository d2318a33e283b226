module FluxMPIB200ComponentArraysExt

using ComponentArrays, FluxMPIB200

# One broadcast / all-reduce of the flat backing vector
# (reference: ext/FluxMPIComponentArraysExt.jl:6-9 for synchronize!; the all-reduce method is
# the capability gap named in SURVEY §2.1).
function FluxMPIB200.synchronize!(x::ComponentArray; root_rank::Integer=0)
    d = FluxMPIB200.bcast!(getdata(x), root_rank)
    return ComponentArray(d, getaxes(x))
end

function FluxMPIB200.allreduce_gradients(x::ComponentArray; on_gpu::Bool=true)
    d = FluxMPIB200.allreduce!(getdata(x), +)
    return ComponentArray(d, getaxes(x))
end

end
