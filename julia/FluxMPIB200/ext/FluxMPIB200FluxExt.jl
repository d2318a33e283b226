module FluxMPIB200FluxExt

using Flux, FluxMPIB200, Functors

# Walk the wrapped Flux model and hand back the bare model
# (behaviour of /root/reference/ext/FluxMPIFluxExt.jl:6-8).
function FluxMPIB200.synchronize!(wrapped::FluxMPIFluxModel; root_rank::Integer=0)
    each(leaf) = FluxMPIB200.synchronize!(leaf; root_rank=root_rank)
    return fmap(each, wrapped.model)
end

end
