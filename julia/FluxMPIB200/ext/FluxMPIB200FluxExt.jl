module FluxMPIB200FluxExt

using Flux, FluxMPIB200, Functors

# reference: ext/FluxMPIFluxExt.jl:6-8 — returns the unwrapped model
function FluxMPIB200.synchronize!(x::FluxMPIFluxModel; root_rank::Integer=0)
    return fmap(y -> FluxMPIB200.synchronize!(y; root_rank), x.model)
end

end
