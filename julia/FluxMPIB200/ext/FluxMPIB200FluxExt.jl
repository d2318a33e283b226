module FluxMPIB200FluxExt

using Flux, FluxMPIB200, Functors

const Opt = FluxMPIB200.Optimisers

# the arrays `Flux.destructure` flattens (Optimisers.trainables from 0.3.2 on, Flux.params before)
_trainable_arrays(m) = isdefined(Opt, :trainables) ? Opt.trainables(m) : collect(Flux.params(m))

"""
    synchronize!(FluxMPIFluxModel(model); root_rank=0) -> model

Make every worker's copy of a Flux model equal to `root_rank`'s and return the *bare* model
(the wrapper only exists to select this method; behaviour of
/root/reference/ext/FluxMPIFluxExt.jl:6-8).

Trainable parameters travel as ONE message: `Flux.destructure` flattens them, the flat vector is
broadcast (host `Array`s through the shared-memory world, `CuArray`s by the NVLink broadcast
kernel) and the model is rebuilt from it. Whatever `destructure` does not cover — BatchNorm
running statistics, integer buffers, scalars — is then walked leaf by leaf with the generic
`synchronize!` methods; arrays already handled are skipped by identity.
"""
function FluxMPIB200.synchronize!(wrapped::FluxMPIFluxModel; root_rank::Integer=0)
    model = wrapped.model
    FluxMPIB200.total_workers() == 1 && return model
    flat, rebuild = Flux.destructure(model)
    if !isempty(flat)
        FluxMPIB200.bcast!(flat, root_rank)
        model = rebuild(flat)
    end
    done = IdDict{Any, Bool}()
    for p in _trainable_arrays(model)
        done[p] = true
    end
    rest(leaf) = haskey(done, leaf) ? leaf : FluxMPIB200.synchronize!(leaf; root_rank=root_rank)
    return fmap(rest, model)
end

end
