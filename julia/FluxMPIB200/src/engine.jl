# FlatEngine: the fused gradient all-reduce + optimiser step from Julia.
#
# A Lux / ComponentArrays script keeps its parameters in one flat vector. FlatEngine places that
# vector and a gradient vector of the same length in a symmetric segment (every peer GPU maps
# it), and `step!` is ONE call into `fmpi_fused_step` per bucket: the owner of each slice sums
# the W gradient copies over NVLink, applies Adam / AdamW / Descent / Momentum / Nesterov in
# registers and writes the new parameters into every replica. This replaces the reference's
# `Optimisers.update` with a `DistributedOptimizer` rule (one blocking, host-staged MPI
# all-reduce per leaf and the rule's elementwise kernels, src/optimizer.jl:20-23).
#
#   eng = FlatEngine(getdata(ps_flat), Optimisers.Adam(1f-3))       # copies ps in, syncs from rank 0
#   ps  = ComponentArray(params(eng), getaxes(ps_flat))             # views of the symmetric vector
#   gs  = gradient(loss, ps)[1]                                     # any flat gradient
#   copyto!(grads(eng), getdata(gs)); step!(eng)                    # ps is updated in place on all ranks
#
# Gradients are SUMMED (reference semantics); pass `average=true` for the mean.

# typedef struct { int32 rule; float eta, beta1, beta2, eps, lambda, rho, grad_scale;
#                  int32 zero_grad; uint64 eta_dev; } fmpi_hyper;        (48 bytes)
struct Hyper
    rule::Int32
    eta::Float32
    beta1::Float32
    beta2::Float32
    eps::Float32
    lambda::Float32
    rho::Float32
    grad_scale::Float32
    zero_grad::Int32
    _pad::Int32
    eta_dev::UInt64
end

const RULE_DESCENT, RULE_MOMENTUM, RULE_NESTEROV, RULE_ADAM, RULE_ADAMW = Int32(0), Int32(1), Int32(2), Int32(3), Int32(4)

function _hyper(rule, scale::Float32)
    mk(r; eta=0f0, b1=0.9f0, b2=0.999f0, eps=1f-8, lambda=0f0, rho=0f0) =
        Hyper(r, eta, b1, b2, eps, lambda, rho, scale, Int32(1), Int32(0), UInt64(0))
    if rule isa Optimisers.Descent
        return mk(RULE_DESCENT; eta=Float32(rule.eta))
    elseif rule isa Optimisers.Momentum
        return mk(RULE_MOMENTUM; eta=Float32(rule.eta), rho=Float32(rule.rho))
    elseif rule isa Optimisers.Nesterov
        return mk(RULE_NESTEROV; eta=Float32(rule.eta), rho=Float32(rule.rho))
    elseif rule isa Optimisers.Adam
        return mk(RULE_ADAM; eta=Float32(rule.eta), b1=Float32(rule.beta[1]), b2=Float32(rule.beta[2]),
                  eps=Float32(rule.epsilon))
    end
    throw(FluxMPIError("FlatEngine: $(typeof(rule)) has no fused kernel (use DistributedOptimizer)"))
end

mutable struct FlatEngine{T}
    seg::Ptr{Cvoid}                 # fmpi_symm* of the [P | G] segment
    comm::DevComm
    n::Int                          # elements (padded to 64)
    p_off::UInt64
    g_off::UInt64
    P::CuVector{T}
    G::CuVector{T}
    m::CuVector{Float32}
    v::CuVector{Float32}
    master::Union{Nothing, CuVector{Float32}}
    step::CuVector{Int64}
    hyper::Hyper
    bucket::Int                     # elements per bucket (multiple of 64)
    nvls::Bool
    oneshot::Bool                   # whole model <= 2 MB: push + one barrier + replicated rule
    stage_off::UInt64               # push slots [2 parities][W][n * sizeof(T)] behind [P | G]
    stage_bytes::UInt64
end

params(e::FlatEngine) = e.P
grads(e::FlatEngine) = e.G

# A second symmetric segment next to the base one; same handle exchange as `_device_bringup`.
function _alloc_segment(w::World, dev::Integer, nbytes::Integer, tagno::Integer)
    h = Ref{Ptr{Cvoid}}(C_NULL)
    if w.size == 1
        ccheck(ccall(csym(:fmpi_symm_create), Cint, (Cint, Csize_t, Cint, Cint, Cint, Ptr{Ptr{Cvoid}}),
                     dev, nbytes, 2, 0, 1, h), "fmpi_symm_create")
        ccall(csym(:fmpi_symm_set_self), Cint, (Ptr{Cvoid}, Cint), h[], 0)
        return h[], C_NULL
    end
    mc = (w.device::DeviceWorld).multicast
    ccheck(ccall(csym(:fmpi_symm_create), Cint, (Cint, Csize_t, Cint, Cint, Cint, Ptr{Ptr{Cvoid}}),
                 dev, nbytes, 0, mc, w.size, h), "fmpi_symm_create")
    ccall(csym(:fmpi_symm_set_self), Cint, (Ptr{Cvoid}, Cint), h[], w.rank)
    tag = string(w.job, "-jl", tagno)
    sock, fd = Ref{Cint}(-1), Ref{Cint}(-1)
    ccheck(ccall(csym(:fmpi_symm_export_fd), Cint, (Ptr{Cvoid}, Ptr{Cint}), h[], fd), "export_fd")
    hcheck(ccall(hsym(:fmpi_fdx_open), Cint, (Cstring, Cint, Ptr{Cint}), tag, w.rank, sock), "fdx_open")
    host_barrier()
    fds = Vector{Cint}(undef, w.size)
    hcheck(ccall(hsym(:fmpi_fdx_exchange), Cint, (Cint, Cstring, Cint, Cint, Cint, Ptr{Cint}, Cdouble),
                 sock[], tag, w.rank, w.size, fd[], fds, 60.0), "fdx_exchange")
    for q in 0:(w.size - 1)
        q == w.rank && continue
        ccheck(ccall(csym(:fmpi_symm_import_fd), Cint, (Ptr{Cvoid}, Cint, Cint), h[], q, fds[q + 1]), "import_fd")
    end
    host_barrier()
    mcptr = C_NULL
    if mc
        mfd = Ref{Cint}(-1)
        ok = w.rank != 0 || ccall(csym(:fmpi_symm_mc_create), Cint, (Ptr{Cvoid}, Ptr{Cint}), h[], mfd) == 0
        if _agree(ok)
            ok = ccall(hsym(:fmpi_fdx_bcast), Cint, (Cint, Cstring, Cint, Cint, Cint, Ptr{Cint}, Cdouble),
                       sock[], tag, w.rank, w.size, 0, mfd, 60.0) == 0
            ok = ok && (w.rank == 0 || ccall(csym(:fmpi_symm_mc_import), Cint, (Ptr{Cvoid}, Cint), h[], mfd[]) == 0)
            ok = _agree(ok) && _agree(ccall(csym(:fmpi_symm_mc_add_device), Cint, (Ptr{Cvoid},), h[]) == 0)
            ok = ok && _agree(ccall(csym(:fmpi_symm_mc_bind_map), Cint, (Ptr{Cvoid},), h[]) == 0)
            ok && (mcptr = ccall(csym(:fmpi_symm_mc_ptr), Ptr{Cvoid}, (Ptr{Cvoid},), h[]))
        end
    end
    ccall(hsym(:fmpi_fdx_close), Cint, (Cint, Cstring, Cint), sock[], tag, w.rank)
    return h[], mcptr
end

const _segment_counter = Ref(1)

"""
    FlatEngine(ps::CuVector{T}, rule; average=false, bucket_mb=32, channel=2)

Collective. `T` is `Float32`, `Float16` or `BFloat16` (16-bit parameters get an fp32 master
copy). The parameters are copied into symmetric memory and broadcast from rank 0.
"""
function FlatEngine(ps::CuVector{T}, rule; average::Bool=false, bucket_mb::Real=32, channel::Integer=2) where {T}
    w = _require_world()
    w.device === nothing && throw(FluxMPIError("FlatEngine needs Init() on a CUDA machine"))
    dev = CUDA.deviceid(CUDA.device())
    n = cld(length(ps), 64) * 64
    es = sizeof(T)
    buf = cld(n * es, 1 << 21) << 21
    no = _segment_counter[]
    _segment_counter[] += 1
    # a Quick-Start-sized model is latency-bound: one-shot flavour (csrc/kernels/fused_step.cu)
    oneshot = w.size > 1 && n * es <= parse(Int, get(ENV, "FMPI_FUSED_ONESHOT_KB", "2048")) * 1024
    stage = oneshot ? 2 * w.size * (cld(n * es, 256) * 256) : 0
    seg, mcptr = _alloc_segment(w, dev, 2 * buf + stage, no)
    base = (w.device::DeviceWorld).comms[1]
    peers = [ccall(csym(:fmpi_symm_peer_ptr), Ptr{Cvoid}, (Ptr{Cvoid}, Cint), seg, q) for q in 0:(w.size - 1)]
    pad(v) = ntuple(i -> i <= length(v) ? v[i] : C_NULL, FMPI_MAX_RANKS)
    comm = DevComm(w.rank, w.size, 0, channel, pad(peers), base.sig, mcptr,
                   UInt64(ccall(csym(:fmpi_symm_size), Csize_t, (Ptr{Cvoid},), seg)), base.ctrl, base.timeout_ns)
    loc = ccall(csym(:fmpi_symm_local_ptr), Ptr{Cvoid}, (Ptr{Cvoid},), seg)
    P = unsafe_wrap(CuArray, reinterpret(CuPtr{T}, UInt(loc)), n)
    G = unsafe_wrap(CuArray, reinterpret(CuPtr{T}, UInt(loc) + UInt(buf)), n)
    fill!(P, zero(T)); fill!(G, zero(T))
    copyto!(P, 1, ps, 1, length(ps))
    scale = average ? 1f0 / w.size : 1f0
    master = T === Float32 ? nothing : CuVector{Float32}(P)
    bucket = max(64, (round(Int, bucket_mb * (1 << 20) / es) ÷ 64) * 64)
    eng = FlatEngine{T}(seg, comm, n, UInt64(0), UInt64(buf), P, G, CUDA.zeros(Float32, n), CUDA.zeros(Float32, n),
                        master, CUDA.zeros(Int64, 1), _hyper(rule, scale), oneshot ? n : bucket,
                        mcptr != C_NULL && w.size >= 4, oneshot, UInt64(2 * buf), UInt64(stage))
    if w.size > 1                       # replicas start from rank 0's values, in one kernel
        ccheck(ccall(csym(:fmpi_bcast_symm), Cint, (Ptr{DevComm}, UInt64, UInt64, Cint, Ptr{Launch}, Ptr{Cvoid}),
                     Ref(comm), 0, n * es, 0, Ref(AUTO), _stream()), "bcast_symm")
        master === nothing || copyto!(master, CuVector{Float32}(P))
    end
    return eng
end

"""
    step!(eng)

Reduce `grads(eng)` over all workers and apply the rule; `params(eng)` is updated in place on
every rank and `grads(eng)` is cleared. Stream-ordered on the current CUDA.jl stream.
"""
function step!(e::FlatEngine{T}) where {T}
    code = dtype_code(T)
    one(x) = x === nothing ? C_NULL : Ref{Ptr{Cvoid}}(Base.unsafe_convert(Ptr{Cvoid}, pointer(x)))
    la = Ref(Launch(0, 0, e.nvls ? 3 : 2))       # 0 CTAs: the library's measured default
    if e.oneshot
        ccheck(ccall(csym(:fmpi_fused_step_oneshot), Cint,
                     (Ptr{DevComm}, UInt64, UInt64, Cint, Ptr{Ptr{Cvoid}}, Ptr{Ptr{Cvoid}}, Ptr{Ptr{Cvoid}},
                      UInt64, UInt64, Ptr{Hyper}, Ptr{Ptr{Cvoid}}, UInt64, UInt64, Ptr{Launch}, Ptr{Cvoid}),
                     Ref(e.comm), e.g_off, e.p_off, code, one(e.m), one(e.v), one(e.master),
                     0, e.n, Ref(e.hyper), one(e.step), e.stage_off, e.stage_bytes, Ref(AUTO), _stream()),
               "fused_step_oneshot")
        ccheck(ccall(csym(:fmpi_step_advance), Cint, (Ptr{Cvoid}, Ptr{Cvoid}),
                     Base.unsafe_convert(Ptr{Cvoid}, pointer(e.step)), _stream()), "step_advance")
        return e
    end
    lo = 0
    while lo < e.n
        hi = min(lo + e.bucket, e.n)
        ccheck(ccall(csym(:fmpi_fused_step), Cint,
                     (Ptr{DevComm}, UInt64, Cint, UInt64, Cint, Ptr{Ptr{Cvoid}}, Ptr{Ptr{Cvoid}}, Ptr{Ptr{Cvoid}},
                      UInt64, UInt64, Ptr{Hyper}, Ptr{Ptr{Cvoid}}, Ptr{Launch}, Ptr{Cvoid}),
                     Ref(e.comm), e.g_off, code, e.p_off, code, one(e.m), one(e.v), one(e.master),
                     lo, hi, Ref(e.hyper), one(e.step), la, _stream()), "fused_step")
        lo = hi
    end
    ccheck(ccall(csym(:fmpi_step_advance), Cint, (Ptr{Cvoid}, Ptr{Cvoid}),
                 Base.unsafe_convert(Ptr{Cvoid}, pointer(e.step)), _stream()), "step_advance")
    return e
end

close!(e::FlatEngine) = ccall(csym(:fmpi_symm_destroy), Cint, (Ptr{Cvoid},), e.seg)
