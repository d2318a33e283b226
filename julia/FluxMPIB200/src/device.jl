# Device world: one symmetric segment per rank (signal pads | all-reduce staging | broadcast
# staging), mapped into every peer — the Julia twin of fluxmpi_b200/parallel/device_world.py.
# CuArray pointers are passed straight to the kernels; nothing is staged through the host
# (the reference's CuArray methods do exactly that, src/mpi_extensions.jl:97-106).

mutable struct DeviceWorld
    handle::Ptr{Cvoid}                 # fmpi_symm*
    comms::Vector{DevComm}             # one per channel
    ar_off::UInt64
    ar_bytes::UInt64
    bc_off::UInt64
    bc_bytes::UInt64
    multicast::Bool
end

const _MB = UInt64(1) << 20

_stream() = Base.unsafe_convert(Ptr{Cvoid}, CUDA.stream().handle)

function _agree(ok::Bool)
    flags = host_allgather(UInt8[ok ? 0x01 : 0x00])
    return all(f -> f[1] == 0x01, flags)
end

function _device_bringup(w::World, dev::Integer)
    lib = cudalib()
    sig = Int(ccall(csym(:fmpi_sigpad_bytes), Csize_t, ()))
    sig = cld(sig, 1 << 16) << 16
    ar = parse(Int, get(ENV, "FMPI_AR_STAGE_MB", "160")) * Int(_MB)
    bc = parse(Int, get(ENV, "FMPI_BCAST_STAGE_MB", "64")) * Int(_MB)
    need = w.size == 1 ? sig + 4 * Int(_MB) : sig + ar + bc
    w.size == 1 && ((ar, bc) = (2 * Int(_MB), 2 * Int(_MB)))
    h = Ref{Ptr{Cvoid}}(C_NULL)
    mode = w.size == 1 ? 2 : (get(ENV, "FMPI_SYMM_MODE", "vmm") == "ipc" ? 1 : 0)
    mc = mode == 0 && ccall(csym(:fmpi_symm_mc_supported), Cint, (Cint,), dev) != 0
    mc = mode == 0 ? _agree(mc) : false
    ccheck(ccall(csym(:fmpi_symm_create), Cint, (Cint, Csize_t, Cint, Cint, Cint, Ptr{Ptr{Cvoid}}),
                 dev, need, mode, mc, w.size, h), "fmpi_symm_create")
    ccall(csym(:fmpi_symm_set_self), Cint, (Ptr{Cvoid}, Cint), h[], w.rank)
    mcptr = C_NULL
    if mode == 0
        tag = string(w.job, "-jl0")
        sock = Ref{Cint}(-1)
        fd = Ref{Cint}(-1)
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
    elseif mode == 1
        blob = Vector{UInt8}(undef, 64)
        ccheck(ccall(csym(:fmpi_symm_export_ipc), Cint, (Ptr{Cvoid}, Ptr{UInt8}), h[], blob), "export_ipc")
        blobs = host_allgather(blob)
        for q in 0:(w.size - 1)
            q == w.rank && continue
            ccheck(ccall(csym(:fmpi_symm_import_ipc), Cint, (Ptr{Cvoid}, Cint, Ptr{UInt8}), h[], q, blobs[q + 1]), "import_ipc")
        end
        host_barrier()
    end
    size = ccall(csym(:fmpi_symm_size), Csize_t, (Ptr{Cvoid},), h[])
    peers = [ccall(csym(:fmpi_symm_peer_ptr), Ptr{Cvoid}, (Ptr{Cvoid}, Cint), h[], q) for q in 0:(w.size - 1)]
    pad(v) = ntuple(i -> i <= length(v) ? v[i] : C_NULL, FMPI_MAX_RANKS)
    data = pad([p + sig for p in peers])
    sigs = pad(peers)
    mcd = mcptr == C_NULL ? C_NULL : mcptr + sig
    hp = Ref{Ptr{Cvoid}}(C_NULL)
    dp = Ref{Ptr{Cvoid}}(C_NULL)
    ccheck(ccall(csym(:fmpi_health_page), Cint, (Ptr{Ptr{Cvoid}}, Ptr{Ptr{Cvoid}}), hp, dp), "fmpi_health_page")
    budget = UInt64(round(parse(Float64, get(ENV, "FMPI_DEVICE_TIMEOUT_S", get(ENV, "FMPI_TIMEOUT_S", "300"))) * 1e9))
    comms = [DevComm(w.rank, w.size, 0, ch, data, sigs, mcd, UInt64(size - sig), dp[], budget) for ch in 0:(FMPI_NUM_CHANNELS - 1)]
    return DeviceWorld(h[], comms, 0, ar, ar, bc, mcptr != C_NULL)
end

_device_close(d::DeviceWorld) = ccall(csym(:fmpi_symm_destroy), Cint, (Ptr{Cvoid},), d.handle)

function _dev_allreduce!(v::CuArray{T}, op, root::Integer=-1; channel::Integer=0) where {T}
    w = WORLD[]
    (w.size == 1 || isempty(v)) && return v
    if w.device === nothing || !(op isa NativeOp)
        # P2P kernels switched off, a job over several nodes, or a custom reduction function:
        # through host memory, the way the reference does without CUDA-aware MPI
        # (src/mpi_extensions.jl:97-106)
        h = Array(v)
        out = root < 0 ? _host_allreduce!(h, op) : _host_reduce!(h, op, root)
        (root < 0 || w.rank == root) && copyto!(v, out)
        return v
    end
    d = w.device::DeviceWorld
    p = Ref{Ptr{Cvoid}}(Base.unsafe_convert(Ptr{Cvoid}, pointer(v)))
    comm = Ref(d.comms[channel + 1])
    la = Ref(AUTO)
    if root < 0
        ccheck(ccall(csym(:fmpi_allreduce_any), Cint,
                     (Ptr{DevComm}, Ptr{Ptr{Cvoid}}, Ptr{Ptr{Cvoid}}, UInt64, Cint, Cint, UInt64, UInt64, Ptr{Launch}, Ptr{Cvoid}),
                     comm, p, p, length(v), dtype_code(T), op_code(op), d.ar_off, d.ar_bytes, la, _stream()), "allreduce!")
    else
        ccheck(ccall(csym(:fmpi_reduce_any), Cint,
                     (Ptr{DevComm}, Ptr{Ptr{Cvoid}}, Ptr{Ptr{Cvoid}}, UInt64, Cint, Cint, Cint, UInt64, UInt64, Ptr{Launch}, Ptr{Cvoid}),
                     comm, p, p, length(v), dtype_code(T), op_code(op), root, d.ar_off, d.ar_bytes, la, _stream()), "reduce!")
    end
    return v
end

function _dev_bcast!(v::CuArray{T}, root::Integer; channel::Integer=0) where {T}
    w = WORLD[]
    (w.size == 1 || isempty(v)) && return v
    if w.device === nothing                 # see _dev_allreduce!
        h = Array(v)
        _host_bcast!(h, root)
        w.rank == root || copyto!(v, h)
        return v
    end
    d = w.device::DeviceWorld
    p = Ref{Ptr{Cvoid}}(Base.unsafe_convert(Ptr{Cvoid}, pointer(v)))
    ccheck(ccall(csym(:fmpi_bcast_any), Cint,
                 (Ptr{DevComm}, Ptr{Ptr{Cvoid}}, UInt64, Cint, UInt64, UInt64, Ptr{Launch}, Ptr{Cvoid}),
                 Ref(d.comms[channel + 1]), p, sizeof(v), root, d.bc_off, d.bc_bytes, Ref(AUTO), _stream()), "bcast!")
    return v
end

# ---- public primitives (reference: src/mpi_extensions.jl) ---------------------------------

"""
    allreduce!(v, op, comm=nothing)

In-place all-reduce; returns `v`. `Array`s use the shm host backend, `CuArray`s the NVLink
kernels on the current CUDA.jl stream (reference: src/mpi_extensions.jl:97-111).
"""
allreduce!(v::Array, op, comm=nothing) = _host_allreduce!(v, op)
allreduce!(v::CuArray, op, comm=nothing) = _dev_allreduce!(v, op)

"""    bcast!(v, root, comm=nothing)   (reference: src/mpi_extensions.jl:119-133)"""
bcast!(v::Array, root::Integer, comm=nothing) = _host_bcast!(v, root)
bcast!(v::CuArray, root::Integer, comm=nothing) = _dev_bcast!(v, root)

"""    reduce!(v, op, root, comm=nothing) — non-root buffers untouched (src/mpi_extensions.jl:141-155)"""
reduce!(v::Array, op, root::Integer, comm=nothing) = _host_reduce!(v, op, root)
reduce!(v::CuArray, op, root::Integer, comm=nothing) = _dev_allreduce!(v, op, root)

"""
A request returned by `Iallreduce!` / `Ibcast!`; `wait(req)` ≙ `MPI.Wait!`. Device requests wrap
a CUDA event recorded after the kernel; host requests wrap a `Task`.
"""
struct Request
    task::Union{Nothing, Task}
    event::Union{Nothing, CuEvent}
    buffer::Any
end
function Base.wait(r::Request)
    r.task === nothing || Base.wait(r.task)
    r.event === nothing || CUDA.wait(r.event)
    return r.buffer
end
waitall(rs) = foreach(wait, rs)

"""    Iallreduce!(sendbuf, recvbuf, op, comm=nothing) / Iallreduce!(buf, op, comm=nothing)
(reference: src/mpi_extensions.jl:26-60)"""
function Iallreduce!(send::CuArray, recv::CuArray, op, comm=nothing)
    send === recv || copyto!(recv, send)
    _dev_allreduce!(recv, op; channel=1)
    ev = CuEvent()
    CUDA.record(ev)
    return recv, Request(nothing, ev, recv)
end
function Iallreduce!(send::Array, recv::Array, op, comm=nothing)
    send === recv || copyto!(recv, send)
    # host collectives must keep issue order across ranks: run them on one sticky task chain
    t = @async _host_allreduce!(recv, op)
    return recv, Request(t, nothing, recv)
end
Iallreduce!(buf::AbstractArray, op, comm=nothing) = Iallreduce!(buf, buf, op, comm)

"""    Ibcast!(buf, root, comm=nothing)  (reference: src/mpi_extensions.jl:70-88)"""
function Ibcast!(buf::CuArray, root::Integer, comm=nothing)
    _dev_bcast!(buf, root; channel=1)
    ev = CuEvent()
    CUDA.record(ev)
    return buf, Request(nothing, ev, buf)
end
function Ibcast!(buf::Array, root::Integer, comm=nothing)
    t = @async _host_bcast!(buf, root)
    return buf, Request(t, nothing, buf)
end

# device movers (reference: src/mpi_extensions.jl:5-8)
cpu(x) = x
gpu(x) = x
cpu(x::AbstractArray) = Array(x)
gpu(x::AbstractArray) = CuArray(x)
