# Library loading and mirrors of the C types in include/fmpi.h.

const FMPI_MAX_RANKS = 8
const FMPI_NUM_CHANNELS = 4

const _libdir = Ref{String}("")
const _host = Ref{Ptr{Cvoid}}(C_NULL)
const _cuda = Ref{Ptr{Cvoid}}(C_NULL)

function _lib_dir()
    if isempty(_libdir[])
        _libdir[] = get(ENV, "FLUXMPI_B200_LIB",
                        normpath(joinpath(@__DIR__, "..", "..", "..", "fluxmpi_b200", "_lib")))
    end
    return _libdir[]
end

function hostlib()
    _host[] == C_NULL && (_host[] = Libdl.dlopen(joinpath(_lib_dir(), "libfmpi_host.so")))
    return _host[]
end

function cudalib()
    _cuda[] == C_NULL && (_cuda[] = Libdl.dlopen(joinpath(_lib_dir(), "libfmpi_cuda.so")))
    return _cuda[]
end

hsym(name::Symbol) = Libdl.dlsym(hostlib(), name)
csym(name::Symbol) = Libdl.dlsym(cudalib(), name)

struct FluxMPIError <: Exception
    msg::String
end
Base.showerror(io::IO, e::FluxMPIError) = print(io, "FluxMPIB200: ", e.msg)

function hcheck(rc::Cint, what)
    rc == 0 && return
    throw(FluxMPIError(string(what, ": ", unsafe_string(ccall(hsym(:fmpi_last_error), Cstring, ())))))
end
function ccheck(rc::Cint, what)
    rc == 0 && return
    throw(FluxMPIError(string(what, ": ", unsafe_string(ccall(csym(:fmpi_cuda_last_error), Cstring, ())))))
end

# fmpi_dtype / fmpi_op
dtype_code(::Type{Float32}) = Cint(0)
dtype_code(::Type{Float64}) = Cint(1)
dtype_code(::Type{Float16}) = Cint(2)
dtype_code(::Type{Int32}) = Cint(4)
dtype_code(::Type{Int64}) = Cint(5)
dtype_code(::Type{UInt8}) = Cint(6)
dtype_code(::Type{Int8}) = Cint(7)
dtype_code(::Type{Int16}) = Cint(8)
dtype_code(::Type{Bool}) = Cint(6)
dtype_code(T::Type) = string(T) == "BFloat16" ? Cint(3) : throw(FluxMPIError("unsupported eltype $T"))

# The reference accepts any Julia function and wraps it in MPI.Op (src/mpi_extensions.jl:52-54);
# here the four reductions the kernels implement.
op_code(::typeof(+)) = Cint(0)
op_code(::typeof(*)) = Cint(1)
op_code(::typeof(min)) = Cint(2)
op_code(::typeof(max)) = Cint(3)
op_code(f) = throw(FluxMPIError("unsupported reduction $f (use +, *, min, max)"))

# typedef struct { int32 rank, world, emulate, channel; void* data[8]; uint32* sig[8];
#                  void* mc_data; uint64 data_bytes; uint32* ctrl; uint64 timeout_ns; } fmpi_devcomm;
# (ctrl = device view of the health page: a lost peer ends a wait instead of hanging it)
struct DevComm
    rank::Int32
    world::Int32
    emulate::Int32
    channel::Int32
    data::NTuple{FMPI_MAX_RANKS, Ptr{Cvoid}}
    sig::NTuple{FMPI_MAX_RANKS, Ptr{Cvoid}}
    mc_data::Ptr{Cvoid}
    data_bytes::UInt64
    ctrl::Ptr{Cvoid}
    timeout_ns::UInt64
end

# typedef struct { int32 blocks, threads, algo; } fmpi_launch;
struct Launch
    blocks::Int32
    threads::Int32
    algo::Int32
end
const AUTO = Launch(0, 0, 0)
