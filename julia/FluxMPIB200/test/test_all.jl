# The reference's assertions (test/test_common.jl, test_data.jl, test_mpi_extensions.jl,
# test_optimizer.jl, test_synchronize.jl) against this binding, on host arrays.
using FluxMPIB200, Optimisers, Test
const F = FluxMPIB200

F.Init(; verbose=true)
r, w = local_rank(), total_workers()
@test F.Initialized() && r < w

x = ones(4)
@test F.allreduce!(copy(x), +) == x .* w
@test F.allreduce!(copy(x), *) == x
b = r == 0 ? ones(2, 3) : zeros(2, 3)
@test F.bcast!(b, 0) == ones(2, 3)
y = F.reduce!(copy(x), +, 0)
@test y == (r == 0 ? x .* w : x)
y = zeros(4); _, req = F.Iallreduce!(x, y, +); wait(req)
@test y == x .* w

data = randn(Float32, 10)
ddc = DistributedDataContainer(data)
@test length(ddc) == (r != w - 1 ? cld(10, w) : 10 - (w - 1) * cld(10, w))

ps = (a=zeros(4), b=zeros(4))
st = Optimisers.setup(DistributedOptimizer(Adam(0.001)), ps)
@test st.a.state == Optimisers.setup(Adam(0.001), ps).a.state
_, p1 = Optimisers.update(st, ps, (a=ones(4), b=ones(4)))
_, p2 = Optimisers.update(Optimisers.setup(Adam(0.001), ps), ps, (a=fill(Float64(w), 4), b=fill(Float64(w), 4)))
@test isapprox(p1.a, p2.a; atol=1e-5)
gs = allreduce_gradients((a=ones(4), b=ones(4)); on_gpu=false)
@test gs.a == fill(Float64(w), 4)

t = (a=(b=(r == 0 ? ones(2, 3) : zeros(2, 3)), c=zeros(2, 3) .+ (r == 0)), d=zeros(2, 3) .+ (r == 0))
t = F.synchronize!(t; root_rank=0)
@test t.a.b == ones(2, 3) && t.d == ones(2, 3)
@test F.synchronize!(nothing) === nothing
@test F.synchronize!(r == 0 ? :random : :other) == (r == 0 ? :random : :other)
@test F.synchronize!(local_rank(); root_rank=0) == 0
fluxmpi_println("ok")
