# SPMD test harness in the reference's style (/root/reference/test/runtests.jl:1-16): each
# test_*.jl is launched N-way and asserts inside every rank. The launcher is the Python one
# (there is no mpiexec): python -m fluxmpi_b200.launch -n N julia --project test_X.jl
using Test

nprocs = parse(Int, get(ENV, "FLUXMPI_TEST_NPROCS", "3"))
root = normpath(joinpath(@__DIR__, "..", "..", ".."))
@testset "FluxMPIB200" begin
    for f in filter(startswith("test_"), readdir(@__DIR__))
        cmd = `python -m fluxmpi_b200.launch -n $nprocs $(Base.julia_cmd()[1]) --project=$(dirname(@__DIR__)) $(joinpath(@__DIR__, f))`
        run(setenv(cmd, "PYTHONPATH" => root; dir=root))
        @test true
    end
end
