# test_F_cuda.jl — the reference's MATLAB comparison (test/F_test.jl:29-69) run through libdcegm_cuda.so.
#
# WRITTEN, NOT EXECUTED: there is no `julia` binary in the build image or on the GPU box (SURVEY.md §8c).  The same
# comparison runs from Python against the same 100 CSV files (tests/test_gpu_parity_f.py::
# test_F_test_model_matches_oracle_and_golden, fixture tests/golden/matlab_retirement_na500.npz).
#
# Usage, inside a checkout of floswald/DCEGM.jl with this repository's julia/ directory next to it:
#   ENV["DCEGM_CUDA_LIB"] = "/path/to/dcegm.jl_b200/csrc/libdcegm_cuda.so"
#   include("src/DCEGM.jl"); include("julia/DCEGMCuda.jl"); include("julia/test_F_cuda.jl")
using Test, DelimitedFiles
using .DCEGM: Param, FModel, getx, gety
import .DCEGMCuda

@testset "MATLAB comparison through libdcegm_cuda (test/F_test.jl)" begin
    mldir = joinpath(dirname(pathof(DCEGM)), "..", "matlab")   # test/F_test.jl:31
    pd = Dict(:na => 500, :beta => 0.95, :sigma => 0.35, :R => 1.05, :lambda => 0.000002, :delta => 0.0)
    pd[:beta] = 1 / pd[:R]                                      # test/F_test.jl:34-40

    # runf (src/dc_algo.jl:335-341) with dc_EGM! replaced by the library call
    p = Param(par = pd)
    m = FModel(p)
    DCEGMCuda.dc_EGM!(m, p)

    for i in 1:p.nD, it in p.nT:-1:1                            # test/F_test.jl:47-69
        mpo = readdlm(joinpath(mldir, "output", "policy_$(i)_$(it).csv"), ',')
        mvf = readdlm(joinpath(mldir, "output", "value_$(i)_$(it).csv"), ',')
        @testset "period $it choice $i" begin
            @test length(getx(m.c[i, it].env)) == length(mpo[1, :])
            @test length(getx(m.v[i, it].env)) == length(mvf[1, :])
            fedsort = sortperm(mpo[1, :])                       # the MATLAB policy is not always sorted
            @test maximum(abs.(getx(m.c[i, it].env) .- mpo[1, fedsort])) < 1e-4
            @test maximum(abs.(gety(m.c[i, it].env) .- mpo[2, fedsort])) < 1e-4
            if it == p.nT
                @test isapprox(getx(m.v[i, it].env)[1], mvf[1, 1], atol = 1e-4)
                @test isapprox(gety(m.v[i, it].env)[1], mvf[2, 1], atol = 1e-4)
            else
                @test maximum(abs.(getx(m.v[i, it].env) .- mvf[1, :])) < 1e-4
                @test maximum(abs.(gety(m.v[i, it].env) .- mvf[2, :])) < 5e-4   # ASCII export of the MATLAB values
            end
        end
    end

    # against the reference's own CPU solve, where both are available: discrete structure exact, ordinates 1e-10
    mref = FModel(p)
    DCEGM.dc_EGM!(mref, p)
    for i in 1:p.nD, it in 1:p.nT
        @test length(m.c[i, it].env.v) == length(mref.c[i, it].env.v)
        @test length(m.v[i, it].env.v) == length(mref.v[i, it].env.v)
        @test m.v[i, it].removed == mref.v[i, it].removed
        it < p.nT && @test isapprox(gety(m.c[i, it].env), gety(mref.c[i, it].env), rtol = 1e-10)
    end
end
