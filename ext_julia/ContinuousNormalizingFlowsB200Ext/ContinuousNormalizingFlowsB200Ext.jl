# ContinuousNormalizingFlowsB200Ext — reference-side binding of libcnfb200.so.
#
# UNTESTED HERE: Julia is not installed in the build environment.  It is the `ccall` form of what
# continuousnf.jl_b200/api.py does through ctypes (that Python mirror is what the parity tests
# drive).  Drop this directory into the reference's `ext/`, add to its Project.toml
#     [weakdeps]    CNFB200 = "<uuid of the tiny stub package that defines B200Libs>"
#     [extensions]  ContinuousNormalizingFlowsB200Ext = "CNFB200"
# (same mechanism as ext/ContinuousNormalizingFlowsCUDAExt, Project.toml:35-39) and callers change
# one keyword:  construct(RNODE, nn, nvars, naugs; resource = CNFB200.B200Libs()).
module ContinuousNormalizingFlowsB200Ext

import ContinuousNormalizingFlows as CNF
import ComputationalResources, ChainRulesCore, Lux, Random
import CNFB200: B200Libs   # struct B200Libs <: ComputationalResources.AbstractResource; device::Int; libpath::String; end

const MAXL = 8

# mirror of `cnf_config` / `cnf_stats` (include/cnfb200.h)
struct CnfConfig
    D::Int32; C::Int32; nvars::Int32; n_layers::Int32
    dims::NTuple{MAXL + 1, Int32}
    trace_mode::Int32; flags::UInt32
    lambda1::Float32; lambda2::Float32; lambda3::Float32
    reltol::Float32; abstol::Float32
    maxiters::Int64
    dt0::Float32; dtmax::Float32; dtmin::Float32
    device::Int32; kernel_path::Int32
end
mutable struct CnfStats
    naccept::Int64; nreject::Int64; nf::Int64
    retcode::Int32; kernel_path::Int32
    t_final::Float32; dt_last::Float32
    ms_total::Float64; ms_solve::Float64; gpu_launches::Int64
    ms_stage::Float64; n_stage::Int64
    CnfStats() = new(0, 0, 0, 0, 0, 0.0f0, 0.0f0, 0.0, 0.0, 0, 0.0, 0)
end

const B200ICNF{CM, INPLACE, COND, AUG, STEER, NZ, NJ, NZA} = CNF.ICNF{
    Float32, CM, INPLACE, COND, AUG, STEER, NZ, NJ, NZA,
    <:Any, <:Any, <:B200Libs,
}

trace_mode(::CNF.DIVecJacMatrixMode) = Int32(0)
trace_mode(::CNF.DIJacVecMatrixMode) = Int32(1)
trace_mode(::Any) = throw(ArgumentError("B200Libs supports DIVecJacMatrixMode / DIJacVecMatrixMode only"))

# Dense(tanh) chain -> layer widths; anything else is rejected (no CPU fallback in this path)
function layer_dims(nn)
    layers = nn isa Lux.Chain ? collect(values(nn.layers)) : [nn]
    all(l -> l isa Lux.Dense && l.activation in (tanh, Lux.NNlib.tanh_fast), layers) ||
        throw(ArgumentError("B200Libs supports Chain(Dense(in => out, tanh)...) only"))
    Int32[first(layers).in_dims; [l.out_dims for l in layers]...]
end

# per-layer weight/bias offsets inside the flat ComponentArray (0-based, in floats)
function offsets(dims)
    w = Int64[]; b = Int64[]; o = 0
    for l in 1:(length(dims) - 1)
        push!(w, o); o += dims[l] * dims[l + 1]
        push!(b, o); o += dims[l + 1]
    end
    w, b, o
end

const HANDLES = Dict{UInt, Ptr{Cvoid}}()
const HLOCK = ReentrantLock()

function handle(icnf::B200ICNF)
    lock(HLOCK) do
        get!(HANDLES, objectid(icnf)) do
            dims = layer_dims(icnf.nn)
            D = icnf.nvars + icnf.naugmented
            C = Int(dims[1]) - D
            kw = icnf.sol_kwargs
            haskey(kw, :alg) && !(nameof(typeof(kw.alg)) === :Tsit5) &&
                throw(ArgumentError("B200Libs implements the Tsit5 branch only"))
            flags = UInt32(0)
            C > 0 && (flags |= 0x01); icnf.naugmented != 0 && (flags |= 0x02)
            !iszero(icnf.λ₁) && (flags |= 0x04); !iszero(icnf.λ₂) && (flags |= 0x08); !iszero(icnf.λ₃) && (flags |= 0x10)
            cfg = CnfConfig(D, C, icnf.nvars, length(dims) - 1,
                ntuple(i -> i <= length(dims) ? dims[i] : Int32(0), MAXL + 1),
                trace_mode(icnf.compute_mode), flags, icnf.λ₁, icnf.λ₂, icnf.λ₃,
                get(kw, :reltol, 0.0f0), get(kw, :abstol, 0.0f0), min(get(kw, :maxiters, 0), typemax(Int64) ÷ 2),
                get(kw, :dt, 0.0f0), get(kw, :dtmax, 0.0f0), get(kw, :dtmin, 0.0f0),
                icnf.resource.device, 0)
            h = Ref{Ptr{Cvoid}}(C_NULL)
            rc = ccall((:cnfb200_create, icnf.resource.libpath), Int32, (Ref{CnfConfig}, Ref{Ptr{Cvoid}}), cfg, h)
            rc == 0 || error(unsafe_string(ccall((:cnfb200_last_error, icnf.resource.libpath), Cstring, (Ptr{Cvoid},), h[])))
            h[]
        end
    end
end

function check(icnf, h, rc, what)
    rc < 0 && error(what * ": " * unsafe_string(ccall((:cnfb200_last_error, icnf.resource.libpath), Cstring, (Ptr{Cvoid},), h)))
    rc > 0 && @warn "$what: solver retcode $rc (1 MaxIters, 2 DtLessThanMin, 3 Unstable)"
    nothing
end

function set_params!(icnf, h, ps)
    θ = convert(Vector{Float32}, collect(ps))          # ComponentArray data vector (mlj_ext/core_icnf.jl:37-38)
    w, b, n = offsets(layer_dims(icnf.nn))
    GC.@preserve θ w b check(icnf, h, ccall((:cnfb200_set_params, icnf.resource.libpath), Int32,
        (Ptr{Cvoid}, Ptr{Float32}, Int64, Ptr{Int64}, Ptr{Int64}), h, θ, n, w, b), "cnfb200_set_params")
end

# ---- inference(icnf, mode, xs[, ys], ps, st): replaces base_icnf.jl:407-426 for RESOURCE <: B200Libs
function _inference(icnf::B200ICNF, mode::CNF.Mode, xs::AbstractMatrix, ys, ps)
    h = handle(icnf); set_params!(icnf, h, ps)
    D = icnf.nvars + icnf.naugmented; B = size(xs, 2)
    train = mode isa CNF.TrainMode
    x = convert(Matrix{Float32}, xs)                    # column-major nvars x B == what the ABI expects
    y = ys === nothing ? Float32[] : convert(Matrix{Float32}, ys)
    ϵ = Matrix{Float32}(undef, D, B); train && Random.rand!(icnf.rng, icnf.epsdist, ϵ)   # base_icnf.jl:277-278
    t₀, t₁ = CNF.steer_tspan(icnf, mode)                # base_icnf.jl:108-121
    logp = Vector{Float32}(undef, B); regs = Matrix{Float32}(undef, train ? 3 : 1, B); st = CnfStats()
    GC.@preserve x y ϵ logp regs st begin
        rc = ccall((:cnfb200_logdensity, icnf.resource.libpath), Int32,
            (Ptr{Cvoid}, Int32, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Int64, Float32, Float32,
             Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Ref{CnfStats}),
            h, train ? 1 : 0, x, ys === nothing ? C_NULL : pointer(y), train ? pointer(ϵ) : C_NULL, B, t₀, t₁,
            logp, regs, C_NULL, st)
        check(icnf, h, rc, "cnfb200_logdensity")
    end
    (logp, eachrow(regs))                               # same shape as inference_sol's return, base_icnf.jl:188
end
CNF.inference(icnf::B200ICNF{<:CNF.MatrixMode}, mode::CNF.Mode, xs::AbstractMatrix{<:Real}, ps::Any, st::NamedTuple) =
    _inference(icnf, mode, xs, nothing, ps)
CNF.inference(icnf::B200ICNF{<:CNF.MatrixMode}, mode::CNF.Mode, xs::AbstractMatrix{<:Real}, ys::AbstractMatrix{<:Real}, ps::Any, st::NamedTuple) =
    _inference(icnf, mode, xs, ys, ps)

# ---- fused loss + discrete-adjoint gradient; rrule so AutoZygote works, EnzymeRules analogous
function loss_grad(icnf::B200ICNF, xs, ys, ps; want_gx = false, mode::CNF.Mode = CNF.TrainMode())
    h = handle(icnf); set_params!(icnf, h, ps)
    D = icnf.nvars + icnf.naugmented; B = size(xs, 2)
    x = convert(Matrix{Float32}, xs); y = ys === nothing ? Float32[] : convert(Matrix{Float32}, ys)
    ϵ = Matrix{Float32}(undef, D, B); Random.rand!(icnf.rng, icnf.epsdist, ϵ)
    t₀, t₁ = CNF.steer_tspan(icnf, mode)
    ℓ = Ref{Float32}(0); g = zeros(Float32, length(ps)); gx = want_gx ? similar(x) : Float32[]; st = CnfStats()
    GC.@preserve x y ϵ g gx st begin
        rc = ccall((:cnfb200_loss_grad_mode, icnf.resource.libpath), Int32,
            (Ptr{Cvoid}, Int32, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Int64, Float32, Float32, Ref{Float32},
             Ptr{Float32}, Ptr{Float32}, Ref{CnfStats}),
            h, mode isa CNF.TrainMode ? 1 : 0, x, ys === nothing ? C_NULL : pointer(y), ϵ, B, t₀, t₁, ℓ, g,
            want_gx ? pointer(gx) : C_NULL, st)
        check(icnf, h, rc, "cnfb200_loss_grad_mode")
    end
    ℓ[], g, gx
end
CNF.loss(icnf::B200ICNF{<:CNF.MatrixMode}, ::CNF.TrainMode, xs::AbstractMatrix{<:Real}, ps::Any, st::NamedTuple) =
    first(loss_grad(icnf, xs, nothing, ps))
CNF.loss(icnf::B200ICNF{<:CNF.MatrixMode}, ::CNF.TrainMode, xs::AbstractMatrix{<:Real}, ys::AbstractMatrix{<:Real}, ps::Any, st::NamedTuple) =
    first(loss_grad(icnf, xs, ys, ps))
# (TestMode: loss = -mean(logp̂x), base_icnf.jl:489-508; same entry point, exact-trace reverse sweep)
function ChainRulesCore.rrule(::typeof(CNF.loss), icnf::B200ICNF{<:CNF.MatrixMode}, mode::CNF.Mode, xs::AbstractMatrix{<:Real}, ps::Any, st::NamedTuple)
    ℓ, g, gx = loss_grad(icnf, xs, nothing, ps; want_gx = true, mode = mode)
    back(ℓ̄) = (ChainRulesCore.NoTangent(), ChainRulesCore.NoTangent(), ChainRulesCore.NoTangent(), ℓ̄ .* gx, ℓ̄ .* g, ChainRulesCore.NoTangent())
    ℓ, back
end

# ---- generate(icnf, mode, [ys,] ps, st, n): replaces base_icnf.jl:447-466
function _generate(icnf::B200ICNF, mode::CNF.Mode, ys, ps, n::Int)
    h = handle(icnf); set_params!(icnf, h, ps)
    D = icnf.nvars + icnf.naugmented; train = mode isa CNF.TrainMode
    z0 = Matrix{Float32}(undef, D, n); Random.rand!(icnf.rng, icnf.basedist, z0)         # base_icnf.jl:367-368
    ϵ = Matrix{Float32}(undef, D, n); train && Random.rand!(icnf.rng, icnf.epsdist, ϵ)
    y = ys === nothing ? Float32[] : convert(Matrix{Float32}, ys)
    t₀, t₁ = CNF.steer_tspan(icnf, mode)
    out = Matrix{Float32}(undef, icnf.nvars, n); st = CnfStats()
    GC.@preserve z0 ϵ y out st begin
        rc = ccall((:cnfb200_generate, icnf.resource.libpath), Int32,
            (Ptr{Cvoid}, Int32, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Int64, Float32, Float32, Ptr{Float32}, Ref{CnfStats}),
            h, train ? 1 : 0, z0, ys === nothing ? C_NULL : pointer(y), train ? pointer(ϵ) : C_NULL, n, t₁, t₀, out, st)  # reversed tspan, :377
        check(icnf, h, rc, "cnfb200_generate")
    end
    out
end
CNF.generate(icnf::B200ICNF{<:CNF.MatrixMode}, mode::CNF.Mode, ps::Any, st::NamedTuple, n::Int) = _generate(icnf, mode, nothing, ps, n)
CNF.generate(icnf::B200ICNF{<:CNF.MatrixMode}, mode::CNF.Mode, ys::AbstractMatrix{<:Real}, ps::Any, st::NamedTuple, n::Int) = _generate(icnf, mode, ys, ps, n)

# rng_AT / base_AT stay the host defaults (base_icnf.jl:123-135): MLJ fit/transform and ICNFDist
# keep working unchanged because host arrays stay host arrays.
end
