# ContinuousNormalizingFlowsB200Ext — reference-side binding of libcnfb200.so (include/cnfb200.h).
#
# UNTESTED HERE: Julia is not installed in the build environment.  It is the `ccall` form of what
# continuousnf.jl_b200/api.py does through ctypes (that Python mirror is what the parity tests drive);
# tests/test_cabi.py parses every `ccall` below against the prototypes of include/cnfb200.h (argument count
# and types), so the two cannot drift apart silently.
#
# Drop this directory into the reference's `ext/`, add to its Project.toml
#     [weakdeps]    CNFB200 = "<uuid of the stub package that defines B200Libs>"
#     [extensions]  ContinuousNormalizingFlowsB200Ext = "CNFB200"
# (same mechanism as ext/ContinuousNormalizingFlowsCUDAExt, Project.toml:35-39) and callers change one keyword:
#     construct(RNODE, nn, nvars, naugs; resource = CNFB200.B200Libs())
#
# What dispatches here (RESOURCE <: B200Libs, T == Float32):
#   inference / generate / loss          src/base_icnf.jl:407-466, src/icnf.jl:481-502, src/base_icnf.jl:489-508
#   d(loss)/d(ps), d(loss)/d(xs)         EnzymeRules (the default adtype of ICNFModel, src/exts/mlj_ext/core_icnf.jl:19-22)
#                                        and ChainRulesCore.rrule (AutoZygote), both loss methods (with and without ys)
#   MLJModelInterface.fit (row N2)       src/exts/mlj_ext/core_icnf.jl:30-92 -> cnfb200_fit_epochs when the optimisers are Lion
#   PlanarLayer networks (row N3)        src/layers/planar_layer.jl:70-78 -> CNF_FLAG_LAST_LINEAR
#   VectorMode (row N4)                  src/base_icnf.jl:145-165, src/exts/dist_ext/core_icnf.jl:13-31 -> option "vector_mode"
# Multi-GPU: one Julia process per GPU (MPI.jl / Distributed), each with `B200Libs(device = r)`; rank 0 calls
# `comm_unique_id()`, ships the 128 bytes to the others, every rank calls `comm_init!(icnf, nranks, rank, id)`.
# A single process does not drive several devices (the handle owns one device); see INTEGRATION.md "Multi-GPU".
module ContinuousNormalizingFlowsB200Ext

import ContinuousNormalizingFlows as CNF
import ComputationalResources, ChainRulesCore, ComponentArrays, Distributions, EnzymeCore, Lux, LuxCore
import MLJModelInterface, Optimisers, Random
import EnzymeCore: EnzymeRules
import CNFB200: B200Libs   # struct B200Libs <: ComputationalResources.AbstractResource; device::Int; libpath::String; kernel_path::Int; end

const MAXL = 8
const FLAG_COND, FLAG_AUGMENTED, FLAG_NORM_Z, FLAG_NORM_J, FLAG_NORM_Z_AUG, FLAG_LAST_LINEAR =
    UInt32(1), UInt32(2), UInt32(4), UInt32(8), UInt32(16), UInt32(32)

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
    stiff_max::Float32; stiff_run::Int32
    CnfStats() = new(0, 0, 0, 0, 0, 0.0f0, 0.0f0, 0.0, 0.0, 0, 0.0, 0, -1.0f0, 0)
end

const B200ICNF{CM} = CNF.ICNF{
    Float32, CM, <:Any, <:Any, <:Any, <:Any, <:Any, <:Any, <:Any, <:Any, <:Any, <:B200Libs,
}
lib(icnf) = icnf.resource.libpath

trace_mode(::Union{CNF.DIVecJacMatrixMode, CNF.DIVecJacVectorMode}) = Int32(0)
trace_mode(::Union{CNF.DIJacVecMatrixMode, CNF.DIJacVecVectorMode}) = Int32(1)
trace_mode(::Any) = throw(ArgumentError("B200Libs supports the DIVecJac* / DIJacVec* compute modes only"))

# ---- network description --------------------------------------------------------------------------------------------
# Dense(tanh) chain -> layer widths; a PlanarLayer(nvars, tanh; n_cond) is Dense(d0 => 1, tanh) followed by a
# bias-free linear product (CNF_FLAG_LAST_LINEAR, row N3); anything else is rejected (no CPU fallback in this path)
is_tanh(f) = f === tanh || f === Lux.NNlib.tanh_fast
function describe_nn(nn)
    if nn isa CNF.PlanarLayer
        (is_tanh(nn.activation) && nn isa CNF.PlanarLayer{true}) ||
            throw(ArgumentError("B200Libs supports PlanarLayer(nvars, tanh; use_bias = true) only"))
        return Int32[nn.nvars + nn.n_cond, 1, nn.nvars], true
    end
    layers = nn isa Lux.Chain ? collect(values(nn.layers)) : [nn]
    (length(layers) == 1 && first(layers) isa CNF.PlanarLayer) && return describe_nn(first(layers))
    all(l -> l isa Lux.Dense && is_tanh(l.activation) && l.use_bias isa Lux.True, layers) ||
        throw(ArgumentError("B200Libs supports Chain(Dense(in => out, tanh)...) and PlanarLayer(nvars, tanh) only"))
    Int32[first(layers).in_dims; [l.out_dims for l in layers]...], false
end

# per-layer weight/bias offsets inside the flat ComponentArray (0-based, in floats).  Dense: (weight, bias) per layer;
# PlanarLayer: (u[nvars], w[nvars + n_cond], b[1]) (src/layers/planar_layer.jl:38-60)
function offsets(dims, planar)
    if planar
        d0, D = Int(dims[1]), Int(dims[3])
        return Int64[D, 0], Int64[D + d0, 0], D + d0 + 1
    end
    w = Int64[]; b = Int64[]; o = 0
    for l in 1:(length(dims) - 1)
        push!(w, o); o += dims[l] * dims[l + 1]
        push!(b, o); o += dims[l + 1]
    end
    w, b, o
end

# ---- handle cache: one library handle per ICNF object, destroyed by a finalizer ---------------------------------------
mutable struct HandleBox
    ptr::Ptr{Cvoid}
    libpath::String
    function HandleBox(ptr, libpath)
        box = new(ptr, libpath)
        finalizer(box) do b
            b.ptr == C_NULL || ccall((:cnfb200_destroy, b.libpath), Int32, (Ptr{Cvoid},), b.ptr)
            b.ptr = C_NULL
        end
        box
    end
end
const HANDLES = Dict{UInt, HandleBox}()
const HORDER = UInt[]            # least recently created first
const HMAX = 16                  # evicted boxes are destroyed by their finalizer at the next GC
const HLOCK = ReentrantLock()

last_error(libpath, h) = unsafe_string(ccall((:cnfb200_last_error, libpath), Cstring, (Ptr{Cvoid},), h))
version(libpath) = unsafe_string(ccall((:cnfb200_version, libpath), Cstring, ()))

function make_config(icnf::B200ICNF)
    dims, planar = describe_nn(icnf.nn)
    D = icnf.nvars + icnf.naugmented
    C = Int(dims[1]) - D
    kw = icnf.sol_kwargs
    haskey(kw, :alg) && !(nameof(typeof(kw.alg)) === :Tsit5) &&
        throw(ArgumentError("B200Libs implements the Tsit5 branch only"))
    flags = UInt32(0)
    C > 0 && (flags |= FLAG_COND)
    icnf.naugmented != 0 && (flags |= FLAG_AUGMENTED)
    !iszero(icnf.λ₁) && (flags |= FLAG_NORM_Z)
    !iszero(icnf.λ₂) && (flags |= FLAG_NORM_J)
    !iszero(icnf.λ₃) && (flags |= FLAG_NORM_Z_AUG)
    planar && (flags |= FLAG_LAST_LINEAR)
    CnfConfig(D, C, icnf.nvars, length(dims) - 1,
        ntuple(i -> i <= length(dims) ? dims[i] : Int32(0), MAXL + 1),
        trace_mode(icnf.compute_mode), flags, icnf.λ₁, icnf.λ₂, icnf.λ₃,
        get(kw, :reltol, 0.0f0), get(kw, :abstol, 0.0f0), min(get(kw, :maxiters, 0), typemax(Int64) ÷ 2),
        get(kw, :dt, 0.0f0), get(kw, :dtmax, 0.0f0), get(kw, :dtmin, 0.0f0),
        icnf.resource.device, icnf.resource.kernel_path)
end

function handle(icnf::B200ICNF)
    lock(HLOCK) do
        key = hash(make_config(icnf), objectid(icnf))
        box = get(HANDLES, key, nothing)
        if box === nothing
            cfg = make_config(icnf)
            h = Ref{Ptr{Cvoid}}(C_NULL)
            rc = ccall((:cnfb200_create, lib(icnf)), Int32, (Ref{CnfConfig}, Ref{Ptr{Cvoid}}), cfg, h)
            if rc != 0
                msg = h[] == C_NULL ? "cnfb200_create failed" : last_error(lib(icnf), h[])
                h[] == C_NULL || ccall((:cnfb200_destroy, lib(icnf)), Int32, (Ptr{Cvoid},), h[])
                error(msg)
            end
            box = HandleBox(h[], lib(icnf))
            HANDLES[key] = box
            push!(HORDER, key)
            length(HORDER) > HMAX && delete!(HANDLES, popfirst!(HORDER))
            if icnf.compute_mode isa CNF.VectorMode      # row N4: every column its own adaptive solve
                set_option!(icnf, box.ptr, "vector_mode", 1)
            end
        end
        box.ptr
    end
end

"drop the cached handle of `icnf` now (otherwise its finalizer does it)"
function release!(icnf::B200ICNF)
    lock(HLOCK) do
        key = hash(make_config(icnf), objectid(icnf))
        box = pop!(HANDLES, key, nothing)
        filter!(!=(key), HORDER)
        box === nothing || finalize(box)
    end
    nothing
end

function check(icnf, h, rc, what)
    rc < 0 && error(what * ": " * last_error(lib(icnf), h))
    rc > 0 && @warn "$what: solver retcode $rc (1 MaxIters, 2 DtLessThanMin, 3 Unstable)"
    nothing
end

set_option!(icnf, h, key::String, value::Integer) =
    check(icnf, h, ccall((:cnfb200_set_option, lib(icnf)), Int32, (Ptr{Cvoid}, Cstring, Int64), h, key, value), "cnfb200_set_option")

function set_params!(icnf, h, ps)
    θ = convert(Vector{Float32}, collect(ComponentArrays.getdata(ps)))     # flat ComponentArray data (mlj_ext/core_icnf.jl:37-38)
    w, b, n = offsets(describe_nn(icnf.nn)...)
    length(θ) == n || throw(ArgumentError("ps has $(length(θ)) parameters, the network needs $n"))
    GC.@preserve θ w b check(icnf, h, ccall((:cnfb200_set_params, lib(icnf)), Int32,
        (Ptr{Cvoid}, Ptr{Float32}, Int64, Ptr{Int64}, Ptr{Int64}), h, θ, n, w, b), "cnfb200_set_params")
end

# ---- multi-GPU: one process per GPU -----------------------------------------------------------------------------------
function comm_unique_id(libpath::String)
    id = Vector{UInt8}(undef, 128)
    rc = ccall((:cnfb200_comm_unique_id, libpath), Int32, (Ptr{UInt8},), id)
    rc == 0 || error("cnfb200_comm_unique_id failed (rc = $rc)")
    id
end
function comm_init!(icnf::B200ICNF, nranks::Integer, rank::Integer, id::Vector{UInt8})
    h = handle(icnf)
    GC.@preserve id check(icnf, h, ccall((:cnfb200_comm_init, lib(icnf)), Int32, (Ptr{Cvoid}, Int32, Int32, Ptr{UInt8}),
        h, nranks, rank, id), "cnfb200_comm_init")
end
"true when the per-attempt error-norm exchange runs over NVLink peer memory inside the controller kernel"
comm_p2p(icnf::B200ICNF) = ccall((:cnfb200_comm_p2p, lib(icnf)), Int32, (Ptr{Cvoid},), handle(icnf)) != 0

# ---- row N1: device-side draws -----------------------------------------------------------------------------------------
"from now on eps (TrainMode) and z0 (generate) are drawn by the library's Philox stream instead of icnf.rng"
function seed_device_rng!(icnf::B200ICNF, seed::Integer, sample_offset::Integer = 0)
    h = handle(icnf)
    check(icnf, h, ccall((:cnfb200_seed_rng, lib(icnf)), Int32, (Ptr{Cvoid}, UInt64, UInt64), h, seed, sample_offset), "cnfb200_seed_rng")
    DEVICE_RNG[objectid(icnf)] = true
    nothing
end
const DEVICE_RNG = Dict{UInt, Bool}()
device_rng(icnf) = get(DEVICE_RNG, objectid(icnf), false)
function last_eps(icnf::B200ICNF, B::Integer)
    ϵ = Matrix{Float32}(undef, icnf.nvars + icnf.naugmented, B); h = handle(icnf)
    GC.@preserve ϵ check(icnf, h, ccall((:cnfb200_last_eps, lib(icnf)), Int32, (Ptr{Cvoid}, Ptr{Float32}, Int64), h, ϵ, B), "cnfb200_last_eps")
    ϵ
end

# page-locked host matrices for inputs / outputs (copies inside the entry points then run at full PCIe rate)
function pinned_matrix(icnf::B200ICNF, rows::Integer, cols::Integer)
    p = Ref{Ptr{Cvoid}}(C_NULL)
    rc = ccall((:cnfb200_host_alloc, lib(icnf)), Int32, (Csize_t, Ref{Ptr{Cvoid}}), rows * cols * sizeof(Float32), p)
    rc == 0 || error("cnfb200_host_alloc failed (rc = $rc)")
    A = unsafe_wrap(Array, Ptr{Float32}(p[]), (Int(rows), Int(cols)); own = false)
    libpath = lib(icnf)
    finalizer(a -> ccall((:cnfb200_host_free, libpath), Int32, (Ptr{Cvoid},), pointer(a)), A)
    A
end

"sustained FP32-FMA rate of the device in TFLOP/s (roofline denominator of the SIMT kernels; bench only)"
function measure_fp32_peak(icnf::B200ICNF)
    t = Ref{Float64}(0.0)
    ccall((:cnfb200_measure_fp32_peak, lib(icnf)), Int32, (Int32, Ref{Float64}), icnf.resource.device, t)
    t[]
end

# eps ~ epsdist, D x B, once per call on the host (src/base_icnf.jl:277-278); TestMode never reads it
function draw_eps(icnf, train, D, B)
    (!train || device_rng(icnf)) && return nothing
    ϵ = Matrix{Float32}(undef, D, B)
    Random.rand!(icnf.rng, icnf.epsdist, ϵ)
    ϵ
end
ptr_or_null(A) = A === nothing ? Ptr{Float32}(C_NULL) : pointer(A)
as_f32(A) = A === nothing ? nothing : convert(Matrix{Float32}, A)      # column-major rows x B == what the ABI expects

# ---- one evaluation of augmented_f (src/icnf.jl:148-164, :318-350, :384-420): used by tests --------------------------
function augmented_f(icnf::B200ICNF, mode::CNF.Mode, u::AbstractMatrix, ys, ϵ, ps)
    h = handle(icnf); set_params!(icnf, h, ps)
    U = as_f32(u); y = as_f32(ys); e = as_f32(ϵ); du = similar(U)
    GC.@preserve U y e du check(icnf, h, ccall((:cnfb200_augmented_f, lib(icnf)), Int32,
        (Ptr{Cvoid}, Int32, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Int64, Ptr{Float32}),
        h, mode isa CNF.TrainMode ? 1 : 0, U, ptr_or_null(y), ptr_or_null(e), size(U, 2), du), "cnfb200_augmented_f")
    du
end

# the same evaluation without any device allocation inside the call (no reference counterpart; csrc/cnf_noalloc.cu):
# u, ys, ϵ, du are DEVICE pointers (e.g. `pointer(::CuArray)` reinterpreted), (2D + C + R) * round_up(B, 256) floats
# must fit CNFB200_STATIC_SCRATCH_BYTES = 16 MiB
function augmented_f_static!(du::Ptr{Float32}, icnf::B200ICNF, mode::CNF.Mode, u::Ptr{Float32}, ys::Ptr{Float32},
        ϵ::Ptr{Float32}, B::Integer, ps)
    h = handle(icnf); set_params!(icnf, h, ps)
    check(icnf, h, ccall((:cnfb200_augmented_f_static, lib(icnf)), Int32,
        (Ptr{Cvoid}, Int32, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Int64, Ptr{Float32}),
        h, mode isa CNF.TrainMode ? 1 : 0, u, ys, ϵ, B, du), "cnfb200_augmented_f_static")
    du
end

# ---- inference(icnf, mode, xs[, ys], ps, st): replaces base_icnf.jl:407-426 for RESOURCE <: B200Libs --------------------
function _inference(icnf::B200ICNF, mode::CNF.Mode, xs::AbstractMatrix, ys, ps)
    h = handle(icnf); set_params!(icnf, h, ps)
    D = icnf.nvars + icnf.naugmented; B = size(xs, 2)
    train = mode isa CNF.TrainMode
    x = as_f32(xs); y = as_f32(ys); ϵ = draw_eps(icnf, train, D, B)
    t₀, t₁ = CNF.steer_tspan(icnf, mode)                # base_icnf.jl:108-121
    logp = Vector{Float32}(undef, B); regs = Matrix{Float32}(undef, train ? 3 : 1, B); st = CnfStats()
    GC.@preserve x y ϵ logp regs st begin
        rc = ccall((:cnfb200_logdensity, lib(icnf)), Int32,
            (Ptr{Cvoid}, Int32, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Int64, Float32, Float32,
             Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Ref{CnfStats}),
            h, train ? 1 : 0, x, ptr_or_null(y), ptr_or_null(ϵ), B, t₀, t₁, logp, regs, C_NULL, st)
        check(icnf, h, rc, "cnfb200_logdensity")
    end
    (logp, eachrow(regs))                               # same shape as inference_sol's return, base_icnf.jl:188
end
CNF.inference(icnf::B200ICNF{<:CNF.MatrixMode}, mode::CNF.Mode, xs::AbstractMatrix{<:Real}, ps::Any, st::NamedTuple) =
    _inference(icnf, mode, xs, nothing, ps)
CNF.inference(icnf::B200ICNF{<:CNF.MatrixMode}, mode::CNF.Mode, xs::AbstractMatrix{<:Real}, ys::AbstractMatrix{<:Real}, ps::Any, st::NamedTuple) =
    _inference(icnf, mode, xs, ys, ps)

# row N4, VectorMode: a vector is one column; the result is (logp̂x::Real, (Ȧ,)) like inference_sol's vector method
# (base_icnf.jl:145-165).  A matrix of columns goes down in ONE call (every column its own step control) instead of the
# reference's broadcast over eachcol (dist_ext/core_icnf.jl:24-25).
function CNF.inference(icnf::B200ICNF{<:CNF.VectorMode}, mode::CNF.Mode, xs::AbstractVector{<:Real}, ps::Any, st::NamedTuple)
    logp, regs = _inference(icnf, mode, reshape(xs, :, 1), nothing, ps)
    (only(logp), map(only, regs))
end
function Distributions._logpdf(d::CNF.ICNFDist{<:B200ICNF{<:CNF.VectorMode}}, A::AbstractMatrix{<:Real})
    first(_inference(d.m, d.mode, A, nothing, d.ps))
end

# ---- loss and its discrete-adjoint gradient ---------------------------------------------------------------------------
# want_g / want_gx = false pass NULL: the forward solve alone (no step tape, no reverse sweep)
function loss_grad(icnf::B200ICNF, mode::CNF.Mode, xs, ys, ps; want_g::Bool = true, want_gx::Bool = false)
    h = handle(icnf); set_params!(icnf, h, ps)
    D = icnf.nvars + icnf.naugmented; B = size(xs, 2)
    train = mode isa CNF.TrainMode
    x = as_f32(xs); y = as_f32(ys); ϵ = draw_eps(icnf, train, D, B)
    t₀, t₁ = CNF.steer_tspan(icnf, mode)
    ℓ = Ref{Float32}(0)
    g = want_g ? zeros(Float32, length(ComponentArrays.getdata(ps))) : nothing
    gx = want_gx ? similar(x) : nothing
    st = CnfStats()
    GC.@preserve x y ϵ g gx st begin
        rc = ccall((:cnfb200_loss_grad_mode, lib(icnf)), Int32,
            (Ptr{Cvoid}, Int32, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Int64, Float32, Float32, Ref{Float32},
             Ptr{Float32}, Ptr{Float32}, Ref{CnfStats}),
            h, train ? 1 : 0, x, ptr_or_null(y), ptr_or_null(ϵ), B, t₀, t₁, ℓ, ptr_or_null(g), ptr_or_null(gx), st)
        check(icnf, h, rc, "cnfb200_loss_grad_mode")
    end
    ℓ[], g, gx
end
# the TrainMode-only entry point of the header (cnfb200_loss_grad == cnfb200_loss_grad_mode with mode = 1)
function loss_grad_train(icnf::B200ICNF, xs, ys, ϵ, ps, t₀::Float32, t₁::Float32)
    h = handle(icnf); set_params!(icnf, h, ps)
    x = as_f32(xs); y = as_f32(ys); e = as_f32(ϵ)
    ℓ = Ref{Float32}(0); g = zeros(Float32, length(ComponentArrays.getdata(ps))); st = CnfStats()
    GC.@preserve x y e g st check(icnf, h, ccall((:cnfb200_loss_grad, lib(icnf)), Int32,
        (Ptr{Cvoid}, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Int64, Float32, Float32, Ref{Float32}, Ptr{Float32}, Ptr{Float32}, Ref{CnfStats}),
        h, x, ptr_or_null(y), ptr_or_null(e), size(x, 2), t₀, t₁, ℓ, g, C_NULL, st), "cnfb200_loss_grad")
    ℓ[], g
end

const MM = CNF.MatrixMode
# plain loss: forward only (TrainMode: src/icnf.jl:481-502; TestMode: -mean(logp̂x), src/base_icnf.jl:489-508)
CNF.loss(icnf::B200ICNF{<:MM}, mode::CNF.Mode, xs::AbstractMatrix{<:Real}, ps::Any, st::NamedTuple) =
    first(loss_grad(icnf, mode, xs, nothing, ps; want_g = false))
CNF.loss(icnf::B200ICNF{<:MM}, mode::CNF.Mode, xs::AbstractMatrix{<:Real}, ys::AbstractMatrix{<:Real}, ps::Any, st::NamedTuple) =
    first(loss_grad(icnf, mode, xs, ys, ps; want_g = false))

# -- ChainRules (AutoZygote and every ChainRules-based backend)
function ChainRulesCore.rrule(::typeof(CNF.loss), icnf::B200ICNF{<:MM}, mode::CNF.Mode, xs::AbstractMatrix{<:Real}, ps::Any, st::NamedTuple)
    ℓ, g, gx = loss_grad(icnf, mode, xs, nothing, ps; want_gx = true)
    NT = ChainRulesCore.NoTangent()
    back(ℓ̄) = (NT, NT, NT, ℓ̄ .* gx, ℓ̄ .* g, NT)
    ℓ, back
end
function ChainRulesCore.rrule(::typeof(CNF.loss), icnf::B200ICNF{<:MM}, mode::CNF.Mode, xs::AbstractMatrix{<:Real}, ys::AbstractMatrix{<:Real}, ps::Any, st::NamedTuple)
    ℓ, g, gx = loss_grad(icnf, mode, xs, ys, ps; want_gx = true)
    NT = ChainRulesCore.NoTangent()
    back(ℓ̄) = (NT, NT, NT, ℓ̄ .* gx, NT, ℓ̄ .* g, NT)     # ys is a constant input of the flow (CondLayer, src/layers/cond_layer.jl:1-9)
    ℓ, back
end

# -- Enzyme (ICNFModel's default adtype: AutoEnzyme(Reverse, runtime activity), src/exts/mlj_ext/core_icnf.jl:19-22;
#    test/call_tests.jl:239-251 differentiates w.r.t. ps and w.r.t. xs through DifferentiationInterface.gradient).
#    The rule sits on `loss` itself, so Enzyme never sees the ccall.  The gradient is computed in the augmented primal
#    (one forward solve + one reverse sweep) and kept on the tape; `reverse` scales it by the incoming cotangent.
function _enzyme_aug(config, icnf, mode, xs, ys, ps)
    want_gx = !(xs isa EnzymeCore.Const)
    ℓ, g, gx = loss_grad(icnf.val, mode.val, xs.val, ys === nothing ? nothing : ys.val, ps.val;
        want_g = !(ps isa EnzymeCore.Const), want_gx = want_gx)
    primal = EnzymeRules.needs_primal(config) ? ℓ : nothing
    EnzymeRules.AugmentedReturn(primal, nothing, (g, gx))
end
function _enzyme_rev(dret, tape, xs, ps)
    g, gx = tape
    if ps isa EnzymeCore.Duplicated || ps isa EnzymeCore.DuplicatedNoNeed
        ComponentArrays.getdata(ps.dval) .+= dret.val .* g
    end
    if xs isa EnzymeCore.Duplicated || xs isa EnzymeCore.DuplicatedNoNeed
        xs.dval .+= dret.val .* gx
    end
    nothing
end
function EnzymeRules.augmented_primal(config::EnzymeRules.RevConfig, ::EnzymeCore.Const{typeof(CNF.loss)}, ::Type{<:EnzymeCore.Active},
        icnf::EnzymeCore.Const{<:B200ICNF{<:MM}}, mode::EnzymeCore.Const{<:CNF.Mode}, xs::EnzymeCore.Annotation{<:AbstractMatrix{<:Real}},
        ps::EnzymeCore.Annotation, st::EnzymeCore.Const{<:NamedTuple})
    _enzyme_aug(config, icnf, mode, xs, nothing, ps)
end
function EnzymeRules.reverse(config::EnzymeRules.RevConfig, ::EnzymeCore.Const{typeof(CNF.loss)}, dret::EnzymeCore.Active, tape,
        icnf::EnzymeCore.Const{<:B200ICNF{<:MM}}, mode::EnzymeCore.Const{<:CNF.Mode}, xs::EnzymeCore.Annotation{<:AbstractMatrix{<:Real}},
        ps::EnzymeCore.Annotation, st::EnzymeCore.Const{<:NamedTuple})
    _enzyme_rev(dret, tape, xs, ps)
    (nothing, nothing, nothing, nothing, nothing)
end
function EnzymeRules.augmented_primal(config::EnzymeRules.RevConfig, ::EnzymeCore.Const{typeof(CNF.loss)}, ::Type{<:EnzymeCore.Active},
        icnf::EnzymeCore.Const{<:B200ICNF{<:MM}}, mode::EnzymeCore.Const{<:CNF.Mode}, xs::EnzymeCore.Annotation{<:AbstractMatrix{<:Real}},
        ys::EnzymeCore.Annotation{<:AbstractMatrix{<:Real}}, ps::EnzymeCore.Annotation, st::EnzymeCore.Const{<:NamedTuple})
    _enzyme_aug(config, icnf, mode, xs, ys, ps)
end
function EnzymeRules.reverse(config::EnzymeRules.RevConfig, ::EnzymeCore.Const{typeof(CNF.loss)}, dret::EnzymeCore.Active, tape,
        icnf::EnzymeCore.Const{<:B200ICNF{<:MM}}, mode::EnzymeCore.Const{<:CNF.Mode}, xs::EnzymeCore.Annotation{<:AbstractMatrix{<:Real}},
        ys::EnzymeCore.Annotation{<:AbstractMatrix{<:Real}}, ps::EnzymeCore.Annotation, st::EnzymeCore.Const{<:NamedTuple})
    _enzyme_rev(dret, tape, xs, ps)
    (nothing, nothing, nothing, nothing, nothing, nothing)
end

# ---- generate(icnf, mode, [ys,] ps, st, n): replaces base_icnf.jl:447-466 -----------------------------------------------
function _generate(icnf::B200ICNF, mode::CNF.Mode, ys, ps, n::Int)
    h = handle(icnf); set_params!(icnf, h, ps)
    D = icnf.nvars + icnf.naugmented; train = mode isa CNF.TrainMode
    z0 = nothing
    if !device_rng(icnf)
        z0 = Matrix{Float32}(undef, D, n); Random.rand!(icnf.rng, icnf.basedist, z0)      # base_icnf.jl:367-368
    end
    ϵ = draw_eps(icnf, train, D, n); y = as_f32(ys)
    t₀, t₁ = CNF.steer_tspan(icnf, mode)
    out = Matrix{Float32}(undef, icnf.nvars, n); st = CnfStats()
    GC.@preserve z0 ϵ y out st begin
        rc = ccall((:cnfb200_generate, lib(icnf)), Int32,
            (Ptr{Cvoid}, Int32, Ptr{Float32}, Ptr{Float32}, Ptr{Float32}, Int64, Float32, Float32, Ptr{Float32}, Ref{CnfStats}),
            h, train ? 1 : 0, ptr_or_null(z0), ptr_or_null(y), ptr_or_null(ϵ), n, t₁, t₀, out, st)      # reversed tspan, base_icnf.jl:377
        check(icnf, h, rc, "cnfb200_generate")
    end
    out
end
CNF.generate(icnf::B200ICNF{<:MM}, mode::CNF.Mode, ps::Any, st::NamedTuple, n::Int) = _generate(icnf, mode, nothing, ps, n)
CNF.generate(icnf::B200ICNF{<:MM}, mode::CNF.Mode, ys::AbstractMatrix{<:Real}, ps::Any, st::NamedTuple, n::Int) = _generate(icnf, mode, ys, ps, n)

# ---- row N2: MLJModelInterface.fit as one library call ---------------------------------------------------------------------
# The reference's fit (src/exts/mlj_ext/core_icnf.jl:30-92) runs DataLoader(shuffle = true, partial = true) x n_epochs of
# loss + gradient + optimiser step.  With Optimisers.Lion optimisers (the default) the whole loop runs inside
# cnfb200_fit_epochs: data, theta, momentum and loss history stay on the device; eps and the steered t1 are drawn per
# mini-batch by the library (cnfb200_seed_rng, cnfb200_set_steer).  Other optimisers take the reference's own loop, whose
# loss calls land on the methods above.
lion_params(o::Optimisers.Lion) = (Float32(o.eta), Float32(o.beta[1]), Float32(o.beta[2]))
function fit_epochs!(icnf::B200ICNF, x::AbstractMatrix, y, ps, batch_size::Integer, n_epochs::Integer, opt::Optimisers.Lion;
        steer_rate::Real = 0)
    h = handle(icnf); set_params!(icnf, h, ps)
    device_rng(icnf) || seed_device_rng!(icnf, rand(icnf.rng, UInt64))
    check(icnf, h, ccall((:cnfb200_set_steer, lib(icnf)), Int32, (Ptr{Cvoid}, Float32), h, steer_rate), "cnfb200_set_steer")
    X = as_f32(x); Y = as_f32(y)
    η, β₁, β₂ = lion_params(opt)
    t₀, t₁ = icnf.tspan
    θ = Vector{Float32}(undef, length(ComponentArrays.getdata(ps))); hist = Vector{Float32}(undef, n_epochs); st = CnfStats()
    GC.@preserve X Y θ hist st begin
        rc = ccall((:cnfb200_fit_epochs, lib(icnf)), Int32,
            (Ptr{Cvoid}, Ptr{Float32}, Ptr{Float32}, Int64, Int64, Int32, Ptr{Int64}, Float32, Float32, Float32, Float32, Float32,
             Ptr{Float32}, Ptr{Float32}, Ref{CnfStats}),
            h, X, ptr_or_null(Y), size(X, 2), batch_size, n_epochs, C_NULL, t₀, t₁, η, β₁, β₂, θ, hist, st)
        check(icnf, h, rc, "cnfb200_fit_epochs")
    end
    ComponentArrays.getdata(ps) .= θ
    hist
end
steer_rate_of(icnf) = icnf.steerdist isa Distributions.Uniform ? Float32(maximum(icnf.steerdist)) : 0.0f0
function MLJModelInterface.fit(model::CNF.ICNFModel{<:B200ICNF{<:MM}}, verbosity, X)
    all(o -> o isa Optimisers.Lion, model.optimizers) ||
        return invoke(MLJModelInterface.fit, Tuple{CNF.ICNFModel, Any, Any}, model, verbosity, X)
    x = collect(transpose(MLJModelInterface.matrix(X)))
    ps, st = LuxCore.setup(model.m.rng, model.m)
    ps = ComponentArrays.ComponentArray(ps)
    bs = model.use_batch ? model.batch_size : size(x, 2)
    tst = @timed for opt in model.optimizers
        fit_epochs!(model.m, x, nothing, ps, bs, model.n_epochs, opt; steer_rate = steer_rate_of(model.m))
    end
    ((ps, st), nothing, (stats = tst,))
end
function MLJModelInterface.fit(model::CNF.CondICNFModel{<:B200ICNF{<:MM}}, verbosity, XY)
    all(o -> o isa Optimisers.Lion, model.optimizers) ||
        return invoke(MLJModelInterface.fit, Tuple{CNF.CondICNFModel, Any, Any}, model, verbosity, XY)
    X, Y = XY
    x = collect(transpose(MLJModelInterface.matrix(X))); y = collect(transpose(MLJModelInterface.matrix(Y)))
    ps, st = LuxCore.setup(model.m.rng, model.m)
    ps = ComponentArrays.ComponentArray(ps)
    bs = model.use_batch ? model.batch_size : size(x, 2)
    tst = @timed for opt in model.optimizers
        fit_epochs!(model.m, x, y, ps, bs, model.n_epochs, opt; steer_rate = steer_rate_of(model.m))
    end
    ((ps, st), nothing, (stats = tst,))
end

# rng_AT / base_AT stay the host defaults (base_icnf.jl:123-135): MLJ transform and ICNFDist keep working unchanged
# because host arrays stay host arrays.
end
