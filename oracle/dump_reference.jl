# dump_reference.jl — run the REAL ContinuousNormalizingFlows.jl (CPU) on the golden configs and write
# `<case>_ref.npz` fixtures that tests/test_golden.py consumes when present (closes "parity unpinned").
# TEST INFRASTRUCTURE.  It cannot be executed in the build environment (no Julia); it is written to run as is:
#
#   python tests/golden/export_inputs.py tests/golden/ref_inputs          # *_xs/_ys/_eps/_theta.npy from the fixtures
#   julia --project=<reference checkout> oracle/dump_reference.jl tests/golden/ref_inputs tests/golden
#
# Needs, next to the reference's own dependencies: NPZ, OrdinaryDiffEqTsit5 (both registered packages).
# Per case it writes: logp_ref, regs_ref (inference), loss_ref, grad_theta_ref, grad_x_ref (DifferentiationInterface
# gradient with the package's default AutoEnzyme adtype), naccept_ref / nreject_ref (sol.stats of the same problem solved
# through `inference_prob`), and — for the solve WITHOUT `alg`, which is what the package does (src/base_icnf.jl:141) —
# alg_default_ref: the algorithm choices of the default composite solver (all 1 = it stayed on Tsit5, SURVEY A.1) and
# logp_default_ref.  Inputs, weights and eps draws are the bytes the CUDA path and the oracle saw.
import ContinuousNormalizingFlows as CNF
import ADTypes, ComponentArrays, DifferentiationInterface, Distributions, Enzyme, Lux, NPZ, OrdinaryDiffEqTsit5, Random, SciMLBase

# an RNG that replays a stored eps matrix: inference_prob draws eps with rand!(icnf.rng, icnf.epsdist, ϵ)
# (src/base_icnf.jl:277-278); every draw of a D x B matrix returns OUR eps, so the reference integrates the same field
mutable struct ReplayEps <: Random.AbstractRNG
    eps::Matrix{Float32}
end
function Random.rand!(rng::ReplayEps, ::Distributions.Sampleable, A::AbstractMatrix)
    size(A) == size(rng.eps) || error("ReplayEps: asked for $(size(A)), holds $(size(rng.eps))")
    copyto!(A, rng.eps)
end
Random.rand!(rng::ReplayEps, A::AbstractMatrix) = copyto!(A, rng.eps)
Random.rand(::ReplayEps, ::Distributions.Uniform{T}) where {T} = zero(T)          # steer draw: r = 0 (parity runs)

# (id, B, mode, trace, nvars, naug, C, dims, tspan, λ₁, λ₂, λ₃, reltol, abstol) — the cases of tests/golden/make_golden.py
const CASES = [
    ("C0", 16, :train, :vjp, 8, 8, 0, [16, 16], (0.0f0, 13.0f0), 1.0f-2, 1.0f-2, 1.0f-2, nothing, nothing),
    ("C1", 64, :train, :vjp, 1, 1, 0, [2, 6, 2], (0.0f0, 13.0f0), 1.0f-2, 1.0f-2, 1.0f-2, sqrt(eps(Float32)), eps(Float32)),
    ("C2", 128, :test, :jvp, 2, 0, 0, [2, 64, 64, 2], (0.0f0, 1.0f0), 0.0f0, 0.0f0, 0.0f0, nothing, nothing),
    ("C3", 8, :train, :vjp, 16, 16, 0, [32, 256, 256, 32], (0.0f0, 1.0f0), 1.0f-2, 1.0f-2, 0.0f0, nothing, nothing),
    ("C4", 32, :train, :vjp, 8, 0, 8, [16, 128, 128, 8], (0.0f0, 1.0f0), 1.0f-2, 1.0f-2, 0.0f0, nothing, nothing),
    ("C4", 32, :test, :jvp, 8, 0, 8, [16, 128, 128, 8], (0.0f0, 1.0f0), 1.0f-2, 1.0f-2, 0.0f0, nothing, nothing),
]
case_name(id, B, mode, trace) = "$(id)_B$(B)_mode$(mode === :train ? 1 : 0)_tm$(trace === :jvp ? 1 : 0)"

function build(case, eps, alg)
    id, B, mode, trace, nv, na, C, dims, tspan, l1, l2, l3, rt, at = case
    nn = Lux.Chain([Lux.Dense(dims[i] => dims[i + 1], tanh) for i in 1:(length(dims) - 1)]...)
    rev = ADTypes.AutoEnzyme(; mode = Enzyme.set_runtime_activity(Enzyme.Reverse), function_annotation = Enzyme.Const)
    fwd = ADTypes.AutoEnzyme(; mode = Enzyme.set_runtime_activity(Enzyme.Forward), function_annotation = Enzyme.Const)
    cm = trace === :vjp ? CNF.DIVecJacMatrixMode(rev) : CNF.DIJacVecMatrixMode(fwd)
    kw = rt === nothing ? (;) : (; reltol = rt, abstol = at)
    alg === nothing || (kw = merge(kw, (; alg = alg)))
    tag = C > 0 ? CNF.CondRNODE : CNF.RNODE
    CNF.construct(tag, nn, nv, na; compute_mode = cm, tspan = tspan, λ₁ = l1, λ₂ = l2, λ₃ = l3, sol_kwargs = kw,
        rng = ReplayEps(eps))
end

function run_case(dir, out, case)
    id, B, mode, trace, nv, na, C = case
    name = case_name(id, B, mode, trace)
    xs = NPZ.npzread(joinpath(dir, "$(name)_xs.npy"))            # nvars x B (column = sample)
    eps = NPZ.npzread(joinpath(dir, "$(name)_eps.npy"))          # D x B
    theta = NPZ.npzread(joinpath(dir, "$(name)_theta.npy"))
    ys = C > 0 ? NPZ.npzread(joinpath(dir, "$(name)_ys.npy")) : nothing
    m = mode === :train ? CNF.TrainMode() : CNF.TestMode()
    res = Dict{String, Any}()
    for (tagname, alg) in (("", OrdinaryDiffEqTsit5.Tsit5()), ("_default", nothing))
        icnf = build(case, eps, alg)
        ps, st = Lux.setup(Random.default_rng(), icnf)
        ps = ComponentArrays.ComponentArray(ps)
        length(ps) == length(theta) || error("$name: theta has $(length(theta)) entries, the model $(length(ps))")
        ComponentArrays.getdata(ps) .= theta
        args = ys === nothing ? (xs,) : (xs, ys)
        logp, regs = CNF.inference(icnf, m, args..., ps, st)
        res["logp$(tagname)_ref"] = collect(Float32, logp)
        # the same problem solved by hand for the solver statistics (src/base_icnf.jl:137-143 keeps only the last state)
        prob = CNF.inference_prob(icnf, m, args..., ps, st)
        sol = SciMLBase.solve(prob; icnf.sol_kwargs...)
        if alg === nothing
            res["alg_default_ref"] = hasproperty(sol, :alg_choice) && sol.alg_choice !== nothing ? collect(Int32, sol.alg_choice) : Int32[0]
        else
            res["regs_ref"] = reduce(hcat, collect.(regs))       # B x (3 | 1)
            res["naccept_ref"] = Int64(sol.stats.naccept)
            res["nreject_ref"] = Int64(sol.stats.nreject)
            res["nf_ref"] = Int64(sol.stats.nf)
            res["loss_ref"] = Float32(CNF.loss(icnf, m, args..., ps, st))
            ad = ADTypes.AutoEnzyme(; mode = Enzyme.set_runtime_activity(Enzyme.Reverse), function_annotation = Enzyme.Const)
            lθ = ys === nothing ? (p -> CNF.loss(icnf, m, xs, p, st)) : (p -> CNF.loss(icnf, m, xs, ys, p, st))
            lx = ys === nothing ? (x -> CNF.loss(icnf, m, x, ps, st)) : (x -> CNF.loss(icnf, m, x, ys, ps, st))
            res["grad_theta_ref"] = collect(Float32, ComponentArrays.getdata(DifferentiationInterface.gradient(lθ, ad, ps)))
            res["grad_x_ref"] = collect(Float32, DifferentiationInterface.gradient(lx, ad, xs))       # nvars x B
        end
    end
    NPZ.npzwrite(joinpath(out, "$(name)_ref.npz"), res)
    @info "wrote $(name)_ref.npz" naccept = res["naccept_ref"] stayed_on_tsit5 = all(==(1), res["alg_default_ref"])
end

length(ARGS) == 2 || error("usage: julia dump_reference.jl <inputs dir> <out dir>")
foreach(c -> run_case(ARGS[1], ARGS[2], c), CASES)
