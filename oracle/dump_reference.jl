# dump_reference.jl — run the REAL ContinuousNormalizingFlows.jl (CPU) on the golden configs and
# write fixtures in the layout of tests/golden/*.npz.  Cannot be executed in the build environment
# (no Julia); provided so that anyone with Julia can close "parity unpinned".
#   julia --project=<reference> oracle/dump_reference.jl tests/golden/inputs_dir out_dir
# Inputs (xs, ys, eps, theta as .npy, column-major = the bytes of the committed fixtures) are read
# so that inputs, weights and eps draws are identical to what the CUDA path and the oracle saw.
using ContinuousNormalizingFlows, Lux, ComponentArrays, NPZ, SciMLBase, ADTypes, Enzyme, DifferentiationInterface

const CASES = [  # (id, nvars, naug, C, dims, tspan, trace, λ1, λ2, λ3, mode, reltol, abstol)
    ("C0", 8, 8, 0, [16, 16], (0.0f0, 13.0f0), :vjp, 1f-2, 1f-2, 1f-2, :train, nothing, nothing),
    ("C1", 1, 1, 0, [2, 6, 2], (0.0f0, 13.0f0), :vjp, 1f-2, 1f-2, 1f-2, :train, sqrt(eps(Float32)), eps(Float32)),
    ("C2", 2, 0, 0, [2, 64, 64, 2], (0.0f0, 1.0f0), :jvp, 0f0, 0f0, 0f0, :test, nothing, nothing),
    ("C3", 16, 16, 0, [32, 256, 256, 32], (0.0f0, 1.0f0), :vjp, 1f-2, 1f-2, 0f0, :train, nothing, nothing),
    ("C4", 8, 0, 8, [16, 128, 128, 8], (0.0f0, 1.0f0), :vjp, 1f-2, 1f-2, 0f0, :train, nothing, nothing),
]

# an rng whose rand!(rng, MvNormal, ϵ) returns the stored eps, so the reference consumes OUR draws
mutable struct ReplayEps <: Random.AbstractRNG; eps::Matrix{Float32}; end
# (left to the runner: overload Random.rand!(::ReplayEps, ::MvNormal, A) = copyto!(A, rng.eps))

function run_case(dir, out, (id, nv, na, C, dims, tspan, trace, l1, l2, l3, mode, rt, at))
    xs = npzread(joinpath(dir, "$(id)_xs.npy")); eps = npzread(joinpath(dir, "$(id)_eps.npy"))
    theta = npzread(joinpath(dir, "$(id)_theta.npy"))
    nn = Chain([Dense(dims[i] => dims[i + 1], tanh) for i in 1:(length(dims) - 1)]...)
    ad = AutoEnzyme(; mode = Enzyme.set_runtime_activity(Enzyme.Reverse), function_annotation = Enzyme.Const)
    cm = trace === :vjp ? DIVecJacMatrixMode(ad) : DIJacVecMatrixMode(AutoEnzyme(; mode = Enzyme.set_runtime_activity(Enzyme.Forward), function_annotation = Enzyme.Const))
    kw = rt === nothing ? (; alg = Tsit5()) : (; alg = Tsit5(), reltol = rt, abstol = at)   # pin the Tsit5 branch
    tag = C > 0 ? CondRNODE : RNODE
    icnf = construct(tag, nn, nv, na; compute_mode = cm, tspan, λ₁ = l1, λ₂ = l2, λ₃ = l3, sol_kwargs = kw, rng = ReplayEps(eps))
    ps, st = Lux.setup(icnf.rng, icnf); ps = ComponentArray(ps); ps .= theta
    m = mode === :train ? TrainMode() : TestMode()
    logp, regs = C > 0 ? inference(icnf, m, xs, npzread(joinpath(dir, "$(id)_ys.npy")), ps, st) : inference(icnf, m, xs, ps, st)
    npzwrite(joinpath(out, "$(id)_ref.npz"), Dict("logp_ref" => collect(logp), "regs_ref" => reduce(hcat, collect.(regs))))
end

foreach(c -> run_case(ARGS[1], ARGS[2], c), CASES)
