"""CPU tests of the oracle itself.  The reference holds no golden vector for
this path (parity unpinned), so the oracle is pinned by closed-form answers, by an independent
float64 restatement (oracle/ref_f64.py), by finite differences, and by committed fixtures."""
import math

import numpy as np
import pytest

from helpers import O, oracle_cfg, relerr, relnorm, wl
from oracle import ref_f64 as F


def test_k1_zero_weights_closed_form():
    """W = 0: zdot = tanh(b) constant, J = 0 -> z(t1) = x + T tanh(b), dlogp = 0,
    E = T ||tanh b||, n = 0; EEst is rounding noise so every step is accepted and dt grows fast."""
    dims = [3, 5, 3]
    cfg = O.make_config(2, 1, dims, lambda1=1e-2, lambda2=1e-2)
    w, b, n = O.offsets(dims)
    theta = np.zeros(n, np.float32)
    rng = np.random.default_rng(0)
    theta[b[1]:b[1] + 3] = rng.normal(0, 0.5, 3)
    xs = rng.normal(0, 1, (37, 2)).astype(np.float32)
    eps = rng.normal(0, 1, (37, 3)).astype(np.float32)
    T = 2.0
    lp, regs, z, st = O.logdensity(cfg, theta, 1, xs, None, eps, 0.0, T)
    tb = np.tanh(theta[b[1]:b[1] + 3].astype(np.float64))
    zref = np.concatenate([xs, np.zeros((37, 1))], 1) + T * tb
    assert st["nreject"] == 0 and st["retcode"] == 0
    assert np.allclose(z, zref, atol=2e-6)
    assert np.allclose(regs[:, 0], T * np.linalg.norm(tb), rtol=1e-5)
    assert np.all(regs[:, 1] == 0)
    logpz = -0.5 * (zref ** 2).sum(1) - 1.5 * math.log(2 * math.pi)
    assert np.allclose(lp, logpz, rtol=2e-6, atol=2e-6)
    dts = O.last_dts()
    for a, bb in zip(dts[:-2], dts[1:-1]):
        assert bb / a > 2.0   # EEst is rounding-level, so dt grows fast (up to qmax = 10)


def test_k2_linear_regime_matches_expm():
    """tiny weights: tanh is linear, z' = A z + c with A = W2 W1; dlogp = -tr(A) T"""
    dims = [2, 4, 2]
    cfg = O.make_config(2, 0, dims, reltol=1e-6, abstol=1e-9)
    w, b, n = O.offsets(dims)
    rng = np.random.default_rng(1)
    theta = (rng.normal(0, 1, n) * 1e-2).astype(np.float32)
    theta[b[0]:b[0] + 4] = 0
    theta[b[1]:b[1] + 2] = 0
    W1 = theta[w[0]:w[0] + 8].reshape(2, 4).T.astype(np.float64)
    W2 = theta[w[1]:w[1] + 8].reshape(4, 2).T.astype(np.float64)
    A = W2 @ W1
    xs = rng.normal(0, 0.1, (16, 2)).astype(np.float32)
    lp, regs, z, st = O.logdensity(cfg, theta, 0, xs, None, None, 0.0, 1.0)
    from scipy.linalg import expm
    zref = xs.astype(np.float64) @ expm(A).T
    assert np.allclose(z, zref, atol=5e-6)
    dlogp = (-0.5 * (z.astype(np.float64) ** 2).sum(1) - math.log(2 * math.pi)) - lp
    assert np.allclose(dlogp, -np.trace(A), atol=5e-6)


def test_k3_scalar_hutchinson_equals_exact():
    """D = 1, eps = +-1: eps J eps = J, so TrainMode's l-row equals TestMode's and VJP == JVP norms"""
    dims = [1, 7, 1]
    rng = np.random.default_rng(2)
    w, b, n = O.offsets(dims)
    theta = rng.normal(0, 0.7, n).astype(np.float32)
    B = 33
    u_test = rng.normal(0, 1, (B, 2)).astype(np.float32)
    u_train = np.concatenate([u_test, np.zeros((B, 2), np.float32)], 1)
    eps = rng.choice([-1.0, 1.0], (B, 1)).astype(np.float32)
    out = {}
    for tm in (0, 1):
        cfg = O.make_config(1, 0, dims, trace_mode=tm, lambda1=1e-2, lambda2=1e-2)
        out[tm, 0] = O.augmented_f(cfg, theta, 0, u_test, None, None)
        out[tm, 1] = O.augmented_f(cfg, theta, 1, u_train, None, eps)
        assert np.allclose(out[tm, 0][:, :2], out[tm, 1][:, :2], rtol=1e-6, atol=1e-7)
    assert np.allclose(out[0, 1], out[1, 1], rtol=1e-6, atol=1e-7)


@pytest.mark.parametrize("cid,B", [("C0", 32), ("C2", 64), ("C4", 32)])
def test_k4_exact_trace_vjp_equals_jvp(cid, B):
    inp = wl.make_inputs(cid, B=B)
    D = wl.CONFIGS[cid]["nvars"] + wl.CONFIGS[cid]["naug"]
    u = np.random.default_rng(3).normal(0, 1, (B, D + 1)).astype(np.float32)
    a = O.augmented_f(oracle_cfg(cid, trace_mode=0), inp["theta"], 0, u, inp["ys"], None)
    b = O.augmented_f(oracle_cfg(cid, trace_mode=1), inp["theta"], 0, u, inp["ys"], None)
    assert np.allclose(a, b, rtol=1e-5, atol=2e-6)


@pytest.mark.parametrize("cid,B,mode", [("C1", 64, 1), ("C2", 128, 0), ("C3", 16, 1), ("C4", 32, 1), ("C4", 32, 0)])
def test_k5_fp32_oracle_vs_float64_restatement(cid, B, mode):
    """two structurally different implementations; the f64 one replays the fp32 run's accepted steps,
    so the comparison isolates arithmetic from step-size decisions"""
    c = wl.CONFIGS[cid]
    inp = wl.make_inputs(cid, B=B)
    cfg = oracle_cfg(cid)
    lp, regs, z, st = O.logdensity(cfg, inp["theta"], mode, inp["xs"], inp["ys"], inp["eps"], *c["tspan"])
    dts = [float(d) for d in O.last_dts()]
    r = F.run(cfg, inp["theta"], mode, inp["xs"], inp["ys"], inp["eps"], *c["tspan"], frozen_dts=dts)
    assert relerr(lp, r["logp"]) < 2e-5
    assert relnorm(z, r["z"]) < 2e-5
    assert np.allclose(regs, r["regs"], rtol=1e-4, atol=1e-5)
    # and the adaptive f64 solve lands within the solver tolerance of it (true-solution sanity)
    r2 = F.run(cfg, inp["theta"], mode, inp["xs"], inp["ys"], inp["eps"], *c["tspan"])
    assert relerr(lp, r2["logp"]) < 2e-2


@pytest.mark.parametrize("cid,B,tm", [("C0", 16, 0), ("C0", 16, 1), ("C1", 64, 0), ("C4", 16, 0), ("C4", 16, 1)])
def test_k6_discrete_gradient_vs_float64_autograd(cid, B, tm):
    c = wl.CONFIGS[cid]
    inp = wl.make_inputs(cid, B=B)
    cfg = oracle_cfg(cid, trace_mode=tm, lambda3=(1e-2 if c["naug"] else 0.0))
    val, g, gx, st = O.loss_grad(cfg, inp["theta"], inp["xs"], inp["ys"], inp["eps"], *c["tspan"])
    dts = [float(d) for d in O.last_dts()]
    r = F.run(cfg, inp["theta"], 1, inp["xs"], inp["ys"], inp["eps"], *c["tspan"], frozen_dts=dts,
              requires_grad=True)
    assert abs(val - r["loss"]) / abs(r["loss"]) < 1e-5
    assert relnorm(g, r["grad_theta"]) < 1e-4
    assert relnorm(gx, r["grad_x"]) < 1e-4


@pytest.mark.parametrize("cid,B,tm", [("C2", 16, 1), ("C2", 16, 0), ("C4", 8, 0)])
def test_k6c_testmode_loss_gradient_vs_float64_autograd(cid, B, tm):
    """loss(icnf, TestMode(), ...) = -mean(logp) (reference src/base_icnf.jl:489-508) is differentiated by
    the reference's tests too (test/call_tests.jl:239-251): exact-trace reverse sweep vs torch autograd"""
    c = wl.CONFIGS[cid]
    inp = wl.make_inputs(cid, B=B)
    cfg = oracle_cfg(cid, trace_mode=tm)
    val, g, gx, st = O.loss_grad(cfg, inp["theta"], inp["xs"], inp["ys"], None, *c["tspan"], mode=0)
    dts = [float(d) for d in O.last_dts()]
    r = F.run(cfg, inp["theta"], 0, inp["xs"], inp["ys"], None, *c["tspan"], frozen_dts=dts, requires_grad=True)
    assert abs(val - r["loss"]) / abs(r["loss"]) < 1e-5
    assert relnorm(g, r["grad_theta"]) < 1e-4
    assert relnorm(gx, r["grad_x"]) < 1e-4


def test_k6b_float64_gradient_vs_finite_differences():
    cid, B = "C1", 8
    c = wl.CONFIGS[cid]
    inp = wl.make_inputs(cid, B=B)
    cfg = oracle_cfg(cid)
    O.loss_grad(cfg, inp["theta"], inp["xs"], None, inp["eps"], *c["tspan"], want_grad=False)
    dts = [float(d) for d in O.last_dts()]
    th = inp["theta"].astype(np.float64)
    r = F.run(cfg, th, 1, inp["xs"], None, inp["eps"], *c["tspan"], frozen_dts=dts, requires_grad=True)
    h = 1e-6
    for k in (0, 5, 13, 20, 31):
        tp, tm_ = th.copy(), th.copy()
        tp[k] += h
        tm_[k] -= h
        fp = F.run(cfg, tp, 1, inp["xs"], None, inp["eps"], *c["tspan"], frozen_dts=dts)["loss"]
        fm = F.run(cfg, tm_, 1, inp["xs"], None, inp["eps"], *c["tspan"], frozen_dts=dts)["loss"]
        fd = (fp - fm) / (2 * h)
        assert abs(fd - r["grad_theta"][k]) < 1e-5 * max(1.0, abs(fd))


def test_k7_generate_inverts_inference():
    cid, B = "C2", 256
    c = wl.CONFIGS[cid]
    inp = wl.make_inputs(cid, B=B)
    cfg = oracle_cfg(cid)
    lp, regs, z, st = O.logdensity(cfg, inp["theta"], 0, inp["xs"], None, None, 0.0, 1.0)
    xr, st2 = O.generate(cfg, inp["theta"], 0, z, None, None, 1.0, 0.0)
    assert np.max(np.abs(xr - inp["xs"])) < 2e-2


def test_k8_train_and_test_rows_differ_only_through_steps():
    """same z dynamics; the two extra rows enter the whole-matrix RMS norm (F8) so the step
    sequences differ, the z solution agrees to solver tolerance"""
    cid, B = "C4", 64
    c = wl.CONFIGS[cid]
    inp = wl.make_inputs(cid, B=B)
    cfg = oracle_cfg(cid)
    _, _, z0, st0 = O.logdensity(cfg, inp["theta"], 0, inp["xs"], inp["ys"], None, 0.0, 1.0)
    d0 = O.last_dts()
    _, _, z1, st1 = O.logdensity(cfg, inp["theta"], 1, inp["xs"], inp["ys"], inp["eps"], 0.0, 1.0)
    d1 = O.last_dts()
    assert len(d0) != len(d1) or not np.array_equal(d0, d1)
    assert np.max(np.abs(z0 - z1)) < 5e-3


def test_empty_and_ragged_batches():
    cid = "C1"
    c = wl.CONFIGS[cid]
    cfg = oracle_cfg(cid)
    full = wl.make_inputs(cid, B=40)
    for B in (1, 15, 16, 17, 33):   # around the 16-sample tile
        lp, _, _, st = O.logdensity(cfg, full["theta"], 1, full["xs"][:B], None, full["eps"][:B], *c["tspan"])
        assert lp.shape == (B,) and np.all(np.isfinite(lp)) and st["retcode"] == 0


def test_reverse_time_and_maxiters():
    cid = "C1"
    inp = wl.make_inputs(cid, B=8)
    cfg = oracle_cfg(cid)
    cfg.maxiters = 3
    lp, _, _, st = O.logdensity(cfg, inp["theta"], 0, inp["xs"], None, None, 0.0, 13.0)
    assert st["retcode"] == 1 and st["naccept"] + st["nreject"] == 3


def test_n1_philox_known_answers_and_normal_moments():
    """N1: the generator behind cnfb200_seed_rng.  Known-answer vectors of Philox4x32-10 from the
    Random123 distribution (kat_vectors: 'philox4x32 10 ...'); then the Box-Muller output is a
    standard normal matrix whose draws / rows / samples are distinct streams."""
    assert O.philox4x32_10([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert O.philox4x32_10([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert O.philox4x32_10([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]
    n = O.philox_normal(99, 0, 0, 6, 100000)
    assert n.shape == (6, 100000) and np.isfinite(n).all()
    assert abs(n.mean()) < 5e-3 and abs(n.std() - 1) < 5e-3 and abs((n ** 4).mean() - 3) < 5e-2
    c = np.corrcoef(n)
    assert np.abs(c - np.eye(6)).max() < 1.5e-2
    # sharding: columns [b0, b0 + m) of the global draw = the shard's own draw with sample_offset = b0
    np.testing.assert_array_equal(O.philox_normal(99, 1000, 0, 6, 500), n[:, 1000:1500])
    assert not np.array_equal(O.philox_normal(99, 0, 1, 6, 500), n[:, :500])      # next draw
    assert not np.array_equal(O.philox_normal(98, 0, 0, 6, 500), n[:, :500])      # other seed


def _planar_case(nvars, C, B, seed=7, **kw):
    """PlanarLayer(nvars, tanh; n_cond = C) as the vector field (test/call_tests.jl:98-127)"""
    rng = np.random.default_rng(seed)
    d0 = nvars + C
    cfg = O.make_config(nvars, 0, [d0, 1, nvars], C_=C, planar=True, **kw)
    theta = np.concatenate([rng.normal(0, 0.8, nvars), rng.normal(0, 0.8, d0), rng.normal(0, 0.3, 1)]).astype(np.float32)
    xs = rng.normal(0, 1, (B, nvars)).astype(np.float32)
    ys = rng.normal(0, 1, (B, C)).astype(np.float32) if C else None
    eps = rng.normal(0, 1, (B, nvars)).astype(np.float32)
    return cfg, theta, xs, ys, eps


@pytest.mark.parametrize("C", [0, 3])
@pytest.mark.parametrize("tm", [0, 1])
def test_n3_planar_layer_closed_form_trace(C, tm):
    """N3: with nn = PlanarLayer the field is u * tanh(w'[z; ys] + b) (src/layers/planar_layer.jl:70-88)
    and the exact trace has the closed form act'(.) * (w[1:D] . u)."""
    nv, B = 5, 33
    cfg, theta, xs, ys, eps = _planar_case(nv, C, B, trace_mode=tm)
    u = np.concatenate([xs, np.zeros((B, 1), np.float32)], axis=1)
    du = O.augmented_f(cfg, theta, 0, u, ys, None)
    uu, w, b = theta[:nv].astype(np.float64), theta[nv:nv + nv + C].astype(np.float64), float(theta[-1])
    inp = xs.astype(np.float64) if C == 0 else np.concatenate([xs, ys], axis=1).astype(np.float64)
    a = np.tanh(inp @ w + b)
    np.testing.assert_allclose(du[:, :nv], a[:, None] * uu[None, :], rtol=2e-5, atol=2e-6)
    np.testing.assert_allclose(du[:, nv], -(1 - a * a) * float(w[:nv] @ uu), rtol=2e-5, atol=2e-6)


@pytest.mark.parametrize("C,tm,mode", [(0, 0, 1), (0, 1, 1), (2, 0, 1), (0, 0, 0), (2, 1, 0)])
def test_n3_planar_loss_gradient_vs_float64_autograd(C, tm, mode):
    nv, B = 4, 12
    cfg, theta, xs, ys, eps = _planar_case(nv, C, B, trace_mode=tm, lambda1=1e-2, lambda2=1e-2)
    val, g, gx, st = O.loss_grad(cfg, theta, xs, ys, eps if mode else None, 0.0, 1.0, mode=mode)
    dts = [float(d) for d in O.last_dts()]
    r = F.run(cfg, theta, mode, xs, ys, eps if mode else None, 0.0, 1.0, frozen_dts=dts, requires_grad=True)
    assert g.shape == (nv + nv + C + 1,)
    assert abs(val - r["loss"]) / abs(r["loss"]) < 1e-5
    assert relnorm(g, r["grad_theta"]) < 1e-4
    assert relnorm(gx, r["grad_x"]) < 1e-4
