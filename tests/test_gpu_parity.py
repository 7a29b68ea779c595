"""GPU parity: libcnfb200 (through the C ABI, via the host mirror) vs the CPU oracle on the same
seeded inputs.  Tolerances are BASELINE.json's north_star: log-density 1e-4 relative, loss 1e-4
relative, gradients 1e-3 relative, accepted steps equal or +-1."""
import numpy as np
import pytest

from helpers import O, args_of, cnf, make_icnf, oracle_cfg, relerr, relnorm, wl

pytestmark = pytest.mark.gpu

SMALL = [("C0", 64), ("C1", 1024), ("C2", 4096), ("C3", 512), ("C4", 2048)]


@pytest.mark.parametrize("cid,B", SMALL)
@pytest.mark.parametrize("trace_mode", [0, 1])
@pytest.mark.parametrize("mode", [0, 1])
def test_augmented_f(cid, B, trace_mode, mode):
    inp = wl.make_inputs(cid, B=B)
    c = wl.CONFIGS[cid]
    D = c["nvars"] + c["naug"]
    R = D + (3 if mode else 1)
    rng = np.random.default_rng(5)
    u = rng.normal(0, 1, (B, R)).astype(np.float32)
    icnf = make_icnf(cid, trace_mode=trace_mode)
    m = cnf.TrainMode() if mode else cnf.TestMode()
    du = np.asarray(cnf.augmented_f(icnf, m, u.T, *args_of(inp), eps=inp["eps"].T)).T
    ref = O.augmented_f(oracle_cfg(cid, trace_mode=trace_mode), inp["theta"], mode, u, inp["ys"], inp["eps"])
    scale = np.maximum(np.abs(ref).max(axis=0, keepdims=True), 1e-6)
    # tensor path (wide nets, TrainMode) computes the layer GEMMs with BF16x3 operand splitting
    tol = 1e-4 if (mode == 1 and cid in ("C3", "C4")) else 2e-5
    assert np.max(np.abs(du - ref) / scale) < tol


# ---- bookkeeping of the two documented widenings of north_star's tolerances ("What's weak" 2):
# (a) the conditioning floor of the tspan = (0, 13) configs, (b) a +-1 difference of the accepted-step count.
# Every case records what it actually used; test_zz_tolerance_report prints the table and bounds both.
FLOORS = {}      # test id -> tolerance applied to the relative log-density / loss error
FLIPS = {}       # test id -> True when the accepted-step counts differed by one
FLOOR_MAX = 1e-3  # no case may need a floor above this


def _logdensity_case(name, cid, B, trace_mode, mode, inp=None, kernel_path=0):
    inp = wl.make_inputs(cid, B=B) if inp is None else inp
    c = wl.CONFIGS[cid]
    icnf = make_icnf(cid, trace_mode=trace_mode, kernel_path=kernel_path)
    m = cnf.TrainMode() if mode else cnf.TestMode()
    lp, regs, z = cnf.inference(icnf, m, inp["xs"].T, *args_of(inp), eps=inp["eps"].T, return_z=True)
    st = cnf.last_stats(icnf)
    rlp, rregs, rz, rst = O.logdensity(oracle_cfg(cid, trace_mode=trace_mode), inp["theta"], mode,
                                       inp["xs"], inp["ys"], inp["eps"], *c["tspan"])
    assert st["retcode"] == 0 and rst["retcode"] == 0
    assert abs(st["naccept"] - rst["naccept"]) <= 1, (st, rst)
    # conditioning floor: how far the oracle itself moves under a ~1 ulp change of every tanh
    # (long tspans amplify rounding; a second FP32 implementation cannot beat this spread)
    O.set_perturb(True)
    plp, _, pz, pst = O.logdensity(oracle_cfg(cid, trace_mode=trace_mode), inp["theta"], mode,
                                   inp["xs"], inp["ys"], inp["eps"], *c["tspan"])
    O.set_perturb(False)
    floor = 4 * relerr(plp, rlp) if pst["naccept"] == rst["naccept"] else 0.0
    tol = max(1e-4, floor)
    FLIPS[name] = st["naccept"] != rst["naccept"]
    err = relerr(lp, rlp)
    # the SAME tolerance whether or not the accepted-step counts agree: north_star asks for 1e-4 AND steps +-1.
    # (Where the counts differ it is by one step fewer on the tensor-core paths, systematically: their f has less
    # rounding noise than an FP32 FMA chain, and on these problems the abstol = 1e-6 rows of the error norm turn that
    # noise into the step-size sequence — DESIGN.md §6.  Both solves are far more accurate than the solver tolerance.)
    FLOORS[name] = tol
    print(f"{name}: max rel logp err {err:.2e}, tolerance applied {tol:.2e} (floor {floor:.2e}), naccept {st['naccept']} "
          f"vs oracle {rst['naccept']}")
    assert tol <= FLOOR_MAX, (name, tol)
    assert err < tol
    assert relnorm(np.asarray(z).T, rz) < tol
    assert np.allclose(np.stack(regs, 1), rregs, rtol=1e-4, atol=1e-5)
    return st


@pytest.mark.parametrize("cid,B", SMALL)
@pytest.mark.parametrize("trace_mode", [0, 1])
@pytest.mark.parametrize("mode", [0, 1])
def test_logdensity(cid, B, trace_mode, mode):
    _logdensity_case(f"logdensity[{cid},B={B},tm={trace_mode},mode={mode}]", cid, B, trace_mode, mode)


def test_logdensity_c2_at_the_benchmarked_size():
    """C2 at B = 2^20, the size bench.py times: ntiles >> grid, the persistent tile loop of k_stage_narrow_tc and the
    whole-batch RMS norm over 3 x 2^20 entries against the oracle's solve of the same batch"""
    st = _logdensity_case("logdensity[C2,B=2^20]", "C2", 1 << 20, 1, 0)
    assert st["kernel_path"] == 6


def test_logdensity_c5_shape():
    """C5 (128 -> 1024 -> 1024 -> 128, tensor-bound layer GEMMs): per-sample log-density, not only loss/gradient"""
    st = _logdensity_case("logdensity[C5,B=256]", "C5", 256, 0, 1)
    assert st["kernel_path"] == 3


def test_logdensity_c4_hutchinson_large():
    """C4 Hutchinson inference at 2^16 samples on the fused multi-layer kernel (one 128-sample tile per CTA pass, many
    passes per CTA); the full 2^20 of BASELINE.json's config runs with CNFB200_SLOW=1 (the oracle needs minutes)"""
    import os
    B = (1 << 20) if os.environ.get("CNFB200_SLOW") == "1" else (1 << 16)
    st = _logdensity_case(f"logdensity[C4,B={B},hutchinson]", "C4", B, 0, 1)
    assert st["kernel_path"] == 7


@pytest.mark.parametrize("cid,B", [("C0", 64), ("C1", 1024), ("C3", 256), ("C4", 1024), ("C5", 64)])
@pytest.mark.parametrize("trace_mode", [0, 1])
def test_loss_grad(cid, B, trace_mode):
    inp = wl.make_inputs(cid, B=B)
    c = wl.CONFIGS[cid]
    icnf = make_icnf(cid, trace_mode=trace_mode)
    val, g, gx = cnf.loss_and_grad(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp),
                                   eps=inp["eps"].T, want_grad_x=True)
    st = cnf.last_stats(icnf)
    rval, rg, rgx, rst = O.loss_grad(oracle_cfg(cid, trace_mode=trace_mode), inp["theta"], inp["xs"],
                                     inp["ys"], inp["eps"], *c["tspan"])
    assert abs(st["naccept"] - rst["naccept"]) <= 1, (st, rst)
    O.set_perturb(True)
    pval, pg, _, pst = O.loss_grad(oracle_cfg(cid, trace_mode=trace_mode), inp["theta"], inp["xs"],
                                   inp["ys"], inp["eps"], *c["tspan"])
    O.set_perturb(False)
    same = pst["naccept"] == rst["naccept"]
    fl_l = 4 * abs(pval - rval) / abs(rval) if same else 0.0
    fl_g = 4 * relnorm(pg, rg) if same else 0.0
    tol_l, tol_g = max(1e-4, fl_l), max(1e-3, fl_g)     # the same whether or not the step counts agree (see _logdensity_case)
    name = f"loss_grad[{cid},B={B},tm={trace_mode}]"
    FLIPS[name] = st["naccept"] != rst["naccept"]
    FLOORS[name] = tol_l
    assert tol_l <= FLOOR_MAX and tol_g <= 10 * FLOOR_MAX, (name, tol_l, tol_g)
    print(f"{name}: loss rel err {abs(val - rval) / abs(rval):.2e} (tol {tol_l:.1e}), grad rel err {relnorm(g, rg):.2e} "
          f"(tol {tol_g:.1e}), naccept {st['naccept']} vs {rst['naccept']}")
    assert abs(val - rval) / abs(rval) < tol_l
    assert relnorm(g, rg) < tol_g
    assert relnorm(np.asarray(gx).T, rgx) < tol_g


@pytest.mark.parametrize("cid,B,tm", [("C2", 1024, 1), ("C2", 512, 0), ("C1", 256, 0), ("C4", 128, 0), ("C4", 384, 1)])
def test_loss_grad_testmode(cid, B, tm):
    """gradient of loss(icnf, TestMode(), ...) = -mean(logp) w.r.t. ps and xs (reference
    test/call_tests.jl:239-251 differentiates it): exact-trace reverse sweep vs the oracle's"""
    inp = wl.make_inputs(cid, B=B)
    c = wl.CONFIGS[cid]
    icnf = make_icnf(cid, trace_mode=tm)
    val, g, gx = cnf.loss_and_grad(icnf, cnf.TestMode(), inp["xs"].T, *args_of(inp), want_grad_x=True)
    st = cnf.last_stats(icnf)
    rval, rg, rgx, rst = O.loss_grad(oracle_cfg(cid, trace_mode=tm), inp["theta"], inp["xs"], inp["ys"], None,
                                     *c["tspan"], mode=0)
    assert abs(st["naccept"] - rst["naccept"]) <= 1, (st, rst)
    O.set_perturb(True)
    pval, pg, _, pst = O.loss_grad(oracle_cfg(cid, trace_mode=tm), inp["theta"], inp["xs"], inp["ys"], None,
                                   *c["tspan"], mode=0)
    O.set_perturb(False)
    same = pst["naccept"] == rst["naccept"]
    fl_l = 4 * abs(pval - rval) / abs(rval) if same else 0.0
    fl_g = 4 * relnorm(pg, rg) if same else 0.0
    tol_l, tol_g = max(1e-4, fl_l), max(1e-3, fl_g)
    print(f"loss_grad_testmode[{cid},B={B},tm={tm}]: loss rel err {abs(val - rval) / abs(rval):.2e}, grad rel err {relnorm(g, rg):.2e}, "
          f"naccept {st['naccept']} vs {rst['naccept']}, path {st['kernel_path']}")
    assert abs(val - rval) / abs(rval) < tol_l
    assert relnorm(g, rg) < tol_g
    assert relnorm(np.asarray(gx).T, rgx) < tol_g
    # and the loss agrees with the TestMode inference path
    assert abs(val - cnf.loss(icnf, cnf.TestMode(), inp["xs"].T, *args_of(inp))) < 1e-5 * abs(val)


@pytest.mark.parametrize("cid,B", [("C1", 512), ("C2", 2048), ("C4", 512)])
def test_generate_roundtrip(cid, B):
    """K7: generate from z(t1) returns x within solver tolerance; and parity with the oracle."""
    inp = wl.make_inputs(cid, B=B)
    c = wl.CONFIGS[cid]
    icnf = make_icnf(cid)
    m = cnf.TestMode()
    lp, regs, z = cnf.inference(icnf, m, inp["xs"].T, *args_of(inp), return_z=True)
    z = np.asarray(z)
    xr = np.asarray(cnf.generate(icnf, m, *args_of(inp), B, z0=z)).T
    assert np.max(np.abs(xr - inp["xs"])) < 2e-2
    ref, _ = O.generate(oracle_cfg(cid), inp["theta"], 0, z.T, inp["ys"], None, c["tspan"][1], c["tspan"][0])
    assert np.max(np.abs(xr - ref)) < 1e-3


def test_device_pointers_match_host_buffers():
    import torch
    cid, B = "C2", 8192
    inp = wl.make_inputs(cid, B=B)
    icnf = make_icnf(cid)
    lp_h, _ = cnf.inference(icnf, cnf.TestMode(), inp["xs"].T, inp["theta"], {})
    xd = torch.from_numpy(inp["xs"]).cuda().t()
    lp_d, _ = cnf.inference(icnf, cnf.TestMode(), xd, inp["theta"], {})
    assert lp_d.is_cuda
    assert np.array_equal(lp_d.cpu().numpy(), lp_h)


@pytest.mark.parametrize("cid,B", [("C3", 256), ("C4", 1024)])
def test_forced_generic_path_on_wide_nets(cid, B):
    """kernel_path = 1 keeps the FP32 SIMT kernels reachable for the wide configs (and checks that the
    tensor path and the generic path agree with each other through a whole solve)"""
    inp = wl.make_inputs(cid, B=B)
    c = wl.CONFIGS[cid]
    out = {}
    for kp in (1, 3):
        icnf = make_icnf(cid, kernel_path=kp)
        val, g = cnf.loss_and_grad(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp), eps=inp["eps"].T)
        st = cnf.last_stats(icnf)
        assert st["kernel_path"] == kp
        out[kp] = (val, g, st["naccept"])
    assert abs(out[1][2] - out[3][2]) <= 1
    if out[1][2] == out[3][2]:
        assert abs(out[1][0] - out[3][0]) / abs(out[1][0]) < 1e-4
        assert relnorm(out[3][1], out[1][1]) < 1e-3


def test_tensor_path_is_selected_for_wide_train_and_narrow_for_c2():
    inp = wl.make_inputs("C3", B=256)
    icnf = make_icnf("C3")
    cnf.inference(icnf, cnf.TrainMode(), inp["xs"].T, inp["theta"], {}, eps=inp["eps"].T)
    assert cnf.last_stats(icnf)["kernel_path"] == 3
    inp = wl.make_inputs("C2", B=256)
    icnf = make_icnf("C2")
    cnf.inference(icnf, cnf.TestMode(), inp["xs"].T, inp["theta"], {})
    assert cnf.last_stats(icnf)["kernel_path"] == 6     # narrow net, layer 2 on tcgen05
    icnf = make_icnf("C2", kernel_path=2)
    cnf.inference(icnf, cnf.TestMode(), inp["xs"].T, inp["theta"], {})
    assert cnf.last_stats(icnf)["kernel_path"] == 2     # register-resident SIMT kernel stays reachable
    # the reference's default compute mode (VecJac) builds the same Jacobian row by row: same fast path
    icnf = make_icnf("C2", trace_mode=0)
    lp0, _ = cnf.inference(icnf, cnf.TestMode(), inp["xs"].T, inp["theta"], {})
    assert cnf.last_stats(icnf)["kernel_path"] == 6
    rlp, _, _, rst = O.logdensity(oracle_cfg("C2", trace_mode=0), inp["theta"], 0, inp["xs"], None, None, 0.0, 1.0)
    assert abs(cnf.last_stats(icnf)["naccept"] - rst["naccept"]) <= 1
    if cnf.last_stats(icnf)["naccept"] == rst["naccept"]:
        assert relerr(lp0, rlp) < 1e-4


def _narrow_model(nvars, kernel_path, tspan=(0.0, 1.0)):
    nn = cnf.Chain(cnf.Dense(nvars, 64, "tanh"), cnf.Dense(64, 64, "tanh"), cnf.Dense(64, nvars, "tanh"))
    return cnf.construct(cnf.FFJORD, nn, nvars, 0, compute_mode=cnf.DIJacVecMatrixMode(), tspan=tspan,
                         resource=cnf.B200Libs(0, kernel_path), rng=np.random.default_rng(1))


@pytest.mark.parametrize("nvars,B", [(2, 1), (2, 129), (2, 1000), (2, 40000), (1, 77), (1, 5000)])
def test_narrow_tensor_path_matches_oracle_and_simt_kernel(nvars, B):
    """k_stage_narrow_tc (kernel_path 6: layer 2 as FP16 hi/lo x3 tcgen05 MMAs) against the oracle and the
    register-resident FP32 kernel (kernel_path 2): ragged batches, one tile to several tiles per group,
    D = 1 and D = 2"""
    dims = [nvars, 64, 64, nvars]
    rng = np.random.default_rng(100 + B)
    theta = wl.glorot_theta(dims, rng)
    theta = (theta + 0.05 * rng.normal(size=theta.shape)).astype(np.float32)   # non-zero biases
    xs = rng.normal(0, 1.5, (B, nvars)).astype(np.float32)
    out = {}
    for kp in (0, 2):
        icnf = _narrow_model(nvars, kp)
        lp, _ = cnf.inference(icnf, cnf.TestMode(), xs.T, theta, {})
        st = cnf.last_stats(icnf)
        assert st["kernel_path"] == (6 if kp == 0 else 2) and st["retcode"] == 0
        out[kp] = (np.asarray(lp), st["naccept"], st["nreject"])
    assert abs(out[0][1] - out[2][1]) <= 1
    if out[0][1:] == out[2][1:]:
        assert relerr(out[0][0], out[2][0]) < 1e-4
    if B <= 5000:
        cfg = O.make_config(nvars, 0, dims, trace_mode=1)
        rlp, _, _, rst = O.logdensity(cfg, theta, 0, xs, None, None, 0.0, 1.0)
        assert abs(out[0][1] - rst["naccept"]) <= 1
        if out[0][1] == rst["naccept"]:
            assert relerr(out[0][0], rlp) < 1e-4


def test_narrow_tensor_path_generate_roundtrip_and_nan():
    """reverse-time solve (generate) on the narrow tensor path returns the data; a NaN sample poisons the
    whole-batch error norm exactly like the reference's matrix-mode solve (retcode != 0, no hang)"""
    inp = wl.make_inputs("C2", B=3000)
    icnf = make_icnf("C2")
    lp, _, z = cnf.inference(icnf, cnf.TestMode(), inp["xs"].T, inp["theta"], {}, return_z=True)
    assert cnf.last_stats(icnf)["kernel_path"] == 6
    xs2 = cnf.generate(icnf, cnf.TestMode(), inp["theta"], {}, 3000, z0=np.asarray(z))
    assert cnf.last_stats(icnf)["kernel_path"] == 6
    assert np.max(np.abs(np.asarray(xs2).T - inp["xs"])) < 5e-3
    bad = inp["xs"].copy()
    bad[17, 0] = np.nan
    lpb, _ = cnf.inference(icnf, cnf.TestMode(), bad.T, inp["theta"], {})
    assert cnf.last_stats(icnf)["retcode"] != 0 or not np.isfinite(np.asarray(lpb)[17])


@pytest.mark.parametrize("nvars,naug,C,hidden,B", [(5, 0, 3, (160, 96), 300), (3, 2, 0, (64, 32, 96), 1000), (20, 0, 0, (320,), 130)])
def test_tensor_path_odd_shapes(nvars, naug, C, hidden, B):
    """tensor path forced on shapes that exercise every padding rule (widths that are not multiples
    of 64, D < 16, condition rows, 2- and 4-layer chains, ragged batches) against the generic path"""
    D = nvars + naug
    dims = [D + C, *hidden, D]
    rng = np.random.default_rng(11)
    theta = wl.glorot_theta(dims, rng)
    theta[theta == 0] = 0.05 * rng.standard_normal(int((theta == 0).sum())).astype(np.float32)   # non-zero biases
    xs = rng.normal(0, 1, (B, nvars)).astype(np.float32)
    ys = rng.normal(0, 1, (B, C)).astype(np.float32) if C else None
    eps = rng.normal(0, 1, (B, D)).astype(np.float32)
    nn = cnf.Chain(*[cnf.Dense(dims[i], dims[i + 1], "tanh") for i in range(len(dims) - 1)])
    tag = cnf.CondRNODE if C else cnf.RNODE
    out = {}
    for kp in (1, 3):
        icnf = cnf.construct(tag, nn, nvars, naug, compute_mode=cnf.DIVecJacMatrixMode(), tspan=(0.0, 1.0),
                             lambda1=1e-2, lambda2=1e-2, lambda3=(1e-2 if naug else 0.0),
                             resource=cnf.B200Libs(0, kp), rng=np.random.default_rng(1))
        tail = ((ys.T,) if C else ()) + (theta, {})
        val, g, gx = cnf.loss_and_grad(icnf, cnf.TrainMode(), xs.T, *tail, eps=eps.T, want_grad_x=True)
        st = cnf.last_stats(icnf)
        assert st["kernel_path"] == kp and st["retcode"] == 0
        out[kp] = (val, g, np.asarray(gx), st["naccept"])
    assert abs(out[1][3] - out[3][3]) <= 1
    if out[1][3] == out[3][3]:
        assert abs(out[1][0] - out[3][0]) / abs(out[1][0]) < 1e-4
        assert relnorm(out[3][1], out[1][1]) < 1e-3
        assert relnorm(out[3][2], out[1][2]) < 1e-3


@pytest.mark.gpu
@pytest.mark.parametrize("cid,B", [("C3", 300), ("C0", 1024)])
def test_n1_device_eps_matches_oracle_stream_and_host_eps_path(cid, B):
    """N1: eps drawn on the device (Philox4x32-10 + Box-Muller) = the oracle's restatement of the
    stream; a call with the drawn eps passed back from the host gives the same bits; the next call
    uses the next draw; a shard with sample_offset reproduces its columns of the global draw."""
    inp = wl.make_inputs(cid, B=B)
    icnf = make_icnf(cid)
    cnf.seed_device_rng(icnf, 4242)
    v1, g1 = cnf.loss_and_grad(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp))
    e1 = np.array(cnf.last_eps(icnf, B))
    ref = O.philox_normal(4242, 0, 0, icnf.D, B)
    np.testing.assert_allclose(e1, ref, rtol=0, atol=3e-6)
    v2, g2 = cnf.loss_and_grad(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp))
    e2 = np.array(cnf.last_eps(icnf, B))
    np.testing.assert_allclose(e2, O.philox_normal(4242, 0, 1, icnf.D, B), rtol=0, atol=3e-6)
    assert v1 != v2
    host = make_icnf(cid)
    v3, g3 = cnf.loss_and_grad(host, cnf.TrainMode(), inp["xs"].T, *args_of(inp), eps=e1)
    assert v3 == v1
    assert relnorm(np.asarray(g3), np.asarray(g1)) < 1e-5   # parameter gradients are accumulated with atomics
    # against the oracle on the same eps
    ov, og, _, ost = O.loss_grad(oracle_cfg(cid), inp["theta"], inp["xs"], inp["ys"], e1.T.copy(), *wl.CONFIGS[cid]["tspan"])
    same = cnf.last_stats(host)["naccept"] == ost["naccept"]
    assert abs(v1 - ov) <= (1e-4 if same else 5e-3) * abs(ov)
    assert relnorm(np.asarray(g1), og) < (1e-3 if same else 5e-2)
    # shard invariance
    sh = make_icnf(cid)
    cnf.seed_device_rng(sh, 4242, sample_offset=100)
    cnf.loss_and_grad(sh, cnf.TrainMode(), inp["xs"].T[:, 100:228], *[a[:, 100:228] if getattr(a, "ndim", 0) == 2 and a.shape[1] == B else a for a in args_of(inp)])
    np.testing.assert_array_equal(np.array(cnf.last_eps(sh, 128)), e1[:, 100:228])


@pytest.mark.gpu
def test_n1_generate_draws_base_sample_on_device():
    """generate with z0 = None on a seeded handle: z0 is draw 0 of the stream, so passing the
    oracle's draw as z0 to an unseeded handle must give (nearly) the same samples."""
    cid, n = "C2", 2000
    inp = wl.make_inputs(cid, B=n)
    icnf = make_icnf(cid)
    cnf.seed_device_rng(icnf, 7)
    xs = np.array(cnf.generate(icnf, cnf.TestMode(), *args_of(inp), n))
    z0 = O.philox_normal(7, 0, 0, icnf.D, n)
    host = make_icnf(cid)
    xs2 = np.array(cnf.generate(host, cnf.TestMode(), *args_of(inp), n, z0=z0))
    assert xs.shape == (icnf.nvars, n)
    np.testing.assert_allclose(xs, xs2, rtol=1e-4, atol=1e-4)   # z0 agrees to 3e-6; the flow amplifies it


def _oracle_fit(cid, inp, N, bs, n_epochs, perm, seed, eta, beta, t1s=None):
    """the reference's fit loop restated on the host with the oracle's loss gradient, the oracle's
    Philox eps stream (one draw per mini-batch) and Lion (tests only)"""
    cfg, c = oracle_cfg(cid), wl.CONFIGS[cid]
    ps = inp["theta"].copy()
    m = np.zeros_like(ps)
    D = c["nvars"] + c["naug"]
    hist, draw = [], 0
    for e in range(n_epochs):
        tot = 0.0
        for lo in range(0, N, bs):
            idx = perm[e, lo:lo + bs]
            eps = O.philox_normal(seed, 0, draw, D, len(idx)).T.copy()
            draw += 1
            ys = inp["ys"][idx] if inp["ys"] is not None else None
            t1 = c["tspan"][1] if t1s is None else t1s[draw - 1]
            val, g, _, _ = O.loss_grad(cfg, ps, inp["xs"][idx], ys, eps, c["tspan"][0], t1)
            cc = beta[0] * m + (1 - beta[0]) * g
            ps = (ps - eta * np.sign(cc)).astype(np.float32)
            m = (beta[1] * m + (1 - beta[1]) * g).astype(np.float32)
            tot += val * len(idx)
        hist.append(tot / N)
    return ps, np.array(hist)


@pytest.mark.parametrize("cid,N,bs,n_epochs", [("C1", 300, 128, 2), ("C4", 512, 256, 2), ("C3", 384, 128, 1)])
def test_n2_fit_epochs_device_loop_vs_oracle_loop(cid, N, bs, n_epochs):
    """N2: the device-resident epoch loop (gather, device eps, loss+grad, Lion) against the same loop
    on the host with the oracle's gradients: per-epoch losses to 1e-4; parameters equal except where
    a Lion sign flips on a near-zero momentum (each flip moves a coordinate by 2 eta)."""
    inp = wl.make_inputs(cid, B=N)
    eta, beta, seed = 1e-3, (0.9, 0.999), 31337
    perm = np.stack([np.random.default_rng(5 + e).permutation(N) for e in range(n_epochs)])
    icnf = make_icnf(cid)
    cnf.seed_device_rng(icnf, seed)
    ys = inp["ys"].T if inp["ys"] is not None else None
    ps, hist = cnf.fit_epochs(icnf, inp["xs"].T, ys, inp["theta"], batch_size=bs, n_epochs=n_epochs, perm=perm,
                              eta=eta, beta=beta)
    rps, rhist = _oracle_fit(cid, inp, N, bs, n_epochs, perm, seed, eta, beta)
    steps = n_epochs * -(-N // bs)
    np.testing.assert_allclose(hist, rhist, rtol=2e-4)
    d = np.abs(ps - rps)
    assert d.max() <= 2 * eta * steps + 1e-6
    assert (d > 1e-6).mean() < 0.03, (d > 1e-6).mean()
    # the fitted parameters stay in the handle: a loss call without re-uploading sees them
    assert not np.array_equal(ps, inp["theta"])


def test_n2_fit_epochs_steers_t1_per_minibatch():
    """the reference draws the steered t1 inside every TrainMode loss call of its fit
    (src/base_icnf.jl:108-121, :283), so with steer_rate > 0 every mini-batch integrates to its own t1.
    The device loop draws t1 per mini-batch from the documented stream (cnf.steer_sequence); the same loop
    on the host with the oracle and those t1 values must agree, and the t1 values must really differ."""
    cid, N, bs, n_epochs = "C1", 300, 128, 2
    inp = wl.make_inputs(cid, B=N)
    eta, beta, seed, rate = 1e-3, (0.9, 0.999), 4242, 0.1
    perm = np.stack([np.random.default_rng(9 + e).permutation(N) for e in range(n_epochs)])
    icnf = make_icnf(cid)
    icnf.steer_rate = rate
    cnf.seed_device_rng(icnf, seed)
    ps, hist = cnf.fit_epochs(icnf, inp["xs"].T, None, inp["theta"], batch_size=bs, n_epochs=n_epochs, perm=perm,
                              eta=eta, beta=beta)
    steps = n_epochs * -(-N // bs)
    t1s = cnf.steer_sequence(seed, steps, rate, wl.CONFIGS[cid]["tspan"])
    t0, t1 = wl.CONFIGS[cid]["tspan"]
    assert len(set(t1s)) == steps and all(abs(t - t1) <= rate * abs(t1 - t0) * 1.0001 for t in t1s)
    assert abs(cnf.last_stats(icnf)["t_final"] - t1s[-1]) < 1e-6      # the last mini-batch stopped at ITS t1
    rps, rhist = _oracle_fit(cid, inp, N, bs, n_epochs, perm, seed, eta, beta, t1s=t1s)
    np.testing.assert_allclose(hist, rhist, rtol=2e-4)
    d = np.abs(ps - rps)
    assert d.max() <= 2 * eta * steps + 1e-6
    assert (d > 1e-6).mean() < 0.03, (d > 1e-6).mean()
    # and an unsteered fit of the same data gives a different history
    plain = make_icnf(cid)
    cnf.seed_device_rng(plain, seed)
    _, hist0 = cnf.fit_epochs(plain, inp["xs"].T, None, inp["theta"], batch_size=bs, n_epochs=n_epochs, perm=perm,
                              eta=eta, beta=beta)
    assert np.max(np.abs(hist0 - hist)) > 1e-6


def test_n2_adapter_fit_uses_device_loop_and_matches_per_batch_calls():
    """fit(ICNFModel) with a seeded device stream runs the one-call loop; the same model fitted
    through per-mini-batch loss_and_grad calls (same permutations, same device eps draws) agrees."""
    rng = np.random.default_rng(3)
    X = rng.beta(2.0, 4.0, size=(256, 1)).astype(np.float32)
    def mk():
        nn = cnf.Chain(cnf.Dense(1, 8, "tanh"), cnf.Dense(8, 1, "tanh"))
        return cnf.construct(cnf.RNODE, nn, 1, 0, tspan=(0.0, 4.0), resource=cnf.B200Libs(0), rng=np.random.default_rng(11))
    a, b = mk(), mk()
    cnf.seed_device_rng(a, 77)
    cnf.seed_device_rng(b, 77)
    ma = cnf.ICNFModel(a, n_epochs=2, batch_size=64)
    (psa, _), _, ra = cnf.fit(ma, 0, X)
    mb = cnf.ICNFModel(b, n_epochs=2, batch_size=64, optimizers=(cnf.Lion(),))
    mb.optimizers[0].m = np.zeros_like(psa)      # a used optimiser state forces the per-batch path
    (psb, _), _, rb = cnf.fit(mb, 0, X)
    np.testing.assert_allclose(ra["loss_history"], rb["loss_history"], rtol=1e-4)
    d = np.abs(psa - psb)
    assert d.max() <= 2e-3 * 8 + 1e-6 and (d > 1e-6).mean() < 0.05


def _vector_icnf(cid, mode_cls):
    c = wl.CONFIGS[cid]
    dims = c["dims"]
    nn = cnf.Chain(*[cnf.Dense(dims[i], dims[i + 1], "tanh") for i in range(len(dims) - 1)])
    return cnf.construct(cnf.FFJORD, nn, c["nvars"], c["naug"], compute_mode=mode_cls(), tspan=c["tspan"],
                         resource=cnf.B200Libs(0), rng=np.random.default_rng(1))


@pytest.mark.parametrize("mode_cls", ["DIJacVecVectorMode", "DIVecJacVectorMode"])
def test_n4_vector_mode_matches_per_sample_oracle_solves(mode_cls):
    """N4: VectorMode = every column its own adaptive solve (own dt, own error norm over its R
    entries).  Oracle: the CPU solve called once per column with B = 1.  logp to 1e-4, z to 1e-4,
    the largest accepted-step count within 1; and the whole batch is one kernel launch."""
    cid, B = "C2", 333
    inp = wl.make_inputs(cid, B=B)
    c = wl.CONFIGS[cid]
    icnf = _vector_icnf(cid, getattr(cnf, mode_cls))
    lp, regs, z = cnf.inference(icnf, cnf.TestMode(), inp["xs"].T, *args_of(inp), return_z=True)
    st = cnf.last_stats(icnf)
    lp, z = np.asarray(lp), np.asarray(z).T
    cfg = oracle_cfg(cid)
    rlp, rz, nacc = np.empty(B, np.float32), np.empty_like(z), []
    for b in range(B):
        out = O.logdensity(cfg, inp["theta"], 0, inp["xs"][b:b + 1], None, None, *c["tspan"])
        rlp[b], rz[b] = out[0][0], out[2][0]
        nacc.append(out[-1]["naccept"])
    assert np.max(np.abs(lp - rlp) / np.maximum(1.0, np.abs(rlp))) < 1e-4
    assert np.max(np.abs(z - rz)) < 1e-4
    assert abs(st["naccept"] - max(nacc)) <= 1, (st, max(nacc))
    assert len(set(nacc)) > 1          # the samples really take different numbers of steps
    assert st["kernel_path"] == 4 and st["gpu_launches"] <= 4
    # MatrixMode on the same data couples the columns through one error norm: a different answer
    mat = make_icnf(cid)
    lpm, _ = cnf.inference(mat, cnf.TestMode(), inp["xs"].T, *args_of(inp))
    assert np.max(np.abs(np.asarray(lpm) - lp)) > 0
    assert np.max(np.abs(np.asarray(lpm) - lp)) < 5e-2


def test_n4_vector_mode_generate_roundtrip():
    cid, B = "C2", 500
    inp = wl.make_inputs(cid, B=B)
    icnf = _vector_icnf(cid, cnf.DIJacVecVectorMode)
    m = cnf.TestMode()
    lp, regs, z = cnf.inference(icnf, m, inp["xs"].T, *args_of(inp), return_z=True)
    xr = np.asarray(cnf.generate(icnf, m, *args_of(inp), B, z0=np.asarray(z))).T
    assert np.max(np.abs(xr - inp["xs"])) < 2e-2
    # (TrainMode and the shapes outside the narrow list run under the per-sample controller: test_n4_vector_mode_any_net_and_mode)


def _planar_pair(nvars, C, B, trace_mode, seed=7):
    """(product model, oracle config, inputs) for nn = PlanarLayer(nvars, tanh; n_cond = C)"""
    rng = np.random.default_rng(seed)
    d0 = nvars + C
    cfg = O.make_config(nvars, 0, [d0, 1, nvars], C_=C, planar=True, trace_mode=trace_mode, lambda1=1e-2, lambda2=1e-2)
    theta = np.concatenate([rng.normal(0, 0.8, nvars), rng.normal(0, 0.8, d0), rng.normal(0, 0.3, 1)]).astype(np.float32)
    xs = rng.normal(0, 1, (B, nvars)).astype(np.float32)
    ys = rng.normal(0, 1, (B, C)).astype(np.float32) if C else None
    eps = rng.normal(0, 1, (B, nvars)).astype(np.float32)
    cm = cnf.DIVecJacMatrixMode() if trace_mode == 0 else cnf.DIJacVecMatrixMode()
    tag = cnf.CondPlanar if C else cnf.Planar
    icnf = cnf.construct(tag, cnf.PlanarLayer(nvars, "tanh", n_cond=C), nvars, 0, compute_mode=cm, tspan=(0.0, 1.0),
                         lambda1=1e-2, lambda2=1e-2, resource=cnf.B200Libs(0), rng=np.random.default_rng(1))
    tail = ((ys.T,) if C else ()) + (theta, {})
    return icnf, cfg, theta, xs, ys, eps, tail


@pytest.mark.parametrize("C", [0, 3])
@pytest.mark.parametrize("trace_mode", [0, 1])
def test_n3_planar_flow_matches_oracle(C, trace_mode):
    """N3: the Planar / CondPlanar tags (nn = PlanarLayer) on the new resource: log-density in both
    modes, loss and gradients (TrainMode Hutchinson and TestMode exact trace), generate round trip."""
    nv, B = 6, 777
    icnf, cfg, theta, xs, ys, eps, tail = _planar_pair(nv, C, B, trace_mode)
    for mode, m in ((0, cnf.TestMode()), (1, cnf.TrainMode())):
        lp, regs, z = cnf.inference(icnf, m, xs.T, *tail, eps=eps.T if mode else None, return_z=True)
        st = cnf.last_stats(icnf)
        rlp, rregs, rz, rst = O.logdensity(cfg, theta, mode, xs, ys, eps if mode else None, 0.0, 1.0)
        assert abs(st["naccept"] - rst["naccept"]) <= 1
        tol = 1e-4 if st["naccept"] == rst["naccept"] else 5e-3
        assert np.max(np.abs(np.asarray(lp) - rlp) / np.maximum(1.0, np.abs(rlp))) < tol
        assert np.max(np.abs(np.asarray(z).T - rz)) < tol
        val, g, gx = cnf.loss_and_grad(icnf, m, xs.T, *tail, eps=eps.T if mode else None, want_grad_x=True)
        st = cnf.last_stats(icnf)
        rval, rg, rgx, rst = O.loss_grad(cfg, theta, xs, ys, eps if mode else None, 0.0, 1.0, mode=mode)
        same = st["naccept"] == rst["naccept"]
        assert np.asarray(g).shape == (2 * nv + C + 1,)
        assert abs(val - rval) <= (1e-4 if same else 5e-3) * abs(rval)
        assert relnorm(np.asarray(g), rg) < (1e-3 if same else 5e-2)
        assert relnorm(np.asarray(gx).T, rgx) < (1e-3 if same else 5e-2)
    lp, regs, z = cnf.inference(icnf, cnf.TestMode(), xs.T, *tail, return_z=True)
    xr = np.asarray(cnf.generate(icnf, cnf.TestMode(), *tail, B, z0=np.asarray(z))).T
    assert np.max(np.abs(xr - xs)) < 2e-2


@pytest.mark.parametrize("cid,B,tol", [("C1", 700, 1e-5), ("C0", 64, 1e-5), ("C3", 300, 3e-4), ("C4", 1500, 3e-4)])
def test_stage_tape_matches_stage_recompute(cid, B, tol):
    """the reverse sweep reads the stage derivatives k_1..k_5 the forward solve taped (default) or
    recomputes them from u_n (stage_tape = 0): same gradient"""
    inp = wl.make_inputs(cid, B=B)
    out = []
    for tape in (1, 0):
        icnf = make_icnf(cid)
        cnf.set_option(icnf, "stage_tape", tape)
        val, g, gx = cnf.loss_and_grad(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp), eps=inp["eps"].T, want_grad_x=True)
        out.append((val, np.asarray(g), np.asarray(gx), cnf.last_stats(icnf)["gpu_launches"]))
    assert out[0][0] == out[1][0]
    assert relnorm(out[0][1], out[1][1]) < tol
    assert relnorm(out[0][2], out[1][2]) < tol
    assert out[0][3] < out[1][3]          # fewer launches: no recomputed evaluations


@pytest.mark.parametrize("cid,B", [("C1", 32), ("C0", 20), ("C1", 7), ("C1", 1000), ("C0", 64), ("C1", 5000)])
def test_small_batches_run_in_whole_solve_kernels_with_identical_results(cid, B):
    """batches whose tiles of the generic kernel are all resident at once (one tile = the reference's
    default batch_size = 32: block barriers; more tiles: a cooperative launch with grid barriers) run the
    attempt loop and the reverse sweep inside ONE kernel each: same device code, so the results equal
    the launch-per-stage path, in a handful of launches instead of hundreds"""
    inp = wl.make_inputs(cid, B=B)
    res = []
    for tile in (1, 0):
        icnf = make_icnf(cid)
        cnf.set_option(icnf, "tile_solve", tile)
        lp, regs, z = cnf.inference(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp), eps=inp["eps"].T, return_z=True)
        st_i = cnf.last_stats(icnf)
        val, g, gx = cnf.loss_and_grad(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp), eps=inp["eps"].T, want_grad_x=True)
        st_g = cnf.last_stats(icnf)
        lpt, _ = cnf.inference(icnf, cnf.TestMode(), inp["xs"].T, *args_of(inp))
        vt, gt = cnf.loss_and_grad(icnf, cnf.TestMode(), inp["xs"].T, *args_of(inp))
        res.append((np.asarray(lp), np.asarray(z), val, np.asarray(g), np.asarray(gx), np.asarray(lpt), vt, np.asarray(gt), st_i, st_g))
    a, b = res
    np.testing.assert_array_equal(a[0], b[0])
    np.testing.assert_array_equal(a[1], b[1])
    assert a[2] == b[2] and a[6] == b[6]
    np.testing.assert_array_equal(a[5], b[5])
    assert relnorm(a[3], b[3]) < 1e-5 and relnorm(a[4], b[4]) < 1e-5 and relnorm(a[7], b[7]) < 1e-5
    assert a[8]["kernel_path"] == 5 and b[8]["kernel_path"] == 1
    assert a[8]["naccept"] == b[8]["naccept"] and a[9]["naccept"] == b[9]["naccept"]
    assert a[9]["gpu_launches"] < 30 < b[9]["gpu_launches"]


@pytest.mark.parametrize("cid,B", [("C4", 700), ("C3", 300)])
@pytest.mark.parametrize("trace_mode", [0, 1])
def test_tensor_path_exact_trace_testmode(cid, B, trace_mode):
    """TestMode (exact trace) on wide nets runs on the tensor path as D one-hot tangent chains over the same
    layer GEMMs: one evaluation against the oracle, and a whole solve against the oracle and the generic path"""
    inp = wl.make_inputs(cid, B=B)
    c = wl.CONFIGS[cid]
    D = c["nvars"] + c["naug"]
    u = np.random.default_rng(9).normal(0, 1, (B, D + 1)).astype(np.float32)
    icnf = make_icnf(cid, trace_mode=trace_mode)
    du = np.asarray(cnf.augmented_f(icnf, cnf.TestMode(), u.T, *args_of(inp))).T
    ref = O.augmented_f(oracle_cfg(cid, trace_mode=trace_mode), inp["theta"], 0, u, inp["ys"], inp["eps"])
    scale = np.maximum(np.abs(ref).max(axis=0, keepdims=True), 1e-6)
    assert np.max(np.abs(du - ref) / scale) < 1e-4
    out = {}
    for kp in (0, 1):
        m = make_icnf(cid, kernel_path=kp, trace_mode=trace_mode)
        lp, _ = cnf.inference(m, cnf.TestMode(), inp["xs"].T, *args_of(inp))
        st = cnf.last_stats(m)
        assert st["kernel_path"] == (3 if kp == 0 else 1) and st["retcode"] == 0
        out[kp] = (np.asarray(lp), st["naccept"])
    rlp, _, _, rst = O.logdensity(oracle_cfg(cid, trace_mode=trace_mode), inp["theta"], 0, inp["xs"], inp["ys"], None,
                                  *c["tspan"])
    assert abs(out[0][1] - rst["naccept"]) <= 1 and abs(out[0][1] - out[1][1]) <= 1
    if out[0][1] == rst["naccept"]:
        assert relerr(out[0][0], rlp) < 1e-4


def _vector_icnf_any(cid, tm):
    c = wl.CONFIGS[cid]
    dims = c["dims"]
    nn = cnf.Chain(*[cnf.Dense(dims[i], dims[i + 1], "tanh") for i in range(len(dims) - 1)])
    cm = cnf.DIJacVecVectorMode() if tm else cnf.DIVecJacVectorMode()
    tag = {"RNODE": cnf.RNODE, "ICNF": cnf.FFJORD, "CondRNODE": cnf.CondRNODE}[c["tag"]]
    sk = {k: c[k] for k in ("reltol", "abstol") if c[k]}
    return cnf.construct(tag, nn, c["nvars"], c["naug"], compute_mode=cm, tspan=c["tspan"], sol_kwargs=sk,
                         lambda1=c["lambda1"], lambda2=c["lambda2"], lambda3=c["lambda3"],
                         resource=cnf.B200Libs(0), rng=np.random.default_rng(1))


@pytest.mark.parametrize("cid,B,tm,mode,path", [("C1", 48, 0, 1, 8), ("C0", 24, 1, 1, 8), ("C4", 40, 0, 1, 8), ("C4", 24, 1, 0, 8),
                                                 ("C3", 12, 0, 1, 8), ("C2", 64, 0, 1, 8)])
def test_n4_vector_mode_any_net_and_mode(cid, B, tm, mode, path):
    """N4, the rest: VectorMode for TrainMode (Hutchinson), conditional flows, 1/2/3-layer nets and the tensor-core
    widths — the MatrixMode stage kernels under a per-sample controller.  Oracle: the CPU solve called once per column
    with B = 1 (its own dt, error norm and step count)."""
    inp = wl.make_inputs(cid, B=B)
    c = wl.CONFIGS[cid]
    icnf = _vector_icnf_any(cid, tm)
    m = cnf.TrainMode() if mode else cnf.TestMode()
    lp, regs, z = cnf.inference(icnf, m, inp["xs"].T, *args_of(inp), eps=inp["eps"].T, return_z=True)
    st = cnf.last_stats(icnf)
    lp, z, regs = np.asarray(lp), np.asarray(z).T, np.stack([np.asarray(r) for r in regs], 1)
    cfg = oracle_cfg(cid, trace_mode=tm)
    rlp, rz, rregs, nacc = np.empty(B, np.float32), np.empty_like(z), np.empty_like(regs), []
    for b in range(B):
        ys = inp["ys"][b:b + 1] if inp["ys"] is not None else None
        out = O.logdensity(cfg, inp["theta"], mode, inp["xs"][b:b + 1], ys, inp["eps"][b:b + 1], *c["tspan"])
        rlp[b], rregs[b], rz[b] = out[0][0], out[1][0], out[2][0]
        nacc.append(out[-1]["naccept"])
    err = np.max(np.abs(lp - rlp) / np.maximum(1.0, np.abs(rlp)))
    print(f"vector[{cid},B={B},tm={tm},mode={mode}]: max err {err:.2e}, naccept max {st['naccept']} vs oracle {max(nacc)} "
          f"(min {min(nacc)}), path {st['kernel_path']}")
    tol = 1e-4 if cid not in ("C0", "C1") else 1e-3      # tspan (0, 13): rounding is amplified along the flow
    assert err < tol
    assert np.max(np.abs(z - rz)) < 10 * tol
    assert np.allclose(regs, rregs, rtol=10 * tol, atol=1e-4)
    assert abs(st["naccept"] - max(nacc)) <= 1, (st, max(nacc))
    assert st["kernel_path"] == path
    assert st["retcode"] == 0


def test_n4_vector_mode_gradient_is_refused():
    inp = wl.make_inputs("C4", B=32)
    icnf = _vector_icnf_any("C4", 0)
    with pytest.raises(cnf._cabi.CnfError):
        cnf.loss_and_grad(icnf, cnf.TrainMode(), inp["xs"].T, *args_of(inp), eps=inp["eps"].T)


def test_zz_tolerance_report():
    """runs last in this file: the tolerance every parity case above actually used.  north_star asks for 1e-4 / 1e-4 /
    1e-3 AND steps +-1.  One documented widening remains (the conditioning floor of the tspan = (0, 13) configs) and is
    bounded here; cases whose accepted-step count differs by one are held to the same tolerance as the others and are
    listed."""
    if not FLIPS:
        pytest.skip("no parity case ran in this session")
    widened = {k: v for k, v in FLOORS.items() if v > 1e-4}
    flips = [k for k, v in FLIPS.items() if v]
    print(f"{len(FLIPS)} parity cases; tolerance floor above 1e-4 in {len(widened)}: {widened}")
    print(f"accepted-step count differed by one in {len(flips)}: {flips}")
    assert all(v <= FLOOR_MAX for v in FLOORS.values())
