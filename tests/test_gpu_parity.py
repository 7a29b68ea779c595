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


@pytest.mark.parametrize("cid,B", SMALL)
@pytest.mark.parametrize("trace_mode", [0, 1])
@pytest.mark.parametrize("mode", [0, 1])
def test_logdensity(cid, B, trace_mode, mode):
    inp = wl.make_inputs(cid, B=B)
    c = wl.CONFIGS[cid]
    icnf = make_icnf(cid, trace_mode=trace_mode)
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
    if st["naccept"] == rst["naccept"]:
        assert relerr(lp, rlp) < tol
        assert relnorm(np.asarray(z).T, rz) < tol
        assert np.allclose(np.stack(regs, 1), rregs, rtol=1e-4, atol=1e-5)
    else:  # a borderline accept/reject flip: both are valid solves, compare at solver tolerance
        assert relerr(lp, rlp) < 5e-3


@pytest.mark.parametrize("cid,B", [("C0", 64), ("C1", 1024), ("C3", 256), ("C4", 1024), ("C5", 64)])
@pytest.mark.parametrize("trace_mode", [0, 1])
def test_loss_grad(cid, B, trace_mode):
    if cid == "C5" and trace_mode == 1:
        pytest.skip("JacVec training at width 1024: the tensor reverse sweep covers VecJac; the generic one does not fit")
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
    tol_l, tol_g = (max(1e-4, fl_l), max(1e-3, fl_g)) if st["naccept"] == rst["naccept"] else (5e-3, 5e-2)
    assert abs(val - rval) / abs(rval) < tol_l
    assert relnorm(g, rg) < tol_g
    assert relnorm(np.asarray(gx).T, rgx) < tol_g


@pytest.mark.parametrize("cid,B,tm", [("C2", 1024, 1), ("C2", 512, 0), ("C1", 256, 0), ("C4", 128, 0)])
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
    tol_l, tol_g = (max(1e-4, fl_l), max(1e-3, fl_g)) if st["naccept"] == rst["naccept"] else (5e-3, 5e-2)
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
    assert cnf.last_stats(icnf)["kernel_path"] == 2


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
