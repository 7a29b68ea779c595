"""Committed fixtures (tests/golden/*.npz, made by tests/golden/make_golden.py): the FP32 oracle
must still reproduce them on CPU, and the CUDA path must match them on the GPU."""
import glob
import os

import numpy as np
import pytest

from helpers import O, cnf, make_icnf, oracle_cfg, relerr, relnorm, wl

FILES = sorted(f for f in glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*.npz")) if not f.endswith("_ref.npz"))
# outputs of the REAL reference (oracle/dump_reference.jl, run where Julia exists); none are committed yet, so parity
# stays "unpinned" and the tests below skip.  Dropping the files into tests/golden/ turns them on.
REF_FILES = sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "*_ref.npz")))


def _case(path):
    name = os.path.basename(path)[:-4]
    cid, b, m, t = name.split("_")
    return cid, int(b[1:]), int(m[4:]), int(t[2:])


@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f) for f in FILES])
def test_oracle_reproduces_golden(path):
    cid, B, mode, tm = _case(path)
    g = np.load(path)
    c = wl.CONFIGS[cid]
    ys = g["ys"] if "ys" in g else None
    lp, regs, z, st = O.logdensity(oracle_cfg(cid, trace_mode=tm), g["theta"], mode, g["xs"], ys,
                                   g["eps"], *c["tspan"])
    assert st["naccept"] == int(g["naccept"]) and st["nreject"] == int(g["nreject"])
    assert np.allclose(O.last_dts(), g["dts"], rtol=1e-6)
    assert relerr(lp, g["logp32"]) < 1e-6          # against its own earlier output
    assert relerr(lp, g["logp64"]) < 5e-5          # against the float64 restatement
    if mode == 1:
        val, gth, gx, _ = O.loss_grad(oracle_cfg(cid, trace_mode=tm), g["theta"], g["xs"], ys,
                                      g["eps"], *c["tspan"])
        assert abs(val - float(g["loss64"])) / abs(float(g["loss64"])) < 1e-5
        assert relnorm(gth, g["grad_theta64"]) < 1e-4
        assert relnorm(gx, g["grad_x64"]) < 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("path", FILES, ids=[os.path.basename(f) for f in FILES])
def test_cuda_matches_golden(path):
    cid, B, mode, tm = _case(path)
    g = np.load(path)
    ys = g["ys"] if "ys" in g else None
    icnf = make_icnf(cid, trace_mode=tm)
    m = cnf.TrainMode() if mode else cnf.TestMode()
    tail = ((ys.T,) if ys is not None else ()) + (g["theta"], {})
    lp, regs = cnf.inference(icnf, m, g["xs"].T, *tail, eps=g["eps"].T)
    st = cnf.last_stats(icnf)
    assert abs(st["naccept"] - int(g["naccept"])) <= 1
    if st["naccept"] == int(g["naccept"]):
        tol = 1e-4 if cid not in ("C0", "C1") else 1e-3   # tspan (0,13): rounding is amplified
        assert relerr(lp, g["logp64"]) < tol
        if mode == 1:
            val, gth, gx = cnf.loss_and_grad(icnf, m, g["xs"].T, *tail, eps=g["eps"].T, want_grad_x=True)
            assert abs(val - float(g["loss64"])) / abs(float(g["loss64"])) < tol
            assert relnorm(gth, g["grad_theta64"]) < 10 * tol
            assert relnorm(np.asarray(gx).T, g["grad_x64"]) < 10 * tol


def _ref_pairs():
    return [(f, f[:-len("_ref.npz")] + ".npz") for f in REF_FILES]


def test_reference_fixtures_status():
    """states, in the test log, whether parity is pinned by outputs of the real reference"""
    if not REF_FILES:
        pytest.skip("parity unpinned: no tests/golden/*_ref.npz (run oracle/dump_reference.jl on a machine with Julia)")
    for ref, own in _ref_pairs():
        assert os.path.exists(own), f"{ref} has no matching input fixture"


@pytest.mark.parametrize("pair", _ref_pairs(), ids=[os.path.basename(f) for f in REF_FILES])
def test_oracle_matches_reference_dump(pair):
    """the CPU restatement against the real ContinuousNormalizingFlows.jl on the same inputs, weights and eps:
    north_star tolerances (log-density 1e-4, loss 1e-4, accepted steps +-1).  The reference's gradient is a continuous
    adjoint, the oracle's is the discrete one: they agree to O(solver tolerance), reported not asserted."""
    ref, own = pair
    cid, B, mode, tm = _case(own)
    g, r = np.load(own), np.load(ref)
    c = wl.CONFIGS[cid]
    ys = g["ys"] if "ys" in g else None
    lp, regs, z, st = O.logdensity(oracle_cfg(cid, trace_mode=tm), g["theta"], mode, g["xs"], ys, g["eps"], *c["tspan"])
    assert abs(st["naccept"] - int(r["naccept_ref"])) <= 1, (st["naccept"], int(r["naccept_ref"]))
    if "alg_default_ref" in r:
        assert np.all(r["alg_default_ref"] == 1), "the reference's default algorithm left Tsit5 on this case (SURVEY A.1)"
    if st["naccept"] == int(r["naccept_ref"]):
        assert relerr(lp, r["logp_ref"]) < 1e-4
        if mode == 1:
            val, gth, gx, _ = O.loss_grad(oracle_cfg(cid, trace_mode=tm), g["theta"], g["xs"], ys, g["eps"], *c["tspan"])
            assert abs(val - float(r["loss_ref"])) / abs(float(r["loss_ref"])) < 1e-4
            print(f"{cid}: discrete vs reference (continuous-adjoint) gradient, relative: "
                  f"theta {relnorm(gth, r['grad_theta_ref']):.2e}, x {relnorm(gx, r['grad_x_ref'].T):.2e}")


@pytest.mark.gpu
@pytest.mark.parametrize("pair", _ref_pairs(), ids=[os.path.basename(f) for f in REF_FILES])
def test_cuda_matches_reference_dump(pair):
    ref, own = pair
    cid, B, mode, tm = _case(own)
    g, r = np.load(own), np.load(ref)
    ys = g["ys"] if "ys" in g else None
    icnf = make_icnf(cid, trace_mode=tm)
    m = cnf.TrainMode() if mode else cnf.TestMode()
    tail = ((ys.T,) if ys is not None else ()) + (g["theta"], {})
    lp, regs = cnf.inference(icnf, m, g["xs"].T, *tail, eps=g["eps"].T)
    st = cnf.last_stats(icnf)
    assert abs(st["naccept"] - int(r["naccept_ref"])) <= 1
    if st["naccept"] == int(r["naccept_ref"]):
        assert relerr(lp, r["logp_ref"]) < 1e-4
