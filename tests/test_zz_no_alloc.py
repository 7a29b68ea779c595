"""cnfb200_augmented_f_static (csrc/cnf_noalloc.cu): the allocation-free variant of cnfb200_augmented_f, whose work
planes live in one file-scope device array, against the existing entry point — through the C ABI, on two calls in a
row (the second with a different, smaller batch, so anything the first call left in the shared array would show).
Collected last on purpose; passes alone in a fresh process and after the others."""
import numpy as np
import pytest

from helpers import args_of, cnf, make_icnf, wl

pytestmark = pytest.mark.gpu

# (config, mode, tolerance relative to the row maximum): narrow nets run the same generic FP32 kernels under both
# entry points (bit-identical); the existing entry point evaluates wide TrainMode nets on the tensor-core path
# (operand splitting), so those carry the tolerance of test_gpu_parity.test_augmented_f
CASES = [("C1", 0, 0.0), ("C1", 1, 0.0), ("C2", 0, 0.0), ("C0", 1, 0.0), ("C4", 1, 1e-4), ("C4", 0, 2e-5)]


def _one_call(torch, icnf, cid, mode, B, seed):
    c = wl.CONFIGS[cid]
    D = c["nvars"] + c["naug"]
    R = D + (3 if mode else 1)
    inp = wl.make_inputs(cid, B=B, seed=seed)
    rng = np.random.default_rng(100 + seed)
    u = rng.normal(0, 1, (B, R)).astype(np.float32)
    m = cnf.TrainMode() if mode else cnf.TestMode()
    want = np.asarray(cnf.augmented_f(icnf, m, u.T, *args_of(inp), eps=inp["eps"].T)).T   # the existing entry point
    dev = torch.device("cuda:0")
    ud = torch.from_numpy(u).to(dev)
    ed = torch.from_numpy(np.ascontiguousarray(inp["eps"], dtype=np.float32)).to(dev) if mode else None
    yd = torch.from_numpy(np.ascontiguousarray(inp["ys"], dtype=np.float32)).to(dev) if inp["ys"] is not None else None
    out = torch.full((B, R), float("nan"), dtype=torch.float32, device=dev)
    tail = ((yd,) if yd is not None else ()) + (inp["theta"], {})
    got = cnf.augmented_f_static(icnf, m, ud, *tail, eps=ed, out=out).cpu().numpy()
    return got, want


@pytest.mark.parametrize("cid,mode,tol", CASES)
def test_static_scratch_variant_matches_augmented_f_twice_in_a_row(cid, mode, tol):
    import torch
    icnf = make_icnf(cid)
    for B, seed in ((1000, 1), (300, 2)):       # ragged sizes: padding columns of the 256-sample stride are exercised
        got, want = _one_call(torch, icnf, cid, mode, B, seed)
        assert np.isfinite(got).all()
        if tol == 0.0:
            assert np.array_equal(got, want), f"{cid} mode {mode} B {B}: max diff {np.abs(got - want).max()}"
        else:
            scale = np.maximum(np.abs(want).max(axis=0, keepdims=True), 1e-6)
            assert np.max(np.abs(got - want) / scale) < tol


def test_static_scratch_variant_rejects_what_it_cannot_serve():
    import torch
    icnf = make_icnf("C1")
    inp = wl.make_inputs("C1", B=8)
    m = cnf.TestMode()
    dev = torch.device("cuda:0")
    R = 3
    # beyond the stated maximum: (2D + C + R) * round_up(B, 256) floats > 16 MiB
    per = 2 * 2 + 0 + R
    Bbig = cnf._cabi.STATIC_SCRATCH_BYTES // 4 // per + 256
    u = torch.zeros((Bbig, R), dtype=torch.float32, device=dev)
    out = torch.empty_like(u)
    with pytest.raises(cnf._cabi.CnfError, match="exceeds the static scratch"):
        cnf.augmented_f_static(icnf, m, u, inp["theta"], {}, out=out)
    # a host pointer would need a staging allocation: refused by the library itself
    lib = cnf._cabi.load()
    h = icnf.handle()
    host = np.zeros((8, R), np.float32)
    rc = lib.cnfb200_augmented_f_static(h, 0, host.ctypes.data, None, None, 8, out.data_ptr())
    assert rc == -5 and b"device pointers" in lib.cnfb200_last_error(h)
    # and the largest batch that fits still works
    Bfit = cnf._cabi.STATIC_SCRATCH_BYTES // 4 // per // 256 * 256
    u = torch.randn((Bfit, R), dtype=torch.float32, device=dev)
    out = torch.empty_like(u)
    got = cnf.augmented_f_static(icnf, m, u, inp["theta"], {}, out=out)
    want = cnf.augmented_f(icnf, m, u.t(), inp["theta"], {})
    assert torch.equal(got, want.t().contiguous() if hasattr(want, "t") else torch.as_tensor(np.asarray(want).T))
