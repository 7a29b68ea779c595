"""Host-side mirror of the reference interface (no GPU needed): construct's defaults and
validation (src/base_icnf.jl:1-77), parameter layout, steer_tspan, array marshalling, sharding."""
import numpy as np
import pytest

from helpers import cnf, wl


def chain(*dims):
    return cnf.Chain(*[cnf.Dense(dims[i], dims[i + 1], "tanh") for i in range(len(dims) - 1)])


def test_tag_defaults_like_the_reference():
    r = cnf.construct(cnf.RNODE, chain(2, 6, 2), 1, 1)
    assert r.lambda1 == pytest.approx(1e-2) and r.lambda2 == pytest.approx(1e-2) and r.lambda3 == 0
    assert not r.cond and r.tspan == (0.0, 1.0)
    f = cnf.construct(cnf.FFJORD, chain(2, 6, 2), 2)
    assert f.lambda1 == 0 and f.lambda2 == 0
    c = cnf.construct(cnf.CondRNODE, chain(4, 6, 2), 2)
    assert c.cond and c.lambda1 == pytest.approx(1e-2)
    cfg = c._config()
    assert cfg.C == 2 and cfg.D == 2 and cfg.flags & cnf._cabi.FLAG_COND
    assert isinstance(r.compute_mode, cnf.DIVecJacMatrixMode)


def test_flags_follow_type_parameters():
    i = cnf.construct(cnf.RNODE, chain(3, 5, 3), 2, 1, lambda3=1e-2, steer_rate=0.1)
    fl = i._config().flags
    for bit in (cnf._cabi.FLAG_AUGMENTED, cnf._cabi.FLAG_NORM_Z, cnf._cabi.FLAG_NORM_J, cnf._cabi.FLAG_NORM_Z_AUG):
        assert fl & bit
    assert not fl & cnf._cabi.FLAG_COND


def test_unsupported_combinations_raise():
    with pytest.raises(ValueError):
        cnf.construct(cnf.RNODE, chain(2, 2), 2, data_type=np.float64)
    with pytest.raises(ValueError):
        cnf.construct(cnf.RNODE, chain(2, 2), 2, compute_mode=object())
    with pytest.raises(ValueError):
        cnf.construct(cnf.RNODE, chain(2, 2), 2, resource=object())
    with pytest.raises(ValueError):
        cnf.Dense(2, 2, "relu")
    with pytest.raises(ValueError):
        cnf.construct(cnf.RNODE, chain(3, 3), 2)            # widths do not fit nvars
    with pytest.raises(ValueError):
        cnf.construct(cnf.RNODE, chain(2, 2), 2, sol_kwargs={"alg": "Vern7"})._config()
    with pytest.raises(ValueError):
        cnf.construct(cnf.RNODE, chain(2, 2), 2, sol_kwargs={"bogus": 1})._config()


def test_sol_kwargs_are_read_and_ignorable_ones_ignored():
    i = cnf.construct(cnf.RNODE, chain(2, 2), 2, sol_kwargs=dict(reltol=1e-4, abstol=1e-7, maxiters=2 ** 31 - 1,
                                                                  progress=True, save_everystep=False, alg="Tsit5"))
    c = i._config()
    assert c.reltol == pytest.approx(1e-4) and c.abstol == pytest.approx(1e-7) and c.maxiters == 2 ** 31 - 1


def test_param_layout_is_lux_componentarray_order():
    w, b, n = cnf.param_offsets([2, 6, 2])
    assert list(w) == [0, 18] and list(b) == [12, 30] and n == 32
    i = cnf.construct(cnf.RNODE, chain(2, 6, 2), 1, 1)
    ps, st = cnf.setup(np.random.default_rng(0), i)
    assert ps.dtype == np.float32 and ps.size == 32 and st == {}
    assert np.all(ps[12:18] == 0) and np.all(ps[30:] == 0)
    assert wl.nparams([128, 1024, 1024, 128]) == 1312896


def test_steer_tspan_only_in_train_mode():
    i = cnf.construct(cnf.RNODE, chain(2, 2), 2, tspan=(0.0, 13.0), steer_rate=0.1, rng=np.random.default_rng(3))
    assert cnf.steer_tspan(i, cnf.TestMode()) == (0.0, 13.0)
    t0, t1 = cnf.steer_tspan(i, cnf.TrainMode())
    assert t0 == 0.0 and 13.0 * 0.9 <= t1 <= 13.0 * 1.1 and t1 != 13.0
    j = cnf.construct(cnf.RNODE, chain(2, 2), 2, tspan=(0.0, 13.0))
    assert cnf.steer_tspan(j, cnf.TrainMode()) == (0.0, 13.0)


def test_matrix_marshalling_is_column_major():
    from cnfb200.api import _Mat
    a = np.arange(6, dtype=np.float32).reshape(2, 3)      # rows x B, Julia orientation
    m = _Mat(a, 2, "a")
    assert m.B == 3 and list(m.keep.ravel()) == [0, 3, 1, 4, 2, 5]   # column-major memory image
    with pytest.raises(ValueError):
        _Mat(a, 3, "a")


def test_shard_ranges_cover_the_batch():
    for B, W in ((10, 3), (1 << 20, 8), (7, 8)):
        r = [cnf.parallel.shard_range(B, k, W) for k in range(W)]
        assert r[0][0] == 0 and r[-1][1] == B
        assert all(r[k][1] == r[k + 1][0] for k in range(W - 1))


def test_workload_flop_model_matches_survey():
    assert wl.flops_per_eval("C2") == 26112
    assert wl.flops_per_eval("C3") == 327680
    assert wl.flops_per_eval("C4") == 75776
    assert wl.flops_per_eval("C5") == 5242880
    assert wl.flops_backward_per_eval("C3") == 655360


def test_pinned_output_pool_recycles_buffers_only_after_every_view_is_gone():
    """host mirror: results land in page-locked buffers from a pool (cnfb200_host_alloc); a buffer may only be
    handed out again once the array AND its views (the `rows x B` orientation returned to the caller) are dead"""
    import gc
    pool = cnf.api._PinnedPool()
    backing = [np.zeros(1 << 16, np.float32) for _ in range(2)]   # stand-ins for cudaHostAlloc memory
    handed = []

    def take(nbytes):
        free = pool.free.get(nbytes)
        if free:
            return free.pop()
        handed.append(backing[len(handed)].ctypes.data)
        return handed[-1]

    pool.take = take
    a = pool.empty((1 << 14, 4))
    a[:] = 3.0
    assert a.flags["C_CONTIGUOUS"] and backing[0][0] == 3.0 and a.ctypes.data == handed[0]
    view = a.T
    del a
    gc.collect()
    assert not pool.free                       # the view still owns the buffer
    b = pool.empty((1 << 14, 4))
    assert b.ctypes.data == handed[1]          # so a second buffer is handed out
    del view
    gc.collect()
    assert pool.free == {1 << 18: [handed[0]]}
    c = pool.empty((1 << 14, 4))
    assert c.ctypes.data == handed[0]          # recycled
    assert pool._cls(3 << 16) == 1 << 18 and pool._cls(1 << 18) == 1 << 18   # size classes: powers of two
    pool.MAX_FREE_BYTES = 0                    # over the cap a returned buffer is freed, not kept
    freed = []
    orig_load = cnf._cabi.load

    class _Lib:
        def cnfb200_host_free(self, addr):
            freed.append(addr)

    cnf._cabi.load = lambda: _Lib()
    try:
        pool.give(1 << 18, 12345)
    finally:
        cnf._cabi.load = orig_load
    assert freed == [12345] and 12345 not in pool.free.get(1 << 18, [])
    small = pool.empty((8, 4))                 # small results stay ordinary numpy arrays
    assert small.base is None and len(handed) == 2
    # without a CUDA device cnfb200_host_alloc refuses and the pool falls back to pageable memory
    assert cnf.api._pinned.empty((1 << 14, 4)).shape == (1 << 14, 4)


def test_algorithmic_bytes_of_the_tensor_path():
    """bench.py's HBM roofline figures: the exact-trace evaluation is the forward chain plus D tangent chains"""
    hutch, exact = wl.tc_bytes_per_eval("C4"), wl.tc_bytes_per_eval_exact("C4")
    assert hutch == 6304 and exact == 27520
    assert exact > 8 * 2 * 4 * 128             # at least the tau_1 planes of the 8 passes
