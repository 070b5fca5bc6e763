"""CPU-side checks of the C-ABI library: it loads, and exports every symbol include/tmog_b200.h declares."""
import ctypes

import numpy as np


def test_library_loads_and_exports_every_declared_symbol():
    from transmogrifai_b200 import _lib as L
    lib = L.lib()
    declared = L.declared_symbols()
    assert "tmog_real_fit_partials_dev" in declared and "tmog_last_error" in declared
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, f"declared in the header but not exported: {missing}"
    bound = set(L._SIGNATURES) | {"tmog_last_error", "tmog_abi_version"}
    assert set(declared) == bound, f"header and ctypes binding disagree: {set(declared) ^ bound}"
    assert lib.tmog_abi_version() >= 1


def test_host_only_entry_points_work_without_gpu():
    """tmog_real_partials_merge / tmog_real_finish are host code: exercise them against hand-computed sums."""
    from transmogrifai_b200 import _lib as L
    # one column: x = [4, null, 2], y = [1, 0, 1]; shifts 0
    part = np.array([2, 6, 20, 6, 2, 2, 4, 3,      # column: n_valid, sum d, sum d^2, sum d*e, sum_valid e, min, max, n
                     3, 2, 2, 2, 2, 0, 1, 3], dtype=np.float64)   # label
    zero = part.copy()
    zero[[0, 1, 2, 3, 4, 7, 8, 9, 10, 11, 12, 15]] = 0
    zero[[5, 13]] = np.inf
    zero[[6, 14]] = -np.inf
    L.call("tmog_real_partials_merge", zero.ctypes.data, part.ctypes.data, 1)
    assert np.array_equal(zero, part)
    shifts = np.zeros(2)
    fills, mean, var, mn, mx, cov = (np.zeros(k) for k in (1, 2, 2, 2, 2, 2))
    lab, nls = np.zeros(5), np.zeros(1)
    L.call("tmog_real_finish", part.ctypes.data, shifts.ctypes.data, 1, 1, 0.0, 1, fills.ctypes.data,
           mean.ctypes.data, var.ctypes.data, mn.ctypes.data, mx.ctypes.data, cov.ctypes.data, lab.ctypes.data,
           nls.ctypes.data)
    assert fills[0] == 3.0
    f = np.array([4.0, 3.0, 2.0]); q = np.array([0.0, 1.0, 0.0]); y = np.array([1.0, 0.0, 1.0])
    np.testing.assert_allclose(mean, [f.mean(), q.mean()], rtol=1e-15)
    np.testing.assert_allclose(var, [f.var(ddof=1), q.var(ddof=1)], rtol=1e-15)
    assert mn.tolist() == [2.0, 0.0] and mx.tolist() == [4.0, 1.0]
    np.testing.assert_allclose(cov, [((f - f.mean()) * (y - y.mean())).sum(), ((q - q.mean()) * (y - y.mean())).sum()],
                               rtol=1e-14, atol=1e-15)
    np.testing.assert_allclose(lab, [3, y.mean(), y.var(ddof=1), 0, 1], rtol=1e-15)
    assert nls[0] == 0.0
    rc = L.lib().tmog_real_partials_merge(None, None, 1)
    assert rc == -1 and b"bad arguments" in L.lib().tmog_last_error()


def test_sanity_decide_wide_vector_matches_the_rules():
    """tmog_sanity_decide on a vector wide enough for its threaded path (host code, no GPU): the variance and
    correlation rules per column and the parent-level correlation rule, against numpy."""
    import numpy as np
    from transmogrifai_b200 import _lib as L
    rng = np.random.default_rng(0)
    d, nparents = 70_000, 700
    variance = rng.random(d)
    variance[rng.integers(0, d, 500)] = 0.0
    corr = rng.uniform(-1, 1, d + 1)
    corr[rng.integers(0, d, 300)] = np.nan
    parent = (np.arange(d) // 100).astype(np.int32)               # runs of 100 columns share a parent
    offs = np.arange(d + 1, dtype=np.int32)
    pos = np.arange(d, dtype=np.int32)
    gid = np.full(d, -1, dtype=np.int32)
    tsh = np.zeros(d, dtype=np.uint8)
    cv = np.full(d, np.nan)
    rcount = np.full(d, -1, dtype=np.int32)
    rconf, rsup = np.zeros(2 * d), np.zeros(2 * d)
    params = np.array([1e-5, 0.1, 0.9, 0.99, 0.95, 1.0, 1.0, 1.0, 0.0])
    drop = np.zeros(d, dtype=np.uint8)
    reasons = np.zeros(d, dtype=np.uint32)
    pc, pv = np.zeros(d), np.zeros(d)
    L.call("tmog_sanity_decide", d, variance.ctypes.data, corr.ctypes.data, None, 0, pos.ctypes.data, offs.ctypes.data,
           parent.ctypes.data, gid.ctypes.data, tsh.ctypes.data, cv.ctypes.data, rcount.ctypes.data, rconf.ctypes.data,
           rsup.ctypes.data, params.ctypes.data, drop.ctypes.data, reasons.ctypes.data, pc.ctypes.data, pv.ctypes.data)
    ac = np.where(np.isnan(corr[:d]), 0.0, np.abs(corr[:d]))
    exp_pc = np.repeat(ac.reshape(nparents, 100).max(axis=1), 100)
    assert np.array_equal(pc, exp_pc) and np.isnan(pv).all()
    with np.errstate(invalid="ignore"):
        exp = (variance <= 1e-5) * 1 | (np.abs(corr[:d]) < 0.1) * 2 | (np.abs(corr[:d]) > 0.9) * 4 | (exp_pc > 0.9) * 256
    assert np.array_equal(reasons, exp.astype(np.uint32))
    assert np.array_equal(drop, (exp != 0).astype(np.uint8))
