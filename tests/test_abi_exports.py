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
