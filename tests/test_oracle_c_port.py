"""The C + OpenMP restatement (oracle/c) agrees with the Python oracle, for 1 and several threads."""
import numpy as np
import pytest

from oracle import c_port as CP
from oracle import sanity as S
from oracle import vectorizers as V
from tests.util_data import make_real_config, pack_validity


@pytest.mark.parametrize("threads", [1, 3])
@pytest.mark.parametrize("n,ncols", [(1, 2), (257, 6), (5000, 14)])
def test_c_port_real_pipeline_matches_python_oracle(n, ncols, threads):
    x, valid, y = make_real_config(n, ncols, seed=n)
    if ncols > 3:
        valid[3] = False
    vw = pack_validity(valid)
    fills, rows, mean, var, mn, mx, corr = CP.real_pipeline(np.ascontiguousarray(x), vw, y, threads)
    cols = [[float(v) if ok else None for v, ok in zip(c, vc)] for c, vc in zip(x, valid)]
    exp_fill = V.real_vectorizer_fit(cols)
    np.testing.assert_allclose(fills, exp_fill, rtol=1e-13, atol=0)
    exp = V.real_vectorizer_transform(cols, list(fills), True)
    assert np.array_equal(rows[:, :-1], exp) and np.array_equal(rows[:, -1], y)
    st = S.col_stats(rows)
    np.testing.assert_allclose(mean, st.mean, rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(var, st.variance, rtol=1e-10, atol=1e-20)
    assert np.array_equal(mn, st.min) and np.array_equal(mx, st.max)
    if n > 1:
        exp_corr = S.correlations_with_label(rows, st)
        ok = ~np.isnan(exp_corr)
        np.testing.assert_allclose(corr[ok], exp_corr[ok], rtol=1e-9, atol=1e-12)


def test_c_port_label_sums():
    rng = np.random.default_rng(0)
    x = (rng.random((500, 7)) < 0.3).astype(np.float64)
    y = rng.integers(0, 3, 500).astype(np.float64)
    got = CP.label_sums(x, y, np.array([0.0, 1.0, 2.0]), threads=2)
    for k in range(3):
        assert np.array_equal(got[k, :-1], x[y == k].sum(axis=0)) and got[k, -1] == (y == k).sum()
