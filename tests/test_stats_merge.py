"""VectorStats.merge / LabelInfo.merge: the statistics of a vector over two disjoint row sets combine to those over
their union (what the sampled SanityChecker statistics of a chunk-streamed dataset are built from)."""
import numpy as np


def _stats(x, y):
    from transmogrifai_b200.stats import LabelInfo, VectorStats
    n = len(y)
    mx, my = x.mean(0), y.mean()
    cl = np.unique(y)
    cnt = np.array([(y == c).sum() for c in cl])
    cs = np.stack([x[y == c].sum(0) for c in cl])
    cs[:, 1] = np.nan                                     # a column whose per-class sums are not tracked
    var = x.var(0, ddof=1) if n > 1 else np.zeros(x.shape[1])
    lab = LabelInfo("y", n, np.array([n, my, y.var(ddof=1) if n > 1 else 0.0, y.min(), y.max()]), cl, cnt)
    return VectorStats(lab, mx, var, x.min(0), x.max(0), ((x - mx) * (y - my)[:, None]).sum(0), cs)


def test_merge_of_row_chunks_equals_the_statistics_of_all_rows():
    rng = np.random.default_rng(0)
    x = rng.normal(3, 2, (1000, 4))
    x[:, 3] = (rng.random(1000) < 0.1)
    y = rng.integers(0, 3, 1000).astype(float)
    y[:300] = np.minimum(y[:300], 1)                      # the first chunks do not see every class
    acc = None
    for a, b in ((0, 1), (1, 300), (300, 301), (301, 1000)):
        st = _stats(x[a:b], y[a:b])
        acc = st if acc is None else acc.merge(st)
    ref = _stats(x, y)
    for k in ("mean", "variance", "min", "max", "cov_label"):
        np.testing.assert_allclose(getattr(acc, k), getattr(ref, k), rtol=1e-11, atol=1e-12, err_msg=k)
    np.testing.assert_allclose(acc.label.stats, ref.label.stats, rtol=1e-12)
    assert np.array_equal(acc.label.classes, ref.label.classes)
    assert np.array_equal(acc.label.class_counts, ref.label.class_counts)
    np.testing.assert_allclose(acc.class_sums[:, [0, 2, 3]], ref.class_sums[:, [0, 2, 3]])
    assert np.isnan(acc.class_sums[:, 1]).all()
