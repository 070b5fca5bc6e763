"""RawFeatureFilter: the oracle (oracle/raw_feature_filter.py) AND the product's driver-side logic
(transmogrifai_b200/filters.py: no GPU needed for it) against the reference's own vectors:
  core/src/test/scala/com/salesforce/op/filters/FeatureDistributionTest.scala:50-175
  core/src/test/scala/com/salesforce/op/filters/RawFeatureFilterTest.scala:156-268
  core/src/test/scala/com/salesforce/op/filters/FiltersTestData.scala:38-60"""
import math

import numpy as np
import pytest

from oracle import raw_feature_filter as O

INF = float("inf")
EPS = 1e-2          # FiltersTestData.scala:36

# FiltersTestData.scala:38-60: (name, key, count, nulls, distribution)
TRAIN = [("A", None, 10, 1, [1, 4, 0, 0, 6]), ("B", None, 20, 20, [2, 8, 0, 0, 12]),
         ("C", "1", 10, 1, [1, 4, 0, 0, 6]), ("C", "2", 20, 19, [2, 8, 0, 0, 12]),
         ("D", "1", 10, 9, [1, 4, 0, 0, 6]), ("D", "2", 20, 19, [2, 8, 0, 0, 12])]
SCORE = [("A", None, 10, 8, [1, 4, 0, 0, 6]), ("B", None, 20, 20, [2, 8, 0, 0, 12]),
         ("C", "1", 10, 1, [0, 0, 10, 10, 0]), ("C", "2", 20, 19, [2, 8, 0, 0, 12]),
         ("D", "1", 0, 0, [0, 0, 0, 0, 0]), ("D", "2", 0, 0, [0, 0, 0, 0, 0])]


class _OracleImpl:
    name = "oracle"

    @staticmethod
    def fd(name, key, count, nulls, dist, type_="Training"):
        return O.FeatureDistribution(name, key, count, nulls, [float(x) for x in dist], [], type_)

    @staticmethod
    def exclude(train, score, corr, bins, min_fill, max_fill_diff, max_fill_ratio, max_js, max_corr):
        p = O.FilterParams(bins=bins, min_fill=min_fill, max_fill_difference=max_fill_diff,
                           max_fill_ratio_diff=max_fill_ratio, max_js_divergence=max_js, max_correlation=max_corr)
        m, r, drop, keys = O.get_features_to_exclude(train, score, corr, p)
        metrics = [(x.name, x.key, x.training_fill_rate, x.training_null_label_absolute_corr, x.scoring_fill_rate,
                    x.js_divergence, x.fill_rate_diff, x.fill_ratio_diff) for x in m]
        reasons = [(x.name, x.key, x.training_unfilled_state, x.training_null_label_leaker, x.scoring_unfilled_state,
                    x.js_divergence_mismatch, x.fill_rate_diff_mismatch, x.fill_ratio_diff_mismatch, x.excluded)
                   for x in r]
        return metrics, reasons, drop, keys

    fill_rate = staticmethod(lambda d: d.fill_rate())
    rel_rate = staticmethod(lambda a, b: a.relative_fill_rate(b))
    rel_ratio = staticmethod(lambda a, b: a.relative_fill_ratio(b))
    js = staticmethod(lambda a, b: a.js_divergence(b))
    reduce = staticmethod(lambda a, b: a.reduce(b))


class _ProductImpl:
    name = "product"

    @staticmethod
    def fd(name, key, count, nulls, dist, type_="Training"):
        from transmogrifai_b200.filters import FeatureDistribution
        return FeatureDistribution(name, key, count, nulls, np.asarray(dist, dtype=np.float64), np.zeros(0), type_)

    @staticmethod
    def exclude(train, score, corr, bins, min_fill, max_fill_diff, max_fill_ratio, max_js, max_corr):
        from transmogrifai_b200.filters import RawFeatureFilter
        f = RawFeatureFilter(bins, min_fill, max_fill_diff, max_fill_ratio, max_js, max_corr)
        m, r, drop, keys = f.getFeaturesToExclude(train, score, corr)
        metrics = [(x.name, x.key, x.trainingFillRate, x.trainingNullLabelAbsoluteCorr, x.scoringFillRate,
                    x.jsDivergence, x.fillRateDiff, x.fillRatioDiff) for x in m]
        reasons = [(x.name, x.key, x.trainingUnfilledState, x.trainingNullLabelLeaker, x.scoringUnfilledState,
                    x.jsDivergenceMismatch, x.fillRateDiffMismatch, x.fillRatioDiffMismatch, x.excluded) for x in r]
        return metrics, reasons, drop, keys

    fill_rate = staticmethod(lambda d: d.fillRate())
    rel_rate = staticmethod(lambda a, b: a.relativeFillRate(b))
    rel_ratio = staticmethod(lambda a, b: a.relativeFillRatio(b))
    js = staticmethod(lambda a, b: a.jsDivergence(b))
    reduce = staticmethod(lambda a, b: a.reduce(b))


IMPLS = [_OracleImpl, _ProductImpl]


def _sets(impl):
    return ([impl.fd(*t) for t in TRAIN], [impl.fd(*t, "Scoring") for t in SCORE])


@pytest.mark.parametrize("impl", IMPLS, ids=lambda i: i.name)
def test_fill_rates_js_divergence_reduce(impl):
    """FeatureDistributionTest.scala:140-175."""
    train, score = _sets(impl)
    fd1, fd2 = impl.fd("A", None, 10, 1, []), impl.fd("A", None, 20, 20, [])
    assert impl.rel_rate(fd1, fd2) == 0.9
    assert impl.rel_ratio(train[0], score[0]) == 4.5
    assert impl.rel_ratio(train[2], score[2]) == 1.0
    assert impl.rel_ratio(fd1, impl.fd("A", None, 20, 19, [])) == 18.0
    a, b = impl.fd("A", None, 10, 1, [1, 4, 0, 0, 6]), impl.fd("A", None, 20, 20, [2, 8, 0, 0, 12])
    assert impl.js(a, b) < EPS
    fd3 = impl.fd("A", None, 10, 1, [0, 0, 1000, 1000, 0])
    assert impl.js(fd3, fd3) < EPS
    fd4 = impl.fd("A", None, 20, 20, [200, 800, 0, 0, 1200])
    assert abs(impl.js(fd3, fd4) - 1.0) < EPS
    r = impl.reduce(a, b)
    assert (r.count, r.nulls, list(r.distribution)) == (30, 21, [3.0, 12.0, 0.0, 0.0, 18.0])


@pytest.mark.parametrize("impl", IMPLS, ids=lambda i: i.name)
def test_metrics_on_the_reference_summaries(impl):
    """RawFeatureFilterTest.scala:156-196."""
    train, score = _sets(impl)
    m, _r, _d, _k = impl.exclude(train, [], {}, 10, 0.2, 1.0, INF, 1.0, 1.0)
    assert [x[2] for x in m] == [0.9, 0.0, 0.9, 0.05, 0.1, 0.05]
    assert all(x[3] is None and x[4] is None and x[5] is None and x[6] is None and x[7] is None for x in m)
    m, _r, _d, _k = impl.exclude(train, score, {}, 10, 0.2, 1.0, INF, 1.0, 1.0)
    assert [x[0] for x in m] == ["A", "B", "C", "C", "D", "D"]
    assert [x[1] for x in m] == [None, None, "1", "2", "1", "2"]
    assert [x[2] for x in m] == [0.9, 0.0, 0.9, 0.05, 0.1, 0.05]
    assert [x[4] for x in m] == [0.2, 0.0, 0.9, 0.05, 0.0, 0.0]
    assert [x[5] for x in m[:4]] == pytest.approx([0.0, 0.0, 1.0, 0.0], abs=1e-12)
    assert all(math.isnan(x[5]) for x in m[4:])
    assert [x[6] for x in m] == pytest.approx([0.7, 0.0, 0.0, 0.0, 0.1, 0.05], abs=1e-15)
    assert [x[7] for x in m] == [4.5, INF, 1.0, 1.0, INF, INF]


@pytest.mark.parametrize("impl", IMPLS, ids=lambda i: i.name)
def test_exclusions_on_the_reference_summaries(impl):
    """RawFeatureFilterTest.scala:198-268."""
    train, score = _sets(impl)
    # training fill rate only (:198-210)
    _m, r, drop, keys = impl.exclude(train, [], {}, 10, 0.2, 1.0, INF, 1.0, 1.0)
    assert set(drop) == {"B", "D"} and keys == {"C": {"2"}}
    assert {x[0] for x in r if x[2]} == {"B", "C", "D"}
    assert [x[0] for x in r] == ["A", "B", "C", "C", "D", "D"]
    # training and scoring fill rate (:212-224)
    _m, r, drop, keys = impl.exclude(train, score, {}, 10, 0.2, 1.0, INF, 1.0, 1.0)
    assert set(drop) == {"B", "D"} and keys == {"C": {"2"}}
    assert {x[0] for x in r if x[2] or x[4]} == {"B", "C", "D"}
    # relative fill rate (:226-235)
    _m, r, drop, keys = impl.exclude(train, score, {}, 10, 0.0, 0.5, INF, 1.0, 1.0)
    assert set(drop) == {"A"} and keys == {}
    assert {x[0] for x in r if x[6]} == {"A"}
    # fill rate ratio (:237-245)
    _m, r, drop, keys = impl.exclude(train, score, {}, 10, 0.0, 1.0, 2.0, 1.0, 1.0)
    assert set(drop) == {"D", "A", "B"} and keys == {}
    assert {x[0] for x in r if x[7]} == {"A", "B", "D"}
    # JS distance (:247-257)
    _m, r, drop, keys = impl.exclude(train, score, {}, 10, 0.0, 1.0, INF, 0.5, 1.0)
    assert drop == [] and keys == {"C": {"1"}}
    assert {x[0] for x in r if x[8]} == {"C"}
    # everything (:259-268)
    _m, r, drop, keys = impl.exclude(train, score, {}, 10, 0.1, 0.5, INF, 0.5, 1.0)
    assert set(drop) == {"A", "B", "C", "D"} and keys == {}
    assert {x[0] for x in r if x[8]} == {"A", "B", "C", "D"}


def test_oracle_from_summary_on_the_reference_rows():
    """FeatureDistributionTest.scala:50-89 (features) and :111-138 (the numeric and boolean map entries)."""
    bins = 10
    values = [("R", [1.0]), None, ("L", ["male", "female"]), None, ("R", [1.0, 3.0, 5.0])]
    summaries = [O.Summary(0.0, 1.0, 6.0, 10), O.Summary(-1.6, 10.6, 3.0, 10), O.Summary(0.0, 3.0, 7.0, 10),
                 O.Summary(0.0, 0.0, 5.0, 10), O.Summary(1.0, 5.0, 10.0, 10)]
    ds = [O.from_summary((f"f{i}", None), s, v, bins, "Training") for i, (s, v) in enumerate(zip(summaries, values))]
    for d in ds:
        assert d.key is None and d.count == 1 and len(d.distribution) == bins
    assert ds[0].nulls == 0 and ds[1].nulls == 1 and sum(ds[1].distribution) == 0
    assert sum(ds[2].distribution) == 2 and sorted(ds[2].summary_info) == sorted([0.0, 3.0, 7.0, 10.0])
    assert sum(ds[3].distribution) == 0
    assert sum(ds[4].distribution) == 3 and len(ds[4].summary_info) == bins
    d = O.from_summary(("numericMap", "A"), O.Summary(-1.6, 10.6, 30.0, 10), ("R", [1.0]), bins, "Scoring")
    assert len(d.summary_info) == bins and sum(d.distribution) == 1 and len(d.distribution) == bins
    d = O.from_summary(("booleanMap", "B"), O.Summary(0.0, 0.0, 0.0, 10), ("R", [0.0]), bins, "Scoring")
    assert len(d.distribution) == 2 and d.distribution == [1.0, 0.0] and len(d.summary_info) == 4
    # text: 100 bins, every token lands in exactly one of them (:91-109)
    rng = np.random.default_rng(5)
    toks = ["".join(rng.choice(list("abcdefghij"), size=rng.integers(1, 10))) for _ in range(10000)]
    d = O.from_summary(("description", None), O.Summary(1000.0, 50000.0, 70000.0, 10), ("L", toks), 100, "Training")
    assert len(d.distribution) == 100 and sum(d.distribution) == 10000


def test_oracle_bucketize_and_summary():
    """NumericBucketizer.scala:231-258 (left-inclusive, invalid bucket) and SummaryTest.scala:43-52."""
    splits = [0.0, 1.0, 2.0, 3.0]
    assert [O.bucketize_left(splits, v) for v in (0.0, 0.5, 1.0, 2.999, 3.0, -0.1, float("nan"), INF)] == \
        [0, 0, 1, 2, 3, 3, 3, 3]
    s = O.summary_of(("R", [0.1, -0.2, 3.0]))
    assert (s.min, s.max, s.count) == (-0.2, 3.0, 3.0) and abs(s.sum - 2.9) < 1e-15
    s = O.summary_of(("L", ["a", "b", "c", "d"]))
    assert (s.min, s.max, s.sum, s.count) == (4.0, 4.0, 4.0, 1.0)
    assert O.Summary().is_empty()


def test_host_functions_of_the_library_match_the_oracle():
    """tmog_rff_numeric_splits / _js_divergence / _null_label_corr against the oracle's restatement."""
    import ctypes as C
    from transmogrifai_b200 import _lib as L
    rng = np.random.default_rng(11)
    for bins in (2, 3, 10, 100, 1001):
        for _ in range(20):
            lo = float(rng.normal(0, 10.0 ** rng.integers(-3, 6)))
            hi = lo + float(abs(rng.normal(0, 10.0 ** rng.integers(-6, 6)))) + 1e-300
            out = np.zeros(bins)
            L.call("tmog_rff_numeric_splits", lo, hi, bins, out.ctypes.data)
            with np.errstate(all="ignore"):
                step = np.float64(hi - lo) / np.float64(bins - 2.0)
                exp = np.array([np.float64(lo) + step * np.float64(b) for b in range(bins)])
            assert np.array_equal(out, exp, equal_nan=True)
    for _ in range(50):
        k = int(rng.integers(1, 40))
        a = rng.integers(0, 5, k).astype(np.float64) * (rng.random(k) < 0.7)
        b = rng.integers(0, 50, k).astype(np.float64) * (rng.random(k) < 0.7)
        out = C.c_double()
        L.call("tmog_rff_js_divergence", a.ctypes.data, b.ctypes.data, k, C.byref(out))
        exp = O.FeatureDistribution("x", None, 1, 0, list(a), []).js_divergence(
            O.FeatureDistribution("x", None, 1, 0, list(b), []))
        assert (math.isnan(exp) and math.isnan(out.value)) or abs(out.value - exp) < 1e-12
    # correlation of null indicators with the response from sums
    n = 500
    y = rng.normal(3, 2, n)
    ind = [(rng.random(n) < p).astype(np.float64) for p in (0.0, 0.1, 0.5, 1.0)]
    ind.append((y > 4).astype(np.float64))
    exp = O.pearson_columns([list(y)] + [list(c) for c in ind])
    shift = 2.5
    nc = np.array([c.sum() for c in ind])
    ny = np.array([((y - shift) * c).sum() for c in ind])
    out = np.zeros(len(ind))
    L.call("tmog_rff_null_label_corr", float(n), float((y - shift).sum()), float(((y - shift) ** 2).sum()), len(ind),
           nc.ctypes.data, ny.ctypes.data, out.ctypes.data)
    for j in range(len(ind)):
        e = exp[0][j + 1]
        e = e if math.isnan(e) else min(abs(e), 1.0)
        assert (math.isnan(e) and math.isnan(out[j])) or abs(out[j] - e) < 1e-9


def test_js_divergence_properties():
    """Size-independent properties of tmog_rff_js_divergence (FeatureDistribution.jsDivergence, :154-163): symmetric,
    within [0, 1] (log base 2), 0 for proportional histograms, 1 for disjoint ones, blind to bins empty on both sides."""
    import ctypes as C
    from transmogrifai_b200 import _lib as L
    rng = np.random.default_rng(23)

    def js(a, b):
        a, b = np.ascontiguousarray(a, dtype=np.float64), np.ascontiguousarray(b, dtype=np.float64)
        out = C.c_double()
        L.call("tmog_rff_js_divergence", a.ctypes.data, b.ctypes.data, len(a), C.byref(out))
        return out.value

    for _ in range(200):
        k = int(rng.integers(2, 300))
        a = rng.integers(0, 1000, k).astype(np.float64) * (rng.random(k) < 0.8)
        b = rng.integers(0, 1000, k).astype(np.float64) * (rng.random(k) < 0.8)
        if a.sum() == 0 or b.sum() == 0:
            continue
        d = js(a, b)
        assert -1e-12 <= d <= 1.0 + 1e-12
        assert abs(d - js(b, a)) < 1e-12
        assert abs(js(a, 7.0 * a)) < 1e-12
        pad = np.zeros(5)
        assert abs(js(np.concatenate([a, pad]), np.concatenate([b, pad])) - d) < 1e-12
        lo, hi = np.concatenate([a + 1, np.zeros(k)]), np.concatenate([np.zeros(k), b + 1])
        assert abs(js(lo, hi) - 1.0) < 1e-12
