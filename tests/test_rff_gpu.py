"""RawFeatureFilter on the GPU (transmogrifai_b200/filters.py over csrc/rff.cu and the tokenise + hash kernels) against
the oracle's per-row restatement (oracle/raw_feature_filter.py) on the same seeded rows: summaries, distributions
(bit-exact counts), summaryInfo, the null-indicator / response correlations, metrics and exclusions -- one batch and
uneven row chunks, few and many bins (shared-memory and global histograms, both tokenise + hash paths), NaN / Inf values,
scoring values outside the training range; and at a size the oracle cannot reach, against numpy and by properties."""
import numpy as np
import pytest

from oracle import raw_feature_filter as O
from tests.test_rff_host_logic import check_info, chunked, features_of, make_data

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("bins,sizes", [(10, [300]), (100, [700, 1, 999, 300]), (600, [256, 144]), (2500, [150]),
                                        (2, [90, 30])])
def test_feature_stats_and_exclusions_match_the_oracle(bins, sizes):
    from transmogrifai_b200.filters import RawFeatureFilter
    n = sum(sizes)
    train, score = make_data(n, 3), make_data(n, 4, shift=True)
    feats = features_of(train)
    kw = dict(minFill=0.2, maxFillDifference=0.3, maxFillRatioDiff=2.0, maxJSDivergence=0.5, maxCorrelation=0.6)
    f = RawFeatureFilter(bins=bins, minScoringRows=10, **kw)
    tr_chunks, sc_chunks = chunked(train, sizes), chunked(score, sizes)
    got_t = f.computeFeatureStats(tr_chunks if len(sizes) > 1 else tr_chunks[0], feats, "Training")
    exp_t = O.compute_feature_stats(train, ["y"], bins, "Training")
    check_info(got_t, exp_t, n)
    got_s = f.computeFeatureStats(sc_chunks, feats, "Scoring", got_t)
    exp_s = O.compute_feature_stats(score, ["y"], bins, "Scoring", exp_t)
    check_info(got_s, exp_s, n)
    out = f.generateFilteredRaw(tr_chunks, feats, sc_chunks)
    p = O.FilterParams(bins=bins, min_fill=0.2, max_fill_difference=0.3, max_fill_ratio_diff=2.0,
                       max_js_divergence=0.5, max_correlation=0.6, min_scoring_rows=10)
    _ts, _ss, m, r, drop, _keys = O.generate_filtered_raw(train, score, ["y"], p)
    assert sorted(x.name for x in out.featuresToDrop) == sorted(drop) and "r1" in drop
    for g, e in zip(out.rawFeatureFilterResults.exclusionReasons, r):
        assert (g.name, g.trainingUnfilledState, g.trainingNullLabelLeaker, g.scoringUnfilledState,
                g.jsDivergenceMismatch, g.fillRateDiffMismatch, g.fillRatioDiffMismatch, g.excluded) == \
            (e.name, e.training_unfilled_state, e.training_null_label_leaker, e.scoring_unfilled_state,
             e.js_divergence_mismatch, e.fill_rate_diff_mismatch, e.fill_ratio_diff_mismatch, e.excluded)
    for g, e in zip(out.rawFeatureFilterResults.rawFeatureFilterMetrics, m):
        a, b = g.jsDivergence, e.js_divergence
        assert (np.isnan(a) and np.isnan(b)) or abs(a - b) < 1e-9, g.name
    kept = {nm for c in out.cleanedData for b in c.blocks for nm in b.names}
    assert kept == set(train) - set(drop)


def test_device_resident_batch_equals_host_chunks():
    """The same rows as one batch already in HBM and as pinned host chunks: identical distributions."""
    from transmogrifai_b200.columnar import ColumnarBatch
    from transmogrifai_b200.filters import RawFeatureFilter
    train = make_data(1500, 21)
    feats = features_of(train)
    f = RawFeatureFilter(bins=64)
    a = f.computeFeatureStats(ColumnarBatch.from_columns(train, response="y").to("cuda"), feats)
    b = f.computeFeatureStats(chunked(train, [500, 1000]), feats)
    for da, db in zip(a.predictorDistributions + a.responseDistributions,
                      b.predictorDistributions + b.responseDistributions):
        assert da.name == db.name and (da.count, da.nulls) == (db.count, db.nulls)
        assert np.array_equal(da.distribution, db.distribution)
        assert np.array_equal(da.summaryInfo, db.summaryInfo, equal_nan=True)


def test_numeric_histograms_at_scale_match_numpy():
    """2 M rows x 6 numeric columns (f64, int64, u8): the histogram kernel against a vectorised numpy restatement of
    NumericBucketizer's left-inclusive buckets over the same splits; every valid row lands in exactly one bucket."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch, NumericBlock
    from transmogrifai_b200.filters import RawFeatureFilter
    n, bins = 2_000_000, 100
    rng = np.random.default_rng(5)
    ld = (n + 15) // 16 * 16
    words = (n + 63) // 64

    def bitmap(p):
        ok = rng.random(words * 64) >= p
        ok[n:] = False
        return ok[:n], torch.from_numpy(np.packbits(ok, bitorder="little").view(np.int64).copy())

    real = np.zeros((3, ld))
    real[0, :n] = rng.normal(0, 1, n)
    real[1, :n] = rng.exponential(1e6, n)
    real[2, :n] = np.round(rng.normal(0, 3, n))                 # many ties on bucket borders
    oks, maps = zip(*[bitmap(p) for p in (0.1, 0.0, 0.5)])
    ints = np.zeros((2, ld), dtype=np.int64)
    ints[0, :n] = rng.integers(-1000, 1000, n)
    ints[1, :n] = rng.integers(0, 3, n) * (10 ** 12)
    ok_i, map_i = zip(*[bitmap(p) for p in (0.2, 0.01)])
    bools = np.zeros((1, ld), dtype=np.uint8)
    bools[0, :n] = rng.random(n) < 0.3
    blocks = [NumericBlock(["r0", "r1", "r2"], F.REAL, torch.from_numpy(real), torch.stack(maps), n),
              NumericBlock(["i0", "i1"], F.INTEGRAL, torch.from_numpy(ints), torch.stack(map_i), n),
              NumericBlock(["b0"], F.BINARY, torch.from_numpy(bools), None, n)]
    batch = ColumnarBatch(n, blocks, {}, None).to("cuda")
    feats = [F.Feature(nm, b.type_name, False) for b in blocks for nm in b.names]
    info = RawFeatureFilter(bins=bins).computeFeatureStats(batch, feats)
    cols = {"r0": (real[0, :n], oks[0]), "r1": (real[1, :n], oks[1]), "r2": (real[2, :n], oks[2]),
            "i0": (ints[0, :n].astype(np.float64), ok_i[0]), "i1": (ints[1, :n].astype(np.float64), ok_i[1]),
            "b0": (bools[0, :n].astype(np.float64), np.ones(n, dtype=bool))}
    assert [d.name for d in info.predictorDistributions] == list(cols)
    for d in info.predictorDistributions:
        v, ok = cols[d.name]
        v = v[ok]
        s = info.predictorSummaries[(d.name, None)]
        assert (s.min, s.max, s.count) == (v.min(), v.max(), len(v)) and abs(s.sum - v.sum()) <= 1e-9 * abs(v).sum()
        assert d.count == n and d.nulls == n - len(v)
        step = (s.max - s.min) / (bins - 2.0)
        splits = np.array([s.min + step * b for b in range(bins)])
        assert np.array_equal(d.summaryInfo, splits)
        b = np.searchsorted(splits, v, side="right") - 1
        b[(b < 0) | (b >= bins - 1)] = bins - 1
        assert np.array_equal(d.distribution, np.bincount(b, minlength=bins).astype(np.float64)), d.name
        assert d.distribution.sum() == len(v) and d.distribution[bins - 1] == 0
