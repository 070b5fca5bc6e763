"""GPU parity of IntegralVectorizer (mode / constant fill, null indicators, statistics sidecar) against the oracle and
the reference's golden vectors."""
import ctypes as C

import numpy as np
import pytest

from oracle import sanity as S
from oracle import vectorizers as V

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

A, B, CC, DD = [4, 4, 2, None, None], [2, None, 4, 2, None], [2, 1, 1, 2, None], [None] * 5
NAMES = ["inA", "inB", "inC", "inD"]


def _batch(cols, response=None, device="cuda"):
    from transmogrifai_b200.columnar import ColumnarBatch
    b = ColumnarBatch.from_columns(cols, response=response)
    return b.to(device) if device else b


def _feats(names):
    from transmogrifai_b200 import features as F
    return [F.FeatureBuilder.Integral(nm).asPredictor() for nm in names]


@pytest.mark.parametrize("device", [None, "cuda"])
def test_integral_vectorizer_reference_golden(device):
    # IntegralVectorizerTest.scala:48-56, :93-118, :120-149, :159-185
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.integral import IntegralVectorizer
    batch = _batch({n: (F.INTEGRAL, c) for n, c in zip(NAMES, [A, B, CC, DD])}, device=device)
    est = IntegralVectorizer().setInput(*_feats(NAMES))
    with pytest.raises(KeyError):
        IntegralVectorizer().getOutput()
    m = est.setFillWithConstant(3).setTrackNulls(False).fit(batch)
    got = m.transform(batch).vectors[est.getOutput().name].to_numpy()
    assert got.tolist() == [[4.0, 2.0, 2.0, 3.0], [4.0, 3.0, 1.0, 3.0], [2.0, 4.0, 1.0, 3.0], [3.0, 2.0, 2.0, 3.0],
                            [3.0, 3.0, 3.0, 3.0]]
    m = est.setFillWithMode().setTrackNulls(False).fit(batch)
    assert m.fillValues == [4, 2, 1, 0]
    got = m.transform(batch).vectors[est.getOutput().name].to_numpy()
    assert got.tolist() == [[4.0, 2.0, 2.0, 0.0], [4.0, 2.0, 1.0, 0.0], [2.0, 4.0, 1.0, 0.0], [4.0, 2.0, 2.0, 0.0],
                            [4.0, 2.0, 1.0, 0.0]]
    m = est.setFillWithConstant(0).setTrackNulls(True).fit(batch)
    got = m.transform(batch).vectors[est.getOutput().name].to_numpy()
    assert got.tolist() == [[4.0, 0.0, 2.0, 0.0, 2.0, 0.0, 0.0, 1.0], [4.0, 0.0, 0.0, 1.0, 1.0, 0.0, 0.0, 1.0],
                            [2.0, 0.0, 4.0, 0.0, 1.0, 0.0, 0.0, 1.0], [0.0, 1.0, 2.0, 0.0, 2.0, 0.0, 0.0, 1.0],
                            [0.0, 1.0, 0.0, 1.0, 0.0, 1.0, 0.0, 1.0]]
    cols = est.getMetadata().columns
    assert [c.indicatorValue for c in cols] == [None, "NullIndicatorValue"] * 4
    assert [c.parentFeatureType for c in cols][0] == ("Integral",)


@pytest.mark.parametrize("type_name", ["Date", "DateTime"])
def test_integral_vectorizer_on_date_types(type_name):
    # IntegralVectorizerTest.scala:57-74 (data), :240-305 (DateTime and Date columns vectorize like Integral ones)
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.integral import IntegralVectorizer
    batch = _batch({n: (type_name, c) for n, c in zip(NAMES, [A, B, CC, DD])})
    feats = [F.FeatureBuilder.of(type_name, nm).asPredictor() for nm in NAMES]
    est = IntegralVectorizer().setInput(*feats)
    m = est.setFillWithMode().setTrackNulls(False).fit(batch)
    assert m.fillValues == [4, 2, 1, 0]
    got = m.transform(batch).vectors[est.getOutput().name].to_numpy()
    assert got.tolist() == [[4.0, 2.0, 2.0, 0.0], [4.0, 2.0, 1.0, 0.0], [2.0, 4.0, 1.0, 0.0], [4.0, 2.0, 2.0, 0.0],
                            [4.0, 2.0, 1.0, 0.0]]
    assert est.getMetadata().columns[0].parentFeatureType == (type_name,)


def _random_columns(rng, n):
    cols = []
    cols.append([None if rng.random() < 0.1 else int(rng.integers(0, 7)) for _ in range(n)])           # few values
    cols.append([None if rng.random() < 0.3 else int(rng.integers(-50, 50)) * 10 ** 12 for _ in range(n)])  # > 2^32
    z = rng.zipf(1.3, n)
    cols.append([int(v) - 5 for v in z])                                                                 # skewed
    cols.append([int(i) for i in rng.permutation(n)])                     # all distinct: mode = the smallest value
    cols.append([None] * n)                                                                              # empty
    cols.append([-(1 << 63) if i % 3 else int(rng.integers(0, 3)) for i in range(n)])                    # INT64_MIN
    return cols


@pytest.mark.parametrize("n", [1, 257, 20_000, 120_000])
def test_integral_modes_and_transform_match_oracle(n):
    """Modes bit-exact (ties to the smaller value, INT64_MIN, an empty column, a column whose distinct values
    outgrow the first table size), transform values exact."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.integral import IntegralVectorizer
    rng = np.random.default_rng(n)
    cols = _random_columns(rng, n)
    names = [f"i{j}" for j in range(len(cols))]
    batch = _batch({nm: (F.INTEGRAL, c) for nm, c in zip(names, cols)})
    est = IntegralVectorizer().setInput(*_feats(names)).setFillWithMode()
    model = est.fit(batch)
    assert model.fillValues == V.mode_seq_null_int(cols)
    got = model.transform(batch).vectors[est.getOutput().name].to_numpy()
    exp = V.integral_vectorizer_transform(cols, model.fillValues, True)
    assert np.array_equal(got, exp)


def test_integral_in_transmogrify_with_sanity_checker():
    """transmogrify() routes Integral features to IntegralVectorizer (mode fill, null tracking); the statistics
    sidecar of the scan must give SanityChecker what the oracle computes from the materialised vector."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.dsl import fit_and_transform, transmogrify
    rng = np.random.default_rng(3)
    n = 6000
    y = (rng.random(n) < 0.35).astype(np.float64)
    i0 = [None if rng.random() < 0.15 else int(rng.integers(0, 9) + 3 * yy) for yy in y]
    i1 = [None if rng.random() < 0.05 else int(rng.integers(-1000, 1000)) for _ in y]
    i2 = [7] * n                                                                      # constant: dropped
    r0 = [None if rng.random() < 0.1 else float(rng.normal(yy, 1.0)) for yy in y]
    batch = _batch({"i0": (F.INTEGRAL, i0), "i1": (F.INTEGRAL, i1), "i2": (F.INTEGRAL, i2), "r0": (F.REAL, r0),
                    "y": (F.REALNN, y.tolist())}, response="y")
    FB = F.FeatureBuilder
    feats = [FB.Integral("i0").asPredictor(), FB.Real("r0").asPredictor(), FB.Integral("i1").asPredictor(),
             FB.Integral("i2").asPredictor()]
    label = FB.RealNN("y").asResponse()
    checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True, sampleLowerLimit=0,
                                sampleUpperLimit=10 ** 9, featureFeatureCorrLevel="Off")
    out, fitted = fit_and_transform(batch, checked)
    est = checked.origin_stage
    icols = [i0, i1, i2]
    modes = V.integral_vectorizer_fit(icols, fill_with_mode=True)
    xi = V.integral_vectorizer_transform(icols, modes, True)
    fills = V.real_vectorizer_fit([r0])
    xr = V.real_vectorizer_transform([r0], fills, True)
    xo, metas = V.combine([xi, xr], [V.numeric_metadata(["i0", "i1", "i2"], "Integral", True),
                                     V.numeric_metadata(["r0"], "Real", True)])
    p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                       feature_feature_corr_level="off")
    ref = S.sanity_fit(y, xo, metas, p, label_name="y")
    assert est.summary.names[:-1] == [m.col_name() for m in metas]
    assert est.summary.dropped == ref.dropped and "i2_4" in ref.dropped
    assert fitted[-1].indicesToKeep == ref.indices_to_keep
    np.testing.assert_allclose(est.summary.mean, ref.col_stats.mean, rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(est.summary.variance, ref.col_stats.variance, rtol=1e-9, atol=1e-18)
    assert np.array_equal(est.summary.min, ref.col_stats.min) and np.array_equal(est.summary.max, ref.col_stats.max)
    np.testing.assert_allclose(est.summary.correlationsWithLabel, ref.corrs_with_label, rtol=1e-9, atol=1e-12,
                               equal_nan=True)
    got_g = {g.group: g for g in est.summary.categoricalStats}
    assert got_g.keys() == {g.group for g in ref.cat_stats}
    for g in ref.cat_stats:
        assert np.array_equal(got_g[g.group].contingencyMatrix, g.contingency_matrix)
        np.testing.assert_allclose(got_g[g.group].cramersV, g.cramers_v, rtol=1e-9, equal_nan=True)
    got = out.vectors[checked.name].to_numpy()
    exp = xo[:, ref.indices_to_keep]
    np.testing.assert_allclose(got, np.asarray(exp.todense()) if hasattr(exp, "todense") else exp, rtol=1e-12)


def test_integral_mode_table_overflow_is_reported():
    """The C ABI call with a table too small for the column's distinct values sets the overflow flag (the stage then
    calls again with a larger table)."""
    from transmogrifai_b200 import _lib as L
    n, cap = 4096, 64
    x = torch.arange(n, dtype=torch.int64, device="cuda").reshape(1, n)
    nb = C.c_size_t()
    L.call("tmog_integral_mode_scratch_bytes", 1, cap, C.byref(nb))
    scratch = torch.empty(nb.value, dtype=torch.uint8, device="cuda")
    modes = torch.zeros(1, dtype=torch.int64, device="cuda")
    counts = torch.zeros(1, dtype=torch.int64, device="cuda")
    L.call("tmog_integral_mode_dev", x.data_ptr(), n, None, 0, n, 1, cap, scratch.data_ptr(), nb.value,
           modes.data_ptr(), counts.data_ptr(), torch.cuda.current_stream().cuda_stream)
    offs = (C.c_size_t * 5)()
    L.call("tmog_integral_mode_layout", 1, cap, offs)
    assert int(scratch[offs[4]:offs[4] + 4].view(torch.int32)[0].item()) == 1
