"""GPU parity of the reference-facing stage API (setInput / fit / transform / getMetadata) against the oracle."""
import math

import numpy as np
import pytest

from oracle import sanity as S
from oracle import vectorizers as V
from oracle.metadata import ColumnMeta
from tests.util_data import make_real_config

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _batch(cols, response=None, device=None):
    from transmogrifai_b200.columnar import ColumnarBatch
    b = ColumnarBatch.from_columns(cols, response=response)
    return b.to(device) if device else b


def _lists(x, valid):
    return [[float(v) if ok else None for v, ok in zip(c, vc)] for c, vc in zip(x, valid)]


@pytest.mark.parametrize("device", [None, "cuda"])
def test_real_vectorizer_stage_reference_goldens(device):
    # core/src/test/scala/com/salesforce/op/stages/impl/feature/RealVectorizerTest.scala:47-53,118-214
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.numeric import RealVectorizer
    inA, inB, inC = (F.FeatureBuilder.Real(n).asPredictor() for n in ("inA", "inB", "inC"))
    batch = _batch({"inA": (F.REAL, [4.0, None, 2.0]), "inB": (F.REAL, [2.0, 2.0, None]),
                    "inC": (F.REAL, [None, None, None])}, device=device)
    with pytest.raises(KeyError):
        RealVectorizer().getOutput()                                         # :77-79
    vec = RealVectorizer().setInput(inA, inB, inC)
    out = vec.getOutput()
    assert out.type_name == F.OPVECTOR and not out.is_response and out.parents == (inA, inB, inC)   # :81-89
    m = vec.setFillWithMean().setTrackNulls(False).fit(batch)
    assert m.fillValues == [3.0, 2.0, 0.0]
    assert m.transform(batch).vectors[out.name].to_numpy().tolist() == [[4, 2, 0], [3, 2, 0], [2, 2, 0]]
    m = vec.setFillWithConstant(4.2).setTrackNulls(False).fit(batch)
    assert m.transform(batch).vectors[out.name].to_numpy().tolist() == [[4, 2, 4.2], [4.2, 2, 4.2], [2, 4.2, 4.2]]
    m = vec.setFillWithMean().setTrackNulls(True).fit(batch)
    assert m.transform(batch).vectors[out.name].to_numpy().tolist() == \
        [[4, 0, 2, 0, 0, 1], [3, 1, 2, 0, 0, 1], [2, 0, 2, 1, 0, 1]]
    meta = m.getMetadata()
    assert [(c.grouping, c.indicatorValue) for c in meta.columns] == \
        [(None, None), ("inA", "NullIndicatorValue"), (None, None), ("inB", "NullIndicatorValue"), (None, None),
         ("inC", "NullIndicatorValue")]
    assert meta.columnNames()[1] == "inA_inA_NullIndicatorValue_1"


def test_binary_and_realnn_stages():
    # BinaryVectorizerTest.scala:45-70, RealNNVectorizerTest.scala:45
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.numeric import BinaryVectorizer, RealNNVectorizer
    f1 = [False, False, True, True, None, None, False, True, None]
    f2 = [False, True, False, True, False, True, None, None, None]
    batch = _batch({"f1": (F.BINARY, f1), "f2": (F.BINARY, f2), "r": (F.REALNN, [float(i) for i in range(9)])})
    st = BinaryVectorizer().setInput(F.FeatureBuilder.Binary("f1").asPredictor(),
                                     F.FeatureBuilder.Binary("f2").asPredictor())
    got = st.transform(batch).vectors[st.getOutput().name].to_numpy()
    assert np.array_equal(got, V.binary_vectorizer_transform([f1, f2]))
    assert [c.indicatorValue for c in st.getMetadata().columns] == [None, "NullIndicatorValue"] * 2
    rn = RealNNVectorizer().setInput(F.FeatureBuilder.RealNN("r").asPredictor())
    assert rn.transform(batch).vectors[rn.getOutput().name].to_numpy()[:, 0].tolist() == [float(i) for i in range(9)]
    with pytest.raises(ValueError):
        _batch({"r": (F.REALNN, [1.0, None])})


SC_ROWS = [(32, 5.0, 0, 1, 0.5, 0), (32, 4.0, 1, 0, 0, 0.1), (32, 6.0, 1, 1, 0.5, 0), (32, 5.5, 0, 1, 0.5, 0),
           (32, 5.4, 1, 0, 0, 0.1), (32, 5.4, 1, 0, 0, 0.1)]
SC_NAMES = ["age", "height", "height_null", "gender", "testFeatNegCor"]


def _sc_batch():
    """SanityCheckerTest.scala:100-128: a hand-made OPVector column with metadata but no statistics sidecar."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch, DensePart, NumericBlock, VectorColumn
    from transmogrifai_b200.metadata import OpVectorColumnMetadata, OpVectorMetadata
    x = np.array([[r[0], r[1], r[2], r[4], r[5]] for r in SC_ROWS], dtype=np.float64)
    y = [float(r[3]) for r in SC_ROWS]
    blk = NumericBlock.from_lists(["isBlueEye"], F.REALNN, [y])
    meta = OpVectorMetadata("features", [OpVectorColumnMetadata((n,), (F.REAL,)) for n in SC_NAMES])
    vec = VectorColumn("features", 6, meta, [DensePart(torch.from_numpy(x).cuda())])
    batch = ColumnarBatch(6, [blk], {"features": vec}, response="isBlueEye")
    label = F.Feature("isBlueEye", F.REALNN, True)
    fv = F.Feature("features", F.OPVECTOR, False)
    return batch, label, fv, x, y


def test_sanity_checker_reference_drop_set():
    # SanityCheckerTest.scala:133-164
    batch, label, fv, x, y = _sc_batch()
    checked = label.sanityCheck(fv, maxCorrelation=0.99, minVariance=0.0, checkSample=1.0, removeBadFeatures=True,
                                sampleLowerLimit=0)
    est = checked.origin_stage
    model = est.fit(batch)
    assert est.summary.dropped == ["age_0", "gender_3", "testFeatNegCor_4"]
    assert model.indicesToKeep == [1, 2]
    corr = dict(zip(est.summary.names, est.summary.correlationsWithLabel))
    assert math.isnan(corr["age_0"]) and not any(math.isnan(corr[k]) for k in
                                                 ("height_1", "height_null_2", "gender_3", "testFeatNegCor_4"))
    assert est.summary.max[-1] == 1.0 and est.summary.mean[-1] == 0.5        # :659-660
    out = model.transform(batch).vectors[checked.name]
    assert np.array_equal(out.to_numpy(), x[:, [1, 2]])
    assert out.metadata.columnNames() == ["height_0", "height_null_1"]
    # oracle agreement on every statistic
    metas = [ColumnMeta((n,), ("Real",), index=i) for i, n in enumerate(SC_NAMES)]
    p = S.SanityParams(max_correlation=0.99, min_variance=0.0, remove_bad_features=True, sample_lower_limit=0)
    ref = S.sanity_fit(y, x, metas, p)
    np.testing.assert_allclose(est.summary.mean[:-1], ref.col_stats.mean[:-1], rtol=1e-12)
    np.testing.assert_allclose(est.summary.variance, ref.col_stats.variance, rtol=1e-9, atol=1e-18)
    np.testing.assert_allclose(est.summary.featureFeatureCorr, ref.corr_matrix, rtol=1e-9, atol=1e-12, equal_nan=True)


def test_sanity_checker_feature_feature_and_errors():
    # SanityCheckerTest.scala:473-500 (later duplicate dropped), :349-357 (everything dropped -> error)
    batch, label, fv, x, y = _sc_batch()
    est = label.sanityCheck(fv, maxCorrelation=1.1, minVariance=-0.1, removeBadFeatures=True,
                            removeFeatureGroup=False, sampleLowerLimit=0).origin_stage
    est.setMinCorrelation(-0.1).setMaxFeatureCorrelation(0.9).setMaxCramersV(1.0).setCorrelationExclusion("HashedText")
    est.fit(batch)
    assert est.summary.dropped == ["testFeatNegCor_4"]
    est.setFeatureFeatureCorrLevel("Off").fit(batch)
    assert est.summary.dropped == []
    bad = label.sanityCheck(fv, minVariance=1e9, removeBadFeatures=True, sampleLowerLimit=0).origin_stage
    with pytest.raises(ValueError, match="dropped all of your features"):
        bad.fit(batch)
    from transmogrifai_b200.stages.sanity import SanityChecker
    for v in (-1.0, 0.0, 2.0):                                              # :166-171
        with pytest.raises(ValueError):
            SanityChecker().setCheckSample(v)


@pytest.mark.parametrize("device", [None, "cuda"])
@pytest.mark.parametrize("n,ncols,corr_level", [(3000, 14, "Off"), (2500, 20, "Computed")])
def test_transmogrify_sanity_check_real_matches_oracle(device, n, ncols, corr_level):
    """Config-2 shape at oracle size: transmogrify(Real columns) + sanityCheck(removeBadFeatures) -- fused sidecar
    path -- against oracle.sanity_fit on the oracle's own vectorized matrix."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.dsl import fit_and_transform, transmogrify
    x, valid, y = make_real_config(n, ncols, seed=77)
    cols = _lists(x, valid)
    names = [f"c{j:03d}" for j in range(ncols)]
    data = {nm: (F.REAL, c) for nm, c in zip(names, cols)}
    data["y"] = (F.REALNN, y.tolist())
    batch = _batch(data, response="y", device=device)
    feats = [F.FeatureBuilder.Real(nm).asPredictor() for nm in names]
    label = F.FeatureBuilder.RealNN("y").asResponse()
    vec = transmogrify(feats)
    checked = label.sanityCheck(vec, removeBadFeatures=True, sampleLowerLimit=0, sampleUpperLimit=10 ** 9,
                                featureFeatureCorrLevel=corr_level)
    out_batch, fitted = fit_and_transform(batch, checked)
    est = checked.origin_stage
    # oracle
    fills = V.real_vectorizer_fit(cols)
    xo = V.real_vectorizer_transform(cols, fills, True)
    metas = V.numeric_metadata(names, "Real", True)
    p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                       feature_feature_corr_level=corr_level.lower())
    ref = S.sanity_fit(y, xo, metas, p, label_name="y")
    assert est.summary.dropped == ref.dropped                                  # removed-feature set: bit-exact
    assert fitted[-1].indicesToKeep == ref.indices_to_keep
    np.testing.assert_allclose(est.summary.mean, ref.col_stats.mean, rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(est.summary.variance, ref.col_stats.variance, rtol=1e-9, atol=1e-18)
    assert np.array_equal(est.summary.min, ref.col_stats.min) and np.array_equal(est.summary.max, ref.col_stats.max)
    np.testing.assert_allclose(est.summary.correlationsWithLabel, ref.corrs_with_label, rtol=1e-9, atol=1e-12,
                               equal_nan=True)                                 # north_star: corr within 1e-9
    got_v = {g.group: g.cramersV for g in est.summary.categoricalStats}
    ref_v = {g.group: g.cramers_v for g in ref.cat_stats}
    assert got_v.keys() == ref_v.keys()
    for k in ref_v:
        np.testing.assert_allclose(got_v[k], ref_v[k], rtol=1e-9, equal_nan=True)   # Cramer's V within 1e-9
    got = out_batch.vectors[checked.name].to_numpy()
    np.testing.assert_allclose(got, xo[:, ref.indices_to_keep], rtol=1e-12)
    assert out_batch.vectors[checked.name].metadata.columnNames() == \
        [f"{m.col_name().rsplit('_', 1)[0]}_{i}" for i, m in enumerate(metas[j] for j in ref.indices_to_keep)]


@pytest.mark.parametrize("device", [None, "cuda"])
def test_transmogrify_all_types_matches_oracle(device):
    """Config-5 shape in miniature: Binary + PickList + Real + RealNN + Text through transmogrify() + sanityCheck();
    type groups in type-name order; every statistic from the fused sidecars against the oracle."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.dsl import fit_and_transform, transmogrify
    rng = np.random.default_rng(31)
    n = 2000
    y = (rng.random(n) < 0.3).astype(np.float64)
    opt = lambda v, p: None if rng.random() < p else v   # noqa: E731
    b0 = [opt(bool(rng.random() < 0.4 + 0.3 * yy), 0.1) for yy in y]
    b1 = [opt(bool(rng.random() < 0.5), 0.05) for _ in y]
    p0 = [opt(["x", "y", "z", "X ", "w"][int(rng.integers(5))], 0.08) for _ in y]
    r0 = [opt(float(rng.normal(2 + yy, 1)), 0.1) for yy in y]
    r1 = [opt(float(rng.normal(0, 3)), 0.02) for _ in y]
    nn = [float(v) for v in rng.normal(5, 2, n)]
    words = ["".join(rng.choice(list("abcdefgh"), size=rng.integers(3, 7))) for _ in range(150)]
    t0 = [opt(" ".join(words[i] for i in rng.integers(0, 150, rng.poisson(8))), 0.06) for _ in y]
    data = {"b0": (F.BINARY, b0), "b1": (F.BINARY, b1), "p0": (F.PICKLIST, p0), "r0": (F.REAL, r0), "r1": (F.REAL, r1),
            "nn": (F.REALNN, nn), "t0": (F.TEXT, t0), "y": (F.REALNN, y.tolist())}
    batch = _batch(data, response="y", device=device)
    FB = F.FeatureBuilder
    feats = [FB.Text("t0").asPredictor(), FB.Real("r0").asPredictor(), FB.Binary("b0").asPredictor(),
             FB.RealNN("nn").asPredictor(), FB.PickList("p0").asPredictor(), FB.Real("r1").asPredictor(),
             FB.Binary("b1").asPredictor()]
    label = FB.RealNN("y").asResponse()
    checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True, sampleLowerLimit=0,
                                sampleUpperLimit=10 ** 9, featureFeatureCorrLevel="Off")
    out_batch, fitted = fit_and_transform(batch, checked)
    est = checked.origin_stage
    # oracle, Transmogrifier.scala:114 order: Binary < PickList < Real < RealNN < Text
    xb = V.binary_vectorizer_transform([b0, b1])
    tops = V.pivot_fit([p0], 20, 10, True)
    xp = V.pivot_transform([p0], tops, True, True)
    fills = V.real_vectorizer_fit([r0, r1])
    xr = V.real_vectorizer_transform([r0, r1], fills, True)
    xn = V.realnn_vectorizer_transform([nn])
    methods, ttops = V.smart_text_fit([t0], max_cardinality=30, top_k=20, min_support=10)
    xt = V.smart_text_transform([t0], ["t0"], methods, ttops, num_features=512)
    xo, metas = V.combine(
        [xb, xp, xr, xn, xt],
        [V.numeric_metadata(["b0", "b1"], "Binary", True), V.pivot_metadata(["p0"], "PickList", tops, True),
         V.numeric_metadata(["r0", "r1"], "Real", True), V.numeric_metadata(["nn"], "RealNN", False),
         V.smart_text_metadata(["t0"], "Text", methods, ttops, 512, True)])
    p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                       feature_feature_corr_level="off")
    ref = S.sanity_fit(y, xo, metas, p, label_name="y")
    assert est.summary.names[:-1] == [m.col_name() for m in metas]
    assert est.summary.dropped == ref.dropped
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
    got = out_batch.vectors[checked.name].to_scipy()
    exp = xo.tocsc()[:, ref.indices_to_keep].tocsr()
    assert got.shape == exp.shape and abs(got - exp).max() < 1e-9


def test_transmogrify_all_types_continuous_label_matches_oracle():
    """A regression label (every value distinct): no categorical tests, and the label correlations of the one-hot,
    multiplicity, null-indicator and hashed columns come from per-column label sums taken in the fused scans."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.dsl import fit_and_transform, transmogrify
    rng = np.random.default_rng(41)
    n = 2000
    z = rng.normal(0, 1, n)
    y = 100.0 + 3.0 * z + rng.normal(0, 0.5, n)
    opt = lambda v, p: None if rng.random() < p else v   # noqa: E731
    b0 = [opt(bool(zz + rng.normal(0, 1) > 0), 0.1) for zz in z]
    p0 = [opt(["x", "y", "z", "X ", "w"][int(np.clip(round(zz + 2), 0, 4))], 0.08) for zz in z]
    r0 = [opt(float(zz * 2 + rng.normal(0, 1)), 0.1) for zz in z]
    nn = [float(v) for v in rng.normal(5, 2, n)]
    words = ["".join(rng.choice(list("abcdefgh"), size=rng.integers(3, 7))) for _ in range(150)]
    t0 = [opt(" ".join([words[i] for i in rng.integers(0, 150, rng.poisson(8))] + (["zq zq"] if zz > 0.5 else [])), 0.06)
          for zz in z]
    t1 = [opt(["red", "Green", "blue", "blue!"][int(rng.integers(4))], 0.05) for _ in z]
    s0 = [opt([["a", "A "], ["b"], ["a", "c", "d"], []][int(rng.integers(4))] + (["hi"] if zz > 0 else []), 0.07)
          for zz in z]
    data = {"b0": (F.BINARY, b0), "p0": (F.PICKLIST, p0), "r0": (F.REAL, r0), "nn": (F.REALNN, nn), "t0": (F.TEXT, t0),
            "t1": (F.TEXT, t1), "s0": (F.MULTIPICKLIST, s0), "y": (F.REALNN, y.tolist())}
    batch = _batch(data, response="y", device="cuda")
    FB = F.FeatureBuilder
    feats = [FB.Text("t0").asPredictor(), FB.Real("r0").asPredictor(), FB.Binary("b0").asPredictor(),
             FB.RealNN("nn").asPredictor(), FB.PickList("p0").asPredictor(), FB.Text("t1").asPredictor(),
             FB.MultiPickList("s0").asPredictor()]
    label = FB.RealNN("y").asResponse()
    checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True, sampleLowerLimit=0,
                                sampleUpperLimit=10 ** 9, featureFeatureCorrLevel="Off")
    out_batch, fitted = fit_and_transform(batch, checked)
    est = checked.origin_stage
    # oracle; Transmogrifier.scala:114 order: Binary < MultiPickList < PickList < Real < RealNN < Text
    xb = V.binary_vectorizer_transform([b0])
    stops = V.pivot_fit([s0], 20, 10, True)
    xs = V.pivot_transform([s0], stops, True, True)
    tops = V.pivot_fit([p0], 20, 10, True)
    xp = V.pivot_transform([p0], tops, True, True)
    fills = V.real_vectorizer_fit([r0])
    xr = V.real_vectorizer_transform([r0], fills, True)
    xn = V.realnn_vectorizer_transform([nn])
    methods, ttops = V.smart_text_fit([t0, t1], max_cardinality=30, top_k=20, min_support=10)
    assert methods == ["Hash", "Pivot"]
    xt = V.smart_text_transform([t0, t1], ["t0", "t1"], methods, ttops, num_features=512)
    xo, metas = V.combine(
        [xb, xs, xp, xr, xn, xt],
        [V.numeric_metadata(["b0"], "Binary", True), V.pivot_metadata(["s0"], "MultiPickList", stops, True),
         V.pivot_metadata(["p0"], "PickList", tops, True), V.numeric_metadata(["r0"], "Real", True),
         V.numeric_metadata(["nn"], "RealNN", False), V.smart_text_metadata(["t0", "t1"], "Text", methods, ttops, 512, True)])
    p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                       feature_feature_corr_level="off")
    ref = S.sanity_fit(y, xo, metas, p, label_name="y")
    assert not ref.cat_stats and not list(est.summary.categoricalStats)
    assert est.summary.names[:-1] == [m.col_name() for m in metas]
    assert est.summary.dropped == ref.dropped
    assert fitted[-1].indicesToKeep == ref.indices_to_keep
    np.testing.assert_allclose(est.summary.mean, ref.col_stats.mean, rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(est.summary.variance, ref.col_stats.variance, rtol=1e-9, atol=1e-18)
    assert np.array_equal(est.summary.min, ref.col_stats.min) and np.array_equal(est.summary.max, ref.col_stats.max)
    np.testing.assert_allclose(est.summary.correlationsWithLabel, ref.corrs_with_label, rtol=1e-9, atol=1e-12,
                               equal_nan=True)
    got = out_batch.vectors[checked.name].to_scipy()
    import scipy.sparse as sp
    exp = sp.csr_matrix(xo)[:, ref.indices_to_keep]
    assert got.shape == exp.shape and abs(got - exp).max() < 1e-12


def test_min_variance_filter_reference_drop_set():
    # MinVarianceFilterTest.scala:73-124: minVariance 0.1 drops age, gender and testFeatNegCor, keeps the two heights
    from transmogrifai_b200.stages.min_variance import MinVarianceFilter
    batch, _label, fv, x, _y = _sc_batch()
    checked = fv.filterMinVariance(minVariance=0.1, removeBadFeatures=True)
    est = checked.origin_stage
    assert isinstance(est, MinVarianceFilter)
    model = est.fit(batch)
    assert est.summary.dropped == ["age_0", "gender_3", "testFeatNegCor_4"]
    assert model.indicesToKeep == [1, 2]
    out = model.transform(batch).vectors[checked.name]
    assert np.array_equal(out.to_numpy(), x[:, [1, 2]])
    assert out.metadata.columnNames() == ["height_0", "height_null_1"]
    st = S.col_stats(x)
    assert est.summary.count == 6
    np.testing.assert_allclose(est.summary.mean, st.mean, rtol=1e-12)
    np.testing.assert_allclose(est.summary.variance, st.variance, rtol=1e-9, atol=1e-18)
    assert np.array_equal(est.summary.min, st.min) and np.array_equal(est.summary.max, st.max)
    # defaults (MinVarianceFilter.scala:155-158): nothing is removed, the vector passes through
    dflt = MinVarianceFilter().setInput(fv)
    assert dflt.minVariance == 1e-5 and dflt.removeBadFeatures is False
    m2 = dflt.fit(batch)
    assert m2.indicesToKeep == [0, 1, 2, 3, 4] and dflt.summary.dropped == []
    assert np.array_equal(m2.transform(batch).vectors[dflt.getOutput().name].to_numpy(), x)
    with pytest.raises(ValueError, match="dropped all of your features"):
        MinVarianceFilter().setInput(fv).setMinVariance(10.0).setRemoveBadFeatures(True).fit(batch)


@pytest.mark.parametrize("with_label", [True, False])
def test_min_variance_filter_after_transmogrify(with_label):
    """On the output of transmogrify(): with a response in the batch the statistics come from the fused vectorizer
    scans, without one from the dense statistics kernel; both must agree with the oracle's colStats and variance rule."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.dsl import fit_and_transform, transmogrify
    rng = np.random.default_rng(11)
    n = 5000
    cols = {}
    for j in range(6):
        v = rng.normal(j, 1.0 + j, n)
        if j == 2:
            v[:] = 3.25                                     # constant: variance 0
        if j == 4:
            v = rng.normal(0.0, 1e-3, n)                    # variance 1e-6 <= 1e-5
        cols[f"r{j}"] = (F.REAL, [None if rng.random() < 0.1 else float(t) for t in v])
    cols["y"] = (F.REALNN, [float(rng.random() < 0.4) for _ in range(n)])
    batch = _batch(cols, response="y" if with_label else None).to("cuda")
    feats = [F.FeatureBuilder.Real(f"r{j}").asPredictor() for j in range(6)]
    checked = transmogrify(feats).filterMinVariance(removeBadFeatures=True)
    out, fitted = fit_and_transform(batch, checked)
    est = checked.origin_stage
    raw = [cols[f"r{j}"][1] for j in range(6)]
    fills = V.real_vectorizer_fit(raw)
    exp = V.real_vectorizer_transform(raw, fills, True)
    st = S.col_stats(exp)
    np.testing.assert_allclose(est.summary.mean, st.mean, rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(est.summary.variance, st.variance, rtol=1e-9, atol=1e-18)
    keep = [i for i in range(exp.shape[1]) if not st.variance[i] <= 1e-5]
    assert fitted[-1].indicesToKeep == keep and 4 not in keep and 8 not in keep
    got = out.vectors[checked.name].to_numpy()
    np.testing.assert_allclose(got, exp[:, keep], rtol=1e-12, atol=0)


def test_sanity_checker_sampling_matches_oracle_on_the_same_rows():
    """fraction() < 1 (SanityChecker.scala:356-387): the statistics, tests and the drop decision are those of the
    sampled rows of the vector that the models fitted on ALL rows produce.  The sample itself is ours (seeded device
    generator), so the oracle is given the same rows."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.dsl import fit_and_transform, transmogrify
    rng = np.random.default_rng(51)
    n = 6000
    y = (rng.random(n) < 0.3).astype(np.float64)
    opt = lambda v, p: None if rng.random() < p else v   # noqa: E731
    b0 = [opt(bool(rng.random() < 0.4 + 0.3 * yy), 0.1) for yy in y]
    p0 = [opt(["x", "y", "z", "X ", "w"][int(rng.integers(5))], 0.08) for _ in y]
    r0 = [opt(float(rng.normal(2 + yy, 1)), 0.1) for yy in y]
    i0 = [opt(int(rng.integers(0, 6)), 0.1) for _ in y]
    words = ["".join(rng.choice(list("abcdefgh"), size=rng.integers(3, 7))) for _ in range(150)]
    t0 = [opt(" ".join(words[i] for i in rng.integers(0, 150, rng.poisson(8))), 0.06) for _ in y]
    data = {"b0": (F.BINARY, b0), "p0": (F.PICKLIST, p0), "r0": (F.REAL, r0), "i0": (F.INTEGRAL, i0),
            "t0": (F.TEXT, t0), "y": (F.REALNN, y.tolist())}
    batch = _batch(data, response="y", device="cuda")
    FB = F.FeatureBuilder
    feats = [FB.Text("t0").asPredictor(), FB.Real("r0").asPredictor(), FB.Binary("b0").asPredictor(),
             FB.PickList("p0").asPredictor(), FB.Integral("i0").asPredictor()]
    label = FB.RealNN("y").asResponse()
    # reference defaults for the limits, scaled down: upper limit 2000 rows of 6000 -> fraction 1/3
    checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True, checkSample=1.0, sampleLowerLimit=100,
                                sampleUpperLimit=2000, sampleSeed=7, featureFeatureCorrLevel="Computed")
    est = checked.origin_stage
    frac = est.fraction(n)
    assert abs(frac - 2000 / 6000) < 1e-12
    out_batch, fitted = fit_and_transform(batch, checked)
    g = torch.Generator(device="cuda").manual_seed(7)
    idx = (torch.rand(n, device="cuda", generator=g) < frac).nonzero().squeeze(1).cpu().numpy()
    assert 1700 < len(idx) < 2300 and est.summary.count == len(idx) and est.summary.sampleFraction == frac
    # the vector over all rows, from models fitted on all rows
    xb = V.binary_vectorizer_transform([b0])
    modes = V.integral_vectorizer_fit([i0], fill_with_mode=True)
    xi = V.integral_vectorizer_transform([i0], modes, True)
    tops = V.pivot_fit([p0], 20, 10, True)
    xp = V.pivot_transform([p0], tops, True, True)
    fills = V.real_vectorizer_fit([r0])
    xr = V.real_vectorizer_transform([r0], fills, True)
    methods, ttops = V.smart_text_fit([t0], max_cardinality=30, top_k=20, min_support=10)
    xt = V.smart_text_transform([t0], ["t0"], methods, ttops, num_features=512)
    xo, metas = V.combine(
        [xb, xi, xp, xr, xt],
        [V.numeric_metadata(["b0"], "Binary", True), V.numeric_metadata(["i0"], "Integral", True),
         V.pivot_metadata(["p0"], "PickList", tops, True), V.numeric_metadata(["r0"], "Real", True),
         V.smart_text_metadata(["t0"], "Text", methods, ttops, 512, True)])
    p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                       feature_feature_corr_level="computed")
    ref = S.sanity_fit(y[idx], xo.tocsr()[idx], metas, p, label_name="y")
    assert est.summary.dropped == ref.dropped
    assert fitted[-1].indicesToKeep == ref.indices_to_keep
    np.testing.assert_allclose(est.summary.mean, ref.col_stats.mean, rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(est.summary.variance, ref.col_stats.variance, rtol=1e-9, atol=1e-18)
    np.testing.assert_allclose(est.summary.correlationsWithLabel, ref.corrs_with_label, rtol=1e-9, atol=1e-12,
                               equal_nan=True)
    # the model still transforms every row
    got = out_batch.vectors[checked.name].to_scipy()
    exp = xo.tocsc()[:, ref.indices_to_keep].tocsr()
    assert got.shape == exp.shape and abs(got - exp).max() < 1e-9


def test_sanity_checker_default_limits_sample_a_large_batch():
    """The reference's defaults (checkSample 1.0, sampleUpperLimit 1e6) on 1.2 M rows: five sixths of the rows are
    sampled, nothing raises, and the sample statistics are close to those of all rows."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch, NumericBlock
    from transmogrifai_b200.dsl import fit_and_transform, transmogrify
    n, ncols = 1_200_000, 3
    g = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn((ncols, n), dtype=torch.float64, device="cuda", generator=g) * 2.0 + 5.0
    y = (torch.rand(n, device="cuda", generator=g) < 0.3).double()
    x[1] += y                                                    # a correlated column
    names = [f"c{j}" for j in range(ncols)]
    blocks = [NumericBlock(names, F.REAL, x, None, n), NumericBlock(["y"], F.REALNN, y.reshape(1, n), None, n)]
    batch = ColumnarBatch(n, blocks, {}, "y")
    feats = [F.FeatureBuilder.Real(nm).asPredictor() for nm in names]
    label = F.FeatureBuilder.RealNN("y").asResponse()
    checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True)        # every default
    out, fitted = fit_and_transform(batch, checked)
    est = checked.origin_stage
    assert abs(est.summary.sampleFraction - 1e6 / n) < 1e-12
    assert abs(est.summary.count - 1_000_000) < 3000
    np.testing.assert_allclose(est.summary.mean[0:6:2], x.mean(dim=1).cpu().numpy(), rtol=5e-3)
    corr = dict(zip(est.summary.names, est.summary.correlationsWithLabel))
    assert corr["c1_2"] > 0.15 and abs(corr["c0_0"]) < 0.01
    assert out.vectors[checked.name].n == n


def test_host_batch_raw_blocks_cross_the_link_once(monkeypatch):
    """A host batch through fit + transform: every PickList / Text / MultiPickList block is copied to the GPU once (the
    estimator's fit and its model's transform share the copy), and the result equals the device batch's."""
    from transmogrifai_b200 import columnar as CB
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.dsl import fit_and_transform, transmogrify
    rng = np.random.default_rng(5)
    n = 1500
    y = (rng.random(n) < 0.4).astype(np.float64)
    opt = lambda v, p: None if rng.random() < p else v   # noqa: E731
    p0 = [opt("abcde"[int(rng.integers(5))], 0.1) for _ in y]
    p1 = [opt("uvw"[int(rng.integers(3))], 0.05) for _ in y]
    words = ["".join(rng.choice(list("abcdefgh"), size=rng.integers(3, 7))) for _ in range(120)]
    t0 = [opt(" ".join(words[i] for i in rng.integers(0, 120, rng.poisson(6))), 0.05) for _ in y]
    s0 = [set(rng.choice(list("klmn"), size=rng.integers(0, 3), replace=False).tolist()) for _ in y]
    data = {"p0": (F.PICKLIST, p0), "p1": (F.PICKLIST, p1), "t0": (F.TEXT, t0), "s0": (F.MULTIPICKLIST, s0),
            "y": (F.REALNN, y.tolist())}
    FB = F.FeatureBuilder
    feats = [FB.PickList("p0").asPredictor(), FB.PickList("p1").asPredictor(), FB.Text("t0").asPredictor(),
             FB.MultiPickList("s0").asPredictor()]
    label = FB.RealNN("y").asResponse()

    def run(batch):
        checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True, sampleLowerLimit=0,
                                    sampleUpperLimit=10 ** 9, featureFeatureCorrLevel="Off")
        out, fitted = fit_and_transform(batch, checked)
        return out.vectors[checked.name].to_scipy(), fitted[-1].indicesToKeep

    copies = []
    for cls in (CB.DictBlock, CB.TextBlock, CB.SetBlock):
        orig = cls.to
        monkeypatch.setattr(cls, "to", lambda self, device, _o=orig, _c=cls: (copies.append(_c.__name__),
                                                                             _o(self, device))[1])
    host = _batch(data, response="y")
    got_h, keep_h = run(host)
    assert sorted(copies) == ["DictBlock", "SetBlock", "TextBlock"]
    got_d, keep_d = run(_batch(data, response="y", device="cuda"))
    assert keep_h == keep_d and got_h.shape == got_d.shape and abs(got_h - got_d).max() == 0


def test_vector_udt_export_matches_spark_compressed():
    """OPVector -> VectorUDT arrays (wire.py): every row carries exactly its non-zeros, in ascending index order, as a
    SparseVector when 1.5 (nnz + 1) < size and as a DenseVector otherwise (Spark's Vector.compressed, which the
    reference applies at the end of every stage: Transmogrifier.scala:466-475, RichVector.scala:120-134)."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.dsl import fit_and_transform, transmogrify
    from transmogrifai_b200.wire import to_vector_udt
    rng = np.random.default_rng(5)
    n = 1500
    y = (rng.random(n) < 0.3).astype(np.float64)
    opt = lambda v, p: None if rng.random() < p else v   # noqa: E731
    words = ["".join(rng.choice(list("abcdefgh"), size=rng.integers(3, 7))) for _ in range(60)]
    data = {"b0": (F.BINARY, [opt(bool(rng.random() < 0.5), 0.1) for _ in y]),
            "p0": (F.PICKLIST, [opt(["x", "y", "z"][int(rng.integers(3))], 0.1) for _ in y]),
            "r0": (F.REAL, [opt(float(rng.normal(0, 1)) if rng.random() < 0.7 else 0.0, 0.1) for _ in y]),
            "t0": (F.TEXT, [opt(" ".join(words[i] for i in rng.integers(0, 60, rng.poisson(6))), 0.1) for _ in y]),
            "y": (F.REALNN, y.tolist())}
    batch = _batch(data, response="y", device="cuda")
    FB = F.FeatureBuilder
    feats = [FB.Text("t0").asPredictor(), FB.Real("r0").asPredictor(), FB.Binary("b0").asPredictor(),
             FB.PickList("p0").asPredictor()]
    narrow = transmogrify(feats[1:3])                     # 4 columns: rows with two or more non-zeros come out dense
    wide = FB.RealNN("y").asResponse().sanityCheck(transmogrify(feats), removeBadFeatures=True, sampleLowerLimit=0,
                                                   featureFeatureCorrLevel="Off")
    for result in (narrow, wide):
        out, _ = fit_and_transform(batch, result)
        vec = out.vectors[result.name]
        exp = np.asarray(vec.to_scipy().todense())
        udt = to_vector_udt(vec)
        assert udt.width == exp.shape[1]
        types = udt.type.cpu().numpy()
        nnz = (exp != 0).sum(axis=1)
        assert np.array_equal(types == 0, 1.5 * (nnz + 1.0) < exp.shape[1])
        if result is narrow:
            assert (types == 1).any() and (types == 0).any()
        for r in list(range(0, n, 37)) + [n - 1]:
            tp, size, idx, val = udt.row(r)
            if tp == 0:
                assert size == exp.shape[1] and np.array_equal(idx, np.nonzero(exp[r])[0])
                assert np.array_equal(val, exp[r][idx])
            else:
                assert size == -1 and len(idx) == 0 and np.array_equal(val, exp[r])
        assert int(udt.idx_offsets[-1]) == int(nnz[types == 0].sum())
        assert int(udt.val_offsets[-1]) == int(nnz[types == 0].sum() + (types == 1).sum() * exp.shape[1])
