"""GPU parity of the pivot vectorizer (OpTextPivotVectorizer on PickList / Text) and of SanityChecker's categorical
statistics on pivoted columns, against the oracle and the reference's golden vectors."""
import numpy as np
import pytest
import scipy.sparse as sp

from oracle import sanity as S
from oracle import vectorizers as V

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

TEXT1 = ["hello world", "hello world", "good evening", "hello world", None]
TEXT2 = ["Hello world!", "What's up", "How are you doing, my friend?", "Not bad, my friend.", None]


def _batch(cols, response=None, device=None):
    from transmogrifai_b200.columnar import ColumnarBatch
    b = ColumnarBatch.from_columns(cols, response=response)
    return b.to(device) if device else b


def test_host_string_functions_match_reference_goldens():
    # utils/src/test/scala/com/salesforce/op/utils/text/TextUtilsTest.scala:56-58 and the tie-break of
    # core/src/test/.../OpSetVectorizerTest.scala:138-162 (uppercase before lowercase)
    from transmogrifai_b200 import _lib as L
    from transmogrifai_b200.stages.pivot import clean_and_group, top_k
    ids, distinct = clean_and_group(["A string wit#h %bad pun&ctuation mark<=>s", "a", "A.", "C ", "b", "c", ""], True)
    assert distinct == ["AStringWitHBadPunCtuationMarkS", "A", "C", "B", ""]
    assert ids.tolist() == [0, 1, 1, 2, 3, 2, 4]
    assert top_k(["a", "A.", "C ", "b", "c"], np.array([2, 1, 1, 1, 1]), 10, 0) == [0, 1, 2, 3, 4]
    assert top_k(["x", "y", "Z", "z"], np.array([2, 2, 1, 1]), 10, 0) == [0, 1, 2, 3]
    assert top_k(["b", "a", "c"], np.array([5, 5, 9]), 2, 6) == [2]
    with pytest.raises(L.TmogInvalidArgument, match="non-ASCII"):
        clean_and_group(["café"], True)
    ids, distinct = clean_and_group(["café", "x"], False)
    assert distinct == ["café", "x"]
    # UTF-16 order: a supplementary character (surrogates D83D..) sorts before U+FF5E
    assert top_k(["～", "\U0001F600"], np.array([1, 1]), 2, 0) == [1, 0]


@pytest.mark.parametrize("device", [None, "cuda"])
def test_text_pivot_vectorizer_reference_golden(device):
    # core/src/test/scala/com/salesforce/op/stages/impl/feature/OpTextPivotVectorizerTest.scala:45-73
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.pivot import OpTextPivotVectorizer
    batch = _batch({"text1": (F.TEXT, TEXT1), "text2": (F.TEXT, TEXT2)}, device=device)
    f1, f2 = F.FeatureBuilder.Text("text1").asPredictor(), F.FeatureBuilder.Text("text2").asPredictor()
    est = OpTextPivotVectorizer().setMinSupport(1).setTopK(2).setInput(f1, f2)
    model = est.fit(batch)
    got = model.transform(batch).vectors[est.getOutput().name].to_numpy()
    exp = np.zeros((5, 8))
    for r, idx in enumerate([(0, 4), (0, 6), (1, 5), (0, 6), (3, 7)]):
        exp[r, list(idx)] = 1.0
    assert np.array_equal(got, exp)
    meta = model.getMetadata()
    assert [c.indicatorValue for c in meta.columns[:4]] == ["HelloWorld", "GoodEvening", "OTHER", "NullIndicatorValue"]
    assert all(c.grouping == "text1" for c in meta.columns[:4])
    with pytest.raises(ValueError):
        est.setTopK(0)


def _random_picklists(n, ncols, seed):
    rng = np.random.default_rng(seed)
    cols = []
    for j in range(ncols):
        k = [3, 12, 40, 200][j % 4]
        vocab = [f"v{j}_{r}" for r in range(k)] + [f"V{j} {0}", f"v{j}.{1}!"]       # aliases that clean alike
        p = 1.0 / np.arange(1, len(vocab) + 1) ** 1.1
        p /= p.sum()
        draw = rng.choice(len(vocab), size=n, p=p)
        null = rng.random(n) < 0.07
        cols.append([None if null[i] else vocab[draw[i]] for i in range(n)])
    cols[0][5] = ""     # the empty string is a value, not a null
    return cols


@pytest.mark.parametrize("device", [None, "cuda"])
@pytest.mark.parametrize("clean,track,topk,minsup", [(True, True, 20, 10), (False, True, 5, 0), (True, False, 3, 50)])
def test_pivot_matches_oracle(device, clean, track, topk, minsup):
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.pivot import OpTextPivotVectorizer
    n, ncols = 4000, 6
    cols = _random_picklists(n, ncols, 3)
    names = [f"p{j}" for j in range(ncols)]
    batch = _batch({nm: (F.PICKLIST, c) for nm, c in zip(names, cols)}, device=device)
    est = OpTextPivotVectorizer().setInput(*[F.FeatureBuilder.PickList(nm).asPredictor() for nm in names]) \
        .setTopK(topk).setMinSupport(minsup).setCleanText(clean).setTrackNulls(track)
    model = est.fit(batch)
    exp_tops = V.pivot_fit(cols, topk, minsup, clean)
    assert model.topValues == exp_tops                                        # vocabulary + tie-break: bit-exact
    got = model.transform(batch).vectors[est.getOutput().name].to_scipy()
    exp = V.pivot_transform(cols, exp_tops, clean, track)
    assert got.shape == exp.shape and (got != exp).nnz == 0                   # one-hot positions: bit-exact
    exp_meta = V.pivot_metadata(names, "PickList", exp_tops, track)
    assert [(c.grouping, c.indicatorValue) for c in model.getMetadata().columns] == \
        [(m.grouping, m.indicator) for m in exp_meta]
    # unseen values at transform time go to OTHER (OpSetVectorizerTest.scala:197-229)
    other = _batch({nm: (F.PICKLIST, ["never seen", None, c[0]]) for nm, c in zip(names, cols)}, device=device)
    got2 = model.transform(other).vectors[est.getOutput().name].to_scipy()
    exp2 = V.pivot_transform([["never seen", None, c[0]] for c in cols], exp_tops, clean, track)
    assert (got2 != exp2).nnz == 0


@pytest.mark.parametrize("device", [None, "cuda"])
def test_transmogrify_sanity_check_picklist_and_real_matches_oracle(device):
    """PickList + Real through transmogrify() + sanityCheck(removeBadFeatures): vocabularies, contingency matrices,
    Cramer's V, correlations and the dropped set against the oracle on the oracle's own vectorized matrix."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.dsl import fit_and_transform, transmogrify
    n = 3000
    rng = np.random.default_rng(11)
    y = (rng.random(n) < 0.4).astype(np.float64)
    cols = _random_picklists(n, 4, 5)
    leak = [("yes" if (v and rng.random() < 0.98) else "no") if rng.random() > 0.03 else None for v in y]   # leaky
    cols.append(leak)
    pnames = [f"p{j}" for j in range(5)]
    real = [[None if rng.random() < 0.1 else float(rng.normal(3 + 2 * yy, 1.0)) for yy in y],
            [float(v) for v in rng.normal(0, 1, n)]]
    data = {nm: (F.PICKLIST, c) for nm, c in zip(pnames, cols)}
    data.update({"r0": (F.REAL, real[0]), "r1": (F.REAL, real[1]), "y": (F.REALNN, y.tolist())})
    batch = _batch(data, response="y", device=device)
    feats = [F.FeatureBuilder.PickList(nm).asPredictor() for nm in pnames] + \
        [F.FeatureBuilder.Real("r0").asPredictor(), F.FeatureBuilder.Real("r1").asPredictor()]
    label = F.FeatureBuilder.RealNN("y").asResponse()
    checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True, sampleLowerLimit=0,
                                sampleUpperLimit=10 ** 9, featureFeatureCorrLevel="Off", maxCramersV=0.8)
    out_batch, fitted = fit_and_transform(batch, checked)
    est = checked.origin_stage
    # oracle: PickList block first (type-name order), then Real
    tops = V.pivot_fit(cols, 20, 10, True)
    xp = V.pivot_transform(cols, tops, True, True)
    fills = V.real_vectorizer_fit(real)
    xr = V.real_vectorizer_transform(real, fills, True)
    xo, metas = V.combine([xp, xr], [V.pivot_metadata(pnames, "PickList", tops, True),
                                     V.numeric_metadata(["r0", "r1"], "Real", True)])
    p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                       feature_feature_corr_level="off", max_cramers_v=0.8)
    ref = S.sanity_fit(y, xo, metas, p, label_name="y")
    assert est.summary.names[:-1] == [m.col_name() for m in metas]
    assert est.summary.dropped == ref.dropped and len(ref.dropped) > 0
    assert fitted[-1].indicesToKeep == ref.indices_to_keep
    np.testing.assert_allclose(est.summary.mean, ref.col_stats.mean, rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(est.summary.variance, ref.col_stats.variance, rtol=1e-9, atol=1e-18)
    assert np.array_equal(est.summary.min, ref.col_stats.min) and np.array_equal(est.summary.max, ref.col_stats.max)
    np.testing.assert_allclose(est.summary.correlationsWithLabel, ref.corrs_with_label, rtol=1e-9, atol=1e-12,
                               equal_nan=True)
    got_g = {g.group: g for g in est.summary.categoricalStats}
    assert got_g.keys() == {g.group for g in ref.cat_stats}
    for g in ref.cat_stats:
        assert np.array_equal(got_g[g.group].contingencyMatrix, g.contingency_matrix)      # counts: bit-exact
        np.testing.assert_allclose(got_g[g.group].cramersV, g.cramers_v, rtol=1e-9, equal_nan=True)
        np.testing.assert_allclose(got_g[g.group].mutualInfo, g.mutual_info, rtol=1e-9, atol=1e-15, equal_nan=True)
    got = out_batch.vectors[checked.name].to_scipy()
    exp = xo.tocsc()[:, ref.indices_to_keep].tocsr()
    assert got.shape == exp.shape and abs(got - exp).max() < 1e-9


# core/src/test/scala/com/salesforce/op/stages/impl/feature/OpSetVectorizerTest.scala:57-76 (data and expected vectors)
SET_A = [["a", "b"], ["a"], None, ["c"], ["A", "e"]]
SET_B = [["x", "y"], ["x", "x"], ["z"], None, ["y", "Y"]]


def _set_batch(cols, response=None, y=None, device="cuda"):
    from transmogrifai_b200 import features as F
    data = {nm: (F.MULTIPICKLIST, c) for nm, c in cols.items()}
    if y is not None:
        data["y"] = (F.REALNN, list(y))
    return _batch(data, response=response, device=device)


@pytest.mark.parametrize("clean,track,topk,minsup", [(True, True, 10, 0), (False, True, 2, 1), (True, False, 1, 2)])
def test_set_vectorizer_matches_oracle(clean, track, topk, minsup):
    """OpSetVectorizer on MultiPickList columns: vocabulary, multiplicities of values that clean alike, OTHER sums and
    null indicators bit-exact against the oracle (which is pinned to the reference's OpSetVectorizerTest goldens)."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.sets import OpSetVectorizer
    rng = np.random.default_rng(3)
    vocab = ["red", "Red", "green", "blue", "blue!", "b l u e", "teal", "x-1", "X 1", "zz"]
    big = []
    for _ in range(400):
        u = rng.random()
        big.append(None if u < 0.1 else [] if u < 0.15 else
                   [vocab[i] for i in rng.choice(len(vocab), size=rng.integers(1, 6), replace=False)])
    cols = {"sa": SET_A * 80, "sb": SET_B * 80, "sc": big}
    batch = _set_batch(cols)
    feats = [F.FeatureBuilder.MultiPickList(nm).asPredictor() for nm in cols]
    est = OpSetVectorizer().setInput(*feats).setCleanText(clean).setTrackNulls(track).setTopK(topk) \
        .setMinSupport(minsup)
    model = est.fit(batch)
    lists = list(cols.values())
    tops = V.pivot_fit(lists, topk, minsup, clean)
    assert model.topValues == tops
    got = model.transform(batch).vectors[est.getOutput().name].to_scipy()
    exp = V.pivot_transform(lists, tops, clean, track)
    assert got.shape == exp.shape and (got != exp).nnz == 0
    exp_meta = V.pivot_metadata(list(cols), "MultiPickList", tops, track)
    assert [(c.parentFeatureName, c.grouping, c.indicatorValue) for c in model.getMetadata().columns] == \
        [(m.parent_names, m.grouping, m.indicator) for m in exp_meta]


def test_transmogrify_sanity_check_multipicklist_matches_oracle():
    """MultiPickList through transmogrify() + sanityCheck(): column statistics of the multiplicity-valued columns, the
    clipped contingency tables and the max-over-choices Cramer's V against the oracle."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.dsl import fit_and_transform, transmogrify
    rng = np.random.default_rng(9)
    n = 1500
    y = (rng.random(n) < 0.4).astype(np.float64)
    vocab = ["red", "Red", "green", "blue", "blue!", "teal", "x-1", "X 1"]
    sets = []
    for yy in y:
        u = rng.random()
        if u < 0.08:
            sets.append(None)
            continue
        k = int(rng.integers(1, 5))
        pick = list(rng.choice(len(vocab), size=k, replace=False))
        if yy > 0 and rng.random() < 0.7 and 2 not in pick:
            pick.append(2)                                   # "green" leaks the label
        sets.append([vocab[i] for i in pick])
    r0 = [None if rng.random() < 0.1 else float(rng.normal(yy, 1)) for yy in y]
    data = {"s0": (F.MULTIPICKLIST, sets), "r0": (F.REAL, r0), "y": (F.REALNN, y.tolist())}
    batch = _batch(data, response="y", device="cuda")
    FB = F.FeatureBuilder
    feats = [FB.MultiPickList("s0").asPredictor(), FB.Real("r0").asPredictor()]
    label = FB.RealNN("y").asResponse()
    checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True, sampleLowerLimit=0,
                                sampleUpperLimit=10 ** 9, featureFeatureCorrLevel="Off")
    out_batch, fitted = fit_and_transform(batch, checked)
    est = checked.origin_stage
    tops = V.pivot_fit([sets], 20, 10, True)
    xs = V.pivot_transform([sets], tops, True, True)
    fills = V.real_vectorizer_fit([r0])
    xr = V.real_vectorizer_transform([r0], fills, True)
    xo, metas = V.combine([xs, xr], [V.pivot_metadata(["s0"], "MultiPickList", tops, True),
                                     V.numeric_metadata(["r0"], "Real", True)])
    p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                       feature_feature_corr_level="off")
    ref = S.sanity_fit(y, xo, metas, p, label_name="y")
    assert est.summary.names[:-1] == [m.col_name() for m in metas]
    assert est.summary.dropped == ref.dropped
    assert fitted[-1].indicesToKeep == ref.indices_to_keep
    np.testing.assert_allclose(est.summary.mean, ref.col_stats.mean, rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(est.summary.variance, ref.col_stats.variance, rtol=1e-9, atol=1e-18)
    assert np.array_equal(est.summary.min, ref.col_stats.min) and np.array_equal(est.summary.max, ref.col_stats.max)
    assert ref.col_stats.max.max() >= 2.0                   # "red" + "Red" in one set: a multiplicity of 2
    np.testing.assert_allclose(est.summary.correlationsWithLabel, ref.corrs_with_label, rtol=1e-9, atol=1e-12,
                               equal_nan=True)
    got_g = {g.group: g for g in est.summary.categoricalStats}
    assert got_g.keys() == {g.group for g in ref.cat_stats}
    for g in ref.cat_stats:
        assert np.array_equal(got_g[g.group].contingencyMatrix, g.contingency_matrix)
        np.testing.assert_allclose(got_g[g.group].cramersV, g.cramers_v, rtol=1e-9, equal_nan=True)
    got = out_batch.vectors[checked.name].to_scipy()
    exp = sp.csr_matrix(xo)[:, ref.indices_to_keep]
    assert got.shape == exp.shape and abs(got - exp).max() < 1e-12


def test_set_vectorizer_reference_goldens():
    # core/src/test/scala/com/salesforce/op/stages/impl/feature/OpSetVectorizerTest.scala:57-76 (data, expected
    # vectors), :104-118 (vocabulary order A,C,B / X,Y,Z), :169-179 (topK = 1), :197-229 (unseen -> OTHER, empty -> null)
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.sets import OpSetVectorizer
    top = [["a", "b"], ["a"], ["c"], ["C ", "A."]]
    bot = [["x", "x"], ["z", "y", "z", "z", "y"], ["x", "y"], ["Z", "Z", "Z"]]
    feats = [F.FeatureBuilder.MultiPickList(nm).asPredictor() for nm in ("top", "bot")]
    batch = _set_batch({"top": top, "bot": bot})
    est = OpSetVectorizer().setInput(*feats).setTopK(10).setMinSupport(0)
    model = est.fit(batch)
    assert model.topValues == [["A", "C", "B"], ["X", "Y", "Z"]]
    got = model.transform(batch).vectors[est.getOutput().name].to_numpy()
    exp = np.zeros((4, 10))
    for r, idx in enumerate([[0, 2, 5], [0, 6, 7], [1, 5, 6], [0, 1, 7]]):
        exp[r, idx] = 1.0
    assert np.array_equal(got, exp)
    assert [c.indicatorValue for c in model.getMetadata().columns] == \
        ["A", "C", "B", "OTHER", "NullIndicatorValue", "X", "Y", "Z", "OTHER", "NullIndicatorValue"]
    est1 = OpSetVectorizer().setInput(*feats).setTopK(1).setMinSupport(0)
    got1 = est1.fit(batch).transform(batch).vectors[est1.getOutput().name].to_numpy()
    assert got1.tolist() == [[1, 1, 0, 1, 0, 0], [1, 0, 0, 0, 2, 0], [0, 1, 0, 1, 1, 0], [1, 1, 0, 0, 1, 0]]
    # a model fitted on other data: unseen values go to OTHER, the empty set to the null indicator
    small = _set_batch({"top": [["a", "b"], ["a"], []], "bot": [[], [], []]})
    m2 = OpSetVectorizer().setInput(*feats).setTopK(10).setMinSupport(0).fit(small)
    assert m2.topValues == [["A", "B"], []]
    assert m2.transform(small).vectors[m2.getOutput().name].to_numpy().tolist() == \
        [[1, 1, 0, 0, 0, 1], [1, 0, 0, 0, 0, 1], [0, 0, 0, 1, 0, 1]]
    assert m2.transform(batch).vectors[m2.getOutput().name].to_numpy().tolist() == \
        [[1, 1, 0, 0, 1, 0], [1, 0, 0, 0, 2, 0], [0, 0, 1, 0, 2, 0], [1, 0, 1, 0, 1, 0]]
