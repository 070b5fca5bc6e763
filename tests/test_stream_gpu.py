"""A dataset streamed in row chunks (transmogrifai_b200/streaming.py) must give what one resident batch gives and what
the oracle gives: the folds of FitStagesUtil.scala:212-290 / OpOneHotVectorizer.scala:95-97 /
SmartTextVectorizer.scala:88-91,231-235 do not depend on how the rows are cut into partitions (the TextStats fold is
taken in row order, as the oracle takes it)."""
import numpy as np
import pytest
import scipy.sparse as sp

from oracle import sanity as S
from oracle import vectorizers as V

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

SIZES = [100, 700, 1, 999, 300, 650, 250]          # seven uneven chunks


def _data(n, seed, continuous=False):
    from transmogrifai_b200 import features as F
    rng = np.random.default_rng(seed)
    z = rng.normal(0, 1, n)
    y = (100.0 + 3.0 * z + rng.normal(0, 0.5, n)) if continuous else (rng.random(n) < 0.3).astype(np.float64)
    opt = lambda v, p: None if rng.random() < p else v   # noqa: E731
    b0 = [opt(bool(rng.random() < 0.4 + 0.3 * (yy > y.mean())), 0.1) for yy in y]
    b1 = [opt(bool(rng.random() < 0.5), 0.05) for _ in y]
    p0 = [opt(["x", "y", "z", "X ", "w"][int(rng.integers(5))], 0.08) for _ in y]
    # values that only show up late (other chunks, other dictionaries) and aliases that clean alike
    p1 = [opt(f"v{int(rng.zipf(1.3)) % 60}" if i > 900 else ["a", "A.", "b"][int(rng.integers(3))], 0.05)
          for i in range(n)]
    r0 = [opt(float(rng.normal(2 + (yy > y.mean()), 1)), 0.1) for yy in y]
    r1 = [opt(float(rng.normal(0, 3)), 0.02) for _ in y]
    nn = [float(v) for v in rng.normal(5, 2, n)]
    words = ["".join(rng.choice(list("abcdefgh"), size=rng.integers(3, 7))) for _ in range(150)]
    t0 = [opt(" ".join(words[i] for i in rng.integers(0, 150, rng.poisson(8))), 0.06) for _ in y]
    t1 = [opt(["red", "Green", "blue", "blue!"][int(rng.integers(4))], 0.05) for _ in y]
    # 25 distinct values at first; from row 1300 on (inside the fourth chunk) new ones appear and the map passes the
    # cap of 30 there: the fold freezes in the middle of a chunk
    t2 = [opt(f"k{int(rng.integers(25))}" if i < 1300 else f"k{int(rng.integers(400))}", 0.04) for i in range(n)]
    cols = {"b0": (F.BINARY, b0), "b1": (F.BINARY, b1), "p0": (F.PICKLIST, p0), "p1": (F.PICKLIST, p1),
            "r0": (F.REAL, r0), "r1": (F.REAL, r1), "nn": (F.REALNN, nn), "t0": (F.TEXT, t0), "t1": (F.TEXT, t1),
            "t2": (F.TEXT, t2), "y": (F.REALNN, y.tolist())}
    return cols, y


def _plan():
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.dsl import transmogrify
    FB = F.FeatureBuilder
    feats = [FB.Text("t0").asPredictor(), FB.Real("r0").asPredictor(), FB.Binary("b0").asPredictor(),
             FB.RealNN("nn").asPredictor(), FB.PickList("p0").asPredictor(), FB.Real("r1").asPredictor(),
             FB.Binary("b1").asPredictor(), FB.Text("t1").asPredictor(), FB.PickList("p1").asPredictor(),
             FB.Text("t2").asPredictor()]
    label = FB.RealNN("y").asResponse()
    return label.sanityCheck(transmogrify(feats), removeBadFeatures=True, sampleLowerLimit=0,
                             sampleUpperLimit=10 ** 9, featureFeatureCorrLevel="Off")


def _oracle(cols, y):
    c = {k: v[1] for k, v in cols.items()}
    xb = V.binary_vectorizer_transform([c["b0"], c["b1"]])
    tops = V.pivot_fit([c["p0"], c["p1"]], 20, 10, True)
    xp = V.pivot_transform([c["p0"], c["p1"]], tops, True, True)
    fills = V.real_vectorizer_fit([c["r0"], c["r1"]])
    xr = V.real_vectorizer_transform([c["r0"], c["r1"]], fills, True)
    xn = V.realnn_vectorizer_transform([c["nn"]])
    texts = [c["t0"], c["t1"], c["t2"]]
    methods, ttops = V.smart_text_fit(texts, max_cardinality=30, top_k=20, min_support=10)
    xt = V.smart_text_transform(texts, ["t0", "t1", "t2"], methods, ttops, num_features=512)
    xo, metas = V.combine(
        [xb, xp, xr, xn, xt],
        [V.numeric_metadata(["b0", "b1"], "Binary", True), V.pivot_metadata(["p0", "p1"], "PickList", tops, True),
         V.numeric_metadata(["r0", "r1"], "Real", True), V.numeric_metadata(["nn"], "RealNN", False),
         V.smart_text_metadata(["t0", "t1", "t2"], "Text", methods, ttops, 512, True)])
    p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                       feature_feature_corr_level="off")
    return S.sanity_fit(y, sp.csr_matrix(xo), metas, p, label_name="y"), sp.csr_matrix(xo), metas, methods, tops


def _chunks(cols, mode):
    from transmogrifai_b200.columnar import ColumnarBatch
    n = len(cols["y"][1])
    bounds = np.concatenate([[0], np.cumsum(SIZES)])
    assert bounds[-1] == n
    if mode == "own_dictionaries_host":
        # every chunk exported on its own: pinned host buffers, a dictionary per chunk (first-appearance order)
        return [ColumnarBatch.from_columns({k: (t, v[a:b]) for k, (t, v) in cols.items()}, response="y")
                for a, b in zip(bounds[:-1], bounds[1:])]
    full = ColumnarBatch.from_columns(cols, response="y").to("cuda")        # shared dictionaries, rows on the device
    return [full.take_rows(torch.arange(a, b, device="cuda")) for a, b in zip(bounds[:-1], bounds[1:])]


def _check(est, keep, rows, ref, xo, metas):
    assert est.summary.names[:-1] == [m.col_name() for m in metas]
    assert est.summary.dropped == ref.dropped
    assert list(keep) == ref.indices_to_keep
    assert est.summary.count == xo.shape[0]
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
    exp = xo.tocsc()[:, ref.indices_to_keep].tocsr()
    assert rows.shape == exp.shape and abs(rows - exp).max() < 1e-12


@pytest.mark.parametrize("mode", ["shared_dictionaries_device", "own_dictionaries_host"])
@pytest.mark.parametrize("continuous,cells", [(False, 0), (True, 0), (False, 64 << 20), (True, 64 << 20)])
def test_seven_uneven_chunks_equal_one_batch_equal_oracle(mode, continuous, cells):
    from transmogrifai_b200.columnar import ColumnarBatch, VectorColumn
    from transmogrifai_b200.dsl import fit_and_transform
    from transmogrifai_b200.streaming import fit_and_transform_stream
    cols, y = _data(sum(SIZES), 77, continuous)
    ref, xo, metas, methods, tops = _oracle(cols, y)
    assert methods[:2] == ["Hash", "Pivot"]
    chunks = _chunks(cols, mode)
    checked = _plan()
    out = []

    def sink(k, vec, parts):
        out.append(VectorColumn(vec.name, vec.n, vec.metadata, parts).to_scipy())

    # host chunks stay resident in HBM between the fit and the transform when they fit (here: all seven)
    host = mode == "own_dictionaries_host"
    # cells: the text cells of every chunk stay packed in HBM after pass 1 and pass 2 reads them instead of the text
    res = fit_and_transform_stream(chunks, checked, sink, cache_bytes=None if host and not cells else 0,
                                   cell_cache_bytes=cells)
    est = checked.origin_stage
    assert res.rows == sum(SIZES) and res.chunks == len(SIZES) and len(out) == len(SIZES)
    assert res.cache_hits == (len(SIZES) if host and not cells else 0)
    assert res.cells_cached == (len(SIZES) if cells else 0)
    _check(est, res.fitted[-1].indicesToKeep, sp.vstack(out).tocsr(), ref, xo, metas)
    streamed = (list(est.summary.dropped), est.summary.mean.copy(), est.summary.variance.copy(),
                est.summary.correlationsWithLabel.copy())
    # the same plan over one resident batch
    checked1 = _plan()
    batch = ColumnarBatch.from_columns(cols, response="y").to("cuda")
    out_batch, fitted = fit_and_transform(batch, checked1)
    est1 = checked1.origin_stage
    assert streamed[0] == est1.summary.dropped
    np.testing.assert_allclose(streamed[1], est1.summary.mean, rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(streamed[2], est1.summary.variance, rtol=1e-9, atol=1e-18)
    np.testing.assert_allclose(streamed[3], est1.summary.correlationsWithLabel, rtol=1e-9, atol=1e-12, equal_nan=True)
    one = out_batch.vectors[checked1.name].to_scipy()
    assert abs(one - sp.vstack(out).tocsr()).max() < 1e-12


def test_stream_transform_of_other_rows_and_empty_input():
    """The fitted stages of a streamed fit transform rows they were not fitted on; an empty stream is an error."""
    from transmogrifai_b200.columnar import VectorColumn
    from transmogrifai_b200.streaming import fit_stream, transform_stream
    cols, y = _data(sum(SIZES), 78)
    chunks = _chunks(cols, "own_dictionaries_host")
    checked = _plan()
    fit = fit_stream(chunks[:5], checked)
    out = []
    rows = transform_stream(chunks[5:], checked, fit, lambda k, vec, parts: out.append(
        VectorColumn(vec.name, vec.n, vec.metadata, parts).to_scipy()))
    assert rows == sum(SIZES[5:])
    # against the oracle fitted on the first five chunks' rows and applied to the last two
    m = sum(SIZES[:5])
    head = {k: (t, v[:m]) for k, (t, v) in cols.items()}
    ref, _xo, _metas, methods, tops = _oracle(head, y[:m])
    c = {k: v[1] for k, v in cols.items()}
    hc = {k: v[1] for k, v in head.items()}
    xb = V.binary_vectorizer_transform([c["b0"], c["b1"]])
    xp = V.pivot_transform([c["p0"], c["p1"]], tops, True, True)
    xr = V.real_vectorizer_transform([c["r0"], c["r1"]], V.real_vectorizer_fit([hc["r0"], hc["r1"]]), True)
    xn = V.realnn_vectorizer_transform([c["nn"]])
    texts = [c["t0"], c["t1"], c["t2"]]
    _m, ttops = V.smart_text_fit([hc["t0"], hc["t1"], hc["t2"]], max_cardinality=30, top_k=20, min_support=10)
    xt = V.smart_text_transform(texts, ["t0", "t1", "t2"], methods, ttops, num_features=512)
    full = sp.hstack([sp.csr_matrix(x) for x in (xb, xp, xr, xn, xt)]).tocsr()
    exp = full[m:][:, ref.indices_to_keep]
    got = sp.vstack(out).tocsr()
    assert got.shape == exp.shape and abs(got - exp).max() < 1e-12
    with pytest.raises(ValueError):
        fit_stream([], _plan())


@pytest.mark.parametrize("arena,slow", [(32 << 20, False), (32 << 20, True), (300000, False)])
def test_packed_cells_give_the_rows_the_text_gives(arena, slow, monkeypatch):
    """Pass 2 from the packed cells of pass 1 (tmog_text_hash_transform_packed_dev) against pass 2 from the text: the
    same sparse rows, entry for entry -- with text lengths, an ignored column and dropped columns; with every cell on
    the slow path (slots); and with an arena that holds only the first chunks (the others fall back to their text)."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch, VectorColumn
    from transmogrifai_b200.stages.smart_text import SmartTextVectorizer
    from transmogrifai_b200.streaming import fit_and_transform_stream
    from tests.test_text_paths_gpu import SPECIAL
    rng = np.random.default_rng(11)
    n = 6000
    words = ["".join(rng.choice(list("abcdefghij"), size=rng.integers(2, 9))) for _ in range(400)] + ["the", "and", "it's"]
    def cell(i):
        u = rng.random()
        if u < 0.05:
            return None
        if u < 0.08:
            return SPECIAL[i % len(SPECIAL)]
        return " ".join(words[k] for k in rng.zipf(1.2, size=rng.poisson(25)) % len(words))
    cols = {f"t{j}": (F.TEXT, [cell(i) for i in range(n)]) for j in range(3)}
    cols["id"] = (F.TEXT, [f"row {i}" for i in range(n)])                       # ignored: every value differs
    y = (rng.random(n) < 0.4).astype(np.float64)
    cols["y"] = (F.REALNN, y.tolist())
    bounds = [0, 1000, 1001, 3500, 6000]
    chunks = [ColumnarBatch.from_columns({k: (t, v[a:b]) for k, (t, v) in cols.items()}, response="y")
              for a, b in zip(bounds[:-1], bounds[1:])]

    def run(cell_bytes):
        feats = [F.FeatureBuilder.Text(nm).asPredictor() for nm in ("t0", "t1", "t2", "id")]
        label = F.FeatureBuilder.RealNN("y").asResponse()
        vec = SmartTextVectorizer().setMaxCardinality(5).setNumFeatures(512).setTrackTextLen(True) \
            .setMinLengthStdDev(0.5).setInput(*feats).getOutput()
        checked = label.sanityCheck(vec, removeBadFeatures=True, sampleLowerLimit=0, sampleUpperLimit=10 ** 9,
                                    featureFeatureCorrLevel="Off", minVariance=1e-3)
        out = []
        res = fit_and_transform_stream(chunks, checked, lambda k, v, parts: out.append(
            VectorColumn(v.name, v.n, v.metadata, parts).to_scipy()), cell_cache_bytes=cell_bytes)
        return sp.vstack(out).tocsr(), res, checked.origin_stage.summary.dropped

    if slow:
        monkeypatch.setenv("TMOG_TEXT_FORCE_SLOW", "1")
    ref, res0, dropped0 = run(0)
    got, res1, dropped1 = run(arena)
    assert res0.cells_cached == 0 and dropped0 == dropped1 and len(dropped0) > 0
    assert 0 < res1.cells_cached <= len(chunks) and (res1.cells_cached == len(chunks)) == (arena > (1 << 20))
    assert got.shape == ref.shape and abs(got - ref).max() == 0
    assert np.array_equal(got.indptr, ref.indptr) and np.array_equal(got.indices, ref.indices)


def test_compact_return_format_round_trip():
    """HostSink(compact=True): 0 / 1 columns of dense blocks and term frequencies cross PCIe as bytes; expand_compact
    gives back exactly the vector the plain sink delivers (and the oracle's)."""
    from transmogrifai_b200.columnar import VectorColumn
    from transmogrifai_b200.streaming import HostSink, expand_compact, fit_and_transform_stream, part_widths
    cols, y = _data(sum(SIZES), 79)
    ref, xo, metas, _methods, _tops = _oracle(cols, y)
    chunks = _chunks(cols, "own_dictionaries_host")
    plain, widths = [], []

    def sink(k, vec, parts):
        plain.append(VectorColumn(vec.name, vec.n, vec.metadata, parts).to_scipy())
        widths.append(part_widths(parts))

    fit_and_transform_stream(chunks, _plan(), sink)
    got, nbytes = [], [0, 0]
    order = []

    def consume(k, vec, named):
        order.append(k)
        got.append(expand_compact(vec, widths[k], named, SIZES[k]))
        nbytes[1] += sum(t.numel() * t.element_size() for _nm, t in named)

    hs = HostSink(depth=2, consume=consume, compact=True)
    fit_and_transform_stream(chunks, _plan(), hs, cell_cache_bytes=32 << 20)
    hs0 = HostSink(depth=2)
    fit_and_transform_stream(chunks, _plan(), hs0)
    assert order == list(range(len(SIZES)))
    a, b = sp.vstack(got).tocsr(), sp.vstack(plain).tocsr()
    assert a.shape == b.shape and abs(a - b).max() == 0
    exp = xo.tocsc()[:, ref.indices_to_keep].tocsr()
    assert abs(a - exp).max() < 1e-12
    assert hs.bytes == nbytes[1] and hs.bytes < 0.8 * hs0.bytes


@pytest.mark.parametrize("with_mode", [True, False])
def test_integral_vectorizer_over_chunks(with_mode):
    """IntegralVectorizer fitted over row chunks (IntegralVectorizer.scala:81-113, SequenceAggregators.scala:100-131): the
    value -> count maps of the chunks add up key-wise (mode, ties to the smaller value), the statistics merge like
    RealVectorizer's; against the oracle and against one resident batch."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch, VectorColumn
    from transmogrifai_b200.dsl import fit_and_transform
    from transmogrifai_b200.stages.integral import IntegralVectorizer
    from transmogrifai_b200.streaming import fit_and_transform_stream
    rng = np.random.default_rng(21)
    n = sum(SIZES)
    y = (rng.random(n) < 0.4).astype(np.float64)
    # i0: the mode changes from chunk to chunk and ties at the end (7 and 3 equally often -> 3); i1: all null; i2: wide
    i0 = [None if rng.random() < 0.1 else int(v) for v in rng.choice([3, 7, 11, -5], size=n, p=[0.3, 0.3, 0.25, 0.15])]
    c3, c7 = sum(v == 3 for v in i0), sum(v == 7 for v in i0)
    major, todo = (3, c3 - c7) if c3 > c7 else (7, c7 - c3)          # force the tie: surplus of the larger one -> 11
    for k in range(n):
        if todo and i0[k] == major:
            i0[k] = 11
            todo -= 1
    assert sum(v == 3 for v in i0) == sum(v == 7 for v in i0) > sum(v == 11 for v in i0)
    i1 = [None] * n
    i2 = [None if rng.random() < 0.3 else int(v) for v in rng.integers(-10 ** 12, 10 ** 12, size=n)]
    cols = {"i0": (F.INTEGRAL, i0), "i1": (F.INTEGRAL, i1), "i2": (F.INTEGRAL, i2), "y": (F.REALNN, y.tolist())}
    fills = V.integral_vectorizer_fit([i0, i1, i2], fill_with_mode=with_mode)
    if with_mode:
        assert fills[0] == 3 and fills[1] == 0
    exp = sp.csr_matrix(V.integral_vectorizer_transform([i0, i1, i2], fills, True))

    def plan():
        FB = F.FeatureBuilder
        est = IntegralVectorizer().setInput(FB.Integral("i0").asPredictor(), FB.Integral("i1").asPredictor(),
                                            FB.Integral("i2").asPredictor())
        if with_mode:
            est.setFillWithMode()
        label = FB.RealNN("y").asResponse()
        return est, label.sanityCheck(est.getOutput(), removeBadFeatures=False, sampleLowerLimit=0,
                                      sampleUpperLimit=10 ** 9, featureFeatureCorrLevel="Off")

    bounds = np.concatenate([[0], np.cumsum(SIZES)])
    chunks = [ColumnarBatch.from_columns({k_: (t, v[a:b]) for k_, (t, v) in cols.items()}, response="y")
              for a, b in zip(bounds[:-1], bounds[1:])]
    est, checked = plan()
    out = []
    res = fit_and_transform_stream(chunks, checked, lambda i, vec, parts: out.append(
        VectorColumn(vec.name, vec.n, vec.metadata, parts).to_scipy()))
    got = sp.vstack(out).tocsr()
    model = [m for m in res.fitted if hasattr(m, "fillValues")][0]
    assert model.fillValues == [int(f) for f in fills]
    assert got.shape == exp.shape and abs(got - exp).max() == 0
    est1, checked1 = plan()
    batch = ColumnarBatch.from_columns(cols, response="y").to("cuda")
    fit_and_transform(batch, checked1)
    s0, s1 = checked.origin_stage.summary, checked1.origin_stage.summary
    np.testing.assert_allclose(s0.mean, s1.mean, rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(s0.variance, s1.variance, rtol=1e-9, atol=1e-18)
    np.testing.assert_allclose(s0.correlationsWithLabel, s1.correlationsWithLabel, rtol=1e-9, atol=1e-12, equal_nan=True)
    assert np.array_equal(s0.min, s1.min) and np.array_equal(s0.max, s1.max)


@pytest.mark.parametrize("continuous", [False, True])
def test_set_vectorizer_over_chunks(continuous):
    """OpSetVectorizer fitted over row chunks with a dictionary per chunk (OpOneHotVectorizer.scala:75-124, :416-437): the
    counts of pass 0 add up per cleaned value, the statistics of the multiplicity-valued columns come from pass 1;
    against the oracle and against one resident batch."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch, VectorColumn
    from transmogrifai_b200.dsl import fit_and_transform
    from transmogrifai_b200.stages.sets import OpSetVectorizer
    from transmogrifai_b200.streaming import fit_and_transform_stream
    rng = np.random.default_rng(33)
    n = sum(SIZES)
    z = rng.normal(0, 1, n)
    y = (100.0 + 3.0 * z + rng.normal(0, 0.5, n)) if continuous else (z > 0.2).astype(np.float64)
    opt = lambda v, p: None if rng.random() < p else v   # noqa: E731
    base = [["a", "A "], ["b"], ["a", "c", "d"], [], ["b", "B.", "b"]]
    s0 = [opt(list(base[int(rng.integers(5))]) + (["hi"] if zz > 0 else []) + ([f"rare{i}"] if i > 2500 and i % 7 == 0 else []),
              0.07) for i, zz in enumerate(z)]
    s1 = [opt([f"v{int(k)}" for k in rng.zipf(1.4, size=rng.integers(0, 4)) % 30], 0.1) for _ in range(n)]
    cols = {"s0": (F.MULTIPICKLIST, s0), "s1": (F.MULTIPICKLIST, s1), "y": (F.REALNN, y.tolist())}
    tops = V.pivot_fit([s0, s1], 20, 10, True)
    xo = sp.csr_matrix(V.pivot_transform([s0, s1], tops, True, True))
    metas = V.pivot_metadata(["s0", "s1"], "MultiPickList", tops, True)
    p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                       feature_feature_corr_level="off")
    ref = S.sanity_fit(y, xo, metas, p, label_name="y")

    def plan():
        FB = F.FeatureBuilder
        est = OpSetVectorizer().setTopK(20).setMinSupport(10).setInput(FB.MultiPickList("s0").asPredictor(),
                                                                        FB.MultiPickList("s1").asPredictor())
        label = FB.RealNN("y").asResponse()
        return label.sanityCheck(est.getOutput(), removeBadFeatures=True, sampleLowerLimit=0, sampleUpperLimit=10 ** 9,
                                 featureFeatureCorrLevel="Off")

    bounds = np.concatenate([[0], np.cumsum(SIZES)])
    chunks = [ColumnarBatch.from_columns({k_: (t, v[a:b]) for k_, (t, v) in cols.items()}, response="y")
              for a, b in zip(bounds[:-1], bounds[1:])]
    checked = plan()
    out = []
    res = fit_and_transform_stream(chunks, checked, lambda i, vec, parts: out.append(
        VectorColumn(vec.name, vec.n, vec.metadata, parts).to_scipy()))
    est = checked.origin_stage
    model = [m for m in res.fitted if hasattr(m, "topValues")][0]
    assert [list(t) for t in model.topValues] == [list(t) for t in tops]
    _check(est, res.fitted[-1].indicesToKeep, sp.vstack(out).tocsr(), ref, xo, metas)
    checked1 = plan()
    batch = ColumnarBatch.from_columns(cols, response="y").to("cuda")
    out1, _f = fit_and_transform(batch, checked1)
    assert checked1.origin_stage.summary.dropped == est.summary.dropped
    assert abs(out1.vectors[checked1.name].to_scipy() - sp.vstack(out).tocsr()).max() == 0


def test_sanity_checker_sample_over_chunks():
    """SanityChecker with fraction() < 1 over a chunk-streamed dataset (SanityChecker.scala:356-387): the statistics are
    those of a row sample drawn chunk by chunk, taken by the fitted stages (fills and vocabularies of ALL rows) and merged;
    against the oracle on exactly the sampled rows, and the transform of all rows with the kept columns."""
    from transmogrifai_b200.columnar import VectorColumn
    from transmogrifai_b200.streaming import fit_and_transform_stream
    cols, y = _data(sum(SIZES), 91)
    chunks = _chunks(cols, "own_dictionaries_host")
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.dsl import transmogrify
    FB = F.FeatureBuilder
    feats = [FB.Text("t0").asPredictor(), FB.Real("r0").asPredictor(), FB.Binary("b0").asPredictor(),
             FB.RealNN("nn").asPredictor(), FB.PickList("p0").asPredictor(), FB.Real("r1").asPredictor(),
             FB.Binary("b1").asPredictor(), FB.Text("t1").asPredictor(), FB.PickList("p1").asPredictor(),
             FB.Text("t2").asPredictor()]
    label = FB.RealNN("y").asResponse()
    checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True, checkSample=0.4, sampleLowerLimit=0,
                                sampleUpperLimit=10 ** 9, featureFeatureCorrLevel="Off")
    out = []
    res = fit_and_transform_stream(chunks, checked, lambda k, vec, parts: out.append(
        VectorColumn(vec.name, vec.n, vec.metadata, parts).to_scipy()))
    est = checked.origin_stage
    assert res.sample_rows is not None and len(res.sample_rows) == len(SIZES)
    bounds = np.concatenate([[0], np.cumsum(SIZES)])
    rows = np.concatenate([r.cpu().numpy() + a for r, a in zip(res.sample_rows, bounds[:-1])])
    assert 0.3 * sum(SIZES) < len(rows) < 0.5 * sum(SIZES) and est.summary.count == len(rows)
    assert abs(est.summary.sampleFraction - 0.4) < 1e-12
    # the oracle: vectorizers fitted on all rows, SanityChecker's statistics on the sampled ones
    _ref_all, xo, metas, _methods, _tops = _oracle(cols, y)
    p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                       feature_feature_corr_level="off")
    ref = S.sanity_fit(y[rows], xo[rows], metas, p, label_name="y")
    assert est.summary.dropped == ref.dropped
    assert list(res.fitted[-1].indicesToKeep) == ref.indices_to_keep
    np.testing.assert_allclose(est.summary.mean, ref.col_stats.mean, rtol=1e-10, atol=1e-300)
    np.testing.assert_allclose(est.summary.variance, ref.col_stats.variance, rtol=1e-9, atol=1e-18)
    assert np.array_equal(est.summary.min, ref.col_stats.min) and np.array_equal(est.summary.max, ref.col_stats.max)
    np.testing.assert_allclose(est.summary.correlationsWithLabel, ref.corrs_with_label, rtol=1e-9, atol=1e-12,
                               equal_nan=True)
    got_g = {g.group: g for g in est.summary.categoricalStats}
    for g in ref.cat_stats:
        assert np.array_equal(got_g[g.group].contingencyMatrix, g.contingency_matrix)
        np.testing.assert_allclose(got_g[g.group].cramersV, g.cramers_v, rtol=1e-9, equal_nan=True)
    exp = xo.tocsc()[:, ref.indices_to_keep].tocsr()
    got = sp.vstack(out).tocsr()
    assert got.shape == exp.shape and abs(got - exp).max() < 1e-12


@pytest.mark.parametrize("sampled", [False, True])
def test_feature_feature_correlations_over_chunks(sampled):
    """CorrelationLevel.Computed (the reference's default, SanityChecker.scala:467-470) over a chunk-streamed dataset: one
    more pass builds the checker's input vector chunk by chunk (of the sampled rows when the checker samples) and adds
    every chunk's rows to one Gramian; the full Pearson matrix, the drop set and the kept columns against the oracle on
    the same rows."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import VectorColumn
    from transmogrifai_b200.dsl import transmogrify
    from transmogrifai_b200.streaming import fit_and_transform_stream
    cols, y = _data(sum(SIZES), 58, continuous=True)
    chunks = _chunks(cols, "own_dictionaries_host")
    FB = F.FeatureBuilder
    feats = [FB.Text("t0").asPredictor(), FB.Real("r0").asPredictor(), FB.Binary("b0").asPredictor(),
             FB.RealNN("nn").asPredictor(), FB.PickList("p0").asPredictor(), FB.Real("r1").asPredictor(),
             FB.Binary("b1").asPredictor(), FB.Text("t1").asPredictor(), FB.PickList("p1").asPredictor(),
             FB.Text("t2").asPredictor()]
    label = FB.RealNN("y").asResponse()
    checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True, checkSample=0.5 if sampled else 1.0,
                                sampleLowerLimit=0, sampleUpperLimit=10 ** 9, featureFeatureCorrLevel="Computed")
    out = []
    res = fit_and_transform_stream(chunks, checked, lambda k, vec, parts: out.append(
        VectorColumn(vec.name, vec.n, vec.metadata, parts).to_scipy()), cell_cache_bytes=64 << 20)
    est = checked.origin_stage
    n = sum(SIZES)
    if sampled:
        bounds = np.concatenate([[0], np.cumsum(SIZES)])
        rows = np.concatenate([r.cpu().numpy() + a for r, a in zip(res.sample_rows, bounds[:-1])])
        assert 0.4 * n < len(rows) < 0.6 * n
    else:
        assert res.sample_rows is None
        rows = np.arange(n)
    assert est.summary.count == len(rows)
    _ref_all, xo, metas, _methods, _tops = _oracle(cols, y)
    p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                       feature_feature_corr_level="computed")
    ref = S.sanity_fit(y[rows], xo[rows], metas, p, label_name="y")
    m = est.summary.featureFeatureCorr
    assert m is not None and m.shape == ref.corr_matrix.shape == (xo.shape[1] + 1, xo.shape[1] + 1)
    np.testing.assert_allclose(m, ref.corr_matrix, rtol=1e-9, atol=1e-12, equal_nan=True)
    np.testing.assert_allclose(est.summary.correlationsWithLabel, ref.corrs_with_label, rtol=1e-9, atol=1e-12,
                               equal_nan=True)
    assert est.summary.dropped == ref.dropped
    assert list(res.fitted[-1].indicesToKeep) == ref.indices_to_keep
    exp = xo.tocsc()[:, ref.indices_to_keep].tocsr()
    got = sp.vstack(out).tocsr()
    assert got.shape == exp.shape and abs(got - exp).max() < 1e-12
