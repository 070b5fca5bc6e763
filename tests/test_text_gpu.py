"""GPU parity of SmartTextVectorizer (TextStats decision, pivot of categorical text, tokenise + murmur3 hashing, null
indicators) against the oracle and the reference's golden vectors."""
import numpy as np
import pytest

from oracle import sanity as S
from oracle import text as T
from oracle import vectorizers as V

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

TEXT1 = ["hello world", "hello world", "good evening", "hello world", None]
TEXT2 = ["Hello world!", "What's up", "How are you doing, my friend?", "Not bad, my friend.", None]


def _batch(cols, response=None, device=None):
    from transmogrifai_b200.columnar import ColumnarBatch
    b = ColumnarBatch.from_columns(cols, response=response)
    return b.to(device) if device else b


def _feats(names):
    from transmogrifai_b200 import features as F
    return [F.FeatureBuilder.Text(nm).asPredictor() for nm in names]


@pytest.mark.parametrize("device", [None, "cuda"])
def test_smart_text_vectorizer_reference_golden(device):
    # core/src/test/scala/com/salesforce/op/stages/impl/feature/SmartTextVectorizerTest.scala:55-75
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.smart_text import SmartTextVectorizer
    batch = _batch({"text1": (F.TEXT, TEXT1), "text2": (F.TEXT, TEXT2)}, device=device)
    est = SmartTextVectorizer().setMaxCardinality(2).setNumFeatures(4).setMinSupport(1).setTopK(2) \
        .setPrependFeatureName(False).setHashSpaceStrategy("Shared").setInput(*_feats(["text1", "text2"]))
    model = est.fit(batch)
    assert model.vectorizationMethods == ["Pivot", "Hash"]
    got = model.transform(batch).vectors[est.getOutput().name].to_numpy()
    exp = np.zeros((5, 9))
    for r, (idx, vals) in enumerate([((0, 4, 6), (1, 1, 1)), ((0, 8), (1, 1)), ((1, 6), (1, 1)), ((0, 6), (1, 2)),
                                     ((3, 8), (1, 1))]):
        exp[r, list(idx)] = vals
    assert np.array_equal(got, exp)
    meta = model.getMetadata()
    assert [c.indicatorValue for c in meta.columns[:4]] == ["HelloWorld", "GoodEvening", "OTHER", "NullIndicatorValue"]
    assert all(c.grouping is None and c.indicatorValue is None and c.parentFeatureName == ("text2",)
               for c in meta.columns[4:8])
    assert meta.columns[8].indicatorValue == "NullIndicatorValue" and meta.columns[8].grouping == "text2"


def test_hashing_reference_golden_indices():
    # core/src/test/.../OPCollectionHashingVectorizerTest.scala:47-58: feature textList1, tokens x y, 512 buckets,
    # prepended feature-name hash -> {107, 224}
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.smart_text import SmartTextVectorizer, murmur3_index
    assert murmur3_index("textList1", 512) == 364
    batch = _batch({"textList1": (F.TEXT, ["x y", "zebra"])})
    est = SmartTextVectorizer().setMaxCardinality(0).setNumFeatures(512).setTrackNulls(False) \
        .setInput(*_feats(["textList1"]))
    model = est.fit(batch)
    assert model.vectorizationMethods == ["Hash"]
    got = model.transform(batch).vectors[est.getOutput().name].to_scipy()
    assert sorted(got[0].indices.tolist()) == [107, 224] and got[0].data.tolist() == [1.0, 1.0]


TRICKY = [
    "I've got a lovely bunch of coconuts", "There they are, all standing in a row",
    "<body>Big ones, small <h1>ones</h1>, some as big as your head</body>", "", "   ", "What's up", None,
    "www.google.com and foo_bar9 cost 3.14 or 1,000 units; a:b a,b 9:3 x.'y _ __ _a_ 'tis don't.", "ALL CAPS Words",
    "trailing.", ".leading", "a" * 300 + " tail", "tab\tsep\nnewline", "it's its its' 'its", "x" * 31 + " " + "y" * 33,
    "e-mail me@site.org 10:30pm 2019-01-02 v1.2.3 a.b.c.d", "under_score__double ___ 9_9",
]


def _random_text(rng, vocab, n, null_frac=0.05):
    out = []
    for _ in range(n):
        u = rng.random()
        if u < null_frac:
            out.append(None)
        elif u < null_frac + 0.01:
            out.append("")
        else:
            k = rng.poisson(12)
            toks = [vocab[i] for i in rng.zipf(1.3, size=k) % len(vocab)]
            s = " ".join(toks)
            if rng.random() < 0.1:
                s = s.title()
            if rng.random() < 0.1:
                s = s.replace(" ", ", ", 2) + "."
            out.append(s)
    return out


@pytest.mark.parametrize("device", [None, "cuda"])
@pytest.mark.parametrize("num_features,strategy,prepend,binary,track,min_len",
                         [(512, "Auto", True, False, True, 1), (16, "Shared", False, False, True, 1),
                          (64, "Separate", True, True, False, 3)])
def test_hash_columns_match_oracle(device, num_features, strategy, prepend, binary, track, min_len):
    """Tokeniser + stop filter + legacy murmur3 + term frequencies + null indicators: bit-exact against the oracle on
    hand-made tricky strings and random Zipf text."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.smart_text import SmartTextVectorizer
    rng = np.random.default_rng(5)
    vocab = ["".join(rng.choice(list("abcdefghijklmnopqrstuvwxyz"), size=rng.integers(2, 9))) for _ in range(400)]
    vocab += ["the", "and", "you", "9", "42", "x1", "don't", "u.s.a", "A_b"]
    n = 600
    cols = [TRICKY + _random_text(rng, vocab, n - len(TRICKY)), _random_text(rng, vocab, n),
            _random_text(rng, vocab, n, null_frac=0.5)]
    names = ["alpha", "beta", "gamma_text"]
    batch = _batch({nm: (F.TEXT, c) for nm, c in zip(names, cols)}, device=device)
    est = SmartTextVectorizer().setMaxCardinality(0).setNumFeatures(num_features).setHashSpaceStrategy(strategy) \
        .setPrependFeatureName(prepend).setBinaryFreq(binary).setTrackNulls(track).setMinTokenLength(min_len) \
        .setInput(*_feats(names))
    model = est.fit(batch)
    assert model.vectorizationMethods == ["Hash"] * 3
    got = model.transform(batch).vectors[est.getOutput().name].to_scipy()
    exp = V.smart_text_transform(cols, names, ["Hash"] * 3, [[], [], []], num_features=num_features,
                                 track_nulls=track, prepend_feature_name=prepend, binary_freq=binary,
                                 hash_space_strategy=strategy, min_token_length=min_len)
    assert got.shape == exp.shape
    diff = (got != exp)
    assert diff.nnz == 0, f"{diff.nnz} entries differ, first rows {sorted(set(diff.nonzero()[0]))[:5]}"
    exp_meta = V.smart_text_metadata(names, "Text", ["Hash"] * 3, [[], [], []], num_features, track,
                                     hash_space_strategy=strategy)
    assert [(c.parentFeatureName, c.grouping, c.indicatorValue) for c in model.getMetadata().columns] == \
        [(m.parent_names, m.grouping, m.indicator) for m in exp_meta]


@pytest.mark.parametrize("device", [None, "cuda"])
def test_smart_text_decision_and_pivot_match_oracle(device):
    """Mixed columns: low-cardinality text is pivoted (vocabulary + tie-break bit-exact), free text is hashed, a
    column just over the cardinality cap follows the reference's frozen left fold."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.smart_text import SmartTextVectorizer
    rng = np.random.default_rng(9)
    n = 1500
    cat = [None if rng.random() < 0.1 else ["Red", "green", "BLUE ", "red.", "Green!", "yellow"][int(rng.integers(6))]
           for _ in range(n)]
    vocab = ["".join(rng.choice(list("abcdefghij"), size=rng.integers(3, 8))) for _ in range(300)]
    free = _random_text(rng, vocab, n)
    ids = [f"id{int(rng.integers(0, 40)):03d}" if rng.random() > 0.05 else None for _ in range(n)]   # 40 > maxCard 30
    heavy = ["usa" if rng.random() < 0.97 else f"c{int(rng.integers(0, 60))}" for _ in range(n)]      # covered by top-K
    cols = [cat, free, ids, heavy]
    names = ["cat", "free", "ids", "heavy"]
    batch = _batch({nm: (F.TEXT, c) for nm, c in zip(names, cols)}, device=device)
    est = SmartTextVectorizer().setMaxCardinality(30).setNumFeatures(128).setTopK(20).setMinSupport(10) \
        .setInput(*_feats(names))
    model = est.fit(batch)
    methods, tops = V.smart_text_fit(cols, max_cardinality=30, top_k=20, min_support=10)
    assert model.vectorizationMethods == methods
    assert methods[0] == "Pivot" and methods[1] == "Hash"
    assert [model.topValues[j] for j in range(4) if methods[j] == "Pivot"] == \
        [tops[j] for j in range(4) if methods[j] == "Pivot"]
    got = model.transform(batch).vectors[est.getOutput().name].to_scipy()
    exp = V.smart_text_transform(cols, names, methods, tops, num_features=128)
    assert got.shape == exp.shape and (got != exp).nnz == 0
    exp_meta = V.smart_text_metadata(names, "Text", methods, tops, 128, True)
    assert [(c.parentFeatureName, c.grouping, c.indicatorValue) for c in model.getMetadata().columns] == \
        [(m.parent_names, m.grouping, m.indicator) for m in exp_meta]


@pytest.mark.parametrize("device", [None, "cuda"])
def test_transmogrify_sanity_check_text_matches_oracle(device):
    """Text + Real through transmogrify() + sanityCheck(): statistics of hashed columns and null indicators from the
    fused bucket sums against the oracle's colStats / label correlations / Cramer's V on the oracle's own matrix."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.dsl import fit_and_transform, transmogrify
    rng = np.random.default_rng(21)
    n = 1200
    y = (rng.random(n) < 0.35).astype(np.float64)
    vocab = ["".join(rng.choice(list("abcdefghijklmnop"), size=rng.integers(3, 8))) for _ in range(200)]
    t0 = _random_text(rng, vocab, n)
    t1 = [("spam offer " if yy and rng.random() < 0.9 else "") + (s or "") if rng.random() > 0.04 else None
          for yy, s in zip(y, _random_text(rng, vocab, n))]
    cat = [None if rng.random() < 0.1 else ["a", "b", "c"][int(rng.integers(3))] for _ in range(n)]
    r0 = [None if rng.random() < 0.1 else float(v) for v in rng.normal(1.0, 2.0, n)]
    data = {"t0": (F.TEXT, t0), "t1": (F.TEXT, t1), "cat": (F.TEXT, cat), "r0": (F.REAL, r0),
            "y": (F.REALNN, y.tolist())}
    batch = _batch(data, response="y", device=device)
    feats = _feats(["t0", "t1", "cat"]) + [F.FeatureBuilder.Real("r0").asPredictor()]
    label = F.FeatureBuilder.RealNN("y").asResponse()
    checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True, sampleLowerLimit=0,
                                sampleUpperLimit=10 ** 9, featureFeatureCorrLevel="Off")
    out_batch, fitted = fit_and_transform(batch, checked)
    est = checked.origin_stage
    # oracle: Real block first (type-name order Real < Text), then the SmartText block (maxCard 30, 512 hashes)
    fills = V.real_vectorizer_fit([r0])
    xr = V.real_vectorizer_transform([r0], fills, True)
    tcols, tnames = [t0, t1, cat], ["t0", "t1", "cat"]
    methods, tops = V.smart_text_fit(tcols, max_cardinality=30, top_k=20, min_support=10)
    assert methods == ["Hash", "Hash", "Pivot"]
    xt = V.smart_text_transform(tcols, tnames, methods, tops, num_features=512)
    xo, metas = V.combine([xr, xt], [V.numeric_metadata(["r0"], "Real", True),
                                     V.smart_text_metadata(tnames, "Text", methods, tops, 512, True)])
    p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                       feature_feature_corr_level="off")
    ref = S.sanity_fit(y, xo, metas, p, label_name="y")
    assert est.summary.names[:-1] == [m.col_name() for m in metas]
    assert est.summary.dropped == ref.dropped and len(ref.dropped) > 100       # empty hash buckets have no variance
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


def _assert_rows_sorted(vec):
    """Sparse rows come out with strictly increasing indices (what Spark's SparseVector requires)."""
    from transmogrifai_b200.columnar import SparseRowsPart
    for part in vec.parts:
        part = part.materialize(None) if hasattr(part, "materialize") else part
        if not isinstance(part, SparseRowsPart):
            continue
        start, nnz, idx = part.row_start.cpu().numpy(), part.row_nnz.cpu().numpy(), part.indices.cpu().numpy()
        for r in range(len(start)):
            row = idx[start[r]:start[r] + nnz[r]]
            assert (np.diff(row) > 0).all(), f"row {r} is not sorted"


def _hash_case(cols, names, num_features, strategy, **kw):
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.smart_text import SmartTextVectorizer
    batch = _batch({nm: (F.TEXT, c) for nm, c in zip(names, cols)}, device="cuda")
    est = SmartTextVectorizer().setMaxCardinality(0).setNumFeatures(num_features).setHashSpaceStrategy(strategy) \
        .setInput(*_feats(names))
    model = est.fit(batch)
    vec = model.transform(batch).vectors[est.getOutput().name]
    _assert_rows_sorted(vec)
    got = vec.to_scipy()
    exp = V.smart_text_transform(cols, names, ["Hash"] * len(names), [[] for _ in names], num_features=num_features,
                                 track_nulls=True, prepend_feature_name=True, binary_freq=False,
                                 hash_space_strategy=strategy, min_token_length=1)
    assert got.shape == exp.shape
    diff = (got != exp)
    assert diff.nnz == 0, f"{diff.nnz} entries differ, first rows {sorted(set(diff.nonzero()[0]))[:5]}"
    return got


def test_hash_long_cells_and_crowded_rows():
    """Cells longer than the kernel's text buffer (staged in pieces cut at a word break), rows with more distinct
    buckets than the fast table holds (redone by the second launch), runs longer than maxTokenLength, '_' runs."""
    rng = np.random.default_rng(11)
    words = ["".join(rng.choice(list("abcdefghijklmnopqrstuvwxyz0123456789"), size=rng.integers(1, 12)))
             for _ in range(6000)]
    long_cell = " ".join(words[i] for i in rng.integers(0, 6000, size=2500))          # ~16 KB, ~2000 distinct
    seps = [" ", ", ", ".", "'", ":", ";", "_", " - ", "\n"]
    messy = "".join(words[i] + seps[i % len(seps)] for i in rng.integers(0, 300, size=1500))
    col_a = [long_cell, "short one", None, messy, "b" * 255 + " " + "c" * 256 + " " + "d" * 700 + " end",
             "_" * 10 + " " + "_" * 300 + "z" + " _x_ __", " ".join(words[:4000]), "", "z" * 2046, "y" * 2015 + " q"]
    col_b = [" ".join(words[i] for i in rng.integers(0, 6000, size=rng.integers(0, 900))) for _ in col_a]
    col_b[2] = None
    for strategy in ("Shared", "Separate"):
        _hash_case([col_a, col_b], ["long_a", "long_b"], 1 << 12, strategy)


def test_hash_token_run_longer_than_buffer_is_refused():
    with pytest.raises(ValueError, match="longer than"):
        _hash_case([["ok text", "a" * 9000 + " tail"], ["x", "y"]], ["t", "u"], 64, "Shared")
    # separate hash spaces of <= 512 buckets run the per-cell kernels, whose byte-wise slow path has no such limit
    _hash_case([["ok text", "a" * 9000 + " tail"], ["x", "y"]], ["t", "u"], 64, "Separate")


def test_hash_many_columns_wide_keys_and_count_overflow():
    """More than 32 (several cell passes) and more than 64 columns (prefix state not cached) with separate spaces wide
    enough for the 64-bit table; and a bucket met > 255 times in one row with 2^23 buckets, where the packed 32-bit
    slot overflows and the row is redone with 64-bit slots."""
    rng = np.random.default_rng(13)
    vocab = ["".join(rng.choice(list("abcdefghijklmnopqrstuvwxyz"), size=rng.integers(2, 9))) for _ in range(300)]
    names = [f"col{j}" for j in range(70)]
    cols = [_random_text(rng, vocab, 40, null_frac=0.2) for _ in names]
    _hash_case(cols, names, 1 << 18, "Separate")
    _hash_case(cols[:40], names[:40], 1 << 10, "Separate")
    rep = [" ".join(["repeat"] * 700 + ["other"] * 3), "repeat once", " ".join(["w%d" % (i % 7) for i in range(2000)])]
    got = _hash_case([rep], ["rep"], 1 << 23, "Shared")
    assert got[0].data.max() == 700


def test_text_stats_values_entry_point_matches_the_full_scan():
    """tmog_text_stats_values_dev (no lengths past the cardinality cap) against tmog_text_stats_dev: identical value
    tables, distinct counts and non-null counts; the length sums agree for a column that stays under the cap and,
    for a capped one, cover at most what the full scan covers."""
    import ctypes as C
    from transmogrifai_b200 import _lib as L
    from transmogrifai_b200.stages.smart_text import _col_desc
    rng = np.random.default_rng(5)
    n = 20000
    low = [None if rng.random() < 0.1 else f"Val-{rng.integers(0, 12)}!" for _ in range(n)]     # 12 distinct values
    high = [None if rng.random() < 0.1 else f"free text {i} with, punctuation." for i in range(n)]
    dev = _batch({"low": ("Text", low), "high": ("Text", high)}, device="cuda").texts(["low", "high"])
    slots, cap = 128, 30
    head, per = C.c_size_t(), C.c_size_t()
    L.call("tmog_text_stats_layout", 2, slots, 1, C.byref(head), C.byref(per))
    desc = _col_desc(dev)
    out = {}
    for name in ("tmog_text_stats_dev", "tmog_text_stats_values_dev"):
        tables = torch.zeros(head.value + per.value * 2, dtype=torch.uint8, device="cuda")
        flags = torch.zeros(4, dtype=torch.int32, device="cuda")
        L.call(name, desc.data_ptr(), 2, n, None, 1, 1, slots, cap, tables.data_ptr(), flags.data_ptr(),
               torch.cuda.current_stream().cuda_stream)
        assert int(flags[0].item()) == 0
        out[name] = tables[head.value:].cpu().numpy().reshape(2, per.value)
    full, vals = out["tmog_text_stats_dev"], out["tmog_text_stats_values_dev"]
    tail = 16 * slots + ((4 * slots + 7) & ~7)
    assert np.array_equal(full[0], vals[0])                                   # under the cap: everything agrees
    nd = [int(t[1, tail:tail + 4].view(np.int32)[0]) for t in (full, vals)]
    assert nd[0] > cap and nd[1] > cap
    nn_full, nn_vals = (int(t[1, tail + 24:tail + 32].view(np.uint64)[0]) for t in (full, vals))
    assert 0 < nn_vals <= nn_full                                             # the capped column is not scanned on
    ls_full, ls_vals = full[1, tail + 8:tail + 24].view(np.uint64), vals[1, tail + 8:tail + 24].view(np.uint64)
    assert (ls_vals <= ls_full).all() and ls_full[0] > 0


def test_non_ascii_text_in_a_capped_column_is_still_refused():
    """A column past the cardinality cap is no longer read by the TextStats scan; the hashing pass must then be the one
    that refuses non-ASCII text (the tokenizer is restated for ASCII only)."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.smart_text import SmartTextVectorizer
    col = [f"distinct value number {i}" for i in range(300)] + ["café au lait"]
    batch = _batch({"t": (F.TEXT, col)}, device="cuda")
    est = SmartTextVectorizer().setMaxCardinality(30).setNumFeatures(128).setInput(*_feats(["t"]))
    with pytest.raises(ValueError, match="non-ASCII"):
        model = est.fit(batch)
        model.transform(batch).vectors[est.getOutput().name].to_scipy()


def _ignore_route_data(seed=11, n=1000):
    rng = np.random.default_rng(seed)
    opt = lambda v: None if rng.random() < 0.2 else v   # noqa: E731
    places = [" ".join("".join(rng.choice(list("abcdefghijklmnop"), size=rng.integers(3, 9))).capitalize()
                       for _ in range(rng.integers(1, 4))) for _ in range(250)]
    country = [opt(places[int(rng.integers(len(places)))]) for _ in range(n)]
    categorical = [opt("ABCDEF"[int(rng.integers(6))]) for _ in range(n)]
    text_id = [opt("%06d" % (40231 + int(rng.integers(1000)))) for _ in range(n)]
    text = [opt("".join(rng.choice(list("abc xyz,. QRS"), size=rng.integers(0, 101)))) for _ in range(n)]
    return [country, categorical, text_id, text], ["country", "categorical", "textId", "text"]


@pytest.mark.parametrize("tokens,min_std,max_card", [(False, 0.5, 10), (True, 0.5, 300), (False, 0.5, 60)])
@pytest.mark.parametrize("chunked", [False, True])
def test_smart_text_ignore_route_and_text_lengths(tokens, min_std, max_card, chunked):
    """SmartTextVectorizerTest.scala:213-280: machine-generated ids (fixed length) are ignored -- tokenised for their
    null indicator and text length only; layout [pivot | hash | lengths (hashed, then ignored) | nulls]; lengthCounts
    under the maxCardinality cap of the reference's fold (:231-235) and with textLengthType = Tokens (:85)."""
    import scipy.sparse as sp
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch, VectorColumn
    from transmogrifai_b200.stages.smart_text import SmartTextVectorizer
    cols, names = _ignore_route_data()
    methods, tops = V.smart_text_fit(cols, max_cardinality=max_card, top_k=3, min_support=10,
                                     min_length_std_dev=min_std, tokenize_for_lengths=tokens)
    if max_card == 10:
        assert methods == [V.HASH, V.PIVOT, V.IGNORE, V.HASH]
    exp = V.smart_text_transform(cols, names, methods, tops, num_features=5, track_nulls=True, track_text_len=True)
    meta = V.smart_text_metadata(names, "Text", methods, tops, num_features=5, track_nulls=True, track_text_len=True)
    est = SmartTextVectorizer().setMaxCardinality(min(max_card, 3000)).setTopK(3).setMinSupport(10).setNumFeatures(5) \
        .setMinLengthStdDev(min_std).setTextLengthType("Tokens" if tokens else "FullEntry").setTrackTextLen(True) \
        .setInput(*_feats(names))
    if not chunked:
        batch = _batch({nm: (F.TEXT, c) for nm, c in zip(names, cols)}, device="cuda")
        model = est.fit(batch)
        got = model.transform(batch).vectors[est.getOutput().name].to_scipy()
    else:
        from transmogrifai_b200.streaming import fit_stream, transform_stream
        bounds = [0, 37, 400, 401, 1000]
        chunks = [ColumnarBatch.from_columns({nm: (F.TEXT, c[a:b]) for nm, c in zip(names, cols)})
                  for a, b in zip(bounds[:-1], bounds[1:])]
        fit = fit_stream(chunks, est.getOutput())
        model = fit.fitted[-1]
        out = []
        transform_stream(chunks, est.getOutput(), fit, lambda k, vec, parts: out.append(
            VectorColumn(vec.name, vec.n, vec.metadata, parts).to_scipy()))
        got = sp.vstack(out).tocsr()
    assert model.vectorizationMethods == methods
    assert [list(t) for t in model.topValues] == [list(t) for t in tops]
    assert got.shape == exp.shape and abs(got - sp.csr_matrix(exp)).max() == 0
    cm = model.getMetadata().columns
    assert [(c.parentFeatureName, c.grouping, c.indicatorValue, c.descriptorValue) for c in cm] == \
        [(m.parent_names, m.grouping, m.indicator, m.descriptor) for m in meta]


def test_smart_text_lengths_and_ignored_columns_through_sanity_checker():
    """The text-length columns and the null indicators of ignored columns carry SanityChecker statistics like any other
    column (fused sidecar), against the oracle."""
    import scipy.sparse as sp
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.smart_text import SmartTextVectorizer
    cols, names = _ignore_route_data(seed=12)
    rng = np.random.default_rng(3)
    y = (rng.random(len(cols[0])) < 0.4).astype(np.float64)
    methods, tops = V.smart_text_fit(cols, max_cardinality=10, top_k=3, min_support=10, min_length_std_dev=0.5)
    xo = sp.csr_matrix(V.smart_text_transform(cols, names, methods, tops, num_features=5, track_text_len=True))
    metas = V.smart_text_metadata(names, "Text", methods, tops, num_features=5, track_nulls=True, track_text_len=True)
    p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                       feature_feature_corr_level="off")
    ref = S.sanity_fit(y, xo, metas, p, label_name="y")
    data = {nm: (F.TEXT, c) for nm, c in zip(names, cols)}
    data["y"] = (F.REALNN, y.tolist())
    batch = _batch(data, response="y", device="cuda")
    est = SmartTextVectorizer().setMaxCardinality(10).setTopK(3).setMinSupport(10).setNumFeatures(5) \
        .setMinLengthStdDev(0.5).setTrackTextLen(True).setInput(*_feats(names))
    label = F.FeatureBuilder.RealNN("y").asResponse()
    checked = label.sanityCheck(est.getOutput(), removeBadFeatures=True, sampleLowerLimit=0, sampleUpperLimit=10 ** 9,
                                featureFeatureCorrLevel="Off")
    from transmogrifai_b200.dsl import fit_and_transform
    out, fitted = fit_and_transform(batch, checked)
    sm = checked.origin_stage.summary
    assert sm.dropped == ref.dropped
    np.testing.assert_allclose(sm.mean, ref.col_stats.mean, rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(sm.variance, ref.col_stats.variance, rtol=1e-9, atol=1e-18)
    assert np.array_equal(sm.min, ref.col_stats.min) and np.array_equal(sm.max, ref.col_stats.max)
    np.testing.assert_allclose(sm.correlationsWithLabel, ref.corrs_with_label, rtol=1e-9, atol=1e-12, equal_nan=True)
    got = out.vectors[checked.name].to_scipy()
    exp = xo.tocsc()[:, ref.indices_to_keep].tocsr()
    assert got.shape == exp.shape and abs(got - exp).max() == 0
