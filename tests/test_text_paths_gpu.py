"""The two tokenise + hash implementations must agree bit for bit: the per-cell kernels (text_cells.cuh: separate hash
spaces of <= 512 buckets), their byte-wise slow path for the cells the fast path sets aside, and the per-row table kernel
(text_hash_kernel) -- on text of the bench generator's shape and on hand-made cells that hit every special case
(TextTokenizer.scala:160-220, OPCollectionHashingVectorizer.scala:244-272)."""
import os

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

SPECIAL = [
    "What's up www.google.com 3.14 1,000 a,b foo_bar9 ___ _a_ x:y 10:30 I've don't. 'tis", None, "", "   ",
    "the and of to a in is it you that",                                  # stop words only -> null indicator
    "z" * 300 + " tail " + "q" * 255 + " " + "w" * 256,                   # tokens longer than maxTokenLength
    "m" * 44 + " " + "n" * 45 + " " + "o" * 47 + " p" * 3,                # around the ring's reach
    " ".join(f"w{i} w{i}" for i in range(40)),                            # 40 repeated buckets: more than a record holds
    " ".join(["same"] * 500),                                             # one bucket 500 times
    "ALL CAPS AND Mixed Case, with; punct: u.s.a. 1.2.3.4 a.1 1.a _._ a_.b",
    "x" * 15 + " " + "y" * 16 + " " + "z" * 17 + " " + "a" * 31 + " " + "b" * 32 + " " + "c" * 33,
    "tab\tsep\nnewline\rcr end", "trailing.", ".leading", "a", "ab", "9", "_",
    # term frequencies 4, 5, 6 and 1 in one cell (the packed cells' trailer), and two of them in a row
    "aa aa aa aa bb bb bb bb bb cc cc cc cc cc cc dd", "qq qq qq qq qq qq qq qq qq rr rr rr rr", "solo",
]


def _model_outputs(batch, names, num_features, binary, min_len, track, label):
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.stages.smart_text import SmartTextVectorizer
    feats = [F.FeatureBuilder.Text(nm).asPredictor() for nm in names]
    est = SmartTextVectorizer().setMaxCardinality(0).setNumFeatures(num_features).setHashSpaceStrategy("Separate") \
        .setBinaryFreq(binary).setMinTokenLength(min_len).setTrackNulls(track).setInput(*feats)
    for k in ("_label_cache", "_label_dev", "_dev_cache"):
        batch.__dict__.pop(k, None)
    model = est.fit(batch)
    st = model._fit_stats
    rows = model.transform(batch).vectors[est.getOutput().name].to_scipy()
    # and the transform-only pass (no rows left over from the fit)
    from transmogrifai_b200.columnar import ColumnarBatch
    other = ColumnarBatch(batch.n, batch.blocks, {}, None)
    rows2 = model.transform(other).vectors[est.getOutput().name].to_scipy()
    return st, rows, rows2


@pytest.mark.parametrize("num_features,binary,min_len,track", [(512, False, 1, True), (100, True, 2, False)])
def test_cell_path_slow_path_and_row_table_path_agree(num_features, binary, min_len, track):
    import benchdata
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch, NumericBlock, TextBlock
    n = 40_000
    g = torch.Generator(device="cuda").manual_seed(123)
    y = (torch.rand(benchdata._ld(n), device="cuda", generator=g) < 0.3).double()
    tb = benchdata.text_block(g, n, 5)
    rng = np.random.default_rng(9)
    special = [SPECIAL[i % len(SPECIAL)] if rng.random() < 0.5 else
               " ".join(rng.choice(["alpha", "Beta", "the", "gamma9", "x_y", "it's", "3.5", "end."], size=rng.integers(0, 60)))
               for i in range(n)]
    sb = TextBlock.from_lists(["s0"], F.TEXT, [special]).to("cuda")
    block = TextBlock(tb.names + sb.names, F.TEXT, tb.data + sb.data, tb.offsets + sb.offsets, tb.valid + sb.valid, n)
    label = NumericBlock(["label"], F.REALNN, y.view(1, -1), None, n)
    batch = ColumnarBatch(n, [block, label], {}, response="label")
    names = block.names
    out = {}
    try:
        for mode, env in (("span", {}), ("cells", {"TMOG_TEXT_PATH": "cells"}), ("slow", {"TMOG_TEXT_FORCE_SLOW": "1"}),
                          ("rows", {"TMOG_TEXT_PATH": "rows"})):
            for k in ("TMOG_TEXT_FORCE_SLOW", "TMOG_TEXT_PATH"):
                os.environ.pop(k, None)
            os.environ.update(env)
            out[mode] = _model_outputs(batch, names, num_features, binary, min_len, track, y)
    finally:
        for k in ("TMOG_TEXT_FORCE_SLOW", "TMOG_TEXT_PATH"):
            os.environ.pop(k, None)
    ref_st, ref_rows, ref_rows2 = out["rows"]
    assert abs(ref_rows - ref_rows2).max() == 0
    for mode in ("span", "cells", "slow"):
        st, rows, rows2 = out[mode]
        for k in ("mean", "variance", "min", "max", "cov_label"):
            assert np.array_equal(getattr(st, k), getattr(ref_st, k)), (mode, k)
        assert np.array_equal(np.nan_to_num(st.class_sums, nan=-1.0), np.nan_to_num(ref_st.class_sums, nan=-1.0)), mode
        assert rows.shape == ref_rows.shape and abs(rows - ref_rows).max() == 0, mode
        assert abs(rows2 - ref_rows).max() == 0, mode


def test_cell_path_matches_oracle_on_special_cells():
    from oracle import vectorizers as V
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch
    from transmogrifai_b200.stages.smart_text import SmartTextVectorizer
    cols = [SPECIAL, list(reversed(SPECIAL))]
    names = ["first_col", "second"]
    batch = ColumnarBatch.from_columns({nm: (F.TEXT, c) for nm, c in zip(names, cols)}).to("cuda")
    feats = [F.FeatureBuilder.Text(nm).asPredictor() for nm in names]
    for nf in (512, 37):
        est = SmartTextVectorizer().setMaxCardinality(0).setNumFeatures(nf).setHashSpaceStrategy("Separate") \
            .setInput(*feats)
        model = est.fit(batch)
        got = model.transform(batch).vectors[est.getOutput().name].to_scipy()
        exp = V.smart_text_transform(cols, names, ["Hash"] * 2, [[], []], num_features=nf, hash_space_strategy="Separate")
        import scipy.sparse as sp
        assert abs(got - sp.csr_matrix(exp)).max() == 0


def _fuzz_cells(rng, n, heavy_punct):
    """Random ASCII cells: letters, digits, '_', the Mid* characters and separators in changing proportions, lengths
    from 0 to a few thousand bytes, runs of token characters of every length around the fast path's limits."""
    letters = "abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ"
    alphabet = letters * 3 + "0123456789" * (3 if heavy_punct else 1) + "_" * (4 if heavy_punct else 1) + \
        (".':,;" * 6 if heavy_punct else ".,") + "   \t\n-/()!?\"&" * (2 if heavy_punct else 4)
    alphabet = np.frombuffer(alphabet.encode(), dtype=np.uint8)
    out = []
    for i in range(n):
        u = rng.random()
        if u < 0.04:
            out.append(None)
        elif u < 0.07:
            out.append("")
        else:
            length = int(rng.choice([1, 2, 3, 15, 16, 17, 31, 32, 33, 47, 48, 63, 64, 65, 200, 511, 512, 513, 1500, 3000])
                         if rng.random() < 0.3 else rng.integers(1, 400))
            b = alphabet[rng.integers(0, len(alphabet), size=length)].copy()
            if rng.random() < 0.1:                      # a long run of token characters somewhere
                k = int(rng.integers(40, 60))
                at = int(rng.integers(0, max(1, length - k)))
                b[at:at + k] = ord("x")
            out.append(b.tobytes().decode())
    return out


@pytest.mark.parametrize("heavy_punct,num_features,min_len", [(True, 512, 1), (False, 100, 2)])
def test_fuzz_span_kernel_against_row_table_kernel_and_oracle(heavy_punct, num_features, min_len):
    """Random cells (see _fuzz_cells): the span kernel, the lane-per-cell kernel and the per-row table kernel give the
    same statistics and rows; a sample of the rows against the oracle."""
    from oracle import vectorizers as V
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch, NumericBlock, TextBlock
    rng = np.random.default_rng(2024 + int(heavy_punct))
    n = 6011                                            # not a multiple of 32
    cols = [_fuzz_cells(rng, n, heavy_punct) for _ in range(3)]
    names = ["fz0", "fz1", "fz2"]
    block = TextBlock.from_lists(names, F.TEXT, cols).to("cuda")
    y = torch.from_numpy((rng.random(n) < 0.5).astype(np.float64)).cuda()
    label = NumericBlock(["label"], F.REALNN, y.view(1, -1), None, n)
    batch = ColumnarBatch(n, [block, label], {}, response="label")
    out = {}
    try:
        for mode, env in (("span", {}), ("cells", {"TMOG_TEXT_PATH": "cells"}), ("rows", {"TMOG_TEXT_PATH": "rows"})):
            for k in ("TMOG_TEXT_FORCE_SLOW", "TMOG_TEXT_PATH"):
                os.environ.pop(k, None)
            os.environ.update(env)
            out[mode] = _model_outputs(batch, names, num_features, False, min_len, True, y)
    finally:
        for k in ("TMOG_TEXT_FORCE_SLOW", "TMOG_TEXT_PATH"):
            os.environ.pop(k, None)
    ref_st, ref_rows, _ = out["rows"]
    for mode in ("span", "cells"):
        st, rows, rows2 = out[mode]
        for k in ("mean", "variance", "min", "max", "cov_label"):
            assert np.array_equal(getattr(st, k), getattr(ref_st, k)), (mode, k)
        assert abs(rows - ref_rows).max() == 0 and abs(rows2 - ref_rows).max() == 0, mode
    take = 700
    exp = V.smart_text_transform([c[:take] for c in cols], names, ["Hash"] * 3, [[], [], []], num_features=num_features,
                                 hash_space_strategy="Separate", min_token_length=min_len)
    import scipy.sparse as sp
    assert abs(out["span"][1][:take] - sp.csr_matrix(exp)).max() == 0


@pytest.mark.parametrize("binary_freq", [False, True])
def test_packed_cells_with_more_than_32_text_columns(binary_freq):
    """The packed row kernel walks the cells of a row 32 at a time: 40 hashed columns, against the rows from the text
    (term frequencies, and binaryFreq = true)."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch, VectorColumn
    from transmogrifai_b200.stages.smart_text import SmartTextVectorizer
    from transmogrifai_b200.streaming import fit_and_transform_stream
    import scipy.sparse as sp
    rng = np.random.default_rng(5)
    n, k = 700, 40
    words = ["".join(rng.choice(list("abcdefgh"), size=rng.integers(2, 8))) for _ in range(200)]
    cols = {f"c{j:02d}": (F.TEXT, [None if rng.random() < 0.1 else " ".join(words[i] for i in rng.integers(0, 200, rng.poisson(6)))
                                   for _ in range(n)]) for j in range(k)}
    cols["y"] = (F.REALNN, (rng.random(n) < 0.5).astype(np.float64).tolist())
    chunks = [ColumnarBatch.from_columns({nm: (t, v[a:b]) for nm, (t, v) in cols.items()}, response="y")
              for a, b in ((0, 333), (333, 700))]

    def run(cell_bytes):
        feats = [F.FeatureBuilder.Text(f"c{j:02d}").asPredictor() for j in range(k)]
        label = F.FeatureBuilder.RealNN("y").asResponse()
        vec = SmartTextVectorizer().setMaxCardinality(0).setNumFeatures(64).setBinaryFreq(binary_freq) \
            .setInput(*feats).getOutput()
        checked = label.sanityCheck(vec, removeBadFeatures=True, sampleLowerLimit=0, sampleUpperLimit=10 ** 9,
                                    featureFeatureCorrLevel="Off", minVariance=2e-3)
        out = []
        res = fit_and_transform_stream(chunks, checked, lambda i, v, parts: out.append(
            VectorColumn(v.name, v.n, v.metadata, parts).to_scipy()), cell_cache_bytes=cell_bytes)
        return sp.vstack(out).tocsr(), res

    ref, _r0 = run(0)
    got, r1 = run(8 << 20)
    assert r1.cells_cached == 2 and got.shape == ref.shape and got.shape[1] < k * 64 + k
    assert abs(got - ref).max() == 0 and np.array_equal(got.indices, ref.indices)


def test_packed_cells_group_too_large_for_its_offsets_falls_back_to_the_text():
    """32 consecutive cells that hold (nearly) every one of 512 buckets need more bytes than a cell's 14-bit offset inside
    its group can address: the chunk is then not packed and its transform comes from the text -- same rows."""
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch, VectorColumn
    from transmogrifai_b200.stages.smart_text import SmartTextVectorizer
    from transmogrifai_b200.streaming import fit_and_transform_stream
    import scipy.sparse as sp
    from transmogrifai_b200.stages.smart_text import murmur3_index
    rng = np.random.default_rng(8)
    n = 96
    # one token per bucket: a cell with all 512 buckets and no repeat stays on the fast path and takes 512 bytes
    prefix = f"{murmur3_index('big', 512)}_"
    by_bucket = {}
    i = 0
    while len(by_bucket) < 512:
        by_bucket.setdefault(murmur3_index(prefix + f"w{i}", 512), f"w{i}")
        i += 1
    words = list(by_bucket.values())
    big = [" ".join(words[k] for k in rng.permutation(512)) for _ in range(n)]
    small = ["a few words here" for _ in range(n)]
    y = (rng.random(n) < 0.5).astype(np.float64).tolist()
    chunks = [ColumnarBatch.from_columns({"big": (F.TEXT, big), "small": (F.TEXT, small), "y": (F.REALNN, y)}, response="y"),
              ColumnarBatch.from_columns({"big": (F.TEXT, small), "small": (F.TEXT, small), "y": (F.REALNN, y)}, response="y")]

    def run(cell_bytes):
        feats = [F.FeatureBuilder.Text("big").asPredictor(), F.FeatureBuilder.Text("small").asPredictor()]
        label = F.FeatureBuilder.RealNN("y").asResponse()
        vec = SmartTextVectorizer().setMaxCardinality(0).setNumFeatures(512).setInput(*feats).getOutput()
        checked = label.sanityCheck(vec, removeBadFeatures=False, sampleLowerLimit=0, sampleUpperLimit=10 ** 9,
                                    featureFeatureCorrLevel="Off")
        out = []
        res = fit_and_transform_stream(chunks, checked, lambda i, v, parts: out.append(
            VectorColumn(v.name, v.n, v.metadata, parts).to_scipy()), cell_cache_bytes=cell_bytes)
        return sp.vstack(out).tocsr(), res

    ref, _r0 = run(0)
    got, r1 = run(16 << 20)
    assert r1.cells_cached == 1                       # the second chunk's cells are small
    assert abs(got - ref).max() == 0 and np.array_equal(got.indices, ref.indices)
