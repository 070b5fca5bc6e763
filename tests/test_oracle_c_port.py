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


TRICKY = [
    "I've got a lovely bunch of coconuts", "There they are, all standing in a row",
    "<body>Big ones, small <h1>ones</h1>, some as big as your head</body>", "", "   ", "What's up", None,
    "www.google.com and foo_bar9 cost 3.14 or 1,000 units; a:b a,b 9:3 x.'y _ __ _a_ 'tis don't.", "ALL CAPS Words",
    "trailing.", ".leading", "a" * 300 + " tail", "tab\tsep\nnewline", "it's its its' 'its", "x" * 31 + " " + "y" * 33,
    "e-mail me@site.org 10:30pm 2019-01-02 v1.2.3 a.b.c.d", "under_score__double ___ 9_9", "_" * 300 + "z",
]


@pytest.mark.parametrize("threads", [1, 3])
@pytest.mark.parametrize("num_features,strategy,prepend,binary,track,min_len",
                         [(512, "Separate", True, False, True, 1), (16, "Shared", False, False, True, 1),
                          (64, "Shared", True, True, False, 3)])
def test_c_port_text_hash_matches_python_oracle(threads, num_features, strategy, prepend, binary, track, min_len):
    from transmogrifai_b200.stages.smart_text import murmur3_index
    rng = np.random.default_rng(5)
    vocab = ["".join(rng.choice(list("abcdefghijklmnopqrstuvwxyz"), size=rng.integers(2, 9))) for _ in range(200)]
    vocab += ["the", "and", "you", "9", "42", "x1", "don't", "u.s.a", "A_b"]

    def rand_col(n, null_frac):
        out = []
        for _ in range(n):
            if rng.random() < null_frac:
                out.append(None)
            else:
                out.append(" ".join(vocab[i] for i in rng.zipf(1.3, size=rng.poisson(10)) % len(vocab)))
        return out
    n = 300
    cols = [TRICKY + rand_col(n - len(TRICKY), 0.05), rand_col(n, 0.05), rand_col(n, 0.5)]
    names = ["alpha", "beta", "gamma_text"]
    shared = strategy == "Shared"
    enc = CP.encode_text_columns(cols)
    prefixes = [f"{murmur3_index(nm, num_features)}_" if prepend else "" for nm in names]
    y = rng.integers(0, 2, n).astype(np.uint8)
    res = CP.text_hash(enc, n, prefixes, num_features, shared, binary, min_len, track, class_ids=y, nclasses=2,
                       threads=threads)
    exp = V.smart_text_transform(cols, names, ["Hash"] * 3, [[], [], []], num_features=num_features,
                                 track_nulls=True, prepend_feature_name=prepend, binary_freq=binary,
                                 hash_space_strategy=strategy, min_token_length=min_len).tocsr()
    width = res["width"]
    import scipy.sparse as sp
    indptr = np.concatenate([[0], np.cumsum(res["row_nnz"])])
    assert np.array_equal(res["row_start"], indptr[:-1])
    got = sp.csr_matrix((res["val"].astype(np.float64), res["idx"], indptr), shape=(n, width))
    if not track:
        exp = exp[:, :width - 3]
        got = got[:, :width - 3]
    assert (got != exp).nnz == 0
    for r in range(n):
        assert (np.diff(res["idx"][indptr[r]:indptr[r + 1]]) > 0).all()
    dense = np.asarray(got.todense())
    w = dense.shape[1]
    assert np.array_equal(res["class_sum"][:w, 1], dense[y == 1].sum(axis=0))
    assert np.array_equal(res["sq_sum"][:w], (dense ** 2).sum(axis=0))
    assert np.array_equal(res["nnz"][:w], (dense > 0).sum(axis=0))
    assert np.array_equal(res["vmax"][:w], dense.max(axis=0))


def test_c_port_clean_string_and_text_stats():
    from oracle import text as T
    for s in ["A string wit#h %bad pun&ctuation mark<=>s", "hello  world", " x ", "a.b", "Tab\tkept", "", "ALL CAPS"]:
        assert CP.clean_string(s) == T.clean_string(s), s
    cols = [["red", "Red ", "blue", None, "re-d", "blue"], ["a", "b", "c", "d", "e", "f"]]
    enc = CP.encode_text_columns(cols)
    distinct, nonnull, ls, lq = CP.text_stats(enc, 6, True, 3, threads=2)
    assert distinct.tolist() == [3, 4] and nonnull.tolist() == [5, 6]       # 4 = "more than maxCardinality"
    assert ls[0] == sum(len(T.clean_string(v)) for v in cols[0] if v is not None)


def test_c_port_pivot_counts_and_transform():
    rng = np.random.default_rng(2)
    n, ld = 1000, 1000
    codes = rng.integers(-1, 7, size=(3, ld)).astype(np.int32)
    y = rng.integers(0, 3, n).astype(np.uint8)
    got = CP.pivot_counts(codes, n, np.array([7, 7, 7]), y, 3, threads=3)
    for c in range(3):
        for code in range(-1, 7):
            for k in range(3):
                assert got[c][code + 1, k] == ((codes[c, :n] == code) & (y == k)).sum()
    tables = [np.arange(8, dtype=np.int32) + 10 * c for c in range(3)]
    out = CP.pivot_transform(codes, n, tables, threads=2)
    for c in range(3):
        assert np.array_equal(out[:, c], np.where(codes[c, :n] < 0, 7, codes[c, :n]) + 10 * c)
