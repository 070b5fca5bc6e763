"""Pins the CPU oracle against every golden vector / known answer the reference's own tests hold for the hot path.

Each test names the reference test file:line it transcribes (paths relative to /root/reference;
CT = core/src/test/scala/com/salesforce/op, UT = utils/src/test/scala/com/salesforce/op/utils).
"""
import math

import numpy as np
import pytest

from oracle import sanity as S
from oracle import text as T
from oracle import vectorizers as V
from oracle.metadata import ColumnMeta, NULL_STRING, OTHER_STRING, TEXT_LEN_STRING


def dense(m):
    return np.asarray(m.todense()) if hasattr(m, "todense") else np.asarray(m)


def sparse_row(width, idx, vals):
    out = np.zeros(width)
    out[list(idx)] = vals
    return out


# ------------------------------------------------------------------------------------------------ text utils
def test_clean_string_golden():
    # UT/text/TextUtilsTest.scala:56-58
    assert T.clean_string("A string wit#h %bad pun&ctuation mark<=>s") == "AStringWitHBadPunCtuationMarkS"
    # CT/stages/impl/feature/SmartTextVectorizerTest.scala:102-103,584-590
    assert T.clean_string("I have got a LovEly buncH of cOcOnuts. Here they are ALL standing in a row.") == \
        "IHaveGotALovelyBunchOfCoconutsHereTheyAreAllStandingInARow"
    assert T.clean_string("") == ""
    assert T.clean_string(" .,; ") == ""
    assert T.clean_string("  leading") == "Leading"


def test_tokenizer_golden():
    # CT/stages/impl/feature/TextTokenizerTest.scala:44-85
    assert T.tokenize("I've got a lovely bunch of coconuts") == ["got", "lovely", "bunch", "coconuts"]
    assert T.tokenize("There they are, all standing in a row") == ["standing", "row"]
    assert T.tokenize("Big ones, small ones, some as big as your head") == \
        ["big", "ones", "small", "ones", "big", "head"]
    assert T.tokenize("<body>Big ones, small <h1>ones</h1>, some as big as your head</body>") == \
        ["body", "big", "ones", "small", "h1", "ones", "h1", "big", "head", "body"]
    assert T.tokenize("") == []
    assert T.tokenize(None) == []
    # :178 -- dots between letters keep a host name in one token
    assert "www.google.com" in T.tokenize("<a href=www.google.com>x</a>")
    # SmartTextVectorizerTest.scala:57,70 -- every token a stop word => no tokens
    assert T.tokenize("What's up") == []


def test_hashing_tf_golden():
    # CT/stages/impl/feature/OPCollectionHashingVectorizerTest.scala:47-58 : textList1 / (x, y) / 512 -> {107, 224}
    name_hash = T.hashing_tf_index("textList1", 512)
    assert name_hash == 364
    got = sorted(T.hashing_tf_index(f"{name_hash}_{t}", 512) for t in ("x", "y"))
    assert got == [107, 224]
    ent = V.hash_token_lists([["x", "y"]], ["textList1"], 512, shared=False)
    assert ent == [(107, 1.0), (224, 1.0)]
    # canonical murmur3 would give 227 / {136, 489}: the legacy tail handling matters
    assert T.hashing_tf_index("hello", 4) == 0 and T.hashing_tf_index("world", 4) == 2
    assert T.hashing_tf_index("friend", 4) == 2 and T.hashing_tf_index("bad", 4) == 2


# ------------------------------------------------------------------------------------------------ numeric
REAL_DATA = [[4.0, None, 2.0], [2.0, 2.0, None], [None, None, None]]  # inA, inB, inC  (RealVectorizerTest.scala:47-53)


def test_real_vectorizer_golden():
    # CT/stages/impl/feature/RealVectorizerTest.scala:118-145 (means 3.0 / 2.0 / all-null -> 0.0)
    fills = V.real_vectorizer_fit(REAL_DATA)
    assert fills == [3.0, 2.0, 0.0]
    out = V.real_vectorizer_transform(REAL_DATA, fills, track_nulls=False)
    assert out.tolist() == [[4.0, 2.0, 0.0], [3.0, 2.0, 0.0], [2.0, 2.0, 0.0]]
    # :182-214 interleaved [value, isNull]
    out = V.real_vectorizer_transform(REAL_DATA, fills, track_nulls=True)
    assert out.tolist() == [[4.0, 0.0, 2.0, 0.0, 0.0, 1.0], [3.0, 1.0, 2.0, 0.0, 0.0, 1.0],
                            [2.0, 0.0, 2.0, 1.0, 0.0, 1.0]]
    # :85-116 constant fill 4.2 ; :148-180 constant 0.0 tracked
    fills = V.real_vectorizer_fit(REAL_DATA, fill_with_mean=False, fill_value=4.2)
    out = V.real_vectorizer_transform(REAL_DATA, fills, track_nulls=False)
    assert out.tolist() == [[4.0, 2.0, 4.2], [4.2, 2.0, 4.2], [2.0, 4.2, 4.2]]
    out = V.real_vectorizer_transform(REAL_DATA, [0.0] * 3, track_nulls=True)
    assert out.tolist() == [[4.0, 0.0, 2.0, 0.0, 0.0, 1.0], [0.0, 1.0, 2.0, 0.0, 0.0, 1.0],
                            [2.0, 0.0, 0.0, 1.0, 0.0, 1.0]]
    meta = V.numeric_metadata(["inA", "inB", "inC"], "Real", True)
    assert [(m.grouping, m.indicator) for m in meta] == [(None, None), ("inA", NULL_STRING), (None, None),
                                                          ("inB", NULL_STRING), (None, None), ("inC", NULL_STRING)]


def test_binary_vectorizer_golden():
    # CT/stages/impl/feature/BinaryVectorizerTest.scala:45-70
    f1 = [False, False, True, True, None, None, False, True, None]
    f2 = [False, True, False, True, False, True, None, None, None]
    exp = [[0, 0, 0, 0], [0, 0, 1, 0], [1, 0, 0, 0], [1, 0, 1, 0], [0, 1, 0, 0], [0, 1, 1, 0], [0, 0, 0, 1],
           [1, 0, 0, 1], [0, 1, 0, 1]]
    assert V.binary_vectorizer_transform([f1, f2]).tolist() == exp


def test_realnn_vectorizer():
    # CT/stages/impl/feature/RealNNVectorizerTest.scala:45
    out = V.realnn_vectorizer_transform([[1.0, 2.0], [3.0, 4.0]])
    assert out.tolist() == [[1.0, 3.0], [2.0, 4.0]]
    with pytest.raises(ValueError):
        V.realnn_vectorizer_transform([[1.0, None]])


# ------------------------------------------------------------------------------------------------ pivot
SET_TOP = [["a", "b"], ["a"], ["c"], ["C ", "A."]]          # OpSetVectorizerTest.scala:57-62
SET_BOT = [["x", "x"], ["z", "y", "z", "z", "y"], ["x", "y"], ["Z", "Z", "Z"]]


def test_opset_vectorizer_golden():
    # CT/stages/impl/feature/OpSetVectorizerTest.scala:57-76,104-118 -- vocab order A,C,B / X,Y,Z
    tops = V.pivot_fit([SET_TOP, SET_BOT], top_k=10, min_support=0)
    assert tops == [["A", "C", "B"], ["X", "Y", "Z"]]
    out = dense(V.pivot_transform([SET_TOP, SET_BOT], tops))
    exp = [sparse_row(10, [0, 2, 5], 1.0), sparse_row(10, [0, 6, 7], 1.0), sparse_row(10, [1, 5, 6], 1.0),
           sparse_row(10, [0, 1, 7], 1.0)]
    assert np.array_equal(out, np.stack(exp))
    meta = V.pivot_metadata(["top", "bot"], "MultiPickList", tops, True)
    assert [m.indicator for m in meta] == ["A", "C", "B", OTHER_STRING, NULL_STRING, "X", "Y", "Z", OTHER_STRING,
                                            NULL_STRING]
    assert all(m.grouping == ("top" if i < 5 else "bot") for i, m in enumerate(meta))


def test_opset_vectorizer_no_clean_golden():
    # OpSetVectorizerTest.scala:138-162 -- uppercase sorts before lowercase on ties
    tops = V.pivot_fit([SET_TOP, SET_BOT], top_k=10, min_support=0, clean_text=False)
    assert tops == [["a", "A.", "C ", "b", "c"], ["x", "y", "Z", "z"]]
    out = dense(V.pivot_transform([SET_TOP, SET_BOT], tops, clean_text=False))
    exp = [sparse_row(13, [0, 3, 7], 1.0), sparse_row(13, [0, 8, 10], 1.0), sparse_row(13, [4, 7, 8], 1.0),
           sparse_row(13, [1, 2, 9], 1.0)]
    assert np.array_equal(out, np.stack(exp))


def test_opset_vectorizer_topk_minsupport_golden():
    # OpSetVectorizerTest.scala:169-179 (topK = 1), expectedData :71-76
    tops = V.pivot_fit([SET_TOP, SET_BOT], top_k=1, min_support=0)
    out = dense(V.pivot_transform([SET_TOP, SET_BOT], tops))
    assert out.tolist() == [[1, 1, 0, 1, 0, 0], [1, 0, 0, 0, 2, 0], [0, 1, 0, 1, 1, 0], [1, 1, 0, 0, 1, 0]]
    # :181-195 (minSupport = 3)
    tops = V.pivot_fit([SET_TOP, SET_BOT], top_k=10, min_support=3)
    out = dense(V.pivot_transform([SET_TOP, SET_BOT], tops))
    assert out.tolist() == [[1, 1, 0, 1, 0], [1, 0, 0, 2, 0], [0, 1, 0, 2, 0], [1, 1, 0, 1, 0]]


def test_opset_vectorizer_unseen_golden():
    # OpSetVectorizerTest.scala:197-229 -- unseen values -> OTHER, empty -> null indicator
    top = [["a", "b"], ["a"], []]
    bot = [[], [], []]
    tops = V.pivot_fit([top, bot], top_k=10, min_support=0)
    assert tops == [["A", "B"], []]
    out = dense(V.pivot_transform([top, bot], tops))
    assert out.tolist() == [[1, 1, 0, 0, 0, 1], [1, 0, 0, 0, 0, 1], [0, 0, 0, 1, 0, 1]]
    out2 = dense(V.pivot_transform([SET_TOP, SET_BOT], tops))
    assert out2.tolist() == [[1, 1, 0, 0, 1, 0], [1, 0, 0, 0, 2, 0], [0, 0, 1, 0, 2, 0], [1, 0, 1, 0, 1, 0]]


TEXT1 = ["hello world", "hello world", "good evening", "hello world", None]   # SmartTextVectorizerTest.scala:55-63
TEXT2 = ["Hello world!", "What's up", "How are you doing, my friend?", "Not bad, my friend.", None]


def test_text_pivot_vectorizer_golden():
    # CT/stages/impl/feature/OpTextPivotVectorizerTest.scala:45-73
    tops = V.pivot_fit([TEXT1, TEXT2], top_k=2, min_support=1)
    out = dense(V.pivot_transform([TEXT1, TEXT2], tops))
    exp = [sparse_row(8, [0, 4], 1.0), sparse_row(8, [0, 6], 1.0), sparse_row(8, [1, 5], 1.0),
           sparse_row(8, [0, 6], 1.0), sparse_row(8, [3, 7], 1.0)]
    assert np.array_equal(out, np.stack(exp))


def test_smart_text_vectorizer_golden():
    # CT/stages/impl/feature/SmartTextVectorizerTest.scala:55-75
    methods, tops = V.smart_text_fit([TEXT1, TEXT2], max_cardinality=2, top_k=2, min_support=1)
    assert methods == [V.PIVOT, V.HASH]
    out = dense(V.smart_text_transform([TEXT1, TEXT2], ["text1", "text2"], methods, tops, num_features=4,
                                       prepend_feature_name=False, hash_space_strategy="shared"))
    exp = [sparse_row(9, [0, 4, 6], [1.0, 1.0, 1.0]), sparse_row(9, [0, 8], [1.0, 1.0]),
           sparse_row(9, [1, 6], [1.0, 1.0]), sparse_row(9, [0, 6], [1.0, 2.0]), sparse_row(9, [3, 8], [1.0, 1.0])]
    assert np.array_equal(out, np.stack(exp))
    meta = V.smart_text_metadata(["text1", "text2"], "Text", methods, tops, num_features=4,
                                 hash_space_strategy="shared")
    assert len(meta) == 9
    assert [m.indicator for m in meta[:4]] == ["HelloWorld", "GoodEvening", OTHER_STRING, NULL_STRING]
    assert all(m.grouping is None and m.indicator is None for m in meta[4:8])
    assert meta[8].indicator == NULL_STRING and meta[8].grouping == "text2"


def test_smart_text_equals_pivot_plus_hash_composition():
    # SmartTextVectorizerTest.scala:110-186: smart == pivot (+) hash (+) null, per input
    methods, tops = V.smart_text_fit([TEXT1, TEXT2], max_cardinality=10, top_k=2, min_support=1)
    assert methods == [V.PIVOT, V.PIVOT]
    a = dense(V.smart_text_transform([TEXT1, TEXT2], ["text1", "text2"], methods, tops, num_features=4))
    b = dense(V.pivot_transform([TEXT1, TEXT2], tops))
    assert np.array_equal(a, b)
    methods, tops = V.smart_text_fit([TEXT1, TEXT2], max_cardinality=0, top_k=2, min_support=1)
    assert methods == [V.HASH, V.HASH]


def test_shared_hash_space_rule():
    # OPCollectionHashingVectorizer.scala:191-199 ; C4 (2^18 x 20) shared, C5 (512 x 30) separate
    assert V.is_shared_hash_space(1 << 18, 20)
    assert not V.is_shared_hash_space(512, 30)
    assert not V.is_shared_hash_space(512, 256) and V.is_shared_hash_space(512, 257)


def test_combiner_golden():
    # CT/stages/impl/feature/VectorsCombinerTest.scala:64-70 -- dense + sparse concat with running offset
    import scipy.sparse as sp
    a = np.array([[1.0, 0.0, 3.0]])
    b = sp.csr_matrix(([2.0], ([0], [1])), shape=(1, 4))
    out, meta = V.combine([a, b], [[ColumnMeta(("a",), ("Real",))] * 3, [ColumnMeta(("b",), ("Real",))] * 4])
    assert dense(out).tolist() == [[1.0, 0.0, 3.0, 0.0, 2.0, 0.0, 0.0]]
    assert [m.index for m in meta] == list(range(7))


# ------------------------------------------------------------------------------------------------ statistics
def colmajor(r, c, vals):
    return np.asarray(vals, dtype=np.float64).reshape(c, r).T


def cramers_v(m):
    return S.chi_squared_on_filtered(S.filter_empties(m))[0]


def test_cramers_v_golden():
    # UT/stats/OpStatisticsTest.scala:47-105 (tolerance 1e-3 there; the exact values were cross-checked with scipy)
    assert abs(cramers_v(colmajor(2, 2, [757, 726, 731, 2621])) - 0.2921) < 1e-3
    assert abs(cramers_v(colmajor(4, 4, [5, 15, 20, 68, 29, 54, 84, 119, 14, 14, 17, 26, 16, 10, 94, 7])) - 0.279) \
        < 1e-3
    assert abs(cramers_v(colmajor(3, 6, [192, 221, 229, 185, 202, 194, 62, 199, 97, 78, 30, 44, 78, 29, 53, 21, 18,
                                         18])) - 0.1815) < 1e-3
    assert math.isnan(cramers_v(colmajor(5, 2, [0, 0, 0, 0, 0, 5, 6, 1, 2, 7])))
    assert abs(cramers_v(colmajor(6, 2, [0, 757, 0, 0, 726, 0, 0, 731, 0, 0, 2621, 0])) - 0.2921) < 1e-3
    assert abs(cramers_v(colmajor(4, 4, [0, 0, 0, 0, 0, 757, 726, 0, 0, 0, 0, 0, 0, 731, 2621, 0])) - 0.2921) < 1e-3
    assert math.isnan(cramers_v(np.zeros((0, 0))))
    r1 = cramers_v(colmajor(3, 2, [100, 0, 50, 0, 100, 50]))
    r2 = cramers_v(colmajor(4, 2, [100, 0, 50, 50, 0, 100, 50, 50]))
    assert r1 > r2
    from scipy.stats import chi2_contingency
    m = colmajor(2, 2, [757, 726, 731, 2621])
    chi2 = chi2_contingency(m, correction=False)[0]
    assert abs(cramers_v(m) - math.sqrt(chi2 / m.sum())) < 1e-14


def test_pmi_and_confidence_golden():
    # OpStatisticsTest.scala:107-145
    pmi, mi = S.mutual_info(S.filter_empties(colmajor(2, 2, [10, 15, 70, 5])))
    assert np.allclose(pmi["0"], [-1.0, 1.5850], atol=1e-3) and np.allclose(pmi["1"], [0.2224, -1.5850], atol=1e-3)
    assert abs(mi - 0.2142) < 1e-3
    pmi, mi = S.mutual_info(S.filter_empties(colmajor(5, 2, [0, 0, 0, 0, 0, 5, 6, 1, 2, 7])))
    assert all(v == 0 for vs in pmi.values() for v in vs) and mi == 0
    # :147-154
    st = S.contingency_stats(np.zeros((0, 0)))
    assert math.isnan(st.cramers_v) and st.pmi == {} and math.isnan(st.mutual_info)
    # :156-189
    conf, sup = S.max_confidences(colmajor(2, 2, [0.0, 500.0, 121.0, 688.0]))
    assert np.allclose(conf, [1.0, 0.5791], atol=1e-3) and np.allclose(sup, [0.0924, 0.9076], atol=1e-3)
    conf, sup = S.max_confidences(colmajor(6, 4, [132.0, 0.0, 189.0, 688.0, 321.0, 0.0, 98.0, 0.0, 823.0, 223.0, 0.0,
                                                  366.0, 123.0, 14.0, 0.0, 119.0, 0.0, 482.0, 0.0, 0.0, 443.0, 18.0,
                                                  321.0, 0.0]))
    assert np.allclose(conf, [0.3739, 1.0, 0.5656, 0.6564, 0.5, 0.5684], atol=1e-3)
    assert np.allclose(sup, [0.0810, 0.0032, 0.3337, 0.2404, 0.1472, 0.1945], atol=1e-3)


def test_label_corr_matches_full_matrix():
    # UT/stats/OpStatisticsTest.scala:191-206 -- both routes agree to 1e-12 relative on N(0,1) data
    rng = np.random.default_rng(7)
    x = rng.standard_normal((1000, 10))
    st = S.col_stats(x)
    a = S.correlations_with_label(x, st)
    b = S.pearson_matrix(x)[:, -1]
    assert np.allclose(a, b, rtol=1e-12, atol=0)


SC_ROWS = [  # CT/stages/impl/preparators/SanityCheckerTest.scala:100-116 : (age, height, height_null, isBlueEyed,
    (32, 5.0, 0, 1, 0.5, 0),   # gender, testFeatNegCor)
    (32, 4.0, 1, 0, 0, 0.1),
    (32, 6.0, 1, 1, 0.5, 0),
    (32, 5.5, 0, 1, 0.5, 0),
    (32, 5.4, 1, 0, 0, 0.1),
    (32, 5.4, 1, 0, 0, 0.1),
]
SC_NAMES = ["age", "height", "height_null", "gender", "testFeatNegCor"]


def sc_fixture():
    labels = [r[3] for r in SC_ROWS]
    x = np.array([[r[0], r[1], r[2], r[4], r[5]] for r in SC_ROWS], dtype=np.float64)
    metas = [ColumnMeta((n,), ("Real",), index=i) for i, n in enumerate(SC_NAMES)]
    return labels, x, metas


def test_sanity_checker_drop_set_golden():
    # SanityCheckerTest.scala:133-164: maxCorrelation 0.99, minVariance 0 -> dropped age_0, gender_3, testFeatNegCor_4
    labels, x, metas = sc_fixture()
    p = S.SanityParams(max_correlation=0.99, min_variance=0.0, check_sample=1.0, remove_bad_features=True,
                       sample_lower_limit=0)
    res = S.sanity_fit(labels, x, metas, p)
    assert res.dropped == ["age_0", "gender_3", "testFeatNegCor_4"]
    assert res.indices_to_keep == [1, 2]
    by = {s.name: s for s in res.stats}
    assert math.isnan(by["age_0"].corr_label)            # expectedCorrFeatNamesIsNan :131
    assert all(not math.isnan(by[n].corr_label) for n in ["height_1", "height_null_2", "gender_3",
                                                           "testFeatNegCor_4"])
    # :659-660 label summary: max 1, mean 0.5
    assert by["label"].max == 1.0 and by["label"].mean == 0.5
    out = S.sanity_transform(x, res.indices_to_keep, True)
    assert out.tolist() == x[:, [1, 2]].tolist()


def test_sanity_checker_feature_feature_rule():
    # SanityCheckerTest.scala:473-500: the LATER of two mutually correlated features is dropped
    labels, x, metas = sc_fixture()
    p = S.SanityParams(max_correlation=1.1, min_variance=-0.1, max_feature_corr=0.9, remove_bad_features=True,
                       sample_lower_limit=0)
    res = S.sanity_fit(labels, x, metas, p)
    assert "testFeatNegCor_4" in res.dropped and "gender_3" not in res.dropped
    p.feature_feature_corr_level = "off"
    res = S.sanity_fit(labels, x, metas, p)
    assert res.dropped == []


def test_sanity_checker_all_dropped_error():
    # SanityCheckerTest.scala:349-357
    labels, x, metas = sc_fixture()
    p = S.SanityParams(min_variance=1e9, remove_bad_features=True, sample_lower_limit=0)
    with pytest.raises(ValueError, match="dropped all of your features"):
        S.sanity_fit(labels, x, metas, p)


def test_fraction():
    # SanityChecker.scala:356-361 with the defaults :561-564
    p = S.SanityParams()
    assert S.fraction(p, 500) == 1.0 and S.fraction(p, 10_000_000) == 0.1
    assert S.fraction(S.SanityParams(check_sample=0.5), 100_000) == 0.5


def test_integral_vectorizer_golden():
    # IntegralVectorizerTest.scala:48-56 (data), :93-118 (constant fill 3), :120-149 (mode: 4, 2, 1 -- the tie between
    # 1 and 2 in inC goes to the smaller value -- and 0 for the all-null inD)
    a, b, c, d = [4, 4, 2, None, None], [2, None, 4, 2, None], [2, 1, 1, 2, None], [None] * 5
    assert V.mode_seq_null_int([a, b, c, d]) == [4, 2, 1, 0]
    const = V.integral_vectorizer_fit([a, b, c, d], fill_with_mode=False, fill_value=3)
    got = V.integral_vectorizer_transform([a, b, c, d], const, track_nulls=False)
    assert got.tolist() == [[4.0, 2.0, 2.0, 3.0], [4.0, 3.0, 1.0, 3.0], [2.0, 4.0, 1.0, 3.0], [3.0, 2.0, 2.0, 3.0],
                            [3.0, 3.0, 3.0, 3.0]]
    mode = V.integral_vectorizer_fit([a, b, c, d], fill_with_mode=True)
    got = V.integral_vectorizer_transform([a, b, c, d], mode, track_nulls=False)
    assert got.tolist() == [[4.0, 2.0, 2.0, 0.0], [4.0, 2.0, 1.0, 0.0], [2.0, 4.0, 1.0, 0.0], [4.0, 2.0, 2.0, 0.0],
                            [4.0, 2.0, 1.0, 0.0]]
    # with null tracking: [value, isNull] interleaved (IntegralVectorizerTest.scala:151-180)
    got = V.integral_vectorizer_transform([a, b], [4, 2], track_nulls=True)
    assert got[3].tolist() == [4.0, 1.0, 2.0, 0.0] and got[4].tolist() == [4.0, 1.0, 2.0, 1.0]


@pytest.mark.parametrize("tokens,min_std", [(True, 0.5), (False, 1.0)])
def test_smart_text_ignore_route_layout(tokens, min_std):
    """SmartTextVectorizerTest.scala:213-280 ("detect and ignore fields that looks like machine-generated IDs"), with
    data of the same four shapes (:85-97): the country column is hashed, the small-domain column is pivoted although
    its lengths are fixed (the categorical check comes first), the fixed-length id column is ignored, the free text is
    hashed; vector = [pivot 3+2 | 2 x 5 hash buckets | 3 lengths | 3 null indicators] = 21 columns."""
    rng = np.random.default_rng(11)
    n = 1000
    opt = lambda v: None if rng.random() < 0.2 else v   # noqa: E731
    places = [" ".join("".join(rng.choice(list("abcdefghijklmnop"), size=rng.integers(3, 9))).capitalize()
                       for _ in range(rng.integers(1, 4))) for _ in range(250)]
    country = [opt(places[int(rng.integers(len(places)))]) for _ in range(n)]
    categorical = [opt("ABCDEF"[int(rng.integers(6))]) for _ in range(n)]
    text_id = [opt("%06d" % (40231 + int(rng.integers(1000)))) for _ in range(n)]
    text = [opt("".join(rng.choice(list("abc xyz,. QRS"), size=rng.integers(0, 101)))) for _ in range(n)]
    cols, names = [country, categorical, text_id, text], ["country", "categorical", "textId", "text"]
    methods, tops = V.smart_text_fit(cols, max_cardinality=10, top_k=3, min_support=10, min_length_std_dev=min_std,
                                     tokenize_for_lengths=tokens)
    assert methods == [V.HASH, V.PIVOT, V.IGNORE, V.HASH]
    x = V.smart_text_transform(cols, names, methods, tops, num_features=5, track_nulls=True, track_text_len=True,
                               to_lowercase=False, min_token_length=1)
    meta = V.smart_text_metadata(names, "Text", methods, tops, num_features=5, track_nulls=True, track_text_len=True)
    assert x.shape == (n, 21) and len(meta) == 21
    assert all(m.grouping == "categorical" for m in meta[0:5])
    assert all(m.parent_names == ("country",) for m in meta[5:10])
    assert all(m.parent_names == ("text",) for m in meta[10:15])
    assert [m.parent_names[0] for m in meta[15:18]] == ["country", "text", "textId"]        # hashed, then ignored
    assert all(m.descriptor == TEXT_LEN_STRING for m in meta[15:18])
    assert all(m.indicator == NULL_STRING for m in meta[18:21])
    dense = np.asarray(x.todense())
    for r in (0, 1, 2, 500, 999):
        toks = [T.tokenize(c[r], False, 1) for c in (country, text, text_id)]
        assert dense[r, 15:18].tolist() == [float(sum(len(t) for t in tk)) for tk in toks]
        assert dense[r, 18:21].tolist() == [0.0 if tk else 1.0 for tk in toks]
    # the ignored column contributes nothing but its length and its null indicator
    assert np.count_nonzero(dense[:, 5:15].sum(axis=1) == 0) <= sum(1 for a, b in zip(country, text)
                                                                      if not T.tokenize(a, False, 1)
                                                                      and not T.tokenize(b, False, 1))


def _country_column(seed=23):
    """The shape of SmartTextVectorizerTest.scala:107-109 (`categoricalCountryData`): 1000 x "United States" shuffled
    with 1000 draws (20 % empty) from a list of ~450 country names."""
    rng = np.random.default_rng(seed)
    names = sorted({" ".join("".join(rng.choice(list("abcdefghijklmnoprstuvw"), size=rng.integers(3, 9))).capitalize()
                             for _ in range(rng.integers(1, 3))) for _ in range(470)})[:450]
    draws = [None if rng.random() < 0.2 else names[int(rng.integers(len(names)))] for _ in range(1000)]
    col = ["United States"] * 1000 + draws
    return [col[i] for i in rng.permutation(len(col))]


@pytest.mark.parametrize("case,kw,num_features,method,width", [
    # SmartTextVectorizerTest.scala:283-290: coverage near 0 -> pivot, TopK + OTHER + null
    ("coverage~0", dict(coverage_pct=1e-10, max_cardinality=100, min_support=1), 512, "Pivot", 20 + 2),
    # :291-298: coverage near 1 -> hash, DefaultNumOfFeatures + length + null
    ("coverage~1", dict(coverage_pct=1.0 - 1e-10, max_cardinality=100, min_support=1), 512, "Hash", 512 + 2),
    # :300-312: high cardinality, pivoted through the coverage
    ("coverage", dict(max_cardinality=100, top_k=10, min_support=1, coverage_pct=0.5, clean_text=False), 512,
     "Pivot", 10 + 2),
    # :314-328: a minimum support nothing reaches -> coverage 0 -> hash
    ("minSupport 99999", dict(max_cardinality=100, top_k=10, min_support=99999, coverage_pct=0.5, clean_text=False),
     5, "Hash", 5 + 2),
    # :330-344: minimum support 100 leaves the one frequent value, which alone covers half the rows -> pivot of width 3
    ("minSupport 100", dict(max_cardinality=100, top_k=10, min_support=100, coverage_pct=0.5, clean_text=False), 5,
     "Pivot", 3),
    # :346-359: cardinality <= topK -> the coverage rule does not apply -> hash
    ("topK too high", dict(max_cardinality=100, top_k=1000000, coverage_pct=0.5, clean_text=False), 5, "Hash", 5 + 2),
])
def test_smart_text_coverage_decisions(case, kw, num_features, method, width):
    col = _country_column()
    assert len({v for v in col if v is not None}) > 100
    methods, tops = V.smart_text_fit([col], **kw)
    assert methods == [method], case
    meta = V.smart_text_metadata(["rawCatCountry"], "Text", methods, tops, num_features=num_features, track_nulls=True,
                                 track_text_len=True)
    x = V.smart_text_transform([col[:50]], ["rawCatCountry"], methods, tops, num_features=num_features,
                               clean_text=kw.get("clean_text", True), track_nulls=True, track_text_len=True)
    assert len(meta) == width and x.shape == (50, width), case


def test_smart_text_uniform_countries_stay_text_despite_the_coverage():
    # SmartTextVectorizerTest.scala:361-373: no dominant value -> top-10 of the capped map covers less than half
    col = [v for v in _country_column(4) if v != "United States"]
    assert len({v for v in col if v is not None}) > 100
    methods, tops = V.smart_text_fit([col], max_cardinality=100, top_k=10, min_support=1, coverage_pct=0.5,
                                     clean_text=False)
    assert methods == ["Hash"]
    meta = V.smart_text_metadata(["country"], "Text", methods, tops, num_features=5, track_nulls=True,
                                 track_text_len=True)
    assert len(meta) == 5 + 2
