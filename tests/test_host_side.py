"""Host-side pieces that need no GPU: row sampling of a columnar batch, the dictionary digest, and the two-step
dictionary finish of the pivot vectorizer against the one-call form and the oracle."""
import ctypes as C

import numpy as np
import pytest

from oracle import text as T
from oracle import vectorizers as V

torch = pytest.importorskip("torch")


def test_take_rows_matches_a_batch_built_from_the_same_rows():
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch
    rng = np.random.default_rng(0)
    n = 300
    opt = lambda v: None if rng.random() < 0.2 else v   # noqa: E731
    cols = {"r": (F.REAL, [opt(float(i)) for i in range(n)]), "b": (F.BINARY, [opt(bool(i % 2)) for i in range(n)]),
            "i": (F.INTEGRAL, [opt(int(i * 7)) for i in range(n)]), "p": (F.PICKLIST, [opt("abc"[i % 3]) for i in range(n)]),
            "t": (F.TEXT, [opt("w%d x" % i) for i in range(n)]),
            "m": (F.MULTIPICKLIST, [opt(["a", "b", "c"][:i % 4]) for i in range(n)])}
    batch = ColumnarBatch.from_columns(cols, pin=False)
    idx = torch.tensor(sorted(rng.choice(n, 130, replace=False)), dtype=torch.int64)
    sub = batch.take_rows(idx)
    exp = ColumnarBatch.from_columns({k: (tp, [v[j] for j in idx.tolist()]) for k, (tp, v) in cols.items()}, pin=False)
    assert sub.n == 130
    for name in ("r", "b", "i"):
        a, e = sub.numeric([name]), exp.numeric([name])
        assert torch.equal(a.valid, e.valid)
        ok = torch.tensor([cols[name][1][j] is not None for j in idx.tolist()])
        assert torch.equal(a.values[0, :130][ok], e.values[0, :130][ok])
    a, e = sub.texts(["t"]), exp.texts(["t"])
    assert torch.equal(a.offsets[0], e.offsets[0]) and torch.equal(a.valid[0], e.valid[0])
    assert torch.equal(a.data[0][:int(a.offsets[0][-1])], e.data[0][:int(e.offsets[0][-1])])
    a, e = sub.sets(["m"]), exp.sets(["m"])
    assert torch.equal(a.offsets[0], e.offsets[0])
    dec = lambda blk: [None if c < 0 else blk.dicts[0][c] for c in blk.codes[0, :130].tolist()]   # noqa: E731
    assert dec(sub.strings(["p"])) == dec(exp.strings(["p"]))


def test_dictionary_digest_sees_every_byte():
    from transmogrifai_b200.columnar import encode_strings
    from transmogrifai_b200.stages.pivot import _dict_signature
    data, offs = encode_strings([f"value_{i}" for i in range(5000)])
    sig = _dict_signature(data, offs)
    assert sig == _dict_signature(data.copy(), offs.copy())
    for k in (0, 7, 12345, len(data) - 1):
        d2 = data.copy()
        d2[k] ^= 1
        assert _dict_signature(d2, offs) != sig
    o2 = offs.copy()
    o2[77] += 1
    assert _dict_signature(data, o2) != sig


def test_pivot_dictionary_prepare_and_finish_match_the_one_call_form_and_the_oracle():
    from transmogrifai_b200 import _lib as L
    from transmogrifai_b200.columnar import decode_strings, encode_strings
    rng = np.random.default_rng(3)
    raw = ["red", "Red ", "RED!", "green", "Green.", "blue", "teal", "b-lue", "x", "X", "y_y", "yY"] + \
          [f"v{i}" for i in range(200)]
    counts = rng.integers(0, 40, size=(len(raw) + 1, 2)).astype(np.int64)
    counts[1 + 200:] = 0                                   # some entries never occur
    data, offs = encode_strings(raw)
    flat = np.ascontiguousarray(counts.reshape(-1))
    k, min_support = 7, 10

    def outputs():
        return (C.c_int(0), np.zeros(len(raw), dtype=np.int32), np.zeros((k + 2) * 2, dtype=np.int64),
                np.zeros(len(data), dtype=np.uint8), np.zeros(k + 1, dtype=np.int64))

    n1, c1, p1, t1, o1 = outputs()
    L.call("tmog_pivot_column_finish", data.ctypes.data, offs.ctypes.data, len(raw), 1, flat.ctypes.data, 2, k,
           min_support, C.byref(n1), c1.ctypes.data, p1.ctypes.data, t1.ctypes.data, len(t1), o1.ctypes.data)
    h = C.c_void_p()
    L.call("tmog_pivot_dict_prepare", data.ctypes.data, offs.ctypes.data, len(raw), 1, C.byref(h))
    n2, c2, p2, t2, o2 = outputs()
    L.call("tmog_pivot_dict_finish", h, flat.ctypes.data, 2, k, min_support, C.byref(n2), c2.ctypes.data,
           p2.ctypes.data, t2.ctypes.data, len(t2), o2.ctypes.data)
    L.call("tmog_pivot_dict_free", h)
    assert n1.value == n2.value and np.array_equal(c1, c2) and np.array_equal(p1, p2) and np.array_equal(o1, o2)
    top = decode_strings(t2, o2[:n2.value + 1])
    # oracle: counts per cleaned value, count >= minSupport, (count desc, value asc), topK
    merged = {}
    for s, c in zip(raw, counts[1:].sum(axis=1)):
        if c > 0:
            merged[T.clean_string(s)] = merged.get(T.clean_string(s), 0) + int(c)
    assert top == V.top_values(merged, k, min_support)
    assert p2.reshape(k + 2, 2)[n2.value + 1].tolist() == counts[0].tolist()          # the null row
    assert p2.sum() == counts.sum()


def test_large_dictionary_grouping_on_threads_matches_the_oracle():
    """Dictionaries of 32768 entries and more are grouped by several host threads (strings_host.cu group_strings_mt):
    same vocabulary, same code -> column table and the same per-class counts as a plain fold over the entries."""
    from transmogrifai_b200 import _lib as L
    from transmogrifai_b200.columnar import decode_strings, encode_strings
    rng = np.random.default_rng(17)
    nbase = 60000
    raw = [f"v7_{i}" for i in range(nbase)]
    raw += [f"V7 {i}!" for i in rng.integers(0, nbase, 9000)]            # clean to the value of an earlier entry
    raw += [f"v7.{i}" for i in rng.integers(0, nbase, 9000)]
    raw += ["", " ", "__"]                                               # clean to the empty string
    order = rng.permutation(len(raw))
    raw = [raw[i] for i in order]
    counts = rng.integers(0, 6, size=(len(raw) + 1, 2)).astype(np.int64)
    hot = rng.integers(0, len(raw), 40)
    counts[1 + hot] += rng.integers(50, 500, size=(40, 2))
    data, offs = encode_strings(raw)
    flat = np.ascontiguousarray(counts.reshape(-1))
    k, min_support = 25, 10
    h = C.c_void_p()
    L.call("tmog_pivot_dict_prepare", data.ctypes.data, offs.ctypes.data, len(raw), 1, C.byref(h))
    n_top, code_to_out = C.c_int(0), np.zeros(len(raw), dtype=np.int32)
    per_class = np.zeros((k + 2) * 2, dtype=np.int64)
    top_data, top_offs = np.zeros(len(data) + 1, dtype=np.uint8), np.zeros(k + 1, dtype=np.int64)
    L.call("tmog_pivot_dict_finish", h, flat.ctypes.data, 2, k, min_support, C.byref(n_top), code_to_out.ctypes.data,
           per_class.ctypes.data, top_data.ctypes.data, len(top_data), top_offs.ctypes.data)
    L.call("tmog_pivot_dict_free", h)
    top = decode_strings(top_data, top_offs[:n_top.value + 1])
    cleaned = [T.clean_string(s) for s in raw]
    merged, per_value = {}, {}
    for s, c in zip(cleaned, counts[1:]):
        if c.sum() > 0:
            merged[s] = merged.get(s, 0) + int(c.sum())
        per_value[s] = per_value.get(s, np.zeros(2, dtype=np.int64)) + c
    exp_top = V.top_values(merged, k, min_support)
    assert top == exp_top and n_top.value == k
    col_of = {v: i for i, v in enumerate(exp_top)}
    assert code_to_out.tolist() == [col_of.get(s, k) for s in cleaned]                  # k = the OTHER column
    rows = per_class.reshape(k + 2, 2)
    for i, v in enumerate(exp_top):
        assert rows[i].tolist() == per_value[v].tolist()
    assert rows[k + 1].tolist() == counts[0].tolist() and rows.sum() == counts.sum()
