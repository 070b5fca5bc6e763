"""The host-buffer entry points of the string paths, driven the way a JNI shim would drive them: plain ctypes, numpy
host buffers, a tmog_ctx -- no torch anywhere in this module (include/tmog_b200.h).  Results
against the oracle."""
import ctypes as C

import numpy as np
import pytest

from oracle import text as T
from oracle import vectorizers as V

pytestmark = pytest.mark.gpu


def _lib():
    from transmogrifai_b200 import _lib as L
    return L


def _ctx(L):
    h = C.c_void_p()
    L.call("tmog_ctx_create", 0, C.byref(h))
    return h


def _arrow(strings):
    enc = [s.encode("utf-8") for s in strings]
    offs = np.zeros(len(enc) + 1, dtype=np.int64)
    offs[1:] = np.cumsum([len(e) for e in enc])
    data = np.frombuffer(b"".join(enc) + b"\0", dtype=np.uint8).copy()
    return data, offs


def _ptrs(arrays):
    return (C.c_void_p * len(arrays))(*[a.ctypes.data if a is not None else None for a in arrays])


def test_pivot_fit_and_transform_over_host_buffers_without_torch():
    L = _lib()
    rng = np.random.default_rng(3)
    n, top_k, min_support = 20_000, 4, 10
    raw = [["x", "y", "z", "X ", "w", "rare"], [f"v{i}" for i in range(300)] + ["V 0", "v.1!"]]
    probs = [np.array([0.3, 0.25, 0.2, 0.1, 0.1499, 0.0001]), None]
    cols, codes = [], np.full((2, n), -1, dtype=np.int32)
    for j, d in enumerate(raw):
        p = probs[j] if probs[j] is not None else (1.0 / np.arange(1, len(d) + 1)) / (1.0 / np.arange(1, len(d) + 1)).sum()
        c = rng.choice(len(d), size=n, p=p)
        null = rng.random(n) < 0.07
        codes[j] = np.where(null, -1, c)
        cols.append([None if null[i] else d[c[i]] for i in range(n)])
    y = (rng.random(n) < 0.3).astype(np.float64)
    dicts = [_arrow(d) for d in raw]
    ncodes = np.array([len(d) for d in raw], dtype=np.int64)
    ctx = _ctx(L)
    try:
        ncls = C.c_int(0)
        classes, class_counts = np.zeros(100), np.zeros(100, dtype=np.uint64)
        n_top = np.zeros(2, dtype=np.int32)
        cap = 2 * 4096
        top_data, top_offs = np.zeros(cap, dtype=np.uint8), np.zeros(2 * (top_k + 1), dtype=np.int64)
        code_to_out = np.zeros(int(ncodes.sum()), dtype=np.int32)
        per_class = np.zeros(2 * (top_k + 2) * 2, dtype=np.int64)
        L.call("tmog_pivot_fit_host", ctx, codes.ctypes.data, n, n, 2, _ptrs([d[0] for d in dicts]),
               _ptrs([d[1] for d in dicts]), ncodes.ctypes.data, y.ctypes.data, 4096, 1, top_k, min_support, 100,
               C.byref(ncls), classes.ctypes.data, class_counts.ctypes.data, n_top.ctypes.data, top_data.ctypes.data, cap,
               top_offs.ctypes.data, code_to_out.ctypes.data, per_class.ctypes.data)
        assert ncls.value == 2 and classes[:2].tolist() == [0.0, 1.0]
        assert class_counts[:2].tolist() == [int((y == 0).sum()), int((y == 1).sum())]
        exp_tops = V.pivot_fit(cols, top_k, min_support, True)
        got_tops = []
        for j in range(2):
            o = top_offs[j * (top_k + 1):(j + 1) * (top_k + 1)]
            seg = top_data[j * (cap // 2):]
            got_tops.append([bytes(seg[o[i]:o[i + 1]]).decode() for i in range(n_top[j])])
        assert got_tops == exp_tops
        # per-class counts of every emitted column against the oracle's one-hot matrix
        x = np.asarray(V.pivot_transform(cols, exp_tops, True, True).todense())
        off = 0
        for j in range(2):
            w = len(exp_tops[j]) + 2
            got = per_class[j * (top_k + 2) * 2:(j * (top_k + 2) + w) * 2].reshape(w, 2)
            exp = np.stack([x[y == 0, off:off + w].sum(axis=0), x[y == 1, off:off + w].sum(axis=0)], axis=1)
            assert np.array_equal(got, exp.astype(np.int64))
            off += w
        # transform: the position of every row's 1.0, all columns kept and with a dropped column
        out = np.full((n, 2), -7, dtype=np.int32)
        L.call("tmog_pivot_transform_host", ctx, codes.ctypes.data, n, n, 2, ncodes.ctypes.data, code_to_out.ctypes.data,
               n_top.ctypes.data, 1, None, 4096, out.ctypes.data, 2, 0)
        assert np.array_equal(out, np.argwhere(x == 1.0)[:, 1].reshape(n, 2))
        width = x.shape[1]
        newpos = np.arange(width, dtype=np.int32)
        newpos[1] = -1
        newpos[2:] -= 1
        L.call("tmog_pivot_transform_host", ctx, codes.ctypes.data, n, n, 2, ncodes.ctypes.data, code_to_out.ctypes.data,
               n_top.ctypes.data, 1, newpos.ctypes.data, 0, out.ctypes.data, 2, 0)
        assert np.array_equal(out, newpos[np.argwhere(x == 1.0)[:, 1]].reshape(n, 2))
    finally:
        L.call("tmog_ctx_destroy", ctx)


def test_text_hash_statistics_and_rows_over_host_buffers_without_torch():
    L = _lib()
    rng = np.random.default_rng(5)
    n, nf = 3000, 512
    vocab = ["".join(rng.choice(list("abcdefghijklmnopqrstuvwxyz"), size=rng.integers(2, 9))) for _ in range(300)]
    vocab += ["the", "and", "don't", "u.s.a", "x_1", "42"]
    cols = []
    for _ in range(3):
        col = []
        for _ in range(n):
            u = rng.random()
            col.append(None if u < 0.06 else "" if u < 0.08 else
                       " ".join(vocab[i] for i in rng.zipf(1.3, size=rng.poisson(10)) % len(vocab)).title())
        cols.append(col)
    names = ["alpha", "beta", "gamma"]
    y = (rng.random(n) < 0.4).astype(np.float64)
    enc = [_arrow(["" if v is None else v for v in c]) for c in cols]
    valid = []
    for c in cols:
        bits = np.zeros((n + 63) // 64 * 64, dtype=np.uint8)
        bits[:n] = [v is not None for v in c]
        valid.append(np.packbits(bits, bitorder="little").view(np.uint64).copy())
    prefix = np.zeros((3, 16), dtype=np.uint8)
    plen = np.zeros(3, dtype=np.int32)
    for j, nm in enumerate(names):
        b = f"{T.hashing_tf_index(nm, nf)}_".encode()
        prefix[j, :len(b)] = np.frombuffer(b, dtype=np.uint8)
        plen[j] = len(b)
    params = np.array([nf, 0, 0, 1, 1, 1, 3 * nf, 0, 0, 0, 0, 0], dtype=np.int32)
    w = 3 * nf + 3
    exp = np.asarray(V.smart_text_transform(cols, names, ["Hash"] * 3, [[], [], []], num_features=nf).todense())
    ctx = _ctx(L)
    try:
        ncls = C.c_int(0)
        classes, class_counts = np.zeros(100), np.zeros(100, dtype=np.uint64)
        cs, rows, sq = np.zeros(w * 2, dtype=np.int64), np.zeros(w, dtype=np.int64), np.zeros(w, dtype=np.int64)
        mx, mn = np.zeros(w, dtype=np.int32), np.zeros(w, dtype=np.int32)
        L.call("tmog_text_hash_stats_host", ctx, 3, n, _ptrs([e[0] for e in enc]), _ptrs([e[1] for e in enc]),
               _ptrs(valid), prefix.ctypes.data, plen.ctypes.data, params.ctypes.data, y.ctypes.data, 100, C.byref(ncls),
               classes.ctypes.data, class_counts.ctypes.data, cs.ctypes.data, rows.ctypes.data, sq.ctypes.data,
               mx.ctypes.data, mn.ctypes.data)
        assert ncls.value == 2
        assert np.array_equal(cs.reshape(w, 2), np.stack([exp[y == 0].sum(axis=0), exp[y == 1].sum(axis=0)], axis=1))
        assert np.array_equal(rows, (exp > 0).sum(axis=0)) and np.array_equal(sq, (exp * exp).sum(axis=0))
        assert np.array_equal(mx, exp.max(axis=0))
        held = np.where(exp > 0, exp, np.inf).min(axis=0)
        assert np.array_equal(mn, np.where(np.isfinite(held), held, 0))
        # rows, with one bucket and one null indicator dropped
        newpos = np.arange(w, dtype=np.int32)
        drop = [int(np.argmax(exp.sum(axis=0))), w - 2]
        keep = [i for i in range(w) if i not in drop]
        newpos[:] = -1
        newpos[keep] = np.arange(len(keep), dtype=np.int32)
        nnz = C.c_int64(0)
        L.call("tmog_text_hash_transform_host", ctx, 3, n, _ptrs([e[0] for e in enc]), _ptrs([e[1] for e in enc]),
               _ptrs(valid), prefix.ctypes.data, plen.ctypes.data, params.ctypes.data, newpos.ctypes.data, None, None,
               None, 0, C.byref(nnz))
        ref = exp[:, keep]
        assert nnz.value == int((ref != 0).sum())
        indptr = np.zeros(n + 1, dtype=np.int64)
        idx, val = np.zeros(nnz.value, dtype=np.int32), np.zeros(nnz.value, dtype=np.int32)
        L.call("tmog_text_hash_transform_host", ctx, 3, n, _ptrs([e[0] for e in enc]), _ptrs([e[1] for e in enc]),
               _ptrs(valid), prefix.ctypes.data, plen.ctypes.data, params.ctypes.data, newpos.ctypes.data,
               indptr.ctypes.data, idx.ctypes.data, val.ctypes.data, nnz.value, C.byref(nnz))
        import scipy.sparse as sp
        got = sp.csr_matrix((val.astype(np.float64), idx, indptr), shape=(n, len(keep)))
        assert abs(got - sp.csr_matrix(ref)).max() == 0
        assert all(np.all(np.diff(idx[indptr[r]:indptr[r + 1]]) > 0) for r in range(0, n, 97))
    finally:
        L.call("tmog_ctx_destroy", ctx)
