"""ctypes wrapper of oracle/c/liboracle_hot.so -- the C + OpenMP restatement of the reference's hot loops.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): the checker for sizes the Python oracle cannot reach, and the
timed CPU baseline of bench.py.  Built by oracle/c/Makefile (``__graft_entry__.build()`` runs it).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "c", "liboracle_hot.so")
_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} missing: run `make -C oracle/c`")
        _lib = C.CDLL(LIB_PATH)
        _lib.orc_max_threads.restype = C.c_int
    return _lib


def max_threads() -> int:
    return int(lib().orc_max_threads())


def _p(a: Optional[np.ndarray]):
    return None if a is None else C.c_void_p(a.ctypes.data)


def real_fit(values: np.ndarray, valid_words: Optional[np.ndarray], threads: int = 1) -> np.ndarray:
    """values [ncols, ld] f64 C-contiguous (n = ld), valid_words [ncols, ldv] uint64 or None -> fills [ncols]"""
    ncols, ld = values.shape
    fills = np.zeros(ncols)
    ldv = 0 if valid_words is None else valid_words.shape[1]
    lib().orc_real_fit(_p(values), C.c_int64(ld), _p(valid_words), C.c_int64(ldv), C.c_int64(ld), C.c_int(ncols),
                       C.c_int(threads), _p(fills))
    return fills


def real_transform(values: np.ndarray, valid_words: Optional[np.ndarray], fills: np.ndarray, track_nulls: bool = True,
                   threads: int = 1, out: Optional[np.ndarray] = None, extra_cols: int = 0) -> np.ndarray:
    """-> dense row-major [n, width + extra_cols] (extra columns left for the caller, e.g. the label)"""
    ncols, ld = values.shape
    w = ncols * (2 if track_nulls else 1)
    if out is None:
        out = np.empty((ld, w + extra_cols))
    ldv = 0 if valid_words is None else valid_words.shape[1]
    lib().orc_real_transform(_p(values), C.c_int64(ld), _p(valid_words), C.c_int64(ldv), C.c_int64(ld),
                             C.c_int(ncols), _p(fills), C.c_int(int(track_nulls)), C.c_int(threads), _p(out),
                             C.c_int64(out.shape[1]))
    return out


def col_stats(x: np.ndarray, threads: int = 1) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
    """x [n, d] row-major dense -> (mean, variance, min, max) as Spark's MultivariateOnlineSummarizer"""
    n, d = x.shape
    mean, var, mn, mx = (np.zeros(d) for _ in range(4))
    lib().orc_col_stats(_p(x), C.c_int64(x.strides[0] // 8), C.c_int64(n), C.c_int(d), C.c_int(threads), _p(mean),
                        _p(var), _p(mn), _p(mx))
    return mean, var, mn, mx


def label_cov(x: np.ndarray, mean: np.ndarray, threads: int = 1) -> np.ndarray:
    n, d = x.shape
    cov = np.zeros(d)
    lib().orc_label_cov(_p(x), C.c_int64(x.strides[0] // 8), C.c_int64(n), C.c_int(d), _p(mean), C.c_int(threads),
                        _p(cov))
    return cov


def label_sums(x: np.ndarray, y: np.ndarray, labels: np.ndarray, threads: int = 1) -> np.ndarray:
    n, d = x.shape
    out = np.zeros((len(labels), d + 1))
    lib().orc_label_sums(_p(x), C.c_int64(x.strides[0] // 8), _p(y), C.c_int64(n), C.c_int(d), _p(labels),
                         C.c_int(len(labels)), C.c_int(threads), _p(out))
    return out


def real_pipeline(values: np.ndarray, valid_words: Optional[np.ndarray], y: np.ndarray, threads: int = 1):
    """The reference's job sequence for config 2: RealVectorizer.fit -> transform (materialised rows, label appended
    as SanityChecker.scala:397-404 does) -> Statistics.colStats -> label correlations (CorrelationLevel.Off route).
    Returns (fills, rows [n, 2*ncols+1], mean, variance, min, max, corr)."""
    ncols, n = values.shape
    fills = real_fit(values, valid_words, threads)
    rows = real_transform(values, valid_words, fills, True, threads, extra_cols=1)
    rows[:, -1] = y
    mean, var, mn, mx = col_stats(rows, threads)
    cov = label_cov(rows, mean, threads)
    sd = np.sqrt(var)
    with np.errstate(divide="ignore", invalid="ignore"):
        corr = cov / (sd * sd[-1] * (n - 1))
    return fills, rows, mean, var, mn, mx, corr


# ---------------------------------------------------------------------------------------------------------------------
# string paths (oracle/c/oracle_text.c)
# ---------------------------------------------------------------------------------------------------------------------
def encode_text_columns(columns):
    """Python columns (str / None cells) -> per column (bytes uint8, offsets int64 [n+1], validity uint64 words)."""
    out = []
    for col in columns:
        n = len(col)
        enc = [b"" if v is None else v.encode("utf-8") for v in col]
        offs = np.zeros(n + 1, dtype=np.int64)
        offs[1:] = np.cumsum([len(e) for e in enc])
        data = np.frombuffer(b"".join(enc) or b"\0", dtype=np.uint8).copy()
        bits = np.zeros((n + 63) // 64 * 64, dtype=np.uint8)
        bits[:n] = [v is not None for v in col]
        valid = np.packbits(bits, bitorder="little").view(np.uint64).copy()
        out.append((data, offs, valid))
    return out


def _ptr_arrays(cols):
    data = np.array([c[0].ctypes.data for c in cols], dtype=np.uintp)
    offs = np.array([c[1].ctypes.data for c in cols], dtype=np.uintp)
    valid = np.array([0 if c[2] is None else c[2].ctypes.data for c in cols], dtype=np.uintp)
    return data, offs, valid


def text_hash(cols, n: int, prefixes, num_features: int, shared: bool, binary_freq: bool = False,
              min_token_length: int = 1, track_nulls: bool = True, class_ids: Optional[np.ndarray] = None,
              nclasses: int = 1, threads: int = 1, want_stats: bool = True, want_rows: bool = True):
    """cols: [(data, offsets, valid)] per hashed column; prefixes: the "<nameHash>_" bytes per column ("" = none).
    -> dict(class_sum [width, nclasses], sq_sum, nnz, vmax, vmin, row_start, row_nnz, idx, val)"""
    ncols = len(cols)
    hash_width = num_features if shared else num_features * ncols
    null_base = hash_width
    width = null_base + ncols
    pre = np.zeros((max(ncols, 1), 16), dtype=np.uint8)
    plen = np.zeros(max(ncols, 1), dtype=np.int32)
    for j, p in enumerate(prefixes):
        b = p.encode("ascii")
        pre[j, :len(b)] = np.frombuffer(b, dtype=np.uint8)
        plen[j] = len(b)
    dp, op, vp = _ptr_arrays(cols)
    res = {}
    cs = sq = nz = mx = mn = None
    if want_stats:
        cs = np.zeros((width, nclasses), dtype=np.int64)
        sq, nz = np.zeros(width, dtype=np.int64), np.zeros(width, dtype=np.int64)
        mx = np.zeros(width, dtype=np.int32)
        mn = np.full(width, np.iinfo(np.int32).max, dtype=np.int32)
    rs = np.zeros(max(n, 1), dtype=np.int64)
    rn = np.zeros(max(n, 1), dtype=np.int32)
    oi, ov, tot = C.c_void_p(), C.c_void_p(), C.c_int64(0)
    f = lib().orc_text_hash
    f.restype = C.c_int
    rc = f(_p(dp), _p(op), _p(vp), C.c_int(ncols), C.c_int64(n), _p(pre), _p(plen), C.c_int(num_features),
           C.c_int(int(shared)), C.c_int(int(binary_freq)), C.c_int(min_token_length), C.c_int(int(track_nulls)),
           C.c_int(null_base), _p(class_ids), C.c_int(nclasses), C.c_int(threads), _p(cs), _p(sq), _p(nz), _p(mx),
           _p(mn), C.c_int(int(want_rows)), _p(rs), _p(rn), C.byref(oi), C.byref(ov), C.byref(tot))
    if rc != 0:
        raise MemoryError("orc_text_hash ran out of memory")
    res.update(class_sum=cs, sq_sum=sq, nnz=nz, vmax=mx, vmin=mn, width=width)
    if want_rows:
        t = tot.value
        res["row_start"], res["row_nnz"] = rs[:n], rn[:n]
        res["idx"] = np.ctypeslib.as_array(C.cast(oi, C.POINTER(C.c_int32)), shape=(max(t, 1),))[:t].copy()
        res["val"] = np.ctypeslib.as_array(C.cast(ov, C.POINTER(C.c_int32)), shape=(max(t, 1),))[:t].copy()
        lib().orc_free(oi)
        lib().orc_free(ov)
    return res


def text_stats(cols, n: int, clean: bool, max_card: int, threads: int = 1):
    ncols = len(cols)
    dp, op, vp = _ptr_arrays(cols)
    distinct = np.zeros(ncols, dtype=np.int32)
    nonnull, ls, lq = (np.zeros(ncols, dtype=np.int64) for _ in range(3))
    f = lib().orc_text_stats
    f.restype = C.c_int
    if f(_p(dp), _p(op), _p(vp), C.c_int(ncols), C.c_int64(n), C.c_int(int(clean)), C.c_int(max_card),
         C.c_int(threads), _p(distinct), _p(nonnull), _p(ls), _p(lq)) != 0:
        raise MemoryError("orc_text_stats ran out of memory")
    return distinct, nonnull, ls, lq


def clean_string(s: str) -> str:
    b = np.frombuffer(s.encode("ascii") or b"\0", dtype=np.uint8).copy()
    out = np.zeros(max(len(b), 1), dtype=np.uint8)
    f = lib().orc_clean_string
    f.restype = C.c_int64
    m = f(_p(b), C.c_int64(len(s)), _p(out))
    return out[:m].tobytes().decode("ascii")


def pivot_counts(codes: np.ndarray, n: int, ncodes: np.ndarray, class_ids: Optional[np.ndarray], nclasses: int,
                 threads: int = 1):
    """codes [ncols, ld] int32 -> list per column of counts [(1 + ncodes_c), nclasses] (row 0 = nulls)"""
    ncols, ld = codes.shape
    sizes = (np.asarray(ncodes, dtype=np.int64) + 1) * nclasses
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
    counts = np.zeros(int(off[-1]), dtype=np.int64)
    f = lib().orc_pivot_counts
    f.restype = C.c_int
    if f(_p(codes), C.c_int64(ld), C.c_int64(n), C.c_int(ncols), _p(off), C.c_int64(int(off[-1])), _p(class_ids),
         C.c_int(nclasses), C.c_int(threads), _p(counts)) != 0:
        raise MemoryError("orc_pivot_counts ran out of memory")
    return [counts[off[c]:off[c + 1]].reshape(-1, nclasses) for c in range(ncols)]


def pivot_transform(codes: np.ndarray, n: int, pos_tables, threads: int = 1) -> np.ndarray:
    """pos_tables[c]: int32 [ncodes_c + 1] (last entry = the null indicator's index) -> out [n, ncols] int32"""
    ncols, ld = codes.shape
    ncodes = np.array([len(t) - 1 for t in pos_tables], dtype=np.int32)
    off = np.concatenate([[0], np.cumsum([len(t) for t in pos_tables])]).astype(np.int64)
    pos = np.ascontiguousarray(np.concatenate(pos_tables), dtype=np.int32)
    out = np.zeros((n, ncols), dtype=np.int32)
    lib().orc_pivot_transform(_p(codes), C.c_int64(ld), C.c_int64(n), C.c_int(ncols), _p(pos), _p(off), _p(ncodes),
                              C.c_int(threads), _p(out))
    return out


def clean_strings(values):
    """cleanString of every entry of a list of ASCII strings."""
    enc = [v.encode("ascii") for v in values]
    offs = np.zeros(len(enc) + 1, dtype=np.int64)
    offs[1:] = np.cumsum([len(e) for e in enc])
    data = np.frombuffer(b"".join(enc) or b"\0", dtype=np.uint8).copy()
    out = np.zeros(max(int(offs[-1]), 1), dtype=np.uint8)
    ooffs = np.zeros(len(enc) + 1, dtype=np.int64)
    lib().orc_clean_strings(_p(data), _p(offs), C.c_int64(len(enc)), _p(out), _p(ooffs))
    raw = out.tobytes()
    return [raw[ooffs[i]:ooffs[i + 1]].decode("ascii") for i in range(len(enc))]
