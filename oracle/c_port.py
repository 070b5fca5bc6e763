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
