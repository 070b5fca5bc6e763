"""Seeded synthetic inputs shared by the tests, smoke() and bench.py (shapes of BASELINE.json's configs)."""
from __future__ import annotations

import numpy as np


def pack_validity(valid: np.ndarray) -> np.ndarray:
    """bool [ncols, n] -> Arrow-style bitmap words uint64 [ncols, ceil(n/64)] (bit r%64 of word r/64, 1 = present)."""
    ncols, n = valid.shape
    nw = (n + 63) // 64
    padded = np.zeros((ncols, nw * 64), dtype=np.uint8)
    padded[:, :n] = valid
    bits = np.packbits(padded.reshape(ncols, nw, 64), axis=2, bitorder="little")  # [ncols, nw, 8] bytes
    return bits.reshape(ncols, nw * 8).view("<u8").reshape(ncols, nw).copy()


def unpack_validity(words: np.ndarray, n: int) -> np.ndarray:
    ncols = words.shape[0]
    b = np.unpackbits(words.view(np.uint8).reshape(ncols, -1), axis=1, bitorder="little")
    return b[:, :n].astype(bool)


def make_real_config(n: int, ncols: int, seed: int = 20260001, null_frac: float = 0.05):
    """Config 2 of BASELINE.json: N(mu_j, sigma_j) columns, mu_j = j, sigma_j = 1 + j/10, 5 % nulls,
    Bernoulli(0.3) label, 5 columns shifted by +0.5*y, one constant column, one duplicated column.
    Returns (values [ncols, n] f64, valid [ncols, n] bool, label [n] f64)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    y = (rng.random(n) < 0.3).astype(np.float64)
    x = np.empty((ncols, n), dtype=np.float64)
    for j in range(ncols):
        x[j] = rng.standard_normal(n) * (1.0 + j / 10.0) + j
    for j in range(0, min(ncols, 50), 10):
        x[j] += 0.5 * y
    if ncols > 7:
        x[7] = 3.25                      # constant column -> variance rule
    if ncols > 12:
        x[12] = x[11]                    # duplicate -> feature-feature rule
    valid = rng.random((ncols, n)) >= null_frac
    if ncols > 12:
        valid[12] = valid[11]
    return x, valid, y
