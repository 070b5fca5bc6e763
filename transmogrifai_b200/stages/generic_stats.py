"""SanityChecker kernels over an already materialised OPVector (dense parts): used when the input vector carries no
statistics sidecar (it was not produced by the fused vectorizer scans), for CorrelationLevel.Computed, and for the
model's indicesToKeep gather."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np
import torch

from .. import _lib as L
from .. import dist as DD
from .. import runtime as R
from ..columnar import (ColumnarBatch, CsrPart, DensePart, LazyDensePart, LazyOneHotPart, LazyPart, OneHotPart,
                        VectorColumn, WidthPart)
from ..stats import LabelInfo, VectorStats
from .numeric import label_info_dev


def _device_label(batch: ColumnarBatch, label_name: str) -> torch.Tensor:
    cache = batch.__dict__.setdefault("_label_dev", {})
    if label_name not in cache:
        y = batch.label(label_name)[:batch.n]
        cache[label_name] = y if y.is_cuda else y.cuda()
    return cache[label_name]


DENSIFY_LIMIT_BYTES = 64 << 30
GRAM_TILE_BYTES = 4 << 30          # a vector larger than this goes through the Gramian kernel in row tiles


def _materialized(vec: VectorColumn) -> list:
    return [p.materialize(None) if isinstance(p, (LazyPart, LazyOneHotPart)) else p for p in vec.parts]


def _dense_block(vec: VectorColumn, parts: Optional[list] = None, a: int = 0, b: Optional[int] = None) -> torch.Tensor:
    """Rows [a, b) of all parts as one device row-major block: a view for a single dense part, else a copy in which
    sparse parts are densified (what Spark's RowMatrix does for Statistics.corr)."""
    from ..columnar import SparseRowsPart
    parts = _materialized(vec) if parts is None else parts
    b = vec.n if b is None else b
    if len(parts) == 1 and isinstance(parts[0], (DensePart, LazyDensePart)):
        d = parts[0].data[a:b]
        return d if d.is_cuda else d.cuda()
    n, w = b - a, sum(p.width for p in parts)
    if n * w * 8 > DENSIFY_LIMIT_BYTES:
        raise ValueError(f"densifying a {n} x {w} vector needs {n * w * 8 >> 30} GiB: use featureFeatureCorrLevel = Off "
                         "for wide hashed / pivoted vectors (the reference cannot exceed 65535 columns either)")
    out = torch.zeros((n, w), dtype=torch.float64, device="cuda")
    off = 0
    for p in parts:
        if isinstance(p, (DensePart, LazyDensePart)):
            d = p.data[a:b]
            out[:, off:off + p.width] = d if d.is_cuda else d.cuda()
        elif isinstance(p, OneHotPart):
            idx = p.indices[a:b]
            idx = idx if idx.is_cuda else idx.cuda()
            L.call("tmog_onehot_densify_dev", idx.data_ptr(), idx.stride(0), n, idx.shape[1], out.data_ptr(), w, off,
                   R.stream())
        elif isinstance(p, SparseRowsPart):
            rs, rn = p.row_start[a:b], p.row_nnz[a:b]
            L.call("tmog_sparse_rows_densify_dev", rs.data_ptr(), rn.data_ptr(), p.indices.data_ptr(),
                   p.values.data_ptr(), n, out.data_ptr(), w, off, R.stream())
        else:
            raise NotImplementedError(f"cannot densify a {type(p).__name__}")
        off += p.width
    return out


def vector_stats(batch: ColumnarBatch, vec: VectorColumn, label_name: str) -> VectorStats:
    R.require_cuda()
    x = _dense_block(vec)
    n, d = x.shape
    y = _device_label(batch, label_name)
    classes, counts, ids = label_info_dev(batch, y, label_name)
    cat = np.array([c.grouping is not None and c.indicatorValue is not None for c in vec.metadata.columns],
                   dtype=np.uint8)
    d_cat = torch.from_numpy(cat).cuda()
    nb = C.c_size_t()
    L.call("tmog_dense_stats_scratch_bytes", n, d, C.byref(nb))
    scratch = R.scratch("dense_stats", nb.value)
    shifts = torch.zeros(d + 1, dtype=torch.float64, device="cuda")
    partials = torch.zeros((d + 1) * L.REAL_NSTAT, dtype=torch.float64, device="cuda")
    ncls = 0 if classes is None else len(classes)
    class_sums = torch.zeros((max(ncls, 1), d), dtype=torch.float64, device="cuda") if ncls else None
    L.call("tmog_dense_pilot_shift_dev", x.data_ptr(), x.stride(0), n, d, y.data_ptr(), shifts.data_ptr(), R.stream())
    if DD.world()[1] > 1:                                  # partials only merge under common shifts (rank 0's)
        torch.distributed.broadcast(shifts, src=0)
    L.call("tmog_dense_stats_dev", x.data_ptr(), x.stride(0), n, d, y.data_ptr(), ids.data_ptr() if ncls else None,
           ncls, d_cat.data_ptr(), shifts.data_ptr(), scratch.data_ptr(), scratch.numel(), partials.data_ptr(),
           R.ptr(class_sums), R.stream())
    p, _ = DD.gather_fit_results(partials)                 # rows of all ranks
    if DD.world()[1] > 1 and class_sums is not None:
        torch.distributed.all_reduce(class_sums)
    s = shifts.cpu().numpy()
    mean, var, mn, mx, cov = (np.zeros(d) for _ in range(5))
    lab = np.zeros(5)
    L.call("tmog_real_finish", p.ctypes.data, s.ctypes.data, d, 1, 0.0, 0, None, mean.ctypes.data, var.ctypes.data,
           mn.ctypes.data, mx.ctypes.data, cov.ctypes.data, lab.ctypes.data, None)
    cs = None
    if ncls:
        cs = class_sums.cpu().numpy()
        cs[:, cat == 0] = np.nan
    st = VectorStats(LabelInfo(label_name, int(lab[0]), lab, classes, counts), mean, var, mn, mx, cov, cs)
    st._shifted_sums = (s, p.reshape(d + 1, L.REAL_NSTAT)[:, 1].copy())
    return st


class PearsonAccumulator:
    """Gramian and column sums of the shifted rows ([selected columns, label]) of one batch or of many: what
    Statistics.corr(vectorRowsForCorr, "pearson") needs (SanityChecker.scala:467-470).  Rows go through the Gramian
    kernel in tiles (the kernel adds to the matrix), so the limit is the (d + 1)^2 matrix itself -- the reference's
    65535 columns -- not rows x columns.  `shifts` ([d + 1], common to every rank and batch: e.g. the column means
    already known from the fit scans) conditions the sums; without it the first rows seen give a pilot shift."""

    def __init__(self, d_all: int, corr_indices: Sequence[int], shifts: Optional[np.ndarray] = None):
        R.require_cuda()
        feat = [int(i) for i in corr_indices if i < d_all]
        if len(feat) + 1 > 65535:
            raise ValueError(f"requirement failed: Argument with more than 65535 cols: {len(feat) + 1}")
        self.d_all, self.d, self.D = d_all, len(feat), len(feat) + 1
        self.select = None if feat == list(range(d_all)) else torch.tensor(feat, dtype=torch.int32, device="cuda")
        D = self.D
        self.primed = shifts is not None
        if shifts is None:
            self.shifts = torch.zeros(D, dtype=torch.float64, device="cuda")
        else:
            s = np.nan_to_num(np.asarray(shifts, dtype=np.float64), nan=0.0, posinf=0.0, neginf=0.0)
            if s.shape != (D,):
                raise ValueError(f"shifts: expected {D} values, got {s.shape}")
            self.shifts = torch.from_numpy(np.ascontiguousarray(s)).cuda()
        self.gram = torch.zeros((D, D), dtype=torch.float64, device="cuda")
        self.sums = torch.zeros(D * L.REAL_NSTAT, dtype=torch.float64, device="cuda")
        self.partials = torch.zeros(D * L.REAL_NSTAT, dtype=torch.float64, device="cuda")

    def update(self, batch: ColumnarBatch, vec: VectorColumn, label_name: str) -> None:
        n, d_all, d = vec.n, self.d_all, self.d
        if n <= 0:
            return
        parts = _materialized(vec)
        if sum(p.width for p in parts) != d_all:
            raise ValueError(f"the vector has {sum(p.width for p in parts)} columns, expected {d_all}")
        y = _device_label(batch, label_name)
        tile = n if n * d_all * 8 <= GRAM_TILE_BYTES else max(1024, GRAM_TILE_BYTES // (d_all * 8) // 64 * 64)
        nb = C.c_size_t()
        L.call("tmog_dense_stats_scratch_bytes", min(tile, n), d, C.byref(nb))
        scratch = R.scratch("dense_stats", nb.value)
        for a in range(0, n, tile):
            b = min(n, a + tile)
            x = _dense_block(vec, parts, a, b)
            if self.select is not None:
                sub = torch.empty((b - a, d), dtype=torch.float64, device="cuda")
                L.call("tmog_dense_gather_dev", x.data_ptr(), x.stride(0), b - a, self.select.data_ptr(), d,
                       sub.data_ptr(), d, R.stream())
                x = sub
            ya = y[a:]
            if not self.primed:
                L.call("tmog_dense_pilot_shift_dev", x.data_ptr(), x.stride(0), b - a, d, ya.data_ptr(),
                       self.shifts.data_ptr(), R.stream())
                if DD.world()[1] > 1:
                    torch.distributed.broadcast(self.shifts, src=0)
                self.primed = True
            L.call("tmog_dense_stats_dev", x.data_ptr(), x.stride(0), b - a, d, ya.data_ptr(), None, 0, None,
                   self.shifts.data_ptr(), scratch.data_ptr(), scratch.numel(), self.partials.data_ptr(), None,
                   R.stream())
            self.sums += self.partials                     # only the shifted sums ([.][1]) and the row count are used
            L.call("tmog_dense_gram_dev", x.data_ptr(), x.stride(0), b - a, d, ya.data_ptr(), self.shifts.data_ptr(),
                   self.gram.data_ptr(), R.stream())
            del x

    def finish(self) -> np.ndarray:
        """-> the [d + 1, d + 1] Pearson matrix over the rows of every update of every rank."""
        D = self.D
        if DD.world()[1] > 1:                              # Gramian and sums of the shifted rows add up over the ranks
            torch.distributed.all_reduce(self.gram)
            torch.distributed.all_reduce(self.sums)
        g = self.gram.cpu().numpy()
        merged = self.sums.cpu().numpy().reshape(D, L.REAL_NSTAT)
        col_sums = np.ascontiguousarray(merged[:, 1])
        n_all = float(merged[D - 1, 7])
        corr = np.zeros((D, D))
        L.call("tmog_pearson_from_gram", D, n_all, g.ctypes.data, col_sums.ctypes.data, corr.ctypes.data)
        return corr


def pearson_matrix(batch: ColumnarBatch, vec: VectorColumn, label_name: str, corr_indices: Sequence[int],
                   st: Optional[VectorStats]) -> np.ndarray:
    """Statistics.corr(vectorRowsForCorr, "pearson") over the selected columns + label of one resident batch
    (SanityChecker.scala:467-470)."""
    acc = PearsonAccumulator(sum(p.width for p in vec.parts), corr_indices)
    acc.update(batch, vec, label_name)
    return acc.finish()


def select_columns(vec: VectorColumn, keep: List[int]) -> list:
    """removeFeatures: the kept columns, in order (DerivedFeatureFilterUtils.scala:291-305).  Lazy parts are written
    once with just their kept columns; materialised dense parts go through the gather kernel."""
    R.require_cuda()
    out, off = [], 0
    keep = np.asarray(keep, dtype=np.int64)
    for p in vec.parts:
        local = keep[(keep >= off) & (keep < off + p.width)] - off
        off += p.width
        if len(local) == 0:
            continue
        whole = len(local) == p.width and bool((local == np.arange(p.width)).all())
        if isinstance(p, WidthPart):
            out.append(WidthPart(len(local)))
        elif isinstance(p, LazyDensePart):
            out.append(DensePart(p.materialize(local)))
        elif isinstance(p, DensePart):
            if whole:
                out.append(p)
                continue
            x = p.data if p.data.is_cuda else p.data.cuda()
            n = x.shape[0]
            idx = torch.from_numpy(local.astype(np.int32)).cuda()
            dst = torch.empty((n, len(local)), dtype=torch.float64, device="cuda")
            L.call("tmog_dense_gather_dev", x.data_ptr(), x.stride(0), n, idx.data_ptr(), len(local),
                   dst.data_ptr(), len(local), R.stream())
            out.append(DensePart(dst))
        elif isinstance(p, (LazyOneHotPart, LazyPart)):
            out.append(p.materialize(local))
        elif isinstance(p, OneHotPart):
            if whole:
                out.append(p)
                continue
            newpos = torch.full((p.width + 1,), -1, dtype=torch.int32, device=p.indices.device)
            newpos[torch.from_numpy(local).to(p.indices.device)] = torch.arange(
                len(local), dtype=torch.int32, device=p.indices.device)
            idx = p.indices.long()
            out.append(OneHotPart(newpos[torch.where(idx < 0, torch.full_like(idx, p.width), idx)], len(local)))
        else:
            raise NotImplementedError("column selection on CSR parts")
    return out
