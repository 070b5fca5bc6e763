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


def _dense_block(vec: VectorColumn) -> torch.Tensor:
    """All parts as one device row-major block: a view for a single dense part, else a copy in which sparse parts are
    densified (what Spark's RowMatrix does for Statistics.corr)."""
    from ..columnar import SparseRowsPart
    parts = [p.materialize(None) if isinstance(p, (LazyPart, LazyOneHotPart)) else p for p in vec.parts]
    if len(parts) == 1 and isinstance(parts[0], (DensePart, LazyDensePart)):
        d = parts[0].data
        return d if d.is_cuda else d.cuda()
    n, w = vec.n, sum(p.width for p in parts)
    if n * w * 8 > DENSIFY_LIMIT_BYTES:
        raise ValueError(f"densifying a {n} x {w} vector needs {n * w * 8 >> 30} GiB: use featureFeatureCorrLevel = Off "
                         "for wide hashed / pivoted vectors (the reference cannot exceed 65535 columns either)")
    out = torch.zeros((n, w), dtype=torch.float64, device="cuda")
    off = 0
    for p in parts:
        if isinstance(p, (DensePart, LazyDensePart)):
            d = p.data
            out[:, off:off + p.width] = d if d.is_cuda else d.cuda()
        elif isinstance(p, OneHotPart):
            idx = p.indices if p.indices.is_cuda else p.indices.cuda()
            L.call("tmog_onehot_densify_dev", idx.data_ptr(), idx.stride(0), n, idx.shape[1], out.data_ptr(), w, off,
                   R.stream())
        elif isinstance(p, SparseRowsPart):
            L.call("tmog_sparse_rows_densify_dev", p.row_start.data_ptr(), p.row_nnz.data_ptr(), p.indices.data_ptr(),
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


def pearson_matrix(batch: ColumnarBatch, vec: VectorColumn, label_name: str, corr_indices: Sequence[int],
                   st: Optional[VectorStats]) -> np.ndarray:
    """Statistics.corr(vectorRowsForCorr, "pearson") over the selected columns + label (SanityChecker.scala:467-470)."""
    R.require_cuda()
    x = _dense_block(vec)
    n, d_all = x.shape
    feat = [i for i in corr_indices if i < d_all]
    if len(feat) + 1 > 65535:
        raise ValueError(f"requirement failed: Argument with more than 65535 cols: {len(feat) + 1}")
    if feat != list(range(d_all)):
        idx = torch.tensor(feat, dtype=torch.int32, device="cuda")
        sub = torch.empty((n, len(feat)), dtype=torch.float64, device="cuda")
        L.call("tmog_dense_gather_dev", x.data_ptr(), x.stride(0), n, idx.data_ptr(), len(feat), sub.data_ptr(),
               len(feat), R.stream())
        x = sub
    d = x.shape[1]
    y = _device_label(batch, label_name)
    D = d + 1
    dist_world = DD.world()[1]
    shifts = torch.zeros(D, dtype=torch.float64, device="cuda")
    L.call("tmog_dense_pilot_shift_dev", x.data_ptr(), x.stride(0), n, d, y.data_ptr(), shifts.data_ptr(), R.stream())
    if dist_world > 1:
        torch.distributed.broadcast(shifts, src=0)
    nb = C.c_size_t()
    L.call("tmog_dense_stats_scratch_bytes", n, d, C.byref(nb))
    scratch = R.scratch("dense_stats", nb.value)
    partials = torch.zeros(D * L.REAL_NSTAT, dtype=torch.float64, device="cuda")
    L.call("tmog_dense_stats_dev", x.data_ptr(), x.stride(0), n, d, y.data_ptr(), None, 0, None, shifts.data_ptr(),
           scratch.data_ptr(), scratch.numel(), partials.data_ptr(), None, R.stream())
    gram = torch.zeros((D, D), dtype=torch.float64, device="cuda")
    L.call("tmog_dense_gram_dev", x.data_ptr(), x.stride(0), n, d, y.data_ptr(), shifts.data_ptr(), gram.data_ptr(),
           R.stream())
    if dist_world > 1:                                     # Gramian and sums of the shifted rows add up over the ranks
        torch.distributed.all_reduce(gram)
    g = gram.cpu().numpy()
    merged, _ = DD.gather_fit_results(partials)
    merged = merged.reshape(D, L.REAL_NSTAT)
    sums = np.ascontiguousarray(merged[:, 1])
    n_all = float(merged[D - 1, 7]) if dist_world > 1 else float(n)
    corr = np.zeros((D, D))
    L.call("tmog_pearson_from_gram", D, n_all, g.ctypes.data, sums.ctypes.data, corr.ctypes.data)
    return corr


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
