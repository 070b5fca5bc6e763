"""RealVectorizer / RealNNVectorizer / BinaryVectorizer on the GPU.

Reference (core/src/main/scala/com/salesforce/op/stages/impl/feature/): RealVectorizer.scala:44-124,
RealNNVectorizer.scala:44-57, BinaryVectorizer.scala:56-93; metadata Transmogrifier.scala:378-394.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import numpy as np
import torch

from .. import _lib as L
from .. import features as F
from .. import runtime as R
from ..columnar import ColumnarBatch, DensePart, NumericBlock, VectorColumn, round_up
from ..metadata import OpVectorMetadata, feature_column
from ..stats import LabelInfo, VectorStats
from .base import OpEstimator, OpModel, OpTransformer

MAX_LABEL_CLASSES = 100   # SanityChecker.scala:279 -- categorical iff distinct labels < min(100, 0.1 n)


def _numeric_metadata(stage, track_nulls: bool) -> OpVectorMetadata:
    cols = []
    for f in stage.inN:
        cols.append(feature_column(f.name, f.type_name))
        if track_nulls:
            cols.append(feature_column(f.name, f.type_name, is_null=True))
    return OpVectorMetadata(stage.getOutputFeatureName(), cols, stage.history())


def label_info_dev(batch: ColumnarBatch, y: torch.Tensor, name: str):
    """Distinct label classes + per-row class ids on the device (cached on the batch)."""
    cache = batch.__dict__.setdefault("_label_cache", {})
    if name in cache:
        return cache[name]
    n = batch.n
    nb = C.c_size_t()
    L.call("tmog_label_classes_scratch_bytes", C.byref(nb))
    scratch = R.scratch("label", nb.value)
    classes = torch.zeros(256, dtype=torch.float64, device="cuda")
    counts = torch.zeros(256, dtype=torch.int64, device="cuda")
    ncls = torch.zeros(1, dtype=torch.int32, device="cuda")
    ids = torch.empty(max(n, 1), dtype=torch.uint8, device="cuda")
    L.call("tmog_label_classes_dev", y.data_ptr(), n, MAX_LABEL_CLASSES, classes.data_ptr(), ncls.data_ptr(),
           ids.data_ptr(), counts.data_ptr(), scratch.data_ptr(), scratch.numel(), R.stream())
    k = int(ncls.item())
    out = (classes[:k].cpu().numpy() if k > 0 else None, counts[:k].cpu().numpy() if k > 0 else None, ids)
    cache[name] = out
    return out


class RealVectorizer(OpEstimator):
    """Fill nulls with the mean (or a constant) and optionally emit a null indicator per input."""

    input_types = (F.REAL,)

    def __init__(self, uid: Optional[str] = None):
        super().__init__("vecReal", uid)
        self.fillValue = 0.0
        self.withConstant = True      # reference default: setDefault(withConstant, true)
        self.trackNulls = True        # TransmogrifierDefaults.TrackNulls
        self.chunkRows = 1 << 20

    def setFillWithConstant(self, value: float):  # noqa: N802
        self.fillValue, self.withConstant = float(value), True
        return self

    def setFillWithMean(self):  # noqa: N802
        self.withConstant = False
        return self

    def setTrackNulls(self, v: bool):  # noqa: N802
        self.trackNulls = bool(v)
        return self

    def onSetInput(self) -> None:  # noqa: N802
        self._metadata = _numeric_metadata(self, False)     # VectorizerDefaults.onSetInput

    def fit(self, batch: ColumnarBatch) -> "RealVectorizerModel":
        R.require_cuda()
        names = [f.name for f in self.inN]
        block = batch.numeric(names)
        ncols, n = len(names), batch.n
        if self.trackNulls:
            self._metadata = _numeric_metadata(self, True)
        label_name = batch.response
        NS = L.REAL_NSTAT
        partials = np.zeros((ncols + 1) * NS)
        shifts = np.zeros(ncols + 1)
        classes = counts = null_counts = None
        if block.is_cuda:
            y = batch.label() if label_name else torch.zeros(max(n, 2), dtype=torch.float64, device="cuda")
            nb = C.c_size_t()
            L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nb))
            scratch = R.scratch("real_fit", nb.value)
            d_shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
            d_part = torch.zeros((ncols + 1) * NS, dtype=torch.float64, device="cuda")
            L.call("tmog_real_pilot_shift_dev", block.values.data_ptr(), block.ld, R.ptr(block.valid), block.ldv,
                   y.data_ptr(), n, ncols, d_shifts.data_ptr(), R.stream())
            L.call("tmog_real_fit_partials_dev", block.values.data_ptr(), block.ld, R.ptr(block.valid), block.ldv,
                   y.data_ptr(), n, ncols, d_shifts.data_ptr(), scratch.data_ptr(), scratch.numel(),
                   d_part.data_ptr(), R.stream())
            if label_name:
                classes, counts, ids = label_info_dev(batch, y, label_name)
                if classes is not None and self.trackNulls:
                    d_nc = torch.zeros((ncols, len(classes)), dtype=torch.int64, device="cuda")
                    L.call("tmog_null_class_counts_dev", R.ptr(block.valid), block.ldv, ids.data_ptr(), n, ncols,
                           len(classes), 0, d_nc.data_ptr(), R.stream())
                    null_counts = d_nc.cpu().numpy()
            partials = d_part.cpu().numpy()
            shifts = d_shifts.cpu().numpy()
        else:
            y = batch.label() if label_name else torch.zeros(max(n, 2), dtype=torch.float64)
            cls = np.zeros(256)
            cnt = np.zeros(256, dtype=np.uint64)
            ncls = C.c_int(0)
            nc = np.zeros((ncols, MAX_LABEL_CLASSES), dtype=np.uint64).reshape(-1)
            L.call("tmog_real_fit_host", R.ctx(), block.values.data_ptr(), block.ld, R.ptr(block.valid), block.ldv,
                   y.data_ptr(), n, ncols, self.chunkRows, 1, None, shifts.ctypes.data, partials.ctypes.data,
                   MAX_LABEL_CLASSES if label_name else 0, cls.ctypes.data, C.byref(ncls), cnt.ctypes.data,
                   nc.ctypes.data if self.trackNulls else None)
            if label_name and ncls.value > 0:
                k = ncls.value
                classes, counts = cls[:k].copy(), cnt[:k].astype(np.int64)
                if self.trackNulls:
                    null_counts = nc[:ncols * k].reshape(ncols, k).astype(np.int64)
        w = ncols * (2 if self.trackNulls else 1)
        fills = np.zeros(ncols)
        mean, var, mn, mx, cov = (np.zeros(w) for _ in range(5))
        lab = np.zeros(5)
        L.call("tmog_real_finish", partials.ctypes.data, shifts.ctypes.data, ncols, int(not self.withConstant),
               float(self.fillValue), int(self.trackNulls), fills.ctypes.data, mean.ctypes.data, var.ctypes.data,
               mn.ctypes.data, mx.ctypes.data, cov.ctypes.data, lab.ctypes.data, None)
        stats = None
        if label_name:
            class_sums = None
            if classes is not None:
                class_sums = np.full((len(classes), w), np.nan)
                if self.trackNulls and null_counts is not None:
                    class_sums[:, 1::2] = null_counts.T
            stats = VectorStats(LabelInfo(label_name, n, lab, classes, counts), mean, var, mn, mx, cov, class_sums)
        model = RealVectorizerModel(self, fills.tolist(), self.trackNulls)
        model._fit_stats = stats
        model._fit_key = (block.values.data_ptr(), n)
        return model


class RealVectorizerModel(OpModel):
    def __init__(self, parent: RealVectorizer, fillValues: List[float], trackNulls: bool):  # noqa: N803
        super().__init__(parent)
        self.fillValues = list(fillValues)
        self.trackNulls = trackNulls
        self.chunkRows = parent.chunkRows
        self._fit_stats: Optional[VectorStats] = None
        self._fit_key = None

    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        R.require_cuda()
        names = [f.name for f in self.inN]
        block = batch.numeric(names)
        out = _transform_numeric(block, np.asarray(self.fillValues), self.trackNulls, self.chunkRows)
        same = self._fit_key == (block.values.data_ptr(), batch.n)
        return VectorColumn(self.getOutput().name, batch.n, self.getMetadata(), [DensePart(out)],
                            self._fit_stats if same else None)


def _transform_numeric(block: NumericBlock, fills: Optional[np.ndarray], track_nulls: bool, chunk_rows: int,
                       fill_value: bool = False) -> torch.Tensor:
    n, ncols = block.n, len(block.names)
    w = ncols * (2 if track_nulls else 1)
    binary = block.values.dtype == torch.uint8
    if block.is_cuda:
        out = torch.empty((n, round_up(w, 2)), dtype=torch.float64, device="cuda")
        if binary:
            L.call("tmog_binary_transform_dev", block.values.data_ptr(), block.ld, R.ptr(block.valid), block.ldv, n,
                   ncols, int(fill_value), int(track_nulls), out.data_ptr(), out.shape[1], 0, R.stream())
        else:
            d_fills = None if fills is None else torch.from_numpy(fills).cuda()
            L.call("tmog_real_transform_dev", block.values.data_ptr(), block.ld, R.ptr(block.valid), block.ldv, n,
                   ncols, R.ptr(d_fills), int(track_nulls), out.data_ptr(), out.shape[1], 0, R.stream())
        return out[:, :w]
    if binary:
        dev = block.to("cuda")
        return _transform_numeric(dev, fills, track_nulls, chunk_rows, fill_value).cpu()
    out = torch.empty((n, w), dtype=torch.float64)
    try:
        out = out.pin_memory()
    except RuntimeError:
        pass
    L.call("tmog_real_transform_host", R.ctx(), block.values.data_ptr(), block.ld, R.ptr(block.valid), block.ldv, n,
           ncols, None if fills is None else fills.ctypes.data, int(track_nulls), chunk_rows, out.data_ptr(), w, 0)
    return out


class RealNNVectorizer(OpTransformer):
    """Dense copy of non-nullable reals (RealNNVectorizer.scala:52-55)."""

    input_types = (F.REALNN,)

    def __init__(self, uid: Optional[str] = None):
        super().__init__("vecNum", uid)
        self.chunkRows = 1 << 20

    def onSetInput(self) -> None:  # noqa: N802
        self._metadata = _numeric_metadata(self, False)

    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        R.require_cuda()
        block = batch.numeric([f.name for f in self.inN])
        if block.valid is not None:
            raise ValueError("None.get: a RealNN feature holds a null")
        out = _transform_numeric(block, None, False, self.chunkRows)
        return VectorColumn(self.getOutput().name, batch.n, self.getMetadata(), [DensePart(out)], None)


class BinaryVectorizer(OpTransformer):
    """[value or fillValue, isNull] per Binary input (BinaryVectorizer.scala:70-92); no fit."""

    input_types = (F.BINARY,)

    def __init__(self, uid: Optional[str] = None):
        super().__init__("vecBin", uid)
        self.trackNulls = True
        self.fillValue = False
        self.chunkRows = 1 << 20

    def setTrackNulls(self, v: bool):  # noqa: N802
        self.trackNulls = bool(v)
        return self

    def setFillValue(self, v: bool):  # noqa: N802
        self.fillValue = bool(v)
        return self

    def onSetInput(self) -> None:  # noqa: N802
        self._metadata = _numeric_metadata(self, False)

    def onGetMetadata(self) -> None:  # noqa: N802
        self._metadata = _numeric_metadata(self, self.trackNulls)

    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        R.require_cuda()
        block = batch.numeric([f.name for f in self.inN])
        out = _transform_numeric(block, None, self.trackNulls, self.chunkRows, self.fillValue)
        return VectorColumn(self.getOutput().name, batch.n, self.getMetadata(), [DensePart(out)], None)
