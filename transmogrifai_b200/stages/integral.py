"""IntegralVectorizer on the GPU: fill nulls with the mode (or a constant), optional null indicators, and the
SanityChecker statistics of the emitted columns from the same scan of the raw int64 columns.

Reference: core/src/main/scala/com/salesforce/op/stages/impl/feature/IntegralVectorizer.scala:47-117 (fitFn :81-90,
transformFn :103-113), utils/src/main/scala/com/salesforce/op/utils/spark/SequenceAggregators.scala:100-131
(ModeSeqNullInt: most frequent value, the smallest on ties, 0 for an all-null column); Transmogrifier defaults
FillWithMode = true, TrackNulls = true (Transmogrifier.scala:61-70,258-262).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import numpy as np
import torch

from .. import _lib as L
from .. import dist as D
from .. import features as F
from .. import runtime as R
from ..columnar import batch_key, ColumnarBatch, LazyDensePart, NumericBlock, VectorColumn
from ..stats import LabelInfo, VectorStats
from .base import OpEstimator, OpModel
from .numeric import MAX_LABEL_CLASSES, _numeric_metadata, _transform_numeric, label_info_dev

INT64_MIN = -(1 << 63)


def column_modes(dev: NumericBlock, n: int) -> List[int]:
    """ModeSeqNullInt of every column of a device block over the rows of all ranks."""
    ncols = len(dev.names)
    if ncols == 0:
        return []
    size = D.world()[1]
    cap_max = 64
    while cap_max < 2 * max(n, 1):
        cap_max <<= 1
    cap = min(1 << 16, cap_max)
    while True:
        nb = C.c_size_t()
        L.call("tmog_integral_mode_scratch_bytes", ncols, cap, C.byref(nb))
        scratch = torch.empty(nb.value, dtype=torch.uint8, device="cuda")
        modes = torch.zeros(ncols, dtype=torch.int64, device="cuda")
        counts = torch.zeros(ncols, dtype=torch.int64, device="cuda")
        L.call("tmog_integral_mode_dev", dev.values.data_ptr(), dev.ld, R.ptr(dev.valid), dev.ldv, n, ncols, cap,
               scratch.data_ptr(), nb.value, modes.data_ptr(), counts.data_ptr(), R.stream())
        offs = (C.c_size_t * 5)()
        L.call("tmog_integral_mode_layout", ncols, cap, offs)
        overflow = scratch[offs[4]:offs[4] + 4 * ncols].view(torch.int32)
        full = bool(overflow.any().item())
        if size > 1:
            full = bool(D.allreduce_max(np.array([int(full)], dtype=np.int64))[0])
        if not full:
            break
        if cap >= cap_max:
            raise RuntimeError("integral mode table overflow at its largest size")     # cannot happen: cap >= 2 n
        cap = min(cap * 16, cap_max)
    if size == 1:
        return [int(v) for v in modes.cpu().numpy()]
    # several ranks: their value -> count maps add up key-wise (the Algebird map monoid of ModeSeqNullInt.merge)
    keys = scratch[offs[0]:offs[0] + 8 * ncols * cap].view(torch.int64).view(ncols, cap)
    cnts = scratch[offs[1]:offs[1] + 8 * ncols * cap].view(torch.int64).view(ncols, cap)
    nmin = scratch[offs[2]:offs[2] + 8 * ncols].view(torch.int64).cpu().numpy()
    local = []
    for c in range(ncols):
        used = cnts[c] > 0
        k, w = keys[c][used].cpu().numpy(), cnts[c][used].cpu().numpy()
        if nmin[c] > 0:
            k, w = np.append(k, INT64_MIN), np.append(w, nmin[c])
        local.append((k, w))
    out = []
    for c, per_rank in enumerate(zip(*D.all_gather_object(local))):
        k = np.concatenate([p[0] for p in per_rank])
        w = np.concatenate([p[1] for p in per_rank])
        if len(k) == 0:
            out.append(0)
            continue
        uk, inv = np.unique(k, return_inverse=True)
        tot = np.zeros(len(uk), dtype=np.int64)
        np.add.at(tot, inv, w)
        out.append(int(uk[np.lexsort((uk, -tot))[0]]))
    return out


class IntegralVectorizer(OpEstimator):
    """Fill nulls with the mode (or a constant) and optionally emit a null indicator per input."""

    input_types = F.INTEGRAL_TYPES        # IntegralVectorizer[T <: Integral]: Integral, Date, DateTime

    def __init__(self, uid: Optional[str] = None):
        super().__init__("vecInt", uid)
        self.fillValue = 0
        self.withConstant = True      # reference default: setDefault(withConstant, true)
        self.trackNulls = True        # TransmogrifierDefaults.TrackNulls

    def setFillWithConstant(self, value: int):  # noqa: N802
        self.fillValue, self.withConstant = int(value), True
        return self

    def setFillWithMode(self):  # noqa: N802
        self.withConstant = False
        return self

    def setTrackNulls(self, v: bool):  # noqa: N802
        self.trackNulls = bool(v)
        return self

    def onSetInput(self) -> None:  # noqa: N802
        self._metadata = _numeric_metadata(self, False)     # VectorizerDefaults.onSetInput

    def fit(self, batch: ColumnarBatch) -> "IntegralVectorizerModel":
        R.require_cuda()
        names = [f.name for f in self.inN]
        block = batch.numeric(names)
        if block.values.dtype != torch.int64:
            raise ValueError("Integral columns must be int64")
        dev = block if block.is_cuda else block.to("cuda")
        ncols, n = len(names), batch.n
        if self.trackNulls:
            self._metadata = _numeric_metadata(self, True)
        fills = [self.fillValue] * ncols if self.withConstant else column_modes(dev, n)
        model = IntegralVectorizerModel(self, fills, self.trackNulls)
        label_name = batch.response
        if label_name:
            model._fit_stats = self._sidecar(batch, dev, n, ncols, np.asarray(fills, dtype=np.float64), label_name)
            model._fit_key = batch_key(batch)
        return model

    def _sidecar(self, batch: ColumnarBatch, dev: NumericBlock, n: int, ncols: int, fills: np.ndarray,
                 label_name: str) -> VectorStats:
        """Statistics of the emitted columns from one scan of the raw columns (as RealVectorizer.fit does for Real)."""
        y = batch.label()
        y = y if y.is_cuda else y.cuda()
        NS = L.REAL_NSTAT
        nb = C.c_size_t()
        L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nb))
        scratch = R.scratch("real_fit", nb.value)
        d_shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
        d_part = torch.zeros((ncols + 1) * NS, dtype=torch.float64, device="cuda")
        L.call("tmog_integral_pilot_shift_dev", dev.values.data_ptr(), dev.ld, R.ptr(dev.valid), dev.ldv, y.data_ptr(),
               n, ncols, d_shifts.data_ptr(), R.stream())
        if D.world()[1] > 1:
            torch.distributed.broadcast(d_shifts, src=0)
        L.call("tmog_integral_fit_partials_dev", dev.values.data_ptr(), dev.ld, R.ptr(dev.valid), dev.ldv,
               y.data_ptr(), n, ncols, d_shifts.data_ptr(), scratch.data_ptr(), scratch.numel(), d_part.data_ptr(),
               R.stream())
        d_nc = None

        def count_nulls(ids):
            nonlocal d_nc
            if self.trackNulls:
                d_nc = torch.zeros((ncols, MAX_LABEL_CLASSES), dtype=torch.int64, device="cuda")
                L.call("tmog_null_class_counts_dev", R.ptr(dev.valid), dev.ldv, ids.data_ptr(), n, ncols,
                       MAX_LABEL_CLASSES, 0, d_nc.data_ptr(), R.stream())

        classes, counts, _ids = label_info_dev(batch, y, label_name, after_launch=count_nulls)
        want_nc = classes is not None and d_nc is not None
        partials, tabs = D.gather_fit_results(d_part, [d_nc[:, :len(classes)]] if want_nc else [])
        shifts = d_shifts.cpu().numpy()
        w = ncols * (2 if self.trackNulls else 1)
        mean, var, mn, mx, cov = (np.zeros(w) for _ in range(5))
        lab = np.zeros(5)
        L.call("tmog_real_finish_with_fills", partials.ctypes.data, shifts.ctypes.data, ncols, fills.ctypes.data,
               int(self.trackNulls), mean.ctypes.data, var.ctypes.data, mn.ctypes.data, mx.ctypes.data,
               cov.ctypes.data, lab.ctypes.data, None)
        class_sums = None
        if classes is not None:
            class_sums = np.full((len(classes), w), np.nan)
            if want_nc:
                class_sums[:, 1::2] = tabs[0].T
        return VectorStats(LabelInfo(label_name, int(lab[0]), lab, classes, counts), mean, var, mn, mx, cov, class_sums)


class IntegralVectorizerModel(OpModel):
    def __init__(self, parent: IntegralVectorizer, fillValues: List[int], trackNulls: bool):  # noqa: N803
        super().__init__(parent)
        self.fillValues = [int(v) for v in fillValues]
        self.trackNulls = trackNulls
        self._fit_stats: Optional[VectorStats] = None
        self._fit_key = None

    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        R.require_cuda()
        names = [f.name for f in self.inN]
        block = batch.numeric(names)
        fills, track = np.asarray(self.fillValues, dtype=np.float64), self.trackNulls       # .toDouble (:107,:111)
        part = LazyDensePart(len(names) * (2 if track else 1),
                             lambda keep: _transform_numeric(block, fills, track, 1 << 20, keep=keep))
        same = self._fit_key is not None and self._fit_key == batch_key(batch)
        return VectorColumn(self.getOutput().name, batch.n, self.getMetadata(), [part],
                            self._fit_stats if same else None)
