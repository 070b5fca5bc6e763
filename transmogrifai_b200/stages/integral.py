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


def _value_tables(dev: NumericBlock, n: int, sync_ranks: bool = True):
    """tmog_integral_mode_dev over a device block: -> (modes of the block, per column (values, counts) as numpy arrays:
    the block's value -> count map, INT64_MIN counted separately by the kernel).  The table grows until it holds every
    distinct value."""
    ncols = len(dev.names)
    size = D.world()[1] if sync_ranks else 1
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

    def tables():
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
        return local

    return modes, tables


def _merge_tables(per_source) -> List[int]:
    """Value -> count maps of several sources (ranks, chunks) added key-wise, per column (the Algebird map monoid of
    ModeSeqNullInt.merge, SequenceAggregators.scala:100-131) -> the modes, ties to the smaller value; 0 for no value."""
    out = []
    for per_col in zip(*per_source):
        k = np.concatenate([p[0] for p in per_col])
        w = np.concatenate([p[1] for p in per_col])
        if len(k) == 0:
            out.append(0)
            continue
        uk, inv = np.unique(k, return_inverse=True)
        tot = np.zeros(len(uk), dtype=np.int64)
        np.add.at(tot, inv, w)
        out.append(int(uk[np.lexsort((uk, -tot))[0]]))
    return out


def _add_tables(acc, local):
    """acc += local, column by column (value -> count maps as (values, counts) arrays)."""
    if acc is None:
        return local
    out = []
    for (k0, w0), (k1, w1) in zip(acc, local):
        uk, inv = np.unique(np.concatenate([k0, k1]), return_inverse=True)
        tot = np.zeros(len(uk), dtype=np.int64)
        np.add.at(tot, inv, np.concatenate([w0, w1]))
        out.append((uk, tot))
    return out


def column_modes(dev: NumericBlock, n: int) -> List[int]:
    """ModeSeqNullInt of every column of a device block over the rows of all ranks."""
    if len(dev.names) == 0:
        return []
    modes, tables = _value_tables(dev, n)
    if D.world()[1] == 1:
        return [int(v) for v in modes.cpu().numpy()]
    return _merge_tables(D.all_gather_object(tables()))


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
        return self._stats_from(partials, d_shifts.cpu().numpy(), fills, classes, counts, tabs[0] if want_nc else None,
                                label_name)

    # -- row chunks (streaming.py): partial sums under the first chunk's shifts, value -> count maps added chunk by chunk
    def stream_begin(self, label):
        if self.trackNulls:
            self._metadata = _numeric_metadata(self, True)
        return {"label": label, "shifts": None, "parts": [], "nc": None, "tables": None}

    def stream_update(self, st, chunk: ColumnarBatch) -> None:
        names = [f.name for f in self.inN]
        block = chunk.numeric(names)
        if block.values.dtype != torch.int64:
            raise ValueError("Integral columns must be int64")
        dev = chunk.on_device("numeric", names, block)
        ncols, n = len(names), chunk.n
        lab = st["label"]
        if not self.withConstant:            # the chunk's value -> count maps (one device-to-host copy per chunk)
            _modes, tables = _value_tables(dev, n, sync_ranks=False)
            st["tables"] = _add_tables(st["tables"], tables())
        if lab is None:
            return
        y = lab.device_label(chunk)
        if st["shifts"] is None:
            d_shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
            L.call("tmog_integral_pilot_shift_dev", dev.values.data_ptr(), dev.ld, R.ptr(dev.valid), dev.ldv,
                   y.data_ptr(), n, ncols, d_shifts.data_ptr(), R.stream())
            if D.world()[1] > 1:
                torch.distributed.broadcast(d_shifts, src=0)
            st["shifts"] = d_shifts
            nb = C.c_size_t()
            L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nb))
            st["scratch"] = torch.empty(max(nb.value, 256), dtype=torch.uint8, device="cuda")
        d_part = torch.empty((ncols + 1) * L.REAL_NSTAT, dtype=torch.float64, device="cuda")
        L.call("tmog_integral_fit_partials_dev", dev.values.data_ptr(), dev.ld, R.ptr(dev.valid), dev.ldv, y.data_ptr(),
               n, ncols, st["shifts"].data_ptr(), st["scratch"].data_ptr(), st["scratch"].numel(), d_part.data_ptr(),
               R.stream())
        st["parts"].append(d_part)
        if lab.classes is not None and self.trackNulls and dev.valid is not None:
            if st["nc"] is None:
                st["nc"] = torch.zeros((ncols, MAX_LABEL_CLASSES), dtype=torch.int64, device="cuda")
            ids = lab.class_ids(chunk)
            L.call("tmog_null_class_counts_dev", dev.valid.data_ptr(), dev.ldv, ids.data_ptr(), n, ncols,
                   MAX_LABEL_CLASSES, 1, st["nc"].data_ptr(), R.stream())

    def stream_finish(self, st) -> "IntegralVectorizerModel":
        ncols = len(self.inN)
        if self.withConstant:
            fills = [self.fillValue] * ncols
        else:
            local = st["tables"] if st["tables"] is not None else [(np.zeros(0, np.int64), np.zeros(0, np.int64))] * ncols
            fills = _merge_tables(D.all_gather_object(local) if D.world()[1] > 1 else [local])
        model = IntegralVectorizerModel(self, fills, self.trackNulls)
        lab = st["label"]
        if lab is None:
            return model
        if st["shifts"] is None:
            raise ValueError("requirement failed: Input dataset cannot be empty")
        host = torch.stack(st["parts"]).cpu().numpy()
        merged = np.ascontiguousarray(host[0])
        for k in range(1, host.shape[0]):
            other = np.ascontiguousarray(host[k])
            L.call("tmog_real_partials_merge", merged.ctypes.data, other.ctypes.data, ncols)
        classes = counts = None
        extra = []
        if lab.classes is not None:
            classes, counts = lab.classes, lab.class_counts
            if self.trackNulls:
                nc = st["nc"] if st["nc"] is not None else torch.zeros((ncols, MAX_LABEL_CLASSES), dtype=torch.int64,
                                                                       device="cuda")
                extra = [nc[:, :len(classes)]]
        partials, tabs = D.gather_fit_results(merged, extra)
        model._fit_stats = self._stats_from(partials, st["shifts"].cpu().numpy(), np.asarray(fills, dtype=np.float64),
                                            classes, counts, tabs[0] if extra else None, lab.name)
        return model

    def _stats_from(self, partials, shifts, fills, classes, counts, null_counts, label_name) -> VectorStats:
        ncols = len(self.inN)
        w = ncols * (2 if self.trackNulls else 1)
        mean, var, mn, mx, cov = (np.zeros(w) for _ in range(5))
        lab = np.zeros(5)
        L.call("tmog_real_finish_with_fills", partials.ctypes.data, shifts.ctypes.data, ncols, fills.ctypes.data,
               int(self.trackNulls), mean.ctypes.data, var.ctypes.data, mn.ctypes.data, mx.ctypes.data,
               cov.ctypes.data, lab.ctypes.data, None)
        class_sums = None
        if classes is not None:
            class_sums = np.full((len(classes), w), np.nan)
            if null_counts is not None:
                class_sums[:, 1::2] = null_counts.T
        return VectorStats(LabelInfo(label_name, int(lab[0]), lab, classes, counts), mean, var, mn, mx, cov, class_sums)


class IntegralVectorizerModel(OpModel):
    def __init__(self, parent: IntegralVectorizer, fillValues: List[int], trackNulls: bool):  # noqa: N803
        super().__init__(parent)
        self.fillValues = [int(v) for v in fillValues]
        self.trackNulls = trackNulls
        self._fit_stats: Optional[VectorStats] = None
        self._fit_key = None

    def stats_for(self, batch: ColumnarBatch) -> Optional[VectorStats]:
        """SanityChecker statistics of this model's columns over another batch (a row sample of the training data)."""
        if not batch.response:
            return None
        names = [f.name for f in self.inN]
        block = batch.numeric(names)
        dev = batch.on_device("numeric", names, block)
        return self.parent._sidecar(batch, dev, batch.n, len(names), np.asarray(self.fillValues, dtype=np.float64),
                                    batch.response)

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
