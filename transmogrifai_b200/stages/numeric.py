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
from .. import dist as D
from .. import features as F
from .. import runtime as R
from ..columnar import batch_key, ColumnarBatch, DensePart, LazyDensePart, NumericBlock, VectorColumn, round_up, to_host
from ..metadata import OpVectorMetadata, feature_column
from ..stats import LabelInfo, VectorStats
from .base import OpEstimator, OpModel, OpTransformer

MAX_LABEL_CLASSES = 100   # SanityChecker.scala:279 -- categorical iff distinct labels < min(100, 0.1 n)


def _numeric_metadata(stage, track_nulls: bool) -> OpVectorMetadata:
    key = (tuple(f.name for f in stage.inN), track_nulls)
    cache = stage.__dict__.setdefault("_meta_cache", {})
    if key not in cache:
        cache[key] = _build_numeric_metadata(stage, track_nulls)
    return cache[key]


def _build_numeric_metadata(stage, track_nulls: bool) -> OpVectorMetadata:
    cols = []
    for f in stage.inN:
        cols.append(feature_column(f.name, f.type_name))
        if track_nulls:
            cols.append(feature_column(f.name, f.type_name, is_null=True))
    return OpVectorMetadata(stage.getOutputFeatureName(), cols, stage.history())


def label_info_dev(batch: ColumnarBatch, y: torch.Tensor, name: str, after_launch=None):
    """Distinct label classes (global over all ranks), their global row counts, and per-row class ids of the local
    rows on the device.  Cached on the batch.  `after_launch(ids)` is called once the class ids are queued on the
    stream and before the host waits for the classes, so that kernels consuming the ids run back to back with them."""
    cache = batch.__dict__.setdefault("_label_cache", {})
    if name in cache:
        if after_launch is not None:
            after_launch(cache[name][2])
        return cache[name]
    n = batch.n
    distributed = D.world()[1] > 1
    nb = C.c_size_t()
    L.call("tmog_label_classes_scratch_bytes", C.byref(nb))
    scratch = R.scratch("label", nb.value)
    classes = torch.zeros(256, dtype=torch.float64, device="cuda")
    counts = torch.zeros(256, dtype=torch.int64, device="cuda")
    ncls = torch.zeros(1, dtype=torch.int32, device="cuda")
    ids = torch.empty(max(n, 1), dtype=torch.uint8, device="cuda")
    L.call("tmog_label_classes_dev", y.data_ptr(), n, MAX_LABEL_CLASSES, classes.data_ptr(), ncls.data_ptr(),
           ids.data_ptr(), counts.data_ptr(), scratch.data_ptr(), scratch.numel(), R.stream())
    if after_launch is not None:
        after_launch(ids)
    if distributed:
        # Every rank maps its rows against its own class list right away (the kernels that consume the ids are queued
        # behind it without waiting for the host); one exchange then tells whether all lists were the same -- the
        # usual case -- and only otherwise are the rows mapped again, against the union.
        union, same, g_counts = D.gather_label_classes(ncls, classes, counts, MAX_LABEL_CLASSES)
        if union is None or len(union) == 0:
            out = (None, None, ids)
        elif same:
            out = (union, g_counts, ids)
        else:
            classes[:len(union)] = torch.from_numpy(union).cuda()
            ncls.fill_(len(union))
            L.call("tmog_label_class_ids_dev", y.data_ptr(), n, classes.data_ptr(), ncls.data_ptr(),
                   MAX_LABEL_CLASSES, ids.data_ptr(), counts.data_ptr(), R.stream())
            if after_launch is not None:
                after_launch(ids)
            g_counts = counts[:len(union)].clone()
            torch.distributed.all_reduce(g_counts)
            out = (union, g_counts.cpu().numpy(), ids)
    else:
        # one device-to-host copy: {number of classes, classes, rows per class}
        host = torch.cat([ncls.double(), classes, counts.view(torch.float64)]).cpu().numpy()
        k = int(host[0])
        out = (host[1:1 + k].copy(), host[257:257 + k].view(np.int64).copy(), ids) if k > 0 else (None, None, ids)
    cache[name] = out
    return out


def label_moments_dev(batch: ColumnarBatch, y: torch.Tensor, name: str):
    """A continuous label (more distinct values than label classes are tracked): its count, mean, variance, min, max
    over all ranks, and the shift every label-sum kernel subtracts (rank 0's pilot mean).  Cached on the batch."""
    cache = batch.__dict__.setdefault("_label_cache", {})
    key = (name, "moments")
    if key in cache:
        return cache[key]
    n = batch.n
    nb = C.c_size_t()
    L.call("tmog_real_fit_scratch_bytes", 0, C.byref(nb))
    scratch = R.scratch("real_fit", nb.value)
    d_shift = torch.zeros(1, dtype=torch.float64, device="cuda")
    d_part = torch.zeros(L.REAL_NSTAT, dtype=torch.float64, device="cuda")
    ld = round_up(max(n, 1), 2)
    L.call("tmog_real_pilot_shift_dev", None, ld, None, 0, y.data_ptr(), n, 0, d_shift.data_ptr(), R.stream())
    if D.world()[1] > 1:
        torch.distributed.broadcast(d_shift, src=0)
    L.call("tmog_real_fit_partials_dev", None, ld, None, 0, y.data_ptr(), n, 0, d_shift.data_ptr(),
           scratch.data_ptr(), scratch.numel(), d_part.data_ptr(), R.stream())
    partials = D.allreduce_real_partials(d_part.cpu().numpy())
    shift = d_shift.cpu().numpy()
    lab = np.zeros(5)
    L.call("tmog_real_finish", partials.ctypes.data, shift.ctypes.data, 0, 1, 0.0, 0, None, None, None, None, None,
           None, lab.ctypes.data, None)
    out = (lab, float(shift[0]))
    cache[key] = out
    return out


def stats_from_sums_continuous(sv: np.ndarray, sv2: np.ndarray, nnz: np.ndarray, vmax: np.ndarray, vmin: np.ndarray,
                               wsum: np.ndarray, lab: np.ndarray, shift: float, label_name: str) -> VectorStats:
    """Statistics of non-negative integer-valued columns (indicators, term frequencies) against a continuous label:
    sv = sum v, sv2 = sum v^2, nnz = rows with v > 0, vmax / vmin = max / min of v over those rows,
    wsum = sum v (y - shift); cov(v, y) = wsum - sv (mean_y - shift)   (U/stats/OpStatistics.scala:86-90)."""
    n = int(lab[0])
    mean = sv / n if n else np.zeros_like(sv)
    var = np.maximum((sv2 - sv * sv / n) / (n - 1), 0.0) if n > 1 else np.zeros_like(sv)
    mx = np.where(nnz > 0, vmax, 0.0).astype(np.float64)
    mn = np.where(nnz >= n, vmin, 0.0).astype(np.float64)
    cov = wsum - sv * (lab[1] - shift)
    return VectorStats(LabelInfo(label_name, n, lab, None, None), mean, var, mn, mx, cov, None)


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
        if self.trackNulls:
            self._metadata = _numeric_metadata(self, True)
        partials, shifts, classes, counts, null_counts, label_name = self._scan_batch(batch)
        model = self._finish(partials, shifts, classes, counts, null_counts, label_name)
        model._fit_key = batch_key(batch)
        return model

    def _scan_batch(self, batch: ColumnarBatch):
        """The fused scan of one batch (all ranks): merged partial sums, their shifts, label classes and the per-class
        null counts."""
        names = [f.name for f in self.inN]
        block = batch.numeric(names)
        ncols, n = len(names), batch.n
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
            if D.world()[1] > 1:        # partials only merge when every rank uses rank 0's shifts
                torch.distributed.broadcast(d_shifts, src=0)
            L.call("tmog_real_fit_partials_dev", block.values.data_ptr(), block.ld, R.ptr(block.valid), block.ldv,
                   y.data_ptr(), n, ncols, d_shifts.data_ptr(), scratch.data_ptr(), scratch.numel(),
                   d_part.data_ptr(), R.stream())
            if label_name:
                d_nc = None

                def count_nulls(ids):       # queued right behind the class-id kernel, counted against 255 classes
                    nonlocal d_nc
                    if self.trackNulls:
                        d_nc = torch.zeros((ncols, MAX_LABEL_CLASSES), dtype=torch.int64, device="cuda")
                        L.call("tmog_null_class_counts_dev", R.ptr(block.valid), block.ldv, ids.data_ptr(), n, ncols,
                               MAX_LABEL_CLASSES, 0, d_nc.data_ptr(), R.stream())

                classes, counts, ids = label_info_dev(batch, y, label_name, after_launch=count_nulls)
            # partial sums and null counts of all ranks in one exchange and one device-to-host copy
            want_nc = label_name and classes is not None and d_nc is not None
            # (the shifts ride along as raw bits; every rank holds rank 0's, so their "sum" is not used)
            extra = [d_shifts.view(torch.int64)] + ([d_nc[:, :len(classes)]] if want_nc else [])
            if D.world()[1] == 1:
                partials, tabs = D.gather_fit_results(d_part, extra)
                shifts = tabs[0].view(np.float64)
                tabs = tabs[1:]
            else:
                partials, tabs = D.gather_fit_results(d_part, extra[1:])
                shifts = d_shifts.cpu().numpy()
            if want_nc:
                null_counts = tabs[0]
        else:
            y = batch.label() if label_name else torch.zeros(max(n, 2), dtype=torch.float64)
            cls = np.zeros(256)
            cnt = np.zeros(256, dtype=np.uint64)
            ncls = C.c_int(0)
            nc = np.zeros((ncols, MAX_LABEL_CLASSES), dtype=np.uint64).reshape(-1)
            shifts_in = None
            if D.world()[1] > 1:
                shifts_in = D.broadcast(_host_pilot_shifts(block, y, n, ncols))
            L.call("tmog_real_fit_host", R.ctx(), block.values.data_ptr(), block.ld, R.ptr(block.valid), block.ldv,
                   y.data_ptr(), n, ncols, self.chunkRows, 1, None if shifts_in is None else shifts_in.ctypes.data,
                   shifts.ctypes.data, partials.ctypes.data,
                   MAX_LABEL_CLASSES if label_name else 0, cls.ctypes.data, C.byref(ncls), cnt.ctypes.data,
                   nc.ctypes.data if self.trackNulls else None)
            if label_name and ncls.value > 0:
                k = ncls.value
                classes, counts = cls[:k].copy(), cnt[:k].astype(np.int64)
                if self.trackNulls:
                    null_counts = nc[:ncols * k].reshape(ncols, k).astype(np.int64)
            if D.world()[1] > 1:
                partials = D.allreduce_real_partials(partials)
                union = D.union_classes(classes if label_name else np.zeros(0), MAX_LABEL_CLASSES)
                if label_name and union is not None and len(union) > 0:
                    ubits = union.view(np.int64)
                    pos = np.array([int(np.nonzero(ubits == b)[0][0]) for b in classes.view(np.int64)],
                                   dtype=np.int64) if classes is not None else np.zeros(0, dtype=np.int64)
                    g_counts = np.zeros(len(union), dtype=np.int64)
                    g_null = np.zeros((ncols, len(union)), dtype=np.int64)
                    if classes is not None:
                        g_counts[pos] = counts
                        if null_counts is not None:
                            g_null[:, pos] = null_counts
                    classes, counts = union, D.allreduce_sum(g_counts)
                    null_counts = D.allreduce_sum(g_null) if self.trackNulls else None
                else:
                    classes = counts = null_counts = None
        return partials, shifts, classes, counts, null_counts, label_name

    def _finish(self, partials, shifts, classes, counts, null_counts, label_name) -> "RealVectorizerModel":
        """Fill values and the statistics sidecar from the merged partial sums (host)."""
        ncols = len(self.inN)
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
            stats = VectorStats(LabelInfo(label_name, int(lab[0]), lab, classes, counts), mean, var, mn, mx, cov,
                                class_sums)
        model = RealVectorizerModel(self, fills.tolist(), self.trackNulls)
        model._fit_stats = stats
        return model

    # -- row chunks (streaming.py): partial sums of every chunk under the first chunk's shifts, merged in chunk order
    def stream_begin(self, label: Optional["StreamLabel"]):
        if self.trackNulls:
            self._metadata = _numeric_metadata(self, True)
        return {"label": label, "shifts": None, "parts": [], "nc": None}

    def stream_update(self, st, chunk: ColumnarBatch) -> None:
        names = [f.name for f in self.inN]
        block = chunk.numeric(names)
        dev = chunk.on_device("numeric", names, block)
        ncols, n = len(names), chunk.n
        lab = st["label"]
        y = lab.device_label(chunk) if lab is not None else torch.zeros(max(n, 2), dtype=torch.float64, device="cuda")
        if st["shifts"] is None:
            d_shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
            L.call("tmog_real_pilot_shift_dev", dev.values.data_ptr(), dev.ld, R.ptr(dev.valid), dev.ldv,
                   y.data_ptr(), n, ncols, d_shifts.data_ptr(), R.stream())
            if D.world()[1] > 1:
                torch.distributed.broadcast(d_shifts, src=0)
            st["shifts"] = d_shifts
            nb = C.c_size_t()
            L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nb))
            st["scratch"] = torch.empty(max(nb.value, 256), dtype=torch.uint8, device="cuda")
        d_part = torch.empty((ncols + 1) * L.REAL_NSTAT, dtype=torch.float64, device="cuda")
        L.call("tmog_real_fit_partials_dev", dev.values.data_ptr(), dev.ld, R.ptr(dev.valid), dev.ldv, y.data_ptr(),
               n, ncols, st["shifts"].data_ptr(), st["scratch"].data_ptr(), st["scratch"].numel(), d_part.data_ptr(),
               R.stream())
        st["parts"].append(d_part)
        if lab is not None and lab.classes is not None and self.trackNulls and dev.valid is not None:
            if st["nc"] is None:
                st["nc"] = torch.zeros((ncols, MAX_LABEL_CLASSES), dtype=torch.int64, device="cuda")
            ids = lab.class_ids(chunk)
            L.call("tmog_null_class_counts_dev", dev.valid.data_ptr(), dev.ldv, ids.data_ptr(), n, ncols,
                   MAX_LABEL_CLASSES, 1, st["nc"].data_ptr(), R.stream())

    def stream_finish(self, st) -> "RealVectorizerModel":
        lab = st["label"]
        ncols = len(self.inN)
        if st["shifts"] is None:
            raise ValueError("requirement failed: Input dataset cannot be empty")
        host = torch.stack(st["parts"]).cpu().numpy()                  # one copy for all chunks
        merged = np.ascontiguousarray(host[0])
        for k in range(1, host.shape[0]):
            other = np.ascontiguousarray(host[k])
            L.call("tmog_real_partials_merge", merged.ctypes.data, other.ctypes.data, ncols)
        classes = counts = null_counts = None
        extra = []
        if lab is not None and lab.classes is not None:
            classes, counts = lab.classes, lab.class_counts
            if self.trackNulls:
                nc = st["nc"] if st["nc"] is not None else torch.zeros((ncols, MAX_LABEL_CLASSES), dtype=torch.int64,
                                                                       device="cuda")
                extra = [nc[:, :len(classes)]]
        partials, tabs = D.gather_fit_results(merged, extra)
        if extra:
            null_counts = tabs[0]
        shifts = st["shifts"].cpu().numpy()
        return self._finish(partials, shifts, classes, counts, null_counts, lab.name if lab is not None else None)


def _host_pilot_shifts(block: NumericBlock, y: torch.Tensor, n: int, ncols: int) -> np.ndarray:
    """Pilot shifts of a host block: upload its first rows and run the pilot kernel (rank 0's result is broadcast)."""
    rows = min(n, 4096)
    ld = round_up(max(rows, 1), 2)
    v = torch.zeros((max(ncols, 1), ld), dtype=torch.float64)
    v[:ncols, :rows] = block.values[:, :rows]
    dv = v.cuda()
    dm = None if block.valid is None else block.valid[:, :(rows + 63) // 64].contiguous().cuda()
    dy = y[:max(rows, 1)].cuda()
    sh = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
    L.call("tmog_real_pilot_shift_dev", dv.data_ptr(), ld, R.ptr(dm), 0 if dm is None else dm.shape[1], dy.data_ptr(),
           rows, ncols, sh.data_ptr(), R.stream())
    return sh.cpu().numpy()


class RealVectorizerModel(OpModel):
    def __init__(self, parent: RealVectorizer, fillValues: List[float], trackNulls: bool):  # noqa: N803
        super().__init__(parent)
        self.fillValues = list(fillValues)
        self.trackNulls = trackNulls
        self.chunkRows = parent.chunkRows
        self._fit_stats: Optional[VectorStats] = None
        self._fit_key = None

    def stats_for(self, batch: ColumnarBatch) -> Optional[VectorStats]:
        """SanityChecker statistics of this model's columns over another batch (a row sample of the training data):
        the same fused scan, finished with the model's fill values instead of the batch's own means."""
        if not batch.response:
            return None
        partials, shifts, classes, counts, null_counts, label_name = self.parent._scan_batch(batch)
        ncols = len(self.inN)
        w = ncols * (2 if self.trackNulls else 1)
        fills = np.asarray(self.fillValues, dtype=np.float64)
        mean, var, mn, mx, cov = (np.zeros(w) for _ in range(5))
        lab = np.zeros(5)
        L.call("tmog_real_finish_with_fills", partials.ctypes.data, shifts.ctypes.data, ncols, fills.ctypes.data,
               int(self.trackNulls), mean.ctypes.data, var.ctypes.data, mn.ctypes.data, mx.ctypes.data, cov.ctypes.data,
               lab.ctypes.data, None)
        class_sums = None
        if classes is not None:
            class_sums = np.full((len(classes), w), np.nan)
            if self.trackNulls and null_counts is not None:
                class_sums[:, 1::2] = null_counts.T
        return VectorStats(LabelInfo(label_name, int(lab[0]), lab, classes, counts), mean, var, mn, mx, cov, class_sums)

    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        R.require_cuda()
        names = [f.name for f in self.inN]
        block = _stream_block(batch, names)
        fills, track, chunk = np.asarray(self.fillValues), self.trackNulls, self.chunkRows
        part = LazyDensePart(len(names) * (2 if track else 1),
                             lambda keep: _transform_numeric(block, fills, track, chunk, keep=keep))
        same = self._fit_key is not None and self._fit_key == batch_key(batch)
        return VectorColumn(self.getOutput().name, batch.n, self.getMetadata(), [part],
                            self._fit_stats if same else None)


def _stream_block(batch: ColumnarBatch, names) -> NumericBlock:
    """The numeric block of these columns; for a row chunk of a streamed dataset, its device copy (the chunk's output
    stays on the device and the caller's sink moves it)."""
    block = batch.numeric(names)
    if batch.__dict__.get("_stream") is not None and not block.is_cuda:
        block = batch.on_device("numeric", names, block)
    return block


_DEV_CONST: dict = {}


def _dev_const(kind: str, arr: np.ndarray) -> torch.Tensor:
    """Small constant arrays of a transform (keep bit masks, fill values) on the device, uploaded once per content:
    a model transforms many row chunks with the same ones, and a pageable upload per call stalls the stream."""
    key = (kind, torch.cuda.current_device(), arr.dtype.str, arr.tobytes())
    t = _DEV_CONST.get(key)
    if t is None:
        if len(_DEV_CONST) > 64:
            _DEV_CONST.clear()
        t = _DEV_CONST[key] = torch.from_numpy(arr.copy()).cuda()
    return t


def _transform_numeric(block: NumericBlock, fills: Optional[np.ndarray], track_nulls: bool, chunk_rows: int,
                       fill_value: bool = False, keep: Optional[List[int]] = None) -> torch.Tensor:
    """Run the transform kernel.  keep = the output columns wanted, in order (None = all): the kernel then writes
    only those (SanityChecker's indicesToKeep fused into the vectorizer's transform)."""
    n, ncols = block.n, len(block.names)
    w_full = ncols * (2 if track_nulls else 1)
    binary = block.values.dtype == torch.uint8
    integral = block.values.dtype == torch.int64
    kp = None
    w = w_full
    if keep is not None:
        kp = np.asarray(keep, dtype=np.int32)
        if len(kp) > 1 and not (np.diff(kp) > 0).all():
            raise ValueError("kept columns must be ascending")
        w = len(kp)
    if block.is_cuda:
        out = torch.empty((n, max(w, 1)), dtype=torch.float64, device="cuda")   # dense rows: no unwritten padding
        d_keep = None
        if kp is not None:
            bits = np.zeros(round_up(w_full, 64), dtype=np.uint8)
            bits[kp] = 1
            d_keep = _dev_const("keep", np.packbits(bits, bitorder="little").view("<i8"))
        if integral:
            d_fills = None if fills is None else _dev_const("fills", np.asarray(fills, dtype=np.float64))
            if kp is None:
                L.call("tmog_integral_transform_dev", block.values.data_ptr(), block.ld, R.ptr(block.valid), block.ldv,
                       n, ncols, R.ptr(d_fills), int(track_nulls), out.data_ptr(), out.shape[1], 0, R.stream())
            else:
                L.call("tmog_integral_transform_select_dev", block.values.data_ptr(), block.ld, R.ptr(block.valid),
                       block.ldv, n, ncols, R.ptr(d_fills), int(track_nulls), d_keep.data_ptr(), w, out.data_ptr(),
                       out.shape[1], 0, R.stream())
        elif binary:
            if kp is None:
                L.call("tmog_binary_transform_dev", block.values.data_ptr(), block.ld, R.ptr(block.valid), block.ldv,
                       n, ncols, int(fill_value), int(track_nulls), out.data_ptr(), out.shape[1], 0, R.stream())
            else:
                L.call("tmog_binary_transform_select_dev", block.values.data_ptr(), block.ld, R.ptr(block.valid),
                       block.ldv, n, ncols, int(fill_value), int(track_nulls), d_keep.data_ptr(), w, out.data_ptr(),
                       out.shape[1], 0, R.stream())
        else:
            d_fills = None if fills is None else _dev_const("fills", np.asarray(fills, dtype=np.float64))
            if kp is None:
                L.call("tmog_real_transform_dev", block.values.data_ptr(), block.ld, R.ptr(block.valid), block.ldv,
                       n, ncols, R.ptr(d_fills), int(track_nulls), out.data_ptr(), out.shape[1], 0, R.stream())
            else:
                L.call("tmog_real_transform_select_dev", block.values.data_ptr(), block.ld, R.ptr(block.valid),
                       block.ldv, n, ncols, R.ptr(d_fills), int(track_nulls), d_keep.data_ptr(), w, out.data_ptr(),
                       out.shape[1], 0, R.stream())
        return out[:, :w]
    if binary or integral:
        dev = block.to("cuda")
        return to_host([_transform_numeric(dev, fills, track_nulls, chunk_rows, fill_value, keep)])[0]   # pinned
    out = torch.empty((n, w), dtype=torch.float64, pin_memory=True)
    L.call("tmog_real_transform_host", R.ctx(), block.values.data_ptr(), block.ld, R.ptr(block.valid), block.ldv, n,
           ncols, None if fills is None else fills.ctypes.data, int(track_nulls), chunk_rows,
           None if kp is None else kp.ctypes.data, w, out.data_ptr(), w, 0)
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
        block = _stream_block(batch, [f.name for f in self.inN])
        if block.valid is not None:
            raise ValueError("None.get: a RealNN feature holds a null")
        chunk = self.chunkRows
        part = LazyDensePart(len(block.names), lambda keep: _transform_numeric(block, None, False, chunk, keep=keep))
        stats = _realnn_stats(batch, block) if batch.response else None
        return VectorColumn(self.getOutput().name, batch.n, self.getMetadata(), [part], stats)

    # -- row chunks: the statistics sidecar of a dataset that arrives in chunks (the transformer itself has no fit)
    def stream_begin(self, label):
        return {"label": label, "shifts": None, "parts": []}

    def stream_update(self, st, chunk: ColumnarBatch) -> None:
        if st["label"] is None:
            return
        block = chunk.numeric([f.name for f in self.inN])
        if block.valid is not None:
            raise ValueError("None.get: a RealNN feature holds a null")
        _realnn_scan(st, chunk, block, st["label"].device_label(chunk))

    def stream_finish(self, st):
        lab = st["label"]
        if lab is None or not st["parts"]:
            return None
        return _realnn_finish(st, len(self.inN), lab.name, lab.classes, lab.class_counts)


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
        block = _stream_block(batch, [f.name for f in self.inN])
        track, chunk, fv = self.trackNulls, self.chunkRows, self.fillValue
        part = LazyDensePart(len(block.names) * (2 if track else 1),
                             lambda keep: _transform_numeric(block, None, track, chunk, fv, keep=keep))
        stats = _binary_stats(batch, block, track, fv) if batch.response else None
        return VectorColumn(self.getOutput().name, batch.n, self.getMetadata(), [part], stats)

    # -- row chunks
    def stream_begin(self, label):
        return {"label": label, "acc": {}, "n": 0}

    def stream_update(self, st, chunk: ColumnarBatch) -> None:
        lab = st["label"]
        if lab is None:
            return
        block = chunk.numeric([f.name for f in self.inN])
        ids = lab.class_ids(chunk) if lab.classes is not None else None
        _binary_scan(st["acc"], chunk, block, lab.device_label(chunk), lab.classes, ids, lab.continuous())
        st["n"] += chunk.n

    def stream_finish(self, st):
        lab = st["label"]
        if lab is None or not st["acc"]:
            return None
        return _binary_finish(st["acc"], len(self.inN), self.trackNulls, self.fillValue, lab.name, lab.classes,
                              lab.class_counts, lab.continuous(), st["n"])


def _realnn_stats(batch: ColumnarBatch, block: NumericBlock) -> Optional[VectorStats]:
    """Statistics sidecar of a RealNN block: the fused fit scan without nulls (the vectorizer itself has no fit)."""
    st = {"shifts": None, "parts": []}
    _realnn_scan(st, batch, block, batch.label())
    classes, counts, _ids = label_info_dev(batch, st["y"], batch.response)
    return _realnn_finish(st, len(block.names), batch.response, classes, counts)


def _realnn_scan(st: dict, batch: ColumnarBatch, block: NumericBlock, y: torch.Tensor) -> None:
    dev = block if block.is_cuda else batch.on_device("numeric", block.names, block)
    y = y if y.is_cuda else y.cuda()
    st["y"] = y
    n, ncols = batch.n, len(block.names)
    if st["shifts"] is None:
        nb = C.c_size_t()
        L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nb))
        st["scratch"] = R.scratch("real_fit", nb.value)
        st["shifts"] = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
        L.call("tmog_real_pilot_shift_dev", dev.values.data_ptr(), dev.ld, None, 0, y.data_ptr(), n, ncols,
               st["shifts"].data_ptr(), R.stream())
        if D.world()[1] > 1:
            torch.distributed.broadcast(st["shifts"], src=0)
    d_part = torch.zeros((ncols + 1) * L.REAL_NSTAT, dtype=torch.float64, device="cuda")
    L.call("tmog_real_fit_partials_dev", dev.values.data_ptr(), dev.ld, None, 0, y.data_ptr(), n, ncols,
           st["shifts"].data_ptr(), st["scratch"].data_ptr(), st["scratch"].numel(), d_part.data_ptr(), R.stream())
    st["parts"].append(d_part)


def _realnn_finish(st: dict, ncols: int, label_name: str, classes, counts) -> VectorStats:
    host = torch.stack(st["parts"]).cpu().numpy()
    merged = np.ascontiguousarray(host[0])
    for k in range(1, host.shape[0]):
        other = np.ascontiguousarray(host[k])
        L.call("tmog_real_partials_merge", merged.ctypes.data, other.ctypes.data, ncols)
    partials = D.allreduce_real_partials(merged)
    shifts = st["shifts"].cpu().numpy()
    mean, var, mn, mx, cov = (np.zeros(ncols) for _ in range(5))
    lab = np.zeros(5)
    L.call("tmog_real_finish", partials.ctypes.data, shifts.ctypes.data, ncols, 1, 0.0, 0, None, mean.ctypes.data,
           var.ctypes.data, mn.ctypes.data, mx.ctypes.data, cov.ctypes.data, lab.ctypes.data, None)
    cs = None if classes is None else np.full((len(classes), ncols), np.nan)
    return VectorStats(LabelInfo(label_name, int(lab[0]), lab, classes, counts), mean, var, mn, mx, cov, cs)


def _binary_stats(batch: ColumnarBatch, block: NumericBlock, track_nulls: bool, fill_value: bool
                  ) -> Optional[VectorStats]:
    """Statistics sidecar of BinaryVectorizer's [value, isNull] columns from count[{null, false, true}][label class]
    (a continuous label adds the label sums per {null, false, true})."""
    label_name = batch.response
    y = batch.label()
    y = y if y.is_cuda else y.cuda()
    classes, class_counts, ids = label_info_dev(batch, y, label_name)
    cont = None if classes is not None else label_moments_dev(batch, y, label_name)
    st: dict = {}
    _binary_scan(st, batch, block, y, classes, ids, cont)
    return _binary_finish(st, len(block.names), track_nulls, fill_value, label_name, classes, class_counts, cont,
                          batch.n)


def _binary_scan(st: dict, batch: ColumnarBatch, block: NumericBlock, y: torch.Tensor, classes, ids, cont) -> None:
    """One batch / row chunk into the accumulators of `st` (count tables widened to int64 per call)."""
    dev = block if block.is_cuda else batch.on_device("numeric", block.names, block)
    n, ncols, ncls = batch.n, len(block.names), (len(classes) if classes is not None else 1)
    ldc = round_up(max(n, 1), 4)
    codes = torch.empty((ncols, ldc), dtype=torch.int32, device="cuda")
    L.call("tmog_binary_codes_dev", dev.values.data_ptr(), dev.ld, R.ptr(dev.valid), dev.ldv, n, ncols,
           codes.data_ptr(), ldc, R.stream())
    if "desc" not in st:
        desc = np.zeros((ncols, 4), dtype=np.int64)
        for c in range(ncols):
            desc[c] = [3, 0, c * 3 * ncls, 2]          # ids: 0 null, 1 false, 2 true
        st["desc"] = torch.from_numpy(desc).cuda()
        st["acc"] = torch.zeros(ncols * 3 * ncls, dtype=torch.int64, device="cuda")
        st["bad"] = torch.zeros(1, dtype=torch.int32, device="cuda")
        st["ws"] = torch.zeros(ncols * 3, dtype=torch.float64, device="cuda") if cont is not None else None
        st["woff"] = torch.arange(0, ncols * 3, 3, dtype=torch.int64, device="cuda")
    d_counts = torch.zeros(ncols * 3 * ncls, dtype=torch.int32, device="cuda")
    L.call("tmog_pivot_count_dev", codes.data_ptr(), ldc, n, ncols, ids.data_ptr() if classes is not None else None,
           ncls, st["desc"].data_ptr(), None, 3, d_counts.data_ptr(), st["bad"].data_ptr(), R.stream())
    st["acc"] += d_counts.to(torch.int64) & 0xFFFFFFFF
    if cont is not None:                             # continuous label: label sums per {null, false, true}
        L.call("tmog_pivot_label_sum_dev", codes.data_ptr(), ldc, n, ncols, st["desc"].data_ptr(), None,
               st["woff"].data_ptr(), y.data_ptr(), cont[1], st["ws"].data_ptr(), R.stream())


def _binary_finish(st: dict, ncols: int, track_nulls: bool, fill_value: bool, label_name: str, classes,
                   class_counts, cont, n_local: int) -> VectorStats:
    from .pivot import _stats_from_class_counts
    ncls = len(classes) if classes is not None else 1
    cnt = D.allreduce_sum(st["acc"].cpu().numpy()).reshape(ncols, 3, ncls)
    if classes is None:
        lab, shift = cont
        ws = D.allreduce_sum(st["ws"].cpu().numpy()).reshape(ncols, 3)
        c1, w1 = [], []
        for c in range(ncols):
            c1.append(cnt[c, 2, 0] + (cnt[c, 0, 0] if fill_value else 0))
            w1.append(ws[c, 2] + (ws[c, 0] if fill_value else 0.0))
            if track_nulls:
                c1.append(cnt[c, 0, 0])
                w1.append(ws[c, 0])
        sv = np.asarray(c1, dtype=np.float64)
        ones = np.ones_like(sv)
        return stats_from_sums_continuous(sv, sv, sv, ones, ones, np.asarray(w1), lab, shift, label_name)
    n_global = int(np.asarray(class_counts).sum())
    rows = []
    for c in range(ncols):
        ones = cnt[c, 2] + (cnt[c, 0] if fill_value else 0)
        rows.append(ones)
        if track_nulls:
            rows.append(cnt[c, 0])
    stt = _stats_from_class_counts(np.stack(rows).astype(np.float64), classes, class_counts, n_global, label_name)
    if stt.class_sums is not None:                    # only the null indicators are categorical columns
        keep = np.zeros(stt.width, dtype=bool)
        if track_nulls:
            keep[1::2] = True
        stt.class_sums[:, ~keep] = np.nan
    return stt
