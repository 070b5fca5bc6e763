"""OpTextPivotVectorizer (PickList / Text pivot: top-K one-hot + OTHER + null indicator) on the GPU.

Reference: core/src/main/scala/com/salesforce/op/stages/impl/feature/OpOneHotVectorizer.scala:61-124 (fit),
:203-247 (OpTextPivotVectorizer + model), :388-399 (metadata), :416-437 (pivotFn); params PivotParams /
MinSupportParam / CleanTextFun in Transmogrifier.scala:484-575; utils/.../text/TextUtils.scala:39-47.

Per-row work is integer only (dictionary codes); strings are touched once per dictionary entry on the host
(tmog_clean_strings / tmog_dedupe_strings / tmog_pivot_topk).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .. import _lib as L
from .. import dist as D
from .. import features as F
from .. import runtime as R
from ..columnar import (ColumnarBatch, DictBlock, LazyOneHotPart, OneHotPart, VectorColumn, decode_strings,
                        encode_strings, round_up)
from ..metadata import NullString, OpVectorColumnMetadata, OpVectorMetadata, OtherString
from ..stats import LabelInfo, VectorStats
from .base import OpEstimator, OpModel
from .numeric import label_info_dev


def clean_and_group(strings: Sequence[str], clean: bool) -> Tuple[np.ndarray, List[str]]:
    """Raw dictionary -> (id of the cleaned value per raw entry, the distinct cleaned values in id order)."""
    data, offs = encode_strings(strings)
    n = len(strings)
    if clean and n:
        out = np.zeros(max(len(data), 1), dtype=np.uint8)
        ooffs = np.zeros(n + 1, dtype=np.int64)
        L.call("tmog_clean_strings", data.ctypes.data, offs.ctypes.data, n, out.ctypes.data, len(out),
               ooffs.ctypes.data)
        data, offs = out, ooffs
    ids = np.zeros(max(n, 1), dtype=np.int32)
    nd = np.zeros(1, dtype=np.int64)
    L.call("tmog_dedupe_strings", data.ctypes.data, offs.ctypes.data, n, ids.ctypes.data, nd.ctypes.data)
    ids = ids[:n]
    cleaned = decode_strings(data, offs)
    distinct: List[str] = [""] * int(nd[0])
    for i, k in enumerate(ids):
        distinct[k] = cleaned[i]
    return ids, distinct


def top_k(values: Sequence[str], counts: np.ndarray, k: int, min_support: int) -> List[int]:
    data, offs = encode_strings(values)
    out = np.zeros(max(k, 1), dtype=np.int32)
    n_out = np.zeros(1, dtype=np.int32)
    cnt = np.ascontiguousarray(counts, dtype=np.int64)
    L.call("tmog_pivot_topk", data.ctypes.data, offs.ctypes.data, cnt.ctypes.data, len(values), int(k),
           int(min_support), out.ctypes.data, n_out.ctypes.data)
    return out[:int(n_out[0])].tolist()


class _Plan:
    """Device tables of one fit / transform call over a DictBlock."""

    def __init__(self, block: DictBlock, clean: bool):
        self.ids: List[np.ndarray] = []
        self.values: List[List[str]] = []
        for d in block.dicts:
            ids, distinct = clean_and_group(d, clean)
            self.ids.append(ids)
            self.values.append(distinct)
        ncols = len(block.names)
        self.ncodes = np.array([len(d) for d in block.dicts], dtype=np.int64)
        self.nids = self.ncodes + 1                                                    # raw codes + the null id 0
        self.slot_offset = np.concatenate([[0], np.cumsum(self.ncodes)[:-1]]).astype(np.int64) if ncols else \
            np.zeros(0, dtype=np.int64)

    def desc(self, nclasses: int) -> np.ndarray:
        count_offset = np.concatenate([[0], np.cumsum(self.nids * nclasses)[:-1]]).astype(np.int64)
        self.count_offset = count_offset
        return np.ascontiguousarray(np.stack([self.nids, self.slot_offset, count_offset, self.ncodes], axis=1))


class OpTextPivotVectorizer(OpEstimator):
    input_types = (F.PICKLIST, F.TEXT, F.TEXTAREA)

    def __init__(self, uid: Optional[str] = None):
        super().__init__("pivotText", uid)
        self.topK = 20                  # TransmogrifierDefaults.TopK
        self.minSupport = 10            # TransmogrifierDefaults.MinSupport
        self.cleanText = True
        self.trackNulls = True
        self.unseenName = OtherString
        self.maxPctCardinality = 1.0

    def setTopK(self, v: int):  # noqa: N802
        if v <= 0:
            raise ValueError(f"requirement failed: topK given invalid value {v}")   # OpSetVectorizerTest.scala:164-167
        self.topK = int(v)
        return self

    def setMinSupport(self, v: int):  # noqa: N802
        if v < 0:
            raise ValueError(f"requirement failed: minSupport given invalid value {v}")
        self.minSupport = int(v)
        return self

    def setCleanText(self, v: bool):  # noqa: N802
        self.cleanText = bool(v)
        return self

    def setTrackNulls(self, v: bool):  # noqa: N802
        self.trackNulls = bool(v)
        return self

    def setUnseenName(self, v: str):  # noqa: N802
        self.unseenName = v
        return self

    def setMaxPctCardinality(self, v: float):  # noqa: N802
        if not (0.0 < v <= 1.0):
            raise ValueError(f"requirement failed: maxPctCardinality given invalid value {v}")
        if v != 1.0:
            raise NotImplementedError("maxPctCardinality < 1 (HyperLogLog cardinality filter) is not on the GPU path")
        self.maxPctCardinality = float(v)
        return self

    def _metadata_for(self, tops: List[List[str]]) -> OpVectorMetadata:
        cols = []
        for f, top in zip(self.inN, tops):
            vals = list(top) + [self.unseenName] + ([NullString] if self.trackNulls else [])
            cols += [OpVectorColumnMetadata((f.name,), (f.type_name,), grouping=f.name, indicatorValue=v)
                     for v in vals]
        return OpVectorMetadata(self.getOutputFeatureName(), cols, self.history())

    def fit(self, batch: ColumnarBatch) -> "OpTextPivotVectorizerModel":
        R.require_cuda()
        names = [f.name for f in self.inN]
        block = batch.strings(names)
        dev = block if block.is_cuda else block.to("cuda")
        n, ncols = batch.n, len(names)
        plan = _Plan(block, self.cleanText)
        label_name = batch.response
        classes = counts_per_class = ids = None
        if label_name:
            y = batch.label()
            y = y if y.is_cuda else y.cuda()
            classes, counts_per_class, ids = label_info_dev(batch, y, label_name)
        ncls = len(classes) if classes is not None else 1
        desc = plan.desc(ncls)
        total_bins = int((plan.nids * ncls).sum())
        d_desc = torch.from_numpy(desc).cuda()
        d_counts = torch.zeros(max(total_bins, 1), dtype=torch.int32, device="cuda")     # uint32 bins
        d_bad = torch.zeros(1, dtype=torch.int32, device="cuda")
        L.call("tmog_pivot_count_dev", dev.codes.data_ptr(), dev.ld, n, ncols,
               ids.data_ptr() if classes is not None else None, ncls, d_desc.data_ptr(), None,
               int(plan.nids.max()) if ncols else 1, d_counts.data_ptr(), d_bad.data_ptr(), R.stream())
        counts = D.allreduce_sum(d_counts.cpu().numpy().astype(np.int64) & 0xFFFFFFFF)
        if int(d_bad.item()):
            raise ValueError("a dictionary code outside its column's dictionary was found in the batch")
        n_global = int(D.allreduce_sum(np.array([n], dtype=np.int64))[0])
        tops: List[List[str]] = []
        col_stats = []
        for c in range(ncols):
            k = int(plan.nids[c])
            raw = counts[plan.count_offset[c]:plan.count_offset[c] + k * ncls].reshape(k, ncls)
            # bins are per RAW code (row 0 = null): merge the codes whose cleaned values coincide
            tab = np.zeros((len(plan.values[c]), ncls), dtype=np.int64)
            if len(plan.ids[c]):
                np.add.at(tab, plan.ids[c], raw[1:])
            totals = tab.sum(axis=1)
            chosen = top_k(plan.values[c], totals, self.topK, self.minSupport)
            tops.append([plan.values[c][i] for i in chosen])
            rest = np.ones(len(tab), dtype=bool)
            rest[chosen] = False
            rows = [tab[i] for i in chosen] + [tab[rest].sum(axis=0)]
            if self.trackNulls:
                rows.append(raw[0])
            col_stats.append(np.stack(rows).astype(np.float64))            # [emitted columns, classes]
        self._metadata = self._metadata_for(tops)
        stats = None
        if label_name and classes is not None:
            stats = _stats_from_class_counts(np.concatenate(col_stats, axis=0), classes, counts_per_class, n_global,
                                             label_name)
        model = OpTextPivotVectorizerModel(self, tops, self.cleanText, self.trackNulls)
        model._fit_stats = stats
        model._fit_key = (block.codes.data_ptr(), n)
        return model


def _stats_from_class_counts(per_class: np.ndarray, classes: np.ndarray, class_counts: np.ndarray, n: int,
                             label_name: str) -> VectorStats:
    """Statistics of 0/1 indicator columns from their per-label-class counts:
    mean c/n, variance c(n-c)/(n(n-1)), cov(., y) = sum_l y_l c_l - c mean_y."""
    cc = class_counts.astype(np.float64)
    mean_y = float((classes * cc).sum() / n) if n else 0.0
    var_y = float((cc * (classes - mean_y) ** 2).sum() / (n - 1)) if n > 1 else 0.0
    lab = np.array([n, mean_y, var_y, classes.min() if n else 0.0, classes.max() if n else 0.0])
    c = per_class.sum(axis=1)
    mean = c / n if n else np.zeros_like(c)
    var = c * (n - c) / n / (n - 1) if n > 1 else np.zeros_like(c)
    mn = np.where(c >= n, 1.0, 0.0) if n else np.zeros_like(c)
    mx = np.where(c > 0, 1.0, 0.0)
    cov = per_class @ (classes - mean_y)
    return VectorStats(LabelInfo(label_name, n, lab, classes, class_counts), mean, var, mn, mx, cov,
                       np.ascontiguousarray(per_class.T))


class OpTextPivotVectorizerModel(OpModel):
    def __init__(self, parent, topValues: List[List[str]], shouldCleanText: bool,  # noqa: N803
                 shouldTrackNulls: bool):  # noqa: N803
        super().__init__(parent)
        self.topValues = [list(t) for t in topValues]
        self.shouldCleanText = shouldCleanText
        self.shouldTrackNulls = shouldTrackNulls
        self._fit_stats: Optional[VectorStats] = None
        self._fit_key = None

    def _widths(self) -> List[int]:
        return [len(t) + (2 if self.shouldTrackNulls else 1) for t in self.topValues]

    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        R.require_cuda()
        names = [f.name for f in self.inN]
        block = batch.strings(names)
        width = sum(self._widths())
        part = LazyOneHotPart(width, lambda keep: self._materialize(block, keep))
        same = self._fit_key == (block.codes.data_ptr(), batch.n)
        return VectorColumn(self.getOutput().name, batch.n, self.getMetadata(), [part],
                            self._fit_stats if same else None)

    def _materialize(self, block: DictBlock, keep: Optional[List[int]]) -> OneHotPart:
        dev = block if block.is_cuda else block.to("cuda")
        n, ncols = block.n, len(block.names)
        widths = self._widths()
        width = sum(widths)
        newpos = np.arange(width, dtype=np.int32)
        out_width = width
        if keep is not None:
            newpos = np.full(width, -1, dtype=np.int32)
            newpos[np.asarray(keep, dtype=np.int64)] = np.arange(len(keep), dtype=np.int32)
            out_width = len(keep)
        plan = _Plan(block, self.shouldCleanText)
        tables = []
        off = 0
        for c in range(ncols):
            top = self.topValues[c]
            index_of = {v: i for i, v in enumerate(top)}
            local = np.array([index_of.get(v, len(top)) for v in plan.values[c]], dtype=np.int64)   # OTHER = len(top)
            per_code = local[plan.ids[c]] if len(plan.ids[c]) else np.zeros(0, dtype=np.int64)
            tab = np.empty(len(per_code) + 1, dtype=np.int32)
            tab[:-1] = newpos[off + per_code]
            tab[-1] = newpos[off + len(top) + 1] if self.shouldTrackNulls else -1   # null -> indicator (or nothing)
            tables.append(tab)
            off += widths[c]
        pos = np.concatenate(tables) if tables else np.zeros(1, dtype=np.int32)
        desc = plan.desc(1)
        d_desc = torch.from_numpy(desc).cuda()
        d_pos = torch.from_numpy(pos).cuda()
        out = torch.empty((n, max(ncols, 1)), dtype=torch.int32, device="cuda")   # dense rows: no unwritten padding
        L.call("tmog_pivot_transform_dev", dev.codes.data_ptr(), dev.ld, n, ncols, d_desc.data_ptr(),
               d_pos.data_ptr(), out.data_ptr(), out.shape[1], 0, R.stream())
        return OneHotPart(out[:, :ncols], out_width)
