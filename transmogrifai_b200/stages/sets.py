"""OpSetVectorizer (MultiPickList): top-K pivot of set-valued columns with OTHER and null columns
(core/src/main/scala/com/salesforce/op/stages/impl/feature/OpOneHotVectorizer.scala:75-124 fit, :160-195 the stage,
:416-437 pivotFn).  Same parameters, vocabulary rule (count >= minSupport, count descending then value ascending, top K)
and metadata as OpTextPivotVectorizer; what differs is the data: a row holds a set of values, so a row contributes to
several counts, a top value's column holds the NUMBER of raw values of the row that clean to it, and OTHER the number
of the remaining ones.

All string work is per dictionary entry on the host (clean, group, order); the GPU sees integer codes
(csrc/sets.cu).  The SanityChecker statistics of the emitted columns come from a second pass over the codes once the
vocabulary is known (the values are not 0/1, so they do not follow from the fit counts)."""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import numpy as np
import torch

from .. import _lib as L
from .. import dist as D
from .. import features as F
from .. import runtime as R
from ..columnar import batch_key, ColumnarBatch, LazyPart, SetBlock, SparseRowsPart, VectorColumn
from ..stats import LabelInfo, VectorStats
from .base import OpModel
from .numeric import label_info_dev, label_moments_dev, stats_from_sums_continuous
from .pivot import OpTextPivotVectorizer, clean_and_group, top_k


class OpSetVectorizer(OpTextPivotVectorizer):
    """SequenceEstimator[MultiPickList, OPVector]; parameters as OpTextPivotVectorizer (OpOneHotVectorizer's)."""

    input_types = (F.MULTIPICKLIST,)

    def __init__(self, uid: Optional[str] = None):
        super().__init__(uid, "vecSet")

    def fit(self, batch: ColumnarBatch) -> "OpSetVectorizerModel":
        R.require_cuda()
        names = [f.name for f in self.inN]
        block = batch.sets(names)
        dev = batch.on_device("sets", names, block)
        n, ncols = batch.n, len(names)
        label_name = batch.response
        classes = class_counts = ids = None
        if label_name:
            y = batch.label()
            y = y if y.is_cuda else y.cuda()
            classes, class_counts, ids = label_info_dev(batch, y, label_name)
        ncls = len(classes) if classes is not None else 1
        n_all = int(D.allreduce_sum(np.array([n], dtype=np.int64))[0])
        # every rank works against the union dictionary (rank order), so that groups and counts line up
        g_dicts, g_maps = self._union_dictionaries(block)
        tops: List[List[str]] = []
        pos_tables: List[torch.Tensor] = []
        d_bad = torch.zeros(1, dtype=torch.int32, device="cuda")
        for c in range(ncols):
            gid_u, distinct = clean_and_group(g_dicts[c], self.cleanText)          # over the union dictionary
            local_to_u = np.arange(len(block.dicts[c]), dtype=np.int64) if g_maps[c] is None else g_maps[c]
            gid = np.ascontiguousarray(gid_u[local_to_u], dtype=np.int32) if len(local_to_u) else \
                np.zeros(1, dtype=np.int32)
            d_gid = torch.from_numpy(gid).cuda()
            ngroups = len(distinct)
            d_counts = torch.zeros((1 + ngroups) * ncls, dtype=torch.int32, device="cuda")
            L.call("tmog_set_count_dev", dev.offsets[c].data_ptr(), dev.codes[c].data_ptr(), n, d_gid.data_ptr(),
                   len(block.dicts[c]), ngroups, ids.data_ptr() if classes is not None else None, ncls,
                   d_counts.data_ptr(), d_bad.data_ptr(), R.stream())
            counts = D.allreduce_sum(d_counts.cpu().numpy().astype(np.int64) & 0xFFFFFFFF).reshape(1 + ngroups, ncls)
            per_value = counts[1:].sum(axis=1)
            present = [g for g in range(ngroups) if per_value[g] > 0]
            order = top_k([distinct[g] for g in present], per_value[present], self.topK, self.minSupport) \
                if present else []
            top_groups = [present[i] for i in order]
            tops.append([distinct[g] for g in top_groups])
            slot_of_group = np.full(max(ngroups, 1), len(top_groups), dtype=np.int32)      # default: OTHER
            slot_of_group[top_groups] = np.arange(len(top_groups), dtype=np.int32)
            pos_tables.append(torch.from_numpy(np.ascontiguousarray(slot_of_group[gid])).cuda())
        if int(d_bad.item()):
            raise ValueError("a dictionary code outside its column's dictionary was found in the batch")
        self._metadata = self._metadata_for(tops)
        model = OpSetVectorizerModel(self, tops, self.cleanText, self.trackNulls)
        model._pos_cache = {id(block.dicts[c]): (block.dicts[c], pos_tables[c]) for c in range(ncols)}
        model._fit_key = batch_key(batch)
        if label_name:
            cont = None if classes is not None else label_moments_dev(batch, y, label_name) + (y,)
            model._fit_stats = model._sidecar(dev, n, n_all, classes, class_counts, ids, label_name, cont)
        return model


    # -- row chunks (streaming.py).  The statistics of the emitted columns need the vocabulary (a column holds how many
    #    of a row's values clean to its top value), so the counting is the pre-pass (pass 0, all chunks) and the
    #    statistics are the model's pass 1.  Chunks may come with their own dictionaries: the counts are kept per
    #    cleaned value.
    def stream_pre_begin(self):
        return {"counts": [dict() for _ in self.inN]}

    def stream_pre_update(self, st, chunk: ColumnarBatch) -> bool:
        names = [f.name for f in self.inN]
        block = chunk.sets(names)
        dev = chunk.on_device("sets", names, block)
        d_bad = torch.zeros(1, dtype=torch.int32, device="cuda")
        for c in range(len(names)):
            gid, distinct = clean_and_group(block.dicts[c], self.cleanText)
            ngroups = len(distinct)
            d_gid = torch.from_numpy(np.ascontiguousarray(gid, dtype=np.int32) if len(gid) else
                                     np.zeros(1, dtype=np.int32)).cuda()
            d_counts = torch.zeros(1 + ngroups, dtype=torch.int32, device="cuda")
            L.call("tmog_set_count_dev", dev.offsets[c].data_ptr(), dev.codes[c].data_ptr(), chunk.n, d_gid.data_ptr(),
                   len(block.dicts[c]), ngroups, None, 1, d_counts.data_ptr(), d_bad.data_ptr(), R.stream())
            per_value = d_counts.cpu().numpy().astype(np.int64)[1:] & 0xFFFFFFFF
            acc = st["counts"][c]
            for g in np.nonzero(per_value)[0]:
                acc[distinct[g]] = acc.get(distinct[g], 0) + int(per_value[g])
        if int(d_bad.item()):
            raise ValueError("a dictionary code outside its column's dictionary was found in the batch")
        return False                                   # every chunk counts

    def stream_pre_finish(self, st, chunks) -> "OpSetVectorizerModel":
        tops: List[List[str]] = []
        for c, per_rank in enumerate(zip(*D.all_gather_object(st["counts"]))):
            total: dict = {}
            for d in per_rank:
                for v, k in d.items():
                    total[v] = total.get(v, 0) + k
            values = list(total)
            order = top_k(values, np.array([total[v] for v in values], dtype=np.int64), self.topK, self.minSupport) \
                if values else []
            tops.append([values[i] for i in order])
        self._metadata = self._metadata_for(tops)
        return OpSetVectorizerModel(self, tops, self.cleanText, self.trackNulls)


class OpSetVectorizerModel(OpModel):
    def __init__(self, parent: OpSetVectorizer, topValues: List[List[str]], shouldCleanText: bool,  # noqa: N803
                 shouldTrackNulls: bool):  # noqa: N803
        super().__init__(parent)
        self.topValues = [list(t) for t in topValues]
        self.shouldCleanText = shouldCleanText
        self.shouldTrackNulls = shouldTrackNulls
        self.widths = [len(t) + (2 if shouldTrackNulls else 1) for t in self.topValues]
        self._pos_cache = {}
        self._fit_stats: Optional[VectorStats] = None
        self._fit_key = None

    # code -> output slot of a column's dictionary (top index, or len(top) = OTHER)
    def _pos_table(self, block: SetBlock, c: int) -> torch.Tensor:
        d = block.dicts[c]
        hit = self._pos_cache.get(id(d))
        if hit is not None and hit[0] is d:
            return hit[1]
        gid, distinct = clean_and_group(d, self.shouldCleanText)
        top = self.topValues[c]
        index = {v: i for i, v in enumerate(top)}
        slot_of_group = np.array([index.get(v, len(top)) for v in distinct] or [len(top)], dtype=np.int32)
        table = torch.from_numpy(np.ascontiguousarray(slot_of_group[gid]) if len(d) else
                                 np.zeros(1, dtype=np.int32)).cuda()
        self._pos_cache[id(d)] = (d, table)
        return table

    def _sidecar(self, dev: SetBlock, n: int, n_all: int, classes, class_counts, ids, label_name: str,
                 cont=None) -> VectorStats:
        """cont = (label moments, shift, label on the device) for a continuous label (then classes is None)."""
        st = self._side_begin(classes, cont is not None)
        self._side_scan(st, dev, n, ids, cont)
        return self._side_finish(st, n_all, classes, class_counts, label_name, cont)

    def _side_begin(self, classes, continuous: bool) -> dict:
        """Accumulators of the statistics scans (one batch, or the row chunks of a dataset one after the other)."""
        ncls = len(classes) if classes is not None else 1
        ws_ = [len(t) + 2 for t in self.topValues]
        return {"ncls": ncls, "bad": torch.zeros(1, dtype=torch.int32, device="cuda"),
                "cs": [np.zeros((w, ncls), dtype=np.int64) for w in ws_],
                "pr": [np.zeros((w, ncls), dtype=np.int64) for w in ws_],
                "ex": [np.zeros(w, dtype=np.int64) for w in ws_],
                "ax": [np.zeros((w, 3), dtype=np.int64) for w in ws_],
                "ws": [np.zeros(w) for w in ws_] if continuous else None}

    def _side_scan(self, st: dict, dev: SetBlock, n: int, ids, cont) -> None:
        ncls = st["ncls"]
        for c in range(len(dev.names)):
            k = len(self.topValues[c])
            w = k + 2
            pos = self._pos_table(dev, c)
            cs = torch.zeros(w * ncls, dtype=torch.int32, device="cuda")
            pr = torch.zeros(w * ncls, dtype=torch.int32, device="cuda")
            ex = torch.zeros(w, dtype=torch.int64, device="cuda")
            ax = torch.zeros(3 * w, dtype=torch.int32, device="cuda")
            ws = torch.zeros(w, dtype=torch.float64, device="cuda") if cont is not None else None
            L.call("tmog_set_stats_dev", dev.offsets[c].data_ptr(), dev.codes[c].data_ptr(), n, pos.data_ptr(),
                   len(dev.dicts[c]), k, int(self.shouldTrackNulls), ids.data_ptr() if ids is not None else None,
                   ncls, cs.data_ptr(), pr.data_ptr(), ex.data_ptr(), ax.data_ptr(),
                   cont[2].data_ptr() if cont is not None else None, cont[1] if cont is not None else 0.0,
                   ws.data_ptr() if cont is not None else None, st["bad"].data_ptr(), R.stream())
            u32 = lambda t: t.cpu().numpy().astype(np.int64) & 0xFFFFFFFF   # noqa: E731
            st["cs"][c] += u32(cs).reshape(w, ncls)
            st["pr"][c] += u32(pr).reshape(w, ncls)
            st["ex"][c] += ex.cpu().numpy()
            a = u32(ax).reshape(w, 3)
            acc = st["ax"][c]
            acc[:, 0] = np.maximum(acc[:, 0], a[:, 0])
            acc[:, 1] = np.maximum(acc[:, 1], a[:, 1])
            acc[:, 2] += a[:, 2]
            if cont is not None:
                st["ws"][c] += ws.cpu().numpy()

    def _side_finish(self, st: dict, n_all: int, classes, class_counts, label_name: str, cont) -> VectorStats:
        keep = self.widths                             # without the null slot when nulls are not tracked
        cut = lambda arrs: np.concatenate([a[:keep[c]] for c, a in enumerate(arrs)])   # noqa: E731
        cs = D.allreduce_sum(cut(st["cs"])).astype(np.float64)
        pr = D.allreduce_sum(cut(st["pr"])).astype(np.float64)
        ex = D.allreduce_sum(cut(st["ex"])).astype(np.float64)
        ax = cut(st["ax"])
        if D.world()[1] > 1:
            ax = np.stack([D.allreduce_max(ax[:, 0].copy()), D.allreduce_max(ax[:, 1].copy()),
                           D.allreduce_sum(ax[:, 2].copy())], axis=1)
        if int(st["bad"].item()):
            raise ValueError("a dictionary code outside its column's dictionary was found in the batch")
        if cont is not None:
            sv, nnz = cs[:, 0], pr[:, 0]
            vmax = np.maximum(ax[:, 0], 1).astype(np.float64)
            vmin = np.where(ax[:, 2] >= nnz, (~ax[:, 1]) & 0xFFFFFFFF, 1).astype(np.float64)
            return stats_from_sums_continuous(sv, sv + ex, nnz, vmax, vmin, D.allreduce_sum(cut(st["ws"])),
                                              cont[0], cont[1], label_name)
        cc = class_counts.astype(np.float64)
        mean_y = float((classes * cc).sum() / n_all)
        var_y = float((cc * (classes - mean_y) ** 2).sum() / (n_all - 1)) if n_all > 1 else 0.0
        lab = np.array([n_all, mean_y, var_y, classes.min(), classes.max()])
        sv = cs.sum(axis=1)
        nnz = pr.sum(axis=1)
        mean = sv / n_all
        var = np.maximum((sv + ex - sv * sv / n_all) / (n_all - 1), 0.0) if n_all > 1 else np.zeros_like(sv)
        mx = np.where(nnz > 0, np.maximum(ax[:, 0], 1), 0).astype(np.float64)
        min_gt1 = (~ax[:, 1]) & 0xFFFFFFFF
        mn = np.where(nnz >= n_all, np.where(ax[:, 2] >= nnz, min_gt1, 1), 0).astype(np.float64)
        cov = cs @ (classes - mean_y)
        # the per-label sums SanityChecker's contingency tables use clip MultiPickList columns to <= 1 (:258-268)
        return VectorStats(LabelInfo(label_name, n_all, lab, classes, class_counts), mean, var, mn, mx, cov,
                           np.ascontiguousarray(pr.T))

    def stats_for(self, batch: ColumnarBatch) -> Optional[VectorStats]:
        """SanityChecker statistics of this model's columns over another batch (a row sample of the training data)."""
        label_name = batch.response
        if not label_name:
            return None
        names = [f.name for f in self.inN]
        block = batch.sets(names)
        dev = batch.on_device("sets", names, block)
        y = batch.label()
        y = y if y.is_cuda else y.cuda()
        classes, class_counts, ids = label_info_dev(batch, y, label_name)
        n_all = int(D.allreduce_sum(np.array([batch.n], dtype=np.int64))[0])
        cont = None if classes is not None else label_moments_dev(batch, y, label_name) + (y,)
        return self._sidecar(dev, batch.n, n_all, classes, class_counts, ids, label_name, cont)

    # -- row chunks (streaming.py): the statistics sidecar accumulated over the chunks (pass 1)
    def stream_begin(self, label):
        if label is None:
            return {"label": None}
        st = self._side_begin(label.classes, label.classes is None)
        st["label"] = label
        return st

    def stream_update(self, st, chunk: ColumnarBatch) -> None:
        lab = st["label"]
        if lab is None:
            return
        names = [f.name for f in self.inN]
        block = chunk.sets(names)
        dev = chunk.on_device("sets", names, block)
        ids = lab.class_ids(chunk) if lab.classes is not None else None
        cont = None if lab.classes is not None else lab.continuous() + (lab.device_label(chunk),)
        self._side_scan(st, dev, chunk.n, ids, cont)

    def stream_finish(self, st):
        lab = st["label"]
        if lab is not None:
            cont = None if lab.classes is not None else lab.continuous() + (None,)
            self._fit_stats = self._side_finish(st, lab.n, lab.classes, lab.class_counts, lab.name, cont)
        return self

    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        R.require_cuda()
        names = [f.name for f in self.inN]
        block = batch.sets(names)
        n = batch.n
        same = bool(names) and self._fit_key is not None and self._fit_key == batch_key(batch)
        dev_of = batch.device_thunk("sets", names, block)
        parts = [LazyPart(self.widths[c], (lambda keep, c=c: self._materialize(dev_of(), c, n, keep)))
                 for c in range(len(names))]
        return VectorColumn(self.getOutput().name, n, self.getMetadata(), parts, self._fit_stats if same else None)

    def _materialize(self, block: SetBlock, c: int, n: int, keep) -> SparseRowsPart:
        dev = block if block.is_cuda else block.to("cuda")
        k = len(self.topValues[c])
        width = self.widths[c]
        newpos = np.full(k + 2, -1, dtype=np.int32)
        if keep is None:
            newpos[:width] = np.arange(width, dtype=np.int32)
            out_width = width
        else:
            newpos[np.asarray(keep, dtype=np.int64)] = np.arange(len(keep), dtype=np.int32)
            out_width = len(keep)
        d_newpos = torch.from_numpy(newpos).cuda()
        pos = self._pos_table(dev, c)
        row_nnz = torch.zeros(max(n, 1), dtype=torch.int32, device="cuda")
        d_bad = torch.zeros(1, dtype=torch.int32, device="cuda")
        args = (dev.offsets[c].data_ptr(), dev.codes[c].data_ptr(), n, pos.data_ptr(), len(dev.dicts[c]), k,
                int(self.shouldTrackNulls), d_newpos.data_ptr())
        L.call("tmog_set_transform_dev", *args, None, row_nnz.data_ptr(), None, None, d_bad.data_ptr(), R.stream())
        incl = torch.cumsum(row_nnz.to(torch.int64), 0)
        row_start = (incl - row_nnz).contiguous()
        total = int(incl[-1].item()) if n else 0
        idx = torch.empty(max(total, 1), dtype=torch.int32, device="cuda")
        val = torch.empty(max(total, 1), dtype=torch.int32, device="cuda")
        L.call("tmog_set_transform_dev", *args, row_start.data_ptr(), row_nnz.data_ptr(), idx.data_ptr(),
               val.data_ptr(), d_bad.data_ptr(), R.stream())
        if int(d_bad.item()):
            raise ValueError("a dictionary code outside its column's dictionary was found in the batch")
        return SparseRowsPart(row_start[:n], row_nnz[:n], idx, val, out_width)
