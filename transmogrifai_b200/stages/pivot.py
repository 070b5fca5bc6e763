"""OpTextPivotVectorizer (PickList / Text pivot: top-K one-hot + OTHER + null indicator) on the GPU.

Reference: core/src/main/scala/com/salesforce/op/stages/impl/feature/OpOneHotVectorizer.scala:61-124 (fit),
:203-247 (OpTextPivotVectorizer + model), :388-399 (metadata), :416-437 (pivotFn); params PivotParams /
MinSupportParam / CleanTextFun in Transmogrifier.scala:484-575; utils/.../text/TextUtils.scala:39-47.

Per-row work is integer only (dictionary codes); strings are touched once per dictionary entry on the host
(tmog_clean_strings / tmog_dedupe_strings / tmog_pivot_topk).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .. import _lib as L
from .. import dist as D
from .. import features as F
from .. import runtime as R
from ..columnar import (batch_key, ColumnarBatch, DictBlock, LazyOneHotPart, OneHotPart, VectorColumn, decode_strings,
                        encode_strings, round_up)
from ..metadata import NullString, OpVectorColumnMetadata, OpVectorMetadata, OtherString
from ..stats import LabelInfo, VectorStats
from .base import OpEstimator, OpModel
from .numeric import label_info_dev, label_moments_dev, stats_from_sums_continuous


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


_COLUMN_DESCRIPTIONS: Dict[tuple, OpVectorColumnMetadata] = {}


class OpTextPivotVectorizer(OpEstimator):
    input_types = (F.PICKLIST, F.TEXT, F.TEXTAREA)

    def __init__(self, uid: Optional[str] = None, operationName: str = "pivotText"):  # noqa: N803
        super().__init__(operationName, uid)
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
            for v in vals:
                key = (f.name, f.type_name, v)
                col = _COLUMN_DESCRIPTIONS.get(key)
                if col is None:                 # immutable descriptions, shared between fits
                    if len(_COLUMN_DESCRIPTIONS) > 1_000_000:
                        _COLUMN_DESCRIPTIONS.clear()
                    col = OpVectorColumnMetadata((f.name,), (f.type_name,), grouping=f.name, indicatorValue=v)
                    _COLUMN_DESCRIPTIONS[key] = col
                cols.append(col)
        return OpVectorMetadata(self.getOutputFeatureName(), cols, self.history())

    @staticmethod
    def _union_dictionaries(block):
        """-> (dictionary per column valid on every rank, per column the map local code -> code in it, or None where
        all ranks already share one dictionary)."""
        ncols = len(block.dicts)
        if D.world()[1] == 1:
            return block.dicts, [None] * ncols
        local = list(_pool().map(lambda c: _dict_signature(*block.dict_arrays(c)), range(ncols)))
        sigs = D.all_gather_object(local)
        differ = [c for c in range(ncols) if any(s[c] != sigs[0][c] for s in sigs)]
        dicts, maps = list(block.dicts), [None] * ncols
        if differ:
            gathered = D.all_gather_object([block.dicts[c] for c in differ])
            for t, c in enumerate(differ):
                index = {}
                for per_rank in gathered:
                    for v in per_rank[t]:
                        index.setdefault(v, len(index))
                dicts[c] = list(index)
                maps[c] = np.array([index[v] for v in block.dicts[c]], dtype=np.int64)
        return dicts, maps

    def fit(self, batch: ColumnarBatch) -> "OpTextPivotVectorizerModel":
        R.require_cuda()
        names = [f.name for f in self.inN]
        block = batch.strings(names)
        dev = batch.on_device("strings", names, block)
        n = batch.n
        label_name = batch.response
        # One process: the dictionaries are final, so their cleaning starts now, largest first, and runs on host
        # threads while the label classes and the count kernel are on the GPU.
        # Ranks may have encoded their rows against different dictionaries (one per exported partition): the count
        # tables are then re-indexed against the union dictionary (rank order) before they are summed.  Where all ranks
        # share their dictionaries (compared by digest) the cleaning can start at once as well.
        g_dicts, g_maps = self._union_dictionaries(block)
        early = self._start_prepare(block, g_maps)
        classes = counts_per_class = ids = y = cont = None
        if label_name:
            y = batch.label()
            y = y if y.is_cuda else y.cuda()
            classes, counts_per_class, ids = label_info_dev(batch, y, label_name)
            if classes is None:
                cont = label_moments_dev(batch, y, label_name)
        st = self._scan_begin(block.dicts, len(classes) if classes is not None else 1, cont is not None)
        self._scan(st, dev, n, ids if classes is not None else None, y, cont)
        model = self._finish_counts(block, g_dicts, g_maps, early, [st], classes, counts_per_class, n, label_name, cont)
        model._fit_key = batch_key(batch)
        return model

    def _start_prepare(self, block, g_maps):
        """The counts-independent half of every column's finish (clean + group its dictionary) on the pool threads,
        largest dictionary first -- only when the dictionaries are final (no union across ranks pending)."""
        ncols = len(block.dicts)
        if ncols == 0 or not all(m is None for m in g_maps):
            return None
        clean = int(self.cleanText)
        order = sorted(range(ncols), key=lambda c: -len(block.dicts[c]))
        return {c: _pool().submit(_prepare_dict, *block.dict_arrays(c), len(block.dicts[c]), clean) for c in order}

    @staticmethod
    def _scan_begin(dicts, ncls: int, continuous: bool) -> dict:
        """Accumulators of the count scan over one set of dictionaries (`dicts`: one list per column)."""
        ncols = len(dicts)
        ncodes = np.array([len(d) for d in dicts], dtype=np.int64)
        nids = ncodes + 1                                   # raw codes + the null id 0
        zeros = np.zeros(ncols, dtype=np.int64)
        count_offset = np.concatenate([[0], np.cumsum(nids * ncls)[:-1]]).astype(np.int64) if ncols else zeros
        desc = np.ascontiguousarray(np.stack([nids, zeros, count_offset, ncodes], axis=1)) if ncols else \
            np.zeros((0, 4), dtype=np.int64)
        total_bins = int((nids * ncls).sum())
        st = dict(dicts=list(dicts), ncls=ncls, ncodes=ncodes, nids=nids, count_offset=count_offset,
                  d_desc=torch.from_numpy(desc).cuda(), total_bins=total_bins,
                  acc=None, bad=torch.zeros(1, dtype=torch.int32, device="cuda"), ws=None, calls=0)
        if continuous:
            w_off = np.concatenate([[0], np.cumsum(nids)[:-1]]).astype(np.int64) if ncols else zeros
            st["w_off"] = w_off
            st["d_woff"] = torch.from_numpy(w_off).cuda()
            st["ws"] = torch.zeros(max(int(nids.sum()), 1), dtype=torch.float64, device="cuda")
        return st

    @staticmethod
    def _scan(st: dict, dev, n: int, ids, y, cont) -> None:
        """count[code][class] of one batch / row chunk (uint32 bins per call, widened into the int64 accumulator)."""
        ncols = len(st["dicts"])
        d_counts = torch.zeros(max(st["total_bins"], 1), dtype=torch.int32, device="cuda")     # uint32 bins
        L.call("tmog_pivot_count_dev", dev.codes.data_ptr(), dev.ld, n, ncols,
               ids.data_ptr() if ids is not None else None, st["ncls"], st["d_desc"].data_ptr(), None,
               int(st["nids"].max()) if ncols else 1, d_counts.data_ptr(), st["bad"].data_ptr(), R.stream())
        if st["calls"] == 0:
            st["first"] = d_counts                    # one batch: the uint32 table goes to the host as it is
        else:
            if st["acc"] is None:
                st["acc"] = st.pop("first").to(torch.int64) & 0xFFFFFFFF
            st["acc"].add_(d_counts.to(torch.int64) & 0xFFFFFFFF)
        if cont is not None:
            # continuous label: the label sums of every raw code next to the plain counts (one class)
            L.call("tmog_pivot_label_sum_dev", dev.codes.data_ptr(), dev.ld, n, ncols, st["d_desc"].data_ptr(), None,
                   st["d_woff"].data_ptr(), y.data_ptr(), cont[1], st["ws"].data_ptr(), R.stream())
        st["calls"] += 1

    def _finish_counts(self, block, g_dicts, g_maps, early, scans, classes, counts_per_class, n_local: int,
                       label_name, cont) -> "OpTextPivotVectorizerModel":
        """Everything after the count scans: `scans` = the accumulators of this rank, one per set of dictionaries its
        rows were encoded against (scans[0] is the one `block`, `g_dicts` and `g_maps` describe; further ones come
        with `maps`: per column the table of their codes in block.dicts)."""
        ncols = len(block.dicts)
        ncls = len(classes) if classes is not None else 1
        zeros = np.zeros(ncols, dtype=np.int64)
        st0 = scans[0]
        nids, count_offset = st0["nids"], st0["count_offset"]
        shared_dicts = all(m is None for m in g_maps)
        def host_counts(st) -> np.ndarray:
            """int64 counts of one accumulator on the host (a single call's uint32 table through pinned memory)"""
            if st.get("acc") is None:
                d32 = st["first"]
                pinned = torch.empty(d32.shape, dtype=torch.int32, pin_memory=True)
                pinned.copy_(d32, non_blocking=True)
                torch.cuda.current_stream().synchronize()
                return pinned.numpy().view(np.uint32).astype(np.int64)
            return st["acc"].cpu().numpy().copy()

        def device_counts(st) -> torch.Tensor:
            return st["acc"] if st.get("acc") is not None else st["first"].to(torch.int64) & 0xFFFFFFFF

        if len(scans) == 1 and D.world()[1] > 1 and shared_dicts:
            # every rank counted against the same dictionaries: the tables add up on the device, one copy back
            g_dev = device_counts(st0).clone()
            torch.distributed.all_reduce(g_dev)
            counts = g_dev.cpu().numpy()
            summed = True
        else:
            counts = host_counts(st0)
            summed = False
        if any(int(st["bad"].item()) for st in scans):
            raise ValueError("a dictionary code outside its column's dictionary was found in the batch")
        ws = None if cont is None else st0["ws"].cpu().numpy().copy()
        for st in scans[1:]:                                 # row chunks that came with other dictionaries
            loc = host_counts(st)
            lws = None if cont is None else st["ws"].cpu().numpy()
            for c in range(ncols):
                m = st["maps"][c]
                src = loc[st["count_offset"][c]:st["count_offset"][c] + st["nids"][c] * ncls].reshape(-1, ncls)
                dst = counts[count_offset[c]:count_offset[c] + nids[c] * ncls].reshape(-1, ncls)
                dst[0] += src[0]
                np.add.at(dst, 1 + m, src[1:])
                if ws is not None:
                    ws[st0["w_off"][c]] += lws[st["w_off"][c]]
                    np.add.at(ws, st0["w_off"][c] + 1 + m, lws[st["w_off"][c] + 1:st["w_off"][c] + st["nids"][c]])
        if classes is not None:
            n_global = int(np.asarray(counts_per_class).sum())          # rows of all ranks, from the label classes
        elif cont is not None:
            n_global = int(cont[0][0])
        else:
            n_global = int(D.allreduce_sum(np.array([n_local], dtype=np.int64))[0])
        g_ncodes = np.array([len(d) for d in g_dicts], dtype=np.int64)
        g_offset = np.concatenate([[0], np.cumsum((g_ncodes + 1) * ncls)[:-1]]).astype(np.int64) if ncols else zeros
        if not shared_dicts:
            g_counts = np.zeros(int(((g_ncodes + 1) * ncls).sum()), dtype=np.int64)
            for c in range(ncols):
                loc = counts[count_offset[c]:count_offset[c] + nids[c] * ncls].reshape(int(nids[c]), ncls)
                dst = g_counts[g_offset[c]:g_offset[c] + (g_ncodes[c] + 1) * ncls].reshape(int(g_ncodes[c]) + 1, ncls)
                if g_maps[c] is None:
                    dst[:] = loc
                else:
                    dst[0] = loc[0]                                   # the null id
                    dst[1 + g_maps[c]] = loc[1:]
            counts = g_counts
        if not summed:
            counts = D.allreduce_sum(counts)
        clean = int(self.cleanText)

        def finish(c: int):
            nc = int(g_ncodes[c])
            if early is not None:
                handle, _keep = early[c].result()
            else:
                handle, _keep = _prepare_dict(*(block.dict_arrays(c) if g_maps[c] is None
                                                else encode_strings(g_dicts[c])), nc, clean)
            try:
                raw = np.ascontiguousarray(counts[g_offset[c]:g_offset[c] + (nc + 1) * ncls])
                k = self.topK
                n_top = C.c_int(0)
                code_to_out = np.zeros(max(nc, 1), dtype=np.int32)
                per_class = np.zeros((k + 2) * ncls, dtype=np.int64)
                top_data = np.zeros(max(int(_keep[1][-1]) if len(_keep[1]) else 0, 1), dtype=np.uint8)
                top_offs = np.zeros(k + 1, dtype=np.int64)
                L.call("tmog_pivot_dict_finish", handle, raw.ctypes.data, ncls, k, self.minSupport, C.byref(n_top),
                       code_to_out.ctypes.data, per_class.ctypes.data, top_data.ctypes.data, len(top_data),
                       top_offs.ctypes.data)
            finally:
                L.call("tmog_pivot_dict_free", handle)
            t = n_top.value
            top = decode_strings(top_data, top_offs[:t + 1])
            rows = per_class.reshape(k + 2, ncls)[:t + 2]
            if not self.trackNulls:
                rows = rows[:t + 1]
            local = code_to_out[:nc] if g_maps[c] is None else code_to_out[:nc][g_maps[c]]
            return top, local, rows.astype(np.float64)

        try:
            if ncols > 1:                  # the largest dictionaries first: they are the critical path
                order = sorted(range(ncols), key=lambda c: -int(g_ncodes[c]))
                done = dict(zip(order, _pool().map(finish, order)))
                results = [done[c] for c in range(ncols)]
            else:
                results = [finish(c) for c in range(ncols)]
        except Exception:
            if early is not None:          # handles of columns whose finish never ran
                for f in early.values():
                    try:
                        L.call("tmog_pivot_dict_free", f.result()[0])
                    except Exception:
                        pass
            raise
        tops = [r[0] for r in results]
        self._metadata = self._metadata_for(tops)
        stats = None
        if label_name and classes is not None:
            stats = _stats_from_class_counts(np.concatenate([r[2] for r in results], axis=0), classes,
                                             counts_per_class, n_global, label_name)
        elif label_name:
            # continuous label: the plain counts above (one class) + the label sums of every raw code, folded into the
            # emitted columns (top values, OTHER, null) with the code -> column table of the finish
            lab, shift = cont
            w_off = st0["w_off"]
            cnt_cols, ws_cols = [], []
            for c in range(ncols):
                k = len(tops[c])
                w = ws[w_off[c]:w_off[c] + nids[c]]
                folded = np.bincount(results[c][1], weights=w[1:], minlength=k + 1)[:k + 1]
                folded = np.append(folded, w[0])                       # the null indicator
                ws_cols.append(folded[:results[c][2].shape[0]])
                cnt_cols.append(results[c][2][:, 0])
            sv = np.concatenate(cnt_cols)
            ones = np.ones_like(sv)
            stats = stats_from_sums_continuous(sv, sv, sv, ones, ones, D.allreduce_sum(np.concatenate(ws_cols)), lab,
                                               shift, label_name)
        model = OpTextPivotVectorizerModel(self, tops, self.cleanText, self.trackNulls)
        model._fit_stats = stats
        model._fit_tables = {id(block.dicts[c]): (block.dicts[c], results[c][1]) for c in range(ncols)}
        for st in scans[1:]:               # the other dictionaries this rank met: their codes through block.dicts
            for c in range(ncols):
                model._fit_tables[id(st["dicts"][c])] = (st["dicts"][c], results[c][1][st["maps"][c]])
        return model

    # -- row chunks (streaming.py): one accumulator per set of dictionaries the chunks were encoded against
    def stream_begin(self, label):
        return {"label": label, "groups": {}, "order": [], "n": 0}

    def stream_update(self, st, chunk: ColumnarBatch) -> None:
        names = [f.name for f in self.inN]
        block = chunk.strings(names)
        dev = chunk.on_device("strings", names, block)
        lab = st["label"]
        sig = tuple(id(d) for d in block.dicts)
        g = st["groups"].get(sig)
        if g is None:
            ncls = len(lab.classes) if lab is not None and lab.classes is not None else 1
            g = st["groups"][sig] = self._scan_begin(block.dicts, ncls, lab is not None and lab.classes is None)
            st["order"].append(sig)
            st["names"], st["type_name"] = list(block.names), block.type_name
        ids = y = cont = None
        if lab is not None:
            if lab.classes is not None:
                ids = lab.class_ids(chunk)
            else:
                y, cont = lab.device_label(chunk), lab.continuous()
        self._scan(g, dev, chunk.n, ids, y, cont)
        st["n"] += chunk.n

    def stream_finish(self, st) -> "OpTextPivotVectorizerModel":
        lab = st["label"]
        scans = [st["groups"][sig] for sig in st["order"]]
        if not scans:
            raise ValueError("requirement failed: Input dataset cannot be empty")
        ncols = len(scans[0]["dicts"])
        # dictionary of this rank: the first set, extended by the values only later chunks brought (first-seen order)
        dicts = [list(d) for d in scans[0]["dicts"]] if len(scans) > 1 else scans[0]["dicts"]
        if len(scans) > 1:
            index = [{v: i for i, v in enumerate(d)} for d in dicts]
            for sc in scans[1:]:
                sc["maps"] = []
                for c in range(ncols):
                    ix = index[c]
                    for v in sc["dicts"][c]:
                        if v not in ix:
                            ix[v] = len(dicts[c])
                            dicts[c].append(v)
                    sc["maps"].append(np.array([ix[v] for v in sc["dicts"][c]], dtype=np.int64))
            # the first accumulator is re-laid out for the extended dictionaries (its own codes keep their places)
            ncls = scans[0]["ncls"]
            first = scans[0]
            wide = self._scan_begin(dicts, ncls, first["ws"] is not None)
            acc = torch.zeros(max(wide["total_bins"], 1), dtype=torch.int64, device="cuda")
            first_acc = first["acc"] if first.get("acc") is not None else first["first"].to(torch.int64) & 0xFFFFFFFF
            for c in range(ncols):
                a, k = int(first["count_offset"][c]), int(first["nids"][c]) * ncls
                b = int(wide["count_offset"][c])
                acc[b:b + k] = first_acc[a:a + k]
                if first["ws"] is not None:
                    wa, wb, kk = int(first["w_off"][c]), int(wide["w_off"][c]), int(first["nids"][c])
                    wide["ws"][wb:wb + kk] = first["ws"][wa:wa + kk]
            wide["acc"], wide["bad"], wide["calls"] = acc, first["bad"], max(first["calls"], 2)
            first_tables = first["dicts"]
            scans = [wide] + scans[1:]
        block = DictBlock(st["names"], st["type_name"], torch.zeros((ncols, 4), dtype=torch.int32), dicts, 0)
        g_dicts, g_maps = self._union_dictionaries(block)
        early = self._start_prepare(block, g_maps)
        classes = counts = cont = None
        if lab is not None:
            classes, counts, cont = lab.classes, lab.class_counts, lab.continuous()
        model = self._finish_counts(block, g_dicts, g_maps, early, scans, classes, counts, st["n"],
                                    lab.name if lab is not None else None, cont)
        if len(st["order"]) > 1:           # codes of the first chunk's own dictionaries are a prefix of the extended ones
            for c in range(ncols):
                d0 = first_tables[c]
                model._fit_tables[id(d0)] = (d0, model._fit_tables[id(dicts[c])][1][:len(d0)])
        return model


def _prepare_dict(data, offs, nc: int, clean: int):
    """tmog_pivot_dict_prepare of one column's dictionary (Arrow utf8 arrays); the arrays stay alive with the handle."""
    h = C.c_void_p()
    L.call("tmog_pivot_dict_prepare", data.ctypes.data, offs.ctypes.data, nc, clean, C.byref(h))
    return h, (data, offs)


_POOL = None


def _dict_signature(data: np.ndarray, offs: np.ndarray) -> str:
    """Digest of a dictionary in Arrow utf8 layout."""
    out = []
    for arr, seed in ((data, 1), (offs, 2)):
        h = C.c_uint64(0)
        a = np.ascontiguousarray(arr)
        L.call("tmog_hash_bytes", a.ctypes.data, a.nbytes, seed, C.byref(h))
        out.append(h.value)
    return "%016x%016x" % tuple(out)


def _pool():
    """Threads for the per-column driver-side finish (the C functions release the GIL)."""
    global _POOL
    if _POOL is None:
        import os
        from concurrent.futures import ThreadPoolExecutor
        _POOL = ThreadPoolExecutor(max_workers=min(32, os.cpu_count() or 4))
    return _POOL


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
        dev_of = batch.device_thunk("strings", names, block)
        part = LazyOneHotPart(width, lambda keep: self._materialize(block, keep, dev_of))
        same = self._fit_key is not None and self._fit_key == batch_key(batch)
        return VectorColumn(self.getOutput().name, batch.n, self.getMetadata(), [part],
                            self._fit_stats if same else None)

    def stats_for(self, batch: ColumnarBatch) -> Optional[VectorStats]:
        """SanityChecker statistics of this model's one-hot columns over another batch (a row sample): counts per raw
        code, folded into the emitted columns with the model's vocabulary; every rank folds its own counts first, so
        ranks with different dictionaries need no union."""
        label_name = batch.response
        if not label_name:
            return None
        est = self.parent
        names = [f.name for f in self.inN]
        block = batch.strings(names)
        dev = batch.on_device("strings", names, block)
        y = batch.label()
        y = y if y.is_cuda else y.cuda()
        classes, counts_per_class, ids = label_info_dev(batch, y, label_name)
        cont = None if classes is not None else label_moments_dev(batch, y, label_name)
        ncls = len(classes) if classes is not None else 1
        st = est._scan_begin(block.dicts, ncls, cont is not None)
        est._scan(st, dev, batch.n, ids if classes is not None else None, y, cont)
        if int(st["bad"].item()):
            raise ValueError("a dictionary code outside its column's dictionary was found in the batch")
        acc = (st["acc"] if st.get("acc") is not None else st["first"].to(torch.int64) & 0xFFFFFFFF).cpu().numpy()
        ws = None if cont is None else st["ws"].cpu().numpy()
        rows, wrows = [], []
        for c in range(len(names)):
            k = len(self.topValues[c])
            local = self._code_to_out(block, c).astype(np.int64)
            tab = acc[st["count_offset"][c]:st["count_offset"][c] + st["nids"][c] * ncls].reshape(-1, ncls)
            out = np.zeros((k + 2, ncls), dtype=np.int64)
            np.add.at(out, local, tab[1:])
            out[k + 1] = tab[0]
            rows.append(out if self.shouldTrackNulls else out[:k + 1])
            if ws is not None:
                w = ws[st["w_off"][c]:st["w_off"][c] + st["nids"][c]]
                wr = np.zeros(k + 2)
                np.add.at(wr, local, w[1:])
                wr[k + 1] = w[0]
                wrows.append(wr if self.shouldTrackNulls else wr[:k + 1])
        per_class = D.allreduce_sum(np.concatenate(rows, axis=0)).astype(np.float64)
        if classes is not None:
            return _stats_from_class_counts(per_class, classes, counts_per_class, int(np.asarray(counts_per_class).sum()),
                                            label_name)
        sv = per_class[:, 0]
        ones = np.ones_like(sv)
        return stats_from_sums_continuous(sv, sv, sv, ones, ones, D.allreduce_sum(np.concatenate(wrows)), cont[0], cont[1],
                                          label_name)

    def _code_to_out(self, block: DictBlock, c: int) -> np.ndarray:
        """Emitted-column index (top position or OTHER) of every code of column c's dictionary."""
        d = block.dicts[c]
        hit = getattr(self, "_fit_tables", {}).get(id(d))
        if hit is not None and hit[0] is d:              # the dictionary the model was fitted on
            return hit[1]
        data, offs = block.dict_arrays(c)
        top_data, top_offs = encode_strings(self.topValues[c])
        out = np.zeros(max(len(d), 1), dtype=np.int32)
        L.call("tmog_pivot_column_map", data.ctypes.data, offs.ctypes.data, len(d), int(self.shouldCleanText),
               top_data.ctypes.data, top_offs.ctypes.data, len(self.topValues[c]), out.ctypes.data)
        return out[:len(d)]

    def _materialize(self, block: DictBlock, keep: Optional[List[int]], on_device=None) -> OneHotPart:
        dev = on_device() if on_device is not None else (block if block.is_cuda else block.to("cuda"))
        n, ncols = block.n, len(block.names)
        # position tables depend on the dictionaries and the kept columns only: row chunks encoded against the same
        # dictionaries share them
        key = (torch.cuda.current_device(), tuple(id(d) for d in block.dicts),
               None if keep is None else np.asarray(keep).tobytes())
        plans = self.__dict__.setdefault("_plans", {})
        plan = plans.get(key)
        if plan is None or any(a is not b for a, b in zip(plan[3], block.dicts)):
            if len(plans) > 8:
                plans.clear()
            plan = plans[key] = self._position_tables(block, keep) + (list(block.dicts),)
        d_desc, d_pos, out_width, _dicts = plan
        out = torch.empty((n, max(ncols, 1)), dtype=torch.int32, device="cuda")   # dense rows: no unwritten padding
        L.call("tmog_pivot_transform_dev", dev.codes.data_ptr(), dev.ld, n, ncols, d_desc.data_ptr(),
               d_pos.data_ptr(), out.data_ptr(), out.shape[1], 0, R.stream())
        return OneHotPart(out[:, :ncols], out_width)

    def _position_tables(self, block: DictBlock, keep):
        """Per column the table code -> output position (top value, OTHER, null indicator; -1 = dropped by
        SanityChecker): the column index of every code is laid out once per set of dictionaries, the kept-column
        renumbering is a gather on the device."""
        ncols = len(block.names)
        widths = self._widths()
        width = sum(widths)
        key = (torch.cuda.current_device(), tuple(id(d) for d in block.dicts))
        flats = self.__dict__.setdefault("_flat_tables", {})
        hit = flats.get(key)
        if hit is None or any(a is not b for a, b in zip(hit[2], block.dicts)):
            offs = np.concatenate([[0], np.cumsum(widths)[:-1]]).astype(np.int64) if ncols else np.zeros(0, np.int64)
            pieces = []
            for c in range(ncols):
                local = self._code_to_out(block, c)
                tab = np.empty(len(local) + 1, dtype=np.int64)
                tab[:-1] = offs[c] + local
                tab[-1] = offs[c] + len(self.topValues[c]) + 1 if self.shouldTrackNulls else width   # width = "nothing"
                pieces.append(tab)
            flat = np.concatenate(pieces) if pieces else np.zeros(1, dtype=np.int64)
            ncodes = np.array([len(d) for d in block.dicts], dtype=np.int64)
            slot_off = np.concatenate([[0], np.cumsum(ncodes)[:-1]]).astype(np.int64) if ncols else ncodes
            desc = np.ascontiguousarray(np.stack([ncodes + 1, slot_off, np.zeros_like(ncodes), ncodes], axis=1))
            if len(flats) > 4:
                flats.clear()
            hit = flats[key] = (torch.from_numpy(flat).cuda(), torch.from_numpy(desc).cuda(), list(block.dicts))
        d_flat, d_desc, _dicts = hit
        newpos = np.full(width + 1, -1, dtype=np.int32)
        if keep is None:
            newpos[:width] = np.arange(width, dtype=np.int32)
            out_width = width
        else:
            newpos[np.asarray(keep, dtype=np.int64)] = np.arange(len(keep), dtype=np.int32)
            out_width = len(keep)
        d_pos = torch.from_numpy(newpos).cuda().index_select(0, d_flat)
        return d_desc, d_pos, out_width
