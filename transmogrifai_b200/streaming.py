"""Fit + transform of a dataset that arrives in row chunks (a dataset larger than one GPU's HBM, or the exported
partitions of a Spark DataFrame one after the other).

Reference: OpWorkflow.train()'s fitAndTransformDAG (core/src/main/scala/com/salesforce/op/utils/stages/
FitStagesUtil.scala:212-290): every estimator of a layer is fitted by a fold over the partitions
(OpOneHotVectorizer.scala:95-97 RDD.fold of per-row maps, SmartTextVectorizer.scala:88-91 reduce(_ + _) of TextStats,
SequenceAggregators.scala:76-88 for the means, SanityChecker.scala:397-412 treeAggregate / reduceByKey), then the
layer's transformers map over the partitions again (:96-119).  Here the partitions are ColumnarBatch chunks:

  pass 0   label only: distinct label values, rows per value, label moments (what SanityChecker keys everything on);
           SmartTextVectorizer's TextStats fold rides along until every text column has passed the cardinality cap
  pass 1   every vectorizer's fused fit + statistics scan, chunk by chunk, into accumulators that merge by addition
           (partial sums under shared shifts, int64 count tables, per-bucket sums) -- the same merges dist.py uses
           for ranks; with several ranks each rank streams its own chunks and the accumulators are all-reduced once
  decide   VectorsCombiner / SanityChecker / MinVarianceFilter are fitted from the merged statistics sidecar alone
  pass 2   the fitted models transform chunk by chunk; only the columns SanityChecker kept are written, and `sink`
           receives every chunk's output (device parts) as soon as it is queued

No vector of the whole dataset is ever held: a chunk's raw blocks are dropped from the device once the chunk is done.
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, Dict, Iterable, List, Optional, Sequence

import numpy as np
import torch

from . import _lib as L
from . import dist as D
from . import features as F
from . import runtime as R
from .columnar import ColumnarBatch, DensePart, VectorColumn, WidthPart, round_up
from .dsl import stages_of
from .stages.base import OpEstimator, OpPipelineStage
from .stages.numeric import MAX_LABEL_CLASSES


class StreamLabel:
    """The response over all chunks and ranks: sorted distinct values and rows per value (None when there are more
    than MAX_LABEL_CLASSES: a continuous label), count / mean / variance / min / max, and the shift the label-sum
    kernels subtract.  Hands every chunk its label on the device and its class ids against the global class list."""

    def __init__(self, name: str):
        self.name = name
        self.n = 0
        self.classes: Optional[np.ndarray] = None
        self.class_counts: Optional[np.ndarray] = None
        self.stats = np.zeros(5)
        self.shift = 0.0
        self._d_classes = None
        self._d_ncls = None
        # pass 0 accumulators
        self._found: List[torch.Tensor] = []
        self._parts: List[torch.Tensor] = []
        self._d_shift = None

    # -- pass 0 ------------------------------------------------------------------------------------------------------
    def device_label(self, chunk: ColumnarBatch) -> torch.Tensor:
        cache = chunk.__dict__.setdefault("_label_dev", {})
        y = cache.get(self.name)
        if y is None:
            y = chunk.label(self.name)
            y = y if y.is_cuda else y.to("cuda", non_blocking=True)
            cache[self.name] = y
        return y

    def scan(self, chunk: ColumnarBatch) -> None:
        n = chunk.n
        if n == 0:
            return
        y = self.device_label(chunk)
        nb = C.c_size_t()
        L.call("tmog_label_classes_scratch_bytes", C.byref(nb))
        scratch = R.scratch("label", nb.value)
        found = torch.zeros(1 + 256 + 256, dtype=torch.float64, device="cuda")    # {n classes, classes, rows per class}
        ncls = torch.zeros(1, dtype=torch.int32, device="cuda")
        L.call("tmog_label_classes_dev", y.data_ptr(), n, MAX_LABEL_CLASSES, found[1:].data_ptr(), ncls.data_ptr(), None,
               found[257:].data_ptr(), scratch.data_ptr(), scratch.numel(), R.stream())
        found[0] = ncls[0]
        self._found.append(found)
        # moments: the fit scan with no feature columns, under the first chunk's pilot shift
        ld = round_up(max(n, 1), 2)
        if self._d_shift is None:
            self._d_shift = torch.zeros(1, dtype=torch.float64, device="cuda")
            L.call("tmog_real_pilot_shift_dev", None, ld, None, 0, y.data_ptr(), n, 0, self._d_shift.data_ptr(),
                   R.stream())
            if D.world()[1] > 1:
                torch.distributed.broadcast(self._d_shift, src=0)
        nb2 = C.c_size_t()
        L.call("tmog_real_fit_scratch_bytes", 0, C.byref(nb2))
        sc2 = R.scratch("real_fit", nb2.value)
        part = torch.zeros(L.REAL_NSTAT, dtype=torch.float64, device="cuda")
        L.call("tmog_real_fit_partials_dev", None, ld, None, 0, y.data_ptr(), n, 0, self._d_shift.data_ptr(),
               sc2.data_ptr(), sc2.numel(), part.data_ptr(), R.stream())
        self._parts.append(part)

    def finish(self) -> None:
        if not self._found:
            raise ValueError("requirement failed: Sample size cannot be zero")
        host = torch.stack(self._found).cpu().numpy()
        lists, continuous = [], False
        for h in host:
            k = int(h[0])
            if k < 0:
                continuous = True
                break
            lists.append((h[1:1 + k].copy(), h[257:257 + k].view(np.int64).copy()))
        local = None
        if not continuous:
            allv = np.concatenate([v for v, _ in lists])
            bits = np.unique(allv.view(np.int64))                 # distinct by bit pattern (java.lang.Double.equals)
            key = np.where(bits < 0, ~bits.view(np.uint64), bits.view(np.uint64) | np.uint64(1 << 63))
            local = bits.view(np.float64)[np.argsort(key, kind="stable")]
            if len(local) > MAX_LABEL_CLASSES:
                local = None
        union = D.union_classes(local, MAX_LABEL_CLASSES) if D.world()[1] > 1 else local
        if union is not None and len(union) > 0:
            counts = np.zeros(len(union), dtype=np.int64)
            index = {int(b): i for i, b in enumerate(union.view(np.int64))}
            for v, c in lists:
                for b, cnt in zip(v.view(np.int64), c):
                    counts[index[int(b)]] += int(cnt)
            self.classes, self.class_counts = union, D.allreduce_sum(counts)
            self._d_classes = torch.zeros(256, dtype=torch.float64, device="cuda")
            self._d_classes[:len(union)] = torch.from_numpy(union).cuda()
            self._d_ncls = torch.tensor([len(union)], dtype=torch.int32, device="cuda")
        parts = torch.stack(self._parts).cpu().numpy()
        merged = np.ascontiguousarray(parts[0])
        for k in range(1, parts.shape[0]):
            other = np.ascontiguousarray(parts[k])
            L.call("tmog_real_partials_merge", merged.ctypes.data, other.ctypes.data, 0)
        merged = D.allreduce_real_partials(merged)
        shift = self._d_shift.cpu().numpy()
        lab = np.zeros(5)
        L.call("tmog_real_finish", merged.ctypes.data, shift.ctypes.data, 0, 1, 0.0, 0, None, None, None, None, None,
               None, lab.ctypes.data, None)
        self.stats, self.shift, self.n = lab, float(shift[0]), int(lab[0])
        self._found, self._parts = [], []

    # -- passes 1 / 2 ------------------------------------------------------------------------------------------------
    def continuous(self):
        """(label moments, shift) when the label is continuous, else None"""
        return None if self.classes is not None else (self.stats, self.shift)

    def class_ids(self, chunk: ColumnarBatch) -> torch.Tensor:
        """One byte per row of the chunk: index of its label in the global class list (cached on the chunk)."""
        cache = chunk.__dict__.setdefault("_label_cache", {})
        hit = cache.get(self.name)
        if hit is not None:
            return hit[2]
        y = self.device_label(chunk)
        ids = torch.empty(max(chunk.n, 1), dtype=torch.uint8, device="cuda")
        L.call("tmog_label_class_ids_dev", y.data_ptr(), chunk.n, self._d_classes.data_ptr(), self._d_ncls.data_ptr(),
               MAX_LABEL_CLASSES, ids.data_ptr(), None, R.stream())
        cache[self.name] = (self.classes, self.class_counts, ids)
        return ids

    def prime(self, chunk: ColumnarBatch) -> None:
        """Make the stage code that asks a batch for its label classes / moments see the global ones."""
        cache = chunk.__dict__.setdefault("_label_cache", {})
        if self.classes is not None:
            self.class_ids(chunk)
        else:
            cache[self.name] = (None, None, None)
            cache[(self.name, "moments")] = (self.stats, self.shift)


def _block_tensors(block):
    for name in ("values", "valid", "codes"):
        t = getattr(block, name, None)
        if isinstance(t, torch.Tensor):
            yield t
    for name in ("data", "offsets", "valid", "codes"):
        ts = getattr(block, name, None)
        if isinstance(ts, (list, tuple)):
            for t in ts:
                if isinstance(t, torch.Tensor):
                    yield t


class ChunkCache:
    """Uploaded chunks kept in HBM between pass 1 and pass 2 (by their position in the stream) as long as they fit a
    byte budget: a partition that is still resident when the transform comes needs no second trip over PCIe -- the
    streamed counterpart of ColumnarBatch.on_device sharing one upload between an estimator's fit and its model's
    transform.  budget = None: half of the device memory that is free when the cache is made."""

    def __init__(self, budget: Optional[int] = None):
        if budget is None:
            free, _total = torch.cuda.mem_get_info()
            budget = free // 2
        self.budget, self.used, self.hits = int(budget), 0, 0
        self.items: Dict[int, ColumnarBatch] = {}

    @staticmethod
    def nbytes(chunk: ColumnarBatch) -> int:
        return sum(t.numel() * t.element_size() for b in chunk.blocks for t in _block_tensors(b))

    def offer(self, k: int, dev: ColumnarBatch) -> None:
        nb = self.nbytes(dev)
        if k not in self.items and self.used + nb <= self.budget:
            self.items[k] = dev
            self.used += nb

    def take(self, k: int) -> Optional[ColumnarBatch]:
        dev = self.items.pop(k, None)
        if dev is not None:
            self.used -= self.nbytes(dev)
            self.hits += 1
        return dev


def device_chunks(chunks: Iterable[ColumnarBatch], depth: int = 1, cache: Optional[ChunkCache] = None,
                  fill: bool = False, needed: Optional[Callable[[int, ColumnarBatch, object], bool]] = None):
    """The chunks in HBM, in order.  Chunks that are there already pass through; a chunk in (pinned) host memory is
    copied on a side stream while the chunks before it are being worked on, `depth` chunks ahead -- partition k + 1
    crosses the link under partition k's kernels.  At most depth + 2 uploaded chunks exist at a time: the copy of a
    further chunk waits for the work on chunk k - 1 to finish.  The device copy keeps the host chunk's token.
    needed(index, chunk, block) -> False leaves a host block where it is (no stage will read it for this chunk)."""
    main = torch.cuda.current_stream()
    copy = None
    pending = []                       # [(device chunk, upload event)]
    done: List[torch.cuda.Event] = []  # work on earlier yielded chunks finished

    index = [0]

    def upload(chunk: ColumnarBatch):
        nonlocal copy
        k = index[0]
        index[0] += 1
        if all(getattr(b, "is_cuda", True) for b in chunk.blocks):
            return chunk, None
        if cache is not None and not fill:
            kept = cache.take(k)                      # still resident from pass 1
            if kept is not None:
                return kept, None
        if copy is None:
            copy = torch.cuda.Stream()
        if len(done) > depth + 1:
            done[-(depth + 2)].synchronize()
        with torch.cuda.stream(copy):
            blocks = [b if b.is_cuda or (needed is not None and not needed(k, chunk, b)) else b.to("cuda")
                      for b in chunk.blocks]
            ev = torch.cuda.Event()
            ev.record(copy)
        dev = ColumnarBatch(chunk.n, blocks, {}, chunk.response, chunk.token)
        if cache is not None and fill:
            cache.offer(k, dev)
        return dev, ev

    def hand_over(item):
        dev, ev = item
        if ev is not None:
            main.wait_event(ev)
            for b in dev.blocks:
                for t in _block_tensors(b):
                    if t.is_cuda:
                        t.record_stream(main)
        return dev

    for chunk in chunks:
        pending.append(upload(chunk))
        if len(pending) > depth:
            yield hand_over(pending.pop(0))
            ev = torch.cuda.Event()
            ev.record(main)
            done.append(ev)
            del done[:-(depth + 3)]
    while pending:
        yield hand_over(pending.pop(0))


class HostSink:
    """A sink for transform_stream that moves every chunk's output to pinned host memory on a side stream, `depth`
    chunks in flight (the device-to-host copy of chunk k runs under the kernels of chunk k + 1 and, over PCIe, beside
    the upload of chunk k + 2).  consume(index, vector, host_tensors) is called once a chunk's copies have landed, in
    order; the pinned buffers are reused after that."""

    def __init__(self, depth: int = 2, consume=None, compact: bool = False):
        """compact: the chunk crosses PCIe in the compact return format (compact_parts): the 0 / 1 columns of dense
        blocks and the term frequencies of sparse rows as bytes; consume then gets [(name, host tensor)] and
        expand_compact() turns them back into the vector."""
        self.depth, self.consume, self.compact = depth, consume, compact
        self.stream = torch.cuda.Stream()
        self.slots = [dict(bufs=[], ev=None, item=None) for _ in range(depth)]
        self.bytes = 0
        self.chunks = 0
        self._dest: Dict[tuple, tuple] = {}
        self._flags = None

    @staticmethod
    def _tensors(parts):
        out = []
        for p in parts:
            for name in ("data", "indices", "values", "row_start", "row_nnz"):
                t = getattr(p, name, None)
                if isinstance(t, torch.Tensor):
                    out.append(t)
        return out

    def _drain(self, slot):
        if slot["ev"] is not None:
            slot["ev"].synchronize()
            if self.consume is not None:
                k, vec, views = slot["item"]
                self.consume(k, vec, views)
            slot["ev"], slot["item"] = None, None

    def __call__(self, k: int, vec: VectorColumn, parts) -> None:
        slot = self.slots[k % self.depth]
        self._drain(slot)
        names = None
        if self.compact:
            if self._flags is None:
                self._flags = torch.zeros(4, dtype=torch.int32, device="cuda")
            named = compact_parts(vec, parts, self._dest, self._flags)
            names, tensors = [nm for nm, _t in named], [t for _nm, t in named]
        else:
            tensors = self._tensors(parts)
        main = torch.cuda.current_stream()
        self.stream.wait_stream(main)
        views = []
        with torch.cuda.stream(self.stream):
            for i, t in enumerate(tensors):
                nbytes = t.numel() * t.element_size()
                if i >= len(slot["bufs"]):
                    slot["bufs"].append(None)
                buf = slot["bufs"][i]
                if buf is None or buf.numel() < nbytes:
                    buf = slot["bufs"][i] = torch.empty(max(int(nbytes * 1.05), 1), dtype=torch.uint8, pin_memory=True)
                src = t if t.is_contiguous() else t.contiguous()
                dst = buf[:nbytes].view(t.dtype).view(src.shape)
                dst.copy_(src, non_blocking=True)
                src.record_stream(self.stream)
                views.append(dst)
                self.bytes += nbytes
            ev = torch.cuda.Event()
            ev.record(self.stream)
        slot["ev"], slot["item"] = ev, (k, vec, views if names is None else list(zip(names, views)))
        self.chunks += 1

    def finish(self) -> None:
        order = sorted((s for s in self.slots if s["item"] is not None), key=lambda s: s["item"][0])
        for s in order:
            self._drain(s)
        if self._flags is not None and int(self._flags[0].item()):
            raise ValueError("a null-indicator / Binary column holds a value other than 0 and 1: use compact=False")
        if self._flags is not None and int(self._flags[1].item()):
            raise ValueError("a column index did not fit 16 bits: use compact=False")


def _unit_columns(vec: VectorColumn, col0: int, width: int) -> np.ndarray:
    """Which columns [col0, col0 + width) of the vector are 0 / 1 by construction: null indicators and the values of
    Binary features."""
    cols = vec.metadata.columns[col0:col0 + width]
    return np.array([bool(c.isNullIndicator) or (len(c.parentFeatureType) == 1 and c.parentFeatureType[0] == F.BINARY)
                     for c in cols], dtype=bool)


def compact_parts(vec: VectorColumn, parts, dest_cache: Optional[dict] = None, flags: Optional[torch.Tensor] = None):
    """The materialised parts of a chunk's vector in the compact return format, as [(name, device tensor)]:
      dense block p  ->  "p.values" f64 [n][nv] (its other columns, in order) + "p.units" u8 [n][nu] (its 0 / 1 columns,
                         in order; _unit_columns says which they are -- the receiver applies the same rule to the
                         metadata);
      one-hot block  ->  "p.indices" int32 [n][k] as it is;
      sparse rows    ->  "p.row_start", "p.row_nnz" as they are, "p.indices" as uint16 (stored in an int16 tensor) when the
                         block has at most 65 536 columns, "p.values" u8 (255 = look the entry up in "p.over_pos" /
                         "p.over_val", "p.over_count" of them)."""
    from . import _lib as L
    dest_cache = {} if dest_cache is None else dest_cache
    flags = torch.zeros(4, dtype=torch.int32, device="cuda") if flags is None else flags
    out = []
    col0 = 0
    for i, p in enumerate(parts):
        data = p if isinstance(p, torch.Tensor) else getattr(p, "data", None)
        if isinstance(data, torch.Tensor) and data.dtype == torch.float64 and data.dim() == 2:
            n, w = data.shape
            key = (i, col0, w)
            hit = dest_cache.get(key)
            if hit is None:
                unit = _unit_columns(vec, col0, w)
                dest = np.where(unit, ~(np.cumsum(unit) - 1), np.cumsum(~unit) - 1).astype(np.int32)
                hit = dest_cache[key] = (torch.from_numpy(dest).cuda(), int((~unit).sum()), int(unit.sum()))
            d_dest, nv, nu = hit
            values = torch.empty((n, nv), dtype=torch.float64, device="cuda")
            units = torch.empty((n, nu), dtype=torch.uint8, device="cuda")
            L.call("tmog_dense_split_dev", data.data_ptr(), data.stride(0), n, w, d_dest.data_ptr(), R.ptr(values), nv,
                   R.ptr(units), nu, flags.data_ptr(), R.stream())
            out += [(f"{i}.values", values), (f"{i}.units", units)]
            col0 += w
        elif hasattr(p, "row_start"):
            nnz = p.values.numel()
            v8 = torch.empty(max(nnz, 4), dtype=torch.uint8, device="cuda")
            over_pos = torch.zeros(256, dtype=torch.int64, device="cuda")
            over_val = torch.zeros(256, dtype=torch.int32, device="cuda")
            over_count = torch.zeros(1, dtype=torch.int32, device="cuda")
            used = getattr(p, "used", None)
            L.call("tmog_values_narrow_dev", p.values.data_ptr(), nnz, R.ptr(used), v8.data_ptr(), over_pos.data_ptr(),
                   over_val.data_ptr(), 256, over_count.data_ptr(), R.stream())
            indices = p.indices
            if p.width <= 65536 and nnz:           # the block's column numbers fit 16 bits
                i16 = torch.empty(max(nnz, 8), dtype=torch.int16, device="cuda")
                L.call("tmog_indices_narrow_dev", p.indices.data_ptr(), nnz, R.ptr(used), i16.data_ptr(), flags.data_ptr(),
                       R.stream())
                indices = i16[:nnz]
            out += [(f"{i}.row_start", p.row_start), (f"{i}.row_nnz", p.row_nnz), (f"{i}.indices", indices),
                    (f"{i}.values", v8[:nnz]), (f"{i}.over_pos", over_pos), (f"{i}.over_val", over_val),
                    (f"{i}.over_count", over_count)]
            col0 += p.width
        else:
            out.append((f"{i}.indices", p.indices))
            col0 += p.width
    return out


def part_widths(parts) -> List[int]:
    return [int(p.shape[1]) if isinstance(p, torch.Tensor) else int(p.width) for p in parts]


def expand_compact(vec: VectorColumn, widths: Sequence[int], named, n: int):
    """The scipy CSR matrix of a chunk from what HostSink(compact=True) delivered (what the receiving side does);
    widths = part_widths(parts): the columns of every part, which the receiver knows from the vector's metadata."""
    import scipy.sparse as sp
    by: Dict[int, dict] = {}
    for name, t in named:
        i, field = name.split(".")
        by.setdefault(int(i), {})[field] = t.cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)
    mats = []
    col0 = 0
    for i, w in enumerate(widths):
        f = by[i]
        if "units" in f:
            unit = _unit_columns(vec, col0, w)
            dense = np.zeros((n, w))
            dense[:, ~unit] = f["values"]
            dense[:, unit] = f["units"]
            mats.append(sp.csr_matrix(dense))
        elif "row_start" in f:
            val = f["values"].astype(np.float64)
            cnt = int(f["over_count"][0])
            if cnt > len(f["over_pos"]):
                raise ValueError("more term frequencies above 254 than the overflow list holds")
            val[f["over_pos"][:cnt]] = f["over_val"][:cnt]
            nnz = f["row_nnz"].astype(np.int64)
            indptr = np.zeros(n + 1, dtype=np.int64)
            indptr[1:] = np.cumsum(nnz)
            take = np.repeat(f["row_start"] - indptr[:-1], nnz) + np.arange(indptr[-1])
            idx = f["indices"]
            idx = idx.view(np.uint16) if idx.dtype == np.int16 else idx
            mats.append(sp.csr_matrix((val[take], idx[take].astype(np.int64), indptr), shape=(n, w)))
        else:
            idx = f["indices"]
            k = idx.shape[1]
            flat = idx.reshape(-1)
            ok = flat >= 0
            m = sp.csr_matrix((np.ones(int(ok.sum())), (np.repeat(np.arange(n), k)[ok], flat[ok].astype(np.int64))),
                              shape=(n, w))
            m.sum_duplicates()
            mats.append(m)
        col0 += w
    return sp.hstack(mats).tocsr()


def _release(chunk: ColumnarBatch) -> None:
    """Drop what a pass left on a chunk (device copies of host blocks, label, class ids)."""
    for k in ("_dev_cache", "_label_dev", "_label_cache", "_dict_cache"):
        chunk.__dict__.pop(k, None)


def _is_raw_stage(st: OpPipelineStage) -> bool:
    return all(f.type_name != F.OPVECTOR for f in st.inN)


class StreamResult:
    def __init__(self, fitted, label: Optional[StreamLabel], n_local: int, n_chunks: int, vectors: Dict[str, VectorColumn]):
        self.fitted = fitted
        self.label = label
        self.rows, self.chunks = n_local, n_chunks
        self.vectors = vectors              # the vectors of the whole dataset: metadata + statistics, no rows


def fit_stream(chunks: Iterable[ColumnarBatch], result: F.Feature, response: Optional[str] = None,
               on_chunk: Optional[Callable[[int, int, ColumnarBatch], None]] = None,
               cache: Optional[ChunkCache] = None, cell_cache_bytes: int = 0) -> StreamResult:
    """Passes 0 and 1 and the decisions: every stage of `result`'s DAG fitted over the chunks (`chunks` is iterated
    several times and must yield the same chunks in the same order each time).  on_chunk(pass, index, chunk) is called
    after a chunk has been queued (for callers that recycle staging buffers).  cell_cache_bytes: HBM that stages may
    use to keep what pass 1 found in a chunk for the chunk's transform (SmartTextVectorizer: the packed cells, so that
    pass 2 neither uploads nor tokenises the text again)."""
    R.require_cuda()
    layers = stages_of(result)
    stages = [st for layer in layers for st in layer]
    raw = [st for st in stages if _is_raw_stage(st)]
    for st in raw:
        if not (hasattr(st, "stream_begin") or hasattr(st, "stream_pre_begin")):
            raise NotImplementedError(f"{type(st).__name__} cannot be fitted over row chunks yet")
    label: Optional[StreamLabel] = None
    pre = [(st, st.stream_pre_begin()) for st in raw if hasattr(st, "stream_pre_begin")]
    pre_done = [False] * len(pre)
    n_local = n_chunks = 0
    # ---- pass 0
    for k, chunk in enumerate(chunks):
        name = response or chunk.response
        if name and label is None:
            label = StreamLabel(name)
        if label is not None:
            label.scan(chunk)
        for i, (st, acc) in enumerate(pre):
            if not pre_done[i]:
                pre_done[i] = st.stream_pre_update(acc, chunk)
        n_local += chunk.n
        n_chunks += 1
        if on_chunk is not None:
            on_chunk(0, k, chunk)
        _release(chunk)
        if label is None and all(pre_done):
            break
    if n_chunks == 0:
        raise ValueError("requirement failed: Input dataset cannot be empty")
    if label is not None:
        label.finish()
    fitters: Dict[int, object] = {}
    for st, acc in pre:
        fitters[id(st)] = st.stream_pre_finish(acc, chunks)       # e.g. the SmartTextVectorizer model, no statistics yet
    # ---- pass 1
    accs = []
    cachers = [w for w in (fitters.get(id(st), st) for st in raw) if hasattr(w, "enable_cell_cache")]
    for w in cachers:
        w.enable_cell_cache(int(cell_cache_bytes) // len(cachers) if cell_cache_bytes else 0)
    for st in raw:
        worker = fitters.get(id(st), st)
        accs.append((st, worker, worker.stream_begin(label)))
    for k, chunk in enumerate(device_chunks(chunks, cache=cache, fill=True)):
        if label is not None:
            label.prime(chunk)
        for _st, worker, acc in accs:
            worker.stream_update(acc, chunk)
        if on_chunk is not None:
            on_chunk(1, k, chunk)
        _release(chunk)
    models: Dict[int, object] = {}
    stats_of: Dict[int, object] = {}
    for st, worker, acc in accs:
        out = worker.stream_finish(acc)
        if isinstance(st, OpEstimator):
            models[id(st)] = out
            stats_of[id(st)] = getattr(out, "_fit_stats", None)
        else:                                   # a transformer: stream_finish returns its statistics sidecar
            models[id(st)] = st
            stats_of[id(st)] = out
    # ---- decisions: the vector stages see a batch without rows whose vectors carry metadata + statistics
    n_all = label.n if label is not None else int(D.allreduce_sum(np.array([n_local], dtype=np.int64))[0])
    virtual = ColumnarBatch(n_all, [], {}, label.name if label is not None else None)
    # ---- pass 1s, only when a SanityChecker samples (fraction() < 1, its default above a million rows,
    #      SanityChecker.scala:356-387): the statistics it reads are those of a row sample, taken by the fitted stages
    #      over the sampled rows of every chunk (the same fused scans, finished with the fills / vocabularies of ALL rows)
    #      and merged chunk by chunk
    sample_rows = None
    samplers = [st for st in stages if not _is_raw_stage(st) and hasattr(st, "fraction") and label is not None]
    frac = min((st.fraction(n_all) for st in samplers), default=1.0)
    if frac < 1.0:
        seed = int(getattr(samplers[0], "sampleSeed", 42))
        stats_of, sample_rows = _sampled_stats(chunks, [models[id(st)] for st in raw], raw, label.name, frac, seed)
    fitted = []
    marks = {"_stream_corr": _stream_corr(chunks, fitted, sample_rows)}    # re-applied: a new batch object per vector
    if frac < 1.0:
        marks["_presampled"] = frac
    for st in stages:
        if _is_raw_stage(st):
            model = models[id(st)]
            meta = model.getMetadata()
            vec = VectorColumn(model.getOutput().name, n_all, meta, [WidthPart(meta.size)], stats_of[id(st)], model)
            virtual = virtual.with_vector(vec)
            virtual.__dict__.update(marks)
        else:
            model = st.fit(virtual) if isinstance(st, OpEstimator) else st
            vec = model.transformColumn(virtual) if not hasattr(model, "stream_transform_virtual") else \
                model.stream_transform_virtual(virtual)
            vec.producer = model
            virtual = virtual.with_vector(vec)
            virtual.__dict__.update(marks)
        fitted.append(model)
    res = StreamResult(fitted, label, n_local, n_chunks, virtual.vectors)
    res.sample_rows = sample_rows            # per chunk: the rows a sampling SanityChecker looked at (None: all rows)
    return res


def _stream_corr(chunks, fitted: list, picked: Optional[list]):
    """The SanityChecker's full Pearson matrix (CorrelationLevel.Computed, SanityChecker.scala:467-470) of a
    chunk-streamed dataset: one more pass in which the stages fitted so far (`fitted`, read when the matrix is asked for)
    produce the checker's input vector chunk by chunk -- of the sampled rows `picked[k]` when the checker samples -- and the
    Gramian kernel adds each chunk's rows to one (d + 1)^2 matrix; the ranks' matrices add.  The shifts are the column
    means the fit scans have found already (the same on every rank and chunk)."""
    def matrix(vec_name: str, label_name: str, corr_indices, stats) -> np.ndarray:
        from .stages.generic_stats import PearsonAccumulator
        d_all = len(stats.mean)
        feat = [int(i) for i in corr_indices if i < d_all]
        shifts = np.append(np.asarray(stats.mean, dtype=np.float64)[feat], float(stats.label.stats[1]))
        acc = PearsonAccumulator(d_all, corr_indices, shifts)
        models = list(fitted)
        for k, chunk in enumerate(device_chunks(chunks)):
            rows = chunk
            if picked is not None:
                rows = chunk.take_rows(picked[k]) if picked[k].numel() else None
            if rows is not None and rows.n:
                batch = ColumnarBatch(rows.n, rows.blocks, {}, None)                   # no response: no statistics
                with D.local_only():          # this rank's rows only: finish() adds the ranks' matrices
                    for model in models:
                        batch = model.transform(batch)
                    acc.update(batch, batch.vectors[vec_name], label_name)
                batch = None
            rows = None
            _release(chunk)
        return acc.finish()
    return matrix


def _sampled_stats(chunks, models, raw, label_name: str, frac: float, seed: int):
    """Statistics sidecars of every raw stage's vector over a Bernoulli(frac) sample of the rows of all chunks."""
    rank = D.world()[0]
    acc: Dict[int, object] = {}
    picked = []
    for k, chunk in enumerate(device_chunks(chunks)):
        g = torch.Generator(device="cuda").manual_seed(seed + 7919 * k + 104729 * rank)
        idx = (torch.rand(chunk.n, device="cuda", generator=g) < frac).nonzero().squeeze(1)
        picked.append(idx)
        if idx.numel():
            sub = chunk.take_rows(idx)
            sub.response = label_name
            with D.local_only():              # this rank's rows only: the ranks' statistics are merged below
                for st, model in zip(raw, models):
                    vec = model.transformColumn(sub)
                    stats = vec.stats if vec.stats is not None else model.stats_for(sub)
                    if stats is None:
                        raise NotImplementedError(f"{type(model).__name__} cannot give statistics of a row sample")
                    acc[id(st)] = stats if id(st) not in acc else acc[id(st)].merge(stats)
        _release(chunk)
    if D.world()[1] > 1:                      # the ranks' samples merge like the chunks' do, in rank order
        merged: Dict[int, object] = {}
        for per_rank in D.all_gather_object([acc.get(id(st)) for st in raw]):
            for st, stats in zip(raw, per_rank):
                if stats is not None:
                    merged[id(st)] = stats if id(st) not in merged else merged[id(st)].merge(stats)
        acc = merged
    if not acc:
        raise ValueError("requirement failed: Sample size cannot be zero")
    return acc, picked


def transform_stream(chunks: Iterable[ColumnarBatch], result: F.Feature, fit: StreamResult,
                     sink: Optional[Callable[[int, VectorColumn, list], None]] = None,
                     on_chunk: Optional[Callable[[int, int, ColumnarBatch], None]] = None,
                     cache: Optional[ChunkCache] = None) -> int:
    """Pass 2: the fitted stages applied chunk by chunk.  sink(index, vector, parts) receives the materialised parts of
    `result`'s vector for every chunk (device tensors; they may be overwritten once sink returns only if the sink
    copied or consumed them on the current stream).  -> rows transformed."""
    R.require_cuda()
    flags = torch.zeros(4, dtype=torch.int32, device="cuda")
    flags_clean = torch.zeros(4, dtype=torch.int32, device="cuda")
    rows = 0
    # a host block is uploaded unless every stage that reads its columns has what it needs of this chunk already
    readers: Dict[str, list] = {}
    for model in fit.fitted:
        for f in getattr(model, "inN", ()):
            if f.type_name != F.OPVECTOR:
                readers.setdefault(f.name, []).append(model)

    def needed(index: int, chunk: ColumnarBatch, block) -> bool:
        for name in getattr(block, "names", ()):
            for model in readers.get(name, ()):
                if not (hasattr(model, "stream_cached") and model.stream_cached(index, chunk.token)):
                    return True
        return False

    for k, chunk in enumerate(device_chunks(chunks, cache=cache, needed=needed)):
        batch = ColumnarBatch(chunk.n, chunk.blocks, {}, None, chunk.token)          # no response: no statistics
        batch.__dict__["_dev_cache"] = chunk.__dict__.setdefault("_dev_cache", {})
        batch.__dict__["_stream"] = {"index": k, "token": chunk.token, "flags": flags, "flags_clean": flags_clean}
        for model in fit.fitted:
            nb = model.transform(batch)
            nb.__dict__["_stream"] = batch.__dict__["_stream"]
            batch = nb
        vec = batch.vectors[result.name]
        parts = [p.materialize(None) if hasattr(p, "materialize") else p for p in vec.parts]
        parts = [DensePart(p) if isinstance(p, torch.Tensor) else p for p in parts]     # a dense block as a part
        for p in parts:
            if hasattr(p, "data") and not hasattr(p, "materialize"):
                _ = p.data
        if sink is not None:
            sink(k, vec, parts)
        rows += chunk.n
        if on_chunk is not None:
            on_chunk(2, k, chunk)
        _release(chunk)
        parts = vec = batch = nb = p = _ = None   # the chunk's output is the sink's now: not kept alive under the next chunk
    if hasattr(sink, "finish"):
        sink.finish()
    from .stages.smart_text import NON_ASCII_CLEAN, _check_hash_flags
    _check_hash_flags(flags)
    if int(flags_clean[0].item()):
        raise ValueError(NON_ASCII_CLEAN)
    return rows


def fit_and_transform_stream(chunks: Iterable[ColumnarBatch], result: F.Feature, sink=None,
                             response: Optional[str] = None, on_chunk=None, cache_bytes: Optional[int] = 0,
                             cell_cache_bytes: int = 0) -> StreamResult:
    """OpWorkflow.train() over row chunks: fit_stream, then transform_stream of the same chunks.  cache_bytes: HBM
    budget for host chunks that stay resident between the two (0 = none, None = half of the free device memory);
    cell_cache_bytes: HBM budget for what pass 1 keeps of every chunk for pass 2 (fit_stream)."""
    cache = None if cache_bytes == 0 else ChunkCache(cache_bytes)
    fit = fit_stream(chunks, result, response, on_chunk, cache, cell_cache_bytes)
    fit.cache_hits = 0
    try:
        transform_stream(chunks, result, fit, sink, on_chunk, cache)
    finally:
        fit.cells_cached = 0
        for model in fit.fitted:
            if hasattr(model, "release_cell_cache"):
                fit.cells_cached += len((getattr(model, "_cells", None) or {}).get("chunks", {}))
                model.release_cell_cache()
    if cache is not None:
        fit.cache_hits = cache.hits
        cache.items.clear()
    return fit
