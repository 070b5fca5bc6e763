"""SmartTextVectorizer on the GPU: per text column decide pivot / hash / ignore from its TextStats, pivot the
categorical ones, tokenise + hash the others, add null indicators.

Reference: core/src/main/scala/com/salesforce/op/stages/impl/feature/SmartTextVectorizer.scala:60-205 (estimator,
metadata), :207-299 (TextStats), :301-376 (model); OPCollectionHashingVectorizer.scala:184-305 (HashingFun);
TextTokenizer.scala:160-220.  Output order: [pivot block | hash block | null indicators]  (:354).
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
from ..columnar import batch_key, ColumnarBatch, LazyPart, OneHotPart, SparseRowsPart, TextBlock, VectorColumn, encode_strings
from ..metadata import NullString, OpVectorColumnMetadata, OpVectorMetadata, OtherString, TextLenString, feature_column
from ..stats import LabelInfo, VectorStats
from .base import OpEstimator, OpModel
from .numeric import label_info_dev, label_moments_dev, stats_from_sums_continuous
from .pivot import _stats_from_class_counts, clean_and_group, top_k

PIVOT, HASH, IGNORE = "Pivot", "Hash", "Ignore"
NON_ASCII_CLEAN = ("non-ASCII text met with cleanText = true: Java's Unicode case mapping is not restated on "
                   "this path")
MAX_NUM_OF_FEATURES = 1 << 17     # TransmogrifierDefaults.MaxNumOfFeatures
M64 = (1 << 64) - 1


def cell_hash(cleaned: bytes) -> int:
    """The 64-bit hash text_stats_kernel / text_code_kernel compute over a cleaned cell (FNV-1a + avalanche)."""
    h = 1469598103934665603
    for b in cleaned:
        h = ((h ^ b) * 1099511628211) & M64
    h ^= h >> 32
    h = (h * 0xd6e8feb86659fd93) & M64
    h ^= h >> 32
    return 0x1234567 if h == M64 else h


def murmur3_index(term: str, num_features: int) -> int:
    """hasher.indexOf(term) of Spark 2.4.5 HashingTF (legacy murmur3_x86_32, seed 42) -- needed on the host only for
    the feature-name prefix (OPCollectionHashingVectorizer.scala:252)."""
    data = term.encode("utf-8")
    m32 = 0xFFFFFFFF
    rotl = lambda x, r: ((x << r) | (x >> (32 - r))) & m32  # noqa: E731

    def mix_k1(k):
        k = (k * 0xcc9e2d51) & m32
        return (rotl(k, 15) * 0x1b873593) & m32

    def mix_h1(h, k):
        return (rotl(h ^ k, 13) * 5 + 0xe6546b64) & m32

    h, n = 42, len(data)
    for i in range(0, n - n % 4, 4):
        h = mix_h1(h, mix_k1(int.from_bytes(data[i:i + 4], "little")))
    for b in data[n - n % 4:]:
        h = mix_h1(h, mix_k1((b - 256 if b >= 128 else b) & m32))
    h ^= n
    h ^= h >> 16
    h = (h * 0x85ebca6b) & m32
    h ^= h >> 13
    h = (h * 0xc2b2ae35) & m32
    h ^= h >> 16
    s = h - (1 << 32) if h >= (1 << 31) else h
    r = abs(s) % num_features
    return ((-r if s < 0 else r) + num_features) % num_features


def _col_desc(block: TextBlock) -> torch.Tensor:
    d = np.zeros((len(block.names), 4), dtype=np.int64)
    for j in range(len(block.names)):
        d[j, 0] = block.data[j].data_ptr()
        d[j, 1] = block.offsets[j].data_ptr()
        d[j, 2] = 0 if block.valid[j] is None else block.valid[j].data_ptr()
    return torch.from_numpy(d).cuda()


class _HashLayout:
    """Geometry of the text part of the output vector: [hash block | text lengths | null indicators]
    (SmartTextVectorizer.scala:154-186, :333-356).  The kernels index their statistics arrays by vector index inside
    this part, with room for the null indicators of every text column whether or not they are emitted."""

    def __init__(self, names: Sequence[str], hash_idx: List[int], ignore_idx: List[int], num_features: int,
                 num_inputs: int, strategy: str, prepend: bool, track_nulls: bool, track_len: bool = False):
        s = strategy.lower()
        self.shared = True if s == "shared" else False if s == "separate" else \
            num_features * num_inputs > MAX_NUM_OF_FEATURES          # isSharedHashSpace :191-199
        self.num_features = num_features
        self.hash_idx, self.ignore_idx = hash_idx, ignore_idx
        self.text_idx = hash_idx + ignore_idx                       # lengths / null indicators: hash columns, then ignored
        self.hash_width = 0 if not hash_idx else (num_features if self.shared else num_features * len(hash_idx))
        self.len_width = len(self.text_idx) if track_len else 0
        self.null_width = len(self.text_idx) if track_nulls else 0
        self.len_base = self.hash_width
        self.null_base = self.hash_width + self.len_width
        self.kernel_width = self.null_base + len(self.text_idx)
        self.width = self.hash_width + self.len_width + self.null_width
        # the reference zips the NON-PIVOT features (hash and ignore, input order) with the hash columns (:348,358-361)
        non_pivot = sorted(self.text_idx)
        self.prefix_names = [names[non_pivot[k]] for k in range(len(hash_idx))]
        self.prepend = prepend

    def stats_columns(self, *arrays):
        """The emitted columns of arrays indexed by kernel vector index (rows)."""
        return tuple(a[:self.width] for a in arrays)

    def stats_columns_1d(self, a):
        return a[:self.width]


class SmartTextVectorizer(OpEstimator):
    input_types = (F.TEXT, F.TEXTAREA)

    def __init__(self, uid: Optional[str] = None):
        super().__init__("smartTxtVec", uid)
        self.maxCardinality = 1000           # SmartTextVectorizer.MaxCardinality (transmogrify() passes 30)
        self.minLengthStdDev = 0.0
        self.coveragePct = 0.90
        self.topK = 20
        self.minSupport = 10
        self.cleanText = True
        self.trackNulls = True
        self.trackTextLen = False
        self.textLengthType = "FullEntry"
        self.unseenName = OtherString
        self.numFeatures = 512
        self.hashSpaceStrategy = "Auto"
        self.prependFeatureName = True
        self.binaryFreq = False
        self.hashWithIndex = False
        self.minTokenLength = 1
        self.toLowercase = True
        self.autoDetectLanguage = False

    # -- params --------------------------------------------------------------------------------------------------------
    def setMaxCardinality(self, v):  # noqa: N802
        if not (0 <= v <= 3000):
            raise ValueError(f"requirement failed: maxCardinality given invalid value {v}")
        self.maxCardinality = int(v)
        return self

    def setCoveragePct(self, v):  # noqa: N802
        if not (0.0 < v <= 1.0):
            raise ValueError(f"requirement failed: coveragePct given invalid value {v}")
        self.coveragePct = float(v)
        return self

    def setMinLengthStdDev(self, v):  # noqa: N802
        self.minLengthStdDev = float(v)
        return self

    def setTopK(self, v):  # noqa: N802
        if v <= 0:
            raise ValueError(f"requirement failed: topK given invalid value {v}")
        self.topK = int(v)
        return self

    def setMinSupport(self, v):  # noqa: N802
        if v < 0:
            raise ValueError(f"requirement failed: minSupport given invalid value {v}")
        self.minSupport = int(v)
        return self

    def setCleanText(self, v):  # noqa: N802
        self.cleanText = bool(v)
        return self

    def setTrackNulls(self, v):  # noqa: N802
        self.trackNulls = bool(v)
        return self

    def setTrackTextLen(self, v):  # noqa: N802
        self.trackTextLen = bool(v)
        return self

    def setTextLengthType(self, v):  # noqa: N802  (TextLengthType: FullEntry | Tokens, SmartTextVectorizer.scala:410-417)
        if str(v).lower() not in ("fullentry", "tokens"):
            raise ValueError(f"requirement failed: textLengthType given invalid value {v}")
        self.textLengthType = "Tokens" if str(v).lower() == "tokens" else "FullEntry"
        return self

    def setNumFeatures(self, v):  # noqa: N802
        if not (0 < v <= 10_000_000):      # OPCollectionHashingVectorizer.scala:101-108
            raise ValueError(f"requirement failed: numFeatures given invalid value {v}")
        self.numFeatures = int(v)
        return self

    def setHashSpaceStrategy(self, v):  # noqa: N802
        if str(v).lower() not in ("auto", "shared", "separate"):
            raise ValueError(f"requirement failed: hashSpaceStrategy given invalid value {v}")
        self.hashSpaceStrategy = str(v)
        return self

    def setPrependFeatureName(self, v):  # noqa: N802
        self.prependFeatureName = bool(v)
        return self

    def setBinaryFreq(self, v):  # noqa: N802
        self.binaryFreq = bool(v)
        return self

    def setMinTokenLength(self, v):  # noqa: N802
        self.minTokenLength = int(v)
        return self

    def setToLowercase(self, v):  # noqa: N802
        self.toLowercase = bool(v)      # tokens are lower-cased by the analyzer either way (LowerCaseFilter)
        return self

    def setAutoDetectLanguage(self, v):  # noqa: N802
        if v:
            raise NotImplementedError("language detection is not on the GPU path (default false)")
        return self

    # -- fit -------------------------------------------------------------------------------------------------------
    def fit(self, batch: ColumnarBatch) -> "SmartTextVectorizerModel":
        R.require_cuda()
        if batch.n == 0:
            raise ValueError("requirement failed: Input dataset cannot be empty")
        names = [f.name for f in self.inN]
        block = batch.texts(names)
        dev = batch.on_device("texts", names, block)
        n, ncols = batch.n, len(names)
        self._prefix = {}                          # cleaned first rows of the columns past the cap, of THIS batch
        self._length_sources = lambda: [(dev, n, D.world()[0] << 40)]
        # 1. TextStats per column (merged over the ranks like the reference's reduce over partitions)
        values, totals = self._text_stats(dev, n)
        values = self._merge_text_stats(values, dev, n)
        methods, tops = [], []
        # columns past the cardinality cap need the first rows on the host (the reference's fold freezes there): fetch
        # them for all such columns at once
        over = [c for c in range(ncols) if values[c] is None]
        if over:
            self._fetch_prefix(dev, over, n)
        for c in range(ncols):
            m, top = self._decide(values[c], totals[c], dev, c, n)
            methods.append(m)
            tops.append(top)
        self._length_sources = None                # the estimator does not hold on to the batch
        model = SmartTextVectorizerModel(self, methods, tops)
        self._metadata = model._metadata = model._build_metadata()
        # 2. statistics sidecar (needs a categorical label)
        label_name = batch.response
        if label_name:
            y = batch.label()
            y = y if y.is_cuda else y.cuda()
            classes, class_counts, ids = label_info_dev(batch, y, label_name)
            n_all = int(D.allreduce_sum(np.array([n], dtype=np.int64))[0])
            cont = None if classes is not None else label_moments_dev(batch, y, label_name) + (y,)
            st = model._side_begin(classes, cont is not None)
            model._side_scan(st, dev, n, ids, cont, fused_rows=True)
            model._fit_stats = model._side_finish(st, n_all, classes, class_counts, label_name, cont)
            # the sparse rows the fused pass wrote belong to this batch, not to the model: they are dropped with it
            batch.__dict__.setdefault("_dev_cache", {})[("fit_rows", model.uid)] = (st.get("rows"), st.get("local_nnz"))
        model._fit_key = batch_key(batch)
        return model

    # -- row chunks (streaming.py) ------------------------------------------------------------------------------------
    # Pre-pass: TextStats of every column folded over the chunks in row order (the reference's reduce over partitions
    # taken as one left fold, like _merge_text_stats does for ranks).  A column is done once its map has passed
    # maxCardinality -- free text does so within its first rows, so the pre-pass usually ends with the first chunk.
    def stream_pre_begin(self):
        ncols = len(self.inN)
        self._prefix = {}
        return {"maps": [[] for _ in range(ncols)], "acc": [dict() for _ in range(ncols)],
                "frozen": [False] * ncols, "k": 0}

    def stream_pre_update(self, st, chunk: ColumnarBatch) -> bool:
        """-> True when no column needs anything from later chunks"""
        names = [f.name for f in self.inN]
        live = [c for c in range(len(names)) if not st["frozen"][c]]
        if live and chunk.n > 0:
            sub_names = [names[c] for c in live]
            block = chunk.texts(sub_names)
            dev = chunk.on_device("texts", sub_names, block)
            values, _totals = self._text_stats(dev, chunk.n)
            for t, c in enumerate(live):
                vc, acc = values[t], st["acc"][c]
                st["maps"][c].append(vc)
                if vc is not None and len(set(acc) | set(vc)) <= self.maxCardinality:
                    for k, v in vc.items():
                        acc[k] = acc.get(k, 0) + v
                else:
                    st["frozen"][c] = True
        else:
            for c in live:
                st["maps"][c].append({})
        st["k"] += 1
        return all(st["frozen"])

    def stream_pre_finish(self, st, chunks) -> "SmartTextVectorizerModel":
        """Global fold (ranks in order, chunks in order within a rank), decisions, model without statistics."""
        if st["k"] == 0:
            raise ValueError("requirement failed: Input dataset cannot be empty")
        names = [f.name for f in self.inN]
        ncols = len(names)
        rank, size = D.world()
        local = [None if st["frozen"][c] else st["acc"][c] for c in range(ncols)]
        gathered = D.all_gather_object(local)
        merged, frozen_by = [], {}
        for c in range(ncols):
            acc: Dict[str, int] = {}
            for r in range(size):
                vc = gathered[r][c]
                if vc is not None and len(set(acc) | set(vc)) <= self.maxCardinality:
                    for k, v in vc.items():
                        acc[k] = acc.get(k, 0) + v
                else:
                    frozen_by[c] = r
                    break
            merged.append(acc)
        mine = [c for c, r in frozen_by.items() if r == rank]
        res: Dict[int, Dict[str, int]] = {}
        need: Dict[int, List[int]] = {}
        for c in mine:                       # continue the fold over this rank's chunks from the ranks before
            acc = dict(merged[c])
            at = None
            for k, vc in enumerate(st["maps"][c]):
                if vc is not None and len(set(acc) | set(vc)) <= self.maxCardinality:
                    for kk, v in vc.items():
                        acc[kk] = acc.get(kk, 0) + v
                else:
                    at = k
                    break
            res[c] = acc
            if at is not None:
                need.setdefault(at, []).append(c)
        if need:
            last = max(need)
            for k, chunk in enumerate(chunks):
                if k in need:
                    cols = need[k]
                    sub_names = [names[c] for c in cols]
                    block = chunk.texts(sub_names)
                    dev = chunk.on_device("texts", sub_names, block)
                    self._prefix = {}
                    self._fetch_prefix(dev, list(range(len(cols))), chunk.n)
                    for t, c in enumerate(cols):
                        res[c] = self._frozen_counts(dev, t, chunk.n, seed=res[c])
                    chunk.__dict__.pop("_dev_cache", None)
                if k >= last:
                    break
        for per_rank in D.all_gather_object(res):
            for c, vc in per_rank.items():
                merged[c] = vc
        def sources():
            key = rank << 40
            for chunk in chunks:
                block = chunk.texts(names)
                yield chunk.on_device("texts", names, block), chunk.n, key
                key += chunk.n
                chunk.__dict__.pop("_dev_cache", None)

        self._length_sources = sources
        methods, tops = [], []
        for c in range(ncols):
            m, top = self._decide(merged[c], 0, None, c, 0)
            methods.append(m)
            tops.append(top)
        self._length_sources = None                # the estimator does not hold on to the chunks
        model = SmartTextVectorizerModel(self, methods, tops)
        self._metadata = model._metadata = model._build_metadata()
        return model

    def _text_stats(self, dev: TextBlock, n: int):
        """-> per column ({cleaned value: count} or None when the cardinality cap was exceeded, non-null count)"""
        ncols = len(dev.names)
        slots = 64
        while slots < 4 * (self.maxCardinality + 2):
            slots <<= 1
        head, per = C.c_size_t(), C.c_size_t()
        L.call("tmog_text_stats_layout", ncols, slots, 1, C.byref(head), C.byref(per))
        tables = torch.zeros(head.value + per.value * ncols, dtype=torch.uint8, device="cuda")
        flags = torch.zeros(4, dtype=torch.int32, device="cuda")
        desc = _col_desc(dev)
        # lengths of capped columns matter to minLengthStdDev > 0 only (which this path does not take)
        L.call("tmog_text_stats_values_dev", desc.data_ptr(), ncols, n, None, 1, int(self.cleanText), slots,
               self.maxCardinality, tables.data_ptr(), flags.data_ptr(), R.stream())
        if int(flags[0].item()) and self.cleanText:
            raise ValueError("non-ASCII text met with cleanText = true: Java's Unicode case mapping is not restated on "
                             "this path")
        raw = tables[head.value:].cpu().numpy()
        out_vals, out_tot = [], []
        for c in range(ncols):
            q = raw[c * per.value:(c + 1) * per.value]
            hashes = q[:8 * slots].view(np.uint64)
            first = q[8 * slots:16 * slots].view(np.uint64)
            counts = q[16 * slots:16 * slots + 4 * slots].view(np.uint32)
            tail = 16 * slots + ((4 * slots + 7) & ~7)
            ndistinct = int(q[tail:tail + 4].view(np.int32)[0])
            nonnull = int(q[tail + 24:tail + 32].view(np.uint64)[0])
            out_tot.append(nonnull)
            if ndistinct > self.maxCardinality:
                out_vals.append(None)
                continue
            used = np.nonzero(hashes != np.uint64(M64))[0]
            rows = first[used].astype(np.int64)
            raws = dev.cells(c, rows)
            ids, distinct = clean_and_group(raws, self.cleanText)
            vc: Dict[str, int] = {}
            for k, i in enumerate(ids):
                vc[distinct[i]] = vc.get(distinct[i], 0) + int(counts[used[k]])
            out_vals.append(vc)
        return out_vals, out_tot

    def _merge_text_stats(self, values, dev: TextBlock, n: int):
        """Row shards are the partitions of one dataset in rank order.  The value counts of a column add up over the
        ranks as long as the merged map stays within maxCardinality; the rank in whose rows it would grow past the cap
        continues the fold over its own rows in row order (that is where the reference's left fold freezes) and
        tells the others."""
        rank, size = D.world()
        if size == 1:
            return values
        gathered = D.all_gather_object(values)
        merged, frozen_by = [], {}
        for c in range(len(values)):
            acc: Dict[str, int] = {}
            for r in range(size):
                vc = gathered[r][c]
                if vc is not None and len(set(acc) | set(vc)) <= self.maxCardinality:
                    for k, v in vc.items():
                        acc[k] = acc.get(k, 0) + v
                else:
                    frozen_by[c] = r
                    break
            merged.append(acc)
        if frozen_by:
            # every rank continues the fold for the columns that freeze in its rows (their first rows fetched and
            # cleaned in one go), then ONE exchange hands all the frozen maps round
            mine = [c for c, r in frozen_by.items() if r == rank]
            if mine:
                self._fetch_prefix(dev, mine, n)
            res = {c: self._frozen_counts(dev, c, n, seed=merged[c]) for c in mine}
            for per_rank in D.all_gather_object(res):
                for c, vc in per_rank.items():
                    merged[c] = vc
        return merged

    def _fetch_prefix(self, dev: TextBlock, cols: List[int], n: int) -> None:
        """The cleaned values of the first rows of several columns with one fetch and one cleaning call."""
        if not hasattr(self, "_prefix"):
            self._prefix = {}
        rows0 = min(self._prefix_rows(), n)
        fetched = dev.prefix_cells_many(cols, rows0)
        flat = [v for c in cols for v in fetched[c] if v is not None]
        ids, distinct = clean_and_group(flat, self.cleanText)
        at = 0
        for c in cols:
            k = sum(v is not None for v in fetched[c])
            self._prefix[c] = (rows0, [distinct[i] for i in ids[at:at + k]])
            at += k

    def _prefix_rows(self) -> int:
        """Rows fetched at first for a column past the cap: free text shows maxCardinality + 1 distinct values within
        about that many rows; a column that needs more is fetched again, four times as far."""
        return 2 * (self.maxCardinality + 8)

    def _frozen_counts(self, dev: TextBlock, c: int, n: int, seed: Optional[Dict[str, int]] = None) -> Dict[str, int]:
        """TextStats of a column whose cardinality exceeds maxCardinality: the reference's left fold stops adding once
        its map holds more than maxCardinality keys (SmartTextVectorizer.scala:231-235), so only a prefix of the rows
        counts.  Emulated in row order on a prefix fetched to the host (`seed`: the map of the ranks before)."""
        rows = self._prefix_rows()
        first = True
        while True:
            rows = min(rows, n)
            pre = getattr(self, "_prefix", {}).get(c) if first else None
            first = False
            if pre is not None and pre[0] == rows:
                cleaned = pre[1]                     # the cleaned values of the first rows, fetched with the others
            else:
                present = [v for v in dev.prefix_cells(c, rows) if v is not None]
                ids, distinct = clean_and_group(present, self.cleanText)
                cleaned = [distinct[i] for i in ids]
            vc: Dict[str, int] = dict(seed or {})
            for v in cleaned:
                if len(vc) > self.maxCardinality:
                    return vc
                vc[v] = vc.get(v, 0) + 1
            if len(vc) > self.maxCardinality or rows >= n:
                return vc
            rows *= 4

    LENGTH_BINS = 4096

    def _length_std_dev(self, c: int) -> float:
        """TextStats.lengthStdDev of column c (population standard deviation of lengthCounts, :207-220) over the row
        sources the current fit registered in self._length_sources: a callable yielding (device TextBlock, rows,
        row key of its first row) in fold order."""
        lengths = self._length_counts(c)
        size = sum(lengths.values())
        if size == 0:
            return float("nan")
        mean = sum(k * v for k, v in lengths.items()) / size
        var = sum(v * (k - mean) * (k - mean) for k, v in lengths.items()) / size
        return float(np.sqrt(var))

    def _length_counts(self, c: int) -> Dict[int, int]:
        nb = self.LENGTH_BINS
        tokens = int(self.textLengthType == "Tokens")
        big = 1 << 62

        def scan(limit_key: Optional[int]):
            hist = torch.zeros(nb, dtype=torch.int64, device="cuda")
            first = torch.full((nb,), big, dtype=torch.int64, device="cuda")
            flags = torch.zeros(4, dtype=torch.int32, device="cuda")
            for dev, n, key0 in self._length_sources():
                m = n if limit_key is None else max(0, min(n, limit_key - key0 + 1))
                if m > 0:
                    desc = _col_desc(dev.rows([c]))
                    L.call("tmog_text_length_hist_dev", desc.data_ptr(), 1, m, key0, int(self.cleanText), tokens,
                           int(self.minTokenLength), nb, hist.data_ptr(), first.data_ptr(), flags.data_ptr(), R.stream())
                if limit_key is not None and key0 + n > limit_key:
                    break
            if int(flags[0].item()) and self.cleanText:
                raise ValueError(NON_ASCII_CLEAN)
            return D.allreduce_sum(hist.cpu().numpy()), -D.allreduce_max(-first.cpu().numpy())

        h, f = scan(None)
        if h[nb - 1] > 0:
            raise NotImplementedError(f"text lengths of {nb - 1} and more with minLengthStdDev > 0")
        used = np.nonzero(h)[0]
        if len(used) > self.maxCardinality:
            if tokens:
                raise NotImplementedError("textLengthType = Tokens with more distinct token lengths than maxCardinality "
                                          "(the reference's per-row fold is order dependent there)")
            # the fold stops adding after the row that brought the map past maxCardinality lengths (:231-235)
            r_star = int(np.sort(f[used])[self.maxCardinality])
            h, _ = scan(r_star)
            used = np.nonzero(h)[0]
        return {int(k): int(h[k]) for k in used}

    def _decide(self, vc: Optional[Dict[str, int]], nonnull: int, dev: TextBlock, c: int, n: int
                ) -> Tuple[str, List[str]]:
        """SmartTextVectorizer.fitFn :93-120"""
        if vc is None:
            vc = self._frozen_counts(dev, c, n)
        total = sum(vc.values())
        filtered = {k: v for k, v in vc.items() if v >= self.minSupport}
        sorted_counts = sorted(filtered.values(), reverse=True)
        cum = np.cumsum(sorted_counts) if sorted_counts else np.zeros(0)
        k = min(self.topK, len(cum)) - 1
        num = float(cum[k]) if 0 <= k < len(cum) else 0.0
        coverage = num / total if total else float("nan")
        card = len(vc)
        if card > self.maxCardinality and card > self.topK and coverage >= self.coveragePct:
            method = PIVOT
        elif card <= self.maxCardinality:
            method = PIVOT
        elif self.minLengthStdDev > 0 and self._length_std_dev(c) < self.minLengthStdDev:
            method = IGNORE                                     # reads like machine-generated ids (:113-115)
        else:
            method = HASH
        vals = list(filtered)
        order = top_k(vals, np.array([filtered[v] for v in vals], dtype=np.int64), self.topK, self.minSupport) \
            if vals else []
        return method, [vals[i] for i in order]


class SmartTextVectorizerModel(OpModel):
    def __init__(self, parent: SmartTextVectorizer, vectorizationMethods: List[str],  # noqa: N803
                 topValues: List[List[str]]):  # noqa: N803
        super().__init__(parent)
        self.vectorizationMethods = list(vectorizationMethods)
        self.topValues = [list(t) for t in topValues]
        self.shouldCleanText = parent.cleanText
        self.shouldTrackNulls = parent.trackNulls
        p = parent
        self.hashingParams = dict(numFeatures=p.numFeatures, numInputs=len(parent.inN), binaryFreq=p.binaryFreq,
                                  prependFeatureName=p.prependFeatureName, hashSpaceStrategy=p.hashSpaceStrategy,
                                  minTokenLength=p.minTokenLength)
        self.unseenName = p.unseenName
        self._fit_stats: Optional[VectorStats] = None
        self._fit_key = None
        names = [f.name for f in self.inN]
        self.pivot_idx = [j for j, m in enumerate(self.vectorizationMethods) if m == PIVOT]
        hash_idx = [j for j, m in enumerate(self.vectorizationMethods) if m == HASH]
        ignore_idx = [j for j, m in enumerate(self.vectorizationMethods) if m == IGNORE]
        self.shouldTrackLen = bool(getattr(p, "trackTextLen", False))
        self.layout = _HashLayout(names, hash_idx, ignore_idx, p.numFeatures, len(names), p.hashSpaceStrategy,
                                  p.prependFeatureName, p.trackNulls, self.shouldTrackLen)
        self.pivot_widths = [len(self.topValues[j]) + (2 if self.shouldTrackNulls else 1) for j in self.pivot_idx]
        self.pivot_width = sum(self.pivot_widths)

    @property
    def categoricalTopValues(self) -> List[List[str]]:  # noqa: N802
        return [self.topValues[j] for j in self.pivot_idx]

    # -- metadata (SmartTextVectorizer.makeVectorMetadata :154-186) ----------------------------------------------------
    def _build_metadata(self) -> OpVectorMetadata:
        feats = self.inN
        runs: List[Tuple[OpVectorColumnMetadata, int]] = []
        for j in self.pivot_idx:
            f = feats[j]
            vals = list(self.topValues[j]) + [self.unseenName] + ([NullString] if self.shouldTrackNulls else [])
            runs += [(OpVectorColumnMetadata((f.name,), (f.type_name,), grouping=f.name, indicatorValue=v), 1)
                     for v in vals]
        lay = self.layout
        if lay.text_idx:
            hf = [feats[j] for j in lay.hash_idx]
            if lay.hash_idx:
                if lay.shared:
                    runs.append((OpVectorColumnMetadata(tuple(f.name for f in hf), tuple(f.type_name for f in hf)),
                                 lay.num_features))
                else:
                    runs += [(feature_column(f.name, f.type_name), lay.num_features) for f in hf]
            if self.shouldTrackLen:             # :168-172: lengths of the hashed columns, then of the ignored ones
                runs += [(OpVectorColumnMetadata((feats[j].name,), (feats[j].type_name,), grouping=feats[j].name,
                                                 descriptorValue=TextLenString), 1) for j in lay.text_idx]
            # the reference appends the null-indicator columns to the metadata whether or not nulls are tracked
            # (SmartTextVectorizer.scala:176-183); kept as is, SanityChecker then rejects the width mismatch
            runs += [(feature_column(feats[j].name, feats[j].type_name, is_null=True), 1) for j in lay.text_idx]
        return OpVectorMetadata(self.parent.getOutputFeatureName(), history=self.parent.history(), runs=runs)

    # -- device tables -------------------------------------------------------------------------------------------------
    def _pivot_table_build(self, ncols: int):
        """Small hash tables holding just the top values of every pivoted column (built once per model and device)."""
        key = (torch.cuda.current_device(), ncols)
        hit = self.__dict__.get("_pt_cache")
        if hit is not None and hit[0] == key:
            return hit[1]
        slots = 64
        while slots < 4 * (max((len(self.topValues[j]) for j in self.pivot_idx), default=0) + 2):
            slots <<= 1
        head, per = C.c_size_t(), C.c_size_t()
        L.call("tmog_text_stats_layout", ncols, slots, 1, C.byref(head), C.byref(per))
        host = np.zeros(head.value + per.value * ncols, dtype=np.uint8)
        # struct TextStatsTable {hash*, first_row*, counts*, ndistinct*, len_sum*, nonnull*} per column
        tables = torch.zeros(len(host), dtype=torch.uint8, device="cuda")
        base = tables.data_ptr() + head.value
        ptrs = np.zeros((ncols, 6), dtype=np.int64)
        slot_of: List[Dict[str, int]] = []
        for c in range(ncols):
            q0 = base + per.value * c
            tail = 16 * slots + ((4 * slots + 7) & ~7)
            ptrs[c] = [q0, q0 + 8 * slots, q0 + 16 * slots, q0 + tail, q0 + tail + 8, q0 + tail + 24]
            hashes = np.full(slots, M64, dtype=np.uint64)
            mapping: Dict[str, int] = {}
            if c in self.pivot_idx:
                for v in self.topValues[c]:
                    h = cell_hash(v.encode("utf-8"))
                    s = (h >> 20) & (slots - 1)
                    while hashes[s] != np.uint64(M64):
                        s = (s + 1) & (slots - 1)
                    hashes[s] = np.uint64(h)
                    mapping[v] = s
            off = head.value + per.value * c
            host[off:off + 8 * slots] = hashes.view(np.uint8)
            slot_of.append(mapping)
        host[:ptrs.nbytes] = ptrs.view(np.uint8).reshape(-1)
        tables.copy_(torch.from_numpy(host))
        sel = torch.tensor(self.pivot_idx or [0], dtype=torch.int32, device="cuda")
        built = (tables, slots, slot_of, sel)
        self.__dict__["_pt_cache"] = (key, built)
        return built

    def _pivot_tables(self, dev: TextBlock, n: int, flags: Optional[torch.Tensor] = None):
        """The dictionary codes of all rows against the top-value tables (code -2 = a value outside the vocabulary ->
        OTHER).  With `flags` given the non-ASCII check is left to the caller (row chunks: checked once at the end)."""
        k = len(self.pivot_idx)
        tables, slots, slot_of, sel = self._pivot_table_build(len(dev.names))
        ld = (n + 3) // 4 * 4
        codes = torch.empty((max(k, 1), ld), dtype=torch.int32, device="cuda")
        desc = _col_desc(dev)
        own = flags is None
        if own:
            flags = torch.zeros(4, dtype=torch.int32, device="cuda")
        L.call("tmog_text_codes_dev", desc.data_ptr(), sel.data_ptr(), k, n, int(self.shouldCleanText), slots,
               tables.data_ptr(), codes.data_ptr(), ld, flags.data_ptr(), R.stream())
        if own and k and self.shouldCleanText and int(flags[0].item()):
            raise ValueError(NON_ASCII_CLEAN)
        return codes, ld, slots, slot_of

    def _pivot_pos_tables(self, slots: int, slot_of, newpos: np.ndarray):
        """pos tables for tmog_pivot_transform_dev over the slot codes: entry s = output index of the value in slot s,
        every other slot (and code -2, clamped to `slots`) = OTHER... the null code -1 -> null indicator."""
        tables, desc = [], []
        off, slot_off = 0, 0
        for t, j in enumerate(self.pivot_idx):
            top = self.topValues[j]
            tab = np.full(slots + 2, newpos[off + len(top)], dtype=np.int32)           # default: OTHER
            for i, v in enumerate(top):
                tab[slot_of[j][v]] = newpos[off + i]
            # layout expected by the kernel: ncodes + 1 entries, the last one for a null (code < 0 or >= ncodes)
            tab[slots + 1] = newpos[off + len(top) + 1] if self.shouldTrackNulls else -1
            tables.append(tab)
            desc.append([slots + 2, slot_off, 0, slots + 1])
            slot_off += slots + 1
            off += self.pivot_widths[t]
        return np.concatenate(tables) if tables else np.zeros(1, dtype=np.int32), \
            np.asarray(desc, dtype=np.int64).reshape(-1, 4)

    def _hash_args(self, dev: TextBlock):
        lay = self.layout
        hp = self.hashingParams
        k = len(lay.hash_idx)
        key = torch.cuda.current_device()
        hit = self.__dict__.get("_ha_cache")
        if hit is None or hit[0] != key:
            rows = max(len(lay.text_idx), k, 1)        # one row per column of the call; the ignored ones stay empty
            prefix = np.zeros((rows, 16), dtype=np.uint8)
            plen = np.zeros(rows, dtype=np.int32)
            for t in range(k):
                s = f"{murmur3_index(lay.prefix_names[t], lay.num_features)}_".encode("ascii")
                prefix[t, :len(s)] = np.frombuffer(s, dtype=np.uint8)
                plen[t] = len(s)
            hit = self.__dict__["_ha_cache"] = (key, torch.from_numpy(prefix).cuda(), torch.from_numpy(plen).cuda())
        sub = dev.rows(lay.text_idx)      # the ignored columns ride behind the hashed ones
        params = np.array([lay.num_features, int(lay.shared), int(hp["binaryFreq"]), hp["minTokenLength"],
                           int(self.shouldTrackNulls), int(hp["prependFeatureName"]), lay.null_base, 0,
                           len(lay.hash_idx) if lay.ignore_idx else 0, lay.len_base if lay.len_width else 0, 0, 0],
                          dtype=np.int32)
        return sub, hit[1], hit[2], params

    # -- statistics sidecar --------------------------------------------------------------------------------------------
    def _sidecar(self, dev: TextBlock, n: int, n_all: int, classes, class_counts, ids, label_name: str,
                 cont=None) -> VectorStats:
        """n = local rows, n_all = rows of all ranks; the count tables and bucket sums are all-reduced.
        cont = (label moments, shift, label on the device) for a continuous label (then classes is None)."""
        st = self._side_begin(classes, cont is not None)
        self._side_scan(st, dev, n, ids, cont)
        return self._side_finish(st, n_all, classes, class_counts, label_name, cont)

    def _side_begin(self, classes, continuous: bool) -> dict:
        """Accumulators of the statistics scans (one batch, or the row chunks of a dataset one after the other)."""
        ncls = len(classes) if classes is not None else 1
        st: dict = {"ncls": ncls, "flags": torch.zeros(4, dtype=torch.int32, device="cuda"),
                    "pflags": torch.zeros(4, dtype=torch.int32, device="cuda"), "chunk_nnz": [], "scans": 0}
        if self.pivot_idx:
            k = len(self.pivot_idx)
            _tables, slots, _slot_of, _sel = self._pivot_table_build(len(self.inN))
            nids = slots + 2
            desc = np.zeros((k, 4), dtype=np.int64)
            for t in range(k):
                desc[t] = [nids, 0, t * nids * ncls, slots + 1]
            st.update(p_desc=torch.from_numpy(desc).cuda(), p_nids=nids,
                      p_acc=torch.zeros(k * nids * ncls, dtype=torch.int64, device="cuda"),
                      p_bad=torch.zeros(1, dtype=torch.int32, device="cuda"),
                      p_ws=torch.zeros(k * nids, dtype=torch.float64, device="cuda") if continuous else None,
                      p_woff=torch.arange(0, k * nids, nids, dtype=torch.int64, device="cuda"))
        lay = self.layout
        if lay.text_idx:
            w = lay.kernel_width
            st.update(w=w, class_sum=torch.zeros(w * ncls, dtype=torch.int64, device="cuda"),
                      dup=torch.zeros(w, dtype=torch.int64, device="cuda"),
                      extra=torch.zeros(w, dtype=torch.int64, device="cuda"),
                      aux=torch.zeros((w, 3), dtype=torch.int64, device="cuda"),
                      wsum=torch.zeros(w, dtype=torch.float64, device="cuda") if continuous else None)
        return st

    def _side_scan(self, st: dict, dev: TextBlock, n: int, ids, cont, fused_rows: bool = False) -> None:
        """One batch / row chunk into the accumulators.  fused_rows: the same tokenise + hash pass also writes the sparse
        rows of the batch (st["rows"]; a workflow transforms its training rows right after the fit)."""
        ncls = st["ncls"]
        ids_ptr = ids.data_ptr() if ids is not None else None
        if self.pivot_idx:
            codes, ld, slots, _slot_of = self._pivot_tables(dev, n, st["pflags"])
            # count per slot code: remap -2 (unknown value) to slot index `slots` so that it lands in OTHER
            codes = torch.where(codes == -2, torch.full_like(codes, slots), codes)
            k = len(self.pivot_idx)
            d_counts = torch.zeros(k * st["p_nids"] * ncls, dtype=torch.int32, device="cuda")
            L.call("tmog_pivot_count_dev", codes.data_ptr(), ld, n, k, ids_ptr, ncls, st["p_desc"].data_ptr(), None,
                   st["p_nids"], d_counts.data_ptr(), st["p_bad"].data_ptr(), R.stream())
            st["p_acc"] += d_counts.to(torch.int64) & 0xFFFFFFFF
            if cont is not None:
                L.call("tmog_pivot_label_sum_dev", codes.data_ptr(), ld, n, k, st["p_desc"].data_ptr(), None,
                       st["p_woff"].data_ptr(), cont[2].data_ptr(), cont[1], st["p_ws"].data_ptr(), R.stream())
        lay = self.layout
        if lay.text_idx:
            sub, d_prefix, d_plen, params = self._hash_args(dev)
            params[7] = ncls
            params[6] = lay.null_base
            w = st["w"]
            class_sum = torch.zeros(w * ncls, dtype=torch.int32, device="cuda")
            dup = torch.zeros(w, dtype=torch.int32, device="cuda")
            aux = torch.zeros(3 * w, dtype=torch.int32, device="cuda")
            extra = st["extra"]                       # 64-bit sums: accumulated in place
            flags = st["flags"]
            scratch = _hash_scratch(n, len(sub.names), lay, fused_rows)
            desc = _col_desc(sub)
            wsum = st["wsum"]
            k = len(sub.names)
            label_ptr = cont[2].data_ptr() if cont is not None else None
            shift = cont[1] if cont is not None else 0.0
            cells = getattr(self, "_cells", None) if st.get("pack_token") is not None else None
            if not fused_rows and cells is not None and not cells.get("full"):
                import ctypes
                need = ctypes.c_size_t(0)
                L.call("tmog_text_cells_pack_bytes", n, k, int(lay.len_width > 0), ctypes.byref(need))
                meta = torch.empty(max(need.value, 16), dtype=torch.uint8, device="cuda")
                pflag = torch.zeros(1, dtype=torch.int32, device="cuda")
                L.call("tmog_text_hash_stats_pack_dev", desc.data_ptr(), k, n, d_prefix.data_ptr(), d_plen.data_ptr(),
                       params.ctypes.data, ids_ptr, class_sum.data_ptr(), dup.data_ptr(), extra.data_ptr(),
                       aux.data_ptr(), label_ptr, shift, R.ptr(wsum), flags.data_ptr(), scratch.data_ptr(),
                       scratch.numel(), meta.data_ptr(), cells["arena"].data_ptr(), cells["arena"].numel(),
                       cells["cursor"].data_ptr(), pflag.data_ptr(), R.stream())
                cells["pending"].append((st["pack_token"], meta, pflag, n, k))
                # every 8 chunks: has the arena run full?  (one small read; chunks after that are not packed)
                if len(cells["pending"]) % 8 == 0 and int(cells["cursor"].item()) >= cells["arena"].numel():
                    cells["full"] = True
            elif not fused_rows:
                L.call("tmog_text_hash_stats_dev", desc.data_ptr(), k, n, d_prefix.data_ptr(), d_plen.data_ptr(),
                       params.ctypes.data, ids_ptr, class_sum.data_ptr(), dup.data_ptr(), extra.data_ptr(),
                       aux.data_ptr(), label_ptr, shift, R.ptr(wsum), flags.data_ptr(), scratch.data_ptr(),
                       scratch.numel(), R.stream())
            else:
                # One pass over the text gives the statistics AND the sparse rows of this batch: the output is sized
                # from a sample of the first rows; if the estimate turns out too small the rows are dropped and the
                # transform tokenises again.
                ident = torch.arange(w, dtype=torch.int32, device="cuda")
                row_start = torch.empty(max(n, 1), dtype=torch.int64, device="cuda")
                row_nnz = torch.empty(max(n, 1), dtype=torch.int32, device="cuda")
                cursor = torch.zeros(1, dtype=torch.int64, device="cuda")
                m = min(n, 16384)
                params1 = params.copy()
                params1[7] = 1
                L.call("tmog_text_hash_transform_dev", desc.data_ptr(), k, m, d_prefix.data_ptr(), d_plen.data_ptr(),
                       params1.ctypes.data, ident.data_ptr(), row_start.data_ptr(), row_nnz.data_ptr(), None, None,
                       cursor.data_ptr(), 0, flags.data_ptr(), scratch.data_ptr(), scratch.numel(), R.stream())
                capacity = int(int(cursor.item()) / max(m, 1) * n * 1.1) + 65536
                cursor.zero_()
                flags.zero_()
                idx = torch.empty(capacity, dtype=torch.int32, device="cuda")
                val = torch.empty(capacity, dtype=torch.int32, device="cuda")
                L.call("tmog_text_hash_fit_transform_dev", desc.data_ptr(), k, n, d_prefix.data_ptr(),
                       d_plen.data_ptr(), params.ctypes.data, ids_ptr, class_sum.data_ptr(), dup.data_ptr(),
                       extra.data_ptr(), aux.data_ptr(), label_ptr, shift, R.ptr(wsum), ident.data_ptr(),
                       row_start.data_ptr(), row_nnz.data_ptr(), idx.data_ptr(), val.data_ptr(), cursor.data_ptr(),
                       capacity, flags.data_ptr(), scratch.data_ptr(), scratch.numel(), R.stream())
                too_small = int(flags[3].item()) != 0
                flags[3] = 0
                st["rows"] = None if too_small else SparseRowsPart(row_start[:n], row_nnz[:n], idx, val, w)
            # unsigned 32-bit words of this scan widened into the 64-bit accumulators
            u32 = lambda t: t.to(torch.int64) & 0xFFFFFFFF   # noqa: E731
            cs64, dp64, ax64 = u32(class_sum), u32(dup), u32(aux).view(w, 3)
            st["class_sum"] += cs64
            st["dup"] += dp64
            acc = st["aux"]
            torch.maximum(acc[:, 0], ax64[:, 0], out=acc[:, 0])
            torch.maximum(acc[:, 1], ax64[:, 1], out=acc[:, 1])
            acc[:, 2] += ax64[:, 2]
            # entries this scan would write as sparse rows (all columns): sizes the transform of the same chunk
            st["chunk_nnz"].append(cs64.sum() - dp64.sum())
        st["scans"] += 1

    def _side_finish(self, st: dict, n_all: int, classes, class_counts, label_name: str, cont) -> VectorStats:
        ncls = st["ncls"]
        parts = []
        if self.pivot_idx:
            if self.shouldCleanText and int(st["pflags"][0].item()):
                raise ValueError(NON_ASCII_CLEAN)
            k = len(self.pivot_idx)
            _tables, slots, slot_of, _sel = self._pivot_table_build(len(self.inN))
            nids = st["p_nids"]
            cnt = D.allreduce_sum(st["p_acc"].cpu().numpy()).reshape(k, nids, ncls)
            ws = None
            if cont is not None:
                ws = D.allreduce_sum(st["p_ws"].cpu().numpy()).reshape(k, nids)
            rows, wrows = [], []
            for t, j in enumerate(self.pivot_idx):
                top = self.topValues[j]
                tab = cnt[t]                       # row 0 = null, row 1 + s = slot s, row 1 + slots = unknown
                chosen = [1 + slot_of[j][v] for v in top]
                rest = np.ones(nids, dtype=bool)
                rest[0] = False
                rest[chosen] = False
                r = [tab[i] for i in chosen] + [tab[rest].sum(axis=0)]
                if self.shouldTrackNulls:
                    r.append(tab[0])
                rows.append(np.stack(r).astype(np.float64))
                if ws is not None:
                    wr = [ws[t][i] for i in chosen] + [ws[t][rest].sum()]
                    if self.shouldTrackNulls:
                        wr.append(ws[t][0])
                    wrows.append(np.asarray(wr))
            if cont is None:
                parts.append(_stats_from_class_counts(np.concatenate(rows, axis=0), classes, class_counts, n_all,
                                                      label_name))
            else:
                sv = np.concatenate(rows, axis=0)[:, 0]
                ones = np.ones_like(sv)
                parts.append(stats_from_sums_continuous(sv, sv, sv, ones, ones, np.concatenate(wrows), cont[0], cont[1],
                                                        label_name))
        lay = self.layout
        if lay.text_idx:
            _check_hash_flags(st["flags"])
            w = st["w"]
            # the bucket arrays of all scans brought back in one copy
            packed = torch.cat([st["class_sum"].reshape(-1), st["dup"].reshape(-1), st["extra"].reshape(-1),
                                st["aux"].reshape(-1)]).cpu().numpy()
            k = w * ncls
            cs = packed[:k].reshape(w, ncls).astype(np.float64)
            # rows holding an index = sum of its term frequencies - (tf - 1) of the rows with tf > 1
            nz = cs.sum(axis=1) - packed[k:k + w].astype(np.float64)
            ex = packed[k + w:k + 2 * w].astype(np.float64)
            ax = packed[k + 2 * w:].reshape(w, 3)
            st["local_nnz"] = nz                     # of the local rows: sizes the sparse output of the same batch
            if D.world()[1] > 1:
                # two exchanges in all: everything that adds (counts below 2^53, exact in f64), everything that maxes
                sums = D.allreduce_sum(np.concatenate([cs.reshape(-1), nz, ex, ax[:, 2].astype(np.float64)]))
                maxs = D.allreduce_max(np.concatenate([ax[:, 0], ax[:, 1]]))
                k, m = cs.size, len(nz)
                cs, nz, ex = sums[:k].reshape(cs.shape), sums[k:k + m], sums[k + m:k + 2 * m]
                ax = np.stack([maxs[:m], maxs[m:], sums[k + 2 * m:].astype(np.int64)], axis=1)
            stats = lay.stats_columns(cs, nz, ex, ax)
            if cont is None:
                parts.append(_stats_from_bucket_sums(*stats, classes, class_counts, n_all, label_name, lay))
            else:
                cs, nz, ex, ax = stats
                wv = lay.stats_columns_1d(D.allreduce_sum(st["wsum"].cpu().numpy()))
                sv = cs[:, 0]
                vmax = np.maximum(ax[:, 0], 1).astype(np.float64)
                vmin = np.where(ax[:, 2] >= nz, ((~ax[:, 1]) & 0xFFFFFFFF), 1).astype(np.float64)
                parts.append(stats_from_sums_continuous(sv, sv + ex, nz, vmax, vmin, wv, cont[0], cont[1], label_name))
        return VectorStats.concat(parts)

    def stats_for(self, batch: ColumnarBatch) -> Optional[VectorStats]:
        """SanityChecker statistics of this model's columns over another batch (a row sample of the training data)."""
        label_name = batch.response
        if not label_name:
            return None
        names = [f.name for f in self.inN]
        block = batch.texts(names)
        dev = batch.on_device("texts", names, block)
        y = batch.label()
        y = y if y.is_cuda else y.cuda()
        classes, class_counts, ids = label_info_dev(batch, y, label_name)
        n_all = int(D.allreduce_sum(np.array([batch.n], dtype=np.int64))[0])
        cont = None if classes is not None else label_moments_dev(batch, y, label_name) + (y,)
        return self._sidecar(dev, batch.n, n_all, classes, class_counts, ids, label_name, cont)

    # -- packed cells: what the statistics pass found in every cell of a chunk, kept in HBM for the chunk's transform
    def enable_cell_cache(self, nbytes: int) -> bool:
        """Give the streamed fit an arena of `nbytes` of device memory for the packed cells of the chunks
        (tmog_text_hash_stats_pack_dev): the transform of a chunk whose cells fitted reads them instead of the text."""
        lay = self.layout
        self._cells = None
        if nbytes < (1 << 16) or not lay.text_idx or lay.shared or lay.num_features > 512:
            return False
        nbytes -= nbytes % 256
        self._cells = {"arena": torch.empty(nbytes, dtype=torch.uint8, device="cuda"),
                       "cursor": torch.zeros(1, dtype=torch.int64, device="cuda"), "chunks": {}, "pending": []}
        return True

    def stream_cached(self, index: int, token) -> bool:
        """Chunk number `index` of the stream (with this token) needs none of its text columns for the transform."""
        cells = getattr(self, "_cells", None)
        return bool(cells) and not self.pivot_idx and (index, token) in cells["chunks"]

    def release_cell_cache(self) -> None:
        self._cells = None

    # -- row chunks (streaming.py): the statistics sidecar accumulated over the chunks
    def stream_begin(self, label):
        if label is None:
            return {"label": None}
        st = self._side_begin(label.classes, label.classes is None)
        st["label"], st["n"] = label, 0
        return st

    def stream_update(self, st, chunk: ColumnarBatch) -> None:
        lab = st["label"]
        if lab is None:
            return
        names = [f.name for f in self.inN]
        block = chunk.texts(names)
        dev = chunk.on_device("texts", names, block)
        ids = lab.class_ids(chunk) if lab.classes is not None else None
        cont = None if lab.classes is not None else lab.continuous() + (lab.device_label(chunk),)
        st["pack_token"] = (st.get("stream_index", 0), chunk.token)      # the chunk's place in the stream and its token
        st["stream_index"] = st.get("stream_index", 0) + 1
        self._side_scan(st, dev, chunk.n, ids, cont)
        st["pack_token"] = None
        st.setdefault("tokens", []).append(chunk.token)
        st["n"] += chunk.n

    def stream_finish(self, st):
        lab = st["label"]
        if lab is None:
            return self
        cont = None if lab.classes is not None else lab.continuous() + (None,)
        self._fit_stats = self._side_finish(st, lab.n, lab.classes, lab.class_counts, lab.name, cont)
        if st["chunk_nnz"]:             # sparse entries of every chunk's rows, by the chunk's token: sizes its transform
            nnz = torch.stack(st["chunk_nnz"]).cpu().numpy().astype(np.int64)
            self._stream_nnz = dict(zip(st.get("tokens", []), nnz.tolist()))
        cells = getattr(self, "_cells", None)
        if cells and cells["pending"]:  # the chunks whose cells fitted the arena
            bad = torch.cat([f for _t, _m, f, _n, _k in cells["pending"]]).cpu().numpy()
            for (token, meta, _f, n, k), b in zip(cells["pending"], bad):
                if not b:
                    cells["chunks"][token] = (meta, n, k)
            cells["pending"] = []
        return self

    # -- transform -----------------------------------------------------------------------------------------------------
    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        R.require_cuda()
        names = [f.name for f in self.inN]
        block = batch.texts(names)
        n = batch.n
        parts = []
        same = self._fit_key is not None and self._fit_key == batch_key(batch)
        dev_of = batch.device_thunk("texts", names, block)
        # what the fit pass left on this very batch (sparse rows of all columns, entries per vector index)
        fitted = batch.__dict__.get("_dev_cache", {}).get(("fit_rows", self.uid)) if same else None
        stream = batch.__dict__.get("_stream")        # row chunk of a streamed dataset: (chunk index, shared flags)
        if self.pivot_idx:
            parts.append(LazyPart(self.pivot_width, lambda keep: self._materialize_pivot(dev_of(), n, keep, stream)))
        if self.layout.text_idx:
            parts.append(LazyPart(self.layout.width,
                                  lambda keep: self._materialize_hash(dev_of, n, keep, fitted, stream)))
        return VectorColumn(self.getOutput().name, n, self.getMetadata(), parts, self._fit_stats if same else None)

    def _plan(self, kind: str, keep, build):
        """Device tables of a transform that depend on the kept columns only: built once per keep set."""
        key = (kind, torch.cuda.current_device(), None if keep is None else np.asarray(keep).tobytes())
        cache = self.__dict__.setdefault("_plans", {})
        if key not in cache:
            if len(cache) > 8:
                cache.clear()
            cache[key] = build()
        return cache[key]

    def _materialize_pivot(self, block: TextBlock, n: int, keep, stream=None) -> OneHotPart:
        dev = block if block.is_cuda else block.to("cuda")
        width = self.pivot_width
        k = len(self.pivot_idx)

        def build():
            newpos = np.arange(width, dtype=np.int32)
            out_width = width
            if keep is not None:
                newpos = np.full(width, -1, dtype=np.int32)
                newpos[np.asarray(keep, dtype=np.int64)] = np.arange(len(keep), dtype=np.int32)
                out_width = len(keep)
            _tables, slots, slot_of, _sel = self._pivot_table_build(len(self.inN))
            pos, desc = self._pivot_pos_tables(slots, slot_of, newpos)
            return torch.from_numpy(desc).cuda(), torch.from_numpy(pos).cuda(), out_width

        d_desc, d_pos, out_width = self._plan("pivot", keep, build)
        codes, ld, slots, _slot_of = self._pivot_tables(dev, n, None if stream is None else stream["flags_clean"])
        codes = torch.where(codes == -2, torch.full_like(codes, slots), codes)
        out = torch.empty((n, max(k, 1)), dtype=torch.int32, device="cuda")
        L.call("tmog_pivot_transform_dev", codes.data_ptr(), ld, n, k, d_desc.data_ptr(), d_pos.data_ptr(),
               out.data_ptr(), out.shape[1], 0, R.stream())
        return OneHotPart(out[:, :k], out_width)

    def _materialize_hash(self, block, n: int, keep, fitted=None, stream=None) -> SparseRowsPart:
        """block: the text block, or a thunk that fetches it (not called when the chunk's packed cells are at hand)."""
        lay = self.layout
        width = lay.width
        full = lay.kernel_width

        def build():
            newpos = np.full(full, -1, dtype=np.int32)
            if keep is None:
                newpos[:width] = np.arange(width, dtype=np.int32)
                out_width = width
            else:
                newpos[np.asarray(keep, dtype=np.int64)] = np.arange(len(keep), dtype=np.int32)
                out_width = len(keep)
            return newpos, torch.from_numpy(newpos).cuda(), out_width

        newpos, d_newpos, out_width = self._plan("hash", keep, build)
        cells = getattr(self, "_cells", None)
        packed = cells["chunks"].get((stream.get("index"), stream.get("token"))) \
            if cells and stream is not None and fitted is None else None
        if packed is not None and packed[1] == n:
            return self._rows_from_cells(cells, packed, n, d_newpos, out_width, stream)
        if callable(block):
            block = block()
        dev = block if block.is_cuda else block.to("cuda")
        rows, local_nnz = fitted if fitted is not None else (None, None)
        if rows is not None:
            # the fit pass wrote the rows of this batch: keep the wanted columns of them
            if keep is None or (len(keep) == width and bool((np.asarray(keep) == np.arange(width)).all())):
                return SparseRowsPart(rows.row_start, rows.row_nnz, rows.indices, rows.values, width)
            out_nnz = torch.empty(max(n, 1), dtype=torch.int32, device="cuda")
            args = (rows.row_start.data_ptr(), rows.row_nnz.data_ptr(), rows.indices.data_ptr(),
                    rows.values.data_ptr(), n, d_newpos.data_ptr())
            L.call("tmog_sparse_rows_select_dev", *args, None, out_nnz.data_ptr(), None, None, R.stream())
            incl = torch.cumsum(out_nnz.to(torch.int64), 0)
            out_start = (incl - out_nnz).contiguous()
            total = int(incl[n - 1].item()) if n else 0
            oi = torch.empty(max(total, 1), dtype=torch.int32, device="cuda")
            ov = torch.empty(max(total, 1), dtype=torch.int32, device="cuda")
            L.call("tmog_sparse_rows_select_dev", *args, out_start.data_ptr(), out_nnz.data_ptr(), oi.data_ptr(),
                   ov.data_ptr(), R.stream())
            return SparseRowsPart(out_start[:n], out_nnz[:n], oi, ov, out_width)
        sub, d_prefix, d_plen, params = self._hash_args(dev)
        params[7] = 1
        params[6] = lay.null_base
        desc = _col_desc(sub)
        k = len(sub.names)
        row_start = torch.empty(n, dtype=torch.int64, device="cuda")
        row_nnz = torch.empty(n, dtype=torch.int32, device="cuda")
        own_flags = stream is None
        flags = torch.zeros(4, dtype=torch.int32, device="cuda") if own_flags else stream["flags"]
        cursor = torch.zeros(1, dtype=torch.int64, device="cuda")
        scratch = _hash_scratch(n, k, lay, True)
        chunk_nnz = getattr(self, "_stream_nnz", None)
        if local_nnz is not None:              # exact size known from the fit pass over the same rows
            capacity = int(local_nnz[newpos[:len(local_nnz)] >= 0].sum())
        elif stream is not None and chunk_nnz is not None and stream.get("token") in chunk_nnz:
            capacity = int(chunk_nnz[stream["token"]])     # entries of these very rows over all columns (fit pass)
        else:                                  # sizing pass
            L.call("tmog_text_hash_transform_dev", desc.data_ptr(), k, n, d_prefix.data_ptr(),
                   d_plen.data_ptr(), params.ctypes.data, d_newpos.data_ptr(), row_start.data_ptr(),
                   row_nnz.data_ptr(), None, None, cursor.data_ptr(), 0, flags.data_ptr(), scratch.data_ptr(),
                   scratch.numel(), R.stream())
            capacity = int(cursor.item())
            cursor.zero_()
            flags[3] = 0
        idx = torch.empty(max(capacity, 1), dtype=torch.int32, device="cuda")
        val = torch.empty(max(capacity, 1), dtype=torch.int32, device="cuda")
        L.call("tmog_text_hash_transform_dev", desc.data_ptr(), k, n, d_prefix.data_ptr(),
               d_plen.data_ptr(), params.ctypes.data, d_newpos.data_ptr(), row_start.data_ptr(), row_nnz.data_ptr(),
               idx.data_ptr(), val.data_ptr(), cursor.data_ptr(), capacity, flags.data_ptr(), scratch.data_ptr(),
               scratch.numel(), R.stream())
        if own_flags:
            _check_hash_flags(flags)
        return SparseRowsPart(row_start, row_nnz, idx, val, out_width, cursor)


    def _rows_from_cells(self, cells, packed, n: int, d_newpos, out_width: int, stream) -> SparseRowsPart:
        """The sparse rows of a chunk from its packed cells (tmog_text_hash_transform_packed_dev)."""
        meta, _n, k = packed
        lay = self.layout
        hp = self.hashingParams
        params = np.array([lay.num_features, int(lay.shared), int(hp["binaryFreq"]), hp["minTokenLength"],
                           int(self.shouldTrackNulls), int(hp["prependFeatureName"]), lay.null_base, 1,
                           len(lay.hash_idx) if lay.ignore_idx else 0, lay.len_base if lay.len_width else 0, 0, 0],
                          dtype=np.int32)
        row_start = torch.empty(n, dtype=torch.int64, device="cuda")
        row_nnz = torch.empty(n, dtype=torch.int32, device="cuda")
        flags = stream["flags"]
        cursor = torch.zeros(1, dtype=torch.int64, device="cuda")
        chunk_nnz = getattr(self, "_stream_nnz", None) or {}
        args = (k, n, params.ctypes.data, meta.data_ptr(), cells["arena"].data_ptr(), cells["arena"].numel(),
                d_newpos.data_ptr(), row_start.data_ptr(), row_nnz.data_ptr())
        if stream.get("token") in chunk_nnz:
            capacity = int(chunk_nnz[stream["token"]])
        else:
            L.call("tmog_text_hash_transform_packed_dev", *args, None, None, cursor.data_ptr(), 0, flags.data_ptr(),
                   R.stream())
            capacity = int(cursor.item())
            cursor.zero_()
            flags[3] = 0
        idx = torch.empty(max(capacity, 1), dtype=torch.int32, device="cuda")
        val = torch.empty(max(capacity, 1), dtype=torch.int32, device="cuda")
        L.call("tmog_text_hash_transform_packed_dev", *args, idx.data_ptr(), val.data_ptr(), cursor.data_ptr(), capacity,
               flags.data_ptr(), R.stream())
        return SparseRowsPart(row_start, row_nnz, idx, val, out_width, cursor)


def _hash_scratch(n: int, ncols: int = 0, lay: Optional[_HashLayout] = None, sparse_out: bool = True) -> torch.Tensor:
    """Device scratch of the tokenise + hash entry points for n rows of ncols text columns."""
    import ctypes
    need = ctypes.c_size_t(0)
    if lay is None:
        L.call("tmog_text_hash_scratch_bytes", n, ctypes.byref(need))
    else:
        params = np.array([lay.num_features, int(lay.shared), 0, 0, 0, 0, 0, 0,
                           len(lay.hash_idx) if lay.ignore_idx else 0, 0, 0, 0], dtype=np.int32)
        L.call("tmog_text_hash_scratch_bytes2", n, ncols, params.ctypes.data, int(sparse_out), ctypes.byref(need))
    return torch.empty(max(need.value, 8), dtype=torch.uint8, device="cuda")


def _check_hash_flags(flags: torch.Tensor) -> None:
    f = flags.cpu().numpy()
    if f[0]:
        raise ValueError("non-ASCII text met: Lucene's Unicode word-break rules are restated for ASCII only")
    if f[1]:
        raise ValueError("a run of letters / digits longer than ~4000 bytes: beyond the GPU tokenizer's text buffer")
    if f[2]:
        raise ValueError("a row holds more than ~12000 distinct hashed tokens: beyond the GPU path's per-row table")
    if f[3]:
        raise RuntimeError("sparse output capacity exceeded")


def _stats_from_bucket_sums(class_sum: np.ndarray, nnz: np.ndarray, sq_extra: np.ndarray, aux: np.ndarray, classes,
                            class_counts, n: int, label_name: str, lay: _HashLayout) -> VectorStats:
    """Column statistics of term-frequency (and text-length) columns from per-class sums:
    sum v = sum_c S_c, sum v^2 = sum v + extra, cov(v, y) = sum_c y_c S_c - mean_y sum v."""
    cc = class_counts.astype(np.float64)
    mean_y = float((classes * cc).sum() / n)
    var_y = float((cc * (classes - mean_y) ** 2).sum() / (n - 1)) if n > 1 else 0.0
    lab = np.array([n, mean_y, var_y, classes.min(), classes.max()])
    # values are >= 1 where present; aux = [max v, ~min v among v > 1, rows with v > 1]
    w = class_sum.shape[0]
    cs64 = np.ascontiguousarray(class_sum, dtype=np.float64)
    nz64, ex64 = np.ascontiguousarray(nnz, dtype=np.float64), np.ascontiguousarray(sq_extra, dtype=np.float64)
    ax64 = np.ascontiguousarray(aux, dtype=np.int64)
    cl64 = np.ascontiguousarray(classes, dtype=np.float64)
    mean, var, mn, mx, cov = (np.empty(w) for _ in range(5))
    L.call("tmog_bucket_stats", w, len(cl64), cs64.ctypes.data, nz64.ctypes.data, ex64.ctypes.data, ax64.ctypes.data,
           cl64.ctypes.data, float(n), mean_y, mean.ctypes.data, var.ctypes.data, mn.ctypes.data, mx.ctypes.data,
           cov.ctypes.data)
    cs = np.full((len(classes), w), np.nan)
    cs[:, lay.null_base:] = class_sum[lay.null_base:].T     # only the null indicators are categorical columns
    return VectorStats(LabelInfo(label_name, n, lab, classes, class_counts), mean, var, mn, mx, cov, cs)
