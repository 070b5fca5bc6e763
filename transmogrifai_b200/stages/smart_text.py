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
from ..columnar import ColumnarBatch, LazyPart, OneHotPart, SparseRowsPart, TextBlock, VectorColumn, encode_strings
from ..metadata import NullString, OpVectorColumnMetadata, OpVectorMetadata, OtherString, feature_column
from ..stats import LabelInfo, VectorStats
from .base import OpEstimator, OpModel
from .numeric import label_info_dev, label_moments_dev, stats_from_sums_continuous
from .pivot import _stats_from_class_counts, clean_and_group, top_k

PIVOT, HASH, IGNORE = "Pivot", "Hash", "Ignore"
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
    """Geometry of the hash + null-indicator part of the output vector."""

    def __init__(self, names: Sequence[str], hash_idx: List[int], ignore_idx: List[int], num_features: int,
                 num_inputs: int, strategy: str, prepend: bool, track_nulls: bool):
        s = strategy.lower()
        self.shared = True if s == "shared" else False if s == "separate" else \
            num_features * num_inputs > MAX_NUM_OF_FEATURES          # isSharedHashSpace :191-199
        self.num_features = num_features
        self.hash_idx, self.ignore_idx = hash_idx, ignore_idx
        self.text_idx = hash_idx + ignore_idx                       # null indicators: hash columns, then ignored
        self.hash_width = 0 if not hash_idx else (num_features if self.shared else num_features * len(hash_idx))
        self.null_width = len(self.text_idx) if track_nulls else 0
        self.width = self.hash_width + self.null_width
        # the reference zips the NON-PIVOT features (hash and ignore, input order) with the hash columns (:348,358-361)
        non_pivot = sorted(self.text_idx)
        self.prefix_names = [names[non_pivot[k]] for k in range(len(hash_idx))]
        self.prepend = prepend


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
        if v:
            raise NotImplementedError("trackTextLen is not on the GPU path (default false)")
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
            model._fit_stats = model._sidecar(dev, n, n_all, classes, class_counts, ids, label_name, cont)
        model._fit_key = (block.offsets[0].data_ptr(), n)
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
        elif self.minLengthStdDev > 0:
            raise NotImplementedError("minLengthStdDev > 0 (the Ignore route) is not on the GPU path")
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
        self.layout = _HashLayout(names, hash_idx, ignore_idx, p.numFeatures, len(names), p.hashSpaceStrategy,
                                  p.prependFeatureName, p.trackNulls)
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
            # the reference appends the null-indicator columns to the metadata whether or not nulls are tracked
            # (SmartTextVectorizer.scala:176-183); kept as is, SanityChecker then rejects the width mismatch
            runs += [(feature_column(feats[j].name, feats[j].type_name, is_null=True), 1) for j in lay.text_idx]
        return OpVectorMetadata(self.parent.getOutputFeatureName(), history=self.parent.history(), runs=runs)

    # -- device tables -------------------------------------------------------------------------------------------------
    def _pivot_tables(self, dev: TextBlock, n: int):
        """Small hash tables holding just the top values of every pivoted column, and the dictionary codes of all
        rows against them (code -2 = a value outside the vocabulary -> OTHER)."""
        k = len(self.pivot_idx)
        slots = 64
        while slots < 4 * (max((len(self.topValues[j]) for j in self.pivot_idx), default=0) + 2):
            slots <<= 1
        head, per = C.c_size_t(), C.c_size_t()
        ncols = len(dev.names)
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
        ld = (n + 3) // 4 * 4
        codes = torch.empty((max(k, 1), ld), dtype=torch.int32, device="cuda")
        sel = torch.tensor(self.pivot_idx or [0], dtype=torch.int32, device="cuda")
        desc = _col_desc(dev)
        flags = torch.zeros(4, dtype=torch.int32, device="cuda")
        L.call("tmog_text_codes_dev", desc.data_ptr(), sel.data_ptr(), k, n, int(self.shouldCleanText), slots,
               tables.data_ptr(), codes.data_ptr(), ld, flags.data_ptr(), R.stream())
        if k and self.shouldCleanText and int(flags[0].item()):
            raise ValueError("non-ASCII text met with cleanText = true: Java's Unicode case mapping is not restated on "
                             "this path")
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
        prefix = np.zeros((max(k, 1), 16), dtype=np.uint8)
        plen = np.zeros(max(k, 1), dtype=np.int32)
        for t in range(k):
            s = f"{murmur3_index(lay.prefix_names[t], lay.num_features)}_".encode("ascii")
            prefix[t, :len(s)] = np.frombuffer(s, dtype=np.uint8)
            plen[t] = len(s)
        sub = dev.rows(lay.hash_idx)
        params = np.array([lay.num_features, int(lay.shared), int(hp["binaryFreq"]), hp["minTokenLength"],
                           int(self.shouldTrackNulls), int(hp["prependFeatureName"]), lay.hash_width, 0],
                          dtype=np.int32)
        return sub, torch.from_numpy(prefix).cuda(), torch.from_numpy(plen).cuda(), params

    # -- statistics sidecar --------------------------------------------------------------------------------------------
    def _sidecar(self, dev: TextBlock, n: int, n_all: int, classes, class_counts, ids, label_name: str,
                 cont=None) -> VectorStats:
        """n = local rows, n_all = rows of all ranks; the count tables and bucket sums are all-reduced.
        cont = (label moments, shift, label on the device) for a continuous label (then classes is None)."""
        ncls = len(classes) if classes is not None else 1
        ids_ptr = ids.data_ptr() if classes is not None else None
        parts = []
        if self.pivot_idx:
            codes, ld, slots, slot_of = self._pivot_tables(dev, n)
            # count per slot code: remap -2 (unknown value) to slot index `slots` so that it lands in OTHER
            codes = torch.where(codes == -2, torch.full_like(codes, slots), codes)
            k = len(self.pivot_idx)
            nids = slots + 2
            desc = np.zeros((k, 4), dtype=np.int64)
            for t in range(k):
                desc[t] = [nids, 0, t * nids * ncls, slots + 1]
            d_desc = torch.from_numpy(desc).cuda()
            d_counts = torch.zeros(k * nids * ncls, dtype=torch.int32, device="cuda")
            d_bad = torch.zeros(1, dtype=torch.int32, device="cuda")
            L.call("tmog_pivot_count_dev", codes.data_ptr(), ld, n, k, ids_ptr, ncls, d_desc.data_ptr(), None,
                   nids, d_counts.data_ptr(), d_bad.data_ptr(), R.stream())
            cnt = D.allreduce_sum(d_counts.cpu().numpy().astype(np.int64) & 0xFFFFFFFF).reshape(k, nids, ncls)
            ws = None
            if cont is not None:
                d_ws = torch.zeros(k * nids, dtype=torch.float64, device="cuda")
                d_woff = torch.arange(0, k * nids, nids, dtype=torch.int64, device="cuda")
                L.call("tmog_pivot_label_sum_dev", codes.data_ptr(), ld, n, k, d_desc.data_ptr(), None,
                       d_woff.data_ptr(), cont[2].data_ptr(), cont[1], d_ws.data_ptr(), R.stream())
                ws = D.allreduce_sum(d_ws.cpu().numpy()).reshape(k, nids)
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
            if lay.ignore_idx:
                raise NotImplementedError("ignored text columns are not on the GPU path")
            sub, d_prefix, d_plen, params = self._hash_args(dev)
            params[7] = ncls
            w = lay.hash_width + len(lay.hash_idx)
            class_sum = torch.zeros(w * ncls, dtype=torch.int32, device="cuda")
            dup = torch.zeros(w, dtype=torch.int32, device="cuda")
            extra = torch.zeros(w, dtype=torch.int64, device="cuda")
            scratch = _hash_scratch(n)
            aux = torch.zeros(3 * w, dtype=torch.int32, device="cuda")
            flags = torch.zeros(4, dtype=torch.int32, device="cuda")
            desc = _col_desc(sub)
            wsum = torch.zeros(w, dtype=torch.float64, device="cuda") if cont is not None else None
            # One pass over the text gives the statistics AND the sparse rows of this batch (the workflow transforms
            # its training rows right after the fit): the output is sized from a sample of the first rows; if the
            # estimate turns out too small the rows are dropped and the transform tokenises again.
            k = len(lay.hash_idx)
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
            L.call("tmog_text_hash_fit_transform_dev", desc.data_ptr(), k, n, d_prefix.data_ptr(), d_plen.data_ptr(),
                   params.ctypes.data, ids_ptr, class_sum.data_ptr(), dup.data_ptr(), extra.data_ptr(), aux.data_ptr(),
                   cont[2].data_ptr() if cont is not None else None, cont[1] if cont is not None else 0.0,
                   wsum.data_ptr() if cont is not None else None, ident.data_ptr(), row_start.data_ptr(),
                   row_nnz.data_ptr(), idx.data_ptr(), val.data_ptr(), cursor.data_ptr(), capacity, flags.data_ptr(),
                   scratch.data_ptr(), scratch.numel(), R.stream())
            too_small = int(flags[3].item()) != 0
            flags[3] = 0
            _check_hash_flags(flags)
            self._fit_rows = None if too_small else SparseRowsPart(row_start[:n], row_nnz[:n], idx, val, w)
            # the four bucket arrays widened on the device and brought back in one copy (unsigned 32-bit words read
            # as int64, the 64-bit extra sums as they are): a 262 144-wide block costs milliseconds of host casts
            u32 = lambda t: t.view(torch.int32).to(torch.int64) & 0xFFFFFFFF   # noqa: E731
            packed = torch.cat([u32(class_sum).reshape(-1), u32(dup).reshape(-1), extra.view(torch.int64).reshape(-1),
                                u32(aux).reshape(-1)]).cpu().numpy()
            k = w * ncls
            cs = packed[:k].reshape(w, ncls).astype(np.float64)
            # rows holding an index = sum of its term frequencies - (tf - 1) of the rows with tf > 1
            nz = cs.sum(axis=1) - packed[k:k + w].astype(np.float64)
            ex = packed[k + w:k + 2 * w].astype(np.float64)
            ax = packed[k + 2 * w:].reshape(w, 3)
            if not self.shouldTrackNulls:
                cs, nz, ex, ax = cs[:lay.hash_width], nz[:lay.hash_width], ex[:lay.hash_width], ax[:lay.hash_width]
            self._hash_nnz = nz                      # of the local rows: sizes the sparse output of the same batch
            if D.world()[1] > 1:
                # two exchanges in all: everything that adds (counts below 2^53, exact in f64), everything that maxes
                sums = D.allreduce_sum(np.concatenate([cs.reshape(-1), nz, ex, ax[:, 2].astype(np.float64)]))
                maxs = D.allreduce_max(np.concatenate([ax[:, 0], ax[:, 1]]))
                k, m = cs.size, len(nz)
                cs, nz, ex = sums[:k].reshape(cs.shape), sums[k:k + m], sums[k + m:k + 2 * m]
                ax = np.stack([maxs[:m], maxs[m:], sums[k + 2 * m:].astype(np.int64)], axis=1)
            if cont is None:
                parts.append(_stats_from_bucket_sums(cs, nz, ex, ax, classes, class_counts, n_all, label_name,
                                                     lay.hash_width))
            else:
                wv = D.allreduce_sum(wsum.cpu().numpy()[:len(nz)])
                sv = cs[:, 0]
                vmax = np.maximum(ax[:, 0], 1).astype(np.float64)
                vmin = np.where(ax[:, 2] >= nz, ((~ax[:, 1]) & 0xFFFFFFFF), 1).astype(np.float64)
                parts.append(stats_from_sums_continuous(sv, sv + ex, nz, vmax, vmin, wv, cont[0], cont[1], label_name))
        return VectorStats.concat(parts)

    # -- transform -----------------------------------------------------------------------------------------------------
    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        R.require_cuda()
        names = [f.name for f in self.inN]
        block = batch.texts(names)
        n = batch.n
        parts = []
        same = self._fit_key == (block.offsets[0].data_ptr(), n)
        dev_of = batch.device_thunk("texts", names, block)
        if self.pivot_idx:
            parts.append(LazyPart(self.pivot_width, lambda keep: self._materialize_pivot(dev_of(), n, keep)))
        if self.layout.text_idx:
            parts.append(LazyPart(self.layout.width, lambda keep: self._materialize_hash(dev_of(), n, keep, same)))
        return VectorColumn(self.getOutput().name, n, self.getMetadata(), parts, self._fit_stats if same else None)

    def _materialize_pivot(self, block: TextBlock, n: int, keep) -> OneHotPart:
        dev = block if block.is_cuda else block.to("cuda")
        width = self.pivot_width
        newpos = np.arange(width, dtype=np.int32)
        out_width = width
        if keep is not None:
            newpos = np.full(width, -1, dtype=np.int32)
            newpos[np.asarray(keep, dtype=np.int64)] = np.arange(len(keep), dtype=np.int32)
            out_width = len(keep)
        codes, ld, slots, slot_of = self._pivot_tables(dev, n)
        codes = torch.where(codes == -2, torch.full_like(codes, slots), codes)
        pos, desc = self._pivot_pos_tables(slots, slot_of, newpos)
        k = len(self.pivot_idx)
        out = torch.empty((n, max(k, 1)), dtype=torch.int32, device="cuda")
        d_desc, d_pos = torch.from_numpy(desc).cuda(), torch.from_numpy(pos).cuda()
        L.call("tmog_pivot_transform_dev", codes.data_ptr(), ld, n, k, d_desc.data_ptr(), d_pos.data_ptr(),
               out.data_ptr(), out.shape[1], 0, R.stream())
        return OneHotPart(out[:, :k], out_width)

    def _materialize_hash(self, block: TextBlock, n: int, keep, same_batch: bool) -> SparseRowsPart:
        dev = block if block.is_cuda else block.to("cuda")
        lay = self.layout
        if lay.ignore_idx:
            raise NotImplementedError("ignored text columns are not on the GPU path")
        width = lay.width
        full = lay.hash_width + len(lay.hash_idx)
        newpos = np.full(full, -1, dtype=np.int32)
        if keep is None:
            newpos[:width] = np.arange(width, dtype=np.int32)
            out_width = width
        else:
            newpos[np.asarray(keep, dtype=np.int64)] = np.arange(len(keep), dtype=np.int32)
            out_width = len(keep)
        d_newpos = torch.from_numpy(newpos).cuda()
        rows = getattr(self, "_fit_rows", None)
        if same_batch and rows is not None:
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
        desc = _col_desc(sub)
        row_start = torch.empty(n, dtype=torch.int64, device="cuda")
        row_nnz = torch.empty(n, dtype=torch.int32, device="cuda")
        flags = torch.zeros(4, dtype=torch.int32, device="cuda")
        cursor = torch.zeros(1, dtype=torch.int64, device="cuda")
        scratch = _hash_scratch(n)
        nz = getattr(self, "_hash_nnz", None)
        if same_batch and nz is not None:      # exact size known from the fit pass over the same rows
            capacity = int(nz[newpos[:len(nz)] >= 0].sum())
        else:                                  # sizing pass
            L.call("tmog_text_hash_transform_dev", desc.data_ptr(), len(lay.hash_idx), n, d_prefix.data_ptr(),
                   d_plen.data_ptr(), params.ctypes.data, d_newpos.data_ptr(), row_start.data_ptr(),
                   row_nnz.data_ptr(), None, None, cursor.data_ptr(), 0, flags.data_ptr(), scratch.data_ptr(),
                   scratch.numel(), R.stream())
            capacity = int(cursor.item())
            cursor.zero_()
            flags.zero_()
        idx = torch.empty(max(capacity, 1), dtype=torch.int32, device="cuda")
        val = torch.empty(max(capacity, 1), dtype=torch.int32, device="cuda")
        L.call("tmog_text_hash_transform_dev", desc.data_ptr(), len(lay.hash_idx), n, d_prefix.data_ptr(),
               d_plen.data_ptr(), params.ctypes.data, d_newpos.data_ptr(), row_start.data_ptr(), row_nnz.data_ptr(),
               idx.data_ptr(), val.data_ptr(), cursor.data_ptr(), capacity, flags.data_ptr(), scratch.data_ptr(),
               scratch.numel(), R.stream())
        _check_hash_flags(flags)
        return SparseRowsPart(row_start, row_nnz, idx, val, out_width)


def _hash_scratch(n: int) -> torch.Tensor:
    import ctypes
    need = ctypes.c_size_t(0)
    L.call("tmog_text_hash_scratch_bytes", n, ctypes.byref(need))
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
                            class_counts, n: int, label_name: str, hash_width: int) -> VectorStats:
    """Column statistics of term-frequency columns from per-class sums: sum v = sum_c S_c,
    sum v^2 = sum v + extra, cov(v, y) = sum_c y_c S_c - mean_y sum v."""
    cc = class_counts.astype(np.float64)
    mean_y = float((classes * cc).sum() / n)
    var_y = float((cc * (classes - mean_y) ** 2).sum() / (n - 1)) if n > 1 else 0.0
    lab = np.array([n, mean_y, var_y, classes.min(), classes.max()])
    # term frequencies are >= 1 where present; aux = [max tf, ~min tf among tf > 1, rows with tf > 1]
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
    cs[:, hash_width:] = class_sum[hash_width:].T          # null indicators are categorical columns
    return VectorStats(LabelInfo(label_name, n, lab, classes, class_counts), mean, var, mn, mx, cov, cs)
