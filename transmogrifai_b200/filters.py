"""RawFeatureFilter: the pre-DAG check of the raw features' distributions, training against scoring
data -- core/src/main/scala/com/salesforce/op/filters/RawFeatureFilter.scala, FeatureDistribution.scala, Summary.scala,
PreparedFeatures.scala -- with the reference's names, argument meaning, defaults (OpWorkflow.withRawFeatureFilter,
core/.../OpWorkflow.scala:537-552) and `require` messages.

The reference maps every row to per-feature Summaries, then every row again to one FeatureDistribution per feature, and
reduces; here every feature block of a batch (or of every row chunk of a dataset) goes through one column scan per
pass (csrc/rff.cu): numeric summaries, numeric histograms over the training summary's splits, hashed-token
distributions of Text / TextArea columns (the tokenise + hash kernels of csrc/text.cu, then a per-bucket sum) and of
PickList columns (rows counted per dictionary code, the dictionary's entries tokenised once), and the sums behind the
null-indicator / response correlation.  What is left for the host is what the reference's driver does: metrics,
exclusion reasons, the names to drop.

Covered feature types: Real, RealNN, Integral (Date, DateTime as plain numbers: timePeriod = None), Binary, Text,
TextArea, PickList, MultiPickList.  Not covered (raise): maps, lists, geolocations, vectors, a custom textBinsFormula,
timePeriod.  Not produced: FeatureDistribution.moments / cardEstimate (reported by the reference, read by no rule).
Order of the distributions: responses, then predictors, in the order of `features` (the reference's is the iteration
order of a Scala Map: unspecified beyond four features).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

from . import _lib as L
from . import features as F

FeatureKey = Tuple[str, Optional[str]]
MAX_BINS = 100000                      # FeatureDistribution.MaxBins (FeatureDistribution.scala:192)
MIN_SCORING_ROWS_DEFAULT = 500         # RawFeatureFilter.minScoringRowsDefault (RawFeatureFilter.scala:590)


@dataclass
class Summary:
    """Summary.scala:43: min / max / sum / count of a numeric feature's values, or of a text's token counts."""
    min: float = float("inf")
    max: float = float("-inf")
    sum: float = 0.0
    count: float = 0.0

    def as_array(self) -> np.ndarray:
        return np.array([self.min, self.max, self.sum, self.count], dtype=np.float64)


@dataclass
class FeatureDistribution:
    """FeatureDistribution.scala:59-69."""
    name: str
    key: Optional[str]
    count: int
    nulls: int
    distribution: np.ndarray
    summaryInfo: np.ndarray  # noqa: N815
    type: str = "Training"

    def __post_init__(self):
        self.distribution = np.asarray(self.distribution, dtype=np.float64)
        self.summaryInfo = np.asarray(self.summaryInfo, dtype=np.float64)

    @property
    def featureKey(self) -> FeatureKey:  # noqa: N802
        return (self.name, self.key)

    def _check_match(self, fd: "FeatureDistribution") -> None:             # :82-88
        for fld, a, b in (("Name", self.name, fd.name), ("Key", self.key, fd.key)):
            if a != b:
                raise ValueError(f"requirement failed: {fld} must match to compare or combine feature distributions: "
                                 f"{a} != {b}")

    def fillRate(self) -> float:  # noqa: N802  (:99)
        return 0.0 if self.count == 0 else (self.count - self.nulls) / float(self.count)

    def reduce(self, fd: "FeatureDistribution") -> "FeatureDistribution":  # :107-122
        self._check_match(fd)
        a, b = self.distribution, fd.distribution
        if len(a) < len(b):
            a, b = b, a
        dist = a.copy()
        dist[:len(b)] += b                                   # algebird's array monoid pads the shorter one
        info = self.summaryInfo if len(self.summaryInfo) > len(fd.summaryInfo) else fd.summaryInfo
        return FeatureDistribution(self.name, self.key, self.count + fd.count, self.nulls + fd.nulls, dist, info,
                                   self.type)

    def relativeFillRatio(self, fd: "FeatureDistribution") -> float:  # noqa: N802  (:130-135)
        self._check_match(fd)
        a, b = self.fillRate(), fd.fillRate()
        small, large = (a, b) if a < b else (b, a)
        return float("inf") if small == 0.0 else large / small

    def relativeFillRate(self, fd: "FeatureDistribution") -> float:  # noqa: N802  (:143-146)
        self._check_match(fd)
        return abs(self.fillRate() - fd.fillRate())

    def jsDivergence(self, fd: "FeatureDistribution") -> float:  # noqa: N802  (:154-163)
        self._check_match(fd)
        m = min(len(self.distribution), len(fd.distribution))          # zip stops at the shorter one
        a = np.ascontiguousarray(self.distribution[:m])
        b = np.ascontiguousarray(fd.distribution[:m])
        out = C.c_double(0.0)
        L.call("tmog_rff_js_divergence", a.ctypes.data, b.ctypes.data, m, C.byref(out))
        return out.value


@dataclass
class AllFeatureInformation:
    """AllFeatureInformation.scala:44-52."""
    responseSummaries: Dict[FeatureKey, Summary]  # noqa: N815
    responseDistributions: List[FeatureDistribution]  # noqa: N815
    predictorSummaries: Dict[FeatureKey, Summary]  # noqa: N815
    predictorDistributions: List[FeatureDistribution]  # noqa: N815
    correlationInfo: Dict[FeatureKey, Dict[FeatureKey, float]]  # noqa: N815


@dataclass
class RawFeatureFilterMetrics:
    """RawFeatureFilterResults.scala:117-127."""
    name: str
    key: Optional[str]
    trainingFillRate: float  # noqa: N815
    trainingNullLabelAbsoluteCorr: Optional[float]  # noqa: N815
    scoringFillRate: Optional[float]  # noqa: N815
    jsDivergence: Optional[float]  # noqa: N815
    fillRateDiff: Optional[float]  # noqa: N815
    fillRatioDiff: Optional[float]  # noqa: N815


@dataclass
class ExclusionReasons:
    """RawFeatureFilterResults.scala:146-157."""
    name: str
    key: Optional[str]
    trainingUnfilledState: bool  # noqa: N815
    trainingNullLabelLeaker: bool  # noqa: N815
    scoringUnfilledState: bool  # noqa: N815
    jsDivergenceMismatch: bool  # noqa: N815
    fillRateDiffMismatch: bool  # noqa: N815
    fillRatioDiffMismatch: bool  # noqa: N815
    excluded: bool


@dataclass
class RawFeatureFilterResults:
    """RawFeatureFilterResults.scala:52-58."""
    rawFeatureFilterConfig: Dict[str, object] = field(default_factory=dict)  # noqa: N815
    rawFeatureDistributions: List[FeatureDistribution] = field(default_factory=list)  # noqa: N815
    rawFeatureFilterMetrics: List[RawFeatureFilterMetrics] = field(default_factory=list)  # noqa: N815
    exclusionReasons: List[ExclusionReasons] = field(default_factory=list)  # noqa: N815


@dataclass
class FilteredRawData:
    """RawFeatureFilter.scala:603-622: the cleaned training data, what was dropped and why."""
    cleanedData: object  # noqa: N815
    featuresToDrop: List[F.Feature]  # noqa: N815
    mapKeysToDrop: Dict[str, set]  # noqa: N815
    rawFeatureFilterResults: RawFeatureFilterResults  # noqa: N815

    @property
    def trainingFeatureDistributions(self) -> List[FeatureDistribution]:  # noqa: N802
        return [d for d in self.rawFeatureFilterResults.rawFeatureDistributions if d.type == "Training"]

    @property
    def scoringFeatureDistributions(self) -> List[FeatureDistribution]:  # noqa: N802
        return [d for d in self.rawFeatureFilterResults.rawFeatureDistributions if d.type == "Scoring"]


class RawFeatureFilter:
    """RawFeatureFilter.scala:92-107.  The readers of the reference are the data themselves here: `generateFilteredRaw`
    takes the training and (optionally) the scoring data as ColumnarBatches, or as iterables of row chunks."""

    stageName = "RawFeatureFilter"  # noqa: N815
    uid = "RawFeatureFilter_100000000000"

    def __init__(self, bins: int = 100, minFill: float = 0.001, maxFillDifference: float = 0.90,  # noqa: N803
                 maxFillRatioDiff: float = 20.0, maxJSDivergence: float = 0.90, maxCorrelation: float = 0.95,  # noqa: N803
                 correlationType: str = "pearson", jsDivergenceProtectedFeatures: Iterable[str] = (),  # noqa: N803
                 protectedFeatures: Iterable[str] = (), minScoringRows: int = MIN_SCORING_ROWS_DEFAULT):  # noqa: N803
        def req(ok: bool, msg: str) -> None:
            if not ok:
                raise ValueError("requirement failed: " + msg)

        req(1 < bins <= MAX_BINS, f"Invalid bin size {bins}, bins must be between 1 and {MAX_BINS}")
        req(0.0 <= minFill <= 1.0, f"Invalid minFill size {minFill}, minFill must be between 0 and 1")
        req(0.0 <= maxFillDifference <= 1.0, f"Invalid maxFillDifference size {maxFillDifference}, "
                                             "maxFillDifference must be between 0 and 1")
        req(maxFillRatioDiff >= 0.0, f"Invalid maxFillRatioDiff size {maxFillRatioDiff}, "
                                     "maxFillRatioDiff must be greater than 0.0")
        req(0.0 <= maxJSDivergence <= 1.0, f"Invalid maxJSDivergence size {maxJSDivergence}, "
                                           "maxJSDivergence must be between 0 and 1")
        req(minScoringRows >= 0, f"minRowsForScoringSet must be >= 0, but was set to {minScoringRows}")
        if str(correlationType).lower() != "pearson":
            raise NotImplementedError("correlationType: only Pearson is built (Spearman ranks are not)")
        self.bins = int(bins)
        self.minFill = float(minFill)
        self.maxFillDifference = float(maxFillDifference)
        self.maxFillRatioDiff = float(maxFillRatioDiff)
        self.maxJSDivergence = float(maxJSDivergence)
        self.maxCorrelation = float(maxCorrelation)
        self.correlationType = "Pearson"
        self.jsDivergenceProtectedFeatures = set(jsDivergenceProtectedFeatures)
        self.protectedFeatures = set(protectedFeatures)
        self.minScoringRows = int(minScoringRows)

    # -- metrics and decisions (driver side in the reference too) ------------------------------------------------------
    def getRawFeatureFilterMetrics(self, trainingDistribs: Sequence[FeatureDistribution],  # noqa: N802,N803
                                   scoringDistribs: Sequence[FeatureDistribution],  # noqa: N803
                                   correlationInfo: Dict[FeatureKey, Dict[FeatureKey, float]]  # noqa: N803
                                   ) -> List[RawFeatureFilterMetrics]:
        """RawFeatureFilter.scala:211-296."""
        k = len(trainingDistribs)
        fills = [d.fillRate() for d in trainingDistribs]
        if not correlationInfo:
            corrs: List[Optional[float]] = [None] * k
        else:
            # one entry per (distribution, response), flattened, then zipped with the distributions (:225-231): with a
            # single response -- the only case in which the reference lines them up -- entry i belongs to feature i
            corrs = [per_response.get(d.featureKey) for d in trainingDistribs
                     for per_response in correlationInfo.values()]
        if not scoringDistribs:
            sf = js = fd = fr = [None] * k
        else:
            if len(scoringDistribs) != k:
                raise ValueError("requirement failed: scoring and training features must match")
            pairs = list(zip(trainingDistribs, scoringDistribs))
            sf = [s.fillRate() for s in scoringDistribs]
            js = [t.jsDivergence(s) for t, s in pairs]
            fd = [t.relativeFillRate(s) for t, s in pairs]
            fr = [t.relativeFillRatio(s) for t, s in pairs]
        return [RawFeatureFilterMetrics(d.name, d.key, fills[i], corrs[i], sf[i], js[i], fd[i], fr[i])
                for i, d in enumerate(trainingDistribs) if i < len(corrs)]

    def getRawFeatureFilterExclusionReasons(self, trainingDistribs, scoringDistribs,  # noqa: N802,N803
                                            rawFeatureFilterMetrics: Sequence[RawFeatureFilterMetrics]  # noqa: N803
                                            ) -> List[ExclusionReasons]:
        """RawFeatureFilter.scala:307-427."""
        out = []
        for d, m in zip(trainingDistribs, rawFeatureFilterMetrics):
            unfilled = m.trainingFillRate < self.minFill
            c = m.trainingNullLabelAbsoluteCorr
            leaker = c is not None and c > self.maxCorrelation
            s_unfilled = js_mis = fd_mis = fr_mis = False
            if scoringDistribs:
                s_unfilled = m.scoringFillRate is not None and m.scoringFillRate < self.minFill
                js_mis = (m.name not in self.jsDivergenceProtectedFeatures and m.jsDivergence is not None
                          and m.jsDivergence > self.maxJSDivergence)
                fd_mis = m.fillRateDiff is not None and m.fillRateDiff > self.maxFillDifference
                fr_mis = m.fillRatioDiff is not None and m.fillRatioDiff > self.maxFillRatioDiff
            out.append(ExclusionReasons(d.name, d.key, unfilled, leaker, s_unfilled, js_mis, fd_mis, fr_mis,
                                        any((unfilled, leaker, s_unfilled, js_mis, fd_mis, fr_mis))))
        return out

    def getFeaturesToExclude(self, trainingDistribs, scoringDistribs, correlationInfo):  # noqa: N802,N803
        """RawFeatureFilter.scala:440-470 -> (metrics, exclusion reasons, names to drop, map keys to drop)."""
        metrics = self.getRawFeatureFilterMetrics(trainingDistribs, scoringDistribs, correlationInfo)
        reasons = self.getRawFeatureFilterExclusionReasons(trainingDistribs, scoringDistribs, metrics)
        dropped: Dict[str, List[FeatureDistribution]] = {}
        kept: Dict[str, List[FeatureDistribution]] = {}
        for d, r in zip(trainingDistribs, reasons):
            (dropped if r.excluded else kept).setdefault(d.name, []).append(d)
        map_keys = set(kept) & set(dropped)
        to_drop = [k for k in dropped if k not in map_keys]
        to_drop_keys = {k: {d.key for d in v if d.key is not None} for k, v in dropped.items() if k in map_keys}
        return metrics, reasons, to_drop, to_drop_keys

    def config(self) -> Dict[str, object]:
        """RawFeatureFilterConfig (RawFeatureFilterResults.scala:84-94)."""
        return {"minFill": self.minFill, "maxFillDifference": self.maxFillDifference,
                "maxFillRatioDiff": self.maxFillRatioDiff, "maxJSDivergence": self.maxJSDivergence,
                "maxCorrelation": self.maxCorrelation, "correlationType": self.correlationType,
                "jsDivergenceProtectedFeatures": sorted(self.jsDivergenceProtectedFeatures),
                "protectedFeatures": sorted(self.protectedFeatures)}

    # -- the scans (csrc/rff.cu; every method below queues kernels on the current stream) ---------------------------
    _device = "cuda"

    @staticmethod
    def _dtype_code(t) -> int:
        import torch
        return {torch.float64: 0, torch.int64: 1, torch.uint8: 2}[t.dtype]

    def _k_summary(self, sub, n: int, out) -> None:
        """out [ncols][5] += Summary of every column of the numeric block `sub` (device)."""
        from . import runtime as R
        L.call("tmog_rff_numeric_summary_dev", sub.values.data_ptr(), self._dtype_code(sub.values),
               sub.values.stride(0),
               R.ptr(sub.valid), 0 if sub.valid is None else sub.valid.stride(0), n, len(sub.names), out.data_ptr(),
               R.stream())

    def _k_hist(self, sub, n: int, splits, mode, same, hist, vsum, nonnull) -> None:
        from . import runtime as R
        L.call("tmog_rff_numeric_hist_dev", sub.values.data_ptr(), self._dtype_code(sub.values), sub.values.stride(0),
               R.ptr(sub.valid), 0 if sub.valid is None else sub.valid.stride(0), n, len(sub.names), self.bins,
               splits.data_ptr(), mode.data_ptr(), same.data_ptr(), hist.data_ptr(), vsum.data_ptr(),
               nonnull.data_ptr(), R.stream())

    def _k_text_rows(self, sub, n: int):
        """Tokenise + hash the text columns of `sub` (device TextBlock) into separate spaces of `bins` buckets
        (TextTokenizer.Analyzer.analyze, PreparedFeatures.scala:201; HashingTF(bins), FeatureDistribution.scala:317-320)
        -> sparse rows (row_start, row_nnz, idx, val) with index = column * bins + bucket."""
        import torch
        from . import runtime as R
        from .stages.smart_text import _check_hash_flags, _col_desc
        k, bins = len(sub.names), self.bins
        params = np.array([bins, 0, 0, 1, 1, 0, k * bins, 1, 0, 0, 0, 0], dtype=np.int32)
        prefix = torch.zeros((k, 16), dtype=torch.uint8, device="cuda")
        plen = torch.zeros(k, dtype=torch.int32, device="cuda")
        newpos = torch.from_numpy(np.concatenate([np.arange(k * bins), np.full(k, -1)]).astype(np.int32)).cuda()
        desc = _col_desc(sub)
        need = C.c_size_t(0)
        if bins <= 512:
            L.call("tmog_text_hash_scratch_bytes2", n, k, params.ctypes.data, 1, C.byref(need))
        else:
            L.call("tmog_text_hash_scratch_bytes", n, C.byref(need))
        scratch = torch.empty(max(need.value, 8), dtype=torch.uint8, device="cuda")
        row_start = torch.empty(max(n, 1), dtype=torch.int64, device="cuda")
        row_nnz = torch.empty(max(n, 1), dtype=torch.int32, device="cuda")
        cursor = torch.zeros(1, dtype=torch.int64, device="cuda")
        flags = torch.zeros(4, dtype=torch.int32, device="cuda")
        args = (desc.data_ptr(), k, n, prefix.data_ptr(), plen.data_ptr(), params.ctypes.data, newpos.data_ptr(),
                row_start.data_ptr(), row_nnz.data_ptr())
        L.call("tmog_text_hash_transform_dev", *args, None, None, cursor.data_ptr(), 0, flags.data_ptr(),
               scratch.data_ptr(), scratch.numel(), R.stream())                      # sizing pass
        capacity = int(cursor.item())
        cursor.zero_()
        flags[3] = 0
        idx = torch.empty(max(capacity, 1), dtype=torch.int32, device="cuda")
        val = torch.empty(max(capacity, 1), dtype=torch.int32, device="cuda")
        L.call("tmog_text_hash_transform_dev", *args, idx.data_ptr(), val.data_ptr(), cursor.data_ptr(), capacity,
               flags.data_ptr(), scratch.data_ptr(), scratch.numel(), R.stream())
        _check_hash_flags(flags)
        return row_start, row_nnz, idx, val

    def _k_rows_hist(self, rows, n: int, ncols: int, valid_ptrs, weight, hist, tok) -> None:
        from . import runtime as R
        row_start, row_nnz, idx, val = rows
        L.call("tmog_rff_rows_hist_dev", row_start.data_ptr(), row_nnz.data_ptr(), idx.data_ptr(), val.data_ptr(), n,
               ncols, self.bins, R.ptr(valid_ptrs), R.ptr(weight), hist.data_ptr(), tok.data_ptr(), R.stream())

    def _k_code_count(self, sub, n: int, col_off, ncodes, counts, bad, y, y_valid, shift: float, null_y) -> None:
        from . import runtime as R
        L.call("tmog_rff_code_count_dev", sub.codes.data_ptr(), sub.codes.stride(0), n, len(sub.names),
               col_off.data_ptr(), ncodes.data_ptr(), counts.data_ptr(), bad.data_ptr(), R.ptr(y), R.ptr(y_valid),
               float(shift), R.ptr(null_y), R.stream())

    def _k_null_label(self, valid_ptrs, ncols: int, n: int, y, y_valid, shift: float, out) -> None:
        from . import runtime as R
        L.call("tmog_rff_null_label_sums_dev", R.ptr(valid_ptrs), ncols, n, y.data_ptr(), R.ptr(y_valid), float(shift),
               out.data_ptr(), R.stream())

    @staticmethod
    def _groups(chunk, feats: Sequence[F.Feature]):
        """The chunk's blocks that hold wanted features: [(block, columns inside the block, positions in `feats`)]."""
        by_block: Dict[int, list] = {}
        for i, f in enumerate(feats):
            b, c = chunk.find(f.name)
            by_block.setdefault(id(b), [b, [], []])
            by_block[id(b)][1].append(c)
            by_block[id(b)][2].append(i)
        return [tuple(v) for v in by_block.values()]

    def _dict_rows(self, cache: dict, block, col: int):
        """The hashed rows of a dictionary's entries (every entry tokenised once per dictionary object)."""
        import torch
        from .columnar import TextBlock
        d = block.dicts[col]
        hit = cache.get(id(d))
        if hit is None or hit[0] is not d:
            data, offsets = block.dict_arrays(col)
            buf = torch.zeros(len(data) + 16, dtype=torch.uint8, device=self._device)
            if len(data):
                buf[:len(data)] = torch.from_numpy(np.ascontiguousarray(data)).to(self._device)
            entries = TextBlock(["dictionary"], F.TEXT, [buf],
                                [torch.from_numpy(np.ascontiguousarray(offsets)).to(self._device)], [None], len(d))
            hit = (d, self._k_text_rows(entries, len(d)) if len(d) else None)
            cache[id(d)] = hit
        return hit[1]

    def _set_dict_rows(self, cache: dict, d: list):
        """One sparse row per dictionary value of a MultiPickList column: (HashingTF bucket of the whole value, 1)."""
        import torch
        from .stages.smart_text import murmur3_index
        hit = cache.get(("set", id(d)))
        if hit is None or hit[0] is not d:
            m = len(d)
            buckets = np.array([murmur3_index(v, self.bins) for v in d], dtype=np.int32)
            mk = lambda a: torch.from_numpy(a).to(self._device)   # noqa: E731
            hit = (d, (mk(np.arange(m, dtype=np.int64)), mk(np.ones(m, dtype=np.int32)), mk(buckets),
                       mk(np.ones(m, dtype=np.int32))))
            cache[("set", id(d))] = hit
        return hit[1]

    def computeFeatureStats(self, data, features: Sequence[F.Feature],  # noqa: N802
                            featureDistributionType: str = "Training",  # noqa: N803
                            allFeatureInfo: Optional[AllFeatureInformation] = None  # noqa: N803
                            ) -> AllFeatureInformation:
        """RawFeatureFilter.computeFeatureStats (:138-200).  data: a ColumnarBatch, or a list of row chunks (read twice
        for training data: summaries, then distributions); with several ranks every rank passes its own rows."""
        import torch
        from . import dist as D
        from . import runtime as R
        from .columnar import ColumnarBatch, DictBlock, NumericBlock, SetBlock, TextBlock
        R.require_cuda()
        dev = self._device
        chunks = [data] if isinstance(data, ColumnarBatch) else data
        bins = self.bins
        resp = [f for f in features if f.isResponse and f.type_name in F.NUMERIC_TYPES]      # :146-155
        pred = [f for f in features if not f.isResponse]
        feats = list(resp) + list(pred)
        for f in feats:
            if f.type_name not in F.NUMERIC_TYPES + F.TEXT_TYPES + (F.PICKLIST, F.MULTIPICKLIST):
                raise RuntimeError(f"Feature type {f.type_name} is not supported in RawFeatureFilter")
        k, nr = len(feats), len(resp)
        training = allFeatureInfo is None
        is_num = np.array([f.type_name in F.NUMERIC_TYPES for f in feats], dtype=bool)

        def device_block(block):
            return block if block.is_cuda else block.to(dev)

        # ---- pass A: the numeric Summaries (training data only: scoring data are binned over the training summaries)
        summ = np.tile(np.array([np.inf, -np.inf, 0.0, 0.0, 0.0]), (k, 1))
        if training:
            d_summ = torch.from_numpy(summ.copy()).to(dev)
            for chunk in chunks:
                for block, cols, at in self._groups(chunk, feats):
                    if not isinstance(block, NumericBlock) or chunk.n == 0:
                        continue
                    sub = device_block(block).rows(cols)
                    tmp = d_summ[at].contiguous()
                    self._k_summary(sub, chunk.n, tmp)
                    d_summ[at] = tmp
            parts = D.all_gather_object(d_summ.cpu().numpy())
            summ = parts[0].copy()
            for p in parts[1:]:
                summ[:, 0] = np.minimum(summ[:, 0], p[:, 0])
                summ[:, 1] = np.maximum(summ[:, 1], p[:, 1])
                summ[:, 2:4] += p[:, 2:4]
                summ[:, 4] = np.maximum(summ[:, 4], p[:, 4])
            nan = summ[:, 4] > 0                         # java Math.min / max: one NaN value makes both NaN
            summ[nan, 0] = summ[nan, 1] = np.nan
        else:
            known = dict(allFeatureInfo.responseSummaries)
            known.update(allFeatureInfo.predictorSummaries)
            for i, f in enumerate(feats):
                s = known.get((f.name, None))
                if s is not None:
                    summ[i, :4] = [s.min, s.max, s.sum, s.count]
        # ---- how every numeric feature is binned (FeatureDistribution.scala:328-350)
        splits = np.zeros((k, bins))
        mode = np.full(k, 2, dtype=np.int32)
        same = np.zeros(k)
        for i in np.nonzero(is_num)[0]:
            mn, mx = summ[i, 0], summ[i, 1]
            if mn < mx:
                mode[i] = 1
                L.call("tmog_rff_numeric_splits", float(mn), float(mx), bins, splits[i].ctypes.data)
            else:
                same[i] = mx
        d_splits, d_mode, d_same = (torch.from_numpy(a).to(dev) for a in (splits, mode, same))
        # ---- pass B: distributions, empty cells, and (training) the response summed over every predictor's empty rows
        hist = torch.zeros((k, bins), dtype=torch.int64, device=dev)
        vsum = torch.zeros(k, dtype=torch.float64, device=dev)
        nonnull = torch.zeros(k, dtype=torch.int64, device=dev)
        tok = torch.tensor([[np.iinfo(np.int64).max, -1, 0]] * max(k, 1), dtype=torch.int64, device=dev)[:k]
        bad = torch.zeros(1, dtype=torch.int32, device=dev)
        shifts = [summ[i, 2] / summ[i, 3] if summ[i, 3] > 0 and np.isfinite(summ[i, 2]) else 0.0 for i in range(nr)]
        nl = torch.zeros((max(nr, 1), k + 2), dtype=torch.float64, device=dev)
        dict_cache: dict = {}
        n_local = 0
        for chunk in chunks:
            n = chunk.n
            n_local += n
            if n == 0:
                continue
            groups = [(device_block(b), cols, at) for b, cols, at in self._groups(chunk, feats)]
            ys = []
            if training:
                for f in resp:                            # the response as f64, an empty one counting as 0.0 (:63-66)
                    b, c = chunk.find(f.name)
                    b = device_block(b)
                    ys.append((b.values[c].to(torch.float64), None if b.valid is None else b.valid[c].contiguous()))
            vptr = np.zeros(k, dtype=np.int64)
            for block, cols, at in groups:
                m = len(cols)
                if isinstance(block, NumericBlock):
                    sub = block.rows(cols)
                    h = torch.zeros((m, bins), dtype=torch.int64, device=dev)
                    vs, nn = vsum[at].contiguous(), nonnull[at].contiguous()
                    self._k_hist(sub, n, d_splits[at].contiguous(), d_mode[at].contiguous(), d_same[at].contiguous(), h,
                                 vs, nn)
                    hist[at] += h
                    vsum[at], nonnull[at] = vs, nn
                    if block.valid is not None:
                        for c, i in zip(cols, at):
                            vptr[i] = block.valid[c].data_ptr()
                elif isinstance(block, TextBlock):
                    sub = block.rows(cols)
                    rows = self._k_text_rows(sub, n)
                    ptrs = np.array([0 if v is None else v.data_ptr() for v in sub.valid], dtype=np.int64)
                    d_ptrs = torch.from_numpy(ptrs).to(dev)
                    h = torch.zeros(m * bins, dtype=torch.int64, device=dev)
                    t = tok[at].contiguous()
                    self._k_rows_hist(rows, n, m, d_ptrs, None, h, t)
                    hist[at] += h.view(m, bins)
                    tok[at] = t
                    for p, i in zip(ptrs, at):
                        vptr[i] = p
                elif isinstance(block, DictBlock):
                    sub = block.rows(cols)
                    sizes = np.array([len(d) for d in sub.dicts], dtype=np.int64)
                    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
                    d_off, d_sizes = torch.from_numpy(offs[:-1].copy()).to(dev), torch.from_numpy(sizes).to(dev)
                    counts = torch.zeros(max(int(offs[-1]), 1), dtype=torch.int64, device=dev)
                    for r in range(max(len(ys), 1)):      # the codes are counted once; every response has its own sums
                        y, yv = ys[r] if ys else (None, None)
                        ny = torch.zeros(m, dtype=torch.float64, device=dev) if ys else None
                        self._k_code_count(sub, n, d_off, d_sizes, counts if r == 0 else torch.zeros_like(counts),
                                           bad, y, yv, shifts[r] if ys else 0.0, ny)
                        if ys:
                            nl[r, at] += ny
                    for j, i in enumerate(at):
                        if sizes[j] == 0:
                            continue
                        rows = self._dict_rows(dict_cache, sub, j)
                        h = torch.zeros(bins, dtype=torch.int64, device=dev)
                        t = tok[i:i + 1].contiguous()
                        self._k_rows_hist(rows, int(sizes[j]), 1, None, counts[offs[j]:offs[j + 1]].contiguous(), h, t)
                        hist[i] += h
                        tok[i:i + 1] = t
                elif isinstance(block, SetBlock):
                    # MultiPickList: Left(values), hashed whole (PreparedFeatures.scala:179); an empty set is an empty
                    # feature.  Rows per dictionary code by the count kernel over the flat code list, then every
                    # dictionary entry's bucket weighted by its count; set sizes are the "token" counts
                    sub = block.rows(cols)
                    totals = torch.stack([o[n] for o in sub.offsets]).cpu().numpy()
                    big = int(np.iinfo(np.int64).max)
                    for j, i in enumerate(at):
                        d = sub.dicts[j]
                        sizes = sub.offsets[j][1:n + 1] - sub.offsets[j][:n]
                        filled = sizes > 0
                        tok[i, 0] = torch.minimum(tok[i, 0], torch.where(filled, sizes, big).min())
                        tok[i, 1] = torch.maximum(tok[i, 1], torch.where(filled, sizes, -1).max())
                        tok[i, 2] += filled.sum()
                        for r, (y, yv) in enumerate(ys):
                            yy = y[:n]
                            if yv is not None:
                                rows_i = torch.arange(n, device=dev)
                                held = ((yv[rows_i >> 6] >> (rows_i & 63)) & 1) != 0
                                yy = torch.where(held, yy, torch.zeros_like(yy))
                            nl[r, i] += ((yy - shifts[r]) * (~filled)).sum()
                        total = int(totals[j])
                        if total == 0 or len(d) == 0:
                            continue
                        flat = DictBlock([sub.names[j]], sub.type_name, sub.codes[j][:total].view(1, total), [d], total)
                        counts = torch.zeros(len(d), dtype=torch.int64, device=dev)
                        self._k_code_count(flat, total, torch.zeros(1, dtype=torch.int64, device=dev),
                                           torch.tensor([len(d)], dtype=torch.int64, device=dev), counts, bad, None,
                                           None, 0.0, None)
                        h = torch.zeros(bins, dtype=torch.int64, device=dev)
                        unused = torch.tensor([[big, -1, 0]], dtype=torch.int64, device=dev)
                        self._k_rows_hist(self._set_dict_rows(dict_cache, d), len(d), 1, None, counts, h, unused)
                        hist[i] += h
                else:
                    raise RuntimeError(f"{type(block).__name__} columns are not supported in RawFeatureFilter")
            if ys:
                d_vptr = torch.from_numpy(vptr).to(dev)
                for r, (y, yv) in enumerate(ys):
                    out = torch.zeros(k + 2, dtype=torch.float64, device=dev)
                    self._k_null_label(d_vptr, k, n, y, yv, shifts[r], out)
                    nl[r] += out
            del groups, ys
        if int(bad.item()):
            raise ValueError("a dictionary code beyond its dictionary")
        # the text / PickList cells that hold a value: counted by the row kernel
        is_tok = ~is_num
        if is_tok.any():
            sel = torch.from_numpy(np.nonzero(is_tok)[0]).to(dev)
            nonnull[sel] = tok[sel, 2]
        # ---- the ranks' sums add
        n_all = int(D.allreduce_sum(np.array([n_local], dtype=np.int64))[0])
        h_hist = D.allreduce_sum(hist.cpu().numpy())
        h_nonnull = D.allreduce_sum(nonnull.cpu().numpy())
        h_nl = D.allreduce_sum(nl.cpu().numpy())
        h_tok = tok.cpu().numpy()
        if D.world()[1] > 1:
            parts = D.all_gather_object(h_tok)
            h_tok = parts[0].copy()
            for p in parts[1:]:
                h_tok[:, 0] = np.minimum(h_tok[:, 0], p[:, 0])
                h_tok[:, 1] = np.maximum(h_tok[:, 1], p[:, 1])
        return self._assemble(feats, nr, is_num, summ, splits, mode, h_hist, h_nonnull, h_tok, h_nl, n_all,
                              featureDistributionType, allFeatureInfo)

    def _assemble(self, feats, nr: int, is_num, summ, splits, mode, hist, nonnull, tok, nl, n_all: int, type_: str,
                  info: Optional[AllFeatureInformation]) -> AllFeatureInformation:
        """What the reduce of the per-row FeatureDistributions leaves (FeatureDistribution.scala:107-122, :217-248):
        a feature that is empty in every training row has no Summary and therefore no distribution at all."""
        bins = self.bins
        training = info is None
        if training:
            summaries: Dict[FeatureKey, Summary] = {}
            for i, f in enumerate(feats):
                if is_num[i]:
                    if summ[i, 3] > 0:
                        summaries[(f.name, None)] = Summary(*[float(v) for v in summ[i, :4]])
                elif nonnull[i] > 0:
                    summaries[(f.name, None)] = Summary(float(tok[i, 0]), float(tok[i, 1]), float(hist[i].sum()),
                                                        float(nonnull[i]))
        else:
            summaries = dict(info.responseSummaries)
            summaries.update(info.predictorSummaries)
        dists: Dict[FeatureKey, FeatureDistribution] = {}
        for i, f in enumerate(feats):
            s = summaries.get((f.name, None))
            if s is None:
                continue
            nulls = int(n_all - nonnull[i])
            info4 = s.as_array()
            dist = hist[i].astype(np.float64)
            summary_info = info4
            if is_num[i]:
                if mode[i] == 1:
                    # the rows' summaryInfo: the splits for a row with a value, [min, max, sum, count] for an empty
                    # one; the reduce keeps the longer array (on a tie -- bins = 4 -- the later row's: the splits here)
                    if nonnull[i] > 0 and (bins >= 4 or nulls == 0):
                        summary_info = splits[i].copy()
                else:
                    # [same, other]: two entries, padded to `bins` by the array monoid once an empty row joins in
                    dist = dist[:2] if nulls == 0 else dist
            dists[(f.name, None)] = FeatureDistribution(f.name, None, n_all, nulls, dist, summary_info, type_)
        rkeys = [(f.name, None) for f in feats[:nr] if (f.name, None) in dists]
        pkeys = [(f.name, None) for f in feats[nr:] if (f.name, None) in dists]
        if training:
            at = {(f.name, None): i for i, f in enumerate(feats)}
            corr_info: Dict[FeatureKey, Dict[FeatureKey, float]] = {}
            for rk in rkeys:
                r = at[rk]
                cols = [at[pk] for pk in pkeys]
                null_count = np.ascontiguousarray((n_all - nonnull[cols]).astype(np.float64))
                null_y = np.ascontiguousarray(nl[r, cols].astype(np.float64))
                out = np.zeros(len(cols))
                L.call("tmog_rff_null_label_corr", float(n_all), float(nl[r, len(feats)]), float(nl[r, len(feats) + 1]),
                       len(cols), null_count.ctypes.data, null_y.ctypes.data, out.ctypes.data)
                corr_info[rk] = {pk: float(out[j]) for j, pk in enumerate(pkeys)}
        else:
            corr_info = info.correlationInfo
        return AllFeatureInformation({k_: summaries[k_] for k_ in rkeys}, [dists[k_] for k_ in rkeys],
                                     {k_: summaries[k_] for k_ in pkeys}, [dists[k_] for k_ in pkeys], corr_info)

    def generateFilteredRaw(self, trainData, rawFeatures: Sequence[F.Feature],  # noqa: N802,N803
                            scoreData=None) -> FilteredRawData:  # noqa: N803
        """RawFeatureFilter.generateFilteredRaw (:482-566): the readers' data frames are the arguments here.  The
        cleaned data are the training batch (or chunks) without the dropped features' columns."""
        from . import dist as D
        from .columnar import ColumnarBatch

        def rows_of(data) -> int:
            local = data.n if isinstance(data, ColumnarBatch) else sum(c.n for c in data)
            return int(D.allreduce_sum(np.array([local], dtype=np.int64))[0])

        if rows_of(trainData) <= 0:
            raise ValueError("requirement failed: RawFeatureFilter cannot work with empty training data")
        training = self.computeFeatureStats(trainData, rawFeatures, "Training")
        scoring = None
        if scoreData is not None and rows_of(scoreData) >= self.minScoringRows:
            scoring = self.computeFeatureStats(scoreData, rawFeatures, "Scoring", training)
        keep = lambda ds: [d for d in ds if d.name not in self.protectedFeatures]   # noqa: E731
        metrics, reasons, to_drop, to_drop_keys = self.getFeaturesToExclude(
            keep(training.predictorDistributions), keep(scoring.predictorDistributions) if scoring is not None else [],
            training.correlationInfo)
        features_to_drop = [f for f in rawFeatures if f.name in to_drop]
        features_to_keep = [f for f in rawFeatures if f.name not in to_drop]
        if not any(not f.isResponse for f in features_to_keep):
            raise ValueError("requirement failed: The raw feature filter has dropped all of your features, check your "
                             "input data quality")
        dropped = set(to_drop)

        def clean(batch: ColumnarBatch) -> ColumnarBatch:
            blocks = []
            for b in batch.blocks:
                left = [i for i, nm in enumerate(b.names) if nm not in dropped]
                if len(left) == len(b.names):
                    blocks.append(b)
                elif left:
                    blocks.append(b.rows(left))
            return ColumnarBatch(batch.n, blocks, {}, batch.response)

        cleaned = clean(trainData) if isinstance(trainData, ColumnarBatch) else [clean(c) for c in trainData]
        distributions = training.responseDistributions + training.predictorDistributions
        if scoring is not None:
            distributions = distributions + scoring.responseDistributions + scoring.predictorDistributions
        results = RawFeatureFilterResults(self.config(), distributions, metrics, reasons)
        return FilteredRawData(cleaned, features_to_drop, to_drop_keys, results)
