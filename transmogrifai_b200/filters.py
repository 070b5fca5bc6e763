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
TextArea, PickList.  Not covered (raise): maps, MultiPickList, lists, geolocations, vectors, a custom textBinsFormula,
timePeriod.  Not produced: FeatureDistribution.moments / cardEstimate (reported by the reference, read by no rule).
Order of the distributions: responses, then predictors, in the order of `features` (the reference's is the iteration
order of a Scala Map: unspecified beyond four features).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Sequence, Tuple, Union

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
            corrs = [per_response.get(d.featureKey) for d in trainingDistribs for per_response in correlationInfo.values()]
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
