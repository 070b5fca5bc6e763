"""SanityChecker on the GPU path.

Reference: core/src/main/scala/com/salesforce/op/stages/impl/preparators/SanityChecker.scala:58-222 (params),
:252-348 (categoricalTests), :356-361 (fraction), :367-541 (fitFn), :544-559 (model), :561-581 (defaults);
DerivedFeatureFilterUtils.scala:95-305; SanityCheckerMetadata.scala (summary).

The statistics come from the sidecar the fused vectorizer scans attach to the input vector (stats.py); the
contingency statistics and the drop decision run in libtmog_b200's host functions.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np
import torch

from .. import _lib as L
from .. import dist as D
from .. import features as F
from ..columnar import ColumnarBatch, CsrPart, DensePart, VectorColumn
from ..metadata import OpVectorMetadata
from ..stats import VectorStats
from .base import OpEstimator, OpModel

REASONS = {1: "variance lower than min variance", 2: "correlation lower than min correlation",
           4: "correlation higher than max correlation", 8: "feature-feature correlation higher than max",
           16: "Cramer's V higher than max Cramer's V", 32: "max association rule confidence above threshold",
           64: "other feature in indicator group flagged for removal via rule confidence checks",
           128: "Cramer's V for something in parent feature set higher than max Cramer's V",
           256: "correlation for something in parent feature set higher than max correlation"}


@dataclass
class CategoricalGroupStats:
    group: str
    categoricalFeatures: List[str]
    contingencyMatrix: np.ndarray        # rows = choices, cols = label classes ascending
    cramersV: float
    mutualInfo: float
    pointwiseMutualInfo: np.ndarray
    maxRuleConfidences: np.ndarray
    supports: np.ndarray


@dataclass
class SanityCheckerSummary:
    names: List[str]
    count: int
    mean: np.ndarray
    variance: np.ndarray
    min: np.ndarray
    max: np.ndarray
    correlationsWithLabel: np.ndarray
    corrIndices: List[int]
    categoricalStats: List[CategoricalGroupStats]
    dropped: List[str]
    reasons: Dict[str, List[str]]
    sampleFraction: float
    featureFeatureCorr: Optional[np.ndarray] = None


class SanityChecker(OpEstimator):
    """BinaryEstimator[RealNN, OPVector, OPVector]: setInput(label, featureVector)."""

    def __init__(self, uid: Optional[str] = None):
        super().__init__("SanityChecker", uid)
        # defaults: SanityChecker.scala:561-581
        self.checkSample = 1.0
        self.sampleLowerLimit = 1000
        self.sampleUpperLimit = 1000000
        self.maxCorrelation = 0.95
        self.maxFeatureCorrelation = 0.99
        self.minCorrelation = 0.0
        self.minVariance = 1e-5
        self.maxCramersV = 0.95
        self.removeBadFeatures = False
        self.removeFeatureGroup = True
        self.protectTextSharedHash = False
        self.maxRuleConfidence = 1.0
        self.minRequiredRuleSupport = 1.0
        self.featureFeatureCorrLevel = "Computed"
        self.correlationExclusion = "NoExclusion"
        self.categoricalLabel: Optional[bool] = None
        self.sampleSeed = 0

    # -- params (same names and validation as SanityCheckerParams) ---------------------------------------------------
    def _set(self, name, value, ok=True, what=""):
        if not ok:
            raise ValueError(f"requirement failed: {name} given invalid value {value}{what}")
        setattr(self, name, value)
        return self

    def setCheckSample(self, v):  # noqa: N802
        return self._set("checkSample", float(v), 0.0 < v <= 1.0)

    def setSampleLowerLimit(self, v):  # noqa: N802
        return self._set("sampleLowerLimit", int(v), v >= 0)

    def setSampleUpperLimit(self, v):  # noqa: N802
        return self._set("sampleUpperLimit", int(v), v >= 0)

    def setMaxCorrelation(self, v):  # noqa: N802
        return self._set("maxCorrelation", float(v), -0.1 <= v <= 1.1)

    def setMinCorrelation(self, v):  # noqa: N802
        return self._set("minCorrelation", float(v), -0.1 <= v <= 1.1)

    def setMaxFeatureCorrelation(self, v):  # noqa: N802
        return self._set("maxFeatureCorrelation", float(v), -0.1 <= v <= 1.1)

    def setMinVariance(self, v):  # noqa: N802
        return self._set("minVariance", float(v))

    def setMaxCramersV(self, v):  # noqa: N802
        return self._set("maxCramersV", float(v), 0.0 <= v <= 1.0)

    def setRemoveBadFeatures(self, v):  # noqa: N802
        return self._set("removeBadFeatures", bool(v))

    def setRemoveFeatureGroup(self, v):  # noqa: N802
        return self._set("removeFeatureGroup", bool(v))

    def setProtectTextSharedHash(self, v):  # noqa: N802
        return self._set("protectTextSharedHash", bool(v))

    def setMaxRuleConfidence(self, v):  # noqa: N802
        return self._set("maxRuleConfidence", float(v), 0.0 <= v <= 1.0)

    def setMinRequiredRuleSupport(self, v):  # noqa: N802
        return self._set("minRequiredRuleSupport", float(v), 0.0 <= v <= 1.0)

    def setFeatureFeatureCorrLevel(self, v: str):  # noqa: N802
        return self._set("featureFeatureCorrLevel", v, v in ("Computed", "Off"))

    def setCorrelationExclusion(self, v: str):  # noqa: N802
        return self._set("correlationExclusion", v, v in ("NoExclusion", "HashedText"))

    def setCategoricalLabel(self, v: bool):  # noqa: N802
        return self._set("categoricalLabel", bool(v))

    def setSampleSeed(self, v: int):  # noqa: N802
        return self._set("sampleSeed", int(v))

    def setInput(self, label: F.Feature, features: F.Feature):  # noqa: N802
        if not label.is_response:
            raise ValueError("requirement failed: The first input to SanityChecker must be a response feature")
        if label.type_name != F.REALNN or features.type_name != F.OPVECTOR:
            raise ValueError("requirement failed: SanityChecker inputs must be (RealNN label, OPVector features)")
        self._inputs = [label, features]
        self._output = None
        return self

    def outputIsResponse(self) -> bool:  # noqa: N802
        return False

    def getOutputFeatureName(self) -> str:  # noqa: N802
        names = "-".join(f.name for f in self.inN)
        return f"{names}_{len(self.inN)}-stagesApplied_OPVector_{self.uid[-12:]}"

    def fraction(self, total: int) -> float:
        min_f = min(1.0, self.sampleLowerLimit / total)
        max_f = max(0.0, self.sampleUpperLimit / total)
        return max(min(self.checkSample, max_f), min_f)

    # -- fit ---------------------------------------------------------------------------------------------------------
    def fit(self, batch: ColumnarBatch) -> "SanityCheckerModel":
        label_f, vec_f = self.inN
        vec = batch.vectors[vec_f.name]
        n = int(D.allreduce_sum(np.array([batch.n], dtype=np.int64))[0])    # rows over all ranks
        if n <= 0:
            raise ValueError("requirement failed: Sample size cannot be zero")
        frac = self.fraction(n)
        if frac < 1.0:
            raise NotImplementedError(
                f"SanityChecker sampling (fraction {frac:.4g} < 1) is not on the GPU path: set checkSample = 1.0 and "
                "sampleUpperLimit >= the row count (the reference default down-samples to 1e6 rows at random)")
        meta = vec.metadata
        d = vec.width
        if d <= 0:
            raise ValueError("requirement failed: Feature vector passed in is empty, check your vectorizers")
        if d != meta.size:
            raise ValueError(f"requirement failed: Number of columns in vector metadata ({meta.size}) did not match "
                             f"number of columns in data({d}), check your vectorizers")
        st: Optional[VectorStats] = vec.stats
        if st is None or st.label.name != label_f.name or st.label.n != n:
            from .generic_stats import vector_stats
            st = vector_stats(batch, vec, label_f.name)
        lab = st.label
        cols = meta.columns
        # correlations with the label ---------------------------------------------------------------------------
        if self.correlationExclusion == "HashedText":
            hashed = {c.index for c in cols if c.grouping is None and c.indicatorValue is None
                      and c.hasParentOfType(F.TEXT, F.TEXTAREA, "TextMap", "TextAreaMap")}
            corr_indices = [i for i in range(d + 1) if i not in hashed]
        else:
            corr_indices = list(range(d + 1))
        feat_idx = [i for i in corr_indices if i < d]
        corr = np.zeros(len(corr_indices))
        if n <= 1:
            raise ValueError(f"requirement failed: computeCorrelationsWithLabel called on matrix with only {n} rows.")
        cov = np.ascontiguousarray(st.cov_label[feat_idx])
        var = np.ascontiguousarray(st.variance[feat_idx])
        sub = np.zeros(len(feat_idx))
        L.call("tmog_pearson_from_cov", len(feat_idx), float(n), cov.ctypes.data, var.ctypes.data,
               float(lab.stats[2]), sub.ctypes.data)
        corr[:len(feat_idx)] = sub
        corr[-1] = 1.0 if lab.stats[2] > 0 else np.nan
        corr_matrix = None
        if self.featureFeatureCorrLevel == "Computed":
            from .generic_stats import pearson_matrix
            corr_matrix = pearson_matrix(batch, vec, label_f.name, corr_indices, st)
            corr = corr_matrix[:, -1].copy()
        # categorical tests -------------------------------------------------------------------------------------
        cat_stats: List[CategoricalGroupStats] = []
        is_cat = self.categoricalLabel is True or (lab.classes is not None and len(lab.classes) < min(100.0, n * 0.1))
        if self.categoricalLabel is not False and is_cat:
            if lab.classes is None or st.class_sums is None:
                raise NotImplementedError("categoricalLabel = true with more than 255 distinct label values")
            cat_stats = self._categorical_tests(meta, st)
        # decision ----------------------------------------------------------------------------------------------
        drop, reasons, pc, pv = self._decide(meta, st, corr, corr_indices, corr_matrix, cat_stats)
        names = meta.columnNames()
        to_drop = [names[i] for i in range(d) if drop[i]] if self.removeBadFeatures else []
        keep = [i for i in range(d) if not (drop[i] and self.removeBadFeatures)]
        out_meta = OpVectorMetadata(self.getOutputFeatureName(), [cols[i] for i in keep], meta.history)
        self._metadata = out_meta
        self.summary = SanityCheckerSummary(
            names=names + [label_f.name], count=n, mean=np.append(st.mean, lab.stats[1]),
            variance=np.append(st.variance, lab.stats[2]), min=np.append(st.min, lab.stats[3]),
            max=np.append(st.max, lab.stats[4]), correlationsWithLabel=corr, corrIndices=corr_indices,
            categoricalStats=cat_stats, dropped=to_drop,
            reasons={names[i]: [t for b, t in REASONS.items() if reasons[i] & b] for i in range(d) if reasons[i]},
            sampleFraction=frac, featureFeatureCorr=corr_matrix)
        self.parentCorr, self.parentCramersV = pc, pv
        if not keep:
            raise ValueError("requirement failed: The sanity checker has dropped all of your features, check your "
                             "input data quality")
        return SanityCheckerModel(self, keep, self.removeBadFeatures)

    def _categorical_tests(self, meta: OpVectorMetadata, st: VectorStats) -> List[CategoricalGroupStats]:
        counts = st.label.class_counts.astype(np.float64)
        groups: Dict[str, list] = {}
        for c in meta.columns:
            if c.grouping is not None and c.indicatorValue is not None:
                groups.setdefault(c.featureGroup(), []).append(c)
        out = []
        for g, cols in groups.items():
            seen, cleaned = set(), []
            for c in cols:                       # duplicate indicator values: only the first is used (:295-298)
                if c.indicatorValue not in seen:
                    seen.add(c.indicatorValue)
                    cleaned.append(c)
            idx = [c.index for c in cleaned]
            sums = st.class_sums[:, idx]         # [L, k]
            if np.isnan(sums).any():
                raise ValueError(f"no per-label sums were tracked for categorical group {g}")
            if any(c.isMultiPickList for c in cleaned):
                raise NotImplementedError("MultiPickList contingency statistics are not on the GPU path yet")
            if len(idx) == 1:
                mat = np.stack([sums[:, 0], counts - sums[:, 0]])       # 2 x L
            else:
                mat = sums.T.copy()                                     # k x L
            mat = np.ascontiguousarray(mat, dtype=np.float64)
            r, c_ = mat.shape
            v, chi, mi = (np.zeros(1) for _ in range(3))
            pmi = np.zeros(r * c_)
            conf, sup = np.zeros(r), np.zeros(r)
            rc = L.lib().tmog_contingency_stats(mat.ctypes.data, r, c_, v.ctypes.data, chi.ctypes.data,
                                                mi.ctypes.data, pmi.ctypes.data, conf.ctypes.data, sup.ctypes.data)
            if rc < 0:
                L.check(rc)
            if rc == 1:
                conf, sup = np.zeros(0), np.zeros(0)
            out.append(CategoricalGroupStats(g, [c.makeColName() for c in cleaned], mat, float(v[0]), float(mi[0]),
                                             pmi.reshape(r, c_), conf, sup))
        return out

    def _decide(self, meta, st, corr, corr_indices, corr_matrix, cat_stats):
        d = meta.size
        pos = {ci: k for k, ci in enumerate(corr_indices)}
        corr_pos = np.array([pos.get(i, -1) for i in range(d)], dtype=np.int32)
        parent_ids: Dict[str, int] = {}
        group_ids: Dict[str, int] = {}
        offs, pids = [0], []
        gid = np.full(d, -1, dtype=np.int32)
        tsh = np.zeros(d, dtype=np.uint8)
        for c in meta.columns:
            for pn in c.parentNamesWithMapKeys():
                pids.append(parent_ids.setdefault(pn, len(parent_ids)))
            offs.append(len(pids))
            g = c.featureGroup()
            if g is not None:
                gid[c.index] = group_ids.setdefault(g, len(group_ids))
            tsh[c.index] = c.isTextSharedHash
        name_to_idx = {c.makeColName(): c.index for c in meta.columns}
        cram = np.full(d, np.nan)
        rule_n = np.full(d, -1, dtype=np.int32)
        rconf, rsup = np.zeros(2 * d), np.zeros(2 * d)
        for g in cat_stats:
            lone = len(g.categoricalFeatures) == 1
            for k, nm in enumerate(g.categoricalFeatures):
                i = name_to_idx[nm]
                cram[i] = g.cramersV
                if lone:
                    m = min(2, len(g.maxRuleConfidences))
                    rule_n[i] = m
                    rconf[2 * i:2 * i + m] = g.maxRuleConfidences[:m]
                    rsup[2 * i:2 * i + m] = g.supports[:m]
                else:
                    has = k < len(g.maxRuleConfidences)
                    rule_n[i] = 1 if has else 0
                    if has:
                        rconf[2 * i], rsup[2 * i] = g.maxRuleConfidences[k], g.supports[k]
        params = np.array([self.minVariance, self.minCorrelation, self.maxCorrelation, self.maxFeatureCorrelation,
                           self.maxCramersV, self.maxRuleConfidence, self.minRequiredRuleSupport,
                           float(self.removeFeatureGroup), float(self.protectTextSharedHash)])
        offs = np.asarray(offs, dtype=np.int32)
        pids = np.asarray(pids, dtype=np.int32)
        drop = np.zeros(d, dtype=np.uint8)
        reasons = np.zeros(d, dtype=np.uint32)
        pc, pv = np.zeros(d), np.zeros(d)
        var = np.ascontiguousarray(st.variance, dtype=np.float64)
        corr = np.ascontiguousarray(corr, dtype=np.float64)
        cm = None if corr_matrix is None else np.ascontiguousarray(corr_matrix, dtype=np.float64)
        L.call("tmog_sanity_decide", d, var.ctypes.data, corr.ctypes.data, None if cm is None else cm.ctypes.data,
               0 if cm is None else cm.shape[0], corr_pos.ctypes.data, offs.ctypes.data, pids.ctypes.data,
               gid.ctypes.data, tsh.ctypes.data, cram.ctypes.data, rule_n.ctypes.data, rconf.ctypes.data,
               rsup.ctypes.data, params.ctypes.data, drop.ctypes.data, reasons.ctypes.data, pc.ctypes.data,
               pv.ctypes.data)
        return drop, reasons, pc, pv


class SanityCheckerModel(OpModel):
    def __init__(self, parent: SanityChecker, indicesToKeep: List[int], removeBadFeatures: bool):  # noqa: N803
        super().__init__(parent)
        self.indicesToKeep = list(indicesToKeep)
        self.removeBadFeatures = removeBadFeatures

    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        vec = batch.vectors[self.inN[1].name]
        name = self.getOutput().name
        if not self.removeBadFeatures:
            return VectorColumn(name, batch.n, self.getMetadata(), vec.parts, vec.stats)
        from .generic_stats import select_columns
        parts = select_columns(vec, self.indicesToKeep)
        stats = None if vec.stats is None else vec.stats.select(self.indicesToKeep)
        return VectorColumn(name, batch.n, self.getMetadata(), parts, stats)
