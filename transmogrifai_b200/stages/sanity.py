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
from ..metadata import OpVectorColumnMetadata, OpVectorMetadata
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


class _GroupStatsList:
    """categoricalStats as a read-only sequence over the packed arrays of tmog_contingency_stats_batch; the per-group
    objects are only built when somebody looks at them (a wide vector has thousands of lone null-indicator groups)."""

    def __init__(self, plan, mats, v, mi, pm, conf, sup, status):
        self.plan, self.mats, self.v, self.mi, self.pm, self.conf, self.sup, self.status = \
            plan, mats, v, mi, pm, conf, sup, status
        self._built: Dict[int, CategoricalGroupStats] = {}

    def __len__(self) -> int:
        return len(self.plan.groups)

    def __getitem__(self, g):
        if isinstance(g, slice):
            return [self[i] for i in range(*g.indices(len(self)))]
        if g < 0:
            g += len(self)
        got = self._built.get(g)
        if got is None:
            name, _names, idx = self.plan.groups[g]
            col_names = self.plan.meta.names_of(idx)
            a, b = self.plan.row_offsets[g], self.plan.row_offsets[g] + self.plan.rows[g]
            empty = self.status[g] == 1
            got = CategoricalGroupStats(name, col_names, self.mats[a:b], float(self.v[g]), float(self.mi[g]),
                                        self.pm[a:b], np.zeros(0) if empty else self.conf[a:b],
                                        np.zeros(0) if empty else self.sup[a:b])
            self._built[g] = got
        return got

    def __iter__(self):
        return (self[g] for g in range(len(self)))

    def __eq__(self, other):
        return list(self) == list(other)


@dataclass
class SanityCheckerSummary:
    meta: OpVectorMetadata               # metadata of the checked vector (column names are expanded on demand)
    labelName: str
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

    @property
    def names(self) -> List[str]:
        """featureNames :+ label name (SanityChecker.scala:519)"""
        return self.meta.columnNames() + [self.labelName]


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
        sidecar = vec.stats
        if sidecar is not None and sidecar.label.name == label_f.name:
            n = int(sidecar.label.n)                                          # rows over all ranks, counted by the fit scans
        else:
            n = int(D.allreduce_sum(np.array([batch.n], dtype=np.int64))[0])
        if n <= 0:
            raise ValueError("requirement failed: Sample size cannot be zero")
        frac = self.fraction(n)
        pre = batch.__dict__.get("_presampled")
        if pre is not None:
            # a chunk-streamed dataset: the sample has been drawn chunk by chunk and the sidecar is the sample's
            # (streaming.fit_stream); batch.n is the row count of the whole dataset
            frac = float(pre)
        elif frac < 1.0:
            batch, vec, n = self._sample(batch, vec, frac)
            if n <= 0:
                raise ValueError("requirement failed: Sample size cannot be zero")
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
        plan = self._plan(meta)
        # correlations with the label ---------------------------------------------------------------------------
        if self.correlationExclusion == "HashedText":
            corr_indices = np.concatenate([np.nonzero(~plan.hashed_text)[0], [d]]).astype(np.int64)
        else:
            corr_indices = np.arange(d + 1, dtype=np.int64)
        feat_idx = corr_indices[:-1]
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
            stream_corr = batch.__dict__.get("_stream_corr")
            if stream_corr is not None:
                # a chunk-streamed dataset: the rows are not in `batch`; one more pass over the chunks accumulates the
                # Gramian (streaming.fit_stream)
                corr_matrix = stream_corr(vec_f.name, label_f.name, [int(i) for i in corr_indices], st)
            else:
                from .generic_stats import pearson_matrix
                corr_matrix = pearson_matrix(batch, vec, label_f.name, [int(i) for i in corr_indices], st)
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
        drop_idx = np.nonzero(drop)[0] if self.removeBadFeatures else np.zeros(0, dtype=np.int64)
        to_drop = meta.names_of(drop_idx)
        keep = np.nonzero(drop == 0)[0] if self.removeBadFeatures else np.arange(d)
        key = (id(meta), keep.tobytes())
        if getattr(self, "_out_meta_key", None) != key:     # same input metadata and keep set: same output metadata
            self._out_meta = meta.select(self.getOutputFeatureName(), keep)
            self._out_meta_key = key
        self._metadata = self._out_meta
        flagged = np.nonzero(reasons)[0]
        self.summary = SanityCheckerSummary(
            meta=meta, labelName=label_f.name, count=n, mean=np.append(st.mean, lab.stats[1]),
            variance=np.append(st.variance, lab.stats[2]), min=np.append(st.min, lab.stats[3]),
            max=np.append(st.max, lab.stats[4]), correlationsWithLabel=corr,
            corrIndices=corr_indices, categoricalStats=cat_stats, dropped=to_drop,
            reasons={nm: [t for b, t in REASONS.items() if reasons[i] & b]
                     for nm, i in zip(meta.names_of(flagged), flagged)},
            sampleFraction=frac, featureFeatureCorr=corr_matrix)
        self.parentCorr, self.parentCramersV = pc, pv
        if len(keep) == 0:
            raise ValueError("requirement failed: The sanity checker has dropped all of your features, check your "
                             "input data quality")
        return SanityCheckerModel(self, keep, self.removeBadFeatures)

    def _sample(self, batch: ColumnarBatch, vec, frac: float):
        """data.sample(withReplacement = false, fraction, seed) (SanityChecker.scala:381-387): every row is kept with
        probability `frac` (a counter-based generator on the device, seeded with sampleSeed -- Spark's per-partition
        XORShift stream is not restated, so the sample is a different one with the same law).  The sampled raw rows
        form a batch of their own, the upstream transformers run again on it, and the statistics are those of the
        resulting vector (no statistics sidecar applies to a sample)."""
        if not batch.blocks:
            raise NotImplementedError("SanityChecker sampling (fraction < 1) over a dataset streamed in row chunks: set "
                                      "sampleUpperLimit >= the row count (statistics then come from the fused scans)")
        dev = "cuda" if any(getattr(b, "is_cuda", False) for b in batch.blocks) else "cpu"
        g = torch.Generator(device=dev).manual_seed(int(self.sampleSeed) + D.world()[0])   # a stream per row shard
        idx = (torch.rand(batch.n, device=dev, generator=g) < frac).nonzero().squeeze(1)
        on_dev = lambda b: b if getattr(b, "is_cuda", dev == "cuda") == (dev == "cuda") else b.to(dev)  # noqa: E731
        full = ColumnarBatch(batch.n, [on_dev(b) for b in batch.blocks], batch.vectors, batch.response)
        sub = full.take_rows(idx)
        svec = vec.recompute(batch, sub)          # with its statistics sidecar where the upstream stages provide one
        m = int(D.allreduce_sum(np.array([idx.numel()], dtype=np.int64))[0])      # sampled rows of all ranks
        return sub.with_vector(svec), svec, m

    @staticmethod
    def _plan(meta: OpVectorMetadata) -> "_MetaPlan":
        plan = meta.__dict__.get("_sanity_plan")
        if plan is None:
            # the plan depends on the column descriptions only: vectors with equal metadata (a refit on the next
            # batch, the folds of a cross-validation) share it
            key = tuple(meta.runs)
            plan = _PLAN_CACHE.get(key)
            if plan is None:
                if len(_PLAN_CACHE) >= 64:
                    _PLAN_CACHE.clear()
                plan = _PLAN_CACHE[key] = _MetaPlan(meta)
            meta.__dict__["_sanity_plan"] = plan
        return plan

    def _categorical_tests(self, meta: OpVectorMetadata, st: VectorStats) -> List[CategoricalGroupStats]:
        """SanityChecker.categoricalTests :252-348 -- contingency matrices from the per-class column sums, then one
        batched tmog_contingency_stats call."""
        plan = self._plan(meta)
        if not plan.groups:
            return []
        counts = st.label.class_counts.astype(np.float64)
        L_ = len(counts)
        sums = st.class_sums[:, plan.cat_cols]                    # [L, total categorical columns]
        if np.isnan(sums).any():
            raise ValueError("no per-label sums were tracked for some categorical group")
        # rows of every group, packed: lone indicators contribute [indicator; complement]
        data = np.empty(plan.total_rows * L_)
        mats = data.reshape(plan.total_rows, L_)
        mats[plan.row_of_col] = sums.T
        if len(plan.lone_rows):
            mats[plan.lone_rows + 1] = counts[None, :] - mats[plan.lone_rows]
        ng = len(plan.groups)
        cols = np.full(ng, L_, dtype=np.int32)
        offsets = plan.row_offsets * L_
        v, chi, mi = np.zeros(ng), np.zeros(ng), np.zeros(ng)
        pmi = np.zeros(plan.total_rows * L_)
        conf, sup = np.zeros(plan.total_rows), np.zeros(plan.total_rows)
        status = np.zeros(ng, dtype=np.int32)
        L.call("tmog_contingency_stats_batch", ng, plan.rows.ctypes.data, cols.ctypes.data, offsets.ctypes.data,
               plan.row_offsets.ctypes.data, data.ctypes.data, v.ctypes.data, chi.ctypes.data, mi.ctypes.data,
               pmi.ctypes.data, conf.ctypes.data, sup.ctypes.data, status.ctypes.data)
        # MultiPickList groups (OpStatistics.contingencyStatsFromMultiPickList :346-382): Cramer's V is the largest V of
        # the 2 x nLabels tables [choice; rest] over the non-empty choices; everything else stays the full table's
        for g in plan.mpl_groups:
            a, b = plan.row_offsets[g], plan.row_offsets[g] + plan.rows[g]
            m = mats[a:b]
            m = m[m.sum(axis=1) != 0.0]
            if len(m) and not (m.sum(axis=0) != 0.0).all():
                # filterEmpties drops the label column, and the reference then builds its 2 x numLabels matrices from
                # rows that are too short: DenseMatrix's size check fails (OpStatistics.scala:352-357)
                raise ValueError(f"requirement failed: The values array must have length {2 * m.shape[1]}: no row of "
                                 f"the MultiPickList group '{plan.groups[g][0]}' holds one of the label values")
            best = None
            for row in m:
                two = np.ascontiguousarray(np.stack([row, counts[:len(row)] - row]))
                v1, chi1, mi1 = np.zeros(1), np.zeros(1), np.zeros(1)
                L.call("tmog_contingency_stats", two.ctypes.data, 2, two.shape[1], v1.ctypes.data, chi1.ctypes.data,
                       mi1.ctypes.data, None, None, None)
                if best is None or v1[0] > best:          # maxBy keeps the first maximum; NaN never compares greater
                    best = float(v1[0])
            if best is not None:
                v[g] = best
        return _GroupStatsList(plan, mats, v, mi, pmi.reshape(plan.total_rows, L_), conf, sup, status)

    def _decide(self, meta, st, corr, corr_indices, corr_matrix, cat_stats):
        d = meta.size
        plan = self._plan(meta)
        if len(corr_indices) == d + 1:
            corr_pos = plan.identity_pos
        else:
            corr_pos = np.full(d, -1, dtype=np.int32)
            corr_pos[corr_indices[:-1]] = np.arange(len(corr_indices) - 1, dtype=np.int32)
        cram = np.full(d, np.nan)
        rule_n = np.full(d, -1, dtype=np.int32)
        rconf, rsup = np.zeros(2 * d), np.zeros(2 * d)
        if len(cat_stats):
            cs = cat_stats
            cram[plan.cat_cols] = cs.v[plan.group_of_col]
            # lone indicators (one column, two rows: indicator and complement): both rule confidences count
            ok = cs.status[plan.lone_groups] != 1
            lc, lr = plan.lone_cols, plan.lone_rows
            rule_n[lc] = np.where(ok, 2, 0)
            rconf[2 * lc], rconf[2 * lc + 1] = np.where(ok, cs.conf[lr], 0.0), np.where(ok, cs.conf[lr + 1], 0.0)
            rsup[2 * lc], rsup[2 * lc + 1] = np.where(ok, cs.sup[lr], 0.0), np.where(ok, cs.sup[lr + 1], 0.0)
            for g in plan.multi_groups:
                ia = np.asarray(plan.groups[g][2])
                a = plan.row_offsets[g]
                k = 0 if cs.status[g] == 1 else int(plan.rows[g])
                rule_n[ia] = 0
                rule_n[ia[:k]] = 1
                rconf[2 * ia[:k]] = cs.conf[a:a + k]
                rsup[2 * ia[:k]] = cs.sup[a:a + k]
        params = np.array([self.minVariance, self.minCorrelation, self.maxCorrelation, self.maxFeatureCorrelation,
                           self.maxCramersV, self.maxRuleConfidence, self.minRequiredRuleSupport,
                           float(self.removeFeatureGroup), float(self.protectTextSharedHash)])
        drop = np.zeros(d, dtype=np.uint8)
        reasons = np.zeros(d, dtype=np.uint32)
        pc, pv = np.zeros(d), np.zeros(d)
        var = np.ascontiguousarray(st.variance, dtype=np.float64)
        corr = np.ascontiguousarray(corr, dtype=np.float64)
        cm = None if corr_matrix is None else np.ascontiguousarray(corr_matrix, dtype=np.float64)
        L.call("tmog_sanity_decide", d, var.ctypes.data, corr.ctypes.data, None if cm is None else cm.ctypes.data,
               0 if cm is None else cm.shape[0], corr_pos.ctypes.data, plan.offs.ctypes.data, plan.pids.ctypes.data,
               plan.gid.ctypes.data, plan.tsh.ctypes.data, cram.ctypes.data, rule_n.ctypes.data, rconf.ctypes.data,
               rsup.ctypes.data, params.ctypes.data, drop.ctypes.data, reasons.ctypes.data, pc.ctypes.data,
               pv.ctypes.data)
        return drop, reasons, pc, pv


_RUN_INFO: Dict[OpVectorColumnMetadata, tuple] = {}
_PLAN_CACHE: Dict[tuple, "_MetaPlan"] = {}


class _MetaPlan:
    """Column metadata of one vector reduced to the integer arrays the C ABI wants (built once per metadata object,
    run by run): parent / group ids for tmog_sanity_decide, and the categorical groups of categoricalTests."""

    def __init__(self, meta: OpVectorMetadata):
        d = meta.size
        self.meta = meta
        parent_ids: Dict[str, int] = {}
        group_ids: Dict[str, int] = {}
        run_np, run_gid, run_tsh, run_hashed = [], [], [], []
        pid_list: List[int] = []
        long_runs = []                      # (position in pid_list, parent ids, repeat) of runs longer than one column
        groups: Dict[str, list] = {}
        counts = np.array([k for _, k in meta.runs], dtype=np.int64)
        for (c, k), s0 in zip(meta.runs, meta.run_starts[:-1]):
            info = _RUN_INFO.get(c)
            if info is None:                # per column description (immutable, shared between fits)
                if len(_RUN_INFO) > 1_000_000:
                    _RUN_INFO.clear()
                info = (tuple(c.parentNamesWithMapKeys()), c.featureGroup(), c.isTextSharedHash,
                        c.grouping is None and c.indicatorValue is None
                        and c.hasParentOfType(F.TEXT, F.TEXTAREA, "TextMap", "TextAreaMap"))
                _RUN_INFO[c] = info
            parents, g, tsh, hashed = info
            pids = [parent_ids.setdefault(pn, len(parent_ids)) for pn in parents]
            run_np.append(len(pids))
            if k == 1:
                pid_list.extend(pids)
            else:                           # a hashed block: hundreds of thousands of identical columns
                long_runs.append((len(pid_list), pids, k))
            run_gid.append(group_ids.setdefault(g, len(group_ids)) if g is not None else -1)
            run_tsh.append(tsh)
            run_hashed.append(hashed)
            if g is not None and c.indicatorValue is not None:
                for i in range(k):
                    groups.setdefault(g, []).append((c, int(s0) + i))
        per_col_np = np.repeat(np.asarray(run_np, dtype=np.int64), counts)
        self.offs = np.concatenate([[0], np.cumsum(per_col_np)]).astype(np.int32)
        pieces, at = [], 0
        for where, pids, k in long_runs:
            pieces.append(np.asarray(pid_list[at:where], dtype=np.int32))
            pieces.append(np.tile(np.asarray(pids, dtype=np.int32), k))
            at = where
        pieces.append(np.asarray(pid_list[at:], dtype=np.int32))
        self.pids = np.concatenate(pieces)
        self.gid = np.repeat(np.asarray(run_gid, dtype=np.int32), counts)
        self.tsh = np.repeat(np.asarray(run_tsh, dtype=np.uint8), counts)
        self.hashed_text = np.repeat(np.asarray(run_hashed, dtype=bool), counts)
        self.identity_pos = np.arange(d, dtype=np.int32)
        # categorical groups: duplicate indicator values inside a group -> only the first is used (:295-298)
        self.groups = []
        self.mpl_groups: List[int] = []
        cat_cols, row_of_col, rows, lone_rows = [], [], [], []
        r = 0
        for g, cols in groups.items():
            seen, cleaned = set(), []
            for c, i in cols:
                if c.indicatorValue not in seen:
                    seen.add(c.indicatorValue)
                    cleaned.append((c, i))
            idx = [i for _, i in cleaned]
            if any(c.isMultiPickList for c, _ in cleaned):
                self.mpl_groups.append(len(self.groups))
            self.groups.append((g, None, idx))        # the column names are looked up when somebody asks
            nrows = 2 if len(idx) == 1 else len(idx)
            if len(idx) == 1:
                lone_rows.append(r)
            cat_cols.extend(idx)
            row_of_col.extend(range(r, r + len(idx)))
            rows.append(nrows)
            r += nrows
        self.group_of_col = np.repeat(np.arange(len(self.groups), dtype=np.int64),
                                      [len(g[2]) for g in self.groups]) if self.groups else np.zeros(0, dtype=np.int64)
        self.lone_groups = np.asarray([g for g, grp in enumerate(self.groups) if len(grp[2]) == 1], dtype=np.int64)
        self.lone_cols = np.asarray([self.groups[g][2][0] for g in self.lone_groups], dtype=np.int64)
        self.multi_groups = [g for g, grp in enumerate(self.groups) if len(grp[2]) != 1]
        self.cat_cols = np.asarray(cat_cols, dtype=np.int64)
        self.row_of_col = np.asarray(row_of_col, dtype=np.int64)
        self.lone_rows = np.asarray(lone_rows, dtype=np.int64)
        self.rows = np.asarray(rows, dtype=np.int32)
        self.row_offsets = np.concatenate([[0], np.cumsum(rows)[:-1]]).astype(np.int64) if rows else \
            np.zeros(0, dtype=np.int64)
        self.total_rows = r


class SanityCheckerModel(OpModel):
    def __init__(self, parent: SanityChecker, indicesToKeep: List[int], removeBadFeatures: bool):  # noqa: N803
        super().__init__(parent)
        self._keep = np.asarray(indicesToKeep, dtype=np.int64)
        self._keep_list: Optional[List[int]] = None
        self.removeBadFeatures = removeBadFeatures

    @property
    def indicesToKeep(self) -> List[int]:  # noqa: N802 -- a Python list on demand (262 k entries for a hashed block)
        if self._keep_list is None:
            self._keep_list = self._keep.tolist()
        return self._keep_list

    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        vec = batch.vectors[self.inN[1].name]
        name = self.getOutput().name
        if not self.removeBadFeatures:
            return VectorColumn(name, batch.n, self.getMetadata(), vec.parts, vec.stats)
        from .generic_stats import select_columns
        parts = select_columns(vec, self._keep)
        stats = None if vec.stats is None else vec.stats.select(self._keep)
        return VectorColumn(name, batch.n, self.getMetadata(), parts, stats)
