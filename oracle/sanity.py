"""Oracle: SanityChecker statistics and the removeBadFeatures decision (TEST INFRASTRUCTURE ONLY).

Follows
  * core/src/main/scala/com/salesforce/op/stages/impl/preparators/SanityChecker.scala:252-348 (categoricalTests),
    :356-361 (fraction), :367-541 (fitFn), :556-558 (model transform), :561-581 (defaults)
  * core/.../preparators/DerivedFeatureFilterUtils.scala:95-214 (makeColumnStatistics), :234-281 (getFeaturesToDrop),
    :291-305 (removeFeatures), :350-417 (reasonsToRemove / isTextSharedHash)
  * utils/src/main/scala/com/salesforce/op/utils/stats/OpStatistics.scala:71-97 (label correlations), :141-175
    (filterEmpties), :202-211 (Cramer's V), :234-271 (mutual info), :280-290 (confidences), :300-326, :346-382
and restates the Spark 2.4.5 pieces that are not under /root/reference:
  * mllib.stat.Statistics.colStats (MultivariateOnlineSummarizer: count, mean, unbiased variance, min, max)
  * mllib.stat.Statistics.corr "pearson" (RowMatrix.computeCovariance + PearsonCorrelation: NaN where
    cov_ii <= 1e-12, diagonal 1.0)
  * mllib.stat.Statistics.chiSqTest(Matrix) (Pearson chi-squared, no continuity correction)
Sampling (:378-394) is not restated: every parity run uses settings for which fraction() == 1.0.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import scipy.sparse as sp

from .metadata import ColumnMeta, TEXT_TYPES


@dataclass
class SanityParams:
    """SanityChecker defaults, SanityChecker.scala:561-581"""
    check_sample: float = 1.0
    sample_lower_limit: int = 1000
    sample_upper_limit: int = 1000000
    max_correlation: float = 0.95
    max_feature_corr: float = 0.99
    min_correlation: float = 0.0
    min_variance: float = 1e-5
    max_cramers_v: float = 0.95
    remove_bad_features: bool = False
    remove_feature_group: bool = True
    protect_text_shared_hash: bool = False
    max_rule_confidence: float = 1.0
    min_required_rule_support: float = 1.0
    feature_feature_corr_level: str = "computed"   # "computed" | "off"
    correlation_exclusion: str = "none"            # "none" | "hashed_text"
    categorical_label: Optional[bool] = None


def fraction(p: SanityParams, total: int) -> float:
    """SanityChecker.fraction :356-361"""
    min_f = min(1.0, p.sample_lower_limit / total)
    max_f = max(0.0, p.sample_upper_limit / total)
    return max(min(p.check_sample, max_f), min_f)


# ----------------------------------------------------------------------------------------------------------------
# Spark Statistics.colStats / corr
# ----------------------------------------------------------------------------------------------------------------
@dataclass
class ColStats:
    count: int
    mean: np.ndarray
    variance: np.ndarray
    min: np.ndarray
    max: np.ndarray


def col_stats(x) -> ColStats:
    """Unbiased sample variance over all n rows (zeros included); 0 when n <= 1.  Computed two-pass in float64."""
    x = np.asarray(x.todense()) if sp.issparse(x) else np.asarray(x, dtype=np.float64)
    n = x.shape[0]
    mean = x.mean(axis=0)
    if n > 1:
        d = x - mean
        var = (d * d).sum(axis=0) / (n - 1)
    else:
        var = np.zeros(x.shape[1])
    return ColStats(n, mean, var, x.min(axis=0), x.max(axis=0))


def correlations_with_label(x: np.ndarray, st: ColStats) -> np.ndarray:
    """OpStatistics.computeCorrelationsWithLabel :71-97; label = last column; 0/0 -> NaN kept."""
    n = x.shape[0]
    if n <= 1:
        raise ValueError(f"computeCorrelationsWithLabel called on matrix with only {n} rows.")
    d = x - st.mean
    cov = (d * d[:, -1:]).sum(axis=0)
    sd = np.sqrt(st.variance)
    with np.errstate(divide="ignore", invalid="ignore"):
        return cov / (sd * sd[-1] * (n - 1))


def pearson_matrix(x: np.ndarray) -> np.ndarray:
    """Statistics.corr(rows, "pearson"): covariance / (sigma_i sigma_j); sigma := 0 where cov_ii <= 1e-12 -> NaN
    row/column; diagonal forced to 1.0."""
    n, d = x.shape
    if d > 65535:
        raise ValueError(f"Argument with more than 65535 cols: {d}")
    mean = x.mean(axis=0)
    c = x - mean
    cov = (c.T @ c) / (n - 1)
    sig = np.sqrt(np.where(np.abs(np.diag(cov)) <= 1e-12, 0.0, np.diag(cov)))
    with np.errstate(divide="ignore", invalid="ignore"):
        corr = cov / np.outer(sig, sig)
    corr[sig == 0.0, :] = np.nan
    corr[:, sig == 0.0] = np.nan
    np.fill_diagonal(corr, 1.0)
    return corr


# ----------------------------------------------------------------------------------------------------------------
# Contingency statistics (OpStatistics)
# ----------------------------------------------------------------------------------------------------------------
def filter_empties(m: np.ndarray) -> np.ndarray:
    """:141-175 -- drop all-zero rows, then all-zero columns."""
    if m.size == 0:   # Matrix.numActives <= 0 (a dense matrix counts every stored entry)
        return m
    rows = m.sum(axis=1) != 0.0
    fr = m[rows, :]
    cols = fr.sum(axis=0) != 0.0
    return fr[:, cols]


@dataclass
class ContingencyStats:
    cramers_v: float
    chi_squared: float
    pmi: Dict[str, List[float]]
    mutual_info: float
    max_confidences: List[float]
    supports: List[float]
    matrix: np.ndarray


def chi_squared_on_filtered(f: np.ndarray) -> Tuple[float, float]:
    """:202-211 + Spark ChiSqTest.chiSquaredMatrix: stat = sum (O-E)^2/E, E = row*col/N"""
    if f.shape[0] > 1 and f.shape[1] > 1:
        total = f.sum()
        row = f.sum(axis=1, keepdims=True)
        col = f.sum(axis=0, keepdims=True)
        e = row * col / total
        stat = float((((f - e) ** 2) / e).sum())
        phi2 = stat / total
        return math.sqrt(phi2 / min(f.shape[0] - 1, f.shape[1] - 1)), stat
    return float("nan"), float("nan")


def mutual_info(m: np.ndarray) -> Tuple[Dict[str, List[float]], float]:
    """:234-271 on the UNFILTERED matrix (as contingencyStats :314 does)"""
    rows, cols = m.shape
    row_sum = m.sum(axis=1)
    col_sum = m.sum(axis=0)
    total = row_sum.sum()
    pmi = np.zeros((rows, cols))
    for j in range(cols):
        for i in range(rows):
            v = m[i, j]
            if v == 0 or row_sum[i] == 0 or col_sum[j] == 0:
                pmi[i, j] = 0.0
            else:
                pmi[i, j] = math.log(max(v, 1e-99) * total / (row_sum[i] * col_sum[j])) / math.log(2.0)
    mi = float((pmi * m / total).sum())
    return {str(j): list(pmi[:, j]) for j in range(cols)}, mi


def max_confidences(m: np.ndarray) -> Tuple[List[float], List[float]]:
    """:280-290"""
    row_sum = m.sum(axis=1)
    total = row_sum.sum()
    supports = [float(r / total) for r in row_sum]
    conf = [0.0 if r == 0 else float(m[i].max() / r) for i, r in enumerate(row_sum)]
    return conf, supports


def contingency_stats(m: np.ndarray) -> ContingencyStats:
    """:300-326"""
    m = np.asarray(m, dtype=np.float64)
    f = filter_empties(m)
    if f.size == 0:   # everything was filtered away
        return ContingencyStats(float("nan"), float("nan"), {}, float("nan"), [], [], m)
    pmi, mi = mutual_info(m)
    v, stat = chi_squared_on_filtered(f)
    conf, sup = max_confidences(m)
    return ContingencyStats(v, stat, pmi, mi, conf, sup, m)


def contingency_stats_multipicklist(m: np.ndarray, label_counts: np.ndarray) -> ContingencyStats:
    """:346-382 -- max over choices of the 2 x nLabels Cramer's V"""
    m = np.asarray(m, dtype=np.float64)
    f = filter_empties(m)
    singles = []
    for row in f:
        two = np.stack([row, np.asarray(label_counts[:len(row)]) - row])
        singles.append(contingency_stats(two))
    full = contingency_stats(m)
    if not singles:
        return full
    # Scala maxBy keeps the first maximum; a NaN never compares greater
    best = singles[0]
    for s in singles[1:]:
        if s.cramers_v > best.cramers_v:
            best = s
    return ContingencyStats(best.cramers_v, best.chi_squared, full.pmi, full.mutual_info, full.max_confidences,
                            full.supports, m)


@dataclass
class CategoricalGroupStats:
    group: str
    categorical_features: List[str]
    contingency_matrix: np.ndarray       # rows = choices, cols = labels (ascending label value)
    cramers_v: float
    mutual_info: float
    pmi: Dict[str, List[float]]
    max_rule_confidences: List[float]
    supports: List[float]


def categorical_tests(labels: np.ndarray, x, metas: Sequence[ColumnMeta], p: SanityParams
                      ) -> List[CategoricalGroupStats]:
    """SanityChecker.categoricalTests :252-348"""
    n, d = x.shape
    xs = sp.csr_matrix(x)
    mpl_idx = [c.index for c in metas if c.has_parent_of_type("MultiPickList")]
    if mpl_idx:
        xs = xs.tolil(copy=True)
        for i in mpl_idx:
            col = xs[:, i].toarray().ravel()
            xs[:, i] = np.minimum(col, 1.0).reshape(-1, 1)
        xs = xs.tocsr()
    distinct = np.unique(labels)  # ascending == sortBy(_._1)
    if not ((p.categorical_label is True) or len(distinct) < min(100.0, n * 0.1)):
        return []
    # per label: column sums + row count in slot d
    cont = np.zeros((len(distinct), d + 1))
    for li, lv in enumerate(distinct):
        sel = np.nonzero(labels == lv)[0]
        cont[li, :d] = np.asarray(xs[sel].sum(axis=0)).ravel()
        cont[li, d] = len(sel)
    with_ind = [c for c in metas if c.grouping is not None and c.indicator is not None]
    groups: Dict[str, List[ColumnMeta]] = {}
    for c in with_ind:
        groups.setdefault(c.feature_group(), []).append(c)
    out: List[CategoricalGroupStats] = []
    for g, cols in groups.items():
        seen = set()
        cleaned = []
        for c in cols:  # duplicate indicator values inside a group: only the first is used (:295-298)
            if c.indicator in seen:
                continue
            seen.add(c.indicator)
            cleaned.append(c)
        idx = [c.index for c in cleaned]
        is_mpl = any(c.has_parent_of_type("MultiPickList") for c in cleaned)
        if len(idx) == 1:
            ind = cont[:, idx[0]]
            mat = np.stack([ind, cont[:, d] - ind])            # 2 x nLabels
        else:
            mat = cont[:, idx].T.copy()                          # nChoices x nLabels
        st = contingency_stats_multipicklist(mat, cont[:, d]) if is_mpl else contingency_stats(mat)
        out.append(CategoricalGroupStats(g, [c.col_name() for c in cleaned], mat, st.cramers_v, st.mutual_info,
                                         st.pmi, st.max_confidences, st.supports))
    return out


# ----------------------------------------------------------------------------------------------------------------
# Column statistics and the drop decision (DerivedFeatureFilterUtils)
# ----------------------------------------------------------------------------------------------------------------
@dataclass
class ColumnStatistics:
    name: str
    column: Optional[ColumnMeta]
    is_label: bool
    count: int
    mean: float
    min: float
    max: float
    variance: float
    corr_label: Optional[float]
    cramers_v: Optional[float]
    parent_corr: Optional[float]
    parent_cramers_v: Optional[float]
    feature_corrs: List[float] = field(default_factory=list)
    max_rule_confidences: List[float] = field(default_factory=list)
    supports: List[float] = field(default_factory=list)


def _max_by_parent(pairs: Sequence[Tuple[str, float]]) -> Dict[str, float]:
    out: Dict[str, float] = {}
    seen: Dict[str, List[float]] = {}
    for k, v in pairs:
        seen.setdefault(k, []).append(v)
    for k, vs in seen.items():
        m = 0.0
        for v in vs:
            if not math.isnan(v):
                m = max(m, abs(v))
        out[k] = m
    return out


def make_column_statistics(metas: Sequence[ColumnMeta], st: ColStats, label_name: str, corrs_with_label: np.ndarray,
                           corr_indices: Sequence[int], cat_stats: Sequence[CategoricalGroupStats],
                           corr_matrix: Optional[np.ndarray]) -> List[ColumnStatistics]:
    """:95-214"""
    corr_pos = {ci: k for k, ci in enumerate(corr_indices)}
    cramers_map: Dict[str, float] = {}
    for f in cat_stats:
        for c in f.categorical_features:
            cramers_map[c] = f.cramers_v
    corr_parent = _max_by_parent([(pn, float(corrs_with_label[corr_pos[c.index]]))
                                  for c in metas if c.index in corr_pos for pn in c.parent_names_with_map_keys()])
    parent_of = {}
    for c in metas:
        for pn in c.parent_names_with_map_keys():
            parent_of[c.col_name()] = pn  # .toMap keeps the LAST parent for a multi-parent column
    cramers_parent = _max_by_parent([(parent_of[k], v) for k, v in cramers_map.items()])
    support_map: Dict[str, List[float]] = {}
    conf_map: Dict[str, List[float]] = {}
    for f in cat_stats:
        if len(f.categorical_features) == 1:
            support_map[f.categorical_features[0]] = list(f.supports)
            conf_map[f.categorical_features[0]] = list(f.max_rule_confidences)
        else:
            for nm, s in zip(f.categorical_features, f.supports):
                support_map[nm] = [s]
            for nm, s in zip(f.categorical_features, f.max_rule_confidences):
                conf_map[nm] = [s]

    def parent_value(col: ColumnMeta, check: Dict[str, float]) -> Optional[float]:
        vals = [check[k] for k in col.parent_names_with_map_keys() if k in check]
        return max(vals) if vals else None

    d = len(metas)
    out: List[ColumnStatistics] = []
    for col in metas:
        i = col.index
        name = col.col_name()
        k = corr_pos.get(i)
        fc: List[float] = []
        if k is not None and corr_matrix is not None:
            fc = list(corr_matrix[:, k][:-1])  # every row's k-th entry, label row dropped
        out.append(ColumnStatistics(
            name=name, column=col, is_label=False, count=st.count, mean=float(st.mean[i]), min=float(st.min[i]),
            max=float(st.max[i]), variance=float(st.variance[i]),
            corr_label=None if k is None else float(corrs_with_label[k]),
            cramers_v=cramers_map.get(name), parent_corr=parent_value(col, corr_parent),
            parent_cramers_v=parent_value(col, cramers_parent), feature_corrs=fc,
            max_rule_confidences=conf_map.get(name, []), supports=support_map.get(name, [])))
    label = ColumnStatistics(name=label_name, column=None, is_label=True, count=st.count, mean=float(st.mean[d]),
                             min=float(st.min[d]), max=float(st.max[d]), variance=float(st.variance[d]),
                             corr_label=None, cramers_v=None, parent_corr=None, parent_cramers_v=None)
    return [label] + out


def is_text_shared_hash(c: ColumnMeta) -> bool:
    """:410-414"""
    return c.has_parent_of_type(*TEXT_TYPES) and c.grouping is None and c.indicator is None


def reasons_to_remove(cs: ColumnStatistics, p: SanityParams, removed_groups: Sequence[str]) -> List[str]:
    """ColumnStatistics.reasonsToRemove :350-405 (reason strings shortened; only emptiness matters)"""
    if cs.is_label:
        return []
    r: List[str] = []
    if cs.variance <= p.min_variance:
        r.append("variance")
    if cs.corr_label is not None and abs(cs.corr_label) < p.min_correlation:
        r.append("minCorrelation")
    if cs.corr_label is not None and abs(cs.corr_label) > p.max_correlation:
        r.append("maxCorrelation")
    if cs.column is not None and any(abs(v) > p.max_feature_corr for v in cs.feature_corrs[:cs.column.index]):
        r.append("maxFeatureCorr")
    if cs.cramers_v is not None and cs.cramers_v > p.max_cramers_v:
        r.append("maxCramersV")
    if any(c > p.max_rule_confidence and s > p.min_required_rule_support
           for c, s in zip(cs.max_rule_confidences, cs.supports)):
        r.append("ruleConfidence")
    if cs.column is not None and cs.column.feature_group() is not None \
            and cs.column.feature_group() in removed_groups:
        r.append("ruleConfidenceGroup")
    col_is_shared = True if cs.column is None else is_text_shared_hash(cs.column)
    if p.remove_feature_group and ((not col_is_shared) or (not p.protect_text_shared_hash)):
        if cs.parent_cramers_v is not None and cs.parent_cramers_v > p.max_cramers_v:
            r.append("parentCramersV")
        if cs.parent_corr is not None and cs.parent_corr > p.max_correlation:
            r.append("parentCorr")
    return r


def features_to_drop(stats: Sequence[ColumnStatistics], p: SanityParams) -> List[ColumnStatistics]:
    """getFeaturesToDrop :234-281"""
    by_group: Dict[Optional[str], List[ColumnStatistics]] = {}
    for s in stats:
        g = s.column.feature_group() if s.column is not None else None
        by_group.setdefault(g, []).append(s)
    removed_groups = [g for g, cols in by_group.items() if g is not None and any(
        any(c > p.max_rule_confidence and sup > p.min_required_rule_support
            for c, sup in zip(f.max_rule_confidences, f.supports)) for f in cols)]
    return [s for s in stats if reasons_to_remove(s, p, removed_groups)]


@dataclass
class SanityResult:
    indices_to_keep: List[int]
    dropped: List[str]
    stats: List[ColumnStatistics]
    cat_stats: List[CategoricalGroupStats]
    col_stats: ColStats
    corrs_with_label: np.ndarray
    corr_matrix: Optional[np.ndarray]
    corr_indices: List[int]


def sanity_fit(labels: Sequence[float], x, metas: Sequence[ColumnMeta], p: SanityParams,
               label_name: str = "label") -> SanityResult:
    """SanityChecker.fitFn :367-541 with fraction() == 1.0 (no sampling)."""
    labels = np.asarray(labels, dtype=np.float64)
    n = len(labels)
    if fraction(p, n) < 1.0:
        raise NotImplementedError("oracle runs only un-sampled configurations (fraction() == 1.0)")
    dense = np.asarray(x.todense()) if sp.issparse(x) else np.asarray(x, dtype=np.float64)
    if n == 0:
        raise ValueError("requirement failed: Sample size cannot be zero")
    d = dense.shape[1]
    if d == 0:
        raise ValueError("requirement failed: Feature vector passed in is empty, check your vectorizers")
    if d != len(metas):
        raise ValueError(f"requirement failed: Number of columns in vector metadata ({len(metas)}) did not match "
                         f"number of columns in data({d})")
    full = np.concatenate([dense, labels[:, None]], axis=1)
    st = col_stats(full)
    if p.correlation_exclusion == "hashed_text":
        hashed = {c.index for c in metas if c.grouping is None and c.indicator is None
                  and c.has_parent_of_type(*TEXT_TYPES)}
        corr_indices = [i for i in range(d + 1) if i not in hashed]
    else:
        corr_indices = list(range(d + 1))
    sub = full[:, corr_indices]
    if p.feature_feature_corr_level == "off":
        sub_st = ColStats(st.count, st.mean[corr_indices], st.variance[corr_indices], st.min[corr_indices],
                          st.max[corr_indices])
        corr_matrix = None
        corrs = correlations_with_label(sub, sub_st)
    else:
        corr_matrix = pearson_matrix(sub)
        corrs = corr_matrix[:, -1].copy()
    cat = [] if p.categorical_label is False else categorical_tests(labels, dense, metas, p)
    stats = make_column_statistics(metas, st, label_name, corrs, corr_indices, cat, corr_matrix)
    to_drop = features_to_drop(stats, p) if p.remove_bad_features else []
    drop_idx = {s.column.index for s in to_drop if s.column is not None}
    keep = [c.index for c in metas if c.index not in drop_idx]
    if not keep:
        raise ValueError("requirement failed: The sanity checker has dropped all of your features, "
                         "check your input data quality")
    return SanityResult(keep, [s.name for s in to_drop], stats, cat, st, corrs, corr_matrix, corr_indices)


def sanity_transform(x, indices_to_keep: Sequence[int], remove_bad_features: bool):
    """SanityCheckerModel.transformFn -> removeFeatures :291-305"""
    if not remove_bad_features:
        return x
    return x[:, list(indices_to_keep)]
