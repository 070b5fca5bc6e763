"""Oracle: RawFeatureFilter's feature distributions, metrics and exclusions (TEST INFRASTRUCTURE ONLY -- nothing under
transmogrifai_b200/ may import this; the product path is transmogrifai_b200/filters.py + csrc/rff.cu).

A per-row restatement, statement by statement, of
  * core/src/main/scala/com/salesforce/op/filters/Summary.scala:43-65
  * core/src/main/scala/com/salesforce/op/filters/PreparedFeatures.scala:52-72 (summaries, null-label vector),
    :156-193 (prepareFeature: numerics -> Right(Seq(d)), Text -> Left(tokens), MultiPickList -> Left(values))
  * core/src/main/scala/com/salesforce/op/filters/FeatureDistribution.scala:94-163 (fillRate, reduce, fill ratio /
    difference, jsDivergence), :217-248 (fromSummary), :307-351 (histValues)
  * core/src/main/scala/com/salesforce/op/stages/impl/feature/NumericBucketizer.scala:219-264 (bucketize, Inclusion.Left,
    trackInvalid)
  * core/src/main/scala/com/salesforce/op/filters/RawFeatureFilter.scala:138-200 (computeFeatureStats), :211-296
    (metrics), :307-427 (exclusion reasons), :440-470 (getFeaturesToExclude)
Third-party pieces restated: algebird's Array monoid (element-wise sum, the shorter array padded with zeros), Spark
mllib HashingTF (oracle/text.py), Spark Statistics.corr (Pearson: NaN where a variance is <= 1e-12).
Not restated: `moments` and `cardEstimate` (reported only, no rule reads them), maps, dates, geolocations, vectors.

Pinned by tests/test_rff_oracle_golden.py against the reference's own vectors (FeatureDistributionTest.scala:50-175,
RawFeatureFilterTest.scala:156-268, FiltersTestData.scala:38-60).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

from . import text as T

INF = float("inf")
NUMERIC = ("Real", "RealNN", "Binary", "Integral", "Percent", "Currency")
TEXT = ("Text", "TextArea")
ProcessedSeq = Tuple[str, list]          # ("L", [strings]) | ("R", [doubles])
FeatureKey = Tuple[str, Optional[str]]


@dataclass
class Summary:
    """Summary.scala:43: min / max / sum / count of a numeric feature's values, or of a text's token counts."""
    min: float = INF
    max: float = -INF
    sum: float = 0.0
    count: float = 0.0

    def plus(self, r: "Summary") -> "Summary":                       # Summary.scala:49-54 (java Math.min/max: NaN wins)
        return Summary(_jmin(self.min, r.min), _jmax(self.max, r.max), self.sum + r.sum, self.count + r.count)

    def is_empty(self) -> bool:
        return self.min == INF and self.max == -INF and self.sum == 0.0 and self.count == 0.0


def _jmin(a: float, b: float) -> float:
    return float("nan") if (a != a or b != b) else min(a, b)


def _jmax(a: float, b: float) -> float:
    return float("nan") if (a != a or b != b) else max(a, b)


def summary_of(seq: ProcessedSeq) -> Summary:                        # Summary.scala:60-65
    kind, v = seq
    if kind == "L":
        return Summary(float(len(v)), float(len(v)), float(len(v)), 1.0)
    out = Summary()
    for d in v:
        out = out.plus(Summary(d, d, d, 1.0))
    return out


def prepare_feature(type_name: str, value) -> Optional[ProcessedSeq]:
    """PreparedFeatures.prepareFeature (:156-193) for the types of the path; None = the feature is empty in this row."""
    if value is None:
        return None
    if type_name in TEXT or type_name == "PickList":
        return ("L", T.tokenize(value))                 # TextTokenizer.Analyzer.analyze(s, Language.Unknown) :201
    if type_name == "MultiPickList":
        vals = list(dict.fromkeys(value))
        return ("L", vals) if vals else None            # an empty set is an empty feature
    if type_name == "Binary":
        return ("R", [1.0 if value else 0.0])
    if type_name in NUMERIC:
        return ("R", [float(value)])
    raise RuntimeError(f"Feature type {type_name} is not supported in RawFeatureFilter")


def array_plus(a: List[float], b: List[float]) -> List[float]:
    """algebird ArrayMonoid.plus: element-wise, the shorter one padded with zeros."""
    if len(a) < len(b):
        a, b = b, a
    return [x + (b[i] if i < len(b) else 0.0) for i, x in enumerate(a)]


def bucketize_left(splits: Sequence[float], v: float) -> int:
    """NumericBucketizer.bucketize with trackInvalid = true, trackNulls = false, Inclusion.Left (:219-264): the index
    of the 1.0 in a vector of len(splits) entries (the last one = invalid: NaN, +-Inf, outside the splits)."""
    nb = len(splits) - 1
    idx = -1
    for i in range(nb):
        if splits[i] <= v < splits[i + 1]:
            idx = i
            break
    if idx < 0 or math.isnan(v) or math.isinf(v):
        return nb
    return idx


@dataclass
class FeatureDistribution:
    """FeatureDistribution.scala:59-69 (without moments / cardEstimate)."""
    name: str
    key: Optional[str]
    count: int
    nulls: int
    distribution: List[float]
    summary_info: List[float]
    type: str = "Training"

    @property
    def feature_key(self) -> FeatureKey:
        return (self.name, self.key)

    def fill_rate(self) -> float:                                             # :99
        return 0.0 if self.count == 0 else (self.count - self.nulls) / float(self.count)

    def reduce(self, fd: "FeatureDistribution") -> "FeatureDistribution":     # :107-122
        assert self.name == fd.name and self.key == fd.key
        info = self.summary_info if len(self.summary_info) > len(fd.summary_info) else fd.summary_info
        return FeatureDistribution(self.name, self.key, self.count + fd.count, self.nulls + fd.nulls,
                                   array_plus(self.distribution, fd.distribution), info, self.type)

    def relative_fill_ratio(self, fd: "FeatureDistribution") -> float:        # :130-135
        a, b = self.fill_rate(), fd.fill_rate()
        small, large = (a, b) if a < b else (b, a)
        return INF if small == 0.0 else large / small

    def relative_fill_rate(self, fd: "FeatureDistribution") -> float:         # :143-146
        return abs(self.fill_rate() - fd.fill_rate())

    def js_divergence(self, fd: "FeatureDistribution") -> float:              # :154-163
        pairs = [(a, b) for a, b in zip(self.distribution, fd.distribution) if not (a == 0.0 and b == 0.0)]
        this_count = that_count = 0.0
        for a, b in pairs:
            this_count, that_count = this_count + a, that_count + b
        probs = [(_div(a, this_count), _div(b, that_count)) for a, b in pairs]

        def log2(x: float) -> float:
            return _log10(x) / math.log10(2.0)

        def kl(a: float, b: float) -> float:
            return 0.0 if a == 0.0 else a * log2(_div(a, b))

        total = 0.0
        for a, b in probs:
            m = (a + b) / 2
            total += 0.5 * kl(a, m) + 0.5 * kl(b, m)
        return total


def _div(a: float, b: float) -> float:
    """JVM double division (x / 0.0 = +-Inf or NaN instead of an exception)."""
    if b == 0.0:
        return float("nan") if (a == 0.0 or a != a) else math.copysign(INF, a)
    return a / b


def _log10(x: float) -> float:
    if x != x or x < 0:
        return float("nan")
    if x == 0.0:
        return -INF
    return math.log10(x)


def hist_values(seq: ProcessedSeq, summary: Summary, bins: int, text_bins_formula=None
                ) -> Tuple[List[float], List[float]]:
    """FeatureDistribution.histValues (:307-351) -> (summaryInfo, distribution) of ONE row's values."""
    kind, v = seq
    info4 = [summary.min, summary.max, summary.sum, summary.count]
    if kind == "L":
        num_bins = bins if text_bins_formula is None else text_bins_formula(summary, bins)
        dist = [0.0] * num_bins
        idxs, vals = T.hashing_tf_transform(v, num_bins, binary=False)
        for idx, tf in zip(idxs, vals):
            dist[idx] = float(tf)
        return info4, dist
    if summary.is_empty():
        return [summary.min, summary.max], [float(x) for x in v]
    if summary.min < summary.max:
        step = _div(summary.max - summary.min, bins - 2.0)            # bins = 2: x / 0.0 = Infinity on the JVM
        splits = [summary.min + step * b for b in range(bins)]
        hist = [0.0] * bins
        for x in v:
            hist[bucketize_left(splits, x)] += 1.0
        return splits, hist
    same = float(sum(1 for x in v if x == summary.max))
    other = float(sum(1 for x in v if x != summary.max))
    return info4, [same, other]


def from_summary(key: FeatureKey, summary: Summary, value: Optional[ProcessedSeq], bins: int, type_: str,
                 text_bins_formula=None) -> FeatureDistribution:
    """FeatureDistribution.fromSummary (:217-248): the distribution of ONE row."""
    if value is None:
        nulls, info, dist = 1, [summary.min, summary.max, summary.sum, summary.count], [0.0] * bins
    else:
        nulls = 0
        info, dist = hist_values(value, summary, bins, text_bins_formula)
    return FeatureDistribution(key[0], key[1], 1, nulls, dist, info, type_)


@dataclass
class AllFeatureInformation:                                      # AllFeatureInformation.scala:44-52
    response_summaries: Dict[FeatureKey, Summary]
    response_distributions: List[FeatureDistribution]
    predictor_summaries: Dict[FeatureKey, Summary]
    predictor_distributions: List[FeatureDistribution]
    correlation_info: Dict[FeatureKey, Dict[FeatureKey, float]]


def pearson_columns(cols: List[List[float]]) -> List[List[float]]:
    """Spark Statistics.corr(rows, "pearson"): covariance from the Gramian with n - 1, sigma = 0 where the variance is
    <= 1e-12, correlation NaN where a sigma is 0, diagonal 1.0."""
    n = len(cols[0]) if cols else 0
    d = len(cols)
    mean = [math.fsum(c) / n for c in cols]
    cov = [[(math.fsum(a * b for a, b in zip(cols[i], cols[j])) - n * mean[i] * mean[j]) / (n - 1) for j in range(d)]
           for i in range(d)]
    sigma = [0.0 if abs(cov[i][i]) <= 1e-12 else math.sqrt(cov[i][i]) for i in range(d)]
    out = [[float("nan")] * d for _ in range(d)]
    for i in range(d):
        for j in range(d):
            if i == j:
                out[i][j] = 1.0
            elif sigma[i] != 0.0 and sigma[j] != 0.0:
                out[i][j] = cov[i][j] / (sigma[i] * sigma[j])
    return out


def compute_feature_stats(columns: Dict[str, Tuple[str, Sequence]], responses: Sequence[str], bins: int,
                          type_: str = "Training", info: Optional[AllFeatureInformation] = None,
                          text_bins_formula=None) -> AllFeatureInformation:
    """RawFeatureFilter.computeFeatureStats (:138-200).  columns: {name: (type name, python values, None = empty)} in
    feature order; responses: names of the response features (only numeric ones are used, :146-155)."""
    names = list(columns)
    n = len(columns[names[0]][1]) if names else 0
    resp = [nm for nm in names if nm in responses and columns[nm][0] in NUMERIC]
    pred = [nm for nm in names if nm not in responses]
    prepared = {nm: [prepare_feature(columns[nm][0], v) for v in columns[nm][1]] for nm in resp + pred}

    def summaries(group):
        out: Dict[FeatureKey, Summary] = {}
        for r in range(n):                                          # preparedFeatures.map(_.summaries).reduce(_ + _)
            for nm in group:
                seq = prepared[nm][r]
                if seq is not None:                                 # an empty feature has no entry in the row's map
                    out[(nm, None)] = out[(nm, None)].plus(summary_of(seq)) if (nm, None) in out else summary_of(seq)
        # the reference's order is a Scala Map's iteration order (unspecified beyond four keys): feature order here
        return {(nm, None): out[(nm, None)] for nm in group if (nm, None) in out}

    if info is not None:
        rs, ps = info.response_summaries, info.predictor_summaries
    else:
        rs, ps = summaries(resp), summaries(pred)

    def distributions(summ: Dict[FeatureKey, Summary]) -> List[FeatureDistribution]:
        out = []
        for key, s in summ.items():
            acc = None
            for r in range(n):
                fd = from_summary(key, s, prepared[key[0]][r], bins, type_, text_bins_formula)
                acc = fd if acc is None else acc.reduce(fd)
            out.append(acc)
        return out

    rd, pd = distributions(rs), distributions(ps)
    if info is not None:
        corr_info = info.correlation_info
    else:
        rkeys, pkeys = list(rs), list(ps)
        cols = []
        for key in rkeys:                                           # getNullLabelLeakageVector (:62-70)
            cols.append([seq[1][0] if (seq is not None and seq[0] == "R" and len(seq[1]) == 1) else 0.0
                         for seq in prepared[key[0]]])
        for key in pkeys:
            cols.append([0.0 if seq is not None else 1.0 for seq in prepared[key[0]]])
        m = pearson_columns(cols) if cols and n > 0 else []
        corr_info = {rk: {pk: _jmin(abs(m[i][j + len(rkeys)]), 1.0) for j, pk in enumerate(pkeys)}
                     for i, rk in enumerate(rkeys)}
    return AllFeatureInformation(rs, rd, ps, pd, corr_info)


@dataclass
class RawFeatureFilterMetrics:                                    # RawFeatureFilterResults.scala (metrics record)
    name: str
    key: Optional[str]
    training_fill_rate: float
    training_null_label_absolute_corr: Optional[float]
    scoring_fill_rate: Optional[float]
    js_divergence: Optional[float]
    fill_rate_diff: Optional[float]
    fill_ratio_diff: Optional[float]


@dataclass
class ExclusionReasons:
    name: str
    key: Optional[str]
    training_unfilled_state: bool
    training_null_label_leaker: bool
    scoring_unfilled_state: bool
    js_divergence_mismatch: bool
    fill_rate_diff_mismatch: bool
    fill_ratio_diff_mismatch: bool
    excluded: bool


@dataclass
class FilterParams:
    bins: int = 100
    min_fill: float = 0.001
    max_fill_difference: float = 0.90
    max_fill_ratio_diff: float = 20.0
    max_js_divergence: float = 0.90
    max_correlation: float = 0.95
    js_divergence_protected: Sequence[str] = ()
    protected: Sequence[str] = ()
    min_scoring_rows: int = 500


def get_metrics(train: Sequence[FeatureDistribution], score: Sequence[FeatureDistribution],
                corr_info: Dict[FeatureKey, Dict[FeatureKey, float]]) -> List[RawFeatureFilterMetrics]:
    """RawFeatureFilter.getRawFeatureFilterMetrics (:211-296)."""
    k = len(train)
    fills = [d.fill_rate() for d in train]
    if not corr_info:
        corrs: List[Optional[float]] = [None] * k
    else:
        # `for (distrib <- trainingDistribs) yield correlationInfo.values.map(_.get(key))`, flattened (:225-231):
        # one entry per (distribution, response); zipped with the distributions it only lines up for ONE response
        corrs = []
        for d in train:
            for per_response in corr_info.values():
                corrs.append(per_response.get(d.feature_key))
    if not score:
        sf = js = fd = fr = [None] * k
    else:
        assert len(score) == k, "scoring and training features must match"
        sf = [s.fill_rate() for s in score]
        js = [t.js_divergence(s) for t, s in zip(train, score)]
        fd = [t.relative_fill_rate(s) for t, s in zip(train, score)]
        fr = [t.relative_fill_ratio(s) for t, s in zip(train, score)]
    return [RawFeatureFilterMetrics(d.name, d.key, fills[i], corrs[i], sf[i], js[i], fd[i], fr[i])
            for i, d in enumerate(train) if i < len(corrs)]           # zip truncates to the shortest


def get_exclusion_reasons(train, score, metrics: Sequence[RawFeatureFilterMetrics], p: FilterParams
                          ) -> List[ExclusionReasons]:
    """RawFeatureFilter.getRawFeatureFilterExclusionReasons (:307-427)."""
    out = []
    for d, m in zip(train, metrics):
        unfilled = m.training_fill_rate < p.min_fill
        c = m.training_null_label_absolute_corr
        leaker = c is not None and c > p.max_correlation
        if not score:
            s_unfilled = js_mis = fd_mis = fr_mis = False
        else:
            s_unfilled = m.scoring_fill_rate is not None and m.scoring_fill_rate < p.min_fill
            js_mis = (m.name not in p.js_divergence_protected and m.js_divergence is not None
                      and m.js_divergence > p.max_js_divergence)
            fd_mis = m.fill_rate_diff is not None and m.fill_rate_diff > p.max_fill_difference
            fr_mis = m.fill_ratio_diff is not None and m.fill_ratio_diff > p.max_fill_ratio_diff
        out.append(ExclusionReasons(d.name, d.key, unfilled, leaker, s_unfilled, js_mis, fd_mis, fr_mis,
                                    any([unfilled, leaker, s_unfilled, js_mis, fd_mis, fr_mis])))
    return out


def get_features_to_exclude(train, score, corr_info, p: FilterParams):
    """RawFeatureFilter.getFeaturesToExclude (:440-470) -> (metrics, reasons, names to drop, map keys to drop)."""
    metrics = get_metrics(train, score, corr_info)
    reasons = get_exclusion_reasons(train, score, metrics, p)
    dropped: Dict[str, List[FeatureDistribution]] = {}
    kept: Dict[str, List[FeatureDistribution]] = {}
    for d, r in zip(train, reasons):
        (dropped if r.excluded else kept).setdefault(d.name, []).append(d)
    map_keys = set(kept) & set(dropped)
    to_drop = [k for k in dropped if k not in map_keys]
    to_drop_keys = {k: {d.key for d in v if d.key is not None} for k, v in dropped.items() if k in map_keys}
    return metrics, reasons, to_drop, to_drop_keys


def generate_filtered_raw(train_cols, score_cols, responses: Sequence[str], p: FilterParams):
    """RawFeatureFilter.generateFilteredRaw (:482-566) up to the names of the features to drop."""
    n_train = len(next(iter(train_cols.values()))[1]) if train_cols else 0
    if n_train <= 0:
        raise ValueError("requirement failed: RawFeatureFilter cannot work with empty training data")
    ts = compute_feature_stats(train_cols, responses, p.bins, "Training")
    ss = None
    if score_cols is not None:
        n_score = len(next(iter(score_cols.values()))[1]) if score_cols else 0
        if n_score >= p.min_scoring_rows:
            ss = compute_feature_stats(score_cols, responses, p.bins, "Scoring", ts)
    keep = lambda ds: [d for d in ds if d.name not in p.protected]   # noqa: E731
    metrics, reasons, to_drop, to_drop_keys = get_features_to_exclude(
        keep(ts.predictor_distributions), keep(ss.predictor_distributions) if ss is not None else [],
        ts.correlation_info, p)
    predictors_left = [nm for nm in train_cols if nm not in responses and nm not in to_drop]
    if not predictors_left:
        raise ValueError("requirement failed: The raw feature filter has dropped all of your features, check your "
                         "input data quality")
    return ts, ss, metrics, reasons, to_drop, to_drop_keys
