"""Oracle: the transmogrify() vectorizers, row by row as the reference does them (TEST INFRASTRUCTURE ONLY).

Plain Python / numpy restatement.  Inputs are Python lists with ``None`` for a missing value (Spark null); outputs are
``scipy.sparse.csr_matrix`` (logical vector content -- Spark's ``.compressed`` dense/sparse choice is representation
only) plus the column metadata of ``oracle.metadata``.

Follows (all under core/src/main/scala/com/salesforce/op/stages/impl/feature/ unless a path is given)
  * RealVectorizer.scala:85-123, utils/.../spark/SequenceAggregators.scala:76-88      (mean fill + null indicators)
  * RealNNVectorizer.scala:52-55, BinaryVectorizer.scala:70-92
  * OpOneHotVectorizer.scala:75-124 (fit), :171-177/:219-225 (row -> map), :388-399 (metadata), :416-437 (pivotFn)
  * Transmogrifier.scala:437-475 (reindex / makeSparseVector)
  * SmartTextVectorizer.scala:79-152 (fit + decision), :207-299 (TextStats), :333-375 (transform)
  * OPCollectionHashingVectorizer.scala:191-305 (HashingFun)
  * VectorsCombiner.scala:64-88, features/.../utils/spark/RichVector.scala:120-134
"""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np
import scipy.sparse as sp

from .metadata import (ColumnMeta, NULL_STRING, OTHER_STRING, TEXT_LEN_STRING, feature_column, flatten, reindex)
from .text import clean_text_fn, hashing_tf_index, tokenize

Entries = List[Tuple[int, float]]  # one row's (index, value) pairs


def rows_to_csr(rows: Sequence[Entries], width: int) -> sp.csr_matrix:
    """Logical content of a Seq[(Int, Double)] per row; explicit zeros are dropped (RichVector.combine :126 and
    Vector.compressed both discard them), duplicates summed."""
    indptr = [0]
    idx: List[int] = []
    val: List[float] = []
    for r in rows:
        acc: Dict[int, float] = {}
        for i, v in r:
            acc[i] = acc.get(i, 0.0) + v
        for i in sorted(acc):
            if acc[i] != 0.0:
                idx.append(i)
                val.append(acc[i])
        indptr.append(len(idx))
    return sp.csr_matrix((np.asarray(val, dtype=np.float64), np.asarray(idx, dtype=np.int64),
                          np.asarray(indptr, dtype=np.int64)), shape=(len(rows), width))


# ----------------------------------------------------------------------------------------------------------------
# Real / RealNN / Binary
# ----------------------------------------------------------------------------------------------------------------
def mean_seq_null_num(columns: Sequence[Sequence[Optional[float]]]) -> List[float]:
    """SequenceAggregators.MeanSeqNullNum (:76-88): (sum, Int count) of the valid values; s / c if c > 0 else s."""
    out = []
    for col in columns:
        vals = [v for v in col if v is not None]
        s = math.fsum(vals)
        c = len(vals)
        out.append(s / c if c > 0 else s)
    return out


def real_vectorizer_fit(columns: Sequence[Sequence[Optional[float]]], fill_with_mean: bool = True,
                        fill_value: float = 0.0) -> List[float]:
    """RealVectorizer.fitFn :91-97 -> fillValues"""
    return mean_seq_null_num(columns) if fill_with_mean else [float(fill_value)] * len(columns)


def real_vectorizer_transform(columns: Sequence[Sequence[Optional[float]]], fill_values: Sequence[float],
                              track_nulls: bool = True) -> np.ndarray:
    """RealVectorizerModel.transformFn :112-123 -- dense, [value-or-fill, isNull] interleaved per column."""
    n = len(columns[0]) if columns else 0
    w = (2 if track_nulls else 1) * len(columns)
    out = np.zeros((n, w), dtype=np.float64)
    for j, (col, m) in enumerate(zip(columns, fill_values)):
        for i, v in enumerate(col):
            if track_nulls:
                out[i, 2 * j] = m if v is None else v
                out[i, 2 * j + 1] = 1.0 if v is None else 0.0
            else:
                out[i, j] = m if v is None else v
    return out


def mode_seq_null_int(columns: Sequence[Sequence[Optional[int]]]) -> List[int]:
    """SequenceAggregators.ModeSeqNullInt (utils/src/main/scala/com/salesforce/op/utils/spark/SequenceAggregators.scala
    :100-131): per column the value -> count map of the non-null rows; finish = minBy(-count, value), 0 if empty."""
    out = []
    for col in columns:
        m: Dict[int, int] = {}
        for v in col:
            if v is not None:
                m[int(v)] = m.get(int(v), 0) + 1
        out.append(min(m.items(), key=lambda kv: (-kv[1], kv[0]))[0] if m else 0)
    return out


def integral_vectorizer_fit(columns: Sequence[Sequence[Optional[int]]], fill_with_mode: bool = False,
                            fill_value: int = 0) -> List[int]:
    """IntegralVectorizer.fitFn (core/src/main/scala/com/salesforce/op/stages/impl/feature/IntegralVectorizer.scala
    :81-90): withConstant defaults to true."""
    return mode_seq_null_int(columns) if fill_with_mode else [int(fill_value)] * len(columns)


def integral_vectorizer_transform(columns: Sequence[Sequence[Optional[int]]], fill_values: Sequence[int],
                                  track_nulls: bool = True) -> np.ndarray:
    """IntegralVectorizerModel.transformFn :103-113: value-or-fill .toDouble, [v, isNull] interleaved."""
    cols = [[None if v is None else float(int(v)) for v in c] for c in columns]
    return real_vectorizer_transform(cols, [float(int(m)) for m in fill_values], track_nulls)


def numeric_metadata(names: Sequence[str], type_name: str, track_nulls: bool) -> List[ColumnMeta]:
    """VectorizerDefaults.vectorMetadata, Transmogrifier.scala:378-384"""
    cols: List[ColumnMeta] = []
    for nm in names:
        cols.append(feature_column(nm, type_name))
        if track_nulls:
            cols.append(feature_column(nm, type_name, is_null=True))
    return reindex(cols)


def realnn_vectorizer_transform(columns: Sequence[Sequence[float]]) -> np.ndarray:
    """RealNNVectorizer.transformFn :52-55 (value.get throws on a null)"""
    for col in columns:
        if any(v is None for v in col):
            raise ValueError("None.get: RealNN feature holds a null")
    return np.asarray(columns, dtype=np.float64).T.copy() if columns else np.zeros((0, 0))


def binary_vectorizer_transform(columns: Sequence[Sequence[Optional[bool]]], track_nulls: bool = True,
                                fill_value: bool = False) -> np.ndarray:
    """BinaryVectorizer.transformFn :70-82"""
    n = len(columns[0]) if columns else 0
    w = (2 if track_nulls else 1) * len(columns)
    out = np.zeros((n, w), dtype=np.float64)
    for j, col in enumerate(columns):
        for i, v in enumerate(col):
            if track_nulls:
                out[i, 2 * j] = float(fill_value) if v is None else float(bool(v))
                out[i, 2 * j + 1] = 1.0 if v is None else 0.0
            else:
                out[i, j] = float(fill_value) if v is None else float(bool(v))
    return out


# ----------------------------------------------------------------------------------------------------------------
# Pivot (OpOneHotVectorizer family)
# ----------------------------------------------------------------------------------------------------------------
def _as_set(cell) -> List[str]:
    """convertToSet: Text/PickList -> Option.toSet ; MultiPickList -> the Set itself."""
    if cell is None:
        return []
    if isinstance(cell, str):
        return [cell]
    return list(OrderedDict.fromkeys(cell))  # a Scala Set: duplicates collapse


def java_string_key(s: str):
    """java.lang.String.compareTo order = UTF-16 code-unit order."""
    return s.encode("utf-16-be", "surrogatepass")


def top_values(counts: Dict[str, int], top_k: int, min_support: int) -> List[str]:
    """filter(count >= minSupport).sortBy(-count -> value).take(topK)   (OpOneHotVectorizer.scala:100-103)"""
    kept = [(v, c) for v, c in counts.items() if c >= min_support]
    kept.sort(key=lambda vc: (-vc[1], java_string_key(vc[0])))
    return [v for v, _ in kept[:top_k]]


def pivot_value_counts(column: Sequence, clean_text: bool) -> Dict[str, int]:
    """convertToSeqOfMaps + the Algebird map-monoid fold (:95-97, :171-177, :219-225): every row contributes
    ``cleaned -> 1`` for each distinct cleaned value it holds (``.toMap`` collapses raw values that clean alike)."""
    counts: Dict[str, int] = {}
    for cell in column:
        keys = {clean_text_fn(str(v), clean_text) for v in _as_set(cell)}
        for k in keys:
            counts[k] = counts.get(k, 0) + 1
    return counts


def pivot_fit(columns: Sequence[Sequence], top_k: int = 20, min_support: int = 10,
              clean_text: bool = True) -> List[List[str]]:
    """OpOneHotVectorizer.fitFn with maxPctCardinality == 1.0 (default; HLL branch skipped)"""
    return [top_values(pivot_value_counts(c, clean_text), top_k, min_support) for c in columns]


def pivot_row_entries(cell, top: Sequence[str], clean_text: bool, track_nulls: bool) -> Entries:
    """One column's part of OneHotModelFun.pivotFn (:420-433)."""
    these: Dict[str, int] = {}
    for v in _as_set(cell):
        k = clean_text_fn(str(v), clean_text)
        these[k] = these.get(k, 0) + 1
    ent: Entries = [(i, float(these[c])) for i, c in enumerate(top) if c in these]
    top_set = set(top)
    other = float(sum(cnt for k, cnt in these.items() if k not in top_set))
    ent.append((len(top), other))
    if track_nulls:
        ent.append((len(top) + 1, 1.0 if not these else 0.0))
    return ent


def reindex_entries(seq: Sequence[Entries]) -> Entries:
    """VectorizerUtils.reindex, Transmogrifier.scala:437-449"""
    acc: Entries = []
    nxt = 0
    for curr in seq:
        ind = nxt
        for i, v in curr:
            ind = i + nxt
            acc.append((ind, v))
        nxt = ind + 1
    return acc


def pivot_width(tops: Sequence[Sequence[str]], track_nulls: bool) -> int:
    return sum(len(t) + (2 if track_nulls else 1) for t in tops)


def pivot_transform(columns: Sequence[Sequence], tops: Sequence[Sequence[str]], clean_text: bool = True,
                    track_nulls: bool = True) -> sp.csr_matrix:
    n = len(columns[0]) if columns else 0
    rows = []
    for r in range(n):
        rows.append(reindex_entries([pivot_row_entries(col[r], top, clean_text, track_nulls)
                                     for col, top in zip(columns, tops)]))
    return rows_to_csr(rows, pivot_width(tops, track_nulls))


def pivot_metadata(names: Sequence[str], type_name: str, tops: Sequence[Sequence[str]], track_nulls: bool,
                   unseen: str = OTHER_STRING) -> List[ColumnMeta]:
    """OneHotFun.makeVectorColumnMetadata (:388-399): grouping = feature name, indicator = value / OTHER / null."""
    cols: List[ColumnMeta] = []
    for nm, top in zip(names, tops):
        vals = list(top) + [unseen] + ([NULL_STRING] if track_nulls else [])
        for v in vals:
            cols.append(ColumnMeta((nm,), (type_name,), grouping=nm, indicator=v))
    return reindex(cols)


# ----------------------------------------------------------------------------------------------------------------
# Hashing (HashingFun)
# ----------------------------------------------------------------------------------------------------------------
MAX_NUM_OF_FEATURES = 1 << 17  # TransmogrifierDefaults.MaxNumOfFeatures


def is_shared_hash_space(num_features: int, num_inputs: int, strategy: str = "auto",
                         max_num_of_features: int = MAX_NUM_OF_FEATURES) -> bool:
    """HashingFun.isSharedHashSpace :191-199"""
    s = strategy.lower()
    if s == "shared":
        return True
    if s == "separate":
        return False
    return num_features * num_inputs > max_num_of_features


def hash_token_lists(token_lists: Sequence[Sequence[str]], feature_names: Sequence[str], num_features: int,
                     shared: bool, prepend_feature_name: bool = True, binary_freq: bool = False) -> Entries:
    """HashingFun.hash (:244-272) for ONE row.  ``feature_names`` is the ``features`` argument (zipped with ``in`` in
    order -- the reference's own alignment, including its quirk when Ignore columns exist)."""
    if not token_lists:
        return []
    name_hashes = [hashing_tf_index(nm, num_features) for nm in feature_names]
    blocks: List[Entries] = []
    shared_tf: Dict[int, float] = {}
    for name_hash, toks in zip(name_hashes, token_lists):
        terms = [f"{name_hash}_{t}" if prepend_feature_name else t for t in toks]
        tf = shared_tf if shared else {}
        for t in terms:
            i = hashing_tf_index(t, num_features)
            tf[i] = 1.0 if binary_freq else tf.get(i, 0.0) + 1.0
        if not shared:
            blocks.append(sorted(tf.items()))
    if shared:
        return sorted(shared_tf.items())
    # separate spaces: RichVector.combine with running offset numFeatures per input
    out: Entries = []
    for j, b in enumerate(blocks):
        out.extend((i + j * num_features, v) for i, v in b)
    return out


def hash_metadata(names: Sequence[str], type_name: str, num_features: int, shared: bool) -> List[ColumnMeta]:
    """HashingFun.makeVectorColumnMetadata :210-230"""
    if shared:
        return [ColumnMeta(tuple(names), tuple(type_name for _ in names)) for _ in range(num_features)]
    return [feature_column(nm, type_name) for nm in names for _ in range(num_features)]


# ----------------------------------------------------------------------------------------------------------------
# SmartTextVectorizer
# ----------------------------------------------------------------------------------------------------------------
PIVOT, HASH, IGNORE = "Pivot", "Hash", "Ignore"


def _addition_helper(total: Dict, to_add: Dict, max_card: int) -> Dict:
    """TextStats.additionHelper :231-235 (order dependent once a map exceeds maxCardinality)"""
    if len(total) > max_card:
        return total
    if len(to_add) > max_card:
        return to_add
    out = dict(total)
    for k, v in to_add.items():
        out[k] = out.get(k, 0) + v
    return out


def text_stats(column: Sequence[Optional[str]], clean_text: bool, max_card: int,
               tokenize_for_lengths: bool = False) -> Tuple[Dict[str, int], Dict[int, int]]:
    """Row-order left fold of TextStats (:79-91, :207-299)."""
    values: Dict[str, int] = {}
    lengths: Dict[int, int] = {}
    for cell in column:
        if cell is None:
            continue
        cleaned = clean_text_fn(cell, clean_text)
        if tokenize_for_lengths:
            lm: Dict[int, int] = {}
            for t in tokenize(cell):
                lm = _addition_helper(lm, {len(t): 1}, max_card)
        else:
            lm = {java_length(cleaned): 1}
        values = _addition_helper(values, {cleaned: 1}, max_card)
        lengths = _addition_helper(lengths, lm, max_card)
    return values, lengths


def java_length(s: str) -> int:
    """String.length = number of UTF-16 code units"""
    return len(s.encode("utf-16-le", "surrogatepass")) // 2


def length_std_dev(lengths: Dict[int, int]) -> float:
    size = sum(lengths.values())
    if size == 0:
        return float("nan")
    mean = sum(k * v for k, v in lengths.items()) / size
    var = sum(v * (k - mean) * (k - mean) for k, v in lengths.items()) / size
    return math.sqrt(var)


def smart_text_fit(columns: Sequence[Sequence[Optional[str]]], max_cardinality: int = 1000, top_k: int = 20,
                   min_support: int = 10, clean_text: bool = True, coverage_pct: float = 0.90,
                   min_length_std_dev: float = 0.0, tokenize_for_lengths: bool = False
                   ) -> Tuple[List[str], List[List[str]]]:
    """SmartTextVectorizer.fitFn :79-124 -> (vectorizationMethods, topValues)"""
    if not columns or len(columns[0]) == 0:
        raise ValueError("requirement failed: Input dataset cannot be empty")
    methods: List[str] = []
    tops: List[List[str]] = []
    for col in columns:
        values, lengths = text_stats(col, clean_text, max_cardinality, tokenize_for_lengths)
        total = sum(values.values())
        filtered = {k: c for k, c in values.items() if c >= min_support}
        sorted_counts = sorted(filtered.values(), reverse=True)
        cum = list(np.cumsum(sorted_counts)) if sorted_counts else []
        k = min(top_k, len(cum)) - 1
        if 0 <= k < len(cum):
            coverage = (cum[k] * 1.0 / total) if total != 0 else float("nan")
        else:
            coverage = (0.0 / total) if total != 0 else float("nan")
        card = len(values)
        if card > max_cardinality and card > top_k and coverage >= coverage_pct:
            m = PIVOT
        elif card <= max_cardinality:
            m = PIVOT
        elif length_std_dev(lengths) < min_length_std_dev:
            m = IGNORE
        else:
            m = HASH
        methods.append(m)
        tops.append(top_values(filtered, top_k, min_support))
    return methods, tops


def smart_text_transform(columns: Sequence[Sequence[Optional[str]]], names: Sequence[str], methods: Sequence[str],
                         tops: Sequence[Sequence[str]], num_features: int = 512, clean_text: bool = True,
                         track_nulls: bool = True, track_text_len: bool = False, prepend_feature_name: bool = True,
                         binary_freq: bool = False, hash_space_strategy: str = "auto", to_lowercase: bool = True,
                         min_token_length: int = 1) -> sp.csr_matrix:
    """SmartTextVectorizerModel.transformFn :333-356: [pivot | hash | len | null] blocks."""
    n = len(columns[0]) if columns else 0
    piv = [j for j, m in enumerate(methods) if m == PIVOT]
    hsh = [j for j, m in enumerate(methods) if m == HASH]
    ign = [j for j, m in enumerate(methods) if m == IGNORE]
    non_pivot_names = [names[j] for j, m in enumerate(methods) if m != PIVOT]  # getTextTransientFeatures
    shared = is_shared_hash_space(num_features, len(columns), hash_space_strategy)
    piv_tops = [tops[j] for j in piv]
    piv_w = pivot_width(piv_tops, track_nulls) if piv else 0
    hash_w = 0 if not hsh else (num_features if shared else num_features * len(hsh))
    txt = hsh + ign
    len_w = len(txt) if track_text_len else 0
    null_w = len(txt) if track_nulls else 0
    rows: List[Entries] = []
    for r in range(n):
        ent: Entries = []
        if piv:
            ent.extend(reindex_entries([pivot_row_entries(columns[j][r], tops[j], clean_text, track_nulls)
                                        for j in piv]))
        tok_h = [tokenize(columns[j][r], to_lowercase, min_token_length) for j in hsh]
        tok_i = [tokenize(columns[j][r], to_lowercase, min_token_length) for j in ign]
        off = piv_w
        ent.extend((i + off, v) for i, v in hash_token_lists(tok_h, non_pivot_names, num_features, shared,
                                                             prepend_feature_name, binary_freq))
        off += hash_w
        if track_text_len:
            ent.extend((off + k, float(sum(java_length(t) for t in toks))) for k, toks in enumerate(tok_h + tok_i))
            off += len_w
        if track_nulls:
            ent.extend((off + k, 1.0 if not toks else 0.0) for k, toks in enumerate(tok_h + tok_i))
        rows.append(ent)
    return rows_to_csr(rows, piv_w + hash_w + len_w + null_w)


def smart_text_metadata(names: Sequence[str], type_name: str, methods: Sequence[str],
                        tops: Sequence[Sequence[str]], num_features: int = 512, track_nulls: bool = True,
                        track_text_len: bool = False, hash_space_strategy: str = "auto") -> List[ColumnMeta]:
    """SmartTextVectorizer.makeVectorMetadata :154-186"""
    piv = [j for j, m in enumerate(methods) if m == PIVOT]
    hsh = [j for j, m in enumerate(methods) if m == HASH]
    ign = [j for j, m in enumerate(methods) if m == IGNORE]
    cols: List[ColumnMeta] = []
    if piv:
        cols += pivot_metadata([names[j] for j in piv], type_name, [tops[j] for j in piv], track_nulls)
    txt = [names[j] for j in hsh + ign]
    if txt:
        shared = is_shared_hash_space(num_features, len(names), hash_space_strategy)
        cols += hash_metadata([names[j] for j in hsh], type_name, num_features, shared) if hsh else []
        if track_text_len:
            cols += [ColumnMeta((nm,), (type_name,), grouping=nm, descriptor=TEXT_LEN_STRING) for nm in txt]
        cols += [feature_column(nm, type_name, is_null=True) for nm in txt]
    return reindex(cols)


# ----------------------------------------------------------------------------------------------------------------
# VectorsCombiner
# ----------------------------------------------------------------------------------------------------------------
def combine(blocks: Sequence, metas: Sequence[Sequence[ColumnMeta]]) -> Tuple[sp.csr_matrix, List[ColumnMeta]]:
    """VectorsCombinerModel.transformFn + updateMetadata: concat in input order with a running offset."""
    mats = [sp.csr_matrix(b) for b in blocks]
    for m, meta in zip(mats, metas):
        if m.shape[1] != len(meta):
            raise ValueError(f"vector width {m.shape[1]} does not match its metadata ({len(meta)} columns)")
    out = sp.hstack(mats, format="csr") if mats else sp.csr_matrix((0, 0))
    out.eliminate_zeros()
    return out, flatten(metas)
