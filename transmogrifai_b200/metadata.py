"""OpVectorColumnMetadata / OpVectorMetadata: the output contract of every vectorizer (column lineage, grouping,
indicator values) that SanityChecker and ModelInsights read.

Reference: features/src/main/scala/com/salesforce/op/utils/spark/OpVectorColumnMetadata.scala:67-171,
features/src/main/scala/com/salesforce/op/utils/spark/OpVectorMetadata.scala:51-277.
"""
from __future__ import annotations

from dataclasses import dataclass, replace
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

from .features import MULTIPICKLIST, TEXT_TYPES

NullString = "NullIndicatorValue"
TextLenString = "TextLenValue"
OtherString = "OTHER"


@dataclass(frozen=True)
class OpVectorColumnMetadata:
    parentFeatureName: Tuple[str, ...]
    parentFeatureType: Tuple[str, ...]
    grouping: Optional[str] = None
    indicatorValue: Optional[str] = None
    descriptorValue: Optional[str] = None
    index: int = 0

    def __post_init__(self):
        if not self.parentFeatureName:
            raise ValueError("requirement failed: must provide parent feature name")
        if len(self.parentFeatureName) != len(self.parentFeatureType):
            raise ValueError("requirement failed: must provide both type and name for every parent feature")
        if self.indicatorValue is not None and self.descriptorValue is not None:
            raise ValueError("requirement failed: cannot have both indicatorValue and descriptorValue")

    def makeColName(self) -> str:  # noqa: N802
        s = "_".join(self.parentFeatureName)
        for part in (self.grouping, self.indicatorValue, self.descriptorValue):
            if part is not None:
                s += "_" + part
        return f"{s}_{self.index}"

    def featureGroup(self) -> Optional[str]:  # noqa: N802
        return None if self.grouping is None else "_".join(self.parentFeatureName) + "_" + self.grouping

    def parentNamesWithMapKeys(self) -> Tuple[str, ...]:  # noqa: N802  (no OPMap parents on this path)
        return self.parentFeatureName

    def hasParentOfType(self, *types: str) -> bool:  # noqa: N802
        return any(t in types for t in self.parentFeatureType)

    @property
    def isNullIndicator(self) -> bool:  # noqa: N802
        return self.indicatorValue == NullString

    @property
    def isOtherIndicator(self) -> bool:  # noqa: N802
        return self.indicatorValue == OtherString

    @property
    def isTextSharedHash(self) -> bool:  # noqa: N802  (DerivedFeatureFilterUtils.scala:410-414)
        return self.hasParentOfType(*TEXT_TYPES, "TextMap", "TextAreaMap") and self.grouping is None \
            and self.indicatorValue is None

    @property
    def isMultiPickList(self) -> bool:  # noqa: N802
        return self.hasParentOfType(MULTIPICKLIST)


def feature_column(name: str, type_name: str, is_null: bool = False) -> OpVectorColumnMetadata:
    """TransientFeature.toColumnMetaData(isNull) -- features/.../TransientFeature.scala:118-124"""
    return OpVectorColumnMetadata((name,), (type_name,), grouping=name if is_null else None,
                                  indicatorValue=NullString if is_null else None)


class OpVectorMetadata:
    """Metadata of an OPVector: one OpVectorColumnMetadata per column, re-indexed 0..size-1 (OpVectorMetadata.scala:62).

    Hashed text blocks repeat one column description hundreds of thousands of times, so the columns are held as runs
    (description, count) and only expanded on demand; everything SanityChecker needs is computed per run."""

    def __init__(self, name: str, columns: Optional[Iterable[OpVectorColumnMetadata]] = None,
                 history: Optional[Dict[str, Sequence[str]]] = None,
                 runs: Optional[Sequence[Tuple[OpVectorColumnMetadata, int]]] = None):
        self.name = name
        self.history = {k: tuple(v) for k, v in (history or {}).items()}
        merged: List[List] = []
        source = runs if runs is not None else ((c, 1) for c in (columns or ()))
        for c, k in source:
            if k <= 0:
                continue
            if merged and merged[-1][0] is c:
                merged[-1][1] += k
            else:
                merged.append([c, k])
        self.runs: List[Tuple[OpVectorColumnMetadata, int]] = [(c, k) for c, k in merged]
        self.run_starts = np.concatenate([[0], np.cumsum([k for _, k in self.runs])]).astype(np.int64)
        self._columns: Optional[List[OpVectorColumnMetadata]] = None
        self._names: Optional[List[str]] = None

    @property
    def size(self) -> int:
        return int(self.run_starts[-1])

    @property
    def columns(self) -> List[OpVectorColumnMetadata]:
        if self._columns is None:
            out: List[OpVectorColumnMetadata] = []
            for (c, k), s0 in zip(self.runs, self.run_starts[:-1]):
                s0 = int(s0)
                out.extend(c if (k == 1 and c.index == s0) else replace(c, index=s0 + i) for i in range(k))
            self._columns = out
        return self._columns

    def column(self, i: int) -> OpVectorColumnMetadata:
        r = int(np.searchsorted(self.run_starts, i, side="right")) - 1
        return replace(self.runs[r][0], index=int(i))

    def columnNames(self) -> List[str]:  # noqa: N802
        if self._names is None:
            out: List[str] = []
            for (c, k), s0 in zip(self.runs, self.run_starts[:-1]):
                stem = c.makeColName().rsplit("_", 1)[0]
                out.extend(f"{stem}_{int(s0) + i}" for i in range(k))
            self._names = out
        return self._names

    def names_of(self, indices: Sequence[int]) -> List[str]:
        if self._names is not None:
            return [self._names[int(i)] for i in indices]
        out = []
        for i in indices:
            r = int(np.searchsorted(self.run_starts, i, side="right")) - 1
            out.append(f"{self.runs[r][0].makeColName().rsplit('_', 1)[0]}_{int(i)}")
        return out

    def select(self, name: str, keep: Sequence[int]) -> "OpVectorMetadata":
        """Metadata of the vector restricted to the columns `keep` (ascending), runs preserved."""
        keep = np.asarray(keep, dtype=np.int64)
        lo = np.searchsorted(keep, self.run_starts[:-1], side="left")
        hi = np.searchsorted(keep, self.run_starts[1:], side="left")
        return OpVectorMetadata(name, history=self.history,
                                runs=[(c, int(h - l)) for (c, _), l, h in zip(self.runs, lo, hi) if h > l])

    def __eq__(self, other) -> bool:
        return isinstance(other, OpVectorMetadata) and self.name == other.name and self.size == other.size and \
            [(replace(c, index=0), k) for c, k in self.runs] == [(replace(c, index=0), k) for c, k in other.runs]

    @staticmethod
    def flatten(name: str, vectors: Sequence["OpVectorMetadata"]) -> "OpVectorMetadata":
        """OpVectorMetadata.scala:270-275"""
        runs: List[Tuple[OpVectorColumnMetadata, int]] = []
        hist: Dict[str, Tuple[str, ...]] = {}
        for v in vectors:
            runs.extend(v.runs)
            hist.update(v.history)
        return OpVectorMetadata(name, history=hist, runs=runs)
