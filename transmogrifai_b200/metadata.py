"""OpVectorColumnMetadata / OpVectorMetadata: the output contract of every vectorizer (column lineage, grouping,
indicator values) that SanityChecker and ModelInsights read.

Reference: features/src/main/scala/com/salesforce/op/utils/spark/OpVectorColumnMetadata.scala:67-171,
features/src/main/scala/com/salesforce/op/utils/spark/OpVectorMetadata.scala:51-277.
"""
from __future__ import annotations

from dataclasses import dataclass, replace
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

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


@dataclass
class OpVectorMetadata:
    name: str
    columns: List[OpVectorColumnMetadata]
    history: Dict[str, Tuple[str, ...]]

    def __init__(self, name: str, columns: Iterable[OpVectorColumnMetadata],
                 history: Optional[Dict[str, Sequence[str]]] = None):
        self.name = name
        self.columns = [c if c.index == i else replace(c, index=i) for i, c in enumerate(columns)]  # :62
        self.history = {k: tuple(v) for k, v in (history or {}).items()}

    @property
    def size(self) -> int:
        return len(self.columns)

    def columnNames(self) -> List[str]:  # noqa: N802
        return [c.makeColName() for c in self.columns]

    @staticmethod
    def flatten(name: str, vectors: Sequence["OpVectorMetadata"]) -> "OpVectorMetadata":
        """OpVectorMetadata.scala:270-275"""
        cols: List[OpVectorColumnMetadata] = []
        hist: Dict[str, Tuple[str, ...]] = {}
        for v in vectors:
            cols.extend(v.columns)
            hist.update(v.history)
        return OpVectorMetadata(name, cols, hist)
