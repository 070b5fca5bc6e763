"""Oracle: OpVectorColumnMetadata / OpVectorMetadata restatement (TEST INFRASTRUCTURE ONLY).

Follows
  * features/src/main/scala/com/salesforce/op/utils/spark/OpVectorColumnMetadata.scala:67-171
  * features/src/main/scala/com/salesforce/op/utils/spark/OpVectorMetadata.scala:51-277 (re-indexing :62, flatten :270-275)
  * features/src/main/scala/com/salesforce/op/features/TransientFeature.scala:118-139 (toColumnMetaData)
"""
from __future__ import annotations

from dataclasses import dataclass, replace
from typing import List, Optional, Sequence, Tuple

NULL_STRING = "NullIndicatorValue"   # OpVectorColumnMetadata.scala:169
TEXT_LEN_STRING = "TextLenValue"     # :170
OTHER_STRING = "OTHER"               # :171

TEXT_TYPES = ("Text", "TextArea", "TextMap", "TextAreaMap")


@dataclass(frozen=True)
class ColumnMeta:
    parent_names: Tuple[str, ...]
    parent_types: Tuple[str, ...]
    grouping: Optional[str] = None
    indicator: Optional[str] = None
    descriptor: Optional[str] = None
    index: int = 0

    def col_name(self) -> str:
        """makeColName, OpVectorColumnMetadata.scala:125-127"""
        s = "_".join(self.parent_names)
        if self.grouping is not None:
            s += "_" + self.grouping
        if self.indicator is not None:
            s += "_" + self.indicator
        if self.descriptor is not None:
            s += "_" + self.descriptor
        return f"{s}_{self.index}"

    def feature_group(self) -> Optional[str]:
        """featureGroup, :156-157"""
        if self.grouping is None:
            return None
        return "_".join(self.parent_names) + "_" + self.grouping

    def parent_names_with_map_keys(self) -> Tuple[str, ...]:
        """:146-149 -- no OPMap parents exist on this path."""
        return self.parent_names

    def has_parent_of_type(self, *types: str) -> bool:
        return any(t in types for t in self.parent_types)

    @property
    def is_null_indicator(self) -> bool:
        return self.indicator == NULL_STRING

    @property
    def is_other_indicator(self) -> bool:
        return self.indicator == OTHER_STRING


def feature_column(name: str, type_name: str, is_null: bool = False) -> ColumnMeta:
    """TransientFeature.toColumnMetaData(isNull), TransientFeature.scala:118-124"""
    return ColumnMeta((name,), (type_name,),
                      grouping=name if is_null else None,
                      indicator=NULL_STRING if is_null else None)


def reindex(cols: Sequence[ColumnMeta]) -> List[ColumnMeta]:
    """OpVectorMetadata re-indexes its columns on construction, OpVectorMetadata.scala:62"""
    return [replace(c, index=i) for i, c in enumerate(cols)]


def flatten(vectors: Sequence[Sequence[ColumnMeta]]) -> List[ColumnMeta]:
    """OpVectorMetadata.flatten, OpVectorMetadata.scala:270-275"""
    out: List[ColumnMeta] = []
    for v in vectors:
        out.extend(v)
    return reindex(out)
