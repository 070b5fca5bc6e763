"""Feature types and typed feature handles (host-side mirror of the reference's feature layer, just enough to
address columns and carry lineage through the vectorizer stages).

Reference: features/src/main/scala/com/salesforce/op/features/types/*.scala (type names),
features/src/main/scala/com/salesforce/op/features/FeatureBuilder.scala (asPredictor / asResponse).
"""
from __future__ import annotations

import itertools
from dataclasses import dataclass, field
from typing import Optional, Tuple

REAL, REALNN, BINARY, INTEGRAL = "Real", "RealNN", "Binary", "Integral"
DATE, DATETIME = "Date", "DateTime"                 # Integral subtypes (epoch milliseconds as int64)
INTEGRAL_TYPES = (INTEGRAL, DATE, DATETIME)
PICKLIST, MULTIPICKLIST = "PickList", "MultiPickList"
TEXT, TEXTAREA = "Text", "TextArea"
OPVECTOR = "OPVector"

NUMERIC_TYPES = (REAL, REALNN, BINARY, INTEGRAL, DATE, DATETIME)
TEXT_TYPES = (TEXT, TEXTAREA)
ALL_TYPES = (REAL, REALNN, BINARY, INTEGRAL, DATE, DATETIME, PICKLIST, MULTIPICKLIST, TEXT, TEXTAREA, OPVECTOR)

_uid_counter = itertools.count(1)


def make_uid(prefix: str) -> str:
    """UID[...] of the reference: ClassName_00000000000x"""
    return f"{prefix}_{next(_uid_counter):012x}"


@dataclass(frozen=True)
class Feature:
    name: str
    type_name: str
    is_response: bool = False
    origin_stage: Optional[object] = field(default=None, compare=False, repr=False)
    parents: Tuple["Feature", ...] = field(default=(), compare=False, repr=False)

    @property
    def isResponse(self) -> bool:  # noqa: N802 (reference spelling)
        return self.is_response

    @property
    def typeName(self) -> str:  # noqa: N802
        return self.type_name

    # DSL entry points (core/src/main/scala/com/salesforce/op/dsl/*), wired in dsl.py to avoid an import cycle
    def sanityCheck(self, feature_vector: "Feature", **kwargs) -> "Feature":  # noqa: N802
        from .dsl import sanity_check
        return sanity_check(self, feature_vector, **kwargs)

    def filterMinVariance(self, minVariance: float = 1e-5, removeBadFeatures: bool = False) -> "Feature":  # noqa: N802,N803
        """RichVectorFeature.filterMinVariance (core/src/main/scala/com/salesforce/op/dsl/RichVectorFeature.scala:153-163)"""
        from .stages.min_variance import MinVarianceFilter
        return MinVarianceFilter().setInput(self).setMinVariance(minVariance) \
            .setRemoveBadFeatures(removeBadFeatures).getOutput()


class _Builder:
    def __init__(self, name: str, type_name: str):
        self._name, self._type = name, type_name

    def asPredictor(self) -> Feature:  # noqa: N802
        return Feature(self._name, self._type, False)

    def asResponse(self) -> Feature:  # noqa: N802
        return Feature(self._name, self._type, True)


class FeatureBuilder:
    """FeatureBuilder.Real[Passenger].extract(...).asPredictor -- extraction itself is the columnar export."""

    @staticmethod
    def of(type_name: str, name: str) -> _Builder:
        if type_name not in ALL_TYPES:
            raise ValueError(f"unsupported feature type {type_name}")
        return _Builder(name, type_name)

    Real = staticmethod(lambda name: _Builder(name, REAL))
    RealNN = staticmethod(lambda name: _Builder(name, REALNN))
    Binary = staticmethod(lambda name: _Builder(name, BINARY))
    Integral = staticmethod(lambda name: _Builder(name, INTEGRAL))
    Date = staticmethod(lambda name: _Builder(name, DATE))
    DateTime = staticmethod(lambda name: _Builder(name, DATETIME))
    PickList = staticmethod(lambda name: _Builder(name, PICKLIST))
    MultiPickList = staticmethod(lambda name: _Builder(name, MULTIPICKLIST))
    Text = staticmethod(lambda name: _Builder(name, TEXT))
    TextArea = staticmethod(lambda name: _Builder(name, TEXTAREA))
