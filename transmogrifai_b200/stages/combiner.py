"""VectorsCombiner: concatenate OPVectors and their metadata in input order.

Reference: core/src/main/scala/com/salesforce/op/stages/impl/feature/VectorsCombiner.scala:48-89,
features/src/main/scala/com/salesforce/op/utils/spark/RichVector.scala:120-134.  No arithmetic and no copy: the
combined vector is the ordered list of the inputs' dense / CSR parts; column offsets are implied by the order.
"""
from __future__ import annotations

from typing import Optional

from .. import features as F
from ..columnar import ColumnarBatch, VectorColumn
from ..metadata import OpVectorMetadata
from ..stats import VectorStats
from .base import OpEstimator, OpModel


class VectorsCombiner(OpEstimator):
    input_types = (F.OPVECTOR,)

    def __init__(self, uid: Optional[str] = None):
        super().__init__("combineVector", uid)

    def fit(self, batch: ColumnarBatch) -> "VectorsCombinerModel":
        vecs = [batch.vectors[f.name] for f in self.inN]
        key = tuple(id(v.metadata) for v in vecs)
        if getattr(self, "_flat_key", None) != key:      # same input metadata objects: same flattened metadata
            self._flat = OpVectorMetadata.flatten(self.getOutputFeatureName(), [v.metadata for v in vecs])
            self._flat_key, self._flat_src = key, [v.metadata for v in vecs]   # keep the sources alive (ids stay valid)
        self._metadata = self._flat
        return VectorsCombinerModel(self)


class VectorsCombinerModel(OpModel):
    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        vecs = [batch.vectors[f.name] for f in self.inN]
        meta = self.getMetadata()
        parts = [p for v in vecs for p in v.parts]
        stats = VectorStats.concat([v.stats for v in vecs]) if all(v.stats is not None for v in vecs) else None
        return VectorColumn(self.getOutput().name, batch.n, meta, parts, stats)
