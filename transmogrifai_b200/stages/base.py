"""Stage base classes: the host-side mirror of the reference's OpPipelineStage / SequenceEstimator / SequenceModel /
BinaryEstimator API (setInput, getOutput, fit, transform, getMetadata), kept to what the hot path's stages need.

Reference: features/src/main/scala/com/salesforce/op/stages/OpPipelineStages.scala:55-200,
features/src/main/scala/com/salesforce/op/stages/base/sequence/SequenceEstimator.scala:57-106,
features/src/main/scala/com/salesforce/op/stages/base/binary/BinaryEstimator.scala:61-114.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

from .. import features as F
from ..columnar import ColumnarBatch, VectorColumn
from ..metadata import OpVectorMetadata


class OpPipelineStage:
    """setInput / getOutput / getMetadata plumbing shared by estimators, models and transformers."""

    input_types: Sequence[str] = ()      # accepted input feature types (empty = any)
    output_type: str = F.OPVECTOR

    def __init__(self, operationName: str, uid: Optional[str] = None):  # noqa: N803
        self.operationName = operationName
        self.uid = uid or F.make_uid(type(self).__name__)
        self._inputs: List[F.Feature] = []
        self._metadata: Optional[OpVectorMetadata] = None
        self._output: Optional[F.Feature] = None

    @property
    def stageName(self) -> str:  # noqa: N802
        return f"{self.operationName}_{self.uid}"

    # -- inputs / outputs ------------------------------------------------------------------------------------------
    def setInput(self, *features: F.Feature):  # noqa: N802
        feats = list(features[0]) if len(features) == 1 and isinstance(features[0], (list, tuple)) else list(features)
        if not feats:
            raise ValueError("requirement failed: Inputs cannot be empty")
        for f in feats:
            if self.input_types and f.type_name not in self.input_types:
                raise ValueError(f"requirement failed: {type(self).__name__} does not accept input feature "
                                 f"'{f.name}' of type {f.type_name}")
        self._inputs = feats
        self._output = None
        self.onSetInput()
        return self

    def onSetInput(self) -> None:  # noqa: N802
        pass

    def getInputFeatures(self) -> List[F.Feature]:  # noqa: N802
        return list(self._inputs)

    @property
    def inN(self) -> List[F.Feature]:  # noqa: N802
        if not self._inputs:
            raise KeyError("NoSuchElementException: inputs are not set")   # RealVectorizerTest.scala:77-79
        return self._inputs

    def getOutputFeatureName(self) -> str:  # noqa: N802
        names = "-".join(f.name for f in self.inN)
        return f"{names}_{len(self.inN)}-stagesApplied_{self.output_type}_{self.uid[-12:]}"

    def outputIsResponse(self) -> bool:  # noqa: N802
        return all(f.is_response for f in self.inN)

    def getOutput(self) -> F.Feature:  # noqa: N802
        if self._output is None:
            self._output = F.Feature(self.getOutputFeatureName(), self.output_type, self.outputIsResponse(), self,
                                     tuple(self.inN))
        return self._output

    # -- metadata --------------------------------------------------------------------------------------------------
    def setMetadata(self, m: OpVectorMetadata):  # noqa: N802
        self._metadata = m
        return self

    def getMetadata(self) -> OpVectorMetadata:  # noqa: N802
        self.onGetMetadata()
        if self._metadata is None:
            raise ValueError(f"stage {self.stageName} has no metadata yet (fit it first)")
        return self._metadata

    def onGetMetadata(self) -> None:  # noqa: N802
        pass

    def history(self) -> dict:
        """Transmogrifier.inputFeaturesToHistory: origin features + stages applied, per input feature."""
        return {f.name: (self.stageName,) for f in self.inN}


class OpTransformer(OpPipelineStage):
    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        raise NotImplementedError

    def transform(self, batch: ColumnarBatch) -> ColumnarBatch:
        vec = self.transformColumn(batch)
        vec.producer = self
        return batch.with_vector(vec)


class OpModel(OpTransformer):
    """A fitted stage: shares uid, inputs and output feature with the estimator that produced it."""

    def __init__(self, parent: "OpEstimator"):
        super().__init__(parent.operationName, parent.uid)
        self.parent = parent
        self._inputs = list(parent._inputs)
        self._output = parent.getOutput()
        self._metadata = parent._metadata


class OpEstimator(OpPipelineStage):
    def fit(self, batch: ColumnarBatch) -> OpModel:
        raise NotImplementedError
