"""DSL entry points: transmogrify() and sanityCheck().

Reference: core/src/main/scala/com/salesforce/op/dsl/RichFeaturesCollection.scala:52,69-70 (transmogrify + combine),
core/src/main/scala/com/salesforce/op/stages/impl/feature/Transmogrifier.scala:102-352 (type dispatch, defaults
:52-88), core/src/main/scala/com/salesforce/op/dsl/RichNumericFeature.scala:469-514 (sanityCheck).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

from . import features as F
from .columnar import ColumnarBatch
from .stages.base import OpEstimator, OpPipelineStage
from .stages.combiner import VectorsCombiner
from .stages.numeric import BinaryVectorizer, RealNNVectorizer, RealVectorizer
from .stages.sanity import SanityChecker


class TransmogrifierDefaults:
    """Transmogrifier.scala:52-88"""
    TopK = 20
    MinSupport = 10
    DefaultNumOfFeatures = 512
    MaxNumOfFeatures = 1 << 17
    FillValue = 0
    BinaryFillValue = False
    PrependFeatureName = True
    HashSpaceStrategy = "Auto"
    CleanText = True
    BinaryFreq = False
    FillWithMean = True
    FillWithMode = True
    TrackNulls = True
    TrackTextLen = False
    MaxCategoricalCardinality = 30
    MinTokenLength = 1
    ToLowercase = True


def _vectorizer_for(type_name: str, feats: List[F.Feature], d=TransmogrifierDefaults) -> F.Feature:
    if type_name == F.BINARY:
        return BinaryVectorizer().setInput(*feats).setTrackNulls(d.TrackNulls).setFillValue(d.BinaryFillValue) \
            .getOutput()
    if type_name == F.REAL:
        st = RealVectorizer().setInput(*feats).setTrackNulls(d.TrackNulls)
        st = st.setFillWithMean() if d.FillWithMean else st.setFillWithConstant(d.FillValue)
        return st.getOutput()
    if type_name == F.REALNN:
        return RealNNVectorizer().setInput(*feats).getOutput()
    if type_name == F.INTEGRAL:                      # Transmogrifier.scala:258-262
        from .stages.integral import IntegralVectorizer
        st = IntegralVectorizer().setInput(*feats).setTrackNulls(d.TrackNulls)
        st = st.setFillWithMode() if d.FillWithMode else st.setFillWithConstant(d.FillValue)
        return st.getOutput()
    if type_name in (F.PICKLIST, F.MULTIPICKLIST, F.TEXT, F.TEXTAREA):
        from .stages.text import vectorizer_for_text_type
        return vectorizer_for_text_type(type_name, feats, d)
    raise NotImplementedError(f"transmogrify(): no default vectorizer for feature type {type_name} on this path")


def transmogrify(features: Sequence[F.Feature], label: Optional[F.Feature] = None,
                 defaults=TransmogrifierDefaults) -> F.Feature:
    """Seq(features).transmogrify(): one vectorizer per feature TYPE (types visited in type-name order, members in
    input order), then VectorsCombiner."""
    if not features:
        raise ValueError("requirement failed: transmogrify() needs at least one feature")
    groups: Dict[str, List[F.Feature]] = {}
    for f in features:
        groups.setdefault(f.type_name, []).append(f)
    vectors = [f for tp in sorted(groups) for f in ([_vectorizer_for(tp, groups[tp], defaults)] if tp != F.OPVECTOR
                                                     else groups[tp])]
    return VectorsCombiner().setInput(*vectors).getOutput()


def sanity_check(label: F.Feature, feature_vector: F.Feature, checkSample: float = 1.0,  # noqa: N803
                 sampleSeed: int = 0, sampleLowerLimit: int = 1000, sampleUpperLimit: int = 1000000,
                 maxCorrelation: float = 0.95, minCorrelation: float = 0.0, maxCramersV: float = 0.95,
                 minVariance: float = 1e-5, removeBadFeatures: bool = False, removeFeatureGroup: bool = True,
                 protectTextSharedHash: bool = False, maxRuleConfidence: float = 1.0,
                 minRequiredRuleSupport: float = 1.0, featureFeatureCorrLevel: str = "Computed",
                 categoricalLabel: Optional[bool] = None) -> F.Feature:
    """label.sanityCheck(featureVector, ...) -- RichNumericFeature.scala:469-514"""
    st = SanityChecker().setCheckSample(checkSample).setSampleSeed(sampleSeed) \
        .setSampleLowerLimit(sampleLowerLimit).setSampleUpperLimit(sampleUpperLimit) \
        .setMaxCorrelation(maxCorrelation).setMinCorrelation(minCorrelation).setMaxCramersV(maxCramersV) \
        .setMinVariance(minVariance).setRemoveBadFeatures(removeBadFeatures) \
        .setRemoveFeatureGroup(removeFeatureGroup).setProtectTextSharedHash(protectTextSharedHash) \
        .setMaxRuleConfidence(maxRuleConfidence).setMinRequiredRuleSupport(minRequiredRuleSupport) \
        .setFeatureFeatureCorrLevel(featureFeatureCorrLevel)
    if categoricalLabel is not None:
        st.setCategoricalLabel(categoricalLabel)
    return st.setInput(label, feature_vector).getOutput()


def stages_of(result: F.Feature) -> List[List[OpPipelineStage]]:
    """Layers of the stage DAG by distance from the result, farthest first (FitStagesUtil.computeDAG :173-198)."""
    dist: Dict[int, int] = {}
    stage_of: Dict[int, OpPipelineStage] = {}

    def walk(f: F.Feature, depth: int):
        st = f.origin_stage
        if st is None:
            return
        key = id(st)
        stage_of[key] = st
        dist[key] = max(dist.get(key, 0), depth)
        for p in f.parents:
            walk(p, depth + 1)

    walk(result, 0)
    layers: Dict[int, List[OpPipelineStage]] = {}
    for key, dd in dist.items():
        layers.setdefault(dd, []).append(stage_of[key])
    return [layers[k] for k in sorted(layers, reverse=True)]


def fit_and_transform(batch: ColumnarBatch, result: F.Feature):
    """Minimal stand-in for OpWorkflow.train()'s fitAndTransformDAG (FitStagesUtil.scala:212-290): fit every
    estimator of a layer, then apply the layer's transformers.  Returns (batch with the new columns, fitted stages)."""
    fitted = []
    for layer in stages_of(result):
        for st in layer:
            model = st.fit(batch) if isinstance(st, OpEstimator) else st
            batch = model.transform(batch)
            fitted.append(model)
    return batch, fitted
