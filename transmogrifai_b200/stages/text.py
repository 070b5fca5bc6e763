"""Default vectorizers of the string-typed features (Transmogrifier.scala:275-317): PickList -> pivot,
MultiPickList -> OpSetVectorizer, Text / TextArea -> SmartTextVectorizer."""
from __future__ import annotations

from typing import List

from .. import features as F
from .pivot import OpTextPivotVectorizer


def vectorizer_for_text_type(type_name: str, feats: List[F.Feature], d) -> F.Feature:
    if type_name == F.PICKLIST:
        return OpTextPivotVectorizer().setInput(*feats).setTopK(d.TopK).setMinSupport(d.MinSupport) \
            .setCleanText(d.CleanText).setTrackNulls(d.TrackNulls).getOutput()
    if type_name == F.MULTIPICKLIST:
        from .sets import OpSetVectorizer
        return OpSetVectorizer().setInput(*feats).setTopK(d.TopK).setMinSupport(d.MinSupport) \
            .setCleanText(d.CleanText).setTrackNulls(d.TrackNulls).getOutput()
    if type_name in (F.TEXT, F.TEXTAREA):
        from .smart_text import SmartTextVectorizer
        return SmartTextVectorizer().setInput(*feats).setMaxCardinality(d.MaxCategoricalCardinality) \
            .setTrackNulls(d.TrackNulls).setNumFeatures(d.DefaultNumOfFeatures) \
            .setHashSpaceStrategy(d.HashSpaceStrategy).setMinTokenLength(d.MinTokenLength) \
            .setToLowercase(d.ToLowercase).setPrependFeatureName(d.PrependFeatureName).setTopK(d.TopK) \
            .setMinSupport(d.MinSupport).setCleanText(d.CleanText).getOutput()
    raise NotImplementedError(f"no default vectorizer for feature type {type_name} on this path")
