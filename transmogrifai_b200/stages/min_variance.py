"""MinVarianceFilter on the GPU: the label-free subset of SanityChecker -- column statistics of the feature vector and
the variance rule -- and its model's column gather.

Reference: core/src/main/scala/com/salesforce/op/stages/impl/preparators/MinVarianceFilter.scala:59-159 (fitFn :72-139,
defaults :155-158), DerivedFeatureFilterUtils.scala:95-214 (makeColumnStatistics), :234-281 (getFeaturesToDrop with
every threshold but minVariance at its neutral default), :350-365 (variance <= minVariance), :291-305 (removeFeatures);
summary MinVarianceFilterMetadata.scala:47-68; DSL RichVectorFeature.scala:153-163 (filterMinVariance).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import List, Optional

import numpy as np
import torch

from .. import _lib as L
from .. import dist as D
from .. import runtime as R
from ..columnar import ColumnarBatch, VectorColumn
from ..metadata import OpVectorMetadata
from .base import OpEstimator, OpModel


@dataclass
class MinVarianceSummary:
    """MinVarianceFilterMetadata.scala:47-52 (featuresStatistics = SummaryStatistics(colStats, sample = 1.0))."""
    meta: OpVectorMetadata
    count: int
    mean: np.ndarray
    variance: np.ndarray
    min: np.ndarray
    max: np.ndarray
    dropped: List[str]
    sampleFraction: float = 1.0

    @property
    def names(self) -> List[str]:
        return self.meta.columnNames()


def column_stats(batch: ColumnarBatch, vec: VectorColumn):
    """Statistics.colStats of the vector over the rows of all ranks -> (n, mean, variance, min, max).  A vector that
    comes with the statistics sidecar of the fused vectorizer scans is not read again; any other vector is scanned by
    the dense statistics kernel (sparse parts densified, as Spark's colStats sees them)."""
    n = int(D.allreduce_sum(np.array([batch.n], dtype=np.int64))[0])
    st = vec.stats
    if st is not None and st.label.n == n:
        return n, st.mean, st.variance, st.min, st.max
    R.require_cuda()
    from .generic_stats import _dense_block
    x = _dense_block(vec)
    rows, d = x.shape
    y = torch.zeros(max(rows, 2), dtype=torch.float64, device="cuda")          # the kernel's label slot, unused here
    nb = C.c_size_t()
    L.call("tmog_dense_stats_scratch_bytes", rows, d, C.byref(nb))
    scratch = R.scratch("dense_stats", nb.value)
    shifts = torch.zeros(d + 1, dtype=torch.float64, device="cuda")
    partials = torch.zeros((d + 1) * L.REAL_NSTAT, dtype=torch.float64, device="cuda")
    L.call("tmog_dense_pilot_shift_dev", x.data_ptr(), x.stride(0), rows, d, y.data_ptr(), shifts.data_ptr(),
           R.stream())
    if D.world()[1] > 1:                                   # partials only merge under common shifts
        torch.distributed.broadcast(shifts, src=0)
    L.call("tmog_dense_stats_dev", x.data_ptr(), x.stride(0), rows, d, y.data_ptr(), None, 0, None,
           shifts.data_ptr(), scratch.data_ptr(), scratch.numel(), partials.data_ptr(), None, R.stream())
    p, _ = D.gather_fit_results(partials)
    s = shifts.cpu().numpy()
    mean, var, mn, mx, cov = (np.zeros(d) for _ in range(5))
    lab = np.zeros(5)
    L.call("tmog_real_finish", p.ctypes.data, s.ctypes.data, d, 1, 0.0, 0, None, mean.ctypes.data, var.ctypes.data,
           mn.ctypes.data, mx.ctypes.data, cov.ctypes.data, lab.ctypes.data, None)
    return n, mean, var, mn, mx


class MinVarianceFilter(OpEstimator):
    """UnaryEstimator[OPVector, OPVector]: setInput(featureVector)."""

    def __init__(self, uid: Optional[str] = None):
        super().__init__("MinVarianceFilter", uid)
        self.minVariance = 1e-5            # MinVarianceFilter.MinVariance
        self.removeBadFeatures = False     # MinVarianceFilter.RemoveBadFeatures
        self.summary: Optional[MinVarianceSummary] = None

    def setMinVariance(self, v: float):  # noqa: N802
        self.minVariance = float(v)
        return self

    def setRemoveBadFeatures(self, v: bool):  # noqa: N802
        self.removeBadFeatures = bool(v)
        return self

    def fit(self, batch: ColumnarBatch) -> "MinVarianceFilterModel":
        vec = batch.vectors[self.inN[0].name]
        meta = vec.metadata
        n, mean, var, mn, mx = column_stats(batch, vec)
        if n <= 0:
            raise ValueError("requirement failed: Sample size cannot be zero")
        d = vec.width
        if d <= 0:
            raise ValueError("requirement failed: Feature vector passed in is empty, check your vectorizers")
        if d != meta.size:
            raise ValueError(f"requirement failed: Number of columns in vector metadata ({meta.size}) did not match "
                             f"number of columns in data({d}), check your vectorizers")
        # getFeaturesToDrop(stats, minVariance): without a label no other rule can fire
        drop = (var <= self.minVariance) if self.removeBadFeatures else np.zeros(d, dtype=bool)
        drop_idx = np.nonzero(drop)[0]
        keep = np.nonzero(~drop)[0]
        self._metadata = meta.select(self.getOutputFeatureName(), keep)
        self.summary = MinVarianceSummary(meta=meta, count=n, mean=mean, variance=var, min=mn, max=mx,
                                          dropped=meta.names_of(drop_idx))
        if len(keep) == 0:
            raise ValueError("requirement failed: The minimum variance filter has dropped all of your features, check "
                             "your input data or your threshold")
        return MinVarianceFilterModel(self, keep, self.removeBadFeatures)


class MinVarianceFilterModel(OpModel):
    def __init__(self, parent: MinVarianceFilter, indicesToKeep, removeBadFeatures: bool):  # noqa: N803
        super().__init__(parent)
        self._keep = np.asarray(indicesToKeep, dtype=np.int64)
        self.removeBadFeatures = removeBadFeatures

    @property
    def indicesToKeep(self) -> List[int]:  # noqa: N802
        return self._keep.tolist()

    def transformColumn(self, batch: ColumnarBatch) -> VectorColumn:  # noqa: N802
        vec = batch.vectors[self.inN[0].name]
        name = self.getOutput().name
        if not self.removeBadFeatures:
            return VectorColumn(name, batch.n, self.getMetadata(), vec.parts, vec.stats)
        from .generic_stats import select_columns
        parts = select_columns(vec, self._keep)
        stats = None if vec.stats is None else vec.stats.select(self._keep)
        return VectorColumn(name, batch.n, self.getMetadata(), parts, stats)
