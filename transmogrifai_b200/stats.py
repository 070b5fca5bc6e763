"""Statistics sidecar of an OPVector column: everything SanityChecker needs about the vectorized columns, produced
by the fused vectorizer scans (one read of the raw columns) instead of five passes over the materialised vector
(SanityChecker.scala:378, :407, :465-468, :265-272).  Sidecars of the vectors that VectorsCombiner concatenates
simply concatenate."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np


@dataclass
class LabelInfo:
    name: str
    n: int
    stats: np.ndarray                      # [5] n, mean, variance (n-1), min, max
    classes: Optional[np.ndarray]          # sorted distinct values, None if more than 255 of them
    class_counts: Optional[np.ndarray]     # rows per class

    def same_as(self, other: "LabelInfo") -> bool:
        return self.name == other.name and self.n == other.n


@dataclass
class VectorStats:
    label: LabelInfo
    mean: np.ndarray                       # [w] Statistics.colStats of every output column
    variance: np.ndarray
    min: np.ndarray
    max: np.ndarray
    cov_label: np.ndarray                  # [w] sum (x - mean_x)(y - mean_y)
    class_sums: Optional[np.ndarray]       # [n_classes, w] per-label-class column sums (NaN where not tracked)

    @property
    def width(self) -> int:
        return len(self.mean)

    @staticmethod
    def concat(parts: Sequence["VectorStats"]) -> "VectorStats":
        first = parts[0]
        for p in parts[1:]:
            if not p.label.same_as(first.label):
                raise ValueError("statistics sidecars were computed against different labels")
        cs = None
        if all(p.class_sums is not None for p in parts):
            cs = np.concatenate([p.class_sums for p in parts], axis=1)
        cat = lambda k: np.concatenate([getattr(p, k) for p in parts])  # noqa: E731
        return VectorStats(first.label, cat("mean"), cat("variance"), cat("min"), cat("max"), cat("cov_label"), cs)

    def select(self, idx: List[int]) -> "VectorStats":
        idx = np.asarray(idx, dtype=np.int64)
        if len(idx) == self.width and bool((idx == np.arange(self.width)).all()):
            return self
        cs = None if self.class_sums is None else self.class_sums[:, idx]
        return VectorStats(self.label, self.mean[idx], self.variance[idx], self.min[idx], self.max[idx],
                           self.cov_label[idx], cs)
