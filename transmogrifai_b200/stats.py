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

    def merge(self, other: "LabelInfo") -> "LabelInfo":
        """The label over the rows of both (disjoint row sets): moments by the pairwise update, classes united."""
        na, nb = float(self.n), float(other.n)
        n = na + nb
        d = other.stats[1] - self.stats[1]
        mean = self.stats[1] + d * nb / n
        m2 = self.stats[2] * max(na - 1, 0) + other.stats[2] * max(nb - 1, 0) + d * d * na * nb / n
        st = np.array([n, mean, m2 / (n - 1) if n > 1 else 0.0, min(self.stats[3], other.stats[3]),
                       max(self.stats[4], other.stats[4])])
        classes = counts = None
        if self.classes is not None and other.classes is not None:
            bits = np.union1d(self.classes.view(np.int64), other.classes.view(np.int64))
            classes = np.sort(bits.view(np.float64))
            key = {int(b): i for i, b in enumerate(classes.view(np.int64))}
            counts = np.zeros(len(classes), dtype=np.int64)
            for src in (self, other):
                for b, k in zip(src.classes.view(np.int64), src.class_counts):
                    counts[key[int(b)]] += int(k)
        return LabelInfo(self.name, int(n), st, classes, counts)


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

    def merge(self, other: "VectorStats") -> "VectorStats":
        """The statistics over the rows of both (two disjoint row sets of the same vector layout: the row chunks of a
        sample): pairwise update of the moments and co-moments, per-class sums added class by class."""
        na, nb = float(self.label.n), float(other.label.n)
        if nb == 0:
            return self
        if na == 0:
            return other
        n = na + nb
        lab = self.label.merge(other.label)
        d = other.mean - self.mean
        mean = self.mean + d * (nb / n)
        m2 = self.variance * max(na - 1, 0) + other.variance * max(nb - 1, 0) + d * d * (na * nb / n)
        var = m2 / (n - 1) if n > 1 else np.zeros_like(m2)
        dy = other.label.stats[1] - self.label.stats[1]
        cov = self.cov_label + other.cov_label + d * dy * (na * nb / n)
        cs = None
        if self.class_sums is not None and other.class_sums is not None and lab.classes is not None:
            key = {int(b): i for i, b in enumerate(lab.classes.view(np.int64))}
            cs = np.zeros((len(lab.classes), self.width))
            nan = np.zeros(self.width, dtype=bool)
            for src in (self, other):
                nan |= np.isnan(src.class_sums).all(axis=0) if len(src.class_sums) else False
                for b, row in zip(src.label.classes.view(np.int64), src.class_sums):
                    cs[key[int(b)]] += np.nan_to_num(row, nan=0.0)
            cs[:, nan] = np.nan                        # columns whose per-class sums are not tracked stay untracked
        return VectorStats(lab, mean, var, np.minimum(self.min, other.min), np.maximum(self.max, other.max), cov, cs)

    def select(self, idx: List[int]) -> "VectorStats":
        idx = np.asarray(idx, dtype=np.int64)
        if len(idx) == self.width and bool((idx == np.arange(self.width)).all()):
            return self
        cs = None if self.class_sums is None else self.class_sums[:, idx]
        return VectorStats(self.label, self.mean[idx], self.variance[idx], self.min[idx], self.max[idx],
                           self.cov_label[idx], cs)
