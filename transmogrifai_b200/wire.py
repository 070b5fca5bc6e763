"""OPVector -> Spark VectorUDT, columnar (the wire format either side of the path): what a
mapPartitions transform returns for the rows of a partition.

Reference: `.compressed` at the end of every vectorizer / combiner / SanityCheckerModel
(core/.../Transmogrifier.scala:466-475, features/.../RichVector.scala:120-134,
core/.../DerivedFeatureFilterUtils.scala:291-305) and Spark 2.4.5 ml.linalg.VectorUDT's struct
(type, size, indices, values)."""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from . import _lib as L
from . import runtime as R
from .columnar import DensePart, LazyDensePart, OneHotPart, SparseRowsPart, VectorColumn


@dataclass
class VectorUDTColumn:
    """Arrow-style arrays of a VectorUDT column: type u8 [n] (0 sparse, 1 dense), size i32 [n] (-1 = null: dense rows),
    indices = idx_offsets i64 [n + 1] + i32 [], values = val_offsets i64 [n + 1] + f64 []."""
    type: torch.Tensor
    size: torch.Tensor
    idx_offsets: torch.Tensor
    indices: torch.Tensor
    val_offsets: torch.Tensor
    values: torch.Tensor
    width: int

    def row(self, r: int):
        """Row r as (type, size, indices, values) -- numpy, for tests."""
        a, b = int(self.idx_offsets[r]), int(self.idx_offsets[r + 1])
        c, d = int(self.val_offsets[r]), int(self.val_offsets[r + 1])
        return int(self.type[r]), int(self.size[r]), self.indices[a:b].cpu().numpy(), self.values[c:d].cpu().numpy()


def to_vector_udt(vec: VectorColumn) -> VectorUDTColumn:
    """The rows of an OPVector column as VectorUDT arrays on the device (two HBM-bound kernels and a prefix sum)."""
    R.require_cuda()
    parts = [p if isinstance(p, (DensePart, LazyDensePart)) or not hasattr(p, "materialize") else p.materialize(None)
             for p in vec.parts]
    n, width = vec.n, sum(p.width for p in parts)
    desc = np.zeros((max(len(parts), 1), 9), dtype=np.int64)
    keep = []
    col0 = 0
    for i, p in enumerate(parts):
        if isinstance(p, (DensePart, LazyDensePart)):
            x = p.data if p.data.is_cuda else p.data.cuda()
            if x.stride(1) != 1:
                x = x.contiguous()
            keep.append(x)
            desc[i] = [0, p.width, col0, x.stride(0), 0, x.data_ptr(), 0, 0, 0]
        elif isinstance(p, OneHotPart):
            x = p.indices if p.indices.is_cuda else p.indices.cuda()
            if x.stride(1) != 1:
                x = x.contiguous()
            keep.append(x)
            desc[i] = [1, p.width, col0, x.stride(0), x.shape[1], x.data_ptr(), 0, 0, 0]
        elif isinstance(p, SparseRowsPart):
            desc[i] = [2, p.width, col0, 0, 0, p.row_start.data_ptr(), p.row_nnz.data_ptr(), p.indices.data_ptr(),
                       p.values.data_ptr()]
        else:
            raise NotImplementedError(f"cannot export a {type(p).__name__}")
        col0 += p.width
    d_desc = torch.from_numpy(desc).cuda()
    tp = torch.empty(max(n, 1), dtype=torch.uint8, device="cuda")
    sz = torch.empty(max(n, 1), dtype=torch.int32, device="cuda")
    il = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
    vl = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
    L.call("tmog_vector_export_count_dev", d_desc.data_ptr(), len(parts), n, width, tp.data_ptr(), sz.data_ptr(),
           il[1:].data_ptr(), vl[1:].data_ptr(), R.stream())
    io, vo = torch.cumsum(il, 0), torch.cumsum(vl, 0)
    ni, nv = int(io[-1].item()), int(vo[-1].item())
    idx = torch.empty(max(ni, 1), dtype=torch.int32, device="cuda")
    val = torch.empty(max(nv, 1), dtype=torch.float64, device="cuda")
    L.call("tmog_vector_export_write_dev", d_desc.data_ptr(), len(parts), n, width, tp.data_ptr(), io.data_ptr(),
           vo.data_ptr(), idx.data_ptr(), val.data_ptr(), R.stream())
    del keep
    return VectorUDTColumn(tp[:n], sz[:n], io, idx[:ni], vo, val[:nv], width)
