"""Row-sharded multi-GPU support: one process per GPU, rows split contiguously across ranks, and only the small
partial results cross NVLink -- the B200 counterpart of the reference's Spark aggregations (SURVEY.md section 2,
"collective call sites"): Aggregator.merge / treeAggregate / reduceByKey become one NCCL all-reduce of
[ncols] x {count, sums, min, max} (f64) and of the per-label-class count tables (i64).  No data-path collective.

torch.distributed is the transport (NCCL on GPUs; gloo in the CPU tests)."""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist

from ._lib import REAL_NSTAT


_LOCAL_ONLY = [0]


def world() -> Tuple[int, int]:
    if _LOCAL_ONLY[0]:
        return 0, 1
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


class local_only:
    """Inside this block the stage layer sees a single rank: fits and statistics cover this rank's rows only and make no
    exchange (for work whose pieces are merged explicitly afterwards, e.g. the per-chunk statistics of a row sample)."""

    def __enter__(self):
        _LOCAL_ONLY[0] += 1
        return self

    def __exit__(self, *exc):
        _LOCAL_ONLY[0] -= 1
        return False


def _device() -> torch.device:
    return torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")


def allreduce_real_partials(partials: np.ndarray) -> np.ndarray:
    """Merge [(ncols+1) * NSTAT] partial blocks of all ranks (they must share their shifts): sums and counts add,
    [5] is a min, [6] a max -- tmog_real_partials_merge across ranks."""
    if world()[1] == 1:
        return partials
    p = torch.from_numpy(partials.reshape(-1, REAL_NSTAT).copy()).to(_device())
    sums = p[:, [0, 1, 2, 3, 4, 7]].contiguous()
    ext = torch.stack([p[:, 5], -p[:, 6]], dim=1).contiguous()
    dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    dist.all_reduce(ext, op=dist.ReduceOp.MIN)
    p[:, [0, 1, 2, 3, 4, 7]] = sums
    p[:, 5] = ext[:, 0]
    p[:, 6] = -ext[:, 1]
    return p.cpu().numpy().reshape(-1)


def gather_fit_results(partials, tables=()):
    """Everything one fit scan leaves behind, merged over the ranks in ONE exchange: the [(ncols+1) * NSTAT] partial
    block (device tensor or numpy; merged in rank order by tmog_real_partials_merge, so the result does not depend on
    the reduction tree) and any number of int64 count tables (summed).  The blocks travel as raw bytes in a single
    all-gather followed by a single device-to-host copy.  -> (partials, [tables]) as numpy arrays."""
    from . import _lib as L
    as_t = lambda a: a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))  # noqa: E731
    p = as_t(partials).reshape(-1)
    ts = [as_t(t) for t in tables]
    shapes = [tuple(t.shape) for t in ts]
    if world()[1] == 1:
        if not ts or not p.is_cuda:
            return p.cpu().numpy(), [t.cpu().numpy() for t in ts]
        # one device-to-host copy for all of it
        flat = torch.cat([p] + [t.contiguous().reshape(-1).view(torch.float64) for t in ts]).cpu().numpy()
        outs, at = [], p.numel()
        for shp in shapes:
            k = int(np.prod(shp)) if len(shp) else 1
            outs.append(flat[at:at + k].view(np.int64).reshape(shp).copy())
            at += k
        return flat[:p.numel()].copy(), outs
    dev = _device()
    flat = torch.cat([p.to(dev)] + [t.to(dev).contiguous().reshape(-1).view(torch.float64) for t in ts])
    out = torch.empty(world()[1] * flat.numel(), dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(out, flat)
    host = out.cpu().numpy().reshape(world()[1], flat.numel())
    npart = p.numel()
    ncols = npart // REAL_NSTAT - 1
    merged = host[0, :npart].copy()
    for r in range(1, world()[1]):
        other = np.ascontiguousarray(host[r, :npart])
        L.call("tmog_real_partials_merge", merged.ctypes.data, other.ctypes.data, ncols)
    outs, at = [], npart
    for shp in shapes:
        k = int(np.prod(shp)) if len(shp) else 1
        outs.append(np.ascontiguousarray(host[:, at:at + k]).view(np.int64).sum(axis=0).reshape(shp))
        at += k
    return merged, outs


def broadcast(a: np.ndarray, src: int = 0) -> np.ndarray:
    if world()[1] == 1:
        return a
    t = torch.from_numpy(np.ascontiguousarray(a)).to(_device())
    dist.broadcast(t, src=src)
    return t.cpu().numpy()


def allreduce_sum(a: np.ndarray) -> np.ndarray:
    if world()[1] == 1:
        return a
    t = torch.from_numpy(np.ascontiguousarray(a)).to(_device())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.cpu().numpy()


def allreduce_max(a: np.ndarray) -> np.ndarray:
    if world()[1] == 1:
        return a
    t = torch.from_numpy(np.ascontiguousarray(a)).to(_device())
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.cpu().numpy()


def all_gather_object(obj) -> list:
    """Small driver-side objects of every rank, in rank order (TextStats maps: what Spark's reduce would merge)."""
    if world()[1] == 1:
        return [obj]
    out = [None] * world()[1]
    dist.all_gather_object(out, obj)
    return out


def broadcast_object(obj, src: int = 0):
    if world()[1] == 1:
        return obj
    box = [obj if world()[0] == src else None]
    dist.broadcast_object_list(box, src=src, device=_device())
    return box[0]


def union_classes(local: Optional[np.ndarray], cap: int = 255) -> Optional[np.ndarray]:
    """Union of every rank's sorted distinct label values; None if any rank (or the union) exceeds `cap`."""
    if world()[1] == 1:
        return local
    buf = np.full(cap + 1, np.nan)
    k = -1 if local is None else len(local)
    buf[0] = k
    if k > 0:
        buf[1:1 + k] = local
    t = torch.from_numpy(buf).to(_device())
    out = torch.empty(world()[1] * t.numel(), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, t)
    vals = []
    for o in out.cpu().numpy().reshape(world()[1], -1):      # one device-to-host copy for all ranks
        kk = int(o[0])
        if kk < 0:
            return None
        vals.append(o[1:1 + kk])
    allv = np.concatenate(vals) if vals else np.zeros(0)
    # distinct by bit pattern (java.lang.Double.equals), ascending by value with -0.0 < 0.0 and NaN last
    bits = np.unique(allv.view(np.int64))
    u = bits.view(np.float64)
    key = np.where(bits < 0, ~bits.view(np.uint64), bits.view(np.uint64) | np.uint64(1 << 63))
    u = u[np.argsort(key, kind="stable")]
    return None if len(u) > cap else u


def gather_label_classes(ncls: torch.Tensor, classes: torch.Tensor, counts: torch.Tensor, cap: int = 255):
    """Label classes of all ranks in one exchange.  ncls (int32 [1]: number of local classes, < 0 = more than the cap),
    classes (f64 [256], sorted) and counts (int64 [256], rows per local class) are device tensors as
    tmog_label_classes_dev leaves them.  -> (union, same, counts): the union of the class lists (None if any rank or
    the union exceeds `cap`), whether every rank's own list already equals it (then the class ids each rank computed
    against its own list are valid as they are), and in that case the rows per class summed over the ranks."""
    size = world()[1]
    dev = _device()
    flat = torch.cat([ncls.to(dev).double(), classes.to(dev), counts.to(dev).view(torch.float64)])
    out = torch.empty(size * flat.numel(), dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(out, flat)
    host = out.cpu().numpy().reshape(size, flat.numel())
    w = classes.numel()
    ks = [int(h[0]) for h in host]
    if any(k < 0 for k in ks):
        return None, False, None
    lists = [host[r, 1:1 + ks[r]] for r in range(size)]
    allv = np.concatenate(lists) if lists else np.zeros(0)
    # distinct by bit pattern (java.lang.Double.equals), ascending by value with -0.0 < 0.0 and NaN last
    bits = np.unique(allv.view(np.int64))
    key = np.where(bits < 0, ~bits.view(np.uint64), bits.view(np.uint64) | np.uint64(1 << 63))
    union = bits.view(np.float64)[np.argsort(key, kind="stable")]
    if len(union) > cap:
        return None, False, None
    same = all(len(v) == len(union) and np.array_equal(v.view(np.int64), union.view(np.int64)) for v in lists)
    total = None
    if same:
        total = np.ascontiguousarray(host[:, 1 + w:1 + w + len(union)]).view(np.int64).sum(axis=0)
    return union, same, total


def shard_rows(n: int, rank: int, size: int, align: int = 2048) -> Tuple[int, int]:
    """Contiguous row range of `rank` (starts aligned so that bitmap words and 16-byte loads stay aligned)."""
    per = (n + size - 1) // size
    per = (per + align - 1) // align * align
    lo = min(n, rank * per)
    return lo, min(n, lo + per)
