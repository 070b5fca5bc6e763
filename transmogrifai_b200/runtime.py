"""Process-wide runtime: one tmog_ctx per device, scratch tensors, stream handles.  torch provides device memory and
streams (plumbing); all arithmetic is in libtmog_b200."""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional

import numpy as np
import torch

from . import _lib as L

_ctx: Dict[int, C.c_void_p] = {}
_scratch: Dict[tuple, torch.Tensor] = {}


def require_cuda() -> None:
    if not torch.cuda.is_available():
        raise RuntimeError("transmogrifai_b200 needs a CUDA device: there is no CPU fallback")


def ctx(device: Optional[int] = None) -> C.c_void_p:
    require_cuda()
    dev = torch.cuda.current_device() if device is None else device
    if dev not in _ctx:
        h = C.c_void_p()
        L.call("tmog_ctx_create", dev, C.byref(h))
        _ctx[dev] = h
    return _ctx[dev]


def release() -> None:
    for h in _ctx.values():
        L.call("tmog_ctx_destroy", h)
    _ctx.clear()
    _scratch.clear()


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def scratch(tag: str, nbytes: int) -> torch.Tensor:
    """Grow-only device scratch, one buffer per (device, tag)."""
    key = (torch.cuda.current_device(), tag)
    t = _scratch.get(key)
    if t is None or t.numel() < nbytes:
        t = torch.empty(max(nbytes, 256), dtype=torch.uint8, device="cuda")
        _scratch[key] = t
    return t


def ptr(t) -> Optional[int]:
    """Raw pointer of a torch tensor / numpy array (None passes NULL)."""
    if t is None:
        return None
    if isinstance(t, np.ndarray):
        return t.ctypes.data
    return t.data_ptr()


def clear_caches() -> None:
    """Drop the content-keyed caches of the stage layer (metadata plans, column descriptions, small device constants).
    They hold nothing that depends on row data; a caller that wants every fit to start from nothing calls this."""
    from .stages import numeric, pivot, sanity
    sanity._PLAN_CACHE.clear()
    sanity._RUN_INFO.clear()
    pivot._COLUMN_DESCRIPTIONS.clear()
    numeric._DEV_CONST.clear()
