"""Columnar batches: the form in which Spark DataFrame partitions reach the library (SURVEY.md 8b "data at the
boundary") -- Arrow-style column blocks in pinned host memory or, once uploaded, in HBM.

* numeric block : several same-typed columns stored column-major, values [ncols, ld] (f64 for Real/RealNN, u8 for
                  Binary), validity bitmaps [ncols, ldv] (bit r%64 of 64-bit word r/64, 1 = present) or None
* vector column : an OPVector -- ordered parts (dense row-major blocks and CSR blocks) plus OpVectorMetadata, and an
                  optional statistics sidecar filled by the fused vectorizer scans (see stages/sanity.py)

torch tensors are the storage (CPU pinned or CUDA): plumbing only, every computation goes through libtmog_b200.
"""
from __future__ import annotations

import itertools
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import features as F
from .metadata import OpVectorMetadata


def round_up(v: int, m: int) -> int:
    return (v + m - 1) // m * m


def pack_validity(valid: np.ndarray) -> np.ndarray:
    """bool [ncols, n] -> bitmap words [ncols, ceil(n/64)] as int64 (bit pattern of uint64)."""
    ncols, n = valid.shape
    nw = max((n + 63) // 64, 1)
    padded = np.zeros((ncols, nw * 64), dtype=np.uint8)
    padded[:, :n] = valid
    bits = np.packbits(padded.reshape(ncols, nw, 64), axis=2, bitorder="little")
    return bits.reshape(ncols, nw * 8).view("<i8").reshape(ncols, nw).copy()


def _pin(t: torch.Tensor) -> torch.Tensor:
    if t.is_cuda or not torch.cuda.is_available():
        return t
    try:
        return t.pin_memory()
    except RuntimeError:
        return t


def to_host(tensors, piece_bytes: Optional[int] = None) -> list:
    """Device tensors -> pinned host tensors: every copy is queued asynchronously on the current stream and the
    stream is waited for once (what a reader of the transformed partition does with the kept columns).  Pinned
    buffers come from torch's caching host allocator, so repeated reads of same-sized results reuse them; a
    pageable destination would halve the PCIe rate and fault in fresh pages on every call.  With `piece_bytes` a
    tensor larger than that comes back as a list of flat pinned pieces in order (the allocator rounds a request up
    to a power of two, which a 13 GB array cannot afford)."""
    out, pending = [], False

    def pull(t):
        h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        h.copy_(t, non_blocking=True)
        return h

    for t in tensors:
        if t is None or not t.is_cuda:
            out.append(t)
            continue
        pending = True
        nbytes = t.numel() * t.element_size()
        if piece_bytes is None or nbytes <= piece_bytes or not t.is_contiguous():
            out.append(pull(t))
        else:
            flat, step = t.reshape(-1), max(piece_bytes // t.element_size(), 1)
            out.append([pull(flat[i:i + step]) for i in range(0, flat.numel(), step)])
    if pending:
        torch.cuda.current_stream().synchronize()
    return out


def take_bitmap(valid: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """Arrow validity words [..., ldv] of the rows `idx` (int64, ascending or not), repacked: [..., ceil(m / 64)]."""
    m = idx.numel()
    nw = max((m + 63) // 64, 1)
    bits = (valid.index_select(-1, idx >> 6) >> (idx & 63)) & 1
    pad = torch.zeros(valid.shape[:-1] + (nw * 64 - m,), dtype=torch.int64, device=valid.device)
    bits = torch.cat([bits, pad], dim=-1).reshape(valid.shape[:-1] + (nw, 64))
    weights = torch.ones(64, dtype=torch.int64, device=valid.device) << torch.arange(64, device=valid.device)
    return (bits * weights).sum(dim=-1)


def take_ragged(data: torch.Tensor, offsets: torch.Tensor, idx: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Rows `idx` of a ragged column (Arrow offsets int64 [n + 1] into `data`): (data', offsets' [m + 1])."""
    a = offsets.index_select(0, idx)
    lens = offsets.index_select(0, idx + 1) - a
    new_off = torch.zeros(idx.numel() + 1, dtype=torch.int64, device=offsets.device)
    new_off[1:] = torch.cumsum(lens, 0)
    total = int(new_off[-1].item())
    if total == 0:
        return torch.zeros(1, dtype=data.dtype, device=data.device), new_off
    src = torch.repeat_interleave(a - new_off[:-1], lens) + torch.arange(total, device=data.device)
    return data.index_select(0, src), new_off


@dataclass
class NumericBlock:
    names: List[str]
    type_name: str
    values: torch.Tensor                 # [ncols, ld] float64 (Real, RealNN), uint8 (Binary) or int64 (Integral)
    valid: Optional[torch.Tensor]        # [ncols, ldv] int64 bitmap words, or None (no nulls)
    n: int

    @property
    def is_cuda(self) -> bool:
        return self.values.is_cuda

    @property
    def ld(self) -> int:
        return self.values.shape[1]

    @property
    def ldv(self) -> int:
        return 0 if self.valid is None else self.valid.shape[1]

    def rows(self, idx: Sequence[int]) -> "NumericBlock":
        """Sub-block of the given columns: a view when they are a contiguous run, else a gathered copy."""
        idx = list(idx)
        if idx == list(range(idx[0], idx[0] + len(idx))):
            v = self.values[idx[0]:idx[0] + len(idx)]
            m = None if self.valid is None else self.valid[idx[0]:idx[0] + len(idx)]
        else:
            sel = torch.as_tensor(idx, device=self.values.device)
            v = self.values.index_select(0, sel)
            m = None if self.valid is None else self.valid.index_select(0, sel)
        return NumericBlock([self.names[i] for i in idx], self.type_name, v, m, self.n)

    def take_rows(self, idx: torch.Tensor) -> "NumericBlock":
        """The rows `idx` (int64 tensor on the block's device) as a block of their own."""
        m = idx.numel()
        ld = round_up(max(m, 1), 16 if self.values.dtype == torch.uint8 else 2)
        vals = torch.zeros((self.values.shape[0], ld), dtype=self.values.dtype, device=self.values.device)
        vals[:, :m] = self.values.index_select(1, idx)
        valid = None if self.valid is None else take_bitmap(self.valid, idx)
        return NumericBlock(self.names, self.type_name, vals, valid, m)

    def to(self, device) -> "NumericBlock":
        return NumericBlock(self.names, self.type_name, self.values.to(device, non_blocking=True),
                            None if self.valid is None else self.valid.to(device, non_blocking=True), self.n)

    @staticmethod
    def from_lists(names: Sequence[str], type_name: str, columns: Sequence[Sequence], pin: bool = True
                   ) -> "NumericBlock":
        n = len(columns[0]) if columns else 0
        binary, integral = type_name == F.BINARY, type_name in F.INTEGRAL_TYPES
        ld = round_up(max(n, 1), 16 if binary else 2)
        vals = np.zeros((len(columns), ld), dtype=np.uint8 if binary else np.int64 if integral else np.float64)
        ok = np.zeros((len(columns), n), dtype=bool)
        for j, col in enumerate(columns):
            if len(col) != n:
                raise ValueError("columns of one batch must have the same length")
            for i, v in enumerate(col):
                if v is not None:
                    ok[j, i] = True
                    vals[j, i] = (1 if v else 0) if binary else int(v) if integral else float(v)
        if type_name == F.REALNN and not ok.all():
            raise ValueError("None.get: a RealNN feature holds a null")   # RealNNVectorizer.scala:53
        valid = None if ok.all() else torch.from_numpy(pack_validity(ok))
        v = torch.from_numpy(vals)
        if pin:
            v, valid = _pin(v), (None if valid is None else _pin(valid))
        return NumericBlock(list(names), type_name, v, valid, n)


def encode_strings(strings: Sequence[str]) -> Tuple[np.ndarray, np.ndarray]:
    """Arrow utf8 layout: (bytes uint8 [total], offsets int64 [count + 1])."""
    enc = [t.encode("utf-8") for t in strings]
    offsets = np.zeros(len(enc) + 1, dtype=np.int64)
    if enc:
        offsets[1:] = np.cumsum([len(e) for e in enc])
    data = np.frombuffer(b"".join(enc), dtype=np.uint8).copy() if enc else np.zeros(0, dtype=np.uint8)
    return data, offsets


def decode_strings(data: np.ndarray, offsets: np.ndarray) -> List[str]:
    b = data.tobytes()
    return [b[offsets[i]:offsets[i + 1]].decode("utf-8") for i in range(len(offsets) - 1)]


_DICT_ARRAYS: Dict[int, tuple] = {}     # id(dictionary list) -> (list, bytes, offsets)


@dataclass
class DictBlock:
    """Dictionary-encoded string columns (PickList / Text): codes [ncols, ld] int32 column-major (code < 0 = null,
    ld a multiple of 4) and one dictionary (list of distinct raw strings) per column -- the Arrow DictionaryArray the
    columnar export produces."""
    names: List[str]
    type_name: str
    codes: torch.Tensor
    dicts: List[List[str]]
    n: int

    @property
    def is_cuda(self) -> bool:
        return self.codes.is_cuda

    @property
    def ld(self) -> int:
        return self.codes.shape[1]

    def dict_arrays(self, col: int) -> Tuple[np.ndarray, np.ndarray]:
        """Column `col`'s dictionary in Arrow utf8 layout (bytes, int64 offsets) -- the form the columnar export
        delivers; built once per dictionary object."""
        d = self.dicts[col]
        cache = _DICT_ARRAYS.get(id(d))
        if cache is None or cache[0] is not d:
            cache = (d,) + encode_strings(d)
            _DICT_ARRAYS[id(d)] = cache
        return cache[1], cache[2]

    def rows(self, idx: Sequence[int]) -> "DictBlock":
        idx = list(idx)
        if idx == list(range(idx[0], idx[0] + len(idx))):
            c = self.codes[idx[0]:idx[0] + len(idx)]
        else:
            c = self.codes.index_select(0, torch.as_tensor(idx, device=self.codes.device))
        return DictBlock([self.names[i] for i in idx], self.type_name, c, [self.dicts[i] for i in idx], self.n)

    def take_rows(self, idx: torch.Tensor) -> "DictBlock":
        m = idx.numel()
        codes = torch.full((self.codes.shape[0], round_up(max(m, 1), 4)), -1, dtype=torch.int32,
                           device=self.codes.device)
        codes[:, :m] = self.codes.index_select(1, idx)
        return DictBlock(self.names, self.type_name, codes, self.dicts, m)

    def to(self, device) -> "DictBlock":
        return DictBlock(self.names, self.type_name, self.codes.to(device, non_blocking=True), self.dicts, self.n)

    @staticmethod
    def from_lists(names: Sequence[str], type_name: str, columns: Sequence[Sequence[Optional[str]]],
                   pin: bool = True) -> "DictBlock":
        n = len(columns[0]) if columns else 0
        ld = round_up(max(n, 1), 4)
        codes = np.full((len(columns), ld), -1, dtype=np.int32)
        dicts: List[List[str]] = []
        for j, col in enumerate(columns):
            seen: Dict[str, int] = {}
            for i, v in enumerate(col):
                if v is not None:
                    codes[j, i] = seen.setdefault(v, len(seen))     # first-appearance order
            dicts.append(list(seen))
        t = torch.from_numpy(codes)
        return DictBlock(list(names), type_name, _pin(t) if pin else t, dicts, n)


@dataclass
class SetBlock:
    """Set-valued string columns (MultiPickList) in Arrow List<dictionary> layout, per column: offsets int64 [n + 1],
    codes int32 [offsets[n]] (row r holds codes[offsets[r]:offsets[r + 1]], no duplicates; an empty range is the empty
    set, which is also what a null becomes), and the dictionary of distinct raw strings."""
    names: List[str]
    type_name: str
    offsets: List[torch.Tensor]
    codes: List[torch.Tensor]
    dicts: List[List[str]]
    n: int

    @property
    def is_cuda(self) -> bool:
        return bool(self.offsets) and self.offsets[0].is_cuda

    def dict_arrays(self, col: int) -> Tuple[np.ndarray, np.ndarray]:
        d = self.dicts[col]
        cache = _DICT_ARRAYS.get(id(d))
        if cache is None or cache[0] is not d:
            cache = (d,) + encode_strings(d)
            _DICT_ARRAYS[id(d)] = cache
        return cache[1], cache[2]

    def rows(self, idx: Sequence[int]) -> "SetBlock":
        idx = list(idx)
        return SetBlock([self.names[i] for i in idx], self.type_name, [self.offsets[i] for i in idx],
                        [self.codes[i] for i in idx], [self.dicts[i] for i in idx], self.n)

    def take_rows(self, idx: torch.Tensor) -> "SetBlock":
        pairs = [take_ragged(c, o, idx) for c, o in zip(self.codes, self.offsets)]
        return SetBlock(self.names, self.type_name, [p[1] for p in pairs], [p[0] for p in pairs], self.dicts,
                        idx.numel())

    def to(self, device) -> "SetBlock":
        mv = lambda t: t.to(device, non_blocking=True)  # noqa: E731
        return SetBlock(self.names, self.type_name, [mv(t) for t in self.offsets], [mv(t) for t in self.codes],
                        self.dicts, self.n)

    @staticmethod
    def from_lists(names: Sequence[str], type_name: str, columns: Sequence[Sequence], pin: bool = True) -> "SetBlock":
        """Cells: None, or any iterable of strings (duplicates collapse, first-seen order kept: a Scala Set)."""
        n = len(columns[0]) if columns else 0
        offsets, codes, dicts = [], [], []
        for col in columns:
            seen: Dict[str, int] = {}
            offs = np.zeros(n + 1, dtype=np.int64)
            flat: List[int] = []
            for i, cell in enumerate(col):
                if cell is not None:
                    for v in dict.fromkeys(cell):
                        flat.append(seen.setdefault(v, len(seen)))
                offs[i + 1] = len(flat)
            to = torch.from_numpy(offs)
            tc = torch.from_numpy(np.asarray(flat if flat else [0], dtype=np.int32)[:max(len(flat), 1)])
            offsets.append(_pin(to) if pin else to)
            codes.append(_pin(tc) if pin else tc)
            dicts.append(list(seen))
        return SetBlock(list(names), type_name, offsets, codes, dicts, n)


@dataclass
class TextBlock:
    """Raw utf8 string columns (Text / TextArea) in Arrow layout, one (data, offsets, validity) triple per column:
    data uint8 [total bytes], offsets int64 [n + 1], valid int64 bitmap words [ceil(n/64)] or None."""
    names: List[str]
    type_name: str
    data: List[torch.Tensor]
    offsets: List[torch.Tensor]
    valid: List[Optional[torch.Tensor]]
    n: int

    @property
    def is_cuda(self) -> bool:
        return self.offsets[0].is_cuda if self.offsets else False

    def rows(self, idx: Sequence[int]) -> "TextBlock":
        idx = list(idx)
        return TextBlock([self.names[i] for i in idx], self.type_name, [self.data[i] for i in idx],
                         [self.offsets[i] for i in idx], [self.valid[i] for i in idx], self.n)

    def take_rows(self, idx: torch.Tensor) -> "TextBlock":
        pairs = [take_ragged(d, o, idx) for d, o in zip(self.data, self.offsets)]
        valid = [None if v is None else take_bitmap(v, idx) for v in self.valid]
        return TextBlock(self.names, self.type_name, [p[0] for p in pairs], [p[1] for p in pairs], valid, idx.numel())

    def to(self, device) -> "TextBlock":
        mv = lambda t: None if t is None else t.to(device, non_blocking=True)  # noqa: E731
        return TextBlock(self.names, self.type_name, [mv(t) for t in self.data], [mv(t) for t in self.offsets],
                         [mv(t) for t in self.valid], self.n)

    def cell(self, col: int, row: int) -> Optional[str]:
        """One cell as a Python string (host round trip; used for the handful of vocabulary rows only)."""
        v = self.valid[col]
        if v is not None and not (int(v[row >> 6].item()) >> (row & 63)) & 1:
            return None
        a, b = int(self.offsets[col][row].item()), int(self.offsets[col][row + 1].item())
        return bytes(self.data[col][a:b].cpu().numpy().tobytes()).decode("utf-8")

    def prefix_cells(self, col: int, rows: int) -> List[Optional[str]]:
        """The first `rows` cells of a column with three host round trips in all (offsets, bytes, validity)."""
        rows = min(rows, self.n)
        off = self.offsets[col][:rows + 1].cpu().numpy()
        raw = self.data[col][int(off[0]):int(off[-1])].cpu().numpy().tobytes() if rows else b""
        v = self.valid[col]
        ok = np.ones(rows, dtype=bool) if v is None else np.unpackbits(
            v[:(rows + 63) // 64].cpu().numpy().view(np.uint8), bitorder="little")[:rows].astype(bool)
        base = int(off[0])
        return [raw[int(off[r]) - base:int(off[r + 1]) - base].decode("utf-8") if ok[r] else None
                for r in range(rows)]

    def prefix_cells_many(self, cols: Sequence[int], rows: int) -> Dict[int, List[Optional[str]]]:
        """prefix_cells of several columns with three host round trips in all."""
        rows = min(rows, self.n)
        cols = list(cols)
        if not cols or rows == 0:
            return {c: [] for c in cols}
        offs = torch.stack([self.offsets[c][:rows + 1] for c in cols]).cpu().numpy()
        raw = torch.cat([self.data[c][int(o[0]):int(o[-1])] for c, o in zip(cols, offs)]).cpu().numpy().tobytes()
        words = (rows + 63) // 64
        with_valid = [c for c in cols if self.valid[c] is not None]
        bits = {}
        if with_valid:
            vw = torch.stack([self.valid[c][:words] for c in with_valid]).cpu().numpy()
            for c, w in zip(with_valid, vw):
                bits[c] = np.unpackbits(w.view(np.uint8), bitorder="little")[:rows].astype(bool)
        out, at = {}, 0
        for c, o in zip(cols, offs):
            ok = bits.get(c)
            base = int(o[0]) - at
            out[c] = [raw[int(o[r]) - base:int(o[r + 1]) - base].decode("utf-8") if (ok is None or ok[r]) else None
                      for r in range(rows)]
            at += int(o[-1]) - int(o[0])
        return out

    def cells(self, col: int, rows: Sequence[int]) -> List[Optional[str]]:
        """Arbitrary cells of a column: one round trip for their offsets, one per cell for the bytes."""
        if not len(rows):
            return []
        idx = torch.as_tensor(np.asarray(rows, dtype=np.int64), device=self.offsets[col].device)
        a = self.offsets[col].index_select(0, idx).cpu().numpy()
        b = self.offsets[col].index_select(0, idx + 1).cpu().numpy()
        out = []
        for r, lo, hi in zip(rows, a, b):
            v = self.valid[col]
            if v is not None and not (int(v[int(r) >> 6].item()) >> (int(r) & 63)) & 1:
                out.append(None)
            else:
                out.append(self.data[col][int(lo):int(hi)].cpu().numpy().tobytes().decode("utf-8"))
        return out

    @staticmethod
    def from_lists(names: Sequence[str], type_name: str, columns: Sequence[Sequence[Optional[str]]],
                   pin: bool = True) -> "TextBlock":
        n = len(columns[0]) if columns else 0
        data, offsets, valid = [], [], []
        for col in columns:
            enc = [b"" if v is None else v.encode("utf-8") for v in col]
            off = np.zeros(n + 1, dtype=np.int64)
            if n:
                off[1:] = np.cumsum([len(e) for e in enc])
            buf = np.frombuffer(b"".join(enc), dtype=np.uint8).copy() if off[-1] else np.zeros(1, dtype=np.uint8)
            ok = np.array([[v is not None for v in col]], dtype=bool)
            d, o = torch.from_numpy(buf), torch.from_numpy(off)
            m = None if ok.all() else torch.from_numpy(pack_validity(ok)[0])
            if pin:
                d, o, m = _pin(d), _pin(o), (None if m is None else _pin(m))
            data.append(d)
            offsets.append(o)
            valid.append(m)
        return TextBlock(list(names), type_name, data, offsets, valid, n)


def _is_identity(keep, width: int) -> bool:
    k = np.asarray(keep)
    return len(k) == width and bool((k == np.arange(width)).all())


@dataclass
class DensePart:
    data: torch.Tensor                   # [n, width] float64 row-major (may be a column slice of a wider block)

    @property
    def width(self) -> int:
        return self.data.shape[1]

    def to_scipy(self):
        import scipy.sparse as sp
        return sp.csr_matrix(self.data.cpu().numpy())


class LazyDensePart:
    """A dense block that is only written when somebody needs it.  The vectorizer models return these, so that
    SanityChecker -- which fits from the statistics sidecar, not from the data -- can decide first and the transform
    kernel then writes just the kept columns, once (``materialize(keep)``)."""

    def __init__(self, width: int, materialize):
        self.width = width
        self._materialize = materialize      # (keep: Optional[List[int]]) -> tensor [n, len(keep) or width]
        self._data = None

    def materialize(self, keep=None) -> torch.Tensor:
        if keep is None or _is_identity(keep, self.width):
            return self.data
        if self._data is not None:
            idx = torch.as_tensor(np.asarray(keep, dtype=np.int64), device=self._data.device)
            return self._data.index_select(1, idx)
        return self._materialize(np.asarray(keep, dtype=np.int64))

    @property
    def data(self) -> torch.Tensor:
        if self._data is None:
            self._data = self._materialize(None)
        return self._data

    def to_scipy(self):
        import scipy.sparse as sp
        return sp.csr_matrix(self.data.cpu().numpy())


@dataclass
class CsrPart:
    indptr: torch.Tensor                 # [n+1] int64
    indices: torch.Tensor                # [nnz] int32, sorted per row, local to this part
    values: torch.Tensor                 # [nnz] float64
    width: int

    def to_scipy(self):
        import scipy.sparse as sp
        n = self.indptr.shape[0] - 1
        m = sp.csr_matrix((self.values.cpu().numpy(), self.indices.cpu().numpy().astype(np.int64),
                           self.indptr.cpu().numpy()), shape=(n, self.width))
        m.eliminate_zeros()
        return m


@dataclass
class OneHotPart:
    """Sparse block with at most `k` unit entries per row: indices [n, k] int32 are column positions local to the
    part (-1 = no entry; the value 1.0 is implicit).  What the pivot vectorizers emit for single-valued columns."""
    indices: torch.Tensor
    width: int

    def to_scipy(self):
        import scipy.sparse as sp
        idx = self.indices.cpu().numpy()
        n, k = idx.shape
        rows = np.repeat(np.arange(n), k)
        flat = idx.reshape(-1)
        ok = flat >= 0
        m = sp.csr_matrix((np.ones(int(ok.sum())), (rows[ok], flat[ok].astype(np.int64))), shape=(n, self.width))
        m.sum_duplicates()
        return m


class LazyOneHotPart:
    """OneHotPart written on demand (see LazyDensePart): materialize(keep) emits indices renumbered to the kept
    columns, -1 for dropped ones."""

    def __init__(self, width: int, materialize):
        self.width = width
        self._materialize = materialize      # (keep: Optional[List[int]]) -> OneHotPart
        self._part = None

    def materialize(self, keep=None) -> OneHotPart:
        if keep is None or _is_identity(keep, self.width):
            if self._part is None:
                self._part = self._materialize(None)
            return self._part
        return self._materialize(np.asarray(keep, dtype=np.int64))

    @property
    def data(self):
        return self.materialize(None).indices

    def to_scipy(self):
        return self.materialize(None).to_scipy()


@dataclass
class SparseRowsPart:
    """Sparse block in 4-array CSR form: row r owns indices / values [row_start[r], row_start[r] + row_nnz[r]),
    indices ascending and local to the part.  What the hashing vectorizer emits (rows are written by whichever CTA
    finishes them, so their storage order is not the row order)."""
    row_start: torch.Tensor              # [n] int64
    row_nnz: torch.Tensor                # [n] int32
    indices: torch.Tensor                # [nnz] int32
    values: torch.Tensor                 # [nnz] int32 (term frequencies)
    width: int
    used: Optional[torch.Tensor] = None  # device int64 [1]: entries written (indices / values may be longer), or None

    def to_scipy(self):
        import scipy.sparse as sp
        start = self.row_start.cpu().numpy()
        nnz = self.row_nnz.cpu().numpy().astype(np.int64)
        idx = self.indices.cpu().numpy()
        val = self.values.cpu().numpy()
        n = len(start)
        indptr = np.zeros(n + 1, dtype=np.int64)
        indptr[1:] = np.cumsum(nnz)
        take = np.repeat(start - indptr[:-1], nnz) + np.arange(indptr[-1])
        m = sp.csr_matrix((val[take].astype(np.float64), idx[take].astype(np.int64), indptr), shape=(n, self.width))
        m.eliminate_zeros()
        return m


class LazyPart:
    """A part of any kind written on demand: materialize(keep) returns a concrete part restricted to `keep`."""

    def __init__(self, width: int, materialize):
        self.width = width
        self._materialize = materialize
        self._part = None

    def materialize(self, keep=None):
        if keep is None or _is_identity(keep, self.width):
            if self._part is None:
                self._part = self._materialize(None)
            return self._part
        return self._materialize(np.asarray(keep, dtype=np.int64))

    def to_scipy(self):
        return self.materialize(None).to_scipy()


class WidthPart:
    """Stand-in for the parts of a vector that is never materialised as a whole (a dataset streamed in row chunks):
    VectorsCombiner and SanityChecker only need its width, the statistics come from the sidecar."""

    def __init__(self, width: int):
        self.width = width

    def materialize(self, keep=None):
        raise ValueError("the vector of a streamed dataset exists chunk by chunk only")


@dataclass
class VectorColumn:
    name: str
    n: int
    metadata: OpVectorMetadata
    parts: List[object]                  # DensePart | CsrPart, in output-column order
    stats: Optional[object] = None       # stages.sanity.VectorStats sidecar
    producer: Optional[object] = field(default=None, compare=False, repr=False)   # the transformer that made it

    @property
    def width(self) -> int:
        return sum(p.width for p in self.parts)

    def recompute(self, full: "ColumnarBatch", sub: "ColumnarBatch") -> "VectorColumn":
        """This vector over another batch of raw columns (`sub`, e.g. a row sample of `full`): the transformer that
        produced it is run again, after the vectors it takes as input have been recomputed the same way."""
        if self.producer is None:
            raise ValueError(f"vector '{self.name}' does not know the stage that produced it")
        for f in self.producer.inN:
            if f.name in full.vectors and f.name not in sub.vectors:
                sub = sub.with_vector(full.vectors[f.name].recompute(full, sub))
        out = self.producer.transformColumn(sub)
        out.producer = self.producer
        if out.stats is None and sub.response and hasattr(self.producer, "stats_for"):
            # the statistics sidecar of the new rows from the same fused scans that serve a fit (no dense vector)
            out.stats = self.producer.stats_for(sub)
        return out

    def to_scipy(self):
        import scipy.sparse as sp
        if not self.parts:
            return sp.csr_matrix((self.n, 0))
        return sp.hstack([p.to_scipy() for p in self.parts], format="csr")

    def to_numpy(self) -> np.ndarray:
        return np.asarray(self.to_scipy().todense())


_batch_tokens = itertools.count(1)


def batch_key(batch: "ColumnarBatch", block=None) -> int:
    """Identity of the raw rows a stage was fitted on: the batch's token (never reused, unlike a device address, which
    the allocator hands out again once a batch is freed).  A batch is immutable once built; refilling its buffers in
    place is not supported."""
    return batch.token


@dataclass
class ColumnarBatch:
    n: int
    blocks: List[object] = field(default_factory=list)     # NumericBlock | DictBlock
    vectors: Dict[str, VectorColumn] = field(default_factory=dict)
    response: Optional[str] = None       # name of the response (label) column, a RealNN
    token: int = field(default_factory=lambda: next(_batch_tokens), compare=False)   # unique per set of raw rows

    def find(self, name: str):
        index = self.__dict__.get("_name_index")
        if index is None or index[0] != len(self.blocks):
            table = {}
            for b in self.blocks:
                for i, nm in enumerate(b.names):
                    table.setdefault(nm, (b, i))
            index = (len(self.blocks), table)
            self.__dict__["_name_index"] = index
        try:
            return index[1][name]
        except KeyError:
            raise KeyError(f"no column named {name!r} in the batch") from None

    def has(self, name: str) -> bool:
        return name in self.vectors or any(name in b.names for b in self.blocks)

    def strings(self, names: Sequence[str]) -> "DictBlock":
        """The dictionary-encoded block holding exactly these columns, in this order."""
        located = [self.find(nm) for nm in names]
        first = located[0][0]
        if all(isinstance(b, TextBlock) for b, _ in located):
            # raw text handed to a pivot stage: dictionary-encode it (host side, cached per column set)
            cache = self.__dict__.setdefault("_dict_cache", {})
            key = tuple(names)
            if key not in cache:
                cols = [[b.cell(i, r) for r in range(self.n)] for b, i in located]
                blk = DictBlock.from_lists(list(names), first.type_name, cols)
                cache[key] = blk.to("cuda") if first.is_cuda else blk
            return cache[key]
        if not all(isinstance(b, DictBlock) for b, _ in located):
            raise ValueError("not string columns")
        if all(b is first for b, _ in located):
            return first.rows([i for _, i in located])
        subs = [b.rows([i]) for b, i in located]
        return DictBlock(list(names), subs[0].type_name, torch.cat([t.codes for t in subs], dim=0),
                         [t.dicts[0] for t in subs], self.n)

    def sets(self, names: Sequence[str]) -> "SetBlock":
        located = [self.find(nm) for nm in names]
        if not all(isinstance(b, SetBlock) for b, _ in located):
            raise ValueError("not set-valued columns")
        first = located[0][0]
        return SetBlock(list(names), first.type_name, [b.offsets[i] for b, i in located],
                        [b.codes[i] for b, i in located], [b.dicts[i] for b, i in located], self.n)

    def texts(self, names: Sequence[str]) -> "TextBlock":
        located = [self.find(nm) for nm in names]
        if not all(isinstance(b, TextBlock) for b, _ in located):
            raise ValueError("not raw text columns")
        first = located[0][0]
        if all(b is first for b, _ in located):
            return first.rows([i for _, i in located])
        return TextBlock(list(names), first.type_name, [b.data[i] for b, i in located],
                         [b.offsets[i] for b, i in located], [b.valid[i] for b, i in located], self.n)

    def numeric(self, names: Sequence[str]) -> NumericBlock:
        """The block holding exactly these columns, in this order (zero-copy when they are a run of one block)."""
        located = [self.find(nm) for nm in names]
        first = located[0][0]
        if all(b is first for b, _ in located):
            return first.rows([i for _, i in located])
        subs = [b.rows([i]) for b, i in located]
        ld = max(s.ld for s in subs)
        if any(s.ld != ld or s.type_name != subs[0].type_name for s in subs):
            raise ValueError("columns of different layouts cannot form one block")
        vals = torch.cat([s.values for s in subs], dim=0)
        valid = None
        if any(s.valid is not None for s in subs):
            full = torch.full((1, (self.n + 63) // 64 or 1), -1, dtype=torch.int64, device=vals.device)
            valid = torch.cat([s.valid if s.valid is not None else full for s in subs], dim=0)
        return NumericBlock(list(names), subs[0].type_name, vals, valid, self.n)

    def label(self, name: Optional[str] = None) -> torch.Tensor:
        nm = name or self.response
        if nm is None:
            raise KeyError("the batch has no response column")
        b, i = self.find(nm)
        if b.type_name != F.REALNN or b.valid is not None:
            raise ValueError("Sanity checker input missing label for row")   # SanityChecker.scala:391-393
        return b.values[i]

    def with_vector(self, v: VectorColumn) -> "ColumnarBatch":
        out = ColumnarBatch(self.n, self.blocks, dict(self.vectors), self.response, self.token)   # same raw rows
        out.vectors[v.name] = v
        out.__dict__["_dev_cache"] = self.__dict__.setdefault("_dev_cache", {})   # same raw blocks, same copies
        return out

    def on_device(self, kind: str, names: Sequence[str], block):
        """`block` (what strings / sets / texts returned for `names`) in HBM: itself when it is there already, else
        ONE host-to-device copy shared by every stage that works on these columns of this batch (an estimator's fit
        and its model's transform read the same rows; a second copy over PCIe costs more than both kernels)."""
        if block.is_cuda:
            return block
        cache = self.__dict__.setdefault("_dev_cache", {})
        key = (kind, tuple(names))
        if key not in cache:
            cache[key] = block.to("cuda")
        return cache[key]

    def device_thunk(self, kind: str, names: Sequence[str], block):
        """on_device, deferred to the first call (lazy parts move their input only when they are read)."""
        return lambda: self.on_device(kind, names, block)

    def take_rows(self, idx: torch.Tensor) -> "ColumnarBatch":
        """The raw columns of the rows `idx` (int64 tensor on the blocks' device) as a batch of their own (what
        Dataset.sample leaves); vector columns are not carried over."""
        return ColumnarBatch(int(idx.numel()), [b.take_rows(idx) for b in self.blocks], {}, self.response)

    def to(self, device) -> "ColumnarBatch":
        return ColumnarBatch(self.n, [b.to(device) for b in self.blocks], dict(self.vectors), self.response)

    @staticmethod
    def from_columns(columns: Dict[str, Tuple[str, Sequence]], response: Optional[str] = None, pin: bool = True
                     ) -> "ColumnarBatch":
        """{name: (type_name, python values with None for null)} -> one block per numeric type, in first-seen order
        (the TestFeatureBuilder of this package)."""
        n = None
        groups: Dict[str, List[str]] = {}
        for name, (tp, vals) in columns.items():
            if n is None:
                n = len(vals)
            groups.setdefault(tp, []).append(name)
        blocks = []
        for tp, names in groups.items():
            if tp in F.NUMERIC_TYPES:
                blocks.append(NumericBlock.from_lists(names, tp, [columns[nm][1] for nm in names], pin))
            elif tp == F.PICKLIST:
                blocks.append(DictBlock.from_lists(names, tp, [columns[nm][1] for nm in names], pin))
            elif tp in (F.TEXT, F.TEXTAREA):
                blocks.append(TextBlock.from_lists(names, tp, [columns[nm][1] for nm in names], pin))
            elif tp == F.MULTIPICKLIST:
                blocks.append(SetBlock.from_lists(names, tp, [columns[nm][1] for nm in names], pin))
            else:
                raise NotImplementedError(f"columns of type {tp} are not supported yet")
        return ColumnarBatch(n or 0, blocks, {}, response)
