"""Synthetic workloads of BASELINE.json's configs, generated on the device with seeded torch RNG.

Each make_* returns (ColumnarBatch resident in HBM, predictor features, label feature, info dict with the
algorithmic bytes per row).  Used by bench.py and tools/; no part of the product imports this.
"""
from __future__ import annotations

from typing import List

import numpy as np
import torch

from transmogrifai_b200 import features as F
from transmogrifai_b200.columnar import ColumnarBatch, DictBlock, NumericBlock, TextBlock

_W64 = None


def _pack_bits(ok: torch.Tensor) -> torch.Tensor:
    """bool [rows (multiple of 64)] -> int64 bitmap words"""
    global _W64
    if _W64 is None or _W64.device != ok.device:
        _W64 = (torch.ones(64, dtype=torch.int64, device=ok.device) << torch.arange(64, device=ok.device)).view(1, 64)
    return (ok.view(-1, 64).long() * _W64).sum(dim=1)


def _label(g, ld: int) -> torch.Tensor:
    return (torch.rand(ld, device="cuda", generator=g) < 0.3).double()


def real_block(g, rows: int, ncols: int, y: torch.Tensor, prefix: str = "c") -> NumericBlock:
    """N(j, 1 + j/10) columns, 5 % nulls, 5 label-shifted columns, one constant and one duplicated column."""
    ld = y.numel()
    nw = ld // 64
    vals = torch.empty((ncols, ld), dtype=torch.float64, device="cuda")
    valid = torch.empty((ncols, nw), dtype=torch.int64, device="cuda")
    for j in range(ncols):
        vals[j].normal_(float(j), 1.0 + j / 10.0, generator=g)
        valid[j] = _pack_bits(torch.rand(ld, device="cuda", generator=g) >= 0.05)
    for j in range(0, min(ncols, 50), 10):
        vals[j] += 0.5 * y
    if ncols > 12:
        vals[7] = 3.25
        vals[12] = vals[11]
        valid[12] = valid[11]
    return NumericBlock([f"{prefix}{j:03d}" for j in range(ncols)], F.REAL, vals, valid, rows)


def binary_block(g, rows: int, ncols: int, y: torch.Tensor) -> NumericBlock:
    ld = y.numel()
    vals = torch.empty((ncols, ld), dtype=torch.uint8, device="cuda")
    valid = torch.empty((ncols, ld // 64), dtype=torch.int64, device="cuda")
    for j in range(ncols):
        p = 0.1 + 0.8 * j / max(ncols - 1, 1)
        vals[j] = (torch.rand(ld, device="cuda", generator=g) < p).to(torch.uint8)
        valid[j] = _pack_bits(torch.rand(ld, device="cuda", generator=g) >= 0.05)
    return NumericBlock([f"b{j:03d}" for j in range(ncols)], F.BINARY, vals, valid, rows)


def picklist_block(g, rows: int, ncols: int, y: torch.Tensor, max_card: int = 100_000) -> DictBlock:
    """Cardinalities log-spaced 10 .. max_card, ranks Zipf(1.1), 5 % nulls; dictionary entry r of column j is
    "v<j>_<r>" (the underscore is \\p{Punct}: cleaning changes every value), plus two aliases per column that clean to
    the same value as ranks 0 and 1; the first three columns leak the label."""
    ld = y.numel()
    cards = np.round(np.logspace(1, np.log10(max_card), ncols)).astype(np.int64) if ncols > 1 else np.array([max_card])
    codes = torch.empty((ncols, ld), dtype=torch.int32, device="cuda")
    dicts: List[List[str]] = []
    for j, k in enumerate(cards):
        p = 1.0 / torch.arange(1, k + 1, dtype=torch.float64, device="cuda") ** 1.1
        cdf = torch.cumsum(p / p.sum(), 0)
        u = torch.rand(ld, dtype=torch.float64, device="cuda", generator=g)
        if j < 3:                                       # leaky: positives are pushed towards the head of the list
            u = torch.where(y > 0, u * 0.5, u)
        c = torch.searchsorted(cdf, u).clamp_(max=int(k) - 1).int()
        alias = torch.rand(ld, device="cuda", generator=g) < 0.02
        c = torch.where(alias & (c < 2), c + int(k), c)  # codes k, k+1 = aliases of ranks 0, 1
        null = torch.rand(ld, device="cuda", generator=g) < 0.05
        codes[j] = torch.where(null, torch.full_like(c, -1), c)
        dicts.append([f"v{j}_{r}" for r in range(int(k))] + [f"V{j} 0", f"v{j}.1!"])
    return DictBlock([f"p{j:03d}" for j in range(ncols)], F.PICKLIST, codes, dicts, rows)


_VOCAB = None


def _vocab(n_words: int = 200_000):
    """[a-z]{3,10} words that are not stop words (none of the 174 stop words can appear: they are re-drawn)."""
    global _VOCAB
    if _VOCAB is None:
        rng = np.random.default_rng(20260003)
        wl = rng.integers(3, 11, n_words)
        off = np.concatenate([[0], np.cumsum(wl)])
        by = rng.integers(97, 123, off[-1]).astype(np.uint8)
        stop = {"the", "and", "you", "for", "was", "are", "his", "her", "its", "our", "who", "all", "any", "few", "not",
                "nor", "own", "too", "did", "had", "has", "him", "she", "off", "out", "how", "why", "but", "can't"}
        for i in np.nonzero(wl <= 3)[0]:
            while by[off[i]:off[i + 1]].tobytes().decode() in stop:
                by[off[i]:off[i + 1]] = rng.integers(97, 123, wl[i])
        _VOCAB = (torch.from_numpy(by).cuda(), torch.from_numpy(off).cuda(), torch.from_numpy(wl).cuda())
    return _VOCAB


def text_block(g, rows: int, ncols: int, mean_tokens: float = 30.0) -> TextBlock:
    """Cells of Poisson(mean_tokens) words drawn Zipf(1.05) from a 200 000-word vocabulary, single-space separated,
    5 % nulls (the 1 % mixed-case / punctuation variants are exercised by the parity tests instead)."""
    wbytes, woff, wl = _vocab()
    v = wl.numel()
    p = 1.0 / torch.arange(1, v + 1, dtype=torch.float64, device="cuda") ** 1.05
    cdf = torch.cumsum(p / p.sum(), 0)
    data, offsets, valid = [], [], []
    nw = (rows + 63) // 64
    for _ in range(ncols):
        ntok = torch.poisson(torch.full((rows,), mean_tokens, device="cuda"), generator=g).long()
        null = torch.rand(rows, device="cuda", generator=g) < 0.05
        ntok = torch.where(null, torch.zeros_like(ntok), ntok)
        t = int(ntok.sum().item())
        w = torch.searchsorted(cdf, torch.rand(t, dtype=torch.float64, device="cuda", generator=g)).clamp_(max=v - 1)
        tl = wl[w] + 1                                  # word + separator
        cs = torch.cumsum(tl, 0)
        tstart = cs - tl
        last = torch.cumsum(ntok, 0) - 1
        ends = torch.where(ntok > 0, cs[last.clamp(min=0)], torch.zeros_like(last))
        off = torch.zeros(rows + 1, dtype=torch.int64, device="cuda")
        off[1:] = torch.cummax(ends, 0).values
        buf = torch.full((max(int(cs[-1].item()) if t else 0, 1),), 32, dtype=torch.uint8, device="cuda")
        lens = wl[w]
        rep = torch.repeat_interleave(torch.arange(t, device="cuda"), lens)
        inner = torch.arange(rep.numel(), device="cuda") - torch.repeat_interleave(torch.cumsum(lens, 0) - lens, lens)
        buf[tstart[rep] + inner] = wbytes[woff[w][rep] + inner]
        ok = torch.zeros(nw * 64, dtype=torch.bool, device="cuda")
        ok[:rows] = ~null
        data.append(buf)
        offsets.append(off)
        valid.append(_pack_bits(ok))
        del rep, inner, w, tl, cs, tstart
    return TextBlock([f"t{j:03d}" for j in range(ncols)], F.TEXT, data, offsets, valid, rows)


def _features(block) -> List[F.Feature]:
    return [F.Feature(nm, block.type_name, False) for nm in block.names]


def _ld(rows: int) -> int:
    return (rows + 2047) // 2048 * 2048


def text_bytes_per_row(block: TextBlock) -> float:
    return sum(int(o[-1].item()) for o in block.offsets) / max(block.n, 1)


WORKLOADS = ("c2", "c3", "c4", "c5")
DEFAULT_ROWS = {"c2": 10_000_000, "c3": 50_000_000, "c4": 5_000_000, "c5": 4_000_000}


def make_workload(name: str, rows: int, seed: int):
    """-> (batch, predictor features, label feature, info).  info: description, sanity-checker settings, algorithmic
    bytes per row of the fit scan and of the transform, SmartTextVectorizer numFeatures."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    ld = _ld(rows)
    y = _label(g, ld)
    label_block = NumericBlock(["label"], F.REALNN, y.view(1, ld), None, rows)
    label = F.Feature("label", F.REALNN, True)
    blocks, feats = [], []
    info = {"num_features": 512}
    if name == "c2":
        rb = real_block(g, rows, 100, y)
        blocks, feats = [rb], _features(rb)
        info.update(
            desc=f"C2: synthetic {rows} rows x 100 Real columns (5% nulls) per GPU: RealVectorizer mean-fill + null "
                 "indicators fused with SanityChecker (column moments, label correlations, Cramer's V of the null "
                 "indicators, removeBadFeatures decision, featureFeatureCorrLevel=Off), then the transform writing "
                 "the kept columns",
            fit_bytes=100 * 8 + 100 / 8 + 8, transform_bytes=100 * 8 + 100 / 8 + 100 * 16)
    elif name == "c3":
        pb = picklist_block(g, rows, 50, y)
        blocks, feats = [pb], _features(pb)
        info.update(
            desc=f"C3: synthetic {rows} rows x 50 PickList columns (cardinality 10..100k, Zipf 1.1, 5% nulls, "
                 "dictionary-encoded int32 codes) per GPU: OpTextPivotVectorizer top-K pivot / one-hot fused with "
                 "SanityChecker (count[code][label] -> vocabulary, contingency, Cramer's V, moments, correlations), "
                 "then the one-hot transform of the kept columns",
            fit_bytes=50 * 4 + 1 + 8, transform_bytes=50 * 4 + 50 * 4)
    elif name == "c4":
        tb = text_block(g, rows, 20)
        blocks, feats = [tb], _features(tb)
        bpr = text_bytes_per_row(tb)
        info.update(
            num_features=1 << 18,
            desc=f"C4 (scaled to {rows} rows per GPU; the config names 20M): {rows} rows x 20 Text columns (~30 tokens "
                 "per cell, 5% nulls): SmartTextVectorizer -> tokenise + Spark-legacy murmur3 into 2^18 shared buckets + "
                 "null indicators fused with SanityChecker bucket statistics, then the sparse-row transform of the "
                 "kept columns",
            fit_bytes=bpr + 20 * 8 + 20 / 8 + 1 + 8, transform_bytes=bpr + 20 * 8 + 20 / 8)
    elif name == "c5":
        bb, pb = binary_block(g, rows, 20, y), picklist_block(g, rows, 50, y)
        rb, tb = real_block(g, rows, 100, y), text_block(g, rows, 30)
        blocks = [bb, pb, rb, tb]
        feats = _features(bb) + _features(pb) + _features(rb) + _features(tb)
        bpr = text_bytes_per_row(tb)
        info.update(
            desc=f"C5 shape (scaled to {rows} rows per GPU; the config names 100M): 100 Real + 50 PickList + 30 Text "
                 "(512 hashes, separate spaces) + 20 Binary columns: full transmogrify() + sanityCheck fit + transform",
            fit_bytes=(100 * 8 + 100 / 8) + (50 * 4) + (bpr + 30 * 8 + 30 / 8) + 20 * (1 + 1 / 8) + 8 + 1,
            transform_bytes=(100 * 8 + 100 / 8 + 100 * 16) + (50 * 4 + 50 * 4) + (bpr + 30 * 8) + 20 * (1 + 1 / 8 + 16))
    else:
        raise ValueError(f"unknown workload {name}")
    batch = ColumnarBatch(rows, blocks + [label_block], {}, response="label")
    return batch, feats, label, info


def host_copy(batch: ColumnarBatch) -> ColumnarBatch:
    """The same batch in pinned host memory (what the columnar export of a Spark partition hands over)."""
    def pin(t):
        if t is None:
            return None
        h = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
        h.copy_(t)
        return h
    out = []
    for b in batch.blocks:
        if isinstance(b, NumericBlock):
            out.append(NumericBlock(b.names, b.type_name, pin(b.values), pin(b.valid), b.n))
        elif isinstance(b, DictBlock):
            out.append(DictBlock(b.names, b.type_name, pin(b.codes), b.dicts, b.n))
        else:
            out.append(TextBlock(b.names, b.type_name, [pin(t) for t in b.data], [pin(t) for t in b.offsets],
                                 [pin(t) for t in b.valid], b.n))
    return ColumnarBatch(batch.n, out, {}, batch.response)


def text_bytes(batch: ColumnarBatch) -> int:
    """Bytes of the text blocks of a chunk (data, offsets, validity)."""
    tot = 0
    for b in batch.blocks:
        if not isinstance(b, (NumericBlock, DictBlock)):
            tot += sum(t.numel() for t in b.data) + sum(t.numel() * 8 for t in b.offsets) + \
                sum(0 if t is None else t.numel() * 8 for t in b.valid)
    return tot


def input_bytes(batch: ColumnarBatch) -> int:
    tot = 0
    for b in batch.blocks:
        if isinstance(b, NumericBlock):
            tot += b.values.numel() * b.values.element_size() + (0 if b.valid is None else b.valid.numel() * 8)
        elif isinstance(b, DictBlock):
            tot += b.codes.numel() * 4
        else:
            tot += sum(t.numel() for t in b.data) + sum(t.numel() * 8 for t in b.offsets) + \
                sum(0 if t is None else t.numel() * 8 for t in b.valid)
    return tot
