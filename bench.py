#!/usr/bin/env python
"""bench.py -- rows/sec of transmogrify() + sanityCheck() fit + transform on synthetic data (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c5|c2|c3|c4] [--rows R] [--impl tmog|reference]

The default workload is the one BASELINE.json quotes its metric on: config 5, 100 M rows x 200 mixed columns,
streamed in row chunks (transmogrifai_b200/streaming.py): the rows are split over the N ranks (strong scaling), every
rank folds its chunks (pass 0 label, pass 1 fused fit + statistics scans), the accumulators are all-reduced once (NCCL),
the SanityChecker decision is taken, and pass 2 transforms chunk by chunk.  100 M rows are ~740 GB of raw columns, so
the chunks are cycled from a ring of distinct synthetic chunks resident in HBM (`config.distinct_chunks`).  c2 / c3 / c4
run one resident batch (weak scaling: R rows per GPU); at N = 1 the default run appends their lines as `others`.

A "step" is one full pass of the hot path over the workload: fit of every stage, the drop decision, and the transform
that writes the kept columns.  Prints ONE JSON line (see DESIGN.md "Measurement").
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SEED = 20260001
METRIC = "rows/sec transmogrify+sanityCheck fit+transform"
METADATA_NOTE = ("schema-level metadata (interned column descriptions, the SanityChecker parent / group plan of a vector "
                 "layout) is kept between steps; label classes, count tables, dictionaries' cleaning, statistics, "
                 "decisions and every device buffer are recomputed each step")


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons during the timed region (NVML, ~20 Hz)."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                     "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                     "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                     "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
            while not self._stop_evt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
                time.sleep(0.05)
        except Exception as e:  # pragma: no cover
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}




ALL_CPUS = os.sched_getaffinity(0)


def bind_to_gpu_cpus(index: int):
    """Run this process on the CPUs NVML reports as local to its GPU (the NUMA node behind the GPU's PCIe root), so that
    the pinned host buffers it allocates afterwards -- first touch -- and the threads that fill them sit on that node:
    with several ranks on one box each GPU then pulls its chunks over its own root complex instead of all of them
    through node 0.  -> the CPU list, or None when NVML gives none."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(index)
        ncpu = os.cpu_count() or 1
        words = nv.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = [64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1 and 64 * w + b < ncpu]
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if allowed:
            os.sched_setaffinity(0, allowed)
            return allowed
    except Exception:
        pass
    return None


def build_plan(feats, label, info, reference_defaults: bool = False):
    from transmogrifai_b200.dsl import TransmogrifierDefaults, transmogrify

    class Defaults(TransmogrifierDefaults):
        DefaultNumOfFeatures = info["num_features"]

    vec = transmogrify(feats, defaults=Defaults)
    if reference_defaults:
        # every SanityChecker default of the reference (SanityChecker.scala:561-581): above 1 M rows a sample of 1 M rows
        # is checked, and the full feature-feature Pearson matrix is computed (the Gramian on the fp64 tensor cores)
        return label.sanityCheck(vec, removeBadFeatures=True, sampleSeed=SEED)
    return label.sanityCheck(vec, removeBadFeatures=True, checkSample=1.0, sampleLowerLimit=0,
                             sampleUpperLimit=2 ** 31 - 1, featureFeatureCorrLevel="Off")


def run_step(batch, checked):
    """One step over one resident batch: fit every stage (fused vectorizer + SanityChecker scans, label classes,
    decision) and materialise the kept columns of the output vector.  The per-batch caches are dropped; the library's
    content-keyed, data-independent metadata caches survive steps (METADATA_NOTE, config.metadata_caches), as they do
    in the chunk-streamed step."""
    from transmogrifai_b200.dsl import fit_and_transform
    for k in ("_label_cache", "_label_dev", "_dict_cache", "_dev_cache"):
        batch.__dict__.pop(k, None)
    out, fitted = fit_and_transform(batch, checked)
    vec = out.vectors[checked.name]
    parts = [p.materialize(None) if hasattr(p, "materialize") else p for p in vec.parts]   # lazy until read: read it
    for p in parts:
        if hasattr(p, "data") and not hasattr(p, "materialize"):
            _ = p.data
    return vec, fitted, parts


def run_stream_step(chunks, feats, label, info, sink=None, cache_bytes=0, cell_cache_bytes=0):
    """One step over a chunked dataset: a fresh plan (new stage objects), passes 0 / 1, decision, pass 2.
    cell_cache_bytes: HBM in which pass 1 leaves the packed text cells of the chunks for pass 2 (allocated, filled and
    released inside the step)."""
    from transmogrifai_b200.streaming import fit_and_transform_stream
    checked = build_plan(feats, label, info)
    res = fit_and_transform_stream(chunks, checked, sink, cache_bytes=cache_bytes, cell_cache_bytes=cell_cache_bytes)
    return checked, res


def cell_cache_budget(arg_gb: float, reserve_gb: float) -> int:
    """Bytes of HBM for the packed cells: what is free now minus a reserve for the step's own buffers."""
    import torch
    if arg_gb >= 0:
        return int(arg_gb * 2 ** 30)
    torch.cuda.empty_cache()
    return max(0, int(torch.cuda.mem_get_info()[0] - reserve_gb * 2 ** 30))


class CountingSink:
    """Pass-2 sink of the device-timed step: the chunk's output stays in HBM and is dropped (its buffers are reused
    by the next chunk); only the bytes are counted."""

    def __init__(self):
        self.bytes = 0

    def __call__(self, k, vec, parts):
        self.bytes += output_bytes(parts)


def output_bytes(parts) -> int:
    tot = 0
    for p in parts:
        for name in ("data", "indices", "values", "row_start", "row_nnz"):
            t = getattr(p, name, None)
            if t is not None and hasattr(t, "numel"):
                tot += t.numel() * t.element_size()
    return tot


def _time(f, reps):
    import torch
    for _ in range(2):
        f()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        f()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def kernel_times(workload, batch, fitted, info, reps: int = 5, text_offset: int = 0):
    """Device time of the dominant kernels of the step on this very batch (CUDA events on the launch stream), with the
    algorithmic bytes each one moves per row.  -> list of dicts, the slowest first."""
    import ctypes as C
    import torch
    from transmogrifai_b200 import _lib as L
    from transmogrifai_b200 import runtime as R
    from transmogrifai_b200.columnar import DictBlock, NumericBlock, TextBlock
    st = R.stream()
    n = batch.n
    out = []
    for blk in batch.blocks:
        if isinstance(blk, NumericBlock) and blk.type_name == "Real":
            ncols = len(blk.names)
            y = batch.label()
            nb = C.c_size_t()
            L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nb))
            scratch = R.scratch("real_fit", nb.value)
            shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
            part = torch.zeros((ncols + 1) * 8, dtype=torch.float64, device="cuda")
            fills = torch.zeros(ncols, dtype=torch.float64, device="cuda")
            keep_local = [i for i in range(2 * ncols) if i != 14]          # what SanityChecker keeps of this block
            bits = np.zeros((2 * ncols + 63) // 64 * 64, dtype=np.uint8)
            bits[keep_local] = 1
            mask = torch.from_numpy(np.packbits(bits, bitorder="little").view("<i8").copy()).cuda()
            dst = torch.empty((n, len(keep_local)), dtype=torch.float64, device="cuda")
            L.call("tmog_real_pilot_shift_dev", blk.values.data_ptr(), blk.ld, blk.valid.data_ptr(), blk.ldv,
                   y.data_ptr(), n, ncols, shifts.data_ptr(), st)
            t_fit = _time(lambda: L.call("tmog_real_fit_partials_dev", blk.values.data_ptr(), blk.ld,
                                         blk.valid.data_ptr(), blk.ldv, y.data_ptr(), n, ncols, shifts.data_ptr(),
                                         scratch.data_ptr(), scratch.numel(), part.data_ptr(), st), reps)
            t_tr = _time(lambda: L.call("tmog_real_transform_select_dev", blk.values.data_ptr(), blk.ld,
                                        blk.valid.data_ptr(), blk.ldv, n, ncols, fills.data_ptr(), 1, mask.data_ptr(),
                                        len(keep_local), dst.data_ptr(), len(keep_local), 0, st), reps)
            out.append({"kernel": "vec_transform_pack_kernel<double, true>", "ms": t_tr,
                        "bytes_per_row": ncols * 8 + ncols / 8 + len(keep_local) * 8})
            out.append({"kernel": "real_fit_tma_kernel", "ms": t_fit, "bytes_per_row": ncols * 8 + ncols / 8 + 8})
            del dst
        elif isinstance(blk, DictBlock):
            ncols = len(blk.names)
            ncodes = np.array([len(d) for d in blk.dicts], dtype=np.int64)
            nids, ncls = ncodes + 1, 2
            zeros = np.zeros(ncols, dtype=np.int64)
            coff = np.concatenate([[0], np.cumsum(nids * ncls)[:-1]]).astype(np.int64)
            soff = np.concatenate([[0], np.cumsum(ncodes)[:-1]]).astype(np.int64)
            desc = torch.from_numpy(np.ascontiguousarray(np.stack([nids, soff, coff, ncodes], axis=1))).cuda()
            counts = torch.zeros(int((nids * ncls).sum()), dtype=torch.int32, device="cuda")
            bad = torch.zeros(1, dtype=torch.int32, device="cuda")
            ids = (batch.label()[:max(n, 4)] > 0).to(torch.uint8)
            pos = torch.zeros(int(ncodes.sum()) + ncols, dtype=torch.int32, device="cuda")
            dst = torch.empty((n, ncols), dtype=torch.int32, device="cuda")
            t_cnt = _time(lambda: L.call("tmog_pivot_count_dev", blk.codes.data_ptr(), blk.ld, n, ncols, ids.data_ptr(),
                                         ncls, desc.data_ptr(), None, int(nids.max()), counts.data_ptr(),
                                         bad.data_ptr(), st), reps)
            t_tr = _time(lambda: L.call("tmog_pivot_transform_dev", blk.codes.data_ptr(), blk.ld, n, ncols,
                                        desc.data_ptr(), pos.data_ptr(), dst.data_ptr(), ncols, 0, st), reps)
            out.append({"kernel": "pivot_transform_kernel", "ms": t_tr, "bytes_per_row": ncols * 8})
            out.append({"kernel": "pivot_count_kernel", "ms": t_cnt, "bytes_per_row": ncols * 4 + 1})
            del dst
        elif isinstance(blk, TextBlock):
            model = [m for m in fitted if hasattr(m, "vectorizationMethods")][0]
            lay = model.layout
            ids = (batch.label()[:max(n, 4)] > 0).to(torch.uint8)
            dev = batch.texts([f.name for f in model.inN])
            sst = model._side_begin(np.array([0.0, 1.0]), False)
            t_s = _time(lambda: model._side_scan(sst, dev, n, ids, None), max(reps // 2, 1))
            keep_all = fitted[-1]._keep
            # the text part's kept columns, local to the part
            off = text_offset
            keep_local = keep_all[(keep_all >= off + model.pivot_width) & (keep_all < off + model.pivot_width + lay.width)] \
                - off - model.pivot_width
            flags = torch.zeros(4, dtype=torch.int32, device="cuda")
            stream = {"index": 0, "token": batch.token, "flags": flags, "flags_clean": flags}
            holder = {}

            def tr():
                holder["p"] = model._materialize_hash(dev, n, keep_local, None, stream)
            t_t = _time(tr, max(reps // 2, 1))
            nnz = int(holder["p"].indices.numel())
            bpr = sum(int(dev.offsets[j][-1].item()) for j in range(len(dev.names))) / max(n, 1)
            k = len(lay.text_idx)
            out.append({"kernel": "text tokenise + hash: sparse rows of the kept columns (pass 2)", "ms": t_t,
                        "bytes_per_row": bpr + k * 8 + k / 8 + 12 + 8.0 * nnz / max(n, 1),
                        "note": "instruction-bound, reported against HBM anyway"})
            out.append({"kernel": "text tokenise + hash: bucket statistics (pass 1)", "ms": t_s,
                        "bytes_per_row": bpr + k * 8 + k / 8 + 1,
                        "note": "instruction-bound, reported against HBM anyway"})
            # the same two passes with packed cells: pass 1 also leaves every cell's buckets in HBM, pass 2 reads them
            if model.enable_cell_cache(int(4 * 2 ** 30 * max(n, 1) / 1e6) + (64 << 20)):
                cells = model._cells
                sst["pack_token"] = (0, batch.token)

                def scan_pack():
                    cells["cursor"].zero_()
                    cells["pending"].clear()
                    model._side_scan(sst, dev, n, ids, None)
                t_sp = _time(scan_pack, max(reps // 2, 1))
                token, meta, pflag, _n, kk = cells["pending"][-1]
                if int(pflag.item()) == 0:
                    cells["chunks"][token] = (meta, n, kk)
                    packed_bytes = int(cells["cursor"].item()) + meta.numel()
                    t_tp = _time(tr, max(reps // 2, 1))
                    out.append({"kernel": "text statistics + packed cells (pass 1 when the cells are kept)", "ms": t_sp,
                                "bytes_per_row": bpr + k * 8 + k / 8 + 1 + packed_bytes / max(n, 1),
                                "note": "instruction-bound, reported against HBM anyway"})
                    out.append({"kernel": "text_rows_packed_kernel: sparse rows from packed cells (pass 2)", "ms": t_tp,
                                "bytes_per_row": packed_bytes / max(n, 1) + 12 + 8.0 * nnz / max(n, 1)})
                model.release_cell_cache()
            del holder
    return sorted(out, key=lambda d: -d["ms"])


# DRAM bytes per row of the config-2 kernels, from `ncu --set full` (profiles/r01_c2_full_raw.csv: 10 M rows, 199 of
# 200 columns kept): pack transform 8.125 GB read + 15.862 GB written, fit scan 8.286 GB read + 0.004 GB written
NCU_DRAM_BYTES_PER_ROW = {"vec_transform_pack_kernel<double, true>": 2398.8, "real_fit_tma_kernel": 829.0}
# config 5's text calls, from profiles/r02_c5_text_raw.csv (1 M rows x 30 text columns): span kernel 7.02 GB read +
# 3.81 GB of records written, text_rows_kernel 3.91 GB read + 5.76 GB written; fit + pack 7.02 + 0.84 GB; packed rows
# 2.00 + 5.75 GB; fit without packing 7.04 GB read (r02_span captures at 500 k rows)
NCU_DRAM_BYTES_PER_ROW_C5 = {
    "text tokenise + hash: sparse rows of the kept columns (pass 2)": 20503.0,
    "text statistics + packed cells (pass 1 when the cells are kept)": 7860.0,
    "text tokenise + hash: bucket statistics (pass 1)": 7050.0,
    "text_rows_packed_kernel: sparse rows from packed cells (pass 2)": 7744.0,
}


def traffic_gb(workload: str, kernel: str, rows: int):
    """Measured DRAM traffic of one launch of `kernel` in GB (None where no capture exists for the workload)."""
    per_row = NCU_DRAM_BYTES_PER_ROW.get(kernel) if workload == "c2" else \
        NCU_DRAM_BYTES_PER_ROW_C5.get(kernel) if workload == "c5" else None
    return None if per_row is None else per_row * rows / 1e9


_C2_SAMPLE: dict = {}


def cpu_reference_c2(rows: int, threads: int):
    """The reference's CPU job sequence for config 2 (oracle/c: C + OpenMP) on `rows` rows -> seconds."""
    from oracle import c_port as CP
    from tests.util_data import make_real_config, pack_validity
    if _C2_SAMPLE.get("rows") != rows:              # drawn once per process, before any clock starts
        x, valid, y = make_real_config(rows, 100, seed=SEED)
        _C2_SAMPLE.update(rows=rows, x=np.ascontiguousarray(x), valid=valid, vw=pack_validity(valid), y=y)
    x, valid, vw, y = (_C2_SAMPLE[k] for k in ("x", "valid", "vw", "y"))
    CP.real_pipeline(x[:, :2048].copy(), pack_validity(valid[:, :2048]), y[:2048].copy(), threads)   # page in
    t0 = time.perf_counter()
    fills, rows_m, mean, var, mn, mx, corr = CP.real_pipeline(x, vw, y, threads)
    CP.label_sums(rows_m[:, :-1], y, np.array([0.0, 1.0]), threads)      # categoricalTests' reduceByKey
    return time.perf_counter() - t0


_CPU_SAMPLE: dict = {}


def cpu_sample_generic(workload: str, rows: int):
    """A bounded sample of the synthetic workload as host numpy arrays (drawn once per process, before any clock)."""
    key = (workload, rows)
    if _CPU_SAMPLE.get("key") == key:
        return _CPU_SAMPLE["work"], _CPU_SAMPLE["y"], _CPU_SAMPLE["info"]
    import torch
    import benchdata
    from transmogrifai_b200.columnar import DictBlock, NumericBlock, TextBlock
    batch, feats, label, info = benchdata.make_workload(workload, rows, SEED)
    y = np.ascontiguousarray(batch.label()[:rows].cpu().numpy())
    work = []
    for b in batch.blocks:
        if isinstance(b, NumericBlock):
            if b.names == ["label"]:
                continue
            v = np.ascontiguousarray(b.values[:, :rows].cpu().numpy().astype(np.float64))
            vw = None if b.valid is None else np.ascontiguousarray(
                b.valid[:, :(rows + 63) // 64].cpu().numpy().view(np.uint64))
            work.append(("num", b.type_name, v, vw))
        elif isinstance(b, DictBlock):
            work.append(("dict", np.ascontiguousarray(b.codes[:, :rows].cpu().numpy()), b.dicts))
        elif isinstance(b, TextBlock):
            cols = []
            for j in range(len(b.names)):
                offs = b.offsets[j][:rows + 1].cpu().numpy().astype(np.int64)
                data = b.data[j][:max(int(offs[-1]), 1)].cpu().numpy()
                valid = None if b.valid[j] is None else b.valid[j][:(rows + 63) // 64].cpu().numpy().view(np.uint64)
                cols.append((np.ascontiguousarray(data), np.ascontiguousarray(offs),
                             None if valid is None else np.ascontiguousarray(valid)))
            work.append(("text", list(b.names), cols))
    del batch
    torch.cuda.empty_cache()
    _CPU_SAMPLE.update(key=key, work=work, y=y, info=info)
    return work, y, info


def cpu_reference_generic(workload: str, rows: int, threads: int):
    """The reference's job sequence on a sample of the same synthetic workload, as oracle/c (C + OpenMP, row
    partitions like Spark local[N]): per type group the vectorizer's fit and transform, then SanityChecker's sums
    (column moments, label covariance, per-label column sums).  The string work is done once per dictionary entry and
    the hashed columns' statistics are taken in the hashing pass -- both cheaper than what the reference does per
    row, so this baseline errs on the fast side.  -> seconds (the sample is drawn before the clock starts)."""
    from oracle import c_port as CP
    from oracle import vectorizers as V
    from transmogrifai_b200.stages.smart_text import murmur3_index
    work, y, info = cpu_sample_generic(workload, rows)
    cls = (y > 0).astype(np.uint8)
    labels = np.array([0.0, 1.0])
    t0 = time.perf_counter()
    for item in work:
        if item[0] == "num":
            _, tp, v, vw = item
            if tp == "Real":
                _, rows_m, _, _, _, _, _ = CP.real_pipeline(v, vw, y, threads)
            else:                                           # Binary / RealNN: constant fill, then the same passes
                rows_m = CP.real_transform(v, vw, np.zeros(v.shape[0]), tp == "Binary", threads, extra_cols=1)
                rows_m[:, -1] = y
                mean, _, _, _ = CP.col_stats(rows_m, threads)
                CP.label_cov(rows_m, mean, threads)
            CP.label_sums(rows_m[:, :-1], y, labels, threads)
        elif item[0] == "dict":
            _, codes, dicts = item
            ncodes = np.array([len(d) for d in dicts])
            counts = CP.pivot_counts(codes, rows, ncodes, cls, 2, threads)
            tables, off = [], 0
            for c, d in enumerate(dicts):
                merged = {}
                cleaned = CP.clean_strings(d)
                tot = counts[c][1:].sum(axis=1)
                for k, cv in enumerate(cleaned):
                    merged[cv] = merged.get(cv, 0) + int(tot[k])
                top = V.top_values(merged, 20, 10)
                index = {v: i for i, v in enumerate(top)}
                tab = np.array([off + index.get(cv, len(top)) for cv in cleaned] + [off + len(top) + 1], dtype=np.int32)
                tables.append(tab)
                off += len(top) + 2
            CP.pivot_transform(codes, rows, tables, threads)     # one-hot rows; their sums are the counts above
        else:
            _, names, cols = item
            CP.text_stats(cols, rows, True, 30, threads)
            nf = info["num_features"]
            shared = nf * len(names) > (1 << 17)
            prefixes = [f"{murmur3_index(nm, nf)}_" for nm in names]
            CP.text_hash(cols, rows, prefixes, nf, shared, class_ids=cls, nclasses=2, threads=threads)
    return time.perf_counter() - t0


CPU_ROWS = {"c3": 5_000_000, "c4": 500_000, "c5": 400_000}


def cpu_baseline(workload: str, cores: int, cpu_rows: int):
    if workload == "c2":
        secs = cpu_reference_c2(cpu_rows, cores)
        return secs, cpu_rows, cores, (f"{cpu_rows} rows x 100 Real columns of the same workload; the reference's job "
                                       f"sequence (fit, transform, colStats, label correlations, per-label sums) as "
                                       f"oracle/c (C + OpenMP) on {cores} threads")
    rows = CPU_ROWS[workload]
    secs = cpu_reference_generic(workload, rows, cores)
    return secs, rows, cores, (f"{rows} rows of the same workload (one pass: the CPU rate does not depend on the row "
                               f"count); the reference's job sequence per type group (fit, transform, SanityChecker "
                               f"sums) as oracle/c (C + OpenMP) on {cores} threads, string work per dictionary entry")


def make_ring(rows_per_chunk: int, distinct: int, seed: int):
    """`distinct` synthetic config-5 chunks resident in HBM; the PickList dictionaries are shared objects, as the
    columnar export of one DataFrame with a global dictionary delivers them."""
    import benchdata
    from transmogrifai_b200.columnar import DictBlock
    ring, feats, label, info = [], None, None, None
    for i in range(distinct):
        b, feats, label, info = benchdata.make_workload("c5", rows_per_chunk, seed + 7919 * i)
        if ring:
            for blk, first in zip(b.blocks, ring[0].blocks):
                if isinstance(blk, DictBlock):
                    assert all(a == c for a, c in zip(blk.dicts, first.dicts))
                    blk.dicts = first.dicts
        ring.append(b)
    return ring, feats, label, info


def rank_chunks(ring, total_rows: int, chunk_rows: int, rank: int, world: int):
    """The chunks of this rank's share of `total_rows` (contiguous share; the last chunk may be shorter), cycled from the
    ring."""
    import torch
    per = total_rows // world + (1 if rank < total_rows % world else 0)
    full, rem = per // chunk_rows, per % chunk_rows
    chunks = [ring[i % len(ring)] for i in range(full)]
    if rem:
        chunks.append(ring[full % len(ring)].take_rows(torch.arange(rem, device="cuda")))
    return chunks, per


def parity_check_stream(ring, feats, label, info, world: int):
    """Untimed self-check over the distinct chunks: the streamed fit against torch f64 reductions of the raw columns
    (fills, null counts, class counts, PickList group totals), the sparse rows pass 2 writes -- once from the packed
    cells, once from the text -- against the per-column sums pass 1 took (different kernels over the same text), and
    the same dropped set on every rank."""
    import torch
    from transmogrifai_b200.columnar import DictBlock, NumericBlock
    from transmogrifai_b200.streaming import fit_stream, transform_stream
    from transmogrifai_b200 import runtime as R
    R.clear_caches()
    checked = build_plan(feats, label, info)
    # pass 1 also packs the text cells (as the timed step does): pass 2 is then checked twice, from the packed cells and
    # from the text
    fit = fit_stream(ring, checked, cell_cache_bytes=int(sum(b.n for b in ring) * 1400) + (64 << 20))
    est = checked.origin_stage
    sm = est.summary
    names = sm.meta.columnNames()
    col = {nm: i for i, nm in enumerate(names)}
    n = sum(b.n for b in ring)
    problems = []
    # with several ranks the statistics are global: the local reductions below are all-reduced the same way
    def allsum(t):
        if world > 1:
            import torch.distributed as dist
            dist.all_reduce(t)
        return t
    n_all = int(allsum(torch.tensor([n], dtype=torch.int64, device="cuda")).item())
    if sm.count != n_all:
        problems.append(f"count {sm.count} != {n_all}")
    for bi, blk in enumerate(ring[0].blocks):
        if isinstance(blk, NumericBlock) and blk.type_name == "Real":
            ncols = len(blk.names)
            tot = torch.zeros(ncols, dtype=torch.float64, device="cuda")
            cnt = torch.zeros(ncols, dtype=torch.float64, device="cuda")
            for b in ring:
                rb = b.blocks[bi]
                bits = ((rb.valid[:, :, None] >> torch.arange(64, device="cuda")) & 1).reshape(ncols, -1)[:, :b.n].bool()
                v = rb.values[:, :b.n]
                tot += torch.where(bits, v, torch.zeros_like(v)).sum(dim=1)
                cnt += bits.sum(dim=1).double()
            tot, cnt = allsum(tot), allsum(cnt)
            mean = (tot / cnt).cpu().numpy()
            nulls = (n_all - cnt).cpu().numpy()
            vcol = {c.parentFeatureName[0]: i for i, c in enumerate(sm.meta.columns) if c.indicatorValue is None}
            ncol = {c.parentFeatureName[0]: i for i, c in enumerate(sm.meta.columns)
                    if c.indicatorValue == "NullIndicatorValue"}
            fills = np.array([sm.mean[vcol[nm]] for nm in blk.names])
            if not np.allclose(fills, mean, rtol=1e-10, atol=0):
                problems.append("Real fill values differ from the torch f64 means")
            nmean = np.array([sm.mean[ncol[nm]] for nm in blk.names])
            if not np.array_equal(np.round(nmean * n_all), nulls):
                problems.append("Real null-indicator counts differ from the bitmap popcounts")
        elif isinstance(blk, DictBlock):
            for j, nm in enumerate(blk.names[:8]):
                cnull = torch.zeros(1, dtype=torch.int64, device="cuda")
                for b in ring:
                    cnull += (b.blocks[bi].codes[j, :b.n] < 0).sum()
                cnull = int(allsum(cnull).item())
                grp = [g for g in sm.categoricalStats if g.group == f"{nm}_{nm}"]
                if not grp:
                    problems.append(f"no categorical group for {nm}")
                    continue
                m = grp[0].contingencyMatrix
                if int(m.sum()) != n_all or int(m[-1].sum()) != cnull:
                    problems.append(f"PickList {nm}: contingency totals {int(m.sum())}, nulls {int(m[-1].sum())} != "
                                    f"{n_all}, {cnull}")
    # pass 2 against pass 1: column sums of the written rows == mean * n of the fit statistics
    keep = np.asarray(fit.fitted[-1]._keep)
    width = len(keep)
    sums = torch.zeros(width, dtype=torch.float64, device="cuda")

    def sink(k, vec, parts):
        off = 0
        for p_ in parts:
            if hasattr(p_, "row_nnz"):           # sparse rows
                used = torch.zeros(p_.indices.numel(), dtype=torch.bool, device="cuda")
                # every stored entry belongs to exactly one row: starts + lengths tile the array
                if int(p_.row_nnz.sum().item()) != p_.indices.numel():
                    problems.append("sparse rows do not tile their storage")
                sums[off:off + p_.width] += torch.bincount(p_.indices.long(), weights=p_.values.double(),
                                                            minlength=p_.width)[:p_.width]
                del used
            elif hasattr(p_, "indices"):         # one-hot
                idx = p_.indices.reshape(-1).long()
                idx = idx[idx >= 0]
                sums[off:off + p_.width] += torch.bincount(idx, minlength=p_.width)[:p_.width].double()
            else:
                sums[off:off + p_.width] += p_.data.sum(dim=0)
            off += p_.width

    row_nnz = []
    inner = sink

    def sink_nnz(k, vec, parts):
        row_nnz.append([p_.row_nnz.clone() for p_ in parts if hasattr(p_, "row_nnz")])
        inner(k, vec, parts)

    exp = sm.mean[:-1][keep] * n_all
    packed_models = [m for m in fit.fitted if getattr(m, "_cells", None)]
    n_packed = sum(len(m._cells["chunks"]) for m in packed_models)
    for how in ("packed cells", "text"):
        if how == "text":
            if not packed_models:
                break
            for m in packed_models:
                m.release_cell_cache()
        elif n_packed != len(ring) * len(packed_models):
            problems.append(f"only {n_packed} chunks were packed")
        sums.zero_()
        first_nnz, row_nnz = row_nnz, []
        transform_stream(ring, checked, fit, sink_nnz)
        got = allsum(sums.clone()).cpu().numpy()
        if not np.allclose(got, exp, rtol=1e-9, atol=1e-6):
            bad = np.nonzero(~np.isclose(got, exp, rtol=1e-9, atol=1e-6))[0]
            problems.append(f"pass-2 column sums (from the {how}) differ from the pass-1 statistics in {len(bad)} "
                            f"columns (first {bad[:3]})")
        if how == "text" and not all(torch.equal(a, b) for x, y in zip(first_nnz, row_nnz) for a, b in zip(x, y)):
            problems.append("rows from the packed cells and rows from the text differ in their lengths")
    if world > 1:
        import torch.distributed as dist
        box = [None] * world
        dist.all_gather_object(box, (tuple(sm.dropped), tuple(keep.tolist()[:50])))
        if any(b != box[0] for b in box):
            problems.append("ranks disagree on the dropped set")
    return "ok" if not problems else "FAILED: " + "; ".join(problems)


def bench_single(wl: str, rows: int, steps: int, warmup: int, rank: int, world: int, local_rank: int, barrier,
                 max_over_ranks, peak: float, reference_defaults: bool = False):
    """One resident batch per GPU (c2 / c3 / c4, or the c5 shape at a size that fits): weak scaling."""
    import torch
    import benchdata
    from transmogrifai_b200 import _lib as L
    batch, feats, label, info = benchdata.make_workload(wl, rows, SEED + rank)
    checked = build_plan(feats, label, info, reference_defaults)
    if reference_defaults:
        info = dict(info, desc=info["desc"] + " -- with every SanityChecker default of the reference: a sample of 1e6 "
                                              "rows is checked and featureFeatureCorrLevel = Computed (full Pearson "
                                              "matrix, Gramian on the fp64 tensor cores)")
    for _ in range(warmup):
        vec, fitted, parts = run_step(batch, checked)
        del vec, parts
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    calls0 = L.DEV_CALLS[0]
    e0.record()
    for _ in range(steps):
        vec, fitted, parts = run_step(batch, checked)
        del vec, parts
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1) / steps)
    launches = L.DEV_CALLS[0] - calls0
    clocks = sampler.stop()
    kt = kernel_times(wl, batch, fitted, info, text_offset=_text_offset(fitted))
    for k in kt:
        k["achieved"] = rows * k["bytes_per_row"] / (k["ms"] * 1e-3) / 1e9
        k["frac"] = k["achieved"] / peak
    if reference_defaults and wl == "c2":
        # the Gramian of the checked sample (fp64 tensor cores): upper-triangle flops 2 * n * D (D + 1) / 2
        from transmogrifai_b200 import runtime as R
        ns, dcol = 1_000_000, 200
        x = torch.randn((ns, dcol), dtype=torch.float64, device="cuda")
        yy = torch.randn(ns, dtype=torch.float64, device="cuda")
        sh = torch.zeros(dcol + 1, dtype=torch.float64, device="cuda")
        gram = torch.zeros((dcol + 1, dcol + 1), dtype=torch.float64, device="cuda")
        t_g = _time(lambda: L.call("tmog_dense_gram_dev", x.data_ptr(), dcol, ns, dcol, yy.data_ptr(), sh.data_ptr(),
                                   gram.data_ptr(), R.stream()), 3)
        flops = float(ns) * (dcol + 1) * (dcol + 2)
        kt.append({"kernel": "dense_gram_dmma_kernel (1e6 sampled rows x 201, fp64 mma.sync m8n8k4)", "ms": t_g,
                   "bound": "tensor (fp64)", "tflops_upper_triangle": flops / (t_g * 1e-3) / 1e12,
                   "bytes_per_row": (dcol + 1) * 8.0, "achieved": ns * (dcol + 1) * 8.0 / (t_g * 1e-3) / 1e9,
                   "frac": None})
        del x, yy, gram
    est = checked.origin_stage
    line = {"workload": info["desc"], "rows_per_gpu": rows, "ms_per_step": ms, "value": rows * world / (ms * 1e-3),
            "unit": "rows/s", "steps": steps, "warmup": warmup,
            "vector_width": len(est.summary.names) - 1, "kept_columns": len(fitted[-1].indicesToKeep),
            "dropped_columns": len(est.summary.dropped),
            "algorithmic_bytes_per_row": {"fit": info["fit_bytes"], "transform": info["transform_bytes"]},
            "step_hbm_frac": (info["fit_bytes"] + info["transform_bytes"]) * rows / (ms * 1e-3) / 1e9 / peak,
            "kernels": kt, "gpu_launches": launches, "clocks": clocks}
    return line, batch, checked, fitted, feats, label, info


def _text_offset(fitted) -> int:
    off = 0
    for m in fitted:
        if hasattr(m, "vectorizationMethods"):
            return off
        if hasattr(m, "indicesToKeep") or type(m).__name__ == "VectorsCombinerModel":
            break
        off += m.getMetadata().size
    return off


def main():
    import benchdata
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c5", choices=list(benchdata.WORKLOADS))
    ap.add_argument("--rows", type=int, default=0,
                    help="c5: rows of the whole job over all GPUs (default 100 M); c2 / c3 / c4: rows per GPU")
    ap.add_argument("--chunk-rows", type=int, default=1_000_000, help="rows per streamed chunk (c5)")
    ap.add_argument("--distinct-chunks", type=int, default=8, help="distinct synthetic chunks resident in HBM (c5)")
    ap.add_argument("--impl", default="tmog", choices=["tmog", "reference"])
    ap.add_argument("--cpu-rows", type=int, default=4_000_000, help="rows of the bounded CPU-baseline sample (c2)")
    ap.add_argument("--cell-cache-gb", type=float, default=-1.0,
                    help="HBM for the packed text cells kept between pass 1 and pass 2 (c5); -1 = what is free minus a "
                         "reserve, 0 = none")
    ap.add_argument("--full-width-return", action="store_true",
                    help="end-to-end arm: copy the output back full width instead of the compact return format")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-others", action="store_true", help="skip the c2 / c3 / c4 lines of the default run")
    ap.add_argument("--no-check", action="store_true", help="skip the untimed parity self-check")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    steps, warmup = max(args.steps, 1), max(args.warmup, 3 if args.impl == "tmog" else 1)
    cores = os.cpu_count() or 1
    wl = args.workload
    streamed = wl == "c5"
    total_rows = args.rows or (100_000_000 if streamed else benchdata.DEFAULT_ROWS[wl])

    if args.impl == "reference":
        if rank != 0:
            return
        import torch
        if wl != "c2":
            torch.cuda.set_device(0)      # the sample of the string workloads is drawn by the device generator
        secs, used = [], 1
        for i in range(warmup + steps):
            t, srows, used, sample = cpu_baseline(wl, cores, args.cpu_rows)
            if i >= warmup:
                secs.append(t)
        v = srows * len(secs) / sum(secs)
        what = (f"C5: {total_rows} rows x 200 mixed columns, chunk-streamed" if streamed else
                f"{wl}: {total_rows} rows per GPU") + f" -- timed on a bounded sample of {srows} rows per step"
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": v, "unit": "rows/s", "n_gpus": args.gpus, "steps": steps,
            "warmup": warmup, "ms_per_step": 1e3 * sum(secs) / len(secs), "higher_is_better": True,
            "scaling": "strong" if streamed else "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": what, "rows_per_step": srows},
            "cpu_baseline": {"value": v, "unit": "rows/s", "cores": used, "kind": "port",
                             "sample": sample + "; the JVM reference cannot run on this box (no JVM)"},
            "e2e": {"value": v, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import torch
    import torch.distributed as dist
    numa = bind_to_gpu_cpus(local_rank)      # before any pinned allocation: first touch decides the NUMA node
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from transmogrifai_b200 import _lib as L
    L.lib()   # fail loudly here if the CUDA library is missing
    peak, peak_src = peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v: float) -> float:
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if not streamed:
        line, batch, checked, fitted, feats, label, info = bench_single(
            wl, total_rows, steps, warmup, rank, world, local_rank, barrier, max_over_ranks, peak)
        rows = total_rows
        e2e = None
        if not args.no_e2e:
            from transmogrifai_b200.columnar import to_host
            hbatch = benchdata.host_copy(batch)
            h2d = benchdata.input_bytes(hbatch)
            del batch
            torch.cuda.empty_cache()

            def e2e_step():
                vec, _f, parts = run_step(hbatch, checked)
                dev = []
                for p in parts:                              # results that are still on the device go back to the host
                    for name in ("data", "indices", "values", "row_start", "row_nnz"):
                        t = getattr(p, name, None)
                        if t is not None and hasattr(t, "is_cuda"):
                            dev.append(t)
                return parts, to_host(dev, piece_bytes=1 << 30)  # pinned destinations, one wait

            for _ in range(2):
                parts, host = e2e_step()
            barrier()
            t0 = time.perf_counter()
            k_e2e = max(2, min(steps, 5))
            for _ in range(k_e2e):
                parts, host = e2e_step()
            barrier()
            dt = max_over_ranks((time.perf_counter() - t0) / k_e2e)
            e2e = {"value": rows * world / dt, "unit": "rows/s", "h2d_bytes_per_step": int(h2d),
                   "d2h_bytes_per_step": int(sum(t.numel() * t.element_size() for h in host
                                                 for t in (h if isinstance(h, list) else [h]))),
                   "ms_per_step": dt * 1e3, "steps": k_e2e,
                   "api": "fit_and_transform(pinned host ColumnarBatch, label.sanityCheck(transmogrify(features)))"}
            del hbatch, parts, host
        cpu = None
        if rank == 0 and world == 1 and not args.no_cpu:
            os.sched_setaffinity(0, ALL_CPUS)      # the CPU baseline runs on every host core
            secs, srows, used, sample = cpu_baseline(wl, cores, args.cpu_rows)
            cpu = {"value": srows / secs, "unit": "rows/s", "cores": used, "kind": "port",
                   "sample": f"{sample}, {secs:.2f} s"}
        if rank == 0:
            kt = line["kernels"]
            top = kt[0]
            print(json.dumps({
                "metric": METRIC, "value": line["value"], "unit": "rows/s", "n_gpus": world, "steps": steps,
                "warmup": warmup, "ms_per_step": line["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": line["workload"], "rows_per_gpu": rows,
                           "l2": "inputs exceed the 126 MB L2 many times over; no flush needed",
                           "metadata_caches": METADATA_NOTE,
                           "vector_width": line["vector_width"], "kept_columns": line["kept_columns"],
                           "dropped_columns": line["dropped_columns"],
                           "algorithmic_bytes_per_row": line["algorithmic_bytes_per_row"]},
                "roofline": {"bound": "hbm", "kernel": top["kernel"], "achieved": top["achieved"], "peak": peak,
                             "unit": "GB/s", "frac": top["frac"], "traffic": traffic_gb(wl, top["kernel"], rows),
                             "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full "
                                               "capture at 10 M rows, scaled by rows (profiles/r01_c2_full_raw.csv)",
                             "peak_source": peak_src, "step_frac": line["step_hbm_frac"],
                             "ms": top["ms"], "bytes_per_row": top["bytes_per_row"], "other": kt[1:]},
                "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": line["gpu_launches"], "clocks": line["clocks"]}))
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- config 5: 100 M rows x 200 mixed columns, chunk-streamed; the rows are split over the ranks (strong scaling)
    chunk_rows = min(args.chunk_rows, max(total_rows // world, 1))
    ring, feats, label, info = make_ring(chunk_rows, args.distinct_chunks, SEED + 100003 * rank)
    chunks, rows_rank = rank_chunks(ring, total_rows, chunk_rows, rank, world)
    sink = CountingSink()
    run_stream_step(chunks[:min(3, len(chunks))], feats, label, info, sink)     # allocator warm-up before sizing the arena
    cells_budget = cell_cache_budget(args.cell_cache_gb, 36.0)
    for _ in range(warmup):
        checked, res = run_stream_step(chunks, feats, label, info, sink, cell_cache_bytes=cells_budget)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    calls0 = L.DEV_CALLS[0]
    sink.bytes = 0
    e0.record()
    for _ in range(steps):
        checked, res = run_stream_step(chunks, feats, label, info, sink, cell_cache_bytes=cells_budget)
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1) / steps)
    launches = L.DEV_CALLS[0] - calls0
    out_bytes_step = sink.bytes // steps
    clocks = sampler.stop()
    value = total_rows / (ms * 1e-3)
    est = checked.origin_stage
    fitted = res.fitted
    summary = {"vector_width": len(est.summary.names) - 1, "kept_columns": len(fitted[-1].indicesToKeep),
               "dropped_columns": len(est.summary.dropped)}
    cells_cached = int(getattr(res, "cells_cached", 0))
    peak_alloc = int(torch.cuda.max_memory_allocated())

    # the kernels of one chunk, timed live on the launch stream (CUDA events), slowest first
    one = ring[0]
    one.__dict__.pop("_label_cache", None)
    kt = kernel_times("c5", one, fitted, info, reps=4, text_offset=_text_offset(fitted))
    for k in kt:
        k["achieved"] = one.n * k["bytes_per_row"] / (k["ms"] * 1e-3) / 1e9
        k["frac"] = k["achieved"] / peak
    # which of these calls the timed step made: with every chunk's cells packed, pass 2 never tokenises
    for k in kt:
        name = k["kernel"]
        if name.startswith("text tokenise + hash"):
            k["in_timed_step"] = cells_cached < res.chunks
        elif name.startswith("text statistics + packed") or name.startswith("text_rows_packed"):
            k["in_timed_step"] = cells_cached > 0
        else:
            k["in_timed_step"] = True
    kt = [k for k in kt if k["in_timed_step"]] + [k for k in kt if not k["in_timed_step"]]
    bytes_row = info["fit_bytes"] + info["transform_bytes"]
    step_frac = bytes_row * total_rows / world / (ms * 1e-3) / 1e9 / peak

    check = None
    if not args.no_check:
        check = parity_check_stream(ring[:min(len(ring), 4)], feats, label, info, world)

    # end to end: the same job from pinned host chunks -- every chunk crosses PCIe in pass 1 and again in pass 2 (on a
    # side stream, a chunk ahead), its output goes back to pinned buffers on a third stream
    e2e = None
    if not args.no_e2e:
        from transmogrifai_b200.streaming import HostSink
        # pinned host memory: at most a third of what the box has free, shared by the ranks (distinct host chunks: up to 4)
        try:
            avail = next(int(ln.split()[1]) * 1024 for ln in open("/proc/meminfo") if ln.startswith("MemAvailable"))
        except (OSError, StopIteration):
            avail = 64 << 30
        chunk_host_bytes = benchdata.input_bytes(ring[0]) + 8 * chunk_rows
        n_host = int(max(1, min(len(ring), 4, (avail // 3 // world - 2 * chunk_host_bytes) // chunk_host_bytes)))
        hring = [benchdata.host_copy(b) for b in ring[:n_host]]
        for h, b in zip(hring, ring):
            h.token = b.token
        del h, b
        in_bytes = [benchdata.input_bytes(h) for h in hring]
        label_bytes = 8 * chunk_rows
        del ring, chunks, one
        torch.cuda.empty_cache()
        hchunks = [hring[i % n_host] for i in range(rows_rank // chunk_rows)]
        rows_e2e = len(hchunks) * chunk_rows
        hs = HostSink(depth=2, compact=not args.full_width_return)
        run_stream_step(hchunks[:min(4, len(hchunks))], feats, label, info, hs)      # warm-up: allocator, pinned buffers
        hs.bytes = 0
        torch.cuda.empty_cache()
        # HBM between the passes: the packed text cells of as many chunks as fit (their text is not uploaded again);
        # without them, whole uploaded chunks up to half of the free memory
        if os.environ.get("TMOG_BENCH_DEBUG"):
            print(f"[e2e] allocated {torch.cuda.memory_allocated() / 2**30:.1f} GiB, free "
                  f"{torch.cuda.mem_get_info()[0] / 2**30:.1f} GiB", file=sys.stderr)
        e_cells = cell_cache_budget(args.cell_cache_gb, 64.0)
        budget = 0 if e_cells else int(torch.cuda.mem_get_info()[0] * 0.5)
        text_bytes = [benchdata.text_bytes(h) for h in hring]
        barrier()
        t0 = time.perf_counter()
        _c, eres = run_stream_step(hchunks, feats, label, info, hs, cache_bytes=budget, cell_cache_bytes=e_cells)
        barrier()
        dt = max_over_ranks(time.perf_counter() - t0)
        kept = int(getattr(eres, "cache_hits", 0))
        packed = int(getattr(eres, "cells_cached", 0))
        # pass 1 uploads every chunk; pass 2 uploads it again unless it stayed resident, and without its text when its
        # cells are packed (the arena fills in chunk order)
        h2d = sum(in_bytes[i % n_host] + label_bytes + (0 if i < kept else in_bytes[i % n_host] -
                                                        (text_bytes[i % n_host] if i < packed else 0))
                  for i in range(len(hchunks))) + in_bytes[0]
        e2e = {"value": rows_e2e * world / dt, "unit": "rows/s", "h2d_bytes_per_step": int(h2d) * world,
               "d2h_bytes_per_step": int(hs.bytes) * world,      # every rank moves its own (equal) share
               "ms_per_step": dt * 1e3, "steps": 1, "rows": rows_e2e * world,
               "distinct_host_chunks": n_host, "chunks_kept_in_hbm_between_passes": kept,
               "hbm_cache_budget_bytes": budget, "chunks_with_packed_cells": packed, "cell_cache_bytes": e_cells,
               "return_format": ("full width: f64 dense blocks, int32 term frequencies" if args.full_width_return else
                                 "compact (streaming.compact_parts): null indicators and Binary values as u8, term "
                                 "frequencies as u8 with an overflow list, column indices of the text rows as u16; "
                                 "row offsets, one-hot indices and Real values unchanged"),
               "api": "fit_and_transform_stream(pinned host chunks, label.sanityCheck(transmogrify(features)), HostSink)"}
        del hring, hchunks, hs
        torch.cuda.empty_cache()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        os.sched_setaffinity(0, ALL_CPUS)          # the CPU baseline runs on every host core
        secs, srows, used, sample = cpu_baseline("c5", cores, args.cpu_rows)
        cpu = {"value": srows / secs, "unit": "rows/s", "cores": used, "kind": "port",
               "sample": f"{sample}, {secs:.2f} s"}

    others = []
    if world == 1 and not args.no_others:
        for owl, osteps, refdef in (("c2", 10, False), ("c2", 5, True), ("c3", 5, False), ("c4", 3, False)):
            torch.cuda.empty_cache()
            oline, ob, _c, _f, _fe, _l, _i = bench_single(owl, benchdata.DEFAULT_ROWS[owl], osteps, 3, rank, world,
                                                           local_rank, barrier, max_over_ranks, peak, refdef)
            del ob, _c, _f
            others.append(oline)

    if rank == 0:
        top = kt[0]
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": "rows/s", "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"C5: {total_rows} rows x 200 mixed columns (100 Real + 50 PickList + 30 Text with 512 "
                                   f"hashes per column in separate spaces + 20 Binary), chunk-streamed (chunk = "
                                   f"{chunk_rows} rows): full transmogrify() + sanityCheck(removeBadFeatures) fit + "
                                   f"transform; rows split over {world} GPU(s)",
                       "rows_total": total_rows, "rows_per_gpu": rows_rank, "chunk_rows": chunk_rows,
                       "chunks_per_gpu": res.chunks, "distinct_chunks": args.distinct_chunks,
                       "chunk_source": "cycled from a ring of distinct synthetic chunks resident in HBM (100 M rows are "
                                       "~740 GB of raw columns); every chunk is read in pass 1 and again in pass 2",
                       "l2": "a chunk (~7 GB) exceeds the 126 MB L2 many times over; no flush needed",
                       "metadata_caches": METADATA_NOTE,
                       "host_cpus": (f"{len(numa)} CPUs local to the GPU (NVML affinity)" if numa else "not bound"),
                       "output_bytes_per_step": int(out_bytes_step),
                       "algorithmic_bytes_per_row": {"fit": info["fit_bytes"], "transform": info["transform_bytes"]},
                       "hbm_peak_allocated_bytes": peak_alloc,
                       "packed_cells": {"hbm_budget_bytes": int(cells_budget), "chunks_packed_per_gpu": cells_cached,
                                        "note": "pass 1 keeps every text cell's buckets and term frequencies (~0.9 KB "
                                                "per row) for as many chunks as the budget holds; pass 2 of those chunks "
                                                "reads them instead of tokenising the text again; filled and released "
                                                "inside every step"},
                       **summary},
            "roofline": {"bound": "hbm", "kernel": top["kernel"], "achieved": top["achieved"], "peak": peak,
                         "unit": "GB/s", "frac": top["frac"], "traffic": traffic_gb("c5", top["kernel"], chunk_rows),
                         "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture at "
                                           "1 M rows, scaled by rows (profiles/r02_c5_text_raw.csv); for the pass-2 call "
                                           "it includes the 128-byte cell records written by one kernel and read by the "
                                           "next (7.7 KB per row beyond the algorithmic bytes)",
                         "peak_source": peak_src,
                         "ms": top["ms"], "bytes_per_row": top["bytes_per_row"], "rows_per_launch": chunk_rows,
                         "step_frac": step_frac, "other": kt[1:]},
            "parity_check": check, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "others": others}))
    if world > 1:
        dist.destroy_process_group()
    if check is not None and check != "ok":
        sys.exit(3)


if __name__ == "__main__":
    main()
