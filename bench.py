#!/usr/bin/env python
"""bench.py -- rows/sec of transmogrify() + sanityCheck() fit + transform on synthetic data (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--rows R] [--impl tmog|reference]

One process per GPU (torchrun for N > 1); rows are sharded across ranks (weak scaling: R rows per GPU) and the only
exchange is the NCCL all-reduce of the per-column partial sums / count tables.  A "step" is one full pass of the hot
path over the resident batch: fused vectorizer + SanityChecker scan (fit), the drop decision, and the transform that
writes the kept columns.  Prints ONE JSON line (see DESIGN.md "Measurement").
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

WORKLOAD = ("C2: synthetic {rows} rows x {ncols} Real columns (5% nulls) per GPU: RealVectorizer mean-fill + null "
            "indicators fused with SanityChecker (column moments, label correlations, Cramer's V of the null "
            "indicators, removeBadFeatures decision, featureFeatureCorrLevel=Off), then the transform writing the "
            "kept columns")
NCOLS = 100
SEED = 20260001
# algorithmic bytes per row: fit = one read of values + bitmaps + label; transform = the same read
# without the label + 16 bytes per (row, input column) written
FIT_BYTES_PER_ROW = NCOLS * 8 + NCOLS / 8 + 8
TRANSFORM_BYTES_PER_ROW = NCOLS * 8 + NCOLS / 8 + NCOLS * 16


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


def make_device_batch(rows: int, seed: int):
    """Config-2 data generated on the device (torch RNG, seeded): N(j, 1 + j/10) columns, 5 % nulls, Bernoulli(0.3)
    label, 5 label-shifted columns, one constant column, one duplicated column."""
    import torch
    from transmogrifai_b200 import features as F
    from transmogrifai_b200.columnar import ColumnarBatch, NumericBlock
    g = torch.Generator(device="cuda").manual_seed(seed)
    ld = (rows + 2047) // 2048 * 2048
    nw = ld // 64
    y = (torch.rand(ld, device="cuda", generator=g) < 0.3).double()
    vals = torch.empty((NCOLS, ld), dtype=torch.float64, device="cuda")
    valid = torch.empty((NCOLS, nw), dtype=torch.int64, device="cuda")
    w = (torch.ones(64, dtype=torch.int64, device="cuda") << torch.arange(64, device="cuda")).view(1, 64)
    for j in range(NCOLS):
        vals[j].normal_(float(j), 1.0 + j / 10.0, generator=g)
        ok = (torch.rand(ld, device="cuda", generator=g) >= 0.05).view(nw, 64).long()
        valid[j] = (ok * w).sum(dim=1)
    for j in range(0, 50, 10):
        vals[j] += 0.5 * y
    vals[7] = 3.25
    vals[12] = vals[11]
    valid[12] = valid[11]
    names = [f"c{j:03d}" for j in range(NCOLS)]
    blocks = [NumericBlock(names, F.REAL, vals, valid, rows),
              NumericBlock(["label"], F.REALNN, y.view(1, ld), None, rows)]
    feats = [F.FeatureBuilder.Real(nm).asPredictor() for nm in names]
    label = F.FeatureBuilder.RealNN("label").asResponse()
    return ColumnarBatch(rows, blocks, {}, response="label"), feats, label


def host_copy(batch):
    import torch
    from transmogrifai_b200.columnar import ColumnarBatch, NumericBlock
    blocks = []
    for b in batch.blocks:
        v = torch.empty(b.values.shape, dtype=b.values.dtype, pin_memory=True)
        v.copy_(b.values)
        m = None
        if b.valid is not None:
            m = torch.empty(b.valid.shape, dtype=b.valid.dtype, pin_memory=True)
            m.copy_(b.valid)
        blocks.append(NumericBlock(b.names, b.type_name, v, m, b.n))
    return ColumnarBatch(batch.n, blocks, {}, batch.response)


def build_plan(feats, label):
    from transmogrifai_b200.dsl import transmogrify
    vec = transmogrify(feats)
    return label.sanityCheck(vec, removeBadFeatures=True, checkSample=1.0, sampleLowerLimit=0,
                             sampleUpperLimit=2 ** 31 - 1, featureFeatureCorrLevel="Off")


def run_step(batch, checked):
    from transmogrifai_b200.dsl import fit_and_transform
    batch.__dict__.pop("_label_cache", None)   # label classes are part of the step: never reuse a previous step's
    batch.__dict__.pop("_label_dev", None)
    out, fitted = fit_and_transform(batch, checked)
    vec = out.vectors[checked.name]
    for p in vec.parts:           # the transform is lazy until somebody reads the vector: read it
        _ = p.data
    return vec, fitted


def kernel_times(batch, reps: int = 5):
    """Average device time of the two dominant kernels at this batch size (CUDA events on the launch stream)."""
    import ctypes as C
    import torch
    from transmogrifai_b200 import _lib as L
    from transmogrifai_b200 import runtime as R
    blk = batch.numeric(batch.blocks[0].names)
    y = batch.label()
    n, ncols = batch.n, len(blk.names)
    nb = C.c_size_t()
    L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nb))
    scratch = R.scratch("real_fit", nb.value)
    shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
    part = torch.zeros((ncols + 1) * 8, dtype=torch.float64, device="cuda")
    fills = torch.zeros(ncols, dtype=torch.float64, device="cuda")
    out = torch.empty((n, 2 * ncols), dtype=torch.float64, device="cuda")
    st = R.stream()
    L.call("tmog_real_pilot_shift_dev", blk.values.data_ptr(), blk.ld, blk.valid.data_ptr(), blk.ldv, y.data_ptr(), n,
           ncols, shifts.data_ptr(), st)

    def fit():
        L.call("tmog_real_fit_partials_dev", blk.values.data_ptr(), blk.ld, blk.valid.data_ptr(), blk.ldv,
               y.data_ptr(), n, ncols, shifts.data_ptr(), scratch.data_ptr(), scratch.numel(), part.data_ptr(), st)

    def tr():
        L.call("tmog_real_transform_dev", blk.values.data_ptr(), blk.ld, blk.valid.data_ptr(), blk.ldv, n, ncols,
               fills.data_ptr(), 1, out.data_ptr(), 2 * ncols, 0, st)

    res = {}
    for name, f in (("fit", fit), ("transform", tr)):
        for _ in range(2):
            f()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            f()
        e1.record()
        torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / reps
    del out
    return res


def cpu_reference(rows: int, threads: int):
    """The reference's CPU job sequence for this workload (oracle/c port, OpenMP) on `rows` rows: returns seconds."""
    from oracle import c_port as CP
    from tests.util_data import make_real_config, pack_validity
    x, valid, y = make_real_config(rows, NCOLS, seed=SEED)
    vw = pack_validity(valid)
    x = np.ascontiguousarray(x)
    CP.real_pipeline(x[:, :2048].copy(), pack_validity(valid[:, :2048]), y[:2048].copy(), threads)   # page in
    t0 = time.perf_counter()
    fills, rows_m, mean, var, mn, mx, corr = CP.real_pipeline(x, vw, y, threads)
    CP.label_sums(rows_m[:, :-1], y, np.array([0.0, 1.0]), threads)      # categoricalTests' reduceByKey
    return time.perf_counter() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--rows", type=int, default=10_000_000, help="rows per GPU")
    ap.add_argument("--impl", default="tmog", choices=["tmog", "reference"])
    ap.add_argument("--cpu-rows", type=int, default=1_000_000, help="rows of the bounded CPU-baseline sample")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    steps, warmup = max(args.steps, 1), max(args.warmup, 3 if args.impl == "tmog" else 1)
    cores = os.cpu_count() or 1

    if args.impl == "reference":
        if rank != 0:
            return
        secs = []
        for i in range(warmup + steps):
            t = cpu_reference(args.cpu_rows, cores)
            if i >= warmup:
                secs.append(t)
        v = args.cpu_rows * len(secs) / sum(secs)
        sample = (f"{args.cpu_rows} rows x {NCOLS} Real columns of the same workload per step; the reference's job "
                  f"sequence (fit, transform, colStats, label correlations, per-label sums) as oracle/c "
                  f"(C + OpenMP) on {cores} threads; the JVM reference cannot run on this box")
        print(json.dumps({
            "impl": "reference", "metric": "rows/sec transmogrify+sanityCheck fit+transform", "value": v,
            "unit": "rows/s", "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
            "ms_per_step": 1e3 * sum(secs) / len(secs), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD.format(rows=args.cpu_rows, ncols=NCOLS)},
            "cpu_baseline": {"value": v, "unit": "rows/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from transmogrifai_b200 import _lib as L
    L.lib()   # fail loudly here if the CUDA library is missing

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    rows = args.rows
    batch, feats, label = make_device_batch(rows, SEED + rank)
    checked = build_plan(feats, label)

    for _ in range(warmup):
        vec, fitted = run_step(batch, checked)
        del vec
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(steps):
        vec, fitted = run_step(batch, checked)
        del vec
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / steps
    clocks = sampler.stop()
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = rows * world / (ms * 1e-3)

    # dominant kernels, measured live on this batch
    kt = kernel_times(batch)
    peak, peak_src = peaks()
    tr_gbs = rows * TRANSFORM_BYTES_PER_ROW / (kt["transform"] * 1e-3) / 1e9
    fit_gbs = rows * FIT_BYTES_PER_ROW / (kt["fit"] * 1e-3) / 1e9

    # end to end through the stage API with pinned host buffers (H2D of the raw columns, D2H of the kept columns)
    e2e = None
    if not args.no_e2e:
        hbatch = host_copy(batch)
        est = checked.origin_stage
        for _ in range(2):
            vec, _f = run_step(hbatch, checked)
        barrier()
        t0 = time.perf_counter()
        k_e2e = max(2, min(steps, 5))
        for _ in range(k_e2e):
            vec, _f = run_step(hbatch, checked)
        barrier()
        dt = (time.perf_counter() - t0) / k_e2e
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        h2d = sum(b.values.numel() * b.values.element_size() + (0 if b.valid is None else b.valid.numel() * 8)
                  for b in hbatch.blocks)
        d2h = sum(p.data.numel() * 8 for p in vec.parts)
        e2e = {"value": rows * world / dt, "unit": "rows/s", "h2d_bytes_per_step": int(h2d),
               "d2h_bytes_per_step": int(d2h), "ms_per_step": dt * 1e3, "steps": k_e2e,
               "api": "fit_and_transform(host ColumnarBatch, label.sanityCheck(transmogrify(features))) -> "
                      "tmog_real_fit_host / tmog_real_transform_host"}
        del hbatch, vec

    cpu = None
    if rank == 0 and world == 1:
        secs = cpu_reference(args.cpu_rows, cores)
        cpu = {"value": args.cpu_rows / secs, "unit": "rows/s", "cores": cores, "kind": "port",
               "sample": f"{args.cpu_rows} rows x {NCOLS} columns of the same workload, oracle/c (C + OpenMP) on "
                         f"{cores} threads, {secs:.2f} s"}
    if rank == 0:
        est = checked.origin_stage
        launches_per_step = 9   # pilot, fit, fit-reduce, label init/distinct/sort/class-id, null-class-count, transform
        print(json.dumps({
            "metric": "rows/sec transmogrify+sanityCheck fit+transform", "value": value, "unit": "rows/s",
            "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": WORKLOAD.format(rows=rows, ncols=NCOLS), "rows_per_gpu": rows,
                       "l2": "inputs (8.1 GB per GPU) exceed the 126 MB L2; no flush needed",
                       "kept_columns": len(fitted[-1].indicesToKeep), "dropped": est.summary.dropped},
            "roofline": {"bound": "hbm", "kernel": "vec_transform_kernel<double, true>", "achieved": tr_gbs,
                         "peak": peak, "unit": "GB/s", "frac": tr_gbs / peak, "traffic": None,
                         "peak_source": peak_src, "ms": kt["transform"],
                         "bytes_per_row": TRANSFORM_BYTES_PER_ROW,
                         "other": {"kernel": "real_fit_kernel", "achieved": fit_gbs, "frac": fit_gbs / peak,
                                   "ms": kt["fit"], "bytes_per_row": FIT_BYTES_PER_ROW}},
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches_per_step * steps, "clocks": clocks}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
