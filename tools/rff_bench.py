"""RawFeatureFilter on one GPU: (1) the two numeric scans on config 2's block (10 M rows x 100 Real, 5 % nulls) timed
alone with CUDA events against the measured HBM peak; (2) computeFeatureStats (training pass: summaries, distributions,
null-indicator / response correlations) over one 1 M-row chunk of config 5's shape (100 Real + 50 PickList + 30 Text
+ 20 Binary), host glue included; (3) the oracle's per-row restatement on a small sample of the same rows (one core).
    python tools/rff_bench.py [--rows-c2 10000000] [--rows-c5 1000000] > profiles/r02_rff_bench.json"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402

import benchdata  # noqa: E402
from transmogrifai_b200.filters import RawFeatureFilter  # noqa: E402


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows-c2", type=int, default=10_000_000)
    ap.add_argument("--rows-c5", type=int, default=1_000_000)
    ap.add_argument("--sample", type=int, default=300)
    args = ap.parse_args()
    peak = 6543.4
    try:
        peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    out = {"hbm_peak_gbps": peak}
    f = RawFeatureFilter(bins=100)
    # ---- (1) the numeric scans alone
    batch, feats, label, _info = benchdata.make_workload("c2", args.rows_c2, 1)
    block = batch.blocks[0]
    n, k = batch.n, len(block.names)
    nbytes = n * k * 8 + (0 if block.valid is None else n * k / 8)
    summ = torch.tensor([[np.inf, -np.inf, 0.0, 0.0, 0.0]] * k, dtype=torch.float64, device="cuda")
    ms = timed(lambda: f._k_summary(block, n, summ))
    out["numeric_summary"] = {"rows": n, "cols": k, "ms": ms, "gbps": nbytes / ms / 1e6, "frac": nbytes / ms / 1e6 / peak}
    s = summ.cpu().numpy()
    splits = np.stack([s[i, 0] + (s[i, 1] - s[i, 0]) / 98.0 * np.arange(100) for i in range(k)])
    d_splits = torch.from_numpy(splits).cuda()
    mode = torch.ones(k, dtype=torch.int32, device="cuda")
    same = torch.zeros(k, dtype=torch.float64, device="cuda")
    hist = torch.zeros((k, 100), dtype=torch.int64, device="cuda")
    vsum = torch.zeros(k, dtype=torch.float64, device="cuda")
    nonnull = torch.zeros(k, dtype=torch.int64, device="cuda")
    ms = timed(lambda: f._k_hist(block, n, d_splits, mode, same, hist, vsum, nonnull))
    out["numeric_hist"] = {"rows": n, "cols": k, "bins": 100, "ms": ms, "gbps": nbytes / ms / 1e6,
                           "frac": nbytes / ms / 1e6 / peak}
    del batch, block, summ, hist
    torch.cuda.empty_cache()
    # ---- (2) one chunk of config 5's shape through computeFeatureStats
    batch, feats, label, _info = benchdata.make_workload("c5", args.rows_c5, 2)
    every = [label] + list(feats)
    f.computeFeatureStats(batch, every)                       # warm-up (dictionaries tokenised, allocator primed)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    info = f.computeFeatureStats(batch, every)
    torch.cuda.synchronize()
    sec = time.perf_counter() - t0
    out["c5_chunk_training_stats"] = {"rows": batch.n, "features": len(feats), "seconds": sec,
                                      "rows_per_s": batch.n / sec, "input_gb": benchdata.input_bytes(batch) / 1e9,
                                      "distributions": len(info.predictorDistributions)}
    # ---- (3) the oracle on the first rows of the same chunk (one core, per-row Python as the reference is per-row JVM)
    from oracle import raw_feature_filter as O
    m = args.sample
    cols = {}
    host = benchdata.host_copy(batch)
    for b in host.blocks:
        for j, nm in enumerate(b.names):
            if hasattr(b, "codes"):
                codes = b.codes[j][:m].numpy()
                cols[nm] = (b.type_name, [None if c < 0 else b.dicts[j][int(c)] for c in codes])
            elif hasattr(b, "offsets"):
                cols[nm] = (b.type_name, b.prefix_cells(j, m))
            else:
                v = b.values[j][:m].numpy()
                ok = np.ones(m, dtype=bool) if b.valid is None else np.unpackbits(
                    b.valid[j].numpy().view(np.uint8), bitorder="little")[:m].astype(bool)
                cols[nm] = (b.type_name, [float(x) if o else None for x, o in zip(v, ok)])
    t0 = time.perf_counter()
    O.compute_feature_stats(cols, ["label"], 100, "Training")
    sec = time.perf_counter() - t0
    out["oracle_port"] = {"rows": m, "seconds": sec, "rows_per_s": m / sec, "cores": 1}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
