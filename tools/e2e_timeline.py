"""Where an end-to-end streamed step spends its time: per pass, host time at which each chunk was queued and the PCIe rates
(development aid).  python tools/e2e_timeline.py [chunks] [chunk_rows]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import benchdata  # noqa: E402
from transmogrifai_b200.streaming import HostSink, fit_and_transform_stream  # noqa: E402

nchunks = int(sys.argv[1]) if len(sys.argv) > 1 else 8
rows = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
ring, feats, label, info = bench.make_ring(rows, 2, 5)
hring = [benchdata.host_copy(b) for b in ring]
in_bytes = benchdata.input_bytes(hring[0])
del ring
torch.cuda.empty_cache()
chunks = [hring[i % 2] for i in range(nchunks)]
# raw link rates with the same buffers: one direction at a time, then both
blk = hring[0].blocks[2]            # the Real block: one 800 MB tensor
dev = blk.values.cuda()
host = torch.empty(blk.values.shape, dtype=blk.values.dtype, pin_memory=True)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def rate(h2d, d2h, reps=8):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if h2d:
            with torch.cuda.stream(s1):
                dev.copy_(blk.values, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2):
                host.copy_(dev2, non_blocking=True)
    torch.cuda.synchronize()
    return reps * dev.numel() * 8 / (time.perf_counter() - t0) / 1e9
dev2 = dev.clone()
print("H2D alone %.1f GB/s, D2H alone %.1f GB/s, both at once %.1f GB/s each" % (rate(1, 0), rate(0, 1), rate(1, 1)), flush=True)
del dev, dev2, host
for trial in range(2):
    stamps = []
    def on_chunk(p, k, c):
        stamps.append((p, k, time.perf_counter()))
    sink = HostSink(depth=2)
    checked = bench.build_plan(feats, label, info)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    fit_and_transform_stream(chunks, checked, sink, on_chunk=on_chunk)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    print(f"trial {trial}: {t1 - t0:.3f} s for {nchunks} chunks of {rows} rows = {nchunks * rows / (t1 - t0) / 1e6:.2f} M rows/s; "
          f"input {in_bytes / 1e9:.2f} GB per chunk, output {sink.bytes / nchunks / 1e9:.2f} GB per chunk", flush=True)
    for p in (0, 1, 2):
        ts = [t for q, k, t in stamps if q == p]
        if ts:
            print(f"  pass {p}: first chunk queued at {ts[0] - t0:.3f} s, last at {ts[-1] - t0:.3f} s")
