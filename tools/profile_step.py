"""Host-side profile of one bench step (development aid): python tools/profile_step.py <workload> [rows]"""
import cProfile, pstats, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench, benchdata
wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
rows = int(sys.argv[2]) if len(sys.argv) > 2 else benchdata.DEFAULT_ROWS[wl]
batch, feats, label, info = benchdata.make_workload(wl, rows, 1)
checked = bench.build_plan(feats, label, info)
for _ in range(3):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(3):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
print("ms/step", (time.perf_counter() - t0) / 3 * 1e3)
# per-entry-point time (meaningful with CUDA_LAUNCH_BLOCKING=1)
from transmogrifai_b200 import _lib
from collections import defaultdict
acc = defaultdict(lambda: [0, 0.0])
_orig = _lib.call
def _timed(name, *a):
    t = time.perf_counter()
    _orig(name, *a)
    torch.cuda.synchronize()
    e = acc[name]; e[0] += 1; e[1] += time.perf_counter() - t
_lib.call = _timed
for _ in range(3):
    bench.run_step(batch, checked)
_lib.call = _orig
for k, (c, t) in sorted(acc.items(), key=lambda kv: -kv[1][1]):
    print(f"  {k:40s} calls/step {c / 3:5.1f}  ms/step {t / 3 * 1e3:8.2f}")
pr = cProfile.Profile()
pr.enable()
for _ in range(3):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(32)
