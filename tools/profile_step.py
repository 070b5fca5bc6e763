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
pr = cProfile.Profile()
pr.enable()
for _ in range(3):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(32)
