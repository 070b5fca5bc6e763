"""cProfile of bench steps sorted by own time (development aid): python tools/hostprof_step.py <workload>"""
import cProfile, pstats, sys, os
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
pr = cProfile.Profile()
pr.enable()
for _ in range(10):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(40)
