"""Host-side profile of one bench step (development aid)."""
import cProfile, pstats, sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
rows = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
batch, feats, label = bench.make_device_batch(rows, 1)
checked = bench.build_plan(feats, label)
for _ in range(3):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
print("ms/step", (time.perf_counter() - t0) / 5 * 1e3)
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(35)
