import cProfile, pstats, sys, os
sys.path.insert(0, "/root/repo")
import torch
import bench, benchdata
batch, feats, label, info = benchdata.make_workload(sys.argv[1] if len(sys.argv) > 1 else "c4", int(sys.argv[2]) if len(sys.argv) > 2 else 2000000, 1)
checked = bench.build_plan(feats, label, info)
for _ in range(3):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(int(sys.argv[3]) if len(sys.argv) > 3 else 5):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(45)
