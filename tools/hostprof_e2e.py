"""cProfile of the end-to-end step (pinned host batch in, kept columns back to pinned host memory):
python tools/hostprof_e2e.py c5 1500000 3"""
import cProfile, pstats, sys
sys.path.insert(0, "/root/repo")
import torch
import bench, benchdata
from transmogrifai_b200.columnar import to_host
batch, feats, label, info = benchdata.make_workload(sys.argv[1] if len(sys.argv) > 1 else "c5",
                                                    int(sys.argv[2]) if len(sys.argv) > 2 else 1500000, 1)
checked = bench.build_plan(feats, label, info)
hbatch = benchdata.host_copy(batch)
del batch
torch.cuda.empty_cache()


def step():
    vec, _f, parts = bench.run_step(hbatch, checked)
    dev = [t for p in parts for t in (getattr(p, nm, None) for nm in ("data", "indices", "values", "row_start",
                                                                      "row_nnz")) if t is not None and hasattr(t, "is_cuda")]
    return to_host(dev, piece_bytes=1 << 30)


for _ in range(2):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(int(sys.argv[3]) if len(sys.argv) > 3 else 3):
    step()
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(60)
