"""cProfile (cumulative) of bench steps on rank 0 under torchrun (development aid):
    torchrun --nproc-per-node 2 tools/hostprof_dist.py <workload> <rows per GPU> [steps]"""
import cProfile, pstats, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import bench, benchdata
rank, local = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
wl, rows = sys.argv[1], int(sys.argv[2])
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
batch, feats, label, info = benchdata.make_workload(wl, rows, 20260001 + rank)
checked = bench.build_plan(feats, label, info)
for _ in range(3):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
dist.barrier()
pr = cProfile.Profile()
pr.enable()
for _ in range(steps):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
pr.disable()
if rank == 0:
    pstats.Stats(pr).sort_stats("cumulative").print_stats(40)
dist.destroy_process_group()
