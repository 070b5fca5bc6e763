"""Aggregate host<->device copy rate of all ranks of one node at once (what bounds the end-to-end arm at N GPUs):
every rank binds to the CPUs local to its GPU (as bench.py does), pins a buffer, and copies H2D only, D2H only, and
both ways at once; rank 0 prints the per-rank and summed GB/s.
   python -m torch.distributed.run --nproc-per-node N tools/pcie_aggregate.py [GiB per buffer]"""
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
numa = bench.bind_to_gpu_cpus(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
gib = float(sys.argv[1]) if len(sys.argv) > 1 else 2.0
n = int(gib * 2 ** 30)
h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_in.fill_(1)
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.ones(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()


def run(up: bool, down: bool, reps: int = 8) -> float:
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        if up:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if down:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    barrier()
    return reps * n / (time.perf_counter() - t0) / 1e9


run(True, True, 2)
res = torch.tensor([run(True, False), run(False, True), run(True, True)], dtype=torch.float64, device="cuda")
box = [torch.zeros_like(res) for _ in range(world)]
if world > 1:
    dist.all_gather(box, res)
else:
    box = [res]
if rank == 0:
    rows = torch.stack(box).cpu().numpy()
    print(f"{world} rank(s), {gib:g} GiB pinned buffers, CPUs bound: {len(numa) if numa else 'no'}")
    for name, col in (("H2D only", 0), ("D2H only", 1), ("both ways (each direction)", 2)):
        print(f"{name:28s} per rank min {rows[:, col].min():6.1f} max {rows[:, col].max():6.1f} GB/s   sum {rows[:, col].sum():7.1f} GB/s")
if world > 1:
    dist.destroy_process_group()
