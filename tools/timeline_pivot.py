"""Wall-clock timeline of the C-ABI calls of one config-3 step (which thread, when, how long): shows what the host
does between the count kernel and the transform.  python tools/timeline_pivot.py [rows]"""
import sys, threading, time
sys.path.insert(0, "/root/repo")
import torch
import bench, benchdata
from transmogrifai_b200 import _lib as L

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
batch, feats, label, info = benchdata.make_workload("c3", rows, 1)
checked = bench.build_plan(feats, label, info)
for _ in range(3):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
events, orig, lock = [], L.call, threading.Lock()


def timed(name, *a):
    t0 = time.perf_counter()
    try:
        return orig(name, *a)
    finally:
        t1 = time.perf_counter()
        with lock:
            events.append((t0, t1, threading.get_ident(), name))


L.call = timed
for mod in list(sys.modules.values()):
    if getattr(mod, "L", None) is L:
        pass                                     # the stages call L.call through the module attribute
t_start = time.perf_counter()
bench.run_step(batch, checked)
torch.cuda.synchronize()
t_end = time.perf_counter()
L.call = orig
tids = {}
print(f"step {1e3 * (t_end - t_start):.2f} ms, {len(events)} calls")
agg = {}
for t0, t1, tid, name in sorted(events):
    k = tids.setdefault(tid, len(tids))
    a = agg.setdefault(name, [0, 0.0, 1e9, 0.0])
    a[0] += 1
    a[1] += t1 - t0
    a[2] = min(a[2], t0 - t_start)
    a[3] = max(a[3], t1 - t_start)
for name, (cnt, tot, first, last) in sorted(agg.items(), key=lambda kv: kv[1][2]):
    print(f"{name:36s} n={cnt:3d} sum={1e3 * tot:7.2f} ms  first start {1e3 * first:6.2f}  last end {1e3 * last:6.2f}")
big = sorted(events, key=lambda e: e[0] - e[1])[:12]
for t0, t1, tid, name in big:
    print(f"  longest: {name:30s} thread {tids[tid]:2d} {1e3 * (t0 - t_start):6.2f} -> {1e3 * (t1 - t_start):6.2f}")
