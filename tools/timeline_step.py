"""Host timeline of one bench step: when every C-ABI call is issued and when the host blocks (development aid).
    python tools/timeline_step.py <workload>"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench, benchdata
from transmogrifai_b200 import _lib
wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
rows = int(sys.argv[2]) if len(sys.argv) > 2 else benchdata.DEFAULT_ROWS[wl]
batch, feats, label, info = benchdata.make_workload(wl, rows, 1)
checked = bench.build_plan(feats, label, info)
for _ in range(3):
    bench.run_step(batch, checked)
torch.cuda.synchronize()
ev = []
_orig = _lib.call
def _call(name, *a):
    t = time.perf_counter(); _orig(name, *a); ev.append((name, t, time.perf_counter()))
_lib.call = _call
_cpu = torch.Tensor.cpu
def cpu(self, *a, **k):
    t = time.perf_counter(); r = _cpu(self, *a, **k); ev.append(("<.cpu()>", t, time.perf_counter())); return r
torch.Tensor.cpu = cpu
_item = torch.Tensor.item
def item(self):
    t = time.perf_counter(); r = _item(self); ev.append(("<.item()>", t, time.perf_counter())); return r
torch.Tensor.item = item
t0 = time.perf_counter()
bench.run_step(batch, checked)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
for name, a, b in ev:
    if b - a > 2e-5 or name.endswith("_dev"):
        print(f"{(a - t0) * 1e3:8.3f} ms  +{(b - a) * 1e3:7.3f}  {name}")
print(f"host returns at {(t1 - t0) * 1e3:.3f} ms, GPU done at {(t2 - t0) * 1e3:.3f} ms")
