"""Timing of the Gramian kernel of CorrelationLevel.Computed (development aid): python tools/microbench_gram.py [n] [d]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from transmogrifai_b200 import _lib as L
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 200
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn((n, d), dtype=torch.float64, device="cuda", generator=g)
y = torch.randn(n, dtype=torch.float64, device="cuda", generator=g)
shifts = torch.zeros(d + 1, dtype=torch.float64, device="cuda")
gram = torch.zeros((d + 1) * (d + 1), dtype=torch.float64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
def run():
    gram.zero_()
    L.call("tmog_dense_gram_dev", x.data_ptr(), d, n, d, y.data_ptr(), shifts.data_ptr(), gram.data_ptr(), st)
for _ in range(2): run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): run()
e1.record(); torch.cuda.synchronize()
t = e0.elapsed_time(e1) / 3
D = d + 1
print(f"gram {n} x {D}: {t:.2f} ms  {n * D * (D + 1) / t / 1e9:.2f} TFLOP/s (upper triangle)  {n * d * 8 / t / 1e6:.0f} GB/s")
m = min(n, 20000)
xs = torch.cat([x[:m], y[:m, None]], dim=1)
gram.zero_()
L.call("tmog_dense_gram_dev", x.data_ptr(), d, m, d, y.data_ptr(), shifts.data_ptr(), gram.data_ptr(), st)
ref = (xs.T @ xs)
got = gram.view(D, D)
iu = torch.triu_indices(D, D)
err = ((got[iu[0], iu[1]] - ref[iu[0], iu[1]]).abs() / ref[iu[0], iu[1]].abs().clamp_min(1.0)).max().item()
print("max rel err vs torch fp64 matmul:", err)
