"""Fit-kernel timing at config-2 size (development aid).  TMOG_FIT_VARIANT picks a kernel configuration when the
library was built with the variant switch; prints the time of tmog_real_fit_partials_dev alone and a checksum."""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from transmogrifai_b200 import _lib as L

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
ncols = int(sys.argv[2]) if len(sys.argv) > 2 else 100
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn((ncols, n), dtype=torch.float64, device="cuda", generator=g)
y = (torch.rand(n, device="cuda", generator=g) < 0.3).double()
vw = torch.full((ncols, (n + 63) // 64), -1, dtype=torch.int64, device="cuda")
r = torch.randint(0, 2**62, (ncols, (n + 63) // 64, 3), dtype=torch.int64, device="cuda", generator=g)
sparse = ~(r[..., 0] & r[..., 1] & r[..., 2]) | (r[..., 0] >> 1)      # mostly ones
vw = sparse
nb = C.c_size_t(); L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nb))
scratch = torch.empty(nb.value, dtype=torch.uint8, device="cuda")
shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
part = torch.zeros((ncols + 1) * 8, dtype=torch.float64, device="cuda")
st = torch.cuda.current_stream().cuda_stream
L.call("tmog_real_pilot_shift_dev", x.data_ptr(), n, vw.data_ptr(), vw.shape[1], y.data_ptr(), n, ncols, shifts.data_ptr(), st)

def fit():
    L.call("tmog_real_fit_partials_dev", x.data_ptr(), n, vw.data_ptr(), vw.shape[1], y.data_ptr(), n, ncols,
           shifts.data_ptr(), scratch.data_ptr(), nb.value, part.data_ptr(), st)

for _ in range(3): fit()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): fit()
e1.record(); torch.cuda.synchronize()
t = e0.elapsed_time(e1) / 20
b = n * (ncols * 8 + ncols / 8 + 8)
p = part.cpu().numpy().reshape(-1, 8)
print(f"variant {os.environ.get('TMOG_FIT_VARIANT', '0')}: fit {t:.3f} ms  {b / t / 1e6:.1f} GB/s  "
      f"checksum {p[:, 0].sum():.0f} {p[:, 2].sum():.6f} {p[:, 3].sum():.9f}", flush=True)
