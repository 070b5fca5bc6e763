"""Kernel-level timing of the Real path at config-2 size (development aid; bench.py is the contract)."""
import ctypes as C
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from transmogrifai_b200 import _lib as L

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
ncols = int(sys.argv[2]) if len(sys.argv) > 2 else 100
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.randn((ncols, n), dtype=torch.float64, device="cuda", generator=g)
y = (torch.rand(n, device="cuda", generator=g) < 0.3).double()
vw = torch.randint(-2**63, 2**63 - 1, (ncols, (n + 63) // 64), dtype=torch.int64, device="cuda", generator=g)
nb = C.c_size_t(); L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nb))
scratch = torch.empty(nb.value, dtype=torch.uint8, device="cuda")
shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
part = torch.zeros((ncols + 1) * 8, dtype=torch.float64, device="cuda")
fills = torch.zeros(ncols, dtype=torch.float64, device="cuda")
out = torch.empty((n, 2 * ncols), dtype=torch.float64, device="cuda")
st = torch.cuda.current_stream().cuda_stream

def fit():
    L.call("tmog_real_pilot_shift_dev", x.data_ptr(), n, vw.data_ptr(), vw.shape[1], y.data_ptr(), n, ncols, shifts.data_ptr(), st)
    L.call("tmog_real_fit_partials_dev", x.data_ptr(), n, vw.data_ptr(), vw.shape[1], y.data_ptr(), n, ncols,
           shifts.data_ptr(), scratch.data_ptr(), nb.value, part.data_ptr(), st)
def tr():
    L.call("tmog_real_transform_dev", x.data_ptr(), n, vw.data_ptr(), vw.shape[1], n, ncols, fills.data_ptr(), 1,
           out.data_ptr(), 2 * ncols, 0, st)

def timeit(f, reps=10):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

fit_b = n * (ncols * 8 + ncols / 8 + 8)
tr_b = n * (ncols * 8 + ncols / 8) + n * ncols * 16
tf, tt = timeit(fit), timeit(tr)
print(f"fit: {tf:.3f} ms  {fit_b / tf / 1e6:.1f} GB/s   transform: {tt:.3f} ms  {tr_b / tt / 1e6:.1f} GB/s", flush=True)
import numpy as np
_bits = np.ones((2 * ncols + 63) // 64 * 64, dtype=np.uint8); _bits[14] = 0; _bits[2 * ncols:] = 0
keep = torch.from_numpy(np.packbits(_bits, bitorder="little").view("<i8").copy()).cuda()
out2 = torch.empty((n, 2 * ncols - 1), dtype=torch.float64, device="cuda")
def trs():
    L.call("tmog_real_transform_select_dev", x.data_ptr(), n, vw.data_ptr(), vw.shape[1], n, ncols, fills.data_ptr(), 1,
           keep.data_ptr(), 2 * ncols - 1, out2.data_ptr(), 2 * ncols - 1, 0, st)
ts = timeit(trs)
print(f"transform-select(199): {ts:.3f} ms  {(tr_b - 8 * n) / ts / 1e6:.1f} GB/s", flush=True)
