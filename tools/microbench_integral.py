"""Kernel-level timing of the Integral path (development aid): mode tables, fit scan and transform over int64 columns."""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from transmogrifai_b200 import _lib as L

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
ncols = int(sys.argv[2]) if len(sys.argv) > 2 else 20
g = torch.Generator(device="cuda").manual_seed(1)
st = torch.cuda.current_stream().cuda_stream
y = (torch.rand(n, device="cuda", generator=g) < 0.3).double()
vw = torch.randint(-2**63, 2**63 - 1, (ncols, (n + 63) // 64), dtype=torch.int64, device="cuda", generator=g) | \
    torch.randint(-2**63, 2**63 - 1, (ncols, (n + 63) // 64), dtype=torch.int64, device="cuda", generator=g)


def timeit(f, reps=5):
    for _ in range(2): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for name, hi in (("100 distinct values", 100), ("65 k distinct values", 65_000), ("Zipf-like (squares)", 0)):
    if hi:
        x = torch.randint(0, hi, (ncols, n), dtype=torch.int64, device="cuda", generator=g)
    else:
        x = (torch.rand((ncols, n), device="cuda", generator=g) ** 8 * 1000).long()
    cap = 1 << 18
    nb = C.c_size_t(); L.call("tmog_integral_mode_scratch_bytes", ncols, cap, C.byref(nb))
    scratch = torch.empty(nb.value, dtype=torch.uint8, device="cuda")
    modes = torch.zeros(ncols, dtype=torch.int64, device="cuda")
    counts = torch.zeros(ncols, dtype=torch.int64, device="cuda")
    f = lambda: L.call("tmog_integral_mode_dev", x.data_ptr(), n, vw.data_ptr(), vw.shape[1], n, ncols, cap,
                       scratch.data_ptr(), nb.value, modes.data_ptr(), counts.data_ptr(), st)
    t = timeit(f)
    b = n * ncols * (8 + 1 / 8)
    print(f"mode tables, {name}: {t:.3f} ms  {b / t / 1e6:.1f} GB/s", flush=True)

nbf = C.c_size_t(); L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nbf))
fs = torch.empty(nbf.value, dtype=torch.uint8, device="cuda")
shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
part = torch.zeros((ncols + 1) * 8, dtype=torch.float64, device="cuda")
L.call("tmog_integral_pilot_shift_dev", x.data_ptr(), n, vw.data_ptr(), vw.shape[1], y.data_ptr(), n, ncols,
       shifts.data_ptr(), st)
t = timeit(lambda: L.call("tmog_integral_fit_partials_dev", x.data_ptr(), n, vw.data_ptr(), vw.shape[1], y.data_ptr(), n,
                          ncols, shifts.data_ptr(), fs.data_ptr(), nbf.value, part.data_ptr(), st))
print(f"fit scan (int64): {t:.3f} ms  {n * (ncols * 8 + ncols / 8 + 8) / t / 1e6:.1f} GB/s", flush=True)
fills = torch.zeros(ncols, dtype=torch.float64, device="cuda")
out = torch.empty((n, 2 * ncols), dtype=torch.float64, device="cuda")
t = timeit(lambda: L.call("tmog_integral_transform_dev", x.data_ptr(), n, vw.data_ptr(), vw.shape[1], n, ncols,
                          fills.data_ptr(), 1, out.data_ptr(), 2 * ncols, 0, st))
print(f"transform (int64): {t:.3f} ms  {n * (ncols * 8 + ncols / 8 + ncols * 16) / t / 1e6:.1f} GB/s", flush=True)
