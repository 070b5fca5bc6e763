"""Kernel-level timing of the pivot path at config-3 size (development aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from transmogrifai_b200 import _lib as L

n = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
ncols = int(sys.argv[2]) if len(sys.argv) > 2 else 50
ncls = 2
g = torch.Generator(device="cuda").manual_seed(2)
ld = (n + 3) // 4 * 4
cards = np.unique(np.round(np.logspace(1, 5, ncols)).astype(np.int64), return_index=False)
cards = np.round(np.logspace(1, 5, ncols)).astype(np.int64)
codes = torch.empty((ncols, ld), dtype=torch.int32, device="cuda")
for j, K in enumerate(cards):
    p = 1.0 / torch.arange(1, K + 1, dtype=torch.float64, device="cuda") ** 1.1
    cdf = torch.cumsum(p / p.sum(), 0)
    u = torch.rand(ld, dtype=torch.float64, device="cuda", generator=g)
    c = torch.searchsorted(cdf, u).clamp_(max=int(K) - 1).int()
    null = torch.rand(ld, device="cuda", generator=g) < 0.05
    codes[j] = torch.where(null, torch.full_like(c, -1), c)
cls = (torch.rand(ld, device="cuda", generator=g) < 0.3).to(torch.uint8)
nids = cards + 1
slot_off = np.concatenate([[0], np.cumsum(cards)[:-1]])
cnt_off = np.concatenate([[0], np.cumsum(nids * ncls)[:-1]])
desc = torch.from_numpy(np.ascontiguousarray(np.stack([nids, slot_off, cnt_off, cards], axis=1))).cuda()
slots = torch.from_numpy(np.concatenate([np.arange(K, dtype=np.int32) for K in cards])).cuda()
counts = torch.zeros(int((nids * ncls).sum()), dtype=torch.int32, device="cuda")
bad = torch.zeros(1, dtype=torch.int32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
SL = None if os.environ.get("NOSLOT", "1") == "1" else slots.data_ptr()
pos = torch.from_numpy(np.concatenate([np.minimum(np.arange(K + 1), 21).astype(np.int32) + 22 * j for j, K in enumerate(cards)])).cuda()
PADW = int(os.environ.get("PADW", "8")); out = torch.empty((n, (ncols + PADW - 1) // PADW * PADW), dtype=torch.int32, device="cuda")

def count():
    L.call("tmog_pivot_count_dev", codes.data_ptr(), ld, n, ncols, cls.data_ptr(), ncls, desc.data_ptr(), SL,
           int(nids.max()), counts.data_ptr(), bad.data_ptr(), st)
def tr():
    L.call("tmog_pivot_transform_dev", codes.data_ptr(), ld, n, ncols, desc.data_ptr(), pos.data_ptr(), out.data_ptr(),
           out.shape[1], 0, st)
def timeit(f, reps=5):
    for _ in range(2): f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
tc, tt = timeit(count), timeit(tr)
cb = n * (ncols * 4 + 1); tb = n * ncols * 8
print(f"count: {tc:.3f} ms {cb / tc / 1e6:.1f} GB/s   transform: {tt:.3f} ms {tb / tt / 1e6:.1f} GB/s", flush=True)
# sanity: totals
tot = counts.cpu().numpy().astype(np.int64) & 0xffffffff
print("rows counted per column / n:", float(tot[:int(nids[0] * ncls)].sum()) / n / 7, "nulls", float(tot[:2].sum()) / n / 7)
# per-column timing of the count kernel
for j in [0, 10, 25, 40, 49]:
    d1 = desc[j:j + 1].clone(); d1[0, 1] = int(slot_off[j]); d1[0, 2] = int(cnt_off[j])
    f = lambda: L.call("tmog_pivot_count_dev", codes[j].data_ptr(), ld, n, 1, cls.data_ptr(), ncls, d1.data_ptr(), SL,
                       int(nids[j]), counts.data_ptr(), bad.data_ptr(), st)
    t = timeit(f, 3)
    print(f"  col {j} K={cards[j]}: {t:.3f} ms  {n * 5 / t / 1e6:.1f} GB/s", flush=True)
