"""Kernel-level timing of the text hashing path at (a fraction of) config-4 size (development aid)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from transmogrifai_b200 import _lib as L

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
ncols = int(sys.argv[2]) if len(sys.argv) > 2 else 20
numf = int(sys.argv[3]) if len(sys.argv) > 3 else (1 << 18)
rng = np.random.default_rng(3)
V = 200_000
# vocabulary of [a-z]{3,10} words laid out in one byte buffer with offsets
wl = rng.integers(3, 11, V)
woff = np.concatenate([[0], np.cumsum(wl)])
wbytes = rng.integers(97, 123, woff[-1]).astype(np.uint8)
d_wbytes = torch.from_numpy(wbytes).cuda(); d_woff = torch.from_numpy(woff).cuda(); d_wl = torch.from_numpy(wl).cuda()
g = torch.Generator(device="cuda").manual_seed(3)
p = 1.0 / torch.arange(1, V + 1, dtype=torch.float64, device="cuda") ** 1.05
cdf = torch.cumsum(p / p.sum(), 0)
cols = []
tot_bytes = 0
for j in range(ncols):
    ntok = torch.poisson(torch.full((n,), 30.0, device="cuda"), generator=g).long()
    null = torch.rand(n, device="cuda", generator=g) < 0.05
    ntok = torch.where(null, torch.zeros_like(ntok), ntok)
    T = int(ntok.sum().item())
    w = torch.searchsorted(cdf, torch.rand(T, dtype=torch.float64, device="cuda", generator=g)).clamp_(max=V - 1)
    tl = d_wl[w] + 1                                   # token + separator
    tstart = torch.cumsum(tl, 0) - tl                  # byte offset of every token
    row_tok0 = torch.cumsum(ntok, 0) - ntok
    total = int(tl.sum().item())
    cell_end = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
    cs = torch.cumsum(tl, 0)
    last = row_tok0 + ntok - 1
    has = ntok > 0
    ends = torch.where(has, cs[last.clamp(min=0)], torch.zeros_like(last))
    cell_end[1:] = torch.cummax(ends, 0).values        # offsets (each cell keeps a trailing space)
    data = torch.full((total,), 32, dtype=torch.uint8, device="cuda")
    # scatter token bytes
    rep = torch.repeat_interleave(torch.arange(T, device="cuda"), d_wl[w])
    inner = torch.arange(rep.numel(), device="cuda") - torch.repeat_interleave(torch.cumsum(d_wl[w], 0) - d_wl[w], d_wl[w])
    data[tstart[rep] + inner] = d_wbytes[d_woff[w][rep] + inner]
    nw = (n + 63) // 64
    okb = torch.zeros(nw * 64, dtype=torch.int64, device="cuda"); okb[:n] = (~null).long()
    wts = (torch.ones(64, dtype=torch.int64, device="cuda") << torch.arange(64, device="cuda")).view(1, 64)
    valid = (okb.view(nw, 64) * wts).sum(dim=1)
    cols.append((data, cell_end, valid))
    tot_bytes += total
    del rep, inner, w, tl, tstart
desc = np.zeros((ncols, 4), dtype=np.int64)
for j, (d, o, v) in enumerate(cols):
    desc[j] = [d.data_ptr(), o.data_ptr(), v.data_ptr(), 0]
d_desc = torch.from_numpy(desc).cuda()
shared = numf * ncols > (1 << 17)
hw = numf if shared else numf * ncols
prefix = np.zeros((ncols, 16), dtype=np.uint8); plen = np.zeros(ncols, dtype=np.int32)
for j in range(ncols):
    s = f"{(j * 37) % numf}_".encode(); prefix[j, :len(s)] = np.frombuffer(s, dtype=np.uint8); plen[j] = len(s)
d_prefix, d_plen = torch.from_numpy(prefix).cuda(), torch.from_numpy(plen).cuda()
ncls = 2
params = np.array([numf, int(shared), 0, 1, 1, 1, hw, ncls], dtype=np.int32)
cls = (torch.rand(n, device="cuda", generator=g) < 0.3).to(torch.uint8)
w_all = hw + ncols
class_sum = torch.zeros(w_all * ncls, dtype=torch.int32, device="cuda"); nnz = torch.zeros(w_all, dtype=torch.int32, device="cuda")
extra = torch.zeros(w_all, dtype=torch.int64, device="cuda"); aux = torch.zeros(3 * w_all, dtype=torch.int32, device="cuda")
flags = torch.zeros(4, dtype=torch.int32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
import ctypes
_need = ctypes.c_size_t(0); L.call("tmog_text_hash_scratch_bytes", n, ctypes.byref(_need))
scratch = torch.empty(_need.value, dtype=torch.uint8, device="cuda")
def stats():
    L.call("tmog_text_hash_stats_dev", d_desc.data_ptr(), ncols, n, d_prefix.data_ptr(), d_plen.data_ptr(), params.ctypes.data,
           cls.data_ptr(), class_sum.data_ptr(), nnz.data_ptr(), extra.data_ptr(), aux.data_ptr(), None, 0.0, None, flags.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
def timeit(f, reps=3):
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
t = timeit(stats)
print(f"rows {n} cols {ncols} bytes {tot_bytes/1e9:.2f} GB  flags {flags.cpu().tolist()}")
print(f"hash-stats: {t:.2f} ms  {n / t * 1e3 / 1e6:.2f} M rows/s  {tot_bytes / t / 1e6:.1f} GB/s  tokens/s {nnz.numel() and float(class_sum.long().sum()) / 4 / t / 1e6:.1f} G")
params2 = params.copy(); params2[7] = 1
newpos = torch.arange(w_all, dtype=torch.int32, device="cuda")
row_start = torch.empty(n, dtype=torch.int64, device="cuda"); row_nnz = torch.empty(n, dtype=torch.int32, device="cuda")
cap = int(class_sum.long().sum().item()) // 4 + 16   # upper bound: every hit distinct
idx = torch.empty(cap, dtype=torch.int32, device="cuda"); val = torch.empty(cap, dtype=torch.int32, device="cuda")
cursor = torch.zeros(1, dtype=torch.int64, device="cuda")
def tr():
    cursor.zero_()
    L.call("tmog_text_hash_transform_dev", d_desc.data_ptr(), ncols, n, d_prefix.data_ptr(), d_plen.data_ptr(), params2.ctypes.data,
           newpos.data_ptr(), row_start.data_ptr(), row_nnz.data_ptr(), idx.data_ptr(), val.data_ptr(), cursor.data_ptr(), cap, flags.data_ptr(), scratch.data_ptr(), scratch.numel(), st)
t2 = timeit(tr)
print("deferred rows:", scratch[24:32].view(torch.int32).tolist())
print(f"hash-transform: {t2:.2f} ms  {n / t2 * 1e3 / 1e6:.2f} M rows/s  nnz {int(cursor.item())} cap {cap} flags {flags.cpu().tolist()}")
