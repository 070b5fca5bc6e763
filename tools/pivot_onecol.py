"""One pivot-count launch on a single high-cardinality column (for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from transmogrifai_b200 import _lib as L
n, K, ncls = 20_000_000, int(os.environ.get("K", "100000")), 2
g = torch.Generator(device="cuda").manual_seed(2)
p = 1.0 / torch.arange(1, K + 1, dtype=torch.float64, device="cuda") ** 1.1
cdf = torch.cumsum(p / p.sum(), 0)
codes = torch.searchsorted(cdf, torch.rand(n, dtype=torch.float64, device="cuda", generator=g)).clamp_(max=K - 1).int()
cls = (torch.rand(n, device="cuda", generator=g) < 0.3).to(torch.uint8)
desc = torch.tensor([[K + 1, 0, 0, K]], dtype=torch.int64, device="cuda")
counts = torch.zeros((K + 1) * ncls, dtype=torch.int32, device="cuda")
bad = torch.zeros(1, dtype=torch.int32, device="cuda")
for _ in range(3):
    L.call("tmog_pivot_count_dev", codes.data_ptr(), n, n, 1, cls.data_ptr(), ncls, desc.data_ptr(), None, K + 1,
           counts.data_ptr(), bad.data_ptr(), None)
torch.cuda.synchronize()
print("ok", int(counts.sum().item()) // 3, n)
