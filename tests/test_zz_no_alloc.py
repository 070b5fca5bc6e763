"""The allocation-free fit entry point (tmog_real_fit_partials_noalloc_dev).

It serves hosts that cannot call cudaMalloc -- e.g. a JNI shim in a Spark executor whose GPU memory is owned by
another component's pool -- by keeping the kernel's per-CTA scratch in an array built into the library.  The test
drives it through the C ABI exactly as such a host would (inputs and outputs only, no scratch argument) and checks
it against the ordinary entry point, which takes caller-provided scratch.  Two calls in a row, on different data,
because the built-in array is shared by every call and still holds the previous call's contents.

This file sorts last on purpose: a run that stops at the first failure has then already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _inputs(n, ncols, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    ld = (n + 1) // 2 * 2
    x = torch.randn((ncols, ld), dtype=torch.float64, device="cuda", generator=g) * 4.0 + 10.0
    y = (torch.rand(ld, device="cuda", generator=g) < 0.3).double()
    nw = (n + 63) // 64
    a = torch.randint(-2 ** 63, 2 ** 63 - 1, (ncols, nw), dtype=torch.int64, device="cuda", generator=g)
    b = torch.randint(-2 ** 63, 2 ** 63 - 1, (ncols, nw), dtype=torch.int64, device="cuda", generator=g)
    return x, ld, a | b, nw, y          # ~25 % nulls


def _both(n, ncols, seed):
    from transmogrifai_b200 import _lib as L
    x, ld, vw, nw, y = _inputs(n, ncols, seed)
    st = torch.cuda.current_stream().cuda_stream
    shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
    L.call("tmog_real_pilot_shift_dev", x.data_ptr(), ld, vw.data_ptr(), nw, y.data_ptr(), n, ncols,
           shifts.data_ptr(), st)
    nb = C.c_size_t()
    L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nb))
    scratch = torch.empty(nb.value, dtype=torch.uint8, device="cuda")
    ref = torch.zeros((ncols + 1) * L.REAL_NSTAT, dtype=torch.float64, device="cuda")
    L.call("tmog_real_fit_partials_dev", x.data_ptr(), ld, vw.data_ptr(), nw, y.data_ptr(), n, ncols,
           shifts.data_ptr(), scratch.data_ptr(), nb.value, ref.data_ptr(), st)
    got = torch.full_like(ref, float("nan"))
    L.call("tmog_real_fit_partials_noalloc_dev", x.data_ptr(), ld, vw.data_ptr(), nw, y.data_ptr(), n, ncols,
           shifts.data_ptr(), got.data_ptr(), st)
    torch.cuda.synchronize()
    # the same sums in numpy from the raw inputs, so that a fault shared by both entry points would show too
    xs, ys, sh = x[:, :n].cpu().numpy(), y[:n].cpu().numpy(), shifts.cpu().numpy()
    ok = np.unpackbits(vw.cpu().numpy().view(np.uint8), axis=1, bitorder="little")[:, :n].astype(bool)
    e = ys - sh[ncols]
    exp = np.zeros((ncols + 1, L.REAL_NSTAT))
    for j in range(ncols):
        d = xs[j][ok[j]] - sh[j]
        exp[j] = [ok[j].sum(), d.sum(), (d * d).sum(), (d * e[ok[j]]).sum(), e[ok[j]].sum(),
                  xs[j][ok[j]].min(), xs[j][ok[j]].max(), n]
    exp[ncols] = [n, e.sum(), (e * e).sum(), (e * e).sum(), e.sum(), ys.min(), ys.max(), n]
    return (got.cpu().numpy().reshape(-1, L.REAL_NSTAT), ref.cpu().numpy().reshape(-1, L.REAL_NSTAT), exp)


def test_noalloc_fit_matches_the_scratch_taking_entry_point_twice_in_a_row():
    # second call: other shape and data, so stale scratch from the first call would show
    for n, ncols, seed in ((300_001, 40, 1), (70_000, 9, 2)):
        got, ref, exp = _both(n, ncols, seed)
        assert np.array_equal(got[:, [0, 5, 6, 7]], ref[:, [0, 5, 6, 7]])          # counts, min, max, rows: exact
        np.testing.assert_allclose(got[:, 1:5], ref[:, 1:5], rtol=1e-9, atol=1e-6)  # sums: as test_real_gpu.py
        assert np.array_equal(got[:, [0, 5, 6, 7]], exp[:, [0, 5, 6, 7]])          # and against numpy on the raw inputs
        np.testing.assert_allclose(got[:, 1:5], exp[:, 1:5], rtol=1e-9, atol=1e-6)


def test_noalloc_fit_rejects_blocks_beyond_its_built_in_scratch():
    from transmogrifai_b200 import _lib as L
    ncols, n = 2000, 64           # far more columns than 4 MiB of per-CTA partials can hold
    x = torch.zeros((ncols, n), dtype=torch.float64, device="cuda")
    y = torch.zeros(n, dtype=torch.float64, device="cuda")
    shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
    out = torch.zeros((ncols + 1) * L.REAL_NSTAT, dtype=torch.float64, device="cuda")
    with pytest.raises(L.TmogInvalidArgument, match="built-in scratch"):
        L.call("tmog_real_fit_partials_noalloc_dev", x.data_ptr(), n, None, 0, y.data_ptr(), n, ncols,
               shifts.data_ptr(), out.data_ptr(), None)
