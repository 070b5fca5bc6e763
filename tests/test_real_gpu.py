"""GPU parity: Real / RealNN / Binary vectorizers + fused SanityChecker column scan vs the CPU oracle, through the
C ABI (device-pointer entry points)."""
import ctypes as C

import numpy as np
import pytest

from oracle import sanity as S
from oracle import vectorizers as V
from tests.util_data import make_real_config, pack_validity

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _pad_cols(x, mult):
    ncols, n = x.shape
    ld = (n + mult - 1) // mult * mult
    out = np.zeros((ncols, ld), dtype=x.dtype)
    out[:, :n] = x
    return out, ld


def run_fit(x, valid, y, fill_with_mean=True, fill_const=0.0, track_nulls=True):
    from transmogrifai_b200 import _lib as L
    ncols, n = x.shape
    xp, ld = _pad_cols(x, 2)
    xd = _dev(xp)
    vw = pack_validity(valid) if valid is not None else None
    vd = _dev(vw.view(np.int64)) if vw is not None else None
    ldv = vw.shape[1] if vw is not None else 0
    yd = _dev(y)
    shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
    partials = torch.zeros((ncols + 1) * L.REAL_NSTAT, dtype=torch.float64, device="cuda")
    nbytes = C.c_size_t()
    L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nbytes))
    scratch = torch.empty(max(nbytes.value, 8), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    vptr = vd.data_ptr() if vd is not None else None
    L.call("tmog_real_pilot_shift_dev", xd.data_ptr(), ld, vptr, ldv, yd.data_ptr(), n, ncols, shifts.data_ptr(), st)
    L.call("tmog_real_fit_partials_dev", xd.data_ptr(), ld, vptr, ldv, yd.data_ptr(), n, ncols, shifts.data_ptr(),
           scratch.data_ptr(), nbytes.value, partials.data_ptr(), st)
    torch.cuda.synchronize()
    p = partials.cpu().numpy()
    s = shifts.cpu().numpy()
    w = ncols * (2 if track_nulls else 1)
    fills = np.zeros(ncols)
    out = {k: np.zeros(w) for k in ("mean", "variance", "min", "max", "cov")}
    lab = np.zeros(5)
    nls = np.zeros(ncols)
    L.call("tmog_real_finish", p.ctypes.data, s.ctypes.data, ncols, int(fill_with_mean), float(fill_const),
           int(track_nulls), fills.ctypes.data, out["mean"].ctypes.data, out["variance"].ctypes.data,
           out["min"].ctypes.data, out["max"].ctypes.data, out["cov"].ctypes.data, lab.ctypes.data, nls.ctypes.data)
    return dict(fills=fills, label=lab, null_label_sum=nls, xd=xd, ld=ld, vptr=vptr, ldv=ldv, keep=(vd, yd), **out)


def run_transform(res, n, ncols, track_nulls=True, out_ld=None, out_col0=0):
    from transmogrifai_b200 import _lib as L
    w = ncols * (2 if track_nulls else 1)
    out_ld = out_ld or w
    out = torch.full((n, out_ld), -7.0, dtype=torch.float64, device="cuda")
    fills = _dev(res["fills"])
    L.call("tmog_real_transform_dev", res["xd"].data_ptr(), res["ld"], res["vptr"], res["ldv"], n, ncols,
           fills.data_ptr(), int(track_nulls), out.data_ptr(), out_ld, out_col0,
           torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return out.cpu().numpy()


def to_lists(x, valid):
    return [[float(v) if ok else None for v, ok in zip(col, vcol)] for col, vcol in zip(x, valid)]


@pytest.mark.parametrize("n,ncols", [(3, 3), (1, 1), (64, 5), (2048, 9), (2049, 33), (10007, 40), (50000, 100)])
def test_real_fit_transform_matches_oracle(n, ncols):
    x, valid, y = make_real_config(n, ncols, seed=100 + n)
    if ncols >= 3:
        valid[2] = False  # an all-null column -> fill 0.0
    res = run_fit(x, valid, y)
    cols = to_lists(x, valid)
    exp_fill = V.real_vectorizer_fit(cols)
    np.testing.assert_allclose(res["fills"], exp_fill, rtol=1e-12, atol=0)     # north_star: fills within 1e-12 rel
    got = run_transform(res, n, ncols)
    exp = V.real_vectorizer_transform(cols, exp_fill, True)
    # values pass through untouched; only the fill carries the 1e-12 tolerance
    np.testing.assert_allclose(got, exp, rtol=1e-12, atol=0)
    assert np.array_equal(got[:, 1::2], exp[:, 1::2])                            # null indicators bit-exact
    if n > 1:
        full = np.concatenate([exp, y[:, None]], axis=1)
        st = S.col_stats(full)
        np.testing.assert_allclose(res["mean"], st.mean[:-1], rtol=1e-12, atol=1e-300)
        np.testing.assert_allclose(res["variance"], st.variance[:-1], rtol=1e-9, atol=1e-18)
        assert np.array_equal(res["min"], st.min[:-1]) and np.array_equal(res["max"], st.max[:-1])
        np.testing.assert_allclose(res["label"], [n, st.mean[-1], st.variance[-1], st.min[-1], st.max[-1]],
                                   rtol=1e-12, atol=1e-300)
        d = full - st.mean
        cov = (d[:, :-1] * d[:, -1:]).sum(axis=0)
        scale = np.sqrt((d[:, :-1] ** 2).sum(axis=0) * (d[:, -1] ** 2).sum())
        np.testing.assert_allclose(res["cov"], cov, rtol=1e-9, atol=1e-9 * max(scale.max(), 1e-300))


@pytest.mark.parametrize("n,ncols,nulls", [(148 * 4096 * 2 + 4096 * 5 + 2 * 1024 + 77, 11, True),
                                           (4096 * 300 + 1024, 3, True), (1024 * 7, 17, False), (1023, 9, True),
                                           (5000, 400, True)])   # 400 columns: the register-streaming kernel
def test_real_fit_partials_over_many_super_tiles(n, ncols, nulls):
    """Raw partials of the fit scan when every CTA walks several super-tiles (the label ring wraps, both column
    buffers of a warp are reused) plus a ragged tail, against numpy on the same shifted sums."""
    from transmogrifai_b200 import _lib as L
    rng = np.random.default_rng(n + ncols)
    x = rng.normal(3.0, 2.0, size=(ncols, n))
    valid = rng.random((ncols, n)) > 0.2 if nulls else None
    y = (rng.random(n) < 0.3).astype(np.float64)
    res = run_fit(x, valid, y)
    xd, vd, yd = res["xd"], res["keep"][0], res["keep"][1]
    shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
    partials = torch.zeros((ncols + 1) * L.REAL_NSTAT, dtype=torch.float64, device="cuda")
    nb = C.c_size_t()
    L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nb))
    scratch = torch.empty(nb.value, dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    L.call("tmog_real_pilot_shift_dev", xd.data_ptr(), res["ld"], res["vptr"], res["ldv"], yd.data_ptr(), n, ncols,
           shifts.data_ptr(), st)
    L.call("tmog_real_fit_partials_dev", xd.data_ptr(), res["ld"], res["vptr"], res["ldv"], yd.data_ptr(), n, ncols,
           shifts.data_ptr(), scratch.data_ptr(), nb.value, partials.data_ptr(), st)
    torch.cuda.synchronize()
    got = partials.cpu().numpy().reshape(ncols + 1, L.REAL_NSTAT)
    s = shifts.cpu().numpy()
    e = y - s[ncols]
    cols = np.concatenate([x, y[None, :]], axis=0)
    ok = np.ones((ncols + 1, n), dtype=bool)
    if valid is not None:
        ok[:ncols] = valid
    d = np.where(ok, cols - s[:, None], 0.0)
    exp = np.stack([ok.sum(axis=1), d.sum(axis=1), (d * d).sum(axis=1), (d * e).sum(axis=1),
                    np.where(ok, e, 0.0).sum(axis=1), np.where(ok, cols, np.inf).min(axis=1),
                    np.where(ok, cols, -np.inf).max(axis=1), np.full(ncols + 1, n)], axis=1)
    assert np.array_equal(got[:, [0, 5, 6, 7]], exp[:, [0, 5, 6, 7]])      # counts, min, max: exact
    scale = np.sqrt(n) * 10.0
    np.testing.assert_allclose(got[:, 1:5], exp[:, 1:5], rtol=1e-11, atol=1e-12 * scale * scale)
    del vd


def test_real_reference_golden_vectors():
    # core/src/test/scala/com/salesforce/op/stages/impl/feature/RealVectorizerTest.scala:118-145,182-214
    x = np.array([[4.0, 0.0, 2.0], [2.0, 2.0, 0.0], [0.0, 0.0, 0.0]])
    valid = np.array([[1, 0, 1], [1, 1, 0], [0, 0, 0]], dtype=bool)
    res = run_fit(x, valid, np.zeros(3))
    assert res["fills"].tolist() == [3.0, 2.0, 0.0]
    got = run_transform(res, 3, 3)
    assert got.tolist() == [[4.0, 0.0, 2.0, 0.0, 0.0, 1.0], [3.0, 1.0, 2.0, 0.0, 0.0, 1.0],
                            [2.0, 0.0, 2.0, 1.0, 0.0, 1.0]]
    got = run_transform(res, 3, 3, track_nulls=False)
    assert got.tolist() == [[4.0, 2.0, 0.0], [3.0, 2.0, 0.0], [2.0, 2.0, 0.0]]


def test_real_transform_into_combined_block():
    n, ncols = 1000, 7
    x, valid, y = make_real_config(n, ncols, seed=5)
    res = run_fit(x, valid, y)
    got = run_transform(res, n, ncols, out_ld=30, out_col0=10)
    exp = V.real_vectorizer_transform(to_lists(x, valid), res["fills"], True)
    assert np.array_equal(got[:, 10:24], exp)
    assert np.all(got[:, :10] == -7.0) and np.all(got[:, 24:] == -7.0)      # neighbours untouched


def test_realnn_and_binary_transform():
    from transmogrifai_b200 import _lib as L
    n, ncols = 777, 5
    rng = np.random.default_rng(3)
    x = rng.standard_normal((ncols, n))
    xp, ld = _pad_cols(x, 2)
    xd = _dev(xp)
    out = torch.zeros((n, ncols), dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    L.call("tmog_real_transform_dev", xd.data_ptr(), ld, None, 0, n, ncols, None, 0, out.data_ptr(), ncols, 0, st)
    assert np.array_equal(out.cpu().numpy(), V.realnn_vectorizer_transform(x.tolist()))
    b = (rng.random((ncols, n)) < 0.4)
    valid = rng.random((ncols, n)) >= 0.2
    bp, ldb = _pad_cols(b.astype(np.uint8), 16)
    bd = _dev(bp)
    vw = pack_validity(valid)
    vd = _dev(vw.view(np.int64))
    out = torch.zeros((n, 2 * ncols), dtype=torch.float64, device="cuda")
    L.call("tmog_binary_transform_dev", bd.data_ptr(), ldb, vd.data_ptr(), vw.shape[1], n, ncols, 0, 1,
           out.data_ptr(), 2 * ncols, 0, st)
    cols = [[bool(v) if ok else None for v, ok in zip(c, vc)] for c, vc in zip(b, valid)]
    assert np.array_equal(out.cpu().numpy(), V.binary_vectorizer_transform(cols))


def test_real_invalid_arguments_raise():
    from transmogrifai_b200 import _lib as L
    x = torch.zeros(10, dtype=torch.float64, device="cuda")
    with pytest.raises(L.TmogInvalidArgument):
        L.call("tmog_real_transform_dev", x.data_ptr(), 3, None, 0, 3, 1, None, 0, x.data_ptr(), 1, 0, None)
    nbytes = C.c_size_t()
    L.call("tmog_real_fit_scratch_bytes", 4, C.byref(nbytes))
    with pytest.raises(L.TmogError, match="scratch too small"):
        L.call("tmog_real_fit_partials_dev", x.data_ptr(), 2, None, 0, x.data_ptr(), 2, 1, x.data_ptr(),
               x.data_ptr(), 8, x.data_ptr(), None)


def test_real_full_size_properties():
    """Config 2 at full size (10M x 100): size-independent checks -- chunk-merge additivity, counts, and the fill
    against a float64 numpy mean of the same data."""
    from transmogrifai_b200 import _lib as L
    n, ncols = 10_000_000, 100
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.randn((ncols, n), dtype=torch.float64, device="cuda", generator=g) * 3.0 + 50.0
    y = (torch.rand(n, device="cuda", generator=g) < 0.3).double()
    vw = torch.randint(-2**63, 2**63 - 1, (ncols, (n + 63) // 64), dtype=torch.int64, device="cuda", generator=g) | \
        torch.randint(-2**63, 2**63 - 1, (ncols, (n + 63) // 64), dtype=torch.int64, device="cuda", generator=g)
    nbytes = C.c_size_t()
    L.call("tmog_real_fit_scratch_bytes", ncols, C.byref(nbytes))
    scratch = torch.empty(nbytes.value, dtype=torch.uint8, device="cuda")
    shifts = torch.zeros(ncols + 1, dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    L.call("tmog_real_pilot_shift_dev", x.data_ptr(), n, vw.data_ptr(), vw.shape[1], y.data_ptr(), n, ncols,
           shifts.data_ptr(), st)

    def partials(r0, r1):
        p = torch.zeros((ncols + 1) * 8, dtype=torch.float64, device="cuda")
        assert r0 % 64 == 0
        L.call("tmog_real_fit_partials_dev", x.data_ptr() + 8 * r0, n, vw.data_ptr() + 8 * (r0 // 64), vw.shape[1],
               y.data_ptr() + 8 * r0, r1 - r0, ncols, shifts.data_ptr(), scratch.data_ptr(), nbytes.value,
               p.data_ptr(), st)
        torch.cuda.synchronize()
        return p.cpu().numpy()

    whole = partials(0, n)
    cut = 6_400_000
    a, b = partials(0, cut), partials(cut, n)
    L.call("tmog_real_partials_merge", a.ctypes.data, b.ctypes.data, ncols)
    w, m = whole.reshape(-1, 8), a.reshape(-1, 8)
    assert np.array_equal(w[:, [0, 5, 6, 7]], m[:, [0, 5, 6, 7]])               # counts, min, max: exact
    np.testing.assert_allclose(w[:, 1:5], m[:, 1:5], rtol=1e-9, atol=1e-6)      # sums: order-of-summation only
    # counts against popcount of the bitmaps, fills against torch float64
    bits = torch.stack([(vw >> k) & 1 for k in range(64)], dim=2).reshape(ncols, -1)[:, :n].bool()
    assert np.array_equal(w[:-1, 0], bits.sum(dim=1).cpu().numpy().astype(np.float64))
    s = shifts.cpu().numpy()
    fills = np.zeros(ncols)
    L.call("tmog_real_finish", whole.ctypes.data, s.ctypes.data, ncols, 1, 0.0, 1, fills.ctypes.data, None, None,
           None, None, None, None, None)
    ref = (torch.where(bits, x, torch.zeros_like(x)).sum(dim=1) / bits.sum(dim=1)).cpu().numpy()
    np.testing.assert_allclose(fills, ref, rtol=1e-12)


@pytest.mark.parametrize("track", [True, False])
def test_real_transform_select_writes_only_kept_columns(track):
    """tmog_real_transform_select_dev == transform followed by SanityCheckerModel's column gather."""
    from transmogrifai_b200 import _lib as L
    n, ncols = 3001, 37
    x, valid, y = make_real_config(n, ncols, seed=9)
    res = run_fit(x, valid, y, track_nulls=track)
    full = V.real_vectorizer_transform(to_lists(x, valid), res["fills"], track)
    rng = np.random.default_rng(1)
    w = full.shape[1]
    for keep in ([0], list(range(w)), sorted(rng.choice(w, size=w // 2, replace=False).tolist()), [w - 1],
                 [i for i in range(w) if i != 14]):
        bits = np.zeros((w + 63) // 64 * 64, dtype=np.uint8)
        bits[keep] = 1
        kd = torch.from_numpy(np.packbits(bits, bitorder="little").view("<i8").copy()).cuda()
        out = torch.full((n, len(keep) + 3), -7.0, dtype=torch.float64, device="cuda")
        fills = _dev(res["fills"])
        L.call("tmog_real_transform_select_dev", res["xd"].data_ptr(), res["ld"], res["vptr"], res["ldv"], n, ncols,
               fills.data_ptr(), int(track), kd.data_ptr(), len(keep), out.data_ptr(), len(keep) + 3, 2,
               torch.cuda.current_stream().cuda_stream)
        got = out.cpu().numpy()
        assert np.array_equal(got[:, 2:2 + len(keep)], full[:, keep])
        assert np.all(got[:, :2] == -7.0) and np.all(got[:, 2 + len(keep):] == -7.0)
