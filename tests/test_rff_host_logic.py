"""RawFeatureFilter's host logic without a GPU: computeFeatureStats / generateFilteredRaw (transmogrifai_b200/filters.py)
with the six kernel calls replaced by numpy restatements of what csrc/rff.cu and the tokenise + hash kernels compute
(reading the same tensors through the same pointers), against the oracle.  What this pins is everything around the
kernels: grouping features by block, accumulation over row chunks, the all-empty-feature rule, array lengths,
summaryInfo, correlation assembly, exclusions.  The kernels themselves are checked on the GPU (tests/test_rff_gpu.py)."""
import ctypes

import numpy as np
import pytest

from oracle import raw_feature_filter as O
from oracle import text as OT

torch = pytest.importorskip("torch")


def _words(ptr, count):
    return np.ctypeslib.as_array((ctypes.c_uint64 * count).from_address(int(ptr)))


def _valid_mask(words, n):
    if words is None:
        return np.ones(n, dtype=bool)
    w = np.asarray(words).astype(np.uint64)
    return np.unpackbits(w.view(np.uint8), bitorder="little")[:n].astype(bool)


class _Emulated:
    """numpy stand-ins of RawFeatureFilter's _k_* methods (bound onto an instance by `emulate`)."""

    @staticmethod
    def summary(self, sub, n, out):
        o = out.numpy()
        for c in range(len(sub.names)):
            ok = _valid_mask(None if sub.valid is None else sub.valid[c].numpy(), n)
            v = sub.values[c].numpy()[:n][ok].astype(np.float64)
            if len(v):
                o[c, 0] = min(o[c, 0], np.nanmin(v) if not np.isnan(v).all() else np.inf)
                o[c, 1] = max(o[c, 1], np.nanmax(v) if not np.isnan(v).all() else -np.inf)
                o[c, 2] += v.sum()
                o[c, 3] += len(v)
                o[c, 4] = max(o[c, 4], float(np.isnan(v).any()))

    @staticmethod
    def hist(self, sub, n, splits, mode, same, hist, vsum, nonnull):
        bins = self.bins
        for c in range(len(sub.names)):
            ok = _valid_mask(None if sub.valid is None else sub.valid[c].numpy(), n)
            v = sub.values[c].numpy()[:n][ok].astype(np.float64)
            nonnull[c] += len(v)
            sp = splits[c].numpy()
            for x in v:
                if int(mode[c]) == 1:
                    hit = np.nonzero((sp[:-1] <= x) & (x < sp[1:]))[0]
                    b = int(hit[0]) if len(hit) and np.isfinite(x) else bins - 1
                    hist[c, b] += 1
                elif int(mode[c]) == 2:
                    hist[c, 0 if x == float(same[c]) else 1] += 1
                else:
                    vsum[c] += x

    @staticmethod
    def text_rows(self, sub, n):
        bins = self.bins
        starts, nnz, idx, val = [], [], [], []
        for r in range(n):
            starts.append(len(idx))
            for c in range(len(sub.names)):
                ok = _valid_mask(None if sub.valid[c] is None else sub.valid[c].numpy(), n)
                if not ok[r]:
                    continue
                off = sub.offsets[c].numpy()
                s = bytes(sub.data[c].numpy()[off[r]:off[r + 1]]).decode("utf-8")
                ii, vv = OT.hashing_tf_transform(OT.tokenize(s), bins, binary=False)
                for i, t in sorted(zip(ii, vv)):
                    idx.append(c * bins + int(i))
                    val.append(int(t))
            nnz.append(len(idx) - starts[-1])
        mk = lambda a, dt: torch.tensor(a if a else [0], dtype=dt)   # noqa: E731
        return mk(starts, torch.int64), mk(nnz, torch.int32), mk(idx, torch.int32), mk(val, torch.int32)

    @staticmethod
    def rows_hist(self, rows, n, ncols, valid_ptrs, weight, hist, tok):
        bins = self.bins
        row_start, row_nnz, idx, val = (t.numpy() for t in rows)
        h, t = hist.numpy().reshape(-1), tok.numpy()
        masks = []
        for c in range(ncols):
            p = 0 if valid_ptrs is None else int(valid_ptrs[c])
            masks.append(_valid_mask(None if p == 0 else _words(p, (n + 63) // 64), n))
        for r in range(n):
            w = 1 if weight is None else int(weight[r])
            if w == 0:
                continue
            per_col = np.zeros(ncols, dtype=np.int64)
            for p in range(row_start[r], row_start[r] + row_nnz[r]):
                if idx[p] < ncols * bins:
                    h[idx[p]] += w * val[p]
                    per_col[idx[p] // bins] += val[p]
            for c in range(ncols):
                if masks[c][r]:
                    t[c, 0] = min(t[c, 0], per_col[c])
                    t[c, 1] = max(t[c, 1], per_col[c])
                    t[c, 2] += w

    @staticmethod
    def code_count(self, sub, n, col_off, ncodes, counts, bad, y, y_valid, shift, null_y):
        for c in range(len(sub.names)):
            codes = sub.codes[c].numpy()[:n]
            for r, code in enumerate(codes):
                if code >= int(ncodes[c]):
                    bad[0] = 1
                elif code >= 0:
                    counts[int(col_off[c]) + int(code)] += 1
                if y is not None and (code < 0 or code >= int(ncodes[c])):
                    ok = _valid_mask(None if y_valid is None else y_valid.numpy(), n)[r]
                    null_y[c] += (float(y[r]) if ok else 0.0) - shift

    @staticmethod
    def null_label(self, valid_ptrs, ncols, n, y, y_valid, shift, out):
        ok_y = _valid_mask(None if y_valid is None else y_valid.numpy(), n)
        d = np.where(ok_y, y.numpy()[:n], 0.0) - shift
        o = out.numpy()
        for c in range(ncols):
            p = int(valid_ptrs[c])
            if p:
                o[c] += d[~_valid_mask(_words(p, (n + 63) // 64), n)].sum()
        o[ncols] += d.sum()
        o[ncols + 1] += (d * d).sum()


def emulate(f, monkeypatch):
    import types
    from transmogrifai_b200 import runtime as R
    monkeypatch.setattr(R, "require_cuda", lambda: None)
    f._device = "cpu"
    for name, fn in (("_k_summary", _Emulated.summary), ("_k_hist", _Emulated.hist),
                     ("_k_text_rows", _Emulated.text_rows), ("_k_rows_hist", _Emulated.rows_hist),
                     ("_k_code_count", _Emulated.code_count), ("_k_null_label", _Emulated.null_label)):
        setattr(f, name, types.MethodType(fn, f))
    return f


def make_data(n, seed, shift=False):
    """Mixed raw features with the cases the rules look at: an all-empty feature, a constant one, one whose emptiness
    follows the response, a scoring set that drifts."""
    from transmogrifai_b200 import features as F
    rng = np.random.default_rng(seed)
    y = (rng.random(n) < 0.4).astype(np.float64)
    opt = lambda v, p: None if rng.random() < p else v   # noqa: E731
    words = ["alpha", "beta", "gamma", "delta", "the", "of", "Epsilon", "zeta9", "eta's", "x"]
    cols = {
        "y": (F.REALNN, y.tolist()),
        "r0": (F.REAL, [opt(float(rng.normal(5 if shift else 0, 1)), 0.1) for _ in range(n)]),
        "r1": (F.REAL, [None if yy > 0 and rng.random() < 0.9 else float(rng.normal()) for yy in y]),    # leaks the label
        "rnull": (F.REAL, [None] * n),
        "rconst": (F.REAL, [opt(3.5, 0.3) for _ in range(n)]),
        "rfull": (F.REAL, [7.0] * n),
        "rnan": (F.REAL, [opt(float("nan") if rng.random() < 0.02 else float(rng.normal()), 0.1) for _ in range(n)]),
        "rinf": (F.REAL, [opt(float("inf") if i == 17 else float(rng.normal()), 0.1) for i in range(n)]),
        "rwide": (F.REAL, [opt(float(rng.normal(0, 3 if shift else 1)) * 1e6, 0.1) for _ in range(n)]),
        "i0": (F.INTEGRAL, [opt(int(rng.integers(-5, 50)), 0.6 if shift else 0.2) for _ in range(n)]),
        "b0": (F.BINARY, [opt(bool(rng.random() < 0.3), 0.05) for _ in range(n)]),
        "t0": (F.TEXT, [opt(" ".join(words[i] for i in rng.integers(0, 5 if shift else 10, rng.integers(0, 7))), 0.15)
                        for _ in range(n)]),
        "t1": (F.TEXT, [opt(["red", "Green", "blue!", ""][int(rng.integers(4))], 0.5) for _ in range(n)]),
        "p0": (F.PICKLIST, [opt(["x", "y z", "Z", "w-w", ""][int(rng.integers(5))], 0.25) for _ in range(n)]),
        "m0": (F.MULTIPICKLIST, [opt([["a b", "c", "D.", "e", "a b"][int(k)] for k in rng.integers(0, 5, rng.integers(0, 4))],
                                     0.2) for _ in range(n)]),
    }
    return cols


def features_of(cols):
    from transmogrifai_b200 import features as F
    return [F.Feature(nm, tp, nm == "y") for nm, (tp, _v) in cols.items()]


def chunked(cols, sizes):
    from transmogrifai_b200.columnar import ColumnarBatch
    bounds = np.concatenate([[0], np.cumsum(sizes)])
    return [ColumnarBatch.from_columns({k: (t, v[a:b]) for k, (t, v) in cols.items()}, response="y", pin=False)
            for a, b in zip(bounds[:-1], bounds[1:])]


def check_info(got, exp, n):
    """AllFeatureInformation of the product against the oracle's."""
    for gs, es in ((got.responseSummaries, exp.response_summaries), (got.predictorSummaries, exp.predictor_summaries)):
        assert list(gs) == list(es)
        for key in es:
            g, e = gs[key], es[key]
            assert (g.min == e.min or (np.isnan(g.min) and np.isnan(e.min))) and \
                (g.max == e.max or (np.isnan(g.max) and np.isnan(e.max))) and g.count == e.count, key
            assert (np.isnan(g.sum) and np.isnan(e.sum)) or g.sum == e.sum or \
                abs(g.sum - e.sum) <= 1e-9 * max(1.0, abs(e.sum)), key
    for gd, ed in ((got.responseDistributions, exp.response_distributions),
                   (got.predictorDistributions, exp.predictor_distributions)):
        assert [d.name for d in gd] == [d.name for d in ed]
        for g, e in zip(gd, ed):
            assert (g.count, g.nulls, g.type) == (e.count, e.nulls, e.type), g.name
            assert list(g.distribution) == list(e.distribution), g.name
            assert len(g.summaryInfo) == len(e.summary_info), g.name
            np.testing.assert_allclose(g.summaryInfo, e.summary_info, rtol=1e-12, atol=0, equal_nan=True,
                                       err_msg=g.name)
    assert list(got.correlationInfo) == list(exp.correlation_info)
    for rk, per in exp.correlation_info.items():
        assert list(got.correlationInfo[rk]) == list(per)
        for pk, e in per.items():
            g = got.correlationInfo[rk][pk]
            assert (np.isnan(e) and np.isnan(g)) or abs(g - e) < 1e-9, (rk, pk, g, e)


@pytest.mark.parametrize("bins,sizes", [(10, [300]), (100, [100, 1, 150, 49]), (3, [120, 180]), (2, [90, 30])])
def test_feature_stats_and_exclusions_match_the_oracle(monkeypatch, bins, sizes):
    from transmogrifai_b200.filters import RawFeatureFilter
    n = sum(sizes)
    train, score = make_data(n, 3), make_data(n, 4, shift=True)
    feats = features_of(train)
    f = emulate(RawFeatureFilter(bins=bins, minFill=0.2, maxFillDifference=0.3, maxFillRatioDiff=2.0,
                                 maxJSDivergence=0.5, maxCorrelation=0.6, minScoringRows=10), monkeypatch)
    tr_chunks, sc_chunks = chunked(train, sizes), chunked(score, sizes)
    got_t = f.computeFeatureStats(tr_chunks if len(sizes) > 1 else tr_chunks[0], feats, "Training")
    exp_t = O.compute_feature_stats(train, ["y"], bins, "Training")
    check_info(got_t, exp_t, n)
    assert ("rnull", None) not in got_t.predictorSummaries           # never a value: no summary, no distribution
    got_s = f.computeFeatureStats(sc_chunks, feats, "Scoring", got_t)
    exp_s = O.compute_feature_stats(score, ["y"], bins, "Scoring", exp_t)
    check_info(got_s, exp_s, n)
    out = f.generateFilteredRaw(tr_chunks, feats, sc_chunks)
    p = O.FilterParams(bins=bins, min_fill=0.2, max_fill_difference=0.3, max_fill_ratio_diff=2.0,
                       max_js_divergence=0.5, max_correlation=0.6, min_scoring_rows=10)
    _ts, _ss, m, r, drop, keys = O.generate_filtered_raw(train, score, ["y"], p)
    assert sorted(x.name for x in out.featuresToDrop) == sorted(drop) and out.mapKeysToDrop == keys == {}
    assert len(drop) >= 2 and "r1" in drop                            # the rules bite on this data
    for g, e in zip(out.rawFeatureFilterResults.exclusionReasons, r):
        assert (g.name, g.trainingUnfilledState, g.trainingNullLabelLeaker, g.scoringUnfilledState,
                g.jsDivergenceMismatch, g.fillRateDiffMismatch, g.fillRatioDiffMismatch, g.excluded) == \
            (e.name, e.training_unfilled_state, e.training_null_label_leaker, e.scoring_unfilled_state,
             e.js_divergence_mismatch, e.fill_rate_diff_mismatch, e.fill_ratio_diff_mismatch, e.excluded)
    for g, e in zip(out.rawFeatureFilterResults.rawFeatureFilterMetrics, m):
        for a, b in ((g.trainingFillRate, e.training_fill_rate), (g.scoringFillRate, e.scoring_fill_rate),
                     (g.jsDivergence, e.js_divergence), (g.fillRateDiff, e.fill_rate_diff),
                     (g.fillRatioDiff, e.fill_ratio_diff),
                     (g.trainingNullLabelAbsoluteCorr, e.training_null_label_absolute_corr)):
            assert (a is None and b is None) or (np.isnan(a) and np.isnan(b)) or abs(a - b) < 1e-9 or a == b, g.name
    kept = {nm for c in out.cleanedData for b in c.blocks for nm in b.names}
    assert kept == set(train) - set(drop)
    assert len(out.trainingFeatureDistributions) == len(out.scoringFeatureDistributions) == len(train) - 1


def test_too_few_scoring_rows_and_requirements(monkeypatch):
    from transmogrifai_b200.filters import RawFeatureFilter
    train = make_data(200, 8)
    feats = features_of(train)
    f = emulate(RawFeatureFilter(bins=10, minScoringRows=500), monkeypatch)
    out = f.generateFilteredRaw(chunked(train, [200])[0], feats, chunked(make_data(200, 9, True), [200])[0])
    assert out.scoringFeatureDistributions == []                      # below minScoringRows: training checks only
    assert all(m.scoringFillRate is None and m.jsDivergence is None
               for m in out.rawFeatureFilterResults.rawFeatureFilterMetrics)
    for kw, msg in (({"bins": 1}, "Invalid bin size 1"), ({"minFill": 1.5}, "Invalid minFill size 1.5"),
                    ({"maxJSDivergence": -0.1}, "Invalid maxJSDivergence"), ({"minScoringRows": -1}, "minRowsForScoringSet")):
        with pytest.raises(ValueError, match=msg):
            RawFeatureFilter(**kw)
    g = emulate(RawFeatureFilter(bins=10, minFill=1.0, maxFillDifference=0.0), monkeypatch)
    only_sparse = {k: v for k, v in train.items() if k in ("y", "r0", "t1")}
    with pytest.raises(ValueError, match="dropped all of your features"):
        g.generateFilteredRaw(chunked(only_sparse, [200])[0], features_of(only_sparse))


def _rank_worker(rank, size, port, q):
    """Every rank passes its own rows (two chunks each); the summaries merge and the sums add (gloo, CPU)."""
    import os
    import types
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=size)
    try:
        from transmogrifai_b200 import runtime as R
        from transmogrifai_b200.filters import RawFeatureFilter

        class _Patch:
            def setattr(self, obj, name, value):
                setattr(obj, name, value)

        n, bins = 400, 20
        train, score = make_data(n, 13), make_data(n, 14, shift=True)
        feats = features_of(train)
        f = emulate(RawFeatureFilter(bins=bins, minFill=0.2, maxFillDifference=0.3, maxFillRatioDiff=2.0,
                                     maxJSDivergence=0.5, maxCorrelation=0.6, minScoringRows=10), _Patch())
        lo, hi = (0, 150) if rank == 0 else (150, n)
        mine = lambda cols: chunked({k: (t, v[lo:hi]) for k, (t, v) in cols.items()},   # noqa: E731
                                    [(hi - lo) // 3, (hi - lo) - (hi - lo) // 3])
        got_t = f.computeFeatureStats(mine(train), feats, "Training")
        exp_t = O.compute_feature_stats(train, ["y"], bins, "Training")
        check_info(got_t, exp_t, n)
        got_s = f.computeFeatureStats(mine(score), feats, "Scoring", got_t)
        check_info(got_s, O.compute_feature_stats(score, ["y"], bins, "Scoring", exp_t), n)
        out = f.generateFilteredRaw(mine(train), feats, mine(score))
        p = O.FilterParams(bins=bins, min_fill=0.2, max_fill_difference=0.3, max_fill_ratio_diff=2.0,
                           max_js_divergence=0.5, max_correlation=0.6, min_scoring_rows=10)
        drop = O.generate_filtered_raw(train, score, ["y"], p)[4]
        assert sorted(x.name for x in out.featuresToDrop) == sorted(drop)
        q.put((rank, "ok"))
    except Exception:  # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_two_ranks_add_up(tmp_path):
    import socket
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], "\n".join(str(r[1]) for r in res)


def test_protected_features_and_js_protection(monkeypatch):
    """RawFeatureFilterTest.scala:310-356: protected features never enter the checks; JS-protected ones skip only the
    divergence rule.  Against the oracle with the same settings."""
    from transmogrifai_b200.filters import RawFeatureFilter
    n = 400
    train, score = make_data(n, 3), make_data(n, 4, shift=True)
    feats = features_of(train)
    kw = dict(bins=10, minFill=0.0, maxFillDifference=1.0, maxFillRatioDiff=float("inf"), maxJSDivergence=0.0,
              maxCorrelation=1.0, minScoringRows=0)
    tr, sc = chunked(train, [n])[0], chunked(score, [n])[0]
    base = emulate(RawFeatureFilter(**kw), monkeypatch).generateFilteredRaw(tr, feats, sc)
    dropped = {f.name for f in base.featuresToDrop}
    assert {"r0", "t0", "i0"} <= dropped                      # any drift at all exceeds maxJSDivergence = 0
    prot = emulate(RawFeatureFilter(protectedFeatures={"r0"}, jsDivergenceProtectedFeatures={"t0"}, **kw), monkeypatch)
    out = prot.generateFilteredRaw(tr, feats, sc)
    got = {f.name for f in out.featuresToDrop}
    assert got == dropped - {"r0", "t0"}
    names = [m.name for m in out.rawFeatureFilterResults.rawFeatureFilterMetrics]
    assert "r0" not in names and "t0" in names               # protected: not even measured; JS-protected: measured
    t0 = [r for r in out.rawFeatureFilterResults.exclusionReasons if r.name == "t0"][0]
    assert not t0.jsDivergenceMismatch and not t0.excluded
    p = O.FilterParams(bins=10, min_fill=0.0, max_fill_difference=1.0, max_fill_ratio_diff=float("inf"),
                       max_js_divergence=0.0, max_correlation=1.0, js_divergence_protected=["t0"], protected=["r0"],
                       min_scoring_rows=0)
    assert sorted(got) == sorted(O.generate_filtered_raw(train, score, ["y"], p)[4])
    # the distributions of protected features are still reported (RawFeatureFilter.scala:547-549)
    assert "r0" in [d.name for d in out.trainingFeatureDistributions]
