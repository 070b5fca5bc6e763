"""Two ranks on one GPU (gloo for the small host-side exchanges, both processes computing on cuda:0): the row-sharded
fit of every vectorizer + SanityChecker gives the statistics, vocabularies, decisions and dropped set of the whole
dataset, and each rank's transform gives its own rows."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _make_data(n):
    from transmogrifai_b200 import features as F
    rng = np.random.default_rng(77)
    y = (rng.random(n) < 0.3).astype(np.float64)
    opt = lambda v, p: None if rng.random() < p else v   # noqa: E731
    b0 = [opt(bool(rng.random() < 0.4 + 0.3 * yy), 0.1) for yy in y]
    p0 = [opt(["x", "y", "z", "X ", "w"][int(rng.integers(5))], 0.08) for _ in y]
    r0 = [opt(float(rng.normal(2 + yy, 1)), 0.1) for yy in y]
    nn = [float(v) for v in rng.normal(5, 2, n)]
    words = ["".join(rng.choice(list("abcdefgh"), size=rng.integers(3, 7))) for _ in range(150)]
    t0 = [opt(" ".join(words[i] for i in rng.integers(0, 150, rng.poisson(8))), 0.06) for _ in y]
    # low-cardinality text: pivoted; its values arrive in a different order on the two shards
    t1 = [opt(["red", "Green", "blue", "blue!", "teal"][int(rng.integers(5 if i > n // 2 else 3))], 0.05)
          for i in range(n)]
    # "the the ..." in every row: a hashed bucket met twice or more in all rows (column minimum 2)
    t2 = ["zq zq " * int(rng.integers(1, 3)) + words[int(rng.integers(150))] for _ in y]
    # an Integral column whose mode differs between the two shards (5 on the first rows, 2 on the rest; 2 overall)
    i0 = [opt(int(5 if (i < n // 3 and rng.random() < 0.6) else rng.integers(0, 4) if rng.random() < 0.5 else 2), 0.1)
          for i in range(n)]
    return {"i0": (F.INTEGRAL, i0), "b0": (F.BINARY, b0), "p0": (F.PICKLIST, p0), "r0": (F.REAL, r0), "nn": (F.REALNN, nn), "t0": (F.TEXT, t0),
            "t1": (F.TEXT, t1), "t2": (F.TEXT, t2), "y": (F.REALNN, y.tolist())}, y


def _init(rank, size, port, backend):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    if backend == "nccl":                       # a GPU per rank, the exchanges over NCCL
        torch.cuda.set_device(rank)
        dist.init_process_group("nccl", rank=rank, world_size=size, device_id=torch.device("cuda", rank))
    else:
        dist.init_process_group("gloo", rank=rank, world_size=size)


def _worker(rank, size, port, q, backend="gloo"):
    _init(rank, size, port, backend)
    try:
        from oracle import sanity as S
        from oracle import vectorizers as V
        from transmogrifai_b200 import dist as D
        from transmogrifai_b200 import features as F
        from transmogrifai_b200.columnar import ColumnarBatch
        from transmogrifai_b200.dsl import fit_and_transform, transmogrify
        n = 5000
        data, y = _make_data(n)
        lo, hi = D.shard_rows(n, rank, size)
        assert hi > lo
        shard = {k: (t, v[lo:hi]) for k, (t, v) in data.items()}
        batch = ColumnarBatch.from_columns(shard, response="y").to("cuda")
        FB = F.FeatureBuilder
        feats = [FB.Text("t0").asPredictor(), FB.Real("r0").asPredictor(), FB.Binary("b0").asPredictor(),
                 FB.RealNN("nn").asPredictor(), FB.PickList("p0").asPredictor(), FB.Text("t1").asPredictor(),
                 FB.Text("t2").asPredictor(), FB.Integral("i0").asPredictor()]
        label = FB.RealNN("y").asResponse()
        checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True, sampleLowerLimit=0,
                                    sampleUpperLimit=10 ** 9, featureFeatureCorrLevel="Off")
        out_batch, fitted = fit_and_transform(batch, checked)
        est = checked.origin_stage
        # oracle on the whole dataset
        col = lambda k: data[k][1]   # noqa: E731
        xb = V.binary_vectorizer_transform([col("b0")])
        modes = V.integral_vectorizer_fit([col("i0")], fill_with_mode=True)
        xi = V.integral_vectorizer_transform([col("i0")], modes, True)
        tops = V.pivot_fit([col("p0")], 20, 10, True)
        xp = V.pivot_transform([col("p0")], tops, True, True)
        fills = V.real_vectorizer_fit([col("r0")])
        xr = V.real_vectorizer_transform([col("r0")], fills, True)
        xn = V.realnn_vectorizer_transform([col("nn")])
        tcols, tnames = [col("t0"), col("t1"), col("t2")], ["t0", "t1", "t2"]
        methods, ttops = V.smart_text_fit(tcols, max_cardinality=30, top_k=20, min_support=10)
        assert methods == ["Hash", "Pivot", "Hash"]
        xt = V.smart_text_transform(tcols, tnames, methods, ttops, num_features=512)
        xo, metas = V.combine(
            [xb, xi, xp, xr, xn, xt],
            [V.numeric_metadata(["b0"], "Binary", True), V.numeric_metadata(["i0"], "Integral", True),
             V.pivot_metadata(["p0"], "PickList", tops, True),
             V.numeric_metadata(["r0"], "Real", True), V.numeric_metadata(["nn"], "RealNN", False),
             V.smart_text_metadata(tnames, "Text", methods, ttops, 512, True)])
        p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                           feature_feature_corr_level="off")
        ref = S.sanity_fit(y, xo, metas, p, label_name="y")
        exp_names = [m.col_name() for m in metas]
        got_names = est.summary.names[:-1]
        assert got_names == exp_names, [(i, a, b) for i, (a, b) in enumerate(zip(got_names, exp_names)) if a != b][:6] + \
            [len(got_names), len(exp_names)]
        assert est.summary.dropped == ref.dropped
        assert fitted[-1].indicesToKeep == ref.indices_to_keep
        bad = np.nonzero(~np.isclose(est.summary.mean, ref.col_stats.mean, rtol=1e-12, atol=1e-300))[0]
        assert len(bad) == 0, [(int(i), (exp_names + ["y"])[i], est.summary.mean[i], ref.col_stats.mean[i]) for i in bad[:6]]
        np.testing.assert_allclose(est.summary.variance, ref.col_stats.variance, rtol=1e-9, atol=1e-18)
        assert np.array_equal(est.summary.min, ref.col_stats.min) and np.array_equal(est.summary.max, ref.col_stats.max)
        assert ref.col_stats.min.max() >= 2.0          # the "zq zq" bucket: met at least twice in every row
        np.testing.assert_allclose(est.summary.correlationsWithLabel, ref.corrs_with_label, rtol=1e-9, atol=1e-12,
                                   equal_nan=True)
        got_g = {g.group: g for g in est.summary.categoricalStats}
        assert got_g.keys() == {g.group for g in ref.cat_stats}
        for g in ref.cat_stats:
            assert np.array_equal(got_g[g.group].contingencyMatrix, g.contingency_matrix)
            np.testing.assert_allclose(got_g[g.group].cramersV, g.cramers_v, rtol=1e-9, equal_nan=True)
        got = out_batch.vectors[checked.name].to_scipy()
        exp = xo[lo:hi][:, ref.indices_to_keep]
        assert got.shape == exp.shape
        np.testing.assert_allclose(np.asarray(got.todense()), np.asarray(exp.todense()) if hasattr(exp, "todense") else exp,
                                   rtol=1e-12)
        q.put((rank, "ok"))
    except Exception:  # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def _run(target, *extra):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=target, args=(r, 2, port, q) + extra) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], "\n".join(str(r[1]) for r in res)


def test_two_ranks_all_types_match_oracle():
    _run(_worker)


needs_two_gpus = pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (NCCL)")


@needs_two_gpus
def test_two_ranks_over_nccl_all_types_match_oracle():
    """The same rank merges with a GPU per rank and NCCL as the transport (what bench.py --gpus N runs)."""
    _run(_worker, "nccl")


def _worker_streamed(rank, size, port, q, backend="gloo", sample=False, corr="Off"):
    """Every rank streams its row shard in three chunks (own dictionaries per chunk): passes 0 / 1 fold the chunks, the
    accumulators are merged over the ranks once, pass 2 transforms chunk by chunk."""
    _init(rank, size, port, backend)
    try:
        import scipy.sparse as sp
        from oracle import sanity as S
        from oracle import vectorizers as V
        from transmogrifai_b200 import dist as D
        from transmogrifai_b200 import features as F
        from transmogrifai_b200.columnar import ColumnarBatch, VectorColumn
        from transmogrifai_b200.dsl import transmogrify
        from transmogrifai_b200.streaming import fit_and_transform_stream
        n = 5000
        data, y = _make_data(n)
        data.pop("i0")
        lo, hi = D.shard_rows(n, rank, size)
        cuts = [lo, lo + (hi - lo) // 5, lo + (hi - lo) // 2, hi]
        chunks = [ColumnarBatch.from_columns({k: (t, v[a:b]) for k, (t, v) in data.items()}, response="y")
                  for a, b in zip(cuts[:-1], cuts[1:])]
        FB = F.FeatureBuilder
        feats = [FB.Text("t0").asPredictor(), FB.Real("r0").asPredictor(), FB.Binary("b0").asPredictor(),
                 FB.RealNN("nn").asPredictor(), FB.PickList("p0").asPredictor(), FB.Text("t1").asPredictor(),
                 FB.Text("t2").asPredictor()]
        label = FB.RealNN("y").asResponse()
        checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True, sampleLowerLimit=0,
                                    sampleUpperLimit=10 ** 9, featureFeatureCorrLevel=corr,
                                    checkSample=0.5 if sample else 1.0)
        out = []
        # rank 0 keeps the packed text cells of its chunks for pass 2, rank 1 transforms from the text: same rows
        res = fit_and_transform_stream(chunks, checked, lambda k, vec, parts: out.append(
            VectorColumn(vec.name, vec.n, vec.metadata, parts).to_scipy()), cell_cache_bytes=(16 << 20) if rank == 0 else 0)
        assert res.cells_cached == (len(chunks) if rank == 0 else 0)
        est = checked.origin_stage
        col = lambda k: data[k][1]   # noqa: E731
        xb = V.binary_vectorizer_transform([col("b0")])
        tops = V.pivot_fit([col("p0")], 20, 10, True)
        xp = V.pivot_transform([col("p0")], tops, True, True)
        fills = V.real_vectorizer_fit([col("r0")])
        xr = V.real_vectorizer_transform([col("r0")], fills, True)
        xn = V.realnn_vectorizer_transform([col("nn")])
        tcols, tnames = [col("t0"), col("t1"), col("t2")], ["t0", "t1", "t2"]
        methods, ttops = V.smart_text_fit(tcols, max_cardinality=30, top_k=20, min_support=10)
        xt = V.smart_text_transform(tcols, tnames, methods, ttops, num_features=512)
        xo, metas = V.combine(
            [xb, xp, xr, xn, xt],
            [V.numeric_metadata(["b0"], "Binary", True), V.pivot_metadata(["p0"], "PickList", tops, True),
             V.numeric_metadata(["r0"], "Real", True), V.numeric_metadata(["nn"], "RealNN", False),
             V.smart_text_metadata(tnames, "Text", methods, ttops, 512, True)])
        p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                           feature_feature_corr_level=corr.lower())
        rows = np.arange(n)
        if sample:          # SanityChecker looked at a row sample: every rank's sampled rows, gathered
            mine = np.concatenate([r.cpu().numpy() + a for r, a in zip(res.sample_rows, cuts[:-1])])
            rows = np.sort(np.concatenate(D.all_gather_object(mine)))
            assert 0.4 * n < len(rows) < 0.6 * n
        ref = S.sanity_fit(y[rows], sp.csr_matrix(xo)[rows], metas, p, label_name="y")
        assert est.summary.count == len(rows)
        assert est.summary.names[:-1] == [m.col_name() for m in metas]
        assert est.summary.dropped == ref.dropped
        assert res.fitted[-1].indicesToKeep == ref.indices_to_keep
        np.testing.assert_allclose(est.summary.mean, ref.col_stats.mean, rtol=1e-10 if sample else 1e-12, atol=1e-300)
        np.testing.assert_allclose(est.summary.variance, ref.col_stats.variance, rtol=1e-9, atol=1e-18)
        assert np.array_equal(est.summary.min, ref.col_stats.min) and np.array_equal(est.summary.max, ref.col_stats.max)
        np.testing.assert_allclose(est.summary.correlationsWithLabel, ref.corrs_with_label, rtol=1e-9, atol=1e-12,
                                   equal_nan=True)
        if corr == "Computed":      # every rank's chunks add to its Gramian, the ranks' Gramians add
            np.testing.assert_allclose(est.summary.featureFeatureCorr, ref.corr_matrix, rtol=1e-9, atol=1e-12,
                                       equal_nan=True)
        got_g = {g.group: g for g in est.summary.categoricalStats}
        for g in ref.cat_stats:
            assert np.array_equal(got_g[g.group].contingencyMatrix, g.contingency_matrix)
            np.testing.assert_allclose(got_g[g.group].cramersV, g.cramers_v, rtol=1e-9, equal_nan=True)
        got = sp.vstack(out).tocsr()
        exp = sp.csr_matrix(xo)[lo:hi][:, ref.indices_to_keep]
        assert got.shape == exp.shape and abs(got - exp).max() < 1e-12
        q.put((rank, "ok"))
    except Exception:  # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_two_ranks_streamed_chunks_match_oracle():
    _run(_worker_streamed)


@needs_two_gpus
def test_two_ranks_over_nccl_streamed_chunks_match_oracle():
    _run(_worker_streamed, "nccl")


def test_two_ranks_streamed_chunks_sampled_sanity_checker_matches_oracle():
    """fraction() < 1 over streamed chunks on two ranks: the ranks' sample statistics merge in rank order."""
    _run(_worker_streamed, "gloo", True)


@pytest.mark.parametrize("sample", [False, True])
def test_two_ranks_streamed_chunks_full_correlation_matrix_matches_oracle(sample):
    """CorrelationLevel.Computed over streamed chunks on two ranks (all rows, and a row sample)."""
    _run(_worker_streamed, "gloo", sample, "Computed")


def _worker_sampled(rank, size, port, q):
    _init(rank, size, port, "gloo")
    try:
        from oracle import sanity as S
        from oracle import vectorizers as V
        from transmogrifai_b200 import dist as D
        from transmogrifai_b200 import features as F
        from transmogrifai_b200.columnar import ColumnarBatch
        from transmogrifai_b200.dsl import fit_and_transform, transmogrify
        n, seed = 8192, 11
        data, y = _make_data(n)
        keep = ("b0", "p0", "r0", "nn", "y")
        lo, hi = D.shard_rows(n, rank, size)
        shard = {k: (data[k][0], data[k][1][lo:hi]) for k in keep}
        batch = ColumnarBatch.from_columns(shard, response="y").to("cuda")
        FB = F.FeatureBuilder
        feats = [FB.Real("r0").asPredictor(), FB.Binary("b0").asPredictor(), FB.RealNN("nn").asPredictor(),
                 FB.PickList("p0").asPredictor()]
        label = FB.RealNN("y").asResponse()
        checked = label.sanityCheck(transmogrify(feats), removeBadFeatures=True, checkSample=0.5, sampleLowerLimit=0,
                                    sampleUpperLimit=10 ** 9, sampleSeed=seed, featureFeatureCorrLevel="Computed")
        out_batch, fitted = fit_and_transform(batch, checked)
        est = checked.origin_stage
        # the rows every rank sampled from its shard (its own stream: seed + rank), in global numbering
        idx = []
        for r in range(size):
            a, b = D.shard_rows(n, r, size)
            g = torch.Generator(device="cuda").manual_seed(seed + r)
            idx.append(a + (torch.rand(b - a, device="cuda", generator=g) < 0.5).nonzero().squeeze(1).cpu().numpy())
        idx = np.concatenate(idx)
        col = lambda k: data[k][1]   # noqa: E731
        xb = V.binary_vectorizer_transform([col("b0")])
        tops = V.pivot_fit([col("p0")], 20, 10, True)
        xp = V.pivot_transform([col("p0")], tops, True, True)
        fills = V.real_vectorizer_fit([col("r0")])
        xr = V.real_vectorizer_transform([col("r0")], fills, True)
        xn = V.realnn_vectorizer_transform([col("nn")])
        xo, metas = V.combine(
            [xb, xp, xr, xn],
            [V.numeric_metadata(["b0"], "Binary", True), V.pivot_metadata(["p0"], "PickList", tops, True),
             V.numeric_metadata(["r0"], "Real", True), V.numeric_metadata(["nn"], "RealNN", False)])
        p = S.SanityParams(remove_bad_features=True, sample_lower_limit=0, sample_upper_limit=10 ** 9,
                           feature_feature_corr_level="computed")
        xs = xo.tocsr()[idx] if hasattr(xo, "tocsr") else xo[idx]
        ref = S.sanity_fit(y[idx], xs, metas, p, label_name="y")
        assert est.summary.count == len(idx)
        assert est.summary.dropped == ref.dropped
        assert fitted[-1].indicesToKeep == ref.indices_to_keep
        np.testing.assert_allclose(est.summary.mean, ref.col_stats.mean, rtol=1e-12, atol=1e-300)
        np.testing.assert_allclose(est.summary.variance, ref.col_stats.variance, rtol=1e-9, atol=1e-18)
        np.testing.assert_allclose(est.summary.correlationsWithLabel, ref.corrs_with_label, rtol=1e-9, atol=1e-12,
                                   equal_nan=True)
        q.put((rank, "ok"))
    except Exception:  # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


def test_two_ranks_sampled_sanity_checker_matches_oracle():
    """fraction() < 1 over two row shards: every rank samples its shard, the statistics of the samples are merged."""
    _run(_worker_sampled)


def _worker_comm(rank, size, port, q):
    """tmog_comm_* through ctypes only (no torch.distributed): the NCCL id travels through a file, as a Spark driver
    would broadcast it."""
    try:
        import ctypes as C
        import time
        from transmogrifai_b200 import _lib as L
        path = f"/tmp/tmog_comm_id_{port}"
        ident = (C.c_uint8 * 128)()
        if rank == 0:
            L.call("tmog_comm_unique_id", ident)
            with open(path + ".tmp", "wb") as f:
                f.write(bytes(ident))
            os.replace(path + ".tmp", path)
        else:
            for _ in range(600):
                if os.path.exists(path):
                    break
                time.sleep(0.05)
            ident = (C.c_uint8 * 128).from_buffer_copy(open(path, "rb").read())
        comm = C.c_void_p()
        L.call("tmog_comm_init", ident, rank, size, rank, C.byref(comm))
        counts = np.arange(1000, dtype=np.int64) * (rank + 1)
        L.call("tmog_comm_allreduce_host", comm, counts.ctypes.data, len(counts), 0, 0)
        assert np.array_equal(counts, np.arange(1000, dtype=np.int64) * 3)
        sums = np.array([0.5 + rank, -2.0 * rank, 1e-3], dtype=np.float64)
        mx = sums.copy()
        L.call("tmog_comm_allreduce_host", comm, sums.ctypes.data, 3, 1, 0)
        L.call("tmog_comm_allreduce_host", comm, mx.ctypes.data, 3, 1, 1)
        assert np.allclose(sums, [2.0, -2.0, 2e-3]) and np.array_equal(mx, [1.5, 0.0, 1e-3])
        L.call("tmog_comm_destroy", comm)
        q.put((rank, "ok"))
    except Exception:  # pragma: no cover
        import traceback
        q.put((rank, traceback.format_exc()))


@needs_two_gpus
def test_comm_entry_points_allreduce_over_nccl():
    _run(_worker_comm)
