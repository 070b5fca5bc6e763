"""World-size-2 gloo tests (CPU) of the row-sharded multi-GPU plumbing: partial merge, shared shifts, label-class
union, count-table reduction.  The data path needs no collective; these are the only exchanges."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist  # noqa: E402
import torch.multiprocessing as mp  # noqa: E402


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, size, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=size)
    try:
        from transmogrifai_b200 import _lib as L
        from transmogrifai_b200 import dist as D
        assert D.world() == (rank, size)
        # 1. partial merge == the library's own host merge
        rng = np.random.default_rng(5)
        parts = [rng.standard_normal((4, 8)) for _ in range(size)]
        for p in parts:
            p[:, 0] = rng.integers(0, 100, 4)
            p[:, 7] = 100
        merged = D.allreduce_real_partials(parts[rank].reshape(-1).copy())
        ref = parts[0].reshape(-1).copy()
        for p in parts[1:]:
            L.call("tmog_real_partials_merge", ref.ctypes.data, p.reshape(-1).copy().ctypes.data, 3)
        np.testing.assert_allclose(merged, ref, rtol=1e-15)
        # 1b. the fused exchange: partials merged in rank order + count tables summed, one all-gather
        tabs = [np.arange(6, dtype=np.int64).reshape(2, 3) * (rank + 1), np.array([rank + 5], dtype=np.int64)]
        merged2, sums = D.gather_fit_results(parts[rank].reshape(-1).copy(), tabs)
        assert np.array_equal(merged2, ref)
        assert np.array_equal(sums[0], np.arange(6).reshape(2, 3) * 3) and sums[1].tolist() == [11]
        # 2. shifts come from rank 0
        sh = D.broadcast(np.full(5, float(rank + 1)))
        assert np.array_equal(sh, np.full(5, 1.0))
        # 3. label-class union: sorted, distinct by bit pattern, -0.0 before 0.0; None if any rank is continuous
        local = np.array([0.0, 2.0]) if rank == 0 else np.array([-0.0, 1.0, 2.0])
        u = D.union_classes(local)
        assert u.tolist() == [0.0, 0.0, 1.0, 2.0] and np.signbit(u[0]) and not np.signbit(u[1])
        assert D.union_classes(None if rank == 1 else local) is None
        # 3b. the same union from the device-side tensors in one exchange, with the per-class row counts when the
        #     lists agree
        def dev_lists(vals, cnts):
            c = torch.zeros(256, dtype=torch.float64)
            c[:len(vals)] = torch.tensor(vals, dtype=torch.float64)
            k = torch.zeros(256, dtype=torch.int64)
            k[:len(cnts)] = torch.tensor(cnts, dtype=torch.int64)
            return torch.tensor([len(vals)], dtype=torch.int32), c, k
        u, same, tot = D.gather_label_classes(*dev_lists([0.0, 1.0], [3 + rank, 7]))
        assert u.tolist() == [0.0, 1.0] and same and tot.tolist() == [7, 14]
        u, same, tot = D.gather_label_classes(*dev_lists([0.0, 2.0] if rank == 0 else [1.0, 2.0], [1, 1]))
        assert u.tolist() == [0.0, 1.0, 2.0] and not same and tot is None
        nc, c, k = dev_lists([0.0], [1])
        if rank == 1:
            nc[0] = -1                                # more classes than tracked on one rank: a continuous label
        assert D.gather_label_classes(nc, c, k)[0] is None
        # 4. count tables add
        c = D.allreduce_sum(np.arange(6, dtype=np.int64).reshape(2, 3) * (rank + 1))
        assert np.array_equal(c, np.arange(6).reshape(2, 3) * 3)
        # 5. shards tile the rows, starts aligned
        lo, hi = D.shard_rows(10_000, rank, size)
        assert lo % 2048 == 0 and (lo, hi) == ((0, 6144) if rank == 0 else (6144, 10_000))
        # 6. streamed fits over ranks: IntegralVectorizer's value -> count maps add key-wise (mode, ties to the smaller
        #    value), OpSetVectorizer's counts per cleaned value add up before the top-K (host side of stream_*_finish)
        from oracle import text as T
        from oracle import vectorizers as V
        from transmogrifai_b200 import features as F
        from transmogrifai_b200.stages.integral import _merge_tables
        from transmogrifai_b200.stages.sets import OpSetVectorizer
        rng = np.random.default_rng(9)
        col = [None if rng.random() < 0.2 else int(v) for v in rng.choice([4, 8, -1], size=400, p=[0.4, 0.4, 0.2])]
        lo2, hi2 = (0, 150) if rank == 0 else (150, 400)
        vals = np.array([v for v in col[lo2:hi2] if v is not None], dtype=np.int64)
        k, w = np.unique(vals, return_counts=True)
        assert _merge_tables(D.all_gather_object([(k, w.astype(np.int64))])) == V.mode_seq_null_int([col])
        sets = [[["a", "A "], ["b"], ["a", "c"], [], ["b", "b"]][int(rng.integers(5))] for _ in range(400)]
        est = OpSetVectorizer().setTopK(2).setMinSupport(5).setInput(F.FeatureBuilder.MultiPickList("s").asPredictor())
        local = {}
        for row in sets[lo2:hi2]:
            for v in row:
                cv = T.clean_string(v)
                local[cv] = local.get(cv, 0) + 1
        model = est.stream_pre_finish({"counts": [local]}, None)
        assert [list(t) for t in model.topValues] == [list(t) for t in V.pivot_fit([sets], 2, 5, True)]
        q.put((rank, "ok"))
    except Exception as e:  # pragma: no cover
        q.put((rank, repr(e)))
    finally:
        dist.destroy_process_group()


def test_gloo_world_size_2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res
