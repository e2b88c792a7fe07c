"""GPU suite, multi-GPU part: ShardedFlatIndex over NCCL with 2 ranks (skipped on a 1-GPU box;
the gloo tests in test_sharded_gloo.py cover the host logic everywhere)."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out, fused):
    os.environ["B200R_FUSED_EXCHANGE"] = fused
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from generativeaiexamples_b200.sharded import ShardedFlatIndex

        rng = np.random.default_rng(11)
        n, d = 50001, 256
        x = rng.standard_normal((n, d), dtype=np.float32)
        q = rng.standard_normal((40, d), dtype=np.float32)
        res = {}
        for metric in (0, 1):
            idx = ShardedFlatIndex(d, metric, device=rank)
            idx.add(x)
            assert idx.ntotal == n
            for nq, k in ((1, 10), (3, 4), (40, 10)):   # scan path and GEMM path under the merge
                D, I = idx.search(q[:nq], k)
                res[(metric, nq, k)] = (D, I)
            Dd, Id = idx.search(torch.from_numpy(q[:2]).cuda(), 5)  # device in -> device out
            res[(metric, "dev")] = (Dd.cpu().numpy(), Id.cpu().numpy())
            # a run of back-to-back small searches: exercises the two-epoch buffering of the fused
            # exchange (device outputs are only read after the whole run)
            outs = [idx.search(torch.from_numpy(q[i:i + 1]).cuda(), 10) for i in range(12)]
            res[(metric, "run")] = (np.concatenate([o[0].cpu().numpy() for o in outs]),
                                    np.concatenate([o[1].cpu().numpy() for o in outs]))
            res[(metric, "fused")] = idx._xchg is not None
            # the corpus keeps growing and shrinking like one faiss index: a later batch (appended to the last
            # rank), deletions across both blocks, then searches and a reconstruct on the new positions
            extra = np.random.default_rng(12).standard_normal((777, d), dtype=np.float32)
            idx.add(extra)
            gone = np.array([0, 5, 24999, 25001, 50000, 50001 + 3, 50001 + 776, 10 ** 9], dtype=np.int64)
            assert idx.remove_ids(gone) == 7 and idx.ntotal == n + 777 - 7
            res[(metric, "mut")] = idx.search(q[:6], 10) + (idx.reconstruct(n + 100),)
            # one faiss file for the whole sharded corpus; loads back re-levelled over the ranks and, on rank 0,
            # into a plain single-GPU index -- same answers either way
            path = os.path.join(os.environ["B200R_TEST_TMP"], f"sharded_{metric}_{fused}.faiss")
            idx.write(path)
            again = ShardedFlatIndex.read(path, device=rank)
            assert again.ntotal == idx.ntotal and again.metric_type == metric
            Da, Ia = again.search(q[:6], 10)
            np.testing.assert_array_equal(Ia, res[(metric, "mut")][1])
            np.testing.assert_array_equal(Da, res[(metric, "mut")][0])
            if rank == 0:
                from generativeaiexamples_b200 import B200IndexFlat

                one = B200IndexFlat.read(path, device=rank)
                D1, I1 = one.search(q[:6], 10)
                np.testing.assert_array_equal(I1, Ia)
                np.testing.assert_array_equal(D1, Da)
                one.close()
            # re-level the blocks (the later batch sits on the last rank): rows move over NCCL, positions stay
            assert idx._counts[0] != idx._counts[1]
            assert idx.rebalance() is True and abs(idx._counts[0] - idx._counts[1]) <= 1
            Db, Ib = idx.search(q[:6], 10)
            np.testing.assert_array_equal(Ib, res[(metric, "mut")][1])
            np.testing.assert_array_equal(Db, res[(metric, "mut")][0])
            np.testing.assert_array_equal(idx.reconstruct(n + 100), res[(metric, "mut")][2])
            assert idx.rebalance() is False
        if rank == 0:
            out.put(res)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("fused", ["0", "1"])
def test_sharded_nccl_two_ranks(built_lib, fused, tmp_path):
    import torch
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    os.environ["B200R_TEST_TMP"] = str(tmp_path)
    from oracle import assert_topk_parity, truth_topk

    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out, fused)) for r in range(2)]
    for p in procs:
        p.start()
    res = out.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    rng = np.random.default_rng(11)
    x = rng.standard_normal((50001, 256), dtype=np.float32)
    q = rng.standard_normal((40, 256), dtype=np.float32)
    for (metric, nq, k), (D, I) in ((kk, v) for kk, v in res.items() if len(kk) == 3):
        Dt, It = truth_topk("l2" if metric else "ip", q[:nq], x, k + 8)
        assert_topk_parity(D, I, Dt, It, what=f"sharded nccl metric={metric} nq={nq} k={k}")
    for metric in (0, 1):
        D, I = res[(metric, "dev")]
        Dt, It = truth_topk("l2" if metric else "ip", q[:2], x, 13)
        assert_topk_parity(D, I, Dt, It, what="sharded nccl device tensors")
        D, I = res[(metric, "run")]
        Dt, It = truth_topk("l2" if metric else "ip", q[:12], x, 18)
        assert_topk_parity(D, I, Dt, It, what="sharded back-to-back run")
        D, I, row = res[(metric, "mut")]
        extra = np.random.default_rng(12).standard_normal((777, 256), dtype=np.float32)
        gone = [0, 5, 24999, 25001, 50000, 50001 + 3, 50001 + 776]
        xm = np.delete(np.concatenate([x, extra]), gone, axis=0)
        Dt, It = truth_topk("l2" if metric else "ip", q[:6], xm, 18)
        assert_topk_parity(D, I, Dt, It, what="sharded index after add + remove_ids")
        np.testing.assert_array_equal(row, xm[50001 + 100])
    assert res[(0, "fused")] == (fused == "1")
