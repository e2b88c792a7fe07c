"""CPU suite, part 4: the N>1 host logic over gloo, world_size 2 and 3 (row blocks, id
offsets, packed allgather layout, merge ordering).  The two device steps are replaced by
the CPU oracle, which is allowed in tests only."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from generativeaiexamples_b200.sharded import ShardedFlatIndex, _packed_layout, shard_bounds


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class _LocalRef:
    """IndexFlatRef + the id_base the device index applies in its epilogue."""

    def __init__(self, d, metric):
        from oracle.flat_ref import IndexFlatRef

        self.ix, self.base = IndexFlatRef(d, metric), 0

    ntotal = property(lambda self: self.ix.ntotal)

    def add(self, x):
        self.ix.add(np.asarray(x, dtype=np.float32))

    def remove_ids(self, ids):
        return self.ix.remove_ids(np.asarray(ids, dtype=np.int64))

    def reconstruct(self, key):
        return self.ix.reconstruct(int(key))

    def reset(self):
        self.ix.reset()

    def set_id_base(self, b):
        self.base = b

    def search(self, q, k):
        D, I = self.ix.search(q, k)
        return D, np.where(I >= 0, I + self.base, -1)


def _merge_ref(metric, cand_d, cand_i, k):
    """[G][nq][k] -> [nq][k]; same rule as the K4 kernel: skip id<0, ties to the lower rank."""
    G, nq, kin = cand_d.shape
    D = np.full((nq, k), np.finfo(np.float32).max if metric == 1 else -np.finfo(np.float32).max, dtype=np.float32)
    I = np.full((nq, k), -1, dtype=np.int64)
    for qi in range(nq):
        d = cand_d[:, qi, :].reshape(-1)
        i = cand_i[:, qi, :].reshape(-1)
        pos = np.arange(d.size)
        ok = i >= 0
        d, i, pos = d[ok], i[ok], pos[ok]
        order = np.lexsort((pos, d if metric == 1 else -d))[:k]
        D[qi, : len(order)] = d[order]
        I[qi, : len(order)] = i[order]
    return D, I


def _worker(rank, world, port, metric, n, nq, k, q_out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(5)
        d = 32
        x = rng.standard_normal((n, d), dtype=np.float32)
        q = rng.standard_normal((nq, d), dtype=np.float32)
        idx = ShardedFlatIndex(d, metric, local_index=_LocalRef(d, metric), merge_fn=_merge_ref)
        idx.add(x)
        lo, hi = shard_bounds(n, world, rank)
        assert idx.local.ntotal == hi - lo and idx.id_base == lo and idx.ntotal == n
        D, I = idx.search(q, k)
        if rank == 0:
            q_out.put((D, I))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("metric", [0, 1])
def test_sharded_search_matches_single_index(world, metric):
    from oracle import FlatRef

    n, nq, k = 1001, 5, 10
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, metric, n, nq, k, out)) for r in range(world)]
    for p in procs:
        p.start()
    D, I = out.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rng = np.random.default_rng(5)
    x = rng.standard_normal((n, 32), dtype=np.float32)
    q = rng.standard_normal((nq, 32), dtype=np.float32)
    Dr, Ir = FlatRef.knn(metric, q, x, k)
    np.testing.assert_array_equal(I, Ir)
    np.testing.assert_allclose(D, Dr, rtol=1e-6)


def test_shard_bounds_and_layout():
    assert [shard_bounds(10, 4, r) for r in range(4)] == [(0, 3), (3, 6), (6, 9), (9, 10)]
    assert [shard_bounds(2, 4, r) for r in range(4)] == [(0, 1), (1, 2), (2, 2), (2, 2)]
    assert shard_bounds(0, 2, 1) == (0, 0)
    d_bytes, block = _packed_layout(1, 10)
    assert d_bytes == 40 and block == 120
    d_bytes, block = _packed_layout(3, 3)  # 36 -> padded to 40 so the ids stay 8-byte aligned
    assert d_bytes == 40 and block == 40 + 72


def test_single_process_world1():
    idx = ShardedFlatIndex(8, 1, local_index=_LocalRef(8, 1), merge_fn=_merge_ref)
    x = np.eye(8, dtype=np.float32)
    idx.add(x)
    D, I = idx.search(x[:2], 3)
    assert I[:, 0].tolist() == [0, 1] and idx.ntotal == 8


def _mutating_worker(rank, world, port, q_out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle.flat_ref import IndexFlatRef

        rng = np.random.default_rng(11)
        d = 16
        idx = ShardedFlatIndex(d, 1, local_index=_LocalRef(d, 1), merge_fn=_merge_ref)
        ref = IndexFlatRef(d, 1)
        log = []
        for step in range(6):
            x = rng.standard_normal((int(rng.integers(5, 60)), d)).astype(np.float32)
            idx.add(x)          # the first batch is split into blocks, later ones go to the last rank
            ref.add(x)
            if step % 2:
                ids = rng.integers(-2, ref.ntotal + 2, 9).astype(np.int64)
                assert idx.remove_ids(ids) == ref.remove_ids(ids)
            assert idx.ntotal == ref.ntotal
            key = int(rng.integers(0, ref.ntotal))
            np.testing.assert_array_equal(idx.reconstruct(key), ref.reconstruct(key))
            q = rng.standard_normal((3, d)).astype(np.float32)
            D, I = idx.search(q, 7)
            Dr, Ir = ref.search(q, 7)
            np.testing.assert_array_equal(I, Ir)
            np.testing.assert_allclose(D, Dr, rtol=1e-6)
            log.append(int(idx.ntotal))
        idx.reset()
        assert idx.ntotal == 0 and (idx.search(q, 2)[1] == -1).all()
        if rank == 0:
            q_out.put(log)
    finally:
        dist.destroy_process_group()


def test_sharded_index_is_mutable_like_faiss():
    """add after the first batch, remove_ids, reconstruct and reset on the row-sharded index (world 3, gloo): ids
    stay faiss positions -- later rows shift down on removal -- because the blocks stay contiguous in rank order."""
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_mutating_worker, args=(r, 3, port, out)) for r in range(3)]
    for p in procs:
        p.start()
    log = out.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert len(log) == 6 and log[-1] > 0


def _persist_worker(rank, world, port, path, q_out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(21)
        d, n = 24, 517
        x = rng.standard_normal((n, d), dtype=np.float32)
        q = rng.standard_normal((4, d), dtype=np.float32)
        idx = ShardedFlatIndex(d, 0, local_index=_LocalRef(d, 0), merge_fn=_merge_ref)
        idx.add(x[:400])
        idx.add(x[400:])          # a later batch: goes to the last rank, positions 400..516
        idx.remove_ids([3, 250, 516])
        idx.write(path, chunk_rows=100)
        D0, I0 = idx.search(q, 7)
        again = ShardedFlatIndex.read(path, local_index_factory=lambda d_, m_: _LocalRef(d_, m_), merge_fn=_merge_ref)
        assert again.ntotal == n - 3 and again.d == d and again.metric_type == 0
        lo, hi = shard_bounds(n - 3, world, rank)
        assert again.local.ntotal == hi - lo and again.id_base == lo   # re-levelled on load
        D1, I1 = again.search(q, 7)
        np.testing.assert_array_equal(I0, I1)
        np.testing.assert_array_equal(D0, D1)
        with pytest.raises(RuntimeError):
            again.read_rows_from(path)  # not empty
        if rank == 0:
            q_out.put("ok")
    finally:
        dist.destroy_process_group()


def test_sharded_write_read_faiss_file(tmp_path):
    """One faiss flat-index file for the whole sharded corpus: the bytes a single index would write, loadable
    by any number of ranks (and by a single-GPU index)."""
    import struct

    world = 3
    path = str(tmp_path / "sharded.faiss")
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_persist_worker, args=(r, world, port, path, out)) for r in range(world)]
    for p in procs:
        p.start()
    assert out.get(timeout=120) == "ok"
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rng = np.random.default_rng(21)
    x = rng.standard_normal((517, 24), dtype=np.float32)
    keep = np.delete(x, [3, 250, 516], axis=0)
    raw = open(path, "rb").read()
    head = struct.pack("<4siqqqBiQ", b"IxFI", 24, 514, 1 << 20, 1 << 20, 1, 0, 514 * 24)
    assert raw[:45] == head
    assert raw[45:] == keep.tobytes()


def _rebalance_worker(rank, world, port, q_out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(31)
        d = 16
        x = rng.standard_normal((900, d), dtype=np.float32)
        q = rng.standard_normal((5, d), dtype=np.float32)
        idx = ShardedFlatIndex(d, 1, local_index=_LocalRef(d, 1), merge_fn=_merge_ref)
        idx.add(x[:90])
        assert idx.rebalance() is False          # level already
        for lo in range(90, 900, 90):            # nine later batches: all land on the last rank
            idx.add(x[lo:lo + 90])
        first_lo, first_hi = shard_bounds(90, world, world - 1)
        assert idx._counts[-1] == (first_hi - first_lo) + 810 and idx.ntotal == 900
        idx.remove_ids([0, 1, 2, 899])
        D0, I0 = idx.search(q, 9)
        rows0 = [idx.reconstruct(k_) for k_ in (0, 29, 30, 500, 895)]
        assert idx.rebalance() is True
        lo, hi = shard_bounds(896, world, rank)
        assert idx.local.ntotal == hi - lo and idx.id_base == lo and idx.ntotal == 896
        D1, I1 = idx.search(q, 9)
        np.testing.assert_array_equal(I0, I1)   # positions did not change
        np.testing.assert_array_equal(D0, D1)
        for k_, r0 in zip((0, 29, 30, 500, 895), rows0):
            np.testing.assert_array_equal(idx.reconstruct(k_), r0)
        assert idx.rebalance() is False
        if rank == 0:
            q_out.put("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_rebalance_moves_rows_not_positions(world):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rebalance_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    assert out.get(timeout=120) == "ok"
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
