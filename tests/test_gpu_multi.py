"""GPU suite: the multi-GPU index that lives in ONE process (b200r_config.n_devices >= 2,
csrc/multi.cu) against the single-GPU index and the float64 truth.

The reference's chain server is a single process (RAG/src/chain_server/server.py:203-238) with an
incremental corpus (chains.py:83 add_documents per file, utils.py:546-560 delete per file); the
multi-GPU index must therefore answer exactly like faiss's one IndexFlat would: ids are positions,
deletions shift later ids down, ties go to the lowest id.  Listing a device twice puts two shards
on that GPU, so every code path here also runs on a 1-GPU box.
"""
import numpy as np
import pytest

from generativeaiexamples_b200 import _lib

pytestmark = pytest.mark.gpu


def _device_sets():
    sets = [[0, 0], [0, 0, 0]]
    try:
        import torch

        n = torch.cuda.device_count()
        if n >= 2:
            sets.append([0, 1])
        if n >= 4:
            sets.append([0, 1, 2, 3])
        if n >= 8:
            sets.append(list(range(8)))
    except Exception:
        pass
    return sets


DEVICE_SETS = _device_sets()
_ids = lambda v: "dev" + "".join(map(str, v))  # noqa: E731


def _pair(d, metric, devices, **kw):
    from generativeaiexamples_b200 import B200IndexFlat

    return B200IndexFlat(d, metric, devices=devices, **kw), B200IndexFlat(d, metric, **kw)


@pytest.mark.parametrize("devices", DEVICE_SETS, ids=_ids)
@pytest.mark.parametrize("metric", [0, 1])
def test_multi_equals_single(built_lib, devices, metric, monkeypatch):
    """Same rows added in the same order: every search answers bit-identically to the one-GPU index,
    on the fused small-batch route, the event route and the GEMM route, including ties that straddle
    shards (duplicated rows) and k beyond the shard sizes."""
    from oracle import assert_topk_parity, truth_topk

    monkeypatch.setenv("B200R_MULTI_MIN_PIECE", "50")
    rng = np.random.default_rng(11 + metric)
    for d in (64, 1024, 100):
        multi, single = _pair(d, metric, devices)
        x = rng.standard_normal((6000, d), dtype=np.float32)
        x[4000:4200] = x[100:300]  # exact duplicates far apart: equal scores on different shards
        cuts = [0, 1, 8, 700, 701, 2600, 2601, 2700, 6000]  # tiny, medium and large batches
        for a, b in zip(cuts[:-1], cuts[1:]):
            multi.add(x[a:b])
            single.add(x[a:b])
        st = multi.stats()
        assert multi.ntotal == single.ntotal == 6000 and st["ntotal"] == 6000
        assert st["n_devices"] == len(devices) and st["n_segments"] >= len(devices)
        q = rng.standard_normal((200, d), dtype=np.float32)
        q[::2] = x[rng.integers(0, 6000, 100)] + 0.02 * q[::2]
        q[1] = x[150]  # a duplicated row itself: two exact ties at the top
        for nq, k in ((1, 10), (2, 1), (3, 4), (4, 128), (1, 300), (17, 10), (200, 10), (40, 100), (5, 1024)):
            Dm, Im = multi.search(q[:nq], k)
            Ds, Is = single.search(q[:nq], k)
            np.testing.assert_array_equal(Im, Is, err_msg=f"ids d={d} nq={nq} k={k}")
            np.testing.assert_array_equal(Dm, Ds, err_msg=f"scores d={d} nq={nq} k={k}")
        Dt, It = truth_topk("l2" if metric else "ip", q[:8], x, 18)
        Dm, Im = multi.search(q[:8], 10)
        assert_topk_parity(Dm, Im, Dt, It, what=f"multi-GPU vs f64 truth d={d}")
        assert Im[1, 0] == 150 and Im[1, 1] == 4050  # tie across shards: the lower position first
        multi.close()
        single.close()


@pytest.mark.parametrize("devices", DEVICE_SETS[:3], ids=_ids)
def test_multi_mutations_follow_faiss_positions(built_lib, devices, monkeypatch):
    """remove_ids / reconstruct / reconstruct_batch / reset on the segment map: positions behave like
    faiss's (later rows shift down) although no row ever changes GPU."""
    from generativeaiexamples_b200 import B200IndexFlat

    monkeypatch.setenv("B200R_MULTI_MIN_PIECE", "32")
    rng = np.random.default_rng(5)
    d = 96
    ix = B200IndexFlat(d, 1, devices=devices)
    model = np.zeros((0, d), dtype=np.float32)
    for step in range(12):
        m = int(rng.choice([3, 40, 400, 2500]))
        x = rng.standard_normal((m, d), dtype=np.float32)
        ix.add(x)
        model = np.concatenate([model, x])
        if step % 2:
            ids = rng.integers(-3, model.shape[0] + 3, int(rng.choice([1, 30, 700])))
            valid = np.unique(ids[(ids >= 0) & (ids < model.shape[0])])
            assert ix.remove_ids(ids.astype(np.int64)) == len(valid)
            model = np.delete(model, valid, axis=0)
        assert ix.ntotal == model.shape[0]
        keys = rng.integers(0, model.shape[0], 50)
        np.testing.assert_array_equal(ix.reconstruct_batch(keys), model[keys])
        j = int(rng.integers(0, model.shape[0]))
        np.testing.assert_array_equal(ix.reconstruct(j), model[j])
        D, I = ix.search(model[keys[:5]], 1)
        # a stored row finds itself (or an identical earlier row) at distance 0
        assert (D[:, 0] == 0).all() and (I[:, 0] <= keys[:5]).all()
    with pytest.raises(RuntimeError):
        ix.reconstruct(model.shape[0])
    assert ix.remove_ids(np.arange(model.shape[0], dtype=np.int64)) == model.shape[0]
    assert ix.ntotal == 0 and (ix.search(model[:2], 3)[1] == -1).all()
    ix.add(model[:10])
    ix.reset()
    assert ix.ntotal == 0 and ix.stats()["n_segments"] == 0
    ix.add(model[:777])
    np.testing.assert_array_equal(ix.reconstruct_batch(np.arange(777)), model[:777])
    ix.close()


@pytest.mark.parametrize("devices", DEVICE_SETS[:3], ids=_ids)
def test_multi_device_tensors_save_load_merge(built_lib, devices, tmp_path, monkeypatch):
    """CUDA tensors in and out (asynchronously, back to back), faiss file round trip, merge_from in
    every combination of single- and multi-GPU indexes, bf16 storage and the bf16 shadow."""
    import torch

    from generativeaiexamples_b200 import B200IndexFlat
    from oracle import assert_topk_parity, truth_topk

    monkeypatch.setenv("B200R_MULTI_MIN_PIECE", "100")
    rng = np.random.default_rng(23)
    d = 200
    x = rng.standard_normal((5000, d), dtype=np.float32)
    q = x[rng.integers(0, 5000, 64)] + 0.05 * rng.standard_normal((64, d), dtype=np.float32)
    dev = torch.device("cuda", devices[-1])  # tensors on the LAST listed GPU: peers read them in place
    multi, single = _pair(d, 0, devices, shadow_bf16=True)
    xt = torch.from_numpy(x).to(dev)
    multi.add(xt[:3000])
    multi.add(xt[3000:])
    single.add(x)
    qt = torch.from_numpy(q).to(dev)
    outs = [multi.search(qt[:nq], k) for nq, k in ((1, 10), (4, 5), (64, 10), (1, 10))]  # no sync in between
    torch.cuda.synchronize(dev)
    for (Dm, Im), (nq, k) in zip(outs, ((1, 10), (4, 5), (64, 10), (1, 10))):
        Ds, Is = single.search(q[:nq], k)
        np.testing.assert_array_equal(Im.cpu().numpy(), Is)
        np.testing.assert_array_equal(Dm.cpu().numpy(), Ds)
    # host queries, device results
    path = str(tmp_path / "index.faiss")
    multi.write(path)
    again = B200IndexFlat.read(path, devices=devices)
    one = B200IndexFlat.read(path)
    assert again.ntotal == one.ntotal == 5000 and again.d == d and again.metric_type == 0
    np.testing.assert_array_equal(again.reconstruct_batch(np.arange(0, 5000, 97)), x[::97])
    np.testing.assert_array_equal(again.search(q[:3], 7)[1], one.search(q[:3], 7)[1])
    # merge_from: multi <- single, single <- multi, multi <- multi
    extra = rng.standard_normal((1500, d), dtype=np.float32)
    src = B200IndexFlat(d, 0)
    src.add(extra)
    again.merge_from(src)
    assert again.ntotal == 6500 and src.ntotal == 0
    np.testing.assert_array_equal(again.reconstruct(5000 + 123), extra[123])
    Dt, It = truth_topk("ip", q[:4], np.concatenate([x, extra]), 18)
    assert_topk_parity(*again.search(q[:4], 10), Dt, It, what="multi <- single merge")
    one.merge_from(again)
    assert one.ntotal == 5000 + 6500 and again.ntotal == 0
    np.testing.assert_array_equal(one.reconstruct(5000 + 5000 + 9), extra[9])
    again.add(extra[:300])
    other = B200IndexFlat(d, 0, devices=devices)
    other.add(x[:2100])
    again.merge_from(other)
    assert again.ntotal == 2400 and other.ntotal == 0
    np.testing.assert_array_equal(again.reconstruct_batch(np.array([0, 299, 300, 2399])),
                                  np.stack([extra[0], extra[299], x[0], x[2099]]))
    with pytest.raises(RuntimeError):
        again.merge_from(B200IndexFlat(d, 1, devices=devices))
    # bf16 storage spread over the shards: recall mode behaves as on one GPU
    mb, sb = _pair(d, 1, devices, storage="bf16")
    mb.add(x)
    sb.add(x)
    for nq in (1, 40):
        np.testing.assert_array_equal(mb.search(q[:nq], 10)[1], sb.search(q[:nq], 10)[1])
    for ix in (multi, single, again, one, src, other, mb, sb):
        ix.close()


@pytest.mark.parametrize("devices", DEVICE_SETS[:3], ids=_ids)
def test_multi_routes_timing_and_errors(built_lib, devices, monkeypatch):
    """The fused route (flags written by the scan kernels), the event route and the inline (no worker
    threads) issue mode give the same answers; timing splits kernel time from exchange time."""
    from generativeaiexamples_b200 import B200IndexFlat

    rng = np.random.default_rng(31)
    d = 128
    x = rng.standard_normal((30000, d), dtype=np.float32)
    q = rng.standard_normal((4, d), dtype=np.float32)
    ref = B200IndexFlat(d, 1)
    ref.add(x)
    want = ref.search(q, 10)
    for fused, workers in (("1", "1"), ("0", "1"), ("1", "0"), ("0", "0")):
        monkeypatch.setenv("B200R_MULTI_FUSED", fused)
        monkeypatch.setenv("B200R_MULTI_WORKERS", workers)
        _lib.reload_env()
        ix = B200IndexFlat(d, 1, devices=devices)
        ix.add(x)
        for _ in range(3):
            D, I = ix.search(q, 10)
            np.testing.assert_array_equal(I, want[1])
            np.testing.assert_array_equal(D, want[0])
        ix.set_timing(True)
        ix.search(q[:1], 10)
        st = ix.stats()
        assert st["last_kernel_us"] > 0 and st["last_exchange_us"] >= 0 and st["last_search_ms"] > 0
        assert st["bytes_scanned"] >= 30000 * d * 4 and st["n_kernel_launches"] > 0
        ix.set_timing(False)
        with pytest.raises(AssertionError):
            ix.search(q, 0)
        ix.close()
    with pytest.raises(RuntimeError):
        B200IndexFlat(d, 1, devices=[0, 99])


@pytest.mark.parametrize("devices", DEVICE_SETS[:3], ids=_ids)
def test_multi_concurrent_callers(built_lib, devices):
    """Searches from many threads on one multi-GPU index (the Morpheus pattern, faiss_vdb_service.py:169-174: B
    executor threads with one query each) serialise inside the library and all return the right rows; a writer
    between two waves of readers is seen by the second wave."""
    from concurrent.futures import ThreadPoolExecutor

    from generativeaiexamples_b200 import B200IndexFlat

    rng = np.random.default_rng(41)
    d = 256
    x = rng.standard_normal((20000, d), dtype=np.float32)
    ix = B200IndexFlat(d, 1, devices=devices)
    ix.add(x[:15000])

    def one(i):
        D, I = ix.search(x[i: i + 1], 3)
        return int(I[0, 0]), float(D[0, 0])

    rows = rng.integers(0, 15000, 200).tolist()
    with ThreadPoolExecutor(16) as ex:
        got = list(ex.map(one, rows))
    assert [g[0] for g in got] == rows and all(g[1] == 0.0 for g in got)
    ix.add(x[15000:])
    assert ix.remove_ids(np.arange(0, 100, dtype=np.int64)) == 100  # positions shift down by 100
    rows = rng.integers(100, 20000, 200).tolist()
    with ThreadPoolExecutor(16) as ex:
        got = list(ex.map(one, rows))
    assert [g[0] for g in got] == [r - 100 for r in rows] and all(g[1] == 0.0 for g in got)
    ix.close()
