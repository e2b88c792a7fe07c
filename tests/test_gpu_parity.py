"""GPU suite: parity of the CUDA path (through the C ABI) with the oracle.

Bar (BASELINE.json north_star): fp32 mode -- top-k ids identical to the reference's except
ties within 1e-5 relative score, scores within 1e-5 (abs-or-rel, oracle/compare.py);
integer outputs (ids, ntotal, remove_ids) bit-exact.  bf16 corpus mode: recall@k.
"""
import ctypes
import os

import numpy as np
import pytest

from generativeaiexamples_b200 import _lib

pytestmark = pytest.mark.gpu


def _device_sets():
    """Index placements every mutable-corpus test runs on: one GPU, two shards on one GPU (the
    multi-GPU code path -- segment map, per-shard workers, gather + final select -- on a 1-GPU box)
    and, when the box has them, two real GPUs."""
    sets = [None, [0, 0]]
    try:
        import torch

        if torch.cuda.device_count() >= 2:
            sets.append([0, 1])
    except Exception:
        pass
    return sets


DEVICE_SETS = _device_sets()

HERE = os.path.dirname(os.path.abspath(__file__))
FLT_MAX = np.finfo(np.float32).max


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "seeded_topk.npz"))


def _mk(metric, d, **kw):
    from generativeaiexamples_b200 import B200IndexFlat

    return B200IndexFlat(d, 1 if metric == "l2" else 0, **kw)


def _check(metric, D, I, q, x, k, what):
    from oracle import FlatRef, assert_topk_parity, truth_topk

    Dt, It = truth_topk(metric, q, x, k + 8)
    assert_topk_parity(D, I, Dt, It, what=what + " vs f64 truth")
    Dr, Ir = FlatRef.knn(1 if metric == "l2" else 0, q, x, k)
    scale = 1.0
    if metric == "l2" and q.shape[0] >= 20:
        # faiss's nq >= 20 path computes ||q||^2 + ||x||^2 - 2 q.x in fp32: its own error is
        # ~1e-7 of those magnitudes, however small the distance.  (The f64 check above is strict.)
        scale = float((q.astype(np.float64) ** 2).sum(1).max() + (x.astype(np.float64) ** 2).sum(1).max())
    assert_topk_parity(D, I, Dr, Ir, what=what + " vs faiss restatement", scale=scale)
    # structural: sorted best-first
    s = D if metric == "l2" else -D
    assert (np.diff(s, axis=1) >= 0).all(), what + ": scores not sorted"


def test_kat_through_abi(built_lib):
    import json

    kat = json.load(open(os.path.join(HERE, "golden", "kat_d4.json")))
    for case in kat["cases"]:
        ix = _mk(case["metric"], 4)
        ix.add(np.array(case["corpus"], dtype=np.float32))
        D, I = ix.search(np.array([case["query"]], dtype=np.float32), case["k"])
        pad = FLT_MAX if case["metric"] == "l2" else -FLT_MAX
        np.testing.assert_array_equal(I[0], np.array(case["ids"]), err_msg=case["name"])
        np.testing.assert_array_equal(D[0], np.array([pad if s is None else s for s in case["scores"]],
                                                     dtype=np.float32), err_msg=case["name"])
        ix.close()
    # normalize_L2 and remove_ids known answers
    from generativeaiexamples_b200 import normalize_L2

    x = np.array(kat["normalize"]["rows"], dtype=np.float32)
    normalize_L2(x)
    np.testing.assert_allclose(x, np.array(kat["normalize"]["expect"], dtype=np.float32), rtol=1e-6, atol=0)
    r = kat["remove_ids"]
    ix = _mk("l2", 4)
    ix.add(np.array(r["rows"], dtype=np.float32))
    assert ix.remove_ids(np.array(r["ids"], dtype=np.int64)) == r["n_removed"]
    assert ix.ntotal == len(r["expect"])
    for i, row in enumerate(r["expect"]):
        np.testing.assert_array_equal(ix.reconstruct(i), np.array(row, dtype=np.float32))
    ix.close()


def test_golden_fixtures(built_lib, gold):
    """Committed float64 results generated before any kernel existed."""
    from oracle import assert_topk_parity
    from tests.golden.make_fixtures import CASES, make_case

    for name, (n, d, nq, k, metric, normalized, seed) in CASES.items():
        x, q = make_case(*CASES[name])
        ix = _mk(metric, d)
        ix.add(x)
        D, I = ix.search(q, k)
        assert_topk_parity(D, I, gold[name + "__D"], gold[name + "__I"], what=name)
        ix.close()


@pytest.mark.parametrize("metric", ["ip", "l2"])
@pytest.mark.parametrize("d", [1024, 384, 100, 4, 2048, 8192])
def test_scan_parity_dims(built_lib, metric, d):
    rng = np.random.default_rng(d)
    n = 30011 if d <= 1024 else (9001 if d <= 2048 else 2003)  # ragged: not a multiple of the tile height
    x = rng.standard_normal((n, d), dtype=np.float32)
    q = rng.standard_normal((6, d), dtype=np.float32)
    q[::2] = x[rng.integers(0, n, 3)] + 0.05 * q[::2]
    ix = _mk(metric, d)
    ix.add(x)
    assert ix.ntotal == n
    for nq, k in ((1, 1), (1, 10), (3, 4), (4, 32), (6, 33), (2, 128)):
        D, I = ix.search(q[:nq], k)
        _check(metric, D, I, q[:nq], x, k, f"{metric} d={d} nq={nq} k={k}")
    ix.close()


@pytest.mark.parametrize("metric", ["ip", "l2"])
def test_edge_cases(built_lib, metric):
    d = 64
    ix = _mk(metric, d)
    pad = FLT_MAX if metric == "l2" else -FLT_MAX
    # empty index: all padding (faiss: heap initial values)
    D, I = ix.search(np.ones((2, d), dtype=np.float32), 5)
    assert (I == -1).all() and (D == pad).all()
    # fewer rows than k; incremental adds keep positions
    rng = np.random.default_rng(3)
    x = rng.standard_normal((7, d), dtype=np.float32)
    ix.add(x[:3])
    ix.add(x[3:])
    D, I = ix.search(x[:1], 10)
    assert (I[0, 7:] == -1).all() and (D[0, 7:] == pad).all() and sorted(I[0, :7].tolist()) == list(range(7))
    _check(metric, D[:, :7], I[:, :7], x[:1], x, 7, f"{metric} k>ntotal")
    # exact duplicates: ties resolved by lowest id
    ix.reset()
    assert ix.ntotal == 0
    dup = np.repeat(x[:1], 50, axis=0)
    ix.add(dup)
    D, I = ix.search(x[:1], 10)
    assert I[0].tolist() == list(range(10)) and len(set(D[0].tolist())) == 1
    # argument errors mirror the faiss wrapper
    with pytest.raises(AssertionError):
        ix.search(np.ones((1, d + 1), dtype=np.float32), 3)
    with pytest.raises(AssertionError):
        ix.search(np.ones((1, d), dtype=np.float32), 0)
    with pytest.raises(AssertionError):
        ix.add(np.ones((1, d - 1), dtype=np.float32))
    with pytest.raises(RuntimeError):
        ix.reconstruct(10 ** 6)
    ix.close()


def test_remove_ids_parity(built_lib):
    from oracle.flat_ref import IndexFlatRef

    rng = np.random.default_rng(9)
    n, d = 50000, 128
    x = rng.standard_normal((n, d), dtype=np.float32)
    ix, ref = _mk("l2", d), IndexFlatRef(d, 1)
    ix.add(x)
    ref.add(x)
    for ids in (np.array([0]), np.array([n - 2]), rng.integers(0, n, 5000), np.arange(100, 20000, 3),
                np.array([-5, 10 ** 9])):
        ids = ids.astype(np.int64)
        assert ix.remove_ids(ids) == ref.remove_ids(ids)
        assert ix.ntotal == ref.ntotal
        for key in (0, 1, ref.ntotal // 2, ref.ntotal - 1):
            np.testing.assert_array_equal(ix.reconstruct(key), ref.reconstruct(key))
        q = x[rng.integers(0, n, 2)]
        D, I = ix.search(q, 5)
        Dr, Ir = ref.search(q, 5)
        np.testing.assert_array_equal(I, Ir)
    ix.close()


def test_normalize_fused_and_standalone(built_lib):
    """C5: ingest-time normalisation + normalised queries (rag_service.py:194,292)."""
    from generativeaiexamples_b200 import B200IndexFlat, normalize_L2
    from oracle import FlatRef

    rng = np.random.default_rng(21)
    n, d = 20000, 1024
    x = (rng.standard_normal((n, d), dtype=np.float32) * rng.uniform(0.1, 10, (n, 1)).astype(np.float32))
    x[17] = 0  # zero row stays zero
    q = x[rng.integers(0, n, 3)] + 0.1 * rng.standard_normal((3, d), dtype=np.float32)
    xn, qn = x.copy(), q.copy()
    FlatRef.normalize_L2(xn)
    FlatRef.normalize_L2(qn)
    x2 = x.copy()
    normalize_L2(x2)
    np.testing.assert_allclose(x2, xn, rtol=2e-6, atol=1e-9)
    ix = B200IndexFlat(d, 1, normalize_on_add=True, normalize_queries=True)
    ix.add(x)
    np.testing.assert_allclose(ix.reconstruct(5), xn[5], rtol=2e-6, atol=1e-9)
    assert not ix.reconstruct(17).any()
    D, I = ix.search(q, 10)
    _check("l2", D, I, qn, xn, 10, "fused normalise")
    ix.close()


def test_device_tensors_roundtrip(built_lib):
    """torch hand-off: CUDA tensors in, CUDA tensors out, on the current stream."""
    import torch

    rng = np.random.default_rng(33)
    n, d = 40000, 1024
    x = rng.standard_normal((n, d), dtype=np.float32)
    q = x[:4] + 0.01
    ix = _mk("ip", d)
    ix.add(torch.from_numpy(x).cuda())
    D, I = ix.search(torch.from_numpy(q).cuda(), 10)
    assert D.is_cuda and I.is_cuda and I.dtype == torch.int64
    _check("ip", D.cpu().numpy(), I.cpu().numpy(), q, x, 10, "device tensors")
    Dh, Ih = ix.search(q, 10)
    np.testing.assert_array_equal(Ih, I.cpu().numpy())
    np.testing.assert_array_equal(Dh, D.cpu().numpy())
    ix.close()


def test_bf16_recall(built_lib):
    """Optional bf16 corpus mode: reported as recall@k against the fp32 truth."""
    from oracle import recall_at_k, truth_topk

    rng = np.random.default_rng(44)
    n, d, k = 50000, 1024, 100
    x = rng.standard_normal((n, d), dtype=np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    q = x[rng.integers(0, n, 4)] + 0.3 * rng.standard_normal((4, d), dtype=np.float32) / np.sqrt(d)
    ix = _mk("ip", d, storage="bf16")
    ix.add(x)
    D, I = ix.search(q, k)
    _, It = truth_topk("ip", q, x, k)
    r = recall_at_k(I, It)
    assert r >= 0.95, r
    # and exact w.r.t. the bf16-rounded corpus it actually stores
    import torch

    xb = torch.from_numpy(x).to(torch.bfloat16).to(torch.float32).numpy()
    np.testing.assert_array_equal(ix.reconstruct(123), xb[123])
    _check("ip", D, I, q, xb, k, "bf16 storage vs rounded corpus")
    ix.close()


def test_save_load_faiss_file(built_lib, tmp_path):
    import struct

    from generativeaiexamples_b200 import B200IndexFlat

    rng = np.random.default_rng(55)
    x = rng.standard_normal((3000, 100), dtype=np.float32)
    ix = _mk("l2", 100)
    ix.add(x)
    p = str(tmp_path / "index.faiss")
    ix.write(p)
    raw = open(p, "rb").read()
    assert raw[:4] == b"IxF2"
    d, nt, _, _, trained, metric, nfl = struct.unpack("<iqqqBiQ", raw[4:4 + 4 + 8 + 8 + 8 + 1 + 4 + 8])
    assert (d, nt, trained, metric, nfl) == (100, 3000, 1, 1, 300000)
    np.testing.assert_array_equal(np.frombuffer(raw[45:], dtype=np.float32).reshape(3000, 100), x)
    ix2 = B200IndexFlat.read(p)
    assert ix2.ntotal == 3000 and ix2.d == 100 and ix2.metric_type == 1
    D1, I1 = ix.search(x[:3], 5)
    D2, I2 = ix2.search(x[:3], 5)
    np.testing.assert_array_equal(I1, I2)
    np.testing.assert_array_equal(D1, D2)
    ix.close()
    ix2.close()


def test_concurrent_searches(built_lib):
    """faiss_vdb_service.py:169-174: B threads each running an nq=1 search."""
    from concurrent.futures import ThreadPoolExecutor

    rng = np.random.default_rng(66)
    n, d = 30000, 256
    x = rng.standard_normal((n, d), dtype=np.float32)
    ix = _mk("l2", d)
    ix.add(x)
    rows = rng.integers(0, n, 32)
    with ThreadPoolExecutor(8) as ex:
        res = list(ex.map(lambda r: ix.search(x[r:r + 1], 3)[1][0, 0], rows))
    assert res == rows.tolist()
    ix.close()


@pytest.mark.parametrize("devices", DEVICE_SETS, ids=lambda v: "dev" + "".join(map(str, v or [0])))
def test_vectorstore_end_to_end(built_lib, devices, monkeypatch):
    """basic_rag call sequence on the real engine against the LangChain restatement."""
    import warnings
    import zlib

    from generativeaiexamples_b200 import (B200FlatVectorStore, create_vectorstore_langchain,
                                           del_docs_vectorstore_langchain, get_docs_vectorstore_langchain,
                                           get_vectorstore)

    if devices:
        monkeypatch.setenv("B200R_DEVICES", ",".join(map(str, devices)))
        monkeypatch.setenv("B200R_MULTI_MIN_PIECE", "16")
    from oracle.flat_ref import IndexFlatRef
    from oracle.lc_faiss_shim import RefDocstore, RefFAISS

    d = 1024

    class Emb:
        def _v(self, t):
            r = np.random.default_rng(zlib.crc32(t.encode()))
            base = r.standard_normal(d)
            if t.startswith("about "):  # a query near its document: cos ~ 0.8
                base = np.asarray(Emb()._v(t[6:])) + 0.75 * base / np.linalg.norm(base)
            return (base / np.linalg.norm(base)).astype(np.float32).tolist()

        def embed_documents(self, texts):
            return [self._v(t) for t in texts]

        def embed_query(self, t):
            return self._v(t)

    monkeypatch.setenv("APP_EMBEDDINGS_DIMENSIONS", str(d))
    vs = get_vectorstore(None, Emb())  # utils.py:335-347: created lazily on first use ...
    assert isinstance(vs, B200FlatVectorStore) and vs.index.d == d
    assert get_vectorstore(vs, Emb()) is vs  # ... and handed back afterwards
    assert vs.index.stats()["n_devices"] == (len(devices) if devices else 1)
    other = create_vectorstore_langchain(Emb(), name="b200flat", dimensions=d)
    assert isinstance(other, B200FlatVectorStore) and other is not vs
    ref = RefFAISS(Emb(), IndexFlatRef(d, 1), RefDocstore(), {})
    texts = [f"chunk {i}" for i in range(300)]
    metas = [{"source": f"/tmp-data/uploaded_files/f{i % 6}.pdf"} for i in range(300)]
    ids = [f"id{i}" for i in range(300)]
    vs.add_texts(texts, metas, ids=list(ids))
    ref.add_texts(texts, metas, ids=list(ids))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for query in ("chunk 5", "about chunk 17", "nothing like it"):
            r = vs.as_retriever(search_type="similarity_score_threshold", search_kwargs={"score_threshold": 0.25, "k": 4})
            got = r.get_relevant_documents(query)
            want = ref.retrieve(query, "similarity_score_threshold", {"score_threshold": 0.25, "k": 4})
            assert [g.page_content for g in got] == [w.page_content for w in want], query
            a = vs.similarity_search_with_score(query, k=25)
            b = ref.similarity_search_with_score(query, k=25)
            assert [x.page_content for x, _ in a] == [x.page_content for x, _ in b]
            np.testing.assert_allclose([s for _, s in a], [s for _, s in b], rtol=1e-5, atol=1e-5)
    assert [g.page_content for g in vs.as_retriever().get_relevant_documents("about chunk 17")][0] == "chunk 17"
    vs.delete([ids[17], ids[5]])
    ref.delete([ids[17], ids[5]])
    assert vs.index.ntotal == 298 and vs.index_to_docstore_id == ref.index_to_docstore_id
    a = vs.similarity_search("about chunk 18", k=3)
    b = ref.similarity_search("about chunk 18", k=3)
    assert [x.page_content for x in a] == [x.page_content for x in b]
    # the chain server's document listing / deletion helpers (utils.py:492-603, FAISS branches)
    assert sorted(get_docs_vectorstore_langchain(vs)) == [f"f{i}.pdf" for i in range(6)]
    assert del_docs_vectorstore_langchain(vs, ["f3.pdf"]) is True
    ref.delete([i_ for i_, m_ in zip(ids, metas) if m_["source"].endswith("f3.pdf") and i_ not in (ids[17], ids[5])])
    assert vs.index.ntotal == ref.index.ntotal == 248 and vs.index_to_docstore_id == ref.index_to_docstore_id
    assert sorted(get_docs_vectorstore_langchain(vs)) == [f"f{i}.pdf" for i in (0, 1, 2, 4, 5)]
    assert del_docs_vectorstore_langchain(vs, ["never-uploaded.pdf"]) is False
    a = vs.similarity_search_with_score("about chunk 21", k=6)
    b = ref.similarity_search_with_score("about chunk 21", k=6)
    assert [x.page_content for x, _ in a] == [x.page_content for x, _ in b]
    # a chunk without metadata['source'] makes the reference log an error and answer [] / False
    vs.add_texts(["orphan"], [{}])
    assert get_docs_vectorstore_langchain(vs) == []
    assert del_docs_vectorstore_langchain(vs, ["f0.pdf"]) is False


def test_merge_kernel(built_lib):
    """K4 against a numpy merge with the same tie rule."""
    import torch

    from generativeaiexamples_b200 import _lib
    from tests.test_sharded_gloo import _merge_ref

    rng = np.random.default_rng(77)
    for metric, G, nq, k in ((0, 8, 1, 10), (1, 4, 5, 33), (0, 2, 256, 10), (1, 8, 3, 128), (0, 3, 2, 300), (1, 2, 1, 1024)):
        cd = rng.standard_normal((G, nq, k)).astype(np.float32)
        cd = np.sort(cd, axis=2) if metric == 1 else -np.sort(-cd, axis=2)
        cd[:, :, 1] = cd[:, :, 0]  # force ties
        ci = rng.permutation(G * nq * k).reshape(G, nq, k).astype(np.int64)
        ci[-1, :, k - 2:] = -1  # padding in the last shard
        De, Ie = _merge_ref(metric, cd, ci, k)
        tcd, tci = torch.from_numpy(cd).cuda(), torch.from_numpy(ci).cuda()
        D = torch.empty((nq, k), dtype=torch.float32, device="cuda")
        I = torch.empty((nq, k), dtype=torch.int64, device="cuda")
        _lib.check(built_lib.b200r_merge_topk(metric, ctypes.c_void_p(tcd.data_ptr()), ctypes.c_void_p(tci.data_ptr()),
                                              nq, G, k, nq * k, nq * k, k, ctypes.c_void_p(D.data_ptr()),
                                              ctypes.c_void_p(I.data_ptr()), 0,
                                              ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
        torch.cuda.synchronize()
        np.testing.assert_array_equal(I.cpu().numpy(), Ie)
        np.testing.assert_array_equal(D.cpu().numpy(), De)


# ----------------------------------------------------------------------------------------
# K3: tcgen05 GEMM filter + exact re-score (nq >= 20, faiss's BLAS-path regime)
# ----------------------------------------------------------------------------------------
@pytest.mark.parametrize("metric", ["ip", "l2"])
@pytest.mark.parametrize("d,n", [(1024, 40000), (384, 30011), (100, 9001), (2048, 5000)])
def test_gemm_parity(built_lib, metric, d, n):
    """AUTO dispatch sends nq >= 5 to the tensor cores; results must still be fp32-exact."""
    from generativeaiexamples_b200 import _lib

    rng = np.random.default_rng(1000 + d)
    x = rng.standard_normal((n, d), dtype=np.float32)
    if metric == "l2":
        x *= rng.uniform(0.5, 1.5, (n, 1)).astype(np.float32)  # un-normalised rows: true L2
    q = rng.standard_normal((260, d), dtype=np.float32)
    q[::2] = x[rng.integers(0, n, 130)] + 0.05 * q[::2]
    ix = _mk(metric, d)
    ix.add(x)
    for nq, k in ((5, 10), (20, 10), (64, 4), (129, 1), (260, 10), (33, 100), (200, 128)):
        D, I = ix.search(q[:nq], k)
        assert ix.stats()["last_path"] == _lib.PATH_GEMM
        _check(metric, D, I, q[:nq], x, k, f"gemm {metric} d={d} nq={nq} k={k}")
    # below the threshold the scan path answers, bit-identically for the shared queries
    D20, I20 = ix.search(q[:20], 10)
    D4, I4 = ix.search(q[:4], 10)
    assert ix.stats()["last_path"] == _lib.PATH_SCAN
    np.testing.assert_array_equal(I20[:4], I4)
    np.testing.assert_array_equal(D20[:4], D4)  # re-score uses the scan kernel's summation order
    ix.close()


def test_gemm_forced_small_batches_and_tiny_corpus(built_lib):
    from generativeaiexamples_b200 import B200IndexFlat, _lib

    rng = np.random.default_rng(5)
    d = 256
    for n in (1, 7, 255, 256, 257, 3000):
        x = rng.standard_normal((n, d), dtype=np.float32)
        q = rng.standard_normal((3, d), dtype=np.float32)
        for metric in ("ip", "l2"):
            ix = B200IndexFlat(d, 1 if metric == "l2" else 0, search_path=_lib.PATH_GEMM)
            ix.add(x)
            for k in (1, 10, 50):
                D, I = ix.search(q, k)
                kk = min(k, n)
                assert (I[:, kk:] == -1).all()
                _check(metric, D[:, :kk], I[:, :kk], q, x, kk, f"forced gemm n={n} {metric} k={k}")
            ix.close()


def test_gemm_containment_check_falls_back(built_lib):
    """Rows closer together than the tf32 error bound cannot be separated by the filter: the
    containment check must notice and the scan path must deliver the exact answer."""
    from generativeaiexamples_b200 import B200IndexFlatIP

    rng = np.random.default_rng(6)
    n, d = 20000, 1024
    x = rng.standard_normal((n, d), dtype=np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    centre = x[123].copy()
    x[1000:1200] = centre + 1e-4 * rng.standard_normal((200, d), dtype=np.float32)  # 200 near-duplicates
    q = np.repeat(centre[None, :], 24, axis=0) + 1e-3 * rng.standard_normal((24, d), dtype=np.float32)
    ix = B200IndexFlatIP(d)
    ix.add(x)
    D, I = ix.search(q, 10)
    assert ix.stats()["n_gemm_fallbacks"] > 0
    _check("ip", D, I, q, x, 10, "gemm fallback")
    ix.close()


def test_gemm_bf16_storage(built_lib):
    from oracle import recall_at_k, truth_topk
    import torch

    rng = np.random.default_rng(8)
    n, d, k = 60000, 1024, 100
    x = rng.standard_normal((n, d), dtype=np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    q = x[rng.integers(0, n, 64)] + 0.3 * rng.standard_normal((64, d), dtype=np.float32) / np.sqrt(d)
    ix = _mk("ip", d, storage="bf16")
    ix.add(x)
    D, I = ix.search(q, k)
    _, It = truth_topk("ip", q, x, k)
    assert recall_at_k(I, It) >= 0.95
    xb = torch.from_numpy(x).to(torch.bfloat16).to(torch.float32).numpy()
    _check("ip", D, I, q, xb, k, "gemm bf16 vs rounded corpus")
    ix.close()


@pytest.mark.parametrize("metric", ["ip", "l2"])
def test_large_k(built_lib, metric):
    """128 < k <= 1024: single-query wide-list scan variants and the GEMM path with a larger K'
    ("typically 1000 for real world", RAG/notebooks/langchain/langchain_basic_RAG.ipynb:445)."""
    from generativeaiexamples_b200 import _lib

    rng = np.random.default_rng(77)
    n, d = 30000, 256
    x = rng.standard_normal((n, d), dtype=np.float32)
    q = rng.standard_normal((40, d), dtype=np.float32)
    q[::2] = x[rng.integers(0, n, 20)] + 0.1 * q[::2]
    ix = _mk(metric, d)
    ix.add(x)
    for nq, k, path in ((1, 129, _lib.PATH_SCAN), (1, 1000, _lib.PATH_SCAN), (3, 300, _lib.PATH_SCAN),
                        (8, 500, _lib.PATH_GEMM), (40, 1024, _lib.PATH_GEMM)):
        D, I = ix.search(q[:nq], k)
        assert ix.stats()["last_path"] == path
        _check(metric, D, I, q[:nq], x, k, f"large k {metric} nq={nq} k={k}")
    with pytest.raises(RuntimeError):
        ix.search(q[:1], 1025)
    # k larger than the corpus: padding
    small = _mk(metric, d)
    small.add(x[:300])
    D, I = small.search(q[:2], 700)
    assert (I[:, 300:] == -1).all() and sorted(I[0, :300].tolist()) == list(range(300))
    _check(metric, D[:, :300], I[:, :300], q[:2], x[:300], 300, "large k > ntotal")
    ix.close()
    small.close()


def test_coalesced_concurrent_searches(built_lib):
    """f2 on the real engine: 64 threads, one query each, served by a handful of batched launches."""
    from concurrent.futures import ThreadPoolExecutor

    from generativeaiexamples_b200.batching import SearchCoalescer

    rng = np.random.default_rng(67)
    n, d = 200000, 256
    x = rng.standard_normal((n, d), dtype=np.float32)
    ix = _mk("l2", d)
    ix.add(x)
    co = SearchCoalescer(ix)
    rows = rng.integers(0, n, 64)
    calls0 = ix.stats()["n_search_calls"]
    with ThreadPoolExecutor(64) as ex:
        res = list(ex.map(lambda r: co.search(x[r:r + 1], 3), rows))
    assert [int(I[0, 0]) for _, I in res] == rows.tolist()
    launches = ix.stats()["n_search_calls"] - calls0
    assert launches == co.n_launches and launches < 64
    single = [ix.search(x[r:r + 1], 3) for r in rows[:8]]
    for (D1, I1), (D2, I2) in zip(res[:8], single):
        np.testing.assert_array_equal(I1, I2)
        np.testing.assert_array_equal(D1, D2)
    ix.close()


def test_incremental_growth_and_interleaved_ops(built_lib):
    """Ingest the way the chain server does -- a few hundred chunks per uploaded file, searches in
    between, a delete now and then -- so that the HBM buffer re-allocates several times."""
    from oracle.flat_ref import IndexFlatRef

    rng = np.random.default_rng(91)
    d = 384
    ix, ref = _mk("l2", d), IndexFlatRef(d, 1)
    for step in range(12):
        m = int(rng.integers(50, 900))
        x = rng.standard_normal((m, d), dtype=np.float32)
        ix.add(x)
        ref.add(x)
        assert ix.ntotal == ref.ntotal
        if step % 4 == 3:
            ids = rng.integers(0, ref.ntotal, 40).astype(np.int64)
            assert ix.remove_ids(ids) == ref.remove_ids(ids)
        q = ref._codes[rng.integers(0, ref.ntotal, 6)] + 0.01
        for nq in (1, 6):
            D, I = ix.search(q[:nq], 5)
            _check("l2", D, I, q[:nq], ref._codes, 5, f"growth step {step} nq={nq}")
    assert ix.stats()["capacity_rows"] >= ix.ntotal > 1024  # grew past the initial allocation
    np.testing.assert_array_equal(ix.reconstruct(ix.ntotal - 1), ref.reconstruct(ref.ntotal - 1))
    ix.close()


def test_reconstruct_batch_and_mmr_on_engine(built_lib):
    from generativeaiexamples_b200 import B200FlatVectorStore

    rng = np.random.default_rng(92)
    n, d = 5000, 1024
    for storage in ("f32", "bf16"):
        x = rng.standard_normal((n, d), dtype=np.float32)
        ix = _mk("l2", d, storage=storage)
        ix.add(x)
        keys = rng.integers(0, n, 37)
        got = ix.reconstruct_batch(keys)
        want = np.stack([ix.reconstruct(int(i)) for i in keys])
        np.testing.assert_array_equal(got, want)
        with pytest.raises(RuntimeError):
            ix.reconstruct_batch(np.array([0, n]))
        ix.close()
    vs = B200FlatVectorStore(lambda t: [0.0] * d, d)
    vs.add_embeddings([(f"t{i}", x[i].tolist()) for i in range(200)])
    docs = vs.max_marginal_relevance_search_by_vector(x[5].tolist(), k=4, fetch_k=20)
    assert docs[0].page_content == "t5" and len(docs) == 4 and len({d_.page_content for d_ in docs}) == 4


def test_cross_stream_ordering(built_lib):
    """add() with a CUDA tensor is asynchronous on the caller's stream; a host search right after
    it runs on the index's own stream and must still see the rows (the library orders them)."""
    import torch

    d = 1024
    n = 400_000
    side = torch.cuda.Stream()
    for _ in range(3):
        ix = _mk("ip", d, reserve_rows=n)
        with torch.cuda.stream(side):
            x = torch.randn((n, d), device="cuda")
            x /= x.norm(dim=1, keepdim=True)
            ix.add(x)                                   # queued behind the generation kernels
            q = x[n - 5: n - 4].cpu().numpy()           # (this D2H also runs on `side`)
        D, I = ix.search(q, 1)                          # host path, the index's own stream
        assert I[0, 0] == n - 5 and D[0, 0] > 0.999
        np.testing.assert_array_equal(ix.reconstruct(n - 5), q[0])
        ix.close()


def test_adds_on_two_streams_with_growth(built_lib):
    """Two device adds on different streams, the second one growing the buffer: the growth copies the rows the
    first ingest kernel writes, so the second add must be ordered behind the first (it is: b200r_add waits for
    the last writer's event before anything else).  Every row must be there afterwards."""
    import torch

    d = 1024
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for rep in range(3):
        ix = _mk("l2", d)  # no reserve_rows: the second add re-allocates
        with torch.cuda.stream(s1):
            a = torch.randn((300_000, d), device="cuda")
            a = a * 1.0  # a few more kernels in front of the ingest on s1
            ix.add(a)
        with torch.cuda.stream(s2):
            b = torch.randn((200_000, d), device="cuda")
            ix.add(b)  # grows 300k -> 500k rows while the first ingest may still be running
        host = torch.randn((1000, d)).numpy()
        ix.add(host)   # and a host add on the index's own stream on top
        assert ix.ntotal == 501_000
        torch.cuda.synchronize()
        ah, bh = a.cpu().numpy(), b.cpu().numpy()
        keys = np.array([0, 1, 149_999, 299_999, 300_000, 300_001, 499_999, 500_000, 500_999], dtype=np.int64)
        want = np.stack([ah[0], ah[1], ah[149_999], ah[299_999], bh[0], bh[1], bh[199_999], host[0], host[999]])
        np.testing.assert_array_equal(ix.reconstruct_batch(keys), want)
        D, I = ix.search(np.stack([ah[77], bh[77], host[77]]), 1)
        assert I[:, 0].tolist() == [77, 300_077, 500_077] and (D[:, 0] == 0).all()
        ix.close()


def test_gemm_search_right_after_an_add_on_another_stream(built_lib):
    """The GEMM route's containment check uses the largest row norm of the index, refreshed lazily by a small
    kernel.  A search on one stream right after a device add of LARGE-norm rows on another stream must not compute
    that maximum from norms the ingest kernel has not written yet (the search now waits for the last writer before
    it enqueues anything)."""
    import torch

    d, n = 256, 200_000
    side = torch.cuda.Stream()
    for rep in range(3):
        ix = _mk("ip", d, search_path=_lib.PATH_GEMM, reserve_rows=2 * n)
        small = torch.randn((n, d), device="cuda") * 0.01
        ix.add(small)
        q = torch.randn((40, d), device="cuda")
        ix.search(q, 10)  # the max-norm word now holds the small rows' maximum
        with torch.cuda.stream(side):
            big = torch.randn((n, d), device="cuda") * 30.0  # norms 3000x larger
            big = big + 0.0
            ix.add(big)
        D, I = ix.search(q, 10)  # default stream: must be ordered behind the ingest on `side`
        torch.cuda.synchronize()
        x = np.concatenate([small.cpu().numpy(), big.cpu().numpy()])
        _check("ip", D.cpu().numpy(), I.cpu().numpy(), q.cpu().numpy(), x, 10, "GEMM search after a side-stream add")
        assert (I.cpu().numpy() >= n).all()  # every winner is one of the large rows
        ix.close()


# bf16 shadow of an fp32 corpus: the filter streams the shadow, the answers stay fp32-exact
# ----------------------------------------------------------------------------------------
@pytest.mark.parametrize("metric", ["ip", "l2"])
@pytest.mark.parametrize("d,n", [(1024, 40000), (100, 9001), (388, 20011)])
def test_shadow_filter_parity(built_lib, metric, d, n):
    from generativeaiexamples_b200 import _lib

    rng = np.random.default_rng(4000 + d)
    x = rng.standard_normal((n, d), dtype=np.float32)
    if metric == "l2":
        x *= rng.uniform(0.5, 1.5, (n, 1)).astype(np.float32)
    q = rng.standard_normal((260, d), dtype=np.float32)
    q[::2] = x[rng.integers(0, n, 130)] + 0.05 * q[::2]
    ix = _mk(metric, d, shadow_bf16=True)
    plain = _mk(metric, d)
    half = n // 2
    ix.add(x[:half])
    ix.add(x[half:])  # the shadow follows growth
    plain.add(x)
    for nq, k in ((5, 10), (64, 4), (129, 1), (260, 10), (33, 100), (200, 128)):
        b0 = ix.stats()["bytes_scanned"]
        D, I = ix.search(q[:nq], k)
        st = ix.stats()
        assert st["last_path"] == _lib.PATH_GEMM
        # the filter pass streamed 2-byte elements (fallback scans, if any, add 4-byte passes)
        d_pad = (d + 7) // 8 * 8
        blocks = (nq + 255) // 256
        assert (st["bytes_scanned"] - b0 - blocks * n * d_pad * 2) % (n * d_pad * 4) == 0
        _check(metric, D, I, q[:nq], x, k, f"shadow {metric} d={d} nq={nq} k={k}")
        Dp, Ip = plain.search(q[:nq], k)
        np.testing.assert_array_equal(I, Ip)
        np.testing.assert_array_equal(D, Dp)  # same fp32 re-score either way
    # small batches still take the fp32 scan
    D4, I4 = ix.search(q[:4], 10)
    assert ix.stats()["last_path"] == _lib.PATH_SCAN
    Dp, Ip = plain.search(q[:4], 10)
    np.testing.assert_array_equal(I4, Ip)
    np.testing.assert_array_equal(D4, Dp)
    # run-time switch: filter on the fp32 rows instead
    ix.set_shadow_filter(False)
    b0 = ix.stats()["bytes_scanned"]
    D, I = ix.search(q[:64], 10)
    assert ix.stats()["bytes_scanned"] - b0 >= n * ((d + 7) // 8 * 8) * 4
    Dp, Ip = plain.search(q[:64], 10)
    np.testing.assert_array_equal(I, Ip)
    np.testing.assert_array_equal(D, Dp)
    ix.close()
    plain.close()


def test_shadow_filter_remove_and_fallback(built_lib):
    from generativeaiexamples_b200 import B200IndexFlat, B200IndexFlatIP

    rng = np.random.default_rng(41)
    n, d = 20000, 1024
    x = rng.standard_normal((n, d), dtype=np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    centre = x[123].copy()
    # 600 rows closer together than bf16 can tell apart: the filter cannot rank them
    x[1000:1600] = centre + 1e-4 * rng.standard_normal((600, d), dtype=np.float32)
    q = np.repeat(centre[None, :], 24, axis=0) + 1e-3 * rng.standard_normal((24, d), dtype=np.float32)
    ix = B200IndexFlatIP(d, shadow_bf16=True)
    ix.add(x)
    D, I = ix.search(q, 10)
    assert ix.stats()["n_gemm_fallbacks"] > 0
    _check("ip", D, I, q, x, 10, "shadow fallback")
    # delete: rows, norms and shadow must move together
    gone = np.unique(np.concatenate([rng.integers(0, n, 3000), np.arange(1000, 1590)]))
    assert ix.remove_ids(gone) == len(gone)
    keep = np.setdiff1d(np.arange(n), gone)
    qq = x[keep[rng.integers(0, len(keep), 48)]] + 0.02 * rng.standard_normal((48, d), dtype=np.float32)
    D, I = ix.search(qq, 10)
    _check("ip", D, I, qq, x[keep], 10, "shadow after remove_ids")
    ix.close()
    with pytest.raises(Exception):
        B200IndexFlat(d, 0, storage="bf16", shadow_bf16=True)
    plain = B200IndexFlatIP(d)
    with pytest.raises(Exception):
        plain.set_shadow_filter(True)
    plain.close()


@pytest.mark.parametrize("metric", ["ip", "l2"])
def test_shadow_small_batches(built_lib, metric, monkeypatch):
    """nq <= 4 on a shadow index: the scan on small corpora, the shadow GEMM filter on large ones (forced here
    through the size knob); the answers equal the plain index's bit for bit either way."""
    from generativeaiexamples_b200 import _lib

    rng = np.random.default_rng(77)
    for d, n in ((1024, 30000), (100, 9001), (384, 50), (64, 7)):
        x = rng.standard_normal((n, d), dtype=np.float32)
        if metric == "l2":
            x *= rng.uniform(0.5, 1.5, (n, 1)).astype(np.float32)
        q = rng.standard_normal((4, d), dtype=np.float32)
        q[::2] = x[rng.integers(0, n, 2)] + 0.05 * q[::2]
        ix = _mk(metric, d, shadow_bf16=True)
        plain = _mk(metric, d)
        ix.add(x)
        plain.add(x)
        d_pad = (d + 7) // 8 * 8
        for big in (False, True):
            if big:
                monkeypatch.setenv("B200R_GEMM_MIN_BYTES", "1")  # treat this corpus as a large one
                _lib.reload_env()
            else:
                monkeypatch.delenv("B200R_GEMM_MIN_BYTES", raising=False)
                _lib.reload_env()
            for nq, k in ((1, 10), (4, 10), (3, 1), (2, 14), (1, 100), (2, 200)):
                b0 = ix.stats()["bytes_scanned"]
                D, I = ix.search(q[:nq], k)
                st = ix.stats()
                if big:
                    assert st["last_path"] == _lib.PATH_GEMM
                    assert (st["bytes_scanned"] - b0 - n * d_pad * 2) % (n * d_pad * 4) == 0, (d, n, nq, k)
                else:
                    assert st["last_path"] == _lib.PATH_SCAN
                kk = min(k, n)
                assert (I[:, kk:] == -1).all()
                _check(metric, D[:, :kk], I[:, :kk], q[:nq], x, kk, f"shadow small {metric} d={d} n={n} nq={nq} k={k}")
                Dp, Ip = plain.search(q[:nq], k)
                np.testing.assert_array_equal(I, Ip)
                np.testing.assert_array_equal(D, Dp)
        ix.close()
        plain.close()


def test_small_batch_dispatch_by_corpus_size(built_lib, monkeypatch):
    """2..4 queries: scan on a small corpus, GEMM from ~200k rows (here: via the size knob); same answers."""
    from generativeaiexamples_b200 import B200IndexFlatL2, _lib

    rng = np.random.default_rng(79)
    n, d = 20000, 256
    x = rng.standard_normal((n, d), dtype=np.float32)
    q = x[:4] + 0.1 * rng.standard_normal((4, d), dtype=np.float32)
    ix = B200IndexFlatL2(d)
    ix.add(x)
    D0, I0 = ix.search(q, 10)
    assert ix.stats()["last_path"] == _lib.PATH_SCAN
    monkeypatch.setenv("B200R_GEMM_MIN_BYTES", "1")
    _lib.reload_env()
    for nq in (2, 3, 4):
        D1, I1 = ix.search(q[:nq], 10)
        assert ix.stats()["last_path"] == _lib.PATH_GEMM
        np.testing.assert_array_equal(I1, I0[:nq])
        np.testing.assert_array_equal(D1, D0[:nq])
    D1, I1 = ix.search(q[:1], 10)
    assert ix.stats()["last_path"] == _lib.PATH_SCAN  # a single query without a shadow always scans
    np.testing.assert_array_equal(I1, I0[:1])
    _check("l2", D0, I0, q, x, 10, "dispatch")
    ix.close()


def test_shadow_single_query_fallback(built_lib, monkeypatch):
    from generativeaiexamples_b200 import B200IndexFlatIP, _lib

    monkeypatch.setenv("B200R_GEMM_MIN_BYTES", "1")
    _lib.reload_env()
    rng = np.random.default_rng(78)
    n, d = 20000, 1024
    x = rng.standard_normal((n, d), dtype=np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    centre = x[123].copy()
    x[1000:1600] = centre + 1e-4 * rng.standard_normal((600, d), dtype=np.float32)
    q = np.repeat(centre[None, :], 3, axis=0) + 1e-3 * rng.standard_normal((3, d), dtype=np.float32)
    ix = B200IndexFlatIP(d, shadow_bf16=True)
    ix.add(x)
    for nq in (1, 3):
        f0 = ix.stats()["n_gemm_fallbacks"]
        D, I = ix.search(q[:nq], 10)
        assert ix.stats()["last_path"] == _lib.PATH_GEMM
        assert ix.stats()["n_gemm_fallbacks"] > f0
        _check("ip", D, I, q[:nq], x, 10, "shadow single-query fallback")
    f0 = ix.stats()["n_gemm_fallbacks"]
    q2 = x[5000:5001] + 0.05 * rng.standard_normal((1, d), dtype=np.float32)
    D, I = ix.search(q2, 10)
    assert ix.stats()["n_gemm_fallbacks"] == f0  # an ordinary query needs none
    _check("ip", D, I, q2, x, 10, "shadow single query")
    ix.close()


@pytest.mark.parametrize("shadow", [False, True])
def test_gemm_massive_ties(built_lib, shadow):
    """Thousands of identical rows: whatever the filter keeps, the answer is the lowest ids (faiss keeps the
    first-seen on ties) with exactly equal scores; the check must send these queries to the scan path."""
    from generativeaiexamples_b200 import B200IndexFlat

    rng = np.random.default_rng(91)
    n, d, k = 12000, 256, 10
    x = rng.standard_normal((n, d), dtype=np.float32)
    x[2000:9000] = x[2000]  # 7000 copies of one row
    q = np.repeat(x[2000][None, :], 8, axis=0)
    q[4:] = rng.standard_normal((4, d), dtype=np.float32)  # and four ordinary queries
    for metric in (0, 1):
        ix = B200IndexFlat(d, metric, shadow_bf16=shadow)
        ix.add(x)
        D, I = ix.search(q, k)
        assert ix.stats()["last_path"] == 2
        for i in range(4):
            np.testing.assert_array_equal(I[i], np.arange(2000, 2000 + k))
            assert (D[i] == D[i, 0]).all()
        _check("l2" if metric == 1 else "ip", D[4:], I[4:], q[4:], x, k, f"ties metric={metric} shadow={shadow}")
        ix.close()


def test_shadow_survives_load_and_reset(built_lib, tmp_path, monkeypatch):
    from generativeaiexamples_b200 import B200IndexFlat, B200IndexFlatIP, _lib

    monkeypatch.setenv("B200R_GEMM_MIN_BYTES", "1")
    _lib.reload_env()
    rng = np.random.default_rng(93)
    n, d = 9000, 200
    x = rng.standard_normal((n, d), dtype=np.float32)
    q = x[:6] + 0.05 * rng.standard_normal((6, d), dtype=np.float32)
    plain = B200IndexFlatIP(d)
    plain.add(x)
    path = str(tmp_path / "index.faiss")
    plain.write(path)
    ix = B200IndexFlat.read(path, shadow_bf16=True)  # the shadow is rebuilt while the file streams in
    assert ix.ntotal == n
    b0 = ix.stats()["bytes_scanned"]
    D, I = ix.search(q, 10)
    assert ix.stats()["last_path"] == _lib.PATH_GEMM
    assert (ix.stats()["bytes_scanned"] - b0) % (n * 200 * 4) == n * 200 * 2  # the filter streamed 2-byte rows
    Dp, Ip = plain.search(q, 10)
    np.testing.assert_array_equal(I, Ip)
    np.testing.assert_array_equal(D, Dp)
    ix.reset()
    assert ix.ntotal == 0
    ix.add(x[:500])
    D, I = ix.search(q, 10)
    _check("ip", D, I, q, x[:500], 10, "shadow after reset")
    ix.close()
    plain.close()


@pytest.mark.parametrize("seed", range(128))
def test_randomised_differential(built_lib, seed, monkeypatch):
    """Random shapes, metrics, routes and index options against the float64 truth: whatever the dispatch picks
    (scan, tf32 / bf16-shadow GEMM filter, fallbacks), ids and scores must satisfy the north-star tolerance."""
    from generativeaiexamples_b200 import B200IndexFlat, _lib
    from oracle import assert_topk_parity, truth_topk

    rng = np.random.default_rng(5000 + seed)
    d = int(rng.choice([1, 3, 4, 7, 8, 31, 64, 100, 257, 384, 768, 1024, 1100]))
    n = int(rng.choice([1, 2, 31, 255, 256, 257, 1000, 4099, 12000, 20000]))
    nq = int(rng.choice([1, 2, 3, 4, 5, 19, 20, 64, 129, 200]))
    k = int(rng.choice([1, 2, 4, 10, 32, 33, 100, 128, 129, 300]))
    metric = "l2" if rng.random() < 0.5 else "ip"
    shadow = bool(rng.random() < 0.4)
    path = int(rng.choice([_lib.PATH_AUTO, _lib.PATH_AUTO, _lib.PATH_SCAN, _lib.PATH_GEMM]))
    norm_add = bool(rng.random() < 0.3)
    bf16 = bool(not shadow and not norm_add and rng.random() < 0.3)  # stored rows = bf16(x): compare with those
    norm_q = bool(rng.random() < 0.3)
    if rng.random() < 0.5:
        monkeypatch.setenv("B200R_GEMM_MIN_BYTES", "1")
        _lib.reload_env()
    x = rng.standard_normal((n, d), dtype=np.float32) * rng.uniform(0.3, 3.0, (n, 1)).astype(np.float32)
    q = rng.standard_normal((nq, d), dtype=np.float32)
    src = rng.integers(0, n, nq)
    near = rng.random(nq) < 0.5
    q[near] = x[src[near]] + 0.05 * q[near]
    if bf16:
        import torch

        x = torch.from_numpy(x).to(torch.bfloat16).to(torch.float32).numpy()
    ix = B200IndexFlat(d, 1 if metric == "l2" else 0, shadow_bf16=shadow, search_path=path,
                       normalize_on_add=norm_add, normalize_queries=norm_q, storage="bf16" if bf16 else "f32")
    cut = int(rng.integers(0, n + 1))
    ix.add(x[:cut])
    ix.add(x[cut:])
    D, I = ix.search(q, k)

    def unit(a):
        nrm = np.sqrt((a.astype(np.float32) ** 2).sum(1, keepdims=True, dtype=np.float32))
        return np.where(nrm > 0, a / np.where(nrm > 0, nrm, 1), a).astype(np.float32)

    xe = unit(x) if norm_add else x
    qe = unit(q) if norm_q else q
    kk = min(k, n)
    what = f"seed={seed} d={d} n={n} nq={nq} k={k} {metric} shadow={shadow} bf16={bf16} path={path} norm={norm_add},{norm_q}"
    assert (I[:, kk:] == -1).all(), what
    Dt, It = truth_topk(metric, qe, xe, min(n, kk + 8))
    # fp32 rounding is relative to the magnitude of the operands, not of the result: a mid-ranked inner product
    # of two long vectors can be ~0 while carrying ~1e-7 * |q| * |x| of rounding.  The tolerance floor is
    # therefore |q|*|x| (the reference's own fp32 kernels are no closer to the float64 truth than that).
    qn = float(np.sqrt((qe.astype(np.float64) ** 2).sum(1).max()))
    xn = float(np.sqrt((xe.astype(np.float64) ** 2).sum(1).max())) if n else 0.0
    scale = max(1.0, qn * xn) if metric == "ip" else max(1.0, qn * qn + xn * xn)
    assert_topk_parity(D[:, :kk], I[:, :kk], Dt, It, what=what, scale=scale)
    s = D[:, :kk] if metric == "l2" else -D[:, :kk]
    assert (np.diff(s, axis=1) >= 0).all(), what + ": not sorted"
    ix.close()


@pytest.mark.parametrize("devices", DEVICE_SETS, ids=lambda v: "dev" + "".join(map(str, v or [0])))
@pytest.mark.parametrize("seed", range(8))
def test_randomised_mutation_sequences(built_lib, seed, devices, monkeypatch):
    """add / remove_ids / reset / search in random order against a numpy model of faiss's semantics
    (positions are ids; remove_ids compacts and shifts later ids down)."""
    from generativeaiexamples_b200 import B200IndexFlat
    from oracle import assert_topk_parity, truth_topk

    rng = np.random.default_rng(7000 + seed)
    d = int(rng.choice([8, 100, 256, 1024]))
    metric = "l2" if seed % 2 else "ip"
    shadow = bool(seed % 3 == 0)
    if seed % 4 < 2:
        monkeypatch.setenv("B200R_GEMM_MIN_BYTES", "1")
        _lib.reload_env()
    monkeypatch.setenv("B200R_MULTI_MIN_PIECE", "64")  # several segments per shard even at these sizes
    ix = B200IndexFlat(d, 1 if metric == "l2" else 0, shadow_bf16=shadow, devices=devices)
    model = np.zeros((0, d), dtype=np.float32)
    for step in range(14):
        op = rng.choice(["add", "add", "remove", "search", "search", "reset"], p=[0.3, 0.2, 0.2, 0.15, 0.1, 0.05])
        if op == "add" or model.shape[0] == 0:
            m = int(rng.choice([1, 5, 300, 3000]))
            x = rng.standard_normal((m, d), dtype=np.float32)
            ix.add(x)
            model = np.concatenate([model, x])
        elif op == "remove":
            ids = rng.integers(-5, model.shape[0] + 5, int(rng.choice([1, 10, 500])))
            valid = np.unique(ids[(ids >= 0) & (ids < model.shape[0])])
            assert ix.remove_ids(ids.astype(np.int64)) == len(valid)
            model = np.delete(model, valid, axis=0)
        elif op == "reset":
            ix.reset()
            model = np.zeros((0, d), dtype=np.float32)
        assert ix.ntotal == model.shape[0]
        if model.shape[0] == 0:
            D, I = ix.search(rng.standard_normal((2, d), dtype=np.float32), 3)
            assert (I == -1).all()
            continue
        nq = int(rng.choice([1, 3, 8, 40]))
        k = int(rng.choice([1, 10, 50]))
        q = rng.standard_normal((nq, d), dtype=np.float32)
        pick = rng.integers(0, model.shape[0], nq)
        q[::2] = model[pick[::2]] + 0.05 * q[::2]
        D, I = ix.search(q, k)
        kk = min(k, model.shape[0])
        assert (I[:, kk:] == -1).all()
        Dt, It = truth_topk(metric, q, model, min(model.shape[0], kk + 8))
        assert_topk_parity(D[:, :kk], I[:, :kk], Dt, It, what=f"seed={seed} step={step} op={op} n={model.shape[0]} nq={nq} k={k}")
        j = int(rng.integers(0, model.shape[0]))
        np.testing.assert_array_equal(ix.reconstruct(j), model[j])
    ix.close()


def test_merge_from(built_lib, monkeypatch):
    """index.merge_from / FAISS.merge_from (vectorstore_updater.py:217) against the restatement."""
    import zlib

    from generativeaiexamples_b200 import B200FlatVectorStore, B200IndexFlat
    from oracle import IndexFlatRef
    from oracle.lc_faiss_shim import RefDocstore, RefFAISS

    rng = np.random.default_rng(95)
    for d, kw_dst, kw_src in ((100, dict(shadow_bf16=True), dict()), (64, dict(), dict(shadow_bf16=True)),
                              (257, dict(storage="bf16"), dict(storage="bf16")), (1024, dict(), dict())):
        xa = rng.standard_normal((3000, d), dtype=np.float32)
        xb = rng.standard_normal((1234, d), dtype=np.float32)
        if kw_dst.get("storage") == "bf16":
            import torch

            xa = torch.from_numpy(xa).to(torch.bfloat16).to(torch.float32).numpy()
            xb = torch.from_numpy(xb).to(torch.bfloat16).to(torch.float32).numpy()
        for metric in (0, 1):
            a, b = B200IndexFlat(d, metric, **kw_dst), B200IndexFlat(d, metric, **kw_src)
            ra, rb = IndexFlatRef(d, metric), IndexFlatRef(d, metric)
            for ix, x in ((a, xa), (ra, xa), (b, xb), (rb, xb)):
                ix.add(x)
            a.merge_from(b)
            ra.merge_from(rb)
            assert a.ntotal == ra.ntotal == 4234 and b.ntotal == rb.ntotal == 0
            np.testing.assert_array_equal(a.reconstruct(3000 + 77), xb[77])
            q = np.concatenate([xa[:3], xb[:3]]) + 0.01 * rng.standard_normal((6, d), dtype=np.float32)
            for nq in (1, 6):  # scan and GEMM routes
                D, I = a.search(q[:nq], 5)
                _check("l2" if metric == 1 else "ip", D, I, q[:nq], np.concatenate([xa, xb]), 5, f"merge_from d={d} {kw_dst}")
            b.add(xa[:10])  # the emptied index stays usable
            assert b.search(xa[:1], 1)[1][0, 0] == 0
            a.close()
            b.close()
    x = B200IndexFlat(16, 1)
    with pytest.raises(RuntimeError):
        x.merge_from(B200IndexFlat(16, 0))
    with pytest.raises(RuntimeError):
        x.merge_from(B200IndexFlat(17, 1))
    with pytest.raises(RuntimeError):
        x.merge_from(B200IndexFlat(16, 1, storage="bf16"))
    with pytest.raises(RuntimeError):
        x.merge_from(x)

    # the LangChain surface
    def emb(t):
        return np.random.default_rng(zlib.crc32(t.encode())).standard_normal(64).astype(np.float32).tolist()

    class E:
        def embed_documents(self, ts):
            return [emb(t) for t in ts]

        def embed_query(self, t):
            return emb(t)

    s1, s2 = B200FlatVectorStore(E(), 64), B200FlatVectorStore(E(), 64)
    r1, r2 = RefFAISS(emb, IndexFlatRef(64), RefDocstore(), {}), RefFAISS(emb, IndexFlatRef(64), RefDocstore(), {})
    t1, t2 = [f"first {i}" for i in range(40)], [f"second {i}" for i in range(25)]
    for s, t in ((s1, t1), (r1, t1), (s2, t2), (r2, t2)):
        s.add_texts(t)
    s1.merge_from(s2)
    r1.merge_from(r2)
    assert s1.index.ntotal == 65 and sorted(s1.index_to_docstore_id) == list(range(65))
    for qt in ("second 7", "first 3", "something else"):
        got = [(d_.page_content, float(sc)) for d_, sc in s1.similarity_search_with_score(qt, k=4)]
        want = [(d_.page_content, float(sc)) for d_, sc in r1.similarity_search_with_score(qt, k=4)]
        assert [g[0] for g in got] == [w[0] for w in want]
        np.testing.assert_allclose([g[1] for g in got], [w[1] for w in want], rtol=1e-5, atol=1e-5)


def test_service_mirror_on_engine(built_lib, tmp_path):
    """faiss_vdb_service.py:63-66,147-174 on the real engine: load_local, then one batched launch per call."""
    import asyncio
    import zlib

    from generativeaiexamples_b200 import B200FlatVectorStore, B200VectorDBResourceService

    def emb(t):
        v = np.random.default_rng(zlib.crc32(t.encode())).standard_normal(96)
        return (v / np.linalg.norm(v)).tolist()

    class E:
        def embed_documents(self, ts):
            return [emb(t) for t in ts]

        def embed_query(self, t):
            return emb(t)

    vs = B200FlatVectorStore.from_texts([f"doc {i}" for i in range(500)], E(), metadatas=[{"i": i} for i in range(500)])
    vs.save_local(str(tmp_path), index_name="code")
    svc = B200VectorDBResourceService.load(str(tmp_path), E(), name="code")
    assert svc.ntotal == 500
    res = asyncio.run(svc.similarity_search([emb(f"doc {i}") for i in (5, 250, 499)] + [emb("nothing like it")], k=4))
    assert [r[0]["page_content"] for r in res[:3]] == ["doc 5", "doc 250", "doc 499"]
    assert res[1][0]["metadata"] == {"i": 250} and all(len(r) == 4 for r in res)
    single = [d.page_content for d in vs.similarity_search_by_vector(emb("nothing like it"), k=4)]
    assert [d["page_content"] for d in res[3]] == single


def test_async_searches_on_different_streams(built_lib):
    """Device-tensor searches return before the kernels run; issued on several streams they share the index's
    workspace (query slots, candidate scratch, tickets), so the library must order them."""
    import torch

    d, n = 512, 300_000
    ix = _mk("ip", d, reserve_rows=n)
    x = torch.randn((n, d), device="cuda")
    x /= x.norm(dim=1, keepdim=True)
    ix.add(x)
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream() for _ in range(4)]
    outs = []
    for rep in range(24):
        s = streams[rep % 4]
        nq = (1, 3, 4, 16)[rep % 4]  # scan and GEMM routes interleaved
        with torch.cuda.stream(s):
            rows = torch.arange(rep * 100, rep * 100 + nq, device="cuda")
            q = x[rows] + 0.01 * torch.randn((nq, d), device="cuda")
            D, I = ix.search(q, 5)
            outs.append((rows, I, D))
    torch.cuda.synchronize()
    for rows, I, D in outs:
        assert (I[:, 0] == rows).all(), (rows.tolist(), I[:, 0].tolist())
        assert (D[:, 0] > 0.9).all()
    ix.close()


def test_many_queries_and_many_flagged(built_lib):
    """Thousands of queries in one call (many 256-query blocks), hundreds of which fail the containment check:
    the device-side fallback loops over all of them in one cooperative launch."""
    from generativeaiexamples_b200 import B200IndexFlatIP, _lib

    rng = np.random.default_rng(97)
    n, d = 30000, 256
    x = rng.standard_normal((n, d), dtype=np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    centre = x[77].copy()
    x[5000:5800] = centre + 1e-4 * rng.standard_normal((800, d), dtype=np.float32)  # 800 near-duplicates
    nq = 3001
    q = rng.standard_normal((nq, d), dtype=np.float32)
    q[::5] = centre + 1e-3 * rng.standard_normal((len(q[::5]), d), dtype=np.float32)  # 601 queries that must fall back
    ix = B200IndexFlatIP(d)
    ix.add(x)
    f0 = ix.stats()["n_gemm_fallbacks"]
    D, I = ix.search(q, 10)
    st = ix.stats()
    assert st["last_path"] == _lib.PATH_GEMM
    assert st["n_gemm_fallbacks"] - f0 >= 500
    _check("ip", D, I, q, x, 10, "3001 queries, ~600 flagged")
    # and again with device tensors, twice in a row without synchronising in between
    import torch

    qd = torch.from_numpy(q).cuda()
    D1, I1 = ix.search(qd, 10)
    D2, I2 = ix.search(qd[:1500], 10)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(I1.cpu().numpy(), I)
    np.testing.assert_array_equal(I2.cpu().numpy(), I[:1500])
    np.testing.assert_array_equal(D1.cpu().numpy(), D)
    ix.close()


@pytest.mark.parametrize("metric", ["ip", "l2"])
@pytest.mark.parametrize("d", [4096, 8192, 16384])
def test_gemm_parity_wide_rows(built_lib, metric, d):
    """The tensor-core route on the widest rows the index accepts (a row may be up to 64 KB): the containment
    bound's accumulation slack grows with the K depth, so the filter + exact re-score is checked against the
    float64 truth at d = 4096, 8192 and 16384, for fp32 rows, bf16 rows and the bf16 shadow of fp32 rows."""
    rng = np.random.default_rng(d + (1 if metric == "l2" else 0))
    n, nq, k = 5000, 40, 10
    x = rng.standard_normal((n, d), dtype=np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    q = rng.standard_normal((nq, d), dtype=np.float32)
    q[::2] = x[rng.integers(0, n, nq // 2)] + 0.3 * q[::2] / np.sqrt(d)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    for kw in (dict(), dict(shadow_bf16=True), dict(storage="bf16")):
        if kw.get("storage") == "bf16" and d * 2 > 64 * 1024:
            continue
        ix = _mk(metric, d, search_path=_lib.PATH_GEMM, **kw)
        ix.add(x)
        D, I = ix.search(q, k)
        st = ix.stats()
        assert st["last_path"] == _lib.PATH_GEMM
        if kw.get("storage") == "bf16":
            import torch

            xr = torch.from_numpy(x).to(torch.bfloat16).to(torch.float32).numpy()  # exact vs the rounded corpus
            _check(metric, D, I, q, xr, k, f"wide rows d={d} bf16 storage")
        else:
            _check(metric, D, I, q, x, k, f"wide rows d={d} {kw}")
            D1, I1 = _mk_scan_result(metric, d, x, q[:3], k)
            np.testing.assert_array_equal(I[:3], I1)
            np.testing.assert_array_equal(D[:3], D1)  # both routes end in the same fp32 scoring code
        ix.close()


def _mk_scan_result(metric, d, x, q, k):
    ix = _mk(metric, d, search_path=_lib.PATH_SCAN)
    ix.add(x)
    out = ix.search(q, k)
    ix.close()
    return out


def test_route_query_and_exchange_status(built_lib):
    """The two small helpers of the one-process-per-GPU form: b200r_auto_route answers for a shard of a given size
    (every rank asks with the largest shard, so all of them pick the same exchange), b200r_exchange_status reads
    and clears the "a peer did not deliver" word of an exchange buffer."""
    import torch

    ix = _mk("ip", 64)
    ix.add(np.random.default_rng(0).standard_normal((1000, 64), dtype=np.float32))
    route = ctypes.c_int32(-1)

    def ask(nq, rows):
        _lib.check(built_lib.b200r_auto_route(ix._h, nq, rows, ctypes.byref(route)))
        return route.value

    assert ask(1, -1) == _lib.PATH_SCAN and ask(4, -1) == _lib.PATH_SCAN and ask(5, -1) == _lib.PATH_GEMM
    assert ask(3, 100_000_000) == _lib.PATH_GEMM and ask(1, 100_000_000) == _lib.PATH_SCAN
    assert ask(64, 0) == _lib.PATH_SCAN  # an empty shard has nothing to multiply
    nbytes = int(built_lib.b200r_exchange_buffer_bytes())
    buf = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    status = ctypes.c_int32(-1)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(built_lib.b200r_exchange_status(ctypes.c_void_p(buf.data_ptr()), 0, stream, ctypes.byref(status)))
    assert status.value == 0
    buf[-256:-252] = torch.tensor([3, 0, 0, 0], dtype=torch.uint8, device="cuda")  # "rank 2 never arrived"
    _lib.check(built_lib.b200r_exchange_status(ctypes.c_void_p(buf.data_ptr()), 0, stream, ctypes.byref(status)))
    assert status.value == 3
    _lib.check(built_lib.b200r_exchange_status(ctypes.c_void_p(buf.data_ptr()), 0, stream, ctypes.byref(status)))
    assert status.value == 0  # cleared by the previous call
    ix.close()


def test_scan_tile_assignment_modes_agree(built_lib, monkeypatch):
    """K2 deals its tiles statically (B200R_SCAN_DYNAMIC=0), from a global counter (1), or statically with a dynamic
    tail (2, the default on corpora of >= 32 rounds of tiles): which CTA scores a row must never show in the answer.
    350k x 256 fp32 rows = 5469 tiles of 64 rows over 148 CTAs, so mode 2 really draws its last rounds dynamically."""
    import torch

    from generativeaiexamples_b200 import _lib

    dev = torch.device("cuda", 0)
    n, d = 350_000, 256
    g = torch.Generator(device=dev)
    g.manual_seed(99)
    x = torch.randn((n, d), generator=g, device=dev)
    x[1000:1040] = x[7]          # exact duplicates: ties must go to the lowest id in every mode
    q = torch.randn((4, d), generator=g, device=dev)
    q[0] = x[7]
    xh, qh = x.cpu().numpy(), q.cpu().numpy()
    for metric in ("ip", "l2"):
        ix = _mk(metric, d, search_path=_lib.PATH_SCAN, reserve_rows=n)
        ix.add(x)
        seen = {}
        for mode in ("0", "1", "2", "0"):
            monkeypatch.setenv("B200R_SCAN_DYNAMIC", mode)
            _lib.reload_env()
            for nq, k in ((1, 10), (1, 100), (4, 10), (3, 50), (1, 300)):
                D, I = ix.search(q[:nq], k)
                D, I = D.cpu().numpy(), I.cpu().numpy()
                if (nq, k) in seen:
                    np.testing.assert_array_equal(I, seen[(nq, k)][1], err_msg=f"{metric} mode {mode} nq={nq} k={k}")
                    np.testing.assert_array_equal(D, seen[(nq, k)][0], err_msg=f"{metric} mode {mode} nq={nq} k={k}")
                else:
                    seen[(nq, k)] = (D, I)
                    _check(metric, D, I, qh[:nq], xh, k, f"tile modes {metric} nq={nq} k={k}")
            for rep in range(5):     # back to back: the counter is reset by every launch's final writer
                D2, I2 = ix.search(q[:1], 10)
                np.testing.assert_array_equal(I2.cpu().numpy(), seen[(1, 10)][1])
        assert seen[(1, 10)][1][0, 0] == 7   # the duplicate with the lowest id
        ix.close()
    monkeypatch.delenv("B200R_SCAN_DYNAMIC", raising=False)
    _lib.reload_env()
