"""CPU suite, part 1: the oracle itself.

The reference holds no golden vector or test for this path (SURVEY.md sections 4, 8c), so the
oracle is pinned from two independent sides: the hand-computed known answers in
tests/golden/kat_d4.json (exact rational arithmetic) and the float64 brute force.
"""
import json
import math
import os
import warnings

import numpy as np
import pytest

from oracle import FlatRef, IndexFlatRef, assert_topk_parity, truth_topk
from oracle.f64_truth import FLT_MAX
from oracle.flat_ref import METRIC_INNER_PRODUCT, METRIC_L2, load_flat_ref
from oracle.lc_faiss_shim import (COSINE, EUCLIDEAN_DISTANCE, MAX_INNER_PRODUCT, RefDocstore, RefDocument, RefFAISS)

HERE = os.path.dirname(os.path.abspath(__file__))
KAT = json.load(open(os.path.join(HERE, "golden", "kat_d4.json")))
M = {"ip": METRIC_INNER_PRODUCT, "l2": METRIC_L2}


def _expect(case):
    pad = FLT_MAX if case["metric"] == "l2" else -FLT_MAX
    return (np.array([[pad if s is None else s for s in case["scores"]]], dtype=np.float64),
            np.array([case["ids"]], dtype=np.int64))


@pytest.mark.parametrize("case", KAT["cases"], ids=[c["name"] for c in KAT["cases"]])
@pytest.mark.parametrize("path", [0, 1, 2])
def test_kat_search(case, path):
    x = np.array(case["corpus"], dtype=np.float32)
    q = np.array([case["query"]], dtype=np.float32)
    D, I = FlatRef.knn(M[case["metric"]], q, x, case["k"], path=path)
    De, Ie = _expect(case)
    if path == 2 and case["metric"] == "l2":
        # BLAS expansion ||q||^2+||x||^2-2qx is exact on these small integers / quarters too
        pass
    np.testing.assert_array_equal(I, Ie)
    np.testing.assert_allclose(D, De, rtol=0, atol=0)


@pytest.mark.parametrize("case", KAT["cases"], ids=[c["name"] for c in KAT["cases"]])
def test_kat_truth(case):
    x = np.array(case["corpus"], dtype=np.float32)
    q = np.array([case["query"]], dtype=np.float32)
    D, I = truth_topk(case["metric"], q, x, case["k"])
    De, Ie = _expect(case)
    np.testing.assert_array_equal(I, Ie)
    np.testing.assert_allclose(D, De, rtol=0, atol=0)


def test_kat_normalize():
    x = np.array(KAT["normalize"]["rows"], dtype=np.float32)
    FlatRef.normalize_L2(x)
    np.testing.assert_allclose(x, np.array(KAT["normalize"]["expect"], dtype=np.float32), rtol=1e-7, atol=0)


def test_kat_remove_ids():
    r = KAT["remove_ids"]
    idx = IndexFlatRef(4, METRIC_L2)
    idx.add(np.array(r["rows"], dtype=np.float32))
    assert idx.remove_ids(np.array(r["ids"], dtype=np.int64)) == r["n_removed"]
    assert idx.ntotal == len(r["expect"])
    for i, row in enumerate(r["expect"]):
        np.testing.assert_array_equal(idx.reconstruct(i), np.array(row, dtype=np.float32))


def test_kat_relevance():
    rel = KAT["relevance"]
    for row in rel["table"]:
        d = row["distance"]
        assert RefFAISS._euclidean_relevance_score_fn(d) == pytest.approx(row["euclidean"], abs=1e-15)
        assert RefFAISS._cosine_relevance_score_fn(d) == pytest.approx(row["cosine"], abs=1e-15)
        assert RefFAISS._max_inner_product_relevance_score_fn(d) == pytest.approx(row["max_inner_product"], abs=1e-15)
    # squared L2 between unit vectors: relevance = 1 - sqrt(2)*(1-cos); thr 0.25 <=> cos >= 0.4697
    assert 1.0 - (2 - 2 * 0.4697) / math.sqrt(2) == pytest.approx(0.25, abs=1e-4)


@pytest.mark.parametrize("metric", ["ip", "l2"])
@pytest.mark.parametrize("nq", [1, 7, 19, 20, 64])
def test_restatement_matches_f64_truth(metric, nq):
    """T1 (both faiss code paths: nq < 20 sequential, nq >= 20 BLAS expansion) against T0."""
    rng = np.random.default_rng(100 + nq)
    x = rng.standard_normal((6000, 96), dtype=np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    q = rng.standard_normal((nq, 96), dtype=np.float32)
    q[::2] = x[rng.integers(0, 6000, len(q[::2]))] + 0.05 * q[::2]
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    for k in (1, 10, 100):
        D, I = FlatRef.knn(M[metric], q, x, k)
        Dt, It = truth_topk(metric, q, x, k + 8)
        assert_topk_parity(D, I, Dt, It, what=f"{metric} nq={nq} k={k}")
        D2, I2 = FlatRef.knn_allcores(M[metric], q, x, k)
        assert_topk_parity(D2, I2, Dt, It, what=f"allcores {metric} nq={nq} k={k}")


def test_golden_fixtures_reproduce():
    """The committed seeded fixtures are what the oracle computes today."""
    from tests.golden.make_fixtures import CASES, make_case

    gold = np.load(os.path.join(HERE, "golden", "seeded_topk.npz"))
    for name in ("odd_dim_ip", "batch17_l2"):  # the small ones; the rest run in the gpu suite
        n, d, nq, k, metric, normalized, seed = CASES[name]
        x, q = make_case(*CASES[name])
        D, I = FlatRef.knn(M[metric], q, x, k)
        assert_topk_parity(D, I, gold[name + "__D"], gold[name + "__I"], what=name)


def test_empty_and_ragged():
    idx = IndexFlatRef(8, METRIC_L2)
    D, I = idx.search(np.zeros((2, 8), dtype=np.float32), 3)
    assert (I == -1).all() and (D == np.float32(FLT_MAX)).all()
    idx.add(np.ones((1, 8), dtype=np.float32))
    D, I = idx.search(np.zeros((1, 8), dtype=np.float32), 3)
    assert I.tolist() == [[0, -1, -1]] and D[0, 0] == 8.0
    with pytest.raises(AssertionError):
        idx.search(np.zeros((1, 7), dtype=np.float32), 3)
    with pytest.raises(AssertionError):
        idx.search(np.zeros((1, 8), dtype=np.float32), 0)
    with pytest.raises(AssertionError):
        idx.add(np.zeros((1, 9), dtype=np.float32))


class _Emb:
    """Deterministic stand-in for the embedding NIM: text -> seeded unit vector."""

    def __init__(self, d=64):
        self.d = d

    def _v(self, t):
        import zlib

        rng = np.random.default_rng(zlib.crc32(t.encode()))
        v = rng.standard_normal(self.d)
        return (v / np.linalg.norm(v)).tolist()

    def embed_documents(self, texts):
        return [self._v(t) for t in texts]

    def embed_query(self, t):
        return self._v(t)


def test_shim_threshold_and_delete():
    emb = _Emb()
    vs = RefFAISS(emb, IndexFlatRef(64, METRIC_L2), RefDocstore(), {})
    ids = vs.add_texts([f"chunk {i}" for i in range(50)], metadatas=[{"source": f"f{i % 5}.txt"} for i in range(50)])
    assert len(ids) == 50 and vs.index.ntotal == 50
    # identical text -> distance 0 -> relevance 1.0, everything else is far below 0.25
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        docs = vs.retrieve("chunk 7", "similarity_score_threshold", {"score_threshold": 0.25, "k": 4})
    assert [d.page_content for d in docs] == ["chunk 7"]
    with pytest.raises(ValueError):
        vs.retrieve("chunk 7", "similarity_score_threshold", {"score_threshold": 1, "k": 4})  # int, not float
    assert len(vs.retrieve("chunk 7", "similarity", {})) == 4
    # delete shifts positions
    vs.delete([ids[7], ids[0]])
    assert vs.index.ntotal == 48 and vs.index_to_docstore_id[0] == ids[1] and vs.index_to_docstore_id[6] == ids[8]
    with pytest.raises(ValueError):
        vs.delete(["nope"])
    assert isinstance(vs.docstore.search(ids[1]), RefDocument)


def test_oracle_build_is_loadable():
    lib = load_flat_ref()
    assert lib.oracle_max_threads() >= 1


@pytest.mark.parametrize("nq", [3, 40])
def test_restatement_matches_independent_bruteforce(nq):
    """Third-party cross-check (not the reference: faiss is absent from this image): scikit-learn's brute-force
    k-NN must agree with the restatement on ids, and on distances after squaring (faiss returns squared L2)."""
    sk = pytest.importorskip("sklearn.neighbors")
    rng = np.random.default_rng(2024 + nq)
    x = rng.standard_normal((3000, 96)).astype(np.float32)
    q = rng.standard_normal((nq, 96)).astype(np.float32)
    nn = sk.NearestNeighbors(n_neighbors=12, algorithm="brute", metric="euclidean").fit(x.astype(np.float64))
    dist, ids = nn.kneighbors(q.astype(np.float64))
    D, I = FlatRef.knn(METRIC_L2, q, x, 12)
    scale = float((q.astype(np.float64) ** 2).sum(1).max() + (x.astype(np.float64) ** 2).sum(1).max()) if nq >= 20 else 1.0
    assert_topk_parity(D, I, (dist ** 2).astype(np.float32), ids.astype(np.int64), what="restatement vs sklearn brute force",
                       scale=scale)


def test_merge_from_restatement():
    """faiss IndexFlatCodes::merge_from + LangChain FAISS.merge_from (vectorstore_updater.py:217)."""
    rng = np.random.default_rng(12)
    a, b = IndexFlatRef(8, METRIC_L2), IndexFlatRef(8, METRIC_L2)
    xa, xb = rng.standard_normal((5, 8)).astype(np.float32), rng.standard_normal((3, 8)).astype(np.float32)
    a.add(xa)
    b.add(xb)
    a.merge_from(b)
    assert a.ntotal == 8 and b.ntotal == 0
    np.testing.assert_array_equal(a.reconstruct(6), xb[1])
    with pytest.raises(RuntimeError):
        a.merge_from(IndexFlatRef(8, METRIC_INNER_PRODUCT))
    with pytest.raises(RuntimeError):
        a.merge_from(IndexFlatRef(9, METRIC_L2))

    def emb(t):
        return np.random.default_rng(zlib_crc(t)).standard_normal(8).astype(np.float32).tolist()

    import zlib

    def zlib_crc(t):
        return zlib.crc32(t.encode())

    s1 = RefFAISS(emb, IndexFlatRef(8), RefDocstore(), {})
    s2 = RefFAISS(emb, IndexFlatRef(8), RefDocstore(), {})
    s1.add_texts(["a", "b", "c"])
    s2.add_texts(["d", "e"])
    s1.merge_from(s2)
    assert s1.index.ntotal == 5 and sorted(s1.index_to_docstore_id) == [0, 1, 2, 3, 4]
    assert s1.similarity_search("e", k=1)[0].page_content == "e"
    assert s1.similarity_search("a", k=1)[0].page_content == "a"


def test_restatement_against_a_real_faiss_when_one_is_installed():
    """The pin this oracle is waiting for: faiss-cpu is the reference's engine, but there is
    no wheel for it in this image, so this test skips here.  On a machine that has faiss (any 1.7.x / 1.8.x), the
    restatement must return faiss's own ids and scores on both code paths (nq < 20 sequential, nq >= 20 BLAS),
    both metrics, k > ntotal padding, normalize_L2 and remove_ids."""
    import os
    import sys

    ref_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref")
    if os.path.isdir(ref_dir) and ref_dir not in sys.path:
        sys.path.append(ref_dir)
    faiss = pytest.importorskip("faiss")
    from oracle import FlatRef, IndexFlatRef, assert_topk_parity

    rng = np.random.default_rng(2024)
    d, n = 96, 3000
    x = rng.standard_normal((n, d), dtype=np.float32)
    for metric, ctor in ((1, faiss.IndexFlatL2), (0, faiss.IndexFlatIP)):
        fx = ctor(d)
        fx.add(x)
        for nq in (1, 7, 19, 20, 64):
            q = rng.standard_normal((nq, d), dtype=np.float32)
            q[::2] = x[rng.integers(0, n, (nq + 1) // 2)] + 0.05 * q[::2]
            for k in (1, 10, 100):
                Df, If = fx.search(q, k)
                Dr, Ir = FlatRef.knn(metric, q, x, k)
                scale = float((q ** 2).sum(1).max() + (x ** 2).sum(1).max()) if (metric == 1 and nq >= 20) else 1.0
                assert_topk_parity(Dr, Ir, Df, If, what=f"faiss m={metric} nq={nq} k={k}", scale=scale)
        Df, If = fx.search(x[:2], n + 5)  # k > ntotal: -1 padding, +-FLT_MAX
        Dr, Ir = FlatRef.knn(metric, x[:2], x, n + 5)
        assert (If[:, n:] == -1).all() and (Ir[:, n:] == -1).all()
        np.testing.assert_array_equal(Df[:, n:], Dr[:, n:])
        ours = IndexFlatRef(d, metric)
        ours.add(x)
        ids = np.array([5, 17, 17, 2999, 4000], dtype=np.int64)
        assert fx.remove_ids(ids) == ours.remove_ids(ids)
        np.testing.assert_array_equal(fx.reconstruct(16), ours.reconstruct(16))
    y = x[:50].copy()
    y[3] = 0
    z = y.copy()
    faiss.normalize_L2(y)
    FlatRef.normalize_L2(z)
    np.testing.assert_allclose(y, z, rtol=1e-6, atol=0)
