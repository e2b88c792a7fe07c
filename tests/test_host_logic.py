"""CPU suite, part 3: the host-side mirror of the LangChain surface, with the CPU oracle
injected as the index object (the product never does that itself).  The same call sequence
runs through oracle.lc_faiss_shim.RefFAISS and B200FlatVectorStore and must agree."""
import asyncio
import os
import warnings
import zlib

import numpy as np
import pytest

from generativeaiexamples_b200.factory import del_docs_vectorstore_langchain, get_docs_vectorstore_langchain
from generativeaiexamples_b200.vectorstore import (B200FlatVectorStore, DistanceStrategy, Document, InMemoryDocstore,
                                                   VectorStoreRetriever)
from oracle.flat_ref import METRIC_INNER_PRODUCT, METRIC_L2, IndexFlatRef
from oracle.lc_faiss_shim import RefDocstore, RefFAISS

D = 64


class Emb:
    def _v(self, t):
        rng = np.random.default_rng(zlib.crc32(t.encode()))
        v = rng.standard_normal(D)
        return (v / np.linalg.norm(v)).tolist()

    def embed_documents(self, texts):
        return [self._v(t) for t in texts]

    def embed_query(self, t):
        return self._v(t)


def _pair(metric=METRIC_L2, strategy="EUCLIDEAN_DISTANCE", **kw):
    ours = B200FlatVectorStore(Emb(), IndexFlatRef(D, metric), InMemoryDocstore(), {},
                               distance_strategy=DistanceStrategy(strategy), **kw)
    ref = RefFAISS(Emb(), IndexFlatRef(D, metric), RefDocstore(), {}, distance_strategy=strategy, **kw)
    return ours, ref


def _ingest(ours, ref, n=60):
    texts = [f"chunk {i}" for i in range(n)]
    metas = [{"source": f"/tmp-data/uploaded_files/doc{i % 4}.pdf"} for i in range(n)]
    ids = [f"id-{i}" for i in range(n)]
    assert ours.add_texts(texts, metadatas=metas, ids=list(ids)) == ids
    assert ref.add_texts(texts, metadatas=metas, ids=list(ids)) == ids
    return texts, ids


def _same(a, b):
    assert [(d.page_content, d.metadata) for d in a] == [(d.page_content, d.metadata) for d in b]


def test_basic_rag_call_sequence():
    """chains.py:83 (add_documents) then chains.py:156-163 (threshold retriever)."""
    ours, ref = _pair()
    docs = [Document(page_content=f"chunk {i}", metadata={"source": f"/x/doc{i % 3}.txt"}) for i in range(40)]
    ours.add_documents(docs)
    ref.add_texts([d.page_content for d in docs], [d.metadata for d in docs])
    assert ours.index.ntotal == 40 and len(ours.docstore._dict) == 40
    seen = []

    class CB:
        def on_retriever_start(self, serialized, query, **kw):
            seen.append(("start", query))

        def on_retriever_end(self, documents, **kw):
            seen.append(("end", len(documents)))

    retriever = ours.as_retriever(search_type="similarity_score_threshold",
                                  search_kwargs={"score_threshold": 0.25, "k": 4})
    assert isinstance(retriever, VectorStoreRetriever)
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        got = retriever.get_relevant_documents("chunk 5", callbacks=[CB()])
        want = ref.retrieve("chunk 5", "similarity_score_threshold", {"score_threshold": 0.25, "k": 4})
    _same(got, want)
    assert [d.page_content for d in got] == ["chunk 5"]
    assert seen == [("start", "chunk 5"), ("end", 1)]
    # fallback branch chains.py:166
    _same(ours.as_retriever().get_relevant_documents("chunk 5"), ref.retrieve("chunk 5"))
    _same(ours.as_retriever().invoke("chunk 9", config={"callbacks": []}), ref.retrieve("chunk 9"))


@pytest.mark.parametrize("metric,strategy", [(METRIC_L2, "EUCLIDEAN_DISTANCE"),
                                             (METRIC_INNER_PRODUCT, "MAX_INNER_PRODUCT")])
def test_scores_filters_thresholds(metric, strategy):
    ours, ref = _pair(metric, strategy)
    _ingest(ours, ref)
    for k in (1, 4, 25, 100):  # 100 > ntotal: -1 padding is skipped
        a = ours.similarity_search_with_score("chunk 3", k=k)
        b = ref.similarity_search_with_score("chunk 3", k=k)
        _same([d for d, _ in a], [d for d, _ in b])
        assert [float(s) for _, s in a] == [float(s) for _, s in b]
        assert all(isinstance(s, np.float32) for _, s in a)
    flt = {"source": "/tmp-data/uploaded_files/doc1.pdf"}
    _same(ours.similarity_search("chunk 3", k=3, filter=flt, fetch_k=30),
          ref.similarity_search("chunk 3", k=3, filter=flt, fetch_k=30))
    # a callable over the metadata (langchain-community >= 0.0.21) selects the same documents as the dict form
    _same(ours.similarity_search("chunk 3", k=3, filter=lambda md: md.get("source") == flt["source"], fetch_k=30),
          ref.similarity_search("chunk 3", k=3, filter=flt, fetch_k=30))
    thr = 1.0 if metric == METRIC_L2 else 0.2
    a = ours.similarity_search_with_score_by_vector(Emb()._v("chunk 3"), k=10, score_threshold=thr)
    b = ref.similarity_search_with_score_by_vector(Emb()._v("chunk 3"), k=10, score_threshold=thr)
    _same([d for d, _ in a], [d for d, _ in b])
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        a = ours.similarity_search_with_relevance_scores("chunk 3", k=4, score_threshold=0.25)
        b = ref.similarity_search_with_relevance_scores("chunk 3", k=4, score_threshold=0.25)
    assert [(d.page_content, float(s)) for d, s in a] == [(d.page_content, float(s)) for d, s in b]
    if metric == METRIC_L2:  # orthogonal-ish unit vectors: d ~ 2 -> relevance 1 - sqrt(2) < 0
        assert any("Relevance scores must be between 0 and 1" in str(x.message) for x in w)


def test_threshold_boundary_is_inclusive():
    ours, _ = _pair(relevance_score_fn=lambda d: 0.25)
    ours.add_texts(["a", "b"])
    assert len(ours.similarity_search_with_relevance_scores("a", k=2, score_threshold=0.25)) == 2
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        assert ours.similarity_search_with_relevance_scores("a", k=2, score_threshold=0.2500001) == []
    assert any("No relevant docs were retrieved" in str(x.message) for x in w)


def test_retriever_validation():
    ours, _ = _pair()
    with pytest.raises(ValueError):
        ours.as_retriever(search_type="similarity_score_threshold", search_kwargs={"k": 4})
    with pytest.raises(ValueError):
        ours.as_retriever(search_type="similarity_score_threshold", search_kwargs={"score_threshold": 1, "k": 4})
    with pytest.raises(ValueError):
        ours.as_retriever(search_type="nearest")


def test_delete_and_docstore_helpers():
    """utils.py:506-510 (list) and :546-560 (delete by file name)."""
    ours, ref = _pair()
    texts, ids = _ingest(ours, ref)
    assert sorted(get_docs_vectorstore_langchain(ours)) == [f"doc{i}.pdf" for i in range(4)]
    assert del_docs_vectorstore_langchain(ours, ["doc2.pdf"]) is True
    ref.delete([i for n, i in enumerate(ids) if n % 4 == 2])
    assert ours.index.ntotal == ref.index.ntotal == 45
    assert ours.index_to_docstore_id == ref.index_to_docstore_id
    assert del_docs_vectorstore_langchain(ours, ["missing.pdf"]) is False
    _same(ours.similarity_search("chunk 7", k=5), ref.similarity_search("chunk 7", k=5))
    with pytest.raises(ValueError):
        ours.delete(["not-there"])
    with pytest.raises(ValueError):
        ours.delete(None)
    with pytest.raises(ValueError):  # duplicate id on add
        ours.add_texts(["again"], ids=[ids[0]])


def test_add_embeddings_and_async():
    ours, ref = _pair()
    e = Emb()
    pairs = [(f"t{i}", e._v(f"t{i}")) for i in range(20)]
    ours.add_embeddings(pairs, ids=[f"i{i}" for i in range(20)])
    ref.add_embeddings(pairs, ids=[f"i{i}" for i in range(20)])
    v = e._v("t4")
    _same(ours.similarity_search_by_vector(v, k=3), ref.similarity_search_by_vector(v, k=3))

    async def go():  # faiss_vdb_service.py:169-174: gather of per-query async searches
        return await asyncio.gather(*[ours.asimilarity_search_by_vector(e._v(f"t{i}"), k=2) for i in range(5)])

    res = asyncio.run(go())
    assert [r[0].page_content for r in res] == [f"t{i}" for i in range(5)]
    Dm, Im = ours.search_by_vectors(np.array([e._v(f"t{i}") for i in range(5)], dtype=np.float32), k=2)
    assert Im[:, 0].tolist() == list(range(5))


def test_empty_store_and_bad_embedding_function(caplog):
    vs = B200FlatVectorStore(lambda t: [0.0] * D, None)
    assert vs.similarity_search_with_score_by_vector([0.0] * D) == []
    assert vs.max_marginal_relevance_search("q") == []


def test_mmr_matches_restatement():
    """search_type="mmr" (langchain allowed_search_types): engine fetches, LangChain's greedy
    diversity selection runs on the <= fetch_k reconstructed vectors."""
    ours, ref = _pair()
    rng = np.random.default_rng(3)
    base = rng.standard_normal((8, D))
    texts, embs = [], []
    for c in range(8):  # 8 clusters of 10 near-duplicates: MMR must spread over clusters
        for j in range(10):
            v = base[c] + 0.05 * rng.standard_normal(D)
            texts.append(f"c{c}-{j}")
            embs.append((v / np.linalg.norm(v)).astype(np.float32).tolist())
    ours.add_embeddings(list(zip(texts, embs)), ids=list(texts))
    ref.add_embeddings(list(zip(texts, embs)), ids=list(texts))
    qv = (base[2] + 0.3 * base[5]).astype(np.float32)
    qv = (qv / np.linalg.norm(qv)).tolist()
    for lam in (0.5, 0.1, 0.9):
        got = ours.max_marginal_relevance_search_by_vector(qv, k=5, fetch_k=30, lambda_mult=lam)
        want = ref.max_marginal_relevance_search_by_vector(qv, k=5, fetch_k=30, lambda_mult=lam)
        assert [d.page_content for d in got] == [d.page_content for d in want], lam
    plain = [d.page_content for d in ours.similarity_search_by_vector(qv, k=5)]
    mmr = [d.page_content for d in ours.max_marginal_relevance_search_by_vector(qv, k=5, fetch_k=30, lambda_mult=0.1)]
    assert len({t.split("-")[0] for t in mmr}) > len({t.split("-")[0] for t in plain})
    r = ours.as_retriever(search_type="mmr", search_kwargs={"k": 3, "fetch_k": 20})
    assert len(r.get_relevant_documents("c2-0")) == 3


def test_coalescer_batches_concurrent_callers():
    """f2: B threads with one query each (faiss_vdb_service.py:169-174) end up in few launches."""
    import threading
    import time

    from generativeaiexamples_b200.batching import SearchCoalescer

    class SlowIndex:
        def __init__(self):
            self.ix = IndexFlatRef(D, METRIC_L2)
            self.calls = []

        def search(self, x, k):
            self.calls.append(x.shape[0])
            time.sleep(0.02)  # a search in flight: later arrivals must queue up behind it
            return self.ix.search(x, k)

    rng = np.random.default_rng(1)
    x = rng.standard_normal((500, D)).astype(np.float32)
    slow = SlowIndex()
    slow.ix.add(x)
    co = SearchCoalescer(slow)
    res = [None] * 32

    def work(i):
        res[i] = co.search(x[i:i + 1], 3 if i % 8 else 5)  # two different k values in the mix

    th = [threading.Thread(target=work, args=(i,)) for i in range(32)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert all(r[1][0, 0] == i for i, r in enumerate(res))
    assert all(r[0].shape == (1, 3 if i % 8 else 5) for i, r in enumerate(res))
    assert sum(slow.calls) == 32 and len(slow.calls) < 16 and max(slow.calls) > 4
    assert co.n_requests == 32 and co.n_launches == len(slow.calls)


def test_coalescer_hands_leadership_over_under_sustained_load():
    """A caller that found the coalescer idle serves batches only until its OWN request is answered, then a
    waiting caller takes over: under a steady stream of requests nobody is kept serving strangers."""
    import threading
    import time

    from generativeaiexamples_b200.batching import SearchCoalescer

    class SlowIndex:
        def __init__(self):
            self.ix = IndexFlatRef(D, METRIC_L2)

        def search(self, x, k):
            time.sleep(0.005)
            return self.ix.search(x, k)

    rng = np.random.default_rng(2)
    x = rng.standard_normal((300, D)).astype(np.float32)
    slow = SlowIndex()
    slow.ix.add(x)
    co = SearchCoalescer(slow)
    stop = time.perf_counter() + 0.6
    worst = [0.0] * 8
    served = [0] * 8

    def work(t):
        i = t
        while time.perf_counter() < stop:
            t0 = time.perf_counter()
            _, I = co.search(x[i % 300: i % 300 + 1], 2)
            worst[t] = max(worst[t], time.perf_counter() - t0)
            assert I[0, 0] == i % 300
            served[t] += 1
            i += 8

    th = [threading.Thread(target=work, args=(t,)) for t in range(8)]
    for t in th:
        t.start()
    for t in th:
        t.join(timeout=10)
    assert not any(t.is_alive() for t in th)
    assert co.n_handoffs > 0 and min(served) > 5
    # the queue never ran dry for 0.6 s, yet no single request took anywhere near that long
    assert max(worst) < 0.2, worst
    assert not co._busy and not co._pending


def test_store_with_coalescing_and_async_gather():
    ours = B200FlatVectorStore(Emb(), IndexFlatRef(D, METRIC_L2), InMemoryDocstore(), {}, coalesce=True)
    ours.add_texts([f"t{i}" for i in range(50)])
    e = Emb()

    async def go():
        return await asyncio.gather(*[ours.asimilarity_search_by_vector(e._v(f"t{i}"), k=2) for i in range(20)])

    res = asyncio.run(go())
    assert [r[0].page_content for r in res] == [f"t{i}" for i in range(20)]
    assert ours._coalescer.n_requests == 20


def test_multi_turn_rag_surface():
    """f1: multi_turn_rag keeps two stores (documents + conversation), adds a turn at a time with
    add_texts and fetches k=40 candidates for the reranker
    (RAG/examples/advanced_rag/multi_turn_rag/chains.py:63-68,138,147-172)."""
    docstore, _ = _pair()
    conv_store, ref_conv = _pair()
    docs = [f"doc chunk {i}" for i in range(100)]
    docstore.add_texts(docs, metadatas=[{"source": "a.pdf"}] * 100)
    for turn in range(6):  # one small add per turn
        t = [f"user said {turn}", f"assistant said {turn}"]
        conv_store.add_texts(t)
        ref_conv.add_texts(t)
    r = docstore.as_retriever(search_kwargs={"k": 40})
    got = r.get_relevant_documents("doc chunk 7")
    assert len(got) == 40 and got[0].page_content == "doc chunk 7"
    _same(conv_store.as_retriever(search_kwargs={"k": 4}).get_relevant_documents("user said 3"),
          ref_conv.retrieve("user said 3", "similarity", {"k": 4}))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        thr = conv_store.as_retriever(search_type="similarity_score_threshold",
                                      search_kwargs={"score_threshold": 0.25, "k": 40})
        assert [d.page_content for d in thr.get_relevant_documents("user said 3")] == ["user said 3"]


def test_to_matrix_fast_paths_match_numpy():
    from generativeaiexamples_b200.vectorstore import _to_matrix

    rng = np.random.default_rng(3)
    rows = (rng.standard_normal((5, 37)) * 1e3).tolist()
    for inp in (rows[0], tuple(rows[0]), rows, [tuple(r) for r in rows], np.asarray(rows), np.asarray(rows, dtype=np.float32)):
        got = _to_matrix(inp)
        want = np.asarray(inp, dtype=np.float32)
        want = want[None, :] if want.ndim == 1 else want
        assert got.dtype == np.float32 and got.flags["C_CONTIGUOUS"] and got.shape == want.shape
        np.testing.assert_array_equal(got, want)
    with pytest.raises(ValueError):
        _to_matrix([[1.0, 2.0], [3.0]])
    np.testing.assert_array_equal(_to_matrix([[1, 2], [3, 4]]), np.array([[1, 2], [3, 4]], dtype=np.float32))


def test_morpheus_service_mirror(tmp_path):
    """faiss_vdb_service.py:147-174: async similarity_search(embeddings, k) -> list[list[Document.dict()]];
    every other operation raises NotImplementedError, as in the reference."""
    from generativeaiexamples_b200 import B200VectorDBResourceService

    ours, ref = _pair()
    e = Emb()
    pairs = [(f"t{i}", e._v(f"t{i}")) for i in range(30)]
    ours.add_embeddings(pairs, metadatas=[{"n": i} for i in range(30)])
    ref.add_embeddings(pairs, metadatas=[{"n": i} for i in range(30)])
    svc = B200VectorDBResourceService(ours)
    embs = [e._v(f"t{i}") for i in (3, 7, 11)] + [e._v("unseen")]
    got = asyncio.run(svc.similarity_search(embs, k=3))
    want = [[{"page_content": d.page_content, "metadata": d.metadata, "type": "Document"}
             for d in ref.similarity_search_by_vector(v, k=3)] for v in embs]
    assert got == want
    assert got[0][0]["page_content"] == "t3" and got[1][0]["metadata"] == {"n": 7}
    assert asyncio.run(svc.similarity_search([], k=3)) == []
    assert svc.ntotal == 30
    for call in (lambda: svc.insert([[0.0]]), lambda: svc.insert_dataframe(None), svc.describe, lambda: svc.query("x"),
                 lambda: svc.update([]), lambda: svc.delete_by_keys(1), lambda: svc.delete("x"),
                 lambda: svc.retrieve_by_keys(1), svc.count, svc.drop):
        with pytest.raises(NotImplementedError, match="not supported in FAISS"):
            call()
    empty = B200VectorDBResourceService(B200FlatVectorStore(Emb(), None))
    assert asyncio.run(empty.similarity_search(embs[:2], k=3)) == [[], []]


def test_exchange_epochs_never_zero_and_alternate():
    """The fused exchange buffers two epochs by parity and treats 0 as "no wait": the uint32 counter must
    skip 0 and keep alternating when it wraps."""
    from generativeaiexamples_b200.sharded import _next_epoch

    assert _next_epoch(0) == 1 and _next_epoch(1) == 2
    e = 0xFFFFFFFE
    seen = []
    for _ in range(4):
        e = _next_epoch(e)
        seen.append(e)
    assert seen == [0xFFFFFFFF, 2, 3, 4]
    assert all(0 < v <= 0xFFFFFFFF for v in seen)
    assert all((a ^ b) & 1 for a, b in zip(seen, seen[1:]))


def test_faiss_compat_exports_the_names_direct_callers_use():
    """rag_service.py:88,194,376,399 -- module-level names only; nothing is created without a GPU."""
    import generativeaiexamples_b200.faiss_compat as faiss

    for name in ("IndexFlatL2", "IndexFlatIP", "IndexFlat", "normalize_L2", "write_index", "read_index"):
        assert callable(getattr(faiss, name)), name
    assert faiss.METRIC_L2 == 1 and faiss.METRIC_INNER_PRODUCT == 0
    with pytest.raises(TypeError):
        faiss.write_index(object(), "/nonexistent")
