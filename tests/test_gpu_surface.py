"""GPU suite: the callers either side of index.search on the real engine,
against the LangChain restatement (oracle/lc_faiss_shim.py) and numpy.

* a4  FAISS.from_embeddings            (industries/.../video_preprocess.py:192: the only in-tree call site)
* f1  multi_turn_rag surface           (RAG/examples/advanced_rag/multi_turn_rag/chains.py:63-68,147-172)
* f4  relevance + score_threshold and the MMR selection computed on the device
"""
import warnings
import zlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

D_EMB = 384  # all-MiniLM-L6-v2 shape, as in video_preprocess.py:76


def _vec(text, d=D_EMB):
    r = np.random.default_rng(zlib.crc32(text.encode()))
    base = r.standard_normal(d)
    if text.startswith("about "):  # a query near its document: cos ~ 0.8
        base = np.asarray(_vec(text[6:], d)) + 0.75 * base / np.linalg.norm(base)
    return (base / np.linalg.norm(base)).astype(np.float32).tolist()


class Emb:
    def embed_documents(self, texts):
        return [_vec(t) for t in texts]

    def embed_query(self, t):
        return _vec(t)


def _ref(strategy="EUCLIDEAN_DISTANCE", normalize=False):
    from oracle.flat_ref import IndexFlatRef
    from oracle.lc_faiss_shim import RefDocstore, RefFAISS

    return RefFAISS(Emb(), IndexFlatRef(D_EMB, 0 if strategy == "MAX_INNER_PRODUCT" else 1), RefDocstore(), {},
                    normalize_L2=normalize, distance_strategy=strategy)


def _same_docs(got, want):
    assert [g.page_content for g in got] == [w.page_content for w in want]


def test_from_embeddings_like_video_preprocess(built_lib):
    """a4: FAISS.from_embeddings(text_embeddings=list(zip(texts, embeddings)), embedding=model, metadatas=...)"""
    from generativeaiexamples_b200 import B200FlatVectorStore
    from oracle.lc_faiss_shim import RefFAISS

    texts = [f"caption of frame {i}" for i in range(500)]
    embs = [_vec(t) for t in texts]
    metas = [{"timestamp": i * 0.5} for i in range(500)]
    ids = [f"f{i}" for i in range(500)]
    vs = B200FlatVectorStore.from_embeddings(text_embeddings=list(zip(texts, embs)), embedding=Emb(), metadatas=metas,
                                             ids=list(ids))
    ref = RefFAISS.from_embeddings(list(zip(texts, embs)), Emb(), metadatas=metas, ids=list(ids))
    assert vs.index.ntotal == 500 and vs.index.d == D_EMB and vs.index_to_docstore_id == ref.index_to_docstore_id
    for q in ("about caption of frame 42", "caption of frame 7", "unrelated"):
        a = vs.similarity_search_with_score(q, k=6)
        b = ref.similarity_search_with_score(q, k=6)
        _same_docs([x for x, _ in a], [x for x, _ in b])
        np.testing.assert_allclose([s for _, s in a], [s for _, s in b], rtol=1e-5, atol=1e-5)
        assert a[0][0].metadata == b[0][0].metadata
    more = vs.add_embeddings([("late frame", _vec("late frame"))], metadatas=[{"timestamp": 999.0}])
    assert len(more) == 1 and vs.index.ntotal == 501
    assert vs.similarity_search("late frame", k=1)[0].metadata == {"timestamp": 999.0}


@pytest.mark.parametrize("devices", [None, [0, 0]], ids=["dev0", "dev00"])
def test_multi_turn_rag_two_live_stores(built_lib, devices, monkeypatch):
    """f1: a document store and a conversation store alive on the GPU at once; one small add_texts per turn on
    the conversation store; k=40 candidates for the reranker, then k=4; the threshold retriever on both."""
    from generativeaiexamples_b200 import create_vectorstore_langchain

    if devices:
        monkeypatch.setenv("B200R_DEVICES", ",".join(map(str, devices)))
        monkeypatch.setenv("B200R_MULTI_MIN_PIECE", "8")
    docstore = create_vectorstore_langchain(Emb(), name="b200flat", dimensions=D_EMB)
    conv_store = create_vectorstore_langchain(Emb(), name="b200flat", dimensions=D_EMB)
    ref_doc, ref_conv = _ref(), _ref()
    docs = [f"doc chunk {i}" for i in range(400)]
    for s in (docstore, ref_doc):
        s.add_texts(docs, [{"source": f"f{i % 5}.pdf"} for i in range(400)])
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for turn in range(12):
            t = [f"user said {turn}", f"assistant said {turn}"]
            conv_store.add_texts(t)  # chains.py:65,86: one add per turn
            ref_conv.add_texts(t)
            assert conv_store.index.ntotal == ref_conv.index.ntotal == 2 * (turn + 1)
            q = f"about user said {turn // 2}"
            _same_docs(conv_store.as_retriever(search_kwargs={"k": 4}).get_relevant_documents(q),
                       ref_conv.retrieve(q, "similarity", {"k": 4}))
            # interleaved with searches on the other store
            q2 = f"about doc chunk {turn * 7}"
            got40 = docstore.as_retriever(search_kwargs={"k": 40}).get_relevant_documents(q2)
            _same_docs(got40, ref_doc.retrieve(q2, "similarity", {"k": 40}))
            assert len(got40) == 40 and got40[0].page_content == f"doc chunk {turn * 7}"
            for store, ref, qq in ((docstore, ref_doc, q2), (conv_store, ref_conv, q)):
                kw = {"score_threshold": 0.25, "k": 40}
                _same_docs(store.as_retriever(search_type="similarity_score_threshold", search_kwargs=kw)
                           .get_relevant_documents(qq), ref.retrieve(qq, "similarity_score_threshold", dict(kw)))


@pytest.mark.parametrize("strategy,normalize", [("EUCLIDEAN_DISTANCE", False), ("EUCLIDEAN_DISTANCE", True),
                                                ("MAX_INNER_PRODUCT", False), ("COSINE", False)])
def test_threshold_on_device_matches_langchain(built_lib, strategy, normalize):
    """f4: relevance scores and `score >= score_threshold` computed by b200r_search_threshold; the LangChain
    surface returns what the host-side restatement returns, warnings included."""
    from generativeaiexamples_b200 import B200FlatVectorStore

    texts = [f"chunk {i}" for i in range(300)]
    vs = B200FlatVectorStore(Emb(), D_EMB, normalize_L2=normalize, distance_strategy=strategy)
    ref = _ref(strategy, normalize)
    for s in (vs, ref):
        s.add_texts(texts)
    calls0 = vs.index.stats()["n_search_calls"]
    for q in ("chunk 5", "about chunk 17", "nothing like it", "about chunk 250"):
        for thr in (None, 0.0, 0.25, 0.6, 0.999):
            for k in (1, 4, 40, 400):  # 400 > ntotal: padded hits are skipped
                kw = {} if thr is None else {"score_threshold": thr}
                with warnings.catch_warnings(record=True) as w_got:
                    warnings.simplefilter("always")
                    got = vs.similarity_search_with_relevance_scores(q, k=k, **kw)
                with warnings.catch_warnings(record=True) as w_want:
                    warnings.simplefilter("always")
                    want = ref.similarity_search_with_relevance_scores(q, k=k, **kw)
                _same_docs([g for g, _ in got], [x for x, _ in want])
                np.testing.assert_allclose([s for _, s in got], [s for _, s in want], rtol=1e-5, atol=1e-5)
                kinds = lambda ws: sorted(str(x.message)[:30] for x in ws)  # noqa: E731
                assert kinds(w_got) == kinds(w_want), (q, thr, k)
    assert vs.index.stats()["n_search_calls"] > calls0
    # raw engine call, a batch: against numpy on the plain search
    x = np.asarray([_vec(t) for t in texts], dtype=np.float32)
    qs = np.asarray([_vec(f"about chunk {i}") for i in range(17)], dtype=np.float32)
    if normalize:
        return
    D, I = vs.index.search(qs, 25)
    kind = strategy
    R, D2, I2, counts, oor = vs.index.search_threshold(qs, 25, kind, 0.3)
    fn = {"EUCLIDEAN_DISTANCE": lambda d_: 1.0 - d_ / np.sqrt(2.0), "COSINE": lambda d_: 1.0 - d_,
          "MAX_INNER_PRODUCT": lambda d_: np.where(d_ > 0, 1.0 - d_, -d_)}[kind]
    rel = fn(D.astype(np.float64))
    for i in range(17):
        keep = rel[i] >= 0.3
        assert counts[i] == keep.sum() and oor[i] == ((rel[i] < 0) | (rel[i] > 1)).sum()
        np.testing.assert_array_equal(I2[i, :counts[i]], I[i][keep])
        np.testing.assert_array_equal(D2[i, :counts[i]], D[i][keep])
        np.testing.assert_array_equal(R[i, :counts[i]], rel[i][keep])
        np.testing.assert_array_equal(I2[i, counts[i]:], I[i][~keep])  # dropped hits follow, in order
    assert x.shape[0] == 300


@pytest.mark.parametrize("devices", [None, [0, 0]], ids=["dev0", "dev00"])
def test_mmr_on_device_matches_langchain(built_lib, devices, monkeypatch):
    """f4: the greedy MMR selection runs on the GPU over the gathered candidate rows; the picks equal
    LangChain's maximal_marginal_relevance on the same candidates (numpy restatement in vectorstore.py), also
    with duplicated rows (ties) and a zero vector (similarity forced to 0)."""
    from generativeaiexamples_b200 import B200FlatVectorStore, B200IndexFlat
    from generativeaiexamples_b200.vectorstore import maximal_marginal_relevance

    monkeypatch.setenv("B200R_MULTI_MIN_PIECE", "16")
    rng = np.random.default_rng(17)
    for d, storage in ((64, "f32"), (1024, "f32"), (200, "bf16")):
        ix = B200IndexFlat(d, 1, storage=storage, devices=devices)
        x = rng.standard_normal((3000, d), dtype=np.float32)
        centre = rng.standard_normal(d).astype(np.float32)
        x[:200] = centre + 0.4 * x[:200]  # a cluster: redundancy matters
        x[50] = x[10]
        x[77] = 0.0
        ix.add(x)
        stored = ix.reconstruct_batch(np.arange(3000))
        for m, k, lam in ((20, 4, 0.5), (50, 10, 0.2), (200, 30, 0.9), (3, 5, 0.5), (1, 1, 0.0), (120, 7, 1.0)):
            q = centre + 0.3 * rng.standard_normal(d).astype(np.float32)
            keys = np.concatenate([[10, 50, 77], rng.permutation(200)[: max(0, m - 3)]])[:m].astype(np.int64)
            got = ix.mmr_select(q, keys, k, lam)
            want = maximal_marginal_relevance(np.array([q], dtype=np.float32), stored[keys], lambda_mult=lam, k=k)
            assert got.tolist() == list(want), (d, storage, m, k, lam)
        assert ix.mmr_select(centre, np.zeros(0, dtype=np.int64), 4).size == 0
        with pytest.raises(RuntimeError):
            ix.mmr_select(centre, np.array([3000], dtype=np.int64), 1)
        ix.close()
    # through the LangChain surface
    vs = B200FlatVectorStore(Emb(), D_EMB)
    ref = _ref()
    texts = [f"chunk {i}" for i in range(400)]
    for s in (vs, ref):
        s.add_texts(texts)
    for q in ("about chunk 3", "chunk 77", "far away"):
        got = vs.as_retriever(search_type="mmr", search_kwargs={"k": 5, "fetch_k": 30, "lambda_mult": 0.3}) \
            .get_relevant_documents(q)
        want = ref.max_marginal_relevance_search_by_vector(_vec(q), k=5, fetch_k=30, lambda_mult=0.3)
        _same_docs(got, want)


def test_save_local_load_local_across_gpu_layouts(built_lib, tmp_path, monkeypatch):
    """FAISS.save_local / load_local (llm-prompt-design-helper/vector_store/faiss_vector_store.py:123-133): a store saved
    from one GPU loads into an index spread over two shards and the other way round; answers and ids do not change."""
    from generativeaiexamples_b200 import B200FlatVectorStore

    monkeypatch.setenv("B200R_MULTI_MIN_PIECE", "16")
    texts = [f"chunk {i}" for i in range(500)]
    one = B200FlatVectorStore(Emb(), D_EMB)
    one.add_texts(texts, [{"source": f"f{i % 3}.pdf"} for i in range(500)])
    one.delete([one.index_to_docstore_id[7]])
    one.save_local(str(tmp_path / "a"))
    two = B200FlatVectorStore.load_local(str(tmp_path / "a"), Emb(), allow_dangerous_deserialization=True, devices=[0, 0])
    assert two.index.stats()["n_devices"] == 2 and two.index.ntotal == 499
    two.add_texts(["late chunk"])
    two.save_local(str(tmp_path / "b"))
    back = B200FlatVectorStore.load_local(str(tmp_path / "b"), Emb(), allow_dangerous_deserialization=True)
    assert back.index.stats()["n_devices"] == 1 and back.index.ntotal == 500
    one.add_texts(["late chunk"])
    for q in ("about chunk 3", "late chunk", "chunk 7"):
        want = [(d.page_content, float(s)) for d, s in one.similarity_search_with_score(q, k=5)]
        for store in (two, back):
            got = [(d.page_content, float(s)) for d, s in store.similarity_search_with_score(q, k=5)]
            assert got == want, q


def test_direct_faiss_caller_flow_through_faiss_compat(built_lib, tmp_path):
    """The RAGService path (community/chat-llama-nemotron/backend-rag/src/rag_service.py:88,190-197,289-301,
    364-422) with ``import generativeaiexamples_b200.faiss_compat as faiss`` as its only change:
    normalise -> add -> normalise the query -> search -> write_index -> read_index -> same answers;
    checked against the faiss restatement (T1) fed the same normalised rows."""
    import generativeaiexamples_b200.faiss_compat as faiss
    from oracle.flat_ref import IndexFlatRef

    rng = np.random.default_rng(77)
    dim, n = 384, 5000
    index = faiss.IndexFlatL2(dim)
    embeddings = rng.standard_normal((n, dim)).astype("float32") * rng.uniform(0.2, 3.0, (n, 1)).astype("float32")
    faiss.normalize_L2(embeddings)
    assert np.allclose(np.linalg.norm(embeddings, axis=1), 1.0, atol=1e-5)
    index.add(np.array(embeddings).astype("float32"))
    assert index.ntotal == n and index.d == dim and index.is_trained

    query_embedding = (embeddings[123] + 0.05 * rng.standard_normal(dim)).astype("float32")
    faiss.normalize_L2(query_embedding.reshape(1, -1))
    search_k = 5 * 3
    distances, indices = index.search(np.array([query_embedding]).astype("float32"), search_k)
    assert distances.shape == (1, search_k) and indices.dtype == np.int64 and indices[0, 0] == 123

    ref = IndexFlatRef(dim, 1)
    ref.add(embeddings)
    Dr, Ir = ref.search(query_embedding.reshape(1, -1), search_k)
    np.testing.assert_array_equal(indices, Ir)
    np.testing.assert_allclose(distances, Dr, rtol=1e-5, atol=1e-5)

    path = tmp_path / "faiss.index"
    faiss.write_index(index, str(path))
    with open(path, "rb") as f:
        assert f.read(4) == b"IxF2"
    again = faiss.read_index(str(path))
    assert again.ntotal == n and again.d == dim
    D2, I2 = again.search(np.array([query_embedding]).astype("float32"), search_k)
    np.testing.assert_array_equal(I2, indices)
    np.testing.assert_array_equal(D2, distances)
    ip = faiss.IndexFlatIP(dim)
    ip.add(embeddings)
    Di, Ii = ip.search(query_embedding.reshape(1, -1), 3)
    assert Ii[0, 0] == 123 and Di[0, 0] > Di[0, 1]
    with pytest.raises(ValueError):
        faiss.IndexFlat(dim, 23)
    for ix in (index, again, ip):
        ix.close()
