"""``B200FlatVectorStore`` -- the LangChain ``VectorStore`` surface of the reference's FAISS
store, served from HBM by libb200r.

It is what ``create_vectorstore_langchain`` returns for ``APP_VECTORSTORE_NAME=b200flat``
(see factory.py) in place of

    FAISS(document_embedder, IndexFlatL2(config.embeddings.dimensions), InMemoryDocstore(), {})

(/root/reference/RAG/src/chain_server/utils.py:305-306), and it answers the calls the
chain server makes on that object:

* ``add_documents`` / ``add_texts``                         chains.py:83, multi_turn_rag/chains.py:65
* ``add_embeddings`` / ``from_embeddings``                  vision_workflows/.../video_preprocess.py:192
* ``as_retriever(search_type="similarity_score_threshold", search_kwargs={score_threshold,k})``
  then ``get_relevant_documents(query, callbacks=[...])``   chains.py:156-163, 207-211
* ``as_retriever()`` (k=4, no threshold)                    chains.py:166; query_decomposition_rag/chains.py:288
* ``similarity_search*`` / ``*_by_vector`` (+ async)        faiss_vdb_service.py:170
* ``delete(ids)`` and ``docstore._dict``                    utils.py:507, 547-559
* ``save_local`` / ``load_local``                           llm-prompt-design-helper/.../faiss_vector_store.py:124-133

Method names, argument meaning, returned types (raw faiss scores as numpy float32:
squared L2 ascending / inner product descending) and error behaviour follow
langchain-community's ``FAISS`` and langchain-core's ``VectorStore`` at the versions the
reference pins (requirements.txt:4,20).  When langchain-core is importable the class
derives from its ``VectorStore`` so isinstance checks and callback plumbing are the real
ones; otherwise a local base with the same methods is used (langchain is not installed in
this image).

Extension beyond LangChain: ``search_by_vectors(Q, k)`` runs ONE batched launch for a
whole query matrix (numpy or CUDA tensor) instead of one call per query.
"""
from __future__ import annotations

import array
import asyncio
import logging
import math
import operator
import struct
import os
import pickle
import uuid
import warnings
from enum import Enum
from functools import partial
from typing import Any, Callable, Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np

from .flat_index import B200IndexFlat, METRIC_INNER_PRODUCT, METRIC_L2

logger = logging.getLogger(__name__)

# ---------------------------------------------------------------------------
# LangChain types: the real ones when available, equivalents otherwise
# ---------------------------------------------------------------------------
try:  # pragma: no cover - langchain is absent from this image
    from langchain_core.documents import Document
    from langchain_core.vectorstores import VectorStore as _VSBase
    from langchain_core.vectorstores import VectorStoreRetriever

    _HAVE_LANGCHAIN = True
except Exception:
    _HAVE_LANGCHAIN = False

    class Document:
        """``langchain_core.documents.Document``: page_content + metadata."""

        def __init__(self, page_content: str, metadata: Optional[dict] = None, **kwargs: Any):
            self.page_content = page_content
            self.metadata = metadata if metadata is not None else {}
            for key, value in kwargs.items():
                setattr(self, key, value)

        def __eq__(self, other):
            return (isinstance(other, Document) and self.page_content == other.page_content
                    and self.metadata == other.metadata)

        def __repr__(self):
            return f"Document(page_content={self.page_content!r}, metadata={self.metadata!r})"

        def dict(self) -> dict:
            """pydantic's ``Document.dict()`` (used by faiss_vdb_service.py:172)."""
            return {"page_content": self.page_content, "metadata": self.metadata, "type": "Document"}

    class _VSBase:
        pass

    class VectorStoreRetriever:
        """``langchain_core.vectorstores.VectorStoreRetriever`` (sync + async entry points)."""

        allowed_search_types = ("similarity", "similarity_score_threshold", "mmr")

        def __init__(self, vectorstore, search_type: str = "similarity", search_kwargs: Optional[dict] = None,
                     tags: Optional[List[str]] = None, **kwargs: Any):
            search_kwargs = dict(search_kwargs or {})
            if search_type not in self.allowed_search_types:
                raise ValueError(f"search_type of {search_type} not allowed. Valid values are: "
                                 f"{self.allowed_search_types}")
            if search_type == "similarity_score_threshold":
                score_threshold = search_kwargs.get("score_threshold")
                if (score_threshold is None) or (not isinstance(score_threshold, float)):
                    raise ValueError("`score_threshold` is not specified with a float value(0~1) in `search_kwargs`.")
            self.vectorstore = vectorstore
            self.search_type = search_type
            self.search_kwargs = search_kwargs
            self.tags = tags or []

        def _get_relevant_documents(self, query: str) -> List[Document]:
            vs, kw = self.vectorstore, self.search_kwargs
            if self.search_type == "similarity":
                return vs.similarity_search(query, **kw)
            if self.search_type == "similarity_score_threshold":
                return [doc for doc, _ in vs.similarity_search_with_relevance_scores(query, **kw)]
            if self.search_type == "mmr":
                return vs.max_marginal_relevance_search(query, **kw)
            raise ValueError(f"search_type of {self.search_type} not allowed.")

        def get_relevant_documents(self, query: str, *, callbacks: Any = None, tags: Any = None,
                                   metadata: Any = None, run_name: Any = None, **kwargs: Any) -> List[Document]:
            handlers = list(callbacks) if isinstance(callbacks, (list, tuple)) else []
            run_id = uuid.uuid4() if handlers else None  # (os.urandom costs ~3 us: only when somebody listens)
            for h in handlers:  # same hook names the OpenTelemetry handler implements
                fn = getattr(h, "on_retriever_start", None)
                if fn:
                    fn({"name": type(self).__name__}, query, run_id=run_id, parent_run_id=None, tags=tags,
                       metadata=metadata)
            try:
                docs = self._get_relevant_documents(query)
            except Exception as exc:
                for h in handlers:
                    fn = getattr(h, "on_retriever_error", None)
                    if fn:
                        fn(exc, run_id=run_id, parent_run_id=None)
                raise
            for h in handlers:
                fn = getattr(h, "on_retriever_end", None)
                if fn:
                    fn(docs, run_id=run_id, parent_run_id=None)
            return docs

        def invoke(self, input: str, config: Optional[dict] = None, **kwargs: Any) -> List[Document]:
            config = config or {}
            return self.get_relevant_documents(input, callbacks=config.get("callbacks"), tags=config.get("tags"),
                                               metadata=config.get("metadata"))

        async def aget_relevant_documents(self, query: str, **kwargs: Any) -> List[Document]:
            loop = asyncio.get_running_loop()
            return await loop.run_in_executor(None, partial(self.get_relevant_documents, query, **kwargs))

        async def ainvoke(self, input: str, config: Optional[dict] = None, **kwargs: Any) -> List[Document]:
            loop = asyncio.get_running_loop()
            return await loop.run_in_executor(None, partial(self.invoke, input, config))


class DistanceStrategy(str, Enum):
    """``langchain_community.vectorstores.utils.DistanceStrategy`` (the members FAISS accepts)."""

    EUCLIDEAN_DISTANCE = "EUCLIDEAN_DISTANCE"
    MAX_INNER_PRODUCT = "MAX_INNER_PRODUCT"
    DOT_PRODUCT = "DOT_PRODUCT"
    JACCARD = "JACCARD"
    COSINE = "COSINE"


class InMemoryDocstore:
    """``langchain_community.docstore.in_memory.InMemoryDocstore``; the chain server reads
    ``._dict`` directly (utils.py:507,547)."""

    def __init__(self, _dict: Optional[Dict[str, Document]] = None):
        self._dict = _dict if _dict is not None else {}

    def add(self, texts: Dict[str, Document]) -> None:
        overlapping = set(texts).intersection(self._dict)
        if overlapping:
            raise ValueError(f"Tried to add ids that already exist: {overlapping}")
        self._dict = {**self._dict, **texts}

    def delete(self, ids: List) -> None:
        overlapping = set(ids).intersection(self._dict)
        if not overlapping:
            raise ValueError(f"Tried to delete ids that does not  exist: {ids}")
        for _id in ids:
            self._dict.pop(_id)

    def search(self, search: str):
        if search not in self._dict:
            return f"ID {search} not found."
        return self._dict[search]


_PACKERS: Dict[int, struct.Struct] = {}

# distance strategy -> the engine's relevance function (b200r_search_threshold)
_DEVICE_RELEVANCE_KINDS = {"EUCLIDEAN_DISTANCE": "EUCLIDEAN_DISTANCE", "COSINE": "COSINE",
                           "MAX_INNER_PRODUCT": "MAX_INNER_PRODUCT"}


def _pack_row(row) -> np.ndarray:
    """One Python list / tuple of floats -> float32 vector.  struct.pack converts C doubles to floats (round to
    nearest, like numpy) about twice as fast as array.array('f') and three times as fast as numpy's object path:
    11 vs 20 vs 37 us for the 1024 values of one query embedding, which is a third of a 10k-chunk search."""
    n = len(row)
    packer = _PACKERS.get(n)
    if packer is None:
        if len(_PACKERS) > 64:
            _PACKERS.clear()
        packer = _PACKERS[n] = struct.Struct(f"{n}f")
    try:
        return np.frombuffer(packer.pack(*row), dtype=np.float32)
    except (OverflowError, struct.error, TypeError):  # values beyond float32 (numpy gives inf), ints mixed in ...
        return np.frombuffer(array.array("f", [float(v) for v in row]), dtype=np.float32) \
            if all(abs(float(v)) < 3.0e38 or v != v for v in row) else np.asarray(row, dtype=np.float32)


def _to_matrix(embeddings) -> np.ndarray:
    """list-of-lists / ndarray -> C-contiguous float32 (n, d); the host marshal LangChain
    does with ``np.array(embeddings, dtype=np.float32)``."""
    if isinstance(embeddings, (list, tuple)) and len(embeddings) > 0:
        first = embeddings[0]
        if isinstance(first, float):
            return _pack_row(embeddings)[None, :]
        if isinstance(first, (list, tuple)) and len(first) > 0 and isinstance(first[0], float):
            d = len(first)
            m = np.empty((len(embeddings), d), dtype=np.float32)
            for i, row in enumerate(embeddings):
                if len(row) != d:
                    raise ValueError(f"embedding {i} has {len(row)} values, expected {d}")
                m[i] = _pack_row(row)
            return m
    m = np.asarray(embeddings, dtype=np.float32)
    if m.ndim == 1:
        m = m[None, :]
    return np.ascontiguousarray(m)


class B200FlatVectorStore(_VSBase):
    """Drop-in for ``langchain_community.vectorstores.FAISS`` over a ``B200IndexFlat``."""

    def __init__(self, embedding_function, index, docstore: Optional[InMemoryDocstore] = None,
                 index_to_docstore_id: Optional[Dict[int, str]] = None,
                 relevance_score_fn: Optional[Callable[[float], float]] = None, normalize_L2: bool = False,
                 distance_strategy: DistanceStrategy = DistanceStrategy.EUCLIDEAN_DISTANCE, *, storage: str = "f32",
                 device: int = 0, reserve_rows: int = 0, coalesce: bool = False, devices=None):
        """``index`` is a ``B200IndexFlat`` or, for lazy construction, the embedding
        dimension (int) / None (sized from the first vectors added, like ``FAISS.__from``)."""
        if not (hasattr(embedding_function, "embed_documents") and hasattr(embedding_function, "embed_query")):
            logger.warning("`embedding_function` is expected to be an Embeddings object, support "
                           "for passing in a function will soon be removed.")
        self.embedding_function = embedding_function
        self.docstore = docstore if docstore is not None else InMemoryDocstore()
        self.index_to_docstore_id = index_to_docstore_id if index_to_docstore_id is not None else {}
        self.distance_strategy = DistanceStrategy(distance_strategy)
        self.override_relevance_score_fn = relevance_score_fn
        self._normalize_L2 = bool(normalize_L2)
        # devices=[0, 1, ...]: a lazily created index spans those GPUs of this process (csrc/multi.cu)
        self._opts = dict(storage=storage, device=device, reserve_rows=reserve_rows, devices=devices)
        # coalesce=True: concurrent single-query searches (e.g. asyncio.gather of
        # asimilarity_search_by_vector, faiss_vdb_service.py:169-174) share one batched launch
        self._coalesce = bool(coalesce)
        self._coalescer = None
        if self.distance_strategy != DistanceStrategy.EUCLIDEAN_DISTANCE and self._normalize_L2:
            warnings.warn(f"Normalizing L2 is not applicable for metric type: {self.distance_strategy}")
        if index is None:
            self.index = None
        elif isinstance(index, B200IndexFlat) or hasattr(index, "search"):
            self.index = index  # any faiss-IndexFlat-shaped object (tests inject the CPU oracle here)
        else:
            self.index = self._new_index(int(index))

    # -- plumbing -------------------------------------------------------------
    @property
    def embeddings(self):
        return self.embedding_function if hasattr(self.embedding_function, "embed_query") else None

    def _metric(self) -> int:
        return METRIC_INNER_PRODUCT if self.distance_strategy == DistanceStrategy.MAX_INNER_PRODUCT else METRIC_L2

    def _new_index(self, d: int) -> B200IndexFlat:
        # normalisation is fused into the ingest / query-prep kernels on the device
        return B200IndexFlat(d, self._metric(), normalize_on_add=self._normalize_L2,
                             normalize_queries=self._normalize_L2, **self._opts)

    def _embed_documents(self, texts: List[str]):
        if hasattr(self.embedding_function, "embed_documents"):
            return self.embedding_function.embed_documents(texts)
        return [self.embedding_function(t) for t in texts]

    def _embed_query(self, text: str):
        if hasattr(self.embedding_function, "embed_query"):
            return self.embedding_function.embed_query(text)
        return self.embedding_function(text)

    # -- ingest (FAISS.__add) ---------------------------------------------------
    def _append(self, texts: Sequence[str], embeddings, metadatas: Optional[List[dict]] = None,
                ids: Optional[List[str]] = None) -> List[str]:
        if not hasattr(self.docstore, "add"):
            raise ValueError("If trying to add texts, the underlying docstore should support "
                             f"adding items, which {self.docstore} does not")
        texts = list(texts)
        docs = [Document(page_content=t, metadata=(metadatas[i] if metadatas else {})) for i, t in enumerate(texts)]
        if ids is None:
            ids = [str(uuid.uuid4()) for _ in texts]
        vectors = embeddings if _is_cuda_tensor(embeddings) else _to_matrix(embeddings)
        if self.index is None:
            self.index = self._new_index(int(vectors.shape[1]))
        self.index.add(vectors)  # normalisation, cast and row norms happen in the ingest kernel
        self.docstore.add(dict(zip(ids, docs)))
        base = len(self.index_to_docstore_id)
        self.index_to_docstore_id.update({base + j: id_ for j, id_ in enumerate(ids)})
        return ids

    def add_texts(self, texts: Iterable[str], metadatas: Optional[List[dict]] = None,
                  ids: Optional[List[str]] = None, **kwargs: Any) -> List[str]:
        texts = list(texts)
        return self._append(texts, self._embed_documents(texts), metadatas=metadatas, ids=ids)

    def add_embeddings(self, text_embeddings: Iterable[Tuple[str, List[float]]],
                       metadatas: Optional[List[dict]] = None, ids: Optional[List[str]] = None,
                       **kwargs: Any) -> List[str]:
        texts, embeddings = zip(*text_embeddings)
        return self._append(texts, list(embeddings), metadatas=metadatas, ids=ids)

    def add_documents(self, documents: List[Document], **kwargs: Any) -> List[str]:
        return self.add_texts([d.page_content for d in documents], [d.metadata for d in documents], **kwargs)

    # -- search -----------------------------------------------------------------
    def search_by_vectors(self, queries, k: int = 4):
        """Batched search: (nq, d) numpy array or CUDA tensor -> raw (D, I), one launch group."""
        if self.index is None:
            raise ValueError("the store is empty and has no dimension yet")
        return self.index.search(queries if _is_cuda_tensor(queries) else _to_matrix(queries), k)

    def _hits(self, scores, indices, k: int, filter: Optional[Dict[str, Any]], kwargs: dict):
        out = []
        wanted = None
        accept = None  # langchain-community >= 0.0.21 also takes a callable over the metadata
        if callable(filter):
            accept = filter
        elif filter is not None:
            wanted = {key: (value if isinstance(value, list) else [value]) for key, value in filter.items()}
        for score, pos in zip(scores, indices):
            if pos == -1:  # fewer than k rows in the index
                continue
            _id = self.index_to_docstore_id[int(pos)]
            doc = self.docstore.search(_id)
            if not isinstance(doc, Document):
                raise ValueError(f"Could not find document for id {_id}, got {doc}")
            if accept is not None:
                if accept(doc.metadata):
                    out.append((doc, score))
            elif wanted is None or all(doc.metadata.get(key) in value for key, value in wanted.items()):
                out.append((doc, score))
        score_threshold = kwargs.get("score_threshold")
        if score_threshold is not None:
            keep = operator.ge if self.distance_strategy in (DistanceStrategy.MAX_INNER_PRODUCT,
                                                             DistanceStrategy.JACCARD) else operator.le
            out = [(doc, s) for doc, s in out if keep(s, score_threshold)]
        return out[:k]

    def similarity_search_with_score_by_vector(self, embedding: List[float], k: int = 4,
                                               filter: Optional[Dict[str, Any]] = None, fetch_k: int = 20,
                                               **kwargs: Any) -> List[Tuple[Document, float]]:
        if self.index is None:
            return []
        want = k if filter is None else fetch_k
        if self._coalesce:
            if self._coalescer is None or self._coalescer.index is not self.index:
                from .batching import SearchCoalescer

                self._coalescer = SearchCoalescer(self.index)
            scores, indices = self._coalescer.search(_to_matrix(embedding), want)
        else:
            scores, indices = self.index.search(_to_matrix(embedding), want)
        return self._hits(scores[0], indices[0], k, filter, kwargs)

    def similarity_search_with_score(self, query: str, k: int = 4, filter: Optional[Dict[str, Any]] = None,
                                     fetch_k: int = 20, **kwargs: Any) -> List[Tuple[Document, float]]:
        return self.similarity_search_with_score_by_vector(self._embed_query(query), k, filter=filter,
                                                           fetch_k=fetch_k, **kwargs)

    def similarity_search_by_vector(self, embedding: List[float], k: int = 4, filter: Optional[Dict[str, Any]] = None,
                                    fetch_k: int = 20, **kwargs: Any) -> List[Document]:
        return [doc for doc, _ in self.similarity_search_with_score_by_vector(embedding, k, filter=filter,
                                                                              fetch_k=fetch_k, **kwargs)]

    def similarity_search(self, query: str, k: int = 4, filter: Optional[Dict[str, Any]] = None, fetch_k: int = 20,
                          **kwargs: Any) -> List[Document]:
        return [doc for doc, _ in self.similarity_search_with_score(query, k, filter=filter, fetch_k=fetch_k, **kwargs)]

    # async variants: LangChain runs the sync method in the default executor
    async def asimilarity_search_with_score_by_vector(self, embedding, k=4, filter=None, fetch_k=20, **kwargs):
        loop = asyncio.get_running_loop()
        return await loop.run_in_executor(None, partial(self.similarity_search_with_score_by_vector, embedding, k=k,
                                                        filter=filter, fetch_k=fetch_k, **kwargs))

    async def asimilarity_search_by_vector(self, embedding, k=4, filter=None, fetch_k=20, **kwargs):
        return [doc for doc, _ in await self.asimilarity_search_with_score_by_vector(embedding, k=k, filter=filter,
                                                                                    fetch_k=fetch_k, **kwargs)]

    async def asimilarity_search_with_score(self, query, k=4, filter=None, fetch_k=20, **kwargs):
        loop = asyncio.get_running_loop()
        return await loop.run_in_executor(None, partial(self.similarity_search_with_score, query, k=k, filter=filter,
                                                        fetch_k=fetch_k, **kwargs))

    async def asimilarity_search(self, query, k=4, filter=None, fetch_k=20, **kwargs):
        return [doc for doc, _ in await self.asimilarity_search_with_score(query, k=k, filter=filter,
                                                                          fetch_k=fetch_k, **kwargs)]

    # -- relevance scores (langchain_core VectorStore) ----------------------------
    @staticmethod
    def _euclidean_relevance_score_fn(distance: float) -> float:
        # LangChain feeds faiss's SQUARED distance here; kept as is for parity
        return 1.0 - distance / math.sqrt(2)

    @staticmethod
    def _cosine_relevance_score_fn(distance: float) -> float:
        return 1.0 - distance

    @staticmethod
    def _max_inner_product_relevance_score_fn(distance: float) -> float:
        return 1.0 - distance if distance > 0 else -1.0 * distance

    def _select_relevance_score_fn(self) -> Callable[[float], float]:
        """Present so that ``search_type="similarity_score_threshold"`` works (the Milvus store
        raises NotImplementedError here and chains.py:164 falls back; FAISS does not)."""
        if self.override_relevance_score_fn is not None:
            return self.override_relevance_score_fn
        table = {
            DistanceStrategy.MAX_INNER_PRODUCT: self._max_inner_product_relevance_score_fn,
            DistanceStrategy.EUCLIDEAN_DISTANCE: self._euclidean_relevance_score_fn,
            DistanceStrategy.COSINE: self._cosine_relevance_score_fn,
        }
        if self.distance_strategy not in table:
            raise ValueError("Unknown distance strategy, must be cosine, max_inner_product, or euclidean")
        return table[self.distance_strategy]

    def _similarity_search_with_relevance_scores(self, query: str, k: int = 4, filter: Optional[Dict[str, Any]] = None,
                                                 fetch_k: int = 20, **kwargs: Any) -> List[Tuple[Document, float]]:
        fn = self._select_relevance_score_fn()
        if fn is None:
            raise ValueError("normalize_score_fn must be provided to FAISS constructor to normalize scores")
        hits = self.similarity_search_with_score(query, k=k, filter=filter, fetch_k=fetch_k, **kwargs)
        return [(doc, fn(score)) for doc, score in hits]

    def _device_relevance(self, query: str, k: int, score_threshold: Optional[float]):
        """Relevance scores and the threshold computed by the engine (``b200r_search_threshold``): the host
        only walks the docstore for hits that passed.  ``None`` when this store's configuration needs
        LangChain's host path (custom relevance function, coalesced searches, a foreign index object)."""
        if (self.override_relevance_score_fn is not None or self._coalesce or self.index is None
                or not hasattr(self.index, "search_threshold")):
            return None
        kind = _DEVICE_RELEVANCE_KINDS.get(self.distance_strategy)
        if kind is None:
            return None
        R, _, I, counts, oor = self.index.search_threshold(_to_matrix(self._embed_query(query)), k, kind,
                                                           score_threshold)
        ids, rel = I[0].tolist(), R[0].tolist()  # one conversion each instead of a numpy scalar per hit
        id_of, find = self.index_to_docstore_id, self.docstore.search

        def docs(lo, hi):
            out = []
            for j in range(lo, hi):
                _id = id_of[ids[j]]
                doc = find(_id)
                if not isinstance(doc, Document):
                    raise ValueError(f"Could not find document for id {_id}, got {doc}")
                out.append((doc, rel[j]))
            return out

        n_kept = int(counts[0])
        kept = docs(0, n_kept)
        if oor[0] > 0:  # LangChain's warning lists every hit, also the ones the threshold drops
            n_valid = sum(1 for v in ids if v >= 0)
            warnings.warn(f"Relevance scores must be between 0 and 1, got {kept + docs(n_kept, n_valid)}")
        return kept

    def similarity_search_with_relevance_scores(self, query: str, k: int = 4,
                                                **kwargs: Any) -> List[Tuple[Document, float]]:
        score_threshold = kwargs.pop("score_threshold", None)
        if not kwargs:
            scored = self._device_relevance(query, k, score_threshold)
            if scored is not None:
                if score_threshold is not None and len(scored) == 0:
                    warnings.warn("No relevant docs were retrieved using the relevance score threshold "
                                  f"{score_threshold}")
                return scored
        scored = self._similarity_search_with_relevance_scores(query, k=k, **kwargs)
        if any(s < 0.0 or s > 1.0 for _, s in scored):
            warnings.warn(f"Relevance scores must be between 0 and 1, got {scored}")
        if score_threshold is not None:
            scored = [(doc, s) for doc, s in scored if s >= score_threshold]
            if len(scored) == 0:
                warnings.warn(f"No relevant docs were retrieved using the relevance score threshold {score_threshold}")
        return scored

    # -- MMR (langchain_community FAISS.max_marginal_relevance_search*) ---------------------
    def max_marginal_relevance_search_with_score_by_vector(self, embedding: List[float], *, k: int = 4,
                                                           fetch_k: int = 20, lambda_mult: float = 0.5,
                                                           filter: Optional[Dict[str, Any]] = None,
                                                           ) -> List[Tuple[Document, float]]:
        """The engine fetches ``fetch_k`` candidates (one search + one gather of their rows); the
        greedy diversity selection over those <= fetch_k vectors is LangChain's own host-side
        ``maximal_marginal_relevance`` (cosine similarities), restated below."""
        if self.index is None:
            return []
        scores, indices = self.index.search(_to_matrix(embedding), fetch_k if filter is None else fetch_k * 2)
        scores, indices = scores[0], indices[0]
        if filter is not None:
            if callable(filter):
                accept = filter
            else:
                wanted = {key: (value if isinstance(value, list) else [value]) for key, value in filter.items()}

                def accept(md):
                    return all(md.get(key) in value for key, value in wanted.items())
            keep = []
            for pos in indices:
                if pos == -1:
                    continue
                doc = self.docstore.search(self.index_to_docstore_id[int(pos)])
                if not isinstance(doc, Document):
                    raise ValueError(f"Could not find document for id {self.index_to_docstore_id[int(pos)]}, got {doc}")
                if accept(doc.metadata):
                    keep.append(int(pos))
            order = {int(p_): j for j, p_ in enumerate(indices)}
            scores = np.array([scores[order[p_]] for p_ in keep], dtype=np.float32)
            indices = np.array(keep, dtype=np.int64)
        valid = [int(i) for i in indices if i != -1]  # -1 happens when not enough docs are returned
        if hasattr(self.index, "mmr_select") and len(valid) <= 2048:
            # the candidates' rows are gathered and compared on the GPU; only the picks come back
            selected = [int(j) for j in self.index.mmr_select(np.asarray(embedding, dtype=np.float32),
                                                              np.array(valid, dtype=np.int64), k, lambda_mult)]
        else:
            if hasattr(self.index, "reconstruct_batch"):
                embeddings = self.index.reconstruct_batch(np.array(valid, dtype=np.int64))
            else:
                embeddings = (np.stack([self.index.reconstruct(i) for i in valid]) if valid
                              else np.zeros((0, 1), np.float32))
            selected = maximal_marginal_relevance(np.array([embedding], dtype=np.float32), embeddings, k=k,
                                                  lambda_mult=lambda_mult)
        out = []
        for i in selected:
            if indices[i] == -1:
                continue
            _id = self.index_to_docstore_id[int(indices[i])]
            doc = self.docstore.search(_id)
            if not isinstance(doc, Document):
                raise ValueError(f"Could not find document for id {_id}, got {doc}")
            out.append((doc, scores[i]))
        return out

    def max_marginal_relevance_search_by_vector(self, embedding: List[float], k: int = 4, fetch_k: int = 20,
                                                lambda_mult: float = 0.5, filter: Optional[Dict[str, Any]] = None,
                                                **kwargs: Any) -> List[Document]:
        return [doc for doc, _ in self.max_marginal_relevance_search_with_score_by_vector(
            embedding, k=k, fetch_k=fetch_k, lambda_mult=lambda_mult, filter=filter)]

    def max_marginal_relevance_search(self, query: str, k: int = 4, fetch_k: int = 20, lambda_mult: float = 0.5,
                                      filter: Optional[Dict[str, Any]] = None, **kwargs: Any) -> List[Document]:
        return self.max_marginal_relevance_search_by_vector(self._embed_query(query), k=k, fetch_k=fetch_k,
                                                            lambda_mult=lambda_mult, filter=filter, **kwargs)

    # -- retriever ------------------------------------------------------------------
    def as_retriever(self, **kwargs: Any) -> VectorStoreRetriever:
        if _HAVE_LANGCHAIN:  # pragma: no cover
            return super().as_retriever(**kwargs)
        tags = list(kwargs.pop("tags", None) or []) + [type(self).__name__]
        if self.embeddings is not None:
            tags.append(type(self.embeddings).__name__)
        return VectorStoreRetriever(vectorstore=self, tags=tags, **kwargs)

    # -- delete (FAISS.delete) ---------------------------------------------------------
    def delete(self, ids: Optional[List[str]] = None, **kwargs: Any) -> Optional[bool]:
        if ids is None:
            raise ValueError("No ids provided to delete.")
        known = set(self.index_to_docstore_id.values())
        missing_ids = set(ids).difference(known)
        if missing_ids:
            raise ValueError("Some specified ids do not exist in the current store. Ids not found: "
                             f"{missing_ids}")
        position_of = {id_: pos for pos, id_ in self.index_to_docstore_id.items()}
        doomed = {position_of[id_] for id_ in ids}
        self.index.remove_ids(np.fromiter(doomed, dtype=np.int64, count=len(doomed)))
        self.docstore.delete(ids)
        survivors = [id_ for pos, id_ in sorted(self.index_to_docstore_id.items()) if pos not in doomed]
        self.index_to_docstore_id = dict(enumerate(survivors))
        return True

    # -- constructors ---------------------------------------------------------------------
    def merge_from(self, target: "B200FlatVectorStore") -> None:
        """LangChain ``FAISS.merge_from`` (reference caller: community/oran-chatbot-multimodal/vectorstore/
        vectorstore_updater.py:217): the target's vectors and documents are appended; its index is left empty."""
        if target.index is None or target.index.ntotal == 0:
            return
        if self.index is None:
            self.index = self._new_index(int(target.index.d))
        starting_len = len(self.index_to_docstore_id)
        self.index.merge_from(target.index)
        full_info = []
        for i, target_id in target.index_to_docstore_id.items():
            doc = target.docstore.search(target_id)
            if not isinstance(doc, Document):
                raise ValueError("Document should be returned")
            full_info.append((starting_len + i, target_id, doc))
        self.docstore.add({_id: doc for _, _id, doc in full_info})
        self.index_to_docstore_id.update({index: _id for index, _id, _ in full_info})

    @classmethod
    def from_embeddings(cls, text_embeddings: Iterable[Tuple[str, List[float]]], embedding,
                        metadatas: Optional[List[dict]] = None, ids: Optional[List[str]] = None,
                        **kwargs: Any) -> "B200FlatVectorStore":
        texts, embeddings = zip(*text_embeddings)
        store = cls(embedding, None, InMemoryDocstore(), {}, **kwargs)
        store._append(texts, list(embeddings), metadatas=metadatas, ids=ids)
        return store

    @classmethod
    def from_texts(cls, texts: List[str], embedding, metadatas: Optional[List[dict]] = None,
                   ids: Optional[List[str]] = None, **kwargs: Any) -> "B200FlatVectorStore":
        store = cls(embedding, None, InMemoryDocstore(), {}, **kwargs)
        store._append(texts, store._embed_documents(list(texts)), metadatas=metadatas, ids=ids)
        return store

    @classmethod
    def from_documents(cls, documents: List[Document], embedding, **kwargs: Any) -> "B200FlatVectorStore":
        return cls.from_texts([d.page_content for d in documents], embedding,
                              metadatas=[d.metadata for d in documents], **kwargs)

    # -- persistence (FAISS.save_local / load_local) ------------------------------------------
    def save_local(self, folder_path: str, index_name: str = "index") -> None:
        os.makedirs(folder_path, exist_ok=True)
        if self.index is None:
            raise ValueError("nothing to save: the store has no index yet")
        self.index.write(os.path.join(folder_path, f"{index_name}.faiss"))
        with open(os.path.join(folder_path, f"{index_name}.pkl"), "wb") as f:
            pickle.dump((self.docstore, self.index_to_docstore_id), f)

    @classmethod
    def load_local(cls, folder_path: str, embeddings, index_name: str = "index", *,
                   allow_dangerous_deserialization: bool = False, **kwargs: Any) -> "B200FlatVectorStore":
        if not allow_dangerous_deserialization:
            raise ValueError("The de-serialization relies loading a pickle file. Pickle files can be modified to "
                             "deliver a malicious payload. Set `allow_dangerous_deserialization` to `True` only "
                             "if you trust the source of the file.")
        storage = kwargs.pop("storage", "f32")
        device = kwargs.pop("device", 0)
        devices = kwargs.pop("devices", None)  # [0, 1, ...]: load straight into an index spread over those GPUs
        if devices is None and os.getenv("B200R_DEVICES"):
            devices = [int(t) for t in os.getenv("B200R_DEVICES").replace(" ", "").split(",") if t != ""]
        normalize = bool(kwargs.get("normalize_L2", False))
        index = B200IndexFlat.read(os.path.join(folder_path, f"{index_name}.faiss"), storage=storage, device=device,
                                   devices=devices, normalize_on_add=normalize, normalize_queries=normalize)
        with open(os.path.join(folder_path, f"{index_name}.pkl"), "rb") as f:
            docstore, index_to_docstore_id = pickle.load(f)
        if "distance_strategy" not in kwargs and index.metric_type == METRIC_INNER_PRODUCT:
            kwargs["distance_strategy"] = DistanceStrategy.MAX_INNER_PRODUCT
        return cls(embeddings, index, docstore, index_to_docstore_id, storage=storage, device=device, devices=devices,
                   **kwargs)


def _cosine_similarity(X: np.ndarray, Y: np.ndarray) -> np.ndarray:
    """langchain_community.utils.math.cosine_similarity: row-wise, zero norms give 0."""
    if len(X) == 0 or len(Y) == 0:
        return np.array([])
    X = np.array(X)
    Y = np.array(Y)
    if X.shape[1] != Y.shape[1]:
        raise ValueError(f"Number of columns in X and Y must be the same. X has shape {X.shape} and Y has shape "
                         f"{Y.shape}.")
    X_norm = np.linalg.norm(X, axis=1)
    Y_norm = np.linalg.norm(Y, axis=1)
    with np.errstate(divide="ignore", invalid="ignore"):
        similarity = np.dot(X, Y.T) / np.outer(X_norm, Y_norm)
    similarity[np.isnan(similarity) | np.isinf(similarity)] = 0.0
    return similarity


def maximal_marginal_relevance(query_embedding: np.ndarray, embedding_list, lambda_mult: float = 0.5,
                               k: int = 4) -> List[int]:
    """langchain_community.vectorstores.utils.maximal_marginal_relevance (greedy MMR)."""
    if min(k, len(embedding_list)) <= 0:
        return []
    if query_embedding.ndim == 1:
        query_embedding = np.expand_dims(query_embedding, axis=0)
    similarity_to_query = _cosine_similarity(query_embedding, embedding_list)[0]
    most_similar = int(np.argmax(similarity_to_query))
    idxs = [most_similar]
    selected = np.array([embedding_list[most_similar]])
    while len(idxs) < min(k, len(embedding_list)):
        best_score = -np.inf
        idx_to_add = -1
        similarity_to_selected = _cosine_similarity(embedding_list, selected)
        for i, query_score in enumerate(similarity_to_query):
            if i in idxs:
                continue
            redundant_score = max(similarity_to_selected[i])
            equation_score = lambda_mult * query_score - (1 - lambda_mult) * redundant_score
            if equation_score > best_score:
                best_score = equation_score
                idx_to_add = i
        idxs.append(idx_to_add)
        selected = np.append(selected, [embedding_list[idx_to_add]], axis=0)
    return idxs


def _is_cuda_tensor(x) -> bool:
    return type(x).__module__.startswith("torch") and getattr(x, "is_cuda", False)
