"""T2: restatement of LangChain's ``FAISS`` vector store on top of T1 (IndexFlatRef).

TEST INFRASTRUCTURE (see oracle/__init__.py).  Follows, method by method, the
published behaviour of ``langchain_community/vectorstores/faiss.py`` (0.0.2x, resolved
from ``langchain==0.1.9``) and ``langchain_core/vectorstores.py`` (0.1.29) -- the
versions /root/reference/RAG/src/chain_server/requirements.txt:4,20 pin -- as the
reference reaches them from RAG/src/chain_server/utils.py:306,507,547-559 and
RAG/examples/basic_rag/langchain/chains.py:83,156-167,207-215.  Neither package is in
/root/reference or importable here; PARITY UNPINNED.

It deliberately shares no code with generativeaiexamples_b200/vectorstore.py: the
tests run the same call sequence through both and compare what comes back.
"""
from __future__ import annotations

import math
import operator
import uuid
import warnings
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, Iterable, List, Optional, Tuple

import numpy as np

from .flat_ref import FlatRef, IndexFlatRef, METRIC_INNER_PRODUCT, METRIC_L2

EUCLIDEAN_DISTANCE = "EUCLIDEAN_DISTANCE"
MAX_INNER_PRODUCT = "MAX_INNER_PRODUCT"
COSINE = "COSINE"


@dataclass
class RefDocument:
    page_content: str
    metadata: dict = field(default_factory=dict)


class RefDocstore:
    """langchain_community.docstore.in_memory.InMemoryDocstore"""

    def __init__(self, _dict: Optional[Dict[str, RefDocument]] = None):
        self._dict = _dict if _dict is not None else {}

    def add(self, texts: Dict[str, RefDocument]) -> None:
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


class RefFAISS:
    def __init__(self, embedding_function, index: IndexFlatRef, docstore: RefDocstore,
                 index_to_docstore_id: Dict[int, str], relevance_score_fn: Optional[Callable] = None,
                 normalize_L2: bool = False, distance_strategy: str = EUCLIDEAN_DISTANCE):
        self.embedding_function = embedding_function
        self.index = index
        self.docstore = docstore
        self.index_to_docstore_id = index_to_docstore_id
        self.distance_strategy = distance_strategy
        self.override_relevance_score_fn = relevance_score_fn
        self._normalize_L2 = normalize_L2

    # -- embedding helpers ------------------------------------------------
    def _embed_documents(self, texts):
        if hasattr(self.embedding_function, "embed_documents"):
            return self.embedding_function.embed_documents(texts)
        return [self.embedding_function(t) for t in texts]

    def _embed_query(self, text):
        if hasattr(self.embedding_function, "embed_query"):
            return self.embedding_function.embed_query(text)
        return self.embedding_function(text)

    # -- add ----------------------------------------------------------------
    def _add(self, texts, embeddings, metadatas=None, ids=None):
        documents = []
        for i, text in enumerate(texts):
            metadata = metadatas[i] if metadatas else {}
            documents.append(RefDocument(page_content=text, metadata=metadata))
        if ids is None:
            ids = [str(uuid.uuid4()) for _ in texts]
        vector = np.array(embeddings, dtype=np.float32)
        if self._normalize_L2:
            FlatRef.normalize_L2(vector)
        self.index.add(vector)
        self.docstore.add({id_: doc for id_, doc in zip(ids, documents)})
        starting_len = len(self.index_to_docstore_id)
        self.index_to_docstore_id.update({starting_len + j: id_ for j, id_ in enumerate(ids)})
        return ids

    def add_texts(self, texts: Iterable[str], metadatas=None, ids=None, **kwargs):
        texts = list(texts)
        return self._add(texts, self._embed_documents(texts), metadatas=metadatas, ids=ids)

    def add_embeddings(self, text_embeddings: Iterable[Tuple[str, List[float]]], metadatas=None, ids=None, **kwargs):
        texts, embeddings = zip(*text_embeddings)
        return self._add(texts, embeddings, metadatas=metadatas, ids=ids)

    def add_documents(self, documents, **kwargs):
        return self.add_texts([d.page_content for d in documents], [d.metadata for d in documents], **kwargs)

    # -- search -------------------------------------------------------------
    def similarity_search_with_score_by_vector(self, embedding, k: int = 4, filter: Optional[Dict[str, Any]] = None,
                                               fetch_k: int = 20, **kwargs):
        vector = np.array([embedding], dtype=np.float32)
        if self._normalize_L2:
            FlatRef.normalize_L2(vector)
        scores, indices = self.index.search(vector, k if filter is None else fetch_k)
        docs = []
        for j, i in enumerate(indices[0]):
            if i == -1:
                continue
            _id = self.index_to_docstore_id[i]
            doc = self.docstore.search(_id)
            if not isinstance(doc, RefDocument):
                raise ValueError(f"Could not find document for id {_id}, got {doc}")
            if filter is not None:
                flt = {key: [value] if not isinstance(value, list) else value for key, value in filter.items()}
                if all(doc.metadata.get(key) in value for key, value in flt.items()):
                    docs.append((doc, scores[0][j]))
            else:
                docs.append((doc, scores[0][j]))
        score_threshold = kwargs.get("score_threshold")
        if score_threshold is not None:
            cmp = operator.ge if self.distance_strategy == MAX_INNER_PRODUCT else operator.le
            docs = [(doc, similarity) for doc, similarity in docs if cmp(similarity, score_threshold)]
        return docs[:k]

    def similarity_search_with_score(self, query, k=4, filter=None, fetch_k=20, **kwargs):
        return self.similarity_search_with_score_by_vector(self._embed_query(query), k, filter=filter,
                                                           fetch_k=fetch_k, **kwargs)

    def similarity_search_by_vector(self, embedding, k=4, filter=None, fetch_k=20, **kwargs):
        return [d for d, _ in self.similarity_search_with_score_by_vector(embedding, k, filter=filter,
                                                                          fetch_k=fetch_k, **kwargs)]

    def similarity_search(self, query, k=4, filter=None, fetch_k=20, **kwargs):
        return [d for d, _ in self.similarity_search_with_score(query, k, filter=filter, fetch_k=fetch_k, **kwargs)]

    # -- relevance (langchain_core VectorStore) -------------------------------
    @staticmethod
    def _euclidean_relevance_score_fn(distance: float) -> float:
        return 1.0 - distance / math.sqrt(2)

    @staticmethod
    def _cosine_relevance_score_fn(distance: float) -> float:
        return 1.0 - distance

    @staticmethod
    def _max_inner_product_relevance_score_fn(distance: float) -> float:
        if distance > 0:
            return 1.0 - distance
        return -1.0 * distance

    def _select_relevance_score_fn(self):
        if self.override_relevance_score_fn is not None:
            return self.override_relevance_score_fn
        if self.distance_strategy == MAX_INNER_PRODUCT:
            return self._max_inner_product_relevance_score_fn
        if self.distance_strategy == EUCLIDEAN_DISTANCE:
            return self._euclidean_relevance_score_fn
        if self.distance_strategy == COSINE:
            return self._cosine_relevance_score_fn
        raise ValueError("Unknown distance strategy, must be cosine, max_inner_product, or euclidean")

    def similarity_search_with_relevance_scores(self, query, k=4, **kwargs):
        score_threshold = kwargs.pop("score_threshold", None)
        fn = self._select_relevance_score_fn()
        docs_and_scores = self.similarity_search_with_score(query, k=k, **kwargs)
        docs_and_similarities = [(doc, fn(score)) for doc, score in docs_and_scores]
        if any(s < 0.0 or s > 1.0 for _, s in docs_and_similarities):
            warnings.warn(f"Relevance scores must be between 0 and 1, got {docs_and_similarities}")
        if score_threshold is not None:
            docs_and_similarities = [(d, s) for d, s in docs_and_similarities if s >= score_threshold]
            if len(docs_and_similarities) == 0:
                warnings.warn(f"No relevant docs were retrieved using the relevance score threshold {score_threshold}")
        return docs_and_similarities

    # -- MMR ----------------------------------------------------------------
    def max_marginal_relevance_search_by_vector(self, embedding, k=4, fetch_k=20, lambda_mult=0.5):
        q = np.asarray(embedding, dtype=np.float64)
        scores, indices = self.index.search(np.array([embedding], dtype=np.float32), fetch_k)
        cand = [int(i) for i in indices[0] if i != -1]
        if not cand or k <= 0:
            return []
        E = np.stack([self.index.reconstruct(i) for i in cand]).astype(np.float32)

        def cos(a, b):
            na, nb = np.linalg.norm(a), np.linalg.norm(b)
            return 0.0 if na == 0 or nb == 0 else float(np.dot(a, b) / (na * nb))

        to_q = [cos(np.float32(q), e) for e in E]
        chosen = [int(np.argmax(to_q))]
        while len(chosen) < min(k, len(cand)):
            best, best_i = -np.inf, -1
            for i in range(len(cand)):
                if i in chosen:
                    continue
                red = max(cos(E[i], E[j]) for j in chosen)
                sc = lambda_mult * to_q[i] - (1 - lambda_mult) * red
                if sc > best:
                    best, best_i = sc, i
            chosen.append(best_i)
        return [self.docstore.search(self.index_to_docstore_id[cand[i]]) for i in chosen]

    # -- delete -------------------------------------------------------------
    def delete(self, ids: Optional[List[str]] = None, **kwargs):
        if ids is None:
            raise ValueError("No ids provided to delete.")
        missing_ids = set(ids).difference(self.index_to_docstore_id.values())
        if missing_ids:
            raise ValueError(f"Some specified ids do not exist in the current store. Ids not found: {missing_ids}")
        reversed_index = {id_: idx for idx, id_ in self.index_to_docstore_id.items()}
        index_to_delete = [reversed_index[id_] for id_ in ids]
        self.index.remove_ids(np.array(index_to_delete, dtype=np.int64))
        self.docstore.delete(ids)
        remaining_ids = [id_ for i, id_ in sorted(self.index_to_docstore_id.items()) if i not in index_to_delete]
        self.index_to_docstore_id = {i: id_ for i, id_ in enumerate(remaining_ids)}
        return True

    # -- merge_from (langchain_community FAISS.merge_from; reference caller:
    #    community/oran-chatbot-multimodal/vectorstore/vectorstore_updater.py:217) --------------------
    def merge_from(self, target: "RefFAISS") -> None:
        starting_len = len(self.index_to_docstore_id)
        self.index.merge_from(target.index)
        full_info = []
        for i, target_id in target.index_to_docstore_id.items():
            doc = target.docstore.search(target_id)
            if not isinstance(doc, RefDocument):
                raise ValueError("Document should be returned")
            full_info.append((starting_len + i, target_id, doc))
        self.docstore.add({_id: doc for _, _id, doc in full_info})
        self.index_to_docstore_id.update({index: _id for index, _id, _ in full_info})

    # -- retriever ------------------------------------------------------------
    def retrieve(self, query, search_type="similarity", search_kwargs=None):
        """VectorStoreRetriever._get_relevant_documents"""
        search_kwargs = dict(search_kwargs or {})
        if search_type == "similarity":
            return self.similarity_search(query, **search_kwargs)
        if search_type == "similarity_score_threshold":
            thr = search_kwargs.get("score_threshold")
            if thr is None or not isinstance(thr, float):
                raise ValueError("`score_threshold` is not specified with a float value(0~1) "
                                 "in `search_kwargs`.")
            return [d for d, _ in self.similarity_search_with_relevance_scores(query, **search_kwargs)]
        raise ValueError(f"search_type of {search_type} not allowed.")

    @classmethod
    def from_embeddings(cls, text_embeddings, embedding, metadatas=None, ids=None,
                        distance_strategy=EUCLIDEAN_DISTANCE, normalize_L2=False):
        texts, embeddings = zip(*text_embeddings)
        d = len(embeddings[0])
        index = IndexFlatRef(d, METRIC_INNER_PRODUCT if distance_strategy == MAX_INNER_PRODUCT else METRIC_L2)
        vs = cls(embedding, index, RefDocstore(), {}, normalize_L2=normalize_L2, distance_strategy=distance_strategy)
        vs._add(texts, embeddings, metadatas=metadatas, ids=ids)
        return vs
