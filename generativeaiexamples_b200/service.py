"""Mirror of the reference's Morpheus vector-DB resource over a FAISS store
(``community/event-driven-rag-cve-analysis/cyber_dev_day/faiss_vdb_service.py:42-277``, same class in
``vision_workflows/.../video_agentic_rag_with_morpheus_riva/src/faiss_vdb_service.py``).

The reference implements exactly one operation there -- ``async similarity_search(embeddings, k)`` -- as an
``asyncio.gather`` of one ``asimilarity_search_by_vector`` per embedding (B executor threads, each an nq=1
faiss search, faiss_vdb_service.py:169-174).  Here the B embeddings go to the GPU as ONE nq=B launch
(``B200FlatVectorStore.search_by_vectors``) and the documents are looked up afterwards; the results are the
same lists of ``Document.dict()``.  Every other operation raises ``NotImplementedError`` with the
reference's message.
"""
from __future__ import annotations

import asyncio
import functools
from typing import Any, List

import numpy as np

from .vectorstore import B200FlatVectorStore, Document, _to_matrix


class B200VectorDBResourceService:
    def __init__(self, store: B200FlatVectorStore, *, name: str = "index") -> None:
        self._index = store
        self._name = name

    @classmethod
    def load(cls, local_dir: str, embeddings, *, name: str = "index", **kwargs: Any) -> "B200VectorDBResourceService":
        """faiss_vdb_service.py:63-66: ``FAISS.load_local(folder_path, embeddings, index_name, allow_dangerous_...=True)``."""
        store = B200FlatVectorStore.load_local(local_dir, embeddings, index_name=name,
                                               allow_dangerous_deserialization=True, **kwargs)
        return cls(store, name=name)

    def _search_sync(self, embeddings, k: int) -> List[List[dict]]:
        if len(embeddings) == 0:
            return []
        if self._index.index is None or self._index.index.ntotal == 0:
            return [[] for _ in embeddings]
        q = _to_matrix(embeddings)
        _, I = self._index.search_by_vectors(q, k)
        I = np.asarray(I.cpu() if hasattr(I, "cpu") else I)
        out = []
        for row in I:
            docs = []
            for i in row:
                if i == -1:  # fewer than k vectors in the store
                    continue
                doc = self._index.docstore.search(self._index.index_to_docstore_id[int(i)])
                if not isinstance(doc, Document):
                    raise ValueError(f"Could not find document for id {self._index.index_to_docstore_id[int(i)]}, got {doc}")
                docs.append(doc.dict())
            out.append(docs)
        return out

    async def similarity_search(self, embeddings: List[List[float]], k: int = 4, **kwargs: Any) -> List[List[dict]]:
        loop = asyncio.get_running_loop()
        return await loop.run_in_executor(None, functools.partial(self._search_sync, embeddings, k))

    @property
    def ntotal(self) -> int:
        """Extension (the reference's ``count`` is unsupported): vectors in the store."""
        return 0 if self._index.index is None else int(self._index.index.ntotal)

    def count(self, **kwargs: Any) -> int:
        raise NotImplementedError("Count operation is not supported in FAISS")

    # -- everything else is unsupported in the reference too (faiss_vdb_service.py:68-277) ------------
    def insert(self, data, **kwargs: Any) -> dict:
        raise NotImplementedError("Insert operation is not supported in FAISS")

    def insert_dataframe(self, df, **kwargs: Any) -> dict:
        raise NotImplementedError("Insert operation is not supported in FAISS")

    def describe(self, **kwargs: Any) -> dict:
        raise NotImplementedError("Describe operation is not supported in FAISS")

    def query(self, query: str, **kwargs: Any) -> Any:
        raise NotImplementedError("Query operation is not supported in FAISS")

    def update(self, data, **kwargs: Any) -> dict:
        raise NotImplementedError("Update operation is not supported in FAISS")

    def delete_by_keys(self, keys, **kwargs: Any) -> Any:
        raise NotImplementedError("Delete by keys operation is not supported in FAISS")

    def delete(self, expr: str, **kwargs: Any) -> dict:
        raise NotImplementedError("Delete operation is not supported in FAISS")

    def retrieve_by_keys(self, keys, **kwargs: Any) -> list:
        raise NotImplementedError("Retrieve by keys operation is not supported in FAISS")

    def drop(self, **kwargs: Any) -> None:
        raise NotImplementedError("Drop operation is not supported in FAISS")
