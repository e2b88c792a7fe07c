"""The factory branch a maintainer adds to the chain server.

Mirrors ``create_vectorstore_langchain`` / ``get_vectorstore`` of
/root/reference/RAG/src/chain_server/utils.py:288-347 for the one store this package
provides.  Configuration follows the reference's ``APP_<SECTION>_<FIELD>`` environment
convention (``RAG/src/chain_server/configuration_wizard.py:38``,
``configuration.py:20-44,109-113``): ``APP_VECTORSTORE_NAME=b200flat`` selects the store,
``APP_EMBEDDINGS_DIMENSIONS`` (default 1024) sizes it; ``faiss`` is accepted as an alias
so an existing FAISS deployment switches over by installing this package.

Engine options (no reference counterpart) use the ``B200R_`` prefix:
``B200R_DEVICE`` (CUDA ordinal, default 0), ``B200R_DEVICES`` (comma-separated ordinals, e.g.
``0,1,2,3,4,5,6,7``: ONE store whose rows are spread over those GPUs of this process -- the chain
server is a single process, ``RAG/src/chain_server/server.py:203-238``), ``B200R_STORAGE``
(``f32`` | ``bf16``), ``B200R_RESERVE_ROWS`` (rows of HBM reserved up front).
"""
from __future__ import annotations

import logging
import os
from typing import Optional

from .flat_index import B200IndexFlatL2
from .vectorstore import B200FlatVectorStore, InMemoryDocstore

logger = logging.getLogger(__name__)

SUPPORTED_NAMES = ("b200flat", "faiss")


def create_vectorstore_langchain(document_embedder, collection_name: str = "", *, name: Optional[str] = None,
                                 dimensions: Optional[int] = None) -> B200FlatVectorStore:
    """utils.py:288-332, FAISS branch (:305-306): an empty squared-L2 flat store of
    ``config.embeddings.dimensions`` columns with an in-memory docstore."""
    name = name if name is not None else os.getenv("APP_VECTORSTORE_NAME", "b200flat")
    if not collection_name:
        collection_name = os.getenv("COLLECTION_NAME", "vector_db")
    if name not in SUPPORTED_NAMES:
        raise ValueError(f"{name} vector database is not supported")
    d = int(dimensions if dimensions is not None else os.getenv("APP_EMBEDDINGS_DIMENSIONS", "1024"))
    storage = os.getenv("B200R_STORAGE", "f32")
    devices = [int(t) for t in os.getenv("B200R_DEVICES", "").replace(" ", "").split(",") if t != ""]
    index = B200IndexFlatL2(d, device=int(os.getenv("B200R_DEVICE", "0")), devices=devices or None, storage=storage,
                            reserve_rows=int(os.getenv("B200R_RESERVE_ROWS", "0")),
                            # opt-in bf16 copy of the fp32 rows for the filter pass (+50 % HBM, same answers)
                            shadow_bf16=storage == "f32" and os.getenv("B200R_SHADOW_BF16", "0") == "1")
    vectorstore = B200FlatVectorStore(document_embedder, index, InMemoryDocstore(), {})
    logger.info("Vector store created and saved.")
    return vectorstore


def get_vectorstore(vectorstore: Optional[B200FlatVectorStore], document_embedder) -> B200FlatVectorStore:
    """utils.py:335-347: hand back the module-global store, creating it on first use."""
    if vectorstore is None:
        return create_vectorstore_langchain(document_embedder)
    return vectorstore


def get_docs_vectorstore_langchain(vectorstore: B200FlatVectorStore):
    """utils.py:492-529, FAISS branch (:506-510): file names of the ingested documents.  As in the
    reference any failure (a document without ``metadata['source']``, a store without a docstore)
    is logged and yields ``[]``."""
    try:
        in_memory_docstore = vectorstore.docstore._dict
        filenames = [os.path.basename(doc.metadata["source"]) for doc in in_memory_docstore.values()]
        return list(dict.fromkeys(filenames))  # the reference returns list(set(...)): same members
    except Exception as e:
        logger.error(f"Error occurred while retrieving documents: {e}")
    return []


def del_docs_vectorstore_langchain(vectorstore: B200FlatVectorStore, filenames) -> bool:
    """utils.py:532-603, FAISS branch (:546-560): delete every chunk of the named files; ``False`` when
    a file is unknown or, as in the reference, when anything raises on the way."""
    try:
        in_memory_docstore = vectorstore.docstore._dict
        for filename in filenames:
            ids_list = [doc_id for doc_id, doc_data in in_memory_docstore.items()
                        if os.path.basename(doc_data.metadata["source"]) == filename]
            if not ids_list:
                logger.info("File does not exist in the vectorstore")
                return False
            vectorstore.delete(ids_list)
            logger.info(f"Deleted documents with filenames {filename}")
    except Exception as e:
        logger.error(f"Error occurred while deleting documents: {e}")
        return False
    return True
