"""generativeaiexamples_b200 -- B200-native exact dense retrieval behind the LangChain
``VectorStore`` surface the reference's RAG chain server uses.

Scope: only the vector-store add/search step that
``RAG/src/chain_server/utils.py:288-347`` constructs and
``RAG/examples/basic_rag/langchain/chains.py:83,156-167,207-215`` calls.  Everything
that computes runs in ``libb200r.so`` (hand-written CUDA for sm_100a, C ABI declared in
``include/b200r.h``); there is no CPU fallback and importing the index classes without
the built library raises.
"""
from .flat_index import (  # noqa: F401
    B200IndexFlat, B200IndexFlatIP, B200IndexFlatL2, METRIC_INNER_PRODUCT, METRIC_L2, normalize_L2,
)
from .vectorstore import (  # noqa: F401
    B200FlatVectorStore, DistanceStrategy, Document, InMemoryDocstore, VectorStoreRetriever,
)
from .factory import (  # noqa: F401
    create_vectorstore_langchain, del_docs_vectorstore_langchain, get_docs_vectorstore_langchain, get_vectorstore,
)
from .service import B200VectorDBResourceService  # noqa: F401

__version__ = "0.2.0"
