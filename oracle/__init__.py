"""CPU oracle for the flat-search hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product package
(``generativeaiexamples_b200``) never does; it fails loudly when its CUDA
library is missing.

PARITY UNPINNED: the arithmetic of this path lives in faiss-cpu==1.7.4 and
langchain-community/-core (pinned at
/root/reference/RAG/src/chain_server/requirements.txt:4,20,25), none of which is
vendored in /root/reference or installable in this image, and the reference has
no test or golden vector for the path (SURVEY.md sections 4, 8c).  The pieces:

* ``f64_truth``      T0: float64 brute force, exact top-k, lowest-id tie break.
* ``flat_ref`` (C)   T1: fp32 restatement of faiss 1.7.4 ``IndexFlat`` search/add/
                     remove_ids/normalize_L2 (both the nq<20 and the BLAS path).
* ``lc_faiss_shim``  T2: restatement of LangChain ``FAISS`` + ``VectorStore``
                     relevance / threshold logic on top of T1.
* ``compare``        the tie-aware comparator the north star defines.
"""
from .flat_ref import FlatRef, IndexFlatRef, load_flat_ref  # noqa: F401
from .f64_truth import truth_topk  # noqa: F401
from .compare import assert_topk_parity, recall_at_k  # noqa: F401
