"""Developer scratch tool (GPU box): a small pass over every kernel for compute-sanitizer."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from generativeaiexamples_b200 import B200IndexFlat, normalize_L2
rng = np.random.default_rng(0)
for d, n in ((1024, 3000), (100, 777)):
    x = rng.standard_normal((n, d), dtype=np.float32)
    q = rng.standard_normal((9, d), dtype=np.float32)
    for metric in (0, 1):
        for storage in ("f32", "bf16"):
            ix = B200IndexFlat(d, metric, storage=storage, normalize_on_add=(metric == 1), normalize_queries=(metric == 1))
            ix.add(x[: n // 2]); ix.add(x[n // 2:])
            for nq, k in ((1, 10), (3, 100), (9, 10), (9, 200), (1, 300)):
                D, I = ix.search(q[:nq], k)
                assert I.shape == (nq, k)
            ix.remove_ids(np.arange(0, n, 7, dtype=np.int64))
            ix.search(q[:2], 5)
            ix.reconstruct(3)
            ix.close()
y = x.copy(); normalize_L2(y)
print("sanitize pass done")
