"""Writes tests/golden/seeded_topk.npz: expected top-k (ids + float64 scores) for seeded
synthetic corpora, computed with the float64 brute force (oracle/f64_truth.py) in THIS
container.  The corpora are regenerated from their seeds by the tests (numpy PCG64, same
numpy build in the image), so the file stays a few KB.  It pins the GPU results on the
GPU box to values that were fixed before any kernel existed.

Run:  python tests/golden/make_fixtures.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle.f64_truth import truth_topk  # noqa: E402

# name: (n, d, nq, k, metric, normalized, seed)
CASES = {
    "c1_shape_l2": (10000, 1024, 1, 4, "l2", True, 11),       # basic_rag: 10k x 1024, k=4
    "c1_unnorm_l2": (10000, 1024, 3, 4, "l2", False, 12),
    "ip_k10": (20000, 1024, 4, 10, "ip", True, 13),
    "minilm_l2": (30000, 384, 2, 10, "l2", True, 14),          # all-MiniLM-L6-v2 shape
    "odd_dim_ip": (5000, 100, 3, 7, "ip", False, 15),          # d not a multiple of 4
    "k100_ip": (20000, 256, 2, 100, "ip", True, 16),
    "batch17_l2": (8000, 128, 17, 10, "l2", True, 17),         # below faiss's BLAS threshold (20)
}


def make_case(n, d, nq, k, metric, normalized, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    x = rng.standard_normal((n, d), dtype=np.float32)
    if normalized:
        x /= np.linalg.norm(x, axis=1, keepdims=True)
    else:
        x *= rng.uniform(0.5, 1.5, size=(n, 1)).astype(np.float32)
    q = rng.standard_normal((nq, d), dtype=np.float32)
    src = rng.integers(0, n, nq)
    for i in range(0, nq, 2):  # every other query is a perturbed corpus row
        q[i] = x[src[i]] + 0.1 * q[i] / np.sqrt(d)
    if normalized:
        q /= np.linalg.norm(q, axis=1, keepdims=True)
    return np.ascontiguousarray(x), np.ascontiguousarray(q)


if __name__ == "__main__":
    out = {}
    for name, (n, d, nq, k, metric, normalized, seed) in CASES.items():
        x, q = make_case(n, d, nq, k, metric, normalized, seed)
        D, I = truth_topk(metric, q, x, k + 8)  # 8 extra columns adjudicate ties at the k-th place
        out[name + "__D"] = D
        out[name + "__I"] = I
        print(name, I[0, :4], D[0, :4])
    np.savez_compressed(os.path.join(HERE, "seeded_topk.npz"), **out)
