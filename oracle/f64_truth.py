"""T0 ground truth: float64 brute force with exact top-k and lowest-id tie break.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Defines what "the right answer" is
for the semantics of faiss ``IndexFlat.search``: L2
returns SQUARED distance ascending, IP returns the dot product descending,
``k > ntotal`` pads with id -1 and +FLT_MAX / -FLT_MAX.
"""
from __future__ import annotations

import numpy as np

FLT_MAX = np.finfo(np.float32).max


def scores_f64(metric: str, q: np.ndarray, x: np.ndarray) -> np.ndarray:
    q64 = np.asarray(q, dtype=np.float64)
    x64 = np.asarray(x, dtype=np.float64)
    ip = q64 @ x64.T
    if metric == "ip":
        return ip
    if metric == "l2":
        # explicit differences are O(nq*N*d) memory; the f64 expansion is exact
        # to ~1e-13 for the magnitudes used here
        qn = np.einsum("ij,ij->i", q64, q64)[:, None]
        xn = np.einsum("ij,ij->i", x64, x64)[None, :]
        return np.maximum(qn + xn - 2.0 * ip, 0.0)
    raise ValueError(metric)


def truth_topk(metric: str, q: np.ndarray, x: np.ndarray, k: int, chunk: int = 262144):
    """Returns (D float64 [nq,k], I int64 [nq,k]); best first, ties by lowest id."""
    q = np.atleast_2d(q)
    nq, n = q.shape[0], x.shape[0]
    sign = 1.0 if metric == "l2" else -1.0  # sort key ascending
    best_s = np.full((nq, 0), 0.0)
    best_i = np.full((nq, 0), 0, dtype=np.int64)
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        s = sign * scores_f64(metric, q, x[lo:hi])
        ids = np.broadcast_to(np.arange(lo, hi, dtype=np.int64), s.shape)
        s = np.concatenate([best_s, s], axis=1)
        ids = np.concatenate([best_i, ids], axis=1)
        order = np.lexsort((ids, s), axis=1)[:, :k]
        best_s = np.take_along_axis(s, order, axis=1)
        best_i = np.take_along_axis(ids, order, axis=1)
    D = np.full((nq, k), FLT_MAX if metric == "l2" else -FLT_MAX, dtype=np.float64)
    I = np.full((nq, k), -1, dtype=np.int64)
    m = best_s.shape[1]
    D[:, :m] = sign * best_s
    I[:, :m] = best_i
    return D, I
