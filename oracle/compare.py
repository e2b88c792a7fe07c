"""Tie-aware comparator (TEST INFRASTRUCTURE, see oracle/__init__.py).

The acceptance rule from BASELINE.json's north_star: in fp32 mode the top-k ids are
identical to the reference's except for ties within 1e-5 relative score, and scores
agree within 1e-5.  Following SURVEY.md section 7 the tolerance is absolute-or-
relative: ``|a-b| <= tol * max(1, |b|)``.
"""
from __future__ import annotations

import numpy as np

TOL = 1e-5


def _close(a, b, tol, scale=1.0):
    return np.abs(a - b) <= tol * np.maximum(scale, np.abs(b))


def assert_topk_parity(D, I, D_ref, I_ref, tol: float = TOL, what: str = "", scale: float = 1.0):
    """D/I: candidate result [nq,k]; D_ref/I_ref: truth [nq,k'] with k' >= k
    (extra truth columns help adjudicate ties at the k-th place).

    ``scale``: magnitude floor of the tolerance, ``|a-b| <= tol * max(scale, |b|)``.  1.0 for
    everything except comparisons against faiss's BLAS expansion ``||q||^2 + ||x||^2 - 2 q.x``
    (nq >= 20), whose fp32 cancellation error is relative to ``||q||^2 + ||x||^2``, not to the
    (possibly tiny) distance itself -- there the caller passes that magnitude."""
    D = np.asarray(D, dtype=np.float64)
    D_ref = np.asarray(D_ref, dtype=np.float64)
    I = np.asarray(I)
    I_ref = np.asarray(I_ref)
    nq, k = I.shape
    assert D.shape == (nq, k) and I_ref.shape[0] == nq and I_ref.shape[1] >= k, what
    # 1. rank-wise scores agree
    ok = _close(D, D_ref[:, :k], tol, scale)
    assert ok.all(), f"{what}: scores differ beyond {tol}: max abs diff " \
                     f"{np.abs(D - D_ref[:, :k]).max():.3e} at {np.argwhere(~ok)[:4].tolist()}"
    # 2. ids agree except inside tie groups
    for qi in range(nq):
        if np.array_equal(I[qi], I_ref[qi, :k]):
            continue
        ref_score = {int(i): s for i, s in zip(I_ref[qi], D_ref[qi]) if i >= 0}
        assert len(set(I[qi][I[qi] >= 0].tolist())) == int((I[qi] >= 0).sum()), \
            f"{what}: duplicate ids in query {qi}: {I[qi]}"
        assert (I[qi] >= 0).sum() == (I_ref[qi, :k] >= 0).sum(), f"{what}: padding differs q={qi}"
        kth = D_ref[qi, k - 1]
        for r in range(k):
            got, want = int(I[qi, r]), int(I_ref[qi, r])
            if got == want:
                continue
            if got in ref_score:
                # swapped within truth list: the two truth scores must tie
                assert _close(ref_score[got], D_ref[qi, r], tol, scale), \
                    f"{what}: q={qi} rank={r} id {got} (truth score {ref_score[got]}) " \
                    f"vs id {want} (truth {D_ref[qi, r]}) is not a tie"
            else:
                # id from beyond the truth window: only legal as a tie with the k-th
                assert _close(D[qi, r], kth, tol, scale), \
                    f"{what}: q={qi} rank={r} id {got} score {D[qi, r]} not in truth top-{I_ref.shape[1]} " \
                    f"and not tied with k-th {kth}"


def recall_at_k(I, I_ref) -> float:
    I = np.asarray(I)
    I_ref = np.asarray(I_ref)
    hits = 0
    for a, b in zip(I, I_ref):
        hits += len(set(a[a >= 0].tolist()) & set(b[b >= 0].tolist()))
    return hits / max(1, int((I_ref >= 0).sum()))
