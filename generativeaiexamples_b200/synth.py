"""Seeded synthetic embeddings for the benchmark configs.

There is no network for real corpora, so every config runs on Gaussian vectors of the
NV-EmbedQA-E5-v5 / NV-CLIP shape (d = 1024), L2-normalised like those models' outputs.
Host generators use numpy PCG64 (chunk c of 1M rows has seed 1234 + c, so any shard can be
regenerated without materialising the whole corpus); device generators use torch's Philox
on the GPU for the 10M+ row configs, again one seed per 1M-row chunk.  Half of the
queries are perturbed corpus rows so that true near neighbours exist (pure Gaussians at
d = 1024 have |cos| < 0.15 and would all fall under the reference's 0.25 threshold).
"""
from __future__ import annotations

import numpy as np

CHUNK_ROWS = 1_000_000
CORPUS_SEED = 1234
QUERY_SEED = 4321


def corpus_chunk_host(chunk_idx: int, rows: int, d: int, normalize: bool = True, scale_rows: bool = False) -> np.ndarray:
    rng = np.random.Generator(np.random.PCG64(CORPUS_SEED + chunk_idx))
    x = rng.standard_normal((rows, d), dtype=np.float32)
    if normalize:
        x /= np.linalg.norm(x, axis=1, keepdims=True)
    if scale_rows:  # un-normalised HF-embedder case (utils.py:417): exercise true L2
        x *= rng.uniform(0.5, 1.5, size=(rows, 1)).astype(np.float32)
    return x


def corpus_host(n: int, d: int, normalize: bool = True, scale_rows: bool = False) -> np.ndarray:
    out = np.empty((n, d), dtype=np.float32)
    for c, lo in enumerate(range(0, n, CHUNK_ROWS)):
        hi = min(n, lo + CHUNK_ROWS)
        out[lo:hi] = corpus_chunk_host(c, hi - lo, d, normalize, scale_rows)
    return out


def queries_host(nq: int, d: int, corpus: np.ndarray, noise: float = 0.1, normalize: bool = True):
    """Returns (Q, planted): planted[i] is the corpus row query i was derived from, or -1."""
    rng = np.random.Generator(np.random.PCG64(QUERY_SEED))
    q = rng.standard_normal((nq, d), dtype=np.float32)
    planted = np.full(nq, -1, dtype=np.int64)
    n = corpus.shape[0]
    for i in range(nq):
        if i % 2 == 0 and n > 0:  # even queries: a perturbed corpus row (nq = 1 -> planted)
            j = int(rng.integers(0, n))
            base = corpus[j] / max(np.linalg.norm(corpus[j]), 1e-30)
            q[i] = base + noise * q[i] / np.sqrt(d)
            planted[i] = j
    if normalize:
        q /= np.linalg.norm(q, axis=1, keepdims=True)
    return q, planted


def corpus_chunk_device(chunk_idx: int, rows: int, d: int, device, normalize: bool = True):
    """Same layout on the GPU (torch Philox; not bit-identical to the host generator)."""
    import torch

    g = torch.Generator(device=device)
    g.manual_seed(CORPUS_SEED + chunk_idx)
    x = torch.randn((rows, d), generator=g, device=device, dtype=torch.float32)
    if normalize:
        x /= x.norm(dim=1, keepdim=True)
    return x


def queries_device(nq: int, d: int, index_rows_fn, n: int, device, noise: float = 0.1):
    """index_rows_fn(ids: LongTensor) -> (len, d) tensor of corpus rows (for planting)."""
    import torch

    g = torch.Generator(device=device)
    g.manual_seed(QUERY_SEED)
    q = torch.randn((nq, d), generator=g, device=device, dtype=torch.float32)
    planted = torch.full((nq,), -1, dtype=torch.int64)
    cpu_rng = np.random.Generator(np.random.PCG64(QUERY_SEED))
    for i in range(0, nq, 2):
        if n > 0:
            planted[i] = int(cpu_rng.integers(0, n))
    sel = (planted >= 0).nonzero().flatten()
    if len(sel):
        base = index_rows_fn(planted[sel])
        base = base / base.norm(dim=1, keepdim=True).clamp_min(1e-30)
        q[sel.to(device)] = base + noise * q[sel.to(device)] / (d ** 0.5)
    q /= q.norm(dim=1, keepdim=True)
    return q, planted
