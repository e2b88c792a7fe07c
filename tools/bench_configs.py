"""Secondary benchmark (GPU box, 1 GPU): the other BASELINE.json configs at their per-GPU shard
sizes, each checked for correctness before timing.  Prints one JSON line per config.

  C1  10k x 1024 fp32, L2, k=4, nq=1 through the LangChain-surface store (threshold 0.25)
  C2  1M x 1024 fp32, IP, k=10, nq=1
  C4s 12.5M x 1024 bf16 (= C4's per-GPU shard of 100M/8), IP, k=100, nq=1024; recall@100 vs fp32 truth
      is measured on a 1M-row slice where the truth is affordable
  C5s 5M x 1024 fp32, L2 with ingest-time normalisation, k=10, nq=64 (full C5 on one GPU)
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from generativeaiexamples_b200 import B200FlatVectorStore, B200IndexFlat, synth

dev = torch.device("cuda", 0)
D = 1024
HBM = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"] \
    if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6650.0


def timed(fn, steps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def build(n, metric, storage="f32", **kw):
    ix = B200IndexFlat(D, metric, storage=storage, reserve_rows=n, **kw)
    for c, lo in enumerate(range(0, n, 1_000_000)):
        ix.add(synth.corpus_chunk_device(c, min(1_000_000, n - lo), D, dev))
    torch.cuda.synchronize()
    return ix


def truth_device(q, X, k, l2=False):
    S = q.double() @ X.double().T
    if l2:
        S = (q.double() ** 2).sum(1, keepdim=True) + (X.double() ** 2).sum(1)[None, :] - 2 * S
        return torch.topk(S, k, dim=1, largest=False)
    return torch.topk(S, k, dim=1)


def c1():
    class Emb:
        def embed_documents(self, t):
            return [self.embed_query(x) for x in t]

        def embed_query(self, t):
            r = np.random.default_rng(abs(hash(t)) % 2 ** 32).standard_normal(D)
            return (r / np.linalg.norm(r)).tolist()

    x = synth.corpus_host(10_000, D)
    vs = B200FlatVectorStore(Emb(), D)
    vs.add_embeddings([(f"chunk {i}", x[i]) for i in range(10_000)])
    q, planted = synth.queries_host(8, D, x)
    docs = vs.similarity_search_with_score_by_vector(q[0].tolist(), k=4)
    assert docs[0][0].page_content == f"chunk {planted[0]}"
    t0 = time.perf_counter()
    n = 200
    for i in range(n):
        vs.similarity_search_with_score_by_vector(q[i % 8].tolist(), k=4)
    lat = (time.perf_counter() - t0) / n
    print(json.dumps({"config": "C1 10k x 1024 fp32 L2 k=4 nq=1 via B200FlatVectorStore (list[float] in, Documents out)",
                      "latency_us": lat * 1e6, "qps": 1 / lat,
                      "note": "41 MB corpus is L2-resident: not an HBM-roofline number"}))


def c2():
    n = 1_000_000
    ix = build(n, 0)
    X = torch.cat([synth.corpus_chunk_device(0, n, D, dev)])
    q = X[:64] + 0.01 * torch.randn(64, D, device=dev)
    Dg, Ig = ix.search(q[:4].contiguous(), 10)
    Dt, It = truth_device(q[:4], X, 10)
    assert (Ig == It).all()
    i = [0]

    def step():
        i[0] += 1
        ix.search(q[i[0] % 64: i[0] % 64 + 1], 10)

    ms = timed(step, 200, 20)
    print(json.dumps({"config": "C2 1M x 1024 fp32 IP k=10 nq=1 (scan path)", "ms": ms, "qps": 1000 / ms,
                      "hbm_gbs": n * 4096 / ms / 1e6, "hbm_frac": n * 4096 / ms / 1e6 / HBM}))
    ix.close()


def c5():
    n = 5_000_000
    ix = build(n, 1, normalize_on_add=True, normalize_queries=True)
    X1 = synth.corpus_chunk_device(0, 1_000_000, D, dev)
    q = (X1[:64] + 0.02 * torch.randn(64, D, device=dev)).contiguous() * 3.0  # un-normalised queries
    Dg, Ig = ix.search(q, 10)
    assert (Ig[:, 0].cpu() == torch.arange(64)).all() and ix.stats()["last_path"] == 2
    ms = timed(lambda: ix.search(q, 10), 20)
    print(json.dumps({"config": "C5 5M x 1024 fp32 L2, normalised at ingest + query, k=10 nq=64 (GEMM path, 1 GPU)",
                      "ms_per_batch": ms, "qps": 64 * 1000 / ms, "hbm_gbs": n * 4096 / ms / 1e6,
                      "hbm_frac": n * 4096 / ms / 1e6 / HBM, "tflops": 2 * 64 * n * D / ms / 1e9,
                      "fallbacks": ix.stats()["n_gemm_fallbacks"]}))
    ix.close()


def c4():
    # recall on a 1M-row slice
    n1 = 1_000_000
    X = synth.corpus_chunk_device(0, n1, D, dev)
    ixs = B200IndexFlat(D, 0, storage="bf16", reserve_rows=n1)
    ixs.add(X)
    q = (X[torch.randint(0, n1, (1024,), device=dev)] + 0.5 * torch.randn(1024, D, device=dev) / D ** 0.5).contiguous()
    q /= q.norm(dim=1, keepdim=True)
    Dg, Ig = ixs.search(q, 100)
    hits = 0
    for lo in range(0, 1024, 128):
        _, It = truth_device(q[lo:lo + 128], X, 100)
        for a, b in zip(Ig[lo:lo + 128].tolist(), It.tolist()):
            hits += len(set(a) & set(b))
    recall = hits / (1024 * 100)
    ixs.close()
    del X
    n = 12_500_000
    ix = build(n, 0, storage="bf16")
    ms = timed(lambda: ix.search(q, 100), 5, 2)
    print(json.dumps({"config": "C4 shard: 12.5M x 1024 bf16 IP k=100 nq=1024 (1/8 of the 100M corpus, GEMM path)",
                      "ms_per_batch": ms, "qps_per_gpu_shard": 1024 * 1000 / ms, "tflops": 2 * 1024 * n * D / ms / 1e9,
                      "recall_at_100_vs_fp32_truth_1M_slice": recall, "fallbacks": ix.stats()["n_gemm_fallbacks"]}))
    ix.close()


def k1k5():
    """K1 ingest and K5 remove_ids throughput on resident data (bytes = read + write of the rows touched)."""
    n = 2_000_000
    x = torch.cat([synth.corpus_chunk_device(c, 1_000_000, D, dev, normalize=False) for c in range(2)])
    for storage, norm in (("f32", False), ("f32", True), ("bf16", True)):
        best = 1e9
        for _ in range(3):
            ix = B200IndexFlat(D, 1, storage=storage, reserve_rows=n, normalize_on_add=norm)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ix.add(x)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
            ix.close()
        es = 2 if storage == "bf16" else 4
        byts = n * D * 4 + n * D * es + n * 4
        print(json.dumps({"config": f"K1 ingest 2M x 1024 fp32 -> {storage}, normalise={norm} (device input)", "ms": best,
                          "rows_per_s": n / best * 1e3, "gbs_read_plus_write": byts / best / 1e6,
                          "hbm_frac": byts / best / 1e6 / HBM}))
    ix = B200IndexFlat(D, 1, reserve_rows=n)
    ix.add(x)
    ids = np.arange(0, n, 1000, dtype=np.int64)  # 2000 scattered rows: everything after row 0 moves
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ix.remove_ids(ids)
    dt = time.perf_counter() - t0
    print(json.dumps({"config": "K5 remove_ids: 2000 scattered ids out of 2M x 1024 fp32 (8.2 GB compacted in place)",
                      "ms": dt * 1e3, "moved_gbs": (n - len(ids)) * D * 4 / dt / 1e9,
                      "note": "gather -> 64 MB bounce -> copy back: 4 bytes of traffic per byte moved"}))
    ix.close()


if __name__ == "__main__":
    which = sys.argv[1:] or ["c1", "c2", "c5", "c4", "k1k5"]
    for w in which:
        globals()[w]()
