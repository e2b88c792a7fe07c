"""Developer scratch tool (GPU box): scan-path time for bf16 vs fp32 rows across k and nq (4M x 1024)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat, synth, _lib

dev = torch.device("cuda", 0)
d, n, chunk = 1024, 4_000_000, 500_000


def timed(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for storage in ("bf16", "f32"):
    ix = B200IndexFlat(d, 0, reserve_rows=n, storage=storage, search_path=_lib.PATH_SCAN)
    for c in range(n // chunk):
        ix.add(synth.corpus_chunk_device(c, chunk, d, dev))
    es = 2 if storage == "bf16" else 4
    for nq in (1, 4):
        q = torch.randn(nq, d, device=dev)
        q /= q.norm(dim=1, keepdim=True)
        for k in (10, 32, 64, 128):
            ms = timed(lambda: ix.search(q, k))
            print(json.dumps({"storage": storage, "nq": nq, "k": k, "ms": round(ms, 3),
                              "tbs": round(n * d * es / ms / 1e9, 2)}), flush=True)
    ix.close()
    del ix
    torch.cuda.empty_cache()
