"""Developer scratch tool (GPU box): bf16-storage index, nq 1..4, scan vs GEMM route at several sizes."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat, synth
from generativeaiexamples_b200 import _lib as _b200r_lib
dev = torch.device("cuda", 0)
d = 1024
def timed(fn, reps=20, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
for n in (50_000, 100_000, 250_000, 500_000, 4_000_000):
    ix = B200IndexFlat(d, 0, reserve_rows=n, storage="bf16")
    ch = min(n, 500_000)
    for c in range(n // ch): ix.add(synth.corpus_chunk_device(c, ch, d, dev))
    for nq in (1, 2, 4):
        q = torch.randn(nq, d, device=dev); q /= q.norm(dim=1, keepdim=True)
        rec = {"n": n, "nq": nq}
        for name, m in (("scan", 99), ("gemm", 1)):
            os.environ["B200R_GEMM_MIN_NQ"] = str(m); _b200r_lib.reload_env()
            rec[name] = round(timed(lambda: ix.search(q, 10)), 3)
        os.environ.pop("B200R_GEMM_MIN_NQ"); _b200r_lib.reload_env()
        rec["auto"] = round(timed(lambda: ix.search(q, 10)), 3)
        print(json.dumps(rec), flush=True)
    ix.close(); del ix; torch.cuda.empty_cache()
