"""Developer scratch tool (GPU box): 256-query block, bf16 and fp32 rows, with the epilogue's selection switched off
(B200R_GEMM_DEBUG_NOSELECT=1: scores are read from TMEM but nothing is kept; =2: TMEM is not even read) -- the
ceiling of the TMA + MMA main loop.  Results of those runs are garbage by construction."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B200R_GEMM_NOBOOST"] = "1"
import torch
from generativeaiexamples_b200 import B200IndexFlat, synth
from generativeaiexamples_b200 import _lib as _b200r_lib
dev = torch.device("cuda", 0)
d, n, chunk = 1024, 4_000_000, 500_000
def timed(fn, reps=8, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
for storage in ("bf16", "f32"):
    ix = B200IndexFlat(d, 0, reserve_rows=n, storage=storage)
    for c in range(n // chunk): ix.add(synth.corpus_chunk_device(c, chunk, d, dev))
    q = torch.randn(256, d, device=dev); q /= q.norm(dim=1, keepdim=True)
    for mode in ("0", "1", "2", "0"):
        os.environ["B200R_GEMM_DEBUG_NOSELECT"] = mode; _b200r_lib.reload_env()
        ms = timed(lambda: ix.search(q, 10))
        print(json.dumps({"storage": storage, "noselect": mode, "ms": round(ms, 3), "tflops": round(2 * n * d * 256 / ms / 1e9, 1)}), flush=True)
    os.environ["B200R_GEMM_DEBUG_NOSELECT"] = "0"; _b200r_lib.reload_env()
    ix.close(); del ix; torch.cuda.empty_cache()
