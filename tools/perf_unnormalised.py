"""Developer scratch tool (GPU box): un-normalised rows (norms spread over 3x, the HF-embedder case of
utils.py:417): fallbacks and time of a 256-query block, tf32 and bf16-shadow filters, IP and L2."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat, synth
dev = torch.device("cuda", 0)
d, n, chunk = 1024, 2_000_000, 500_000
def timed(fn, reps=6, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
for metric in (0, 1):
    ix = B200IndexFlat(d, metric, reserve_rows=n, shadow_bf16=True)
    g = torch.Generator(device=dev).manual_seed(5)
    for c in range(n // chunk):
        x = synth.corpus_chunk_device(c, chunk, d, dev)
        x *= (0.5 + torch.rand((chunk, 1), device=dev, generator=g))
        ix.add(x)
    q = torch.randn(256, d, device=dev, generator=g)
    q[::2] = x[:128] + 0.1 * q[::2] / d ** 0.5
    for shadow in (False, True):
        ix.set_shadow_filter(shadow)
        f0 = ix.stats()["n_gemm_fallbacks"]
        ms = timed(lambda: ix.search(q, 10))
        fb = (ix.stats()["n_gemm_fallbacks"] - f0) / 8
        print(json.dumps({"metric": "l2" if metric else "ip", "shadow": shadow, "ms": round(ms, 3), "fallbacks_per_call": fb}), flush=True)
    ix.close(); del ix; torch.cuda.empty_cache()
