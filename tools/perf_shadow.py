"""Developer scratch tool (GPU box): batch search on an fp32 index with and without the bf16 shadow
filter, plus the C4-shaped bf16 search, with fallback counts."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat, synth

dev = torch.device("cuda", 0)
d = 1024
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 4_000_000
chunk = 500_000


def timed(fn, reps=8, warm=3):
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


out = []
ix = B200IndexFlat(d, 0, reserve_rows=n, shadow_bf16=True)
for c in range(n // chunk):
    ix.add(synth.corpus_chunk_device(c, chunk, d, dev))
g = torch.Generator(device=dev).manual_seed(7)
for nq, k in ((256, 10), (64, 10), (1024, 10), (256, 100)):
    q = torch.randn(nq, d, device=dev, generator=g)
    q /= q.norm(dim=1, keepdim=True)
    res = {}
    for on in (True, False):
        ix.set_shadow_filter(on)
        f0 = ix.stats()["n_gemm_fallbacks"]
        ms = timed(lambda: ix.search(q, k))
        fb = (ix.stats()["n_gemm_fallbacks"] - f0) / 11
        D, I = ix.search(q, k)
        res[on] = (ms, fb, D.clone(), I.clone())
    same = bool((res[True][3] == res[False][3]).all() and (res[True][2] == res[False][2]).all())
    rec = {"n": n, "nq": nq, "k": k, "shadow_ms": res[True][0], "fp32_ms": res[False][0],
           "shadow_qps": nq / res[True][0] * 1e3, "fp32_qps": nq / res[False][0] * 1e3,
           "shadow_tflops": 2 * n * d * nq / res[True][0] / 1e9, "fp32_tflops": 2 * n * d * nq / res[False][0] / 1e9,
           "shadow_fallbacks_per_call": res[True][1], "fp32_fallbacks_per_call": res[False][1], "identical": same}
    print(json.dumps(rec), flush=True)
    out.append(rec)
ix.close()
del ix
torch.cuda.empty_cache()

# bf16 storage with the rigorous 2^-8 bound
ix = B200IndexFlat(d, 0, storage="bf16", reserve_rows=n)
for c in range(n // chunk):
    ix.add(synth.corpus_chunk_device(c, chunk, d, dev))
for nq, k in ((1024, 100), (256, 10)):
    q = torch.randn(nq, d, device=dev, generator=g)
    q /= q.norm(dim=1, keepdim=True)
    f0 = ix.stats()["n_gemm_fallbacks"]
    ms = timed(lambda: ix.search(q, k))
    fb = (ix.stats()["n_gemm_fallbacks"] - f0) / 11
    rec = {"storage": "bf16", "n": n, "nq": nq, "k": k, "ms": ms, "tflops": 2 * n * d * nq / ms / 1e9,
           "fallbacks_per_call": fb}
    print(json.dumps(rec), flush=True)
    out.append(rec)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/perf_shadow.json", "w"), indent=1)
