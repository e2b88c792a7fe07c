"""Developer scratch tool (GPU box): time one 256-query block against K' (B200R_GEMM_KPRIME) for the
bf16-storage and bf16-shadow filters; prints fallbacks per call."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat, synth
from generativeaiexamples_b200 import _lib as _b200r_lib

dev = torch.device("cuda", 0)
d, n, chunk = 1024, 4_000_000, 500_000
mode = sys.argv[1]  # bf16 | shadow | f32
os.environ["B200R_GEMM_NOBOOST"] = "1"; _b200r_lib.reload_env()


def timed(fn, reps=6, warm=2):
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


kw = {"storage": "bf16"} if mode == "bf16" else ({"shadow_bf16": True} if mode == "shadow" else {})
ix = B200IndexFlat(d, 0, reserve_rows=n, **kw)
for c in range(n // chunk):
    ix.add(synth.corpus_chunk_device(c, chunk, d, dev))
g = torch.Generator(device=dev).manual_seed(7)
q = torch.randn(256, d, device=dev, generator=g)
q /= q.norm(dim=1, keepdim=True)
for k, kps in ((10, (0, 32, 48, 64, 96, 128, 192)), (100, (224, 280, 324, 400, 488, 600))):
    for kp in kps:
        if kp == 0:
            os.environ["B200R_GEMM_KPCAP"] = "32"; _b200r_lib.reload_env()
            os.environ.pop("B200R_GEMM_KPRIME", None); _b200r_lib.reload_env()
        else:
            os.environ["B200R_GEMM_KPCAP"] = "256"; _b200r_lib.reload_env()
            os.environ["B200R_GEMM_KPRIME"] = str(kp); _b200r_lib.reload_env()
        f0 = ix.stats()["n_gemm_fallbacks"]
        ms = timed(lambda: ix.search(q, k))
        fb = (ix.stats()["n_gemm_fallbacks"] - f0) / 8
        print(json.dumps({"mode": mode, "k": k, "kprime": kp or 32, "kp_cap": 32 if kp == 0 else 256, "ms": round(ms, 3),
                          "tflops": round(2 * n * d * 256 / ms / 1e9, 1), "fallbacks_per_call": fb}), flush=True)
