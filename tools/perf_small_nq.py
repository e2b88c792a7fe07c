"""Developer scratch tool (GPU box): small batches (nq 1..8) on every route -- fp32 scan, tf32 GEMM filter,
bf16-shadow scan, bf16-shadow GEMM filter -- to place the dispatch thresholds.  argv: rows (default 4M)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat, synth
from generativeaiexamples_b200 import _lib as _b200r_lib

dev = torch.device("cuda", 0)
d = 1024
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 4_000_000
chunk = min(n, 500_000)


def timed(fn, reps=20, warm=3):
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


ix = B200IndexFlat(d, 0, reserve_rows=n, shadow_bf16=True)
for c in range(n // chunk):
    ix.add(synth.corpus_chunk_device(c, chunk, d, dev))
for nq in (1, 2, 3, 4, 5, 8):
    q = torch.randn(nq, d, device=dev)
    q /= q.norm(dim=1, keepdim=True)
    rec = {"n": n, "nq": nq}
    for name, shadow, min_nq in (("scan_f32", False, 99), ("gemm_tf32", False, 1), ("scan_shadow", True, 99),
                                 ("gemm_shadow", True, 1)):
        ix.set_shadow_filter(shadow)
        os.environ["B200R_GEMM_MIN_NQ"] = str(min_nq); _b200r_lib.reload_env()
        rec[name] = round(timed(lambda: ix.search(q, 10)), 3)
    print(json.dumps(rec), flush=True)
