"""Developer tool (GPU box): the batch-1 scan on an 8-GPU-sized shard (1.25M x 1024 fp32 = 5.12 GB) and on
the full 10M rows -- static vs dynamic tile assignment, and where a launch's time goes (B200R_SCAN_TRACE).
argv: rows (default 1.25M)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat, _lib, synth

dev = torch.device("cuda", 0)
d, k = 1024, 10
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_250_000
ix = B200IndexFlat(d, 0, reserve_rows=n)
for c, lo in enumerate(range(0, n, 500_000)):
    ix.add(synth.corpus_chunk_device(c, min(500_000, n - lo), d, dev))
Q = torch.randn(64, d, device=dev)
Q /= Q.norm(dim=1, keepdim=True)
D = torch.empty(1, k, device=dev)
I = torch.empty(1, k, dtype=torch.int64, device=dev)
torch.cuda.synchronize()


def timed(reps=200, warm=20):
    for i in range(warm):
        ix.search(Q[i % 64: i % 64 + 1], k, D=D, I=I)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        ix.search(Q[i % 64: i % 64 + 1], k, D=D, I=I)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


ref = None
for rnd in range(3):
    for dyn in ("0", "1", "2"):
        os.environ["B200R_SCAN_DYNAMIC"] = dyn
        _lib.reload_env()
        us = timed()
        ix.search(Q[:1], k, D=D, I=I)
        torch.cuda.synchronize()
        if ref is None:
            ref = (D.clone(), I.clone())
        assert torch.equal(I, ref[1]) and torch.equal(D, ref[0])
        print(json.dumps({"rows": n, "dynamic_tiles": dyn, "us_per_query": round(us, 1),
                          "GBps": round(n * 4096 / us / 1e3, 1)}), flush=True)
for dyn in ("0", "2"):
    os.environ["B200R_SCAN_DYNAMIC"] = dyn
    os.environ["B200R_SCAN_TRACE"] = "1"
    _lib.reload_env()
    print("trace, dynamic_tiles =", dyn, flush=True)
    for i in range(4):
        ix.search(Q[i: i + 1], k, D=D, I=I)
    os.environ["B200R_SCAN_TRACE"] = "0"
    _lib.reload_env()
