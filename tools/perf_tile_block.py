"""Developer tool (GPU box): the scan's static tile assignment -- B consecutive 64 KB tiles per CTA and round
(B200R_SCAN_TILE_BLOCK) against the cyclic walk (B = 1); answers must not change."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat, synth
from generativeaiexamples_b200 import _lib

dev = torch.device("cuda", 0)
d, k = 1024, 10
out = []
for n in (1_250_000, 10_000_000):
    ix = B200IndexFlat(d, 0, reserve_rows=n)
    for lo in range(0, n, 1_000_000):
        ix.add(synth.corpus_chunk_device(lo // 1_000_000, min(1_000_000, n - lo), d, dev))
    Q = torch.randn(64, d, device=dev)
    Q /= Q.norm(dim=1, keepdim=True)
    D = torch.empty(1, k, device=dev)
    I = torch.empty(1, k, dtype=torch.int64, device=dev)
    ref = None
    for B in ((1, 4, 1, 4, 1, 4, 1, 4) if len(sys.argv) > 1 and sys.argv[1] == "ab" else (1, 2, 4, 16, 64, 1)):
        os.environ["B200R_SCAN_TILE_BLOCK"] = str(B)
        _lib.reload_env()
        Dq, Iq = ix.search(Q[:4], k)
        got = (Dq.cpu(), Iq.cpu())
        if ref is None:
            ref = got
        assert torch.equal(got[0], ref[0]) and torch.equal(got[1], ref[1]), f"answers changed at B={B}"
        for i in range(10):
            ix.search(Q[i:i + 1], k, D=D, I=I)
        torch.cuda.synchronize()
        best = 1e9
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(40):
                ix.search(Q[i:i + 1], k, D=D, I=I)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 40 * 1e3)
        rec = {"rows": n, "tile_block": B, "us": round(best, 1), "GBps": round(n * 4096 / best / 1e3, 1)}
        out.append(rec)
        print(rec, flush=True)
    ix.close()
os.makedirs("gpurun_out", exist_ok=True)
with open("gpurun_out/tile_block.jsonl", "w") as f:
    for r in out:
        f.write(json.dumps(r) + "\n")
