"""Developer scratch tool (GPU box): three 256-query searches (mode bf16|shadow|f32, k from argv) for an ncu launch list."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat, synth
dev = torch.device("cuda", 0)
n, d = 2_000_000, 1024
mode, k = sys.argv[1], int(sys.argv[2])
kw = {"storage": "bf16"} if mode == "bf16" else ({"shadow_bf16": True} if mode == "shadow" else {})
ix = B200IndexFlat(d, 0, reserve_rows=n, **kw)
for c in range(4): ix.add(synth.corpus_chunk_device(c, 500_000, d, dev))
q = torch.randn(256, d, device=dev); q /= q.norm(dim=1, keepdim=True)
for _ in range(3): ix.search(q, k)
torch.cuda.synchronize(); print("done", ix.stats()["n_gemm_fallbacks"])
