"""Developer scratch tool (GPU box): one C5-shaped search (fp32, L2, k=10, nq=64) on 2M rows for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat, synth
dev = torch.device("cuda", 0)
n, d = 2_000_000, 1024
ix = B200IndexFlat(d, 1, reserve_rows=n)
for c in range(2): ix.add(synth.corpus_chunk_device(c, 1_000_000, d, dev))
q = torch.randn(64, d, device=dev); q /= q.norm(dim=1, keepdim=True)
for _ in range(2): ix.search(q, 10)
torch.cuda.synchronize(); print("done")
