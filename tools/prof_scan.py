"""Developer tool (GPU box): a handful of batch-1 scan searches for an ncu capture.  argv: rows [nq] [k]."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat, _lib, synth
dev = torch.device("cuda", 0)
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_250_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 1
k = int(sys.argv[3]) if len(sys.argv) > 3 else 10
d = 1024
ix = B200IndexFlat(d, 0, reserve_rows=n, search_path=_lib.PATH_SCAN)
for c, lo in enumerate(range(0, n, 500_000)):
    ix.add(synth.corpus_chunk_device(c, min(500_000, n - lo), d, dev))
Q = torch.randn(8 * nq, d, device=dev)
Q /= Q.norm(dim=1, keepdim=True)
for i in range(6):
    ix.search(Q[i * nq:(i + 1) * nq], k)
torch.cuda.synchronize()
print("ok", ix.stats()["last_path"])
