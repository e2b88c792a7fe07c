"""Developer scratch tool (GPU box): five single-query searches on a small shadow index via the GEMM route (launch list)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["B200R_GEMM_MIN_BYTES"] = "1"
import torch
from generativeaiexamples_b200 import B200IndexFlat, synth
from generativeaiexamples_b200 import _lib as _b200r_lib
dev = torch.device("cuda", 0)
n, d = int(float(sys.argv[1])) if len(sys.argv) > 1 else 20000, 1024
ix = B200IndexFlat(d, 0, reserve_rows=n, shadow_bf16=True)
ix.add(synth.corpus_chunk_device(0, n, d, dev))
q = torch.randn(1, d, device=dev); q /= q.norm(dim=1, keepdim=True)
for _ in range(5): ix.search(q, 10)
torch.cuda.synchronize(); print("done")
