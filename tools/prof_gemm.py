"""Developer scratch tool (GPU box): see DESIGN.md section 4 for the numbers it produced."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, sys
from generativeaiexamples_b200 import B200IndexFlat, synth
dev = torch.device("cuda",0)
n, d = 2_000_000, 1024
nq = int(sys.argv[1]); storage = sys.argv[2] if len(sys.argv) > 2 else "f32"
ix = B200IndexFlat(d, 0, storage=storage, reserve_rows=n)
for c in range(n//1_000_000): ix.add(synth.corpus_chunk_device(c, 1_000_000, d, dev))
Q = torch.randn(nq, d, device=dev); Q /= Q.norm(dim=1,keepdim=True)
for _ in range(3): ix.search(Q, 10)
torch.cuda.synchronize()
print("ok", ix.stats()["last_path"])
