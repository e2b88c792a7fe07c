"""Developer scratch tool (GPU box): see DESIGN.md section 4 for the numbers it produced."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, os, sys
from generativeaiexamples_b200 import B200IndexFlat, synth
from generativeaiexamples_b200 import _lib as _b200r_lib
dev = torch.device("cuda",0)
n, d = 4_000_000, 1024
def build(storage):
    ix = B200IndexFlat(d, 0, storage=storage, reserve_rows=n)
    for c in range(n//1_000_000): ix.add(synth.corpus_chunk_device(c, 1_000_000, d, dev))
    return ix
def perf(ix, nq, k, storage, tag):
    Q = torch.randn(nq, d, device=dev); Q /= Q.norm(dim=1,keepdim=True)
    torch.cuda.synchronize()
    for _ in range(2): ix.search(Q, k)
    torch.cuda.synchronize()
    e0,e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): ix.search(Q, k)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)/5
    es = 2 if storage=="bf16" else 4
    print(f"{tag} nq={nq} k={k} {storage}: {ms:.3f} ms/batch  {2*nq*n*d/ms/1e9:.1f} TFLOP/s  {n*d*es/ms/1e6:.0f} GB/s corpus  fb={ix.stats()['n_gemm_fallbacks']}", flush=True)
for storage in ("f32","bf16"):
    ix = build(storage)
    for dbg in ("0","1","2"):
        os.environ["B200R_GEMM_DEBUG_NOSELECT"]=dbg; _b200r_lib.reload_env()
        for nq,k in ((32,10),(128,10),(256,10)):
            perf(ix, nq, k, storage, f"noselect={dbg}")
    ix.close()
os.environ["B200R_GEMM_DEBUG_NOSELECT"]="0"; _b200r_lib.reload_env()
for storage in ("f32","bf16"):
    ix = build(storage)
    for nq,k in ((64,100),(256,100),(1024,100),(1024,10)):
        perf(ix, nq, k, storage, "k100")
    ix.close()
