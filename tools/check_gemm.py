"""Developer scratch tool (GPU box): see DESIGN.md section 4 for the numbers it produced."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np, os, sys, time
from generativeaiexamples_b200 import B200IndexFlat, _lib
dev = torch.device("cuda",0)
torch.manual_seed(0)
def run(n, d, nq, k, metric, storage="f32"):
    X = torch.randn(n, d, device=dev); X /= X.norm(dim=1, keepdim=True)
    Q = torch.randn(nq, d, device=dev); Q[::2] = X[torch.randint(0,n,(len(Q[::2]),),device=dev)] + 0.3*Q[::2]/d**0.5; Q /= Q.norm(dim=1,keepdim=True)
    ix = B200IndexFlat(d, metric, storage=storage, search_path=_lib.PATH_GEMM)
    ix.add(X); torch.cuda.synchronize()
    D, I = ix.search(Q, k); torch.cuda.synchronize()
    Xs = X.to(torch.bfloat16).float() if storage=="bf16" else X
    S = Q.double() @ Xs.double().T
    if metric == 1:
        S = (Q.double()**2).sum(1,keepdim=True) + (Xs.double()**2).sum(1)[None,:] - 2*S
        Dt, It = torch.topk(S, k, dim=1, largest=False)
    else:
        Dt, It = torch.topk(S, k, dim=1)
    st = ix.stats()
    same = (I == It).float().mean().item()
    err = (D.double()-Dt).abs().max().item()
    print(f"n={n} d={d} nq={nq} k={k} metric={metric} {storage}: id-match={same:.4f} max|dD|={err:.2e} fallbacks={st['n_gemm_fallbacks']} path={st['last_path']}", flush=True)
    if same < 0.999 or err > 1e-5:
        bad = (I != It).any(dim=1).nonzero().flatten()[:3]
        for b in bad.tolist():
            print("  q", b, "got", I[b].tolist(), D[b].tolist()); print("     want", It[b].tolist(), Dt[b].tolist())
    ix.close()
cases = [(50000,1024,64,10,0),(50000,1024,200,10,0),(100000,1024,256,10,0),(50000,1024,64,10,1),(30000,384,33,4,1),(20000,100,21,7,0),(60000,1024,130,100,0),(1000,64,20,10,0),(100,1024,256,10,0)]
for c in cases: run(*c)
run(50000,1024,64,10,0,"bf16"); run(50000,1024,200,100,1,"bf16")
# scale test: 4M x 1024, nq=256
def perf(n, nq, k, storage="f32"):
    from generativeaiexamples_b200 import synth
    d=1024
    ix = B200IndexFlat(d, 0, storage=storage, reserve_rows=n)
    for c in range(n//1_000_000):
        ix.add(synth.corpus_chunk_device(c, 1_000_000, d, dev))
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
    print(f"perf n={n} nq={nq} k={k} {storage}: {ms:.3f} ms/batch  {2*nq*n*d/ms/1e9:.1f} TFLOP/s  {n*d*es/ms/1e6:.0f} GB/s corpus  fallbacks={ix.stats()['n_gemm_fallbacks']}", flush=True)
    ix.close()
perf(4_000_000, 256, 10); perf(4_000_000, 128, 10); perf(4_000_000, 64, 10); perf(4_000_000, 32, 10)
perf(4_000_000, 256, 10, "bf16"); perf(4_000_000, 256, 100, "bf16"); perf(4_000_000, 1024, 100, "bf16")
