"""Developer scratch tool (GPU box): parity sweep over unusual dimensions on both search paths."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from generativeaiexamples_b200 import B200IndexFlat
dev = torch.device("cuda", 0)
bad = 0
for d in (1, 3, 8, 12, 33, 64, 768, 1536, 3072, 4096, 8192, 12000):
    n = 20011 if d <= 4096 else 3001
    for metric in (0, 1):
        for storage in ("f32", "bf16"):
            torch.manual_seed(d * 7 + metric)
            X = torch.randn(n, d, device=dev)
            Q = torch.randn(40, d, device=dev); Q[::2] = X[:20] + 0.05 * Q[::2]
            try:
                ix = B200IndexFlat(d, metric, storage=storage)
                ix.add(X)
                Xs = X.to(torch.bfloat16).float() if storage == "bf16" else X
                for nq, k in ((1, 10), (3, 5), (40, 10), (9, 130)):
                    D, I = ix.search(Q[:nq].contiguous(), k)
                    S = Q[:nq].double() @ Xs.double().T
                    if metric == 1:
                        S = (Q[:nq].double() ** 2).sum(1, keepdim=True) + (Xs.double() ** 2).sum(1)[None] - 2 * S
                        Dt, It = torch.topk(S, k, dim=1, largest=False)
                    else:
                        Dt, It = torch.topk(S, k, dim=1)
                    err = ((D.double() - Dt).abs() / Dt.abs().clamp_min(1)).max().item()
                    match = (I == It).float().mean().item()
                    if err > 1e-5 or match < 0.99:
                        bad += 1
                        print(f"BAD d={d} metric={metric} {storage} nq={nq} k={k}: relerr={err:.2e} idmatch={match:.3f}", flush=True)
                ix.close()
            except Exception as e:
                print(f"d={d} metric={metric} {storage}: {type(e).__name__}: {str(e)[:150]}", flush=True)
print("done, bad =", bad)
