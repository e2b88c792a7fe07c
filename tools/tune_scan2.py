"""Developer scratch tool (GPU box): KPL=1 vs KPL=4 scan variants at several shard sizes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat
dev = torch.device("cuda", 0)
d = 1024
for n in (1_250_000, 2_500_000, 10_000_000):
    ix = B200IndexFlat(d, 0, reserve_rows=n)
    for lo in range(0, n, 1_000_000): ix.add(torch.randn(min(1_000_000, n - lo), d, device=dev))
    Q = torch.randn(64, d, device=dev)
    for k in (10, 33, 10, 33):
        D = torch.empty(1, k, device=dev); I = torch.empty(1, k, dtype=torch.int64, device=dev)
        for i in range(10): ix.search(Q[i:i+1], k, D=D, I=I)
        torch.cuda.synchronize()
        best = 1e9
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(40): ix.search(Q[i:i+1], k, D=D, I=I)
            e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1) / 40 * 1e3)
        print(f"n={n:>8} k={k:>3}: {best:8.1f} us  {n*4096/best/1e6:7.1f} GB/s", flush=True)
    ix.close()
