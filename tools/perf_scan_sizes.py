"""Developer scratch tool (GPU box): scan-path latency vs shard size, to separate the fixed cost
(launch, pipeline ramp, per-CTA and last-CTA merges) from the streaming part."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat
dev = torch.device("cuda", 0)
d = 1024
for n in (1184, 9472, 100_000, 1_250_000, 5_000_000):
    for k in (10, 100):
        ix = B200IndexFlat(d, 0, reserve_rows=n)
        for lo in range(0, n, 1_000_000):
            m = min(1_000_000, n - lo)
            ix.add(torch.randn(m, d, device=dev))
        Q = torch.randn(64, d, device=dev)
        torch.cuda.synchronize()
        D = torch.empty(1, k, device=dev); I = torch.empty(1, k, dtype=torch.int64, device=dev)
        for i in range(10): ix.search(Q[i:i+1], k, D=D, I=I)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(50): ix.search(Q[i:i+1], k, D=D, I=I)
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 50 * 1e3
        print(f"n={n:>8} k={k:>3}: {us:8.1f} us/query   stream-only at 7.2 TB/s = {n*4096/7.2e6:8.1f} us   overhead = {us - n*4096/7.2e6:6.1f} us", flush=True)
        ix.close()
