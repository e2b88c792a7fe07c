"""Developer scratch tool (GPU box): scan-kernel tuning sweep over stage size / depth / L2 policy."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from generativeaiexamples_b200 import B200IndexFlat
from generativeaiexamples_b200 import _lib as _b200r_lib
dev = torch.device("cuda", 0)
d = 1024
def run(ix, n, k, tag):
    Q = torch.randn(64, d, device=dev)
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
    print(f"{tag:32s} n={n:>8} k={k}: {best:8.1f} us  {n*4096/best/1e6:7.1f} GB/s", flush=True)
for n in (1_250_000, 10_000_000):
    ix = B200IndexFlat(d, 0, reserve_rows=n)
    for lo in range(0, n, 1_000_000): ix.add(torch.randn(min(1_000_000, n - lo), d, device=dev))
    for kb, st in ((32, 6), (16, 12), (64, 3), (32, 4), (24, 8), (48, 4), (8, 16)):
        os.environ["B200R_SCAN_STAGE_KB"] = str(kb); _b200r_lib.reload_env(); os.environ["B200R_SCAN_STAGES"] = str(st); _b200r_lib.reload_env()
        for ev in ("1", "0"):
            os.environ["B200R_EVICT_FIRST"] = ev; _b200r_lib.reload_env()
            run(ix, n, 10, f"stage={kb}KB x{st} evict_first={ev}")
    ix.close()
