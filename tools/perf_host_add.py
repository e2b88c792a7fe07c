"""Developer scratch tool (GPU box): host-buffer ingest throughput (pageable numpy -> pinned ring -> H2D -> K1)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from generativeaiexamples_b200 import B200IndexFlat
n, d = 500_000, 1024
x = np.random.default_rng(0).standard_normal((n, d), dtype=np.float32)
for rep in range(3):
    ix = B200IndexFlat(d, 1, reserve_rows=n)
    t0 = time.perf_counter(); ix.add(x); dt = time.perf_counter() - t0
    print(f"host add {n} x {d}: {dt*1e3:.1f} ms  {n*d*4/dt/1e9:.2f} GB/s  {n/dt/1e6:.2f} M rows/s", flush=True)
    ix.close()
ix = B200IndexFlat(d, 1)
t0 = time.perf_counter()
for i in range(0, 20000, 200): ix.add(x[i:i+200])
dt = time.perf_counter() - t0
print(f"100 adds of 200 rows (chunk-per-document pattern, growing buffer): {dt*1e3:.1f} ms total, {dt*10:.2f} ms per add")
