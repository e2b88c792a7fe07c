import os, sys
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from generativeaiexamples_b200 import B200IndexFlat, _lib
rng = np.random.default_rng(0)
for n in (10_000, 100_000):
    x = rng.standard_normal((n, 1024), dtype=np.float32)
    ix = B200IndexFlat(1024, 1); ix.add(x)
    for i in range(20): ix.search(x[i:i+1], 4)
    os.environ["B200R_SCAN_TRACE"] = "1"; _lib.reload_env()
    for i in range(4): ix.search(x[i:i+1], 4)
    os.environ["B200R_SCAN_TRACE"] = "0"; _lib.reload_env()
