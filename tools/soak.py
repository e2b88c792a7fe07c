"""Developer scratch tool (GPU box): create / add / search / remove / destroy cycles; free device memory must return to
its starting value (no leaked allocations), and results stay correct throughout.  argv[1] = "multi": every other cycle
builds the multi-GPU index (two shards on GPU 0, or GPUs 0 and 1 when there are two) with its worker threads, gather
buffers, segment tables and re-ranking scratch (threshold + MMR calls included)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from generativeaiexamples_b200 import B200IndexFlat
torch.cuda.init()
rng = np.random.default_rng(0)
d = 256
x = rng.standard_normal((30000, d), dtype=np.float32)
q = x[:40] + 0.01 * rng.standard_normal((40, d), dtype=np.float32)
free0 = None
multi = len(sys.argv) > 1 and sys.argv[1] == "multi"
devs = [0, 1] if torch.cuda.device_count() >= 2 else [0, 0]
for cyc in range(120):
    kw = [dict(), dict(shadow_bf16=True), dict(storage="bf16")][cyc % 3]
    if multi and cyc % 2:
        kw = dict(kw, devices=devs)
    ix = B200IndexFlat(d, cyc % 2, **kw)
    for lo in range(0, 30000, 7000):
        ix.add(x[lo:lo + 7000])
    for nq, k in ((1, 10), (4, 5), (40, 10), (7, 200)):
        D, I = ix.search(q[:nq], k)
        assert (I[:, 0] == np.arange(nq)).all(), (cyc, nq, k, I[:, 0])
    if multi:
        R, Dt, It, cnt, oor = ix.search_threshold(q[:3], 5, "EUCLIDEAN_DISTANCE" if cyc % 2 else "MAX_INNER_PRODUCT", 0.1)
        assert (It[:, 0] == np.arange(3)).all() or cnt.min() == 0
        assert ix.mmr_select(q[0], np.arange(50, dtype=np.int64), 5).size == 5
    ix.remove_ids(np.arange(100, 5000, 3, dtype=np.int64))
    D, I = ix.search(q[:8], 10)
    assert (I[:, 0] == np.arange(8)).all()
    ix.close()
    del ix
    if cyc == 5:
        torch.cuda.synchronize(); free0 = sum(torch.cuda.mem_get_info(g)[0] for g in set(devs))
torch.cuda.synchronize()
free1 = sum(torch.cuda.mem_get_info(g)[0] for g in set(devs))
print("free after warm-up cycles: %.1f MB, at the end: %.1f MB, delta %.1f MB" % (free0 / 2**20, free1 / 2**20, (free0 - free1) / 2**20))
assert abs(free0 - free1) < 64 * 2**20, "device memory leak"
print("soak ok")
