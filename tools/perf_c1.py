"""Developer tool (GPU box): where a C1-sized search (10k x 1024 fp32 rows, one query, k=4) spends its time --
wall clock per call of the host-buffer search, of the threshold variant and of a device-buffer search, and the
device time of the kernels alone (events)."""
import os, sys, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from generativeaiexamples_b200 import B200IndexFlat, _lib

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 10_000
d, k = 1024, 4
rng = np.random.default_rng(0)
x = rng.standard_normal((n, d), dtype=np.float32)
x /= np.linalg.norm(x, axis=1, keepdims=True)
q = x[:64] + 0.05 * rng.standard_normal((64, d), dtype=np.float32)
ix = B200IndexFlat(d, 1)
ix.add(x)
qd = torch.from_numpy(q).cuda()


def wall(fn, reps=2000, warm=200):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(reps):
        fn(i)
    torch.cuda.synchronize()
    return round((time.perf_counter() - t0) / reps * 1e6, 2)


rec = {"rows": n}
rec["host_search_us"] = wall(lambda i: ix.search(q[i % 64: i % 64 + 1], k))
rec["host_threshold_us"] = wall(lambda i: ix.search_threshold(q[i % 64: i % 64 + 1], k, "EUCLIDEAN_DISTANCE", 0.25))
D = torch.empty((1, k), device="cuda")
I = torch.empty((1, k), dtype=torch.int64, device="cuda")
rec["device_async_us"] = wall(lambda i: ix.search(qd[i % 64: i % 64 + 1], k, D=D, I=I))


def dev_sync(i):
    ix.search(qd[i % 64: i % 64 + 1], k, D=D, I=I)
    torch.cuda.synchronize()


rec["device_synced_us"] = wall(dev_sync)
# query already on the device, results to the host: what the host-to-device copy of the query costs
import ctypes
Dh = np.empty((1, k), dtype=np.float32)
Ih = np.empty((1, k), dtype=np.int64)
L = _lib.load()


def dev_q_host_out(i):
    _lib.check(L.b200r_search(ix._h, ctypes.c_void_p(qd[i % 64: i % 64 + 1].data_ptr()), 1, k, _lib.DEVICE,
                              _lib.addr(Dh), _lib.addr(Ih), _lib.HOST, None))


torch.cuda.synchronize()
rec["device_query_host_results_us"] = wall(dev_q_host_out)
os.environ["B200R_ZERO_COPY_OUT"] = "0"
_lib.reload_env()
rec["host_search_with_d2h_copy_us"] = wall(lambda i: ix.search(q[i % 64: i % 64 + 1], k))
os.environ["B200R_ZERO_COPY_OUT"] = "1"
_lib.reload_env()
ix.set_timing(True)
ks = []
for i in range(200):
    ix.search(q[i % 64: i % 64 + 1], k)
    ks.append(ix.stats()["last_search_ms"] * 1000)
ix.set_timing(False)
rec["kernels_device_us_p50"] = round(sorted(ks)[100], 2)
# the Python call alone (argument checks, ctypes marshalling): an index with no rows launches nothing heavy
empty = B200IndexFlat(d, 1)
rec["host_search_empty_index_us"] = wall(lambda i: empty.search(q[i % 64: i % 64 + 1], k))
print(json.dumps(rec))
