"""Developer tool (GPU box, >= 2 GPUs): per-call cost of the single-process multi-GPU search on a small
corpus (the scan itself is ~35 us per shard, so what is left is issue + exchange overhead), for host and
device buffers, fused / event route, worker threads on / off; plus a bare cross-device event ping-pong."""
import os, sys, json, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from generativeaiexamples_b200 import B200IndexFlat, _lib

G = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
rows_per_gpu = int(float(sys.argv[2])) if len(sys.argv) > 2 else 100_000
d, n, k = 1024, rows_per_gpu * G, 10
x = torch.randn(min(n, 1_000_000), d, device="cuda:0")
q = torch.randn(64, d, device="cuda:0")
qh = q.cpu().numpy()

# bare cross-device event hop: stream on GPU0 -> stream on GPU1 -> back
s0, s1 = torch.cuda.Stream(0), torch.cuda.Stream(1 % G)
e0, e1 = torch.cuda.Event(), torch.cuda.Event()
a0 = torch.zeros(8, device="cuda:0")
a1 = torch.zeros(8, device=f"cuda:{1 % G}")
torch.cuda.synchronize(0)
t0 = time.perf_counter()
for i in range(200):
    with torch.cuda.stream(s0):
        a0.add_(1)
        e0.record(s0)
    s1.wait_event(e0)
    with torch.cuda.stream(s1):
        a1.add_(1)
        e1.record(s1)
    s0.wait_event(e1)
s0.synchronize()
s1.synchronize()
print(json.dumps({"event_round_trip_us": round((time.perf_counter() - t0) / 200 * 1e6, 1)}), flush=True)

for fused in ("1", "0"):
    for workers in ("1", "0"):
        os.environ["B200R_MULTI_FUSED"] = fused
        os.environ["B200R_MULTI_WORKERS"] = workers
        _lib.reload_env()
        ix = B200IndexFlat(d, 0, devices=list(range(G)))
        for lo in range(0, n, x.shape[0]):
            ix.add(x[: min(x.shape[0], n - lo)])
        rec = {"gpus": G, "rows_per_gpu": rows_per_gpu, "fused": fused, "workers": workers}
        for name, arg in (("host", qh), ("device", q)):
            for i in range(20):
                ix.search(arg[i: i + 1], k)
            torch.cuda.synchronize(0)
            t0 = time.perf_counter()
            for i in range(300):
                r = ix.search(arg[i % 64: i % 64 + 1], k)
            for g in range(G):
                torch.cuda.synchronize(g)
            rec[name + "_us"] = round((time.perf_counter() - t0) / 300 * 1e6, 1)
        t0 = time.perf_counter()
        for i in range(100):
            r = ix.search(q[i % 64: i % 64 + 1], k)
            torch.cuda.synchronize(0)
        rec["device_synced_each_us"] = round((time.perf_counter() - t0) / 100 * 1e6, 1)
        ix.set_timing(True)
        ix.search(qh[:1], k)
        st = ix.stats()
        rec["kernel_us"], rec["exchange_us"] = round(st["last_kernel_us"], 1), round(st["last_exchange_us"], 1)
        print(json.dumps(rec), flush=True)
        ix.close()
one = B200IndexFlat(d, 0)
for lo in range(0, rows_per_gpu, x.shape[0]):
    one.add(x[: min(x.shape[0], rows_per_gpu - lo)])
for name, arg in (("host", qh), ("device", q)):
    for i in range(20):
        one.search(arg[i: i + 1], k)
    torch.cuda.synchronize(0)
    t0 = time.perf_counter()
    for i in range(300):
        one.search(arg[i % 64: i % 64 + 1], k)
    torch.cuda.synchronize(0)
    print(json.dumps({"single_gpu_shard_sized": name, "us": round((time.perf_counter() - t0) / 300 * 1e6, 1)}), flush=True)
