"""Developer tool (GPU box, 1 GPU): where the gap between the device-resident step and the host-buffer (e2e) step of
the fused-exchange search comes from, on one 8-GPU-sized shard (1.25M x 1024 fp32) with world = 1."""
import ctypes, os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from generativeaiexamples_b200 import B200IndexFlat, _lib, synth

dev = torch.device("cuda", 0)
n, d, k = 1_250_000, 1024, 10
ix = B200IndexFlat(d, 0, reserve_rows=n)
for c, lo in enumerate(range(0, n, 500_000)):
    ix.add(synth.corpus_chunk_device(c, min(500_000, n - lo), d, dev))
L = _lib.load()
xb = torch.zeros(int(L.b200r_exchange_buffer_bytes()), dtype=torch.uint8, device=dev)
ptrs = (ctypes.c_void_p * 1)(xb.data_ptr())
Q = torch.randn(64, d, device=dev)
Q /= Q.norm(dim=1, keepdim=True)
Qh = Q.cpu().numpy()
D = torch.empty(1, k, device=dev)
I = torch.empty(1, k, dtype=torch.int64, device=dev)
Dh = np.empty((1, k), dtype=np.float32)
Ih = np.empty((1, k), dtype=np.int64)
epoch = [0]
stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def dev_step(i):
    epoch[0] += 1
    _lib.check(L.b200r_search_exchange(ix._h, ctypes.c_void_p(Q[i % 64: i % 64 + 1].data_ptr()), 1, k, _lib.DEVICE, ptrs, 1, 0,
                                       epoch[0], ctypes.c_void_p(D.data_ptr()), ctypes.c_void_p(I.data_ptr()), _lib.DEVICE, stream))


qptr = [Qh[i: i + 1].ctypes.data_as(ctypes.c_void_p) for i in range(64)]
dptr, iptr = Dh.ctypes.data_as(ctypes.c_void_p), Ih.ctypes.data_as(ctypes.c_void_p)


def host_step(i):
    epoch[0] += 1
    _lib.check(L.b200r_search_exchange(ix._h, qptr[i % 64], 1, k, _lib.HOST, ptrs, 1, 0, epoch[0], dptr, iptr, _lib.HOST, None))


def plain_host_step(i):
    ix.search(Qh[i % 64: i % 64 + 1], k)


def loop(fn, reps=300, warm=30):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(reps):
        fn(i)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e6


rec = {}
for rnd in range(2):
    rec[f"device_resident_us_{rnd}"] = round(loop(dev_step), 1)
    rec[f"host_buffers_c_call_us_{rnd}"] = round(loop(host_step), 1)
    rec[f"host_buffers_python_search_us_{rnd}"] = round(loop(plain_host_step), 1)
os.environ["B200R_ZERO_COPY_OUT"] = "0"
_lib.reload_env()
rec["host_buffers_c_call_d2h_copy_us"] = round(loop(host_step), 1)
print(json.dumps(rec))
