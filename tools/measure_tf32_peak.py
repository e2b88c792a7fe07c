"""Measures the dense TF32 GEMM throughput of this GPU with cuBLAS (torch.matmul, allow_tf32),
the same way the driver measured the bf16 figure in MEASURED_PEAKS.json (8192^3, best of 10 for
the burst number, back to back for ~3 s for the sustained one).  Output: one JSON line; the
committed copy lives in profiles/tf32_peak.json and is the denominator bench.py uses for the
tensor-pipe fraction of the fp32 (tf32-filter) GEMM path."""
import json
import time

import torch

torch.backends.cuda.matmul.allow_tf32 = True
n = 8192
a = torch.randn(n, n, device="cuda")
b = torch.randn(n, n, device="cuda")
for _ in range(3):
    a @ b
torch.cuda.synchronize()
best = 1e9
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    a @ b
    e1.record()
    torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1))
flops = 2.0 * n ** 3
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.time()
iters = 0
e0.record()
while time.time() - t0 < 3.0:
    for _ in range(10):
        a @ b
    iters += 10
    torch.cuda.synchronize()
e1.record()
torch.cuda.synchronize()
sustained = flops * iters / (e0.elapsed_time(e1) * 1e-3) / 1e12
print(json.dumps({"tf32_tflops": flops / (best * 1e-3) / 1e12, "tf32_tflops_sustained": sustained,
                  "how": "torch.matmul fp32 8192^3 with allow_tf32 (cuBLAS), best of 10 / 3 s back to back",
                  "gpu": torch.cuda.get_device_name(0)}))
