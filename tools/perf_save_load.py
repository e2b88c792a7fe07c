import sys, time, os
sys.path.insert(0, "/root/repo")
import torch
from generativeaiexamples_b200 import B200IndexFlat, synth
dev = torch.device("cuda", 0)
for storage, d in (("bf16", 1024), ("f32", 1022)):
    ix = B200IndexFlat(d, 0, storage=storage, reserve_rows=1_000_000)
    ix.add(torch.randn(1_000_000, d, device=dev))
    torch.cuda.synchronize()
    path = "/dev/shm/x.faiss" if os.path.isdir("/dev/shm") else "/tmp/x.faiss"
    t0 = time.perf_counter(); ix.write(path); t1 = time.perf_counter()
    again = B200IndexFlat.read(path, storage=storage); t2 = time.perf_counter()
    print(storage, d, "save %.2f s, load %.2f s, %.1f GB" % (t1 - t0, t2 - t1, os.path.getsize(path) / 1e9), again.ntotal)
    os.unlink(path)
