"""Secondary benchmark (GPU box, N GPUs under torch.distributed.run): the two BASELINE.json configs that are
defined on several GPUs, at full size, through ShardedFlatIndex (row shards + NCCL allgather + K4 select).

  C4  100M x 1024 bf16 (12.5M rows = 25.6 GB per GPU on 8), IP, k=100, nq=1024
  C5  5M x 1024 fp32, L2 with ingest-time and query normalisation, k=10, nq=64

Rows are generated on the device chunk by chunk (synth.corpus_chunk_device), queries are perturbed copies of
rows of rank 0's shard so the expected first hit is known.  One JSON line per config on rank 0.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/bench_configs_multi.py [c4] [c5]
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from generativeaiexamples_b200 import synth
from generativeaiexamples_b200.sharded import ShardedFlatIndex, shard_bounds

D = 1024
world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local_rank = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local_rank)
dev = torch.device("cuda", local_rank)
json_fd = os.dup(1)
os.dup2(2, 1)  # native libraries' stdout chatter -> stderr
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def emit(rec):
    if rank == 0:
        os.write(json_fd, (json.dumps(rec) + "\n").encode())


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn, steps, warm):
    for _ in range(warm):
        fn()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def build(n_total, metric, **kw):
    lo, hi = shard_bounds(n_total, world, rank)
    index = ShardedFlatIndex(D, metric, device=local_rank, reserve_rows=hi - lo, **kw)
    probe = None
    row = lo
    while row < hi:
        c = row // synth.CHUNK_ROWS
        c_lo, c_hi = c * synth.CHUNK_ROWS, min((c + 1) * synth.CHUNK_ROWS, n_total)
        x = synth.corpus_chunk_device(c, c_hi - c_lo, D, dev)
        part = x[max(lo, c_lo) - c_lo: min(hi, c_hi) - c_lo]
        if probe is None:
            probe = part[:2048].clone()
        index.local.add(part)
        row = min(hi, c_hi)
        del x, part
    index._sync_counts()
    torch.cuda.synchronize()
    assert index.ntotal == n_total, (index.ntotal, n_total)
    src = probe if rank == 0 else torch.zeros_like(probe)
    if world > 1:
        dist.broadcast(src, 0)
    return index, src  # src = global rows 0..2047


def queries(src, nq, noise):
    g = torch.Generator(device=dev)
    g.manual_seed(synth.QUERY_SEED + 7)
    q = torch.randn((nq, D), generator=g, device=dev)
    planted = torch.full((nq,), -1, dtype=torch.int64)
    for i in range(0, nq, 2):
        j = (i * 61) % src.shape[0]
        q[i] = src[j] + noise * q[i] / D ** 0.5
        planted[i] = j
    return q, planted


def c4():
    n = 100_000_000 if world >= 8 else 12_500_000 * world  # 12.5M rows per GPU either way
    index, src = build(n, 0, storage="bf16")
    q, planted = queries(src, 1024, 0.3)
    q /= q.norm(dim=1, keepdim=True)
    Dg, Ig = index.search(q, 100)
    torch.cuda.synchronize()
    I0 = Ig[:, 0].cpu()
    ok = bool((I0[::2] == planted[::2]).all())
    srt = bool((Dg[:, 1:] <= Dg[:, :-1]).all())
    ms = timed(lambda: index.search(q, 100), 5, 2)
    flops = 2.0 * 1024 * n * D
    emit({"config": f"C4: {n} x 1024 bf16 IP k=100 nq=1024 on {world} GPU(s) (12.5M rows = 25.6 GB per GPU)",
          "n_gpus": world, "ms_per_batch": ms, "qps": 1024 * 1000 / ms, "tflops_aggregate": flops / ms / 1e9,
          "tflops_per_gpu": flops / ms / 1e9 / world, "planted_first_hits_ok": ok, "sorted": srt,
          "fallbacks_rank0": index.local.stats()["n_gemm_fallbacks"],
          "note": "recall@100 vs fp32 truth is measured by tools/bench_configs.py c4 on a 1M-row slice (0.997)"})
    assert ok and srt
    index.local.close()


def c5():
    n = 5_000_000
    index, src = build(n, 1, normalize_on_add=True, normalize_queries=True)
    q, planted = queries(src, 64, 0.1)
    q = q * 3.0  # un-normalised queries: the index normalises them
    Dg, Ig = index.search(q, 10)
    torch.cuda.synchronize()
    I0 = Ig[:, 0].cpu()
    ok = bool((I0[::2] == planted[::2]).all())
    ms = timed(lambda: index.search(q, 10), 50, 5)
    emit({"config": f"C5: 5M x 1024 fp32 L2, normalised at ingest and on query, k=10 nq=64 on {world} GPU(s)",
          "n_gpus": world, "ms_per_batch": ms, "qps": 64 * 1000 / ms, "hbm_gbs_aggregate": n * 4096 / ms / 1e6,
          "hbm_gbs_per_gpu": n * 4096 / ms / 1e6 / world, "planted_first_hits_ok": ok,
          "fallbacks_rank0": index.local.stats()["n_gemm_fallbacks"]})
    assert ok
    index.local.close()


if __name__ == "__main__":
    which = [a for a in sys.argv[1:] if not a.startswith("-")] or ["c5", "c4"]
    for w in which:
        {"c4": c4, "c5": c5}[w]()
        torch.cuda.empty_cache()
    if world > 1:
        dist.destroy_process_group()
