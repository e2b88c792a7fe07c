#!/usr/bin/env python
"""bench.py -- headline benchmark: exact top-10 QPS over 10M x 1024 fp32 embeddings.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--devices N]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
           --master-port P bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json metric; config C3 of SURVEY.md section 8): corpus of 10M x 1024
fp32 unit vectors (40.96 GB -- fits one 180 GB B200), inner product, top-10, batch-1
queries.  A "step" is one batch-1 query.  Two ways to use N GPUs, same answers:

* `--gpus N` under torch.distributed.run: one process per GPU, rows in contiguous blocks, each
  rank searches its shard, candidates merged by one NCCL allgather + the K4 select
  (generativeaiexamples_b200/sharded.py).  STRONG scaling: total rows fixed.
* `--devices N` in ONE process: the index itself spans the GPUs (b200r_config.n_devices,
  csrc/multi.cu) -- what the reference's single-process chain server would use.

Data is synthetic (torch Philox on the device, one seed per 1M-row chunk;
generativeaiexamples_b200/synth.py); every step uses a different query and the per-GPU corpus
(>= 5 GB) is far larger than the 126 MB L2, so no flush is needed.

Before anything is timed, EVERY result path that is about to be timed is checked against a float64
brute force computed chunk by chunk on the device during ingest (`parity_check`): 64 batch-1 queries
-- half of them planted next to rows spread over every GPU's shard -- and 32 queries of the
batch-256 block.  A mismatch ends the run with a non-zero exit code.

JSON line: `value` = device-resident QPS (query already in HBM, results left in HBM),
`e2e` = the same through the public API with HOST query/result buffers (H2D + D2H inside
the timed region), `roofline` = scan kernel vs measured HBM bandwidth, `cpu_baseline` = the
faiss-1.7.4 restatement (oracle/) timed on this box's host cores on a bounded sample,
`batch256` = the batch-256 variant of the same search, `c1` / `c2` / `c5` = the other BASELINE
configs that fit this run.

`--impl reference` times only the CPU restatement (the reference's engine, faiss-cpu 1.7.4,
is not installable here: DESIGN.md "Oracle") on the same config and prints the same line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_TOTAL = 10_000_000
DIM = 1024
TOPK = 10
CHUNK = 1_000_000
POOL = 64          # batch-1 query pool (every second one planted next to a corpus row)
N_CHECK_256 = 32   # queries of the batch-256 block that are checked against the float64 truth
METRIC_NAME = "top10_qps_10Mx1024_fp32_ip_batch1"
WORKLOAD = "C3: 10M x 1024 fp32 corpus, inner product, top-10, batch-1 queries, rows sharded over N GPUs"
TOL = 1e-5


# ----------------------------------------------------------------------------- helpers
def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def _tensor_peaks():
    """(tf32, bf16) dense TFLOP/s denominators: bf16 from MEASURED_PEAKS.json (sustained), tf32 from
    profiles/tf32_peak.json (tools/measure_tf32_peak.py on this pool), else nominal 1.1 PF."""
    bf16, tf32, src = 1400.0, 1100.0, "nominal"
    try:
        bf16 = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"])
    except Exception:
        pass
    try:
        tf32 = float(json.load(open(os.path.join(ROOT, "profiles", "tf32_peak.json")))["tf32_tflops_sustained"])
        src = "measured (profiles/tf32_peak.json, cuBLAS tf32 8192^3 sustained)"
    except Exception:
        pass
    return tf32, bf16, src


def _traffic():
    """dram bytes per launch of the scan kernel from the committed ncu capture, if any."""
    p = os.path.join(ROOT, "profiles", "scan_traffic.json")
    if os.path.exists(p):
        try:
            return json.load(open(p))
        except Exception:
            return None
    return None


class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap,power.limit")

    def __init__(self, gpu_index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "20", "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        sm, mx, pw, pl, reasons = [], [], [], [], set()
        for r in rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                try:
                    pw.append(float(r[3]))
                    pl.append(float(r[8]))
                except (ValueError, IndexError):
                    pass
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.strip().lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        if pw and pl:
            pw.sort()
            out.update(power_w=pw[len(pw) // 2], power_limit_w=max(pl))
        return out


# ----------------------------------------------------------------------------- CPU arm
def _mem_available_bytes():
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable:"):
                return int(ln.split()[1]) * 1024
    except OSError:
        pass
    return 0


def cpu_reference(steps: int, warmup: int, sample_rows: int = 0, budget_s: float = 25.0, full_if_possible: bool = False):
    """faiss-1.7.4 restatement on the host cores, nq=1.

    Rows scanned: the full 10M x 1024 corpus when `full_if_possible` and the host
    has >= 1.5 x 41 GB of free RAM; else a slice of at least 1M rows, with QPS scaled linearly to 10M rows
    (an exhaustive scan is exactly linear in N) and `extrapolated: true` in the record.  The OpenMP team
    is set explicitly to the cores this process may run on, so a launcher's OMP_NUM_THREADS=1
    (torch.distributed.run exports it) cannot turn "all cores" into one core."""
    import numpy as np

    from oracle import FlatRef, load_flat_ref

    lib = load_flat_ref()
    cores = len(os.sched_getaffinity(0))
    lib.oracle_set_threads(cores)
    threads = int(lib.oracle_max_threads())
    need_full = int(1.5 * N_TOTAL * DIM * 4)
    if sample_rows <= 0:
        sample_rows = N_TOTAL if (full_if_possible and _mem_available_bytes() >= need_full) else 1_000_000
    sample_rows = min(sample_rows, N_TOTAL)
    extrapolated = sample_rows < N_TOTAL

    # rows: seeded Gaussian unit vectors like synth.py's (numpy PCG64, one seed per 62,500-row piece so
    # that the pieces can be filled by all cores at once: numpy's generators release the GIL)
    from concurrent.futures import ThreadPoolExecutor

    t_gen = time.perf_counter()
    x = np.empty((sample_rows, DIM), dtype=np.float32)
    piece = 62_500

    def fill(c):
        lo, hi = c * piece, min(sample_rows, (c + 1) * piece)
        r = np.random.Generator(np.random.PCG64(1234 + c))
        blk = r.standard_normal((hi - lo, DIM), dtype=np.float32)
        blk /= np.linalg.norm(blk, axis=1, keepdims=True)
        x[lo:hi] = blk

    n_pieces = (sample_rows + piece - 1) // piece
    with ThreadPoolExecutor(max(1, min(cores, n_pieces))) as ex:
        list(ex.map(fill, range(n_pieces)))
    gen_s = time.perf_counter() - t_gen
    rng = np.random.Generator(np.random.PCG64(4321))
    q = rng.standard_normal((64, DIM), dtype=np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)

    def run(fn, n_warm, n_steps, budget):
        for i in range(n_warm):
            fn(q[i % 64: i % 64 + 1])
        t0 = time.perf_counter()
        done = 0
        for i in range(n_steps):
            fn(q[i % 64: i % 64 + 1])
            done += 1
            if time.perf_counter() - t0 > budget:
                break
        dt = time.perf_counter() - t0
        return done / dt, done, dt

    scale = sample_rows / N_TOTAL
    # Opportunistic: a real faiss on the box (never seen on this pool) is timed as the
    # reference proper and checked against the restatement; its absence is the expected case.
    live = {"faiss_importable": False}
    try:
        ref_dir = os.path.join(ROOT, "baseline", "_ref")
        if os.path.isdir(ref_dir) and ref_dir not in sys.path:
            sys.path.append(ref_dir)
        import faiss  # noqa: F401

        fidx = faiss.IndexFlatIP(DIM)
        fidx.add(x)
        qps_faiss, nf, _ = run(lambda v: fidx.search(v, TOPK), min(warmup, 2), steps, budget_s * 0.3)
        Df, If = fidx.search(q[:8], TOPK)
        Dr, Ir = FlatRef.knn(0, q[:8], x, TOPK)
        live = {"faiss_importable": True, "faiss_version": getattr(faiss, "__version__", "?"),
                "value": qps_faiss * scale, "steps": nf, "ids_equal_to_restatement": bool((If == Ir).all()),
                "max_abs_score_diff_vs_restatement": float(np.abs(Df - Dr).max())}
    except ImportError:
        pass
    # (i) faithful: faiss 1.7.4 runs nq < 20 on ONE thread (OpenMP over queries only)
    qps_faithful, n1, _ = run(lambda v: FlatRef.knn(0, v, x, TOPK), 1 if warmup else 0, max(1, min(steps, 4)),
                              budget_s * 0.4)
    # (ii) the same scan row-parallel over all cores (not a faiss code path; stronger baseline)
    qps_all, n2, dt_all = run(lambda v: FlatRef.knn_allcores(0, v, x, TOPK), min(warmup, 3), steps, budget_s * 0.6)
    cpu_model = ""
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                cpu_model = ln.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    common = {
        "unit": "queries/s", "cores": threads, "threads_used": threads, "host_cores_visible": cores,
        "cpu_model": cpu_model, "sample_rows": sample_rows, "extrapolated": extrapolated,
        "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS", "unset"),
        "steps_timed": n2, "seconds_timed": dt_all, "generate_s": round(gen_s, 2),
        "faithful_1thread_value": qps_faithful * scale, "faithful_1thread_steps": n1,
        "host_mem_available_gb": round(_mem_available_bytes() / 2 ** 30, 1),
        "live_faiss": live,
    }
    what = (f"nq=1 top-{TOPK} scans over " +
            (f"the full {N_TOTAL} x {DIM} fp32 corpus" if not extrapolated else
             f"a {sample_rows} x {DIM} fp32 slice, scaled linearly to {N_TOTAL} rows (extrapolated)") +
            f" ({n2} all-core steps on {threads} threads, {n1} single-thread steps)")
    if live["faiss_importable"]:  # the real thing: report it, keep the restatement's numbers beside it
        return dict(common, value=live["value"], kind="reference", port_allcores_value=qps_all * scale,
                    sample="faiss.IndexFlatIP.search; " + what)
    return dict(common, value=qps_all * scale, kind="port",
                blas="OpenBLAS bundled with numpy (not on the nq=1 path)", sample=what,
                note="value = all-cores row-parallel scan; faithful_1thread_value = faiss 1.7.4's own nq<20 path (1 thread)")


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0  # rank 0 alone runs the CPU arm
    cb = cpu_reference(args.steps, args.warmup, sample_rows=args.cpu_sample_rows, budget_s=90.0, full_if_possible=True)
    line = {
        "impl": "reference", "metric": METRIC_NAME, "value": cb["value"], "unit": "queries/s", "n_gpus": args.gpus,
        "steps": cb["steps_timed"], "warmup": args.warmup,
        # what one step of THIS run took (a scan of sample_rows rows); value is scaled to 10M rows when extrapolated
        "ms_per_step": 1000.0 * cb["seconds_timed"] / max(1, cb["steps_timed"]),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "rows": N_TOTAL, "dim": DIM, "k": TOPK, "nq": 1,
                   "sample_rows": cb["sample_rows"], "extrapolated": cb["extrapolated"]},
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ----------------------------------------------------------------------------- float64 checker
class TruthTopK:
    """Running exact top-k' (float64, lowest id first among equals) of `q` against the rows shown to
    update(): the checker of this benchmark.  Inner product (l2=False) or squared L2."""

    def __init__(self, q, kk, l2=False):
        import torch

        self.q = q.double()
        self.kk, self.l2 = kk, l2
        nq = q.shape[0]
        self.s = torch.empty((nq, 0), dtype=torch.float64, device=q.device)
        self.i = torch.empty((nq, 0), dtype=torch.int64, device=q.device)

    def update(self, x, row0, sub=250_000):
        import torch

        for a in range(0, x.shape[0], sub):
            xs = x[a: a + sub].double()
            s = self.q @ xs.T
            if self.l2:
                s = (self.q * self.q).sum(1, keepdim=True) + (xs * xs).sum(1)[None, :] - 2.0 * s
                s = -s  # keep "larger is better" inside
            ids = torch.arange(row0 + a, row0 + a + xs.shape[0], device=x.device).expand(s.shape[0], -1)
            cs, ci = torch.cat([self.s, s], 1), torch.cat([self.i, ids], 1)
            ts, ti = torch.topk(cs, min(self.kk, cs.shape[1]), dim=1)
            self.s, self.i = ts, torch.gather(ci, 1, ti)
            del xs, s, cs, ci

    def merged(self, world, dist):
        """(scores, ids) as numpy over the union of every rank's rows."""
        import torch

        s, i = self.s, self.i
        if world > 1:
            pad = self.kk - s.shape[1]
            if pad > 0:
                s = torch.cat([s, torch.full((s.shape[0], pad), -float("inf"), dtype=s.dtype, device=s.device)], 1)
                i = torch.cat([i, torch.full((i.shape[0], pad), -1, dtype=i.dtype, device=i.device)], 1)
            gs = [torch.empty_like(s) for _ in range(world)]
            gi = [torch.empty_like(i) for _ in range(world)]
            dist.all_gather(gs, s.contiguous())
            dist.all_gather(gi, i.contiguous())
            cs, ci = torch.cat(gs, 1), torch.cat(gi, 1)
            # order by (score desc, id asc): sort by id first, then a stable sort by score
            o = torch.argsort(ci, dim=1, stable=True)
            cs, ci = torch.gather(cs, 1, o), torch.gather(ci, 1, o)
            o = torch.argsort(cs, dim=1, descending=True, stable=True)[:, : self.kk]
            s, i = torch.gather(cs, 1, o), torch.gather(ci, 1, o)
        s = -s if self.l2 else s
        return s.cpu().numpy(), i.cpu().numpy()


def parity_errors(D, I, Dt, It, what):
    """North-star rule: scores within 1e-5 (abs-or-rel) of the float64 truth, ids identical except
    among ties within that tolerance.  Returns a list of messages (empty = pass)."""
    import numpy as np

    D, Dt = np.asarray(D, dtype=np.float64), np.asarray(Dt, dtype=np.float64)
    I, It = np.asarray(I), np.asarray(It)
    nq, k = I.shape
    errs = []
    close = lambda a, b: np.abs(a - b) <= TOL * np.maximum(1.0, np.abs(b))  # noqa: E731
    bad = ~close(D, Dt[:, :k])
    if bad.any():
        errs.append(f"{what}: scores differ beyond {TOL}: max |diff| {np.abs(D - Dt[:, :k]).max():.3e}")
    for qi in range(nq):
        if np.array_equal(I[qi], It[qi, :k]):
            continue
        truth = {int(i): s for i, s in zip(It[qi], Dt[qi])}
        if len(set(I[qi].tolist())) != k:
            errs.append(f"{what}: duplicate ids for query {qi}: {I[qi].tolist()}")
            continue
        for r in range(k):
            got, want = int(I[qi, r]), int(It[qi, r])
            if got == want:
                continue
            ref = truth.get(got, None)
            tie = close(ref, Dt[qi, r]) if ref is not None else close(D[qi, r], Dt[qi, k - 1])
            if not tie:
                errs.append(f"{what}: query {qi} rank {r}: id {got} instead of {want} and not a tie")
    return errs


# ----------------------------------------------------------------------------- GPU arm
def main_gpu(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    from generativeaiexamples_b200 import B200IndexFlat, _lib, synth
    from generativeaiexamples_b200.flat_index import METRIC_INNER_PRODUCT, METRIC_L2
    from generativeaiexamples_b200.sharded import ShardedFlatIndex, shard_bounds

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    n_dev = args.devices  # > 1: ONE process, the index spans the GPUs
    if world == 1 and args.gpus > 1 and n_dev <= 1:
        n_dev = args.gpus  # `python bench.py --gpus N` without torch.distributed.run: the one-process form
    if n_dev > 1 and world > 1:
        raise SystemExit("--devices is the single-process mode: do not combine it with torch.distributed.run")
    if n_dev > torch.cuda.device_count():
        raise SystemExit(f"--devices {n_dev} but only {torch.cuda.device_count()} GPUs are visible")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    # stdout carries exactly one JSON line.  Native libraries write there too (NCCL prints its
    # version banner on stdout whatever NCCL_DEBUG_FILE says), so file descriptor 1 points at stderr
    # for the whole run and the JSON line goes to the saved descriptor at the end.
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus or n_dev > 1, \
        f"--gpus {args.gpus} but WORLD_SIZE={world} (launch with torch.distributed.run, or use --devices N)"
    n_gpus = n_dev if n_dev > 1 else world

    n_total = args.rows
    lo, hi = shard_bounds(n_total, world, rank)
    # the optional bf16 shadow (+50% HBM) only serves the `batch256_shadow` line; it is switched off
    # for every other measurement, which therefore streams the fp32 rows as BASELINE.json's metric says
    use_shadow = args.batch256_steps > 0 and not args.no_shadow
    if n_dev > 1:
        index = B200IndexFlat(DIM, METRIC_INNER_PRODUCT, devices=list(range(n_dev)), reserve_rows=n_total,
                              shadow_bf16=use_shadow)
        local = index
    else:
        index = ShardedFlatIndex(DIM, METRIC_INNER_PRODUCT, device=local_rank, reserve_rows=hi - lo,
                                 shadow_bf16=use_shadow)
        local = index.local
    if use_shadow:
        local.set_shadow_filter(False)

    # ---- queries, fixed before ingest so that the float64 checker can run alongside it ---------------
    # batch-1 pool: 64 queries, every second one planted next to a corpus row; the planted rows are
    # spread over the WHOLE corpus, i.e. over every GPU's shard.  Every rank regenerates the chunks
    # that hold them (the generator is seeded per chunk), so all ranks build identical queries.
    g = torch.Generator(device=dev)
    g.manual_seed(synth.QUERY_SEED)
    Q = torch.randn((POOL, DIM), generator=g, device=dev)
    planted = np.full(POOL, -1, dtype=np.int64)
    n_plant = POOL // 2
    for j in range(n_plant):
        planted[2 * j] = (((j * 13) % n_plant) * n_total) // n_plant + (j * 4099) % max(1, n_total // n_plant)
    by_chunk = {}
    for i in range(0, POOL, 2):
        by_chunk.setdefault(int(planted[i]) // CHUNK, []).append(i)
    for c, qs in sorted(by_chunk.items()):
        xc = synth.corpus_chunk_device(c, min(CHUNK, n_total - c * CHUNK), DIM, dev)
        for i in qs:
            Q[i] = xc[int(planted[i]) - c * CHUNK] + 0.1 * Q[i] / DIM ** 0.5
        del xc
    Q /= Q.norm(dim=1, keepdim=True)
    g.manual_seed(synth.QUERY_SEED + 1)
    Q256 = torch.randn((256, DIM), generator=g, device=dev)
    Q256 /= Q256.norm(dim=1, keepdim=True)
    Qh = Q.cpu().numpy()
    truth = TruthTopK(torch.cat([Q, Q256[:N_CHECK_256]]), TOPK + 8)

    # ---- ingest (untimed here; K1 throughput is reported separately) -------------------
    t_ing = time.perf_counter()
    row = lo
    while row < hi:
        c = row // CHUNK
        c_lo, c_hi = c * CHUNK, min((c + 1) * CHUNK, n_total)
        x = synth.corpus_chunk_device(c, c_hi - c_lo, DIM, dev)
        part = x[max(lo, c_lo) - c_lo: min(hi, c_hi) - c_lo]
        local.add(part)
        truth.update(part, max(lo, c_lo))
        row = min(hi, c_hi)
        del x, part
    if n_dev <= 1:
        index._sync_counts()
    torch.cuda.synchronize()
    ingest_s = time.perf_counter() - t_ing
    assert index.ntotal == n_total, (index.ntotal, n_total)
    Dt_all, It_all = truth.merged(world, dist)
    Dt, It = Dt_all[:POOL], It_all[:POOL]
    Dt256, It256 = Dt_all[POOL:], It_all[POOL:]
    del truth
    shard_rows = (n_total + n_gpus - 1) // n_gpus
    shards_hit = sorted({int(It[i, 0]) // shard_rows for i in range(0, POOL, 2)})

    def barrier():
        if world > 1:
            dist.barrier()
        for d_ in range(max(1, n_dev)):
            torch.cuda.synchronize(d_ if n_dev > 1 else local_rank)

    def timed(fn, steps, warmup):
        for i in range(warmup):
            fn(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for i in range(steps):
            fn(warmup + i)
        e1.record()
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        barrier()
        ms = e0.elapsed_time(e1)
        # host-blocking calls (e2e) are bounded by wall clock, device-async ones by events
        ms = max(ms, 0.0)
        t = torch.tensor([ms, wall * 1000.0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1])

    results = {}

    def dev_step(i):
        results["last"] = index.search(Q[i % POOL: i % POOL + 1], TOPK)

    def e2e_step(i):
        results["last_h"] = index.search(Qh[i % POOL: i % POOL + 1], TOPK)

    # ---- parity gate: every path that is timed below, against the float64 truth, before any timing -----
    parity = {"queries": 0, "paths": [], "ranks_with_planted_hits": len(shards_hit), "n_gpus": n_gpus,
              "tolerance": TOL, "truth": "float64 brute force on the device, chunk by chunk during ingest, "
                                         "all-gathered over the ranks"}
    errors = []

    def check_batch1(search, name, host):
        Dd = np.empty((POOL, TOPK), dtype=np.float32)
        Ii = np.empty((POOL, TOPK), dtype=np.int64)
        for i in range(POOL):
            d_, i_ = search((Qh if host else Q)[i: i + 1], TOPK)
            if not host:
                torch.cuda.synchronize()
                d_, i_ = d_.cpu().numpy(), i_.cpu().numpy()
            Dd[i], Ii[i] = d_[0], i_[0]
        e = parity_errors(Dd, Ii, Dt, It, name)
        for i in range(0, POOL, 2):  # the planted row must come first, from whichever shard holds it
            if Ii[i, 0] != planted[i] or Dd[i, 0] < 0.98:
                e.append(f"{name}: planted row {planted[i]} of query {i} not found first: {Ii[i, :3]} {Dd[i, :3]}")
        parity["paths"].append({"path": name, "queries": POOL, "ok": not e})
        parity["queries"] += POOL
        errors.extend(e)

    # clocks and throttle reasons are sampled from here to the end of the batch-1 measurements: the parity gate
    # runs the same searches as the timed regions (on 8 GPUs the device-timed region alone lasts ~15 ms, less than
    # nvidia-smi needs to start), so every sample is taken under the load that is being measured
    sampler = ClockSampler(local_rank) if rank == 0 else None
    check_batch1(index.search, "batch-1, device buffers", host=False)
    check_batch1(index.search, "batch-1, host buffers (e2e)", host=True)
    # the other exchange over the same shards (NCCL allgather + K4 when the fused NVLink exchange is the
    # one in use, and the other way round): checked too, and timed as a secondary line
    alt, alt_name, main_name = None, None, None
    if world > 1:
        main_fused = index._exchange() is not None
        main_name = "fused NVLink exchange" if main_fused else "NCCL allgather + K4"
        keep = os.environ.get("B200R_FUSED_EXCHANGE")
        os.environ["B200R_FUSED_EXCHANGE"] = "0" if main_fused else "1"
        alt = ShardedFlatIndex(DIM, METRIC_INNER_PRODUCT, device=local_rank, local_index=index.local)
        alt._sync_counts()
        alt_name = "NCCL allgather + K4" if main_fused else "fused NVLink exchange"
        if not main_fused and alt._exchange() is None:
            alt = None
        if keep is None:
            os.environ.pop("B200R_FUSED_EXCHANGE")
        else:
            os.environ["B200R_FUSED_EXCHANGE"] = keep
        if alt is not None:
            check_batch1(alt.search, "batch-1, " + alt_name + ", device buffers", host=False)
            check_batch1(alt.search, "batch-1, " + alt_name + ", host buffers", host=True)

    launches0 = local.stats()["n_kernel_launches"]
    # one GPU, one process: the step IS the scan launch, so every launch of the timed region gets its own event pair
    # on the launching stream and roofline.kernel_ms comes from the timed region itself
    step_pairs = None
    timed_step = dev_step
    if world == 1 and n_dev <= 1:
        step_pairs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                      for _ in range(args.steps + args.warmup)]

        def timed_step(i):
            e0, e1 = step_pairs[i]
            e0.record()
            dev_step(i)
            e1.record()

    ms_dev, _ = timed(timed_step, args.steps, args.warmup)
    launches = local.stats()["n_kernel_launches"] - launches0
    launches = launches * args.steps // (args.steps + args.warmup)
    if world > 1 and not getattr(index, "_xchg", None):
        launches += args.steps  # K4 merge, one per step (the fused route's select kernel is counted by the library)
    _, ms_e2e_wall = timed(e2e_step, args.steps, args.warmup)
    # per-call latency distribution of the host-buffer path (wall clock around each call)
    lat = []
    for i in range(min(args.steps, 200)):
        t0 = time.perf_counter()
        e2e_step(i)
        lat.append((time.perf_counter() - t0) * 1e6)
    lat.sort()
    e2e_lat = {"p50_us": lat[len(lat) // 2], "p99_us": lat[min(len(lat) - 1, int(len(lat) * 0.99))], "calls": len(lat)}
    # (the sampler covers the device-timed region, the end-to-end region and the per-call latency loop: on 8 GPUs
    # the device-timed region alone is a few milliseconds, shorter than nvidia-smi's sampling period)
    clocks = sampler.stop() if sampler else None
    if clocks is not None:
        clocks["window"] = "parity gate + device-timed + end-to-end + per-call latency loop (batch-1 searches throughout)"
    alt_line = None
    if alt is not None:
        ms_f, _ = timed(lambda i: results.__setitem__("last_f", alt.search(Q[i % POOL: i % POOL + 1], TOPK)),
                        args.steps, args.warmup)
        _, ms_fw = timed(lambda i: results.__setitem__("last_f", alt.search(Qh[i % POOL: i % POOL + 1], TOPK)),
                         args.steps, args.warmup)
        alt_line = {"exchange": alt_name, "value": args.steps * 1000.0 / ms_f, "unit": "queries/s",
                    "ms_per_step": ms_f / args.steps, "e2e": {"value": args.steps * 1000.0 / ms_fw, "unit": "queries/s"},
                    "note": "the same shards through the other exchange (B200R_FUSED_EXCHANGE=0 selects NCCL): fused = the "
                            "scan kernels store their top-k into every rank over NVLink and the select kernel waits on "
                            "flags; NCCL = one all_gather_into_tensor of 120 B per rank + the K4 select kernel"}

    # ---- scan kernel alone: CUDA events on the launching stream, live -----------------------
    kms, xus = [], []
    if n_dev > 1:
        # one process over several GPUs: the scans run on the shards' own streams, the library times them
        local.set_timing(True)
        for i in range(args.steps):
            local.search(Q[i % POOL: i % POOL + 1], TOPK)
            st_ = local.stats()
            kms.append(st_["last_kernel_us"] / 1000.0)
            xus.append(st_["last_exchange_us"])
        local.set_timing(False)
    elif step_pairs is not None:
        kms = [e0.elapsed_time(e1) for e0, e1 in step_pairs[args.warmup:]]
        xus = [0.0]
    else:
        # this GPU's shard alone (no exchange), launches back to back as in the timed region, one event pair per
        # launch on the stream the kernel is launched on (torch's current stream: device-buffer searches run there)
        pairs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for i in range(3):
            local.search(Q[i % POOL: i % POOL + 1], TOPK)
        for i, (e0, e1) in enumerate(pairs):
            e0.record()
            local.search(Q[i % POOL: i % POOL + 1], TOPK)
            e1.record()
        torch.cuda.synchronize()
        kms = [e0.elapsed_time(e1) for e0, e1 in pairs]
        xus = [0.0]
    kernel_ms = float(np.mean(kms))
    exchange_us = float(np.mean(xus))
    kt = torch.tensor([kernel_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(kt, op=dist.ReduceOp.MAX)
    kernel_ms = float(kt[0])
    max_rows = shard_rows
    algo_bytes = max_rows * DIM * 4 + DIM * 4 + TOPK * 12
    peak, peak_src = _peaks()
    achieved = algo_bytes / (kernel_ms * 1e-3) / 1e9
    traffic = _traffic()

    # ---- batch-256 variant (secondary) ------------------------------------------------------
    b256 = None
    if args.batch256_steps > 0:
        def b256_step(i):
            results["b256"] = index.search(Q256, TOPK)

        b256_step(0)
        torch.cuda.synchronize()
        D_f, I_f = results["b256"]
        D_f, I_f = D_f.clone(), I_f.clone()
        e = parity_errors(D_f[:N_CHECK_256].cpu().numpy(), I_f[:N_CHECK_256].cpu().numpy(), Dt256, It256, "batch-256")
        parity["paths"].append({"path": "batch-256, device buffers", "queries": N_CHECK_256, "ok": not e})
        parity["queries"] += N_CHECK_256
        errors.extend(e)
        s256 = ClockSampler(local_rank) if rank == 0 else None
        ms256, _ = timed(b256_step, args.batch256_steps, 3)
        clocks256 = s256.stop() if s256 else None
        per = ms256 / args.batch256_steps
        flops = 2.0 * 256 * n_total * DIM
        tf32_peak, _, tf32_src = _tensor_peaks()
        per_gpu_tflops = flops / n_gpus / (per * 1e-3) / 1e12
        st_ = local.stats()
        b256 = {"value": 256 * 1000.0 / per, "unit": "queries/s", "ms_per_batch": per, "steps": args.batch256_steps,
                "tensor_frac_tf32": per_gpu_tflops / tf32_peak, "tensor_peak_tflops": tf32_peak,
                "tensor_peak_source": tf32_src,
                "path": "scan x64 passes" if st_["last_path"] == 1
                else "tcgen05 tf32 GEMM filter + exact fp32 re-score", "gemm_fallbacks": st_["n_gemm_fallbacks"],
                "achieved_tflops": flops / (per * 1e-3) / 1e12,
                "hbm_frac": (max_rows * DIM * 4 / (per * 1e-3) / 1e9) / peak,
                "note": "crossover config (128 flop/B): both fractions reported; flops = 2*nq*N*d, bytes = N*d*4",
                "clocks": clocks256}
        if use_shadow:
            # same batch, filter pass on the bf16 shadow of the rows; the fp32 re-score makes the answers identical
            local.set_shadow_filter(True)
            fb0 = local.stats()["n_gemm_fallbacks"]
            ms256s, _ = timed(b256_step, args.batch256_steps, 3)
            D_s, I_s = results["b256"]
            same = bool(torch.equal(I_s, I_f) and torch.equal(D_s, D_f))
            local.set_shadow_filter(False)
            pers = ms256s / args.batch256_steps
            bf16_peak = _tensor_peaks()[1]
            b256["shadow_bf16"] = {
                "value": 256 * 1000.0 / pers, "unit": "queries/s", "ms_per_batch": pers,
                "achieved_tflops": flops / (pers * 1e-3) / 1e12,
                "tensor_frac_bf16": flops / n_gpus / (pers * 1e-3) / 1e12 / bf16_peak, "tensor_peak_tflops": bf16_peak,
                "hbm_frac": (max_rows * DIM * 2 / (pers * 1e-3) / 1e9) / peak,
                "identical_to_fp32_filter": same,
                "gemm_fallbacks": local.stats()["n_gemm_fallbacks"] - fb0,
                "note": "opt-in index mode (shadow_bf16=True, +50% HBM): filter streams N*d*2 bytes of bf16 rows at the "
                        "bf16 MMA rate, survivors are re-scored on the fp32 rows; not used for `value`/`e2e`"}
            if not same:
                errors.append("batch-256: shadow-filter results differ from the fp32-filter results")

    # ---- batch-1 on the bf16 shadow (opt-in index mode; secondary) ---------------------------------
    b1s = None
    if use_shadow:
        ref = [index.search(Q[i: i + 1], TOPK) for i in range(8)]
        ref = [(d_.clone(), i_.clone()) for d_, i_ in ref]
        local.set_shadow_filter(True)
        fb0 = local.stats()["n_gemm_fallbacks"]
        same = True
        for i in range(8):
            d_, i_ = index.search(Q[i: i + 1], TOPK)
            same = same and bool(torch.equal(i_, ref[i][1]) and torch.equal(d_, ref[i][0]))
        steps1 = max(20, args.steps // 2)
        ms_s, _ = timed(dev_step, steps1, 5)
        _, ms_s_wall = timed(e2e_step, steps1, 5)
        fbs = local.stats()["n_gemm_fallbacks"] - fb0
        local.set_shadow_filter(False)
        b1s = {"value": steps1 * 1000.0 / ms_s, "unit": "queries/s", "ms_per_step": ms_s / steps1,
               "e2e": {"value": steps1 * 1000.0 / ms_s_wall, "unit": "queries/s"},
               "hbm_frac_bf16_bytes": (max_rows * DIM * 2 / (ms_s / steps1 * 1e-3) / 1e9) / peak,
               "identical_to_fp32_scan": same, "path": "tcgen05 bf16 GEMM filter on the shadow + exact fp32 re-score",
               "gemm_fallbacks": fbs, "steps": steps1,
               "note": "opt-in index mode (shadow_bf16=True, +50% HBM): the tcgen05 filter streams N*d*2 bytes of bf16 "
                       "rows for the K'=112 best approximate scores, which are re-scored on the fp32 rows and checked "
                       "for containment; not used for `value`/`e2e`/`roofline`"}
        if not same:
            errors.append("batch-1: shadow-filter results differ from the fp32-scan results")

    # ---- service view: 64 concurrent single-query callers through the coalescer (N=1 only) -------
    conc = None
    if n_gpus == 1 and args.concurrent > 0:
        from concurrent.futures import ThreadPoolExecutor

        from generativeaiexamples_b200.batching import SearchCoalescer

        co = SearchCoalescer(local)
        nreq = args.concurrent * 24

        def one(i):
            return co.search(Qh[i % POOL: i % POOL + 1], TOPK)

        with ThreadPoolExecutor(args.concurrent) as ex:
            list(ex.map(one, range(args.concurrent)))  # warm-up
            co.n_requests = co.n_launches = 0
            t0 = time.perf_counter()
            list(ex.map(one, range(nreq)))
            dt = time.perf_counter() - t0
        conc = {"value": nreq / dt, "unit": "queries/s", "callers": args.concurrent, "requests": nreq,
                "launches": co.n_launches,
                "note": "threads each calling an nq=1 host search (the pattern of faiss_vdb_service.py:169-174); "
                        "SearchCoalescer merges the queued ones into one launch (wall clock, host buffers)"}

    # ---- the other BASELINE configs that fit this run (secondary lines) ---------------------------------
    extra = {}
    if not args.no_configs:
        extra = other_configs(args, world, rank, local_rank, n_dev, dev, dist, timed, errors, parity, peak)

    parity["ok"] = not errors
    okt = torch.tensor([0 if errors else 1], dtype=torch.int32, device=dev)
    if world > 1:
        dist.all_reduce(okt, op=dist.ReduceOp.MIN)
    if int(okt[0]) != 1:
        for e in errors[:20]:
            print("PARITY FAILURE:", e, file=sys.stderr)
        if world > 1:
            dist.destroy_process_group()
        raise SystemExit("bench: results differ from the float64 truth (see stderr); no number is reported")

    if rank == 0:
        qps = args.steps * 1000.0 / ms_dev
        e2e_qps = args.steps * 1000.0 / ms_e2e_wall
        cb = cpu_reference(20, 2, sample_rows=args.cpu_sample_rows) if (n_gpus == 1 and not args.no_cpu_baseline) else None
        ncu_peak = 8180.0  # GB/s: ncu's own DRAM ceiling on this part (profiles/r01_scan_full_final_2M.csv: 7.02 TB/s = 85.8 %)
        if n_dev > 1:
            par = f"one process, index over {n_dev} GPUs (b200r_config.n_devices): per-shard scan, top-k stored into the " \
                  "primary GPU's gather buffer over NVLink, translating select kernel"
            api = "B200IndexFlat(devices=[...]).search(numpy) = b200r_search(HOST query, HOST results) on the multi-GPU index"
        elif world > 1:
            par = f"row-shard x{world} + {main_name}"
            api = ("ShardedFlatIndex.search(numpy) = b200r_search_exchange(HOST query, HOST results): pinned H2D -> scan "
                   "-> NVLink stores into every rank -> waiting select -> one packed D2H") if getattr(index, "_xchg", None) \
                else ("ShardedFlatIndex.search(numpy): pinned H2D -> b200r_search(DEVICE) -> NCCL allgather -> "
                      "b200r_merge_topk -> one packed D2H")
        else:
            par, api = "single GPU", "B200IndexFlat.search(numpy) = b200r_search(HOST query, HOST results)"
        line = {
            "metric": METRIC_NAME, "value": qps, "unit": "queries/s", "n_gpus": n_gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD if n_total == N_TOTAL else f"REDUCED ({n_total} rows) " + WORKLOAD,
                       "rows": n_total, "rows_per_gpu": max_rows, "dim": DIM, "k": TOPK, "nq": 1,
                       "parallelism": par,
                       "l2_policy": "no flush: per-GPU corpus (>=5 GB) >> 126 MB L2, new query every step",
                       "ingest_s": round(ingest_s, 3)},
            "e2e": {"value": e2e_qps, "unit": "queries/s", "h2d_bytes_per_step": DIM * 4,
                    "d2h_bytes_per_step": TOPK * 12, "latency": e2e_lat, "api": api},
            "gpu_launches": int(launches),
            "parity_check": parity,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "frac_of_ncu_dram_peak": achieved / ncu_peak, "ncu_dram_peak": ncu_peak,
                         "traffic": (algo_bytes * traffic["traffic_over_algorithmic"]) if traffic else None,
                         "traffic_source": ("ncu dram__bytes_read+write per launch = %.4f x algorithmic bytes, measured on a "
                                            "%d-row shard (%s), scaled to this shard" %
                                            (traffic["traffic_over_algorithmic"], traffic.get("rows", 0),
                                             traffic["source"])) if traffic else None,
                         "peak_source": peak_src,
                         "kernel": "scan_topk_kernel<1,1,IP,f32>", "kernel_ms": kernel_ms,
                         # what a step costs beyond its slowest scan launch: exchange over NVLink + final select (+ the
                         # skew between the GPUs); one process: measured with events inside the library
                         "exchange_us": exchange_us if n_dev > 1 else None,
                         "step_minus_kernel_us": (ms_dev / args.steps - kernel_ms) * 1000.0,
                         "algorithmic_bytes_per_launch": algo_bytes,
                         "note": "kernel_ms = mean duration of the scan launch on this GPU's shard, one CUDA event pair per launch on the launching stream, launches back to back; "
                                 "frac can exceed 1 because `peak` is a copy (read+write) figure while the scan only reads: "
                                 "frac_of_ncu_dram_peak is the same rate against ncu's DRAM ceiling"},
            "cpu_baseline": cb,
            "batch256": b256,
            "batch1_shadow_bf16": b1s,
            "alt_exchange": alt_line,
            "concurrent_callers": conc,
            "clocks": clocks,
        }
        line.update(extra)
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()
    return 0


def other_configs(args, world, rank, local_rank, n_dev, dev, dist, timed, errors, parity, peak):
    """C2 (1M x 1024 IP nq=1), C5 (5M x 1024 L2, normalised at ingest and on the query, nq=64) and C1
    (10k rows through the LangChain surface, k=4, threshold 0.25) -- each checked against a float64
    truth before it is timed.  C2 and C1 are single-GPU configs (rank 0 of a 1-GPU run); C5 runs at
    every N like the headline."""
    import numpy as np
    import torch

    from generativeaiexamples_b200 import B200FlatVectorStore, B200IndexFlat, synth
    from generativeaiexamples_b200.flat_index import METRIC_INNER_PRODUCT, METRIC_L2
    from generativeaiexamples_b200.sharded import ShardedFlatIndex, shard_bounds

    out = {}
    n_gpus = n_dev if n_dev > 1 else world
    g = torch.Generator(device=dev)

    # ---- C5: 5M x 1024, L2 with ingest-time normalisation, top-10, batch 64 ---------------------------
    n5 = args.c5_rows
    if n5 > 0:
        lo, hi = shard_bounds(n5, world, rank)
        kw = dict(normalize_on_add=True, normalize_queries=True)
        if n_dev > 1:
            ix5 = B200IndexFlat(DIM, METRIC_L2, devices=list(range(n_dev)), reserve_rows=n5, **kw)
            loc5 = ix5
        else:
            ix5 = ShardedFlatIndex(DIM, METRIC_L2, device=local_rank, reserve_rows=hi - lo, **kw)
            loc5 = ix5.local
        g.manual_seed(synth.QUERY_SEED + 5)
        q5 = 3.0 * torch.randn((64, DIM), generator=g, device=dev)  # un-normalised on purpose
        # half of them near rows 0..31 of the corpus (its first chunk, generated exactly as in the ingest loop)
        xc = synth.corpus_chunk_device(100, min(CHUNK, n5), DIM, dev, normalize=False)
        q5[::2] = xc[:32] * 2.5 + 0.3 * q5[::2]
        del xc
        qn = q5 / q5.norm(dim=1, keepdim=True)
        tr = TruthTopK(qn, TOPK + 8, l2=True)
        row = lo
        while row < hi:
            c = row // CHUNK
            c_lo, c_hi = c * CHUNK, min((c + 1) * CHUNK, n5)
            x = 1.7 * synth.corpus_chunk_device(100 + c, c_hi - c_lo, DIM, dev, normalize=False)  # rag_service.py:194 normalises
            part = x[max(lo, c_lo) - c_lo: min(hi, c_hi) - c_lo]
            loc5.add(part)
            tr.update(part / part.norm(dim=1, keepdim=True), max(lo, c_lo))
            row = min(hi, c_hi)
            del x, part
        if n_dev <= 1:
            ix5._sync_counts()
        torch.cuda.synchronize()
        Dt5, It5 = tr.merged(world, dist)
        res = {}

        def c5_step(i):
            res["r"] = ix5.search(q5, TOPK)

        c5_step(0)
        torch.cuda.synchronize()
        e = parity_errors(res["r"][0].cpu().numpy(), res["r"][1].cpu().numpy(), Dt5, It5, "C5 nq=64")
        parity["paths"].append({"path": "C5: L2 + fused normalisation, nq=64", "queries": 64, "ok": not e})
        parity["queries"] += 64
        errors.extend(e)
        steps5 = max(10, args.batch256_steps)
        ms5, _ = timed(c5_step, steps5, 3)
        per = ms5 / steps5
        rows_gpu = (n5 + n_gpus - 1) // n_gpus
        out["c5"] = {"workload": f"C5: {n5} x 1024 fp32, squared L2, rows normalised at ingest (K1), queries normalised in "
                                 "the prep kernel, top-10, batch 64",
                     "value": 64 * 1000.0 / per, "unit": "queries/s", "ms_per_batch": per, "steps": steps5,
                     "hbm_frac": (rows_gpu * DIM * 4 / (per * 1e-3) / 1e9) / peak,
                     "gemm_fallbacks": loc5.stats()["n_gemm_fallbacks"],
                     "path": "tcgen05 tf32 GEMM filter + exact fp32 re-score" if loc5.stats()["last_path"] == 2 else "scan"}
        loc5.close() if hasattr(loc5, "close") else None
        del ix5, loc5
        torch.cuda.empty_cache()

    # ---- C4 (--c4): 12.5M x 1024 bf16 rows per GPU (100M on 8), top-100, batch 1024; recall@100 vs fp32 truth -------
    if args.c4:
        n4 = 12_500_000 * n_gpus
        lo, hi = shard_bounds(n4, world, rank)
        if n_dev > 1:
            ix4 = B200IndexFlat(DIM, METRIC_INNER_PRODUCT, devices=list(range(n_dev)), reserve_rows=n4, storage="bf16")
            loc4 = ix4
        else:
            ix4 = ShardedFlatIndex(DIM, METRIC_INNER_PRODUCT, device=local_rank, reserve_rows=hi - lo, storage="bf16")
            loc4 = ix4.local
        g.manual_seed(synth.QUERY_SEED + 4)
        q4 = torch.randn((1024, DIM), generator=g, device=dev)
        xc = synth.corpus_chunk_device(300, min(CHUNK, n4), DIM, dev)
        q4[::2] = xc[:512] + 0.3 * q4[::2] / DIM ** 0.5  # half of the queries near rows 0..511
        del xc
        q4 /= q4.norm(dim=1, keepdim=True)
        n_truth = 64
        tr = TruthTopK(q4[:n_truth], 100)  # on the fp32 ORIGINALS: the index stores them rounded to bf16
        row = lo
        while row < hi:
            c = row // CHUNK
            c_lo, c_hi = c * CHUNK, min((c + 1) * CHUNK, n4)
            x = synth.corpus_chunk_device(300 + c, c_hi - c_lo, DIM, dev)
            part = x[max(lo, c_lo) - c_lo: min(hi, c_hi) - c_lo]
            loc4.add(part)
            tr.update(part, max(lo, c_lo))
            row = min(hi, c_hi)
            del x, part
        if n_dev <= 1:
            ix4._sync_counts()
        torch.cuda.synchronize()
        _, It4 = tr.merged(world, dist)
        res = {}

        def c4_step(i):
            res["r"] = ix4.search(q4, 100)

        c4_step(0)
        torch.cuda.synchronize()
        I4 = res["r"][1][:n_truth].cpu().numpy()
        hits = sum(len(set(a.tolist()) & set(b.tolist())) for a, b in zip(I4, It4))
        recall = hits / float(n_truth * 100)
        planted_ok = bool((res["r"][1][:1024:2, 0].cpu() == torch.arange(512)).all())
        if recall < 0.95 or not planted_ok:
            errors.append(f"C4: recall@100 {recall:.4f} vs fp32 truth (planted rows first: {planted_ok})")
        parity["paths"].append({"path": "C4: bf16 corpus, recall@100 vs fp32 truth", "queries": n_truth,
                                "ok": recall >= 0.95 and planted_ok})
        parity["queries"] += n_truth
        ms4, _ = timed(c4_step, 5, 2)
        per = ms4 / 5
        flops = 2.0 * 1024 * n4 * DIM
        bf16_peak = _tensor_peaks()[1]
        out["c4"] = {"workload": f"C4: {n4} x 1024 bf16 rows (25.6 GB per GPU), inner product, top-100, batch 1024",
                     "value": 1024 * 1000.0 / per, "unit": "queries/s", "ms_per_batch": per, "steps": 5,
                     "recall_at_100_vs_fp32_truth": recall, "recall_queries": n_truth,
                     "achieved_tflops": flops / (per * 1e-3) / 1e12,
                     "tensor_frac_bf16": flops / n_gpus / (per * 1e-3) / 1e12 / bf16_peak, "tensor_peak_tflops": bf16_peak,
                     "gemm_fallbacks": loc4.stats()["n_gemm_fallbacks"]}
        loc4.close() if hasattr(loc4, "close") else None
        del ix4, loc4
        torch.cuda.empty_cache()

    if n_gpus > 1 or rank != 0:
        return out

    # ---- C2: 1M x 1024 fp32, inner product, top-10, batch 1, one GPU -----------------------------------
    n2 = 1_000_000
    ix2 = B200IndexFlat(DIM, METRIC_INNER_PRODUCT, device=local_rank, reserve_rows=n2)
    x = synth.corpus_chunk_device(200, n2, DIM, dev)
    g.manual_seed(synth.QUERY_SEED + 2)
    q2 = torch.randn((32, DIM), generator=g, device=dev)
    q2[::2] = x[1000:1016] + 0.1 * q2[::2] / DIM ** 0.5
    q2 /= q2.norm(dim=1, keepdim=True)
    ix2.add(x)
    tr = TruthTopK(q2, TOPK + 8)
    tr.update(x, 0)
    Dt2, It2 = tr.merged(1, None)
    del x
    D2 = np.empty((32, TOPK), dtype=np.float32)
    I2 = np.empty((32, TOPK), dtype=np.int64)
    q2h = q2.cpu().numpy()
    for i in range(32):
        D2[i], I2[i] = [a[0] for a in ix2.search(q2h[i: i + 1], TOPK)]
    e = parity_errors(D2, I2, Dt2, It2, "C2 nq=1")
    parity["paths"].append({"path": "C2: 1M rows, nq=1, host buffers", "queries": 32, "ok": not e})
    parity["queries"] += 32
    errors.extend(e)
    res = {}
    steps2 = max(50, args.steps)
    ms2, _ = timed(lambda i: res.__setitem__("r", ix2.search(q2[i % 32: i % 32 + 1], TOPK)), steps2, 5)
    _, ms2w = timed(lambda i: res.__setitem__("r", ix2.search(q2h[i % 32: i % 32 + 1], TOPK)), steps2, 5)
    ix2.set_timing(True)
    k2 = []
    for i in range(20):
        ix2.search(q2[i % 32: i % 32 + 1], TOPK)
        k2.append(ix2.stats()["last_search_ms"])
    ix2.set_timing(False)
    out["c2"] = {"workload": "C2: 1M x 1024 fp32, inner product, top-10, batch-1, one GPU",
                 "value": steps2 * 1000.0 / ms2, "unit": "queries/s", "ms_per_step": ms2 / steps2,
                 "e2e": {"value": steps2 * 1000.0 / ms2w, "unit": "queries/s"},
                 "kernel_ms": float(np.mean(k2)),
                 "hbm_frac": (n2 * DIM * 4 / (float(np.mean(k2)) * 1e-3) / 1e9) / peak, "steps": steps2}
    ix2.close()
    del ix2
    torch.cuda.empty_cache()

    # ---- C1: basic_rag shape -- 10k chunks through the LangChain surface, k=4, threshold 0.25 ----------
    n1 = 10_000
    xs = synth.corpus_host(n1, DIM)
    q1, _ = synth.queries_host(64, DIM, xs, noise=0.6)  # cos ~ 0.86 to the planted row: passes the threshold

    class Emb:  # the embedder is a remote NIM in the reference: here a lookup, so only the store is timed
        def __init__(self):
            self.table = {}

        def embed_documents(self, texts):
            return [self.table[t] for t in texts]

        def embed_query(self, t):
            return self.table[t]

    emb = Emb()
    texts = [f"chunk {i}" for i in range(n1)]
    for t, v in zip(texts, xs):
        emb.table[t] = v.tolist()  # List[float], as LangChain embedders return
    qtexts = [f"query {i}" for i in range(64)]
    for t, v in zip(qtexts, q1):
        emb.table[t] = v.tolist()
    from generativeaiexamples_b200 import create_vectorstore_langchain

    vs = create_vectorstore_langchain(emb, name="b200flat", dimensions=DIM)
    assert isinstance(vs, B200FlatVectorStore)
    vs.add_texts(texts, [{"source": f"f{i % 20}.pdf"} for i in range(n1)])
    retr = vs.as_retriever(search_type="similarity_score_threshold", search_kwargs={"score_threshold": 0.25, "k": 4})
    # truth: squared L2 in float64, relevance 1 - d/sqrt(2) >= 0.25 (LangChain _euclidean_relevance_score_fn)
    d2 = ((q1.astype(np.float64)[:, None, :] - xs.astype(np.float64)[None, :, :]) ** 2).sum(-1) if n1 <= 2000 else None
    if d2 is None:
        qq = (q1.astype(np.float64) ** 2).sum(1)[:, None]
        xx = (xs.astype(np.float64) ** 2).sum(1)[None, :]
        d2 = qq + xx - 2.0 * q1.astype(np.float64) @ xs.astype(np.float64).T
    import warnings

    bad = 0
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for i in range(64):
            order = np.argsort(d2[i], kind="stable")[:4]
            want = [texts[j] for j in order if 1.0 - d2[i, j] / np.sqrt(2.0) >= 0.25 + 1e-6]
            maybe = [texts[j] for j in order if 1.0 - d2[i, j] / np.sqrt(2.0) >= 0.25 - 1e-6]
            got = [d_.page_content for d_ in retr.get_relevant_documents(qtexts[i])]
            if not (got[: len(want)] == want and set(got) <= set(maybe)):
                bad += 1
                errors.append(f"C1: query {i}: retriever returned {got}, float64 truth {want}")
        parity["paths"].append({"path": "C1: LangChain retriever, k=4, threshold 0.25", "queries": 64, "ok": bad == 0})
        parity["queries"] += 64
        # (the benchmark's own embedding table is 10M Python floats: keep the cyclic collector from walking it
        # in the middle of a 70 us call)
        import gc

        gc.collect()
        gc.freeze()
        lat = []
        for i in range(300):
            t0 = time.perf_counter()
            retr.get_relevant_documents(qtexts[i % 64])
            lat.append((time.perf_counter() - t0) * 1e6)
        lat = sorted(lat[50:])
        vec = [emb.table[t] for t in qtexts]
        lat2 = []
        for i in range(300):
            t0 = time.perf_counter()
            vs.similarity_search_with_score_by_vector(vec[i % 64], k=4)
            lat2.append((time.perf_counter() - t0) * 1e6)
        lat2 = sorted(lat2[50:])
        gc.unfreeze()
    out["c1"] = {"workload": "C1: 10k x 1024 fp32 chunks, squared L2, k=4, score_threshold 0.25, single query, through "
                             "B200FlatVectorStore.as_retriever(similarity_score_threshold).get_relevant_documents "
                             "(List[float] in, Documents out)",
                 "value": 1e6 / (sum(lat) / len(lat)), "unit": "queries/s",
                 "p50_us": lat[len(lat) // 2], "p99_us": lat[int(len(lat) * 0.99)],
                 "by_vector_p50_us": lat2[len(lat2) // 2], "by_vector_p99_us": lat2[int(len(lat2) * 0.99)],
                 "calls": len(lat),
                 "note": "41 MB corpus: L2-resident, not measured against the HBM roofline (SURVEY.md section 8d)"}
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--devices", type=int, default=0,
                    help="N > 1: ONE process whose index spans N GPUs (b200r_config.n_devices) instead of N ranks")
    ap.add_argument("--rows", type=int, default=N_TOTAL, help="total corpus rows (default: the 10M headline config)")
    ap.add_argument("--batch256-steps", type=int, default=20)
    ap.add_argument("--no-shadow", action="store_true", help="skip the bf16-shadow batch-256 line (saves 50%% HBM)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the secondary C1 / C2 / C5 lines")
    ap.add_argument("--c5-rows", type=int, default=5_000_000)
    ap.add_argument("--c4", action="store_true",
                    help="also run C4: 12.5M x 1024 bf16 rows per GPU (100M on 8 GPUs), top-100, batch 1024, with "
                         "recall@100 against the fp32 truth of 64 queries")
    ap.add_argument("--cpu-sample-rows", type=int, default=0,
                    help="rows the CPU arm scans (0 = the full corpus when host RAM allows for --impl reference, "
                         "else 1M; scaled linearly to 10M rows and flagged `extrapolated`)")
    ap.add_argument("--concurrent", type=int, default=64, help="callers for the coalesced service-view number (0 = skip)")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup
    sys.exit(main_reference(a) if a.impl == "reference" else main_gpu(a))
