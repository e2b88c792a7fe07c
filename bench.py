#!/usr/bin/env python
"""bench.py -- headline benchmark: exact top-10 QPS over 10M x 1024 fp32 embeddings.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
           --master-port P bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json metric; config C3 of SURVEY.md section 8): corpus of 10M x 1024
fp32 unit vectors (40.96 GB -- fits one 180 GB B200), inner product, top-10, batch-1
queries.  The corpus is row-sharded over the N ranks (STRONG scaling: total rows fixed), each
rank searches its shard, candidates are merged with one NCCL allgather + the K4 select.
A "step" is one batch-1 query.  Data is synthetic (torch Philox on the device, one seed
per 1M-row chunk; generativeaiexamples_b200/synth.py); every step uses a different query
and the per-rank corpus (>= 5 GB) is far larger than the 126 MB L2, so no flush is needed.

JSON line: `value` = device-resident QPS (query already in HBM, results left in HBM),
`e2e` = the same through the public API with HOST query/result buffers (H2D + D2H inside
the timed region), `roofline` = scan kernel vs measured HBM bandwidth, `cpu_baseline` = the
faiss-1.7.4 restatement (oracle/) timed on this box's host cores on a bounded sample, and
`batch256` = the batch-256 variant of the same search.

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
METRIC_NAME = "top10_qps_10Mx1024_fp32_ip_batch1"
WORKLOAD = "C3: 10M x 1024 fp32 corpus, inner product, top-10, batch-1 queries, rows sharded over N GPUs"


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
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100", "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
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
        sm, mx, reasons = [], [], set()
        for r in rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.strip().lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


# ----------------------------------------------------------------------------- CPU arm
def cpu_reference(steps: int, warmup: int, sample_rows: int = 400_000, budget_s: float = 25.0):
    """faiss-1.7.4 restatement on the host cores, nq=1, on a bounded slice of the workload.
    Exhaustive scan is exactly linear in N, so QPS at 10M rows = QPS(sample) * sample/10M."""
    import numpy as np

    from oracle import FlatRef, load_flat_ref

    lib = load_flat_ref()
    threads = int(lib.oracle_max_threads())
    rng = np.random.Generator(np.random.PCG64(1234))
    x = rng.standard_normal((sample_rows, DIM), dtype=np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
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
        return done / (time.perf_counter() - t0), done

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
        qps_faiss, nf = run(lambda v: fidx.search(v, TOPK), min(warmup, 2), steps, budget_s * 0.3)
        Df, If = fidx.search(q[:8], TOPK)
        Dr, Ir = FlatRef.knn(0, q[:8], x, TOPK)
        live = {"faiss_importable": True, "faiss_version": getattr(faiss, "__version__", "?"),
                "value": qps_faiss * scale, "steps": nf, "ids_equal_to_restatement": bool((If == Ir).all()),
                "max_abs_score_diff_vs_restatement": float(np.abs(Df - Dr).max())}
    except ImportError:
        pass
    # (i) faithful: faiss 1.7.4 runs nq < 20 on ONE thread (OpenMP over queries only)
    qps_faithful, n1 = run(lambda v: FlatRef.knn(0, v, x, TOPK), min(warmup, 2), steps, budget_s * 0.4)
    # (ii) the same scan row-parallel over all cores (not a faiss code path; stronger baseline)
    qps_all, n2 = run(lambda v: FlatRef.knn_allcores(0, v, x, TOPK), min(warmup, 3), steps, budget_s * 0.6)
    cpu_model = ""
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                cpu_model = ln.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    if live["faiss_importable"]:  # the real thing: report it, keep the restatement's numbers beside it
        return {"value": live["value"], "unit": "queries/s", "cores": threads, "kind": "reference", "cpu_model": cpu_model,
                "sample": f"faiss.IndexFlatIP.search, nq=1 top-{TOPK} over a {sample_rows} x {DIM} slice, scaled to {N_TOTAL} rows",
                "port_allcores_value": qps_all * scale, "faithful_1thread_value": qps_faithful * scale,
                "host_cores_visible": len(os.sched_getaffinity(0)), "live_faiss": live}
    return {
        "value": qps_all * scale, "unit": "queries/s", "cores": threads, "kind": "port", "cpu_model": cpu_model,
        "omp_num_threads": os.environ.get("OMP_NUM_THREADS", "unset (OpenMP default = all cores)"),
        "blas": "OpenBLAS bundled with numpy (not on the nq=1 path)",
        "sample": (f"nq=1 top-{TOPK} scans over a {sample_rows} x {DIM} fp32 slice ({n2} all-core steps, {n1} "
                   f"single-thread steps), scaled linearly to {N_TOTAL} rows"),
        "faithful_1thread_value": qps_faithful * scale,
        "note": "value = all-cores row-parallel scan; faithful_1thread_value = faiss 1.7.4's own nq<20 path (1 thread)",
        "host_cores_visible": len(os.sched_getaffinity(0)),
        "live_faiss": live,
    }


def main_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0  # rank 0 alone runs the CPU arm
    cb = cpu_reference(args.steps, args.warmup, sample_rows=args.cpu_sample_rows)
    line = {
        "impl": "reference", "metric": METRIC_NAME, "value": cb["value"], "unit": "queries/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 / cb["value"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "rows": N_TOTAL, "dim": DIM, "k": TOPK, "nq": 1},
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ----------------------------------------------------------------------------- GPU arm
def main_gpu(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    from generativeaiexamples_b200 import synth
    from generativeaiexamples_b200.flat_index import METRIC_INNER_PRODUCT
    from generativeaiexamples_b200.sharded import ShardedFlatIndex, shard_bounds

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
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
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world} (launch with torch.distributed.run)"

    n_total = args.rows
    lo, hi = shard_bounds(n_total, world, rank)
    # the optional bf16 shadow (+50% HBM) only serves the `batch256_shadow` line; it is switched off
    # for every other measurement, which therefore streams the fp32 rows as BASELINE.json's metric says
    use_shadow = args.batch256_steps > 0 and not args.no_shadow
    index = ShardedFlatIndex(DIM, METRIC_INNER_PRODUCT, device=local_rank, reserve_rows=hi - lo,
                             shadow_bf16=use_shadow)
    if use_shadow:
        index.local.set_shadow_filter(False)
    # ---- ingest (untimed here; K1 throughput is reported separately) -------------------
    t_ing = time.perf_counter()
    row = lo
    probe_rows = None
    while row < hi:
        c = row // CHUNK
        c_lo, c_hi = c * CHUNK, min((c + 1) * CHUNK, n_total)
        x = synth.corpus_chunk_device(c, c_hi - c_lo, DIM, dev)
        part = x[max(lo, c_lo) - c_lo: min(hi, c_hi) - c_lo]
        if probe_rows is None:
            probe_rows = part[:4096].clone()  # planted-neighbour sources (global ids lo .. lo+4095)
        index.local.add(part)
        row = min(hi, c_hi)
        del x, part
    index._sync_counts()
    torch.cuda.synchronize()
    ingest_s = time.perf_counter() - t_ing
    assert index.ntotal == n_total, (index.ntotal, n_total)

    # ---- queries: pool of 64, half planted near rows of rank 0's shard --------------------
    g = torch.Generator(device=dev)
    g.manual_seed(synth.QUERY_SEED)
    pool = 64
    Q = torch.randn((pool, DIM), generator=g, device=dev)
    planted = torch.full((pool,), -1, dtype=torch.int64)
    src = probe_rows if rank == 0 else torch.zeros_like(probe_rows)
    if world > 1:
        dist.broadcast(src, 0)
    for i in range(0, pool, 2):
        j = (i * 61) % src.shape[0]
        Q[i] = src[j] + 0.1 * Q[i] / DIM ** 0.5
        planted[i] = j
    Q /= Q.norm(dim=1, keepdim=True)
    Qh = Q.cpu().numpy()
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

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
        results["last"] = index.search(Q[i % pool: i % pool + 1], TOPK)

    def e2e_step(i):
        results["last_h"] = index.search(Qh[i % pool: i % pool + 1], TOPK)

    # correctness of what is being timed (planted ids, sorted scores) before any timing
    D, I = index.search(Q[:8], TOPK)
    torch.cuda.synchronize()
    Dc, Ic = D.cpu().numpy(), I.cpu().numpy()
    for i in range(0, 8, 2):
        assert Ic[i, 0] == int(planted[i]) and Dc[i, 0] > 0.98, ("planted neighbour not found", i, Ic[i], Dc[i])
    assert (np.diff(-Dc, axis=1) >= 0).all()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    launches0 = index.local.stats()["n_kernel_launches"]
    ms_dev, _ = timed(dev_step, args.steps, args.warmup)
    launches = index.local.stats()["n_kernel_launches"] - launches0
    launches = launches * args.steps // (args.steps + args.warmup)
    if world > 1:
        launches += args.steps  # K4 merge, one per step
    clocks = sampler.stop() if sampler else None
    _, ms_e2e_wall = timed(e2e_step, args.steps, args.warmup)
    # per-call latency distribution of the host-buffer path (wall clock around each call)
    lat = []
    for i in range(min(args.steps, 200)):
        t0 = time.perf_counter()
        e2e_step(i)
        lat.append((time.perf_counter() - t0) * 1e6)
    lat.sort()
    e2e_lat = {"p50_us": lat[len(lat) // 2], "p99_us": lat[min(len(lat) - 1, int(len(lat) * 0.99))], "calls": len(lat)}

    # ---- scan kernel alone: CUDA events on the launching stream, live -----------------------
    index.local.set_timing(True)
    kms = []
    for i in range(args.steps):
        index.local.search(Q[i % pool: i % pool + 1], TOPK)
        kms.append(index.local.stats()["last_search_ms"])
    index.local.set_timing(False)
    kernel_ms = float(np.mean(kms))
    kt = torch.tensor([kernel_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(kt, op=dist.ReduceOp.MAX)
    kernel_ms = float(kt[0])
    local_rows = hi - lo
    max_rows = (n_total + world - 1) // world
    algo_bytes = max_rows * DIM * 4 + DIM * 4 + TOPK * 12
    peak, peak_src = _peaks()
    achieved = algo_bytes / (kernel_ms * 1e-3) / 1e9
    traffic = _traffic()

    # ---- batch-256 variant (secondary) ------------------------------------------------------
    b256 = None
    if args.batch256_steps > 0:
        g.manual_seed(synth.QUERY_SEED + 1)
        Q256 = torch.randn((256, DIM), generator=g, device=dev)
        Q256 /= Q256.norm(dim=1, keepdim=True)

        def b256_step(i):
            results["b256"] = index.search(Q256, TOPK)

        ms256, _ = timed(b256_step, args.batch256_steps, 3)
        per = ms256 / args.batch256_steps
        flops = 2.0 * 256 * n_total * DIM
        tf32_peak, _, tf32_src = _tensor_peaks()
        per_gpu_tflops = flops / world / (per * 1e-3) / 1e12
        b256 = {"value": 256 * 1000.0 / per, "unit": "queries/s", "ms_per_batch": per, "steps": args.batch256_steps,
                "tensor_frac_tf32": per_gpu_tflops / tf32_peak, "tensor_peak_tflops": tf32_peak,
                "tensor_peak_source": tf32_src,
                "path": "scan x64 passes" if index.local.stats()["last_path"] == 1
                else "tcgen05 tf32 GEMM filter + exact fp32 re-score", "gemm_fallbacks": index.local.stats()["n_gemm_fallbacks"], "achieved_tflops": flops / (per * 1e-3) / 1e12,
                "hbm_frac": (max_rows * DIM * 4 / (per * 1e-3) / 1e9) / peak,
                "note": "crossover config (128 flop/B): both fractions reported; flops = 2*nq*N*d, bytes = N*d*4"}
        if use_shadow:
            # same batch, filter pass on the bf16 shadow of the rows; the fp32 re-score makes the answers identical
            D_f, I_f = results["b256"]
            D_f, I_f = D_f.clone(), I_f.clone()
            index.local.set_shadow_filter(True)
            fb0 = index.local.stats()["n_gemm_fallbacks"]
            ms256s, _ = timed(b256_step, args.batch256_steps, 3)
            D_s, I_s = results["b256"]
            same = bool(torch.equal(I_s, I_f) and torch.equal(D_s, D_f))
            index.local.set_shadow_filter(False)
            pers = ms256s / args.batch256_steps
            bf16_peak = _tensor_peaks()[1]
            b256["shadow_bf16"] = {
                "value": 256 * 1000.0 / pers, "unit": "queries/s", "ms_per_batch": pers,
                "achieved_tflops": flops / (pers * 1e-3) / 1e12,
                "tensor_frac_bf16": flops / world / (pers * 1e-3) / 1e12 / bf16_peak, "tensor_peak_tflops": bf16_peak,
                "hbm_frac": (max_rows * DIM * 2 / (pers * 1e-3) / 1e9) / peak,
                "identical_to_fp32_filter": same,
                "gemm_fallbacks": index.local.stats()["n_gemm_fallbacks"] - fb0,
                "note": "opt-in index mode (shadow_bf16=True, +50% HBM): filter streams N*d*2 bytes of bf16 rows at the "
                        "bf16 MMA rate, survivors are re-scored on the fp32 rows; not used for `value`/`e2e`"}
            if not same:
                raise SystemExit("bench: shadow-filter results differ from the fp32-filter results")

    # ---- batch-1 on the bf16 shadow (opt-in index mode; secondary) ---------------------------------
    b1s = None
    if use_shadow:
        ref = [index.search(Q[i: i + 1], TOPK) for i in range(8)]
        ref = [(d_.clone(), i_.clone()) for d_, i_ in ref]
        index.local.set_shadow_filter(True)
        fb0 = index.local.stats()["n_gemm_fallbacks"]
        same = True
        for i in range(8):
            d_, i_ = index.search(Q[i: i + 1], TOPK)
            same = same and bool(torch.equal(i_, ref[i][1]) and torch.equal(d_, ref[i][0]))
        steps1 = max(20, args.steps // 2)
        ms_s, _ = timed(dev_step, steps1, 5)
        _, ms_s_wall = timed(e2e_step, steps1, 5)
        fbs = index.local.stats()["n_gemm_fallbacks"] - fb0
        index.local.set_shadow_filter(False)
        b1s = {"value": steps1 * 1000.0 / ms_s, "unit": "queries/s", "ms_per_step": ms_s / steps1,
               "e2e": {"value": steps1 * 1000.0 / ms_s_wall, "unit": "queries/s"},
               "hbm_frac_bf16_bytes": (max_rows * DIM * 2 / (ms_s / steps1 * 1e-3) / 1e9) / peak,
               "identical_to_fp32_scan": same, "path": "tcgen05 bf16 GEMM filter on the shadow + exact fp32 re-score", "gemm_fallbacks": fbs, "steps": steps1,
               "note": "opt-in index mode (shadow_bf16=True, +50% HBM): the tcgen05 filter streams N*d*2 bytes of bf16 "
                       "rows for the K'=112 best approximate scores, which are re-scored on the fp32 rows and checked "
                       "for containment; not used for `value`/`e2e`/`roofline`"}
        if not same:
            raise SystemExit("bench: shadow-scan results differ from the fp32-scan results")

    # ---- service view: 64 concurrent single-query callers through the coalescer (N=1 only) -------
    conc = None
    if world == 1 and args.concurrent > 0:
        from concurrent.futures import ThreadPoolExecutor

        from generativeaiexamples_b200.batching import SearchCoalescer

        co = SearchCoalescer(index.local)
        nreq = args.concurrent * 24

        def one(i):
            return co.search(Qh[i % pool: i % pool + 1], TOPK)

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

    if rank == 0:
        qps = args.steps * 1000.0 / ms_dev
        e2e_qps = args.steps * 1000.0 / ms_e2e_wall
        cb = cpu_reference(20, 2, sample_rows=args.cpu_sample_rows) if (world == 1 and not args.no_cpu_baseline) else None
        line = {
            "metric": METRIC_NAME, "value": qps, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD if n_total == N_TOTAL else f"REDUCED ({n_total} rows) " + WORKLOAD,
                       "rows": n_total, "rows_per_gpu": max_rows, "dim": DIM, "k": TOPK, "nq": 1,
                       "parallelism": (f"row-shard x{world} + " + ("fused NVLink peer-store exchange" if getattr(index, "_xchg", None)
                                                                  else "NCCL allgather") + " + K4 select") if world > 1 else "single GPU",
                       "l2_policy": "no flush: per-GPU corpus (>=5 GB) >> 126 MB L2, new query every step",
                       "ingest_s": round(ingest_s, 3)},
            "e2e": {"value": e2e_qps, "unit": "queries/s", "h2d_bytes_per_step": DIM * 4,
                    "d2h_bytes_per_step": TOPK * 12, "latency": e2e_lat,
                    "api": ("B200IndexFlat.search(numpy) = b200r_search(HOST query, HOST results)" if world == 1 else
                            "ShardedFlatIndex.search(numpy): pinned H2D -> b200r_search(DEVICE) -> NCCL allgather -> "
                            "b200r_merge_topk -> one packed D2H")},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": (algo_bytes * traffic["traffic_over_algorithmic"]) if traffic else None,
                         "traffic_source": ("ncu dram__bytes_read+write per launch = %.4f x algorithmic bytes, measured on a "
                                            "2M-row shard (%s), scaled to this shard" %
                                            (traffic["traffic_over_algorithmic"], traffic["source"])) if traffic else None,
                         "peak_source": peak_src,
                         "kernel": "scan_topk_kernel<1,1,IP,f32>", "kernel_ms": kernel_ms,
                         "algorithmic_bytes_per_launch": algo_bytes,
                         "note": "kernel_ms = prep_queries + scan launch, CUDA events on the launching stream"},
            "cpu_baseline": cb,
            "batch256": b256,
            "batch1_shadow_bf16": b1s,
            "concurrent_callers": conc,
            "clocks": clocks,
        }
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=N_TOTAL, help="total corpus rows (default: the 10M headline config)")
    ap.add_argument("--batch256-steps", type=int, default=20)
    ap.add_argument("--no-shadow", action="store_true", help="skip the bf16-shadow batch-256 line (saves 50%% HBM)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample-rows", type=int, default=400_000,
                    help="rows of the bounded sample the CPU arm / cpu_baseline scans (scaled linearly to --rows)")
    ap.add_argument("--concurrent", type=int, default=64, help="callers for the coalesced service-view number (0 = skip)")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3) if a.impl == "b200" else a.warmup
    sys.exit(main_reference(a) if a.impl == "reference" else main_gpu(a))
