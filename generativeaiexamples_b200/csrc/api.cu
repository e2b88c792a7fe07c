// The single-GPU index ("leaf") behind the C ABI of libb200r.so: HBM corpus buffer, pinned
// staging rings, per-index search workspace, error plumbing.  include/b200r.h names the reference
// call each operation replaces; abi.cu exports them.  No CPU compute path exists here.
#include "index.h"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

using namespace b200r;

// ---------------------------------------------------------------------------
// errors
// ---------------------------------------------------------------------------
static thread_local std::string g_err;

namespace b200r {
int fail(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
void set_last_error(const char* msg) { g_err = msg ? msg : ""; }
const char* last_error_cstr() { return g_err.c_str(); }

// ---------------------------------------------------------------------------
// environment switches
// ---------------------------------------------------------------------------
static Knobs g_knobs;
static std::once_flag g_knobs_once;
static std::mutex g_knobs_mu;

static void read_knobs_locked() {
    Knobs k;
    auto geti = [](const char* name, long long dflt) -> long long {
        const char* ev = getenv(name);
        return (ev && *ev) ? atoll(ev) : dflt;
    };
    k.stage_threads = (int)std::max<long long>(1, geti("B200R_STAGE_THREADS", 8));
    k.gemm_min_nq = geti("B200R_GEMM_MIN_NQ", -1);
    k.gemm_min_bytes = geti("B200R_GEMM_MIN_BYTES", -1);
    k.evict_first = (int)geti("B200R_EVICT_FIRST", -1);
    k.gemm_noboost = getenv("B200R_GEMM_NOBOOST") != nullptr;
    k.debug_fails = getenv("B200R_DEBUG_FAILS") != nullptr;
    k.scan_stage_kb = (int)geti("B200R_SCAN_STAGE_KB", 0);
    k.scan_stages = (int)geti("B200R_SCAN_STAGES", 0);
    k.scan_dynamic = (int)geti("B200R_SCAN_DYNAMIC", -1);
    k.scan_trace = (int)geti("B200R_SCAN_TRACE", 0);
    k.scan_tile_block = (int)geti("B200R_SCAN_TILE_BLOCK", 1);
    k.zero_copy_out = (int)geti("B200R_ZERO_COPY_OUT", 1);
    k.gemm_kpcap = (int)geti("B200R_GEMM_KPCAP", 0);
    k.gemm_kprime = (int)geti("B200R_GEMM_KPRIME", 0);
    k.gemm_cluster = (int)geti("B200R_GEMM_CLUSTER", 0);
    k.gemm_pair = (int)geti("B200R_GEMM_PAIR", -1);
    k.gemm_prefetch = (int)geti("B200R_GEMM_PREFETCH", 0);
    k.gemm_noselect = (int)geti("B200R_GEMM_DEBUG_NOSELECT", 0);
    k.gemm_warmup = (int)geti("B200R_GEMM_WARMUP", -1);
    k.multi_workers = (int)geti("B200R_MULTI_WORKERS", -1);
    k.multi_spin_us = (int)geti("B200R_MULTI_SPIN_US", 2000);
    k.multi_fused = (int)geti("B200R_MULTI_FUSED", -1);
    g_knobs = k;
}
const Knobs& knobs() {
    std::call_once(g_knobs_once, [] {
        std::lock_guard<std::mutex> lk(g_knobs_mu);
        read_knobs_locked();
    });
    return g_knobs;
}
void reload_knobs() {
    knobs();
    std::lock_guard<std::mutex> lk(g_knobs_mu);
    read_knobs_locked();
}
}  // namespace b200r

// corpus sizes from which small batches go to the GEMM path (measured crossovers with the device-side
// fallback in place, tools/perf_small_nq.py / perf_bf16_small.py).  2..4 queries: the four-rows-per-task scan
// (round 2) matches the GEMM route up to ~1M x 1024 fp32 rows (0.60-0.66 ms against 0.62) and loses beyond
// (4M rows: 2.7-2.85 ms against 2.4: it is bound by the SM clock, which drops under the power cap on long
// launches), so the switch sits at 3 GB; a single query on the bf16 shadow or on bf16 rows from ~250k / ~500k rows
constexpr int64_t kGemmSmallBatchBytes = 3072ll << 20;
constexpr int64_t kGemmShadowSingleBytes = 1024ll << 20;
constexpr int64_t kGemmBf16SmallBatchBytes = 64ll << 20;
constexpr int64_t kGemmBf16SingleBytes = 1024ll << 20;

namespace b200r {

static int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }
static void fold_fallback_counts(b200r_index* ix, bool wait);

// pageable -> pinned staging copy; one core moves ~12 GB/s, PCIe Gen5 takes ~50, so large
// copies are split over a few threads (at most half the cores)
static void staged_memcpy(void* dst, const void* src, size_t bytes) {
    constexpr size_t kMinPerThread = 4u << 20;
    unsigned hw = std::thread::hardware_concurrency();
    const int max_threads = knobs().stage_threads;  // measured on the 16-core GPU box: 18 GB/s with 4, 24 with 8, 21 with 12
    int nt = (int)std::min<size_t>(std::min<unsigned>(hw ? std::max(1u, hw / 2) : 1, (unsigned)max_threads), bytes / kMinPerThread);
    if (nt <= 1) {
        memcpy(dst, src, bytes);
        return;
    }
    std::vector<std::thread> th;
    const size_t per = (bytes / nt + 4095) & ~(size_t)4095;
    for (int t = 1; t < nt; t++) {
        const size_t lo = std::min(bytes, per * t), hi = std::min(bytes, per * (t + 1));
        if (hi > lo)
            th.emplace_back([=] { memcpy((uint8_t*)dst + lo, (const uint8_t*)src + lo, hi - lo); });
    }
    memcpy(dst, src, std::min(bytes, per));
    for (auto& t : th) t.join();
}

static int ensure_capacity(b200r_index* ix, int64_t need, cudaStream_t st) {
    if (need <= ix->capacity) return B200R_OK;
    if (need > kMaxRows) return fail(B200R_E_UNSUPPORTED, "index is limited to %lld rows per GPU", (long long)kMaxRows);
    int64_t cap = std::max<int64_t>(need, ix->capacity + ix->capacity / 2);
    cap = std::max<int64_t>(cap, 1024);
    uint8_t* nr = nullptr;
    float* nn = nullptr;
    cudaError_t e = cudaMalloc(&nr, (size_t)cap * ix->row_bytes);
    if (e != cudaSuccess && cap > need) {  // retry without head-room
        cudaGetLastError();
        cap = need;
        e = cudaMalloc(&nr, (size_t)cap * ix->row_bytes);
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(B200R_E_NOMEM, "cudaMalloc of %.2f GB for %lld rows failed: %s", cap * (double)ix->row_bytes / 1e9,
                    (long long)cap, cudaGetErrorString(e));
    }
    e = cudaMalloc(&nn, (size_t)cap * sizeof(float));
    if (e != cudaSuccess) {
        cudaGetLastError();
        cudaFree(nr);
        return fail(B200R_E_NOMEM, "cudaMalloc of row norms failed: %s", cudaGetErrorString(e));
    }
    uint8_t* ns = nullptr;
    if (ix->cfg.shadow_bf16) {
        e = cudaMalloc(&ns, (size_t)cap * ix->d_pad * 2);
        if (e != cudaSuccess) {
            cudaGetLastError();
            cudaFree(nr);
            cudaFree(nn);
            return fail(B200R_E_NOMEM, "cudaMalloc of %.2f GB for the bf16 shadow failed: %s",
                        cap * (double)ix->d_pad * 2 / 1e9, cudaGetErrorString(e));
        }
        if (ix->ntotal > 0)
            CU(cudaMemcpyAsync(ns, ix->shadow, (size_t)ix->ntotal * ix->d_pad * 2, cudaMemcpyDeviceToDevice, st));
    }
    if (ix->ntotal > 0) {
        CU(cudaMemcpyAsync(nr, ix->rows, (size_t)ix->ntotal * ix->row_bytes, cudaMemcpyDeviceToDevice, st));
        CU(cudaMemcpyAsync(nn, ix->norms, (size_t)ix->ntotal * sizeof(float), cudaMemcpyDeviceToDevice, st));
    }
    CU(cudaStreamSynchronize(st));
    // asynchronous searches on other streams may still be reading the old buffers
    if (ix->rows) CU(cudaDeviceSynchronize());
    if (ix->rows) cudaFree(ix->rows);
    if (ix->norms) cudaFree(ix->norms);
    if (ix->shadow) cudaFree(ix->shadow);
    ix->rows = nr;
    ix->norms = nn;
    ix->shadow = ns;
    ix->capacity = cap;
    return B200R_OK;
}

void leaf_free(b200r_index* ix) {
    if (!ix) return;
    DeviceGuard g(ix->device);
    if (ix->stream) cudaStreamSynchronize(ix->stream);
    cudaFree(ix->rows);
    cudaFree(ix->norms);
    cudaFree(ix->shadow);
    cudaFreeHost(ix->pin_add);
    cudaFree(ix->dev_add);
    cudaFreeHost(ix->q_pin);
    cudaFree(ix->q_raw);
    cudaFree(ix->q_dev);
    cudaFree(ix->scratch);
    cudaFree(ix->counter);
    cudaFree(ix->out_buf);
    cudaFreeHost(ix->out_pin);
    cudaFree(ix->q_bf16);
    cudaFree(ix->gemm_cand);
    cudaFree(ix->max_norm_bits);
    cudaFree(ix->gemm_aux);
    cudaFree(ix->fail_dev);
    cudaFreeHost(ix->fail_pin);
    cudaFree(ix->flag_dev);
    cudaFreeHost(ix->fb_pin);
    cudaFree(ix->recon_dev);
    cudaFreeHost(ix->recon_pin);
    cudaFree(ix->bounce);
    cudaFree(ix->adj_dev);
    cudaFree(ix->trace_dev);
    free_rerank_ws(ix);
    for (auto& e : ix->fb_ev)
        if (e) cudaEventDestroy(e);
    for (auto& e : ix->add_ev)
        if (e) cudaEventDestroy(e);
    if (ix->data_ev) cudaEventDestroy(ix->data_ev);
    if (ix->ws_ev) cudaEventDestroy(ix->ws_ev);
    if (ix->t0) cudaEventDestroy(ix->t0);
    if (ix->t1) cudaEventDestroy(ix->t1);
    if (ix->stream) cudaStreamDestroy(ix->stream);
    cudaGetLastError();
    delete ix;
}

int leaf_create(const b200r_config* cfg, b200r_index** out) {
    if (!cfg || !out) return fail(B200R_E_INVALID, "b200r_create: null argument");
    *out = nullptr;
    if (cfg->dim <= 0) return fail(B200R_E_INVALID, "dim must be positive, got %d", cfg->dim);
    if (cfg->metric != B200R_METRIC_INNER_PRODUCT && cfg->metric != B200R_METRIC_L2)
        return fail(B200R_E_INVALID, "metric must be B200R_METRIC_INNER_PRODUCT or B200R_METRIC_L2");
    if (cfg->storage != B200R_F32 && cfg->storage != B200R_BF16)
        return fail(B200R_E_INVALID, "storage must be B200R_F32 or B200R_BF16");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(B200R_E_CUDA, "no CUDA device available (%s); libb200r has no CPU path",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    }
    if (cfg->device < 0 || cfg->device >= ndev)
        return fail(B200R_E_INVALID, "device %d out of range (%d visible)", cfg->device, ndev);
    DeviceGuard g(cfg->device);
    if (!g.ok) return fail(B200R_E_CUDA, "cudaSetDevice(%d) failed", cfg->device);
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major != 10)
        return fail(B200R_E_UNSUPPORTED, "device %d is sm_%d%d; libb200r is built for sm_100a (B200) only", cfg->device,
                    prop.major, prop.minor);

    b200r_index* ix = new b200r_index();
    ix->cfg = *cfg;
    ix->d = cfg->dim;
    ix->bf16 = cfg->storage == B200R_BF16;
    ix->l2 = cfg->metric == B200R_METRIC_L2;
    if (cfg->shadow_bf16 && ix->bf16) {
        delete ix;
        return fail(B200R_E_INVALID, "shadow_bf16 needs fp32 storage (a bf16 index is its own shadow)");
    }
    ix->cfg.shadow_bf16 = cfg->shadow_bf16 ? 1 : 0;
    ix->shadow_on = ix->cfg.shadow_bf16 != 0;
    ix->d_pad = (int)round_up(cfg->dim, (ix->bf16 || ix->cfg.shadow_bf16) ? 8 : 4);
    ix->row_bytes = ix->d_pad * (ix->bf16 ? 2 : 4);
    ix->device = cfg->device;
    ix->num_sms = prop.multiProcessorCount;
    ix->smem_optin = prop.sharedMemPerBlockOptin;
    ix->st.num_sms = ix->num_sms;
    if (ix->row_bytes > 64 * 1024) {
        delete ix;
        return fail(B200R_E_UNSUPPORTED, "dim %d too large: a row must fit 64 KB", cfg->dim);
    }
#define CU_OR_FREE(call)                                                                          \
    do {                                                                                          \
        cudaError_t e__ = (call);                                                                 \
        if (e__ != cudaSuccess) {                                                                 \
            leaf_free(ix);                                                                       \
            cudaGetLastError();                                                                   \
            return fail(e__ == cudaErrorMemoryAllocation ? B200R_E_NOMEM : B200R_E_CUDA, "%s: %s", #call, \
                        cudaGetErrorString(e__));                                                 \
        }                                                                                         \
    } while (0)
    CU_OR_FREE(cudaStreamCreateWithFlags(&ix->stream, cudaStreamNonBlocking));
    CU_OR_FREE(cudaEventCreateWithFlags(&ix->add_ev[0], cudaEventDisableTiming));
    CU_OR_FREE(cudaEventCreateWithFlags(&ix->add_ev[1], cudaEventDisableTiming));
    CU_OR_FREE(cudaEventCreateWithFlags(&ix->data_ev, cudaEventDisableTiming));
    CU_OR_FREE(cudaEventCreateWithFlags(&ix->ws_ev, cudaEventDisableTiming));
    CU_OR_FREE(cudaEventCreate(&ix->t0));
    CU_OR_FREE(cudaEventCreate(&ix->t1));
    CU_OR_FREE(cudaMalloc(&ix->scratch, scan_scratch_bytes(ix->num_sms)));
    CU_OR_FREE(cudaMalloc(&ix->counter, 256));
    CU_OR_FREE(cudaMemset(ix->counter, 0, 256));
#undef CU_OR_FREE
    if (cfg->reserve_rows > 0) {
        int rc = ensure_capacity(ix, cfg->reserve_rows, ix->stream);
        if (rc != B200R_OK) {
            leaf_free(ix);
            return rc;
        }
    }
    *out = ix;
    return B200R_OK;
}

int leaf_stats(b200r_index* ix, b200r_stats* out) {
    std::lock_guard<std::mutex> lk(ix->mu);
    {
        DeviceGuard g(ix->device);
        fold_fallback_counts(ix, true);
    }
    ix->st.ntotal = ix->ntotal;
    ix->st.capacity_rows = ix->capacity;
    ix->st.last_kernel_us = ix->st.last_search_ms * 1000.0;
    ix->st.last_exchange_us = 0.0;
    ix->st.n_devices = 1;
    *out = ix->st;
    return B200R_OK;
}

int leaf_reset(b200r_index* ix) {
    std::lock_guard<std::mutex> lk(ix->mu);
    ix->ntotal = 0;
    ix->max_norm_dirty = true;
    return B200R_OK;
}

// ---------------------------------------------------------------------------
// add
// ---------------------------------------------------------------------------
int leaf_add(b200r_index* ix, const void* x, int64_t n, int dtype, int mem_kind, void* stream, bool allow_normalize) {
    if (n < 0) return fail(B200R_E_INVALID, "b200r_add: n = %lld", (long long)n);
    if (n == 0) return B200R_OK;
    if (!x) return fail(B200R_E_INVALID, "b200r_add: null data");
    if (dtype != B200R_F32) return fail(B200R_E_UNSUPPORTED, "b200r_add: input vectors must be fp32");
    if (mem_kind != B200R_HOST && mem_kind != B200R_DEVICE) return fail(B200R_E_INVALID, "b200r_add: bad mem_kind");
    std::lock_guard<std::mutex> lk(ix->mu);
    DeviceGuard g(ix->device);
    // device input: ordered on the caller's stream exactly as given (NULL is CUDA's default
    // stream, e.g. torch's); host input: the index's own non-blocking stream unless told otherwise
    cudaStream_t st = (mem_kind == B200R_DEVICE || stream) ? (cudaStream_t)stream : ix->stream;
    // an earlier device add may still be writing rows on another stream: this add is ordered after it
    // (a growth below copies those rows, and data_ev is about to be re-recorded on `st` alone)
    if (ix->data_ev_set && ix->data_ev_stream != st) CU(cudaStreamWaitEvent(st, ix->data_ev, 0));
    int rc = ensure_capacity(ix, ix->ntotal + n, st);
    if (rc != B200R_OK) return rc;
    const bool norm = allow_normalize && ix->cfg.normalize_on_add != 0;
    uint8_t* dst = ix->rows + (size_t)ix->ntotal * ix->row_bytes;
    float* dstn = ix->norms + ix->ntotal;
    if (mem_kind == B200R_DEVICE) {
        CU(launch_ingest((const float*)x, n, ix->d, ix->d_pad, norm, ix->bf16, dst, dstn, st));
        ix->st.n_kernel_launches++;
        if (ix->shadow) {
            CU(launch_ingest((const float*)x, n, ix->d, ix->d_pad, norm, true,
                             ix->shadow + (size_t)ix->ntotal * ix->d_pad * 2, nullptr, st));
            ix->st.n_kernel_launches++;
        }
    } else {
        if (!ix->pin_add) {
            CU(cudaMallocHost(&ix->pin_add, 2 * kStageHalfBytes));
            CU(cudaMalloc(&ix->dev_add, 2 * kStageHalfBytes));
        }
        const size_t in_row = (size_t)ix->d * sizeof(float);
        const int64_t chunk = std::max<int64_t>(1, (int64_t)(kStageHalfBytes / in_row));
        int half = 0;
        bool used[2] = {false, false};
        for (int64_t r0 = 0; r0 < n; r0 += chunk, half ^= 1) {
            const int64_t m = std::min(chunk, n - r0);
            if (used[half]) CU(cudaEventSynchronize(ix->add_ev[half]));
            uint8_t* ph = ix->pin_add + half * kStageHalfBytes;
            uint8_t* dh = ix->dev_add + half * kStageHalfBytes;
            staged_memcpy(ph, (const uint8_t*)x + (size_t)r0 * in_row, (size_t)m * in_row);
            CU(cudaMemcpyAsync(dh, ph, (size_t)m * in_row, cudaMemcpyHostToDevice, st));
            CU(launch_ingest((const float*)dh, m, ix->d, ix->d_pad, norm, ix->bf16, dst + (size_t)r0 * ix->row_bytes,
                             dstn + r0, st));
            if (ix->shadow) {
                CU(launch_ingest((const float*)dh, m, ix->d, ix->d_pad, norm, true,
                                 ix->shadow + (size_t)(ix->ntotal + r0) * ix->d_pad * 2, nullptr, st));
                ix->st.n_kernel_launches++;
            }
            CU(cudaEventRecord(ix->add_ev[half], st));
            used[half] = true;
            ix->st.n_kernel_launches++;
        }
        CU(cudaStreamSynchronize(st));
    }
    ix->ntotal += n;
    ix->max_norm_dirty = true;
    ix->st.n_rows_added += n;
    if (mem_kind == B200R_DEVICE) {  // still in flight on the caller's stream
        CU(cudaEventRecord(ix->data_ev, st));
        ix->data_ev_set = true;
        ix->data_ev_stream = st;
    } else {
        ix->data_ev_set = false;  // host adds return after a stream synchronize
    }
    return B200R_OK;
}

// ---------------------------------------------------------------------------
// search
// ---------------------------------------------------------------------------
static int ensure_search_ws(b200r_index* ix, int64_t nq_slots, int64_t nq, int k) {
    if (nq_slots > ix->q_cap) {
        cudaFreeHost(ix->q_pin);
        cudaFree(ix->q_raw);
        cudaFree(ix->q_dev);
        cudaFree(ix->q_bf16);
        ix->q_pin = ix->q_raw = ix->q_dev = nullptr;
        ix->q_bf16 = nullptr;
        ix->q_cap = 0;
        const int64_t cap = std::max<int64_t>(nq_slots, 256);
        CU(cudaMallocHost(&ix->q_pin, (size_t)cap * ix->d * sizeof(float)));
        CU(cudaMalloc(&ix->q_raw, (size_t)cap * ix->d * sizeof(float)));
        CU(cudaMalloc(&ix->q_dev, (size_t)cap * ix->d_pad * sizeof(float)));
        if (ix->bf16 || ix->cfg.shadow_bf16) CU(cudaMalloc(&ix->q_bf16, (size_t)cap * ix->d_pad * 2));
        ix->q_cap = cap;
    }
    const int64_t need = nq * k;
    if (need > ix->out_cap) {
        cudaFree(ix->out_buf);
        cudaFreeHost(ix->out_pin);
        ix->out_buf = nullptr;
        ix->out_d = nullptr;
        ix->out_i = nullptr;
        ix->out_pin = nullptr;
        ix->out_cap = 0;
        const int64_t cap = std::max<int64_t>(need, 4096);
        CU(cudaMalloc(&ix->out_buf, (size_t)cap * (sizeof(float) + sizeof(int64_t)) + 64));
        ix->out_d = reinterpret_cast<float*>(ix->out_buf);
        ix->out_i = reinterpret_cast<int64_t*>(ix->out_buf + round_up((size_t)cap * sizeof(float), 16));
        CU(cudaMallocHost(&ix->out_pin, (size_t)cap * (sizeof(float) + sizeof(int64_t)) + 64));
        ix->out_cap = cap;
    }
    return B200R_OK;
}

static int ensure_gemm_ws(b200r_index* ix, int64_t nq, int k, cudaStream_t st) {
    const int filter = ix->bf16 ? kFilterBf16 : (ix->shadow && ix->shadow_on ? kFilterShadow : kFilterTf32);
    const size_t need = gemm_cand_bytes(ix->num_sms, kGemmMaxBlock, gemm_kp_cap(k, filter, ix->kp_boost));
    if (need > ix->gemm_cand_bytes) {
        cudaFree(ix->gemm_cand);
        ix->gemm_cand = nullptr;
        ix->gemm_cand_bytes = 0;
        CU(cudaMalloc(&ix->gemm_cand, need));
        ix->gemm_cand_bytes = need;
    }
    if (!ix->max_norm_bits) {
        CU(cudaMallocHost(&ix->fb_pin, b200r_index::kFbRing * sizeof(int)));
        for (auto& e : ix->fb_ev) CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        CU(cudaMalloc(&ix->max_norm_bits, 256));
        CU(cudaMalloc(&ix->gemm_aux, gemm_aux_bytes()));
        ix->max_norm_dirty = true;
    }
    if (ix->max_norm_dirty) {
        CU(launch_max_norm(ix->norms, ix->ntotal, ix->max_norm_bits, st));
        ix->st.n_kernel_launches++;
        ix->max_norm_dirty = false;
    }
    if (nq > ix->fail_cap) {
        cudaFree(ix->fail_dev);
        cudaFree(ix->flag_dev);
        cudaFreeHost(ix->fail_pin);
        ix->fail_dev = nullptr;
        ix->flag_dev = nullptr;
        ix->fail_pin = nullptr;
        ix->fail_cap = 0;
        const int64_t cap = std::max<int64_t>(nq, 1024);
        CU(cudaMalloc(&ix->fail_dev, (size_t)(cap + 1) * sizeof(int)));
        CU(cudaMalloc(&ix->flag_dev, (size_t)(cap + 1) * sizeof(int)));
        CU(cudaMallocHost(&ix->fail_pin, (size_t)(cap + 1) * sizeof(int)));
        ix->fail_cap = cap;
    }
    return B200R_OK;
}

// Flagged-query counts of earlier searches that have arrived by now (or all of them, waiting):
// statistics, and the K' boost for the searches that follow.
static void fold_fallback_counts(b200r_index* ix, bool wait) {
    for (int sl = 0; sl < b200r_index::kFbRing; sl++) {
        if (!ix->fb_pending[sl]) continue;
        if (wait) {
            cudaEventSynchronize(ix->fb_ev[sl]);
        } else if (cudaEventQuery(ix->fb_ev[sl]) != cudaSuccess) {
            cudaGetLastError();  // cudaErrorNotReady is not an error
            continue;
        }
        const int cnt = ix->fb_pin[sl];
        ix->fb_pending[sl] = false;
        if (cnt <= 0) continue;
        ix->st.n_gemm_fallbacks += cnt;
        const int per = std::max(1, ix->fb_per_pass[sl]);
        ix->st.bytes_scanned += (int64_t)((cnt + per - 1) / per) * ix->fb_pass_bytes[sl];
        if (!knobs().gemm_noboost)  // (tuning knob: keep K' fixed)
            ix->kp_boost = std::min(400, ix->kp_boost * 3 / 2);
    }
}

// Two-way dispatch like faiss (scan for small batches, GEMM for large ones).  faiss switches
// at nq = 20 (distance_compute_blas_threshold).  Here one GEMM pass streams the corpus at
// ~6.9 TB/s whatever nq <= 128 is, while the scan kernel is FMA-bound beyond one query
// (10M x 1024 fp32: 5.7 ms for nq = 1, 9.5-10.1 ms for nq = 2..4; GEMM 5.9-6.2 ms), so on a
// corpus large enough to hide the GEMM path's extra launches the GEMM
// takes over at nq = 2 -- and at nq = 1 when it can filter on the bf16 shadow (3.0 ms).
// Small corpora keep the scan up to nq = 4.  Both paths end in the same fp32 scoring code, so
// the choice never changes a result.
static bool auto_uses_gemm(const b200r_index* ix, int64_t nq, int64_t rows = -1) {
    const Knobs& kn = knobs();
    if (rows < 0) rows = ix->ntotal;
    const int64_t corpus_bytes = rows * ix->row_bytes;
    const bool eb = kn.gemm_min_bytes >= 0;
    int64_t gemm_min = 5;
    if (corpus_bytes >= (eb ? kn.gemm_min_bytes : kGemmSmallBatchBytes)) gemm_min = 2;
    if (ix->shadow && ix->shadow_on && corpus_bytes >= (eb ? kn.gemm_min_bytes : kGemmShadowSingleBytes)) gemm_min = 1;
    if (ix->bf16) {
        // bf16 rows: the scan unpacks every element (5.5 TB/s for one query, 1.5 TB/s for four),
        // the tensor cores take bf16 as it is (6.8 TB/s)
        if (corpus_bytes >= (eb ? kn.gemm_min_bytes : kGemmBf16SmallBatchBytes)) gemm_min = 2;
        if (corpus_bytes >= (eb ? kn.gemm_min_bytes : kGemmBf16SingleBytes)) gemm_min = 1;
    }
    if (kn.gemm_min_nq >= 0) gemm_min = kn.gemm_min_nq;
    return rows > 0 && (ix->cfg.search_path == B200R_PATH_GEMM ||
                        (ix->cfg.search_path == B200R_PATH_AUTO && nq >= gemm_min));
}

int leaf_route(b200r_index* ix, int64_t nq, int64_t rows) {
    std::lock_guard<std::mutex> lk(ix->mu);
    return auto_uses_gemm(ix, nq, rows) ? B200R_PATH_GEMM : B200R_PATH_SCAN;
}

int leaf_search(b200r_index* ix, const void* q, int64_t nq, int k, int q_mem_kind, float* D, int64_t* I,
                int out_mem_kind, void* stream, int route, const ScanExchange* xchg) {
    if (nq < 0) return fail(B200R_E_INVALID, "b200r_search: nq = %lld", (long long)nq);
    if (k <= 0) return fail(B200R_E_INVALID, "b200r_search: k must be > 0, got %d", k);  // faiss: assert k > 0
    if (nq == 0) return B200R_OK;
    if (!q || ((!D || !I) && !xchg)) return fail(B200R_E_INVALID, "b200r_search: null buffer");
    if ((q_mem_kind != B200R_HOST && q_mem_kind != B200R_DEVICE) ||
        (out_mem_kind != B200R_HOST && out_mem_kind != B200R_DEVICE))
        return fail(B200R_E_INVALID, "b200r_search: bad mem_kind");
    if (k > kMaxK) return fail(B200R_E_UNSUPPORTED, "b200r_search: k = %d exceeds the supported maximum %d", k, kMaxK);
    if (xchg && (nq > kScanMaxNQ || k > kScanMaxK || out_mem_kind != B200R_DEVICE))
        return fail(B200R_E_UNSUPPORTED, "fused exchange serves nq <= %d, k <= %d with device results", kScanMaxNQ, kScanMaxK);

    std::lock_guard<std::mutex> lk(ix->mu);
    DeviceGuard g(ix->device);
    const Knobs& kn = knobs();
    const bool any_dev = q_mem_kind == B200R_DEVICE || out_mem_kind == B200R_DEVICE;
    cudaStream_t st = (any_dev || stream) ? (cudaStream_t)stream : ix->stream;
    fold_fallback_counts(ix, false);  // may raise the K' boost used below
    const bool use_gemm = !xchg && ix->ntotal > 0 &&
                          (route == B200R_PATH_GEMM || (route == B200R_PATH_AUTO && auto_uses_gemm(ix, nq)));
    const int64_t nq_slots = use_gemm ? round_up(nq, 128) : round_up(nq, kScanMaxNQ);
    const bool filtered = use_gemm;  // answers come from a filter + re-score + check
    int rc = ensure_search_ws(ix, nq_slots, nq, k);
    if (rc != B200R_OK) return rc;
    // this search is ordered after the last writer of the rows and the last user of the workspace
    // BEFORE anything of it is enqueued (the max-norm refresh below reads the norms and rewrites a
    // word that an earlier asynchronous search may still be reading)
    if (ix->data_ev_set && ix->data_ev_stream != st) CU(cudaStreamWaitEvent(st, ix->data_ev, 0));
    if (ix->ws_ev_set && ix->ws_ev_stream != st) CU(cudaStreamWaitEvent(st, ix->ws_ev, 0));
    if (filtered) {
        rc = ensure_gemm_ws(ix, nq, k, st);
        if (rc != B200R_OK) return rc;
    }

    if (ix->timing) CU(cudaEventRecord(ix->t0, st));
    const float* q_src = (const float*)q;
    // queries that need no padding, normalisation or zero-filled spare slots skip the prep kernel:
    // device ones are used in place, host ones are copied straight into the prepared-query buffer
    const bool q_plain = !use_gemm && ix->d == ix->d_pad && !ix->cfg.normalize_queries &&
                         (nq == 1 || nq % kScanMaxNQ == 0);
    const bool q_in_place = q_plain && q_mem_kind == B200R_DEVICE && (reinterpret_cast<uintptr_t>(q) & 15) == 0;
    const bool q_direct = q_plain && q_mem_kind == B200R_HOST;
    if (q_mem_kind == B200R_HOST) {
        memcpy(ix->q_pin, q, (size_t)nq * ix->d * sizeof(float));
        float* dst = q_direct ? ix->q_dev : ix->q_raw;
        CU(cudaMemcpyAsync(dst, ix->q_pin, (size_t)nq * ix->d * sizeof(float), cudaMemcpyHostToDevice, st));
        q_src = dst;
    }
    const float* q_ready = ix->q_dev;
    if (q_in_place) {
        q_ready = (const float*)q;
    } else if (q_direct) {
        // nothing to do: q_dev holds the queries
    } else {
        CU(launch_prep_queries(q_src, nq, ix->d, ix->d_pad, nq_slots, ix->cfg.normalize_queries != 0, ix->q_dev, st));
        ix->st.n_kernel_launches++;
    }

    // host outputs: scores and ids packed for this nq*k so that one copy brings both back
    const size_t ids_off = round_up((size_t)nq * k * sizeof(float), 16);
    // Small host results are written by the final writer straight into the pinned host buffer (mapped
    // memory, posted stores over PCIe): no separate device-to-host copy sits between the kernel and the
    // caller.  Larger ones go through device memory and one DMA.  (B200R_ZERO_COPY_OUT=0 switches it off.)
    const bool zc_out = out_mem_kind == B200R_HOST && !use_gemm && kn.zero_copy_out != 0 &&
                        (size_t)nq * k * 12 <= (size_t)16 << 10;
    uint8_t* host_stage = zc_out ? ix->out_pin : ix->out_buf;
    float* od = out_mem_kind == B200R_DEVICE ? D : reinterpret_cast<float*>(host_stage);
    int64_t* oi = out_mem_kind == B200R_DEVICE ? I : reinterpret_cast<int64_t*>(host_stage + ids_off);
    if (xchg) {  // results go to the exchange slots; out_d / out_i stay unused placeholders
        od = ix->out_d;
        oi = ix->out_i;
    }

    const bool use_shadow = use_gemm && ix->shadow && ix->shadow_on;
    int* fail_count = filtered ? ix->fail_dev + ix->fail_cap : nullptr;
    if (filtered) CU(cudaMemsetAsync(fail_count, 0, sizeof(int), st));
    if (use_gemm) {
        if (ix->bf16 || use_shadow) {
            CU(launch_f32_to_bf16(ix->q_dev, ix->q_bf16, nq_slots * ix->d_pad, st));
            ix->st.n_kernel_launches++;
        }
        for (int64_t q0 = 0; q0 < nq; q0 += kGemmMaxBlock) {
            GemmSearchArgs a{};
            a.corpus = ix->rows;
            a.norms = ix->norms;
            a.n_rows = ix->ntotal;
            a.d_pad = ix->d_pad;
            a.bf16 = ix->bf16;
            a.l2 = ix->l2;
            a.q_f32 = ix->q_dev + (size_t)q0 * ix->d_pad;
            a.shadow = use_shadow ? ix->shadow : nullptr;
            a.kp_boost_pct = ix->kp_boost;
            a.q_mma = (ix->bf16 || use_shadow) ? (const void*)((const uint8_t*)ix->q_bf16 + (size_t)q0 * ix->d_pad * 2)
                                               : (const void*)a.q_f32;
            a.nq = (int)std::min<int64_t>(kGemmMaxBlock, nq - q0);
            a.k = k;
            a.cand = ix->gemm_cand;
            a.max_norm_bits = ix->max_norm_bits;
            gemm_aux_layout(ix->gemm_aux, &a.hist, &a.qscale, &a.gthr, &a.list_counts, &a.overflow);
            a.out_d = od + (size_t)q0 * k;
            a.out_i = oi + (size_t)q0 * k;
            a.id_base = ix->cfg.id_base;
            a.fail_flags = ix->fail_dev + q0;
            a.fail_count = fail_count;
            a.num_sms = std::min(ix->num_sms, kGemmMaxGrid);
            CU(launch_gemm_search(a, st));
            ix->st.n_kernel_launches += 3;
            ix->st.bytes_scanned += (int64_t)ix->ntotal * (use_shadow ? ix->d_pad * 2 : ix->row_bytes);
        }
        ix->st.last_path = B200R_PATH_GEMM;
        ix->st.last_grid = ix->num_sms;
    }
    ScanParams p{};
    p.corpus = ix->rows;
    p.n_rows = ix->ntotal;
    p.d = ix->d_pad;
    p.row_bytes = ix->row_bytes;
    p.k = k;
    p.evict_first = ((size_t)ix->ntotal * ix->row_bytes > (size_t)96 << 20) ? 1 : 0;
    if (kn.evict_first >= 0) p.evict_first = kn.evict_first;  // tuning knob
    p.scratch = ix->scratch;
    p.counter = ix->counter;
    // Tile assignment (B200R_SCAN_DYNAMIC; tools/perf_scan_tail.py).  0: static, tile t -> CTA t mod grid.
    // 1: every tile is drawn from a global counter (second word group of the ticket buffer): the CTAs finish
    // within 2 us of each other instead of 20 (1.25M rows) .. 110 us (10M rows), but the stream itself gets
    // 1.7 % slower on 10M rows.  2 (default): static for the first 7/8 of the tiles, the tail from the counter:
    // levels the finish without the cost -- 1.25M rows 716 -> 706 us, 10M rows 5573 -> 5558 us.
    p.tile_counter = scan_dynamic_mode() >= 1 ? ix->counter + 16 : nullptr;
    p.dyn_from = -1;  // resolved by launch_scan once the tile count is known (scan_dynamic == 2: static head)
    p.tile_block = kn.scan_tile_block;
    p.id_base = ix->cfg.id_base;
    if (kn.scan_trace) {
        if (!ix->trace_dev) CU(cudaMalloc(&ix->trace_dev, 256 * 8 * sizeof(unsigned long long)));
        CU(cudaMemsetAsync(ix->trace_dev, 0, 256 * 8 * sizeof(unsigned long long), st));
        p.trace = ix->trace_dev;
    }
    if (xchg) {
        p.xchg_ntargets = xchg->n_targets;
        p.xchg_rank = xchg->slot;
        p.xchg_epoch = xchg->epoch;
        for (int r = 0; r < xchg->n_targets; r++) p.xchg_peer[r] = xchg->target[r];
    }
    // one query per launch for k > 128 (wide lists) and for rows so wide that four queries do not
    // fit in shared memory beside the stages
    const int scan_step = k > kScanMaxK ? 1 : scan_queries_per_launch(ix->d_pad, ix->row_bytes, ix->smem_optin);
    if (xchg && nq > scan_step)
        return fail(B200R_E_UNSUPPORTED, "fused exchange: rows of %d bytes leave room for one query per launch", ix->row_bytes);
    if (!filtered) {
        for (int64_t q0 = 0; q0 < nq; q0 += scan_step) {
            const int nqg = (int)std::min<int64_t>(scan_step, nq - q0);
            p.q = q_ready + (size_t)q0 * ix->d_pad;
            p.out_d = od + (size_t)q0 * k;
            p.out_i = oi + (size_t)q0 * k;
            int grid = 0;
            CU(launch_scan(p, nqg, ix->l2, ix->bf16, ix->num_sms, ix->smem_optin, st, &grid));
            ix->st.n_kernel_launches++;
            ix->st.last_grid = grid;
            ix->st.bytes_scanned += (int64_t)ix->ntotal * ix->row_bytes;
            if (kn.scan_trace) {  // diagnosis: where the time of this launch went (synchronises)
                std::vector<unsigned long long> tr((size_t)grid * 8);
                CU(cudaStreamSynchronize(st));
                CU(cudaMemcpy(tr.data(), ix->trace_dev, tr.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
                unsigned long long t0 = ~0ull, t1max = 0, t2min = ~0ull, t2max = 0, t3max = 0, t4 = 0;
                for (int c = 0; c < grid; c++) {
                    t0 = std::min(t0, tr[c * 8]);
                    t1max = std::max(t1max, tr[c * 8 + 1]);
                    t2min = std::min(t2min, tr[c * 8 + 2]);
                    t2max = std::max(t2max, tr[c * 8 + 2]);
                    t3max = std::max(t3max, tr[c * 8 + 3]);
                    t4 = std::max(t4, tr[c * 8 + 4]);
                }
                fprintf(stderr,
                        "[b200r scan trace] grid %d rows %lld: queries staged +%.1f us, first CTA out of rows +%.1f, last "
                        "+%.1f, last CTA list stored +%.1f, result written +%.1f\n",
                        grid, (long long)ix->ntotal, (t1max - t0) / 1e3, (t2min - t0) / 1e3, (t2max - t0) / 1e3,
                        (t3max - t0) / 1e3, (t4 - t0) / 1e3);
            }
        }
    } else {
        // Queries whose containment check failed are re-run on the scan kernel WITHOUT a trip to the
        // host: a compaction kernel lists them and one cooperative launch serves them scan_step per
        // corpus pass (it returns at once when the list is empty, the common case).  How many there
        // were comes back asynchronously, for the statistics and the K' boost only.
        p.q = q_ready;
        p.out_d = od;
        p.out_i = oi;
        CU(launch_scan_flagged(p, ix->fail_dev, nq, ix->flag_dev, ix->flag_dev + ix->fail_cap, ix->l2, ix->bf16,
                               ix->num_sms, ix->smem_optin, st));
        ix->st.n_kernel_launches += 2;
        int sl = ix->fb_next;
        ix->fb_next = (sl + 1) % b200r_index::kFbRing;
        if (ix->fb_pending[sl]) {  // the ring is full of unread counts: take the oldest now
            cudaEventSynchronize(ix->fb_ev[sl]);
            fold_fallback_counts(ix, false);
        }
        CU(cudaMemcpyAsync(ix->fb_pin + sl, fail_count, sizeof(int), cudaMemcpyDeviceToHost, st));
        CU(cudaEventRecord(ix->fb_ev[sl], st));
        ix->fb_pending[sl] = true;
        ix->fb_pass_bytes[sl] = (int64_t)ix->ntotal * ix->row_bytes;
        ix->fb_per_pass[sl] = scan_step;
        if (kn.debug_fails) {  // diagnosis: read the per-query flags back (synchronises)
            CU(cudaMemcpyAsync(ix->fail_pin, ix->fail_dev, (size_t)(ix->fail_cap + 1) * sizeof(int), cudaMemcpyDeviceToHost,
                               st));
            CU(cudaStreamSynchronize(st));
            if (ix->fail_pin[ix->fail_cap]) {
                int cause[4] = {0, 0, 0, 0};
                for (int64_t i = 0; i < nq; i++) cause[ix->fail_pin[i] & 3]++;
                fprintf(stderr, "[b200r] gemm fallbacks: check=%d bucket=%d slice=%d; first failing queries:", cause[1],
                        cause[2], cause[3]);
                int shown = 0;
                for (int64_t i = 0; i < nq && shown < 40; i++)
                    if (ix->fail_pin[i]) {
                        fprintf(stderr, " %lld", (long long)i);
                        shown++;
                    }
                fprintf(stderr, "\n");
            }
        }
    }
    if (!filtered) ix->st.last_path = B200R_PATH_SCAN;
    if (ix->timing) CU(cudaEventRecord(ix->t1, st));

    if (out_mem_kind == B200R_HOST) {
        float* pd = (float*)ix->out_pin;
        int64_t* pi = (int64_t*)(ix->out_pin + ids_off);
        // out_buf / out_pin hold cap*(4+8)+64 bytes, so the 16-byte rounding always fits
        if (!zc_out)
            CU(cudaMemcpyAsync(ix->out_pin, ix->out_buf, ids_off + (size_t)nq * k * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        memcpy(D, pd, (size_t)nq * k * sizeof(float));
        memcpy(I, pi, (size_t)nq * k * sizeof(int64_t));
    }
    if (ix->timing && !xchg && !ix->timing_deferred) {
        CU(cudaEventSynchronize(ix->t1));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, ix->t0, ix->t1));
        ix->st.last_search_ms = ms;
    }
    if (out_mem_kind == B200R_DEVICE) {  // still in flight on the caller's stream
        // (host queries were staged through the index's pinned buffer: it must be free before the next call)
        // (with an exchange the caller synchronises after its own select kernel)
        if (q_mem_kind == B200R_HOST && !xchg) CU(cudaStreamSynchronize(st));
        CU(cudaEventRecord(ix->ws_ev, st));
        ix->ws_ev_set = true;
        ix->ws_ev_stream = st;
    } else {
        ix->ws_ev_set = false;  // host results: the stream was synchronised above
    }
    ix->st.n_search_calls++;
    ix->st.n_queries += nq;
    return B200R_OK;
}

// with timing on and results in flight (multi-GPU search): device time of the last search once it has finished
int leaf_last_search_ms(b200r_index* ix, double* ms_out) {
    std::lock_guard<std::mutex> lk(ix->mu);
    DeviceGuard g(ix->device);
    CU(cudaEventSynchronize(ix->t1));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, ix->t0, ix->t1));
    ix->st.last_search_ms = ms;
    *ms_out = ms;
    return B200R_OK;
}

// ---------------------------------------------------------------------------
// remove_ids / reconstruct
// ---------------------------------------------------------------------------
int leaf_remove_ids(b200r_index* ix, const int64_t* ids, int64_t n, int64_t* n_removed) {
    if (n < 0 || (n > 0 && !ids)) return fail(B200R_E_INVALID, "b200r_remove_ids: bad ids");
    if (n_removed) *n_removed = 0;
    std::lock_guard<std::mutex> lk(ix->mu);
    std::vector<int64_t> del;
    del.reserve((size_t)n);
    for (int64_t i = 0; i < n; i++)
        if (ids[i] >= 0 && ids[i] < ix->ntotal) del.push_back(ids[i]);
    std::sort(del.begin(), del.end());
    del.erase(std::unique(del.begin(), del.end()), del.end());
    const int64_t m = (int64_t)del.size();
    if (m == 0) return B200R_OK;
    const int64_t new_total = ix->ntotal - m;
    const int64_t first = del[0];  // rows before it stay where they are
    std::vector<int64_t> adj((size_t)m);
    for (int64_t i = 0; i < m; i++) adj[(size_t)i] = del[(size_t)i] - i;

    DeviceGuard g(ix->device);
    cudaStream_t st = ix->stream;
    CU(cudaDeviceSynchronize());  // rows move in place: nothing may be reading or writing them
    ix->data_ev_set = false;
    // the bounce buffer and the id list stay with the index (a chain server deletes document by document)
    const int64_t chunk_rows = std::max<int64_t>(1, std::min<int64_t>((int64_t)((64u << 20) / ix->row_bytes),
                                                                       std::max<int64_t>(1, new_total - first)));
    const size_t need_bounce = (size_t)chunk_rows * ix->row_bytes;
    if (need_bounce > ix->bounce_bytes) {
        cudaFree(ix->bounce);
        ix->bounce = nullptr;
        ix->bounce_bytes = 0;
        cudaError_t e = cudaMalloc(&ix->bounce, need_bounce);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return fail(B200R_E_NOMEM, "b200r_remove_ids: bounce buffer: %s", cudaGetErrorString(e));
        }
        ix->bounce_bytes = need_bounce;
    }
    if (m > ix->adj_cap) {
        cudaFree(ix->adj_dev);
        ix->adj_dev = nullptr;
        ix->adj_cap = 0;
        const int64_t cap = std::max<int64_t>(m, 1024);
        CU(cudaMalloc(&ix->adj_dev, (size_t)cap * sizeof(int64_t)));
        ix->adj_cap = cap;
    }
    uint8_t* bounce = ix->bounce;
    int64_t* adj_dev = ix->adj_dev;
    CU(cudaMemcpyAsync(adj_dev, adj.data(), (size_t)m * sizeof(int64_t), cudaMemcpyHostToDevice, st));
    for (int64_t o0 = first; o0 < new_total; o0 += chunk_rows) {
        const int64_t o1 = std::min(new_total, o0 + chunk_rows);
        // rows: gather the survivors of this output range, then put them in place.
        // Sources of later ranges lie above o1, so they are still intact.
        CU(launch_gather_rows(ix->rows, bounce, o0, o1, ix->row_bytes, adj_dev, m, st));
        CU(cudaMemcpyAsync(ix->rows + (size_t)o0 * ix->row_bytes, bounce, (size_t)(o1 - o0) * ix->row_bytes,
                           cudaMemcpyDeviceToDevice, st));
        if (ix->shadow) {
            CU(launch_gather_rows(ix->shadow, bounce, o0, o1, ix->d_pad * 2, adj_dev, m, st));
            CU(cudaMemcpyAsync(ix->shadow + (size_t)o0 * ix->d_pad * 2, bounce, (size_t)(o1 - o0) * ix->d_pad * 2,
                               cudaMemcpyDeviceToDevice, st));
        }
        CU(launch_gather_rows(ix->norms, bounce, o0, o1, (int)sizeof(float), adj_dev, m, st));
        CU(cudaMemcpyAsync(ix->norms + o0, bounce, (size_t)(o1 - o0) * sizeof(float), cudaMemcpyDeviceToDevice, st));
        ix->st.n_kernel_launches += 2;
    }
    CU(cudaStreamSynchronize(st));
    ix->ntotal = new_total;
    ix->max_norm_dirty = true;
    ix->st.n_rows_removed += m;
    if (n_removed) *n_removed = m;
    return B200R_OK;
}

int leaf_reconstruct(b200r_index* ix, int64_t key, float* out) {
    std::lock_guard<std::mutex> lk(ix->mu);
    if (key < 0 || key >= ix->ntotal)
        return fail(B200R_E_INVALID, "b200r_reconstruct: key %lld out of range [0, %lld)", (long long)key,
                    (long long)ix->ntotal);
    DeviceGuard g(ix->device);
    if (ix->data_ev_set) CU(cudaStreamWaitEvent(ix->stream, ix->data_ev, 0));
    if (!ix->recon_dev) {  // one row of device + pinned staging, kept with the index
        CU(cudaMalloc(&ix->recon_dev, (size_t)ix->d * sizeof(float)));
        CU(cudaMallocHost(&ix->recon_pin, (size_t)ix->d * sizeof(float)));
    }
    CU(launch_row_to_f32(ix->rows + (size_t)key * ix->row_bytes, ix->d, ix->bf16, ix->recon_dev, ix->stream));
    ix->st.n_kernel_launches++;
    CU(cudaMemcpyAsync(ix->recon_pin, ix->recon_dev, (size_t)ix->d * sizeof(float), cudaMemcpyDeviceToHost, ix->stream));
    CU(cudaStreamSynchronize(ix->stream));
    memcpy(out, ix->recon_pin, (size_t)ix->d * sizeof(float));
    return B200R_OK;
}

int leaf_reconstruct_batch(b200r_index* ix, const int64_t* keys, int64_t n, float* out) {
    if (n < 0 || (n > 0 && (!keys || !out))) return fail(B200R_E_INVALID, "b200r_reconstruct_batch: bad arguments");
    if (n == 0) return B200R_OK;
    std::lock_guard<std::mutex> lk(ix->mu);
    for (int64_t i = 0; i < n; i++)
        if (keys[i] < 0 || keys[i] >= ix->ntotal)
            return fail(B200R_E_INVALID, "b200r_reconstruct_batch: key %lld out of range [0, %lld)", (long long)keys[i],
                        (long long)ix->ntotal);
    DeviceGuard g(ix->device);
    if (ix->data_ev_set) cudaStreamWaitEvent(ix->stream, ix->data_ev, 0);
    int64_t* kd = nullptr;
    float* od = nullptr;
    cudaError_t e = cudaMalloc(&kd, (size_t)n * sizeof(int64_t));
    if (e == cudaSuccess) e = cudaMalloc(&od, (size_t)n * ix->d * sizeof(float));
    if (e == cudaSuccess) e = cudaMemcpyAsync(kd, keys, (size_t)n * sizeof(int64_t), cudaMemcpyHostToDevice, ix->stream);
    if (e == cudaSuccess) e = launch_rows_to_f32(ix->rows, ix->row_bytes, ix->d, ix->bf16, kd, n, od, ix->stream);
    if (e == cudaSuccess)
        e = cudaMemcpyAsync(out, od, (size_t)n * ix->d * sizeof(float), cudaMemcpyDeviceToHost, ix->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ix->stream);
    cudaFree(kd);
    cudaFree(od);
    ix->st.n_kernel_launches++;
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(e == cudaErrorMemoryAllocation ? B200R_E_NOMEM : B200R_E_CUDA, "b200r_reconstruct_batch: %s",
                    cudaGetErrorString(e));
    }
    return B200R_OK;
}

// rows [r0, r0+n) as fp32 [n][d] into a device buffer on this GPU; synchronised on return
int leaf_rows_to_device_f32(b200r_index* ix, int64_t r0, int64_t n, float* dev_out) {
    std::lock_guard<std::mutex> lk(ix->mu);
    if (r0 < 0 || n < 0 || r0 + n > ix->ntotal) return fail(B200R_E_INVALID, "row range out of bounds");
    if (n == 0) return B200R_OK;
    DeviceGuard g(ix->device);
    if (ix->data_ev_set) CU(cudaStreamWaitEvent(ix->stream, ix->data_ev, 0));
    CU(launch_rows_to_f32(ix->rows + (size_t)r0 * ix->row_bytes, ix->row_bytes, ix->d, ix->bf16, nullptr, n, dev_out,
                          ix->stream));
    ix->st.n_kernel_launches++;
    CU(cudaStreamSynchronize(ix->stream));
    return B200R_OK;
}

// rows [r0, r0+n) as fp32 without the padding columns into a host buffer, in bounded chunks
int leaf_read_rows(b200r_index* ix, int64_t r0, int64_t n, float* out) {
    std::lock_guard<std::mutex> lk(ix->mu);
    if (r0 < 0 || n < 0 || r0 + n > ix->ntotal) return fail(B200R_E_INVALID, "row range out of bounds");
    if (n == 0) return B200R_OK;
    DeviceGuard g(ix->device);
    if (ix->data_ev_set) CU(cudaStreamWaitEvent(ix->stream, ix->data_ev, 0));
    const int d = ix->d;
    const bool direct = !ix->bf16 && ix->d_pad == ix->d;  // stored rows are already packed fp32
    const int64_t chunk = std::max<int64_t>(1, (int64_t)((32u << 20) / ((size_t)d * 4)));
    float* dev = nullptr;
    cudaError_t e = direct ? cudaSuccess : cudaMalloc(&dev, (size_t)std::min(chunk, n) * d * 4);
    for (int64_t c0 = 0; e == cudaSuccess && c0 < n; c0 += chunk) {
        const int64_t m = std::min(chunk, n - c0);
        const uint8_t* src = ix->rows + (size_t)(r0 + c0) * ix->row_bytes;
        if (direct) {
            e = cudaMemcpyAsync(out + (size_t)c0 * d, src, (size_t)m * d * 4, cudaMemcpyDeviceToHost, ix->stream);
        } else {
            e = launch_rows_to_f32(src, ix->row_bytes, d, ix->bf16, nullptr, m, dev, ix->stream);
            ix->st.n_kernel_launches++;
            if (e == cudaSuccess)
                e = cudaMemcpyAsync(out + (size_t)c0 * d, dev, (size_t)m * d * 4, cudaMemcpyDeviceToHost, ix->stream);
        }
        if (e == cudaSuccess) e = cudaStreamSynchronize(ix->stream);
    }
    cudaFree(dev);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(e == cudaErrorMemoryAllocation ? B200R_E_NOMEM : B200R_E_CUDA, "reading rows back: %s",
                    cudaGetErrorString(e));
    }
    return B200R_OK;
}

// ---------------------------------------------------------------------------
// merge_from: append another index's rows, empty the other index
// ---------------------------------------------------------------------------
int leaf_merge_from(b200r_index* dst, b200r_index* src) {
    // faiss check_compatible_for_merge: same d, metric and index type (here also the storage and the GPU)
    if (dst->d != src->d || dst->l2 != src->l2 || dst->bf16 != src->bf16)
        return fail(B200R_E_INVALID, "b200r_merge_from: indexes are not compatible (dim %d/%d, metric or storage differ)",
                    dst->d, src->d);
    if (dst->device != src->device)
        return fail(B200R_E_UNSUPPORTED, "b200r_merge_from: indexes live on different GPUs (%d, %d)", dst->device, src->device);
    std::scoped_lock lk(dst->mu, src->mu);
    const int64_t n = src->ntotal;
    if (n == 0) return B200R_OK;
    DeviceGuard g(dst->device);
    cudaStream_t st = dst->stream;
    if (src->data_ev_set) CU(cudaStreamWaitEvent(st, src->data_ev, 0));
    if (dst->data_ev_set) CU(cudaStreamWaitEvent(st, dst->data_ev, 0));
    int rc = ensure_capacity(dst, dst->ntotal + n, st);
    if (rc != B200R_OK) return rc;
    // The stored rows go through fp32 and the ingest kernel again (without normalisation): exact
    // for both storages (bf16 -> fp32 -> bf16 is the identity), it re-pads rows whose padding
    // differs between the two indexes and fills the norms and the bf16 shadow in the same pass.
    const int64_t chunk = std::max<int64_t>(1, (int64_t)((64u << 20) / ((size_t)dst->d * sizeof(float))));
    float* tmp = nullptr;
    cudaError_t e = cudaMalloc(&tmp, (size_t)std::min(chunk, n) * dst->d * sizeof(float));
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(B200R_E_NOMEM, "b200r_merge_from: staging buffer: %s", cudaGetErrorString(e));
    }
    for (int64_t r0 = 0; r0 < n && e == cudaSuccess; r0 += chunk) {
        const int64_t m = std::min(chunk, n - r0);
        e = launch_rows_to_f32(src->rows + (size_t)r0 * src->row_bytes, src->row_bytes, src->d, src->bf16, nullptr, m, tmp, st);
        if (e == cudaSuccess)
            e = launch_ingest(tmp, m, dst->d, dst->d_pad, false, dst->bf16,
                              dst->rows + (size_t)(dst->ntotal + r0) * dst->row_bytes, dst->norms + dst->ntotal + r0, st);
        if (e == cudaSuccess && dst->shadow)
            e = launch_ingest(tmp, m, dst->d, dst->d_pad, false, true,
                              dst->shadow + (size_t)(dst->ntotal + r0) * dst->d_pad * 2, nullptr, st);
        dst->st.n_kernel_launches += dst->shadow ? 3 : 2;
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    cudaFree(tmp);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return fail(B200R_E_CUDA, "b200r_merge_from: %s", cudaGetErrorString(e));
    }
    dst->ntotal += n;
    dst->max_norm_dirty = true;
    dst->st.n_rows_added += n;
    dst->data_ev_set = false;
    src->ntotal = 0;  // faiss: the other index is reset
    src->max_norm_dirty = true;
    src->data_ev_set = false;
    return B200R_OK;
}

}  // namespace b200r
