// C ABI of libb200r.so (see include/b200r.h for the reference call each entry
// point replaces).  Host-side runtime: HBM corpus buffer, pinned staging rings,
// per-index search workspace, error plumbing.  No CPU compute path exists here.
#include "../../include/b200r.h"
#include "common.cuh"
#include "kernels.h"

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

static int fail(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

namespace b200r {
void set_last_error(const char* msg) { g_err = msg ? msg : ""; }
}  // namespace b200r

#define CU(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e__ = (call);                                                                    \
        if (e__ != cudaSuccess)                                                                      \
            return fail(e__ == cudaErrorMemoryAllocation ? B200R_E_NOMEM : B200R_E_CUDA, "%s: %s (%s:%d)", #call, \
                        cudaGetErrorString(e__), __FILE__, __LINE__);                                \
    } while (0)

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

// ---------------------------------------------------------------------------
// the index
// ---------------------------------------------------------------------------
constexpr size_t kStageHalfBytes = 32u << 20;  // pinned / device staging, per half
constexpr int64_t kMaxRows = 0xFFFFFFFEll;     // ids are 32-bit inside the kernels
// corpus sizes from which small batches go to the GEMM path (measured crossovers with the device-side
// fallback in place, tools/perf_small_nq.py / perf_bf16_small.py: 2..4 queries from ~110k x 1024 fp32 rows;
// a single query on the bf16 shadow or on bf16 rows from ~250k / ~500k rows)
constexpr int64_t kGemmSmallBatchBytes = 448ll << 20;
constexpr int64_t kGemmShadowSingleBytes = 1024ll << 20;
constexpr int64_t kGemmBf16SmallBatchBytes = 64ll << 20;
constexpr int64_t kGemmBf16SingleBytes = 1024ll << 20;

struct b200r_index {
    b200r_config cfg{};
    int d = 0, d_pad = 0, row_bytes = 0;
    bool bf16 = false, l2 = false;
    int device = 0, num_sms = 0;
    size_t smem_optin = 0;

    uint8_t* rows = nullptr;  // [capacity][row_bytes]
    float* norms = nullptr;   // [capacity] squared norms of the stored rows
    // optional (cfg.shadow_bf16, fp32 storage): the same rows rounded to bf16, [capacity][d_pad];
    // batches on the GEMM path filter on it (half the bytes, bf16 MMA rate) and re-score on `rows`
    uint8_t* shadow = nullptr;
    bool shadow_on = false;   // b200r_set_shadow_filter
    int64_t ntotal = 0, capacity = 0;

    cudaStream_t stream = nullptr;  // the index's own stream
    // last writer of rows/norms: searches on another stream wait for it (add() with a device tensor
    // runs asynchronously on the caller's stream)
    cudaEvent_t data_ev = nullptr;
    bool data_ev_set = false;
    cudaStream_t data_ev_stream = nullptr;
    // last user of the search workspace (query slots, candidate scratch, tickets, ...): a search
    // enqueued on another stream waits for it, so asynchronous searches on different streams
    // never overlap on the shared buffers
    cudaEvent_t ws_ev = nullptr;
    bool ws_ev_set = false;
    cudaStream_t ws_ev_stream = nullptr;
    // add(): double-buffered host->device staging
    uint8_t* pin_add = nullptr;
    uint8_t* dev_add = nullptr;
    cudaEvent_t add_ev[2] = {nullptr, nullptr};
    // search(): workspace
    float* q_pin = nullptr;
    float* q_raw = nullptr;   // device, as given
    float* q_dev = nullptr;   // device, padded / normalised
    int64_t q_cap = 0;        // query slots allocated
    Cand* scratch = nullptr;
    unsigned* counter = nullptr;
    uint8_t* out_buf = nullptr;  // host-output searches: [scores | pad to 16 B | ids], one D2H per search
    float* out_d = nullptr;      // = out_buf
    int64_t* out_i = nullptr;    // ids of a full-capacity result (only a placeholder target otherwise)
    uint8_t* out_pin = nullptr;
    int64_t out_cap = 0;  // nq*k entries allocated
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    bool timing = false;
    // GEMM path workspace
    void* q_bf16 = nullptr;        // bf16 copy of q_dev (bf16 corpus only), q_cap rows
    Cand* gemm_cand = nullptr;
    size_t gemm_cand_bytes = 0;
    unsigned* max_norm_bits = nullptr;
    void* gemm_aux = nullptr;      // histograms, bucket maps, grid-wide thresholds
    bool max_norm_dirty = true;
    int* fail_dev = nullptr;       // [fail_cap] flags + 1 counter at the end
    int* fail_pin = nullptr;
    int64_t fail_cap = 0;
    // K' multiplier in percent: raised after a search whose containment check failed for some
    // query (corpus with many near-ties), so the next batches filter with more slack
    int kp_boost = 100;
    // device-side fallback of flagged queries (launch_scan_flagged): compacted positions + their number
    int* flag_dev = nullptr;       // [fail_cap] list + 1 count at the end
    // the number of flagged queries of recent searches comes back asynchronously (statistics and the
    // K' boost only): a small ring of pinned words, each with the event that says it has landed
    static constexpr int kFbRing = 8;
    int* fb_pin = nullptr;         // [kFbRing]
    cudaEvent_t fb_ev[kFbRing] = {};
    bool fb_pending[kFbRing] = {};
    int64_t fb_pass_bytes[kFbRing] = {};  // corpus bytes one fallback pass streams
    int fb_per_pass[kFbRing] = {};        // flagged queries one pass serves
    int fb_next = 0;

    std::mutex mu;
    b200r_stats st{};
};

static int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }
static void fold_fallback_counts(b200r_index* ix, bool wait);

// pageable -> pinned staging copy; one core moves ~12 GB/s, PCIe Gen5 takes ~50, so large
// copies are split over a few threads (at most half the cores)
static void staged_memcpy(void* dst, const void* src, size_t bytes) {
    constexpr size_t kMinPerThread = 4u << 20;
    unsigned hw = std::thread::hardware_concurrency();
    static const int max_threads = [] {
        const char* ev = getenv("B200R_STAGE_THREADS");  // tuning knob
        return ev ? std::max(1, atoi(ev)) : 8;  // measured on the 16-core GPU box: 18 GB/s with 4, 24 with 8, 21 with 12
    }();
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

static void free_index(b200r_index* ix) {
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

extern "C" int b200r_version(void) { return B200R_VERSION; }
extern "C" const char* b200r_last_error(void) { return g_err.c_str(); }

extern "C" int b200r_create(const b200r_config* cfg, b200r_index** out) {
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
            free_index(ix);                                                                       \
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
            free_index(ix);
            return rc;
        }
    }
    *out = ix;
    return B200R_OK;
}

extern "C" int b200r_destroy(b200r_index* ix) {
    free_index(ix);
    return B200R_OK;
}

extern "C" int b200r_ntotal(const b200r_index* ix, int64_t* out) {
    if (!ix || !out) return fail(B200R_E_INVALID, "b200r_ntotal: null argument");
    *out = ix->ntotal;
    return B200R_OK;
}

extern "C" int b200r_set_id_base(b200r_index* ix, int64_t id_base) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_set_id_base: null index");
    std::lock_guard<std::mutex> lk(ix->mu);
    ix->cfg.id_base = id_base;
    return B200R_OK;
}

extern "C" int b200r_set_shadow_filter(b200r_index* ix, int32_t enabled) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_set_shadow_filter: null index");
    if (!ix->cfg.shadow_bf16) return fail(B200R_E_INVALID, "index was created without shadow_bf16");
    std::lock_guard<std::mutex> lk(ix->mu);
    ix->shadow_on = enabled != 0;
    return B200R_OK;
}

extern "C" int b200r_set_timing(b200r_index* ix, int32_t enabled) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_set_timing: null index");
    ix->timing = enabled != 0;
    return B200R_OK;
}

extern "C" int b200r_stats_get(b200r_index* ix, b200r_stats* out) {
    if (!ix || !out) return fail(B200R_E_INVALID, "b200r_stats_get: null argument");
    std::lock_guard<std::mutex> lk(ix->mu);
    {
        DeviceGuard g(ix->device);
        fold_fallback_counts(ix, true);
    }
    ix->st.ntotal = ix->ntotal;
    ix->st.capacity_rows = ix->capacity;
    *out = ix->st;
    return B200R_OK;
}

extern "C" int b200r_reset(b200r_index* ix) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_reset: null index");
    std::lock_guard<std::mutex> lk(ix->mu);
    ix->ntotal = 0;
    ix->max_norm_dirty = true;
    return B200R_OK;
}

// ---------------------------------------------------------------------------
// add
// ---------------------------------------------------------------------------
extern "C" int b200r_add(b200r_index* ix, const void* x, int64_t n, int dtype, int mem_kind, void* stream) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_add: null index");
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
    int rc = ensure_capacity(ix, ix->ntotal + n, st);
    if (rc != B200R_OK) return rc;
    const bool norm = ix->cfg.normalize_on_add != 0;
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
        if (!getenv("B200R_GEMM_NOBOOST"))  // (tuning knob: keep K' fixed)
            ix->kp_boost = std::min(400, ix->kp_boost * 3 / 2);
    }
}

extern "C" int b200r_search(b200r_index* ix, const void* q, int64_t nq, int k, int q_mem_kind, float* D, int64_t* I,
                            int out_mem_kind, void* stream) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_search: null index");
    if (nq < 0) return fail(B200R_E_INVALID, "b200r_search: nq = %lld", (long long)nq);
    if (k <= 0) return fail(B200R_E_INVALID, "b200r_search: k must be > 0, got %d", k);  // faiss: assert k > 0
    if (nq == 0) return B200R_OK;
    if (!q || !D || !I) return fail(B200R_E_INVALID, "b200r_search: null buffer");
    if ((q_mem_kind != B200R_HOST && q_mem_kind != B200R_DEVICE) ||
        (out_mem_kind != B200R_HOST && out_mem_kind != B200R_DEVICE))
        return fail(B200R_E_INVALID, "b200r_search: bad mem_kind");
    if (k > kMaxK) return fail(B200R_E_UNSUPPORTED, "b200r_search: k = %d exceeds the supported maximum %d", k, kMaxK);

    std::lock_guard<std::mutex> lk(ix->mu);
    DeviceGuard g(ix->device);
    const bool any_dev = q_mem_kind == B200R_DEVICE || out_mem_kind == B200R_DEVICE;
    cudaStream_t st = (any_dev || stream) ? (cudaStream_t)stream : ix->stream;
    // Two-way dispatch like faiss (scan for small batches, GEMM for large ones).  faiss switches
    // at nq = 20 (distance_compute_blas_threshold).  Here one GEMM pass streams the corpus at
    // ~6.9 TB/s whatever nq <= 128 is, while the scan kernel is FMA-bound beyond one query
    // (10M x 1024 fp32: 5.7 ms for nq = 1, 9.5-10.1 ms for nq = 2..4; GEMM 5.9-6.2 ms), so on a
    // corpus large enough to hide the GEMM path's extra launches the GEMM
    // takes over at nq = 2 -- and at nq = 1 when it can filter on the bf16 shadow (3.0 ms).
    // Small corpora keep the scan up to nq = 4.  Both paths end in the same fp32 scoring code, so
    // the choice never changes a result.
    const char* env_thr = getenv("B200R_GEMM_MIN_NQ");
    const char* env_bytes = getenv("B200R_GEMM_MIN_BYTES");
    const int64_t corpus_bytes = (int64_t)ix->ntotal * ix->row_bytes;
    int64_t gemm_min = 5;
    if (corpus_bytes >= (env_bytes ? atoll(env_bytes) : kGemmSmallBatchBytes)) gemm_min = 2;
    if (ix->shadow && ix->shadow_on && corpus_bytes >= (env_bytes ? atoll(env_bytes) : kGemmShadowSingleBytes)) gemm_min = 1;
    if (ix->bf16) {
        // bf16 rows: the scan unpacks every element (5.5 TB/s for one query, 1.5 TB/s for four),
        // the tensor cores take bf16 as it is (6.8 TB/s)
        if (corpus_bytes >= (env_bytes ? atoll(env_bytes) : kGemmBf16SmallBatchBytes)) gemm_min = 2;
        if (corpus_bytes >= (env_bytes ? atoll(env_bytes) : kGemmBf16SingleBytes)) gemm_min = 1;
    }
    if (env_thr) gemm_min = atoll(env_thr);
    fold_fallback_counts(ix, false);  // may raise the K' boost used below
    const bool use_gemm = ix->ntotal > 0 && (ix->cfg.search_path == B200R_PATH_GEMM ||
                                             (ix->cfg.search_path == B200R_PATH_AUTO && nq >= gemm_min));
    const int64_t nq_slots = use_gemm ? round_up(nq, 128) : round_up(nq, kScanMaxNQ);
    const bool filtered = use_gemm;  // answers come from a filter + re-score + check
    int rc = ensure_search_ws(ix, nq_slots, nq, k);
    if (rc != B200R_OK) return rc;
    if (filtered) {
        rc = ensure_gemm_ws(ix, nq, k, st);
        if (rc != B200R_OK) return rc;
    }

    if (ix->data_ev_set && ix->data_ev_stream != st) CU(cudaStreamWaitEvent(st, ix->data_ev, 0));
    if (ix->ws_ev_set && ix->ws_ev_stream != st) CU(cudaStreamWaitEvent(st, ix->ws_ev, 0));
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
    float* od = out_mem_kind == B200R_DEVICE ? D : reinterpret_cast<float*>(ix->out_buf);
    int64_t* oi = out_mem_kind == B200R_DEVICE ? I : reinterpret_cast<int64_t*>(ix->out_buf + ids_off);

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
    if (const char* ev = getenv("B200R_EVICT_FIRST")) p.evict_first = atoi(ev);  // tuning knob
    p.scratch = ix->scratch;
    p.counter = ix->counter;
    p.id_base = ix->cfg.id_base;
    // one query per launch for k > 128 (wide lists) and for rows so wide that four queries do not
    // fit in shared memory beside the stages
    const int scan_step = k > kScanMaxK ? 1 : scan_queries_per_launch(ix->d_pad, ix->row_bytes, ix->smem_optin);
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
        if (getenv("B200R_DEBUG_FAILS")) {  // diagnosis: read the per-query flags back (synchronises)
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
        CU(cudaMemcpyAsync(ix->out_pin, ix->out_buf, ids_off + (size_t)nq * k * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        memcpy(D, pd, (size_t)nq * k * sizeof(float));
        memcpy(I, pi, (size_t)nq * k * sizeof(int64_t));
    }
    if (ix->timing) {
        CU(cudaEventSynchronize(ix->t1));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, ix->t0, ix->t1));
        ix->st.last_search_ms = ms;
    }
    if (out_mem_kind == B200R_DEVICE) {  // still in flight on the caller's stream
        // (host queries were staged through the index's pinned buffer: it must be free before the next call)
        if (q_mem_kind == B200R_HOST) CU(cudaStreamSynchronize(st));
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

// ---------------------------------------------------------------------------
// fused search + NVLink exchange (multi-GPU small batches)
// ---------------------------------------------------------------------------
extern "C" int64_t b200r_exchange_buffer_bytes(void) { return kXchgBytes; }

extern "C" int b200r_search_exchange(b200r_index* ix, const float* q, int64_t nq, int k, void* const* peer_bufs,
                                     int32_t world, int32_t rank, uint32_t epoch, float* D, int64_t* I, void* stream) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_search_exchange: null index");
    if (!q || !D || !I || !peer_bufs) return fail(B200R_E_INVALID, "b200r_search_exchange: null buffer");
    if (nq < 1 || nq > kScanMaxNQ)
        return fail(B200R_E_UNSUPPORTED, "b200r_search_exchange: nq %lld outside 1..%d", (long long)nq, kScanMaxNQ);
    if (k < 1 || k > kScanMaxK) return fail(B200R_E_UNSUPPORTED, "b200r_search_exchange: k %d outside 1..%d", k, kScanMaxK);
    if (world < 1 || world > kXchgMaxWorld || rank < 0 || rank >= world)
        return fail(B200R_E_INVALID, "b200r_search_exchange: bad world/rank %d/%d", world, rank);
    if (epoch == 0) return fail(B200R_E_INVALID, "b200r_search_exchange: epoch must start at 1");
    for (int r = 0; r < world; r++)
        if (!peer_bufs[r]) return fail(B200R_E_INVALID, "b200r_search_exchange: null peer buffer %d", r);
    std::lock_guard<std::mutex> lk(ix->mu);
    DeviceGuard g(ix->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t nq_slots = round_up(nq, kScanMaxNQ);
    int rc = ensure_search_ws(ix, nq_slots, nq, k);
    if (rc != B200R_OK) return rc;
    if (ix->data_ev_set && ix->data_ev_stream != st) CU(cudaStreamWaitEvent(st, ix->data_ev, 0));
    if (ix->ws_ev_set && ix->ws_ev_stream != st) CU(cudaStreamWaitEvent(st, ix->ws_ev, 0));
    CU(launch_prep_queries(q, nq, ix->d, ix->d_pad, nq_slots, ix->cfg.normalize_queries != 0, ix->q_dev, st));
    ScanParams p{};
    p.corpus = ix->rows;
    p.n_rows = ix->ntotal;
    p.d = ix->d_pad;
    p.row_bytes = ix->row_bytes;
    p.k = k;
    p.evict_first = ((size_t)ix->ntotal * ix->row_bytes > (size_t)96 << 20) ? 1 : 0;
    p.scratch = ix->scratch;
    p.counter = ix->counter;
    p.id_base = ix->cfg.id_base;
    p.q = ix->q_dev;
    p.out_d = ix->out_d;  // unused in exchange mode
    p.out_i = ix->out_i;
    p.xchg_world = world;
    p.xchg_rank = rank;
    p.xchg_epoch = epoch;
    for (int r = 0; r < world; r++) p.xchg_peer[r] = reinterpret_cast<uint8_t*>(peer_bufs[r]);
    int grid = 0;
    CU(launch_scan(p, (int)nq, ix->l2, ix->bf16, ix->num_sms, ix->smem_optin, st, &grid));
    CU(launch_merge_wait(ix->l2, reinterpret_cast<uint8_t*>(peer_bufs[rank]), world, (int)nq, k, epoch, D, I, st));
    CU(cudaEventRecord(ix->ws_ev, st));
    ix->ws_ev_set = true;
    ix->ws_ev_stream = st;
    ix->st.n_kernel_launches += 3;
    ix->st.last_grid = grid;
    ix->st.last_path = B200R_PATH_SCAN;
    ix->st.bytes_scanned += (int64_t)ix->ntotal * ix->row_bytes;
    ix->st.n_search_calls++;
    ix->st.n_queries += nq;
    return B200R_OK;
}

// ---------------------------------------------------------------------------
// remove_ids / reconstruct
// ---------------------------------------------------------------------------
extern "C" int b200r_remove_ids(b200r_index* ix, const int64_t* ids, int64_t n, int64_t* n_removed) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_remove_ids: null index");
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
    int64_t* adj_dev = nullptr;
    uint8_t* bounce = nullptr;
    const int64_t chunk_rows = std::max<int64_t>(1, (int64_t)((64u << 20) / ix->row_bytes));
    CU(cudaMalloc(&adj_dev, (size_t)m * sizeof(int64_t)));
    cudaError_t e = cudaMalloc(&bounce, (size_t)chunk_rows * ix->row_bytes);
    if (e != cudaSuccess) {
        cudaFree(adj_dev);
        cudaGetLastError();
        return fail(B200R_E_NOMEM, "b200r_remove_ids: bounce buffer: %s", cudaGetErrorString(e));
    }
    int rc = B200R_OK;
    auto body = [&]() -> int {
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
        return B200R_OK;
    };
    rc = body();
    cudaFree(adj_dev);
    cudaFree(bounce);
    if (rc != B200R_OK) return rc;
    ix->ntotal = new_total;
    ix->max_norm_dirty = true;
    ix->st.n_rows_removed += m;
    if (n_removed) *n_removed = m;
    return B200R_OK;
}

extern "C" int b200r_reconstruct(b200r_index* ix, int64_t key, float* out) {
    if (!ix || !out) return fail(B200R_E_INVALID, "b200r_reconstruct: null argument");
    std::lock_guard<std::mutex> lk(ix->mu);
    if (key < 0 || key >= ix->ntotal)
        return fail(B200R_E_INVALID, "b200r_reconstruct: key %lld out of range [0, %lld)", (long long)key,
                    (long long)ix->ntotal);
    DeviceGuard g(ix->device);
    if (ix->data_ev_set) CU(cudaStreamWaitEvent(ix->stream, ix->data_ev, 0));
    float* tmp = nullptr;
    CU(cudaMalloc(&tmp, (size_t)ix->d * sizeof(float)));
    cudaError_t e = launch_row_to_f32(ix->rows + (size_t)key * ix->row_bytes, ix->d, ix->bf16, tmp, ix->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, tmp, (size_t)ix->d * sizeof(float), cudaMemcpyDeviceToHost, ix->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ix->stream);
    cudaFree(tmp);
    ix->st.n_kernel_launches++;
    if (e != cudaSuccess) return fail(B200R_E_CUDA, "b200r_reconstruct: %s", cudaGetErrorString(e));
    return B200R_OK;
}

extern "C" int b200r_reconstruct_batch(b200r_index* ix, const int64_t* keys, int64_t n, float* out) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_reconstruct_batch: null index");
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

// ---------------------------------------------------------------------------
// merge_from: append another index's rows, empty the other index
// ---------------------------------------------------------------------------
extern "C" int b200r_merge_from(b200r_index* dst, b200r_index* src) {
    if (!dst || !src) return fail(B200R_E_INVALID, "b200r_merge_from: null index");
    if (dst == src) return fail(B200R_E_INVALID, "b200r_merge_from: an index cannot be merged into itself");
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

// ---------------------------------------------------------------------------
// stand-alone normalise and merge
// ---------------------------------------------------------------------------
extern "C" int b200r_normalize_l2(float* x, int64_t n, int32_t d, int mem_kind, int32_t device, void* stream) {
    if (n < 0 || d <= 0) return fail(B200R_E_INVALID, "b200r_normalize_l2: bad shape");
    if (n == 0) return B200R_OK;
    if (!x) return fail(B200R_E_INVALID, "b200r_normalize_l2: null data");
    DeviceGuard g(device);
    if (!g.ok) return fail(B200R_E_CUDA, "cudaSetDevice(%d) failed", device);
    cudaStream_t st = (cudaStream_t)stream;
    if (mem_kind == B200R_DEVICE) {
        CU(launch_normalize(x, n, d, st));
        return B200R_OK;
    }
    if (mem_kind != B200R_HOST) return fail(B200R_E_INVALID, "b200r_normalize_l2: bad mem_kind");
    const size_t row = (size_t)d * sizeof(float);
    const int64_t chunk = std::max<int64_t>(1, (int64_t)((64u << 20) / row));
    float* dev = nullptr;
    CU(cudaMalloc(&dev, (size_t)std::min(chunk, n) * row));
    cudaError_t e = cudaSuccess;
    for (int64_t r0 = 0; r0 < n && e == cudaSuccess; r0 += chunk) {
        const int64_t m = std::min(chunk, n - r0);
        e = cudaMemcpyAsync(dev, x + r0 * d, (size_t)m * row, cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = launch_normalize(dev, m, d, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(x + r0 * d, dev, (size_t)m * row, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    }
    cudaFree(dev);
    if (e != cudaSuccess) return fail(B200R_E_CUDA, "b200r_normalize_l2: %s", cudaGetErrorString(e));
    return B200R_OK;
}

extern "C" int b200r_merge_topk(int32_t metric, const float* cand_D, const int64_t* cand_I, int64_t nq, int32_t n_lists,
                                int32_t k_in, int64_t stride_D, int64_t stride_I, int32_t k_out, float* D, int64_t* I,
                                int32_t device, void* stream) {
    if (metric != B200R_METRIC_INNER_PRODUCT && metric != B200R_METRIC_L2)
        return fail(B200R_E_INVALID, "b200r_merge_topk: bad metric");
    if (nq < 0 || n_lists < 1 || k_in < 1 || k_out < 1 || stride_D < 0 || stride_I < 0)
        return fail(B200R_E_INVALID, "b200r_merge_topk: bad shape");
    if (k_out > kMaxK) return fail(B200R_E_UNSUPPORTED, "b200r_merge_topk: k_out %d > %d", k_out, kMaxK);
    if (nq == 0) return B200R_OK;
    if (!cand_D || !cand_I || !D || !I) return fail(B200R_E_INVALID, "b200r_merge_topk: null buffer");
    DeviceGuard g(device);
    if (!g.ok) return fail(B200R_E_CUDA, "cudaSetDevice(%d) failed", device);
    CU(launch_merge(metric == B200R_METRIC_L2, cand_D, cand_I, nq, n_lists, k_in, stride_D, stride_I, k_out, D, I,
                    (cudaStream_t)stream));
    return B200R_OK;
}

// ---------------------------------------------------------------------------
// save / load: faiss flat-index file ("IxF2" / "IxFI"), fp32 payload
//   u32 fourcc | i32 d | i64 ntotal | i64 dummy(1<<20) x2 | u8 is_trained | i32 metric |
//   u64 n_floats | fp32 data          (faiss/impl/index_write.cpp write_index_header +
//   WRITEXBVECTOR(codes))
// ---------------------------------------------------------------------------
static uint32_t fourcc(const char* s) {
    return (uint32_t)(uint8_t)s[0] | ((uint32_t)(uint8_t)s[1] << 8) | ((uint32_t)(uint8_t)s[2] << 16) |
           ((uint32_t)(uint8_t)s[3] << 24);
}

extern "C" int b200r_save(b200r_index* ix, const char* path) {
    if (!ix || !path) return fail(B200R_E_INVALID, "b200r_save: null argument");
    std::lock_guard<std::mutex> lk(ix->mu);
    DeviceGuard g(ix->device);
    if (ix->data_ev_set) cudaStreamWaitEvent(ix->stream, ix->data_ev, 0);
    FILE* f = fopen(path, "wb");
    if (!f) return fail(B200R_E_IO, "b200r_save: cannot open %s", path);
    const uint32_t h = fourcc(ix->l2 ? "IxF2" : "IxFI");
    const int32_t d = ix->d;
    const int64_t nt = ix->ntotal, dummy = 1 << 20;
    const uint8_t trained = 1;
    const int32_t metric = ix->cfg.metric;
    const uint64_t nfloats = (uint64_t)nt * (uint64_t)d;
    bool ok = fwrite(&h, 4, 1, f) == 1 && fwrite(&d, 4, 1, f) == 1 && fwrite(&nt, 8, 1, f) == 1 &&
              fwrite(&dummy, 8, 1, f) == 1 && fwrite(&dummy, 8, 1, f) == 1 && fwrite(&trained, 1, 1, f) == 1 &&
              fwrite(&metric, 4, 1, f) == 1 && fwrite(&nfloats, 8, 1, f) == 1;
    // rows go out as fp32 without the padding columns, in bounded chunks
    const int64_t chunk = std::max<int64_t>(1, (int64_t)((32u << 20) / ((size_t)d * 4)));
    float* dev = nullptr;
    float* host = nullptr;
    cudaError_t e = cudaMalloc(&dev, (size_t)chunk * d * 4);
    if (e == cudaSuccess) e = cudaMallocHost(&host, (size_t)chunk * d * 4);
    for (int64_t r0 = 0; ok && e == cudaSuccess && r0 < nt; r0 += chunk) {
        const int64_t m = std::min(chunk, nt - r0);
        if (!ix->bf16 && ix->d_pad == ix->d) {
            e = cudaMemcpyAsync(host, ix->rows + (size_t)r0 * ix->row_bytes, (size_t)m * d * 4, cudaMemcpyDeviceToHost,
                                ix->stream);
        } else {
            for (int64_t r = 0; r < m && e == cudaSuccess; r++)
                e = launch_row_to_f32(ix->rows + (size_t)(r0 + r) * ix->row_bytes, d, ix->bf16, dev + r * d, ix->stream);
            if (e == cudaSuccess)
                e = cudaMemcpyAsync(host, dev, (size_t)m * d * 4, cudaMemcpyDeviceToHost, ix->stream);
        }
        if (e == cudaSuccess) e = cudaStreamSynchronize(ix->stream);
        if (e == cudaSuccess) ok = fwrite(host, (size_t)d * 4, (size_t)m, f) == (size_t)m;
    }
    cudaFree(dev);
    cudaFreeHost(host);
    ok = (fclose(f) == 0) && ok;
    if (e != cudaSuccess) return fail(B200R_E_CUDA, "b200r_save: %s", cudaGetErrorString(e));
    if (!ok) return fail(B200R_E_IO, "b200r_save: write to %s failed", path);
    return B200R_OK;
}

extern "C" int b200r_load(const char* path, const b200r_config* cfg, b200r_index** out) {
    if (!path || !cfg || !out) return fail(B200R_E_INVALID, "b200r_load: null argument");
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    if (!f) return fail(B200R_E_IO, "b200r_load: cannot open %s", path);
    uint32_t h = 0;
    int32_t d = 0, metric = 0;
    int64_t nt = 0, dummy = 0;
    uint8_t trained = 0;
    uint64_t nfloats = 0;
    bool ok = fread(&h, 4, 1, f) == 1 && fread(&d, 4, 1, f) == 1 && fread(&nt, 8, 1, f) == 1 &&
              fread(&dummy, 8, 1, f) == 1 && fread(&dummy, 8, 1, f) == 1 && fread(&trained, 1, 1, f) == 1 &&
              fread(&metric, 4, 1, f) == 1;
    if (ok && metric > 1) {
        float metric_arg;
        ok = fread(&metric_arg, 4, 1, f) == 1;
    }
    ok = ok && fread(&nfloats, 8, 1, f) == 1;
    if (!ok || (h != fourcc("IxF2") && h != fourcc("IxFI") && h != fourcc("IxFl"))) {
        fclose(f);
        return fail(B200R_E_IO, "b200r_load: %s is not a faiss flat index file", path);
    }
    if (metric != B200R_METRIC_INNER_PRODUCT && metric != B200R_METRIC_L2) {
        fclose(f);
        return fail(B200R_E_UNSUPPORTED, "b200r_load: metric type %d is not supported", metric);
    }
    if (d <= 0 || nt < 0 || nfloats != (uint64_t)nt * (uint64_t)d) {
        fclose(f);
        return fail(B200R_E_IO, "b200r_load: inconsistent header in %s", path);
    }
    b200r_config c = *cfg;
    c.dim = d;
    c.metric = metric;
    c.reserve_rows = std::max<int64_t>(c.reserve_rows, nt);
    const int norm_on_add = c.normalize_on_add;
    c.normalize_on_add = 0;  // stored rows are loaded verbatim
    b200r_index* ix = nullptr;
    int rc = b200r_create(&c, &ix);
    if (rc != B200R_OK) {
        fclose(f);
        return rc;
    }
    const int64_t chunk = std::max<int64_t>(1, (int64_t)((32u << 20) / ((size_t)d * 4)));
    std::vector<float> buf((size_t)std::min<int64_t>(chunk, std::max<int64_t>(nt, 1)) * d);
    for (int64_t r0 = 0; r0 < nt; r0 += chunk) {
        const int64_t m = std::min(chunk, nt - r0);
        if (fread(buf.data(), (size_t)d * 4, (size_t)m, f) != (size_t)m) {
            fclose(f);
            free_index(ix);
            return fail(B200R_E_IO, "b200r_load: %s is truncated", path);
        }
        rc = b200r_add(ix, buf.data(), m, B200R_F32, B200R_HOST, nullptr);
        if (rc != B200R_OK) {
            fclose(f);
            free_index(ix);
            return rc;
        }
    }
    fclose(f);
    ix->cfg.normalize_on_add = norm_on_add;
    ix->st.n_rows_added = 0;
    *out = ix;
    return B200R_OK;
}
