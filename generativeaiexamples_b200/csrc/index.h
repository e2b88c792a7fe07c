// Host-side types shared by the .cu files that implement the C ABI: the single-GPU index
// ("leaf", api.cu), the multi-GPU index built from leaves (multi.cu) and the exported entry
// points that pick between them (abi.cu).  Nothing here is visible through include/b200r.h.
#pragma once
#include "../../include/b200r.h"
#include "common.cuh"
#include "kernels.h"

#include <mutex>
#include <string>

namespace b200r {

// ---------------------------------------------------------------------------
// errors (thread-local message behind b200r_last_error)
// ---------------------------------------------------------------------------
int fail(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
const char* last_error_cstr();

#define CU(call)                                                                                     \
    do {                                                                                             \
        cudaError_t e__ = (call);                                                                    \
        if (e__ != cudaSuccess)                                                                      \
            return ::b200r::fail(e__ == cudaErrorMemoryAllocation ? B200R_E_NOMEM : B200R_E_CUDA,    \
                                 "%s: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
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
// environment switches: read once per process (first use) and again on b200r_reload_env()
// ---------------------------------------------------------------------------
struct Knobs {
    int stage_threads = 8;          // B200R_STAGE_THREADS
    long long gemm_min_nq = -1;     // B200R_GEMM_MIN_NQ      (-1 = unset)
    long long gemm_min_bytes = -1;  // B200R_GEMM_MIN_BYTES   (-1 = unset)
    int evict_first = -1;           // B200R_EVICT_FIRST      (-1 = unset)
    bool gemm_noboost = false;      // B200R_GEMM_NOBOOST
    bool debug_fails = false;       // B200R_DEBUG_FAILS
    int scan_stage_kb = 0;          // B200R_SCAN_STAGE_KB    (0 = default)
    int scan_stages = 0;            // B200R_SCAN_STAGES      (0 = default)
    int scan_dynamic = -1;          // B200R_SCAN_DYNAMIC     (-1 = default = 2: static head, dynamic tail)
    int scan_trace = 0;             // B200R_SCAN_TRACE       1: per-CTA timestamps of every scan launch to stderr
    int scan_tile_block = 1;        // B200R_SCAN_TILE_BLOCK  (consecutive tiles per CTA and round)
    int zero_copy_out = 1;          // B200R_ZERO_COPY_OUT    0: small host results also go through a device-to-host copy
    int gemm_kpcap = 0;             // B200R_GEMM_KPCAP       (0 = unset)
    int gemm_kprime = 0;            // B200R_GEMM_KPRIME      (0 = unset)
    int gemm_cluster = 0;           // B200R_GEMM_CLUSTER     (0 = unset)
    int gemm_pair = -1;             // B200R_GEMM_PAIR        (-1 = unset)
    int gemm_prefetch = 0;          // B200R_GEMM_PREFETCH
    int gemm_noselect = 0;          // B200R_GEMM_DEBUG_NOSELECT
    int gemm_warmup = -1;           // B200R_GEMM_WARMUP      (-1 = unset)
    int multi_workers = -1;         // B200R_MULTI_WORKERS    (-1 = default: on for >= 2 shards)
    int multi_spin_us = 2000;       // B200R_MULTI_SPIN_US    how long an idle shard worker polls before it sleeps
    int multi_fused = -1;           // B200R_MULTI_FUSED      (-1 = default)
};
const Knobs& knobs();
inline int scan_dynamic_mode() { return knobs().scan_dynamic < 0 ? 2 : knobs().scan_dynamic; }
void reload_knobs();

// ---------------------------------------------------------------------------
// the single-GPU index
// ---------------------------------------------------------------------------
constexpr size_t kStageHalfBytes = 32u << 20;  // pinned / device staging, per half
constexpr int64_t kMaxRows = 0xFFFFFFFEll;     // ids are 32-bit inside the kernels

struct MultiIndex;

}  // namespace b200r

struct b200r_index {
    b200r_config cfg{};
    b200r::MultiIndex* multi = nullptr;  // non-null: this handle is a multi-GPU index and nothing below is used
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
    b200r::Cand* scratch = nullptr;
    unsigned* counter = nullptr;
    uint8_t* out_buf = nullptr;  // host-output searches: [scores | pad to 16 B | ids], one D2H per search
    float* out_d = nullptr;      // = out_buf
    int64_t* out_i = nullptr;    // ids of a full-capacity result (only a placeholder target otherwise)
    uint8_t* out_pin = nullptr;
    int64_t out_cap = 0;  // nq*k entries allocated
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    bool timing = false;
    bool timing_deferred = false;  // shard of a multi-GPU index: the owner reads t0/t1 after its own synchronisation
    // GEMM path workspace
    void* q_bf16 = nullptr;        // bf16 copy of q_dev (bf16 corpus only), q_cap rows
    b200r::Cand* gemm_cand = nullptr;
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
    // small scratch kept with the index so that per-call entry points allocate nothing
    float* recon_dev = nullptr;    // [d] one reconstructed row
    float* recon_pin = nullptr;
    uint8_t* bounce = nullptr;     // remove_ids compaction buffer
    size_t bounce_bytes = 0;
    int64_t* adj_dev = nullptr;    // remove_ids: adjusted deleted positions
    int64_t adj_cap = 0;
    unsigned long long* trace_dev = nullptr;  // B200R_SCAN_TRACE
    // threshold / MMR entry points (abi.cu): staging on this index's (primary) GPU
    std::mutex rr_mu;
    uint8_t* rr_dev = nullptr;
    uint8_t* rr_pin = nullptr;
    size_t rr_cap = 0;
    cudaStream_t rr_stream = nullptr;

    std::mutex mu;
    b200r_stats st{};
};

namespace b200r {

// Where the scan kernel's final writer puts the result instead of out_d / out_i (multi-GPU
// search: one slot per shard in a gather buffer that lives on the primary GPU; plain stores,
// over NVLink when the buffer is a peer's).
struct ScanExchange {
    int n_targets = 0;        // 0 = off
    int slot = 0;             // this shard's slot
    uint32_t epoch = 0;
    uint8_t* target[8] = {};
};

// leaf operations (api.cu); the extern "C" entry points in abi.cu call these, or their
// multi-GPU counterparts (multi.cu) when the handle carries a MultiIndex
int leaf_create(const b200r_config* cfg, b200r_index** out);
void leaf_free(b200r_index* ix);
// allow_normalize = false: store the rows as given whatever cfg.normalize_on_add says (merge / load)
int leaf_add(b200r_index* ix, const void* x, int64_t n, int dtype, int mem_kind, void* stream, bool allow_normalize);
// route: B200R_PATH_AUTO = the index's own dispatch; xchg (may be null) only with the scan route
int leaf_search(b200r_index* ix, const void* q, int64_t nq, int k, int q_mem_kind, float* D, int64_t* I,
                int out_mem_kind, void* stream, int route, const ScanExchange* xchg);
// the route AUTO would take for nq queries right now (B200R_PATH_SCAN or B200R_PATH_GEMM)
int leaf_route(b200r_index* ix, int64_t nq, int64_t rows = -1);  // rows < 0: the index's own row count
int leaf_remove_ids(b200r_index* ix, const int64_t* ids, int64_t n, int64_t* n_removed);
int leaf_reconstruct(b200r_index* ix, int64_t key, float* out);
int leaf_reconstruct_batch(b200r_index* ix, const int64_t* keys, int64_t n, float* out);
// rows [r0, r0+n) as fp32 without padding into a HOST buffer of n*d floats
int leaf_read_rows(b200r_index* ix, int64_t r0, int64_t n, float* out);
// rows [r0, r0+n) as fp32 [n][d] into a DEVICE buffer on the index's GPU, on its own stream (synchronised)
int leaf_rows_to_device_f32(b200r_index* ix, int64_t r0, int64_t n, float* dev_out);
int leaf_reset(b200r_index* ix);
int leaf_merge_from(b200r_index* dst, b200r_index* src);
int leaf_stats(b200r_index* ix, b200r_stats* out);
int leaf_last_search_ms(b200r_index* ix, double* ms_out);

// re-ranking kernels (rerank.cu)
constexpr int kMmrMaxCandidates = 2048;
cudaError_t launch_relevance_filter(const float* d_in, const int64_t* i_in, int nq, int k, int kind, double threshold,
                                    bool apply_threshold, double* r_out, float* d_out, int64_t* i_out, int32_t* counts,
                                    int32_t* out_of_range, cudaStream_t st);
cudaError_t launch_mmr(const float* emb, const float* q, int m, int d, int k, double lambda_mult, int32_t* selected,
                       int32_t* n_selected, cudaStream_t st);
void free_rerank_ws(b200r_index* ix);

// multi-GPU index (multi.cu)
int multi_create(const b200r_config* cfg, b200r_index** out);
void multi_free(b200r_index* ix);
int multi_add(b200r_index* ix, const void* x, int64_t n, int dtype, int mem_kind, void* stream, bool allow_normalize);
int multi_search(b200r_index* ix, const void* q, int64_t nq, int k, int q_mem_kind, float* D, int64_t* I,
                 int out_mem_kind, void* stream);
int multi_remove_ids(b200r_index* ix, const int64_t* ids, int64_t n, int64_t* n_removed);
int multi_reconstruct_batch(b200r_index* ix, const int64_t* keys, int64_t n, float* out);
int multi_read_rows(b200r_index* ix, int64_t r0, int64_t n, float* out);
int multi_reset(b200r_index* ix);
int multi_merge_from(b200r_index* dst, b200r_index* src);
int multi_stats(b200r_index* ix, b200r_stats* out);
int multi_set_shadow_filter(b200r_index* ix, int enabled);
int multi_set_timing(b200r_index* ix, int enabled);
int64_t multi_ntotal(const b200r_index* ix);

}  // namespace b200r
