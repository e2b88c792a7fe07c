// The extern "C" entry points of libb200r.so (include/b200r.h names the reference call each one
// replaces).  A handle is either a single-GPU index (api.cu) or a multi-GPU index built from
// single-GPU shards (multi.cu); every entry point checks its arguments and hands over to one of them.
#include "index.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <vector>

using namespace b200r;

extern "C" int b200r_version(void) { return B200R_VERSION; }
extern "C" const char* b200r_last_error(void) { return last_error_cstr(); }
extern "C" int b200r_reload_env(void) {
    reload_knobs();
    return B200R_OK;
}

static bool is_multi_cfg(const b200r_config* cfg) { return cfg->n_devices > 1; }

extern "C" int b200r_create(const b200r_config* cfg, b200r_index** out) {
    if (!cfg || !out) return fail(B200R_E_INVALID, "b200r_create: null argument");
    *out = nullptr;
    if (cfg->dim <= 0) return fail(B200R_E_INVALID, "dim must be positive, got %d", cfg->dim);
    if (cfg->metric != B200R_METRIC_INNER_PRODUCT && cfg->metric != B200R_METRIC_L2)
        return fail(B200R_E_INVALID, "metric must be B200R_METRIC_INNER_PRODUCT or B200R_METRIC_L2");
    if (cfg->storage != B200R_F32 && cfg->storage != B200R_BF16)
        return fail(B200R_E_INVALID, "storage must be B200R_F32 or B200R_BF16");
    if (cfg->n_devices < 0 || cfg->n_devices > B200R_MAX_DEVICES)
        return fail(B200R_E_INVALID, "n_devices must be 0..%d, got %d", B200R_MAX_DEVICES, cfg->n_devices);
    if (is_multi_cfg(cfg)) return multi_create(cfg, out);
    b200r_config c = *cfg;
    if (cfg->n_devices == 1) c.device = cfg->device_ids[0];
    return leaf_create(&c, out);
}

extern "C" int b200r_destroy(b200r_index* ix) {
    if (!ix) return B200R_OK;
    if (ix->multi)
        multi_free(ix);
    else
        leaf_free(ix);
    return B200R_OK;
}

extern "C" int b200r_ntotal(const b200r_index* ix, int64_t* out) {
    if (!ix || !out) return fail(B200R_E_INVALID, "b200r_ntotal: null argument");
    *out = ix->multi ? multi_ntotal(ix) : ix->ntotal;
    return B200R_OK;
}

extern "C" int b200r_set_id_base(b200r_index* ix, int64_t id_base) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_set_id_base: null index");
    std::lock_guard<std::mutex> lk(ix->mu);
    ix->cfg.id_base = id_base;  // a multi-GPU index adds it in its final select
    return B200R_OK;
}

extern "C" int b200r_set_shadow_filter(b200r_index* ix, int32_t enabled) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_set_shadow_filter: null index");
    if (!ix->cfg.shadow_bf16) return fail(B200R_E_INVALID, "index was created without shadow_bf16");
    if (ix->multi) return multi_set_shadow_filter(ix, enabled);
    std::lock_guard<std::mutex> lk(ix->mu);
    ix->shadow_on = enabled != 0;
    return B200R_OK;
}

extern "C" int b200r_set_timing(b200r_index* ix, int32_t enabled) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_set_timing: null index");
    if (ix->multi) return multi_set_timing(ix, enabled);
    ix->timing = enabled != 0;
    return B200R_OK;
}

extern "C" int b200r_stats_get(b200r_index* ix, b200r_stats* out) {
    if (!ix || !out) return fail(B200R_E_INVALID, "b200r_stats_get: null argument");
    return ix->multi ? multi_stats(ix, out) : leaf_stats(ix, out);
}

extern "C" int b200r_reset(b200r_index* ix) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_reset: null index");
    return ix->multi ? multi_reset(ix) : leaf_reset(ix);
}

extern "C" int b200r_add(b200r_index* ix, const void* x, int64_t n, int dtype, int mem_kind, void* stream) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_add: null index");
    return ix->multi ? multi_add(ix, x, n, dtype, mem_kind, stream, true) : leaf_add(ix, x, n, dtype, mem_kind, stream, true);
}

extern "C" int b200r_search(b200r_index* ix, const void* q, int64_t nq, int k, int q_mem_kind, float* D, int64_t* I,
                            int out_mem_kind, void* stream) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_search: null index");
    if (ix->multi) return multi_search(ix, q, nq, k, q_mem_kind, D, I, out_mem_kind, stream);
    return leaf_search(ix, q, nq, k, q_mem_kind, D, I, out_mem_kind, stream, B200R_PATH_AUTO, nullptr);
}

extern "C" int b200r_auto_route(b200r_index* ix, int64_t nq, int64_t n_rows, int32_t* route) {
    if (!ix || !route) return fail(B200R_E_INVALID, "b200r_auto_route: null argument");
    if (ix->multi) return fail(B200R_E_UNSUPPORTED, "b200r_auto_route: ask a single-GPU index");
    *route = leaf_route(ix, nq, n_rows);
    return B200R_OK;
}

// ---------------------------------------------------------------------------
// fused search + NVLink exchange between processes (one rank per GPU, small batches)
// ---------------------------------------------------------------------------
extern "C" int64_t b200r_exchange_buffer_bytes(void) { return kXchgBytes; }

extern "C" int b200r_search_exchange(b200r_index* ix, const float* q, int64_t nq, int k, int q_mem_kind,
                                     void* const* peer_bufs, int32_t world, int32_t rank, uint32_t epoch, float* D,
                                     int64_t* I, int out_mem_kind, void* stream) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_search_exchange: null index");
    if (ix->multi) return fail(B200R_E_UNSUPPORTED, "b200r_search_exchange: the index already spans several GPUs");
    if (!q || !D || !I || !peer_bufs) return fail(B200R_E_INVALID, "b200r_search_exchange: null buffer");
    if ((q_mem_kind != B200R_HOST && q_mem_kind != B200R_DEVICE) ||
        (out_mem_kind != B200R_HOST && out_mem_kind != B200R_DEVICE))
        return fail(B200R_E_INVALID, "b200r_search_exchange: bad mem_kind");
    if (nq < 1 || nq > kScanMaxNQ)
        return fail(B200R_E_UNSUPPORTED, "b200r_search_exchange: nq %lld outside 1..%d", (long long)nq, kScanMaxNQ);
    if (k < 1 || k > kScanMaxK) return fail(B200R_E_UNSUPPORTED, "b200r_search_exchange: k %d outside 1..%d", k, kScanMaxK);
    if (world < 1 || world > kXchgMaxWorld || rank < 0 || rank >= world)
        return fail(B200R_E_INVALID, "b200r_search_exchange: bad world/rank %d/%d", world, rank);
    if (epoch == 0) return fail(B200R_E_INVALID, "b200r_search_exchange: epoch must start at 1");
    ScanExchange x{};
    x.n_targets = world;
    x.slot = rank;
    x.epoch = epoch;
    for (int r = 0; r < world; r++) {
        if (!peer_bufs[r]) return fail(B200R_E_INVALID, "b200r_search_exchange: null peer buffer %d", r);
        x.target[r] = reinterpret_cast<uint8_t*>(peer_bufs[r]);
    }
    // as in b200r_search: with host buffers only, NULL selects the index's own stream
    const bool any_dev = q_mem_kind == B200R_DEVICE || out_mem_kind == B200R_DEVICE;
    cudaStream_t st = (any_dev || stream) ? (cudaStream_t)stream : ix->stream;
    int rc = leaf_search(ix, q, nq, k, q_mem_kind, nullptr, nullptr, B200R_DEVICE, st, B200R_PATH_SCAN, &x);
    if (rc != B200R_OK) return rc;
    DeviceGuard g(ix->device);
    std::lock_guard<std::mutex> lk(ix->mu);
    // host results: the select kernel writes [scores | ids | status] into the index's staging buffer,
    // ONE copy brings all of it back
    const size_t ids_off = ((size_t)nq * k * sizeof(float) + 15) / 16 * 16;
    const size_t status_off = (ids_off + (size_t)nq * k * sizeof(int64_t) + 15) / 16 * 16;
    const bool host_out = out_mem_kind == B200R_HOST;
    // (nq <= 4, k <= 128: the block is at most 6 KB, so the select kernel writes it straight into the pinned
    // host buffer -- mapped memory -- unless B200R_ZERO_COPY_OUT=0 asks for a device-to-host copy)
    const bool zc = host_out && knobs().zero_copy_out != 0;
    uint8_t* stage = zc ? ix->out_pin : ix->out_buf;
    float* od = host_out ? reinterpret_cast<float*>(stage) : D;
    int64_t* oi = host_out ? reinterpret_cast<int64_t*>(stage + ids_off) : I;
    int* status_dev = host_out ? reinterpret_cast<int*>(stage + status_off) : nullptr;
    CU(launch_merge_wait(ix->l2, reinterpret_cast<uint8_t*>(peer_bufs[rank]), world, (int)nq, k, epoch, od, oi, status_dev,
                         st));
    ix->st.n_kernel_launches++;
    if (host_out) {
        if (!zc) CU(cudaMemcpyAsync(ix->out_pin, ix->out_buf, status_off + sizeof(int), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        ix->ws_ev_set = false;
        const int status = *reinterpret_cast<const int*>(ix->out_pin + status_off);
        if (status != 0) {
            CU(cudaMemsetAsync(reinterpret_cast<uint8_t*>(peer_bufs[rank]) + kXchgStatusOffset, 0, sizeof(int), st));
            CU(cudaStreamSynchronize(st));
            return fail(B200R_E_CUDA, "b200r_search_exchange: rank %d did not deliver its candidates within 2 s", status - 1);
        }
        memcpy(D, ix->out_pin, (size_t)nq * k * sizeof(float));
        memcpy(I, ix->out_pin + ids_off, (size_t)nq * k * sizeof(int64_t));
        return B200R_OK;
    }
    if (q_mem_kind == B200R_HOST) CU(cudaStreamSynchronize(st));  // the pinned query buffer is free again
    CU(cudaEventRecord(ix->ws_ev, st));
    ix->ws_ev_set = true;
    ix->ws_ev_stream = st;
    return B200R_OK;
}

extern "C" int b200r_exchange_status(void* my_buf, int32_t device, void* stream, int32_t* status) {
    if (!my_buf || !status) return fail(B200R_E_INVALID, "b200r_exchange_status: null argument");
    DeviceGuard g(device);
    if (!g.ok) return fail(B200R_E_CUDA, "cudaSetDevice(%d) failed", device);
    int32_t word = 0;
    CU(cudaMemcpyAsync(&word, reinterpret_cast<uint8_t*>(my_buf) + kXchgStatusOffset, sizeof word, cudaMemcpyDeviceToHost,
                       (cudaStream_t)stream));
    CU(cudaStreamSynchronize((cudaStream_t)stream));
    if (word != 0) {  // clear it so that the next search starts clean
        CU(cudaMemsetAsync(reinterpret_cast<uint8_t*>(my_buf) + kXchgStatusOffset, 0, sizeof word, (cudaStream_t)stream));
        CU(cudaStreamSynchronize((cudaStream_t)stream));
    }
    *status = word;
    return B200R_OK;
}

// ---------------------------------------------------------------------------
// remove_ids / reconstruct / merge_from
// ---------------------------------------------------------------------------
extern "C" int b200r_remove_ids(b200r_index* ix, const int64_t* ids, int64_t n, int64_t* n_removed) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_remove_ids: null index");
    return ix->multi ? multi_remove_ids(ix, ids, n, n_removed) : leaf_remove_ids(ix, ids, n, n_removed);
}

extern "C" int b200r_reconstruct(b200r_index* ix, int64_t key, float* out) {
    if (!ix || !out) return fail(B200R_E_INVALID, "b200r_reconstruct: null argument");
    if (ix->multi) {
        if (key < 0 || key >= multi_ntotal(ix))
            return fail(B200R_E_INVALID, "b200r_reconstruct: key %lld out of range [0, %lld)", (long long)key,
                        (long long)multi_ntotal(ix));
        return multi_reconstruct_batch(ix, &key, 1, out);
    }
    return leaf_reconstruct(ix, key, out);
}

extern "C" int b200r_reconstruct_batch(b200r_index* ix, const int64_t* keys, int64_t n, float* out) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_reconstruct_batch: null index");
    return ix->multi ? multi_reconstruct_batch(ix, keys, n, out) : leaf_reconstruct_batch(ix, keys, n, out);
}

extern "C" int b200r_merge_from(b200r_index* dst, b200r_index* src) {
    if (!dst || !src) return fail(B200R_E_INVALID, "b200r_merge_from: null index");
    if (dst == src) return fail(B200R_E_INVALID, "b200r_merge_from: an index cannot be merged into itself");
    if (dst->multi || src->multi) return multi_merge_from(dst, src);
    return leaf_merge_from(dst, src);
}

// ---------------------------------------------------------------------------
// relevance threshold and MMR on the device (rerank.cu)
// ---------------------------------------------------------------------------
namespace b200r {
void free_rerank_ws(b200r_index* ix) {
    if (ix->rr_stream) cudaStreamSynchronize(ix->rr_stream);
    cudaFree(ix->rr_dev);
    cudaFreeHost(ix->rr_pin);
    if (ix->rr_stream) cudaStreamDestroy(ix->rr_stream);
    ix->rr_dev = ix->rr_pin = nullptr;
    ix->rr_stream = nullptr;
    ix->rr_cap = 0;
}
}  // namespace b200r

static int ensure_rerank_ws(b200r_index* ix, size_t bytes) {
    if (!ix->rr_stream) CU(cudaStreamCreateWithFlags(&ix->rr_stream, cudaStreamNonBlocking));
    if (bytes <= ix->rr_cap) return B200R_OK;
    CU(cudaStreamSynchronize(ix->rr_stream));
    cudaFree(ix->rr_dev);
    cudaFreeHost(ix->rr_pin);
    ix->rr_dev = ix->rr_pin = nullptr;
    ix->rr_cap = 0;
    const size_t cap = std::max<size_t>(bytes, 1u << 20);
    CU(cudaMalloc(&ix->rr_dev, cap));
    CU(cudaMallocHost(&ix->rr_pin, cap));
    ix->rr_cap = cap;
    return B200R_OK;
}

static size_t align16(size_t v) { return (v + 15) / 16 * 16; }

extern "C" int b200r_search_threshold(b200r_index* ix, const float* q, int64_t nq, int k, int32_t relevance_kind,
                                      double threshold, int32_t apply_threshold, double* R, float* D, int64_t* I,
                                      int32_t* counts, int32_t* out_of_range) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_search_threshold: null index");
    if (nq < 0) return fail(B200R_E_INVALID, "b200r_search_threshold: nq = %lld", (long long)nq);
    if (k <= 0) return fail(B200R_E_INVALID, "b200r_search_threshold: k must be > 0, got %d", k);
    if (nq == 0) return B200R_OK;
    if (!q || !R || !D || !I || !counts) return fail(B200R_E_INVALID, "b200r_search_threshold: null buffer");
    if (relevance_kind < 0 || relevance_kind > 2)
        return fail(B200R_E_INVALID, "b200r_search_threshold: relevance_kind must be 0 (euclidean), 1 (cosine) or 2 "
                                     "(max inner product)");
    std::lock_guard<std::mutex> lk(ix->rr_mu);
    DeviceGuard g(ix->device);
    const int d = ix->cfg.dim;
    const size_t nk = (size_t)nq * k;
    // device: [queries | raw D | raw I] then the block that travels back: [R | D | I | counts | out-of-range]
    const size_t o_q = 0, o_rd = align16(o_q + (size_t)nq * d * 4), o_ri = align16(o_rd + nk * 4);
    const size_t o_back = align16(o_ri + nk * 8);
    const size_t b_r = 0, b_d = align16(b_r + nk * 8), b_i = align16(b_d + nk * 4), b_c = align16(b_i + nk * 8),
                 b_o = align16(b_c + (size_t)nq * 4), back_bytes = align16(b_o + (size_t)nq * 4);
    int rc = ensure_rerank_ws(ix, o_back + back_bytes);
    if (rc != B200R_OK) return rc;
    cudaStream_t st = ix->rr_stream;
    uint8_t* dv = ix->rr_dev;
    memcpy(ix->rr_pin, q, (size_t)nq * d * 4);
    CU(cudaMemcpyAsync(dv + o_q, ix->rr_pin, (size_t)nq * d * 4, cudaMemcpyHostToDevice, st));
    float* raw_d = reinterpret_cast<float*>(dv + o_rd);
    int64_t* raw_i = reinterpret_cast<int64_t*>(dv + o_ri);
    rc = ix->multi ? multi_search(ix, dv + o_q, nq, k, B200R_DEVICE, raw_d, raw_i, B200R_DEVICE, st)
                   : leaf_search(ix, dv + o_q, nq, k, B200R_DEVICE, raw_d, raw_i, B200R_DEVICE, st, B200R_PATH_AUTO, nullptr);
    if (rc != B200R_OK) return rc;
    // small results: the filter kernel writes straight into the pinned host block (mapped memory); larger
    // ones go through device memory and one copy
    const bool zc = back_bytes <= ((size_t)32 << 10) && knobs().zero_copy_out != 0;
    uint8_t* back = zc ? ix->rr_pin + o_back : dv + o_back;
    CU(launch_relevance_filter(raw_d, raw_i, (int)nq, k, relevance_kind, threshold, apply_threshold != 0,
                               reinterpret_cast<double*>(back + b_r), reinterpret_cast<float*>(back + b_d),
                               reinterpret_cast<int64_t*>(back + b_i), reinterpret_cast<int32_t*>(back + b_c),
                               reinterpret_cast<int32_t*>(back + b_o), st));
    if (!zc) CU(cudaMemcpyAsync(ix->rr_pin + o_back, back, back_bytes, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const uint8_t* hb = ix->rr_pin + o_back;
    const int32_t* hc = reinterpret_cast<const int32_t*>(hb + b_c);
    memcpy(counts, hc, (size_t)nq * 4);
    if (out_of_range) memcpy(out_of_range, hb + b_o, (size_t)nq * 4);
    // only the passing hits are meaningful: rows are copied whole (k is small), the tail is left as it is
    memcpy(R, hb + b_r, nk * 8);
    memcpy(D, hb + b_d, nk * 4);
    memcpy(I, hb + b_i, nk * 8);
    return B200R_OK;
}

extern "C" int b200r_mmr(b200r_index* ix, const float* query, const int64_t* keys, int64_t m, int k, double lambda_mult,
                         int32_t* selected, int32_t* n_selected) {
    if (!ix) return fail(B200R_E_INVALID, "b200r_mmr: null index");
    if (m < 0 || k <= 0) return fail(B200R_E_INVALID, "b200r_mmr: bad sizes");
    if (!query || !selected || !n_selected || (m > 0 && !keys)) return fail(B200R_E_INVALID, "b200r_mmr: null buffer");
    if (m > kMmrMaxCandidates)
        return fail(B200R_E_UNSUPPORTED, "b200r_mmr: %lld candidates, at most %d are supported", (long long)m, kMmrMaxCandidates);
    *n_selected = 0;
    if (m == 0) return B200R_OK;
    std::lock_guard<std::mutex> lk(ix->rr_mu);
    DeviceGuard g(ix->device);
    const int d = ix->cfg.dim;
    const size_t o_q = 0, o_k = align16((size_t)d * 4), o_e = align16(o_k + (size_t)m * 8),
                 o_s = align16(o_e + (size_t)m * d * 4), total = align16(o_s + ((size_t)k + 1) * 4);
    int rc = ensure_rerank_ws(ix, total);
    if (rc != B200R_OK) return rc;
    cudaStream_t st = ix->rr_stream;
    uint8_t* dv = ix->rr_dev;
    float* emb = reinterpret_cast<float*>(dv + o_e);
    if (ix->multi) {
        // the candidates live on several GPUs: fetch them through the segment map, then select on the primary GPU
        rc = multi_reconstruct_batch(ix, keys, m, reinterpret_cast<float*>(ix->rr_pin + o_e));
        if (rc != B200R_OK) return rc;
        memcpy(ix->rr_pin + o_q, query, (size_t)d * 4);
        CU(cudaMemcpyAsync(dv + o_q, ix->rr_pin + o_q, (size_t)d * 4, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(emb, ix->rr_pin + o_e, (size_t)m * d * 4, cudaMemcpyHostToDevice, st));
    } else {
        std::lock_guard<std::mutex> li(ix->mu);
        for (int64_t i = 0; i < m; i++)
            if (keys[i] < 0 || keys[i] >= ix->ntotal)
                return fail(B200R_E_INVALID, "b200r_mmr: key %lld out of range [0, %lld)", (long long)keys[i],
                            (long long)ix->ntotal);
        memcpy(ix->rr_pin + o_q, query, (size_t)d * 4);
        memcpy(ix->rr_pin + o_k, keys, (size_t)m * 8);
        CU(cudaMemcpyAsync(dv + o_q, ix->rr_pin + o_q, o_k + (size_t)m * 8, cudaMemcpyHostToDevice, st));
        if (ix->data_ev_set) CU(cudaStreamWaitEvent(st, ix->data_ev, 0));
        // the candidates' rows never leave the GPU: gathered as fp32 next to the query
        CU(launch_rows_to_f32(ix->rows, ix->row_bytes, d, ix->bf16, reinterpret_cast<const int64_t*>(dv + o_k), m, emb, st));
        ix->st.n_kernel_launches++;
    }
    int32_t* sel_dev = reinterpret_cast<int32_t*>(dv + o_s);
    CU(launch_mmr(emb, reinterpret_cast<const float*>(dv + o_q), (int)m, d, k, lambda_mult, sel_dev + 1, sel_dev, st));
    CU(cudaMemcpyAsync(ix->rr_pin + o_s, sel_dev, ((size_t)std::min<int64_t>(k, m) + 1) * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const int32_t* hs = reinterpret_cast<const int32_t*>(ix->rr_pin + o_s);
    *n_selected = hs[0];
    memcpy(selected, hs + 1, (size_t)hs[0] * 4);
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
    FILE* f = fopen(path, "wb");
    if (!f) return fail(B200R_E_IO, "b200r_save: cannot open %s", path);
    const bool l2 = ix->cfg.metric == B200R_METRIC_L2;
    const uint32_t h = fourcc(l2 ? "IxF2" : "IxFI");
    const int32_t d = ix->cfg.dim;
    const int64_t nt = ix->multi ? multi_ntotal(ix) : ix->ntotal, dummy = 1 << 20;
    const uint8_t trained = 1;
    const int32_t metric = ix->cfg.metric;
    const uint64_t nfloats = (uint64_t)nt * (uint64_t)d;
    bool ok = fwrite(&h, 4, 1, f) == 1 && fwrite(&d, 4, 1, f) == 1 && fwrite(&nt, 8, 1, f) == 1 &&
              fwrite(&dummy, 8, 1, f) == 1 && fwrite(&dummy, 8, 1, f) == 1 && fwrite(&trained, 1, 1, f) == 1 &&
              fwrite(&metric, 4, 1, f) == 1 && fwrite(&nfloats, 8, 1, f) == 1;
    // rows go out as fp32 without the padding columns, in bounded chunks (one conversion launch per chunk)
    const int64_t chunk = std::max<int64_t>(1, (int64_t)((64u << 20) / ((size_t)d * 4)));
    std::vector<float> host((size_t)std::min<int64_t>(chunk, std::max<int64_t>(nt, 1)) * d);
    int rc = B200R_OK;
    for (int64_t r0 = 0; ok && rc == B200R_OK && r0 < nt; r0 += chunk) {
        const int64_t m = std::min(chunk, nt - r0);
        rc = ix->multi ? multi_read_rows(ix, r0, m, host.data()) : leaf_read_rows(ix, r0, m, host.data());
        if (rc == B200R_OK) ok = fwrite(host.data(), (size_t)d * 4, (size_t)m, f) == (size_t)m;
    }
    ok = (fclose(f) == 0) && ok;
    if (rc != B200R_OK) return rc;
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
            b200r_destroy(ix);
            return fail(B200R_E_IO, "b200r_load: %s is truncated", path);
        }
        // stored rows are loaded verbatim (no normalisation)
        rc = ix->multi ? multi_add(ix, buf.data(), m, B200R_F32, B200R_HOST, nullptr, false)
                       : leaf_add(ix, buf.data(), m, B200R_F32, B200R_HOST, nullptr, false);
        if (rc != B200R_OK) {
            fclose(f);
            b200r_destroy(ix);
            return rc;
        }
    }
    fclose(f);
    if (!ix->multi) ix->st.n_rows_added = 0;
    *out = ix;
    return B200R_OK;
}
