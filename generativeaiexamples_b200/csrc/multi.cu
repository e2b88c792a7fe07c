// Multi-GPU index inside ONE process (b200r_config.n_devices >= 2).
//
// Why it exists: the reference's chain server is a single process that builds its vector store once
// (RAG/src/chain_server/server.py:203-238, RAG/examples/basic_rag/langchain/chains.py:40-49), adds
// documents file by file (chains.py:83), deletes them one by one (utils.py:546-560) and queries
// synchronously (chains.py:156-167).  To let that process use every GPU of the box the index itself
// spans them: one single-GPU index ("shard", api.cu) per device, and on top of it
//
//  * a SEGMENT MAP that keeps faiss's "id == position, later rows shift down on delete" contract
//    without ever moving a row between GPUs.  The global row order is a sequence of segments
//    (shard, first local row, length).  add() appends segments (small batches go to the emptiest
//    shard, large ones are spread so that the shards level out), remove_ids() compacts every shard
//    locally (stable, K5) and only shortens segments.  Within a shard the local order always equals
//    the global order, so a shard's "lowest local row first" tie-break is the global one too;
//  * a search that runs the per-shard search on every GPU at once -- a host worker thread per shard
//    issues its copy + kernels so the GPUs start within microseconds of each other -- lets every
//    shard put its top-k straight into its slot of a gather buffer on the primary GPU (plain stores
//    over NVLink) and finishes with one select kernel there that translates local rows to global
//    positions (merge.cu: merge_shards_kernel).  Small batches on the scan route signal "slot
//    filled" from inside the scan kernel (flags in the gather buffer) and the select kernel waits
//    for the flags; every other case uses one CUDA event per shard.
//
// Only plain loads and stores ever touch another GPU's memory here (query vectors read by the scan /
// prep kernels, result slots written by the final writers).  Every TMA tensor map, multicast or
// not, is built by the shard over its own cudaMalloc'ed buffers.
#include "index.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>

namespace b200r {

struct Segment {
    int shard;
    int64_t lstart;  // first local row on the shard
    int64_t len;
};

struct Shard {
    int device = 0;
    b200r_index* leaf = nullptr;
    cudaStream_t stream = nullptr;   // searches of this shard
    cudaEvent_t done_ev = nullptr;   // this shard's slot is filled (event route)
    float* q_dev = nullptr;          // query copy on this GPU
    int64_t q_cap = 0;
};

// host threads that issue per-shard work in parallel
struct Workers {
    std::vector<std::thread> th;
    std::atomic<uint64_t> seq{0};
    std::atomic<int> pending{0};
    std::atomic<int> sleepers{0};
    std::atomic<bool> stop{false};
    std::function<void(int)> job;
    std::mutex mu;
    std::condition_variable cv;
};

struct MultiIndex {
    int G = 0;
    std::vector<Shard> sh;
    std::vector<Segment> segs;       // global order
    int64_t ntotal = 0;
    int primary = 0;                 // device of shard 0: gather buffer, final select, host staging
    int64_t min_piece = 1024;        // smallest run of rows worth its own segment

    cudaStream_t pstream = nullptr;  // primary GPU: final select, result copy
    cudaEvent_t merge_ev = nullptr;  // final select done
    bool prev_async = false;         // the last search left work in flight (device results)
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    bool timing = false;
    cudaEvent_t caller_ev[64] = {};  // per caller device: "inputs ready" marker for device callers

    uint8_t* gather = nullptr;       // [G] slots of (scores | ids), generic route
    size_t gather_bytes = 0;
    uint8_t* xbuf = nullptr;         // fused route: exchange buffer (kernels.h layout), zero-filled once
    uint32_t epoch = 0;
    // segment tables on the primary GPU
    int* seg_off_dev = nullptr;
    int64_t* seg_tab_dev = nullptr;  // lstart[n] then gbase[n]
    int64_t seg_tab_cap = 0;
    // host staging
    float* q_pin = nullptr;
    int64_t q_pin_cap = 0;
    uint8_t* out_dev = nullptr;
    uint8_t* out_pin = nullptr;
    int64_t out_cap = 0;

    Workers* workers = nullptr;
    std::vector<int> rc;             // per-shard status of the last parallel job
    std::vector<std::string> err;

    b200r_stats st{};
};

static int64_t round_up64(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

// ---------------------------------------------------------------------------
// workers
// ---------------------------------------------------------------------------
static void worker_main(MultiIndex* m, int g) {
    cudaSetDevice(m->sh[g].device);
    Workers* w = m->workers;
    uint64_t seen = 0;
    for (;;) {
        const int spin_us = knobs().multi_spin_us;
        auto t_start = std::chrono::steady_clock::now();
        int polls = 0;
        while (w->seq.load(std::memory_order_acquire) == seen && !w->stop.load(std::memory_order_acquire)) {
            if ((++polls & 63) == 0 &&
                std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t_start).count() >
                    spin_us) {
                std::unique_lock<std::mutex> lk(w->mu);
                w->sleepers.fetch_add(1);
                w->cv.wait(lk, [&] { return w->seq.load() != seen || w->stop.load(); });
                w->sleepers.fetch_sub(1);
                break;
            }
#if defined(__x86_64__)
            __builtin_ia32_pause();
#endif
        }
        if (w->stop.load(std::memory_order_acquire)) return;
        seen = w->seq.load(std::memory_order_acquire);
        w->job(g);
        w->pending.fetch_sub(1, std::memory_order_release);
    }
}

// runs fn(g) for every shard: shard 0 on the calling thread, the others on their workers (or inline)
template <typename Fn>
static int run_on_shards(MultiIndex* m, Fn&& fn) {
    for (int g = 0; g < m->G; g++) {
        m->rc[g] = B200R_OK;
        m->err[g].clear();
    }
    auto wrapped = [&](int g) {
        const int rc = fn(g);
        m->rc[g] = rc;
        if (rc != B200R_OK) m->err[g] = last_error_cstr();
    };
    Workers* w = m->workers;
    if (w) {
        w->job = wrapped;
        w->pending.store(m->G - 1, std::memory_order_release);
        w->seq.fetch_add(1);
        if (w->sleepers.load() > 0) {
            std::lock_guard<std::mutex> lk(w->mu);
            w->cv.notify_all();
        }
        {
            DeviceGuard g0(m->sh[0].device);
            wrapped(0);
        }
        while (w->pending.load(std::memory_order_acquire) != 0) {
#if defined(__x86_64__)
            __builtin_ia32_pause();
#endif
        }
    } else {
        for (int g = 0; g < m->G; g++) {
            DeviceGuard dg(m->sh[g].device);
            wrapped(g);
        }
    }
    for (int g = 0; g < m->G; g++)
        if (m->rc[g] != B200R_OK) return fail(m->rc[g], "GPU %d: %s", m->sh[g].device, m->err[g].c_str());
    return B200R_OK;
}

// ---------------------------------------------------------------------------
// segment map
// ---------------------------------------------------------------------------
static void coalesce_segments(MultiIndex* m) {
    std::vector<Segment> out;
    out.reserve(m->segs.size());
    for (const Segment& s : m->segs) {
        if (s.len == 0) continue;
        if (!out.empty() && out.back().shard == s.shard && out.back().lstart + out.back().len == s.lstart)
            out.back().len += s.len;
        else
            out.push_back(s);
    }
    m->segs.swap(out);
}

// per-shard tables (segments by local start, with their global start) -> primary GPU
static int upload_segments(MultiIndex* m) {
    const int n = (int)m->segs.size();
    std::vector<int> off(m->G + 1, 0);
    for (const Segment& s : m->segs) off[s.shard + 1]++;
    for (int g = 0; g < m->G; g++) off[g + 1] += off[g];
    std::vector<int64_t> tab((size_t)2 * std::max(n, 1));
    std::vector<int> fill(off.begin(), off.end() - 1);
    int64_t gpos = 0;
    for (const Segment& s : m->segs) {  // global order == local order within a shard
        const int at = fill[s.shard]++;
        tab[at] = s.lstart;
        tab[(size_t)std::max(n, 1) + at] = gpos;
        gpos += s.len;
    }
    DeviceGuard dg(m->primary);
    if (n > m->seg_tab_cap || !m->seg_tab_dev) {
        cudaFree(m->seg_tab_dev);
        m->seg_tab_dev = nullptr;
        m->seg_tab_cap = 0;
        const int64_t cap = std::max<int64_t>(2 * (int64_t)n, 256);
        CU(cudaMalloc(&m->seg_tab_dev, (size_t)2 * cap * sizeof(int64_t)));
        m->seg_tab_cap = cap;
    }
    // a search still in flight may be reading the old tables
    CU(cudaStreamSynchronize(m->pstream));
    if (!m->sh.empty() && m->sh[0].stream) CU(cudaStreamSynchronize(m->sh[0].stream));
    CU(cudaMemcpyAsync(m->seg_off_dev, off.data(), (size_t)(m->G + 1) * sizeof(int), cudaMemcpyHostToDevice, m->pstream));
    const int64_t nn = std::max(n, 1);
    CU(cudaMemcpyAsync(m->seg_tab_dev, tab.data(), (size_t)nn * sizeof(int64_t), cudaMemcpyHostToDevice, m->pstream));
    CU(cudaMemcpyAsync(m->seg_tab_dev + m->seg_tab_cap, tab.data() + nn, (size_t)nn * sizeof(int64_t),
                       cudaMemcpyHostToDevice, m->pstream));
    CU(cudaStreamSynchronize(m->pstream));
    m->st.n_segments = n;
    return B200R_OK;
}

struct Located {
    int shard;
    int64_t local;
};

// global position -> (shard, local row); prefix[i] = global start of segment i
static std::vector<int64_t> segment_prefix(const MultiIndex* m) {
    std::vector<int64_t> prefix(m->segs.size() + 1, 0);
    for (size_t i = 0; i < m->segs.size(); i++) prefix[i + 1] = prefix[i] + m->segs[i].len;
    return prefix;
}
static Located locate(const MultiIndex* m, const std::vector<int64_t>& prefix, int64_t pos, size_t* seg_out = nullptr) {
    const size_t i = (size_t)(std::upper_bound(prefix.begin(), prefix.end(), pos) - prefix.begin()) - 1;
    if (seg_out) *seg_out = i;
    return {m->segs[i].shard, m->segs[i].lstart + (pos - prefix[i])};
}

// ---------------------------------------------------------------------------
// create / destroy
// ---------------------------------------------------------------------------
void multi_free(b200r_index* ix) {
    if (!ix) return;
    MultiIndex* m = ix->multi;
    if (m) {
        if (m->workers) {
            m->workers->stop.store(true);
            {
                std::lock_guard<std::mutex> lk(m->workers->mu);
                m->workers->cv.notify_all();
            }
            for (auto& t : m->workers->th) t.join();
            delete m->workers;
        }
        for (Shard& s : m->sh) {
            DeviceGuard dg(s.device);
            if (s.stream) cudaStreamSynchronize(s.stream);
            cudaFree(s.q_dev);
            if (s.done_ev) cudaEventDestroy(s.done_ev);
            if (s.stream) cudaStreamDestroy(s.stream);
            if (s.leaf) leaf_free(s.leaf);
        }
        {
            DeviceGuard dg(m->primary);
            if (m->pstream) cudaStreamSynchronize(m->pstream);
            cudaFree(m->gather);
            cudaFree(m->xbuf);
            cudaFree(m->seg_off_dev);
            cudaFree(m->seg_tab_dev);
            cudaFreeHost(m->q_pin);
            cudaFree(m->out_dev);
            cudaFreeHost(m->out_pin);
            if (m->merge_ev) cudaEventDestroy(m->merge_ev);
            if (m->t0) cudaEventDestroy(m->t0);
            if (m->t1) cudaEventDestroy(m->t1);
            if (m->pstream) cudaStreamDestroy(m->pstream);
        }
        for (int d = 0; d < 64; d++)
            if (m->caller_ev[d]) {
                DeviceGuard dg(d);
                cudaEventDestroy(m->caller_ev[d]);
            }
        cudaGetLastError();
        delete m;
    }
    {
        DeviceGuard dg(ix->device);
        free_rerank_ws(ix);
    }
    delete ix;
}

static int enable_peer(int from, int to) {
    if (from == to) return B200R_OK;
    // remembered per process: asking the driver again costs far more than a search step
    static std::mutex mu;
    static bool enabled[64][64] = {};
    if (from < 0 || to < 0 || from >= 64 || to >= 64) return fail(B200R_E_INVALID, "device ordinal out of range");
    std::lock_guard<std::mutex> lk(mu);
    if (enabled[from][to]) return B200R_OK;
    int can = 0;
    CU(cudaDeviceCanAccessPeer(&can, from, to));
    if (!can) return fail(B200R_E_UNSUPPORTED, "GPU %d cannot address GPU %d's memory (no peer access)", from, to);
    DeviceGuard dg(from);
    cudaError_t e = cudaDeviceEnablePeerAccess(to, 0);
    if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
        e = cudaSuccess;
    }
    CU(e);
    enabled[from][to] = true;
    return B200R_OK;
}

int multi_create(const b200r_config* cfg, b200r_index** out) {
    const int G = cfg->n_devices;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(B200R_E_CUDA, "no CUDA device available (%s); libb200r has no CPU path",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    }
    // (a device may be listed more than once: every entry is one shard, so two entries put two
    // shards on that GPU -- of no use for speed, but it runs this whole code path on a 1-GPU box)
    for (int g = 0; g < G; g++)
        if (cfg->device_ids[g] < 0 || cfg->device_ids[g] >= ndev)
            return fail(B200R_E_INVALID, "device_ids[%d] = %d out of range (%d visible)", g, cfg->device_ids[g], ndev);
    b200r_index* ix = new b200r_index();
    MultiIndex* m = new MultiIndex();
    ix->multi = m;
    ix->cfg = *cfg;
    ix->cfg.shadow_bf16 = cfg->shadow_bf16 ? 1 : 0;
    ix->d = cfg->dim;
    ix->bf16 = cfg->storage == B200R_BF16;
    ix->l2 = cfg->metric == B200R_METRIC_L2;
    ix->device = cfg->device_ids[0];
    m->G = G;
    m->primary = cfg->device_ids[0];
    m->sh.resize(G);
    m->rc.assign(G, B200R_OK);
    m->err.resize(G);
    if (const char* ev = getenv("B200R_MULTI_MIN_PIECE")) m->min_piece = std::max<long long>(1, atoll(ev));
    int rc = B200R_OK;
    auto body = [&]() -> int {
        for (int a = 0; a < G; a++)
            for (int b = 0; b < G; b++) {
                int r = enable_peer(cfg->device_ids[a], cfg->device_ids[b]);
                if (r != B200R_OK) return r;
            }
        for (int g = 0; g < G; g++) {
            Shard& s = m->sh[g];
            s.device = cfg->device_ids[g];
            b200r_config c = *cfg;
            c.n_devices = 0;
            c.device = s.device;
            c.id_base = 0;  // shards answer in local row numbers
            c.reserve_rows = cfg->reserve_rows > 0 ? (cfg->reserve_rows + G - 1) / G + m->min_piece : 0;
            int r = leaf_create(&c, &s.leaf);
            if (r != B200R_OK) return r;
            s.leaf->timing_deferred = true;
            DeviceGuard dg(s.device);
            CU(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
            CU(cudaEventCreateWithFlags(&s.done_ev, cudaEventDisableTiming));
        }
        DeviceGuard dg(m->primary);
        CU(cudaStreamCreateWithFlags(&m->pstream, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&m->merge_ev, cudaEventDisableTiming));
        CU(cudaEventCreate(&m->t0));
        CU(cudaEventCreate(&m->t1));
        CU(cudaMalloc(&m->xbuf, kXchgBytes));
        CU(cudaMemset(m->xbuf, 0, kXchgBytes));
        CU(cudaMalloc(&m->seg_off_dev, (size_t)(B200R_MAX_DEVICES + 1) * sizeof(int)));
        CU(cudaMemset(m->seg_off_dev, 0, (size_t)(B200R_MAX_DEVICES + 1) * sizeof(int)));
        return upload_segments(m);
    };
    rc = body();
    if (rc != B200R_OK) {
        const std::string keep = last_error_cstr();
        multi_free(ix);
        return fail(rc, "%s", keep.c_str());
    }
    const int want_workers = knobs().multi_workers;
    if (want_workers != 0) {
        m->workers = new Workers();
        for (int g = 1; g < G; g++) m->workers->th.emplace_back(worker_main, m, g);
    }
    m->st.n_devices = G;
    m->st.num_sms = m->sh[0].leaf->num_sms;
    *out = ix;
    return B200R_OK;
}

int64_t multi_ntotal(const b200r_index* ix) { return ix->multi->ntotal; }

// ---------------------------------------------------------------------------
// add
// ---------------------------------------------------------------------------
// how many of the n new rows each shard takes: small batches whole to the emptiest shard, large
// ones so that the shards level out ("water filling"), no piece below min_piece
static std::vector<int64_t> place_rows(const MultiIndex* m, int64_t n) {
    const int G = m->G;
    std::vector<int64_t> cnt(G), piece(G, 0);
    for (int g = 0; g < G; g++) cnt[g] = m->sh[g].leaf->ntotal;
    if (n < 2 * m->min_piece) {
        int best = 0;
        for (int g = 1; g < G; g++)
            if (cnt[g] < cnt[best]) best = g;
        piece[best] = n;
        return piece;
    }
    int64_t total = n;
    for (int g = 0; g < G; g++) total += cnt[g];
    // the level L with sum max(0, L - cnt) >= n
    int64_t lo = total / G, hi = total;  // hi: everything on one shard is always enough
    auto need_at = [&](int64_t L) {
        int64_t s = 0;
        for (int g = 0; g < G; g++) s += std::max<int64_t>(0, L - cnt[g]);
        return s;
    };
    while (lo < hi) {
        const int64_t mid = lo + (hi - lo) / 2;
        if (need_at(mid) >= n)
            hi = mid;
        else
            lo = mid + 1;
    }
    int64_t left = n;
    for (int g = 0; g < G && left > 0; g++) {
        const int64_t want = std::min(left, std::max<int64_t>(0, lo - cnt[g]));
        piece[g] = want;
        left -= want;
    }
    // fold pieces that are too small into the largest one
    int big = 0;
    for (int g = 1; g < G; g++)
        if (piece[g] > piece[big]) big = g;
    piece[big] += left;
    for (int g = 0; g < G; g++)
        if (g != big && piece[g] > 0 && piece[g] < m->min_piece) {
            piece[big] += piece[g];
            piece[g] = 0;
        }
    return piece;
}

static int device_of_pointer(const void* p, int* dev_out) {
    cudaPointerAttributes pa{};
    if (cudaPointerGetAttributes(&pa, p) != cudaSuccess || pa.type != cudaMemoryTypeDevice) {
        cudaGetLastError();
        return fail(B200R_E_INVALID, "a buffer passed as B200R_DEVICE is not device memory");
    }
    *dev_out = pa.device;
    return B200R_OK;
}

static cudaEvent_t caller_event(MultiIndex* m, int dev) {
    if (dev < 0 || dev >= 64) return nullptr;
    if (!m->caller_ev[dev]) {
        DeviceGuard dg(dev);
        if (cudaEventCreateWithFlags(&m->caller_ev[dev], cudaEventDisableTiming) != cudaSuccess) {
            cudaGetLastError();
            m->caller_ev[dev] = nullptr;
        }
    }
    return m->caller_ev[dev];
}

static int multi_add_locked(b200r_index* ix, const void* x, int64_t n, int dtype, int mem_kind, void* stream,
                            bool allow_normalize) {
    MultiIndex* m = ix->multi;
    if (n < 0) return fail(B200R_E_INVALID, "b200r_add: n = %lld", (long long)n);
    if (n == 0) return B200R_OK;
    if (!x) return fail(B200R_E_INVALID, "b200r_add: null data");
    if (dtype != B200R_F32) return fail(B200R_E_UNSUPPORTED, "b200r_add: input vectors must be fp32");
    if (mem_kind != B200R_HOST && mem_kind != B200R_DEVICE) return fail(B200R_E_INVALID, "b200r_add: bad mem_kind");
    const std::vector<int64_t> piece = place_rows(m, n);
    int caller_dev = -1;
    cudaEvent_t in_ev = nullptr;
    if (mem_kind == B200R_DEVICE) {
        int src_dev = -1;
        int rc = device_of_pointer(x, &src_dev);
        if (rc != B200R_OK) return rc;
        for (int g = 0; g < m->G; g++) {
            rc = enable_peer(m->sh[g].device, src_dev);
            if (rc != B200R_OK) return rc;
        }
        // the caller's stream belongs to the caller's current device: everything it has enqueued so far
        // (the producer of x) precedes the ingest on every shard
        CU(cudaGetDevice(&caller_dev));
        in_ev = caller_event(m, caller_dev);
        if (!in_ev) return fail(B200R_E_CUDA, "cannot create an event on device %d", caller_dev);
        CU(cudaEventRecord(in_ev, (cudaStream_t)stream));
    }
    const size_t in_row = (size_t)ix->d * sizeof(float);
    int64_t off = 0;
    for (int g = 0; g < m->G; g++) {
        if (piece[g] == 0) continue;
        Shard& s = m->sh[g];
        const int64_t before = s.leaf->ntotal;
        const uint8_t* src = reinterpret_cast<const uint8_t*>(x) + (size_t)off * in_row;
        int rc;
        if (mem_kind == B200R_DEVICE) {
            DeviceGuard dg(s.device);
            CU(cudaStreamWaitEvent(s.stream, in_ev, 0));
            // the ingest kernel of this shard reads x where it lies (plain loads, over NVLink for a peer)
            rc = leaf_add(s.leaf, src, piece[g], dtype, B200R_DEVICE, s.stream, allow_normalize);
        } else {
            rc = leaf_add(s.leaf, src, piece[g], dtype, B200R_HOST, nullptr, allow_normalize);
        }
        if (rc != B200R_OK) {
            // earlier pieces stay added (a prefix of x): publish the map of what succeeded, then report
            const std::string keep = last_error_cstr();
            coalesce_segments(m);
            m->st.n_rows_added += off;
            upload_segments(m);
            return fail(rc, "%s", keep.c_str());
        }
        m->segs.push_back({g, before, piece[g]});
        m->ntotal += piece[g];
        off += piece[g];
    }
    if (mem_kind == B200R_DEVICE) {
        // the caller may reuse x once its stream has passed this point
        for (int g = 0; g < m->G; g++) {
            if (piece[g] == 0) continue;
            Shard& s = m->sh[g];
            DeviceGuard dg(s.device);
            CU(cudaEventRecord(s.done_ev, s.stream));
            DeviceGuard dc(caller_dev);
            CU(cudaStreamWaitEvent((cudaStream_t)stream, s.done_ev, 0));
        }
    }
    coalesce_segments(m);
    m->st.n_rows_added += n;
    return upload_segments(m);
}

int multi_add(b200r_index* ix, const void* x, int64_t n, int dtype, int mem_kind, void* stream, bool allow_normalize) {
    std::lock_guard<std::mutex> lk(ix->mu);
    return multi_add_locked(ix, x, n, dtype, mem_kind, stream, allow_normalize);
}

static int quiesce(MultiIndex* m);

// ---------------------------------------------------------------------------
// search
// ---------------------------------------------------------------------------
static int ensure_search_buffers(MultiIndex* m, int d, int64_t nq, int k) {
    DeviceGuard dg(m->primary);
    if (nq > m->q_pin_cap) {
        cudaFreeHost(m->q_pin);
        m->q_pin = nullptr;
        m->q_pin_cap = 0;
        const int64_t cap = std::max<int64_t>(nq, 256);
        CU(cudaMallocHost(&m->q_pin, (size_t)cap * d * sizeof(float)));
        m->q_pin_cap = cap;
    }
    const int64_t need = nq * k;
    if (need > m->out_cap) {
        cudaFree(m->out_dev);
        cudaFreeHost(m->out_pin);
        cudaFree(m->gather);
        m->out_dev = m->out_pin = m->gather = nullptr;
        m->out_cap = 0;
        const int64_t cap = std::max<int64_t>(need, 4096);
        const size_t slot = (size_t)cap * 12 + 64;
        CU(cudaMalloc(&m->out_dev, slot));
        CU(cudaMallocHost(&m->out_pin, slot));
        CU(cudaMalloc(&m->gather, slot * m->G));
        m->gather_bytes = slot * m->G;
        m->out_cap = cap;
    }
    return B200R_OK;
}

int multi_search(b200r_index* ix, const void* q, int64_t nq, int k, int q_mem_kind, float* D, int64_t* I,
                 int out_mem_kind, void* stream) {
    MultiIndex* m = ix->multi;
    if (nq < 0) return fail(B200R_E_INVALID, "b200r_search: nq = %lld", (long long)nq);
    if (k <= 0) return fail(B200R_E_INVALID, "b200r_search: k must be > 0, got %d", k);  // faiss: assert k > 0
    if (nq == 0) return B200R_OK;
    if (!q || !D || !I) return fail(B200R_E_INVALID, "b200r_search: null buffer");
    if ((q_mem_kind != B200R_HOST && q_mem_kind != B200R_DEVICE) ||
        (out_mem_kind != B200R_HOST && out_mem_kind != B200R_DEVICE))
        return fail(B200R_E_INVALID, "b200r_search: bad mem_kind");
    if (k > kMaxK) return fail(B200R_E_UNSUPPORTED, "b200r_search: k = %d exceeds the supported maximum %d", k, kMaxK);
    std::lock_guard<std::mutex> lk(ix->mu);
    const int G = m->G;
    const int d = ix->d;
    int rc = ensure_search_buffers(m, d, nq, k);
    if (rc != B200R_OK) return rc;

    const bool any_dev = q_mem_kind == B200R_DEVICE || out_mem_kind == B200R_DEVICE;
    int caller_dev = -1;
    cudaEvent_t in_ev = nullptr;
    if (any_dev) {
        int pd = -1;
        if (q_mem_kind == B200R_DEVICE) {
            rc = device_of_pointer(q, &pd);
            if (rc != B200R_OK) return rc;
            for (int g = 0; g < G; g++)
                if ((rc = enable_peer(m->sh[g].device, pd)) != B200R_OK) return rc;
        }
        if (out_mem_kind == B200R_DEVICE) {
            rc = device_of_pointer(D, &pd);
            if (rc == B200R_OK) rc = enable_peer(m->primary, pd);
            if (rc == B200R_OK) rc = device_of_pointer(I, &pd);
            if (rc == B200R_OK) rc = enable_peer(m->primary, pd);
            if (rc != B200R_OK) return rc;
        }
        CU(cudaGetDevice(&caller_dev));
        in_ev = caller_event(m, caller_dev);
        if (!in_ev) return fail(B200R_E_CUDA, "cannot create an event on device %d", caller_dev);
        CU(cudaEventRecord(in_ev, (cudaStream_t)stream));
    }
    if (q_mem_kind == B200R_HOST) memcpy(m->q_pin, q, (size_t)nq * d * sizeof(float));

    // route: small batches that every shard would scan take the fused exchange (the scan kernel's
    // final writer fills the slot and raises a flag); everything else fills the slot from whatever
    // kernels the shard's own dispatch picks, and signals with an event
    const Knobs& kn = knobs();
    bool fused = kn.multi_fused != 0 && nq <= kScanMaxNQ && k <= kScanMaxK;
    if (fused)
        for (int g = 0; g < G && fused; g++) {
            b200r_index* lf = m->sh[g].leaf;
            if (leaf_route(lf, nq) != B200R_PATH_SCAN) fused = false;
            if (nq > 1 && scan_queries_per_launch(lf->d_pad, lf->row_bytes, lf->smem_optin) < kScanMaxNQ) fused = false;
        }
    const size_t ids_off = (size_t)round_up64(nq * k * (int64_t)sizeof(float), 16);
    const size_t slot_bytes = ids_off + (size_t)nq * k * sizeof(int64_t);
    // epochs are never 0 and alternate in parity (two are buffered): after 0xFFFFFFFF comes 2
    if (fused) m->epoch = m->epoch == 0xFFFFFFFFu ? 2u : m->epoch + 1u;
    const uint32_t epoch = fused ? m->epoch : 0;
    const bool timing = m->timing;
    const bool wait_prev = m->prev_async;

    // The final select runs on the primary GPU.  On the fused route it goes on the primary shard's OWN
    // stream, behind that shard's scan: launched ahead of the scan (a caller that runs ahead of the GPU
    // does that) it would sit on an SM polling for flags and keep one of the scan's persistent CTAs
    // from becoming resident until the rest of the grid has finished (measured: 2.3 -> 3.4 ms per step).
    cudaStream_t fin = fused ? m->sh[0].stream : m->pstream;
    if (timing) {
        DeviceGuard dg(m->primary);
        CU(cudaEventRecord(m->t0, fin));
    }
    rc = run_on_shards(m, [&](int g) -> int {
        Shard& s = m->sh[g];
        // (the calling thread / the worker has this shard's device current)
        if (wait_prev) CU(cudaStreamWaitEvent(s.stream, m->merge_ev, 0));  // the slots of the last search have been read
        if (in_ev) CU(cudaStreamWaitEvent(s.stream, in_ev, 0));
        const float* qg = reinterpret_cast<const float*>(q);
        if (q_mem_kind == B200R_HOST) {
            if (nq > s.q_cap) {
                CU(cudaStreamSynchronize(s.stream));
                cudaFree(s.q_dev);
                s.q_dev = nullptr;
                s.q_cap = 0;
                const int64_t cap = std::max<int64_t>(nq, 256);
                CU(cudaMalloc(&s.q_dev, (size_t)cap * d * sizeof(float)));
                s.q_cap = cap;
            }
            CU(cudaMemcpyAsync(s.q_dev, m->q_pin, (size_t)nq * d * sizeof(float), cudaMemcpyHostToDevice, s.stream));
            qg = s.q_dev;
        }
        s.leaf->timing = timing;
        if (fused) {
            ScanExchange x{};
            x.n_targets = 1;
            x.slot = g;
            x.epoch = epoch;
            x.target[0] = m->xbuf;
            return leaf_search(s.leaf, qg, nq, k, B200R_DEVICE, nullptr, nullptr, B200R_DEVICE, s.stream, B200R_PATH_SCAN, &x);
        }
        uint8_t* slot = m->gather + (size_t)g * slot_bytes;
        int r = leaf_search(s.leaf, qg, nq, k, B200R_DEVICE, reinterpret_cast<float*>(slot),
                            reinterpret_cast<int64_t*>(slot + ids_off), B200R_DEVICE, s.stream, B200R_PATH_AUTO, nullptr);
        if (r != B200R_OK) return r;
        CU(cudaEventRecord(s.done_ev, s.stream));
        return B200R_OK;
    });
    if (rc != B200R_OK) return rc;

    DeviceGuard dg(m->primary);
    // small host results: the select kernel writes straight into the pinned host block (mapped memory)
    const bool zc = out_mem_kind == B200R_HOST && kn.zero_copy_out != 0 && slot_bytes <= ((size_t)16 << 10);
    uint8_t* stage = zc ? m->out_pin : m->out_dev;
    float* od = out_mem_kind == B200R_DEVICE ? D : reinterpret_cast<float*>(stage);
    int64_t* oi = out_mem_kind == B200R_DEVICE ? I : reinterpret_cast<int64_t*>(stage + ids_off);
    const int* seg_off = m->seg_off_dev;
    const int64_t* seg_ls = m->seg_tab_dev;
    const int64_t* seg_gb = m->seg_tab_dev + m->seg_tab_cap;
    // host results carry one more word: did every shard deliver (fused route)?
    const size_t status_off = (size_t)round_up64((int64_t)slot_bytes, 16);
    int* status_dev = out_mem_kind == B200R_HOST ? reinterpret_cast<int*>(stage + status_off) : nullptr;
    if (fused) {
        const size_t par = (size_t)(epoch & 1u) * kXchgParityBytes;
        CU(launch_merge_shards(ix->l2, m->xbuf + par, kXchgSlotBytes, kXchgIdOffset, G, (int)nq, k, k, seg_off, seg_ls,
                               seg_gb, ix->cfg.id_base, od, oi, m->xbuf, epoch, status_dev, fin));
    } else {
        for (int g = 0; g < G; g++) CU(cudaStreamWaitEvent(fin, m->sh[g].done_ev, 0));
        CU(launch_merge_shards(ix->l2, m->gather, (int64_t)slot_bytes, (int64_t)ids_off, G, (int)nq, k, k, seg_off, seg_ls,
                               seg_gb, ix->cfg.id_base, od, oi, nullptr, 0, status_dev, fin));
    }
    m->st.n_kernel_launches++;
    if (timing) CU(cudaEventRecord(m->t1, fin));
    CU(cudaEventRecord(m->merge_ev, fin));
    if (out_mem_kind == B200R_HOST) {
        if (!zc) CU(cudaMemcpyAsync(m->out_pin, m->out_dev, status_off + sizeof(int), cudaMemcpyDeviceToHost, fin));
        CU(cudaStreamSynchronize(fin));
        m->prev_async = false;
        const int status = *reinterpret_cast<const int*>(m->out_pin + status_off);
        if (status != 0) {
            // a shard's result never arrived (its GPU stalled for seconds): let everything drain so
            // that the next search starts from a clean state, and report it
            quiesce(m);
            return fail(B200R_E_CUDA, "multi-GPU search: GPU %d did not deliver its candidates within 2 s",
                        m->sh[std::min(G - 1, status - 1)].device);
        }
        memcpy(D, m->out_pin, (size_t)nq * k * sizeof(float));
        memcpy(I, m->out_pin + ids_off, (size_t)nq * k * sizeof(int64_t));
    } else {
        // the caller's stream continues after the final select; host queries went through the one
        // pinned buffer, which must be free again before the next call
        DeviceGuard dc(caller_dev);
        CU(cudaStreamWaitEvent((cudaStream_t)stream, m->merge_ev, 0));
        if (q_mem_kind == B200R_HOST) {
            DeviceGuard dp(m->primary);
            CU(cudaStreamSynchronize(fin));
            m->prev_async = false;
        } else {
            m->prev_async = true;
        }
    }
    if (timing) {
        DeviceGuard dp(m->primary);
        CU(cudaEventSynchronize(m->t1));
        float total_ms = 0.f;
        CU(cudaEventElapsedTime(&total_ms, m->t0, m->t1));
        double kmax = 0.0;
        for (int g = 0; g < G; g++) {
            double ms = 0.0;
            if (leaf_last_search_ms(m->sh[g].leaf, &ms) == B200R_OK) kmax = std::max(kmax, ms);
        }
        m->st.last_search_ms = total_ms;
        m->st.last_kernel_us = kmax * 1000.0;
        m->st.last_exchange_us = std::max(0.0, (double)total_ms - kmax) * 1000.0;
    }
    m->st.n_search_calls++;
    m->st.n_queries += nq;
    m->st.last_path = m->sh[0].leaf->st.last_path;
    return B200R_OK;
}

// ---------------------------------------------------------------------------
// remove_ids / reconstruct / reset / read rows
// ---------------------------------------------------------------------------
static int quiesce(MultiIndex* m) {
    for (Shard& s : m->sh) {
        DeviceGuard dg(s.device);
        CU(cudaStreamSynchronize(s.stream));
    }
    DeviceGuard dg(m->primary);
    CU(cudaStreamSynchronize(m->pstream));
    m->prev_async = false;
    return B200R_OK;
}

int multi_remove_ids(b200r_index* ix, const int64_t* ids, int64_t n, int64_t* n_removed) {
    MultiIndex* m = ix->multi;
    if (n < 0 || (n > 0 && !ids)) return fail(B200R_E_INVALID, "b200r_remove_ids: bad ids");
    if (n_removed) *n_removed = 0;
    std::lock_guard<std::mutex> lk(ix->mu);
    std::vector<int64_t> del;
    del.reserve((size_t)n);
    for (int64_t i = 0; i < n; i++)
        if (ids[i] >= 0 && ids[i] < m->ntotal) del.push_back(ids[i]);
    std::sort(del.begin(), del.end());
    del.erase(std::unique(del.begin(), del.end()), del.end());
    if (del.empty()) return B200R_OK;
    int rc = quiesce(m);
    if (rc != B200R_OK) return rc;
    const std::vector<int64_t> prefix = segment_prefix(m);
    std::vector<std::vector<int64_t>> local(m->G);
    std::vector<int64_t> seg_removed(m->segs.size(), 0);
    for (int64_t pos : del) {
        size_t si = 0;
        const Located at = locate(m, prefix, pos, &si);
        local[at.shard].push_back(at.local);
        seg_removed[si]++;
    }
    // every shard compacts its own rows (stable): no row crosses a GPU boundary
    rc = run_on_shards(m, [&](int g) -> int {
        if (local[g].empty()) return B200R_OK;
        int64_t got = 0;
        int r = leaf_remove_ids(m->sh[g].leaf, local[g].data(), (int64_t)local[g].size(), &got);
        if (r == B200R_OK && got != (int64_t)local[g].size())
            return fail(B200R_E_CUDA, "shard removed %lld of %lld rows", (long long)got, (long long)local[g].size());
        return r;
    });
    // a shard that failed kept its rows: the map follows what each shard actually did, so the index
    // stays consistent, and the error is reported after the map has been published
    const std::string keep = rc != B200R_OK ? last_error_cstr() : "";
    std::vector<int64_t> shard_pos(m->G, 0);  // rows of each shard seen so far, after the removal
    int64_t gone = 0;
    for (size_t i = 0; i < m->segs.size(); i++) {
        Segment& s = m->segs[i];
        if (m->rc[s.shard] == B200R_OK) {
            s.len -= seg_removed[i];
            gone += seg_removed[i];
        }
        s.lstart = shard_pos[s.shard];
        shard_pos[s.shard] += s.len;
    }
    coalesce_segments(m);
    m->ntotal -= gone;
    m->st.n_rows_removed += gone;
    if (n_removed) *n_removed = gone;
    const int up = upload_segments(m);
    if (rc != B200R_OK) return fail(rc, "%s", keep.c_str());
    return up;
}

int multi_reconstruct_batch(b200r_index* ix, const int64_t* keys, int64_t n, float* out) {
    MultiIndex* m = ix->multi;
    if (n < 0 || (n > 0 && (!keys || !out))) return fail(B200R_E_INVALID, "b200r_reconstruct_batch: bad arguments");
    if (n == 0) return B200R_OK;
    std::lock_guard<std::mutex> lk(ix->mu);
    for (int64_t i = 0; i < n; i++)
        if (keys[i] < 0 || keys[i] >= m->ntotal)
            return fail(B200R_E_INVALID, "b200r_reconstruct_batch: key %lld out of range [0, %lld)", (long long)keys[i],
                        (long long)m->ntotal);
    const std::vector<int64_t> prefix = segment_prefix(m);
    std::vector<std::vector<int64_t>> local(m->G), where(m->G);
    for (int64_t i = 0; i < n; i++) {
        const Located at = locate(m, prefix, keys[i]);
        local[at.shard].push_back(at.local);
        where[at.shard].push_back(i);
    }
    const size_t row = (size_t)ix->d * sizeof(float);
    std::vector<std::vector<float>> tmp(m->G);
    int rc = run_on_shards(m, [&](int g) -> int {
        if (local[g].empty()) return B200R_OK;
        tmp[g].resize(local[g].size() * (size_t)ix->d);
        return leaf_reconstruct_batch(m->sh[g].leaf, local[g].data(), (int64_t)local[g].size(), tmp[g].data());
    });
    if (rc != B200R_OK) return rc;
    for (int g = 0; g < m->G; g++)
        for (size_t j = 0; j < where[g].size(); j++)
            memcpy(reinterpret_cast<uint8_t*>(out) + (size_t)where[g][j] * row, tmp[g].data() + j * (size_t)ix->d, row);
    return B200R_OK;
}

// rows [r0, r0+n) of the global order as fp32 into a host buffer
static int multi_read_rows_locked(b200r_index* ix, int64_t r0, int64_t n, float* out) {
    MultiIndex* m = ix->multi;
    if (r0 < 0 || n < 0 || r0 + n > m->ntotal) return fail(B200R_E_INVALID, "row range out of bounds");
    int64_t gpos = 0;
    for (const Segment& s : m->segs) {
        const int64_t lo = std::max(r0, gpos), hi = std::min(r0 + n, gpos + s.len);
        if (lo < hi) {
            int rc = leaf_read_rows(m->sh[s.shard].leaf, s.lstart + (lo - gpos), hi - lo, out + (size_t)(lo - r0) * ix->d);
            if (rc != B200R_OK) return rc;
        }
        gpos += s.len;
        if (gpos >= r0 + n) break;
    }
    return B200R_OK;
}

int multi_read_rows(b200r_index* ix, int64_t r0, int64_t n, float* out) {
    std::lock_guard<std::mutex> lk(ix->mu);
    return multi_read_rows_locked(ix, r0, n, out);
}

int multi_reset(b200r_index* ix) {
    MultiIndex* m = ix->multi;
    std::lock_guard<std::mutex> lk(ix->mu);
    int rc = quiesce(m);
    if (rc != B200R_OK) return rc;
    for (Shard& s : m->sh) leaf_reset(s.leaf);
    m->segs.clear();
    m->ntotal = 0;
    return upload_segments(m);
}

// ---------------------------------------------------------------------------
// merge_from: either side (or both) spans several GPUs.  The source's rows, in its global order,
// are turned back into fp32 on their own GPU and ingested by the destination from there (plain
// loads over NVLink where the GPUs differ); then the source is emptied, as faiss does.
// ---------------------------------------------------------------------------
int multi_merge_from(b200r_index* dst, b200r_index* src) {
    if (dst->cfg.dim != src->cfg.dim || dst->cfg.metric != src->cfg.metric || dst->cfg.storage != src->cfg.storage)
        return fail(B200R_E_INVALID, "b200r_merge_from: indexes are not compatible (dim %d/%d, metric or storage differ)",
                    dst->cfg.dim, src->cfg.dim);
    // only the handles of multi-GPU indexes are locked here: single-GPU ones lock inside their own calls
    std::unique_lock<std::mutex> l1, l2;
    {
        b200r_index* a = dst->multi ? dst : nullptr;
        b200r_index* b = src->multi ? src : nullptr;
        if (a && b && b < a) std::swap(a, b);
        if (a) l1 = std::unique_lock<std::mutex>(a->mu);
        if (b) l2 = std::unique_lock<std::mutex>(b->mu);
    }
    struct Run {
        b200r_index* leaf;
        int64_t r0, len;
    };
    std::vector<Run> runs;
    if (src->multi) {
        int rc = quiesce(src->multi);
        if (rc != B200R_OK) return rc;
        for (const Segment& s : src->multi->segs) runs.push_back({src->multi->sh[s.shard].leaf, s.lstart, s.len});
    } else if (src->ntotal > 0) {
        runs.push_back({src, 0, src->ntotal});
    }
    const int d = dst->cfg.dim;
    const int64_t chunk = std::max<int64_t>(1, (int64_t)((64u << 20) / ((size_t)d * sizeof(float))));
    for (const Run& r : runs) {
        DeviceGuard dg(r.leaf->device);
        float* tmp = nullptr;
        cudaError_t e = cudaMalloc(&tmp, (size_t)std::min(chunk, r.len) * d * sizeof(float));
        if (e != cudaSuccess) {
            cudaGetLastError();
            return fail(B200R_E_NOMEM, "b200r_merge_from: staging buffer: %s", cudaGetErrorString(e));
        }
        int rc = B200R_OK;
        for (int64_t c0 = 0; rc == B200R_OK && c0 < r.len; c0 += chunk) {
            const int64_t mrows = std::min(chunk, r.len - c0);
            rc = leaf_rows_to_device_f32(r.leaf, r.r0 + c0, mrows, tmp);
            if (rc != B200R_OK) break;
            if (dst->multi) {
                rc = multi_add_locked(dst, tmp, mrows, B200R_F32, B200R_DEVICE, r.leaf->stream, false);
                if (rc == B200R_OK) rc = quiesce(dst->multi);
            } else {
                rc = enable_peer(dst->device, r.leaf->device);
                if (rc == B200R_OK) rc = leaf_add(dst, tmp, mrows, B200R_F32, B200R_DEVICE, dst->stream, false);
                if (rc == B200R_OK) {
                    DeviceGuard dd(dst->device);
                    if (cudaStreamSynchronize(dst->stream) != cudaSuccess) rc = fail(B200R_E_CUDA, "b200r_merge_from: sync failed");
                }
            }
            // tmp is rewritten by the next chunk: everything that read it has finished above
            if (rc == B200R_OK && cudaStreamSynchronize(r.leaf->stream) != cudaSuccess)
                rc = fail(B200R_E_CUDA, "b200r_merge_from: sync failed");
        }
        cudaFree(tmp);
        if (rc != B200R_OK) return rc;
    }
    if (src->multi) {
        for (Shard& s : src->multi->sh) leaf_reset(s.leaf);
        src->multi->segs.clear();
        src->multi->ntotal = 0;
        return upload_segments(src->multi);
    }
    return leaf_reset(src);  // faiss: the other index is reset
}

// ---------------------------------------------------------------------------
// switches and statistics
// ---------------------------------------------------------------------------
int multi_set_shadow_filter(b200r_index* ix, int enabled) {
    std::lock_guard<std::mutex> lk(ix->mu);
    for (Shard& s : ix->multi->sh) {
        std::lock_guard<std::mutex> ll(s.leaf->mu);
        s.leaf->shadow_on = enabled != 0;
    }
    return B200R_OK;
}

int multi_set_timing(b200r_index* ix, int enabled) {
    std::lock_guard<std::mutex> lk(ix->mu);
    ix->multi->timing = enabled != 0;
    for (Shard& s : ix->multi->sh) s.leaf->timing = enabled != 0;
    return B200R_OK;
}

int multi_stats(b200r_index* ix, b200r_stats* out) {
    MultiIndex* m = ix->multi;
    std::lock_guard<std::mutex> lk(ix->mu);
    b200r_stats t = m->st;
    t.ntotal = m->ntotal;
    t.capacity_rows = 0;
    t.bytes_scanned = 0;
    t.n_gemm_fallbacks = 0;
    int64_t launches = m->st.n_kernel_launches;
    for (Shard& s : m->sh) {
        b200r_stats ls{};
        int rc = leaf_stats(s.leaf, &ls);
        if (rc != B200R_OK) return rc;
        t.capacity_rows += ls.capacity_rows;
        t.bytes_scanned += ls.bytes_scanned;
        t.n_gemm_fallbacks += ls.n_gemm_fallbacks;
        launches += ls.n_kernel_launches;
        t.last_grid = ls.last_grid;
        t.last_path = ls.last_path;
    }
    t.n_kernel_launches = launches;
    t.n_devices = m->G;
    t.n_segments = (int)m->segs.size();
    *out = t;
    return B200R_OK;
}

}  // namespace b200r
