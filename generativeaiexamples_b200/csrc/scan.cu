// K2: small-batch exact scorer with fused top-k ("scan" path, nq <= 4 per launch).
//
// Replaces, on the GPU, what faiss 1.7.4 runs for nq < 20 in
// IndexFlat::search -> knn_L2sqr / knn_inner_product -> exhaustive_*_seq
// (faiss/utils/distances.cpp): one fvec_inner_product / fvec_L2sqr per corpus row and a
// heap of the k best.  Reached in the reference from
// RAG/examples/basic_rag/langchain/chains.py:163 (retriever) via
// langchain FAISS.similarity_search_with_score_by_vector -> index.search(v, k).
//
// Design (HBM-bound; roofline = N * d * sizeof(elem) bytes per launch):
//  * persistent grid, one CTA per SM; tile t -> CTA (t mod grid), a tile is
//    `rows_per_tile` consecutive rows = one contiguous <= 64 KB block;
//  * warp 0 is the producer: one lane issues cp.async.bulk (TMA engine, SASS UBLKCP)
//    into a 3-deep shared-memory ring (192 KB in flight per SM), completion on mbarriers;
//  * 16 consumer warps: a warp owns whole rows, reads them with 128-bit LDS, keeps
//    fp32 partial sums, butterfly-reduces, and offers the score to its register
//    top-k list (WarpTopK).  Direct (q-x)^2 for L2, as faiss's nq<20 path does;
//  * epilogue: lists -> shared memory -> per-CTA merge -> global scratch; the last
//    CTA to finish (atomic ticket) merges all CTA lists and writes D/I.  The score
//    vector is never written to HBM.
#include "index.h"

#include <cstdlib>

namespace b200r {

// one launcher per <queries per launch, list slots per lane> group; defined in scan_inst_*.cu
cudaError_t scan_launch_1_1(const ScanParams& p, bool l2, bool bf16, int grid, size_t smem, cudaStream_t st, bool flagged);
cudaError_t scan_launch_1_4(const ScanParams& p, bool l2, bool bf16, int grid, size_t smem, cudaStream_t st, bool flagged);
cudaError_t scan_launch_1_8(const ScanParams& p, bool l2, bool bf16, int grid, size_t smem, cudaStream_t st, bool flagged);
cudaError_t scan_launch_1_16(const ScanParams& p, bool l2, bool bf16, int grid, size_t smem, cudaStream_t st, bool flagged);
cudaError_t scan_launch_1_32(const ScanParams& p, bool l2, bool bf16, int grid, size_t smem, cudaStream_t st, bool flagged);
cudaError_t scan_launch_4_1(const ScanParams& p, bool l2, bool bf16, int grid, size_t smem, cudaStream_t st, bool flagged);
cudaError_t scan_launch_4_4(const ScanParams& p, bool l2, bool bf16, int grid, size_t smem, cudaStream_t st, bool flagged);
constexpr int kScanMaxThreads = 32 * 17;  // the one-query shape: 1 producer + 16 consumer warps (scan_kernel.cuh)
constexpr int kScanMaxStages = 16;

size_t scan_scratch_bytes(int max_grid) {
    const size_t per_cta = kScanMaxNQ * kScanMaxK > kScanMaxKWide ? kScanMaxNQ * kScanMaxK : kScanMaxKWide;
    return (size_t)max_grid * per_cta * sizeof(Cand);
}

// Fills in the tiling fields of `p` and launches.  p.q must hold nq_slots(nq) rows.
// how many queries one launch can serve for rows this wide: the 4-query variants keep 4 x d floats
// of queries in shared memory beside at least two stages
int scan_queries_per_launch(int d_pad, int row_bytes, size_t smem_limit) {
    const size_t stage = row_bytes > 64 * 1024 ? (size_t)(row_bytes + 127) / 128 * 128 : (size_t)64 * 1024;
    const size_t fixed = 1024 + 2 * 16 * sizeof(uint64_t) + 2 * stage;
    return (fixed + (size_t)kScanMaxNQ * d_pad * sizeof(float) <= smem_limit) ? kScanMaxNQ : 1;
}

static cudaError_t launch_scan_impl(ScanParams p, int nq, bool l2, bool bf16, int num_sms, size_t smem_limit,
                                    cudaStream_t st, int* grid_out, bool flagged) {
    const int NQ = nq <= 1 ? 1 : 4;
    // list slots per lane: 32*KPL >= k.  Above 128 only the single-query variants exist.
    const int KPL = p.k <= 32 ? 1 : p.k <= 128 ? 4 : p.k <= 256 ? 8 : p.k <= 512 ? 16 : 32;
    if (nq < 1 || nq > kScanMaxNQ || p.k < 1 || p.k > kScanMaxKWide) return cudaErrorInvalidValue;
    if (p.k > kScanMaxK && nq != 1) return cudaErrorInvalidValue;
    smem_limit -= 1024;  // static shared + slack
    const size_t q_bytes = (size_t)NQ * p.d * sizeof(float);
    const size_t bar_bytes = 2 * 16 * sizeof(uint64_t);
    const size_t merge_bytes = (size_t)NQ * (kScanMaxThreads / 32) * KPL * 32 * sizeof(Cand);  // <= 139 KB (NQ=1, KPL=32)
    // stage: as many whole rows as fit in 32 KB (at least one)
    // 64 KB stages: 16 rows at d = 1024 fp32, i.e. one row per consumer warp per stage, and fewer,
    // larger bulk copies.  Measured on B200 (tools/tune_scan.py, 10M rows): 64 KB x 3 = 5.66 ms,
    // 48 KB x 4 = 5.69, 32 KB x 6 = 5.74, 16 KB x 12 = 6.39, 8 KB x 16 = 8.00.
    int stage_bytes = 64 * 1024;
    int max_stages = 3;
    if (knobs().scan_stage_kb > 0) stage_bytes = knobs().scan_stage_kb * 1024;  // tuning knobs
    if (knobs().scan_stages > 0) max_stages = knobs().scan_stages;
    if (p.row_bytes > stage_bytes) stage_bytes = (p.row_bytes + 127) / 128 * 128;
    int R = stage_bytes / p.row_bytes;
    if (smem_limit < q_bytes + bar_bytes + 2 * (size_t)stage_bytes) return cudaErrorInvalidValue;
    int nst = (int)((smem_limit - q_bytes - bar_bytes) / stage_bytes);
    if (nst > kScanMaxStages) nst = kScanMaxStages;
    if (nst > max_stages) nst = max_stages;
    if ((size_t)nst * stage_bytes < merge_bytes) return cudaErrorInvalidValue;
    p.nstages = nst;
    p.stage_bytes = stage_bytes;
    p.rows_per_tile = R;
    p.n_tiles = (p.n_rows + R - 1) / R;
    p.nq_active = nq;
    // mode 2 (the default) is for streams of many rounds; a corpus of a few rounds (L2-resident sizes, where a
    // search costs tens of microseconds) keeps the static walk and pays for no counter round trips
    if (p.tile_counter && scan_dynamic_mode() == 2 && p.n_tiles < 32ll * num_sms) p.tile_counter = nullptr;
    if (p.tile_counter) {
        // mode 2: the last eighth of the tiles (at least four rounds) is drawn dynamically
        long long tail = p.n_tiles / 8;
        const long long min_tail = 4ll * num_sms;
        if (tail < min_tail) tail = min_tail;
        p.dyn_from = scan_dynamic_mode() == 2 && p.n_tiles > tail ? p.n_tiles - tail : 0;
    }
    int grid = (int)(p.n_tiles < num_sms ? p.n_tiles : num_sms);
    if (grid < 1) grid = 1;
    if (grid_out) *grid_out = grid;
    const size_t smem = (size_t)nst * stage_bytes + q_bytes + bar_bytes;
    if (KPL == 8) return scan_launch_1_8(p, l2, bf16, grid, smem, st, flagged);
    if (KPL == 16) return scan_launch_1_16(p, l2, bf16, grid, smem, st, flagged);
    if (KPL == 32) return scan_launch_1_32(p, l2, bf16, grid, smem, st, flagged);
    if (NQ == 1)
        return KPL == 1 ? scan_launch_1_1(p, l2, bf16, grid, smem, st, flagged) : scan_launch_1_4(p, l2, bf16, grid, smem, st, flagged);
    return KPL == 1 ? scan_launch_4_1(p, l2, bf16, grid, smem, st, flagged) : scan_launch_4_4(p, l2, bf16, grid, smem, st, flagged);
}

cudaError_t launch_scan(ScanParams p, int nq, bool l2, bool bf16, int num_sms, size_t smem_limit, cudaStream_t st,
                        int* grid_out) {
    p.flag_list = nullptr;
    p.flag_count = nullptr;
    return launch_scan_impl(p, nq, l2, bf16, num_sms, smem_limit, st, grid_out, false);
}

// fail_flags[0..n) (non-zero = flagged) -> ascending list of flagged positions + their number
__global__ void compact_flags_kernel(const int* __restrict__ fail_flags, int64_t n, int* flag_list, int* flag_count) {
    __shared__ int s_base;
    __shared__ int s_warp[32];
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int64_t i0 = 0; i0 < n; i0 += blockDim.x) {
        const int64_t i = i0 + threadIdx.x;
        const bool f = i < n && fail_flags[i] != 0;
        const unsigned m = __ballot_sync(0xffffffffu, f);
        if (lane == 0) s_warp[warp] = __popc(m);
        __syncthreads();
        int off = s_base;
        for (int w = 0; w < warp; w++) off += s_warp[w];
        if (f) flag_list[off + __popc(m & ((1u << lane) - 1u))] = (int)i;
        __syncthreads();
        if (threadIdx.x == 0) {
            int t = 0;
            for (int w = 0; w < nw; w++) t += s_warp[w];
            s_base += t;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) *flag_count = s_base;
}

cudaError_t launch_scan_flagged(ScanParams p, const int* fail_flags, int64_t nq_total, int* flag_list, int* flag_count,
                                bool l2, bool bf16, int num_sms, size_t smem_limit, cudaStream_t st) {
    if (nq_total < 1 || nq_total > 0x7fffffff) return cudaErrorInvalidValue;
    compact_flags_kernel<<<1, 256, 0, st>>>(fail_flags, nq_total, flag_list, flag_count);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    p.flag_list = flag_list;
    p.flag_count = flag_count;
    p.xchg_ntargets = 0;
    // four flagged queries per corpus pass where the four-query variants exist and fit
    const int per_pass = p.k > kScanMaxK ? 1 : scan_queries_per_launch(p.d, p.row_bytes, smem_limit);
    return launch_scan_impl(p, per_pass, l2, bf16, num_sms, smem_limit, st, nullptr, true);
}

}  // namespace b200r

// ===========================================================================
// b200r_search_noalloc -- the scan search for callers that must not allocate.
//
// Who needs this: a caller whose device memory is already spoken for -- an inference
// server that handed the whole HBM to its own pool at start-up, so a cudaMalloc from a
// plug-in either fails or eats into someone else's budget -- and a caller inside a
// latency-critical serving loop, where cudaMalloc / cudaFree stall every stream of the
// device while they run.  b200r_search needs an index object, and that object owns
// cudaMalloc'ed workspace.  This entry point owns nothing: the corpus, the queries and
// the result arrays are the caller's device buffers, and the only extra global memory the
// kernel touches (the per-CTA candidate lists, the finish ticket and the zero-padded query
// slots) lives in the file-scope array below, which is part of the loaded module.
//
// Cost and limits of that choice: every GPU that loads libb200r.so carries the array
// (672,000 bytes) until the process exits; sizes are capped at what it was laid out for
// (dim <= 4096, nq <= 4, k <= 128, at most 148 CTAs); and there is exactly one array per
// device, so two calls whose kernels overlap on the same GPU (different streams or
// threads) race on it -- callers must serialise them.  The array keeps whatever the last
// call left; the ticket and the query slots are cleared on the stream before each launch,
// the candidate lists are fully rewritten by every launch before they are read.
//
// The first call on a device looks up the array's device address (cudaGetSymbolAddress
// with that device current; in host code the symbol's name is only a stand-in, not an
// address the GPU can use) and reads two device attributes.  Make that first call before
// any stream capture begins.
// ===========================================================================
#include "../../include/b200r.h"
#include <cstdio>
#include <mutex>

namespace b200r {

constexpr int kNoAllocMaxDim = 4096;
constexpr int kNoAllocMaxGrid = 148;
constexpr size_t kNoAllocTicketBytes = 256;
constexpr size_t kNoAllocCandBytes = (size_t)kNoAllocMaxGrid * kScanMaxNQ * kScanMaxK * sizeof(Cand);  // 606,208
constexpr size_t kNoAllocQueryBytes = (size_t)kScanMaxNQ * kNoAllocMaxDim * sizeof(float);             //  65,536
static_assert(kNoAllocCandBytes % 256 == 0, "sections of the scratch array stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocTicketBytes + kNoAllocCandBytes +
                                                                 kNoAllocQueryBytes];

struct NoAllocDevice {
    bool ready = false;
    unsigned char* base = nullptr;  // DEVICE address of g_noalloc_scratch on this GPU
    int num_sms = 0;
    size_t smem_optin = 0;
};
constexpr int kNoAllocMaxDevices = 64;
static NoAllocDevice g_noalloc_dev[kNoAllocMaxDevices];
static std::mutex g_noalloc_mu;

// Once per device, with that device current.
static cudaError_t noalloc_setup(int device, NoAllocDevice** out) {
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    NoAllocDevice& nd = g_noalloc_dev[device];
    if (!nd.ready) {
        unsigned char* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        int sms = 0, smem = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        if (e != cudaSuccess) return e;
        e = cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, device);
        if (e != cudaSuccess) return e;
        nd.base = ptr;
        nd.num_sms = sms < kNoAllocMaxGrid ? sms : kNoAllocMaxGrid;
        nd.smem_optin = (size_t)smem;
        nd.ready = true;
    }
    *out = &nd;
    return cudaSuccess;
}

static int noalloc_fail(int code, const char* fmt, long long a = 0, long long b = 0) {
    char buf[512];
    snprintf(buf, sizeof buf, fmt, a, b);
    set_last_error(buf);
    return code;
}

}  // namespace b200r

extern "C" int b200r_search_noalloc(const void* corpus, int64_t n_rows, int32_t dim, int32_t storage, int32_t metric,
                                    const float* q, int32_t nq, int32_t k, float* D, int64_t* I, int64_t id_base,
                                    int32_t device, void* stream) {
    using namespace b200r;
    if (storage != B200R_F32 && storage != B200R_BF16)
        return noalloc_fail(B200R_E_INVALID, "b200r_search_noalloc: storage must be B200R_F32 or B200R_BF16");
    if (metric != B200R_METRIC_INNER_PRODUCT && metric != B200R_METRIC_L2)
        return noalloc_fail(B200R_E_INVALID, "b200r_search_noalloc: bad metric");
    const bool bf16 = storage == B200R_BF16;
    if (dim <= 0 || n_rows < 0 || !q || !D || !I || (n_rows > 0 && !corpus))
        return noalloc_fail(B200R_E_INVALID, "b200r_search_noalloc: null buffer or negative size");
    if (dim > kNoAllocMaxDim)
        return noalloc_fail(B200R_E_UNSUPPORTED,
                            "b200r_search_noalloc: dim %lld exceeds the built-in scratch (max %lld); use b200r_search",
                            dim, kNoAllocMaxDim);
    if (nq < 1 || nq > kScanMaxNQ)
        return noalloc_fail(B200R_E_UNSUPPORTED,
                            "b200r_search_noalloc: nq %lld outside 1..%lld, the built-in scratch holds no more", nq,
                            kScanMaxNQ);
    if (k < 1 || k > kScanMaxK)
        return noalloc_fail(B200R_E_UNSUPPORTED,
                            "b200r_search_noalloc: k %lld outside 1..%lld, the built-in scratch holds no more", k,
                            kScanMaxK);
    if (dim % (bf16 ? 8 : 4) != 0 || (reinterpret_cast<uintptr_t>(corpus) & 15) != 0)
        return noalloc_fail(B200R_E_UNSUPPORTED,
                            "b200r_search_noalloc: rows must be 16-byte multiples at a 16-byte aligned address "
                            "(dim %lld); there is no room to re-pack them without allocating", dim);
    if (n_rows > 0xFFFFFFFEll) return noalloc_fail(B200R_E_UNSUPPORTED, "b200r_search_noalloc: too many rows");
    if (device < 0 || device >= kNoAllocMaxDevices)
        return noalloc_fail(B200R_E_INVALID, "b200r_search_noalloc: device %lld out of range", device);

    int prev = -1;
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != device && cudaSetDevice(device) != cudaSuccess)
        return noalloc_fail(B200R_E_CUDA, "b200r_search_noalloc: cudaSetDevice(%lld) failed", device);
    cudaStream_t st = (cudaStream_t)stream;
    NoAllocDevice* nd = nullptr;
    cudaError_t e = noalloc_setup(device, &nd);
    if (e == cudaSuccess) {
        unsigned* ticket = reinterpret_cast<unsigned*>(nd->base);
        Cand* cand = reinterpret_cast<Cand*>(nd->base + kNoAllocTicketBytes);
        float* qslots = reinterpret_cast<float*>(nd->base + kNoAllocTicketBytes + kNoAllocCandBytes);
        const int slots = nq <= 1 ? 1 : kScanMaxNQ;
        // what must start at zero: the ticket (a failed earlier launch may have left it
        // mid-count) and the query slots beyond nq
        e = cudaMemsetAsync(ticket, 0, kNoAllocTicketBytes, st);
        if (e == cudaSuccess) e = cudaMemsetAsync(qslots, 0, (size_t)slots * dim * sizeof(float), st);
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(qslots, q, (size_t)nq * dim * sizeof(float), cudaMemcpyDeviceToDevice, st);
        if (e == cudaSuccess) {
            ScanParams p{};
            p.corpus = corpus;
            p.n_rows = n_rows;
            p.d = dim;
            p.row_bytes = dim * (bf16 ? 2 : 4);
            p.q = qslots;
            p.k = k;
            p.evict_first = ((size_t)n_rows * p.row_bytes > (size_t)96 << 20) ? 1 : 0;
            p.scratch = cand;
            p.counter = ticket;
            p.out_d = D;
            p.out_i = I;
            p.id_base = id_base;
            e = launch_scan(p, nq, metric == B200R_METRIC_L2, bf16, nd->num_sms, nd->smem_optin, st, nullptr);
        }
    }
    if (prev >= 0 && prev != device) cudaSetDevice(prev);
    if (e != cudaSuccess) {
        set_last_error(cudaGetErrorString(e));
        return B200R_E_CUDA;
    }
    return B200R_OK;
}
