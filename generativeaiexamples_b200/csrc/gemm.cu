// K3: large-batch search = tcgen05/TMEM inner-product GEMM with a fused top-K' filter,
// followed by an exact fp32 re-score of the survivors ("gemm" path).
//
// Replaces what faiss 1.7.4 runs for nq >= 20: IndexFlat::search -> knn_* ->
// exhaustive_inner_product_blas / exhaustive_L2sqr_blas (faiss/utils/distances.cpp): sgemm
// over 4096 x 1024 blocks, L2 through ||q||^2 + ||x||^2 - 2 q.x, heap top-k.  In the
// reference it is reached by batched callers of index.search (the only in-tree one is
// community/event-driven-rag-cve-analysis/cyber_dev_day/faiss_vdb_service.py:147-174).
//
// Why two kernels.  The tensor cores have no fp32 inputs: kind::tf32 keeps 10 mantissa
// bits of each operand, kind::f16 (bf16) keeps 8.  The north star wants fp32 results
// (scores within 1e-5), so the GEMM is used as a FILTER with a proven error bound, and the
// few survivors are re-scored exactly on the CUDA cores:
//   1. gemm_filter_kernel: approx scores s~ for every (query, row); per query keep the
//      K' = k + slack best approx scores (never written to HBM beyond those lists);
//   2. rescore_kernel: exact fp32 score of each survivor (same arithmetic and summation
//      order as the scan path), exact top-k among them, and the check
//            exact_kth  >  approx_K'th + eps,      eps = c * ||q|| * max_row ||x||
//      (|s~ - s| <= eps for every row, by Cauchy-Schwarz on the per-product rounding
//      error).  If it holds, no row outside the survivors can belong to the exact top-k.
//      A query that fails the check is flagged and the host re-runs it on the scan path.
//
// gemm_filter_kernel layout (one CTA per SM, persistent over 256-row corpus tiles):
//   D[128 queries x 256 rows] (+ a second 128-query M tile when nq > 128) in TMEM;
//   A = queries, B = corpus rows, both K-major, 128-byte swizzled K blocks staged by TMA
//   (cp.async.bulk.tensor.2d, SASS UTMALDG) into a 3/4-stage mbarrier ring;
//   warp 0 = TMA producer, warp 1 = MMA issuer (one lane, tcgen05.mma cta_group::1),
//   warp 2 = TMEM allocator, warps 4.. = epilogue: tcgen05.ld 32 columns at a time, each
//   THREAD owns one query (TMEM lane), keeps its running K'-th best in a register and its
//   sorted survivor list in local memory (touched only on the rare insert).
#include "index.h"

#include <cuda.h>
#include <cstdlib>

namespace b200r {

// ---------------------------------------------------------------------------
// tcgen05 / TMA PTX wrappers
// ---------------------------------------------------------------------------
__device__ __forceinline__ void tma_load_2d(void* dst_smem, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
            smem_u32(dst_smem)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}

__device__ __forceinline__ void tma_load_2d_hint(void* dst_smem, const CUtensorMap* map, int c0, int c1, uint64_t* bar,
                                                 uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, "
        "%4}], [%2], %5;" ::"r"(smem_u32(dst_smem)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
        : "memory");
}

// contiguous global -> L2 prefetch (no shared-memory destination)
__device__ __forceinline__ void bulk_prefetch_l2(const void* gmem, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gmem), "r"(bytes) : "memory");
}

// multicast: the tile lands at the same CTA-relative offset in every CTA of `mask` and completes
// bytes on the mbarrier at the same offset in each of them
__device__ __forceinline__ void tma_load_2d_mc(void* dst_smem, const CUtensorMap* map, int c0, int c1, uint64_t* bar,
                                               uint16_t mask) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, "
        "{%3, %4}], [%2], %5;" ::"r"(smem_u32(dst_smem)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask)
        : "memory");
}

// 2-CTA (cta_group::2) TMA load: the bytes land in THIS CTA's shared memory, the transaction is
// counted on the LEADER CTA's mbarrier (bit 24 of a shared address selects the odd CTA of a pair)
__device__ __forceinline__ void tma_load_2d_pair(void* dst_smem, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, "
        "%4}], [%2];" ::"r"(smem_u32(dst_smem)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1)
        : "memory");
}

__device__ __forceinline__ void tma_load_2d_pair_hint(void* dst_smem, const CUtensorMap* map, int c0, int c1,
                                                      uint64_t* bar, uint64_t policy) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
        "[%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(dst_smem)),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar) & 0xFEFFFFFFu), "r"(c0), "r"(c1), "l"(policy)
        : "memory");
}

// arrive on the mbarrier at this offset in CTA `rank` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t rank) {
    asm volatile(
        "{\n\t.reg .b32 rem;\n\tmapa.shared::cluster.u32 rem, %0, %1;\n\t"
        "mbarrier.arrive.shared::cluster.b64 _, [rem];\n\t}" ::"r"(smem_u32(bar)),
        "r"(rank)
        : "memory");
}

__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__device__ __forceinline__ uint32_t cluster_cta_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}

__device__ __forceinline__ void prefetch_tmap(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}

__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

__device__ __forceinline__ void tmem_alloc_pair(uint32_t* dst_smem, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// cta_group::2 commit: arrives (once the pair's MMAs are done) on the barrier at this offset in
// every CTA of `mask`
__device__ __forceinline__ void tc_commit_pair(uint64_t* bar, uint16_t mask) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
            smem_u32(bar)),
        "h"(mask)
        : "memory");
}

// one M=256 MMA over a CTA pair: each CTA's tensor core computes its 128 rows of D from its own A
// half and BOTH halves of B (N/2 rows in each CTA's shared memory)
template <bool BF16>
__device__ __forceinline__ void umma_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
    if (BF16) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
            "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
            : "memory");
    } else {
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
            "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
            : "memory");
    }
}

// mbarrier arrives once all tcgen05.mma issued so far by this thread have completed
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}

// same, arriving on the barrier at this offset in every CTA of `mask`
__device__ __forceinline__ void tc_commit_mc(uint64_t* bar, uint16_t mask) {
    asm volatile(
        "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
            smem_u32(bar)),
        "h"(mask)
        : "memory");
}

template <bool BF16>
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                     uint32_t accumulate) {
    if (BF16) {
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
            "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
            : "memory");
    } else {
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
            "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
            : "memory");
    }
}

// 32 lanes x 32 consecutive columns: thread t of the warp gets row (lane) t
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

#ifndef B200R_GEMM_KBYTES
#define B200R_GEMM_KBYTES 128  // bytes of K per pipeline stage = width of the swizzle atom (128 or 64)
#endif

// K-major swizzled operand tile: rows of KBYTES bytes, 8-row groups 8*KBYTES apart.
// start>>4 | LBO=1 (unused for swizzled K-major) | SBO=(8*KBYTES)>>4 | version=1 | layout
// (SWIZZLE_128B = 2, SWIZZLE_64B = 4)
__device__ __forceinline__ uint64_t make_kmajor_sw128_desc(const void* smem_tile) {
    const uint64_t addr = (uint64_t)((smem_u32(smem_tile) & 0x3FFFFu) >> 4);
    constexpr uint64_t sbo = (8ull * B200R_GEMM_KBYTES) >> 4;
    constexpr uint64_t layout = B200R_GEMM_KBYTES == 128 ? 2ull : 4ull;
    return addr | (1ull << 16) | (sbo << 32) | (1ull << 46) | (layout << 61);
}

// ---------------------------------------------------------------------------
constexpr int kGemmTileN = 256;        // corpus rows per tile (UMMA N)
constexpr int kGemmTileM = 128;        // queries per M tile (UMMA M)
constexpr int kGemmKBytes = B200R_GEMM_KBYTES;  // one swizzle atom of K per stage
constexpr int kGemmBBytes = kGemmTileN * kGemmKBytes;   // 32 KB
constexpr int kGemmABytes = kGemmTileM * kGemmKBytes;   // 16 KB per M tile
constexpr int kGemmBuckets = 512;      // score histogram resolution for the grid-wide threshold

struct GemmParams {
    int64_t n_rows;
    int nq;            // active queries (<= 128 * MT)
    int d_pad;         // elements per row
    int kprime;        // survivors kept per query (<= KP)
    int num_kb;        // K blocks of 128 bytes
    int64_t n_tiles;
    const float* norms;   // [n_rows] squared norms (L2 only)
    Cand* cand;           // [grid][128*MT][KP]
    int* gthr;            // [128*MT] grid-wide thresholds, ordered-int encoded, pre-set to the worst value
    const void* corpus;   // base of the row-major corpus (for the L2 prefetch stream)
    int prefetch_tiles;   // tiles of look-ahead for that stream (0 = off)
    unsigned* hist;       // [128*MT][kGemmBuckets] rows inserted so far per goodness bucket (zeroed per launch)
    const float* qscale;  // [128*MT][2] bucket map per query: goodness lo, buckets per unit
    int warmup;           // 1: first tile of each CTA only feeds the histogram and is replayed at the end
    int* list_counts;     // [grid][128*MT] valid entries in each survivor list
    int* overflow;        // [128*MT] set when some CTA's slice for the query overflowed
    int debug_no_select;  // timing experiments only: 1 = skip the selection, 2 = also skip the TMEM loads
    int evict_first;
};

// monotone float <-> int encoding, so that atomicMax / atomicMin order floats correctly
__device__ __forceinline__ int enc_ordered(float f) {
    const int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__device__ __forceinline__ float dec_ordered(int o) { return __int_as_float(o >= 0 ? o : o ^ 0x7FFFFFFF); }

// v[j] for a run-time j without sending v to local memory: a 5-level select tree (31 SELs)
__device__ __forceinline__ float pick32(const float (&v)[32], int j) {
    float a[16], b[8], c[4];
#pragma unroll
    for (int i = 0; i < 16; i++) a[i] = (j & 1) ? v[2 * i + 1] : v[2 * i];
#pragma unroll
    for (int i = 0; i < 8; i++) b[i] = (j & 2) ? a[2 * i + 1] : a[2 * i];
#pragma unroll
    for (int i = 0; i < 4; i++) c[i] = (j & 4) ? b[2 * i + 1] : b[2 * i];
    const float d0 = (j & 8) ? c[1] : c[0];
    const float d1 = (j & 8) ? c[3] : c[2];
    return (j & 16) ? d1 : d0;
}

template <bool SMALL>
__device__ __forceinline__ float better_of(float a, float b) {
    return SMALL ? fminf(a, b) : fmaxf(a, b);
}

// CS = cluster size.  CS > 1: the CTAs of a cluster walk their (different) corpus tiles in lock
// step, and every K block of the query operand -- identical for all of them -- is fetched ONCE per
// cluster: CTA r loads rows [r*128/CS, (r+1)*128/CS) of each M tile and TMA-multicasts them into
// all CS shared memories.  That removes (CS-1)/CS of the query re-reads from the L2->SM traffic,
// which is what bounds the 256-query variants (operands at MMA rate need ~14 TB/s of it).
//
// PAIR = 2-CTA MMA (cta_group::2, cluster of 2, MT must be 1, CS must be 1): 256 queries x 256 rows
// per tile over a CTA PAIR.  CTA r stages queries [128r, 128r+128) and corpus rows [128r, 128r+128)
// of the tile (32 KB per K block instead of 64 KB -> 6 stages), the leader CTA issues one M=256
// tcgen05.mma.cta_group::2 per K step, and each CTA's TMEM receives its 128 queries x 256 rows --
// 256 columns per accumulator stage, so the accumulator is DOUBLE-buffered and the ~4k-clock TMEM
// drain of a tile overlaps the next tile's MMAs (with two M tiles in one CTA all 512 columns are one
// stage and the tensor pipe waits for every drain).
template <bool BF16, int MT, int KP, bool L2, int CS, bool PAIR>
__global__ void __launch_bounds__(128 + 128 * MT, 1)
gemm_filter_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_x,
                   const GemmParams p) {
    static_assert(!PAIR || (MT == 1 && CS == 1), "the 2-CTA mode has one M tile per CTA and no query multicast");
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    constexpr int B_ROWS = PAIR ? kGemmTileN / 2 : kGemmTileN;  // corpus rows this CTA stages per tile
    constexpr int B_BYTES = B_ROWS * kGemmKBytes;
    constexpr int STAGES = (PAIR ? 6 : (MT == 1 ? 4 : 3)) * (128 / kGemmKBytes);
    constexpr int STAGE_BYTES = B_BYTES + MT * kGemmABytes;
    constexpr int ACC_STAGES = MT == 1 ? 2 : 1;      // 512 TMEM columns in total
    constexpr int ACC_COLS = MT * kGemmTileN;        // columns per accumulator stage
    constexpr int UMMA_K_BYTES = 32;                 // 8 tf32 / 16 bf16
    constexpr int N_EPI_WARPS = 4 * MT;

    // carve-up (1024-byte aligned stages for the 128B swizzle)
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* stages = smem;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tfull_bar = empty_bar + STAGES;
    uint64_t* tempty_bar = tfull_bar + ACC_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + ACC_STAGES);
    volatile int* s_done = reinterpret_cast<volatile int*>(tmem_slot + 1);  // epilogue warps that have finished
    float* s_norms = reinterpret_cast<float*>(tmem_slot + 4);  // [N_EPI_WARPS][256] (L2 only)

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], CS);  // one MMA commit from every CTA whose multicast fills this stage
        }
        for (int a = 0; a < ACC_STAGES; a++) {
            mbar_init(&tfull_bar[a], 1);
            // PAIR: the leader's barrier collects the epilogue warps of both CTAs
            mbar_init(&tempty_bar[a], PAIR ? 2 * N_EPI_WARPS : N_EPI_WARPS);
        }
        *s_done = 0;
        fence_mbar_init();
        prefetch_tmap(&map_q);
        prefetch_tmap(&map_x);
    }
    if (PAIR) cluster_sync_all();  // both CTAs of the pair are resident before the paired TMEM allocation
    if (warp == 2) {
        if (PAIR)
            tmem_alloc_pair(tmem_slot, 512);
        else
            tmem_alloc(tmem_slot, 512);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t cta_rank = (CS > 1 || PAIR) ? cluster_cta_rank() : 0u;
    constexpr uint16_t kClusterMask = (uint16_t)((1u << CS) - 1u);
    if (CS > 1 || PAIR) cluster_sync_all();  // every CTA's barriers exist before any remote arrive / multicast

    const int64_t n_tiles = p.n_tiles;
    // tiles of this CTA: blockIdx.x, +grid, ...; with warm-up the first one is visited once more at the
    // end.  In a cluster every CTA runs the same number of iterations (tiles past the end are all
    // out-of-bounds rows: TMA zero-fills them and the epilogue masks them).
    // work unit: a CTA, or a CTA pair in 2-CTA mode (tiles are dealt to pairs)
    const int64_t unit = PAIR ? (int64_t)(blockIdx.x >> 1) : (int64_t)blockIdx.x;
    const int64_t n_units = PAIR ? (int64_t)(gridDim.x >> 1) : (int64_t)gridDim.x;
    const int64_t n_my = (CS > 1 || PAIR) ? (n_tiles + n_units - 1) / n_units : (n_tiles - unit + n_units - 1) / n_units;
    const int64_t n_iter = n_my + (p.warmup ? 1 : 0);
    auto tile_of = [&](int64_t i) -> int64_t { return unit + (i < n_my ? i : 0) * n_units; };

    if (warp == 0) {
        // ---------------- TMA producer ----------------
        if (lane == 0) {
            const uint64_t pol = policy_evict_first();
            int s = 0;
            uint32_t ph = 0;
            const int kstep = kGemmKBytes / (BF16 ? 2 : 4);  // elements per K block
            // The K-blocked tile loads touch 256 rows x 128 B each: scattered for DRAM.  Stream the
            // tile's rows (one contiguous block) into L2 a few tiles ahead with bulk prefetches, so
            // that DRAM sees long sequential reads and the TMA tile loads hit L2.
            const int row_bytes = p.d_pad * (BF16 ? 2 : 4);
            const uint8_t* xg = reinterpret_cast<const uint8_t*>(p.corpus);
            auto prefetch_tile = [&](int64_t tt) {
                if (tt >= n_tiles) return;
                const int64_t r0 = tt * kGemmTileN;
                const int64_t rows = (p.n_rows - r0) < kGemmTileN ? (p.n_rows - r0) : kGemmTileN;
                const int64_t bytes = rows * row_bytes;
                const uint8_t* base = xg + r0 * row_bytes;
                for (int64_t o = 0; o < bytes; o += 32768) {
                    const int64_t len = (bytes - o) < 32768 ? (bytes - o) : 32768;
                    bulk_prefetch_l2(base + o, (uint32_t)len);
                }
            };
            for (int pf = 0; pf < p.prefetch_tiles; pf++) prefetch_tile(unit + (int64_t)pf * n_units);
            for (int64_t i = 0; i < n_iter; i++) {
                const int64_t t = tile_of(i);
                const int row0 = (int)(t * kGemmTileN);
                if (p.prefetch_tiles > 0) prefetch_tile(t + (int64_t)p.prefetch_tiles * n_units);
                for (int kb = 0; kb < p.num_kb; kb++) {
                    mbar_wait(&empty_bar[s], ph ^ 1);
                    uint8_t* st = stages + (size_t)s * STAGE_BYTES;
                    if (PAIR) {
                        // both CTAs' loads are counted on the leader's barrier
                        if (cta_rank == 0) mbar_arrive_expect_tx(&full_bar[s], 2 * STAGE_BYTES);
                        const int xr = row0 + (int)cta_rank * B_ROWS;
                        if (p.evict_first)
                            tma_load_2d_pair_hint(st, &map_x, kb * kstep, xr, &full_bar[s], pol);
                        else
                            tma_load_2d_pair(st, &map_x, kb * kstep, xr, &full_bar[s]);
                        tma_load_2d_pair(st + B_BYTES, &map_q, kb * kstep, (int)cta_rank * kGemmTileM, &full_bar[s]);
                        if (++s == STAGES) {
                            s = 0;
                            ph ^= 1;
                        }
                        continue;
                    }
                    mbar_arrive_expect_tx(&full_bar[s], STAGE_BYTES);
                    if (p.evict_first)
                        tma_load_2d_hint(st, &map_x, kb * kstep, row0, &full_bar[s], pol);
                    else
                        tma_load_2d(st, &map_x, kb * kstep, row0, &full_bar[s]);
#pragma unroll
                    for (int mt = 0; mt < MT; mt++) {
                        if (CS == 1)
                            tma_load_2d(st + B_BYTES + mt * kGemmABytes, &map_q, kb * kstep, mt * kGemmTileM,
                                        &full_bar[s]);
                        else
                            tma_load_2d_mc(st + B_BYTES + mt * kGemmABytes + cta_rank * (kGemmABytes / CS), &map_q,
                                           kb * kstep, mt * kGemmTileM + (int)cta_rank * (kGemmTileM / CS), &full_bar[s],
                                           kClusterMask);
                    }
                    if (++s == STAGES) {
                        s = 0;
                        ph ^= 1;
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ---------------- MMA issuer (in 2-CTA mode: the leader CTA's, for the pair) ----------------
        if (lane == 0 && (!PAIR || cta_rank == 0)) {
            // idesc: D=f32 (1<<4) | A,B format (tf32=2 / bf16=1) | K-major both | N>>3 | M>>4
            const uint32_t fmt = BF16 ? 1u : 2u;
            const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(kGemmTileN >> 3) << 17) |
                                   ((uint32_t)((PAIR ? 2 * kGemmTileM : kGemmTileM) >> 4) << 24);
            int s = 0;
            uint32_t ph = 0;
            for (int64_t it = 0; it < n_iter; ++it) {
                const int as = (int)(it % ACC_STAGES);
                const uint32_t aph = (uint32_t)((it / ACC_STAGES) & 1);
                mbar_wait(&tempty_bar[as], aph ^ 1);
                tc_fence_after();
                for (int kb = 0; kb < p.num_kb; kb++) {
                    mbar_wait(&full_bar[s], ph);
                    tc_fence_after();
                    const uint8_t* st = stages + (size_t)s * STAGE_BYTES;
                    const uint64_t db = make_kmajor_sw128_desc(st);
#pragma unroll
                    for (int mt = 0; mt < MT; mt++) {
                        const uint64_t da = make_kmajor_sw128_desc(st + B_BYTES + mt * kGemmABytes);
                        const uint32_t td = tmem_base + (uint32_t)(as * ACC_COLS + mt * kGemmTileN);
#pragma unroll
                        for (int k4 = 0; k4 < kGemmKBytes / UMMA_K_BYTES; k4++) {
                            const uint64_t adv = (uint64_t)((k4 * UMMA_K_BYTES) >> 4);
                            if (PAIR)
                                umma_pair<BF16>(td, da + adv, db + adv, idesc, (uint32_t)((kb | k4) != 0));
                            else
                                umma<BF16>(td, da + adv, db + adv, idesc, (uint32_t)((kb | k4) != 0));
                        }
                    }
                    // frees the smem stage when these MMAs are done (in every CTA that fills / reads it)
                    if (PAIR)
                        tc_commit_pair(&empty_bar[s], 0x3);
                    else if (CS == 1)
                        tc_commit(&empty_bar[s]);
                    else
                        tc_commit_mc(&empty_bar[s], kClusterMask);
                    if (kb == p.num_kb - 1) {
                        if (PAIR)
                            tc_commit_pair(&tfull_bar[as], 0x3);  // both CTAs drain their half of D
                        else
                            tc_commit(&tfull_bar[as]);
                    }
                    if (++s == STAGES) {
                        s = 0;
                        ph ^= 1;
                    }
                }
            }
        }
        __syncwarp();
    } else if (warp == 3) {
        // ---------------- grid-wide threshold refresher ----------------
        // hist[q][b] counts the rows that entered some CTA's list for query q with goodness in
        // bucket b (goodness = score for IP, -value for L2).  The lower edge of the highest bucket
        // whose suffix count reaches K' is a valid lower bound of the final K'-th best, and a far
        // tighter one than any single CTA's own K'-th best.  This warp keeps publishing it.
        const float sgn = L2 ? -1.f : 1.f;
        int round = 0;
        while (*s_done < N_EPI_WARPS) {
            // round r: the grid's CTAs cover queries {c + r*grid} -- together they sweep all queries
            // every ceil(nq/grid) rounds (a few microseconds), so no query's threshold goes stale
            const int q = (int)(((int64_t)blockIdx.x + (int64_t)round * gridDim.x) % p.nq);
            ++round;
            const uint4* h4 = reinterpret_cast<const uint4*>(p.hist + (size_t)q * kGemmBuckets) + lane * 4;
            unsigned c[16];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const uint4 t4 = __ldcg(h4 + u);
                c[4 * u] = t4.x;
                c[4 * u + 1] = t4.y;
                c[4 * u + 2] = t4.z;
                c[4 * u + 3] = t4.w;
            }
            unsigned tot = 0;
#pragma unroll
            for (int u = 0; u < 16; u++) tot += c[u];
            // suffix sums across lanes: above = rows in the buckets of all higher lanes
            unsigned incl = tot;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned up = __shfl_down_sync(0xffffffffu, incl, o);
                if (lane + o < 32) incl += up;
            }
            const unsigned above = incl - tot;
            const unsigned need = (unsigned)p.kprime;
            const unsigned reach = __ballot_sync(0xffffffffu, incl >= need);
            if (reach != 0) {
                const int lstar = 31 - __clz(reach);  // highest lane whose suffix count reaches K'
                if (lane == lstar) {
                    unsigned run = above;
                    int bstar = lane * 16;
#pragma unroll
                    for (int u = 15; u >= 0; u--) {
                        run += c[u];
                        if (run >= need) {
                            bstar = lane * 16 + u;
                            break;
                        }
                    }
                    // one bucket of safety margin against rounding in the edge computation
                    const float lo = p.qscale[2 * q], per_unit = p.qscale[2 * q + 1];
                    if (bstar >= 1 && per_unit > 0.f) {
                        const float edge = lo + (float)(bstar - 1) / per_unit;
                        const float tval = sgn * edge;
                        const float cur = dec_ordered(*reinterpret_cast<volatile int*>(p.gthr + q));
                        if (better_val<L2>(tval, cur)) {
                            if (L2)
                                atomicMin(p.gthr + q, enc_ordered(tval));
                            else
                                atomicMax(p.gthr + q, enc_ordered(tval));
                        }
                    }
                }
            }
            __syncwarp();
        }
    } else if (warp >= 4) {
        // ---------------- epilogue: one query per thread ----------------
        const int ew = warp - 4;
        const int mt = ew >> 2;
        const int quad = warp & 3;  // TMEM lane quadrant this warp may read
        // query of this thread inside the launch's block of queries; Q_STRIDE = queries per work unit
        constexpr int Q_STRIDE = PAIR ? 2 * kGemmTileM : kGemmTileM * MT;
        const int qidx = (PAIR ? (int)cta_rank * kGemmTileM : mt * kGemmTileM) + quad * 32 + lane;
        const bool active = qidx < p.nq;
        // KP == 32 (k <= 10): sorted list held entirely in registers -- no memory traffic on insert
        // (the L1 left over beside ~200 KB of pipeline smem is too small for local-memory lists).
        // KP > 32: the launch runs in warm-up mode, so a tight grid-wide threshold exists before
        // the first row is kept and only ~K'/grid survivors reach each thread: they are simply
        // APPENDED to the thread's slice of the global candidate buffer (one fire-and-forget
        // store, nothing to read back).  A slice that overflows (adversarial row order) flags
        // its query for the scan path.
        constexpr int NREG = KP == 32 ? 32 : 1;
        float lv[NREG];
        uint32_t li[NREG];
        const int kp = p.kprime;
        int count = 0;
        Cand* my_list = p.cand + ((size_t)unit * Q_STRIDE + qidx) * KP;
        if (KP == 32) {
#pragma unroll
            for (int j = 0; j < KP; j++) {
                lv[j] = worst_val<L2>();
                li[j] = kEmptyId;
            }
        }
        float lthr = worst_val<L2>();   // what a score must beat to enter THIS thread's list
        float thr = worst_val<L2>();    // effective filter: better of lthr and the grid-wide threshold
        int* gthr = p.gthr + qidx;      // grid-wide K'-th best so far for this query (ordered-int encoded)
        unsigned* my_hist = p.hist + (size_t)qidx * kGemmBuckets;
        const float h_lo = active ? p.qscale[2 * qidx] : 0.f;
        const float h_per_unit = active ? p.qscale[2 * qidx + 1] : 0.f;
        float* my_norms = s_norms + ew * kGemmTileN;
        for (int64_t it = 0; it < n_iter; ++it) {
            const int64_t t = tile_of(it);
            const bool hist_only = p.warmup && it == 0;     // feed the histogram, keep nothing
            const bool replay = p.warmup && it == n_my;     // second visit: rows are already counted
            const int as = (int)(it % ACC_STAGES);
            const uint32_t aph = (uint32_t)((it / ACC_STAGES) & 1);
            const int64_t row0 = t * kGemmTileN;
            // Any CTA's full-list worst is a lower bound of the final K'-th best, so filtering with
            // the best such bound published by any CTA never drops a row of the global top-K'.
            float g = worst_val<L2>();
            int g_pend = active ? *reinterpret_cast<volatile int*>(gthr) : 0;  // consumed one step later
            if (L2) {
                // stage ||x||^2 of the tile's rows; rows past the end never win
                __syncwarp();
#pragma unroll
                for (int u = 0; u < kGemmTileN / 32; u++) {
                    const int64_t r = row0 + u * 32 + lane;
                    my_norms[u * 32 + lane] = r < p.n_rows ? p.norms[r] : FLT_MAX;
                }
                __syncwarp();
            }
            mbar_wait(&tfull_bar[as], aph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + (uint32_t)(as * ACC_COLS + mt * kGemmTileN);
            bool improved = false;
#pragma unroll 1
            for (int c = 0; c < kGemmTileN / 32; c++) {
                uint32_t r[32];
                if (p.debug_no_select == 2) continue;
                // the grid-wide threshold is re-read every chunk, but the value is only used one chunk
                // later, so its L2 round trip (~700 clk) never stalls the accumulator drain
                const int g_enc = g_pend;
                if (active) g_pend = *reinterpret_cast<volatile int*>(gthr);
                tmem_ld32(taddr + c * 32, r);
                if (active) {
                    g = better_of<L2>(g, dec_ordered(g_enc));
                    thr = better_of<L2>(lthr, g);
                }
                if (active && p.debug_no_select == 0) {
                    float v[32];
                    const int64_t rbase = row0 + c * 32;
                    if (L2) {
                        // ||x||^2 - 2 q.x   (+ ||q||^2 is added at re-score time; order is unaffected)
#pragma unroll
                        for (int j = 0; j < 32; j++) v[j] = fmaf(-2.f, __uint_as_float(r[j]), my_norms[c * 32 + j]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; j++) v[j] = __uint_as_float(r[j]);
                    }
                    if (rbase + 32 > p.n_rows) {  // last tile only: rows past the end never win
#pragma unroll
                        for (int j = 0; j < 32; j++)
                            if (rbase + j >= p.n_rows) v[j] = worst_val<L2>();
                    }
                    float best = v[0];
#pragma unroll
                    for (int j = 1; j < 32; j++) best = better_of<L2>(best, v[j]);
                    if (hist_only) {
#pragma unroll
                        for (int j = 0; j < 32; j++)
                            if (rbase + j < p.n_rows) {
                                const float good = L2 ? -v[j] : v[j];
                                int b = (int)((good - h_lo) * h_per_unit);
                                b = b < 0 ? 0 : (b > kGemmBuckets - 1 ? kGemmBuckets - 1 : b);
                                atomicAdd(my_hist + b, 1u);
                            }
                    } else if (better_val<L2>(best, thr)) {
                        // A chunk with a survivor is rare.  The survivors' columns go into a bit mask and
                        // ONE copy of the insert code walks it: unrolling the insert per column made
                        // ~60 KB of code whose instruction-cache misses were the top stall (r01 notes).
                        uint32_t hit = 0;
#pragma unroll
                        for (int j = 0; j < 32; j++) hit |= better_val<L2>(v[j], thr) ? (1u << j) : 0u;
                        while (hit) {
                            const int j = __ffs(hit) - 1;
                            hit &= hit - 1;
                            const float vj = pick32(v, j);
                            if (!better_val<L2>(vj, thr)) continue;  // the threshold moved meanwhile
                            if (KP == 32) {
                                // compare-exchange chain over the 32 register slots (static indices)
                                float cv = vj;
                                uint32_t ci = (uint32_t)(rbase + j);
#pragma unroll
                                for (int e = 0; e < 32; e++) {
                                    const bool sw = better_val<L2>(cv, lv[e]);
                                    const float tv = lv[e];
                                    const uint32_t ti = li[e];
                                    lv[e] = sw ? cv : tv;
                                    li[e] = sw ? ci : ti;
                                    cv = sw ? tv : cv;
                                    ci = sw ? ti : ci;
                                }
                                lthr = lv[31];
                                count = 32;  // sentinel slots make the list "full" from the start
                            } else {
                                if (count == KP) {
                                    // slice full: entries admitted under an older, weaker threshold that
                                    // the current one rules out can go (it is a lower bound of the final
                                    // K'-th best, so they cannot be among the K' best)
                                    g = better_of<L2>(g, dec_ordered(*reinterpret_cast<volatile int*>(gthr)));
                                    int w = 0;
                                    for (int e = 0; e < KP; e++) {
                                        const Cand c2 = my_list[e];
                                        if (better_val<L2>(c2.v, g)) my_list[w++] = c2;
                                    }
                                    count = w;
                                }
                                if (count < KP && better_val<L2>(vj, g)) {
                                    Cand cnd;
                                    cnd.v = vj;
                                    cnd.id = (uint32_t)(rbase + j);
                                    my_list[count++] = cnd;
                                } else if (count >= KP) {
                                    count = KP + 1;  // still full after compaction: overflow
                                }
                            }
                            thr = better_of<L2>(lthr, g);
                            improved = true;
                            if (!replay) {
                                const float good = L2 ? -vj : vj;
                                int b = (int)((good - h_lo) * h_per_unit);
                                b = b < 0 ? 0 : (b > kGemmBuckets - 1 ? kGemmBuckets - 1 : b);
                                atomicAdd(my_hist + b, 1u);
                            }
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (PAIR && cta_rank != 0)
                    mbar_arrive_remote(&tempty_bar[as], 0);  // the leader's MMA issuer owns the accumulator hand-off
                else
                    mbar_arrive(&tempty_bar[as]);
            }
            if (hist_only && active) {
                // give the refresher warps (any CTA's) a moment to turn the first counts into a
                // threshold; bounded wait, the result is correct with or without it
                const long long t_end = clock64() + 40000;
                while (dec_ordered(*reinterpret_cast<volatile int*>(gthr)) == worst_val<L2>() && clock64() < t_end) {
                }
            }
            // publish: only a FULL list's worst is a valid bound
            if (KP == 32 && active && improved && count >= kp && lthr != worst_val<L2>() && better_val<L2>(lthr, g)) {
                if (L2)
                    atomicMin(gthr, enc_ordered(lthr));
                else
                    atomicMax(gthr, enc_ordered(lthr));
            }
        }
        __syncwarp();
        if (lane == 0) atomicAdd(const_cast<int*>(s_done), 1);
        // survivors -> global, [cta][query][KP] (unsorted; unused slots empty)
        if (active) {
            Cand* dst = p.cand + ((size_t)unit * Q_STRIDE + qidx) * KP;
            if (KP == 32) {
                int nv = 0;
#pragma unroll
                for (int j = 0; j < 32; j++) {
                    Cand cnd;
                    cnd.v = lv[j];
                    cnd.id = li[j];
                    if (li[j] != kEmptyId) {  // sorted: the valid entries come first
                        dst[j] = cnd;
                        nv++;
                    }
                }
                p.list_counts[(size_t)unit * Q_STRIDE + qidx] = nv;
            } else {
                p.list_counts[(size_t)unit * Q_STRIDE + qidx] = count < KP ? count : KP;
                if (count > KP) p.overflow[qidx] = 1;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (CS > 1 || PAIR) cluster_sync_all();  // no CTA leaves while a peer may still arrive on its barriers
    if (warp == 2) {
        tc_fence_after();
        if (PAIR)
            tmem_dealloc_pair(tmem_base, 512);
        else
            tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------
// rescore_kernel: one CTA per query.
// ---------------------------------------------------------------------------
constexpr int kRescoreWarps = 8;        // per CTA when many queries are re-scored at once
constexpr int kRescoreWarpsFew = 32;    // a handful of queries: one CTA per query is all the parallelism there is

struct RescoreParams {
    const void* corpus;
    int64_t n_rows;
    int d_pad, row_bytes;
    const float* q;        // [nq][d_pad] fp32 (prepared)
    const Cand* cand;      // [grid][q_stride][KP] approx survivors
    const int* list_counts; // [grid][q_stride]
    const int* overflow;    // [q_stride]
    const float* qscale;    // [q_stride][2] the filter's bucket map
    int grid, q_stride, kp_stride, kprime, k;
    const unsigned* max_norm_bits;  // max_i ||x_i||^2 as float bits
    float err_coeff;       // c in eps = c * ||q|| * max||x||
    float* out_d;
    int64_t* out_i;
    int64_t id_base;
    int* fail_flags;       // [nq]
    int* fail_count;
};

// block-wide bitonic sort of n (a power of two) candidates in shared memory, best first;
// ties on the score go to the lower id, as everywhere else
template <bool L2>
__device__ void block_sort_best_first(Cand* a, int n) {
    for (int k = 2; k <= n; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < n; i += blockDim.x) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const Cand x = a[i], y = a[ixj];
                    const bool first_half = (i & k) == 0;  // this sub-sequence is sorted best-first
                    const bool swap = first_half ? better<L2>(y.v, y.id, x.v, x.id) : better<L2>(x.v, x.id, y.v, y.id);
                    if (swap) {
                        a[i] = y;
                        a[ixj] = x;
                    }
                }
            }
            __syncthreads();
        }
    }
}

// Sort by counting: every candidate's position is the number of candidates that beat it (the
// (score, id) order is total because ids are unique).  n^2 / blockDim comparisons on broadcast
// shared-memory reads and ONE barrier, against ~log^2(n)/2 barriers of the bitonic network: for
// the few hundred survivors of a query this is several times faster.  src and dst must not overlap.
template <bool L2>
__device__ void block_rank_sort(const Cand* src, Cand* dst, int n) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const Cand me = src[i];
        int r = 0;
        for (int j = 0; j < n; j++) {
            const Cand o = src[j];
            r += better<L2>(o.v, o.id, me.v, me.id) ? 1 : 0;
        }
        dst[r] = me;
    }
    __syncthreads();
}
constexpr int kRankSortMax = 1024;  // beyond this the bitonic network wins

__device__ __forceinline__ int next_pow2(int v, int lo) {
    int p2 = lo;
    while (p2 < v) p2 <<= 1;
    return p2;
}

constexpr int kRescoreCap = 4096;    // selected candidates (K' + one boundary bucket) that fit the sort buffer
constexpr int kRescoreMaxKp = 2048;  // >= the largest K' (k = 1024 -> 1536)

template <bool L2, bool BF16, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) rescore_kernel(const RescoreParams p) {
    extern __shared__ __align__(16) uint8_t rs_smem[];
    float* qs = reinterpret_cast<float*>(rs_smem);              // [d_pad]
    Cand* buf = reinterpret_cast<Cand*>(qs + p.d_pad);          // [kRescoreCap] approx survivors, then sorted
    Cand* ex = buf + kRescoreCap;                               // [kRescoreMaxKp] exact scores of the K' best
    __shared__ float s_red[WARPS];
    __shared__ float s_qnorm2;
    __shared__ int s_off[kGemmMaxGrid + 1];
    __shared__ unsigned s_hist[kGemmBuckets];
    __shared__ int s_nsel, s_bstar;

    const int qi = blockIdx.x;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* qg = p.q + (size_t)qi * p.d_pad;
    float part = 0.f;
    for (int i = threadIdx.x; i < p.d_pad; i += blockDim.x) {
        const float v = qg[i];
        qs[i] = v;
        part = fmaf(v, v, part);
    }
    part = warp_sum(part);
    if (lane == 0) s_red[warp] = part;
    // survivors per CTA list -> offsets
    for (int g = threadIdx.x; g < p.grid; g += blockDim.x) {
        int n_e = p.list_counts[(size_t)g * p.q_stride + qi];
        s_off[g + 1] = n_e < p.kp_stride ? n_e : p.kp_stride;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float s = 0.f;
        for (int w = 0; w < WARPS; w++) s += s_red[w];
        s_qnorm2 = s;
        int run = 0;
        s_off[0] = 0;
        for (int g = 0; g < p.grid; g++) {
            run += s_off[g + 1];
            s_off[g + 1] = run;
        }
    }
    __syncthreads();
    const int total = s_off[p.grid];
    // Select the K' best approx scores among the `total` survivors of all CTAs without sorting
    // them all: histogram over the filter's own bucket map -> boundary bucket b* -> collect the
    // candidates at or above b* (K' plus one bucket's worth) -> sort only those.
    const float h_lo = p.qscale[2 * qi], h_per_unit = p.qscale[2 * qi + 1];
    auto bucket_of = [&](float v) -> int {
        const float good = L2 ? -v : v;
        int bk = (int)((good - h_lo) * h_per_unit);
        return bk < 0 ? 0 : (bk > kGemmBuckets - 1 ? kGemmBuckets - 1 : bk);
    };
    for (int i = threadIdx.x; i < kGemmBuckets; i += blockDim.x) s_hist[i] = 0u;
    if (threadIdx.x == 0) s_nsel = 0;
    __syncthreads();
    int bstar = 0;
    if (total > p.kprime) {
        for (int g = warp; g < p.grid; g += WARPS) {
            const Cand* src = p.cand + ((size_t)g * p.q_stride + qi) * p.kp_stride;
            const int n_e = s_off[g + 1] - s_off[g];
            for (int e = lane; e < n_e; e += 32) atomicAdd(&s_hist[bucket_of(src[e].v)], 1u);
        }
        __syncthreads();
        if (warp == 0) {
            unsigned c16[16], tot = 0;
#pragma unroll
            for (int u = 0; u < 16; u++) {
                c16[u] = s_hist[lane * 16 + u];
                tot += c16[u];
            }
            unsigned incl = tot;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned up = __shfl_down_sync(0xffffffffu, incl, o);
                if (lane + o < 32) incl += up;
            }
            const unsigned need = (unsigned)p.kprime;
            const unsigned reach = __ballot_sync(0xffffffffu, incl >= need);
            const int lstar = 31 - __clz(reach);  // reach != 0 because total > K'
            if (lane == lstar) {
                unsigned run = incl - tot;
                int bs = lane * 16;
#pragma unroll
                for (int u = 15; u >= 0; u--) {
                    run += c16[u];
                    if (run >= need) {
                        bs = lane * 16 + u;
                        break;
                    }
                }
                s_bstar = bs;
            }
        }
        __syncthreads();
        bstar = s_bstar;
    }
    for (int g = warp; g < p.grid; g += WARPS) {
        const Cand* src = p.cand + ((size_t)g * p.q_stride + qi) * p.kp_stride;
        const int n_e = s_off[g + 1] - s_off[g];
        for (int e = lane; e < n_e; e += 32) {
            const Cand c = src[e];
            if (bucket_of(c.v) >= bstar) {
                const int pos = atomicAdd(&s_nsel, 1);
                if (pos < kRescoreCap) buf[pos] = c;
            }
        }
    }
    __syncthreads();
    const int n_sel = s_nsel;
    if (n_sel > kRescoreCap || p.overflow[qi]) {
        // one bucket holds thousands of candidates (massive ties): let the scan path answer
        if (threadIdx.x == 0) {
            p.fail_flags[qi] = p.overflow[qi] ? 3 : 2;  // cause codes: 1 check, 2 boundary bucket too full, 3 slice overflow
            atomicAdd(p.fail_count, 1);
        }
        return;
    }
    const bool rank_mode = n_sel <= kRankSortMax;
    const Cand* sorted = buf;  // approx survivors, best first
    if (rank_mode) {
        block_rank_sort<L2>(buf, buf + kRescoreCap / 2, n_sel);
        sorted = buf + kRescoreCap / 2;
    } else {
        const int P = next_pow2(n_sel, 32);
        for (int i = n_sel + threadIdx.x; i < P; i += blockDim.x) {
            Cand c;
            c.v = worst_val<L2>();
            c.id = kEmptyId;
            buf[i] = c;
        }
        __syncthreads();
        block_sort_best_first<L2>(buf, P);
    }
    const int n_valid = n_sel < p.kprime ? n_sel : p.kprime;   // the K' best approx scores: sorted[0..n_valid)

    // all warps: exact fp32 score of every survivor, same lane mapping / summation order as
    // the scan kernel (4 partial sums per lane, butterfly)
    const int nvec = BF16 ? (p.d_pad >> 3) : (p.d_pad >> 2);
    for (int c = warp; c < n_valid; c += WARPS) {
        const uint32_t id = sorted[c].id;
        const uint4* xr = reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(p.corpus) +
                                                         (size_t)id * p.row_bytes);
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        for (int i = lane; i < nvec; i += 32) {
            const uint4 xv = __ldg(xr + i);
            if (!BF16) {
                const float4 qv = reinterpret_cast<const float4*>(qs)[i];
                const float x0 = __uint_as_float(xv.x), x1 = __uint_as_float(xv.y), x2 = __uint_as_float(xv.z),
                            x3 = __uint_as_float(xv.w);
                if (L2) {
                    const float t0 = qv.x - x0, t1 = qv.y - x1, t2 = qv.z - x2, t3 = qv.w - x3;
                    a0 = fmaf(t0, t0, a0);
                    a1 = fmaf(t1, t1, a1);
                    a2 = fmaf(t2, t2, a2);
                    a3 = fmaf(t3, t3, a3);
                } else {
                    a0 = fmaf(qv.x, x0, a0);
                    a1 = fmaf(qv.y, x1, a1);
                    a2 = fmaf(qv.z, x2, a2);
                    a3 = fmaf(qv.w, x3, a3);
                }
            } else {
                const float4 qa = reinterpret_cast<const float4*>(qs)[2 * i];
                const float4 qb = reinterpret_cast<const float4*>(qs)[2 * i + 1];
                const float qq[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
                const float x[8] = {__uint_as_float(xv.x << 16), __uint_as_float(xv.x & 0xffff0000u),
                                    __uint_as_float(xv.y << 16), __uint_as_float(xv.y & 0xffff0000u),
                                    __uint_as_float(xv.z << 16), __uint_as_float(xv.z & 0xffff0000u),
                                    __uint_as_float(xv.w << 16), __uint_as_float(xv.w & 0xffff0000u)};
                float acc[4] = {a0, a1, a2, a3};
#pragma unroll
                for (int e = 0; e < 8; e++) {
                    if (L2) {
                        const float tt = qq[e] - x[e];
                        acc[e & 3] = fmaf(tt, tt, acc[e & 3]);
                    } else {
                        acc[e & 3] = fmaf(qq[e], x[e], acc[e & 3]);
                    }
                }
                a0 = acc[0];
                a1 = acc[1];
                a2 = acc[2];
                a3 = acc[3];
            }
        }
        const float sc = warp_sum((a0 + a2) + (a1 + a3));
        if (lane == 0) {
            Cand c2;
            c2.v = sc;
            c2.id = id;
            ex[c] = c2;
        }
    }
    const int P2 = next_pow2(n_valid, 32);
    if (rank_mode) {
        // exact scores sorted into the (now free) lower half of buf
        __syncthreads();
        block_rank_sort<L2>(ex, buf, n_valid);
        for (int i = n_valid + threadIdx.x; i < P2; i += blockDim.x) {
            Cand c;
            c.v = worst_val<L2>();
            c.id = kEmptyId;
            buf[i] = c;
        }
        __syncthreads();
        ex = buf;
    } else {
        for (int i = n_valid + threadIdx.x; i < P2; i += blockDim.x) {
            Cand c;
            c.v = worst_val<L2>();
            c.id = kEmptyId;
            ex[i] = c;
        }
        __syncthreads();
        block_sort_best_first<L2>(ex, P2);
    }

    // exact top-k = ex[0..k) + the containment check
    for (int e = threadIdx.x; e < p.k; e += blockDim.x) {
        Cand c;
        c.v = worst_val<L2>();
        c.id = kEmptyId;
        if (e < P2) c = ex[e];
        p.out_d[(size_t)qi * p.k + e] = c.v;
        p.out_i[(size_t)qi * p.k + e] = (c.id == kEmptyId) ? -1 : (int64_t)c.id + p.id_base;
    }
    if (threadIdx.x == 0) {
        bool ok = true;
        // every row outside buf[0..n_valid) has an approx score no better than buf[n_valid-1]
        // (only meaningful when K' survivors exist; otherwise every row seen IS a survivor)
        if (total >= p.kprime && (int64_t)n_valid < p.n_rows) {
            const float approx_worst = sorted[n_valid - 1].v;
            const float qn = sqrtf(s_qnorm2);
            const float xn = sqrtf(__uint_as_float(*p.max_norm_bits));
            float eps = p.err_coeff * qn * xn;
            const Cand kth = (p.k - 1 < P2) ? ex[p.k - 1] : ex[P2 - 1];
            if (L2) {
                // approx value is ||x||^2 - 2 q.x~ ; exact is sum (q-x)^2 = ||q||^2 + ||x||^2 - 2 q.x
                eps = 2.f * eps + 1e-6f * (s_qnorm2 + xn * xn);
                ok = (kth.v - s_qnorm2) < approx_worst - eps;
            } else {
                ok = kth.v > approx_worst + eps;
            }
            if (kth.id == kEmptyId) ok = false;  // fewer than k finite survivors: be safe
        }
        p.fail_flags[qi] = ok ? 0 : 1;
        if (!ok) atomicAdd(p.fail_count, 1);
    }
}

// One warp per query slot: bucket map for the score histogram, zeroed histogram row, and the
// grid-wide threshold preset to the worst value.
//   IP : goodness = score      in [-B, B],               B = ||q|| * max||x||
//   L2 : goodness = -(||x||^2 - 2 q.x) in [-(maxn + 2B), 2B]
__global__ void gemm_prepare_kernel(const float* __restrict__ q, int d_pad, int nq, const unsigned* max_norm_bits,
                                    int l2, unsigned* hist, float* qscale, int* gthr, int* overflow) {
    const int qi = blockIdx.x;
    const int lane = threadIdx.x;
    float s = 0.f;
    if (qi < nq)
        for (int i = lane; i < d_pad; i += 32) {
            const float v = q[(size_t)qi * d_pad + i];
            s = fmaf(v, v, s);
        }
    s = warp_sum(s);
    for (int b = lane; b < kGemmBuckets; b += 32) hist[(size_t)qi * kGemmBuckets + b] = 0u;
    if (lane == 0) {
        const float maxn = __uint_as_float(*max_norm_bits);
        const float B = sqrtf(s) * sqrtf(maxn) * 1.001f + 1e-30f;
        const float lo = l2 ? -(maxn + 2.f * B) : -B;
        const float hi = l2 ? 2.f * B : B;
        qscale[2 * qi] = lo;
        qscale[2 * qi + 1] = (hi > lo) ? (float)kGemmBuckets / (hi - lo) : 0.f;
        gthr[qi] = l2 ? 0x7F7FFFFF : (int)(0xFF7FFFFFu ^ 0x7FFFFFFFu);  // enc(+FLT_MAX) / enc(-FLT_MAX)
        overflow[qi] = 0;
    }
}

// max over rows of the squared norm, as float bits (non-negative floats order like uints)
__global__ void max_norm_kernel(const float* __restrict__ norms, int64_t n, unsigned* __restrict__ out_bits) {
    float m = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        m = fmaxf(m, norms[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) atomicMax(out_bits, __float_as_uint(m));
}

cudaError_t launch_max_norm(const float* norms, int64_t n, unsigned* out_bits, cudaStream_t st) {
    cudaError_t e = cudaMemsetAsync(out_bits, 0, sizeof(unsigned), st);
    if (e != cudaSuccess || n <= 0) return e;
    int g = (int)((n + 1023) / 1024);
    if (g > 148 * 4) g = 148 * 4;
    max_norm_kernel<<<g, 256, 0, st>>>(norms, n, out_bits);
    return cudaGetLastError();
}

// fp32 [nq_slots][d_pad] -> bf16 (queries for the bf16-corpus GEMM); RNE
__global__ void f32_to_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int64_t n) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = __float2bfloat16_rn(src[i]);
}

cudaError_t launch_f32_to_bf16(const float* src, void* dst, int64_t n, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    int g = (int)((n + 255) / 256);
    if (g > 148 * 8) g = 148 * 8;
    f32_to_bf16_kernel<<<g, 256, 0, st>>>(src, reinterpret_cast<__nv_bfloat16*>(dst), n);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(sym);
    }
    return fn;
}

// [rows][d_pad] row-major matrix, box = 128 bytes of K x box_rows rows, 128B swizzle
static bool make_map(CUtensorMap* map, const void* base, int64_t rows, int d_pad, bool bf16, int box_rows) {
    EncodeTiledFn enc = get_encode();
    if (!enc) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)d_pad, (cuuint64_t)(rows > 0 ? rows : 1)};
    const cuuint64_t strides[1] = {(cuuint64_t)d_pad * (bf16 ? 2 : 4)};
    const cuuint32_t box[2] = {(cuuint32_t)(kGemmKBytes / (bf16 ? 2 : 4)), (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = enc(map, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                           const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           kGemmKBytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

size_t gemm_cand_bytes(int grid, int nq_pad, int kp_cap) { return (size_t)grid * nq_pad * kp_cap * sizeof(Cand); }
size_t gemm_aux_bytes() {
    return (size_t)kGemmMaxBlock * (kGemmBuckets * sizeof(unsigned) + 2 * sizeof(float) + sizeof(int)) +
           (size_t)kGemmMaxGrid * kGemmMaxBlock * sizeof(int) + (size_t)kGemmMaxBlock * sizeof(int) + 1024;
}
void gemm_aux_layout(void* aux, unsigned** hist, float** qscale, int** gthr, int** list_counts, int** overflow) {
    uint8_t* b = reinterpret_cast<uint8_t*>(aux);
    *hist = reinterpret_cast<unsigned*>(b);
    *qscale = reinterpret_cast<float*>(b + (size_t)kGemmMaxBlock * kGemmBuckets * sizeof(unsigned));
    *gthr = reinterpret_cast<int*>(b + (size_t)kGemmMaxBlock * (kGemmBuckets * sizeof(unsigned) + 2 * sizeof(float)));
    *overflow = *gthr + kGemmMaxBlock;
    *list_counts = *overflow + kGemmMaxBlock;
}

// survivors kept per query: the slack beyond k must cover the rows whose approx score lies
// within 2*eps of the k-th best; their number grows with k (and with N), so it scales with k
// filter: kFilterTf32 (fp32 rows through tf32 MMA), kFilterBf16 (bf16 rows, bf16-rounded queries),
// kFilterShadow (bf16 shadow of fp32 rows: both operands rounded).  The error bound grows in that
// order (see launch_gemm_search), and so must the slack.  boost_pct (>= 100) widens the selection
// after searches whose containment check failed (api.cu raises it; data with many near-ties).
// Measured on 4M x 1024 unit-norm rows, 256 queries (r01 notes): the filter kernel slows by ~25%
// per doubling of K' (more chunks take the append path), one scan-path fallback costs about as
// much as the whole block, so K' is the smallest value that gave no fallback plus a margin.
int gemm_kp_cap(int k, int filter, int boost_pct) {
    if (knobs().gemm_kpcap > 0) {  // tuning knob
        if (knobs().gemm_kpcap == 256 || k + kGemmSlack > 32) return 256;
        return 32;
    }
    if (filter != kFilterTf32 || boost_pct > 100) return 256;
    return (k + kGemmSlack <= 32) ? 32 : 256;
}
int gemm_kprime(int k, int filter, int boost_pct) {
    if (gemm_kp_cap(k, filter, boost_pct) == 32) return 32;  // the register list below always holds 32
    // selection size K' (the per-thread slices stay at 256 entries: with warm-up thresholds only
    // ~K'/grid survivors reach a thread)
    int kp = 2 * k + 24;
    if (k + kGemmSlack <= 32) kp = 32;
    if (filter == kFilterBf16) kp = 3 * k + 34;
    if (filter == kFilterShadow) kp = 4 * k + 72;
    if (knobs().gemm_kprime > 0) kp = knobs().gemm_kprime > k ? knobs().gemm_kprime : k + 1;  // tuning knob
    if (boost_pct > 100) kp = (int)((int64_t)kp * boost_pct / 100);
    const int cap = k + 512;
    return kp < cap ? kp : cap;
}

// the opt-in to > 48 KB of dynamic shared memory is per kernel and per device: asked for once
static cudaError_t ensure_smem_optin_gemm(const void* kern, size_t smem, size_t (&granted)[64]) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (smem <= granted[dev]) return cudaSuccess;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) granted[dev] = smem;
    return e;
}

template <bool BF16, int MT, int KP, bool L2, int CS, bool PAIR = false>
static cudaError_t launch_filter_one(const CUtensorMap& mq, const CUtensorMap& mx, const GemmParams& p, int grid,
                                     cudaStream_t st) {
    constexpr int STAGES = (PAIR ? 6 : (MT == 1 ? 4 : 3)) * (128 / kGemmKBytes);
    constexpr int STAGE_BYTES = (PAIR ? kGemmBBytes / 2 : kGemmBBytes) + MT * kGemmABytes;
    const size_t smem = 1024 + (size_t)STAGES * STAGE_BYTES + 256 + (L2 ? (size_t)4 * MT * kGemmTileN * sizeof(float) : 0);
    auto kern = gemm_filter_kernel<BF16, MT, KP, L2, CS, PAIR>;
    static size_t granted[64] = {};
    cudaError_t e = ensure_smem_optin_gemm(reinterpret_cast<const void*>(kern), smem, granted);
    if (e != cudaSuccess) return e;
    if (CS == 1 && !PAIR) {
        kern<<<grid, 128 + 128 * MT, smem, st>>>(mq, mx, p);
        return cudaGetLastError();
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(128 + 128 * MT);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = PAIR ? 2 : CS;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, mq, mx, p);
}

// mode: 1 = plain CTAs, 2 = CTA pairs with query multicast, 3 = 2-CTA MMA (MT must be 1)
template <bool BF16, int MT, int KP>
static cudaError_t launch_filter_l2(bool l2, int cs, const CUtensorMap& mq, const CUtensorMap& mx, const GemmParams& p,
                                    int grid, cudaStream_t st) {
    if constexpr (MT == 1) {
        if (cs == 3)
            return l2 ? launch_filter_one<BF16, 1, KP, true, 1, true>(mq, mx, p, grid, st)
                      : launch_filter_one<BF16, 1, KP, false, 1, true>(mq, mx, p, grid, st);
    }
    if (cs == 2)
        return l2 ? launch_filter_one<BF16, MT, KP, true, 2>(mq, mx, p, grid, st)
                  : launch_filter_one<BF16, MT, KP, false, 2>(mq, mx, p, grid, st);
    return l2 ? launch_filter_one<BF16, MT, KP, true, 1>(mq, mx, p, grid, st)
              : launch_filter_one<BF16, MT, KP, false, 1>(mq, mx, p, grid, st);
}

template <bool L2, bool BF16, int WARPS>
static cudaError_t launch_rescore_one(const RescoreParams& rp, int nq, size_t smem, cudaStream_t st) {
    auto kern = rescore_kernel<L2, BF16, WARPS>;
    static size_t granted[64] = {};
    cudaError_t e = ensure_smem_optin_gemm(reinterpret_cast<const void*>(kern), smem, granted);
    if (e != cudaSuccess) return e;
    kern<<<nq, WARPS * 32, smem, st>>>(rp);
    return cudaGetLastError();
}

template <bool L2>
static cudaError_t launch_rescore_bf(bool bf16, const RescoreParams& rp, int nq, size_t smem, cudaStream_t st) {
    // up to 32 queries: 32 warps per query (the lists of all CTAs are walked and K' rows fetched
    // by one CTA; measured 64 -> 22 us for a single query with K' = 112)
    const bool few = nq <= 32;
    if (bf16)
        return few ? launch_rescore_one<L2, true, kRescoreWarpsFew>(rp, nq, smem, st)
                   : launch_rescore_one<L2, true, kRescoreWarps>(rp, nq, smem, st);
    return few ? launch_rescore_one<L2, false, kRescoreWarpsFew>(rp, nq, smem, st)
               : launch_rescore_one<L2, false, kRescoreWarps>(rp, nq, smem, st);
}

// One block of up to 256 queries against the whole shard.
//   q_f32:   [nq_pad][d_pad] prepared fp32 queries (nq_pad = 128 or 256 rows, unused rows zero)
//   q_mma:   the same matrix in the MMA input type (== q_f32 for an fp32 corpus, bf16 copy otherwise)
cudaError_t launch_gemm_search(const GemmSearchArgs& a, cudaStream_t st) {
    if (a.nq < 1 || a.nq > 2 * kGemmTileM || a.k < 1 || a.k > kMaxK) return cudaErrorInvalidValue;
    const int MT = a.nq <= kGemmTileM ? 1 : 2;
    const int filter = a.shadow ? kFilterShadow : (a.bf16 ? kFilterBf16 : kFilterTf32);
    const int boost = a.kp_boost_pct < 100 ? 100 : a.kp_boost_pct;
    const int kp_cap = gemm_kp_cap(a.k, filter, boost);
    const int kprime = gemm_kprime(a.k, filter, boost);
    const int64_t n_tiles_all = (a.n_rows + kGemmTileN - 1) / kGemmTileN;
    int cs = 2;  // query-operand multicast over CTA pairs
    if (knobs().gemm_cluster > 0) cs = knobs().gemm_cluster == 1 ? 1 : 2;
    if (n_tiles_all < 2 || a.num_sms < 2) cs = 1;
    if (cs == 2) {
        // TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load
        // from another GPU's memory hung the GPU.  The multicast operand is the index's own query
        // buffer; anything that is not plain device memory of the current GPU takes the
        // non-multicast variant.
        cudaPointerAttributes pa{};
        int cur = -1;
        if (cudaGetDevice(&cur) != cudaSuccess || cudaPointerGetAttributes(&pa, a.q_mma) != cudaSuccess ||
            pa.type != cudaMemoryTypeDevice || pa.device != cur) {
            cudaGetLastError();
            cs = 1;
        }
    }
    // more than 128 queries: 2-CTA MMA over CTA pairs (double-buffered accumulators) unless turned off
    bool pair = MT == 2 && cs == 2;
    if (knobs().gemm_pair >= 0) pair = pair && knobs().gemm_pair != 0;
    if (pair) cs = 3;
    // what the filter pass streams: the corpus itself, or its bf16 shadow
    const bool fbf16 = a.bf16 || a.shadow != nullptr;
    const void* fcorpus = a.shadow ? a.shadow : a.corpus;
    if (a.shadow && (a.d_pad % 8) != 0) return cudaErrorInvalidValue;
    CUtensorMap mq, mx;
    if (!make_map(&mq, a.q_mma, (int64_t)MT * kGemmTileM, a.d_pad, fbf16, pair ? kGemmTileM : kGemmTileM / cs))
        return cudaErrorNotSupported;
    if (!make_map(&mx, fcorpus, a.n_rows, a.d_pad, fbf16, pair ? kGemmTileN / 2 : kGemmTileN)) return cudaErrorNotSupported;
    GemmParams p{};
    p.n_rows = a.n_rows;
    p.nq = a.nq;
    p.d_pad = a.d_pad;
    p.kprime = kprime;
    p.num_kb = (a.d_pad * (fbf16 ? 2 : 4) + kGemmKBytes - 1) / kGemmKBytes;
    p.n_tiles = (a.n_rows + kGemmTileN - 1) / kGemmTileN;
    p.norms = a.norms;
    p.cand = a.cand;
    p.gthr = a.gthr;
    p.corpus = fcorpus;
    p.prefetch_tiles = knobs().gemm_prefetch;    // tuning knob; measured: the extra L2 fill stream hurts (r01 notes)
    p.debug_no_select = knobs().gemm_noselect;
    p.hist = a.hist;
    p.list_counts = a.list_counts;
    p.overflow = a.overflow;
    p.qscale = a.qscale;
    gemm_prepare_kernel<<<MT * kGemmTileM, 32, 0, st>>>(a.q_f32, a.d_pad, a.nq, a.max_norm_bits, a.l2 ? 1 : 0, a.hist,
                                                        a.qscale, a.gthr, a.overflow);
    p.evict_first = ((size_t)a.n_rows * a.d_pad * (fbf16 ? 2 : 4) > ((size_t)96 << 20)) ? 1 : 0;
    int grid = (int)(p.n_tiles < a.num_sms ? p.n_tiles : a.num_sms);
    if (grid < 1) grid = 1;
    if (pair && grid < 2) grid = 2;
    grid -= grid % (pair ? 2 : cs);
    // TMA coordinates are 32-bit: rows (including the padded tail tiles of a cluster launch) must fit
    if (a.n_rows + (int64_t)kGemmTileN * (grid + 1) >= (1ll << 31)) return cudaErrorNotSupported;
    {
        // warm-up pays one extra tile per CTA; worth it when the lists are large (memory heaps)
        // and every CTA has several tiles
        p.warmup = kp_cap > 32 ? 1 : 0;  // the append-only lists of the large-K' variant rely on it
        if (knobs().gemm_warmup >= 0 && kp_cap == 32) p.warmup = knobs().gemm_warmup;
    }
    cudaError_t e;
    if (fbf16) {
        if (MT == 1)
            e = kp_cap == 32 ? launch_filter_l2<true, 1, 32>(a.l2, cs, mq, mx, p, grid, st)
                             : launch_filter_l2<true, 1, 256>(a.l2, cs, mq, mx, p, grid, st);
        else if (pair)
            e = kp_cap == 32 ? launch_filter_l2<true, 1, 32>(a.l2, 3, mq, mx, p, grid, st)
                             : launch_filter_l2<true, 1, 256>(a.l2, 3, mq, mx, p, grid, st);
        else
            e = kp_cap == 32 ? launch_filter_l2<true, 2, 32>(a.l2, cs, mq, mx, p, grid, st)
                             : launch_filter_l2<true, 2, 256>(a.l2, cs, mq, mx, p, grid, st);
    } else {
        if (MT == 1)
            e = kp_cap == 32 ? launch_filter_l2<false, 1, 32>(a.l2, cs, mq, mx, p, grid, st)
                             : launch_filter_l2<false, 1, 256>(a.l2, cs, mq, mx, p, grid, st);
        else if (pair)
            e = kp_cap == 32 ? launch_filter_l2<false, 1, 32>(a.l2, 3, mq, mx, p, grid, st)
                             : launch_filter_l2<false, 1, 256>(a.l2, 3, mq, mx, p, grid, st);
        else
            e = kp_cap == 32 ? launch_filter_l2<false, 2, 32>(a.l2, cs, mq, mx, p, grid, st)
                             : launch_filter_l2<false, 2, 256>(a.l2, cs, mq, mx, p, grid, st);
    }
    if (e != cudaSuccess) return e;

    RescoreParams rp{};
    rp.corpus = a.corpus;
    rp.n_rows = a.n_rows;
    rp.d_pad = a.d_pad;
    rp.row_bytes = a.d_pad * (a.bf16 ? 2 : 4);
    rp.q = a.q_f32;
    rp.cand = a.cand;
    rp.list_counts = a.list_counts;
    rp.overflow = a.overflow;
    rp.qscale = a.qscale;
    rp.grid = pair ? grid / 2 : grid;  // survivor lists per query: one per work unit (CTA or CTA pair)
    rp.q_stride = MT * kGemmTileM;
    rp.kp_stride = kp_cap;
    rp.kprime = kprime;
    rp.k = a.k;
    rp.max_norm_bits = a.max_norm_bits;
    // |q.x - (q.x)~| <= c * ||q|| * ||x|| (Cauchy-Schwarz over the per-product relative errors):
    //   tf32: both operands truncated to 11 significant bits, 2 * 2^-10 = 1.95e-3
    //   bf16 rows: the query is rounded to 8 significant bits, 2^-8 = 3.91e-3, the rows are exact
    //   bf16 shadow of fp32 rows: both sides rounded, 2 * 2^-8 = 7.81e-3
    // plus 5e-4 for the fp32 accumulation inside the tensor core.
    rp.err_coeff = filter == kFilterShadow ? 8.4e-3f : (filter == kFilterBf16 ? 4.5e-3f : 2.5e-3f);
    rp.out_d = a.out_d;
    rp.out_i = a.out_i;
    rp.id_base = a.id_base;
    rp.fail_flags = a.fail_flags;
    rp.fail_count = a.fail_count;
    const size_t rs_smem = (size_t)a.d_pad * sizeof(float) + (size_t)(kRescoreCap + kRescoreMaxKp) * sizeof(Cand);
    e = a.l2 ? launch_rescore_bf<true>(a.bf16, rp, a.nq, rs_smem, st) : launch_rescore_bf<false>(a.bf16, rp, a.nq, rs_smem, st);
    return e;
}

}  // namespace b200r
