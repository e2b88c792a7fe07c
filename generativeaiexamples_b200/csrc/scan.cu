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

#include <cooperative_groups.h>

#include <cstdlib>

namespace b200r {

// Consumer warps and rows per warp task.  One query: 16 warps, a warp owns whole rows one at a time
// (x and q are both read from shared memory: 2 bytes of LDS per corpus byte, 39 % of the LSU peak at
// HBM speed).  Several queries: a warp takes FOUR consecutive rows at a time and holds each query
// vector's slice in registers across them, so shared-memory traffic is (4 + NQ) / 4 bytes per corpus
// byte instead of 1 + NQ (at NQ = 4 that was 5x, which made the four-query variants LSU-bound at half
// the HBM rate); the 64 accumulators of 4 rows x 4 queries need more registers, hence 8 warps.
template <int NQ>
struct ScanShape {
#ifndef B200R_SCAN_NQ4_WARPS
#define B200R_SCAN_NQ4_WARPS 8
#endif
    static constexpr int kConsumerWarps = NQ == 1 ? 16 : B200R_SCAN_NQ4_WARPS;
    static constexpr int kRowsPerTask = NQ == 1 ? 1 : 4;
    static constexpr int kThreads = 32 * (1 + kConsumerWarps);
};
constexpr int kScanMaxThreads = ScanShape<1>::kThreads;
constexpr int kScanMaxStages = 16;

__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// FLAGGED: the cooperative re-run of the queries that failed the GEMM route's containment check.
// The same body runs once per group of NQ flagged queries, with a grid-wide barrier between groups
// (the candidate scratch, the ticket and the shared-memory ring are reused); the pipeline position
// (stage, phase) simply carries on from one group to the next.
template <int NQ, int KPL, bool L2, bool BF16, bool FLAGGED>
__global__ void __launch_bounds__(ScanShape<NQ>::kThreads, 1) scan_topk_kernel(const ScanParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    constexpr int NCW = ScanShape<NQ>::kConsumerWarps;
    constexpr int RPW = ScanShape<NQ>::kRowsPerTask;
    constexpr int kScanThreads = ScanShape<NQ>::kThreads;
    constexpr int KCAP = KPL * 32;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int nst = p.nstages;

    uint8_t* stage_base = smem;
    float* qs = reinterpret_cast<float*>(smem + (size_t)nst * p.stage_bytes);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(qs + (size_t)NQ * p.d);
    uint64_t* empty_bar = full_bar + nst;
    __shared__ int s_is_last;
    // tile held by each stage (-1: the stream has ended).  With a global tile counter the CTAs draw
    // tiles as they go, so the slowest SM no longer decides when the kernel ends; the producer tells
    // the consumers which tile a stage holds through this array (written before the barrier arrive).
    __shared__ long long s_tile[kScanMaxStages];
    const bool dyn = !FLAGGED && p.tile_counter != nullptr;
    if (p.trace && threadIdx.x == 0) p.trace[blockIdx.x * 8 + 0] = global_ns();

    int n_groups = 1, n_flagged = 0;
    if constexpr (FLAGGED) {
        n_flagged = *p.flag_count;
        if (n_flagged == 0) return;  // the common case: every query passed its check
        n_groups = (n_flagged + NQ - 1) / NQ;
    }

    // ---- set-up ---------------------------------------------------------
    if (threadIdx.x == 0) {
        for (int s = 0; s < nst; s++) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], NCW);
        }
        fence_mbar_init();
    }
    // position in the stage ring: the producer lane and every consumer warp advance it identically
    int s = 0;
    uint32_t ph = 0;
    const int64_t n_tiles = p.n_tiles;
    const int R = p.rows_per_tile;
    // producer state (thread 0 only): the next tile of this CTA
    int64_t t_next = blockIdx.x;
    auto draw_tile = [&]() -> int64_t {
        if (dyn) return (int64_t)atomicAdd(p.tile_counter, 1u);
        const int64_t t = t_next;
        t_next += gridDim.x;
        return t;
    };
    auto issue_tile = [&](int64_t t, uint64_t pol) {
        const uint8_t* g = reinterpret_cast<const uint8_t*>(p.corpus);
        if (t >= n_tiles) {  // end of the stream: an empty stage that says so
            s_tile[s] = -1;
            mbar_arrive(&full_bar[s]);
            return;
        }
        s_tile[s] = t;
        const int64_t row0 = t * R;
        const int64_t rem = p.n_rows - row0;
        const uint32_t rows = (uint32_t)(rem < R ? rem : R);
        const uint32_t bytes = rows * (uint32_t)p.row_bytes;
        mbar_arrive_expect_tx(&full_bar[s], bytes);
        if (p.evict_first)
            bulk_g2s_hint(stage_base + (size_t)s * p.stage_bytes, g + row0 * (int64_t)p.row_bytes, bytes, &full_bar[s], pol);
        else
            bulk_g2s(stage_base + (size_t)s * p.stage_bytes, g + row0 * (int64_t)p.row_bytes, bytes, &full_bar[s]);
    };
    // The first stages are requested before the queries are staged: thread 0 initialised the barriers
    // itself, so its bulk copies can leave while the other threads still fetch the query vectors.
    int pro_s = 0;          // producer's own ring position (it runs ahead of the consumers' s / ph)
    uint32_t pro_ph = 0;
    int64_t t_pending = -2; // tile drawn ahead (-2: none)
    int prologue = 0;
    if (!FLAGGED && threadIdx.x == 0) {
        const uint64_t pol = policy_evict_first();
        for (; prologue < nst; prologue++) {
            const int64_t t = draw_tile();
            const int keep_s = s;
            s = pro_s;
            issue_tile(t, pol);
            s = keep_s;
            if (++pro_s == nst) {
                pro_s = 0;
                pro_ph ^= 1;
            }
            if (t >= n_tiles) {
                prologue = -1;  // the stream ended inside the prologue
                break;
            }
        }
        if (prologue >= 0 && dyn) t_pending = draw_tile();
    }
  for (int grp = 0; grp < n_groups; grp++) {
    const int nq_active = FLAGGED ? (n_flagged - grp * NQ < NQ ? n_flagged - grp * NQ : NQ) : p.nq_active;
    // queries -> shared memory (fp32, zero padded to NQ rows by the host)
    if constexpr (FLAGGED) {
        float4* q4 = reinterpret_cast<float4*>(qs);
        const int d4 = p.d / 4;
        for (int qi = 0; qi < NQ; qi++) {
            const bool have = qi < nq_active;
            const float4* qg = reinterpret_cast<const float4*>(p.q + (size_t)(have ? p.flag_list[grp * NQ + qi] : 0) * p.d);
            for (int i = threadIdx.x; i < d4; i += blockDim.x)
                q4[(size_t)qi * d4 + i] = have ? qg[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    } else {
        const float4* qg = reinterpret_cast<const float4*>(p.q);
        float4* q4 = reinterpret_cast<float4*>(qs);
        const int n4 = NQ * p.d / 4;
        for (int i = threadIdx.x; i < n4; i += blockDim.x) q4[i] = qg[i];
    }
    __syncthreads();
    if (p.trace && threadIdx.x == 0) p.trace[blockIdx.x * 8 + 1] = global_ns();

    WarpTopK<KPL, L2> top[NQ];
    // rows that beat the running threshold wait here (lane i holds the i-th of them) until 32 have
    // gathered, then enter the list in one sort + bitonic merge instead of 32 serial inserts
    float pend_v[NQ];
    uint32_t pend_i[NQ];
    int npend[NQ];
#pragma unroll
    for (int qi = 0; qi < NQ; qi++) {
        top[qi].init(p.k);
        pend_v[qi] = worst_val<L2>();
        pend_i[qi] = kEmptyId;
        npend[qi] = 0;
    }

    if (warp == 0) {
        // ---- producer ---------------------------------------------------
        if (lane == 0) {
            const uint64_t pol = policy_evict_first();
            if (FLAGGED) {  // every group starts a fresh walk over the tiles
                t_next = blockIdx.x;
                pro_s = s;
                pro_ph = ph;
                prologue = 0;
            }
            bool ended = prologue < 0;
            while (!ended) {
                // the next tile is drawn before waiting for a free stage, so the counter's round trip
                // to L2 is hidden behind that wait
                const int64_t t = (t_pending != -2) ? t_pending : draw_tile();
                t_pending = -2;
                if (dyn && t < n_tiles) t_pending = draw_tile();
                mbar_wait(&empty_bar[pro_s], pro_ph ^ 1);
                const int keep_s = s;
                s = pro_s;
                issue_tile(t, pol);
                s = keep_s;
                if (++pro_s == nst) {
                    pro_s = 0;
                    pro_ph ^= 1;
                }
                ended = t >= n_tiles;
            }
            // leave the consumers' ring position where they will be: one past the end-of-stream stage
            s = pro_s;
            ph = pro_ph;
        }
        __syncwarp();
    } else {
        // ---- consumers --------------------------------------------------
        const int cw = warp - 1;
        int64_t seq = 0;  // rows this CTA has streamed before the current tile
        const int nvec = BF16 ? (p.d >> 3) : (p.d >> 2);  // 16-byte vectors per row
        for (;;) {
            mbar_wait(&full_bar[s], ph);
            const int64_t t = s_tile[s];
            if (t < 0) {  // end of the stream: release the marker stage and step past it
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty_bar[s]);
                if (++s == nst) {
                    s = 0;
                    ph ^= 1;
                }
                break;
            }
            const int64_t row0 = t * R;
            const int64_t rem = p.n_rows - row0;
            const int rows = (int)(rem < R ? rem : R);
            const uint8_t* tile = stage_base + (size_t)s * p.stage_bytes;
            if constexpr (RPW > 1) {
                // round-robin ownership of groups of RPW consecutive rows across the CTA's row stream
                const int ngroups = (rows + RPW - 1) / RPW;
                int g = (int)(((int64_t)cw - seq) % NCW);
                if (g < 0) g += NCW;
                for (; g < ngroups; g += NCW) {
                    const int r0 = g * RPW;
                    const int nr = rows - r0 < RPW ? rows - r0 : RPW;
                    const uint4* xr[RPW];
#pragma unroll
                    for (int rr = 0; rr < RPW; rr++)  // rows past the end of the tile repeat the last one (never offered)
                        xr[rr] = reinterpret_cast<const uint4*>(tile + (size_t)(r0 + (rr < nr ? rr : nr - 1)) * p.row_bytes);
                    float acc[RPW][NQ][4];
#pragma unroll
                    for (int rr = 0; rr < RPW; rr++)
#pragma unroll
                        for (int qi = 0; qi < NQ; qi++) acc[rr][qi][0] = acc[rr][qi][1] = acc[rr][qi][2] = acc[rr][qi][3] = 0.f;
                    // same element order per (row, query) as the one-row path: scores are bit-identical
                    for (int i0 = 0; i0 < nvec; i0 += 32 * 8) {
#pragma unroll
                        for (int u = 0; u < 8; u++) {
                            const int i = i0 + u * 32 + lane;
                            if (i < nvec) {
                                if (!BF16) {
                                    float4 qv[NQ];
#pragma unroll
                                    for (int qi = 0; qi < NQ; qi++) qv[qi] = reinterpret_cast<const float4*>(qs + (size_t)qi * p.d)[i];
#pragma unroll
                                    for (int rr = 0; rr < RPW; rr++) {
                                        const uint4 xv = xr[rr][i];
                                        const float x0 = __uint_as_float(xv.x), x1 = __uint_as_float(xv.y),
                                                    x2 = __uint_as_float(xv.z), x3 = __uint_as_float(xv.w);
#pragma unroll
                                        for (int qi = 0; qi < NQ; qi++) {
                                            // packed fp32 (FFMA2 / FADD2): same IEEE results as the scalar code of the
                                            // one-row path, half the issue slots -- these variants are issue-bound
                                            if (L2) {
                                                float t0, t1, t2, t3;
                                                sub2(t0, t1, qv[qi].x, qv[qi].y, x0, x1);
                                                sub2(t2, t3, qv[qi].z, qv[qi].w, x2, x3);
                                                fma2(acc[rr][qi][0], acc[rr][qi][1], t0, t1, t0, t1);
                                                fma2(acc[rr][qi][2], acc[rr][qi][3], t2, t3, t2, t3);
                                            } else {
                                                fma2(acc[rr][qi][0], acc[rr][qi][1], qv[qi].x, qv[qi].y, x0, x1);
                                                fma2(acc[rr][qi][2], acc[rr][qi][3], qv[qi].z, qv[qi].w, x2, x3);
                                            }
                                        }
                                    }
                                } else {
                                    float x[RPW][8];
#pragma unroll
                                    for (int rr = 0; rr < RPW; rr++) {
                                        const uint4 xv = xr[rr][i];  // 8 bf16 -> fp32 (exact: bf16 is the top half of an fp32)
                                        x[rr][0] = __uint_as_float(xv.x << 16);
                                        x[rr][1] = __uint_as_float(xv.x & 0xffff0000u);
                                        x[rr][2] = __uint_as_float(xv.y << 16);
                                        x[rr][3] = __uint_as_float(xv.y & 0xffff0000u);
                                        x[rr][4] = __uint_as_float(xv.z << 16);
                                        x[rr][5] = __uint_as_float(xv.z & 0xffff0000u);
                                        x[rr][6] = __uint_as_float(xv.w << 16);
                                        x[rr][7] = __uint_as_float(xv.w & 0xffff0000u);
                                    }
#pragma unroll
                                    for (int qi = 0; qi < NQ; qi++) {
                                        const float4* qp = reinterpret_cast<const float4*>(qs + (size_t)qi * p.d) + 2 * i;
                                        const float4 qa = qp[0], qb = qp[1];
                                        const float qq[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
#pragma unroll
                                        for (int rr = 0; rr < RPW; rr++)
#pragma unroll
                                            for (int e = 0; e < 8; e += 2) {  // acc[e & 3] takes element e, then e + 4
                                                if (L2) {
                                                    float t0, t1;
                                                    sub2(t0, t1, qq[e], qq[e + 1], x[rr][e], x[rr][e + 1]);
                                                    fma2(acc[rr][qi][e & 3], acc[rr][qi][(e & 3) + 1], t0, t1, t0, t1);
                                                } else {
                                                    fma2(acc[rr][qi][e & 3], acc[rr][qi][(e & 3) + 1], qq[e], qq[e + 1],
                                                         x[rr][e], x[rr][e + 1]);
                                                }
                                            }
                                    }
                                }
                            }
                        }
                    }
                    // all RPW x NQ butterfly reductions advance together: one reduction is five DEPENDENT
                    // shuffle + add steps (~30 cycles each), and done one after the other -- with a
                    // data-dependent branch behind each -- they were two thirds of a task's time (ncu: the
                    // FADDs behind the shuffles held 34 % of all stall samples)
                    float scs[RPW][NQ];
#pragma unroll
                    for (int rr = 0; rr < RPW; rr++)
#pragma unroll
                        for (int qi = 0; qi < NQ; qi++)
                            scs[rr][qi] = (acc[rr][qi][0] + acc[rr][qi][2]) + (acc[rr][qi][1] + acc[rr][qi][3]);
#pragma unroll
                    for (int o = 16; o >= 1; o >>= 1)
#pragma unroll
                        for (int rr = 0; rr < RPW; rr++)
#pragma unroll
                            for (int qi = 0; qi < NQ; qi++) scs[rr][qi] += __shfl_xor_sync(0xffffffffu, scs[rr][qi], o);
#pragma unroll
                    for (int rr = 0; rr < RPW; rr++) {
                        if (rr < nr) {
                            const uint32_t row = (uint32_t)(row0 + r0 + rr);
#pragma unroll
                            for (int qi = 0; qi < NQ; qi++) {
                                const float sc = scs[rr][qi];
                                if (better_val<L2>(sc, top[qi].thr)) {  // row ids only grow: a tie never displaces
                                    if (lane == npend[qi]) {
                                        pend_v[qi] = sc;
                                        pend_i[qi] = row;
                                    }
                                    if (++npend[qi] == 32) {
                                        top[qi].offer_batch(pend_v[qi], pend_i[qi]);
                                        pend_v[qi] = worst_val<L2>();
                                        pend_i[qi] = kEmptyId;
                                        npend[qi] = 0;
                                    }
                                }
                            }
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&empty_bar[s]);
                seq += ngroups;  // (counts row groups on this path)
                if (++s == nst) {
                    s = 0;
                    ph ^= 1;
                }
                continue;
            }
            // round-robin row ownership across the CTA's row stream
            int r = (int)(((int64_t)cw - seq) % NCW);
            if (r < 0) r += NCW;
            for (; r < rows; r += NCW) {
                const uint4* xr = reinterpret_cast<const uint4*>(tile + (size_t)r * p.row_bytes);
                float acc[NQ][4];
#pragma unroll
                for (int qi = 0; qi < NQ; qi++) acc[qi][0] = acc[qi][1] = acc[qi][2] = acc[qi][3] = 0.f;
                for (int i0 = 0; i0 < nvec; i0 += 32 * 8) {
                    uint4 xv[8];
#pragma unroll
                    for (int u = 0; u < 8; u++) {
                        const int i = i0 + u * 32 + lane;
                        xv[u] = (i < nvec) ? xr[i] : make_uint4(0, 0, 0, 0);
                    }
#pragma unroll
                    for (int u = 0; u < 8; u++) {
                        const int i = i0 + u * 32 + lane;
                        if (i < nvec) {
                            if (!BF16) {
                                const float x0 = __uint_as_float(xv[u].x), x1 = __uint_as_float(xv[u].y),
                                            x2 = __uint_as_float(xv[u].z), x3 = __uint_as_float(xv[u].w);
#pragma unroll
                                for (int qi = 0; qi < NQ; qi++) {
                                    const float4 qv = reinterpret_cast<const float4*>(qs + (size_t)qi * p.d)[i];
                                    if (L2) {
                                        const float t0 = qv.x - x0, t1 = qv.y - x1, t2 = qv.z - x2, t3 = qv.w - x3;
                                        acc[qi][0] = fmaf(t0, t0, acc[qi][0]);
                                        acc[qi][1] = fmaf(t1, t1, acc[qi][1]);
                                        acc[qi][2] = fmaf(t2, t2, acc[qi][2]);
                                        acc[qi][3] = fmaf(t3, t3, acc[qi][3]);
                                    } else {
                                        acc[qi][0] = fmaf(qv.x, x0, acc[qi][0]);
                                        acc[qi][1] = fmaf(qv.y, x1, acc[qi][1]);
                                        acc[qi][2] = fmaf(qv.z, x2, acc[qi][2]);
                                        acc[qi][3] = fmaf(qv.w, x3, acc[qi][3]);
                                    }
                                }
                            } else {
                                // 8 bf16 -> fp32 (exact: bf16 is the top half of an fp32)
                                float x[8];
                                x[0] = __uint_as_float(xv[u].x << 16);
                                x[1] = __uint_as_float(xv[u].x & 0xffff0000u);
                                x[2] = __uint_as_float(xv[u].y << 16);
                                x[3] = __uint_as_float(xv[u].y & 0xffff0000u);
                                x[4] = __uint_as_float(xv[u].z << 16);
                                x[5] = __uint_as_float(xv[u].z & 0xffff0000u);
                                x[6] = __uint_as_float(xv[u].w << 16);
                                x[7] = __uint_as_float(xv[u].w & 0xffff0000u);
#pragma unroll
                                for (int qi = 0; qi < NQ; qi++) {
                                    const float4* qp = reinterpret_cast<const float4*>(qs + (size_t)qi * p.d) + 2 * i;
                                    const float4 qa = qp[0], qb = qp[1];
                                    const float qq[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
#pragma unroll
                                    for (int e = 0; e < 8; e++) {
                                        if (L2) {
                                            const float tt = qq[e] - x[e];
                                            acc[qi][e & 3] = fmaf(tt, tt, acc[qi][e & 3]);
                                        } else {
                                            acc[qi][e & 3] = fmaf(qq[e], x[e], acc[qi][e & 3]);
                                        }
                                    }
                                }
                            }
                        }
                    }
                }
                const uint32_t row = (uint32_t)(row0 + r);
#pragma unroll
                for (int qi = 0; qi < NQ; qi++) {
                    const float sc = warp_sum((acc[qi][0] + acc[qi][2]) + (acc[qi][1] + acc[qi][3]));
                    if (better_val<L2>(sc, top[qi].thr)) {  // row ids only grow: a tie never displaces
                        if (lane == npend[qi]) {
                            pend_v[qi] = sc;
                            pend_i[qi] = row;
                        }
                        if (++npend[qi] == 32) {
                            top[qi].offer_batch(pend_v[qi], pend_i[qi]);
                            pend_v[qi] = worst_val<L2>();
                            pend_i[qi] = kEmptyId;
                            npend[qi] = 0;
                        }
                    }
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[s]);
            seq += rows;
            if (++s == nst) {
                s = 0;
                ph ^= 1;
            }
        }
    }

    // ---- epilogue: per-CTA merge, then last-CTA grid merge ----------------
    if (p.trace && threadIdx.x == 32) p.trace[blockIdx.x * 8 + 2] = global_ns();
    if (warp > 0) {
#pragma unroll
        for (int qi = 0; qi < NQ; qi++) top[qi].offer_batch(pend_v[qi], pend_i[qi]);  // stragglers
    }
    __syncthreads();  // every stage consumed; the ring is free for reuse
    Cand* sm_lists = reinterpret_cast<Cand*>(smem);  // [NQ][NCW][KCAP]
    if (warp > 0) {
#pragma unroll
        for (int qi = 0; qi < NQ; qi++) top[qi].store(sm_lists + ((size_t)qi * NCW + (warp - 1)) * KCAP);
    }
    __syncthreads();
    // the NCW warp lists of every query fold into one by a tree of pairwise merges (log2 NCW rounds)
    tree_merge_lists<KPL, L2>(sm_lists, NCW, NQ, (size_t)NCW * KCAP, p.k);
    if (warp < NQ) {
        const Cand* src = sm_lists + (size_t)warp * NCW * KCAP;
        Cand* dst = p.scratch + ((size_t)blockIdx.x * NQ + warp) * KCAP;
#pragma unroll
        for (int j = 0; j < KPL; j++) dst[j * 32 + lane] = src[j * 32 + lane];
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        if (p.trace) p.trace[blockIdx.x * 8 + 3] = global_ns();
        const unsigned ticket = atomicAdd(p.counter, 1u);
        s_is_last = (ticket == gridDim.x - 1);
    }
    __syncthreads();
    if (s_is_last) {
    __threadfence();

    // phase A: every warp folds a strided subset of the CTA lists, per query
    constexpr int NW = kScanThreads / 32;  // every warp of this kernel shape
    for (int qi = 0; qi < NQ; qi++) {
        WarpTopK<KPL, L2> m;
        m.init(p.k);
        m.template merge_from<true>(p.scratch + (size_t)qi * KCAP, warp, (int)gridDim.x, NW, (size_t)NQ * KCAP);
        m.store(sm_lists + ((size_t)qi * NW + warp) * KCAP);
    }
    __syncthreads();
    // phase B: the NW partial lists of every query fold into one (tree of pairwise merges); one warp per
    // query writes the result
    tree_merge_lists<KPL, L2>(sm_lists, NW, NQ, (size_t)NW * KCAP, p.k);
    if (warp < NQ && warp < nq_active) {
        WarpTopK<KPL, L2> m;
        m.load(sm_lists + (size_t)warp * NW * KCAP, p.k);
        if (FLAGGED || p.xchg_ntargets == 0) {
            // result row: the query's own position (flagged re-run) or its slot in this launch
            const size_t orow = FLAGGED ? (size_t)p.flag_list[grp * NQ + warp] : (size_t)warp;
#pragma unroll
            for (int j = 0; j < KPL; j++) {
                const int e = j * 32 + lane;
                if (e < p.k) {
                    p.out_d[orow * p.k + e] = m.v[j];
                    p.out_i[orow * p.k + e] = (m.id[j] == kEmptyId) ? -1 : (int64_t)m.id[j] + p.id_base;
                }
            }
        } else {
            // fused exchange: this shard's top-k goes straight into slot [my rank] of EVERY rank's
            // exchange buffer (plain stores to peer-mapped addresses = NVLink writes)
            const size_t slot = (size_t)(p.xchg_epoch & 1u) * kXchgParityBytes + (size_t)p.xchg_rank * kXchgSlotBytes;
            for (int r = 0; r < p.xchg_ntargets; r++) {
                float* pd = reinterpret_cast<float*>(p.xchg_peer[r] + slot);
                int64_t* pi = reinterpret_cast<int64_t*>(p.xchg_peer[r] + slot + kXchgIdOffset);
#pragma unroll
                for (int j = 0; j < KPL; j++) {
                    const int e = j * 32 + lane;
                    if (e < p.k) {
                        pd[(size_t)warp * p.k + e] = m.v[j];
                        pi[(size_t)warp * p.k + e] = (m.id[j] == kEmptyId) ? -1 : (int64_t)m.id[j] + p.id_base;
                    }
                }
            }
            __threadfence_system();
        }
    }
    if (!FLAGGED && p.xchg_ntargets > 0) {
        __syncthreads();  // every result store of this CTA is fenced; now raise the flags
        if (threadIdx.x < p.xchg_ntargets) {
            uint32_t* flag = reinterpret_cast<uint32_t*>(p.xchg_peer[threadIdx.x] + kXchgFlagOffset +
                                                        ((p.xchg_epoch & 1u) * kXchgMaxWorld + p.xchg_rank) * 128);
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(p.xchg_epoch) : "memory");
        }
    }
    if (threadIdx.x == 0) {
        *p.counter = 0;  // ready for the next group / the next launch on this stream
        if (p.tile_counter) *p.tile_counter = 0;
        if (p.trace) p.trace[blockIdx.x * 8 + 4] = global_ns();
    }
    }  // s_is_last
    if constexpr (FLAGGED) {
        __threadfence();
        fence_proxy_async();  // the ring held candidate lists (generic stores); the next group's bulk copies write it again
        cooperative_groups::this_grid().sync();  // scratch, ticket and the shared-memory ring are free again
    }
  }  // groups
}

// ---------------------------------------------------------------------------
// The opt-in to > 48 KB of dynamic shared memory is per kernel and per device: asked for once (for
// the largest size seen so far), not on every launch.
template <typename K>
static cudaError_t ensure_smem_optin(K kern, size_t smem, size_t (&granted)[64]) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (smem <= granted[dev]) return cudaSuccess;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) granted[dev] = smem;
    return e;
}

template <int NQ, int KPL, bool L2, bool BF16>
static cudaError_t launch_one(const ScanParams& p, int grid, size_t smem, cudaStream_t st, bool flagged) {
    if (flagged) {
        // cooperative: the kernel loops over the flagged queries with grid-wide barriers in between
        auto kern = scan_topk_kernel<NQ, KPL, L2, BF16, true>;
        static size_t granted[64] = {};
        cudaError_t e = ensure_smem_optin(kern, smem, granted);
        if (e != cudaSuccess) return e;
        ScanParams pc = p;
        void* args[] = {&pc};
        return cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kern), dim3(grid), dim3(ScanShape<NQ>::kThreads), args, smem,
                                           st);
    }
    auto kern = scan_topk_kernel<NQ, KPL, L2, BF16, false>;
    static size_t granted[64] = {};
    cudaError_t e = ensure_smem_optin(kern, smem, granted);
    if (e != cudaSuccess) return e;
    kern<<<grid, ScanShape<NQ>::kThreads, smem, st>>>(p);
    return cudaGetLastError();
}

template <int NQ, int KPL>
static cudaError_t launch_mk(const ScanParams& p, bool l2, bool bf16, int grid, size_t smem, cudaStream_t st,
                             bool flagged) {
    if (l2) return bf16 ? launch_one<NQ, KPL, true, true>(p, grid, smem, st, flagged)
                        : launch_one<NQ, KPL, true, false>(p, grid, smem, st, flagged);
    return bf16 ? launch_one<NQ, KPL, false, true>(p, grid, smem, st, flagged)
                : launch_one<NQ, KPL, false, false>(p, grid, smem, st, flagged);
}

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
    int grid = (int)(p.n_tiles < num_sms ? p.n_tiles : num_sms);
    if (grid < 1) grid = 1;
    if (grid_out) *grid_out = grid;
    const size_t smem = (size_t)nst * stage_bytes + q_bytes + bar_bytes;
    if (KPL == 8) return launch_mk<1, 8>(p, l2, bf16, grid, smem, st, flagged);
    if (KPL == 16) return launch_mk<1, 16>(p, l2, bf16, grid, smem, st, flagged);
    if (KPL == 32) return launch_mk<1, 32>(p, l2, bf16, grid, smem, st, flagged);
    if (NQ == 1)
        return KPL == 1 ? launch_mk<1, 1>(p, l2, bf16, grid, smem, st, flagged) : launch_mk<1, 4>(p, l2, bf16, grid, smem, st, flagged);
    return KPL == 1 ? launch_mk<4, 1>(p, l2, bf16, grid, smem, st, flagged) : launch_mk<4, 4>(p, l2, bf16, grid, smem, st, flagged);
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
