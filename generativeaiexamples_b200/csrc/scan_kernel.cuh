// The scan kernel and its launch templates, shared by the translation units that instantiate it
// (scan_inst_*.cu: one group of <NQ, KPL> variants each, so that they compile in parallel -- in one
// file the 56 instances took four minutes).  See scan.cu for the design notes.
#pragma once
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
    // (the tail of) their tiles as they go, so the slowest SM no longer decides when the kernel ends; the
    // producer tells the consumers which tile a stage holds through this array (written before the arrive).
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
    // static assignment: runs of tile_block consecutive tiles per CTA while whole rounds of grid * tile_block
    // tiles remain, the rest dealt one by one (tile_block = 1: tile t -> CTA t mod grid)
    int64_t t_next = 0;  // how many tiles this CTA has drawn
    const int64_t tb = p.tile_block > 1 ? p.tile_block : 1;
    const int64_t round_tiles = (int64_t)gridDim.x * tb;
    const int64_t blocked_tiles = (n_tiles / round_tiles) * round_tiles;  // tiles covered by whole rounds
    const int64_t blocked_mine = blocked_tiles / gridDim.x;              // of which this CTA's
    // hybrid (dyn_from > 0): tiles below dyn_from are dealt statically (t mod grid), the tail is drawn from
    // the counter, so the CTAs level out at the end without a counter round trip per tile before that
    const int64_t dyn_from = dyn ? (p.dyn_from < n_tiles ? p.dyn_from : n_tiles) : 0;
    auto draw_tile = [&]() -> int64_t {
        if (dyn) {
            if (dyn_from > 0) {
                const int64_t t = (t_next++) * gridDim.x + blockIdx.x;
                if (t < dyn_from) return t;
            }
            return dyn_from + (int64_t)atomicAdd(p.tile_counter, 1u);
        }
        const int64_t i = t_next++;
        if (i < blocked_mine) return (i / tb) * round_tiles + (int64_t)blockIdx.x * tb + (i % tb);
        return blocked_tiles + (i - blocked_mine) * gridDim.x + blockIdx.x;
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
                t_next = 0;
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
static inline cudaError_t launch_mk(const ScanParams& p, bool l2, bool bf16, int grid, size_t smem, cudaStream_t st,
                             bool flagged) {
    if (l2) return bf16 ? launch_one<NQ, KPL, true, true>(p, grid, smem, st, flagged)
                        : launch_one<NQ, KPL, true, false>(p, grid, smem, st, flagged);
    return bf16 ? launch_one<NQ, KPL, false, true>(p, grid, smem, st, flagged)
                : launch_one<NQ, KPL, false, false>(p, grid, smem, st, flagged);
}


}  // namespace b200r
