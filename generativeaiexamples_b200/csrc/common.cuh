// Shared device/host helpers for libb200r (sm_100a only).
//
// Nothing here has a counterpart in the reference: the reference's search step is
// faiss-cpu's IndexFlat::search (see include/b200r.h for the call sites).  The
// candidate lists below implement the same *selection* as faiss's binary heap
// (faiss/utils/Heap.h) -- keep the k best, replace only on improvement -- but as
// warp-distributed sorted lists so that the score matrix never leaves the SM.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <float.h>

namespace b200r {

constexpr uint32_t kEmptyId = 0xFFFFFFFFu;

// ---------------------------------------------------------------------------
// PTX wrappers: mbarrier + 1-D bulk async copy (TMA engine, SASS UBLKCP)
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}

__device__ __forceinline__ void fence_mbar_init() {
    // make the inits visible to the async proxy (TMA) before first use
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(
                     smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}

// global -> shared bulk copy; completes `bytes` on `bar`.  src, dst and bytes
// must be multiples of 16.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// same, with an L2 eviction policy (streamed-once corpus: evict_first)
__device__ __forceinline__ void bulk_g2s_hint(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar,
                                              uint64_t policy) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::
            "r"(smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
        : "memory");
}

__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
    return p;
}

// Packed fp32 arithmetic (sm_100: FFMA2 / FADD2): two independent IEEE operations per instruction, the
// same results as two scalar ones at half the issue slots.
__device__ __forceinline__ void fma2(float& a0, float& a1, float x0, float x1, float y0, float y1) {
    asm("{\n\t.reg .b64 ra, rx, ry;\n\tmov.b64 ra, {%0, %1};\n\tmov.b64 rx, {%2, %3};\n\tmov.b64 ry, {%4, %5};\n\t"
        "fma.rn.f32x2 ra, rx, ry, ra;\n\tmov.b64 {%0, %1}, ra;\n\t}"
        : "+f"(a0), "+f"(a1)
        : "f"(x0), "f"(x1), "f"(y0), "f"(y1));
}
__device__ __forceinline__ void sub2(float& t0, float& t1, float a0, float a1, float b0, float b1) {
    asm("{\n\t.reg .b64 ra, rx, ry;\n\tmov.b64 rx, {%2, %3};\n\tmov.b64 ry, {%4, %5};\n\t"
        "sub.rn.f32x2 ra, rx, ry;\n\tmov.b64 {%0, %1}, ra;\n\t}"
        : "=f"(t0), "=f"(t1)
        : "f"(a0), "f"(a1), "f"(b0), "f"(b1));
}

__device__ __forceinline__ float warp_sum(float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    return v;
}

// ---------------------------------------------------------------------------
// ordering: SMALL = keep the smallest (squared L2), else keep the largest (IP).
// Ties on the score go to the lowest id, which makes every result deterministic
// (faiss leaves the order of equal distances unspecified).
// ---------------------------------------------------------------------------
template <bool SMALL>
__device__ __forceinline__ bool better_val(float a, float b) {
    return SMALL ? (a < b) : (a > b);
}

template <bool SMALL>
__device__ __forceinline__ bool better(float a, uint32_t ia, float b, uint32_t ib) {
    return better_val<SMALL>(a, b) || (a == b && ia < ib);
}

template <bool SMALL>
__device__ __forceinline__ float worst_val() {
    return SMALL ? FLT_MAX : -FLT_MAX;  // faiss pads with +FLT_MAX (L2) / -FLT_MAX (IP)
}

struct __align__(8) Cand {
    float v;
    uint32_t id;
};

// ---------------------------------------------------------------------------
// Warp-distributed sorted candidate list of capacity KPL*32, best first.
// Entry e lives in slot e/32 of lane e%32.  All methods are warp-collective.
// ---------------------------------------------------------------------------
template <int KPL, bool SMALL>
struct WarpTopK {
    float v[KPL];
    uint32_t id[KPL];
    float thr;        // score of entry k-1 (uniform across the warp)
    uint32_t thr_id;  // its id
    int k;

    __device__ __forceinline__ void init(int k_) {
        k = k_;
#pragma unroll
        for (int j = 0; j < KPL; j++) {
            v[j] = worst_val<SMALL>();
            id[j] = kEmptyId;
        }
        thr = worst_val<SMALL>();
        thr_id = kEmptyId;
    }

    __device__ __forceinline__ void refresh_thr() {
        // entry k-1 sits in slot (k-1)/32 of lane (k-1)%32.  Shuffle every slot first and choose
        // among the results: choosing the slot first (v[e >> 5]) would be a runtime index into the
        // register array and push the whole list into local memory.
        const int e = k - 1;
        float tv = __shfl_sync(0xffffffffu, v[0], e & 31);
        uint32_t ti = __shfl_sync(0xffffffffu, id[0], e & 31);
#pragma unroll
        for (int j = 1; j < KPL; j++) {
            const float sv = __shfl_sync(0xffffffffu, v[j], e & 31);
            const uint32_t si = __shfl_sync(0xffffffffu, id[j], e & 31);
            const bool pick = (e >> 5) == j;
            tv = pick ? sv : tv;
            ti = pick ? si : ti;
        }
        thr = tv;
        thr_id = ti;
    }

    // s/sid uniform across the warp.  Caller has checked that (s,sid) beats thr.
    __device__ __forceinline__ void insert(float s, uint32_t sid) {
        const int lane = threadIdx.x & 31;
        int pos = 0;
#pragma unroll
        for (int j = 0; j < KPL; j++) pos += __popc(__ballot_sync(0xffffffffu, better<SMALL>(v[j], id[j], s, sid)));
#pragma unroll
        for (int j = KPL - 1; j >= 0; j--) {
            float up_v = __shfl_up_sync(0xffffffffu, v[j], 1);
            uint32_t up_i = __shfl_up_sync(0xffffffffu, id[j], 1);
            if (j > 0) {
                float cv = __shfl_sync(0xffffffffu, v[j - 1], 31);
                uint32_t ci = __shfl_sync(0xffffffffu, id[j - 1], 31);
                if (lane == 0) {
                    up_v = cv;
                    up_i = ci;
                }
            }
            const int e = j * 32 + lane;
            if (e > pos) {
                v[j] = up_v;
                id[j] = up_i;
            } else if (e == pos) {
                v[j] = s;
                id[j] = sid;
            }
        }
        refresh_thr();
    }

    // value-only test for the streaming loop, where ids only ever increase
    __device__ __forceinline__ void offer_stream(float s, uint32_t sid) {
        if (better_val<SMALL>(s, thr)) insert(s, sid);
    }

    // each lane brings one candidate; survivors are inserted one by one
    __device__ __forceinline__ void offer_lanes(float cv, uint32_t cid) {
        unsigned m = __ballot_sync(0xffffffffu, cid != kEmptyId && better<SMALL>(cv, cid, thr, thr_id));
        while (m) {
            const int src = __ffs(m) - 1;
            m &= m - 1;
            const float s = __shfl_sync(0xffffffffu, cv, src);
            const uint32_t sid = __shfl_sync(0xffffffffu, cid, src);
            if (better<SMALL>(s, sid, thr, thr_id)) insert(s, sid);
        }
    }

    // ---- bitonic machinery -------------------------------------------------------------
    // Merge with another SORTED list of the same layout (bv/bi, best first): afterwards this
    // list holds the KPL*32 best of the union, sorted.  max(A[e], B[KCAP-1-e]) is a bitonic
    // sequence containing exactly those entries; log2(KCAP) half-cleaner stages sort it.
    __device__ __forceinline__ void merge_sorted(const float (&bv)[KPL], const uint32_t (&bi)[KPL]) {
        const int lane = threadIdx.x & 31;
#pragma unroll
        for (int j = 0; j < KPL; j++) {
            const float rv = __shfl_sync(0xffffffffu, bv[KPL - 1 - j], 31 - lane);
            const uint32_t ri = __shfl_sync(0xffffffffu, bi[KPL - 1 - j], 31 - lane);
            if (better<SMALL>(rv, ri, v[j], id[j])) {
                v[j] = rv;
                id[j] = ri;
            }
        }
#pragma unroll
        for (int s = KPL / 2; s >= 1; s >>= 1) {  // partner lives s slots away, same lane
#pragma unroll
            for (int j = 0; j < KPL; j++)
                if ((j & s) == 0) {
                    if (better<SMALL>(v[j + s], id[j + s], v[j], id[j])) {
                        const float tv = v[j];
                        const uint32_t ti = id[j];
                        v[j] = v[j + s];
                        id[j] = id[j + s];
                        v[j + s] = tv;
                        id[j + s] = ti;
                    }
                }
        }
#pragma unroll
        for (int s = 16; s >= 1; s >>= 1) {  // partner lives s lanes away, same slot
            const bool low = (lane & s) == 0;
#pragma unroll
            for (int j = 0; j < KPL; j++) {
                const float pv = __shfl_xor_sync(0xffffffffu, v[j], s);
                const uint32_t pi = __shfl_xor_sync(0xffffffffu, id[j], s);
                const bool p_better = better<SMALL>(pv, pi, v[j], id[j]);
                if (low == p_better) {  // the lower lane keeps the better one, the upper lane the worse one
                    v[j] = pv;
                    id[j] = pi;
                }
            }
        }
        refresh_thr();
    }

    // sort one candidate per lane, best first (bitonic network over the 32 lanes)
    __device__ __forceinline__ static void sort32(float& cv, uint32_t& ci) {
        const int lane = threadIdx.x & 31;
#pragma unroll
        for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
            for (int s = k >> 1; s >= 1; s >>= 1) {
                const float pv = __shfl_xor_sync(0xffffffffu, cv, s);
                const uint32_t pi = __shfl_xor_sync(0xffffffffu, ci, s);
                const bool low = (lane & s) == 0;
                const bool desc = (lane & k) == 0;  // this k-block is being sorted best-first
                const bool p_better = better<SMALL>(pv, pi, cv, ci);
                if ((low == desc) == p_better) {
                    cv = pv;
                    ci = pi;
                }
            }
        }
    }

    // each lane brings at most one candidate (kEmptyId = none): a few are inserted one by one,
    // a crowd is sorted and merged in one go
    __device__ __forceinline__ void offer_batch(float cv, uint32_t cid) {
        const bool pass = cid != kEmptyId && better<SMALL>(cv, cid, thr, thr_id);
        const unsigned m = __ballot_sync(0xffffffffu, pass);
        if (m == 0) return;
        if (__popc(m) <= 4) {
            offer_lanes(cv, cid);
            return;
        }
        if (!pass) {
            cv = worst_val<SMALL>();
            cid = kEmptyId;
        }
        sort32(cv, cid);
        float bv[KPL];
        uint32_t bi[KPL];
#pragma unroll
        for (int j = 0; j < KPL; j++) {
            bv[j] = j == 0 ? cv : worst_val<SMALL>();
            bi[j] = j == 0 ? cid : kEmptyId;
        }
        merge_sorted(bv, bi);
    }

    // merge `nlists` stored SORTED lists (each KPL*32 entries, `stride` entries apart); the next
    // list is fetched while the current one is being merged
    template <bool GLOBAL>
    __device__ __forceinline__ void merge_from(const Cand* lists, int first, int nlists, int step, size_t stride) {
        const int lane = threadIdx.x & 31;
        float bv[KPL];
        uint32_t bi[KPL];
        auto fetch = [&](int l) {
            const Cand* src = lists + (size_t)l * stride;
#pragma unroll
            for (int j = 0; j < KPL; j++) {
                if (GLOBAL) {
                    // written by other CTAs of this launch: read through L2
                    const unsigned long long raw =
                        __ldcg(reinterpret_cast<const unsigned long long*>(src + j * 32 + lane));
                    bv[j] = __uint_as_float((uint32_t)(raw & 0xffffffffull));
                    bi[j] = (uint32_t)(raw >> 32);
                } else {
                    const Cand c = src[j * 32 + lane];
                    bv[j] = c.v;
                    bi[j] = c.id;
                }
            }
        };
        if (GLOBAL && KPL == 1) {
            // short lists written by other CTAs: fetch a batch of them at once, so that one trip to L2
            // covers nine lists (148 CTA lists over 17 warps) instead of one
            constexpr int B = 9;
            for (int l0 = first; l0 < nlists; l0 += step * B) {
                float gv[B];
                uint32_t gi[B];
#pragma unroll
                for (int b = 0; b < B; b++) {
                    const int l = l0 + b * step;
                    gv[b] = worst_val<SMALL>();
                    gi[b] = kEmptyId;
                    if (l < nlists) {
                        const unsigned long long raw =
                            __ldcg(reinterpret_cast<const unsigned long long*>(lists + (size_t)l * stride + lane));
                        gv[b] = __uint_as_float((uint32_t)(raw & 0xffffffffull));
                        gi[b] = (uint32_t)(raw >> 32);
                    }
                }
#pragma unroll
                for (int b = 0; b < B; b++) {
                    const float head_v = __shfl_sync(0xffffffffu, gv[b], 0);
                    const uint32_t head_i = __shfl_sync(0xffffffffu, gi[b], 0);
                    if (head_i != kEmptyId && better<SMALL>(head_v, head_i, thr, thr_id)) {
                        float cv[KPL];
                        uint32_t ci[KPL];
                        cv[0] = gv[b];
                        ci[0] = gi[b];
                        merge_sorted(cv, ci);
                    }
                }
            }
            return;
        }
        if (first < nlists) fetch(first);
        for (int l = first; l < nlists; l += step) {
            float cv[KPL];
            uint32_t ci[KPL];
#pragma unroll
            for (int j = 0; j < KPL; j++) {
                cv[j] = bv[j];
                ci[j] = bi[j];
            }
            if (l + step < nlists) fetch(l + step);
            // a list whose best entry cannot enter changes nothing
            const float head_v = __shfl_sync(0xffffffffu, cv[0], 0);
            const uint32_t head_i = __shfl_sync(0xffffffffu, ci[0], 0);
            if (head_i != kEmptyId && better<SMALL>(head_v, head_i, thr, thr_id)) merge_sorted(cv, ci);
        }
    }

    // take over a stored sorted list (the inverse of store)
    __device__ __forceinline__ void load(const Cand* src, int k_) {
        const int lane = threadIdx.x & 31;
        k = k_;
#pragma unroll
        for (int j = 0; j < KPL; j++) {
            const Cand c = src[j * 32 + lane];
            v[j] = c.v;
            id[j] = c.id;
        }
        refresh_thr();
    }

    __device__ __forceinline__ void store(Cand* dst) const {
        const int lane = threadIdx.x & 31;
#pragma unroll
        for (int j = 0; j < KPL; j++) {
            Cand c;
            c.v = v[j];
            c.id = id[j];
            dst[j * 32 + lane] = c;
        }
    }
};

// Folds the n sorted lists of each of `nsets` independent sets (list l of set s at lists + s*set_stride + l*KPL*32,
// shared memory) into list 0 of the set by pairwise bitonic merges: ceil(log2 n) rounds with a CTA barrier after
// each, the pairs of a round spread over the warps -- instead of one warp folding the n lists one after the other
// (a merge is ~14 dependent shuffles: 16 of them in a row were most of a CTA's epilogue).  Every thread of the CTA
// must call it (barriers inside).
template <int KPL, bool SMALL>
__device__ __forceinline__ void tree_merge_lists(Cand* lists, int n, int nsets, size_t set_stride, int k) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    constexpr int KCAP = KPL * 32;
    for (int step = 1; step < n; step <<= 1) {
        const int pairs = (n + 2 * step - 1) / (2 * step);
        for (int task = warp; task < nsets * pairs; task += nwarps) {
            const int sidx = task / pairs, pi = task - sidx * pairs;
            const int left = pi * 2 * step, right = left + step;
            if (right < n) {
                Cand* a = lists + (size_t)sidx * set_stride + (size_t)left * KCAP;
                const Cand* b = lists + (size_t)sidx * set_stride + (size_t)right * KCAP;
                WarpTopK<KPL, SMALL> m;
                m.load(a, k);
                float bv[KPL];
                uint32_t bi[KPL];
#pragma unroll
                for (int j = 0; j < KPL; j++) {
                    const Cand c = b[j * 32 + lane];
                    bv[j] = c.v;
                    bi[j] = c.id;
                }
                // a list whose best entry cannot enter changes nothing
                const float head_v = __shfl_sync(0xffffffffu, bv[0], 0);
                const uint32_t head_i = __shfl_sync(0xffffffffu, bi[0], 0);
                if (head_i != kEmptyId && better<SMALL>(head_v, head_i, m.thr, m.thr_id)) {
                    m.merge_sorted(bv, bi);
                    m.store(a);
                }
            }
        }
        __syncthreads();
    }
}

}  // namespace b200r
