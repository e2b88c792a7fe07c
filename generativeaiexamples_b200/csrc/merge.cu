// K4: final select over per-shard candidates (multi-GPU merge step).
//
// No counterpart in the reference (its FAISS index lives in one process on the CPU).
// After each rank's local search, the (score, id) candidates are exchanged with an
// NCCL allgather (torch.distributed, see generativeaiexamples_b200/sharded.py) into one
// block per rank, each [nq][k_in]; this kernel keeps the k_out best per query -- the same
// selection faiss's heap makes, because top-k of a union is top-k of the per-shard
// top-k's.  Ties go to the earlier list position, which equals the lowest global id
// for contiguous row-block sharding with rank-ordered lists.
#include "common.cuh"
#include "kernels.h"

namespace b200r {

template <int KPL, bool L2>
__global__ void __launch_bounds__(128) merge_topk_kernel(const float* __restrict__ cand_d,
                                                         const int64_t* __restrict__ cand_i, int64_t nq, int n_lists,
                                                         int k_in, int64_t stride_d, int64_t stride_i, int k_out,
                                                         float* __restrict__ out_d, int64_t* __restrict__ out_i) {
    const int lane = threadIdx.x & 31;
    const int64_t q = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (q >= nq) return;
    // candidate (list l, entry e) lives at cand_x[l * stride_x + q * k_in + e]; its
    // identity inside the kernel is pos = l * k_in + e
    const int64_t total = (int64_t)n_lists * k_in;
    WarpTopK<KPL, L2> m;
    m.init(k_out);
    for (int64_t base = 0; base < total; base += 32) {
        const int64_t pos = base + lane;
        float v = worst_val<L2>();
        uint32_t id = kEmptyId;
        if (pos < total) {
            const int64_t l = pos / k_in, e = pos - l * k_in;
            if (cand_i[l * stride_i + q * k_in + e] >= 0) {
                v = cand_d[l * stride_d + q * k_in + e];
                id = (uint32_t)pos;
            }
        }
        m.offer_lanes(v, id);
    }
#pragma unroll
    for (int j = 0; j < KPL; j++) {
        const int e = j * 32 + lane;
        if (e < k_out) {
            int64_t gid = -1;
            if (m.id[j] != kEmptyId) {
                const int64_t l = m.id[j] / k_in, ee = m.id[j] - l * k_in;
                gid = cand_i[l * stride_i + q * k_in + ee];
            }
            out_d[q * k_out + e] = m.v[j];
            out_i[q * k_out + e] = gid;
        }
    }
}

// Fused-exchange variant: first wait until every rank's scan kernel has delivered this epoch's
// result into the local exchange buffer (flags written with st.release.sys over NVLink), then run
// the same select over the `world` slots.  The wait is bounded: if a peer never shows up the status
// word is set and the output is padding, instead of a kernel that spins forever.
template <int KPL, bool L2>
__global__ void __launch_bounds__(128) merge_wait_kernel(uint8_t* buf, int world, int nq, int k, uint32_t epoch,
                                                         float* __restrict__ out_d, int64_t* __restrict__ out_i,
                                                         int* status_out) {
    __shared__ int s_ok;
    const uint32_t parity = epoch & 1u;
    if (threadIdx.x == 0) s_ok = 1;
    __syncthreads();
    if (threadIdx.x < world) {
        const uint32_t* flag =
            reinterpret_cast<const uint32_t*>(buf + kXchgFlagOffset + (parity * kXchgMaxWorld + threadIdx.x) * 128);
        const long long t_end = clock64() + 4000000000ll;  // ~2 s
        uint32_t seen;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(flag) : "memory");
        } while (seen != epoch && clock64() < t_end);
        if (seen != epoch) {
            s_ok = 0;
            *reinterpret_cast<int*>(buf + kXchgStatusOffset) = 1 + (int)threadIdx.x;
        }
    }
    __syncthreads();
    if (status_out && threadIdx.x == 0) *status_out = s_ok ? 0 : *reinterpret_cast<int*>(buf + kXchgStatusOffset);
    const int lane = threadIdx.x & 31;
    const int q = threadIdx.x >> 5;
    if (q >= nq) return;
    const uint8_t* base = buf + (size_t)parity * kXchgParityBytes;
    const int total = world * k;
    WarpTopK<KPL, L2> m;
    m.init(k);
    if (s_ok) {
        for (int b0 = 0; b0 < total; b0 += 32) {
            const int pos = b0 + lane;
            float v = worst_val<L2>();
            uint32_t id = kEmptyId;
            if (pos < total) {
                const int l = pos / k, e = pos - l * k;
                const uint8_t* slot = base + (size_t)l * kXchgSlotBytes;
                if (reinterpret_cast<const int64_t*>(slot + kXchgIdOffset)[q * k + e] >= 0) {
                    v = reinterpret_cast<const float*>(slot)[q * k + e];
                    id = (uint32_t)pos;
                }
            }
            m.offer_lanes(v, id);
        }
    }
#pragma unroll
    for (int j = 0; j < KPL; j++) {
        const int e = j * 32 + lane;
        if (e < k) {
            int64_t gid = -1;
            if (m.id[j] != kEmptyId) {
                const int l = m.id[j] / k, ee = m.id[j] - l * k;
                gid = reinterpret_cast<const int64_t*>(base + (size_t)l * kXchgSlotBytes + kXchgIdOffset)[q * k + ee];
            }
            out_d[q * k + e] = m.v[j];
            out_i[q * k + e] = gid;
        }
    }
}

// ---------------------------------------------------------------------------
// Final select of the multi-GPU index that lives in ONE process (multi.cu).
//
// Each of the n_lists shards delivers its local top-k for every query -- sorted best first, ties by
// the lower LOCAL row number, padding (id -1) at the end -- into its slot of a gather buffer on this
// GPU.  A shard's rows are a set of runs ("segments") of the global row order, in order, so a local
// row number maps to its global position (the faiss id) through the shard's segment table, and the
// map is monotone: every list is also sorted by (score, global id).  The rank of a candidate in the
// merged order is therefore the sum over the lists of "entries better than it", each found by a
// binary search; candidates with rank < k_out go straight to their output position.  64-bit ids
// and exact lowest-global-id tie breaking come for free; no list is ever re-sorted.
//
//   slot l:  scores fp32 [nq][k_in] at base + l*slot_stride,  ids int64 [nq][k_in] at + ids_off
//   seg_off [n_lists+1], seg_lstart / seg_gbase [seg_off[n_lists]]: segments of shard l, by lstart
//   wait_epoch != 0: the slots are filled by the scan kernels' final writers (fused exchange); wait
//   for the n_lists flags of this epoch first (bounded, ~2 s; a miss is reported through status_out
//   and yields padding).
// ---------------------------------------------------------------------------
struct MergeShardsParams {
    const uint8_t* base;
    int64_t slot_stride, ids_off;
    int n_lists, k_in, k_out, nq;
    const int* seg_off;
    const int64_t* seg_lstart;
    const int64_t* seg_gbase;
    int64_t id_base;
    float* out_d;
    int64_t* out_i;
    uint8_t* xbuf;        // exchange buffer holding the flags (fused exchange only)
    uint32_t wait_epoch;
    int* status_out;      // optional: 0, or 1 + the first shard whose slot never arrived
};

template <bool L2>
__global__ void __launch_bounds__(256) merge_shards_kernel(const MergeShardsParams p) {
    extern __shared__ __align__(16) uint8_t ms_smem[];
    const int total = p.n_lists * p.k_in;
    int64_t* s_gid = reinterpret_cast<int64_t*>(ms_smem);        // [n_lists][k_in] global id, -1 = padding
    float* s_val = reinterpret_cast<float*>(s_gid + total);      // [n_lists][k_in]
    __shared__ int s_valid[8];  // valid entries per list
    __shared__ int s_ok;
    const int q = blockIdx.x;
    if (threadIdx.x == 0) s_ok = 1;
    if (threadIdx.x < 8) s_valid[threadIdx.x] = 0;
    __syncthreads();
    if (p.wait_epoch != 0 && threadIdx.x < p.n_lists) {
        const uint32_t* flag = reinterpret_cast<const uint32_t*>(
            p.xbuf + kXchgFlagOffset + ((p.wait_epoch & 1u) * kXchgMaxWorld + threadIdx.x) * 128);
        const long long t_end = clock64() + 4000000000ll;  // ~2 s
        uint32_t seen;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(flag) : "memory");
        } while (seen != p.wait_epoch && clock64() < t_end);
        if (seen != p.wait_epoch) atomicMin(&s_ok, -(int)threadIdx.x);  // 0, -1, ...: the first missing shard
    }
    __syncthreads();
    const bool ok = s_ok > 0;
    if (p.status_out && q == 0 && threadIdx.x == 0) *p.status_out = ok ? 0 : 1 - s_ok;
    // phase 1: translate every candidate's local row number to its global position
    for (int c = threadIdx.x; c < total; c += blockDim.x) {
        const int l = c / p.k_in, e = c - l * p.k_in;
        const uint8_t* slot = p.base + (int64_t)l * p.slot_stride;
        int64_t lid = -1;
        float v = worst_val<L2>();
        if (ok) {
            // written by other GPUs / other kernels: read through L2
            lid = __ldcg(reinterpret_cast<const int64_t*>(slot + p.ids_off) + (int64_t)q * p.k_in + e);
            v = __ldcg(reinterpret_cast<const float*>(slot) + (int64_t)q * p.k_in + e);
        }
        int64_t gid = -1;
        if (lid >= 0) {
            int lo = p.seg_off[l], hi = p.seg_off[l + 1];  // last segment with lstart <= lid
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (p.seg_lstart[mid] <= lid)
                    lo = mid;
                else
                    hi = mid;
            }
            gid = p.seg_gbase[lo] + (lid - p.seg_lstart[lo]);
            atomicAdd(&s_valid[l], 1);
        }
        s_gid[c] = gid;
        s_val[c] = v;
    }
    __syncthreads();
    int n_valid = 0;
    for (int l = 0; l < p.n_lists; l++) n_valid += s_valid[l];
    // phase 2: rank = number of candidates that beat this one (score, then lower global id)
    for (int c = threadIdx.x; c < total; c += blockDim.x) {
        const int64_t gid = s_gid[c];
        if (gid < 0) continue;
        const float v = s_val[c];
        const int l = c / p.k_in;
        int rank = c - l * p.k_in;  // its own list: everything before it
        for (int l2 = 0; l2 < p.n_lists; l2++) {
            if (l2 == l) continue;
            const int64_t* g2 = s_gid + l2 * p.k_in;
            const float* v2 = s_val + l2 * p.k_in;
            int lo = 0, hi = s_valid[l2];  // first entry of list l2 that does NOT beat (v, gid)
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                const float mv = v2[mid];
                const bool beats = better_val<L2>(mv, v) || (mv == v && g2[mid] < gid);
                if (beats)
                    lo = mid + 1;
                else
                    hi = mid;
            }
            rank += lo;
        }
        if (rank < p.k_out) {
            p.out_d[(int64_t)q * p.k_out + rank] = v;
            p.out_i[(int64_t)q * p.k_out + rank] = gid + p.id_base;
        }
    }
    for (int e = n_valid + threadIdx.x; e < p.k_out; e += blockDim.x) {  // fewer than k rows in the whole index
        p.out_d[(int64_t)q * p.k_out + e] = worst_val<L2>();
        p.out_i[(int64_t)q * p.k_out + e] = -1;
    }
}

cudaError_t launch_merge_shards(bool l2, const uint8_t* base, int64_t slot_stride, int64_t ids_off, int n_lists, int nq,
                                int k_in, int k_out, const int* seg_off, const int64_t* seg_lstart,
                                const int64_t* seg_gbase, int64_t id_base, float* out_d, int64_t* out_i, uint8_t* xbuf,
                                uint32_t wait_epoch, int* status_out, cudaStream_t st) {
    if (nq <= 0) return cudaSuccess;
    if (n_lists < 1 || n_lists > 8 || k_in < 1 || k_out < 1 || k_out > kMaxK || k_in > kMaxK) return cudaErrorInvalidValue;
    MergeShardsParams p{};
    p.base = base;
    p.slot_stride = slot_stride;
    p.ids_off = ids_off;
    p.n_lists = n_lists;
    p.k_in = k_in;
    p.k_out = k_out;
    p.nq = nq;
    p.seg_off = seg_off;
    p.seg_lstart = seg_lstart;
    p.seg_gbase = seg_gbase;
    p.id_base = id_base;
    p.out_d = out_d;
    p.out_i = out_i;
    p.xbuf = xbuf;
    p.wait_epoch = wait_epoch;
    p.status_out = status_out;
    const size_t smem = (size_t)n_lists * k_in * (sizeof(int64_t) + sizeof(float));
    if (l2) {
        if (smem > 48 * 1024) {
            cudaError_t e = cudaFuncSetAttribute(merge_shards_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
        }
        merge_shards_kernel<true><<<nq, 256, smem, st>>>(p);
    } else {
        if (smem > 48 * 1024) {
            cudaError_t e = cudaFuncSetAttribute(merge_shards_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
        }
        merge_shards_kernel<false><<<nq, 256, smem, st>>>(p);
    }
    return cudaGetLastError();
}

cudaError_t launch_merge_wait(bool l2, uint8_t* my_buf, int world, int nq, int k, uint32_t epoch, float* out_d,
                              int64_t* out_i, int* status_out, cudaStream_t st) {
    if (world < 1 || world > kXchgMaxWorld || nq < 1 || nq > kScanMaxNQ || k < 1 || k > kScanMaxK)
        return cudaErrorInvalidValue;
    if (k <= 32) {
        if (l2) merge_wait_kernel<1, true><<<1, 128, 0, st>>>(my_buf, world, nq, k, epoch, out_d, out_i, status_out);
        else merge_wait_kernel<1, false><<<1, 128, 0, st>>>(my_buf, world, nq, k, epoch, out_d, out_i, status_out);
    } else {
        if (l2) merge_wait_kernel<4, true><<<1, 128, 0, st>>>(my_buf, world, nq, k, epoch, out_d, out_i, status_out);
        else merge_wait_kernel<4, false><<<1, 128, 0, st>>>(my_buf, world, nq, k, epoch, out_d, out_i, status_out);
    }
    return cudaGetLastError();
}

cudaError_t launch_merge(bool l2, const float* cand_d, const int64_t* cand_i, int64_t nq, int n_lists, int k_in,
                         int64_t stride_d, int64_t stride_i, int k_out, float* out_d, int64_t* out_i,
                         cudaStream_t st) {
    if (nq <= 0) return cudaSuccess;
    if (k_out < 1 || k_out > kMaxK || n_lists < 1 || k_in < 1) return cudaErrorInvalidValue;
    if ((int64_t)n_lists * k_in >= 0xFFFFFFFFll) return cudaErrorInvalidValue;
    const int wpb = 4;
    const int g = (int)((nq + wpb - 1) / wpb);
#define B200R_MERGE(KPLV)                                                                                        \
    do {                                                                                                         \
        if (l2)                                                                                                  \
            merge_topk_kernel<KPLV, true><<<g, wpb * 32, 0, st>>>(cand_d, cand_i, nq, n_lists, k_in, stride_d,    \
                                                                  stride_i, k_out, out_d, out_i);                \
        else                                                                                                     \
            merge_topk_kernel<KPLV, false><<<g, wpb * 32, 0, st>>>(cand_d, cand_i, nq, n_lists, k_in, stride_d,   \
                                                                   stride_i, k_out, out_d, out_i);               \
    } while (0)
    if (k_out <= 32)
        B200R_MERGE(1);
    else if (k_out <= 128)
        B200R_MERGE(4);
    else if (k_out <= 256)
        B200R_MERGE(8);
    else if (k_out <= 512)
        B200R_MERGE(16);
    else
        B200R_MERGE(32);
#undef B200R_MERGE
    return cudaGetLastError();
}

}  // namespace b200r
