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
                                                         float* __restrict__ out_d, int64_t* __restrict__ out_i) {
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

cudaError_t launch_merge_wait(bool l2, uint8_t* my_buf, int world, int nq, int k, uint32_t epoch, float* out_d,
                              int64_t* out_i, cudaStream_t st) {
    if (world < 1 || world > kXchgMaxWorld || nq < 1 || nq > kScanMaxNQ || k < 1 || k > kScanMaxK)
        return cudaErrorInvalidValue;
    if (k <= 32) {
        if (l2) merge_wait_kernel<1, true><<<1, 128, 0, st>>>(my_buf, world, nq, k, epoch, out_d, out_i);
        else merge_wait_kernel<1, false><<<1, 128, 0, st>>>(my_buf, world, nq, k, epoch, out_d, out_i);
    } else {
        if (l2) merge_wait_kernel<4, true><<<1, 128, 0, st>>>(my_buf, world, nq, k, epoch, out_d, out_i);
        else merge_wait_kernel<4, false><<<1, 128, 0, st>>>(my_buf, world, nq, k, epoch, out_d, out_i);
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
