// K1 ingest (normalise / cast / row norms), query preparation, K5 compaction gather.
//
// Reference behaviour restated here (none of it is GPU code in the reference):
//  * faiss IndexFlatCodes::add -- append n*d fp32 values, no transform
//    (call sites: langchain FAISS.__add -> index.add(vector), reached from
//    RAG/examples/basic_rag/langchain/chains.py:83; direct:
//    community/chat-llama-nemotron/backend-rag/src/rag_service.py:197,450);
//  * faiss.normalize_L2 / fvec_renorm_L2 -- per row, s = sum x^2 in fp32; if s > 0 the
//    row is scaled by 1/sqrt(s); zero rows stay zero (rag_service.py:194,292,439);
//  * faiss IndexFlatCodes::remove_ids -- stable compaction (utils.py:559 via FAISS.delete).
// All three are pure HBM streaming: bytes = read + write of the rows touched.
#include "common.cuh"
#include "kernels.h"

namespace b200r {

constexpr int kRowWarps = 8;  // warps (= rows) per CTA

__device__ __forceinline__ float ld_src(const float* row, int j, int d_in) { return j < d_in ? row[j] : 0.f; }

// One warp per row.  VEC: row starts and d_in are multiples of 4 -> 128-bit accesses.
template <bool VEC, bool TO_BF16>
__global__ void __launch_bounds__(kRowWarps * 32) ingest_kernel(const float* src, int64_t n, int d_in, int d_pad,
                                                                 int normalize, void* dst, float* norms) {
    // no __restrict__: launch_normalize runs this in place (src == dst)
    const int lane = threadIdx.x & 31;
    const int64_t warps_total = (int64_t)gridDim.x * kRowWarps;
    for (int64_t row = (int64_t)blockIdx.x * kRowWarps + (threadIdx.x >> 5); row < n; row += warps_total) {
        const float* s = src + row * d_in;
        if (VEC && d_pad <= 32 * 8 * 4) {
            // rows of up to 1024 floats: the whole row lives in registers (8 x 128-bit per lane), so
            // normalising costs no second read and all eight loads are in flight at once
            const float4* s4 = reinterpret_cast<const float4*>(s);
            const int n4_in = d_in >> 2, n4_pad = d_pad >> 2;
            float4 v[8];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int i = u * 32 + lane;
                v[u] = i < n4_in ? s4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            float sc = 1.f;
            if (normalize) {
                float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
                for (int u = 0; u < 8; u++) {
                    a0 = fmaf(v[u].x, v[u].x, a0);
                    a1 = fmaf(v[u].y, v[u].y, a1);
                    a2 = fmaf(v[u].z, v[u].z, a2);
                    a3 = fmaf(v[u].w, v[u].w, a3);
                }
                const float nr = warp_sum((a0 + a2) + (a1 + a3));
                if (nr > 0.f) sc = 1.0f / sqrtf(nr);
            }
            float c0 = 0.f, c1 = 0.f, c2 = 0.f, c3 = 0.f;
#pragma unroll
            for (int u = 0; u < 8; u++) {
                const int i = u * 32 + lane;
                float4 w = v[u];
                w.x *= sc;
                w.y *= sc;
                w.z *= sc;
                w.w *= sc;
                if (TO_BF16) {
                    const __nv_bfloat162 lo = __floats2bfloat162_rn(w.x, w.y);
                    const __nv_bfloat162 hi = __floats2bfloat162_rn(w.z, w.w);
                    w.x = __low2float(lo);
                    w.y = __high2float(lo);
                    w.z = __low2float(hi);
                    w.w = __high2float(hi);
                    if (i < n4_pad) {
                        uint2 pk;
                        pk.x = *reinterpret_cast<const uint32_t*>(&lo);
                        pk.y = *reinterpret_cast<const uint32_t*>(&hi);
                        reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(dst) + row * d_pad)[i] = pk;
                    }
                } else if (i < n4_pad) {
                    reinterpret_cast<float4*>(reinterpret_cast<float*>(dst) + row * d_pad)[i] = w;
                }
                c0 = fmaf(w.x, w.x, c0);
                c1 = fmaf(w.y, w.y, c1);
                c2 = fmaf(w.z, w.z, c2);
                c3 = fmaf(w.w, w.w, c3);
            }
            const float nn1 = warp_sum((c0 + c2) + (c1 + c3));
            if (lane == 0 && norms) norms[row] = nn1;
            continue;
        }
        float scale = 1.f;
        if (normalize) {
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
            if (VEC) {
                const float4* s4 = reinterpret_cast<const float4*>(s);
                for (int i = lane; i < (d_in >> 2); i += 32) {
                    const float4 v = s4[i];
                    a0 = fmaf(v.x, v.x, a0);
                    a1 = fmaf(v.y, v.y, a1);
                    a2 = fmaf(v.z, v.z, a2);
                    a3 = fmaf(v.w, v.w, a3);
                }
            } else {
                for (int j = lane; j < d_in; j += 32) a0 = fmaf(s[j], s[j], a0);
            }
            const float nr = warp_sum((a0 + a2) + (a1 + a3));
            if (nr > 0.f) scale = 1.0f / sqrtf(nr);
        }
        // second pass (row is L1/L2 resident): scale, store, norm of what is stored
        float b0 = 0.f, b1 = 0.f, b2 = 0.f, b3 = 0.f;
        if (VEC) {
            const float4* s4 = reinterpret_cast<const float4*>(s);
            const int n4_in = d_in >> 2, n4_pad = d_pad >> 2;
            for (int i = lane; i < n4_pad; i += 32) {
                float4 v = i < n4_in ? s4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
                v.x *= scale;
                v.y *= scale;
                v.z *= scale;
                v.w *= scale;
                if (TO_BF16) {
                    const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y);
                    const __nv_bfloat162 hi = __floats2bfloat162_rn(v.z, v.w);
                    v.x = __low2float(lo);
                    v.y = __high2float(lo);
                    v.z = __low2float(hi);
                    v.w = __high2float(hi);
                    uint2 pk;
                    pk.x = *reinterpret_cast<const uint32_t*>(&lo);
                    pk.y = *reinterpret_cast<const uint32_t*>(&hi);
                    reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(dst) + row * d_pad)[i] = pk;
                } else {
                    reinterpret_cast<float4*>(reinterpret_cast<float*>(dst) + row * d_pad)[i] = v;
                }
                b0 = fmaf(v.x, v.x, b0);
                b1 = fmaf(v.y, v.y, b1);
                b2 = fmaf(v.z, v.z, b2);
                b3 = fmaf(v.w, v.w, b3);
            }
        } else {
            for (int j = lane; j < d_pad; j += 32) {
                float v = ld_src(s, j, d_in) * scale;
                if (TO_BF16) {
                    const __nv_bfloat16 h = __float2bfloat16_rn(v);
                    v = __bfloat162float(h);
                    (reinterpret_cast<__nv_bfloat16*>(dst) + row * d_pad)[j] = h;
                } else {
                    (reinterpret_cast<float*>(dst) + row * d_pad)[j] = v;
                }
                b0 = fmaf(v, v, b0);
            }
        }
        const float nn = warp_sum((b0 + b2) + (b1 + b3));
        if (lane == 0 && norms) norms[row] = nn;
    }
}

static int row_grid(int64_t n) {
    int64_t g = (n + kRowWarps - 1) / kRowWarps;
    const int64_t cap = 148 * 8;  // 8 CTAs of 256 threads per SM
    if (g > cap) g = cap;
    if (g < 1) g = 1;
    return (int)g;
}

cudaError_t launch_ingest(const float* src, int64_t n, int d_in, int d_pad, bool normalize, bool to_bf16,
                          void* dst_rows, float* dst_norms, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    const bool vec = (d_in % 4 == 0) && (d_pad % 4 == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
    const int g = row_grid(n);
    const int t = kRowWarps * 32;
    if (vec) {
        if (to_bf16)
            ingest_kernel<true, true><<<g, t, 0, st>>>(src, n, d_in, d_pad, normalize, dst_rows, dst_norms);
        else
            ingest_kernel<true, false><<<g, t, 0, st>>>(src, n, d_in, d_pad, normalize, dst_rows, dst_norms);
    } else {
        if (to_bf16)
            ingest_kernel<false, true><<<g, t, 0, st>>>(src, n, d_in, d_pad, normalize, dst_rows, dst_norms);
        else
            ingest_kernel<false, false><<<g, t, 0, st>>>(src, n, d_in, d_pad, normalize, dst_rows, dst_norms);
    }
    return cudaGetLastError();
}

cudaError_t launch_normalize(float* x, int64_t n, int d, cudaStream_t st) {
    // in place: src == dst, d_in == d_pad; rows are read fully before being rewritten
    // by the same warp, element by element at identical offsets
    return launch_ingest(x, n, d, d, true, false, x, nullptr, st);
}

// queries: pad to d_pad, zero the unused slots, optional normalise (one warp per slot)
__global__ void __launch_bounds__(kRowWarps * 32) prep_queries_kernel(const float* __restrict__ src, int64_t nq,
                                                                       int d_in, int d_pad, int64_t nq_slots,
                                                                       int normalize, float* __restrict__ dst) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * kRowWarps + (threadIdx.x >> 5);
    if (row >= nq_slots) return;
    float* o = dst + row * d_pad;
    if (row >= nq) {
        for (int j = lane; j < d_pad; j += 32) o[j] = 0.f;
        return;
    }
    const float* s = src + row * d_in;
    float scale = 1.f;
    if (normalize) {
        float a = 0.f;
        for (int j = lane; j < d_in; j += 32) a = fmaf(s[j], s[j], a);
        a = warp_sum(a);
        if (a > 0.f) scale = 1.0f / sqrtf(a);
    }
    for (int j = lane; j < d_pad; j += 32) o[j] = ld_src(s, j, d_in) * scale;
}

cudaError_t launch_prep_queries(const float* src, int64_t nq, int d_in, int d_pad, int64_t nq_slots, bool normalize,
                                float* dst, cudaStream_t st) {
    if (nq_slots <= 0) return cudaSuccess;
    const int g = (int)((nq_slots + kRowWarps - 1) / kRowWarps);
    prep_queries_kernel<<<g, kRowWarps * 32, 0, st>>>(src, nq, d_in, d_pad, nq_slots, normalize, dst);
    return cudaGetLastError();
}

// K5: out[o - o0] = src[o + #{i : del_adj[i] <= o}],  del_adj[i] = del[i] - i for the
// sorted, de-duplicated deleted positions.  One warp per row, 16-byte or 4-byte words.
__global__ void __launch_bounds__(kRowWarps * 32) gather_rows_kernel(const uint8_t* __restrict__ src,
                                                                      uint8_t* __restrict__ out, int64_t o0, int64_t o1,
                                                                      int row_bytes, const int64_t* __restrict__ del_adj,
                                                                      int64_t n_del) {
    const int lane = threadIdx.x & 31;
    const int64_t warps_total = (int64_t)gridDim.x * kRowWarps;
    for (int64_t o = o0 + (int64_t)blockIdx.x * kRowWarps + (threadIdx.x >> 5); o < o1; o += warps_total) {
        int64_t lo = 0, hi = n_del;  // first index with del_adj > o
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (del_adj[mid] <= o)
                lo = mid + 1;
            else
                hi = mid;
        }
        const uint8_t* s = src + (o + lo) * row_bytes;
        uint8_t* d = out + (o - o0) * row_bytes;
        if ((row_bytes & 15) == 0) {
            const uint4* s4 = reinterpret_cast<const uint4*>(s);
            uint4* d4 = reinterpret_cast<uint4*>(d);
            for (int i = lane; i < (row_bytes >> 4); i += 32) d4[i] = s4[i];
        } else {
            const uint32_t* s1 = reinterpret_cast<const uint32_t*>(s);
            uint32_t* d1 = reinterpret_cast<uint32_t*>(d);
            for (int i = lane; i < (row_bytes >> 2); i += 32) d1[i] = s1[i];
        }
    }
}

cudaError_t launch_gather_rows(const void* src, void* out, int64_t o0, int64_t o1, int row_bytes,
                               const int64_t* del_adj, int64_t n_del, cudaStream_t st) {
    if (o1 <= o0) return cudaSuccess;
    gather_rows_kernel<<<row_grid(o1 - o0), kRowWarps * 32, 0, st>>>(reinterpret_cast<const uint8_t*>(src),
                                                                     reinterpret_cast<uint8_t*>(out), o0, o1, row_bytes,
                                                                     del_adj, n_del);
    return cudaGetLastError();
}

__global__ void row_to_f32_kernel(const void* row, int d, int bf16, float* out) {
    for (int j = threadIdx.x; j < d; j += blockDim.x)
        out[j] = bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(row)[j])
                      : reinterpret_cast<const float*>(row)[j];
}

// rows[ids[i]] -> out[i][0..d) as fp32, one warp per requested row (ids already validated)
__global__ void __launch_bounds__(kRowWarps * 32) rows_to_f32_kernel(const uint8_t* __restrict__ rows, int row_bytes,
                                                                      int d, int bf16, const int64_t* __restrict__ ids,
                                                                      int64_t n, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * kRowWarps + (threadIdx.x >> 5);
    if (i >= n) return;
    const uint8_t* r = rows + (ids ? ids[i] : i) * (int64_t)row_bytes;  // no id list: rows 0..n-1 in order
    for (int j = lane; j < d; j += 32)
        out[i * d + j] = bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(r)[j])
                              : reinterpret_cast<const float*>(r)[j];
}

cudaError_t launch_rows_to_f32(const void* rows, int row_bytes, int d, bool bf16, const int64_t* ids, int64_t n,
                               float* out, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    const int g = (int)((n + kRowWarps - 1) / kRowWarps);
    rows_to_f32_kernel<<<g, kRowWarps * 32, 0, st>>>(reinterpret_cast<const uint8_t*>(rows), row_bytes, d, bf16 ? 1 : 0,
                                                     ids, n, out);
    return cudaGetLastError();
}

cudaError_t launch_row_to_f32(const void* row, int d, bool bf16, float* out, cudaStream_t st) {
    row_to_f32_kernel<<<1, 256, 0, st>>>(row, d, bf16 ? 1 : 0, out);
    return cudaGetLastError();
}

}  // namespace b200r
