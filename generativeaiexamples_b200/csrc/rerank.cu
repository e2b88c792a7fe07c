// The two steps LangChain runs on the HOST right after index.search, moved onto the GPU (SURVEY.md
// section 8, row f4):
//
//  * relevance + score_threshold of VectorStore.similarity_search_with_relevance_scores
//    (langchain_core/vectorstores.py: _euclidean / _cosine / _max_inner_product_relevance_score_fn,
//    then `score >= score_threshold`), reached in the reference through
//    as_retriever(search_type="similarity_score_threshold") at
//    RAG/examples/basic_rag/langchain/chains.py:156-163 and multi_turn_rag/chains.py:157-170;
//  * the greedy selection of maximal_marginal_relevance (langchain_community/vectorstores/utils.py),
//    reached through as_retriever(search_type="mmr").
//
// Both are tiny (k, fetch_k <= a few hundred) and latency-bound; what they save is the host's walk
// over hits that the threshold drops, and the fetch_k x d floats an MMR search would otherwise copy
// to the host.
#include "index.h"

namespace b200r {

// relevance kinds: numeric values are part of the C ABI (include/b200r.h)
//   0 EUCLIDEAN          rel = 1 - d / sqrt(2)     (d = faiss's squared L2; LangChain feeds it as is)
//   1 COSINE             rel = 1 - d
//   2 MAX_INNER_PRODUCT  rel = d > 0 ? 1 - d : -d
// Python evaluates these in double on the float32 score, so does this kernel.
__device__ __forceinline__ double relevance_of(int kind, float d) {
    const double x = (double)d;
    if (kind == 0) return 1.0 - x / 1.4142135623730951;
    if (kind == 1) return 1.0 - x;
    return x > 0.0 ? 1.0 - x : -1.0 * x;
}

// one warp per query: relevance of every hit; the passing hits go to the front (in order), the other
// valid hits follow them (in order; the host only looks at those when it has to word LangChain's
// "scores must be between 0 and 1" warning), the rest of the row is padding (id -1)
__global__ void relevance_filter_kernel(const float* __restrict__ d_in, const int64_t* __restrict__ i_in, int nq, int k,
                                        int kind, double threshold, int apply_threshold, double* __restrict__ r_out,
                                        float* __restrict__ d_out, int64_t* __restrict__ i_out,
                                        int32_t* __restrict__ counts, int32_t* __restrict__ out_of_range) {
    const int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (q >= nq) return;
    const size_t row = (size_t)q * k;
    int n_pass = 0, n_valid = 0, n_oor = 0;
    for (int j0 = 0; j0 < k; j0 += 32) {  // pass 1: how many pass
        const int j = j0 + lane;
        bool valid = false, pass = false;
        double r = 0.0;
        if (j < k && i_in[row + j] >= 0) {  // -1 pads a search with fewer than k rows
            valid = true;
            r = relevance_of(kind, d_in[row + j]);
            pass = !apply_threshold || r >= threshold;
        }
        n_pass += __popc(__ballot_sync(0xffffffffu, pass));
        n_valid += __popc(__ballot_sync(0xffffffffu, valid));
        n_oor += __popc(__ballot_sync(0xffffffffu, valid && (r < 0.0 || r > 1.0)));
    }
    int at_pass = 0, at_fail = n_pass;
    for (int j0 = 0; j0 < k; j0 += 32) {  // pass 2: stable partition
        const int j = j0 + lane;
        bool valid = false, pass = false;
        float dv = 0.f;
        int64_t iv = -1;
        double r = 0.0;
        if (j < k) {
            dv = d_in[row + j];
            iv = i_in[row + j];
            valid = iv >= 0;
            if (valid) {
                r = relevance_of(kind, dv);
                pass = !apply_threshold || r >= threshold;
            }
        }
        const unsigned pm = __ballot_sync(0xffffffffu, pass);
        const unsigned fm = __ballot_sync(0xffffffffu, valid && !pass);
        const unsigned below = (1u << lane) - 1u;
        if (valid) {
            const int at = pass ? at_pass + __popc(pm & below) : at_fail + __popc(fm & below);
            r_out[row + at] = r;
            d_out[row + at] = dv;
            i_out[row + at] = iv;
        }
        at_pass += __popc(pm);
        at_fail += __popc(fm);
    }
    for (int j = n_valid + lane; j < k; j += 32) {
        r_out[row + j] = 0.0;
        d_out[row + j] = 0.f;
        i_out[row + j] = -1;
    }
    if (lane == 0) {
        counts[q] = n_pass;
        out_of_range[q] = n_oor;  // LangChain warns when any relevance falls outside [0, 1]
    }
}

cudaError_t launch_relevance_filter(const float* d_in, const int64_t* i_in, int nq, int k, int kind, double threshold,
                                    bool apply_threshold, double* r_out, float* d_out, int64_t* i_out, int32_t* counts,
                                    int32_t* out_of_range, cudaStream_t st) {
    if (nq <= 0) return cudaSuccess;
    const int wpb = 4;
    relevance_filter_kernel<<<(nq + wpb - 1) / wpb, wpb * 32, 0, st>>>(d_in, i_in, nq, k, kind, threshold,
                                                                      apply_threshold ? 1 : 0, r_out, d_out, i_out, counts,
                                                                      out_of_range);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// maximal marginal relevance over m candidate rows (one CTA).
//   emb [m][d] fp32 (gathered rows), q [d]; LangChain's loop:
//     sim_q[i] = cos(q, x_i);  first pick = argmax sim_q
//     repeat: red[i] = max over picked j of cos(x_i, x_j)
//             score[i] = lambda * sim_q[i] - (1 - lambda) * red[i]   (double, like the Python floats)
//             pick the first i not yet picked with the largest score (strict >)
//   until min(k, m) are picked.  A NaN / inf similarity (zero vector) counts as 0, as in LangChain.
// ---------------------------------------------------------------------------
constexpr int kMmrThreads = 512;
constexpr int kMmrMaxM = 2048;  // 16 bytes of shared memory per candidate

__device__ __forceinline__ float clean_sim(float s) { return (isnan(s) || isinf(s)) ? 0.f : s; }

__global__ void __launch_bounds__(kMmrThreads) mmr_kernel(const float* __restrict__ emb, const float* __restrict__ q, int m,
                                                          int d, int k, double lambda_mult, int32_t* __restrict__ selected,
                                                          int32_t* __restrict__ n_selected) {
    extern __shared__ __align__(16) uint8_t mmr_smem[];
    float* s_norm = reinterpret_cast<float*>(mmr_smem);  // [m]
    float* s_simq = s_norm + m;                          // [m]
    float* s_red = s_simq + m;                           // [m] max similarity to anything picked so far
    int* s_taken = reinterpret_cast<int*>(s_red + m);    // [m]
    __shared__ double s_best[kMmrThreads / 32];
    __shared__ int s_besti[kMmrThreads / 32];
    __shared__ int s_pick;
    __shared__ float s_qnorm;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;

    if (warp == 0) {
        float a = 0.f;
        for (int j = lane; j < d; j += 32) a = fmaf(q[j], q[j], a);
        a = warp_sum(a);
        if (lane == 0) s_qnorm = sqrtf(a);
    }
    __syncthreads();
    for (int i = warp; i < m; i += nw) {
        const float* x = emb + (size_t)i * d;
        float a = 0.f, b = 0.f;
        for (int j = lane; j < d; j += 32) {
            const float v = x[j];
            a = fmaf(v, v, a);
            b = fmaf(v, q[j], b);
        }
        a = warp_sum(a);
        b = warp_sum(b);
        if (lane == 0) {
            const float nrm = sqrtf(a);
            s_norm[i] = nrm;
            s_simq[i] = clean_sim(b / (s_qnorm * nrm));
            s_red[i] = -FLT_MAX;
            s_taken[i] = 0;
        }
    }
    __syncthreads();
    const int want = k < m ? k : m;
    for (int round = 0; round < want; round++) {
        // argmax (first index among equals) of sim_q in round 0, of the MMR score afterwards
        double best = -1e300;
        int besti = 0x7fffffff;
        for (int i = threadIdx.x; i < m; i += blockDim.x) {
            if (s_taken[i]) continue;
            const double sc = round == 0 ? (double)s_simq[i]
                                         : lambda_mult * (double)s_simq[i] - (1.0 - lambda_mult) * (double)s_red[i];
            if (sc > best || (sc == best && i < besti)) {
                best = sc;
                besti = i;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oi = __shfl_xor_sync(0xffffffffu, besti, o);
            if (ob > best || (ob == best && oi < besti)) {
                best = ob;
                besti = oi;
            }
        }
        if (lane == 0) {
            s_best[warp] = best;
            s_besti[warp] = besti;
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            double b2 = s_best[0];
            int i2 = s_besti[0];
            for (int w = 1; w < nw; w++)
                if (s_best[w] > b2 || (s_best[w] == b2 && s_besti[w] < i2)) {
                    b2 = s_best[w];
                    i2 = s_besti[w];
                }
            s_pick = i2;
            selected[round] = i2;
            s_taken[i2] = 1;
        }
        __syncthreads();
        const int pick = s_pick;
        if (round + 1 == want) break;
        // similarity of every candidate to the new pick
        const float* xp = emb + (size_t)pick * d;
        for (int i = warp; i < m; i += nw) {
            const float* x = emb + (size_t)i * d;
            float b = 0.f;
            for (int j = lane; j < d; j += 32) b = fmaf(x[j], xp[j], b);
            b = warp_sum(b);
            if (lane == 0) {
                const float s = clean_sim(b / (s_norm[i] * s_norm[pick]));
                s_red[i] = fmaxf(s_red[i], s);
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) *n_selected = want;
}

cudaError_t launch_mmr(const float* emb, const float* q, int m, int d, int k, double lambda_mult, int32_t* selected,
                       int32_t* n_selected, cudaStream_t st) {
    if (m < 0 || m > kMmrMaxM || k < 1 || d < 1) return cudaErrorInvalidValue;
    if (m == 0) return cudaMemsetAsync(n_selected, 0, sizeof(int32_t), st);
    const size_t smem = (size_t)m * (3 * sizeof(float) + sizeof(int));
    mmr_kernel<<<1, kMmrThreads, smem, st>>>(emb, q, m, d, k, lambda_mult, selected, n_selected);
    return cudaGetLastError();
}

}  // namespace b200r
