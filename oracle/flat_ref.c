/*
 * oracle/flat_ref.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement (T1, "fp32-faithful") of the arithmetic the reference's hot
 * path runs inside faiss-cpu==1.7.4 (pinned at
 * /root/reference/RAG/src/chain_server/requirements.txt:25), reached from
 *   RAG/src/chain_server/utils.py:306          FAISS(..., IndexFlatL2(dim), ...)
 *   RAG/examples/basic_rag/langchain/chains.py:83,156-167   add / search
 *   community/chat-llama-nemotron/backend-rag/src/rag_service.py:194,197,292,301
 *
 * faiss itself is NOT vendored under /root/reference and is not installable in
 * this image (no wheel, no network), and the reference holds no golden vector
 * or test for this path.  PARITY UNPINNED: this file
 * restates the published faiss 1.7.4 algorithm from its documented behaviour:
 *   faiss/utils/distances.cpp   knn_L2sqr / knn_inner_product,
 *                               exhaustive_*_seq (nq < 20) and
 *                               exhaustive_*_blas (nq >= 20, blocks 4096x1024,
 *                               ||q||^2 + ||x||^2 - 2 q.x clamped at 0),
 *                               fvec_renorm_L2
 *   faiss/utils/Heap.h          binary heap top-k, strict-improvement replace,
 *                               final reorder, -1 / +-FLT_MAX padding
 *   faiss/IndexFlatCodes.cpp    add (append), remove_ids (stable compaction),
 *                               reconstruct
 * It is validated against the float64 brute force in oracle/f64_truth.py and the
 * hand-computed cases in tests/golden/.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.
 *
 * Build: see oracle/Makefile  (gcc -O3 -march=native -fopenmp -shared -fPIC).
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_METRIC_IP 0 /* faiss METRIC_INNER_PRODUCT */
#define ORACLE_METRIC_L2 1 /* faiss METRIC_L2 (squared) */

/* faiss distance_compute_blas_threshold (distances.cpp) */
#define BLAS_THRESHOLD 20
/* faiss distance_compute_blas_query_bs / database_bs */
#define BLAS_QUERY_BS 4096
#define BLAS_DB_BS 1024

/* ------------------------------------------------------------------ */
/* optional external sgemm (OpenBLAS bundled with numpy, ILP64 cblas) */
/* ------------------------------------------------------------------ */
typedef void (*cblas_sgemm64_fn)(int order, int transa, int transb, int64_t m,
                                 int64_t n, int64_t k, float alpha,
                                 const float* a, int64_t lda, const float* b,
                                 int64_t ldb, float beta, float* c, int64_t ldc);
static cblas_sgemm64_fn g_sgemm = 0;

void oracle_set_sgemm(void* fn) { g_sgemm = (cblas_sgemm64_fn)fn; }
int oracle_has_blas(void) { return g_sgemm != 0; }

int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* Size of the OpenMP team used from now on.  A launcher may export
 * OMP_NUM_THREADS=1 (torch.distributed.run does), which would silently turn
 * the "all cores" baseline into one core; bench.py states the team instead. */
void oracle_set_threads(int n) {
#ifdef _OPENMP
    if (n >= 1) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* ------------------------------------------------------------------ */
/* elementary distances: faiss fvec_inner_product / fvec_L2sqr /       */
/* fvec_norm_L2sqr.  8 partial sums, as the 8-lane SIMD versions keep. */
/* ------------------------------------------------------------------ */
float oracle_fvec_inner_product(const float* x, const float* y, int64_t d) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int64_t i = 0;
    for (; i + 8 <= d; i += 8)
        for (int l = 0; l < 8; l++) acc[l] += x[i + l] * y[i + l];
    float s = ((acc[0] + acc[4]) + (acc[2] + acc[6])) +
              ((acc[1] + acc[5]) + (acc[3] + acc[7]));
    for (; i < d; i++) s += x[i] * y[i];
    return s;
}

float oracle_fvec_L2sqr(const float* x, const float* y, int64_t d) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int64_t i = 0;
    for (; i + 8 <= d; i += 8)
        for (int l = 0; l < 8; l++) {
            float t = x[i + l] - y[i + l];
            acc[l] += t * t;
        }
    float s = ((acc[0] + acc[4]) + (acc[2] + acc[6])) +
              ((acc[1] + acc[5]) + (acc[3] + acc[7]));
    for (; i < d; i++) {
        float t = x[i] - y[i];
        s += t * t;
    }
    return s;
}

float oracle_fvec_norm_L2sqr(const float* x, int64_t d) {
    return oracle_fvec_inner_product(x, x, d);
}

/* faiss fvec_renorm_L2: rows with zero norm are left untouched. */
void oracle_renorm_L2(int64_t d, int64_t nx, float* x) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < nx; i++) {
        float* xi = x + i * d;
        float nr = oracle_fvec_norm_L2sqr(xi, d);
        if (nr > 0) {
            const float inv_nr = 1.0f / sqrtf(nr);
            for (int64_t j = 0; j < d; j++) xi[j] *= inv_nr;
        }
    }
}

/* ------------------------------------------------------------------ */
/* top-k heap (faiss Heap.h).  "worse" is the order the heap root keeps */
/* on top: for L2 the largest distance, for IP the smallest product.    */
/* Ties on the value are ordered by id so the result is deterministic.  */
/* ------------------------------------------------------------------ */
typedef struct {
    float* val;
    int64_t* ids;
    int64_t k;
    int keep_small; /* 1: L2 (keep smallest), 0: IP (keep largest) */
} heap_t;

/* a is "worse or equal" than b => a belongs nearer the root */
static inline int worse(const heap_t* h, float va, int64_t ia, float vb,
                        int64_t ib) {
    if (va != vb) return h->keep_small ? (va > vb) : (va < vb);
    return ia > ib;
}

static void heap_init(heap_t* h) {
    for (int64_t i = 0; i < h->k; i++) {
        h->val[i] = h->keep_small ? FLT_MAX : -FLT_MAX;
        h->ids[i] = -1;
    }
}

static void heap_sift_down(heap_t* h, int64_t n, int64_t i, float v,
                           int64_t id) {
    for (;;) {
        int64_t l = 2 * i + 1, r = l + 1, c;
        if (l >= n) break;
        c = (r < n && worse(h, h->val[r], h->ids[r], h->val[l], h->ids[l])) ? r
                                                                            : l;
        if (worse(h, v, id, h->val[c], h->ids[c])) break;
        h->val[i] = h->val[c];
        h->ids[i] = h->ids[c];
        i = c;
    }
    h->val[i] = v;
    h->ids[i] = id;
}

/* replace only on STRICT improvement over the current k-th (faiss
 * HeapResultHandler::add_result: if (C::cmp(heap_dis_tab[0], dis))). */
static inline void heap_offer(heap_t* h, float v, int64_t id) {
    float top = h->val[0];
    int better = h->keep_small ? (v < top) : (v > top);
    if (better) heap_sift_down(h, h->k, 0, v, id);
}

/* faiss heap_reorder: sort best-first in place; padding (-1) goes last. */
static void heap_reorder(heap_t* h) {
    int64_t n = h->k;
    /* pop the root repeatedly into the tail => tail holds worst..., i.e.
     * array ends best-first after the loop */
    while (n > 1) {
        float v = h->val[0];
        int64_t id = h->ids[0];
        float lv = h->val[n - 1];
        int64_t li = h->ids[n - 1];
        n--;
        heap_sift_down(h, n, 0, lv, li);
        h->val[n] = v;
        h->ids[n] = id;
    }
    /* array is now worst-last?  popping the worst first places the worst at
     * the end: best-first order achieved. padding entries (id -1) compare as
     * the worst value and, among equals, id -1 < any id: push them last. */
    int64_t w = 0;
    for (int64_t i = 0; i < h->k; i++)
        if (h->ids[i] >= 0) {
            if (w != i) {
                float tv = h->val[i];
                int64_t ti = h->ids[i];
                for (int64_t j = i; j > w; j--) {
                    h->val[j] = h->val[j - 1];
                    h->ids[j] = h->ids[j - 1];
                }
                h->val[w] = tv;
                h->ids[w] = ti;
            }
            w++;
        }
}

/* ------------------------------------------------------------------ */
/* fallback sgemm: C[m][n] = A[m][kk] . B[n][kk]^T  (row-major, both    */
/* operands K-contiguous, as faiss calls sgemm_("T","N",...)).          */
/* ------------------------------------------------------------------ */
static void sgemm_nt_fallback(int64_t m, int64_t n, int64_t kk, const float* a,
                              const float* b, float* c, int64_t ldc) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < m; i++)
        for (int64_t j = 0; j < n; j++)
            c[i * ldc + j] = oracle_fvec_inner_product(a + i * kk, b + j * kk, kk);
}

static void sgemm_nt(int64_t m, int64_t n, int64_t kk, const float* a,
                     const float* b, float* c, int64_t ldc) {
    if (g_sgemm) /* CblasRowMajor=101, NoTrans=111, Trans=112 */
        g_sgemm(101, 111, 112, m, n, kk, 1.0f, a, kk, b, kk, 0.0f, c, ldc);
    else
        sgemm_nt_fallback(m, n, kk, a, b, c, ldc);
}

/* ------------------------------------------------------------------ */
/* exhaustive_*_seq: nq < 20.  OpenMP over QUERIES only, so nq == 1 is  */
/* one thread streaming the whole corpus (faiss 1.7.4 behaviour).       */
/* ------------------------------------------------------------------ */
static void knn_seq(int metric, const float* x, int64_t nq, const float* y,
                    int64_t ny, int64_t d, int64_t k, float* D, int64_t* I) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < nq; i++) {
        heap_t h = {D + i * k, I + i * k, k, metric == ORACLE_METRIC_L2};
        heap_init(&h);
        const float* xi = x + i * d;
        if (metric == ORACLE_METRIC_L2) {
            for (int64_t j = 0; j < ny; j++)
                heap_offer(&h, oracle_fvec_L2sqr(xi, y + j * d, d), j);
        } else {
            for (int64_t j = 0; j < ny; j++)
                heap_offer(&h, oracle_fvec_inner_product(xi, y + j * d, d), j);
        }
        heap_reorder(&h);
    }
}

/* ------------------------------------------------------------------ */
/* exhaustive_*_blas: nq >= 20.                                         */
/* ------------------------------------------------------------------ */
static void knn_blas(int metric, const float* x, int64_t nq, const float* y,
                     int64_t ny, int64_t d, int64_t k, float* D, int64_t* I) {
    if (nq == 0) return;
    const int l2 = metric == ORACLE_METRIC_L2;
    float* ip = (float*)malloc(sizeof(float) * BLAS_QUERY_BS * BLAS_DB_BS);
    float* xn = 0;
    float* yn = 0;
    if (l2) {
        xn = (float*)malloc(sizeof(float) * (nq > 0 ? nq : 1));
        yn = (float*)malloc(sizeof(float) * (ny > 0 ? ny : 1));
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < nq; i++)
            xn[i] = oracle_fvec_norm_L2sqr(x + i * d, d);
#pragma omp parallel for schedule(static)
        for (int64_t j = 0; j < ny; j++)
            yn[j] = oracle_fvec_norm_L2sqr(y + j * d, d);
    }
    for (int64_t i = 0; i < nq; i++) {
        heap_t h = {D + i * k, I + i * k, k, l2};
        heap_init(&h);
    }
    for (int64_t i0 = 0; i0 < nq; i0 += BLAS_QUERY_BS) {
        int64_t i1 = i0 + BLAS_QUERY_BS < nq ? i0 + BLAS_QUERY_BS : nq;
        for (int64_t j0 = 0; j0 < ny; j0 += BLAS_DB_BS) {
            int64_t j1 = j0 + BLAS_DB_BS < ny ? j0 + BLAS_DB_BS : ny;
            int64_t nyi = j1 - j0;
            sgemm_nt(i1 - i0, nyi, d, x + i0 * d, y + j0 * d, ip, nyi);
#pragma omp parallel for schedule(static)
            for (int64_t i = i0; i < i1; i++) {
                heap_t h = {D + i * k, I + i * k, k, l2};
                const float* line = ip + (i - i0) * nyi;
                if (l2) {
                    for (int64_t j = j0; j < j1; j++) {
                        float dis = xn[i] + yn[j] - 2 * line[j - j0];
                        if (dis < 0) dis = 0; /* negative: rounding error */
                        heap_offer(&h, dis, j);
                    }
                } else {
                    for (int64_t j = j0; j < j1; j++)
                        heap_offer(&h, line[j - j0], j);
                }
            }
        }
    }
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < nq; i++) {
        heap_t h = {D + i * k, I + i * k, k, l2};
        heap_reorder(&h);
    }
    free(ip);
    free(xn);
    free(yn);
}

/* IndexFlat::search -> knn_L2sqr / knn_inner_product.
 * path: 0 = faithful dispatch (nq < 20 seq, else blas), 1 = force seq,
 *       2 = force blas. Returns 0, or -1 on bad arguments (the python
 *       wrapper asserts k > 0 and d == index.d before calling). */
int oracle_knn(int metric, const float* x, int64_t nq, const float* y,
               int64_t ny, int64_t d, int64_t k, float* D, int64_t* I,
               int path) {
    if (k <= 0 || d <= 0 || nq < 0 || ny < 0) return -1;
    if (metric != ORACLE_METRIC_IP && metric != ORACLE_METRIC_L2) return -1;
    int use_blas = path == 2 || (path == 0 && nq >= BLAS_THRESHOLD);
    if (use_blas)
        knn_blas(metric, x, nq, y, ny, d, k, D, I);
    else
        knn_seq(metric, x, nq, y, ny, d, k, D, I);
    return 0;
}

/* Variant (ii) of the CPU baseline: the same scan,
 * row-parallel over all cores for small nq, so a GPU/CPU ratio is not
 * inflated by faiss 1.7.4's single-threaded nq == 1 path.  Not a faiss code
 * path; results are identical to knn_seq up to tie order. */
int oracle_knn_allcores(int metric, const float* x, int64_t nq, const float* y,
                        int64_t ny, int64_t d, int64_t k, float* D,
                        int64_t* I) {
    if (k <= 0 || d <= 0 || nq < 0 || ny < 0) return -1;
    const int l2 = metric == ORACLE_METRIC_L2;
    int nt = oracle_max_threads();
    float* pd = (float*)malloc(sizeof(float) * nt * nq * k);
    int64_t* pi = (int64_t*)malloc(sizeof(int64_t) * nt * nq * k);
#pragma omp parallel num_threads(nt)
    {
#ifdef _OPENMP
        int t = omp_get_thread_num();
#else
        int t = 0;
#endif
        int64_t lo = ny * t / nt, hi = ny * (t + 1) / nt;
        for (int64_t i = 0; i < nq; i++) {
            heap_t h = {pd + (t * nq + i) * k, pi + (t * nq + i) * k, k, l2};
            heap_init(&h);
            const float* xi = x + i * d;
            for (int64_t j = lo; j < hi; j++) {
                float s = l2 ? oracle_fvec_L2sqr(xi, y + j * d, d)
                             : oracle_fvec_inner_product(xi, y + j * d, d);
                heap_offer(&h, s, j);
            }
        }
    }
    for (int64_t i = 0; i < nq; i++) {
        heap_t h = {D + i * k, I + i * k, k, l2};
        heap_init(&h);
        for (int t = 0; t < nt; t++)
            for (int64_t r = 0; r < k; r++) {
                int64_t id = pi[(t * nq + i) * k + r];
                if (id >= 0) heap_offer(&h, pd[(t * nq + i) * k + r], id);
            }
        heap_reorder(&h);
    }
    free(pd);
    free(pi);
    return 0;
}

/* IndexFlatCodes::remove_ids with an IDSelectorBatch: stable compaction of
 * the rows whose position is NOT in ids; duplicates and out-of-range ids are
 * ignored.  Returns the number of rows removed; *ntotal is updated. */
int64_t oracle_remove_ids(float* codes, int64_t* ntotal, int64_t d,
                          const int64_t* ids, int64_t n) {
    int64_t nt = *ntotal;
    unsigned char* del = (unsigned char*)calloc((size_t)(nt > 0 ? nt : 1), 1);
    for (int64_t i = 0; i < n; i++)
        if (ids[i] >= 0 && ids[i] < nt) del[ids[i]] = 1;
    int64_t j = 0;
    for (int64_t i = 0; i < nt; i++) {
        if (del[i]) continue;
        if (i > j) memmove(codes + j * d, codes + i * d, sizeof(float) * d);
        j++;
    }
    free(del);
    *ntotal = j;
    return nt - j;
}
