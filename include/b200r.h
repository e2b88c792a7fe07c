/*
 * b200r.h -- C ABI of libb200r.so, the B200-native exact flat-search engine.
 *
 * This is the drop-in boundary for the vector-store search step of the reference's
 * RAG chain server.  The reference reaches that step through faiss-cpu 1.7.4's SWIG
 * wrapper (faiss/python/class_wrappers.py), from LangChain's FAISS vector store, which
 * RAG/src/chain_server/utils.py:305-306 constructs:
 *
 *     FAISS(document_embedder, IndexFlatL2(config.embeddings.dimensions), InMemoryDocstore(), {})
 *
 * Every entry point below names the reference-side call it replaces.  All pointers are
 * plain C; no torch types.  Every function returns 0 on success or a B200R_E_* code;
 * b200r_last_error() then describes the failure (thread-local).  There is NO CPU
 * fallback: creating an index without a usable CUDA device fails.
 *
 * Threading: searches on one index may be issued from many threads (they serialise on
 * an internal lock, as the Morpheus service in
 * community/event-driven-rag-cve-analysis/cyber_dev_day/faiss_vdb_service.py:169-174
 * does with faiss); add/remove/reset must not overlap a search on the same index
 * (faiss has the same rule).
 */
#ifndef B200R_H
#define B200R_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200R_VERSION 200

/* status codes */
#define B200R_OK 0
#define B200R_E_INVALID 1     /* bad argument (faiss wrapper: AssertionError) */
#define B200R_E_CUDA 2        /* CUDA runtime/driver failure (faiss: RuntimeError) */
#define B200R_E_UNSUPPORTED 3 /* valid request this build cannot serve */
#define B200R_E_NOMEM 4       /* device or pinned allocation failed */
#define B200R_E_IO 5          /* save/load */

/* metric: numeric values equal faiss::MetricType */
#define B200R_METRIC_INNER_PRODUCT 0 /* IndexFlatIP: larger is better, scores descending */
#define B200R_METRIC_L2 1            /* IndexFlatL2: SQUARED distance, ascending */

/* element type of the corpus in HBM / of vectors handed to add() */
#define B200R_F32 0
#define B200R_BF16 1

/* where a caller-provided buffer lives */
#define B200R_HOST 0
#define B200R_DEVICE 1

/* search path selection (b200r_config.search_path); AUTO mirrors faiss's own
 * dispatch: small batches scan (faiss: exhaustive_*_seq for nq < 20), large batches
 * go through the GEMM expansion (faiss: exhaustive_*_blas for nq >= 20). */
#define B200R_PATH_AUTO 0
#define B200R_PATH_SCAN 1
#define B200R_PATH_GEMM 2

#define B200R_MAX_DEVICES 8

typedef struct b200r_index b200r_index; /* opaque */

typedef struct b200r_config {
    int32_t dim;               /* IndexFlat(d) */
    int32_t metric;            /* B200R_METRIC_* */
    int32_t storage;           /* B200R_F32 (exact mode) or B200R_BF16 (recall mode) */
    int32_t normalize_on_add;  /* fuse faiss.normalize_L2 into ingest (LangChain normalize_L2=True,
                                  rag_service.py:194) */
    int32_t normalize_queries; /* same for queries (rag_service.py:292) */
    int32_t device;            /* CUDA device ordinal */
    int32_t search_path;       /* B200R_PATH_* */
    int32_t shadow_bf16;       /* fp32 storage only: also keep the rows rounded to bf16 (+50% HBM).
                                  Batches on the GEMM path then filter on the shadow (half the bytes,
                                  bf16 tensor-core rate) and re-score the survivors on the fp32 rows:
                                  results stay exact, guarded by the same containment check.  No
                                  faiss counterpart.  0 = off (default). */
    int64_t reserve_rows;      /* HBM reserved up front (0 = grow on demand) */
    int64_t id_base;           /* added to every returned id: global position of local row 0
                                  when the corpus is row-sharded over several processes */
    int32_t n_devices;         /* 0 or 1: the index lives on `device`.  2..8: ONE index whose rows are spread
                                  over the GPUs device_ids[0..n_devices) of this process (SURVEY.md section 8e:
                                  the reference's chain server is a single process,
                                  RAG/src/chain_server/server.py:203-238).  Rows keep faiss's "id == position"
                                  contract; every entry point below works on such an index.  (A device may be
                                  listed more than once -- every entry is one shard -- which is how the tests run
                                  this code path on a 1-GPU box.) */
    int32_t device_ids[B200R_MAX_DEVICES];
} b200r_config;

typedef struct b200r_stats {
    int64_t ntotal;
    int64_t capacity_rows;
    int64_t n_search_calls;
    int64_t n_queries;
    int64_t n_kernel_launches;   /* kernels of this library launched so far */
    int64_t bytes_scanned;       /* algorithmic corpus bytes streamed by searches */
    int64_t n_rows_added;
    int64_t n_rows_removed;
    int32_t last_grid;           /* CTAs of the last scoring launch */
    int32_t last_path;           /* B200R_PATH_SCAN / B200R_PATH_GEMM */
    int32_t num_sms;
    int32_t reserved0;
    double last_search_ms;       /* device time of the last search (events), 0 if timing is off */
    int64_t n_gemm_fallbacks;    /* queries of the GEMM path whose containment check failed and
                                    were re-run on the scan path (results are exact either way) */
    double last_kernel_us;       /* with timing on: device time of the scoring kernels of the last search
                                    (multi-GPU: the slowest shard) */
    double last_exchange_us;     /* multi-GPU, timing on: what the last search took beyond that -- the
                                    candidate exchange over NVLink plus the final select (0 on one GPU) */
    int32_t n_devices;           /* GPUs this index is spread over */
    int32_t n_segments;          /* multi-GPU: contiguous runs of rows in the position -> (GPU, slot) map */
} b200r_stats;

/* faiss.IndexFlatL2(d) / IndexFlatIP(d)   [utils.py:306; rag_service.py:88] */
int b200r_create(const b200r_config* cfg, b200r_index** out);
int b200r_destroy(b200r_index* idx); /* NULL is a no-op */

/* IndexFlat.add(x)  -- faiss replacement_add: x is (n, d) C-contiguous; appended in
 * order, row i of the call gets position ntotal+i.   [langchain FAISS.__add ->
 * chains.py:83; rag_service.py:197,450].   dtype: element type of x (B200R_F32 only for
 * host input).  stream: with DEVICE input the work is ordered on this cudaStream_t exactly
 * as given (NULL = CUDA's default stream, which is also torch's default); with HOST input
 * NULL selects the index's own non-blocking stream.  HOST input may be pageable; the call
 * returns when x has been consumed. */
int b200r_add(b200r_index* idx, const void* x, int64_t n, int dtype, int mem_kind, void* stream);

/* IndexFlat.search(x, k) -- faiss replacement_search: q is (nq, d) fp32; D is (nq, k)
 * fp32, I is (nq, k) int64, both caller-allocated.  L2 returns squared distances
 * ascending, IP dot products descending; when k > ntotal the tail is id -1 with
 * +FLT_MAX (L2) / -FLT_MAX (IP).  Ties are broken by lowest id.
 * [langchain FAISS.similarity_search_with_score_by_vector -> chains.py:163,211;
 * rag_service.py:301].  q and D/I may independently live on the host or the device;
 * with device q AND device outputs the call returns without waiting for the GPU, on every
 * route (queries whose tensor-core filter could not be proven exact are re-run by a
 * device-side loop, not by the host); host q with device outputs waits once, because the
 * queries pass through the index's own pinned buffer.  Searches issued on different streams
 * are ordered after one another on the GPU (they share the index's workspace).  Stream rule
 * as for add(): if any buffer is DEVICE memory the given stream is used as is (NULL = default
 * stream). */
int b200r_search(b200r_index* idx, const void* q, int64_t nq, int k, int q_mem_kind, float* D, int64_t* I,
                 int out_mem_kind, void* stream);

/* index.search followed, ON THE DEVICE, by what LangChain's VectorStore.similarity_search_with_relevance_scores
 * does on the host: every hit's distance d becomes a relevance score (relevance_kind 0 EUCLIDEAN: 1 - d/sqrt(2);
 * 1 COSINE: 1 - d; 2 MAX_INNER_PRODUCT: d > 0 ? 1 - d : -d -- evaluated in double like the Python functions) and,
 * with apply_threshold != 0, only hits with relevance >= threshold are kept, moved to the front in their order.
 * [as_retriever(search_type="similarity_score_threshold") -> chains.py:156-163; multi_turn_rag/chains.py:157-170]
 * q is a HOST array (nq, d); R (nq, k) double, D (nq, k) fp32, I (nq, k) int64, counts (nq) and out_of_range
 * (nq, may be NULL) are HOST arrays: the first counts[i] entries of row i are the kept hits; the hits that the
 * threshold dropped follow them (so that a caller can still word LangChain's warning), then padding with id -1;
 * out_of_range[i] = how many of the query's hits (kept or not) had a relevance outside [0, 1] (LangChain warns
 * about those). */
int b200r_search_threshold(b200r_index* idx, const float* q, int64_t nq, int k, int32_t relevance_kind, double threshold,
                           int32_t apply_threshold, double* R, float* D, int64_t* I, int32_t* counts,
                           int32_t* out_of_range);

/* The greedy selection of LangChain's maximal_marginal_relevance(query, embeddings, lambda_mult, k)
 * [search_type="mmr": FAISS.max_marginal_relevance_search_with_score_by_vector] over the rows keys[0..m) of
 * the index: the rows are gathered and compared on the device (cosine similarities, the score
 * lambda*sim(q, x) - (1 - lambda)*max sim(x, picked) in double), only the min(k, m) selected POSITIONS INTO
 * keys come back.  query, keys, selected (k entries) and n_selected are HOST arrays; m <= 2048. */
int b200r_mmr(b200r_index* idx, const float* query, const int64_t* keys, int64_t m, int k, double lambda_mult,
              int32_t* selected, int32_t* n_selected);

/* IndexFlat.remove_ids(ids) -- IDSelectorBatch semantics: positions in ids are
 * removed, later rows shift down (stable), duplicates / out-of-range ignored.
 * [langchain FAISS.delete -> utils.py:559].  ids is a host array. */
int b200r_remove_ids(b200r_index* idx, const int64_t* ids, int64_t n, int64_t* n_removed);

/* IndexFlat.ntotal */
int b200r_ntotal(const b200r_index* idx, int64_t* out);

/* IndexFlat.reconstruct(key): the stored row as fp32 into a host buffer of d floats */
int b200r_reconstruct(b200r_index* idx, int64_t key, float* out);

/* IndexFlat.reconstruct_batch(keys) / the per-id reconstruct loop of LangChain FAISS
 * max_marginal_relevance_search_with_score_by_vector: rows keys[0..n) as fp32 into a host
 * buffer of n*d floats, one gather launch.  Any key out of range fails the whole call. */
int b200r_reconstruct_batch(b200r_index* idx, const int64_t* keys, int64_t n, float* out);

/* IndexFlat.reset() */
int b200r_reset(b200r_index* idx);

/* faiss.normalize_L2(x): in place on (n, d) fp32, host or device memory.
 * [rag_service.py:194,292,439; langchain FAISS with normalize_L2=True] */
int b200r_normalize_l2(float* x, int64_t n, int32_t d, int mem_kind, int32_t device, void* stream);

/* faiss.write_index / faiss.read_index for flat indexes ("IxF2" / "IxFI" files, the
 * index.faiss that LangChain FAISS.save_local/load_local writes).
 * [community/llm-prompt-design-helper/vector_store/faiss_vector_store.py:124-133;
 * rag_service.py:364-422].  cfg->dim/metric are taken from the file on load. */
int b200r_save(b200r_index* idx, const char* path);
int b200r_load(const char* path, const b200r_config* cfg, b200r_index** out);

/* No faiss counterpart: the final select of the multi-GPU search.  The NCCL-allgathered
 * per-shard results sit in DEVICE memory as n_lists blocks in rank order; block l holds
 * scores [nq][k_in] at cand_D + l*stride_D (floats) and ids [nq][k_in] at
 * cand_I + l*stride_I (int64s).  Writes the k_out best per query to DEVICE arrays D / I
 * ([nq][k_out]); ids < 0 are padding and are skipped; ties go to the lower rank. */
int b200r_merge_topk(int32_t metric, const float* cand_D, const int64_t* cand_I, int64_t nq, int32_t n_lists,
                     int32_t k_in, int64_t stride_D, int64_t stride_I, int32_t k_out, float* D, int64_t* I,
                     int32_t device, void* stream);

/* Has no counterpart in the reference.  The same small-batch scan search as b200r_search,
 * for callers that cannot allocate device memory: a plug-in inside a server that gave all
 * of the HBM to its own pool at start-up (a cudaMalloc there fails or takes someone else's
 * budget), or a serving loop that cannot afford the device-wide stall of cudaMalloc /
 * cudaFree.  It takes no index object.  corpus ([n_rows][dim], fp32 or bf16, row-major),
 * q ([nq][dim] fp32), D ([nq][k] fp32) and I ([nq][k] int64) are all the caller's DEVICE
 * buffers; results follow b200r_search exactly (no normalisation is applied).  The kernel's
 * scratch is a fixed 672,000-byte array built into the library -- every GPU that loads the
 * library holds it until exit -- hence the limits, rejected with B200R_E_UNSUPPORTED and a
 * message: dim <= 4096 and a multiple of 4 (fp32) / 8 (bf16), corpus 16-byte aligned,
 * 1 <= nq <= 4, 1 <= k <= 128.  It calls no cudaMalloc / cudaFree or allocating routine.
 * All calls on one GPU share that array: calls whose kernels may overlap on the GPU
 * (different streams or threads) race and must be serialised by the caller.  The first
 * call on each device also performs a one-time address look-up; issue it before any
 * stream capture starts.  The work is ordered on `stream` (NULL = default stream). */
int b200r_search_noalloc(const void* corpus, int64_t n_rows, int32_t dim, int32_t storage, int32_t metric,
                         const float* q, int32_t nq, int32_t k, float* D, int64_t* I, int64_t id_base,
                         int32_t device, void* stream);

/* No faiss counterpart: the multi-GPU search for small batches (nq <= 4, k <= 128) with the
 * exchange FUSED into the kernels instead of an NCCL call.  Every rank owns one exchange buffer of
 * b200r_exchange_buffer_bytes() bytes, zero-filled once, in memory that all ranks can address
 * (CUDA IPC / torch symmetric memory); peer_bufs[r] is rank r's buffer as mapped in THIS process.
 * The scan kernel's final writer stores this shard's top-k straight into slot [rank] of every
 * rank's buffer (NVLink stores) and raises a per-(epoch parity, rank) flag there; the select kernel
 * that follows on the same stream waits (bounded, ~2 s) for the `world` flags of this epoch in its
 * own buffer and writes the global top-k to the DEVICE arrays D / I.  All ranks must call it
 * collectively with the same nq, k and epoch; epoch starts at 1 and grows by 1 per call, and
 * 0xFFFFFFFF is followed by 2 (never 0, parity alternates: two epochs are buffered, which the
 * stream order of scan -> select on every rank makes sufficient).
 * q (nq, d) and D / I may each live on the host or the device, as for b200r_search; with device buffers
 * the call is asynchronous on `stream`, with host results it returns when they have arrived (one packed
 * copy) and fails if a peer's candidates did not show up within the bound. */
int64_t b200r_exchange_buffer_bytes(void);
int b200r_search_exchange(b200r_index* idx, const float* q, int64_t nq, int k, int q_mem_kind, void* const* peer_bufs,
                          int32_t world, int32_t rank, uint32_t epoch, float* D, int64_t* I, int out_mem_kind,
                          void* stream);
/* After a b200r_search_exchange: synchronises `stream` and reports whether the select kernel gave up
 * waiting for a peer (*status = 0: all arrived; r + 1: rank r's candidates did not arrive within ~2 s
 * and the results of that call are padding).  A non-zero word is cleared.  my_buf = this rank's buffer. */
int b200r_exchange_status(void* my_buf, int32_t device, void* stream, int32_t* status);

/* index.merge_from(other)   [faiss 1.7.4 IndexFlatCodes::merge_from, reached through LangChain's
 * FAISS.merge_from: community/oran-chatbot-multimodal/vectorstore/vectorstore_updater.py:217]
 * Appends src's rows to dst (positions dst.ntotal ...) and empties src.  Both must have the same
 * dim, metric and storage (faiss: check_compatible_for_merge) and live on the same GPU. */
int b200r_merge_from(b200r_index* dst, b200r_index* src);

/* The route (B200R_PATH_SCAN / B200R_PATH_GEMM) the AUTO dispatch takes for a batch of nq queries on a shard of
 * n_rows rows (n_rows < 0: the index's current size).  Ranks of a row-sharded index ask it with the LARGEST
 * shard's size, so that all of them choose the same exchange for a batch.  No faiss counterpart. */
int b200r_auto_route(b200r_index* idx, int64_t nq, int64_t n_rows, int32_t* route);

/* row-sharding support: change the id offset after the shard layout moved */
int b200r_set_id_base(b200r_index* idx, int64_t id_base);

/* an index created with shadow_bf16 filters on the shadow by default; this turns that off/on at
 * run time (no faiss counterpart) */
int b200r_set_shadow_filter(b200r_index* idx, int32_t enabled);

/* observability (no faiss counterpart) */
int b200r_stats_get(b200r_index* idx, b200r_stats* out);
int b200r_set_timing(b200r_index* idx, int32_t enabled);

/* The B200R_* environment switches (README.md) are read once per process; this reads them again
 * (tests and tuning tools change them between searches).  No faiss counterpart. */
int b200r_reload_env(void);

const char* b200r_last_error(void);
int b200r_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B200R_H */
