// Internal launch interfaces between the .cu files of libb200r.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200r {

struct Cand;

constexpr int kScanMaxNQ = 4;    // queries per scan launch
constexpr int kScanMaxK = 128;   // k served by the 4-query scan variants (and by b200r_search_noalloc)
constexpr int kScanMaxKWide = 1024;  // k served by the single-query scan variants (lists of 8..32 slots per lane)
constexpr int kMaxK = kScanMaxKWide; // largest k b200r_search accepts

struct ScanParams {
    const void* corpus;   // [n_rows][d] fp32 or bf16, row-major, rows 16-byte multiples
    int64_t n_rows;
    int d;                // padded dimension (multiple of 4 fp32 / 8 bf16)
    int row_bytes;
    const float* q;       // [NQ][d] fp32 on the device, unused slots zero
    int nq_active;
    int k;
    int nstages, stage_bytes, rows_per_tile;
    int64_t n_tiles;
    int evict_first;      // corpus larger than L2: stream it with evict_first
    Cand* scratch;        // [grid][NQ][KCAP]
    unsigned* counter;    // zero on entry, zero on exit
    float* out_d;         // [nq][k]
    int64_t* out_i;       // [nq][k]
    int64_t id_base;      // added to every returned id (row-sharded multi-GPU)
    // fused exchange (xchg_ntargets > 0): the final writer stores this shard's result straight into slot
    // [xchg_rank] of every target exchange buffer (over NVLink when it is a peer's) and raises a flag
    // there, instead of writing out_d / out_i.  Targets = every rank (allgather) or the primary GPU alone.
    int xchg_ntargets, xchg_rank;
    uint32_t xchg_epoch;
    uint8_t* xchg_peer[8];  // peer-mapped base of each rank's exchange buffer (layout below)
    // flagged re-run (launch_scan_flagged): the queries to serve are q[flag_list[0..*flag_count)], their
    // results go to row flag_list[i] of out_d / out_i; the launch is cooperative and loops over them
    const int* flag_list;
    const int* flag_count;
    // diagnosis only (B200R_SCAN_TRACE=1): per-CTA timestamps [grid][8] (globaltimer ns), else null
    unsigned long long* trace;
    // dynamic tile assignment (tile_counter != null): CTAs draw tiles from a global counter instead of
    // walking t = blockIdx.x, +grid, ...; zero on entry, reset by the final writer
    unsigned* tile_counter;
    int tile_block;            // static assignment: consecutive tiles per CTA and round (<= 1: cyclic)
    long long dyn_from;        // with a tile counter: tiles below this index are dealt statically (0: all dynamic)
};

// Exchange buffer of one rank (symmetric: same layout on every rank, peer-mapped over NVLink):
//   [parity 0|1][slot of source rank 0..7] of kXchgSlotBytes: scores fp32 [nq][k] at +0, ids int64 [nq][k] at +2048
//   then flags: one uint32 per (parity, source rank), 128 bytes apart; then one status word
constexpr int kXchgMaxWorld = 8;
constexpr int kXchgSlotBytes = 8192;
constexpr int kXchgIdOffset = 2048;
constexpr int kXchgParityBytes = kXchgMaxWorld * kXchgSlotBytes;
constexpr int kXchgFlagOffset = 2 * kXchgParityBytes;
constexpr int kXchgStatusOffset = kXchgFlagOffset + 2 * kXchgMaxWorld * 128;
constexpr int kXchgBytes = kXchgStatusOffset + 256;

cudaError_t launch_merge_wait(bool l2, uint8_t* my_buf, int world, int nq, int k, uint32_t epoch, float* out_d,
                              int64_t* out_i, int* status_out, cudaStream_t st);

size_t scan_scratch_bytes(int max_grid);
int scan_queries_per_launch(int d_pad, int row_bytes, size_t smem_limit);
// Device-side fallback of the GEMM route: compacts fail_flags[0..nq_total) into flag_list / flag_count and
// re-runs exactly those queries through the scan kernel in ONE cooperative launch (which returns at
// once when nothing is flagged), so the host never has to read the flags back.
cudaError_t launch_scan_flagged(ScanParams p, const int* fail_flags, int64_t nq_total, int* flag_list, int* flag_count,
                                bool l2, bool bf16, int num_sms, size_t smem_limit, cudaStream_t st);
cudaError_t launch_scan(ScanParams p, int nq, bool l2, bool bf16, int num_sms, size_t smem_limit, cudaStream_t st,
                        int* grid_out);

// K1 ingest: rows [n][d_in] fp32 -> corpus rows [n][d_pad] (fp32 or bf16), optional
// L2 normalisation (faiss fvec_renorm_L2 semantics), squared norms of what was stored.
cudaError_t launch_ingest(const float* src, int64_t n, int d_in, int d_pad, bool normalize, bool to_bf16, void* dst_rows,
                          float* dst_norms, cudaStream_t st);
// in-place faiss.normalize_L2 on device memory
cudaError_t launch_normalize(float* x, int64_t n, int d, cudaStream_t st);
// query prep: [nq][d_in] fp32 -> [nq_slots][d_pad] fp32, zero padded, optional normalise
cudaError_t launch_prep_queries(const float* src, int64_t nq, int d_in, int d_pad, int64_t nq_slots, bool normalize,
                                float* dst, cudaStream_t st);

// K5 compaction gather: out row o (o0 <= o < o1) <- src row o + #{deleted ids before it}
cudaError_t launch_gather_rows(const void* src, void* out, int64_t o0, int64_t o1, int row_bytes,
                               const int64_t* del_adj, int64_t n_del, cudaStream_t st);

// K4 merge: n_lists candidate blocks, block l at cand_x + l * stride_x holding [nq][k_in]
// -> [nq][k_out]
cudaError_t launch_merge(bool l2, const float* cand_d, const int64_t* cand_i, int64_t nq, int n_lists, int k_in,
                         int64_t stride_d, int64_t stride_i, int k_out, float* out_d, int64_t* out_i,
                         cudaStream_t st);

// final select of the single-process multi-GPU index: n_lists sorted per-shard lists in the slots of a
// gather buffer, local row numbers translated to global positions through per-shard segment tables
cudaError_t launch_merge_shards(bool l2, const uint8_t* base, int64_t slot_stride, int64_t ids_off, int n_lists, int nq,
                                int k_in, int k_out, const int* seg_off, const int64_t* seg_lstart,
                                const int64_t* seg_gbase, int64_t id_base, float* out_d, int64_t* out_i, uint8_t* xbuf,
                                uint32_t wait_epoch, int* status_out, cudaStream_t st);

// K3 (gemm.cu): tcgen05 GEMM filter + exact re-score for one block of <= 256 queries
constexpr int kGemmSlack = 22;       // survivors kept beyond k when k + slack <= 32
constexpr int kGemmMaxBlock = 256;   // queries per GEMM launch (two 128-row M tiles)
constexpr int kGemmMaxGrid = 256;    // CTAs of a GEMM launch never exceed the SM count (148 on B200)
struct GemmSearchArgs {
    const void* corpus;
    const float* norms;
    int64_t n_rows;
    int d_pad;
    bool bf16, l2;
    const float* q_f32;   // [128 or 256][d_pad] prepared queries, unused rows zero
    const void* q_mma;    // same rows in the filter's MMA input type (fp32: == q_f32; bf16: bf16 copy)
    // optional bf16 shadow of an fp32 corpus ([n_rows][d_pad] bf16, d_pad % 8 == 0): the filter pass
    // streams it instead of `corpus` (half the bytes, bf16 MMA rate); the re-score still reads `corpus`
    const void* shadow;
    int kp_boost_pct;     // >= 100: widens the selection size K' (raised after failed containment checks)
    int nq, k;
    Cand* cand;           // >= gemm_cand_bytes(num_sms, 256, gemm_kp_cap(k, filter, boost))
    int* gthr;            // [256] scratch: grid-wide per-query thresholds
    unsigned* hist;       // [256][512] scratch: score histograms
    float* qscale;        // [256][2] scratch: bucket maps
    int* list_counts;     // [grid][256] scratch: survivors per (CTA, query)
    int* overflow;        // [256] scratch: survivor slice overflowed
    const unsigned* max_norm_bits;
    float* out_d;
    int64_t* out_i;
    int64_t id_base;
    int* fail_flags;      // [nq]: 1 = containment check failed, re-run on the scan path
    int* fail_count;
    int num_sms;
};
cudaError_t launch_gemm_search(const GemmSearchArgs& a, cudaStream_t st);
cudaError_t launch_max_norm(const float* norms, int64_t n, unsigned* out_bits, cudaStream_t st);
cudaError_t launch_f32_to_bf16(const float* src, void* dst, int64_t n, cudaStream_t st);
size_t gemm_cand_bytes(int grid, int nq_pad, int kp_cap);
size_t gemm_aux_bytes();
void gemm_aux_layout(void* aux, unsigned** hist, float** qscale, int** gthr, int** list_counts, int** overflow);
enum { kFilterTf32 = 0, kFilterBf16 = 1, kFilterShadow = 2 };
int gemm_kp_cap(int k, int filter, int boost_pct);

// lets entry points outside api.cu fill b200r_last_error()
void set_last_error(const char* msg);

// reconstruct helper: corpus row -> fp32
cudaError_t launch_row_to_f32(const void* row, int d, bool bf16, float* out, cudaStream_t st);
cudaError_t launch_rows_to_f32(const void* rows, int row_bytes, int d, bool bf16, const int64_t* ids, int64_t n,
                               float* out, cudaStream_t st);

}  // namespace b200r
