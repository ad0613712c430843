// Internal launch interfaces between the translation units of libb2flat.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b2_flat.h"

namespace b2 {

// b2_scan.cu ---------------------------------------------------------------
size_t scan_smem_bytes(int nq_tile, int nchunks, bool bf16, int kpad);
int scan_pass(const void* rows, int64_t n, size_t pitch_bytes, int d, int dtype, int metric, const float* q_dev,
              int nq_valid, int nq_tile, int k, int num_sms, cudaStream_t stream, uint64_t** partial, int* grid,
              const uint32_t* run_if = nullptr, const void* resid = nullptr);  // resid: residual plane of bf16 rows

void prof_enable(int on);
void* prof_begin(cudaStream_t stream);                       // nullptr when profiling is off
void prof_end(void* token, int kind, cudaStream_t stream);   // kind 0 = scan kernel, 1 = tensor-core kernel
int prof_read(int kind, double* ms, int64_t* n, int reset);

// b2_index.cu -------------------------------------------------------------
// Read-only snapshot of an index handle for entry points implemented next to their kernels.
struct IndexView {
    const void* rows;
    const void* resid;  // residual plane of bf16 rows (same pitch), or null
    int64_t ntotal;
    size_t pitch;
    int d, dtype, metric, device, num_sms;
};
// Takes the handle's reader lock and fills *v; pair with index_unlock_shared.
int index_view_lock_shared(b2_index* h, IndexView* v);
void index_unlock_shared(b2_index* h);

// Peer-memory exchange of the row-sharded search: where the final lists of this rank go instead of local D / I.
// block[g] = peer g's receive block for THIS rank's lists ([I: nq_total x k int64 | D: nq_total x k fp32]),
// flag[g] = peer g's flag words for this rank (one per query); after a query's k results are stored to every
// peer, flag[g][q] = epoch is released system-wide.  Pointers are NVLink peer mappings (or local for g == self).
constexpr int kMaxPushPeers = 16;
struct PeerPush {
    int nparts;
    uint32_t epoch;
    int64_t nq_total;
    int64_t q_offset;  // first query of a merge launch within the call (multi-pass scans)
    char* block[kMaxPushPeers];
    uint32_t* flag[kMaxPushPeers];
};

// b2_merge.cu --------------------------------------------------------------
// partial: [nq][nlists][k] candidates (pos = local row id) -> D/I [nq][k] with I = id_base + pos.
// qnorm (optional, L2 only): keys are 2 q.x - |x|^2 and D = max(0, qnorm[q] - key) (the BLAS-path form).
// only_if (optional, device, [nq]): query q is merged only when only_if[q] != 0 (fix-up launches).
int merge_partials(const uint64_t* partial, int nq, int nlists, int k, int metric, int64_t id_base, float* D,
                   int64_t* I, size_t out_stride, cudaStream_t stream, const float* qnorm = nullptr,
                   const PeerPush* push = nullptr, const uint32_t* only_if = nullptr);
// refine mode of the tensor-core kernel: merge by approximate key, re-score the candidates near the cut exactly
// (fp32 FMA dot of q_dev[q] with the stored row), emit the exact top-k.  partial: [nq][nlists][list_len].
int merge_rescore(const uint64_t* partial, int nq, int nlists, int list_len, int k, int metric, float* D, int64_t* I,
                  size_t out_stride, cudaStream_t stream, const float* qnorm, const float* q_dev, int d, const void* rows,
                  size_t pitch_bytes, int dtype, const float* norms, const float* band, const PeerPush* push = nullptr,
                  uint32_t* overflow = nullptr, bool l2_direct = false, const void* resid = nullptr);
// resid: residual plane of bf16 rows (same pitch): candidates are re-scored against hi + lo
// overflow (optional, device, [nq + 1], zeroed by the caller): the lists are sorted best-first and may have dropped
// candidates beyond list_len; a query whose cut reaches into a full list's last entry gets overflow[q] = 1 (and
// overflow[nq] = 1) and NO output: the caller redoes it exactly.  Needs nlists * list_len <= kMergeSelectMax.
constexpr int kMergeSelectMax = 24 * 512;  // candidates per query the register-resident merge holds (24 per thread)
// merge of exchanged blocks that first waits (per query) for every part's flag to reach `epoch`
int merge_shards_wait(int metric, int nq, int k, int nparts, const char* blocks, size_t part_stride_bytes,
                      const uint32_t* flags, size_t flag_stride_elems, uint32_t epoch, const int64_t* id_bases, float* D,
                      int64_t* I, cudaStream_t stream);

// query block -> every rank's receive buffer + flag (peer mappings), and the wait for it on the receiving stream
int query_push(const float* q_dev, size_t nbytes, int nparts, void* const* bufs, uint32_t* const* flags, uint32_t epoch,
               cudaStream_t stream);
int flag_wait(const uint32_t* flag, uint32_t epoch, cudaStream_t stream);

// b2_filter.cu -------------------------------------------------------------
// langchain's post-search selection (metadata filter as a row bitmask, raw / relevance thresholds, cut at k) over
// device results [nq][fetch_k] -> [nq][k] + counts[nq]; see include/b2_flat.h: b2_index_search_filtered.
int filter_results(const float* D_in, const int64_t* I_in, int64_t nq, int fetch_k, int k, int metric,
                   const uint32_t* allow, int64_t allow_bits, int use_raw, double raw_thr, int rel_kind, double rel_thr,
                   float* D_out, int64_t* I_out, int32_t* counts, cudaStream_t stream);

int build_shadow_rows(const void* rows, size_t pitch, void* shadow, size_t spitch, int64_t row0, int64_t n, int d,
                      uint32_t* err_ord, cudaStream_t stream);

// b2_gemm.cu ---------------------------------------------------------------
// Tensor-core search for batches of >= 20 queries (faiss's BLAS threshold), k <= 128.
// q [nq][d] fp32 -> A [na][nq_pad][kp] operand parts (bf16 hi/lo, or fp32 words holding tf32 hi/lo), zero padded;
// optional qn[row] = |q|^2 and tau[row] = initial admission bound.
// mixed_slice > 0 (fp32 operands only): part 1 becomes the bf16 correction operand, per mixed_slice-element K slice
// [bf16(q) | bf16(q - tf32(q))] (same bytes per row).
int prep_queries(const float* q_dev, int nq, int d, int nq_pad, int kp, int na, bool bf16, void* A, float* qn,
                 uint32_t* tau, cudaStream_t stream, int mixed_slice = 0);
bool gemm_path_supported(int64_t n, int d, int dtype, int k, int64_t nq);
// A bf16 copy of fp32 rows for the approximate pass of filter-and-refine (b2_ingest.cu: build_shadow_rows).
struct ApproxRows {
    const void* rows;         // [n][pitch] bf16
    size_t pitch;             // bytes
    const uint32_t* err_ord;  // device: ordered image of max over rows of |x - bf16(x)|^2
};
// Internal status (never crosses the C ABI): the tensor-core search could not finish rigorously and the caller redoes
// the batch with the exact FFMA scan (rows with a residual plane whose candidate lists overflowed).
#define B2_INTERNAL_REDO_EXACT 100
// Residual plane of a bf16 row store: the index holds hi + lo, the tensor cores read hi only.
struct ResidualRows {
    const void* rows;         // [n][pitch] bf16, same pitch as the main plane
    const uint32_t* err_ord;  // device: ordered image of max over rows of |lo|^2
};
int gemm_search(const void* rows, int64_t n, size_t pitch_bytes, int d, int dtype, int metric, const float* norms,
                const float* q_dev, int64_t nq, int k, int num_sms, float* D, int64_t* I, cudaStream_t stream,
                const PeerPush* push = nullptr, int query_parts = 0, const uint32_t* max_norm_ord = nullptr,
                const ApproxRows* approx = nullptr, const ResidualRows* resid = nullptr);
// band[q] = 2 x a rigorous bound on |approximate score - exact score| for one-product tensor-core scoring of query q
// (bf16 or tf32 operands, `steps` accumulation steps, rows bounded by max_norm_ord, optional row rounding error).
int compute_band(const float* q_dev, int nq, int d, int nq_pad, int steps, bool bf16_operands, bool l2,
                 const uint32_t* max_norm_ord, const uint32_t* row_err_ord, float* band, cudaStream_t stream);
// true when gemm_search with these options would run filter-and-refine (and so can use a bf16 shadow of fp32 rows)
bool gemm_uses_refine(int query_parts, const uint32_t* max_norm_ord);
// (D_parts, I_parts): [nparts][nq][k] -> [nq][k]
// part g's lists start d_stride_bytes / i_stride_bytes after part g-1's (0 = natural [nparts][nq][k] arrays);
// id_bases (host, nparts entries, optional): ids in the lists are local to their part, global = base + local.
int merge_shards(int metric, int nq, int k, int nparts, const float* D_parts, const int64_t* I_parts, float* D,
                 int64_t* I, cudaStream_t stream, size_t d_stride_bytes = 0, size_t i_stride_bytes = 0,
                 const int64_t* id_bases = nullptr);

// b2_ingest.cu -------------------------------------------------------------
// src: n x d fp32 dense (device) -> dst rows at `pitch_bytes`, fp32 or bf16, padding zeroed;
// optional L2 normalisation; norms[i] = sum of squares of the STORED row.
// resid (bf16 rows only, optional): second plane of the same pitch receiving bf16(x - bf16(x)); the row the index
// holds is then hi + lo, norms[i] its squared norm, and *resid_err_ord the running maximum of sum lo^2 (ordered image).
int ingest_rows(const float* src, int64_t n, int d, void* dst, size_t pitch_bytes, int dtype, bool normalize,
                float* norms, cudaStream_t stream, void* resid = nullptr, uint32_t* resid_err_ord = nullptr);
int normalize_rows_inplace(float* x, int64_t n, int d, cudaStream_t stream);
// out[j] = row ids[j] as fp32 (n x d dense); negative / out-of-range ids give zero rows
int gather_rows(const void* rows, int64_t ntotal, size_t pitch_bytes, int d, int dtype, const int64_t* ids, int64_t n,
                float* out, cudaStream_t stream, const void* resid = nullptr);
// keep[i] != 0 rows are compacted to the front preserving order; dst_index = exclusive prefix of keep
// (norms may be null: a plane without its own norm array)
int compact_rows(void* rows, size_t pitch_bytes, float* norms, const int64_t* src_of_dst, int64_t n_dst,
                 int64_t first_moved, cudaStream_t stream);

// remove_ids on the device: from n ids (device) the list src[j] = row that moves to (first + j), the number of
// distinct rows dropped and the first of them (ntotal if none).  src_dev holds up to ntotal entries; scratch of
// keep_list_scratch_bytes(ntotal) bytes.  Synchronises `stream` (the two counts come back to the host).
size_t keep_list_scratch_bytes(int64_t ntotal);
int build_keep_list(const int64_t* ids_dev, int64_t n, int64_t ntotal, int64_t* src_dev, void* scratch, size_t scratch_bytes,
                    int64_t* removed, int64_t* first, cudaStream_t stream);

// max_norm_ord = max(max_norm_ord, ordered image of max(norms[0..n)))
int update_max_norm(const float* norms, int64_t n, uint32_t* max_norm_ord, cudaStream_t stream);

// b2_mmr.cu ----------------------------------------------------------------
int mmr_select(const float* q, int d, int nq, const float* cand_rows /*[nq][fetch_k][d]*/, const int64_t* cand_ids,
               const float* cand_scores, int fetch_k, int k, float lambda_mult, int64_t* out_ids, float* out_scores,
               cudaStream_t stream);

// b2_scan_tc.cu ------------------------------------------------------------
// Tensor-core scan for 3..16 queries (inner product): the corpus is the M operand, HBM-bound.
// approx: the filter-and-refine mode is available (fp32 rows; needed for L2).
// residual: bf16 rows with a residual plane (inner product then rides the approximate mode too, like L2)
bool scan_tc_supported(int64_t n, int d, int dtype, int metric, int k, int64_t nq, bool approx = true, bool residual = false);
// the approximate mode's register-resident merge holds (CTAs launched x list length) candidates per query
bool scan_tc_approx_feasible(int64_t n, int k, int num_sms);
// max_norm_ord (optional): enables the approximate (filter-and-refine) mode on fp32 rows.  metric L2 (fp32 rows,
// approximate mode only) takes the rows' squared norms.
int scan_tc_search(const void* rows, int64_t n, size_t pitch_bytes, int d, int dtype, const float* q_dev, int nq, int k,
                   int num_sms, float* D, int64_t* I, cudaStream_t stream, const PeerPush* push = nullptr,
                   const uint32_t* max_norm_ord = nullptr, int metric = B2_METRIC_IP, const float* norms = nullptr,
                   const ResidualRows* resid = nullptr);

// b2_synth.cu --------------------------------------------------------------
int synth_fill(float* out, int64_t row0, int64_t nrows, int d, uint64_t seed, cudaStream_t stream);

}  // namespace b2
