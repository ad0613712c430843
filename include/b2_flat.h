/*
 * b2_flat.h — C ABI of the B200 flat (exact) vector index.
 *
 * This is the drop-in boundary for the one hot path of april-yyt/NV_GenAI that
 * this library replaces: the faiss IndexFlatL2 / IndexFlatIP object that
 * langchain's FAISS vector store drives.  The reference reaches it at
 *   RetrievalAugmentedGeneration/common/utils.py:85   (from faiss import IndexFlatL2)
 *   RetrievalAugmentedGeneration/common/utils.py:217  (FAISS(embedder, IndexFlatL2(dim), InMemoryDocstore(), {}))
 *   examples/5_mins_rag_no_gpu/main.py:90             (FAISS.from_documents -> IndexFlatL2(d) -> index.add)
 *   examples/5_mins_rag_no_gpu/main.py:125-126        (retriever.invoke -> index.search(1 x 1024, k=4))
 * faiss itself (faiss-cpu==1.7.4, RetrievalAugmentedGeneration/requirements.txt:25)
 * is bound through SWIG; every entry point below names the faiss / langchain
 * routine whose SWIG binding it replaces.
 *
 * Conventions
 *   - plain C: opaque handle, pointers and sizes; no C++/torch types cross it.
 *   - every function returns a b2_status (0 = ok); b2_last_error() gives the
 *     thread-local message of the last failure.  No exception crosses the ABI.
 *   - "where" arguments say whether a caller buffer lives in host or device
 *     memory (B2_HOST / B2_DEVICE).  Host buffers are copied inside the call.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).
 *     Work is enqueued on it; calls that return results in HOST buffers
 *     synchronise that stream before returning, device-output calls do not.
 *   - the caller owns x/q/D/I; the library owns the row store and any scratch
 *     it allocates (freed by b2_index_destroy).
 *   - many readers / one writer: searches may run concurrently from different
 *     host threads; add / remove / merge / reset must not overlap anything.
 *   - there is no CPU fallback: without a CUDA device every call fails.
 */
#ifndef B2_FLAT_H
#define B2_FLAT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define B2_API __attribute__((visibility("default")))
#else
#define B2_API
#endif

typedef struct b2_index b2_index;

typedef enum {
    B2_OK = 0,
    B2_ERR_INVALID = 1,   /* bad argument: dimension mismatch, k <= 0, null pointer ... */
    B2_ERR_CUDA = 2,      /* a CUDA runtime / driver call failed */
    B2_ERR_OOM = 3,       /* device or host allocation failed */
    B2_ERR_UNSUPPORTED = 4, /* valid request outside what the kernels cover (e.g. k > B2_MAX_K) */
    B2_ERR_NO_DEVICE = 5  /* no usable CUDA device: the library has no CPU path */
} b2_status;

/* faiss MetricType values (faiss/MetricType.h): METRIC_INNER_PRODUCT = 0, METRIC_L2 = 1 */
#define B2_METRIC_IP 0
#define B2_METRIC_L2 1

/* row-store element type */
#define B2_DTYPE_F32 0
#define B2_DTYPE_BF16 1

#define B2_HOST 0
#define B2_DEVICE 1

/* b2_index_add flags */
#define B2_ADD_NORMALIZE 1u /* faiss.normalize_L2 fused into the ingest kernel */

#define B2_MAX_K 2048

/* Replaces the constructors faiss.IndexFlatL2(d) / faiss.IndexFlatIP(d)
 * (common/utils.py:217; created inside FAISS.from_documents at main.py:90).
 * `device` is the CUDA ordinal that will hold the row store. */
B2_API int b2_index_create(int d, int metric, int dtype, int device, b2_index** out);

/* Replaces the SWIG destructor of the index object. */
B2_API int b2_index_destroy(b2_index* h);

/* Replace the attributes index.ntotal, index.d, index.metric_type
 * (langchain reads ntotal in FAISS.__add; callers at common/utils.py:342). */
B2_API int64_t b2_index_ntotal(const b2_index* h);
B2_API int b2_index_d(const b2_index* h);
B2_API int b2_index_metric(const b2_index* h);
B2_API int b2_index_dtype(const b2_index* h);
B2_API int b2_index_device(const b2_index* h);

/* Replaces index.add(x) -> IndexFlatCodes::add, plus faiss.normalize_L2(x)
 * when B2_ADD_NORMALIZE is set (langchain FAISS.__add; reached from
 * add_documents at examples/nvidia_api_catalog/chains.py:61, add_texts at
 * examples/multi_turn_rag/chains.py:62, from_documents at main.py:90).
 * x: n x d fp32, C-contiguous, host or device.  Rows get ids ntotal..ntotal+n-1. */
B2_API int b2_index_add(b2_index* h, int64_t n, const float* x, int x_where, unsigned flags, void* stream);

/* Replaces index.search(x, k) -> IndexFlat::search (knn_L2sqr / knn_inner_product);
 * every retriever call of the reference ends here.
 * q: nq x d fp32.  D: nq x k fp32, I: nq x k int64 (both host or both device,
 * per out_where).  L2 = squared distance ascending, IP = dot product descending,
 * unfilled slots I = -1 and D = +FLT_MAX (L2) / -FLT_MAX (IP); equal scores are
 * ordered lowest id first. */
B2_API int b2_index_search(b2_index* h, int64_t nq, const float* q, int q_where, int k, float* D, int64_t* I,
                    int out_where, void* stream);

/* No counterpart in the reference: b2_index_search for callers that must not allocate.
 * A service that gives all of HBM to its own memory manager at start-up, or a request loop
 * that cannot risk the device-wide synchronisation an allocator call may cause, cannot use
 * b2_index_search, which takes scratch from cudaMallocAsync.  This variant runs the same scan
 * and merge kernels with the per-CTA partial lists in a fixed 2 MiB array inside the library
 * image (held by every GPU that loads the library, until the process exits) and calls no
 * cudaMalloc / cudaFree / allocating routine.
 *   - q, D, I are DEVICE pointers (nq x d fp32, nq x k fp32, nq x k int64); nq <= 16;
 *     work is enqueued on `stream`, nothing is synchronised.
 *   - limit: query tile (nq rounded up to a power of two) x CTAs launched x k must not exceed
 *     256 Ki candidates, otherwise B2_ERR_UNSUPPORTED (on a 148-SM B200: nq = 16 with k = 100
 *     and nq = 1 with k = 10 fit; nq = 1 with k = 1024 does not).
 *   - set-up happens inside the first calls, not at load time: the first call on each device looks up the
 *     array's device address (cudaGetSymbolAddress, cached per device), and the first call of each SHAPE - query
 *     tile (nq rounded up to a power of two), k and d, per metric and row type - makes the CUDA runtime load that
 *     kernel variant and sets its shared-memory limit (cudaFuncSetAttribute, an occupancy query).  Loading a
 *     kernel may itself take device memory and synchronise.  So run every shape the service will use once during
 *     start-up - before the memory is handed over and before any stream capture - or start the process with
 *     CUDA_MODULE_LOADING=EAGER; after that a call touches neither the allocator nor the module loader.
 *   - all calls share the one array and it keeps the last call's contents.  Nothing in it has
 *     to start at zero (every slot a call reads, the same call has written first), but two
 *     calls whose kernels overlap on the GPU (different streams, or threads) race on it:
 *     serialise them.
 * It always runs the FFMA scan kernel.  Results are bit-identical to b2_index_search whenever that runs the same
 * kernel: on a row store under 8 GB, up to 4 queries on fp32 rows (2 on bf16 rows), or k > 64.  From 5 queries (3 on
 * bf16 rows) with k <= 64, and for any batch on a store of 8 GB and more, b2_index_search uses the tensor-core scan
 * for both metrics (re-scored in fp32): the two then agree within the 1e-5 relative score tolerance, and ids inside
 * a run of scores that tie within it may swap. */
B2_API int b2_index_search_static(b2_index* h, int64_t nq, const float* q_dev, int k, float* D_dev, int64_t* I_dev,
                           void* stream);

/* Replaces index.reconstruct(i) (used by langchain's MMR search), batched:
 * out[j] = row ids[j] as fp32 (n x d).  Negative ids give zero rows. */
B2_API int b2_index_reconstruct_batch(b2_index* h, int64_t n, const int64_t* ids, int ids_where, float* out,
                               int out_where, void* stream);

/* Replaces index.remove_ids(np.int64 array) -> IndexFlatCodes::remove_ids
 * (langchain FAISS.delete; called at common/utils.py:376): in-place
 * compaction preserving order; ids shift down.  ids: host int64. */
B2_API int b2_index_remove_ids(b2_index* h, int64_t n, const int64_t* ids, int64_t* n_removed);

/* Replaces index.merge_from(other) (langchain FAISS.merge_from; called at
 * experimental/oran-chatbot-multimodal/vectorstore/vectorstore_updater.py:216):
 * appends other's rows, then empties other. */
B2_API int b2_index_merge_from(b2_index* dst, b2_index* src);

/* Relevance-score functions of langchain's FAISS (_euclidean_relevance_score_fn,
 * _cosine_relevance_score_fn, _max_inner_product_relevance_score_fn), selected there by distance_strategy. */
#define B2_REL_NONE 0
#define B2_REL_EUCLIDEAN 1         /* 1 - d / sqrt(2), d = squared L2 distance as returned */
#define B2_REL_COSINE 2            /* 1 - d */
#define B2_REL_MAX_INNER_PRODUCT 3 /* d > 0 ? 1 - d : -d */

/* Replaces the selection langchain's FAISS.similarity_search_with_score_by_vector and
 * similarity_search_with_relevance_scores run on the host after index.search (thresholded retrievers:
 * RetrievalAugmentedGeneration/examples/nvidia_api_catalog/chains.py:122,162,
 * examples/multi_turn_rag/chains.py:160-165, experimental/oran-chatbot-multimodal/retriever/retriever.py:69-71;
 * the threshold comes from RetrievalAugmentedGeneration/common/configuration.py:146):
 * searches fetch_k results per query (the caller passes fetch_k = k when there is no metadata filter, as upstream
 * does), then, in search order,
 *   1. drops padding (-1) and rows whose bit in `allow` is clear (allow = NULL: no metadata filter; bit i of word
 *      i / 32 = row i; allow_bits = rows covered, rows beyond it are dropped),
 *   2. if use_raw_threshold: keeps score >= raw_threshold (inner product) or score <= raw_threshold (L2),
 *   3. cuts the list at k,
 *   4. if relevance_kind != B2_REL_NONE: keeps relevance(score) >= relevance_threshold,
 * and writes the survivors' RAW scores and ids to D / I (nq x k, padded with -1 / the metric's worst score) and
 * their number to counts (nq, int32; may be NULL).  Comparisons are in double on the widened fp32 score, like
 * the numpy 1.26 the reference pins.  q, allow, D, I and counts each follow their *_where argument. */
B2_API int b2_index_search_filtered(b2_index* h, int64_t nq, const float* q, int q_where, int k, int fetch_k,
                                    const uint32_t* allow, int64_t allow_bits, int allow_where,
                                    int use_raw_threshold, double raw_threshold, int relevance_kind,
                                    double relevance_threshold, float* D, int64_t* I, int32_t* counts, int out_where,
                                    void* stream);

/* The same selection applied to results already on the device (nq x fetch_k, as b2_index_search wrote them):
 * all pointers are device pointers.  metric: B2_METRIC_IP / B2_METRIC_L2. */
B2_API int b2_filter_results(int metric, int64_t nq, int fetch_k, int k, const float* D_in, const int64_t* I_in,
                             const uint32_t* allow, int64_t allow_bits, int use_raw_threshold, double raw_threshold,
                             int relevance_kind, double relevance_threshold, float* D_out, int64_t* I_out,
                             int32_t* counts, int device, void* stream);

/* Engine options without a faiss counterpart.  "gemm_query_parts": how the query enters the tensor-core kernel
 * for batches of >= 20 queries.  0 (default) = filter and refine: one approximate tensor-core product per score,
 * every candidate within a rigorous error band of the k-th best is re-scored exactly in fp32, so results equal
 * the all-exact arithmetic; 2 = all-exact tensor-core arithmetic (query split into hi + lo parts; tf32 main +
 * bf16 correction product on fp32 rows); 1 = single part without re-scoring: scores then carry the bf16 (or tf32)
 * rounding of the query, recall@100 0.9952 instead of 0.9970 at 10M x 1024 bf16.
 * "bf16_shadow" (fp32 rows; default 1): keep a bf16 copy of the rows (+2 bytes per element of HBM, built by the
 * first batch of >= 20 queries, kept current across add / remove_ids) that the approximate product of
 * filter-and-refine reads instead of the fp32 rows: twice the tensor rate, same results (the fp32 rows decide).
 * 0 releases the copy.  Skipped silently when the device has no room for it.
 * "bf16_residual" (bf16 rows; default 0; can only be changed while the index is EMPTY): store every row as two bf16
 * planes, hi = bf16(x) and lo = bf16(x - hi), 4 bytes per element.  The index then stands for the rows hi + lo
 * (~16 mantissa bits; b2_index_reconstruct_batch returns them): the tensor-core passes stream the main plane only,
 * as on a plain bf16 store, with the error band widened by the residual's norm, and every returned score is
 * re-computed against both planes; batches that do not take a filter-and-refine path run the FFMA scan over both
 * planes (the traffic of fp32 rows).  Measured at 10M x 1024: recall@100 against the fp32 rows 1.000 (plain bf16:
 * 0.997) at the plain bf16 store's time + 2 %.  Indexes merged with b2_index_merge_from must agree on it. */
B2_API int b2_index_set_option(b2_index* h, const char* name, int64_t value);

/* Replaces index.reset(). */
B2_API int b2_index_reset(b2_index* h);

/* Replaces the MMR step of langchain
 * FAISS.max_marginal_relevance_search_with_score_by_vector: index.search(q, fetch_k)
 * -> index.reconstruct for each hit -> maximal_marginal_relevance(q, rows, lambda, k)
 * (langchain_community/vectorstores/utils.py), for a batch of queries, all on
 * the device.  out_ids: nq x k int64 in pick order (-1 padded); out_scores:
 * nq x k fp32 = the search score of each pick. */
B2_API int b2_index_mmr(b2_index* h, int64_t nq, const float* q, int q_where, int k, int fetch_k, float lambda_mult,
                 int64_t* out_ids, float* out_scores, int out_where, void* stream);

/* Row-sharded search, exchange step: merges G per-shard result lists
 * (D_parts: G x nq x k fp32, I_parts: G x nq x k int64 GLOBAL ids, device
 * memory, shard g's rows all have lower ids than shard g+1's) into the global
 * top-k (D, I: nq x k, device).  Has no single faiss counterpart on the CPU
 * path of the reference; it restates what faiss IndexShards does after its
 * per-shard searches and is what the NCCL allgather feeds. */
B2_API int b2_merge_shard_lists(int metric, int64_t nq, int k, int nparts, const float* D_parts, const int64_t* I_parts,
                         float* D, int64_t* I, int device, void* stream);

/* The same exchange step for the layout the all-gather produces directly: part g's lists are at
 * D_parts + g * part_stride_bytes and I_parts + g * part_stride_bytes (each rank sends one packed
 * block [I: nq x k int64 | D: nq x k fp32], so D_parts = I_parts + nq * k * 8 bytes), ids are LOCAL to
 * their shard and id_bases[g] (host array, nparts entries) is the global id of shard g's first row.
 * At most B2_MAX_SHARDS parts. */
#define B2_MAX_SHARDS 64
B2_API int b2_merge_shard_lists_packed(int metric, int64_t nq, int k, int nparts, const float* D_parts,
                                const int64_t* I_parts, int64_t part_stride_bytes, const int64_t* id_bases, float* D,
                                int64_t* I, int device, void* stream);

/* Row-sharded search, exchange over NVLink peer memory instead of a collective (no counterpart in the
 * reference).  b2_index_search_push runs the local search and, in the kernel that produces the final lists,
 * stores them straight into every rank's receive block: peer_blocks[g] = the address, mapped into THIS process,
 * of rank g's block for this rank's lists (layout [I: nq x k int64 local ids | D: nq x k fp32]); when a query's
 * k results are stored on all ranks, peer_flags[g][q] = epoch is released system-wide (peer_flags[g] = rank g's
 * nq flag words for this rank).  Nothing is written locally except through the entry for this rank itself.
 * b2_merge_shard_lists_wait is the receiving side: per query it waits until flags[g * flag_stride + q] == epoch
 * for all g, then merges the nparts blocks (blocks + g * part_stride_bytes) like b2_merge_shard_lists_packed.
 * The caller owns the peer-mapped memory (e.g. torch symmetric memory / cudaIpc), alternates between two sets of
 * blocks and flags from one search to the next, and increases `epoch` every search; every rank must call both
 * functions for every search (the wait spins on the GPU until all peers have delivered, traps after ~20 s).
 * At most 16 peers. */
B2_API int b2_index_search_push(b2_index* h, int64_t nq, const float* q_dev, int k, int nparts, void* const* peer_blocks,
                         uint32_t* const* peer_flags, uint32_t epoch, void* stream);
B2_API int b2_merge_shard_lists_wait(int metric, int64_t nq, int k, int nparts, const void* blocks,
                              int64_t part_stride_bytes, const uint32_t* flags, int64_t flag_stride, uint32_t epoch,
                              const int64_t* id_bases, float* D, int64_t* I, int device, void* stream);

/* Row-sharded search, query distribution over the same peer memory (no counterpart in the reference; it takes the
 * place of the ncclBroadcast SURVEY.md 8e names as step 1 of a sharded search).  The rank that owns the query batch
 * calls b2_query_push: one CTA per rank stores the nbytes of q_dev into peer_bufs[g] (rank g's receive buffer, mapped
 * into this process; the entry for the calling rank is its own local buffer) and then releases *peer_flags[g] = epoch
 * system-wide.  Every rank queues b2_flag_wait(its own flag, epoch) before the search that reads its receive buffer:
 * a one-warp kernel that returns when the flag shows the epoch (traps after ~20 s).  Both are asynchronous on
 * `stream`.  The caller alternates between two buffer / flag sets from one search to the next, as for the lists.
 * nbytes: a multiple of 16, at most B2_QUERY_PUSH_MAX_BYTES (1024 queries of 1024 fp32); <= 16 ranks. */
#define B2_QUERY_PUSH_MAX_BYTES (4 * 1024 * 1024)
B2_API int b2_query_push(const float* q_dev, int64_t nbytes, int nparts, void* const* peer_bufs, uint32_t* const* peer_flags,
                  uint32_t epoch, int device, void* stream);
B2_API int b2_flag_wait(const uint32_t* flag, uint32_t epoch, int device, void* stream);

/* A flat index sharded by rows over several GPUs of THIS process (b2_shard_group_create /
 * b2_shard_search; no counterpart in the reference, whose index lives on one host - the role of faiss.IndexShards
 * over per-GPU flat indexes).  Shard g lives on devices[g] (a device may appear more than once) and holds the
 * contiguous id range after shard g - 1's: rows are appended to shard g until it holds rows_per_shard rows, then to
 * shard g + 1; the last shard takes whatever follows.  Ids are faiss's implicit positions, results equal the single
 * index's bit for bit (same kernels per row, lowest id first among equal scores).  devices[0] is the ROOT: device
 * inputs / outputs of add and search live there.  Every pair of distinct devices needs peer access.  A search runs
 * the peer-memory protocol described above (query push, per-shard search with its lists stored into the root's
 * receive blocks, flag-waiting merge) on one stream per shard; calls on one group are enqueued one at a time.
 * b2_shard_group_remove_ids takes GLOBAL ids and renumbers like faiss (rows after a removed one move down). */
typedef struct b2_shard_group b2_shard_group;
B2_API int b2_shard_group_create(int ndev, const int* devices, int d, int metric, int dtype, int64_t rows_per_shard,
                          b2_shard_group** out);
B2_API void b2_shard_group_destroy(b2_shard_group* g);
B2_API int64_t b2_shard_group_ntotal(const b2_shard_group* g);
B2_API int b2_shard_group_nshards(const b2_shard_group* g);
B2_API int64_t b2_shard_group_shard_ntotal(const b2_shard_group* g, int shard);
B2_API int b2_shard_group_add(b2_shard_group* g, int64_t n, const float* x, int x_where, unsigned flags);
B2_API int b2_shard_group_remove_ids(b2_shard_group* g, int64_t n, const int64_t* ids, int64_t* n_removed);
B2_API int b2_shard_group_reset(b2_shard_group* g);
/* index.reconstruct over the group: out[j] = row ids[j] (GLOBAL ids, host arrays, n x d fp32); ids outside the store
 * give zero rows. */
B2_API int b2_shard_group_reconstruct_batch(b2_shard_group* g, int64_t n, const int64_t* ids, float* out);
/* b2_index_set_option on every shard (e.g. "bf16_residual" on an empty bf16 group, "bf16_shadow", "gemm_query_parts"). */
B2_API int b2_shard_group_set_option(b2_shard_group* g, const char* name, int64_t value);
B2_API int b2_shard_search(b2_shard_group* g, int64_t nq, const float* q, int q_where, int k, float* D, int64_t* I,
                    int out_where, void* stream);

/* faiss.normalize_L2(x) on a caller-owned device array (n x d fp32, in place). */
B2_API int b2_normalize_l2(float* x, int64_t n, int d, int device, void* stream);

/* Counter-based synthetic embeddings, written straight
 * into device memory: rows row0..row0+nrows-1 of stream `seed`, n x d fp32.
 * Bit-identical to oracle/flat_oracle.c b2o_synth_fill.  Bench/test support. */
B2_API int b2_synth_fill(float* out, int64_t row0, int64_t nrows, int d, uint64_t seed, int device, void* stream);

/* Kernel launches issued by this library (process-wide) since the last
 * reset (bench.py's gpu_launches). */
B2_API int64_t b2_launch_count(void);
B2_API void b2_launch_count_reset(void);

/* Measurement support for bench.py (no counterpart in the reference): when enabled, every
 * launch of the dominant scan kernel is bracketed by CUDA events on its own stream;
 * b2_prof_read synchronises those events and returns the summed kernel time and the number
 * of launches since the last reset. */
B2_API int b2_prof_enable(int on);
B2_API int b2_prof_read(double* scan_ms, int64_t* scan_launches, int reset);
/* The same per kernel: B2_PROF_SCAN = the small-batch scan kernel, B2_PROF_GEMM = the tensor-core kernel. */
#define B2_PROF_SCAN 0
#define B2_PROF_GEMM 1
B2_API int b2_prof_read_kernel(int kernel, double* ms, int64_t* launches, int reset);

/* Thread-local message of the most recent failure on this thread ("" if none). */
B2_API const char* b2_last_error(void);

/* Library version string. */
B2_API const char* b2_version(void);

#ifdef __cplusplus
}
#endif
#endif /* B2_FLAT_H */
