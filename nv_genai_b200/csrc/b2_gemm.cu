// Large-batch exact search on the 5th-generation tensor cores (tcgen05 + TMEM), top-k fused
// into the accumulator read-out so the nq x ntotal score matrix never reaches HBM.
//
// Replaces faiss exhaustive_inner_product_blas / exhaustive_L2sqr_blas + the heap / reservoir
// result handlers (faiss/utils/distances.cpp, the nq >= distance_compute_blas_threshold = 20
// path of IndexFlat::search; reached from every batched index.search of the reference, e.g.
// /root/reference/RetrievalAugmentedGeneration/examples/nvidia_api_catalog/chains.py:122-123
// when requests are batched, and the hard-negative mining matmul at
// /root/reference/experimental/synthetic-data-retriever-customization/retriever_customization.ipynb cell 21).
// Like faiss's BLAS path, L2 is |q|^2 + |x|^2 - 2 q.x clamped at 0 (the |x|^2 come from the
// ingest kernel's side array), IP is the plain dot product.
//
// Default mode = FILTER AND REFINE (gemm_search_refine below): ONE approximate product per score on the tensor cores
//   (bf16 rows: bf16(q).x; fp32 rows: bf16(q).bf16(x) over a bf16 shadow of the rows, or tf32 straight from the raw
//   fp32 tile when there is no shadow), a rigorous per-query error band (band_kernel), every candidate within the band
//   of the running k-th best logged, and merge_rescore (b2_merge.cu) re-scoring the candidates near the final cut in
//   fp32, so ids and scores equal the all-exact arithmetic.  Lists that cannot hold a band are flagged and redone by
//   a second launch with the ALL-EXACT arithmetic described next (also selected by B2_GEMM_REFINE=0 or the
//   gemm_query_parts option).
//
// Mapping (one persistent CTA per SM, warp-specialised: TMA producer warp, MMA issuer warp, four
// epilogue warps, and for the exact fp32 arithmetic eight transform warps):
//   M = 128 queries per CTA (TMEM lanes), N = 256 corpus rows (TMEM columns), K = the row,
//   streamed in 64- or 128-byte slices.  Both operands are K-major, swizzled, brought in by TMA:
//     A = the query tile, prepared by prep_queries_kernel,
//     B = corpus rows straight from the row store (or its bf16 shadow).
//   bf16 rows (kind::f16, 2 products):  D += q_hi.x + q_lo.x with q split into bf16 hi + lo, so q is
//     carried to ~16 mantissa bits and the scores are exact with respect to the stored rows.  CTAs run
//     as PAIRS (cta_group::2): one M = 256 instruction stream issued by the leader CTA, each CTA holds
//     its 128 queries, half of the corpus tile and its own 128 TMEM lanes (see the kernel comment).
//     Accumulators are double-buffered in TMEM (2 x 256 columns).
//   fp32 rows (fp32-faithful, "3xTF32" with the corrections moved to the bf16 pipe):
//     main  D1 += x_hi.q_hi   kind::tf32; the raw fp32 tile IS x_hi (the tensor core ignores the low
//                             13 mantissa bits of a tf32 operand), q_hi = tf32(q);
//     corr  D2 += [x_lo | x_hi] . [q_hi | q_lo]   ONE kind::f16 MMA over a doubled K: eight transform
//                             warps write, per K slice and row, bf16(x - trunc_tf32(x)) | bf16(x) in the
//                             same swizzled layout; the correction terms are 2^-11 of the main term, so
//                             bf16 operands cost 2^-20 relative.  (Plain 3xTF32, B2_GEMM_MIXED=0: 422 ms
//                             vs 388 ms.)  The correction tiles live in their own 2-slot ring, so a TMA
//                             stage is only raw A + raw B and one more stage fits: the fp32 path is bound
//                             by the depth of the TMA pipeline, not by the tensor pipe.
//     The tensor core adds each product into the fp32 accumulator with truncation, a bias that grows
//     with the number of accumulation steps (measured: -1e-5 relative at d = 1536 with all three
//     products in one accumulator), hence the separate accumulators: columns [0,256) main, [256,512)
//     corrections, added in fp32 registers by the epilogue.  That uses all 512 TMEM columns, so the
//     fp32 path is single-buffered (tile time ~33k cycles vs ~2k for the read-out).
//   Epilogue: tcgen05.ld (software-pipelined over 32-column chunks), one thread = one query = one TMEM lane.  A max
//   tree rejects the chunks without a candidate for any of the warp's 32 queries; survivors are appended by the
//   query's own lane to its key RESERVOIR in shared memory (rcap > k slots, faiss's ReservoirResultHandler scheme) and
//   to the query's candidate log in global memory; a full reservoir is reduced to its k best by the warp (bisection on
//   the key bits) and the k-th best becomes the query's threshold.
//   Work item = (query tile, corpus segment); items are dealt round-robin to the CTAs with the
//   query tile varying fastest, so the CTAs that run at the same time read the same corpus
//   segment and the corpus is streamed from HBM about once.  A finished item publishes its
//   k-th best key per query (atomicMax); later items of the same query tile start from that
//   bound, and running items trade bounds every 32 tiles.  Every item emits k (refine: k + slack) candidates per
//   query; merge_partials_kernel / merge_rescore finish.
//   For k >= 32 on a long corpus the refine search is TWO launches: a threshold pre-pass over every 32nd corpus tile,
//   tau_union_kernel (the k-th best key of the union of the pre-pass lists: a lower bound of the final k-th, since it is
//   the k-th best of a subset of the rows), then the main pass over the other tiles, whose items all start from that
//   bound instead of -inf (see gemm_search_refine).
//
// Tie rule: equal keys keep the lower row id (a key has to be strictly better than the running
// k-th to enter, rows arrive in id order, and bounds imported from other segments admit ties).
//
// Algorithmic work per launch: 2 * nq_pad * ntotal * K flop per product (x1 filter-and-refine; all-exact: x2 bf16, x3 fp32);
// corpus bytes ntotal * pitch read ~once from HBM, query tile re-read from L2 per corpus tile.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "b2_kernels.h"
#include "b2_tc.cuh"

namespace b2 {

using namespace tc;

namespace {

constexpr int kTileM = 128;                        // queries per tile
constexpr int kTileN = 256;                        // corpus rows per tile
constexpr int kMaxStages = 8;
constexpr int kMaxCorrSlots = 4;                   // fp32 rows: ring of correction-operand tiles, decoupled from the TMA stages
constexpr int kCandCap = 1024;                     // candidate log entries per query per CTA
constexpr int kGemmMaxK = 128;                     // heap slots per query (k <= 128 on this path)
constexpr int kEpiThreads = 128;
constexpr int kXfThreads = 256;                    // fp32 rows: transform threads (warps 6..13)
constexpr int kThreadsBf16 = 192, kThreadsFp32 = 192 + kXfThreads;

struct GemmParams {
    int nq, nq_pad;
    uint32_t n;
    int k;
    int kstages;        // K slices per tile
    int metric;
    int n_qt, n_seg, tiles_per_seg, n_tiles, n_items;  // n_tiles = LOGICAL tiles of this launch (see tile_of)
    int tile_mul;       // logical tile u -> corpus tile u * tile_mul            (threshold pre-pass: every tile_mul-th tile)
    int tile_skip;      // != 0: logical tile u -> u + u / (tile_skip - 1) + 1   (main pass: all but every tile_skip-th tile)
    int n_lists;        // candidate lists per query in `partial`; this launch writes lists list_base .. list_base + n_seg - 1
    int list_base;
    int num_stages;
    int corr_slots;     // fp32 rows: correction tiles in the ring (2..4)
    int trunc_hi;       // fp32 rows: 1 = leave the raw tile as the hi operand (hardware truncation), 0 = round it in place
    int mixed;          // fp32 rows: 1 = correction products on the bf16 pipe (see kernel comment), 0 = plain 3xTF32
    int refine;         // 1 = single approximate product; log everything within `band` of the running k-th, exact
                        //     re-scoring happens in merge_rescore_kernel (see "Filter and refine" above)
    int rcap;           // slots of a query's key reservoir in shared memory (k < rcap <= kMaxRcap, row stride rcap + 1)
    int list_len;       // entries emitted per (query, segment): k, or k + slack when refine
    int only_flagged;   // fix-up launch: process only query groups whose flag is set
    const float* band;  // [nq_pad] refine: 2 x (bound on |approximate key - exact key|), per query
    unsigned long long* stats;  // diagnostics (B2_GEMM_DEBUG): [0] MMA cycles waiting for a free accumulator, [1] for operands,
                        // [2] epilogue warp 0 cycles waiting for a full accumulator, [3] reservoir admissions, [4] log entries,
                        // [5] log compactions, [6] MMA warp cycles in total, [7] as [0] but only inside a unit's first item,
                        // [8] reservoir selections; nullptr = off
    uint32_t* flags;    // [n_qt] refine: set when a query of the group had more in-band candidates than fit
    const float* norms; // |x|^2 per stored row
    uint64_t* cand;     // [grid][kCandCap][128]
    uint64_t* partial;  // [nq_pad][n_lists][list_len]
    uint32_t* tau;      // [nq_pad] ordered image of an exclusive lower bound of the final k-th key
};


// Register-only v[c] for a run-time c (a dynamically indexed array would live in local memory).
__device__ __forceinline__ float select32(const float (&v)[32], int c) {
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = (c & 16) ? v[16 + i] : v[i];
    float b[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) b[i] = (c & 8) ? a[8 + i] : a[i];
    float e[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) e[i] = (c & 4) ? b[4 + i] : b[i];
    const float f0 = (c & 2) ? e[2] : e[0], f1 = (c & 2) ? e[3] : e[1];
    return (c & 1) ? f1 : f0;
}

// The running k best keys of a query live in a RESERVOIR of `rcap` > k slots in shared memory (the scheme of faiss's
// ReservoirResultHandler, which faiss itself switches to at k >= 100): a key above the query's threshold is appended
// by the query's own lane - one shared-memory store, no warp cooperation - and only when the reservoir is full does
// the warp select the k-th largest of its entries, keep the k best and raise the threshold to the k-th.  The
// threshold therefore lags the true running k-th by at most rcap - k admissions (10-15 % more admissions than a
// threshold updated on every insertion), and the cost per admitted key drops from a 400-600 cycle dependent chain
// (replace the minimum, re-scan k entries for the new one) to a store plus 1 / (rcap - k) of a selection.
//
// coop_compact: all 32 lanes; R[0..n) are the reservoir entries of ONE query (n >= k, every entry > -inf).  Finds the
// k-th largest by bisection on the order-preserving integer image of the keys, two bits per step (three counts packed
// into one redux), starting below the leading bits all entries share; then moves the k best to R[0..k) (the entries
// equal to the k-th fill up to exactly k) and returns the k-th largest.
constexpr int kResPerLane = 6;   // rcap <= 192
constexpr int kMaxRcap = 32 * kResPerLane;
__device__ __forceinline__ float coop_compact(float* __restrict__ R, int n, int k, int lane) {
    uint32_t e[kResPerLane];
    uint32_t mx = 0u, mn = 0xffffffffu;
#pragma unroll
    for (int i = 0; i < kResPerLane; ++i) {
        const int idx = lane + 32 * i;
        e[i] = idx < n ? f2ord(R[idx]) : 0u;  // images of keys > -inf are > 0: 0 marks "no entry"
        mx = max(mx, e[i]);
        if (e[i] != 0u) mn = min(mn, e[i]);
    }
    mx = __reduce_max_sync(kFull, mx);
    mn = __reduce_min_sync(kFull, mn);
    uint32_t prefix = mx;  // all entries equal
    const uint32_t diff = mx ^ mn;
    if (diff != 0u) {
        int bit = 31 - __clz(diff);                               // highest bit in which two entries differ
        prefix = (bit == 31) ? 0u : (mx & ~((2u << bit) - 1u));   // the bits above it are common to all entries
        for (; bit >= 1; bit -= 2) {
            const int sh = bit - 1;
            const uint32_t c1 = prefix | (1u << sh), c2 = prefix | (2u << sh), c3 = prefix | (3u << sh);
            uint32_t cnt = 0u;
#pragma unroll
            for (int i = 0; i < kResPerLane; ++i)
                cnt += (e[i] >= c1 ? 1u : 0u) + (e[i] >= c2 ? (1u << 10) : 0u) + (e[i] >= c3 ? (1u << 20) : 0u);
            const uint32_t tot = __reduce_add_sync(kFull, cnt);
            if ((int)(tot >> 20) >= k) prefix = c3;
            else if ((int)((tot >> 10) & 1023u) >= k) prefix = c2;
            else if ((int)(tot & 1023u) >= k) prefix = c1;
        }
        if (bit == 0) {
            const uint32_t c1 = prefix | 1u;
            uint32_t cnt = 0u;
#pragma unroll
            for (int i = 0; i < kResPerLane; ++i) cnt += e[i] >= c1 ? 1u : 0u;
            if ((int)__reduce_add_sync(kFull, cnt) >= k) prefix = c1;
        }
    }
    // prefix = image of the k-th largest entry.  All entries are in registers: overwrite the row in place.
    __syncwarp();
    const uint32_t lt = (1u << lane) - 1u;
    int off = 0;
#pragma unroll
    for (int i = 0; i < kResPerLane; ++i) {
        const bool g = e[i] > prefix;
        const uint32_t b = __ballot_sync(kFull, g);
        if (g) R[off + __popc(b & lt)] = ord2f(e[i]);
        off += __popc(b);
    }
#pragma unroll
    for (int i = 0; i < kResPerLane; ++i) {
        const bool q = e[i] == prefix;
        const uint32_t b = __ballot_sync(kFull, q);
        const int pos = off + __popc(b & lt);
        if (q && pos < k) R[pos] = ord2f(prefix);
        off += __popc(b);
    }
    __syncwarp();
    return ord2f(prefix);
}

// Keep exactly the current top-k of this thread's log, in arrival order: keys above tau, plus the earliest
// entries equal to tau.  Called by all lanes of a warp together (the log is [entry][128 queries], so lanes
// read neighbouring words); lanes with !active only keep the loops uniform.
__device__ __forceinline__ void compact_log(uint64_t* __restrict__ log, int& cnt, float tau, int k, bool active) {
    int cmax = active ? cnt : 0;
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) cmax = max(cmax, __shfl_xor_sync(kFull, cmax, s));
    const uint32_t to = f2ord(tau);
    const int n = active ? cnt : 0;
    int n_gt = 0;
    for (int j0 = 0; j0 < cmax; j0 += 8) {
        uint64_t e[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) e[u] = (j0 + u < n) ? log[(size_t)(j0 + u) * kEpiThreads] : 0ull;
#pragma unroll
        for (int u = 0; u < 8; ++u) n_gt += (j0 + u < n && (uint32_t)(e[u] >> 32) > to) ? 1 : 0;
    }
    int eq_left = k - n_gt;
    int w = 0;
    for (int j0 = 0; j0 < cmax; j0 += 8) {
        uint64_t e[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) e[u] = (j0 + u < n) ? log[(size_t)(j0 + u) * kEpiThreads] : 0ull;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            if (j0 + u < n) {
                const uint32_t ko = (uint32_t)(e[u] >> 32);
                bool keep = ko > to;
                if (ko == to && eq_left > 0) {
                    keep = true;
                    --eq_left;
                }
                if (keep) {
                    log[(size_t)w * kEpiThreads] = e[u];  // w <= j0 + u: never overtakes the reads
                    ++w;
                }
            }
        }
    }
    if (active) cnt = w;
}

// Refine mode: keep the log entries whose approximate key is above `thr_log` (arrival order preserved).  Called by
// all lanes of a warp together, like compact_log.
__device__ __forceinline__ void compact_band(uint64_t* __restrict__ log, int& cnt, float thr_log, bool active) {
    int cmax = active ? cnt : 0;
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) cmax = max(cmax, __shfl_xor_sync(kFull, cmax, s));
    const uint32_t to = f2ord(thr_log);
    const int n = active ? cnt : 0;
    int w = 0;
    for (int j0 = 0; j0 < cmax; j0 += 8) {
        uint64_t e[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) e[u] = (j0 + u < n) ? log[(size_t)(j0 + u) * kEpiThreads] : 0ull;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            if (j0 + u < n && (uint32_t)(e[u] >> 32) > to) {
                log[(size_t)w * kEpiThreads] = e[u];
                ++w;
            }
        }
    }
    if (active) cnt = w;
}

// Corpus tile of a launch's logical tile u.  One launch covers either every tile, or every tile_mul-th tile (the
// threshold pre-pass over a sample of the rows), or all the others (the main pass after such a pre-pass).
__device__ __forceinline__ int tile_of(const GemmParams& p, int u) {
    return p.tile_skip ? u + u / (p.tile_skip - 1) + 1 : u * p.tile_mul;
}

// ------------------------------------------------------------------ the kernel
// KIND 0: bf16 rows, kind::f16.  KIND 1: fp32 rows, fp32-faithful (tf32 main + bf16 correction, transform warps).
// KIND 2: fp32 rows, ONE tf32 product straight from the raw tile (approximate; only with refine).
// NA: query parts (1 = hi only, 2 = hi + lo).
// SWB: bytes of K per pipeline stage = the shared-memory swizzle width (64 or 128).
// CG: CTAs per MMA (tcgen05 cta_group).  CG = 2 launches clusters of two CTAs on the two SMs of a TPC: the pair
// computes a 256-query x 256-row tile with ONE M = 256 instruction stream issued by the leader CTA; each CTA holds its
// own 128 queries (A), HALF of the corpus tile (B, 128 rows) and its own 128 TMEM lanes of the accumulators.  Per CTA
// that halves the B bytes pulled from L2, the transform work and the shared-memory traffic of the MMAs (the fp32 path
// was shared-memory-bandwidth-bound with one CTA per MMA) and makes room for more pipeline stages.
template <int KIND, int NA, int SWB, int CG>
__global__ void __launch_bounds__(KIND == 1 ? kThreadsFp32 : kThreadsBf16, 1)
gemm_topk_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmParams p) {
    constexpr int ELEM = (KIND == 0) ? 2 : 4;
    constexpr int kSwizzleBytes = SWB;
    constexpr int kAPartBytes = kTileM * SWB;
    constexpr int kBRows = kTileN / CG;                   // corpus rows of the tile held by this CTA
    constexpr int kBTileBytes = kBRows * SWB;
    constexpr int KSLICE = kSwizzleBytes / ELEM;          // elements per stage along K
    constexpr int STAGE_BYTES = NA * kAPartBytes + kBTileBytes;  // what TMA fills; fp32 rows add kCorrSlots correction tiles
    constexpr uint32_t STAGE_TX = NA * kAPartBytes + kBTileBytes;
    constexpr int B_OFF = NA * kAPartBytes;
    constexpr uint32_t NBUF = (KIND == 1) ? 1u : 2u;  // accumulator buffers (KIND 1: main + correction use all of TMEM)
    constexpr int MK = (KIND == 0) ? 0 : 1;          // MMA kind: f16 or tf32
    // instruction descriptor: D fp32, A/B bf16 or tf32, both K-major, N = 256, M = 128
    constexpr uint32_t FMT = (KIND == 0) ? 1u : 2u;
    constexpr uint32_t IDESC = (1u << 4) | (FMT << 7) | (FMT << 10) | ((uint32_t)(kTileN >> 3) << 17) |
                               ((uint32_t)((kTileM * CG) >> 4) << 24);
    constexpr uint32_t IDESC_BF16 = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kTileN >> 3) << 17) |
                                    ((uint32_t)((kTileM * CG) >> 4) << 24);
    // bf16 rows, CG = 2: both CTAs' TMA loads complete on the LEADER's full barrier (no transform stage in between)
    constexpr bool PAIR_FULL = (CG == 2 && KIND != 1);

    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t pad = ((raw + 1023u) & ~1023u) - raw;
    uint8_t* smem = smem_raw + pad;
    const int S = p.num_stages;
    uint8_t* stages = smem;
    uint8_t* corr_tiles = stages + (size_t)S * STAGE_BYTES;                              // [kCorrSlots][B tile] (fp32 rows)
    float* heap_s = reinterpret_cast<float*>(corr_tiles + (KIND == 1 ? p.corr_slots * kBTileBytes : 0));  // [128 queries][rcap + 1] key reservoirs
    float* nrm_s = heap_s + (size_t)(p.rcap + 1) * kEpiThreads;                           // [2][256]
    uint64_t* bars = reinterpret_cast<uint64_t*>(nrm_s + 2 * kTileN);
    uint64_t* full_bar = bars;                    // [S] TMA -> MMA / transform
    uint64_t* empty_bar = bars + kMaxStages;      // [S] MMA -> TMA
    uint64_t* xf_bar = bars + 2 * kMaxStages;     // [S] transform -> MMA (fp32 rows)
    uint64_t* tfull_bar = bars + 3 * kMaxStages;  // [2] MMA -> epilogue
    uint64_t* tempty_bar = tfull_bar + 2;         // [2] epilogue -> MMA
    uint64_t* pxf_bar = tempty_bar + 2;           // [S] CTA pairs, fp32 rows: the peer's transform is done (leader's copy)
    uint64_t* corr_empty_bar = pxf_bar + kMaxStages;  // [kCorrSlots] MMA -> transform (fp32 rows)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(corr_empty_bar + kMaxCorrSlots);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t cta_rank = (CG == 2) ? cluster_ctarank() : 0u;  // 0 = leader (issues the MMAs)
    const int unit = blockIdx.x / CG, n_units = gridDim.x / CG;    // a unit = one CTA, or one CTA pair

    if (tid == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(smem_u32(full_bar + s), 1);
            mbar_init(smem_u32(empty_bar + s), 1);
            mbar_init(smem_u32(xf_bar + s), kXfThreads / 32);  // this CTA's transform warps
            mbar_init(smem_u32(pxf_bar + s), 1);               // one forwarded arrival from the peer CTA
        }
        for (int s = 0; s < kMaxCorrSlots; ++s) mbar_init(smem_u32(corr_empty_bar + s), 1);
        for (int b = 0; b < 2; ++b) {  // fp32 rows use only buffer 0
            mbar_init(smem_u32(tfull_bar + b), 1);
            mbar_init(smem_u32(tempty_bar + b), CG * 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 5) {
        if (CG == 2) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                         "r"(512u)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                         "r"(512u)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync_all();  // the peer's barriers are initialised before anything is signalled across
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 4) {
        // ================= TMA producer (whole warp loops, one elected lane issues; see tc::elect_one)
        int stage = 0;
        uint32_t phase = 0;
        for (int item = unit; item < p.n_items; item += n_units) {
            if (p.only_flagged && __ldcg(p.flags + item % p.n_qt) == 0u) continue;  // fix-up launch: flagged groups only
            const int qt = (item % p.n_qt) * CG + (int)cta_rank, seg = item / p.n_qt;
            const int t0 = seg * p.tiles_per_seg;
            const int t1 = min(p.n_tiles, t0 + p.tiles_per_seg);
            for (int u = t0; u < t1; ++u) {
                const int t = tile_of(p, u);
                for (int ks = 0; ks < p.kstages; ++ks) {
                    mbar_wait(smem_u32(empty_bar + stage), phase ^ 1u);
                    if (elect_one()) {
                        const uint32_t fb = smem_u32(full_bar + stage);
                        const uint32_t sb = smem_u32(stages + (size_t)stage * STAGE_BYTES);
                        const int x = ks * KSLICE;
                        const int brow = t * kTileN + (int)cta_rank * kBRows;
                        if (PAIR_FULL) {
                            const uint32_t lfb = map_to_cta(fb, 0);  // the leader's barrier counts both CTAs' bytes
                            if (cta_rank == 0) mbar_expect_tx(fb, 2 * STAGE_TX);
                            tma_load_2d_pair(sb, &tmA, x, qt * kTileM, lfb);
                            if (NA == 2) tma_load_2d_pair(sb + kAPartBytes, &tmA, x, p.nq_pad + qt * kTileM, lfb);
                            tma_load_2d_pair(sb + B_OFF, &tmB, x, brow, lfb);
                        } else {
                            mbar_expect_tx(fb, STAGE_TX);
                            tma_load_2d(sb, &tmA, x, qt * kTileM, fb);
                            if (NA == 2) tma_load_2d(sb + kAPartBytes, &tmA, x, p.nq_pad + qt * kTileM, fb);
                            tma_load_2d(sb + B_OFF, &tmB, x, brow, fb);
                        }
                    }
                    __syncwarp();
                    if (++stage == S) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
        }
    } else if (warp == 5) {
        // ================= MMA issuer (whole warp loops, one elected lane issues; CG = 2: the leader CTA only)
        int stage = 0;
        uint32_t phase = 0;
        uint32_t tc = 0;
        int cslot = 0;
        long long w_acc = 0, w_ops = 0, w_acc_first = 0;
        const long long c_begin = p.stats ? clock64() : 0;
        for (int item = unit; cta_rank == 0 && item < p.n_items; item += n_units) {
            if (p.only_flagged && __ldcg(p.flags + item % p.n_qt) == 0u) continue;  // fix-up launch: flagged groups only
            const int seg = item / p.n_qt;
            const int t0 = seg * p.tiles_per_seg;
            const int t1 = min(p.n_tiles, t0 + p.tiles_per_seg);
            for (int t = t0; t < t1; ++t, ++tc) {
                const uint32_t buf = tc % NBUF, par = (tc / NBUF) & 1u;
                const long long c0 = p.stats ? clock64() : 0;
                mbar_wait(smem_u32(tempty_bar + buf), par ^ 1u);
                if (p.stats) {
                    w_acc += clock64() - c0;
                    if (item == unit) w_acc_first += clock64() - c0;
                }
                tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + buf * kTileN;
                const uint32_t c_tmem = tmem_base + kTileN;  // fp32 rows: correction accumulator
                for (int ks = 0; ks < p.kstages; ++ks) {
                    const long long c1 = p.stats ? clock64() : 0;
                    mbar_wait(smem_u32((KIND == 1 ? xf_bar : full_bar) + stage), phase);
                    if (CG == 2 && KIND == 1) mbar_wait(smem_u32(pxf_bar + stage), phase);
                    if (p.stats) w_ops += clock64() - c1;
                    tcgen05_fence_after();
                    if (elect_one()) {
                        const uint32_t sb = smem_u32(stages + (size_t)stage * STAGE_BYTES);
                        const uint64_t a_hi0 = smem_desc<SWB>(sb);
                        const uint64_t a_lo0 = smem_desc<SWB>(sb + kAPartBytes);
                        const uint64_t b_hi0 = smem_desc<SWB>(sb + B_OFF);
                        const uint64_t b_lo0 = smem_desc<SWB>(smem_u32(corr_tiles + (size_t)cslot * kBTileBytes));
#pragma unroll
                        for (int j = 0; j < kSwizzleBytes / 32; ++j) {
                            const uint32_t acc = (ks | j) ? 1u : 0u;
                            const uint64_t dj = (uint64_t)(j * 2);  // +32 bytes in the descriptor's address field
                            if (CG == 2) {
                                umma_pair<MK>(d_tmem, a_hi0 + dj, b_hi0 + dj, IDESC, acc);
                                if (KIND != 1) {
                                    if (NA == 2) umma_pair<MK>(d_tmem, a_lo0 + dj, b_hi0 + dj, IDESC, 1u);
                                } else if (p.mixed) {
                                    if (NA == 2) umma_pair<0>(c_tmem, a_lo0 + dj, b_lo0 + dj, IDESC_BF16, acc);
                                } else {
                                    umma_pair<MK>(c_tmem, a_hi0 + dj, b_lo0 + dj, IDESC, acc);
                                    if (NA == 2) umma_pair<MK>(c_tmem, a_lo0 + dj, b_hi0 + dj, IDESC, 1u);
                                }
                                continue;
                            }
                            umma<MK>(d_tmem, a_hi0 + dj, b_hi0 + dj, IDESC, acc);
                            if (KIND != 1) {
                                if (NA == 2) umma<MK>(d_tmem, a_lo0 + dj, b_hi0 + dj, IDESC, 1u);
                            } else if (p.mixed) {
                                // both correction products in ONE bf16 MMA: [q_hi | q_lo] . [x_lo | x_hi] over a doubled K
                                if (NA == 2) umma<0>(c_tmem, a_lo0 + dj, b_lo0 + dj, IDESC_BF16, acc);
                            } else {
                                umma<MK>(c_tmem, a_hi0 + dj, b_lo0 + dj, IDESC, acc);
                                if (NA == 2) umma<MK>(c_tmem, a_lo0 + dj, b_hi0 + dj, IDESC, 1u);
                            }
                        }
                        if (CG == 2) tcgen05_commit_pair(smem_u32(empty_bar + stage));
                        else tcgen05_commit(smem_u32(empty_bar + stage));
                        if (KIND == 1) {
                            if (CG == 2) tcgen05_commit_pair(smem_u32(corr_empty_bar + cslot));
                            else tcgen05_commit(smem_u32(corr_empty_bar + cslot));
                        }
                    }
                    __syncwarp();
                    if (KIND == 1 && ++cslot == p.corr_slots) cslot = 0;
                    if (++stage == S) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
                if (elect_one()) {
                    if (CG == 2) tcgen05_commit_pair(smem_u32(tfull_bar + buf));
                    else tcgen05_commit(smem_u32(tfull_bar + buf));
                }
                __syncwarp();
            }
        }
        if (p.stats && cta_rank == 0 && lane == 0) {
            atomicAdd(p.stats + 0, (unsigned long long)w_acc);
            atomicAdd(p.stats + 1, (unsigned long long)w_ops);
            atomicAdd(p.stats + 6, (unsigned long long)(clock64() - c_begin));
            atomicAdd(p.stats + 7, (unsigned long long)w_acc_first);
        }
        // CTA pairs, fp32 rows: the peer's MMA warp has nothing to issue; it forwards "my transform warps are done with
        // this K slice" to the leader with ONE cluster-scope arrival per slice (sixteen cluster-scope arrivals per
        // slice, one per transform warp of both CTAs, made the paired fp32 path slower than the unpaired one)
        if (CG == 2 && KIND == 1 && cta_rank != 0) {
            int fstage = 0;
            uint32_t fphase = 0;
            for (int item = unit; item < p.n_items; item += n_units) {
                if (p.only_flagged && __ldcg(p.flags + item % p.n_qt) == 0u) continue;  // fix-up launch: flagged groups only
                const int seg = item / p.n_qt;
                const int t0 = seg * p.tiles_per_seg;
                const int t1 = min(p.n_tiles, t0 + p.tiles_per_seg);
                for (int t = t0; t < t1; ++t) {
                    for (int ks = 0; ks < p.kstages; ++ks) {
                        mbar_wait(smem_u32(xf_bar + fstage), fphase);
                        if (elect_one()) mbar_arrive_cluster(map_to_cta(smem_u32(pxf_bar + fstage), 0));
                        __syncwarp();
                        if (++fstage == S) {
                            fstage = 0;
                            fphase ^= 1u;
                        }
                    }
                }
            }
        }
    } else if (warp >= 6) {
        // ================= fp32 rows: split the landed B slice into tf32 hi (in place) + lo
        if (KIND == 1) {
            const int xt = tid - 192;  // 0..kXfThreads-1
            int stage = 0;
            uint32_t phase = 0;
            int cslot = 0;
            uint32_t cphase = 0;
            for (int item = unit; item < p.n_items; item += n_units) {
                if (p.only_flagged && __ldcg(p.flags + item % p.n_qt) == 0u) continue;  // fix-up launch: flagged groups only
                const int seg = item / p.n_qt;
                const int t0 = seg * p.tiles_per_seg;
                const int t1 = min(p.n_tiles, t0 + p.tiles_per_seg);
                for (int t = t0; t < t1; ++t) {
                    for (int ks = 0; ks < p.kstages; ++ks) {
                        mbar_wait(smem_u32(full_bar + stage), phase);
                        mbar_wait(smem_u32(corr_empty_bar + cslot), cphase ^ 1u);
                        float4* b_hi = reinterpret_cast<float4*>(stages + (size_t)stage * STAGE_BYTES + B_OFF);
                        float4* b_lo = reinterpret_cast<float4*>(corr_tiles + (size_t)cslot * kBTileBytes);
                        if (p.mixed) {
                            // correction operand: per row, [bf16(x - trunc_tf32(x)) | bf16(x)] over this K slice, written
                            // in the same swizzled K-major layout (16-byte chunk index XOR a row-dependent mask)
                            constexpr int ITEMS_PER_ROW = SWB / 32;  // 8 floats in, one lo chunk + one hi chunk out
#pragma unroll
                            for (int i = 0; i < kBRows * ITEMS_PER_ROW / kXfThreads; ++i) {
                                const int w = i * kXfThreads + xt;
                                const int r = w / ITEMS_PER_ROW, g = w % ITEMS_PER_ROW;
                                const int sw = (SWB == 128) ? (r & 7) : ((r >> 1) & 3);
                                const int rowc = r * (SWB / 16);  // first chunk of the row
                                const float4 x0 = b_hi[rowc + ((2 * g) ^ sw)];
                                const float4 x1 = b_hi[rowc + ((2 * g + 1) ^ sw)];
                                const float xs[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
                                uint32_t lo[4], hi[4];
#pragma unroll
                                for (int e = 0; e < 4; ++e) {
                                    const float a = xs[2 * e], b = xs[2 * e + 1];
                                    const float la = a - __uint_as_float(__float_as_uint(a) & 0xffffe000u);
                                    const float lb = b - __uint_as_float(__float_as_uint(b) & 0xffffe000u);
                                    __nv_bfloat162 pl = __floats2bfloat162_rn(la, lb), ph = __floats2bfloat162_rn(a, b);
                                    lo[e] = *reinterpret_cast<uint32_t*>(&pl);
                                    hi[e] = *reinterpret_cast<uint32_t*>(&ph);
                                }
                                b_lo[rowc + (g ^ sw)] = make_float4(__uint_as_float(lo[0]), __uint_as_float(lo[1]),
                                                                    __uint_as_float(lo[2]), __uint_as_float(lo[3]));
                                b_lo[rowc + ((ITEMS_PER_ROW + g) ^ sw)] =
                                    make_float4(__uint_as_float(hi[0]), __uint_as_float(hi[1]), __uint_as_float(hi[2]),
                                                __uint_as_float(hi[3]));
                            }
                        } else {
#pragma unroll
                        for (int i = 0; i < kBTileBytes / 16 / kXfThreads; ++i) {
                            const int idx = i * kXfThreads + xt;
                            const float4 x = b_hi[idx];
                            float4 h, l;
                            if (p.trunc_hi) {
                                // the tensor core ignores the low 13 mantissa bits of a tf32 operand, so the raw
                                // fp32 tile already IS the hi part; only lo = x - trunc(x) has to be written
                                h.x = __uint_as_float(__float_as_uint(x.x) & 0xffffe000u);
                                h.y = __uint_as_float(__float_as_uint(x.y) & 0xffffe000u);
                                h.z = __uint_as_float(__float_as_uint(x.z) & 0xffffe000u);
                                h.w = __uint_as_float(__float_as_uint(x.w) & 0xffffe000u);
                            } else {
                                uint32_t u;
                                asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x.x)); h.x = __uint_as_float(u);
                                asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x.y)); h.y = __uint_as_float(u);
                                asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x.z)); h.z = __uint_as_float(u);
                                asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x.w)); h.w = __uint_as_float(u);
                                b_hi[idx] = h;
                            }
                            l.x = x.x - h.x; l.y = x.y - h.y; l.z = x.z - h.z; l.w = x.w - h.w;
                            b_lo[idx] = l;
                        }
                        }
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) mbar_arrive(smem_u32(xf_bar + stage));  // local; the peer's is forwarded (pxf_bar)
                        if (++cslot == p.corr_slots) {
                            cslot = 0;
                            cphase ^= 1u;
                        }
                        if (++stage == S) {
                            stage = 0;
                            phase ^= 1u;
                        }
                    }
                }
            }
        }
    } else {
        // ================= epilogue: warps 0..3, thread = query = TMEM lane
        constexpr float kNegInf = -__builtin_huge_valf();
        const int ql = tid;  // 0..127
        const int k = p.k;
        const int rcap = p.rcap, rstride = p.rcap + 1;       // odd row stride: lane-parallel appends spread over the banks
        float* keys_w = heap_s + (size_t)(warp * 32) * rstride;  // this warp's 32 key reservoirs
        float* R = keys_w + (size_t)lane * rstride;              // this lane's (= this query's) reservoir
        uint64_t* log = p.cand + (size_t)blockIdx.x * kCandCap * kEpiThreads + ql;
        const bool l2 = p.metric == B2_METRIC_L2;
        uint32_t tc = 0;
        long long w_full = 0;
        unsigned n_ins = 0, n_log = 0, n_cmp = 0, n_sel = 0;
        for (int item = unit; item < p.n_items; item += n_units) {
            if (p.only_flagged && __ldcg(p.flags + item % p.n_qt) == 0u) continue;  // fix-up launch: flagged groups only
            const int qt = (item % p.n_qt) * CG + (int)cta_rank, seg = item / p.n_qt;
            const int t0 = seg * p.tiles_per_seg;
            const int t1 = min(p.n_tiles, t0 + p.tiles_per_seg);
            const int q = qt * kTileM + ql;
            const bool active = q < p.nq;
            // bound imported from finished items of this query (exclusive); +inf parks padding lanes
            float thr0 = active ? ord2f(__ldcg(p.tau + q)) : -kNegInf;
            float thr = thr0;      // a key enters the sorted list if it is above this
            // refine: approximate keys; everything within `band` below the running k-th is logged as well, and the
            // imported bound is already such a log threshold
            const bool refine = p.refine != 0;
            const float band = (refine && active) ? __ldg(p.band + q) : 0.f;
            float thr_log = thr0;  // a key is logged if it is above this (== thr unless refine)
            bool overflow = false;
            int cnt = 0;
            float kth_own = kNegInf;  // k-th best key of this lane's query at its reservoir's last selection (-inf before the first)
            int rc = 0;               // entries in this lane's reservoir
            __syncwarp();
            for (int u = t0; u < t1; ++u, ++tc) {
                const int t = tile_of(p, u);
                const uint32_t buf = tc % NBUF, par = (tc / NBUF) & 1u;
                const uint32_t row0 = (uint32_t)t * kTileN;
                // Every 32 tiles the segments of a query that run at the same time (on other SMs) trade bounds: publish
                // this list's k-th, pick up the best one published so far.  Without it every segment re-learns its
                // threshold from -inf (measured at k = 100: 300 insertions per (query, segment) where ~100 suffice).
                if (((u - t0) & 31) == 31 && active) {
                    const float kth_now = kth_own;
                    if (kth_now > kNegInf) atomicMax(p.tau + q, f2ord(refine ? kth_now - band : kth_now) - 1u);
                    thr0 = fmaxf(thr0, ord2f(__ldcg(p.tau + q)));
                    thr = fmaxf(thr, thr0);
                    thr_log = fmaxf(thr_log, thr0);
                }
                if (l2) {
                    float* ns = nrm_s + (tc & 1u) * kTileN;  // always double-buffered: a faster warp may already be one tile ahead
                    for (int c = ql; c < kTileN; c += kEpiThreads) {
                        const uint32_t r = row0 + c;
                        ns[c] = (r < p.n) ? p.norms[r] : 0.f;
                    }
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                }
                // a whole tile of survivors must fit behind cnt: compact first if any lane is close (warp-uniform)
                if (__any_sync(kFull, cnt > kCandCap - kTileN)) {
                    ++n_cmp;
                    if (refine) {
                        compact_band(log, cnt, thr_log, active);
                        if (cnt > kCandCap - kTileN && !overflow) {  // too many in-band candidates: leave it to the fix-up pass
                            overflow = true;
                            atomicOr(p.flags + item % p.n_qt, 1u);
                        }
                    } else {
                        compact_log(log, cnt, kth_own, k, cnt > k);
                    }
                }
                const long long c2 = p.stats ? clock64() : 0;
                mbar_wait(smem_u32(tfull_bar + buf), par);
                if (p.stats) w_full += clock64() - c2;
                tcgen05_fence_after();
                const int limit = (int)min((uint32_t)kTileN, p.n - row0);
                const uint32_t taddr0 = tmem_base + buf * kTileN + ((uint32_t)(warp * 32) << 16);
                // one 32-column chunk of the tile: raw32 = this thread's 32 accumulator values (already loaded)
                auto process_chunk = [&](int ch, uint32_t (&raw32)[32]) {
                    float key[32];
                    if (l2) {
                        const float4* n4 = reinterpret_cast<const float4*>(nrm_s + (tc & 1u) * kTileN + ch * 32);
#pragma unroll
                        for (int c4 = 0; c4 < 8; ++c4) {
                            const float4 nn = n4[c4];
                            key[4 * c4 + 0] = fmaf(2.f, __uint_as_float(raw32[4 * c4 + 0]), -nn.x);
                            key[4 * c4 + 1] = fmaf(2.f, __uint_as_float(raw32[4 * c4 + 1]), -nn.y);
                            key[4 * c4 + 2] = fmaf(2.f, __uint_as_float(raw32[4 * c4 + 2]), -nn.z);
                            key[4 * c4 + 3] = fmaf(2.f, __uint_as_float(raw32[4 * c4 + 3]), -nn.w);
                        }
                    } else {
#pragma unroll
                        for (int c = 0; c < 32; ++c) key[c] = __uint_as_float(raw32[c]) + 0.0f;
                    }
                    // Fast reject: most chunks hold no candidate for any of the warp's 32 queries; a max tree (log depth,
                    // independent operations) answers that without building any mask.  m8[c] = max over the columns
                    // c, c + 8, c + 16, c + 24.
                    float m8[8];
#pragma unroll
                    for (int c = 0; c < 8; ++c)
                        m8[c] = fmaxf(fmaxf(key[c], key[c + 8]), fmaxf(key[c + 16], key[c + 24]));
                    {
                        const float mx = fmaxf(fmaxf(fmaxf(m8[0], m8[1]), fmaxf(m8[2], m8[3])),
                                               fmaxf(fmaxf(m8[4], m8[5]), fmaxf(m8[6], m8[7])));
                        if (!__any_sync(kFull, mx > thr_log)) return;  // thr_log <= thr always
                    }
                    // the mask of columns above the log threshold, only for the column classes whose maximum passes
                    // (a chunk that gets here typically holds one candidate for one or two of the 32 queries)
                    uint32_t ml = 0;
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        if (m8[c] > thr_log) {
                            ml |= ((key[c] > thr_log) ? (1u << c) : 0u) | ((key[c + 8] > thr_log) ? (1u << (c + 8)) : 0u) |
                                  ((key[c + 16] > thr_log) ? (1u << (c + 16)) : 0u) |
                                  ((key[c + 24] > thr_log) ? (1u << (c + 24)) : 0u);
                        }
                    }
                    const int left = limit - ch * 32;
                    const uint32_t valid = (left >= 32) ? 0xffffffffu : ((left <= 0) ? 0u : ((1u << left) - 1u));
                    ml &= valid;
                    uint32_t m = 0;  // columns above the admission threshold that still have to enter the reservoir
                    if (refine) {
                        while (ml) {  // lane-local: log every in-band candidate of this chunk, in row order
                            const int c = __ffs(ml) - 1;
                            ml &= ml - 1u;
                            const float kv = select32(key, c);
                            if (cnt < kCandCap) {
                                log[(size_t)cnt * kEpiThreads] = pack_cand(kv, row0 + ch * 32 + c);
                                ++cnt;
                                ++n_log;
                            }
                            if (kv > thr) {
                                if (rc < rcap) {
                                    R[rc++] = kv;
                                    ++n_ins;
                                } else {
                                    m |= 1u << c;  // reservoir full: after the selection below
                                }
                            }
                        }
                    } else {
                        m = ml;  // thr_log == thr
                    }
                    // Reservoirs that are full get their k best selected (warp-cooperative, one query at a time; the
                    // query's threshold rises to its k-th best), then the lanes append what they still hold.
                    unsigned busy = __ballot_sync(kFull, m != 0u || rc >= rcap);
                    while (busy) {
                        unsigned fullm = __ballot_sync(kFull, rc >= rcap);
                        while (fullm) {
                            const int src = __ffs(fullm) - 1;
                            fullm &= fullm - 1u;
                            const float kth = coop_compact(keys_w + (size_t)src * rstride, rcap, k, lane);
                            if (lane == src) {
                                rc = k;
                                kth_own = kth;
                                thr = fmaxf(thr0, kth);
                                thr_log = fmaxf(thr0, kth - band);
                                ++n_sel;
                            }
                        }
                        while (m != 0u && rc < rcap) {
                            const int c = __ffs(m) - 1;
                            m &= m - 1u;
                            const float kv = select32(key, c);
                            if (kv > thr) {  // the threshold may have risen since the mask was built
                                R[rc++] = kv;
                                ++n_ins;
                                if (!refine) {
                                    log[(size_t)cnt * kEpiThreads] = pack_cand(kv, row0 + ch * 32 + c);
                                    ++cnt;
                                }
                            }
                        }
                        busy = __ballot_sync(kFull, m != 0u);  // whoever still holds survivors filled its reservoir
                    }
                };
                if (KIND == 1) {
                    // exact fp32 arithmetic: main + correction accumulators, loaded chunk by chunk
#pragma unroll 1
                    for (int ch = 0; ch < kTileN / 32; ++ch) {
                        uint32_t raw32[32], cor32[32];
                        tmem_ld32(taddr0 + ch * 32, raw32);
                        tmem_ld32(taddr0 + kTileN + ch * 32, cor32);
                        tmem_ld_wait();
#pragma unroll
                        for (int c = 0; c < 32; ++c)
                            raw32[c] = __float_as_uint(__uint_as_float(raw32[c]) + __uint_as_float(cor32[c]));
                        process_chunk(ch, raw32);
                    }
                } else {
                    // software pipeline over the chunks: the TMEM load of chunk ch + 1 is in flight while chunk ch is
                    // examined (the load latency was most of a chunk's cost: one epilogue warp per scheduler, nothing
                    // else to issue)
                    uint32_t ra[32], rb[32];
                    tmem_ld32(taddr0, ra);
#pragma unroll 1
                    for (int ch = 0; ch < kTileN / 32; ch += 2) {
                        tmem_ld_wait_regs(ra);
                        tmem_ld32(taddr0 + (ch + 1) * 32, rb);
                        process_chunk(ch, ra);
                        tmem_ld_wait_regs(rb);
                        if (ch + 2 < kTileN / 32) tmem_ld32(taddr0 + (ch + 2) * 32, ra);
                        process_chunk(ch + 1, rb);
                    }
                }
                tcgen05_fence_before();
                __syncwarp();
                if (lane == 0) {
                    if (CG == 2) mbar_arrive_cluster(map_to_cta(smem_u32(tempty_bar + buf), 0));  // the leader's MMA waits
                    else mbar_arrive(smem_u32(tempty_bar + buf));
                }
            }
            // ---- item done: the exact k-th best of the segment (one last selection where the reservoir holds more than
            // k keys, or exactly k that were never selected), emit this query's k best, publish the bound
            {
                unsigned todo = __ballot_sync(kFull, active && rc >= k && (rc > k || kth_own == kNegInf));
                while (todo) {
                    const int src = __ffs(todo) - 1;
                    todo &= todo - 1u;
                    const int n_src = __shfl_sync(kFull, rc, src);
                    const float kth = coop_compact(keys_w + (size_t)src * rstride, n_src, k, lane);
                    if (lane == src) {
                        rc = k;
                        kth_own = kth;
                        thr_log = fmaxf(thr0, kth - band);
                        ++n_sel;
                    }
                }
            }
            const float tau = kth_own;
            const bool full = active && tau > kNegInf;
            const int L = p.list_len;
            if (refine) {
                compact_band(log, cnt, thr_log, active);  // everything that can still be in the exact top-k
                if (active && (cnt > L || overflow)) atomicOr(p.flags + item % p.n_qt, 1u);
                if (full) atomicMax(p.tau + q, f2ord(tau - band) - 1u);  // a log threshold for the other segments
            } else {
                compact_log(log, cnt, tau, k, full);
                if (full) atomicMax(p.tau + q, f2ord(tau) - 1u);
            }
            if (active) {
                uint64_t* out = p.partial + ((size_t)q * p.n_lists + p.list_base + seg) * L;
                for (int i0 = 0; i0 < L; i0 += 8) {
                    uint64_t e[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        e[u] = (i0 + u < cnt) ? log[(size_t)(i0 + u) * kEpiThreads] : B2_SENTINEL;
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        if (i0 + u < L) out[i0 + u] = e[u];
                }
            }
            __syncwarp();
        }
        if (p.stats) {
            if (warp == 0 && lane == 0) atomicAdd(p.stats + 2, (unsigned long long)w_full);
            atomicAdd(p.stats + 3, (unsigned long long)n_ins);
            atomicAdd(p.stats + 4, (unsigned long long)n_log);
            if (lane == 0) atomicAdd(p.stats + 5, (unsigned long long)n_cmp);
            atomicAdd(p.stats + 8, (unsigned long long)n_sel);
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync_all();  // neither CTA may leave (or free TMEM) while its peer can still signal it
    if (warp == 5) {
        tcgen05_fence_after();
        if (CG == 2)
            asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
        else
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

// ------------------------------------------------------------------ query preparation
// q [nq][d] fp32 -> A [NA][nq_pad][kp] (bf16 or fp32-holding-tf32), zero padded; qn[q] = |q|^2.
template <int KIND>
__global__ void prep_queries_kernel(const float* __restrict__ q, int nq, int d, int nq_pad, int kp, int na, void* A,
                                    float* __restrict__ qn, uint32_t* __restrict__ tau, int mixed_slice) {
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= nq_pad) return;
    float ss = 0.f;
    for (int c = lane; c < kp; c += 32) {
        const float x = (row < nq && c < d) ? q[(size_t)row * d + c] : 0.f;
        ss = fmaf(x, x, ss);
        if (KIND == 0) {
            __nv_bfloat16* a = static_cast<__nv_bfloat16*>(A);
            const __nv_bfloat16 h = __float2bfloat16_rn(x);
            a[(size_t)row * kp + c] = h;
            if (na == 2) a[((size_t)nq_pad + row) * kp + c] = __float2bfloat16_rn(x - __bfloat162float(h));
        } else {
            float* a = static_cast<float*>(A);
            uint32_t u;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
            const float h = __uint_as_float(u);
            a[(size_t)row * kp + c] = h;
            if (na == 2 && mixed_slice > 0) {
                // part 1 = the bf16 correction operand: per K slice [bf16(q) | bf16(q - tf32(q))], same bytes per row
                __nv_bfloat16* a1 = reinterpret_cast<__nv_bfloat16*>(a + ((size_t)nq_pad + row) * kp);
                const int sl = c / mixed_slice, j = c - sl * mixed_slice;
                a1[sl * 2 * mixed_slice + j] = __float2bfloat16_rn(x);
                a1[sl * 2 * mixed_slice + mixed_slice + j] = __float2bfloat16_rn(x - h);
            } else if (na == 2) {
                a[((size_t)nq_pad + row) * kp + c] = x - h;
            }
        }
    }
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) ss += __shfl_xor_sync(kFull, ss, s);
    if (lane == 0) {
        if (qn != nullptr) qn[row] = ss;
        if (tau != nullptr) tau[row] = 0x00800000u;  // image of -FLT_MAX as the exclusive bound: admits every key > -FLT_MAX
    }
}

// Refine mode: band[q] = 2 x (a rigorous bound on |approximate key - exact key| over all stored rows), per query.
//   bf16 rows:  approximate = bf16(q) . x         -> |error| <= |q - bf16(q)| |x|
//   fp32 rows:  approximate = tf32(q) . trunc(x)   -> |error| <= (|q - tf32(q)| + 2^-10 |tf32(q)|) |x|
// plus the tensor core's truncating accumulation (<= steps x 2^-21 |q||x|, measured ~2^-24 per step), with |x| <=
// sqrt(max_norm) from the ingest side array; x2 again for L2 keys (2 q.x - |x|^2).
template <int KIND>
__global__ void band_kernel(const float* __restrict__ q, int nq, int d, int nq_pad, int steps, int l2,
                            const uint32_t* __restrict__ max_norm_ord, const uint32_t* __restrict__ row_err_ord,
                            float* __restrict__ band) {
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= nq_pad) return;
    float ss = 0.f, se = 0.f;
    if (row < nq) {
        for (int c = lane; c < d; c += 32) {
            const float x = q[(size_t)row * d + c];
            float h;
            if (KIND == 0) {
                h = __bfloat162float(__float2bfloat16_rn(x));
            } else {
                uint32_t u;
                asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
                h = __uint_as_float(u);
            }
            ss = fmaf(h, h, ss);
            se = fmaf(x - h, x - h, se);
        }
    }
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
        ss += __shfl_xor_sync(kFull, ss, s);
        se += __shfl_xor_sync(kFull, se, s);
    }
    if (lane == 0)
        band[row] = band_from_sums(ss, se, KIND != 0, steps, l2 != 0, d, ord2f(*max_norm_ord),
                                   row_err_ord != nullptr ? ord2f(*row_err_ord) : 0.f);
}

// After the threshold pre-pass: tau[q] = an exclusive lower bound of the final k-th key of query q = the k-th largest
// key among ALL candidates the pre-pass emitted for q (the union of its parts' lists, each of which holds its part's k
// best), less the band in refine mode.  The k-th best of any subset of the rows bounds the k-th best of all rows
// from below, so every work item of the main pass starts with a useful threshold instead of -inf: without it each
// of the items that start cold admits k (1 + ln(rows / k)) keys and runs the slow path of the epilogue for most chunks.
// One warp per query; the keys are staged in shared memory, bisection on the key bits (two per step).
constexpr int kUnionWarps = 4;
__global__ void __launch_bounds__(kUnionWarps * 32)
tau_union_kernel(const uint64_t* __restrict__ partial, int nq, int n_lists, int list_base, int n_parts, int list_len, int k,
                 const float* __restrict__ band, uint32_t* __restrict__ tau) {
    extern __shared__ uint32_t ukeys[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * kUnionWarps + warp;
    if (q >= nq) return;
    const int n = n_parts * list_len;
    uint32_t* K = ukeys + (size_t)warp * n;
    const uint64_t* src = partial + ((size_t)q * n_lists + list_base) * list_len;  // the parts' lists are contiguous
    int valid = 0;
    uint32_t mx = 0u, mn = 0xffffffffu;
    for (int i = lane; i < n; i += 32) {
        const uint64_t c = src[i];
        const uint32_t o = (c == B2_SENTINEL) ? 0u : (uint32_t)(c >> 32);  // 0 = no entry (images of real keys are larger)
        K[i] = o;
        if (o != 0u) {
            ++valid;
            mx = max(mx, o);
            mn = min(mn, o);
        }
    }
    valid = __reduce_add_sync(kFull, valid);
    if (valid < k) return;  // fewer than k candidates: no bound (tau keeps its initial value)
    mx = __reduce_max_sync(kFull, mx);
    mn = __reduce_min_sync(kFull, mn);
    __syncwarp();
    uint32_t prefix = mx;
    const uint32_t diff = mx ^ mn;
    if (diff != 0u) {
        int bit = 31 - __clz(diff);
        prefix = (bit == 31) ? 0u : (mx & ~((2u << bit) - 1u));
        for (; bit >= 0; bit -= 2) {
            const int sh = max(bit - 1, 0);
            const uint32_t step = (bit >= 1) ? 1u : 0u;  // a last single bit: c1 = c2 = c3 candidates collapse to one
            const uint32_t c1 = prefix | (1u << sh), c2 = step ? (prefix | (2u << sh)) : 0xffffffffu,
                           c3 = step ? (prefix | (3u << sh)) : 0xffffffffu;
            uint32_t n1 = 0, n2 = 0, n3 = 0;
            for (int i = lane; i < n; i += 32) {
                const uint32_t o = K[i];
                n1 += o >= c1;
                n2 += o >= c2;
                n3 += o >= c3;
            }
            n1 = __reduce_add_sync(kFull, n1);
            n2 = __reduce_add_sync(kFull, n2);
            n3 = __reduce_add_sync(kFull, n3);
            if (step && (int)n3 >= k) prefix = c3;
            else if (step && (int)n2 >= k) prefix = c2;
            else if ((int)n1 >= k) prefix = c1;
        }
    }
    if (lane == 0) {
        const float kth = ord2f(prefix);
        const float b = band != nullptr ? band[q] : 0.f;
        atomicMax(tau + q, f2ord(kth - b) - 1u);
    }
}

}  // namespace

int compute_band(const float* q_dev, int nq, int d, int nq_pad, int steps, bool bf16_operands, bool l2,
                 const uint32_t* max_norm_ord, const uint32_t* row_err_ord, float* band, cudaStream_t stream) {
    const int blocks = (nq_pad + 7) / 8;
    if (bf16_operands)
        band_kernel<0><<<blocks, 256, 0, stream>>>(q_dev, nq, d, nq_pad, steps, l2 ? 1 : 0, max_norm_ord, row_err_ord, band);
    else
        band_kernel<1><<<blocks, 256, 0, stream>>>(q_dev, nq, d, nq_pad, steps, l2 ? 1 : 0, max_norm_ord, row_err_ord, band);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

namespace {

// smallest reservoir for k results: room for at least 8 admissions between two selections
int min_rcap(int k) { return (k + 8 + 7) / 8 * 8; }

// `slots` = floats per query in the key reservoirs: rcap + 1
size_t gemm_smem_slots(bool fp32_rows, int na, int swb, int stages, int slots, int cg, int corr_slots) {
    const size_t stage = (size_t)na * kTileM * swb + (size_t)(kTileN / cg) * swb;
    const size_t corr = fp32_rows ? (size_t)corr_slots * (kTileN / cg) * swb : 0;
    return 1024 + stages * stage + corr + (size_t)slots * kEpiThreads * 4 + 2 * kTileN * 4 +
           (4 * kMaxStages + kMaxCorrSlots + 4) * 8 + 16;
}
// with the smallest reservoir of that k: what the pipeline depth is chosen against
size_t gemm_smem_bytes(bool fp32_rows, int na, int swb, int stages, int k, int cg = 1, int corr_slots = 2) {
    return gemm_smem_slots(fp32_rows, na, swb, stages, min_rcap(k) + 1, cg, corr_slots);
}
// The largest reservoir (multiple of 8, at most 2k and kMaxRcap) that fits beside an already chosen pipeline: a longer
// reservoir means fewer selections and a threshold that lags less in relative terms.
int pick_rcap(bool fp32_rows, int na, int swb, int stages, int k, int cg, int corr_slots) {
    int rcap = min_rcap(k);
    const int cap = std::min(kMaxRcap, std::max(32, 2 * k));
    while (rcap + 8 <= cap && gemm_smem_slots(fp32_rows, na, swb, stages, rcap + 8 + 1, cg, corr_slots) <= 227 * 1024) rcap += 8;
    return rcap;
}

template <int KIND, int NA, int SWB, int CG>
int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int grid, size_t smem,
                cudaStream_t stream) {
    auto kern = gemm_topk_kernel<KIND, NA, SWB, CG>;
    static std::mutex mu;
    static size_t attr[64] = {0};
    int dev = 0;
    B2_CUDA_TRY(cudaGetDevice(&dev));
    {
        std::lock_guard<std::mutex> lk(mu);
        if (smem > attr[dev & 63]) {
            B2_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr[dev & 63] = smem;
        }
    }
    void* tok = p.only_flagged ? nullptr : prof_begin(stream);  // the fix-up launch is not the profiled kernel
    if (CG == 2) {
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof(cfg));
        cfg.gridDim = dim3((unsigned)grid);
        cfg.blockDim = dim3(KIND == 1 ? kThreadsFp32 : kThreadsBf16);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tmA, tmB, p);
        prof_end(tok, 1, stream);
        B2_COUNT_LAUNCH();
        if (e != cudaSuccess) {
            cudaGetLastError();
            set_error("cluster launch of the tensor-core kernel failed: %s", cudaGetErrorString(e));
            return B2_ERR_UNSUPPORTED;  // gemm_search retries with one CTA per MMA
        }
        return B2_OK;
    }
    kern<<<grid, KIND == 1 ? kThreadsFp32 : kThreadsBf16, smem, stream>>>(tmA, tmB, p);
    prof_end(tok, 1, stream);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

}  // namespace

int prep_queries(const float* q_dev, int nq, int d, int nq_pad, int kp, int na, bool bf16, void* A, float* qn,
                 uint32_t* tau, cudaStream_t stream, int mixed_slice) {
    const int wpb = 8;
    const int blocks = (nq_pad + wpb - 1) / wpb;
    if (bf16)
        prep_queries_kernel<0><<<blocks, wpb * 32, 0, stream>>>(q_dev, nq, d, nq_pad, kp, na, A, qn, tau, 0);
    else
        prep_queries_kernel<1><<<blocks, wpb * 32, 0, stream>>>(q_dev, nq, d, nq_pad, kp, na, A, qn, tau, mixed_slice);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

bool gemm_path_supported(int64_t n, int d, int dtype, int k, int64_t nq) {
    if (getenv("B2_NO_TENSOR") != nullptr) return false;
    const int elem = dtype == B2_DTYPE_BF16 ? 2 : 4;
    return nq >= 20 && k <= kGemmMaxK && n >= 4096 && n < ((int64_t)1 << 31) && (int64_t)d * elem >= 64;
}

static int gemm_search_cg(const void* rows, int64_t n, size_t pitch_bytes, int d, int dtype, int metric,
                          const float* norms, const float* q_dev, int64_t nq, int k, int num_sms, float* D, int64_t* I,
                          cudaStream_t stream, const PeerPush* push, int cg, int query_parts);
static int gemm_search_refine(const void* rows, int64_t n, size_t pitch_bytes, int d, int dtype, int metric,
                              const float* norms, const uint32_t* max_norm_ord, const float* q_dev, int64_t nq, int k,
                              int num_sms, float* D, int64_t* I, cudaStream_t stream, const PeerPush* push, int cg,
                              const ApproxRows* approx, const ResidualRows* resid);

bool gemm_uses_refine(int query_parts, const uint32_t* max_norm_ord) {
    bool refine = max_norm_ord != nullptr && query_parts == 0 && getenv("B2_GEMM_QUERY_PARTS") == nullptr;
    if (const char* e = getenv("B2_GEMM_REFINE")) refine = refine && atoi(e) != 0;
    return refine;
}

int gemm_search(const void* rows, int64_t n, size_t pitch_bytes, int d, int dtype, int metric, const float* norms,
                const float* q_dev, int64_t nq, int k, int num_sms, float* D, int64_t* I, cudaStream_t stream,
                const PeerPush* push, int query_parts, const uint32_t* max_norm_ord, const ApproxRows* approx,
                const ResidualRows* resid) {
    // Filter and refine (default): one approximate product on the tensor cores, exact re-scoring of the few
    // candidates near the cut.  An explicit number of query parts (option / environment) selects the all-exact
    // tensor-core arithmetic instead, as does B2_GEMM_REFINE=0.
    const bool refine = gemm_uses_refine(query_parts, max_norm_ord);
    if (dtype == B2_DTYPE_BF16) approx = nullptr;
    if (dtype != B2_DTYPE_BF16) resid = nullptr;
    if (resid != nullptr && !refine) {
        set_error("tensor-core search: rows with a residual plane need the filter-and-refine mode");
        return B2_ERR_UNSUPPORTED;
    }
    if (refine) {
        int cgr = (num_sms % 2 == 0) ? 2 : 1;
        if (const char* e = getenv("B2_GEMM_CTA_GROUP")) cgr = (atoi(e) == 2 && num_sms % 2 == 0) ? 2 : 1;
        int rc = gemm_search_refine(rows, n, pitch_bytes, d, dtype, metric, norms, max_norm_ord, q_dev, nq, k, num_sms, D,
                                    I, stream, push, cgr, approx, resid);
        if (rc == B2_ERR_UNSUPPORTED && cgr == 2)
            rc = gemm_search_refine(rows, n, pitch_bytes, d, dtype, metric, norms, max_norm_ord, q_dev, nq, k, num_sms,
                                    D, I, stream, push, 1, approx, resid);
        return rc;
    }
    // CTA pairs (cta_group::2) for bf16 rows when the SM count is even.  Measured at 10M x 1024, nq = 4096, k = 100:
    // bf16 rows 132.7 ms with pairs vs 141.0 ms without; fp32 rows 536 ms with pairs vs 388 ms without (every K slice
    // then needs 16 cluster-scope arrivals from the two CTAs' transform warps before the leader may issue), so fp32
    // rows keep one CTA per MMA.  B2_GEMM_CTA_GROUP overrides.
    int cg = (dtype == B2_DTYPE_BF16 && num_sms % 2 == 0) ? 2 : 1;
    if (const char* e = getenv("B2_GEMM_CTA_GROUP")) cg = (atoi(e) == 2 && num_sms % 2 == 0) ? 2 : 1;
    int rc = gemm_search_cg(rows, n, pitch_bytes, d, dtype, metric, norms, q_dev, nq, k, num_sms, D, I, stream, push, cg,
                           query_parts);
    if (rc == B2_ERR_UNSUPPORTED && cg == 2)  // no cluster launch here (e.g. a partitioned GPU): one CTA per MMA
        rc = gemm_search_cg(rows, n, pitch_bytes, d, dtype, metric, norms, q_dev, nq, k, num_sms, D, I, stream, push, 1,
                            query_parts);
    return rc;
}

static int gemm_search_cg(const void* rows, int64_t n, size_t pitch_bytes, int d, int dtype, int metric,
                          const float* norms, const float* q_dev, int64_t nq, int k, int num_sms, float* D, int64_t* I,
                          cudaStream_t stream, const PeerPush* push, int cg, int query_parts) {
    const bool bf16 = dtype == B2_DTYPE_BF16;
    const int elem = bf16 ? 2 : 4;
    int na = 2;
    if (const char* e = getenv("B2_GEMM_QUERY_PARTS")) na = (atoi(e) == 1) ? 1 : 2;
    if (query_parts == 1 || query_parts == 2) na = query_parts;  // b2_index_set_option("gemm_query_parts")
    // stage width: 128-byte swizzle when at least three stages of it fit beside the k-slot heaps, else 64
    int swb = (gemm_smem_bytes(!bf16, na, 128, 3, k, cg) <= 227 * 1024) ? 128 : 64;
    if (const char* e = getenv("B2_GEMM_SWIZZLE")) swb = (atoi(e) == 128) ? 128 : 64;  // tuning override
    if (swb == 128 && gemm_smem_bytes(!bf16, na, 128, 2, k, cg) > 227 * 1024) swb = 64;
    const size_t kbytes = (((size_t)d * elem + swb - 1) / swb) * swb;
    const int kp = (int)(kbytes / elem);

    GemmParams p;
    p.nq = (int)nq;
    p.nq_pad = (int)((nq + kTileM * cg - 1) / (kTileM * cg)) * kTileM * cg;
    p.n = (uint32_t)n;
    p.k = k;
    p.kstages = (int)(kbytes / swb);
    p.metric = metric;
    p.n_qt = p.nq_pad / (kTileM * cg);  // query-tile groups: one per CTA, or per CTA pair
    p.n_tiles = (int)((n + kTileN - 1) / kTileN);
    p.norms = norms;

    // segments: make (query tiles x segments) a near multiple of the CTA count, segments >= 8 tiles
    const int grid_max = num_sms / cg;  // units that run at once: CTAs, or CTA pairs
    int best_s = 1;
    double best_cost = 1e30;
    const int max_s = std::max(1, std::min(p.n_tiles / 8, 4 * grid_max));
    for (int s = 1; s <= max_s; ++s) {
        const int tps = (p.n_tiles + s - 1) / s;
        const int s_eff = (p.n_tiles + tps - 1) / tps;
        const long long items = (long long)p.n_qt * s_eff;
        const long long waves = (items + grid_max - 1) / grid_max;
        // time ~ waves x tiles per segment (+ a per-item warm-up cost of about 6 tiles)
        const double cost = (double)waves * (tps + 6.0);
        if (cost < best_cost * 0.995) {
            best_cost = cost;
            best_s = s_eff;
        }
    }
    p.tiles_per_seg = (p.n_tiles + best_s - 1) / best_s;
    p.n_seg = (p.n_tiles + p.tiles_per_seg - 1) / p.tiles_per_seg;
    p.n_items = p.n_qt * p.n_seg;
    p.tile_mul = 1;
    p.tile_skip = 0;
    p.n_lists = p.n_seg;
    p.list_base = 0;
    const int grid = (int)std::min<long long>(grid_max, p.n_items) * cg;

    // pipeline depth from the shared-memory budget
    int stages = kMaxStages;
    while (stages > 2 && gemm_smem_bytes(!bf16, na, swb, stages, k, cg) > 227 * 1024) --stages;
    // fp32 rows: (TMA stages, correction tiles) from what was measured at 10M x 1024, nq = 4096: (4, 4) 352 ms,
    // (4, 2) 366 ms, (3, 3) 398 ms, (5, 3) 420 ms - a fifth stage does not pay, a ring shorter than the stages does
    int corr_slots = 2;
    if (!bf16) {
        static const int cand[5][2] = {{4, 4}, {4, 2}, {3, 3}, {3, 2}, {2, 2}};
        for (int c = 0; c < 5; ++c) {
            if (gemm_smem_bytes(true, na, swb, cand[c][0], k, cg, cand[c][1]) <= 227 * 1024) {
                stages = cand[c][0];
                corr_slots = cand[c][1];
                break;
            }
        }
        if (const char* e = getenv("B2_GEMM_CORR_SLOTS")) {
            corr_slots = std::max(2, std::min(kMaxCorrSlots, atoi(e)));
            stages = kMaxStages;
            while (stages > 2 && gemm_smem_bytes(true, na, swb, stages, k, cg, corr_slots) > 227 * 1024) --stages;
        }
    }
    p.rcap = pick_rcap(!bf16, na, swb, stages, k, cg, corr_slots);
    const size_t smem = gemm_smem_slots(!bf16, na, swb, stages, p.rcap + 1, cg, corr_slots);
    p.corr_slots = corr_slots;
    if (smem > 227 * 1024) {
        set_error("tensor-core search: k=%d does not fit shared memory", k);
        return B2_ERR_UNSUPPORTED;
    }
    p.num_stages = stages;
    p.trunc_hi = 1;  // measured on B200: identical accuracy to rounding in place (recall_check), one smem write less
    if (const char* e = getenv("B2_GEMM_TRUNC")) p.trunc_hi = atoi(e) ? 1 : 0;
    p.mixed = (!bf16 && na == 2) ? 1 : 0;
    if (const char* e = getenv("B2_GEMM_MIXED")) p.mixed = (!bf16 && na == 2 && atoi(e)) ? 1 : 0;
    p.refine = 0;
    p.only_flagged = 0;
    p.list_len = k;
    p.band = nullptr;
    p.flags = nullptr;
    p.stats = nullptr;

    // scratch: [A parts][qn][tau][partial][candidate logs]
    auto al = [](size_t v) { return (v + 1023) / 1024 * 1024; };
    const size_t a_bytes = al((size_t)na * p.nq_pad * kbytes);
    const size_t qn_bytes = al((size_t)p.nq_pad * 4);
    const size_t tau_bytes = al((size_t)p.nq_pad * 4);
    const size_t part_bytes = al((size_t)p.nq_pad * p.n_seg * k * 8);
    const size_t cand_bytes = al((size_t)grid * kCandCap * kEpiThreads * 8);
    char* blk = nullptr;
    B2_CUDA_TRY(cudaMallocAsync(&blk, a_bytes + qn_bytes + tau_bytes + part_bytes + cand_bytes, stream));
    void* A = blk;
    float* qn = reinterpret_cast<float*>(blk + a_bytes);
    p.tau = reinterpret_cast<uint32_t*>(blk + a_bytes + qn_bytes);
    p.partial = reinterpret_cast<uint64_t*>(blk + a_bytes + qn_bytes + tau_bytes);
    p.cand = reinterpret_cast<uint64_t*>(blk + a_bytes + qn_bytes + tau_bytes + part_bytes);

    int rc = B2_OK;
    CUtensorMap tmA, tmB;
    rc = make_map(&tmA, bf16, swb, A, (uint64_t)kp, (uint64_t)na * p.nq_pad, kbytes, kTileM);
    if (rc == B2_OK) rc = make_map(&tmB, bf16, swb, rows, (uint64_t)(pitch_bytes / elem), (uint64_t)n, pitch_bytes, kTileN / cg);
    if (rc == B2_OK) rc = prep_queries(q_dev, (int)nq, d, p.nq_pad, kp, na, bf16, A, qn, p.tau, stream, p.mixed ? swb / 4 : 0);
    if (rc == B2_OK) {
        const int variant = (cg == 2 ? 8 : 0) + (bf16 ? 0 : 4) + (na == 2 ? 2 : 0) + (swb == 128 ? 1 : 0);
        switch (variant) {
            case 0: rc = launch_gemm<0, 1, 64, 1>(tmA, tmB, p, grid, smem, stream); break;
            case 1: rc = launch_gemm<0, 1, 128, 1>(tmA, tmB, p, grid, smem, stream); break;
            case 2: rc = launch_gemm<0, 2, 64, 1>(tmA, tmB, p, grid, smem, stream); break;
            case 3: rc = launch_gemm<0, 2, 128, 1>(tmA, tmB, p, grid, smem, stream); break;
            case 4: rc = launch_gemm<1, 1, 64, 1>(tmA, tmB, p, grid, smem, stream); break;
            case 5: rc = launch_gemm<1, 1, 128, 1>(tmA, tmB, p, grid, smem, stream); break;
            case 6: rc = launch_gemm<1, 2, 64, 1>(tmA, tmB, p, grid, smem, stream); break;
            case 7: rc = launch_gemm<1, 2, 128, 1>(tmA, tmB, p, grid, smem, stream); break;
            case 8: rc = launch_gemm<0, 1, 64, 2>(tmA, tmB, p, grid, smem, stream); break;
            case 9: rc = launch_gemm<0, 1, 128, 2>(tmA, tmB, p, grid, smem, stream); break;
            case 10: rc = launch_gemm<0, 2, 64, 2>(tmA, tmB, p, grid, smem, stream); break;
            case 11: rc = launch_gemm<0, 2, 128, 2>(tmA, tmB, p, grid, smem, stream); break;
            case 12: rc = launch_gemm<1, 1, 64, 2>(tmA, tmB, p, grid, smem, stream); break;
            case 13: rc = launch_gemm<1, 1, 128, 2>(tmA, tmB, p, grid, smem, stream); break;
            case 14: rc = launch_gemm<1, 2, 64, 2>(tmA, tmB, p, grid, smem, stream); break;
            default: rc = launch_gemm<1, 2, 128, 2>(tmA, tmB, p, grid, smem, stream); break;
        }
    }
    if (rc == B2_OK)
        rc = merge_partials(p.partial, (int)nq, p.n_seg, k, metric, 0, D, I, (size_t)k, stream,
                            metric == B2_METRIC_L2 ? qn : nullptr, push);
    cudaFreeAsync(blk, stream);
    return rc;
}

// ------------------------------------------------------------------ filter and refine
// One approximate product per score on the tensor cores (bf16 rows: bf16(q) . x; fp32 rows: tf32(q) . trunc_tf32(x)
// straight from the raw tile, no transform warps), every candidate within a rigorous error band of the running k-th
// is kept, and merge_rescore_kernel re-scores the handful of candidates near the final cut exactly (fp32 FMA, the
// small-batch scan's arithmetic) before it selects the top k.  Exactness does not depend on the data: the band
// comes from |q - q~| |x| (Cauchy-Schwarz) with |x| <= the largest stored norm.  If a query collects more in-band
// candidates than its list holds (dense score ties), its query group is flagged and a second launch of the
// all-exact kernel (2 query parts / tf32 + bf16 correction) redoes exactly those groups.
template <int KIND, int NA, int SWB>
static int launch_gemm_cg(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, int grid, size_t smem,
                          cudaStream_t stream, int cg) {
    return cg == 2 ? launch_gemm<KIND, NA, SWB, 2>(tmA, tmB, p, grid, smem, stream)
                   : launch_gemm<KIND, NA, SWB, 1>(tmA, tmB, p, grid, smem, stream);
}

static int gemm_search_refine(const void* rows, int64_t n, size_t pitch_bytes, int d, int dtype, int metric,
                              const float* norms, const uint32_t* max_norm_ord, const float* q_dev, int64_t nq, int k,
                              int num_sms, float* D, int64_t* I, cudaStream_t stream, const PeerPush* push, int cg,
                              const ApproxRows* approx, const ResidualRows* resid) {
    const bool bf16 = dtype == B2_DTYPE_BF16;
    const int elem = bf16 ? 2 : 4;
    // the approximate pass reads bf16 operands when the rows are bf16 or a bf16 shadow of the fp32 rows exists
    const bool abf16 = bf16 || approx != nullptr;
    const int aelem = abf16 ? 2 : 4;
    const void* arows = approx != nullptr ? approx->rows : rows;
    const size_t apitch = approx != nullptr ? approx->pitch : pitch_bytes;
    auto fits = [&](bool fp32_exact, int na, int swb, int stages, int corr) {
        return gemm_smem_bytes(fp32_exact, na, swb, stages, k, cg, corr) <= 227 * 1024;
    };
    // ---- the approximate launch: one query part, no correction tiles
    int swb = fits(false, 1, 128, 3, 0) ? 128 : 64;
    if (const char* e = getenv("B2_GEMM_SWIZZLE")) swb = (atoi(e) == 128 && fits(false, 1, 128, 2, 0)) ? 128 : 64;
    const size_t kbytes = (((size_t)d * aelem + swb - 1) / swb) * swb;
    const int kp = (int)(kbytes / aelem);
    int stages = kMaxStages;
    if (const char* e = getenv("B2_GEMM_STAGES")) stages = std::max(2, std::min(kMaxStages, atoi(e)));
    while (stages > 2 && !fits(false, 1, swb, stages, 0)) --stages;
    if (!fits(false, 1, swb, stages, 0)) {
        set_error("tensor-core search: k=%d does not fit shared memory", k);
        return B2_ERR_UNSUPPORTED;
    }
    // ---- the exact fix-up launch (same grid, same items): two query parts
    const int swb_x = 64;
    int stages_x = kMaxStages, corr_x = 2;
    if (bf16) {
        while (stages_x > 2 && !fits(false, 2, swb_x, stages_x, 0)) --stages_x;
    } else {
        static const int cand[5][2] = {{4, 4}, {4, 2}, {3, 3}, {3, 2}, {2, 2}};
        stages_x = 2;
        for (int c = 0; c < 5; ++c)
            if (fits(true, 2, swb_x, cand[c][0], cand[c][1])) {
                stages_x = cand[c][0];
                corr_x = cand[c][1];
                break;
            }
    }
    const size_t kbytes_x = (((size_t)d * elem + swb_x - 1) / swb_x) * swb_x;
    const int kp_x = (int)(kbytes_x / elem);
    const int rcap_x = pick_rcap(!bf16, 2, swb_x, stages_x, k, cg, corr_x);
    const size_t smem_x = gemm_smem_slots(!bf16, 2, swb_x, stages_x, rcap_x + 1, cg, corr_x);
    if (smem_x > 227 * 1024) {
        set_error("tensor-core search: k=%d does not fit shared memory", k);
        return B2_ERR_UNSUPPORTED;
    }

    GemmParams p;
    p.nq = (int)nq;
    p.nq_pad = (int)((nq + kTileM * cg - 1) / (kTileM * cg)) * kTileM * cg;
    p.n = (uint32_t)n;
    p.k = k;
    p.kstages = (int)(kbytes / swb);
    p.metric = metric;
    p.n_qt = p.nq_pad / (kTileM * cg);
    p.n_tiles = (int)((n + kTileN - 1) / kTileN);
    p.norms = norms;
    // Room for the in-band candidates of a segment beyond its k best.  Their number grows with k (about
    // k * z * band / sigma for scores with a Gaussian tail: ~0.4 k on 1024-d unit vectors with bf16 queries); a
    // segment that overflows the list is flagged and redone by the exact fix-up launch, so this is a speed knob only.
    int slack = std::max(48, 32 + 5 * k / 4);
    if (approx != nullptr || resid != nullptr) slack = std::max(64, 32 + 3 * k);  // the band also covers the rows' bf16 rounding: ~2.3x wider
    if (const char* e = std::getenv("B2_GEMM_LIST_SLACK")) slack = std::max(8, std::min(1024 - k - 8, atoi(e)));
    p.list_len = ((k + slack) + 7) / 8 * 8;

    const int grid_max = num_sms / cg;
    // segments: (query-tile groups x segments) a near multiple of the units that run at once, segments >= 8 tiles
    auto plan = [&](int tiles, int* tps_out, int* nseg_out) {
        int best_s = 1;
        double best_cost = 1e30;
        const int max_s = std::max(1, std::min(tiles / 8, 4 * grid_max));
        for (int s = 1; s <= max_s; ++s) {
            const int tps = (tiles + s - 1) / s;
            const int s_eff = (tiles + tps - 1) / tps;
            const long long items = (long long)p.n_qt * s_eff;
            const long long waves = (items + grid_max - 1) / grid_max;
            const double cost = (double)waves * (tps + 6.0);
            if (cost < best_cost * 0.995) {
                best_cost = cost;
                best_s = s_eff;
            }
        }
        *tps_out = (tiles + best_s - 1) / best_s;
        *nseg_out = (tiles + *tps_out - 1) / *tps_out;
    };
    // Threshold pre-pass (two launches of the same kernel): first every `sstride`-th corpus tile - a sample of the rows
    // spread over the whole store - split into as many parts as there are idle units per query group; tau_union_kernel
    // turns the parts' lists into one lower bound of the final k-th key per query; the main launch then covers the
    // other tiles and every one of its work items starts from that bound.  Without it the items of the first wave start
    // from -inf, admit k (1 + ln(rows / k)) keys each and keep the epilogue on its slow path for most chunks (measured at
    // 10M x 1024, 4096 queries, k = 100: 11 of the 14 Mcycles the MMA warp waited for a free accumulator fell into
    // each unit's first item).  Worth it from k = 32 and a corpus of >= 64 sample tiles per launch wave.
    const int total_tiles = p.n_tiles;
    int sstride = 32;
    if (const char* e = std::getenv("B2_GEMM_SAMPLE")) sstride = atoi(e);  // 0 = single pass; else the sample stride (>= 2)
    const bool two_pass = sstride >= 2 && k >= 32 && total_tiles >= 64 * sstride;
    const int n_sample = two_pass ? (total_tiles + sstride - 1) / sstride : 0;
    int tps_s = 0, nseg_s = 0, tps_m = 0, nseg_m = 0;
    if (two_pass) plan(n_sample, &tps_s, &nseg_s);
    plan(total_tiles - n_sample, &tps_m, &nseg_m);
    // merge_rescore's fast kernel holds all candidates of a query in registers (<= kMergeSelectMax): when the pre-pass lists do
    // not fit beside the main ones, the main launch covers the sample tiles again (3 % more products) and the pre-pass
    // lists only feed the threshold union
    const bool keep_sample_lists = two_pass && (size_t)(nseg_m + nseg_s) * p.list_len <= (size_t)kMergeSelectMax;
    if (two_pass && !keep_sample_lists) plan(total_tiles, &tps_m, &nseg_m);
    p.n_lists = keep_sample_lists ? nseg_m + nseg_s : nseg_m;  // what merge_rescore reads
    // the main launch (when there is no pre-pass: the only one)
    p.n_tiles = keep_sample_lists ? total_tiles - n_sample : total_tiles;
    p.tiles_per_seg = tps_m;
    p.n_seg = nseg_m;
    p.n_items = p.n_qt * p.n_seg;
    p.tile_mul = 1;
    p.tile_skip = keep_sample_lists ? sstride : 0;
    p.list_base = 0;
    const int grid = (int)std::min<long long>(grid_max, p.n_items) * cg;
    p.num_stages = stages;
    p.corr_slots = 2;
    p.trunc_hi = 1;
    p.mixed = 0;
    p.refine = 1;
    p.only_flagged = 0;
    // scratch: [A (1 part)][A exact (2 parts)][qn][tau][band][flags][partial][candidate logs]
    auto al = [](size_t v) { return (v + 1023) / 1024 * 1024; };
    const size_t a_bytes = al((size_t)p.nq_pad * kbytes);
    const size_t ax_bytes = al((size_t)2 * p.nq_pad * kbytes_x);
    const size_t qn_bytes = al((size_t)p.nq_pad * 4);
    const size_t fl_bytes = al((size_t)p.n_qt * 4) + 1024;  // + diagnostics counters
    const size_t part_bytes = al((size_t)p.nq_pad * p.n_lists * p.list_len * 8);
    const size_t spart_bytes = (two_pass && !keep_sample_lists) ? al((size_t)p.nq_pad * nseg_s * p.list_len * 8) : 0;
    const size_t cand_bytes = al((size_t)grid_max * cg * kCandCap * kEpiThreads * 8);  // any of the launches below
    char* blk = nullptr;
    B2_CUDA_TRY(cudaMallocAsync(&blk, a_bytes + ax_bytes + 3 * qn_bytes + fl_bytes + part_bytes + spart_bytes + cand_bytes, stream));
    char* cur = blk;
    void* A = cur;
    cur += a_bytes;
    void* Ax = cur;
    cur += ax_bytes;
    float* qn = reinterpret_cast<float*>(cur);
    cur += qn_bytes;
    p.tau = reinterpret_cast<uint32_t*>(cur);
    cur += qn_bytes;
    float* band = reinterpret_cast<float*>(cur);
    cur += qn_bytes;
    p.flags = reinterpret_cast<uint32_t*>(cur);
    static const bool want_stats = std::getenv("B2_GEMM_DEBUG") != nullptr;
    p.stats = want_stats ? reinterpret_cast<unsigned long long*>(cur + fl_bytes - 1024) : nullptr;
    cur += fl_bytes;
    p.partial = reinterpret_cast<uint64_t*>(cur);
    cur += part_bytes;
    uint64_t* spartial = reinterpret_cast<uint64_t*>(cur);  // pre-pass lists that are not merged (see keep_sample_lists)
    cur += spart_bytes;
    p.cand = reinterpret_cast<uint64_t*>(cur);
    p.band = band;

    int rc = B2_OK;
    CUtensorMap tmA, tmB, tmAx, tmBx;
    rc = make_map(&tmA, abf16, swb, A, (uint64_t)kp, (uint64_t)p.nq_pad, kbytes, kTileM);
    if (rc == B2_OK) rc = make_map(&tmB, abf16, swb, arows, (uint64_t)(apitch / aelem), (uint64_t)n, apitch, kTileN / cg);
    if (rc == B2_OK) rc = make_map(&tmAx, bf16, swb_x, Ax, (uint64_t)kp_x, (uint64_t)2 * p.nq_pad, kbytes_x, kTileM);
    if (rc == B2_OK) rc = make_map(&tmBx, bf16, swb_x, rows, (uint64_t)(pitch_bytes / elem), (uint64_t)n, pitch_bytes, kTileN / cg);
    if (rc == B2_OK && cudaMemsetAsync(p.flags, 0, fl_bytes, stream) != cudaSuccess) {
        set_error("tensor-core search: cudaMemsetAsync failed");
        rc = B2_ERR_CUDA;
    }
    if (rc == B2_OK) rc = prep_queries(q_dev, (int)nq, d, p.nq_pad, kp, 1, abf16, A, qn, p.tau, stream, 0);
    if (rc == B2_OK) rc = prep_queries(q_dev, (int)nq, d, p.nq_pad, kp_x, 2, bf16, Ax, nullptr, nullptr, stream, bf16 ? 0 : swb_x / 4);
    if (rc == B2_OK)
        rc = compute_band(q_dev, (int)nq, d, p.nq_pad, abf16 ? kp / 16 : kp / 8, abf16, metric == B2_METRIC_L2, max_norm_ord,
                          approx != nullptr ? approx->err_ord : (resid != nullptr ? resid->err_ord : nullptr), band, stream);
    p.rcap = pick_rcap(false, 1, swb, stages, k, cg, 0);
    const size_t smem = gemm_smem_slots(false, 1, swb, stages, p.rcap + 1, cg, 0);
    static const bool debug_t = std::getenv("B2_GEMM_DEBUG") != nullptr;
    cudaEvent_t dbg_ev0 = nullptr;
    if (debug_t) {
        cudaEventCreate(&dbg_ev0);
        cudaEventRecord(dbg_ev0, stream);
    }
    auto launch_approx_smem = [&](const GemmParams& gp, int g, size_t sm) {
        if (abf16)
            return swb == 128 ? launch_gemm_cg<0, 1, 128>(tmA, tmB, gp, g, sm, stream, cg)
                              : launch_gemm_cg<0, 1, 64>(tmA, tmB, gp, g, sm, stream, cg);
        return swb == 128 ? launch_gemm_cg<2, 1, 128>(tmA, tmB, gp, g, sm, stream, cg)
                          : launch_gemm_cg<2, 1, 64>(tmA, tmB, gp, g, sm, stream, cg);
    };
    auto launch_approx = [&](const GemmParams& gp, int g) { return launch_approx_smem(gp, g, smem); };
    if (rc == B2_OK && two_pass) {
        GemmParams ps = p;
        ps.n_tiles = n_sample;
        ps.tiles_per_seg = tps_s;
        ps.n_seg = nseg_s;
        ps.n_items = ps.n_qt * ps.n_seg;
        ps.tile_mul = sstride;
        ps.tile_skip = 0;
        ps.list_base = keep_sample_lists ? nseg_m : 0;
        // every item of the pre-pass starts cold and is short, so it is bound by reservoir selections, not by the
        // operand pipeline: the longest reservoir that leaves three pipeline stages (a selection per rcap - k admissions)
        int stages_s = stages;
        int rcap_s = p.rcap;
        {
            const int want = std::min(kMaxRcap, std::max(32, 2 * k)) / 8 * 8;
            for (int st = stages; st >= 3; --st) {
                int rc_try = want;
                while (rc_try > p.rcap && gemm_smem_slots(false, 1, swb, st, rc_try + 1, cg, 0) > 227 * 1024) rc_try -= 8;
                if (rc_try > rcap_s) {
                    rcap_s = rc_try;
                    stages_s = st;
                }
                if (rc_try >= want) break;
            }
        }
        ps.rcap = rcap_s;
        ps.num_stages = stages_s;
        if (!keep_sample_lists) {
            ps.partial = spartial;
            ps.n_lists = nseg_s;
        }
        ps.stats = nullptr;
        rc = launch_approx_smem(ps, (int)std::min<long long>(grid_max, ps.n_items) * cg,
                                gemm_smem_slots(false, 1, swb, stages_s, rcap_s + 1, cg, 0));
        if (rc == B2_OK) {
            const size_t us = (size_t)kUnionWarps * nseg_s * p.list_len * 4;
            static std::mutex umu;
            static size_t uattr[64] = {0};
            int dev = 0;
            cudaGetDevice(&dev);
            {
                std::lock_guard<std::mutex> lk(umu);
                if (us > 48 * 1024 && us > uattr[dev & 63]) {
                    if (cudaFuncSetAttribute(tau_union_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)us) != cudaSuccess) {
                        cudaGetLastError();
                        set_error("tensor-core search: shared memory for the threshold union");
                        rc = B2_ERR_CUDA;
                    }
                    uattr[dev & 63] = us;
                }
            }
            if (rc == B2_OK) {
                tau_union_kernel<<<(unsigned)((nq + kUnionWarps - 1) / kUnionWarps), kUnionWarps * 32, us, stream>>>(
                    ps.partial, (int)nq, ps.n_lists, ps.list_base, nseg_s, p.list_len, k, band, p.tau);
                B2_COUNT_LAUNCH();
                if (cudaGetLastError() != cudaSuccess) {
                    set_error("tensor-core search: tau_union_kernel launch failed");
                    rc = B2_ERR_CUDA;
                }
            }
        }
    }
    cudaEvent_t dbg_evp = nullptr;
    if (debug_t && two_pass) {
        cudaEventCreate(&dbg_evp);
        cudaEventRecord(dbg_evp, stream);
    }
    if (rc == B2_OK) rc = launch_approx(p, grid);
    static const bool debug = std::getenv("B2_GEMM_DEBUG") != nullptr;
    cudaEvent_t dbg_ev[4] = {nullptr, nullptr, nullptr, nullptr};
    if (debug) {
        for (auto& e : dbg_ev) cudaEventCreate(&e);
        cudaEventRecord(dbg_ev[1], stream);
    }
    if (rc == B2_OK && debug) {  // diagnostics only: how many query groups the approximate pass handed to the fix-up
        std::vector<uint32_t> fl(p.n_qt);
        std::vector<float> bd(std::min(p.nq, 4));
        cudaStreamSynchronize(stream);
        cudaMemcpy(fl.data(), p.flags, fl.size() * 4, cudaMemcpyDeviceToHost);
        cudaMemcpy(bd.data(), band, bd.size() * 4, cudaMemcpyDeviceToHost);
        int nf = 0;
        for (uint32_t f : fl) nf += f != 0u;
        std::fprintf(stderr, "[b2 gemm refine] k=%d list_len=%d n_seg=%d (+%d pre-pass parts, stride %d) n_qt=%d flagged=%d band[0]=%g\n", k, p.list_len,
                     p.n_seg, nseg_s, two_pass ? sstride : 0, p.n_qt, nf, bd.empty() ? 0.0 : (double)bd[0]);
        unsigned long long st[9] = {0};
        cudaMemcpy(st, p.stats, sizeof(st), cudaMemcpyDeviceToHost);
        const double units = (double)(grid / cg), lanes = (double)p.nq_pad * p.n_seg;
        std::fprintf(stderr,
                     "[b2 gemm refine] per MMA warp: %.1f Mcycles, waiting for a free accumulator %.1f, for operands %.1f; "
                     "epilogue warp 0 waiting for a full accumulator %.1f (per CTA); per (query, segment): %.1f insertions, "
                     "%.1f log entries, %.2f reservoir selections (rcap %d); %.2f compactions per warp-item; accumulator wait inside the first item of each unit %.1f Mcycles\n",
                     st[6] / units / 1e6, st[0] / units / 1e6, st[1] / units / 1e6, st[2] / (double)grid / 1e6,
                     st[3] / lanes, st[4] / lanes, st[8] / lanes, p.rcap, st[5] / ((double)p.n_items * cg * 4), st[7] / units / 1e6);
    }
    if (rc == B2_OK && resid != nullptr) {
        // Rows with a residual plane have no all-exact tensor arithmetic (it would score against the main plane alone):
        // if a list overflowed - dense score ties, normally never - the caller redoes the batch with the two-plane
        // FFMA scan.  Costs one synchronisation per search in this mode.
        std::vector<uint32_t> fl((size_t)p.n_qt);
        cudaError_t e = cudaMemcpyAsync(fl.data(), p.flags, fl.size() * 4, cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) {
            set_error("tensor-core search: reading the overflow flags failed: %s", cudaGetErrorString(e));
            rc = B2_ERR_CUDA;
        } else {
            for (uint32_t f : fl)
                if (f != 0u) rc = B2_INTERNAL_REDO_EXACT;
        }
    }
    if (rc == B2_OK && resid == nullptr) {  // fix-up: the all-exact kernel over the flagged query groups only (normally none)
        GemmParams px = p;
        px.refine = 0;
        px.only_flagged = 1;
        px.stats = nullptr;
        // every corpus tile, spread over all n_lists list slots (slots beyond the last tile get empty lists), so that
        // the lists of a redone query group are replaced completely
        px.n_tiles = total_tiles;
        px.tile_mul = 1;
        px.tile_skip = 0;
        px.list_base = 0;
        px.n_seg = p.n_lists;
        px.tiles_per_seg = (total_tiles + p.n_lists - 1) / p.n_lists;
        px.n_items = px.n_qt * px.n_seg;
        px.kstages = (int)(kbytes_x / swb_x);
        px.num_stages = stages_x;
        px.corr_slots = corr_x;
        px.rcap = rcap_x;
        px.mixed = bf16 ? 0 : 1;
        const int grid_x = (int)std::min<long long>(grid_max, px.n_items) * cg;
        rc = bf16 ? launch_gemm_cg<0, 2, 64>(tmAx, tmBx, px, grid_x, smem_x, stream, cg)
                  : launch_gemm_cg<1, 2, 64>(tmAx, tmBx, px, grid_x, smem_x, stream, cg);
    }
    if (debug) cudaEventRecord(dbg_ev[2], stream);
    if (rc == B2_OK)
        rc = merge_rescore(p.partial, (int)nq, p.n_lists, p.list_len, k, metric, D, I, (size_t)k, stream,
                           metric == B2_METRIC_L2 ? qn : nullptr, q_dev, d, rows, pitch_bytes, dtype, norms, band, push, nullptr,
                           false, resid != nullptr ? resid->rows : nullptr);
    if (debug) {
        cudaEventRecord(dbg_ev[3], stream);
        cudaStreamSynchronize(stream);
        float t_approx = 0.f, t_fix = 0.f, t_merge = 0.f;
        cudaEventElapsedTime(&t_approx, dbg_ev0, dbg_ev[1]);
        cudaEventElapsedTime(&t_fix, dbg_ev[1], dbg_ev[2]);
        cudaEventElapsedTime(&t_merge, dbg_ev[2], dbg_ev[3]);
        float t_pre = 0.f;
        if (dbg_evp != nullptr) {
            cudaEventElapsedTime(&t_pre, dbg_ev0, dbg_evp);
            cudaEventDestroy(dbg_evp);
        }
        std::fprintf(stderr, "[b2 gemm refine] approx %.2f ms (of which threshold pre-pass + union %.2f ms; stages %d, swizzle %d), fix-up %.2f ms (incl. the flag read-back), merge %.2f ms\n",
                     t_approx, t_pre, stages, swb, t_fix, t_merge);
        for (auto& e : dbg_ev) cudaEventDestroy(e);
        cudaEventDestroy(dbg_ev0);
    }
    cudaFreeAsync(blk, stream);
    return rc;
}

}  // namespace b2
