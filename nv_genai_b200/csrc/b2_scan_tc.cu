// Small-batch exact scan on the tensor cores: up to 32 queries (up to 32 for k <= 64, 16 for k <= 128) in one pass over
// the rows; from 1 query on a row store of 8 GB or more, from 5 (fp32 rows) / 3 (bf16 rows) on a smaller one.
//
// Replaces faiss exhaustive_inner_product_seq / exhaustive_L2sqr_seq + HeapResultHandler (faiss/utils/distances.cpp,
// the nq < 20 path of IndexFlat::search), the same routines b2_scan.cu replaces, for the query counts
// where the FP32 pipe cannot keep up with HBM: 16 queries x 10M x 1024 rows are 1.6e11 FMAs, 4.7 ms
// at the FFMA peak against 5.8 ms of HBM time, and the FFMA scan reaches 13.7 ms.  Reached from
// index.search at /root/reference/RetrievalAugmentedGeneration/examples/nvidia_api_catalog/chains.py:122-123
// (BASELINE.json configs[1], batch 1 and 16).
//
// Default on fp32 rows with k <= 64 = FILTER AND REFINE (template parameter APX, scan_tc_search_approx): ONE tf32
// product per score straight from the raw tile -- no q_lo half, no transform warps, the kernel only streams the rows
// (measured at 10M x 1024: 5.5 ms at 1 query, 6.0 ms at 16 = 1.13 / 1.04 of the copy peak) -- every CTA emits its
// k + slack best approximate candidates per query and merge_rescore re-scores everything within the rigorous error
// band of the k-th best in fp32 (inner product: the FFMA scan's dot product; L2: faiss's direct form sum (x - q)^2,
// the filter keys being 2 q.x - |x|^2).  Queries whose band reaches the end of a full CTA list are redone exactly.
// The all-exact arithmetic below serves bf16 rows, k > 64 and B2_SCAN_TC_EXACT=1 (inner product only).
//
// Mapping: the CORPUS is the M operand.  A tile is 128 consecutive rows (TMEM lanes), the 16 queries
// are the N operand, K = the row, streamed by TMA in 128-byte swizzled slices straight from the row
// store; the query slice ([q_hi; q_lo], 32 x 128 B) rides along with every row slice (L2-resident).
//   bf16 rows:  D1[128 x 32] += A . [q_hi; q_lo]^T                      (kind::f16; q kept to ~16 bits)
//   fp32 rows:  D1[128 x 32] += A_hi . [q_hi; q_lo]^T,  D2[128 x 16] += A_lo . q_hi^T   (3xTF32)
//               A_hi is the raw tile (the tensor core ignores the low 13 mantissa bits),
//               A_lo = x - trunc_tf32(x) is computed by eight transform warps (thread = row) and stored
//               straight into TENSOR MEMORY (tcgen05.st, 4-slot ring of 32 columns); the second MMA
//               takes its A operand from TMEM, so the remainder never costs shared-memory bandwidth
//               (with A_lo in shared memory the kernel was smem-bound: 8.4 ms vs 5.8 ms of HBM time).
//   score(row, j) = sum of D1[row, j] + (sum of D1[row, 16 + j] + D2[row, j]) over four accumulator chains
//   (one per 32-byte K step of a slice), added in fp32 registers.
// MMA time per slice is 64-96 cycles against ~650 cycles of HBM time per 16 KB slice and SM, so the
// kernel is HBM-bound like the FFMA scan at batch 1; the tensor pipe is ~15 % busy by design.
// Selection: one thread = one row, per-(warp, query) lists in shared memory guarded by a threshold, CTA
// merge at the end, the FFMA scan's partial-list format, merge kernel and tie rule (equal scores -> lowest
// id).  Lists of 32 slots that keep at most 28 entries (k <= ~20) run as reservoirs: the lanes whose row
// passes append side by side in one step and only a full reservoir is ranked (all pairs, by shuffles);
// longer lists replace their worst entry per admission (warp_list_insert), as the FFMA scan does.
//
// Algorithmic HBM bytes per launch = n * pitch_bytes (each row read exactly once).
#include "b2_kernels.h"
#include "b2_tc.cuh"

#ifndef B2_SCAN_TC_CHAINS4
#define B2_SCAN_TC_CHAINS4 0
#endif

namespace b2 {

using namespace tc;

namespace {

constexpr int kTileRows = 128;
constexpr int kMaxNQ = 32;                       // queries per pass: 16 or 32 (padded; template parameter NQ).  A tile of 64 was
                                                 // measured slower than the GEMM kernel (fp32 12.3 vs 11.0 ms, bf16 8.0 vs 6.3 ms)
constexpr int kSlice = 128;                      // bytes of K per stage (swizzle width)
constexpr int kATile = kTileRows * kSlice;       // 16 KB
constexpr int kLoSlots = 4;                       // fp32 rows: A_lo slices resident in TMEM
constexpr int kMaxStages = 10;
constexpr int kEpi = 128;                        // epilogue threads
constexpr int kEpiWarps = 4;
constexpr int kXfWarps = 8;                       // fp32 rows: transform warps (256 threads, 4 x 16 B each per slice)
constexpr int kThreadsBf16 = 192, kThreadsFp32 = 192 + kXfWarps * 32;

struct ScanTcParams {
    uint32_t n;
    int nq, k, kpad, kslices, n_tiles, num_stages;
    int sps;                  // K slices per pipeline stage (1, 2 or 4; kslices % sps == 0; the exact fp32 variant uses 1)
    int klist;                // entries each (warp, query) list keeps and each CTA emits: k, or kpad in the approximate mode
    uint64_t* partial;        // [nq][gridDim.x][klist] sorted best-first
    const uint32_t* run_if;   // not null: the whole launch is skipped unless *run_if != 0 (fix-up launch of the approximate mode)
    const float* norms;       // not null (approximate mode only): L2, keys are 2 q.x - |x|^2
};

__device__ __forceinline__ void epi_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// bitonic sort (descending) by the 128 epilogue threads, named barrier 1
__device__ __forceinline__ void epi_bitonic_desc(uint64_t* buf, int n, int tid) {
    for (int size = 2; size <= n; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            epi_sync();
            for (int t = tid; t < (n >> 1); t += kEpi) {
                const int lo = 2 * t - (t & (stride - 1));
                const int hi = lo + stride;
                const bool desc = ((lo & size) == 0);
                const uint64_t a = buf[lo], b = buf[hi];
                if ((a < b) == desc) {
                    buf[lo] = b;
                    buf[hi] = a;
                }
            }
        }
    }
    epi_sync();
}

// KIND 0: bf16 rows (kind::f16).  KIND 1: fp32 rows (kind::tf32, 3xTF32).  NQ: queries per pass (16, 32, 64).
// APX (fp32 rows only): filter-and-refine.  ONE tf32 product per score straight from the raw tile -- no q_lo half, no
// transform warps, no A_lo in TMEM -- so the kernel has nothing left to do but stream the rows; each CTA emits its
// kpad (> k) best approximate candidates per query, and merge_rescore re-scores in fp32 everything within the error
// band of the k-th best (b2_merge.cu).  A query whose band does not fit a CTA's list is redone by the exact kernel.
template <int KIND, int NQ, bool APX>
__global__ void __launch_bounds__((KIND == 1 && !APX) ? kThreadsFp32 : kThreadsBf16, 1)
scan_tc_kernel(const __grid_constant__ CUtensorMap tmRows, const __grid_constant__ CUtensorMap tmQ, const ScanTcParams p) {
    if (p.run_if != nullptr && *p.run_if == 0u) return;  // uniform over the grid, before anything is allocated
    constexpr int ELEM = (KIND == 0) ? 2 : 4;
    constexpr int KSLICE = kSlice / ELEM;
    constexpr uint32_t FMT = (KIND == 0) ? 1u : 2u;
    constexpr uint32_t IDESC_BASE = (1u << 4) | (FMT << 7) | (FMT << 10) | ((uint32_t)(kTileRows >> 4) << 24);
    constexpr int kNQ = NQ;
    constexpr int kBTile = (APX ? 1 : 2) * NQ * kSlice;  // [q_hi; q_lo]: 4 / 8 / 16 KB (q_hi alone when APX)
    constexpr int kStage = kATile + kBTile;
    constexpr uint32_t IDESC_N32 = IDESC_BASE | (((uint32_t)(2 * NQ) >> 3) << 17);  // D1: N = 2 NQ
    constexpr uint32_t IDESC_N16 = IDESC_BASE | (((uint32_t)NQ >> 3) << 17);        // D2: N = NQ
    // TMEM: 2 tile buffers x kChains accumulator chains x 3 NQ columns (D1: 2 NQ, D2: NQ), then 4 x 32 columns of A_lo
    // slices.  With -DB2_SCAN_TC_CHAINS4=1 the four 32-byte K steps of a slice accumulate into four independent
    // chains that the epilogue adds up (shorter chains halve the truncation bias); measured speed is the same
    // (7.35 / 3.57 ms at 16 queries either way), so the default is one chain and three TMEM loads per tile.
    constexpr uint32_t kChains = (NQ == 16 && B2_SCAN_TC_CHAINS4 && !APX) ? kSlice / 32 : 1;  // more queries: one chain (TMEM has 512 columns)
    constexpr uint32_t kChainCols = APX ? (uint32_t)NQ : 3u * NQ;  // D1 (2 NQ) + D2 (NQ); APX: D1 (NQ)
    constexpr uint32_t kBufCols = kChains * kChainCols;         // 192 / 96 / 192
    constexpr uint32_t kTmemCols = 512;
    constexpr uint32_t kLoCol0 = 2 * kBufCols;    // 384

    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    uint8_t* smem = smem_raw + (((raw + 1023u) & ~1023u) - raw);
    const int S = p.num_stages;
    const int SPS = p.sps;
    const uint32_t stage_bytes = (uint32_t)SPS * kStage;  // a stage = SPS consecutive K slices of the same 128 rows
    uint8_t* stages = smem;
    uint64_t* lists = reinterpret_cast<uint64_t*>(stages + (size_t)S * stage_bytes);  // [kNQ][4 warps][kpad]
    uint64_t* worst_s = lists + (size_t)kNQ * kEpiWarps * p.kpad;                                    // [kNQ][4]
    int* worstpos_s = reinterpret_cast<int*>(worst_s + kNQ * kEpiWarps);                             // [kNQ][4]
    uint64_t* bars = reinterpret_cast<uint64_t*>(worstpos_s + kNQ * kEpiWarps);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + kMaxStages;
    uint64_t* xf_bar = bars + 2 * kMaxStages;
    uint64_t* lo_empty_bar = bars + 3 * kMaxStages;
    uint64_t* tfull_bar = lo_empty_bar + kLoSlots;
    uint64_t* tempty_bar = tfull_bar + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
        for (int s = 0; s < S; ++s) {
            mbar_init(smem_u32(full_bar + s), 1);
            mbar_init(smem_u32(empty_bar + s), 1);
            mbar_init(smem_u32(xf_bar + s), kXfWarps);
        }
        for (int s = 0; s < kLoSlots; ++s) mbar_init(smem_u32(lo_empty_bar + s), 1);
        for (int b = 0; b < 2; ++b) {
            mbar_init(smem_u32(tfull_bar + b), 1);
            mbar_init(smem_u32(tempty_bar + b), kEpiWarps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 5) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (warp < kEpiWarps) {
        for (int i = tid; i < kNQ * kEpiWarps * p.kpad; i += kEpi) lists[i] = B2_SENTINEL;
        for (int i = tid; i < kNQ * kEpiWarps; i += kEpi) {
            worst_s[i] = B2_SENTINEL;
            worstpos_s[i] = 0;
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 4) {
        // ================= TMA producer (whole warp loops, one elected lane issues)
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
            for (int ks = 0; ks < p.kslices; ks += SPS) {
                mbar_wait(smem_u32(empty_bar + stage), phase ^ 1u);
                if (elect_one()) {
                    const uint32_t fb = smem_u32(full_bar + stage);
                    mbar_expect_tx(fb, stage_bytes);
                    const uint32_t sb = smem_u32(stages + (size_t)stage * stage_bytes);
                    // the SPS slices of a row are neighbours in memory and are requested back to back, so DRAM sees
                    // SPS x 128 contiguous bytes per row instead of 128 bytes from each of 128 different pages
                    for (int u = 0; u < SPS; ++u) {
                        tma_load_2d(sb + u * kStage, &tmRows, (ks + u) * KSLICE, tile * kTileRows, fb);
                        tma_load_2d(sb + u * kStage + kATile, &tmQ, (ks + u) * KSLICE, 0, fb);
                    }
                }
                __syncwarp();
                if (++stage == S) {
                    stage = 0;
                    phase ^= 1u;
                }
            }
        }
    } else if (warp == 5) {
        // ================= MMA issuer (whole warp loops, one elected lane issues)
        int stage = 0, lo_slot = 0;
        uint32_t phase = 0, tc = 0;
        for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++tc) {
            const uint32_t buf = tc & 1u, par = (tc >> 1) & 1u;
            mbar_wait(smem_u32(tempty_bar + buf), par ^ 1u);
            tcgen05_fence_after();
            const uint32_t dbase = tmem_base + buf * kBufCols;
            for (int ks = 0; ks < p.kslices; ks += SPS) {
                mbar_wait(smem_u32(((KIND == 1 && !APX) ? xf_bar : full_bar) + stage), phase);
                tcgen05_fence_after();
                if (elect_one()) {
                    for (int u = 0; u < SPS; ++u) {
                    const uint32_t a = smem_u32(stages + (size_t)stage * stage_bytes + (size_t)u * kStage);
                    const uint64_t ad = smem_desc<kSlice>(a);
                    const uint64_t bd = smem_desc<kSlice>(a + kATile);
                    const uint32_t alo = tmem_base + kLoCol0 + (uint32_t)lo_slot * 32u;  // TMEM columns of this slice's A_lo
#pragma unroll
                    for (int j = 0; j < kSlice / 32; ++j) {
                        const uint32_t acc = (kChains == 1 ? (ks | u | j) : ks) ? 1u : 0u;
                        const uint32_t d1 = dbase + (kChains == 1 ? 0u : j * kChainCols), d2 = d1 + 2u * NQ;
                        umma<KIND>(d1, ad + (uint64_t)(j * 2), bd + (uint64_t)(j * 2), APX ? IDESC_N16 : IDESC_N32, acc);  // +32 B in the address field
                        if (KIND == 1 && !APX) umma_tf32_ts(d2, alo + j * 8, bd + (uint64_t)(j * 2), IDESC_N16, acc);
                    }
                    }
                    tcgen05_commit(smem_u32(empty_bar + stage));
                    if (KIND == 1 && !APX) tcgen05_commit(smem_u32(lo_empty_bar + lo_slot));
                }
                __syncwarp();
                if (KIND == 1 && !APX && ++lo_slot == kLoSlots) lo_slot = 0;
                if (++stage == S) {
                    stage = 0;
                    phase ^= 1u;
                }
            }
            if (elect_one()) tcgen05_commit(smem_u32(tfull_bar + buf));
            __syncwarp();
        }
    } else if (warp >= 6) {
        // ================= fp32 rows: A_lo = x - trunc_tf32(x), thread = row, written to TMEM
        if (KIND == 1 && !APX) {
            const int quarter = warp & 3;                 // the TMEM lanes this warp may touch: 32 * (warp % 4) ..
            const int half = (warp - 6) >> 2;             // which 64 bytes of the 128-byte row slice
            const int r = quarter * 32 + lane;            // row of the tile = TMEM lane
            int stage = 0, lo_slot = 0;
            uint32_t phase = 0, lo_phase = 0;
            for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
                for (int ks = 0; ks < p.kslices; ++ks) {
                    mbar_wait(smem_u32(full_bar + stage), phase);
                    mbar_wait(smem_u32(lo_empty_bar + lo_slot), lo_phase ^ 1u);
                    tcgen05_fence_after();
                    const uint8_t* rowp = stages + (size_t)stage * stage_bytes + r * kSlice;  // SPS == 1 in this variant
                    uint32_t lo[16];
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const int chunk = (half * 4 + c) ^ (r & 7);  // 128-byte swizzle: 16-byte chunk index XOR row % 8
                        const float4 x = *reinterpret_cast<const float4*>(rowp + chunk * 16);
                        lo[4 * c + 0] = __float_as_uint(x.x - __uint_as_float(__float_as_uint(x.x) & 0xffffe000u));
                        lo[4 * c + 1] = __float_as_uint(x.y - __uint_as_float(__float_as_uint(x.y) & 0xffffe000u));
                        lo[4 * c + 2] = __float_as_uint(x.z - __uint_as_float(__float_as_uint(x.z) & 0xffffe000u));
                        lo[4 * c + 3] = __float_as_uint(x.w - __uint_as_float(__float_as_uint(x.w) & 0xffffe000u));
                    }
                    tmem_st16(tmem_base + kLoCol0 + (uint32_t)lo_slot * 32u + (uint32_t)half * 16u + ((uint32_t)(quarter * 32) << 16), lo);
                    tmem_st_wait();
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(smem_u32(xf_bar + stage));
                    if (++lo_slot == kLoSlots) {
                        lo_slot = 0;
                        lo_phase ^= 1u;
                    }
                    if (++stage == S) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
        }
    } else {
        // ================= epilogue: warps 0..3, thread = corpus row = TMEM lane
        uint32_t tc = 0;
        // lists of 32 slots that keep at most 28: run them as reservoirs (see below); longer lists keep the insertion path
        const bool reservoir = p.kpad == 32 && p.klist <= 28;
        for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++tc) {
            const uint32_t buf = tc & 1u, par = (tc >> 1) & 1u;
            mbar_wait(smem_u32(tfull_bar + buf), par);
            tcgen05_fence_after();
            const uint32_t taddr = tmem_base + buf * kBufCols + ((uint32_t)(warp * 32) << 16);
            const uint32_t row = (uint32_t)tile * kTileRows + warp * 32 + lane;
            const float nrm = (APX && p.norms != nullptr && row < p.n) ? __ldg(p.norms + row) : 0.f;
            // queries in groups of 16: columns [g*16, +16) of the q_hi block, of the q_lo block and of D2
#pragma unroll 1
            for (int g = 0; g < NQ / 16; ++g) {
                float big[16], small[16];  // main products / correction products, summed over the chains
#pragma unroll
                for (int j = 0; j < 16; ++j) big[j] = small[j] = 0.f;
#pragma unroll
                for (int c = 0; c < (int)kChains; ++c) {
                    uint32_t dh[16], dl[16], d2[16];
                    tmem_ld16(taddr + c * kChainCols + g * 16u, dh);
                    if (!APX) tmem_ld16(taddr + c * kChainCols + NQ + g * 16u, dl);
                    if (KIND == 1 && !APX) tmem_ld16(taddr + c * kChainCols + 2u * NQ + g * 16u, d2);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        big[j] += __uint_as_float(dh[j]);
                        if (!APX) small[j] += __uint_as_float(dl[j]);
                        if (KIND == 1 && !APX) small[j] += __uint_as_float(d2[j]);
                    }
                }
                if (g == NQ / 16 - 1) {
                    tcgen05_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(smem_u32(tempty_bar + buf));  // the last group is in registers: TMEM is free
                }
#pragma unroll
                for (int jj = 0; jj < 16; ++jj) {
                    const int j = g * 16 + jj;
                    if (j < p.nq) {  // warp-uniform
                        const float key = (APX && p.norms != nullptr) ? fmaf(2.f, big[jj], -nrm) : (big[jj] + small[jj]) + 0.0f;
                        const uint64_t cand = pack_cand(key, row);
                        const int li = j * kEpiWarps + warp;
                        if (reservoir) {
                            // 32-slot reservoir per (warp, query): passing lanes append side by side in one step (slot =
                            // count + rank among the passing lanes); only a full reservoir is reduced to its klist best
                            // - every lane holds one entry and ranks it against the other 31 by shuffles - and the
                            // klist-th best becomes the threshold.  An admission costs a store instead of the
                            // replace-the-worst-and-rescan chain of warp_list_insert, which is what the thresholds'
                            // settling phase of every CTA (~k (1 + ln(rows per CTA / k)) admissions per query, all
                            // 148 CTAs at once) used to cost: ~300 us per launch at 16 queries, whatever the row count.
                            uint64_t* L = lists + (size_t)li * p.kpad;
                            uint64_t thr = worst_s[li];
                            int cnt = worstpos_s[li];  // entries in the reservoir
                            bool pass = row < p.n && key > -FLT_MAX && cand > thr;
                            unsigned m = __ballot_sync(kFull, pass);
                            if (m) {
                                const unsigned lt = (1u << lane) - 1u;
                                while (m) {
                                    const int room = 32 - cnt;
                                    if (room == 0) {
                                        const uint64_t e = L[lane];
                                        int rank = 0;
#pragma unroll 8
                                        for (int s2 = 0; s2 < 32; ++s2) rank += (shfl_u64(e, s2) > e) ? 1 : 0;  // entries are distinct
                                        __syncwarp();
                                        if (rank < p.klist) L[rank] = e;  // sorted best-first as a by-product
                                        if (lane >= p.klist) L[lane] = B2_SENTINEL;  // the slots behind them are free again
                                        const int src = __ffs(__ballot_sync(kFull, rank == p.klist - 1)) - 1;
                                        thr = shfl_u64(e, src);
                                        cnt = p.klist;
                                        __syncwarp();
                                        pass = pass && cand > thr;
                                        m = __ballot_sync(kFull, pass);
                                        continue;
                                    }
                                    const int r = __popc(m & lt);
                                    if (pass && r < room) {
                                        L[cnt + r] = cand;
                                        pass = false;
                                    }
                                    cnt += min(__popc(m), room);
                                    m = __ballot_sync(kFull, pass);
                                }
                                __syncwarp();
                                if (lane == 0) {
                                    worst_s[li] = thr;
                                    worstpos_s[li] = cnt;
                                }
                                __syncwarp();
                            }
                            continue;
                        }
                        const bool pass = row < p.n && key > -FLT_MAX && cand > worst_s[li];
                        unsigned m = __ballot_sync(kFull, pass);
                        while (m) {
                            const int src = __ffs(m) - 1;
                            m &= m - 1;
                            const uint64_t c = shfl_u64(cand, src);
                            warp_list_insert(lists + (size_t)li * p.kpad, p.klist, c, worst_s + li, worstpos_s + li, lane);
                        }
                    }
                }
            }
        }
        // ---- CTA merge: per query, sort the four warps' lists (contiguous) and emit the best k
        const int region = kEpiWarps * p.kpad;
        for (int qi = 0; qi < p.nq; ++qi) {
            uint64_t* buf = lists + (size_t)qi * region;
            epi_bitonic_desc(buf, region, tid);
            uint64_t* out = p.partial + ((size_t)qi * gridDim.x + blockIdx.x) * p.klist;
            for (int i = tid; i < p.klist; i += kEpi) out[i] = buf[i];
        }
    }

    tcgen05_fence_before();
    __syncthreads();
    if (warp == 5) {
        tcgen05_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

size_t scan_tc_smem_bytes(int nqt, int stages, int kpad, int parts = 2, int sps = 1) {
    const size_t stage = ((size_t)kATile + (size_t)parts * nqt * kSlice) * sps;
    return 1024 + (size_t)stages * stage + (size_t)nqt * kEpiWarps * kpad * 8 + (size_t)nqt * kEpiWarps * 12 +
           (3 * kMaxStages + kLoSlots + 4) * 8 + 16;
}

// queries one pass can take at this k: the per-(warp, query) lists (4 x kpad x 8 B each) get at most 64 KB
int scan_tc_max_queries(int k) {
    const int kpad = next_pow2(k < 32 ? 32 : k);
    if (kpad > 128) return 0;
    return std::min(kMaxNQ, 2048 / kpad);  // kpad 32, 64 -> 32 queries, 128 -> 16
}

template <int KIND, int NQ, bool APX = false>
int launch_scan_tc(const CUtensorMap& tmRows, const CUtensorMap& tmQ, const ScanTcParams& p, int grid, size_t smem,
                   cudaStream_t stream) {
    auto kern = scan_tc_kernel<KIND, NQ, APX>;
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
    void* tok = p.run_if != nullptr ? nullptr : prof_begin(stream);  // the (normally skipped) fix-up launch is not the profiled kernel
    kern<<<grid, (KIND == 1 && !APX) ? kThreadsFp32 : kThreadsBf16, smem, stream>>>(tmRows, tmQ, p);
    prof_end(tok, 0, stream);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

// Everything the approximate scan needs from the queries, in ONE launch (it used to be a memset and three kernels:
// at one query the fixed cost of the launch chain, not the row stream, is what a 5 GB shard notices): the rounded
// query operand Q (tf32 on fp32 rows, bf16 on bf16 rows), the two-part operand Qx of the exact fix-up scan (fp32
// rows: tf32 hi | x - hi), the error band, and the overflow flags cleared.  One warp per row of the query tile.
template <bool BF16>
__global__ void scan_tc_prep_kernel(const float* __restrict__ q, int nq, int d, int nqt, int kp, void* __restrict__ Q,
                                    float* __restrict__ Qx, int steps, int l2, const uint32_t* __restrict__ max_norm_ord,
                                    float* __restrict__ band, uint32_t* __restrict__ flags,
                                    const uint32_t* __restrict__ row_err_ord = nullptr) {
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (blockIdx.x == 0)
        for (int i = threadIdx.x; i <= nq; i += blockDim.x) flags[i] = 0u;
    if (row >= nqt) return;
    float ss = 0.f, se = 0.f;
    for (int c = lane; c < kp; c += 32) {
        const float x = (row < nq && c < d) ? q[(size_t)row * d + c] : 0.f;
        float h;
        if (BF16) {
            const __nv_bfloat16 hb = __float2bfloat16_rn(x);
            static_cast<__nv_bfloat16*>(Q)[(size_t)row * kp + c] = hb;
            h = __bfloat162float(hb);
        } else {
            uint32_t u;
            asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
            h = __uint_as_float(u);
            static_cast<float*>(Q)[(size_t)row * kp + c] = h;
            if (Qx != nullptr) {
                Qx[(size_t)row * kp + c] = h;
                Qx[((size_t)nqt + row) * kp + c] = x - h;
            }
        }
        ss = fmaf(h, h, ss);
        se = fmaf(x - h, x - h, se);
    }
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) {
        ss += __shfl_xor_sync(kFull, ss, s);
        se += __shfl_xor_sync(kFull, se, s);
    }
    // row_err: the rows the scores are finally computed against differ from what the tensor core reads (residual plane)
    if (lane == 0)
        band[row] = band_from_sums(ss, se, !BF16, steps, l2 != 0, d, ord2f(*max_norm_ord),
                                   row_err_ord != nullptr ? ord2f(*row_err_ord) : 0.f);
}

}  // namespace

bool scan_tc_approx_feasible(int64_t n, int k, int num_sms) {
    return (long long)std::min<long long>(num_sms, (n + kTileRows - 1) / kTileRows) * next_pow2(k < 32 ? 32 : k) <= kMergeSelectMax;
}

bool scan_tc_supported(int64_t n, int d, int dtype, int metric, int k, int64_t nq, bool approx, bool residual) {
    if (getenv("B2_NO_TENSOR") != nullptr || getenv("B2_NO_SCAN_TC") != nullptr) return false;
    if (residual && (k > 64 || !approx || getenv("B2_SCAN_TC_EXACT") != nullptr)) return false;
    const int elem = dtype == B2_DTYPE_BF16 ? 2 : 4;
    // L2 rides the approximate mode only (k <= 64; bf16 rows too: one bf16 product): keys 2 q.x - |x|^2 for the filter, the direct form
    // sum (x - q)^2 -- faiss's arithmetic below 20 queries -- for the candidates that are re-scored
    static const bool exact_only = getenv("B2_SCAN_TC_EXACT") != nullptr;
    // (from 20 queries on, L2 takes the GEMM kernel and the |q|^2 + |x|^2 - 2 q.x form, as faiss does)
    if (metric == B2_METRIC_L2 && (k > 64 || !approx || exact_only || nq >= 20)) return false;
    // Measured at 10M x 1024 on B200 (kernel time, one pass): fp32 rows, FFMA scan 5.7-5.9 ms at 1 query, 6.1 ms at 4,
    // 7.2 ms at 5..8; this kernel 5.5 ms at 1, 5.7 ms at 2..4, 6.0 ms at 16, 6.8 ms at 32.  bf16 rows: FFMA scan
    // 2.9-3.3 ms at 1..2 queries; this kernel 2.76 ms at 1, 2.83 ms at 2, 3.2 ms at 16.  So on a large index every
    // inner-product batch up to the tile size takes this kernel; on a small one (1M x 1024: 0.61 vs 0.62 ms at one
    // query) the FFMA scan's two launches win until the FP32 pipe falls behind (5 queries fp32, 3 bf16).
    // (fp32 rows with k > 64 run the all-exact variant, 7.3 ms: the small thresholds stay)
    int min_nq = (dtype == B2_DTYPE_BF16) ? 3 : 5;
    static const double min_bytes = getenv("B2_SCAN_TC_MIN_BYTES") ? atof(getenv("B2_SCAN_TC_MIN_BYTES")) : 8e9;
    if ((double)n * d * elem >= min_bytes && (dtype == B2_DTYPE_BF16 || k <= 64)) min_nq = 1;
    if (const char* e = getenv("B2_SCAN_TC_MIN_NQ")) min_nq = atoi(e);
    // up to 32 queries ride along in one pass when k is small (gemm path at 20..128 queries: 10-11.6 ms fp32,
    // 5.9-6.7 ms bf16 at 10M x 1024, against the 5.8 / 3.1 ms it takes to stream the rows once)
    return nq >= min_nq && nq <= scan_tc_max_queries(k) && n >= 4096 && n < ((int64_t)1 << 31) &&
           (int64_t)d * elem >= kSlice;
}

// K slices per pipeline stage (see the producer loop) and the number of stages that fit next to the lists.  Measured at
// 10M x 1024, 16 queries: 1 / 2 / 4 slices per stage -> 6.87 / 6.79 / 6.64 ms on fp32 rows, 3.61 / 3.53 / 3.43 ms on bf16
// rows (each row then gets its requests 512 contiguous bytes at a time).  The exact fp32 variant needs 1.
static void pick_pipeline(int kslices, bool sps_allowed, int nqt, int kpad, int parts, int* sps_out, int* stages_out) {
    int sps = 4;
    if (const char* e = getenv("B2_SCAN_TC_SPS")) sps = atoi(e);
    if (!sps_allowed || (sps != 2 && sps != 4)) sps = 1;
    for (;; sps >>= 1) {
        int stages = kMaxStages;
        while (stages > 2 && scan_tc_smem_bytes(nqt, stages, kpad, parts, sps) > 227 * 1024) --stages;
        const bool fits = scan_tc_smem_bytes(nqt, stages, kpad, parts, sps) <= 227 * 1024;
        if ((fits && kslices % sps == 0) || sps == 1) {
            *sps_out = sps;
            *stages_out = fits ? stages : 0;
            return;
        }
    }
}

// k <= 64: filter and refine (see scan_tc_kernel, APX) - fp32 rows for both metrics, bf16 rows for L2 and, with a
// residual plane, for inner product too.  Launches: scan_tc_prep_kernel (operands, band, cleared flags) + approximate
// scan + merge_rescore, then the exact fix-up pair -- exact tensor-core scan (fp32 rows, inner product) or FFMA scan
// passes (L2; bf16 rows, both planes) + merge -- whose kernels return at once unless merge_rescore raised a query's
// overflow flag (a CTA would have to hold a full list of candidates within the error band of the k-th best: runs of
// near-duplicate rows do that, random data never).
static int scan_tc_search_approx(const void* rows, int64_t n, size_t pitch_bytes, int d, const float* q_dev, int nq, int k,
                                 int num_sms, float* D, int64_t* I, cudaStream_t stream, const PeerPush* push,
                                 const uint32_t* max_norm_ord, int metric, const float* norms, int dtype,
                                 const ResidualRows* resid) {
    const bool l2 = metric == B2_METRIC_L2;
    const bool bf16 = dtype == B2_DTYPE_BF16;  // bf16 rows come here for L2 only: one bf16 product, the stored rows are exact
    const int elem = bf16 ? 2 : 4;
    if (l2 && norms == nullptr) {
        set_error("tensor-core scan: L2 needs the rows' squared norms");
        return B2_ERR_INVALID;
    }
    const size_t kbytes = (((size_t)d * elem + kSlice - 1) / kSlice) * kSlice;
    const int kp = (int)(kbytes / elem);
    ScanTcParams p;
    p.n = (uint32_t)n;
    p.nq = nq;
    p.k = k;
    p.kpad = next_pow2(k < 32 ? 32 : k);
    // every list keeps k plus a little room for the candidates inside the error band (a CTA sees 1/148 of the rows: a
    // full list whose last entry is still inside the band is the rare overflow case).  Longer lists cost insertions
    // while the thresholds settle: with kpad = 32 entries at k = 10 the start-up phase cost 0.7 ms at 16 queries.
    p.klist = std::min(p.kpad, k + std::max(4, k / 4));
    p.kslices = (int)(kbytes / kSlice);
    p.n_tiles = (int)((n + kTileRows - 1) / kTileRows);
    p.run_if = nullptr;
    p.norms = l2 ? norms : nullptr;
    const int nqt = nq <= 16 ? 16 : 32;
    int stages = 0, stages_x = 0, sps_x = 1;
    pick_pipeline(p.kslices, true, nqt, p.kpad, 1, &p.sps, &stages);
    pick_pipeline(p.kslices, false, nqt, p.kpad, 2, &sps_x, &stages_x);
    const size_t smem = scan_tc_smem_bytes(nqt, stages, p.kpad, 1, p.sps), smem_x = scan_tc_smem_bytes(nqt, stages_x, p.kpad, 2);
    if (stages == 0 || stages_x == 0) {
        set_error("tensor-core scan: k=%d does not fit shared memory", k);
        return B2_ERR_UNSUPPORTED;
    }
    p.num_stages = stages;
    const int grid = (int)std::min<long long>(num_sms, p.n_tiles);

    auto al = [](size_t v) { return (v + 1023) / 1024 * 1024; };
    const size_t q_bytes = al((size_t)nqt * kbytes), qx_bytes = al((size_t)2 * nqt * kbytes);
    const size_t part_bytes = al((size_t)nq * grid * std::max(p.klist, k) * 8);
    const size_t band_bytes = al((size_t)nqt * 4), flag_bytes = al((size_t)(nq + 1) * 4);
    char* blk = nullptr;
    B2_CUDA_TRY(cudaMallocAsync(&blk, q_bytes + qx_bytes + part_bytes + band_bytes + flag_bytes, stream));
    void* Q = blk;
    void* Qx = blk + q_bytes;
    p.partial = reinterpret_cast<uint64_t*>(blk + q_bytes + qx_bytes);
    float* band = reinterpret_cast<float*>(blk + q_bytes + qx_bytes + part_bytes);
    uint32_t* flags = reinterpret_cast<uint32_t*>(blk + q_bytes + qx_bytes + part_bytes + band_bytes);

    CUtensorMap tmRows, tmQ, tmQx;
    int promo = 256;
    if (const char* e = getenv("B2_SCAN_TC_L2PROMO")) promo = atoi(e);
    int rc = make_map(&tmRows, bf16, kSlice, rows, (uint64_t)(pitch_bytes / elem), (uint64_t)n, pitch_bytes, kTileRows, promo);
    if (rc == B2_OK) rc = make_map(&tmQ, bf16, kSlice, Q, (uint64_t)kp, (uint64_t)nqt, kbytes, nqt);
    if (rc == B2_OK) rc = make_map(&tmQx, bf16, kSlice, Qx, (uint64_t)kp, (uint64_t)2 * nqt, kbytes, 2 * nqt);
    if (rc == B2_OK) {
        const int steps = bf16 ? kp / 16 : kp / 8;
        if (bf16)
            scan_tc_prep_kernel<true><<<(nqt + 7) / 8, 256, 0, stream>>>(q_dev, nq, d, nqt, kp, Q, nullptr, steps, l2 ? 1 : 0,
                                                                        max_norm_ord, band, flags,
                                                                        resid != nullptr ? resid->err_ord : nullptr);
        else
            scan_tc_prep_kernel<false><<<(nqt + 7) / 8, 256, 0, stream>>>(q_dev, nq, d, nqt, kp, Q, l2 ? nullptr : static_cast<float*>(Qx),
                                                                         steps, l2 ? 1 : 0, max_norm_ord, band, flags);
        B2_COUNT_LAUNCH();
        if (cudaGetLastError() != cudaSuccess) {
            set_error("tensor-core scan: query preparation launch failed");
            rc = B2_ERR_CUDA;
        }
    }
    if (rc == B2_OK) {
        if (bf16)
            rc = nqt == 16 ? launch_scan_tc<0, 16, true>(tmRows, tmQ, p, grid, smem, stream)
                           : launch_scan_tc<0, 32, true>(tmRows, tmQ, p, grid, smem, stream);
        else
            rc = nqt == 16 ? launch_scan_tc<1, 16, true>(tmRows, tmQ, p, grid, smem, stream)
                           : launch_scan_tc<1, 32, true>(tmRows, tmQ, p, grid, smem, stream);
    }
    if (rc == B2_OK)
        rc = merge_rescore(p.partial, nq, grid, p.klist, k, metric, D, I, (size_t)k, stream, nullptr, q_dev, d, rows,
                           pitch_bytes, dtype, norms, band, push, flags, l2, resid != nullptr ? resid->rows : nullptr);
    // ---- exact fix-up for the flagged queries (each kernel returns immediately when there is none)
    if (rc == B2_OK && (l2 || bf16)) {
        // L2, and bf16 rows (with a residual plane: both planes): the FFMA scan, a pass per query tile: 16 queries, or fewer when rows are so long that 16
        // query rows do not fit its shared memory (d = 4096: 8)
        int tmax = 16;
        while (tmax > 1 && scan_smem_bytes(tmax, (int)(pitch_bytes / 512), bf16, next_pow2(k < 32 ? 32 : k)) > 200 * 1024) tmax >>= 1;
        for (int off = 0; off < nq && rc == B2_OK; off += tmax) {
            const int valid = std::min(tmax, nq - off);
            uint64_t* part = nullptr;
            int g2 = 0;
            rc = scan_pass(rows, n, pitch_bytes, d, dtype, metric, q_dev + (size_t)off * d, valid,
                           next_pow2(valid), k, num_sms, stream, &part, &g2, flags + nq,
                           resid != nullptr ? resid->rows : nullptr);
            if (rc != B2_OK) break;
            PeerPush pp;
            if (push != nullptr) {
                pp = *push;
                pp.q_offset += off;
            }
            rc = merge_partials(part, valid, g2, k, metric, 0, D ? D + (size_t)off * k : nullptr,
                                I ? I + (size_t)off * k : nullptr, (size_t)k, stream, nullptr, push != nullptr ? &pp : nullptr,
                                flags + off);
            cudaFreeAsync(part, stream);
        }
        cudaFreeAsync(blk, stream);
        return rc;
    }
    if (rc == B2_OK) {  // Qx was written by scan_tc_prep_kernel
        ScanTcParams px = p;
        px.klist = k;
        px.sps = 1;
        px.num_stages = stages_x;
        px.run_if = flags + nq;
        px.norms = nullptr;
        rc = nqt == 16 ? launch_scan_tc<1, 16>(tmRows, tmQx, px, grid, smem_x, stream)
                       : launch_scan_tc<1, 32>(tmRows, tmQx, px, grid, smem_x, stream);
    }
    if (rc == B2_OK)
        rc = merge_partials(p.partial, nq, grid, k, B2_METRIC_IP, 0, D, I, (size_t)k, stream, nullptr, push, flags);
    cudaFreeAsync(blk, stream);
    return rc;
}

int scan_tc_search(const void* rows, int64_t n, size_t pitch_bytes, int d, int dtype, const float* q_dev, int nq, int k,
                   int num_sms, float* D, int64_t* I, cudaStream_t stream, const PeerPush* push,
                   const uint32_t* max_norm_ord, int metric, const float* norms, const ResidualRows* resid) {
    const bool bf16 = dtype == B2_DTYPE_BF16;
    if (!bf16) resid = nullptr;
    const int elem = bf16 ? 2 : 4;
    static const bool no_approx = getenv("B2_SCAN_TC_EXACT") != nullptr;
    if ((!bf16 || metric == B2_METRIC_L2 || resid != nullptr) && max_norm_ord != nullptr && !no_approx && k <= 64 &&
        scan_tc_approx_feasible(n, k, num_sms))
        return scan_tc_search_approx(rows, n, pitch_bytes, d, q_dev, nq, k, num_sms, D, I, stream, push, max_norm_ord,
                                     metric, norms, dtype, resid);
    if (resid != nullptr) {
        set_error("tensor-core scan: rows with a residual plane need the approximate mode (k <= 64)");
        return B2_ERR_UNSUPPORTED;
    }
    if (metric != B2_METRIC_IP) {
        set_error("tensor-core scan: L2 needs the approximate mode (fp32 rows, k <= 64)");
        return B2_ERR_UNSUPPORTED;
    }
    const size_t kbytes = (((size_t)d * elem + kSlice - 1) / kSlice) * kSlice;
    const int kp = (int)(kbytes / elem);

    ScanTcParams p;
    p.n = (uint32_t)n;
    p.nq = nq;
    p.k = k;
    p.klist = k;
    p.run_if = nullptr;
    p.norms = nullptr;
    p.kpad = next_pow2(k < 32 ? 32 : k);
    p.kslices = (int)(kbytes / kSlice);
    p.n_tiles = (int)((n + kTileRows - 1) / kTileRows);
    const int nqt = nq <= 16 ? 16 : 32;  // query tile of the pass
    int stages = 0;
    pick_pipeline(p.kslices, bf16, nqt, p.kpad, 2, &p.sps, &stages);
    const size_t smem = scan_tc_smem_bytes(nqt, stages, p.kpad, 2, p.sps);
    if (stages == 0) {
        set_error("tensor-core scan: k=%d does not fit shared memory", k);
        return B2_ERR_UNSUPPORTED;
    }
    p.num_stages = stages;
    const int grid = (int)std::min<long long>(num_sms, p.n_tiles);

    auto al = [](size_t v) { return (v + 1023) / 1024 * 1024; };
    const size_t q_bytes = al((size_t)2 * nqt * kbytes);
    const size_t part_bytes = al((size_t)nq * grid * k * 8);
    char* blk = nullptr;
    B2_CUDA_TRY(cudaMallocAsync(&blk, q_bytes + part_bytes, stream));
    void* Q = blk;
    p.partial = reinterpret_cast<uint64_t*>(blk + q_bytes);

    CUtensorMap tmRows, tmQ;
    int promo = 256;  // consecutive K slices of a row are fetched back to back: let L2 bring in 256 B per miss
    if (const char* e = getenv("B2_SCAN_TC_L2PROMO")) promo = atoi(e);
    int rc = make_map(&tmRows, bf16, kSlice, rows, (uint64_t)(pitch_bytes / elem), (uint64_t)n, pitch_bytes, kTileRows, promo);
    if (rc == B2_OK) rc = make_map(&tmQ, bf16, kSlice, Q, (uint64_t)kp, (uint64_t)2 * nqt, kbytes, 2 * nqt);
    if (rc == B2_OK) rc = prep_queries(q_dev, nq, d, nqt, kp, 2, bf16, Q, nullptr, nullptr, stream);
    if (rc == B2_OK) {
        switch ((bf16 ? 0 : 2) + (nqt == 16 ? 0 : 1)) {
            case 0: rc = launch_scan_tc<0, 16>(tmRows, tmQ, p, grid, smem, stream); break;
            case 1: rc = launch_scan_tc<0, 32>(tmRows, tmQ, p, grid, smem, stream); break;
            case 2: rc = launch_scan_tc<1, 16>(tmRows, tmQ, p, grid, smem, stream); break;
            default: rc = launch_scan_tc<1, 32>(tmRows, tmQ, p, grid, smem, stream); break;
        }
    }
    if (rc == B2_OK) rc = merge_partials(p.partial, nq, grid, k, B2_METRIC_IP, 0, D, I, (size_t)k, stream, nullptr, push);
    cudaFreeAsync(blk, stream);
    return rc;
}

}  // namespace b2
