// Block merge of partial top-k lists -> final (D, I).
//
// Replaces the tail of faiss's result handlers (HeapResultHandler::end -> heap_reorder,
// faiss/impl/ResultHandler.h, faiss/utils/Heap.h): the final sort of the k survivors and the
// -1 / +-FLT_MAX padding, reached from index.search.  The same kernel does
// the exchange step of the row-sharded search (per-shard lists after the NCCL allgather).
//
// One CTA per query.  The running best `kpad` candidates sit at the front of a shared-memory
// buffer of SORT_N u64; the rest of the buffer is refilled with the next chunk of candidates
// and the whole buffer is bitonic-sorted, until the input is consumed.
#include <algorithm>
#include <cstdlib>

#include "b2_common.cuh"
#include "b2_kernels.h"

namespace b2 {

constexpr int kMergeThreads = 512;

// Peer stores and flags (row-sharded search over NVLink peer memory, see PeerPush in b2_kernels.h)
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void emit_result(uint64_t c, int metric, int64_t id_base, float* D, int64_t* I,
                                            const float* qn = nullptr) {
    const uint32_t pos = cand_pos(c);
    if (pos == kNoPos) {
        *I = -1;
        *D = (metric == B2_METRIC_IP) ? -FLT_MAX : FLT_MAX;
    } else {
        *I = id_base + (int64_t)pos;
        const float key = cand_key(c);
        if (qn != nullptr)
            *D = fmaxf(*qn - key, 0.f);  // faiss exhaustive_L2sqr_blas: |q|^2 + |x|^2 - 2 q.x, negatives clamped
        else
            *D = (metric == B2_METRIC_IP) ? key : -key;
    }
}

// Result i of query q (index within this launch): to local D / I, or to every peer's receive block.
__device__ __forceinline__ void emit_out(uint64_t c, int q, int i, int k, int metric, int64_t id_base, float* D,
                                         int64_t* I, size_t out_stride, const float* qnorm, const PeerPush& push) {
    if (push.nparts == 0) {
        emit_result(c, metric, id_base, D + (size_t)q * out_stride + i, I + (size_t)q * out_stride + i,
                    qnorm ? qnorm + q : nullptr);
        return;
    }
    float d;
    int64_t id;
    emit_result(c, metric, id_base, &d, &id, qnorm ? qnorm + q : nullptr);
    const size_t off = (size_t)(push.q_offset + q) * k + i;
    const size_t d_base = (size_t)push.nq_total * k * sizeof(int64_t);
    for (int g = 0; g < push.nparts; ++g) {
        reinterpret_cast<int64_t*>(push.block[g])[off] = id;
        reinterpret_cast<float*>(push.block[g] + d_base)[off] = d;
    }
}
// After a query's results are stored everywhere: make them visible system-wide, then raise the query's flag on
// every peer.  Called by all threads of the CTA.
__device__ __forceinline__ void push_signal(int q, const PeerPush& push) {
    if (push.nparts == 0) return;
    __threadfence_system();
    __syncthreads();
    if ((int)threadIdx.x < push.nparts) st_release_sys(push.flag[threadIdx.x] + push.q_offset + q, push.epoch);
}

__global__ void __launch_bounds__(kMergeThreads) merge_partials_kernel(const uint64_t* __restrict__ partial,
                                                                      int total, int k, int kpad, int sort_n,
                                                                      int metric, int64_t id_base, float* D,
                                                                      int64_t* I, size_t out_stride,
                                                                      const float* __restrict__ qnorm,
                                                                      const PeerPush push,
                                                                      const uint32_t* __restrict__ only_if) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* buf = reinterpret_cast<uint64_t*>(smem_raw);
    const int q = blockIdx.x, tid = threadIdx.x;
    if (only_if != nullptr && only_if[q] == 0u) return;
    const uint64_t* in = partial + (size_t)q * total;
    const int cap = sort_n - kpad;
    for (int i = tid; i < kpad; i += kMergeThreads) buf[i] = B2_SENTINEL;
    for (int off = 0; off < total; off += cap) {
        for (int i = tid; i < cap; i += kMergeThreads) buf[kpad + i] = (off + i < total) ? in[off + i] : B2_SENTINEL;
        block_bitonic_desc(buf, sort_n, tid, kMergeThreads);
    }
    __syncthreads();
    for (int i = tid; i < k; i += kMergeThreads) emit_out(buf[i], q, i, k, metric, id_base, D, I, out_stride, qnorm, push);
    push_signal(q, push);
}

// ---------------------------------------------------------------- k <= 32: register merge
// The shared-memory sort above is latency-bound when there is one CTA per query and few queries (72 us for a
// single query's 592 x 10 candidates).  For k <= 32 each warp instead keeps its running best 32 in registers
// (one candidate per lane, sorted across lanes, best in lane 0), takes the candidates 32 at a time, skips a
// chunk whose maximum cannot enter (the usual case after the first few chunks), otherwise sorts it with a
// shuffle bitonic network and merges; the 16 warps' lists are merged by warp 0 at the end.
__device__ __forceinline__ uint64_t warp_sort_desc(uint64_t v, int lane) {
#pragma unroll
    for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
        for (int j = size >> 1; j > 0; j >>= 1) {
            const uint64_t o = shfl_xor_u64(v, j);
            const bool keep_max = (((lane & j) == 0) == ((lane & size) == 0));
            v = keep_max ? (v > o ? v : o) : (v < o ? v : o);
        }
    }
    return v;
}
// best, chunk: sorted descending across lanes -> the 32 largest of the union, sorted descending
__device__ __forceinline__ uint64_t warp_merge_desc(uint64_t best, uint64_t chunk, int lane) {
    const uint64_t rev = shfl_u64(chunk, 31 - lane);
    uint64_t v = best > rev ? best : rev;  // bitonic: holds the top 32 of the union
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) {
        const uint64_t o = shfl_xor_u64(v, j);
        v = ((lane & j) == 0) ? (v > o ? v : o) : (v < o ? v : o);
    }
    return v;
}

__global__ void __launch_bounds__(kMergeThreads) merge_partials_small_kernel(const uint64_t* __restrict__ partial,
                                                                            int total, int k, int metric,
                                                                            int64_t id_base, float* D, int64_t* I,
                                                                            size_t out_stride,
                                                                            const float* __restrict__ qnorm,
                                                                            const PeerPush push,
                                                                            const uint32_t* __restrict__ only_if) {
    __shared__ uint64_t lists[kMergeThreads / 32][32];
    const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (only_if != nullptr && only_if[q] == 0u) return;
    constexpr int kWarps = kMergeThreads / 32;
    const uint64_t* in = partial + (size_t)q * total;
    uint64_t best = B2_SENTINEL;
    for (int c0 = warp * 32; c0 < total; c0 += kWarps * 32) {
        const int i = c0 + lane;
        uint64_t c = (i < total) ? in[i] : B2_SENTINEL;
        const uint64_t worst = shfl_u64(best, 31);
        if (__any_sync(kFull, c > worst)) best = warp_merge_desc(best, warp_sort_desc(c, lane), lane);
    }
    lists[warp][lane] = best;
    __syncthreads();
    if (warp == 0) {
        best = lists[0][lane];
        for (int w = 1; w < kWarps; ++w) {
            const uint64_t c = lists[w][lane];
            const uint64_t worst = shfl_u64(best, 31);
            if (__any_sync(kFull, c > worst)) best = warp_merge_desc(best, c, lane);
        }
        if (lane < k) emit_out(best, q, lane, k, metric, id_base, D, I, out_stride, qnorm, push);
    }
    push_signal(q, push);
}

int merge_partials(const uint64_t* partial, int nq, int nlists, int k, int metric, int64_t id_base, float* D,
                   int64_t* I, size_t out_stride, cudaStream_t stream, const float* qnorm, const PeerPush* push_in,
                   const uint32_t* only_if) {
    const int kpad = next_pow2(k < 32 ? 32 : k);
    const int total = nlists * k;
    PeerPush push;
    if (push_in != nullptr) {
        push = *push_in;
    } else {
        push.nparts = 0;
        push.epoch = 0;
        push.nq_total = push.q_offset = 0;
        for (int g = 0; g < kMaxPushPeers; ++g) {
            push.block[g] = nullptr;
            push.flag[g] = nullptr;
        }
    }
    if (kpad == 32) {
        merge_partials_small_kernel<<<nq, kMergeThreads, 0, stream>>>(partial, total, k, metric, id_base, D, I,
                                                                      out_stride, qnorm, push, only_if);
        B2_AFTER_LAUNCH();
        return B2_OK;
    }
    int sort_n = 2 * kpad;
    const int want = next_pow2(total + kpad);
    while (sort_n < want && sort_n < 4096) sort_n <<= 1;
    const size_t smem = (size_t)sort_n * sizeof(uint64_t);
    merge_partials_kernel<<<nq, kMergeThreads, smem, stream>>>(partial, total, k, kpad, sort_n, metric, id_base, D, I,
                                                               out_stride, qnorm, push, only_if);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

// ---------------------------------------------------------------- merge + exact re-scoring (filter and refine)
// The tensor-core kernel in refine mode scores with ONE approximate product and emits, per (query, segment), every
// candidate whose approximate key is within `band` of the segment's k-th best approximate key, where band = 2 x a
// rigorous bound on |approximate - exact|.  The exact top-k of a query is then contained in
//     { candidates with approximate key > (k-th best approximate key over all segments) - band },
// typically k + a dozen entries.  One CTA per query: pass A finds the k-th best approximate key (chunked bitonic
// sort as above); pass B re-scores exactly the candidates above the cut - one warp per candidate: fp32 FMA dot of
// the fp32 query with the stored row, the arithmetic of the small-batch scan - and keeps the best k exact keys.
// Replaces the same faiss routines as merge_partials_kernel; the scores it returns are fp32-exact with respect to
// the stored rows on both row types.
struct RescoreCtx {
    const float* q;        // [nq][d] fp32 queries (device)
    const char* rows;      // row store
    const char* resid;     // bf16 rows with a residual plane (same pitch): the stored row is hi + lo; null = none
    const float* norms;    // |x|^2 per row (L2 keys)
    const float* band;     // [nq]
    size_t pitch_bytes;
    int d;
    int bf16;
    int l2_direct;  // re-score with sum (x - q)^2, faiss's form below 20 queries (exactly 0 for a stored copy of the query)
};

// sum (x - q)^2 over the row in the FFMA scan's order
__device__ __forceinline__ float warp_exact_l2(const RescoreCtx& c, const float* __restrict__ qrow, uint32_t row, int lane) {
    const char* r = c.rows + (size_t)row * c.pitch_bytes;
    float acc = 0.f;
    if (c.bf16) {
        for (int e0 = lane * 8; e0 < c.d; e0 += 256) {
            const uint4 w = *reinterpret_cast<const uint4*>(r + (size_t)e0 * 2);
            float x[8] = {bf16lo(w.x), bf16hi(w.x), bf16lo(w.y), bf16hi(w.y),
                          bf16lo(w.z), bf16hi(w.z), bf16lo(w.w), bf16hi(w.w)};
            if (c.resid != nullptr) {
                const uint4 y = *reinterpret_cast<const uint4*>(c.resid + (size_t)row * c.pitch_bytes + (size_t)e0 * 2);
                x[0] += bf16lo(y.x); x[1] += bf16hi(y.x); x[2] += bf16lo(y.y); x[3] += bf16hi(y.y);
                x[4] += bf16lo(y.z); x[5] += bf16hi(y.z); x[6] += bf16lo(y.w); x[7] += bf16hi(y.w);
            }
#pragma unroll
            for (int e = 0; e < 8; ++e)
                if (e0 + e < c.d) {
                    const float t = x[e] - __ldg(qrow + e0 + e);
                    acc = fmaf(t, t, acc);
                }
        }
    } else {
        for (int e0 = lane * 4; e0 < c.d; e0 += 128) {
            const float4 w = *reinterpret_cast<const float4*>(r + (size_t)e0 * 4);
            const float x[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (e0 + e < c.d) {
                    const float t = x[e] - __ldg(qrow + e0 + e);
                    acc = fmaf(t, t, acc);
                }
        }
    }
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) acc += __shfl_xor_sync(kFull, acc, s);
    return acc;
}

__device__ __forceinline__ float warp_exact_dot(const RescoreCtx& c, const float* __restrict__ qrow, uint32_t row, int lane) {
    const char* r = c.rows + (size_t)row * c.pitch_bytes;
    float acc = 0.f;
    if (c.bf16) {
        for (int e0 = lane * 8; e0 < c.d; e0 += 256) {
            const uint4 w = *reinterpret_cast<const uint4*>(r + (size_t)e0 * 2);  // 8 bf16; padding columns are zero
            float x[8] = {bf16lo(w.x), bf16hi(w.x), bf16lo(w.y), bf16hi(w.y),
                          bf16lo(w.z), bf16hi(w.z), bf16lo(w.w), bf16hi(w.w)};
            if (c.resid != nullptr) {
                const uint4 y = *reinterpret_cast<const uint4*>(c.resid + (size_t)row * c.pitch_bytes + (size_t)e0 * 2);
                x[0] += bf16lo(y.x); x[1] += bf16hi(y.x); x[2] += bf16lo(y.y); x[3] += bf16hi(y.y);
                x[4] += bf16lo(y.z); x[5] += bf16hi(y.z); x[6] += bf16lo(y.w); x[7] += bf16hi(y.w);
            }
#pragma unroll
            for (int e = 0; e < 8; ++e)
                if (e0 + e < c.d) acc = fmaf(x[e], __ldg(qrow + e0 + e), acc);
        }
    } else {
        for (int e0 = lane * 4; e0 < c.d; e0 += 128) {
            const float4 w = *reinterpret_cast<const float4*>(r + (size_t)e0 * 4);
            const float x[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int e = 0; e < 4; ++e)
                if (e0 + e < c.d) acc = fmaf(x[e], __ldg(qrow + e0 + e), acc);
        }
    }
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) acc += __shfl_xor_sync(kFull, acc, s);
    return acc;
}

__global__ void __launch_bounds__(kMergeThreads) merge_rescore_kernel(const uint64_t* __restrict__ partial, int total,
                                                                     int k, int kpad, int sort_n, int metric,
                                                                     float* D, int64_t* I, size_t out_stride,
                                                                     const float* __restrict__ qnorm,
                                                                     const RescoreCtx ctx, const PeerPush push) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* buf = reinterpret_cast<uint64_t*>(smem_raw);
    const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t* in = partial + (size_t)q * total;
    const int cap = sort_n - kpad;
    // ---- pass A: the k-th best APPROXIMATE key
    for (int i = tid; i < kpad; i += kMergeThreads) buf[i] = B2_SENTINEL;
    for (int off = 0; off < total; off += cap) {
        for (int i = tid; i < cap; i += kMergeThreads) buf[kpad + i] = (off + i < total) ? in[off + i] : B2_SENTINEL;
        block_bitonic_desc(buf, sort_n, tid, kMergeThreads);
    }
    __syncthreads();
    const uint64_t kth = buf[k - 1];
    // fewer than k candidates: everything is re-scored
    const float cut = (cand_pos(kth) == kNoPos) ? -__builtin_huge_valf() : cand_key(kth) - ctx.band[q];
    __syncthreads();
    // ---- pass B: exact keys for the candidates above the cut, running best k of those
    const float* qrow = ctx.q + (size_t)q * ctx.d;
    for (int i = tid; i < kpad; i += kMergeThreads) buf[i] = B2_SENTINEL;
    for (int off = 0; off < total; off += cap) {
        __syncthreads();
        for (int base = warp * 32; base < cap; base += kMergeThreads) {
            const int i = base + lane;
            const uint64_t c = (off + i < total) ? in[off + i] : B2_SENTINEL;
            const uint32_t row = cand_pos(c);
            const bool need = row != kNoPos && cand_key(c) > cut;
            uint64_t outc = B2_SENTINEL;
            unsigned m = __ballot_sync(kFull, need);
            while (m) {
                const int src = __ffs(m) - 1;
                m &= m - 1;
                const uint32_t r = __shfl_sync(kFull, row, src);
                const float s = warp_exact_dot(ctx, qrow, r, lane);
                if (lane == src) {
                    const float key = (metric == B2_METRIC_IP) ? (s + 0.0f) : fmaf(2.f, s, -ctx.norms[r]);
                    if (key > -FLT_MAX) outc = pack_cand(key, r);  // NaN / -FLT_MAX never enter, as everywhere else
                }
            }
            buf[kpad + i] = outc;
        }
        block_bitonic_desc(buf, sort_n, tid, kMergeThreads);
    }
    __syncthreads();
    for (int i = tid; i < k; i += kMergeThreads) emit_out(buf[i], q, i, k, metric, 0, D, I, out_stride, qnorm, push);
    push_signal(q, push);
}

// The same merge for candidate sets that fit the CTA's registers (<= kSelMaxPerThread per thread; the usual case:
// segments x (k + slack) is a few thousand).  No sort of the approximate keys at all: the k-th best approximate key is
// found by bisection on the 32 key bits (one block-wide count per bit), the few candidates above the cut are compacted
// into shared memory, re-scored one per warp (balanced over the 16 warps), and only those are sorted.  Measured at
// 10M x 1024, 4096 queries, k = 100 (23 x 264 candidates per query): 6.9 ms -> see DESIGN.md for the chunked-sort kernel above.
constexpr int kSelMaxPerThread = kMergeSelectMax / kMergeThreads;  // 24

// MINB = CTAs per SM the compiler plans for.  1: the candidates stay in registers (102 registers, no spills) - the
// latency of a search of a few queries, where a CTA per query cannot fill the GPU anyway (36 -> 30 us at one query).
// 2: 64 registers with the candidate array partly in local memory, but two CTAs per SM - faster when thousands of
// queries are merged (2.3 vs 1.7 ms at 4096 queries, k = 100).
template <int MINB>
__global__ void __launch_bounds__(kMergeThreads, MINB) merge_rescore_select_kernel(const uint64_t* __restrict__ partial, int total,
                                                                            int k, int kpad, int sort_n, int metric,
                                                                            float* D, int64_t* I, size_t out_stride,
                                                                            const float* __restrict__ qnorm,
                                                                            const RescoreCtx ctx, const PeerPush push,
                                                                            uint32_t* __restrict__ overflow, int list_len) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* buf = reinterpret_cast<uint64_t*>(smem_raw);  // [sort_n]: running best kpad, then this round's re-scored candidates
    __shared__ int cnt_s[2][kMergeThreads / 32];
    __shared__ int n_need_s, left_s;
    const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t* in = partial + (size_t)q * total;
    const int cap = sort_n - kpad;
    uint64_t e[kSelMaxPerThread];
#pragma unroll
    for (int r = 0; r < kSelMaxPerThread; ++r) {
        const int i = tid + r * kMergeThreads;
        e[r] = (i < total) ? __ldcs(reinterpret_cast<const unsigned long long*>(in) + i) : B2_SENTINEL;
    }
    // ---- the k-th best approximate key: largest T with #(key >= T) >= k, one bit at a time (sentinels never count)
    uint32_t prefix = 0;
    for (int bit = 31; bit >= 0; --bit) {
        const uint32_t cand = prefix | (1u << bit);
        int c = 0;
#pragma unroll
        for (int r = 0; r < kSelMaxPerThread; ++r) c += ((uint32_t)(e[r] >> 32) >= cand && (uint32_t)e[r] != 0u) ? 1 : 0;
        c = __reduce_add_sync(kFull, c);
        if (lane == 0) cnt_s[bit & 1][warp] = c;
        __syncthreads();
        int tot = 0;
#pragma unroll
        for (int w = 0; w < kMergeThreads / 32; ++w) tot += cnt_s[bit & 1][w];
        if (tot >= k) prefix = cand;
    }
    // prefix == 0: fewer than k candidates, everything is re-scored
    const float cut = (prefix == 0u) ? -__builtin_huge_valf() : ord2f(prefix) - ctx.band[q];
    uint32_t pending = 0;  // bit r: entry r still has to be re-scored
    bool cut_in_last = false;  // a sorted, full list whose LAST entry is above the cut may have dropped candidates
#pragma unroll
    for (int r = 0; r < kSelMaxPerThread; ++r)
        if ((uint32_t)e[r] != 0u && cand_key(e[r]) > cut) {
            pending |= 1u << r;
            if (overflow != nullptr && (tid + r * kMergeThreads) % list_len == list_len - 1) cut_in_last = true;
        }
    if (overflow != nullptr && __syncthreads_or(cut_in_last ? 1 : 0)) {
        if (tid == 0) {
            overflow[q] = 1u;
            overflow[gridDim.x] = 1u;
        }
        return;  // no output for this query: the caller's exact fix-up launch produces it
    }
    const float* qrow = ctx.q + (size_t)q * ctx.d;
    for (int i = tid; i < kpad; i += kMergeThreads) buf[i] = B2_SENTINEL;
    // ---- rounds of at most `cap` candidates (one round unless the data is full of near-ties)
    for (;;) {
        if (tid == 0) {
            n_need_s = 0;
            left_s = 0;
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < kSelMaxPerThread; ++r) {
            if (pending & (1u << r)) {
                const int slot = atomicAdd(&n_need_s, 1);
                if (slot < cap) {
                    buf[kpad + slot] = e[r];
                    pending &= ~(1u << r);
                } else {
                    left_s = 1;
                }
            }
        }
        __syncthreads();
        const int n = min(n_need_s, cap);
        const bool more = left_s != 0;
        for (int j = warp; j < n; j += kMergeThreads / 32) {
            const uint32_t row = cand_pos(buf[kpad + j]);
            const float s = ctx.l2_direct ? warp_exact_l2(ctx, qrow, row, lane) : warp_exact_dot(ctx, qrow, row, lane);
            if (lane == 0) {
                const float key = ctx.l2_direct ? -s : ((metric == B2_METRIC_IP) ? (s + 0.0f) : fmaf(2.f, s, -ctx.norms[row]));
                buf[kpad + j] = (key > -FLT_MAX) ? pack_cand(key, row) : B2_SENTINEL;  // NaN / -FLT_MAX never enter
            }
        }
        int sort_m = 2 * kpad;
        while (sort_m < kpad + n) sort_m <<= 1;
        for (int i = kpad + n + tid; i < sort_m; i += kMergeThreads) buf[i] = B2_SENTINEL;
        block_bitonic_desc(buf, sort_m, tid, kMergeThreads);
        __syncthreads();
        if (!more) break;
    }
    for (int i = tid; i < k; i += kMergeThreads) emit_out(buf[i], q, i, k, metric, 0, D, I, out_stride, qnorm, push);
    push_signal(q, push);
}

int merge_rescore(const uint64_t* partial, int nq, int nlists, int list_len, int k, int metric, float* D, int64_t* I,
                  size_t out_stride, cudaStream_t stream, const float* qnorm, const float* q_dev, int d, const void* rows,
                  size_t pitch_bytes, int dtype, const float* norms, const float* band, const PeerPush* push_in,
                  uint32_t* overflow, bool l2_direct, const void* resid) {
    const int kpad = next_pow2(k < 32 ? 32 : k);
    const int total = nlists * list_len;
    PeerPush push;
    if (push_in != nullptr) {
        push = *push_in;
    } else {
        push.nparts = 0;
        push.epoch = 0;
        push.nq_total = push.q_offset = 0;
        for (int g = 0; g < kMaxPushPeers; ++g) {
            push.block[g] = nullptr;
            push.flag[g] = nullptr;
        }
    }
    RescoreCtx ctx;
    ctx.q = q_dev;
    ctx.rows = static_cast<const char*>(rows);
    ctx.resid = dtype == B2_DTYPE_BF16 ? static_cast<const char*>(resid) : nullptr;
    ctx.norms = norms;
    ctx.band = band;
    ctx.pitch_bytes = pitch_bytes;
    ctx.d = d;
    ctx.bf16 = dtype == B2_DTYPE_BF16;
    ctx.l2_direct = l2_direct ? 1 : 0;
    int sort_n = 2 * kpad;
    const int want = next_pow2(total + kpad);
    while (sort_n < want && sort_n < 4096) sort_n <<= 1;
    const size_t smem = (size_t)sort_n * sizeof(uint64_t);
    static const bool no_select = getenv("B2_MERGE_NO_SELECT") != nullptr;
    if (total <= kSelMaxPerThread * kMergeThreads && !no_select) {
        // a round holds up to sort_n - kpad re-scored candidates; 1024 is plenty (k + in-band entries, typically < 300)
        const int sel_sort_n = std::max(2 * kpad, 1024);
        if (nq <= 148)
            merge_rescore_select_kernel<1><<<nq, kMergeThreads, (size_t)sel_sort_n * sizeof(uint64_t), stream>>>(
                partial, total, k, kpad, sel_sort_n, metric, D, I, out_stride, qnorm, ctx, push, overflow, list_len);
        else
            merge_rescore_select_kernel<2><<<nq, kMergeThreads, (size_t)sel_sort_n * sizeof(uint64_t), stream>>>(
                partial, total, k, kpad, sel_sort_n, metric, D, I, out_stride, qnorm, ctx, push, overflow, list_len);
        B2_AFTER_LAUNCH();
        return B2_OK;
    }
    if (overflow != nullptr || l2_direct) {
        set_error("merge_rescore: %d candidates per query are too many for the register-resident merge", total);
        return B2_ERR_UNSUPPORTED;
    }
    merge_rescore_kernel<<<nq, kMergeThreads, smem, stream>>>(partial, total, k, kpad, sort_n, metric, D, I, out_stride,
                                                              qnorm, ctx, push);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

// ---------------------------------------------------------------- shard lists
// Candidate position = part * k + rank: parts are in global id order and each part's list
// is already ordered lowest-id-first among equal keys, so "lower position first" is exactly
// "lower global id first".  Part g's lists start at D_parts + g * d_stride and I_parts +
// g * i_stride (bytes); ids are local to the part when `bases` is given (global id =
// bases.v[g] + local id), global otherwise.
struct ShardBases {
    int64_t v[B2_MAX_SHARDS];
};

__global__ void __launch_bounds__(kMergeThreads) merge_shards_kernel(int metric, int k, int nparts, int nq,
                                                                    const char* __restrict__ Dp,
                                                                    const char* __restrict__ Ip, size_t d_stride, size_t i_stride,
                                                                    const ShardBases bases, int use_bases, int sort_n,
                                                                    float* D, int64_t* I,
                                                                    const uint32_t* __restrict__ flags,
                                                                    size_t flag_stride, uint32_t epoch) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* buf = reinterpret_cast<uint64_t*>(smem_raw);
    const int q = blockIdx.x, tid = threadIdx.x;
    const int total = nparts * k;
    if (flags != nullptr) {
        // peer-memory exchange: part g's lists for this query are complete once its flag shows this epoch
        if (tid < nparts) {
            const uint32_t* f = flags + (size_t)tid * flag_stride + q;
            const unsigned long long t0 = clock64();
            while (ld_acquire_sys(f) != epoch) {
                __nanosleep(64);
                if (clock64() - t0 > 40000000000ull) __trap();  // ~20 s: a peer never delivered
            }
        }
        __syncthreads();
        __threadfence_system();
    }
    for (int i = tid; i < sort_n; i += kMergeThreads) {
        uint64_t c = B2_SENTINEL;
        if (i < total) {
            const int part = i / k, r = i - part * k;
            const size_t off = (size_t)q * k + r;
            const int64_t id = __ldcg(reinterpret_cast<const long long*>(Ip + (size_t)part * i_stride) + off);
            if (id >= 0) {
                const float d = __ldcg(reinterpret_cast<const float*>(Dp + (size_t)part * d_stride) + off);
                c = pack_cand(metric == B2_METRIC_IP ? d : -d, (uint32_t)i);
            }
        }
        buf[i] = c;
    }
    block_bitonic_desc(buf, sort_n, tid, kMergeThreads);
    for (int i = tid; i < k; i += kMergeThreads) {
        const uint64_t c = buf[i];
        const uint32_t pos = cand_pos(c);
        float* Dq = D + (size_t)q * k + i;
        int64_t* Iq = I + (size_t)q * k + i;
        if (pos == kNoPos) {
            *Iq = -1;
            *Dq = (metric == B2_METRIC_IP) ? -FLT_MAX : FLT_MAX;
        } else {
            const int part = pos / k, r = pos - part * k;
            const size_t off = (size_t)q * k + r;
            const int64_t id = __ldcg(reinterpret_cast<const long long*>(Ip + (size_t)part * i_stride) + off);
            *Iq = use_bases ? id + bases.v[part] : id;
            *Dq = __ldcg(reinterpret_cast<const float*>(Dp + (size_t)part * d_stride) + off);
        }
    }
}

static int merge_shards_impl(int metric, int nq, int k, int nparts, const float* D_parts, const int64_t* I_parts,
                             float* D, int64_t* I, cudaStream_t stream, size_t d_stride_bytes, size_t i_stride_bytes,
                             const int64_t* id_bases, const uint32_t* flags, size_t flag_stride, uint32_t epoch) {
    const long long total = (long long)nparts * k;
    const int sort_n = next_pow2((int)(total < 32 ? 32 : total));
    const size_t smem = (size_t)sort_n * sizeof(uint64_t);
    if (smem > 200 * 1024 || nparts > B2_MAX_SHARDS) {
        set_error("merge_shards: nparts=%d x k=%d too large", nparts, k);
        return B2_ERR_UNSUPPORTED;
    }
    if (d_stride_bytes == 0) d_stride_bytes = (size_t)nq * k * sizeof(float);
    if (i_stride_bytes == 0) i_stride_bytes = (size_t)nq * k * sizeof(int64_t);
    ShardBases bases;
    for (int g = 0; g < B2_MAX_SHARDS; ++g) bases.v[g] = (id_bases != nullptr && g < nparts) ? id_bases[g] : 0;
    static size_t attr = 0;
    if (smem > attr) {
        B2_CUDA_TRY(cudaFuncSetAttribute(merge_shards_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr = smem;
    }
    merge_shards_kernel<<<nq, kMergeThreads, smem, stream>>>(metric, k, nparts, nq, reinterpret_cast<const char*>(D_parts),
                                                             reinterpret_cast<const char*>(I_parts), d_stride_bytes,
                                                             i_stride_bytes, bases, id_bases != nullptr, sort_n, D, I,
                                                             flags, flag_stride, epoch);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

// ---------------------------------------------------------------- query distribution over peer memory
// One CTA per destination: copies the query block into that rank's receive buffer (an NVLink peer mapping, or local
// memory for the sending rank itself), makes it visible system-wide and releases the rank's flag.
struct QueryPushDst {
    float4* buf[kMaxPushPeers];
    uint32_t* flag[kMaxPushPeers];
};
__global__ void __launch_bounds__(1024) query_push_kernel(const float4* __restrict__ q, size_t n16, const QueryPushDst dst,
                                                          uint32_t epoch) {
    float4* out = dst.buf[blockIdx.x];
    for (size_t i = threadIdx.x; i < n16; i += blockDim.x) out[i] = q[i];
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) st_release_sys(dst.flag[blockIdx.x], epoch);
}
// The receiving side: returns once the flag shows this epoch, so that the kernels queued behind it on the stream read
// the delivered block.
__global__ void flag_wait_kernel(const uint32_t* __restrict__ flag, uint32_t epoch) {
    if (threadIdx.x == 0) {
        const unsigned long long t0 = clock64();
        while (ld_acquire_sys(flag) != epoch) {
            __nanosleep(32);
            if (clock64() - t0 > 40000000000ull) __trap();  // ~20 s: the sender never delivered
        }
        __threadfence_system();
    }
}

int query_push(const float* q_dev, size_t nbytes, int nparts, void* const* bufs, uint32_t* const* flags, uint32_t epoch,
               cudaStream_t stream) {
    QueryPushDst dst;
    for (int g = 0; g < kMaxPushPeers; ++g) {
        dst.buf[g] = g < nparts ? static_cast<float4*>(bufs[g]) : nullptr;
        dst.flag[g] = g < nparts ? flags[g] : nullptr;
    }
    query_push_kernel<<<nparts, nbytes >= 16384 ? 1024 : 256, 0, stream>>>(reinterpret_cast<const float4*>(q_dev), nbytes / 16, dst, epoch);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

int flag_wait(const uint32_t* flag, uint32_t epoch, cudaStream_t stream) {
    flag_wait_kernel<<<1, 32, 0, stream>>>(flag, epoch);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

int merge_shards(int metric, int nq, int k, int nparts, const float* D_parts, const int64_t* I_parts, float* D,
                 int64_t* I, cudaStream_t stream, size_t d_stride_bytes, size_t i_stride_bytes,
                 const int64_t* id_bases) {
    return merge_shards_impl(metric, nq, k, nparts, D_parts, I_parts, D, I, stream, d_stride_bytes, i_stride_bytes,
                             id_bases, nullptr, 0, 0);
}

int merge_shards_wait(int metric, int nq, int k, int nparts, const char* blocks, size_t part_stride_bytes,
                      const uint32_t* flags, size_t flag_stride_elems, uint32_t epoch, const int64_t* id_bases, float* D,
                      int64_t* I, cudaStream_t stream) {
    const int64_t* Ip = reinterpret_cast<const int64_t*>(blocks);
    const float* Dp = reinterpret_cast<const float*>(blocks + (size_t)nq * k * sizeof(int64_t));
    return merge_shards_impl(metric, nq, k, nparts, Dp, Ip, D, I, stream, part_stride_bytes, part_stride_bytes, id_bases,
                             flags, flag_stride_elems, epoch);
}

}  // namespace b2
