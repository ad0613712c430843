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
#include "b2_common.cuh"
#include "b2_kernels.h"

namespace b2 {

constexpr int kMergeThreads = 512;

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

__global__ void __launch_bounds__(kMergeThreads) merge_partials_kernel(const uint64_t* __restrict__ partial,
                                                                      int total, int k, int kpad, int sort_n,
                                                                      int metric, int64_t id_base, float* D,
                                                                      int64_t* I, size_t out_stride,
                                                                      const float* __restrict__ qnorm) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* buf = reinterpret_cast<uint64_t*>(smem_raw);
    const int q = blockIdx.x, tid = threadIdx.x;
    const uint64_t* in = partial + (size_t)q * total;
    const int cap = sort_n - kpad;
    for (int i = tid; i < kpad; i += kMergeThreads) buf[i] = B2_SENTINEL;
    for (int off = 0; off < total; off += cap) {
        for (int i = tid; i < cap; i += kMergeThreads) buf[kpad + i] = (off + i < total) ? in[off + i] : B2_SENTINEL;
        block_bitonic_desc(buf, sort_n, tid, kMergeThreads);
    }
    __syncthreads();
    for (int i = tid; i < k; i += kMergeThreads)
        emit_result(buf[i], metric, id_base, D + (size_t)q * out_stride + i, I + (size_t)q * out_stride + i,
                    qnorm ? qnorm + q : nullptr);
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
                                                                            const float* __restrict__ qnorm) {
    __shared__ uint64_t lists[kMergeThreads / 32][32];
    const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
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
        if (lane < k)
            emit_result(best, metric, id_base, D + (size_t)q * out_stride + lane, I + (size_t)q * out_stride + lane,
                        qnorm ? qnorm + q : nullptr);
    }
}

int merge_partials(const uint64_t* partial, int nq, int nlists, int k, int metric, int64_t id_base, float* D,
                   int64_t* I, size_t out_stride, cudaStream_t stream, const float* qnorm) {
    const int kpad = next_pow2(k < 32 ? 32 : k);
    const int total = nlists * k;
    if (kpad == 32) {
        merge_partials_small_kernel<<<nq, kMergeThreads, 0, stream>>>(partial, total, k, metric, id_base, D, I,
                                                                      out_stride, qnorm);
        B2_AFTER_LAUNCH();
        return B2_OK;
    }
    int sort_n = 2 * kpad;
    const int want = next_pow2(total + kpad);
    while (sort_n < want && sort_n < 4096) sort_n <<= 1;
    const size_t smem = (size_t)sort_n * sizeof(uint64_t);
    merge_partials_kernel<<<nq, kMergeThreads, smem, stream>>>(partial, total, k, kpad, sort_n, metric, id_base, D, I,
                                                               out_stride, qnorm);
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
                                                                    float* D, int64_t* I) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* buf = reinterpret_cast<uint64_t*>(smem_raw);
    const int q = blockIdx.x, tid = threadIdx.x;
    const int total = nparts * k;
    for (int i = tid; i < sort_n; i += kMergeThreads) {
        uint64_t c = B2_SENTINEL;
        if (i < total) {
            const int part = i / k, r = i - part * k;
            const size_t off = (size_t)q * k + r;
            const int64_t id = reinterpret_cast<const int64_t*>(Ip + (size_t)part * i_stride)[off];
            if (id >= 0) {
                const float d = reinterpret_cast<const float*>(Dp + (size_t)part * d_stride)[off];
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
            const int64_t id = reinterpret_cast<const int64_t*>(Ip + (size_t)part * i_stride)[off];
            *Iq = use_bases ? id + bases.v[part] : id;
            *Dq = reinterpret_cast<const float*>(Dp + (size_t)part * d_stride)[off];
        }
    }
}

int merge_shards(int metric, int nq, int k, int nparts, const float* D_parts, const int64_t* I_parts, float* D,
                 int64_t* I, cudaStream_t stream, size_t d_stride_bytes, size_t i_stride_bytes,
                 const int64_t* id_bases) {
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
                                                             i_stride_bytes, bases, id_bases != nullptr, sort_n, D, I);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

}  // namespace b2
