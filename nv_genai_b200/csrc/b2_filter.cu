// Post-search selection on the device: the part of langchain's FAISS glue that runs between index.search and the
// Document list (call sites RetrievalAugmentedGeneration/examples/nvidia_api_catalog/chains.py:122,162,
// examples/multi_turn_rag/chains.py:160-165):
//
//     scores, indices = index.search(vector, k if filter is None else fetch_k)
//     skip -1; keep the hits whose Document passes the metadata filter;
//     keep score >= thr (inner product) / score <= thr (L2)              <- raw score_threshold kwarg
//     docs[:k]
//     keep relevance(score) >= thr                                      <- similarity_search_with_relevance_scores
//
// The metadata filter arrives as one bit per row (the host evaluates the predicate once per distinct filter and keeps
// the bitmask on the device), so a batch of thresholded / filtered retrievals returns k results and a count per query
// instead of fetch_k results per query and a Python loop.  One warp per query; results stay in search order.
//
// Comparisons are done in double on the widened fp32 score, which is what the reference's numpy (1.26.4,
// RetrievalAugmentedGeneration/frontend/requirements.txt:5) does for `np.float32 <= python float`.
#include <math.h>

#include "b2_common.cuh"
#include "b2_kernels.h"

namespace b2 {
namespace {

__device__ __forceinline__ double relevance_of(int kind, float score) {
    const double d = (double)score;
    switch (kind) {
        case B2_REL_EUCLIDEAN: return 1.0 - d / 1.4142135623730951;  // math.sqrt(2)
        case B2_REL_COSINE: return 1.0 - d;
        case B2_REL_MAX_INNER_PRODUCT: return d > 0.0 ? 1.0 - d : -d;
        default: return d;
    }
}

__global__ void filter_results_kernel(const float* __restrict__ D_in, const int64_t* __restrict__ I_in, int64_t nq,
                                      int fetch_k, int k, int metric, const uint32_t* __restrict__ allow,
                                      int64_t allow_bits, int use_raw, double raw_thr, int rel_kind, double rel_thr,
                                      float* __restrict__ D_out, int64_t* __restrict__ I_out,
                                      int32_t* __restrict__ counts) {
    const int lane = threadIdx.x & 31;
    const int64_t q = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (q >= nq) return;
    const float* din = D_in + q * fetch_k;
    const int64_t* iin = I_in + q * fetch_k;
    float* dout = D_out + q * k;
    int64_t* iout = I_out + q * k;
    const unsigned lt = (1u << lane) - 1u;
    int taken = 0;  // hits that passed the filter and the raw threshold so far (the list docs[:k] is cut from)
    int out = 0;    // of those, inside the first k and above the relevance threshold
    for (int base = 0; base < fetch_k && taken < k; base += 32) {
        const int i = base + lane;
        float d = 0.f;
        int64_t id = -1;
        if (i < fetch_k) {
            d = din[i];
            id = iin[i];
        }
        bool keep = id >= 0;
        if (keep && allow != nullptr) keep = id < allow_bits && ((allow[id >> 5] >> (id & 31)) & 1u) != 0u;
        if (keep && use_raw) keep = (metric == B2_METRIC_IP) ? ((double)d >= raw_thr) : ((double)d <= raw_thr);
        const unsigned b1 = __ballot_sync(0xffffffffu, keep);
        const int pos1 = taken + __popc(b1 & lt);
        bool emit = keep && pos1 < k;
        if (emit && rel_kind != B2_REL_NONE) emit = relevance_of(rel_kind, d) >= rel_thr;
        const unsigned b2 = __ballot_sync(0xffffffffu, emit);
        if (emit) {
            const int pos2 = out + __popc(b2 & lt);
            dout[pos2] = d;
            iout[pos2] = id;
        }
        taken += __popc(b1);
        out += __popc(b2);
    }
    // pad like index.search pads a short result: id -1, score = the metric's worst value
    const float pad = (metric == B2_METRIC_IP) ? -FLT_MAX : FLT_MAX;
    for (int j = out + lane; j < k; j += 32) {
        dout[j] = pad;
        iout[j] = -1;
    }
    if (lane == 0 && counts != nullptr) counts[q] = out;
}

}  // namespace

int filter_results(const float* D_in, const int64_t* I_in, int64_t nq, int fetch_k, int k, int metric,
                   const uint32_t* allow, int64_t allow_bits, int use_raw, double raw_thr, int rel_kind, double rel_thr,
                   float* D_out, int64_t* I_out, int32_t* counts, cudaStream_t stream) {
    if (nq == 0) return B2_OK;
    const int warps = 8;
    const int64_t blocks = (nq + warps - 1) / warps;
    filter_results_kernel<<<(unsigned)blocks, warps * 32, 0, stream>>>(D_in, I_in, nq, fetch_k, k, metric, allow,
                                                                      allow_bits, use_raw, raw_thr, rel_kind, rel_thr,
                                                                      D_out, I_out, counts);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

}  // namespace b2
