// Maximal-marginal-relevance re-ranking on the device.
//
// Replaces langchain_community.vectorstores.utils.maximal_marginal_relevance +
// langchain_community.utils.math.cosine_similarity as driven by
// FAISS.max_marginal_relevance_search_with_score_by_vector (the
// same cosine_similarity is imported at
// /root/reference/experimental/fm-asr-streaming-rag/chain-server/common.py:25):
//   sim_q[i] = cos(q, e_i);  first pick = argmax sim_q (first index on ties)
//   repeat: pick argmax over unpicked i of  lambda*sim_q[i] - (1-lambda)*max_{j picked} cos(e_i, e_j)
//   with strict '>' (first index on ties); NaN/Inf cosines count as 0.
// One CTA per query; a warp per candidate row for the dot products.  Instead of the full
// fetch_k x fetch_k Gram matrix only the column of the newest pick is computed per round and
// folded into a running max, which is the same arithmetic the reference evaluates.
#include "b2_common.cuh"
#include "b2_kernels.h"

namespace b2 {

constexpr int kMmrThreads = 256;

__device__ __forceinline__ float warp_dot(const float* a, const float* b, int d, int lane) {
    float s = 0.f;
    if ((d & 3) == 0 && ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0) {
        const float4* a4 = reinterpret_cast<const float4*>(a);
        const float4* b4 = reinterpret_cast<const float4*>(b);
        for (int c = lane; c < (d >> 2); c += 32) {
            const float4 x = a4[c], y = b4[c];
            s = fmaf(x.x, y.x, s);
            s = fmaf(x.y, y.y, s);
            s = fmaf(x.z, y.z, s);
            s = fmaf(x.w, y.w, s);
        }
    } else {
        for (int c = lane; c < d; c += 32) s = fmaf(a[c], b[c], s);
    }
#pragma unroll
    for (int m = 16; m >= 1; m >>= 1) s += __shfl_xor_sync(kFull, s, m);
    return s;
}

__device__ __forceinline__ float clean_cos(float dot, float na, float nb) {
    const float v = dot / (na * nb);
    return (isnan(v) || isinf(v)) ? 0.f : v;
}

__global__ void __launch_bounds__(kMmrThreads) mmr_kernel(const float* __restrict__ q, int d,
                                                          const float* __restrict__ rows,
                                                          const int64_t* __restrict__ ids,
                                                          const float* __restrict__ scores, int fetch_k, int k,
                                                          float lambda_mult, int64_t* __restrict__ out_ids,
                                                          float* __restrict__ out_scores) {
    extern __shared__ float sm[];
    float* simq = sm;                 // [fetch_k]
    float* red = simq + fetch_k;      // [fetch_k] running max cos to the picked set
    float* nrm = red + fetch_k;       // [fetch_k]
    int* picked = reinterpret_cast<int*>(nrm + fetch_k);  // [fetch_k]
    __shared__ int s_nvalid, s_last;
    __shared__ float s_qn;

    const int qi = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float* qv = q + (size_t)qi * d;
    const float* E = rows + (size_t)qi * fetch_k * d;
    const int64_t* id = ids + (size_t)qi * fetch_k;

    if (tid == 0) {
        int nv = 0;
        for (int i = 0; i < fetch_k; ++i)
            if (id[i] >= 0) nv = i + 1;  // hits are a prefix; -1 only pads the tail
        s_nvalid = nv;
    }
    for (int i = tid; i < fetch_k; i += kMmrThreads) {
        picked[i] = 0;
        red[i] = -INFINITY;
    }
    if (warp == 0) {
        const float qq = warp_dot(qv, qv, d, lane);
        if (lane == 0) s_qn = sqrtf(qq);
    }
    __syncthreads();
    const int nvalid = s_nvalid;
    const int npick = min(k, nvalid);
    for (int i = warp; i < nvalid; i += kMmrThreads / 32) {
        const float* e = E + (size_t)i * d;
        const float ee = warp_dot(e, e, d, lane);
        const float qe = warp_dot(qv, e, d, lane);
        if (lane == 0) {
            const float n = sqrtf(ee);
            nrm[i] = n;
            simq[i] = clean_cos(qe, s_qn, n);
        }
    }
    __syncthreads();

    for (int round = 0; round < npick; ++round) {
        if (tid == 0) {
            int best = -1;
            float bs = -INFINITY;
            if (round == 0) {
                best = 0;
                bs = simq[0];
                for (int i = 1; i < nvalid; ++i)
                    if (simq[i] > bs) {  // np.argmax: first maximum
                        bs = simq[i];
                        best = i;
                    }
            } else {
                for (int i = 0; i < nvalid; ++i) {
                    if (picked[i]) continue;
                    const float sc = lambda_mult * simq[i] - (1.0f - lambda_mult) * red[i];
                    if (sc > bs) {
                        bs = sc;
                        best = i;
                    }
                }
                if (best < 0) best = nvalid - 1;  // reference quirk: idx_to_add stays -1 -> last element
            }
            picked[best] = 1;
            s_last = best;
            out_ids[(size_t)qi * k + round] = id[best];
            out_scores[(size_t)qi * k + round] = scores[(size_t)qi * fetch_k + best];
        }
        __syncthreads();
        const int last = s_last;
        if (round + 1 < npick) {
            const float* es = E + (size_t)last * d;
            for (int i = warp; i < nvalid; i += kMmrThreads / 32) {
                const float dt = warp_dot(E + (size_t)i * d, es, d, lane);
                if (lane == 0) red[i] = fmaxf(red[i], clean_cos(dt, nrm[i], nrm[last]));
            }
        }
        __syncthreads();
    }
    for (int r = npick + tid; r < k; r += kMmrThreads) {
        out_ids[(size_t)qi * k + r] = -1;
        out_scores[(size_t)qi * k + r] = 0.f;
    }
}

int mmr_select(const float* q, int d, int nq, const float* cand_rows, const int64_t* cand_ids,
               const float* cand_scores, int fetch_k, int k, float lambda_mult, int64_t* out_ids, float* out_scores,
               cudaStream_t stream) {
    if (nq <= 0) return B2_OK;
    const size_t smem = (size_t)fetch_k * 16;
    if (smem > 96 * 1024) {
        set_error("mmr: fetch_k=%d too large", fetch_k);
        return B2_ERR_UNSUPPORTED;
    }
    if (smem > 48 * 1024)
        B2_CUDA_TRY(cudaFuncSetAttribute(mmr_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    mmr_kernel<<<nq, kMmrThreads, smem, stream>>>(q, d, cand_rows, cand_ids, cand_scores, fetch_k, k, lambda_mult,
                                                  out_ids, out_scores);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

}  // namespace b2
