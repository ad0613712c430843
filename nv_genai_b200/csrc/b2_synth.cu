// Counter-based synthetic embeddings, generated straight into device memory
// (BASELINE.md section 4): Philox-4x32-10, counter = (row_lo, row_hi, col/4, 0),
// key = (seed_lo, seed_hi); each output word w -> ((w & 0xffff) + (w >> 16) - 65535) * scale.
// Integer arithmetic plus one exact int->float convert and one fp32 multiply, so the values
// equal oracle/flat_oracle.c b2o_synth_fill bit for bit.  Bench/test support only; it stands
// in for the remote embedding service of the reference
// (/root/reference/RetrievalAugmentedGeneration/common/utils.py:292-318).
#include "b2_common.cuh"
#include "b2_kernels.h"

namespace b2 {

__device__ __forceinline__ void philox4x32_10(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
        const uint32_t n0 = hi1 ^ c[1] ^ k0, n1 = lo1, n2 = hi0 ^ c[3] ^ k1, n3 = lo0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
}

__device__ __forceinline__ float synth_word(uint32_t w) {
    const int s = (int)(w & 0xffffu) + (int)(w >> 16) - 65535;
    return (float)s * (1.0f / 26754.58f);
}

__global__ void __launch_bounds__(256) synth_kernel(float* __restrict__ out, int64_t row0, int64_t nrows, int d,
                                                    int nc4, uint32_t k0, uint32_t k1) {
    const int64_t total = nrows * (int64_t)nc4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / nc4;
        const int c4 = (int)(i - r * nc4);
        const uint64_t row = (uint64_t)(row0 + r);
        uint32_t c[4] = {(uint32_t)row, (uint32_t)(row >> 32), (uint32_t)c4, 0u};
        philox4x32_10(c, k0, k1);
        float* o = out + r * (int64_t)d + c4 * 4;
        if ((d & 3) == 0) {
            // rows are 16-byte aligned when d % 4 == 0 and the base is
            if ((reinterpret_cast<uintptr_t>(out) & 15) == 0) {
                *reinterpret_cast<float4*>(o) =
                    make_float4(synth_word(c[0]), synth_word(c[1]), synth_word(c[2]), synth_word(c[3]));
                continue;
            }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (c4 * 4 + j < d) o[j] = synth_word(c[j]);
    }
}

int synth_fill(float* out, int64_t row0, int64_t nrows, int d, uint64_t seed, cudaStream_t stream) {
    if (nrows <= 0) return B2_OK;
    const int nc4 = (d + 3) / 4;
    const int64_t total = nrows * (int64_t)nc4;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 16) blocks = 148LL * 16;
    synth_kernel<<<(int)blocks, 256, 0, stream>>>(out, row0, nrows, d, nc4, (uint32_t)seed, (uint32_t)(seed >> 32));
    B2_AFTER_LAUNCH();
    return B2_OK;
}

}  // namespace b2
