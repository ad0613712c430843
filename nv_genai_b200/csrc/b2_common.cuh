// Shared device/host helpers for the B200 flat index kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/b2_flat.h"

namespace b2 {

// ---------------------------------------------------------------- errors
void set_error(const char* fmt, ...);
extern std::atomic<long long> g_launches;

#define B2_CUDA_TRY(expr)                                                                          \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess) {                                                                   \
            b2::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return _e == cudaErrorMemoryAllocation ? B2_ERR_OOM : B2_ERR_CUDA;                     \
        }                                                                                          \
    } while (0)

#define B2_TRY(expr)                   \
    do {                               \
        int _s = (expr);               \
        if (_s != B2_OK) return _s;    \
    } while (0)

#define B2_COUNT_LAUNCH() (b2::g_launches.fetch_add(1, std::memory_order_relaxed))

// After a <<<>>> launch: count it and surface configuration errors.
#define B2_AFTER_LAUNCH()                 \
    do {                                  \
        B2_COUNT_LAUNCH();                \
        B2_CUDA_TRY(cudaGetLastError()); \
    } while (0)

// ---------------------------------------------------------------- constants
constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;

// ---------------------------------------------------------------- candidate encoding
// A candidate is one u64: high word = order-preserving image of the key
// (larger key is better: IP score, or minus the L2 distance), low word = ~pos,
// where pos is a row id / list position for which "lower comes first" among
// equal keys.  Larger u64 == better candidate; selection is plain integer max.
__host__ __device__ __forceinline__ uint32_t f2ord(float f) {
#ifdef __CUDA_ARCH__
    uint32_t u = __float_as_uint(f);
#else
    union { float f; uint32_t u; } c; c.f = f; uint32_t u = c.u;
#endif
    return u ^ ((u >> 31) ? 0xffffffffu : 0x80000000u);
}
__host__ __device__ __forceinline__ float ord2f(uint32_t o) {
    uint32_t u = o ^ ((o >> 31) ? 0x80000000u : 0xffffffffu);
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}
__host__ __device__ __forceinline__ uint64_t pack_cand(float key, uint32_t pos) {
    return ((uint64_t)f2ord(key) << 32) | (uint64_t)(0xffffffffu - pos);
}
__host__ __device__ __forceinline__ float cand_key(uint64_t c) { return ord2f((uint32_t)(c >> 32)); }
__host__ __device__ __forceinline__ uint32_t cand_pos(uint64_t c) { return 0xffffffffu - (uint32_t)c; }
// Empty slot: key -FLT_MAX, pos 0xffffffff.  A row is admitted only if key > -FLT_MAX
// (strictly; NaN fails), exactly faiss's heap seeded with C::neutral().
#define B2_SENTINEL ((((uint64_t)0x00800000u) << 32))  // f2ord(-FLT_MAX) = ~0xff7fffff = 0x00800000
constexpr uint32_t kNoPos = 0xffffffffu;

__host__ __device__ __forceinline__ int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

#ifdef __CUDACC__
// ---------------------------------------------------------------- loads
// 128-bit streaming load: read-only path, do not allocate in L1 (corpus rows are read once).
__device__ __forceinline__ float4 ld_stream_f4(const void* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ uint4 ld_stream_u4(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ float bf16lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }

__device__ __forceinline__ uint64_t shfl_xor_u64(uint64_t v, int m) {
    uint32_t lo = __shfl_xor_sync(kFull, (uint32_t)v, m);
    uint32_t hi = __shfl_xor_sync(kFull, (uint32_t)(v >> 32), m);
    return ((uint64_t)hi << 32) | lo;
}
__device__ __forceinline__ uint64_t shfl_u64(uint64_t v, int src) {
    uint32_t lo = __shfl_sync(kFull, (uint32_t)v, src);
    uint32_t hi = __shfl_sync(kFull, (uint32_t)(v >> 32), src);
    return ((uint64_t)hi << 32) | lo;
}

// Replace the worst entry of a warp-private list by c (if c still beats it) and recompute
// the worst.  Warp-uniform; `c` must be identical in all lanes.
__device__ __forceinline__ void warp_list_insert(uint64_t* L, int k, uint64_t c, volatile uint64_t* worst,
                                                 volatile int* worstpos, int lane) {
    if (c <= *worst) return;
    if (lane == 0) L[*worstpos] = c;
    __syncwarp();
    uint64_t mn = ~0ull;
    int mp = 0x7fffffff;
    for (int i = lane; i < k; i += kWarp) {
        const uint64_t e = L[i];
        if (e < mn) {
            mn = e;
            mp = i;
        }
    }
    // warp minimum of a 64-bit value with two 32-bit redux operations (high word, then low word among the lanes that
    // hold the minimal high word); entries are unique except for sentinels, where any position will do
    const uint32_t hi = (uint32_t)(mn >> 32), lo = (uint32_t)mn;
    const uint32_t hmin = __reduce_min_sync(kFull, hi);
    const uint32_t lmin = __reduce_min_sync(kFull, hi == hmin ? lo : 0xffffffffu);
    const int src = __ffs(__ballot_sync(kFull, hi == hmin && lo == lmin)) - 1;
    mp = __shfl_sync(kFull, mp, src);
    mn = ((uint64_t)hmin << 32) | lmin;
    if (lane == 0) {
        *worst = mn;
        *worstpos = mp;
    }
    __syncwarp();
}

// ---------------------------------------------------------------- block bitonic sort (descending)
// Sorts buf[0..n) (n a power of two, in shared memory) so that buf[0] is the
// best candidate.  All `nthreads` threads of the CTA must call it.
__device__ __forceinline__ void block_bitonic_desc(uint64_t* buf, int n, int tid, int nthreads) {
    for (int size = 2; size <= n; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            for (int t = tid; t < (n >> 1); t += nthreads) {
                int lo = 2 * t - (t & (stride - 1));
                int hi = lo + stride;
                bool desc = ((lo & size) == 0);
                uint64_t a = buf[lo], b = buf[hi];
                if ((a < b) == desc) {
                    buf[lo] = b;
                    buf[hi] = a;
                }
            }
        }
    }
    __syncthreads();
}
#endif  // __CUDACC__

}  // namespace b2
