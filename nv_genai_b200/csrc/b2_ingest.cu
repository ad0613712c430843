// Row-store ingest and row movement kernels.
//
// ingest_rows replaces index.add -> IndexFlatCodes::add (a host memcpy in faiss) fused with
// faiss.normalize_L2 -> fvec_renorm_L2 (faiss/utils/distances.cpp): per row nr = sum x^2; if
// nr > 0 the row is scaled by 1/sqrtf(nr), zero rows stay zero.  Reached from langchain
// FAISS.__add (add_documents at
// /root/reference/RetrievalAugmentedGeneration/examples/nvidia_api_catalog/chains.py:61,
// from_documents at /root/reference/examples/5_mins_rag_no_gpu/main.py:90).
// One warp per row: 128-bit loads, shuffle-reduced sum of squares, scale, store as fp32 or
// bf16 into the 512-byte-pitched row store together with the squared norm of the stored row.
// Algorithmic bytes per row: 4 d read + (4 | 2) d written.
//
// gather_rows replaces index.reconstruct (IndexFlatCodes::reconstruct); copy_rows_indexed is
// the mover behind index.remove_ids (IndexFlatCodes::remove_ids compaction).
#include <algorithm>

#include "b2_common.cuh"
#include "b2_kernels.h"

namespace b2 {

constexpr int kIngestWarps = 8;
constexpr int kIngestThreads = kIngestWarps * kWarp;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) v += __shfl_xor_sync(kFull, v, s);
    return v;
}

// VEC: rows are 16-byte aligned and d % 4 == 0.  REGS: the row fits 8 float4 per lane (d <= 1024).
template <bool VEC, bool REGS, bool BF16>
__global__ void __launch_bounds__(kIngestThreads) ingest_kernel(const float* src, int64_t n, int d,
                                                               size_t src_pitch_elems, char* dst,
                                                               size_t pitch_bytes, int pitch_elems, bool normalize,
                                                               float* __restrict__ norms, char* resid = nullptr,
                                                               uint32_t* __restrict__ resid_err_ord = nullptr) {
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (int64_t)blockIdx.x * kIngestWarps + (threadIdx.x >> 5);
    const int64_t wstride = (int64_t)gridDim.x * kIngestWarps;
    for (int64_t row = warp0; row < n; row += wstride) {
        const float* s = src + row * src_pitch_elems;
        char* o = dst + row * pitch_bytes;
        // bf16 rows with a residual plane: the row is stored as hi = bf16(x) plus lo = bf16(x - hi) in a second plane
        // of the same pitch; what the index holds is then hi + lo (exact in fp32), its norm is the one recorded, and
        // sum lo^2 feeds the running maximum the error band of the filter-and-refine paths needs
        char* ores = (BF16 && resid != nullptr) ? resid + row * pitch_bytes : nullptr;
        float se = 0.f;
        float ss = 0.f;
        float4 keep[REGS ? 8 : 1];
        if (VEC) {
            const float4* s4 = reinterpret_cast<const float4*>(s);
            const int nv = d >> 2;
            if (REGS) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int c = j * 32 + lane;
                    keep[j] = (c < nv) ? __ldcs(s4 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
                    ss = fmaf(keep[j].x, keep[j].x, ss);
                    ss = fmaf(keep[j].y, keep[j].y, ss);
                    ss = fmaf(keep[j].z, keep[j].z, ss);
                    ss = fmaf(keep[j].w, keep[j].w, ss);
                }
            } else {
                for (int c = lane; c < nv; c += 32) {
                    const float4 v = s4[c];
                    ss = fmaf(v.x, v.x, ss);
                    ss = fmaf(v.y, v.y, ss);
                    ss = fmaf(v.z, v.z, ss);
                    ss = fmaf(v.w, v.w, ss);
                }
            }
        } else {
            for (int c = lane; c < d; c += 32) ss = fmaf(s[c], s[c], ss);
        }
        const float nr = warp_sum(ss);
        const float scale = (normalize && nr > 0.f) ? (1.0f / sqrtf(nr)) : 1.0f;

        float ss2 = 0.f;
        if (VEC) {
            const float4* s4 = reinterpret_cast<const float4*>(s);
            const int nv = d >> 2;
            const int nvp = pitch_elems >> 2;
            auto put = [&](int c, float4 v) {
                v.x *= scale; v.y *= scale; v.z *= scale; v.w *= scale;
                if (BF16) {
                    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
                    float2 fa = __bfloat1622float2(a), fb = __bfloat1622float2(b);
                    if (ores != nullptr) {
                        const __nv_bfloat162 la = __floats2bfloat162_rn(v.x - fa.x, v.y - fa.y),
                                             lb = __floats2bfloat162_rn(v.z - fb.x, v.w - fb.y);
                        const float2 ga = __bfloat1622float2(la), gb = __bfloat1622float2(lb);
                        se = fmaf(ga.x, ga.x, se); se = fmaf(ga.y, ga.y, se);
                        se = fmaf(gb.x, gb.x, se); se = fmaf(gb.y, gb.y, se);
                        fa.x += ga.x; fa.y += ga.y; fb.x += gb.x; fb.y += gb.y;
                        uint2 wl;
                        wl.x = *reinterpret_cast<const uint32_t*>(&la);
                        wl.y = *reinterpret_cast<const uint32_t*>(&lb);
                        reinterpret_cast<uint2*>(ores)[c] = wl;
                    }
                    ss2 = fmaf(fa.x, fa.x, ss2); ss2 = fmaf(fa.y, fa.y, ss2);
                    ss2 = fmaf(fb.x, fb.x, ss2); ss2 = fmaf(fb.y, fb.y, ss2);
                    uint2 w;
                    w.x = *reinterpret_cast<uint32_t*>(&a);
                    w.y = *reinterpret_cast<uint32_t*>(&b);
                    reinterpret_cast<uint2*>(o)[c] = w;  // (streaming stores measured slower here: 2.15 vs 1.55 ms / 1M rows)
                } else {
                    ss2 = fmaf(v.x, v.x, ss2); ss2 = fmaf(v.y, v.y, ss2);
                    ss2 = fmaf(v.z, v.z, ss2); ss2 = fmaf(v.w, v.w, ss2);
                    __stcs(reinterpret_cast<float4*>(o) + c, v);
                }
            };
            if (REGS) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const int c = j * 32 + lane;
                    if (c < nvp) put(c, keep[j]);  // keep[j] is already zero past d
                }
            } else {
                for (int c = lane; c < nvp; c += 32) put(c, (c < nv) ? s4[c] : make_float4(0.f, 0.f, 0.f, 0.f));
            }
        } else {
            for (int c = lane; c < pitch_elems; c += 32) {
                float v = (c < d) ? s[c] * scale : 0.f;
                if (BF16) {
                    const __nv_bfloat16 b = __float2bfloat16_rn(v);
                    const float hi = __bfloat162float(b);
                    reinterpret_cast<__nv_bfloat16*>(o)[c] = b;
                    if (ores != nullptr) {
                        const __nv_bfloat16 l = __float2bfloat16_rn(v - hi);
                        const float lo = __bfloat162float(l);
                        reinterpret_cast<__nv_bfloat16*>(ores)[c] = l;
                        se = fmaf(lo, lo, se);
                        v = hi + lo;
                    } else {
                        v = hi;
                    }
                } else {
                    reinterpret_cast<float*>(o)[c] = v;
                }
                ss2 = fmaf(v, v, ss2);
            }
        }
        ss2 = warp_sum(ss2);
        if (norms != nullptr && lane == 0) norms[row] = ss2;
        if (ores != nullptr && resid_err_ord != nullptr) {
            se = warp_sum(se);
            if (lane == 0 && se > 0.f) atomicMax(resid_err_ord, f2ord(se * 1.0001f));
        }
    }
}

template <bool BF16>
static int launch_ingest(const float* src, int64_t n, int d, size_t src_pitch_elems, void* dst, size_t pitch_bytes,
                         bool normalize, float* norms, cudaStream_t stream, void* resid = nullptr,
                         uint32_t* resid_err_ord = nullptr) {
    if (n <= 0) return B2_OK;
    const int pitch_elems = (int)(pitch_bytes / (BF16 ? 2 : 4));
    const bool vec = (d % 4 == 0) && (src_pitch_elems % 4 == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0) &&
                     ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) && (pitch_bytes % 16 == 0) &&
                     (pitch_elems % 4 == 0);
    const bool regs = vec && pitch_elems <= 1024;
    long long blocks = (n + kIngestWarps - 1) / kIngestWarps;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    char* o = static_cast<char*>(dst);
    char* rs = static_cast<char*>(resid);
    if (vec && regs)
        ingest_kernel<true, true, BF16><<<(int)blocks, kIngestThreads, 0, stream>>>(src, n, d, src_pitch_elems, o,
                                                                                  pitch_bytes, pitch_elems, normalize, norms, rs, resid_err_ord);
    else if (vec)
        ingest_kernel<true, false, BF16><<<(int)blocks, kIngestThreads, 0, stream>>>(src, n, d, src_pitch_elems, o,
                                                                                   pitch_bytes, pitch_elems, normalize, norms, rs, resid_err_ord);
    else
        ingest_kernel<false, false, BF16><<<(int)blocks, kIngestThreads, 0, stream>>>(src, n, d, src_pitch_elems, o,
                                                                                    pitch_bytes, pitch_elems, normalize, norms, rs, resid_err_ord);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

int ingest_rows(const float* src, int64_t n, int d, void* dst, size_t pitch_bytes, int dtype, bool normalize,
                float* norms, cudaStream_t stream, void* resid, uint32_t* resid_err_ord) {
    if (dtype == B2_DTYPE_BF16)
        return launch_ingest<true>(src, n, d, (size_t)d, dst, pitch_bytes, normalize, norms, stream, resid, resid_err_ord);
    return launch_ingest<false>(src, n, d, (size_t)d, dst, pitch_bytes, normalize, norms, stream);
}

int normalize_rows_inplace(float* x, int64_t n, int d, cudaStream_t stream) {
    // dense in place: destination pitch == source pitch == d floats, no padding columns
    if (n <= 0) return B2_OK;
    const bool vec = (d % 4 == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0);
    long long blocks = (n + kIngestWarps - 1) / kIngestWarps;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    char* o = reinterpret_cast<char*>(x);
    if (vec && d <= 1024)
        ingest_kernel<true, true, false><<<(int)blocks, kIngestThreads, 0, stream>>>(x, n, d, (size_t)d, o, (size_t)d * 4, d, true, nullptr);
    else if (vec)
        ingest_kernel<true, false, false><<<(int)blocks, kIngestThreads, 0, stream>>>(x, n, d, (size_t)d, o, (size_t)d * 4, d, true, nullptr);
    else
        ingest_kernel<false, false, false><<<(int)blocks, kIngestThreads, 0, stream>>>(x, n, d, (size_t)d, o, (size_t)d * 4, d, true, nullptr);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

// ------------------------------------------------------------------ gather (reconstruct)
__global__ void __launch_bounds__(256) gather_rows_kernel(const char* __restrict__ rows, int64_t ntotal,
                                                          size_t pitch_bytes, int d, bool bf16,
                                                          const int64_t* __restrict__ ids, int64_t n,
                                                          float* __restrict__ out, const char* __restrict__ resid) {
    const int lane = threadIdx.x & 31;
    const int64_t w = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (w >= n) return;
    const int64_t id = ids[w];
    float* o = out + w * (int64_t)d;
    if (id < 0 || id >= ntotal) {
        for (int c = lane; c < d; c += 32) o[c] = 0.f;
        return;
    }
    const char* r = rows + (size_t)id * pitch_bytes;
    if (bf16) {
        const __nv_bfloat16* rb = reinterpret_cast<const __nv_bfloat16*>(r);
        if (resid != nullptr) {  // the stored row is hi + lo
            const __nv_bfloat16* rl = reinterpret_cast<const __nv_bfloat16*>(resid + (size_t)id * pitch_bytes);
            for (int c = lane; c < d; c += 32) o[c] = __bfloat162float(rb[c]) + __bfloat162float(rl[c]);
        } else {
            for (int c = lane; c < d; c += 32) o[c] = __bfloat162float(rb[c]);
        }
    } else {
        const float* rf = reinterpret_cast<const float*>(r);
        for (int c = lane; c < d; c += 32) o[c] = rf[c];
    }
}

int gather_rows(const void* rows, int64_t ntotal, size_t pitch_bytes, int d, int dtype, const int64_t* ids, int64_t n,
                float* out, cudaStream_t stream, const void* resid) {
    if (n <= 0) return B2_OK;
    const long long blocks = (n + 7) / 8;
    gather_rows_kernel<<<(unsigned)blocks, 256, 0, stream>>>(static_cast<const char*>(rows), ntotal, pitch_bytes, d,
                                                            dtype == B2_DTYPE_BF16, ids, n, out,
                                                            dtype == B2_DTYPE_BF16 ? static_cast<const char*>(resid) : nullptr);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

// ------------------------------------------------------------------ compaction mover
// tmp row j  <-  row src_of_dst[j]  (raw 16-byte copies), and the same for norms.
__global__ void __launch_bounds__(256) copy_rows_indexed_kernel(const char* __restrict__ rows, size_t pitch_bytes,
                                                                const float* __restrict__ norms,
                                                                const int64_t* __restrict__ src_idx, int64_t n,
                                                                char* __restrict__ tmp_rows,
                                                                float* __restrict__ tmp_norms) {
    const int lane = threadIdx.x & 31;
    const int64_t w = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (w >= n) return;
    const int64_t s = src_idx[w];
    const uint4* a = reinterpret_cast<const uint4*>(rows + (size_t)s * pitch_bytes);
    uint4* b = reinterpret_cast<uint4*>(tmp_rows + (size_t)w * pitch_bytes);
    const int nv = (int)(pitch_bytes >> 4);
    for (int c = lane; c < nv; c += 32) b[c] = a[c];
    if (lane == 0 && norms != nullptr) tmp_norms[w] = norms[s];
}

// Moves rows so that dst row (first_moved + j) = old row src_of_dst[j], j < n_dst, where
// src_of_dst is increasing and src_of_dst[j] >= first_moved + j.  Chunked through a
// temporary so that no row is overwritten before it has been read.
int compact_rows(void* rows, size_t pitch_bytes, float* norms, const int64_t* src_of_dst, int64_t n_dst,
                 int64_t first_moved, cudaStream_t stream) {
    if (n_dst <= 0) return B2_OK;
    int64_t chunk = (int64_t)((64ull << 20) / pitch_bytes);
    if (chunk < 1) chunk = 1;
    if (chunk > n_dst) chunk = n_dst;
    char* tmp = nullptr;
    float* tmpn = nullptr;
    B2_CUDA_TRY(cudaMallocAsync(&tmp, (size_t)chunk * pitch_bytes, stream));
    cudaError_t e = cudaMallocAsync(&tmpn, (size_t)chunk * sizeof(float), stream);
    if (e != cudaSuccess) {
        cudaFreeAsync(tmp, stream);
        set_error("compact_rows: %s", cudaGetErrorString(e));
        return B2_ERR_OOM;
    }
    char* base = static_cast<char*>(rows);
    int rc = B2_OK;
    for (int64_t c0 = 0; c0 < n_dst && rc == B2_OK; c0 += chunk) {
        const int64_t m = (n_dst - c0 < chunk) ? (n_dst - c0) : chunk;
        copy_rows_indexed_kernel<<<(unsigned)((m + 7) / 8), 256, 0, stream>>>(base, pitch_bytes, norms,
                                                                            src_of_dst + c0, m, tmp, tmpn);
        B2_COUNT_LAUNCH();
        e = cudaGetLastError();
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(base + (size_t)(first_moved + c0) * pitch_bytes, tmp, (size_t)m * pitch_bytes,
                                cudaMemcpyDeviceToDevice, stream);
        if (e == cudaSuccess && norms != nullptr)
            e = cudaMemcpyAsync(norms + first_moved + c0, tmpn, (size_t)m * sizeof(float), cudaMemcpyDeviceToDevice,
                                stream);
        if (e != cudaSuccess) {
            set_error("compact_rows: %s", cudaGetErrorString(e));
            rc = B2_ERR_CUDA;
        }
    }
    cudaFreeAsync(tmp, stream);
    cudaFreeAsync(tmpn, stream);
    return rc;
}

// ------------------------------------------------------------------ keep list of remove_ids, built on the device
// IndexFlatCodes::remove_ids walks all rows on the host; here the ids (a few KB) are the only thing that crosses PCIe:
// a bitmap of the rows to drop, a prefix sum over the kept rows of each 32-row word (block scan + scan of the block
// totals), and the list "destination first + j takes row src[j]" that compact_rows consumes.
// stats[0] = rows dropped (distinct, in range), stats[1] = first dropped row (ntotal if none).
__global__ void mark_drop_kernel(const int64_t* __restrict__ ids, int64_t n, int64_t ntotal, uint32_t* __restrict__ drop,
                                 unsigned long long* __restrict__ stats) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t id = ids[i];
        if (id < 0 || id >= ntotal) continue;
        const uint32_t bit = 1u << (id & 31);
        if ((atomicOr(drop + (id >> 5), bit) & bit) == 0u) {  // first time this row is named
            atomicAdd(stats, 1ull);
            atomicMin(stats + 1, (unsigned long long)id);
        }
    }
}

constexpr int kKeepThreads = 1024;  // one thread per 32-row word: a block covers 32768 rows

__device__ __forceinline__ uint32_t keep_word(const uint32_t* drop, int64_t w, int64_t nwords, int64_t ntotal) {
    if (w >= nwords) return 0u;
    uint32_t keep = ~drop[w];
    const int64_t left = ntotal - w * 32;
    if (left < 32) keep &= (1u << left) - 1u;
    return keep;
}

// pass 1: kept rows per block
__global__ void __launch_bounds__(kKeepThreads) keep_count_kernel(const uint32_t* __restrict__ drop, int64_t nwords,
                                                                 int64_t ntotal, uint32_t* __restrict__ block_tot) {
    __shared__ uint32_t wsum[kKeepThreads / 32];
    const int64_t w = (int64_t)blockIdx.x * kKeepThreads + threadIdx.x;
    uint32_t c = __popc(keep_word(drop, w, nwords, ntotal));
    c = __reduce_add_sync(kFull, c);
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t v = wsum[threadIdx.x];
        v = __reduce_add_sync(kFull, v);
        if (threadIdx.x == 0) block_tot[blockIdx.x] = v;
    }
}

// pass 2: exclusive scan of the block totals, 64-bit (one block; a few hundred entries at 10M rows)
__global__ void __launch_bounds__(1024) keep_scan_blocks_kernel(const uint32_t* __restrict__ block_tot, int64_t nblocks,
                                                               unsigned long long* __restrict__ block_off) {
    __shared__ unsigned long long carry_s;
    __shared__ unsigned long long wtot[32];
    if (threadIdx.x == 0) carry_s = 0ull;
    __syncthreads();
    for (int64_t base = 0; base < nblocks; base += 1024) {
        const int64_t i = base + threadIdx.x;
        const unsigned long long v = i < nblocks ? block_tot[i] : 0ull;
        unsigned long long incl = v;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
            const unsigned long long t = __shfl_up_sync(kFull, incl, s);
            if ((threadIdx.x & 31) >= s) incl += t;
        }
        if ((threadIdx.x & 31) == 31) wtot[threadIdx.x >> 5] = incl;
        __syncthreads();
        unsigned long long woff = 0ull;
        for (int ww = 0; ww < (int)(threadIdx.x >> 5); ++ww) woff += wtot[ww];
        const unsigned long long carry = carry_s;
        if (i < nblocks) block_off[i] = carry + woff + incl - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry_s = carry + woff + incl;
        __syncthreads();
    }
}

// pass 3: src[kept rows before r - first] = r for every kept row r >= first
__global__ void __launch_bounds__(kKeepThreads) keep_emit_kernel(const uint32_t* __restrict__ drop, int64_t nwords,
                                                                int64_t ntotal,
                                                                const unsigned long long* __restrict__ block_off,
                                                                const unsigned long long* __restrict__ stats,
                                                                int64_t* __restrict__ src) {
    __shared__ uint32_t wsum[kKeepThreads / 32];
    const int64_t first = (int64_t)stats[1];
    const int64_t w = (int64_t)blockIdx.x * kKeepThreads + threadIdx.x;
    if ((int64_t)(blockIdx.x + 1) * kKeepThreads * 32 <= first) return;  // the whole block lies before the first dropped row
    uint32_t keep = keep_word(drop, w, nwords, ntotal);
    const uint32_t c = __popc(keep);
    uint32_t incl = c;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
        const uint32_t t = __shfl_up_sync(kFull, incl, s);
        if ((threadIdx.x & 31) >= s) incl += t;
    }
    if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (int ww = 0; ww < (int)(threadIdx.x >> 5); ++ww) woff += wsum[ww];
    int64_t pos = (int64_t)block_off[blockIdx.x] + woff + incl - c;  // kept rows before this word
    while (keep) {
        const int b = __ffs(keep) - 1;
        keep &= keep - 1u;
        const int64_t r = w * 32 + b;
        if (r >= first) src[pos - first] = r;
        ++pos;
    }
}

// ids_dev: n ids on the device.  Outputs (device): src list (capacity ntotal entries), stats[2]; host: *removed, *first.
int build_keep_list(const int64_t* ids_dev, int64_t n, int64_t ntotal, int64_t* src_dev, void* scratch, size_t scratch_bytes,
                    int64_t* removed, int64_t* first, cudaStream_t stream) {
    const int64_t nwords = (ntotal + 31) / 32;
    const int64_t nblocks = (nwords + kKeepThreads - 1) / kKeepThreads;
    const size_t need = keep_list_scratch_bytes(ntotal);
    if (scratch_bytes < need) {
        set_error("build_keep_list: scratch too small");
        return B2_ERR_INVALID;
    }
    char* base = static_cast<char*>(scratch);
    unsigned long long* stats = reinterpret_cast<unsigned long long*>(base);
    unsigned long long* block_off = stats + 2;
    uint32_t* block_tot = reinterpret_cast<uint32_t*>(block_off + nblocks);
    uint32_t* drop = block_tot + ((nblocks + 1) & ~(int64_t)1);
    B2_CUDA_TRY(cudaMemsetAsync(drop, 0, (size_t)nwords * 4, stream));
    const unsigned long long init[2] = {0ull, (unsigned long long)ntotal};
    B2_CUDA_TRY(cudaMemcpyAsync(stats, init, sizeof(init), cudaMemcpyHostToDevice, stream));
    mark_drop_kernel<<<(unsigned)std::min<int64_t>((n + 255) / 256, 1184), 256, 0, stream>>>(ids_dev, n, ntotal, drop, stats);
    B2_AFTER_LAUNCH();
    keep_count_kernel<<<(unsigned)nblocks, kKeepThreads, 0, stream>>>(drop, nwords, ntotal, block_tot);
    B2_AFTER_LAUNCH();
    keep_scan_blocks_kernel<<<1, 1024, 0, stream>>>(block_tot, nblocks, block_off);
    B2_AFTER_LAUNCH();
    keep_emit_kernel<<<(unsigned)nblocks, kKeepThreads, 0, stream>>>(drop, nwords, ntotal, block_off, stats, src_dev);
    B2_AFTER_LAUNCH();
    unsigned long long st[2] = {0ull, 0ull};
    B2_CUDA_TRY(cudaMemcpyAsync(st, stats, sizeof(st), cudaMemcpyDeviceToHost, stream));
    B2_CUDA_TRY(cudaStreamSynchronize(stream));
    *removed = (int64_t)st[0];
    *first = (int64_t)st[1];
    return B2_OK;
}

size_t keep_list_scratch_bytes(int64_t ntotal) {
    const int64_t nwords = (ntotal + 31) / 32;
    const int64_t nblocks = (nwords + kKeepThreads - 1) / kKeepThreads;
    return 16 + (size_t)nblocks * 8 + (size_t)((nblocks + 1) & ~(int64_t)1) * 4 + (size_t)nwords * 4 + 64;
}

// ------------------------------------------------------------------ running maximum of the squared row norms
// (ordered-uint image, atomicMax): the refine mode of the tensor-core search bounds its scoring error with it.
__global__ void max_norm_kernel(const float* __restrict__ norms, int64_t n, uint32_t* __restrict__ out) {
    float m = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const float v = norms[i];
        m = (v > m) ? v : m;  // NaN norms are ignored here: such rows never enter a result anyway
    }
#pragma unroll
    for (int s = 16; s >= 1; s >>= 1) m = fmaxf(m, __shfl_xor_sync(kFull, m, s));
    if ((threadIdx.x & 31) == 0) atomicMax(out, f2ord(m));
}

int update_max_norm(const float* norms, int64_t n, uint32_t* max_norm_ord, cudaStream_t stream) {
    if (n <= 0) return B2_OK;
    const int blocks = (int)std::min<int64_t>((n + 255) / 256, 148 * 8);
    max_norm_kernel<<<blocks, 256, 0, stream>>>(norms, n, max_norm_ord);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

// ------------------------------------------------------------------ bf16 shadow of fp32 rows
// The tensor-core search of fp32 rows filters with ONE bf16 product per score when a bf16 copy of the rows exists
// (half the operand bytes, twice the tensor rate of tf32) and re-scores the candidates near the cut from the fp32
// rows.  One warp per row: read the fp32 row once, write round-to-nearest bf16, and fold sum (x - bf16(x))^2 into a
// running maximum (ordered-uint image) -- the rigorous bound on |q . (x - x~)| the error band is built from.
__global__ void shadow_rows_kernel(const char* __restrict__ rows, size_t pitch, char* __restrict__ shadow, size_t spitch,
                                   int64_t row0, int64_t n, int d, int dpad, uint32_t* __restrict__ err_ord) {
    const int lane = threadIdx.x & 31;
    const int64_t w = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nw = (int64_t)gridDim.x * (blockDim.x >> 5);
    float emax = 0.f;
    for (int64_t r = w; r < n; r += nw) {
        const float4* src = reinterpret_cast<const float4*>(rows + (size_t)(row0 + r) * pitch);
        uint2* dst = reinterpret_cast<uint2*>(shadow + (size_t)(row0 + r) * spitch);
        float e = 0.f;
        for (int c4 = lane; c4 * 4 < dpad; c4 += 32) {  // the pitch covers dpad (zero padded by the ingest kernel)
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c4 * 4 < d) v = __ldcs(src + c4);
            if (c4 * 4 + 1 >= d) v.y = 0.f;
            if (c4 * 4 + 2 >= d) v.z = 0.f;
            if (c4 * 4 + 3 >= d) v.w = 0.f;
            const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
            const float2 fl = __bfloat1622float2(lo), fh = __bfloat1622float2(hi);
            const float dx = v.x - fl.x, dy = v.y - fl.y, dz = v.z - fh.x, dw = v.w - fh.y;
            e = fmaf(dx, dx, fmaf(dy, dy, fmaf(dz, dz, fmaf(dw, dw, e))));
            uint2 o;
            o.x = *reinterpret_cast<const uint32_t*>(&lo);
            o.y = *reinterpret_cast<const uint32_t*>(&hi);
            dst[c4] = o;
        }
#pragma unroll
        for (int s = 16; s >= 1; s >>= 1) e += __shfl_xor_sync(kFull, e, s);
        emax = (e > emax) ? e : emax;  // NaN rows: ignored, they never enter a result
    }
    if (lane == 0 && emax > 0.f) atomicMax(err_ord, f2ord(emax * 1.0001f));
}

int build_shadow_rows(const void* rows, size_t pitch, void* shadow, size_t spitch, int64_t row0, int64_t n, int d,
                      uint32_t* err_ord, cudaStream_t stream) {
    if (n <= 0) return B2_OK;
    const int dpad = (int)(spitch / 2);  // the whole shadow row, padding included, is written (zeros beyond d)
    const int blocks = (int)std::min<int64_t>((n + 7) / 8, 148 * 8);
    shadow_rows_kernel<<<blocks, 256, 0, stream>>>(static_cast<const char*>(rows), pitch, static_cast<char*>(shadow),
                                                  spitch, row0, n, d, dpad, err_ord);
    B2_AFTER_LAUNCH();
    return B2_OK;
}

}  // namespace b2
