// Small-batch exact scan (nq <= 16 per pass): the HBM-bound regime of IndexFlat::search.
//
// Replaces faiss exhaustive_inner_product_seq / exhaustive_L2sqr_seq + HeapResultHandler
// (faiss/utils/distances.cpp, the nq < 20 path reached from index.search at
// /root/reference/examples/5_mins_rag_no_gpu/main.py:126 and every retriever call of
// /root/reference/RetrievalAugmentedGeneration/examples/nvidia_api_catalog/chains.py:122-123).
//
// Layout: the row store is row-major with a row pitch that is a multiple of 512 B, so one
// warp reads one row as `nchunks` fully coalesced 512-byte requests (lane l takes bytes
// [16 l, 16 l + 16) of each chunk) with 128-bit streaming loads.  Each warp owns groups of R
// consecutive rows, interleaved across the whole grid, keeps R x NQ fp32 accumulators in
// registers, reads the query tile from shared memory, reduces across lanes with a
// transposing butterfly and feeds a per-(warp, query) top-k list in shared memory that is
// guarded by its current worst entry, so the common case is one compare per score.  The CTA
// merges its warps' lists at the end; the score matrix never exists in HBM.
//
// Algorithmic HBM bytes per launch = n * pitch_bytes (each row read exactly once).
#include "b2_common.cuh"
#include "b2_kernels.h"

#include <mutex>
#include <vector>

namespace b2 {

// ---- optional per-launch timing of the dominant kernels (bench.py's roofline figures)
// kind 0 = scan_topk_kernel, kind 1 = gemm_topk_kernel.  Events are recorded on the launching stream.
static std::atomic<int> g_prof_on{0};
static std::mutex g_prof_mu;
struct ProfEvents { cudaEvent_t a, b; int kind; };
static std::vector<ProfEvents> g_prof_events;
static double g_prof_ms[2] = {0.0, 0.0};
static long long g_prof_n[2] = {0, 0};

void prof_enable(int on) { g_prof_on.store(on ? 1 : 0); }

void* prof_begin(cudaStream_t stream) {
    if (g_prof_on.load() == 0) return nullptr;
    cudaEvent_t* ev = new cudaEvent_t[2];
    if (cudaEventCreate(&ev[0]) != cudaSuccess || cudaEventCreate(&ev[1]) != cudaSuccess) {
        cudaGetLastError();
        delete[] ev;
        return nullptr;
    }
    cudaEventRecord(ev[0], stream);
    return ev;
}

void prof_end(void* token, int kind, cudaStream_t stream) {
    if (token == nullptr) return;
    cudaEvent_t* ev = static_cast<cudaEvent_t*>(token);
    cudaEventRecord(ev[1], stream);
    {
        std::lock_guard<std::mutex> lk(g_prof_mu);
        g_prof_events.push_back({ev[0], ev[1], kind});
    }
    delete[] ev;
}

int prof_read(int kind, double* ms, int64_t* n, int reset) {
    if (kind < 0 || kind > 1) kind = 0;
    std::lock_guard<std::mutex> lk(g_prof_mu);
    for (auto& ev : g_prof_events) {
        float t = 0.f;
        if (cudaEventSynchronize(ev.b) == cudaSuccess && cudaEventElapsedTime(&t, ev.a, ev.b) == cudaSuccess) {
            g_prof_ms[ev.kind] += t;
            ++g_prof_n[ev.kind];
        }
        cudaEventDestroy(ev.a);
        cudaEventDestroy(ev.b);
    }
    cudaGetLastError();
    g_prof_events.clear();
    if (ms) *ms = g_prof_ms[kind];
    if (n) *n = g_prof_n[kind];
    if (reset) {
        g_prof_ms[kind] = 0.0;
        g_prof_n[kind] = 0;
    }
    return B2_OK;
}

constexpr int kScanWarps = 8;
constexpr int kScanThreads = kScanWarps * kWarp;

struct ScanParams {
    const char* rows;      // row store base
    uint32_t n;            // rows
    uint32_t pitch_bytes;  // row pitch (multiple of 512)
    int nchunks;           // pitch_bytes / 512
    const float* q;        // [nq_valid][d] fp32, device
    int d;
    int nq_valid;          // <= NQ
    int k, kpad;
    uint64_t* partial;     // [NQ][gridDim.x][k] sorted best-first
    const uint32_t* run_if;  // not null: the launch does nothing unless *run_if != 0 (fix-up pass of the tensor-core scan)
    long long resid_delta;   // bf16 rows with a residual plane: byte offset from a row to its residual row (0 = none)
};

// Transposing butterfly: V per-lane partial sums -> max(V/32,1) complete sums per lane.
// On return v[0..VL) are totals of flat indices `base + i`; for V < 32 lanes that share the
// same upper lane bits hold the same totals.
template <int V>
__device__ __forceinline__ int warp_transpose_reduce(float (&v)[V], int lane) {
    int base = 0;
    int n = V;
#pragma unroll
    for (int step = 16; step >= 1; step >>= 1) {
        if (n > 1) {
            const int h = n >> 1;
            const bool upper = (lane & step) != 0;
#pragma unroll
            for (int i = 0; i < V / 2; ++i) {
                if (i < h) {
                    const float send = upper ? v[i] : v[i + h];
                    const float keep = upper ? v[i + h] : v[i];
                    v[i] = keep + __shfl_xor_sync(kFull, send, step);
                }
            }
            base += upper ? h : 0;
            n = h;
        } else {
            v[0] += __shfl_xor_sync(kFull, v[0], step);
        }
    }
    return base;
}

template <int METRIC, bool BF16, int NQ, int R>
__global__ void __launch_bounds__(kScanThreads) scan_topk_kernel(const ScanParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    if (p.run_if != nullptr && *p.run_if == 0u) return;
    constexpr int CH = BF16 ? 256 : 128;  // elements per 512-byte chunk
    constexpr int V = R * NQ;
    constexpr int VL = (V >= 32) ? V / 32 : 1;

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const int nchunks = p.nchunks;
    const int qpitch = nchunks * CH;

    float* q_s = reinterpret_cast<float*>(smem_raw);                                   // [NQ][qpitch]
    uint64_t* lists = reinterpret_cast<uint64_t*>(q_s + (size_t)NQ * qpitch);          // [NQ][warps][kpad]
    uint64_t* worst_s = lists + (size_t)NQ * kScanWarps * p.kpad;                      // [NQ][warps]
    int* worstpos_s = reinterpret_cast<int*>(worst_s + NQ * kScanWarps);               // [NQ][warps]

    // ---- stage the query tile (zero padded; bf16 rows use a lane-interleaved layout so that
    //      both 128-bit query reads of a lane are bank-conflict free)
    for (int i = tid; i < NQ * qpitch; i += kScanThreads) {
        const int qi = i / qpitch, c = i - qi * qpitch;
        float v = 0.f;
        if (qi < p.nq_valid && c < p.d) v = p.q[(size_t)qi * p.d + c];
        int dst;
        if (BF16) {
            const int j = c >> 8, l = (c & 255) >> 3, e = c & 7;
            dst = ((((qi * nchunks + j) * 2 + (e >> 2)) * 32 + l) << 2) + (e & 3);
        } else {
            dst = qi * qpitch + c;
        }
        q_s[dst] = v;
    }
    for (int i = tid; i < NQ * kScanWarps * p.kpad; i += kScanThreads) lists[i] = B2_SENTINEL;
    for (int i = tid; i < NQ * kScanWarps; i += kScanThreads) {
        worst_s[i] = B2_SENTINEL;
        worstpos_s[i] = 0;
    }
    __syncthreads();

    const float4* q4 = reinterpret_cast<const float4*>(q_s);
    const uint32_t n = p.n;
    const uint32_t ngroups = (n + R - 1) / R;
    const uint32_t W = gridDim.x * kScanWarps;

    for (uint32_t g = blockIdx.x * kScanWarps + warp; g < ngroups; g += W) {
        const uint32_t row0 = g * R;
        const char* rp[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            uint32_t row = row0 + r;
            row = row < n ? row : n - 1;
            rp[r] = p.rows + (size_t)row * p.pitch_bytes + lane * 16;
        }
        float acc[V];
#pragma unroll
        for (int i = 0; i < V; ++i) acc[i] = 0.f;

        // Many queries per pass make the loop FMA-heavy (NQ x R x 4 FFMA per 128-bit load) with few warps
        // resident, so the next chunk's loads are issued before this chunk's arithmetic (register double
        // buffer); with few queries the loop is load-bound and plain unrolling keeps more loads in flight.
        constexpr bool PREFETCH = (!BF16) && (NQ >= 8);
        float4 xnext[PREFETCH ? R : 1];
        if (PREFETCH) {
#pragma unroll
            for (int r = 0; r < R; ++r) xnext[PREFETCH ? r : 0] = ld_stream_f4(rp[r]);
        }
#pragma unroll 2
        for (int j = 0; j < nchunks; ++j) {
            if (!BF16) {
                float4 x[R];
                if (PREFETCH) {
#pragma unroll
                    for (int r = 0; r < R; ++r) x[r] = xnext[PREFETCH ? r : 0];
                    if (j + 1 < nchunks) {
#pragma unroll
                        for (int r = 0; r < R; ++r) xnext[PREFETCH ? r : 0] = ld_stream_f4(rp[r] + (size_t)(j + 1) * 512);
                    }
                } else {
#pragma unroll
                    for (int r = 0; r < R; ++r) x[r] = ld_stream_f4(rp[r] + (size_t)j * 512);
                }
#pragma unroll
                for (int qi = 0; qi < NQ; ++qi) {
                    const float4 qv = q4[(qi * nchunks + j) * 32 + lane];
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        float a = acc[qi * R + r];
                        if (METRIC == B2_METRIC_IP) {
                            a = fmaf(x[r].x, qv.x, a);
                            a = fmaf(x[r].y, qv.y, a);
                            a = fmaf(x[r].z, qv.z, a);
                            a = fmaf(x[r].w, qv.w, a);
                        } else {
                            float t;
                            t = x[r].x - qv.x; a = fmaf(t, t, a);
                            t = x[r].y - qv.y; a = fmaf(t, t, a);
                            t = x[r].z - qv.z; a = fmaf(t, t, a);
                            t = x[r].w - qv.w; a = fmaf(t, t, a);
                        }
                        acc[qi * R + r] = a;
                    }
                }
            } else {
                uint4 x[R];
#pragma unroll
                for (int r = 0; r < R; ++r) x[r] = ld_stream_u4(rp[r] + (size_t)j * 512);
                // residual plane (warp-uniform): the stored row is hi + lo, exact in fp32
                const bool two = p.resid_delta != 0;
                uint4 y[R];
#pragma unroll
                for (int r = 0; r < R; ++r)
                    y[r] = two ? ld_stream_u4(rp[r] + p.resid_delta + (size_t)j * 512) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
                for (int qi = 0; qi < NQ; ++qi) {
                    const float4 qa = q4[((qi * nchunks + j) * 2 + 0) * 32 + lane];
                    const float4 qb = q4[((qi * nchunks + j) * 2 + 1) * 32 + lane];
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        float a = acc[qi * R + r];
                        float e0 = bf16lo(x[r].x), e1 = bf16hi(x[r].x), e2 = bf16lo(x[r].y), e3 = bf16hi(x[r].y);
                        float e4 = bf16lo(x[r].z), e5 = bf16hi(x[r].z), e6 = bf16lo(x[r].w), e7 = bf16hi(x[r].w);
                        if (two) {
                            e0 += bf16lo(y[r].x); e1 += bf16hi(y[r].x); e2 += bf16lo(y[r].y); e3 += bf16hi(y[r].y);
                            e4 += bf16lo(y[r].z); e5 += bf16hi(y[r].z); e6 += bf16lo(y[r].w); e7 += bf16hi(y[r].w);
                        }
                        if (METRIC == B2_METRIC_IP) {
                            a = fmaf(e0, qa.x, a); a = fmaf(e1, qa.y, a); a = fmaf(e2, qa.z, a); a = fmaf(e3, qa.w, a);
                            a = fmaf(e4, qb.x, a); a = fmaf(e5, qb.y, a); a = fmaf(e6, qb.z, a); a = fmaf(e7, qb.w, a);
                        } else {
                            float t;
                            t = e0 - qa.x; a = fmaf(t, t, a); t = e1 - qa.y; a = fmaf(t, t, a);
                            t = e2 - qa.z; a = fmaf(t, t, a); t = e3 - qa.w; a = fmaf(t, t, a);
                            t = e4 - qb.x; a = fmaf(t, t, a); t = e5 - qb.y; a = fmaf(t, t, a);
                            t = e6 - qb.z; a = fmaf(t, t, a); t = e7 - qb.w; a = fmaf(t, t, a);
                        }
                        acc[qi * R + r] = a;
                    }
                }
            }
        }

        // ---- lane reduction: afterwards this lane owns totals of flat indices fbase .. fbase+VL-1
        const int fbase = warp_transpose_reduce<V>(acc, lane);
        const bool owner = (V >= 32) || ((lane & ((32 / (V < 32 ? V : 32)) - 1)) == 0);

#pragma unroll
        for (int i = 0; i < VL; ++i) {
            const int f = fbase + i;
            const int qi = f / R, r = f - qi * R;
            const uint32_t row = row0 + r;
            const float key = (METRIC == B2_METRIC_IP) ? (acc[i] + 0.0f) : -acc[i];
            const uint64_t cand = pack_cand(key, row);
            const int li = qi * kScanWarps + warp;
            bool pass = owner && row < n && key > -FLT_MAX && cand > worst_s[li];
            unsigned m = __ballot_sync(kFull, pass);
            while (m) {
                const int src = __ffs(m) - 1;
                m &= m - 1;
                const uint64_t c = shfl_u64(cand, src);
                const int cl = __shfl_sync(kFull, li, src);
                warp_list_insert(lists + (size_t)cl * p.kpad, p.k, c, worst_s + cl, worstpos_s + cl, lane);
            }
        }
    }

    // ---- CTA merge: per query, sort the warps' lists (contiguous) and emit the best k
    __syncthreads();
    const int region = kScanWarps * p.kpad;
    for (int qi = 0; qi < p.nq_valid; ++qi) {
        uint64_t* buf = lists + (size_t)qi * region;
        block_bitonic_desc(buf, region, tid, kScanThreads);
        uint64_t* out = p.partial + ((size_t)qi * gridDim.x + blockIdx.x) * p.k;
        for (int i = tid; i < p.k; i += kScanThreads) out[i] = buf[i];
    }
}

// ------------------------------------------------------------------ launcher
template <int METRIC, bool BF16, int NQ, int R>
static int launch_scan(const ScanParams& p, int num_sms, size_t smem, cudaStream_t stream, int* grid_out) {
    auto kern = scan_topk_kernel<METRIC, BF16, NQ, R>;
    // attribute + occupancy are cached per (instantiation, device, smem): they cost several microseconds
    static std::mutex mu;
    static int c_dev[16];
    static size_t c_smem[16];
    static int c_occ[16];
    static int c_n = 0;
    static size_t c_attr[16] = {0};  // largest dynamic-smem attribute set so far, per device ordinal
    int dev = 0, occ = -1;
    B2_CUDA_TRY(cudaGetDevice(&dev));
    {
        std::lock_guard<std::mutex> lk(mu);
        for (int i = 0; i < c_n; ++i)
            if (c_dev[i] == dev && c_smem[i] == smem) occ = c_occ[i];
    }
    if (occ < 0) {
        std::lock_guard<std::mutex> lk(mu);
        if (smem > c_attr[dev & 15]) {  // the attribute is a maximum: only ever raise it
            B2_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            c_attr[dev & 15] = smem;
        }
        B2_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kScanThreads, smem));
        if (c_n < 16) {
            c_dev[c_n] = dev;
            c_smem[c_n] = smem;
            c_occ[c_n] = occ;
            ++c_n;
        }
    }
    if (occ < 1) {
        set_error("scan kernel does not fit on an SM (smem=%zu)", smem);
        return B2_ERR_UNSUPPORTED;
    }
    const uint32_t ngroups = (p.n + R - 1) / R;
    long long grid = (long long)num_sms * occ;
    const long long need = (ngroups + kScanWarps - 1) / kScanWarps;
    if (grid > need) grid = need;
    if (grid < 1) grid = 1;
    *grid_out = (int)grid;
    return B2_OK;
}

size_t scan_smem_bytes(int nq_tile, int nchunks, bool bf16, int kpad) {
    const size_t qpitch = (size_t)nchunks * (bf16 ? 256 : 128);
    return (size_t)nq_tile * qpitch * 4 + (size_t)nq_tile * kScanWarps * kpad * 8 + (size_t)nq_tile * kScanWarps * 12 + 16;
}

template <int METRIC, bool BF16, int NQ, int R>
static int run_scan(ScanParams p, int num_sms, cudaStream_t stream, uint64_t** partial, int* grid_out) {
    const size_t smem = scan_smem_bytes(NQ, p.nchunks, BF16, p.kpad);
    int grid = 0;
    B2_TRY((launch_scan<METRIC, BF16, NQ, R>(p, num_sms, smem, stream, &grid)));
    uint64_t* buf = nullptr;
    B2_CUDA_TRY(cudaMallocAsync(&buf, (size_t)NQ * grid * p.k * sizeof(uint64_t), stream));
    p.partial = buf;
    void* tok = prof_begin(stream);
    scan_topk_kernel<METRIC, BF16, NQ, R><<<grid, kScanThreads, smem, stream>>>(p);
    B2_COUNT_LAUNCH();
    cudaError_t e = cudaGetLastError();
    prof_end(tok, 0, stream);
    if (e != cudaSuccess) {
        cudaFreeAsync(buf, stream);
        set_error("scan kernel launch failed: %s", cudaGetErrorString(e));
        return B2_ERR_CUDA;
    }
    *partial = buf;
    *grid_out = grid;
    return B2_OK;
}

template <int METRIC, bool BF16>
static int dispatch_nq(int nq_tile, const ScanParams& p, int num_sms, cudaStream_t s, uint64_t** partial, int* grid) {
    switch (nq_tile) {
        case 1: return run_scan<METRIC, BF16, 1, 4>(p, num_sms, s, partial, grid);
        case 2: return run_scan<METRIC, BF16, 2, 4>(p, num_sms, s, partial, grid);
        case 4: return run_scan<METRIC, BF16, 4, 4>(p, num_sms, s, partial, grid);
        case 8: return run_scan<METRIC, BF16, 8, 4>(p, num_sms, s, partial, grid);
        case 16: return run_scan<METRIC, BF16, 16, 4>(p, num_sms, s, partial, grid);
    }
    set_error("internal: bad query tile %d", nq_tile);
    return B2_ERR_INVALID;
}

// One pass over the rows for a tile of <= 16 queries.  Leaves [nq_tile][grid][k] sorted
// partial lists in *partial (stream-ordered allocation; the caller frees it with cudaFreeAsync).
int scan_pass(const void* rows, int64_t n, size_t pitch_bytes, int d, int dtype, int metric, const float* q_dev,
              int nq_valid, int nq_tile, int k, int num_sms, cudaStream_t stream, uint64_t** partial, int* grid,
              const uint32_t* run_if, const void* resid) {
    ScanParams p;
    p.run_if = run_if;
    p.resid_delta = (resid != nullptr && dtype == B2_DTYPE_BF16)
                        ? (long long)(static_cast<const char*>(resid) - static_cast<const char*>(rows)) : 0;
    p.rows = static_cast<const char*>(rows);
    p.n = (uint32_t)n;
    p.pitch_bytes = (uint32_t)pitch_bytes;
    p.nchunks = (int)(pitch_bytes / 512);
    p.q = q_dev;
    p.d = d;
    p.nq_valid = nq_valid;
    p.k = k;
    p.kpad = next_pow2(k < 32 ? 32 : k);
    p.partial = nullptr;
    const bool bf16 = dtype == B2_DTYPE_BF16;
    if (metric == B2_METRIC_IP)
        return bf16 ? dispatch_nq<B2_METRIC_IP, true>(nq_tile, p, num_sms, stream, partial, grid)
                    : dispatch_nq<B2_METRIC_IP, false>(nq_tile, p, num_sms, stream, partial, grid);
    return bf16 ? dispatch_nq<B2_METRIC_L2, true>(nq_tile, p, num_sms, stream, partial, grid)
                : dispatch_nq<B2_METRIC_L2, false>(nq_tile, p, num_sms, stream, partial, grid);
}

}  // namespace b2

// =====================================================================================
// b2_index_search_static: the same scan + merge, for callers that must not allocate.
//
// Some users of this library cannot let a search touch the CUDA allocator: a retrieval
// service that hands the whole HBM to its own memory manager at start-up (any later
// cudaMalloc / cudaMallocAsync would fail or fragment that plan), or a latency-critical
// request loop that cannot afford the device-wide synchronisation an allocator call may
// trigger.  b2_index_search takes its per-CTA partial top-k lists from cudaMallocAsync;
// this variant keeps them in the fixed array below instead and therefore calls neither
// cudaMalloc, cudaFree nor any allocating library routine.
//
// The array is part of the library image: every GPU that loads libb2flat.so holds these
// 2 MiB until the process exits, whether or not this entry point is ever called.
// =====================================================================================
namespace b2 {

constexpr size_t kStaticPartialEntries = (size_t)256 << 10;  // 256 Ki candidates x 8 B = 2 MiB
static __device__ __align__(256) uint64_t g_static_partials[kStaticPartialEntries];

// `g_static_partials` is only a stand-in on the host side: its device address has to be asked
// for once per device, with that device current.  Cached afterwards.
static int static_partials_ptr(int device, uint64_t** out) {
    static std::mutex mu;
    static uint64_t* cache[64] = {nullptr};
    if (device < 0 || device >= 64) {
        set_error("b2_index_search_static: device ordinal %d not supported (0..63)", device);
        return B2_ERR_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lk(mu);
    if (cache[device] == nullptr) {
        void* p = nullptr;
        B2_CUDA_TRY(cudaGetSymbolAddress(&p, g_static_partials));
        cache[device] = static_cast<uint64_t*>(p);
    }
    *out = cache[device];
    return B2_OK;
}

template <int METRIC, bool BF16, int NQ, int R>
static int run_scan_static(ScanParams p, int num_sms, cudaStream_t stream, uint64_t* scratch, int* grid_out) {
    const size_t smem = scan_smem_bytes(NQ, p.nchunks, BF16, p.kpad);
    int grid = 0;
    B2_TRY((launch_scan<METRIC, BF16, NQ, R>(p, num_sms, smem, stream, &grid)));
    const size_t need = (size_t)NQ * grid * p.k;
    if (need > kStaticPartialEntries) {
        set_error("b2_index_search_static: query tile %d x %d CTAs x k=%d = %zu partial candidates exceed the fixed "
                  "scratch of %zu (2 MiB); use b2_index_search",
                  NQ, grid, p.k, need, kStaticPartialEntries);
        return B2_ERR_UNSUPPORTED;
    }
    p.partial = scratch;
    scan_topk_kernel<METRIC, BF16, NQ, R><<<grid, kScanThreads, smem, stream>>>(p);
    B2_AFTER_LAUNCH();
    *grid_out = grid;
    return B2_OK;
}

template <int METRIC, bool BF16>
static int dispatch_nq_static(int nq_tile, const ScanParams& p, int num_sms, cudaStream_t s, uint64_t* scratch, int* grid) {
    switch (nq_tile) {
        case 1: return run_scan_static<METRIC, BF16, 1, 4>(p, num_sms, s, scratch, grid);
        case 2: return run_scan_static<METRIC, BF16, 2, 4>(p, num_sms, s, scratch, grid);
        case 4: return run_scan_static<METRIC, BF16, 4, 4>(p, num_sms, s, scratch, grid);
        case 8: return run_scan_static<METRIC, BF16, 8, 4>(p, num_sms, s, scratch, grid);
        case 16: return run_scan_static<METRIC, BF16, 16, 4>(p, num_sms, s, scratch, grid);
    }
    set_error("internal: bad query tile %d", nq_tile);
    return B2_ERR_INVALID;
}

}  // namespace b2

extern "C" int b2_index_search_static(b2_index* h, int64_t nq, const float* q_dev, int k, float* D_dev, int64_t* I_dev,
                                      void* stream_) {
    using namespace b2;
    if (nq < 0 || k <= 0 || (nq > 0 && (q_dev == nullptr || D_dev == nullptr || I_dev == nullptr))) {
        set_error("b2_index_search_static: bad arguments (nq=%lld, k=%d)", (long long)nq, k);
        return B2_ERR_INVALID;
    }
    if (nq > 16) {
        set_error("b2_index_search_static: nq=%lld exceeds the 16 queries one scan pass holds", (long long)nq);
        return B2_ERR_UNSUPPORTED;
    }
    if (k > B2_MAX_K) {
        set_error("b2_index_search_static: k=%d exceeds B2_MAX_K=%d", k, B2_MAX_K);
        return B2_ERR_UNSUPPORTED;
    }
    IndexView v;
    B2_TRY(index_view_lock_shared(h, &v));
    if (nq == 0) {
        index_unlock_shared(h);
        return B2_OK;
    }
    int prev = -1;
    int rc = B2_OK;
    if (cudaGetDevice(&prev) != cudaSuccess || (prev != v.device && cudaSetDevice(v.device) != cudaSuccess)) {
        cudaGetLastError();
        set_error("cannot select device %d", v.device);
        rc = B2_ERR_CUDA;
    }
    if (rc == B2_OK) {
        cudaStream_t stream = static_cast<cudaStream_t>(stream_);
        ScanParams p;
        p.rows = static_cast<const char*>(v.rows);
        p.n = (uint32_t)v.ntotal;
        p.pitch_bytes = (uint32_t)v.pitch;
        p.nchunks = (int)(v.pitch / 512);
        p.q = q_dev;
        p.d = v.d;
        p.nq_valid = (int)nq;
        p.k = k;
        p.kpad = next_pow2(k < 32 ? 32 : k);
        p.partial = nullptr;
        p.run_if = nullptr;
        p.resid_delta = (v.resid != nullptr && v.dtype == B2_DTYPE_BF16)
                            ? (long long)(static_cast<const char*>(v.resid) - static_cast<const char*>(v.rows)) : 0;
        const bool bf16 = v.dtype == B2_DTYPE_BF16;
        const int tile = next_pow2((int)nq);
        uint64_t* scratch = nullptr;
        int grid = 0;
        if (scan_smem_bytes(tile, p.nchunks, bf16, p.kpad) > 200 * 1024) {
            set_error("b2_index_search_static: %lld queries with d=%d, k=%d do not fit one scan pass; use b2_index_search",
                      (long long)nq, v.d, k);
            rc = B2_ERR_UNSUPPORTED;
        }
        if (rc == B2_OK) rc = static_partials_ptr(v.device, &scratch);
        if (rc == B2_OK) {
            if (v.metric == B2_METRIC_IP)
                rc = bf16 ? dispatch_nq_static<B2_METRIC_IP, true>(tile, p, v.num_sms, stream, scratch, &grid)
                          : dispatch_nq_static<B2_METRIC_IP, false>(tile, p, v.num_sms, stream, scratch, &grid);
            else
                rc = bf16 ? dispatch_nq_static<B2_METRIC_L2, true>(tile, p, v.num_sms, stream, scratch, &grid)
                          : dispatch_nq_static<B2_METRIC_L2, false>(tile, p, v.num_sms, stream, scratch, &grid);
        }
        if (rc == B2_OK) rc = merge_partials(scratch, (int)nq, grid, k, v.metric, 0, D_dev, I_dev, (size_t)k, stream);
    }
    if (prev >= 0 && prev != v.device) cudaSetDevice(prev);
    index_unlock_shared(h);
    return rc;
}
