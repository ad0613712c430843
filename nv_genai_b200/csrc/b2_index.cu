// C ABI of the B200 flat index: handle, device row store, dispatch to the kernels.
// Declarations and the reference interfaces they replace: include/b2_flat.h.
#include <cuda.h>

#include <algorithm>
#include <cstring>
#include <mutex>
#include <shared_mutex>
#include <string>
#include <thread>
#include <vector>

#include "b2_common.cuh"
#include "b2_kernels.h"

namespace b2 {

// ------------------------------------------------------------------ error plumbing
static thread_local char t_err[512] = {0};
std::atomic<long long> g_launches{0};

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof(t_err), fmt, ap);
    va_end(ap);
}

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) {
            ok = false;
            prev = -1;
            return;
        }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

// ------------------------------------------------------------------ driver entry points (no link-time libcuda)
struct DriverApi {
    bool ok = false;
    CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
    CUresult (*MemAddressFree)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long) = nullptr;
    CUresult (*MemRelease)(CUmemGenericAllocationHandle) = nullptr;
    CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
    CUresult (*MemUnmap)(CUdeviceptr, size_t) = nullptr;
    CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
    CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags) = nullptr;
};

static bool load_sym(const char* name, void** fn) {
    cudaDriverEntryPointQueryResult qres;
    *fn = nullptr;
    if (cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &qres) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return qres == cudaDriverEntryPointSuccess && *fn != nullptr;
}

static const DriverApi& driver() {
    static DriverApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        bool ok = true;
        ok &= load_sym("cuMemAddressReserve", (void**)&api.MemAddressReserve);
        ok &= load_sym("cuMemAddressFree", (void**)&api.MemAddressFree);
        ok &= load_sym("cuMemCreate", (void**)&api.MemCreate);
        ok &= load_sym("cuMemRelease", (void**)&api.MemRelease);
        ok &= load_sym("cuMemMap", (void**)&api.MemMap);
        ok &= load_sym("cuMemUnmap", (void**)&api.MemUnmap);
        ok &= load_sym("cuMemSetAccess", (void**)&api.MemSetAccess);
        ok &= load_sym("cuMemGetAllocationGranularity", (void**)&api.MemGetAllocationGranularity);
        api.ok = ok;
    });
    return api;
}

// ------------------------------------------------------------------ row store
// One contiguous virtual range per index, reserved once for the whole HBM and backed by
// physical chunks as rows arrive (cuMemCreate/cuMemMap), so a 41 GB store grows without the
// realloc-and-copy that std::vector<uint8_t> codes does in IndexFlatCodes::add.  If the VMM
// driver API is unavailable the store falls back to cudaMalloc with geometric growth.
struct RowStore {
    int device = 0;
    bool vmm = false;
    CUdeviceptr va = 0;
    size_t reserved = 0, mapped = 0, gran = 0;
    struct Chunk { CUmemGenericAllocationHandle h; size_t off, size; };
    std::vector<Chunk> chunks;
    void* plain = nullptr;  // fallback
    size_t plain_cap = 0;

    char* base() const { return vmm ? reinterpret_cast<char*>(va) : static_cast<char*>(plain); }
    size_t capacity() const { return vmm ? mapped : plain_cap; }

    int init(int dev) {
        device = dev;
        const DriverApi& d = driver();
        if (!d.ok || getenv("B2_NO_VMM") != nullptr) return B2_OK;
        CUmemAllocationProp prop;
        memset(&prop, 0, sizeof(prop));
        prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
        prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
        prop.location.id = dev;
        size_t g = 0;
        if (d.MemGetAllocationGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || g == 0)
            return B2_OK;
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) {
            cudaGetLastError();
            return B2_OK;
        }
        const size_t want = ((total_b + g - 1) / g) * g;
        CUdeviceptr p = 0;
        if (d.MemAddressReserve(&p, want, 0, 0, 0) != CUDA_SUCCESS) return B2_OK;
        va = p;
        reserved = want;
        gran = g;
        vmm = true;
        return B2_OK;
    }

    int ensure(size_t bytes, cudaStream_t stream) {
        if (bytes <= capacity()) return B2_OK;
        if (vmm) {
            const DriverApi& d = driver();
            size_t grow = bytes - mapped;
            const size_t quarter = std::min<size_t>(mapped / 4, (size_t)4 << 30);
            grow = std::max(grow, quarter);
            grow = ((grow + gran - 1) / gran) * gran;
            if (mapped + grow > reserved) grow = reserved - mapped;
            if (mapped + grow < bytes) {
                set_error("row store: %zu bytes exceed the device (%zu reserved)", bytes, reserved);
                return B2_ERR_OOM;
            }
            CUmemAllocationProp prop;
            memset(&prop, 0, sizeof(prop));
            prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
            prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
            prop.location.id = device;
            CUmemGenericAllocationHandle h;
            CUresult r = d.MemCreate(&h, grow, &prop, 0);
            if (r != CUDA_SUCCESS && grow > ((bytes - mapped + gran - 1) / gran) * gran) {
                grow = ((bytes - mapped + gran - 1) / gran) * gran;  // retry with the bare minimum
                r = d.MemCreate(&h, grow, &prop, 0);
            }
            if (r != CUDA_SUCCESS) {
                set_error("row store: cuMemCreate(%zu) failed (%d)", grow, (int)r);
                return B2_ERR_OOM;
            }
            if ((r = d.MemMap(va + mapped, grow, 0, h, 0)) != CUDA_SUCCESS) {
                d.MemRelease(h);
                set_error("row store: cuMemMap failed (%d)", (int)r);
                return B2_ERR_CUDA;
            }
            CUmemAccessDesc acc;
            memset(&acc, 0, sizeof(acc));
            acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
            acc.location.id = device;
            acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
            if ((r = d.MemSetAccess(va + mapped, grow, &acc, 1)) != CUDA_SUCCESS) {
                d.MemUnmap(va + mapped, grow);
                d.MemRelease(h);
                set_error("row store: cuMemSetAccess failed (%d)", (int)r);
                return B2_ERR_CUDA;
            }
            chunks.push_back({h, mapped, grow});
            mapped += grow;
            return B2_OK;
        }
        // fallback: geometric growth with a copy
        size_t ncap = std::max(bytes, plain_cap + plain_cap / 2);
        ncap = std::max<size_t>(ncap, (size_t)1 << 20);
        void* np = nullptr;
        cudaError_t e = cudaMalloc(&np, ncap);
        if (e != cudaSuccess) {
            cudaGetLastError();
            e = cudaMalloc(&np, bytes);
            ncap = bytes;
        }
        if (e != cudaSuccess) {
            cudaGetLastError();
            set_error("row store: cudaMalloc(%zu) failed: %s", ncap, cudaGetErrorString(e));
            return B2_ERR_OOM;
        }
        if (plain != nullptr) {
            // the old contents may still be in use by work on `stream`: copy in stream order, then free
            if (plain_cap) B2_CUDA_TRY(cudaMemcpyAsync(np, plain, plain_cap, cudaMemcpyDeviceToDevice, stream));
            B2_CUDA_TRY(cudaStreamSynchronize(stream));
            cudaFree(plain);
        }
        plain = np;
        plain_cap = ncap;
        return B2_OK;
    }

    void release() {
        if (vmm) {
            const DriverApi& d = driver();
            for (auto& c : chunks) {
                d.MemUnmap(va + c.off, c.size);
                d.MemRelease(c.h);
            }
            chunks.clear();
            if (va) d.MemAddressFree(va, reserved);
            va = 0;
            mapped = reserved = 0;
            vmm = false;
        }
        if (plain) cudaFree(plain);
        plain = nullptr;
        plain_cap = 0;
    }
};

}  // namespace b2

using namespace b2;

struct b2_index {
    int d = 0, metric = 0, dtype = 0, device = 0, num_sms = 148;
    size_t pitch = 0;  // bytes per row, multiple of 512
    int64_t ntotal = 0;
    RowStore store;
    float* norms = nullptr;  // squared norm of each stored row
    int64_t norms_cap = 0;
    uint32_t* max_norm_ord = nullptr;  // device: ordered image of the largest squared row norm seen so far
    int gemm_query_parts = 0;  // 0 = default (filter and refine), 2 = all-exact tensor arithmetic, 1 = single part (approximate)
    // bf16 shadow of fp32 rows for the approximate pass of the tensor-core search: built lazily by the first large
    // batch (under shadow_mu, readers hold only the shared lock), extended after adds, dropped by removals.
    RowStore shadow;
    size_t shadow_pitch = 0;
    int64_t shadow_rows = 0;           // rows [0, shadow_rows) of the shadow are current
    uint32_t* shadow_err_ord = nullptr;  // device: ordered image of max |x - bf16(x)|^2 over the shadowed rows
    cudaEvent_t shadow_ev = nullptr;   // completion of the latest shadow build
    int shadow_enabled = 1;            // option "bf16_shadow"
    int64_t shadow_refused_rows = 0;   // last size the device had no room for (do not retry until the index changes)
    std::mutex shadow_mu;
    // bf16 rows with a residual plane (option "bf16_residual", set on an empty bf16 index): a second bf16 plane of the
    // same pitch holds lo = bf16(x - bf16(x)), the row the index stands for is hi + lo (about 16 mantissa bits).  The
    // tensor-core passes read the main plane only - the scan streams 2 bytes per element as on a plain bf16 store - and
    // every returned score is re-computed against hi + lo, with the band widened by the residual's norm.
    RowStore resid;
    int residual_enabled = 0;
    uint32_t* resid_err_ord = nullptr;  // device: ordered image of max |lo|^2 over the rows
    std::shared_mutex mu;
};

static const void* resid_base(const b2_index* h) { return h->residual_enabled ? h->resid.base() : nullptr; }

// ------------------------------------------------------------------ host ingest: library-owned pinned staging ring
// A caller of the reference API hands over pageable numpy rows (np.array(embeddings, dtype=np.float32) in
// langchain's FAISS.__add).  A cudaMemcpyAsync from pageable memory is staged by the driver through a small pinned
// buffer by one thread and does not overlap anything; here the rows are copied by up to 8 host threads into a ring of
// pinned slots owned by the library (one ring per device, allocated by the first large pageable add), each slot is
// sent by DMA on a copy stream as soon as it is full, and the ingest kernel of chunk i runs on the caller's stream
// while chunk i + 1 is on the wire.  Pinned (or registered) sources skip the ring and are sent from where they are.
namespace {
struct PinnedRing {
    static constexpr int kSlots = 4;
    static constexpr size_t kSlotBytes = (size_t)32 << 20;
    std::mutex mu;  // one host add per device stages through the ring at a time
    char* slot[kSlots] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t drained[kSlots] = {nullptr, nullptr, nullptr, nullptr};  // the DMA out of the slot has finished
    bool tried = false, ok = false;
    bool ensure() {  // with the ring's device current
        if (tried) return ok;
        tried = true;
        ok = true;
        for (int i = 0; i < kSlots && ok; ++i) {
            ok = cudaHostAlloc(reinterpret_cast<void**>(&slot[i]), kSlotBytes, cudaHostAllocPortable) == cudaSuccess &&
                 cudaEventCreateWithFlags(&drained[i], cudaEventDisableTiming) == cudaSuccess;
        }
        if (!ok) {
            cudaGetLastError();
            for (int i = 0; i < kSlots; ++i) {
                if (slot[i]) cudaFreeHost(slot[i]);
                if (drained[i]) cudaEventDestroy(drained[i]);
                slot[i] = nullptr;
                drained[i] = nullptr;
            }
        }
        return ok;
    }
};
PinnedRing g_rings[64];

void parallel_copy(char* dst, const char* src, size_t bytes) {
    static const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const unsigned T = bytes < ((size_t)4 << 20) ? 1u : std::min(8u, std::max(1u, hw / 2));
    if (T == 1) {
        memcpy(dst, src, bytes);
        return;
    }
    std::thread th[8];
    const size_t part = ((bytes / T) + 4095) & ~(size_t)4095;
    for (unsigned t = 1; t < T; ++t) {
        const size_t off = std::min(bytes, part * t), len = std::min(bytes - off, part);
        th[t] = std::thread([=] { if (len) memcpy(dst + off, src + off, len); });
    }
    memcpy(dst, src, std::min(bytes, part));
    for (unsigned t = 1; t < T; ++t) th[t].join();
}
}  // namespace

static size_t row_pitch_bytes(int d, int dtype) {
    const size_t raw = (size_t)d * (dtype == B2_DTYPE_BF16 ? 2 : 4);
    return ((raw + 511) / 512) * 512;
}

static int ensure_norms(b2_index* h, int64_t n, cudaStream_t stream) {
    if (n <= h->norms_cap) return B2_OK;
    int64_t ncap = std::max<int64_t>(n, h->norms_cap + h->norms_cap / 2);
    ncap = std::max<int64_t>(ncap, 1024);
    float* np = nullptr;
    B2_CUDA_TRY(cudaMalloc(&np, (size_t)ncap * sizeof(float)));
    if (h->norms != nullptr) {
        cudaError_t e = cudaSuccess;
        if (h->ntotal > 0)
            e = cudaMemcpyAsync(np, h->norms, (size_t)h->ntotal * sizeof(float), cudaMemcpyDeviceToDevice, stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) {
            cudaFree(np);  // the old array stays in place
            set_error("growing the norm array failed: %s", cudaGetErrorString(e));
            return B2_ERR_CUDA;
        }
        cudaFree(h->norms);
    }
    h->norms = np;
    h->norms_cap = ncap;
    return B2_OK;
}

static int check_handle(const b2_index* h) {
    if (h == nullptr) {
        set_error("null index handle");
        return B2_ERR_INVALID;
    }
    return B2_OK;
}

namespace b2 {
int index_view_lock_shared(b2_index* h, IndexView* v) {
    B2_TRY(check_handle(h));
    h->mu.lock_shared();
    v->rows = h->store.base();
    v->resid = resid_base(h);
    v->ntotal = h->ntotal;
    v->pitch = h->pitch;
    v->d = h->d;
    v->dtype = h->dtype;
    v->metric = h->metric;
    v->device = h->device;
    v->num_sms = h->num_sms;
    return B2_OK;
}
void index_unlock_shared(b2_index* h) { h->mu.unlock_shared(); }
}  // namespace b2

extern "C" {

const char* b2_last_error(void) { return t_err; }
const char* b2_version(void) { return "b2flat 0.1 (sm_100a)"; }
int64_t b2_launch_count(void) { return (int64_t)g_launches.load(); }
void b2_launch_count_reset(void) { g_launches.store(0); }

int b2_prof_enable(int on) {
    prof_enable(on);
    return B2_OK;
}
int b2_prof_read(double* scan_ms, int64_t* scan_launches, int reset) { return prof_read(0, scan_ms, scan_launches, reset); }
int b2_prof_read_kernel(int kernel, double* ms, int64_t* launches, int reset) {
    if (kernel != B2_PROF_SCAN && kernel != B2_PROF_GEMM) {
        set_error("b2_prof_read_kernel: unknown kernel %d", kernel);
        return B2_ERR_INVALID;
    }
    return prof_read(kernel, ms, launches, reset);
}

int b2_index_create(int d, int metric, int dtype, int device, b2_index** out) {
    if (out == nullptr) {
        set_error("b2_index_create: out is null");
        return B2_ERR_INVALID;
    }
    *out = nullptr;
    if (d <= 0 || d > 32768) {
        set_error("b2_index_create: dimension %d out of range [1, 32768]", d);
        return B2_ERR_INVALID;
    }
    if (metric != B2_METRIC_IP && metric != B2_METRIC_L2) {
        set_error("b2_index_create: unknown metric %d", metric);
        return B2_ERR_INVALID;
    }
    if (dtype != B2_DTYPE_F32 && dtype != B2_DTYPE_BF16) {
        set_error("b2_index_create: unknown dtype %d", dtype);
        return B2_ERR_INVALID;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        set_error("no CUDA device available: this library has no CPU path");
        return B2_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= ndev) {
        set_error("b2_index_create: device %d out of range (have %d)", device, ndev);
        return B2_ERR_INVALID;
    }
    DeviceGuard g(device);
    if (!g.ok) {
        set_error("cannot select device %d", device);
        return B2_ERR_CUDA;
    }
    int major = 0, sms = 0;
    B2_CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
    B2_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    if (major != 10) {
        set_error("device %d is sm_%d0: this build targets sm_100a (B200) only", device, major);
        return B2_ERR_NO_DEVICE;
    }
    // keep stream-ordered scratch cached in the pool instead of returning it to the OS
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
        unsigned long long thr = ~0ull;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    cudaGetLastError();
    b2_index* h = new (std::nothrow) b2_index();
    if (h == nullptr) {
        set_error("out of host memory");
        return B2_ERR_OOM;
    }
    h->d = d;
    h->metric = metric;
    h->dtype = dtype;
    h->device = device;
    h->num_sms = sms;
    h->pitch = row_pitch_bytes(d, dtype);
    h->store.init(device);
    if (dtype == B2_DTYPE_F32 && getenv("B2_NO_SHADOW") == nullptr) {
        h->shadow_pitch = row_pitch_bytes(d, B2_DTYPE_BF16);
        h->shadow.init(device);
        if (!h->shadow.vmm) h->shadow_enabled = 0;  // growing a plain allocation would move rows under running searches
    } else {
        h->shadow_enabled = 0;
    }
    h->resid.init(device);
    if (cudaMalloc(&h->max_norm_ord, 3 * sizeof(uint32_t)) != cudaSuccess ||
        cudaMemset(h->max_norm_ord, 0, 3 * sizeof(uint32_t)) != cudaSuccess) {  // 0 orders below every float
        cudaGetLastError();
        delete h;
        set_error("out of device memory");
        return B2_ERR_OOM;
    }
    h->shadow_err_ord = h->max_norm_ord + 1;
    h->resid_err_ord = h->max_norm_ord + 2;
    *out = h;
    return B2_OK;
}

int b2_index_destroy(b2_index* h) {
    if (h == nullptr) return B2_OK;
    {
        DeviceGuard g(h->device);
        cudaDeviceSynchronize();
        h->store.release();
        h->shadow.release();
        h->resid.release();
        if (h->shadow_ev) cudaEventDestroy(h->shadow_ev);
        if (h->norms) cudaFree(h->norms);
        if (h->max_norm_ord) cudaFree(h->max_norm_ord);
        cudaGetLastError();
    }
    delete h;
    return B2_OK;
}

int64_t b2_index_ntotal(const b2_index* h) { return h ? h->ntotal : -1; }
int b2_index_d(const b2_index* h) { return h ? h->d : -1; }
int b2_index_metric(const b2_index* h) { return h ? h->metric : -1; }
int b2_index_dtype(const b2_index* h) { return h ? h->dtype : -1; }
int b2_index_device(const b2_index* h) { return h ? h->device : -1; }

int b2_index_add(b2_index* h, int64_t n, const float* x, int x_where, unsigned flags, void* stream_) {
    B2_TRY(check_handle(h));
    if (n < 0 || (n > 0 && x == nullptr)) {
        set_error("b2_index_add: bad arguments (n=%lld)", (long long)n);
        return B2_ERR_INVALID;
    }
    if (n == 0) return B2_OK;
    if (h->ntotal + n >= (int64_t)0xfffffff0ll) {
        set_error("b2_index_add: more than 2^32-16 rows per index are not supported");
        return B2_ERR_UNSUPPORTED;
    }
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    std::unique_lock<std::shared_mutex> lk(h->mu);
    DeviceGuard g(h->device);
    if (!g.ok) {
        set_error("cannot select device %d", h->device);
        return B2_ERR_CUDA;
    }
    B2_TRY(h->store.ensure((size_t)(h->ntotal + n) * h->pitch, stream));
    if (h->residual_enabled) B2_TRY(h->resid.ensure((size_t)(h->ntotal + n) * h->pitch, stream));
    B2_TRY(ensure_norms(h, h->ntotal + n, stream));
    const bool normalize = (flags & B2_ADD_NORMALIZE) != 0;
    char* dst = h->store.base() + (size_t)h->ntotal * h->pitch;
    char* rdst = h->residual_enabled ? h->resid.base() + (size_t)h->ntotal * h->pitch : nullptr;  // advanced like dst
    uint32_t* rerr = h->residual_enabled ? h->resid_err_ord : nullptr;
    float* nrm = h->norms + h->ntotal;
    if (x_where == B2_DEVICE) {
        B2_TRY(ingest_rows(x, n, h->d, dst, h->pitch, h->dtype, normalize, nrm, stream, rdst, rerr));
    } else {
        const size_t row_bytes = (size_t)h->d * sizeof(float);
        const size_t total_bytes = (size_t)n * row_bytes;
        if (total_bytes <= ((size_t)8 << 20) || row_bytes > PinnedRing::kSlotBytes) {
            // a few documents at a time (the usual langchain add): one staged copy on the caller's stream
            float* sb = nullptr;
            B2_CUDA_TRY(cudaMallocAsync(&sb, total_bytes, stream));
            cudaError_t e = cudaMemcpyAsync(sb, x, total_bytes, cudaMemcpyHostToDevice, stream);
            int rc = B2_OK;
            if (e != cudaSuccess) {
                set_error("b2_index_add: H2D copy failed: %s", cudaGetErrorString(e));
                rc = B2_ERR_CUDA;
            } else {
                rc = ingest_rows(sb, n, h->d, dst, h->pitch, h->dtype, normalize, nrm, stream, rdst, rerr);
            }
            cudaFreeAsync(sb, stream);
            if (rc != B2_OK) return rc;
            B2_CUDA_TRY(cudaStreamSynchronize(stream));  // the caller may reuse its host buffer as soon as we return
        } else {
            // bulk load: chunks of one ring slot, copy stream + compute stream, two device staging buffers
            cudaPointerAttributes pa;
            bool pinned_src = cudaPointerGetAttributes(&pa, x) == cudaSuccess && pa.type == cudaMemoryTypeHost;
            cudaGetLastError();
            PinnedRing& ring = g_rings[h->device & 63];
            std::unique_lock<std::mutex> ring_lk(ring.mu, std::defer_lock);
            bool use_ring = false;
            if (!pinned_src && getenv("B2_NO_PINNED_RING") == nullptr) {
                ring_lk.lock();
                use_ring = ring.ensure();
                if (!use_ring) ring_lk.unlock();  // no pinned memory to be had: the driver's own staging
            }
            const int64_t chunk = std::max<int64_t>(1, (int64_t)(PinnedRing::kSlotBytes / row_bytes));
            float* stage[2] = {nullptr, nullptr};
            cudaEvent_t copied[2] = {nullptr, nullptr}, consumed[2] = {nullptr, nullptr}, ready = nullptr;
            cudaStream_t cs = nullptr;
            int rc = B2_OK;
            auto cu = [&](cudaError_t e, const char* what) {
                if (e != cudaSuccess && rc == B2_OK) {
                    set_error("b2_index_add: %s: %s", what, cudaGetErrorString(e));
                    rc = e == cudaErrorMemoryAllocation ? B2_ERR_OOM : B2_ERR_CUDA;
                }
                return e == cudaSuccess;
            };
            cu(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking), "copy stream");
            for (int b = 0; b < 2 && rc == B2_OK; ++b) {
                cu(cudaMallocAsync(&stage[b], (size_t)chunk * row_bytes, stream), "staging allocation");
                cu(cudaEventCreateWithFlags(&copied[b], cudaEventDisableTiming), "event");
                cu(cudaEventCreateWithFlags(&consumed[b], cudaEventDisableTiming), "event");
            }
            if (rc == B2_OK) {  // the copy stream may touch the staging buffers once their allocation is ordered
                cu(cudaEventCreateWithFlags(&ready, cudaEventDisableTiming), "event");
                if (rc == B2_OK) cu(cudaEventRecord(ready, stream), "event record");
                if (rc == B2_OK) cu(cudaStreamWaitEvent(cs, ready, 0), "stream wait");
            }
            int64_t c = 0;
            for (int64_t off = 0; off < n && rc == B2_OK; off += chunk, ++c) {
                const int64_t m = std::min(chunk, n - off);
                const size_t bytes = (size_t)m * row_bytes;
                const int b = (int)(c & 1);
                const char* src = reinterpret_cast<const char*>(x) + (size_t)off * row_bytes;
                if (c >= 2) cu(cudaStreamWaitEvent(cs, consumed[b], 0), "stream wait");  // chunk c - 2 has been ingested
                if (use_ring) {
                    const int sl = (int)(c % PinnedRing::kSlots);
                    if (c >= PinnedRing::kSlots) cu(cudaEventSynchronize(ring.drained[sl]), "ring wait");
                    parallel_copy(ring.slot[sl], src, bytes);
                    cu(cudaMemcpyAsync(stage[b], ring.slot[sl], bytes, cudaMemcpyHostToDevice, cs), "H2D copy");
                    cu(cudaEventRecord(ring.drained[sl], cs), "event record");
                } else {
                    cu(cudaMemcpyAsync(stage[b], src, bytes, cudaMemcpyHostToDevice, cs), "H2D copy");
                }
                cu(cudaEventRecord(copied[b], cs), "event record");
                cu(cudaStreamWaitEvent(stream, copied[b], 0), "stream wait");
                if (rc == B2_OK)
                    rc = ingest_rows(stage[b], m, h->d, dst + (size_t)off * h->pitch, h->pitch, h->dtype, normalize, nrm + off, stream,
                                     rdst ? rdst + (size_t)off * h->pitch : nullptr, rerr);
                cu(cudaEventRecord(consumed[b], stream), "event record");
            }
            // the caller may reuse its host buffer as soon as we return; the staging buffers are freed in stream order
            if (cs) cudaStreamSynchronize(cs);
            const cudaError_t es = cudaStreamSynchronize(stream);
            for (int b = 0; b < 2; ++b) {
                if (stage[b]) cudaFreeAsync(stage[b], stream);
                if (copied[b]) cudaEventDestroy(copied[b]);
                if (consumed[b]) cudaEventDestroy(consumed[b]);
            }
            if (ready) cudaEventDestroy(ready);
            if (cs) cudaStreamDestroy(cs);
            if (rc != B2_OK) return rc;
            B2_CUDA_TRY(es);
        }
    }
    B2_TRY(update_max_norm(nrm, n, h->max_norm_ord, stream));
    h->ntotal += n;
    return B2_OK;
}

// largest query tile (power of two <= 16) whose shared memory fits for this k
static int max_query_tile(const b2_index* h, int kpad) {
    const int nchunks = (int)(h->pitch / 512);
    for (int t = 16; t >= 1; t >>= 1)
        if (scan_smem_bytes(t, nchunks, h->dtype == B2_DTYPE_BF16, kpad) <= 200 * 1024) return t;
    return 0;
}

__global__ void fill_empty_kernel(float* D, int64_t* I, int64_t n, float dval) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        D[i] = dval;
        I[i] = -1;
    }
}

// The bf16 shadow of fp32 rows, current for all ntotal rows and ordered before later work on `stream`; false when the
// index has none (bf16 rows, option off, no VMM, or no room on the device: the search then runs on the fp32 rows).
// Caller holds the shared lock, so ntotal and the fp32 rows are stable; builders serialise on shadow_mu.
static bool ensure_shadow(b2_index* h, cudaStream_t stream, ApproxRows* out) {
    if (!h->shadow_enabled || h->dtype != B2_DTYPE_F32) return false;
    std::lock_guard<std::mutex> lk(h->shadow_mu);
    const int64_t n = h->ntotal;
    if (h->shadow_rows < n) {
        if (h->shadow_refused_rows == n) return false;
        const size_t need = (size_t)n * h->shadow_pitch;
        if (need > h->shadow.capacity()) {
            size_t free_b = 0, total_b = 0;
            if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) {
                cudaGetLastError();
                return false;
            }
            // leave room for the search scratch and for the application: the shadow is an accelerator, not a need
            const size_t grow = need - h->shadow.capacity();
            if (free_b < grow + grow / 4 + ((size_t)6 << 30) || h->shadow.ensure(need, stream) != B2_OK) {
                h->shadow_refused_rows = n;
                return false;
            }
        }
        if (h->shadow_ev == nullptr && cudaEventCreateWithFlags(&h->shadow_ev, cudaEventDisableTiming) != cudaSuccess) {
            cudaGetLastError();
            h->shadow_ev = nullptr;
            return false;
        }
        // an earlier build may still be running on another stream: order this one behind it
        if (h->shadow_rows > 0) cudaStreamWaitEvent(stream, h->shadow_ev, 0);
        if (build_shadow_rows(h->store.base(), h->pitch, h->shadow.base(), h->shadow_pitch, h->shadow_rows,
                              n - h->shadow_rows, h->d, h->shadow_err_ord, stream) != B2_OK)
            return false;
        cudaEventRecord(h->shadow_ev, stream);
        h->shadow_rows = n;
    } else if (h->shadow_ev != nullptr) {
        cudaStreamWaitEvent(stream, h->shadow_ev, 0);  // built on another stream, possibly still in flight
    }
    out->rows = h->shadow.base();
    out->pitch = h->shadow_pitch;
    out->err_ord = h->shadow_err_ord;
    return true;
}

// Device-side search: q_dev [nq][d], D_dev/I_dev [nq][k]; caller holds the shared lock and the device.
static int search_device(b2_index* h, int64_t nq, const float* q_dev, int k, float* D_dev, int64_t* I_dev,
                         cudaStream_t stream, const PeerPush* push = nullptr) {
    if (h->ntotal == 0 && push != nullptr) {
        // an empty shard still has to deliver its (all-padding) lists and raise its flags
        for (int64_t off = 0; off < nq; off += 1024) {
            PeerPush pp = *push;
            pp.q_offset = off;
            B2_TRY(merge_partials(nullptr, (int)std::min<int64_t>(1024, nq - off), 0, k, h->metric, 0, nullptr, nullptr,
                                  (size_t)k, stream, nullptr, &pp));
        }
        return B2_OK;
    }
    if (h->ntotal == 0) {
        const int64_t tot = nq * k;
        fill_empty_kernel<<<(unsigned)std::min<int64_t>((tot + 255) / 256, 1024), 256, 0, stream>>>(
            D_dev, I_dev, tot, h->metric == B2_METRIC_IP ? -FLT_MAX : FLT_MAX);
        B2_AFTER_LAUNCH();
        return B2_OK;
    }
    const bool tc_approx = h->gemm_query_parts == 0 && h->max_norm_ord != nullptr;
    // bf16 rows with a residual plane: the tensor-core paths only in their filter-and-refine form (the all-exact
    // tensor arithmetic would score against the main plane alone); everything else runs the two-plane FFMA scan below
    ResidualRows rr;
    rr.rows = resid_base(h);
    rr.err_ord = h->resid_err_ord;
    const ResidualRows* resid = h->residual_enabled ? &rr : nullptr;
    const bool tc_ok = resid == nullptr || (tc_approx && k <= 64 && scan_tc_approx_feasible(h->ntotal, k, h->num_sms));
    if (tc_ok && scan_tc_supported(h->ntotal, h->d, h->dtype, h->metric, k, nq, tc_approx, resid != nullptr))
        return scan_tc_search(h->store.base(), h->ntotal, h->pitch, h->d, h->dtype, q_dev, (int)nq, k, h->num_sms,
                              D_dev, I_dev, stream, push, tc_approx ? h->max_norm_ord : nullptr, h->metric, h->norms, resid);
    if (gemm_path_supported(h->ntotal, h->d, h->dtype, k, nq) &&
        (resid == nullptr || gemm_uses_refine(h->gemm_query_parts, h->max_norm_ord))) {
        ApproxRows approx;
        const bool shadowed = gemm_uses_refine(h->gemm_query_parts, h->max_norm_ord) && ensure_shadow(h, stream, &approx);
        const int rc = gemm_search(h->store.base(), h->ntotal, h->pitch, h->d, h->dtype, h->metric, h->norms, q_dev, nq, k,
                                   h->num_sms, D_dev, I_dev, stream, push, h->gemm_query_parts, h->max_norm_ord,
                                   shadowed ? &approx : nullptr, resid);
        if (rc != B2_INTERNAL_REDO_EXACT) return rc;
    }
    const int kpad = next_pow2(k < 32 ? 32 : k);
    const int tmax = max_query_tile(h, kpad);
    if (tmax == 0) {
        set_error("search: d=%d with k=%d does not fit the scan kernel's shared memory", h->d, k);
        return B2_ERR_UNSUPPORTED;
    }
    for (int64_t off = 0; off < nq;) {
        const int64_t rem = nq - off;
        int tile = (int)std::min<int64_t>(next_pow2((int)std::min<int64_t>(rem, 16)), tmax);
        const int valid = (int)std::min<int64_t>(rem, tile);
        uint64_t* partial = nullptr;
        int grid = 0;
        B2_TRY(scan_pass(h->store.base(), h->ntotal, h->pitch, h->d, h->dtype, h->metric, q_dev + (size_t)off * h->d,
                         valid, tile, k, h->num_sms, stream, &partial, &grid, nullptr, resid_base(h)));
        PeerPush pp;
        if (push != nullptr) {
            pp = *push;
            pp.q_offset = off;
        }
        int rc = merge_partials(partial, valid, grid, k, h->metric, 0, D_dev ? D_dev + (size_t)off * k : nullptr,
                                I_dev ? I_dev + (size_t)off * k : nullptr, (size_t)k, stream, nullptr,
                                push != nullptr ? &pp : nullptr);
        cudaFreeAsync(partial, stream);
        B2_TRY(rc);
        off += valid;
    }
    return B2_OK;
}

// Small host inputs / outputs of a search travel through a pinned bounce buffer owned by the calling thread: a copy
// between pageable memory and the device is staged by the driver and synchronous, two of them (D and I) twice; with the
// bounce buffer a host-buffer search is ONE async H2D, the kernels, ONE async D2H of the packed (I | D) block and one
// stream synchronisation.  Grows to what the thread's searches need, up to 1 MiB; larger transfers take the plain path.
namespace {
struct HostBounce {
    char* p = nullptr;
    size_t cap = 0;
    ~HostBounce() {
        if (p) cudaFreeHost(p);
    }
    char* get(size_t bytes) {
        if (bytes > ((size_t)1 << 20)) return nullptr;
        if (bytes > cap) {
            if (p) cudaFreeHost(p);
            p = nullptr;
            cap = 0;
            size_t want = 65536;
            while (want < bytes) want <<= 1;
            if (cudaHostAlloc(reinterpret_cast<void**>(&p), want, cudaHostAllocPortable) != cudaSuccess) {
                cudaGetLastError();
                p = nullptr;
                return nullptr;
            }
            cap = want;
        }
        return p;
    }
};
thread_local HostBounce t_bounce;
}  // namespace

int b2_index_search(b2_index* h, int64_t nq, const float* q, int q_where, int k, float* D, int64_t* I, int out_where,
                    void* stream_) {
    B2_TRY(check_handle(h));
    if (nq < 0 || k <= 0 || (nq > 0 && (q == nullptr || D == nullptr || I == nullptr))) {
        set_error("b2_index_search: bad arguments (nq=%lld, k=%d)", (long long)nq, k);
        return B2_ERR_INVALID;
    }
    if (k > B2_MAX_K) {
        set_error("b2_index_search: k=%d exceeds B2_MAX_K=%d", k, B2_MAX_K);
        return B2_ERR_UNSUPPORTED;
    }
    if (nq == 0) return B2_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    std::shared_lock<std::shared_mutex> lk(h->mu);
    DeviceGuard g(h->device);
    if (!g.ok) {
        set_error("cannot select device %d", h->device);
        return B2_ERR_CUDA;
    }
    const float* q_dev = q;
    float* q_tmp = nullptr;
    float* D_dev = D;
    int64_t* I_dev = I;
    void* out_tmp = nullptr;
    const size_t nres = (size_t)nq * k;
    int rc = B2_OK;
    // one bounce buffer for both directions: [queries | packed results]
    const size_t q_bytes = q_where == B2_HOST ? (size_t)nq * h->d * sizeof(float) : 0;
    const size_t r_bytes = out_where == B2_HOST ? nres * 12 : 0;
    const size_t q_pad = (q_bytes + 255) & ~(size_t)255;
    // only when the call ends with a stream synchronisation (host outputs): the copies out of / into the buffer are
    // then complete before this thread can reuse it
    char* bounce = out_where == B2_HOST ? t_bounce.get(q_pad + r_bytes) : nullptr;
    if (q_where == B2_HOST) {
        B2_CUDA_TRY(cudaMallocAsync(&q_tmp, (size_t)nq * h->d * sizeof(float), stream));
        const void* src = q;
        if (bounce != nullptr) {
            memcpy(bounce, q, q_bytes);
            src = bounce;
        }
        cudaError_t e = cudaMemcpyAsync(q_tmp, src, (size_t)nq * h->d * sizeof(float), cudaMemcpyHostToDevice, stream);
        if (e != cudaSuccess) {
            set_error("b2_index_search: query copy failed: %s", cudaGetErrorString(e));
            rc = B2_ERR_CUDA;
        }
        q_dev = q_tmp;
    }
    if (rc == B2_OK && out_where == B2_HOST) {
        cudaError_t e = cudaMallocAsync(&out_tmp, nres * 12, stream);
        if (e != cudaSuccess) {
            set_error("b2_index_search: result allocation failed: %s", cudaGetErrorString(e));
            rc = B2_ERR_OOM;
        } else {
            I_dev = static_cast<int64_t*>(out_tmp);
            D_dev = reinterpret_cast<float*>(I_dev + nres);
        }
    }
    if (rc == B2_OK) rc = search_device(h, nq, q_dev, k, D_dev, I_dev, stream);
    if (rc == B2_OK && out_where == B2_HOST) {
        cudaError_t e;
        if (bounce != nullptr) {
            e = cudaMemcpyAsync(bounce + q_pad, I_dev, nres * 12, cudaMemcpyDeviceToHost, stream);  // (I | D) are contiguous
        } else {
            e = cudaMemcpyAsync(I, I_dev, nres * 8, cudaMemcpyDeviceToHost, stream);
            if (e == cudaSuccess) e = cudaMemcpyAsync(D, D_dev, nres * 4, cudaMemcpyDeviceToHost, stream);
        }
        if (e != cudaSuccess) {
            set_error("b2_index_search: result copy failed: %s", cudaGetErrorString(e));
            rc = B2_ERR_CUDA;
        }
    }
    if (q_tmp) cudaFreeAsync(q_tmp, stream);
    if (out_tmp) cudaFreeAsync(out_tmp, stream);
    if (rc == B2_OK && out_where == B2_HOST) {
        cudaError_t e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) {
            set_error("b2_index_search: %s", cudaGetErrorString(e));
            rc = B2_ERR_CUDA;
        } else if (bounce != nullptr) {
            memcpy(I, bounce + q_pad, nres * 8);
            memcpy(D, bounce + q_pad + nres * 8, nres * 4);
        }
    }
    return rc;
}

static int check_filter_args(const char* who, int64_t nq, int k, int fetch_k, int relevance_kind) {
    if (nq < 0 || k <= 0 || fetch_k < k) {
        set_error("%s: bad arguments (nq=%lld, k=%d, fetch_k=%d; fetch_k must be >= k)", who, (long long)nq, k, fetch_k);
        return B2_ERR_INVALID;
    }
    if (relevance_kind < B2_REL_NONE || relevance_kind > B2_REL_MAX_INNER_PRODUCT) {
        set_error("%s: unknown relevance function %d", who, relevance_kind);
        return B2_ERR_INVALID;
    }
    if (fetch_k > B2_MAX_K) {
        set_error("%s: fetch_k=%d exceeds B2_MAX_K=%d", who, fetch_k, B2_MAX_K);
        return B2_ERR_UNSUPPORTED;
    }
    return B2_OK;
}

int b2_index_search_filtered(b2_index* h, int64_t nq, const float* q, int q_where, int k, int fetch_k,
                             const uint32_t* allow, int64_t allow_bits, int allow_where, int use_raw_threshold,
                             double raw_threshold, int relevance_kind, double relevance_threshold, float* D, int64_t* I,
                             int32_t* counts, int out_where, void* stream_) {
    B2_TRY(check_handle(h));
    B2_TRY(check_filter_args("b2_index_search_filtered", nq, k, fetch_k, relevance_kind));
    if (nq > 0 && (q == nullptr || D == nullptr || I == nullptr)) {
        set_error("b2_index_search_filtered: null pointer");
        return B2_ERR_INVALID;
    }
    if (allow != nullptr && allow_bits < 0) {
        set_error("b2_index_search_filtered: allow_bits=%lld", (long long)allow_bits);
        return B2_ERR_INVALID;
    }
    if (nq == 0) return B2_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    std::shared_lock<std::shared_mutex> lk(h->mu);
    DeviceGuard g(h->device);
    if (!g.ok) {
        set_error("cannot select device %d", h->device);
        return B2_ERR_CUDA;
    }
    // one block: [queries if host][fetch_k results I, D][k results I, D, counts if host][allow words if host]
    auto al = [](size_t v) { return (v + 255) / 256 * 256; };
    const size_t q_bytes = q_where == B2_HOST ? al((size_t)nq * h->d * sizeof(float)) : 0;
    const size_t nfetch = (size_t)nq * fetch_k, nres = (size_t)nq * k;
    const size_t fi_bytes = al(nfetch * 8), fd_bytes = al(nfetch * 4);
    const size_t oi_bytes = out_where == B2_HOST ? al(nres * 8) : 0, od_bytes = out_where == B2_HOST ? al(nres * 4) : 0;
    const size_t oc_bytes = out_where == B2_HOST ? al((size_t)nq * 4) : 0;
    const size_t allow_words = allow != nullptr ? (size_t)((allow_bits + 31) / 32) : 0;
    const size_t aw_bytes = (allow != nullptr && allow_where == B2_HOST) ? al(allow_words * 4 + 4) : 0;
    char* blk = nullptr;
    B2_CUDA_TRY(cudaMallocAsync(&blk, q_bytes + fi_bytes + fd_bytes + oi_bytes + od_bytes + oc_bytes + aw_bytes + 256, stream));
    char* cur = blk;
    float* q_tmp = reinterpret_cast<float*>(cur);
    cur += q_bytes;
    int64_t* FI = reinterpret_cast<int64_t*>(cur);
    cur += fi_bytes;
    float* FD = reinterpret_cast<float*>(cur);
    cur += fd_bytes;
    int64_t* OI = reinterpret_cast<int64_t*>(cur);
    cur += oi_bytes;
    float* OD = reinterpret_cast<float*>(cur);
    cur += od_bytes;
    int32_t* OC = reinterpret_cast<int32_t*>(cur);
    cur += oc_bytes;
    uint32_t* AW = reinterpret_cast<uint32_t*>(cur);
    int rc = B2_OK;
    cudaError_t e = cudaSuccess;
    const float* q_dev = q;
    if (q_where == B2_HOST) {
        e = cudaMemcpyAsync(q_tmp, q, (size_t)nq * h->d * sizeof(float), cudaMemcpyHostToDevice, stream);
        q_dev = q_tmp;
    }
    const uint32_t* allow_dev = allow;
    if (e == cudaSuccess && aw_bytes != 0) {
        e = cudaMemcpyAsync(AW, allow, allow_words * 4, cudaMemcpyHostToDevice, stream);
        allow_dev = AW;
    }
    if (e != cudaSuccess) {
        set_error("b2_index_search_filtered: input copy failed: %s", cudaGetErrorString(e));
        rc = B2_ERR_CUDA;
    }
    if (rc == B2_OK) rc = search_device(h, nq, q_dev, fetch_k, FD, FI, stream);
    float* D_dev = out_where == B2_HOST ? OD : D;
    int64_t* I_dev = out_where == B2_HOST ? OI : I;
    int32_t* C_dev = out_where == B2_HOST ? OC : counts;
    if (rc == B2_OK)
        rc = filter_results(FD, FI, nq, fetch_k, k, h->metric, allow_dev, allow_bits, use_raw_threshold, raw_threshold,
                            relevance_kind, relevance_threshold, D_dev, I_dev, C_dev, stream);
    if (rc == B2_OK && out_where == B2_HOST) {
        e = cudaMemcpyAsync(I, OI, nres * 8, cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(D, OD, nres * 4, cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess && counts != nullptr) e = cudaMemcpyAsync(counts, OC, (size_t)nq * 4, cudaMemcpyDeviceToHost, stream);
        if (e != cudaSuccess) {
            set_error("b2_index_search_filtered: result copy failed: %s", cudaGetErrorString(e));
            rc = B2_ERR_CUDA;
        }
    }
    cudaFreeAsync(blk, stream);
    if (rc == B2_OK && out_where == B2_HOST) {
        e = cudaStreamSynchronize(stream);
        if (e != cudaSuccess) {
            set_error("b2_index_search_filtered: %s", cudaGetErrorString(e));
            rc = B2_ERR_CUDA;
        }
    }
    return rc;
}

int b2_filter_results(int metric, int64_t nq, int fetch_k, int k, const float* D_in, const int64_t* I_in,
                      const uint32_t* allow, int64_t allow_bits, int use_raw_threshold, double raw_threshold,
                      int relevance_kind, double relevance_threshold, float* D_out, int64_t* I_out, int32_t* counts,
                      int device, void* stream_) {
    B2_TRY(check_filter_args("b2_filter_results", nq, k, fetch_k, relevance_kind));
    if (metric != B2_METRIC_IP && metric != B2_METRIC_L2) {
        set_error("b2_filter_results: unknown metric %d", metric);
        return B2_ERR_INVALID;
    }
    if (nq > 0 && (D_in == nullptr || I_in == nullptr || D_out == nullptr || I_out == nullptr)) {
        set_error("b2_filter_results: null pointer");
        return B2_ERR_INVALID;
    }
    if (nq == 0) return B2_OK;
    DeviceGuard g(device);
    if (!g.ok) {
        set_error("cannot select device %d", device);
        return B2_ERR_CUDA;
    }
    return filter_results(D_in, I_in, nq, fetch_k, k, metric, allow, allow_bits, use_raw_threshold, raw_threshold,
                          relevance_kind, relevance_threshold, D_out, I_out, counts, static_cast<cudaStream_t>(stream_));
}

int b2_index_reconstruct_batch(b2_index* h, int64_t n, const int64_t* ids, int ids_where, float* out, int out_where,
                               void* stream_) {
    B2_TRY(check_handle(h));
    if (n < 0 || (n > 0 && (ids == nullptr || out == nullptr))) {
        set_error("b2_index_reconstruct_batch: bad arguments");
        return B2_ERR_INVALID;
    }
    if (n == 0) return B2_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    std::shared_lock<std::shared_mutex> lk(h->mu);
    DeviceGuard g(h->device);
    const int64_t* ids_dev = ids;
    int64_t* ids_tmp = nullptr;
    float* out_dev = out;
    float* out_tmp = nullptr;
    if (ids_where == B2_HOST) {
        B2_CUDA_TRY(cudaMallocAsync(&ids_tmp, (size_t)n * 8, stream));
        B2_CUDA_TRY(cudaMemcpyAsync(ids_tmp, ids, (size_t)n * 8, cudaMemcpyHostToDevice, stream));
        ids_dev = ids_tmp;
    }
    if (out_where == B2_HOST) {
        B2_CUDA_TRY(cudaMallocAsync(&out_tmp, (size_t)n * h->d * 4, stream));
        out_dev = out_tmp;
    }
    int rc = gather_rows(h->store.base(), h->ntotal, h->pitch, h->d, h->dtype, ids_dev, n, out_dev, stream, resid_base(h));
    if (rc == B2_OK && out_where == B2_HOST) {
        cudaError_t e = cudaMemcpyAsync(out, out_dev, (size_t)n * h->d * 4, cudaMemcpyDeviceToHost, stream);
        if (e != cudaSuccess) {
            set_error("reconstruct: %s", cudaGetErrorString(e));
            rc = B2_ERR_CUDA;
        }
    }
    if (ids_tmp) cudaFreeAsync(ids_tmp, stream);
    if (out_tmp) cudaFreeAsync(out_tmp, stream);
    if (rc == B2_OK && out_where == B2_HOST) B2_CUDA_TRY(cudaStreamSynchronize(stream));
    return rc;
}

int b2_index_remove_ids(b2_index* h, int64_t n, const int64_t* ids, int64_t* n_removed) {
    B2_TRY(check_handle(h));
    if (n_removed) *n_removed = 0;
    if (n < 0 || (n > 0 && ids == nullptr)) {
        set_error("b2_index_remove_ids: bad arguments");
        return B2_ERR_INVALID;
    }
    if (n == 0) return B2_OK;
    std::unique_lock<std::shared_mutex> lk(h->mu);  // ntotal is only read under the writer lock: an add may be running
    if (h->ntotal == 0) return B2_OK;
    DeviceGuard g(h->device);
    const int64_t nt = h->ntotal;
    // the keep list is built on the device (b2_ingest.cu: build_keep_list): only the ids cross PCIe
    cudaStream_t stream = nullptr;
    int64_t removed = 0, first = nt;
    int64_t* ids_dev = nullptr;
    int64_t* src_dev = nullptr;
    void* scratch = nullptr;
    const size_t sbytes = keep_list_scratch_bytes(nt);
    int rc = B2_OK;
    auto cu = [&](cudaError_t e) {
        if (e != cudaSuccess && rc == B2_OK) {
            set_error("remove_ids: %s", cudaGetErrorString(e));
            rc = e == cudaErrorMemoryAllocation ? B2_ERR_OOM : B2_ERR_CUDA;
        }
    };
    cu(cudaMallocAsync(&ids_dev, (size_t)n * 8, stream));
    if (rc == B2_OK) cu(cudaMallocAsync(&src_dev, (size_t)nt * 8, stream));
    if (rc == B2_OK) cu(cudaMallocAsync(&scratch, sbytes, stream));
    if (rc == B2_OK) cu(cudaMemcpyAsync(ids_dev, ids, (size_t)n * 8, cudaMemcpyHostToDevice, stream));
    if (rc == B2_OK) rc = build_keep_list(ids_dev, n, nt, src_dev, scratch, sbytes, &removed, &first, stream);
    if (rc == B2_OK && removed > 0) {
        const int64_t n_dst = nt - first - removed;  // rows behind the first dropped one that stay
        if (n_dst > 0) {
            rc = compact_rows(h->store.base(), h->pitch, h->norms, src_dev, n_dst, first, stream);
            if (rc == B2_OK && h->residual_enabled)
                rc = compact_rows(h->resid.base(), h->pitch, nullptr, src_dev, n_dst, first, stream);
        }
    }
    if (ids_dev) cudaFreeAsync(ids_dev, stream);
    if (src_dev) cudaFreeAsync(src_dev, stream);
    if (scratch) cudaFreeAsync(scratch, stream);
    B2_CUDA_TRY(cudaStreamSynchronize(stream));
    if (rc != B2_OK) return rc;
    if (removed == 0) return B2_OK;
    h->ntotal = nt - removed;
    h->shadow_rows = std::min(h->shadow_rows, first);  // rows from the first removed one on moved: reconvert lazily
    h->shadow_refused_rows = 0;
    if (n_removed) *n_removed = removed;
    return B2_OK;
}

int b2_index_merge_from(b2_index* dst, b2_index* src) {
    B2_TRY(check_handle(dst));
    B2_TRY(check_handle(src));
    if (dst == src) {
        set_error("b2_index_merge_from: cannot merge an index into itself");
        return B2_ERR_INVALID;
    }
    if (dst->d != src->d || dst->dtype != src->dtype || dst->metric != src->metric ||
        dst->residual_enabled != src->residual_enabled) {
        set_error("b2_index_merge_from: indexes are not compatible (d %d vs %d, dtype %d vs %d, metric %d vs %d, residual "
                  "plane %d vs %d)", dst->d, src->d, dst->dtype, src->dtype, dst->metric, src->metric,
                  dst->residual_enabled, src->residual_enabled);
        return B2_ERR_INVALID;
    }
    // both writer locks, taken together without an order: merge(a, b) next to merge(b, a) must not deadlock
    std::unique_lock<std::shared_mutex> l1(dst->mu, std::defer_lock);
    std::unique_lock<std::shared_mutex> l2(src->mu, std::defer_lock);
    std::lock(l1, l2);
    const int64_t n = src->ntotal;
    if (n == 0) return B2_OK;
    if (dst->ntotal + n >= (int64_t)0xfffffff0ll) {
        set_error("b2_index_merge_from: more than 2^32-16 rows per index are not supported");
        return B2_ERR_UNSUPPORTED;
    }
    DeviceGuard g(dst->device);
    cudaStream_t stream = nullptr;
    B2_TRY(dst->store.ensure((size_t)(dst->ntotal + n) * dst->pitch, stream));
    B2_TRY(ensure_norms(dst, dst->ntotal + n, stream));
    if (dst->residual_enabled) {
        B2_TRY(dst->resid.ensure((size_t)(dst->ntotal + n) * dst->pitch, stream));
        char* rto = dst->resid.base() + (size_t)dst->ntotal * dst->pitch;
        if (dst->device == src->device)
            B2_CUDA_TRY(cudaMemcpyAsync(rto, src->resid.base(), (size_t)n * dst->pitch, cudaMemcpyDeviceToDevice, stream));
        else
            B2_CUDA_TRY(cudaMemcpyPeerAsync(rto, dst->device, src->resid.base(), src->device, (size_t)n * dst->pitch, stream));
        // the destination's error bound must cover the rows it receives
        uint32_t e_src = 0, e_dst = 0;
        {
            DeviceGuard gs(src->device);
            B2_CUDA_TRY(cudaMemcpy(&e_src, src->resid_err_ord, 4, cudaMemcpyDeviceToHost));
        }
        B2_CUDA_TRY(cudaMemcpy(&e_dst, dst->resid_err_ord, 4, cudaMemcpyDeviceToHost));
        if (e_src > e_dst) B2_CUDA_TRY(cudaMemcpy(dst->resid_err_ord, &e_src, 4, cudaMemcpyHostToDevice));
    }
    char* to = dst->store.base() + (size_t)dst->ntotal * dst->pitch;
    if (dst->device == src->device) {
        B2_CUDA_TRY(cudaMemcpyAsync(to, src->store.base(), (size_t)n * dst->pitch, cudaMemcpyDeviceToDevice, stream));
        B2_CUDA_TRY(cudaMemcpyAsync(dst->norms + dst->ntotal, src->norms, (size_t)n * 4, cudaMemcpyDeviceToDevice, stream));
    } else {
        B2_CUDA_TRY(cudaMemcpyPeerAsync(to, dst->device, src->store.base(), src->device, (size_t)n * dst->pitch, stream));
        B2_CUDA_TRY(cudaMemcpyPeerAsync(dst->norms + dst->ntotal, dst->device, src->norms, src->device, (size_t)n * 4, stream));
    }
    B2_TRY(update_max_norm(dst->norms + dst->ntotal, n, dst->max_norm_ord, stream));
    B2_CUDA_TRY(cudaStreamSynchronize(stream));
    dst->ntotal += n;
    src->ntotal = 0;  // faiss: other.reset()
    src->shadow_rows = 0;
    dst->shadow_refused_rows = src->shadow_refused_rows = 0;
    return B2_OK;
}

int b2_index_set_option(b2_index* h, const char* name, int64_t value) {
    B2_TRY(check_handle(h));
    if (name == nullptr) {
        set_error("b2_index_set_option: null option name");
        return B2_ERR_INVALID;
    }
    std::unique_lock<std::shared_mutex> lk(h->mu);
    if (strcmp(name, "gemm_query_parts") == 0) {
        if (value != 0 && value != 1 && value != 2) {
            set_error("b2_index_set_option: gemm_query_parts must be 0 (default), 1 or 2");
            return B2_ERR_INVALID;
        }
        h->gemm_query_parts = (int)value;
        return B2_OK;
    }
    if (strcmp(name, "bf16_residual") == 0) {
        if (value != 0 && value != 1) {
            set_error("b2_index_set_option: bf16_residual must be 0 or 1");
            return B2_ERR_INVALID;
        }
        if ((int)value == h->residual_enabled) return B2_OK;
        if (h->dtype != B2_DTYPE_BF16 || h->ntotal != 0) {
            set_error("b2_index_set_option: bf16_residual can only be changed on an EMPTY bf16 index (dtype %d, %lld rows)",
                      h->dtype, (long long)h->ntotal);
            return B2_ERR_INVALID;
        }
        h->residual_enabled = (int)value;
        return B2_OK;
    }
    if (strcmp(name, "bf16_shadow") == 0) {
        if (value != 0 && value != 1) {
            set_error("b2_index_set_option: bf16_shadow must be 0 or 1");
            return B2_ERR_INVALID;
        }
        if (value == 1 && (h->dtype != B2_DTYPE_F32 || !h->shadow.vmm)) {
            if (h->dtype != B2_DTYPE_F32) return B2_OK;  // bf16 rows are their own shadow
            set_error("b2_index_set_option: bf16_shadow needs the virtual-memory row store (B2_NO_VMM / B2_NO_SHADOW is set)");
            return B2_ERR_UNSUPPORTED;
        }
        h->shadow_enabled = (int)value;
        if (value == 0) {  // give the memory back (no search is running: this is the writer lock)
            DeviceGuard g(h->device);
            cudaDeviceSynchronize();
            const int dev = h->device;
            h->shadow.release();
            h->shadow.init(dev);
            h->shadow_rows = 0;
        }
        h->shadow_refused_rows = 0;
        return B2_OK;
    }
    set_error("b2_index_set_option: unknown option '%s'", name);
    return B2_ERR_INVALID;
}

int b2_index_reset(b2_index* h) {
    B2_TRY(check_handle(h));
    std::unique_lock<std::shared_mutex> lk(h->mu);
    h->ntotal = 0;
    h->shadow_rows = 0;
    h->shadow_refused_rows = 0;
    return B2_OK;
}

int b2_index_mmr(b2_index* h, int64_t nq, const float* q, int q_where, int k, int fetch_k, float lambda_mult,
                 int64_t* out_ids, float* out_scores, int out_where, void* stream_) {
    B2_TRY(check_handle(h));
    if (nq < 0 || k <= 0 || fetch_k <= 0 || (nq > 0 && (q == nullptr || out_ids == nullptr || out_scores == nullptr))) {
        set_error("b2_index_mmr: bad arguments (nq=%lld k=%d fetch_k=%d)", (long long)nq, k, fetch_k);
        return B2_ERR_INVALID;
    }
    if (fetch_k > B2_MAX_K) {
        set_error("b2_index_mmr: fetch_k=%d exceeds B2_MAX_K=%d", fetch_k, B2_MAX_K);
        return B2_ERR_UNSUPPORTED;
    }
    if (nq == 0) return B2_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    std::shared_lock<std::shared_mutex> lk(h->mu);
    DeviceGuard g(h->device);
    const size_t nf = (size_t)nq * fetch_k, nk = (size_t)nq * k;
    // one stream-ordered scratch block: [q?][I fetch][D fetch][rows][out ids][out scores]
    const size_t q_bytes = (q_where == B2_HOST) ? (size_t)nq * h->d * 4 : 0;
    const size_t rows_bytes = nf * h->d * 4;
    const size_t out_bytes = (out_where == B2_HOST) ? nk * 12 : 0;
    auto al = [](size_t v) { return (v + 255) / 256 * 256; };
    const size_t total = al(q_bytes) + al(nf * 8) + al(nf * 4) + al(rows_bytes) + al(out_bytes) + 256;
    char* blk = nullptr;
    B2_CUDA_TRY(cudaMallocAsync(&blk, total, stream));
    char* p = blk;
    float* q_dev = const_cast<float*>(q);
    if (q_where == B2_HOST) {
        q_dev = reinterpret_cast<float*>(p);
        p += al(q_bytes);
    }
    int64_t* If = reinterpret_cast<int64_t*>(p);
    p += al(nf * 8);
    float* Df = reinterpret_cast<float*>(p);
    p += al(nf * 4);
    float* rows = reinterpret_cast<float*>(p);
    p += al(rows_bytes);
    int64_t* oi = out_ids;
    float* os = out_scores;
    if (out_where == B2_HOST) {
        oi = reinterpret_cast<int64_t*>(p);
        os = reinterpret_cast<float*>(oi + nk);
    }
    int rc = B2_OK;
    cudaError_t e = cudaSuccess;
    if (q_where == B2_HOST) e = cudaMemcpyAsync(q_dev, q, q_bytes, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) {
        set_error("b2_index_mmr: %s", cudaGetErrorString(e));
        rc = B2_ERR_CUDA;
    }
    if (rc == B2_OK) rc = search_device(h, nq, q_dev, fetch_k, Df, If, stream);
    if (rc == B2_OK) rc = gather_rows(h->store.base(), h->ntotal, h->pitch, h->d, h->dtype, If, (int64_t)nf, rows, stream, resid_base(h));
    if (rc == B2_OK) rc = mmr_select(q_dev, h->d, (int)nq, rows, If, Df, fetch_k, k, lambda_mult, oi, os, stream);
    if (rc == B2_OK && out_where == B2_HOST) {
        e = cudaMemcpyAsync(out_ids, oi, nk * 8, cudaMemcpyDeviceToHost, stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(out_scores, os, nk * 4, cudaMemcpyDeviceToHost, stream);
        if (e != cudaSuccess) {
            set_error("b2_index_mmr: %s", cudaGetErrorString(e));
            rc = B2_ERR_CUDA;
        }
    }
    cudaFreeAsync(blk, stream);
    if (rc == B2_OK && out_where == B2_HOST) B2_CUDA_TRY(cudaStreamSynchronize(stream));
    return rc;
}

int b2_merge_shard_lists(int metric, int64_t nq, int k, int nparts, const float* D_parts, const int64_t* I_parts,
                         float* D, int64_t* I, int device, void* stream_) {
    if (nq < 0 || k <= 0 || nparts <= 0 || (nq > 0 && (!D_parts || !I_parts || !D || !I))) {
        set_error("b2_merge_shard_lists: bad arguments");
        return B2_ERR_INVALID;
    }
    if (nq == 0) return B2_OK;
    DeviceGuard g(device);
    if (!g.ok) {
        set_error("cannot select device %d", device);
        return B2_ERR_CUDA;
    }
    return merge_shards(metric, (int)nq, k, nparts, D_parts, I_parts, D, I, static_cast<cudaStream_t>(stream_));
}

int b2_index_search_push(b2_index* h, int64_t nq, const float* q_dev, int k, int nparts, void* const* peer_blocks,
                         uint32_t* const* peer_flags, uint32_t epoch, void* stream_) {
    B2_TRY(check_handle(h));
    if (nq <= 0 || k <= 0 || q_dev == nullptr || nparts <= 0 || peer_blocks == nullptr || peer_flags == nullptr) {
        set_error("b2_index_search_push: bad arguments (nq=%lld, k=%d, nparts=%d)", (long long)nq, k, nparts);
        return B2_ERR_INVALID;
    }
    if (nparts > kMaxPushPeers || k > B2_MAX_K) {
        set_error("b2_index_search_push: at most %d peers and k <= %d", kMaxPushPeers, B2_MAX_K);
        return B2_ERR_UNSUPPORTED;
    }
    PeerPush push;
    push.nparts = nparts;
    push.epoch = epoch;
    push.nq_total = nq;
    push.q_offset = 0;
    for (int g = 0; g < kMaxPushPeers; ++g) {
        push.block[g] = g < nparts ? static_cast<char*>(peer_blocks[g]) : nullptr;
        push.flag[g] = g < nparts ? peer_flags[g] : nullptr;
        if (g < nparts && (push.block[g] == nullptr || push.flag[g] == nullptr)) {
            set_error("b2_index_search_push: null peer pointer for part %d", g);
            return B2_ERR_INVALID;
        }
    }
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    std::shared_lock<std::shared_mutex> lk(h->mu);
    DeviceGuard g(h->device);
    if (!g.ok) {
        set_error("cannot select device %d", h->device);
        return B2_ERR_CUDA;
    }
    return search_device(h, nq, q_dev, k, nullptr, nullptr, stream, &push);
}

int b2_query_push(const float* q_dev, int64_t nbytes, int nparts, void* const* peer_bufs, uint32_t* const* peer_flags,
                  uint32_t epoch, int device, void* stream_) {
    if (q_dev == nullptr || nbytes <= 0 || nbytes % 16 != 0 || nparts <= 0 || peer_bufs == nullptr || peer_flags == nullptr ||
        (reinterpret_cast<uintptr_t>(q_dev) & 15) != 0) {
        set_error("b2_query_push: bad arguments (nbytes=%lld must be a positive multiple of 16, 16-byte aligned source)",
                  (long long)nbytes);
        return B2_ERR_INVALID;
    }
    if (nparts > kMaxPushPeers || nbytes > B2_QUERY_PUSH_MAX_BYTES) {
        set_error("b2_query_push: at most %d peers and %d bytes", kMaxPushPeers, B2_QUERY_PUSH_MAX_BYTES);
        return B2_ERR_UNSUPPORTED;
    }
    for (int g = 0; g < nparts; ++g)
        if (peer_bufs[g] == nullptr || peer_flags[g] == nullptr) {
            set_error("b2_query_push: null peer pointer for part %d", g);
            return B2_ERR_INVALID;
        }
    DeviceGuard g(device);
    if (!g.ok) {
        set_error("cannot select device %d", device);
        return B2_ERR_CUDA;
    }
    return query_push(q_dev, (size_t)nbytes, nparts, peer_bufs, peer_flags, epoch, static_cast<cudaStream_t>(stream_));
}

int b2_flag_wait(const uint32_t* flag, uint32_t epoch, int device, void* stream_) {
    if (flag == nullptr) {
        set_error("b2_flag_wait: null flag");
        return B2_ERR_INVALID;
    }
    DeviceGuard g(device);
    if (!g.ok) {
        set_error("cannot select device %d", device);
        return B2_ERR_CUDA;
    }
    return flag_wait(flag, epoch, static_cast<cudaStream_t>(stream_));
}

int b2_merge_shard_lists_wait(int metric, int64_t nq, int k, int nparts, const void* blocks, int64_t part_stride_bytes,
                              const uint32_t* flags, int64_t flag_stride, uint32_t epoch, const int64_t* id_bases,
                              float* D, int64_t* I, int device, void* stream_) {
    if (nq <= 0 || k <= 0 || nparts <= 0 || nparts > B2_MAX_SHARDS || part_stride_bytes <= 0 || flag_stride < nq ||
        !blocks || !flags || !id_bases || !D || !I) {
        set_error("b2_merge_shard_lists_wait: bad arguments");
        return B2_ERR_INVALID;
    }
    DeviceGuard g(device);
    if (!g.ok) {
        set_error("cannot select device %d", device);
        return B2_ERR_CUDA;
    }
    return merge_shards_wait(metric, (int)nq, k, nparts, static_cast<const char*>(blocks), (size_t)part_stride_bytes,
                             flags, (size_t)flag_stride, epoch, id_bases, D, I, static_cast<cudaStream_t>(stream_));
}

int b2_merge_shard_lists_packed(int metric, int64_t nq, int k, int nparts, const float* D_parts, const int64_t* I_parts,
                                int64_t part_stride_bytes, const int64_t* id_bases, float* D, int64_t* I, int device,
                                void* stream_) {
    if (nq < 0 || k <= 0 || nparts <= 0 || nparts > B2_MAX_SHARDS || part_stride_bytes <= 0 ||
        (nq > 0 && (!D_parts || !I_parts || !D || !I || !id_bases))) {
        set_error("b2_merge_shard_lists_packed: bad arguments");
        return B2_ERR_INVALID;
    }
    if (nq == 0) return B2_OK;
    DeviceGuard g(device);
    if (!g.ok) {
        set_error("cannot select device %d", device);
        return B2_ERR_CUDA;
    }
    return merge_shards(metric, (int)nq, k, nparts, D_parts, I_parts, D, I, static_cast<cudaStream_t>(stream_),
                        (size_t)part_stride_bytes, (size_t)part_stride_bytes, id_bases);
}

int b2_normalize_l2(float* x, int64_t n, int d, int device, void* stream_) {
    if (n < 0 || d <= 0 || (n > 0 && x == nullptr)) {
        set_error("b2_normalize_l2: bad arguments");
        return B2_ERR_INVALID;
    }
    DeviceGuard g(device);
    if (!g.ok) {
        set_error("cannot select device %d", device);
        return B2_ERR_CUDA;
    }
    return normalize_rows_inplace(x, n, d, static_cast<cudaStream_t>(stream_));
}

int b2_synth_fill(float* out, int64_t row0, int64_t nrows, int d, uint64_t seed, int device, void* stream_) {
    if (nrows < 0 || d <= 0 || (nrows > 0 && out == nullptr)) {
        set_error("b2_synth_fill: bad arguments");
        return B2_ERR_INVALID;
    }
    DeviceGuard g(device);
    if (!g.ok) {
        set_error("cannot select device %d", device);
        return B2_ERR_CUDA;
    }
    return synth_fill(out, row0, nrows, d, seed, static_cast<cudaStream_t>(stream_));
}

}  // extern "C"
