// Row-sharded flat index over several GPUs of ONE process, behind the C ABI (SURVEY.md 8b:
// b2_shard_group_create / b2_shard_search; 8e: contiguous row blocks, queries made visible on every GPU, per-shard
// top-k lists exchanged and merged on the device).
//
// The reference has no multi-GPU code; what this replaces is what a maintainer would otherwise build from
// faiss.IndexShards / index_cpu_to_all_gpus around the flat index of
// /root/reference/RetrievalAugmentedGeneration/common/utils.py:217.  The group is a composition of the single-GPU
// entry points of include/b2_flat.h: every shard is a b2_index on its own device; a search is
//   root:    b2_query_push            the query block -> every shard's receive buffer (peer stores)
//   shard g: b2_index_search_push with ONE destination: the shard's receive block on the root (the shard's merge
//            kernel stores its final lists there over NVLink)
//   root:    b2_merge_shard_lists_wait over the G blocks, adding the shard bases -> global (D, I)
// i.e. the data path nv_genai_b200/sharded.py runs with one process per GPU, here with one host thread, G streams and
// peer access between the devices.  No collective, no host synchronisation inside a search.  Ordering between the
// root's stream and the shards' streams is by CUDA events, not by the flag-spinning kernels of the multi-process
// form: shards may share a device with each other and with the root's merge, and a kernel that spins for a result
// can keep the kernel that produces it off the SMs (the flags are still written, the merge finds them set).
//
// Rows keep faiss's implicit ids (global id = insertion position): shard g holds the contiguous id range
// [base_g, base_g + ntotal_g) and is filled up to `rows_per_shard` before shard g + 1 receives rows (the last shard
// takes whatever follows), so shard order is id order and the merge's lowest-id tie rule holds across shards.
#include <algorithm>
#include <cstring>
#include <mutex>
#include <vector>

#include "b2_common.cuh"
#include "b2_kernels.h"

using namespace b2;

namespace {

constexpr int64_t kMaxResults = 1 << 17;  // nq * k per pass (12 B each -> 1.5 MB per shard block)
constexpr int64_t kMaxNq = 4096;          // flag words per shard
constexpr size_t kBlockBytes = (size_t)kMaxResults * 12;

struct DevGuard {
    int prev = -1;
    bool ok = true;
    explicit DevGuard(int dev) {
        if (cudaGetDevice(&prev) != cudaSuccess) {
            ok = false;
            prev = -1;
            return;
        }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    }
    ~DevGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

}  // namespace

struct b2_shard_group {
    int n = 0, d = 0, metric = 0, dtype = 0;
    int64_t rows_per_shard = 0;
    std::vector<int> dev;
    std::vector<b2_index*> shard;
    std::vector<cudaStream_t> stream;     // one per shard, on its device
    // on the root device (device of shard 0): two alternating sets of G receive blocks + G x kMaxNq flag words
    char* root_blocks = nullptr;          // [2][G][kBlockBytes]
    uint32_t* root_flags = nullptr;       // [2][G][kMaxNq]
    float* root_q = nullptr;              // staging for host queries / the root's copy (B2_QUERY_PUSH_MAX_BYTES)
    float* D_tmp = nullptr;               // staging for host outputs
    // on every shard's device: two query receive buffers + two flag words
    std::vector<char*> qbuf;              // [2][B2_QUERY_PUSH_MAX_BYTES]
    std::vector<uint32_t*> qflag;         // [2] (one 128-byte line each)
    cudaEvent_t done = nullptr;           // root: completion of the previous search's merge
    cudaEvent_t pushed = nullptr;         // root: the query block of the current pass is in every shard's buffer
    std::vector<cudaEvent_t> searched;    // per shard (on its device): its lists of the current pass are on the root
    uint32_t epoch = 0;
    std::mutex mu;                        // searches and mutations of one group are enqueued one at a time
};

static int shard_check(const b2_shard_group* g) {
    if (g == nullptr) {
        set_error("null shard group handle");
        return B2_ERR_INVALID;
    }
    return B2_OK;
}

extern "C" {

int b2_shard_group_create(int ndev, const int* devices, int d, int metric, int dtype, int64_t rows_per_shard,
                          b2_shard_group** out) {
    if (out == nullptr || devices == nullptr || ndev <= 0 || ndev > kMaxPushPeers || rows_per_shard <= 0) {
        set_error("b2_shard_group_create: bad arguments (ndev=%d must be 1..%d, rows_per_shard=%lld > 0)", ndev, kMaxPushPeers,
                  (long long)rows_per_shard);
        return B2_ERR_INVALID;
    }
    *out = nullptr;
    b2_shard_group* g = new b2_shard_group();
    g->n = ndev;
    g->d = d;
    g->metric = metric;
    g->dtype = dtype;
    g->rows_per_shard = rows_per_shard;
    g->dev.assign(devices, devices + ndev);
    g->shard.assign(ndev, nullptr);
    g->stream.assign(ndev, nullptr);
    g->qbuf.assign(ndev, nullptr);
    g->qflag.assign(ndev, nullptr);
    g->searched.assign(ndev, nullptr);
    int rc = B2_OK;
    auto fail = [&](int code) {
        b2_shard_group_destroy(g);
        return code;
    };
    for (int i = 0; i < ndev && rc == B2_OK; ++i) rc = b2_index_create(d, metric, dtype, devices[i], &g->shard[i]);
    if (rc != B2_OK) return fail(rc);
    const int root = devices[0];
    for (int i = 0; i < ndev; ++i) {
        DevGuard dg(devices[i]);
        if (!dg.ok) {
            set_error("b2_shard_group_create: cannot select device %d", devices[i]);
            return fail(B2_ERR_CUDA);
        }
        // peer access in both directions with every other device of the group (the lists go shard -> root, the
        // queries root -> shard; merge_from and friends may use any pair)
        for (int j = 0; j < ndev; ++j) {
            if (devices[j] == devices[i]) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, devices[i], devices[j]) != cudaSuccess || !can) {
                cudaGetLastError();
                set_error("b2_shard_group_create: device %d cannot access device %d (no peer access)", devices[i], devices[j]);
                return fail(B2_ERR_UNSUPPORTED);
            }
            const cudaError_t e = cudaDeviceEnablePeerAccess(devices[j], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                set_error("b2_shard_group_create: cudaDeviceEnablePeerAccess(%d -> %d): %s", devices[i], devices[j],
                          cudaGetErrorString(e));
                return fail(B2_ERR_CUDA);
            }
            cudaGetLastError();
        }
        cudaError_t e = cudaStreamCreateWithFlags(&g->stream[i], cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaMalloc(&g->qbuf[i], 2 * (size_t)B2_QUERY_PUSH_MAX_BYTES);
        if (e == cudaSuccess) e = cudaMalloc(&g->qflag[i], 2 * 128);
        if (e == cudaSuccess) e = cudaMemset(g->qflag[i], 0, 2 * 128);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g->searched[i], cudaEventDisableTiming);
        if (e != cudaSuccess) {
            set_error("b2_shard_group_create: device %d: %s", devices[i], cudaGetErrorString(e));
            return fail(e == cudaErrorMemoryAllocation ? B2_ERR_OOM : B2_ERR_CUDA);
        }
    }
    {
        DevGuard dg(root);
        cudaError_t e = cudaMalloc(&g->root_blocks, 2 * (size_t)ndev * kBlockBytes);
        if (e == cudaSuccess) e = cudaMalloc(&g->root_flags, 2 * (size_t)ndev * kMaxNq * 4);
        if (e == cudaSuccess) e = cudaMemset(g->root_flags, 0, 2 * (size_t)ndev * kMaxNq * 4);
        if (e == cudaSuccess) e = cudaMalloc(&g->root_q, (size_t)B2_QUERY_PUSH_MAX_BYTES);
        if (e == cudaSuccess) e = cudaMalloc(&g->D_tmp, kBlockBytes);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g->done, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g->pushed, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
            set_error("b2_shard_group_create: root device %d: %s", root, cudaGetErrorString(e));
            return fail(e == cudaErrorMemoryAllocation ? B2_ERR_OOM : B2_ERR_CUDA);
        }
    }
    *out = g;
    return B2_OK;
}

void b2_shard_group_destroy(b2_shard_group* g) {
    if (g == nullptr) return;
    for (int i = 0; i < g->n; ++i) {
        DevGuard dg(g->dev[i]);
        if (g->stream[i]) {
            cudaStreamSynchronize(g->stream[i]);
            cudaStreamDestroy(g->stream[i]);
        }
        if (g->qbuf[i]) cudaFree(g->qbuf[i]);
        if (g->qflag[i]) cudaFree(g->qflag[i]);
        if (g->searched[i]) cudaEventDestroy(g->searched[i]);
        if (g->shard[i]) b2_index_destroy(g->shard[i]);
    }
    if (g->n > 0) {
        DevGuard dg(g->dev[0]);
        if (g->root_blocks) cudaFree(g->root_blocks);
        if (g->root_flags) cudaFree(g->root_flags);
        if (g->root_q) cudaFree(g->root_q);
        if (g->D_tmp) cudaFree(g->D_tmp);
        if (g->done) cudaEventDestroy(g->done);
        if (g->pushed) cudaEventDestroy(g->pushed);
    }
    delete g;
}

int64_t b2_shard_group_ntotal(const b2_shard_group* g) {
    if (g == nullptr) return -1;
    int64_t t = 0;
    for (int i = 0; i < g->n; ++i) t += b2_index_ntotal(g->shard[i]);
    return t;
}

int b2_shard_group_nshards(const b2_shard_group* g) { return g ? g->n : -1; }

int64_t b2_shard_group_shard_ntotal(const b2_shard_group* g, int shard) {
    if (g == nullptr || shard < 0 || shard >= g->n) return -1;
    return b2_index_ntotal(g->shard[shard]);
}

int b2_shard_group_add(b2_shard_group* g, int64_t n, const float* x, int x_where, unsigned flags) {
    B2_TRY(shard_check(g));
    if (n < 0 || (n > 0 && x == nullptr)) {
        set_error("b2_shard_group_add: bad arguments (n=%lld)", (long long)n);
        return B2_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(g->mu);
    int64_t off = 0;
    for (int i = 0; i < g->n && off < n; ++i) {
        const int64_t have = b2_index_ntotal(g->shard[i]);
        // a shard takes rows only while every later shard is still empty (ids stay contiguous in shard order)
        bool later_empty = true;
        for (int j = i + 1; j < g->n; ++j) later_empty = later_empty && b2_index_ntotal(g->shard[j]) == 0;
        if (!later_empty) continue;
        int64_t room = (i == g->n - 1) ? (n - off) : std::max<int64_t>(0, g->rows_per_shard - have);
        const int64_t m = std::min(room, n - off);
        if (m <= 0) continue;
        const float* src = x + (size_t)off * g->d;
        int rc;
        if (x_where == B2_DEVICE && g->dev[i] != g->dev[0]) {
            // device rows live on the root device: bring this shard's slice over NVLink first
            DevGuard dg(g->dev[i]);
            float* tmp = nullptr;
            const size_t bytes = (size_t)m * g->d * sizeof(float);
            B2_CUDA_TRY(cudaMalloc(&tmp, bytes));
            cudaError_t e = cudaMemcpyPeer(tmp, g->dev[i], src, g->dev[0], bytes);
            if (e != cudaSuccess) {
                cudaFree(tmp);
                set_error("b2_shard_group_add: peer copy to device %d failed: %s", g->dev[i], cudaGetErrorString(e));
                return B2_ERR_CUDA;
            }
            rc = b2_index_add(g->shard[i], m, tmp, B2_DEVICE, flags, nullptr);
            cudaDeviceSynchronize();
            cudaFree(tmp);
        } else {
            rc = b2_index_add(g->shard[i], m, src, x_where, flags, nullptr);
            if (rc == B2_OK && x_where == B2_DEVICE) {
                DevGuard dg(g->dev[i]);
                cudaStreamSynchronize(nullptr);
            }
        }
        if (rc != B2_OK) return rc;
        off += m;
    }
    return B2_OK;
}

int b2_shard_group_set_option(b2_shard_group* g, const char* name, int64_t value) {
    B2_TRY(shard_check(g));
    std::lock_guard<std::mutex> lk(g->mu);
    for (int i = 0; i < g->n; ++i) B2_TRY(b2_index_set_option(g->shard[i], name, value));
    return B2_OK;
}

int b2_shard_group_reset(b2_shard_group* g) {
    B2_TRY(shard_check(g));
    std::lock_guard<std::mutex> lk(g->mu);
    for (int i = 0; i < g->n; ++i) B2_TRY(b2_index_reset(g->shard[i]));
    return B2_OK;
}

int b2_shard_group_remove_ids(b2_shard_group* g, int64_t n, const int64_t* ids, int64_t* n_removed) {
    B2_TRY(shard_check(g));
    if (n < 0 || (n > 0 && ids == nullptr)) {
        set_error("b2_shard_group_remove_ids: bad arguments");
        return B2_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(g->mu);
    int64_t total = 0, base = 0;
    for (int i = 0; i < g->n; ++i) {
        const int64_t cnt = b2_index_ntotal(g->shard[i]);
        std::vector<int64_t> local;
        for (int64_t j = 0; j < n; ++j)
            if (ids[j] >= base && ids[j] < base + cnt) local.push_back(ids[j] - base);
        if (!local.empty()) {
            int64_t r = 0;
            B2_TRY(b2_index_remove_ids(g->shard[i], (int64_t)local.size(), local.data(), &r));
            total += r;
        }
        base += cnt;  // bases are those BEFORE the removal: ids name rows of the store as the caller saw it
    }
    if (n_removed) *n_removed = total;
    return B2_OK;
}

int b2_shard_group_reconstruct_batch(b2_shard_group* g, int64_t n, const int64_t* ids, float* out) {
    B2_TRY(shard_check(g));
    if (n < 0 || (n > 0 && (ids == nullptr || out == nullptr))) {
        set_error("b2_shard_group_reconstruct_batch: bad arguments");
        return B2_ERR_INVALID;
    }
    std::lock_guard<std::mutex> lk(g->mu);
    int64_t base = 0;
    std::vector<bool> done((size_t)n, false);
    for (int i = 0; i < g->n; ++i) {
        const int64_t cnt = b2_index_ntotal(g->shard[i]);
        std::vector<int64_t> local, where;
        for (int64_t j = 0; j < n; ++j)
            if (ids[j] >= base && ids[j] < base + cnt) {
                local.push_back(ids[j] - base);
                where.push_back(j);
            }
        if (!local.empty()) {
            std::vector<float> rows(local.size() * (size_t)g->d);
            B2_TRY(b2_index_reconstruct_batch(g->shard[i], (int64_t)local.size(), local.data(), B2_HOST, rows.data(), B2_HOST,
                                              nullptr));
            for (size_t t = 0; t < local.size(); ++t) {
                memcpy(out + (size_t)where[t] * g->d, rows.data() + t * (size_t)g->d, (size_t)g->d * sizeof(float));
                done[(size_t)where[t]] = true;
            }
        }
        base += cnt;
    }
    for (int64_t j = 0; j < n; ++j)  // ids outside the store: zero rows, as b2_index_reconstruct_batch gives
        if (!done[(size_t)j]) memset(out + (size_t)j * g->d, 0, (size_t)g->d * sizeof(float));
    return B2_OK;
}

int b2_shard_search(b2_shard_group* g, int64_t nq, const float* q, int q_where, int k, float* D, int64_t* I, int out_where,
                    void* stream_) {
    B2_TRY(shard_check(g));
    if (nq < 0 || k <= 0 || (nq > 0 && (q == nullptr || D == nullptr || I == nullptr))) {
        set_error("b2_shard_search: bad arguments (nq=%lld, k=%d)", (long long)nq, k);
        return B2_ERR_INVALID;
    }
    if (k > B2_MAX_K) {
        set_error("b2_shard_search: k=%d exceeds B2_MAX_K=%d", k, B2_MAX_K);
        return B2_ERR_UNSUPPORTED;
    }
    if (nq == 0) return B2_OK;
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    const int G = g->n, root = g->dev[0];
    const size_t row_bytes = (size_t)g->d * sizeof(float);
    if (row_bytes % 16 != 0) {
        set_error("b2_shard_search: d must be a multiple of 4");
        return B2_ERR_UNSUPPORTED;
    }
    // queries per pass: bounded by the receive blocks, the flag words and the query buffers
    int64_t per_pass = std::min<int64_t>(kMaxNq, kMaxResults / k);
    per_pass = std::min<int64_t>(per_pass, (int64_t)(B2_QUERY_PUSH_MAX_BYTES / row_bytes));
    if (per_pass < 1) {
        set_error("b2_shard_search: d=%d is too large for the query buffers", g->d);
        return B2_ERR_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lk(g->mu);
    DevGuard dg(root);
    if (!dg.ok) {
        set_error("cannot select device %d", root);
        return B2_ERR_CUDA;
    }
    int64_t bases[kMaxPushPeers];
    int64_t acc = 0;
    for (int i = 0; i < G; ++i) {
        bases[i] = acc;
        acc += b2_index_ntotal(g->shard[i]);
    }
    // a search on another stream than the previous one must not overtake it: the two buffer sets are only safe to
    // reuse once the merge of the search before last has finished (see the file comment)
    B2_CUDA_TRY(cudaStreamWaitEvent(stream, g->done, 0));
    for (int64_t q0 = 0; q0 < nq; q0 += per_pass) {
        const int64_t m = std::min(per_pass, nq - q0);
        const size_t qbytes = (size_t)m * row_bytes;
        const float* qsrc = q + (size_t)q0 * g->d;
        if (q_where == B2_HOST) {
            // the previous pass's push kernel has read root_q: it ran on `stream` before this copy
            B2_CUDA_TRY(cudaMemcpyAsync(g->root_q, qsrc, qbytes, cudaMemcpyHostToDevice, stream));
            qsrc = g->root_q;
        }
        g->epoch = (g->epoch + 1) & 0xffffffffu;
        if (g->epoch == 0) g->epoch = 1;
        const uint32_t e = g->epoch;
        const int slot = (int)(e & 1u);
        void* bufs[kMaxPushPeers];
        uint32_t* qflags[kMaxPushPeers];
        for (int i = 0; i < G; ++i) {
            bufs[i] = g->qbuf[i] + (size_t)slot * B2_QUERY_PUSH_MAX_BYTES;
            qflags[i] = g->qflag[i] + slot * 32;
        }
        B2_TRY(b2_query_push(qsrc, (int64_t)qbytes, G, bufs, qflags, e, root, stream));
        B2_CUDA_TRY(cudaEventRecord(g->pushed, stream));
        char* blocks = g->root_blocks + (size_t)slot * G * kBlockBytes;
        uint32_t* flags = g->root_flags + (size_t)slot * G * kMaxNq;
        for (int i = 0; i < G; ++i) {
            void* blk = blocks + (size_t)i * kBlockBytes;
            uint32_t* fl = flags + (size_t)i * kMaxNq;
            {
                DevGuard sg(g->dev[i]);
                B2_CUDA_TRY(cudaStreamWaitEvent(g->stream[i], g->pushed, 0));
            }
            B2_TRY(b2_index_search_push(g->shard[i], m, static_cast<const float*>(bufs[i]), k, 1, &blk, &fl, e,
                                        g->stream[i]));
            {
                DevGuard sg(g->dev[i]);
                B2_CUDA_TRY(cudaEventRecord(g->searched[i], g->stream[i]));
            }
            B2_CUDA_TRY(cudaStreamWaitEvent(stream, g->searched[i], 0));
        }
        float* D_dev = D + (size_t)q0 * k;
        int64_t* I_dev = I + (size_t)q0 * k;
        if (out_where == B2_HOST) {
            I_dev = reinterpret_cast<int64_t*>(g->D_tmp);
            D_dev = reinterpret_cast<float*>(I_dev + (size_t)m * k);
        }
        B2_TRY(b2_merge_shard_lists_wait(g->metric, m, k, G, blocks, (int64_t)kBlockBytes, flags, kMaxNq, e, bases, D_dev, I_dev,
                                         root, stream));
        B2_CUDA_TRY(cudaEventRecord(g->done, stream));
        if (out_where == B2_HOST) {
            B2_CUDA_TRY(cudaMemcpyAsync(I + (size_t)q0 * k, I_dev, (size_t)m * k * 8, cudaMemcpyDeviceToHost, stream));
            B2_CUDA_TRY(cudaMemcpyAsync(D + (size_t)q0 * k, D_dev, (size_t)m * k * 4, cudaMemcpyDeviceToHost, stream));
            B2_CUDA_TRY(cudaStreamSynchronize(stream));  // D_tmp / root_q are reused by the next pass and the caller reads D, I
        }
    }
    return B2_OK;
}

}  // extern "C"
