/*
 * flat_oracle.c — CPU restatement of the flat vector-store search path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under nv_genai_b200/ may import, link or
 * execute this file; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker or the
 * timed CPU baseline.
 *
 * PARITY UNPINNED BY THE REFERENCE: the reference repository
 * (april-yyt/NV_GenAI) holds no implementation of this path and no test,
 * golden vector or fixture for it.  It only constructs and calls a
 * third-party index:
 *     RetrievalAugmentedGeneration/common/utils.py:217
 *         FAISS(document_embedder, IndexFlatL2(config.embeddings.dimensions), InMemoryDocstore(), {})
 *     examples/5_mins_rag_no_gpu/main.py:90
 *         FAISS.from_documents(documents, document_embedder)
 * The arithmetic lives in faiss-cpu==1.7.4
 * (RetrievalAugmentedGeneration/requirements.txt:25,
 *  examples/5_mins_rag_no_gpu/requirements.txt:2), which is not vendored under
 * /root/reference and is not installable here (no network, no wheel).  This
 * file restates the published faiss 1.7.4 IndexFlat algorithm
 * (faiss/IndexFlat.cpp, faiss/utils/distances.cpp, distances_simd.cpp,
 * utils/Heap.h, impl/ResultHandler.h) from its documented behaviour:
 *   - IndexFlatL2::search  -> squared L2, results ascending
 *   - IndexFlatIP::search  -> inner product, results descending
 *   - nq < 20 (distance_compute_blas_threshold): one query per OpenMP thread,
 *     sequential scan over all rows with an 8-lane SIMD accumulator, running
 *     top-k that admits a row only when strictly better than the current worst
 *   - fewer than k rows: I = -1, D = +FLT_MAX (L2) / -FLT_MAX (IP)
 *   - faiss.normalize_L2 -> fvec_renorm_L2: nr = sum x^2; if nr > 0 scale by 1/sqrtf(nr)
 * The one place it is pinned against reference-owned code is the inner-product
 * top-k SET, checked in tests/ against golden vectors generated from the
 * reference's own numpy statement of "IP scores + top-k"
 * (experimental/synthetic-data-retriever-customization/retriever_customization.ipynb,
 * cell 21: np.matmul(Q, P.T) ... np.argpartition(row, -k)[-k:]); see
 * tests/golden/make_golden.py.
 *
 * Tie rule (ours, stated in DESIGN.md): among exactly equal scores the lowest
 * id comes first.  For a sequential scan with strict-better admission this is
 * what faiss's heap produces too.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define B2O_METRIC_IP 0 /* faiss METRIC_INNER_PRODUCT = 0 */
#define B2O_METRIC_L2 1 /* faiss METRIC_L2 = 1 */

/* ------------------------------------------------------------------ */
/* Synthetic data: Philox-4x32-10, counter-based, integer-exact.       */
/* One call per 4 consecutive columns of one row:                      */
/*   counter = (row_lo, row_hi, col/4, 0), key = (seed_lo, seed_hi)    */
/* each 32-bit output word w -> ((w & 0xffff) + (w >> 16) - 65535) *   */
/* B2O_SYNTH_SCALE, a triangular variate of unit variance.  Only       */
/* integer ops and one exact int->float convert + one fp32 multiply,   */
/* so the CUDA generator reproduces these values bit for bit.          */
/* ------------------------------------------------------------------ */
#define PHILOX_M0 0xD2511F53u
#define PHILOX_M1 0xCD9E8D57u
#define PHILOX_W0 0x9E3779B9u
#define PHILOX_W1 0xBB67AE85u
#define B2O_SYNTH_SCALE (1.0f / 26754.58f) /* 65536/sqrt(6) */

static inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)PHILOX_M0 * c[0];
        uint64_t p1 = (uint64_t)PHILOX_M1 * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += PHILOX_W0; k1 += PHILOX_W1;
    }
}

static inline float synth_word(uint32_t w) {
    int32_t s = (int32_t)(w & 0xffffu) + (int32_t)(w >> 16) - 65535;
    return (float)s * B2O_SYNTH_SCALE;
}

/* Fill out[nrows][d] with rows row0 .. row0+nrows-1 of the stream `seed`. */
void b2o_synth_fill(float* out, int64_t row0, int64_t nrows, int d, uint64_t seed) {
    const uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < nrows; ++r) {
        uint64_t row = (uint64_t)(row0 + r);
        float* o = out + r * (int64_t)d;
        for (int c4 = 0; c4 * 4 < d; ++c4) {
            uint32_t c[4] = {(uint32_t)row, (uint32_t)(row >> 32), (uint32_t)c4, 0u};
            philox4x32_10(c, k0, k1);
            for (int j = 0; j < 4 && c4 * 4 + j < d; ++j) o[c4 * 4 + j] = synth_word(c[j]);
        }
    }
}

/* ------------------------------------------------------------------ */
/* faiss.normalize_L2 == fvec_renorm_L2 (faiss/utils/distances.cpp).   */
/* Reached from langchain FAISS.__add when normalize_L2=True           */
/* (SURVEY.md Appendix A.1); the reference never sets it               */
/* (common/utils.py:298 normalize_embeddings False) but north_star     */
/* names it (cfg5).                                                    */
/* ------------------------------------------------------------------ */
static inline float dot8(const float* x, const float* y, int d);

void b2o_normalize_l2(float* x, int64_t n, int d) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        float* r = x + i * (int64_t)d;
        float nr = dot8(r, r, d);
        if (nr > 0) {
            const float inv = 1.0f / sqrtf(nr);
            for (int j = 0; j < d; ++j) r[j] *= inv;
        }
    }
}

/* ------------------------------------------------------------------ */
/* Pairwise kernels: 8-lane accumulators + horizontal add, the shape   */
/* of faiss's AVX2 fvec_inner_product / fvec_L2sqr                     */
/* (faiss/utils/distances_simd.cpp), scalar tail.                      */
/* ------------------------------------------------------------------ */
static inline float hsum8(const float a[8]) {
    /* lo128 + hi128, then two hadd steps */
    float b0 = a[0] + a[4], b1 = a[1] + a[5], b2 = a[2] + a[6], b3 = a[3] + a[7];
    return (b0 + b1) + (b2 + b3);
}

static inline float dot8(const float* x, const float* y, int d) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int i = 0;
    for (; i + 8 <= d; i += 8)
        for (int l = 0; l < 8; ++l) acc[l] += x[i + l] * y[i + l];
    float s = hsum8(acc);
    for (; i < d; ++i) s += x[i] * y[i];
    return s;
}

static inline float l2sqr8(const float* x, const float* y, int d) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int i = 0;
    for (; i + 8 <= d; i += 8)
        for (int l = 0; l < 8; ++l) {
            const float t = x[i + l] - y[i + l];
            acc[l] += t * t;
        }
    float s = hsum8(acc);
    for (; i < d; ++i) {
        const float t = x[i] - y[i];
        s += t * t;
    }
    return s;
}

/* ------------------------------------------------------------------ */
/* Running top-k.  key = score for IP, -distance for L2, so "larger    */
/* key is better" for both; kept sorted best-first by (key desc, id    */
/* asc).  Unfilled slots are (-FLT_MAX, -1): a row enters only if its  */
/* key is strictly greater (faiss: heap seeded with C::neutral(),      */
/* admission by strict C::cmp), so NaN and -FLT_MAX never enter.       */
/* ------------------------------------------------------------------ */
typedef struct {
    int k;
    float* key;
    int64_t* id;
} topk_t;

static inline int beats(float key, int64_t id, float wkey, int64_t wid) {
    if (wid < 0) return key > wkey; /* empty slot */
    return key > wkey || (key == wkey && id < wid);
}

static inline void topk_push(topk_t* t, float key, int64_t id) {
    const int k = t->k;
    if (!beats(key, id, t->key[k - 1], t->id[k - 1])) return;
    int p = k - 1;
    while (p > 0 && beats(key, id, t->key[p - 1], t->id[p - 1])) {
        t->key[p] = t->key[p - 1];
        t->id[p] = t->id[p - 1];
        --p;
    }
    t->key[p] = key;
    t->id[p] = id;
}

static void scan_one_query(const float* x, int64_t n0, int64_t n1, int d, const float* q, int metric,
                           topk_t* t) {
    if (metric == B2O_METRIC_IP) {
        for (int64_t i = n0; i < n1; ++i) topk_push(t, dot8(q, x + i * (int64_t)d, d) + 0.0f, i);
    } else {
        for (int64_t i = n0; i < n1; ++i) topk_push(t, -l2sqr8(q, x + i * (int64_t)d, d), i);
    }
}

static void topk_emit(const topk_t* t, int metric, float* D, int64_t* I) {
    for (int j = 0; j < t->k; ++j) {
        I[j] = t->id[j];
        if (t->id[j] < 0)
            D[j] = (metric == B2O_METRIC_IP) ? -FLT_MAX : FLT_MAX;
        else
            D[j] = (metric == B2O_METRIC_IP) ? t->key[j] : -t->key[j];
    }
}

/*
 * IndexFlat::search for nq < 20 (faiss/utils/distances.cpp
 * exhaustive_inner_product_seq / exhaustive_L2sqr_seq): OpenMP over QUERIES
 * only, so one query runs on one core.  Called by every retriever call in the
 * reference.  Returns the number of threads used.
 */
int b2o_search(const float* x, int64_t n, int d, const float* q, int64_t nq, int k, int metric, float* D,
               int64_t* I) {
    int used = 1;
#ifdef _OPENMP
    int maxt = omp_get_max_threads();
    used = (int)(nq < maxt ? nq : maxt);
    if (used < 1) used = 1;
#endif
#pragma omp parallel for schedule(static) num_threads(used)
    for (int64_t qi = 0; qi < nq; ++qi) {
        topk_t t;
        t.k = k;
        t.key = (float*)malloc(sizeof(float) * k);
        t.id = (int64_t*)malloc(sizeof(int64_t) * k);
        for (int j = 0; j < k; ++j) {
            t.key[j] = -FLT_MAX;
            t.id[j] = -1;
        }
        scan_one_query(x, 0, n, d, q + qi * (int64_t)d, metric, &t);
        topk_emit(&t, metric, D + qi * (int64_t)k, I + qi * (int64_t)k);
        free(t.key);
        free(t.id);
    }
    return used;
}

/*
 * NOT faiss behaviour: rows split across all cores for each query, partial
 * lists merged with the same ordering.  Reported next to the faithful figure
 * as a fairer upper bound for batch 1.
 */
int b2o_search_rowpar(const float* x, int64_t n, int d, const float* q, int64_t nq, int k, int metric,
                      float* D, int64_t* I) {
    int T = 1;
#ifdef _OPENMP
    T = omp_get_max_threads();
#endif
    float* pk = (float*)malloc(sizeof(float) * (size_t)T * k);
    int64_t* pi = (int64_t*)malloc(sizeof(int64_t) * (size_t)T * k);
    for (int64_t qi = 0; qi < nq; ++qi) {
#pragma omp parallel num_threads(T)
        {
            int tid = 0;
#ifdef _OPENMP
            tid = omp_get_thread_num();
#endif
            topk_t t = {k, pk + (size_t)tid * k, pi + (size_t)tid * k};
            for (int j = 0; j < k; ++j) {
                t.key[j] = -FLT_MAX;
                t.id[j] = -1;
            }
            int64_t lo = n * tid / T, hi = n * (tid + 1) / T;
            scan_one_query(x, lo, hi, d, q + qi * (int64_t)d, metric, &t);
        }
        topk_t g;
        g.k = k;
        g.key = (float*)malloc(sizeof(float) * k);
        g.id = (int64_t*)malloc(sizeof(int64_t) * k);
        for (int j = 0; j < k; ++j) {
            g.key[j] = -FLT_MAX;
            g.id[j] = -1;
        }
        for (int tt = 0; tt < T; ++tt)
            for (int j = 0; j < k; ++j)
                if (pi[(size_t)tt * k + j] >= 0) topk_push(&g, pk[(size_t)tt * k + j], pi[(size_t)tt * k + j]);
        topk_emit(&g, metric, D + qi * (int64_t)k, I + qi * (int64_t)k);
        free(g.key);
        free(g.id);
    }
    free(pk);
    free(pi);
    return T;
}

/* Streaming variant for corpora that do not fit host RAM: the caller feeds
 * consecutive row blocks and the running lists persist in (key, id). */
void b2o_search_block(const float* xblk, int64_t row0, int64_t nrows, int d, const float* q, int64_t nq, int k,
                      int metric, float* key_state, int64_t* id_state) {
#pragma omp parallel for schedule(static)
    for (int64_t qi = 0; qi < nq; ++qi) {
        topk_t t = {k, key_state + qi * (int64_t)k, id_state + qi * (int64_t)k};
        const float* qq = q + qi * (int64_t)d;
        if (metric == B2O_METRIC_IP) {
            for (int64_t i = 0; i < nrows; ++i) topk_push(&t, dot8(qq, xblk + i * (int64_t)d, d) + 0.0f, row0 + i);
        } else {
            for (int64_t i = 0; i < nrows; ++i) topk_push(&t, -l2sqr8(qq, xblk + i * (int64_t)d, d), row0 + i);
        }
    }
}

/* Round fp32 -> bf16 -> fp32 (round to nearest even), for the bf16 row store. */
void b2o_round_bf16(float* x, int64_t n) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        uint32_t u;
        memcpy(&u, &x[i], 4);
        if ((u & 0x7f800000u) == 0x7f800000u) { /* inf / nan: truncate, keep nan a nan */
            if (u & 0x007fffffu) u |= 0x00400000u;
            u &= 0xffff0000u;
        } else {
            u += 0x7fffu + ((u >> 16) & 1u);
            u &= 0xffff0000u;
        }
        memcpy(&x[i], &u, 4);
    }
}

int b2o_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
