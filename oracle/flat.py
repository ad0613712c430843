"""CPU oracle for the flat vector-store search path (ctypes over flat_oracle.c + numpy).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package
(nv_genai_b200/) never imports it.

PARITY UNPINNED BY THE REFERENCE (see the header of flat_oracle.c): the
reference holds no implementation, test or golden vector for this path; the
algorithm is the un-vendored dependency faiss-cpu==1.7.4
(/root/reference/RetrievalAugmentedGeneration/requirements.txt:25) reached
through langchain's FAISS vector store
(/root/reference/RetrievalAugmentedGeneration/common/utils.py:217,
/root/reference/examples/5_mins_rag_no_gpu/main.py:90).  Each function below
cites the upstream routine it restates and the reference call site reaching it.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import List, Optional, Sequence, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libflat_oracle.so")
_SRC = os.path.join(_HERE, "flat_oracle.c")

METRIC_IP = 0  # faiss.METRIC_INNER_PRODUCT
METRIC_L2 = 1  # faiss.METRIC_L2
FLT_MAX = float(np.finfo(np.float32).max)

_lib = None


def build(force: bool = False) -> str:
    """Compile flat_oracle.c -> libflat_oracle.so (gcc -O3 -march=native -fopenmp)."""
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        cmd = ["gcc", "-O3", "-march=native", "-fopenmp", "-fPIC", "-shared", "-o", _SO, _SRC, "-lm"]
        subprocess.check_call(cmd)
    return _SO


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = ctypes.CDLL(_SO)
        f32p = ctypes.POINTER(ctypes.c_float)
        i64p = ctypes.POINTER(ctypes.c_int64)
        L.b2o_synth_fill.argtypes = [f32p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int, ctypes.c_uint64]
        L.b2o_synth_fill.restype = None
        L.b2o_normalize_l2.argtypes = [f32p, ctypes.c_int64, ctypes.c_int]
        L.b2o_normalize_l2.restype = None
        for name in ("b2o_search", "b2o_search_rowpar"):
            fn = getattr(L, name)
            fn.argtypes = [f32p, ctypes.c_int64, ctypes.c_int, f32p, ctypes.c_int64, ctypes.c_int, ctypes.c_int,
                           f32p, i64p]
            fn.restype = ctypes.c_int
        L.b2o_search_block.argtypes = [f32p, ctypes.c_int64, ctypes.c_int64, ctypes.c_int, f32p, ctypes.c_int64,
                                       ctypes.c_int, ctypes.c_int, f32p, i64p]
        L.b2o_search_block.restype = None
        L.b2o_round_bf16.argtypes = [f32p, ctypes.c_int64]
        L.b2o_round_bf16.restype = None
        L.b2o_max_threads.restype = ctypes.c_int
        _lib = L
    return _lib


def _f32(a: np.ndarray):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _i64(a: np.ndarray):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))


# --------------------------------------------------------------------------
# synthetic data (generator defined in flat_oracle.c)
# --------------------------------------------------------------------------
CORPUS_SEED = 0x5EED0001
QUERY_SEED = 0x5EED0002


def synth(row0: int, nrows: int, d: int, seed: int) -> np.ndarray:
    out = np.empty((nrows, d), dtype=np.float32)
    lib().b2o_synth_fill(_f32(out), row0, nrows, d, seed)
    return out


def synth_numpy(row0: int, nrows: int, d: int, seed: int) -> np.ndarray:
    """Pure-numpy Philox-4x32-10 restatement of b2o_synth_fill (small sizes; cross-checks the C)."""
    M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
    W0, W1 = 0x9E3779B9, 0xBB67AE85
    nc4 = (d + 3) // 4
    rows = (np.arange(nrows, dtype=np.uint64) + np.uint64(row0))[:, None]
    c0 = np.broadcast_to(rows & np.uint64(0xFFFFFFFF), (nrows, nc4)).copy()
    c1 = np.broadcast_to(rows >> np.uint64(32), (nrows, nc4)).copy()
    c2 = np.broadcast_to(np.arange(nc4, dtype=np.uint64)[None, :], (nrows, nc4)).copy()
    c3 = np.zeros((nrows, nc4), dtype=np.uint64)
    k0, k1 = seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF
    mask = np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0 = M0 * c0
        p1 = M1 * c2
        n0 = ((p1 >> np.uint64(32)) ^ c1 ^ np.uint64(k0)) & mask
        n1 = p1 & mask
        n2 = ((p0 >> np.uint64(32)) ^ c3 ^ np.uint64(k1)) & mask
        n3 = p0 & mask
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    w = np.stack([c0, c1, c2, c3], axis=-1).reshape(nrows, nc4 * 4)[:, :d]
    s = (w & np.uint64(0xFFFF)).astype(np.int64) + (w >> np.uint64(16)).astype(np.int64) - 65535
    return (s.astype(np.float32) * np.float32(1.0 / 26754.58)).astype(np.float32)


# --------------------------------------------------------------------------
# faiss.normalize_L2 (fvec_renorm_L2)
# --------------------------------------------------------------------------
def normalize_L2(x: np.ndarray) -> None:
    """In place, like faiss.normalize_L2: zero rows stay zero."""
    assert x.dtype == np.float32 and x.flags.c_contiguous and x.ndim == 2
    lib().b2o_normalize_l2(_f32(x), x.shape[0], x.shape[1])


def round_bf16(x: np.ndarray) -> np.ndarray:
    y = np.ascontiguousarray(x, dtype=np.float32).copy()
    lib().b2o_round_bf16(_f32(y), y.size)
    return y


def round_bf16_inplace(x: np.ndarray) -> None:
    """round_bf16 without the copy (streaming checks over multi-GB blocks)."""
    assert x.dtype == np.float32 and x.flags.c_contiguous
    lib().b2o_round_bf16(_f32(x), x.size)


# --------------------------------------------------------------------------
# IndexFlat.search
# --------------------------------------------------------------------------
BLAS_THRESHOLD = 20  # faiss distance_compute_blas_threshold


def search_seq(x: np.ndarray, q: np.ndarray, k: int, metric: int, rowpar: bool = False):
    """faiss exhaustive_*_seq (nq < 20 path): returns (D, I, threads_used)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    q = np.ascontiguousarray(q, dtype=np.float32)
    assert x.ndim == 2 and q.ndim == 2 and (x.shape[0] == 0 or x.shape[1] == q.shape[1])
    assert k > 0
    nq, d = q.shape
    D = np.empty((nq, k), dtype=np.float32)
    I = np.empty((nq, k), dtype=np.int64)
    fn = lib().b2o_search_rowpar if rowpar else lib().b2o_search
    used = fn(_f32(x), x.shape[0], d, _f32(q), nq, k, metric, _f32(D), _i64(I))
    return D, I, used


def _topk_rows(key: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray]:
    """Row-wise top-k of `key` (larger better), ordered (key desc, id asc); pads with -1."""
    nq, n = key.shape
    outk = np.full((nq, k), -FLT_MAX, dtype=np.float32)
    outi = np.full((nq, k), -1, dtype=np.int64)
    kk = min(k, n)
    if kk == 0:
        return outk, outi
    ids = np.arange(n, dtype=np.int64)
    for r in range(nq):
        row = key[r]
        valid = row > -FLT_MAX  # NaN and -FLT_MAX never enter (strict admission)
        vi = ids[valid]
        vr = row[valid]
        order = np.lexsort((vi, -vr.astype(np.float64)))[:kk]
        m = len(order)
        outk[r, :m] = vr[order]
        outi[r, :m] = vi[order]
    return outk, outi


def search_blas(x: np.ndarray, q: np.ndarray, k: int, metric: int, block: int = 65536):
    """faiss exhaustive_*_blas (nq >= 20): sgemm blocks; L2 = |q|^2 + |x|^2 - 2 q.x clamped at 0."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    q = np.ascontiguousarray(q, dtype=np.float32)
    nq = q.shape[0]
    n = x.shape[0]
    bk = np.full((nq, k), -FLT_MAX, dtype=np.float32)
    bi = np.full((nq, k), -1, dtype=np.int64)
    qn = np.einsum("ij,ij->i", q, q).astype(np.float32)
    for s in range(0, n, block):
        xb = x[s:s + block]
        ip = q @ xb.T
        if metric == METRIC_IP:
            key = ip + np.float32(0.0)
        else:
            xn = np.einsum("ij,ij->i", xb, xb).astype(np.float32)
            dis = qn[:, None] + xn[None, :] - np.float32(2.0) * ip
            dis = np.maximum(dis, np.float32(0.0))
            key = -dis
        ck, ci = _topk_rows(key, k)
        ci = np.where(ci >= 0, ci + s, -1)
        ak = np.concatenate([bk, ck], axis=1)
        ai = np.concatenate([bi, ci], axis=1)
        for r in range(nq):
            valid = ai[r] >= 0
            order = np.lexsort((ai[r][valid], -ak[r][valid].astype(np.float64)))[:k]
            m = len(order)
            bk[r, :] = -FLT_MAX
            bi[r, :] = -1
            bk[r, :m] = ak[r][valid][order]
            bi[r, :m] = ai[r][valid][order]
    D = np.where(bi >= 0, bk if metric == METRIC_IP else -bk, -FLT_MAX if metric == METRIC_IP else FLT_MAX)
    return D.astype(np.float32), bi


def search(x: np.ndarray, q: np.ndarray, k: int, metric: int):
    """IndexFlat::search dispatch (faiss knn_inner_product / knn_L2sqr): seq below 20 queries, BLAS above."""
    q = np.atleast_2d(q)
    if q.shape[0] < BLAS_THRESHOLD:
        D, I, _ = search_seq(x, q, k, metric)
        return D, I
    return search_blas(x, q, k, metric)


def search_f64(x: np.ndarray, q: np.ndarray, k: int, metric: int):
    """Ground truth in float64 (direct forms), same ordering rule; pins the fp32 oracle on small shapes."""
    x64 = np.asarray(x, dtype=np.float64)
    q64 = np.atleast_2d(np.asarray(q, dtype=np.float64))
    if metric == METRIC_IP:
        key = q64 @ x64.T
    else:
        key = -((q64[:, None, :] - x64[None, :, :]) ** 2).sum(-1)
    nq, n = key.shape
    D = np.full((nq, k), -FLT_MAX if metric == METRIC_IP else FLT_MAX, dtype=np.float64)
    I = np.full((nq, k), -1, dtype=np.int64)
    for r in range(nq):
        order = np.lexsort((np.arange(n), -key[r]))[:k]
        m = len(order)
        I[r, :m] = order
        D[r, :m] = key[r][order] if metric == METRIC_IP else -key[r][order]
    return D, I


class StreamingSearch:
    """Scan a corpus that is produced block by block (bench cpu_baseline on a bounded sample)."""

    def __init__(self, q: np.ndarray, k: int, metric: int):
        self.q = np.ascontiguousarray(np.atleast_2d(q), dtype=np.float32)
        self.k, self.metric = k, metric
        nq = self.q.shape[0]
        self.key = np.full((nq, k), -FLT_MAX, dtype=np.float32)
        self.id = np.full((nq, k), -1, dtype=np.int64)

    def feed(self, xblk: np.ndarray, row0: int) -> None:
        xblk = np.ascontiguousarray(xblk, dtype=np.float32)
        lib().b2o_search_block(_f32(xblk), row0, xblk.shape[0], xblk.shape[1], _f32(self.q), self.q.shape[0],
                               self.k, self.metric, _f32(self.key), _i64(self.id))

    def result(self):
        if self.metric == METRIC_IP:
            D = np.where(self.id >= 0, self.key, -FLT_MAX)
        else:
            D = np.where(self.id >= 0, -self.key, FLT_MAX)
        return D.astype(np.float32), self.id.copy()


# --------------------------------------------------------------------------
# langchain glue restated (langchain_community/vectorstores/utils.py, utils/math.py)
# --------------------------------------------------------------------------
def cosine_similarity(X: np.ndarray, Y: np.ndarray) -> np.ndarray:
    """langchain_community.utils.math.cosine_similarity (numpy path): NaN/Inf -> 0."""
    X = np.array(X)
    Y = np.array(Y)
    if X.size == 0 or Y.size == 0:
        return np.array([])
    if X.shape[1] != Y.shape[1]:
        raise ValueError(
            f"Number of columns in X and Y must be the same. X has shape {X.shape} and Y has shape {Y.shape}.")
    Xn = np.linalg.norm(X, axis=1)
    Yn = np.linalg.norm(Y, axis=1)
    with np.errstate(divide="ignore", invalid="ignore"):
        sim = np.dot(X, Y.T) / np.outer(Xn, Yn)
    sim[np.isnan(sim) | np.isinf(sim)] = 0.0
    return sim


def maximal_marginal_relevance(query_embedding: np.ndarray, embedding_list: Sequence[np.ndarray],
                               lambda_mult: float = 0.5, k: int = 4) -> List[int]:
    """langchain_community.vectorstores.utils.maximal_marginal_relevance."""
    if min(k, len(embedding_list)) <= 0:
        return []
    if query_embedding.ndim == 1:
        query_embedding = np.expand_dims(query_embedding, axis=0)
    similarity_to_query = cosine_similarity(query_embedding, embedding_list)[0]
    most_similar = int(np.argmax(similarity_to_query))
    idxs = [most_similar]
    selected = np.array([embedding_list[most_similar]])
    while len(idxs) < min(k, len(embedding_list)):
        best_score = -np.inf
        idx_to_add = -1
        similarity_to_selected = cosine_similarity(embedding_list, selected)
        for i, query_score in enumerate(similarity_to_query):
            if i in idxs:
                continue
            redundant_score = max(similarity_to_selected[i])
            equation_score = lambda_mult * query_score - (1 - lambda_mult) * redundant_score
            if equation_score > best_score:
                best_score = equation_score
                idx_to_add = i
        idxs.append(idx_to_add)
        selected = np.append(selected, [embedding_list[idx_to_add]], axis=0)
    return idxs


def mmr_search(x: np.ndarray, q: np.ndarray, k: int, fetch_k: int, lambda_mult: float, metric: int):
    """FAISS.max_marginal_relevance_search_with_score_by_vector for one query: (ids, scores) in pick order."""
    D, I = search(x, np.atleast_2d(q), fetch_k, metric)
    ids = [int(i) for i in I[0] if i != -1]
    embs = [x[i] for i in ids]
    sel = maximal_marginal_relevance(np.asarray(q, dtype=np.float32).reshape(1, -1), embs, lambda_mult, k)
    return [ids[s] for s in sel], [float(D[0][s]) for s in sel]


def euclidean_relevance(d: float) -> float:
    """langchain VectorStore._euclidean_relevance_score_fn, applied to the squared L2 as-is."""
    return 1.0 - d / np.sqrt(2.0)


def max_inner_product_relevance(d: float) -> float:
    """langchain VectorStore._max_inner_product_relevance_score_fn."""
    return 1.0 - d if d > 0 else -1.0 * d


def cosine_relevance(d: float) -> float:
    """langchain VectorStore._cosine_relevance_score_fn."""
    return 1.0 - d


REL_NONE, REL_EUCLIDEAN, REL_COSINE, REL_MAX_INNER_PRODUCT = 0, 1, 2, 3
_REL_FN = {REL_EUCLIDEAN: euclidean_relevance, REL_COSINE: cosine_relevance,
           REL_MAX_INNER_PRODUCT: max_inner_product_relevance}


def select_hits(D: np.ndarray, I: np.ndarray, k: int, metric: int, allow: Optional[np.ndarray] = None,
                raw_threshold: Optional[float] = None, relevance: int = REL_NONE, relevance_threshold: float = 0.0):
    """What langchain's FAISS does between index.search and the Document list (thresholded retrievers at
    /root/reference/RetrievalAugmentedGeneration/examples/nvidia_api_catalog/chains.py:122,162), on index positions:

        similarity_search_with_score_by_vector: skip -1; keep hits passing the metadata filter (`allow[i]`, a bool per
        row); keep score >= thr (inner product) or <= thr (L2); docs[:k]
        similarity_search_with_relevance_scores: keep relevance(score) >= thr

    Comparisons on float(score), i.e. in double, as numpy 1.26 does for np.float32 against a Python float.
    Returns (D [nq, k] raw scores, I [nq, k], counts [nq]) padded like index.search pads."""
    D = np.atleast_2d(D)
    I = np.atleast_2d(I)
    nq, fetch_k = I.shape
    pad = -FLT_MAX if metric == METRIC_IP else FLT_MAX
    Do = np.full((nq, k), pad, dtype=np.float32)
    Io = np.full((nq, k), -1, dtype=np.int64)
    C = np.zeros((nq,), dtype=np.int32)
    for q in range(nq):
        docs = []
        for j in range(fetch_k):
            i = int(I[q, j])
            if i == -1:
                continue
            if allow is not None and not (i < len(allow) and allow[i]):
                continue
            docs.append((D[q, j], i))
        if raw_threshold is not None:
            if metric == METRIC_IP:
                docs = [(s, i) for s, i in docs if float(s) >= raw_threshold]
            else:
                docs = [(s, i) for s, i in docs if float(s) <= raw_threshold]
        docs = docs[:k]
        if relevance != REL_NONE:
            fn = _REL_FN[relevance]
            docs = [(s, i) for s, i in docs if fn(float(s)) >= relevance_threshold]
        for j, (s, i) in enumerate(docs):
            Do[q, j], Io[q, j] = s, i
        C[q] = len(docs)
    return Do, Io, C
