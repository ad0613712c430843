"""faiss-shaped index objects over the B200 C ABI.

`IndexFlatL2(d)` / `IndexFlatIP(d)` keep the shape of the objects the reference builds at
/root/reference/RetrievalAugmentedGeneration/common/utils.py:217
(`FAISS(document_embedder, IndexFlatL2(config.embeddings.dimensions), InMemoryDocstore(), {})`)
and that langchain's FAISS creates inside `from_documents`
(/root/reference/examples/5_mins_rag_no_gpu/main.py:90): attributes `d`, `ntotal`,
`metric_type`, `is_trained`; methods `add`, `search -> (D, I)`, `reconstruct`, `remove_ids`,
`merge_from`, `reset` with the argument meaning and the AssertionError / ValueError
behaviour of faiss's SWIG wrapper.  The rows live in HBM; every
method is a call through include/b2_flat.h.  numpy in -> numpy out (host buffers, copies
inside the call); torch CUDA tensor in -> torch CUDA tensor out on the current stream.
"""
from __future__ import annotations

import ctypes
import sys
from typing import Optional, Tuple

import numpy as np

from . import _native as N

METRIC_INNER_PRODUCT = N.METRIC_INNER_PRODUCT
METRIC_L2 = N.METRIC_L2

_DTYPES = {"float32": N.DTYPE_F32, "fp32": N.DTYPE_F32, "f32": N.DTYPE_F32, "bfloat16": N.DTYPE_BF16,
           "bf16": N.DTYPE_BF16, "bfloat16+residual": N.DTYPE_BF16, "bf16+residual": N.DTYPE_BF16}
# "bfloat16+residual": bf16 rows plus a second bf16 plane holding x - bf16(x) (4 bytes per element in HBM).  Searches
# stream the main plane only, like a plain bf16 store, and every returned score is re-computed against both planes, so
# recall against the fp32 embeddings is that of ~16 mantissa bits instead of 8 (b2_index_set_option "bf16_residual").


def _torch():
    return sys.modules.get("torch")


def _is_tensor(x) -> bool:
    t = _torch()
    return t is not None and isinstance(x, t.Tensor)


def _resolve_device(device) -> int:
    if device is None:
        t = _torch()
        if t is not None and t.cuda.is_available() and t.cuda.is_initialized():
            return int(t.cuda.current_device())
        return 0
    if isinstance(device, int):
        return device
    s = str(device)
    if s == "cuda":
        return _resolve_device(None)
    if s.startswith("cuda:"):
        return int(s.split(":", 1)[1])
    raise ValueError(f"not a CUDA device: {device!r}")


def _stream_for(device: int) -> Optional[int]:
    """The caller's current torch stream if torch is in use, else the legacy default stream."""
    t = _torch()
    if t is not None and t.cuda.is_available() and t.cuda.is_initialized():
        return int(t.cuda.current_stream(device).cuda_stream) or None
    return None


class _Buf:
    """A C-contiguous fp32 [n, d] view of the caller's data plus where it lives."""

    __slots__ = ("obj", "ptr", "where", "n", "d", "is_tensor")

    def __init__(self, x, d_expected: Optional[int], device: int, what: str):
        if _is_tensor(x):
            t = _torch()
            if x.dim() == 1:
                x = x.unsqueeze(0)
            if x.dim() != 2:
                raise ValueError(f"{what}: expected a 2-D array, got shape {tuple(x.shape)}")
            if x.is_cuda:
                if x.device.index != device:
                    raise ValueError(f"{what}: tensor is on cuda:{x.device.index}, the index is on cuda:{device}")
                x = x.detach().to(t.float32).contiguous()
                self.where = N.DEVICE
            else:
                x = x.detach().to(t.float32).contiguous()
                self.where = N.HOST
            self.obj, self.ptr, self.is_tensor = x, x.data_ptr(), bool(x.is_cuda)
            self.n, self.d = int(x.shape[0]), int(x.shape[1])
        else:
            a = np.asarray(x)
            if a.ndim == 1:
                a = a.reshape(1, -1)
            if a.ndim != 2:
                raise ValueError(f"{what}: expected a 2-D array, got shape {a.shape}")
            a = np.ascontiguousarray(a, dtype=np.float32)
            self.obj, self.ptr, self.where, self.is_tensor = a, a.ctypes.data, N.HOST, False
            self.n, self.d = int(a.shape[0]), int(a.shape[1])
        if d_expected is not None:
            assert self.d == d_expected, f"{what}: dimension {self.d} does not match the index dimension {d_expected}"


class IndexFlat:
    """Exact (brute-force) index with its rows in HBM.  faiss.IndexFlat look-alike."""

    def __init__(self, d: int, metric: int = METRIC_L2, dtype: str = "float32", device=None):
        if dtype not in _DTYPES:
            raise ValueError(f"dtype must be one of {sorted(_DTYPES)}")
        self._h = ctypes.c_void_p()
        residual = dtype.endswith("+residual")
        self._dtype_name = ("bfloat16+residual" if residual else "bfloat16") if _DTYPES[dtype] == N.DTYPE_BF16 else "float32"
        self.device = _resolve_device(device)
        N.check(N.lib.b2_index_create(int(d), int(metric), _DTYPES[dtype], self.device, ctypes.byref(self._h)))
        if residual:
            N.check(N.lib.b2_index_set_option(self._h, b"bf16_residual", 1))
        self.is_trained = True

    # ---- attributes faiss exposes
    @property
    def d(self) -> int:
        return int(N.lib.b2_index_d(self._h))

    @property
    def ntotal(self) -> int:
        return int(N.lib.b2_index_ntotal(self._h))

    @property
    def metric_type(self) -> int:
        return int(N.lib.b2_index_metric(self._h))

    @property
    def dtype(self) -> str:
        return self._dtype_name

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                N.lib.b2_index_destroy(h)
            except Exception:
                pass
            h.value = None

    def __len__(self) -> int:
        return self.ntotal

    # ---- index.add
    def add(self, x, normalize: bool = False) -> None:
        """faiss `index.add(x)`; `normalize=True` fuses faiss.normalize_L2(x) into the ingest kernel."""
        b = _Buf(x, self.d, self.device, "add")
        flags = N.ADD_NORMALIZE if normalize else 0
        N.check(N.lib.b2_index_add(self._h, b.n, b.ptr, b.where, flags, _stream_for(self.device)))

    # ---- index.search
    def search(self, x, k: int):
        """faiss `index.search(x, k) -> (D, I)`: D float32 [n, k], I int64 [n, k]."""
        b = _Buf(x, self.d, self.device, "search")
        k = int(k)
        assert k > 0, "k must be positive"
        if b.is_tensor:
            t = _torch()
            D = t.empty((b.n, k), dtype=t.float32, device=b.obj.device)
            I = t.empty((b.n, k), dtype=t.int64, device=b.obj.device)
            N.check(N.lib.b2_index_search(self._h, b.n, b.ptr, N.DEVICE, k, D.data_ptr(), I.data_ptr(), N.DEVICE,
                                          _stream_for(self.device)))
            return D, I
        D = np.empty((b.n, k), dtype=np.float32)
        I = np.empty((b.n, k), dtype=np.int64)
        N.check(N.lib.b2_index_search(self._h, b.n, b.ptr, b.where, k, D.ctypes.data, I.ctypes.data, N.HOST,
                                      _stream_for(self.device)))
        return D, I

    def search_into(self, x, k: int, D, I) -> None:
        """`search` into caller-provided CUDA tensors D [n, k] float32 and I [n, k] int64 (contiguous), asynchronous on
        the current stream: the batched device path without any allocation on the Python side."""
        b = _Buf(x, self.d, self.device, "search")
        assert b.is_tensor and D.is_cuda and I.is_cuda and D.is_contiguous() and I.is_contiguous()
        assert tuple(D.shape) == (b.n, int(k)) and tuple(I.shape) == (b.n, int(k))
        N.check(N.lib.b2_index_search(self._h, b.n, b.ptr, N.DEVICE, int(k), D.data_ptr(), I.data_ptr(), N.DEVICE,
                                      _stream_for(self.device)))

    def upload_mask(self, words: np.ndarray):
        """A row bitmask (uint32 words, bit i of word i // 32 = row i) as a CUDA int32 tensor on the index's device,
        the form `search_filtered(allow=...)` takes without a copy per call."""
        import torch

        return torch.from_numpy(np.ascontiguousarray(words, dtype=np.uint32).view(np.int32)).to(f"cuda:{self.device}")

    def search_filtered(self, x, k: int, fetch_k: Optional[int] = None, allow=None, score_threshold=None,
                        relevance: int = N.REL_NONE, relevance_threshold: float = 0.0):
        """`search(x, fetch_k)` followed, on the device, by the selection langchain's FAISS runs on the host
        (b2_index_search_filtered in include/b2_flat.h): row bitmask `allow` (uint32 words, numpy or CUDA tensor; None =
        no metadata filter), raw `score_threshold` (>= for inner product, <= for L2), cut at k, then the relevance
        threshold if `relevance` names one of langchain's relevance functions.

        Returns (D [n, k] raw scores, I [n, k], counts [n] int32): the first counts[i] entries of row i are results."""
        b = _Buf(x, self.d, self.device, "search")
        k = int(k)
        fetch_k = k if fetch_k is None else int(fetch_k)
        assert k > 0, "k must be positive"
        use_raw = score_threshold is not None
        raw = float(score_threshold) if use_raw else 0.0
        a_ptr, a_bits, a_where = None, 0, N.DEVICE
        if allow is not None:
            if _is_tensor(allow):
                t = _torch()
                assert allow.is_cuda and allow.dtype in (t.int32, t.uint32) and allow.is_contiguous()
                a_ptr, a_bits, a_where = allow.data_ptr(), allow.numel() * 32, N.DEVICE
            else:
                allow = np.ascontiguousarray(allow, dtype=np.uint32)
                a_ptr, a_bits, a_where = allow.ctypes.data, allow.size * 32, N.HOST
        if b.is_tensor:
            t = _torch()
            D = t.empty((b.n, k), dtype=t.float32, device=b.obj.device)
            I = t.empty((b.n, k), dtype=t.int64, device=b.obj.device)
            C = t.empty((b.n,), dtype=t.int32, device=b.obj.device)
            if a_where == N.HOST and a_ptr is not None:  # device outputs take a device mask
                allow = t.from_numpy(allow.view(np.int32)).to(b.obj.device)
                a_ptr, a_where = allow.data_ptr(), N.DEVICE
            N.check(N.lib.b2_index_search_filtered(self._h, b.n, b.ptr, N.DEVICE, k, fetch_k, a_ptr, a_bits, a_where,
                                                   1 if use_raw else 0, raw, int(relevance), float(relevance_threshold),
                                                   D.data_ptr(), I.data_ptr(), C.data_ptr(), N.DEVICE,
                                                   _stream_for(self.device)))
            return D, I, C
        D = np.empty((b.n, k), dtype=np.float32)
        I = np.empty((b.n, k), dtype=np.int64)
        C = np.empty((b.n,), dtype=np.int32)
        N.check(N.lib.b2_index_search_filtered(self._h, b.n, b.ptr, b.where, k, fetch_k, a_ptr, a_bits, a_where,
                                               1 if use_raw else 0, raw, int(relevance), float(relevance_threshold),
                                               D.ctypes.data, I.ctypes.data, C.ctypes.data, N.HOST,
                                               _stream_for(self.device)))
        return D, I, C

    # ---- index.reconstruct*
    def reconstruct_batch(self, ids):
        if _is_tensor(ids) and ids.is_cuda:
            t = _torch()
            ids = ids.to(t.int64).contiguous()
            out = t.empty((ids.numel(), self.d), dtype=t.float32, device=ids.device)
            N.check(N.lib.b2_index_reconstruct_batch(self._h, ids.numel(), ids.data_ptr(), N.DEVICE, out.data_ptr(),
                                                     N.DEVICE, _stream_for(self.device)))
            return out
        a = np.ascontiguousarray(np.asarray(ids, dtype=np.int64).reshape(-1))
        out = np.empty((a.size, self.d), dtype=np.float32)
        N.check(N.lib.b2_index_reconstruct_batch(self._h, a.size, a.ctypes.data, N.HOST, out.ctypes.data, N.HOST,
                                                 _stream_for(self.device)))
        return out

    def reconstruct(self, key: int) -> np.ndarray:
        key = int(key)
        if key < 0 or key >= self.ntotal:
            raise RuntimeError(f"reconstruct: key {key} out of range [0, {self.ntotal})")
        return self.reconstruct_batch([key])[0]

    def reconstruct_n(self, n0: int = 0, ni: int = -1) -> np.ndarray:
        if ni == -1:
            ni = self.ntotal - n0
        if n0 < 0 or ni < 0 or n0 + ni > self.ntotal:
            raise RuntimeError(f"reconstruct_n: range [{n0}, {n0 + ni}) out of [0, {self.ntotal})")
        return self.reconstruct_batch(np.arange(n0, n0 + ni, dtype=np.int64))

    # ---- index.remove_ids / merge_from / reset
    def remove_ids(self, ids) -> int:
        a = np.ascontiguousarray(np.asarray(ids, dtype=np.int64).reshape(-1))
        removed = ctypes.c_int64(0)
        N.check(N.lib.b2_index_remove_ids(self._h, a.size, a.ctypes.data, ctypes.byref(removed)))
        return int(removed.value)

    def merge_from(self, other: "IndexFlat", add_id: int = 0) -> None:
        if not isinstance(other, IndexFlat):
            raise TypeError("merge_from expects another B200 IndexFlat")
        N.check(N.lib.b2_index_merge_from(self._h, other._h))

    def reset(self) -> None:
        N.check(N.lib.b2_index_reset(self._h))

    def set_option(self, name: str, value: int) -> None:
        """Engine options (no faiss counterpart), e.g. ("gemm_query_parts", 1) for the faster single-part query
        operand of the tensor-core kernel; see b2_index_set_option in include/b2_flat.h."""
        N.check(N.lib.b2_index_set_option(self._h, name.encode(), int(value)))

    # ---- batched MMR on the device (langchain's maximal_marginal_relevance over index.search hits)
    def mmr(self, x, k: int = 4, fetch_k: int = 20, lambda_mult: float = 0.5):
        """For each query: search fetch_k, reconstruct the hits, greedy MMR pick of k.

        Returns (ids [n, k] int64 in pick order, -1 padded; scores [n, k] float32 = search scores)."""
        b = _Buf(x, self.d, self.device, "mmr")
        k, fetch_k = int(k), int(fetch_k)
        assert k > 0 and fetch_k > 0
        if b.is_tensor:
            t = _torch()
            ids = t.empty((b.n, k), dtype=t.int64, device=b.obj.device)
            sc = t.empty((b.n, k), dtype=t.float32, device=b.obj.device)
            N.check(N.lib.b2_index_mmr(self._h, b.n, b.ptr, N.DEVICE, k, fetch_k, float(lambda_mult), ids.data_ptr(),
                                       sc.data_ptr(), N.DEVICE, _stream_for(self.device)))
            return ids, sc
        ids = np.empty((b.n, k), dtype=np.int64)
        sc = np.empty((b.n, k), dtype=np.float32)
        N.check(N.lib.b2_index_mmr(self._h, b.n, b.ptr, b.where, k, fetch_k, float(lambda_mult), ids.ctypes.data,
                                   sc.ctypes.data, N.HOST, _stream_for(self.device)))
        return ids, sc

    # ---- pickling (the reference pickles the whole vector store: examples/5_mins_rag_no_gpu/main.py:93-94)
    def __getstate__(self):
        return {"d": self.d, "metric": self.metric_type, "dtype": self._dtype_name,
                "rows": self.reconstruct_n(0, self.ntotal) if self.ntotal else np.empty((0, self.d), np.float32)}

    def __setstate__(self, st):
        IndexFlat.__init__(self, st["d"], st["metric"], st.get("dtype", "float32"))
        if len(st["rows"]):
            self.add(st["rows"])

    def __repr__(self) -> str:
        name = "IP" if self.metric_type == METRIC_INNER_PRODUCT else "L2"
        return f"<b200 IndexFlat{name} d={self.d} ntotal={self.ntotal} dtype={self._dtype_name} cuda:{self.device}>"


class IndexFlatL2(IndexFlat):
    """Drop-in for `faiss.IndexFlatL2(d)` (squared Euclidean distance, ascending)."""

    def __init__(self, d: int, dtype: str = "float32", device=None):
        super().__init__(d, METRIC_L2, dtype, device)


class IndexFlatIP(IndexFlat):
    """Drop-in for `faiss.IndexFlatIP(d)` (inner product, descending)."""

    def __init__(self, d: int, dtype: str = "float32", device=None):
        super().__init__(d, METRIC_INNER_PRODUCT, dtype, device)


def normalize_L2(x) -> None:
    """`faiss.normalize_L2(x)`: in place, zero rows untouched; runs the device kernel."""
    if _is_tensor(x) and x.is_cuda:
        t = _torch()
        assert x.dtype == t.float32 and x.is_contiguous() and x.dim() == 2
        N.check(N.lib.b2_normalize_l2(x.data_ptr(), x.shape[0], x.shape[1], x.device.index,
                                      _stream_for(x.device.index)))
        return
    assert isinstance(x, np.ndarray) and x.dtype == np.float32 and x.ndim == 2 and x.flags.c_contiguous, \
        "normalize_L2 needs a C-contiguous float32 2-D array"
    if x.shape[0] == 0:
        return
    import torch  # device memory + stream plumbing only

    dev = _resolve_device(None)
    tmp = torch.from_numpy(x).to(f"cuda:{dev}")
    N.check(N.lib.b2_normalize_l2(tmp.data_ptr(), x.shape[0], x.shape[1], dev, _stream_for(dev)))
    x[...] = tmp.cpu().numpy()


def synth_fill(out, row0: int, seed: int) -> None:
    """Fill a CUDA float32 [n, d] tensor with rows row0.. of the synthetic stream `seed`."""
    assert _is_tensor(out) and out.is_cuda and out.is_contiguous() and out.dim() == 2
    N.check(N.lib.b2_synth_fill(out.data_ptr(), int(row0), out.shape[0], out.shape[1], int(seed), out.device.index,
                                _stream_for(out.device.index)))


def merge_shard_lists_packed(metric: int, gathered, nq: int, k: int, bases):
    """`gathered`: uint8 [G, nq*k*12] as the all-gather leaves it, each part = [I local ids int64 | D fp32];
    `bases`: G global ids of the shards' first rows.  Returns the global (D, I) [nq, k]."""
    t = _torch()
    assert gathered.is_cuda and gathered.is_contiguous() and gathered.dim() == 2
    G = gathered.shape[0]
    stride = gathered.shape[1]
    assert stride == nq * k * 12 and len(bases) >= G
    D = t.empty((nq, k), dtype=t.float32, device=gathered.device)
    I = t.empty((nq, k), dtype=t.int64, device=gathered.device)
    arr = (ctypes.c_int64 * G)(*[int(b) for b in bases[:G]])
    dev = gathered.device.index
    base_ptr = gathered.data_ptr()
    N.check(N.lib.b2_merge_shard_lists_packed(int(metric), nq, k, G, base_ptr + nq * k * 8, base_ptr, stride, arr,
                                              D.data_ptr(), I.data_ptr(), dev, _stream_for(dev)))
    return D, I


def merge_shard_lists(metric: int, D_parts, I_parts):
    """[G, nq, k] per-shard lists (global ids, shard order = id order) -> global (D, I) [nq, k]."""
    t = _torch()
    assert D_parts.is_cuda and I_parts.is_cuda and D_parts.dim() == 3
    D_parts = D_parts.contiguous()
    I_parts = I_parts.contiguous()
    G, nq, k = D_parts.shape
    D = t.empty((nq, k), dtype=t.float32, device=D_parts.device)
    I = t.empty((nq, k), dtype=t.int64, device=D_parts.device)
    dev = D_parts.device.index
    N.check(N.lib.b2_merge_shard_lists(int(metric), nq, k, G, D_parts.data_ptr(), I_parts.data_ptr(), D.data_ptr(),
                                       I.data_ptr(), dev, _stream_for(dev)))
    return D, I
