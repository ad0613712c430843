"""ctypes binding of libb2flat.so — the C ABI declared in include/b2_flat.h.

There is no CPU fallback: if the shared library is missing the import of this module
fails loudly, and every entry point fails with B2_ERR_NO_DEVICE when no B200 is visible.
"""
from __future__ import annotations

import ctypes
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
# B2FLAT_SO: load another build of the same ABI (A/B timing of two builds on one box); default = the in-tree build
SO_PATH = os.environ.get("B2FLAT_SO") or os.path.join(_PKG, "lib", "libb2flat.so")

B2_OK, B2_ERR_INVALID, B2_ERR_CUDA, B2_ERR_OOM, B2_ERR_UNSUPPORTED, B2_ERR_NO_DEVICE = range(6)
METRIC_INNER_PRODUCT = 0  # same values as faiss.METRIC_INNER_PRODUCT / METRIC_L2
METRIC_L2 = 1
DTYPE_F32, DTYPE_BF16 = 0, 1
HOST, DEVICE = 0, 1
ADD_NORMALIZE = 1
MAX_K = 2048
REL_NONE, REL_EUCLIDEAN, REL_COSINE, REL_MAX_INNER_PRODUCT = 0, 1, 2, 3  # langchain's relevance-score functions

# Every symbol include/b2_flat.h declares (tests check the library exports all of them).
SYMBOLS = [
    "b2_index_create", "b2_index_destroy", "b2_index_ntotal", "b2_index_d", "b2_index_metric", "b2_index_dtype",
    "b2_index_device", "b2_index_add", "b2_index_search", "b2_index_search_static", "b2_index_search_filtered", "b2_filter_results", "b2_index_reconstruct_batch", "b2_index_remove_ids",
    "b2_index_merge_from", "b2_index_set_option", "b2_index_reset", "b2_index_mmr", "b2_merge_shard_lists", "b2_merge_shard_lists_packed", "b2_index_search_push", "b2_merge_shard_lists_wait", "b2_query_push", "b2_flag_wait", "b2_shard_group_create", "b2_shard_group_destroy", "b2_shard_group_ntotal", "b2_shard_group_nshards", "b2_shard_group_shard_ntotal", "b2_shard_group_add", "b2_shard_group_remove_ids", "b2_shard_group_reset", "b2_shard_group_reconstruct_batch", "b2_shard_group_set_option", "b2_shard_search", "b2_normalize_l2",
    "b2_synth_fill", "b2_launch_count", "b2_launch_count_reset", "b2_prof_enable", "b2_prof_read", "b2_prof_read_kernel", "b2_last_error",
    "b2_version",
]


class B2Error(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(message)
        self.status = status


def _load() -> ctypes.CDLL:
    if not os.path.exists(SO_PATH):
        raise ImportError(
            f"{SO_PATH} is missing: build it with `python -m nv_genai_b200.build` "
            "(nvcc, sm_100a). nv_genai_b200 has no CPU or PyTorch fallback.")
    L = ctypes.CDLL(SO_PATH)
    vp, i64, i32, u32, u64, f32 = (ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_uint, ctypes.c_uint64,
                                   ctypes.c_float)
    L.b2_index_create.argtypes = [i32, i32, i32, i32, ctypes.POINTER(vp)]
    L.b2_index_destroy.argtypes = [vp]
    for name in ("b2_index_d", "b2_index_metric", "b2_index_dtype", "b2_index_device"):
        getattr(L, name).argtypes = [vp]
        getattr(L, name).restype = i32
    L.b2_index_ntotal.argtypes = [vp]
    L.b2_index_ntotal.restype = i64
    L.b2_index_add.argtypes = [vp, i64, vp, i32, u32, vp]
    L.b2_index_search.argtypes = [vp, i64, vp, i32, i32, vp, vp, i32, vp]
    L.b2_index_search_static.argtypes = [vp, i64, vp, i32, vp, vp, vp]
    L.b2_index_search_static.restype = i32
    f64 = ctypes.c_double
    L.b2_index_search_filtered.argtypes = [vp, i64, vp, i32, i32, i32, vp, i64, i32, i32, f64, i32, f64, vp, vp, vp, i32, vp]
    L.b2_index_search_filtered.restype = i32
    L.b2_filter_results.argtypes = [i32, i64, i32, i32, vp, vp, vp, i64, i32, f64, i32, f64, vp, vp, vp, i32, vp]
    L.b2_filter_results.restype = i32
    L.b2_index_reconstruct_batch.argtypes = [vp, i64, vp, i32, vp, i32, vp]
    L.b2_index_remove_ids.argtypes = [vp, i64, vp, ctypes.POINTER(i64)]
    L.b2_index_merge_from.argtypes = [vp, vp]
    L.b2_index_set_option.argtypes = [vp, ctypes.c_char_p, i64]
    L.b2_index_set_option.restype = i32
    L.b2_index_reset.argtypes = [vp]
    L.b2_index_mmr.argtypes = [vp, i64, vp, i32, i32, i32, f32, vp, vp, i32, vp]
    L.b2_merge_shard_lists.argtypes = [i32, i64, i32, i32, vp, vp, vp, vp, i32, vp]
    L.b2_merge_shard_lists_packed.argtypes = [i32, i64, i32, i32, vp, vp, i64, vp, vp, vp, i32, vp]
    L.b2_merge_shard_lists_packed.restype = i32
    L.b2_index_search_push.argtypes = [vp, i64, vp, i32, i32, vp, vp, u32, vp]
    L.b2_index_search_push.restype = i32
    L.b2_merge_shard_lists_wait.argtypes = [i32, i64, i32, i32, vp, i64, vp, i64, u32, vp, vp, vp, i32, vp]
    L.b2_merge_shard_lists_wait.restype = i32
    L.b2_query_push.argtypes = [vp, i64, i32, vp, vp, u32, i32, vp]
    L.b2_query_push.restype = i32
    L.b2_flag_wait.argtypes = [vp, u32, i32, vp]
    L.b2_flag_wait.restype = i32
    L.b2_shard_group_create.argtypes = [i32, ctypes.POINTER(i32), i32, i32, i32, i64, ctypes.POINTER(vp)]
    L.b2_shard_group_create.restype = i32
    L.b2_shard_group_destroy.argtypes = [vp]
    L.b2_shard_group_destroy.restype = None
    L.b2_shard_group_ntotal.argtypes = [vp]
    L.b2_shard_group_ntotal.restype = i64
    L.b2_shard_group_nshards.argtypes = [vp]
    L.b2_shard_group_nshards.restype = i32
    L.b2_shard_group_shard_ntotal.argtypes = [vp, i32]
    L.b2_shard_group_shard_ntotal.restype = i64
    L.b2_shard_group_add.argtypes = [vp, i64, vp, i32, u32]
    L.b2_shard_group_add.restype = i32
    L.b2_shard_group_remove_ids.argtypes = [vp, i64, vp, ctypes.POINTER(i64)]
    L.b2_shard_group_remove_ids.restype = i32
    L.b2_shard_group_reset.argtypes = [vp]
    L.b2_shard_group_reset.restype = i32
    L.b2_shard_group_reconstruct_batch.argtypes = [vp, i64, vp, vp]
    L.b2_shard_group_reconstruct_batch.restype = i32
    L.b2_shard_group_set_option.argtypes = [vp, ctypes.c_char_p, i64]
    L.b2_shard_group_set_option.restype = i32
    L.b2_shard_search.argtypes = [vp, i64, vp, i32, i32, vp, vp, i32, vp]
    L.b2_shard_search.restype = i32
    L.b2_normalize_l2.argtypes = [vp, i64, i32, i32, vp]
    L.b2_synth_fill.argtypes = [vp, i64, i64, i32, u64, i32, vp]
    L.b2_launch_count.argtypes = []
    L.b2_launch_count.restype = i64
    L.b2_launch_count_reset.argtypes = []
    L.b2_launch_count_reset.restype = None
    L.b2_prof_enable.argtypes = [i32]
    L.b2_prof_enable.restype = i32
    L.b2_prof_read.argtypes = [ctypes.POINTER(ctypes.c_double), ctypes.POINTER(i64), i32]
    L.b2_prof_read.restype = i32
    L.b2_prof_read_kernel.argtypes = [i32, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(i64), i32]
    L.b2_prof_read_kernel.restype = i32
    L.b2_last_error.argtypes = []
    L.b2_last_error.restype = ctypes.c_char_p
    L.b2_version.argtypes = []
    L.b2_version.restype = ctypes.c_char_p
    for name in ("b2_index_create", "b2_index_destroy", "b2_index_add", "b2_index_search",
                 "b2_index_reconstruct_batch", "b2_index_remove_ids", "b2_index_merge_from", "b2_index_set_option", "b2_index_reset",
                 "b2_index_mmr", "b2_merge_shard_lists", "b2_merge_shard_lists_packed", "b2_index_search_push", "b2_merge_shard_lists_wait", "b2_normalize_l2", "b2_synth_fill"):
        getattr(L, name).restype = i32
    return L


lib = _load()


def check(status: int) -> None:
    """Map a b2_status to the exception type the reference's callers would see from faiss/SWIG."""
    if status == B2_OK:
        return
    msg = lib.b2_last_error().decode("utf-8", "replace")
    if status == B2_ERR_INVALID:
        raise ValueError(msg)
    if status == B2_ERR_OOM:
        raise MemoryError(msg)
    if status == B2_ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    raise B2Error(status, msg)


def launch_count() -> int:
    return int(lib.b2_launch_count())


def launch_count_reset() -> None:
    lib.b2_launch_count_reset()


def prof_enable(on: bool) -> None:
    lib.b2_prof_enable(1 if on else 0)


PROF_SCAN, PROF_GEMM = 0, 1


def prof_read(reset: bool = True, kernel: int = PROF_SCAN):
    """(summed kernel milliseconds, launches) of the scan or the tensor-core kernel since the last reset."""
    ms, n = ctypes.c_double(0.0), ctypes.c_int64(0)
    lib.b2_prof_read_kernel(kernel, ctypes.byref(ms), ctypes.byref(n), 1 if reset else 0)
    return float(ms.value), int(n.value)


def version() -> str:
    return lib.b2_version().decode()
