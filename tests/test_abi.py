"""CPU: the C-ABI library loads without a GPU and exports every symbol include/b2_flat.h declares."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b2_flat.h")


def declared_symbols():
    src = open(HEADER).read()
    return sorted(set(re.findall(r"B2_API\s+[\w\s\*]+?\b(b2_\w+)\s*\(", src)))


def test_header_declares_the_expected_entry_points():
    syms = declared_symbols()
    for must in ("b2_index_create", "b2_index_add", "b2_index_search", "b2_index_mmr", "b2_merge_shard_lists",
                 "b2_index_remove_ids", "b2_index_merge_from", "b2_index_reconstruct_batch", "b2_last_error"):
        assert must in syms
    text = open(HEADER).read()
    assert "common/utils.py:217" in text and "main.py:90" in text  # every entry cites the interface it replaces


def test_library_exports_every_declared_symbol():
    from nv_genai_b200 import _native

    lib = ctypes.CDLL(_native.SO_PATH)
    for name in declared_symbols():
        assert hasattr(lib, name), f"{name} declared in b2_flat.h but not exported"
    assert sorted(_native.SYMBOLS) == declared_symbols()
    assert "sm_100a" in _native.version()


def test_no_cpu_fallback_without_a_gpu():
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from nv_genai_b200 import _native as N
    from nv_genai_b200 import IndexFlatL2

    h = ctypes.c_void_p()
    st = N.lib.b2_index_create(8, N.METRIC_L2, N.DTYPE_F32, 0, ctypes.byref(h))
    assert st == N.B2_ERR_NO_DEVICE and not h.value
    assert b"no CPU path" in N.lib.b2_last_error()
    with pytest.raises(N.B2Error):
        IndexFlatL2(8)


def test_argument_validation_happens_before_any_device_work():
    from nv_genai_b200 import _native as N

    h = ctypes.c_void_p()
    assert N.lib.b2_index_create(0, N.METRIC_L2, N.DTYPE_F32, 0, ctypes.byref(h)) == N.B2_ERR_INVALID
    assert N.lib.b2_index_create(8, 7, N.DTYPE_F32, 0, ctypes.byref(h)) == N.B2_ERR_INVALID
    assert N.lib.b2_index_create(8, N.METRIC_L2, 9, 0, ctypes.byref(h)) == N.B2_ERR_INVALID
    assert N.lib.b2_index_search(None, 1, None, 0, 1, None, None, 0, None) == N.B2_ERR_INVALID
    assert N.lib.b2_index_ntotal(None) == -1
    # the post-search selection entry points check their arguments first as well
    assert N.lib.b2_index_search_filtered(None, 1, None, 0, 1, 1, None, 0, 0, 0, 0.0, 0, 0.0, None, None, None, 0,
                                          None) == N.B2_ERR_INVALID
    assert N.lib.b2_filter_results(0, 1, 3, 4, None, None, None, 0, 0, 0.0, 0, 0.0, None, None, None, 0, None) \
        == N.B2_ERR_INVALID and b"fetch_k must be >= k" in N.lib.b2_last_error()
    assert N.lib.b2_filter_results(0, 1, 4, 4, None, None, None, 0, 0, 0.0, 7, 0.0, None, None, None, 0, None) \
        == N.B2_ERR_INVALID and b"relevance function" in N.lib.b2_last_error()
    assert N.lib.b2_index_destroy(None) == N.B2_OK


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "nv_genai_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, re.M), f
                assert "flat_oracle" not in text.replace("oracle/flat_oracle.c b2o_synth_fill", ""), f


def test_bench_and_oracle_use_the_same_synthetic_streams():
    """bench.py keeps its own copy of the two seeds (its GPU arm does not import the oracle)."""
    import importlib.util
    import os

    from oracle import flat

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("_bench", os.path.join(root, "bench.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    assert b.CORPUS_SEED == flat.CORPUS_SEED and b.QUERY_SEED == flat.QUERY_SEED
    assert set(b.CONFIGS) >= {"cfg1", "cfg2", "cfg2b16", "cfg3", "cfg3bf16", "cfg4", "cfg5"}


def test_shard_group_and_query_push_reject_bad_arguments_without_a_gpu():
    from nv_genai_b200 import _native as N

    g = ctypes.c_void_p()
    devs = (ctypes.c_int * 2)(0, 0)
    assert N.lib.b2_shard_group_create(0, devs, 8, N.METRIC_L2, N.DTYPE_F32, 100, ctypes.byref(g)) == N.B2_ERR_INVALID
    assert N.lib.b2_shard_group_create(2, devs, 8, N.METRIC_L2, N.DTYPE_F32, 0, ctypes.byref(g)) == N.B2_ERR_INVALID
    assert N.lib.b2_shard_group_create(17, devs, 8, N.METRIC_L2, N.DTYPE_F32, 100, ctypes.byref(g)) == N.B2_ERR_INVALID
    assert N.lib.b2_shard_search(None, 1, None, 0, 1, None, None, 0, None) == N.B2_ERR_INVALID
    assert N.lib.b2_shard_group_ntotal(None) == -1 and N.lib.b2_shard_group_nshards(None) == -1
    assert N.lib.b2_shard_group_add(None, 1, None, 0, 0) == N.B2_ERR_INVALID
    N.lib.b2_shard_group_destroy(None)  # a null handle is a no-op
    assert N.lib.b2_query_push(None, 16, 1, None, None, 1, 0, None) == N.B2_ERR_INVALID
    assert N.lib.b2_flag_wait(None, 1, 0, None) == N.B2_ERR_INVALID
    assert b"b2_flag_wait" in N.lib.b2_last_error()


def test_bench_check_helpers():
    """The comparison bench.py applies to what it timed: the parity rule in short form."""
    import importlib.util

    import numpy as np

    spec = importlib.util.spec_from_file_location("_bench2", os.path.join(ROOT, "bench.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    Dr = np.array([[0.9, 0.8, 0.7, 0.6]], np.float32)
    Ir = np.array([[5, 9, 2, 7]], np.int64)
    assert b.compare_topk(Dr, Ir, Dr, Ir) == "ok"
    assert b.compare_topk(Dr * (1 + 5e-6), Ir, Dr, Ir) == "ok"
    assert b.compare_topk(Dr * 1.001, Ir, Dr, Ir) != "ok"
    swapped = np.array([[9, 5, 2, 7]], np.int64)
    assert b.compare_topk(Dr, swapped, Dr, Ir) != "ok"  # 0.9 and 0.8 do not tie
    tie = np.array([[0.9, 0.9, 0.7, 0.6]], np.float32)
    assert b.compare_topk(tie, swapped, tie, Ir) == "ok"
    assert b.compare_topk(Dr[:, :3], Ir[:, :3], Dr, Ir).startswith("shape")
    ids = b.planted_ids(10_000_000, 8)
    assert len(ids) == 8 and ids[0] == 0 and 9_999_999 in ids and 1_249_999 in ids and 1_250_000 in ids
    assert b.planted_ids(10, 1) == [0, 1, 4, 5, 9, 3, 6, 2]
