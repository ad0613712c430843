"""GPU: b2_index_search_static — the search for callers that must not allocate.

Some users of the library cannot let a search reach the CUDA allocator: a service that hands
all of HBM to its own memory manager at start-up, or a request loop that cannot risk the
device-wide synchronisation an allocator call may cause.  For them the library offers one
extra entry point that keeps the scan kernel's per-CTA partial lists in a fixed 2 MiB array
inside the library image instead of taking them from cudaMallocAsync.

This file checks, through the C ABI only, that the entry point returns what b2_index_search
returns (and what the CPU oracle returns), on two calls in a row — the second call reuses the
array with the first call's contents still in it — that a call really leaves the device's
memory pool and free memory untouched, and that it refuses sizes beyond its fixed scratch
with a clear error.
The entry point always runs the FFMA scan kernel; b2_index_search runs the same kernel for
very small batches on a small index (then the comparison is bit for bit) and the tensor-core
scan for 5..16 queries (3..16 on bf16 rows), where the two agree within the 1e-5 relative
score tolerance every parity test here uses (tests/parity.py).  It is named test_zz_* so
that pytest collects it after every other test file.
"""
import ctypes

import numpy as np
import pytest

from oracle import flat
from parity import assert_parity

pytestmark = pytest.mark.gpu


class _Pool:
    """The device's default stream-ordered memory pool (where cudaMallocAsync takes its memory from), read through the
    CUDA runtime that torch has already loaded."""

    USED_CURRENT, USED_HIGH = 7, 8  # cudaMemPoolAttrUsedMemCurrent / cudaMemPoolAttrUsedMemHigh

    def __init__(self, device=0):
        self.rt = None
        for name in ("libcudart.so.12", "libcudart.so"):
            try:
                self.rt = ctypes.CDLL(name)
                break
            except OSError:
                continue
        if self.rt is None:
            pytest.skip("CUDA runtime library not loadable by name")
        self.pool = ctypes.c_void_p()
        assert self.rt.cudaDeviceGetDefaultMemPool(ctypes.byref(self.pool), device) == 0

    def get(self, attr):
        v = ctypes.c_uint64(0)
        assert self.rt.cudaMemPoolGetAttribute(self.pool, attr, ctypes.byref(v)) == 0
        return int(v.value)

    def reset_high(self):
        v = ctypes.c_uint64(0)
        assert self.rt.cudaMemPoolSetAttribute(self.pool, self.USED_HIGH, ctypes.byref(v)) == 0

    def free_bytes(self):
        free, total = ctypes.c_size_t(0), ctypes.c_size_t(0)
        assert self.rt.cudaMemGetInfo(ctypes.byref(free), ctypes.byref(total)) == 0
        return int(free.value)


def _dev_ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _search_pair(N, h, q, k, stream):
    """(D, I) from b2_index_search and from b2_index_search_static for the same device queries."""
    import torch

    nq = q.shape[0]
    D0 = torch.empty((nq, k), dtype=torch.float32, device=q.device)
    I0 = torch.empty((nq, k), dtype=torch.int64, device=q.device)
    D1 = torch.full((nq, k), float("nan"), dtype=torch.float32, device=q.device)
    I1 = torch.full((nq, k), -7, dtype=torch.int64, device=q.device)
    N.check(N.lib.b2_index_search(h, nq, _dev_ptr(q), N.DEVICE, k, _dev_ptr(D0), _dev_ptr(I0), N.DEVICE, stream))
    N.check(N.lib.b2_index_search_static(h, nq, _dev_ptr(q), k, _dev_ptr(D1), _dev_ptr(I1), stream))
    torch.cuda.synchronize()
    return D0.cpu().numpy(), I0.cpu().numpy(), D1.cpu().numpy(), I1.cpu().numpy()


@pytest.mark.parametrize("metric_name,dtype_name", [("ip", "f32"), ("l2", "f32"), ("ip", "bf16")])
def test_static_scratch_search_equals_the_allocating_search_twice_in_a_row(metric_name, dtype_name):
    import torch

    from nv_genai_b200 import _native as N

    dev = torch.device("cuda", 0)
    n, d = 60_000, 1024
    metric = N.METRIC_INNER_PRODUCT if metric_name == "ip" else N.METRIC_L2
    dtype = N.DTYPE_F32 if dtype_name == "f32" else N.DTYPE_BF16
    h = ctypes.c_void_p()
    N.check(N.lib.b2_index_create(d, metric, dtype, 0, ctypes.byref(h)))
    try:
        x = torch.empty((n, d), dtype=torch.float32, device=dev)
        N.check(N.lib.b2_synth_fill(_dev_ptr(x), 0, n, d, 0x5EED0001, 0, None))
        N.check(N.lib.b2_index_add(h, n, _dev_ptr(x), N.DEVICE, N.ADD_NORMALIZE, None))
        q = torch.empty((16, d), dtype=torch.float32, device=dev)
        N.check(N.lib.b2_synth_fill(_dev_ptr(q), 0, 16, d, 0x5EED0002, 0, None))
        N.check(N.lib.b2_normalize_l2(_dev_ptr(q), 16, d, 0, None))
        torch.cuda.synchronize()
        # the rows as the index holds them, for the oracle
        rows = torch.empty((n, d), dtype=torch.float32, device=dev)
        ids = torch.arange(n, dtype=torch.int64, device=dev)
        N.check(N.lib.b2_index_reconstruct_batch(h, n, _dev_ptr(ids), N.DEVICE, _dev_ptr(rows), N.DEVICE, None))
        torch.cuda.synchronize()
        x_host, q_host = rows.cpu().numpy(), q.cpu().numpy()
        del rows, ids

        # first call: 16 queries, k = 10 (BASELINE config 2 shape); second call: other queries, other k,
        # smaller tile, so it reads a differently shaped region of the array the first call filled
        for nq, k, first in ((16, 10, 0), (3, 100, 5)):
            qs = q[first:first + nq].contiguous()
            D0, I0, D1, I1 = _search_pair(N, h, qs, k, None)
            # below 5 queries on fp32 rows (3 on bf16 rows) b2_index_search runs the FFMA scan as well; from there on the
            # tensor-core scan (fp32 rows: L2 too, re-scored with the FFMA scan's direct form)
            same_kernel = nq < (5 if dtype_name == "f32" else 3)
            if same_kernel:
                assert np.array_equal(I0, I1) and np.array_equal(D0, D1), (nq, k)  # bit for bit
            else:
                assert_parity(D1, I1, D0, I0, metric_ip=(metric_name == "ip"))  # 1e-5 relative, ties may swap
            # and against the CPU oracle on the stored rows
            Dr, Ir, _ = flat.search_seq(x_host, q_host[first:first + nq], k,
                                        flat.METRIC_IP if metric_name == "ip" else flat.METRIC_L2)
            assert_parity(D1, I1, Dr, Ir, x_host, q_host[first:first + nq], metric_ip=(metric_name == "ip"))
            assert (I1 >= 0).all() and (I1 < n).all()
            if metric_name == "ip":
                assert np.all(np.diff(D1, axis=1) <= 0)
            else:
                assert np.all(np.diff(D1, axis=1) >= 0)
    finally:
        N.lib.b2_index_destroy(h)


def test_static_scratch_search_leaves_the_memory_pool_and_free_memory_untouched():
    """The property the entry point exists for, observed from outside: across a call of a shape that has run once
    before, the pool's high-water mark of used bytes does not rise and the device's free memory does not change.
    b2_index_search over the same queries does raise the mark, which shows that the probe sees such allocations."""
    import torch

    from nv_genai_b200 import _native as N

    dev = torch.device("cuda", 0)
    n, d, nq, k = 40_000, 512, 4, 10
    h = ctypes.c_void_p()
    N.check(N.lib.b2_index_create(d, N.METRIC_INNER_PRODUCT, N.DTYPE_F32, 0, ctypes.byref(h)))
    try:
        x = torch.empty((n, d), dtype=torch.float32, device=dev)
        N.check(N.lib.b2_synth_fill(_dev_ptr(x), 0, n, d, 0x5EED0001, 0, None))
        N.check(N.lib.b2_index_add(h, n, _dev_ptr(x), N.DEVICE, N.ADD_NORMALIZE, None))
        q = torch.empty((nq, d), dtype=torch.float32, device=dev)
        N.check(N.lib.b2_synth_fill(_dev_ptr(q), 0, nq, d, 0x5EED0002, 0, None))
        D = torch.empty((nq, k), dtype=torch.float32, device=dev)
        I = torch.empty((nq, k), dtype=torch.int64, device=dev)
        # once before: the first call of a shape loads the kernels and sets their launch attributes
        N.check(N.lib.b2_index_search_static(h, nq, _dev_ptr(q), k, _dev_ptr(D), _dev_ptr(I), None))
        N.check(N.lib.b2_index_search(h, nq, _dev_ptr(q), N.DEVICE, k, _dev_ptr(D), _dev_ptr(I), N.DEVICE, None))
        torch.cuda.synchronize()
        pool = _Pool(0)
        used0, free0 = pool.get(pool.USED_CURRENT), pool.free_bytes()
        pool.reset_high()
        for _ in range(3):
            N.check(N.lib.b2_index_search_static(h, nq, _dev_ptr(q), k, _dev_ptr(D), _dev_ptr(I), None))
        torch.cuda.synchronize()
        assert pool.get(pool.USED_HIGH) <= used0, "b2_index_search_static took memory from the pool"
        assert pool.get(pool.USED_CURRENT) == used0 and pool.free_bytes() == free0
        # the allocating entry point is visible to the same probe
        pool.reset_high()
        N.check(N.lib.b2_index_search(h, nq, _dev_ptr(q), N.DEVICE, k, _dev_ptr(D), _dev_ptr(I), N.DEVICE, None))
        torch.cuda.synchronize()
        assert pool.get(pool.USED_HIGH) > used0
    finally:
        N.lib.b2_index_destroy(h)


def test_static_scratch_search_rejects_what_does_not_fit_and_handles_small_indexes():
    import torch

    from nv_genai_b200 import _native as N

    dev = torch.device("cuda", 0)
    n, d = 70_000, 128
    h = ctypes.c_void_p()
    N.check(N.lib.b2_index_create(d, N.METRIC_L2, N.DTYPE_F32, 0, ctypes.byref(h)))
    try:
        q = torch.empty((17, d), dtype=torch.float32, device=dev)
        N.check(N.lib.b2_synth_fill(_dev_ptr(q), 0, 17, d, 0x5EED0002, 0, None))
        D = torch.empty((17, 2048), dtype=torch.float32, device=dev)
        I = torch.empty((17, 2048), dtype=torch.int64, device=dev)

        # empty index: padding only, as b2_index_search gives
        N.check(N.lib.b2_index_search_static(h, 2, _dev_ptr(q), 3, _dev_ptr(D), _dev_ptr(I), None))
        torch.cuda.synchronize()
        assert (I.view(-1)[:6].cpu().numpy() == -1).all()
        assert (D.view(-1)[:6].cpu().numpy() == np.finfo(np.float32).max).all()

        x = torch.empty((n, d), dtype=torch.float32, device=dev)
        N.check(N.lib.b2_synth_fill(_dev_ptr(x), 0, n, d, 0x5EED0001, 0, None))
        N.check(N.lib.b2_index_add(h, n, _dev_ptr(x), N.DEVICE, 0, None))

        # more than one scan pass of queries
        st = N.lib.b2_index_search_static(h, 17, _dev_ptr(q), 4, _dev_ptr(D), _dev_ptr(I), None)
        assert st == N.B2_ERR_UNSUPPORTED and b"16 queries" in N.lib.b2_last_error()
        # partial lists larger than the fixed scratch: 1 query x (hundreds of CTAs) x k = 1024
        st = N.lib.b2_index_search_static(h, 1, _dev_ptr(q), 1024, _dev_ptr(D), _dev_ptr(I), None)
        assert st == N.B2_ERR_UNSUPPORTED and b"exceed the fixed scratch" in N.lib.b2_last_error()
        # a query tile that does not fit one scan pass at this k
        st = N.lib.b2_index_search_static(h, 8, _dev_ptr(q), 2048, _dev_ptr(D), _dev_ptr(I), None)
        assert st == N.B2_ERR_UNSUPPORTED and b"one scan pass" in N.lib.b2_last_error()
        # bad arguments
        assert N.lib.b2_index_search_static(h, 1, None, 4, _dev_ptr(D), _dev_ptr(I), None) == N.B2_ERR_INVALID
        assert N.lib.b2_index_search_static(None, 1, _dev_ptr(q), 4, _dev_ptr(D), _dev_ptr(I), None) == N.B2_ERR_INVALID

        # and a size that fits still works after the refusals
        D0, I0, D1, I1 = _search_pair(N, h, q[:1].contiguous(), 4, None)
        assert np.array_equal(I0, I1) and np.array_equal(D0, D1)
    finally:
        N.lib.b2_index_destroy(h)
