"""GPU: parity of the tensor-core search (batches of >= 20 queries: tcgen05 GEMM with the top-k fused
into the TMEM read-out, b2_gemm.cu) with the oracle, through the C ABI.

fp32 row store (3xTF32): scores within 1e-5 relative of the oracle's BLAS-path restatement
(faiss exhaustive_*_blas: IP = q.x, L2 = |q|^2 + |x|^2 - 2 q.x clamped at 0) and of the float64 brute
force; ids equal except inside tolerance-wide ties (tests/parity.py).
bf16 row store (bf16 x (bf16 hi + bf16 lo)): the same rule against the oracle run on the bf16-rounded
rows, plus recall@k against the fp32 oracle.
"""
import numpy as np
import pytest

from oracle import flat
from parity import assert_parity, recall_at_k

pytestmark = pytest.mark.gpu


def _mk(metric, d, dtype="float32"):
    from nv_genai_b200 import IndexFlat

    return IndexFlat(d, metric, dtype)


def _data(n, d, nq, normalize, seed=0):
    x = flat.synth(0, n, d, flat.CORPUS_SEED + seed)
    q = flat.synth(0, nq, d, flat.QUERY_SEED + seed)
    if normalize:
        flat.normalize_L2(x)
        flat.normalize_L2(q)
    return x, q


def _launched_tensor_kernel(fn):
    """Run fn() and return (result, kernel launches).  One pass for the whole batch: the tensor-core scan is prep +
    scan + merge = 3 launches, the GEMM path in its default filter-and-refine mode 2 x prep + band + gemm + fix-up +
    merge = 6 (3 with an explicit number of query parts; 7 when the first large batch on fp32 rows also converts the
    rows to their bf16 shadow); ceil(nq/16) FFMA scan passes would be 2 per pass."""
    from nv_genai_b200 import _native

    _native.launch_count_reset()
    out = fn()
    return out, _native.launch_count()


@pytest.mark.parametrize("metric", [flat.METRIC_IP, flat.METRIC_L2])
@pytest.mark.parametrize("nq,k", [(20, 10), (33, 4), (128, 100), (129, 128), (300, 1)])
def test_fp32_store_matches_blas_oracle(metric, nq, k):
    n, d = 30_011, 256  # ragged: last corpus tile has 59 rows
    x, q = _data(n, d, nq, True, seed=nq)
    idx = _mk(metric, d)
    idx.add(x)
    (D, I), launches = _launched_tensor_kernel(lambda: idx.search(q, k))
    assert launches in (3, 5, 6, 7), launches  # one pass for the whole batch, not ceil(nq/16) scan passes
    Dr, Ir = flat.search_blas(x, q, k, metric)
    assert_parity(D, I, Dr, Ir, x, q, metric_ip=(metric == flat.METRIC_IP))
    D64, I64 = flat.search_f64(x, q, k, metric)
    assert_parity(D, I, D64.astype(np.float32), I64, x, q, metric_ip=(metric == flat.METRIC_IP))
    if metric == flat.METRIC_IP:
        assert np.all(np.diff(D, axis=1) <= 0)
    else:
        assert np.all(np.diff(D, axis=1) >= 0)


@pytest.mark.parametrize("d", [16, 100, 384, 1024, 1536])
def test_dimensions_fp32_unnormalised_l2_and_ip(d):
    n, nq, k = 9_000, 40, 10
    x, q = _data(n, d, nq, False, seed=d)
    for metric in (flat.METRIC_IP, flat.METRIC_L2):
        idx = _mk(metric, d)
        idx.add(x)
        D, I = idx.search(q, k)
        Dr, Ir = flat.search_blas(x, q, k, metric)
        assert_parity(D, I, Dr, Ir, x, q, metric_ip=(metric == flat.METRIC_IP))


@pytest.mark.parametrize("metric", [flat.METRIC_IP, flat.METRIC_L2])
def test_bf16_store_matches_oracle_on_rounded_rows(metric):
    n, d, nq, k = 50_000, 1024, 64, 100
    x, q = _data(n, d, nq, True, seed=7)
    idx = _mk(metric, d, "bfloat16")
    idx.add(x)
    (D, I), launches = _launched_tensor_kernel(lambda: idx.search(q, k))
    assert launches in (3, 5, 6, 7)
    xr = flat.round_bf16(x)
    Dr, Ir = flat.search_blas(xr, q, k, metric)
    assert_parity(D, I, Dr, Ir, xr, q, metric_ip=(metric == flat.METRIC_IP))
    _, I32 = flat.search_blas(x, q, k, metric)
    assert recall_at_k(I, I32) >= 0.99


def test_ties_duplicates_keep_the_lowest_ids_across_segments_and_tiles():
    """Planted exact duplicates of the queries at known ids spread over many tiles/segments: all copies tie,
    and the k returned must be the k lowest ids, in id order."""
    n, d, nq, k = 40_000, 128, 24, 8
    x, q = _data(n, d, nq, True, seed=3)
    planted = np.array([5, 231, 256, 487, 9_000, 9_100, 20_000, 25_000, 31_111, 39_900])  # + r, r < 24: all distinct
    for r in range(nq):
        x[planted + r] = q[r]
    idx = _mk(flat.METRIC_IP, d)
    idx.add(x)
    D, I = idx.search(q, k)
    D64, I64 = flat.search_f64(x, q, k, flat.METRIC_IP)
    for r in range(nq):
        want = np.sort(planted + r)[:k]
        got = I[r]
        # all copies have (nearly) the same score; parity accepts reorderings within tolerance, but the SET
        # must be the k lowest ids only if scores tie exactly -- they do: identical rows, identical arithmetic
        assert sorted(got.tolist()) == sorted(want.tolist()), (r, got, want)
        assert np.array_equal(got, want), (r, got, want)  # equal keys -> lowest id first
    assert_parity(D, I, D64.astype(np.float32), I64, x, q, metric_ip=True)


def test_all_equal_rows_and_k_larger_than_survivors():
    n, d, nq, k = 5_000, 64, 32, 16
    x = np.ones((n, d), dtype=np.float32)
    q = np.ones((nq, d), dtype=np.float32)
    idx = _mk(flat.METRIC_IP, d)
    idx.add(x)
    D, I = idx.search(q, k)
    assert np.array_equal(I, np.tile(np.arange(k), (nq, 1)))
    assert np.allclose(D, float(d))


def test_ascending_scores_force_candidate_log_compaction():
    """Scores grow with the row id, so every row beats the running k-th: the per-query candidate log overflows
    and is compacted many times per segment.  The answer is the last k rows, best first."""
    n, d, nq, k = 20_000, 64, 20, 16
    x = np.ones((n, d), dtype=np.float32) * (np.arange(1, n + 1, dtype=np.float32) / n)[:, None]
    q = np.ones((nq, d), dtype=np.float32)
    idx = _mk(flat.METRIC_IP, d)
    idx.add(x)
    D, I = idx.search(q, k)
    Dr, Ir = flat.search_blas(x, q, k, flat.METRIC_IP)
    assert_parity(D, I, Dr, Ir, x, q, metric_ip=True)
    assert set(I[0].tolist()) <= set(range(n - 4 * k, n))


def test_nan_rows_never_enter_and_results_stay_sorted():
    n, d, nq, k = 8_192, 64, 20, 10
    x, q = _data(n, d, nq, True, seed=11)
    x[100] = np.nan
    x[4097] = np.nan
    idx = _mk(flat.METRIC_IP, d)
    idx.add(x)
    D, I = idx.search(q, k)
    assert not np.isin(I, [100, 4097]).any()
    assert np.isfinite(D).all() and np.all(np.diff(D, axis=1) <= 0)


def test_mmr_batch_uses_tensor_path_and_matches_oracle():
    """BASELINE config 5 shape at a small size: batch of 64 queries, fetch_k = 20, k = 4."""
    n, d, nq = 20_000, 256, 64
    x, q = _data(n, d, nq, True, seed=5)
    idx = _mk(flat.METRIC_IP, d)
    idx.add(x)
    ids, _ = idx.mmr(q, k=4, fetch_k=20, lambda_mult=0.5)
    bad = 0
    for r in range(nq):
        want, _ = flat.mmr_search(x, q[r], 4, 20, 0.5, flat.METRIC_IP)
        bad += int(ids[r].tolist() != want)
    assert bad == 0, bad


def test_large_batch_1024_queries_k100_property():
    """BASELINE config 3/4 shape at reduced N: sorted output, ids unique, and equal to the scan path (<= 16
    queries per pass, direct form) on a sample of the queries."""
    n, d, nq, k = 200_000, 1024, 1024, 100
    x, q = _data(n, d, nq, True, seed=9)
    idx = _mk(flat.METRIC_IP, d)
    idx.add(x)
    D, I = idx.search(q, k)
    assert np.all(np.diff(D, axis=1) <= 0)
    assert all(len(set(row.tolist())) == k for row in I)
    sample = [0, 1, 127, 128, 500, 1023]
    Ds, Is = idx.search(q[sample], k)  # 6 queries -> scan kernel
    assert_parity(D[sample], I[sample], Ds, Is, x, q[sample], metric_ip=True)


# ---------------------------------------------------------------------------------------------------------------
# tensor-core scan (b2_scan_tc.cu): 3..16 queries, inner product, corpus as the M operand
@pytest.mark.parametrize("dtype", ["float32", "bfloat16"])
@pytest.mark.parametrize("nq,k", [(6, 10), (9, 1), (16, 10), (16, 100), (5, 128)])
def test_small_batch_tensor_scan_matches_sequential_oracle(dtype, nq, k):
    n, d = 70_001, 1024  # ragged last tile (128-row tiles)
    x, q = _data(n, d, nq, True, seed=100 + nq)
    idx = _mk(flat.METRIC_IP, d, dtype)
    idx.add(x)
    (D, I), launches = _launched_tensor_kernel(lambda: idx.search(q, k))
    # bf16 rows, or k > 64: prep + scan_tc + merge (the FFMA scan is 2).  fp32 rows, k <= 64: filter and refine =
    # one preparation kernel + approximate scan_tc + merge/re-score, then the fix-up pair that returns at once
    assert launches == (5 if dtype == "float32" and k <= 64 else 3), launches
    xr = flat.round_bf16(x) if dtype == "bfloat16" else x
    Dr, Ir, _ = flat.search_seq(xr, q, k, flat.METRIC_IP)
    assert_parity(D, I, Dr, Ir, xr, q, metric_ip=True)
    assert np.all(np.diff(D, axis=1) <= 0)


@pytest.mark.parametrize("dtype", ["float32", "bfloat16"])
@pytest.mark.parametrize("nq,k", [(17, 10), (32, 32), (33, 4), (64, 20), (20, 64), (32, 33)])
def test_medium_batch_rides_one_tensor_scan_pass_when_k_is_small(dtype, nq, k):
    """Up to 32 queries share one pass over the rows for k <= 64 (query tile of 32); beyond that the GEMM kernel."""
    n, d = 50_003, 512
    x, q = _data(n, d, nq, True, seed=200 + nq)
    idx = _mk(flat.METRIC_IP, d, dtype)
    idx.add(x)
    (D, I), launches = _launched_tensor_kernel(lambda: idx.search(q, k))
    assert launches in (3, 5, 6, 7), launches  # one pass over the rows either way
    xr = flat.round_bf16(x) if dtype == "bfloat16" else x
    Dr, Ir = flat.search_f64(xr, q, k, flat.METRIC_IP)
    assert_parity(D, I, Dr.astype(np.float32), Ir, xr, q, metric_ip=True)


def test_small_batch_tensor_scan_ties_and_dims():
    """Planted duplicates (lowest id first) and a dimension that is not a multiple of the 128-byte K slice."""
    n, d, nq, k = 30_000, 200, 7, 6
    x, q = _data(n, d, nq, True, seed=21)
    planted = np.array([3, 121, 128, 4_000, 12_345, 20_000, 29_999 - 7])  # + r, r < 7: all distinct; 121+6 and 128 straddle a tile
    for r in range(nq):
        x[planted + r] = q[r]
    idx = _mk(flat.METRIC_IP, d)
    idx.add(x)
    D, I = idx.search(q, k)
    for r in range(nq):
        assert np.array_equal(I[r], np.sort(planted + r)[:k]), (r, I[r])
    Dr, Ir, _ = flat.search_seq(x, q, k, flat.METRIC_IP)
    assert_parity(D, I, Dr, Ir, x, q, metric_ip=True)


def test_small_batch_l2_keeps_the_direct_form():
    """L2 below 20 queries: faiss's direct form sum (x - q)^2, so a stored copy of the query is at distance 0.0 exactly.
    The tensor-core scan filters with 2 q.x - |x|^2 and the merge re-scores in the direct form; k > 64 stays on the FFMA
    scan."""
    n, d, nq = 10_000, 256, 8
    x, q = _data(n, d, nq, False, seed=31)
    x[77] = q[2]
    idx = _mk(flat.METRIC_L2, d)
    idx.add(x)
    (D, I), launches = _launched_tensor_kernel(lambda: idx.search(q, 4))
    assert launches == 5, launches  # preparation + approximate scan + merge/re-score, FFMA fix-up pair (returns at once)
    assert I[2, 0] == 77 and D[2, 0] == 0.0
    Dr, Ir, _ = flat.search_seq(x, q, 4, flat.METRIC_L2)
    assert_parity(D, I, Dr, Ir, x, q, metric_ip=False)
    assert np.all(np.diff(D, axis=1) >= 0)
    (D, I), launches = _launched_tensor_kernel(lambda: idx.search(q, 100))  # k > 64: FFMA scan
    assert launches == 2 and I[2, 0] == 77 and D[2, 0] == 0.0
    xb = _mk(flat.METRIC_L2, d, "bfloat16")  # bf16 rows: one bf16 product as the filter, same re-scoring
    xb.add(x)
    (D, I), launches = _launched_tensor_kernel(lambda: xb.search(q, 4))
    assert launches == 5 and I[2, 0] == 77
    xr = flat.round_bf16(x)
    Dr, Ir, _ = flat.search_seq(xr, q, 4, flat.METRIC_L2)
    assert_parity(D, I, Dr, Ir, xr, q, metric_ip=False)


@pytest.mark.parametrize("dtype", ["float32", "bfloat16"])
@pytest.mark.parametrize("nq,k", [(5, 10), (16, 10), (19, 64), (13, 1)])
def test_small_batch_l2_on_the_tensor_scan_matches_the_sequential_oracle(nq, k, dtype):
    """Un-normalised rows (norms ~ d), planted copies, ragged d: ids and distances equal the direct-form oracle; more
    near-ties than a CTA list holds take the FFMA fix-up."""
    n, d = 40_000, 200
    x, q = _data(n, d, nq, False, seed=300 + nq)
    x[1234] = q[0]
    x[2000:2090] = q[1]  # 90 copies: overflow -> exact fix-up for that query
    idx = _mk(flat.METRIC_L2, d, dtype)
    idx.add(x)
    D, I = idx.search(q, k)
    xr = flat.round_bf16(x) if dtype == "bfloat16" else x
    assert I[0, 0] == 1234 and np.array_equal(I[1, :min(k, 90)], np.arange(2000, 2000 + min(k, 90)))
    if dtype == "float32":
        assert D[0, 0] == 0.0 and np.all(D[1, :min(k, 90)] == 0.0)
    Dr, Ir, _ = flat.search_seq(xr, q, k, flat.METRIC_L2)
    assert_parity(D, I, Dr, Ir, xr, q, metric_ip=False)
    assert np.all(np.diff(D, axis=1) >= 0)


def test_set_option_single_query_part_is_approximate_but_close():
    n, d, nq, k = 60_000, 1024, 128, 100
    x, q = _data(n, d, nq, True, seed=41)
    idx = _mk(flat.METRIC_IP, d, "bfloat16")
    idx.add(x)
    D2, I2 = idx.search(q, k)
    idx.set_option("gemm_query_parts", 1)
    D1, I1 = idx.search(q, k)
    idx.set_option("gemm_query_parts", 0)
    D0, I0 = idx.search(q, k)
    assert np.array_equal(I0, I2) and np.array_equal(D0, D2)  # 0 restores the default
    assert recall_at_k(I1, I2) >= 0.97
    assert np.allclose(D1, D2, rtol=5e-3)
    assert not np.array_equal(D1, D2)  # it really is a different operand
    with pytest.raises(ValueError):
        idx.set_option("gemm_query_parts", 3)
    with pytest.raises(ValueError):
        idx.set_option("no_such_option", 1)


def test_concurrent_readers_across_all_three_kernels():
    """Many host threads searching one index at once with batch sizes that take the FFMA scan, the tensor-core scan
    and the GEMM kernel (each call has its own scratch; static launch state is guarded)."""
    from concurrent.futures import ThreadPoolExecutor

    n, d = 40_000, 256
    x, q = _data(n, d, 96, True, seed=51)
    idx = _mk(flat.METRIC_IP, d)
    idx.add(x)
    jobs = [(0, 1, 5), (1, 16, 10), (17, 24, 10), (41, 40, 100), (2, 3, 7), (50, 33, 100), (10, 12, 50), (30, 64, 128)] * 3
    serial = [idx.search(q[a:a + m], k) for a, m, k in jobs]
    with ThreadPoolExecutor(8) as ex:
        par = list(ex.map(lambda j: idx.search(q[j[0]:j[0] + j[1]], j[2]), jobs))
    for (Ds, Is), (Dp, Ip) in zip(serial, par):
        assert np.array_equal(Is, Ip) and np.array_equal(Ds, Dp)


@pytest.mark.parametrize("metric", [flat.METRIC_IP, flat.METRIC_L2])
def test_bf16_shadow_of_fp32_rows_filters_and_the_fp32_rows_decide(metric):
    """Large batches on fp32 rows filter with one bf16 product per score over a lazily built bf16 copy of the rows; every
    candidate inside the (wider) error band is re-scored from the fp32 rows, so the results equal those without the
    shadow bit for bit.  The shadow follows add / remove_ids / reset."""
    n, d, nq = 40_000, 384, 70
    x, q = _data(n + 5_000, d, nq, metric == flat.METRIC_IP, seed=77)
    planted = np.array([5, 255, 300, 9_000, 39_990])
    for r in range(3):  # exact copies of a query: score 1.0 / distance ~0 must survive the bf16 filter
        x[planted + r] = q[r]
    idx = _mk(metric, d)
    idx.add(x[:n])

    def both(k):
        idx.set_option("bf16_shadow", 1)
        (D1, I1), l1 = _launched_tensor_kernel(lambda: idx.search(q, k))
        idx.set_option("bf16_shadow", 0)
        (D0, I0), l0 = _launched_tensor_kernel(lambda: idx.search(q, k))
        idx.set_option("bf16_shadow", 1)
        assert l1 == 7 and l0 == 6, (l1, l0)  # shadow rebuilt after it was dropped: + the conversion kernel
        assert np.array_equal(I1, I0) and np.array_equal(D1, D0)
        return D1, I1

    for k in (10, 100):
        D, I = both(k)
        Dr, Ir = flat.search_f64(x[:n], q, k, metric)
        assert_parity(D, I, Dr.astype(np.float32), Ir, x[:n], q, metric_ip=(metric == flat.METRIC_IP))
        for r in range(3):
            assert np.array_equal(np.sort(I[r, :5]), planted + r)
    # a second search reuses the shadow: no conversion launch
    idx.search(q, 10)
    _, launches = _launched_tensor_kernel(lambda: idx.search(q, 10))
    assert launches == 6
    # add: only the new rows are converted, results cover them
    x[n + 17] = q[5]
    idx.add(x[n:])
    (D, I), launches = _launched_tensor_kernel(lambda: idx.search(q, 10))
    assert launches == 7 and I[5, 0] == n + 17
    Dr, Ir = flat.search_f64(x, q, 10, metric)
    assert_parity(D, I, Dr.astype(np.float32), Ir, x, q, metric_ip=(metric == flat.METRIC_IP))
    # remove_ids shifts rows: the shadow is reconverted from the first removed row on
    gone = np.array([100, 101, 20_000, n + 17])
    assert idx.remove_ids(gone) == 4
    keep = np.ones(len(x), bool)
    keep[gone] = False
    (D, I), launches = _launched_tensor_kernel(lambda: idx.search(q, 10))
    assert launches == 7
    Dr, Ir = flat.search_f64(x[keep], q, 10, metric)
    assert_parity(D, I, Dr.astype(np.float32), Ir, x[keep], q, metric_ip=(metric == flat.METRIC_IP))
    idx.reset()
    idx.add(x[:8_000])
    D, I = idx.search(q, 10)
    Dr, Ir = flat.search_f64(x[:8_000], q, 10, metric)
    assert_parity(D, I, Dr.astype(np.float32), Ir, x[:8_000], q, metric_ip=(metric == flat.METRIC_IP))


@pytest.mark.parametrize("dtype,nq", [("float32", 8), ("float32", 70), ("bfloat16", 70)])
def test_more_near_ties_than_the_candidate_lists_hold_take_the_exact_fix_up(dtype, nq):
    """Filter-and-refine keeps k + slack approximate candidates per list.  90 copies of one query packed into
    neighbouring rows put more candidates inside the error band of that query's k-th best than a list holds: the
    query (scan) or its query group (GEMM) is flagged and redone by the exact kernel.  The answer must still be the
    lowest ids among the copies, and the other queries must be unaffected."""
    n, d, k = 30_000, 256, 10
    x, q = _data(n, d, nq, True, seed=91)
    copies = np.arange(1000, 1090)
    x[copies] = q[0]
    x[20_000:20_005] = q[1]  # a second query with a handful of ties: stays on the normal path
    idx = _mk(flat.METRIC_IP, d, dtype)
    idx.add(x)
    D, I = idx.search(q, k)
    assert np.array_equal(I[0], copies[:k]), I[0]
    assert np.array_equal(I[1, :5], np.arange(20_000, 20_005))
    xr = flat.round_bf16(x) if dtype == "bfloat16" else x
    Dr, Ir = flat.search_f64(xr, q, k, flat.METRIC_IP)
    assert_parity(D, I, Dr.astype(np.float32), Ir, xr, q, metric_ip=True)
    # k = 60: still more copies than k, list = 64 (scan) / k + slack (GEMM)
    D, I = idx.search(q, 60)
    assert np.array_equal(I[0], copies[:60])


@pytest.mark.parametrize("dtype", ["float32", "bfloat16", "bfloat16+residual"])
def test_overflowing_lists_in_the_two_launch_search_take_the_fix_up(dtype):
    """k >= 32 on >= 2048 corpus tiles: threshold pre-pass over every 32nd tile, union, main pass.  Hundreds of exact
    copies of one query in neighbouring rows overflow a main-pass list; 256 copies of another query filling a SAMPLE
    tile overflow a pre-pass list.  The flagged query group is redone (all-exact tensor kernel over all tiles and
    all list slots; on a residual-plane store the whole batch by the two-plane FFMA scan) and every query must still
    match the float64 oracle, the copies coming back lowest id first."""
    n, d, nq, k = 600_000, 256, 40, 50
    x, q = _data(n, d, nq, True, seed=93)
    main_copies = np.arange(100_000, 100_400)          # tiles 390..392: not sample tiles
    sample_copies = np.arange(416 * 256, 417 * 256)    # tile 416 = 13 * 32: a tile of the pre-pass
    x[main_copies] = q[0]
    x[sample_copies] = q[1]
    idx = _mk(flat.METRIC_IP, d, dtype)
    idx.add(x)
    D, I = idx.search(q, k)
    assert np.array_equal(I[0], main_copies[:k]), I[0][:8]
    assert np.array_equal(I[1], sample_copies[:k]), I[1][:8]
    if dtype == "float32":
        xr = x
    else:
        xr = flat.round_bf16(x)
        if dtype.endswith("+residual"):
            xr = xr + flat.round_bf16(x - xr)
    Dr, Ir = flat.search_f64(xr, q, k, flat.METRIC_IP)
    assert_parity(D, I, Dr.astype(np.float32), Ir, xr, q, metric_ip=True)
