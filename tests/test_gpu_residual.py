"""GPU: bf16 rows with a residual plane (dtype "bfloat16+residual"): the index stands for hi + lo with hi = bf16(x),
lo = bf16(x - hi).  Every path - two-plane FFMA scan, tensor-core scan (filter with the main plane, re-score with
both), tensor-core GEMM (the same), reconstruct, remove_ids, merge_from - must behave as if the rows hi + lo were
stored in fp32: parity with the oracle on those rows, and recall against the fp32 embeddings >= 0.999 (north_star),
where the plain bf16 store reaches ~0.997."""
import numpy as np
import pytest

from oracle import flat
from parity import assert_parity, recall_at_k

pytestmark = pytest.mark.gpu


def _planes(x):
    hi = flat.round_bf16(x)
    lo = flat.round_bf16(x - hi)
    return hi, lo, hi + lo


@pytest.mark.parametrize("metric", [flat.METRIC_IP, flat.METRIC_L2])
def test_residual_store_behaves_like_fp32_rows_hi_plus_lo(metric):
    from nv_genai_b200 import IndexFlat

    n, d = 200_000, 1024
    x = flat.synth(0, n, d, 41)
    flat.normalize_L2(x)
    q = flat.synth(0, 64, d, 42)
    flat.normalize_L2(q)
    q[7] = x[123_456]
    _, _, eff = _planes(x)
    idx = IndexFlat(d, metric, "bfloat16+residual")
    idx.add(x[:120_000])
    idx.add(x[120_000:])
    assert idx.dtype == "bfloat16+residual"
    got = idx.reconstruct_n(0, n)
    assert np.array_equal(got, eff)  # reconstruct returns hi + lo, exactly
    ip = metric == flat.METRIC_IP
    for nq, k in ((1, 10), (2, 100), (16, 10), (32, 10), (64, 10), (64, 100), (3, 200)):
        D, I = idx.search(q[:nq], k)
        Dr, Ir = flat.search(eff, q[:nq], k, metric)
        assert_parity(D, I, Dr, Ir, eff, q[:nq], metric_ip=ip)
        # against the fp32 embeddings themselves
        Df, If = flat.search(x, q[:nq], k, metric)
        assert recall_at_k(I, If) >= 0.999, (nq, k, recall_at_k(I, If))
    D, I = idx.search(q[:16], 10)
    assert I[7, 0] == 123_456


def test_residual_store_recall_beats_plain_bf16_and_mutations_keep_the_planes_aligned():
    from nv_genai_b200 import IndexFlatIP

    n, d = 300_000, 1024
    x = flat.synth(0, n, d, 43)
    flat.normalize_L2(x)
    q = flat.synth(0, 256, d, 44)
    flat.normalize_L2(q)
    _, _, eff = _planes(x)
    plain = IndexFlatIP(d, dtype="bfloat16")
    plain.add(x)
    res = IndexFlatIP(d, dtype="bfloat16+residual")
    res.add(x)
    Df, If = flat.search(x, q, 100, flat.METRIC_IP)
    _, Ip = plain.search(q, 100)
    _, Ir = res.search(q, 100)
    r_plain, r_res = recall_at_k(Ip, If), recall_at_k(Ir, If)
    assert r_res >= 0.999 and r_res >= r_plain, (r_plain, r_res)
    # remove_ids compacts both planes; merge_from appends both
    drop = np.arange(1000, 5000, 7)
    assert res.remove_ids(drop) == len(drop)
    keep = np.setdiff1d(np.arange(n), drop)
    assert np.array_equal(res.reconstruct_n(0, res.ntotal), eff[keep])
    other = IndexFlatIP(d, dtype="bfloat16+residual")
    other.add(x[:3000])
    res.merge_from(other)
    assert other.ntotal == 0 and res.ntotal == len(keep) + 3000
    assert np.array_equal(res.reconstruct_n(len(keep), 3000), eff[:3000])
    D, I = res.search(q[:16], 10)
    cur = np.concatenate([eff[keep], eff[:3000]])
    Dr, Ir2 = flat.search(cur, q[:16], 10, flat.METRIC_IP)
    assert_parity(D, I, Dr, Ir2, cur, q[:16], metric_ip=True)
    # a plain bf16 index and a residual one do not merge, and the option only flips on an empty index
    with pytest.raises(ValueError):
        res.merge_from(plain)
    with pytest.raises(ValueError):
        res.set_option("bf16_residual", 0)
