"""GPU: the error band of the filter-and-refine paths on ADVERSARIAL data.

The band that decides which candidates are re-scored exactly is rigorous for operand rounding (Cauchy-Schwarz), but
its accumulation term - steps x 2^-21 |q||x| for the tensor core's truncating accumulate - rests on a measurement taken
on benign data.  Here every dot product runs through partial sums about as large as |q||x| that cancel to a small
score: rows are [a w + g, -(a w + g)] + e and queries [w + h, w + h], so after half of the K steps the accumulator
holds a |w|^2 and the final score is only q.e.  If the hardware lost more per step than the band assumes, true
neighbours would be dropped by the filter and never re-scored.  The check is against float64 scores of ALL rows: every
row that beats the k-th best by more than fp32 re-scoring noise must be returned, and nothing worse than the k-th best
by more than that noise may be."""
import numpy as np
import pytest

from oracle import flat

pytestmark = pytest.mark.gpu
N = 100_000


def _adversarial(d, nq, seed):
    rng = np.random.default_rng(seed)
    h = d // 2
    w = rng.standard_normal(h).astype(np.float32) * 4.0
    a = rng.uniform(0.5, 2.0, size=(N, 1)).astype(np.float32) * rng.choice([-1.0, 1.0], size=(N, 1)).astype(np.float32)
    g = rng.standard_normal((N, h)).astype(np.float32) * 0.25
    half = a * w[None, :] + g
    x = np.concatenate([half, -half], axis=1)
    x += rng.standard_normal((N, d)).astype(np.float32) * 0.05
    hq = rng.standard_normal((nq, h)).astype(np.float32) * 0.25
    qh = w[None, :] + hq
    q = np.concatenate([qh, qh], axis=1) + rng.standard_normal((nq, d)).astype(np.float32) * 0.05
    return np.ascontiguousarray(x, dtype=np.float32), np.ascontiguousarray(q, dtype=np.float32)


def _scores64(x_stored, q, metric_ip):
    """float64 keys (larger = better) of every (query, row) pair, plus |q| per query and the largest |x|."""
    q64 = q.astype(np.float64)
    s = np.empty((q.shape[0], x_stored.shape[0]), dtype=np.float64)
    xn2 = np.empty(x_stored.shape[0], dtype=np.float64)
    for r0 in range(0, x_stored.shape[0], 20_000):  # blockwise: the float64 copy of all rows would be 3 GB at d = 4096
        xb = x_stored[r0:r0 + 20_000].astype(np.float64)
        s[:, r0:r0 + 20_000] = q64 @ xb.T
        xn2[r0:r0 + 20_000] = (xb ** 2).sum(1)
    if not metric_ip:
        s = -((q64 ** 2).sum(1)[:, None] + xn2[None, :] - 2.0 * s)
    return s, np.linalg.norm(q64, axis=1), np.sqrt(xn2.max())


def _check(I, s, qn, xn, d, k, metric_ip):
    for r in range(I.shape[0]):
        # fp32 re-scoring noise of a d-term dot product whose terms are as large as |q||x| / sqrt(d)
        tol = 8.0 * np.sqrt(d) * 2.0 ** -24 * qn[r] * xn * (1.0 if metric_ip else 4.0)
        kth = np.partition(s[r], -k)[-k]
        got = I[r]
        assert len(set(got.tolist())) == k and got.min() >= 0
        assert s[r][got].min() >= kth - tol, (r, s[r][got].min(), kth, tol)
        must = np.nonzero(s[r] > kth + tol)[0]
        missing = set(must.tolist()) - set(got.tolist())
        assert not missing, (r, sorted(missing)[:5], kth, tol)


@pytest.mark.parametrize("d", [1024, 4096])
@pytest.mark.parametrize("mode", ["fp32_shadow", "fp32_tf32", "bf16"])
@pytest.mark.parametrize("metric_ip", [True, False])
def test_band_covers_cancelling_partial_sums(d, mode, metric_ip):
    from nv_genai_b200 import IndexFlatIP, IndexFlatL2

    x, q = _adversarial(d, 64, seed=d + (1 if metric_ip else 2))
    # the construction must do what it claims: partial sums far above the final scores
    half = (q[:4, : d // 2].astype(np.float64) @ x[:200, : d // 2].astype(np.float64).T)
    full = q[:4].astype(np.float64) @ x[:200].astype(np.float64).T
    assert np.abs(half).mean() > 50 * np.abs(full).mean()
    dtype = "bfloat16" if mode == "bf16" else "float32"
    idx = (IndexFlatIP if metric_ip else IndexFlatL2)(d, dtype=dtype)
    if mode == "fp32_tf32":
        idx.set_option("bf16_shadow", 0)
    idx.add(x)
    stored = flat.round_bf16(x) if mode == "bf16" else x
    s, qn, xn = _scores64(stored, q, metric_ip)
    for nq, k in ((64, 10), (64, 100), (16, 10), (5, 10)):  # GEMM filter-and-refine, tensor-core scan (approximate)
        D, I = idx.search(q[:nq], k)
        _check(I, s[:nq], qn, xn, d, k, metric_ip)
