"""GPU: MMR re-ranking kernel against the numpy restatement of langchain's maximal_marginal_relevance."""
import numpy as np
import pytest

from oracle import flat

pytestmark = pytest.mark.gpu


def _check_query(x, q, ids, scores, k, fetch_k, lam, metric):
    """Pick-by-pick check that tolerates fp32-vs-numpy near ties in the greedy objective."""
    D, I = flat.search(x, q[None, :], fetch_k, metric)
    cand = [int(i) for i in I[0] if i != -1]
    E = x[cand].astype(np.float64)
    qq = q.astype(np.float64)
    simq = flat.cosine_similarity(qq[None, :], E)[0]
    picked = []
    for r, pid in enumerate([int(i) for i in ids if i != -1]):
        assert pid in cand and pid not in [cand[p] for p in picked]
        pos = cand.index(pid)
        if r == 0:
            obj = simq
        else:
            red = flat.cosine_similarity(E, E[picked]).max(axis=1)
            obj = lam * simq - (1 - lam) * red
            obj[picked] = -np.inf
        assert obj[pos] >= obj.max() - 1e-5, f"pick {r}: {pid} objective {obj[pos]} < best {obj.max()}"
        assert abs(float(scores[r]) - float(D[0][pos])) <= 1e-5 * max(1.0, abs(float(D[0][pos])))
        picked.append(pos)
    assert len(picked) == min(k, len(cand))


@pytest.mark.parametrize("metric", [flat.METRIC_L2, flat.METRIC_IP])
def test_mmr_batch_matches_reference_semantics(metric):
    from nv_genai_b200 import IndexFlat

    rng = np.random.default_rng(1)
    cent = flat.synth(0, 40, 256, 31)
    x = np.ascontiguousarray(cent[rng.integers(0, 40, 5000)] + 0.3 * flat.synth(0, 5000, 256, 32), np.float32)
    q = np.ascontiguousarray(cent[:12] + 0.3 * flat.synth(0, 12, 256, 33), np.float32)
    idx = IndexFlat(256, metric)
    idx.add(x)
    for k, fetch_k, lam in ((4, 20, 0.5), (6, 30, 0.2), (3, 10, 1.0), (5, 5, 0.0)):
        ids, sc = idx.mmr(q, k=k, fetch_k=fetch_k, lambda_mult=lam)
        assert ids.shape == (12, k)
        exact = 0
        for r in range(12):
            _check_query(x, q[r], ids[r], sc[r], k, fetch_k, lam, metric)
            want, _ = flat.mmr_search(x, q[r], k, fetch_k, lam, metric)
            exact += int([int(i) for i in ids[r] if i != -1] == want)
        assert exact >= 11  # identical to the numpy restatement except for a possible near tie


def test_mmr_config5_shape_and_padding():
    """BASELINE config 5b: batch 256, fetch_k=20, k=4, lambda 0.5 (at a small corpus)."""
    import torch

    from nv_genai_b200 import IndexFlat

    x = flat.synth(0, 50_000, 1024, 41)
    flat.normalize_L2(x)
    q = flat.synth(0, 256, 1024, 42)
    flat.normalize_L2(q)
    idx = IndexFlat(1024, flat.METRIC_IP)
    idx.add(x)
    ids, sc = idx.mmr(torch.from_numpy(q).cuda(), k=4, fetch_k=20, lambda_mult=0.5)
    assert ids.is_cuda and ids.shape == (256, 4)
    ids, sc = ids.cpu().numpy(), sc.cpu().numpy()
    for r in range(0, 256, 37):
        _check_query(x, q[r], ids[r], sc[r], 4, 20, 0.5, flat.METRIC_IP)
    small = IndexFlat(1024, flat.METRIC_IP)
    small.add(x[:2])
    ids, sc = small.mmr(q[:3], k=4, fetch_k=20)
    assert np.all(ids[:, 2:] == -1) and np.all(ids[:, :2] >= 0)
    empty = IndexFlat(1024, flat.METRIC_IP)
    ids, sc = empty.mmr(q[:1], k=4, fetch_k=20)
    assert np.all(ids == -1)
