"""GPU: BASELINE.json's full size (10M x 1024 fp32, normalised, inner product) checked through
size-independent properties, since the oracle cannot scan 41 GB in seconds:
 - planted queries (copies of corpus rows) come back at rank 1 with score ~1;
 - a slice of the corpus scanned by the oracle bounds what the full search may return;
 - the union of per-slice searches merged on the device equals the whole search bit for bit;
 - top-(k) is a prefix of top-(k+5)."""
import numpy as np
import pytest

from oracle import flat

pytestmark = pytest.mark.gpu
N, D, CHUNK = 10_000_000, 1024, 1_000_000


@pytest.fixture(scope="module")
def big():
    import torch

    from nv_genai_b200 import IndexFlatIP, synth_fill

    free, _ = torch.cuda.mem_get_info()
    if free < 60 * (1 << 30):
        pytest.skip("needs ~45 GB of free HBM")
    idx = IndexFlatIP(D)
    buf = torch.empty((CHUNK, D), dtype=torch.float32, device="cuda")
    for r0 in range(0, N, CHUNK):
        synth_fill(buf, r0, flat.CORPUS_SEED)
        idx.add(buf, normalize=True)  # fused normalise on ingest
    del buf
    torch.cuda.synchronize()
    assert idx.ntotal == N
    return idx


def _queries(nq):
    q = flat.synth(0, nq, D, flat.QUERY_SEED)
    flat.normalize_L2(q)
    return q


def test_planted_rows_rank_first_with_unit_score(big):
    rows = np.array([0, 1, 4_999_999, 5_000_000, 9_999_999, 7_654_321], dtype=np.int64)
    q = big.reconstruct_batch(rows)
    for nq in (1, 6):
        D_, I_ = big.search(q[:nq], 10)
        assert np.array_equal(I_[:, 0], rows[:nq])
        np.testing.assert_allclose(D_[:, 0], 1.0, rtol=1e-5)
        assert np.all(np.diff(D_, axis=1) <= 0) and np.all(D_[:, 1] < 0.5)


@pytest.mark.parametrize("nq", [1, 16])
def test_full_search_is_consistent_with_an_oracle_slice(big, nq):
    q = _queries(nq)
    D_, I_ = big.search(q, 10)
    assert np.all(I_ >= 0) and np.all(I_ < N)
    lo = 3_000_000
    xs = flat.synth(lo, 400_000, D, flat.CORPUS_SEED)
    flat.normalize_L2(xs)
    Ds, Is, _ = flat.search_seq(xs, q, 10, flat.METRIC_IP)
    for r in range(nq):
        # every returned id inside the slice must be in the slice's own top-10, with the oracle's score
        for j, gid in enumerate(I_[r]):
            if lo <= gid < lo + 400_000:
                pos = np.nonzero(Is[r] == gid - lo)[0]
                assert len(pos) == 1 and abs(Ds[r][pos[0]] - D_[r][j]) <= 1e-5 * abs(Ds[r][pos[0]])
        # and nothing in the slice may beat the k-th full result without being returned
        kth = D_[r, -1]
        better = Is[r][Ds[r] > kth * (1 + 1e-5) if kth > 0 else Ds[r] > kth]
        assert set((better + lo).tolist()) <= set(I_[r].tolist())
    # scores recomputed in float64 from the stored rows
    rows = big.reconstruct_batch(I_.reshape(-1)).reshape(nq, 10, D).astype(np.float64)
    ts = np.einsum("qkd,qd->qk", rows, q.astype(np.float64))
    np.testing.assert_allclose(D_, ts, rtol=1e-5, atol=1e-7)


def test_prefix_property_and_decomposition(big):
    import torch

    from nv_genai_b200 import IndexFlatIP, merge_shard_lists, synth_fill

    q = torch.from_numpy(_queries(16)).cuda()
    D10, I10 = big.search(q, 10)
    D15, I15 = big.search(q, 15)
    assert torch.equal(I15[:, :10], I10) and torch.equal(D15[:, :10], D10)
    # decomposition: four 1M-row shards searched separately then merged == one 4M index
    whole = IndexFlatIP(D)
    parts_D, parts_I = [], []
    buf = torch.empty((CHUNK, D), dtype=torch.float32, device="cuda")
    for g in range(4):
        synth_fill(buf, g * CHUNK, flat.CORPUS_SEED)
        whole.add(buf, normalize=True)
        sh = IndexFlatIP(D)
        sh.add(buf, normalize=True)
        Dg, Ig = sh.search(q, 10)
        parts_D.append(Dg)
        parts_I.append(Ig + g * CHUNK)
        del sh
    Dm, Im = merge_shard_lists(flat.METRIC_IP, torch.stack(parts_D), torch.stack(parts_I))
    Dw, Iw = whole.search(q, 10)
    assert torch.equal(Im, Iw) and torch.equal(Dm, Dw)
