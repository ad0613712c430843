"""CPU: the oracle against its pins — reference-generated golden vectors, float64 brute force,
hand-computable known answers, and properties."""
import json
import os

import numpy as np
import pytest
from hypothesis import given, settings
from hypothesis import strategies as st

from oracle import flat

HERE = os.path.dirname(os.path.abspath(__file__))
FLT_MAX = np.finfo(np.float32).max


def test_synth_c_matches_numpy_restatement():
    for row0, n, d, seed in [(0, 5, 8, 1), (123456789012, 7, 1024, flat.CORPUS_SEED), (3, 4, 10, flat.QUERY_SEED)]:
        a = flat.synth(row0, n, d, seed)
        b = flat.synth_numpy(row0, n, d, seed)
        assert a.dtype == np.float32 and np.array_equal(a, b)
    x = flat.synth(0, 4000, 256, 7)
    assert abs(float(x.mean())) < 0.01 and abs(float(x.std()) - 1.0) < 0.01
    # rows are addressable independently of the block they are generated in
    assert np.array_equal(flat.synth(1000, 10, 256, 7), x[1000:1010])


def test_golden_ip_topk_sets_from_reference_notebook():
    g = json.load(open(os.path.join(HERE, "golden", "ip_topk_reference.json")))
    assert "retriever_customization.ipynb" in g["source"]
    for c in g["cases"]:
        P = flat.synth(0, c["n"], c["d"], g["corpus_seed"])
        Q = flat.synth(0, c["nq"], c["d"], g["query_seed"])
        if c["normalize"]:
            flat.normalize_L2(P)
            flat.normalize_L2(Q)
        D, I = flat.search(P, Q, c["k"], flat.METRIC_IP)
        for r in range(c["nq"]):
            assert sorted(I[r].tolist()) == c["topk_sets"][r], c["name"]
        assert np.all(np.diff(D, axis=1) <= 0)  # IP descending
        if c["normalize"]:
            # unit vectors: |x - q|^2 = 2 - 2 q.x, so the same sets pin IndexFlatL2 (the reference's default index)
            D2, I2 = flat.search(P, Q, c["k"], flat.METRIC_L2)
            for r in range(c["nq"]):
                assert sorted(I2[r].tolist()) == c["topk_sets"][r], c["name"] + " (L2)"
            assert np.all(np.diff(D2, axis=1) >= 0)  # L2 ascending


@pytest.mark.parametrize("metric", [flat.METRIC_IP, flat.METRIC_L2])
@pytest.mark.parametrize("n,d,nq,k", [(1000, 64, 3, 5), (257, 100, 2, 16), (5000, 1024, 2, 10)])
def test_seq_matches_float64_bruteforce(metric, n, d, nq, k):
    x = flat.synth(0, n, d, 11)
    q = flat.synth(0, nq, d, 12)
    D, I, used = flat.search_seq(x, q, k, metric)
    D64, I64 = flat.search_f64(x, q, k, metric)
    assert used == min(nq, flat.lib().b2o_max_threads())  # OpenMP over queries only
    assert np.array_equal(I, I64)
    np.testing.assert_allclose(D, D64, rtol=2e-6, atol=1e-6)
    mono = np.diff(D, axis=1)
    assert np.all(mono <= 0) if metric == flat.METRIC_IP else np.all(mono >= 0)


def test_rowparallel_variant_equals_faithful_one():
    x = flat.synth(0, 3001, 96, 5)
    q = flat.synth(0, 2, 96, 6)
    for metric in (flat.METRIC_IP, flat.METRIC_L2):
        a = flat.search_seq(x, q, 7, metric)
        b = flat.search_seq(x, q, 7, metric, rowpar=True)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_blas_path_agrees_with_seq_path():
    x = flat.synth(0, 4000, 128, 21)
    q = flat.synth(0, 24, 128, 22)
    for metric in (flat.METRIC_IP, flat.METRIC_L2):
        Ds, Is, _ = flat.search_seq(x, q, 10, metric)
        Db, Ib = flat.search_blas(x, q, 10, metric, block=1024)
        assert (Is == Ib).mean() > 0.99
        np.testing.assert_allclose(Ds, Db, rtol=1e-4, atol=1e-3)


def test_known_answers_basis_vectors():
    d = 16
    x = np.eye(d, dtype=np.float32)
    q = (np.arange(d, dtype=np.float32) + 1)[None, :]  # strictly increasing weights
    D, I = flat.search(x, q, 4, flat.METRIC_IP)
    assert I[0].tolist() == [15, 14, 13, 12] and D[0].tolist() == [16.0, 15.0, 14.0, 13.0]
    D, I = flat.search(x, x[3:4], 2, flat.METRIC_L2)
    assert I[0, 0] == 3 and D[0, 0] == 0.0 and D[0, 1] == 2.0  # squared, not rooted


def test_ties_lowest_id_first_and_duplicates():
    x = flat.synth(0, 50, 32, 3)
    x[40] = x[7]
    x[20] = x[7]
    D, I = flat.search(x, x[7:8], 3, flat.METRIC_L2)
    assert I[0].tolist()[:3] == [7, 20, 40] and np.all(D[0] == 0)
    ones = np.ones((9, 8), np.float32)
    D, I = flat.search(ones, ones[:1], 4, flat.METRIC_IP)
    assert I[0].tolist() == [0, 1, 2, 3]


def test_padding_when_fewer_rows_than_k_and_empty_index():
    x = flat.synth(0, 3, 8, 1)
    q = flat.synth(0, 2, 8, 2)
    D, I = flat.search(x, q, 5, flat.METRIC_L2)
    assert np.all(I[:, 3:] == -1) and np.all(D[:, 3:] == FLT_MAX) and np.all(I[:, :3] >= 0)
    D, I = flat.search(x, q, 5, flat.METRIC_IP)
    assert np.all(I[:, 3:] == -1) and np.all(D[:, 3:] == -FLT_MAX)
    D, I = flat.search(np.empty((0, 8), np.float32), q, 2, flat.METRIC_L2)
    assert np.all(I == -1) and np.all(D == FLT_MAX)


def test_nan_rows_never_enter():
    x = flat.synth(0, 10, 8, 1)
    x[4, 0] = np.nan
    D, I = flat.search(x, flat.synth(0, 1, 8, 2), 10, flat.METRIC_IP)
    assert 4 not in I[0].tolist() and I[0, -1] == -1


def test_normalize_l2_semantics():
    x = flat.synth(0, 20, 48, 9)
    x[5] = 0
    y = x.copy()
    flat.normalize_L2(y)
    assert np.all(y[5] == 0)  # zero rows untouched
    nz = np.delete(np.arange(20), 5)
    np.testing.assert_allclose(np.linalg.norm(y[nz], axis=1), 1.0, rtol=1e-6)
    np.testing.assert_allclose(y[nz], x[nz] / np.linalg.norm(x[nz].astype(np.float64), axis=1)[:, None], rtol=1e-6)


def test_round_bf16_is_round_to_nearest_even():
    import torch

    x = flat.synth(0, 64, 64, 4)
    assert np.array_equal(flat.round_bf16(x), torch.from_numpy(x).bfloat16().float().numpy())


def test_streaming_search_equals_one_shot():
    x = flat.synth(0, 5000, 64, 31)
    q = flat.synth(0, 3, 64, 32)
    s = flat.StreamingSearch(q, 9, flat.METRIC_IP)
    for r0 in range(0, 5000, 1300):
        s.feed(x[r0:r0 + 1300], r0)
    D, I = s.result()
    D1, I1, _ = flat.search_seq(x, q, 9, flat.METRIC_IP)
    assert np.array_equal(I, I1) and np.array_equal(D, D1)


def test_mmr_hand_example():
    # e0 is closest to q; e1 is a near-copy of e0 (redundant); e2 is orthogonal to e0 but still relevant
    q = np.array([1.0, 0.2, 0.0], np.float32)
    E = [np.array(v, np.float32) for v in ([1, 0, 0], [0.99, 0.05, 0], [0.3, 1, 0], [-1, 0, 0])]
    assert flat.maximal_marginal_relevance(q, E, lambda_mult=0.5, k=2) == [1, 2] or \
        flat.maximal_marginal_relevance(q, E, lambda_mult=0.5, k=2)[1] == 2
    assert flat.maximal_marginal_relevance(q, E, lambda_mult=1.0, k=3)[:2] in ([1, 0], [0, 1])
    assert flat.maximal_marginal_relevance(q, [], k=3) == []
    assert len(flat.maximal_marginal_relevance(q, E, k=10)) == 4
    z = flat.cosine_similarity(np.zeros((1, 3)), np.ones((2, 3)))
    assert np.all(z == 0)  # NaN -> 0


def test_relevance_score_identities():
    assert flat.euclidean_relevance(0.0) == 1.0
    assert abs(flat.euclidean_relevance(np.sqrt(2.0))) < 1e-12
    assert flat.max_inner_product_relevance(0.25) == 0.75
    assert flat.max_inner_product_relevance(-0.25) == 0.25


@settings(max_examples=25, deadline=None)
@given(st.integers(1, 200), st.integers(1, 40), st.integers(1, 12), st.integers(0, 2 ** 31), st.booleans())
def test_property_permutation_and_k_monotone(n, d, k, seed, ip):
    metric = flat.METRIC_IP if ip else flat.METRIC_L2
    x = flat.synth(0, n, d, seed)
    q = flat.synth(0, 1, d, seed + 1)
    D, I = flat.search(x, q, k, metric)
    # result set is invariant under a row permutation (ids mapped back); scores identical
    perm = np.random.default_rng(seed).permutation(n)
    Dp, Ip = flat.search(x[perm], q, k, metric)
    assert np.array_equal(D, Dp)
    back = np.where(Ip >= 0, perm[np.clip(Ip, 0, n - 1)], -1)
    if len(np.unique(D[0][I[0] >= 0])) == (I[0] >= 0).sum():  # no exact ties: ids map one to one
        assert np.array_equal(back, I)
    # top-(k+1) extends top-k
    D2, I2 = flat.search(x, q, k + 1, metric)
    assert np.array_equal(I2[:, :k], I) and np.array_equal(D2[:, :k], D)


@pytest.mark.parametrize("metric", [flat.METRIC_IP, flat.METRIC_L2])
def test_streaming_search_over_blocks_equals_one_scan(metric):
    """What the full-size GPU tests and bench.py's check rely on: feeding the corpus block by block (ragged blocks,
    rows regenerated per block) gives the same (D, I) as one sequential scan, ids global, ties lowest id first."""
    n, d, nq, k = 5_003, 96, 7, 12
    x = flat.synth(0, n, d, 5)
    x[4000] = x[17]  # a tie across two blocks
    q = flat.synth(0, nq, d, 6)
    q[2] = x[17]
    st_ = flat.StreamingSearch(q, k, metric)
    for r0 in range(0, n, 1_111):
        st_.feed(x[r0:r0 + 1_111], r0)
    D, I = st_.result()
    Dr, Ir, _ = flat.search_seq(x, q, k, metric)
    assert np.array_equal(I, Ir) and np.array_equal(D, Dr)
    assert I[2, 0] == 17 and I[2, 1] == 4000
    # fewer rows than k: padded like index.search
    st2 = flat.StreamingSearch(q, k, metric)
    st2.feed(x[:5], 0)
    D2, I2 = st2.result()
    D2r, I2r, _ = flat.search_seq(x[:5], q, k, metric)
    assert np.array_equal(I2, I2r) and np.array_equal(D2, D2r)
    # the in-place bf16 rounding equals the copying one
    y = x.copy()
    flat.round_bf16_inplace(y)
    assert np.array_equal(y, flat.round_bf16(x))
