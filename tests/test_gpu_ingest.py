"""GPU: ingest (add + fused normalise + bf16 cast), reconstruct, remove_ids, merge_from, persistence."""
import numpy as np
import pytest

from oracle import flat

pytestmark = pytest.mark.gpu


def _idx(d, metric=flat.METRIC_L2, dtype="float32"):
    from nv_genai_b200 import IndexFlat

    return IndexFlat(d, metric, dtype)


@pytest.mark.parametrize("d", [3, 64, 100, 1024, 1100, 4096])
def test_add_reconstruct_roundtrip_is_exact(d):
    x = flat.synth(0, 500, d, 5)
    idx = _idx(d)
    idx.add(x[:200])
    idx.add(x[200:])
    assert np.array_equal(idx.reconstruct_n(0, 500), x)
    assert np.array_equal(idx.reconstruct(499), x[499])
    got = idx.reconstruct_batch([5, -1, 0])
    assert np.array_equal(got[0], x[5]) and np.all(got[1] == 0) and np.array_equal(got[2], x[0])
    with pytest.raises(RuntimeError):
        idx.reconstruct(500)


@pytest.mark.parametrize("d", [7, 128, 1024, 2048])
def test_fused_normalise_matches_faiss_normalize_l2(d):
    x = flat.synth(0, 1000, d, 8)
    x[13] = 0  # zero rows stay zero
    want = x.copy()
    flat.normalize_L2(want)
    idx = _idx(d)
    idx.add(x, normalize=True)
    got = idx.reconstruct_n(0, 1000)
    assert np.all(got[13] == 0)
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-9)
    # module-level faiss.normalize_L2 look-alike, in place on numpy and on CUDA tensors
    import torch

    from nv_genai_b200 import normalize_L2

    y = x.copy()
    normalize_L2(y)
    np.testing.assert_allclose(y, want, rtol=1e-6, atol=1e-9)
    t = torch.from_numpy(x).cuda()
    normalize_L2(t)
    np.testing.assert_allclose(t.cpu().numpy(), want, rtol=1e-6, atol=1e-9)


def test_bf16_store_holds_round_to_nearest_even_values():
    x = flat.synth(0, 300, 1024, 4)
    idx = _idx(1024, dtype="bfloat16")
    idx.add(x)
    assert np.array_equal(idx.reconstruct_n(0, 300), flat.round_bf16(x))
    x2 = flat.synth(0, 50, 100, 4)
    idx = _idx(100, dtype="bfloat16")
    idx.add(x2)
    assert np.array_equal(idx.reconstruct_n(0, 50), flat.round_bf16(x2))


def test_large_host_add_is_chunked_correctly():
    x = flat.synth(0, 70_000, 1024, 12)  # 287 MB: several 64 MB staging chunks
    idx = _idx(1024)
    idx.add(x)
    ids = np.array([0, 16_383, 16_384, 16_385, 32_768, 65_535, 69_999])
    assert np.array_equal(idx.reconstruct_batch(ids), x[ids])


def test_synth_on_device_is_bit_identical_to_the_oracle_generator():
    import torch

    from nv_genai_b200 import synth_fill

    for d in (1024, 100, 6):
        t = torch.empty((257, d), dtype=torch.float32, device="cuda")
        synth_fill(t, 123_456_789, flat.CORPUS_SEED)
        assert np.array_equal(t.cpu().numpy(), flat.synth(123_456_789, 257, d, flat.CORPUS_SEED))


def test_remove_ids_compacts_in_order():
    x = flat.synth(0, 40_000, 256, 3)
    idx = _idx(256)
    idx.add(x)
    rng = np.random.default_rng(0)
    victims = rng.choice(40_000, 9000, replace=False)
    n = idx.remove_ids(np.concatenate([victims, victims[:10], [-5, 99_999]]))  # duplicates / unknown ids ignored
    assert n == 9000 and idx.ntotal == 31_000
    keep = np.ones(40_000, bool)
    keep[victims] = False
    assert np.array_equal(idx.reconstruct_n(0, 31_000), x[keep])
    D, I = idx.search(x[keep][1234:1235], 1)
    assert I[0, 0] == 1234 and D[0, 0] == 0.0
    assert idx.remove_ids([]) == 0
    idx.add(x[:3])
    assert idx.ntotal == 31_003 and np.array_equal(idx.reconstruct(31_000), x[0])


def test_merge_from_appends_and_resets_other():
    a, b = _idx(64), _idx(64)
    xa, xb = flat.synth(0, 100, 64, 1), flat.synth(0, 50, 64, 2)
    a.add(xa)
    b.add(xb)
    a.merge_from(b)
    assert a.ntotal == 150 and b.ntotal == 0
    assert np.array_equal(a.reconstruct_n(0, 150), np.concatenate([xa, xb]))
    with pytest.raises(ValueError):
        a.merge_from(_idx(32))
    a.reset()
    assert a.ntotal == 0


def test_save_load_and_pickle_roundtrip(tmp_path):
    import pickle

    from nv_genai_b200 import read_index, write_index

    x = flat.synth(0, 1234, 96, 6)
    for metric in (flat.METRIC_L2, flat.METRIC_IP):
        idx = _idx(96, metric)
        idx.add(x)
        p = str(tmp_path / f"i{metric}.faiss")
        write_index(idx, p)
        back = read_index(p)
        assert back.metric_type == metric and back.ntotal == 1234 and np.array_equal(back.reconstruct_n(), x)
        again = pickle.loads(pickle.dumps(idx))
        assert again.metric_type == metric and np.array_equal(again.reconstruct_n(), x)


def test_bulk_host_add_through_the_pinned_ring_and_from_pinned_memory():
    """More than 8 MB of host rows take the pipelined path of b2_index_add: pageable rows through the library's pinned
    ring (several slots, chunks that do not divide the input), pinned rows straight from where they are.  Both must
    store exactly what the one-copy path and the device path store."""
    import torch

    from nv_genai_b200 import IndexFlatIP

    n, d = 70_001, 768  # 215 MB: seven 32 MB chunks and a ragged tail; rows of 3 KB do not divide a chunk evenly
    x = flat.synth(0, n, d, 77)
    ref = IndexFlatIP(d)
    ref.add(torch.from_numpy(x).cuda(), normalize=True)  # device rows
    want = ref.reconstruct_n(0, n)
    a = IndexFlatIP(d)
    a.add(x, normalize=True)  # pageable numpy
    assert np.array_equal(a.reconstruct_n(0, n), want)
    pinned = torch.from_numpy(x).pin_memory()
    b = IndexFlatIP(d, dtype="bfloat16")
    b.add(pinned.numpy(), normalize=True)
    ref16 = IndexFlatIP(d, dtype="bfloat16")
    ref16.add(torch.from_numpy(x).cuda(), normalize=True)
    assert np.array_equal(b.reconstruct_n(0, n), ref16.reconstruct_n(0, n))
    # a second bulk add reuses the ring
    a.add(x[:40_000], normalize=False)
    assert a.ntotal == n + 40_000 and np.array_equal(a.reconstruct_n(n, 40_000), x[:40_000])


def test_remove_ids_keep_list_built_on_the_device_matches_numpy_compaction():
    """remove_ids marks, scans and compacts on the device (only the ids are uploaded): random ids with duplicates,
    negative and out-of-range entries, runs that cross the 32-row words and the 32768-row blocks of the scan."""
    from nv_genai_b200 import IndexFlatL2

    n, d = 200_003, 64
    x = flat.synth(0, n, d, 55)
    rng = np.random.default_rng(5)
    for ids in (
        np.array([n - 1]),                                  # only the last row
        np.array([0]),                                      # only the first: everything moves
        np.concatenate([rng.integers(0, n, 7000), [-5, n, n + 10, 31, 32, 33, 32767, 32768, 65535, 65536]]),
        np.arange(40_000, 120_000),                         # a run over several scan blocks
    ):
        idx = IndexFlatL2(d)
        idx.add(x)
        keep = np.ones(n, bool)
        valid = ids[(ids >= 0) & (ids < n)]
        keep[valid] = False
        assert idx.remove_ids(ids) == int((~keep).sum())
        assert idx.ntotal == int(keep.sum())
        assert np.array_equal(idx.reconstruct_n(0, idx.ntotal), x[keep])
        D, I = idx.search(x[keep][:3], 1)
        assert I[:, 0].tolist() == [0, 1, 2] and np.all(D[:, 0] == 0.0)
    assert idx.remove_ids(np.array([10**9, -1])) == 0 and idx.ntotal == int(keep.sum())
