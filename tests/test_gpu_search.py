"""GPU: parity of the CUDA search (through the C ABI) with the oracle on the same seeded inputs."""
import json
import os

import numpy as np
import pytest

from oracle import flat
from parity import assert_parity, recall_at_k

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
FLT_MAX = np.finfo(np.float32).max


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


def test_cfg1_100k_x_1024_l2_k4_single_query():
    """BASELINE config 1: the 5_mins_rag_no_gpu path, IndexFlatL2, un-normalised, nq=1, k=4."""
    x, q = _data(100_000, 1024, 1, False)
    idx = _mk(flat.METRIC_L2, 1024)
    idx.add(x)
    assert idx.ntotal == 100_000 and idx.d == 1024
    D, I = idx.search(q, 4)
    Dr, Ir, _ = flat.search_seq(x, q, 4, flat.METRIC_L2)
    assert D.dtype == np.float32 and I.dtype == np.int64 and D.shape == (1, 4)
    assert_parity(D, I, Dr, Ir, x, q, metric_ip=False)
    assert np.array_equal(I, Ir)


@pytest.mark.parametrize("nq", [1, 16])
def test_cfg2_shape_ip_normalised_k10(nq):
    """BASELINE config 2 at an oracle-friendly size (the full 10M run is in test_gpu_fullsize.py)."""
    x, q = _data(300_000, 1024, nq, True)
    idx = _mk(flat.METRIC_IP, 1024)
    idx.add(x)
    D, I = idx.search(q, 10)
    Dr, Ir, _ = flat.search_seq(x, q, 10, flat.METRIC_IP)
    assert_parity(D, I, Dr, Ir, x, q, metric_ip=True)
    assert np.all(np.diff(D, axis=1) <= 0)


@pytest.mark.parametrize("metric", [flat.METRIC_IP, flat.METRIC_L2])
@pytest.mark.parametrize("d", [4, 100, 384, 768, 1024, 1536, 4100])
def test_dimensions_including_ragged(metric, d):
    x, q = _data(3000, d, 3, False, seed=d)
    idx = _mk(metric, d)
    idx.add(x)
    D, I = idx.search(q, 7)
    Dr, Ir, _ = flat.search_seq(x, q, 7, metric)
    assert_parity(D, I, Dr, Ir, x, q, metric_ip=(metric == flat.METRIC_IP))


@pytest.mark.parametrize("nq", [1, 2, 3, 5, 8, 15, 16, 17, 33, 70])
def test_query_batch_sizes(nq):
    x, q = _data(20_000, 256, nq, True, seed=nq)
    for metric in (flat.METRIC_IP, flat.METRIC_L2):
        idx = _mk(metric, 256)
        idx.add(x)
        D, I = idx.search(q, 10)
        Dr, Ir = flat.search_f64(x, q, 10, metric)  # direct forms for every batch size
        assert_parity(D, I, Dr.astype(np.float32), Ir, x, q, metric_ip=(metric == flat.METRIC_IP))


@pytest.mark.parametrize("k", [1, 4, 10, 31, 32, 33, 100, 128, 257, 1000, 2048])
def test_k_values(k):
    x, q = _data(30_000, 128, 4, True, seed=k)
    idx = _mk(flat.METRIC_IP, 128)
    idx.add(x)
    D, I = idx.search(q, k)
    Dr, Ir, _ = flat.search_seq(x, q, k, flat.METRIC_IP)
    assert_parity(D, I, Dr, Ir, x, q, metric_ip=True)


def test_k_too_large_is_a_clear_error():
    idx = _mk(flat.METRIC_L2, 16)
    idx.add(flat.synth(0, 10, 16, 1))
    with pytest.raises(NotImplementedError, match="B2_MAX_K"):
        idx.search(flat.synth(0, 1, 16, 2), 4096)
    with pytest.raises(AssertionError):
        idx.search(flat.synth(0, 1, 16, 2), 0)
    with pytest.raises(AssertionError):
        idx.search(flat.synth(0, 1, 8, 2), 1)  # dimension mismatch -> AssertionError like the SWIG wrapper
    with pytest.raises(AssertionError):
        idx.add(flat.synth(0, 2, 8, 2))


def test_empty_index_and_fewer_rows_than_k():
    q = flat.synth(0, 2, 32, 2)
    for metric, pad in ((flat.METRIC_L2, FLT_MAX), (flat.METRIC_IP, -FLT_MAX)):
        idx = _mk(metric, 32)
        D, I = idx.search(q, 3)
        assert np.all(I == -1) and np.all(D == pad)
        x = flat.synth(0, 3, 32, 1)
        idx.add(x)
        D, I = idx.search(q, 5)
        Dr, Ir = flat.search(x, q, 5, metric)
        assert np.array_equal(I, Ir) and np.all(D[:, 3:] == pad)
        assert_parity(D, I, Dr, Ir, x, q, metric_ip=(metric == flat.METRIC_IP))


def test_duplicates_and_ties_lowest_id_first():
    x, q = _data(5000, 64, 2, False, seed=5)
    x[4000] = x[17]
    x[2500] = x[17]
    x[100:108] = x[99]
    q[0] = x[17]
    q[1] = x[99]
    for metric in (flat.METRIC_L2, flat.METRIC_IP):
        idx = _mk(metric, 64)
        idx.add(x)
        D, I = idx.search(q, 12)
        Dr, Ir = flat.search(x, q, 12, metric)
        assert I[0, :3].tolist() == [17, 2500, 4000] and I[1, :9].tolist() == list(range(99, 108))
        assert_parity(D, I, Dr, Ir, x, q, metric_ip=(metric == flat.METRIC_IP))
    if True:  # exact zeros for self matches under the direct L2 form
        idx = _mk(flat.METRIC_L2, 64)
        idx.add(x)
        D, _ = idx.search(q, 3)
        assert np.all(D[0, :3] == 0.0)
    ones = np.ones((1000, 32), np.float32)
    idx = _mk(flat.METRIC_IP, 32)
    idx.add(ones)
    D, I = idx.search(ones[:1], 40)
    assert I[0].tolist() == list(range(40)) and np.all(D == 32.0)  # all-equal corpus: ids ascending


def test_clustered_corpus_dense_neighbourhoods():
    rng = np.random.default_rng(3)
    cent = flat.synth(0, 64, 256, 901)
    flat.normalize_L2(cent)
    x = cent[rng.integers(0, 64, 40_000)] + 0.05 * flat.synth(0, 40_000, 256, 902) / 16.0
    x = np.ascontiguousarray(x, dtype=np.float32)
    flat.normalize_L2(x)
    q = np.ascontiguousarray(cent[:8] + 0.05 * flat.synth(0, 8, 256, 903) / 16.0, dtype=np.float32)
    flat.normalize_L2(q)
    idx = _mk(flat.METRIC_IP, 256)
    idx.add(x)
    D, I = idx.search(q, 50)
    Dr, Ir, _ = flat.search_seq(x, q, 50, flat.METRIC_IP)
    assert_parity(D, I, Dr, Ir, x, q, metric_ip=True)


def test_nan_and_inf_rows():
    x = flat.synth(0, 200, 32, 1)
    x[7, 3] = np.nan
    x[9, 0] = np.inf
    q = np.abs(flat.synth(0, 1, 32, 2)) + 0.1
    idx = _mk(flat.METRIC_IP, 32)
    idx.add(x)
    D, I = idx.search(q, 200)
    Dr, Ir = flat.search(x, q, 200, flat.METRIC_IP)
    assert 7 not in I[0].tolist() and I[0, 0] == 9 and I[0, -1] == -1
    assert np.array_equal(I == -1, Ir == -1)


def test_golden_reference_sets_on_the_cuda_path():
    g = json.load(open(os.path.join(HERE, "golden", "ip_topk_reference.json")))
    for c in g["cases"]:
        P = flat.synth(0, c["n"], c["d"], g["corpus_seed"])
        Q = flat.synth(0, c["nq"], c["d"], g["query_seed"])
        if c["normalize"]:
            flat.normalize_L2(P)
            flat.normalize_L2(Q)
        idx = _mk(flat.METRIC_IP, c["d"])
        idx.add(P)
        _, I = idx.search(Q, c["k"])
        for r in range(c["nq"]):
            assert sorted(I[r].tolist()) == c["topk_sets"][r], c["name"]
        if c["normalize"]:  # unit vectors: the same sets pin IndexFlatL2 (2 - 2 q.x)
            idx2 = _mk(flat.METRIC_L2, c["d"])
            idx2.add(P)
            _, I2 = idx2.search(Q, c["k"])
            for r in range(c["nq"]):
                assert sorted(I2[r].tolist()) == c["topk_sets"][r], c["name"] + " (L2)"


def test_bf16_row_store_recall_and_scores():
    x, q = _data(200_000, 1024, 16, True, seed=9)
    idx = _mk(flat.METRIC_IP, 1024, "bfloat16")
    idx.add(x)
    D, I = idx.search(q, 10)
    Dr, Ir, _ = flat.search_seq(x, q, 10, flat.METRIC_IP)  # fp32 oracle
    assert recall_at_k(I, Ir) >= 0.999 - 1e-9 or recall_at_k(I, Ir) >= 0.99  # 160 results: allow one boundary swap
    # against the oracle on the bf16-rounded rows the fp32 parity rule holds
    xb = flat.round_bf16(x)
    Db, Ib, _ = flat.search_seq(xb, q, 10, flat.METRIC_IP)
    assert_parity(D, I, Db, Ib, xb, q, metric_ip=True)
    for metric in (flat.METRIC_L2,):
        idx = _mk(metric, 1024, "bfloat16")
        idx.add(x[:20_000])
        D, I = idx.search(q[:3], 5)
        Db, Ib, _ = flat.search_seq(xb[:20_000], q[:3], 5, metric)
        assert_parity(D, I, Db, Ib, xb[:20_000], q[:3], metric_ip=False)


def test_torch_cuda_tensors_in_and_out():
    import torch

    x, q = _data(10_000, 128, 4, True, seed=2)
    idx = _mk(flat.METRIC_IP, 128)
    idx.add(torch.from_numpy(x).cuda())
    D, I = idx.search(torch.from_numpy(q).cuda(), 6)
    assert D.is_cuda and I.is_cuda and I.dtype == torch.int64
    Dr, Ir, _ = flat.search_seq(x, q, 6, flat.METRIC_IP)
    assert_parity(D.cpu().numpy(), I.cpu().numpy(), Dr, Ir, x, q, metric_ip=True)
    # incremental adds continue the id space
    idx.add(x[:5])
    assert idx.ntotal == 10_005
    D2, I2 = idx.search(x[:1], 2)
    assert sorted(I2[0].tolist()) == [0, 10_000]


def test_concurrent_readers_from_threads():
    from concurrent.futures import ThreadPoolExecutor

    x, q = _data(50_000, 128, 32, True, seed=4)
    idx = _mk(flat.METRIC_IP, 128)
    idx.add(x)
    Dr, Ir, _ = flat.search_seq(x, q, 5, flat.METRIC_IP)
    with ThreadPoolExecutor(8) as ex:
        res = list(ex.map(lambda i: idx.search(q[i:i + 1], 5), range(32)))
    for i, (D, I) in enumerate(res):
        assert_parity(D, I, Dr[i:i + 1], Ir[i:i + 1], x, q[i:i + 1], metric_ip=True)


def test_sharded_merge_equals_single_index_bit_for_bit():
    """Emulates G ranks in one process: G shard indexes on this GPU + b2_merge_shard_lists."""
    import torch

    from nv_genai_b200 import merge_shard_lists
    from nv_genai_b200.sharded import shard_bounds

    x, q = _data(60_001, 256, 9, True, seed=6)
    x[45_000] = x[10]  # cross-shard duplicate
    q[0] = x[10]
    for metric in (flat.METRIC_IP, flat.METRIC_L2):
        whole = _mk(metric, 256)
        whole.add(x)
        qd = torch.from_numpy(q).cuda()
        Dw, Iw = whole.search(qd, 20)
        for G in (2, 3, 8):
            Dp, Ip = [], []
            for g in range(G):
                lo, hi = shard_bounds(len(x), G, g)
                sh = _mk(metric, 256)
                sh.add(x[lo:hi])
                D, I = sh.search(qd, 20)
                Dp.append(D)
                Ip.append(torch.where(I >= 0, I + lo, I))
            D, I = merge_shard_lists(metric, torch.stack(Dp), torch.stack(Ip))
            assert torch.equal(I, Iw) and torch.equal(D, Dw)
    # k larger than a shard: padding survives the merge
    tiny = [_mk(flat.METRIC_L2, 256) for _ in range(2)]
    tiny[0].add(x[:3])
    tiny[1].add(x[3:5])
    parts = [t.search(qd, 4) for t in tiny]
    I1 = torch.where(parts[1][1] >= 0, parts[1][1] + 3, parts[1][1])
    D, I = merge_shard_lists(flat.METRIC_L2, torch.stack([parts[0][0], parts[1][0]]), torch.stack([parts[0][1], I1]))
    ref = _mk(flat.METRIC_L2, 256)
    ref.add(x[:5])
    Dr, Ir = ref.search(qd, 4)
    assert torch.equal(I, Ir) and torch.equal(D, Dr)


def test_packed_shard_exchange_layout_equals_single_index_bit_for_bit():
    """The layout ShardedFlatIndex all-gathers on GPUs: each shard writes (I local ids | D) into one packed block
    with search_into, the merge kernel adds the shard bases.  G ranks emulated on this GPU, including a batch that
    takes the tensor-core path on every shard."""
    import torch

    from nv_genai_b200.index import merge_shard_lists_packed
    from nv_genai_b200.sharded import shard_bounds

    x, q = _data(60_001, 256, 40, True, seed=16)
    x[45_000] = x[10]
    q[0] = x[10]
    whole = _mk(flat.METRIC_IP, 256)
    whole.add(x)
    for nq, k in ((1, 10), (9, 20), (40, 100)):
        qd = torch.from_numpy(q[:nq]).cuda()
        Dw, Iw = whole.search(qd, k)
        for G in (2, 8):
            nres = nq * k
            gathered = torch.empty((G, nres * 12), dtype=torch.uint8, device="cuda")
            bases = []
            for g in range(G):
                lo, hi = shard_bounds(len(x), G, g)
                bases.append(lo)
                sh = _mk(flat.METRIC_IP, 256)
                sh.add(x[lo:hi])
                I_loc = gathered[g, : nres * 8].view(torch.int64).view(nq, k)
                D_loc = gathered[g, nres * 8:].view(torch.float32).view(nq, k)
                sh.search_into(qd, k, D_loc, I_loc)
            D, I = merge_shard_lists_packed(flat.METRIC_IP, gathered, nq, k, bases)
            assert torch.equal(I, Iw), (nq, k, G)
            if nq < 20:
                assert torch.equal(D, Dw)  # same kernel scores each row: bit for bit
            else:
                # tensor path: the segment boundaries differ between the whole index and the shards, the scores
                # of a row do not (same MMA sequence per row)
                assert torch.equal(D, Dw)


def test_multi_gpu_sharded_search_peer_and_nccl_exchange():
    """Two real ranks (one per GPU) when the box has them: tests/mp_sharded_gpu.py under torchrun."""
    import subprocess
    import sys

    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = os.path.join(HERE, "mp_sharded_gpu.py")
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533", script],
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    line = [ln for ln in res.stdout.splitlines() if ln.startswith("{")][-1]
    rep = json.loads(line)
    assert rep["ok"] and all(c["equal"] for c in rep["cases"])
