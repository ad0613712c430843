"""GPU: the C-ABI shard group (b2_shard_group_* / b2_shard_search) against one index holding the same rows and against
the oracle.  Shards may share a device, so the whole protocol - query push, per-shard search storing its lists into the
root's receive blocks, flag-waiting merge with shard bases - runs on a single GPU; with more GPUs visible the same
test spreads the shards over them."""
import numpy as np
import pytest

from oracle import flat
from parity import assert_parity

pytestmark = pytest.mark.gpu


def _devices(n):
    import torch

    g = torch.cuda.device_count()
    return [i % g for i in range(n)]


@pytest.mark.parametrize("metric", [flat.METRIC_IP, flat.METRIC_L2])
@pytest.mark.parametrize("dtype", ["float32", "bfloat16", "bfloat16+residual"])
def test_shard_group_equals_single_index_and_oracle(metric, dtype):
    from nv_genai_b200 import IndexFlat, ShardGroup

    # 4 shards of 12.5k rows: every shard and the single index take the same kernel for a given batch (a shard under
    # 4096 rows would fall back to the FP32-pipe scan, whose L2 arithmetic differs in the last bit from the GEMM form)
    n, d, per = 50_000, 256, 12_500
    x = flat.synth(0, n, d, 21)
    flat.normalize_L2(x)
    q = flat.synth(0, 40, d, 22)
    flat.normalize_L2(q)
    q[5] = x[12_499]  # planted on a shard boundary
    q[6] = x[12_500]
    grp = ShardGroup(d, metric, _devices(4), per, dtype=dtype)
    grp.add(x[:30_000])
    grp.add(x[30_000:])
    assert grp.ntotal == n and [grp.shard_ntotal(i) for i in range(4)] == [12_500] * 4
    one = IndexFlat(d, metric, dtype)
    one.add(x)
    xs = x
    if dtype.startswith("bfloat16"):
        xs = flat.round_bf16(x)
        if dtype.endswith("+residual"):
            xs = xs + flat.round_bf16(x - xs)
    for nq, k in ((1, 4), (16, 10), (40, 100)):
        D, I = grp.search(q[:nq], k)
        D1, I1 = one.search(q[:nq], k)
        assert np.array_equal(I, I1) and np.array_equal(D, D1), (nq, k)
        Dr, Ir = flat.search(xs, q[:nq], k, metric)
        assert_parity(D, I, Dr, Ir, xs, q[:nq], metric_ip=metric == flat.METRIC_IP)
    if metric == flat.METRIC_IP:
        D, I = grp.search(q[:16], 10)
        assert I[5, 0] == 12_499 and I[6, 0] == 12_500


def test_shard_group_device_tensors_many_queries_and_mutation():
    import torch

    from nv_genai_b200 import IndexFlatIP, ShardGroup

    n, d = 30_000, 128
    x = flat.synth(0, n, d, 31)
    q = flat.synth(0, 5000, d, 32)  # more than one pass of 4096 queries
    grp = ShardGroup(d, flat.METRIC_IP, _devices(3), 10_000)
    dev = f"cuda:{grp.devices[0]}"
    grp.add(torch.from_numpy(x).to(dev), normalize=True)
    one = IndexFlatIP(d, device=grp.devices[0])
    one.add(x, normalize=True)
    qt = torch.from_numpy(q).to(dev)
    D, I = grp.search(qt, 7)
    D1, I1 = one.search(qt, 7)
    assert torch.equal(I, I1) and torch.equal(D, D1)
    # two searches back to back on different streams
    s1 = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(s1):
        Da, Ia = grp.search(qt[:8], 5)
    Db, Ib = grp.search(qt[8:16], 5)
    torch.cuda.synchronize()
    assert torch.equal(Ia, I1[:8, :5]) and torch.equal(Ib, I1[8:16, :5])
    # remove_ids takes global ids and renumbers like faiss
    drop = [0, 9_999, 10_000, 25_000]
    assert grp.remove_ids(drop) == 4 and one.remove_ids(drop) == 4
    assert grp.ntotal == n - 4 and [grp.shard_ntotal(i) for i in range(3)] == [9_998, 9_999, 9_999]
    D, I = grp.search(q[:16], 7)
    D1, I1 = one.search(q[:16], 7)
    assert np.array_equal(I, I1) and np.array_equal(D, D1)
    grp.reset()
    assert grp.ntotal == 0
    D, I = grp.search(q[:2], 3)
    assert np.all(I == -1)


def test_shard_group_argument_errors():
    from nv_genai_b200 import ShardGroup

    with pytest.raises(ValueError):
        ShardGroup(64, flat.METRIC_IP, [], 100)
    with pytest.raises(ValueError):
        ShardGroup(64, flat.METRIC_IP, [0], 0)
    grp = ShardGroup(64, flat.METRIC_IP, [0, 0], 100)
    with pytest.raises(AssertionError):
        grp.search(np.zeros((1, 32), np.float32), 3)


def test_langchain_faiss_class_over_a_shard_group(tmp_path):
    """`FAISS(embedder, ShardGroup(...), InMemoryDocstore(), {})`: the reference's constructor shape
    (/root/reference/RetrievalAugmentedGeneration/common/utils.py:217) with the index sharded over GPUs must answer
    like the same store over one IndexFlatL2: search with scores, thresholds, MMR, delete, save_local."""
    from fakes import HashEmbeddings
    from nv_genai_b200 import FAISS, IndexFlatL2, InMemoryDocstore, ShardGroup

    d = 64
    emb = HashEmbeddings(d)
    texts = [f"chunk {i} about topic {i % 7}" for i in range(300)]
    metas = [{"topic": i % 7} for i in range(300)]
    one = FAISS(emb, IndexFlatL2(d), InMemoryDocstore(), {})
    grp = FAISS(emb, ShardGroup(d, flat.METRIC_L2, _devices(3), 100), InMemoryDocstore(), {})
    ids = [f"id-{i}" for i in range(300)]
    one.add_texts(texts, metas, ids=ids)
    grp.add_texts(texts[:150], metas[:150], ids=ids[:150])
    grp.add_texts(texts[150:], metas[150:], ids=ids[150:])
    assert grp.index.ntotal == 300 and [grp.index.shard_ntotal(i) for i in range(3)] == [100, 100, 100]
    for query in ("chunk 17 about topic 3", "something else entirely"):
        a = one.similarity_search_with_score(query, k=5)
        b = grp.similarity_search_with_score(query, k=5)
        assert [(x.page_content, float(s)) for x, s in a] == [(x.page_content, float(s)) for x, s in b]
        a = one.similarity_search_with_score(query, k=4, filter={"topic": 3}, fetch_k=50)
        b = grp.similarity_search_with_score(query, k=4, filter={"topic": 3}, fetch_k=50)
        assert [x.page_content for x, _ in a] == [x.page_content for x, _ in b]
    # MMR with queries that are not stored texts (a stored copy makes every later MMR score an exact tie at zero)
    for query in ("a question about topic 3", "something else entirely"):
        a = one.max_marginal_relevance_search(query, k=4, fetch_k=20)
        b = grp.max_marginal_relevance_search(query, k=4, fetch_k=20)
        assert [x.page_content for x in a] == [x.page_content for x in b]
    assert grp.similarity_search(texts[42], k=1)[0].page_content == texts[42]
    # delete renumbers like faiss across the shard boundaries
    gone = ["id-0", "id-99", "id-100", "id-250"]
    one.delete(gone)
    grp.delete(gone)
    assert grp.index.ntotal == 296 and grp.index_to_docstore_id == one.index_to_docstore_id
    a = one.similarity_search_with_score(texts[101], k=3)
    b = grp.similarity_search_with_score(texts[101], k=3)
    assert [(x.page_content, float(s)) for x, s in a] == [(x.page_content, float(s)) for x, s in b]
    # saved in faiss's flat format: loads back into a single index
    grp.save_local(str(tmp_path))
    back = FAISS.load_local(str(tmp_path), emb, allow_dangerous_deserialization=True)
    assert back.index.ntotal == 296
    assert back.similarity_search(texts[200], k=1)[0].page_content == texts[200]
