"""GPU: the reference's call sites replayed through the FAISS-compatible class on the CUDA index."""
import asyncio
import pickle

import numpy as np
import pytest

from fakes import HashEmbeddings
from oracle import flat

pytestmark = pytest.mark.gpu
D = 1024  # config.embeddings.dimensions default (configuration.py:121-125)
TEXTS = [f"passage {i}: the quick brown fox number {i * 7919 % 1013}" for i in range(300)]


def test_chain_server_flow_l2_threshold_retriever():
    from nv_genai_b200 import FAISS, Document, IndexFlatL2, InMemoryDocstore

    emb = HashEmbeddings(D)
    vs = FAISS(emb, IndexFlatL2(D), InMemoryDocstore(), {})  # common/utils.py:217
    docs = [Document(page_content=t, metadata={"source": f"doc{i % 5}.pdf"}) for i, t in enumerate(TEXTS)]
    vs.add_documents(docs)  # chains.py:61
    assert vs.index.ntotal == 300
    x = np.array(emb.embed_documents(TEXTS), np.float32)
    qv = np.array([emb.embed_query(TEXTS[42])], np.float32)
    Dr, Ir = flat.search(x, qv, 4, flat.METRIC_L2)
    got = vs.similarity_search_with_score(TEXTS[42], k=4)
    assert [d.page_content for d, _ in got] == [TEXTS[i] for i in Ir[0]]
    np.testing.assert_allclose([float(s) for _, s in got], Dr[0], rtol=1e-5, atol=1e-6)
    r = vs.as_retriever(search_type="similarity_score_threshold", search_kwargs={"score_threshold": 0.25, "k": 4})
    with pytest.warns(UserWarning):
        hits = r.get_relevant_documents(TEXTS[42], callbacks=[None])
    assert [h.page_content for h in hits] == [TEXTS[42]]
    # delete by source, as common/utils.py:362-377 does
    victims = [k for k, d in vs.docstore._dict.items() if d.metadata["source"] == "doc2.pdf"]
    assert vs.delete(victims) and vs.index.ntotal == 240
    assert vs.similarity_search(TEXTS[43], k=1)[0].page_content == TEXTS[43]
    assert all(d.metadata["source"] != "doc2.pdf" for d in vs.similarity_search(TEXTS[2], k=50))


def test_five_minute_rag_flow_from_documents_pickle_retriever(tmp_path):
    from nv_genai_b200 import FAISS, Document

    emb = HashEmbeddings(D)
    vs = FAISS.from_documents([Document(page_content=t) for t in TEXTS], emb)  # main.py:90
    with open(tmp_path / "vectorstore.pkl", "wb") as f:
        pickle.dump(vs, f)  # main.py:93-94
    with open(tmp_path / "vectorstore.pkl", "rb") as f:
        vs2 = pickle.load(f)  # main.py:78-79
    docs = vs2.as_retriever().invoke(TEXTS[7])  # main.py:125-126
    assert len(docs) == 4 and docs[0].page_content == TEXTS[7]


def test_save_local_load_local_and_merge(tmp_path):
    from nv_genai_b200 import FAISS

    emb = HashEmbeddings(D)
    a = FAISS.from_texts(TEXTS[:100], emb, metadatas=[{"i": i} for i in range(100)])
    b = FAISS.from_texts(TEXTS[100:180], emb)
    a.save_local(str(tmp_path), index_name="idx")
    c = FAISS.load_local(str(tmp_path), emb, index_name="idx", allow_dangerous_deserialization=True)
    assert c.index.ntotal == 100 and c.similarity_search(TEXTS[5], k=1)[0].metadata == {"i": 5}
    c.merge_from(b)
    assert c.index.ntotal == 180 and b.index.ntotal == 0
    assert c.similarity_search(TEXTS[150], k=1)[0].page_content == TEXTS[150]


def test_inner_product_store_with_array_fast_path_and_batch_search():
    import torch

    from nv_genai_b200 import FAISS, DistanceStrategy

    emb = HashEmbeddings(D)
    x = flat.synth(0, 20_000, D, 77)
    flat.normalize_L2(x)
    texts = [f"row {i}" for i in range(20_000)]
    vs = FAISS.from_embeddings((texts, torch.from_numpy(x).cuda()), emb,
                               distance_strategy=DistanceStrategy.MAX_INNER_PRODUCT)
    assert vs.index.metric_type == 0 and vs.index.ntotal == 20_000
    q = flat.synth(0, 16, D, 78)
    flat.normalize_L2(q)
    Dg, Ig = vs.search_batch(q, 10)
    Dr, Ir, _ = flat.search_seq(x, q, 10, flat.METRIC_IP)
    assert np.array_equal(Ig, Ir)
    pairs = vs.similarity_search_with_score_by_vector(q[0].tolist(), k=3)
    assert [p[0].page_content for p in pairs] == [f"row {i}" for i in Ir[0, :3]]
    rel = vs._select_relevance_score_fn()
    assert abs(rel(0.2) - 0.8) < 1e-12


def test_normalize_l2_store_and_mmr_and_async():
    from nv_genai_b200 import FAISS

    emb = HashEmbeddings(D)
    vs = FAISS.from_texts(TEXTS, emb, normalize_L2=True)
    rows = vs.index.reconstruct_n(0, 10)
    np.testing.assert_allclose(np.linalg.norm(rows, axis=1), 1.0, rtol=1e-6)
    s = vs.similarity_search_with_score(TEXTS[3], k=2)
    assert s[0][0].page_content == TEXTS[3] and abs(float(s[0][1])) < 1e-6 and float(s[1][1]) <= 4.0
    docs = vs.max_marginal_relevance_search(TEXTS[11], k=4, fetch_k=20)
    assert len(docs) == 4 and docs[0].page_content == TEXTS[11]
    # a query that is NOT a stored row: with a self-query lambda*cos(q,e) - (1-lambda)*cos(e,e_self) is
    # identically zero for lambda = 0.5 and rounding noise alone would pick the order
    probe = "a question that was never ingested"
    docs = vs.max_marginal_relevance_search(probe, k=4, fetch_k=20)
    x = vs.index.reconstruct_n(0, 300)
    want, _ = flat.mmr_search(x, np.array(emb.embed_query(probe), np.float32), 4, 20, 0.5, flat.METRIC_L2)
    assert [d.page_content for d in docs] == [TEXTS[i] for i in want]
    filt = vs.max_marginal_relevance_search(TEXTS[11], k=3, fetch_k=20, filter=lambda m: True)
    assert len(filt) == 3 and filt[0].page_content == TEXTS[11]

    async def many():
        qs = [emb.embed_query(TEXTS[i]) for i in range(16)]
        return await asyncio.gather(*[vs.asimilarity_search_by_vector(embedding=q, k=1) for q in qs])

    res = asyncio.run(many())
    assert [r[0].page_content for r in res] == TEXTS[:16]


def test_concurrent_single_vector_searches_are_coalesced_into_batched_scans():
    """faiss_vdb_service.py:169-174 fires many asimilarity_search_by_vector at once: with request batching on, they
    share scans of the corpus (fewer passes over the rows) and return what the one-at-a-time searches return."""
    from nv_genai_b200 import FAISS, IndexFlatIP, InMemoryDocstore, DistanceStrategy, _native

    emb = HashEmbeddings(D)
    n = 20_000
    x = flat.synth(0, n, D, 77)
    flat.normalize_L2(x)
    vs = FAISS(emb, IndexFlatIP(D), InMemoryDocstore(), {}, distance_strategy=DistanceStrategy.MAX_INNER_PRODUCT)
    vs.add_embeddings(([f"row {i}" for i in range(n)], x))
    qs = flat.synth(0, 48, D, 78)
    flat.normalize_L2(qs)
    want = [vs.similarity_search_with_score_by_vector(q.tolist(), k=5) for q in qs]
    _native.prof_enable(True)  # counts the scan kernels = passes over the rows
    _native.prof_read(reset=True)
    for q in qs[:8]:
        vs.similarity_search_with_score_by_vector(q.tolist(), k=5)
    alone = _native.prof_read(reset=True)[1] / 8
    assert alone == 1

    vs.enable_request_batching(max_batch=16, max_wait_us=20_000)
    try:
        async def fire():
            return await asyncio.gather(*[vs.asimilarity_search_with_score_by_vector(q.tolist(), k=5) for q in qs])

        got = asyncio.run(fire())
        passes = _native.prof_read(reset=True)[1]
    finally:
        vs.disable_request_batching()
        _native.prof_enable(False)
    assert passes < alone * len(qs) * 0.6, (passes, alone)  # scans were shared
    for a, b in zip(got, want):
        assert [d.page_content for d, _ in a] == [d.page_content for d, _ in b]
        np.testing.assert_allclose([float(s) for _, s in a], [float(s) for _, s in b], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("name", ["flat_l2_3x4.faiss", "flat_ip_2x3.faiss"])
def test_read_index_loads_a_hand_assembled_faiss_file_and_writes_it_back(tmp_path, name):
    """tests/golden/*.faiss: assembled from faiss's header layout outside io.py (make_faiss_fixture.py)."""
    import os
    import sys

    from nv_genai_b200.io import read_index, write_index

    golden = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    sys.path.insert(0, golden)
    import make_faiss_fixture as fx

    rows = fx.L2_ROWS if "l2" in name else fx.IP_ROWS
    idx = read_index(os.path.join(golden, name))
    assert idx.ntotal == len(rows) and idx.d == len(rows[0])
    assert idx.metric_type == (1 if "l2" in name else 0)
    assert np.array_equal(idx.reconstruct_n(0, idx.ntotal), np.array(rows, dtype=np.float32))
    out = tmp_path / name
    write_index(idx, str(out))
    assert out.read_bytes() == open(os.path.join(golden, name), "rb").read()
