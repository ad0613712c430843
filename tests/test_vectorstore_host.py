"""CPU: host logic of the FAISS-compatible class, replaying the reference's call sites
with a deterministic fake Embeddings and an oracle-backed index (the arithmetic is not under test here;
the GPU tests run the same scenarios through the CUDA index)."""
import asyncio
import os
import pickle
import struct
import warnings

import numpy as np
import pytest

from fakes import HashEmbeddings, OracleIndex
from nv_genai_b200 import FAISS, DistanceStrategy, Document, InMemoryDocstore, VectorStoreRetriever
from nv_genai_b200.index import METRIC_INNER_PRODUCT, METRIC_L2
from nv_genai_b200.io import write_index
from oracle import flat

D = 64
TEXTS = [f"chunk number {i} about topic {i % 7}" for i in range(40)]


def make_store(ip=False, normalize=False, texts=TEXTS, metadatas=None):
    """The shape of common/utils.py:217: FAISS(embedder, IndexFlatL2(dim), InMemoryDocstore(), {})."""
    emb = HashEmbeddings(D)
    vs = FAISS(emb, OracleIndex(D, METRIC_INNER_PRODUCT if ip else METRIC_L2), InMemoryDocstore(), {},
               normalize_L2=normalize,
               distance_strategy=DistanceStrategy.MAX_INNER_PRODUCT if ip else DistanceStrategy.EUCLIDEAN_DISTANCE)
    if texts:
        vs.add_texts(texts, metadatas)
    return vs, emb


def test_constructor_shape_of_the_chain_server_and_add_documents():
    vs, emb = make_store(texts=None)
    assert vs.index.ntotal == 0 and vs.docstore._dict == {} and vs.index_to_docstore_id == {}
    docs = [Document(page_content=t, metadata={"source": f"f{i % 3}.txt"}) for i, t in enumerate(TEXTS)]
    ids = vs.add_documents(docs)  # nvidia_api_catalog/chains.py:61
    assert len(ids) == 40 and vs.index.ntotal == 40 and len(vs.docstore._dict) == 40
    assert [vs.index_to_docstore_id[i] for i in range(40)] == ids
    # common/utils.py:342 walks vectorstore.docstore._dict and reads doc.metadata["source"]
    assert {d.metadata["source"] for d in vs.docstore._dict.values()} == {"f0.txt", "f1.txt", "f2.txt"}


def test_empty_store_search_returns_nothing():  # multi_turn_rag searches a fresh conversation store
    vs, _ = make_store(texts=None)
    assert vs.similarity_search("anything", k=4) == []
    assert vs.as_retriever().invoke("anything") == []


def test_similarity_search_self_hit_and_scores():
    vs, emb = make_store()
    docs = vs.similarity_search(TEXTS[5], k=4)  # main.py:125-126 default retriever path
    assert docs[0].page_content == TEXTS[5] and len(docs) == 4
    pairs = vs.similarity_search_with_score(TEXTS[5], k=3)
    assert pairs[0][1] == 0.0 and pairs[0][1] <= pairs[1][1] <= pairs[2][1]  # squared L2 ascending
    x = np.array(emb.embed_documents(TEXTS), dtype=np.float32)
    Dr, Ir = flat.search(x, np.array([emb.embed_query(TEXTS[5])], np.float32), 3, flat.METRIC_L2)
    assert [p[0].page_content for p in pairs] == [TEXTS[i] for i in Ir[0]]
    assert [float(p[1]) for p in pairs] == [float(v) for v in Dr[0]]


def test_k_larger_than_store_skips_padding():
    vs, _ = make_store(texts=TEXTS[:2])  # 2-row adds happen per turn in multi_turn_rag/chains.py:62-67
    assert len(vs.similarity_search("q", k=4)) == 2


def test_retriever_variants():
    vs, _ = make_store()
    r = vs.as_retriever()
    assert isinstance(r, VectorStoreRetriever) and r.search_type == "similarity"
    assert len(r.get_relevant_documents(TEXTS[0], callbacks=[object()])) == 4  # chains.py:123 passes callbacks
    r3 = vs.as_retriever(search_kwargs={"k": 3})
    assert len(r3.invoke(TEXTS[0])) == 3
    with pytest.raises(ValueError):
        vs.as_retriever(search_type="similarity_score_threshold", search_kwargs={"k": 4})
    with pytest.raises(ValueError):
        vs.as_retriever(search_type="similarity_score_threshold", search_kwargs={"score_threshold": 1})
    with pytest.raises(ValueError):
        vs.as_retriever(search_type="nope")
    mm = vs.as_retriever(search_type="mmr", search_kwargs={"k": 3, "fetch_k": 10})
    assert len(mm.invoke(TEXTS[1])) == 3


def test_relevance_threshold_retriever_chain_server_defaults():
    """nvidia_api_catalog/chains.py:122: search_type=similarity_score_threshold, score_threshold=0.25, k=4;
    relevance = 1 - d/sqrt(2) on the SQUARED distance, keep >= threshold, warn when empty."""
    vs, emb = make_store()
    r = vs.as_retriever(search_type="similarity_score_threshold", search_kwargs={"score_threshold": 0.25, "k": 4})
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        docs = r.get_relevant_documents(TEXTS[9])
    assert [d.page_content for d in docs] == [TEXTS[9]]  # only the exact hit has relevance 1.0; others are < 0
    assert any("between 0 and 1" in str(x.message) for x in w)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        assert r.get_relevant_documents("text that was never ingested") == []
    assert any("No relevant docs were retrieved" in str(x.message) for x in w)
    pairs = vs.similarity_search_with_relevance_scores(TEXTS[9], k=2)
    raw = vs.similarity_search_with_score(TEXTS[9], k=2)
    assert pairs[0][1] == 1.0 and abs(pairs[1][1] - (1.0 - float(raw[1][1]) / np.sqrt(2))) < 1e-4


def test_raw_score_threshold_semantics():
    """oran 2_Evaluation_Metrics.py:217: as_retriever(search_kwargs={"k":3,"score_threshold":0.5}) with the
    default search_type -> raw-distance threshold: L2 keeps dist <= thr, inner product keeps score >= thr."""
    vs, _ = make_store()
    assert [d.page_content for d in vs.as_retriever(search_kwargs={"k": 3, "score_threshold": 0.5}).invoke(TEXTS[2])] \
        == [TEXTS[2]]
    vip, emb = make_store(ip=True, normalize=False)
    q = TEXTS[2]
    self_score = float(np.dot(emb.embed_query(q), emb.embed_query(q)))
    got = vip.similarity_search_with_score(q, k=3, score_threshold=self_score - 1e-3)
    assert [d.page_content for d, _ in got] == [q]
    assert vip._select_relevance_score_fn()(0.25) == 0.75 and vip._select_relevance_score_fn()(-0.5) == 0.5


def test_metadata_filter_dict_list_and_callable():
    metas = [{"source": f"f{i % 3}.txt", "page": i} for i in range(40)]
    vs, _ = make_store(metadatas=metas)
    docs = vs.similarity_search(TEXTS[0], k=4, filter={"source": "f1.txt"}, fetch_k=40)
    assert len(docs) == 4 and all(d.metadata["source"] == "f1.txt" for d in docs)
    docs = vs.similarity_search(TEXTS[0], k=40, filter={"source": ["f1.txt", "f2.txt"]}, fetch_k=40)
    assert len(docs) == 26 and all(d.metadata["source"] != "f0.txt" for d in docs)
    docs = vs.similarity_search(TEXTS[0], k=5, filter=lambda m: m["page"] < 3, fetch_k=40)
    assert sorted(d.metadata["page"] for d in docs) == [0, 1, 2]
    # only fetch_k candidates are inspected (upstream behaviour)
    assert len(vs.similarity_search(TEXTS[0], k=4, filter={"page": 39}, fetch_k=1)) in (0, 1)


def test_add_embeddings_pairs_and_array_fast_path_and_errors():
    vs, emb = make_store(texts=None)
    vecs = emb.embed_documents(TEXTS[:5])
    ids = vs.add_embeddings(list(zip(TEXTS[:5], vecs)), ids=["a", "b", "c", "d", "e"])
    assert ids == ["a", "b", "c", "d", "e"] and vs.index.ntotal == 5
    vs.add_embeddings((TEXTS[5:9], np.array(emb.embed_documents(TEXTS[5:9]), np.float32)))
    assert vs.index.ntotal == 9 and vs.similarity_search(TEXTS[7], k=1)[0].page_content == TEXTS[7]
    with pytest.raises(ValueError, match="Duplicate ids"):
        vs.add_texts(["x", "y"], ids=["z", "z"])
    with pytest.raises(ValueError, match="already exist"):
        vs.add_texts(["x"], ids=["a"])
    with pytest.raises(ValueError, match="equal length"):
        vs.add_texts(["x", "y"], metadatas=[{}])
    assert vs.index.ntotal == 9  # failed adds leave no orphan rows


def test_from_texts_from_embeddings_from_documents_pick_the_index_type():
    emb = HashEmbeddings(D)
    import nv_genai_b200.vectorstore as V

    orig_l2, orig_ip = V._ix.IndexFlatL2, V._ix.IndexFlatIP
    V._ix.IndexFlatL2 = lambda d, **kw: OracleIndex(d, METRIC_L2)
    V._ix.IndexFlatIP = lambda d, **kw: OracleIndex(d, METRIC_INNER_PRODUCT)
    try:
        a = FAISS.from_texts(TEXTS[:6], emb, metadatas=[{"i": i} for i in range(6)])  # notebooks 02/05/06/08
        assert a.index.metric_type == METRIC_L2 and a.index.d == D and a.index.ntotal == 6
        b = FAISS.from_documents([Document(page_content=t) for t in TEXTS[:6]], emb,
                                 distance_strategy=DistanceStrategy.MAX_INNER_PRODUCT)  # main.py:90 shape
        assert b.index.metric_type == METRIC_INNER_PRODUCT
        c = FAISS.from_embeddings(list(zip(TEXTS[:3], emb.embed_documents(TEXTS[:3]))), emb, ids=["1", "2", "3"])
        assert list(c.index_to_docstore_id.values()) == ["1", "2", "3"]
    finally:
        V._ix.IndexFlatL2, V._ix.IndexFlatIP = orig_l2, orig_ip


def test_normalize_l2_flag_normalises_rows_and_queries():
    vs, emb = make_store(normalize=True, texts=None)
    import nv_genai_b200.vectorstore as V

    orig = V._ix.normalize_L2
    V._ix.normalize_L2 = flat.normalize_L2  # query normalisation runs on the device in the product
    try:
        vs.add_texts(TEXTS[:10])
        np.testing.assert_allclose(np.linalg.norm(vs.index._x, axis=1), 1.0, rtol=1e-6)
        pairs = vs.similarity_search_with_score(TEXTS[3], k=2)
        assert pairs[0][0].page_content == TEXTS[3] and abs(pairs[0][1]) < 1e-6 and pairs[1][1] <= 4.0
    finally:
        V._ix.normalize_L2 = orig
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        FAISS(emb, OracleIndex(D, METRIC_INNER_PRODUCT), InMemoryDocstore(), {}, normalize_L2=True,
              distance_strategy=DistanceStrategy.MAX_INNER_PRODUCT)
    assert any("Normalizing L2 is not applicable" in str(x.message) for x in w)


def test_delete_renumbers_like_upstream():
    """common/utils.py:370-376: find ids by metadata in docstore._dict, then vectorstore.delete(ids)."""
    metas = [{"source": f"f{i % 4}.txt"} for i in range(40)]
    vs, emb = make_store(metadatas=metas)
    ids_before = dict(vs.index_to_docstore_id)
    victims = [k for k, d in vs.docstore._dict.items() if d.metadata["source"] == "f2.txt"]
    assert vs.delete(victims) is True
    assert vs.index.ntotal == 30 and len(vs.docstore._dict) == 30
    survivors = [ids_before[i] for i in range(40) if i % 4 != 2]
    assert [vs.index_to_docstore_id[i] for i in range(30)] == survivors  # old order, renumbered 0..n-1
    hit = vs.similarity_search(TEXTS[7], k=1)[0]
    assert hit.page_content == TEXTS[7]  # row/doc association survives compaction
    assert all(d.metadata["source"] != "f2.txt" for d in vs.similarity_search(TEXTS[2], k=30))
    with pytest.raises(ValueError, match="No ids provided"):
        vs.delete(None)
    with pytest.raises(ValueError, match="do not exist"):
        vs.delete(["not-an-id"])


def test_merge_from_offsets_ids_and_empties_target():  # oran vectorstore_updater.py:215-216
    a, _ = make_store(texts=TEXTS[:10])
    b, _ = make_store(texts=TEXTS[10:25])
    b_ids = [b.index_to_docstore_id[i] for i in range(15)]
    a.merge_from(b)
    assert a.index.ntotal == 25 and b.index.ntotal == 0
    assert [a.index_to_docstore_id[10 + i] for i in range(15)] == b_ids
    assert a.similarity_search(TEXTS[20], k=1)[0].page_content == TEXTS[20]


def test_pickle_whole_store_and_serialize_roundtrip():  # main.py:78-79,93-94 pickles the whole object
    vs, emb = make_store()
    vs2 = pickle.loads(pickle.dumps(vs))
    assert vs2.index.ntotal == 40 and vs2.similarity_search(TEXTS[4], k=1)[0].page_content == TEXTS[4]
    blob = vs.serialize_to_bytes()
    vs3 = FAISS.deserialize_from_bytes(blob, emb)
    assert vs3.index_to_docstore_id == vs.index_to_docstore_id


def test_index_file_format_is_faiss_flat_layout(tmp_path):
    vs, _ = make_store(texts=TEXTS[:7])
    p = tmp_path / "index.faiss"
    write_index(vs.index, str(p))
    raw = p.read_bytes()
    assert raw[:4] == b"IxF2"
    d, n, d1, d2, trained, metric = struct.unpack("<iqqqBi", raw[4:37])
    assert (d, n, d1, d2, trained, metric) == (D, 7, 1 << 20, 1 << 20, 1, 1)
    (size,) = struct.unpack("<Q", raw[37:45])
    assert size == 7 * D and len(raw) == 45 + 4 * 7 * D
    assert np.array_equal(np.frombuffer(raw[45:], "<f4").reshape(7, D), vs.index._x)
    with pytest.raises(ValueError, match="allow_dangerous_deserialization"):
        FAISS.load_local(str(tmp_path), HashEmbeddings(D))


def test_async_by_vector_under_gather():  # faiss_vdb_service.py:169-174
    vs, emb = make_store()

    async def many():
        qs = [emb.embed_query(TEXTS[i]) for i in range(8)]
        return await asyncio.gather(*[vs.asimilarity_search_by_vector(embedding=q, k=2) for q in qs])

    res = asyncio.run(many())
    assert [r[0].page_content for r in res] == TEXTS[:8]


def test_mmr_search_paths():
    metas = [{"g": i % 2} for i in range(40)]
    vs, emb = make_store(metadatas=metas)
    docs = vs.max_marginal_relevance_search(TEXTS[3], k=4, fetch_k=12)
    assert len(docs) == 4 and docs[0].page_content == TEXTS[3]
    pairs = vs.max_marginal_relevance_search_with_score_by_vector(emb.embed_query(TEXTS[3]), k=3, fetch_k=8)
    x = np.array(emb.embed_documents(TEXTS), np.float32)
    want_ids, want_sc = flat.mmr_search(x, np.array(emb.embed_query(TEXTS[3]), np.float32), 3, 8, 0.5, flat.METRIC_L2)
    assert [p[0].page_content for p in pairs] == [TEXTS[i] for i in want_ids]
    assert [float(p[1]) for p in pairs] == want_sc


def test_plain_function_embedding_still_works_with_a_warning(caplog):
    emb = HashEmbeddings(D)
    vs = FAISS(emb.embed_query, OracleIndex(D), InMemoryDocstore(), {})
    vs.add_texts(TEXTS[:5])
    assert vs.embeddings is None and vs.similarity_search(TEXTS[1], k=1)[0].page_content == TEXTS[1]


def test_rejects_foreign_index_objects():
    with pytest.raises(TypeError):
        FAISS(HashEmbeddings(D), object(), InMemoryDocstore(), {})


# ---------------------------------------------------------------------------------------------------------------
# request coalescing (nv_genai_b200/batching.py), host logic on CPU with the oracle-backed index
def test_request_batching_coalesces_concurrent_single_vector_searches():
    import threading

    from nv_genai_b200.batching import BatchingIndex

    d, n = 32, 500
    x = flat.synth(0, n, d, 123)
    base = OracleIndex(d, METRIC_L2)
    base.add(x)
    calls = []
    orig = base.search

    def counting_search(q, k):
        calls.append(np.atleast_2d(q).shape[0])
        return orig(q, k)

    base.search = counting_search
    proxy = BatchingIndex(base, max_batch=8, max_wait_us=200_000)  # generous window: the test must coalesce
    try:
        assert proxy.d == d and proxy.ntotal == n  # the faiss-shaped interface passes through
        qs = flat.synth(0, 8, d, 456)
        ks = [4, 1, 7, 4, 4, 2, 4, 3]
        out = [None] * 8

        def worker(i):
            out[i] = proxy.search(qs[i:i + 1], ks[i])

        threads = [threading.Thread(target=worker, args=(i,)) for i in range(8)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=30)
        assert proxy.requests_served == 8 and proxy.batches_run < 8  # at least some were merged
        assert sum(calls) == 8 and max(calls) > 1
        for i in range(8):
            D, I = out[i]
            Dr, Ir = flat.search(x, qs[i:i + 1], ks[i], METRIC_L2)
            assert D.shape == (1, ks[i]) and np.array_equal(I, Ir) and np.allclose(D, Dr, rtol=1e-6)
        # a call that already is a batch, and k validation, bypass / behave like the wrapped index
        D, I = proxy.search(qs[:3], 2)
        assert D.shape == (3, 2)
        with pytest.raises(AssertionError):
            proxy.search(qs[:1], 0)
        with pytest.raises(AssertionError):
            proxy.search(np.zeros((1, d + 1), np.float32), 2)
    finally:
        proxy.close()


def test_vectorstore_with_request_batching_gives_the_same_documents():
    from concurrent.futures import ThreadPoolExecutor

    emb = HashEmbeddings(48)
    texts = [f"document number {i}" for i in range(200)]
    vs = FAISS(emb, OracleIndex(48, METRIC_L2), InMemoryDocstore(), {})
    vs.add_texts(texts)
    want = [vs.similarity_search_with_score(t, k=3) for t in texts[:20]]
    vs.enable_request_batching(max_batch=16, max_wait_us=50_000)
    try:
        with ThreadPoolExecutor(max_workers=20) as ex:
            got = list(ex.map(lambda t: vs.similarity_search_with_score(t, k=3), texts[:20]))
        for a, b in zip(got, want):
            assert [d.page_content for d, _ in a] == [d.page_content for d, _ in b]
            assert np.allclose([s for _, s in a], [s for _, s in b], rtol=1e-6)
        assert vs.index.batches_run < 20
        data = vs.serialize_to_bytes()  # the proxy (a thread) is not what gets pickled
        assert isinstance(data, bytes)
    finally:
        vs.disable_request_batching()
    assert isinstance(vs.index, OracleIndex)


# ---------------------------------------------------------------- selection after the search
def test_select_hits_oracle_known_answers():
    """The restatement of langchain's post-search selection on hand-computable lists."""
    Din = np.array([[0.9, 0.5, 0.2, -0.1, -0.4, -3.4e38]], np.float32)
    Iin = np.array([[7, 3, 12, 5, 9, -1]], np.int64)
    # no options: first k, padding skipped
    Do, Io, C = flat.select_hits(Din, Iin, 8, flat.METRIC_IP)
    assert C.tolist() == [5] and Io[0].tolist() == [7, 3, 12, 5, 9, -1, -1, -1] and Do[0, 5] == -flat.FLT_MAX
    # metadata filter (rows 3 and 9 allowed), then cut at k
    allow = np.zeros(16, bool)
    allow[[3, 9]] = True
    _, Io, C = flat.select_hits(Din, Iin, 1, flat.METRIC_IP, allow)
    assert C.tolist() == [1] and Io[0].tolist() == [3]
    # raw threshold: inner product keeps >=, L2 keeps <=
    _, Io, C = flat.select_hits(Din, Iin, 4, flat.METRIC_IP, raw_threshold=0.2)
    # the comparison is float(np.float32(0.2)) >= 0.2 in double: float32(0.2) = 0.20000000298 passes
    assert Io[0, :C[0]].tolist() == [7, 3, 12]
    Dl2 = np.array([[0.0, 0.3, 0.6, 2.5]], np.float32)
    _, Io, C = flat.select_hits(Dl2, np.array([[4, 2, 8, 1]], np.int64), 4, flat.METRIC_L2, raw_threshold=0.6000001)
    assert Io[0, :C[0]].tolist() == [4, 2, 8]
    # the max-inner-product relevance function is not monotone: 1 - d for d > 0, -d otherwise; the cut at k
    # comes BEFORE the relevance threshold (similarity_search_with_relevance_scores filters the k it got)
    _, Io, C = flat.select_hits(Din, Iin, 3, flat.METRIC_IP, relevance=flat.REL_MAX_INNER_PRODUCT,
                                relevance_threshold=0.45)
    assert Io[0, :C[0]].tolist() == [3, 12]  # relevances 0.1, 0.5, 0.8 | rows 5, 9 are beyond k
    _, Io, C = flat.select_hits(Dl2, np.array([[4, 2, 8, 1]], np.int64), 4, flat.METRIC_L2,
                                relevance=flat.REL_EUCLIDEAN, relevance_threshold=0.25)
    assert Io[0, :C[0]].tolist() == [4, 2, 8]  # 1 - 2.5 / sqrt(2) < 0.25


def test_dict_filters_and_thresholds_select_like_the_per_hit_host_loop():
    """A dict filter runs as a row bitmask + search_filtered; the same predicate as a callable runs upstream's
    per-hit loop on the host.  Both must return the same documents and scores."""
    metas = [{"source": f"f{i % 3}.txt", "page": i} for i in range(40)]
    for ip in (False, True):
        vs, emb = make_store(ip=ip, metadatas=metas)
        for qi in (0, 7, 23):
            q = emb.embed_query(TEXTS[qi])
            full = vs.similarity_search_with_score_by_vector(q, k=40)
            mid = float(full[10][1])
            for k, fetch_k, thr in ((4, 20, None), (4, 40, mid), (10, 12, mid), (3, 3, None)):
                kw = {} if thr is None else {"score_threshold": thr}
                a = vs.similarity_search_with_score_by_vector(q, k=k, filter={"source": ["f1.txt", "f2.txt"]},
                                                              fetch_k=fetch_k, **kw)
                b = vs.similarity_search_with_score_by_vector(q, k=k, filter=lambda m: m["source"] != "f0.txt",
                                                              fetch_k=fetch_k, **kw)
                assert [(d.page_content, float(s)) for d, s in a] == [(d.page_content, float(s)) for d, s in b]
            # threshold without a filter: only k are searched, as upstream does
            a = vs.similarity_search_with_score_by_vector(q, k=12, score_threshold=mid)
            want = [(d.page_content, float(s)) for d, s in full[:12] if (s >= mid if ip else s <= mid)]
            assert [(d.page_content, float(s)) for d, s in a] == want and 0 < len(want) < 12


def test_filter_bitmask_cache_follows_mutations():
    metas = [{"source": f"f{i % 2}.txt"} for i in range(40)]
    vs, emb = make_store(metadatas=metas)
    flt = {"source": "f1.txt"}
    assert len(vs.similarity_search(TEXTS[1], k=40, filter=flt, fetch_k=40)) == 20
    mask = vs._device_mask(flt, build=False)
    assert mask is not None and int(np.unpackbits(mask.view(np.uint8)).sum()) == 20
    # add: the cached mask is stale, the next search rebuilds it
    vs.add_texts(["a new chunk"], [{"source": "f1.txt"}])
    assert vs._device_mask(flt, build=False) is None
    assert len(vs.similarity_search(TEXTS[1], k=41, filter=flt, fetch_k=41)) == 21
    # delete renumbers rows: stale again
    victim = vs.index_to_docstore_id[1]
    vs.delete([victim])
    assert vs._device_mask(flt, build=False) is None
    docs = vs.similarity_search(TEXTS[3], k=40, filter=flt, fetch_k=40)
    assert len(docs) == 20 and all(d.metadata["source"] == "f1.txt" for d in docs)
    # above device_filter_auto_rows a dict filter stays on the host path until it is prepared explicitly
    vs.clear_filter_cache()
    vs.device_filter_auto_rows = 10
    assert len(vs.similarity_search(TEXTS[3], k=40, filter=flt, fetch_k=40)) == 20
    assert vs._device_mask(flt, build=False) is None
    vs.prepare_filter(flt)
    assert vs._device_mask(flt, build=False) is not None
    assert len(vs.similarity_search(TEXTS[3], k=40, filter=flt, fetch_k=40)) == 20
    # the cache is not part of the pickled state
    assert "_filter_masks" not in vs.__getstate__()
    with pytest.raises(ValueError):
        vs.prepare_filter(lambda m: True)


def test_search_batch_with_filter_and_thresholds_returns_counts():
    metas = [{"source": f"f{i % 3}.txt"} for i in range(40)]
    vs, emb = make_store(metadatas=metas)
    X = np.array([emb.embed_query(TEXTS[i]) for i in (0, 5, 11)], np.float32)
    D0, I0 = vs.search_batch(X, 6)
    assert D0.shape == (3, 6) and I0[:, 0].tolist() == [0, 5, 11]
    Dm, Im, Cm = vs.search_batch(X, 4, filter={"source": "f2.txt"}, fetch_k=40)
    assert Cm.tolist() == [4, 4, 4] and all(i % 3 == 2 for i in Im.reshape(-1))
    assert Im[1, 0] == 5 and Im[2, 0] == 11
    # relevance threshold as the chain server's retriever applies it (configuration.py:146: 0.25), per query on the host
    r = vs.as_retriever(search_type="similarity_score_threshold", search_kwargs={"score_threshold": 0.25, "k": 6})
    Dr, Ir, Cr = vs.search_batch(X, 6, relevance_score_threshold=0.25)
    for row, qi in enumerate((0, 5, 11)):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            want = [d.page_content for d in r.invoke(TEXTS[qi])]
        assert [TEXTS[i] for i in Ir[row, :Cr[row]]] == want
        assert (Ir[row, Cr[row]:] == -1).all()
    # raw threshold
    thr = float(D0[0, 3])
    Dt, It, Ct = vs.search_batch(X[:1], 6, score_threshold=thr)
    assert Ct.tolist() == [4] and It[0, :4].tolist() == I0[0, :4].tolist()
    with pytest.raises(ValueError):
        vs.search_batch(X, 4, filter=lambda m: True)


# ------------------------------------------------------------------ on-disk formats pinned from outside io.py
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name,metric,rows", [
    ("flat_l2_3x4.faiss", METRIC_L2, [[1.0, 0.0, -0.5, 2.0], [0.25, 0.75, 3.0, -1.0], [0.0, 0.0, 0.0, 8.0]]),
    ("flat_ip_2x3.faiss", METRIC_INNER_PRODUCT, [[0.5, -0.5, 0.125], [4.0, 1.0, -2.0]]),
])
def test_write_index_reproduces_the_hand_assembled_faiss_file(tmp_path, name, metric, rows):
    """tests/golden/*.faiss were assembled byte by byte from faiss's header layout (make_faiss_fixture.py)."""
    want = open(os.path.join(GOLDEN, name), "rb").read()
    idx = OracleIndex(len(rows[0]), metric)
    idx.add(np.array(rows, dtype=np.float32))
    p = tmp_path / name
    write_index(idx, str(p))
    assert p.read_bytes() == want


class _PydanticV1Doc:
    """What langchain-core 0.1.29's Document does when pickled and unpickled (pydantic v1 BaseModel)."""

    def __getstate__(self):
        return {"__dict__": self.__dict__.copy(), "__fields_set__": set(self.__dict__),
                "__private_attribute_values__": {}}

    def __setstate__(self, state):
        assert set(state) == {"__dict__", "__fields_set__", "__private_attribute_values__"}
        object.__setattr__(self, "__dict__", state["__dict__"])


class _PlainDocstore:
    def __init__(self, d):
        self._dict = d


def _with_stub_langchain(fn):
    """Run fn() with importable stand-ins for langchain's two classes under their upstream module paths."""
    import sys
    import types

    doc = type("Document", (_PydanticV1Doc,), {"__module__": "langchain_core.documents.base"})
    store = type("InMemoryDocstore", (_PlainDocstore,), {"__module__": "langchain_community.docstore.in_memory"})
    mods = {"langchain_core": None, "langchain_core.documents": None, "langchain_core.documents.base": doc,
            "langchain_community": None, "langchain_community.docstore": None,
            "langchain_community.docstore.in_memory": store}
    try:
        for n, cls in mods.items():
            m = types.ModuleType(n)
            if cls is not None:
                setattr(m, cls.__name__, cls)
            sys.modules[n] = m
        return fn(doc, store)
    finally:
        for n in mods:
            sys.modules.pop(n, None)


def test_langchain_written_index_pkl_loads_without_langchain(tmp_path):
    """a pydantic-v1 Document pickle must come back with page_content and metadata."""
    def write(doc, store):
        d1, d2 = doc.__new__(doc), doc.__new__(doc)
        d1.__dict__.update(page_content="alpha", metadata={"source": "a.pdf"}, type="Document")
        d2.__dict__.update(page_content="beta", metadata={}, type="Document")
        return pickle.dumps((store({"id-1": d1, "id-2": d2}), {0: "id-1", 1: "id-2"}))

    blob = _with_stub_langchain(write)
    assert b"langchain_core.documents.base" in blob and b"langchain_community.docstore.in_memory" in blob
    import io as _io

    from nv_genai_b200.vectorstore import _CompatUnpickler

    docstore, id_map = _CompatUnpickler(_io.BytesIO(blob)).load()
    assert id_map == {0: "id-1", 1: "id-2"}
    got = docstore.search("id-1")
    assert got.page_content == "alpha" and got.metadata == {"source": "a.pdf"}
    assert docstore.search("id-2").page_content == "beta" and docstore.search("id-2").metadata == {}


def test_save_local_compat_langchain_loads_upstream(tmp_path):
    """The way back: a store saved here with compat="langchain" unpickles under langchain's own class paths."""
    vs, emb = make_store(texts=TEXTS[:5], metadatas=[{"n": i} for i in range(5)])
    vs.save_local(str(tmp_path), compat="langchain")
    raw = (tmp_path / "index.pkl").read_bytes()
    assert b"nv_genai_b200" not in raw and b"langchain_core.documents.base" in raw

    def load(doc, store):
        docstore, id_map = pickle.loads(raw)  # plain pickle, as upstream FAISS.load_local does
        assert type(docstore) is store and all(type(v) is doc for v in docstore._dict.values())
        return {k: (v.page_content, v.metadata) for k, v in docstore._dict.items()}, id_map

    docs, id_map = _with_stub_langchain(load)
    assert id_map == vs.index_to_docstore_id
    assert docs == {k: (v.page_content, v.metadata) for k, v in vs.docstore._dict.items()}
    assert (tmp_path / "index.faiss").read_bytes()[:4] == b"IxF2"
    with pytest.raises(ValueError):
        vs.save_local(str(tmp_path), compat="nope")
