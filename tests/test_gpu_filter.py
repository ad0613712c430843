"""GPU: b2_index_search_filtered / b2_filter_results — langchain's post-search selection on the device — against the
oracle's restatement (oracle.flat.select_hits) applied to b2_index_search's own results."""
import ctypes

import numpy as np
import pytest

from fakes import HashEmbeddings
from oracle import flat

pytestmark = pytest.mark.gpu


def _pack(allow_bool):
    n = len(allow_bool)
    bits = np.zeros(((n + 31) // 32) * 32, np.uint8)
    bits[:n] = allow_bool
    return np.packbits(bits.reshape(-1, 32), axis=1, bitorder="little").view("<u4").reshape(-1).astype(np.uint32)


@pytest.mark.parametrize("metric", ["ip", "l2"])
def test_search_filtered_equals_search_then_oracle_selection(metric):
    import torch

    from nv_genai_b200 import IndexFlatIP, IndexFlatL2
    from nv_genai_b200 import _native as N

    n, d, nq = 20_000, 256, 37
    x = flat.synth(0, n, d, flat.CORPUS_SEED)
    q = flat.synth(0, nq, d, flat.QUERY_SEED)
    if metric == "ip":
        flat.normalize_L2(x)
        flat.normalize_L2(q)
    idx = (IndexFlatIP if metric == "ip" else IndexFlatL2)(d, device=0)
    idx.add(x)
    m = flat.METRIC_IP if metric == "ip" else flat.METRIC_L2
    rng = np.random.default_rng(5)
    allow = rng.random(n) < 0.3
    words = _pack(allow)
    rel_kinds = (N.REL_MAX_INNER_PRODUCT, N.REL_COSINE) if metric == "ip" else (N.REL_EUCLIDEAN,)
    for k, fetch_k in ((4, 20), (10, 10), (7, 100), (33, 64), (1, 1)):
        D0, I0 = idx.search(q, fetch_k)
        mid = float(np.median(D0[:, fetch_k // 2]))  # a raw threshold that cuts some lists and not others
        cases = [dict(), dict(allow=words), dict(score_threshold=mid), dict(allow=words, score_threshold=mid)]
        for rk in rel_kinds:
            fn = flat._REL_FN[rk]
            rthr = float(np.median([fn(float(s)) for s in D0[:, min(k, fetch_k) // 2]]))
            cases += [dict(relevance=rk, relevance_threshold=rthr),
                      dict(allow=words, score_threshold=mid, relevance=rk, relevance_threshold=rthr)]
        for kw in cases:
            want = flat.select_hits(D0, I0, k, m, allow if "allow" in kw else None, kw.get("score_threshold"),
                                    kw.get("relevance", 0), kw.get("relevance_threshold", 0.0))
            # host buffers in, host buffers out
            got = idx.search_filtered(q, k, fetch_k, **kw)
            for g, w in zip(got, want):
                assert np.array_equal(g, w), (k, fetch_k, kw.keys())
            # device tensors in, device tensors out (mask uploaded once)
            kwd = dict(kw)
            if "allow" in kwd:
                kwd["allow"] = idx.upload_mask(words)
            gd = idx.search_filtered(torch.from_numpy(q).cuda(), k, fetch_k, **kwd)
            for g, w in zip(gd, want):
                assert np.array_equal(g.cpu().numpy(), w), (k, fetch_k, kw.keys())


def test_filter_results_on_crafted_lists_and_bad_arguments():
    import torch

    from nv_genai_b200 import _native as N

    dev = torch.device("cuda", 0)
    Din = np.array([[0.9, 0.5, 0.2, -0.1, -0.4, -3.4e38], [0.3, 0.1, -3.4e38, -3.4e38, -3.4e38, -3.4e38]], np.float32)
    Iin = np.array([[7, 3, 12, 5, 9, -1], [40, 2, -1, -1, -1, -1]], np.int64)
    Dd, Id = torch.from_numpy(Din).to(dev), torch.from_numpy(Iin).to(dev)
    allow = np.zeros(41, bool)
    allow[[3, 9, 40]] = True
    mask = torch.from_numpy(_pack(allow).view(np.int32)).to(dev)

    def run(k, mask_t=None, bits=0, raw=None, rel=0, rthr=0.0, metric=N.METRIC_INNER_PRODUCT):
        Do = torch.empty((2, k), dtype=torch.float32, device=dev)
        Io = torch.empty((2, k), dtype=torch.int64, device=dev)
        C = torch.empty((2,), dtype=torch.int32, device=dev)
        N.check(N.lib.b2_filter_results(metric, 2, 6, k, Dd.data_ptr(), Id.data_ptr(),
                                        mask_t.data_ptr() if mask_t is not None else None, bits,
                                        0 if raw is None else 1, 0.0 if raw is None else raw, rel, rthr,
                                        Do.data_ptr(), Io.data_ptr(), C.data_ptr(), 0, None))
        torch.cuda.synchronize()
        return Do.cpu().numpy(), Io.cpu().numpy(), C.cpu().numpy()

    for kw in (dict(k=6), dict(k=2), dict(k=1, mask_t=mask, bits=64), dict(k=3, mask_t=mask, bits=64),
               dict(k=4, raw=0.2), dict(k=3, rel=N.REL_MAX_INNER_PRODUCT, rthr=0.45),
               dict(k=6, mask_t=mask, bits=64, raw=-0.5, rel=N.REL_MAX_INNER_PRODUCT, rthr=0.45)):
        got = run(**kw)
        want = flat.select_hits(Din, Iin, kw["k"], flat.METRIC_IP, allow if "mask_t" in kw else None, kw.get("raw"),
                                kw.get("rel", 0), kw.get("rthr", 0.0))
        for g, w in zip(got, want):
            assert np.array_equal(g, w), kw
    # rows beyond allow_bits count as filtered out: bits = 32 hides row 40
    _, Io, C = run(6, mask, 32)
    assert C.tolist() == [2, 0] and Io[0, :2].tolist() == [3, 9]

    out = torch.empty((2, 6), dtype=torch.float32, device=dev)
    oi = torch.empty((2, 6), dtype=torch.int64, device=dev)
    args = (Dd.data_ptr(), Id.data_ptr(), None, 0, 0, 0.0)
    assert N.lib.b2_filter_results(0, 2, 6, 7, *args, 0, 0.0, out.data_ptr(), oi.data_ptr(), None, 0, None) == N.B2_ERR_INVALID
    assert b"fetch_k must be >= k" in N.lib.b2_last_error()
    assert N.lib.b2_filter_results(0, 2, 6, 0, *args, 0, 0.0, out.data_ptr(), oi.data_ptr(), None, 0, None) == N.B2_ERR_INVALID
    assert N.lib.b2_filter_results(0, 2, 6, 3, *args, 9, 0.0, out.data_ptr(), oi.data_ptr(), None, 0, None) == N.B2_ERR_INVALID
    assert N.lib.b2_filter_results(5, 2, 6, 3, *args, 0, 0.0, out.data_ptr(), oi.data_ptr(), None, 0, None) == N.B2_ERR_INVALID
    assert N.lib.b2_filter_results(0, 2, 6, 3, None, Id.data_ptr(), None, 0, 0, 0.0, 0, 0.0, out.data_ptr(), oi.data_ptr(),
                                   None, 0, None) == N.B2_ERR_INVALID
    assert N.lib.b2_filter_results(0, 2, 4096, 3, *args, 0, 0.0, out.data_ptr(), oi.data_ptr(), None, 0, None) == N.B2_ERR_UNSUPPORTED


def test_vectorstore_filters_and_thresholds_run_on_the_device_and_match_the_host_loop():
    from nv_genai_b200 import FAISS, DistanceStrategy, IndexFlatIP, IndexFlatL2, InMemoryDocstore
    from nv_genai_b200 import _native as N

    d = 128
    texts = [f"passage {i} about subject {i % 11}" for i in range(500)]
    metas = [{"source": f"doc{i % 4}.pdf", "page": i} for i in range(500)]
    emb = HashEmbeddings(d)
    for ip in (False, True):
        vs = FAISS(emb, (IndexFlatIP if ip else IndexFlatL2)(d, device=0), InMemoryDocstore(), {},
                   distance_strategy=DistanceStrategy.MAX_INNER_PRODUCT if ip else DistanceStrategy.EUCLIDEAN_DISTANCE)
        vs.add_texts(texts, metas)
        for qi in (3, 250):
            q = emb.embed_query(texts[qi])
            full = vs.similarity_search_with_score_by_vector(q, k=60)
            mid = float(full[25][1])
            for k, fetch_k, thr in ((4, 20, None), (8, 60, mid), (5, 5, None)):
                kw = {} if thr is None else {"score_threshold": thr}
                N.launch_count_reset()
                a = vs.similarity_search_with_score_by_vector(q, k=k, filter={"source": ["doc1.pdf", "doc3.pdf"]},
                                                              fetch_k=fetch_k, **kw)
                dev_launches = N.launch_count()
                b = vs.similarity_search_with_score_by_vector(q, k=k, filter=lambda m: m["page"] % 2 == 1,
                                                              fetch_k=fetch_k, **kw)
                assert [(x.page_content, float(s)) for x, s in a] == [(x.page_content, float(s)) for x, s in b]
                assert dev_launches >= 3  # scan + merge + selection kernel
        # batch form: counts per query, relevance threshold like the chain server's retriever
        X = np.array([emb.embed_query(texts[i]) for i in (3, 250, 499)], np.float32)
        Dm, Im, Cm = vs.search_batch(X, 4, filter={"source": "doc2.pdf"}, fetch_k=100)
        assert Cm.tolist() == [4, 4, 4] and all(i % 4 == 2 for i in Im.reshape(-1))
        r = vs.as_retriever(search_type="similarity_score_threshold", search_kwargs={"score_threshold": 0.25, "k": 6})
        import warnings

        Dr, Ir, Cr = vs.search_batch(X, 6, relevance_score_threshold=0.25)
        for row, qi in enumerate((3, 250, 499)):
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                want = [x.page_content for x in r.invoke(texts[qi])]
            assert [texts[i] for i in Ir[row, :Cr[row]]] == want
