"""`FAISS`: the LangChain vector-store API over the B200 flat index.

Host-side mirror of `langchain_community.vectorstores.faiss.FAISS` — the class the reference
constructs at /root/reference/RetrievalAugmentedGeneration/common/utils.py:217 and
/root/reference/examples/5_mins_rag_no_gpu/main.py:90 — with the same method names, argument
meaning, return types and error behaviour (call sites in SURVEY.md
3.5).  Text handling, the docstore and score post-processing are plain Python exactly as
upstream; everything that touches vectors (`index.add`, `faiss.normalize_L2`,
`index.search`, `index.reconstruct`, MMR, `remove_ids`, `merge_from`) runs in the CUDA
library through include/b2_flat.h.

Additions over upstream: `add_embeddings` also accepts
`(texts, ndarray | torch.Tensor)` so that millions of rows never become Python lists, and
`search_batch(X, k) -> (D, I)` is the batched entry point with no per-row Python objects.
"""
from __future__ import annotations

import asyncio
import functools
import logging
import math
import operator
import os
import pickle
import uuid
import warnings
from pathlib import Path
from typing import Any, Callable, Dict, Iterable, List, Optional, Sized, Tuple, Union

import numpy as np

from . import index as _ix
from .documents import HAVE_LANGCHAIN, DistanceStrategy, Document, Embeddings, InMemoryDocstore
from .io import read_index, write_index

logger = logging.getLogger(__name__)


def _len_check_if_sized(x: Any, y: Any, x_name: str, y_name: str) -> None:
    if isinstance(x, Sized) and isinstance(y, Sized) and len(x) != len(y):
        raise ValueError(f"{x_name} and {y_name} expected to be equal length but len({x_name})={len(x)} and "
                         f"len({y_name})={len(y)}")


def _is_array(x: Any) -> bool:
    return isinstance(x, np.ndarray) or _ix._is_tensor(x)


class FAISS:
    """Vector store with the rows in B200 HBM; API of LangChain's `FAISS`."""

    def __init__(self, embedding_function: Union[Callable[[str], List[float]], Embeddings], index: Any,
                 docstore: InMemoryDocstore, index_to_docstore_id: Dict[int, str],
                 relevance_score_fn: Optional[Callable[[float], float]] = None, normalize_L2: bool = False,
                 distance_strategy: DistanceStrategy = DistanceStrategy.EUCLIDEAN_DISTANCE):
        if not hasattr(embedding_function, "embed_documents") and not hasattr(embedding_function, "embed_query"):
            logger.warning("`embedding_function` is expected to be an Embeddings object, support for passing in a "
                           "function will soon be removed.")
        from .sharded import ShardGroup  # a flat index sharded over several GPUs of this process serves as well

        if not isinstance(getattr(index, "wrapped", index), (_ix.IndexFlat, ShardGroup)):
            raise TypeError("index must be a nv_genai_b200 IndexFlatL2 / IndexFlatIP or ShardGroup (there is no CPU faiss path)")
        self.embedding_function = embedding_function
        self.index = index
        self.docstore = docstore
        self.index_to_docstore_id = index_to_docstore_id
        self.distance_strategy = distance_strategy
        self.override_relevance_score_fn = relevance_score_fn
        self._normalize_L2 = normalize_L2
        if self.distance_strategy != DistanceStrategy.EUCLIDEAN_DISTANCE and self._normalize_L2:
            warnings.warn(f"Normalizing L2 is not applicable for metric type: {self.distance_strategy}")

    def __getstate__(self):
        state = dict(self.__dict__)
        state.pop("_filter_masks", None)  # device bitmasks are a cache, not state
        return state

    # ------------------------------------------------------------------ embeddings
    @property
    def embeddings(self) -> Optional[Embeddings]:
        ef = self.embedding_function
        return ef if hasattr(ef, "embed_documents") and hasattr(ef, "embed_query") else None

    def _embed_documents(self, texts: List[str]) -> List[List[float]]:
        if hasattr(self.embedding_function, "embed_documents"):
            return self.embedding_function.embed_documents(texts)
        return [self.embedding_function(text) for text in texts]

    def _embed_query(self, text: str) -> List[float]:
        if hasattr(self.embedding_function, "embed_query"):
            return self.embedding_function.embed_query(text)
        return self.embedding_function(text)

    async def _aembed_documents(self, texts: List[str]) -> List[List[float]]:
        ef = self.embedding_function
        if hasattr(ef, "aembed_documents"):
            return await ef.aembed_documents(texts)
        return await _run_in_executor(self._embed_documents, texts)

    async def _aembed_query(self, text: str) -> List[float]:
        ef = self.embedding_function
        if hasattr(ef, "aembed_query"):
            return await ef.aembed_query(text)
        return await _run_in_executor(self._embed_query, text)

    # ------------------------------------------------------------------ add
    def _add(self, texts: Iterable[str], embeddings: Any, metadatas: Optional[Iterable[dict]] = None,
             ids: Optional[List[str]] = None) -> List[str]:
        """upstream FAISS.__add: checks, np.float32 matrix, optional normalise, index.add, docstore, id map."""
        texts = list(texts) if not isinstance(texts, list) else texts
        n_emb = embeddings.shape[0] if _is_array(embeddings) else None
        if n_emb is None:
            embeddings = list(embeddings) if not isinstance(embeddings, list) else embeddings
            n_emb = len(embeddings)
        if len(texts) != n_emb:
            raise ValueError(f"texts and embeddings expected to be equal length but len(texts)={len(texts)} and "
                             f"len(embeddings)={n_emb}")
        _metadatas = list(metadatas) if metadatas is not None else None
        if _metadatas is not None:
            _len_check_if_sized(texts, _metadatas, "texts", "metadatas")
        else:
            _metadatas = [{} for _ in texts]
        if ids is not None:
            _len_check_if_sized(texts, ids, "texts", "ids")
            if len(ids) != len(set(ids)):
                raise ValueError("Duplicate ids found in the ids list.")
        ids = ids if ids is not None else [str(uuid.uuid4()) for _ in texts]
        documents = [Document(page_content=t, metadata=m) for t, m in zip(texts, _metadatas)]
        # validate against the docstore before touching the device so a failure leaves no orphan rows
        clash = set(ids).intersection(self.docstore._dict)
        if clash:
            raise ValueError(f"Tried to add ids that already exist: {clash}")
        if len(texts) == 0:
            return []
        vector = embeddings if _is_array(embeddings) else np.array(embeddings, dtype=np.float32)
        starting_len = len(self.index_to_docstore_id)
        self.__dict__["_mutations"] = self.__dict__.get("_mutations", 0) + 1  # cached filter bitmasks are stale now
        self.index.add(vector, normalize=self._normalize_L2)
        self.docstore.add({id_: doc for id_, doc in zip(ids, documents)})
        self.index_to_docstore_id.update({starting_len + j: id_ for j, id_ in enumerate(ids)})
        return ids

    def add_texts(self, texts: Iterable[str], metadatas: Optional[List[dict]] = None,
                  ids: Optional[List[str]] = None, **kwargs: Any) -> List[str]:
        texts = list(texts)
        embeddings = self._embed_documents(texts)
        return self._add(texts, embeddings, metadatas=metadatas, ids=ids)

    async def aadd_texts(self, texts: Iterable[str], metadatas: Optional[List[dict]] = None,
                         ids: Optional[List[str]] = None, **kwargs: Any) -> List[str]:
        texts = list(texts)
        embeddings = await self._aembed_documents(texts)
        return self._add(texts, embeddings, metadatas=metadatas, ids=ids)

    def add_embeddings(self, text_embeddings: Any, metadatas: Optional[List[dict]] = None,
                       ids: Optional[List[str]] = None, **kwargs: Any) -> List[str]:
        """Upstream form: iterable of (text, vector) pairs.  Fast form: (texts, ndarray | torch.Tensor)."""
        if isinstance(text_embeddings, tuple) and len(text_embeddings) == 2 and _is_array(text_embeddings[1]):
            texts, embeddings = text_embeddings
        else:
            texts, embeddings = zip(*text_embeddings) if text_embeddings else ((), ())
        return self._add(list(texts), embeddings if _is_array(embeddings) else list(embeddings),
                         metadatas=metadatas, ids=ids)

    def add_documents(self, documents: List[Document], **kwargs: Any) -> List[str]:
        """langchain_core VectorStore.add_documents (called at nvidia_api_catalog/chains.py:61)."""
        texts = [doc.page_content for doc in documents]
        metadatas = [doc.metadata for doc in documents]
        return self.add_texts(texts, metadatas, **kwargs)

    async def aadd_documents(self, documents: List[Document], **kwargs: Any) -> List[str]:
        texts = [doc.page_content for doc in documents]
        metadatas = [doc.metadata for doc in documents]
        return await self.aadd_texts(texts, metadatas, **kwargs)

    # ------------------------------------------------------------------ search
    def _query_matrix(self, embedding: Any, normalize: bool) -> np.ndarray:
        vector = np.array([embedding], dtype=np.float32)
        if normalize and self._normalize_L2:
            _ix.normalize_L2(vector)
        return vector

    @staticmethod
    def _build_filter(filter: Union[Callable, Dict[str, Any]]) -> Callable[[dict], bool]:
        if callable(filter):
            return filter
        if not isinstance(filter, dict):
            raise ValueError(f"filter must be a dict of metadata or a callable, not {type(filter)}")
        conds = {key: (value if isinstance(value, list) else [value]) for key, value in filter.items()}

        def filter_func(metadata: dict) -> bool:
            return all(metadata.get(key) in allowed for key, allowed in conds.items())

        return filter_func

    def _doc_at(self, i: int) -> Document:
        _id = self.index_to_docstore_id[i]
        doc = self.docstore.search(_id)
        if not isinstance(doc, Document):
            raise ValueError(f"Could not find document for id {_id}, got {doc}")
        return doc

    # ---- metadata filters as row bitmasks on the device
    #: dict filters are turned into a device bitmask automatically up to this many rows (one pass of Python over the
    #: docstore per distinct filter); above it only after an explicit `prepare_filter(filter)`.
    device_filter_auto_rows = 200_000

    @staticmethod
    def _filter_key(filter: Any) -> Optional[str]:
        if not isinstance(filter, dict):
            return None  # callables are evaluated on the host, hit by hit, as upstream does
        try:
            return repr(sorted((str(key), repr(value)) for key, value in filter.items()))
        except Exception:
            return None

    def _store_stamp(self) -> Tuple[int, int]:
        return (self.__dict__.get("_mutations", 0), len(self.index_to_docstore_id))

    def clear_filter_cache(self) -> None:
        """Drop the cached row bitmasks (call it after editing Document.metadata in place)."""
        self.__dict__["_filter_masks"] = {}

    def prepare_filter(self, filter: Dict[str, Any]):
        """Evaluate a dict metadata filter once over the whole docstore and keep the result on the device as one bit
        per row; later searches with an equal filter (single or `search_batch`) select on the device.  Returns the
        mask (CUDA int32 tensor of ceil(n / 32) words)."""
        key = self._filter_key(filter)
        if key is None:
            raise ValueError("prepare_filter takes a dict of metadata; callables are evaluated per hit on the host")
        base = getattr(self.index, "wrapped", self.index)
        n = len(self.index_to_docstore_id)
        filter_func = self._build_filter(filter)
        bits = np.zeros(((n + 31) // 32) * 32, dtype=np.uint8)
        docs = self.docstore._dict if hasattr(self.docstore, "_dict") else None
        for i in range(n):
            _id = self.index_to_docstore_id[i]
            doc = docs.get(_id) if docs is not None else self.docstore.search(_id)
            if isinstance(doc, Document) and filter_func(doc.metadata):
                bits[i] = 1
        words = np.packbits(bits.reshape(-1, 32), axis=1, bitorder="little").view("<u4").reshape(-1)
        mask = base.upload_mask(words.astype(np.uint32))
        cache = self.__dict__.setdefault("_filter_masks", {})
        if len(cache) >= 16:
            cache.pop(next(iter(cache)))
        cache[key] = (self._store_stamp(), mask)
        return mask

    def _device_mask(self, filter: Any, build: bool):
        """The cached bitmask of `filter` if it is still valid, building it when allowed; else None."""
        key = self._filter_key(filter)
        if key is None:
            return None
        hit = self.__dict__.get("_filter_masks", {}).get(key)
        if hit is not None and hit[0] == self._store_stamp():
            return hit[1]
        if build and (hit is not None or len(self.index_to_docstore_id) <= self.device_filter_auto_rows):
            return self.prepare_filter(filter)
        return None

    def _raw_threshold_is_ge(self) -> bool:
        return self.distance_strategy in (DistanceStrategy.MAX_INNER_PRODUCT, DistanceStrategy.JACCARD)

    def _device_selection_ok(self, base: Any) -> bool:
        """The device applies a raw threshold as `>=` on inner-product indexes and `<=` on L2 indexes; upstream picks the
        direction from distance_strategy.  They agree except for odd pairings (say COSINE over IndexFlatIP)."""
        return self._raw_threshold_is_ge() == (base.metric_type == _ix.METRIC_INNER_PRODUCT)

    def similarity_search_with_score_by_vector(self, embedding: List[float], k: int = 4,
                                               filter: Optional[Union[Callable, Dict[str, Any]]] = None,
                                               fetch_k: int = 20, **kwargs: Any) -> List[Tuple[Document, float]]:
        """Documents closest to `embedding` with their raw score (squared L2, or inner product)."""
        vector = self._query_matrix(embedding, normalize=True)
        score_threshold = kwargs.get("score_threshold")
        base = self.index
        if isinstance(base, _ix.IndexFlat) and (filter is not None or score_threshold is not None) \
                and (score_threshold is None or self._device_selection_ok(base)):
            # selection on the device: metadata filter as a row bitmask, raw threshold, cut at k
            mask = self._device_mask(filter, build=True) if filter is not None else None
            if filter is None or mask is not None:
                n_fetch = k if filter is None else fetch_k
                if n_fetch >= k:
                    D, I, C = base.search_filtered(vector, k, n_fetch, allow=mask, score_threshold=score_threshold)
                    return [(self._doc_at(int(I[0][j])), D[0][j]) for j in range(int(C[0]))]
        scores, indices = self.index.search(vector, k if filter is None else fetch_k)
        docs: List[Tuple[Document, float]] = []
        filter_func = self._build_filter(filter) if filter is not None else None
        for j, i in enumerate(indices[0]):
            if i == -1:  # fewer than k rows in the index
                continue
            doc = self._doc_at(int(i))
            if filter_func is None or filter_func(doc.metadata):
                docs.append((doc, scores[0][j]))
        score_threshold = kwargs.get("score_threshold")
        if score_threshold is not None:
            cmp = (operator.ge if self.distance_strategy in (DistanceStrategy.MAX_INNER_PRODUCT,
                                                             DistanceStrategy.JACCARD) else operator.le)
            docs = [(doc, similarity) for doc, similarity in docs if cmp(similarity, score_threshold)]
        return docs[:k]

    def similarity_search_with_score(self, query: str, k: int = 4,
                                     filter: Optional[Union[Callable, Dict[str, Any]]] = None, fetch_k: int = 20,
                                     **kwargs: Any) -> List[Tuple[Document, float]]:
        embedding = self._embed_query(query)
        return self.similarity_search_with_score_by_vector(embedding, k, filter=filter, fetch_k=fetch_k, **kwargs)

    def similarity_search_by_vector(self, embedding: List[float], k: int = 4, filter: Optional[Any] = None,
                                    fetch_k: int = 20, **kwargs: Any) -> List[Document]:
        pairs = self.similarity_search_with_score_by_vector(embedding, k, filter=filter, fetch_k=fetch_k, **kwargs)
        return [doc for doc, _ in pairs]

    def similarity_search(self, query: str, k: int = 4, filter: Optional[Any] = None, fetch_k: int = 20,
                          **kwargs: Any) -> List[Document]:
        pairs = self.similarity_search_with_score(query, k, filter=filter, fetch_k=fetch_k, **kwargs)
        return [doc for doc, _ in pairs]

    def enable_request_batching(self, max_batch: int = 16, max_wait_us: float = 200.0) -> None:
        """Coalesce concurrent single-query searches (thread-pool / `asyncio.gather` callers such as
        /root/reference/experimental/event-driven-rag-cve-analysis/cyber_dev_day/faiss_vdb_service.py:169-174) into
        batched scans: one pass over the rows serves up to `max_batch` requests.  See nv_genai_b200/batching.py."""
        from .batching import BatchingIndex

        if not isinstance(self.index, BatchingIndex):
            self.index = BatchingIndex(self.index, max_batch=max_batch, max_wait_us=max_wait_us)

    def disable_request_batching(self) -> None:
        from .batching import BatchingIndex

        if isinstance(self.index, BatchingIndex):
            proxy = self.index
            self.index = proxy.wrapped
            proxy.close()

    def _relevance_kind(self) -> int:
        """The device's code for the relevance function `_select_relevance_score_fn` would pick."""
        from . import _native as N

        if self.override_relevance_score_fn is not None:
            raise NotImplementedError("a custom relevance_score_fn cannot run on the device; use the per-query API")
        if self.distance_strategy == DistanceStrategy.MAX_INNER_PRODUCT:
            return N.REL_MAX_INNER_PRODUCT
        if self.distance_strategy == DistanceStrategy.EUCLIDEAN_DISTANCE:
            return N.REL_EUCLIDEAN
        if self.distance_strategy == DistanceStrategy.COSINE:
            return N.REL_COSINE
        raise ValueError("Unknown distance strategy, must be cosine, max_inner_product, or euclidean")

    def search_batch(self, X: Any, k: int = 4, filter: Optional[Dict[str, Any]] = None, fetch_k: int = 20,
                     score_threshold: Optional[float] = None, relevance_score_threshold: Optional[float] = None):
        """Batched fast path: X [B, d] (numpy or torch CUDA) -> (D [B, k], I [B, k]); no Python per row.

        With `filter` (a dict of metadata, turned into a device bitmask by `prepare_filter`), `score_threshold` (raw
        score, as `similarity_search_with_score` takes it) or `relevance_score_threshold` (as a
        `similarity_score_threshold` retriever takes it) the selection langchain runs per query on the host runs on
        the device for the whole batch, and the call returns (D, I, counts): row i holds counts[i] results, raw scores."""
        if self._normalize_L2:
            if _ix._is_tensor(X) and X.is_cuda:
                X = X.detach().float().contiguous().clone()
            else:
                X = np.array(X, dtype=np.float32, copy=True)
            _ix.normalize_L2(X)
        if filter is None and score_threshold is None and relevance_score_threshold is None:
            return self.index.search(X, k)
        from . import _native as N

        base = getattr(self.index, "wrapped", self.index)
        if score_threshold is not None and not self._device_selection_ok(base):
            raise NotImplementedError("raw score_threshold direction of this distance_strategy does not match the index metric")
        mask = None
        if filter is not None:
            if self._filter_key(filter) is None:
                raise ValueError("search_batch takes a dict metadata filter")
            mask = self._device_mask(filter, build=True)
            if mask is None:
                mask = self.prepare_filter(filter)
        rel = self._relevance_kind() if relevance_score_threshold is not None else N.REL_NONE
        return base.search_filtered(X, k, k if filter is None else fetch_k, allow=mask, score_threshold=score_threshold,
                                    relevance=rel, relevance_threshold=(relevance_score_threshold or 0.0))

    def search(self, query: str, search_type: str, **kwargs: Any) -> List[Document]:
        """langchain_core VectorStore.search dispatch."""
        if search_type == "similarity":
            return self.similarity_search(query, **kwargs)
        if search_type == "similarity_score_threshold":
            return [doc for doc, _ in self.similarity_search_with_relevance_scores(query, **kwargs)]
        if search_type == "mmr":
            return self.max_marginal_relevance_search(query, **kwargs)
        raise ValueError(f"search_type of {search_type} not allowed. Expected search_type to be 'similarity', "
                         f"'similarity_score_threshold' or 'mmr'.")

    # ------------------------------------------------------------------ relevance scores
    @staticmethod
    def _euclidean_relevance_score_fn(distance: float) -> float:
        return 1.0 - distance / math.sqrt(2)

    @staticmethod
    def _cosine_relevance_score_fn(distance: float) -> float:
        return 1.0 - distance

    @staticmethod
    def _max_inner_product_relevance_score_fn(distance: float) -> float:
        if distance > 0:
            return 1.0 - distance
        return -1.0 * distance

    def _select_relevance_score_fn(self) -> Callable[[float], float]:
        if self.override_relevance_score_fn is not None:
            return self.override_relevance_score_fn
        if self.distance_strategy == DistanceStrategy.MAX_INNER_PRODUCT:
            return self._max_inner_product_relevance_score_fn
        if self.distance_strategy == DistanceStrategy.EUCLIDEAN_DISTANCE:
            return self._euclidean_relevance_score_fn
        if self.distance_strategy == DistanceStrategy.COSINE:
            return self._cosine_relevance_score_fn
        raise ValueError("Unknown distance strategy, must be cosine, max_inner_product, or euclidean")

    def _similarity_search_with_relevance_scores(self, query: str, k: int = 4, filter: Optional[Any] = None,
                                                 fetch_k: int = 20, **kwargs: Any) -> List[Tuple[Document, float]]:
        relevance_score_fn = self._select_relevance_score_fn()
        if relevance_score_fn is None:
            raise ValueError("normalize_score_fn must be provided to FAISS constructor to normalize scores")
        docs_and_scores = self.similarity_search_with_score(query, k=k, filter=filter, fetch_k=fetch_k, **kwargs)
        return [(doc, relevance_score_fn(score)) for doc, score in docs_and_scores]

    def similarity_search_with_relevance_scores(self, query: str, k: int = 4,
                                                **kwargs: Any) -> List[Tuple[Document, float]]:
        """langchain_core VectorStore.similarity_search_with_relevance_scores: [0,1] scores, `>=` threshold."""
        score_threshold = kwargs.pop("score_threshold", None)
        docs_and_similarities = self._similarity_search_with_relevance_scores(query, k=k, **kwargs)
        if any(similarity < 0.0 or similarity > 1.0 for _, similarity in docs_and_similarities):
            warnings.warn(f"Relevance scores must be between 0 and 1, got {docs_and_similarities}")
        if score_threshold is not None:
            docs_and_similarities = [(doc, similarity) for doc, similarity in docs_and_similarities
                                     if similarity >= score_threshold]
            if len(docs_and_similarities) == 0:
                warnings.warn(f"No relevant docs were retrieved using the relevance score threshold {score_threshold}")
        return docs_and_similarities

    # ------------------------------------------------------------------ MMR
    def max_marginal_relevance_search_with_score_by_vector(self, embedding: List[float], *, k: int = 4,
                                                           fetch_k: int = 20, lambda_mult: float = 0.5,
                                                           filter: Optional[Any] = None
                                                           ) -> List[Tuple[Document, float]]:
        """MMR pick of k among the fetch_k nearest; the query is NOT re-normalised here (as upstream)."""
        vector = self._query_matrix(embedding, normalize=False)
        if filter is None:
            ids, scores = self.index.mmr(vector, k=k, fetch_k=fetch_k, lambda_mult=lambda_mult)
            return [(self._doc_at(int(i)), scores[0][j]) for j, i in enumerate(ids[0]) if i != -1]
        # filtered: search 2*fetch_k, drop non-matching hits, then MMR over the survivors' rows
        scores, indices = self.index.search(vector, fetch_k * 2)
        filter_func = self._build_filter(filter)
        kept = [(int(i), scores[0][j]) for j, i in enumerate(indices[0])
                if i != -1 and filter_func(self._doc_at(int(i)).metadata)]
        if not kept:
            return []
        rows = self.index.reconstruct_batch([i for i, _ in kept])
        picks = _mmr_on_rows(self.index, vector, rows, k, lambda_mult)
        return [(self._doc_at(kept[p][0]), kept[p][1]) for p in picks]

    def max_marginal_relevance_search_by_vector(self, embedding: List[float], k: int = 4, fetch_k: int = 20,
                                                lambda_mult: float = 0.5, filter: Optional[Any] = None,
                                                **kwargs: Any) -> List[Document]:
        pairs = self.max_marginal_relevance_search_with_score_by_vector(embedding, k=k, fetch_k=fetch_k,
                                                                        lambda_mult=lambda_mult, filter=filter)
        return [doc for doc, _ in pairs]

    def max_marginal_relevance_search(self, query: str, k: int = 4, fetch_k: int = 20, lambda_mult: float = 0.5,
                                      filter: Optional[Any] = None, **kwargs: Any) -> List[Document]:
        embedding = self._embed_query(query)
        return self.max_marginal_relevance_search_by_vector(embedding, k=k, fetch_k=fetch_k, lambda_mult=lambda_mult,
                                                            filter=filter, **kwargs)

    def mmr_batch(self, X: Any, k: int = 4, fetch_k: int = 20, lambda_mult: float = 0.5):
        """Batched MMR: X [B, d] -> (ids [B, k], scores [B, k]) in pick order."""
        return self.index.mmr(X, k=k, fetch_k=fetch_k, lambda_mult=lambda_mult)

    # ------------------------------------------------------------------ mutate
    def delete(self, ids: Optional[List[str]] = None, **kwargs: Any) -> Optional[bool]:
        if ids is None:
            raise ValueError("No ids provided to delete.")
        missing_ids = set(ids).difference(self.index_to_docstore_id.values())
        if missing_ids:
            raise ValueError(f"Some specified ids do not exist in the current store. Ids not found: {missing_ids}")
        reversed_index = {id_: idx for idx, id_ in self.index_to_docstore_id.items()}
        index_to_delete = {reversed_index[id_] for id_ in ids}
        self.__dict__["_mutations"] = self.__dict__.get("_mutations", 0) + 1
        self.index.remove_ids(np.fromiter(index_to_delete, dtype=np.int64))
        self.docstore.delete(ids)
        remaining_ids = [id_ for i, id_ in sorted(self.index_to_docstore_id.items()) if i not in index_to_delete]
        self.index_to_docstore_id = {i: id_ for i, id_ in enumerate(remaining_ids)}
        return True

    def merge_from(self, target: "FAISS") -> None:
        if not isinstance(self.docstore, InMemoryDocstore) and not hasattr(self.docstore, "add"):
            raise ValueError("Cannot merge with this type of docstore")
        starting_len = len(self.index_to_docstore_id)
        full_info = []
        for i, target_id in target.index_to_docstore_id.items():
            doc = target.docstore.search(target_id)
            if not isinstance(doc, Document):
                raise ValueError("Document should be returned")
            full_info.append((starting_len + i, target_id, doc))
        clash = {tid for _, tid, _ in full_info}.intersection(self.docstore._dict)
        if clash:
            raise ValueError(f"Tried to add ids that already exist: {clash}")
        self.__dict__["_mutations"] = self.__dict__.get("_mutations", 0) + 1
        target.__dict__["_mutations"] = target.__dict__.get("_mutations", 0) + 1
        self.index.merge_from(getattr(target.index, "wrapped", target.index))
        self.docstore.add({_id: doc for _, _id, doc in full_info})
        self.index_to_docstore_id.update({index: _id for index, _id, _ in full_info})

    # ------------------------------------------------------------------ constructors
    @classmethod
    def _from(cls, texts: Iterable[str], embeddings: Any, embedding: Embeddings,
              metadatas: Optional[Iterable[dict]] = None, ids: Optional[List[str]] = None,
              normalize_L2: bool = False, distance_strategy: DistanceStrategy = DistanceStrategy.EUCLIDEAN_DISTANCE,
              **kwargs: Any) -> "FAISS":
        if _is_array(embeddings):
            dim = int(embeddings.shape[1])
        else:
            embeddings = list(embeddings)
            dim = len(embeddings[0])
        index_kwargs = {key: kwargs.pop(key) for key in ("dtype", "device") if key in kwargs}
        if distance_strategy == DistanceStrategy.MAX_INNER_PRODUCT:
            index = _ix.IndexFlatIP(dim, **index_kwargs)
        else:
            index = _ix.IndexFlatL2(dim, **index_kwargs)  # the default, as upstream
        docstore = kwargs.pop("docstore", InMemoryDocstore())
        index_to_docstore_id = kwargs.pop("index_to_docstore_id", {})
        vecstore = cls(embedding, index, docstore, index_to_docstore_id, normalize_L2=normalize_L2,
                       distance_strategy=distance_strategy, **kwargs)
        vecstore._add(texts, embeddings, metadatas=metadatas, ids=ids)
        return vecstore

    @classmethod
    def from_texts(cls, texts: List[str], embedding: Embeddings, metadatas: Optional[List[dict]] = None,
                   ids: Optional[List[str]] = None, **kwargs: Any) -> "FAISS":
        embeddings = embedding.embed_documents(texts)
        return cls._from(texts, embeddings, embedding, metadatas=metadatas, ids=ids, **kwargs)

    @classmethod
    async def afrom_texts(cls, texts: List[str], embedding: Embeddings, metadatas: Optional[List[dict]] = None,
                          ids: Optional[List[str]] = None, **kwargs: Any) -> "FAISS":
        if hasattr(embedding, "aembed_documents"):
            embeddings = await embedding.aembed_documents(texts)
        else:
            embeddings = await _run_in_executor(embedding.embed_documents, texts)
        return cls._from(texts, embeddings, embedding, metadatas=metadatas, ids=ids, **kwargs)

    @classmethod
    def from_embeddings(cls, text_embeddings: Any, embedding: Embeddings, metadatas: Optional[Iterable[dict]] = None,
                        ids: Optional[List[str]] = None, **kwargs: Any) -> "FAISS":
        if isinstance(text_embeddings, tuple) and len(text_embeddings) == 2 and _is_array(text_embeddings[1]):
            texts, embeddings = text_embeddings
        else:
            texts, embeddings = zip(*text_embeddings)
            embeddings = list(embeddings)
        return cls._from(list(texts), embeddings, embedding, metadatas=metadatas, ids=ids, **kwargs)

    @classmethod
    def from_documents(cls, documents: List[Document], embedding: Embeddings, **kwargs: Any) -> "FAISS":
        """langchain_core VectorStore.from_documents (examples/5_mins_rag_no_gpu/main.py:90)."""
        texts = [d.page_content for d in documents]
        metadatas = [d.metadata for d in documents]
        return cls.from_texts(texts, embedding, metadatas=metadatas, **kwargs)

    @classmethod
    async def afrom_documents(cls, documents: List[Document], embedding: Embeddings, **kwargs: Any) -> "FAISS":
        texts = [d.page_content for d in documents]
        metadatas = [d.metadata for d in documents]
        return await cls.afrom_texts(texts, embedding, metadatas=metadatas, **kwargs)

    # ------------------------------------------------------------------ persistence
    def save_local(self, folder_path: str, index_name: str = "index", compat: Optional[str] = None) -> None:
        """`<index_name>.faiss` in faiss's own IndexFlat file format + `<index_name>.pkl` (docstore, id map).

        `compat="langchain"` writes the pickle under LangChain's class paths
        (`langchain_community.docstore.in_memory.InMemoryDocstore`, `langchain_core.documents.base.Document` with the
        pydantic-v1 state langchain-core 0.1.x uses), so that the folder loads with upstream
        `FAISS.load_local` + faiss-cpu as well - the way back for the stores the reference's apps keep on disk
        (/root/reference/experimental/oran-chatbot-multimodal/vectorstore/vectorstore_updater.py:184,216,220,
        /root/reference/examples/5_mins_rag_no_gpu/main.py:78-79,93-94).  The default keeps this package's own classes."""
        if compat not in (None, "langchain"):
            raise ValueError("compat must be None or 'langchain'")
        path = Path(folder_path)
        path.mkdir(exist_ok=True, parents=True)
        write_index(getattr(self.index, "wrapped", self.index), str(path / f"{index_name}.faiss"))
        with open(path / f"{index_name}.pkl", "wb") as f:
            if compat == "langchain" and not HAVE_LANGCHAIN:
                f.write(_dumps_as_langchain(self.docstore, self.index_to_docstore_id))
            else:  # with langchain installed the docstore already holds its classes
                pickle.dump((self.docstore, self.index_to_docstore_id), f)

    @classmethod
    def load_local(cls, folder_path: str, embeddings: Embeddings, index_name: str = "index", *,
                   allow_dangerous_deserialization: bool = False, **kwargs: Any) -> "FAISS":
        if not allow_dangerous_deserialization:
            raise ValueError(
                "The de-serialization relies loading a pickle file. Pickle files can be modified to deliver a "
                "malicious payload that results in execution of arbitrary code on your machine.You will need to set "
                "`allow_dangerous_deserialization` to `True` to enable deserialization. If you do this, make sure "
                "that you trust the source of the data. For example, if you are loading a file that you created, and "
                "know that no one else has modified the file, then this is safe to do. Do not set this to `True` if "
                "you are loading a file from an untrusted source (e.g., some random site on the internet.).")
        path = Path(folder_path)
        index_kwargs = {key: kwargs.pop(key) for key in ("dtype", "device") if key in kwargs}
        index = read_index(str(path / f"{index_name}.faiss"), **index_kwargs)
        with open(path / f"{index_name}.pkl", "rb") as f:
            docstore, index_to_docstore_id = _CompatUnpickler(f).load()
        return cls(embeddings, index, docstore, index_to_docstore_id, **kwargs)

    def serialize_to_bytes(self) -> bytes:
        return pickle.dumps((getattr(self.index, "wrapped", self.index), self.docstore, self.index_to_docstore_id))

    @classmethod
    def deserialize_from_bytes(cls, serialized: bytes, embeddings: Embeddings, **kwargs: Any) -> "FAISS":
        index, docstore, index_to_docstore_id = pickle.loads(serialized)
        return cls(embeddings, index, docstore, index_to_docstore_id, **kwargs)

    # ------------------------------------------------------------------ retriever + async
    def as_retriever(self, **kwargs: Any) -> "VectorStoreRetriever":
        tags = kwargs.pop("tags", None) or []
        tags.extend([self.__class__.__name__] + ([self.embeddings.__class__.__name__] if self.embeddings else []))
        return VectorStoreRetriever(vectorstore=self, tags=tags, **kwargs)

    async def asimilarity_search_by_vector(self, embedding: List[float], k: int = 4, filter: Optional[Any] = None,
                                           fetch_k: int = 20, **kwargs: Any) -> List[Document]:
        """Called from many coroutines at once by the reference
        (experimental/event-driven-rag-cve-analysis/cyber_dev_day/faiss_vdb_service.py:169-174)."""
        return await _run_in_executor(self.similarity_search_by_vector, embedding, k, filter=filter, fetch_k=fetch_k,
                                      **kwargs)

    async def asimilarity_search_with_score_by_vector(self, embedding: List[float], k: int = 4,
                                                      filter: Optional[Any] = None, fetch_k: int = 20,
                                                      **kwargs: Any) -> List[Tuple[Document, float]]:
        return await _run_in_executor(self.similarity_search_with_score_by_vector, embedding, k, filter=filter,
                                      fetch_k=fetch_k, **kwargs)

    async def asimilarity_search_with_score(self, query: str, k: int = 4, filter: Optional[Any] = None,
                                            fetch_k: int = 20, **kwargs: Any) -> List[Tuple[Document, float]]:
        embedding = await self._aembed_query(query)
        return await self.asimilarity_search_with_score_by_vector(embedding, k, filter=filter, fetch_k=fetch_k,
                                                                  **kwargs)

    async def asimilarity_search(self, query: str, k: int = 4, filter: Optional[Any] = None, fetch_k: int = 20,
                                 **kwargs: Any) -> List[Document]:
        pairs = await self.asimilarity_search_with_score(query, k, filter=filter, fetch_k=fetch_k, **kwargs)
        return [doc for doc, _ in pairs]

    async def asimilarity_search_with_relevance_scores(self, query: str, k: int = 4,
                                                       **kwargs: Any) -> List[Tuple[Document, float]]:
        return await _run_in_executor(self.similarity_search_with_relevance_scores, query, k, **kwargs)

    async def amax_marginal_relevance_search(self, query: str, k: int = 4, fetch_k: int = 20,
                                             lambda_mult: float = 0.5, filter: Optional[Any] = None,
                                             **kwargs: Any) -> List[Document]:
        return await _run_in_executor(self.max_marginal_relevance_search, query, k=k, fetch_k=fetch_k,
                                      lambda_mult=lambda_mult, filter=filter, **kwargs)

    async def asearch(self, query: str, search_type: str, **kwargs: Any) -> List[Document]:
        return await _run_in_executor(self.search, query, search_type, **kwargs)


def _mmr_on_rows(index: _ix.IndexFlat, query: np.ndarray, rows: np.ndarray, k: int, lambda_mult: float) -> List[int]:
    """MMR over an explicit candidate set: a scratch IP index over the candidates, all picks on the device."""
    tmp = _ix.IndexFlat(index.d, _ix.METRIC_INNER_PRODUCT, "float32", index.device)
    tmp.add(rows)
    # fetch_k = every candidate; order of the candidate list is restored through the returned ids
    ids, _ = tmp.mmr(query, k=k, fetch_k=len(rows), lambda_mult=lambda_mult)
    return [int(i) for i in ids[0] if i != -1]


async def _run_in_executor(func: Callable, *args: Any, **kwargs: Any) -> Any:
    return await asyncio.get_running_loop().run_in_executor(None, functools.partial(func, *args, **kwargs))


def _dumps_as_langchain(docstore: Any, index_to_docstore_id: Dict[int, str]) -> bytes:
    """pickle of (docstore, id map) that names LangChain's classes instead of this package's stand-ins.  pickle stores
    classes by reference (module + name) and checks that the name resolves to the very class being pickled, so the
    stand-in modules exist in sys.modules only for the duration of the dump."""
    import sys
    import types

    doc_cls = type("Document", (), {
        "__module__": "langchain_core.documents.base",
        # the state a pydantic-v1 BaseModel pickles: BaseModel.__setstate__ reads exactly these three keys
        "__getstate__": lambda self: {"__dict__": dict(self._fields), "__fields_set__": set(self._fields),
                                      "__private_attribute_values__": {}},
    })
    store_cls = type("InMemoryDocstore", (), {"__module__": "langchain_community.docstore.in_memory"})
    store = store_cls.__new__(store_cls)
    store._dict = {}
    for key, doc in getattr(docstore, "_dict", {}).items():
        d = doc_cls.__new__(doc_cls)
        fields = {"page_content": doc.page_content, "metadata": dict(doc.metadata or {}), "type": "Document"}
        object.__setattr__(d, "_fields", fields)
        store._dict[key] = d
    names = {"langchain_core": None, "langchain_core.documents": None, "langchain_core.documents.base": doc_cls,
             "langchain_community": None, "langchain_community.docstore": None,
             "langchain_community.docstore.in_memory": store_cls}
    saved = {n: sys.modules.get(n) for n in names}
    try:
        for n, cls in names.items():
            mod = types.ModuleType(n)
            if cls is not None:
                setattr(mod, cls.__name__, cls)
            sys.modules[n] = mod
        return pickle.dumps((store, dict(index_to_docstore_id)), protocol=4)
    finally:
        for n, old in saved.items():
            if old is None:
                sys.modules.pop(n, None)
            else:
                sys.modules[n] = old


class _CompatUnpickler(pickle.Unpickler):
    """Loads `index.pkl` files written by LangChain's FAISS.save_local even when langchain is absent."""

    _MAP = {
        ("langchain_community.docstore.in_memory", "InMemoryDocstore"): InMemoryDocstore,
        ("langchain.docstore.in_memory", "InMemoryDocstore"): InMemoryDocstore,
        ("langchain_core.documents.base", "Document"): Document,
        ("langchain.schema.document", "Document"): Document,
        ("langchain.schema", "Document"): Document,
    }

    def find_class(self, module: str, name: str):
        try:
            return super().find_class(module, name)
        except (ModuleNotFoundError, AttributeError):
            if (module, name) in self._MAP:
                return self._MAP[(module, name)]
            raise


class VectorStoreRetriever:
    """langchain_core.vectorstores.VectorStoreRetriever: `search_type` in similarity |
    similarity_score_threshold | mmr; `search_kwargs` forwarded to the store."""

    allowed_search_types = ("similarity", "similarity_score_threshold", "mmr")

    def __init__(self, vectorstore: FAISS, search_type: str = "similarity",
                 search_kwargs: Optional[dict] = None, tags: Optional[List[str]] = None,
                 metadata: Optional[dict] = None, **_: Any):
        if search_type not in self.allowed_search_types:
            raise ValueError(f"search_type of {search_type} not allowed. Valid values are: "
                             f"{self.allowed_search_types}")
        search_kwargs = dict(search_kwargs or {})
        if search_type == "similarity_score_threshold":
            score_threshold = search_kwargs.get("score_threshold")
            if score_threshold is None or not isinstance(score_threshold, float):
                raise ValueError("`score_threshold` is not specified with a float value(0~1) in `search_kwargs`.")
        self.vectorstore = vectorstore
        self.search_type = search_type
        self.search_kwargs = search_kwargs
        self.tags = tags or []
        self.metadata = metadata or {}

    def _get_relevant_documents(self, query: str) -> List[Document]:
        if self.search_type == "similarity":
            return self.vectorstore.similarity_search(query, **self.search_kwargs)
        if self.search_type == "similarity_score_threshold":
            pairs = self.vectorstore.similarity_search_with_relevance_scores(query, **self.search_kwargs)
            return [doc for doc, _ in pairs]
        return self.vectorstore.max_marginal_relevance_search(query, **self.search_kwargs)

    def get_relevant_documents(self, query: str, *, callbacks: Any = None, tags: Any = None, metadata: Any = None,
                               run_name: Any = None, **kwargs: Any) -> List[Document]:
        return self._get_relevant_documents(query)

    def invoke(self, input: str, config: Any = None, **kwargs: Any) -> List[Document]:
        return self._get_relevant_documents(input)

    async def aget_relevant_documents(self, query: str, **kwargs: Any) -> List[Document]:
        return await _run_in_executor(self._get_relevant_documents, query)

    async def ainvoke(self, input: str, config: Any = None, **kwargs: Any) -> List[Document]:
        return await _run_in_executor(self._get_relevant_documents, input)

    def add_documents(self, documents: List[Document], **kwargs: Any) -> List[str]:
        return self.vectorstore.add_documents(documents, **kwargs)
