"""Test doubles: a deterministic Embeddings and an oracle-backed index for CPU-only host-logic tests."""
import hashlib

import numpy as np

from nv_genai_b200.documents import Embeddings
from nv_genai_b200.index import METRIC_INNER_PRODUCT, METRIC_L2, IndexFlat
from oracle import flat


class HashEmbeddings(Embeddings):
    """Text -> fixed pseudo-random unit-variance vector; identical texts give identical vectors."""

    def __init__(self, d=64):
        self.d = d

    def _one(self, text):
        seed = int.from_bytes(hashlib.sha256(text.encode()).digest()[:8], "little")
        return flat.synth(0, 1, self.d, seed)[0]

    def embed_documents(self, texts):
        return [self._one(t).tolist() for t in texts]

    def embed_query(self, text):
        return self._one(text).tolist()


class OracleIndex(IndexFlat):
    """Same interface as IndexFlat, rows in numpy, arithmetic by the oracle.  TESTS ONLY (host logic on CPU)."""

    def __init__(self, d, metric=METRIC_L2):
        self._d, self._metric = d, metric
        self._x = np.empty((0, d), dtype=np.float32)
        self.is_trained = True
        self.device = 0
        self._dtype_name = "float32"

    d = property(lambda self: self._d)
    ntotal = property(lambda self: self._x.shape[0])
    metric_type = property(lambda self: self._metric)

    def __del__(self):
        pass

    def add(self, x, normalize=False):
        x = np.array(x, dtype=np.float32, copy=True).reshape(-1, self._d)
        assert x.shape[1] == self._d
        if normalize:
            flat.normalize_L2(x)
        self._x = np.concatenate([self._x, x])

    def search(self, x, k):
        x = np.ascontiguousarray(np.atleast_2d(np.asarray(x, dtype=np.float32)))
        assert x.shape[1] == self._d and k > 0
        return flat.search(self._x, x, k, self._metric)

    def upload_mask(self, words):
        return np.ascontiguousarray(words, dtype=np.uint32)

    def search_filtered(self, x, k, fetch_k=None, allow=None, score_threshold=None, relevance=0,
                        relevance_threshold=0.0):
        fetch_k = k if fetch_k is None else fetch_k
        D, I = self.search(x, fetch_k)
        bits = None
        if allow is not None:
            bits = np.unpackbits(np.asarray(allow, dtype="<u4").view(np.uint8), bitorder="little").astype(bool)
        return flat.select_hits(D, I, k, self._metric, bits, score_threshold, relevance, relevance_threshold)

    def reconstruct_batch(self, ids):
        return self._x[np.asarray(ids, dtype=np.int64)]

    def reconstruct_n(self, n0=0, ni=-1):
        ni = self.ntotal - n0 if ni == -1 else ni
        return self._x[n0:n0 + ni]

    def remove_ids(self, ids):
        keep = np.ones(self.ntotal, bool)
        ids = np.asarray(ids, dtype=np.int64)
        ids = ids[(ids >= 0) & (ids < self.ntotal)]
        keep[ids] = False
        removed = int((~keep).sum())
        self._x = self._x[keep]
        return removed

    def merge_from(self, other, add_id=0):
        self._x = np.concatenate([self._x, other._x])
        other._x = np.empty((0, self._d), np.float32)

    def reset(self):
        self._x = np.empty((0, self._d), np.float32)

    def mmr(self, x, k=4, fetch_k=20, lambda_mult=0.5):
        x = np.atleast_2d(np.asarray(x, dtype=np.float32))
        ids = np.full((x.shape[0], k), -1, np.int64)
        sc = np.zeros((x.shape[0], k), np.float32)
        for r in range(x.shape[0]):
            i, s = flat.mmr_search(self._x, x[r], k, fetch_k, lambda_mult, self._metric)
            ids[r, :len(i)] = i
            sc[r, :len(s)] = s
        return ids, sc

    def __getstate__(self):
        return {"d": self._d, "metric": self._metric, "x": self._x}

    def __setstate__(self, st):
        OracleIndex.__init__(self, st["d"], st["metric"])
        self._x = st["x"]
