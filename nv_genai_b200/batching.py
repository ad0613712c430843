"""Request coalescing for concurrent single-vector searches.

The reference's heaviest caller of the flat search fires many single-vector searches at once:
`asyncio.gather(*[asimilarity_search_by_vector(...)])` in
/root/reference/experimental/event-driven-rag-cve-analysis/cyber_dev_day/faiss_vdb_service.py:169-174, and every
chain-server request is one `index.search(1 x d, k)` (chains.py:122-123).  On this engine a scan of the corpus costs
the same HBM time for 1 query as for 16 (the kernels are bandwidth-bound), so searches that arrive within a short
window are worth running as ONE batched `b2_index_search`.  `BatchingIndex` wraps an index object, keeps its
faiss-shaped interface, and funnels concurrent `search(x[1, d], k)` calls through one worker thread that groups up
to `max_batch` of them (waiting at most `max_wait_us` for company) into a single call; each caller gets the rows of
its own query cut at its own k (top-k is a prefix of top-k_max under the engine's deterministic tie rule).  A
coalesced batch may be scored by another kernel than a lone query (the tensor-core scan from 5 queries on fp32
rows, 3 on bf16 rows): scores then agree within the engine's 1e-5 relative parity tolerance, and ids inside a run of
scores that tie within that tolerance may swap.
Calls that already carry a batch, and every other method, go straight to the wrapped index.
"""
from __future__ import annotations

import threading
import time
from typing import Any, List, Optional

import numpy as np


class _Slot:
    __slots__ = ("x", "k", "event", "D", "I", "exc")

    def __init__(self, x: np.ndarray, k: int):
        self.x, self.k = x, k
        self.event = threading.Event()
        self.D = self.I = self.exc = None


class BatchingIndex:
    """faiss-shaped proxy that coalesces concurrent single-vector searches into batched ones."""

    def __init__(self, index: Any, max_batch: int = 16, max_wait_us: float = 200.0):
        if max_batch < 1:
            raise ValueError("max_batch must be >= 1")
        self._index = index
        self.max_batch = int(max_batch)
        self.max_wait_s = float(max_wait_us) * 1e-6
        self._cv = threading.Condition()
        self._pending: List[_Slot] = []
        self._closed = False
        self.batches_run = 0   # batched index.search calls issued by the worker
        self.requests_served = 0
        self._thread = threading.Thread(target=self._run, name="b2-batching", daemon=True)
        self._thread.start()

    # ---- the wrapped index keeps its whole interface
    def __getattr__(self, name: str) -> Any:
        return getattr(self._index, name)

    @property
    def wrapped(self) -> Any:
        return self._index

    def close(self) -> None:
        with self._cv:
            self._closed = True
            self._cv.notify_all()
        self._thread.join(timeout=5)

    # ---- index.search
    def search(self, x: Any, k: int):
        if self._closed or not isinstance(x, (np.ndarray, list, tuple)):
            return self._index.search(x, k)  # device tensors: the caller already batches on the device
        a = np.asarray(x, dtype=np.float32)
        if a.ndim == 1:
            a = a.reshape(1, -1)
        if a.ndim != 2 or a.shape[0] != 1:
            return self._index.search(a, k)
        k = int(k)
        assert k > 0, "k must be positive"
        assert a.shape[1] == self._index.d, \
            f"search: dimension {a.shape[1]} does not match the index dimension {self._index.d}"
        slot = _Slot(np.ascontiguousarray(a), k)
        with self._cv:
            closed = self._closed  # read under the lock: after close() nobody serves the queue any more
            if not closed:
                self._pending.append(slot)
                self._cv.notify_all()
        if closed:
            return self._index.search(a, k)
        while not slot.event.wait(1.0):
            if not self._thread.is_alive():  # the worker died without serving this slot: search directly
                with self._cv:
                    if slot in self._pending:
                        self._pending.remove(slot)
                if not slot.event.is_set():
                    return self._index.search(a, k)
        if slot.exc is not None:
            raise slot.exc
        return slot.D, slot.I

    def _take_batch(self) -> Optional[List[_Slot]]:
        with self._cv:
            while not self._pending and not self._closed:
                self._cv.wait()
            if not self._pending:
                return None
            deadline = time.monotonic() + self.max_wait_s
            while len(self._pending) < self.max_batch and not self._closed:
                left = deadline - time.monotonic()
                if left <= 0:
                    break
                self._cv.wait(left)
            batch = self._pending[: self.max_batch]
            del self._pending[: self.max_batch]
            return batch

    def _run(self) -> None:
        while True:
            batch = self._take_batch()
            if batch is None:
                return
            try:
                kmax = max(s.k for s in batch)
                X = np.concatenate([s.x for s in batch], axis=0)
                D, I = self._index.search(X, kmax)
                self.batches_run += 1
                self.requests_served += len(batch)
                for i, s in enumerate(batch):
                    s.D = np.ascontiguousarray(D[i:i + 1, : s.k])
                    s.I = np.ascontiguousarray(I[i:i + 1, : s.k])
            except BaseException as exc:  # every waiter must be released
                for s in batch:
                    s.exc = exc
            for s in batch:
                s.event.set()
