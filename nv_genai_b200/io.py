"""faiss IndexFlat file format (`index.faiss`) reader/writer.

Replaces `faiss.write_index` / `faiss.read_index` for the two index types of this path, so
that stores saved by the reference's apps load into the B200 engine and back
(`save_local` / `load_local` call sites:
/root/reference/experimental/oran-chatbot-multimodal/vectorstore/vectorstore_updater.py:184,220,295,
/root/reference/experimental/event-driven-rag-cve-analysis/cyber_dev_day/faiss_vdb_service.py:63-66).

Layout (faiss/impl/index_write.cpp, write_index_header + IndexFlat branch):
    fourcc        4 bytes  "IxF2" (L2) | "IxFI" (inner product)
    d             int32
    ntotal        int64
    dummy         int64 x 2   (faiss writes 1 << 20 twice)
    is_trained    uint8
    metric_type   int32       (0 = inner product, 1 = L2)
    size          uint64      number of 4-byte units that follow (= ntotal * d)
    rows          ntotal * d float32, row-major
All little-endian.
"""
from __future__ import annotations

import struct

import numpy as np

from . import index as _ix

_FOURCC = {_ix.METRIC_L2: b"IxF2", _ix.METRIC_INNER_PRODUCT: b"IxFI"}
_CHUNK_ROWS = 1 << 16


def write_index(index: "_ix.IndexFlat", path: str) -> None:
    d, n, metric = index.d, index.ntotal, index.metric_type
    with open(path, "wb") as f:
        f.write(_FOURCC[metric])
        f.write(struct.pack("<iqqqBi", d, n, 1 << 20, 1 << 20, 1, metric))
        f.write(struct.pack("<Q", n * d))
        for r0 in range(0, n, _CHUNK_ROWS):
            rows = index.reconstruct_n(r0, min(_CHUNK_ROWS, n - r0))
            f.write(np.ascontiguousarray(rows, dtype="<f4").tobytes())


def read_index(path: str, dtype: str = "float32", device=None) -> "_ix.IndexFlat":
    with open(path, "rb") as f:
        fourcc = f.read(4)
        if fourcc not in (b"IxF2", b"IxFI", b"IxFl"):
            raise RuntimeError(f"{path}: index type {fourcc!r} is not a flat index (only IxF2 / IxFI are supported)")
        d, n, _d1, _d2, _trained, metric = struct.unpack("<iqqqBi", f.read(4 + 8 + 8 + 8 + 1 + 4))
        if metric not in (_ix.METRIC_L2, _ix.METRIC_INNER_PRODUCT):
            raise RuntimeError(f"{path}: metric_type {metric} is not supported (only L2 and inner product)")
        (size,) = struct.unpack("<Q", f.read(8))
        if size != n * d:
            raise RuntimeError(f"{path}: payload of {size} floats does not match ntotal*d = {n * d}")
        cls = _ix.IndexFlatIP if metric == _ix.METRIC_INNER_PRODUCT else _ix.IndexFlatL2
        index = cls(d, dtype, device)
        for r0 in range(0, n, _CHUNK_ROWS):
            m = min(_CHUNK_ROWS, n - r0)
            buf = f.read(m * d * 4)
            if len(buf) != m * d * 4:
                raise RuntimeError(f"{path}: truncated file")
            index.add(np.frombuffer(buf, dtype="<f4").reshape(m, d))
    return index
