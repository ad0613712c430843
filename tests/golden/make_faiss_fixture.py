#!/usr/bin/env python
"""Writes tests/golden/flat_l2_3x4.faiss and flat_ip_2x3.faiss: two tiny faiss flat-index files assembled byte by byte
from the file layout of faiss 1.7.4 (faiss/impl/index_write.cpp: write_index_header + the IndexFlat branch), NOT
through nv_genai_b200/io.py, so that they pin io.py's reader and writer from outside:

    fourcc "IxF2" | "IxFI", d int32, ntotal int64, dummy int64 = 1 << 20 (twice), is_trained uint8 = 1,
    metric_type int32 (1 = L2, 0 = inner product), size uint64 = ntotal * d (units of 4 bytes), rows float32

The rows are small dyadic numbers so the files are readable in a hex dump."""
import os
import struct

HERE = os.path.dirname(os.path.abspath(__file__))


def flat_file(fourcc, metric, rows):
    d = len(rows[0])
    out = bytearray(fourcc)
    out += struct.pack("<i", d)
    out += struct.pack("<q", len(rows))
    out += struct.pack("<q", 1 << 20) + struct.pack("<q", 1 << 20)
    out += struct.pack("<B", 1)
    out += struct.pack("<i", metric)
    out += struct.pack("<Q", len(rows) * d)
    for r in rows:
        for v in r:
            out += struct.pack("<f", v)
    return bytes(out)


L2_ROWS = [[1.0, 0.0, -0.5, 2.0], [0.25, 0.75, 3.0, -1.0], [0.0, 0.0, 0.0, 8.0]]
IP_ROWS = [[0.5, -0.5, 0.125], [4.0, 1.0, -2.0]]

if __name__ == "__main__":
    with open(os.path.join(HERE, "flat_l2_3x4.faiss"), "wb") as f:
        f.write(flat_file(b"IxF2", 1, L2_ROWS))
    with open(os.path.join(HERE, "flat_ip_2x3.faiss"), "wb") as f:
        f.write(flat_file(b"IxFI", 0, IP_ROWS))
