#!/usr/bin/env python
"""Developer timing of the ingest kernel (device rows -> row store, fused normalise): rows/s and GB/s."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from nv_genai_b200 import IndexFlatIP, synth_fill

rows, d = 1_000_000, 1024
buf = torch.empty((rows, d), dtype=torch.float32, device="cuda:0")
synth_fill(buf, 0, 0x5EED0001)
for dtype in ("float32", "bfloat16"):
    idx = IndexFlatIP(d, dtype=dtype, device=0)
    for _ in range(10):  # grow the row store to its final size first (VMM mapping is host work, not the kernel)
        idx.add(buf, normalize=True)
    idx.reset()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        idx.add(buf, normalize=True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    b = rows * d * (4 + (4 if dtype == "float32" else 2))
    print(json.dumps({"dtype": dtype, "ms_per_1M_rows": ms, "GBps": b / ms / 1e6}))
    del idx
