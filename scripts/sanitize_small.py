#!/usr/bin/env python
"""Small shapes through every kernel, for `compute-sanitizer --tool memcheck python scripts/sanitize_small.py`."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from nv_genai_b200 import IndexFlatIP, IndexFlatL2, normalize_L2, synth_fill

n, d = 9_001, 200
for cls in (IndexFlatIP, IndexFlatL2):
    for dtype in ("float32", "bfloat16"):
        idx = cls(d, dtype=dtype, device=0)
        buf = torch.empty((n, d), dtype=torch.float32, device="cuda:0")
        synth_fill(buf, 0, 1)
        idx.add(buf, normalize=True)
        idx.add(buf[:77].cpu().numpy())          # host path
        q = torch.empty((70, d), dtype=torch.float32, device="cuda:0")
        synth_fill(q, 0, 2)
        normalize_L2(q)
        for nq, k in ((1, 4), (7, 10), (16, 100), (20, 10), (33, 64), (70, 100), (70, 128), (3, 300)):
            D, I = idx.search(q[:nq], k)
            assert int(I.min()) >= 0 and int(I.max()) < n + 77
        ids, sc = idx.mmr(q[:5], k=4, fetch_k=20)
        idx.remove_ids(np.array([0, 5, 9000]))
        D, I = idx.search(q[:2].cpu().numpy(), 5)
        rows = idx.reconstruct_n(0, 10)
        del idx
torch.cuda.synchronize()
print("sanitize_small ok")
