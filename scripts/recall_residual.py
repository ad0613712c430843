#!/usr/bin/env python
"""bf16 rows with a residual plane at full size: recall@k against the fp32 row store (the exact FFMA scan of the fp32
rows, itself pinned against the CPU oracle over all rows by tests/test_gpu_fullsize.py) and time per batch beside the
plain bf16 store's, for the large-batch tensor path (batch 4096, k = 100) and the tensor-core scan (batch 16, k = 10).

    python scripts/recall_residual.py [--rows 10000000] > profiles/r02_recall_residual.json
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--dim", type=int, default=1024)
    a = ap.parse_args()
    import torch

    from nv_genai_b200 import IndexFlatIP, normalize_L2, synth_fill

    stores = {"float32": IndexFlatIP(a.dim, dtype="float32"), "bfloat16": IndexFlatIP(a.dim, dtype="bfloat16"),
              "bfloat16+residual": IndexFlatIP(a.dim, dtype="bfloat16+residual")}
    chunk = min(1_000_000, a.rows)
    buf = torch.empty((chunk, a.dim), dtype=torch.float32, device="cuda:0")
    for r0 in range(0, a.rows, chunk):
        m = min(chunk, a.rows - r0)
        synth_fill(buf[:m], r0, 0x5EED0001)
        for s in stores.values():
            s.add(buf[:m], normalize=True)
    del buf
    q = torch.empty((4096, a.dim), dtype=torch.float32, device="cuda:0")
    synth_fill(q, 0, 0x5EED0002)
    normalize_L2(q)

    def timed(idx, qq, k, reps=3):
        idx.search(qq, k)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = 1e30
        for _ in range(reps):
            e0.record()
            D, I = idx.search(qq, k)
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return D.cpu(), I.cpu(), best

    def recall(I, Iref):
        hits = 0
        for r in range(I.shape[0]):
            hits += len(set(I[r].tolist()) & set(Iref[r].tolist()))
        return hits / I.numel()

    out = {"rows": a.rows, "dim": a.dim, "truth": "exact FFMA scan of the fp32 row store (B2_NO_TENSOR=1), 256 queries"}
    os.environ["B2_NO_TENSOR"] = "1"
    truth = {}
    for k in (100, 10):
        _, I, _ = timed(stores["float32"], q[:256], k, reps=1)
        truth[k] = I
    os.environ.pop("B2_NO_TENSOR")
    for name in ("bfloat16", "bfloat16+residual", "float32"):
        idx = stores[name]
        D, I, ms = timed(idx, q, 100)
        _, I16, ms16 = timed(idx, q[:16], 10, reps=20)
        _, I1, ms1 = timed(idx, q[:1], 10, reps=20)
        out[name] = {"batch4096_k100_ms": ms, "batch4096_k100_qps": 4096 / (ms * 1e-3),
                     "recall@100_first_256_queries": recall(I[:256], truth[100]),
                     "batch16_k10_ms": ms16, "recall@10_batch16": recall(I16, truth[10][:16]),
                     "batch1_k10_ms": ms1, "recall@10_batch1": recall(I1, truth[10][:1])}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
