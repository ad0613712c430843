#!/usr/bin/env python
"""Developer timing of the tensor-core search (BASELINE configs 3/4 shapes) on one GPU.

    python scripts/time_gemm.py [--rows 10000000] [--nq 4096] [--k 100] [--dtype float32|bfloat16] [--reps 3]
Prints one JSON line: ms per batch, queries/s, useful TFLOP/s (2 nq N d) and MMA TFLOP/s (x products).
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
    ap.add_argument("--nq", type=int, default=4096)
    ap.add_argument("--k", type=int, default=100)
    ap.add_argument("--dtype", default="float32")
    ap.add_argument("--metric", default="ip")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--sweep", default="", help="comma list of k:swizzle:parts, e.g. 10:64:2,10:128:2 (env overrides per run)")
    a = ap.parse_args()
    import torch

    from nv_genai_b200 import IndexFlatIP, IndexFlatL2, _native, normalize_L2, synth_fill

    dev = 0
    idx = (IndexFlatIP if a.metric == "ip" else IndexFlatL2)(a.dim, dtype=a.dtype, device=dev)
    chunk = min(1_000_000, a.rows)
    buf = torch.empty((chunk, a.dim), dtype=torch.float32, device="cuda:0")
    for r0 in range(0, a.rows, chunk):
        m = min(chunk, a.rows - r0)
        synth_fill(buf[:m], r0, 0x5EED0001)
        idx.add(buf[:m], normalize=True)
    del buf
    q = torch.empty((a.nq, a.dim), dtype=torch.float32, device="cuda:0")
    synth_fill(q, 0, 0x5EED0002)
    normalize_L2(q)
    torch.cuda.synchronize()
    combos = [c.split(":") for c in a.sweep.split(",") if c] or [[str(a.k), "", ""]]
    for kk, sw, parts in combos:
        a.k = int(kk)
        extra = {}
        if ":" in parts:  # parts may carry more overrides: parts:cg:mixed
            parts, *rest = parts.split(":")
        for name, val in (("B2_GEMM_SWIZZLE", sw), ("B2_GEMM_QUERY_PARTS", parts)):
            if val:
                os.environ[name] = val
            else:
                os.environ.pop(name, None)
        run_one(a, idx, q, torch)


def run_one(a, idx, q, torch):
    D, I = idx.search(q, a.k)  # warm-up
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    times = []
    for _ in range(a.reps):
        e0.record()
        D, I = idx.search(q, a.k)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
    ms = min(times)
    # tensor-core products issued per score: 1 under filter-and-refine (the default), else query parts (+1 on fp32 rows)
    refine = os.environ.get("B2_GEMM_REFINE", "1") != "0" and "B2_GEMM_QUERY_PARTS" not in os.environ
    products = 1 if refine else int(os.environ.get("B2_GEMM_QUERY_PARTS", "2")) + (1 if a.dtype == "float32" else 0)
    nq_pad = (a.nq + 127) // 128 * 128
    useful = 2.0 * a.nq * a.rows * a.dim / (ms * 1e-3) / 1e12
    mma = 2.0 * nq_pad * a.rows * a.dim * products / (ms * 1e-3) / 1e12
    # cross-check a few queries against the scan path
    s = [0, a.nq // 2, a.nq - 1]
    Ds, Is = idx.search(q[s], a.k)
    os.environ["B2_NO_TENSOR"] = "1"
    Ds, Is = idx.search(q[s], a.k)
    os.environ.pop("B2_NO_TENSOR")
    same = bool((Is == I[s]).all().item())
    maxrel = float(((Ds - D[s]).abs() / Ds.abs().clamp_min(1e-6)).max().item())
    print(json.dumps({"rows": a.rows, "dim": a.dim, "nq": a.nq, "k": a.k, "dtype": a.dtype, "metric": a.metric,
                      "ms": ms, "all_ms": times, "qps": a.nq / (ms * 1e-3), "useful_tflops": useful,
                      "mma_tflops": mma, "products": products, "refine": refine, "swizzle": os.environ.get("B2_GEMM_SWIZZLE", "auto"), "mixed": os.environ.get("B2_GEMM_MIXED", "default"), "cta_group": os.environ.get("B2_GEMM_CTA_GROUP", "default"), "ids_equal_scan_sample": same,
                      "max_rel_score_diff_vs_scan": maxrel}), flush=True)


if __name__ == "__main__":
    main()
