#!/usr/bin/env python
"""Developer check at full size: recall@k of the bf16 row store against the exact fp32 scan, and score/id agreement of
the large-batch tensor path with the FFMA scan of the same rows, in its default mode (filter and refine; fp32 rows with
and without the bf16 shadow) and with the all-exact tensor arithmetic."""
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
    ap.add_argument("--nq", type=int, default=256)
    ap.add_argument("--k", type=int, default=100)
    a = ap.parse_args()
    import torch

    from nv_genai_b200 import IndexFlatIP, normalize_L2, synth_fill

    f32 = IndexFlatIP(a.dim, dtype="float32", device=0)
    b16 = IndexFlatIP(a.dim, dtype="bfloat16", device=0)
    chunk = min(1_000_000, a.rows)
    buf = torch.empty((chunk, a.dim), dtype=torch.float32, device="cuda:0")
    for r0 in range(0, a.rows, chunk):
        m = min(chunk, a.rows - r0)
        synth_fill(buf[:m], r0, 0x5EED0001)
        f32.add(buf[:m], normalize=True)
        b16.add(buf[:m], normalize=True)
    del buf
    q = torch.empty((a.nq, a.dim), dtype=torch.float32, device="cuda:0")
    synth_fill(q, 0, 0x5EED0002)
    normalize_L2(q)

    def run(idx, **env):
        for k_, v in env.items():
            os.environ[k_] = v
        D, I = idx.search(q, a.k)
        torch.cuda.synchronize()
        for k_ in env:
            os.environ.pop(k_)
        return D.cpu(), I.cpu()

    def recall(I, Iref, kk):
        hits = 0
        for r in range(I.shape[0]):
            hits += len(set(I[r, :kk].tolist()) & set(Iref[r, :kk].tolist()))
        return hits / (I.shape[0] * kk)

    Dx, Ix = run(f32, B2_NO_TENSOR="1")
    out = {"rows": a.rows, "nq": a.nq, "k": a.k}
    cases = [("f32_refine_bf16_shadow (default)", f32, {}), ("f32_refine_tf32_no_shadow", f32, {"_shadow": "0"}),
             ("f32_all_exact_tf32_plus_bf16_correction", f32, {"B2_GEMM_REFINE": "0"}),
             ("bf16_scan", b16, {"B2_NO_TENSOR": "1"}), ("bf16_refine (default)", b16, {}),
             ("bf16_all_exact_2parts", b16, {"B2_GEMM_REFINE": "0"}), ("bf16_tensor_1part_no_rescoring", b16, {"B2_GEMM_QUERY_PARTS": "1"})]
    for name, idx, env in cases:
        env = dict(env)
        if env.pop("_shadow", None) == "0":
            idx.set_option("bf16_shadow", 0)
        D, I = run(idx, **env)
        out[name] = {"recall@k": recall(I, Ix, a.k), "recall@10": recall(I, Ix, 10),
                     "ids_equal_frac": float((I == Ix).float().mean()),
                     "max_rel_score_diff": float(((D - Dx).abs() / Dx.abs().clamp_min(1e-9)).max()),
                     "mean_rel_score_diff_signed": float(((D - Dx) / Dx.abs().clamp_min(1e-9)).mean())}
        if idx is f32:
            idx.set_option("bf16_shadow", 1)
    Db, Ib = run(b16, B2_NO_TENSOR="1")
    D2, I2 = run(b16)
    out["bf16_refine_vs_bf16_scan"] = {"ids_equal_frac": float((I2 == Ib).float().mean()),
                                              "max_rel_score_diff": float(((D2 - Db).abs() / Db.abs().clamp_min(1e-9)).max())}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
