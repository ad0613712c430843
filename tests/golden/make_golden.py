"""Generate golden top-k vectors from the reference's OWN code (run in the build container only).

The reference has no implementation of the flat search (it calls faiss-cpu, which is not
installable here), but it does hold one numpy statement of "inner-product scores + top-k":
`hard_negative_mining` in
/root/reference/experimental/synthetic-data-retriever-customization/retriever_customization.ipynb
(cell 21): `np.matmul(batch_query_embeddings, passage_embeddings.T)` followed by
`np.argpartition(row, -num_negs)[-num_negs:]`.  This script extracts that cell's source
verbatim from the notebook at run time (nothing is copied into this repository), executes
it on seeded synthetic embeddings with no positives and an infinite mining threshold — which
reduces it to exactly "top-k by inner product" — and stores the resulting index SETS
(argpartition leaves the order unspecified) in tests/golden/ip_topk_reference.json.
On the normalised cases the same sets also pin IndexFlatL2 (the reference's default index,
/root/reference/RetrievalAugmentedGeneration/common/utils.py:217): for unit vectors the squared
L2 distance is 2 - 2 q.x, so the k nearest rows are the k rows of largest inner product.

/root/reference does not exist on the GPU box, so only the JSON travels; tests regenerate
the inputs from the recorded seeds with the integer-exact synthetic generator.

    python tests/golden/make_golden.py
"""
import json
import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import flat  # noqa: E402

NB = "/root/reference/experimental/synthetic-data-retriever-customization/retriever_customization.ipynb"

CASES = [
    # name, n passages, n queries, d, k, normalise
    ("u_2000x64_k10", 2000, 8, 64, 10, True),
    ("u_5000x1024_k4", 5000, 4, 1024, 4, True),
    ("raw_3000x96_k25", 3000, 6, 96, 25, False),
    ("u_1500x384_k100", 1500, 5, 384, 100, True),
    # the tensor-core paths' shapes: >= 20 queries (faiss's BLAS routine; here the GEMM kernel), k = 100, and a corpus
    # long enough for the threshold pre-pass of the GEMM kernel (>= 2048 tiles of 256 rows)
    ("u_200000x1024_k100_q24", 200_000, 24, 1024, 100, True),
    ("u_600000x256_k100_q32", 600_000, 32, 256, 100, True),
    ("u_200000x1024_k10_q8", 200_000, 8, 1024, 10, True),
]


def load_reference_fn():
    nb = json.load(open(NB))
    src = None
    for cell in nb["cells"]:
        s = "".join(cell["source"])
        if cell["cell_type"] == "code" and "def hard_negative_mining" in s:
            src = s
            break
    assert src is not None, "reference cell not found"
    ns = {"np": np, "math": math}
    exec(compile(src, NB + ":cell21", "exec"), ns)
    return ns["hard_negative_mining"]


def main():
    fn = load_reference_fn()
    out = {"source": NB + " cell 21 (hard_negative_mining)", "corpus_seed": flat.CORPUS_SEED,
           "query_seed": flat.QUERY_SEED, "cases": []}
    for name, n, nq, d, k, norm in CASES:
        P = flat.synth(0, n, d, flat.CORPUS_SEED)
        Q = flat.synth(0, nq, d, flat.QUERY_SEED)
        if norm:
            flat.normalize_L2(P)
            flat.normalize_L2(Q)
        idx = fn(Q, P, batch_size=3, margin=0.95, num_negs=k, query_positive_paragraph_idxs=[[] for _ in range(nq)])
        sets = [sorted(int(i) for i in row) for row in idx]
        out["cases"].append({"name": name, "n": n, "nq": nq, "d": d, "k": k, "normalize": norm, "topk_sets": sets})
        print(name, "ok", sets[0][:5])
    with open(os.path.join(HERE, "ip_topk_reference.json"), "w") as f:
        json.dump(out, f)
    print("wrote ip_topk_reference.json")


if __name__ == "__main__":
    main()
