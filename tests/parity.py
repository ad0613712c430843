"""Parity rule between the CUDA path and the oracle.

fp32: scores within RTOL relative; ids equal position for position except inside runs where
neighbouring oracle scores differ by less than RTOL relative (reorderings inside such a run,
and swaps across the k-th boundary with a row whose true score ties the k-th, are accepted).
"""
import numpy as np

RTOL = 1e-5
ATOL = 1e-6  # squared L2 of planted exact duplicates is ~0: relative tolerance alone is meaningless there


def true_scores(x, q, ids, metric_ip):
    r = np.asarray(x[ids], dtype=np.float64)
    qq = np.asarray(q, dtype=np.float64)
    return r @ qq if metric_ip else ((r - qq[None, :]) ** 2).sum(-1)


def assert_parity(D_gpu, I_gpu, D_ref, I_ref, x=None, q=None, metric_ip=True, rtol=RTOL, atol=ATOL):
    D_gpu, I_gpu, D_ref, I_ref = map(np.asarray, (D_gpu, I_gpu, D_ref, I_ref))
    assert D_gpu.shape == D_ref.shape and I_gpu.shape == I_ref.shape
    nq, k = I_ref.shape
    for r in range(nq):
        # padding must agree exactly
        np.testing.assert_array_equal(I_gpu[r] == -1, I_ref[r] == -1, err_msg=f"query {r}: -1 padding differs")
        valid = I_ref[r] != -1
        np.testing.assert_array_equal(D_gpu[r][~valid], D_ref[r][~valid])
        dg, dr = D_gpu[r][valid].astype(np.float64), D_ref[r][valid].astype(np.float64)
        tol = atol + rtol * np.abs(dr)
        assert np.all(np.abs(dg - dr) <= tol), f"query {r}: scores differ: {dg} vs {dr}"
        ig, ir = I_gpu[r][valid], I_ref[r][valid]
        for j in np.nonzero(ig != ir)[0]:
            # the gpu's id must sit at a tie-equivalent position in the oracle list ...
            pos = np.nonzero(ir == ig[j])[0]
            if len(pos):
                assert abs(dr[pos[0]] - dr[j]) <= atol + rtol * abs(dr[j]), \
                    f"query {r} pos {j}: id {ig[j]} is at oracle pos {pos[0]} with a non-tied score"
            else:
                # ... or be an excluded row whose true score ties the boundary
                assert x is not None, f"query {r} pos {j}: id {ig[j]} not in the oracle list"
                ts = true_scores(x, q[r], np.array([ig[j]]), metric_ip)[0]
                assert abs(ts - dr[-1]) <= atol + 10 * rtol * abs(dr[-1]), \
                    f"query {r} pos {j}: id {ig[j]} (score {ts}) does not tie the k-th oracle score {dr[-1]}"
        assert len(set(ig.tolist())) == len(ig), f"query {r}: duplicate ids returned"


def recall_at_k(I_gpu, I_ref):
    hits = tot = 0
    for a, b in zip(np.asarray(I_gpu), np.asarray(I_ref)):
        b = set(int(v) for v in b if v != -1)
        hits += len(b.intersection(int(v) for v in a))
        tot += len(b)
    return hits / max(tot, 1)
