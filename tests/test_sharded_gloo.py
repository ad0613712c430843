"""CPU, world_size 2, gloo: the host side of the row-sharded search (bounds, bases, the single
all-gather, the merge contract).  The local index is the oracle-backed fake and the merge is a numpy
restatement; on GPUs the same class runs over NCCL with the CUDA index and b2_merge_shard_lists."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nv_genai_b200.sharded import shard_bounds


def test_shard_bounds_partition():
    for n in (0, 1, 7, 10, 100, 10_000_001):
        for w in (1, 2, 3, 4, 8):
            spans = [shard_bounds(n, w, r) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def numpy_merge(metric, D_parts, I_parts):
    """Same contract as b2_merge_shard_lists: order by (key desc, global id asc), pad with -1."""
    Dp, Ip = D_parts.numpy(), I_parts.numpy()
    G, nq, k = Dp.shape
    D = np.empty((nq, k), np.float32)
    I = np.empty((nq, k), np.int64)
    fmax = np.finfo(np.float32).max
    for q in range(nq):
        d = Dp[:, q, :].reshape(-1)
        i = Ip[:, q, :].reshape(-1)
        ok = i >= 0
        key = d[ok].astype(np.float64) * (1 if metric == 0 else -1)
        order = np.lexsort((i[ok], -key))[:k]
        m = len(order)
        D[q] = -fmax if metric == 0 else fmax
        I[q] = -1
        D[q, :m] = d[ok][order]
        I[q, :m] = i[ok][order]
    return torch.from_numpy(D), torch.from_numpy(I)


def _worker(rank, world, port, n, d, nq, k, metric, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import sys

        sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
        from fakes import OracleIndex
        from nv_genai_b200.sharded import ShardedFlatIndex
        from oracle import flat

        x = flat.synth(0, n, d, 77)
        x[n // 2 + 3] = x[2]  # a duplicate that straddles the two shards: tie must resolve to the lower id
        q = flat.synth(0, nq, d, 78)
        q[0] = x[2]
        sh = ShardedFlatIndex(d, metric, local_index=OracleIndex(d, metric), merge_fn=numpy_merge)
        sh.add_global(x)
        assert sh.ntotal == n and sh.bases[rank] == shard_bounds(n, world, rank)[0]
        qt = torch.from_numpy(q.copy()) if rank == 0 else torch.zeros(nq, d)
        D, I = sh.search(qt, k, replicated=False, src=0)  # rank 0 broadcasts the queries
        Dr, Ir = flat.search(x, q, k, metric)
        assert np.array_equal(I.numpy(), Ir), (rank, I.numpy(), Ir)
        assert np.array_equal(D.numpy(), Dr)
        # incremental growth keeps ids in rank order
        sh2 = ShardedFlatIndex(d, metric, local_index=OracleIndex(d, metric), merge_fn=numpy_merge)
        lo, hi = shard_bounds(n, world, rank)
        sh2.add_local(x[lo:hi])
        D2, I2 = sh2.search(torch.from_numpy(q), k)
        assert np.array_equal(I2.numpy(), Ir)
        # packed output block [I | D]: what a caller copies back with one transfer
        packed = torch.empty(nq * k * 12, dtype=torch.uint8)
        sh2.search(torch.from_numpy(q), k, out=packed)
        assert np.array_equal(packed[: nq * k * 8].view(torch.int64).view(nq, k).numpy(), Ir)
        assert np.array_equal(packed[nq * k * 8:].view(torch.float32).view(nq, k).numpy(), Dr)
        # k larger than a shard
        small = ShardedFlatIndex(d, metric, local_index=OracleIndex(d, metric), merge_fn=numpy_merge)
        small.add_global(x[:5])
        D3, I3 = small.search(torch.from_numpy(q), 4)
        D3r, I3r = flat.search(x[:5], q, 4, metric)
        assert np.array_equal(I3.numpy(), I3r)
        out.put((rank, "ok"))
    except Exception as e:  # surface the failure in the parent
        import traceback

        out.put((rank, "FAIL " + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("metric", [0, 1])
def test_sharded_search_world2_gloo_equals_single(metric):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 1001, 48, 3, 7, metric, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = [out.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(60)
    assert all(r[1] == "ok" for r in res), res
