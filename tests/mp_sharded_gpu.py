#!/usr/bin/env python
"""Multi-process GPU check of the row-sharded search (run under torchrun, one rank per GPU; not collected by pytest).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
        tests/mp_sharded_gpu.py

Every rank ingests its block of the synthetic corpus, rank 0 additionally holds the whole corpus in one index.  For
several (nq, k) the sharded result — through the NVLink peer-memory exchange and through the NCCL all-gather — must
equal the single-index result bit for bit on every rank.  Prints one JSON line on rank 0 and exits non-zero on failure.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist

    from nv_genai_b200 import IndexFlatIP, normalize_L2, synth_fill
    from nv_genai_b200.sharded import ShardedFlatIndex, shard_bounds

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n, d = 300_001, 1024
    lo, hi = shard_bounds(n, world, rank)
    dev = f"cuda:{local}"

    def rows(a, b):
        buf = torch.empty((b - a, d), dtype=torch.float32, device=dev)
        synth_fill(buf, a, 0x5EED0001)
        return buf

    report = {"world": world, "cases": []}
    ok = True
    for dtype in ("float32", "bfloat16"):
        shard = IndexFlatIP(d, dtype=dtype, device=local)
        shard.add(rows(lo, hi), normalize=True)
        whole = IndexFlatIP(d, dtype=dtype, device=local)
        whole.add(rows(0, n), normalize=True)  # every rank keeps a reference copy (300k rows: 1.2 GB)
        q = torch.empty((300, d), dtype=torch.float32, device=dev)
        synth_fill(q, 0, 0x5EED0002)
        normalize_L2(q)
        for use_peer in (True, False):
            if use_peer:
                os.environ.pop("B2_NO_PEER_PUSH", None)
            else:
                os.environ["B2_NO_PEER_PUSH"] = "1"
            sh = ShardedFlatIndex(d, shard.metric_type, local_index=shard)
            sh._refresh_bases()
            assert sh.ntotal == n
            peer_active = sh._peer is not None
            for nq, k in ((1, 10), (16, 10), (3, 100), (40, 100), (300, 4)):
                for rep in range(3):  # several searches in a row: both block sets and the epoch logic are exercised
                    qq = q[rep:rep + nq].contiguous()
                    D, I = sh.search(qq, k, replicated=True)
                    Dw, Iw = whole.search(qq, k)
                    same = bool(torch.equal(I, Iw)) and bool(torch.equal(D, Dw))
                    flag = torch.tensor([1 if same else 0], device=dev)
                    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
                    same_all = int(flag.item()) == 1
                    ok = ok and same_all
                    if rank == 0 and rep == 0:
                        report["cases"].append({"dtype": dtype, "peer": use_peer, "peer_active": peer_active,
                                                "nq": nq, "k": k, "equal": same_all})
            # non-replicated queries: rank 0 owns the batch (through peer memory when the exchange is active, else by
            # broadcast); several in a row so that both buffer sets are used, and into a packed output block
            for rep in range(5):
                nq1 = (5, 1, 16, 2, 64)[rep]
                qq = q[rep:rep + nq1].contiguous() if rank == 0 else torch.zeros((nq1, d), dtype=torch.float32, device=dev)
                packed = torch.empty(nq1 * 10 * 12, dtype=torch.uint8, device=dev) if rep % 2 == 0 else None
                D, I = sh.search(qq, 10, replicated=False, src=0, out=packed)
                Dw, Iw = whole.search(q[rep:rep + nq1].contiguous(), 10)
                same = bool(torch.equal(I, Iw)) and bool(torch.equal(D, Dw))
                if packed is not None:
                    same = same and bool(torch.equal(packed[: nq1 * 80].view(torch.int64).view(nq1, 10), Iw))
                    same = same and bool(torch.equal(packed[nq1 * 80:].view(torch.float32).view(nq1, 10), Dw))
                if rank == 0:
                    report["cases"].append({"dtype": dtype, "peer": use_peer, "non_replicated_nq": nq1, "equal": same})
                ok = ok and same
            if use_peer and not peer_active and rank == 0:
                report["peer_error"] = getattr(sh, "_peer_error", "disabled")
            del sh
        del shard, whole
        torch.cuda.empty_cache()
    torch.cuda.synchronize()
    dist.barrier()
    if rank == 0:
        report["ok"] = ok
        print(json.dumps(report), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
