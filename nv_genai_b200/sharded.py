"""Row-sharded exact search across GPUs: one process per GPU, torch.distributed for plumbing.

The reference has no multi-GPU code; this is the partitioning the path
admits: rank g of G holds one contiguous block of rows, queries are made
visible on every rank, each rank runs the single-GPU scan, the per-shard sorted
`(score, id)` lists are exchanged with ONE all-gather (NCCL over NVLink on GPUs) and merged
on the device by `b2_merge_shard_lists`.  Top-k is a decomposable reduction, so the result
equals the single-GPU result bit for bit: the same kernels score each row, and the merge
orders equal scores lowest-global-id first because shards are in id order.

The local index and the merge are injectable so that the host logic (bounds, bases,
exchange layout) is testable with the gloo backend on CPU; the defaults are the CUDA ones
and there is no CPU fallback in the product path.
"""
from __future__ import annotations

import ctypes
import os
from typing import Any, Callable, Optional, Tuple


def shard_bounds(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous block partition: rank r holds rows [lo, hi); blocks differ by at most one row."""
    base, rem = divmod(int(n), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class _PeerExchange:
    """NVLink peer-memory exchange for the sharded search (b2_index_search_push / b2_merge_shard_lists_wait):
    one symmetric allocation per rank = 2 alternating sets of `world` receive blocks + per-query flag words,
    mapped into every rank's address space by torch symmetric memory (torch is plumbing: it only provides the
    peer mappings).  The merge kernel that produces a rank's final lists stores them into every rank's block and
    raises the flags; the receiving merge kernel waits on its own flags.  No collective on the data path."""

    MAX_RESULTS = 1 << 17  # nq * k per search (12 B each -> 1.5 MB blocks)
    MAX_NQ = 4096
    MAX_Q_BYTES = 4 * 1024 * 1024  # query blocks up to this size travel through peer memory (B2_QUERY_PUSH_MAX_BYTES)

    def __init__(self, torch, dist, group, device, world: int, rank: int):
        import torch.distributed._symmetric_memory as symm

        self.world, self.rank = world, rank
        self.blk = self.MAX_RESULTS * 12
        self.slot_bytes = world * self.blk
        self.flag_off = 2 * self.slot_bytes
        # after the list flags: two query receive buffers and their two flag words (one cache line each)
        self.q_off = self.flag_off + 2 * world * self.MAX_NQ * 4
        self.qflag_off = self.q_off + 2 * self.MAX_Q_BYTES
        total = self.qflag_off + 2 * 128
        self.buf = symm.empty(total, dtype=torch.uint8, device=device)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        self.hdl = symm.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        if len(self.ptrs) != world or any(p == 0 for p in self.ptrs):
            raise RuntimeError("symmetric memory rendezvous returned no peer pointers")
        self.epoch = 0

    def fits(self, nq: int, k: int) -> bool:
        return nq <= self.MAX_NQ and nq * k <= self.MAX_RESULTS

    def query_slot(self, epoch: int):
        """(receive-buffer pointers on every rank, flag pointers on every rank, local view of this rank's buffer,
        local flag pointer) of the query block of search `epoch`."""
        slot = epoch & 1
        bufs = [p + self.q_off + slot * self.MAX_Q_BYTES for p in self.ptrs]
        flags = [p + self.qflag_off + slot * 128 for p in self.ptrs]
        lo = self.q_off + slot * self.MAX_Q_BYTES
        return bufs, flags, self.buf[lo: lo + self.MAX_Q_BYTES], self.ptrs[self.rank] + self.qflag_off + slot * 128

    def next_slot(self):
        """(epoch, push block pointers, push flag pointers, local blocks, local flags) of the next search."""
        self.epoch = (self.epoch + 1) & 0xFFFFFFFF or 1
        e, w, r = self.epoch, self.world, self.rank
        slot = e & 1
        blocks = [self.ptrs[g] + slot * self.slot_bytes + r * self.blk for g in range(w)]
        flags = [self.ptrs[g] + self.flag_off + (slot * w + r) * self.MAX_NQ * 4 for g in range(w)]
        local_blocks = self.ptrs[r] + slot * self.slot_bytes
        local_flags = self.ptrs[r] + self.flag_off + slot * w * self.MAX_NQ * 4
        return e, blocks, flags, local_blocks, local_flags


class ShardedFlatIndex:
    """IndexFlat sharded by rows over the ranks of a torch.distributed process group."""

    def __init__(self, d: int, metric: int, dtype: str = "float32", group: Any = None, device: Any = None,
                 local_index: Any = None, merge_fn: Optional[Callable] = None):
        import torch
        import torch.distributed as dist

        self._torch, self._dist = torch, dist
        if not dist.is_initialized():
            raise RuntimeError("ShardedFlatIndex needs an initialised torch.distributed process group")
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.d, self.metric = int(d), int(metric)
        if local_index is None:
            from .index import IndexFlat

            local_index = IndexFlat(d, metric, dtype, device)
        self.local = local_index
        self._fast = merge_fn is None and hasattr(local_index, "search_into")  # CUDA path: packed exchange
        if merge_fn is None:
            from .index import merge_shard_lists

            merge_fn = merge_shard_lists
        self._merge = merge_fn
        self._ws = {}  # (nq, k) -> (packed send block, gathered blocks)
        self._peer = None
        if self._fast and dist.get_backend(group) == "nccl" and not os.environ.get("B2_NO_PEER_PUSH"):
            ok = 1
            try:
                dev = torch.device("cuda", getattr(local_index, "device", torch.cuda.current_device()))
                self._peer = _PeerExchange(torch, dist, group, dev, self.world, self.rank)
            except Exception as exc:  # no P2P / symmetric memory on this system: NCCL all-gather path
                self._peer, ok = None, 0
                self._peer_error = repr(exc)
            flag = torch.tensor([ok], dtype=torch.int32, device=self._comm_device())
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)  # all ranks take the same path
            if int(flag.item()) == 0:
                self._peer = None
        self.bases = [0] * (self.world + 1)  # bases[g] = global id of shard g's first row

    # ------------------------------------------------------------------ build
    def _comm_device(self):
        t = self._torch
        if self._dist.get_backend(self.group) == "nccl":
            return t.device("cuda", getattr(self.local, "device", t.cuda.current_device()))
        return t.device("cpu")

    def _refresh_bases(self) -> None:
        t, dist = self._torch, self._dist
        mine = t.tensor([int(self.local.ntotal)], dtype=t.int64, device=self._comm_device())
        counts = [t.zeros_like(mine) for _ in range(self.world)]
        dist.all_gather(counts, mine, group=self.group)
        acc = 0
        self.bases = [0]
        for c in counts:
            acc += int(c.item())
            self.bases.append(acc)

    def add_local(self, x, normalize: bool = False) -> None:
        """Append this rank's own rows to its shard; global ids follow rank order.  Collective."""
        self.local.add(x, normalize=normalize)
        self._refresh_bases()

    def add_global(self, x, normalize: bool = False) -> None:
        """Every rank passes the SAME [n, d] block; each keeps its contiguous slice.  Collective.

        Only valid while the shards are still empty or when called in the same order on all ranks for a store
        that is rebuilt; rows of one call are spread over all ranks, so ids of a later call do not continue the
        earlier ones contiguously (use add_local for incremental growth)."""
        if any(b != 0 for b in self.bases):
            raise RuntimeError("add_global on a non-empty sharded index would break the id order; use add_local")
        lo, hi = shard_bounds(x.shape[0], self.world, self.rank)
        self.local.add(x[lo:hi], normalize=normalize)
        self._refresh_bases()

    @property
    def ntotal(self) -> int:
        return self.bases[-1]

    # ------------------------------------------------------------------ search
    def search(self, q, k: int, replicated: bool = True, src: int = 0, out=None):
        """Global top-k on every rank.  q: [nq, d] tensor on the communication device; if not `replicated`, rank `src`
        owns the batch and the other ranks' `q` is only a receive buffer of the same shape: small batches travel through
        NVLink peer memory (one store kernel on `src`, a flag wait on every rank, no collective), large ones by
        broadcast.  `out` (CUDA path): a uint8 tensor of nq * k * 12 bytes that receives the packed result
        [I: nq x k int64 | D: nq x k fp32], so a caller can bring both back with ONE device-to-host copy.
        Returns (D [nq, k] fp32, I [nq, k] int64 global ids) (views into `out` when given)."""
        t, dist = self._torch, self._dist
        pushed = False
        if (not replicated and self._peer is not None and self._fast and isinstance(q, t.Tensor) and q.is_cuda
                and q.dim() == 2 and q.dtype == t.float32 and q.is_contiguous()
                and q.numel() * 4 <= self._peer.MAX_Q_BYTES and (q.numel() * 4) % 16 == 0
                and self._peer.fits(q.shape[0], k)):
            # the query block goes where the lists go: into every rank's symmetric block, flag-released
            from . import _native as N
            from .index import _stream_for

            e = (self._peer.epoch + 1) & 0xFFFFFFFF or 1  # the epoch next_slot() hands out below
            bufs, flags, local, lflag = self._peer.query_slot(e)
            w, dev = self.world, q.device.index
            stream = _stream_for(dev)
            nbytes = q.numel() * 4
            if self.rank == src:
                N.check(N.lib.b2_query_push(q.data_ptr(), nbytes, w, (ctypes.c_void_p * w)(*bufs),
                                            (ctypes.c_void_p * w)(*flags), e, dev, stream))
            N.check(N.lib.b2_flag_wait(lflag, e, dev, stream))
            q = local[:nbytes].view(t.float32).view(q.shape)
            pushed = True
        if not replicated and not pushed:
            dist.broadcast(q, src=src, group=self.group)
        if self._fast and isinstance(q, t.Tensor) and q.is_cuda:
            # CUDA path: the local search writes (I, D) straight into the block that is all-gathered, ids stay
            # local and the merge kernel adds the shard bases: 4 launches + 1 collective per search, no torch glue
            from .index import merge_shard_lists_packed

            nq = q.shape[0] if q.dim() == 2 else 1
            nres = nq * k
            if self._peer is not None and self._peer.fits(nq, k):
                # peer-memory exchange: the merge kernel stores this rank's lists into every rank's block over
                # NVLink and raises per-query flags; the receiving merge waits on its flags.  3 launches, no collective.
                from . import _native as N
                from .index import _stream_for

                e, blocks, flags, lblocks, lflags = self._peer.next_slot()
                w = self.world
                dev = q.device.index
                stream = _stream_for(dev)
                assert q.dim() == 2 and q.shape[1] == self.d, "query dimension does not match the index"
                qc = q if (q.dtype == t.float32 and q.is_contiguous()) else q.to(t.float32).contiguous()
                N.check(N.lib.b2_index_search_push(self.local._h, nq, qc.data_ptr(), k, w, (ctypes.c_void_p * w)(*blocks),
                                                   (ctypes.c_void_p * w)(*flags), e, stream))
                if out is not None:
                    assert out.is_cuda and out.dtype == t.uint8 and out.is_contiguous() and out.numel() == nres * 12
                    I = out[: nres * 8].view(t.int64).view(nq, k)
                    D = out[nres * 8:].view(t.float32).view(nq, k)
                else:
                    D = t.empty((nq, k), dtype=t.float32, device=q.device)
                    I = t.empty((nq, k), dtype=t.int64, device=q.device)
                bases = (ctypes.c_int64 * w)(*[int(b) for b in self.bases[:w]])
                N.check(N.lib.b2_merge_shard_lists_wait(self.metric, nq, k, w, lblocks, self._peer.blk, lflags,
                                                        self._peer.MAX_NQ, e, bases, D.data_ptr(), I.data_ptr(), dev,
                                                        stream))
                return D, I
            ws = self._ws.get((nq, k))
            if ws is None:
                ws = (t.empty(nres * 12, dtype=t.uint8, device=q.device),
                      t.empty((self.world, nres * 12), dtype=t.uint8, device=q.device))
                if len(self._ws) > 16:
                    self._ws.clear()
                self._ws[(nq, k)] = ws
            packed, gathered = ws
            I_loc = packed[: nres * 8].view(t.int64).view(nq, k)
            D_loc = packed[nres * 8:].view(t.float32).view(nq, k)
            self.local.search_into(q, k, D_loc, I_loc)
            dist.all_gather_into_tensor(gathered, packed, group=self.group)
            D, I = merge_shard_lists_packed(self.metric, gathered, nq, k, self.bases)
            if out is not None:
                out[: nres * 8].view(t.int64).view(nq, k).copy_(I)
                out[nres * 8:].view(t.float32).view(nq, k).copy_(D)
            return D, I
        D_loc, I_loc = self.local.search(q, k)
        if not isinstance(D_loc, t.Tensor):
            D_loc, I_loc = t.from_numpy(D_loc), t.from_numpy(I_loc)
        base = self.bases[self.rank]
        I_glob = t.where(I_loc >= 0, I_loc + base, I_loc)
        nq = D_loc.shape[0]
        # one exchange: [ I (int64) | D (fp32) ] packed as bytes, 12 B per result
        packed = t.cat([I_glob.contiguous().view(t.uint8).reshape(-1), D_loc.contiguous().view(t.uint8).reshape(-1)])
        flat_out = t.empty(self.world * packed.numel(), dtype=t.uint8, device=packed.device)
        dist.all_gather_into_tensor(flat_out, packed, group=self.group)
        gathered = flat_out.view(self.world, packed.numel())
        nb_i = nq * k * 8
        I_parts = gathered[:, :nb_i].contiguous().view(t.int64).reshape(self.world, nq, k)
        D_parts = gathered[:, nb_i:].contiguous().view(t.float32).reshape(self.world, nq, k)
        D, I = self._merge(self.metric, D_parts, I_parts)
        if out is not None:  # packed [I | D], as on the CUDA path
            nres = nq * k
            out[: nres * 8].view(t.int64).view(nq, k).copy_(I)
            out[nres * 8:].view(t.float32).view(nq, k).copy_(D)
        return D, I


class ShardGroup:
    """A flat index sharded by rows over several GPUs of THIS process: a thin caller of the C ABI's shard group
    (b2_shard_group_create / b2_shard_search in include/b2_flat.h).  Same faiss-shaped surface as
    IndexFlat (`d`, `ntotal`, `add`, `search -> (D, I)`, `remove_ids`, `reset`); `devices[0]` is the root device, where
    CUDA-tensor inputs and outputs live.  `rows_per_shard` rows go to one shard before the next one receives any."""

    def __init__(self, d: int, metric: int, devices, rows_per_shard: int, dtype: str = "float32"):
        from . import _native as N
        from .index import _DTYPES

        self._N = N
        self.devices = [int(v) for v in devices]
        self._d, self.metric_type = int(d), int(metric)
        self._h = ctypes.c_void_p()
        arr = (ctypes.c_int * len(self.devices))(*self.devices)
        N.check(N.lib.b2_shard_group_create(len(self.devices), arr, self._d, self.metric_type, _DTYPES[dtype],
                                            int(rows_per_shard), ctypes.byref(self._h)))
        if dtype.endswith("+residual"):
            N.check(N.lib.b2_shard_group_set_option(self._h, b"bf16_residual", 1))
        self.is_trained = True

    def set_option(self, name: str, value: int) -> None:
        self._N.check(self._N.lib.b2_shard_group_set_option(self._h, name.encode(), int(value)))

    d = property(lambda self: self._d)
    ntotal = property(lambda self: int(self._N.lib.b2_shard_group_ntotal(self._h)))

    def shard_ntotal(self, shard: int) -> int:
        return int(self._N.lib.b2_shard_group_shard_ntotal(self._h, int(shard)))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                self._N.lib.b2_shard_group_destroy(h)
            except Exception:
                pass
            h.value = None

    def add(self, x, normalize: bool = False) -> None:
        from .index import _Buf

        b = _Buf(x, self._d, self.devices[0], "add")
        N = self._N
        N.check(N.lib.b2_shard_group_add(self._h, b.n, b.ptr, b.where, N.ADD_NORMALIZE if normalize else 0))

    def search(self, x, k: int):
        import numpy as np

        from .index import _Buf, _stream_for, _torch

        b = _Buf(x, self._d, self.devices[0], "search")
        k = int(k)
        assert k > 0, "k must be positive"
        N = self._N
        stream = _stream_for(self.devices[0])
        if b.is_tensor:
            t = _torch()
            D = t.empty((b.n, k), dtype=t.float32, device=b.obj.device)
            I = t.empty((b.n, k), dtype=t.int64, device=b.obj.device)
            N.check(N.lib.b2_shard_search(self._h, b.n, b.ptr, N.DEVICE, k, D.data_ptr(), I.data_ptr(), N.DEVICE, stream))
            return D, I
        D = np.empty((b.n, k), dtype=np.float32)
        I = np.empty((b.n, k), dtype=np.int64)
        N.check(N.lib.b2_shard_search(self._h, b.n, b.ptr, b.where, k, D.ctypes.data, I.ctypes.data, N.HOST, stream))
        return D, I

    def remove_ids(self, ids) -> int:
        import numpy as np

        a = np.ascontiguousarray(np.asarray(ids, dtype=np.int64).reshape(-1))
        removed = ctypes.c_int64(0)
        self._N.check(self._N.lib.b2_shard_group_remove_ids(self._h, a.size, a.ctypes.data, ctypes.byref(removed)))
        return int(removed.value)

    def reset(self) -> None:
        self._N.check(self._N.lib.b2_shard_group_reset(self._h))

    # ---- what LangChain's FAISS glue needs beyond add / search (vectorstore.py accepts a ShardGroup as its index)
    device = property(lambda self: self.devices[0])

    def reconstruct_batch(self, ids):
        import numpy as np

        a = np.ascontiguousarray(np.asarray(ids, dtype=np.int64).reshape(-1))
        out = np.empty((a.size, self._d), dtype=np.float32)
        self._N.check(self._N.lib.b2_shard_group_reconstruct_batch(self._h, a.size, a.ctypes.data, out.ctypes.data))
        return out

    def reconstruct(self, key: int):
        key = int(key)
        if key < 0 or key >= self.ntotal:
            raise RuntimeError(f"reconstruct: key {key} out of range [0, {self.ntotal})")
        return self.reconstruct_batch([key])[0]

    def reconstruct_n(self, n0: int = 0, ni: int = -1):
        import numpy as np

        if ni == -1:
            ni = self.ntotal - n0
        if n0 < 0 or ni < 0 or n0 + ni > self.ntotal:
            raise RuntimeError(f"reconstruct_n: range [{n0}, {n0 + ni}) out of [0, {self.ntotal})")
        return self.reconstruct_batch(np.arange(n0, n0 + ni, dtype=np.int64))

    def merge_from(self, other, add_id: int = 0) -> None:
        raise NotImplementedError("merge_from into a shard group: add the other store's rows with add()")

    def mmr(self, x, k: int = 4, fetch_k: int = 20, lambda_mult: float = 0.5):
        """langchain's maximal_marginal_relevance over the group's search hits: global search of fetch_k, the hits' rows
        gathered from their shards, the greedy pick on the root device (one scratch index per query: this is the
        API-completeness path, the batched device MMR belongs to a single index)."""
        import numpy as np

        from .index import METRIC_INNER_PRODUCT, IndexFlat

        q = np.ascontiguousarray(np.atleast_2d(np.asarray(x.cpu() if hasattr(x, "cpu") else x, dtype=np.float32)))
        D, I = self.search(q, int(fetch_k))
        ids = np.full((q.shape[0], int(k)), -1, dtype=np.int64)
        sc = np.full((q.shape[0], int(k)), 0.0, dtype=np.float32)
        for r in range(q.shape[0]):
            hits = [int(i) for i in I[r] if i != -1]
            if not hits:
                continue
            rows = self.reconstruct_batch(hits)
            tmp = IndexFlat(self._d, METRIC_INNER_PRODUCT, "float32", self.devices[0])
            tmp.add(rows)
            # the candidates in search order are rows 0.. of the scratch index: fetch them all, pick k
            pick, _ = tmp.mmr(q[r:r + 1], k=int(k), fetch_k=len(hits), lambda_mult=lambda_mult)
            for j, p in enumerate(int(v) for v in pick[0] if v != -1):
                ids[r, j] = hits[p]
                sc[r, j] = D[r][p]
        return ids, sc
