"""Row-sharded embedding tables across the GPUs of one box (BASELINE config 4).

Placement rule: row r of the logical [N, D] table lives on shard r % G at local row r // G
-- the rule the reference's GPU parameter server uses for its per-GPU hash tables
(paddle/fluid/framework/fleet/heter_ps/heter_comm_kernel.cu:85-93), chosen over c_embedding's
contiguous ranges (operators/collective/c_embedding_op.cu:48-49) because it spreads the Zipf
head evenly.

One process per GPU.  Each process allocates its shard through the CUDA virtual-memory-
management API (2 MiB pages), hands a file descriptor for it to its peers, and maps every
peer's shard over NVLink; the fused pooling kernel then fetches remote rows itself with
ordinary loads through those mappings, in sequence order, so sharded results are
bit-identical to the single-table results.  There is no bucketing, no all-to-all and no host
synchronisation on the data path; torch.distributed only carries set-up metadata (any
backend: gloo works).  Large pages matter: with legacy cudaIpc mappings random row reads over
a 12.8 GB peer shard ran 50x slower on B200 (page walks), see profiles/README.md.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np

from . import _lib as B

HANDLE_BYTES = 64


# ---- placement arithmetic (pure host logic; also what the CPU tests exercise) -------------
def owner_of(ids, n_shards: int):
    """Shard that holds each row id."""
    return np.asarray(ids) % n_shards


def local_row_of(ids, n_shards: int):
    """Row index inside the owner's shard."""
    return np.asarray(ids) // n_shards


def shard_num_rows(n_rows: int, n_shards: int, shard: int) -> int:
    """How many logical rows r in [0, n_rows) satisfy r % n_shards == shard."""
    if not 0 <= shard < n_shards:
        raise ValueError("shard out of range")
    return (n_rows - shard + n_shards - 1) // n_shards if n_rows > shard else 0


def global_rows_of_shard(n_rows: int, n_shards: int, shard: int) -> np.ndarray:
    """The logical row ids a shard stores, in local-row order."""
    return np.arange(shard, n_rows, n_shards, dtype=np.int64)


def check_n_shards(n_shards: int) -> None:
    if n_shards not in (2, 4, 8):
        raise ValueError(f"the sharded table mode supports 2, 4 or 8 shards, got {n_shards}")


def exchange_handles(handle: bytes, group=None) -> List[bytes]:
    """All-gather the per-rank 64-byte IPC handles (rank order).  Backend-agnostic."""
    import torch.distributed as dist
    if len(handle) != HANDLE_BYTES:
        raise ValueError("an IPC handle is 64 bytes")
    out: List[Optional[bytes]] = [None] * dist.get_world_size(group)
    dist.all_gather_object(out, bytes(handle), group=group)
    return [bytes(h) for h in out]


def exchange_fds(fd: int, group=None, tag: str = "ctrfe") -> List[int]:
    """Give every rank of the (single-node) group a duplicate of every other rank's file
    descriptor: each rank serves its fd on a Unix socket, peers fetch it with SCM_RIGHTS.
    Returns the fds in rank order (this rank's own fd at its index)."""
    import os
    import socket
    import tempfile
    import threading
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    base = os.path.join(tempfile.gettempdir(), f"{tag}_{os.environ.get('MASTER_PORT', '0')}_{os.getuid()}")
    path = f"{base}_{rank}.sock"
    if os.path.exists(path):
        os.unlink(path)
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(path)
    srv.listen(world)

    def serve():
        for _ in range(world - 1):
            conn, _ = srv.accept()
            with conn:
                socket.send_fds(conn, [b"f"], [fd])

    t = threading.Thread(target=serve, daemon=True)
    t.start()
    dist.barrier(group)  # every socket is listening
    out = []
    for r in range(world):
        if r == rank:
            out.append(fd)
            continue
        with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
            c.connect(f"{base}_{r}.sock")
            _, fds, _, _ = socket.recv_fds(c, 1, 1)
            out.append(fds[0])
    t.join(60)
    dist.barrier(group)
    srv.close()
    os.unlink(path)
    return out


class _RawCudaBuffer:
    """Adapter so torch can view memory this library allocated (no copy)."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (ptr, False),
                                         "version": 2, "strides": None}


class ShardedTable:
    """This rank's view of a logical [n_rows, dim] table sharded over the process group.

    sharing="vmm" (default): shards are cuMemCreate allocations at 2 MiB granularity, passed to
    the peers as POSIX file descriptors -- peer mappings then use large pages, which random row
    reads over multi-GB shards need.  sharing="ipc": legacy cudaIpc handles (small-page peer
    mappings; fine for shards well under a GB)."""

    def __init__(self, n_rows: int, dim: int, dtype, group=None, device=None, sharing: str = "vmm"):
        import os
        import torch
        import torch.distributed as dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.n_shards = dist.get_world_size(group)
        check_n_shards(self.n_shards)
        if sharing not in ("vmm", "ipc"):
            raise ValueError("sharing must be 'vmm' or 'ipc'")
        self.sharing = sharing
        self.n_rows, self.dim, self.dtype = int(n_rows), int(dim), dtype
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.local_rows = shard_num_rows(self.n_rows, self.n_shards, self.rank)
        es = torch.empty(0, dtype=dtype).element_size()
        self.nbytes = max(self.local_rows, 1) * self.dim * es
        L = B.lib()
        p = C.c_void_p()
        self._peer_ptrs: List[Optional[C.c_void_p]] = []
        self._peer_sizes: List[int] = []
        ptrs = []
        if sharing == "vmm":
            fd, mapped = C.c_int(-1), C.c_uint64(0)
            B.check(L.ctrfe_shard_alloc_fd(self.device.index, self.nbytes, C.byref(p), C.byref(fd), C.byref(mapped)))
            self._own, self._own_mapped = p, mapped.value
            sizes: List[Optional[int]] = [None] * self.n_shards
            dist.all_gather_object(sizes, int(mapped.value), group=group)
            fds = exchange_fds(fd.value, group)
            for g in range(self.n_shards):
                if g == self.rank:
                    self._peer_ptrs.append(None)
                    self._peer_sizes.append(0)
                    ptrs.append(self._own.value)
                else:
                    q = C.c_void_p()
                    B.check(L.ctrfe_shard_open_fd(self.device.index, fds[g], sizes[g], C.byref(q)))
                    os.close(fds[g])
                    self._peer_ptrs.append(q)
                    self._peer_sizes.append(sizes[g])
                    ptrs.append(q.value)
            os.close(fd.value)
        else:
            B.check(L.ctrfe_shard_alloc(self.device.index, self.nbytes, C.byref(p)))
            self._own = p
            hb = C.create_string_buffer(HANDLE_BYTES)
            B.check(L.ctrfe_shard_export(self._own, hb))
            handles = exchange_handles(hb.raw, group)
            for g in range(self.n_shards):
                if g == self.rank:
                    self._peer_ptrs.append(None)
                    ptrs.append(self._own.value)
                else:
                    q = C.c_void_p()
                    B.check(L.ctrfe_shard_open(self.device.index, handles[g], C.byref(q)))
                    self._peer_ptrs.append(q)
                    ptrs.append(q.value)
        typestr = {torch.float32: "<f4", torch.float16: "<f2", torch.bfloat16: "<u2"}[dtype]
        t = torch.as_tensor(_RawCudaBuffer(self._own.value, (max(self.local_rows, 1), self.dim), typestr),
                            device=self.device)
        self.local = t.view(torch.bfloat16) if dtype == torch.bfloat16 else t
        self.local = self.local[: self.local_rows]
        self.shard_ptrs = (C.c_void_p * self.n_shards)(*ptrs)
        dist.barrier(group)

    def close(self):
        import torch
        import torch.distributed as dist
        if getattr(self, "_own", None) is None:
            return
        torch.cuda.synchronize(self.device)
        dist.barrier(self.group)  # nobody unmaps or frees while a peer may still be reading
        L = B.lib()
        for g, q in enumerate(self._peer_ptrs):
            if q is not None:
                if self.sharing == "vmm":
                    L.ctrfe_shard_unmap(self.device.index, q, self._peer_sizes[g])
                else:
                    L.ctrfe_shard_close(self.device.index, q)
        dist.barrier(self.group)
        self.local = None
        if self.sharing == "vmm":
            L.ctrfe_shard_unmap(self.device.index, self._own, self._own_mapped)
        else:
            L.ctrfe_shard_free(self.device.index, self._own)
        self._own = None


class LocalShards:
    """Several shards inside ONE process / GPU -- the same kernel path without peers.  Used to
    exercise the sharded addressing on a single-GPU box."""

    def __init__(self, shards: Sequence):
        check_n_shards(len(shards))
        self.n_shards = len(shards)
        self.shards = list(shards)
        self.dim = shards[0].shape[1]
        self.dtype = shards[0].dtype
        self.device = shards[0].device
        self.shard_ptrs = (C.c_void_p * self.n_shards)(*[s.data_ptr() for s in shards])
        self.n_rows = sum(s.shape[0] for s in shards)


def pick_hot_ids(sample_ids, capacity: int):
    """The `capacity` most frequent ids of a sample of look-ups (a torch int64 tensor of any
    shape, on the GPU), most frequent first.  Set-up helper, not on the data path."""
    import torch
    uniq, counts = torch.unique(sample_ids.reshape(-1), return_counts=True)
    order = torch.argsort(counts, descending=True)[:capacity]
    return uniq[order].contiguous()


class HotRowCache:
    """Local replicas of a table's most frequently looked-up rows (ctrfe_hot_cache): beside a sharded
    table (`table.hot = HotRowCache(table, ids)`), or beside one big unsharded table
    (`ops.ctr_fused_emb_seqpool_cvm(W, ..., hot_cache=HotRowCache(W, ids))`).

    Peer memory is not cached by the local L2, so without this every look-up of a popular row
    crosses NVLink again; with Zipf ids a few hundred MB of replicas absorb most look-ups.  The
    replicas are exact copies fetched through the sharded kernel itself (pooltype FIRST over
    one-id sequences), so results do not change by a bit.  The table is read-only at inference
    time; rebuild the cache if the table or the id distribution changes."""

    def __init__(self, table, hot_ids, ctx=None):
        import torch
        from . import ops
        dev = table.device
        hot_ids = hot_ids.to(dev).to(torch.int64).contiguous()
        n = hot_ids.numel()
        if n == 0:
            raise ValueError("no hot ids")
        ctx = ctx or ops.default_context(dev)
        lod = torch.arange(n + 1, dtype=torch.int64, device=dev)
        if isinstance(table, torch.Tensor):  # one unsharded table: the replicas are a compact, TLB-friendly region
            rows = ops.ctr_fused_emb_seqpool_cvm(table, [hot_ids], [lod], "FIRST", use_cvm=None, ctx=ctx)[0]
        else:
            saved, table.hot = getattr(table, "hot", None), None
            try:
                rows = ops.ctr_fused_emb_seqpool_cvm(table, [hot_ids], [lod], "FIRST", use_cvm=None, ctx=ctx)[0]
            finally:
                table.hot = saved
        src = rows.contiguous()
        log2 = max(4, int(np.ceil(np.log2(max(n, 1) * 2.0))))   # load factor <= 0.5
        self.log2_slots = log2
        self.keys = torch.full((1 << log2,), -1, dtype=torch.int32, device=dev)
        # the replica of a row sits at its hash slot: a hit is one 4-byte probe, no index look-up
        self.rows = torch.zeros((1 << log2, src.shape[1]), dtype=src.dtype, device=dev)
        B.check(B.lib().ctrfe_hot_cache_build(ctx.handle, hot_ids.data_ptr(), n, src.data_ptr(),
                                              src.shape[1] * src.element_size(), self.keys.data_ptr(),
                                              self.rows.data_ptr(), log2, torch.cuda.current_stream(dev).cuda_stream))
        torch.cuda.synchronize(dev)
        del src
        self.n_rows = n
        self.hot_ids = hot_ids
        self.struct = B.HotCache(self.rows.data_ptr(), self.keys.data_ptr(), log2, n)
        self.nbytes = self.rows.numel() * self.rows.element_size() + 4 * (1 << log2)


# ---- the bucket / all-to-all router (the exchange the reference's HeterComm performs) -----------------
class RoutedForward:
    """The north-star's shard router, built as the measured comparator of the peer-load kernel: ids
    are bucketed by owner shard, exchanged with NCCL all-to-all, the owners gather the rows, the rows
    come back the same way and the requester pools them in sequence order (so results stay
    bit-identical).  Follows HeterComm's pull (framework/fleet/heter_ps/heter_comm_inl.h:1117-1171
    split_input_to_shard = key % G bucket + sort; :1477-1607 pull_normal_sparse; :1296-1475 the
    de-duplicating pull_merge_sparse) and the counted send/recv group of
    operators/collective/global_scatter_op.cu.cc:109-135.  The bucketing uses torch's device sort and
    scan (the reference uses cub's radix sort); the gather and the pooling are the library's kernels;
    the exchanges are torch.distributed all_to_all_single (NCCL).  One host synchronisation per call
    (the counts), as in the reference."""

    def __init__(self, table: "ShardedTable", dedup: bool = False, gather=None, pool=None):
        """`gather(local_shard, rows)` and `pool(rows_in, back_slots, lods, pooltype, **kw)` default to
        the library's kernels; tests on CPU (gloo) pass stand-ins to drive the routing and the three
        exchanges without a GPU."""
        self.table, self.dedup = table, dedup
        self.G, self.rank = table.n_shards, table.rank
        self._gather, self._pool = gather, pool

    def __call__(self, Ids: Sequence, lods: Sequence, pooltype="SUM", *, use_cvm=True, concat=True, pad_value=0.0,
                 out=None, check: bool = False):
        import torch
        import torch.distributed as dist
        from . import ops
        G, grp = self.G, self.table.group
        sizes = [int(i.numel()) for i in Ids]
        all_ids = torch.cat([i.reshape(-1) for i in Ids]) if len(Ids) > 1 else Ids[0].reshape(-1)
        if self.dedup:  # pull_merge_sparse: every distinct id crosses the link once
            keys, inverse = torch.unique(all_ids, return_inverse=True)
        else:
            keys, inverse = all_ids, None
        owner = keys % G
        order = torch.argsort(owner, stable=True)            # bucket by owner shard
        send_rows = (keys[order] // G).contiguous()           # local row at the owner
        send_counts = torch.bincount(owner, minlength=G)
        recv_counts = torch.empty_like(send_counts)
        dist.all_to_all_single(recv_counts, send_counts, group=grp)
        sc, rc = send_counts.tolist(), recv_counts.tolist()   # the one host synchronisation
        recv_rows = torch.empty(sum(rc), dtype=torch.int64, device=all_ids.device)
        dist.all_to_all_single(recv_rows, send_rows, output_split_sizes=rc, input_split_sizes=sc, group=grp)
        gather = self._gather or (lambda w, r: ops.lookup_table_v2(w, r, check=False))
        pool = self._pool or ops.ctr_fused_emb_seqpool_cvm
        rows_out = gather(self.table.local, recv_rows)                                # owner-side gather
        rows_in = torch.empty((sum(sc), self.table.dim), dtype=self.table.dtype, device=all_ids.device)
        dist.all_to_all_single(rows_in, rows_out, output_split_sizes=sc, input_split_sizes=rc, group=grp)
        # position of every id's row in rows_in: the inverse of the bucket permutation (and of the dedup)
        pos = torch.empty_like(order)
        pos[order] = torch.arange(order.numel(), device=order.device)
        back = pos[inverse] if inverse is not None else pos
        back_slots = list(torch.split(back, sizes))
        return pool(rows_in, back_slots, lods, pooltype, use_cvm=use_cvm, concat=concat,
                    pad_value=pad_value, out=out, check=check)
