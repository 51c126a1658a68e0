"""Row-sharded embedding tables across the GPUs of one box (BASELINE config 4).

Placement rule: row r of the logical [N, D] table lives on shard r % G at local row r // G
-- the rule the reference's GPU parameter server uses for its per-GPU hash tables
(paddle/fluid/framework/fleet/heter_ps/heter_comm_kernel.cu:85-93), chosen over c_embedding's
contiguous ranges (operators/collective/c_embedding_op.cu:48-49) because it spreads the Zipf
head evenly.

One process per GPU.  Each process allocates its shard with plain cudaMalloc (exportable),
publishes a CUDA IPC handle, and maps every peer's shard over NVLink; the fused pooling
kernel then fetches remote rows itself with ordinary loads through those mappings, in
sequence order, so sharded results are bit-identical to the single-table results.  There
is no bucketing, no all-to-all and no host synchronisation on the data path; torch.distributed
only carries the 64-byte handles at set-up (any backend: gloo works).
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


class _RawCudaBuffer:
    """Adapter so torch can view memory this library allocated (no copy)."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (ptr, False),
                                         "version": 2, "strides": None}


class ShardedTable:
    """This rank's view of a logical [n_rows, dim] table sharded over the process group."""

    def __init__(self, n_rows: int, dim: int, dtype, group=None, device=None):
        import torch
        import torch.distributed as dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.n_shards = dist.get_world_size(group)
        check_n_shards(self.n_shards)
        self.n_rows, self.dim, self.dtype = int(n_rows), int(dim), dtype
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.local_rows = shard_num_rows(self.n_rows, self.n_shards, self.rank)
        es = torch.empty(0, dtype=dtype).element_size()
        self.nbytes = max(self.local_rows, 1) * self.dim * es
        p = C.c_void_p()
        B.check(B.lib().ctrfe_shard_alloc(self.device.index, self.nbytes, C.byref(p)))
        self._own = p
        typestr = {torch.float32: "<f4", torch.float16: "<f2", torch.bfloat16: "<u2"}[dtype]
        t = torch.as_tensor(_RawCudaBuffer(p.value, (max(self.local_rows, 1), self.dim), typestr), device=self.device)
        self.local = t.view(torch.bfloat16) if dtype == torch.bfloat16 else t
        self.local = self.local[: self.local_rows]
        # publish / map
        hb = C.create_string_buffer(HANDLE_BYTES)
        B.check(B.lib().ctrfe_shard_export(self._own, hb))
        handles = exchange_handles(hb.raw, group)
        self._peer_ptrs: List[Optional[C.c_void_p]] = []
        ptrs = []
        for g in range(self.n_shards):
            if g == self.rank:
                self._peer_ptrs.append(None)
                ptrs.append(self._own.value)
            else:
                q = C.c_void_p()
                B.check(B.lib().ctrfe_shard_open(self.device.index, handles[g], C.byref(q)))
                self._peer_ptrs.append(q)
                ptrs.append(q.value)
        self.shard_ptrs = (C.c_void_p * self.n_shards)(*ptrs)
        dist.barrier(group)

    def close(self):
        import torch
        import torch.distributed as dist
        if getattr(self, "_own", None) is None:
            return
        torch.cuda.synchronize(self.device)
        dist.barrier(self.group)  # nobody unmaps or frees while a peer may still be reading
        for q in self._peer_ptrs:
            if q is not None:
                B.lib().ctrfe_shard_close(self.device.index, q)
        dist.barrier(self.group)
        B.lib().ctrfe_shard_free(self.device.index, self._own)
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
