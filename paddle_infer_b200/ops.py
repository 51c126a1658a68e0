"""Host-side mirrors of the reference operator interfaces for the CTR sparse front end.

Same op names, argument names/meaning and error behaviour as the reference operators;
each function marshals into the C ABI (include/ctrfe.h) and launches sm_100a kernels.
The module itself needs numpy and the library only.  Arrays may be `device.DeviceArray`s (memory
from the library's own allocator), anything with `__cuda_array_interface__`, or torch CUDA
tensors -- torch is imported only when a torch tensor is passed in, and then outputs are torch
tensors and launches go to torch's current stream.

    lookup_table_v2            paddle/fluid/operators/lookup_table_v2_op.cc:26-104
    sequence_pool              paddle/fluid/operators/sequence_ops/sequence_pool_op.{cc,h}
    cvm                        paddle/fluid/operators/cvm_op.{cc,h}
    fused_seqpool_cvm          paddle/fluid/operators/fused/fused_seqpool_cvm_op.{cc,cu}
    fusion_seqpool_cvm_concat  paddle/fluid/operators/fused/fusion_seqpool_cvm_concat_op.cc
    fused_embedding_seq_pool   paddle/fluid/operators/fused/fused_embedding_seq_pool_op.{cc,h}
    pull_box_sparse            paddle/fluid/operators/pull_box_sparse_op.{cc,h}
    ctr_fused_emb_seqpool_cvm  NEW: the whole front end in one launch (SURVEY.md section 8b,
                               route 1 -- Paddle's custom-op loader refuses to re-register a
                               built-in name, framework/custom_operator.cc:722-725)

LoD arguments are offset-based lists of levels, as LoDTensor::lod() returns them
(see paddle_infer_b200.lod for the length-based conversion the Python API uses).
"""
from __future__ import annotations

import ctypes as C
import sys
from collections import namedtuple
from typing import List, Optional, Sequence

import numpy as np

from . import _lib as B
from .device import Arr, DeviceArray, PinnedArray, as_arr, empty

_FLOATS = ("float32", "float16", "bfloat16")
_IDS = {"int64": B.I64, "int32": B.I32}
_IDS_GATHER = {"int64": B.I64, "int32": B.I32, "int16": B.I16}   # lookup_table_v2 on its own also takes int16 ids


class InvalidArgument(ValueError):
    """platform::errors::InvalidArgument of the reference's enforce checks."""


class Unimplemented(NotImplementedError):
    """platform::errors::Unimplemented."""


# ------------------------------------------------------------------------------------------
# context plumbing
# ------------------------------------------------------------------------------------------
def _device_index(device) -> int:
    if device is None:
        if "torch" in sys.modules:
            import torch
            if torch.cuda.is_available():
                return torch.cuda.current_device()
        return 0
    if isinstance(device, int):
        return device
    idx = getattr(device, "index", None)  # a torch.device
    if idx is None and "torch" in sys.modules:
        import torch
        idx = torch.device(device).index
        if idx is None:
            idx = torch.cuda.current_device() if torch.cuda.is_available() else 0
    return int(idx or 0)


class Context:
    """One ctrfe_ctx per CUDA device (pinned staging ring, device scratch, error records).  Used with
    ONE stream at a time: torch's current stream when torch is in the process, else `stream_handle`
    (the legacy default stream unless set)."""

    def __init__(self, device=None):
        idx = _device_index(device)
        self.device_index = idx
        self.stream_handle: Optional[int] = None
        h = C.c_void_p()
        rc = B.lib().ctrfe_ctx_create(idx, C.byref(h))
        if rc != B.OK:
            raise RuntimeError("paddle_infer_b200 needs a CUDA device (sm_100a); there is no CPU fallback: "
                               + B.lib().ctrfe_last_error().decode("utf-8", "replace"))
        self.handle = h

    @property
    def device(self):
        import torch
        return torch.device("cuda", self.device_index)

    def stream(self) -> int:
        if self.stream_handle is not None:
            return self.stream_handle
        if "torch" in sys.modules:
            import torch
            return torch.cuda.current_stream(self.device_index).cuda_stream
        return 0

    def synchronize(self) -> None:
        B.check(B.lib().ctrfe_stream_synchronize(self.device_index, self.stream()))

    def check_ids(self) -> None:
        """Raise BadIdError if any launch since the last check met an id outside the table
        (what EmbeddingCPUFunctor enforces, phi/kernels/cpu/embedding_kernel.cc:59-76), or
        InvalidArgument if a kernel met a malformed device-resident LoD."""
        s, p, i = C.c_int64(-1), C.c_int64(-1), C.c_int64(-1)
        rc = B.lib().ctrfe_check_ids(self.handle, self.stream(), C.byref(s), C.byref(p), C.byref(i))
        if rc == B.ERR_BAD_ID:
            raise B.BadIdError(B.lib().ctrfe_last_error().decode(), s.value, p.value, i.value)
        if rc == B.ERR_ARG:  # a kernel met a malformed device-resident LoD (framework CheckLoD)
            raise InvalidArgument(B.lib().ctrfe_last_error().decode())
        B.check(rc)

    def close(self):
        if getattr(self, "handle", None):
            B.lib().ctrfe_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_ctxs = {}


def default_context(device=None) -> Context:
    idx = _device_index(device)
    if idx not in _ctxs:
        _ctxs[idx] = Context(idx)
    return _ctxs[idx]


def _last_level(lod, name="X") -> List[int]:
    if lod is None or len(lod) == 0:
        raise InvalidArgument(f"Input({name}) Tensor of SequencePoolOp does not contain LoD information.")
    if len(lod) > 2:
        raise InvalidArgument(f"The lod level of input shall be no more than 2.Received lod level is {len(lod)}.")
    if len(lod) == 2 and lod[0][-1] != len(lod[1]) - 1:
        raise InvalidArgument("The input lod information is illegal.")
    return lod[-1]


def _lod_np(level: Sequence[int]) -> np.ndarray:
    a = np.ascontiguousarray(np.asarray(level, dtype=np.uint64))
    if a.ndim != 1 or a.size < 1:
        raise InvalidArgument("a LoD level needs at least one offset")
    return a


def _float_arr(x, name, allow64=False) -> Arr:
    a = as_arr(x, name)
    if a.dtype not in _FLOATS and not (allow64 and a.dtype == "float64"):
        raise Unimplemented(f"{name}: dtype {a.dtype} has no sm_100a kernel (float32/float16/bfloat16)")
    return a


def _ids_arr(x, name="Ids") -> Arr:
    a = as_arr(x, name)
    if a.dtype not in _IDS:
        raise Unimplemented("emebdding input only support int32 and int64")
    return a


# ------------------------------------------------------------------------------------------
# lookup_table_v2
# ------------------------------------------------------------------------------------------
def lookup_table_v2(W, Ids, padding_idx: int = -1, *, check: bool = True, ctx: Optional[Context] = None):
    """Out = W[Ids] with a zero row where Ids == padding_idx; Out.shape = Ids.shape + [D]
    (lookup_table_v2_op.cc:45-66).  Out shares Ids' LoD (the caller keeps carrying it).
    check=True synchronises and raises BadIdError for ids outside [0, N), as the reference
    CPU kernel does; pass check=False in a serving loop and call ctx.check_ids() at will."""
    w, ids = as_arr(W, "W"), as_arr(Ids, "Ids")
    if len(w.shape) != 2:
        raise InvalidArgument(f"ShapeError: The dimensions of the 'lookup table' must be 2. But received "
                              f"lookup table's dimensions = {len(w.shape)}.")
    if ids.dtype not in _IDS_GATHER:
        raise Unimplemented("emebdding input only support int16, int32 and int64")
    ctx = ctx or default_context(w.device)
    out, o = empty(ids.shape + (w.shape[1],), w.dtype, w)
    B.check(B.lib().ctrfe_embedding(ctx.handle, w.ptr, w.shape[0], w.shape[1] * w.itemsize, ids.ptr, _IDS_GATHER[ids.dtype],
                                    ids.numel, int(padding_idx), o.ptr, ctx.stream()))
    if check:
        ctx.check_ids()
    else:
        ctx._keep = (w.obj, ids.obj)
    return out


# ------------------------------------------------------------------------------------------
# sequence_pool
# ------------------------------------------------------------------------------------------
SeqPoolOut = namedtuple("SeqPoolOut", ["Out", "MaxIndex", "lod"])


def sequence_pool(X, lod, pooltype: str = "AVERAGE", pad_value: float = 0.0, is_test: bool = False, *,
                  ctx: Optional[Context] = None) -> SeqPoolOut:
    """Pools each sequence of the LAST LoD level of X ([K, ...]) into one row.
    Out: [num_seq, ...]; Out's LoD is lod[0] when X has two levels, else none
    (sequence_pool_op.h:42-72).  MaxIndex (int32, Out's shape) is produced for MAX unless
    is_test (the reference skips it for is_test on CPU, sequence_pool_op.h:79-87).
    Empty sequences yield pad_value."""
    x = as_arr(X, "X")
    if pooltype not in B.POOLTYPES:
        raise InvalidArgument(f"unsupported pooling pooltype: {pooltype}. Only support \"AVERAGE\", \"SUM\", "
                              "\"SQRT\", \"LAST\", \"FIRST\" and \"MAX\"")
    if x.dtype not in _FLOATS:
        raise Unimplemented(f"sequence_pool: dtype {x.dtype} has no sm_100a kernel (float32/float16/bfloat16)")
    last = _last_level(lod)
    n_seq = len(last) - 1
    if x.shape[0] < n_seq:
        raise InvalidArgument("The first dimension of Input(X) must be large than batch size."
                              f"But received first dimension of Input(X) is {x.shape[0]}, while batch"
                              f"size is {n_seq}.")
    if last[-1] > x.shape[0]:
        raise InvalidArgument(f"The last LoD offset ({last[-1]}) exceeds the first dimension of Input(X) "
                              f"({x.shape[0]}).")
    ctx = ctx or default_context(x.device)
    width = int(np.prod(x.shape[1:])) if len(x.shape) > 1 else 1
    out, o = empty((n_seq,) + x.shape[1:], x.dtype, x)
    want_index = pooltype == "MAX" and not is_test
    index, ix = empty(o.shape, "int32", x) if want_index else (None, None)
    lod_a = _lod_np(last)
    B.check(B.lib().ctrfe_sequence_pool(ctx.handle, x.ptr, x.code, width, lod_a.ctypes.data, n_seq,
                                        B.POOLTYPES[pooltype], float(pad_value), o.ptr,
                                        ix.ptr if want_index else None, ctx.stream()))
    ctx._keep = (x.obj,)
    out_lod = [list(lod[0])] if len(lod) == 2 else []
    return SeqPoolOut(out, index, out_lod)


# ------------------------------------------------------------------------------------------
# cvm
# ------------------------------------------------------------------------------------------
def cvm(X, CVM=None, use_cvm: bool = True, *, ctx: Optional[Context] = None):
    """Y[:,0] = log(X[:,0]+1), Y[:,1] = log(X[:,1]+1) - Y[:,0], rest copied (use_cvm=True);
    or X[:, 2:] (use_cvm=False).  X must be rank 2 (cvm_op.cc:34-38).  The CVM input is
    accepted for signature parity and, as in the reference forward, never read
    (cvm_op.h:60-99).  Y shares X's LoD.  float32 / float64 / float16 / bfloat16."""
    x = as_arr(X, "X")
    if len(x.shape) != 2:
        raise InvalidArgument(f"Input(X)'s rank should be 2, but got {len(x.shape)}")
    if x.dtype not in _FLOATS + ("float64",):
        raise Unimplemented(f"cvm: dtype {x.dtype} has no sm_100a kernel (float32/float64/float16/bfloat16)")
    ctx = ctx or default_context(x.device)
    rows, w = x.shape
    if w < 2:
        raise InvalidArgument("Input(X) of cvm needs at least the show and click columns")
    y, yo = empty((rows, w if use_cvm else w - 2), x.dtype, x)
    B.check(B.lib().ctrfe_cvm(ctx.handle, x.ptr, x.code, rows, w, int(bool(use_cvm)), yo.ptr, ctx.stream()))
    ctx._keep = (x.obj,)
    return y


# ------------------------------------------------------------------------------------------
# backward of the three ops (fp32; SURVEY 8(f) row 4 -- inference never calls these)
# ------------------------------------------------------------------------------------------
def _f32(x, name) -> Arr:
    a = as_arr(x, name)
    if a.dtype != "float32":
        raise Unimplemented(f"{name}: the gradient kernels are float32 only")
    return a


def sequence_pool_grad(OutGrad, lod, pooltype: str, x_rows: Optional[int] = None, MaxIndex=None, *,
                       ctx: Optional[Context] = None):
    """X@GRAD of sequence_pool (sequence_pool_op.h:99-121): [x_rows, ...] from Out@GRAD
    [num_seq, ...] and X's LoD; MAX needs the forward op's MaxIndex."""
    if pooltype not in B.POOLTYPES:
        raise InvalidArgument(f"unsupported pooling pooltype: {pooltype}")
    g = _f32(OutGrad, "Out@GRAD")
    last = _last_level(lod)
    n_seq = len(last) - 1
    if g.shape[0] != n_seq:
        raise InvalidArgument(f"Out@GRAD has {g.shape[0]} rows, the LoD names {n_seq} sequences")
    rows = last[-1] if x_rows is None else int(x_rows)
    width = int(np.prod(g.shape[1:])) if len(g.shape) > 1 else 1
    idx = None
    if pooltype == "MAX":
        if MaxIndex is None:
            raise InvalidArgument("sequence_pool_grad(MAX) needs MaxIndex")
        idx = as_arr(MaxIndex, "MaxIndex")
        if idx.dtype != "int32":
            raise InvalidArgument("MaxIndex must be int32")
    ctx = ctx or default_context(g.device)
    dx, dxo = empty((rows,) + g.shape[1:], "float32", g)
    lod_a = _lod_np(last)
    B.check(B.lib().ctrfe_sequence_pool_grad(ctx.handle, g.ptr, width, lod_a.ctypes.data, n_seq, rows,
                                             B.POOLTYPES[pooltype], idx.ptr if idx is not None else None,
                                             dxo.ptr, ctx.stream()))
    ctx._keep = (g.obj, idx.obj if idx is not None else None)
    return dx


def cvm_grad(YGrad, CVM, use_cvm: bool = True, lod=None, *, ctx: Optional[Context] = None):
    """X@GRAD of cvm (cvm_op.h:100-140): columns 0-1 take the CVM input (show, click) of the row's
    sequence, the rest Y@GRAD.  lod: X's LoD when it has one (CVM then has one row per sequence),
    else None (one CVM row per row of X)."""
    g, c = _f32(YGrad, "Y@GRAD"), _f32(CVM, "CVM")
    if len(g.shape) != 2 or len(c.shape) != 2 or c.shape[1] != 2:
        raise InvalidArgument("Y@GRAD must be rank 2 and CVM [n, 2]")
    rows = g.shape[0]
    w = g.shape[1] + (0 if use_cvm else 2)
    ctx = ctx or default_context(g.device)
    dx, dxo = empty((rows, w), "float32", g)
    lod_a = None
    if lod is not None and len(lod) > 0:
        lod_a = _lod_np(lod[0] if isinstance(lod[0], (list, tuple, np.ndarray)) else lod)
        if lod_a.size - 1 != c.shape[0]:
            raise InvalidArgument("CVM needs one row per sequence of X")
    B.check(B.lib().ctrfe_cvm_grad(ctx.handle, g.ptr, c.ptr, lod_a.ctypes.data if lod_a is not None else None,
                                   c.shape[0], rows, w, int(bool(use_cvm)), dxo.ptr, ctx.stream()))
    ctx._keep = (g.obj, c.obj)
    return dx


def lookup_table_v2_grad(OutGrad, Ids, table_rows: int, padding_idx: int = -1, *, check: bool = True,
                         ctx: Optional[Context] = None):
    """Dense W@GRAD of lookup_table_v2 (phi embedding_grad): [table_rows, D]; rows of ids equal to
    padding_idx get no gradient.  Atomic adds: equal to the CPU kernel up to rounding."""
    g, ids = _f32(OutGrad, "Out@GRAD"), _ids_arr(Ids)
    d = g.shape[-1]
    if g.numel // d != ids.numel:
        raise InvalidArgument("Out@GRAD must have one row per id")
    ctx = ctx or default_context(g.device)
    dw, dwo = empty((table_rows, d), "float32", g)
    B.check(B.lib().ctrfe_embedding_grad(ctx.handle, g.ptr, ids.ptr, _IDS[ids.dtype], ids.numel,
                                         table_rows, d, int(padding_idx), dwo.ptr, ctx.stream()))
    if check:
        ctx.check_ids()
    else:
        ctx._keep = (g.obj, ids.obj)
    return dw


SelectedRows = namedtuple("SelectedRows", ["rows", "value", "height"])


def lookup_table_v2_grad_sparse(OutGrad, Ids, table_rows: int, *, ctx: Optional[Context] = None) -> SelectedRows:
    """W@GRAD of lookup_table_v2 with is_sparse=True, as the reference's SelectedRows: rows = the ids
    as int64 (one entry per id, not merged), value = Out@GRAD as [n_ids, D] (a copy), height =
    table_rows (phi/kernels/gpu/embedding_grad_kernel.cu:141-214; padding_idx is not applied there)."""
    g, ids = as_arr(OutGrad, "Out@GRAD"), as_arr(Ids, "Ids")
    if ids.dtype not in _IDS_GATHER:
        raise Unimplemented("emebdding input only support int16, int32 and int64")
    d = g.shape[-1]
    if g.numel // max(d, 1) != ids.numel:
        raise InvalidArgument("ShapeError: The shape of lookup_table@Grad and output@Grad should be same.")
    ctx = ctx or default_context(g.device)
    rows, ro = empty((ids.numel,), "int64", g)
    value, vo = empty((ids.numel, d), g.dtype, g)
    B.check(B.lib().ctrfe_embedding_grad_sparse(ctx.handle, g.ptr, d * g.itemsize, ids.ptr, _IDS_GATHER[ids.dtype], ids.numel,
                                                ro.ptr, None, vo.ptr, ctx.stream()))
    ctx._keep = (g.obj, ids.obj)
    return SelectedRows(rows, value, int(table_rows))


# ------------------------------------------------------------------------------------------
# the fused front end
# ------------------------------------------------------------------------------------------
def _as_list(x, n):
    return list(x) if isinstance(x, (list, tuple)) else [x] * n


class FusedPlan:
    """A prepared batch: descriptors + LoD resident on the device; launch() enqueues ONE
    kernel for all slots.  Keeps the arrays it names alive."""

    def __init__(self, ctx: Context, handle, outs, keep):
        self.ctx, self.handle, self.outs, self._keep = ctx, handle, outs, keep

    def launch(self, overlap_previous: bool = False) -> None:
        """overlap_previous=True: this launch does not consume the output of the kernel right
        before it in the stream (back-to-back batches), so it may start while that one drains."""
        if overlap_previous:
            B.check(B.lib().ctrfe_plan_launch_ex(self.handle, self.ctx.stream(), B.LAUNCH_OVERLAP_PREVIOUS))
        else:
            B.check(B.lib().ctrfe_plan_launch(self.handle, self.ctx.stream()))

    def num_kernels(self) -> int:
        return B.lib().ctrfe_plan_num_kernels(self.handle)

    def kernel_names(self) -> List[str]:
        return [B.lib().ctrfe_plan_kernel_name(self.handle, i).decode() for i in range(self.num_kernels())]

    def update_lod(self, lods: Sequence) -> None:
        """New host LoD offsets for the plan's slots (same sequence counts); they reach the device
        with the next launch_many(..., upload_lod=True)."""
        arrs = [_lod_np(l) for l in lods]
        ptrs = (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
        B.check(B.lib().ctrfe_plan_update_lod(self.handle, ptrs))

    def close(self):
        if self.handle:
            B.lib().ctrfe_plan_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def launch_many(plans: Sequence[FusedPlan], overlap_previous: bool = False, upload_lod: bool = False) -> None:
    """Launch several prepared batches back to back with ONE call into the library
    (ctrfe_plan_launch_many): the inner step of a serving loop without a Python call per batch."""
    if not plans:
        return
    arr = (C.c_void_p * len(plans))(*[p.handle.value for p in plans])
    flags = (B.LAUNCH_OVERLAP_PREVIOUS if overlap_previous else 0) | (B.LAUNCH_UPLOAD_LOD if upload_lod else 0)
    B.check(B.lib().ctrfe_plan_launch_many(arr, len(plans), plans[0].ctx.stream(), flags))


class _TableRef:
    """What a slot descriptor needs from a table: a base pointer and a row count."""

    def __init__(self, ptr, rows):
        self.ptr, self.shape = ptr, (rows,)


def _is_host_lod(l) -> bool:
    return isinstance(l, (list, tuple, np.ndarray))


def _build_desc(tables, ids, lods, pooltypes, out_ptrs, out_ld, *, dim, dtype_code, pad_value, padding_idx, cvm_mode,
                cvm_offset, scale_mode, init_mode, lod_on_device, sharded=None, keymap=None, hot_cache=None):
    """tables: Arr / _TableRef per slot; ids: Arr or None per slot; lods: host sequences or device Arrs."""
    S = len(lods)
    slots = (B.Slot * S)()
    keep = []
    if sharded is not None:  # slot.table is ignored; table_rows is the LOGICAL row count
        tables = [_TableRef(None, sharded.n_rows)] * S
    for s in range(S):
        t = tables[s]
        if lod_on_device:
            la = as_arr(lods[s], "lod")
            lod_ptr, n_seg = la.ptr, la.numel - 1
            keep.append(la.obj)
        else:
            a = _lod_np(lods[s])
            keep.append(a)
            lod_ptr, n_seg = a.ctypes.data, a.size - 1
        i = ids[s]
        slots[s] = B.Slot(t.ptr, t.shape[0], i.ptr if i is not None else None, lod_ptr, n_seg,
                          out_ptrs[s], out_ld, B.POOLTYPES[pooltypes[s]],
                          _IDS[i.dtype] if i is not None else B.I64, i.numel if i is not None else 0)
        keep.append((getattr(t, "obj", None), i.obj if i is not None else None))
    desc = B.FusedDesc(S, dtype_code, slots, dim, int(padding_idx), float(pad_value), cvm_mode, cvm_offset, scale_mode,
                       init_mode, B.LOD_DEVICE if lod_on_device else B.LOD_HOST, 0)
    if sharded is not None:
        desc.n_shards = sharded.n_shards
        desc.shard_tables = C.cast(sharded.shard_ptrs, C.POINTER(C.c_void_p))
        hot = getattr(sharded, "hot", None)
        if hot is not None:  # local replicas of the hot rows (shard.HotRowCache)
            desc.hot_cache = C.pointer(hot.struct)
            keep.append(hot)
        keep.append(sharded)
    if hot_cache is not None and sharded is None:  # compact replicas of the hot rows beside ONE unsharded table
        desc.hot_cache = C.pointer(hot_cache.struct)
        keep.append(hot_cache)
    if keymap is not None:  # ids are uint64 feasign keys, resolved in the kernel (KeyMap)
        desc.keymap = C.pointer(keymap.struct)
        keep.append(keymap)
    keep.append(slots)
    return desc, keep


def _is_sharded(W) -> bool:
    return hasattr(W, "shard_ptrs") and hasattr(W, "n_shards")


def _sharded_like(W) -> Arr:
    """An Arr standing for a sharded table's element type and device (outputs are allocated like it)."""
    import torch
    t0 = torch.empty((0, W.dim), dtype=W.dtype, device=W.device)
    return Arr(0, (W.n_rows, W.dim), str(W.dtype).replace("torch.", ""), t0.device.index, "torch", t0)


def ctr_fused_emb_seqpool_cvm(W, Ids: Sequence, lods: Sequence, pooltype="SUM", *,
                              pad_value: float = 0.0, padding_idx: int = -1, use_cvm: Optional[bool] = True,
                              cvm_offset: int = 2, concat: bool = False, scale_mode: int = B.SCALE_DIVIDE,
                              init_mode: int = B.INIT_ZERO, check: bool = True, plan_only: bool = False,
                              out=None, ctx: Optional[Context] = None, keymap: Optional["KeyMap"] = None, hot_cache=None):
    """lookup_table_v2 -> sequence_pool -> cvm for every slot of a CTR batch in one launch.

    W         one table [N, D] shared by all slots, or a list with one table per slot, or a
              shard.ShardedTable: one logical table row-sharded over the GPUs of the box, whose
              remote rows the kernel fetches itself over NVLink (results are bit-identical)
    Ids       per slot: 1-D int64/int32 ids, LoD-packed
    lods      per slot: offsets of the last LoD level -- a host sequence / numpy array, or a
              device uint64-as-int64 array when the offsets are already on the device (then the
              kernels check every pair they read, see ctx.check_ids)
    pooltype  one name or one per slot
    use_cvm   True: cvm(use_cvm=True); False: cvm(use_cvm=False) (drops cvm_offset columns);
              None: no cvm op after the pool
    concat    write all slots into one [B, S*D_out] array (concat(axis=1) fused, as
              fusion_seqpool_cvm_concat does)
    hot_cache a shard.HotRowCache built over W (one table read by every slot): lookups of the hottest rows
              go to a compact replica region that the TLB and L2 can hold -- for a table beyond the
              TLB's reach (> ~64 GB) where every cold row costs a page walk; results are bit-identical
    keymap    a KeyMap: Ids are then uint64 feasign keys (viewed as int64, as pull_box_sparse
              takes them) and W holds one row per key of the map; key 0 and unknown keys pool
              as zero rows
    Returns a list of [B_s, D_out] arrays, or the [B, S*D_out] array when concat=True;
    with plan_only=True returns a FusedPlan instead (nothing launched yet).
    """
    S = len(Ids)
    sharded = W if _is_sharded(W) else None
    pools = _as_list(pooltype, S)
    for p in pools:
        if p not in B.POOLTYPES:
            raise InvalidArgument(f"unsupported pooling pooltype: {p}")
    if sharded is not None:
        tables: List = []
        t0 = _sharded_like(sharded)
    else:
        tl = _as_list(W, S)
        if len(tl) != S:
            raise InvalidArgument("W must be one table or one per slot")
        tables = [as_arr(t, "W") for t in tl]
        t0 = tables[0]
    if t0.dtype not in _FLOATS:
        raise Unimplemented(f"dtype {t0.dtype} has no sm_100a kernel (float32/float16/bfloat16)")
    dim = t0.shape[1]
    for t in tables:
        if len(t.shape) != 2 or t.shape[1] != dim or t.dtype != t0.dtype:
            raise InvalidArgument("all tables must be contiguous [N, D] with equal D and dtype")
    ids_a = [_ids_arr(i) for i in Ids]
    ctx = ctx or default_context(t0.device)
    cvm_mode = B.CVM_NONE if use_cvm is None else (B.CVM_USE if use_cvm else B.CVM_STRIP)
    d_out = dim - cvm_offset if cvm_mode == B.CVM_STRIP else dim
    lod_on_device = not _is_host_lod(lods[0])
    n_segs = [(as_arr(l, "lod").numel if lod_on_device else len(l)) - 1 for l in lods]
    es = t0.itemsize
    if concat:
        if len(set(n_segs)) != 1:
            raise InvalidArgument("Batchsize of all inputs should be equal.")
        if out is None:
            out, oa = empty((n_segs[0], S * d_out), t0.dtype, t0)
        else:
            oa = as_arr(out, "out")
        out_ptrs = [oa.ptr + s * d_out * es for s in range(S)]
        ld = S * d_out
        result = out
    else:
        pairs = [empty((n, d_out), t0.dtype, t0) for n in n_segs]
        out_ptrs = [a.ptr for _, a in pairs]
        ld = d_out
        result = [o for o, _ in pairs]
    desc, keep = _build_desc(tables, ids_a, lods, pools, out_ptrs, ld, dim=dim, dtype_code=t0.code,
                             pad_value=pad_value, padding_idx=padding_idx, cvm_mode=cvm_mode, cvm_offset=cvm_offset,
                             scale_mode=scale_mode, init_mode=init_mode, lod_on_device=lod_on_device, sharded=sharded,
                             keymap=keymap, hot_cache=hot_cache)
    if plan_only:
        h = C.c_void_p()
        B.check(B.lib().ctrfe_plan_create(ctx.handle, C.byref(desc), ctx.stream(), C.byref(h)))
        return FusedPlan(ctx, h, result, (keep, result))
    B.check(B.lib().ctrfe_fused_forward(ctx.handle, C.byref(desc), ctx.stream()))
    if check:
        ctx.check_ids()
    else:
        ctx._keep = keep
    return result


class KeyMap:
    """uint64 feasign key -> table row, resident on the GPU (ctrfe_keymap): the lookup half of
    the reference's per-GPU HashTable (framework/fleet/heter_ps/hashtable_kernel.cu), which
    pull_box_sparse / pull_gpups_sparse consult.  keys[i] maps to row i of the value table the
    caller keeps beside it; keys are non-zero (0 is the padding feasign); of duplicate keys the
    first occurrence wins."""

    def __init__(self, keys, ctx: Optional[Context] = None):
        k = as_arr(keys, "keys")
        if k.dtype != "int64":
            raise InvalidArgument("feasign keys are uint64 values held in an int64 array")
        n = k.numel
        ctx = ctx or default_context(k.device)
        log2 = max(4, int(np.ceil(np.log2(max(n, 1) * 2.0))))  # load factor <= 0.5
        while True:
            if log2 > 31:
                raise InvalidArgument("too many keys for one key map")
            self.keys = DeviceArray((1 << log2,), "int64", k.device)
            self.rows = DeviceArray((1 << log2,), "int32", k.device)
            unplaced = DeviceArray((1,), "int32", k.device)
            zero = np.zeros(1 << log2, np.int64)
            self.keys.fill_bytes_from(zero, ctx.stream())
            unplaced.fill_bytes_from(np.zeros(1, np.int32), ctx.stream())
            B.check(B.lib().ctrfe_keymap_build(ctx.handle, k.ptr, n, self.keys.ptr, self.rows.ptr,
                                               log2, unplaced.ptr, ctx.stream()))
            if int(unplaced.numpy(ctx.stream())[0]) == 0:
                break
            log2 += 1  # a probe chain longer than the kernel follows: spread the keys out
        self.log2_slots, self.n_keys = log2, n
        self.struct = B.KeyMap(self.keys.ptr, self.rows.ptr, log2, 0, n)


def pull_box_sparse(keymap: KeyMap, W, Ids: Sequence, *, ctx: Optional[Context] = None) -> List:
    """pull_box_sparse / pull_gpups_sparse over a resident table (operators/pull_box_sparse_op.h:
    32-62, .cc:62-100): per slot, Ids [K_s(, 1)] uint64 feasign keys -> Out [K_s, D] with
    D = W.shape[1] (the op's `size`); key 0 and unknown keys give zero rows.  One launch for all
    slots.  The fused call (ctr_fused_emb_seqpool_cvm(..., keymap=...)) never writes these
    values; this clone exists for programs that consume them elsewhere."""
    flat = [as_arr(i, "Ids") for i in Ids]
    lods = [np.arange(i.numel + 1, dtype=np.uint64) for i in flat]
    flat1 = [Arr(i.ptr, (i.numel,), i.dtype, i.device, i.kind, i.obj) for i in flat]
    return ctr_fused_emb_seqpool_cvm(W, flat1, lods, "FIRST", use_cvm=None, keymap=keymap, ctx=ctx, check=True)


def _dense_slots(Xs, lods):
    flat = []
    for x in Xs:
        a = _float_arr(x, "X")
        w = a.shape[1] if len(a.shape) == 2 else int(np.prod(a.shape[1:]))
        flat.append(Arr(a.ptr, (a.shape[0], w), a.dtype, a.device, a.kind, a.obj))
    w = flat[0].shape[1] if flat else 0
    for f in flat:
        if f.shape[1] != w or f.dtype != flat[0].dtype:
            raise InvalidArgument("Width of all inputs should be equal.")
    lasts = [_last_level(l) for l in lods]
    for f, l in zip(flat, lasts):
        if l[-1] > f.shape[0]:
            raise InvalidArgument("The last LoD offset exceeds the first dimension of Input(X).")
    return flat, lasts, w


def fused_seqpool_cvm(X: Sequence, lods: Sequence, CVM=None, pooltype: str = "SUM", pad_value: float = 0.0,
                      use_cvm: bool = True, cvm_offset: int = 2, *, ctx: Optional[Context] = None) -> List:
    """PaddleBox's multi-slot op over PRE-PULLED values (fused/fused_seqpool_cvm_op.cc:120-150):
    X is a list of [K_s, D] LoD arrays, one per slot; SUM only; the accumulator STARTS at
    pad_value (fused_seqpool_cvm_op.cu:49), so an empty sequence yields pad_value too;
    use_cvm=False drops cvm_offset leading columns.  One launch instead of the reference's
    two kernels + intermediate buffer."""
    if pooltype != "SUM":
        raise InvalidArgument("fused_seqpool_cvm only supports pooltype SUM")  # op maker InEnum({"SUM"})
    if not X:
        return []
    flat, lasts, w = _dense_slots(X, lods)
    ctx = ctx or default_context(flat[0].device)
    S = len(flat)
    cvm_mode = B.CVM_USE if use_cvm else B.CVM_STRIP
    d_out = w if use_cvm else w - cvm_offset
    pairs = [empty((len(l) - 1, d_out), flat[0].dtype, flat[0]) for l in lasts]
    desc, keep = _build_desc(flat, [None] * S, lasts, ["SUM"] * S, [a.ptr for _, a in pairs], d_out, dim=w,
                             dtype_code=flat[0].code, pad_value=pad_value, padding_idx=-1, cvm_mode=cvm_mode,
                             cvm_offset=cvm_offset, scale_mode=B.SCALE_DIVIDE, init_mode=B.INIT_PAD, lod_on_device=False)
    B.check(B.lib().ctrfe_fused_forward(ctx.handle, C.byref(desc), ctx.stream()))
    ctx._keep = keep
    return [o for o, _ in pairs]


def fusion_seqpool_cvm_concat(X: Sequence, lods: Sequence, CVM=None, pooltype: str = "SUM", use_cvm: bool = True,
                              axis: int = 1, *, ctx: Optional[Context] = None):
    """The CPU predictor's fused op (fused/fusion_seqpool_cvm_concat_op.cc:97-159): per slot
    sequence_pool (SUM / AVERAGE / SQRT, jit kernel: scale by reciprocal) -> cvm(use_cvm=True)
    -> written at column s*w of Out [B, S*w].  axis must be 1, use_cvm must be True, as the
    reference enforces (:58-70)."""
    if axis != 1:
        raise InvalidArgument("fusion_seqpool_cvm_concat only supports concat axis=1 yet.")
    if not use_cvm:
        raise InvalidArgument("fusion_seqpool_cvm_concat only supports use_cvm is true yet.")
    if pooltype not in ("SUM", "AVERAGE", "SQRT"):
        raise InvalidArgument(f"fusion_seqpool_cvm_concat: unsupported pooltype {pooltype}")
    flat, lasts, w = _dense_slots(X, lods)
    S = len(flat)
    bs = len(lasts[0]) - 1
    for l in lasts:
        if len(l) - 1 != bs:
            raise InvalidArgument("Batchsize of all inputs should be equal.")
    ctx = ctx or default_context(flat[0].device)
    out, oa = empty((bs, S * w), flat[0].dtype, flat[0])
    es = flat[0].itemsize
    desc, keep = _build_desc(flat, [None] * S, lasts, [pooltype] * S, [oa.ptr + s * w * es for s in range(S)], S * w, dim=w,
                             dtype_code=flat[0].code, pad_value=0.0, padding_idx=-1, cvm_mode=B.CVM_USE, cvm_offset=2,
                             scale_mode=B.SCALE_RECIPROCAL, init_mode=B.INIT_ZERO, lod_on_device=False)
    B.check(B.lib().ctrfe_fused_forward(ctx.handle, C.byref(desc), ctx.stream()))
    ctx._keep = keep
    return out


def fused_embedding_seq_pool(W, Ids, lod, combiner: str = "sum", padding_idx: int = -1, is_sparse: bool = False, *,
                             check: bool = True, ctx: Optional[Context] = None):
    """Gather + SUM per sequence (fused/fused_embedding_seq_pool_op.h:84-127, the EmbSeqPool
    route: first row copied, the rest added in id order).  Ids: [K, idx_width] int64 with one LoD
    level over its rows; Out: [num_seq, D * idx_width], column j of Ids pooled into columns
    [j*D, (j+1)*D) (jit/refer/refer.h:477-528)."""
    if combiner != "sum":
        raise Unimplemented("fused_embedding_seq_pool only supports combiner 'sum'")
    if len(lod) != 1:
        raise InvalidArgument("The LoD level of Input(Ids) should be 1.")
    w, ids = _float_arr(W, "W"), as_arr(Ids, "Ids")
    if len(w.shape) != 2 or len(ids.shape) > 2:
        raise InvalidArgument("The dimension of W and Ids should be 2.")
    if ids.dtype != "int64":
        raise Unimplemented("fused_embedding_seq_pool: Ids must be int64")
    n_rows = ids.shape[0]
    width = ids.numel // max(n_rows, 1) if n_rows else 1
    last = _lod_np(lod[0])
    n_seg = last.size - 1
    if int(last[-1]) > n_rows:
        raise InvalidArgument("The last LoD offset exceeds the first dimension of Input(Ids).")
    ctx = ctx or default_context(w.device)
    out, oa = empty((n_seg, w.shape[1] * width), w.dtype, w)
    B.check(B.lib().ctrfe_fused_embedding_seq_pool(ctx.handle, w.ptr, w.shape[0], w.shape[1], w.code, ids.ptr, n_rows, width,
                                                   last.ctypes.data, n_seg, int(padding_idx), oa.ptr, ctx.stream()))
    if check:
        ctx.check_ids()
    else:
        ctx._keep = (w.obj, ids.obj)
    return out


def _TORCH_DT():
    import torch
    return {torch.float32: B.F32, torch.float16: B.F16, torch.bfloat16: B.BF16, torch.float64: B.F64}


def _TORCH_IDT():
    import torch
    return {torch.int64: B.I64, torch.int32: B.I32}


# ------------------------------------------------------------------------------------------
# host-buffer doorway (what a predictor feeding from CPU memory calls)
# ------------------------------------------------------------------------------------------
class HostFrontEnd:
    """The fused front end with HOST inputs and outputs: per-slot id arrays and LoD offsets in
    pinned host memory in, the pooled rows in pinned host memory out.  This is the shape of the
    reference's feed path -- ZeroCopyTensor::CopyFromCpu + SetLoD per slot, CopyToCpu on the
    result (inference/api/details/zero_copy_tensor.cc) -- around one fused launch.  run() is
    synchronous on return and yields one [B, D_out] pinned host view per slot (the op's
    Out@VECTOR); concat=True instead fills one [B, S*D_out] block.

    mode="zero_copy" (default): the kernel itself reads the ids and LoD offsets from the pinned
    host buffers and writes the pooled rows into the pinned output, over PCIe, inside the one
    launch -- no staging copies, and the transfers overlap the table gathers by construction.
    The buffers must be pinned (cudaHostAlloc / torch pin_memory) so the device can address
    them; the LoD offsets are then checked by the kernel itself, pair by pair as it reads them
    (CTRFE_LOD_DEVICE: non-decreasing, inside the id buffer), and an offending sequence is pooled
    as empty and reported by ctx.check_ids().  A serving loop that reuses its I/O buffers binds
    them once (bind()) and then only launches; a bound id buffer is used up to its full length.
    mode="copy": explicit H2D copies of the ids, host-validated LoD upload, launch, D2H copy."""

    def __init__(self, W, n_slots: int, batch: int, max_ids_per_slot: int, pooltype="SUM", *, pad_value=0.0,
                 padding_idx=-1, use_cvm: Optional[bool] = True, cvm_offset: int = 2, mode: str = "zero_copy",
                 concat: bool = False, ctx: Optional[Context] = None):
        import torch
        self.sharded = W if _is_sharded(W) else None
        if self.sharded is not None:  # one logical table row-sharded over the GPUs of the box
            t0 = torch.empty((0, W.dim), dtype=W.dtype, device=W.device)
            tables = [_TableRef(None, W.n_rows)] * n_slots
            self._dev = W.device
        else:
            tl = _as_list(W, n_slots)
            t0 = tl[0]
            tables = [as_arr(t, "W") for t in tl]
            self._dev = t0.device
        if mode not in ("zero_copy", "copy"):
            raise InvalidArgument("mode must be 'zero_copy' or 'copy'")
        self.mode, self.concat = mode, concat
        self._dtype = t0.dtype
        self.tables, self.S, self.B = tables, n_slots, batch
        self.pools = _as_list(pooltype, n_slots)
        self.kw = dict(pad_value=pad_value, padding_idx=padding_idx, use_cvm=use_cvm, cvm_offset=cvm_offset)
        self.ctx = ctx or default_context(t0.device)
        self.dim = t0.shape[1]
        self.d_out = t0.shape[1] - (cvm_offset if use_cvm is False else 0)
        shape = (batch, n_slots * self.d_out) if concat else (n_slots, batch, self.d_out)
        self.host_out = torch.empty(shape, dtype=t0.dtype, pin_memory=True)
        if mode == "copy":
            self.dev_ids = [torch.empty(max_ids_per_slot, dtype=torch.int64, device=t0.device) for _ in range(n_slots)]
            self.dev_out = torch.empty(shape, dtype=t0.dtype, device=t0.device)
            self.dev_out_cat = None if concat else torch.empty((batch, n_slots * self.d_out), dtype=t0.dtype, device=t0.device)
        self.h2d_bytes = 0
        self.d2h_bytes = self.host_out.numel() * self.host_out.element_size()

    def _result(self):
        return self.host_out if self.concat else [self.host_out[s] for s in range(self.S)]

    def _out_views(self, block):
        if self.concat:
            return [block[:, s * self.d_out:(s + 1) * self.d_out] for s in range(self.S)], self.S * self.d_out
        return [block[s] for s in range(self.S)], self.d_out

    def _desc(self, ids_host, lods_host):
        import torch
        for t in list(ids_host) + list(lods_host):
            if not (isinstance(t, torch.Tensor) and t.is_pinned()):
                raise InvalidArgument("zero_copy mode needs pinned host tensors for ids and LoD offsets")
        S = self.S
        use_cvm = self.kw["use_cvm"]
        cvm_mode = B.CVM_NONE if use_cvm is None else (B.CVM_USE if use_cvm else B.CVM_STRIP)
        outs, ld = self._out_views(self.host_out)
        slots = (B.Slot * S)()
        for s in range(S):
            t = self.tables[s]
            slots[s] = B.Slot(t.ptr, t.shape[0], ids_host[s].data_ptr(), lods_host[s].data_ptr(),
                              lods_host[s].numel() - 1, outs[s].data_ptr(), ld, B.POOLTYPES[self.pools[s]],
                              _TORCH_IDT()[ids_host[s].dtype], ids_host[s].numel())
        desc = B.FusedDesc(S, _TORCH_DT()[self._dtype], slots, self.dim, int(self.kw["padding_idx"]),
                           float(self.kw["pad_value"]), cvm_mode, self.kw["cvm_offset"], B.SCALE_DIVIDE,
                           B.INIT_ZERO, B.LOD_DEVICE, 0)
        if self.sharded is not None:
            desc.n_shards = self.sharded.n_shards
            desc.shard_tables = C.cast(self.sharded.shard_ptrs, C.POINTER(C.c_void_p))
            if getattr(self.sharded, "hot", None) is not None:
                desc.hot_cache = C.pointer(self.sharded.hot.struct)
        h2d = sum(i.numel() * i.element_size() for i in ids_host) + sum(l.numel() * 8 for l in lods_host)
        return desc, slots, h2d

    def bind(self, ids_host: Sequence[torch.Tensor], lods_host: Sequence[torch.Tensor]) -> "BoundHostCall":
        """zero_copy only: fix the (pinned) input buffers once; each BoundHostCall.run() is then
        one kernel launch + one stream synchronise, whatever the buffers currently hold."""
        if self.mode != "zero_copy":
            raise InvalidArgument("bind() needs mode='zero_copy'")
        import torch
        desc, slots, h2d = self._desc(ids_host, lods_host)
        h = C.c_void_p()
        stream = torch.cuda.current_stream(self._dev)
        B.check(B.lib().ctrfe_plan_create(self.ctx.handle, C.byref(desc), stream.cuda_stream, C.byref(h)))
        return BoundHostCall(self, FusedPlan(self.ctx, h, None, (ids_host, lods_host, slots)), h2d)

    def run(self, ids_host: Sequence[torch.Tensor], lods_host: Sequence, check: bool = False):
        """ids_host: pinned int64/int32 CPU tensors, one per slot.  lods_host: per slot the
        offsets -- numpy uint64 arrays (mode="copy") or pinned int64 CPU tensors (zero_copy)."""
        import torch
        S = self.S
        stream = torch.cuda.current_stream(self._dev)
        if self.mode == "copy":
            ids_dev = []
            for s in range(S):
                d = self.dev_ids[s][:ids_host[s].numel()]
                d.copy_(ids_host[s], non_blocking=True)
                ids_dev.append(d)
            ctr_fused_emb_seqpool_cvm(self.sharded if self.sharded is not None else [t.obj for t in self.tables], ids_dev, lods_host,
                                      self.pools, pad_value=self.kw["pad_value"], padding_idx=self.kw["padding_idx"],
                                      use_cvm=self.kw["use_cvm"], cvm_offset=self.kw["cvm_offset"], concat=True, check=False,
                                      out=self.dev_out if self.concat else self.dev_out_cat, ctx=self.ctx)
            if not self.concat:  # per-slot blocks: un-concatenate on the device
                self.dev_out.copy_(self.dev_out_cat.view(self.B, S, self.d_out).permute(1, 0, 2))
            self.host_out.copy_(self.dev_out, non_blocking=True)
            self.h2d_bytes = sum(i.numel() * i.element_size() for i in ids_host) + sum(l.size * 8 for l in lods_host)
        else:
            desc, slots, self.h2d_bytes = self._desc(ids_host, lods_host)
            B.check(B.lib().ctrfe_fused_forward(self.ctx.handle, C.byref(desc), stream.cuda_stream))
        if check:
            self.ctx.check_ids()  # synchronises
        else:
            stream.synchronize()
        return self._result()


class BoundHostCall:
    """HostFrontEnd with its input buffers bound: run() = one launch + one synchronise."""

    def __init__(self, fe: HostFrontEnd, plan: FusedPlan, h2d_bytes: int):
        import torch
        self.fe, self.plan, self.h2d_bytes, self.d2h_bytes = fe, plan, h2d_bytes, fe.d2h_bytes
        self._stream = torch.cuda.current_stream(fe._dev)

    def run(self):
        self.plan.launch()
        self._stream.synchronize()
        return self.fe._result()

    def submit(self, overlap_previous: bool = False):
        """Asynchronous half of run(): enqueue the launch and return a CUDA event that fires when
        the pooled rows are in the pinned host output.  A serving loop that keeps two requests
        in flight (each bound to its own HostFrontEnd, hence its own output block) submits
        request i+1 before waiting for request i, so the PCIe reads of one overlap the PCIe
        writes of the other; overlap_previous=True additionally lets the kernel start while the
        previous request's kernel drains (the requests are independent)."""
        import torch
        self.plan.launch(overlap_previous)
        done = torch.cuda.Event()
        done.record(self._stream)
        return done

    def result(self, done):
        done.synchronize()
        return self.fe._result()


class HostPipe:
    """The fused front end with HOST inputs and outputs through the library's pipelined host-buffer
    call (ctrfe_host_pipe_*): per slot the ids and LoD offsets in (pinned) host memory in, the
    pooled rows in pinned host memory out.  The shape of the reference's feed / fetch path --
    ZeroCopyTensor::CopyFromCpu + SetLoD per slot, CopyToCpu on the result
    (inference/api/details/zero_copy_tensor.cc) -- around the fused launch.  The slots are cut into
    `n_groups` groups and pipelined over both copy engines (group g goes up while group g-1 is
    pooled and group g-2's rows come down), so both directions of PCIe are busy at once; the LoD
    offsets are validated on the host like any host LoD.  Needs no torch: ids are numpy arrays
    (ideally views of `device.PinnedArray`s) or CPU tensors; run() returns one [B, D_out] numpy view
    of the pinned output block per slot (torch views when W is a torch tensor)."""

    def __init__(self, W, n_slots: int, batch: int, max_ids_per_slot: int, pooltype="SUM", *, pad_value=0.0,
                 padding_idx=-1, use_cvm: Optional[bool] = True, cvm_offset: int = 2, n_groups: int = 4,
                 ids_dtype="int64", direct_out: bool = False, direct_in: bool = False, ramp: bool = False,
                 ctx: Optional[Context] = None):
        """direct_out: the kernels store the pooled rows straight into the pinned host block (PCIe writes
        issued by the SMs) instead of a device block that a copy engine brings down.
        direct_in: the kernels read ids and offsets from the caller's PINNED host arrays themselves
        (`device.PinnedArray`, pinned torch tensors); nothing is uploaded.  Pageable arrays are refused.
        ramp: group sizes grow 1 : 2 : 4 : ... (a small first group starts the download engine early)."""
        sharded = W if _is_sharded(W) else None
        if sharded is not None:
            t0 = _sharded_like(sharded)
            tables = [_TableRef(None, W.n_rows)] * n_slots
        else:
            tables = [as_arr(t, "W") for t in _as_list(W, n_slots)]
            t0 = tables[0]
        ids_dtype = str(ids_dtype).replace("torch.", "")
        self.S, self.B, self.dim = n_slots, batch, t0.shape[1]
        self.ctx = ctx or default_context(t0.device)
        pools = _as_list(pooltype, n_slots)
        cvm_mode = B.CVM_NONE if use_cvm is None else (B.CVM_USE if use_cvm else B.CVM_STRIP)
        self.d_out = self.dim - (cvm_offset if use_cvm is False else 0)
        slots = (B.Slot * n_slots)()
        for s in range(n_slots):
            slots[s] = B.Slot(tables[s].ptr, tables[s].shape[0], None, None, batch, None, self.d_out,
                              B.POOLTYPES[pools[s]], _IDS[ids_dtype], 0)
        desc = B.FusedDesc(n_slots, t0.code, slots, self.dim, int(padding_idx), float(pad_value), cvm_mode, cvm_offset,
                           B.SCALE_DIVIDE, B.INIT_ZERO, B.LOD_HOST, 0)
        if sharded is not None:
            desc.n_shards = sharded.n_shards
            desc.shard_tables = C.cast(sharded.shard_ptrs, C.POINTER(C.c_void_p))
            if getattr(sharded, "hot", None) is not None:
                desc.hot_cache = C.pointer(sharded.hot.struct)
        self._keep = (tables, sharded)
        h = C.c_void_p()
        flags = (B.PIPE_DIRECT_OUT if direct_out else 0) | (B.PIPE_DIRECT_IN if direct_in else 0) | (B.PIPE_RAMP if ramp else 0)
        B.check(B.lib().ctrfe_host_pipe_create_ex(self.ctx.handle, C.byref(desc), int(max_ids_per_slot), int(n_groups),
                                                  flags, C.byref(h)))
        self.direct_in = direct_in
        self.handle = h
        np_dt = {"float32": np.float32, "float16": np.float16, "bfloat16": np.uint16}[t0.dtype]
        self._pinned = PinnedArray((n_slots, batch, self.d_out), np_dt)
        self.host_out = self._pinned.array
        if t0.kind == "torch":
            import torch
            ht = torch.from_numpy(self.host_out)
            self.host_out = ht.view(torch.bfloat16) if t0.dtype == "bfloat16" else ht
        row_bytes = batch * self.d_out * t0.itemsize
        self._out_ptrs = (C.c_void_p * n_slots)(*[self._pinned.ptr + s * row_bytes for s in range(n_slots)])
        self._ids_ptrs = (C.c_void_p * n_slots)()
        self._lod_ptrs = (C.c_void_p * n_slots)()
        self._n_ids = (C.c_int64 * n_slots)()
        self._id_es = 8 if ids_dtype == "int64" else 4
        self.d2h_bytes = n_slots * row_bytes
        self.h2d_bytes = 0

    def prepare(self, ids_host: Sequence, lods_host: Sequence):
        """The pointer tables of one request, built once for buffers that are reused (a serving loop's
        feed buffers): run_prepared() then is one call into the library.  Slots whose id arrays
        follow each other in host memory (views of one feed buffer) go up in one copy per group."""
        S = self.S
        ids_ptrs, lod_ptrs, n_ids = (C.c_void_p * S)(), (C.c_void_p * S)(), (C.c_int64 * S)()
        keep, h2d = [ids_host, lods_host], 0
        for s in range(S):
            i, l = ids_host[s], lods_host[s]
            if isinstance(i, np.ndarray):
                ids_ptrs[s], n_ids[s] = i.ctypes.data, i.size
            else:
                ids_ptrs[s], n_ids[s] = i.data_ptr(), i.numel()
            if isinstance(l, np.ndarray) or _is_host_lod(l):
                a = _lod_np(l)
                keep.append(a)
                lod_ptrs[s] = a.ctypes.data
                n = a.size
            else:
                lod_ptrs[s] = l.data_ptr()
                n = l.numel()
            if n != self.B + 1:
                raise InvalidArgument("Batchsize of all inputs should be equal.")
            h2d += n_ids[s] * self._id_es + n * 8
        return (ids_ptrs, n_ids, lod_ptrs, keep, h2d)

    def run_prepared(self, prep, wait=True):
        """wait=True: returns with the rows on the host; False: the context's stream waits for them;
        None: neither -- call wait() before reading host_out (two pipes used in turn keep two requests
        in flight: one goes up and is pooled while the other's rows come down)."""
        ids_ptrs, n_ids, lod_ptrs, _keep, self.h2d_bytes = prep
        self._in_flight = prep   # the request's arrays stay alive until the next run replaces them
        B.check(B.lib().ctrfe_host_pipe_run(self.handle, ids_ptrs, n_ids, lod_ptrs, self._out_ptrs, self.d_out,
                                            self.ctx.stream(), -1 if wait is None else (1 if wait else 0)))
        return [self.host_out[s] for s in range(self.S)]

    def wait(self):
        B.check(B.lib().ctrfe_host_pipe_wait(self.handle))
        return [self.host_out[s] for s in range(self.S)]

    def run(self, ids_host: Sequence, lods_host: Sequence, wait=True):
        """ids_host[s]: int64 (or int32) numpy array or CPU tensor, pinned for asynchronous copies;
        lods_host[s]: the [B + 1] offsets as a numpy uint64 array or an int64 CPU tensor."""
        return self.run_prepared(self.prepare(ids_host, lods_host), wait)

    def close(self):
        if getattr(self, "handle", None):
            B.lib().ctrfe_host_pipe_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
