"""ctypes binding of libctrfe.so (include/ctrfe.h).  Fails loudly when the library is
absent -- there is deliberately no pure-Python or CPU substitute."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CTRFE_LIB") or os.path.join(_HERE, "libctrfe.so")  # CTRFE_LIB: tuning builds

# enums of include/ctrfe.h
OK, ERR_BAD_ID, ERR_ARG, ERR_CUDA, ERR_UNSUPPORTED, ERR_NCCL = range(6)
F32, F16, BF16, F64 = 0, 1, 2, 3
I64, I32, I16 = 0, 1, 2
POOL_AVERAGE, POOL_SUM, POOL_SQRT, POOL_LAST, POOL_FIRST, POOL_MAX = range(6)
POOLTYPES = {"AVERAGE": POOL_AVERAGE, "SUM": POOL_SUM, "SQRT": POOL_SQRT, "LAST": POOL_LAST,
             "FIRST": POOL_FIRST, "MAX": POOL_MAX}
CVM_NONE, CVM_USE, CVM_STRIP = 0, 1, 2
SCALE_DIVIDE, SCALE_RECIPROCAL = 0, 1
INIT_ZERO, INIT_PAD, INIT_FIRST_ROW = 0, 1, 2
LOD_HOST, LOD_DEVICE = 0, 1


class Slot(C.Structure):
    _fields_ = [("table", C.c_void_p), ("table_rows", C.c_int64), ("ids", C.c_void_p),
                ("lod", C.c_void_p), ("n_seg", C.c_int64), ("out", C.c_void_p),
                ("out_ld", C.c_int64), ("pooltype", C.c_int32), ("ids_dtype", C.c_int32), ("n_ids", C.c_int64)]


class HotCache(C.Structure):
    _fields_ = [("rows", C.c_void_p), ("keys", C.c_void_p), ("log2_slots", C.c_int32),
                ("n_rows", C.c_int32)]


class KeyMap(C.Structure):
    _fields_ = [("keys", C.c_void_p), ("rows", C.c_void_p), ("log2_slots", C.c_int32), ("reserved", C.c_int32),
                ("n_keys", C.c_int64)]


class FusedDesc(C.Structure):
    _fields_ = [("n_slots", C.c_int32), ("dtype", C.c_int32), ("slots", C.POINTER(Slot)),
                ("dim", C.c_int64), ("padding_idx", C.c_int64), ("pad_value", C.c_float),
                ("cvm_mode", C.c_int32), ("cvm_offset", C.c_int32), ("scale_mode", C.c_int32),
                ("init_mode", C.c_int32), ("lod_location", C.c_int32), ("n_shards", C.c_int32),
                ("shard_tables", C.POINTER(C.c_void_p)), ("hot_cache", C.POINTER(HotCache)),
                ("keymap", C.POINTER(KeyMap))]


class CtrfeError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(msg)
        self.code = code


COPY_H2D, COPY_D2H, COPY_D2D = 1, 2, 3
LAUNCH_OVERLAP_PREVIOUS = 1
LAUNCH_UPLOAD_LOD = 2
PIPE_DIRECT_OUT, PIPE_DIRECT_IN, PIPE_RAMP = 1, 2, 4


class BadIdError(CtrfeError, ValueError):
    """InvalidArgument of the reference's embedding bounds check
    (paddle/phi/kernels/cpu/embedding_kernel.cc:59-76)."""

    def __init__(self, msg, slot=-1, position=-1, id_value=-1):
        super().__init__(ERR_BAD_ID, msg)
        self.slot, self.position, self.id_value = slot, position, id_value


_lib = None

# every symbol include/ctrfe.h declares (tests check the .so exports each one)
EXPORTS = [
    "ctrfe_version", "ctrfe_last_error", "ctrfe_malloc", "ctrfe_free", "ctrfe_host_alloc", "ctrfe_host_free",
    "ctrfe_memcpy", "ctrfe_stream_synchronize", "ctrfe_ctx_create", "ctrfe_ctx_destroy", "ctrfe_check_ids",
    "ctrfe_fused_forward", "ctrfe_plan_create", "ctrfe_plan_launch", "ctrfe_plan_launch_ex", "ctrfe_plan_destroy",
    "ctrfe_plan_num_kernels", "ctrfe_fused_embedding_seq_pool", "ctrfe_plan_launch_many", "ctrfe_plan_update_lod", "ctrfe_host_pipe_create", "ctrfe_host_pipe_create_ex", "ctrfe_host_pipe_wait",
    "ctrfe_host_pipe_run", "ctrfe_host_pipe_destroy", "ctrfe_plan_kernel_name", "ctrfe_embedding", "ctrfe_sequence_pool", "ctrfe_cvm",
    "ctrfe_fused_forward_static", "ctrfe_sequence_pool_grad", "ctrfe_cvm_grad", "ctrfe_embedding_grad", "ctrfe_embedding_grad_sparse",
    "ctrfe_shard_alloc", "ctrfe_shard_free", "ctrfe_shard_export", "ctrfe_shard_open", "ctrfe_shard_close",
    "ctrfe_shard_alloc_fd", "ctrfe_shard_open_fd", "ctrfe_shard_unmap", "ctrfe_hot_cache_build", "ctrfe_keymap_build",
]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). paddle_infer_b200 has no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
        L.ctrfe_version.restype = C.c_int
        L.ctrfe_last_error.restype = C.c_char_p
        L.ctrfe_malloc.argtypes = [C.c_int, C.c_uint64, C.POINTER(vp)]
        L.ctrfe_free.argtypes = [C.c_int, vp]
        L.ctrfe_host_alloc.argtypes = [C.c_uint64, C.POINTER(vp)]
        L.ctrfe_host_free.argtypes = [vp]
        L.ctrfe_memcpy.argtypes = [C.c_int, vp, vp, C.c_uint64, C.c_int, vp, C.c_int]
        L.ctrfe_stream_synchronize.argtypes = [C.c_int, vp]
        L.ctrfe_ctx_create.argtypes = [C.c_int, C.POINTER(vp)]
        L.ctrfe_ctx_destroy.argtypes = [vp]
        L.ctrfe_check_ids.argtypes = [vp, vp, C.POINTER(i64), C.POINTER(i64), C.POINTER(i64)]
        L.ctrfe_fused_forward.argtypes = [vp, C.POINTER(FusedDesc), vp]
        L.ctrfe_plan_create.argtypes = [vp, C.POINTER(FusedDesc), vp, C.POINTER(vp)]
        L.ctrfe_plan_launch.argtypes = [vp, vp]
        L.ctrfe_plan_launch_ex.argtypes = [vp, vp, C.c_uint32]
        L.ctrfe_plan_destroy.argtypes = [vp]
        L.ctrfe_plan_num_kernels.argtypes = [vp]
        L.ctrfe_plan_launch_many.argtypes = [C.POINTER(vp), C.c_int, vp, C.c_uint32]
        L.ctrfe_plan_update_lod.argtypes = [vp, C.POINTER(vp)]
        L.ctrfe_host_pipe_create.argtypes = [vp, C.POINTER(FusedDesc), i64, C.c_int, C.POINTER(vp)]
        L.ctrfe_host_pipe_create_ex.argtypes = [vp, C.POINTER(FusedDesc), i64, C.c_int, C.c_uint32, C.POINTER(vp)]
        L.ctrfe_host_pipe_run.argtypes = [vp, C.POINTER(vp), C.POINTER(i64), C.POINTER(vp), C.POINTER(vp), i64, vp, C.c_int]
        L.ctrfe_host_pipe_destroy.argtypes = [vp]
        L.ctrfe_host_pipe_wait.argtypes = [vp]
        L.ctrfe_plan_kernel_name.argtypes = [vp, C.c_int]
        L.ctrfe_plan_kernel_name.restype = C.c_char_p
        L.ctrfe_fused_embedding_seq_pool.argtypes = [vp, vp, i64, i64, i32, vp, i64, i64, vp, i64, i64, vp, vp]
        L.ctrfe_embedding.argtypes = [vp, vp, i64, i64, vp, i32, i64, i64, vp, vp]
        L.ctrfe_sequence_pool.argtypes = [vp, vp, i32, i64, vp, i64, i32, C.c_float, vp, vp, vp]
        L.ctrfe_cvm.argtypes = [vp, vp, i32, i64, i64, i32, vp, vp]
        L.ctrfe_sequence_pool_grad.argtypes = [vp, vp, i64, vp, i64, i64, i32, vp, vp, vp]
        L.ctrfe_cvm_grad.argtypes = [vp, vp, vp, vp, i64, i64, i64, i32, vp, vp]
        L.ctrfe_embedding_grad.argtypes = [vp, vp, vp, i32, i64, i64, i64, i64, vp, vp]
        L.ctrfe_embedding_grad_sparse.argtypes = [vp, vp, i64, vp, i32, i64, vp, vp, vp, vp]
        L.ctrfe_fused_forward_static.argtypes = [C.POINTER(FusedDesc), vp, vp]
        L.ctrfe_shard_alloc.argtypes = [C.c_int, C.c_uint64, C.POINTER(vp)]
        L.ctrfe_shard_free.argtypes = [C.c_int, vp]
        L.ctrfe_shard_export.argtypes = [vp, C.c_char_p]
        L.ctrfe_shard_open.argtypes = [C.c_int, C.c_char_p, C.POINTER(vp)]
        L.ctrfe_shard_close.argtypes = [C.c_int, vp]
        L.ctrfe_shard_alloc_fd.argtypes = [C.c_int, C.c_uint64, C.POINTER(vp), C.POINTER(C.c_int), C.POINTER(C.c_uint64)]
        L.ctrfe_shard_open_fd.argtypes = [C.c_int, C.c_int, C.c_uint64, C.POINTER(vp)]
        L.ctrfe_shard_unmap.argtypes = [C.c_int, vp, C.c_uint64]
        L.ctrfe_keymap_build.argtypes = [vp, vp, i64, vp, vp, i32, vp, vp]
        L.ctrfe_hot_cache_build.argtypes = [vp, vp, i64, vp, i64, vp, vp, i32, vp]
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != OK:
        raise CtrfeError(rc, lib().ctrfe_last_error().decode("utf-8", "replace"))
