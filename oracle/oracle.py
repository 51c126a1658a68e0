"""ctypes doorway onto the CPU oracle (oracle/ctr_oracle.c) and, when it was built,
the reference's own jit::refer kernels (oracle/_ref/libref_jit.so).

TEST INFRASTRUCTURE ONLY.  Importers allowed: tests/, __graft_entry__.smoke(),
bench.py's cpu_baseline / --impl reference legs.  The product package
(paddle_infer_b200/) must never import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_FAST_PATH = os.path.join(_HERE, "liboracle_avx2.so")
_REF_PATH = os.path.join(_HERE, "_ref", "libref_jit.so")

F32, F16, BF16, F64 = 0, 1, 2, 3
AVERAGE, SUM, SQRT, LAST, FIRST, MAX = 0, 1, 2, 3, 4, 5
POOLTYPES = {"AVERAGE": AVERAGE, "SUM": SUM, "SQRT": SQRT, "LAST": LAST, "FIRST": FIRST, "MAX": MAX}
CVM_NONE, CVM_USE, CVM_STRIP = 0, 1, 2
SCALE_DIVIDE, SCALE_RECIPROCAL = 0, 1
INIT_ZERO, INIT_PAD, INIT_FIRST_ROW = 0, 1, 2
LOG_DOUBLE, LOG_FLOAT = 0, 1
OK, ERR_BAD_ID, ERR_ARG = 0, 1, 2

_BASE_FLAGS = ["-std=c11", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off", "-fno-fast-math"]


def build(force: bool = False) -> None:
    """Compile the C restatement (generic x86-64 and an AVX2 build used only for the
    timed CPU baseline) and, when /root/reference is present, oracle/_ref."""
    src = os.path.join(_HERE, "ctr_oracle.c")
    for out, extra in ((_LIB_PATH, ["-O2"]), (_FAST_PATH, ["-O3", "-mavx2"])):
        if force or not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(src):
            subprocess.check_call(["gcc", *_BASE_FLAGS, *extra, "-o", out, src, "-lm"])
    ref_root = os.environ.get("CTRFE_REFERENCE_ROOT", "/root/reference")
    if os.path.isdir(os.path.join(ref_root, "paddle", "fluid", "operators", "jit", "refer")):
        subprocess.check_call(["make", "-s", "-C", os.path.join(_HERE, "ref_build"), f"REF={ref_root}"])


class Slot(C.Structure):
    _fields_ = [("table", C.c_void_p), ("table_rows", C.c_int64), ("ids", C.c_void_p),
                ("lod", C.c_void_p), ("n_seg", C.c_int64), ("out", C.c_void_p),
                ("out_ld", C.c_int64), ("pooltype", C.c_int32), ("ids_is_i32", C.c_int32)]


class FusedAttrs(C.Structure):
    _fields_ = [("dtype", C.c_int32), ("cvm_mode", C.c_int32), ("cvm_offset", C.c_int32),
                ("scale_mode", C.c_int32), ("init_mode", C.c_int32), ("log_mode", C.c_int32),
                ("dim", C.c_int64), ("padding_idx", C.c_int64), ("pad_value", C.c_float),
                ("n_threads", C.c_int32)]


def _cpu_has_avx2() -> bool:
    try:
        with open("/proc/cpuinfo") as f:
            return " avx2" in f.read()
    except OSError:
        return False


_libs = {}


def lib(fast: bool = False):
    key = "fast" if fast and _cpu_has_avx2() else "base"
    if key not in _libs:
        path = _FAST_PATH if key == "fast" else _LIB_PATH
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        L.orc_half_to_float.restype = C.c_float
        L.orc_half_to_float.argtypes = [C.c_uint16]
        L.orc_float_to_half.restype = C.c_uint16
        L.orc_float_to_half.argtypes = [C.c_float]
        L.orc_bf16_to_float.restype = C.c_float
        L.orc_bf16_to_float.argtypes = [C.c_uint16]
        L.orc_float_to_bf16.restype = C.c_uint16
        L.orc_float_to_bf16.argtypes = [C.c_float]
        L.orc_embedding.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int, C.c_int64,
                                    C.c_int64, C.c_void_p, C.POINTER(C.c_int64), C.c_int]
        L.orc_sequence_pool.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int,
                                        C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_sequence_pool_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int,
                                            C.c_double, C.c_int, C.c_void_p, C.c_void_p]
        L.orc_cvm.argtypes = [C.c_int, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_void_p,
                              C.c_int]
        L.orc_fused_forward.argtypes = [C.POINTER(Slot), C.c_int, C.POINTER(FusedAttrs),
                                        C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.orc_unfused_forward.argtypes = [C.POINTER(Slot), C.c_int, C.POINTER(FusedAttrs), C.c_void_p,
                                          C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.orc_sequence_pool_grad.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int,
                                             C.c_void_p, C.c_void_p]
        L.orc_cvm_grad.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int,
                                   C.c_void_p]
        L.orc_embedding_grad.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64,
                                         C.c_void_p]
        L.orc_embedding_grad.restype = C.c_int64
        L.orc_fused_seqpool_cvm.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_float,
                                            C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int64]
        L.orc_max_threads.restype = C.c_int
        _libs[key] = L
    return _libs[key]


def ref_lib():
    """The reference's own refer.h kernels, or None when oracle/_ref was never built."""
    if "ref" not in _libs:
        if not os.path.exists(_REF_PATH):
            _libs["ref"] = None
        else:
            R = C.CDLL(_REF_PATH)
            R.ref_seqpool_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
            R.ref_seqpool_f64.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
            R.ref_embseqpool_f32.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_int64,
                                             C.c_int64, C.c_void_p]
            R.ref_seqpool_lod_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_int,
                                              C.c_void_p, C.c_int64]
            R.ref_embseqpool_lod_f32.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                                 C.c_int64, C.c_void_p, C.c_int64]
            R.ref_fused_lod_f32.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64,
                                            C.c_float, C.c_int, C.c_void_p, C.c_int64, C.c_int]
            _libs["ref"] = R
    return _libs["ref"]


# --------------------------------------------------------------------------------------
# numpy-facing helpers.  16-bit dtypes travel as np.float16 (F16) or np.uint16 bit
# patterns (BF16); numpy has no bfloat16.
# --------------------------------------------------------------------------------------
class BadId(ValueError):
    """Mirrors the InvalidArgument the reference throws (embedding_kernel.cc:59-76)."""


def np_dtype_code(a: np.ndarray, bf16: bool = False) -> int:
    if bf16:
        assert a.dtype == np.uint16
        return BF16
    return {np.dtype(np.float32): F32, np.dtype(np.float16): F16, np.dtype(np.float64): F64}[a.dtype]


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def _lod_arr(lod: Sequence[int]) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(lod, dtype=np.uint64))


def embedding(table: np.ndarray, ids: np.ndarray, padding_idx: int = -1, threads: int = 1) -> np.ndarray:
    table = np.ascontiguousarray(table)
    ids = np.ascontiguousarray(ids)
    assert ids.dtype in (np.int64, np.int32)
    n_rows = table.shape[0]
    row_bytes = table.strides[0]
    out = np.empty(ids.shape + table.shape[1:], dtype=table.dtype)
    bad = C.c_int64(-1)
    rc = lib().orc_embedding(_ptr(table), n_rows, row_bytes, _ptr(ids), int(ids.dtype == np.int32),
                             ids.size, padding_idx, _ptr(out), C.byref(bad), threads)
    if rc == ERR_BAD_ID:
        v = int(ids.reshape(-1)[bad.value])
        raise BadId(f"Variable value (input) of OP(fluid.layers.embedding) expected >= 0 and < "
                    f"{n_rows}, but got {v}. Please check input value.")
    assert rc == OK
    return out


def sequence_pool(x: np.ndarray, lod: Sequence[int], pooltype: str, pad_value: float = 0.0,
                  is_test: bool = True, scale_mode: int = SCALE_DIVIDE, init_mode: int = INIT_ZERO,
                  bf16: bool = False, threads: int = 1):
    """x: [K, ...]; lod: offsets of the LAST LoD level.  Returns out or (out, max_index)."""
    x = np.ascontiguousarray(x)
    lod_a = _lod_arr(lod)
    n_seg = lod_a.size - 1
    w = int(np.prod(x.shape[1:])) if x.ndim > 1 else 1
    out = np.empty((n_seg,) + x.shape[1:], dtype=x.dtype)
    pt = POOLTYPES[pooltype]
    want_idx = pt == MAX and not is_test
    idx = np.zeros(out.shape, dtype=np.int32) if want_idx else None
    if x.dtype == np.float64:
        rc = lib().orc_sequence_pool_f64(_ptr(x), _ptr(lod_a), n_seg, w, pt, pad_value, scale_mode,
                                         _ptr(out), _ptr(idx) if want_idx else None)
    else:
        rc = lib().orc_sequence_pool(np_dtype_code(x, bf16), _ptr(x), _ptr(lod_a), n_seg, w, pt, pad_value,
                                     scale_mode, init_mode, _ptr(out), _ptr(idx) if want_idx else None,
                                     threads)
    assert rc == OK, rc
    return (out, idx) if want_idx else out


def cvm(x: np.ndarray, use_cvm: bool = True, log_mode: int = LOG_DOUBLE, bf16: bool = False,
        threads: int = 1) -> np.ndarray:
    x = np.ascontiguousarray(x)
    rows, w = x.shape
    y = np.empty((rows, w if use_cvm else w - 2), dtype=x.dtype)
    rc = lib().orc_cvm(np_dtype_code(x, bf16), _ptr(x), rows, w, int(use_cvm), log_mode, _ptr(y), threads)
    assert rc == OK, rc
    return y


def fused_seqpool_cvm(x: np.ndarray, lod: Sequence[int], pad_value: float = 0.0, use_cvm: bool = True,
                      cvm_offset: int = 2, log_mode: int = LOG_FLOAT) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32)
    lod_a = _lod_arr(lod)
    n_seg, w = lod_a.size - 1, x.shape[1]
    wo = w if use_cvm else w - cvm_offset
    out = np.empty((n_seg, wo), dtype=np.float32)
    rc = lib().orc_fused_seqpool_cvm(_ptr(x), _ptr(lod_a), n_seg, w, pad_value, int(use_cvm), cvm_offset,
                                     log_mode, _ptr(out), wo)
    assert rc == OK
    return out


def fused_forward(tables: Sequence[np.ndarray], ids: Sequence[np.ndarray], lods: Sequence[Sequence[int]],
                  pooltypes: Sequence[str], *, pad_value: float = 0.0, padding_idx: int = -1,
                  cvm_mode: int = CVM_NONE, cvm_offset: int = 2, scale_mode: int = SCALE_DIVIDE,
                  init_mode: int = INIT_ZERO, log_mode: int = LOG_DOUBLE, bf16: bool = False,
                  concat: bool = False, threads: int = 1, unfused: bool = False, fast: bool = False):
    """Multi-slot gather -> pool -> CVM.  Returns a list of [n_seg, D_out] arrays, or one
    [n_seg, S*D_out] array when concat=True (fusion_seqpool_cvm_concat layout)."""
    S = len(ids)
    assert len(tables) in (1, S)
    tabs = [np.ascontiguousarray(t) for t in tables]
    dim = tabs[0].shape[1]
    dt = np_dtype_code(tabs[0], bf16)
    d_out = dim - cvm_offset if cvm_mode == CVM_STRIP else dim
    lod_as = [_lod_arr(l) for l in lods]
    ids_as = [np.ascontiguousarray(i) for i in ids]
    n_seg = lod_as[0].size - 1
    if concat:
        assert all(l.size - 1 == n_seg for l in lod_as)
        big = np.zeros((n_seg, S * d_out), dtype=tabs[0].dtype)
        outs = [big[:, s * d_out:(s + 1) * d_out] for s in range(S)]
        ld = S * d_out
    else:
        outs = [np.zeros((l.size - 1, d_out), dtype=tabs[0].dtype) for l in lod_as]
        ld = d_out
    slots = (Slot * S)()
    for s in range(S):
        t = tabs[s if len(tabs) > 1 else 0]
        slots[s] = Slot(t.ctypes.data, t.shape[0], ids_as[s].ctypes.data, lod_as[s].ctypes.data,
                        lod_as[s].size - 1, outs[s].ctypes.data, ld, POOLTYPES[pooltypes[s]],
                        int(ids_as[s].dtype == np.int32))
    attrs = FusedAttrs(dt, cvm_mode, cvm_offset, scale_mode, init_mode, log_mode, dim, padding_idx,
                       pad_value, threads)
    bs, bp = C.c_int64(-1), C.c_int64(-1)
    L = lib(fast)
    if unfused:
        max_k = max(int(l[-1] - l[0]) for l in lod_as)
        max_seg = max(l.size - 1 for l in lod_as)
        sr = np.empty((max(max_k, 1), dim), dtype=tabs[0].dtype)
        sp = np.empty((max(max_seg, 1), dim), dtype=tabs[0].dtype)
        rc = L.orc_unfused_forward(slots, S, C.byref(attrs), _ptr(sr), _ptr(sp), C.byref(bs), C.byref(bp))
    else:
        rc = L.orc_fused_forward(slots, S, C.byref(attrs), C.byref(bs), C.byref(bp))
    if rc == ERR_BAD_ID:
        raise BadId(f"slot {bs.value}: id at position {bp.value} outside [0, table_rows)")
    assert rc == OK, rc
    return big if concat else outs


# bf16 helpers for tests (numpy has no bfloat16): bit patterns in uint16
def sequence_pool_grad(dout: np.ndarray, lod: Sequence[int], pooltype: str, k_rows: Optional[int] = None,
                       max_index: Optional[np.ndarray] = None) -> np.ndarray:
    """X@GRAD [K, w] of sequence_pool from Out@GRAD [n_seg, w] (fp32); see orc_sequence_pool_grad."""
    dout = np.ascontiguousarray(dout, np.float32)
    lod_a = _lod_arr(lod)
    n_seg = lod_a.size - 1
    w = int(np.prod(dout.shape[1:]))
    k = int(lod_a[-1]) if k_rows is None else int(k_rows)
    dx = np.empty((k,) + dout.shape[1:], np.float32)
    idx = None if max_index is None else np.ascontiguousarray(max_index, np.int32)
    rc = lib().orc_sequence_pool_grad(_ptr(dout), _ptr(lod_a), n_seg, w, k, POOLTYPES[pooltype],
                                      _ptr(idx) if idx is not None else None, _ptr(dx))
    assert rc == OK, rc
    return dx


def cvm_grad(dy: np.ndarray, cvm_in: np.ndarray, use_cvm: bool = True, lod: Optional[Sequence[int]] = None) -> np.ndarray:
    """X@GRAD of cvm from Y@GRAD and the CVM input [n_seq, 2]; see orc_cvm_grad."""
    dy = np.ascontiguousarray(dy, np.float32)
    cvm_in = np.ascontiguousarray(cvm_in, np.float32)
    rows = dy.shape[0]
    w = dy.shape[1] + (0 if use_cvm else 2)
    dx = np.empty((rows, w), np.float32)
    lod_a = None if lod is None else _lod_arr(lod)
    rc = lib().orc_cvm_grad(_ptr(dy), _ptr(cvm_in), _ptr(lod_a) if lod_a is not None else None,
                            cvm_in.shape[0], rows, w, int(use_cvm), _ptr(dx))
    assert rc == OK, rc
    return dx


def embedding_grad(dout: np.ndarray, ids: np.ndarray, n_rows: int, padding_idx: int = -1) -> np.ndarray:
    """Dense W@GRAD [n_rows, d] of lookup_table_v2; raises IndexError on an id outside the table."""
    ids = np.ascontiguousarray(ids, np.int64).reshape(-1)
    dout = np.ascontiguousarray(dout, np.float32).reshape(ids.size, -1)
    d = dout.shape[1]
    dw = np.empty((n_rows, d), np.float32)
    rc = lib().orc_embedding_grad(_ptr(dout), _ptr(ids), ids.size, n_rows, d, padding_idx, _ptr(dw))
    if rc < 0:
        raise IndexError(f"id at position {-rc - 1} is outside the table")
    return dw


def pull_sparse(keys: np.ndarray, values: np.ndarray, query: np.ndarray) -> np.ndarray:
    """pull_box_sparse / pull_gpups_sparse restated (operators/pull_box_sparse_op.h:32-62 ->
    PSGPUWrapper::PullSparse -> HashTable::get + PullCopy): out[k] = values[i] where
    keys[i] == query[k]; key 0 gives a zero row ("0 key fill zero",
    framework/fleet/heter_ps/feature_value.cu:76-80, feature_value.h:498-514); a key the table
    does not hold gives a zero row too (hashtable_kernel.cu:79-82 leaves the value untouched;
    zeros are this repo's definition).  keys / query: uint64 (or int64 views of them)."""
    rows = pull_sparse_rows(keys, query, zero_row=values.shape[0])
    padded = np.concatenate([values, np.zeros((1, values.shape[1]), values.dtype)])
    return padded[rows]


def pull_sparse_rows(keys: np.ndarray, query: np.ndarray, zero_row: int) -> np.ndarray:
    """Row index of every queried feasign key (see pull_sparse); `zero_row` for key 0 and for
    unknown keys."""
    keys = np.ascontiguousarray(keys).view(np.uint64).reshape(-1)
    query = np.ascontiguousarray(query).view(np.uint64).reshape(-1)
    order = np.argsort(keys, kind="stable")
    sk = keys[order]
    pos = np.searchsorted(sk, query)
    pos_c = np.minimum(pos, max(sk.size - 1, 0))
    hit = (sk.size > 0) & (sk[pos_c] == query) & (query != 0) if sk.size else np.zeros(query.shape, bool)
    rows = np.full(query.shape, zero_row, np.int64)
    rows[hit] = order[pos_c[hit]]
    return rows


def f32_to_bf16_bits(a: np.ndarray) -> np.ndarray:
    """RNE float32 -> bfloat16 bit patterns (vectorised twin of orc_float_to_bf16)."""
    x = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    nan = (x & 0x7FFFFFFF) > 0x7F800000
    r = ((x + (0x7FFF + ((x >> 16) & 1))) >> 16).astype(np.uint16)
    r[nan] = ((x[nan] >> 16) | 0x40).astype(np.uint16)
    return r


def bf16_bits_to_f32(a: np.ndarray) -> np.ndarray:
    return (a.astype(np.uint32) << 16).view(np.float32)
