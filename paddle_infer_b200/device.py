"""Device arrays without PyTorch: memory from the library's own ctrfe_malloc, moved with
ctrfe_memcpy, described to the ops by (pointer, shape, dtype).

The ops in `paddle_infer_b200.ops` accept three kinds of array and never import torch unless they
are handed a torch tensor:
  * `DeviceArray` (this module): numpy in, numpy out, nothing else needed;
  * anything with `__cuda_array_interface__` (CuPy, Numba, ...);
  * torch CUDA tensors (then outputs are torch tensors and the launches go to torch's current stream).
bfloat16 has no numpy dtype: a bf16 DeviceArray travels as uint16 bit patterns with `bf16=True`.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import numpy as np

from . import _lib as B

# dtype names used across the package -> (ctrfe code or None, element size, numpy dtype)
_DTYPES = {
    "float32": (B.F32, 4, np.float32), "float16": (B.F16, 2, np.float16), "bfloat16": (B.BF16, 2, np.uint16),
    "float64": (B.F64, 8, np.float64), "int64": (None, 8, np.int64), "int32": (None, 4, np.int32),
    "uint64": (None, 8, np.uint64), "uint32": (None, 4, np.uint32), "uint8": (None, 1, np.uint8), "int16": (None, 2, np.int16),
}
_TYPESTR = {"float32": "<f4", "float16": "<f2", "bfloat16": "<u2", "float64": "<f8", "int64": "<i8", "int32": "<i4",
            "uint64": "<u8", "uint32": "<u4", "uint8": "|u1", "int16": "<i2"}
_FROM_TYPESTR = {"<f4": "float32", "<f2": "float16", "<f8": "float64", "<i8": "int64", "<i4": "int32", "<u8": "uint64",
                 "<u4": "uint32", "<u2": "bfloat16", "|u1": "uint8", "<i2": "int16"}


def current_device() -> int:
    return 0


class DeviceArray:
    """A contiguous array in the memory of one GPU, owned by this object (or a view of one)."""

    def __init__(self, shape: Sequence[int], dtype: str = "float32", device: int = 0, *, _ptr: Optional[int] = None,
                 _base=None):
        if dtype not in _DTYPES:
            raise TypeError(f"unsupported dtype {dtype}")
        self.shape: Tuple[int, ...] = tuple(int(x) for x in shape)
        self.dtype, self.device = dtype, int(device)
        self.itemsize = _DTYPES[dtype][1]
        self.size = int(np.prod(self.shape)) if self.shape else 1
        self.nbytes = self.size * self.itemsize
        self._base = _base
        if _ptr is not None:
            self.ptr = int(_ptr)
        else:
            p = C.c_void_p()
            B.check(B.lib().ctrfe_malloc(self.device, max(self.nbytes, 1), C.byref(p)))
            self.ptr = int(p.value)

    # ---- construction / transfer ------------------------------------------------------------
    @classmethod
    def from_numpy(cls, a: np.ndarray, device: int = 0, bf16: bool = False, stream: int = 0) -> "DeviceArray":
        a = np.ascontiguousarray(a)
        name = "bfloat16" if bf16 else a.dtype.name
        if bf16 and a.dtype != np.uint16:
            raise TypeError("bf16 arrays travel as uint16 bit patterns")
        d = cls(a.shape, name, device)
        if a.nbytes:
            B.check(B.lib().ctrfe_memcpy(device, d.ptr, a.ctypes.data, a.nbytes, B.COPY_H2D, stream, 1))
        return d

    def numpy(self, stream: int = 0) -> np.ndarray:
        out = np.empty(self.shape, dtype=_DTYPES[self.dtype][2])
        if self.nbytes:
            B.check(B.lib().ctrfe_memcpy(self.device, out.ctypes.data, self.ptr, self.nbytes, B.COPY_D2H, stream, 1))
        return out

    def fill_bytes_from(self, a: np.ndarray, stream: int = 0) -> None:
        a = np.ascontiguousarray(a)
        if a.nbytes != self.nbytes:
            raise ValueError("size mismatch")
        B.check(B.lib().ctrfe_memcpy(self.device, self.ptr, a.ctypes.data, a.nbytes, B.COPY_H2D, stream, 1))

    def view(self, offset_elems: int, shape: Sequence[int]) -> "DeviceArray":
        """A view starting `offset_elems` elements in (no bounds beyond the obvious; no strides)."""
        return DeviceArray(shape, self.dtype, self.device, _ptr=self.ptr + offset_elems * self.itemsize, _base=self)

    @property
    def __cuda_array_interface__(self):
        return {"shape": self.shape, "typestr": _TYPESTR[self.dtype], "data": (self.ptr, False), "version": 2, "strides": None}

    def data_ptr(self) -> int:
        return self.ptr

    def close(self) -> None:
        if self._base is None and getattr(self, "ptr", 0):
            B.lib().ctrfe_free(self.device, self.ptr)
        self.ptr = 0

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PinnedArray:
    """Page-locked host memory (cudaHostAlloc) seen as a numpy array: what the host-buffer calls want
    for asynchronous copies."""

    def __init__(self, shape: Sequence[int], dtype=np.float32):
        dt = np.dtype(dtype)
        self.shape = tuple(int(x) for x in shape)
        n = int(np.prod(self.shape)) * dt.itemsize
        p = C.c_void_p()
        B.check(B.lib().ctrfe_host_alloc(max(n, 1), C.byref(p)))
        self.ptr = int(p.value)
        buf = (C.c_char * max(n, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dt, count=int(np.prod(self.shape))).reshape(self.shape)

    def close(self):
        if getattr(self, "ptr", 0):
            self.array = None
            B.lib().ctrfe_host_free(self.ptr)
            self.ptr = 0

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---- what the ops see ---------------------------------------------------------------------------
class Arr:
    """(pointer, shape, dtype name, device, kind) of an array handed to an op; `obj` keeps it alive."""
    __slots__ = ("ptr", "shape", "dtype", "device", "kind", "obj")

    def __init__(self, ptr, shape, dtype, device, kind, obj):
        self.ptr, self.shape, self.dtype, self.device, self.kind, self.obj = ptr, tuple(shape), dtype, device, kind, obj

    @property
    def numel(self) -> int:
        return int(np.prod(self.shape)) if self.shape else 1

    @property
    def itemsize(self) -> int:
        return _DTYPES[self.dtype][1]

    @property
    def code(self):
        return _DTYPES[self.dtype][0]


def is_torch(x) -> bool:
    return type(x).__module__.split(".")[0] == "torch"


def as_arr(x, name: str = "input") -> Arr:
    """Any supported device array -> Arr.  Raises RuntimeError for host data (no CPU fallback)."""
    if isinstance(x, Arr):
        return x
    if isinstance(x, DeviceArray):
        return Arr(x.ptr, x.shape, x.dtype, x.device, "raw", x)
    if is_torch(x):
        if not x.is_cuda:
            raise RuntimeError(f"{name} must live on the GPU: this op has no CPU kernel (no CPU fallback)")
        x = x.contiguous()
        return Arr(x.data_ptr(), tuple(x.shape), str(x.dtype).replace("torch.", ""), x.device.index, "torch", x)
    cai = getattr(x, "__cuda_array_interface__", None)
    if cai is not None:
        if cai.get("strides") is not None:
            raise ValueError(f"{name}: only contiguous arrays are supported")
        dt = _FROM_TYPESTR.get(cai["typestr"])
        if dt is None:
            raise TypeError(f"{name}: unsupported typestr {cai['typestr']}")
        return Arr(int(cai["data"][0]), tuple(cai["shape"]), dt, getattr(x, "device", 0) if isinstance(getattr(x, "device", 0), int) else 0,
                   "raw", x)
    raise RuntimeError(f"{name} must live on the GPU: this op has no CPU kernel (no CPU fallback)")


def empty(shape, dtype: str, like: Arr):
    """An uninitialised output of the same kind as `like`: a torch tensor beside torch inputs, else a
    DeviceArray.  Returns (object to hand back, Arr)."""
    if like.kind == "torch":
        import torch
        t = torch.empty(tuple(shape), dtype=getattr(torch, dtype), device=like.obj.device)
        return t, Arr(t.data_ptr(), tuple(shape), dtype, like.device, "torch", t)
    d = DeviceArray(shape, dtype, like.device)
    return d, Arr(d.ptr, d.shape, dtype, like.device, "raw", d)


def stream_of(like: Arr) -> int:
    """The stream launches go to: torch's current stream beside torch inputs, else the legacy default."""
    if like.kind == "torch":
        import torch
        return torch.cuda.current_stream(like.obj.device).cuda_stream
    return 0
