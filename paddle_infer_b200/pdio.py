"""Reader / writer for Paddle's persistable-tensor stream (`*.pdiparams`, `save_combine`
files, single-variable `save` files), so an embedding table W saved by Paddle can be put
straight into device memory -- whole, or only this rank's rows of a row-sharded table --
without Paddle installed.

Format, one record per tensor, records simply concatenated
(paddle/phi/core/serialization.cc:26-54, paddle/fluid/framework/tensor_util.cc:654-676):

    u32  LoDTensor version (0)
    u64  number of LoD levels;  per level: u64 byte size, then that many bytes of u64 offsets
    u32  Tensor version (0)
    i32  size of the TensorDesc protobuf;  the protobuf (framework.proto VarType.TensorDesc:
         `required Type data_type = 1; repeated int64 dims = 2;`)
    raw  numel * sizeof(dtype) bytes, row-major

Only what this path needs is implemented: fp32 / fp16 / bf16 / fp64 / int32 / int64 tensors.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass
from typing import BinaryIO, Iterator, List, Optional, Sequence, Tuple

import numpy as np

# framework.proto VarType.Type values (paddle/fluid/framework/framework.proto:118-131)
_PD_TYPES = {2: ("int32", 4), 3: ("int64", 8), 4: ("float16", 2), 5: ("float32", 4), 6: ("float64", 8),
             22: ("bfloat16", 2)}
_PD_CODES = {"int32": 2, "int64": 3, "float16": 4, "float32": 5, "float64": 6, "bfloat16": 22}


# ---- the few bytes of protobuf wire format a TensorDesc needs --------------------------------
def _read_varint(b: bytes, i: int) -> Tuple[int, int]:
    v, shift = 0, 0
    while True:
        c = b[i]
        i += 1
        v |= (c & 0x7F) << shift
        if not c & 0x80:
            return v, i
        shift += 7


def _write_varint(v: int) -> bytes:
    v &= (1 << 64) - 1  # negative int64 -> 10-byte two's complement, as protobuf does
    out = bytearray()
    while True:
        c = v & 0x7F
        v >>= 7
        out.append(c | (0x80 if v else 0))
        if not v:
            return bytes(out)


def parse_tensor_desc(b: bytes) -> Tuple[int, List[int]]:
    """-> (data_type code, dims).  Accepts packed and unpacked `dims`."""
    i, dtype, dims = 0, None, []
    while i < len(b):
        tag, i = _read_varint(b, i)
        field, wire = tag >> 3, tag & 7
        if field == 1 and wire == 0:
            dtype, i = _read_varint(b, i)
        elif field == 2 and wire == 0:
            v, i = _read_varint(b, i)
            dims.append(v - (1 << 64) if v >= (1 << 63) else v)
        elif field == 2 and wire == 2:
            n, i = _read_varint(b, i)
            end = i + n
            while i < end:
                v, i = _read_varint(b, i)
                dims.append(v - (1 << 64) if v >= (1 << 63) else v)
        elif wire == 0:
            _, i = _read_varint(b, i)
        elif wire == 2:
            n, i = _read_varint(b, i)
            i += n
        elif wire == 5:
            i += 4
        elif wire == 1:
            i += 8
        else:
            raise ValueError(f"TensorDesc: unsupported wire type {wire}")
    if dtype is None:
        raise ValueError("TensorDesc without data_type")
    return dtype, dims


def build_tensor_desc(dtype_code: int, dims: Sequence[int]) -> bytes:
    out = b"\x08" + _write_varint(dtype_code)
    for d in dims:  # proto2 repeated scalar: unpacked, one tag per element
        out += b"\x10" + _write_varint(int(d))
    return out


@dataclass
class TensorRecord:
    dtype: str            # numpy-style name; "bfloat16" has no numpy type (raw uint16 bits)
    dims: List[int]
    lod: List[List[int]]
    data_offset: int      # byte offset of the raw data in the file
    nbytes: int

    @property
    def elem_size(self) -> int:
        return _PD_TYPES[_PD_CODES[self.dtype]][1]


def _read_exact(f: BinaryIO, n: int) -> bytes:
    b = f.read(n)
    if len(b) != n:
        raise EOFError("truncated tensor stream")
    return b


def read_record_header(f: BinaryIO) -> Optional[TensorRecord]:
    """Reads one record's header, leaving the file positioned at its raw data.  None at EOF."""
    head = f.read(4)
    if len(head) == 0:
        return None
    if len(head) != 4:
        raise EOFError("truncated tensor stream")
    (version,) = struct.unpack("<I", head)
    if version != 0:
        raise ValueError(f"Tensor version {version} is not supported.")  # serialization.cc:63-74
    (n_levels,) = struct.unpack("<Q", _read_exact(f, 8))
    lod = []
    for _ in range(n_levels):
        (nb,) = struct.unpack("<Q", _read_exact(f, 8))
        lod.append(list(np.frombuffer(_read_exact(f, nb), dtype="<u8").astype(np.int64)))
    (tversion,) = struct.unpack("<I", _read_exact(f, 4))
    if tversion != 0:
        raise ValueError(f"Tensor version {tversion} is not supported.")
    (desc_size,) = struct.unpack("<i", _read_exact(f, 4))
    code, dims = parse_tensor_desc(_read_exact(f, desc_size))
    if code not in _PD_TYPES:
        raise NotImplementedError(f"tensor data_type {code} is not handled by this loader")
    name, es = _PD_TYPES[code]
    numel = int(np.prod(dims)) if dims else 1
    return TensorRecord(name, [int(d) for d in dims], lod, f.tell(), numel * es)


def iter_records(path: str) -> Iterator[TensorRecord]:
    """Headers of all tensors in a combined params file, in file order (save_combine writes the
    persistables sorted by name)."""
    with open(path, "rb") as f:
        while True:
            rec = read_record_header(f)
            if rec is None:
                return
            yield rec
            f.seek(rec.data_offset + rec.nbytes)


def read_tensor(path: str, rec: TensorRecord) -> np.ndarray:
    """The whole tensor on the host (bfloat16 as uint16 bit patterns)."""
    np_dtype = np.uint16 if rec.dtype == "bfloat16" else np.dtype(rec.dtype)
    with open(path, "rb") as f:
        f.seek(rec.data_offset)
        a = np.frombuffer(_read_exact(f, rec.nbytes), dtype=np_dtype)
    return a.reshape(rec.dims)


def write_tensor(f: BinaryIO, array: np.ndarray, lod: Sequence[Sequence[int]] = (), bf16: bool = False) -> None:
    """Appends one record (what SerializeToStream writes).  bf16=True: `array` holds uint16 bit
    patterns of bfloat16 values."""
    a = np.ascontiguousarray(array)
    name = "bfloat16" if bf16 else str(a.dtype)
    if name not in _PD_CODES:
        raise NotImplementedError(f"dtype {name}")
    f.write(struct.pack("<I", 0))
    f.write(struct.pack("<Q", len(lod)))
    for level in lod:
        lv = np.asarray(level, dtype="<u8")
        f.write(struct.pack("<Q", lv.nbytes))
        f.write(lv.tobytes())
    f.write(struct.pack("<I", 0))
    desc = build_tensor_desc(_PD_CODES[name], a.shape)
    f.write(struct.pack("<i", len(desc)))
    f.write(desc)
    f.write(a.tobytes())


# ---- straight into device memory ----------------------------------------------------------------
def load_table_to_device(path: str, rec: TensorRecord, out, *, row0: int = 0, row_stride: int = 1,
                         chunk_rows: int = 1 << 18) -> None:
    """Streams table rows row0, row0 + row_stride, ... of the [N, D] tensor `rec` into the CUDA
    tensor `out` ([rows, D]) through a pinned staging buffer, chunk by chunk -- the whole table
    never sits in host memory.  row_stride > 1 picks one shard of a row-sharded table
    (shard.ShardedTable.local with row0 = rank, row_stride = n_shards)."""
    import torch
    if len(rec.dims) != 2:
        raise ValueError("an embedding table is a rank-2 tensor")
    n_rows, dim = rec.dims
    tdt = {"float32": torch.float32, "float16": torch.float16, "bfloat16": torch.bfloat16}.get(rec.dtype)
    if tdt is None or out.dtype != tdt or out.shape[1] != dim or not out.is_cuda:
        raise ValueError("destination must be a CUDA tensor of the table's dtype and width")
    want = len(range(row0, n_rows, row_stride))
    if out.shape[0] != want:
        raise ValueError(f"destination has {out.shape[0]} rows, the selection has {want}")
    es = rec.elem_size
    view_dt = torch.int16 if rec.dtype == "bfloat16" else tdt
    stage = [torch.empty((chunk_rows, dim), dtype=view_dt).pin_memory() for _ in range(2)]
    events = [torch.cuda.Event(), torch.cuda.Event()]
    used = [False, False]
    dst = out.view(view_dt) if rec.dtype == "bfloat16" else out
    np_dt = np.int16 if rec.dtype == "bfloat16" else np.dtype(rec.dtype)
    written = 0
    with open(path, "rb") as f:
        k = 0
        for first in range(0, n_rows, chunk_rows * row_stride if row_stride > 1 else chunk_rows):
            # rows [first, last) of the file contain the next <= chunk_rows selected rows
            span = chunk_rows * row_stride if row_stride > 1 else chunk_rows
            last = min(first + span, n_rows)
            f.seek(rec.data_offset + first * dim * es)
            block = np.frombuffer(_read_exact(f, (last - first) * dim * es), dtype=np_dt).reshape(last - first, dim)
            # first selected row at or after `first`
            off = (row0 - first) % row_stride if first > row0 else row0 - first
            picked = block[off::row_stride] if off < block.shape[0] else block[:0]
            n = picked.shape[0]
            if n == 0:
                continue
            b = k % 2
            if used[b]:
                events[b].synchronize()
            stage[b][:n].copy_(torch.from_numpy(np.array(picked)))  # a writable copy of the mmap-like view
            dst[written:written + n].copy_(stage[b][:n], non_blocking=True)
            events[b].record()
            used[b] = True
            written += n
            k += 1
    torch.cuda.synchronize()
    if written != want:
        raise RuntimeError(f"loaded {written} rows, expected {want}")
