"""LoD (level-of-detail) helpers: the host-side index that packs variable-length
sequences into one tensor.  Offset-based, as the reference stores it
(paddle/fluid/framework/lod_tensor.h:59): each level is a non-decreasing list starting
at 0; the last entry of level i equals the number of sequences of level i+1 (or the
tensor's dim 0 for the last level)."""
from __future__ import annotations

from typing import List, Sequence

LoD = List[List[int]]


def lengths_to_offsets(lengths: Sequence[Sequence[int]]) -> LoD:
    """Length-based ("recursive_sequence_lengths") -> offset-based LoD.  Same running
    sum as the reference tests' convert_to_offset (test_reorder_lod_tensor.py:26-31)."""
    out: LoD = []
    for level in lengths:
        acc = [0]
        for n in level:
            if n < 0:
                raise ValueError("sequence lengths must be >= 0")
            acc.append(acc[-1] + int(n))
        out.append(acc)
    return out


def offsets_to_lengths(lod: Sequence[Sequence[int]]) -> LoD:
    return [[int(b) - int(a) for a, b in zip(level[:-1], level[1:])] for level in lod]


def check_lod(lod: Sequence[Sequence[int]], tensor_height: int = -1) -> bool:
    """CheckLoD / CheckAbsLoD rules (lod_tensor.h:74-104, lod_tensor.cc): every level has
    >= 2 entries, starts at 0, is non-decreasing; level i's last offset is level i+1's
    sequence count; the last level's last offset is the tensor height (when given)."""
    if len(lod) == 0:
        return True
    for level in lod:
        if len(level) < 2 or level[0] != 0:
            return False
        if any(b < a for a, b in zip(level[:-1], level[1:])):
            return False
    for upper, lower in zip(lod[:-1], lod[1:]):
        if upper[-1] != len(lower) - 1:
            return False
    if tensor_height >= 0 and lod[-1][-1] != tensor_height:
        return False
    return True


def num_sequences(lod: Sequence[Sequence[int]]) -> int:
    return len(lod[-1]) - 1
