"""paddle_infer_b200 -- B200-native (sm_100a) sparse-feature front end for Paddle
Inference CTR models: lookup_table_v2 -> sequence_pool -> cvm, fused.

The product is the C-ABI library `libctrfe.so` (include/ctrfe.h).  This package is the
thin Python host side above it: a ctypes binding (`_lib`), LoD helpers (`lod`) and
mirrors of the reference operator interfaces (`ops`).  There is no CPU fallback: when
the CUDA library is missing, importing `_lib` raises.
"""
from . import lod  # noqa: F401

__version__ = "0.1.0"
