"""B200-native succinct column-segment codec (AdaCom's hot path) -- Python front end.

The product is the C-ABI shared library `csrc/libadacom_gpu.so` (hand-written sm_100a kernels) plus the
C++ host shim in `host/`.  This package binds the C ABI with ctypes for tests, bench.py and smoke().
Importing it does not need a GPU; creating a `Context` does, and fails loudly without one.
"""
from . import _capi  # noqa: F401
from .codec import (BLOCK_SIZE, ROW_GROUP_SIZE, VECTOR_SIZE, AdacomError, AggResult, CodecColumn, Column, Context,  # noqa: F401
                    row_group_segment_counts, scan_columns, segment_capacity)
from . import tpch  # noqa: F401
from ._capi import (CODEC_BITPACKING, CODEC_RLE, CMP_EQ, CMP_GE, CMP_GT, CMP_LE, CMP_LT, CMP_NE, LITERAL, PADDED, SEG_COMPACTED,  # noqa: F401
                    SEG_REPACKED, SEG_ANALYZED, SEG_UNCOMPRESSED, ZERO_EXT_MINMAX)

__all__ = ["Context", "Column", "AdacomError", "AggResult", "row_group_segment_counts", "segment_capacity", "scan_columns",
           "tpch"]
