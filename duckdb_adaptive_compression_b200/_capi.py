"""ctypes binding of libadacom_gpu.so (the C ABI in include/adacom_gpu.h).

This module only declares prototypes; it performs no computation.  If the shared library is missing the
import of the product fails loudly: there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
# ADACOM_GPU_LIB: developer override to A/B two builds of the library in one GPU session
LIB_PATH = os.environ.get("ADACOM_GPU_LIB") or os.path.join(HERE, "csrc", "libadacom_gpu.so")
HEADER_PATH = os.path.join(os.path.dirname(HERE), "include", "adacom_gpu.h")

OK, ERR_INVALID, ERR_CUDA, ERR_NOMEM, ERR_UNSUPPORTED, ERR_STATE, ERR_NO_DEVICE, ERR_CORRUPT = range(8)
STATUS_NAMES = ["OK", "ERR_INVALID", "ERR_CUDA", "ERR_NOMEM", "ERR_UNSUPPORTED", "ERR_STATE", "ERR_NO_DEVICE",
                "ERR_CORRUPT"]

UINT8, INT8, UINT16, INT16, UINT32, INT32, UINT64, INT64 = 2, 3, 4, 5, 6, 7, 8, 9
CMP_EQ, CMP_NE, CMP_LT, CMP_GT, CMP_LE, CMP_GE = range(6)
PADDED, LITERAL, ZERO_EXT_MINMAX = 1, 2, 4
SEG_COMPACTED, SEG_REPACKED, SEG_ANALYZED, SEG_UNCOMPRESSED = 1, 2, 4, 8


class SegmentMeta(C.Structure):
    _fields_ = [("row_start", C.c_uint64), ("count", C.c_uint64), ("word_offset", C.c_uint64), ("words", C.c_uint64),
                ("min_factor", C.c_uint64), ("max_factor", C.c_uint64), ("size_in_bytes", C.c_uint64),
                ("width", C.c_uint8), ("state", C.c_uint8), ("type_size", C.c_uint8), ("is_signed", C.c_uint8),
                ("capacity", C.c_uint32)]


class Agg(C.Structure):
    _fields_ = [("sum_lo", C.c_uint64), ("sum_hi", C.c_uint64), ("count", C.c_uint64), ("reserved", C.c_uint64)]


class Pred(C.Structure):
    _fields_ = [("kind", C.c_int32), ("op", C.c_int32), ("constant", C.c_int64)]


CODEC_BITPACKING, CODEC_RLE = 0, 1
PRED_COMPARE, PRED_IS_NULL, PRED_IS_NOT_NULL = 0, 1, 2
CONJ_AND, CONJ_OR = 0, 1


class ScanTerm(C.Structure):
    _fields_ = [("col", C.c_void_p), ("has_predicate", C.c_int32), ("reserved", C.c_int32), ("lo", C.c_int64),
                ("hi", C.c_int64)]


class SumExpr(C.Structure):
    _fields_ = [("n_factors", C.c_uint32), ("term", C.c_uint32 * 3), ("add", C.c_int64 * 3), ("mul", C.c_int64 * 3)]


class ImageHeader(C.Structure):
    _fields_ = [("magic", C.c_char * 8), ("version", C.c_uint32), ("ptype", C.c_uint32), ("n_segments", C.c_uint64),
                ("n_blocks", C.c_uint64), ("n_values", C.c_uint64), ("blocks_offset", C.c_uint64),
                ("total_bytes", C.c_uint64), ("meta_checksum", C.c_uint64)]


class DataPointer(C.Structure):
    _fields_ = [("row_start", C.c_uint64), ("tuple_count", C.c_uint64), ("block_id", C.c_int64),
                ("offset", C.c_uint32), ("payload_bytes", C.c_uint32), ("min_factor", C.c_uint64),
                ("max_factor", C.c_uint64), ("capacity", C.c_uint32), ("compression_type", C.c_uint8),
                ("width", C.c_uint8), ("state", C.c_uint8), ("type_size", C.c_uint8), ("reserved", C.c_uint64)]


assert C.sizeof(ImageHeader) == 64 and C.sizeof(DataPointer) == 64
assert C.sizeof(SegmentMeta) == 64 and C.sizeof(Agg) == 32 and C.sizeof(ScanTerm) == 32 and C.sizeof(SumExpr) == 64

_vp, _u64, _u32, _i = C.c_void_p, C.c_uint64, C.c_uint32, C.c_int
_PROTOTYPES = {
    "adacom_abi_version": (_i, []),
    "adacom_ctx_create": (_i, [_i, C.POINTER(_vp)]),
    "adacom_ctx_destroy": (None, [_vp]),
    "adacom_last_error": (C.c_char_p, [_vp]),
    "adacom_ctx_synchronize": (_i, [_vp]),
    "adacom_ctx_stream": (_vp, [_vp]),
    "adacom_ctx_launch_count": (_u64, [_vp]),
    "adacom_event_create": (_i, [_vp, C.POINTER(_vp)]),
    "adacom_event_record": (_i, [_vp, _vp]),
    "adacom_event_elapsed_ms": (_i, [_vp, _vp, _vp, C.POINTER(C.c_float)]),
    "adacom_event_destroy": (None, [_vp, _vp]),
    "adacom_host_alloc": (_i, [_vp, C.c_size_t, C.POINTER(_vp)]),
    "adacom_host_free": (None, [_vp, _vp]),
    "adacom_device_alloc": (_i, [_vp, C.c_size_t, C.POINTER(_vp)]),
    "adacom_device_free": (None, [_vp, _vp]),
    "adacom_flush_l2": (_i, [_vp]),
    "adacom_column_create": (_i, [_vp, _i, _u64, _vp, _u32, C.POINTER(_vp)]),
    "adacom_column_destroy": (None, [_vp]),
    "adacom_column_num_values": (_u64, [_vp]),
    "adacom_column_num_segments": (_u64, [_vp]),
    "adacom_column_raw_device_ptr": (_vp, [_vp]),
    "adacom_column_packed_device_ptr": (_vp, [_vp]),
    "adacom_column_upload_raw": (_i, [_vp, _vp, _u64, _u64]),
    "adacom_column_release_raw": (_i, [_vp]),
    "adacom_column_from_packed": (_i, [_vp, _i, _u64, _vp, _vp, _u64, C.POINTER(_vp)]),
    "adacom_column_fill_splitmix": (_i, [_vp, _u64, _u64, _u64]),
    "adacom_analyze": (_i, [_vp, _u64, _u64, _i, _vp, _i]),
    "adacom_encode": (_i, [_vp, _u64, _u64, _i]),
    "adacom_analyze_list": (_i, [_vp, _u64, _vp, _i, _vp]),
    "adacom_encode_list": (_i, [_vp, _u64, _vp, _i]),
    "adacom_uncompact_list": (_i, [_vp, _u64, _vp, _i]),
    "adacom_column_get_meta": (_i, [_vp, _u64, _u64, _vp]),
    "adacom_column_data_size": (_i, [_vp, _u64, _u64, C.POINTER(_u64)]),
    "adacom_column_packed_words": (_u64, [_vp]),
    "adacom_column_download_packed": (_i, [_vp, _u64, _u64, _vp]),
    "adacom_segment_serialize": (_i, [_vp, _u64, _vp, _u64, C.POINTER(_u64)]),
    "adacom_decode": (_i, [_vp, _u64, _u64, _i, _vp, _i]),
    "adacom_scan_partial": (_i, [_vp, _u64, _u64, _u64, _i, _vp, _u64, _i]),
    "adacom_scan_stream_open": (_i, [_vp, _u64, _u64, _u32, _u32, _i, C.POINTER(_vp)]),
    "adacom_scan_stream_segment": (_i, [_vp, _u64, C.POINTER(_vp), C.POINTER(_u64)]),
    "adacom_scan_stream_close": (None, [_vp]),
    "adacom_uncompact": (_i, [_vp, _u64, _u64, _i]),
    "adacom_scan_sum": (_i, [_vp, _u64, _u64, _i, C.POINTER(Agg)]),
    "adacom_scan_sum_async": (_i, [_vp, _u64, _u64, _i, _vp]),
    "adacom_packed_sum_noalloc": (_i, [_vp, _u64, _u32, _u64, _i, _vp, _vp]),
    "adacom_scan_filter": (_i, [_vp, _u64, _u64, _i, _i, C.c_int64, C.POINTER(Agg), _vp]),
    "adacom_scan_select": (_i, [_vp, _u64, _u64, _i, _vp, _u32, _i, _vp, _i, _vp, _vp, C.POINTER(Agg)]),
    "adacom_scan_sum_valid": (_i, [_vp, _u64, _u64, _i, _vp, _i, C.POINTER(Agg)]),
    "adacom_point_filter_batch": (_i, [_vp, _u64, _vp, _vp, _i, _vp, _vp]),
    "adacom_scan_columns": (_i, [_vp, _vp, _u32, _vp, _u32, _u64, _u64, _i, _vp, _vp, _vp]),
    "adacom_scan_columns_async": (_i, [_vp, _vp, _u32, _vp, _u32, _u64, _u64, _i, _vp, _vp, _vp]),
    "adacom_scan_columns_grouped": (_i, [_vp, _vp, _u32, _u32, _u32, _vp, _u32, _u64, _u64, _i, _vp, _vp, _vp, _i]),
    "adacom_fetch_rows": (_i, [_vp, _u64, _vp, _vp, _i, _vp, _i]),
    "adacom_column_checkpoint_size": (_i, [_vp, C.POINTER(_u64)]),
    "adacom_column_checkpoint": (_i, [_vp, _vp, _u64, _i, C.POINTER(_u64)]),
    "adacom_column_restore": (_i, [_vp, _vp, _u64, _i, C.POINTER(_vp)]),
    "adacom_codec_column_create": (_i, [_vp, _i, _i, _u64, _vp, _vp, _vp, C.POINTER(_vp)]),
    "adacom_codec_column_destroy": (None, [_vp]),
    "adacom_codec_column_num_values": (_u64, [_vp]),
    "adacom_codec_column_image_bytes": (_u64, [_vp]),
    "adacom_codec_decode": (_i, [_vp, _u64, _u64, _vp, _i]),
    "adacom_codec_scan_sum": (_i, [_vp, _u64, _u64, C.POINTER(Agg)]),
    "adacom_codec_fetch_rows": (_i, [_vp, _u64, _vp, _vp, _vp, _i]),
    "adacom_compress_host": (_i, [_vp, _i, _vp, _u64, _vp, _u32, _i, _vp, _vp, _u64, C.POINTER(_u64), C.POINTER(_vp)]),
}


def declared_symbols() -> list[str]:
    """Every function name include/adacom_gpu.h declares."""
    text = open(HEADER_PATH).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(adacom_[a-z0-9_]+)\s*\(", text)))


_lib = None


def load() -> C.CDLL:
    """dlopen the C-ABI library and attach prototypes.  Raises if it was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(make -C duckdb_adaptive_compression_b200/csrc). There is no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
