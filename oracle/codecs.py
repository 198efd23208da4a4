"""ctypes front end of oracle/_ref/libadacom_codecs_ref.so -- TEST INFRASTRUCTURE ONLY.

The reference's own bitpacking.cpp / rle.cpp (+ vendored fastpforlib) compiled where they lie by
oracle/l2/build_drivers.sh (see oracle/l2/codecs_ref.cpp): encoder decisions, bit layout and decode arithmetic are the
reference's; only the segment writer / scan-loop glue is restated.  The .so travels to the GPU box with oracle/_ref."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libadacom_codecs_ref.so")
MODE_AUTO, MODE_CONSTANT, MODE_CONSTANT_DELTA, MODE_DELTA_FOR, MODE_FOR = range(5)
BITPACKING, RLE = 0, 1


def available() -> bool:
    return os.path.exists(PATH)


_lib = None


def _load():
    global _lib
    if _lib is None:
        lib = C.CDLL(PATH)
        lib.codecref_compress.restype = C.c_void_p
        lib.codecref_compress.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int]
        lib.codecref_free.argtypes = [C.c_void_p]
        for name in ("codecref_num_segments",):
            getattr(lib, name).restype = C.c_uint64
            getattr(lib, name).argtypes = [C.c_void_p]
        for name in ("codecref_segment_count", "codecref_segment_size"):
            getattr(lib, name).restype = C.c_uint64
            getattr(lib, name).argtypes = [C.c_void_p, C.c_uint64]
        lib.codecref_segment_data.restype = C.c_void_p
        lib.codecref_segment_data.argtypes = [C.c_void_p, C.c_uint64]
        lib.codecref_scan.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p]
        lib.codecref_fetch_row.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p]
        _lib = lib
    return _lib


class RefCodecColumn:
    """values -> the segments DuckDB's BitpackingCompress / RLECompress would write (block images + tuple counts)"""

    def __init__(self, values: np.ndarray, validity: np.ndarray | None = None, mode: int = MODE_AUTO, codec: int = BITPACKING):
        self.lib = _load()
        values = np.ascontiguousarray(values)
        self.dtype = values.dtype
        v8 = None if validity is None else np.ascontiguousarray(validity, dtype=np.uint8)
        self.h = self.lib.codecref_compress(values.ctypes.data, None if v8 is None else v8.ctypes.data, len(values),
                                            values.dtype.itemsize, int(values.dtype.kind == "i"), mode, codec)
        if not self.h:
            raise ValueError("the reference's analyze step rejects this input (Flush returned false)")
        n = self.lib.codecref_num_segments(self.h)
        self.counts = [int(self.lib.codecref_segment_count(self.h, s)) for s in range(n)]
        self.images = []
        for s in range(n):
            size = int(self.lib.codecref_segment_size(self.h, s))
            buf = (C.c_uint8 * size).from_address(self.lib.codecref_segment_data(self.h, s))
            self.images.append(np.frombuffer(bytes(buf), dtype=np.uint8))

    def __del__(self):
        try:
            if self.h:
                self.lib.codecref_free(self.h)
                self.h = None
        except Exception:
            pass

    def scan(self, seg: int, start: int, count: int) -> np.ndarray:
        out = np.zeros(count, dtype=self.dtype)
        self.lib.codecref_scan(self.h, seg, start, count, out.ctypes.data)
        return out

    def fetch_row(self, seg: int, row: int):
        out = np.zeros(1, dtype=self.dtype)
        self.lib.codecref_fetch_row(self.h, seg, row, out.ctypes.data)
        return out[0]

    def modes(self, seg: int) -> list[int]:
        """BitpackingMode of every metadata group of a segment (parsed from the image)"""
        img = self.images[seg]
        end = int(img[:8].view(np.uint64)[0])
        n_groups = (self.counts[seg] + 2047) // 2048
        meta = img[end - 4 * n_groups:end].view(np.uint32)[::-1]
        return [int(m >> 24) for m in meta]
