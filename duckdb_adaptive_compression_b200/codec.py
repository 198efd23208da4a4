"""Thin Python front end over the C ABI (used by tests/, bench.py and __graft_entry__).

Mirrors the reference's vocabulary: a `Column` is a batch of ColumnSegments of one PhysicalType on one GPU;
`analyze` / `encode` / `decode` / `scan_partial` / `scan_sum` / `scan_filter` / `fetch_rows` / `uncompact`
are the succinct CompressionFunction's entry points batched over segment ranges.  All computation happens
in libadacom_gpu.so on the GPU; this file only marshals arguments.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _capi
from ._capi import (CMP_EQ, CMP_GE, CMP_GT, CMP_LE, CMP_LT, CMP_NE, LITERAL, PADDED, SEG_COMPACTED,  # noqa: F401
                    SEG_REPACKED, ZERO_EXT_MINMAX, Agg, SegmentMeta)

BLOCK_SIZE = 262144 - 8  # Storage::BLOCK_SIZE
VECTOR_SIZE = 2048       # STANDARD_VECTOR_SIZE
ROW_GROUP_SIZE = 122880  # rows per row group (storage_info.hpp:19)

_PTYPE = {np.dtype(np.uint8): _capi.UINT8, np.dtype(np.int8): _capi.INT8, np.dtype(np.uint16): _capi.UINT16,
          np.dtype(np.int16): _capi.INT16, np.dtype(np.uint32): _capi.UINT32, np.dtype(np.int32): _capi.INT32,
          np.dtype(np.uint64): _capi.UINT64, np.dtype(np.int64): _capi.INT64}


class AdacomError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"adacom {_capi.STATUS_NAMES[status] if 0 <= status < len(_capi.STATUS_NAMES) else status}: {message}")
        self.status = status


def segment_capacity(dtype) -> int:
    """floor(Storage::BLOCK_SIZE / sizeof(T)) (succinct.cpp:313, column_segment.cpp:102-103)."""
    return BLOCK_SIZE // np.dtype(dtype).itemsize


def row_group_segment_counts(n_rows: int, dtype) -> np.ndarray:
    """Segment counts of a column of n_rows rows the way DuckDB lays it out: row groups of 122 880 rows,
    each filled with segments of floor(262136/S) values (segments never cross row groups)."""
    cap = segment_capacity(dtype)
    counts = []
    left = n_rows
    while left > 0:
        rg = min(left, ROW_GROUP_SIZE)
        left -= rg
        while rg > 0:
            c = min(rg, cap)
            counts.append(c)
            rg -= c
    return np.asarray(counts, dtype=np.uint64)


@dataclass
class AggResult:
    sum: int    # exact Python int (128-bit two's complement decoded)
    count: int


def _agg_to_py(a: Agg, signed: bool) -> AggResult:
    v = (a.sum_hi << 64) | a.sum_lo
    if signed and v >= 1 << 127:
        v -= 1 << 128
    return AggResult(v, int(a.count))


def _np_ptr(a: np.ndarray | None):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Context:
    def __init__(self, device: int = 0):
        self.lib = _capi.load()
        h = C.c_void_p()
        st = self.lib.adacom_ctx_create(device, C.byref(h))
        if st != _capi.OK:
            raise AdacomError(st, (self.lib.adacom_last_error(None) or b"").decode())
        self.h = h
        self.device = device
        self._pinned = {}

    def close(self):
        if getattr(self, "h", None):
            for p in list(self._pinned.values()):
                self.lib.adacom_host_free(self.h, p)
            self._pinned.clear()
            self.lib.adacom_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, st: int):
        if st != _capi.OK:
            raise AdacomError(st, (self.lib.adacom_last_error(self.h) or b"").decode())

    def synchronize(self):
        self.check(self.lib.adacom_ctx_synchronize(self.h))

    def flush_l2(self):
        self.check(self.lib.adacom_flush_l2(self.h))

    @property
    def stream_ptr(self) -> int:
        """cudaStream_t of the context's compute stream (for torch.cuda.ExternalStream)"""
        return int(self.lib.adacom_ctx_stream(self.h) or 0)

    @property
    def launch_count(self) -> int:
        return int(self.lib.adacom_ctx_launch_count(self.h))

    # ---- CUDA events on the compute stream ----
    def event(self):
        ev = C.c_void_p()
        self.check(self.lib.adacom_event_create(self.h, C.byref(ev)))
        return ev

    def record(self, ev):
        self.check(self.lib.adacom_event_record(self.h, ev))

    def elapsed_ms(self, start, stop) -> float:
        ms = C.c_float()
        self.check(self.lib.adacom_event_elapsed_ms(self.h, start, stop, C.byref(ms)))
        return float(ms.value)

    def event_destroy(self, ev):
        self.lib.adacom_event_destroy(self.h, ev)

    # ---- pinned host memory as a numpy array ----
    def pinned_array(self, n: int, dtype) -> np.ndarray:
        dt = np.dtype(dtype)
        p = C.c_void_p()
        self.check(self.lib.adacom_host_alloc(self.h, max(n * dt.itemsize, 1), C.byref(p)))
        buf = (C.c_uint8 * max(n * dt.itemsize, 1)).from_address(p.value)
        arr = np.frombuffer(buf, dtype=dt, count=n)
        self._pinned[arr.ctypes.data] = p  # freed explicitly with free_pinned (or with the context)
        return arr

    def free_pinned(self, arr: np.ndarray):
        p = self._pinned.pop(arr.ctypes.data, None)
        if p is not None:
            self.lib.adacom_host_free(self.h, p)

    # ---- columns ----
    def column(self, dtype, seg_counts, seg_capacity: int = 0) -> "Column":
        dt = np.dtype(dtype)
        counts = np.ascontiguousarray(seg_counts, dtype=np.uint64)
        h = C.c_void_p()
        self.check(self.lib.adacom_column_create(self.h, _PTYPE[dt], len(counts), _np_ptr(counts), seg_capacity,
                                                 C.byref(h)))
        return Column(self, h, dt)

    def column_from_values(self, values: np.ndarray, seg_counts=None, seg_capacity: int = 0) -> "Column":
        values = np.ascontiguousarray(values)
        if seg_counts is None:
            seg_counts = row_group_segment_counts(len(values), values.dtype)
        col = self.column(values.dtype, seg_counts, seg_capacity)
        col.upload_raw(values)
        return col

    def column_from_packed(self, dtype, metas, words: np.ndarray) -> "Column":
        dt = np.dtype(dtype)
        words = np.ascontiguousarray(words, dtype=np.uint64)
        arr = (SegmentMeta * len(metas))(*metas)
        h = C.c_void_p()
        self.check(self.lib.adacom_column_from_packed(self.h, _PTYPE[dt], len(metas), arr, _np_ptr(words), len(words),
                                                      C.byref(h)))
        return Column(self, h, dt)

    def restore(self, image, dtype=None, device_bytes: int | None = None) -> "Column":
        """Rebuild a column from a checkpoint image: a numpy uint8 array, or a device pointer (int) of
        `device_bytes` bytes.  Raises AdacomError(ERR_CORRUPT) when a checksum fails."""
        h = C.c_void_p()
        if isinstance(image, np.ndarray):
            image = np.ascontiguousarray(image, dtype=np.uint8)
            self.check(self.lib.adacom_column_restore(self.h, _np_ptr(image), image.nbytes, 0, C.byref(h)))
            ptype = int(image[12:16].view(np.uint32)[0])
        else:
            self.check(self.lib.adacom_column_restore(self.h, C.c_void_p(int(image)), int(device_bytes), 1, C.byref(h)))
            ptype = None
        if dtype is None:
            if ptype is None:
                raise ValueError("dtype is required for a device image")
            dtype = {v: k for k, v in _PTYPE.items()}[ptype]
        return Column(self, h, np.dtype(dtype))

    def compress_host(self, values: np.ndarray, seg_counts, seg_capacity: int = 0, flags: int = 0,
                      words_out: np.ndarray | None = None):
        """host T values -> (Column, metas, total_words); packed words land in words_out when given."""
        dt = values.dtype
        counts = np.ascontiguousarray(seg_counts, dtype=np.uint64)
        metas = (SegmentMeta * len(counts))()
        total = C.c_uint64()
        h = C.c_void_p()
        self.check(self.lib.adacom_compress_host(self.h, _PTYPE[dt], _np_ptr(values), len(counts), _np_ptr(counts),
                                                 seg_capacity, flags, metas, _np_ptr(words_out),
                                                 0 if words_out is None else len(words_out), C.byref(total),
                                                 C.byref(h)))
        return Column(self, h, dt), metas, int(total.value)


class CodecColumn:
    """DuckDB bitpacking / RLE segments (block images as DuckDB writes them) resident on the GPU: batched scan,
    fused SUM and fetch_row (adacom_codec_*)."""

    def __init__(self, ctx: "Context", codec: int, dtype, images, counts):
        self.ctx, self.dtype = ctx, np.dtype(dtype)
        self.signed = self.dtype.kind == "i"
        images = [np.ascontiguousarray(im, dtype=np.uint8) for im in images]
        n = len(images)
        ptrs = (C.c_void_p * max(n, 1))(*[im.ctypes.data for im in images])
        sizes = np.array([im.nbytes for im in images], dtype=np.uint64)
        cnts = np.ascontiguousarray(counts, dtype=np.uint64)
        h = C.c_void_p()
        ctx.check(ctx.lib.adacom_codec_column_create(ctx.h, codec, _PTYPE[self.dtype], n, ptrs, _np_ptr(sizes), _np_ptr(cnts),
                                                     C.byref(h)))
        self.h, self.n_segments, self.counts = h, n, [int(c) for c in cnts]

    def close(self):
        if getattr(self, "h", None) and self.ctx.h:
            self.ctx.lib.adacom_codec_column_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def n_values(self) -> int:
        return int(self.ctx.lib.adacom_codec_column_num_values(self.h))

    @property
    def image_bytes(self) -> int:
        return int(self.ctx.lib.adacom_codec_column_image_bytes(self.h))

    def decode(self, seg_begin=0, seg_end=None, device_out: int | None = None):
        e = self.n_segments if seg_end is None else seg_end
        if device_out is not None:
            self.ctx.check(self.ctx.lib.adacom_codec_decode(self.h, seg_begin, e, C.c_void_p(device_out), 1))
            return None
        out = np.zeros(sum(self.counts[seg_begin:e]), dtype=self.dtype)
        self.ctx.check(self.ctx.lib.adacom_codec_decode(self.h, seg_begin, e, _np_ptr(out), 0))
        return out

    def scan_sum(self, seg_begin=0, seg_end=None) -> AggResult:
        e = self.n_segments if seg_end is None else seg_end
        a = Agg()
        self.ctx.check(self.ctx.lib.adacom_codec_scan_sum(self.h, seg_begin, e, C.byref(a)))
        return _agg_to_py(a, self.signed)

    def fetch_rows(self, seg_idx, row_in_seg) -> np.ndarray:
        seg_idx = np.ascontiguousarray(seg_idx, dtype=np.uint32)
        row_in_seg = np.ascontiguousarray(row_in_seg, dtype=np.uint32)
        out = np.zeros(len(seg_idx), dtype=self.dtype)
        self.ctx.check(self.ctx.lib.adacom_codec_fetch_rows(self.h, len(seg_idx), _np_ptr(seg_idx), _np_ptr(row_in_seg),
                                                            _np_ptr(out), 0))
        return out


class ScanStream:
    """adacom_scan_stream: segment(seg) returns a numpy VIEW of the pinned window (valid until the stream has moved
    depth-1 windows on or is closed) -- copy it if it has to live longer."""

    def __init__(self, col: "Column", seg_begin: int, seg_end: int, window_segments: int, depth: int, flags: int):
        self.col = col
        self.h = C.c_void_p()
        col.ctx.check(col.lib.adacom_scan_stream_open(col.h, seg_begin, seg_end, window_segments, depth, flags, C.byref(self.h)))

    def segment(self, seg: int) -> np.ndarray:
        ptr, cnt = C.c_void_p(), C.c_uint64()
        self.col.ctx.check(self.col.lib.adacom_scan_stream_segment(self.h, seg, C.byref(ptr), C.byref(cnt)))
        if cnt.value == 0:
            return np.zeros(0, dtype=self.col.dtype)
        buf = (C.c_uint8 * (cnt.value * self.col.dtype.itemsize)).from_address(ptr.value)
        return np.frombuffer(buf, dtype=self.col.dtype)

    def close(self):
        if self.h:
            self.col.lib.adacom_scan_stream_close(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _wrap_i64(v: int) -> int:
    v = int(v)
    return v - (1 << 64) if v >= 1 << 63 else v


def scan_columns(ctx: "Context", terms, sums, row_begin: int = 0, row_end: int | None = None, flags: int = 0,
                 bitmap_in: int | None = None, bitmap_out: int | None = None, device_out: int | None = None,
                 group_by: tuple[int, int] | None = None):
    """Fused multi-column scan (adacom_scan_columns / adacom_scan_columns_grouped).

    group_by: (term_index, n_groups) -- GROUP BY the decoded value (0 .. n_groups-1) of that term in the same
              pass; the result is then a list of n_groups lists of AggResult (device_out: group-major aggregates).

    terms: list of (Column, None) or (Column, (lo, hi)) -- inclusive range predicate on the decoded value;
    sums:  list of lists of (term_index, add, mul): SUM over the selected rows of prod (add + mul * value);
           an empty list is COUNT(*).
    Returns one AggResult per sum (None when device_out is given: asynchronous, results stay on the device)."""
    from ._capi import ScanTerm, SumExpr
    n_t, n_s = len(terms), len(sums)
    t_arr = (ScanTerm * max(n_t, 1))()
    for i, (col, pred) in enumerate(terms):
        t_arr[i].col = col.h.value if hasattr(col.h, "value") else col.h
        t_arr[i].has_predicate = 0 if pred is None else 1
        if pred is not None:
            t_arr[i].lo, t_arr[i].hi = _wrap_i64(pred[0]), _wrap_i64(pred[1])
    s_arr = (SumExpr * max(n_s, 1))()
    for g, factors in enumerate(sums):
        s_arr[g].n_factors = len(factors)
        for k, (ti, add, mul) in enumerate(factors):
            s_arr[g].term[k], s_arr[g].add[k], s_arr[g].mul[k] = ti, _wrap_i64(add), _wrap_i64(mul)
    if row_end is None:
        row_end = terms[0][0].n_values
    b_in = C.c_void_p(bitmap_in) if bitmap_in else None
    b_out = C.c_void_p(bitmap_out) if bitmap_out else None
    if group_by is not None:
        g_term, n_g = group_by
        if device_out is not None:
            ctx.check(ctx.lib.adacom_scan_columns_grouped(ctx.h, t_arr, n_t, g_term, n_g, s_arr, n_s, row_begin, row_end,
                                                          flags, b_in, b_out, C.c_void_p(device_out), 1))
            return None
        out = (Agg * max(n_s * n_g, 1))()
        ctx.check(ctx.lib.adacom_scan_columns_grouped(ctx.h, t_arr, n_t, g_term, n_g, s_arr, n_s, row_begin, row_end,
                                                      flags, b_in, b_out, out, 0))
        return [[_agg_to_py(out[g * n_s + k], True) for k in range(n_s)] for g in range(n_g)]
    if device_out is not None:
        ctx.check(ctx.lib.adacom_scan_columns_async(ctx.h, t_arr, n_t, s_arr, n_s, row_begin, row_end, flags, b_in,
                                                    b_out, C.c_void_p(device_out)))
        return None
    out = (Agg * max(n_s, 1))()
    ctx.check(ctx.lib.adacom_scan_columns(ctx.h, t_arr, n_t, s_arr, n_s, row_begin, row_end, flags, b_in, b_out, out))
    return [_agg_to_py(out[g], True) for g in range(n_s)]


class Column:
    def __init__(self, ctx: Context, handle, dtype: np.dtype):
        self.ctx, self.h, self.dtype = ctx, handle, dtype
        self.lib = ctx.lib
        self.signed = dtype.kind == "i"

    def close(self):
        if getattr(self, "h", None) and self.ctx.h:
            self.lib.adacom_column_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def n_values(self) -> int:
        return int(self.lib.adacom_column_num_values(self.h))

    @property
    def n_segments(self) -> int:
        return int(self.lib.adacom_column_num_segments(self.h))

    @property
    def raw_ptr(self) -> int:
        return int(self.lib.adacom_column_raw_device_ptr(self.h) or 0)

    @property
    def packed_words(self) -> int:
        return int(self.lib.adacom_column_packed_words(self.h))

    def _range(self, seg_begin, seg_end):
        return seg_begin, (self.n_segments if seg_end is None else seg_end)

    def upload_raw(self, values: np.ndarray, first_value: int = 0):
        assert values.dtype == self.dtype and values.flags.c_contiguous
        self.ctx.check(self.lib.adacom_column_upload_raw(self.h, _np_ptr(values), first_value, len(values)))

    def release_raw(self):
        self.ctx.check(self.lib.adacom_column_release_raw(self.h))

    def fill_splitmix(self, seed: int, first_index: int, mask: int):
        self.ctx.check(self.lib.adacom_column_fill_splitmix(self.h, seed & (2**64 - 1), first_index, mask & (2**64 - 1)))

    def analyze(self, seg_begin=0, seg_end=None, flags=0, validity: np.ndarray | None = None):
        b, e = self._range(seg_begin, seg_end)
        if validity is not None:
            assert validity.dtype == np.uint64 and validity.flags.c_contiguous
        self.ctx.check(self.lib.adacom_analyze(self.h, b, e, flags, _np_ptr(validity), 0))

    def encode(self, seg_begin=0, seg_end=None, flags=0):
        b, e = self._range(seg_begin, seg_end)
        self.ctx.check(self.lib.adacom_encode(self.h, b, e, flags))

    def compact(self, seg_begin=0, seg_end=None, flags=0, validity=None):
        """ColumnSegment::Compact for a range: analyze (the append path's min/max) + encode."""
        self.analyze(seg_begin, seg_end, flags, validity)
        self.encode(seg_begin, seg_end, flags)

    # ---- segment-list forms (policy epochs): ids strictly ascending ----
    def compact_list(self, segs, flags=0, analyze=True):
        segs = np.ascontiguousarray(segs, dtype=np.uint32)
        if analyze:
            self.ctx.check(self.lib.adacom_analyze_list(self.h, len(segs), _np_ptr(segs), flags, None))
        self.ctx.check(self.lib.adacom_encode_list(self.h, len(segs), _np_ptr(segs), flags))

    def uncompact_list(self, segs, flags=0):
        segs = np.ascontiguousarray(segs, dtype=np.uint32)
        self.ctx.check(self.lib.adacom_uncompact_list(self.h, len(segs), _np_ptr(segs), flags))

    def meta(self, seg_begin=0, seg_end=None):
        b, e = self._range(seg_begin, seg_end)
        arr = (SegmentMeta * max(e - b, 1))()
        self.ctx.check(self.lib.adacom_column_get_meta(self.h, b, e, arr))
        return list(arr)[: e - b]

    def data_size(self, seg_begin=0, seg_end=None) -> int:
        b, e = self._range(seg_begin, seg_end)
        out = C.c_uint64()
        self.ctx.check(self.lib.adacom_column_data_size(self.h, b, e, C.byref(out)))
        return int(out.value)

    def download_packed(self, first_word=0, n_words=None) -> np.ndarray:
        if n_words is None:
            n_words = self.packed_words - first_word
        out = np.zeros(n_words, dtype=np.uint64)
        self.ctx.check(self.lib.adacom_column_download_packed(self.h, first_word, n_words, _np_ptr(out)))
        return out

    def segment_words(self, seg: int) -> np.ndarray:
        m = self.meta(seg, seg + 1)[0]
        return self.download_packed(m.word_offset, m.words)

    def serialize(self, seg: int) -> bytes:
        n = C.c_uint64()
        self.ctx.check(self.lib.adacom_segment_serialize(self.h, seg, None, 0, C.byref(n)))
        buf = (C.c_uint8 * n.value)()
        self.ctx.check(self.lib.adacom_segment_serialize(self.h, seg, buf, n.value, C.byref(n)))
        return bytes(buf)

    def checkpoint_size(self) -> int:
        n = C.c_uint64()
        self.ctx.check(self.lib.adacom_column_checkpoint_size(self.h, C.byref(n)))
        return int(n.value)

    def checkpoint(self, device_out: int | None = None, capacity: int | None = None):
        """Checkpoint image of the column (include/adacom_gpu.h "checkpoint image"): numpy uint8 array, or written
        to the device pointer `device_out` (then the byte count is returned)."""
        n = C.c_uint64()
        if device_out is not None:
            self.ctx.check(self.lib.adacom_column_checkpoint(self.h, C.c_void_p(int(device_out)), int(capacity), 1,
                                                             C.byref(n)))
            return int(n.value)
        out = np.empty(self.checkpoint_size(), dtype=np.uint8)
        self.ctx.check(self.lib.adacom_column_checkpoint(self.h, _np_ptr(out), out.nbytes, 0, C.byref(n)))
        assert n.value == out.nbytes
        return out

    def decode(self, seg_begin=0, seg_end=None, flags=0, out: np.ndarray | None = None, device_out: int | None = None):
        """Batched decode.  Returns a numpy array unless device_out (a device pointer) is given."""
        b, e = self._range(seg_begin, seg_end)
        if device_out is not None:
            self.ctx.check(self.lib.adacom_decode(self.h, b, e, flags, C.c_void_p(device_out), 1))
            return None
        metas = self.meta(b, e)
        rows = sum(m.count for m in metas)
        if out is None:
            out = np.zeros(rows, dtype=self.dtype)
        assert out.dtype == self.dtype and len(out) >= rows
        self.ctx.check(self.lib.adacom_decode(self.h, b, e, flags, _np_ptr(out), 0))
        return out

    def scan_partial(self, seg: int, start: int, scan_count: int, flags=0, result: np.ndarray | None = None,
                     result_offset: int = 0) -> np.ndarray:
        if result is None:
            result = np.zeros(result_offset + scan_count, dtype=self.dtype)
        self.ctx.check(self.lib.adacom_scan_partial(self.h, seg, start, scan_count, flags, _np_ptr(result),
                                                    result_offset, 0))
        return result

    def scan_vector(self, seg: int, start: int, scan_count: int = VECTOR_SIZE, flags=0) -> np.ndarray:
        return self.scan_partial(seg, start, scan_count, flags)

    def scan_stream(self, seg_begin=0, seg_end=None, window_segments=0, depth=0, flags=0) -> "ScanStream":
        """Pull-model scan (adacom_scan_stream_*): windows of segments decoded ahead into a pinned host ring."""
        b, e = self._range(seg_begin, seg_end)
        return ScanStream(self, b, e, window_segments, depth, flags)

    def uncompact(self, seg_begin=0, seg_end=None, flags=0):
        b, e = self._range(seg_begin, seg_end)
        self.ctx.check(self.lib.adacom_uncompact(self.h, b, e, flags))

    def download_raw(self) -> np.ndarray:
        """Test helper: the current values of every segment as one flat array (one batched decode: segments that are not
        re-packed decode as their slot-width image)."""
        return self.decode()

    def scan_sum(self, seg_begin=0, seg_end=None, flags=0) -> AggResult:
        b, e = self._range(seg_begin, seg_end)
        a = Agg()
        self.ctx.check(self.lib.adacom_scan_sum(self.h, b, e, flags, C.byref(a)))
        return _agg_to_py(a, self.signed)

    def scan_sum_async(self, device_out: int, seg_begin=0, seg_end=None, flags=0):
        b, e = self._range(seg_begin, seg_end)
        self.ctx.check(self.lib.adacom_scan_sum_async(self.h, b, e, flags, C.c_void_p(device_out)))

    def scan_filter(self, op: int, constant: int, seg_begin=0, seg_end=None, flags=0, device_bitmap: int | None = None):
        b, e = self._range(seg_begin, seg_end)
        a = Agg()
        c = int(constant)
        if c >= 1 << 63:
            c -= 1 << 64
        self.ctx.check(self.lib.adacom_scan_filter(self.h, b, e, flags, op, c, C.byref(a),
                                                   C.c_void_p(device_bitmap) if device_bitmap else None))
        return _agg_to_py(a, self.signed)

    def scan_select(self, preds, conj: int = 0, validity=None, seg_begin=0, seg_end=None, flags=0,
                    bitmap_in: int | None = None, bitmap_out: int | None = None, validity_on_device: bool = False):
        """adacom_scan_select: preds = list of ("cmp", op, constant) | ("is_null",) | ("is_not_null",); conj 0 = AND,
        1 = OR; validity = numpy uint64 words of the whole column (or a device pointer with validity_on_device).
        Returns (AggResult over the selected valid rows with count = selected rows, selected valid rows)."""
        from ._capi import Pred
        b, e = self._range(seg_begin, seg_end)
        arr = (Pred * max(len(preds), 1))()
        for i, p in enumerate(preds):
            if p[0] == "cmp":
                arr[i].kind, arr[i].op, arr[i].constant = 0, p[1], _wrap_i64(int(p[2]) & (2**64 - 1))
            else:
                arr[i].kind = 1 if p[0] == "is_null" else 2
        a = Agg()
        if validity is None:
            vp = None
        elif validity_on_device:
            vp = C.c_void_p(int(validity))
        else:
            assert validity.dtype == np.uint64 and validity.flags.c_contiguous
            vp = _np_ptr(validity)
        self.ctx.check(self.lib.adacom_scan_select(self.h, b, e, flags, arr, len(preds), conj, vp,
                                                   1 if validity_on_device else 0,
                                                   C.c_void_p(bitmap_in) if bitmap_in else None,
                                                   C.c_void_p(bitmap_out) if bitmap_out else None, C.byref(a)))
        return _agg_to_py(a, self.signed), int(a.reserved)

    def scan_sum_valid(self, validity, seg_begin=0, seg_end=None, flags=0):
        b, e = self._range(seg_begin, seg_end)
        a = Agg()
        self.ctx.check(self.lib.adacom_scan_sum_valid(self.h, b, e, flags, _np_ptr(validity), 0, C.byref(a)))
        return _agg_to_py(a, self.signed), int(a.reserved)

    def point_filter_batch(self, seg_idx, constants, flags=0):
        seg_idx = np.ascontiguousarray(seg_idx, dtype=np.uint32)
        constants = np.ascontiguousarray(constants, dtype=np.int64)
        n = len(seg_idx)
        counts = np.zeros(n, dtype=np.uint64)
        first = np.zeros(n, dtype=np.uint64)
        self.ctx.check(self.lib.adacom_point_filter_batch(self.h, n, _np_ptr(seg_idx), _np_ptr(constants), flags,
                                                          _np_ptr(counts), _np_ptr(first)))
        return counts, first

    def fetch_rows(self, seg_idx, row_in_seg, flags=0) -> np.ndarray:
        seg_idx = np.ascontiguousarray(seg_idx, dtype=np.uint32)
        row_in_seg = np.ascontiguousarray(row_in_seg, dtype=np.uint32)
        out = np.zeros(len(seg_idx), dtype=self.dtype)
        self.ctx.check(self.lib.adacom_fetch_rows(self.h, len(seg_idx), _np_ptr(seg_idx), _np_ptr(row_in_seg), flags,
                                                  _np_ptr(out), 0))
        return out
