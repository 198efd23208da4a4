"""CPU oracles for the succinct column-segment codec -- TEST INFRASTRUCTURE ONLY.

Only tests/, ``__graft_entry__.smoke()`` and bench.py's ``cpu_baseline`` / ``--impl reference`` legs may
import this package.  The product path (duckdb_adaptive_compression_b200) never does.

Two interchangeable back ends behind one ctypes wrapper (`Oracle`):

* ``kind="port"``      -> oracle/liboracle.so, the SDSL-free C restatement (succinct_oracle.c, "L1").
* ``kind="reference"`` -> oracle/_ref/libadacom_ref.so, ref_harness.cpp linked with the reference's own
  vendored SDSL compiled from /root/reference ("L0").  Present whenever `make -C oracle ref` ran in a
  container that has /root/reference; the built .so travels to the GPU box.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
L1_PATH = os.path.join(HERE, "liboracle.so")
L0_PATH = os.path.join(HERE, "_ref", "libadacom_ref.so")

PADDED = 1
LITERAL = 2

CMP_EQ, CMP_NE, CMP_LT, CMP_GT, CMP_LE, CMP_GE = range(6)

VECTOR_SIZE = 2048       # STANDARD_VECTOR_SIZE (vector_size.hpp:17)
BLOCK_SIZE = 262144 - 8  # Storage::BLOCK_SIZE (constants.hpp:101-106)

_FIELDS = dict(count=0, width=1, bit_size=2, min_factor=3, max_factor=4, compacted=5, repacked=6, function=7,
               words=8, num_elements=9)


def build(force: bool = False) -> None:
    """Compile the L1 port and, when /root/reference is present, the L0 reference harness."""
    if force or not os.path.exists(L1_PATH) or (
            os.path.getmtime(L1_PATH) < os.path.getmtime(os.path.join(HERE, "succinct_oracle.c"))):
        subprocess.check_call(["make", "-C", HERE, "l1"], stdout=subprocess.DEVNULL)
    subprocess.check_call(["make", "-C", HERE, "ref"], stdout=subprocess.DEVNULL)
    build_l2()


def build_l2(force: bool = False) -> None:
    """The whole-reference checkers (oracle/l2): the unmodified reference library, the two l2 drivers, the neighbour
    codecs' reference harness (libadacom_codecs_ref.so) and the TPC-H SF0.1 columns + answers minted by the reference.
    Only where /root/reference is mounted and only when an output is missing (a cold build takes ~5 minutes on 8 cores;
    ADACOM_SKIP_L2=1 skips it).  Failures are reported, not raised: the tests that need these files skip without them."""
    reference = os.environ.get("REFERENCE", "/root/reference")
    if os.environ.get("ADACOM_SKIP_L2") or not os.path.exists(os.path.join(reference, "CMakeLists.txt")):
        return
    ref_dir = os.path.join(HERE, "_ref")
    outputs = [os.path.join(ref_dir, "libadacom_codecs_ref.so"), os.path.join(ref_dir, "l2", "l2_driver_ref"),
               os.path.join(ref_dir, "l2", "l2_driver_gpu")]
    try:
        if force or not all(os.path.exists(o) for o in outputs):
            print("oracle: building the reference out of tree (oracle/l2/build_drivers.sh, one-time, ~5 min)", flush=True)
            subprocess.check_call(["bash", os.path.join(HERE, "l2", "build_drivers.sh")], stdout=subprocess.DEVNULL)
        if not os.path.exists(os.path.join(ref_dir, "l2", "tpch_sf0.1", "answers.json")):
            subprocess.check_call([sys.executable, os.path.join(os.path.dirname(HERE), "tests", "golden", "make_tpch_golden.py"),
                                   "0.1"], stdout=subprocess.DEVNULL)
    except (subprocess.CalledProcessError, OSError) as ex:
        print(f"oracle: L2 checkers not built ({ex}); tests that need them will skip", flush=True)


def have_reference() -> bool:
    return os.path.exists(L0_PATH)


class _BenchResult(C.Structure):
    _fields_ = [("analyze_s", C.c_double), ("encode_s", C.c_double), ("scan_s", C.c_double), ("sum_s", C.c_double), ("uncompact_s", C.c_double),
                ("values", C.c_uint64), ("sum_lo", C.c_uint64), ("sum_hi", C.c_uint64), ("sum_count", C.c_uint64),
                ("packed_bytes", C.c_uint64)]


class _Pred(C.Structure):
    _fields_ = [("kind", C.c_int32), ("op", C.c_int32), ("constant", C.c_int64)]


class _Sum128(C.Structure):
    _fields_ = [("lo", C.c_uint64), ("hi", C.c_uint64), ("count", C.c_uint64)]


@dataclass
class SegmentMeta:
    count: int
    width: int
    bit_size: int
    min_factor: int
    max_factor: int
    compacted: bool
    repacked: bool
    function: int
    words: int
    data_size: int


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    """ctypes front end shared by the L1 port (prefix orc_) and the L0 reference harness (prefix ref_)."""

    def __init__(self, kind: str = "port"):
        if kind == "port":
            if not os.path.exists(L1_PATH):
                build()
            self.lib = C.CDLL(L1_PATH)
            self.p = "orc_"
        elif kind == "reference":
            if not os.path.exists(L0_PATH):
                raise FileNotFoundError(f"{L0_PATH} missing: run `make -C oracle ref` where /root/reference exists")
            self.lib = C.CDLL(L0_PATH)
            self.p = "ref_"
        else:
            raise ValueError(kind)
        self.kind = kind
        f = self._f
        f("segment_new_succinct", C.c_void_p, [C.c_uint8, C.c_int, C.c_uint64])
        f("segment_new_uncompressed", C.c_void_p, [C.c_uint8, C.c_int, C.c_uint64])
        f("segment_delete", None, [C.c_void_p])
        f("segment_append", C.c_uint64, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64])
        f("segment_append_uncompressed", C.c_uint64, [C.c_void_p, C.c_void_p, C.c_uint64])
        f("bit_compress_from_succinct", None, [C.c_void_p, C.c_int])
        f("bit_compress_from_uncompressed", None, [C.c_void_p, C.c_int])
        f("scan_partial", None, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint64, C.c_int])
        f("uncompress_succinct", None, [C.c_void_p, C.c_int])
        f("segment_data_size", C.c_uint64, [C.c_void_p])
        f("segment_get", C.c_uint64, [C.c_void_p, C.c_int])
        f("segment_words", C.c_void_p, [C.c_void_p])
        f("segment_block", C.c_void_p, [C.c_void_p])
        f("segment_serialize", C.c_uint64, [C.c_void_p, C.c_void_p, C.c_uint64])
        f("bits_hi", C.c_uint32, [C.c_uint64])
        f("bench_column", C.c_int, [C.c_void_p, C.c_uint64, C.c_uint8, C.c_int, C.c_uint64, C.c_int, C.c_int,
                                    C.POINTER(_BenchResult)])
        if kind == "reference":
            f("bench_lookups", C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_void_p,
                                         C.c_void_p, C.POINTER(C.c_double)])
            f("bench_fetch", C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p,
                                       C.POINTER(C.c_double)])
            f("bench_epoch", C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.POINTER(C.c_double),
                                       C.POINTER(C.c_double)])
            f("bench_query", C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_int32, C.c_uint32, C.c_uint64,
                                       C.c_int, C.c_int, C.c_void_p, C.POINTER(C.c_double)])
        if kind == "port":
            f("segment_sum", None, [C.c_void_p, C.c_int, C.POINTER(_Sum128)])
            f("segment_filter", None, [C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.POINTER(_Sum128), C.c_void_p])
            f("fetch_row", None, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_int])
            f("segment_select", None, [C.c_void_p, C.c_int, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p, C.c_uint64,
                                       C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(_Sum128), C.POINTER(C.c_uint64)])

    def _f(self, name, restype, argtypes):
        fn = getattr(self.lib, self.p + name)
        fn.restype = restype
        fn.argtypes = argtypes
        setattr(self, "_" + name, fn)

    # ---- segment life cycle ----
    def new_succinct(self, dtype, segment_size=BLOCK_SIZE) -> "OracleSegment":
        dt = np.dtype(dtype)
        h = self._segment_new_succinct(dt.itemsize, int(dt.kind == "i"), segment_size)
        return OracleSegment(self, h, dt, segment_size)

    def new_uncompressed(self, dtype, segment_size=BLOCK_SIZE) -> "OracleSegment":
        dt = np.dtype(dtype)
        h = self._segment_new_uncompressed(dt.itemsize, int(dt.kind == "i"), segment_size)
        return OracleSegment(self, h, dt, segment_size)

    def bits_hi(self, x: int) -> int:
        return int(self._bits_hi(x & 0xFFFFFFFFFFFFFFFF))

    def encode(self, values: np.ndarray, flags: int = 0, validity: np.ndarray | None = None,
               segment_size: int | None = None) -> "OracleSegment":
        """append (2048 values per call, like ColumnData::AppendData) + BitCompressFromSuccinct."""
        values = np.ascontiguousarray(values)
        if segment_size is None:
            segment_size = BLOCK_SIZE
        seg = self.new_succinct(values.dtype, segment_size)
        for off in range(0, len(values), VECTOR_SIZE):
            m = min(VECTOR_SIZE, len(values) - off)
            seg.append(values, off, m, validity=validity)
        seg.bit_compress(flags)
        return seg

    def encode_adaptive(self, values: np.ndarray, flags: int = 0, segment_size: int | None = None) -> "OracleSegment":
        values = np.ascontiguousarray(values)
        if segment_size is None:
            segment_size = BLOCK_SIZE
        seg = self.new_uncompressed(values.dtype, segment_size)
        seg.append_uncompressed(values)
        seg.bit_compress_from_uncompressed(flags)
        return seg

    def bench_column(self, values: np.ndarray, seg_capacity: int, threads: int, copy_per_call: bool = False) -> dict:
        values = np.ascontiguousarray(values)
        res = _BenchResult()
        rc = self._bench_column(_ptr(values), len(values), values.dtype.itemsize, int(values.dtype.kind == "i"),
                                seg_capacity, threads, int(copy_per_call), C.byref(res))
        if rc != 0:
            raise RuntimeError("bench_column failed")
        return dict(analyze_s=res.analyze_s, encode_s=res.encode_s, scan_s=res.scan_s, sum_s=res.sum_s, uncompact_s=res.uncompact_s,
                    values=res.values, sum=(res.sum_hi << 64) | res.sum_lo, packed_bytes=res.packed_bytes)


class _QueryCol(C.Structure):
    _fields_ = [("segs", C.c_void_p), ("n_segs", C.c_uint64), ("seg_start", C.c_void_p), ("has_pred", C.c_int32),
                ("reserved", C.c_int32), ("lo", C.c_int64), ("hi", C.c_int64)]


class _SumExpr(C.Structure):
    _fields_ = [("n_factors", C.c_uint32), ("col", C.c_uint32 * 3), ("add", C.c_int64 * 3), ("mul", C.c_int64 * 3)]


class _Agg(C.Structure):
    _fields_ = [("sum_lo", C.c_uint64), ("sum_hi", C.c_uint64), ("count", C.c_uint64), ("reserved", C.c_uint64)]


class RefColumn:
    """A column the way the reference holds it: DuckDB segmentation (seg_counts), every listed segment appended vector by
    vector and compacted (BitCompressFromSuccinct).  Workload CPU baselines (kind="reference" only) run on these."""

    def __init__(self, oracle: "Oracle", values: np.ndarray, seg_counts, only_segments=None, value_of_rows=None,
                 flags: int = 0):
        self.o = oracle
        counts = np.asarray(seg_counts, dtype=np.int64)
        self.starts = np.concatenate([[0], np.cumsum(counts)]).astype(np.uint64)
        self.segs = [None] * len(counts)
        todo = range(len(counts)) if only_segments is None else sorted(set(int(g) for g in only_segments))
        for g in todo:
            a, b = int(self.starts[g]), int(self.starts[g + 1])
            v = values[a:b] if values is not None else value_of_rows(a, b)
            self.segs[g] = oracle.encode(np.ascontiguousarray(v), flags)
        self.handles = (C.c_void_p * len(counts))(*[(s.h if s is not None else None) for s in self.segs])

    def packed_bytes(self, segs=None) -> int:
        ids = range(len(self.segs)) if segs is None else segs
        return sum(self.segs[g].get("words") * 8 for g in ids if self.segs[g] is not None)


def bench_lookups(col: RefColumn, seg_idx, constants, threads: int, copy_per_call: bool = True):
    """config 1 on the reference's SDSL: (match counts, first rows, seconds)."""
    seg_idx = np.ascontiguousarray(seg_idx, dtype=np.uint32)
    constants = np.ascontiguousarray(constants, dtype=np.int64)
    n = len(seg_idx)
    cnt = np.zeros(n, dtype=np.uint64)
    first = np.zeros(n, dtype=np.uint64)
    secs = C.c_double()
    rc = col.o._bench_lookups(col.handles, _ptr(seg_idx), _ptr(constants), n, threads, int(copy_per_call), _ptr(cnt),
                              _ptr(first), C.byref(secs))
    assert rc == 0
    return cnt, first, float(secs.value)


def bench_fetch(col: RefColumn, seg_idx, rows, dtype, threads: int):
    seg_idx = np.ascontiguousarray(seg_idx, dtype=np.uint32)
    rows = np.ascontiguousarray(rows, dtype=np.uint32)
    out = np.zeros(len(seg_idx), dtype=dtype)
    secs = C.c_double()
    rc = col.o._bench_fetch(col.handles, _ptr(seg_idx), _ptr(rows), len(seg_idx), threads, _ptr(out), C.byref(secs))
    assert rc == 0
    return out, float(secs.value)


def bench_epoch(col: RefColumn, hot, flags: int = 0):
    hot = np.ascontiguousarray(hot, dtype=np.uint32)
    a, b = C.c_double(), C.c_double()
    rc = col.o._bench_epoch(col.handles, _ptr(hot), len(hot), flags, C.byref(a), C.byref(b))
    assert rc == 0
    return float(a.value), float(b.value)


def bench_query(o: "Oracle", cols, sums, n_rows: int, threads: int, group_by=None, copy_per_call: bool = True):
    """configs 4/5 on the reference's SDSL.  cols: list of (RefColumn, None | (lo, hi)); sums: list of lists of
    (col_index, add, mul); group_by: (col_index, n_groups).  Returns ([(sum, count)] group-major, seconds)."""
    qc = (_QueryCol * len(cols))()
    for i, (col, pred) in enumerate(cols):
        qc[i].segs = C.cast(col.handles, C.c_void_p)
        qc[i].n_segs = len(col.segs)
        qc[i].seg_start = col.starts.ctypes.data
        qc[i].has_pred = 0 if pred is None else 1
        if pred is not None:
            qc[i].lo, qc[i].hi = int(pred[0]), int(pred[1])
    se = (_SumExpr * max(len(sums), 1))()
    for s, factors in enumerate(sums):
        se[s].n_factors = len(factors)
        for k, (ci, add, mul) in enumerate(factors):
            se[s].col[k], se[s].add[k], se[s].mul[k] = ci, add, mul
    g_col, n_g = (-1, 0) if group_by is None else group_by
    out = (_Agg * (max(n_g, 1) * max(len(sums), 1)))()
    secs = C.c_double()
    rc = o._bench_query(qc, len(cols), se, len(sums), g_col, n_g, n_rows, threads, int(copy_per_call), out, C.byref(secs))
    assert rc == 0
    res = []
    for a in out:
        v = (a.sum_hi << 64) | a.sum_lo
        if v >= 1 << 127:
            v -= 1 << 128
        res.append((v, int(a.count)))
    return res, float(secs.value)


class OracleSegment:
    def __init__(self, oracle: Oracle, handle, dtype: np.dtype, segment_size: int):
        self.o, self.h, self.dtype, self.segment_size = oracle, handle, dtype, segment_size

    def __del__(self):
        try:
            if self.h:
                self.o._segment_delete(self.h)
                self.h = None
        except Exception:
            pass

    def get(self, field: str) -> int:
        return int(self.o._segment_get(self.h, _FIELDS[field]))

    def meta(self) -> SegmentMeta:
        g = self.get
        return SegmentMeta(g("count"), g("width"), g("bit_size"), g("min_factor"), g("max_factor"),
                           bool(g("compacted")), bool(g("repacked")), g("function"), g("words"), self.data_size())

    def append(self, values: np.ndarray, offset: int, count: int, sel: np.ndarray | None = None,
               validity: np.ndarray | None = None) -> int:
        assert values.dtype == self.dtype and values.flags.c_contiguous
        selp = _ptr(np.ascontiguousarray(sel, dtype=np.uint32)) if sel is not None else None
        valp = _ptr(validity) if validity is not None else None
        if validity is not None:
            assert validity.dtype == np.uint64 and validity.flags.c_contiguous
        return int(self.o._segment_append(self.h, _ptr(values), selp, valp, offset, count))

    def append_uncompressed(self, values: np.ndarray) -> int:
        assert values.dtype == self.dtype and values.flags.c_contiguous
        return int(self.o._segment_append_uncompressed(self.h, _ptr(values), len(values)))

    def bit_compress(self, flags: int = 0) -> None:
        self.o._bit_compress_from_succinct(self.h, flags)

    def bit_compress_from_uncompressed(self, flags: int = 0) -> None:
        self.o._bit_compress_from_uncompressed(self.h, flags)

    def scan_partial(self, start: int, scan_count: int, flags: int = 0, result: np.ndarray | None = None,
                     result_offset: int = 0) -> np.ndarray:
        if result is None:
            result = np.zeros(result_offset + scan_count, dtype=self.dtype)
        self.o._scan_partial(self.h, start, scan_count, _ptr(result), result_offset, flags)
        return result

    def decode_all(self, flags: int = 0) -> np.ndarray:
        n = self.get("count")
        out = np.zeros(n, dtype=self.dtype)
        for off in range(0, n, VECTOR_SIZE):
            m = min(VECTOR_SIZE, n - off)
            self.o._scan_partial(self.h, off, m, _ptr(out), off, flags)
        return out

    def uncompress(self, flags: int = 0) -> np.ndarray:
        n = self.get("bit_size") // max(self.get("width"), 1)
        self.o._uncompress_succinct(self.h, flags)
        raw = (C.c_uint8 * self.segment_size).from_address(self.o._segment_block(self.h))
        return np.frombuffer(bytes(raw), dtype=self.dtype, count=n)

    def data_size(self) -> int:
        return int(self.o._segment_data_size(self.h))

    def words(self) -> np.ndarray:
        n = self.get("words")
        if n == 0:
            return np.zeros(0, dtype=np.uint64)
        buf = (C.c_uint64 * n).from_address(self.o._segment_words(self.h))
        return np.frombuffer(bytes(buf), dtype=np.uint64)

    def serialize(self) -> bytes:
        cap = self.data_size() + 16
        buf = (C.c_uint8 * cap)()
        n = self.o._segment_serialize(self.h, buf, cap)
        return bytes(buf[:n])

    # ---- port-only helpers ----
    def sum(self, flags: int = 0) -> tuple[int, int]:
        r = _Sum128()
        self.o._segment_sum(self.h, flags, C.byref(r))
        return ((r.hi << 64) | r.lo), int(r.count)

    def filter(self, op: int, constant: int, flags: int = 0, want_bitmap: bool = False):
        r = _Sum128()
        n = self.get("count")
        bm = np.zeros((n + 63) // 64, dtype=np.uint64) if want_bitmap else None
        self.o._segment_filter(self.h, flags, op, constant & 0xFFFFFFFFFFFFFFFF, C.byref(r),
                               _ptr(bm) if bm is not None else None)
        return ((r.hi << 64) | r.lo), int(r.count), bm

    def select(self, preds, conj: int = 0, validity: np.ndarray | None = None, validity_row0: int = 0,
               bitmap_in: np.ndarray | None = None, bitmap_out: np.ndarray | None = None, bitmap_row0: int = 0,
               flags: int = 0):
        """FilterSelection + NULL-skipping aggregate.  preds: ("cmp", op, constant) | ("is_null",) | ("is_not_null",).
        Returns (sum, selected rows, selected valid rows); bitmap_out (uint64 words) is OR-ed into."""
        arr = (_Pred * max(len(preds), 1))()
        for i, p in enumerate(preds):
            if p[0] == "cmp":
                c = int(p[2]) & 0xFFFFFFFFFFFFFFFF
                arr[i].kind, arr[i].op, arr[i].constant = 0, p[1], c - (1 << 64) if c >= 1 << 63 else c
            else:
                arr[i].kind = 1 if p[0] == "is_null" else 2
        r = _Sum128()
        nv = C.c_uint64()
        self.o._segment_select(self.h, flags, arr, len(preds), conj, _ptr(validity) if validity is not None else None,
                               validity_row0, _ptr(bitmap_in) if bitmap_in is not None else None,
                               _ptr(bitmap_out) if bitmap_out is not None else None, bitmap_row0, C.byref(r),
                               C.byref(nv))
        return ((r.hi << 64) | r.lo), int(r.count), int(nv.value)

    def fetch_row(self, row_id: int, flags: int = 0):
        out = np.zeros(1, dtype=self.dtype)
        self.o._fetch_row(self.h, row_id, _ptr(out), 0, flags)
        return out[0]
