"""adacom_packed_sum_noalloc (include/adacom_gpu.h): the allocation-free fused decode + SUM over one packed payload.

The entry point has no counterpart in the reference; its answer is pinned
to the existing entry point for the same work, adacom_scan_sum over the segment the payload belongs to -- bit-exact
(128-bit integer sums), on two calls in a row, with differently sized calls in between so the shared static scratch
holds stale partials.  This file is named to be collected LAST.
"""
import ctypes as C

import numpy as np
import pytest

from tests.cases import width_values

SEG_REPACKED = 2
MAX_VALUES = 1 << 26
PTYPE = {"uint8": 2, "int8": 3, "uint16": 4, "int16": 5, "uint32": 6, "int32": 7, "uint64": 8, "int64": 9}


class _Agg(C.Structure):
    _fields_ = [("sum_lo", C.c_uint64), ("sum_hi", C.c_uint64), ("count", C.c_uint64), ("reserved", C.c_uint64)]


def _lib():
    import duckdb_adaptive_compression_b200._capi as capi
    return capi.load()


def _call(lib, words_ptr, count, width, min_factor, ptype, stream, out_ptr):
    return lib.adacom_packed_sum_noalloc(C.c_void_p(words_ptr), C.c_uint64(count), C.c_uint32(width),
                                         C.c_uint64(min_factor & 0xFFFFFFFFFFFFFFFF), C.c_int(ptype), C.c_void_p(stream),
                                         C.c_void_p(out_ptr))


def _read(out, signed):
    a = out.cpu().numpy().view(np.uint64)
    v = (int(a[1]) << 64) | int(a[0])
    if signed and v >= 1 << 127:
        v -= 1 << 128
    return v, int(a[2]), int(a[3])


def test_noalloc_rejects_bad_arguments():
    """argument checks come before any CUDA call: they run without a GPU"""
    lib = _lib()
    out = (C.c_uint64 * 4)()
    words = (C.c_uint64 * 4)()
    st = _call(lib, C.addressof(words), MAX_VALUES + 1, 8, 0, PTYPE["uint32"], 0, C.addressof(out))
    assert st == 1  # ADACOM_ERR_INVALID
    msg = (lib.adacom_last_error(None) or b"").decode()
    assert "ADACOM_NOALLOC_MAX_VALUES" in msg and str(MAX_VALUES) in msg
    assert _call(lib, C.addressof(words), 10, 0, 0, PTYPE["uint32"], 0, C.addressof(out)) == 1    # width 0
    assert _call(lib, C.addressof(words), 10, 33, 0, PTYPE["uint32"], 0, C.addressof(out)) == 1   # wider than T
    assert _call(lib, 0, 10, 8, 0, PTYPE["uint32"], 0, C.addressof(out)) == 1                     # no payload
    assert _call(lib, C.addressof(words), 10, 8, 0, PTYPE["uint32"], 0, 0) == 1                   # no output
    assert _call(lib, C.addressof(words) + 4, 10, 8, 0, PTYPE["uint32"], 0, C.addressof(out)) == 1  # misaligned
    assert _call(lib, C.addressof(words), 10, 8, 0, 11, 0, C.addressof(out)) == 4                 # ADACOM_ERR_UNSUPPORTED


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", ["uint32", "int16", "int64", "uint8"])
def test_noalloc_matches_scan_sum(gpu_ctx, dtype):
    import torch
    dt = np.dtype(dtype)
    bits = 8 * dt.itemsize
    cap = 262136 // dt.itemsize
    base = -(1 << (bits - 2)) if dt.kind == "i" else 5  # signed: every value negative (a segment that mixes signs keeps its slot width)
    # (width, count): full segments (16+ tiles of 4096), a ragged one, tiny ones; widths that shrink, so every
    # payload lives in the packed arena
    cases = [(min(13, bits - 1), cap), (3, 777), (min(21, bits - 2), cap - 1), (1, 5), (min(9, bits - 1), 4097), (2, 1)]
    vals = [width_values(dt, w, n, base, seed=i) for i, (w, n) in enumerate(cases)]
    col = gpu_ctx.column_from_values(np.concatenate(vals), [len(v) for v in vals])
    col.compact()
    metas = col.meta()
    lib = gpu_ctx.lib
    lib.adacom_column_packed_device_ptr.restype = C.c_void_p
    arena = int(lib.adacom_column_packed_device_ptr(col.h))
    assert arena
    stream = gpu_ctx.stream_ptr
    out = torch.full((4,), -1, dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    signed = dt.kind == "i"
    # big, small, big, small ...: the small calls find the big calls' partials in the scratch
    for g in [0, 1, 0, 3, 2, 5, 4, 4]:
        m = metas[g]
        assert m.state & SEG_REPACKED, "the test needs payloads in the packed arena"
        want = col.scan_sum(g, g + 1)
        assert want.count == len(vals[g])
        assert want.sum == int(vals[g].astype(object).sum())
        for _ in range(2):  # two calls in a row
            out.fill_(-1)
            torch.cuda.synchronize()
            st = _call(lib, arena + 8 * int(m.word_offset), int(m.count), int(m.width), int(m.min_factor), PTYPE[dtype],
                       stream, out.data_ptr())
            assert st == 0, (lib.adacom_last_error(None) or b"").decode()
            gpu_ctx.synchronize()
            got_sum, got_count, reserved = _read(out, signed)
            assert (got_sum, got_count, reserved) == (want.sum, want.count, 0), (dtype, g)
    col.close()


@pytest.mark.gpu
def test_noalloc_large_payload_and_limit(gpu_ctx):
    """a payload of 2^22 + 3 values (1025 tiles) packed by numpy, then an empty one; the limit is enforced on the GPU box too"""
    import torch
    n = (1 << 22) + 3
    rng = np.random.default_rng(7)
    fields = rng.integers(0, 1 << 16, size=n, dtype=np.uint64).astype(np.uint16)
    padded = np.zeros((n + 3) // 4 * 4, dtype=np.uint16)
    padded[:n] = fields
    words = torch.from_numpy(padded.view(np.int64).copy()).to("cuda:0")
    out = torch.full((4,), -1, dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    lib = gpu_ctx.lib
    stream = gpu_ctx.stream_ptr
    want = int(fields.astype(np.uint64).sum()) + 1000 * n
    for _ in range(2):
        out.fill_(-1)
        torch.cuda.synchronize()
        assert _call(lib, words.data_ptr(), n, 16, 1000, PTYPE["uint32"], stream, out.data_ptr()) == 0
        gpu_ctx.synchronize()
        assert _read(out, False) == (want, n, 0)
    # count == 0: a defined, zero aggregate
    assert _call(lib, words.data_ptr(), 0, 16, 1000, PTYPE["uint32"], stream, out.data_ptr()) == 0
    gpu_ctx.synchronize()
    assert _read(out, False) == (0, 0, 0)
    assert _call(lib, words.data_ptr(), MAX_VALUES + 1, 16, 0, PTYPE["uint32"], stream, out.data_ptr()) == 1
