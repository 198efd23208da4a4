"""GPU parity of the fused single-column filters against the oracle, bit for bit on the SELECTION they produce.

* adacom_scan_filter's selection bitmap (filter_fields_kernel for payloads <= 32 bits, decode_kernel<filter> for wider
  payloads and raw slots): 6 operators x 8 types x ragged segment counts x sub-ranges whose first row is not a multiple
  of 32 / 64, constants inside and outside every segment's [min, max] (the whole-chunk shortcuts).
* adacom_scan_select / adacom_scan_sum_valid: ColumnSegment::FilterSelection semantics (column_segment.cpp:575-683,
  685-844) -- validity mask, IS [NOT] NULL, AND / OR conjunctions, incoming selection -- and DuckDB's NULL-skipping SUM.
"""
import numpy as np
import pytest

from tests.cases import ALL_DTYPES, sweep_cases, width_values

pytestmark = pytest.mark.gpu
MOD = 1 << 128


def _dev_words(n_words):
    import torch
    return torch.zeros(max(n_words, 1) + 2, dtype=torch.int64, device="cuda:0")


def _sync():
    import torch
    torch.cuda.synchronize()


def _to_np(t, n_words):
    return t.cpu().numpy().view(np.uint64)[:n_words].copy()


def _popcount(words):
    return int(np.unpackbits(words.view(np.uint8)).sum())


def _build(gpu_ctx, port_oracle, dtype, cases, validity_frac=None, seed=0):
    vals = [width_values(dtype, w, n, base, seed=i) for i, (w, n, base) in enumerate(cases)]
    counts = [len(v) for v in vals]
    allv = np.concatenate(vals)
    n = len(allv)
    words = None
    if validity_frac is not None:
        rng = np.random.default_rng(seed)
        valid = rng.random(n) > validity_frac
        # a fully-NULL and a fully-valid segment among them
        valid[:counts[0]] = False
        if len(counts) > 1:
            valid[counts[0]:counts[0] + counts[1]] = True
        words = np.packbits(np.pad(valid, (0, (-n) % 64)), bitorder="little").view(np.uint64).copy()
    col = gpu_ctx.column_from_values(allv, counts)
    col.compact(validity=words)
    segs = []
    off = 0
    for v in vals:
        if words is None:
            segs.append(port_oracle.encode(v, 0))
        else:
            # segment-relative validity words for the append path
            bits = np.unpackbits(words.view(np.uint8), bitorder="little")[off:off + len(v)]
            sw = np.packbits(np.pad(bits, (0, (-len(v)) % 64)), bitorder="little").view(np.uint64).copy()
            segs.append(port_oracle.encode(v, 0, validity=sw))
        off += len(v)
    starts = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    return col, vals, segs, starts, words


def _constants(dtype, segs, rng):
    info = np.iinfo(dtype)
    allv = np.concatenate([s.decode_all() for s in segs])
    picks = [int(allv[int(rng.integers(0, len(allv)))]) for _ in range(2)]
    return picks + [int(allv.min()), int(allv.max()), info.min, info.max, int(np.median(allv))]


def _check_filter_bitmaps(gpu_ctx, port_oracle, dtype, cases, ranges):
    rng = np.random.default_rng(5)
    col, vals, segs, starts, _ = _build(gpu_ctx, port_oracle, dtype, cases)
    for (b, e) in ranges:
        e = len(segs) if e is None else e
        rows = int(starts[e] - starts[b])
        nw = (rows + 63) // 64
        bm = _dev_words(nw)
        for op in range(6):
            for const in _constants(dtype, segs[b:e], rng):
                bm.fill_(-1)  # the call must zero it
                _sync()       # fill_ runs on torch's stream, the codec on its own: order them
                got = col.scan_filter(op, const, b, e, device_bitmap=bm.data_ptr())
                want_bm = np.zeros(nw, dtype=np.uint64)
                wsum = wcnt = 0
                for g in range(b, e):
                    sm, c, _ = segs[g].select([("cmp", op, const)], bitmap_out=want_bm, bitmap_row0=int(starts[g] - starts[b]))
                    wsum += sm
                    wcnt += c
                got_bm = _to_np(bm, nw)
                assert np.array_equal(got_bm, want_bm), (np.dtype(dtype).name, op, const, b, e)
                assert got.count == wcnt == _popcount(got_bm)
                assert got.sum % MOD == wsum % MOD
    col.close()


@pytest.mark.parametrize("dtype", ALL_DTYPES, ids=lambda d: np.dtype(d).name)
def test_scan_filter_bitmap_narrow_payloads(gpu_ctx, port_oracle, dtype):
    """payloads <= 32 bits -> filter_fields_kernel (linear chunks, wrap-around frames and raw slots of narrow types)."""
    S = np.dtype(dtype).itemsize
    counts = [1, 31, 33, 65, 2049, 4097, 5000]
    cases = [c for c in sweep_cases(dtype, counts=counts) if c[0] <= 32 and (c[0] % 3 == 1 or c[0] in (8, 16, 32, 8 * S))]
    cases = cases[:60]
    _check_filter_bitmaps(gpu_ctx, port_oracle, dtype, cases, [(0, None), (3, len(cases) - 2), (5, 6)])


@pytest.mark.parametrize("dtype", [np.int64, np.uint64], ids=lambda d: np.dtype(d).name)
def test_scan_filter_bitmap_wide_payloads(gpu_ctx, port_oracle, dtype):
    """a column with payloads > 32 bits and raw 64-bit slots -> decode_kernel<filter>."""
    counts = [1, 31, 65, 2049, 5000]
    cases = [c for c in sweep_cases(dtype, counts=counts) if c[0] % 5 == 3 or c[0] >= 62]
    _check_filter_bitmaps(gpu_ctx, port_oracle, dtype, cases, [(0, None), (2, len(cases) - 3)])


def test_scan_filter_bitmap_full_chunks(gpu_ctx, port_oracle):
    """segments long enough for the full-chunk fast loops (4096-value chunks) with a ragged tail."""
    dtype = np.int32
    cases = [(13, 65534, 1000), (13, 57346, -50000), (7, 20000, 0), (31, 9000, 5)]
    _check_filter_bitmaps(gpu_ctx, port_oracle, dtype, cases, [(0, None), (1, 3)])


PRED_SETS = [
    ([("cmp", 2, None)], 0),                               # value < c, NULLs excluded
    ([("cmp", 1, None)], 0),                               # <>
    ([("is_null",)], 0),
    ([("is_not_null",)], 0),
    ([("cmp", 5, None), ("cmp", 2, "hi")], 0),             # c <= value < hi  (AND)
    ([("cmp", 2, None), ("cmp", 3, "hi")], 1),             # value < c OR value > hi
    ([("cmp", 0, None), ("is_null",)], 1),                 # value = c OR value IS NULL
    ([("cmp", 3, None), ("is_not_null",)], 0),
    ([], 0),                                               # no filter: NULL-skipping SUM
]


@pytest.mark.parametrize("dtype", ALL_DTYPES, ids=lambda d: np.dtype(d).name)
def test_scan_select_nulls_conjunctions(gpu_ctx, port_oracle, dtype):
    S = np.dtype(dtype).itemsize
    rng = np.random.default_rng(9)
    counts = [40, 100, 2049, 5000]
    cases = [c for c in sweep_cases(dtype, counts=counts) if c[0] % 4 == 1 or c[0] == 8 * S][:40]
    col, vals, segs, starts, words = _build(gpu_ctx, port_oracle, dtype, cases, validity_frac=0.25, seed=1)
    allv = np.concatenate([s.decode_all() for s in segs])
    for (b, e) in [(0, len(segs)), (2, len(segs) - 1)]:
        rows = int(starts[e] - starts[b])
        nw = (rows + 63) // 64
        bm_out = _dev_words(nw)
        # an incoming selection for half of the runs
        import torch
        sel_in = rng.integers(0, 1 << 63, size=nw + 2, dtype=np.int64)
        bm_in = torch.from_numpy(sel_in).to("cuda:0")
        for k, (tmpl, conj) in enumerate(PRED_SETS):
            lo = int(allv[int(rng.integers(0, len(allv)))])
            hi = int(allv[int(rng.integers(0, len(allv)))])
            lo, hi = min(lo, hi), max(lo, hi)
            preds = [(p[0], p[1], lo if p[2] is None else hi) if p[0] == "cmp" else p for p in tmpl]
            use_in = k % 2 == 1
            bm_out.fill_(-1)
            _sync()
            got, got_valid = col.scan_select(preds, conj, validity=words, seg_begin=b, seg_end=e,
                                             bitmap_in=bm_in.data_ptr() if use_in else None,
                                             bitmap_out=bm_out.data_ptr())
            want_bm = np.zeros(nw, dtype=np.uint64)
            wsum = wcnt = wvalid = 0
            in_np = sel_in.view(np.uint64) if use_in else None
            for g in range(b, e):
                sm, c, nv = segs[g].select(preds, conj, validity=words, validity_row0=int(starts[g]), bitmap_in=in_np,
                                           bitmap_out=want_bm, bitmap_row0=int(starts[g] - starts[b]))
                wsum += sm
                wcnt += c
                wvalid += nv
            assert np.array_equal(_to_np(bm_out, nw), want_bm), (np.dtype(dtype).name, preds, conj, b, e)
            assert (got.count, got_valid) == (wcnt, wvalid), (preds, conj)
            assert got.sum % MOD == wsum % MOD, (preds, conj)
    # NULL-aware SUM == DuckDB's SUM over the column: NullValue<T>() slots must not enter
    got, nvalid = col.scan_sum_valid(words)
    bits = np.unpackbits(words.view(np.uint8), bitorder="little")[:len(allv)].astype(bool)
    assert nvalid == int(bits.sum()) and got.count == len(allv)
    assert got.sum % MOD == int(allv[bits].astype(object).sum()) % MOD
    # ... while the validity-less call counts them (documented in include/adacom_gpu.h)
    plain = col.scan_sum()
    assert plain.sum % MOD == int(allv.astype(object).sum()) % MOD
    col.close()


@pytest.mark.parametrize("dtype", [np.int8, np.uint16, np.int32, np.int64, np.uint64], ids=lambda d: np.dtype(d).name)
def test_scan_select_random_conjunctions(gpu_ctx, port_oracle, dtype):
    """Random conjunctions of 1..4 children: the range canonicalisation in front of the filter-words kernel (AND =
    intersection with <> holes, OR = merged ranges / complement of the gap, constants at the ends of the type's domain
    and next to each other) and the per-value fallback must both give the oracle's selection."""
    S = np.dtype(dtype).itemsize
    info = np.iinfo(dtype)
    rng = np.random.default_rng(77 + S)
    counts = [33, 2048, 4100]
    cases = [c for c in sweep_cases(dtype, counts=counts) if c[0] % 5 == 2 or c[0] == 8 * S][:12]
    col, vals, segs, starts, words = _build(gpu_ctx, port_oracle, dtype, cases, validity_frac=0.1, seed=3)
    allv = np.concatenate([s.decode_all() for s in segs])
    b, e = 1, len(segs)
    rows = int(starts[e] - starts[b])
    nw = (rows + 63) // 64
    bm_out = _dev_words(nw)
    import torch
    sel_in = rng.integers(0, 1 << 63, size=nw + 2, dtype=np.int64) | rng.integers(0, 1 << 63, size=nw + 2, dtype=np.int64)
    bm_in = torch.from_numpy(sel_in).to("cuda:0")
    pool = [info.min, info.max, info.min + 1, info.max - 1, int(np.median(allv))]
    for trial in range(60):
        anchor = int(allv[int(rng.integers(0, len(allv)))])
        near = [min(max(anchor + d, info.min), info.max) for d in (-1, 0, 1, 2)]
        preds = []
        for _ in range(int(rng.integers(1, 5))):
            if rng.random() < 0.12:
                preds.append(("is_not_null",))
            else:
                c = (pool + near + [int(allv[int(rng.integers(0, len(allv)))])])[int(rng.integers(0, 10))]
                preds.append(("cmp", int(rng.integers(0, 6)), c))
        conj = int(rng.integers(0, 2))
        use_in = trial % 3 == 0
        use_valid = trial % 4 != 3
        bm_out.fill_(-1)
        _sync()
        got, got_valid = col.scan_select(preds, conj, validity=words if use_valid else None, seg_begin=b, seg_end=e,
                                         bitmap_in=bm_in.data_ptr() if use_in else None, bitmap_out=bm_out.data_ptr())
        want_bm = np.zeros(nw, dtype=np.uint64)
        wsum = wcnt = wvalid = 0
        in_np = sel_in.view(np.uint64) if use_in else None
        for g in range(b, e):
            sm, c, nv = segs[g].select(preds, conj, validity=words if use_valid else None, validity_row0=int(starts[g]),
                                       bitmap_in=in_np, bitmap_out=want_bm, bitmap_row0=int(starts[g] - starts[b]))
            wsum += sm
            wcnt += c
            wvalid += nv
        assert np.array_equal(_to_np(bm_out, nw), want_bm), (np.dtype(dtype).name, preds, conj)
        assert (got.count, got_valid) == (wcnt, wvalid), (preds, conj)
        assert got.sum % MOD == wsum % MOD, (preds, conj)
    col.close()


def test_device_alloc_bitmap_chain(gpu_ctx, port_oracle):
    """adacom_device_alloc / adacom_device_free: a caller without the CUDA runtime chains two selections through a
    bitmap it got from the library (lo <= x AND (x = a OR x > hi)), same answer as one oracle pass per segment."""
    import ctypes as C
    dtype = np.int32
    cases = [(13, 5000, 1000), (9, 2048, -300), (20, 4097, 7)]
    col, vals, segs, starts, _ = _build(gpu_ctx, port_oracle, dtype, cases)
    allv = np.concatenate(vals)
    lo, hi, a = int(np.percentile(allv, 30)), int(np.percentile(allv, 80)), int(allv[17])
    rows = len(allv)
    bm = C.c_void_p()
    gpu_ctx.check(gpu_ctx.lib.adacom_device_alloc(gpu_ctx.h, ((rows + 63) // 64 + 2) * 8, C.byref(bm)))
    assert bm.value
    first, _ = col.scan_select([("cmp", 5, lo)], 0, bitmap_out=bm.value)
    got, got_valid = col.scan_select([("cmp", 0, a), ("cmp", 3, hi)], 1, bitmap_in=bm.value)
    keep = (allv >= lo) & ((allv == a) | (allv > hi))
    assert first.count == int((allv >= lo).sum())
    assert got.count == got_valid == int(keep.sum())
    assert got.sum % MOD == int(allv[keep].astype(object).sum()) % MOD
    gpu_ctx.lib.adacom_device_free(gpu_ctx.h, bm)
    col.close()
