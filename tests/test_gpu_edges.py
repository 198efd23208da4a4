"""Edge cases of the CUDA path against the oracle, through the C ABI: empty and ragged inputs, maximum segment sizes
for every physical type, constant / all-(-1) / all-NULL segments, the literal-quirk and zero-extended modes of the
fused aggregates, re-encode cycles and argument errors.  Bit-exact (integer path)."""
from __future__ import annotations

import numpy as np
import pytest

import duckdb_adaptive_compression_b200 as adacom
from tests.cases import ALL_DTYPES, width_values

pytestmark = pytest.mark.gpu

MOD = 1 << 128
LITERAL, ZERO_EXT = 2, 4


def _validity_words(valid: np.ndarray) -> np.ndarray:
    """bool per row -> DuckDB validity words (bit i of word i / 64 = row i valid)"""
    b = np.packbits(valid, bitorder="little")
    return np.ascontiguousarray(np.concatenate([b, np.zeros((-len(b)) % 8, np.uint8)])).view(np.uint64).copy()


def _oracle_sum(segs, flags=0):
    return sum(s.sum(flags)[0] for s in segs) % MOD, sum(s.sum(flags)[1] for s in segs)


def test_empty_segments_and_empty_ranges(gpu_ctx, port_oracle):
    """Compact() returns early for empty segments (column_segment.cpp:278); empty ranges are no-ops."""
    dtype = np.int32
    parts = [width_values(dtype, 9, 100, 5, 1), np.zeros(0, dtype), width_values(dtype, 3, 2049, -7, 2),
             np.zeros(0, dtype), np.zeros(0, dtype), width_values(dtype, 20, 65, 1000003, 3)]
    counts = [len(p) for p in parts]
    col = gpu_ctx.column_from_values(np.concatenate(parts), counts)
    col.compact()
    metas = col.meta()
    segs = [port_oracle.encode(p, 0) if len(p) else None for p in parts]
    for m, p, s in zip(metas, parts, segs):
        assert m.count == len(p)
        if s is None:
            assert not (m.state & 3) and m.words == 0  # untouched: not compacted, nothing packed
        else:
            om = s.meta()
            assert (m.width, m.min_factor, m.max_factor, m.size_in_bytes) == (om.width, om.min_factor, om.max_factor,
                                                                             om.data_size)
    want = np.concatenate([s.decode_all() for s in segs if s is not None])
    assert np.array_equal(col.decode(), want)
    got = col.scan_sum()
    wsum, wcnt = _oracle_sum([s for s in segs if s is not None])
    assert (got.sum % MOD, got.count) == (wsum, wcnt)
    # ranges made only of empty segments, and empty ranges
    assert (col.scan_sum(3, 5).sum, col.scan_sum(3, 5).count) == (0, 0)
    assert (col.scan_sum(2, 2).sum, col.scan_sum(2, 2).count) == (0, 0)
    assert len(col.decode(3, 5)) == 0 and len(col.decode(4, 4)) == 0
    got = col.scan_filter(2, 10 ** 6, 1, 2)  # LT over an empty segment
    assert (got.sum, got.count) == (0, 0)
    col.encode(1, 2)
    col.uncompact(1, 2)
    col.uncompact()
    assert np.array_equal(col.decode(), want)
    col.close()


def test_completely_empty_column(gpu_ctx):
    col = gpu_ctx.column_from_values(np.zeros(0, np.uint16), [0, 0])
    col.compact()
    assert len(col.decode()) == 0
    assert col.scan_sum().count == 0
    col.uncompact()
    col.close()


@pytest.mark.parametrize("dtype", ALL_DTYPES, ids=lambda d: np.dtype(d).name)
def test_maximum_size_segments(gpu_ctx, port_oracle, dtype):
    """Full blocks (262136 / S values: 16 scan chunks) at the narrowest, the widest packable and the slot width."""
    dt = np.dtype(dtype)
    cap = adacom.segment_capacity(dtype)
    bits = 8 * dt.itemsize
    widths = [1, bits - 1, bits]
    parts = [width_values(dtype, w, cap, 3 if w < bits else 0, 11 + w) for w in widths]
    parts.append(width_values(dtype, 5, cap - 1, 0, 99))  # ragged neighbour
    col = gpu_ctx.column_from_values(np.concatenate(parts), [len(p) for p in parts])
    col.compact()
    segs = [port_oracle.encode(p, 0) for p in parts]
    for g, (m, s) in enumerate(zip(col.meta(), segs)):
        om = s.meta()
        assert (m.width, m.min_factor, m.max_factor, m.size_in_bytes) == (om.width, om.min_factor, om.max_factor,
                                                                         om.data_size), g
        if om.repacked:
            assert col.serialize(g) == s.serialize(), g
    want = np.concatenate([s.decode_all() for s in segs])
    assert np.array_equal(col.decode(), want)
    got = col.scan_sum()
    assert (got.sum % MOD, got.count) == _oracle_sum(segs)
    for op, const in ((0, int(want[12345])), (3, int(want[7])), (4, int(want[len(want) // 2]))):
        got = col.scan_filter(op, const)
        wsum = sum(s.filter(op, const)[0] for s in segs) % MOD
        wcnt = sum(s.filter(op, const)[1] for s in segs)
        assert (got.sum % MOD, got.count) == (wsum, wcnt), (op, const)
    # last value of every full block, and the first of the ragged one
    q_seg = np.array([0, 1, 2, 3, 3])
    q_row = np.array([cap - 1, cap - 1, cap - 1, 0, cap - 2])
    assert np.array_equal(col.fetch_rows(q_seg, q_row), np.array([parts[g][r] for g, r in zip(q_seg, q_row)], dtype))
    col.uncompact()
    assert np.array_equal(col.decode(), np.concatenate(parts))
    col.close()


@pytest.mark.parametrize("dtype", [np.int8, np.uint8, np.int16, np.uint32, np.int64, np.uint64],
                         ids=lambda d: np.dtype(d).name)
def test_constant_minus_one_and_all_null_segments(gpu_ctx, port_oracle, dtype):
    """range 0 -> w = 1 (hi(0) = 0); an all-(-1) signed segment has min_factor == UINT64_MAX == "unset" (Q7);
    an all-NULL segment keeps min unset and stores NullValue<T>()."""
    dt = np.dtype(dtype)
    n = 3000
    const = np.full(n, 42, dtype)
    extreme = np.full(n, -1 if dt.kind == "i" else np.iinfo(dt).max, dtype)
    zeros = np.zeros(n, dtype)
    nulls = np.arange(n).astype(dtype)
    parts = [const, extreme, zeros, nulls]
    validity = np.ones(4 * n, dtype=bool)
    validity[3 * n:] = False
    col = gpu_ctx.column_from_values(np.concatenate(parts), [n] * 4)
    col.compact(validity=_validity_words(validity))  # column-wide words for the C ABI ...
    segs = [port_oracle.encode(p, 0, validity=_validity_words(validity[g * n:(g + 1) * n]))  # ... per segment here
            for g, p in enumerate(parts)]
    for g, (m, s) in enumerate(zip(col.meta(), segs)):
        om = s.meta()
        assert (m.width, m.min_factor, m.max_factor, m.size_in_bytes) == (om.width, om.min_factor, om.max_factor,
                                                                         om.data_size), g
        if om.repacked:
            assert col.serialize(g) == s.serialize(), g
    want = np.concatenate([s.decode_all() for s in segs])
    assert np.array_equal(col.decode(), want)
    got = col.scan_sum()
    assert (got.sum % MOD, got.count) == _oracle_sum(segs)
    col.close()


@pytest.mark.parametrize("dtype", [np.int8, np.int16, np.int32, np.int64, np.uint32],
                         ids=lambda d: np.dtype(d).name)
def test_fused_sum_modes_signs_and_zones(gpu_ctx, port_oracle, dtype):
    """The fused SUM takes three paths (field sums inside a linear zone, raw slots, generic per-value); all must
    agree with the oracle for negative frames, sign-mixed segments (never re-packed by the reference's unsigned
    min/max), the literal Q1 mode and the zero-extended adaptive mode."""
    dt = np.dtype(dtype)
    bits = 8 * dt.itemsize
    info = np.iinfo(dt)
    rng = np.random.default_rng(5)
    n = 5000

    def draw(lo, hi):
        return rng.integers(lo, hi, size=n, dtype=np.int64 if dt.kind == "i" or bits < 64 else np.uint64).astype(dtype)

    parts = [draw(10, 1000)]
    if dt.kind == "i":
        parts += [draw(-1000, -10), draw(-50, 50), draw(int(info.min), int(info.min) + 200),
                  draw(int(info.max) - 200, int(info.max))]
    else:
        parts += [draw(int(info.max) - 200, int(info.max)), draw(0, 3)]
    counts = [len(p) for p in parts]
    for flags in (0, LITERAL, ZERO_EXT):
        col = gpu_ctx.column_from_values(np.concatenate(parts), counts)
        col.compact(flags=flags)
        if flags & ZERO_EXT:
            # BitCompressFromUncompressed (column_segment.cpp:385-456): min/max over zero-extended slots; the build
            # keeps the frame of reference (Q2), which is the oracle's non-literal adaptive mode
            segs = [port_oracle.encode_adaptive(p, 0) for p in parts]
            rflags = 0
        else:
            segs = [port_oracle.encode(p, flags) for p in parts]
            rflags = flags
        for g, (m, s) in enumerate(zip(col.meta(), segs)):
            om = s.meta()
            assert (m.width, bool(m.state & 2), m.min_factor, m.max_factor) == (om.width, om.repacked, om.min_factor,
                                                                               om.max_factor), (flags, g)
            if om.repacked:
                assert np.array_equal(col.segment_words(g), s.words()), (flags, g)
        if flags & ZERO_EXT:
            # Q8: where the width does not shrink the reference copies nothing and its scans return the zero-filled
            # vector; the build keeps the values (re-packed segments: oracle == original values anyway)
            for p, s in zip(parts, segs):
                if s.meta().repacked:
                    assert np.array_equal(s.decode_all(0), p)
            want = np.concatenate(parts)
            want_sum = (sum(int(x) for p in parts for x in p.tolist()) % MOD, sum(counts))
        else:
            want = np.concatenate([s.decode_all(rflags) for s in segs])
            want_sum = _oracle_sum(segs, rflags)
        assert np.array_equal(col.decode(flags=rflags), want), flags
        got = col.scan_sum(flags=rflags)
        assert (got.sum % MOD, got.count) == want_sum, flags
        if flags & ZERO_EXT:
            col.close()
            continue
        for b, e in ((1, 2), (0, len(parts) - 1)):
            got = col.scan_sum(b, e, flags=rflags)
            assert (got.sum % MOD, got.count) == _oracle_sum(segs[b:e], rflags), (flags, b, e)
        col.close()


def test_wide_odd_widths_every_chunk_shape(gpu_ctx, port_oracle):
    """From 43 bits up a scan chunk is exactly 2048 values; below it is 128*floor(1024/w): sums over segments
    whose last chunk is short, one value, or absent, for widths on both sides of the switch."""
    dtype = np.int64
    cases = [(w, n) for w in (41, 42, 43, 44, 47, 53, 59, 61, 63) for n in (1, 2047, 2048, 2049, 4096, 32767)]
    vals = [width_values(dtype, w, n, 17, seed=i) for i, (w, n) in enumerate(cases)]
    col = gpu_ctx.column_from_values(np.concatenate(vals), [len(v) for v in vals])
    col.compact()
    segs = [port_oracle.encode(v, 0) for v in vals]
    for m, s, c in zip(col.meta(), segs, cases):
        assert m.width == s.meta().width == (c[0] if c[1] > 1 else 1)  # a single value has range 0 -> w = 1
    got = col.scan_sum()
    assert (got.sum % MOD, got.count) == _oracle_sum(segs)
    for g in range(0, len(cases), 5):
        got = col.scan_sum(g, g + 1)
        assert (got.sum % MOD, got.count) == _oracle_sum(segs[g:g + 1]), cases[g]
    const = int(vals[7][3])
    got = col.scan_filter(5, const)  # GE
    assert got.count == sum(s.filter(5, const)[1] for s in segs)
    assert got.sum % MOD == sum(s.filter(5, const)[0] for s in segs) % MOD
    col.close()


def test_reencode_cycles_change_widths(gpu_ctx, port_oracle):
    """compact -> uncompact -> overwrite the raw values -> compact again: descriptors, chunk counts and payloads
    follow the new widths (the hot/cold cycle of the adaptive policy, column_segment_catalog.cpp:64-116)."""
    dtype = np.uint32
    n_seg, n = 6, 40000
    col = None
    for rnd, w in enumerate((3, 29, 11, 32, 1)):
        vals = [width_values(dtype, w, n - 7 * g, g, seed=100 * rnd + g) for g in range(n_seg)]
        flat = np.concatenate(vals)
        if col is None:
            col = gpu_ctx.column_from_values(flat, [len(v) for v in vals])
        else:
            col.uncompact()
            col.upload_raw(flat)
        col.compact()
        col.encode()  # already compacted: no-op (column_segment.cpp:278)
        segs = [port_oracle.encode(v, 0) for v in vals]
        for g, (m, s) in enumerate(zip(col.meta(), segs)):
            assert m.width == s.meta().width and m.size_in_bytes == s.meta().data_size, (w, g)
            if s.meta().repacked:
                assert col.serialize(g) == s.serialize()
        assert np.array_equal(col.decode(), flat)
        got = col.scan_sum()
        assert (got.sum % MOD, got.count) == _oracle_sum(segs)
    col.close()


def test_argument_errors_are_reported(gpu_ctx):
    col = gpu_ctx.column_from_values(np.arange(1000, dtype=np.int32), [600, 400])
    col.compact()
    for call in (lambda: col.decode(1, 3), lambda: col.scan_sum(2, 1), lambda: col.scan_partial(0, 590, 20),
                 lambda: col.fetch_rows([2], [0]), lambda: col.fetch_rows([1], [400]), lambda: col.scan_filter(6, 1)):
        with pytest.raises(adacom.AdacomError):
            call()
    with pytest.raises(adacom.AdacomError):
        gpu_ctx.column_from_values(np.arange(10, dtype=np.int64), [10], seg_capacity=5)  # count > capacity
    assert col.scan_sum().count == 1000  # the context is still usable afterwards
    # large batches are validated by the kernel (the lowest offending query is reported): a bad segment, a bad row
    seg = np.zeros(5000, dtype=np.uint32)
    row = (np.arange(5000) % 600).astype(np.uint32)
    for bad_q, (bs, br) in ((4321, (2, 0)), (77, (1, 400)), (4999, (0, 600))):
        s2, r2 = seg.copy(), row.copy()
        s2[bad_q], r2[bad_q] = bs, br
        with pytest.raises(adacom.AdacomError) as e:
            col.fetch_rows(s2, r2)
        assert f"query {bad_q}" in str(e.value)
    assert np.array_equal(col.fetch_rows(seg, row), row.astype(np.int32))  # and the next valid batch is served
    col.close()


def test_concurrent_readers_and_a_recompacting_writer(gpu_ctx, port_oracle):
    """SURVEY 8(b) threading: scans / fetches arrive from N worker threads while the policy thread compacts and
    uncompacts the same segments (column_segment_catalog.cpp:64-116).  Every read must see either representation
    completely -- the values never change."""
    import threading

    dtype = np.int32
    vals = [width_values(dtype, 5 + 3 * g, 20000 + 11 * g, 1000 * g, seed=g) for g in range(8)]
    flat = np.concatenate(vals)
    starts = np.concatenate([[0], np.cumsum([len(v) for v in vals])])
    col = gpu_ctx.column_from_values(flat, [len(v) for v in vals])
    col.compact()
    want_sum = int(flat.astype(np.int64).sum())
    errors, stop = [], threading.Event()

    def reader(seed):
        rng = np.random.default_rng(seed)
        try:
            while not stop.is_set():
                g = int(rng.integers(0, len(vals)))
                start = int(rng.integers(0, len(vals[g]) - 100))
                got = col.scan_partial(g, start, 100)
                if not np.array_equal(got, vals[g][start:start + 100]):
                    errors.append(("scan_partial", g, start))
                rows = rng.integers(0, len(vals[g]), size=64)
                if not np.array_equal(col.fetch_rows(np.full(64, g), rows), vals[g][rows]):
                    errors.append(("fetch_rows", g))
                s = col.scan_sum()
                if (s.sum, s.count) != (want_sum, len(flat)):
                    errors.append(("scan_sum", s.sum, s.count))
                if not np.array_equal(col.decode(g, g + 1), flat[starts[g]:starts[g + 1]]):
                    errors.append(("decode", g))
        except Exception as exc:  # noqa: BLE001 -- reported below
            errors.append(("exception", repr(exc)))

    threads = [threading.Thread(target=reader, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    try:
        for rnd in range(30):
            b = rnd % 5
            col.uncompact(b, b + 3)
            col.compact(b, b + 3)
    finally:
        stop.set()
        for t in threads:
            t.join()
    assert not errors, errors[:5]
    assert np.array_equal(col.decode(), flat)
    col.close()


@pytest.mark.parametrize("dtype", [np.uint8, np.int16, np.uint32, np.int32, np.int64], ids=lambda d: np.dtype(d).name)
def test_point_lookups_hits_misses_duplicates(gpu_ctx, dtype):
    """adacom_point_filter_batch == `value = k` on one segment (the reference's zipf_distribution.cpp look-up):
    byte-aligned widths take the SIMD field compare, the others the generic path; keys that are absent, below the
    frame of reference, beyond the field range or duplicated must all give numpy's count and first row."""
    dt = np.dtype(dtype)
    bits = 8 * dt.itemsize
    rng = np.random.default_rng(9)
    info = np.iinfo(dt)
    segs = []
    for w in sorted({1, 7, 8, 13, 16, 24, 32, bits - 1, bits} & set(range(1, bits + 1))):
        for base in (0, 1000, int(info.min) if dt.kind == "i" else 3):
            n = int(rng.integers(3000, 9000))
            lo = min(max(int(info.min), base), int(info.max) - 1)
            hi = min(int(info.max), lo + (1 << w) - 1) if w < 64 else int(info.max)
            v = rng.integers(lo, hi, size=n, endpoint=True, dtype=np.int64 if bits < 64 or dt.kind == "i" else np.uint64)
            v = v.astype(dtype)
            v[: n // 50] = v[0]  # duplicates
            segs.append(v)
    col = gpu_ctx.column_from_values(np.concatenate(segs), [len(v) for v in segs])
    col.compact()
    q_seg, q_key = [], []
    for g, v in enumerate(segs):
        present = [int(v[0]), int(v[-1]), int(v[len(v) // 2])]
        absent = [int(v.min()) - 1, int(v.max()) + 1, int(info.min), int(info.max), 0]
        for k in present + absent:
            if int(info.min) <= k <= int(info.max):
                q_seg.append(g)
                q_key.append(k)
    cnt, first = col.point_filter_batch(np.array(q_seg), np.array(q_key, dtype=np.int64 if dt.kind == "i" or bits < 64 else np.uint64).astype(np.int64))
    for j, (g, k) in enumerate(zip(q_seg, q_key)):
        hit = np.nonzero(segs[g] == np.array(k).astype(dtype))[0]
        assert cnt[j] == len(hit), (g, k, col.meta(g, g + 1)[0].width)
        if len(hit):
            assert first[j] == hit[0], (g, k)
    col.close()


def test_encode_rolls_back_when_the_extent_allocation_fails(gpu_ctx, port_oracle, monkeypatch):
    """An encode whose packed extent cannot be allocated (ADACOM_FAULT_ENCODE_ALLOC: the allocator "runs dry" after
    layout_kernel has already rewritten the descriptors) must leave every segment as it was -- not compacted, slot
    width, readable from the raw arena -- and the next encode must produce exactly the payloads of a clean one."""
    dtype = np.int32
    vals = [width_values(dtype, w, n, base, seed=g) for g, (w, n, base) in
            enumerate([(5, 40000, 7), (32, 3000, -5), (13, 65534, 1000), (1, 1, 0), (20, 2049, -(1 << 30))])]
    flat = np.concatenate(vals)
    counts = [len(v) for v in vals]
    col = gpu_ctx.column_from_values(flat, counts)
    col.analyze()
    before = [(m.width, m.state, m.size_in_bytes) for m in col.meta()]
    for attempt in ("range", "list"):
        monkeypatch.setenv("ADACOM_FAULT_ENCODE_ALLOC", "1")
        with pytest.raises(adacom.AdacomError) as err:
            if attempt == "range":
                col.encode()
            else:
                col.compact_list([0, 2, 4], analyze=False)
        assert err.value.status == 3 and "rolled back" in str(err.value)  # ADACOM_ERR_NOMEM
        monkeypatch.delenv("ADACOM_FAULT_ENCODE_ALLOC")
        assert [(m.width, m.state, m.size_in_bytes) for m in col.meta()] == before, attempt
        assert np.array_equal(col.decode(), flat)
        got = col.scan_sum()
        assert (got.sum, got.count) == (int(flat.astype(object).sum()), len(flat))
        assert col.scan_filter(2, 100).count == int((flat < 100).sum())
    col.encode()  # the same extent id is used again: nothing stale may be bound
    segs = [port_oracle.encode(v, 0) for v in vals]
    for g, (m, s) in enumerate(zip(col.meta(), segs)):
        assert (m.width, m.size_in_bytes) == (s.meta().width, s.meta().data_size), g
        if s.meta().repacked:
            assert col.serialize(g) == s.serialize(), g
    assert np.array_equal(col.decode(), flat)
    # ... and a failure on an already partly compacted column only reverts the segments that call touched
    col.uncompact(1, 4)
    col.analyze(1, 4)
    mid = [(m.width, m.state) for m in col.meta()]
    monkeypatch.setenv("ADACOM_FAULT_ENCODE_ALLOC", "1")
    with pytest.raises(adacom.AdacomError):
        col.encode(0, 5)
    monkeypatch.delenv("ADACOM_FAULT_ENCODE_ALLOC")
    assert [(m.width, m.state) for m in col.meta()] == mid
    assert np.array_equal(col.decode(), flat)
    col.encode(0, 5)
    assert np.array_equal(col.decode(), flat)
    for g, (m, s) in enumerate(zip(col.meta(), segs)):
        assert (m.width, m.size_in_bytes) == (s.meta().width, s.meta().data_size), g
    col.close()


def test_many_encode_cycles_reuse_extent_slots(gpu_ctx):
    """hot/cold cycles of the policy on one column: the extent table must not grow with the number of encodes and
    the payloads stay right (dead slots are reused; ids repeat)"""
    dtype = np.uint16
    rng = np.random.default_rng(3)
    counts = [5000, 131068, 1, 70000]
    flat = rng.integers(100, 100 + (1 << 9), size=sum(counts)).astype(dtype)
    col = gpu_ctx.column_from_values(flat, counts)
    col.compact()
    want = int(flat.astype(object).sum())
    for it in range(200):
        segs = sorted(rng.choice(len(counts), size=int(rng.integers(1, len(counts) + 1)), replace=False).tolist())
        col.uncompact_list(segs)
        if it % 3 == 0:
            col.meta()  # refresh the mirror: gc of dead extents
        col.compact_list(segs)
        if it % 10 == 0:
            got = col.scan_sum()
            assert (got.sum, got.count) == (want, len(flat)), it
    assert np.array_equal(col.decode(), flat)
    assert all(m.state & 2 for m in col.meta())
    col.close()


def test_from_packed_rejects_malformed_metas(gpu_ctx):
    dtype = np.uint32
    vals = width_values(dtype, 11, 5000, 3, seed=1)
    col = gpu_ctx.column_from_values(vals, [3000, 2000])
    col.compact()
    metas, words = col.meta(), col.download_packed()
    good = gpu_ctx.column_from_packed(dtype, metas, words)
    assert np.array_equal(good.decode(), vals)
    good.close()

    def broken(**kw):
        ms = col.meta()
        for k, v in kw.items():
            setattr(ms[1], k, v)
        return ms

    for ms in (broken(width=0), broken(width=32), broken(width=64), broken(state=0x13), broken(state=2),
               broken(count=70000), broken(word_offset=metas[1].word_offset + 1), broken(word_offset=1 << 40)):
        with pytest.raises(adacom.AdacomError) as err:
            gpu_ctx.column_from_packed(dtype, ms, words)
        assert err.value.status == 1
    assert col.scan_sum().count == 5000
    col.close()
