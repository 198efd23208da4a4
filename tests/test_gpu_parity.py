"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on identical inputs.

Bit-exact bar (integer/byte work): packed payload words, per-segment metadata, decoded values, aggregates.
"""
import numpy as np
import pytest

import oracle as orc
from tests.cases import ALL_DTYPES, sweep_cases, width_values

pytestmark = pytest.mark.gpu

FLAG_SETS = [0, orc.PADDED, orc.LITERAL]  # oracle flag values equal the C ABI's (PADDED=1, LITERAL=2)


def build_column(gpu_ctx, dtype, cases):
    vals = [width_values(dtype, w, n, base, seed=i) for i, (w, n, base) in enumerate(cases)]
    counts = [len(v) for v in vals]
    col = gpu_ctx.column_from_values(np.concatenate(vals), counts)
    return col, vals


@pytest.mark.parametrize("dtype", ALL_DTYPES, ids=lambda d: np.dtype(d).name)
@pytest.mark.parametrize("flags", FLAG_SETS)
def test_encode_payload_meta_decode(gpu_ctx, port_oracle, dtype, flags):
    cases = sweep_cases(dtype)
    col, vals = build_column(gpu_ctx, dtype, cases)
    col.compact(flags=flags)
    metas = col.meta()
    arena = col.download_packed()
    decoded = col.decode(flags=flags)
    off = 0
    for g, v in enumerate(vals):
        seg = port_oracle.encode(v, flags)
        om = seg.meta()
        m = metas[g]
        assert (m.count, m.width, m.min_factor, m.max_factor) == (om.count, om.width, om.min_factor, om.max_factor), \
            (cases[g], m.width, om.width)
        assert bool(m.state & 1) == om.compacted and bool(m.state & 2) == om.repacked
        assert m.size_in_bytes == om.data_size, cases[g]
        if om.repacked:
            assert m.words == om.words
            assert np.array_equal(arena[m.word_offset:m.word_offset + m.words], seg.words()), cases[g]
            assert col.serialize(g) == seg.serialize()
        want = seg.decode_all(flags)
        assert np.array_equal(decoded[off:off + len(v)], want), cases[g]
        off += len(v)
    col.close()


@pytest.mark.parametrize("dtype", [np.int8, np.uint16, np.int32, np.uint32, np.int64, np.uint64],
                         ids=lambda d: np.dtype(d).name)
def test_scan_partial_fetch_sum_filter(gpu_ctx, port_oracle, dtype):
    rng = np.random.default_rng(7)
    S = np.dtype(dtype).itemsize
    cases = [(w, n, b) for (w, n, b) in sweep_cases(dtype, counts=[1, 65, 2049, 5000]) if w % 3 == 1 or w == 8 * S]
    col, vals = build_column(gpu_ctx, dtype, cases)
    col.compact()
    segs = [port_oracle.encode(v, 0) for v in vals]
    # scan_partial with arbitrary start / count / result_offset
    for g in rng.choice(len(vals), size=40, replace=False) if len(vals) > 40 else range(len(vals)):
        n = len(vals[g])
        start = int(rng.integers(0, n))
        cnt = int(rng.integers(0, n - start + 1))
        roff = int(rng.integers(0, 5))
        got = col.scan_partial(g, start, cnt, result_offset=roff)
        want = segs[g].scan_partial(start, cnt, result_offset=roff)
        assert np.array_equal(got, want), (cases[g], start, cnt, roff)
    # fetch_rows
    q_seg = rng.integers(0, len(vals), size=3000)
    q_row = np.array([rng.integers(0, len(vals[g])) for g in q_seg])
    got = col.fetch_rows(q_seg, q_row)
    want = np.array([segs[g].fetch_row(int(r)) for g, r in zip(q_seg, q_row)], dtype=dtype)
    assert np.array_equal(got, want)
    # fused SUM over everything and over sub-ranges
    want_sum = sum(s.sum()[0] for s in segs)
    mod = 1 << 128
    got = col.scan_sum()
    assert got.count == sum(len(v) for v in vals)
    assert got.sum % mod == want_sum % mod
    b, e = 3, min(len(vals), 17)
    got = col.scan_sum(b, e)
    assert got.sum % mod == sum(s.sum()[0] for s in segs[b:e]) % mod
    # fused filter, all six operators, with the selection bitmap checked through counts + sums
    allv = np.concatenate([s.decode_all() for s in segs])
    for op in range(6):
        const = int(allv[int(rng.integers(0, len(allv)))])
        got = col.scan_filter(op, const)
        wsum, wcnt = 0, 0
        for s in segs:
            sm, c, _ = s.filter(op, const)
            wsum += sm
            wcnt += c
        assert got.count == wcnt, (op, const)
        assert got.sum % mod == wsum % mod, (op, const)
    # uncompact: back to flat values, sizes follow the reference accounting
    col.uncompact()
    assert np.array_equal(col.decode(), allv)
    for m in col.meta():
        assert not (m.state & 3)
    col.close()


def test_validity_nulls(gpu_ctx, port_oracle):
    rng = np.random.default_rng(11)
    dtype = np.int32
    n = 6000
    v = (rng.integers(0, 1 << 13, size=n) + 1000).astype(dtype)
    valid = rng.random(n) > 0.2
    words = np.zeros((n + 63) // 64, dtype=np.uint64)
    for i in np.nonzero(valid)[0]:
        words[i >> 6] |= np.uint64(1) << np.uint64(i & 63)
    col = gpu_ctx.column_from_values(v, [n])
    col.compact(validity=words)
    seg = port_oracle.encode(v, 0, validity=words)
    m, om = col.meta()[0], seg.meta()
    assert (m.width, m.min_factor, m.max_factor) == (om.width, om.min_factor, om.max_factor)
    assert np.array_equal(col.segment_words(0), seg.words())
    assert np.array_equal(col.decode(), seg.decode_all())
    col.close()


def test_point_filter_batch_sequential(gpu_ctx):
    """config-1 shape: sequential UINTEGER column, point look-ups `i == k` -> exactly one row each."""
    import duckdb_adaptive_compression_b200 as adacom
    n = 1_000_000
    v = np.arange(n, dtype=np.uint32)
    counts = adacom.row_group_segment_counts(n, np.uint32)
    col = gpu_ctx.column_from_values(v, counts)
    col.compact()
    starts = np.concatenate([[0], np.cumsum(counts)[:-1]]).astype(np.int64)
    rng = np.random.default_rng(3)
    keys = rng.integers(0, n, size=2000)
    seg = np.searchsorted(starts, keys, side="right") - 1
    cnt, first = col.point_filter_batch(seg, keys)
    assert np.all(cnt == 1)
    assert np.array_equal(first.astype(np.int64), keys - starts[seg])
    col.close()
