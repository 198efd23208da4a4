"""Segment-list forms of analyze / encode / uncompact (the access-frequency policy's subsets,
column_segment_catalog.cpp:64-116) through the C ABI: same payloads, metadata and values as the range forms and as
the oracle, one launch set per call.  Bit-exact."""
from __future__ import annotations

import numpy as np
import pytest

import duckdb_adaptive_compression_b200 as adacom
from duckdb_adaptive_compression_b200 import _capi
from tests.cases import width_values

pytestmark = pytest.mark.gpu


def _column(gpu_ctx, dtype, n_seg=23):
    S = np.dtype(dtype).itemsize
    bits = 8 * S
    parts = []
    for g in range(n_seg):
        w = 1 + (g * 5) % bits          # includes w == bits - ... and full-width segments
        n = [2049, 63, 1, 4097, 700][g % 5]
        parts.append(width_values(dtype, w, n, [0, 5, 1000003][g % 3], seed=g + 1))
    return gpu_ctx.column_from_values(np.concatenate(parts), [len(p) for p in parts]), parts


def _meta_tuple(m):
    return (m.count, m.words, m.min_factor, m.max_factor, m.size_in_bytes, m.width, m.state)


@pytest.mark.parametrize("dtype", [np.uint8, np.int16, np.uint32, np.int64])
def test_list_forms_match_range_forms_and_oracle(gpu_ctx, port_oracle, dtype):
    col, parts = _column(gpu_ctx, dtype)
    ref, _ = _column(gpu_ctx, dtype)
    values = np.concatenate(parts)
    cold = [0, 1, 4, 5, 6, 11, 17, 22]
    launches = gpu_ctx.launch_count
    col.compact_list(cold)
    assert gpu_ctx.launch_count - launches <= 4  # analyze + layout + pack + bind, whatever the number of runs
    for g in cold:
        ref.compact(g, g + 1)
    ma, mb = col.meta(), ref.meta()
    for g in range(len(parts)):
        assert _meta_tuple(ma[g]) == _meta_tuple(mb[g]), g
        if g in cold:
            o = port_oracle.encode(parts[g], 0)
            om = o.meta()
            assert (ma[g].width, ma[g].min_factor, ma[g].max_factor, ma[g].size_in_bytes) == (
                om.width, om.min_factor, om.max_factor, om.data_size)
            if om.repacked:
                assert np.array_equal(col.segment_words(g), o.words()), g
        else:
            assert not (ma[g].state & _capi.SEG_COMPACTED)
    assert np.array_equal(col.decode(), values)
    want = col.scan_sum()
    assert (want.sum, want.count) == (int(values.astype(object).sum()), len(values))

    # the rest of the column, then the hottest subset back to slot width in one call
    rest = [g for g in range(len(parts)) if g not in cold]
    col.compact_list(rest)
    assert all(m.state & _capi.SEG_COMPACTED for m in col.meta())
    assert np.array_equal(col.decode(), values)
    hot = [1, 2, 3, 11, 21, 22]
    launches = gpu_ctx.launch_count
    col.uncompact_list(hot)
    assert gpu_ctx.launch_count - launches == 2  # decode + descriptor reset
    ms = col.meta()
    for g in range(len(parts)):
        assert bool(ms[g].state & _capi.SEG_COMPACTED) == (g not in hot), g
        if g in hot:
            assert ms[g].width == 8 * np.dtype(dtype).itemsize and ms[g].words == 0
    assert np.array_equal(col.decode(), values)
    assert np.array_equal(col.decode(1, 4), np.concatenate(parts[1:4]))
    # re-compact the hot ones without a new analyze (min/max survive an uncompact): same payload as before
    col.compact_list(hot, analyze=False)
    full, _ = _column(gpu_ctx, dtype)
    full.compact()
    for a, b in zip(col.meta(), full.meta()):
        assert _meta_tuple(a) == _meta_tuple(b)
    for g in hot:
        if col.meta(g, g + 1)[0].words:
            assert np.array_equal(col.segment_words(g), full.segment_words(g))
    assert np.array_equal(col.decode(), values)
    got = col.scan_sum()
    assert (got.sum, got.count) == (want.sum, want.count)
    for c in (col, ref, full):
        c.close()


def test_list_argument_errors_and_empty(gpu_ctx):
    col, parts = _column(gpu_ctx, np.int32, n_seg=6)
    col.compact_list([])
    col.uncompact_list([])
    for bad in ([3, 3], [4, 2], [0, 6]):
        for fn in (col.compact_list, col.uncompact_list):
            with pytest.raises(adacom.AdacomError) as e:
                fn(bad)
            assert e.value.status == _capi.ERR_INVALID
    col.compact_list([0, 5])
    assert np.array_equal(col.decode(), np.concatenate(parts))
    col.close()


def test_policy_epoch_two_calls(gpu_ctx):
    """One epoch of the 90/10 policy over a 1000-segment column: two calls, a handful of launches, values intact,
    data size == the per-segment accounting."""
    n_seg = 1000
    rng = np.random.default_rng(3)
    counts = [2048] * n_seg
    values = (rng.integers(0, 1 << 11, size=2048 * n_seg, dtype=np.int64) + 77).astype(np.int32)
    col = gpu_ctx.column_from_values(values, counts)
    col.compact()
    size_all = col.data_size()
    reads = rng.permutation(n_seg)
    hot = np.sort(reads[: n_seg // 10])
    launches = gpu_ctx.launch_count
    col.uncompact_list(hot)
    assert gpu_ctx.launch_count - launches == 2
    ms = col.meta()
    assert sum(1 for m in ms if not (m.state & _capi.SEG_COMPACTED)) == len(hot)
    assert col.data_size() == sum(m.size_in_bytes for m in ms) > size_all
    assert np.array_equal(col.decode(), values)
    col.compact_list(hot, analyze=False)
    assert col.data_size() == size_all
    assert np.array_equal(col.decode(), values)
    col.close()


@pytest.mark.parametrize("dtype", [np.int16, np.uint32, np.int32, np.int64], ids=lambda d: np.dtype(d).name)
def test_point_lookups_shared_scan(gpu_ctx, dtype):
    """Batches whose segments all pack to 8 / 16 / 32 bits take the shared scan (queries counting-sorted by segment,
    a 16 KiB chunk read once per slice of 64 queries): hot segments with hundreds of queries in arbitrary order,
    multi-chunk segments, duplicates, misses, keys outside the field range, negative frames of reference --
    numpy's count and first row for every query, ONE kernel launch."""
    dt = np.dtype(dtype)
    bits = 8 * dt.itemsize
    info = np.iinfo(dt)
    rng = np.random.default_rng(11)
    cap = 262136 // dt.itemsize
    segs = []
    for w in [x for x in (8, 16, 32) if x < bits]:
        for base in (0, 1000, int(info.min) if dt.kind == "i" else 7):
            for n in (cap, 5000, 1):
                lo = base
                hi = lo + (1 << w) - 1
                v = rng.integers(lo, hi, size=n, endpoint=True, dtype=np.int64).astype(dtype)
                if n > 2:
                    v[0], v[1] = np.array(lo).astype(dtype), np.array(hi).astype(dtype)  # exact width
                    v[n // 2: n // 2 + 5] = v[0]
                segs.append(v)
    col = gpu_ctx.column_from_values(np.concatenate(segs), [len(v) for v in segs])
    col.compact()
    widths = {m.width for m in col.meta() if m.count > 2}
    assert widths <= {8, 16, 32}
    q_seg, q_key = [], []
    for g, v in enumerate(segs):
        k_here = 300 if g % 4 == 0 else 5
        picks = rng.integers(0, len(v), size=k_here)
        for i in picks:
            q_seg.append(g)
            q_key.append(int(v[i]))
        for k in (int(v.min()) - 1, int(v.max()) + 1, int(info.min), int(info.max), 0, int(v[0])):
            if int(info.min) <= k <= int(info.max):
                q_seg.append(g)
                q_key.append(k)
    order = rng.permutation(len(q_seg))
    q_seg = np.array(q_seg, dtype=np.uint32)[order]
    q_key = np.array(q_key, dtype=np.int64)[order]
    # single-value segments keep their slot width or pack to 1 bit: leave them out of the eligible batch
    keep = np.array([len(segs[g]) > 2 for g in q_seg])
    launches = gpu_ctx.launch_count
    cnt, first = col.point_filter_batch(q_seg[keep], q_key[keep])
    assert gpu_ctx.launch_count - launches == 1
    for j, (g, k) in enumerate(zip(q_seg[keep], q_key[keep])):
        hit = np.nonzero(segs[g] == np.array(k).astype(dtype))[0]
        assert cnt[j] == len(hit), (g, k, j)
        assert first[j] == (hit[0] if len(hit) else (1 << 64) - 1), (g, k, j)
    # a batch that also touches an ineligible segment falls back to the per-query kernel: same answers
    cnt2, first2 = col.point_filter_batch(q_seg, q_key)
    assert np.array_equal(cnt2[keep], cnt) and np.array_equal(first2[keep], first)
    col.close()
