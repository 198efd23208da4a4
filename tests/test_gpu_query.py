"""GPU parity of the fused multi-column scan (adacom_scan_columns) and the TPC-H Q6 / Q1 plans built on it.

Checker: numpy on the uncompressed columns == what DuckDB's FilterSelection + SUM compute on the decoded vectors
(column_segment.cpp:575-658); the per-column payloads are already pinned against the oracle in test_gpu_parity."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

import duckdb_adaptive_compression_b200 as adacom
from duckdb_adaptive_compression_b200 import _capi, tpch


@pytest.fixture(scope="module")
def table(gpu_ctx):
    cols = tpch.lineitem(3, scale_factor=0.05, seed=7)
    n = 3 * adacom.ROW_GROUP_SIZE - 12345  # ragged last row group
    cols = {k: v[:n].copy() for k, v in cols.items()}
    t = tpch.LineitemTable(gpu_ctx, {k: cols[k] for k in set(tpch.LineitemTable.Q6_COLUMNS + tpch.LineitemTable.Q1_COLUMNS)})
    yield cols, t
    t.close()


def test_q6_matches_numpy(table):
    cols, t = table
    want_sum, want_count = tpch.q6_expected(cols)
    got = t.q6()
    assert (got.sum, got.count) == (want_sum, want_count)
    # widths the survey measured on the reference's SF0.1 run: shipdate ~12, discount 4, quantity 6 bits
    assert t.cols["l_discount"].meta()[0].width == 4 and t.cols["l_quantity"].meta()[0].width == 6


def test_grouped_scan_matches_per_group_passes_on_ragged_data(gpu_ctx):
    """GROUP BY in the fused scan: keys outside 0..n_groups-1 are dropped, groups may be empty, the predicate and the
    bitmap still apply; compared with numpy and with one filtered pass per group."""
    rng = np.random.default_rng(3)
    n = 3 * 122880 + 777
    key = rng.integers(0, 7, size=n).astype(np.uint8)          # 7 distinct values, only 0..4 are groups
    key[rng.random(n) < 0.5] = 2                                # skewed; group 4 stays rare
    key[key == 3] = 6                                           # group 3 empty
    a = rng.integers(-1000, 1000, size=n).astype(np.int32)
    b = rng.integers(0, 1 << 40, size=n).astype(np.int64)
    c = rng.integers(0, 50, size=n).astype(np.int16)
    cols = {}
    for name, v in (("key", key), ("a", a), ("b", b), ("c", c)):
        cols[name] = gpu_ctx.column_from_values(v, adacom.row_group_segment_counts(n, v.dtype))
        cols[name].compact()
    terms = [(cols["c"], (5, 40)), (cols["key"], None), (cols["a"], None), (cols["b"], None)]
    sums = [[(2, 0, 1)], [(3, 0, 1), (2, 7, -1)], []]
    got = adacom.scan_columns(gpu_ctx, terms, sums, 0, n, group_by=(1, 5))
    sel = (c >= 5) & (c <= 40)
    for g in range(5):
        m = sel & (key == g)
        want = [int(a[m].astype(object).sum()), int((b[m].astype(object) * (7 - a[m].astype(object))).sum()), int(m.sum())]
        assert [r.sum for r in got[g]] == want, g
        assert all(r.count == int(m.sum()) for r in got[g]), g
        per = adacom.scan_columns(gpu_ctx, [(cols["c"], (5, 40)), (cols["key"], (g, g)), (cols["a"], None), (cols["b"], None)],
                                  sums, 0, n)
        assert [r.sum for r in per] == want and per[0].count == int(m.sum())
    # a sub-range that starts inside the table
    r0, r1 = 2048 * 31, n - 5
    got = adacom.scan_columns(gpu_ctx, terms, sums, r0, r1, group_by=(1, 5))
    for g in range(5):
        m = sel[r0:r1] & (key[r0:r1] == g)
        assert got[g][0].sum == int(a[r0:r1][m].astype(object).sum()) and got[g][2].sum == int(m.sum())
    # a product that needs more than 64 bits switches the whole query to 128-bit products
    wide = sums + [[(3, 0, 1), (3, 5, 1), (2, 0, 1)]]
    got = adacom.scan_columns(gpu_ctx, terms, wide, 0, n, group_by=(1, 5))
    flat = adacom.scan_columns(gpu_ctx, terms, wide, 0, n)
    for g in range(5):
        m = sel & (key == g)
        bo, ao = b[m].astype(object), a[m].astype(object)
        assert got[g][3].sum == int((bo * (5 + bo) * ao).sum()), g
        assert got[g][0].sum == int(ao.sum())
    bo, ao = b[sel].astype(object), a[sel].astype(object)
    assert flat[3].sum == int((bo * (5 + bo) * ao).sum()) and flat[3].count == int(sel.sum())
    with pytest.raises(adacom.AdacomError):
        adacom.scan_columns(gpu_ctx, terms, sums, 0, n, group_by=(1, 9))
    for col in cols.values():
        col.close()


@pytest.mark.parametrize("wide_values", [False, True], ids=["int32_values", "values_beyond_int32"])
def test_grouped_factor_program_chains_and_duplicates(gpu_ctx, wide_values):
    """The dense grouped kernel evaluates the sums as one factor program in which an aggregate that extends another's
    factor list continues its term: shared prefixes in any order, the same product twice, COUNT(*) between them, a
    product over the same column twice, unrelated sums.  Narrow columns so that the dense kernel (32-bit values) runs."""
    rng = np.random.default_rng(11)
    n = 5 * 122880 + 4321
    key = rng.integers(0, 4, size=n).astype(np.uint8)
    # values_beyond_int32: the same kernel in its 64-bit arithmetic mode (values above 2^31, still <= 32-bit payloads)
    p = (rng.integers(900, 100000, size=n) + (3_000_000_000 if wide_values else 0)).astype(np.uint32 if wide_values else np.int32)
    d = rng.integers(0, 11, size=n).astype(np.uint8)
    t = rng.integers(0, 9, size=n).astype(np.uint8)
    s = rng.integers(0, 3000, size=n).astype(np.int16)
    cols = {}
    for name, v in (("key", key), ("p", p), ("d", d), ("t", t), ("s", s)):
        cols[name] = gpu_ctx.column_from_values(v, adacom.row_group_segment_counts(n, v.dtype))
        cols[name].compact()
    terms = [(cols["s"], (0, 2700)), (cols["key"], None), (cols["p"], None), (cols["d"], None), (cols["t"], None)]
    P, D, T = (2, 0, 1), (3, 100, -1), (4, 100, 1)
    sums = [[P, D, T], [], [P], [P, D], [(3, 0, 1)], [P, D], [(3, 0, 1), (3, 0, 1)], [T, P]]
    sel = s <= 2700
    po, do, to = p.astype(object), d.astype(object), t.astype(object)
    exprs = [po * (100 - do) * (100 + to), None, po, po * (100 - do), do, po * (100 - do), do * do, (100 + to) * po]
    for rows in ((0, n), (2048 * 7, n - 3)):
        got = adacom.scan_columns(gpu_ctx, terms, sums, rows[0], rows[1], group_by=(1, 4))
        for g in range(4):
            m = sel & (key == g)
            m[:rows[0]] = False
            m[rows[1]:] = False
            want = [int(m.sum()) if e is None else int(e[m].sum()) for e in exprs]
            assert [r.sum for r in got[g]] == want, (rows, g)
            assert all(r.count == int(m.sum()) for r in got[g])
    for col in cols.values():
        col.close()


def test_q1_matches_numpy(table):
    cols, t = table
    assert t.q1() == tpch.q1_expected(cols)
    assert t.q1_per_group_passes() == tpch.q1_expected(cols)


def test_bitmaps_and_row_ranges(gpu_ctx, table):
    import torch
    cols, t = table
    n = t.n_rows
    lo, hi = 4096, n - 777
    words = (hi - lo + 63) // 64
    bm = torch.zeros(words + 1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    c = t.cols
    # pass 1: shipdate filter -> bitmap; pass 2: bitmap AND discount -> sums
    r1 = adacom.scan_columns(gpu_ctx, [(c["l_shipdate"], (tpch.D_1994_01_01, tpch.D_1995_01_01 - 1))], [[]], lo, hi,
                             bitmap_out=bm.data_ptr())
    sel1 = (cols["l_shipdate"][lo:hi] >= tpch.D_1994_01_01) & (cols["l_shipdate"][lo:hi] < tpch.D_1995_01_01)
    assert r1[0].count == int(sel1.sum()) and r1[0].sum == int(sel1.sum())
    gpu_ctx.synchronize()
    bits = np.unpackbits(bm.cpu().numpy().view(np.uint8), bitorder="little")[: hi - lo].astype(bool)
    assert np.array_equal(bits, sel1)
    r2 = adacom.scan_columns(gpu_ctx, [(c["l_discount"], (5, 7)), (c["l_extendedprice"], None)],
                             [[(1, 0, 1)], [(1, 0, 1), (0, 100, -1)]], lo, hi, bitmap_in=bm.data_ptr())
    sel2 = sel1 & (cols["l_discount"][lo:hi] >= 5) & (cols["l_discount"][lo:hi] <= 7)
    p = cols["l_extendedprice"][lo:hi][sel2].astype(object)
    d = cols["l_discount"][lo:hi][sel2].astype(object)
    assert r2[0].count == int(sel2.sum())
    assert r2[0].sum == int(p.sum()) and r2[1].sum == int((p * (100 - d)).sum())


def test_mixed_types_signs_and_raw_segments(gpu_ctx):
    rng = np.random.default_rng(3)
    n = adacom.ROW_GROUP_SIZE + 50000
    a = rng.integers(-5000, 5000, size=n).astype(np.int16)
    b = (rng.integers(0, 1 << 40, size=n).astype(np.uint64) + np.uint64((1 << 63) + 12345))
    c = rng.integers(-(1 << 62), 1 << 62, size=n).astype(np.int64)  # mixed signs: width 64, never re-packed (Q1 domain)
    d = rng.integers(0, 200, size=n).astype(np.uint8)
    cols = []
    for v in (a, b, c, d):
        col = gpu_ctx.column_from_values(v, adacom.row_group_segment_counts(n, v.dtype))
        cols.append(col)
    cols[0].compact(); cols[1].compact(); cols[2].compact()  # d stays raw (never compacted)
    assert not (cols[2].meta()[0].state & _capi.SEG_REPACKED)
    b_lo, b_hi = (1 << 63) + 12345 + (1 << 38), (1 << 63) + 12345 + (1 << 39)
    terms = [(cols[0], (-100, 2500)), (cols[1], (b_lo, b_hi)), (cols[2], None), (cols[3], (10, 150))]
    sums = [[], [(2, 0, 1)], [(2, 7, -3), (0, 1, 1)], [(0, 0, 1), (3, 0, 1), (2, 0, 1)]]
    got = adacom.scan_columns(gpu_ctx, terms, sums, 2048, n)
    s = slice(2048, n)
    sel = (a[s] >= -100) & (a[s] <= 2500) & (b[s] >= np.uint64(b_lo)) & (b[s] <= np.uint64(b_hi)) & (d[s] >= 10) & (d[s] <= 150)
    ao, co, do = a[s][sel].astype(object), c[s][sel].astype(object), d[s][sel].astype(object)
    def wrap(x):
        x %= 1 << 128
        return x - (1 << 128) if x >= 1 << 127 else x
    assert got[0].count == int(sel.sum()) and got[0].sum == int(sel.sum())
    assert got[1].sum == wrap(int(co.sum()))
    assert got[2].sum == wrap(int(((7 - 3 * co) * (1 + ao)).sum()))
    assert got[3].sum == wrap(int((ao * do * co).sum()))
    for col in cols:
        col.close()


def test_rejects_blocks_spanning_too_many_segments(gpu_ctx):
    v = np.arange(8192, dtype=np.int32)
    col = gpu_ctx.column_from_values(v, [500] * 16 + [192], seg_capacity=500)
    col.compact()
    with pytest.raises(adacom.AdacomError) as e:
        adacom.scan_columns(gpu_ctx, [(col, (0, 10))], [[]], 0, 8192)
    assert e.value.status == _capi.ERR_UNSUPPORTED
    col.close()
