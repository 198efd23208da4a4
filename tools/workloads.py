"""The reference's other workload shapes (BASELINE.json configs[0], [2], [3], [4]) as bench.py legs.

Each function runs the workload on the GPU codec (device-timed with CUDA events on the codec's stream), checks the
answers, and times the SAME workload on the reference's own SDSL code path (oracle/_ref, bench_* entry points of
oracle/ref_harness.cpp) on the box's host cores over a BOUNDED sample.  Returned dicts go into bench.py's `configs`
object: {rate, algorithmic bytes, roofline {achieved, peak, frac}, cpu_baseline {value, unit, cores, kind, sample}}.

config 1  benchmark/micro/succinct/zipf_distribution.cpp:13-61  (10^8 sequential UINTEGER, 10^4 Zipf(q=1) point filters)
config 3  run_diff_zipfian.sh:3,12-14 x zipf_distribution_diff_skews.cpp:13-111 (21 skews x {adaptive, non-adaptive
          succinct, non-succinct}; per epoch (look-ups/s, resident bytes) like the reference prints per second)
config 4  TPC-H SF10 lineitem, Q6 + Q1 fused scans on one GPU
config 5  TPC-H SF100 lineitem, row groups dealt over the ranks, Q6 partials all-gathered over NCCL
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SKEWS = [round(0.1 * i, 1) for i in range(21)]  # run_diff_zipfian.sh:3  `seq 0.0 .1 2.0`


def _ref_oracle():
    import oracle
    return oracle.Oracle("reference") if oracle.have_reference() else None


def _dev_timed(ctx, fn, steps, warmup):
    """average device time of fn() in ms between CUDA events on the codec's stream (fn may also sync)"""
    e0, e1 = ctx.event(), ctx.event()
    for _ in range(warmup):
        fn()
    ctx.synchronize()
    ctx.record(e0)
    for _ in range(steps):
        fn()
    ctx.record(e1)
    ms = ctx.elapsed_ms(e0, e1) / steps
    ctx.event_destroy(e0)
    ctx.event_destroy(e1)
    return ms


def _roof(alg_bytes, ms, peak):
    gbs = alg_bytes / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
    return {"bound": "hbm", "achieved": round(gbs, 1), "peak": peak, "unit": "GB/s", "frac": round(gbs / peak, 4),
            "algorithmic_bytes": int(alg_bytes)}


# ---------------------------------------------------------------------------------------------------------------------
# config 1
# ---------------------------------------------------------------------------------------------------------------------
def config1_zipf(ctx, peak, threads, rows=100_000_000, lookups=10_000, steps=5, warmup=3, cpu_lookups=600):
    import duckdb_adaptive_compression_b200 as adacom
    from duckdb_adaptive_compression_b200.zipf import zipf_keys
    import oracle
    dtype = np.dtype(np.uint32)
    counts = adacom.row_group_segment_counts(rows, dtype)
    col = ctx.column(dtype, counts)
    # sequential values 0..N-1 (zipf_distribution.cpp:21-26), uploaded in slabs
    slab = 1 << 24
    for a in range(0, rows, slab):
        b = min(rows, a + slab)
        col.upload_raw(np.arange(a, b, dtype=dtype), a)
    ctx.synchronize()
    size_before = col.data_size()
    col.compact()
    metas = col.meta()
    size_after = col.data_size()
    starts = np.array([m.row_start for m in metas], dtype=np.int64)
    seg_bytes = np.array([m.words * 8 for m in metas], dtype=np.int64)
    keys = zipf_keys(rows, 1.0, lookups, seed=42) - 1           # values are 0..N-1, Zipf draws 1..N
    seg = (np.searchsorted(starts, keys, side="right") - 1).astype(np.uint32)  # zone map: min <= k <= max
    cnt, first = col.point_filter_batch(seg, keys)
    assert np.all(cnt == 1) and np.array_equal(starts[seg] + first.astype(np.int64), keys), "config 1 answers wrong"
    ms = _dev_timed(ctx, lambda: col.point_filter_batch(seg, keys), steps, warmup)
    alg = int(seg_bytes[seg].sum())   # sum over the look-ups of the packed bytes of their segment
    out = {"workload": "benchmark_runner SuccinctZipfDistribution: UINTEGER 0..N-1, Zipf(q=1) point filters, one batched launch",
           "rows": rows, "segments": len(metas), "widths": sorted({int(m.width) for m in metas}),
           "data_size_before": int(size_before), "data_size_after": int(size_after), "lookups": lookups,
           "ms_per_step": round(ms, 4), "value": round(lookups / (ms * 1e-3), 1), "unit": "look-ups/s",
           "timed": "CUDA events around adacom_point_filter_batch incl. its H2D of the queries and D2H of the answers",
           "roofline": dict(_roof(alg, ms, peak), note="algorithmic bytes = packed bytes of the look-up's segment summed per "
                            "look-up (one scan per look-up, as the reference does); the shared scan reads a chunk once "
                            "per 64 look-ups of its segment, so this logical rate is not DRAM traffic")}
    o = _ref_oracle()
    if o is not None:
        k = min(cpu_lookups, lookups)
        sub = np.linspace(0, lookups - 1, k).astype(np.int64)   # evenly spread sample of the same look-ups
        rc = oracle.RefColumn(o, None, counts, only_segments=seg[sub],
                              value_of_rows=lambda a, b: np.arange(a, b, dtype=dtype))
        c_cnt, c_first, secs = oracle.bench_lookups(rc, seg[sub], keys[sub], threads, copy_per_call=True)
        assert np.all(c_cnt == 1) and np.array_equal(c_first, first[sub]), "CPU baseline disagrees with the GPU answers"
        out["cpu_baseline"] = {"value": round(k / secs, 1), "unit": "look-ups/s", "cores": threads, "kind": "reference",
                               "sample": f"{k} of the {lookups} look-ups (evenly spread), each = SuccinctScanPartial over all "
                                         f"vectors of its segment incl. the per-call deep copy (succinct.cpp:127) + equality "
                                         f"filter, look-ups dealt over {threads} threads",
                               "seconds": round(secs, 3)}
    col.close()
    return out


# ---------------------------------------------------------------------------------------------------------------------
# config 3
# ---------------------------------------------------------------------------------------------------------------------
def config3_skews(ctx, peak, threads, rows=1 << 28, keys_per_skew=100_000, epochs=3, skews=SKEWS, cpu_skews=(0.0, 1.0, 2.0)):
    import duckdb_adaptive_compression_b200 as adacom
    from duckdb_adaptive_compression_b200.zipf import zipf_keys
    import oracle
    dtype = np.dtype(np.uint32)
    counts = adacom.row_group_segment_counts(rows, dtype)
    n_seg = len(counts)
    col = ctx.column(dtype, counts)
    slab = 1 << 24
    for a in range(0, rows, slab):
        b = min(rows, a + slab)
        col.upload_raw(np.arange(a, b, dtype=dtype), a)
    ctx.synchronize()
    starts = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    e0, e1 = ctx.event(), ctx.event()

    def serve(seg, row, keys):
        """one batch of look-ups as fetch_row gathers (row id == key: the values are sequential); device ms"""
        ctx.record(e0)
        got = col.fetch_rows(seg, row)
        ctx.record(e1)
        ms = ctx.elapsed_ms(e0, e1)
        assert np.array_equal(got.astype(np.int64), keys), "config 3 fetch_row answers wrong"
        return ms

    def policy_epoch(seg):
        """ColumnSegmentCatalog::CompactSegments (column_segment_catalog.cpp:64-116): sort by reads, Compact() the coldest
        90 %, Uncompact() the rest -- each subset as ONE launch set (adacom_encode_list / adacom_uncompact_list)"""
        reads = np.bincount(seg, minlength=n_seg)
        order = np.argsort(reads, kind="stable")
        n_cold = int(n_seg * 0.90)
        cold, hot = np.sort(order[:n_cold]), np.sort(order[n_cold:])
        metas = col.meta()
        cold = np.array([g for g in cold if not (metas[g].state & 1)], dtype=np.uint32)       # Compact() returns early
        hot = np.array([g for g in hot if metas[g].state & 1], dtype=np.uint32)               # Uncompact() likewise
        moved = int(sum(int(metas[g].count) * 4 for g in cold) + sum(int(metas[g].count) * 4 for g in hot))
        ctx.record(e0)
        if len(cold):
            col.compact_list(cold)
        if len(hot):
            col.uncompact_list(hot)
        ctx.record(e1)
        return ctx.elapsed_ms(e0, e1), len(cold), len(hot), moved

    sweep = []
    cpu = {}
    o = _ref_oracle()
    for q in skews:
        keys = zipf_keys(rows // 2, q, keys_per_skew, seed=42 + int(round(10 * q))) - 1
        seg = (np.searchsorted(starts, keys, side="right") - 1).astype(np.uint32)
        row = (keys - starts[seg]).astype(np.uint32)
        arms = {}
        # --- non-succinct arm (run_diff_zipfian.sh:14): every segment flat
        col.uncompact()
        serve(seg, row, keys)
        ms = min(serve(seg, row, keys) for _ in range(3))
        arms["non_succinct"] = {"lookups_per_s": round(len(keys) / (ms * 1e-3)), "resident_bytes": int(col.data_size())}
        # --- adaptive arm (:12): starts flat, the policy re-compresses after every batch of look-ups
        series = []
        recompress_ms, recompress_bytes = 0.0, 0
        for ep in range(epochs):
            ms = serve(seg, row, keys)
            series.append([round(len(keys) / (ms * 1e-3)), int(col.data_size())])
            pms, n_cold, n_hot, moved = policy_epoch(seg)
            recompress_ms += pms
            recompress_bytes += moved
        ms = serve(seg, row, keys)
        series.append([round(len(keys) / (ms * 1e-3)), int(col.data_size())])
        arms["adaptive"] = {"series_lookups_per_s_resident_bytes": series,
                            "recompress_GBs": round(recompress_bytes / (recompress_ms * 1e-3) / 1e9, 1) if recompress_ms else None,
                            "recompress_ms_total": round(recompress_ms, 3)}
        # filter-scan service of the same look-ups on the hot/cold mix (config-1 style), a prefix
        k = min(2000, len(keys))
        cnt, first = col.point_filter_batch(seg[:k], keys[:k])
        assert np.all(cnt == 1) and np.array_equal(starts[seg[:k]] + first.astype(np.int64), keys[:k])
        # --- non-adaptive succinct arm (:13): every segment compacted
        col.compact()
        serve(seg, row, keys)
        ms = min(serve(seg, row, keys) for _ in range(3))
        arms["non_adaptive"] = {"lookups_per_s": round(len(keys) / (ms * 1e-3)), "resident_bytes": int(col.data_size())}
        sweep.append({"q": q, "segments_hit": int(len(np.unique(seg))), **arms})
        if o is not None and q in cpu_skews:
            touched = np.unique(seg)
            rc = oracle.RefColumn(o, None, counts, only_segments=touched,
                                  value_of_rows=lambda a, b: np.arange(a, b, dtype=dtype))
            got, secs = oracle.bench_fetch(rc, seg, row, dtype, threads)
            assert np.array_equal(got.astype(np.int64), keys)
            hot = touched[: max(1, len(touched) // 10)].astype(np.uint32)
            un_s, re_s = oracle.bench_epoch(rc, hot)
            hot_bytes = int(sum(int(counts[g]) * 4 for g in hot))
            cpu[str(q)] = {"fetch_per_s": round(len(keys) / secs), "uncompact_GBs": round(hot_bytes / un_s / 1e9, 3),
                           "recompact_GBs": round(hot_bytes / re_s / 1e9, 3), "hot_segments": int(len(hot))}
    ctx.event_destroy(e0)
    ctx.event_destroy(e1)
    col.close()
    mid = sweep[len(sweep) // 2]
    out = {"workload": "run_diff_zipfian.sh shapes: 21 skews x {adaptive, non-adaptive succinct, non-succinct}; fetch_row "
                       "gathers + 90/10 policy epochs (cold segments re-compressed, hot ones uncompacted)",
           "rows": rows, "keys_per_skew": keys_per_skew, "epochs": epochs, "segments": int(n_seg),
           "value": mid["non_adaptive"]["lookups_per_s"], "unit": "fetch_row/s (q = 1.0, non-adaptive succinct, device-timed "
                                                                   "incl. H2D of the ids and D2H of the values)",
           "sweep": sweep,
           "timed": "CUDA events around adacom_fetch_rows (H2D ids, kernel, D2H values) and around the policy epoch's "
                    "adacom_encode_list + adacom_uncompact_list"}
    if cpu:
        out["cpu_baseline"] = {"value": cpu.get("1.0", next(iter(cpu.values())))["fetch_per_s"], "unit": "fetch_row/s",
                               "cores": threads, "kind": "reference",
                               "sample": f"skews {sorted(cpu)}: the same {keys_per_skew} keys as sdsl int_vector random access "
                                         f"(+ min) over {threads} threads; epoch = UncompressSuccinct + "
                                         f"BitCompressFromUncompressed of a tenth of the touched segments on ONE thread "
                                         f"(the reference's compactor thread)",
                               "per_skew": cpu}
    return out


# ---------------------------------------------------------------------------------------------------------------------
# configs 4 and 5
# ---------------------------------------------------------------------------------------------------------------------
def _tiled_lineitem(row_groups: int, distinct: int, names, sf: float, seed_base: int = 0):
    """lineitem columns of `row_groups` row groups: `distinct` row groups minted by the generator, tiled (the scan does
    not care that values repeat; minting 4883 row groups on the host would take minutes)"""
    from duckdb_adaptive_compression_b200 import tpch
    distinct = min(distinct, row_groups)
    base = tpch.lineitem(distinct, sf, row_groups=[seed_base + g for g in range(distinct)], names=names)
    reps, rem = divmod(row_groups, distinct)
    cols = {}
    for k in names:
        parts = [base[k]] * reps
        if rem:
            parts.append(base[k][: rem * 122880])
        cols[k] = np.concatenate(parts)
    return cols, base


def config4_tpch(ctx, peak, threads, sf=10.0, steps=10, warmup=3, distinct=61, cpu_row_groups=24):
    import duckdb_adaptive_compression_b200 as adacom
    from duckdb_adaptive_compression_b200 import tpch
    import oracle
    import torch
    names = sorted(set(tpch.LineitemTable.Q6_COLUMNS + tpch.LineitemTable.Q1_COLUMNS))
    n_rg = max(int(6_000_000 * sf) // adacom.ROW_GROUP_SIZE, 1)
    cols, base = _tiled_lineitem(n_rg, distinct, names, sf)
    table = tpch.LineitemTable(ctx, cols)
    ctx.synchronize()
    n_rows = table.n_rows
    reps, rem = divmod(n_rg, min(distinct, n_rg))
    # expected answers: numpy on the distinct part (pinned to the reference by tests/test_tpch_reference.py), scaled by tiles
    def tile_expect(fn):
        full = fn(base)
        part = fn({k: v[: rem * 122880] for k, v in base.items()}) if rem else None
        return full, part
    (q6_sum, q6_cnt), q6_part = tile_expect(tpch.q6_expected)
    want_sum = q6_sum * reps + (q6_part[0] if q6_part else 0)
    want_cnt = q6_cnt * reps + (q6_part[1] if q6_part else 0)
    got = table.q6()
    assert (got.sum, got.count) == (want_sum, want_cnt), "config 4 Q6 answer wrong"
    q1_full, q1_part = tile_expect(tpch.q1_expected)
    q1_got = table.q1()
    for code, g in q1_full.items():
        for key, v in g.items():
            want = v * reps + (q1_part[code][key] if q1_part and code in q1_part else 0)
            assert q1_got[code][key] == want, ("config 4 Q1 answer wrong", code, key)
    dev = torch.device("cuda", ctx.device)
    agg_dev = torch.zeros(4 * 5 * 4, dtype=torch.int64, device=dev)
    q6_ms = _dev_timed(ctx, lambda: table.q6(device_out=agg_dev.data_ptr()), steps, warmup)
    q1_ms = _dev_timed(ctx, lambda: table.q1(device_out=agg_dev.data_ptr()), steps, warmup)
    q6_bytes = table.packed_bytes(tpch.LineitemTable.Q6_COLUMNS)
    q1_bytes = table.packed_bytes(tpch.LineitemTable.Q1_COLUMNS)
    widths = {k: sorted({int(m.width) for m in table.cols[k].meta()}) for k in names}
    out = {"workload": f"TPC-H SF{sf:g} lineitem ({n_rows} rows = {n_rg} row groups: {min(distinct, n_rg)} minted by the "
                       f"deterministic generator, tiled), integer columns succinct-compressed, Q6 + Q1 fused scans",
           "rows": int(n_rows), "widths": widths,
           "q6": {"ms": round(q6_ms, 4), "rows_per_s": round(n_rows / (q6_ms * 1e-3)), "roofline": _roof(q6_bytes, q6_ms, peak)},
           "q1": {"ms": round(q1_ms, 4), "rows_per_s": round(n_rows / (q1_ms * 1e-3)), "roofline": _roof(q1_bytes, q1_ms, peak)},
           "value": round(n_rows / (q6_ms * 1e-3)), "unit": "rows/s (Q6)",
           "timed": "CUDA events around adacom_scan_columns[_grouped] with device-resident aggregates"}
    table.close()
    o = _ref_oracle()
    if o is not None:
        k_rg = min(cpu_row_groups, min(distinct, n_rg))
        n_cpu = k_rg * adacom.ROW_GROUP_SIZE
        sub = {k: np.ascontiguousarray(base[k][:n_cpu]) for k in names}
        rc = {k: oracle.RefColumn(o, sub[k], adacom.row_group_segment_counts(n_cpu, sub[k].dtype)) for k in names}
        q6_cols = [(rc["l_shipdate"], (tpch.D_1994_01_01, tpch.D_1995_01_01 - 1)), (rc["l_discount"], (5, 7)),
                   (rc["l_quantity"], (-(1 << 31), 23)), (rc["l_extendedprice"], None)]
        res, q6_s = oracle.bench_query(o, q6_cols, [[(3, 0, 1), (1, 0, 1)]], n_cpu, threads)
        assert res[0] == tpch.q6_expected(sub), "CPU Q6 baseline disagrees with numpy"
        q1_cols = [(rc["l_shipdate"], (-(1 << 31), tpch.D_1998_09_02)), (rc["l_rfls"], None), (rc["l_quantity"], None),
                   (rc["l_extendedprice"], None), (rc["l_discount"], None), (rc["l_tax"], None)]
        res1, q1_s = oracle.bench_query(o, q1_cols, tpch.LineitemTable.Q1_SUM_EXPRS, n_cpu, threads, group_by=(1, 4))
        exp1 = tpch.q1_expected(sub)
        for code, g in exp1.items():
            assert res1[code * 5][0] == g["sum_qty"] and res1[code * 5][1] == g["count"], "CPU Q1 baseline disagrees"
        out["cpu_baseline"] = {"value": round(n_cpu / q6_s), "unit": "rows/s (Q6)", "cores": threads, "kind": "reference",
                               "q1_rows_per_s": round(n_cpu / q1_s),
                               "sample": f"the first {k_rg} row groups ({n_cpu} rows): per 2048-row vector SuccinctScanPartial of "
                                         f"every filter column (incl. the per-call deep copy, succinct.cpp:127) + "
                                         f"FilterSelection + HUGEINT sums, one row group per task over {threads} threads",
                               "seconds": {"q6": round(q6_s, 3), "q1": round(q1_s, 3)}}
    return out


def config5_tpch_sharded(ctx, peak, rank, world, sf=100.0, steps=10, warmup=3, distinct=61):
    """row groups dealt round-robin over the ranks (sharding.shard_row_groups); every rank runs the fused Q6 over its own
    row groups; the 24-byte partials are all-gathered over NCCL and combined exactly."""
    import torch
    import torch.distributed as dist
    import duckdb_adaptive_compression_b200 as adacom
    from duckdb_adaptive_compression_b200 import sharding, tpch
    names = list(tpch.LineitemTable.Q6_COLUMNS)
    n_rg = max(int(6_000_000 * sf) // adacom.ROW_GROUP_SIZE, 1)
    mine = sharding.shard_row_groups(n_rg, world, rank)
    cols, base = _tiled_lineitem(len(mine), distinct, names, sf, seed_base=rank * distinct)
    table = tpch.LineitemTable(ctx, cols)
    del cols
    ctx.synchronize()
    dev = torch.device("cuda", ctx.device)
    agg_dev = torch.zeros(4, dtype=torch.int64, device=dev)
    e0, e1 = ctx.event(), ctx.event()
    stream = torch.cuda.ExternalStream(ctx.stream_ptr, device=dev) if hasattr(ctx, "stream_ptr") else None

    def one():
        table.q6(device_out=agg_dev.data_ptr())
        if world > 1:
            # the only exchange: {sum_lo, sum_hi, count} per rank, on the codec's own stream (no host sync in between)
            if stream is not None:
                with torch.cuda.stream(stream):
                    return sharding.all_gather_aggregates(agg_dev)
            ctx.synchronize()
            return sharding.all_gather_aggregates(agg_dev)
        return agg_dev

    for _ in range(warmup):
        one()
    ctx.synchronize()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ctx.record(e0)
    gathered = None
    for _ in range(steps):
        gathered = one()
    ctx.record(e1)
    ms = ctx.elapsed_ms(e0, e1) / steps
    torch.cuda.synchronize()
    total_sum, total_cnt = sharding.combine_aggregates(gathered, 1, True)[0]
    # expected: every rank's numpy answer, summed through a (CPU-side) all-reduce of Python ints via all_gather_object
    reps, rem = divmod(len(mine), min(distinct, max(len(mine), 1))) if len(mine) else (0, 0)
    s, c = tpch.q6_expected(base) if len(mine) else (0, 0)
    ps, pc = tpch.q6_expected({k: v[: rem * 122880] for k, v in base.items()}) if rem else (0, 0)
    local = (s * reps + ps, c * reps + pc, int(table.n_rows), int(table.packed_bytes(names)))
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        everyone = [None] * world
        dist.all_gather_object(everyone, local)
    else:
        everyone = [local]
    want_sum, want_cnt = sum(x[0] for x in everyone), sum(x[1] for x in everyone)
    rows_all, bytes_all = sum(x[2] for x in everyone), sum(x[3] for x in everyone)
    assert (total_sum, total_cnt) == (want_sum, want_cnt), "config 5 Q6 answer wrong"
    table.close()
    roof = _roof(bytes_all, ms, peak * world)
    roof["peak_note"] = f"{world} x measured per-GPU peak"
    return {"workload": f"TPC-H SF{sf:g} lineitem ({rows_all} rows, {n_rg} row groups dealt round-robin over {world} GPU(s)), "
                        f"fused Q6 per GPU + NCCL all-gather of the 24-byte partials inside the timed region",
            "rows": int(rows_all), "n_gpus": world, "q6": {"ms": round(ms, 4), "rows_per_s": round(rows_all / (ms * 1e-3)),
                                                           "roofline": roof, "revenue_scale4": str(total_sum)},
            "value": round(rows_all / (ms * 1e-3)), "unit": "rows/s (Q6, all GPUs)",
            "timed": "CUDA events on the codec's stream around scan_columns + all_gather, max over ranks"}


# ---------------------------------------------------------------------------------------------------------------------
# the neighbour codecs (DuckDB bitpacking / RLE segments) on the same GPU
# ---------------------------------------------------------------------------------------------------------------------
def neighbour_codecs(ctx, peak, threads, log2_values=26, steps=5, warmup=3):
    """bitpacking (FOR groups, 17-bit payloads) and RLE (runs of 1..8) over 2^log2_values INT32 values: batched scan to
    flat vectors and fused SUM, device-timed; CPU baseline = the reference's own BitpackingScanPartial / RLEScanPartial
    loops (oracle/_ref/libadacom_codecs_ref.so) over `threads` host threads, one segment per task."""
    import concurrent.futures as cf
    import torch
    import duckdb_adaptive_compression_b200 as adacom
    from oracle import codecs
    from tests.cases import splitmix64
    if not codecs.available():
        return {"error": "oracle/_ref/libadacom_codecs_ref.so not built (needs /root/reference)"}
    n = 1 << log2_values
    dev = torch.device("cuda", ctx.device)
    flat = torch.empty(n * 4 + 256, dtype=torch.uint8, device=dev)
    out = {"workload": f"DuckDB bitpacking (FOR, 17-bit) and RLE (runs of 1..8) segments, 2^{log2_values} INT32 values, "
                       f"scan -> flat vectors and fused SUM", "values": n}
    r = splitmix64(np.arange(n, dtype=np.uint64) + np.uint64(0xB17)).astype(np.uint64)
    inputs = {"bitpacking": ((r & np.uint64((1 << 17) - 1)).astype(np.int64) + 1000).astype(np.int32),
              "rle": np.repeat(((r[: n // 4] >> np.uint64(20)) & np.uint64(0xFFFFF)).astype(np.int32),
                               (r[: n // 4] & np.uint64(7)).astype(np.int64) + 1)[:n]}
    for name, vals in inputs.items():
        codec = adacom.CODEC_BITPACKING if name == "bitpacking" else adacom.CODEC_RLE
        vals = np.ascontiguousarray(vals)
        ref = codecs.RefCodecColumn(vals, codec=codecs.BITPACKING if name == "bitpacking" else codecs.RLE)
        col = adacom.CodecColumn(ctx, codec, np.int32, ref.images, ref.counts)
        m = len(vals)
        col.decode(device_out=flat.data_ptr())
        ctx.synchronize()
        got = flat[: m * 4].view(torch.int32).cpu().numpy()
        assert np.array_equal(got, vals), f"{name}: GPU decode differs from the input"
        agg = col.scan_sum()
        assert agg.sum == int(vals.astype(np.int64).sum()) and agg.count == m, f"{name}: fused SUM wrong"
        dec_ms = _dev_timed(ctx, lambda: col.decode(device_out=flat.data_ptr()), steps, warmup)
        sum_ms = _dev_timed(ctx, lambda: col.scan_sum(), steps, warmup)
        img = col.image_bytes
        entry = {"segments": len(ref.counts), "image_bytes": int(img),
                 "decode": {"ms": round(dec_ms, 4), "roofline": _roof(img + m * 4, dec_ms, peak)},
                 "scan_sum": {"ms": round(sum_ms, 4), "roofline": _roof(img, sum_ms, peak),
                              "timed": "incl. the D2H of the aggregate (synchronous call)"}}
        col.close()
        # CPU: the reference's scan loop per segment in 2048-value vectors, segments dealt over the host threads
        k_seg = min(len(ref.counts), max(threads * 2, 8))
        def scan_seg(s):
            c = ref.counts[s]
            for off in range(0, c, 2048 * 16):
                ref.scan(s, off, min(2048 * 16, c - off))
            return c
        t0 = time.perf_counter()
        with cf.ThreadPoolExecutor(threads) as ex:
            rows = sum(ex.map(scan_seg, range(k_seg)))
        secs = time.perf_counter() - t0
        entry["cpu_baseline"] = {"value": round(rows * 4 / secs / 1e9, 3), "unit": "GB/s decoded", "cores": threads,
                                 "kind": "reference", "sample": f"the first {k_seg} segments ({rows} values) through the "
                                 f"reference's own scan loop (fresh scan state + Skip per 32768-value slice), one segment per "
                                 f"task over {threads} threads"}
        entry["decode"]["logical_GBs"] = round(m * 4 / (dec_ms * 1e-3) / 1e9, 1)
        out[name] = entry
    out["value"] = out["bitpacking"]["decode"]["logical_GBs"]
    out["unit"] = "GB/s decoded (bitpacking scan -> flat)"
    return out


# ---------------------------------------------------------------------------------------------------------------------
# the drop-in vtable path: SuccinctGpuFun's slots driven like ColumnData drives the reference's
# ---------------------------------------------------------------------------------------------------------------------
def vtable_dropin(threads, log2_values=27, bits=17, cpu_log2_values=22):
    """host/build/vtable_bench (C++: ColumnData::AppendData -> batched compaction, ColumnData::ScanVector pulls of 2048
    values through scan_vector / scan_partial, FetchRow / FetchRows) on config 2's column shape (INT64, 2^27 values,
    one width), HOST buffers on both sides; beside it the reference's SuccinctScanPartial loop incl. its per-call deep
    copy (succinct.cpp:127 -- what the reference really executes per pull) from oracle/_ref on the host cores."""
    import json as _json
    import subprocess as _sp
    exe = os.path.join(ROOT, "duckdb_adaptive_compression_b200", "host", "build", "vtable_bench")
    if not os.path.exists(exe):
        raise RuntimeError("host/build/vtable_bench is not built (run __graft_entry__.build())")
    scan_threads = max(1, min(8, threads // 2))
    runs = {}
    for t in sorted({1, scan_threads}):
        r = _sp.run([exe, str(log2_values), str(bits), str(t)], capture_output=True, text=True, timeout=600)
        if r.returncode != 0:
            raise RuntimeError(f"vtable_bench failed (rc {r.returncode}): {r.stdout[-300:]} {r.stderr[-300:]}")
        runs[t] = _json.loads(r.stdout.strip().splitlines()[-1])
    best = max(runs.values(), key=lambda r: r["scan_GBs"])
    scan_threads = best["scan_threads"]
    out = {"workload": f"SuccinctGpuFun vtable driven like ColumnData (append / ScanVector / FetchRow): INT64, 2^{log2_values} "
                       f"values, {bits}-bit payloads, host buffers on both sides",
           "value": best["scan_GBs"], "unit": "GB/s decoded into host vectors (scan_vector pulls)",
           "scan_threads": scan_threads, "scan_1thread_GBs": runs[1]["scan_GBs"],
           "scan_verified_GBs": best["scan_verify_GBs"], "zero_copy_vectors": best["zero_copy_vectors"],
           "vectors": best["vectors"], "segments": best["segments"], "device_columns": best["device_columns"],
           "append_compact_GBs": best["load_GBs"], "fetch_row_per_s": best["fetch_row_per_s"],
           "fetch_rows_batched_per_s": best["fetch_rows_batched_per_s"],
           # point reads (the scan of the key's 2048-row vector, Zipf-like keys, one host thread) under the three arms of
           # run_diff_zipfian.sh:12-14: every segment packed in HBM / every segment a flat host block / after one policy
           # epoch (coldest 90 % packed in HBM by one launch set, hottest 10 % flat in host memory)
           "point_reads_per_s": best.get("point_reads_per_s"), "adaptive_epoch_s": best.get("adaptive_epoch_s"),
           "adaptive_resident_bytes": best.get("adaptive_resident_bytes"), "resident_bytes_packed": best.get("data_size"),
           "answers_checked": best["scan_ok"] and best["fetch_ok"] and best.get("point_ok", False),
           "timed": "host clock around the pull loops (device work, D2H of the pinned windows and the pulls overlap); load = "
                    "append of host vectors + H2D + analyze + encode, batched per 512 segments"}
    o = _ref_oracle()
    if o is not None:
        from tests.cases import splitmix64
        n = 1 << cpu_log2_values
        vals = (splitmix64(np.arange(n, dtype=np.uint64)) & np.uint64((1 << bits) - 1)).view(np.int64) + 1000003
        one = o.bench_column(vals, 32767, 1, copy_per_call=True)
        allc = o.bench_column(vals, 32767, threads, copy_per_call=True)
        per_thread = n * 8 / one["scan_s"] / 1e9
        out["cpu_baseline"] = {"value": round(n * 8 / allc["scan_s"] / 1e9, 3), "unit": "GB/s decoded", "cores": threads,
                               "kind": "reference", "per_thread_GBs": round(per_thread, 3),
                               "append_compact_GBs": round(n * 8 / (allc["analyze_s"] + allc["encode_s"]) / 1e9, 3),
                               "sample": f"2^{cpu_log2_values} values of the same distribution: SuccinctScanPartial per 2048-value "
                                         f"vector incl. the per-call deep copy of the packed vector (succinct.cpp:127), one "
                                         f"thread and {threads} threads"}
        out["speedup_per_host_thread"] = round(runs[1]["scan_GBs"] / per_thread, 1)
    return out
