#!/usr/bin/env python
"""Zipf look-up workloads on the succinct codec (BASELINE.json configs 1 and 3) -- secondary benchmark.

Config 1 (benchmark/micro/succinct/zipf_distribution.cpp:13-61): UINTEGER column 0..N-1 in row order, DuckDB
segmentation, every segment compacted; 10 000 look-ups `value == k`, k ~ Zipf(N, q = 1), each = zone map -> exactly
one segment -> decode all its vectors + equality filter -> one row.  All look-ups run as ONE batched launch
(adacom_point_filter_batch).  Config 3 (zipf_distribution_diff_skews.cpp:13-111, run_diff_zipfian.sh): keys
~ Zipf(N/2, q) for a sweep of q; the look-ups are served as batched fetch_row gathers (row id == key because the
values are sequential) and as filter scans, and one policy epoch re-compresses: read counters per segment ->
coldest 90 % stay/become compacted, the hottest 10 % are uncompacted (column_segment_catalog.cpp:64-116).
Every answer is checked (count == 1, row == key, fetched value == key)."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=100_000_000)
    ap.add_argument("--lookups", type=int, default=10_000)
    ap.add_argument("--skews", default="0.0,0.5,1.0,1.5,2.0")
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()

    import duckdb_adaptive_compression_b200 as adacom
    from duckdb_adaptive_compression_b200.zipf import zipf_keys

    ctx = adacom.Context(0)
    n = args.rows
    dtype = np.dtype(np.uint32)
    counts = adacom.row_group_segment_counts(n, dtype)
    col = ctx.column_from_values(np.arange(n, dtype=dtype), counts)
    raw_size = col.data_size()
    col.compact()
    metas = col.meta()
    packed_size = col.data_size()
    starts = np.array([m.row_start for m in metas], dtype=np.int64)
    seg_bytes = np.array([m.words * 8 for m in metas], dtype=np.int64)
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0

    def timed(fn):
        fn()
        best = 1e9
        for _ in range(args.reps):
            ctx.synchronize()
            t0 = time.perf_counter()
            fn()
            ctx.synchronize()
            best = min(best, time.perf_counter() - t0)
        return best

    # ---- config 1 ----
    keys = zipf_keys(n, 1.0, args.lookups, seed=42) - 1          # values are 0..N-1, Zipf draws 1..N
    seg = (np.searchsorted(starts, keys, side="right") - 1).astype(np.uint32)  # zone map: min <= k <= max
    cnt, first = col.point_filter_batch(seg, keys)
    assert np.all(cnt == 1) and np.array_equal(starts[seg] + first.astype(np.int64), keys), "look-up answers wrong"
    t = timed(lambda: col.point_filter_batch(seg, keys))
    scanned = int(seg_bytes[seg].sum())
    out = {"config1": {"rows": n, "segments": len(metas), "widths": sorted({m.width for m in metas}),
                       "data_size_before": raw_size, "data_size_after": packed_size,
                       "lookups": args.lookups, "ms_e2e_host_clock": round(t * 1e3, 3),
                       "lookups_per_s": round(args.lookups / t), "logical_scanned_packed_bytes": scanned,
                       "logical_GBs": round(scanned / t / 1e9, 1),
                       "note": "one batched launch incl. H2D of the queries and D2H of the answers; logical bytes = "
                               "the packed bytes of the look-up's segment, summed per look-up (what one scan per look-up "
                               "reads, as the reference does); the shared scan reads a 16 KiB chunk once per 64 look-ups "
                               "of its segment, so the rate is not an HBM fraction"}}

    # ---- config 3: skew sweep, fetch_row + one policy epoch ----
    sweep = []
    for q in [float(x) for x in args.skews.split(",")]:
        keys = zipf_keys(n // 2, q, 100_000, seed=42 + int(round(10 * q))) - 1
        seg = (np.searchsorted(starts, keys, side="right") - 1).astype(np.uint32)
        row = (keys - starts[seg]).astype(np.uint32)
        got = col.fetch_rows(seg, row)
        assert np.array_equal(got.astype(np.int64), keys), "fetch_row answers wrong"
        t_fetch = timed(lambda: col.fetch_rows(seg, row))
        # policy epoch (column_segment_catalog.cpp:64-116): sort by reads, coldest 90 % compacted, rest uncompacted
        reads = np.bincount(seg, minlength=len(metas))
        order = np.argsort(reads, kind="stable")
        n_cold = int(len(order) * 0.90)
        hot = np.sort(order[n_cold:])
        t0 = time.perf_counter()
        runs = np.split(hot, np.where(np.diff(hot) != 1)[0] + 1) if len(hot) else []
        launches0 = ctx.launch_count
        col.uncompact_list(hot)  # the whole hot subset in one launch set (adacom_uncompact_list)
        ctx.synchronize()
        epoch_launches = ctx.launch_count - launches0
        t_epoch = time.perf_counter() - t0
        size_after = col.data_size()
        got = col.fetch_rows(seg, row)
        assert np.array_equal(got.astype(np.int64), keys), "fetch_row after the policy epoch wrong"
        cnt, first = col.point_filter_batch(seg[:2000], keys[:2000])
        assert np.all(cnt == 1) and np.array_equal(starts[seg[:2000]] + first.astype(np.int64), keys[:2000])
        t0 = time.perf_counter()
        col.compact_list(hot, analyze=False)  # next epoch start: the hot subset becomes cold again (adacom_encode_list)
        ctx.synchronize()
        t_recompact = time.perf_counter() - t0
        sweep.append({"q": q, "distinct_segments_hit": int((reads > 0).sum()),
                      "Mfetch_per_s_e2e": round(len(keys) / t_fetch / 1e6, 2),
                      "hot_segments_uncompacted": int(len(hot)), "hot_runs": len(runs), "epoch_launches": int(epoch_launches),
                      "epoch_ms": round(t_epoch * 1e3, 3), "data_size_hot_cold": size_after,
                      "recompact_ms": round(t_recompact * 1e3, 3)})
    out["config3"] = {"rows": n, "keys_per_skew": 100_000, "sweep": sweep}
    print(json.dumps(out), flush=True)
    col.close()
    ctx.close()


if __name__ == "__main__":
    main()
