#!/usr/bin/env python
"""TPC-H lineitem on the succinct codec (BASELINE.json configs 4 and 5) -- secondary benchmark, not bench.py's line.

  python tools/tpch_bench.py --sf 10                         # config 4: Q6 + Q1 on one B200
  torchrun --nproc-per-node 8 tools/tpch_bench.py --sf 100   # config 5: row groups round-robin over the ranks,
                                                             # Q6 aggregate all-gathered over NCCL

Every rank mints its own row groups with the deterministic generator (tpch.lineitem_row_group), compresses the
columns segment by segment and runs the fused multi-column scan.  Reported per query: rows/s, achieved HBM GB/s =
packed bytes of the referenced columns / CUDA-event time, and the fraction of MEASURED_PEAKS.json hbm_gbs.
`--check N` verifies the answers of the first N local row groups against numpy on the uncompressed columns.
"""
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
    ap.add_argument("--sf", type=float, default=1.0)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--check", type=int, default=4, help="local row groups to verify against numpy")
    ap.add_argument("--no-q1", action="store_true")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist
    import duckdb_adaptive_compression_b200 as adacom
    from duckdb_adaptive_compression_b200 import sharding, tpch

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    ctx = adacom.Context(local_rank)

    total_rows = int(6_000_000 * args.sf)
    n_rg = max(total_rows // adacom.ROW_GROUP_SIZE, 1)
    mine = sharding.shard_row_groups(n_rg, world, rank)
    names = sorted(set(tpch.LineitemTable.Q6_COLUMNS + tpch.LineitemTable.Q1_COLUMNS))
    t0 = time.time()
    parts = {k: [] for k in names}
    check_cols = {k: [] for k in names}
    for i, g in enumerate(mine):
        rg = tpch.lineitem_row_group(int(g), args.sf)
        for k in names:
            parts[k].append(rg[k])
            if i < args.check:
                check_cols[k].append(rg[k])
    cols = {k: np.concatenate(v) for k, v in parts.items()}
    del parts
    gen_s = time.time() - t0
    t0 = time.time()
    table = tpch.LineitemTable(ctx, cols)
    ctx.synchronize()
    load_s = time.time() - t0
    n_rows = table.n_rows
    raw_bytes = {k: cols[k].nbytes for k in names}
    widths = {k: sorted({m.width for m in table.cols[k].meta()}) for k in names}
    del cols

    # ---- correctness on the first --check local row groups (same plans over a prefix of the rows) ----
    if args.check and len(mine):
        sub = {k: np.concatenate(v) for k, v in check_cols.items()}
        n_sub = len(sub["l_quantity"])
        c = table.cols
        terms = [(c["l_shipdate"], (tpch.D_1994_01_01, tpch.D_1995_01_01 - 1)), (c["l_discount"], (5, 7)),
                 (c["l_quantity"], (-(1 << 31), 23)), (c["l_extendedprice"], None)]
        got = adacom.scan_columns(ctx, terms, [[(3, 0, 1), (1, 0, 1)]], 0, n_sub)[0]
        want = tpch.q6_expected(sub)
        assert (got.sum, got.count) == want, f"Q6 mismatch on rank {rank}: {got} vs {want}"

    agg_dev = torch.zeros(4 * 8, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()
    e0, e1 = ctx.event(), ctx.event()

    def timed(fn):
        for _ in range(args.warmup):
            fn()
        ctx.synchronize()
        if world > 1:
            dist.barrier()
        ctx.record(e0)
        for _ in range(args.steps):
            fn()
        ctx.record(e1)
        ms = ctx.elapsed_ms(e0, e1) / args.steps
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    q6_ms = timed(lambda: table.q6(device_out=agg_dev.data_ptr()))
    ctx.synchronize()
    t0 = time.perf_counter()
    if world > 1:
        gathered = sharding.all_gather_aggregates(agg_dev[:4])
        torch.cuda.synchronize()
    else:
        gathered = agg_dev[:4]
    nccl_ms = (time.perf_counter() - t0) * 1e3
    q6_sum, q6_count = sharding.combine_aggregates(gathered, 1, True)[0]
    q6_bytes = table.packed_bytes(tpch.LineitemTable.Q6_COLUMNS)

    q1_ms, q1_bytes, q1, q1_passes_ms = None, None, None, None
    if not args.no_q1:
        q1_dev = torch.zeros(4 * 5 * 4, dtype=torch.int64, device=dev)
        q1_ms = timed(lambda: table.q1(device_out=q1_dev.data_ptr()))
        q1_passes_ms = timed(lambda: table.q1_per_group_passes())
        q1_bytes = table.packed_bytes(tpch.LineitemTable.Q1_COLUMNS)
        q1 = table.q1()

    def allsum(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t)
        return float(t.item())

    rows_all = int(allsum(n_rows))
    q6_bytes_all = allsum(q6_bytes)
    q1_bytes_all = allsum(q1_bytes) if q1_bytes is not None else None
    if rank == 0:
        try:
            peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
        except Exception:
            peak = 6650.0
        line = {"workload": f"TPC-H SF{args.sf:g} lineitem, succinct-compressed integer columns", "n_gpus": world,
                "rows": rows_all, "row_groups": int(n_rg), "widths": widths,
                "raw_bytes_per_row": sum(raw_bytes[k] for k in names) / max(n_rows, 1),
                "q6": {"ms": round(q6_ms, 4), "rows_per_s": round(rows_all / (q6_ms * 1e-3), 1),
                       "packed_bytes": int(q6_bytes_all), "GBs": round(q6_bytes_all / (q6_ms * 1e-3) / 1e9, 1),
                       "frac_of_hbm_peak_per_gpu": round(q6_bytes_all / (q6_ms * 1e-3) / 1e9 / peak / world, 4),
                       "revenue_scale4": str(q6_sum), "matching_rows": q6_count,
                       "allgather_ms_host_clock": round(nccl_ms, 3)},
                "generate_s": round(gen_s, 2), "compress_upload_s": round(load_s, 2),
                "checked_row_groups_per_rank": min(args.check, len(mine))}
        if q1_ms is not None:
            line["q1"] = {"ms": round(q1_ms, 4), "rows_per_s": round(rows_all / (q1_ms * 1e-3), 1),
                          "packed_bytes": int(q1_bytes_all), "passes": 1,
                          "GBs": round(q1_bytes_all / (q1_ms * 1e-3) / 1e9, 1),
                          "frac_of_hbm_peak_per_gpu": round(q1_bytes_all / (q1_ms * 1e-3) / 1e9 / peak / world, 4),
                          "ms_with_one_filtered_pass_per_group": round(q1_passes_ms, 4),
                          "groups": {str(k): {kk: str(vv) for kk, vv in v.items()} for k, v in (q1 or {}).items()}
                          if world == 1 else "per-rank partials (rank 0 shown omitted)"}
        print(json.dumps(line), flush=True)
    table.close()
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
