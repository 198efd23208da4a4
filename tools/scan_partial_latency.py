"""Per-call cost of the slowest drop-in level: adacom_scan_partial of one 2048-value vector into a host buffer
(kernel launch + D2H through the pinned staging pair + synchronisation), and adacom_fetch_rows of one row."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import duckdb_adaptive_compression_b200 as adacom  # noqa: E402

ctx = adacom.Context(0)
n = 1 << 22
counts = adacom.row_group_segment_counts(n, np.int64)
col = ctx.column(np.int64, counts)
col.fill_splitmix(1, 0, (1 << 17) - 1)
col.compact()
want = col.decode()
res = np.zeros(2048, dtype=np.int64)
rng = np.random.default_rng(0)
calls = 3000
segs = rng.integers(0, len(counts) - 1, size=calls)
for warm in (True, False):
    t0 = time.perf_counter()
    for s in segs[:200] if warm else segs:
        col.scan_partial(int(s), 2048 * 3, 2048, result=res)
    dt = time.perf_counter() - t0
starts = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
assert np.array_equal(res, want[int(starts[segs[-1]]) + 6144:int(starts[segs[-1]]) + 8192])
t0 = time.perf_counter()
for s in segs:
    col.fetch_rows([int(s)], [17])
df = time.perf_counter() - t0
print(json.dumps({"scan_partial_us_per_call": round(dt / calls * 1e6, 2), "scan_partial_GBs": round(calls * 16384 / dt / 1e9, 3),
                  "fetch_row_us_per_call": round(df / calls * 1e6, 2),
                  "note": "python ctypes loop: includes ~2 us of interpreter overhead per call"}))
