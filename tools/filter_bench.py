"""Fused decode + filter (+ selection bitmap) rates on one INT64 column of 2^27 values: adacom_scan_filter.
Algorithmic bytes = packed payload (+ 1 bit per row for the bitmap)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import duckdb_adaptive_compression_b200 as adacom  # noqa: E402


def main():
    import torch
    logn = int(sys.argv[1]) if len(sys.argv) > 1 else 27
    widths = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [4, 8, 17, 32, 40, 63]
    n = 1 << logn
    ctx = adacom.Context(0)
    counts = adacom.row_group_segment_counts(n, np.int64)
    bitmap = torch.zeros(n // 64 + 8, dtype=torch.int64, device="cuda:0")
    e0, e1 = ctx.event(), ctx.event()

    def timeit(fn, reps=5):
        best = 1e9
        for _ in range(reps):
            ctx.flush_l2()
            ctx.record(e0)
            fn()
            ctx.record(e1)
            best = min(best, ctx.elapsed_ms(e0, e1))
        return best

    for b in widths:
        col = ctx.column(np.int64, counts)
        col.fill_splitmix(0xADAC0000 + b, 0, (1 << b) - 1)
        col.compact()
        P = col.packed_words * 8
        k = (1 << b) // 3
        t_lt = timeit(lambda: col.scan_filter(adacom.CMP_LT, k))
        t_eq = timeit(lambda: col.scan_filter(adacom.CMP_EQ, k))
        t_bm = timeit(lambda: col.scan_filter(adacom.CMP_LT, k, device_bitmap=bitmap.data_ptr()))
        print(json.dumps(dict(b=b, P_MB=round(P / 1e6, 1), lt_ms=round(t_lt, 4), lt_GBs=round(P / t_lt / 1e6, 1),
                              eq_ms=round(t_eq, 4), eq_GBs=round(P / t_eq / 1e6, 1), lt_bitmap_ms=round(t_bm, 4),
                              lt_bitmap_GBs=round((P + n / 8) / t_bm / 1e6, 1))), flush=True)
        col.close()
    ctx.close()


main()
