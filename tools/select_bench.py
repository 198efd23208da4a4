"""General FilterSelection pushed to one column (adacom_scan_select): 2^27 INT64 values, a validity mask with 10 % NULLs
resident in HBM, one comparison / a BETWEEN (AND of two) / an OR / IS NOT NULL, with and without the selection bitmap.
Call times incl. the D2H of the aggregate."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import duckdb_adaptive_compression_b200 as adacom  # noqa: E402


def main():
    import torch
    logn = int(sys.argv[1]) if len(sys.argv) > 1 else 27
    widths = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [8, 17, 40]
    n = 1 << logn
    ctx = adacom.Context(0)
    counts = adacom.row_group_segment_counts(n, np.int64)
    rng = np.random.default_rng(1)
    valid = torch.from_numpy((rng.random(n // 64 + 1) < 2).astype(np.int64)).cuda()  # placeholder, replaced below
    words = rng.integers(0, 1 << 63, size=n // 64 + 1, dtype=np.int64) | rng.integers(0, 1 << 63, size=n // 64 + 1, dtype=np.int64) \
        | rng.integers(0, 1 << 63, size=n // 64 + 1, dtype=np.int64)  # ~87.5 % of the bits set
    valid = torch.from_numpy(words).cuda()
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
        col.fill_splitmix(0x5E1EC7 + b, 0, (1 << b) - 1)
        col.compact()
        P = col.packed_words * 8
        k = (1 << b) // 3
        v = valid.data_ptr()
        cases = {"lt": [("cmp", adacom.CMP_LT, k)], "between": [("cmp", adacom.CMP_GE, k // 2), ("cmp", adacom.CMP_LE, k)],
                 "is_not_null": [("is_not_null",)]}
        out = {"b": b, "P_MB": round(P / 1e6, 1)}
        for name, preds in cases.items():
            t = timeit(lambda: col.scan_select(preds, 0, validity=v, validity_on_device=True))
            out[name + "_ms"] = round(t, 4)
        t = timeit(lambda: col.scan_select([("cmp", adacom.CMP_LT, k), ("cmp", adacom.CMP_GT, 2 * k)], 1, validity=v, validity_on_device=True))
        out["or_ms"] = round(t, 4)
        t = timeit(lambda: col.scan_select(cases["lt"], 0, validity=v, validity_on_device=True, bitmap_out=bitmap.data_ptr()))
        out["lt_bitmap_ms"] = round(t, 4)
        out["lt_GBs"] = round((P + n / 8) / out["lt_ms"] / 1e6, 1)
        print(json.dumps(out), flush=True)
        col.close()
    ctx.close()


main()
