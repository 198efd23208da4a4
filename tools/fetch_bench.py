"""fetch_rows_kernel alone: 10^7 uniformly random (segment, row) fetches on a re-packed column, ids and results resident
in HBM (adacom_fetch_rows with on_device=1), CUDA-event timed.  Algorithmic bytes per fetch: 8 (ids) +
16 (<= 2 payload words) + S (result); the 64-byte descriptor is L2-resident.  Run under ncu for sectors/request."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import duckdb_adaptive_compression_b200 as adacom  # noqa: E402


def main():
    import torch
    n_fetch = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
    cases = [(np.uint32, 17, 28), (np.int64, 40, 27), (np.uint8, 5, 28)]
    ctx = adacom.Context(0)
    e0, e1 = ctx.event(), ctx.event()
    rng = np.random.default_rng(7)
    for dtype, b, logn in cases:
        dtype = np.dtype(dtype)
        n = 1 << logn
        counts = adacom.row_group_segment_counts(n, dtype)
        col = ctx.column(dtype, counts)
        col.fill_splitmix(0xFE7C0000 + b, 0, (1 << b) - 1)
        col.compact()
        cnt = np.asarray(counts, dtype=np.int64)
        seg = rng.integers(0, len(cnt), size=n_fetch, dtype=np.int64)
        row = (rng.random(n_fetch) * cnt[seg]).astype(np.int64)
        d_seg = torch.from_numpy(seg.astype(np.uint32).view(np.int32)).cuda()
        d_row = torch.from_numpy(row.astype(np.uint32).view(np.int32)).cuda()
        d_out = torch.zeros(n_fetch * dtype.itemsize, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(5):
            ctx.flush_l2()
            ctx.record(e0)
            ctx.check(ctx.lib.adacom_fetch_rows(col.h, n_fetch, d_seg.data_ptr(), d_row.data_ptr(), 0, d_out.data_ptr(), 1))
            ctx.record(e1)
            best = min(best, ctx.elapsed_ms(e0, e1))
        # spot check against a host-side fetch of the first 1000
        got = d_out.cpu().numpy().view(dtype)[:1000]
        ref = col.fetch_rows(seg[:1000], row[:1000])
        assert np.array_equal(got, ref)
        alg = n_fetch * (8 + 16 + dtype.itemsize)
        print(json.dumps(dict(dtype=dtype.name, width=b, values=n, fetches=n_fetch, ms=round(best, 4),
                              Mfetch_per_s=round(n_fetch / best / 1e3, 1), algorithmic_GBs=round(alg / best / 1e6, 1),
                              sector_GBs=round(n_fetch * (8 + 32 + dtype.itemsize) / best / 1e6, 1))), flush=True)
        col.close()
    ctx.close()


main()
