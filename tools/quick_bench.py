"""Developer timing loop: per-kernel CUDA-event times for a few widths on one INT64 column of 2^27 values."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import duckdb_adaptive_compression_b200 as adacom

def main():
    logn = int(sys.argv[1]) if len(sys.argv) > 1 else 27
    widths = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 4, 8, 13, 16, 24, 32, 40, 48, 63]
    dtype = np.dtype(sys.argv[3]) if len(sys.argv) > 3 else np.dtype(np.int64)
    n = 1 << logn
    S = dtype.itemsize
    ctx = adacom.Context(0)
    counts = adacom.row_group_segment_counts(n, dtype)
    import torch
    out = torch.empty(n * S + 256, dtype=torch.uint8, device="cuda")
    e0, e1 = ctx.event(), ctx.event()
    def timeit(fn, reps=5):
        best = 1e9
        for _ in range(reps):
            ctx.flush_l2()
            ctx.record(e0); fn(); ctx.record(e1)
            best = min(best, ctx.elapsed_ms(e0, e1))
        return best
    for b in widths:
        col = ctx.column(dtype, counts)
        mask = (1 << b) - 1
        col.fill_splitmix(0xADAC0000 + b, 0, mask)
        ctx.synchronize()
        t_an = timeit(lambda: col.analyze())
        # encode needs a fresh (uncompacted) state each rep
        def enc():
            col.encode()
        col.analyze()
        ts = []
        for _ in range(3):
            ctx.flush_l2()
            ctx.record(e0); col.encode(); ctx.record(e1)
            ts.append(ctx.elapsed_ms(e0, e1))
            m = col.meta(0, 1)[0]
            if _ < 2:
                col.uncompact()
        t_en = min(ts)
        P = col.packed_words * 8
        t_de = timeit(lambda: col.decode(device_out=out.data_ptr()))
        agg = torch.zeros(4, dtype=torch.int64, device='cuda')
        col.scan_sum()
        t_su = timeit(lambda: col.scan_sum_async(agg.data_ptr()))
        GB = 1e-9
        raw = n * S
        print(json.dumps(dict(b=b, w=m.width, raw_GB=raw*GB, P_GB=P*GB,
              analyze_ms=round(t_an,4), analyze_GBs=round(raw*GB/(t_an*1e-3),1),
              encode_ms=round(t_en,4), encode_GBs=round((raw+P)*GB/(t_en*1e-3),1),
              decode_ms=round(t_de,4), decode_GBs=round((raw+P)*GB/(t_de*1e-3),1),
              sum_ms=round(t_su,4), sum_GBs=round(P*GB/(t_su*1e-3),1))), flush=True)
        col.close()
    ctx.close()

main()
