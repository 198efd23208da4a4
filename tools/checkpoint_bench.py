"""Checkpoint / restore throughput of one INT64 column of 2^27 values (BASELINE config 2 shape), image resident in HBM.
Algorithmic bytes: checkpoint = payload read + n_blocks * 262144 written; restore = n_blocks * 262144 read + payload
written.  The CUDA-event bracket covers the whole C-ABI call (host planning of the block placement included)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402

import duckdb_adaptive_compression_b200 as adacom  # noqa: E402


def main():
    import torch
    logn = int(sys.argv[1]) if len(sys.argv) > 1 else 27
    widths = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [4, 17, 40, 63]
    n = 1 << logn
    ctx = adacom.Context(0)
    counts = adacom.row_group_segment_counts(n, np.int64)
    e0, e1 = ctx.event(), ctx.event()
    for b in widths:
        col = ctx.column(np.int64, counts)
        col.fill_splitmix(0xADAC0000 + b, 0, (1 << b) - 1)
        col.compact()
        payload = sum(m.words for m in col.meta()) * 8
        size = col.checkpoint_size()
        buf = torch.empty(size, dtype=torch.uint8, device="cuda:0")
        torch.cuda.synchronize()
        t_ck, t_rs = [], []
        for _ in range(5):
            ctx.flush_l2()
            ctx.record(e0)
            col.checkpoint(device_out=buf.data_ptr(), capacity=size)
            ctx.record(e1)
            t_ck.append(ctx.elapsed_ms(e0, e1))
            ctx.flush_l2()
            ctx.record(e0)
            back = ctx.restore(buf.data_ptr(), dtype=np.int64, device_bytes=size)
            ctx.record(e1)
            t_rs.append(ctx.elapsed_ms(e0, e1))
            back.close()
        blocks = size - 4096 * ((64 + 64 * len(counts) + 4095) // 4096)
        alg = payload + blocks
        print(json.dumps(dict(b=b, segments=len(counts), payload_MB=round(payload / 1e6, 1), image_MB=round(size / 1e6, 1),
                              checkpoint_ms=round(min(t_ck), 4), checkpoint_GBs=round(alg / min(t_ck) / 1e6, 1),
                              restore_ms=round(min(t_rs), 4), restore_GBs=round(alg / min(t_rs) / 1e6, 1))), flush=True)
        col.close()
    ctx.close()


main()
