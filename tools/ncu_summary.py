#!/usr/bin/env python
"""Summarise an .ncu-rep (ncu --set full) into a small text table for profiles/.
usage: tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x_summary.txt"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = [("Kernel Name", "kernel"), ("launch__grid_size", "grid"), ("launch__registers_per_thread", "regs"),
        ("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_conflicts"),
        ("smsp__average_warp_latency_issue_stalled_barrier.ratio", "stall_barrier"),
        ("smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "stall_long_sb"),
        ("smsp__average_warp_latency_issue_stalled_short_scoreboard.ratio", "stall_short_sb"),
        ("smsp__average_warp_latency_issue_stalled_mio_throttle.ratio", "stall_mio"),
        ("smsp__average_warp_latency_issue_stalled_membar.ratio", "stall_membar"),
        ("smsp__average_warp_latency_issue_stalled_wait.ratio", "stall_wait"),
        ("smsp__average_warp_latency_issue_stalled_lg_throttle.ratio", "stall_lg")]
idx = [(hdr.index(k), n) for k, n in want if k in hdr]
print("# source:", rep)
print("# units:", ", ".join(f"{n}={units[i]}" for i, n in idx if units[i]))
print("\t".join(n for _, n in idx))
for r in rows[2:]:
    out = []
    for i, n in idx:
        v = r[i]
        if n == "kernel":
            v = v.replace("void ", "").replace("adacom::<unnamed>::", "")[:48]
        else:
            try:
                v = f"{float(v.replace(',', '')):.4g}"
            except ValueError:
                pass
        out.append(v)
    print("\t".join(out))
