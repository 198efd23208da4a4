"""Summarise an ncu report offline: headline raw metrics + instruction / stall-sample share per source line.
usage: python tools/ncu_src_summary.py <report.ncu-rep> [top_n]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
idx = {h: i for i, h in enumerate(hdr)}
want = ["Kernel Name", "gpu__time_duration.sum", "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__warps_eligible.avg.per_cycle_active",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"]
for r in rows[2:]:
    print({w: r[idx[w]] for w in want if w in idx})
    st = {h: r[i] for h, i in idx.items() if "average_warps_issue_stalled" in h and h.endswith(".ratio") and "not_issued" not in h}
    print("  stalls:", ", ".join(f"{k.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')}={float(v):.2f}"
                                 for k, v in sorted(st.items(), key=lambda kv: -float(kv[1] or 0))[:8]))
src_csv = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"], capture_output=True, text=True).stdout
hdr = None
cur = None
inst, samp, text = collections.Counter(), collections.Counter(), {}
for r in csv.reader(io.StringIO(src_csv)):
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        continue
    if hdr and r and r[0].isdigit():
        key = (cur, int(r[0]))
        try:
            inst[key] += int(r[hdr.index("Instructions Executed")])
        except ValueError:
            pass
        try:
            samp[key] += int(r[hdr.index("# Samples")])
        except ValueError:
            pass
        text[key] = r[1][:100]
ti, ts = sum(inst.values()) or 1, sum(samp.values()) or 1
print("total instructions", ti, "samples", ts)
for k, v in inst.most_common(top):
    print(f"{k[0]}:{k[1]:4d} inst {v / ti * 100:5.1f}% samp {samp[k] / ts * 100:5.1f}% | {text[k]}")
