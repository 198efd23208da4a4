"""Top source lines of one kernel by executed warp-instructions and stall samples, from an .ncu-rep captured with
--import-source on (and -lineinfo).  usage: tools/ncu_src_lines.py rep kernel-regex [top] > profiles/<tag>_src_lines.txt"""
import csv
import io
import subprocess
import sys

rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kre}", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
cur, hdr, agg = None, None, []
for r in csv.reader(io.StringIO(raw)):
    if r and r[0] == "File Path":
        cur = r[1]
    elif r and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) == len(hdr) and r[0].isdigit():
        d = dict(zip(hdr, r))
        try:
            agg.append((int(d["Instructions Executed"]), int(d["# Samples"]), cur.split("/")[-1], int(d["Line No"])))
        except (ValueError, KeyError):
            pass
tot, ts = sum(a[0] for a in agg) or 1, sum(a[1] for a in agg) or 1
print(f"# {rep} kernel ~ {kre}: {tot} warp-instructions, {ts} stall samples")
srcs = {}
for a in sorted(agg, reverse=True)[:top]:
    path = [p for p in ("duckdb_adaptive_compression_b200/csrc/" + a[2],) if True][0]
    try:
        if path not in srcs:
            srcs[path] = open(path).read().splitlines()
        text = srcs[path][a[3] - 1].strip()[:100]
    except OSError:
        text = ""
    print(f"{a[0] / tot * 100:5.1f}% inst {a[1] / ts * 100:5.1f}% samples  {a[2]}:{a[3]}  {text}")
