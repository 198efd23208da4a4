#!/usr/bin/env python
"""DRAM traffic of the dominant kernel per launch, from the ncu CSV that tools/ncu_round.sh writes.

usage: tools/traffic_from_ncu.py gpurun_out/<tag>_traffic.csv profiles/roofline_traffic.json
The capture runs `bench.py --steps 1 --warmup 1`: decode_kernel<8,0,0> launches 0..127 are the warm-up step,
128..255 the timed step (scan + uncompact for each of the 64 widths); the later ones belong to bench.py's
sanity checks.  traffic = mean over the timed step of dram__bytes_read.sum + dram__bytes_write.sum."""
import collections
import csv
import json
import sys

SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1, "us": 1e3, "ms": 1e6}
src, dst = sys.argv[1], sys.argv[2]
rows = [r for r in csv.reader(open(src)) if len(r) > 5]
hdr = next(r for r in rows if r[0] == "ID")
ki, mi, vi, ui = (hdr.index(k) for k in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit"))
per = collections.OrderedDict()
for r in rows:
    if r[0].isdigit():
        d = per.setdefault(int(r[0]), {"name": r[ki]})
        d[r[mi]] = float(r[vi].replace(",", "")) * SCALE.get(r[ui], 1)
flat = [d for d in per.values() if "decode_kernel<8, 0, 0>" in d["name"]]
step = flat[128:256]
assert len(step) == 128, f"expected 256+ flat decode launches, got {len(flat)}"
traffic = sum(d["dram__bytes_read.sum"] + d["dram__bytes_write.sum"] for d in step) / len(step)
out = {"kernel": "decode_kernel<8,false,flat>", "launches": len(step),
       "traffic_bytes_per_launch": round(traffic), "mean_read_bytes": round(sum(d["dram__bytes_read.sum"] for d in step) / len(step)),
       "mean_write_bytes": round(sum(d["dram__bytes_write.sum"] for d in step) / len(step)),
       "mean_duration_us_under_ncu": round(sum(d["gpu__time_duration.sum"] for d in step) / len(step) / 1e3, 2),
       "source": src, "how": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum -k regex:decode_kernel, "
                            "bench.py --steps 1 --warmup 1 (widths 1-64, scan + uncompact launches of the timed step)"}
json.dump(out, open(dst, "w"), indent=1)
print(out)
