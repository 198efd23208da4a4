#!/usr/bin/env python
"""Per-kernel totals and shares from an `ncu --metrics gpu__time_duration.sum --csv` launch list."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
for i, r in enumerate(rows):
    if "Kernel Name" in r:
        h, start = r, i + 1
        break
ki, vi = h.index("Kernel Name"), h.index("Metric Value")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[start:]:
    if len(r) <= vi:
        continue
    k = r[ki].split("(")[0].replace("void ", "").replace("adacom::<unnamed>::", "")[-44:]
    agg[k][0] += 1
    agg[k][1] += float(r[vi].replace(",", ""))
tot = sum(v[1] for v in agg.values())
print(f"# {sys.argv[1]}: cold-cache serialised launch times (compare shares, not absolutes)")
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{k:46s} n={v[0]:4d} total_us={v[1] / 1e3:10.1f} avg_us={v[1] / 1e3 / v[0]:8.1f} share={v[1] / tot:.3f}")
