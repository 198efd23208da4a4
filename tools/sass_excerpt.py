"""SASS evidence for the shipped library (no GPU needed): per kernel, the number of bulk-copy (UBLKCP = cp.async.bulk),
mbarrier (SYNCS), funnel-shift (SHF) and global load/store instructions in the sm_100a cubin of libadacom_gpu.so.
usage: python tools/sass_excerpt.py > profiles/<tag>_sass_excerpt.txt"""
import collections
import os
import re
import subprocess
import sys

so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "duckdb_adaptive_compression_b200", "csrc",
                  "libadacom_gpu.so")
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True, check=True).stdout
arch = sorted(set(re.findall(r"arch = (sm_\w+)", sass)))
counts = collections.OrderedDict()
name = None
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", "-p", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = name.replace("(anonymous namespace)::", "").replace("void ", "").replace("adacom::", "")
        counts[name] = collections.Counter()
        continue
    if name is None:
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        op = m.group(1)
        c = counts[name]
        c["total"] += 1
        for key in ("UBLKCP", "SYNCS", "UTMA", "SHF", "LDG", "STG", "LDS", "STS", "ATOM", "RED", "IMAD", "LOP3", "LDGSTS"):
            if op.startswith(key):
                c[key] += 1
print(f"library: {os.path.relpath(so)}  arch: {', '.join(arch)}  kernels: {len(counts)}")
tot = collections.Counter()
for c in counts.values():
    tot.update(c)
print("whole library: " + "  ".join(f"{k}={tot[k]}" for k in ("total", "UBLKCP", "SYNCS", "UTMA", "LDGSTS", "SHF", "LDG", "STG", "ATOM", "RED")))
print()
# one line per kernel family (template instances folded): instance count, instructions, bulk copies, mbarrier ops
fam = collections.OrderedDict()
for n, c in counts.items():
    f = re.sub(r"<.*", "", n)
    d = fam.setdefault(f, collections.Counter())
    d.update(c)
    d["instances"] += 1
print(f"{'kernel family':40s} {'inst':>5s} {'SASS':>8s} {'UBLKCP':>7s} {'SYNCS':>6s} {'LDGSTS':>7s} {'SHF':>6s} {'LDG':>6s} {'STG':>6s} {'ATOM+RED':>8s}")
for f, d in fam.items():
    print(f"{f[:40]:40s} {d['instances']:5d} {d['total']:8d} {d['UBLKCP']:7d} {d['SYNCS']:6d} {d['LDGSTS']:7d} {d['SHF']:6d} {d['LDG']:6d} {d['STG']:6d} {d['ATOM'] + d['RED']:8d}")
