#!/usr/bin/env python3
"""Per-region dynamic instruction counts of one kernel from `ncu --page source --csv --print-source sass` output (gzip ok).
usage: ncu_source_profile.py file.csv[.gz] kernel-substring [occurrence] [chunk]
Prints, per chunk of `chunk` SASS lines (default 150): warp-instructions executed per warp (first instruction's count = warps
launched), stall samples, and the region's landmarks (LDG/STG/LDS/STS/BAR/CALL counts)."""
import csv
import gzip
import sys
from collections import Counter

path, key = sys.argv[1], sys.argv[2]
occ = int(sys.argv[3]) if len(sys.argv) > 3 else 0
chunk = int(sys.argv[4]) if len(sys.argv) > 4 else 150
op = gzip.open if path.endswith(".gz") else open
rows = list(csv.reader(op(path, "rt")))
secs, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}
        secs.append(cur)
    elif r and r[0] == "Address":
        cur["hdr"] = r
    elif cur is not None and r:
        cur["rows"].append(r)
sel = [s for s in secs if key in s["name"]][occ]
h = sel["hdr"]
iS, iE, iN = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
R = sel["rows"]
warps = float(R[0][iE])
tot = sum(float(r[iE]) for r in R)
print("%s: %d SASS lines, %.0f warps, %.1f warp-instructions per warp, %d samples" % (sel["name"], len(R), warps, tot / warps, sum(int(r[iN]) for r in R)))
for a in range(0, len(R), chunk):
    part = R[a:a + chunk]
    ex = sum(float(r[iE]) for r in part) / warps
    sm = sum(int(r[iN]) for r in part)
    marks = Counter()
    ops = Counter()
    for r in part:
        o = r[iS].split()
        o = [x for x in o if not x.startswith("@")]
        name = o[0] if o else "?"
        w = float(r[iE]) / warps
        ops[name.split(".")[0]] += w
        for m in ("LDG", "STG", "LDS", "STS", "BAR", "CALL", "LDC", "ST.E", "LD.E"):
            if name.startswith(m):
                marks[m] += 1
    top = ", ".join("%s %.0f" % (k, v) for k, v in ops.most_common(6))
    print("lines %5d-%5d  executed/warp %7.1f  samples %6d  %s | %s" % (a, a + len(part) - 1, ex, sm, dict(marks), top))
