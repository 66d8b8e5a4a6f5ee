#!/usr/bin/env python3
"""Opcode mix and a pipe-cycle estimate of one kernel's SASS, with loop trip counts applied to address ranges.
usage: sass_mix.py <binary> <kernel-substring> [lo:hi:trips ...]   (hex byte offsets; default = static count)
Pipe model measured on B200 (profiles/r01_int_pipes.txt): IMAD.WIDE / IMAD.HI 4 cycles of the fmaheavy pipe per warp
instruction per SM sub-partition, other IMAD forms 2, ALU ops (IADD3, LOP3, SHF, LEA, SEL, ISETP, PRMT, MOV) 2, one issue per cycle."""
import re
import subprocess
import sys
from collections import Counter

path, kname = sys.argv[1], sys.argv[2]
ranges = [tuple(int(v, 16) if i < 2 else float(v) for i, v in enumerate(a.split(":"))) for a in sys.argv[3:]]
txt = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
cur, funcs = None, {}
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        funcs[cur] = []
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", line)
    if m and cur:
        funcs[cur].append((int(m.group(1), 16), m.group(2).strip()))
name = [f for f in funcs if kname in f][0]
ins = funcs[name]
print("kernel", name, "static instructions", len(ins))
for a, t in ins:
    if re.search(r"\bBRA\b", t):
        m = re.search(r"(0x[0-9a-f]+)", t)
        if m and int(m.group(1), 16) <= a:
            print("  backward branch at %#x -> %s" % (a, m.group(1)))
c = Counter()
for a, t in ins:
    w = 1.0 if not ranges else 0.0
    for lo, hi, trips in ranges:
        if lo <= a < hi:
            w = max(w, 1.0) * trips if w else trips
    p = t.split()
    op = p[1] if p[0].startswith("@") else p[0]
    if op.startswith("IMAD"):
        k = "IMAD.WIDE" if "WIDE" in op else "IMAD.HI" if ".HI" in op else "IMAD.other"
    else:
        k = op.split(".")[0]
    c[k] += w
tot = sum(c.values())
fma = 4 * (c["IMAD.WIDE"] + c["IMAD.HI"]) + 2 * c["IMAD.other"]
alu_ops = sum(v for k, v in c.items() if k in ("IADD3", "LOP3", "SHF", "LEA", "SEL", "ISETP", "PRMT", "MOV", "IABS", "VIADD", "PLOP3", "P2R", "R2P", "LEA.HI"))
print("weighted instructions %.0f ; fmaheavy cycles %.0f ; ALU cycles %.0f ; issue cycles %.0f" % (tot, fma, 2 * alu_ops, tot))
for k, v in c.most_common(16):
    print("  %-12s %9.0f" % (k, v))
