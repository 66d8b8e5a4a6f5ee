#!/usr/bin/env python3
"""Dynamic instruction estimate for a kernel from its SASS (no GPU needed).

usage: sass_count.py <cubin-or-.so> <kernel-substring> [trip counts of the backward-branch loops, in address order]
Counts instructions per opcode class, weighting each loop body (found from backward BRA) by the trip
count given on the command line (default 1).  Pipe model from B300_MICROARCH.md: IMAD* on the fma pipe,
IADD3/LOP3/SHF/SEL/ISETP/MOV/PRMT... on the alu pipe, both half-rate per SMSP, one issue per cycle.
"""
import re
import subprocess
import sys
from collections import Counter


def main():
    path, kname = sys.argv[1], sys.argv[2]
    trips = [float(x) for x in sys.argv[3:]]
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
    names = [f for f in funcs if kname in f]
    assert names, "no kernel matching " + kname
    ins = funcs[names[0]]
    loops = []
    for addr, text in ins:
        m = re.search(r"\bBRA(?:\.U)?\b.*?(0x[0-9a-f]+)", text)
        if m:
            tgt = int(m.group(1), 16)
            if tgt <= addr:
                loops.append((tgt, addr))
    loops.sort()
    weight = {}
    for i, (lo, hi) in enumerate(loops):
        weight[(lo, hi)] = trips[i] if i < len(trips) else 1.0
    tot, byop = 0.0, Counter()
    for addr, text in ins:
        w = 1.0
        for (lo, hi), t in weight.items():
            if lo <= addr <= hi:
                w *= t
        op = text.split()[0]
        if op.startswith("@"):
            op = text.split()[1]
        base = op.split(".")[0]
        if base in ("NOP",):
            continue
        byop[base] += w
        tot += w
    fma = sum(v for k, v in byop.items() if k in ("IMAD", "FFMA", "FMUL", "FADD", "HFMA2", "IMUL"))
    mem = sum(v for k, v in byop.items() if k in ("LDG", "STG", "LDS", "STS", "LDC", "LDCU", "LD", "ST", "LDL", "STL"))
    alu = tot - fma - mem
    print("kernel:", names[0])
    print("loops (start,end,trip):", [(hex(a), hex(b), weight[(a, b)]) for a, b in loops])
    print("weighted instructions: %.0f   fma-pipe %.0f   alu/other %.0f   mem %.0f" % (tot, fma, alu, mem))
    print("issue-bound cycles/warp >= %.0f ; pipe-bound >= %.0f" % (tot, 2 * max(fma, alu)))
    for k, v in byop.most_common(14):
        print("  %-10s %8.0f" % (k, v))


if __name__ == "__main__":
    main()
