#!/usr/bin/env python3
"""Dynamic instruction estimate of one kernel from its SASS with hand-given region weights.

usage: sass_regions.py <binary> <kernel-substring> start:end:weight [start:end:weight ...]
Addresses in hex (inclusive); instructions outside every region get weight `default` (1).
Prints per-pipe counts using the B200 issue costs measured by tools/ubench/int_pipes.cu:
  IMAD.WIDE / IMAD.HI: 4 cycles of the fma-heavy pipe per warp instruction; other IMAD*: 2;
  IADD3/LOP3/SHF/SEL/MOV/PRMT/LEA/ISETP...: 2 cycles of the alu pipe.
"""
import re
import subprocess
import sys
from collections import Counter


def main():
    path, kname = sys.argv[1], sys.argv[2]
    regions = []
    for a in sys.argv[3:]:
        s, e, w = a.split(":")
        regions.append((int(s, 16), int(e, 16), float(w)))
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
    byop = Counter()
    fma_cyc = alu_cyc = tot = 0.0
    for addr, text in ins:
        w = 1.0
        for s, e, ww in regions:
            if s <= addr <= e:
                w = ww
        toks = text.split()
        op = toks[1] if toks[0].startswith("@") else toks[0]
        base = op.split(".")[0]
        if base == "NOP":
            continue
        tot += w
        key = op if base == "IMAD" else base
        byop[key] += w
        if base == "IMAD":
            fma_cyc += w * (4 if (".WIDE" in op or ".HI" in op) else 2)
        elif base in ("FFMA", "FMUL", "FADD", "HFMA2"):
            fma_cyc += w * 1
        elif base in ("LDG", "STG", "LDS", "STS", "LDC", "LDCU", "BAR", "BRA", "EXIT", "S2R", "S2UR", "UMOV", "UIADD3", "UISETP", "ULEA", "UIMAD", "ULOP3", "USHF", "R2UR", "BSSY", "BSYNC", "WARPSYNC", "CS2R"):
            pass
        else:
            alu_cyc += w * 2
    print("kernel:", names[0], " static instructions:", len(ins), "(%.1f KB)" % (len(ins) * 16 / 1024))
    print("dynamic warp-instructions: %.0f   fma-heavy cycles %.0f   alu cycles %.0f   (sum %.0f)" % (tot, fma_cyc, alu_cyc, fma_cyc + alu_cyc))
    for k, v in byop.most_common(16):
        print("  %-16s %8.0f" % (k, v))


if __name__ == "__main__":
    main()
