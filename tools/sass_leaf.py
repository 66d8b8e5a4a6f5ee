#!/usr/bin/env python3
"""Dynamic instruction mix of ONE Poseidon permutation of k_linear_hash_rows, from the SASS of a binary.

usage: sass_leaf.py <binary> [kernel-substring]

The kernel holds two copies of the permutation (the deferred-wrap one, first in address order, and the out-of-line
exact redo).  Loops are found from the backward branches; in the first copy they are, outermost first: the sponge loop
(one trip = one permutation), the 8 full rounds, and inside it the 12-trip pre-matrix loop and the 22-trip partial-round
loop (executed once, in round 3; the pre-matrix loop computes POS_PRE_UNROLL outputs per trip, found from its IMAD.WIDE count).  Every instruction gets the trip count of its innermost loop; code of the full-round
loop that belongs to the round-3 branch only is found as the span between the pre-matrix loop and the end of the
partial-round loop and counted once, the MDS layer (7 of 8 rounds) is counted 8 times: a < 2 % over-estimate.
Pipe model (tools/ubench/int_pipes.cu): IMAD.WIDE / IMAD.HI 4 cycles of the fmaheavy pipe per warp instruction, other
IMAD 2, DFMA/DADD 2 cycles of the FP64 pipe, ALU instructions 2 cycles, one issue slot each."""
import re
import subprocess
import sys
from collections import Counter

args = [a for a in sys.argv[1:] if not a.startswith("--chunk")]
chunk = [int(a.split("=")[1]) for a in sys.argv[1:] if a.startswith("--chunk=")]
path = args[0]
kname = args[1] if len(args) > 1 else "k_linear_hash_rows"
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
loops = []
for a, t in ins:
    if re.search(r"\bBRA\b", t):
        m = re.search(r"(0x[0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a:
            loops.append((int(m.group(1), 16), a))
loops.sort()
# first copy: sponge, full, pre, partial in order of start address (pre and partial are nested in full, full in sponge)
sponge = loops[0]
inner = [l for l in loops if sponge[0] < l[0] and l[1] < sponge[1]]
full = max(inner, key=lambda l: l[1] - l[0])
nested = sorted(l for l in inner if full[0] < l[0] and l[1] < full[1])
pre, part = nested[0], nested[1]
print("kernel %s: %d static instructions; sponge %#x-%#x full %#x-%#x pre %#x-%#x partial %#x-%#x" %
      (name, len(ins), *sponge, *full, *pre, *part))


# outputs per trip of the pre-matrix loop (POS_PRE_UNROLL): 48 dot-product IMAD.WIDE + 1 in the fold per output
pre_wide = sum(1 for a, t in ins if pre[0] <= a <= pre[1] and "IMAD.WIDE" in t)
pre_trips = 12 // max(1, round(pre_wide / 49.0))


# rounds per trip of the partial-round loop (POS_PR_CHUNK): from --chunk=M, else from the loop's S-box squarings (2 per round)
part_sq = sum(1 for a, t in ins if part[0] <= a <= part[1] and re.search(r"IMAD\.WIDE\.U32 R\d+, (R\d+)(\.reuse)?, \1(\.reuse)?, RZ", t))
part_chunk = chunk[0] if chunk else max(1, round(part_sq / 4.0))      # x^2: a0*a0 and a1*a1 per squaring, two squarings per S-box
# a second update loop behind the partial-round loop = a tail chunk (22 % M rounds); without one the rounds are padded at the front
# to ceil(22 / M) trips (POS_PR_NOTAIL; the padded rounds skip their S-box and dot product: a slight over-count here)
has_tail = len(nested) > 2
part_trips = 22 // part_chunk if (has_tail or 22 % part_chunk == 0) else -(-22 // part_chunk)
print("partial-round loop: %d round(s) per trip, %d trips%s" % (part_chunk, part_trips, " (+ %d in the tail)" % (22 % part_chunk) if has_tail else ""))


# the rolled end-of-chunk update loops (11 trips, one per state word): every other loop of the round-3 branch -- the one inside the
# partial-round loop and the one of the tail chunk
upd = [l for l in nested[2:]] + [l for l in inner if part[0] < l[0] and l[1] < part[1]]
print("update loops (11 trips each):", ["%#x-%#x" % l for l in upd])


def weight(a):
    for l in upd:
        if l[0] <= a <= l[1]:
            return 11 * (part_trips if part[0] <= a <= part[1] else 1)
    if pre[0] <= a <= pre[1]:
        return pre_trips
    if part[0] <= a <= part[1]:
        return part_trips
    if pre[0] <= a <= full[1]:                 # the rest of the round-3 branch (laid out at the end of the loop body), incl. the tail chunk
        return 1
    if full[0] <= a <= full[1]:
        return 8
    if sponge[0] <= a <= sponge[1]:
        return 1
    return 0


c = Counter()
for a, t in ins:
    w = weight(a)
    if not w:
        continue
    p = t.split()
    op = p[1] if p[0].startswith("@") else p[0]
    if op.startswith("IMAD"):
        k = "IMAD.WIDE" if "WIDE" in op else "IMAD.HI" if ".HI" in op else "IMAD." + (op.split(".")[1] if "." in op else "LO")
    else:
        k = op.split(".")[0]
    c[k] += w
tot = sum(c.values())
wide = c["IMAD.WIDE"] + c["IMAD.HI"]
oimad = sum(v for k, v in c.items() if k.startswith("IMAD.") and k not in ("IMAD.WIDE", "IMAD.HI"))
fp64 = c["DFMA"] + c["DADD"] + c["DMUL"]
alu = sum(v for k, v in c.items() if k in ("IADD3", "LOP3", "SHF", "LEA", "SEL", "ISETP", "PRMT", "MOV", "VIMNMX3", "VIMNMX", "VIADD", "PLOP3", "P2R", "R2P", "CS2R"))
print("per permutation: %d warp instructions; fmaheavy cycles %d (IMAD.WIDE %d, IMAD.HI %d, other IMAD %d); FP64 cycles %d; ALU cycles %d"
      % (tot, 4 * wide + 2 * oimad, c["IMAD.WIDE"], c["IMAD.HI"], oimad, 2 * fp64, 2 * alu))
for k, v in c.most_common(24):
    print("  %-12s %7d" % (k, v))
