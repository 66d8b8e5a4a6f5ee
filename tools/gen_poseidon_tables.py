#!/usr/bin/env python3
"""Emit the Poseidon-Goldilocks parameter tables used by the CUDA kernels and by the C oracle.

The numbers are the published Poseidon parameters for the Goldilocks field (width 12, 8 full +
22 partial rounds, x^7) in the "optimised" form the reference evaluates
(/root/reference/src/goldilocks/src/poseidon_goldilocks_constants.hpp:931-2139, non-Montgomery
branch; USE_MONTGOMERY is 0, goldilocks_base_field.hpp:11):
    C[118]  round constants (12 + 4*12 + 22 + 3*12)
    M[12][12]  MDS matrix (circulant + diagonal, entries <= 41)
    P[12][12]  pre-multiplied matrix applied once before the partial rounds
    S[22*23]   sparse partial-round vectors
This script runs only in the build container (the reference tree is absent on the GPU box); its
outputs are committed:  xlayer_prover_b200/csrc/poseidon_tables.cuh  and  oracle/poseidon_constants.h.
It parses numbers only and re-checks the structure (MDS circulant) and the reference's
own known-answer test on a pure-Python restatement before writing anything.
"""
import os
import re
import sys

REF = os.environ.get("REF", "/root/reference")
SRC = os.path.join(REF, "src/goldilocks/src/poseidon_goldilocks_constants.hpp")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = 0xFFFFFFFF00000001


def parse():
    text = open(SRC).read()
    text = text[text.index("#else"):]          # non-Montgomery branch
    out = {}
    for name, count in (("C", 118), ("M", 144), ("P", 144), ("S", 507)):
        m = re.search(r"Element\s+%s(\[\d+\])+\s*=?\s*\{" % name, text)
        body = text[m.end():]
        # the initialiser ends at the first "};" after it starts
        body = body[:body.index("};")]
        vals = [int(x, 16) for x in re.findall(r"0x[0-9a-fA-F]+", body)]
        if name == "S" and len(vals) == 506:
            vals.append(0)                     # declared [507], 22*23 = 506 initialisers
        assert len(vals) == count, (name, len(vals))
        out[name] = vals
    return out


def permute(T, x):
    C, M, Pm, S = T["C"], T["M"], T["P"], T["S"]
    x = [(a + c) % P for a, c in zip(x, C[:12])]
    def mat(x, A): return [sum(A[j * 12 + i] * x[j] for j in range(12)) % P for i in range(12)]
    for r in range(3):
        x = [(pow(a, 7, P) + C[12 * (r + 1) + i]) % P for i, a in enumerate(x)]
        x = mat(x, M)
    x = [(pow(a, 7, P) + C[48 + i]) % P for i, a in enumerate(x)]
    x = mat(x, Pm)
    for r in range(22):
        x[0] = (pow(x[0], 7, P) + C[60 + r]) % P
        s0 = sum(x[i] * S[23 * r + i] for i in range(12)) % P
        for i in range(1, 12):
            x[i] = (x[i] + x[0] * S[23 * r + 11 + i]) % P
        x[0] = s0
    for r in range(3):
        x = [(pow(a, 7, P) + C[82 + 12 * r + i]) % P for i, a in enumerate(x)]
        x = mat(x, M)
    x = [pow(a, 7, P) for a in x]
    return mat(x, M)


def spw(T, dmax=7):
    """Partial rounds taken M at a time (poseidon.cuh POS_PR_CHUNK): the rank-1 updates x_i += y_j * v_i(j) of a chunk are
    deferred to its end, so the dot product of round k picks up y_j * W for every earlier round j of the chunk, with
    W = sum_{i>=1} v_i(j) * S_i(k).  W turns out to depend on d = k - j only and to be SMALL (13, 21, 29, ... bits: entries of
    powers of the MDS matrix) -- both facts are asserted here, the kernels rely on them."""
    S = T["S"]
    out = []
    for d in range(1, dmax + 1):
        vals = {sum(S[23 * j + 11 + i] * S[23 * (j + d) + i] for i in range(1, 12)) % P for j in range(0, 22 - d)}
        assert len(vals) == 1, d
        out.append(vals.pop())
    assert all(S[23 * r] == 25 for r in range(22))       # the diagonal weight of every partial round is M[0][0] = 17 + 8
    assert out[0] < 2 ** 32 and out[1] < 2 ** 32 and out[2] < 2 ** 32
    return out


def permute_chunked(T, x, m=4):
    """permute() with the partial rounds taken m at a time (then the remaining 22 % m), as the kernels do -- checked against permute()"""
    C, M, Pm, S = T["C"], T["M"], T["P"], T["S"]
    W = spw(T)
    x = [(a + c) % P for a, c in zip(x, C[:12])]
    def mat(x, A): return [sum(A[j * 12 + i] * x[j] for j in range(12)) % P for i in range(12)]
    for r in range(3):
        x = [(pow(a, 7, P) + C[12 * (r + 1) + i]) % P for i, a in enumerate(x)]
        x = mat(x, M)
    x = [(pow(a, 7, P) + C[48 + i]) % P for i, a in enumerate(x)]
    x = mat(x, Pm)
    r = 0
    while r < 22:
        mm = min(m, 22 - r)
        ys, x0 = [], x[0]
        for k in range(mm):
            y = (pow(x0, 7, P) + C[60 + r + k]) % P
            ys.append(y)
            x0 = (25 * y + sum(x[i] * S[23 * (r + k) + i] for i in range(1, 12)) + sum(ys[j] * W[k - j - 1] for j in range(k))) % P
        for i in range(1, 12):
            x[i] = (x[i] + sum(ys[k] * S[23 * (r + k) + 11 + i] for k in range(mm))) % P
        x[0] = x0
        r += mm
    for r in range(3):
        x = [(pow(a, 7, P) + C[82 + 12 * r + i]) % P for i, a in enumerate(x)]
        x = mat(x, M)
    x = [pow(a, 7, P) for a in x]
    return mat(x, M)


def fmt(vals, per=4):
    lines = []
    for i in range(0, len(vals), per):
        lines.append("    " + ", ".join("0x%016xULL" % v for v in vals[i:i + per]) + ",")
    return "\n".join(lines)


def main():
    T = parse()
    M = T["M"]
    circ = [17, 15, 41, 16, 2, 28, 13, 13, 39, 18, 34, 20]
    for r in range(12):
        for c in range(12):
            want = circ[(r - c) % 12] + (8 if r == c == 0 else 0)
            assert M[r * 12 + c] == want, (r, c)
    fib = [0, 1]
    for _ in range(10):
        fib.append((fib[-1] + fib[-2]) % P)
    got = permute(T, fib)
    assert got[0] == 0x3095570037F4605D and got[11] == 0xF58A41D4028C66A5, hex(got[0])
    assert permute(T, [0] * 12)[0] == 0x3C18A9786CB0B359
    for m in (2, 3, 4, 7):
        assert permute_chunked(T, fib, m) == got and permute_chunked(T, list(range(5, 17)), m) == permute(T, list(range(5, 17))), m
    rcf = T["C"][12:60] + T["C"][82:118] + [0] * 12
    mrc = []
    for r in range(8):
        A = T["P"] if r == 3 else M
        mrc += [sum(A[j * 12 + i] * rcf[r * 12 + j] for j in range(12)) % P for i in range(12)]
    mrcl = []
    for v in mrc:
        mrcl += [v & 0x3FFFFF, (v >> 22) & 0x3FFFFF, v >> 44]
    hdr = ("// GENERATED by tools/gen_poseidon_tables.py -- do not edit.\n"
           "// Poseidon-Goldilocks parameters (width 12, 4+22+4 rounds, x^7), optimised form; values as in\n"
           "// the reference's poseidon_goldilocks_constants.hpp:931-2139 (non-Montgomery).  Layout:\n"
           "//   RC[118]: 0..11 initial add | 12..59 full rounds 1..4 | 60..81 partial | 82..117 full rounds 5..7\n"
           "//   MDS_CIRC[12]: M[r][c] = MDS_CIRC[(r-c) mod 12] (+8 at r=c=0); state' = M^T state\n"
           "//   PRE[144]: PRE[j*12+i] multiplies old state j into new state i\n"
           "//   SP[22*23]: per partial round: 12 dot-product weights, then 11 rank-1 update weights\n")
    cu = [hdr, "#pragma once", "#include <stdint.h>", "",
          "#define GLP_POSEIDON_RC_N 118", "#define GLP_POSEIDON_SP_N 506", "",
          "#define GLP_POSEIDON_RC_INIT \\\n" + fmt(T["C"]).replace("\n", " \\\n"), "",
          "#define GLP_POSEIDON_PRE_INIT \\\n" + fmt(T["P"]).replace("\n", " \\\n"), "",
          "#define GLP_POSEIDON_SP_INIT \\\n" + fmt(T["S"][:506]).replace("\n", " \\\n"), "",
          "// derived layouts for the rolled kernels: RCF[r*12+i] = constant added after the S-box of full round r\n"
          "// (r = 0..3: RC[12+12r+i], r = 4..6: RC[82+12(r-4)+i], r = 7: 0); PRET[i*12+j] = PRE[j*12+i]",
          "#define GLP_POSEIDON_RCF_INIT \\\n" + fmt(T["C"][12:60] + T["C"][82:118] + [0] * 12).replace("\n", " \\\n"), "",
          "#define GLP_POSEIDON_PRET_INIT \\\n" + fmt([T["P"][j * 12 + i] for i in range(12) for j in range(12)]).replace("\n", " \\\n"), "",
          "// the round constants pushed through the linear layer that follows them (the layer is linear, so\n"
          "// L(x^7 + c) = L(x^7) + L(c)): MRC[r*12+i] = sum_j A_r[j][i] * RCF[r*12+j] mod p, A_r = MDS (PRE for r = 3);\n"
          "// MRCL = the same values split into limbs of 22, 22 and 20 bits (the limb form of the MDS layer starts its\n"
          "// accumulators from them)",
          "#define GLP_POSEIDON_MRC_INIT \\\n" + fmt(mrc).replace("\n", " \\\n"), "",
          "#define GLP_POSEIDON_MRCL_INIT \\\n" + "\n".join("    " + ", ".join("0x%xu" % v for v in mrcl[i:i + 12]) + "," for i in range(0, len(mrcl), 12)).replace("\n", " \\\n"), "",
          "// MRCD[(r*12+i)*2 + h] = 2^52 + (32-bit half h of MRC[r*12+i]) as a double: the start value of the FP64 MDS chains",
          "#define GLP_POSEIDON_MRCD_INIT \\\n" + "\n".join("    " + ", ".join("%d.0" % (2 ** 52 + h) for v in mrc[i:i + 2] for h in (v & 0xFFFFFFFF, v >> 32)) + "," for i in range(0, len(mrc), 2)).replace("\n", " \\\n"), "",
          "// SPW[d-1] = sum_{i=1..11} SP[23j + 11 + i] * SP[23(j+d) + i] mod p (the same for every j): the weight the S-box output of\n"
          "// partial round j gets in the dot product of round j+d when the rank-1 updates of rounds j..j+d-1 are still deferred\n"
          "// (POS_PR_CHUNK, poseidon.cuh).  13, 21, 29, 36, ... bits.  SP[23j] (the diagonal weight) is 25 in every round.",
          "#define GLP_POSEIDON_SPW_INIT " + ", ".join("0x%xULL" % v for v in spw(T)), "#define GLP_POSEIDON_SP_DIAG 25", "",
          "#define GLP_MDS_CIRC_INIT " + ", ".join(str(v) for v in circ),
          "#define GLP_MDS_DIAG0 8", ""]
    open(os.path.join(ROOT, "xlayer_prover_b200/csrc/poseidon_tables.cuh"), "w").write("\n".join(cu))
    oc = ["/* TEST INFRASTRUCTURE ONLY (oracle). " + hdr.replace("//", " *").rstrip() + "\n */",
          "#ifndef GLP_ORACLE_POSEIDON_CONSTANTS_H", "#define GLP_ORACLE_POSEIDON_CONSTANTS_H", "#include <stdint.h>",
          "static const uint64_t ORC_C[118] = {\n" + fmt(T["C"]) + "\n};",
          "static const uint64_t ORC_M[144] = {\n" + fmt(T["M"]) + "\n};",
          "static const uint64_t ORC_P[144] = {\n" + fmt(T["P"]) + "\n};",
          "static const uint64_t ORC_S[507] = {\n" + fmt(T["S"]) + "\n};",
          "#endif", ""]
    open(os.path.join(ROOT, "oracle/poseidon_constants.h"), "w").write("\n".join(oc))
    print("wrote tables; KAT ok; S[506] =", hex(T["S"][506]))


if __name__ == "__main__":
    sys.exit(main())
