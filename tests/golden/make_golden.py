#!/usr/bin/env python3
"""Regenerate tests/golden/kat.json and tests/golden/fixtures.npz.  Runs only in the build container
(it reads /root/reference); its outputs are committed so the GPU box needs neither.

kat.json     : the known-answer vectors the reference's own tests assert for this path, lifted from
               src/goldilocks/tests/tests.cpp (each entry records the TEST name and line range) and
               src/goldilocks/benchs/bench.cpp:440-443.
fixtures.npz : outputs of the compiled reference itself (oracle/_ref/libglref.so, built by oracle/Makefile
               from the unmodified sources) on small seeded inputs -- every extended element and every
               Merkle node, canonicalised with Goldilocks::toU64 like all of the reference's asserts.
"""
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("REF", "/root/reference")
TESTS = os.path.join(REF, "src/goldilocks/tests/tests.cpp")
BENCH = os.path.join(REF, "src/goldilocks/benchs/bench.cpp")
P = 0xFFFFFFFF00000001


def test_block(lines, name):
    """(first_line, last_line, text) of TEST(GOLDILOCKS_TEST, name)."""
    start = next(i for i, l in enumerate(lines) if re.match(r"TEST\(GOLDILOCKS_TEST,\s*%s\)" % name, l))
    end = next(i for i in range(start + 1, len(lines)) if lines[i].startswith("TEST(") or lines[i].startswith("#ifdef") or lines[i].startswith("int main"))
    return start + 1, end, "".join(lines[start:end])


def asserts(text, var):
    """values of ASSERT_EQ(Goldilocks::toU64(var[...]), 0x...) in order of appearance"""
    pat = r"ASSERT_EQ\(Goldilocks::toU64\(%s\[[^\]]*\]\),\s*(0[xX][0-9a-fA-F]+)\)" % re.escape(var)
    return [int(v, 16) for v in re.findall(pat, text)]


def main():
    lines = open(TESTS).readlines()
    kat = {"source": "okx/xlayer-prover src/goldilocks/tests/tests.cpp (values compared through Goldilocks::toU64)"}

    a, b, t = test_block(lines, "poseidon_full_seq")
    kat["poseidon_fibonacci"] = {"test": "poseidon_full_seq", "lines": [a, b], "out": asserts(t, "result")}
    kat["poseidon_zero"] = {"test": "poseidon_full_seq", "lines": [a, b], "out": asserts(t, "result0")}
    assert len(kat["poseidon_fibonacci"]["out"]) == 12 and len(kat["poseidon_zero"]["out"]) == 12

    a, b, t = test_block(lines, "linear_hash_seq")
    kat["linear_hash_fibonacci128"] = {"test": "linear_hash_seq", "lines": [a, b], "out": asserts(t, "result")}
    assert len(kat["linear_hash_fibonacci128"]["out"]) == 4

    a, b, t = test_block(lines, "merkletree_seq")
    r = asserts(t, "root")
    assert len(r) == 12
    kat["merkletree_64x128_root"] = {"test": "merkletree_seq", "lines": [a, b], "out": r[0:4]}
    kat["merkletree_64x0_root"] = {"test": "merkletree_seq", "lines": [a, b], "out": r[4:8]}
    kat["merkletree_131072x0_root"] = {"test": "merkletree_seq", "lines": [a, b], "out": r[8:12]}

    a, b, t = test_block(lines, "merkletree_batch_seq")
    r = asserts(t, "root")
    kat["merkletree_batch_64x128_b32_root"] = {"test": "merkletree_batch_seq", "lines": [a, b], "out": r[0:4]}

    a, b, t = test_block(lines, "LDE")
    kat["lde_16_to_32_shift49"] = {"test": "LDE", "lines": [a, b], "out": asserts(t, "a")}
    assert len(kat["lde_16_to_32_shift49"]["out"]) == 32
    a, b, t = test_block(lines, "LDE_block")
    kat["lde_block_16x8_to_32_col0"] = {"test": "LDE_block", "lines": [a, b], "out": asserts(t, "a")}
    a, b, t = test_block(lines, "extendePol")
    kat["extendpol_16x8_to_32_col0"] = {"test": "extendePol", "lines": [a, b], "out": asserts(t, "a")}
    assert len(kat["extendpol_16x8_to_32_col0"]["out"]) == 32

    btxt = open(BENCH).read()
    m = re.search(r"assert\(Goldilocks::toU64\(root\[0\]\) == (0[xX][0-9a-fA-F]+)\);\s*assert\(Goldilocks::toU64\(root\[1\]\) == (0[xX][0-9a-fA-F]+)\);\s*"
                  r"assert\(Goldilocks::toU64\(root\[2\]\) == (0[xX][0-9a-fA-F]+)\);\s*assert\(Goldilocks::toU64\(root\[3\]\) == (0[xX][0-9a-fA-F]+)\);", btxt)
    if m:
        kat["bench_merkletree_8388608x128_root"] = {"test": "benchs/bench.cpp MERKLETREE_BENCH", "out": [int(x, 16) for x in m.groups()]}
    with open(os.path.join(HERE, "kat.json"), "w") as f:
        json.dump(kat, f, indent=1)

    # ---- fixtures from the compiled reference
    from oracle.refbind import reference
    R = reference()
    assert R is not None, "build oracle/_ref first: make -C oracle ref"
    rng = np.random.default_rng(0x600D)
    fx = {}

    def rnd(*shape, edge=True):
        x = rng.integers(0, 2 ** 64, size=shape, dtype=np.uint64)
        if edge and x.size >= 6:                      # non-canonical and boundary words, as tests.cpp:44-65 does
            flat = x.reshape(-1)
            flat[:6] = np.array([P - 1, P, P + 1, 2 ** 64 - 1, 0, 1], dtype=np.uint64)
        return x
    for name, (n, c, blow) in {"ext_64x5_b4": (64, 5, 4), "ext_256x3_b2": (256, 3, 2), "ext_32x16_b8": (32, 16, 8), "ext_1x4_b2": (1, 4, 2)}.items():
        x = rnd(n, c)
        fx[name + "_in"] = x
        fx[name + "_out"] = R.extend_pol(x, n * blow)
    for name, (n, c) in {"ntt_128x3": (128, 3), "ntt_2x3": (2, 3), "ntt_1024x1": (1024, 1)}.items():
        x = rnd(n, c)
        fx[name + "_in"] = x
        fx[name + "_fwd"] = R.ntt(x)
        fx[name + "_inv"] = R.ntt(x, inverse=True)
        fx[name + "_invext"] = R.ntt(x, inverse=True, extend=True)
    for name, (n, c) in {"mt_32x13": (32, 13), "mt_16x4": (16, 4), "mt_16x3": (16, 3), "mt_8x8": (8, 8), "mt_4x660": (4, 660), "mt_1x9": (1, 9)}.items():
        x = rnd(n, c)
        fx[name + "_in"] = x
        fx[name + "_tree"] = R.merkletree(x)
    x = rnd(7, 12)
    fx["perm_in"] = x
    fx["perm_out"] = np.stack([R.hash_full_result(r) for r in x])
    # full commit with a proof
    x = rnd(64, 9)
    ext = R.extend_pol(x, 128)
    nodes, root, psize, gp = R.merkle_tree_gl(ext)
    fx["commit_64x9_in"], fx["commit_64x9_ext"], fx["commit_64x9_tree"] = x, ext, nodes
    fx["commit_64x9_proof_77"] = gp(77)
    np.savez_compressed(os.path.join(HERE, "fixtures.npz"), **fx)
    print("wrote kat.json (%d entries) and fixtures.npz (%d arrays)" % (len(kat) - 1, len(fx)))


if __name__ == "__main__":
    main()
