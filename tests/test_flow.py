"""The widened rows chained the way a prover chains them, with every large buffer resident on the device: stage commit ->
transcript -> step-4 quotient commit -> FRI layers (fold, regroup, tree, challenge from the transcript) -> query indices
from the transcript -> group proofs gathered from the device-resident trees.  Only roots, challenges, the last FRI
polynomial and the proofs cross to the host.  The oracle runs the same chain on the reference build (host arrays).
The "polynomials" handed from one step to the next are just prefixes of the previous step's output buffer -- the test is
about composition through the C ABI with device pointers, not about a sound proof."""
import os
import subprocess

import numpy as np
import pytest

from cases import P, check_equal, edge_matrix, tree_words

HERE = os.path.dirname(os.path.abspath(__file__))


def group_proof_np(source, nodes, idx):
    """MerkleTreeGL::getGroupProof on host arrays (merkleTreeGL.cpp:12-35): row idx, then the sibling at every level"""
    h, w = source.shape
    out, off, n, i = [source[idx] % np.uint64(P)], 0, h, idx
    while n > 1:
        out.append(nodes[off + (i ^ 1) * 4: off + (i ^ 1) * 4 + 4])
        off, n, i = off + 4 * n, n >> 1, i >> 1
    return np.concatenate(out)


def run_flow(ctx, ref, log_n, c1, steps_tail, nq, seed):
    from xlayer_prover_b200 import Transcript
    rng = np.random.default_rng(seed)
    n, n_ext = 1 << log_n, 2 << log_n
    steps = [log_n + 1] + list(steps_tail)
    a = edge_matrix(rng, n, c1)
    ours, theirs = Transcript(ctx=ctx), ref.transcript()
    # ---- oracle chain (host) ----
    ext = ref.extend_pol(a, n_ext)
    tree1 = ref.merkletree(ext)
    theirs.put(tree1[-4:])
    q = ext.reshape(-1)[: 3 * n_ext].reshape(n_ext, 3).copy()
    cm4, tree4 = ref.quotient_commit(q, n, 2)
    theirs.put(tree4[-4:])
    want = {"roots": [tree1[-4:], tree4[-4:]], "layers": []}
    pol, bits, shift_inv = cm4.reshape(-1)[: 3 * n_ext].reshape(n_ext, 3).copy(), steps[0], pow(7, P - 2, P)
    for si, nb in enumerate(steps):
        x = theirs.get_field()
        if si > 0:
            pol = ref.fri_fold(pol, bits, nb, x, shift_inv)
        if si + 1 < len(steps):
            aux, ltree = ref.fri_group_tree(pol, steps[si + 1])
            theirs.put(ltree[-4:])
            want["layers"].append((aux, ltree))
        else:
            theirs.put(pol.reshape(-1))
        for _ in range(bits - nb):
            shift_inv = shift_inv * shift_inv % P
        bits = nb
    want["final"] = pol
    ys = theirs.get_permutations(nq, steps[0])
    # ---- ours: device-resident ----
    da = ctx.to_device(a)
    dext, dtree1 = ctx.device_alloc(8 * n_ext * c1), ctx.device_alloc(8 * tree_words(n_ext))
    root1 = ctx.commit(dext, dtree1, da, n, n_ext, c1)
    ours.put(root1)
    dcm4, dtree4 = ctx.device_alloc(8 * n_ext * 6), ctx.device_alloc(8 * tree_words(n_ext))
    root4 = ctx.quotient_commit(dcm4, dtree4, dext, n, n_ext, 3, 2)          # q = the first 3*n_ext words of the extension
    ours.put(root4)
    check_equal(root1, want["roots"][0], "stage root")
    check_equal(root4, want["roots"][1], "quotient root")
    dpol, bits, shift_inv, layers, frees = dcm4, steps[0], pow(7, P - 2, P), [], [da, dext, dtree1, dcm4, dtree4]
    for si, nb in enumerate(steps):
        x = ours.getField()
        if si > 0:
            dnew = ctx.device_alloc(8 * 3 << nb)
            ctx.fri_fold(dnew, dpol, bits, nb, x, shift_inv)
            frees.append(dnew)
            dpol = dnew
        if si + 1 < len(steps):
            nxt = steps[si + 1]
            daux, dlt = ctx.device_alloc(8 * 3 << nb), ctx.device_alloc(8 * tree_words(1 << nxt))
            root = ctx.fri_group_tree(daux, dlt, dpol, nb, nxt)
            ours.put(root)
            check_equal(root, want["layers"][si][1][-4:], "FRI layer %d root" % (si + 1))
            layers.append((daux, dlt, 1 << nxt, 3 << (nb - nxt)))
            frees += [daux, dlt]
        else:
            final = ctx.to_host(dpol, (1 << nb, 3))
            check_equal(final, want["final"], "last FRI polynomial")
            ours.put(final.reshape(-1))
        for _ in range(bits - nb):
            shift_inv = shift_inv * shift_inv % P
        bits = nb
    assert ours.getPermutations(nq, steps[0]) == ys, "query indices"
    # ---- queries: friProve.cpp:151-178 ----
    idx = list(ys)
    for si in range(len(steps)):
        for k, y in enumerate(idx):
            if si == 0:
                for (src, nodes, w, hsrc, hnodes) in ((dext, dtree1, c1, ext, tree1), (dcm4, dtree4, 6, cm4, tree4)):
                    proof = np.zeros(w + 4 * steps[0], dtype=np.uint64)
                    ctx.merkle_group_proof(proof, src, nodes, w, n_ext, y)
                    check_equal(proof, group_proof_np(hsrc, hnodes, y), "stage proof %d" % y)
            else:
                daux, dlt, h, w = layers[si - 1]
                proof = np.zeros(w + 4 * (h.bit_length() - 1), dtype=np.uint64)
                ctx.merkle_group_proof(proof, daux, dlt, w, h, y)
                check_equal(proof, group_proof_np(*want["layers"][si - 1], y), "layer %d proof %d" % (si, y))
        if si + 1 < len(steps):
            idx = [y % (1 << steps[si + 1]) for y in idx]
    for p in frees:
        ctx.device_free(p)


def test_flow_emulator(reference):
    subprocess.run([os.path.join(HERE, "cuemu", "build_emu.sh")], check=True, capture_output=True)
    from xlayer_prover_b200._capi import Context, GlpLib
    c = Context(0, GlpLib(os.path.join(HERE, "cuemu", "libglprover_emu.so")))
    try:
        run_flow(c, reference, 5, 11, (4, 2), 4, 7)
    finally:
        c.close()


@pytest.mark.gpu
def test_flow_gpu(ctx, reference):
    run_flow(ctx, reference, 14, 40, (11, 7, 3), 16, 11)
    run_flow(ctx, reference, 10, 3, (6, 0), 8, 12)                 # three columns: q is the whole extension; folds by 2^5 and 2^6
