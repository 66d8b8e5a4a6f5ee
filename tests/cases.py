"""Inputs and checks shared by the emulator tests (CPU) and the GPU parity tests.

The generators restate the input loops of the reference's own tests
(src/goldilocks/tests/tests.cpp: Fibonacci vectors :1597-1610, :1747-1760; column-Fibonacci matrices
:1817-1835, :2260-2275) so the known answers in tests/golden/kat.json apply."""
import os

import numpy as np

P = 0xFFFFFFFF00000001


def fib_vector(n, a0=0, a1=1):
    v = [a0 % P, a1 % P]
    for _ in range(2, n):
        v.append((v[-1] + v[-2]) % P)
    return np.array(v[:n], dtype=np.uint64)


def fib_matrix(nrows, ncols):
    """rows 0 and 1 are 1..ncols, row j = row j-1 + row j-2 (tests.cpp:1821-1835, :2260-2275)."""
    m = np.zeros((nrows, ncols), dtype=object)
    for i in range(ncols):
        m[0, i] = i + 1
        if nrows > 1:
            m[1, i] = i + 1
    for j in range(2, nrows):
        m[j] = (m[j - 1] + m[j - 2]) % P
    return m.astype(np.uint64)


def edge_matrix(rng, nrows, ncols):
    """uniform words over the whole u64 range (so ~2^-32 of them are non-canonical) plus the boundary values"""
    x = rng.integers(0, 2 ** 64, size=(nrows, ncols), dtype=np.uint64)
    flat = x.reshape(-1)
    edges = np.array([P - 1, P, P + 1, 2 ** 64 - 1, 0, 1], dtype=np.uint64)
    k = min(flat.size, edges.size)
    flat[:k] = edges[:k]
    return x


def tree_words(nrows):
    return (2 * nrows - 1) * 4 if nrows else 0


def monomial_columns(n, degrees, w_n):
    """column c = evaluations of x^degrees[c] on the size-n subgroup: in[i,c] = w_n^(i*deg)"""
    out = np.zeros((n, len(degrees)), dtype=np.uint64)
    for c, d in enumerate(degrees):
        g = pow(w_n, d, P)
        v = 1
        for i in range(n):
            out[i, c] = v
            v = v * g % P
    return out


def check_equal(got, want, what=""):
    got = np.asarray(got).reshape(-1)
    want = np.asarray(want).reshape(-1)
    assert got.shape == want.shape, (what, got.shape, want.shape)
    if not np.array_equal(got, want):
        bad = np.nonzero(got != want)[0]
        raise AssertionError("%s: %d of %d words differ; first at %d: got %#x want %#x"
                             % (what, bad.size, got.size, bad[0], int(got[bad[0]]), int(want[bad[0]])))


# ---------------------------------------------------------------------------------------------
# parity checks against an oracle, for any Context (real library on the GPU, or the emulator build)
# ---------------------------------------------------------------------------------------------
def run_ntt(ctx, a, inverse=False, extend=False, device=False, inplace=False):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    n, c = a.shape
    if device:
        src = ctx.to_device(a)
        dst = src if inplace else ctx.device_alloc(max(a.nbytes, 8))
        ctx.ntt(dst, src, n, c, inverse, extend)
        out = ctx.to_host(dst, a.shape)
        ctx.device_free(src)
        if not inplace:
            ctx.device_free(dst)
        return out
    if inplace:
        buf = a.copy()
        ctx.ntt(None, buf, n, c, inverse, extend)
        return buf
    out = np.zeros_like(a)
    ctx.ntt(out, a, n, c, inverse, extend)
    return out


def check_ntt(ctx, oracle, a, **kw):
    for inv, ext in [(False, False), (True, False), (True, True)]:
        check_equal(run_ntt(ctx, a, inv, ext, **kw), oracle.ntt(a, inverse=inv, extend=ext),
                    "ntt %s inv=%d ext=%d %s" % (a.shape, inv, ext, kw))


def run_extend(ctx, a, n_ext, device=False, alias=False):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    n, c = a.shape
    if alias:                                   # output aliases input (tests.cpp:2578)
        buf = np.zeros((n_ext, c), dtype=np.uint64)
        buf.reshape(-1)[: n * c] = a.reshape(-1)
        if device:
            d = ctx.to_device(buf)
            ctx.extend_pol(d, d, n_ext, n, c)
            out = ctx.to_host(d, buf.shape)
            ctx.device_free(d)
            return out
        ctx.extend_pol(buf, buf, n_ext, n, c)
        return buf
    if device:
        src = ctx.to_device(a)
        dst = ctx.device_alloc(n_ext * c * 8)
        ctx.extend_pol(dst, src, n_ext, n, c)
        out = ctx.to_host(dst, (n_ext, c))
        ctx.device_free(src)
        ctx.device_free(dst)
        return out
    out = np.zeros((n_ext, c), dtype=np.uint64)
    ctx.extend_pol(out, a, n_ext, n, c)
    return out


def run_merkletree(ctx, a, nrows=None, ncols=None, device=False, dim=1):
    if a is not None:
        a = np.ascontiguousarray(a, dtype=np.uint64)
        nrows, width = a.shape
        ncols = width // dim
    tree = np.zeros(tree_words(nrows), dtype=np.uint64)
    if device and a is not None and a.size:
        src = ctx.to_device(a)
        dt = ctx.device_alloc(max(tree.nbytes, 8))
        ctx.merkletree(dt, src, ncols, nrows, dim)
        tree = ctx.to_host(dt, tree.shape)
        ctx.device_free(src)
        ctx.device_free(dt)
    else:
        ctx.merkletree(tree, a if (a is not None and a.size) else None, ncols, nrows, dim)
    return tree


def run_commit(ctx, a, n_ext, device=False):
    a = np.ascontiguousarray(a, dtype=np.uint64)
    n, c = a.shape
    if device:
        src = ctx.to_device(a)
        dext = ctx.device_alloc(n_ext * c * 8)
        dtree = ctx.device_alloc(tree_words(n_ext) * 8)
        root = ctx.commit(dext, dtree, src, n, n_ext, c)
        ext, tree = ctx.to_host(dext, (n_ext, c)), ctx.to_host(dtree, (tree_words(n_ext),))
        for p in (src, dext, dtree):
            ctx.device_free(p)
        return ext, tree, root
    ext = np.zeros((n_ext, c), dtype=np.uint64)
    tree = np.zeros(tree_words(n_ext), dtype=np.uint64)
    root = ctx.commit(ext, tree, a, n, n_ext, c)
    return ext, tree, root


def check_commit(ctx, oracle, a, n_ext, **kw):
    ext, tree, root = run_commit(ctx, a, n_ext, **kw)
    want_ext = oracle.extend_pol(a, n_ext)
    want_tree = oracle.merkletree(want_ext)
    check_equal(ext, want_ext, "commit ext %s" % (a.shape,))
    check_equal(tree, want_tree, "commit tree %s" % (a.shape,))
    check_equal(root, want_tree[-4:], "commit root")


NTT_MASK = 0x9E3779B97F4A7C15                      # xlayer_prover_b200/csrc/ntt_kernels.cuh


def wrap_twice_matrix(rng, n, ncols):
    """Inputs that make the butterflies of the INTT's first two stages wrap TWICE: with rows r, r + n/4, r + n/2, r + 3n/4
    of a column set to (p-1, p-1, 2^32-1, 2^32-1) stage 1 leaves 2^64-1 in two registers (no carry, legal but >= p) and
    stage 2 adds them (second carry); with (0, p-1, 0, 2^32-1) stage 2 subtracts 2^64-1 from 0 (second borrow).  The
    optimistic NTT passes (k_ntt_pass_opt) leave those wraps out, flag them, and the LDE is repeated exactly.  (The values
    are what the butterflies see; the matrix holds them divided by the input mask.)"""
    a = rng.integers(0, P, size=(n, ncols), dtype=np.uint64)
    q = n // 4
    inv = pow(NTT_MASK, P - 2, P)                   # the optimistic LDE multiplies its inputs by NTT_MASK first

    def pre(v):
        return (v * inv) % P
    a[1, 0], a[1 + q, 0], a[1 + 2 * q, 0], a[1 + 3 * q, 0] = pre(P - 1), pre(P - 1), pre(2 ** 32 - 1), pre(2 ** 32 - 1)
    c = ncols - 1
    a[q - 1, c], a[2 * q - 1, c], a[3 * q - 1, c], a[4 * q - 1, c] = 0, pre(P - 1), 0, pre(2 ** 32 - 1)
    return a


def check_optimistic_lde(make_ctx, oracle, rng, n=256, ncols=20):
    """device-resident extendPol / commit: exact results whether or not the optimistic passes had to be repeated.
    make_ctx(**env) -> a fresh context created under those environment variables; ncols > 16 (narrower matrices take the
    NARROW lane mapping, which has no optimistic flavour)"""
    n_ext = 2 * n
    honest = rng.integers(0, P, size=(n, ncols), dtype=np.uint64)
    crafted = wrap_twice_matrix(rng, n, ncols)
    ctx = make_ctx()
    try:
        r0 = ctx.ntt_retry_count()
        check_equal(run_extend(ctx, honest, n_ext, device=True), oracle.extend_pol(honest, n_ext), "optimistic LDE, honest data")
        assert ctx.ntt_retry_count() == r0, "honest data must not need the exact passes"
        small = rng.integers(0, 7, size=(n, ncols), dtype=np.uint64)             # a trace: 0, 1, small values and -1, -2
        small[rng.random((n, ncols)) < 0.1] = P - 1
        small[rng.random((n, ncols)) < 0.05] = P - 2
        check_equal(run_extend(ctx, small, n_ext, device=True), oracle.extend_pol(small, n_ext), "optimistic LDE, trace-like data")
        assert ctx.ntt_retry_count() == r0, "trace-like data must not need the exact passes (input mask)"
        check_equal(run_extend(ctx, crafted, n_ext, device=True), oracle.extend_pol(crafted, n_ext), "optimistic LDE, crafted data")
        assert ctx.ntt_retry_count() == r0 + 1, "crafted data must have been repeated with the exact passes"
        check_commit(ctx, oracle, crafted, n_ext, device=True)
        assert ctx.ntt_retry_count() == r0 + 2
        check_equal(run_extend(ctx, crafted, n_ext, device=True, alias=True), oracle.extend_pol(crafted, n_ext), "in place: exact passes directly")
        assert ctx.ntt_retry_count() == r0 + 2
    finally:
        ctx.close()
    ctx = make_ctx(GLP_NTT_FORCE_RETRY=1)
    try:
        check_equal(run_extend(ctx, honest, n_ext, device=True), oracle.extend_pol(honest, n_ext), "forced retry")
        assert ctx.ntt_retry_count() == 1
    finally:
        ctx.close()
    ctx = make_ctx(GLP_NTT_OPT=0)
    try:
        check_equal(run_extend(ctx, crafted, n_ext, device=True), oracle.extend_pol(crafted, n_ext), "optimistic passes off")
        assert ctx.ntt_retry_count() == 0
    finally:
        ctx.close()


# ---- the callers either side of the commit; the oracle is the reference build ----
def check_quotient_commit(ctx, reference, rng, shapes=((16, 2, 3, 2), (64, 4, 3, 2), (32, 8, 3, 4), (128, 2, 1, 2))):
    """Starks::genProof step 4 (starks.cpp:261-291): (n, blowup, qDim, qDeg) -> cm4 and every Merkle node."""
    for n, blow, qdim, qdeg in shapes:
        n_ext = n * blow
        q = edge_matrix(rng, n_ext, qdim)
        want_cm4, want_tree = reference.quotient_commit(q, n, qdeg)
        cm4 = np.zeros((n_ext, qdim * qdeg), dtype=np.uint64)
        tree = np.zeros(tree_words(n_ext), dtype=np.uint64)
        root = ctx.quotient_commit(cm4, tree, q, n, n_ext, qdim, qdeg)
        check_equal(cm4, want_cm4, "cm4 %s" % ((n, blow, qdim, qdeg),))
        check_equal(tree, want_tree, "quotient tree")
        check_equal(root, want_tree[-4:], "quotient root")
        # device pointers, root only
        dq = ctx.to_device(q)
        check_equal(ctx.quotient_commit(None, None, dq, n, n_ext, qdim, qdeg), want_tree[-4:], "quotient root (device q)")
        ctx.device_free(dq)


def check_const_tree(ctx, reference, rng, shapes=((16, 2, 5), (64, 4, 21), (8, 2, 1))):
    """bctree's *.consttree words and MerkleTreeGL(tree) mapping them back (build_const_tree.cpp:555-575, merkleTreeGL.hpp:24-32)."""
    from xlayer_prover_b200 import MerkleTreeGL, build_const_tree
    for n, blow, npols in shapes:
        a = edge_matrix(rng, n, npols)
        want = reference.const_tree(a, n * blow)
        got = build_const_tree(a, n.bit_length() - 1, (n * blow).bit_length() - 1, ctx=ctx)
        check_equal(got, want, "consttree %s" % ((n, blow, npols),))
        mt = MerkleTreeGL(tree=got, ctx=ctx)
        assert (mt.width, mt.height) == (npols, n * blow)
        root = np.zeros(4, dtype=np.uint64)
        mt.getRoot(root)
        check_equal(root, want[-4:], "constRoot")
    check_const_tree_files(ctx, reference, rng)


def check_const_tree_files(ctx, reference, rng, n=64, n_ext=128, npols=9):
    """The artefacts themselves: bctree's buildConstTree (build_const_tree.cpp:533-600, compiled from the reference) reads a
    constant-polynomials file and a starkStruct json and writes the *.consttree image and the verification key; the image
    glp_const_tree produces from the same polynomials must be byte-identical (after canonicalisation) and carry that root."""
    import json
    import tempfile
    from xlayer_prover_b200 import build_const_tree
    a = edge_matrix(rng, n, npols) % np.uint64(P)                       # what a .const file holds: canonical words, row-major
    with tempfile.TemporaryDirectory() as d:
        cf, sf, tf, vf = (os.path.join(d, x) for x in ("x.const", "x.starkstruct.json", "x.consttree", "x.verkey.json"))
        a.tofile(cf)
        with open(sf, "w") as f:
            json.dump({"nBits": n.bit_length() - 1, "nBitsExt": n_ext.bit_length() - 1, "nQueries": 8, "verificationHashType": "GL",
                       "steps": [{"nBits": n_ext.bit_length() - 1}, {"nBits": 3}]}, f)
        reference.bctree_build(cf, sf, tf, vf)
        image = reference.canonical(np.fromfile(tf, dtype=np.uint64))
        with open(vf) as f:
            verkey = json.load(f)
    got = build_const_tree(a, n.bit_length() - 1, n_ext.bit_length() - 1, ctx=ctx)
    check_equal(got, image, "*.consttree image written by the reference's bctree")
    check_equal(got[-4:], np.array(verkey["constRoot"], dtype=np.uint64), "constRoot of the verification key")


def check_transcript(ctx, reference, rng):
    """Transcript put/getField/getPermutations against the reference class on the same call sequence (transcript.cpp:4-91)."""
    from xlayer_prover_b200 import Transcript
    ours, ref = Transcript(ctx=ctx), reference.transcript()
    assert ours.getField() == ref.get_field()                      # squeeze from the empty state
    for size in (4, 8, 3, 13, 1):                                   # a Merkle root, a full block, ragged pieces
        a = rng.integers(0, P, size=size, dtype=np.uint64)
        ours.put(a)
        ref.put(a)
        assert ours.getField() == ref.get_field(), "getField after put(%d)" % size
    ours.put(np.array([P - 1, 0, 5], dtype=np.uint64))
    ref.put(np.array([P - 1, 0, 5], dtype=np.uint64))
    for _ in range(14):                                             # runs through all 12 outputs and re-permutes
        assert ours.getFields1() == ref.get_fields1()
    assert ours.getPermutations(64, 23) == ref.get_permutations(64, 23)   # nQueries x nBitsExt bits -> 24 field elements
    assert ours.getPermutations(3, 5) == ref.get_permutations(3, 5)


def check_fri_layers(ctx, reference, rng, pol_bits=10, steps=(7, 4, 2)):
    """FRIProve::prove's step loop (friProve.cpp:20-148) for si > 0: fold 2^pol_bits -> 2^steps[0] -> ... with the coset shift
    squared reduction_bits times per layer (:143-147), and every layer's regrouped Merkle tree (:110-123)."""
    pol = edge_matrix(rng, 1 << pol_bits, 3)
    shift_inv = pow(7, P - 2, P)
    bits = pol_bits
    for si, nb in enumerate(steps):
        x = rng.integers(0, P, size=3, dtype=np.uint64)
        want = reference.fri_fold(pol, bits, nb, x, shift_inv)
        got = np.zeros((1 << nb, 3), dtype=np.uint64)
        ctx.fri_fold(got, pol, bits, nb, x, shift_inv)
        check_equal(got, want, "fri fold 2^%d -> 2^%d" % (bits, nb))
        if si + 1 < len(steps):
            nxt = steps[si + 1]
            want_aux, want_tree = reference.fri_group_tree(want, nxt)
            aux = np.zeros_like(want_aux)
            tree = np.zeros(tree_words(1 << nxt), dtype=np.uint64)
            root = ctx.fri_group_tree(aux, tree, got, nb, nxt)
            check_equal(aux, want_aux, "fri layer source")
            check_equal(tree, want_tree, "fri layer tree")
            check_equal(root, want_tree[-4:], "fri layer root")
        for _ in range(bits - nb):
            shift_inv = shift_inv * shift_inv % P
        pol, bits = want, nb


def check_deferred_rows(ctx, oracle, rng, nrows=4096, device=False):
    """Rows that make a deferred fold of the leaf hash go wrong, so that the exact redo must run (gl_fold_defer_raw).
    On random data that is one row in 10^5, so drive it.  Form 2 of the fold (the S-boxes) computes t = lo64 + p2*EPS and
    then t - p3 with the borrow left out: a state word y = c * 2^48 (c < 2^16) squares to c^2 * 2^96, i.e. lo64 = 0,
    p2 = 0, p3 = c^2 > t = 0.  Form 1 (rank-1 updates, dot products) leaves out the borrow of lo64 - p3; its flag word is
    the high word of lo64, which the same rows drive to zero in the first S-box layer.  Rows carry such a word at every
    rate position and in several sponge blocks; the results must equal the oracle and every crafted row must be redone."""
    import pathlib
    import re
    src = (pathlib.Path(__file__).resolve().parents[1] / "xlayer_prover_b200" / "csrc" / "poseidon_tables.cuh").read_text()
    rc = [int(v, 16) for v in re.findall(r"0x[0-9a-fA-F]{16}", src.split("GLP_POSEIDON_RC_INIT", 1)[1])[:8]]
    P = 0xFFFFFFFF00000001
    ncols = 100
    a = edge_matrix(rng, nrows, ncols)
    crafted = []
    for k, row in enumerate(range(5, nrows, 97)):
        pos, blk = k % 8, (k // 8) % 12
        y = (0xFFFE - k) << 48                            # canonical, above every round constant, low 48 bits zero
        assert rc[pos] <= y < P
        a[row, 8 * blk + pos] = y - rc[pos]               # + RC[pos] = y exactly
        crafted.append(row)
    want = oracle.linear_hash_rows(a)
    before = ctx.leaf_redo_rows()
    got = np.zeros((nrows, 4), dtype=np.uint64)
    ctx.linear_hash_rows(got, a, ncols, ncols, nrows)
    check_equal(got, want, "linear_hash with rows that take the exact path")
    redone = ctx.leaf_redo_rows() - before
    assert redone >= len(crafted), (redone, len(crafted))
    # the same rows as leaves of a tree and through the sponge continuation (cap_in aliases out)
    tree = run_merkletree(ctx, a, device=device)
    check_equal(tree, oracle.merkletree(a), "merkletree with rows that take the exact path")
    cap = np.zeros((nrows, 4), dtype=np.uint64)
    ctx.linear_hash_absorb(cap, np.ascontiguousarray(a[:, :48]), 48, 48, nrows, first=True)
    ctx.linear_hash_absorb(cap, np.ascontiguousarray(a[:, 48:]), ncols - 48, ncols - 48, nrows, first=False)
    check_equal(cap, want, "absorb in two blocks with rows that take the exact path")
    assert ctx.leaf_redo_rows() - before >= 3 * len(crafted)
    # tree nodes: rows of <= 4 columns are copied into the leaves unhashed, so crafted words reach the level-0 pair hashes
    nb = min(nrows, 1 << 10)
    b = edge_matrix(rng, nb, 4)
    made = 0
    for k, row in enumerate(range(3, nb, 41)):
        pos = (k % 4) + 4 * (row & 1)                      # word of the 8-word pair this row's column k % 4 becomes
        b[row, k % 4] = ((0xFFFE - k) << 48) - rc[pos]
        made += 1
    before = ctx.leaf_redo_rows()
    check_equal(run_merkletree(ctx, b, device=device), oracle.merkletree(b), "merkletree whose level-0 nodes take the exact path")
    assert ctx.leaf_redo_rows() - before >= made


def check_multi(make_multi, oracle, rng, worlds=(2, 4), shapes=((64, 2, 40), (32, 4, 21), (128, 2, 660 // 8), (16, 2, 9), (8, 2, 3))):
    """glp_multi_* (several devices, one process; csrc/multi.cuh): extendPol, merkletree and the fused commit must give the
    single-device / oracle bits for every device count, including shapes too small to split (they run on one device)."""
    for G in worlds:
        m = make_multi(G)
        try:
            assert m.device_count() == G
            for n, blow, ncols in shapes:
                n_ext = n * blow
                a = edge_matrix(rng, n, ncols)
                want_ext = oracle.extend_pol(a, n_ext)
                want_tree = oracle.merkletree(want_ext)
                ext = np.zeros((n_ext, ncols), dtype=np.uint64)
                m.extend_pol(ext, a, n_ext, n, ncols)
                check_equal(ext, want_ext, "multi extend_pol G=%d %dx%d" % (G, n, ncols))
                tree = np.zeros(tree_words(n_ext), dtype=np.uint64)
                m.merkletree(tree, ext, ncols, n_ext)
                check_equal(tree, want_tree, "multi merkletree G=%d %dx%d" % (G, n_ext, ncols))
                ext2 = np.zeros((n_ext, ncols), dtype=np.uint64)
                tree2 = np.zeros(tree_words(n_ext), dtype=np.uint64)
                root = m.commit(ext2, tree2, a, n, n_ext, ncols)
                check_equal(ext2, want_ext, "multi commit ext G=%d %dx%d" % (G, n, ncols))
                check_equal(tree2, want_tree, "multi commit tree G=%d %dx%d" % (G, n, ncols))
                check_equal(root, want_tree[-4:], "multi commit root G=%d" % G)
                root_only = m.commit(None, None, a, n, n_ext, ncols)
                check_equal(root_only, want_tree[-4:], "multi commit, root only, G=%d" % G)
        finally:
            m.close()


def check_group_proofs(ctx, reference, rng, device=True, shapes=((256, 21), (1 << 12, 3), (64, 128), (2, 5), (1, 9))):
    """glp_merkle_group_proofs (all the proofs a tree owes the FRI query phase in one launch / one download) against the
    reference's own MerkleTreeGL::getGroupProof called once per index (merkleTreeGL.cpp:12-35, friProve.cpp:219-250)"""
    for h, w in shapes:
        src = edge_matrix(rng, h, w)
        nodes, root, psize, group_proof = reference.merkle_tree_gl(src)
        canon = reference.canonical(src)
        idx = np.array([0, h - 1, h // 2] + [int(x) for x in rng.integers(0, h, size=29)], dtype=np.uint64)
        want = np.stack([group_proof(int(i)) for i in idx])
        words = w + 4 * psize
        got = np.zeros((idx.size, words), dtype=np.uint64)
        ctx.merkle_group_proofs(got, src, nodes, w, h, idx)                     # host-resident tree: gathered where it is
        check_equal(got, want, "group proofs, host tree %dx%d" % (h, w))
        if device:
            dsrc, dnodes = ctx.to_device(src), ctx.to_device(nodes)
            got2 = np.zeros((idx.size, words), dtype=np.uint64)
            ctx.merkle_group_proofs(got2, dsrc, dnodes, w, h, idx)              # device-resident tree: one launch, one download
            check_equal(got2, want, "group proofs, device tree %dx%d" % (h, w))
            one = np.zeros(words, dtype=np.uint64)
            ctx.merkle_group_proof(one, dsrc, dnodes, w, h, int(idx[3]))
            check_equal(one, want[3], "single proof")
            ctx.device_free(dsrc)
            ctx.device_free(dnodes)
        del canon


def check_group_proofs_static(ctx, rng, shapes=((1 << 10, 21), (64, 128), (2, 5), (1, 9)), limit=4096, at_limit=True):
    """glp_merkle_group_proofs_static (index staging in a fixed file-scope device array, no allocation) against the existing
    glp_merkle_group_proofs on the same device-resident tree: two calls in a row per shape, the second with FEWER indices than
    the first so that the array still holds stale entries of the call before; then the limit and the argument checks."""
    from xlayer_prover_b200 import GlpError
    ctx.merkle_group_proofs_static(None, None, None, 0, 1, np.zeros(0, dtype=np.uint64))      # nidx == 0: set-up only
    for h, w in shapes:
        src = edge_matrix(rng, h, w)
        dsrc = ctx.to_device(src)
        dnodes = ctx.device_alloc(8 * tree_words(h))
        ctx.merkletree(dnodes, dsrc, w, h)
        depth = int(h).bit_length() - 1
        words = w + 4 * depth
        for nidx in (min(limit, 300), 7):
            idx = np.array([0, h - 1, h // 2] + [int(x) for x in rng.integers(0, h, size=nidx - 3)], dtype=np.uint64)
            want = np.zeros((nidx, words), dtype=np.uint64)
            ctx.merkle_group_proofs(want, dsrc, dnodes, w, h, idx)
            dgot = ctx.device_alloc(8 * nidx * words)
            ctx.memcpy(dgot, np.full(nidx * words, 0xDEADBEEF, dtype=np.uint64), 8 * nidx * words)
            ctx.merkle_group_proofs_static(dgot, dsrc, dnodes, w, h, idx)
            check_equal(ctx.to_host(dgot, (nidx, words)), want, "static group proofs %dx%d, %d indices" % (h, w, nidx))
            ctx.device_free(dgot)
        ctx.device_free(dsrc)
        ctx.device_free(dnodes)
    # the limit itself is accepted, one more is refused before anything is touched
    h, w = 1 << 8, 4
    src = edge_matrix(rng, h, w)
    dsrc = ctx.to_device(src)
    dnodes = ctx.device_alloc(8 * tree_words(h))
    ctx.merkletree(dnodes, dsrc, w, h)
    words = w + 4 * 8
    idx = rng.integers(0, h, size=limit + 1, dtype=np.uint64)
    dgot = ctx.device_alloc(8 * (limit + 1) * words)
    want = np.zeros((limit, words), dtype=np.uint64)
    nlim = limit if at_limit else 5                      # (the host-thread emulator takes a minute for 4096 CTAs)
    ctx.merkle_group_proofs(want[:nlim], dsrc, dnodes, w, h, idx[:nlim])
    ctx.merkle_group_proofs_static(dgot, dsrc, dnodes, w, h, idx[:nlim])
    check_equal(ctx.to_host(dgot, (nlim, words)), want[:nlim], "static group proofs at the limit")
    for bad in (lambda: ctx.merkle_group_proofs_static(dgot, dsrc, dnodes, w, h, idx),                       # limit + 1 indices
                lambda: ctx.merkle_group_proofs_static(np.zeros((4, words), dtype=np.uint64), dsrc, dnodes, w, h, idx[:4]),  # host output
                lambda: ctx.merkle_group_proofs_static(dgot, dsrc, dnodes, w, h, np.array([h], dtype=np.uint64))):         # index out of range
        try:
            bad()
        except GlpError as e:
            assert e.code == -1, e
        else:
            raise AssertionError("glp_merkle_group_proofs_static accepted a bad call")
    # a refused call leaves the library usable
    ctx.merkle_group_proofs_static(dgot, dsrc, dnodes, w, h, idx[:5])
    check_equal(ctx.to_host(dgot, (5, words)), want[:5], "static group proofs after refused calls")
    for p_ in (dgot, dsrc, dnodes):
        ctx.device_free(p_)


def check_fri_prove(ctx, reference, rng, pol_bits=9, steps=(9, 6, 3), nq=6, widths=(5, 3, 9, 6, 2)):
    """The reference's WHOLE FRIProve::prove (src/starkpil/fri/friProve.cpp:5-190, compiled from the reference with its own
    Polinomial, FRIProof, Transcript and MerkleTreeGL: oracle/_ref/libfriref.so) against the same protocol driven through this
    library with every polynomial and tree on the device: challenges from the transcript, glp_fri_fold per layer, the layer trees
    (glp_fri_group_tree), the query indices (getPermutations) and ALL query proofs of every tree in one glp_merkle_group_proofs
    call per tree.  Compared: every layer root, the last polynomial, and every word of every proof in the order queryPol emits
    them."""
    from xlayer_prover_b200 import Transcript
    n = 1 << pol_bits
    fri_pol = edge_matrix(rng, n, 3)
    srcs = [edge_matrix(rng, n, w) for w in widths]
    seed = rng.integers(0, P, size=11, dtype=np.uint64)
    want = reference.fri_prove(fri_pol, pol_bits, list(steps), nq, srcs, seed)
    # ---- ours ----
    tr = Transcript(ctx=ctx)
    tr.put(seed)
    dsrc = [ctx.to_device(t) for t in srcs]
    dtrees = []
    for t, d in zip(srcs, dsrc):
        dt = ctx.device_alloc(8 * tree_words(n))
        ctx.merkletree(dt, d, t.shape[1], n)
        dtrees.append(dt)
    dpol, bits, shift_inv = ctx.to_device(fri_pol), pol_bits, pow(7, P - 2, P)
    frees, layers, roots = [dpol] + dsrc + dtrees, [], []
    for si, nb in enumerate(steps):
        x = tr.getField()
        if si > 0 or nb != bits:
            dnew = ctx.device_alloc(8 * 3 << nb)
            ctx.fri_fold(dnew, dpol, bits, nb, x, shift_inv)
            frees.append(dnew)
            dpol = dnew
        if si + 1 < len(steps):
            nxt = steps[si + 1]
            daux, dlt = ctx.device_alloc(8 * 3 << nb), ctx.device_alloc(8 * tree_words(1 << nxt))
            root = ctx.fri_group_tree(daux, dlt, dpol, nb, nxt)
            tr.put(root)
            roots.append(root)
            layers.append((daux, dlt, 1 << nxt, 3 << (nb - nxt)))
            frees += [daux, dlt]
        else:
            final = ctx.to_host(dpol, (1 << nb, 3))
            tr.put(final.reshape(-1))
        for _ in range(bits - nb):
            shift_inv = shift_inv * shift_inv % P
        bits = nb
    check_equal(np.array(roots, dtype=np.uint64).reshape(-1, 4), want["roots"], "FRI layer roots against FRIProve::prove")
    check_equal(final, want["final"], "last FRI polynomial against FRIProve::prove")
    ys = np.array(tr.getPermutations(nq, steps[0]), dtype=np.uint64)
    out = []
    for si in range(len(steps)):
        if si == 0:
            per_tree = []
            for t, d, dt in zip(srcs, dsrc, dtrees):
                w = t.shape[1]
                pr = np.zeros((nq, w + 4 * pol_bits), dtype=np.uint64)
                ctx.merkle_group_proofs(pr, d, dt, w, n, ys)
                per_tree.append(pr)
            for i in range(nq):                                   # queryPol(treesGL[5], ys[i]): the five trees' proofs of one index together
                out += [pr[i] for pr in per_tree]
        else:
            daux, dlt, h, w = layers[si - 1]
            pr = np.zeros((nq, w + 4 * (h.bit_length() - 1)), dtype=np.uint64)
            ctx.merkle_group_proofs(pr, daux, dlt, w, h, ys)
            out += [pr[i] for i in range(nq)]
        if si + 1 < len(steps):
            ys = ys % np.uint64(1 << steps[si + 1])
    check_equal(np.concatenate(out), want["queries"], "every query proof against FRIProve::prove")
    for p in frees:
        ctx.device_free(p)
