"""Parity tests proper: the CUDA path, called through the C ABI (ctypes -> libglprover.so) on a B200, against
the oracle (the compiled reference when oracle/_ref travelled with the snapshot, else the pinned C port),
the reference's own known answers (tests/golden/kat.json) and the reference-generated fixtures.

Bit-exact means equal after Goldilocks::toU64 canonicalisation, for every extended element and every Merkle
node: the library emits canonical words, the oracle side is canonicalised by refbind."""
import os

import numpy as np
import pytest

from cases import (P, check_commit, check_const_tree, check_equal, check_fri_layers, check_ntt, check_quotient_commit,
                   check_transcript,
                   edge_matrix, fib_matrix, fib_vector, monomial_columns, run_commit, run_extend, run_merkletree, run_ntt,
                   tree_words)

pytestmark = pytest.mark.gpu


def env_ctx(**env):
    from xlayer_prover_b200 import Context
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        return Context(0)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


# ---------------------------------------------------------------- field (PTX paths)
def test_field_primitives(ctx, rng):
    edges = np.array([0, 1, 2, P - 2, P - 1, P, P + 1, 2 ** 64 - 1, 2 ** 64 - 2, 2 ** 32 - 1, 2 ** 32, 2 ** 32 + 1,
                      2 ** 63, 0xFFFFFFFF, 0xFFFFFFFF00000000, 0xFFFFFFFEFFFFFFFF], dtype=np.uint64)
    a = np.concatenate([np.repeat(edges, edges.size), rng.integers(0, 2 ** 64, size=1 << 16, dtype=np.uint64)])
    b = np.concatenate([np.tile(edges, edges.size), rng.integers(0, 2 ** 64, size=1 << 16, dtype=np.uint64)])
    # shift amounts 0..95 all get exercised through b % 96
    b[edges.size ** 2: edges.size ** 2 + 96] = np.arange(96, dtype=np.uint64)
    got = ctx.field_selftest(a, b)
    ai = [int(x) for x in a]
    bi = [int(x) for x in b]
    want = np.array([[(x + y) % P, (x - y) % P, (x * y) % P, (x * pow(2, y % 96, P)) % P, x % P, x * x % P, pow(y, 7, P), (x + y) % P]
                     for x, y in zip(ai, bi)],
                    dtype=np.uint64)
    check_equal(got, want, "field selftest")


# ---------------------------------------------------------------- Poseidon
def test_poseidon_kats(ctx, golden, fixtures):
    out = np.zeros(12, dtype=np.uint64)
    ctx.hash_full_result(out, fib_vector(12), 1)
    check_equal(out, golden["poseidon_fibonacci"]["out"], "tests.cpp poseidon_full_seq")
    ctx.hash_full_result(out, np.zeros(12, dtype=np.uint64), 1)
    check_equal(out, golden["poseidon_zero"]["out"], "tests.cpp poseidon_full_seq zero")
    outs = np.zeros_like(fixtures["perm_in"])
    ctx.hash_full_result(outs, fixtures["perm_in"], outs.shape[0])
    check_equal(outs, fixtures["perm_out"])


def test_poseidon_random_vs_oracle(ctx, oracle, rng):
    x = edge_matrix(rng, 4096, 12)
    out = np.zeros_like(x)
    ctx.hash_full_result(out, x, x.shape[0])
    for i in rng.integers(0, 4096, size=64):
        check_equal(out[i], oracle.hash_full_result(x[i]), "perm %d" % i)
    check_equal(out[0], oracle.hash_full_result(x[0]), "perm with boundary words")


def test_linear_hash_kat_and_sizes(ctx, oracle, golden, rng):
    out = np.zeros(4, dtype=np.uint64)
    ctx.linear_hash_rows(out, fib_vector(128), 128, 128, 1)
    check_equal(out, golden["linear_hash_fibonacci128"]["out"], "tests.cpp linear_hash_seq")
    for size in list(range(1, 20)) + [24, 25, 100, 128, 660, 665, 751]:
        a = edge_matrix(rng, 37, size + 3)              # pitch > ncols
        got = np.zeros((37, 4), dtype=np.uint64)
        ctx.linear_hash_rows(got, a, size, size + 3, 37)
        check_equal(got, oracle.linear_hash_rows(np.ascontiguousarray(a[:, :size])), "linear_hash size %d" % size)


def test_linear_hash_deferred_wrap_rows_are_redone(ctx, oracle, rng):
    """The leaf kernel's fast path leaves the rare wrap of every fold uncorrected and, when it may have fired, hashes the
    row again with the exact fold (poseidon_kernels.cuh, POS_DEFER_BORROW): crafted rows, leaves and level-0 nodes."""
    from cases import check_deferred_rows
    check_deferred_rows(ctx, oracle, rng, device=True)


def test_optimistic_lde_is_repeated_exactly_when_flagged(oracle, rng):
    """k_ntt_pass_opt + dev_extend_pol_checked: crafted inputs fire the left-out wraps, the LDE is repeated with the exact
    passes; sizes with one, two and three tile passes"""
    from cases import check_optimistic_lde
    for n, ncols in ((128, 33), (1 << 13, 40), (1 << 15, 20), (1 << 17, 17)):
        check_optimistic_lde(env_ctx, oracle, rng, n=n, ncols=ncols)


def test_merkletree_kats(ctx, golden):
    check_equal(run_merkletree(ctx, fib_matrix(64, 128))[-4:], golden["merkletree_64x128_root"]["out"], "64x128")
    check_equal(run_merkletree(ctx, fib_matrix(64, 128), device=True)[-4:], golden["merkletree_64x128_root"]["out"], "64x128 dev")
    check_equal(run_merkletree(ctx, None, 64, 0)[-4:], golden["merkletree_64x0_root"]["out"], "64x0")
    check_equal(run_merkletree(ctx, None, 1 << 17, 0)[-4:], golden["merkletree_131072x0_root"]["out"], "2^17x0")


def test_merkletree_batch(ctx, oracle, golden, rng):
    from test_emu_kernels import check_merkletree_batch
    check_merkletree_batch(ctx, oracle, golden, rng)
    a = edge_matrix(rng, 1 << 12, 128)
    tree = np.zeros(tree_words(1 << 12), dtype=np.uint64)
    ctx.merkletree_batch(tree, a, 128, 1 << 12, 32)
    check_equal(tree, oracle.merkletree_batch(a, 32), "merkletree_batch 2^12 x 128, batch 32")


def test_merkletree_fixtures(ctx, fixtures):
    for k in ("mt_32x13", "mt_16x4", "mt_16x3", "mt_8x8", "mt_4x660", "mt_1x9"):
        for dev in (False, True):
            check_equal(run_merkletree(ctx, fixtures[k + "_in"], device=dev), fixtures[k + "_tree"], k)


@pytest.mark.parametrize("ncols", [1, 3, 4, 5, 8, 9, 12, 16, 100, 128, 660, 665, 751])
def test_merkletree_every_node(ctx, oracle, rng, ncols):
    for nrows in (1, 2, 8, 512):
        a = edge_matrix(rng, nrows, ncols)
        check_equal(run_merkletree(ctx, a, device=(nrows == 512)), oracle.merkletree(a), "tree %dx%d" % (nrows, ncols))


def test_merkletree_wide_tree_and_dim(ctx, oracle, rng):
    a = edge_matrix(rng, 1 << 14, 24)                   # exercises the per-level kernels and the single-CTA top
    check_equal(run_merkletree(ctx, a, device=True), oracle.merkletree(a), "2^14 x 24")
    a = edge_matrix(rng, 64, 36)
    check_equal(run_merkletree(ctx, a, dim=3), oracle.merkletree(a), "dim=3 (FRI trees, friProve.cpp:120)")


# ---------------------------------------------------------------- NTT
def test_ntt_reference_roundtrips(ctx):
    """tests.cpp ntt / ntt_block: NTT then INTT returns the input for every dst/src combination"""
    from xlayer_prover_b200 import NTT_Goldilocks
    g = NTT_Goldilocks(16, ctx=ctx)
    a = fib_vector(16, 1, 1).reshape(16, 1).copy()
    init = a.copy()
    for _ in range(5):
        g.NTT(a, a, 16)
        g.INTT(a, a, 16)
    check_equal(a, init, "ntt")
    a = fib_matrix(16, 8)
    init = a.copy()
    for _ in range(5):                                  # option 1: dst NULL
        g.NTT(None, a, 16, 8)
        g.INTT(None, a, 16, 8)
    check_equal(a, init, "dst=NULL")
    dst = np.zeros_like(a)
    for _ in range(5):                                  # option 3: dst != src
        g.NTT(dst, a, 16, 8)
        a[:] = 0
        g.INTT(a, dst, 16, 8)
    check_equal(a, init, "dst!=src")
    for _ in range(5):                                  # options 4/5: phases and blocks, also out of range
        g.NTT(a, a, 16, 8, None, 3, 5)
        g.INTT(a, a, 16, 8, None, 4, 3)
        g.NTT(a, a, 16, 8, None, 3, 3000)
        g.INTT(a, a, 16, 8, None, 4, -1)
    check_equal(a, init, "phases/blocks")
    a1 = np.array([[1, 2, 3]], dtype=np.uint64)         # size 1
    b1 = np.zeros_like(a1)
    g.NTT(b1, a1, 1, 3)
    check_equal(b1, a1)
    g.INTT(a1, b1, 1, 3)
    check_equal(a1, [1, 2, 3])
    a2 = np.array([[1, 2, 3], [4, 5, 6]], dtype=np.uint64)   # size 2
    b2 = np.zeros_like(a2)
    g.NTT(b2, a2, 2, 3)
    g.INTT(a2, b2, 2, 3)
    check_equal(a2, [1, 2, 3, 4, 5, 6])
    g.NTT(b2, a2, 0, 3)                                 # size 0 / ncols 0 do not crash
    g.INTT(a2, b2, 1, 0)
    assert NTT_Goldilocks(0).s == 0                     # inert object (starks.hpp:81-82 when !generateProof)


@pytest.mark.parametrize("logn", list(range(0, 15)))
def test_ntt_vs_oracle_all_sizes(ctx, oracle, rng, logn):
    for c in (1, 3, 8):
        check_ntt(ctx, oracle, edge_matrix(rng, 1 << logn, c))
    check_ntt(ctx, oracle, edge_matrix(rng, 1 << logn, 5), device=True, inplace=True)


def test_ntt_fixtures(ctx, fixtures):
    for k in ("ntt_128x3", "ntt_2x3", "ntt_1024x1"):
        a = fixtures[k + "_in"]
        check_equal(run_ntt(ctx, a), fixtures[k + "_fwd"], k)
        check_equal(run_ntt(ctx, a, True), fixtures[k + "_inv"], k)
        check_equal(run_ntt(ctx, a, True, True), fixtures[k + "_invext"], k)


def test_ntt_two_pass_large(ctx, oracle, rng):
    for n, c in [(1 << 16, 3), (1 << 18, 2), (1 << 20, 1)]:
        a = edge_matrix(rng, n, c)
        check_equal(run_ntt(ctx, a, device=True), oracle.ntt(a), "ntt %dx%d" % (n, c))
        check_equal(run_ntt(ctx, a, True, True, device=True), oracle.ntt(a, inverse=True, extend=True), "intt-ext %dx%d" % (n, c))


def test_ntt_three_pass_and_cb2(ctx, oracle, rng):
    """the 3-pass plan (sizes above 2^24 in production) reached at small sizes by capping the tile; and the
    2-column tile variant"""
    for env in ({"GLP_NTT_MAX_LOGQ": 5}, {"GLP_NTT_MAX_LOGQ": 4}, {"GLP_NTT_NO_PASS_TABLES": 1}, {"GLP_NTT_NO_FUSED_LDE": 1}, {"GLP_NTT_NO_PASS_TABLES": 1, "GLP_NTT_MAX_LOGQ": 4}):
        c = env_ctx(**env)
        try:
            for n, w in [(1 << 13, 3), (1 << 11, 8), (1 << 6, 1)]:
                check_ntt(c, oracle, edge_matrix(rng, n, w))
            a = edge_matrix(rng, 1 << 10, 5)
            check_equal(run_extend(c, a, 1 << 12), oracle.extend_pol(a, 1 << 12), "extend %s" % env)
        finally:
            c.close()


def test_root_and_reverse_permutation(ctx, oracle, rng):
    from xlayer_prover_b200 import NTT_Goldilocks
    g = NTT_Goldilocks(1 << 10, ctx=ctx)
    for k in range(0, 33):
        assert g.root(k, 1) == oracle.w(k)
    a = edge_matrix(rng, 64, 11)
    dst = np.zeros((64, 4), dtype=np.uint64)
    g.reversePermutation(dst, a, 64, 5, 4, 11)
    br = [int(format(i, "06b")[::-1], 2) for i in range(64)]
    check_equal(dst, a[br, 5:9], "reversePermutation")
    g2 = NTT_Goldilocks(64, extension=2, ctx=ctx)
    g2.reversePermutation(dst, a, 64, 5, 4, 11)
    want = a[br, 5:9].copy()
    want[[i for i in range(64) if br[i] >= 32]] = 0
    check_equal(dst, want, "reversePermutation with zero fill")
    sq = a.copy()
    g.reversePermutation(sq, sq, 64, 0, 11, 11)
    check_equal(sq, a[br], "reversePermutation in place")


# ---------------------------------------------------------------- extendPol / LDE
def test_extendpol_kat(ctx, golden):
    from xlayer_prover_b200 import NTT_Goldilocks
    ntt = NTT_Goldilocks(16, ctx=ctx)
    a = np.zeros((32, 8), dtype=np.uint64)
    a.reshape(-1)[:128] = fib_matrix(16, 8).reshape(-1)
    ntt.extendPol(a, a, 32, 16, 8, np.zeros((32, 8), dtype=np.uint64))     # output aliases input, tests.cpp:2578
    check_equal(a[:, 0], golden["extendpol_16x8_to_32_col0"]["out"], "tests.cpp extendePol")
    check_equal(a[:, 0], golden["lde_block_16x8_to_32_col0"]["out"], "tests.cpp LDE_block")


def test_lde_shift49_kat(ctx, golden):
    from xlayer_prover_b200 import NTT_Goldilocks
    a = fib_vector(16, 1, 1).reshape(16, 1).copy()
    NTT_Goldilocks(16, ctx=ctx).INTT(a, a, 16)
    padded = np.zeros((32, 1), dtype=np.uint64)
    r = 1
    for i in range(16):
        padded[i, 0] = int(a[i, 0]) * r % P
        r = r * 49 % P
    NTT_Goldilocks(32, ctx=ctx).NTT(padded, padded, 32)
    check_equal(padded, golden["lde_16_to_32_shift49"]["out"], "tests.cpp LDE")


def test_extendpol_fixtures(ctx, fixtures):
    for k in ("ext_64x5_b4", "ext_256x3_b2", "ext_32x16_b8", "ext_1x4_b2"):
        for dev in (False, True):
            check_equal(run_extend(ctx, fixtures[k + "_in"], fixtures[k + "_out"].shape[0], device=dev), fixtures[k + "_out"], k)


@pytest.mark.parametrize("n,c,blow", [(1, 3, 2), (2, 5, 4), (16, 1, 2), (64, 9, 8), (256, 12, 16), (1 << 10, 660, 2),
                                      (1 << 12, 17, 4), (1 << 13, 8, 2), (1 << 14, 5, 8), (1 << 17, 18, 8)])
def test_extendpol_vs_oracle(ctx, oracle, port, rng, n, c, blow):
    a = edge_matrix(rng, n, c)
    o = port if n * blow == 4 else oracle            # the reference asserts at N_ext = 4 (ntt_goldilocks.cpp:302)
    want = o.extend_pol(a, n * blow)
    check_equal(run_extend(ctx, a, n * blow), want, "host")
    check_equal(run_extend(ctx, a, n * blow, device=True), want, "device")
    check_equal(run_extend(ctx, a, n * blow, alias=True, device=(n >= 1024)), want, "aliased")


def test_config_a_full_parity(ctx, oracle):
    """BASELINE.json configs[0]: 2^16 rows x 16 cols, blowup 2 -- every extended element and every node."""
    from xlayer_prover_b200 import synth_matrix
    a = synth_matrix(1 << 16, 16, 0x9E3779B97F4A7C15 + 1)
    check_commit(ctx, oracle, a, 1 << 17)
    check_commit(ctx, oracle, a, 1 << 17, device=True)


def test_commit_fixture_and_proofs(ctx, fixtures):
    from xlayer_prover_b200 import MerkleTreeGL, NTT_Goldilocks
    a = fixtures["commit_64x9_in"]
    ext = np.zeros((128, 9), dtype=np.uint64)
    NTT_Goldilocks(64, ctx=ctx).extendPol(ext, a, 128, 64, 9)
    check_equal(ext, fixtures["commit_64x9_ext"])
    mt = MerkleTreeGL(128, 9, ext, ctx=ctx)
    mt.merkelize()
    check_equal(mt.nodes, fixtures["commit_64x9_tree"])
    check_equal(mt.getRoot(), fixtures["commit_64x9_tree"][-4:])
    assert mt.MerkleProofSize() == 7 and mt.getTreeNumElements() == tree_words(128)
    proof = np.zeros(9 + 4 * 7, dtype=np.uint64)
    check_equal(mt.getGroupProof(proof, 77), fixtures["commit_64x9_proof_77"])
    mt2 = MerkleTreeGL(128, 9, None, ctx=ctx)           # source NULL -> allocated, then copySource (friProve.cpp:120-122)
    mt2.copySource(ext)
    mt2.merkelize()
    check_equal(mt2.nodes, fixtures["commit_64x9_tree"])


def test_merkletreegl_vs_reference_class(ctx, reference, rng):
    from xlayer_prover_b200 import MerkleTreeGL
    src = edge_matrix(rng, 256, 21)
    nodes, root, psize, gp = reference.merkle_tree_gl(src)
    mt = MerkleTreeGL(256, 21, src, ctx=ctx)
    mt.merkelize()
    check_equal(mt.nodes, nodes)
    check_equal(mt.getRoot(), root)
    assert mt.MerkleProofSize() == psize
    for idx in (0, 1, 100, 255):
        check_equal(mt.getGroupProof(np.zeros(21 + 4 * psize, dtype=np.uint64), idx), gp(idx), "proof %d" % idx)


def test_extend_pol_scatter(ctx, oracle, rng):
    """row-sharded LDE output (the fused multi-GPU exchange) with all slabs on this GPU"""
    from test_emu_kernels import check_scatter
    check_scatter(ctx, oracle, rng, shapes=((1 << 10, 2, 37), (1 << 12, 4, 9), (1 << 9, 8, 70), (64, 2, 21), (1, 2, 3)), use_ipc=False)


def test_linear_hash_absorb(ctx, oracle, rng):
    from test_emu_kernels import check_absorb
    check_absorb(ctx, oracle, rng)


def test_commit_streamed_default_blocks(ctx, oracle, rng):
    """default pipeline blocks at 300 columns: 32 (short first block) + 96 + 88 + 84; extension + tree + root, root only,
    and the extension written over the input"""
    a = edge_matrix(rng, 1 << 11, 300)
    check_commit(ctx, oracle, a, 1 << 12)
    want = oracle.merkletree(oracle.extend_pol(a, 1 << 12))
    check_equal(ctx.commit(None, None, a, 1 << 11, 1 << 12, 300), want[-4:], "root only")
    buf = np.zeros((1 << 12, 300), dtype=np.uint64)
    buf.reshape(-1)[: a.size] = a.reshape(-1)
    ctx.commit(buf, None, buf, 1 << 11, 1 << 12, 300)
    check_equal(buf, oracle.extend_pol(a, 1 << 12), "aliased host extension")


def test_commit_streamed_host_path(ctx, oracle, rng):
    a = edge_matrix(rng, 1 << 12, 37)
    ctx.set_scratch_limit(((1 << 13) + (1 << 12)) * 8 * 8 * 2)      # 8-column blocks -> 5 streamed blocks
    try:
        check_commit(ctx, oracle, a, 1 << 13)
        check_commit(ctx, oracle, a, 1 << 13, device=True)
    finally:
        ctx.set_scratch_limit(0)
    # pinned host buffers (what bench.py's e2e leg uses)
    n, c, ne = 1 << 12, 37, 1 << 13
    hin = ctx.host_alloc(n * c * 8)
    hext = ctx.host_alloc(ne * c * 8)
    ctx.memcpy(hin, a, a.nbytes)
    tree = np.zeros(tree_words(ne), dtype=np.uint64)
    root = ctx.commit(hext, tree, hin, n, ne, c)
    ext = np.zeros((ne, c), dtype=np.uint64)
    ctx.memcpy(ext, hext, ext.nbytes)
    want = oracle.extend_pol(a, ne)
    check_equal(ext, want, "pinned ext")
    check_equal(tree, oracle.merkletree(want), "pinned tree")
    check_equal(root, tree[-4:])
    ctx.host_free(hin)
    ctx.host_free(hext)


def test_argument_errors(ctx):
    from xlayer_prover_b200 import GLP_E_ARG, GLP_E_DOMAIN, GlpError, NTT_Goldilocks
    a = np.zeros((12, 2), dtype=np.uint64)
    with pytest.raises(GlpError) as e:
        ctx.ntt(a, a, 12, 2)
    assert e.value.code == GLP_E_ARG
    with pytest.raises(GlpError) as e:
        ctx.ntt(a, a, 1 << 33, 1)
    assert e.value.code == GLP_E_DOMAIN
    with pytest.raises(ValueError):
        NTT_Goldilocks(1 << 33)                         # std::range_error in the reference
    with pytest.raises(GlpError):
        ctx.merkletree(np.zeros(100, dtype=np.uint64), a, 2, 12)
    ctx.merkletree(a, a, 2, 0)                          # num_rows == 0 returns (poseidon_goldilocks.cpp:268-271)


# ---------------------------------------------------------------- size-independent properties at large sizes
def test_large_roundtrip_and_linearity(ctx):
    """2^22 x 16 on the device: INTT(NTT(x)) == x (checksum of every word), and LDE(a+b) == LDE(a)+LDE(b)."""
    n, c = 1 << 22, 16
    words = n * c
    x = ctx.device_alloc(words * 8)
    y = ctx.device_alloc(words * 8)
    ctx.synth_fill(x, n, c, 77)
    before = ctx.checksum(x, words)
    ctx.ntt(y, x, n, c)
    assert ctx.checksum(y, words) != before
    ctx.ntt(y, y, n, c, inverse=True)
    assert ctx.checksum(y, words) == before
    ctx.device_free(y)
    ctx.device_free(x)
    n, c, ne = 1 << 20, 8, 1 << 21
    bufs = [ctx.device_alloc(n * c * 8) for _ in range(2)]
    outs = [ctx.device_alloc(ne * c * 8) for _ in range(2)]
    ctx.synth_fill(bufs[0], n, c, 1)
    ctx.synth_fill(bufs[1], n, c, 2)
    sums = []
    for i in range(2):
        ctx.extend_pol(outs[i], bufs[i], ne, n, c)
        sums.append(ctx.checksum(outs[i], ne * c))
    a = ctx.to_host(bufs[0], (n, c)).astype(object) + ctx.to_host(bufs[1], (n, c)).astype(object)
    ctx.memcpy(bufs[0], (a % P).astype(np.uint64), n * c * 8)
    ctx.extend_pol(outs[0], bufs[0], ne, n, c)
    s = ctx.checksum(outs[0], ne * c)
    assert s == ((sums[0][0] + sums[1][0]) % P, (sums[0][1] + sums[1][1]) % P)
    for p in bufs + outs:
        ctx.device_free(p)


def test_large_lde_of_monomials(ctx, oracle):
    """Closed form at 2^20 -> 2^21: a column holding x^d on the subgroup extends to (shift*w_ext^i)^d."""
    n, ne = 1 << 20, 1 << 21
    degrees = [0, 1, 2, 12345, n - 1]
    a = monomial_columns(n, degrees, oracle.w(20))
    ext = run_extend(ctx, a, ne, device=True)
    wext = oracle.w(21)
    rows = [0, 1, 2, 3, ne // 2, ne - 1] + [int(r) for r in np.random.default_rng(3).integers(0, ne, size=200)]
    for i in rows:
        pt = 7 * pow(wext, i, P) % P
        for c, d in enumerate(degrees):
            assert int(ext[i, c]) == pow(pt, d, P), (i, d)


def test_large_merkle_paths(ctx, oracle):
    """2^20 leaves x 64 columns: recompute sampled leaves with the oracle and walk their paths to the root."""
    n, c = 1 << 20, 64
    src = ctx.device_alloc(n * c * 8)
    tree = ctx.device_alloc(tree_words(n) * 8)
    ctx.synth_fill(src, n, c, 99)
    ctx.merkletree(tree, src, c, n)
    nodes = ctx.to_host(tree, (tree_words(n),))
    from xlayer_prover_b200 import synth_matrix
    for idx in [0, 1, n - 1] + [int(r) for r in np.random.default_rng(4).integers(0, n, size=20)]:
        row = synth_matrix(1, c, 99, row0=idx, ncols_total=c)[0]
        proof = np.zeros(c + 4 * 20, dtype=np.uint64)
        ctx.merkle_group_proof(proof, src, tree, c, n, idx)
        check_equal(proof[:c], row)
        h = oracle.linear_hash(row)
        check_equal(nodes[4 * idx: 4 * idx + 4], h, "leaf %d" % idx)
        i = idx
        for lvl in range(20):
            sib = proof[c + 4 * lvl: c + 4 * lvl + 4]
            pair = np.concatenate([h, sib] if i % 2 == 0 else [sib, h] + [])
            h = oracle.hash_full_result(np.concatenate([pair, np.zeros(4, dtype=np.uint64)]))[:4]
            i //= 2
        check_equal(h, nodes[-4:], "root from leaf %d" % idx)
    ctx.device_free(src)
    ctx.device_free(tree)


# ---------------------------------------------------------------- full-size secondary shapes (BASELINE.json configs[2], FRI/cm4 widths)
def _check_full_size_commit(ctx, oracle, logn, ncols, logb, seed, cols_to_check):
    """Device-resident commit at full size; the oracle re-derives whole columns of the extension (it is column-wise
    independent) and sampled Merkle paths (leaf = linear_hash of the row, then the sibling path up to the root)."""
    from xlayer_prover_b200 import synth_matrix
    n, ne = 1 << logn, 1 << (logn + logb)
    src = ctx.device_alloc(n * ncols * 8)
    ext = ctx.device_alloc(ne * ncols * 8)
    tree = ctx.device_alloc(tree_words(ne) * 8)
    ctx.synth_fill(src, n, ncols, seed)
    root = ctx.commit(ext, tree, src, n, ne, ncols)
    e = ctx.to_host(ext, (ne, ncols))
    nodes = ctx.to_host(tree, (tree_words(ne),))
    for c in cols_to_check:
        a = synth_matrix(n, 1, seed, col0=c, ncols_total=ncols)
        check_equal(e[:, c], oracle.extend_pol(a, ne)[:, 0], "2^%d x %d blowup %d, column %d" % (logn, ncols, 1 << logb, c))
    depth = logn + logb
    for idx in [0, ne - 1] + [int(r) for r in np.random.default_rng(5).integers(0, ne, size=10)]:
        h = oracle.linear_hash(e[idx])
        check_equal(nodes[4 * idx: 4 * idx + 4], h, "leaf %d" % idx)
        off, nn, i = 0, ne, idx
        for _ in range(depth):
            sib = nodes[off + 4 * (i ^ 1): off + 4 * (i ^ 1) + 4]
            pair = np.concatenate([h, sib]) if i % 2 == 0 else np.concatenate([sib, h])
            h = oracle.hash_full_result(np.concatenate([pair, np.zeros(4, dtype=np.uint64)]))[:4]
            off += 4 * nn
            nn //= 2
            i //= 2
        check_equal(h, root, "path of leaf %d" % idx)
    check_equal(nodes[-4:], root)
    for p in (src, ext, tree):
        ctx.device_free(p)


def test_config_c_full_size(ctx, oracle):
    """BASELINE.json configs[2]: 2^20 rows x 200 columns, blowup 8 (nBitsExt = 23)"""
    _check_full_size_commit(ctx, oracle, 20, 200, 3, 0x9E3779B97F4A7C15 + 3, [0, 77, 199])


@pytest.mark.parametrize("logn,ncols,logb", [(20, 1, 1), (20, 3, 2), (18, 6, 1), (17, 18, 3), (16, 12, 4)])
def test_narrow_shapes_full_size(ctx, oracle, logn, ncols, logb):
    """the narrow-matrix lane mapping (<= 16 columns: 32/cw tiles per warp) and the recursive-prover widths"""
    _check_full_size_commit(ctx, oracle, logn, ncols, logb, 0xFACE + ncols, sorted(set([0, ncols - 1])))


def test_narrow_equals_wide_mapping(ctx, rng):
    """same transforms with GLP_NTT_NO_NARROW=1 (lane = column) must give identical bits"""
    wide = env_ctx(GLP_NTT_NO_NARROW=1)
    try:
        for n, c in [(1 << 12, 1), (1 << 16, 3), (1 << 9, 16), (1 << 17, 5)]:
            a = edge_matrix(rng, n, c)
            for inv, ext in [(False, False), (True, False), (True, True)]:
                check_equal(run_ntt(ctx, a, inv, ext), run_ntt(wide, a, inv, ext), "narrow vs wide %s inv=%d ext=%d" % ((n, c), inv, ext))
            check_equal(run_extend(ctx, a, 4 * n), run_extend(wide, a, 4 * n), "narrow vs wide extend %s" % ((n, c),))
    finally:
        wide.close()


@pytest.mark.parametrize("ncols", [37, 48, 56, 60, 80, 88, 165])
def test_remainder_column_group_split(ctx, oracle, rng, ncols):
    """32 k + r columns (a rank's slice of 660 columns on 4 / 8 devices: 165 / 80-88): the remainder columns of every NTT pass run
    as their own launch on the narrow mapping.  Against the oracle (NTT, INTT, extension, commit) and against the one-launch
    layout (GLP_NTT_NO_SPLIT=1), at three-pass sizes."""
    a = edge_matrix(rng, 1 << 13, ncols)
    check_ntt(ctx, oracle, a)
    check_equal(run_extend(ctx, a, 1 << 15, device=True), oracle.extend_pol(a, 1 << 15), "extend %d columns" % ncols)
    # guard bands around the output (compute-sanitizer is not available on the pool): the pointer-walking stores of every pass and
    # launch must stay inside [out, out + n_ext * ncols)
    n, n_ext, guard = 1 << 13, 1 << 15, 1 << 12
    src = ctx.to_device(a)
    buf = ctx.device_alloc(8 * (n_ext * ncols + 2 * guard))
    ctx.memcpy(buf, np.full(n_ext * ncols + 2 * guard, 0xA5A5A5A5A5A5A5A5, dtype=np.uint64), 8 * (n_ext * ncols + 2 * guard))
    ctx.extend_pol(buf + 8 * guard, src, n_ext, n, ncols)
    got = ctx.to_host(buf, (n_ext * ncols + 2 * guard,))
    assert (got[:guard] == 0xA5A5A5A5A5A5A5A5).all() and (got[-guard:] == 0xA5A5A5A5A5A5A5A5).all(), "extendPol wrote outside its output"
    check_equal(got[guard:-guard].reshape(n_ext, ncols), oracle.extend_pol(a, n_ext), "extend into a guarded buffer")
    ctx.device_free(src)
    ctx.device_free(buf)
    b = edge_matrix(rng, 1 << 17, ncols)
    whole = env_ctx(GLP_NTT_NO_SPLIT=1)
    try:
        check_equal(run_extend(ctx, b, 1 << 18, device=True), run_extend(whole, b, 1 << 18, device=True), "split vs one launch, %d columns" % ncols)
        got = run_commit(ctx, b, 1 << 18, device=True)
        want = run_commit(whole, b, 1 << 18, device=True)
        for g, w in zip(got, want):
            check_equal(g, w, "commit, split vs one launch, %d columns" % ncols)
    finally:
        whole.close()


# ---------------------------------------------------------------- the callers either side of the commit
def test_quotient_commit(ctx, reference, rng):
    """Starks::genProof step 4 (starks.cpp:261-291) at test sizes and at a zkEVM-like shape (qDim 3, qDeg 2, blowup 2)"""
    check_quotient_commit(ctx, reference, rng)
    check_quotient_commit(ctx, reference, rng, shapes=((1 << 16, 2, 3, 2), (1 << 13, 8, 3, 4)))


def test_const_tree_and_transcript(ctx, reference, rng):
    check_const_tree(ctx, reference, rng)
    check_const_tree(ctx, reference, rng, shapes=((1 << 12, 2, 44),))
    check_transcript(ctx, reference, rng)


def test_fri_layers(ctx, reference, rng):
    """FRI folding + layer trees at test sizes and at a recursive-prover-like schedule (2^20 -> 2^15 -> 2^10 -> 2^5)"""
    check_fri_layers(ctx, reference, rng)
    check_fri_layers(ctx, reference, rng, pol_bits=7, steps=(1, 0))
    check_fri_layers(ctx, reference, rng, pol_bits=20, steps=(15, 10, 5))


def test_merkletree_streamed_host_source(oracle, rng):
    """MerkleTreeGL::merkelize on a host-resident source: row blocks uploaded under the leaf hash"""
    c = env_ctx(GLP_MERKLE_STREAM_MIN=1 << 20)
    try:
        for n, w in ((1 << 15, 100), (1 << 13, 660), (1 << 16, 5)):
            a = edge_matrix(rng, n, w)
            check_equal(run_merkletree(c, a), oracle.canonical(oracle.merkletree(a)), "streamed merkletree %s" % ((n, w),))
    finally:
        c.close()


def test_host_register(ctx, oracle, rng):
    """glp_host_register: the caller's own (page-aligned here) buffers pinned in place, commit through them, unpinned"""
    n, c = 1 << 12, 24
    a = edge_matrix(rng, n, c)
    ext, tree = np.zeros((2 * n, c), dtype=np.uint64), np.zeros(tree_words(2 * n), dtype=np.uint64)
    for buf in (a, ext, tree):
        ctx.host_register(buf)
    try:
        root = ctx.commit(ext, tree, a, n, 2 * n, c)
    finally:
        for buf in (a, ext, tree):
            ctx.host_unregister(buf)
    want_ext = oracle.extend_pol(a, 2 * n)
    want_tree = oracle.merkletree(want_ext)
    check_equal(ext, want_ext, "registered ext")
    check_equal(tree, want_tree, "registered tree")
    check_equal(root, want_tree[-4:], "registered root")


def test_overlapped_commit_knob(oracle, rng):
    """GLP_CO=1: LDE of column block k+1 (k_ntt_pass_co) under the leaf hash of block k -- off by default (slower,
    profiles/r01_overlap_experiment.txt) but it must give the same bits.  2^14 rows so that the tiles are two-step."""
    c = env_ctx(GLP_CO=1, GLP_CO_COLS=64, GLP_CO_FIRST=32)
    try:
        check_commit(c, oracle, edge_matrix(rng, 1 << 14, 200), 1 << 15, device=True)
    finally:
        c.close()


# ---------------------------------------------------------------- full-size known answers and the headline shape
def test_reference_bench_merkle_root_2p23_x_128(ctx, golden):
    """The reference's own large known answer: MERKLETREE_BENCH hashes the 2^23 x 128 column-Fibonacci matrix and asserts the
    root (src/goldilocks/benchs/bench.cpp:406-443).  The matrix (8.6 GB) is generated on the device."""
    import torch
    nrows, ncols = 1 << 23, 128
    dev = torch.device("cuda", 0)
    a = torch.empty(nrows * ncols, dtype=torch.int64, device=dev)
    ctx.fib_fill(a, nrows, ncols)
    # the generator itself: its first rows and a strip in the middle against the host recurrence
    head = a[: 64 * ncols].cpu().numpy().view(np.uint64).reshape(64, ncols)
    check_equal(head, fib_matrix(64, ncols), "device Fibonacci matrix, first rows")
    j = 5 * 512 - 2                                              # straddles a chunk boundary of the generator
    strip = a[j * ncols: (j + 4) * ncols].cpu().numpy().view(np.uint64).reshape(4, ncols)
    assert all((int(strip[2, c]) == (int(strip[0, c]) + int(strip[1, c])) % P) and (int(strip[3, c]) == (int(strip[1, c]) + int(strip[2, c])) % P)
               for c in range(ncols))
    tree = torch.empty(tree_words(nrows), dtype=torch.int64, device=dev)
    ctx.merkletree(tree, a, ncols, nrows)
    root = tree[-4:].cpu().numpy().view(np.uint64)
    check_equal(root, golden["bench_merkletree_8388608x128_root"]["out"], "bench.cpp MERKLETREE_BENCH root")


def test_config_b_full_size(ctx, oracle):
    """BASELINE.json configs[1] at its real size, 2^23 x 660 -> 2^24 (134 GB resident): whole columns of the extension -- the
    first, the last, and both sides of 32-lane group and 8-column sponge-block edges -- against the oracle's extendPol of the
    same columns, 48 Merkle paths (row -> leaf hash -> every sibling -> root) recomputed with the oracle's hashes, and the root
    against the committed value bench.py compares every N with."""
    import torch
    from xlayer_prover_b200 import synth_matrix
    free, _ = torch.cuda.mem_get_info(0)
    n, n_ext, C = 1 << 23, 1 << 24, 660
    need = 8 * (n * C + n_ext * C) + 8 * tree_words(n_ext) + (14 << 30)
    if free < need:
        pytest.skip("needs %.0f GB of device memory, %.0f GB free" % (need / 1e9, free / 1e9))
    seed = 0x9E3779B97F4A7C15 + 2
    dev = torch.device("cuda", 0)
    din = torch.empty(n * C, dtype=torch.int64, device=dev)
    ctx.synth_fill(din, n, C, seed)
    dext = torch.empty(n_ext * C, dtype=torch.int64, device=dev)
    dtree = torch.empty(tree_words(n_ext), dtype=torch.int64, device=dev)
    retries0 = ctx.ntt_retry_count()
    root = ctx.commit(dext, dtree, din, n, n_ext, C)
    assert ctx.ntt_retry_count() == retries0, "uniform synthetic data must not trip the optimistic LDE"
    del din
    ext = dext.view(n_ext, C)
    cols = [0, 7, 8, 31, 32, 63, 64, 329, 639, 640, 655, 656, 659]
    a = np.concatenate([synth_matrix(n, 1, seed, col0=c, ncols_total=C) for c in cols], axis=1)
    want = oracle.extend_pol(a, n_ext)
    got = ext[:, cols].cpu().numpy().view(np.uint64)
    check_equal(got, want, "whole columns %s of the 2^24 x 660 extension" % cols)
    tree = dtree.cpu().numpy().view(np.uint64)
    depth = 24
    prng = np.random.default_rng(7)
    for idx in [0, 1, n_ext - 1, n_ext // 2 - 1, n_ext // 2] + [int(x) for x in prng.integers(0, n_ext, size=43)]:
        row = ext[idx].cpu().numpy().view(np.uint64)
        h = oracle.linear_hash(row)
        check_equal(tree[4 * idx: 4 * idx + 4], h, "leaf %d" % idx)
        off, nn, i = 0, n_ext, idx
        for _ in range(depth):
            sib = tree[off + 4 * (i ^ 1): off + 4 * (i ^ 1) + 4]
            pair = np.concatenate([h, sib]) if i % 2 == 0 else np.concatenate([sib, h])
            h = oracle.hash_full_result(np.concatenate([pair, np.zeros(4, dtype=np.uint64)]))[:4]
            off += 4 * nn
            nn //= 2
            i //= 2
        check_equal(h, tree[-4:], "path of row %d to the root" % idx)
    check_equal(root, tree[-4:], "root returned by glp_commit")
    known = np.array([0x2954300a93d8e1ed, 0xb89e6e7f7a6d3906, 0x8e88252716f12068, 0xc8cda114b642004f], dtype=np.uint64)
    check_equal(root, known, "root of the headline workload (bench.py KNOWN_ROOTS)")


def test_group_proofs_batched(ctx, reference, rng):
    """glp_merkle_group_proofs on device-resident trees, against the reference's MerkleTreeGL::getGroupProof per index"""
    from cases import check_group_proofs
    check_group_proofs(ctx, reference, rng, shapes=((256, 21), (1 << 12, 3), (64, 128), (2, 5), (1, 9), (1 << 16, 40)))


def test_fri_prove_whole(ctx, reference, rng):
    """FRIProve::prove (the reference's own, compiled: oracle/_ref/libfriref.so) against the device-resident FRI flow"""
    from cases import check_fri_prove
    check_fri_prove(ctx, reference, rng, pol_bits=14, steps=(14, 10, 6, 3), nq=16, widths=(40, 12, 21, 6, 3))
    check_fri_prove(ctx, reference, rng, pol_bits=9, steps=(9, 3), nq=5)


# ---------------------------------------------------------------- several devices
def test_multi_contexts_on_one_device(oracle, rng):
    """glp_multi_* with G = 2 and 4 contexts on device 0 (a host thread each, the slabs in the same memory): the split, the
    scatter, the subtree placement and the top levels, on real kernels"""
    from cases import check_multi
    from xlayer_prover_b200 import MultiContext
    check_multi(lambda G: MultiContext([0] * G), oracle, rng,
                shapes=((1 << 12, 2, 72), (1 << 10, 8, 40), (1 << 14, 2, 660 // 4), (64, 2, 21), (8, 2, 3)))


def test_multi_commit_two_gpus(ctx, oracle, rng):
    """glp_multi_commit on two DISTINCT devices: peer stores over NVLink from the LDE's last pass; every extended row and every
    node against the single-device result of the same library and the oracle"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    from xlayer_prover_b200 import MultiContext, synth_matrix
    for devs in ([0, 1],) + (([0, 1, 2, 3],) if torch.cuda.device_count() >= 4 else ()):
        m = MultiContext(devs)
        try:
            for n, n_ext, ncols in ((1 << 15, 1 << 16, 72), (1 << 12, 1 << 15, 330), (1 << 16, 1 << 17, 24)):
                a = synth_matrix(n, ncols, 0xABCDEF)
                ext1, tree1, root1 = run_commit(ctx, a, n_ext)
                ext = np.zeros((n_ext, ncols), dtype=np.uint64)
                tree = np.zeros(tree_words(n_ext), dtype=np.uint64)
                root = m.commit(ext, tree, a, n, n_ext, ncols)
                check_equal(ext, ext1, "extension, devices %s" % devs)
                check_equal(tree, tree1, "nodes, devices %s" % devs)
                check_equal(root, root1, "root, devices %s" % devs)
                ext2 = np.zeros((n_ext, ncols), dtype=np.uint64)
                m.extend_pol(ext2, a, n_ext, n, ncols)
                check_equal(ext2, ext1, "multi extend_pol, devices %s" % devs)
                tree2 = np.zeros(tree_words(n_ext), dtype=np.uint64)
                m.merkletree(tree2, ext2, ncols, n_ext)
                check_equal(tree2, tree1, "multi merkletree, devices %s" % devs)
            want = oracle.merkletree(oracle.extend_pol(a, n_ext))
            check_equal(tree, want, "nodes against the oracle")
        finally:
            m.close()


def test_sharded_commit_two_gpus(tmp_path):
    """xlayer_prover_b200/sharded.py on real devices: one process per GPU under torch.distributed.run, NCCL, CUDA IPC and the
    fused NVLink scatter; every slab row and every node against the N = 1 device result (tests/sharded_gpu_worker.py)"""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    here = os.path.dirname(os.path.abspath(__file__))
    for world in [2] + ([4] if torch.cuda.device_count() >= 4 else []):
        out = tmp_path / ("w%d" % world)
        out.mkdir()
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
               "--master-port", str(29700 + world), os.path.join(here, "sharded_gpu_worker.py"), str(out), "15", "1", "168"]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
        assert (out / "result.txt").read_text().startswith("slabs_equal=1 tree_equal=1"), (out / "result.txt").read_text()
