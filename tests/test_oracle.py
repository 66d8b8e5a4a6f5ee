"""Pins the oracle: the plain-C restatement (oracle/gl_oracle.c) and, when built, the compiled reference
(oracle/_ref) against every known answer the reference's own tests hold for the path, against the committed
reference-generated fixtures, and against each other on seeded inputs.  CPU only."""
import numpy as np
import pytest

from cases import P, check_equal, edge_matrix, fib_matrix, fib_vector, tree_words


def _impls(request):
    from oracle import refbind
    out = [("port", request.getfixturevalue("port"))]
    if refbind.reference() is not None:
        out.append(("reference", refbind.reference()))
        if refbind.reference(avx512=True) is not None:
            out.append(("reference-avx512", refbind.reference(avx512=True)))
    return out


def test_poseidon_kat(request, golden):
    for name, o in _impls(request):
        check_equal(o.hash_full_result(fib_vector(12)), golden["poseidon_fibonacci"]["out"], name)
        check_equal(o.hash_full_result(np.zeros(12, dtype=np.uint64)), golden["poseidon_zero"]["out"], name)
        check_equal(o.hash_full_result(fib_vector(12), seq=True), golden["poseidon_fibonacci"]["out"], name)


def test_linear_hash_kat(request, golden):
    for name, o in _impls(request):
        check_equal(o.linear_hash(fib_vector(128)), golden["linear_hash_fibonacci128"]["out"], name)


def test_merkletree_kat(request, golden):
    for name, o in _impls(request):
        check_equal(o.merkletree(fib_matrix(64, 128))[-4:], golden["merkletree_64x128_root"]["out"], name)
        check_equal(o.merkletree(None, 64, 0)[-4:], golden["merkletree_64x0_root"]["out"], name)
        check_equal(o.merkletree(None, 1 << 17, 0)[-4:], golden["merkletree_131072x0_root"]["out"], name)
        check_equal(o.merkletree_batch(fib_matrix(64, 128), 32)[-4:], golden["merkletree_batch_64x128_b32_root"]["out"], name)


def test_extendpol_kat(request, golden):
    for name, o in _impls(request):
        ext = o.extend_pol(fib_matrix(16, 8), 32)
        check_equal(ext[:, 0], golden["extendpol_16x8_to_32_col0"]["out"], name)
        check_equal(ext[:, 0], golden["lde_block_16x8_to_32_col0"]["out"], name)


def test_lde_shift49_kat(request, golden):
    """tests.cpp LDE: INTT, scale by 49^i on the host, zero-pad, NTT of twice the size."""
    for name, o in _impls(request):
        a = fib_vector(16, 1, 1).reshape(16, 1)
        coef = o.ntt(a, inverse=True)
        r = 1
        padded = np.zeros((32, 1), dtype=np.uint64)
        for i in range(16):
            padded[i, 0] = int(coef[i, 0]) * r % P
            r = r * 49 % P
        check_equal(o.ntt(padded), golden["lde_16_to_32_shift49"]["out"], name)


def test_ntt_roundtrip_and_edges(request):
    rng = np.random.default_rng(5)
    for name, o in _impls(request):
        for n, c in [(1, 3), (2, 3), (16, 8), (64, 1)]:
            a = edge_matrix(rng, n, c)
            back = o.ntt(o.ntt(a), inverse=True)
            check_equal(back, o.canonical(a), name)
        # size 0 / ncols 0 are no-ops (tests.cpp:2374-2383)
        assert o.ntt(np.zeros((0, 3), dtype=np.uint64)).size == 0


def test_fixtures_match_port(port, fixtures):
    """The committed fixtures came from the compiled reference (tests/golden/make_golden.py)."""
    f = fixtures
    for k in [k[:-3] for k in f.files if k.startswith("ext_") and k.endswith("_in")]:
        check_equal(port.extend_pol(f[k + "_in"], f[k + "_out"].shape[0]), f[k + "_out"], k)
    for k in [k[:-3] for k in f.files if k.startswith("ntt_") and k.endswith("_in")]:
        check_equal(port.ntt(f[k + "_in"]), f[k + "_fwd"], k)
        check_equal(port.ntt(f[k + "_in"], inverse=True), f[k + "_inv"], k)
        check_equal(port.ntt(f[k + "_in"], inverse=True, extend=True), f[k + "_invext"], k)
    for k in [k[:-3] for k in f.files if k.startswith("mt_") and k.endswith("_in")]:
        check_equal(port.merkletree(f[k + "_in"]), f[k + "_tree"], k)
    for i, row in enumerate(f["perm_in"]):
        check_equal(port.hash_full_result(row), f["perm_out"][i], "perm %d" % i)
    check_equal(port.extend_pol(f["commit_64x9_in"], 128), f["commit_64x9_ext"], "commit ext")
    check_equal(port.merkletree(f["commit_64x9_ext"]), f["commit_64x9_tree"], "commit tree")


def test_port_vs_reference_random(port, reference):
    rng = np.random.default_rng(11)
    for k in range(33):
        assert port.w(k) == reference.w(k)
    edges = [0, 1, P - 1, P, P + 1, 2 ** 64 - 1, 2 ** 32, 2 ** 32 - 1]
    for a in edges:
        for b in edges:
            assert port.add(a, b) == reference.to_u64(reference.add(a, b))
            assert port.sub(a, b) == reference.to_u64(reference.sub(a, b))
            assert port.mul(a, b) == reference.to_u64(reference.mul(a, b))
    for n, c, blow in [(32, 7, 2), (128, 3, 4), (8, 12, 8)]:
        a = edge_matrix(rng, n, c)
        check_equal(port.extend_pol(a, n * blow), reference.extend_pol(a, n * blow), "extend %dx%d" % (n, c))
    for n, c in [(16, 9), (8, 4), (8, 5), (2, 660), (64, 17)]:
        a = edge_matrix(rng, n, c)
        check_equal(port.merkletree(a), reference.merkletree(a), "tree %dx%d" % (n, c))
        check_equal(port.merkletree(a), reference.merkle_tree_gl(a)[0], "MerkleTreeGL %dx%d" % (n, c))
    # avx2 and avx512 builds of the reference agree
    from oracle import refbind
    r512 = refbind.reference(avx512=True)
    if r512 is not None:
        a = edge_matrix(rng, 64, 33)
        check_equal(r512.merkletree(a), reference.merkletree(a), "avx512 tree")
        check_equal(r512.extend_pol(a, 128), reference.extend_pol(a, 128), "avx512 extend")


def test_tree_words():
    assert tree_words(64) == 508


def test_reference_bctree_interpolate_is_extend_pol(reference):
    """tools/starkpil/bctree/build_const_tree.cpp:398-531 (`interpolate`, compiled from the reference into
    oracle/_ref/libbctree.so) and NTT_Goldilocks::extendPol are the same map -- which is why glp_const_tree may be glp_commit
    writing into the *.consttree layout.  Checked here on the reference's two implementations, not asserted in a comment."""
    rng = np.random.default_rng(5)
    for n, n_ext, npols in ((16, 32, 5), (64, 256, 3), (1 << 13, 1 << 14, 2), (8, 16, 1)):
        a = rng.integers(0, 0xFFFFFFFF00000001, size=(n, npols), dtype=np.uint64)
        assert np.array_equal(reference.bctree_interpolate(a, n_ext), reference.extend_pol(a, n_ext)), (n, n_ext, npols)


def test_reference_fri_fold_helpers_against_first_principles(reference):
    """FRIProve::polMulAxi / evalPol around the reference INTT (oracle/_ref/libfriref.so), on a fold whose answer is known in
    closed form: a polynomial of degree < nX per coset folds to its evaluation -- here friPol = evaluations of the constant
    polynomial c on the domain, so every folded value is c whatever special_x is."""
    P = 0xFFFFFFFF00000001
    pol = np.zeros((1 << 6, 3), dtype=np.uint64)
    pol[:] = np.array([5, 7, 11], dtype=np.uint64)
    out = reference.fri_fold(pol, 6, 3, np.array([123, 456, 789], dtype=np.uint64), pow(7, P - 2, P))
    assert np.array_equal(out, np.tile(np.array([5, 7, 11], dtype=np.uint64), (8, 1)))
