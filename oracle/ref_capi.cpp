// TEST INFRASTRUCTURE ONLY -- not part of the product, never linked into libglprover.so.
//
// extern "C" face over the UNMODIFIED reference classes, compiled from the sources where
// they lie under $REF (= /root/reference) by oracle/Makefile into oracle/_ref/libglref*.so.
// It is the parity oracle (tests/) and the timed CPU baseline (bench.py --impl reference,
// cpu_baseline leg).  Nothing here is reference source: each function only forwards to
//   NTT_Goldilocks      $REF/src/goldilocks/src/ntt_goldilocks.{hpp,cpp}
//   PoseidonGoldilocks  $REF/src/goldilocks/src/poseidon_goldilocks.{hpp,cpp}
//   MerkleTreeGL        $REF/src/starkpil/merkleTree/merkleTreeGL.{hpp,cpp}
//   Transcript          $REF/src/starkpil/transcript/transcript.{hpp,cpp}
// or, for the two call sequences that live inside Starks::genProof / bctree (which do not compile stand-alone),
// repeats that sequence of reference calls with the line numbers it follows.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <omp.h>
#include "goldilocks_base_field.hpp"
#include "ntt_goldilocks.hpp"
#include "poseidon_goldilocks.hpp"
#include "merklehash_goldilocks.hpp"
#include "merkleTreeGL.hpp"
#include "transcript.hpp"
#include "goldilocks_cubic_extension.hpp"
#include <vector>

typedef Goldilocks::Element E;
static_assert(sizeof(E) == sizeof(uint64_t), "Element must be a bare uint64_t");

extern "C" {

int ref_is_avx512(void)
{
#ifdef __AVX512__
    return 1;
#else
    return 0;
#endif
}
int ref_max_threads(void) { return omp_get_max_threads(); }
void ref_set_threads(int n) { omp_set_num_threads(n); }

// ---- field (goldilocks_base_field_scalar.hpp:16-160, _tools.hpp:146-163) ----
uint64_t ref_add(uint64_t a, uint64_t b) { E x{a}, y{b}; return Goldilocks::add(x, y).fe; }
uint64_t ref_sub(uint64_t a, uint64_t b) { E x{a}, y{b}; return Goldilocks::sub(x, y).fe; }
uint64_t ref_mul(uint64_t a, uint64_t b) { E x{a}, y{b}; return Goldilocks::mul(x, y).fe; }
uint64_t ref_to_u64(uint64_t a) { E x{a}; return Goldilocks::toU64(x); }
uint64_t ref_w(uint32_t k) { return Goldilocks::toU64(Goldilocks::w(k)); }
uint64_t ref_shift(void) { return Goldilocks::toU64(Goldilocks::shift()); }
void ref_canonicalize(uint64_t *a, uint64_t n)
{
#pragma omp parallel for
    for (uint64_t i = 0; i < n; i++) a[i] = Goldilocks::toU64(((E *)a)[i]);
}

// ---- NTT (ntt_goldilocks.cpp:166-352) ----
// returns 0, or -2 when the reference throws range_error (ntt_goldilocks.hpp:94-97)
int ref_ntt(uint64_t *dst, uint64_t *src, uint64_t size, uint64_t ncols, uint64_t nphase, uint64_t nblock, int inverse, int extend)
{
    try {
        NTT_Goldilocks ntt(size);
        if (inverse) {
            if (extend) ntt.computeR((int)size);
            ntt.INTT((E *)dst, (E *)src, size, ncols, NULL, nphase, nblock, extend != 0);
        } else {
            ntt.NTT((E *)dst, (E *)src, size, ncols, NULL, nphase, nblock);
        }
    } catch (const std::range_error &) { return -2; }
    return 0;
}
// an NTT_Goldilocks sized for max_size used on a smaller transform (root(domainPow, idx) striding)
int ref_ntt_maxsize(uint64_t *dst, uint64_t *src, uint64_t max_size, uint64_t size, uint64_t ncols, int inverse)
{
    NTT_Goldilocks ntt(max_size);
    if (inverse) ntt.INTT((E *)dst, (E *)src, size, ncols);
    else ntt.NTT((E *)dst, (E *)src, size, ncols);
    return 0;
}
int ref_extend_pol(uint64_t *output, uint64_t *input, uint64_t n_ext, uint64_t n, uint64_t ncols, uint64_t *buffer, uint64_t nphase, uint64_t nblock)
{
    NTT_Goldilocks ntt(n);
    ntt.extendPol((E *)output, (E *)input, n_ext, n, ncols, (E *)buffer, nphase, nblock);
    return 0;
}
void ref_reverse_permutation(uint64_t *dst, uint64_t *src, uint64_t size, uint64_t offset_cols, uint64_t ncols, uint64_t ncols_all)
{
    NTT_Goldilocks ntt(size);
    ntt.reversePermutation((E *)dst, (E *)src, size, offset_cols, ncols, ncols_all);
}
uint64_t ref_root(uint64_t max_size, uint32_t domain_pow, uint64_t idx)
{
    NTT_Goldilocks ntt(max_size);
    return Goldilocks::toU64(ntt.root(domain_pow, idx));
}

// ---- Poseidon (poseidon_goldilocks.cpp) ----
void ref_hash_full_result_seq(uint64_t *out12, const uint64_t *in12) { PoseidonGoldilocks::hash_full_result_seq((E *)out12, (const E *)in12); }
void ref_hash_full_result(uint64_t *out12, const uint64_t *in12)
{
#ifdef __AVX512__
    // the AVX-512 kernel permutes two states at once (poseidon_goldilocks.cpp:355-431)
    E in2[24], out2[24];
    std::memcpy(in2, in12, 96); std::memcpy(in2 + 12, in12, 96);
    // interleave as the reference's own callers do (linear_hash_avx512, :432-483)
    E tin[24], tout[24];
    for (int g = 0; g < 3; g++) for (int k = 0; k < 4; k++) { tin[g * 8 + k] = in2[g * 4 + k]; tin[g * 8 + 4 + k] = in2[12 + g * 4 + k]; }
    PoseidonGoldilocks::hash_full_result_avx512(tout, tin);
    for (int g = 0; g < 3; g++) for (int k = 0; k < 4; k++) out2[g * 4 + k] = tout[g * 8 + k];
    std::memcpy(out12, out2, 96);
#else
    PoseidonGoldilocks::hash_full_result((E *)out12, (const E *)in12);
#endif
}
void ref_linear_hash_seq(uint64_t *out4, uint64_t *in, uint64_t size) { PoseidonGoldilocks::linear_hash_seq((E *)out4, (E *)in, size); }
void ref_linear_hash(uint64_t *out4, uint64_t *in, uint64_t size) { PoseidonGoldilocks::linear_hash((E *)out4, (E *)in, size); }
void ref_linear_hash_rows(uint64_t *out4, uint64_t *in, uint64_t ncols, uint64_t nrows)
{
#pragma omp parallel for
    for (uint64_t i = 0; i < nrows; i++) PoseidonGoldilocks::linear_hash((E *)out4 + 4 * i, (E *)in + i * ncols, ncols);
}
// the dispatcher MerkleTreeGL::merkelize uses (merkleTreeGL.cpp:37-44): avx512 when built with -D__AVX512__, else avx
void ref_merkletree(uint64_t *tree, uint64_t *in, uint64_t ncols, uint64_t nrows, int nthreads, uint64_t dim)
{
    PoseidonGoldilocks::merkletree((E *)tree, (E *)in, ncols, nrows, nthreads, dim);
}
void ref_merkletree_seq(uint64_t *tree, uint64_t *in, uint64_t ncols, uint64_t nrows, int nthreads, uint64_t dim)
{
    PoseidonGoldilocks::merkletree_seq((E *)tree, (E *)in, ncols, nrows, nthreads, dim);
}
void ref_merkletree_batch(uint64_t *tree, uint64_t *in, uint64_t ncols, uint64_t nrows, uint64_t batch, int nthreads, uint64_t dim)
{
    PoseidonGoldilocks::merkletree_batch((E *)tree, (E *)in, ncols, nrows, batch, nthreads, dim);
}
uint64_t ref_tree_num_elements(uint64_t degree) { return MerklehashGoldilocks::getTreeNumElements(degree); }

// ---- MerkleTreeGL (merkleTreeGL.hpp:9-79, .cpp:5-44) ----
void *ref_mt_new(uint64_t height, uint64_t width, uint64_t *source) { return new MerkleTreeGL(height, width, (E *)source); }
void *ref_mt_map(uint64_t *tree) { return new MerkleTreeGL((E *)tree); }
void ref_mt_free(void *mt) { delete (MerkleTreeGL *)mt; }
void ref_mt_copy_source(void *mt, uint64_t *src) { ((MerkleTreeGL *)mt)->copySource((E *)src); }
void ref_mt_merkelize(void *mt) { ((MerkleTreeGL *)mt)->merkelize(); }
uint64_t *ref_mt_nodes(void *mt) { return (uint64_t *)((MerkleTreeGL *)mt)->nodes; }
uint64_t *ref_mt_source(void *mt) { return (uint64_t *)((MerkleTreeGL *)mt)->source; }
uint64_t ref_mt_num_elements(void *mt) { return ((MerkleTreeGL *)mt)->getTreeNumElements(); }
uint64_t ref_mt_proof_size(void *mt) { return ((MerkleTreeGL *)mt)->MerkleProofSize(); }
void ref_mt_get_root(void *mt, uint64_t *root4) { ((MerkleTreeGL *)mt)->getRoot((E *)root4); }
void ref_mt_get_group_proof(void *mt, uint64_t *proof, uint64_t idx) { ((MerkleTreeGL *)mt)->getGroupProof((E *)proof, idx); }

// ---- the hot path as Starks::genProof runs it (starks.cpp:53-57): extendPol then merkelize ----
// seconds[0] = extendPol, seconds[1] = merkletree.  buffer must hold n_ext*ncols words.
int ref_commit(uint64_t *out_ext, uint64_t *tree, uint64_t *in, uint64_t n, uint64_t n_ext, uint64_t ncols, uint64_t *buffer, double *seconds)
{
    NTT_Goldilocks ntt(n);                 // built once per prover in the reference (starks.hpp:81), so untimed
    double t0 = omp_get_wtime();
    ntt.extendPol((E *)out_ext, (E *)in, n_ext, n, ncols, (E *)buffer);
    double t1 = omp_get_wtime();
    PoseidonGoldilocks::merkletree((E *)tree, (E *)out_ext, ncols, n_ext);
    double t2 = omp_get_wtime();
    if (seconds) { seconds[0] = t1 - t0; seconds[1] = t2 - t1; }
    return 0;
}

// ---- step 4 of Starks::genProof, the quotient commit (starks.cpp:232-233,261-291): INTT of the quotient on the
// extended domain (qDim columns), split into qDeg chunks of N coefficients scaled by shiftIn^p, shiftIn = shift^-N
// (:264-281; Goldilocks3::mul by a base-field scalar is a per-component multiply), forward NTT of the zero-padded
// N x (qDeg*qDim) matrix (:285), Merkle tree over its NExtended rows (:290). ----
int ref_quotient_commit(uint64_t *cm4, uint64_t *tree, uint64_t *q_2ns, uint64_t N, uint64_t NExtended, uint64_t qDim, uint64_t qDeg)
{
    NTT_Goldilocks nttExtended(NExtended);
    std::vector<E> qq1(NExtended * qDim), qq2(NExtended * qDeg * qDim);
    std::memset((void *)qq2.data(), 0, qq2.size() * sizeof(E));
    nttExtended.INTT(qq1.data(), (E *)q_2ns, NExtended, qDim, NULL, 2, 1);
    E shiftIn = Goldilocks::exp(Goldilocks::inv(Goldilocks::shift()), N);
    E curS = Goldilocks::one();
    for (uint64_t p = 0; p < qDeg; p++) {
        for (uint64_t k = 0; k < N; k++)
            for (uint64_t d = 0; d < qDim; d++) qq2[(k * qDeg + p) * qDim + d] = Goldilocks::mul(qq1[(p * N + k) * qDim + d], curS);
        curS = Goldilocks::mul(curS, shiftIn);
    }
    nttExtended.NTT((E *)cm4, qq2.data(), NExtended, qDim * qDeg);
    if (tree) PoseidonGoldilocks::merkletree((E *)tree, (E *)cm4, qDim * qDeg, NExtended);
    return 0;
}

// The constant tree (bctree) and the FRI folding step are NOT restated here: oracle/_ref/libbctree.so and libfriref.so are the
// reference's own build_const_tree.cpp and friProve.cpp compiled where they lie (oracle/ref_bctree_capi.cpp, ref_fri_capi.cpp).

// ---- the FRI layer tree (friProve.cpp:110-123,252-270): getTransposed regroups the folded polynomial into
// nGroups = 2^nextBits rows of groupSize = polN / nGroups extension elements, MerkleTreeGL hashes the rows. ----
int ref_fri_group_tree(uint64_t *aux, uint64_t *tree, uint64_t *pol, uint64_t polN, uint64_t nextBits)
{
    const uint64_t w = 1ull << nextBits, h = polN / w;
    for (uint64_t i = 0; i < w; i++)
        for (uint64_t j = 0; j < h; j++) std::memcpy(aux + (i * h + j) * 3, pol + (j * w + i) * 3, 3 * sizeof(uint64_t));
    MerkleTreeGL mt(w, h * 3, NULL);
    mt.copySource((E *)aux);
    mt.merkelize();
    std::memcpy(tree, mt.nodes, mt.getTreeNumElements() * sizeof(E));
    return 0;
}

// ---- Transcript (transcript.cpp:4-91) ----
void *ref_tr_new(void) { return new Transcript(); }
void ref_tr_free(void *t) { delete (Transcript *)t; }
void ref_tr_put(void *t, uint64_t *in, uint64_t n) { ((Transcript *)t)->put((E *)in, n); }
void ref_tr_get_field(void *t, uint64_t *out3) { ((Transcript *)t)->getField((E *)out3); }
uint64_t ref_tr_get_fields1(void *t) { return Goldilocks::toU64(((Transcript *)t)->getFields1()); }
void ref_tr_get_permutations(void *t, uint64_t *res, uint64_t n, uint64_t nbits) { ((Transcript *)t)->getPermutations(res, n, nbits); }
} // extern "C"
