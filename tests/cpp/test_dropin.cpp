// Exercises the C++ drop-in classes (include/ntt_goldilocks.hpp, include/merkleTreeGL.hpp) the way the
// reference's own gtest file does (src/goldilocks/tests/tests.cpp): same input generators, same known
// answers, compared through Goldilocks::toU64.  Built by tests/test_cpp_dropin.py with plain g++ against
// libglprover.so; needs a GPU to run (there is no CPU path behind the classes).
#include "merkleTreeGL.hpp"
#include "ntt_goldilocks.hpp"
#include "transcript.hpp"
#include <cstdio>
#include <vector>

static const uint64_t P = 0xFFFFFFFF00000001ULL;
static int failures = 0;
#define EXPECT_EQ(a, b)                                                                                                         \
    do                                                                                                                          \
    {                                                                                                                           \
        const uint64_t a_ = (a), b_ = (b);                                                                                      \
        if (a_ != b_)                                                                                                           \
        {                                                                                                                       \
            printf("FAIL %s:%d  %s = %#llx, expected %#llx\n", __FILE__, __LINE__, #a, (unsigned long long)a_, (unsigned long long)b_); \
            failures++;                                                                                                         \
        }                                                                                                                       \
    } while (0)

static uint64_t addp(uint64_t a, uint64_t b) { return (uint64_t)(((unsigned __int128)a + b) % P); }
typedef Goldilocks::Element El;

// column-wise Fibonacci, first two rows = 1 + j (tests.cpp:2563-2577 and :1822-1835)
static std::vector<El> fib_matrix(uint64_t rows, uint64_t cols, uint64_t rows_alloc)
{
    std::vector<El> a(rows_alloc * cols);
    for (uint64_t i = 0; i < 2 && i < rows; i++)
        for (uint64_t j = 0; j < cols; j++)
            a[i * cols + j].fe = 1 + j;
    for (uint64_t i = 2; i < rows; i++)
        for (uint64_t j = 0; j < cols; j++)
            a[i * cols + j].fe = addp(a[(i - 1) * cols + j].fe, a[(i - 2) * cols + j].fe);
    return a;
}

static void test_extendpol_kat()
{
    // tests.cpp:2554-2615: FFT_SIZE 2^4, NUM_COLUMNS 8, BLOWUP_FACTOR 1, in place, with a scratch buffer
    const uint64_t N = 1 << 4, NE = 1 << 5, C = 8;
    std::vector<El> a = fib_matrix(N, C, NE), b(NE * C);
    NTT_Goldilocks ntt(N);
    ntt.extendPol(a.data(), a.data(), NE, N, C, b.data());
    EXPECT_EQ(Goldilocks::toU64(a[0 * C]), 0X3E7CA26D67147C31ULL);
    EXPECT_EQ(Goldilocks::toU64(a[1 * C]), 0X1310720153E0ABE4ULL);
    EXPECT_EQ(Goldilocks::toU64(a[16 * C]), 0XC46EE848F93207D8ULL);
    EXPECT_EQ(Goldilocks::toU64(a[31 * C]), 0XA0705C72DD9F9A2FULL);
}

static void test_ntt_roundtrip()
{
    // tests.cpp:2225-2384: INTT(NTT(a)) == a for dst == src, dst != src, dst == NULL, any nphase/nblock; sizes 0, 1, 2
    const uint64_t N = 1 << 10, C = 7;
    std::vector<El> a = fib_matrix(N, C, N), ref = a, tmp(N * C);
    NTT_Goldilocks g(N);
    g.NTT(tmp.data(), a.data(), N, C);
    g.INTT(NULL, tmp.data(), N, C, NULL, 2, 3);
    for (uint64_t i = 0; i < N * C; i++)
        EXPECT_EQ(Goldilocks::toU64(tmp[i]), ref[i].fe);
    g.NTT(a.data(), a.data(), N, C, tmp.data(), 5, 2);
    g.INTT(a.data(), a.data(), N, C);
    for (uint64_t i = 0; i < N * C; i++)
        EXPECT_EQ(Goldilocks::toU64(a[i]), ref[i].fe);
    El one[3] = {{5}, {P + 2}, {9}};
    g.NTT(one, one, 1, 3); // size 1: identity (canonicalised)
    EXPECT_EQ(one[0].fe, 5);
    EXPECT_EQ(one[1].fe, 2);
    g.NTT(one, one, 0, 3); // size 0: no-op
    g.NTT(one, one, 1, 0); // ncols 0: no-op
    EXPECT_EQ(one[2].fe, 9);
    El two[2] = {{3}, {4}};
    g.NTT(two, two, 2);
    EXPECT_EQ(two[0].fe, 7);
    EXPECT_EQ(two[1].fe, P - 1);
    // root(domainPow, idx): W[2] = 2^48, W[1] = -1
    EXPECT_EQ(Goldilocks::toU64(g.root(2, 1)), 1ULL << 48);
    EXPECT_EQ(Goldilocks::toU64(g.root(1, 1)), P - 1);
    EXPECT_EQ(Goldilocks::toU64(g.root(10, 0)), 1);
}

static void test_errors_and_inert()
{
    NTT_Goldilocks inert(0); // starks.hpp:81-82 when !generateProof: must not touch the GPU
    bool thrown = false;
    try
    {
        NTT_Goldilocks big((u_int64_t)1 << 33);
    }
    catch (const std::range_error &)
    {
        thrown = true;
    }
    EXPECT_EQ(thrown, true);
    thrown = false;
    try
    {
        El x[3] = {{1}, {2}, {3}};
        NTT_Goldilocks g(4);
        g.NTT(x, x, 3); // not a power of two (the reference asserts)
    }
    catch (const std::invalid_argument &)
    {
        thrown = true;
    }
    EXPECT_EQ(thrown, true);
}

static void test_merkletree()
{
    // tests.cpp:1817-1850: 64 rows x 128 columns
    const uint64_t R = 64, C = 128;
    std::vector<El> cols = fib_matrix(R, C, R);
    MerkleTreeGL mt(R, C, cols.data());
    mt.merkelize();
    El root[4];
    mt.getRoot(root);
    EXPECT_EQ(Goldilocks::toU64(root[0]), 0X918F7CD0C3E8701FULL);
    EXPECT_EQ(Goldilocks::toU64(root[1]), 0X83A130E00F961B02ULL);
    EXPECT_EQ(Goldilocks::toU64(root[2]), 0X6921497B364123F8ULL);
    EXPECT_EQ(Goldilocks::toU64(root[3]), 0XBD2B98A57B748BF4ULL);
    EXPECT_EQ(mt.getTreeNumElements(), (2 * R - 1) * 4);
    EXPECT_EQ(mt.MerkleProofSize(), 6);
    // getGroupProof: row idx, then siblings; re-hash the path with one-row trees (a leaf pair is an 8-column row)
    const uint64_t idx = 37;
    std::vector<El> proof(C + 4 * mt.MerkleProofSize());
    mt.getGroupProof(proof.data(), idx);
    for (uint64_t j = 0; j < C; j++)
        EXPECT_EQ(proof[j].fe, cols[idx * C + j].fe);
    El h[4];
    std::memcpy(h, mt.nodes + 4 * idx, sizeof h);
    uint64_t id = idx;
    for (uint64_t l = 0; l < mt.MerkleProofSize(); l++)
    {
        El pair[8];
        std::memcpy(pair + ((id & 1) ? 4 : 0), h, sizeof h);
        std::memcpy(pair + ((id & 1) ? 0 : 4), proof.data() + C + 4 * l, sizeof h);
        MerkleTreeGL one(1, 8, pair); // linear_hash of 8 words = one permutation with zero capacity = the node hash
        one.merkelize();
        one.getRoot(h);
        id >>= 1;
    }
    for (int k = 0; k < 4; k++)
        EXPECT_EQ(h[k].fe, root[k].fe);
    // ncols = 0 (tests.cpp:1852-1879)
    MerkleTreeGL empty(64, 0, NULL);
    empty.merkelize();
    empty.getRoot(root);
    EXPECT_EQ(Goldilocks::toU64(root[0]), 0X25225F1A5D49614AULL);
    EXPECT_EQ(Goldilocks::toU64(root[3]), 0X49561260080D30C3ULL);
    // serialised-tree constructor (merkleTreeGL.hpp:24-32) + copySource
    std::vector<El> ser(2 + R * C + (2 * R - 1) * 4);
    ser[0].fe = C;
    ser[1].fe = R;
    MerkleTreeGL view(ser.data());
    view.copySource(cols.data());
    view.merkelize();
    view.getRoot(root);
    EXPECT_EQ(Goldilocks::toU64(root[0]), 0X918F7CD0C3E8701FULL);
}

static void test_transcript()
{
    // answers produced by the reference's own Transcript (oracle/_ref, transcript.cpp) for put(1..19); getField; getPermutations(8, 5)
    Transcript t;
    std::vector<El> in(19);
    for (uint64_t i = 0; i < 19; i++)
        in[i].fe = i + 1;
    t.put(in.data(), 19);
    El f[3];
    t.getField(f);
    EXPECT_EQ(Goldilocks::toU64(f[0]), 14077115803306752431ULL);
    EXPECT_EQ(Goldilocks::toU64(f[1]), 4970989821272620831ULL);
    EXPECT_EQ(Goldilocks::toU64(f[2]), 17413480707857030579ULL);
    uint64_t perm[8];
    const uint64_t want[8] = {25, 8, 5, 7, 29, 13, 27, 26};
    t.getPermutations(perm, 8, 5);
    for (int i = 0; i < 8; i++)
        EXPECT_EQ(perm[i], want[i]);
}

// Starks::genProof step 1 as starks.cpp:53-57 writes it -- extendPol, then merkelize() -- on a matrix wide and tall enough to be
// split over the devices of GLP_DEVICES; prints the root and checksums so that the caller (tests/test_cpp_dropin.py) can compare
// runs with different device lists: the bits must not depend on the device count.
static void test_commit_sequence()
{
    const uint64_t N = 1 << 12, NE = 1 << 13, C = 40;
    std::vector<El> a = fib_matrix(N, C, N), ext(NE * C);
    NTT_Goldilocks ntt(N);
    ntt.extendPol(ext.data(), a.data(), NE, N, C);
    MerkleTreeGL mt(NE, C, ext.data());
    mt.merkelize();
    El root[4];
    mt.getRoot(root);
    uint64_t ck_ext = 0, ck_nodes = 0;
    for (uint64_t i = 0; i < NE * C; i++)
        ck_ext = ck_ext * 0x9E3779B97F4A7C15ULL + Goldilocks::toU64(ext[i]);
    for (uint64_t i = 0; i < mt.getTreeNumElements(); i++)
        ck_nodes = ck_nodes * 0x9E3779B97F4A7C15ULL + Goldilocks::toU64(mt.nodes[i]);
    printf("commit root %016llx %016llx %016llx %016llx ext %016llx nodes %016llx devices %zu\n", (unsigned long long)Goldilocks::toU64(root[0]),
           (unsigned long long)Goldilocks::toU64(root[1]), (unsigned long long)Goldilocks::toU64(root[2]), (unsigned long long)Goldilocks::toU64(root[3]),
           (unsigned long long)ck_ext, (unsigned long long)ck_nodes, glp::device_list().size());
    // an NTT through an object built with extension > 1 reads the upper rows as zero (ntt_goldilocks.cpp:259-277)
    const uint64_t M = 64, K = 3;
    std::vector<El> x = fib_matrix(M, K, M), z = x, y1(M * K), y2(M * K);
    for (uint64_t i = M / 2; i < M; i++)
        for (uint64_t j = 0; j < K; j++)
            z[i * K + j].fe = 0;
    NTT_Goldilocks ext2(M, 0, 2), plain(M);
    ext2.NTT(y1.data(), x.data(), M, K);
    plain.NTT(y2.data(), z.data(), M, K);
    for (uint64_t i = 0; i < M * K; i++)
        EXPECT_EQ(Goldilocks::toU64(y1[i]), Goldilocks::toU64(y2[i]));
}

int main()
{
    test_errors_and_inert();
    test_extendpol_kat();
    test_ntt_roundtrip();
    test_merkletree();
    test_transcript();
    test_commit_sequence();
    if (failures)
    {
        printf("%d failure(s)\n", failures);
        return 1;
    }
    printf("dropin ok\n");
    return 0;
}
