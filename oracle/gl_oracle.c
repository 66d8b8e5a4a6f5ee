/* TEST INFRASTRUCTURE ONLY -- the parity oracle's plain-C restatement ("port").
 *
 * Not part of the product: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this.  libglprover.so never links or calls it.
 *
 * A CPU restatement of the reference's STARK-commitment hot path, written from the maths the
 * reference implements (each function cites the reference file:line it follows).  Parity is
 * PINNED: tests/test_oracle.py checks this file against every golden vector the reference's own
 * tests hold for the path (tests/golden/kat.json, from src/goldilocks/tests/tests.cpp) and against
 * the compiled reference itself (oracle/_ref/libglref.so) on seeded inputs.
 *
 * All results are canonical (in [0,p)); the reference's are not necessarily
 * (goldilocks_base_field_scalar.hpp:16-160), so comparisons canonicalise the reference side,
 * exactly as the reference's tests do through Goldilocks::toU64 (goldilocks_base_field_tools.hpp:146-163).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <omp.h>
#include "poseidon_constants.h"

typedef unsigned __int128 u128;
#define GL_P 0xFFFFFFFF00000001ULL

/* ---- field: goldilocks_base_field_scalar.hpp:16-35 (add), :62-80 (sub), :103-160 (mul) ---- */
static inline uint64_t fcanon(uint64_t a) { return a >= GL_P ? a - GL_P : a; }
static inline uint64_t fadd(uint64_t a, uint64_t b) { return (uint64_t)(((u128)fcanon(a) + fcanon(b)) % GL_P); }
static inline uint64_t fsub(uint64_t a, uint64_t b) { return (uint64_t)(((u128)fcanon(a) + GL_P - fcanon(b)) % GL_P); }
static inline uint64_t fmul(uint64_t a, uint64_t b) { return (uint64_t)(((u128)a * b) % GL_P); }
static uint64_t fpow(uint64_t b, uint64_t e)
{
    uint64_t r = 1;
    b = fcanon(b);
    while (e) { if (e & 1) r = fmul(r, b); b = fmul(b, b); e >>= 1; }
    return r;
}
static inline uint64_t finv(uint64_t a) { return fpow(a, GL_P - 2); }

uint64_t orc_add(uint64_t a, uint64_t b) { return fadd(a, b); }
uint64_t orc_sub(uint64_t a, uint64_t b) { return fsub(a, b); }
uint64_t orc_mul(uint64_t a, uint64_t b) { return fmul(a, b); }
uint64_t orc_to_u64(uint64_t a) { return fcanon(a); }
/* goldilocks_base_field.cpp:11-51: W[32] = 7277203076849721926 generates the 2^32 roots of unity and
 * W[k] = W[32]^(2^(32-k)) (so W[6] = 8, W[5] = 64, W[4] = 2^12, W[3] = 2^24, W[2] = 2^48); SHIFT = 7 */
#define GL_W32 7277203076849721926ULL
uint64_t orc_w(uint32_t k)
{
    uint64_t w = GL_W32;
    for (uint32_t i = k; i < 32; i++) w = fmul(w, w);
    return w;
}
uint64_t orc_shift(void) { return 7; }
void orc_canonicalize(uint64_t *a, uint64_t n)
{
#pragma omp parallel for
    for (uint64_t i = 0; i < n; i++) a[i] = fcanon(a[i]);
}
int orc_max_threads(void) { return omp_get_max_threads(); }
void orc_set_threads(int n) { omp_set_num_threads(n); }

/* ---- NTT: ntt_goldilocks.cpp:3-10 (BR), :12-164 (NTT_iters), :166-232 (NTT), :307-324 (INTT) ---- */
static uint32_t ilog2(uint64_t n) { uint32_t r = 0; while (n > 1) { n >>= 1; r++; } return r; }
static uint64_t bitrev(uint64_t x, uint32_t bits)
{
    uint64_t r = 0;
    for (uint32_t i = 0; i < bits; i++) r |= ((x >> i) & 1) << (bits - 1 - i);
    return r;
}
/* out[i,:] = sum_k in[k,:] * w^(i k), natural order both sides; `a` is size x ncols row-major, in place.
 * rows >= live are known to be zero on entry only through the caller having written zeros. */
static void ntt_core(uint64_t *a, uint64_t size, uint64_t ncols)
{
    uint32_t lg = ilog2(size);
    /* bit-reversal permutation of whole rows (reversePermutation, ntt_goldilocks.cpp:243-305) */
    uint64_t *row = (uint64_t *)malloc(ncols * sizeof(uint64_t));
    for (uint64_t i = 0; i < size; i++) {
        uint64_t r = bitrev(i, lg);
        if (r > i) {
            memcpy(row, a + i * ncols, ncols * 8);
            memcpy(a + i * ncols, a + r * ncols, ncols * 8);
            memcpy(a + r * ncols, row, ncols * 8);
        }
    }
    free(row);
    /* radix-2 decimation in time: t = w*a[hi]; a[lo] = u + t; a[hi] = u - t (ntt_goldilocks.cpp:102-109) */
    for (uint32_t s = 1; s <= lg; s++) {
        uint64_t m = 1ULL << s, half = m >> 1;
        uint64_t wm = orc_w(s);
        uint64_t *tw = (uint64_t *)malloc(half * sizeof(uint64_t));
        tw[0] = 1;
        for (uint64_t j = 1; j < half; j++) tw[j] = fmul(tw[j - 1], wm);
#pragma omp parallel for schedule(static)
        for (uint64_t b = 0; b < size / 2; b++) {
            uint64_t k = (b / half) * m, j = b % half;
            uint64_t *lo = a + (k + j) * ncols, *hi = a + (k + j + half) * ncols;
            uint64_t w = tw[j];
            for (uint64_t c = 0; c < ncols; c++) {
                uint64_t t = fmul(w, hi[c]), u = fcanon(lo[c]);
                lo[c] = fadd(u, t);
                hi[c] = fsub(u, t);
            }
        }
        free(tw);
    }
}

int orc_ntt(uint64_t *dst, uint64_t *src, uint64_t size, uint64_t ncols, uint64_t nphase, uint64_t nblock, int inverse, int extend)
{
    (void)nphase; (void)nblock;               /* performance hints only (tests.cpp:2316-2338) */
    if (size == 0 || ncols == 0) return 0;    /* ntt_goldilocks.cpp:168-171 */
    if (size & (size - 1)) return -1;         /* assert, ntt_goldilocks.cpp:28 */
    if (size > (1ULL << 32)) return -2;       /* range_error, ntt_goldilocks.hpp:94-97 */
    if (dst == NULL) dst = src;               /* ntt_goldilocks.cpp:15-22, :315-322 */
    uint64_t *a = (uint64_t *)malloc(size * ncols * sizeof(uint64_t));
    memcpy(a, src, size * ncols * 8);
    ntt_core(a, size, ncols);
    if (!inverse) {
        memcpy(dst, a, size * ncols * 8);
    } else {
        /* last inverse phase: out[(N-i)%N] = y[i] * N^-1 (* shift^((N-i)%N) when extend)
         * (ntt_goldilocks.cpp:122-150; r_[i] = shift^i / N, ntt_goldilocks.hpp:153-165) */
        uint64_t ninv = finv(size % GL_P);
        uint64_t *sc = (uint64_t *)malloc(size * sizeof(uint64_t));
        sc[0] = ninv;
        for (uint64_t i = 1; i < size; i++) sc[i] = extend ? fmul(sc[i - 1], 7) : ninv;
#pragma omp parallel for schedule(static)
        for (uint64_t i = 0; i < size; i++) {
            uint64_t d = (size - i) % size;
            for (uint64_t c = 0; c < ncols; c++) dst[d * ncols + c] = fmul(a[i * ncols + c], sc[d]);
        }
        free(sc);
    }
    free(a);
    return 0;
}

/* extendPol: ntt_goldilocks.cpp:326-352.  output[i,c] = P_c(shift * w_ext^i); output may alias input. */
int orc_extend_pol(uint64_t *output, uint64_t *input, uint64_t n_ext, uint64_t n, uint64_t ncols, uint64_t *buffer, uint64_t nphase, uint64_t nblock)
{
    (void)buffer; (void)nphase; (void)nblock;
    if (n == 0 || ncols == 0) return 0;
    uint64_t *a = (uint64_t *)calloc(n_ext * ncols, sizeof(uint64_t));
    /* INTT with the coset shift folded in, coefficients land in rows [0,n) */
    orc_ntt(a, input, n, ncols, 0, 0, 1, 1);
    /* rows [n, n_ext) are the zero padding (reversePermutation zero-fill, ntt_goldilocks.cpp:259-277) */
    ntt_core(a, n_ext, ncols);
    memcpy(output, a, n_ext * ncols * 8);
    free(a);
    return 0;
}

/* ---- Poseidon: poseidon_goldilocks.cpp:5-37 (hash_full_result_seq), poseidon_goldilocks.hpp:91-162 ---- */
static inline uint64_t pow7(uint64_t x)
{
    uint64_t x2 = fmul(x, x), x3 = fmul(x, x2), x4 = fmul(x2, x2);
    return fmul(x3, x4);
}
/* new[i] = sum_j mat[j][i] * old[j]  (mvp_, poseidon_goldilocks.hpp:148-162) */
static void mvp(uint64_t *st, const uint64_t *mat)
{
    uint64_t old[12];
    memcpy(old, st, sizeof old);
    for (int i = 0; i < 12; i++) {
        uint64_t s = 0;
        for (int j = 0; j < 12; j++) s = fadd(s, fmul(mat[j * 12 + i], old[j]));
        st[i] = s;
    }
}
void orc_hash_full_result_seq(uint64_t *out, const uint64_t *in)
{
    uint64_t x[12];
    for (int i = 0; i < 12; i++) x[i] = fadd(in[i], ORC_C[i]);
    for (int r = 0; r < 3; r++) {
        for (int i = 0; i < 12; i++) x[i] = fadd(pow7(x[i]), ORC_C[12 * (r + 1) + i]);
        mvp(x, ORC_M);
    }
    for (int i = 0; i < 12; i++) x[i] = fadd(pow7(x[i]), ORC_C[48 + i]);
    mvp(x, ORC_P);
    for (int r = 0; r < 22; r++) {
        const uint64_t *S = ORC_S + 23 * r;
        x[0] = fadd(pow7(x[0]), ORC_C[60 + r]);
        uint64_t s0 = 0;
        for (int i = 0; i < 12; i++) s0 = fadd(s0, fmul(x[i], S[i]));
        for (int i = 1; i < 12; i++) x[i] = fadd(x[i], fmul(x[0], S[11 + i]));
        x[0] = s0;
    }
    for (int r = 0; r < 3; r++) {
        for (int i = 0; i < 12; i++) x[i] = fadd(pow7(x[i]), ORC_C[82 + 12 * r + i]);
        mvp(x, ORC_M);
    }
    for (int i = 0; i < 12; i++) x[i] = pow7(x[i]);
    mvp(x, ORC_M);
    memcpy(out, x, sizeof x);
}
void orc_hash_full_result(uint64_t *out, const uint64_t *in) { orc_hash_full_result_seq(out, in); }

/* linear_hash: poseidon_goldilocks.cpp:229-265 (sponge, rate 8, capacity 4; size<=4 is a raw copy) */
void orc_linear_hash(uint64_t *out4, const uint64_t *in, uint64_t size)
{
    if (size <= 4) {
        for (uint64_t i = 0; i < 4; i++) out4[i] = i < size ? fcanon(in[i]) : 0;
        return;
    }
    uint64_t st[12];
    uint64_t remaining = size;
    while (remaining) {
        if (remaining == size) memset(st + 8, 0, 4 * 8);
        else memcpy(st + 8, st, 4 * 8);
        uint64_t n = remaining < 8 ? remaining : 8;
        memset(st, 0, 8 * 8);
        memcpy(st, in + (size - remaining), n * 8);
        orc_hash_full_result_seq(st, st);
        remaining -= n;
    }
    memcpy(out4, st, 4 * 8);
}
void orc_linear_hash_rows(uint64_t *out4, const uint64_t *in, uint64_t ncols, uint64_t nrows)
{
#pragma omp parallel for
    for (uint64_t i = 0; i < nrows; i++) orc_linear_hash(out4 + 4 * i, in + i * ncols, ncols);
}
uint64_t orc_tree_num_elements(uint64_t degree) { return degree ? (2 * degree - 1) * 4 : 0; } /* merklehash_goldilocks.hpp:24-27 */

/* tree levels: poseidon_goldilocks.cpp:283-300.  level-major, root last. */
static void build_levels(uint64_t *tree, uint64_t num_rows)
{
    uint64_t pending = num_rows, next_index = 0;
    while (pending > 1) {
        uint64_t next_n = (pending - 1) / 2 + 1;
#pragma omp parallel for
        for (uint64_t i = 0; i < next_n; i++) {
            uint64_t in[12], out[12];
            memset(in, 0, sizeof in);
            memcpy(in, tree + next_index + i * 8, 8 * 8);
            orc_hash_full_result_seq(out, in);
            memcpy(tree + next_index + (pending + i) * 4, out, 4 * 8);
        }
        next_index += pending * 4;
        pending /= 2;
    }
}
/* merkletree: poseidon_goldilocks.cpp:266-301 */
void orc_merkletree(uint64_t *tree, const uint64_t *in, uint64_t ncols, uint64_t nrows, int nthreads, uint64_t dim)
{
    (void)nthreads;
    if (nrows == 0) return;
#pragma omp parallel for
    for (uint64_t i = 0; i < nrows; i++) orc_linear_hash(tree + 4 * i, in + i * ncols * dim, ncols * dim);
    build_levels(tree, nrows);
}
/* merkletree_batch: poseidon_goldilocks.cpp:113-164 (leaf = linear_hash of the per-batch linear_hashes) */
void orc_merkletree_batch(uint64_t *tree, const uint64_t *in, uint64_t ncols, uint64_t nrows, uint64_t batch, int nthreads, uint64_t dim)
{
    (void)nthreads;
    if (nrows == 0) return;
    uint64_t nb = ncols ? (ncols + batch - 1) / batch : 1;
    uint64_t nlast = ncols - (nb - 1) * batch;
#pragma omp parallel for
    for (uint64_t i = 0; i < nrows; i++) {
        uint64_t *buf = (uint64_t *)malloc(nb * 4 * 8);
        for (uint64_t j = 0; j < nb; j++)
            orc_linear_hash(buf + 4 * j, in + i * ncols * dim + j * batch * dim, (j == nb - 1 ? nlast : batch) * dim);
        orc_linear_hash(tree + 4 * i, buf, nb * 4);
        free(buf);
    }
    build_levels(tree, nrows);
}

/* extendPol then merkletree, as Starks::genProof step 1 (starks.cpp:53-57) */
int orc_commit(uint64_t *out_ext, uint64_t *tree, uint64_t *in, uint64_t n, uint64_t n_ext, uint64_t ncols, uint64_t *buffer, double *seconds)
{
    double t0 = omp_get_wtime();
    orc_extend_pol(out_ext, in, n_ext, n, ncols, buffer, 0, 0);
    double t1 = omp_get_wtime();
    orc_merkletree(tree, out_ext, ncols, n_ext, 0, 1);
    double t2 = omp_get_wtime();
    if (seconds) { seconds[0] = t1 - t0; seconds[1] = t2 - t1; }
    return 0;
}
