// K6/K7 device core -- the Poseidon-Goldilocks permutation (width 12 = rate 8 + capacity 4,
// 4 + 22 + 4 rounds, S-box x^7), state in registers, round constants in __constant__ memory.
//
// Computes exactly what the reference's PoseidonGoldilocks::hash_full_result_seq does
// (src/goldilocks/src/poseidon_goldilocks.cpp:5-37; helpers poseidon_goldilocks.hpp:91-162), in the
// "optimised" parametrisation the reference uses (pre-multiplied matrix P before the partial rounds,
// sparse vectors S inside them; poseidon_goldilocks_constants.hpp:931-2139).
//
// B200 notes.  One thread owns one 12-word state (24 registers).  The work is all 32-bit integer:
//  * S-box / dense P layer / partial rounds: 64x64 multiplies = 4 IMAD.WIDE.U32 each (gl_field.cuh).
//  * The 7 MDS layers use that M is circulant(17,15,41,16,2,28,13,13,39,18,34,20) + 8 at [0][0]: with the
//    state split into 32-bit halves every output is two 12-term IMAD.WIDE chains with immediate
//    coefficients (sum of coefficients 264 < 2^9, so a chain never overflows 64 bits) and ONE fold.
//  * Dot products (P layer, partial rounds) accumulate the four partial-product columns unreduced in
//    64-bit + carry-count accumulators and fold once per 12 terms instead of once per term.
#pragma once
#include <math.h>
#include "gl_field.cuh"
#include "poseidon_tables.cuh"

#ifndef GLP_CONSTANT
#define GLP_CONSTANT __constant__
#endif

// the host-thread emulator includes this header once per translation unit as well
GLP_CONSTANT gl_t POS_RC[GLP_POSEIDON_RC_N] = {GLP_POSEIDON_RC_INIT};
GLP_CONSTANT gl_t POS_PRE[144] = {GLP_POSEIDON_PRE_INIT};
GLP_CONSTANT gl_t POS_SP[GLP_POSEIDON_SP_N] = {GLP_POSEIDON_SP_INIT};
GLP_CONSTANT gl_t POS_RCF[96] = {GLP_POSEIDON_RCF_INIT};     // full-round constants, one row of 12 per round
GLP_CONSTANT gl_t POS_PRET[144] = {GLP_POSEIDON_PRET_INIT};  // pre-matrix, transposed: row i = weights of output i
GLP_CONSTANT gl_t POS_MRC[96] = {GLP_POSEIDON_MRC_INIT};     // round constants pushed through the following linear layer
GLP_CONSTANT uint32_t POS_MRCL[288] = {GLP_POSEIDON_MRCL_INIT};  // the same, as limbs of 22/22/20 bits
GLP_CONSTANT double POS_MRCD[192] = {GLP_POSEIDON_MRCD_INIT};    // the same, as 2^52 + 32-bit half (low, high)

#define POS_WIDTH 12
#define POS_RATE 8
#define POS_CAP 4

// ---- unreduced multiply-accumulate: acc += a*b over up to 2^8 terms, one fold at the end ----
struct GlDotAcc {
    uint32_t z0, z1, zc;   // sum of a0*b0            (weight 2^0)
    uint32_t x0, x1, xc;   // sum of a0*b1 + a1*b0    (weight 2^32)
    uint32_t y0, y1, yc;   // sum of a1*b1            (weight 2^64)
};

GL_HD void gl_dot_init(GlDotAcc &s) { s.z0 = s.z1 = s.zc = s.x0 = s.x1 = s.xc = s.y0 = s.y1 = s.yc = 0; }

GL_HD void gl_dot_mac(GlDotAcc &s, gl_t a, gl_t b)
{
#if GL_DEVICE_PTX
    asm("{\n\t"
        ".reg .u32 a0,a1,b0,b1;\n\t"
        "mov.b64 {a0,a1}, %9;\n\t"
        "mov.b64 {b0,b1}, %10;\n\t"
        "mad.lo.cc.u32 %0, a0, b0, %0;\n\t"
        "madc.hi.cc.u32 %1, a0, b0, %1;\n\t"
        "addc.u32 %2, %2, 0;\n\t"
        "mad.lo.cc.u32 %3, a0, b1, %3;\n\t"
        "madc.hi.cc.u32 %4, a0, b1, %4;\n\t"
        "addc.u32 %5, %5, 0;\n\t"
        "mad.lo.cc.u32 %3, a1, b0, %3;\n\t"
        "madc.hi.cc.u32 %4, a1, b0, %4;\n\t"
        "addc.u32 %5, %5, 0;\n\t"
        "mad.lo.cc.u32 %6, a1, b1, %6;\n\t"
        "madc.hi.cc.u32 %7, a1, b1, %7;\n\t"
        "addc.u32 %8, %8, 0;\n\t"
        "}"
        : "+r"(s.z0), "+r"(s.z1), "+r"(s.zc), "+r"(s.x0), "+r"(s.x1), "+r"(s.xc), "+r"(s.y0), "+r"(s.y1), "+r"(s.yc)
        : "l"(a), "l"(b));
#else
    uint64_t a0 = (uint32_t)a, a1 = a >> 32, b0 = (uint32_t)b, b1 = b >> 32;
    uint64_t z = ((uint64_t)s.z1 << 32) | s.z0, x = ((uint64_t)s.x1 << 32) | s.x0, y = ((uint64_t)s.y1 << 32) | s.y0;
    uint64_t t;
    t = z + a0 * b0; s.zc += (t < z); z = t;
    t = x + a0 * b1; s.xc += (t < x); x = t;
    t = x + a1 * b0; s.xc += (t < x); x = t;
    t = y + a1 * b1; s.yc += (t < y); y = t;
    s.z0 = (uint32_t)z; s.z1 = (uint32_t)(z >> 32);
    s.x0 = (uint32_t)x; s.x1 = (uint32_t)(x >> 32);
    s.y0 = (uint32_t)y; s.y1 = (uint32_t)(y >> 32);
#endif
}

// acc += a * w for a 32-bit w: two of the four partial products (the chunked partial rounds' cross weights, POS_PR_CHUNK)
GL_HD void gl_dot_mac_small(GlDotAcc &s, gl_t a, uint32_t w)
{
#if GL_DEVICE_PTX
    asm("{\n\t"
        ".reg .u32 a0,a1;\n\t"
        "mov.b64 {a0,a1}, %6;\n\t"
        "mad.lo.cc.u32 %0, a0, %7, %0;\n\t"
        "madc.hi.cc.u32 %1, a0, %7, %1;\n\t"
        "addc.u32 %2, %2, 0;\n\t"
        "mad.lo.cc.u32 %3, a1, %7, %3;\n\t"
        "madc.hi.cc.u32 %4, a1, %7, %4;\n\t"
        "addc.u32 %5, %5, 0;\n\t"
        "}"
        : "+r"(s.z0), "+r"(s.z1), "+r"(s.zc), "+r"(s.x0), "+r"(s.x1), "+r"(s.xc)
        : "l"(a), "r"(w));
#else
    gl_dot_mac(s, a, (gl_t)w);
#endif
}

#ifndef POS_EPS_ALU_FOLD
#define POS_EPS_ALU_FOLD 0
#endif
#ifndef POS_EPS_ALU_MULADD
#define POS_EPS_ALU_MULADD 0
#endif
#ifndef POS_RED_MODE
#define POS_RED_MODE 0            // fold of the S-box multiplies: 0 IMAD.LO+IMAD.HI, 1 ALU, 2 one IMAD.WIDE (gl_reduce128_m)
#endif
#ifndef POS_MUL4W
#define POS_MUL4W 0               // 1: S-box / rank-1 products as four IMAD.WIDE (gl_mul_wide4)
#endif
#ifndef POS_MULADD_MODE
#define POS_MULADD_MODE (POS_EPS_ALU_MULADD ? 1 : 0)
#endif
#ifndef POS_FOLD_MODE
#define POS_FOLD_MODE (POS_EPS_ALU_FOLD ? 1 : 0)
#endif
// DEFER (the leaf hash, POS_DEFER_BORROW): every fold leaves its 2^-33 wrap uncorrected and counts it in bf instead
// (gl_reduce128_defer); the caller redoes the row with DEFER = false when bf != 0 at the end.
template <bool DEFER>
GL_HD gl_t pos_pow7(gl_t x, GlBorrow &bf)
{
    if (DEFER) return gl_pow7_defer(x, bf);
    return gl_pow7_m<POS_RED_MODE, POS_MUL4W != 0>(x);
}
template <bool DEFER, int MODE>
GL_HD gl_t pos_fold128(gl_t lo, gl_t hi, GlBorrow &bf)
{
    if (DEFER) return gl_reduce128_defer_t<false>(lo, hi, bf);   // dot-product accumulators: a general 128-bit value
    return gl_reduce128_m<MODE>(lo, hi);
}

#ifndef POS_EPS_ALU_FOLD
#define POS_EPS_ALU_FOLD 0        // 1: the p2*EPS term of the dot-product folds on the ALU pipe (experiment)
#endif
#ifndef POS_EPS_ALU_MULADD
#define POS_EPS_ALU_MULADD 0      // 1: same for the rank-1 update multiplies of the partial rounds (experiment)
#endif
// value = Z + X*2^32 + Y*2^64  ->  limbs l0..l4 (weights 2^0..2^128), then
//   (l1:l0) + l2*EPS - l3 - l4*2^32   (2^64 = EPS, 2^96 = -1, 2^128 = -2^32)
#ifndef POS_DEFER_DOTSUB
#define POS_DEFER_DOTSUB 1        // deferred rows: the dot-product folds' last subtraction leaves its (2^-28) wrap to the redo as well
#endif
#ifndef POS_MDS_FOLD_FMA
#define POS_MDS_FOLD_FMA 2        // top*EPS of the MDS fold: 1 = IMAD + IMAD.HI (6 pipe cycles), 2 = one IMAD.WIDE (EPS from the constant bank), 0 = four ALU instructions (8)
#endif                            // (each +0.5 % on the leaf hash, profiles/r01_poseidon_defer.txt)
// r - b for b < 2^40 with the (rare: r < b) wrap left out and counted in bf, like gl_reduce128_defer
GL_HD gl_t gl_sub_small_defer(gl_t r, gl_t b, GlBorrow &bf)
{
#if GL_DEVICE_PTX
    gl_t o;
    asm("{\n\t"
        ".reg .u32 a0,a1,b0,b1;\n\t"
        "mov.b64 {a0,a1}, %2;\n\t"
        "mov.b64 {b0,b1}, %3;\n\t"
        "sub.cc.u32 a0, a0, b0;\n\t"
        "subc.cc.u32 a1, a1, b1;\n\t"
        "subc.u32 %1, %1, 0;\n\t"
        "mov.b64 %0, {a0,a1};\n\t"
        "}" : "=l"(o), "+r"(bf.cnt) : "l"(r), "l"(b));
    return o;
#else
    if (r < b) bf.cnt--;                                          // modelled like the device: counted, value left unfixed
    return r - b;
#endif
}
template <bool DEFER>
GL_HD gl_t gl_dot_fold(const GlDotAcc &s, GlBorrow &bf)
{
    uint64_t t1 = (uint64_t)s.z1 + s.x0;
    uint64_t t2 = (uint64_t)s.zc + s.x1 + s.y0 + (t1 >> 32);
    uint64_t t3 = (uint64_t)s.xc + s.y1 + (t2 >> 32);
    uint64_t l4 = (uint64_t)s.yc + (t3 >> 32);
    gl_t lo = ((t1 & 0xFFFFFFFFULL) << 32) | s.z0;
    gl_t hi = ((t3 & 0xFFFFFFFFULL) << 32) | (t2 & 0xFFFFFFFFULL);
    gl_t r = pos_fold128<DEFER, POS_FOLD_MODE>(lo, hi, bf);
#if POS_DEFER_DOTSUB
    if (DEFER) return gl_sub_small_defer(r, l4 << 32, bf);      // l4 < 2^8: borrows with probability < 2^-24, never twice
#endif
    return gl_sub(r, l4 << 32);
}

#ifndef POS_MULADD_FUSED
#define POS_MULADD_FUSED 0        // 1: gl_mul_add puts c into the accumulator of the a0*b0 IMAD.WIDE
#endif
#ifndef POS_MULADD_V2
#define POS_MULADD_V2 0           // 1: the rank-1 updates of the partial rounds fold with form 2 as well (ptxas then runs out of carry predicates
#endif                            //    in the partial-round loop and packs them into registers: 548 instead of 388 instructions a round)
// a*b + c, one fold (c is added into the 128-bit product: (2^64-1)^2 + 2^64 - 1 < 2^128)
template <bool DEFER>
GL_HD gl_t gl_mul_add(gl_t a, gl_t b, gl_t c, GlBorrow &bf)
{
    gl_t lo, hi;
#if GL_DEVICE_PTX && POS_MULADD_FUSED
    // c rides in the accumulator of the a0*b0 product (one IMAD.WIDE with carry-out) instead of a 128-bit add afterwards
    asm("{\n\t"
        ".reg .u32 a0,a1,b0,b1,c0,c1,x0,x1,s0,s1,s2,y0,y1,z0,z1;\n\t"
        ".reg .u64 X,Y;\n\t"
        "mov.b64 {a0,a1}, %2;\n\t"
        "mov.b64 {b0,b1}, %3;\n\t"
        "mov.b64 {c0,c1}, %4;\n\t"
        "mul.wide.u32 Y, a1, b1;\n\t"
        "mov.b64 {y0,y1}, Y;\n\t"
        "mad.lo.cc.u32 z0, a0, b0, c0;\n\t"
        "madc.hi.cc.u32 z1, a0, b0, c1;\n\t"
        "addc.cc.u32 y0, y0, 0;\n\t"
        "addc.u32 y1, y1, 0;\n\t"
        "mul.wide.u32 X, a0, b1;\n\t"
        "mov.b64 {x0,x1}, X;\n\t"
        "mad.lo.cc.u32 s0, a1, b0, x0;\n\t"
        "madc.hi.cc.u32 s1, a1, b0, x1;\n\t"
        "addc.u32 s2, 0, 0;\n\t"
        "add.cc.u32 z1, z1, s0;\n\t"
        "addc.cc.u32 y0, y0, s1;\n\t"
        "addc.u32 y1, y1, s2;\n\t"
        "mov.b64 %0, {z0,z1};\n\t"
        "mov.b64 %1, {y0,y1};\n\t"
        "}" : "=l"(lo), "=l"(hi) : "l"(a), "l"(b), "l"(c));
    if (DEFER) return gl_reduce128_defer_t<POS_MULADD_V2 != 0>(lo, hi, bf);
    return gl_reduce128_m<POS_MULADD_MODE>(lo, hi);
#endif
    if (POS_MUL4W) gl_mul_wide4(a, b, lo, hi);
    else gl_mul_wide(a, b, lo, hi);
#if GL_DEVICE_PTX
    asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, 0;" : "+l"(lo), "+l"(hi) : "l"(c));
#else
    gl_t t = lo + c;
    hi += (t < lo);
    lo = t;
#endif
    if (DEFER) return gl_reduce128_defer_t<POS_MULADD_V2 != 0>(lo, hi, bf);    // a*b + c with b a canonical constant (see gl_fold_defer_raw)
    return gl_reduce128_m<POS_MULADD_MODE>(lo, hi);
}
// two rank-1 updates whose deferred folds share one flag instruction (GL_BORROW_MIN): xa = a*ba + xa, xb = a*bb + xb
GL_HD void gl_mul_add2_defer(gl_t a, gl_t ba, gl_t bb, gl_t &xa, gl_t &xb, GlBorrow &bf)
{
#if GL_BORROW_MIN
    gl_t lo_a, hi_a, lo_b, hi_b;
    gl_mul_wide(a, ba, lo_a, hi_a);
    gl_mul_wide(a, bb, lo_b, hi_b);
#if GL_DEVICE_PTX
    asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, 0;" : "+l"(lo_a), "+l"(hi_a) : "l"(xa));
    asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, 0;" : "+l"(lo_b), "+l"(hi_b) : "l"(xb));
#else
    gl_t t = lo_a + xa; hi_a += (t < lo_a); lo_a = t;
    t = lo_b + xb; hi_b += (t < lo_b); lo_b = t;
#endif
    uint32_t wa, wb;
    xa = gl_fold_defer_raw<POS_MULADD_V2 != 0>(lo_a, hi_a, wa);
    xb = gl_fold_defer_raw<POS_MULADD_V2 != 0>(lo_b, hi_b, wb);
    gl_borrow_note2(bf, wa, wb);
#else
    xa = gl_mul_add<true>(a, ba, xa, bf);
    xb = gl_mul_add<true>(a, bb, xb, bf);
#endif
}

// ---- MDS layer: state' = M^T state,  M[j][i] = circ[(j-i) mod 12] (+8 at i=j=0) ----
// (mvp_ with PoseidonGoldilocksConstants::M, poseidon_goldilocks.hpp:148-162)
GL_HD gl_t pos_mds_fold(uint64_t L, uint64_t H)
{
    // L + H*2^32 with L,H < 2^42:  lo64 = L + (H<<32), top = (H>>32) + carry (< 2^11);  lo64 + top*EPS
#if GL_DEVICE_PTX
    gl_t r;
    asm("{\n\t"
        ".reg .u32 l0,l1,h0,h1,m;\n\t"
        "mov.b64 {l0,l1}, %1;\n\t"
        "mov.b64 {h0,h1}, %2;\n\t"
        "add.cc.u32 l1, l1, h0;\n\t"
        "addc.u32 h1, h1, 0;\n\t"
#if POS_MDS_FOLD_FMA == 2
        "mad.lo.cc.u32 l0, h1, %3, l0;\n\t"            // EPS from the constant bank: one IMAD.WIDE with carry-out
        "madc.hi.cc.u32 l1, h1, %3, l1;\n\t"
#elif POS_MDS_FOLD_FMA
        GL_PTX_ADD_EPS_MUL_FMA("l0", "l1", "h1")
#else
        GL_PTX_ADD_EPS_MUL_ALU("l0", "l1", "h1")
#endif
        "addc.u32 m, 0, 0;\n\t"
        "sub.cc.u32 l0, l0, m;\n\t"
        "subc.u32 l1, l1, 0;\n\t"
        "add.u32 l1, l1, m;\n\t"
        "mov.b64 %0, {l0,l1};\n\t"
        "}" : "=l"(r) : "l"(L), "l"(H), "r"(GL_EPS_OPAQUE));
    return r;
#else
    uint64_t lo = L + (H << 32);
    uint64_t top = (H >> 32) + (lo < L);
    return gl_reduce128(lo, top);
#endif
}

// acc(64) += a(32) * c(32): written as a carry-chain pair so that ptxas keeps ONE IMAD.WIDE with the 64-bit
// addend (a plain `L += (uint64_t)a * c` is strength-reduced to shifts/adds for small c and its accumulate
// split into IADD3 + IADD3.X, which moves the work onto the ALU pipe)
GL_HD void pos_macw(uint32_t &l0, uint32_t &l1, uint32_t a, uint32_t c)
{
#if GL_DEVICE_PTX
    asm("mad.lo.cc.u32 %0, %2, %3, %0;\n\tmadc.hi.u32 %1, %2, %3, %1;" : "+r"(l0), "+r"(l1) : "r"(a), "r"(c));
#else
    uint64_t t = (((uint64_t)l1 << 32) | l0) + (uint64_t)a * c;
    l0 = (uint32_t)t;
    l1 = (uint32_t)(t >> 32);
#endif
}

GL_HD void pos_mds(gl_t x[12])
{
    const uint32_t circ[12] = {GLP_MDS_CIRC_INIT};
    uint32_t lo[12], hi[12];
#pragma unroll
    for (int j = 0; j < 12; j++) { lo[j] = (uint32_t)x[j]; hi[j] = (uint32_t)(x[j] >> 32); }
#pragma unroll
    for (int i = 0; i < 12; i++) {
        uint32_t L0 = 0, L1 = 0, H0 = 0, H1 = 0;
#pragma unroll
        for (int j = 0; j < 12; j++) {
            const uint32_t c = circ[(j - i + 12) % 12] + ((i == 0 && j == 0) ? GLP_MDS_DIAG0 : 0);
            pos_macw(L0, L1, lo[j], c);
            pos_macw(H0, H1, hi[j], c);
        }
        x[i] = pos_mds_fold(((uint64_t)L1 << 32) | L0, ((uint64_t)H1 << 32) | H0);
    }
}

// ---- MDS layer in limb form, round constants folded in: x <- M^T x + M^T rc ----
// IMAD.WIDE occupies the fmaheavy pipe for 4 cycles per warp instruction, a 32-bit IMAD or an ALU op (LEA, shift-add)
// for 2 (profiles/r01_int_pipes.txt), and the IMAD.WIDE form of this layer (pos_mds) made the kernel fmaheavy-bound
// with the ALU pipe half idle.  Here every state word is cut into limbs of 22, 22 and 20 bits; limb * coefficient
// summed over the 12 inputs stays below 2^31 (sum of a row's coefficients is 280), so each output is three chains of
// 32-bit multiply-adds -- which ptxas is free to issue as IMAD (fma pipe) or, for coefficients like 16, 2, 17, as
// LEA / shift-add (ALU pipe) -- started from the limbs of the folded round constant, then one recombination and fold.
#ifndef POS_MDS_LIMB
#define POS_MDS_LIMB 2          // 0: IMAD.WIDE chains (pos_mds); 1: 22-bit limbs, 32-bit IMAD (pos_mds_limb); 2: FP64 pipe (pos_mds_f64)
#endif
GL_HD gl_t pos_mds_combine(uint32_t a0, uint32_t a1, uint32_t a2)
{
    // a0 + a1*2^22 + a2*2^44 (< 2^74)  =  lo64 + top*2^64,  then  lo64 + top*EPS  (top < 2^10)
    uint64_t lo = (uint64_t)a0 + ((uint64_t)a1 << 22);                 // < 2^54
    const uint64_t hi = (uint64_t)(a2 & 0xFFFFFu) << 44;
    uint32_t top = a2 >> 20;
    lo += hi;
    top += (lo < hi);
    const uint64_t t = ((uint64_t)top << 32) - top;                     // top * EPS < 2^42
    uint64_t r = lo + t;
    if (r < t) r += GL_EPS;                                             // wrapped: r < 2^42, cannot wrap again
    return r;
}
GL_HD void pos_mds_limb(gl_t x[12], const uint32_t *rcl)
{
    const uint32_t circ[12] = {GLP_MDS_CIRC_INIT};
    uint32_t l0[12], l1[12], l2[12];
#pragma unroll
    for (int j = 0; j < 12; j++) {
        const uint32_t lo = (uint32_t)x[j], hi = (uint32_t)(x[j] >> 32);
        l0[j] = lo & 0x3FFFFFu;
        l1[j] = ((lo >> 22) | (hi << 10)) & 0x3FFFFFu;
        l2[j] = hi >> 12;
    }
#pragma unroll
    for (int i = 0; i < 12; i++) {
        uint32_t a0 = rcl[3 * i], a1 = rcl[3 * i + 1], a2 = rcl[3 * i + 2];
#pragma unroll
        for (int j = 0; j < 12; j++) {
            const uint32_t c = circ[(j - i + 12) % 12] + ((i == 0 && j == 0) ? GLP_MDS_DIAG0 : 0);
            a0 += l0[j] * c;
            a1 += l1[j] * c;
            a2 += l2[j] * c;
        }
        x[i] = pos_mds_combine(a0, a1, a2);
    }
}

// ---- MDS layer on the FP64 pipe, round constants folded in: x <- M^T x + M^T rc ----
// The B200 issues DFMA at 64 lanes/clk/SM -- twice the rate of IMAD.WIDE (profiles/r01_int_pipes.txt: DFMA 63, IMAD.WIDE
// 32 lanes/clk/SM) -- and a double holds integers below 2^53 exactly.  The IMAD.WIDE form of this layer (pos_mds) needs
// exactly that range: sum_j c_ij * (32-bit half of x_j) < 2^41.  So each 32-bit half is made a double once (the
// 0x43300000 exponent trick + one DADD, no I2F), every output half is a 12-term DFMA chain started from
// 2^52 + (half of the folded round constant), and the integer sum is read back from the mantissa: same arithmetic,
// half the pipe time, and the ALU pipe only sees the fold.  Every intermediate is an integer < 2^53, so the result is
// exact whatever the rounding mode; the host emulator runs the same expressions in IEEE double.
#ifndef POS_CVT_I2F
#define POS_CVT_I2F 1             // 1 = u32 -> double with the conversion instruction instead of the exponent trick + DADD
#endif
GL_HD double pos_u32_to_f64(uint32_t v)
{
#if GL_DEVICE_PTX && POS_CVT_I2F
    return __uint2double_rn(v);                                              // I2F.F64.U32
#elif GL_DEVICE_PTX
    return __hiloint2double(0x43300000, (int)v) - 4503599627370496.0;        // (2^52 + v) - 2^52
#else
    return (double)v;
#endif
}
GL_HD uint64_t pos_f64_mantissa(double d)                                        // d = 2^52 + k, k < 2^52  ->  k
{
#if GL_DEVICE_PTX
    return (uint64_t)__double_as_longlong(d) & 0x000FFFFFFFFFFFFFULL;
#else
    return (uint64_t)(d - 4503599627370496.0);
#endif
}
GL_HD void pos_mds_f64(gl_t x[12], const double *rcd)
{
    const uint32_t circ[12] = {GLP_MDS_CIRC_INIT};
    uint64_t L[12];
    {
        double d[12];
#pragma unroll
        for (int j = 0; j < 12; j++) d[j] = pos_u32_to_f64((uint32_t)x[j]);
#pragma unroll
        for (int i = 0; i < 12; i++) {
            double acc = rcd[2 * i];
#pragma unroll
            for (int j = 0; j < 12; j++) acc = fma(d[j], (double)(circ[(j - i + 12) % 12] + ((i == 0 && j == 0) ? GLP_MDS_DIAG0 : 0)), acc);
            L[i] = pos_f64_mantissa(acc);
        }
    }
    double d[12];
#pragma unroll
    for (int j = 0; j < 12; j++) d[j] = pos_u32_to_f64((uint32_t)(x[j] >> 32));
#pragma unroll
    for (int i = 0; i < 12; i++) {
        double acc = rcd[2 * i + 1];
#pragma unroll
        for (int j = 0; j < 12; j++) acc = fma(d[j], (double)(circ[(j - i + 12) % 12] + ((i == 0 && j == 0) ? GLP_MDS_DIAG0 : 0)), acc);
        x[i] = pos_mds_fold(L[i], pos_f64_mantissa(acc));
    }
}

// ---- the permutation ----
// Code layout matters as much as the arithmetic here: a fully unrolled permutation is ~125 KB of SASS, four
// times the instruction cache, and with 16 warps of an SM at 16 different places in it the kernel stalls on
// instruction fetch (ncu: stall_no_instruction 6 of 11 cycles per issue).  So the eight full rounds are ONE
// loop body (S-box group + MDS), the dense pre-matrix is a 12-trip loop that shifts its results through a
// register queue, and the 22 partial rounds are a loop: ~25 KB of hot code.  POS_LOCKSTEP additionally
// makes the warps of a CTA meet at a barrier every round so that they fetch the same lines together.
#ifndef POS_SBOX_GROUP
#define POS_SBOX_GROUP 12         // S-boxes per trip of the S-box loop (12 = unrolled; 4 or 6 = rolled with a rotating state: smaller code, measured 2-5 % slower)
#endif
#ifndef POS_ROLL_PRE
#define POS_ROLL_PRE 1            // 1 = 12-trip pre-matrix loop with a register queue: 3 KB instead of 30 KB of code.  With the FP64 MDS layer the
                                  // unrolled form pushes the kernel past the 32 KB L1.5 instruction cache and costs 4.5 % (profiles/r01_poseidon_variants.txt)
#endif

#ifndef POS_PRE_UNROLL
#define POS_PRE_UNROLL 3          // outputs per trip of the rolled pre-matrix loop (1: 22 queue moves per output; 3: 6 per output, +4.5 KB of code)
#endif

template <bool LOCKSTEP>
GL_HD void pos_round_sync()
{
#if GL_DEVICE_PTX
    if (LOCKSTEP) __syncthreads();
#endif
}

// x_i = x_i^7 + rc_i for all 12, as 12/G trips over G lanes with the state rotated by G after each trip
// (rc == NULL: no constant here, it was folded into the linear layer that follows -- POS_MDS_LIMB)
template <bool DEFER>
GL_HD void pos_sbox_layer(gl_t x[12], const gl_t *rc, GlBorrow &bf)
{
    constexpr int G = POS_SBOX_GROUP;
    static_assert(12 % G == 0, "POS_SBOX_GROUP must divide 12");
    if (!rc) {
#pragma unroll
        for (int i = 0; i < 12; i++) x[i] = pos_pow7<DEFER>(x[i], bf);
        return;
    }
    if (G == 12) {
#pragma unroll
        for (int i = 0; i < 12; i++) x[i] = gl_add_c(pos_pow7<DEFER>(x[i], bf), rc[i]);
        return;
    }
#pragma unroll 1
    for (int q = 0; q < 12 / G; q++) {
        gl_t t[G];
#pragma unroll
        for (int i = 0; i < G; i++) t[i] = gl_add_c(pos_pow7<DEFER>(x[i], bf), rc[q * G + i]);
#pragma unroll
        for (int i = 0; i + G < 12; i++) x[i] = x[i + G];
#pragma unroll
        for (int i = 0; i < G; i++) x[12 - G + i] = t[i];
    }
}

// x <- PRE^T x  (the dense pre-multiplied matrix before the partial rounds)
template <bool DEFER>
GL_HD void pos_pre_layer(gl_t x[12], GlBorrow &bf)
{
#if POS_ROLL_PRE
    gl_t out[12];
#pragma unroll
    for (int i = 0; i < 12; i++) out[i] = 0;
    constexpr int U = POS_PRE_UNROLL;                    // outputs per trip: the 12-word register queue shifts by U per trip
    static_assert(12 % U == 0, "POS_PRE_UNROLL must divide 12");
#pragma unroll 1
    for (int i = 0; i < 12; i += U) {
        gl_t t[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const gl_t *col = POS_PRET + 12 * (i + u);
            GlDotAcc acc;
            gl_dot_init(acc);
#if POS_MDS_LIMB
            acc.z0 = (uint32_t)POS_MRC[36 + i + u];
            acc.z1 = (uint32_t)(POS_MRC[36 + i + u] >> 32);
#endif
#pragma unroll
            for (int j = 0; j < 12; j++) gl_dot_mac(acc, x[j], col[j]);
            t[u] = gl_dot_fold<DEFER>(acc, bf);
        }
#pragma unroll
        for (int k = 0; k + U < 12; k++) out[k] = out[k + U];
#pragma unroll
        for (int u = 0; u < U; u++) out[12 - U + u] = t[u];
    }
#pragma unroll
    for (int i = 0; i < 12; i++) x[i] = out[i];
#else
    gl_t y[12];
#pragma unroll
    for (int i = 0; i < 12; i++) y[i] = x[i];
#pragma unroll
    for (int i = 0; i < 12; i++) {
        GlDotAcc acc;
        gl_dot_init(acc);
#if POS_MDS_LIMB
        acc.z0 = (uint32_t)POS_MRC[36 + i];              // + PRE^T rc_3 (see POS_MRC)
        acc.z1 = (uint32_t)(POS_MRC[36 + i] >> 32);
#endif
#pragma unroll
        for (int j = 0; j < 12; j++) gl_dot_mac(acc, y[j], POS_PRE[j * 12 + i]);
        x[i] = gl_dot_fold<DEFER>(acc, bf);
    }
#endif
}

// ---- the partial rounds, M at a time ----
// One round is  y = x0^7 + c;  s = 25 y + sum_{i>=1} x_i S_i;  x_i += y v_i (i >= 1);  x0 = s  (poseidon_goldilocks.cpp:19-32).  The
// eleven rank-1 updates cost a full multiply AND a fold each.  Taken M rounds at a time they become one M-term dot product per
// word with ONE fold (x_i + sum_k y_k v_i(k)), and the dot product of round k makes up for the updates it has not seen with
// y_j * W_(k-j) for every earlier round j of the chunk -- where W_d = sum_i v_i(j) S_i(j+d) is the same for every j and SMALL
// (13, 21, 29 bits for d = 1, 2, 3: two partial products instead of four), as is the diagonal weight 25; both facts are asserted
// by tools/gen_poseidon_tables.py, which also checks this formulation against the plain one on the reference's known answers.
// Per round: 22 + (M-1)/2 multiply-accumulates and 1 + 11/M folds instead of 12 multiply-accumulates, 11 multiplies and 12 folds.
#ifndef POS_PR_NOTAIL
#define POS_PR_NOTAIL 0           // 1 = pad at the front instead of a tail chunk: measured slower (the padded rounds' idle update loop costs
#endif                            // more than the tail's 6 KB of code: 1135 against 1177 M perm/s, profiles/r02_pos_variants.txt)
#ifndef POS_PR_CHUNK
#define POS_PR_CHUNK 5            // 1 = round by round (round 1's form); M > 1: chunks of M rounds, the remaining 22 % M as one shorter chunk.
#endif                            // Measured (M perm/s, profiles/r02_pos_variants.txt).  The chunked forms need the warps of a CTA IN PHASE: they are
                                  // 40-55 KB of hot code, and on long rows (660 columns = 83 permutations) free-running warps spread over the whole
                                  // loop and stop sharing instruction and constant fetches -- M = 4 falls from 1177 (64-column rows, where a CTA's
                                  // warps start together) to 1061, below M = 1's 1135.  With one __syncthreads() per permutation in the leaf kernel
                                  // (POS_LEAF_SYNC, poseidon_kernels.cuh) they are width-independent: M = 1 1136, 3 1143, 4 1176, 5 1186, 6 1182-1193.
GLP_CONSTANT gl_t POS_SPW[8] = {GLP_POSEIDON_SPW_INIT, 0};                 // W_d, d = 1..7; the first three fit 32 bits
// Where state words 1..11 live during the partial rounds (POS_PR_CHUNK > 1): in the chunked form they are only read by the dot
// products and rewritten once per chunk, so they sit in shared memory ([word][thread], conflict-free) and the end-of-chunk
// update is a ROLLED loop over the word index -- one copy of its code instead of eleven (the hot loop stays inside the
// instruction cache) and four carry predicates in flight instead of forty-four (with the update unrolled ptxas packs them into
// a register bit mask: 220 LOP3 per chunk).  22 registers are free meanwhile.  The emulator hands in a thread-private array.
struct PosXs {
    gl_t *p;
    uint32_t stride;
};
GL_HD gl_t &pos_xs_at(const PosXs &xs, int i) { return xs.p[(size_t)i * xs.stride]; }
#if defined(GLP_EMU) || !defined(__CUDACC__)
#define POS_XS_DECL(block) gl_t pos_xs_buf[11]; const PosXs xs = {pos_xs_buf, 1}
#else
#define POS_XS_DECL(block) __shared__ gl_t pos_xs_buf[11 * (block)]; const PosXs xs = {pos_xs_buf + threadIdx.x, (uint32_t)(block)}
#endif
// Rounds p .. p+M-1; rounds with a negative index do nothing (POS_PR_NOTAIL: the 22 rounds are padded at the FRONT to a whole
// number of chunks instead of a second, shorter copy of this code behind the loop -- the hot loop has to fit the instruction
// cache, and 6 KB of tail pushed M = 4 to its edge; the price is 11 * 6 idle multiply-accumulate instructions per padded round).
template <bool DEFER, int M>
GL_HD void pos_partial_chunk(gl_t &x0, const PosXs &xs, const int p, GlBorrow &bf, const int m = M)
{
    static_assert(M >= 1 && M <= 8, "cross weights W_d are tabulated for d <= 7");
    gl_t y[M];
#pragma unroll
    for (int k = 0; k < M; k++) {
        y[k] = 0;
        if (p + k >= 0 && k < m) {                                // uniform; false only for padding rounds (POS_PR_NOTAIL 1: front, 2: back)
            const gl_t *S = POS_SP + 23 * (p + k);
            y[k] = gl_add_c(pos_pow7<DEFER>(x0, bf), POS_RC[60 + p + k]);
            GlDotAcc acc;
            gl_dot_init(acc);
            gl_dot_mac_small(acc, y[k], GLP_POSEIDON_SP_DIAG);
#pragma unroll
            for (int i = 1; i < 12; i++) gl_dot_mac(acc, pos_xs_at(xs, i - 1), S[i]);
#pragma unroll
            for (int j = 0; j < k; j++) {
                if (k - j <= 3) gl_dot_mac_small(acc, y[j], (uint32_t)POS_SPW[k - j - 1]);
                else gl_dot_mac(acc, y[j], POS_SPW[k - j - 1]);
            }
            x0 = gl_dot_fold<DEFER>(acc, bf);
        }
    }
    const gl_t *S0 = POS_SP + 23 * (p < 0 ? 0 : p);              // padded rounds multiply y = 0 by whatever round 0's weights are
    const int koff = p < 0 ? p : 0;
#pragma unroll 1
    for (int i = 1; i < 12; i++) {
        GlDotAcc acc;
        gl_dot_init(acc);
        const gl_t xi = pos_xs_at(xs, i - 1);
        acc.z0 = (uint32_t)xi;
        acc.z1 = (uint32_t)(xi >> 32);
#pragma unroll
        for (int k = 0; k < M; k++)
            if (k < m) gl_dot_mac(acc, y[k], S0[23 * (k + koff < 0 ? 0 : k + koff) + 11 + i]);   // uniform branch; k >= m only in a short last chunk
        pos_xs_at(xs, i - 1) = gl_dot_fold<DEFER>(acc, bf);
    }
}

template <bool LOCKSTEP, bool DEFER>
GL_HD void pos_permute_t(gl_t x[12], GlBorrow &bf, const PosXs &xs)
{
#pragma unroll
    for (int i = 0; i < 12; i++) x[i] = gl_add_c(x[i], POS_RC[i]);

    // 8 full rounds {x^7 + C ; linear layer}; the linear layer of round 3 is the pre-matrix followed by the
    // 22 partial rounds, every other one is the MDS matrix; round 7 adds zero constants
#pragma unroll 1
    for (int r = 0; r < 8; r++) {
        pos_round_sync<LOCKSTEP>();
#if POS_MDS_LIMB
        pos_sbox_layer<DEFER>(x, nullptr, bf);                      // constants ride in the linear layer's accumulators
        if (r != 3) {
#if POS_MDS_LIMB == 2
            pos_mds_f64(x, POS_MRCD + 24 * r);
#else
            pos_mds_limb(x, POS_MRCL + 36 * r);
#endif
        } else {
#else
        pos_sbox_layer<DEFER>(x, POS_RCF + 12 * r, bf);
        if (r != 3) {
            pos_mds(x);
        } else {
#endif
            pos_pre_layer<DEFER>(x, bf);
            // partial rounds: x0 = x0^7 + C;  s0 = <x, S[0..12)>;  x_i += x0 * S[11+i] (i >= 1);  x0 = s0
#if POS_PR_CHUNK > 1
            {
                constexpr int M = POS_PR_CHUNK, TAIL = 22 % POS_PR_CHUNK;
#pragma unroll
                for (int i = 1; i < 12; i++) pos_xs_at(xs, i - 1) = x[i];
#if POS_PR_NOTAIL == 2
#pragma unroll 1
                for (int p = 0; p < 22; p += M) {                  // the last chunk runs the same code with 22 - p < M active rounds
                    pos_round_sync<LOCKSTEP>();
                    pos_partial_chunk<DEFER, M>(x[0], xs, p, bf, 22 - p < M ? 22 - p : M);
                }
#elif POS_PR_NOTAIL
#pragma unroll 1
                for (int p = TAIL ? TAIL - M : 0; p < 22; p += M) {
                    pos_round_sync<LOCKSTEP>();
                    pos_partial_chunk<DEFER, M>(x[0], xs, p, bf);
                }
#else
#pragma unroll 1
                for (int p = 0; p + M <= 22; p += M) {
                    pos_round_sync<LOCKSTEP>();
                    pos_partial_chunk<DEFER, M>(x[0], xs, p, bf);
                }
                if (TAIL) pos_partial_chunk<DEFER, (TAIL ? TAIL : 1)>(x[0], xs, 22 - TAIL, bf);
#endif
#pragma unroll
                for (int i = 1; i < 12; i++) x[i] = pos_xs_at(xs, i - 1);
            }
#else
#pragma unroll 1
            for (int p = 0; p < 22; p++) {
                pos_round_sync<LOCKSTEP>();
                const gl_t *S = POS_SP + 23 * p;
                x[0] = gl_add_c(pos_pow7<DEFER>(x[0], bf), POS_RC[60 + p]);
                GlDotAcc acc;
                gl_dot_init(acc);
#pragma unroll
                for (int i = 0; i < 12; i++) gl_dot_mac(acc, x[i], S[i]);
                if (DEFER && GL_BORROW_MIN && !POS_MUL4W && !POS_MULADD_FUSED) {
#pragma unroll
                    for (int i = 1; i < 11; i += 2) gl_mul_add2_defer(x[0], S[11 + i], S[12 + i], x[i], x[i + 1], bf);
                    x[11] = gl_mul_add<DEFER>(x[0], S[22], x[11], bf);
                } else {
#pragma unroll
                    for (int i = 1; i < 12; i++) x[i] = gl_mul_add<DEFER>(x[0], S[11 + i], x[i], bf);
                }
                x[0] = gl_dot_fold<DEFER>(acc, bf);
            }
#endif
        }
    }
}

GL_HD void pos_permute(gl_t x[12], const PosXs &xs)
{
    GlBorrow bf = gl_borrow_init();
    pos_permute_t<false, false>(x, bf, xs);
}
