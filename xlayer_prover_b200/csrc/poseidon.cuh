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

// value = Z + X*2^32 + Y*2^64  ->  limbs l0..l4 (weights 2^0..2^128), then
//   (l1:l0) + l2*EPS - l3 - l4*2^32   (2^64 = EPS, 2^96 = -1, 2^128 = -2^32)
GL_HD gl_t gl_dot_fold(const GlDotAcc &s)
{
    uint64_t t1 = (uint64_t)s.z1 + s.x0;
    uint64_t t2 = (uint64_t)s.zc + s.x1 + s.y0 + (t1 >> 32);
    uint64_t t3 = (uint64_t)s.xc + s.y1 + (t2 >> 32);
    uint64_t l4 = (uint64_t)s.yc + (t3 >> 32);
    gl_t lo = ((t1 & 0xFFFFFFFFULL) << 32) | s.z0;
    gl_t hi = ((t3 & 0xFFFFFFFFULL) << 32) | (t2 & 0xFFFFFFFFULL);
    gl_t r = gl_reduce128(lo, hi);
    return gl_sub(r, l4 << 32);
}

// a*b + c, one fold (c is added into the 128-bit product: (2^64-1)^2 + 2^64 - 1 < 2^128)
GL_HD gl_t gl_mul_add(gl_t a, gl_t b, gl_t c)
{
    gl_t lo, hi;
    gl_mul_wide(a, b, lo, hi);
#if GL_DEVICE_PTX
    asm("add.cc.u64 %0, %0, %2;\n\taddc.u64 %1, %1, 0;" : "+l"(lo), "+l"(hi) : "l"(c));
#else
    gl_t t = lo + c;
    hi += (t < lo);
    lo = t;
#endif
    return gl_reduce128(lo, hi);
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
        "mad.lo.cc.u32 l0, h1, 0xffffffff, l0;\n\t"
        "madc.hi.cc.u32 l1, h1, 0xffffffff, l1;\n\t"
        "addc.u32 m, 0, 0;\n\t"
        "sub.cc.u32 l0, l0, m;\n\t"
        "subc.u32 l1, l1, 0;\n\t"
        "add.u32 l1, l1, m;\n\t"
        "mov.b64 %0, {l0,l1};\n\t"
        "}" : "=l"(r) : "l"(L), "l"(H));
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
#define POS_ROLL_PRE 0            // 1 = 12-trip pre-matrix loop with a register queue (3 KB instead of 30 KB of code, measured 1-2 % slower)
#endif

template <bool LOCKSTEP>
GL_HD void pos_round_sync()
{
#if GL_DEVICE_PTX
    if (LOCKSTEP) __syncthreads();
#endif
}

// x_i = x_i^7 + rc_i for all 12, as 12/G trips over G lanes with the state rotated by G after each trip
GL_HD void pos_sbox_layer(gl_t x[12], const gl_t *rc)
{
    constexpr int G = POS_SBOX_GROUP;
    static_assert(12 % G == 0, "POS_SBOX_GROUP must divide 12");
    if (G == 12) {
#pragma unroll
        for (int i = 0; i < 12; i++) x[i] = gl_add_c(gl_pow7(x[i]), rc[i]);
        return;
    }
#pragma unroll 1
    for (int q = 0; q < 12 / G; q++) {
        gl_t t[G];
#pragma unroll
        for (int i = 0; i < G; i++) t[i] = gl_add_c(gl_pow7(x[i]), rc[q * G + i]);
#pragma unroll
        for (int i = 0; i + G < 12; i++) x[i] = x[i + G];
#pragma unroll
        for (int i = 0; i < G; i++) x[12 - G + i] = t[i];
    }
}

// x <- PRE^T x  (the dense pre-multiplied matrix before the partial rounds)
GL_HD void pos_pre_layer(gl_t x[12])
{
#if POS_ROLL_PRE
    gl_t out[12];
#pragma unroll
    for (int i = 0; i < 12; i++) out[i] = 0;
#pragma unroll 1
    for (int i = 0; i < 12; i++) {
        const gl_t *col = POS_PRET + 12 * i;
        GlDotAcc acc;
        gl_dot_init(acc);
#pragma unroll
        for (int j = 0; j < 12; j++) gl_dot_mac(acc, x[j], col[j]);
#pragma unroll
        for (int k = 0; k < 11; k++) out[k] = out[k + 1];
        out[11] = gl_dot_fold(acc);
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
#pragma unroll
        for (int j = 0; j < 12; j++) gl_dot_mac(acc, y[j], POS_PRE[j * 12 + i]);
        x[i] = gl_dot_fold(acc);
    }
#endif
}

template <bool LOCKSTEP>
GL_HD void pos_permute_t(gl_t x[12])
{
#pragma unroll
    for (int i = 0; i < 12; i++) x[i] = gl_add_c(x[i], POS_RC[i]);

    // 8 full rounds {x^7 + C ; linear layer}; the linear layer of round 3 is the pre-matrix followed by the
    // 22 partial rounds, every other one is the MDS matrix; round 7 adds zero constants
#pragma unroll 1
    for (int r = 0; r < 8; r++) {
        pos_round_sync<LOCKSTEP>();
        pos_sbox_layer(x, POS_RCF + 12 * r);
        if (r != 3) {
            pos_mds(x);
        } else {
            pos_pre_layer(x);
            // partial rounds: x0 = x0^7 + C;  s0 = <x, S[0..12)>;  x_i += x0 * S[11+i] (i >= 1);  x0 = s0
#pragma unroll 1
            for (int p = 0; p < 22; p++) {
                pos_round_sync<LOCKSTEP>();
                const gl_t *S = POS_SP + 23 * p;
                x[0] = gl_add_c(gl_pow7(x[0]), POS_RC[60 + p]);
                GlDotAcc acc;
                gl_dot_init(acc);
#pragma unroll
                for (int i = 0; i < 12; i++) gl_dot_mac(acc, x[i], S[i]);
#pragma unroll
                for (int i = 1; i < 12; i++) x[i] = gl_mul_add(x[0], S[11 + i], x[i]);
                x[0] = gl_dot_fold(acc);
            }
        }
    }
}

GL_HD void pos_permute(gl_t x[12]) { pos_permute_t<false>(x); }
