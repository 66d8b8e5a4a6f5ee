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
        "mad.lo.cc.u32 l0, m, 0xffffffff, l0;\n\t"
        "madc.hi.u32 l1, m, 0xffffffff, l1;\n\t"
        "mov.b64 %0, {l0,l1};\n\t"
        "}" : "=l"(r) : "l"(L), "l"(H));
    return r;
#else
    uint64_t lo = L + (H << 32);
    uint64_t top = (H >> 32) + (lo < L);
    return gl_reduce128(lo, top);
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
        uint64_t L = 0, H = 0;
#pragma unroll
        for (int j = 0; j < 12; j++) {
            uint32_t c = circ[(j - i + 12) % 12] + ((i == 0 && j == 0) ? GLP_MDS_DIAG0 : 0);
            L += (uint64_t)lo[j] * c;
            H += (uint64_t)hi[j] * c;
        }
        x[i] = pos_mds_fold(L, H);
    }
}

// ---- the permutation ----
GL_HD void pos_permute(gl_t x[12])
{
#pragma unroll
    for (int i = 0; i < 12; i++) x[i] = gl_add_c(x[i], POS_RC[i]);

    // first half: 3 x {pow7 + C, MDS}, then pow7 + C and the dense pre-multiplied matrix
#pragma unroll 1
    for (int r = 0; r < 3; r++) {
#pragma unroll
        for (int i = 0; i < 12; i++) x[i] = gl_add_c(gl_pow7(x[i]), POS_RC[12 * (r + 1) + i]);
        pos_mds(x);
    }
    {
        gl_t y[12];
#pragma unroll
        for (int i = 0; i < 12; i++) y[i] = gl_add_c(gl_pow7(x[i]), POS_RC[48 + i]);
#pragma unroll
        for (int i = 0; i < 12; i++) {
            GlDotAcc acc;
            gl_dot_init(acc);
#pragma unroll
            for (int j = 0; j < 12; j++) gl_dot_mac(acc, y[j], POS_PRE[j * 12 + i]);
            x[i] = gl_dot_fold(acc);
        }
    }

    // 22 partial rounds: x0 = x0^7 + C;  s0 = <x, S[0..12)>;  x_i += x0 * S[11+i] (i >= 1);  x0 = s0
#pragma unroll 1
    for (int r = 0; r < 22; r++) {
        const gl_t *S = POS_SP + 23 * r;
        x[0] = gl_add_c(gl_pow7(x[0]), POS_RC[60 + r]);
        GlDotAcc acc;
        gl_dot_init(acc);
#pragma unroll
        for (int i = 0; i < 12; i++) gl_dot_mac(acc, x[i], S[i]);
#pragma unroll
        for (int i = 1; i < 12; i++) x[i] = gl_mul_add(x[0], S[11 + i], x[i]);
        x[0] = gl_dot_fold(acc);
    }

    // second half: 3 x {pow7 + C, MDS}, then pow7 and MDS
#pragma unroll 1
    for (int r = 0; r < 3; r++) {
#pragma unroll
        for (int i = 0; i < 12; i++) x[i] = gl_add_c(gl_pow7(x[i]), POS_RC[82 + 12 * r + i]);
        pos_mds(x);
    }
#pragma unroll
    for (int i = 0; i < 12; i++) x[i] = gl_pow7(x[i]);
    pos_mds(x);
}
