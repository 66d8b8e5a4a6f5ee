// K0 -- Goldilocks field F_p, p = 2^64 - 2^32 + 1, device arithmetic for sm_100a.
//
// Replaces the reference's scalar field ops (src/goldilocks/src/goldilocks_base_field_scalar.hpp:16-35 add,
// :62-80 sub, :103-160 mul; constants goldilocks_base_field.cpp:5-51).  Same contract as the reference:
// operands and results are raw uint64_t in [0, 2^64) -- NOT necessarily canonical -- and every
// function accepts non-canonical inputs.  gl_canon() is the reference's Goldilocks::toU64
// (goldilocks_base_field_tools.hpp:146-163); every kernel applies it once on the way out, so everything
// this library emits is canonical.
//
// The identities used:  2^64 = 2^32 - 1 =: EPS (mod p),  2^96 = -1 (mod p).
// A 64x64 product is four IMAD.WIDE.U32 (32-bit limbs); the 128-bit result p3:p2:p1:p0 folds as
//     (p1:p0) - p3 + p2*EPS      with one conditional -EPS (borrow) and one conditional +EPS (carry).
// The PTX below pins the carry chains so ptxas emits 4 IMAD.WIDE + 14 integer ops per multiply
// (the plain C++ form compiles to 29).  The portable branch is the same arithmetic in C++ and is what
// the host-thread emulator (tests/cuemu) runs.
#pragma once
#include <stdint.h>

#if defined(__CUDA_ARCH__)
#define GL_DEVICE_PTX 1
#else
#define GL_DEVICE_PTX 0
#endif

#ifndef GL_HD
#if defined(__CUDACC__)
#define GL_HD __host__ __device__ __forceinline__
#else
#define GL_HD static inline
#endif
#endif

typedef unsigned long long gl_t;   // same width as the reference's Goldilocks::Element::fe (uint64_t)

#define GL_P 0xFFFFFFFF00000001ULL
#define GL_EPS 0xFFFFFFFFULL
// generator of the 2^32 roots of unity: the reference's Goldilocks::W[32] (goldilocks_base_field.cpp:44);
// W[k] = GL_W32^(2^(32-k)), which makes W[6] = 8: every 64th root of unity is a power of two.
#define GL_W32 7277203076849721926ULL
#define GL_SHIFT 7ULL              // coset generator, Goldilocks::SHIFT (goldilocks_base_field.cpp:50)

// (lo1:lo0) += h*EPS with the carry flag left for the following addc, as PTX text (h*EPS = (h << 32) - h).
//   _ALU: two subtracts + two adds on the ALU pipe -- for the Poseidon kernels, which are bound by the fmaheavy pipe;
//   _FMA: mad.lo.cc/madc.hi.cc, which ptxas turns into an IMAD.HI (4 fmaheavy cycles) -- for the NTT kernels, which are
//         bound by the ALU pipe.
#define GL_PTX_ADD_EPS_MUL_ALU(lo0, lo1, h)                               \
    "{\n\t.reg .u32 e0,e1;\n\t"                                          \
    "sub.cc.u32 e0, 0, " h ";\n\t"                                        \
    "subc.u32 e1, " h ", 0;\n\t"                                          \
    "add.cc.u32 " lo0 ", " lo0 ", e0;\n\t"                                \
    "addc.cc.u32 " lo1 ", " lo1 ", e1;\n\t}\n\t"
#define GL_PTX_ADD_EPS_MUL_FMA(lo0, lo1, h)                               \
    "mad.lo.cc.u32 " lo0 ", " h ", 0xffffffff, " lo0 ";\n\t"              \
    "madc.hi.cc.u32 " lo1 ", " h ", 0xffffffff, " lo1 ";\n\t"
#ifndef GL_EPS_ALU
#define GL_EPS_ALU 0          // what plain gl_mul / gl_reduce128 use; the NTT kernels ask for the _FMA form explicitly
#endif

GL_HD gl_t gl_canon(gl_t a) { return a >= GL_P ? a - GL_P : a; }

// a + b (mod p).  Handles the double wrap the reference handles (two non-canonical operands).
GL_HD gl_t gl_add(gl_t a, gl_t b)
{
#if GL_DEVICE_PTX
    gl_t r;
    asm("{\n\t"
        ".reg .u32 a0,a1,b0,b1,m;\n\t"
        "mov.b64 {a0,a1}, %1;\n\t"
        "mov.b64 {b0,b1}, %2;\n\t"
        "add.cc.u32 a0, a0, b0;\n\t"
        "addc.cc.u32 a1, a1, b1;\n\t"
        "addc.u32 m, 0, 0;\n\t"                 // carry
        "sub.u32 m, 0, m;\n\t"                  // 0xffffffff on carry.  (A mad by 0xffffffff would do this on the fma pipe,
        "add.cc.u32 a0, a0, m;\n\t"             //  6 cycles of it; and `subc` straight after an add chain reads the carry
        "addc.cc.u32 a1, a1, 0;\n\t"            //  flag with the wrong polarity.)
        "addc.u32 m, 0, 0;\n\t"                 // second wrap (only when both operands >= p-ish)
        "sub.cc.u32 a0, a0, m;\n\t"
        "subc.u32 a1, a1, 0;\n\t"
        "add.u32 a1, a1, m;\n\t"
        "mov.b64 %0, {a0,a1};\n\t"
        "}" : "=l"(r) : "l"(a), "l"(b));
    return r;
#else
    gl_t r = a + b;
    if (r < a) { gl_t r2 = r + GL_EPS; if (r2 < r) r2 += GL_EPS; r = r2; }
    return r;
#endif
}

// a + b where at least one operand is canonical (< p): a single wrap suffices.
GL_HD gl_t gl_add_c(gl_t a, gl_t b)
{
#if GL_DEVICE_PTX
    gl_t r;
    asm("{\n\t"
        ".reg .u32 a0,a1,b0,b1,m;\n\t"
        "mov.b64 {a0,a1}, %1;\n\t"
        "mov.b64 {b0,b1}, %2;\n\t"
        "add.cc.u32 a0, a0, b0;\n\t"
        "addc.cc.u32 a1, a1, b1;\n\t"
        "addc.u32 m, 0, 0;\n\t"
        "sub.cc.u32 a0, a0, m;\n\t"
        "subc.u32 a1, a1, 0;\n\t"
        "add.u32 a1, a1, m;\n\t"
        "mov.b64 %0, {a0,a1};\n\t"
        "}" : "=l"(r) : "l"(a), "l"(b));
    return r;
#else
    gl_t r = a + b;
    if (r < a) r += GL_EPS;
    return r;
#endif
}

// a - b (mod p)
GL_HD gl_t gl_sub(gl_t a, gl_t b)
{
#if GL_DEVICE_PTX
    gl_t r;
    asm("{\n\t"
        ".reg .u32 a0,a1,b0,b1,m;\n\t"
        "mov.b64 {a0,a1}, %1;\n\t"
        "mov.b64 {b0,b1}, %2;\n\t"
        "sub.cc.u32 a0, a0, b0;\n\t"
        "subc.cc.u32 a1, a1, b1;\n\t"
        "subc.u32 m, 0, 0;\n\t"                 // 0xffffffff on borrow
        "sub.cc.u32 a0, a0, m;\n\t"             // - EPS on borrow
        "subc.cc.u32 a1, a1, 0;\n\t"
        "subc.u32 m, 0, 0;\n\t"                 // second wrap
        "sub.cc.u32 a0, a0, m;\n\t"
        "subc.u32 a1, a1, 0;\n\t"
        "mov.b64 %0, {a0,a1};\n\t"
        "}" : "=l"(r) : "l"(a), "l"(b));
    return r;
#else
    gl_t r = a - b;
    if (a < b) { gl_t r2 = r - GL_EPS; if (r2 > r) r2 -= GL_EPS; r = r2; }
    return r;
#endif
}

GL_HD gl_t gl_neg(gl_t a) { return gl_sub(0, a); }

// ---- add / subtract with the SECOND wrap left out and its occurrence counted ----
// gl_add's second wrap needs both operands >= 2^64 - 2^32, gl_sub's needs b > 2^64 - 2^32 + a: 2^-64 per operation on
// operands spread over the field, impossible between canonical operands -- but results in [p, 2^64) are legal here, so
// crafted inputs can reach it.  These forms only COUNT the event (three / two instructions fewer) and leave the value
// wrong when it fires; whoever uses them must discard the results and redo the work with gl_add / gl_sub when
// gl_defer_hit() (the optimistic NTT passes raise a flag that the host checks, capi.cu dev_extend_pol_checked).
// Two counters because the add form can only count up and the subtract form only down: one shared counter could cancel.
struct GlDefer { uint32_t up, dn; };
GL_HD bool gl_defer_hit(const GlDefer &f) { return (f.up | f.dn) != 0; }
GL_HD gl_t gl_add_defer(gl_t a, gl_t b, GlDefer &f)
{
#if GL_DEVICE_PTX
    gl_t r;
    asm("{\n\t"
        ".reg .u32 a0,a1,b0,b1,m;\n\t"
        "mov.b64 {a0,a1}, %2;\n\t"
        "mov.b64 {b0,b1}, %3;\n\t"
        "add.cc.u32 a0, a0, b0;\n\t"
        "addc.cc.u32 a1, a1, b1;\n\t"
        "addc.u32 m, 0, 0;\n\t"
        "sub.u32 m, 0, m;\n\t"                        // + EPS on carry
        "add.cc.u32 a0, a0, m;\n\t"
        "addc.cc.u32 a1, a1, 0;\n\t"
        "addc.u32 %1, %1, 0;\n\t"                     // up += second carry
        "mov.b64 %0, {a0,a1};\n\t"
        "}" : "=l"(r), "+r"(f.up) : "l"(a), "l"(b));
    return r;
#else
    gl_t r = a + b;
    if (r < a) { const gl_t r2 = r + GL_EPS; if (r2 < r) f.up++; r = r2; }
    return r;
#endif
}
GL_HD gl_t gl_sub_defer(gl_t a, gl_t b, GlDefer &f)
{
#if GL_DEVICE_PTX
    gl_t r;
    asm("{\n\t"
        ".reg .u32 a0,a1,b0,b1,m;\n\t"
        "mov.b64 {a0,a1}, %2;\n\t"
        "mov.b64 {b0,b1}, %3;\n\t"
        "sub.cc.u32 a0, a0, b0;\n\t"
        "subc.cc.u32 a1, a1, b1;\n\t"
        "subc.u32 m, 0, 0;\n\t"                       // 0xffffffff on borrow
        "sub.cc.u32 a0, a0, m;\n\t"                   // - EPS on borrow
        "subc.cc.u32 a1, a1, 0;\n\t"
        "subc.u32 %1, %1, 0;\n\t"                     // dn -= second borrow
        "mov.b64 %0, {a0,a1};\n\t"
        "}" : "=l"(r), "+r"(f.dn) : "l"(a), "l"(b));
    return r;
#else
    gl_t r = a - b;
    if (a < b) { const gl_t r2 = r - GL_EPS; if (r2 > r) f.dn--; r = r2; }
    return r;
#endif
}
template <bool DEFER> GL_HD gl_t gl_add_x(gl_t a, gl_t b, GlDefer &f) { return DEFER ? gl_add_defer(a, b, f) : gl_add(a, b); }
template <bool DEFER> GL_HD gl_t gl_sub_x(gl_t a, gl_t b, GlDefer &f) { return DEFER ? gl_sub_defer(a, b, f) : gl_sub(a, b); }

// fold a 128-bit value hi:lo to 64 bits (mod p); EPSALU picks the pipe of the p2*EPS term (see above)
template <bool EPSALU>
GL_HD gl_t gl_reduce128_t(gl_t lo, gl_t hi)
{
#if GL_DEVICE_PTX
    gl_t r;
#define GL_REDUCE128_PTX(EPSMUL)                                                                       \
    asm("{\n\t"                                                                                        \
        ".reg .u32 p0,p1,p2,p3,m;\n\t"                                                                 \
        "mov.b64 {p0,p1}, %1;\n\t"                                                                     \
        "mov.b64 {p2,p3}, %2;\n\t"                                                                     \
        "sub.cc.u32 p0, p0, p3;\n\t" /* (p1:p0) - p3          [2^96 = -1] */                           \
        "subc.cc.u32 p1, p1, 0;\n\t"                                                                   \
        "subc.u32 m, 0, 0;\n\t"                                                                        \
        "sub.cc.u32 p0, p0, m;\n\t" /*   - EPS on borrow */                                            \
        "subc.u32 p1, p1, 0;\n\t"                                                                      \
        EPSMUL("p0", "p1", "p2") /* + p2*EPS        [2^64 = EPS] */                                    \
        "addc.u32 m, 0, 0;\n\t" /* + carry*EPS = - carry + (carry << 32) */                            \
        "sub.cc.u32 p0, p0, m;\n\t"                                                                    \
        "subc.u32 p1, p1, 0;\n\t"                                                                      \
        "add.u32 p1, p1, m;\n\t"                                                                       \
        "mov.b64 %0, {p0,p1};\n\t"                                                                     \
        "}" : "=l"(r) : "l"(lo), "l"(hi))
    if (EPSALU) GL_REDUCE128_PTX(GL_PTX_ADD_EPS_MUL_ALU);
    else GL_REDUCE128_PTX(GL_PTX_ADD_EPS_MUL_FMA);
#undef GL_REDUCE128_PTX
    return r;
#else
    uint32_t h0 = (uint32_t)hi, h1 = (uint32_t)(hi >> 32);
    gl_t t = lo - h1;
    if (lo < h1) t -= GL_EPS;
    gl_t u = ((gl_t)h0 << 32) - h0;
    gl_t r = t + u;
    if (r < u) r += GL_EPS;
    return r;
#endif
}
GL_HD gl_t gl_reduce128(gl_t lo, gl_t hi) { return gl_reduce128_t<GL_EPS_ALU != 0>(lo, hi); }

// The same fold with ONE IMAD.WIDE on the multiplier pipe (4 cycles) instead of the IMAD.LO + IMAD.HI pair (6):
//     u = p2*EPS + (EPS - p3)            no overflow: (p2 + 1)*EPS <= 2^64 - 2^32, and u >= 0
//     t = (p1:p0) + u, carry c           value = t + c*2^64 - EPS  =  c ? t : t - EPS  (mod p)
//     r = c ? t : t - EPS, and once more - EPS when that borrows (then r >= 2^64 - 2*EPS, no third wrap)
// One ALU instruction more than gl_reduce128_t<false>; for kernels bound by the multiplier pipes (Poseidon).
#if defined(__CUDACC__) && !defined(GLP_EMU)
// EPS as a value ptxas cannot see: with the literal it rewrites p2 * 0xffffffff as a negate + IMAD.HI (6 cycles again);
// read from the constant bank the product stays one IMAD.WIDE with a c[][] operand
static __constant__ uint32_t GL_EPS_OPAQUE = 0xffffffffu;
#endif
GL_HD gl_t gl_reduce128_wide(gl_t lo, gl_t hi)
{
#if GL_DEVICE_PTX
    gl_t r;
    asm("{\n\t"
        ".reg .u32 p0,p1,p2,p3,u0,u1,m;\n\t"
        "mov.b64 {p0,p1}, %1;\n\t"
        "mov.b64 {p2,p3}, %2;\n\t"
        "not.b32 p3, p3;\n\t"                         // EPS - p3
        "mad.lo.cc.u32 u0, p2, %3, p3;\n\t"
        "madc.hi.u32 u1, p2, %3, 0;\n\t"
        "add.cc.u32 p0, p0, u0;\n\t"
        "addc.cc.u32 p1, p1, u1;\n\t"
        "addc.u32 m, 0xffffffff, 0;\n\t"              // EPS when there was NO carry, else 0
        "sub.cc.u32 p0, p0, m;\n\t"
        "subc.cc.u32 p1, p1, 0;\n\t"
        "subc.u32 m, 0, 0;\n\t"                       // 0xffffffff on borrow
        "sub.cc.u32 p0, p0, m;\n\t"
        "subc.u32 p1, p1, 0;\n\t"
        "mov.b64 %0, {p0,p1};\n\t"
        "}" : "=l"(r) : "l"(lo), "l"(hi), "r"(GL_EPS_OPAQUE));
    return r;
#else
    return gl_reduce128_t<false>(lo, hi);
#endif
}
// The fold with its RARE wrap deferred.  (p1:p0) - p3 borrows only when the low 64 bits of a product are below p3 < 2^32:
// probability 2^-33 per fold on anything but crafted data, yet its fix-up (three instructions) runs every time.  Here the
// event is only RECORDED in `bf` and the value is left wrong when it happens; the caller must redo its whole computation
// with the exact fold whenever gl_borrow_hit(bf) (k_linear_hash_rows redoes the row).
//
// Recording (GL_BORROW_MIN 1): the fold hands back the 32-bit word `w` that must be ZERO for the event to be possible and
// the caller keeps bf.zmin = min(bf.zmin, w); zmin == 0 at the end is the (conservative, 2^-32 per fold) sign.  Two folds
// share one three-input VIMNMX3 (gl_borrow_note2): half an instruction per fold, and nothing hangs off the fold's carry
// chain.  GL_BORROW_MIN 0 is round 1's form: bf.cnt -= borrow, one carry-chain instruction per fold.
//
// Forms of the fold itself (GL_FOLD_WIDE), SASS per fold next to the 4 IMAD.WIDE + 4 of the product (tools/sass_leaf.py):
//   0 : t = (p1:p0) - p3; t += p2*EPS with the literal 0xffffffff, which ptxas turns into IMAD.HI + IMAD.IADD;
//       + carry*EPS.                                                             2 multiplier-pipe + 5 other (round 1)
//   1 : the same with EPS read from the constant bank: ptxas keeps ONE IMAD.WIDE with carry-out.   1 + 5
//   2 : the multiply first, then both corrections in one subtraction:
//           t = (p1:p0) + p2*EPS, carry m        (t <= 2^64 - 2^33 when m = 1)
//           r = t - (p3 + m) + (m << 32)         (= t - p3 + m*EPS; 2^64 = EPS)                     1 + 4
//       r < 0 needs m = 0 and t < p3, i.e. the high word of t zero: that word is `w`.  p3 + m must not overflow: for a
//       product a*b, and for a*b + c with b < p (the Poseidon constants are canonical), p3 = 2^32 - 1 forces m = 0 (see
//       DESIGN.md K6); a general 128-bit value (the dot-product accumulators) can have both, so PRODUCT = false takes form 1.
// (Dropped in round 1, profiles/r01_poseidon_defer.txt: the gl_reduce128_wide form with ITS 2^-32 borrow deferred, all on
// the ALU pipe, the carry fix-up on the multiplier pipe -- equal or slower.)
#ifndef GL_BORROW_MIN
#define GL_BORROW_MIN 1
#endif
#ifndef GL_FOLD_WIDE
#define GL_FOLD_WIDE 2
#endif
struct GlBorrow { uint32_t cnt, zmin; };
GL_HD GlBorrow gl_borrow_init() { GlBorrow b; b.cnt = 0; b.zmin = 0xffffffffu; return b; }
GL_HD bool gl_borrow_hit(const GlBorrow &b) { return b.cnt != 0 || b.zmin == 0; }
GL_HD uint32_t gl_min3(uint32_t a, uint32_t b, uint32_t c)
{
#if GL_DEVICE_PTX
    return __vimin3_u32(a, b, c);                                 // VIMNMX3.U32
#else
    const uint32_t t = a < b ? a : b;
    return t < c ? t : c;
#endif
}
GL_HD void gl_borrow_note(GlBorrow &bf, uint32_t w) { bf.zmin = bf.zmin < w ? bf.zmin : w; }
GL_HD void gl_borrow_note2(GlBorrow &bf, uint32_t wa, uint32_t wb) { bf.zmin = gl_min3(bf.zmin, wa, wb); }

// hi:lo -> 64 bits, rare wrap left out; w = the word that is zero whenever the wrap was possible (GL_BORROW_MIN)
template <bool PRODUCT>
GL_HD gl_t gl_fold_defer_raw(gl_t lo, gl_t hi, uint32_t &w)
{
    constexpr bool V2 = PRODUCT && GL_FOLD_WIDE >= 2;
#if GL_DEVICE_PTX
    gl_t r;
    if (V2 && GL_FOLD_WIDE == 3) {                               // form 2 with p2*EPS = (p2 << 32) - p2 on the ALU pipe (experiment)
        asm("{\n\t"
            ".reg .u32 p0,p1,p2,p3,m,e0,e1;\n\t"
            "mov.b64 {p0,p1}, %2;\n\t"
            "mov.b64 {p2,p3}, %3;\n\t"
            "sub.cc.u32 e0, 0, p2;\n\t"
            "subc.u32 e1, p2, 0;\n\t"
            "add.cc.u32 p0, p0, e0;\n\t"
            "addc.cc.u32 p1, p1, e1;\n\t"
            "addc.u32 m, 0, 0;\n\t"
            "mov.b32 %1, p1;\n\t"
            "add.u32 p3, p3, m;\n\t"
            "sub.cc.u32 p0, p0, p3;\n\t"
            "subc.u32 p1, p1, 0;\n\t"
            "add.u32 p1, p1, m;\n\t"
            "mov.b64 %0, {p0,p1};\n\t"
            "}" : "=l"(r), "=r"(w) : "l"(lo), "l"(hi));
        return r;
    }
    if (V2) {
        asm("{\n\t"
            ".reg .u32 p0,p1,p2,p3,m;\n\t"
            "mov.b64 {p0,p1}, %2;\n\t"
            "mov.b64 {p2,p3}, %3;\n\t"
            "mad.lo.cc.u32 p0, p2, %4, p0;\n\t"
            "madc.hi.cc.u32 p1, p2, %4, p1;\n\t"
            "addc.u32 m, 0, 0;\n\t"
            "mov.b32 %1, p1;\n\t"
            "add.u32 p3, p3, m;\n\t"
            "sub.cc.u32 p0, p0, p3;\n\t"
            "subc.u32 p1, p1, 0;\n\t"
            "add.u32 p1, p1, m;\n\t"
            "mov.b64 %0, {p0,p1};\n\t"
            "}" : "=l"(r), "=r"(w) : "l"(lo), "l"(hi), "r"(GL_EPS_OPAQUE));
        return r;
    }
    w = (uint32_t)(lo >> 32);
#if GL_FOLD_WIDE
#define GL_DEFER_MUL "mad.lo.cc.u32 p0, p2, %3, p0;\n\t" "madc.hi.cc.u32 p1, p2, %3, p1;\n\t"
#else
#define GL_DEFER_MUL GL_PTX_ADD_EPS_MUL_FMA("p0", "p1", "p2")
#endif
    asm("{\n\t"
        ".reg .u32 p0,p1,p2,p3,m;\n\t"
        "mov.b64 {p0,p1}, %1;\n\t"
        "mov.b64 {p2,p3}, %2;\n\t"
        "sub.cc.u32 p0, p0, p3;\n\t"
        "subc.u32 p1, p1, 0;\n\t"
        GL_DEFER_MUL
        "addc.u32 m, 0, 0;\n\t"
        "sub.cc.u32 p0, p0, m;\n\t"
        "subc.u32 p1, p1, 0;\n\t"
        "add.u32 p1, p1, m;\n\t"
        "mov.b64 %0, {p0,p1};\n\t"
        "}" : "=l"(r) : "l"(lo), "l"(hi), "r"(GL_EPS_OPAQUE));
#undef GL_DEFER_MUL
    return r;
#else
    // the emulator models the deferral (value left unfixed, word handed back) so that the redo paths run on the CPU too
    const uint32_t h0 = (uint32_t)hi, h1 = (uint32_t)(hi >> 32);
    const gl_t u = ((gl_t)h0 << 32) - h0;
    if (V2) {
        const gl_t t = lo + u;
        const uint32_t m = t < u;
        w = (uint32_t)(t >> 32);
        return t - (gl_t)(uint32_t)(h1 + m) + ((gl_t)m << 32);    // no fix-up when this goes below zero
    }
    w = (uint32_t)(lo >> 32);
    const gl_t t = lo - h1;                                       // no -EPS fix-up when this borrows
    gl_t r = t + u;
    if (r < u) r += GL_EPS;
    return r;
#endif
}
// round 1's recording: the borrow is counted on the carry chain
GL_HD gl_t gl_fold_defer_cnt(gl_t lo, gl_t hi, uint32_t &cnt)
{
#if GL_DEVICE_PTX
    gl_t r;
    asm("{\n\t"
        ".reg .u32 p0,p1,p2,p3,m;\n\t"
        "mov.b64 {p0,p1}, %2;\n\t"
        "mov.b64 {p2,p3}, %3;\n\t"
        "sub.cc.u32 p0, p0, p3;\n\t"
        "subc.cc.u32 p1, p1, 0;\n\t"
        "subc.u32 %1, %1, 0;\n\t"                     // cnt -= borrow
        GL_PTX_ADD_EPS_MUL_FMA("p0", "p1", "p2")
        "addc.u32 m, 0, 0;\n\t"
        "sub.cc.u32 p0, p0, m;\n\t"
        "subc.u32 p1, p1, 0;\n\t"
        "add.u32 p1, p1, m;\n\t"
        "mov.b64 %0, {p0,p1};\n\t"
        "}" : "=l"(r), "+r"(cnt) : "l"(lo), "l"(hi));
    return r;
#else
    const uint32_t h0 = (uint32_t)hi, h1 = (uint32_t)(hi >> 32);
    if (lo < h1) cnt--;
    const gl_t t = lo - h1, u = ((gl_t)h0 << 32) - h0;
    gl_t r = t + u;
    if (r < u) r += GL_EPS;
    return r;
#endif
}
template <bool PRODUCT>
GL_HD gl_t gl_reduce128_defer_t(gl_t lo, gl_t hi, GlBorrow &bf)
{
#if GL_BORROW_MIN
    uint32_t w;
    const gl_t r = gl_fold_defer_raw<PRODUCT>(lo, hi, w);
    gl_borrow_note(bf, w);
    return r;
#else
    return gl_fold_defer_cnt(lo, hi, bf.cnt);
#endif
}
GL_HD gl_t gl_reduce128_defer(gl_t lo, gl_t hi, GlBorrow &bf) { return gl_reduce128_defer_t<true>(lo, hi, bf); }

// MODE 0: p2*EPS as IMAD.LO + IMAD.HI; 1: on the ALU pipe; 2: one IMAD.WIDE (gl_reduce128_wide)
template <int MODE>
GL_HD gl_t gl_reduce128_m(gl_t lo, gl_t hi)
{
    if (MODE == 2) return gl_reduce128_wide(lo, hi);
    return gl_reduce128_t<MODE == 1>(lo, hi);
}

// full 64x64 -> 128 product as 32-bit limbs (4 IMAD.WIDE.U32)
GL_HD void gl_mul_wide(gl_t a, gl_t b, gl_t &lo, gl_t &hi)
{
#if GL_DEVICE_PTX
    asm("{\n\t"
        ".reg .u32 a0,a1,b0,b1,x0,x1,s0,s1,s2,y0,y1,z0,z1;\n\t"
        ".reg .u64 X,Y,Z;\n\t"
        "mov.b64 {a0,a1}, %2;\n\t"
        "mov.b64 {b0,b1}, %3;\n\t"
        "mul.wide.u32 Z, a0, b0;\n\t"
        "mul.wide.u32 X, a0, b1;\n\t"
        "mul.wide.u32 Y, a1, b1;\n\t"
        "mov.b64 {x0,x1}, X;\n\t"
        "mad.lo.cc.u32 s0, a1, b0, x0;\n\t"
        "madc.hi.cc.u32 s1, a1, b0, x1;\n\t"
        "addc.u32 s2, 0, 0;\n\t"
        "mov.b64 {z0,z1}, Z;\n\t"
        "mov.b64 {y0,y1}, Y;\n\t"
        "add.cc.u32 z1, z1, s0;\n\t"
        "addc.cc.u32 y0, y0, s1;\n\t"
        "addc.u32 y1, y1, s2;\n\t"
        "mov.b64 %0, {z0,z1};\n\t"
        "mov.b64 %1, {y0,y1};\n\t"
        "}" : "=l"(lo), "=l"(hi) : "l"(a), "l"(b));
#else
    unsigned __int128 w = (unsigned __int128)a * b;
    lo = (gl_t)w;
    hi = (gl_t)(w >> 64);
#endif
}

// The same product as FOUR IMAD.WIDE (16 multiplier-pipe cycles instead of 18): no partial sum can overflow 64 bits,
//     Z = a0*b0;  X = a0*b1 + z1;  W = a1*b0 + x0;  Y = a1*b1 + (x1 + w1);   product = Y : w0 : z0
// ((2^32-1)^2 + 2^33 - 2 < 2^64), so no carry-out of a multiply is needed; the price is ALU moves that build the
// zero-extended 64-bit addends.
GL_HD void gl_mul_wide4(gl_t a, gl_t b, gl_t &lo, gl_t &hi)
{
#if GL_DEVICE_PTX
    asm("{\n\t"
        ".reg .u32 a0,a1,b0,b1,z0,z1,x0,x1,w0,w1,t0,t1,y0,y1;\n\t"
        ".reg .u64 Z;\n\t"
        "mov.b64 {a0,a1}, %2;\n\t"
        "mov.b64 {b0,b1}, %3;\n\t"
        "mul.wide.u32 Z, a0, b0;\n\t"
        "mov.b64 {z0,z1}, Z;\n\t"
        "mad.lo.cc.u32 x0, a0, b1, z1;\n\t"         // the lo.cc / hi pair with one addend is what ptxas keeps as a single
        "madc.hi.u32 x1, a0, b1, 0;\n\t"            // IMAD.WIDE with a 64-bit addend (a mad.wide.u32 gets its add split off)
        "mad.lo.cc.u32 w0, a1, b0, x0;\n\t"
        "madc.hi.u32 w1, a1, b0, 0;\n\t"
        "add.cc.u32 t0, x1, w1;\n\t"
        "addc.u32 t1, 0, 0;\n\t"
        "mad.lo.cc.u32 y0, a1, b1, t0;\n\t"
        "madc.hi.u32 y1, a1, b1, t1;\n\t"
        "mov.b64 %0, {z0,w0};\n\t"
        "mov.b64 %1, {y0,y1};\n\t"
        "}" : "=l"(lo), "=l"(hi) : "l"(a), "l"(b));
#else
    gl_mul_wide(a, b, lo, hi);
#endif
}
// multiply / square with the product form (W4: four IMAD.WIDE) and the fold (MODE, see gl_reduce128_m) chosen by the caller
template <int MODE, bool W4>
GL_HD gl_t gl_mul_m(gl_t a, gl_t b)
{
    gl_t lo, hi;
    if (W4) gl_mul_wide4(a, b, lo, hi);
    else gl_mul_wide(a, b, lo, hi);
    return gl_reduce128_m<MODE>(lo, hi);
}

GL_HD gl_t gl_mul(gl_t a, gl_t b)
{
    gl_t lo, hi;
    gl_mul_wide(a, b, lo, hi);
    return gl_reduce128(lo, hi);
}
// the same with the reduction's p2*EPS on the fma pipe (for ALU-bound kernels: the NTT passes)
GL_HD gl_t gl_mul_fma(gl_t a, gl_t b)
{
    gl_t lo, hi;
    gl_mul_wide(a, b, lo, hi);
    return gl_reduce128_t<false>(lo, hi);
}

// a*a: three IMAD.WIDE instead of four (the cross product a0*a1 is computed once and doubled with shifts on the
// ALU pipe, which has slack in the Poseidon kernels; IMAD.WIDE costs 4 fmaheavy cycles, a shift 2 ALU cycles)
#ifndef GL_SQR3
#define GL_SQR3 1
#endif
GL_HD void gl_sqr_wide(gl_t a, gl_t &lo, gl_t &hi)
{
#if GL_DEVICE_PTX && GL_SQR3
    asm("{\n\t"
        ".reg .u32 a0,a1,x0,x1,d0,d1,d2,y0,y1,z0,z1;\n\t"
        ".reg .u64 X,Y,Z;\n\t"
        "mov.b64 {a0,a1}, %2;\n\t"
        "mul.wide.u32 Z, a0, a0;\n\t"
        "mul.wide.u32 X, a0, a1;\n\t"
        "mul.wide.u32 Y, a1, a1;\n\t"
        "mov.b64 {x0,x1}, X;\n\t"
        "mov.b64 {z0,z1}, Z;\n\t"
        "mov.b64 {y0,y1}, Y;\n\t"
        "shl.b32 d0, x0, 1;\n\t"                      // 2X = d2:d1:d0 (65 bits)
        "shf.l.wrap.b32 d1, x0, x1, 1;\n\t"
        "shr.u32 d2, x1, 31;\n\t"
        "add.cc.u32 z1, z1, d0;\n\t"
        "addc.cc.u32 y0, y0, d1;\n\t"
        "addc.u32 y1, y1, d2;\n\t"
        "mov.b64 %0, {z0,z1};\n\t"
        "mov.b64 %1, {y0,y1};\n\t"
        "}" : "=l"(lo), "=l"(hi) : "l"(a));
#else
    gl_mul_wide(a, a, lo, hi);
#endif
}
GL_HD gl_t gl_sqr(gl_t a)
{
    gl_t lo, hi;
    gl_sqr_wide(a, lo, hi);
    return gl_reduce128(lo, hi);
}

template <int MODE>
GL_HD gl_t gl_sqr_m(gl_t a)
{
    gl_t lo, hi;
    gl_sqr_wide(a, lo, hi);
    return gl_reduce128_m<MODE>(lo, hi);
}
template <int MODE, bool W4>
GL_HD gl_t gl_pow7_m(gl_t x)
{
    gl_t x2 = gl_sqr_m<MODE>(x);
    gl_t x3 = gl_mul_m<MODE, W4>(x, x2);
    gl_t x4 = gl_sqr_m<MODE>(x2);
    return gl_mul_m<MODE, W4>(x3, x4);
}

// the same with the rare wrap of every fold deferred into bf (gl_fold_defer_raw; two folds per flag instruction)
GL_HD gl_t gl_pow7_defer(gl_t x, GlBorrow &bf)
{
    gl_t lo, hi;
#if GL_BORROW_MIN
    uint32_t wa, wb;
    gl_sqr_wide(x, lo, hi);
    const gl_t x2 = gl_fold_defer_raw<true>(lo, hi, wa);
    gl_mul_wide(x, x2, lo, hi);
    const gl_t x3 = gl_fold_defer_raw<true>(lo, hi, wb);
    gl_borrow_note2(bf, wa, wb);
    gl_sqr_wide(x2, lo, hi);
    const gl_t x4 = gl_fold_defer_raw<true>(lo, hi, wa);
    gl_mul_wide(x3, x4, lo, hi);
    const gl_t x7 = gl_fold_defer_raw<true>(lo, hi, wb);
    gl_borrow_note2(bf, wa, wb);
    return x7;
#else
    gl_sqr_wide(x, lo, hi);
    const gl_t x2 = gl_reduce128_defer(lo, hi, bf);
    gl_mul_wide(x, x2, lo, hi);
    const gl_t x3 = gl_reduce128_defer(lo, hi, bf);
    gl_sqr_wide(x2, lo, hi);
    const gl_t x4 = gl_reduce128_defer(lo, hi, bf);
    gl_mul_wide(x3, x4, lo, hi);
    return gl_reduce128_defer(lo, hi, bf);
#endif
}

// x^7, the Poseidon S-box (poseidon_goldilocks.hpp:91-97): 4 multiplies
GL_HD gl_t gl_pow7(gl_t x)
{
    gl_t x2 = gl_sqr(x);
    gl_t x3 = gl_mul(x, x2);
    gl_t x4 = gl_sqr(x2);
    return gl_mul(x3, x4);
}

GL_HD gl_t gl_pow(gl_t b, uint64_t e)
{
    gl_t r = 1;
    while (e) {
        if (e & 1) r = gl_mul(r, b);
        b = gl_mul(b, b);
        e >>= 1;
    }
    return r;
}

GL_HD gl_t gl_inv(gl_t a) { return gl_pow(a, GL_P - 2); }

// lo + h*EPS for h < 2^32 (the fold of a value below 2^96: its p3 is zero, so the "- p3" half of gl_reduce128 drops out)
GL_HD gl_t gl_fold96(gl_t lo, uint32_t h)
{
#if GL_DEVICE_PTX
    gl_t r;
    asm("{\n\t"
        ".reg .u32 p0,p1,m;\n\t"
        "mov.b64 {p0,p1}, %1;\n\t"
        GL_PTX_ADD_EPS_MUL_FMA("p0", "p1", "%2")
        "addc.u32 m, 0, 0;\n\t"
        "sub.cc.u32 p0, p0, m;\n\t"
        "subc.u32 p1, p1, 0;\n\t"
        "add.u32 p1, p1, m;\n\t"
        "mov.b64 %0, {p0,p1};\n\t"
        "}" : "=l"(r) : "l"(lo), "r"(h));
    return r;
#else
    const gl_t t = ((gl_t)h << 32) - h;
    gl_t r = lo + t;
    if (r < t) r += GL_EPS;
    return r;
#endif
}
// w0*EPS - w1 - w2*2^32  ( = (w0 + w1*2^32 + w2*2^64) * 2^64 mod p ) for w2 < 2^31.  w0*EPS <= (2^32-1)^2 and the
// subtrahend (w2:w1) is below 2^63, so one conditional wrap is exact.
GL_HD gl_t gl_fold_times_2p64(uint32_t w0, uint32_t w1, uint32_t w2)
{
#if GL_DEVICE_PTX
    gl_t r;
    asm("{\n\t"
        ".reg .u32 r0,r1,m;\n\t"
        "sub.cc.u32 r0, 0, %1;\n\t"                   // (w0 << 32) - w0
        "subc.u32 r1, %1, 0;\n\t"
        "sub.cc.u32 r0, r0, %2;\n\t"
        "subc.cc.u32 r1, r1, %3;\n\t"
        "subc.u32 m, 0, 0;\n\t"                       // 0xffffffff on borrow
        "sub.cc.u32 r0, r0, m;\n\t"                   // - EPS on borrow
        "subc.u32 r1, r1, 0;\n\t"
        "mov.b64 %0, {r0,r1};\n\t"
        "}" : "=l"(r) : "r"(w0), "r"(w1), "r"(w2));
    return r;
#else
    const gl_t x = ((gl_t)w0 << 32) - w0, y = ((gl_t)w2 << 32) | w1;
    gl_t r = x - y;
    if (x < y) r -= GL_EPS;
    return r;
#endif
}

// a * 2^k (mod p) for 0 <= k < 96: shifts and one fold, no multiplier.  Because W[6] = 8 every twiddle
// of a 64-point (and so of a 16-point) transform is such a power of two.  k is a compile-time constant at every call
// site of the NTT kernels, so only one of the branches survives:
//   k <= 32      : a*2^k < 2^96                    -> gl_fold96                    (9 instructions with the shifts)
//   32 < k <= 64 : the general fold                                               (15)
//   k > 64       : (a * 2^(k-64)) * 2^64, 95 bits  -> gl_fold_times_2p64           (10; was 18)
// GL_POW2_FMA 1 (device; off): the shift by k mod 32 as a multiplication by 2^(k mod 32) -- two IMAD.WIDE on the multiplier
// pipe instead of three or four SHF on the ALU pipe, the busier one in the NTT passes (ALU 70-79 %, fmaheavy 41-47 %).  The
// factor is read from the constant bank so that ptxas cannot turn the product back into shifts.  Measured: no change (LDE 313.4
// against 313.7 ms) -- the passes run at the issue rate of their mix, not at a pipe limit; kept as a knob.
#ifndef GL_POW2_FMA
#define GL_POW2_FMA 0
#endif
#if defined(__CUDACC__) && !defined(GLP_EMU)
static __constant__ uint32_t GL_POW2_OPAQUE[32] = {
    1u << 0,  1u << 1,  1u << 2,  1u << 3,  1u << 4,  1u << 5,  1u << 6,  1u << 7,  1u << 8,  1u << 9,  1u << 10, 1u << 11, 1u << 12, 1u << 13, 1u << 14, 1u << 15,
    1u << 16, 1u << 17, 1u << 18, 1u << 19, 1u << 20, 1u << 21, 1u << 22, 1u << 23, 1u << 24, 1u << 25, 1u << 26, 1u << 27, 1u << 28, 1u << 29, 1u << 30, 1u << 31};
#endif
GL_HD gl_t gl_mul_pow2(gl_t a, unsigned k)
{
    if (k == 0) return a;
#if GL_DEVICE_PTX && GL_POW2_FMA
    if (k % 32 != 0 && k < 96) {
        const uint32_t f = GL_POW2_OPAQUE[k % 32];
        const uint64_t t = (uint64_t)(uint32_t)a * f;
        const uint64_t u = (uint64_t)(uint32_t)(a >> 32) * f + (t >> 32);   // a * 2^(k % 32) = (u_hi : u_lo : t_lo), below 2^95
        const uint32_t w0 = (uint32_t)t, w1 = (uint32_t)u, w2 = (uint32_t)(u >> 32);
        if (k < 32) return gl_fold96(((gl_t)w1 << 32) | w0, w2);
        if (k < 64) return gl_reduce128((gl_t)w0 << 32, ((gl_t)w2 << 32) | w1);
        return gl_fold_times_2p64(w0, w1, w2);
    }
#endif
    if (k <= 32) return gl_fold96(a << k, (uint32_t)(a >> (64 - k)));
    if (k < 64) return gl_reduce128(a << k, a >> (64 - k));
    if (k == 64) return gl_reduce128(0, a);
    const gl_t w = a << (k - 64);
    return gl_fold_times_2p64((uint32_t)w, (uint32_t)(w >> 32), (uint32_t)(a >> (128 - k)));
}

// primitive 2^k-th root of unity, the reference's Goldilocks::w(k)
GL_HD gl_t gl_root_of_unity(unsigned k)
{
    gl_t w = GL_W32;
    for (unsigned i = k; i < 32; i++) w = gl_mul(w, w);
    return w;
}
