// K1-K5 -- Goldilocks NTT / INTT / low-degree extension over row-major [rows x cols] matrices.
//
// Replaces NTT_Goldilocks::{NTT, INTT, extendPol, reversePermutation, computeR} and the twiddle tables of its
// constructor (src/goldilocks/src/ntt_goldilocks.hpp:49-136,153-165; ntt_goldilocks.cpp:12-352).  Same maths
//
//   forward : out[k,:] = sum_i in[i,:] * w^(i k)            (w = W[log2 N], natural order in and out)
//   inverse : out[j,:] = 1/N * sum_k in[k,:] * w^(-j k)     (* shift^j when `extend`, ntt_goldilocks.cpp:122-150)
//
// but a different algorithm, chosen for the B200 (DESIGN.md "LDE"):
//
//  * LANE = COLUMN.  The reference layout (element (i,c) at a[i*ncols+c]) makes adjacent columns adjacent
//    words, and every column is an independent transform.  A warp owns 32 adjacent columns: each row access
//    of a warp is one contiguous 256-byte segment (two full lines), every twiddle factor is the same for all
//    32 lanes (it depends on the row only) and is fetched once per warp, shared-memory accesses are
//    conflict-free by construction, and no transpose or bit-reversal pass over HBM exists at all.
//  * A tile = Q <= 256 rows x 32 columns (64 KB of shared memory), transformed by two in-register radix-2^A /
//    radix-2^B steps (A,B <= 4) with ONE shared-memory exchange between them.  A transform of 2^n rows is
//    ceil(n/8) tile passes; pass t works on rows  b*L + q*(L/Q) + p  (q < Q) and is in place on that row set.
//  * Because W[6] = 8, every twiddle INSIDE a 16-point butterfly is a power of two: shifts and one fold
//    (gl_mul_pow2 with a compile-time exponent), no multiplier.  Real multiplies remain only between the two
//    steps of a tile (w_Q^(b u), from a 256-entry table) and between passes (w_L^(p k), one contiguous 2 KB
//    table row per tile).  The inverse transform is the same code with the output registers relabelled
//    (u -> -u mod 2^A) and the inverse tables.
//  * Low-degree extension without any reordering pass: the INTT runs decimation-in-frequency (natural in,
//    digit-reversed out, twiddle after the butterflies), the size-N_ext forward NTT runs decimation-in-time
//    (digit-reversed in, natural out, twiddle before the butterflies) with the same radices, so the INTT's
//    output IS the NTT's input order.  The coset scaling shift^j/N is applied to the INTT's last pass from a
//    table laid out in that storage order.  The zero padding (reversePermutation's zero-fill,
//    ntt_goldilocks.cpp:259-277) is never materialised: with blowup 2^e each contiguous run of Q3 input rows
//    of the NTT's first pass has its last Q3 - Q3/2^e rows identically zero; they are not read and the
//    butterflies that would touch them are pruned at compile time (LSH template parameter).
#pragma once
#include "cuda_compat.h"
#include "gl_field.cuh"

#define NTT_MAX_LOGQ 8              // tile rows <= 2^8
#define NTT_LOGT 12                 // two-level tables: low 12 bits / the rest
// CTA shape: 8 warps, registers capped for 3 CTAs per SM (80 registers; three 64 KB tiles are what shared memory holds).  Round 1
// ran 16 warps x 2 CTAs at 64 registers (8 x 3 was a tie then); once the table loads were no longer serialised behind branches
// the extra 16 registers (fewer spills in the 16-point transforms) and the third tile in flight won: LDE 312.8 -> 302.9 ms
// (profiles/r02_ntt_variants.txt).
#ifndef NTT_WARPS
#define NTT_WARPS 8
#endif
#ifndef NTT_MINBLOCKS
#define NTT_MINBLOCKS 3              // CTAs per SM the register allocation is capped for
#endif
#define NTT_LANES 32
#ifndef NTT_LOAD_FAST
#define NTT_LOAD_FAST 1              // a second, unpredicated copy of the tile loads for tiles whose inputs are all live (LDE -1 %)
#endif
#define NTT_THREADS (NTT_WARPS * NTT_LANES)
#define NTT_MAX_PASSES 4

enum { NTT_TW_NONE = 0, NTT_TW_PRE = 1, NTT_TW_POST = 2 };
enum { NTT_SC_NONE = 0, NTT_SC_CONST = 1, NTT_SC_TABLE = 2, NTT_SC_TWOLEVEL = 3 };

struct NttPass {
    const gl_t *in;                  // first column of this column block
    gl_t *out;
    uint64_t in_pitch, out_pitch;    // words per row
    uint32_t ncols;                  // columns in this block (lanes past the end idle)
    uint32_t logN, logL;             // whole transform; sub-transform this pass splits (tile stride = L/Q)
    uint32_t log_in_L;               // input rows per sub-transform: logL, or logL - e for the zero-padded first DIT pass
    uint64_t live;                   // input rows >= live of each sub-transform are zero and never read
    uint32_t tw_mode;                // NTT_TW_*
    const gl_t *tw;                  // [L/Q][Q]: w_L^(+-p*i); NULL -> two-level tables below
    const gl_t *tw_lo, *tw_hi;       // w_N^e = tw_hi[e >> 12] * tw_lo[e & 4095]   (forward root; the inverse negates e)
    const gl_t *zeta;                // w_256^i, i < 256  (step-to-step twiddles of a tile)
    uint32_t scale_mode;             // NTT_SC_*
    gl_t scale_const;                // 1/N
    const gl_t *sc;                  // NTT_SC_TABLE: shift^j/N indexed by the storage position of coefficient j
    const gl_t *sc_lo, *sc_hi;       // NTT_SC_TWOLEVEL: shift^j/N = sc_hi[j >> 12] * sc_lo[j & 4095], j natural
    uint32_t scatter;                // last DIF pass: store to the natural row index instead of in place
    uint32_t nprev, prev_log[NTT_MAX_PASSES - 1];   // tile sizes (log2) of the earlier passes, for the natural index
    uint32_t canon_in, canon_out;
    // Row-sharded output (the multi-GPU exchange fused into the transform's last pass): when nparts > 0 output row r
    // goes to parts[r >> log_part_rows] + (r & (2^log_part_rows - 1)) * part_pitch + part_col0 instead of `out`.  The
    // part bases may be peer-GPU memory mapped over NVLink (cudaIpcOpenMemHandle); parts is a small device array.
    const unsigned long long *parts;
    uint32_t nparts, log_part_rows;
    uint64_t part_pitch, part_col0;
    uint32_t log_cw;                 // NARROW kernels: lanes per tile = 2^log_cw >= ncols (32 >> log_cw tiles share a warp)
    uint32_t *defer_flag;            // OPTIMISTIC kernels: set to 1 when a deferred wrap fired (the results are then invalid)
    uint32_t opt;                    // launch the optimistic flavour of this pass (set by run_lde for a masked transform)
    gl_t mask;                       // first pass of an optimistic LDE (any kernel flavour): inputs times this constant instead of canonicalised
};

static __device__ __forceinline__ uint32_t ntt_bitrev(uint32_t x, uint32_t bits) { return bits ? (__brev(x) >> (32 - bits)) : 0; }
static constexpr __host__ __device__ int ntt_cbitrev(int x, int bits)
{
    int r = 0;
    for (int i = 0; i < bits; i++) r |= ((x >> i) & 1) << (bits - 1 - i);
    return r;
}

// ---- in-register 2^LOG-point DIF transform, twiddles are powers of two ----
// v[bitrev(u)] <- sum_m v[m] * w^(m u),  w = W[LOG] = 2^(192 / 2^LOG)   (W[4]=2^12, W[3]=2^24, W[2]=2^48, W[1]=2^96=-1)
// LSH: inputs with index >= (2^LOG >> LSH) are zero (not even initialised); butterflies on them are pruned.
// The optimistic passes are data-independent only on data that is spread over the field.  Real traces are not: they are
// full of 0, 1 and small values, -1 + 5 is the legal non-canonical word p + 4, and 0 - (p + 4) wraps twice -- on such data
// some lane always flags and the whole transform is repeated.  So an optimistic LDE multiplies every input by the fixed
// constant R = NTT_MASK in its first pass (instead of canonicalising it: x*R folds any 64-bit x) and the INTT's last pass
// multiplies by shift^j/(N*R) instead of shift^j/N (one scale table more): the transform is linear, the result is the
// same, and what flows through the butterflies is s*R mod p, spread over the field for the small s of a trace
// (profiles/r01_sweep_data_independence.jsonl).  Crafted data can still fire the flag; that only costs the repeat.
#define NTT_MASK 0x9E3779B97F4A7C15ULL

// OPT (optimistic): butterflies on gl_add_defer / gl_sub_defer (13 instead of 18 instructions); the count of wraps that
// were left out accumulates in f and the caller raises NttPass::defer_flag when it is non-zero.
template <int LOG, int LSH, bool OPT>
static __device__ __forceinline__ void ntt_dft_regs(gl_t *v, GlDefer &f)
{
    constexpr int Q = 1 << LOG;
    constexpr int LIVE = Q >> LSH;
#pragma unroll
    for (int t = 0; t < LOG; t++) {
        const int h = Q >> (t + 1);                 // half size of this stage's butterflies
        const int tshift = (96 / h);                // w_(2h) = 2^(96/h)
#pragma unroll
        for (int g = 0; g < Q; g += 2 * h) {
#pragma unroll
            for (int j = 0; j < h; j++) {
                if (LIVE <= h) {                    // upper input is zero: (a + 0, (a - 0) w^j)
                    if (j < LIVE) {
                        const gl_t a = v[g + j];
                        v[g + j + h] = (j == 0) ? a : gl_mul_pow2(a, (unsigned)(tshift * j));
                    } else {
                        v[g + j] = 0;
                        v[g + j + h] = 0;
                    }
                } else {
                    const gl_t a = v[g + j], b = v[g + j + h];
                    v[g + j] = gl_add_x<OPT>(a, b, f);
                    const gl_t d = gl_sub_x<OPT>(a, b, f);
                    v[g + j + h] = (j == 0) ? d : gl_mul_pow2(d, (unsigned)(tshift * j));
                }
            }
        }
    }
}
// register that holds output u of ntt_dft_regs; the inverse transform's output u is the forward one's -u
template <int LOG, bool INV>
static constexpr __host__ __device__ int ntt_slot(int u)
{
    return ntt_cbitrev(INV ? (((1 << LOG) - u) & ((1 << LOG) - 1)) : u, LOG);
}

// w_N^e from the two-level table
static __device__ __forceinline__ gl_t ntt_tw2(const gl_t *lo, const gl_t *hi, uint64_t e)
{
    const gl_t l = __ldg(lo + (e & ((1u << NTT_LOGT) - 1)));
    const uint64_t h = e >> NTT_LOGT;
    return h ? gl_mul_fma(l, __ldg(hi + h)) : l;
}

// ---- slow paths, kept out of line so that the unrolled fast path stays small (the kernel must fit the
// instruction cache: a first version with these inlined at all 48 load/store sites was 190 KB of SASS and
// spent 5 of every 11 issue cycles waiting for instruction fetch) ----
// factor kinds of the slow path
enum { NTT_F_TW = 0, NTT_F_TW_INV = 1, NTT_F_SCALE2 = 2, NTT_F_CONST = 3 };
struct NttSlow {                       // what the slow path needs, by value (a pointer to the kernel parameter would force a stack copy)
    const gl_t *tw_lo, *tw_hi, *sc_lo, *sc_hi;
    gl_t scale_const;
    uint32_t logN, logL;
};
static __device__ GLP_NOINLINE gl_t ntt_factor_slow(NttSlow a, uint32_t kind, uint64_t p, uint64_t idx)
{
    if (kind == NTT_F_CONST) return a.scale_const;
    if (kind == NTT_F_SCALE2) return ntt_tw2(a.sc_lo, a.sc_hi, idx);              // idx = natural coefficient index
    const uint64_t nmask = (1ull << a.logN) - 1;
    uint64_t e = ((p * idx) << (a.logN - a.logL)) & nmask;                         // w_L^(p i) = w_N^(p i N/L)
    if (kind == NTT_F_TW_INV) e = (0 - e) & nmask;
    return ntt_tw2(a.tw_lo, a.tw_hi, e);
}

// Outputs of a tile pass -> memory.  Output w (w < 2^LOGX) of the last in-register transform sits in
// v[ntt_slot<LOGX, INV>(w)] and is tile output t = t0 + w * tstride.  Each output is multiplied by at most one
// factor (the pass twiddle w_L^(+-p t) of a DIF pass, or the scaling of a last inverse pass), read from a table
// row (fptr[t << fsh]) or produced by the slow path, and stored at row rbase + (t << rsh).
template <int LOGX, bool INV>
static __device__ __forceinline__ void ntt_store_outputs(const NttPass &a, const NttSlow &sl, gl_t *v, uint32_t t0, uint32_t tstride, uint64_t p, bool mul,
                                                         const gl_t *fptr, uint32_t fsh, uint32_t fkind, uint64_t fbase, uint32_t fbsh,
                                                         uint64_t rbase, uint32_t rsh, uint32_t col, bool colok)
{
    constexpr int QX = 1 << LOGX;
    if (mul) {
        if (fptr) {                                              // table row: the pointer walks, no index arithmetic per element
            uint64_t fa = (uint64_t)(fptr + ((uint64_t)t0 << fsh));
            const uint64_t fstepb = ((uint64_t)tstride << fsh) * sizeof(gl_t);
#pragma unroll
            for (int w = 0; w < QX; w++) {
                gl_t &x = v[ntt_slot<LOGX, INV>(w)];
                x = gl_mul_fma(x, __ldg((const gl_t *)fa));
                fa += fstepb;
            }
        } else {
#pragma unroll
            for (int w = 0; w < QX; w++) {
                const uint32_t t = t0 + (uint32_t)w * tstride;
                gl_t &x = v[ntt_slot<LOGX, INV>(w)];
                x = gl_mul_fma(x, ntt_factor_slow(sl, fkind, p, fbase + ((uint64_t)t << fbsh)));
            }
        }
    }
    if (a.canon_out) {
#pragma unroll
        for (int w = 0; w < QX; w++) v[w] = gl_canon(v[w]);
    }
    if (!colok) return;
    if (a.nparts) {                                              // scatter rows to the slabs of their owners
        const uint64_t rmask = (1ull << a.log_part_rows) - 1;
#pragma unroll
        for (int w = 0; w < QX; w++) {
            const uint64_t r = rbase + ((uint64_t)(t0 + (uint32_t)w * tstride) << rsh);
            gl_t *base = (gl_t *)__ldg(a.parts + (r >> a.log_part_rows));
            base[(r & rmask) * a.part_pitch + a.part_col0 + col] = v[ntt_slot<LOGX, INV>(w)];
        }
        return;
    }
    uint64_t oa = (uint64_t)(a.out + (rbase + ((uint64_t)t0 << rsh)) * a.out_pitch + col);
    const uint64_t stepb = ((uint64_t)tstride << rsh) * a.out_pitch * sizeof(gl_t);
#pragma unroll
    for (int w = 0; w < QX; w++) {
        *(gl_t *)oa = v[ntt_slot<LOGX, INV>(w)];
        oa += stepb;
    }
}

// One tile pass.  A, B: log2 radix of the two in-register steps (B == 0: single step, no shared memory).
// grid.x = tiles * ceil(ncols / 32) (column group fastest, so neighbouring CTAs touch neighbouring segments of
// the same rows); B == 0 packs NTT_WARPS tiles into a CTA.
// NARROW (matrices of <= 16 columns: FRI polynomials, the cm4 quotient, single-column transforms): a warp's 32 lanes
// are shared by 32/cw neighbouring tiles of cw = 2^log_cw >= ncols lanes each.  Neighbouring tiles are neighbouring rows
// p, p+1, ..., so a warp still touches one contiguous run of memory per access; the tile index, and with it the twiddle
// row, becomes a per-lane quantity instead of a warp-uniform one.  Everything else is the same code.
// The body works on the tile(s) of virtual block index vbx with `nwarps` warps (k_ntt_pass: vbx = blockIdx.x, NTT_WARPS warps;
// k_ntt_pass_co: a small resident grid walks the block indices, see below).
template <int A, int B, bool INV, int LSH, bool NARROW, bool OPT = false>
static __device__ __forceinline__ void ntt_pass_block(const NttPass &a, gl_t *tile, const uint32_t vbx, const uint32_t nwarps)
{
    constexpr int QA = 1 << A, QB = 1 << B, LOGQ = A + B;
    const uint32_t lane = threadIdx.x % NTT_LANES, warp = threadIdx.x / NTT_LANES;
    const uint32_t ncg = (a.ncols + NTT_LANES - 1) / NTT_LANES;
    const uint64_t ntiles = 1ull << (a.logN - LOGQ);
    const uint32_t cg = NARROW ? 0 : vbx % ncg;
    uint64_t tix = NARROW ? vbx : vbx / ncg;
    if (B == 0) tix = tix * NTT_WARPS + warp;
    if (NARROW) tix = (tix << (5 - a.log_cw)) + (lane >> a.log_cw);
    bool tileok = tix < ntiles;
    if (B == 0 && !NARROW && !tileok) return;                    // no barrier in the single-step kernel
    if (!tileok) tix = 0;                                        // idle lanes of a partly filled narrow warp compute on tile 0 and store nothing
    const uint32_t logS = a.logL - LOGQ;
    const uint64_t p = tix & ((1ull << logS) - 1), bb = tix >> logS;
    const uint32_t col = NARROW ? (lane & ((1u << a.log_cw) - 1)) : cg * NTT_LANES + lane;
    const bool colok = col < a.ncols && tileok;
    const gl_t *twp = a.tw ? a.tw + (p << LOGQ) : nullptr;
    const uint64_t posbase = (bb << a.logL) + p;
    // natural index of tile output t: kbase + (t << ksh); digit k_i of pass i has weight Q_0...Q_(i-1)
    uint64_t kbase = 0;
    uint32_t ksh = 0;
    if (a.scatter || a.scale_mode == NTT_SC_TWOLEVEL) {
        for (uint32_t i = 0; i < a.nprev; i++) ksh += a.prev_log[i];
        uint64_t rem = bb;
        uint32_t sh = ksh;
        for (int i = (int)a.nprev - 1; i >= 0; i--) {
            sh -= a.prev_log[i];
            kbase |= (rem & ((1ull << a.prev_log[i]) - 1)) << sh;
            rem >>= a.prev_log[i];
        }
    }
    // the one factor each output may be multiplied by, and where outputs go
    bool omul = false;
    const gl_t *fptr = nullptr;
    uint32_t fsh = 0, fkind = 0, fbsh = 0;
    uint64_t fbase = 0;
    if (a.tw_mode == NTT_TW_POST && (NARROW || p != 0)) { omul = true; fptr = twp; fkind = INV ? NTT_F_TW_INV : NTT_F_TW; }   // row p = 0 of the table is all ones
    else if (a.scale_mode == NTT_SC_CONST) { omul = true; fkind = NTT_F_CONST; }
    else if (a.scale_mode == NTT_SC_TABLE) { omul = true; fptr = a.sc + posbase; fsh = logS; }
    else if (a.scale_mode == NTT_SC_TWOLEVEL) { omul = true; fkind = NTT_F_SCALE2; fbase = kbase; fbsh = ksh; }
    const uint64_t rbase = a.scatter ? kbase : posbase;
    const uint32_t rsh = a.scatter ? ksh : logS;
    const uint64_t in_step = ((uint64_t)QB << logS) * a.in_pitch;   // between consecutive high digits m of q = m*QB + b
    const bool pre = a.tw_mode == NTT_TW_PRE && (NARROW || p != 0);
    const NttSlow sl = {a.tw_lo, a.tw_hi, a.sc_lo, a.sc_hi, a.scale_const, a.logN, a.logL};

    // ---- step 1: 2^A-point transforms over the high digit of q = m*QB + b ----
#pragma unroll 1
    for (uint32_t b = (B == 0 ? 0 : warp); b < (uint32_t)QB; b += nwarps) {
        gl_t v[QA];
        {
            // input m is row r0 + m * 2^(B + logS) of the sub-transform; rows >= a.live are zero and never read.  The count of
            // live inputs is worked out once and the pointer walks by in_step: a 64-bit compare, a 64-bit multiply and an LEA
            // pair per element otherwise (12 of the 150 instructions per element and pass).
            const uint64_t r0 = ((uint64_t)b << logS) + p;       // input row of m = 0 inside the sub-transform
            const gl_t *in = a.in + ((bb << a.log_in_L) + r0) * a.in_pitch + col;
            uint32_t nlive = 0;
            if (colok && r0 < a.live) {
                const uint64_t k = ((a.live - r0 - 1) >> (B + logS)) + 1;
                nlive = k < (uint64_t)(QA >> LSH) ? (uint32_t)k : (uint32_t)(QA >> LSH);
            }
            uint64_t addr = (uint64_t)in;                        // byte address: one 64-bit add per element (an element index costs an LEA pair more)
            const uint64_t stepb = in_step * sizeof(gl_t);
#if NTT_LOAD_FAST
            if (nlive == (uint32_t)(QA >> LSH)) {                // every input live (all passes but a zero-padded first one): no predicates, no zero-fill
#pragma unroll
                for (int m = 0; m < (QA >> LSH); m++) {
                    v[m] = *(const gl_t *)addr;
                    addr += stepb;
                }
            } else
#endif
            {
#pragma unroll
                for (int m = 0; m < (QA >> LSH); m++) {
                    v[m] = (uint32_t)m < nlive ? *(const gl_t *)addr : 0;   // plain load: middle passes run in place
                    addr += stepb;
                }
            }
        }
        if (a.mask) {                                            // see NTT_MASK: x -> x*R, undone by the INTT's scale table
#pragma unroll
            for (int m = 0; m < (QA >> LSH); m++) v[m] = gl_mul_fma(v[m], a.mask);
        } else if (a.canon_in) {
#pragma unroll
            for (int m = 0; m < (QA >> LSH); m++) v[m] = gl_canon(v[m]);
        }
        if (pre) {                                               // w_L^(+-p q)
            // The table test stays OUTSIDE the loop: with `twp ? load : call` per element every load sits behind a branch, ptxas
            // cannot hoist it, and the 16 table loads of a warp were issued and waited for one after another -- half of the time
            // of a pass with pre-twiddles (ncu source page, profiles/r02_ntt_source_profile.txt).
            if (twp) {
#pragma unroll
                for (int m = 0; m < (QA >> LSH); m++) v[m] = gl_mul_fma(v[m], __ldg(twp + (uint32_t)m * QB + b));
            } else {
#pragma unroll
                for (int m = 0; m < (QA >> LSH); m++)
                    v[m] = gl_mul_fma(v[m], ntt_factor_slow(sl, INV ? NTT_F_TW_INV : NTT_F_TW, p, (uint32_t)m * QB + b));
            }
        }
        GlDefer f = {0, 0};
        ntt_dft_regs<A, LSH, OPT>(v, f);
        if (OPT && gl_defer_hit(f)) *a.defer_flag = 1u;
        if (B == 0) {
            ntt_store_outputs<A, INV>(a, sl, v, 0, 1, p, omul, fptr, fsh, fkind, fbase, fbsh, rbase, rsh, col, colok);
        } else {
            // w_Q^(+-b u) = w_256^(+-b u 256/Q) (zeta[0] = 1 for the b = 0 warp), then to shared memory at q' = u*QB + b
            gl_t *trow = tile + (size_t)b * NTT_LANES + lane;
            trow[0] = v[0];
#pragma unroll
            for (int u = 1; u < QA; u++) {
                uint32_t zi = (b * (uint32_t)u) << (8 - LOGQ);
                if (INV) zi = (256 - zi) & 255;
                trow[(size_t)u * QB * NTT_LANES] = gl_mul_fma(v[ntt_slot<A, INV>(u)], __ldg(a.zeta + zi));
            }
        }
    }
    if (B == 0) return;
    __syncthreads();

    // ---- step 2: 2^B-point transforms over b; tile output t = u + QA*w ----
#pragma unroll 1
    for (uint32_t u = warp; u < (uint32_t)QA; u += nwarps) {
        gl_t v[QB];
        const gl_t *trow = tile + (size_t)u * QB * NTT_LANES + lane;
#pragma unroll
        for (int b = 0; b < QB; b++) v[b] = trow[(size_t)b * NTT_LANES];
        GlDefer f = {0, 0};
        ntt_dft_regs<B, 0, OPT>(v, f);
        if (OPT && gl_defer_hit(f)) *a.defer_flag = 1u;
        ntt_store_outputs<B, INV>(a, sl, v, u, QA, p, omul, fptr, fsh, fkind, fbase, fbsh, rbase, rsh, col, colok);
    }
}

template <int A, int B, bool INV, int LSH, bool NARROW>
__global__ void __launch_bounds__(NTT_THREADS, NTT_MINBLOCKS)
k_ntt_pass(NttPass a)
{
    GLP_DYN_SMEM(gl_t, tile);                                    // [Q][32] when B > 0
    ntt_pass_block<A, B, INV, LSH, NARROW>(a, tile, blockIdx.x, NTT_WARPS);
}
// The optimistic flavour of the wide two-step passes: same tile pass on the deferred add / subtract.  Its results are
// valid only if a.defer_flag is still zero when the whole transform has finished; dev_extend_pol_checked (capi.cu) reads
// the flag and runs the transform again with k_ntt_pass when it is not.  Recovery lives on the host because any recovery
// code inside these 3000-instruction straight-line kernels costs what the shorter butterflies gain
// (profiles/r01_ntt_experiments.txt).
template <int A, int B, bool INV, int LSH>
__global__ void __launch_bounds__(NTT_THREADS, NTT_MINBLOCKS)
k_ntt_pass_opt(NttPass a)
{
    static_assert(B > 0, "optimistic passes are two-step tiles");
    GLP_DYN_SMEM(gl_t, tile);
    ntt_pass_block<A, B, INV, LSH, false, true>(a, tile, blockIdx.x, NTT_WARPS);
}

// Co-resident flavour (two-step tiles only): a grid of a few CTAs per SM, each of blockDim.x/32 <= NTT_WARPS warps, walks
// the `nvb` block indices of k_ntt_pass.  It is launched on a high-priority stream UNDER a running leaf-hash kernel
// (commit_overlapped in capi.cu): its whole grid becomes resident as soon as hash CTAs retire, takes a fixed slice of
// every SM (16 K registers and 64 KB of shared memory at 256 threads) and leaves the rest to the hash, so the LDE of
// column block k+1 runs in the issue slots and ALU cycles that the multiplier-bound hash of block k leaves idle.
template <int A, int B, bool INV, int LSH>
__global__ void __launch_bounds__(NTT_THREADS, 2)
k_ntt_pass_co(NttPass a, uint32_t nvb)
{
    static_assert(B > 0, "co-resident passes are two-step tiles");
    GLP_DYN_SMEM(gl_t, tile);
    const uint32_t nwarps = blockDim.x / NTT_LANES;
    for (uint32_t vbx = blockIdx.x; vbx < nvb; vbx += gridDim.x) {
        ntt_pass_block<A, B, INV, LSH, false>(a, tile, vbx, nwarps);
        __syncthreads();                                         // the tile buffer is reused by the next block index
    }
}

// size-1 "transform": canonical copy
__global__ void k_ntt_copy_canon(gl_t *out, uint64_t out_pitch, const gl_t *in, uint64_t in_pitch, uint64_t nrows, uint64_t ncols)
{
    const uint64_t total = nrows * ncols;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r = i / ncols, c = i % ncols;
        out[r * out_pitch + c] = gl_canon(in[r * in_pitch + c]);
    }
}

// rows of a dense [nrows x ncols] block -> the slabs of their owners (fallback of the fused scatter)
__global__ void k_ntt_scatter_rows(const unsigned long long *parts, uint32_t log_part_rows, uint64_t part_pitch, uint64_t part_col0,
                                   const gl_t *in, uint64_t in_pitch, uint64_t nrows, uint64_t ncols)
{
    const uint64_t total = nrows * ncols, rmask = (1ull << log_part_rows) - 1;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r = i / ncols, c = i % ncols;
        gl_t *base = (gl_t *)__ldg(parts + (r >> log_part_rows));
        base[(r & rmask) * part_pitch + part_col0 + c] = gl_canon(in[r * in_pitch + c]);
    }
}

// ---- table generation (K1) ----
// out[i] = c * g^(i * mul)
__global__ void k_ntt_pow_table(gl_t *out, gl_t c, gl_t g, uint64_t mul, uint64_t n)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = gl_canon(gl_mul_fma(c, gl_pow(g, i * mul)));
}
// pass table: out[p*Q + i] = g^(p*i), p < S, i < Q   (g = w_L or its inverse)
__global__ void k_ntt_pass_table(gl_t *out, gl_t g, uint32_t logQ, uint64_t n)
{
    uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= n) return;
    const uint64_t p = idx >> logQ, i = idx & ((1ull << logQ) - 1);
    out[idx] = gl_canon(gl_pow(g, p * i));
}
// scale table in the storage order of a DIF transform with tile sizes lg[0..npass): position
// (k_0, ..., k_(n-1)) (k_0 most significant) holds coefficient j = k_0 + Q_0 k_1 + Q_0 Q_1 k_2 + ...
struct NttPlanLogs { uint32_t npass, lg[NTT_MAX_PASSES]; };
__global__ void k_ntt_scale_table(gl_t *out, gl_t c, gl_t g, NttPlanLogs pl, uint64_t n)
{
    uint64_t pos = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n) return;
    uint64_t rem = pos, j = 0;
    uint32_t sh = 0;
    for (uint32_t i = 0; i < pl.npass; i++) sh += pl.lg[i];
    for (int i = (int)pl.npass - 1; i >= 0; i--) {
        sh -= pl.lg[i];
        j |= (rem & ((1ull << pl.lg[i]) - 1)) << sh;
        rem >>= pl.lg[i];
    }
    out[pos] = gl_canon(gl_mul_fma(c, gl_pow(g, j)));
}
