// K6 (leaf hash), K7 (tree levels) and the batched bare permutation.
//
//  k_linear_hash_rows : one thread = one leaf row; the reference's PoseidonGoldilocks::linear_hash
//                       (poseidon_goldilocks.cpp:229-265) for every row of a row-major matrix.
//  k_tree_level       : one thread = one parent node; one level of the loop in merkletree_avx
//                       (poseidon_goldilocks.cpp:283-300).
//  k_tree_top         : a single CTA finishes the last levels (<= 2*TOP_THREADS nodes) without relaunching.
//  k_permute_batch    : hash_full_result (poseidon_goldilocks.cpp:166-228) on n independent 12-word states.
//
// Roofline: integer issue.  A leaf row of C columns costs ceil(C/8) permutations (~2.2e4 integer
// instructions each) against 8*C bytes of HBM traffic, i.e. > 300 instructions per byte: the row loads are
// software-prefetched one sponge block ahead and otherwise left to the L1/L2 (each thread streams its own
// 64-byte-aligned chunks, so every 32-byte sector fetched is fully used).
#pragma once
#include "cuda_compat.h"
#include "poseidon.cuh"

#define POS_BLOCK 128
#ifndef POS_LEAF_BLOCK
#define POS_LEAF_BLOCK 128          // threads per CTA of the leaf-hash kernel
#endif
#ifndef POS_LEAF_LOCKSTEP
#define POS_LEAF_LOCKSTEP 0         // 1 = barrier per round so a CTA's warps share instruction fetches (no gain once the code fits: measured -3 %)
#endif

// load 8 consecutive words of a row (zero beyond `n` words)
static __device__ __forceinline__ void pos_load8(gl_t v[8], const gl_t *__restrict__ p, uint32_t n, bool vec16)
{
#if !defined(GLP_EMU)
    if (vec16 && n >= 8) {
        const ulonglong2 *q = reinterpret_cast<const ulonglong2 *>(p);
#pragma unroll
        for (int i = 0; i < 4; i++) { ulonglong2 t = __ldg(q + i); v[2 * i] = t.x; v[2 * i + 1] = t.y; }
        return;
    }
#else
    (void)vec16;
#endif
#pragma unroll
    for (int i = 0; i < 8; i++) v[i] = (uint32_t)i < n ? __ldg(p + i) : 0;
}

// out[row*4 .. +4) = linear_hash(in[row*pitch .. +ncols))
// cap_in != NULL continues a sponge: the capacity starts from cap_in[row*4 .. +4) (the first four outputs of the
// previous block's last permutation, :247-250) instead of zero, so a row can be absorbed one column block at a time
// (every block but the last a multiple of 8 columns); cap_in may alias out.
// LOCKSTEP: every thread of the CTA runs the same number of permutations (rows past the end are clamped to the
// last row and simply not stored), so the per-round barriers inside pos_permute_t are reached by all.
// POS_LEAF_MINBLOCKS: second argument of __launch_bounds__, i.e. a register cap (6 -> 80 registers, 4 -> 128; 0 = none: ptxas then
// settles at 96 registers = 5 CTAs of 128 threads per SM).  4 lets ptxas take 120 registers: no spills, 16 warps per SM,
// 1134 against 1124 M perm/s (profiles/r02_pos_variants.txt) -- the kernel does not need the occupancy.
#ifndef POS_LEAF_MINBLOCKS
#define POS_LEAF_MINBLOCKS 4
#endif
#if POS_LEAF_MINBLOCKS
#define POS_LEAF_BOUNDS __launch_bounds__(POS_LEAF_BLOCK, POS_LEAF_MINBLOCKS)
#else
#define POS_LEAF_BOUNDS __launch_bounds__(POS_LEAF_BLOCK)
#endif
// POS_DEFER_BORROW: the row is hashed with the folds' rare wrap deferred (gl_reduce128_defer: two ALU instructions fewer
// in each of the 748 folds of a permutation); a row in which any fold did wrap (2^-33 per fold: about one row in 10^4 at
// 83 permutations a row) is hashed again by pos_leaf_row_exact, a separate out-of-line copy with the exact folds.
// g_pos_redo_rows counts those rows (glp_leaf_redo_rows; the GPU tests drive crafted rows through it).
#ifndef POS_DEFER_BORROW
#define POS_DEFER_BORROW 1
#endif
#if !defined(GLP_EMU)
static __device__ unsigned long long g_pos_redo_rows;
#define POS_COUNT_REDO() atomicAdd(&g_pos_redo_rows, 1ull)
#else
static unsigned long long g_pos_redo_rows;                       // the emulator's threads are host threads
#define POS_COUNT_REDO() __atomic_fetch_add(&g_pos_redo_rows, 1ull, __ATOMIC_RELAXED)
#endif

// the sponge over one row; returns whether a deferred wrap fired (never when !DEFER)
#ifndef POS_LEAF_SYNC
#define POS_LEAF_SYNC 1             // the warps of a CTA meet at a barrier before every permutation of the sponge: on long rows they would drift apart and
                                    // stop sharing instruction / constant fetches (see POS_PR_CHUNK).  Rows past the end are clamped, not skipped; the exact
                                    // redo of a flagged row runs after the loop and has no barrier in it
#endif
template <bool LOCKSTEP, bool DEFER, bool SYNC = false>
static __device__ __forceinline__ bool pos_leaf_row(gl_t x[12], const gl_t *__restrict__ p, uint64_t ncols, bool vec16, const gl_t *cap, const PosXs &xs)
{
    gl_t nxt[8];
    GlBorrow bf = gl_borrow_init();
    const uint64_t nblk = (ncols + 7) / 8;
    pos_load8(nxt, p, (uint32_t)(ncols < 8 ? ncols : 8), vec16);
#pragma unroll
    for (int i = 8; i < 12; i++) x[i] = cap ? cap[i - 8] : 0;   // capacity starts at zero (:243-246)
#pragma unroll 1
    for (uint64_t b = 0; b < nblk; b++) {
#if !defined(GLP_EMU)
        if (SYNC) __syncthreads();
#endif
#pragma unroll
        for (int i = 0; i < 8; i++) x[i] = nxt[i];
        if (b + 1 < nblk) {                        // prefetch the next 64 bytes of this row under the permutation
            uint64_t rem = ncols - 8 * (b + 1);
            pos_load8(nxt, p + 8 * (b + 1), (uint32_t)(rem < 8 ? rem : 8), vec16);
        }
        pos_permute_t<LOCKSTEP, DEFER>(x, bf, xs);
#pragma unroll
        for (int i = 0; i < 4; i++) x[8 + i] = x[i];   // next capacity = first 4 outputs (:247-250)
    }
    return DEFER && gl_borrow_hit(bf);
}
static __device__ GLP_NOINLINE void pos_leaf_row_exact(gl_t *o, const gl_t *p, uint64_t ncols, int vec16, const gl_t *cap, PosXs xs)
{
    gl_t x[12];
    pos_leaf_row<false, false>(x, p, ncols, vec16 != 0, cap, xs);
#pragma unroll
    for (int i = 0; i < 4; i++) o[i] = gl_canon(x[i]);
    POS_COUNT_REDO();
}

template <bool LOCKSTEP>
__global__ void POS_LEAF_BOUNDS
k_linear_hash_rows(gl_t *out, const gl_t *__restrict__ in, uint64_t ncols, uint64_t pitch, uint64_t nrows, int vec16, const gl_t *cap_in,
                   uint64_t out_pitch)
{
    POS_XS_DECL(POS_LEAF_BLOCK);
    constexpr bool DEFER = POS_DEFER_BORROW && !LOCKSTEP;     // a redone row would miss the other warps' barriers
    uint64_t row = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = row < nrows;
    constexpr bool SYNC = POS_LEAF_SYNC != 0 && !LOCKSTEP;
    if (!live) {
        if (!LOCKSTEP && !SYNC) return;
        row = nrows - 1;
    }
    const gl_t *p = in + row * pitch;
    gl_t *o = out + row * out_pitch;                  // 4 for a tree's leaf array; wider when the hashes of several column batches sit side by side
    if (ncols <= POS_CAP && !cap_in) {            // size <= CAPACITY: copy and zero-pad, no hashing (:234-239)
        if (live) {
#pragma unroll
            for (int i = 0; i < 4; i++) o[i] = (uint64_t)i < ncols ? gl_canon(p[i]) : 0;
        }
        return;
    }
    const gl_t *cap = cap_in ? cap_in + row * out_pitch : nullptr;
    gl_t x[12];
    const bool bf = pos_leaf_row<LOCKSTEP, DEFER, SYNC>(x, p, ncols, vec16 != 0, cap, xs);
    if (DEFER && bf) {                            // before any store: cap_in may alias out
        if (live) pos_leaf_row_exact(o, p, ncols, vec16, cap, xs);
        return;
    }
    if (live) {
#pragma unroll
        for (int i = 0; i < 4; i++) o[i] = gl_canon(x[i]);
    }
}

// parents[i] = H(children[2i] || children[2i+1] || 0000)[0:4]; the same deferred wrap and exact redo as the leaf rows
static __device__ GLP_NOINLINE void pos_hash_pair_exact(gl_t *dst, const gl_t *src, PosXs xs)
{
    gl_t x[12];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = src[i];
#pragma unroll
    for (int i = 8; i < 12; i++) x[i] = 0;
    pos_permute(x, xs);
#pragma unroll
    for (int i = 0; i < 4; i++) dst[i] = gl_canon(x[i]);
    POS_COUNT_REDO();
}
static __device__ __forceinline__ void pos_hash_pair(gl_t *__restrict__ dst, const gl_t *__restrict__ src, const PosXs &xs)
{
    gl_t x[12];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = src[i];
#pragma unroll
    for (int i = 8; i < 12; i++) x[i] = 0;
    GlBorrow bf = gl_borrow_init();
    pos_permute_t<false, POS_DEFER_BORROW != 0>(x, bf, xs);
    if (POS_DEFER_BORROW && gl_borrow_hit(bf)) {
        pos_hash_pair_exact(dst, src, xs);
        return;
    }
#pragma unroll
    for (int i = 0; i < 4; i++) dst[i] = gl_canon(x[i]);
}

__global__ void __launch_bounds__(POS_BLOCK)
k_tree_level(gl_t *__restrict__ parents, const gl_t *__restrict__ children, uint64_t nparents)
{
    POS_XS_DECL(POS_BLOCK);
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nparents) return;
    pos_hash_pair(parents + 4 * i, children + 8 * i, xs);
}

// One CTA walks the remaining levels: `level` holds n nodes (n a power of two, n <= 2*blockDim.x) followed
// in memory by its parents, their parents, ... (the reference's level-major layout).
#define POS_TOP_THREADS 128
__global__ void __launch_bounds__(POS_TOP_THREADS)
k_tree_top(gl_t *level, uint64_t n)
{
    POS_XS_DECL(POS_TOP_THREADS);
    gl_t *cur = level;
    while (n > 1) {
        uint64_t np = n >> 1;
        gl_t *par = cur + 4 * n;
        if (threadIdx.x < np) pos_hash_pair(par + 4 * threadIdx.x, cur + 8 * threadIdx.x, xs);
        __syncthreads();
        cur = par;
        n = np;
    }
}

__global__ void __launch_bounds__(POS_BLOCK)
k_permute_batch(gl_t *__restrict__ out, const gl_t *__restrict__ in, uint64_t n)
{
    POS_XS_DECL(POS_BLOCK);
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    gl_t x[12];
#pragma unroll
    for (int k = 0; k < 12; k++) x[k] = in[12 * i + k];
    pos_permute(x, xs);
#pragma unroll
    for (int k = 0; k < 12; k++) out[12 * i + k] = gl_canon(x[k]);
}
