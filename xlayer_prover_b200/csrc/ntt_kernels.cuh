// K1-K5 -- Goldilocks NTT / INTT / low-degree extension over row-major [rows x cols] matrices.
//
// Replaces NTT_Goldilocks::{NTT, INTT, extendPol, reversePermutation, computeR} and the twiddle tables of its
// constructor (src/goldilocks/src/ntt_goldilocks.hpp:49-136,153-165; ntt_goldilocks.cpp:12-352).  Same maths,
// different algorithm:
//
//   out[k,:] = sum_i in[i,:] * w^(i k)          (natural order in and out, w = W[log2 N])
//   inverse  : the same forward transform, written to row (N-k) mod N and scaled by 1/N
//              (* shift^row when `extend`), exactly the reference's last inverse phase
//              (ntt_goldilocks.cpp:122-150) -- so no inverse twiddles exist anywhere.
//
// B200 design (see DESIGN.md "LDE"):
//  * The reference layout (element (i,c) at a[i*ncols+c]) is already batch-contiguous: adjacent columns are
//    adjacent words.  A tile = Q rows (Q <= 4096) x CB adjacent (virtual) columns, so every global access is a
//    full 32-byte sector per row and NO transposes or bit-reversal passes over HBM exist at all.  A row
//    permutation costs nothing at sector granularity, so bit reversal is folded into where a tile stores.
//  * Decimation in frequency, mixed radix.  log2 N is split into at most three tile passes of <= 12 bits
//    (2^24 = 2^12 * 2^12: two passes; HBM traffic 2 reads + 2 writes per element).  Pass t gathers rows
//    b*L + p + m*(L/Q) (m < Q), runs a Q-point transform in shared memory, multiplies output u by w_L^(p u)
//    and writes it back to row b*L + u*(L/Q) + p -- in place on the row set it read.  The last pass (L == Q)
//    scatters to the natural row index.  "Virtual columns" v = p*wb + col make strided passes look like a
//    plain [Q x (L/Q)*wb] matrix, so narrow inputs (1 column) still move whole sectors.
//  * Inside a tile the Q-point transform is radix-16 steps held in registers (16 values per thread, one
//    shared-memory round trip per 4 butterfly stages).  Because W[6] = 8, every twiddle INSIDE a 16-point
//    butterfly is a power of two: gl_mul_pow2 (shifts + one fold), no multiplier, no table.
//    Only the step-to-step twiddles w_Ls^(p u) are real multiplies; they come from per-step tables laid out
//    [u][p] so that a warp reads consecutive words.  Shared rows are padded by one row per 16 so the last
//    step's stride-16 row accesses stay conflict-free.
//  * Zero padding of the LDE (reversePermutation's zero-fill, ntt_goldilocks.cpp:259-277) is the `live`
//    row bound of the first pass: rows >= live are never read.
//  * Pass-to-pass twiddles w_N^e and the coset scaling shift^j / N come from two-level tables
//    (hi[e >> 12] * lo[e & 4095], 2 x 32 KB for N = 2^24) instead of the reference's N-entry arrays
//    (134 MB at 2^24, rebuilt on every extendPol call, ntt_goldilocks.cpp:328).
#pragma once
#include "cuda_compat.h"
#include "gl_field.cuh"

#define NTT_MAX_LOGQ 12
#define NTT_LOGT 12                 // two-level tables: low 12 bits / the rest
#define NTT_THREADS 256
#define NTT_MAX_STEPS 4

struct NttPass {
    const gl_t *in;
    gl_t *out;
    uint64_t in_pitch, out_pitch;    // words per row
    uint64_t in_col0, out_col0;      // first column of this block inside a row of in / out
    uint64_t live;                   // rows >= live of `in` are zero (first pass of an extension)
    uint32_t wb;                     // columns in this block
    uint32_t logN, logL, logQ;       // whole transform, sub-transform this pass splits, tile size
    uint32_t last;                   // last pass: scatter to natural order (+ inverse handling)
    uint32_t canon_in;               // first pass: inputs may be non-canonical
    uint32_t inverse, extend;
    uint32_t nprev, logq_prev0, logq_prev1;   // tile sizes of the earlier passes (for the output index)
    uint32_t nsteps, step_log[NTT_MAX_STEPS]; // radix (log2) of each in-tile step
    const gl_t *step_tw;             // per-step twiddles, concatenated, step s laid out [u][p]
    const gl_t *tw_lo, *tw_hi;       // w_N^e  = tw_hi[e >> 12] * tw_lo[e & 4095]
    const gl_t *sc_lo, *sc_hi;       // shift^j / N = sc_hi[j >> 12] * sc_lo[j & 4095]   (extend)
    gl_t ninv;                       // 1/N (inverse without extend)
};

// ---- in-register 2^LOG-point DIF transform, twiddles are powers of two ----
// v[i] <- sum_m v[m] * w^(m * bitrev(i)),  w = W[LOG] = 2^(192 / 2^LOG)   (W[4]=2^12, W[3]=2^24, W[2]=2^48, W[1]=2^96=-1)
template <int LOG>
static __device__ __forceinline__ void ntt_dft_regs(gl_t *v)
{
    constexpr int Q = 1 << LOG;
#pragma unroll
    for (int t = 0; t < LOG; t++) {
        const int h = Q >> (t + 1);                 // half size of this stage's butterflies
        const int tshift = (96 / h);                // w_(2h) = 2^(96/h)
#pragma unroll
        for (int g = 0; g < Q; g += 2 * h) {
#pragma unroll
            for (int j = 0; j < h; j++) {
                gl_t a = v[g + j], b = v[g + j + h];
                v[g + j] = gl_add(a, b);
                gl_t d = gl_sub(a, b);
                v[g + j + h] = (j == 0) ? d : gl_mul_pow2(d, (unsigned)(tshift * j));
            }
        }
    }
}

static __device__ __forceinline__ uint32_t ntt_bitrev(uint32_t x, uint32_t bits) { return bits ? (__brev(x) >> (32 - bits)) : 0; }

// physical shared-memory row of logical row r: one pad row per 16
static __device__ __forceinline__ uint32_t ntt_phys(uint32_t r) { return r + (r >> 4); }

template <int LOG, int CB>
static __device__ __forceinline__ void ntt_tile_step(gl_t *data, const gl_t *tw, uint32_t logQ, uint32_t logLs, bool last_step)
{
    constexpr int QS = 1 << LOG;
    const uint32_t logStride = logLs - LOG;          // Ls / Qs
    const uint32_t stride = 1u << logStride;
    const uint32_t ngroups = ((1u << logQ) >> LOG) * CB;
    for (uint32_t g = threadIdx.x; g < ngroups; g += blockDim.x) {
        const uint32_t c = g % CB, rest = g / CB;
        const uint32_t pp = rest & (stride - 1), bb = rest >> logStride;
        const uint32_t base = (bb << logLs) + pp;
        gl_t v[QS];
#pragma unroll
        for (int m = 0; m < QS; m++) v[m] = data[ntt_phys(base + ((uint32_t)m << logStride)) * CB + c];
        ntt_dft_regs<LOG>(v);
#pragma unroll
        for (int i = 0; i < QS; i++) {
            const uint32_t u = ntt_bitrev((uint32_t)i, LOG);   // v[i] holds output u
            gl_t x = v[i];
            if (!last_step && u != 0) x = gl_mul(x, tw[(u << logStride) + pp]);
            data[ntt_phys(base + (u << logStride)) * CB + c] = x;
        }
    }
}

// w_N^e from the two-level table
static __device__ __forceinline__ gl_t ntt_tw2(const gl_t *lo, const gl_t *hi, uint64_t e)
{
    gl_t l = __ldg(lo + (e & ((1u << NTT_LOGT) - 1)));
    uint64_t h = e >> NTT_LOGT;
    return h ? gl_mul(l, __ldg(hi + h)) : l;
}

// One tile pass.  grid.x = (N / Q) * ceil((L/Q) * wb / CB)
template <int CB>
__global__ void __launch_bounds__(NTT_THREADS)
k_ntt_tile(NttPass a)
{
    GLP_DYN_SMEM(gl_t, smem);
    const uint32_t Q = 1u << a.logQ;
    const uint32_t logS = a.logL - a.logQ;                       // stride L/Q between a tile's rows
    const uint64_t S = 1ull << logS;
    const uint64_t nvirt = S * a.wb;                             // virtual columns per sub-block
    const uint64_t nvg = (nvirt + CB - 1) / CB;
    const uint64_t tile = blockIdx.x;
    const uint64_t b = tile / nvg, vg = tile % nvg;
    gl_t *data = smem;                                           // [(Q + Q/16) rows][CB]
    gl_t *stw = smem + (size_t)(Q + (Q >> 4) + 1) * CB;          // step twiddles

    // step twiddles -> shared
    uint32_t ntw = 0;
    {
        uint32_t lg = a.logQ;
        for (uint32_t s = 0; s + 1 < a.nsteps; s++) { ntw += 1u << lg; lg -= a.step_log[s]; }
    }
    for (uint32_t i = threadIdx.x; i < ntw; i += blockDim.x) stw[i] = __ldg(a.step_tw + i);

    // this thread's column is fixed: blockDim.x is a multiple of CB
    const uint32_t c = threadIdx.x % CB;
    const uint64_t v = vg * CB + c;
    const bool colok = v < nvirt;
    const uint64_t p = colok ? v / a.wb : 0;
    const uint32_t col = colok ? (uint32_t)(v % a.wb) : 0;
    const uint64_t row0 = (b << a.logL) + p;

    // gather rows row0 + m*S
    for (uint32_t m = threadIdx.x / CB; m < Q; m += blockDim.x / CB) {
        gl_t x = 0;
        const uint64_t row = row0 + ((uint64_t)m << logS);
        if (colok && row < a.live) {
            x = a.in[row * a.in_pitch + a.in_col0 + col];   // plain load: middle passes run in place
            if (a.canon_in) x = gl_canon(x);
        }
        data[ntt_phys(m) * CB + c] = x;
    }
    __syncthreads();

    // Q-point DIF transform in shared memory, radix 2^step_log[s] per step
    {
        uint32_t lg = a.logQ;
        const gl_t *tw = stw;
        for (uint32_t s = 0; s < a.nsteps; s++) {
            const bool last_step = (s + 1 == a.nsteps);
            switch (a.step_log[s]) {
            case 4: ntt_tile_step<4, CB>(data, tw, a.logQ, lg, last_step); break;
            case 3: ntt_tile_step<3, CB>(data, tw, a.logQ, lg, last_step); break;
            case 2: ntt_tile_step<2, CB>(data, tw, a.logQ, lg, last_step); break;
            default: ntt_tile_step<1, CB>(data, tw, a.logQ, lg, last_step); break;
            }
            tw += 1u << lg;
            lg -= a.step_log[s];
            __syncthreads();
        }
    }

    // scatter: shared position j holds tile output k = digit-reversed(j)
    if (!colok) return;
    const uint64_t N = 1ull << a.logN;
    uint64_t kbase = 0;                                          // natural index contributed by earlier passes
    if (a.last) {
        if (a.nprev == 1) kbase = b;
        else if (a.nprev == 2) kbase = (b >> a.logq_prev1) + ((b & ((1ull << a.logq_prev1) - 1)) << a.logq_prev0);
    }
    const uint32_t logNQ = a.logN - a.logQ;
    for (uint32_t j = threadIdx.x / CB; j < Q; j += blockDim.x / CB) {
        // j = sum_s u_s * (L_s/Q_s)  ->  k = u_0 + Q_0 u_1 + Q_0 Q_1 u_2 ...
        uint32_t k = 0, lg = a.logQ, sh = 0;
        for (uint32_t s = 0; s < a.nsteps; s++) {
            lg -= a.step_log[s];
            const uint32_t u = (j >> lg) & ((1u << a.step_log[s]) - 1);
            k |= u << sh;
            sh += a.step_log[s];
        }
        gl_t x = data[ntt_phys(j) * CB + c];
        if (!a.last) {
            // w_L^(p k) = w_N^(p k N/L)
            const uint64_t e = ((p * k) << (a.logN - a.logL)) & (N - 1);
            if (e) x = gl_mul(x, ntt_tw2(a.tw_lo, a.tw_hi, e));
            const uint64_t row = (b << a.logL) + ((uint64_t)k << logS) + p;
            a.out[row * a.out_pitch + a.out_col0 + col] = x;
        } else {
            uint64_t K = kbase + ((uint64_t)k << logNQ);
            if (a.inverse) {
                K = (N - K) & (N - 1);
                x = gl_mul(x, a.extend ? ntt_tw2(a.sc_lo, a.sc_hi, K) : a.ninv);
            }
            a.out[K * a.out_pitch + a.out_col0 + col] = gl_canon(x);
        }
    }
}

// ---- table generation (K1) ----
// out[i] = c * g^(i * mul)
__global__ void k_ntt_pow_table(gl_t *out, gl_t c, gl_t g, uint64_t mul, uint64_t n)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = gl_canon(gl_mul(c, gl_pow(g, i * mul)));
}
// step table for a sub-transform of length 2^logLs split by radix 2^logQs: out[u*stride + p] = w_Ls^(p u)
__global__ void k_ntt_step_table(gl_t *out, uint32_t logLs, uint32_t logQs)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (1u << logLs)) return;
    const uint32_t logStride = logLs - logQs;
    const uint32_t u = i >> logStride, p = i & ((1u << logStride) - 1);
    out[i] = gl_canon(gl_pow(gl_root_of_unity(logLs), (uint64_t)u * p));
}
