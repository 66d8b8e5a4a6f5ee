// FRI folding step and layer regrouping -- the consumer of the committed polynomials, kept on the
// device so that neither the extended matrix nor the FRI polynomial has to travel to the host between layers.
//
//  k_fri_fold      : FRIProve::prove, the si > 0 branch of the step loop (src/starkpil/fri/friProve.cpp:86-106):
//                    one thread = one output g.  Gathers friPol[i*pol2N + g] (i < nX), interpolates them (INTT of
//                    size nX on the three base-field coordinates), scales coefficient i by sinv_g^i,
//                    sinv_g = shiftInv * w(polBits)^-g (polMulAxi, :192-200), and evaluates at special_x in the cubic
//                    extension F_p[X]/(X^3 - X - 1) (evalPol, :201-217).
//  k_fri_transpose : FRIProve::getTransposed (:252-270), aux[i*h + j] = pol[j*w + i] on three-word elements.
//
// HBM-bound in principle (24 bytes read per input element, once); the gather of a thread is strided by pol2N
// elements but neighbouring threads read neighbouring elements, so every warp access is one contiguous 768-byte run.
#pragma once
#include "cuda_compat.h"
#include "gl_field.cuh"

#define FRI_MAX_RED_BITS 6          // nX <= 64 values per output (the reference's starkStructs fold by 2^3 .. 2^5)

// a * b in F_p[X]/(X^3 - X - 1); the reference's Karatsuba form (goldilocks_cubic_extension.hpp:1873-1886)
static __host__ __device__ __forceinline__ void gl3_mul(gl_t r[3], const gl_t a[3], const gl_t b[3])
{
    const gl_t A = gl_mul(gl_add(a[0], a[1]), gl_add(b[0], b[1]));
    const gl_t B = gl_mul(gl_add(a[0], a[2]), gl_add(b[0], b[2]));
    const gl_t C = gl_mul(gl_add(a[1], a[2]), gl_add(b[1], b[2]));
    const gl_t D = gl_mul(a[0], b[0]), E = gl_mul(a[1], b[1]), F = gl_mul(a[2], b[2]);
    const gl_t G = gl_sub(D, E);
    r[0] = gl_sub(gl_add(C, G), F);
    r[1] = gl_sub(gl_sub(gl_sub(gl_add(A, C), E), E), D);
    r[2] = gl_sub(B, G);
}

struct FriFold {
    gl_t *out;                       // [pol2N x 3]
    const gl_t *in;                  // [2^pol_bits x 3]
    uint32_t pol_bits, red_bits;     // nX = 2^red_bits, pol2N = 2^(pol_bits - red_bits)
    gl_t x[3];                       // special_x
    gl_t shift_inv;                  // the layer's coset shift, inverted
    gl_t wi;                         // w(pol_bits)^-1
    gl_t wnx_inv;                    // w(red_bits)^-1
    gl_t nx_inv;                     // 1 / nX
};

// The nX x 3 values of a thread live in shared memory, element (d, i) of thread t at v[(d*nX + i)*FRI_THREADS + t]
// (consecutive threads in consecutive words: conflict-free), because indexing them with the loop variables of the
// butterfly network would put a per-thread array in local memory -- a first version did, and moved 13x the
// algorithmic bytes through DRAM (profiles/r01_launches_widened.txt).  64 threads x 64 x 3 words = 96 KB at nX = 64.
#define FRI_THREADS 64
static inline size_t fri_fold_smem(uint32_t red_bits) { return ((size_t)3 * FRI_THREADS + 1) * ((size_t)1 << red_bits) * sizeof(gl_t); }

__global__ void __launch_bounds__(FRI_THREADS)
k_fri_fold(FriFold a)
{
    GLP_DYN_SMEM(gl_t, sm);
    const uint64_t pol2n = 1ull << (a.pol_bits - a.red_bits);
    const uint32_t nx = 1u << a.red_bits, t = threadIdx.x;
    gl_t *v = sm + t;                                             // v[(d*nx + i) * FRI_THREADS]
    gl_t *tw = sm + (size_t)3 * nx * FRI_THREADS;                 // w_nX^-j, j < nX/2, shared by the CTA
    uint64_t g = (uint64_t)blockIdx.x * FRI_THREADS + t;
    const bool live = g < pol2n;
    if (!live) g = pol2n - 1;
#define FRI_V(d, i) v[((size_t)(d) * nx + (i)) * FRI_THREADS]
    // gather, bit-reversed, so that the decimation-in-time loop below ends in natural order
    for (uint32_t i = 0; i < nx; i++) {
        const uint32_t r = a.red_bits ? (__brev(i) >> (32 - a.red_bits)) : 0;
        const gl_t *p = a.in + ((uint64_t)i * pol2n + g) * 3;
        FRI_V(0, r) = p[0]; FRI_V(1, r) = p[1]; FRI_V(2, r) = p[2];
    }
    if (t == 0) {
        gl_t w = 1;
        for (uint32_t j = 0; j < nx / 2; j++) { tw[j] = w; w = gl_mul(w, a.wnx_inv); }
    }
    __syncthreads();
    // c[i] = 1/nX * sum_k v[k] * w_nX^(-i k)   (NTT_Goldilocks::INTT, ntt_goldilocks.cpp:307-324)
    for (uint32_t half = 1, s = a.red_bits; half < nx; half <<= 1) {
        s--;                                                      // twiddle stride: w_(2 half)^-j = w_nX^-(j << s)
        for (uint32_t base = 0; base < nx; base += 2 * half)
            for (uint32_t j = 0; j < half; j++) {
                const gl_t w = tw[j << s];
#pragma unroll
                for (int d = 0; d < 3; d++) {
                    const gl_t u = FRI_V(d, base + j), x = gl_mul(FRI_V(d, base + j + half), w);
                    FRI_V(d, base + j) = gl_add(u, x);
                    FRI_V(d, base + j + half) = gl_sub(u, x);
                }
            }
    }
    // coefficient i times (1/nX) * sinv_g^i (polMulAxi), then Horner at special_x from the top coefficient down (evalPol)
    const gl_t sinv = gl_mul(a.shift_inv, gl_pow(a.wi, g));
    gl_t r = a.nx_inv;
    for (uint32_t i = 0; i < nx; i++) {
#pragma unroll
        for (int d = 0; d < 3; d++) FRI_V(d, i) = gl_mul(FRI_V(d, i), r);
        r = gl_mul(r, sinv);
    }
    gl_t res[3] = {FRI_V(0, nx - 1), FRI_V(1, nx - 1), FRI_V(2, nx - 1)};
    for (int i = (int)nx - 2; i >= 0; i--) {
        gl_t m[3];
        gl3_mul(m, res, a.x);
#pragma unroll
        for (int d = 0; d < 3; d++) res[d] = gl_add(m[d], FRI_V(d, i));
    }
#undef FRI_V
    if (live) {
#pragma unroll
        for (int d = 0; d < 3; d++) a.out[g * 3 + d] = gl_canon(res[d]);
    }
}

// aux[(i*h + j)*3 ..] = pol[(j*w + i)*3 ..],  i < w = 2^next_bits, j < h = pol_n / w
__global__ void k_fri_transpose(gl_t *aux, const gl_t *pol, uint64_t pol_n, uint32_t next_bits)
{
    const uint64_t w = 1ull << next_bits, h = pol_n >> next_bits;
    for (uint64_t e = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; e < pol_n; e += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t i = e / h, j = e % h;
        const gl_t *s = pol + (j * w + i) * 3;
        gl_t *d = aux + e * 3;
        d[0] = gl_canon(s[0]); d[1] = gl_canon(s[1]); d[2] = gl_canon(s[2]);
    }
}
