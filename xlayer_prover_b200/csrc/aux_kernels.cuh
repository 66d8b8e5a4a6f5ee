// Small support kernels: synthetic trace generation, checksums, proof gather, field self-test.
#pragma once
#include "cuda_compat.h"
#include "gl_field.cuh"

// counter-based generator: v(i,c) = splitmix64(seed ^ (i*ncols_total + c)) mod p
static __host__ __device__ __forceinline__ gl_t glp_splitmix(uint64_t seed, uint64_t idx)
{
    uint64_t z = (seed ^ idx) + 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z ^= z >> 31;
    return z >= GL_P ? z - GL_P : z;
}

__global__ void k_synth_fill(gl_t *dst, uint64_t nrows, uint64_t ncols, uint64_t row0, uint64_t col0, uint64_t ncols_total, uint64_t seed)
{
    const uint64_t total = nrows * ncols;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r = i / ncols, c = i % ncols;
        dst[i] = glp_splitmix(seed, (row0 + r) * ncols_total + col0 + c);
    }
}

#define GLP_CK_THREADS 256
// partial[2*block] = sum w_i, partial[2*block+1] = sum w_i*(i+1)   (mod p; exact, so order-independent)
__global__ void __launch_bounds__(GLP_CK_THREADS)
k_checksum(const gl_t *src, uint64_t n, gl_t *partial)
{
    __shared__ gl_t s0[GLP_CK_THREADS];
    __shared__ gl_t s1[GLP_CK_THREADS];
    gl_t a = 0, b = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const gl_t w = src[i];
        a = gl_add(a, w);
        b = gl_add(b, gl_mul(w, (i + 1) % GL_P));
    }
    s0[threadIdx.x] = a;
    s1[threadIdx.x] = b;
    __syncthreads();
    for (unsigned h = GLP_CK_THREADS / 2; h > 0; h >>= 1) {
        if (threadIdx.x < h) {
            s0[threadIdx.x] = gl_add(s0[threadIdx.x], s0[threadIdx.x + h]);
            s1[threadIdx.x] = gl_add(s1[threadIdx.x], s1[threadIdx.x + h]);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        partial[2 * blockIdx.x] = gl_canon(s0[0]);
        partial[2 * blockIdx.x + 1] = gl_canon(s1[0]);
    }
}

// MerkleTreeGL::getGroupProof (merkleTreeGL.cpp:12-35): row `idx` of source, then the sibling at every level
__global__ void k_group_proof(gl_t *proof, const gl_t *source, const gl_t *nodes, uint64_t width, uint64_t height, uint64_t idx)
{
    for (uint64_t i = threadIdx.x; i < width; i += blockDim.x) proof[i] = gl_canon(source[idx * width + i]);
    if (threadIdx.x < 4) {
        uint64_t n = height, offset = 0, id = idx, lvl = 0;
        while (n > 1) {
            proof[width + 4 * lvl + threadIdx.x] = nodes[offset + (id ^ 1) * 4 + threadIdx.x];
            offset += 4 * n;
            n >>= 1;
            id >>= 1;
            lvl++;
        }
    }
}

// The same for a batch of query indices (the FRI query phase asks every tree for every query index,
// src/starkpil/fri/friProve.cpp:150-170,219-250): CTA q gathers proof q = [row idx[q] | siblings] into proofs + q * stride.
__global__ void k_group_proofs(gl_t *proofs, uint64_t stride, const gl_t *source, const gl_t *nodes, uint64_t width, uint64_t height,
                               const uint64_t *idx, uint64_t nidx)
{
    const uint64_t q = blockIdx.x;
    if (q >= nidx) return;
    const uint64_t row = idx[q];
    gl_t *proof = proofs + q * stride;
    for (uint64_t i = threadIdx.x; i < width; i += blockDim.x) proof[i] = gl_canon(source[row * width + i]);
    // one thread per (level, word): the path is at most 32 levels x 4 words
    uint64_t depth = 0;
    while (((uint64_t)1 << depth) < height) depth++;
    for (uint64_t t = threadIdx.x; t < 4 * depth; t += blockDim.x) {
        const uint64_t lvl = t >> 2, w = t & 3;
        // offset of level lvl = 4 * (height + height/2 + ... ) = 4 * (2*height - (2*height >> lvl))
        const uint64_t offset = 4 * (2 * height - ((2 * height) >> lvl));
        proof[width + t] = nodes[offset + ((row >> lvl) ^ 1) * 4 + w];
    }
}

// Step 4 of Starks::genProof (src/starkpil/starks.cpp:264-281): the quotient's coefficients qq1 [qdeg*n x qdim] are cut
// into qdeg chunks of n and chunk p, scaled by shift_in^p (shift_in = shift^-n), becomes columns [p*qdim, (p+1)*qdim) of
// the n-row matrix qq2:  qq2[k][p*qdim + d] = qq1[p*n + k][d] * shift_in^p.
__global__ void k_quotient_split(gl_t *qq2, const gl_t *qq1, uint64_t n, uint32_t qdim, uint32_t qdeg, gl_t shift_in)
{
    const uint64_t w = (uint64_t)qdim * qdeg, total = n * w;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t k = i / w, r = i % w, p = r / qdim, d = r % qdim;
        gl_t s = 1;
        for (uint64_t j = 0; j < p; j++) s = gl_mul(s, shift_in);
        qq2[i] = gl_canon(gl_mul(qq1[(p * n + k) * qdim + d], s));
    }
}

// NTT_Goldilocks::reversePermutation (ntt_goldilocks.cpp:243-305), out of place
__global__ void k_reverse_permutation(gl_t *dst, const gl_t *src, uint64_t size, uint32_t logn, uint64_t offset_cols, uint64_t ncols,
                                      uint64_t ncols_all, uint64_t live)
{
    const uint64_t total = size * ncols;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t r = i / ncols, c = i % ncols;
        const uint64_t br = logn ? (uint64_t)(__brev((unsigned)r) >> (32 - logn)) : 0;
        dst[i] = br < live ? src[br * ncols_all + offset_cols + c] : 0;
    }
}

#define GLP_SELFTEST_OUTS 8
__global__ void k_field_selftest(const gl_t *a, const gl_t *b, uint64_t n, gl_t *out)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const gl_t x = a[i], y = b[i];
    out[GLP_SELFTEST_OUTS * i + 0] = gl_canon(gl_add(x, y));
    out[GLP_SELFTEST_OUTS * i + 1] = gl_canon(gl_sub(x, y));
    out[GLP_SELFTEST_OUTS * i + 2] = gl_canon(gl_mul(x, y));
    out[GLP_SELFTEST_OUTS * i + 3] = gl_canon(gl_mul_pow2(x, (unsigned)(y % 96)));
    out[GLP_SELFTEST_OUTS * i + 4] = gl_canon(x);
    out[GLP_SELFTEST_OUTS * i + 5] = gl_canon(gl_sqr(x));
    out[GLP_SELFTEST_OUTS * i + 6] = gl_canon(gl_pow7(y));
    out[GLP_SELFTEST_OUTS * i + 7] = gl_canon(gl_add_c(x, gl_canon(y)));
}

// ---- the reference's Fibonacci test matrix, generated on the device ----
// dst[j*ncols + i] = (i + 1) * F_j mod p with F_0 = F_1 = 1, F_j = F_(j-2) + F_(j-1): the input of the reference's Merkle
// benchmark and tests (src/goldilocks/benchs/bench.cpp:410-427, tests/tests.cpp:1822-1835), whose 2^23 x 128 root is a known
// answer (bench.cpp:440-443).  One thread owns one column of one chunk of rows: it jumps to the chunk's first row with the
// fast-doubling identities fib(2k) = fib(k) (2 fib(k+1) - fib(k)), fib(2k+1) = fib(k)^2 + fib(k+1)^2 and then walks the recurrence.
#define GLP_FIB_CHUNK 512
__global__ void k_fib_fill(gl_t *dst, uint64_t nrows, uint64_t ncols)
{
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t nchunks = (nrows + GLP_FIB_CHUNK - 1) / GLP_FIB_CHUNK;
    if (t >= nchunks * ncols) return;
    const uint64_t col = t % ncols, j0 = (t / ncols) * GLP_FIB_CHUNK;
    // (a, b) = (fib(m), fib(m+1)) for m = j0 + 1  (standard fib(0) = 0, fib(1) = 1; F_j = fib(j + 1))
    const uint64_t m = j0 + 1;
    gl_t a = 0, b = 1;
    for (int bit = 63; bit >= 0; bit--) {
        const gl_t c = gl_mul(a, gl_sub(gl_add(b, b), a));         // fib(2k)
        const gl_t d = gl_add(gl_mul(a, a), gl_mul(b, b));         // fib(2k+1)
        a = c;
        b = d;
        if ((m >> bit) & 1) { const gl_t e = gl_add(a, b); a = b; b = e; }
    }
    const gl_t k = col + 1;
    gl_t v0 = gl_mul(a, k), v1 = gl_mul(b, k);
    const uint64_t j1 = j0 + GLP_FIB_CHUNK < nrows ? j0 + GLP_FIB_CHUNK : nrows;
    for (uint64_t j = j0; j < j1; j++) {
        dst[j * ncols + col] = gl_canon(v0);
        const gl_t nx = gl_add(v0, v1);
        v0 = v1;
        v1 = nx;
    }
}

// ---- integer-pipe peak micro-benchmarks ("the IMAD peak must be measured on the box") ----
// MODE 0: 8 independent IMAD (32-bit multiply-add, same pipe as IMAD.WIDE) chains per thread; 1: 8 independent IADD3 chains; 2: both
// interleaved; 3: 8 independent IMAD.WIDE chains (32x32 -> 64 with a 64-bit accumulate, written as the mad.lo.cc / madc.hi pair
// ptxas keeps as ONE instruction); 4: 8 independent DFMA chains.  3 and 4 are the two instruction classes the algorithmic floor
// of the Poseidon permutation is counted in (bench.py roofline).  Each loop iteration issues 8 (16 in mode 2) dependent-free
// instructions per thread.  MODE is a template parameter: with a run-time mode every flavour shared one register allocation and
// one schedule, and adding a mode changed the numbers of the others.
template <int MODE>
__global__ void k_int_peak(uint64_t *sink, uint32_t iters)
{
    uint32_t a = threadIdx.x * 2654435761u + 12345u, b = (blockIdx.x * 40503u + 977u) ^ (threadIdx.x * 97u);
    uint32_t m0 = a, m1 = b, m2 = a ^ b, m3 = a + b, m4 = a * 3u, m5 = b * 5u, m6 = a * 7u, m7 = b * 11u;
    uint32_t s0 = a, s1 = b, s2 = a ^ 1u, s3 = b ^ 2u, s4 = a ^ 3u, s5 = b ^ 4u, s6 = a ^ 5u, s7 = b ^ 6u;
    double d0 = (double)a, d1 = (double)b, d2 = d0 + 1.0, d3 = d1 + 2.0, d4 = d0 + 3.0, d5 = d1 + 4.0, d6 = d0 + 5.0, d7 = d1 + 6.0;
    const double dm = 1.0000000001, dc = 3.0;
    // every operand is loop-carried, so ptxas can neither hoist the product nor fold the adds
#if GL_DEVICE_PTX
#define GLP_MADW(acc, y) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(acc) : "r"(y), "r"(a))
#define GLP_ADD3(acc, x, y) asm volatile("{ .reg .u32 t; add.u32 t, %1, %2; add.u32 %0, %0, t; }" : "+r"(acc) : "r"(x), "r"(y))
#define GLP_MADWIDE(lo, hi, y) asm volatile("{ mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.u32 %1, %2, %3, %1; }" : "+r"(lo), "+r"(hi) : "r"(y), "r"(a))
#define GLP_DFMA(acc) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(acc) : "d"(dm), "d"(dc))
#else
#define GLP_MADW(acc, y) acc = acc * (y) + a
#define GLP_ADD3(acc, x, y) acc += (x) + (y)
#define GLP_MADWIDE(lo, hi, y) { uint64_t t_ = (((uint64_t)(hi) << 32) | (lo)) + (uint64_t)(y) * a; lo = (uint32_t)t_; hi = (uint32_t)(t_ >> 32); }
#define GLP_DFMA(acc) acc = acc * dm + dc
#endif
    for (uint32_t i = 0; i < iters; i++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
            if (MODE == 0 || MODE == 2) {
                GLP_MADW(m0, b); GLP_MADW(m1, b); GLP_MADW(m2, b); GLP_MADW(m3, b);
                GLP_MADW(m4, b); GLP_MADW(m5, b); GLP_MADW(m6, b); GLP_MADW(m7, b);
            }
            if (MODE == 1 || MODE == 2) {
                GLP_ADD3(s0, s4, s5); GLP_ADD3(s1, s5, s6); GLP_ADD3(s2, s6, s7); GLP_ADD3(s3, s7, s4);
                GLP_ADD3(s4, s0, s1); GLP_ADD3(s5, s1, s2); GLP_ADD3(s6, s2, s3); GLP_ADD3(s7, s3, s0);
            }
            if (MODE == 3) {
                GLP_MADWIDE(m0, s0, b); GLP_MADWIDE(m1, s1, b); GLP_MADWIDE(m2, s2, b); GLP_MADWIDE(m3, s3, b);
                GLP_MADWIDE(m4, s4, b); GLP_MADWIDE(m5, s5, b); GLP_MADWIDE(m6, s6, b); GLP_MADWIDE(m7, s7, b);
            }
            if (MODE == 4) {
                GLP_DFMA(d0); GLP_DFMA(d1); GLP_DFMA(d2); GLP_DFMA(d3); GLP_DFMA(d4); GLP_DFMA(d5); GLP_DFMA(d6); GLP_DFMA(d7);
            }
        }
    }
    const double ds = d0 + d1 + d2 + d3 + d4 + d5 + d6 + d7;
    const uint32_t r = m0 ^ m1 ^ m2 ^ m3 ^ m4 ^ m5 ^ m6 ^ m7 ^ s0 ^ s1 ^ s2 ^ s3 ^ s4 ^ s5 ^ s6 ^ s7;
    if (r == iters * 0x9E3779B1u + 7u || (MODE == 4 && ds == 0.125)) sink[0] = r;   // practically never true; keeps the chains alive
}
