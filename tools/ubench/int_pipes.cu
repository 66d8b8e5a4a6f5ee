// Integer-pipe micro-benchmarks for B200 (sm_100a): what one SM can issue per clock for the instruction
// classes the Goldilocks/Poseidon kernels are made of.  Dependent-free chains, no branches inside the
// unrolled body, 8 independent chains per class per thread, 148*8 CTAs x 256 threads.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o int_pipes int_pipes.cu && ./int_pipes
//
// Output: one line per mode: lane-ops/s and lanes/clk/SM at the SM clock read from nvidia-smi-free
// cudaDeviceProp.clockRate is unreliable on B200, so the clock is measured with clock64() in-kernel.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CHAINS 8
#define REPS 16

template <int MODE>
__global__ void __launch_bounds__(256) k_pipe(uint64_t *sink, uint32_t iters, long long *cycles)
{
    uint32_t a = threadIdx.x * 2654435761u + 12345u, b = (blockIdx.x * 40503u + 977u) ^ (threadIdx.x * 97u);
    uint32_t m[CHAINS], s[CHAINS];
    uint64_t w[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; i++) { m[i] = a * (2 * i + 3); s[i] = b ^ (i * 77u); w[i] = ((uint64_t)a << 32) ^ (b * (i + 1)); }
    long long t0 = clock64();
    for (uint32_t it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < REPS; r++) {
#pragma unroll
            for (int i = 0; i < CHAINS; i++) {
                if (MODE == 0) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(m[i]) : "r"(b), "r"(a));
                if (MODE == 1) asm volatile("{ .reg .u32 l,h; mov.b64 {l,h}, %1; mad.wide.u32 %0, l, %2, %0; }" : "+l"(w[i]) : "l"(w[(i + 3) % CHAINS]), "r"(b));
                if (MODE == 2) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(m[i]) : "r"(b), "r"(a));
                if (MODE == 3) asm volatile("{ .reg .u32 t; add.u32 t, %1, %2; add.u32 %0, %0, t; }" : "+r"(s[i]) : "r"(s[(i + 3) % CHAINS]), "r"(a));
                if (MODE == 4) {   // IMAD.WIDE + IADD3, 1:1
                    asm volatile("{ .reg .u32 l,h; mov.b64 {l,h}, %1; mad.wide.u32 %0, l, %2, %0; }" : "+l"(w[i]) : "l"(w[(i + 3) % CHAINS]), "r"(b));
                    asm volatile("{ .reg .u32 t; add.u32 t, %1, %2; add.u32 %0, %0, t; }" : "+r"(s[i]) : "r"(s[(i + 3) % CHAINS]), "r"(a));
                }
                if (MODE == 5) {   // IMAD.WIDE + 2 IADD3
                    asm volatile("{ .reg .u32 l,h; mov.b64 {l,h}, %1; mad.wide.u32 %0, l, %2, %0; }" : "+l"(w[i]) : "l"(w[(i + 3) % CHAINS]), "r"(b));
                    asm volatile("{ .reg .u32 t; add.u32 t, %1, %2; add.u32 %0, %0, t; }" : "+r"(s[i]) : "r"(s[(i + 3) % CHAINS]), "r"(a));
                    asm volatile("{ .reg .u32 t; add.u32 t, %1, %2; add.u32 %0, %0, t; }" : "+r"(m[i]) : "r"(m[(i + 5) % CHAINS]), "r"(b));
                }
                if (MODE == 6) {   // IMAD.LO + IADD3, 1:1
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(m[i]) : "r"(b), "r"(a));
                    asm volatile("{ .reg .u32 t; add.u32 t, %1, %2; add.u32 %0, %0, t; }" : "+r"(s[i]) : "r"(s[(i + 3) % CHAINS]), "r"(a));
                }
                if (MODE == 7) {   // carry chain pair: add.cc + addc  (64-bit add as the kernels do it)
                    asm volatile("{ add.cc.u32 %0, %0, %2; addc.u32 %1, %1, %3; }" : "+r"(m[i]), "+r"(s[i]) : "r"(a), "r"(b));
                }
                if (MODE == 8) {   // IMAD.WIDE with carry-out + consumer (mad.lo.cc / madc.hi.cc / addc)
                    asm volatile("{ mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.cc.u32 %1, %2, %3, %1; addc.u32 %0, %0, 0; }" : "+r"(m[i]), "+r"(s[i]) : "r"(a), "r"(b));
                }
                if (MODE == 9) {   // LOP3 + SHF (funnel shift) pair
                    asm volatile("{ shf.l.wrap.b32 %0, %0, %1, 7; lop3.b32 %1, %1, %0, %2, 0x96; }" : "+r"(m[i]), "+r"(s[i]) : "r"(a));
                }
                if (MODE == 10) {  // IMAD.WIDE + IMAD.LO 1:1 (both fma pipe?)
                    asm volatile("{ .reg .u32 l,h; mov.b64 {l,h}, %1; mad.wide.u32 %0, l, %2, %0; }" : "+l"(w[i]) : "l"(w[(i + 3) % CHAINS]), "r"(b));
                    asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(m[i]) : "r"(b), "r"(a));
                }
                if (MODE == 11) {  // 64-bit add via IMAD.WIDE x*1 + C  vs nothing else
                    asm volatile("{ .reg .u32 l,h; mov.b64 {l,h}, %0; mad.wide.u32 %0, h, 1, %0; }" : "+l"(w[i]) :);
                }
                if (MODE == 12) {  // FFMA reference (fp32 fma, full rate)
                    float f = __uint_as_float(m[i]);
                    asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f) : "f"(__uint_as_float(b)), "f"(__uint_as_float(a)));
                    m[i] = __float_as_uint(f);
                }
                if (MODE == 14 || MODE == 15 || MODE == 16) {  // DFMA alone / DFMA + IMAD.WIDE / DFMA + IMAD.WIDE + IADD3
                    double dd = __longlong_as_double((long long)w[i]);
                    asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(dd) : "d"(__longlong_as_double(0x3FF0000000000001LL)), "d"(__longlong_as_double(0x3FF0000000000003LL)));
                    w[i] = (uint64_t)__double_as_longlong(dd);
                    if (MODE >= 15) asm volatile("{ mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.u32 %1, %2, %3, %1; }" : "+r"(m[i]), "+r"(s[i]) : "r"(a), "r"(b));
                    if (MODE == 16) asm volatile("{ .reg .u32 t; add.u32 t, %1, %2; add.u32 %0, %0, t; }" : "+r"(m[(i + 1) % CHAINS]) : "r"(s[(i + 3) % CHAINS]), "r"(a));
                }
                if (MODE == 17) {  // IMAD.WIDE via mad.lo.cc/madc.hi (register-pair accumulate), alone
                    asm volatile("{ mad.lo.cc.u32 %0, %2, %3, %0; madc.hi.u32 %1, %2, %3, %1; }" : "+r"(m[i]), "+r"(s[i]) : "r"(a), "r"(b));
                }
                if (MODE == 13) {  // IMAD.WIDE (1) + IADD3 (1) + LOP3 (1)
                    asm volatile("{ .reg .u32 l,h; mov.b64 {l,h}, %1; mad.wide.u32 %0, l, %2, %0; }" : "+l"(w[i]) : "l"(w[(i + 3) % CHAINS]), "r"(b));
                    asm volatile("{ .reg .u32 t; add.u32 t, %1, %2; add.u32 %0, %0, t; }" : "+r"(s[i]) : "r"(s[(i + 3) % CHAINS]), "r"(a));
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(m[i]) : "r"(m[(i + 1) % CHAINS]), "r"(a));
                }
            }
        }
    }
    long long t1 = clock64();
    uint64_t r = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; i++) r ^= m[i] ^ s[i] ^ w[i];
    if ((uint32_t)(r ^ (r >> 32)) == iters * 0x9E3779B1u + 7u) sink[0] = r;
    if (blockIdx.x == 0 && threadIdx.x == 0) cycles[0] = t1 - t0;
}

static const char *NAMES[] = {"IMAD.LO", "IMAD.WIDE", "IMAD.HI", "IADD3", "IMAD.WIDE+IADD3", "IMAD.WIDE+2*IADD3", "IMAD.LO+IADD3",
                              "add.cc+addc", "mad.lo.cc+madc.hi.cc+addc", "SHF+LOP3", "IMAD.WIDE+IMAD.LO", "IMAD.WIDE(x*1+C)", "FFMA",
                              "IMAD.WIDE+IADD3+LOP3", "DFMA", "DFMA+IMAD.WIDE", "DFMA+IMAD.WIDE+IADD3", "IMAD.WIDE(mad.lo.cc/madc.hi)"};
static const int OPS[] = {1, 1, 1, 1, 2, 3, 2, 2, 3, 2, 2, 1, 1, 3, 1, 2, 3, 1};

template <int MODE>
static void run(uint64_t *sink, long long *dcyc, int blocks_per_sm)
{
    const unsigned grid = 148 * blocks_per_sm, block = 256;
    const uint32_t iters = 2048;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    k_pipe<MODE><<<grid, block>>>(sink, 64, dcyc);
    cudaEventRecord(a);
    k_pipe<MODE><<<grid, block>>>(sink, iters, dcyc);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    long long cyc = 0;
    cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost);
    const double ops = (double)iters * REPS * CHAINS * OPS[MODE] * grid * block;
    const double clk_ghz = cyc / (ms * 1e6);   // block 0's resident cycles over the kernel time (approximate when waves > 1)
    printf("%-28s %2d CTA/SM  %8.3f ms  %.3e lane-instr/s  %6.1f lanes/clk/SM @1.965GHz   (block0 cycles %lld, ~%.2f GHz)\n", NAMES[MODE],
           blocks_per_sm, ms, ops / (ms * 1e-3), ops / (ms * 1e-3) / 148 / 1.965e9, cyc, clk_ghz);
    cudaEventDestroy(a);
    cudaEventDestroy(b);
}

int main()
{
    uint64_t *sink;
    long long *dcyc;
    cudaMalloc(&sink, 64);
    cudaMalloc(&dcyc, 64);
    for (int bps : {8, 2}) {
        run<0>(sink, dcyc, bps); run<1>(sink, dcyc, bps); run<2>(sink, dcyc, bps); run<3>(sink, dcyc, bps);
        run<4>(sink, dcyc, bps); run<5>(sink, dcyc, bps); run<6>(sink, dcyc, bps); run<7>(sink, dcyc, bps);
        run<8>(sink, dcyc, bps); run<9>(sink, dcyc, bps); run<10>(sink, dcyc, bps); run<11>(sink, dcyc, bps);
        run<12>(sink, dcyc, bps); run<13>(sink, dcyc, bps);
        run<14>(sink, dcyc, bps); run<15>(sink, dcyc, bps); run<16>(sink, dcyc, bps); run<17>(sink, dcyc, bps);
        printf("\n");
    }
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
    return 0;
}
