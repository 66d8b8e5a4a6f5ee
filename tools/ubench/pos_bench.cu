// Leaf-hash (Poseidon sponge) kernel bench for variant exploration: compiles the product's kernel sources
// with -D knobs (POS_SBOX_GROUP, POS_ROLL_PRE, POS_LEAF_BLOCK, POS_LEAF_LOCKSTEP, ...) and prints
// permutations/s plus an order-sensitive checksum of the output (must be identical across variants).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I../../xlayer_prover_b200/csrc -D... -o pos_bench pos_bench.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "poseidon_kernels.cuh"
#include "aux_kernels.cuh"

int main(int argc, char **argv)
{
    const uint64_t nrows = argc > 1 ? strtoull(argv[1], 0, 0) : (1ull << 20);
    const uint64_t ncols = argc > 2 ? strtoull(argv[2], 0, 0) : 64;
    const char *tag = argc > 3 ? argv[3] : "";
    gl_t *in, *out;
    cudaMalloc(&in, nrows * ncols * 8);
    cudaMalloc(&out, nrows * 4 * 8);
    k_synth_fill<<<148 * 32, 256>>>(in, nrows, ncols, 0, 0, ncols, 0x1234567ULL);
    const unsigned nblk = (unsigned)((nrows + POS_LEAF_BLOCK - 1) / POS_LEAF_BLOCK);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    float best = 1e30f;
    for (int rep = 0; rep < 6; rep++) {
        cudaEventRecord(a);
        k_linear_hash_rows<POS_LEAF_LOCKSTEP != 0><<<nblk, POS_LEAF_BLOCK>>>(out, in, ncols, ncols, nrows, 1, nullptr, 4);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms;
        cudaEventElapsedTime(&ms, a, b);
        if (rep > 0 && ms < best) best = ms;
    }
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s CUDA error: %s\n", tag, cudaGetErrorString(e)); return 1; }
    std::vector<gl_t> h(nrows * 4);
    cudaMemcpy(h.data(), out, nrows * 32, cudaMemcpyDeviceToHost);
    gl_t ck = 0;
    for (uint64_t i = 0; i < nrows * 4; i++) ck = gl_add(gl_mul(ck, 0x9E3779B97F4A7C15ULL % GL_P), h[i]);
    const double perms = (double)nrows * ((ncols + 7) / 8);
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, k_linear_hash_rows<POS_LEAF_LOCKSTEP != 0>);
    unsigned long long redo = 0;
    cudaMemcpyFromSymbol(&redo, g_pos_redo_rows, sizeof redo);
    printf("%-44s block %4d regs %3d  %8.3f ms  %7.1f Mperm/s  checksum %016llx  redone rows %llu (6 launches)\n", tag, POS_LEAF_BLOCK,
           fa.numRegs, best, perms / best / 1e3, (unsigned long long)gl_canon(ck), redo);
    return 0;
}
