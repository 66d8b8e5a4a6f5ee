// Host<->device copy rates on this box for the shapes glp_commit's streamed path uses: contiguous pinned copies,
// strided 2-D copies of a column block out of / into a row-major matrix, both directions alone and together,
// and kernel zero-copy stores into mapped pinned memory.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pcie_probe pcie_probe.cu && ./pcie_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

__global__ void k_store_cols(uint64_t *dst, size_t pitch, const uint64_t *src, size_t w, size_t rows)
{
    const size_t total = rows * w;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x)
        dst[(i / w) * pitch + (i % w)] = src[i];
}

int main()
{
    const size_t rows = 1 << 21, C = 660, w = 96;                 // 11 GB matrix, 1.6 GB column block
    uint64_t *h = nullptr, *d = nullptr, *d2 = nullptr;
    CK(cudaHostAlloc(&h, rows * C * 8, cudaHostAllocMapped));
    CK(cudaMalloc(&d, rows * C * 8));
    CK(cudaMalloc(&d2, rows * w * 8));
    for (size_t i = 0; i < rows * C; i += 512) h[i] = i;
    cudaStream_t s0, s1;
    CK(cudaStreamCreate(&s0));
    CK(cudaStreamCreate(&s1));
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    float ms;
    auto rep = [&](const char *name, double bytes) { printf("%-58s %8.2f ms  %7.2f GB/s\n", name, ms, bytes / ms / 1e6); };
    const size_t big = (size_t)4 << 30;
    for (int it = 0; it < 2; it++) {
        CK(cudaEventRecord(a, s0)); CK(cudaMemcpyAsync(d, h, big, cudaMemcpyHostToDevice, s0)); CK(cudaEventRecord(b, s0)); CK(cudaEventSynchronize(b));
        CK(cudaEventElapsedTime(&ms, a, b)); if (it) rep("H2D contiguous 4 GiB pinned", big);
        CK(cudaEventRecord(a, s0)); CK(cudaMemcpyAsync(h, d, big, cudaMemcpyDeviceToHost, s0)); CK(cudaEventRecord(b, s0)); CK(cudaEventSynchronize(b));
        CK(cudaEventElapsedTime(&ms, a, b)); if (it) rep("D2H contiguous 4 GiB pinned", big);
        CK(cudaEventRecord(a, s0));
        CK(cudaMemcpyAsync(d, h, big, cudaMemcpyHostToDevice, s0));
        CK(cudaMemcpyAsync(h + big / 8, d + big / 8, big, cudaMemcpyDeviceToHost, s1));
        CK(cudaStreamSynchronize(s1)); CK(cudaEventRecord(b, s0)); CK(cudaEventSynchronize(b));
        CK(cudaEventElapsedTime(&ms, a, b)); if (it) rep("H2D + D2H concurrently, 4 GiB each (sum)", 2.0 * big);
        CK(cudaEventRecord(a, s0)); CK(cudaMemcpy2DAsync(d2, w * 8, h + 8, C * 8, w * 8, rows, cudaMemcpyHostToDevice, s0)); CK(cudaEventRecord(b, s0));
        CK(cudaEventSynchronize(b)); CK(cudaEventElapsedTime(&ms, a, b)); if (it) rep("H2D 2-D: 96 of 660 columns, 2^21 rows", (double)rows * w * 8);
        CK(cudaEventRecord(a, s0)); CK(cudaMemcpy2DAsync(h + 8, C * 8, d2, w * 8, w * 8, rows, cudaMemcpyDeviceToHost, s0)); CK(cudaEventRecord(b, s0));
        CK(cudaEventSynchronize(b)); CK(cudaEventElapsedTime(&ms, a, b)); if (it) rep("D2H 2-D: 96 of 660 columns, 2^21 rows", (double)rows * w * 8);
        CK(cudaEventRecord(a, s0)); CK(cudaMemcpy2DAsync(h + 8, C * 8, d + 8, C * 8, w * 8, rows, cudaMemcpyDeviceToHost, s0)); CK(cudaEventRecord(b, s0));
        CK(cudaEventSynchronize(b)); CK(cudaEventElapsedTime(&ms, a, b)); if (it) rep("D2H 2-D strided both sides (pitch 660 -> pitch 660)", (double)rows * w * 8);
        uint64_t *hd = nullptr;
        CK(cudaHostGetDevicePointer(&hd, h, 0));
        CK(cudaEventRecord(a, s0)); k_store_cols<<<148 * 8, 256, 0, s0>>>(hd + 8, C, d2, w, rows); CK(cudaEventRecord(b, s0));
        CK(cudaEventSynchronize(b)); CK(cudaEventElapsedTime(&ms, a, b)); if (it) rep("kernel zero-copy stores into mapped host, 96 of 660 cols", (double)rows * w * 8);
        CK(cudaEventRecord(a, s0)); k_store_cols<<<148 * 8, 256, 0, s0>>>(d2, w, hd + 8 - 8, w, rows); CK(cudaEventRecord(b, s0));
        CK(cudaEventSynchronize(b)); CK(cudaEventElapsedTime(&ms, a, b)); if (it) rep("kernel zero-copy loads from mapped host (contiguous)", (double)rows * w * 8);
    }
    // both directions at once, 2-D, for several block widths (what the streamed commit does in steady state)
    for (size_t wb : {64, 128, 256}) {
        uint64_t *d3 = nullptr;
        CK(cudaMalloc(&d3, rows * wb * 8));
        for (int mode = 0; mode < 3; mode++) {
            CK(cudaDeviceSynchronize());
            CK(cudaEventRecord(a, s0));
            CK(cudaMemcpy2DAsync(d3, wb * 8, h + 8, C * 8, wb * 8, rows, cudaMemcpyHostToDevice, s0));                       // up: columns 8..
            if (mode == 0) CK(cudaMemcpy2DAsync(h + 320, C * 8, d + 320, C * 8, wb * 8, rows, cudaMemcpyDeviceToHost, s1));     // down 2-D: columns 320..
            if (mode == 1) CK(cudaMemcpyAsync(h + (rows / 2) * C, d, rows * wb * 8, cudaMemcpyDeviceToHost, s1));               // down contiguous
            if (mode == 2) { CK(cudaMemcpy2DAsync(h + 320, C * 8, d + 320, C * 8, wb * 8, rows / 2, cudaMemcpyDeviceToHost, s1));
                             CK(cudaMemcpy2DAsync(h + 320 + (rows / 2) * C, C * 8, d + 320 + (rows / 2) * C, C * 8, wb * 8, rows / 2, cudaMemcpyDeviceToHost, s0)); }
            CK(cudaStreamSynchronize(s1)); CK(cudaEventRecord(b, s0)); CK(cudaEventSynchronize(b));
            CK(cudaEventElapsedTime(&ms, a, b));
            char nm[128];
            snprintf(nm, sizeof nm, "%s, %zu cols (sum)", mode == 0 ? "H2D 2-D + D2H 2-D concurrently" : mode == 1 ? "H2D 2-D + D2H contiguous" : "H2D 2-D then D2H half, other half on s1", wb);
            rep(nm, 2.0 * rows * wb * 8);
        }
        cudaFree(d3);
    }
    // one block's download split over k streams (row ranges), alone and under a concurrent 2-D upload
    {
        cudaStream_t ss[4];
        for (int i = 0; i < 4; i++) CK(cudaStreamCreateWithFlags(&ss[i], cudaStreamNonBlocking));
        cudaStream_t sup;
        CK(cudaStreamCreateWithFlags(&sup, cudaStreamNonBlocking));
        const size_t wb = 96;
        for (int withup = 0; withup < 2; withup++)
            for (int k : {1, 2, 4}) {
                CK(cudaDeviceSynchronize());
                CK(cudaEventRecord(a, ss[0]));
                for (int i = 1; i < k; i++) CK(cudaStreamWaitEvent(ss[i], a, 0));
                if (withup) { CK(cudaStreamWaitEvent(sup, a, 0)); CK(cudaMemcpy2DAsync(d2, wb * 8, h + 400, C * 8, wb * 8, rows, cudaMemcpyHostToDevice, sup)); }
                for (int i = 0; i < k; i++) {
                    const size_t r0 = rows / k * i, nr = rows / k;
                    CK(cudaMemcpy2DAsync(h + 8 + r0 * C, C * 8, d + 8 + r0 * C, C * 8, wb * 8, nr, cudaMemcpyDeviceToHost, ss[i]));
                }
                for (int i = 1; i < k; i++) { cudaEvent_t e; CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); CK(cudaEventRecord(e, ss[i])); CK(cudaStreamWaitEvent(ss[0], e, 0)); }
                CK(cudaEventRecord(b, ss[0])); CK(cudaEventSynchronize(b)); CK(cudaEventElapsedTime(&ms, a, b));
                char nm[128];
                snprintf(nm, sizeof nm, "D2H 2-D 96 cols over %d stream(s)%s (D2H bytes only)", k, withup ? " + concurrent 2-D upload" : "");
                rep(nm, (double)rows * wb * 8);
                CK(cudaDeviceSynchronize());
            }
    }
    // pageable
    uint64_t *pg = (uint64_t *)malloc(big);
    for (size_t i = 0; i < big / 8; i += 512) pg[i] = i;
    CK(cudaEventRecord(a, s0)); CK(cudaMemcpyAsync(d, pg, big, cudaMemcpyHostToDevice, s0)); CK(cudaEventRecord(b, s0)); CK(cudaEventSynchronize(b));
    CK(cudaEventElapsedTime(&ms, a, b)); rep("H2D contiguous 4 GiB PAGEABLE", big);
    CK(cudaEventRecord(a, s0)); CK(cudaMemcpyAsync(pg, d, big, cudaMemcpyDeviceToHost, s0)); CK(cudaEventRecord(b, s0)); CK(cudaEventSynchronize(b));
    CK(cudaEventElapsedTime(&ms, a, b)); rep("D2H contiguous 4 GiB PAGEABLE", big);
    return 0;
}
