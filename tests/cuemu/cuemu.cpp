// TEST/DEBUG HARNESS ONLY -- see cuemu.h.
#include "cuemu.h"
#include <pthread.h>
#include <stdio.h>
#include <time.h>
#include <atomic>
#include <map>
#include <mutex>
#include <thread>
#include <vector>

namespace cuemu {
thread_local dim3 t_threadIdx;
thread_local dim3 t_blockIdx;
dim3 g_blockIdx, g_blockDim, g_gridDim;
unsigned char *g_dynsmem = nullptr;
static pthread_barrier_t g_bar;
static std::mutex g_launch_mutex;

void barrier() { pthread_barrier_wait(&g_bar); }

struct BlockArgs { const std::function<void()> *body; unsigned tid; dim3 bidx; };
static void *thread_main(void *p)
{
    BlockArgs *a = (BlockArgs *)p;
    unsigned t = a->tid;
    t_threadIdx = dim3(t % g_blockDim.x, (t / g_blockDim.x) % g_blockDim.y, t / (g_blockDim.x * g_blockDim.y));
    t_blockIdx = a->bidx;
    (*a->body)();
    return nullptr;
}

void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()> &body)
{
    std::lock_guard<std::mutex> lk(g_launch_mutex);
    unsigned nthreads = block.x * block.y * block.z;
    g_blockDim = block;
    g_gridDim = grid;
    std::vector<unsigned char> dyn(smem + 64);
    g_dynsmem = (unsigned char *)(((uintptr_t)dyn.data() + 63) & ~(uintptr_t)63);
    pthread_attr_t attr;
    pthread_attr_init(&attr);
    pthread_attr_setstacksize(&attr, 256 * 1024);
    std::vector<pthread_t> th(nthreads);
    std::vector<BlockArgs> args(nthreads);
    for (unsigned bz = 0; bz < grid.z; bz++)
        for (unsigned by = 0; by < grid.y; by++)
            for (unsigned bx = 0; bx < grid.x; bx++) {
                pthread_barrier_init(&g_bar, nullptr, nthreads);
                for (unsigned t = 0; t < nthreads; t++) {
                    args[t] = BlockArgs{&body, t, dim3(bx, by, bz)};
                    if (pthread_create(&th[t], &attr, thread_main, &args[t])) { fprintf(stderr, "cuemu: pthread_create failed\n"); abort(); }
                }
                for (unsigned t = 0; t < nthreads; t++) pthread_join(th[t], nullptr);
                pthread_barrier_destroy(&g_bar);
            }
    pthread_attr_destroy(&attr);
    g_dynsmem = nullptr;
}

void launch_flat(dim3 grid, dim3 block, const std::function<void()> &body)
{
    std::lock_guard<std::mutex> lk(g_launch_mutex);
    g_blockDim = block;
    g_gridDim = grid;
    uint64_t nblocks = (uint64_t)grid.x * grid.y * grid.z;
    unsigned nthreads = block.x * block.y * block.z;
    unsigned nw = std::min<uint64_t>(std::max(1u, std::thread::hardware_concurrency()), nblocks);
    std::atomic<uint64_t> next(0);
    std::vector<std::thread> ws;
    for (unsigned w = 0; w < nw; w++)
        ws.emplace_back([&]() {
            for (;;) {
                uint64_t b = next.fetch_add(1);
                if (b >= nblocks) break;
                t_blockIdx = dim3(b % grid.x, (b / grid.x) % grid.y, b / ((uint64_t)grid.x * grid.y));
                for (unsigned t = 0; t < nthreads; t++) {
                    t_threadIdx = dim3(t % block.x, (t / block.x) % block.y, t / (block.x * block.y));
                    body();
                }
            }
        });
    for (auto &w : ws) w.join();
}
} // namespace cuemu

static double now_ms() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; }
static std::mutex g_alloc_mutex;
static std::map<uintptr_t, std::pair<size_t, int>> g_allocs; // base -> (size, kind 2=device 1=pinned host)

cudaError_t cudaMalloc(void **p, size_t n)
{
    *p = n ? aligned_alloc(256, (n + 255) & ~(size_t)255) : nullptr;
    if (n && !*p) return cudaErrorMemoryAllocation;
    std::lock_guard<std::mutex> lk(g_alloc_mutex);
    if (*p) g_allocs[(uintptr_t)*p] = {n, 2};
    return cudaSuccess;
}
cudaError_t cudaFree(void *p)
{
    if (!p) return cudaSuccess;
    { std::lock_guard<std::mutex> lk(g_alloc_mutex); g_allocs.erase((uintptr_t)p); }
    free(p);
    return cudaSuccess;
}
cudaError_t cudaHostAlloc(void **p, size_t n, unsigned)
{
    *p = n ? aligned_alloc(256, (n + 255) & ~(size_t)255) : nullptr;
    if (n && !*p) return cudaErrorMemoryAllocation;
    std::lock_guard<std::mutex> lk(g_alloc_mutex);
    if (*p) g_allocs[(uintptr_t)*p] = {n, 1};
    return cudaSuccess;
}
cudaError_t cudaFreeHost(void *p) { return cudaFree(p); }
cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { if (n) memmove(d, s, n); return cudaSuccess; }
cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind k, cudaStream_t) { return cudaMemcpy(d, s, n, k); }
cudaError_t cudaMemcpy2DAsync(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h, cudaMemcpyKind, cudaStream_t)
{
    for (size_t r = 0; r < h; r++) memmove((char *)d + r * dp, (const char *)s + r * sp, w);
    return cudaSuccess;
}
cudaError_t cudaMemset(void *d, int v, size_t n) { if (n) memset(d, v, n); return cudaSuccess; }
cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t) { return cudaMemset(d, v, n); }
cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned) { *s = malloc(1); return cudaSuccess; }
cudaError_t cudaStreamDestroy(cudaStream_t s) { free(s); return cudaSuccess; }
cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned) { return cudaSuccess; }
cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
cudaError_t cudaEventCreate(cudaEvent_t *e) { *e = new cuemuEvent{0}; return cudaSuccess; }
cudaError_t cudaEventCreateWithFlags(cudaEvent_t *e, unsigned) { return cudaEventCreate(e); }
cudaError_t cudaEventDestroy(cudaEvent_t e) { delete e; return cudaSuccess; }
cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t) { e->t = now_ms(); return cudaSuccess; }
cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b) { *ms = (float)(b->t - a->t); return cudaSuccess; }
cudaError_t cudaGetLastError() { return cudaSuccess; }
cudaError_t cudaPeekAtLastError() { return cudaSuccess; }
const char *cudaGetErrorString(cudaError_t e) { return e ? "cuemu error" : "no error"; }
cudaError_t cudaGetDeviceCount(int *n) { *n = 1; return cudaSuccess; }
cudaError_t cudaSetDevice(int) { return cudaSuccess; }
cudaError_t cudaGetDevice(int *d) { *d = 0; return cudaSuccess; }
cudaError_t cudaMemGetInfo(size_t *f, size_t *t) { *f = (size_t)8 << 30; *t = (size_t)8 << 30; return cudaSuccess; }
cudaError_t cudaPointerGetAttributes(cudaPointerAttributes *a, const void *p)
{
    std::lock_guard<std::mutex> lk(g_alloc_mutex);
    a->type = cudaMemoryTypeUnregistered; a->device = 0; a->devicePointer = nullptr; a->hostPointer = (void *)p;
    auto it = g_allocs.upper_bound((uintptr_t)p);
    if (it != g_allocs.begin()) {
        --it;
        if ((uintptr_t)p < it->first + it->second.first) {
            a->type = it->second.second == 2 ? cudaMemoryTypeDevice : cudaMemoryTypeHost;
            a->devicePointer = (void *)p;
        }
    }
    return cudaSuccess;
}
cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t *h, void *p) { memset(h, 0, sizeof *h); memcpy(h->reserved, &p, sizeof p); return cudaSuccess; }
cudaError_t cudaIpcOpenMemHandle(void **p, cudaIpcMemHandle_t h, unsigned) { memcpy(p, h.reserved, sizeof *p); return cudaSuccess; }
cudaError_t cudaIpcCloseMemHandle(void *) { return cudaSuccess; }
cudaError_t cudaGetDeviceProperties(cudaDeviceProp *p, int)
{
    memset(p, 0, sizeof *p);
    strcpy(p->name, "cuemu host-thread emulator");
    p->totalGlobalMem = (size_t)8 << 30; p->multiProcessorCount = 4; p->major = 10; p->minor = 0; p->sharedMemPerBlockOptin = 227 * 1024;
    return cudaSuccess;
}
