// TEST/DEBUG HARNESS ONLY -- not part of the product and never a fallback.
//
// A host-thread stand-in for the few CUDA constructs the kernel sources use (threadIdx/blockIdx,
// __syncthreads, __shared__, <<<>>> through GLP_LAUNCH, a dozen cudaXxx runtime calls), so the very same
// .cu/.cuh files can be compiled with g++ -DGLP_EMU into tests/cuemu/libglprover_emu.so and their index
// arithmetic debugged on a machine without a GPU.  Blocks run one after another; a block's threads are
// real OS threads meeting at a barrier, so __syncthreads and static __shared__ behave as on the device.
// Inline PTX is compiled out (GL_DEVICE_PTX == 0): the PTX paths are only ever checked on the B200.
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stddef.h>
#include <functional>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline __attribute__((always_inline))
#define __shared__ static
#define __restrict__
#define __launch_bounds__(...)
#define __align__(n) __attribute__((aligned(n)))
#define GLP_CONSTANT static const
#define GLP_SET_SMEM(kernel, bytes) ((void)0)
#define GLP_NOINLINE __attribute__((noinline))

struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
namespace cuemu {
extern thread_local dim3 t_threadIdx;
extern dim3 g_blockIdx, g_blockDim, g_gridDim;
extern thread_local dim3 t_blockIdx;      // flat launches run blocks on several workers
extern unsigned char *g_dynsmem;
void barrier();
void launch(dim3 grid, dim3 block, size_t smem, const std::function<void()> &body);
void launch_flat(dim3 grid, dim3 block, const std::function<void()> &body);
}
#define threadIdx (cuemu::t_threadIdx)
#define blockIdx (cuemu::t_blockIdx)
#define blockDim (cuemu::g_blockDim)
#define gridDim (cuemu::g_gridDim)
static inline void __syncthreads() { cuemu::barrier(); }
static inline unsigned __brev(unsigned x)
{
    x = (x >> 16) | (x << 16);
    x = ((x & 0xFF00FF00u) >> 8) | ((x & 0x00FF00FFu) << 8);
    x = ((x & 0xF0F0F0F0u) >> 4) | ((x & 0x0F0F0F0Fu) << 4);
    x = ((x & 0xCCCCCCCCu) >> 2) | ((x & 0x33333333u) << 2);
    return ((x & 0xAAAAAAAAu) >> 1) | ((x & 0x55555555u) << 1);
}
static inline unsigned long long __umul64hi(unsigned long long a, unsigned long long b) { return (unsigned long long)(((unsigned __int128)a * b) >> 64); }
template <class T> static inline T __ldg(const T *p) { return *p; }

#define GLP_LAUNCH(kernel, grid, block, smem, stream, ...) cuemu::launch(dim3(grid), dim3(block), (smem), [=]() { kernel(__VA_ARGS__); })
#define GLP_LAUNCH_FLAT(kernel, grid, block, stream, ...) cuemu::launch_flat(dim3(grid), dim3(block), [=]() { kernel(__VA_ARGS__); })
#define GLP_DYN_SMEM(type, name) type *name = reinterpret_cast<type *>(cuemu::g_dynsmem)

// ---- runtime API subset (everything is host memory) ----
typedef int cudaError_t;
typedef void *cudaStream_t;
typedef struct cuemuEvent { double t; } *cudaEvent_t;
enum { cudaSuccess = 0, cudaErrorInvalidValue = 1, cudaErrorMemoryAllocation = 2 };
enum cudaMemcpyKind { cudaMemcpyHostToHost = 0, cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3, cudaMemcpyDefault = 4 };
enum cudaMemoryType { cudaMemoryTypeUnregistered = 0, cudaMemoryTypeHost = 1, cudaMemoryTypeDevice = 2, cudaMemoryTypeManaged = 3 };
struct cudaPointerAttributes { cudaMemoryType type; int device; void *devicePointer; void *hostPointer; };
enum { cudaStreamNonBlocking = 1, cudaHostAllocDefault = 0, cudaEventDisableTiming = 2, cudaEventDefault = 0 };
enum cudaFuncAttribute { cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
struct cudaDeviceProp { char name[256]; size_t totalGlobalMem; int multiProcessorCount; int major, minor; size_t sharedMemPerBlockOptin; };

cudaError_t cudaMalloc(void **p, size_t n);
cudaError_t cudaFree(void *p);
cudaError_t cudaHostAlloc(void **p, size_t n, unsigned flags);
cudaError_t cudaFreeHost(void *p);
enum { cudaHostRegisterPortable = 1 };
static inline cudaError_t cudaHostRegister(void *, size_t, unsigned) { return cudaSuccess; }
static inline cudaError_t cudaHostUnregister(void *) { return cudaSuccess; }
cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind k);
cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind k, cudaStream_t st);
cudaError_t cudaMemcpy2DAsync(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h, cudaMemcpyKind k, cudaStream_t st);
cudaError_t cudaMemset(void *d, int v, size_t n);
cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t st);
cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned flags);
static inline cudaError_t cudaStreamCreateWithPriority(cudaStream_t *s, unsigned flags, int) { return cudaStreamCreateWithFlags(s, flags); }
static inline cudaError_t cudaDeviceGetStreamPriorityRange(int *lo, int *hi) { *lo = 0; *hi = -1; return cudaSuccess; }
cudaError_t cudaStreamDestroy(cudaStream_t s);
cudaError_t cudaStreamSynchronize(cudaStream_t s);
cudaError_t cudaStreamWaitEvent(cudaStream_t s, cudaEvent_t e, unsigned flags);
cudaError_t cudaDeviceSynchronize();
cudaError_t cudaEventCreate(cudaEvent_t *e);
cudaError_t cudaEventCreateWithFlags(cudaEvent_t *e, unsigned flags);
cudaError_t cudaEventDestroy(cudaEvent_t e);
cudaError_t cudaEventRecord(cudaEvent_t e, cudaStream_t s);
cudaError_t cudaEventSynchronize(cudaEvent_t e);
cudaError_t cudaEventElapsedTime(float *ms, cudaEvent_t a, cudaEvent_t b);
cudaError_t cudaGetLastError();
cudaError_t cudaPeekAtLastError();
const char *cudaGetErrorString(cudaError_t e);
cudaError_t cudaGetDeviceCount(int *n);
cudaError_t cudaSetDevice(int d);
cudaError_t cudaGetDevice(int *d);
cudaError_t cudaMemGetInfo(size_t *free_, size_t *total);
cudaError_t cudaPointerGetAttributes(cudaPointerAttributes *a, const void *p);
// a "device symbol" is a plain host variable here
template <class T> static inline cudaError_t cudaGetSymbolAddress(void **p, T &sym) { *p = (void *)&sym; return cudaSuccess; }
// IPC inside one process: the "handle" is the pointer itself (enough to exercise export/import/scatter on the CPU)
struct cudaIpcMemHandle_t { char reserved[64]; };
enum { cudaIpcMemLazyEnablePeerAccess = 1 };
cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t *h, void *p);
cudaError_t cudaIpcOpenMemHandle(void **p, cudaIpcMemHandle_t h, unsigned flags);
cudaError_t cudaIpcCloseMemHandle(void *p);
cudaError_t cudaGetDeviceProperties(cudaDeviceProp *p, int dev);
// one emulated device: peer access is only ever asked for between distinct devices
enum { cudaErrorPeerAccessAlreadyEnabled = 704 };
static inline cudaError_t cudaDeviceCanAccessPeer(int *can, int, int) { *can = 1; return cudaSuccess; }
static inline cudaError_t cudaDeviceEnablePeerAccess(int, unsigned) { return cudaSuccess; }
template <class F> static inline cudaError_t cudaFuncSetAttribute(F, cudaFuncAttribute, int) { return cudaSuccess; }
