// Build-mode switch for the kernel sources.
//   default          : real CUDA (nvcc, sm_100a).  This is the only mode the product is ever built in.
//   -DGLP_EMU        : tests/cuemu/cuemu.h, a host-thread stand-in for the handful of CUDA constructs the
//                      kernels use, so the SAME kernel sources can be stepped through under gdb/asan on a
//                      box without a GPU.  It builds tests/cuemu/libglprover_emu.so, which only
//                      tests/test_emu_*.py load; the package never does (no CPU fallback).
#pragma once
#ifdef GLP_EMU
#include "cuemu.h"
#else
#include <cuda_runtime.h>
// kernels that use __syncthreads()/__shared__
#define GLP_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
// kernels with fully independent threads
#define GLP_LAUNCH_FLAT(kernel, grid, block, stream, ...) kernel<<<(grid), (block), 0, (stream)>>>(__VA_ARGS__)
#define GLP_DYN_SMEM(type, name) extern __shared__ __align__(16) unsigned char name##_raw[]; type *name = reinterpret_cast<type *>(name##_raw)
#define GLP_CONSTANT __constant__
#define GLP_NOINLINE __noinline__
#define GLP_SET_SMEM(kernel, bytes) CU(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)))
#endif
