// platform.h -- the one place where the kernel sources meet their compiler.
//
// Product build (nvcc, sm_100a): plain CUDA.
// Test build (g++ -DB200JPEG_EMU, tests/emul/ only): the same kernel sources run on a small
// SIMT emulation (threads = std::thread, __syncthreads = std::barrier) so that kernel LOGIC can
// be checked against the oracle in the CPU-only test tier.  The emulation is never compiled into
// libb200jpeg.so; the product has no CPU path.
#pragma once
#include <stdint.h>
#include <stddef.h>

#ifndef B200JPEG_EMU
// ------------------------------------------------------------------ CUDA
#include <cuda_runtime.h>
#define B2_DEVICE __device__ __forceinline__
#define B2_HD __host__ __device__ __forceinline__
#define B2_MEMBER __device__ __forceinline__
#define B2_GLOBAL __global__
#define B2_SHARED __shared__
#define B2_LAUNCH(kernel, grid, block, smem, stream, ...) \
  kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#define B2_DYN_SMEM(type, name) extern __shared__ __align__(16) unsigned char name##_raw_[]; \
  type *name = reinterpret_cast<type *>(name##_raw_)

B2_DEVICE uint32_t b2_ld_nc_u32(const uint32_t *p) { return __ldg(p); }
B2_DEVICE uint4 b2_ld_nc_v4(const uint4 *p) { return __ldg(p); }
// float -> uint8 with truncation toward zero and saturation to [0,255] in one F2I
B2_DEVICE uint32_t b2_f2u8_rz_sat(float x) {
  uint32_t r;
  asm("cvt.rzi.u8.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
B2_DEVICE void b2_st_cs_v4(uint4 *p, uint4 v) { __stcs(p, v); }
#else
// ------------------------------------------------------------------ emulation
#include "../../tests/emul/cuda_emul.h"
#endif
