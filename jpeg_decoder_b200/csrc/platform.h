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
#include <cuda.h>
#include <cuda_runtime.h>
struct B2TensorMap { CUtensorMap map; };                 // kernel parameter (__grid_constant__)
#define B2_GRID_CONSTANT __grid_constant__
#define B2_DEVICE __device__ __forceinline__
#define B2_HD __host__ __device__ __forceinline__
#define B2_MEMBER __device__ __forceinline__
#define B2_COLD __device__ __noinline__            /* rare paths stay out of the hot loops' instruction footprint */
#define B2_GLOBAL __global__
#define B2_SHARED __shared__
#ifdef B200JPEG_TUNING   /* measurement / debugging builds only (scripts/build_variant.sh): B200JPEG_SYNC_EACH=1 waits after every
                            launch and names the kernel a fault belongs to (compute-sanitizer is closed on this pool) */
#include <stdio.h>
#include <stdlib.h>
#define B2_LAUNCH(kernel, grid, block, smem, stream, ...)                                                      \
  do {                                                                                                         \
    kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);                                                \
    if (getenv("B200JPEG_SYNC_EACH")) {                                                                        \
      const cudaError_t e_ = cudaStreamSynchronize(stream);                                                    \
      if (e_ != cudaSuccess) fprintf(stderr, "B2_LAUNCH %s grid %d: %s\n", #kernel, (int)(grid), cudaGetErrorString(e_)); \
    }                                                                                                          \
  } while (0)
#else
#define B2_LAUNCH(kernel, grid, block, smem, stream, ...) \
  kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#endif
#define B2_DYN_SMEM(type, name) extern __shared__ __align__(128) unsigned char name##_raw_[]; \
  type *name = reinterpret_cast<type *>(name##_raw_)
#define B2_CONSTANT __constant__

B2_DEVICE uint32_t b2_ld_nc_u32(const uint32_t *p) { return __ldg(p); }
B2_DEVICE uint4 b2_ld_nc_v4(const uint4 *p) { return __ldg(p); }
// four floats -> truncate toward zero, clamp to [0,255], pack a | b<<8 | c<<16 | d<<24.
// cvt.rzi.s32.f32 + cvt.pack.sat.u8.s32 fuse into two F2IP.U8.F32.TRUNC (two floats each, merged): no PRMT.
B2_DEVICE uint32_t b2_f4_to_u8x4(float a, float b, float c, float d) {
  const int ia = __float2int_rz(a), ib = __float2int_rz(b), ic = __float2int_rz(c), id = __float2int_rz(d);
  uint32_t hi, r;
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, 0;" : "=r"(hi) : "r"(id), "r"(ic));
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(ib), "r"(ia), "r"(hi));
  return r;
}

// ---- two fp32 additions in one instruction (add.rn.f32x2, SASS FADD2; sm_100): each half is an IEEE round-to-nearest
//      add, bit for bit what two FADDs give.  Multiplications stay scalar on purpose: ptxas contracts mul.f32x2 + add.f32x2
//      into FFMA2 even under --fmad=false and explicit .rn, which would change the rounding. ----
typedef unsigned long long b2_f2;
B2_DEVICE b2_f2 b2_f2_pack(float lo, float hi) { b2_f2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
B2_DEVICE void b2_f2_unpack(b2_f2 v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
B2_DEVICE b2_f2 b2_f2_add(b2_f2 a, b2_f2 b) { b2_f2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
B2_DEVICE b2_f2 b2_f2_sub(b2_f2 a, b2_f2 b) { b2_f2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

// ---- cp.async (LDGSTS): global -> shared copies that bypass the register file ----
B2_DEVICE void b2_cp_async4(void *dst, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(__cvta_generic_to_global(src)) : "memory");
}
// predicated 4-byte copy to a shared WINDOW address: no branch around it
B2_DEVICE void b2_cp_async4_if(bool on, uint32_t dst_smem, const void *src) {
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %0, 0;\n@p cp.async.ca.shared.global [%1], [%2], 4;\n}" ::"r"((int)on), "r"(dst_smem),
               "l"(__cvta_generic_to_global(src)) : "memory");
}
// four of them: consecutive words of global memory to shared addresses 128 bytes apart ([word][lane] layouts), one
// address computation for all four
B2_DEVICE void b2_cp_async4x4_if(bool on, uint32_t dst_smem, const void *src) {
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %0, 0;\n"
               "@p cp.async.ca.shared.global [%1], [%2], 4;\n"
               "@p cp.async.ca.shared.global [%1+128], [%2+4], 4;\n"
               "@p cp.async.ca.shared.global [%1+256], [%2+8], 4;\n"
               "@p cp.async.ca.shared.global [%1+384], [%2+12], 4;\n}" ::"r"((int)on), "r"(dst_smem),
               "l"(__cvta_generic_to_global(src)) : "memory");
}
B2_DEVICE void b2_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> B2_DEVICE void b2_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ---- TMA 1-D bulk copy (cp.async.bulk, SASS UBLKCP) + mbarrier, the sm_90+/sm_100 async-copy path ----
B2_DEVICE uint32_t b2_smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
B2_DEVICE void b2_mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(b2_smem_addr(bar)), "r"(count) : "memory");
}
B2_DEVICE void b2_mbar_init_fence() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// orders this thread's earlier generic-proxy accesses of shared memory before later async-proxy (TMA) ones
B2_DEVICE void b2_fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// one thread: arm the barrier with `bytes` and start the global -> shared bulk copy that completes it
B2_DEVICE void b2_bulk_load(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b2_smem_addr(bar)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(b2_smem_addr(dst)), "l"(__cvta_generic_to_global(src)), "r"(bytes), "r"(b2_smem_addr(bar))
               : "memory");
}
// shared -> global bulk copy (TMA store, bulk-group completion); every issuing thread commits and waits itself
B2_DEVICE void b2_bulk_store(void *dst, const void *src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               ::"l"(__cvta_generic_to_global(dst)), "r"(b2_smem_addr(src)), "r"(bytes) : "memory");
}
// shared -> global TILE store through a 3-D tensor map (cp.async.bulk.tensor, SASS UTMASTG): one instruction
// moves a whole box (rows x bytes) and clips it at the tensor's bounds; issued by ONE lane
B2_DEVICE void b2_tensor_store_3d(const void *tmap, const void *src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];"
               ::"l"(tmap), "r"(c0), "r"(c1), "r"(c2), "r"(b2_smem_addr(src)) : "memory");
}
B2_DEVICE void b2_bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
B2_DEVICE void b2_bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
B2_DEVICE void b2_mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "B2_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 0x989680;\n"
      "@p bra B2_DONE;\n"
      "bra B2_WAIT;\n"
      "B2_DONE:\n"
      "}" ::"r"(b2_smem_addr(bar)), "r"(parity) : "memory");
}
// ---- producer/consumer words in shared memory between two warps of one CTA ----
B2_DEVICE uint32_t b2_ld_acquire_smem(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.cta.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(b2_smem_addr(p)) : "memory");
  return v;
}
B2_DEVICE void b2_st_release_smem(uint32_t *p, uint32_t v) {
  asm volatile("st.release.cta.shared::cta.u32 [%0], %1;" ::"r"(b2_smem_addr(p)), "r"(v) : "memory");
}
// a waiting warp must not hammer the LSU of the sub-partition it shares with the warp it waits for
B2_DEVICE void b2_backoff(unsigned ns) { __nanosleep(ns); }
// ---- dynamic shared memory through 32-bit window addresses (register + immediate addressing, no
//      generic-pointer arithmetic in inner loops) ----
B2_DEVICE uint32_t b2_dsm_addr(const void *p) { return b2_smem_addr(p); }
B2_DEVICE uint32_t b2_lds(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
B2_DEVICE void b2_sts(uint32_t a, uint32_t v) { asm volatile("st.shared::cta.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
// load of data that does not change while the caller's loop runs (lookup tables), or whose producer is ordered by
// other means (cp.async groups): NOT volatile and no memory clobber, so the compiler may schedule it freely --
// hoist the next step's loads over this step's stores.  The address must come out of b2_launder() placed after
// whatever made the data visible, or the load could be moved in front of that.
B2_DEVICE uint32_t b2_lds_ro(uint32_t a) {
  uint32_t v;
  asm("ld.shared::cta.u32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
// the same under a predicate: `keep` where the predicate is off (no branch, no select behind the load)
B2_DEVICE uint32_t b2_lds_ro_if(bool on, uint32_t a, uint32_t keep) {
  asm("{\n.reg .pred q;\nsetp.ne.u32 q, %2, 0;\n@q ld.shared::cta.u32 %0, [%1];\n}" : "+r"(keep) : "r"(a), "r"((uint32_t)on));
  return keep;
}
B2_DEVICE void b2_sts_u16(uint32_t a, uint32_t v) { asm volatile("st.shared::cta.u16 [%0], %1;" ::"r"(a), "h"((unsigned short)v) : "memory"); }
B2_DEVICE uint32_t b2_lds_u16(uint32_t a) { unsigned short v; asm volatile("ld.shared::cta.u16 %0, [%1];" : "=h"(v) : "r"(a) : "memory"); return v; }
B2_DEVICE void b2_sts_u8(uint32_t a, uint32_t v) { asm volatile("st.shared::cta.u8 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
B2_DEVICE uint32_t b2_lds_u8(uint32_t a) { uint32_t v; asm volatile("ld.shared::cta.u8 %0, [%1];" : "=r"(v) : "r"(a) : "memory"); return v; }
B2_DEVICE uint4 b2_lds_v4(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared::cta.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
  return v;
}
B2_DEVICE void b2_sts_v4(uint32_t a, uint4 v) {
  asm volatile("st.shared::cta.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// ---- mbarriers by 32-bit window address (the segment decoder's hand-overs) ----
B2_DEVICE void b2_mbar_arrive_a(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory"); }
B2_DEVICE void b2_mbar_init_a(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
// has the phase of that parity completed?  Does not wait: issued early, the answer is there when it is needed
B2_DEVICE bool b2_mbar_test_a(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n.reg .pred p;\nmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
               : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0u;
}
B2_DEVICE void b2_mbar_wait_a(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "B2_WAITA:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 0x989680;\n"   /* suspend-time hint: without it the wait comes back at once and the retry loop eats issue slots */
      "@p bra B2_DONEA;\n"
      "bra B2_WAITA;\n"
      "B2_DONEA:\n"
      "}" ::"r"(bar), "r"(parity) : "memory");
}
// base + 4 * ((v & MASK) >> SHIFT) as two instructions (AND or shift, then multiply-add): written in C the compiler turns
// it back into shift, AND and add -- one more link in a dependent chain
template <uint32_t MASK> B2_DEVICE uint32_t b2_and_x4_add(uint32_t v, uint32_t base) {
  uint32_t r;
  asm("{\n.reg .b32 t;\nand.b32 t, %1, %2;\nmad.lo.u32 %0, t, 4, %3;\n}" : "=r"(r) : "r"(v), "n"(MASK), "r"(base));
  return r;
}
template <int SHIFT> B2_DEVICE uint32_t b2_shr_x4_add(uint32_t v, uint32_t base) {
  uint32_t r;
  asm("{\n.reg .b32 t;\nshr.u32 t, %1, %2;\nmad.lo.u32 %0, t, 4, %3;\n}" : "=r"(r) : "r"(v), "n"(SHIFT), "r"(base));
  return r;
}
B2_DEVICE void b2_stg_v4(void *p, uint4 v) {          // a store the compiler cannot take for a generic one
  asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
template <class T> B2_DEVICE void b2_keep(T *&p) { asm volatile("" : "+l"(p)); }
B2_DEVICE uint32_t b2_launder(uint32_t v) { asm volatile("" : "+r"(v) :: "memory"); return v; }
// shifts that give 0 for counts >= 32 (PTX semantics; undefined in C++)
B2_DEVICE uint32_t b2_shr(uint32_t v, uint32_t n) { uint32_t r; asm("shr.u32 %0, %1, %2;" : "=r"(r) : "r"(v), "r"(n)); return r; }
#else
// ------------------------------------------------------------------ emulation
#include "../../tests/emul/cuda_emul.h"
#endif
