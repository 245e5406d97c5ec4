// tests/emul/cuda_emul.h -- TEST INFRASTRUCTURE ONLY.
//
// A minimal SIMT emulation so that the product's kernel sources (jpeg_decoder_b200/csrc/*.cuh)
// can be compiled with g++ and run in the CPU-only test tier (-m "not gpu") against the oracle.
// One OS thread per CUDA thread of a CTA, CTAs run one after another; __syncthreads is a
// std::barrier, warp shuffles go through a per-warp scratch line.  It checks LOGIC (indexing,
// scans, the self-synchronising decode, IDCT op order); it says nothing about performance and
// is never linked into libb200jpeg.so -- the product has no CPU path.
#pragma once
#include <atomic>
#include <barrier>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <thread>
#include <vector>
#include <cstdio>

struct uint2 { uint32_t x, y; };
struct uint4 { uint32_t x, y, z, w; };
struct int2 { int x, y; };
struct int4 { int x, y, z, w; };
struct float2 { float x, y; };
struct float4 { float x, y, z, w; };
struct dim3 { unsigned x, y, z; dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {} };
static inline uint2 make_uint2(uint32_t x, uint32_t y) { return {x, y}; }
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return {x, y, z, w}; }
static inline int2 make_int2(int x, int y) { return {x, y}; }
static inline int4 make_int4(int x, int y, int z, int w) { return {x, y, z, w}; }
static inline float4 make_float4(float x, float y, float z, float w) { return {x, y, z, w}; }
typedef void *cudaStream_t;

namespace emu {
struct Warp {
  std::unique_ptr<std::barrier<>> bar;
  uint64_t scratch[32];
  int lanes;
};
struct Cta {
  std::unique_ptr<std::barrier<>> bar;
  std::vector<Warp> warps;
  std::atomic<int> vote[2];
  std::atomic<int> count[2];
  std::vector<unsigned char> dyn_smem;
};
inline Cta *&cta() { static Cta *c = nullptr; return c; }
struct Tls { dim3 tid, bid, bdim, gdim; int vote_phase; };
inline Tls &tls() { static thread_local Tls t; return t; }

template <class F>
void launch(dim3 grid, dim3 block, size_t smem, F body) {
  const unsigned nt = block.x * block.y * block.z;
  if (grid.x == 0 || nt == 0) return;
  Cta c; cta() = &c;
  c.dyn_smem.assign(smem + 128, 0xCD);   // poison: a read of shared memory nobody wrote must not look like a valid zero
  c.vote[0] = c.vote[1] = 0; c.count[0] = c.count[1] = 0;
  std::barrier<> outer(nt);
  auto setup = [&]() {
    c.bar.reset(new std::barrier<>(nt));
    const unsigned nw = (nt + 31) / 32;
    c.warps.resize(nw);
    for (unsigned w = 0; w < nw; ++w) {
      c.warps[w].lanes = (int)((w + 1) * 32 <= nt ? 32 : nt - w * 32);
      c.warps[w].bar.reset(new std::barrier<>(c.warps[w].lanes));
    }
  };
  setup();
  auto worker = [&](unsigned t) {
    Tls &T = tls();
    T.tid = dim3(t % block.x, (t / block.x) % block.y, t / (block.x * block.y));
    T.bdim = block; T.gdim = grid;
    for (unsigned b = 0; b < grid.x * grid.y * grid.z; ++b) {
      T.bid = dim3(b % grid.x, (b / grid.x) % grid.y, b / (grid.x * grid.y));
      T.vote_phase = 0;
      body();
      // a thread that returns early stops taking part in the CTA's barriers (as on the GPU)
      c.bar->arrive_and_drop();
      c.warps[t / 32].bar->arrive_and_drop();
      outer.arrive_and_wait();
      if (t == 0) { setup(); c.vote[0] = c.vote[1] = 0; c.count[0] = c.count[1] = 0; }
      outer.arrive_and_wait();
    }
  };
  std::vector<std::thread> th;
  for (unsigned t = 1; t < nt; ++t) th.emplace_back(worker, t);
  worker(0);
  for (auto &x : th) x.join();
  cta() = nullptr;
}
inline unsigned linear_tid() {
  const Tls &T = tls();
  return T.tid.x + T.bdim.x * (T.tid.y + T.bdim.y * T.tid.z);
}
}  // namespace emu

#define threadIdx (emu::tls().tid)
#define blockIdx (emu::tls().bid)
#define blockDim (emu::tls().bdim)
#define gridDim (emu::tls().gdim)

#define B2_DEVICE static inline
#define B2_HD static inline
#define B2_MEMBER inline
#define B2_COLD static
#define B2_GLOBAL static
#define B2_SHARED static
#define B2_LAUNCH(kernel, grid, block, smem, stream, ...) \
  emu::launch(dim3(grid), dim3(block), (smem), [&]() { kernel(__VA_ARGS__); })
#define B2_DYN_SMEM(type, name) \
  type *name = reinterpret_cast<type *>((reinterpret_cast<uintptr_t>(emu::cta()->dyn_smem.data()) + 127) & ~uintptr_t(127))
#define B2_CONSTANT static const
#define __restrict__
#define __launch_bounds__(...)

static inline void __syncthreads() { emu::cta()->bar->arrive_and_wait(); }
static inline void __syncwarp(unsigned = 0xffffffffu) { emu::cta()->warps[emu::linear_tid() / 32].bar->arrive_and_wait(); }
static inline int __syncthreads_or(int p) {
  emu::Cta *c = emu::cta(); int &ph = emu::tls().vote_phase;
  if (p) c->vote[ph].store(1);
  c->bar->arrive_and_wait();
  const int r = c->vote[ph].load();
  c->vote[ph ^ 1].store(0);  // the other phase is idle between two votes
  c->bar->arrive_and_wait();
  ph ^= 1;
  return r;
}
static inline int __syncthreads_count(int p) {
  emu::Cta *c = emu::cta(); int &ph = emu::tls().vote_phase;
  if (p) c->count[ph].fetch_add(1);
  c->bar->arrive_and_wait();
  const int r = c->count[ph].load();
  c->count[ph ^ 1].store(0);
  c->bar->arrive_and_wait();
  ph ^= 1;
  return r;
}

namespace emu {
template <class T>
inline T warp_exchange(T v, int src_lane, bool valid_src) {
  static_assert(sizeof(T) <= 8, "");
  Warp &w = cta()->warps[linear_tid() / 32];
  const int lane = linear_tid() % 32;
  uint64_t raw = 0; memcpy(&raw, &v, sizeof(T));
  w.scratch[lane] = raw;
  w.bar->arrive_and_wait();
  T out = v;
  if (valid_src && src_lane >= 0 && src_lane < w.lanes) { uint64_t r = w.scratch[src_lane]; memcpy(&out, &r, sizeof(T)); }
  w.bar->arrive_and_wait();
  return out;
}
}  // namespace emu
template <class T> static inline T __shfl_sync(unsigned, T v, int src, int width = 32) {
  const int lane = emu::linear_tid() % 32; const int base = lane & ~(width - 1);
  return emu::warp_exchange(v, base + (src & (width - 1)), true);
}
template <class T> static inline T __shfl_up_sync(unsigned, T v, unsigned d, int width = 32) {
  const int lane = emu::linear_tid() % 32; const int base = lane & ~(width - 1);
  return emu::warp_exchange(v, lane - (int)d, lane - (int)d >= base);
}
template <class T> static inline T __shfl_down_sync(unsigned, T v, unsigned d, int width = 32) {
  const int lane = emu::linear_tid() % 32; const int base = lane & ~(width - 1);
  return emu::warp_exchange(v, lane + (int)d, lane + (int)d < base + width);
}
template <class T> static inline T __shfl_xor_sync(unsigned, T v, int m, int width = 32) {
  const int lane = emu::linear_tid() % 32;
  (void)width;
  return emu::warp_exchange(v, lane ^ m, true);
}
static inline unsigned __ballot_sync(unsigned, int p) {
  emu::Warp &w = emu::cta()->warps[emu::linear_tid() / 32];
  const int lane = emu::linear_tid() % 32;
  w.scratch[lane] = p ? 1 : 0;
  w.bar->arrive_and_wait();
  unsigned m = 0;
  for (int i = 0; i < w.lanes; ++i) m |= (unsigned)w.scratch[i] << i;
  w.bar->arrive_and_wait();
  return m;
}
static inline void __threadfence() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
static inline int __any_sync(unsigned m, int p) { return __ballot_sync(m, p) != 0; }
static inline int __all_sync(unsigned m, int p) {
  const unsigned b = __ballot_sync(m, p);
  const int lanes = emu::cta()->warps[emu::linear_tid() / 32].lanes;
  return b == (lanes == 32 ? 0xffffffffu : ((1u << lanes) - 1));
}

// ---- atomics (CTAs are sequential, threads of one CTA are concurrent) ----
template <class T> static inline T atomicAdd(T *p, T v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
template <class T> static inline T atomicOr(T *p, T v) { return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }
template <class T> static inline T atomicExch(T *p, T v) { return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
template <class T> static inline T atomicMin(T *p, T v) {
  T o = __atomic_load_n(p, __ATOMIC_SEQ_CST);
  while (v < o && !__atomic_compare_exchange_n(p, &o, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
  return o;
}
template <class T> static inline T atomicMax(T *p, T v) {
  T o = __atomic_load_n(p, __ATOMIC_SEQ_CST);
  while (v > o && !__atomic_compare_exchange_n(p, &o, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
  return o;
}

// ---- intrinsics ----
template <class T> static inline T __ldg(const T *p) { return *p; }
static inline uint32_t __funnelshift_l(uint32_t lo, uint32_t hi, uint32_t s) {
  s &= 31; return s ? (hi << s) | (lo >> (32 - s)) : hi;
}
static inline uint32_t __funnelshift_r(uint32_t lo, uint32_t hi, uint32_t s) {
  s &= 31; return s ? (lo >> s) | (hi << (32 - s)) : lo;
}
static inline uint32_t __byte_perm(uint32_t a, uint32_t b, uint32_t s) {
  const uint64_t v = ((uint64_t)b << 32) | a; uint32_t r = 0;
  for (int i = 0; i < 4; ++i) r |= (uint32_t)((v >> (8 * ((s >> (4 * i)) & 7))) & 0xff) << (8 * i);
  return r;
}
static inline int __clz(int x) { return x == 0 ? 32 : __builtin_clz((unsigned)x); }
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline int __float2int_rz(float x) { return (int)x; }
static inline float __int2float_rn(int x) { return (float)x; }
// volatile keeps g++ from contracting a*b+c into an FMA even if someone adds -march flags
static inline float __fmul_rn(float a, float b) { volatile float r = a * b; return r; }
static inline float __fadd_rn(float a, float b) { volatile float r = a + b; return r; }
static inline float __fsub_rn(float a, float b) { volatile float r = a - b; return r; }
template <class T> static inline T min(T a, T b) { return a < b ? a : b; }
template <class T> static inline T max(T a, T b) { return a > b ? a : b; }
static inline uint32_t b2_ld_nc_u32(const uint32_t *p) { return *p; }
static inline uint4 b2_ld_nc_v4(const uint4 *p) { return *p; }
static inline uint32_t emu_f2u8_rz_sat(float x) {
  if (!(x == x)) return 0;
  if (x <= 0.0f) return 0;
  if (x >= 255.0f) return 255;
  return (uint32_t)(int)x;
}
static inline uint32_t b2_f4_to_u8x4(float a, float b, float c, float d) {
  return emu_f2u8_rz_sat(a) | (emu_f2u8_rz_sat(b) << 8) | (emu_f2u8_rz_sat(c) << 16) | (emu_f2u8_rz_sat(d) << 24);
}
// mbarrier / bulk copy: the barrier word counts completed phases; the copy is done by the issuing thread
static inline void b2_mbar_init(uint64_t *bar, uint32_t) { __atomic_store_n(bar, (uint64_t)0, __ATOMIC_SEQ_CST); }
static inline void b2_mbar_init_fence() {}
static inline void b2_fence_proxy_async() {}
static inline void b2_bulk_load(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
  memcpy(dst, src, bytes);
  __atomic_fetch_add(bar, (uint64_t)1, __ATOMIC_SEQ_CST);
}
static inline void b2_cp_async4(void *dst, const void *src) { memcpy(dst, src, 4); }
static inline void b2_cp_async_commit() {}
template <int N> static inline void b2_cp_async_wait() {}
static inline void b2_bulk_store(void *dst, const void *src, uint32_t bytes) { memcpy(dst, src, bytes); }
struct B2TensorMap { int unused; };                     // the emulation always takes the row-by-row store
#define B2_GRID_CONSTANT
static inline void b2_tensor_store_3d(const void *, const void *, int, int, int) {}
static inline void b2_bulk_commit() {}
static inline void b2_bulk_wait_read() {}
static inline void b2_mbar_wait(uint64_t *bar, uint32_t parity) {
  while ((__atomic_load_n(bar, __ATOMIC_SEQ_CST) & 1u) == parity) std::this_thread::yield();
}
struct b2_f2 { float lo, hi; };
static inline b2_f2 b2_f2_pack(float lo, float hi) { return {lo, hi}; }
static inline void b2_f2_unpack(b2_f2 v, float &lo, float &hi) { lo = v.lo; hi = v.hi; }
static inline b2_f2 b2_f2_add(b2_f2 a, b2_f2 b) { return {__fadd_rn(a.lo, b.lo), __fadd_rn(a.hi, b.hi)}; }
static inline b2_f2 b2_f2_sub(b2_f2 a, b2_f2 b) { return {__fsub_rn(a.lo, b.lo), __fsub_rn(a.hi, b.hi)}; }
static inline void b2_backoff(unsigned) { std::this_thread::yield(); }
// 32-bit window addresses of the CTA's dynamic shared memory
static inline unsigned char *b2_dsm_base() {
  return reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(emu::cta()->dyn_smem.data()) + 127) & ~uintptr_t(127));
}
static inline uint32_t b2_dsm_addr(const void *p) { return (uint32_t)(reinterpret_cast<const unsigned char *>(p) - b2_dsm_base()); }
// every access through a window address is checked: inside the CTA's dynamic shared memory and naturally aligned (the GPU
// faults with "misaligned address" / reads another CTA's memory where this emulation would silently go on)
static inline unsigned char *b2_dsm_at(uint32_t a, uint32_t bytes, const char *what) {
  const size_t cap = emu::cta()->dyn_smem.size() - 128;
  if ((a & (bytes - 1)) != 0 || (size_t)a + bytes > cap) {
    fprintf(stderr, "emulation: %s of %u bytes at shared window address %u (cta %u thread %u): %s\n", what, bytes, a,
            blockIdx.x, emu::linear_tid(), (a & (bytes - 1)) ? "misaligned" : "outside the CTA's shared memory");
    abort();
  }
  return b2_dsm_base() + a;
}
static inline uint32_t b2_lds(uint32_t a) { uint32_t v; memcpy(&v, b2_dsm_at(a, 4, "ld.shared.u32"), 4); return v; }
static inline void b2_sts(uint32_t a, uint32_t v) { memcpy(b2_dsm_at(a, 4, "st.shared.u32"), &v, 4); }
static inline uint32_t b2_lds_ro(uint32_t a) { return b2_lds(a); }
static inline uint32_t b2_lds_ro_if(bool on, uint32_t a, uint32_t keep) { return on ? b2_lds(a) : keep; }
static inline void b2_sts_u16(uint32_t a, uint32_t v) { const uint16_t h = (uint16_t)v; memcpy(b2_dsm_at(a, 2, "st.shared.u16"), &h, 2); }
static inline uint32_t b2_lds_u16(uint32_t a) { uint16_t h; memcpy(&h, b2_dsm_at(a, 2, "ld.shared.u16"), 2); return h; }
static inline void b2_sts_u8(uint32_t a, uint32_t v) { *b2_dsm_at(a, 1, "st.shared.u8") = (unsigned char)v; }
static inline uint32_t b2_lds_u8(uint32_t a) { return *b2_dsm_at(a, 1, "ld.shared.u8"); }
static inline uint4 b2_lds_v4(uint32_t a) { uint4 v; memcpy(&v, b2_dsm_at(a, 16, "ld.shared.v4"), 16); return v; }
static inline void b2_sts_v4(uint32_t a, uint4 v) { memcpy(b2_dsm_at(a, 16, "st.shared.v4"), &v, 16); }
static inline void b2_stg_v4(void *p, uint4 v) { memcpy(p, &v, 16); }
template <class T> static inline void b2_keep(T *&) {}
static inline uint32_t b2_launder(uint32_t v) { return v; }
template <uint32_t MASK> static inline uint32_t b2_and_x4_add(uint32_t v, uint32_t base) { return base + 4u * (v & MASK); }
template <int SHIFT> static inline uint32_t b2_shr_x4_add(uint32_t v, uint32_t base) { return base + 4u * (v >> SHIFT); }
// mbarriers by window address (arrival counts only): [63:32] completed phases, [31:16] arrivals a phase takes, [15:0] arrivals so far
static inline uint64_t *b2_mbar_ptr(uint32_t a) { return reinterpret_cast<uint64_t *>(b2_dsm_at(a, 8, "mbarrier")); }
static inline void b2_mbar_init_a(uint32_t a, uint32_t count) { __atomic_store_n(b2_mbar_ptr(a), (uint64_t)count << 16, __ATOMIC_SEQ_CST); }
static inline void b2_mbar_arrive_a(uint32_t a) {
  uint64_t *p = b2_mbar_ptr(a);
  uint64_t o = __atomic_load_n(p, __ATOMIC_SEQ_CST), n;
  do {
    const uint64_t need = (o >> 16) & 0xffff, got = (o & 0xffff) + 1;
    n = got == need ? (((o >> 32) + 1) << 32) | (need << 16) : (o & ~(uint64_t)0xffff) | got;
  } while (!__atomic_compare_exchange_n(p, &o, n, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST));
}
static inline bool b2_mbar_test_a(uint32_t a, uint32_t parity) {
  return ((__atomic_load_n(b2_mbar_ptr(a), __ATOMIC_SEQ_CST) >> 32) & 1u) != parity;
}
static inline void b2_mbar_wait_a(uint32_t a, uint32_t parity) {
  while (((__atomic_load_n(b2_mbar_ptr(a), __ATOMIC_SEQ_CST) >> 32) & 1u) == parity) std::this_thread::yield();
}
static inline uint32_t b2_shr(uint32_t v, uint32_t n) { return n >= 32 ? 0u : v >> n; }
static inline void b2_cp_async_src_check(const void *src) {
  if (reinterpret_cast<uintptr_t>(src) & 3u) { fprintf(stderr, "emulation: cp.async from a misaligned global address\n"); abort(); }
}
static inline void b2_cp_async4x4_if(bool on, uint32_t dst_smem, const void *src) {
  if (on) { b2_cp_async_src_check(src); for (int i = 0; i < 4; ++i) memcpy(b2_dsm_at(dst_smem + 128 * i, 4, "cp.async"), static_cast<const unsigned char *>(src) + 4 * i, 4); }
}
static inline void b2_cp_async4_if(bool on, uint32_t dst_smem, const void *src) { if (on) { b2_cp_async_src_check(src); memcpy(b2_dsm_at(dst_smem, 4, "cp.async"), src, 4); } }
static inline uint32_t b2_ld_acquire_smem(const uint32_t *p) {
  std::this_thread::yield();                    // callers spin on it: one OS thread per CUDA thread here
  return __atomic_load_n(p, __ATOMIC_ACQUIRE);
}
static inline void b2_st_release_smem(uint32_t *p, uint32_t v) { __atomic_store_n(p, v, __ATOMIC_RELEASE); }
