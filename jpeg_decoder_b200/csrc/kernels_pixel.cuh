// kernels_pixel.cuh -- stage P: fused dequantise + 8x8 IDCT + chroma upsample + YCbCr->RGB.
// Replaces Image::dequantize (jpeg.cpp:563-590), inverseDCT (jpeg.cpp:592-753) and YCbCrToRGB
// (jpeg.cpp:511-561) plus the pixel gather of display.hpp:19-34 in ONE kernel: coefficients are
// read once (int16, 128 B per coded block), RGB8 is written once (interleaved, 128-bit stores).
//
// Exactness recipe: float32 everywhere, the reference's operation order
// (jpeg.cpp:598-662), no FMA contraction (__fmul_rn/__fadd_rn/__fsub_rn are never fused),
// truncation toward zero after each 1-D pass (the reference assigns float to int) and after the
// colour sums, clamp after truncation, constants as the bit patterns the reference's
// double->float initialisers produce (types.hpp:5-19).  int16 coefficient x 8-bit quantiser
// is < 2^24, so float(c)*float(q) is exact and equals the reference's int product converted.
//
// Mapping: a CTA of 96 threads owns a tile of 96 consecutive coded blocks of one MCU row
// (16 / 24 / 32 / 96 MCUs for 4:2:0 / 4:2:2,4:4:0 / 4:4:4 / grey).  Phase 1: one thread per
// block, all 64 values in registers, de-zig-zag by register naming, column pass, row pass,
// results to shared-memory planes.  Phase 2: one thread per 16 pixels of a row: nearest-neighbour
// chroma (jpeg.cpp:518-520), colour, pack, three 128-bit stores (48 contiguous bytes).
// Algorithmic bytes per tile: 96*128 in + 3 B per pixel out.
#pragma once
#include "platform.h"
#include "dev_types.h"
#include "kernels_unstuff.cuh"   // find_owner

namespace b2 {

constexpr int kPixThreads = 96;

B2_DEVICE float f32_bits(uint32_t u) {
#ifdef B200JPEG_EMU
  float f; memcpy(&f, &u, 4); return f;
#else
  return __uint_as_float(u);
#endif
}
// types.hpp:5-19 as float bit patterns (tests/test_oracle.py re-derives them)
#define B2_M1 f32_bits(0x3fb504f3u)
#define B2_M2 f32_bits(0x3f8a8bd4u)
#define B2_M3 f32_bits(0x3fb504f3u)
#define B2_M4 f32_bits(0x40273d74u)
#define B2_M5 f32_bits(0x3f43ef15u)
#define B2_S0 f32_bits(0x3eb504f3u)
#define B2_S1 f32_bits(0x3efb14beu)
#define B2_S2 f32_bits(0x3eec835eu)
#define B2_S3 f32_bits(0x3ed4db31u)
#define B2_S4 f32_bits(0x3eb504f3u)
#define B2_S5 f32_bits(0x3e8e39dau)
#define B2_S6 f32_bits(0x3e43ef15u)
#define B2_S7 f32_bits(0x3dc7c5c2u)
// jpeg.cpp:521-523
#define B2_CR_R f32_bits(0x3fb374bcu)   /* 1.402f */
#define B2_CB_G f32_bits(0x3eb020c5u)   /* 0.344f */
#define B2_CR_G f32_bits(0x3f36c8b4u)   /* 0.714f */
#define B2_CB_B f32_bits(0x3fe2d0e5u)   /* 1.772f */

// One AAN 1-D pass in place, jpeg.cpp:598-662 / 666-730.
B2_DEVICE void idct8(float &x0, float &x1, float &x2, float &x3, float &x4, float &x5, float &x6, float &x7) {
  const float g0 = __fmul_rn(x0, B2_S0), g1 = __fmul_rn(x4, B2_S4), g2 = __fmul_rn(x2, B2_S2), g3 = __fmul_rn(x6, B2_S6);
  const float g4 = __fmul_rn(x5, B2_S5), g5 = __fmul_rn(x1, B2_S1), g6 = __fmul_rn(x7, B2_S7), g7 = __fmul_rn(x3, B2_S3);
  const float f4 = __fsub_rn(g4, g7), f5 = __fadd_rn(g5, g6), f6 = __fsub_rn(g5, g6), f7 = __fadd_rn(g4, g7);
  const float e2 = __fsub_rn(g2, g3), e3 = __fadd_rn(g2, g3), e5 = __fsub_rn(f5, f7), e7 = __fadd_rn(f5, f7);
  const float e8 = __fadd_rn(f4, f6);
  const float d2 = __fmul_rn(e2, B2_M1), d4 = __fmul_rn(f4, B2_M2), d5 = __fmul_rn(e5, B2_M3);
  const float d6 = __fmul_rn(f6, B2_M4), d8 = __fmul_rn(e8, B2_M5);
  const float c0 = __fadd_rn(g0, g1), c1 = __fsub_rn(g0, g1), c2 = __fsub_rn(d2, e3), c4 = __fadd_rn(d4, d8);
  const float c5 = __fadd_rn(d5, e7), c6 = __fsub_rn(d6, d8), c8 = __fsub_rn(c5, c6);
  const float b0 = __fadd_rn(c0, e3), b1 = __fadd_rn(c1, c2), b2 = __fsub_rn(c1, c2), b3 = __fsub_rn(c0, e3);
  const float b4 = __fsub_rn(c4, c8), b6 = __fsub_rn(c6, e7);
  x0 = truncf(__fadd_rn(b0, e7)); x1 = truncf(__fadd_rn(b1, b6)); x2 = truncf(__fadd_rn(b2, c8));
  x3 = truncf(__fadd_rn(b3, b4)); x4 = truncf(__fsub_rn(b3, b4)); x5 = truncf(__fsub_rn(b2, c8));
  x6 = truncf(__fsub_rn(b1, b6)); x7 = truncf(__fsub_rn(b0, e7));
}

#define B2_ZZ_LIST                                                                                      \
  0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, \
      7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, \
      31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63

template <int HS, int VS, int NC>
struct PixCfg {
  static constexpr int kLuma = NC == 1 ? 1 : HS * VS;
  static constexpr int kBpm = NC == 1 ? 1 : HS * VS + 2;
  static constexpr int kTM = kPixThreads / kBpm;        // MCUs per tile
  static constexpr int kYW = kTM * 8 * HS;              // luma tile width (pixels)
  static constexpr int kYH = 8 * VS;
  static constexpr int kCW = NC == 1 ? 0 : kTM * 8;     // chroma tile width (samples)
  static constexpr int kGroups = kYW / 16;              // 16-pixel groups per tile row
  static constexpr int kItems = kYH * kGroups;
  static constexpr size_t kSmemFloats = (size_t)kYH * kYW + 2 * 8 * (size_t)kCW + 3 * 64;
};

B2_DEVICE uint32_t pack4(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  return __byte_perm(__byte_perm(a, b, 0x0040), __byte_perm(c, d, 0x0040), 0x5410);
}

// Tile lookup arrays are per sampling format: grp_img[g] = image index, grp_pre = exclusive
// prefix of tiles over the group's images (ngrp+1 entries).
template <int HS, int VS, int NC>
B2_GLOBAL void __launch_bounds__(kPixThreads)
k_pixel(const ImageDev *__restrict__ imgs, const int32_t *__restrict__ grp_img, const int32_t *__restrict__ grp_pre,
        int ngrp, const TableSet *__restrict__ tsets, const ImageOut *__restrict__ out,
        const int16_t *__restrict__ coef, uint8_t *__restrict__ rgb) {
  typedef PixCfg<HS, VS, NC> C;
  B2_DYN_SMEM(float, sm);
  float *yp = sm;                               // [kYH][kYW]
  float *cbp = sm + C::kYH * C::kYW;            // [8][kCW]
  float *crp = cbp + 8 * C::kCW;
  float *qf = crp + 8 * C::kCW;                 // [3][64] quantisers as float, zig-zag order
  B2_SHARED int s_g;
  const int tid = threadIdx.x;
  if (tid == 0) s_g = find_owner(grp_pre, ngrp, (int)blockIdx.x);
  __syncthreads();
  const int img = grp_img[s_g];
  const ImageDev im = imgs[img];
  if (out[img].status & (ST_BAD_MARKER | ST_NO_EOI | ST_BAD_RST_COUNT)) return;   // nothing was decoded
  const int t = (int)blockIdx.x - grp_pre[s_g];
  const int tiles_per_row = (im.mcux + C::kTM - 1) / C::kTM;
  const int my = t / tiles_per_row, mx0 = (t - my * tiles_per_row) * C::kTM;
  const int nm = min(C::kTM, im.mcux - mx0);

  constexpr int zz[64] = {B2_ZZ_LIST};
  {
    const TableSet &ts = tsets[im.tset];
    for (int k = tid; k < NC * 64; k += kPixThreads) {
      const int c = k >> 6, i = k & 63;
      int nat = 0;
#pragma unroll
      for (int j = 0; j < 64; ++j) if (j == i) nat = zz[j];
      qf[k] = (float)ts.q[im.tq[c] & 3][nat];
    }
  }
  __syncthreads();

  // ---- phase 1: one thread per coded block ----
  {
    const int m = tid / C::kBpm, b = tid - m * C::kBpm;
    if (m < nm) {
      const int64_t blk = im.coef_base + ((int64_t)my * im.mcux + mx0 + m) * C::kBpm + b;
      const uint4 *src = reinterpret_cast<const uint4 *>(coef + blk * 64);
      uint4 raw[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) raw[k] = b2_ld_nc_v4(src + k);
      const int comp = b < C::kLuma ? 0 : b - C::kLuma + 1;
      const float *q = qf + comp * 64;
      float v[64];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t w[4] = {raw[k].x, raw[k].y, raw[k].z, raw[k].w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int i0 = k * 8 + j * 2;       // position in zig-zag order
          const float c0 = (float)(int)(int16_t)(w[j] & 0xffffu);
          const float c1 = (float)(int)(int16_t)(w[j] >> 16);
          v[zz[i0]] = __fmul_rn(c0, q[i0]);        // jpeg.cpp:567 (exact, see header)
          v[zz[i0 + 1]] = __fmul_rn(c1, q[i0 + 1]);
        }
      }
#pragma unroll
      for (int i = 0; i < 8; ++i)                  // columns, jpeg.cpp:596-663
        idct8(v[0 * 8 + i], v[1 * 8 + i], v[2 * 8 + i], v[3 * 8 + i], v[4 * 8 + i], v[5 * 8 + i], v[6 * 8 + i], v[7 * 8 + i]);
#pragma unroll
      for (int i = 0; i < 8; ++i)                  // rows, jpeg.cpp:664-731
        idct8(v[i * 8 + 0], v[i * 8 + 1], v[i * 8 + 2], v[i * 8 + 3], v[i * 8 + 4], v[i * 8 + 5], v[i * 8 + 6], v[i * 8 + 7]);
      float *dst; int pitch;
      if (b < C::kLuma) {
        const int bv = b / HS, bu = b - bv * HS;
        dst = yp + (bv * 8) * C::kYW + (m * HS + bu) * 8; pitch = C::kYW;
      } else {
        dst = (b == C::kLuma ? cbp : crp) + m * 8; pitch = C::kCW;
      }
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        float4 *d4 = reinterpret_cast<float4 *>(dst + r * pitch);
        d4[0] = make_float4(v[r * 8 + 0], v[r * 8 + 1], v[r * 8 + 2], v[r * 8 + 3]);
        d4[1] = make_float4(v[r * 8 + 4], v[r * 8 + 5], v[r * 8 + 6], v[r * 8 + 7]);
      }
    }
  }
  __syncthreads();

  // ---- phase 2: one thread per 16 pixels of one row ----
  const int width_px = nm * 8 * HS;
  for (int it = tid; it < C::kItems; it += kPixThreads) {
    const int row = it / C::kGroups, x = (it - row * C::kGroups) * 16;
    const int gy = my * C::kYH + row, gx = mx0 * 8 * HS + x;
    if (x >= width_px || gy >= im.height || gx >= im.width) continue;
    float yv[16], cbv[16], crv[16];
    {
      const float4 *s = reinterpret_cast<const float4 *>(yp + row * C::kYW + x);
#pragma unroll
      for (int k = 0; k < 4; ++k) { const float4 a = s[k]; yv[4 * k] = a.x; yv[4 * k + 1] = a.y; yv[4 * k + 2] = a.z; yv[4 * k + 3] = a.w; }
    }
    if (NC == 3) {
      // chroma sample (row/VS, x/HS): the reference's (y/Vs + 4v)*8 + (x/Hs + 4h), jpeg.cpp:518-520
      const int crow = row / VS, cx = x / HS;
      const float4 *sb = reinterpret_cast<const float4 *>(cbp + crow * C::kCW + cx);
      const float4 *sr = reinterpret_cast<const float4 *>(crp + crow * C::kCW + cx);
#pragma unroll
      for (int k = 0; k < 16 / HS / 4; ++k) {
        const float4 a = sb[k], c = sr[k];
        cbv[4 * k] = a.x; cbv[4 * k + 1] = a.y; cbv[4 * k + 2] = a.z; cbv[4 * k + 3] = a.w;
        crv[4 * k] = c.x; crv[4 * k + 1] = c.y; crv[4 * k + 2] = c.z; crv[4 * k + 3] = c.w;
      }
    }
    uint32_t px[48];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      float r, g, b;
      if (NC == 3) {
        const float cb = cbv[k / HS], cr = crv[k / HS];
        r = __fadd_rn(__fadd_rn(yv[k], __fmul_rn(B2_CR_R, cr)), 128.0f);                                   // jpeg.cpp:521
        g = __fadd_rn(__fsub_rn(__fsub_rn(yv[k], __fmul_rn(B2_CB_G, cb)), __fmul_rn(B2_CR_G, cr)), 128.0f); // :522
        b = __fadd_rn(__fadd_rn(yv[k], __fmul_rn(B2_CB_B, cb)), 128.0f);                                   // :523
      } else {
        r = g = b = __fadd_rn(yv[k], 128.0f);   // Cb = Cr = 0: every product is +-0 and drops out
      }
      px[3 * k] = b2_f2u8_rz_sat(r); px[3 * k + 1] = b2_f2u8_rz_sat(g); px[3 * k + 2] = b2_f2u8_rz_sat(b); // :524-538
    }
    uint8_t *o = rgb + im.rgb_off + (uint64_t)gy * im.rgb_pitch + (uint64_t)gx * 3;
    uint4 *o4 = reinterpret_cast<uint4 *>(o);
    const uint32_t room = im.rgb_pitch - (uint32_t)gx * 3;
#pragma unroll
    for (int q = 0; q < 3; ++q) {
      if ((uint32_t)q * 16 < room) {
        const uint4 w = make_uint4(pack4(px[16 * q + 0], px[16 * q + 1], px[16 * q + 2], px[16 * q + 3]),
                                   pack4(px[16 * q + 4], px[16 * q + 5], px[16 * q + 6], px[16 * q + 7]),
                                   pack4(px[16 * q + 8], px[16 * q + 9], px[16 * q + 10], px[16 * q + 11]),
                                   pack4(px[16 * q + 12], px[16 * q + 13], px[16 * q + 14], px[16 * q + 15]));
        b2_st_cs_v4(o4 + q, w);
      }
    }
  }
}

// Optional export for parity checks: scan-order/zig-zag int16 -> the reference's MCU-array
// layout (types.hpp:32-50): int16 [bhR*bwR][3][64], natural order, chroma at anchor positions.
B2_GLOBAL void k_export_coef(ImageDev im, int bwr, const int16_t *__restrict__ coef, int16_t *__restrict__ dst) {
  constexpr int zz[64] = {B2_ZZ_LIST};
  const int k = threadIdx.x & 63;
  int nat = 0;
  for (int j = 0; j < 64; ++j) if (j == k) nat = zz[j];
  const int nluma = im.ncomp == 1 ? 1 : im.hs * im.vs;
  for (int64_t blk = (int64_t)blockIdx.x * (blockDim.x / 64) + threadIdx.x / 64; blk < im.nblocks;
       blk += (int64_t)gridDim.x * (blockDim.x / 64)) {
    const int mcu = (int)(blk / im.bpm), b = (int)(blk - (int64_t)mcu * im.bpm);
    const int my = mcu / im.mcux, mx = mcu - my * im.mcux;
    int comp, by, bx;
    if (b < nluma) { comp = 0; by = my * im.vs + b / im.hs; bx = mx * im.hs + b % im.hs; }
    else { comp = b - nluma + 1; by = my * im.vs; bx = mx * im.hs; }
    if (im.ncomp == 1) { by = my; bx = mx; }
    dst[(((int64_t)by * bwr + bx) * 3 + comp) * 64 + nat] = coef[(im.coef_base + blk) * 64 + k];
  }
}

}  // namespace b2
