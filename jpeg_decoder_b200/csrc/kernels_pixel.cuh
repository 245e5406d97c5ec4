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
// Mapping (warp-autonomous and persistent): every WARP owns a private slice of shared memory and walks the tile list with
// stride (warps in the grid).  A tile is kTM consecutive MCUs of one MCU row = at most 32 consecutive coded blocks
// (30 for 4:2:0 / 4:4:4), i.e. one contiguous run of the coefficient buffer.  Per tile:
//   load   lane 0 starts a TMA 1-D bulk copy (cp.async.bulk -> mbarrier) of the tile's coefficients into one of two
//          stage buffers -- issued one tile AHEAD, so HBM latency hides behind the previous tile's arithmetic and no
//          register is spent on prefetching;
//   idct   one lane per coded block, all 64 values in registers: de-zig-zag by register naming, dequantise, column
//          pass, row pass.  The entropy stage stores each block's eight 16-byte chunks XOR-swizzled by (block & 7) so
//          that the lanes' 128-bit reads of a dense copy are bank-conflict free.  Chroma lanes publish their block as
//          an fp32 plane in shared memory (same XOR on its chunks);
//   colour the 8 * kLuma luma rows of an MCU are shared out over ALL its lanes: a luma lane keeps rows [0, kKeep) and
//          hands the rest to the MCU's two chroma lanes through shared memory (transposed [chunk][slot] layout,
//          pix_spill_slot), so no lane idles; nearest-neighbour chroma from the planes (jpeg.cpp:518-520), the four
//          chroma products once per chroma sample, colour sums, truncate + clamp + pack, RGB rows into the (now free)
//          stage buffer;
//   store  the staged tile leaves through ONE TMA tensor store (uniform batches) or one TMA bulk copy per row.
// Only __syncwarp() separates the phases; there is no CTA-wide barrier in the kernel.
// Algorithmic bytes per tile: 128 B per coded block in + 3 B per pixel out.
#pragma once
#include "platform.h"
#include "dev_types.h"
#include "kernels_unstuff.cuh"   // find_owner

namespace b2 {

constexpr int kPixWarps = 4;                 // warps per CTA (each fully independent)
constexpr int kPixThreads = kPixWarps * 32;
constexpr int kQStride = 68;                 // floats between the per-component quantiser rows (272 B: rows land in different bank groups)

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

#ifndef B2_PIX_MAXCTAS
#define B2_PIX_MAXCTAS 4
#endif
// Truncation toward zero after each 1-D pass (the reference assigns float to int, jpeg.cpp:655-662): truncf = one FRND
// on the quarter-rate XU pipe.  Measured alternatives (profiles/r02_experiments.md): truncation by FADD.RZ against 2^23
// (no XU instruction, three on the FMA/ALU pipes) is 5 % SLOWER, int16 -> float through the mantissa of 1.5 * 2^23
// instead of I2F is neutral -- the kernel is bound by issue slots, not by the XU pipe -- so neither is kept.
#define B2_PIX_TRUNC(x) truncf(x)

// One AAN 1-D pass in place (jpeg.cpp:598-662 / 666-730), for TWO vectors at once (two columns, or two rows): the same
// operations in the same order on both, the additions as packed pairs (FADD2: half the instructions), the
// multiplications scalar (see b2_f2 in platform.h).
#define B2_PMUL(a, b, s) b2_f2_pack(__fmul_rn(a, s), __fmul_rn(b, s))
B2_DEVICE b2_f2 b2_f2_mul(b2_f2 v, float s) { float a, b; b2_f2_unpack(v, a, b); return B2_PMUL(a, b, s); }
B2_DEVICE void idct8x2(float &x0, float &x1, float &x2, float &x3, float &x4, float &x5, float &x6, float &x7,
                       float &y0, float &y1, float &y2, float &y3, float &y4, float &y5, float &y6, float &y7) {
  const b2_f2 g0 = B2_PMUL(x0, y0, B2_S0), g1 = B2_PMUL(x4, y4, B2_S4), g2 = B2_PMUL(x2, y2, B2_S2), g3 = B2_PMUL(x6, y6, B2_S6);
  const b2_f2 g4 = B2_PMUL(x5, y5, B2_S5), g5 = B2_PMUL(x1, y1, B2_S1), g6 = B2_PMUL(x7, y7, B2_S7), g7 = B2_PMUL(x3, y3, B2_S3);
  const b2_f2 f4 = b2_f2_sub(g4, g7), f5 = b2_f2_add(g5, g6), f6 = b2_f2_sub(g5, g6), f7 = b2_f2_add(g4, g7);
  const b2_f2 e2 = b2_f2_sub(g2, g3), e3 = b2_f2_add(g2, g3), e5 = b2_f2_sub(f5, f7), e7 = b2_f2_add(f5, f7);
  const b2_f2 e8 = b2_f2_add(f4, f6);
  const b2_f2 d2 = b2_f2_mul(e2, B2_M1), d4 = b2_f2_mul(f4, B2_M2), d5 = b2_f2_mul(e5, B2_M3);
  const b2_f2 d6 = b2_f2_mul(f6, B2_M4), d8 = b2_f2_mul(e8, B2_M5);
  const b2_f2 c0 = b2_f2_add(g0, g1), c1 = b2_f2_sub(g0, g1), c2 = b2_f2_sub(d2, e3), c4 = b2_f2_add(d4, d8);
  const b2_f2 c5 = b2_f2_add(d5, e7), c6 = b2_f2_sub(d6, d8), c8 = b2_f2_sub(c5, c6);
  const b2_f2 b0 = b2_f2_add(c0, e3), b1 = b2_f2_add(c1, c2), b2 = b2_f2_sub(c1, c2), b3 = b2_f2_sub(c0, e3);
  const b2_f2 b4 = b2_f2_sub(c4, c8), b6 = b2_f2_sub(c6, e7);
  float a, b;
  b2_f2_unpack(b2_f2_add(b0, e7), a, b); x0 = B2_PIX_TRUNC(a); y0 = B2_PIX_TRUNC(b);
  b2_f2_unpack(b2_f2_add(b1, b6), a, b); x1 = B2_PIX_TRUNC(a); y1 = B2_PIX_TRUNC(b);
  b2_f2_unpack(b2_f2_add(b2, c8), a, b); x2 = B2_PIX_TRUNC(a); y2 = B2_PIX_TRUNC(b);
  b2_f2_unpack(b2_f2_add(b3, b4), a, b); x3 = B2_PIX_TRUNC(a); y3 = B2_PIX_TRUNC(b);
  b2_f2_unpack(b2_f2_sub(b3, b4), a, b); x4 = B2_PIX_TRUNC(a); y4 = B2_PIX_TRUNC(b);
  b2_f2_unpack(b2_f2_sub(b2, c8), a, b); x5 = B2_PIX_TRUNC(a); y5 = B2_PIX_TRUNC(b);
  b2_f2_unpack(b2_f2_sub(b1, b6), a, b); x6 = B2_PIX_TRUNC(a); y6 = B2_PIX_TRUNC(b);
  b2_f2_unpack(b2_f2_sub(b0, e7), a, b); x7 = B2_PIX_TRUNC(a); y7 = B2_PIX_TRUNC(b);
}

#define B2_ZZ_LIST                                                                                      \
  0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, \
      7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, \
      31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63
B2_CONSTANT uint8_t c_zigzag[64] = {B2_ZZ_LIST};   // zig-zag position -> natural index (types.hpp:23-31)

// Luma rows handed from the luma lanes to the chroma lanes ("spill"): 16-byte chunk c (row r - kKeep, left / right half)
// of luma block `lb` of MCU m.  Transposed layout [chunk][slot], slots in the order the READERS walk them, so that both
// the 128-bit stores of the luma lanes and the 128-bit loads of the chroma lanes touch consecutive 16-byte words
// (measured before, with [block][row] at 64 / 128-byte lane strides: 10 wavefronts per instruction where 2-3 suffice).
template <int HS, int VS, int TM>
B2_DEVICE int pix_spill_slot(int m, int lb) {
  if (HS * VS == 4) return (lb & 1) * (TM * 2) + m * 2 + (lb >> 1);   // a chroma lane reads block 2*cidx + (i >> 1): (m, cidx) consecutive
  return m * (HS * VS) + lb;                                          // 4:2:2 / 4:4:0: block cidx; 4:4:4: block 0
}

template <int HS, int VS, int NC>
struct PixCfg {
  static constexpr int kLuma = NC == 1 ? 1 : HS * VS;
  static constexpr int kBpm = NC == 1 ? 1 : HS * VS + 2;
  static constexpr int kTM = 32 / kBpm;                  // MCUs per warp tile: 10 / 8 / 8 / 5 / 32
  static constexpr int kLanes = kTM * kBpm;              // lanes that own a block: 30 / 32 / 32 / 30 / 32
  static constexpr int kW = kTM * 8 * HS;                // tile width in pixels
  static constexpr int kH = 8 * VS;
  static constexpr int kRowBytes = 3 * kW;               // multiple of 16 for every format
  static constexpr int kInBytes = kLanes * 128;
  static constexpr int kOutBytes = kH * kRowBytes;
  static constexpr int kStageBytes = ((kInBytes > kOutBytes ? kInBytes : kOutBytes) + 127) / 128 * 128;
  static constexpr int kKeep = NC == 1 ? 8 : (HS * VS == 4 ? 6 : (HS * VS == 2 ? 4 : 3));   // luma rows a luma lane converts itself
  static constexpr int kTake = NC == 1 ? 0 : (HS * VS == 1 ? 3 : 4);                          // row iterations of a chroma lane
  static constexpr int kChromaBytes = NC == 3 ? kTM * 2 * 256 : 0;   // fp32 planes of the chroma blocks only
  static constexpr int kSpillBytes = NC == 3 ? kTM * kLuma * (8 - kKeep) * 32 : 0;   // luma rows handed to the chroma lanes
  static constexpr int kPlaneBytes = kChromaBytes + kSpillBytes;
  static constexpr int kQBytes = 3 * kQStride * 4 + 16;  // 832
  static constexpr int kWarpBytes = (2 * kStageBytes + kPlaneBytes + kQBytes + 16 + 127) / 128 * 128;
  static constexpr size_t kSmemBytes = (size_t)kPixWarps * kWarpBytes;
  // resident CTAs per SM the register budget must allow (shared memory allows this many)
  static constexpr int kMinCtas = (int)(232448 / (kSmemBytes + 1024)) > B2_PIX_MAXCTAS ? B2_PIX_MAXCTAS : (int)(232448 / (kSmemBytes + 1024));
};

struct PixTile {
  int img, my, mx0, nm;
  int width, height, qkey;
  uint32_t pitch, dead;
  int64_t blk0;     // global index (in the coefficient buffer) of the tile's first block
  uint64_t rgb_off;
};

// A sampling-format group whose images all have the same geometry and tables and lie at a constant
// stride in the coefficient / RGB buffers (the normal batch): the tile -> image mapping is then pure
// arithmetic on kernel parameters, with no dependent global loads at the head of a tile.
struct PixUni {
  int32_t enabled;
  int32_t img0;                 // images are img0, img0+1, ...
  int32_t mcux, tpr, tpi;       // MCUs per row, tiles per MCU row, tiles per image
  int32_t width, height, qkey;
  uint32_t pitch;
  uint32_t tmap_ok;             // the launch carries a tensor map of the group's RGB output (whole-tile TMA stores)
  uint64_t magic_tpi, magic_tpr;   // ceil(2^40 / d): n / d == (n * magic) >> 40 for n < 2^24, d < 2^16
  int64_t coef_base0, coef_stride; // blocks
  uint64_t rgb_off0, rgb_stride;   // bytes
};

// An image whose coefficient blocks were not ALL written this batch (stage U gave up on the scan, or the entropy stage
// did not find the blocks the header promises) gets black pixels: the coefficient and RGB arenas are reused from batch
// to batch, and nothing of an earlier batch -- possibly another caller's picture -- may come back for a corrupt upload.
constexpr uint32_t kStDead = ST_BAD_MARKER | ST_NO_EOI | ST_BAD_RST_COUNT | ST_BLOCK_COUNT | ST_OVERRUN | ST_NOT_CONVERGED;
B2_DEVICE uint32_t pix_dead(const ImageOut *__restrict__ out, int img) { return (out[img].status & kStDead) != 0; }
B2_DEVICE int pix_qkey(const ImageDev &im) {
  return (im.tset << 6) | (im.tq[0] & 3) | ((im.tq[1] & 3) << 2) | ((im.tq[2] & 3) << 4);
}

// Tile t of a sampling-format group -> image, MCU row, first MCU, MCU count.  General path: `g` walks
// forward only (a warp's tile numbers increase), so the lookup is O(1) amortised.
template <int TM, int BPM>
B2_DEVICE PixTile pix_locate(int t, int &g, const PixUni &u, const ImageDev *__restrict__ imgs,
                             const int32_t *__restrict__ grp_img, const int32_t *__restrict__ grp_pre,
                             const ImageOut *__restrict__ out) {
  PixTile T;
  if (u.enabled) {
    const int gi = (int)(((uint64_t)(uint32_t)t * u.magic_tpi) >> 40);
    const int lt = t - gi * u.tpi;
    T.img = u.img0 + gi;
    T.my = (int)(((uint64_t)(uint32_t)lt * u.magic_tpr) >> 40);
    T.mx0 = (lt - T.my * u.tpr) * TM;
    T.nm = min(TM, u.mcux - T.mx0);
    T.blk0 = u.coef_base0 + (int64_t)gi * u.coef_stride + ((int64_t)T.my * u.mcux + T.mx0) * BPM;
    T.rgb_off = u.rgb_off0 + (uint64_t)gi * u.rgb_stride;
    T.width = u.width; T.height = u.height; T.pitch = u.pitch; T.qkey = u.qkey;
    T.dead = pix_dead(out, T.img);
    return T;
  }
  while (t >= grp_pre[g + 1]) ++g;
  T.img = grp_img[g];
  const ImageDev &im = imgs[T.img];
  const int mcux = im.mcux;
  const int tpr = (mcux + TM - 1) / TM;
  const int lt = t - grp_pre[g];
  T.my = lt / tpr;
  T.mx0 = (lt - T.my * tpr) * TM;
  T.nm = min(TM, mcux - T.mx0);
  T.blk0 = im.coef_base + ((int64_t)T.my * mcux + T.mx0) * BPM;
  T.rgb_off = im.rgb_off;
  T.width = im.width; T.height = im.height; T.pitch = im.rgb_pitch; T.qkey = pix_qkey(im);
  T.dead = pix_dead(out, T.img);
  return T;
}

// Tile lookup arrays are per sampling format: grp_img[g] = image index, grp_pre = exclusive
// prefix of tiles over the group's images (ngrp+1 entries).
template <int HS, int VS, int NC>
B2_GLOBAL void __launch_bounds__(kPixThreads, (PixCfg<HS, VS, NC>::kMinCtas))
k_pixel(const ImageDev *__restrict__ imgs, const int32_t *__restrict__ grp_img, const int32_t *__restrict__ grp_pre,
        int ngrp, const PixUni uni, const TableSet *__restrict__ tsets, const ImageOut *__restrict__ out,
        const int16_t *__restrict__ coef, uint8_t *__restrict__ rgb, const B2_GRID_CONSTANT B2TensorMap tmap) {
  typedef PixCfg<HS, VS, NC> C;
  B2_DYN_SMEM(unsigned char, smraw);
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  unsigned char *wsm = smraw + (size_t)wib * C::kWarpBytes;
  unsigned char *planes = wsm + 2 * C::kStageBytes;
  unsigned char *spill = planes + C::kChromaBytes;
  float *qf = reinterpret_cast<float *>(planes + C::kPlaneBytes);
  uint64_t *mbar = reinterpret_cast<uint64_t *>(planes + C::kPlaneBytes + C::kQBytes);

  const int ntiles = grp_pre[ngrp];
  const int GW = (int)gridDim.x * kPixWarps;
  int t = (int)blockIdx.x * kPixWarps + wib;
  if (t >= ntiles) return;

  if (lane == 0) { b2_mbar_init(&mbar[0], 1); b2_mbar_init(&mbar[1], 1); b2_mbar_init_fence(); }
  __syncwarp();

  constexpr int zz[64] = {B2_ZZ_LIST};
  int g = 0, q_key = -1;
  PixTile cur = pix_locate<C::kTM, C::kBpm>(t, g, uni, imgs, grp_img, grp_pre, out);
  if (lane == 0) b2_bulk_load(wsm, coef + cur.blk0 * 64, (uint32_t)(cur.nm * C::kBpm * 128), &mbar[0]);

  for (int n = 0;; ++n) {
    const int s = n & 1;
    unsigned char *st_cur = wsm + s * C::kStageBytes, *st_nxt = wsm + (s ^ 1) * C::kStageBytes;
    // ---- start the NEXT tile's copy into the other stage (free since the end of the last iteration)
    const int tn = t + GW;
    PixTile nxt = cur;
    if (tn < ntiles) {
      nxt = pix_locate<C::kTM, C::kBpm>(tn, g, uni, imgs, grp_img, grp_pre, out);
      if (lane == 0) {
        b2_fence_proxy_async();
        b2_bulk_load(st_nxt, coef + nxt.blk0 * 64, (uint32_t)(nxt.nm * C::kBpm * 128), &mbar[s ^ 1]);
      }
    }
    const int img = cur.img;
    const bool dead = cur.dead != 0;
    const int width = cur.width, height = cur.height;

    // ---- quantiser rows (float, zig-zag order), rebuilt only when the TABLES change: consecutive tiles of
    //      a warp usually lie in different images, but images of a batch mostly share one table set
    {
      const int qkey = cur.qkey;
      if (qkey != q_key) {
        const TableSet &ts = tsets[imgs[img].tset];
        for (int k = lane; k < NC * 64; k += 32) {
          const int c = k >> 6, i = k & 63;
          qf[c * kQStride + i] = (float)ts.q[imgs[img].tq[c] & 3][c_zigzag[i]];
        }
        q_key = qkey;
        __syncwarp();
      }
    }

    b2_mbar_wait(&mbar[s], (uint32_t)(n >> 1) & 1u);

    // ---- idct: one lane per coded block; the block stays in registers ----
    const bool has_blk = lane < cur.nm * C::kBpm && !dead;
    const int b = lane % C::kBpm, m = lane / C::kBpm;
    float v[64];
    if (has_blk) {
      const int comp = b < C::kLuma ? 0 : b - C::kLuma + 1;
      const uint32_t key = (uint32_t)(cur.blk0 + lane) & 7u;
      const uint4 *src = reinterpret_cast<const uint4 *>(st_cur + lane * 128);
      const float4 *q4 = reinterpret_cast<const float4 *>(qf + comp * kQStride);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint4 raw = src[(uint32_t)k ^ key];
        const float4 qa = q4[2 * k], qb = q4[2 * k + 1];
        const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
        const float q[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int i0 = k * 8 + j * 2;       // position in zig-zag order
          const float c0 = (float)(int)(int16_t)(w[j] & 0xffffu);
          const float c1 = (float)(int)(int16_t)(w[j] >> 16);
          v[zz[i0]] = __fmul_rn(c0, q[2 * j]);          // jpeg.cpp:567 (exact, see header)
          v[zz[i0 + 1]] = __fmul_rn(c1, q[2 * j + 1]);
        }
      }
#pragma unroll
      for (int i = 0; i < 8; i += 2)               // columns, two at a time, jpeg.cpp:596-663
        idct8x2(v[0 * 8 + i], v[1 * 8 + i], v[2 * 8 + i], v[3 * 8 + i], v[4 * 8 + i], v[5 * 8 + i], v[6 * 8 + i], v[7 * 8 + i],
                v[0 * 8 + i + 1], v[1 * 8 + i + 1], v[2 * 8 + i + 1], v[3 * 8 + i + 1], v[4 * 8 + i + 1], v[5 * 8 + i + 1],
                v[6 * 8 + i + 1], v[7 * 8 + i + 1]);
#pragma unroll
      for (int i = 0; i < 8; i += 2)               // rows, two at a time, jpeg.cpp:664-731
        idct8x2(v[i * 8 + 0], v[i * 8 + 1], v[i * 8 + 2], v[i * 8 + 3], v[i * 8 + 4], v[i * 8 + 5], v[i * 8 + 6], v[i * 8 + 7],
                v[i * 8 + 8], v[i * 8 + 9], v[i * 8 + 10], v[i * 8 + 11], v[i * 8 + 12], v[i * 8 + 13], v[i * 8 + 14],
                v[i * 8 + 15]);
      if (NC == 3 && b < C::kLuma) {               // luma lanes hand rows [kKeep,8) to the chroma lanes
        float4 *sp = reinterpret_cast<float4 *>(spill) + pix_spill_slot<HS, VS, C::kTM>(m, b);
        constexpr int kSlots = C::kTM * C::kLuma;
#pragma unroll
        for (int r = C::kKeep; r < 8; ++r) {
          sp[((r - C::kKeep) * 2) * kSlots] = make_float4(v[r * 8 + 0], v[r * 8 + 1], v[r * 8 + 2], v[r * 8 + 3]);
          sp[((r - C::kKeep) * 2 + 1) * kSlots] = make_float4(v[r * 8 + 4], v[r * 8 + 5], v[r * 8 + 6], v[r * 8 + 7]);
        }
      }
      if (NC == 3 && b >= C::kLuma) {              // chroma lanes publish their block: plane 2m (Cb), 2m+1 (Cr)
        const int ord = m * 2 + (b - C::kLuma);
        float4 *dst = reinterpret_cast<float4 *>(planes + ord * 256);
        const uint32_t pk = (uint32_t)ord & 7u;
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          dst[(uint32_t)(2 * r) ^ pk] = make_float4(v[r * 8 + 0], v[r * 8 + 1], v[r * 8 + 2], v[r * 8 + 3]);
          dst[(uint32_t)(2 * r + 1) ^ pk] = make_float4(v[r * 8 + 4], v[r * 8 + 5], v[r * 8 + 6], v[r * 8 + 7]);
        }
      }
    }
    __syncwarp();

    // ---- colour: the 8*kLuma luma rows of an MCU are shared out over ALL its lanes.  A luma lane keeps
    //      rows [0,kKeep) of its block in registers and spills the rest to shared memory; the two chroma
    //      lanes (whose own block now lives in the planes) pick those rows up into their registers.
    //      Every lane then runs the same kKeep row iterations; RGB goes to the stage buffer the copy
    //      came through ----
    const bool is_luma = b < C::kLuma;
    const int cidx = b - C::kLuma;                  // 0 = Cb lane, 1 = Cr lane
    if (NC == 3) {
      if (has_blk && !is_luma) {
#pragma unroll
        for (int i = 0; i < C::kKeep; ++i) {
          int lb, r;
          if (HS * VS == 4) { lb = 2 * cidx + (i >> 1); r = 6 + (i & 1); }
          else if (HS * VS == 2) { lb = cidx; r = 4 + i; }
          else { lb = 0; r = 3 + 3 * cidx + i; }
          if (i < C::kTake && r < 8) {
            const float4 *sp = reinterpret_cast<const float4 *>(spill) + pix_spill_slot<HS, VS, C::kTM>(m, lb);
            constexpr int kSlots = C::kTM * C::kLuma;
            const float4 a = sp[((r - C::kKeep) * 2) * kSlots], c = sp[((r - C::kKeep) * 2 + 1) * kSlots];
            v[i * 8 + 0] = a.x; v[i * 8 + 1] = a.y; v[i * 8 + 2] = a.z; v[i * 8 + 3] = a.w;
            v[i * 8 + 4] = c.x; v[i * 8 + 5] = c.y; v[i * 8 + 6] = c.z; v[i * 8 + 7] = c.w;
          }
        }
      }
    }
    if (has_blk) {
      const float4 *bp = reinterpret_cast<const float4 *>(planes + (m * 2) * 256);
      const float4 *rp = reinterpret_cast<const float4 *>(planes + (m * 2 + 1) * 256);
      const uint32_t bk = (uint32_t)(m * 2) & 7u, rk = bk | 1u;
      float p_r[8 / HS], p_g1[8 / HS], p_g2[8 / HS], p_b[8 / HS];
#pragma unroll
      for (int i = 0; i < C::kKeep; ++i) {
        // which luma block / row this lane converts in iteration i
        int lb = b, r = i;
        bool valid = true;
        if (NC == 3 && !is_luma) {
          if (HS * VS == 4) { lb = 2 * cidx + (i >> 1); r = 6 + (i & 1); }
          else if (HS * VS == 2) { lb = cidx; r = 4 + i; }
          else { lb = 0; r = 3 + 3 * cidx + i; }
          valid = i < C::kTake && r < 8;
          if (!valid) { lb = 0; r = 0; }
        }
        const int by = NC == 1 ? 0 : lb / HS, bx = NC == 1 ? 0 : lb % HS;
        if (NC == 3 && (i % VS) == 0) {
          // chroma sample (row/VS, x/HS): the reference's (y/Vs + 4v)*8 + (x/Hs + 4h), jpeg.cpp:518-520
          const uint32_t crow = (uint32_t)((by * 8 + (r & 7)) / VS);
          float cbv[8 / HS], crv[8 / HS];
          if (HS == 2) {
            const uint32_t ch = crow * 2u + (uint32_t)bx;
            const float4 a = bp[ch ^ bk], c = rp[ch ^ rk];
            cbv[0] = a.x; cbv[1] = a.y; cbv[2] = a.z; cbv[3] = a.w;
            crv[0] = c.x; crv[1] = c.y; crv[2] = c.z; crv[3] = c.w;
          } else {
            const float4 a0 = bp[(crow * 2u) ^ bk], a1 = bp[(crow * 2u + 1u) ^ bk];
            const float4 c0 = rp[(crow * 2u) ^ rk], c1 = rp[(crow * 2u + 1u) ^ rk];
            cbv[0] = a0.x; cbv[1] = a0.y; cbv[2] = a0.z; cbv[3] = a0.w;
            crv[0] = c0.x; crv[1] = c0.y; crv[2] = c0.z; crv[3] = c0.w;
            if (HS == 1) {
              cbv[(8 / HS) - 4] = a1.x; cbv[(8 / HS) - 3] = a1.y; cbv[(8 / HS) - 2] = a1.z; cbv[(8 / HS) - 1] = a1.w;
              crv[(8 / HS) - 4] = c1.x; crv[(8 / HS) - 3] = c1.y; crv[(8 / HS) - 2] = c1.z; crv[(8 / HS) - 1] = c1.w;
            }
          }
#pragma unroll
          for (int j = 0; j < 8 / HS; ++j) {        // the four products of jpeg.cpp:521-523, once per chroma sample
            p_r[j] = __fmul_rn(B2_CR_R, crv[j]); p_g1[j] = __fmul_rn(B2_CB_G, cbv[j]);
            p_g2[j] = __fmul_rn(B2_CR_G, crv[j]); p_b[j] = __fmul_rn(B2_CB_B, cbv[j]);
          }
        }
        float px[24];
        const b2_f2 c128 = b2_f2_pack(128.0f, 128.0f);
#pragma unroll
        for (int k = 0; k < 8; k += 2) {            // two pixels per packed addition (the sums in the reference's order)
          const b2_f2 y = b2_f2_pack(v[i * 8 + k], v[i * 8 + k + 1]);
          if (NC == 3) {
            const int j0 = k / HS, j1 = (k + 1) / HS;
            const b2_f2 r2 = b2_f2_add(b2_f2_add(y, b2_f2_pack(p_r[j0], p_r[j1])), c128);                                  // jpeg.cpp:521
            const b2_f2 g2 = b2_f2_add(b2_f2_sub(b2_f2_sub(y, b2_f2_pack(p_g1[j0], p_g1[j1])), b2_f2_pack(p_g2[j0], p_g2[j1])), c128);   // :522
            const b2_f2 bl2 = b2_f2_add(b2_f2_add(y, b2_f2_pack(p_b[j0], p_b[j1])), c128);                                 // :523
            b2_f2_unpack(r2, px[3 * k], px[3 * k + 3]);
            b2_f2_unpack(g2, px[3 * k + 1], px[3 * k + 4]);
            b2_f2_unpack(bl2, px[3 * k + 2], px[3 * k + 5]);
          } else {                                  // Cb = Cr = 0: every product is +-0 and drops out
            float a, c;
            b2_f2_unpack(b2_f2_add(y, c128), a, c);
            px[3 * k] = px[3 * k + 1] = px[3 * k + 2] = a;
            px[3 * k + 3] = px[3 * k + 4] = px[3 * k + 5] = c;
          }
        }
        if (valid) {
          uint2 *o = reinterpret_cast<uint2 *>(st_cur + (by * 8 + r) * C::kRowBytes + (m * 8 * HS + bx * 8) * 3);
#pragma unroll
          for (int q = 0; q < 3; ++q)               // truncate, clamp (jpeg.cpp:524-535), pack
            o[q] = make_uint2(b2_f4_to_u8x4(px[8 * q + 0], px[8 * q + 1], px[8 * q + 2], px[8 * q + 3]),
                              b2_f4_to_u8x4(px[8 * q + 4], px[8 * q + 5], px[8 * q + 6], px[8 * q + 7]));
        }
      }
    }
    if (dead) {                                    // see pix_dead: the tile leaves as zeros
      uint4 *z = reinterpret_cast<uint4 *>(st_cur);
      for (int k = lane; k < C::kOutBytes / 16; k += 32) z[k] = make_uint4(0u, 0u, 0u, 0u);
    }
    __syncwarp();

    // ---- store: staged rows -> global, 16 bytes per lane, lanes contiguous ----
    {
      const int gy0 = cur.my * C::kH, gx0 = cur.mx0 * 8 * HS;
      const uint32_t pitch = cur.pitch;
      const int rows = min(C::kH, height - gy0);
      const int room = (int)pitch - gx0 * 3;                                  // multiple of 16
      int rowb = (cur.nm * 8 * HS * 3 + 15) & ~15;
      if (rowb > room) rowb = room;
      uint8_t *gbase = rgb + cur.rgb_off + (uint64_t)gy0 * pitch + (uint64_t)gx0 * 3;
      // one TMA bulk store per row, issued by the lane of that row number; the group is drained
      // (source read) before the stage buffer is handed back to the load pipeline
      b2_fence_proxy_async();
      __syncwarp();
      if (uni.tmap_ok) {
        // the whole tile in one instruction: box kRowBytes x kH at (byte column / 4, row, image); rows past the
        // image's height and columns past its pitch are clipped by the TMA unit
        if (lane == 0 && gx0 < width && rows > 0) {
          b2_tensor_store_3d(&tmap, st_cur, gx0 * 3 / 4, gy0, cur.img - uni.img0);
          b2_bulk_commit();
          b2_bulk_wait_read();
        }
      } else if (gx0 < width && lane < rows && rowb > 0) {
        b2_bulk_store(gbase + (uint64_t)lane * pitch, st_cur + lane * C::kRowBytes, (uint32_t)rowb);
        b2_bulk_commit();
        b2_bulk_wait_read();
      }
    }
    __syncwarp();      // the stage buffer is the target of the copy started at the top of the next iteration

    if (tn >= ntiles) break;
    t = tn;
    cur = nxt;
  }
}

// Optional export for parity checks: scan-order/zig-zag/chunk-swizzled int16 -> the reference's
// MCU-array layout (types.hpp:32-50): int16 [bhR*bwR][3][64], natural order, chroma at anchor positions.
B2_GLOBAL void k_export_coef(ImageDev im, int bwr, const int16_t *__restrict__ coef, int16_t *__restrict__ dst) {
  const int k = threadIdx.x & 63;
  const int nat = c_zigzag[k];
  const int nluma = im.ncomp == 1 ? 1 : im.hs * im.vs;
  for (int64_t blk = (int64_t)blockIdx.x * (blockDim.x / 64) + threadIdx.x / 64; blk < im.nblocks;
       blk += (int64_t)gridDim.x * (blockDim.x / 64)) {
    const int mcu = (int)(blk / im.bpm), b = (int)(blk - (int64_t)mcu * im.bpm);
    const int my = mcu / im.mcux, mx = mcu - my * im.mcux;
    int comp, by, bx;
    if (b < nluma) { comp = 0; by = my * im.vs + b / im.hs; bx = mx * im.hs + b % im.hs; }
    else { comp = b - nluma + 1; by = my * im.vs; bx = mx * im.hs; }
    if (im.ncomp == 1) { by = my; bx = mx; }
    const int64_t gb = im.coef_base + blk;
    const int phys = (((k >> 3) ^ (int)(gb & 7)) << 3) | (k & 7);          // undo the chunk swizzle
    dst[(((int64_t)by * bwr + bx) * 3 + comp) * 64 + nat] = coef[gb * 64 + phys];
  }
}

}  // namespace b2
