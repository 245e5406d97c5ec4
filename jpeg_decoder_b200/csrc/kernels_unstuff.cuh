// kernels_unstuff.cuh -- stage U: byte unstuffing, RSTn stripping and segment discovery.
//
// Replaces JPEGFile::readImageData (include/file.hpp:59-104), the per-byte loop that builds
// `clean_data`.  That loop is a two-state machine (FRESH / AFTER_FF) whose state before byte j
// is AFTER_FF exactly when raw[j-1] == 0xFF (FF FF keeps it in AFTER_FF, file.hpp:87-90), so
// every byte can be classified from (raw[j-1], raw[j]) alone:
//     prev==FF: 00 -> emit FF | D0..D7 -> restart marker, nothing emitted | FF -> nothing
//               D9 -> end of scan | anything else -> error (file.hpp:91-95)
//     prev!=FF: FF -> nothing yet | else emit the byte
// Three kernels: count per 4 KiB tile, scan tile counts per image, scatter.  The unstuffed
// stream is written byte-swapped inside each 32-bit word (address ^ 3) so that a plain word
// load yields MSB-first bit order for the entropy decoder.  Each RSTn found becomes a segment
// start (the reference drops the marker and byte-aligns by MCU count, jpeg.cpp:419-425; for a
// valid stream both put the boundary at the same bit).
//
// HBM traffic: reads raw twice (count, scatter), writes it once: 3 B per scan byte.
#pragma once
#include "platform.h"
#include "dev_types.h"

namespace b2 {

constexpr int kUThreads = 256;
constexpr int kUBytesPerThread = 16;
constexpr int kUTile = kUThreads * kUBytesPerThread;

// largest i in [0,n) with base[i] <= idx (base is an exclusive prefix with n+1 entries)
B2_DEVICE int find_owner(const int32_t *__restrict__ base, int n, int idx) {
  int lo = 0, hi = n - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (base[mid] <= idx) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// Exclusive scan of one u32 per thread over a CTA of kUThreads; returns the prefix, *total = sum.
B2_DEVICE uint32_t cta_excl_scan_u32(uint32_t v, uint32_t *total) {
  B2_SHARED uint32_t warp_sum[32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  uint32_t inc = v;
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31 || (int)threadIdx.x == (int)blockDim.x - 1) warp_sum[wid] = inc;
  __syncthreads();
  uint32_t base = 0, tot = 0;
  for (int w = 0; w < nw; ++w) { const uint32_t s = warp_sum[w]; if (w < wid) base += s; tot += s; }
  __syncthreads();
  *total = tot;
  return base + inc - v;
}

struct UClass {
  uint32_t emit_mask;   // bit k: byte k of this thread emits one unstuffed byte
  uint32_t rst_mask;    // bit k: byte k completes an RSTn marker
  uint32_t end_mask;    // bit k: byte k completes FF D9
  uint32_t err_mask;    // bit k: byte k completes an illegal marker
  uint8_t bytes[16];
  uint8_t prev;
};

B2_DEVICE void u_classify(const uint8_t *__restrict__ raw, uint32_t pos, uint32_t raw_len, UClass &c) {
  c.emit_mask = c.rst_mask = c.end_mask = c.err_mask = 0;
  if (pos >= raw_len) { c.prev = 0; return; }
  const uint4 v = b2_ld_nc_v4(reinterpret_cast<const uint4 *>(raw + pos));
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
  uint8_t prev = pos ? raw[pos - 1] : 0;
  c.prev = prev;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const uint8_t cur = (uint8_t)(w[k >> 2] >> (8 * (k & 3)));
    c.bytes[k] = cur;
    if (pos + k < raw_len) {
      if (prev == 0xff) {
        if (cur == 0x00) c.emit_mask |= 1u << k;
        else if ((cur & 0xf8) == 0xd0) c.rst_mask |= 1u << k;
        else if (cur == 0xd9) c.end_mask |= 1u << k;
        else if (cur != 0xff) c.err_mask |= 1u << k;
      } else if (cur != 0xff) c.emit_mask |= 1u << k;
    }
    prev = cur;
  }
}

// U1: per-tile counts and the first EOI / first illegal marker of each image.
B2_GLOBAL void k_unstuff_count(const ImageDev *__restrict__ imgs, const int32_t *__restrict__ tile_base, int nimg,
                               const uint8_t *__restrict__ d_raw, uint32_t *__restrict__ tile_cnt,
                               ImageOut *__restrict__ out) {
  B2_SHARED int s_img;
  if (threadIdx.x == 0) s_img = find_owner(tile_base, nimg, (int)blockIdx.x);
  __syncthreads();
  const int i = s_img;
  const ImageDev im = imgs[i];
  const uint32_t pos = ((uint32_t)blockIdx.x - (uint32_t)im.tile_base) * kUTile + threadIdx.x * kUBytesPerThread;
  UClass c;
  u_classify(d_raw + im.raw_off, pos, im.raw_len, c);
  if (c.end_mask) atomicMin(&out[i].end_pos, pos + (uint32_t)__ffs((int)c.end_mask) - 1);
  if (c.err_mask) atomicMin(&out[i].err_pos, pos + (uint32_t)__ffs((int)c.err_mask) - 1);
  uint32_t tot;
  cta_excl_scan_u32((uint32_t)__popc(c.emit_mask) | ((uint32_t)__popc(c.rst_mask) << 16), &tot);
  if (threadIdx.x == 0) tile_cnt[blockIdx.x] = tot;   // emits in the low half, RSTs in the high half
}

// U2: one warp per image turns tile counts into exclusive prefixes (emits and RSTs separately).
B2_GLOBAL void k_unstuff_scan(const ImageDev *__restrict__ imgs, int nimg, const uint32_t *__restrict__ tile_cnt,
                              uint2 *__restrict__ tile_pre) {
  const int i = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (i >= nimg) return;
  const int t0 = imgs[i].tile_base, nt = imgs[i].ntiles;
  uint32_t carry_e = 0, carry_r = 0;
  for (int b = 0; b < nt; b += 32) {
    const int t = b + lane;
    const uint32_t c = t < nt ? tile_cnt[t0 + t] : 0;
    uint32_t e = c & 0xffffu, r = c >> 16;
    const uint32_t e0 = e, r0 = r;
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t oe = __shfl_up_sync(0xffffffffu, e, d), orr = __shfl_up_sync(0xffffffffu, r, d);
      if (lane >= d) { e += oe; r += orr; }
    }
    if (t < nt) tile_pre[t0 + t] = make_uint2(carry_e + e - e0, carry_r + r - r0);
    carry_e += __shfl_sync(0xffffffffu, e, 31);
    carry_r += __shfl_sync(0xffffffffu, r, 31);
  }
}

// U3: scatter unstuffed bytes (word-swapped) and segment starts.
B2_GLOBAL void k_unstuff_scatter(const ImageDev *__restrict__ imgs, const int32_t *__restrict__ tile_base, int nimg,
                                 const uint8_t *__restrict__ d_raw, const uint2 *__restrict__ tile_pre,
                                 uint8_t *__restrict__ d_us, uint32_t *__restrict__ seg_start,
                                 ImageOut *__restrict__ out) {
  B2_SHARED int s_img;
  if (threadIdx.x == 0) s_img = find_owner(tile_base, nimg, (int)blockIdx.x);
  __syncthreads();
  const int i = s_img;
  const ImageDev im = imgs[i];
  const uint32_t pos = ((uint32_t)blockIdx.x - (uint32_t)im.tile_base) * kUTile + threadIdx.x * kUBytesPerThread;
  UClass c;
  u_classify(d_raw + im.raw_off, pos, im.raw_len, c);
  uint32_t tot;
  const uint32_t pre = cta_excl_scan_u32((uint32_t)__popc(c.emit_mask) | ((uint32_t)__popc(c.rst_mask) << 16), &tot);
  const uint2 tp = tile_pre[blockIdx.x];
  uint32_t e = tp.x + (pre & 0xffffu), r = tp.y + (pre >> 16);
  const uint32_t end_pos = out[i].end_pos;
  uint8_t *us = d_us + im.raw_off;
  uint8_t prev = c.prev;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const uint32_t bit = 1u << k;
    if (c.emit_mask & bit) { us[e ^ 3u] = prev == 0xff ? (uint8_t)0xff : c.bytes[k]; ++e; }
    if (c.rst_mask & bit) {
      if (pos + k < end_pos && (int)r + 1 < im.nseg) seg_start[im.seg_base + 1 + (int)r] = e * 8u;
      ++r;
    }
    if ((c.end_mask & bit) && pos + k == end_pos) { out[i].us_len = e; out[i].nrst = r; }
    prev = c.bytes[k];
  }
}

// U4: close the segment tables and raise the stream-level status bits.
B2_GLOBAL void k_unstuff_finish(const ImageDev *__restrict__ imgs, int nimg, uint32_t *__restrict__ seg_start,
                                ImageOut *__restrict__ out) {
  const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x);
  if (i >= nimg) return;
  ImageOut o = out[i];
  uint32_t st = o.status;
  if (o.end_pos == kNone) st |= ST_NO_EOI;
  if (o.err_pos < o.end_pos) st |= ST_BAD_MARKER;
  if (st == 0 && (int)o.nrst != imgs[i].nseg - 1) st |= ST_BAD_RST_COUNT;
  if (st) o.us_len = 0;
  seg_start[imgs[i].seg_base] = 0;
  seg_start[imgs[i].seg_base + imgs[i].nseg] = o.us_len * 8u;
  o.status = st;
  out[i] = o;
}

// ImageOut arrives memset to 0xFF (end_pos = err_pos = kNone); clear the counters and status.
B2_GLOBAL void k_reset_out(ImageOut *__restrict__ out, int nimg) {
  const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x);
  if (i >= nimg) return;
  out[i].us_len = 0; out[i].nrst = 0; out[i].status = 0;
}
B2_GLOBAL void k_raise_status(ImageOut *__restrict__ out, int nimg, uint32_t bits) {
  const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x);
  if (i < nimg) out[i].status |= bits;
}

}  // namespace b2
