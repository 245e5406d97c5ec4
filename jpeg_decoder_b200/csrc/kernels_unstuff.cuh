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
// (Tried and dropped, profiles/r02_experiments.md: count and scatter as ONE pass with decoupled look-back between the
// tiles of an image.  Correct, but at 4 KiB per tile the look-back's trips through L2 cost more than the second
// classification saves: 0.43 ms instead of 0.34 ms for C2.)
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
  uint32_t ow[4];       // the 16 bytes as they are emitted (a stuffed 00 after FF already turned into FF)
};

// 0xff in every byte lane where a and b agree
B2_DEVICE uint32_t u_eq4(uint32_t a, uint32_t b) {
#ifdef B200JPEG_EMU
  uint32_t r = 0;
  for (int k = 0; k < 4; ++k) if (((a >> (8 * k)) & 0xffu) == ((b >> (8 * k)) & 0xffu)) r |= 0xffu << (8 * k);
  return r;
#else
  return __vcmpeq4(a, b);
#endif
}
B2_DEVICE uint32_t u_nibble(uint32_t m) { return (((m & 0x01010101u) * 0x01020408u) >> 24) & 0xfu; }   // byte lanes -> 4 bits

// Four bytes at a time with byte-lane compares; called by every thread of the CTA (warp shuffle inside).
B2_DEVICE void u_classify(const uint8_t *__restrict__ raw, uint32_t pos, uint32_t raw_len, UClass &c) {
  c.emit_mask = c.rst_mask = c.end_mask = c.err_mask = 0;
  const bool live = pos < raw_len;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (live) v = b2_ld_nc_v4(reinterpret_cast<const uint4 *>(raw + pos));
  // the byte before this thread's sixteen: the last byte of the previous lane, or one load for lane 0
  uint32_t prev = __shfl_up_sync(0xffffffffu, v.w >> 24, 1);
  if ((threadIdx.x & 31) == 0) prev = (live && pos) ? raw[pos - 1] : 0;
  c.ow[0] = v.x; c.ow[1] = v.y; c.ow[2] = v.z; c.ow[3] = v.w;
  if (!live) return;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t w = c.ow[j];
    const uint32_t pw = (w << 8) | prev;
    prev = w >> 24;
    const uint32_t fc = u_eq4(w, 0xffffffffu), fp = u_eq4(pw, 0xffffffffu), zc = u_eq4(w, 0u);
    const uint32_t rs = u_eq4(w & 0xf8f8f8f8u, 0xd0d0d0d0u), d9 = u_eq4(w, 0xd9d9d9d9u);
    c.emit_mask |= u_nibble((~fp & ~fc) | (fp & zc)) << (4 * j);
    c.rst_mask |= u_nibble(fp & rs) << (4 * j);
    c.end_mask |= u_nibble(fp & d9) << (4 * j);
    c.err_mask |= u_nibble(fp & ~(zc | rs | d9 | fc)) << (4 * j);
    c.ow[j] = w | fp;                       // FF 00 -> the 00 is what gets emitted, as FF
  }
  const uint32_t left = raw_len - pos;
  if (left < 16u) {
    const uint32_t m = (1u << left) - 1u;
    c.emit_mask &= m; c.rst_mask &= m; c.end_mask &= m; c.err_mask &= m;
  }
}

// The counting pass needs only HOW MANY bytes a thread emits and how many RSTn it holds, not which: byte-lane masks and
// two population counts per word instead of the bit masks of u_classify.  Returns false when the sixteen bytes hold
// anything else of interest (FF D9, an illegal marker) or are the ragged end of the scan: the caller classifies in full.
B2_DEVICE bool u_count_fast(const uint8_t *__restrict__ raw, uint32_t pos, uint32_t raw_len, uint32_t &emits, uint32_t &rsts) {
  const bool live = pos < raw_len;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (live) v = b2_ld_nc_v4(reinterpret_cast<const uint4 *>(raw + pos));
  uint32_t prev = __shfl_up_sync(0xffffffffu, v.w >> 24, 1);
  if ((threadIdx.x & 31) == 0) prev = (live && pos) ? raw[pos - 1] : 0;
  emits = rsts = 0;
  if (!live) return true;
  const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
  uint32_t held = 0, marks = 0, other = 0;           // byte lanes: not emitted / RSTn / anything else behind an FF
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t w = w4[j], pw = (w << 8) | prev;
    prev = w >> 24;
    const uint32_t fc = u_eq4(w, 0xffffffffu), fp = u_eq4(pw, 0xffffffffu), zc = u_eq4(w, 0u);
    const uint32_t rs = u_eq4(w & 0xf8f8f8f8u, 0xd0d0d0d0u);
    held += (uint32_t)__popc(fc | (fp & ~zc));
    marks += (uint32_t)__popc(fp & rs);
    other |= fp & ~(zc | rs | fc);
  }
  emits = 16u - (held >> 3);
  rsts = marks >> 3;
  return other == 0u && raw_len - pos >= 16u;
}

// A thread's emitted bytes into the tile's staging buffer (byte stores, already word-swapped); q = where they start.
B2_DEVICE void u_stage(const UClass &c, uint32_t q, uint8_t *sbuf, uint32_t *sbuf32) {
  if (c.emit_mask == 0xffffu) {
    // all sixteen bytes pass (no FF near: 94 % of the threads): whole big-endian words, ragged ends as bytes
    const uint32_t sh = q & 3u, wb = q >> 2;
    uint32_t B[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) B[j] = __byte_perm(c.ow[j], 0, 0x0123);
    if (sh == 0) {
#pragma unroll
      for (int j = 0; j < 4; ++j) sbuf32[wb + j] = B[j];
    } else {
#pragma unroll
      for (int j = 1; j < 4; ++j) sbuf32[wb + j] = __funnelshift_r(B[j], B[j - 1], 8u * sh);
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        if ((uint32_t)k < 4u - sh) sbuf[(q + k) ^ 3u] = (uint8_t)(c.ow[0] >> (8 * k));
        if ((uint32_t)k < sh) sbuf[(q + 16u - sh + k) ^ 3u] = (uint8_t)(c.ow[3] >> (8u * (4u - sh + k)));
      }
    }
  } else {
    uint32_t ee = q;
#pragma unroll
    for (int k = 0; k < 16; ++k)
      if (c.emit_mask & (1u << k)) { sbuf[ee ^ 3u] = (uint8_t)(c.ow[k >> 2] >> (8 * (k & 3))); ++ee; }
  }
}
// The tile's output is the contiguous byte range [lo, hi) of the unstuffed scan, staged from the 16-byte aligned
// position a: whole 16-byte chunks leave as 128-bit stores, the ragged first and last bytes (which neighbouring tiles
// share) byte by byte.
B2_DEVICE void u_drain(uint8_t *us, const uint8_t *sbuf, uint32_t a, uint32_t lo, uint32_t hi) {
  const uint32_t c0 = (lo + 15u) & ~15u, c1 = hi & ~15u;           // whole 16-byte chunks inside it
  for (uint32_t g = c0 + threadIdx.x * 16u; g < c1; g += kUThreads * 16u)
    *reinterpret_cast<uint4 *>(us + g) = *reinterpret_cast<const uint4 *>(sbuf + (g - a));
  if (threadIdx.x < 32) {                                          // ragged edges: at most 15 bytes each
    const uint32_t head_end = c0 < hi ? c0 : hi;
    const uint32_t g0 = lo + threadIdx.x;
    if (g0 < head_end) us[g0 ^ 3u] = sbuf[(g0 - a) ^ 3u];
    const uint32_t tail0 = c1 > head_end ? c1 : head_end;
    const uint32_t g1 = tail0 + threadIdx.x;
    if (g1 < hi) us[g1 ^ 3u] = sbuf[(g1 - a) ^ 3u];
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
  uint32_t emits, rsts;
  const bool plain = u_count_fast(d_raw + im.raw_off, pos, im.raw_len, emits, rsts);
  if (__any_sync(0xffffffffu, !plain)) {             // rare: an end of scan, an illegal marker, the scan's last bytes
    UClass c;
    u_classify(d_raw + im.raw_off, pos, im.raw_len, c);
    if (c.end_mask) atomicMin(&out[i].end_pos, pos + (uint32_t)__ffs((int)c.end_mask) - 1);
    if (c.err_mask) atomicMin(&out[i].err_pos, pos + (uint32_t)__ffs((int)c.err_mask) - 1);
    emits = (uint32_t)__popc(c.emit_mask); rsts = (uint32_t)__popc(c.rst_mask);
  }
  uint32_t tot;
  cta_excl_scan_u32(emits | (rsts << 16), &tot);
  if (threadIdx.x == 0) tile_cnt[blockIdx.x] = tot;   // emits in the low half, RSTs in the high half
}

// U2: one warp per image turns tile counts into exclusive prefixes (emits and RSTs separately).  A lane takes kUScanPer
// consecutive tiles per pass -- their loads are independent, the warp-wide scan runs over the lanes' sums -- so a scan of
// several thousand tiles (one 8192 x 8192 image: 7,000) is some thirty dependent passes, not two hundred.
constexpr int kUScanPer = 8;
B2_GLOBAL void k_unstuff_scan(const ImageDev *__restrict__ imgs, int nimg, const uint32_t *__restrict__ tile_cnt,
                              uint2 *__restrict__ tile_pre) {
  const int i = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (i >= nimg) return;
  const int t0 = imgs[i].tile_base, nt = imgs[i].ntiles;
  uint32_t carry_e = 0, carry_r = 0;
  for (int b = 0; b < nt; b += 32 * kUScanPer) {
    const int t = b + lane * kUScanPer;
    uint32_t c[kUScanPer];
#pragma unroll
    for (int k = 0; k < kUScanPer; ++k) c[k] = t + k < nt ? tile_cnt[t0 + t + k] : 0u;
    uint32_t e = 0, r = 0;                                   // this lane's sums
#pragma unroll
    for (int k = 0; k < kUScanPer; ++k) { e += c[k] & 0xffffu; r += c[k] >> 16; }
    const uint32_t e0 = e, r0 = r;
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t oe = __shfl_up_sync(0xffffffffu, e, d), orr = __shfl_up_sync(0xffffffffu, r, d);
      if (lane >= d) { e += oe; r += orr; }
    }
    uint32_t pe = carry_e + e - e0, pr = carry_r + r - r0;   // exclusive prefixes of the lane's first tile
#pragma unroll
    for (int k = 0; k < kUScanPer; ++k) {
      if (t + k < nt) tile_pre[t0 + t + k] = make_uint2(pe, pr);
      pe += c[k] & 0xffffu; pr += c[k] >> 16;
    }
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
  // The tile's output is the contiguous byte range [tp.x, tp.x + emits) of the unstuffed scan.  It is put
  // together in shared memory (byte stores, already word-swapped) and leaves as 128-bit stores; only the
  // ragged first and last 16 bytes, which neighbouring tiles share, go out byte by byte.
  B2_SHARED uint4 sbuf4[kUTile / 16 + 1];
  uint8_t *sbuf = reinterpret_cast<uint8_t *>(sbuf4);
  const uint32_t a = tp.x & ~15u;                 // 16-byte aligned start of the range in the scan
  uint32_t *sbuf32 = reinterpret_cast<uint32_t *>(sbuf4);
  const uint32_t q = e - a;                       // where this thread's bytes start in the staging buffer
  u_stage(c, q, sbuf, sbuf32);
  for (uint32_t m = c.rst_mask; m; m &= m - 1) {  // rare: a restart marker ends at byte k
    const int k = __ffs((int)m) - 1;
    const uint32_t before = (uint32_t)__popc(c.rst_mask & ((1u << k) - 1u));
    const uint32_t ek = e + (uint32_t)__popc(c.emit_mask & ((1u << k) - 1u));
    if (pos + k < end_pos && (int)(r + before) + 1 < im.nseg) seg_start[im.seg_base + 1 + (int)(r + before)] = ek * 8u;
  }
  if (c.end_mask) {                                // rare: FF D9; the first one of the image closes the scan
    for (uint32_t m = c.end_mask; m; m &= m - 1) {
      const int k = __ffs((int)m) - 1;
      if (pos + k == end_pos) {
        out[i].us_len = e + (uint32_t)__popc(c.emit_mask & ((1u << k) - 1u));
        out[i].nrst = r + (uint32_t)__popc(c.rst_mask & ((1u << k) - 1u));
      }
    }
  }
  __syncthreads();
  u_drain(d_us + im.raw_off, sbuf, a, tp.x, tp.x + (tot & 0xffffu));
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
