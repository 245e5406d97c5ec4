// kernels_segdec.cuh -- stage E, segment-parallel path: one THREAD per restart segment (or per
// whole scan when the file has no RSTn), decoding straight from the raw stuffed bytes into the
// final coefficient blocks in ONE pass.  Replaces, for batches that offer thousands of
// independently decodable segments, JPEGFile::readImageData (file.hpp:59-104: the FF00 -> FF
// unstuffing happens in the bit reader's refill), BitStream (file.hpp:122-165) and
// Image::decodeHuffman / decodeMCUComponent / getNextSymbol (jpeg.cpp:300-446).
//
// Why this shape: a segment starts byte-aligned with the DC predictors reset (jpeg.cpp:419-425)
// and its first block index is known a priori (k * RI * blocks-per-MCU), so nothing has to be
// synchronised or prefix-summed -- the self-synchronising path (kernels_entropy.cuh) pays three
// passes over the stream for that.  The price is one long dependent chain per thread; with
// >= ~1200 segments in flight the chain latency (one shared-memory LUT access per symbol) is what
// bounds the kernel, not the SM count, which is why the planner only takes this path when the
// batch has enough segments (pipeline.h, choose_path).
//
//   M1 k_mark_count / k_unstuff_scan / M2 k_mark_scatter / M3 k_mark_finish: find FF D0..D7 / FF D9 /
//      illegal markers in the raw scan (vectorised 128-bit loads, FF bytes located with SWAR tests)
//      and write the byte offset where every segment starts.
//   D  k_entropy_seg: per thread a 128-bit shifting bit buffer fed 64 raw bits at a time from a
//      register queue of 128-bit loads (one vector prefetched ahead); DC and AC symbols share one
//      flat loop, both tables looked up at once so that the table select is off the critical path;
//      each finished block leaves as one 128-byte line (chunks XOR-swizzled, see kernels_pixel.cuh).
#pragma once
#include "platform.h"
#include "dev_types.h"
#include "kernels_unstuff.cuh"
#include "kernels_huff.cuh"

namespace b2 {

constexpr int kSegThreads = 64;          // segments (threads) per CTA of k_entropy_seg
constexpr uint32_t kSegNoUnit = 0xffffffffu;

// ---------------------------------------------------------------------------------- marker pass
// bit 7/15/23/31 set where the byte of w is 0xFF (exact, no carries between bytes)
B2_DEVICE uint32_t ff_bytes(uint32_t w) { return ((w & 0x7f7f7f7fu) + 0x01010101u) & w & 0x80808080u; }

struct MarkHits { uint32_t rst, end, err; };   // bit k: the FF at byte k of this thread starts such a marker

// The thread owns the FF bytes among its 16 bytes at [pos, pos+16); the byte AFTER an FF decides:
// 00 stuffing | FF fill (file.hpp:87-90) | D0..D7 restart | D9 end | anything else illegal (file.hpp:91-95)
B2_DEVICE MarkHits mark_classify(const uint8_t *__restrict__ raw, uint32_t pos, uint32_t raw_len) {
  MarkHits h; h.rst = h.end = h.err = 0;
  if (pos >= raw_len) return h;
  const uint4 v = b2_ld_nc_v4(reinterpret_cast<const uint4 *>(raw + pos));
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    uint32_t f = ff_bytes(w[j]);
    while (f) {
      const int bit = __ffs((int)f) - 1;      // 7, 15, 23 or 31
      f &= f - 1;
      const int k = j * 4 + (bit >> 3);
      if (pos + k + 1 >= raw_len) continue;
      const uint32_t nx = k < 15 ? ((w[(k + 1) >> 2] >> (8 * ((k + 1) & 3))) & 0xffu) : (uint32_t)raw[pos + 16];
      if (nx == 0x00u || nx == 0xffu) continue;
      if ((nx & 0xf8u) == 0xd0u) h.rst |= 1u << k;
      else if (nx == 0xd9u) h.end |= 1u << k;
      else h.err |= 1u << k;
    }
  }
  return h;
}

// M1: RST count per tile (high half of tile_cnt, the layout k_unstuff_scan sums) + first EOI / first illegal marker.
B2_GLOBAL void k_mark_count(const ImageDev *__restrict__ imgs, const int32_t *__restrict__ tile_base, int nimg,
                            const uint8_t *__restrict__ d_raw, uint32_t *__restrict__ tile_cnt,
                            ImageOut *__restrict__ out) {
  B2_SHARED int s_img;
  if (threadIdx.x == 0) s_img = find_owner(tile_base, nimg, (int)blockIdx.x);
  __syncthreads();
  const int i = s_img;
  const ImageDev im = imgs[i];
  if (im.path != kPathSegment) return;
  const uint32_t pos = ((uint32_t)blockIdx.x - (uint32_t)im.tile_base) * kUTile + threadIdx.x * kUBytesPerThread;
  const MarkHits h = mark_classify(d_raw + im.raw_off, pos, im.raw_len);
  if (h.end) atomicMin(&out[i].end_pos, pos + (uint32_t)__ffs((int)h.end));          // position of the D9
  if (h.err) atomicMin(&out[i].err_pos, pos + (uint32_t)__ffs((int)h.err));
  uint32_t tot;
  cta_excl_scan_u32((uint32_t)__popc(h.rst) << 16, &tot);
  if (threadIdx.x == 0) tile_cnt[blockIdx.x] = tot;
}

// M2: the r-th RSTn before the EOI starts segment r+1 at the byte after the marker.
B2_GLOBAL void k_mark_scatter(const ImageDev *__restrict__ imgs, const int32_t *__restrict__ tile_base, int nimg,
                              const uint8_t *__restrict__ d_raw, const uint2 *__restrict__ tile_pre,
                              uint32_t *__restrict__ seg_pos, ImageOut *__restrict__ out) {
  B2_SHARED int s_img;
  if (threadIdx.x == 0) s_img = find_owner(tile_base, nimg, (int)blockIdx.x);
  __syncthreads();
  const int i = s_img;
  const ImageDev im = imgs[i];
  if (im.path != kPathSegment) return;
  const uint32_t pos = ((uint32_t)blockIdx.x - (uint32_t)im.tile_base) * kUTile + threadIdx.x * kUBytesPerThread;
  const MarkHits h = mark_classify(d_raw + im.raw_off, pos, im.raw_len);
  uint32_t tot;
  const uint32_t pre = cta_excl_scan_u32((uint32_t)__popc(h.rst) << 16, &tot);
  uint32_t r = tile_pre[blockIdx.x].y + (pre >> 16);
  const uint32_t end_pos = out[i].end_pos;
  uint32_t m = h.rst;
  while (m) {
    const int k = __ffs((int)m) - 1;
    m &= m - 1;
    if (pos + k + 1 < end_pos) {
      if ((int)r + 1 < im.nseg) seg_pos[im.seg_base + 1 + (int)r] = pos + (uint32_t)k + 2u;
      ++r;
    }
  }
  if (h.end) {
    const uint32_t k = (uint32_t)__ffs((int)h.end) - 1u;
    if (pos + k + 1 == end_pos) out[i].nrst = r;       // RSTn markers before the EOI (r counts those of this thread before k too:
  }                                                     // a marker after the EOI's FF in the same thread fails the pos test above)
}

// M3: close the segment tables and raise the stream-level status bits.
B2_GLOBAL void k_mark_finish(const ImageDev *__restrict__ imgs, int nimg, uint32_t *__restrict__ seg_pos,
                             ImageOut *__restrict__ out) {
  const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x);
  if (i >= nimg || imgs[i].path != kPathSegment) return;
  ImageOut o = out[i];
  uint32_t st = o.status;
  if (o.end_pos == kNone) st |= ST_NO_EOI;
  if (o.err_pos < o.end_pos) st |= ST_BAD_MARKER;
  if (st == 0 && (int)o.nrst != imgs[i].nseg - 1) st |= ST_BAD_RST_COUNT;
  seg_pos[imgs[i].seg_base] = 0;
  seg_pos[imgs[i].seg_base + imgs[i].nseg] = st ? 0u : o.end_pos - 1u;   // the FF of the EOI
  o.us_len = 0;
  o.status = st;
  out[i] = o;
}

// ---------------------------------------------------------------------------------- bit reader
// MSB-first reader over the RAW (stuffed) bytes of one segment.  A 128-bit buffer (b3 = next 32 bits)
// shifts left as bits are consumed; whenever <= 64 valid bits remain, the next 8 raw bytes are
// appended.  Eight bytes without an FF (97 % of them) go in as they are; otherwise they are walked
// byte by byte: FF 00 -> FF, FF FF.. fill skipped, any other FF xx is a marker and ends the
// segment (zeros are fed from then on and counted, so that an overrun can be reported).
struct SegReader {
  const uint4 *ptr;        // next vector to load
  uint4 q, qn;             // current vector (halves not yet taken) and the prefetched next one
  int nq;                  // 8-byte halves left in q
  uint32_t b3, b2, b1, b0;
  int avail;               // valid bits in b3..b0
  uint32_t after_ff, eos;
  int fill;                // zero bits appended after the end of the segment

  B2_MEMBER void place(uint32_t vhi, uint32_t vlo, int nbits) {      // append nbits (left-aligned in vhi:vlo) at `avail`
    const uint32_t s = (uint32_t)avail & 31u;
    const int k = avail >> 5;
    const uint32_t y2 = vhi >> s, y1 = __funnelshift_r(vlo, vhi, s), y0 = __funnelshift_r(0u, vlo, s);
    if (k == 0) { b3 |= y2; b2 |= y1; b1 |= y0; }
    else if (k == 1) { b2 |= y2; b1 |= y1; b0 |= y0; }
    else { b1 |= y2; b0 |= y1; }
    avail += nbits;
  }
  B2_MEMBER void pop(uint32_t &w0, uint32_t &w1) {
    if (nq == 0) { q = qn; qn = b2_ld_nc_v4(ptr); ++ptr; nq = 2; }
    w0 = q.x; w1 = q.y; q.x = q.z; q.y = q.w; --nq;
  }
  // bytes [first,8) of the group (w0 = bytes 0-3, w1 = bytes 4-7, memory order)
  B2_MEMBER void append_slow(uint32_t w0, uint32_t w1, int first) {
    uint32_t vhi = 0, vlo = 0;
    int nb = 0;
    for (int j = first; j < 8 && !eos; ++j) {
      const uint32_t b = ((j < 4 ? w0 : w1) >> (8 * (j & 3))) & 0xffu;
      if (after_ff) {
        if (b == 0xffu) continue;                    // fill byte
        after_ff = 0;
        if (b == 0x00u) continue;                    // stuffing: the FF before it was data
        eos = 1;                                     // RSTn / EOI / anything else: the segment ends here
        break;
      }
      if (nb < 4) vhi |= b << (24 - 8 * nb); else vlo |= b << (56 - 8 * nb);
      ++nb;
      if (b == 0xffu) after_ff = 1;
    }
    place(vhi, vlo, nb * 8);
  }
  B2_MEMBER void refill() {                          // call when avail <= 64
    if (eos) { avail += 64; fill += 64; return; }
    uint32_t w0, w1;
    pop(w0, w1);
    if (!after_ff && (ff_bytes(w0) | ff_bytes(w1)) == 0u)
      place(__byte_perm(w0, 0u, 0x0123), __byte_perm(w1, 0u, 0x0123), 64);
    else
      append_slow(w0, w1, 0);
  }
  B2_MEMBER void init(const uint8_t *raw, uint32_t start) {
    ptr = reinterpret_cast<const uint4 *>(raw + (start & ~15u));
    q = b2_ld_nc_v4(ptr); qn = b2_ld_nc_v4(ptr + 1); ptr += 2; nq = 2;
    b3 = b2 = b1 = b0 = 0; avail = 0; after_ff = eos = 0; fill = 0;
    uint32_t w0, w1;
    if (start & 8u) pop(w0, w1);
    pop(w0, w1);
    append_slow(w0, w1, (int)(start & 7u));
    while (avail <= 64) refill();
  }
  B2_MEMBER void consume(uint32_t n) {               // n < 32
    b3 = __funnelshift_l(b2, b3, n); b2 = __funnelshift_l(b1, b2, n); b1 = __funnelshift_l(b0, b1, n); b0 <<= (n & 31u);
    avail -= (int)n;
  }
};

B2_DEVICE int seg_extend(uint32_t w, uint32_t len, uint32_t size) {    // jpeg.cpp:335-343, 388-397; size >= 1
  const uint32_t v = (w << len) >> (32u - size);
  return (v >> (size - 1u)) ? (int)v : (int)v - (int)((1u << size) - 1u);
}

// ---------------------------------------------------------------------------------- decoder
B2_HD size_t segdec_smem_bytes() { return sizeof(uint32_t) * (size_t)kHTabsPerSet * kHTab + (size_t)kSegThreads * 128; }

// units[u] = (image, segment); the units of one CTA share a table set (the planner pads each
// table-set group to a multiple of kSegThreads with kSegNoUnit).
B2_GLOBAL void __launch_bounds__(kSegThreads)
k_entropy_seg(const uint2 *__restrict__ units, const ImageDev *__restrict__ imgs, const TableSet *__restrict__ tsets,
              const uint32_t *__restrict__ luts, const uint32_t *__restrict__ seg_pos,
              const uint8_t *__restrict__ d_raw, ImageOut *__restrict__ out, int16_t *__restrict__ coef) {
  B2_DYN_SMEM(uint32_t, lut);                                   // [8][kHTab]
  uint4 *stage4 = reinterpret_cast<uint4 *>(lut + kHTabsPerSet * kHTab);   // chunk q of thread t at stage4[q*kSegThreads + t]
  const int tid = threadIdx.x;
  const uint2 unit = units[(size_t)blockIdx.x * kSegThreads + tid];
  {
    // the CTA's table set: that of its first unit (never padding)
    const int set = imgs[units[(size_t)blockIdx.x * kSegThreads].x].tset;
    const uint4 *src = reinterpret_cast<const uint4 *>(luts + (size_t)set * kHTabsPerSet * kHTab);
    uint4 *dst = reinterpret_cast<uint4 *>(lut);
    for (int t = 0; t < kHTabsPerSet; ++t) {
      const bool used = t < 4 ? tsets[set].dc[t].set : tsets[set].ac[t - 4].set;
      if (!used) continue;
      for (int k = tid; k < kHTab / 4; k += kSegThreads) dst[t * (kHTab / 4) + k] = b2_ld_nc_v4(src + t * (kHTab / 4) + k);
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) stage4[q * kSegThreads + tid] = make_uint4(0, 0, 0, 0);
  }
  __syncthreads();
  if (unit.x == kSegNoUnit) return;
  const int img = (int)unit.x, k = (int)unit.y;
  const ImageDev &im = imgs[img];
  if (out[img].status & (ST_BAD_MARKER | ST_NO_EOI | ST_BAD_RST_COUNT)) return;     // segment table is not valid

  const int bpm = im.bpm, nluma = im.ncomp == 1 ? 1 : im.hs * im.vs;
  const int seg_blocks = im.restart_interval * bpm;                               // 0 when there is one segment
  int bidx = k * seg_blocks;
  const int bend = (k + 1 < im.nseg) ? bidx + seg_blocks : im.nblocks;
  if (bidx >= bend) return;
  // table bases (in u32 entries) per component: DC table td -> slot td, AC table ta -> slot 4+ta
  const uint32_t dcb0 = (uint32_t)(im.td[0] & 3) * kHTab, dcb1 = (uint32_t)(im.td[1] & 3) * kHTab, dcb2 = (uint32_t)(im.td[2] & 3) * kHTab;
  const uint32_t acb0 = (uint32_t)(4 + (im.ta[0] & 3)) * kHTab, acb1 = (uint32_t)(4 + (im.ta[1] & 3)) * kHTab,
                 acb2 = (uint32_t)(4 + (im.ta[2] & 3)) * kHTab;
  const TableSet &ts = tsets[im.tset];
  int16_t *obase = coef + im.coef_base * 64;
  uint16_t *stage = reinterpret_cast<uint16_t *>(stage4 + tid);                   // element i at stage[(i>>3)*kSegThreads*8 + (i&7)]

  SegReader rd;
  rd.init(d_raw + im.raw_off, seg_pos[im.seg_base + k]);

  int pred0 = 0, pred1 = 0, pred2 = 0;
  uint32_t err = 0;
  int blk = 0, comp = 0;                 // block inside the MCU, its component
  uint32_t dcb = dcb0, acb = acb0;
  uint32_t z = 0;
  for (;;) {
    if (rd.avail <= 64) rd.refill();
    const uint32_t w = rd.b3;
    const uint32_t idx = w >> (32 - kHLutBits);
    const uint32_t e_ac = lut[acb + idx], e_dc = lut[dcb + idx];                  // both at once: the select is off the critical path
    const bool is_dc = z == 0;
    uint32_t e = is_dc ? e_dc : e_ac;
    if (e & kHSpecial) {                                                          // rare: long code, or an error entry
      const uint32_t base = is_dc ? dcb : acb;
      if (e & kHLong) e = lut[base + (e & 0xffffu) + ((w >> (32 - kHLutBits - 6)) & 63u)];
      if (e & kHOverflow) e = huff_slow(is_dc ? ts.dc[im.td[comp] & 3] : ts.ac[im.ta[comp] & 3], w >> 16, is_dc);
      if (e & kHBadCode) err |= ST_BAD_CODE;
      if (e & kHBadSize) err |= is_dc ? ST_BAD_DC_LEN : ST_BAD_AC_LEN;
    }
    const uint32_t adv = e & 31u, len = (e >> 5) & 31u, zinc = (e >> 10) & 127u;
    const uint32_t size = adv - len;
    int val = size ? seg_extend(w, len, size) : 0;
    rd.consume(adv);
    const uint32_t zn = z + zinc;
    if (is_dc) {                                                                  // jpeg.cpp:335-345
      int p = comp == 0 ? pred0 : (comp == 1 ? pred1 : pred2);
      p += val;
      if (comp == 0) pred0 = p; else if (comp == 1) pred1 = p; else pred2 = p;
      if (p < -32768 || p > 32767) err |= ST_DC_RANGE;
      stage[0] = (uint16_t)p;
    } else if (size) {
      const uint32_t pos = zn - 1u;                                               // z + run
      if (pos < 64u) stage[(pos >> 3) * (kSegThreads * 8) + (pos & 7u)] = (uint16_t)val;
    }
    z = zn;
    if (z < 64u) continue;
    // ---- block end: EOB, or the index ran to 64.  jpeg.cpp:372-376: a run that reaches 64 is an error ----
    if (!(e & kHEob) && !(size != 0u && z == 64u)) err |= ST_BAD_RUN;
    {
      uint4 *o = reinterpret_cast<uint4 *>(obase + (int64_t)bidx * 64);
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        o[q ^ (bidx & 7)] = stage4[q * kSegThreads + tid];                        // chunk swizzle, see kernels_pixel.cuh
        stage4[q * kSegThreads + tid] = make_uint4(0, 0, 0, 0);
      }
    }
    if (++bidx == bend) break;
    blk = blk + 1 == bpm ? 0 : blk + 1;
    comp = blk < nluma ? 0 : blk - nluma + 1;
    dcb = comp == 0 ? dcb0 : (comp == 1 ? dcb1 : dcb2);
    acb = comp == 0 ? acb0 : (comp == 1 ? acb1 : acb2);
    z = 0;
  }
  if (rd.fill > rd.avail) err |= ST_OVERRUN;                                      // bits were taken from beyond the segment's end
  if (err) atomicOr(&out[img].status, err);
}

}  // namespace b2
