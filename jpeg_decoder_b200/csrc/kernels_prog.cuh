// kernels_prog.cuh -- progressive DCT scans (SOF2, ITU-T T.81 Annex G) into the coefficient buffer of the baseline path.
//
// SURVEY.md 8(f) rank 4.  The reference refuses these files (jpeg.cpp:69-73); there is no reference code to replace, only
// the refusal to lift.  A progressive file delivers the coefficients of Image::decodeHuffman (jpeg.cpp:405-446) in several
// scans: DC first / refinement (G.1.2.1), AC first (G.1.2.2, with end-of-band runs) and AC refinement (G.1.2.3, correction
// bits interleaved with newly non-zero coefficients).  After the last scan the buffer holds what the baseline decoders
// leave there -- int16 coef[block][64], blocks in the interleaved MCU order, zig-zag order inside a block, 16-byte chunks
// XOR-swizzled by (block & 7) -- and the pixel stage (kernels_pixel.cuh) runs unchanged.
//
// Shape: ONE THREAD per (image, scan); the scans of an image are launched one after another (a refinement scan reads what
// the scans before it wrote), the images of a batch side by side.  Nothing here is parallel inside a scan: end-of-band runs
// and correction bits make the position of every symbol depend on coefficient VALUES decoded earlier, so the
// self-synchronising construction of kernels_entropy.cuh does not carry over as is.  Correct, not fast: a first version.
// The bit reader walks the STUFFED bytes (FF00 -> FF inline; RSTn at restart boundaries: file.hpp:78-86), so stage U is
// not needed.  Symbols are found by the canonical per-length search (huffman.hpp:17-29 / jpeg.cpp:303-318).
#pragma once
#include "platform.h"
#include "dev_types.h"
#include "kernels_huff.cuh"

namespace b2 {

struct ProgScanDev {
  int32_t ns, comp[3];
  int32_t dc_table[3], ac_table[3];     // index into the batch's table pool (HuffSpec), -1: none
  int32_t ss, se, ah, al;
  int32_t restart_interval;
  uint32_t offset, length;              // the scan's stuffed bytes, relative to the image's first scan byte in d_raw
};
constexpr int kProgMaxScans = 48;       // B200JPEG_PROG_MAX_SCANS
struct ProgImageDev {
  int32_t nscans;
  int32_t width, height;
  ProgScanDev scans[kProgMaxScans];
};

struct ProgBits {
  const uint8_t *p, *end;
  uint32_t buf;         // left-aligned: the next bit is bit 31
  int cnt;              // valid bits in buf
  B2_MEMBER void init(const uint8_t *b, const uint8_t *e) { p = b; end = e; buf = 0; cnt = 0; }
  B2_MEMBER void fill() {
    while (cnt <= 24) {
      uint32_t byte = 0;
      if (p < end) {
        byte = *p;
        if (byte == 0xffu) {
          const uint32_t nx = p + 1 < end ? p[1] : 0xd9u;
          if (nx == 0u) p += 2;                       // stuffed zero (file.hpp:78-82)
          else { byte = 0; }                          // a marker (RSTn or the end of the scan): feed zeros, stay in front of it
        } else ++p;
      }
      buf |= byte << (24 - cnt);
      cnt += 8;
    }
  }
  B2_MEMBER uint32_t peek16() { fill(); return buf >> 16; }
  B2_MEMBER void skip(int n) { buf <<= n; cnt -= n; }
  B2_MEMBER uint32_t bits(int n) {                    // n <= 16
    if (n == 0) return 0;
    fill();
    const uint32_t v = buf >> (32 - n);
    skip(n);
    return v;
  }
  B2_MEMBER uint32_t bit() { return bits(1); }
  B2_MEMBER void restart() {                          // byte-align (BitStream::align, file.hpp:161-164) and step over RSTn
    buf = 0; cnt = 0;
    if (p + 1 < end && p[0] == 0xffu && p[1] >= 0xd0u && p[1] <= 0xd7u) p += 2;
  }
};

// one Huffman symbol: first match by increasing length (jpeg.cpp:303-318); 0x100 | flags when no code matches
B2_DEVICE uint32_t prog_symbol(ProgBits &B, const HuffSpec &spec) {
  const uint32_t w = B.peek16();
  uint32_t code = 0;
  for (int l = 1; l <= 16; ++l) {
    const uint32_t cnt = (uint32_t)spec.offset[l] - (uint32_t)spec.offset[l - 1];
    const uint32_t k = (w >> (16 - l)) - code;
    if (k < cnt) { B.skip(l); return spec.symbols[spec.offset[l - 1] + k]; }
    code = (code + cnt) << 1;
  }
  B.skip(16);
  return 0x100u;
}
B2_DEVICE int prog_extend(uint32_t v, int s) {        // T.81 F.2.2.1, as jpeg.cpp:335-343
  return s != 0 && v < (1u << (s - 1)) ? (int)v - (int)((1u << s) - 1u) : (int)v;
}

// coefficient k (zig-zag index) of block `gb` (global block index in the coefficient buffer)
B2_DEVICE int16_t *prog_coef(int16_t *coef, int64_t gb, int k) {
  return coef + gb * 64 + ((((k >> 3) ^ (int)(gb & 7)) << 3) | (k & 7));
}

// one block of one scan; returns status bits
B2_DEVICE uint32_t prog_block(ProgBits &B, const ProgScanDev &S, int16_t *coef, int64_t gb, int &pred, const HuffSpec *dct,
                              const HuffSpec *act, int &eobrun) {
  if (S.ss == 0) {                                    // DC scans, G.1.2.1
    if (S.ah == 0) {
      const uint32_t s = prog_symbol(B, *dct);
      if (s > 11u) return s & 0x100u ? ST_BAD_CODE : ST_BAD_DC_LEN;
      pred += prog_extend(B.bits((int)s), (int)s);
      const int v = pred * (1 << S.al);
      *prog_coef(coef, gb, 0) = (int16_t)v;
      return (uint32_t)(v + 32768) >> 16 ? ST_DC_RANGE : 0u;
    }
    if (B.bit()) *prog_coef(coef, gb, 0) |= (int16_t)(1 << S.al);
    return 0;
  }
  if (S.ah == 0) {                                    // AC first scan, G.1.2.2
    if (eobrun > 0) { --eobrun; return 0; }
    for (int k = S.ss; k <= S.se; ++k) {
      const uint32_t rs = prog_symbol(B, *act);
      if (rs & 0x100u) return ST_BAD_CODE;
      const int r = (int)(rs >> 4), sz = (int)(rs & 15u);
      if (sz) {
        k += r;
        if (k > 63) return ST_BAD_RUN;
        *prog_coef(coef, gb, k) = (int16_t)(prog_extend(B.bits(sz), sz) * (1 << S.al));
      } else if (r == 15) k += 15;
      else { eobrun = 1 << r; if (r) eobrun += (int)B.bits(r); --eobrun; break; }
    }
    return 0;
  }
  // AC refinement scan, G.1.2.3
  const int p1 = 1 << S.al, m1 = -(1 << S.al);
  int k = S.ss;
  if (eobrun == 0) {
    for (; k <= S.se; ++k) {
      const uint32_t rs = prog_symbol(B, *act);
      if (rs & 0x100u) return ST_BAD_CODE;
      int r = (int)(rs >> 4), sz = (int)(rs & 15u);
      if (sz) sz = B.bit() ? p1 : m1;                 // a newly non-zero coefficient: its sign
      else if (r != 15) { eobrun = 1 << r; if (r) eobrun += (int)B.bits(r); break; }
      do {                                            // skip r still-zero coefficients; correction bits for the others
        int16_t *q = prog_coef(coef, gb, k);
        const int c = *q;
        if (c != 0) { if (B.bit() && (c & p1) == 0) *q = (int16_t)(c + (c >= 0 ? p1 : m1)); }
        else if (--r < 0) break;
        ++k;
      } while (k <= S.se);
      if (sz && k <= 63) *prog_coef(coef, gb, k) = (int16_t)sz;
    }
  }
  if (eobrun > 0) {                                   // rest of the band: correction bits only
    for (; k <= S.se; ++k) {
      int16_t *q = prog_coef(coef, gb, k);
      const int c = *q;
      if (c != 0 && B.bit() && (c & p1) == 0) *q = (int16_t)(c + (c >= 0 ? p1 : m1));
    }
    --eobrun;
  }
  return 0;
}

// scan number `scan_no` of every image that has one; thread t = image t
B2_GLOBAL void k_prog_scan(const ImageDev *__restrict__ imgs, const ProgImageDev *__restrict__ pimgs, int nimg, int scan_no,
                           const HuffSpec *__restrict__ tables, const uint8_t *__restrict__ d_raw, int16_t *__restrict__ coef,
                           ImageOut *__restrict__ out) {
  const int i = (int)(blockIdx.x * blockDim.x + threadIdx.x);
  if (i >= nimg) return;
  const ProgImageDev &pi = pimgs[i];
  if (scan_no >= pi.nscans) return;
  const ImageDev &im = imgs[i];
  const ProgScanDev &S = pi.scans[scan_no];
  ProgBits B;
  const uint8_t *base = d_raw + im.raw_off;
  B.init(base + S.offset, base + S.offset + S.length);
  const int nluma = im.ncomp == 1 ? 1 : im.hs * im.vs;
  int pred[3] = {0, 0, 0}, eobrun = 0;
  uint32_t err = 0;
  int64_t unit = 0;
  if (S.ns > 1) {                                     // interleaved: MCU by MCU, the block order of jpeg.cpp:426-435
    const int64_t mcus = (int64_t)im.mcux * im.mcuy;
    for (int64_t m = 0; m < mcus && !err; ++m, ++unit) {
      if (S.restart_interval && unit && unit % S.restart_interval == 0) { pred[0] = pred[1] = pred[2] = 0; eobrun = 0; B.restart(); }
      for (int j = 0; j < S.ns && !err; ++j) {
        const int c = S.comp[j];
        const int first = c == 0 ? 0 : nluma + c - 1, cnt = c == 0 ? nluma : 1;
        for (int b = 0; b < cnt && !err; ++b)
          err |= prog_block(B, S, coef, im.coef_base + m * im.bpm + first + b, pred[c],
                            S.dc_table[j] >= 0 ? &tables[S.dc_table[j]] : nullptr, S.ac_table[j] >= 0 ? &tables[S.ac_table[j]] : nullptr, eobrun);
      }
    }
  } else {                                            // one component: its blocks in raster order over ceil(component size / 8) (A.2.3)
    const int c = S.comp[0];
    const bool lum = c == 0;
    const int cw = lum ? im.width : (im.width + im.hs - 1) / im.hs, ch = lum ? im.height : (im.height + im.vs - 1) / im.vs;
    const int nbx = (cw + 7) / 8, nby = (ch + 7) / 8;
    const HuffSpec *dct = S.dc_table[0] >= 0 ? &tables[S.dc_table[0]] : nullptr, *act = S.ac_table[0] >= 0 ? &tables[S.ac_table[0]] : nullptr;
    for (int by = 0; by < nby && !err; ++by)
      for (int bx = 0; bx < nbx && !err; ++bx, ++unit) {
        if (S.restart_interval && unit && unit % S.restart_interval == 0) { pred[0] = pred[1] = pred[2] = 0; eobrun = 0; B.restart(); }
        int64_t blk;
        if (lum && im.ncomp == 3) blk = ((int64_t)(by / im.vs) * im.mcux + bx / im.hs) * im.bpm + (by % im.vs) * im.hs + bx % im.hs;
        else if (lum) blk = (int64_t)by * im.mcux + bx;
        else blk = ((int64_t)by * im.mcux + bx) * im.bpm + nluma + c - 1;
        err |= prog_block(B, S, coef, im.coef_base + blk, pred[c], dct, act, eobrun);
      }
  }
  if (err) atomicOr(&out[i].status, err);
}

}  // namespace b2
