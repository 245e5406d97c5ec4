// kernels_huff.cuh -- device Huffman lookup tables, built ONCE per table set and batch.
//
// The reference decodes a symbol by appending one bit at a time and scanning the codes of that
// length (Image::getNextSymbol, jpeg.cpp:300-320) over the canonical codes that
// HuffmanTable::generateCodes assigns (huffman.hpp:17-29).  Here every DHT table becomes a
// two-level lookup table that returns, for the next 16 bits of the stream, everything
// decodeMCUComponent (jpeg.cpp:322-403) derives from the symbol:
//
//   level 1: kHLutBits (10) bits -> entry; codes longer than that carry kHLong and the base of a
//   level 2: 64-entry block indexed by the following 6 bits.  Canonical codes grow with their
//            length, so the long codes occupy the 10-bit prefixes from P0 = first code of length 11
//            (>> 1) upwards: block = prefix - P0.  The Annex-K tables need 5 blocks; kHL2Blocks (8)
//            are provided, a table that needs more sets kHOverflow for the prefixes beyond them and
//            those (pathological) codes are resolved by the per-length search huff_slow().
//
// entry (u32):  [4:0] adv = code length + magnitude bits (<= 27)   [8:5] size = magnitude bits (code length = adv - size)
//               [15:9] zinc = how far the zig-zag index moves (run+1, run for size 0, 64 for EOB,
//               1 for a DC symbol); bit 15 (zinc = 64) doubles as the EOB flag
//               [19:16] adv2, [24:20] zinc2, [27:25] size2: level-1 entries of AC tables only -- when the symbol is followed,
//               inside the same kHLutBits index bits, by a SECOND complete AC symbol (code + magnitude bits),
//               adv2 != 0 is that symbol's length, zinc2 its zig-zag advance (0 = it is EOB) and size2 its magnitude bits.
//               Two thirds of the symbols of a typical scan pair up, so a decoder that takes both at once makes a third
//               fewer steps, and whoever places the coefficients has everything about both symbols in the one entry.
//               bit 28 overflow   bit 29 bad size (jpeg.cpp:330,381)   bit 30 no such code (jpeg.cpp:319)
//               bit 31 long (level-1 only)
//               A LEVEL-1 entry with any of the bits 28..31 has NOTHING in its low 20 bits: a decoder that adds such an
//               entry's adv / zinc / adv2 to its state stays where it is, without looking at the flags -- that test was
//               two links in the segment decoder's dependent chain.  What the entry has to say is in [24:20]: the
//               level-2 block for a long code (huff_l2_base), the code length for an error entry (huff_l1_error gives
//               the standard form: advance by the code, end the block, the flag).
#pragma once
#include "platform.h"
#include "dev_types.h"

namespace b2 {

constexpr int kHLutBits = 10;
constexpr int kHL1 = 1 << kHLutBits;
constexpr int kHL2Blocks = 8;
constexpr int kHL2 = kHL2Blocks * 64;
constexpr int kHTab = kHL1 + kHL2;          // u32 entries per table (6 KiB)
constexpr int kHTabsPerSet = 8;             // dc0..dc3, ac0..ac3
constexpr uint32_t kHEob = 1u << 15, kHOverflow = 1u << 28, kHBadSize = 1u << 29, kHBadCode = 1u << 30, kHLong = 1u << 31;
constexpr uint32_t kHSpecial = kHOverflow | kHBadSize | kHBadCode | kHLong;

B2_HD uint32_t huff_entry(uint32_t adv, uint32_t len, uint32_t zinc, uint32_t flags) {
  return adv | ((adv - len) << 5) | (zinc << 9) | flags;
}
B2_HD uint32_t huff_adv(uint32_t e) { return e & 31u; }
B2_HD uint32_t huff_size(uint32_t e) { return (e >> 5) & 15u; }
B2_HD uint32_t huff_len(uint32_t e) { return huff_adv(e) - huff_size(e); }
B2_HD uint32_t huff_zinc(uint32_t e) { return (e >> 9) & 127u; }
B2_HD uint32_t huff_adv2(uint32_t e) { return (e >> 16) & 15u; }
B2_HD uint32_t huff_zinc2_raw(uint32_t e) { return (e >> 20) & 31u; }                                  // 0: the second symbol is EOB
B2_HD uint32_t huff_zinc2(uint32_t e) { const uint32_t z = huff_zinc2_raw(e); return z ? z : 64u; }   // only if adv2 != 0
B2_HD uint32_t huff_size2(uint32_t e) { return (e >> 25) & 7u; }
// level-1 entries with a flag (see above)
B2_HD uint32_t huff_l1_long(uint32_t blk) { return kHLong | (blk << 20); }
B2_HD uint32_t huff_l2_base(uint32_t e) { return (uint32_t)kHL1 + ((e >> 20) & 7u) * 64u; }
B2_HD bool huff_is_l1_error(uint32_t e) { return (e & (kHBadCode | kHBadSize)) != 0u && (e & 0xfffffu) == 0u; }
B2_HD uint32_t huff_l1_error(uint32_t e) {
  const uint32_t len = (e >> 20) & 31u;
  return huff_entry(len, len, 64, (e & (kHBadCode | kHBadSize)) | kHEob);
}
// symbol -> entry, following decodeMCUComponent (jpeg.cpp:330-400)
B2_HD uint32_t huff_symbol_entry(uint32_t sym, uint32_t len, bool is_dc) {
  if (is_dc) {
    if (sym > 11) return huff_entry(len, len, 64, kHBadSize | kHEob);      // jpeg.cpp:330-334
    return huff_entry(len + sym, len, 1, 0);
  }
  const uint32_t run = sym >> 4, size = sym & 15;
  if (sym == 0x00) return huff_entry(len, len, 64, kHEob);                 // jpeg.cpp:356-363
  if (size == 0) return huff_entry(len, len, sym == 0xf0 ? 16 : run, 0);   // ZRL / bare run, jpeg.cpp:368-380
  if (size > 10) return huff_entry(len, len, 64, kHBadSize | kHEob);       // jpeg.cpp:381-385
  return huff_entry(len + size, len, run + 1, 0);
}

// Canonical description of one table: first code / count / symbol base per length (huffman.hpp:17-29).
struct HuffCanon {
  uint32_t first[18], count[18], base[18];
};
B2_HD void huff_canon(const HuffSpec &spec, HuffCanon &c) {
  uint32_t code = 0;
  c.first[0] = c.count[0] = c.base[0] = 0;
  for (int l = 1; l <= 16; ++l) {
    c.first[l] = code;
    c.count[l] = (uint32_t)spec.offset[l] - (uint32_t)spec.offset[l - 1];
    c.base[l] = spec.offset[l - 1];
    code = (code + c.count[l]) << 1;
  }
  c.first[17] = code; c.count[17] = 0; c.base[17] = 0;
}
// first match by increasing length over lengths [lo,hi], as getNextSymbol (jpeg.cpp:303-318);
// w16 = the next 16 bits of the stream, right-aligned
B2_HD uint32_t huff_search(const HuffSpec &spec, const HuffCanon &c, uint32_t w16, int lo, int hi, bool is_dc) {
  for (int l = lo; l <= hi; ++l) {
    const uint32_t k = (w16 >> (16 - l)) - c.first[l];
    if (k < c.count[l]) return huff_symbol_entry(spec.symbols[c.base[l] + k], (uint32_t)l, is_dc);
  }
  return huff_entry(16, 16, 64, kHBadCode | kHEob);                         // jpeg.cpp:319 "return -1"
}
// resolves a code the tables could not (kHOverflow): straight from the DHT data in global memory
B2_COLD uint32_t huff_slow(const HuffSpec &spec, uint32_t w16, bool is_dc) {
  HuffCanon c;
  huff_canon(spec, c);
  return huff_search(spec, c, w16, kHLutBits + 1, 16, is_dc);
}

// One CTA per (table set, table): luts[(set*8 + t) * kHTab ...].  t < 4: DC table t, else AC table t-4.
B2_GLOBAL void k_build_luts(const TableSet *__restrict__ tsets, uint32_t *__restrict__ luts) {
  B2_SHARED HuffCanon canon;
  const int set = (int)blockIdx.x / kHTabsPerSet, t = (int)blockIdx.x % kHTabsPerSet;
  const bool is_dc = t < 4;
  const HuffSpec &spec = is_dc ? tsets[set].dc[t] : tsets[set].ac[t - 4];
  uint32_t *lut = luts + (size_t)blockIdx.x * kHTab;
  if (threadIdx.x == 0) huff_canon(spec, canon);
  __syncthreads();
  const uint32_t p0 = canon.first[kHLutBits + 1] >> 1;      // first 10-bit prefix not taken by a short code
  for (int idx = threadIdx.x; idx < kHTab; idx += blockDim.x) {
    uint32_t e;
    if (!spec.set) e = huff_entry(16, 16, 64, kHBadCode | kHEob);
    else if (idx < kHL1) {
      e = huff_search(spec, canon, (uint32_t)idx << (16 - kHLutBits), 1, kHLutBits, is_dc);
      if (e & kHBadCode) {                                   // no short code: a long one, or none at all
        const uint32_t blk = (uint32_t)idx - p0;
        if ((uint32_t)idx >= p0 && blk < (uint32_t)kHL2Blocks) e = huff_l1_long(blk);
        else if ((uint32_t)idx >= p0) e = kHOverflow;
      } else if (!is_dc && !(e & (kHSpecial | kHEob)) && huff_adv(e) < (uint32_t)kHLutBits) {
        // a second AC symbol inside the index bits that are left?  its code AND its magnitude bits must fit
        const uint32_t a1 = huff_adv(e), left = (uint32_t)kHLutBits - a1;
        const uint32_t rest = ((uint32_t)idx << a1) & (uint32_t)(kHL1 - 1);            // the bits after symbol 1, left-aligned in the index
        const uint32_t e2 = huff_search(spec, canon, rest << (16 - kHLutBits), 1, (int)left, false);
        if (!(e2 & kHSpecial) && huff_adv(e2) <= left) {
          const uint32_t z2 = (e2 & kHEob) ? 0u : huff_zinc(e2);                      // 0 stands for EOB
          if (((e2 & kHEob) || (z2 >= 1u && z2 <= 16u)) && huff_size(e2) <= 7u) e |= (huff_adv(e2) << 16) | (z2 << 20) | (huff_size(e2) << 25);
        }
      }
    } else {
      const uint32_t blk = (uint32_t)(idx - kHL1) >> 6, low = (uint32_t)(idx - kHL1) & 63u;
      const uint32_t prefix = p0 + blk;
      e = prefix < (uint32_t)kHL1 ? huff_search(spec, canon, (prefix << 6) | low, kHLutBits + 1, 16, is_dc)
                                  : huff_entry(16, 16, 64, kHBadCode | kHEob);
    }
    // level 1: an error entry keeps its flag and the code length only
    if (idx < kHL1 && (e & (kHBadCode | kHBadSize))) e = (e & (kHBadCode | kHBadSize)) | (huff_len(e) << 20);
    lut[idx] = e;
  }
}

}  // namespace b2
