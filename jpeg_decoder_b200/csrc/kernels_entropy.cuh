// kernels_entropy.cuh -- stage E: parallel Huffman decode of the unstuffed scan into
// quantised DCT coefficients.  Replaces Image::decodeHuffman / decodeMCUComponent /
// getNextSymbol (jpeg.cpp:300-446) and BitStream::getBit/getBitN/align (file.hpp:130-164).
//
// The reference walks the scan bit by bit, one block after another.  Here the unstuffed scan
// of every image is cut into fixed-size subsequences ("subs", kSubBits bits), one thread each:
//
//   sync  : each thread decodes the blocks that START inside its sub.  Its start state is
//           (bit position, block-in-MCU index) of the first block start at/after its sub's
//           beginning -- unknown at first, so it is guessed, and then corrected by Jacobi
//           iteration: thread t re-decodes whenever thread t-1's end state changed.  Huffman
//           codes self-synchronise, so this settles after one or two rounds.  Segment starts
//           (RSTn) are known bit positions and re-synchronise every decoder that reaches them.
//   fix   : the same across CTA boundaries (one more launch per hop; two are issued).
//   scan  : segmented prefix sums over the per-sub block counts and DC-difference sums give
//           every sub its first output block and its DC predictors (jpeg.cpp:344-345, 413).
//   write : final decode from the settled states; each thread assembles whole blocks in shared
//           memory and stores them as full 128-byte lines.
//
// Output layout: int16 coef[block][64], blocks in scan order (MCU by MCU, inside an MCU the
// hs*vs luma blocks row-major, then Cb, then Cr -- jpeg.cpp:426-435), the 64 entries in ZIG-ZAG
// order (the order they arrive in; stage P de-zig-zags for free when it names its registers).
// Algorithmic bytes: scan bytes in (each pass) + 128 B per coded block out.
#pragma once
#include "platform.h"
#include "dev_types.h"
#include "kernels_unstuff.cuh"   // find_owner

namespace b2 {

constexpr int kSubBits = 1024;
constexpr int kSubWords = kSubBits / 32;
constexpr int kCT = 128;                    // subs (= threads) per entropy CTA
constexpr int kLutBits = 10;
constexpr int kLutSize = 1 << kLutBits;
constexpr int kHaloWords = 64;              // a block is at most 64*(16+11) bits = 54 words, + look-ahead
constexpr int kChunkWords = kCT * kSubWords + kHaloWords;
constexpr int kMaxSlots = 6;                // <= 3 DC + 3 AC tables per image

// LUT entry: adv | zinc<<8 | len<<16 | run<<24 ; 0 = code longer than kLutBits (or invalid)
//   adv  = bits consumed by code + magnitude,  len = code length,
//   zinc = advance of the zig-zag index (64 ends the block),  run = zeros before the value.
//   run 0xFF / 0xFE mark symbols the reference rejects (bad size / no such code).
constexpr uint32_t kRunBadSize = 0xffu, kRunBadCode = 0xfeu;
B2_HD uint32_t lut_entry(uint32_t adv, uint32_t zinc, uint32_t len, uint32_t run) {
  return adv | (zinc << 8) | (len << 16) | (run << 24);
}
// symbol -> entry, following decodeMCUComponent (jpeg.cpp:330-400)
B2_HD uint32_t make_entry(uint32_t sym, uint32_t len, bool is_dc) {
  if (is_dc) {
    if (sym > 11) return lut_entry(len, 64, len, kRunBadSize);          // jpeg.cpp:330-334
    return lut_entry(len + sym, 0, len, 0);
  }
  const uint32_t run = sym >> 4, size = sym & 15;
  if (sym == 0x00) return lut_entry(len, 64, len, 0);                   // EOB, jpeg.cpp:356-363
  if (size == 0) {                                                      // ZRL (16) or a bare run, jpeg.cpp:368-380
    const uint32_t z = sym == 0xf0 ? 16 : run;
    return lut_entry(len, z, len, z);
  }
  if (size > 10) return lut_entry(len, 64, len, kRunBadSize);           // jpeg.cpp:381-385
  return lut_entry(len + size, run + 1, len, run);
}

// Per-table data for codes longer than kLutBits: first canonical code and count per length.
struct SlowTab {
  uint32_t first[17];   // first[l] = first code of length l (huffman.hpp:17-29)
  uint32_t count[17];
  uint32_t base[17];    // index of that code's symbol
  uint8_t symbols[176];
};

struct EntropyShared {
  SlowTab slow[kMaxSlots];
  SubState e[kCT];                     // end state per sub
  int8_t dc_slot[kMaxBpm], ac_slot[kMaxBpm], comp_of[kMaxBpm];
  uint32_t us_bits;
  int32_t bpm, seg_lo, seg_hi, nslots;
  SubState cta_in;
  int32_t cta_in_valid;
  int32_t img;
};

B2_DEVICE uint32_t slow_lookup(const SlowTab &t, uint32_t w, bool is_dc) {
  for (int l = kLutBits + 1; l <= 16; ++l) {
    const uint32_t idx = (w >> (32 - l)) - t.first[l];
    if (idx < t.count[l]) return make_entry(t.symbols[t.base[l] + idx], (uint32_t)l, is_dc);
  }
  return lut_entry(16, 64, 16, kRunBadCode);                            // jpeg.cpp:319 "return -1"
}

// Cooperative build of one table's LUT + SlowTab from the DHT counts (canonical codes as
// HuffmanTable::generateCodes, huffman.hpp:17-29; first match by increasing length as
// getNextSymbol, jpeg.cpp:303-318).
B2_DEVICE void build_table(const HuffSpec &spec, bool is_dc, uint32_t *lut, SlowTab &slow) {
  if (threadIdx.x == 0) {
    uint32_t code = 0;
    for (int l = 1; l <= 16; ++l) {
      slow.first[l] = code;
      slow.count[l] = (uint32_t)spec.offset[l] - (uint32_t)spec.offset[l - 1];
      slow.base[l] = spec.offset[l - 1];
      code = (code + slow.count[l]) << 1;
    }
    slow.first[0] = slow.count[0] = slow.base[0] = 0;
  }
  for (int k = threadIdx.x; k < 176; k += blockDim.x) slow.symbols[k] = spec.symbols[k];
  __syncthreads();
  for (int idx = threadIdx.x; idx < kLutSize; idx += blockDim.x) {
    uint32_t e = 0;
    for (int l = 1; l <= kLutBits; ++l) {
      const uint32_t k = ((uint32_t)idx >> (kLutBits - l)) - slow.first[l];
      if (k < slow.count[l]) { e = make_entry(slow.symbols[slow.base[l] + k], (uint32_t)l, is_dc); break; }
    }
    lut[idx] = e;
  }
}

// MSB-first bit reader over a word array (shared-memory chunk with one pad word per 32, or
// global memory); replaces BitStream (file.hpp:122-165).
template <bool PADDED>
struct BitReader {
  const uint32_t *base;   // word 0 = chunk start
  uint64_t buf;
  int avail;
  uint32_t widx;
  B2_MEMBER uint32_t ld(uint32_t i) const { return PADDED ? base[i + (i >> 5)] : base[i]; }
  B2_MEMBER void seek(uint32_t bit) {       // bit position relative to `base`
    widx = bit >> 5;
    const uint32_t o = bit & 31;
    const uint64_t w0 = ld(widx), w1 = ld(widx + 1);
    buf = ((w0 << 32) | w1) << o;
    avail = 64 - (int)o;
    widx += 2;
  }
  B2_MEMBER void refill() {
    if (avail <= 32) { buf |= (uint64_t)ld(widx) << (32 - avail); avail += 32; ++widx; }
  }
  B2_MEMBER uint32_t peek() const { return (uint32_t)(buf >> 32); }
  B2_MEMBER void skip(uint32_t n) { buf <<= n; avail -= (int)n; }
  B2_MEMBER uint32_t pos() const { return (widx << 5) - (uint32_t)avail; }
};

B2_DEVICE int extend(uint32_t w, uint32_t len, uint32_t size) {    // jpeg.cpp:335-343, 388-397
  const uint32_t v = (w << len) >> (32 - size);
  return (v >> (size - 1)) ? (int)v : (int)v - (int)((1u << size) - 1);
}

// Decode one block.  WRITE=false: lengths only (+ the DC difference).  WRITE=true: values go to
// the thread's staging block `stage` (zig-zag order), errors are OR-ed into *err.
template <bool WRITE, class Reader>
B2_DEVICE void decode_block(Reader &r, const uint32_t *__restrict__ dcl, const uint32_t *__restrict__ acl,
                            const SlowTab &dcs, const SlowTab &acs, int &dcdiff, uint16_t *stage, int stage_stride,
                            uint32_t *err) {
  r.refill();
  uint32_t w = r.peek();
  uint32_t e = dcl[w >> (32 - kLutBits)];
  if (e == 0) e = slow_lookup(dcs, w, true);
  {
    const uint32_t adv = e & 0xff, len = (e >> 16) & 0xff, run = e >> 24;
    dcdiff = adv > len ? extend(w, len, adv - len) : 0;
    if (WRITE && run >= kRunBadCode) *err |= run == kRunBadCode ? ST_BAD_CODE : ST_BAD_DC_LEN;
    r.skip(adv);
  }
  uint32_t z = 1;
  while (z < 64) {
    r.refill();
    w = r.peek();
    e = acl[w >> (32 - kLutBits)];
    if (e == 0) e = slow_lookup(acs, w, false);
    const uint32_t adv = e & 0xff, zinc = (e >> 8) & 0xff;
    if (WRITE) {
      const uint32_t len = (e >> 16) & 0xff, run = e >> 24;
      if (run >= kRunBadCode) { *err |= run == kRunBadCode ? ST_BAD_CODE : ST_BAD_AC_LEN; }
      else {
        const uint32_t zr = z + run;
        if (zr > 63) { *err |= ST_BAD_RUN; }                            // jpeg.cpp:372-376
        else if (adv > len) {
          // chunk (zr>>3) of this thread lives at ((zr>>3)*stride + tid) * 8 halfwords
          stage[((zr >> 3) * stage_stride) * 8 + (zr & 7)] = (uint16_t)extend(w, len, adv - len);
        }
      }
    }
    r.skip(adv);
    z += zinc;
  }
}

// True when a block that ended at bit `pe` was the last one of the segment ending at `B`:
// it reached/passed B, or only the 1-bit padding of the last byte is left (T.81 F.1.2.3; a
// complete Huffman code is never all ones, so padding cannot be taken for another block).
template <class Reader>
B2_DEVICE bool segment_done(Reader &r, uint32_t pe, uint32_t B) {
  if (pe >= B) return true;
  const uint32_t n = B - pe;
  if (n > 7) return false;
  r.refill();
  return (r.peek() >> (32 - n)) == ((1u << n) - 1u);
}

// largest k in [lo,hi] with seg[k] <= p
B2_DEVICE int seg_find(const uint32_t *__restrict__ seg, int lo, int hi, uint32_t p) {
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (seg[mid] <= p) lo = mid; else hi = mid - 1;
  }
  return lo;
}

struct SubResult { SubState e; int nb, d0, d1, d2, flag; };

// Decode every block that starts in [start.p, hi) -- lengths only -- and return the state at the
// first block start >= hi together with the block count and DC-difference sums accumulated
// since the last segment start met on the way (flag = one was met).
template <class Reader>
B2_DEVICE SubResult decode_sub_sync(const EntropyShared &S, const uint32_t *__restrict__ luts, Reader r,
                                    uint32_t chunk_bit0, const uint32_t *__restrict__ seg, SubState start,
                                    uint32_t hi) {
  SubResult R; R.e = start; R.nb = R.d0 = R.d1 = R.d2 = 0; R.flag = 0;
  uint32_t p = start.p, blk = start.blk;
  if (hi > S.us_bits) hi = S.us_bits;
  if (p >= hi) return R;
  int k = seg_find(seg, S.seg_lo, S.seg_hi, p);
  uint32_t B = seg[k + 1];
  r.seek(p - chunk_bit0);
  while (p < hi) {
    const int ds = S.dc_slot[blk], as = S.ac_slot[blk];
    int diff;
    decode_block<false>(r, luts + ds * kLutSize, luts + as * kLutSize, S.slow[ds], S.slow[as], diff, nullptr, 0, nullptr);
    const int c = S.comp_of[blk];
    if (c == 0) R.d0 += diff; else if (c == 1) R.d1 += diff; else R.d2 += diff;
    ++R.nb;
    const uint32_t pe = chunk_bit0 + r.pos();
    if (segment_done(r, pe, B)) {   // last block of its segment
      p = B; blk = 0;
      R.nb = R.d0 = R.d1 = R.d2 = 0; R.flag = 1;
      if (B >= S.us_bits) break;
      ++k; B = seg[k + 1];
      r.seek(p - chunk_bit0);
    } else {
      p = pe;
      blk = blk + 1 == (uint32_t)S.bpm ? 0 : blk + 1;
    }
  }
  R.e.p = p; R.e.blk = blk;
  return R;
}

// Segmented-sum monoid over (flag, nb, d0, d1, d2): a flag restarts the sums.
struct Seg5 { int nb, d0, d1, d2, flag; };
B2_DEVICE Seg5 seg5_combine(const Seg5 &a, const Seg5 &b) {
  if (b.flag) return b;
  Seg5 r; r.nb = a.nb + b.nb; r.d0 = a.d0 + b.d0; r.d1 = a.d1 + b.d1; r.d2 = a.d2 + b.d2; r.flag = a.flag;
  return r;
}
B2_DEVICE Seg5 seg5_shfl_up(const Seg5 &v, int d) {
  Seg5 o;
  o.nb = __shfl_up_sync(0xffffffffu, v.nb, d); o.d0 = __shfl_up_sync(0xffffffffu, v.d0, d);
  o.d1 = __shfl_up_sync(0xffffffffu, v.d1, d); o.d2 = __shfl_up_sync(0xffffffffu, v.d2, d);
  o.flag = __shfl_up_sync(0xffffffffu, v.flag, d);
  return o;
}
B2_DEVICE Seg5 seg5_shfl(const Seg5 &v, int src) {
  Seg5 o;
  o.nb = __shfl_sync(0xffffffffu, v.nb, src); o.d0 = __shfl_sync(0xffffffffu, v.d0, src);
  o.d1 = __shfl_sync(0xffffffffu, v.d1, src); o.d2 = __shfl_sync(0xffffffffu, v.d2, src);
  o.flag = __shfl_sync(0xffffffffu, v.flag, src);
  return o;
}
B2_DEVICE Seg5 seg5_warp_incl(Seg5 v) {
  const int lane = threadIdx.x & 31;
  for (int d = 1; d < 32; d <<= 1) {
    const Seg5 o = seg5_shfl_up(v, d);
    if (lane >= d) v = seg5_combine(o, v);
  }
  return v;
}
B2_DEVICE Seg5 seg5_from(const SubInfo &i) {
  Seg5 s; s.nb = i.nb_flag & 0x7fffffff; s.flag = (int)((uint32_t)i.nb_flag >> 31); s.d0 = i.d0; s.d1 = i.d1; s.d2 = i.d2;
  return s;
}
B2_DEVICE SubInfo seg5_to(const Seg5 &s) {
  SubInfo i; i.nb_flag = (int32_t)((uint32_t)s.nb | ((uint32_t)s.flag << 31)); i.d0 = s.d0; i.d1 = s.d1; i.d2 = s.d2;
  return i;
}
// CTA-wide exclusive segmented scan; *total = combine of all threads' values.
B2_DEVICE Seg5 seg5_cta_excl(const Seg5 &v, Seg5 *total) {
  B2_SHARED Seg5 wsum[kCT / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const Seg5 inc = seg5_warp_incl(v);
  if (lane == 31) wsum[wid] = inc;
  __syncthreads();
  Seg5 base; base.nb = base.d0 = base.d1 = base.d2 = base.flag = 0;
  Seg5 tot = base;
  for (int w = 0; w < kCT / 32; ++w) { if (w < wid) base = seg5_combine(base, wsum[w]); tot = seg5_combine(tot, wsum[w]); }
  __syncthreads();
  Seg5 prev = seg5_shfl_up(inc, 1);
  if (lane == 0) { prev.nb = prev.d0 = prev.d1 = prev.d2 = prev.flag = 0; }
  *total = tot;
  return seg5_combine(base, prev);
}

// Common CTA prologue: image lookup, table build, chunk staging.  Returns false (uniformly) when
// the CTA has nothing to decode.
B2_DEVICE bool entropy_prologue(EntropyShared &S, uint32_t *luts, uint32_t *chunk, const ImageDev *__restrict__ imgs,
                                const int32_t *__restrict__ cta_base, int nimg, const TableSet *__restrict__ tsets,
                                const ImageOut *__restrict__ out, const uint8_t *__restrict__ d_us,
                                const uint32_t *__restrict__ seg_start, bool stage_chunk, ImageDev &im,
                                uint32_t &chunk_bit0) {
  if (threadIdx.x == 0) {
    const int i = find_owner(cta_base, nimg, (int)blockIdx.x);
    S.img = i;
    S.us_bits = out[i].us_len * 8u;
  }
  __syncthreads();
  im = imgs[S.img];
  const uint32_t local_cta = (uint32_t)blockIdx.x - (uint32_t)im.cta_base;
  chunk_bit0 = local_cta * (uint32_t)(kCT * kSubBits);
  if (chunk_bit0 >= S.us_bits) return false;
  const TableSet &ts = tsets[im.tset];
  if (threadIdx.x == 0) {
    // one LUT slot per distinct (class, id) used by the scan's components
    int nslots = 0; int slot_of_dc[4] = {-1, -1, -1, -1}, slot_of_ac[4] = {-1, -1, -1, -1};
    int8_t cdc[3], cac[3];
    for (int c = 0; c < im.ncomp; ++c) {
      const int td = im.td[c] & 3, ta = im.ta[c] & 3;
      if (slot_of_dc[td] < 0) slot_of_dc[td] = nslots++;
      if (slot_of_ac[ta] < 0) slot_of_ac[ta] = nslots++;
      cdc[c] = (int8_t)slot_of_dc[td]; cac[c] = (int8_t)slot_of_ac[ta];
    }
    S.nslots = nslots;
    const int nluma = im.ncomp == 1 ? 1 : im.hs * im.vs;
    for (int b = 0; b < im.bpm; ++b) {
      const int c = b < nluma ? 0 : b - nluma + 1;
      S.comp_of[b] = (int8_t)c; S.dc_slot[b] = cdc[c]; S.ac_slot[b] = cac[c];
    }
    S.bpm = im.bpm;
    const uint32_t *seg = seg_start + im.seg_base;
    S.seg_lo = seg_find(seg, 0, im.nseg - 1, chunk_bit0);
    S.seg_hi = seg_find(seg, 0, im.nseg - 1, chunk_bit0 + (uint32_t)(kCT * kSubBits + kHaloWords * 32));
  }
  __syncthreads();
  {
    bool done_dc[4] = {false, false, false, false}, done_ac[4] = {false, false, false, false};
    for (int c = 0; c < im.ncomp; ++c) {
      const int td = im.td[c] & 3, ta = im.ta[c] & 3;
      const int b = c == 0 ? 0 : im.bpm - im.ncomp + c;   // a block of component c
      if (!done_dc[td]) { done_dc[td] = true; build_table(ts.dc[td], true, luts + S.dc_slot[b] * kLutSize, S.slow[S.dc_slot[b]]); }
      if (!done_ac[ta]) { done_ac[ta] = true; build_table(ts.ac[ta], false, luts + S.ac_slot[b] * kLutSize, S.slow[S.ac_slot[b]]); }
    }
  }
  if (stage_chunk) {
    // coalesced 128-bit loads of the chunk (+halo) into shared memory, one pad word per 32
    const uint4 *src = reinterpret_cast<const uint4 *>(d_us + im.raw_off + (uint64_t)local_cta * (kCT * kSubBits / 8));
    for (int v = threadIdx.x; v < kChunkWords / 4; v += blockDim.x) {
      const uint4 q = b2_ld_nc_v4(src + v);
      const uint32_t i0 = (uint32_t)v * 4, a = i0 + (i0 >> 5);   // the 4 words never straddle a pad
      chunk[a] = q.x; chunk[a + 1] = q.y; chunk[a + 2] = q.z; chunk[a + 3] = q.w;
    }
  }
  __syncthreads();
  return true;
}

constexpr int kChunkSmemWords = kChunkWords + kChunkWords / 32 + 2;
B2_HD size_t entropy_smem_bytes(int nslots, bool staging) {
  return sizeof(uint32_t) * ((size_t)nslots * kLutSize + kChunkSmemWords) + (staging ? (size_t)kCT * 128 : 0) + 16;
}

// ---- sync / fix -----------------------------------------------------------------------------
// FIRST=true : round 0 from guessed starts, then Jacobi rounds inside the CTA.
// FIRST=false: take the previous CTA's settled end state (bnd_in) and repair, if it changed.
template <bool FIRST>
B2_GLOBAL void __launch_bounds__(kCT)
k_entropy_sync(const ImageDev *__restrict__ imgs, const int32_t *__restrict__ cta_base, int nimg,
               const TableSet *__restrict__ tsets, const ImageOut *__restrict__ out,
               const uint8_t *__restrict__ d_us, const uint32_t *__restrict__ seg_start,
               SubState *__restrict__ sub_e, SubState *__restrict__ sub_used, SubInfo *__restrict__ sub_info,
               SubInfo *__restrict__ cta_tot, const SubState *__restrict__ bnd_in, SubState *__restrict__ bnd_out,
               uint32_t *__restrict__ changed_flag, int nslots_max) {
  B2_DYN_SMEM(uint32_t, smem);
  B2_SHARED EntropyShared S;
  uint32_t *luts = smem;
  uint32_t *chunk = smem + (size_t)nslots_max * kLutSize;
  const int tid = threadIdx.x;

  if (!FIRST) {
    // cheap uniform exit: nothing upstream of this CTA moved
    B2_SHARED int s_skip;
    if (tid == 0) {
      const int i = find_owner(cta_base, nimg, (int)blockIdx.x);
      const int c = (int)blockIdx.x;
      int skip;
      if (c == imgs[i].cta_base) skip = 1;                 // first CTA of an image starts from the truth
      else {
        const SubState in = bnd_in[c - 1];
        const SubState used = sub_used[imgs[i].sub_base + (c - imgs[i].cta_base) * kCT];
        skip = in.p == used.p && in.blk == used.blk;
      }
      if (skip) bnd_out[c] = bnd_in[c];
      s_skip = skip;
    }
    __syncthreads();
    if (s_skip) return;
  }

  ImageDev im; uint32_t chunk_bit0;
  const bool live = entropy_prologue(S, luts, chunk, imgs, cta_base, nimg, tsets, out, d_us, seg_start, true, im, chunk_bit0);
  const int cta = (int)blockIdx.x, local_cta = cta - im.cta_base;
  const int gsub = im.sub_base + local_cta * kCT + tid;
  if (!live) {
    // beyond the end of this image's unstuffed scan: empty subs
    if (FIRST) {
      SubState z; z.p = S.us_bits; z.blk = 0;
      sub_e[gsub] = z; sub_used[gsub] = z;
      SubInfo zi; zi.nb_flag = zi.d0 = zi.d1 = zi.d2 = 0; sub_info[gsub] = zi;
      if (tid == 0) { cta_tot[cta] = zi; bnd_out[cta] = z; }
    } else if (tid == 0) bnd_out[cta] = bnd_in[cta];
    return;
  }
  const uint32_t *seg = seg_start + im.seg_base;
  const uint32_t lo = chunk_bit0 + (uint32_t)tid * kSubBits, hi = lo + kSubBits;
  BitReader<true> rd; rd.base = chunk;

  SubState used; SubResult R;
  if (FIRST) {
    used.p = lo; used.blk = 0;                                  // guess; exact for the image's first sub
    R = decode_sub_sync(S, luts, rd, chunk_bit0, seg, used, hi);
  } else {
    used = sub_used[gsub];
    R.e = sub_e[gsub];
    const Seg5 s = seg5_from(sub_info[gsub]);
    R.nb = s.nb; R.d0 = s.d0; R.d1 = s.d1; R.d2 = s.d2; R.flag = s.flag;
    if (tid == 0) { S.cta_in = bnd_in[cta - 1]; }
  }
  S.e[tid] = R.e;
  __syncthreads();
  for (int round = 0; round < kCT; ++round) {
    SubState prev;
    bool have = true;
    if (tid == 0) { if (FIRST) have = false; else prev = S.cta_in; }
    else prev = S.e[tid - 1];
    const bool need = have && (prev.p != used.p || prev.blk != used.blk);
    __syncthreads();                                            // every S.e read happens before any write
    if (need) {
      used = prev;
      R = decode_sub_sync(S, luts, rd, chunk_bit0, seg, used, hi);
      S.e[tid] = R.e;
    }
    if (!__syncthreads_or(need ? 1 : 0)) break;
  }
  sub_e[gsub] = R.e; sub_used[gsub] = used;
  Seg5 mine; mine.nb = R.nb; mine.d0 = R.d0; mine.d1 = R.d1; mine.d2 = R.d2; mine.flag = R.flag;
  sub_info[gsub] = seg5_to(mine);
  Seg5 total;
  seg5_cta_excl(mine, &total);
  if (tid == kCT - 1) {
    if (!FIRST) {
      const SubState old = bnd_in[cta];
      if (old.p != R.e.p || old.blk != R.e.blk) atomicOr(changed_flag, 1u);
    }
    bnd_out[cta] = R.e;
  }
  if (tid == 0) cta_tot[cta] = seg5_to(total);
}

// ---- scan over CTA totals: one warp per image ----------------------------------------------
B2_GLOBAL void k_entropy_scan(const ImageDev *__restrict__ imgs, int nimg, const SubInfo *__restrict__ cta_tot,
                              SubInfo *__restrict__ cta_pre) {
  const int i = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (i >= nimg) return;
  const int c0 = imgs[i].cta_base, nc = imgs[i].ncta;
  Seg5 carry; carry.nb = carry.d0 = carry.d1 = carry.d2 = carry.flag = 0;
  for (int b = 0; b < nc; b += 32) {
    const int c = b + lane;
    Seg5 v; v.nb = v.d0 = v.d1 = v.d2 = v.flag = 0;
    if (c < nc) v = seg5_from(cta_tot[c0 + c]);
    const Seg5 inc = seg5_warp_incl(v);
    Seg5 prev = seg5_shfl_up(inc, 1);
    if (lane == 0) { prev.nb = prev.d0 = prev.d1 = prev.d2 = prev.flag = 0; }
    if (c < nc) cta_pre[c0 + c] = seg5_to(seg5_combine(carry, prev));
    carry = seg5_combine(carry, seg5_shfl(inc, 31));
  }
}

// ---- write ------------------------------------------------------------------------------------
B2_GLOBAL void __launch_bounds__(kCT)
k_entropy_write(const ImageDev *__restrict__ imgs, const int32_t *__restrict__ cta_base, int nimg,
                const TableSet *__restrict__ tsets, ImageOut *__restrict__ out, const uint8_t *__restrict__ d_us,
                const uint32_t *__restrict__ seg_start, const SubState *__restrict__ sub_e,
                const SubInfo *__restrict__ sub_info, const SubInfo *__restrict__ cta_pre,
                int16_t *__restrict__ coef, int nslots_max) {
  B2_DYN_SMEM(uint32_t, smem);
  B2_SHARED EntropyShared S;
  uint32_t *luts = smem;
  uint32_t *chunk = smem + (size_t)nslots_max * kLutSize;
  uint4 *stage4 = reinterpret_cast<uint4 *>(chunk + ((kChunkSmemWords + 3) & ~3));
  const int tid = threadIdx.x;
  ImageDev im; uint32_t chunk_bit0;
  if (!entropy_prologue(S, luts, chunk, imgs, cta_base, nimg, tsets, out, d_us, seg_start, true, im, chunk_bit0)) return;
  const int cta = (int)blockIdx.x, local_cta = cta - im.cta_base;
  const int gsub = im.sub_base + local_cta * kCT + tid;
  const uint32_t *seg = seg_start + im.seg_base;
  const uint32_t lo = chunk_bit0 + (uint32_t)tid * kSubBits;
  uint32_t hi = lo + kSubBits;
  if (hi > S.us_bits) hi = S.us_bits;

  // first output block and DC predictors of this sub
  const Seg5 mine = seg5_from(sub_info[gsub]);
  Seg5 total;
  Seg5 pre = seg5_cta_excl(mine, &total);
  pre = seg5_combine(seg5_from(cta_pre[cta]), pre);

  SubState start;
  if (local_cta == 0 && tid == 0) { start.p = 0; start.blk = 0; }
  else start = sub_e[gsub - 1];

  // staging: chunk k (8 coefficients) of thread t at stage4[k*kCT + t]
#pragma unroll
  for (int k = 0; k < 8; ++k) stage4[k * kCT + tid] = make_uint4(0, 0, 0, 0);
  uint16_t *stage = reinterpret_cast<uint16_t *>(stage4 + tid);

  uint32_t p = start.p, blk = start.blk, err = 0;
  if (p < hi) {
    int k = seg_find(seg, S.seg_lo, S.seg_hi, p);
    uint32_t B = seg[k + 1];
    const int seg_blocks = im.restart_interval * im.bpm;            // 0 when there is one segment
    int bidx = k * seg_blocks + pre.nb;                              // first output block
    int limit = (k + 1 < im.nseg) ? (k + 1) * seg_blocks : im.nblocks;
    int pred[3] = {pre.d0, pre.d1, pre.d2};
    BitReader<true> rd; rd.base = chunk;
    rd.seek(p - chunk_bit0);
    int16_t *obase = coef + im.coef_base * 64;
    while (p < hi) {
      const int ds = S.dc_slot[blk], as = S.ac_slot[blk];
      int diff;
      decode_block<true>(rd, luts + ds * kLutSize, luts + as * kLutSize, S.slow[ds], S.slow[as], diff, stage, kCT, &err);
      const int c = S.comp_of[blk];
      const int dc = (pred[c] += diff);
      if (dc < -32768 || dc > 32767) err |= ST_DC_RANGE;
      stage[0] = (uint16_t)dc;
      if (bidx < limit) {                                            // blocks past the segment's quota are trailing garbage
        uint4 *o = reinterpret_cast<uint4 *>(obase + (int64_t)bidx * 64);
#pragma unroll
        for (int q = 0; q < 8; ++q) { o[q] = stage4[q * kCT + tid]; }
      }
#pragma unroll
      for (int q = 0; q < 8; ++q) stage4[q * kCT + tid] = make_uint4(0, 0, 0, 0);
      ++bidx;
      const uint32_t pe = chunk_bit0 + rd.pos();
      if (segment_done(rd, pe, B)) {
        if (pe > B) err |= ST_OVERRUN;
        if (bidx < limit) err |= ST_BLOCK_COUNT;
        if (bidx > limit && k + 1 < im.nseg) err |= ST_BLOCK_COUNT;
        p = B; blk = 0;
        if (B >= S.us_bits) break;
        ++k; B = seg[k + 1];
        bidx = k * seg_blocks; limit = (k + 1 < im.nseg) ? (k + 1) * seg_blocks : im.nblocks;
        pred[0] = pred[1] = pred[2] = 0;
        rd.seek(p - chunk_bit0);
      } else {
        p = pe;
        blk = blk + 1 == (uint32_t)S.bpm ? 0 : blk + 1;
      }
    }
  }
  if (err) atomicOr(&out[S.img].status, err);
}

}  // namespace b2
