// kernels_entropy.cuh -- stage E: parallel Huffman decode of the unstuffed scan into
// quantised DCT coefficients.  Replaces Image::decodeHuffman / decodeMCUComponent /
// getNextSymbol (jpeg.cpp:300-446) and BitStream::getBit/getBitN/align (file.hpp:130-164).
//
// The reference walks the scan bit by bit, one block after another.  Here the unstuffed scan
// of every image is cut into fixed-size subsequences ("subs", kSubBits bits), one thread each:
//
//   sync  : each thread decodes the blocks that START inside its sub.  Its start state is
//           (bit position, block-in-MCU index) of the first block start at/after its sub's
//           beginning.  That state is unknown, so the thread starts kOverlapBits EARLIER from a
//           guess (or from the last segment start, which is exact) and lets the Huffman code
//           self-synchronise on the way: measured on 4:2:0 streams the state -- including the
//           block-in-MCU phase -- is right after 4096 bits for 92 % of starts (4:4:4: > 99.9 %).
//           Thread t then checks its start against thread t-1's end state; the few subs that
//           disagree are re-decoded (Jacobi iteration, compacted so whole warps stay busy).
//           Segment starts (RSTn) re-synchronise every decoder that reaches them.
//   fix   : the same check across CTA boundaries (one more launch per hop; two are issued).
//   scan  : segmented prefix sums over the per-sub block counts and DC-difference sums give
//           every sub its first output block and its DC predictors (jpeg.cpp:344-345, 413).
//   write : final decode from the settled states; each thread assembles whole blocks in shared
//           memory and stores them as full 128-byte lines.
//
// Output layout: int16 coef[block][64], blocks in scan order (MCU by MCU, inside an MCU the
// hs*vs luma blocks row-major, then Cb, then Cr -- jpeg.cpp:426-435), the 64 entries in ZIG-ZAG
// order (the order they arrive in; stage P de-zig-zags for free when it names its registers);
// the eight 16-byte chunks of a block are stored at chunk ^ (block & 7) so that stage P can read a
// dense TMA copy of consecutive blocks from shared memory without bank conflicts.
// Algorithmic bytes: scan bytes in (each pass) + 128 B per coded block out.
#pragma once
#include "platform.h"
#include "dev_types.h"
#include "kernels_unstuff.cuh"   // find_owner

namespace b2 {

#ifndef B2_SUB_BITS
#define B2_SUB_BITS 4096
#endif
#ifndef B2_OVERLAP_BITS
#define B2_OVERLAP_BITS 4096
#endif
constexpr int kSubBits = B2_SUB_BITS;
constexpr int kOverlapBits = B2_OVERLAP_BITS;         // self-synchronisation run-in before a sub (see header)
constexpr int kSubWords = kSubBits / 32;
constexpr int kCT = 128;                    // subs (= threads) per entropy CTA
constexpr int kLutBits = 10;
constexpr int kLutSize = 1 << kLutBits;
constexpr int kChunkWords = 64 + 8;          // read-ahead slack past the end of a scan: one block (<= 54 words) + queue
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
    return lut_entry(len + sym, 1, len, 0);                             // z: 0 -> 1
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
  SubState used[kCT];                  // start state the end state was computed from
  SubInfo info[kCT];                   // blocks / DC sums since the last segment start
  int32_t list[kCT];                   // compacted indices of subs to re-decode
  int32_t tab_of[kMaxBpm];             // per block-in-MCU: dc slot | ac slot << 8 | component << 16
  uint32_t us_bits;
  int32_t bpm, nslots, count;
  SubState cta_in;
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

// MSB-first bit reader over the (word-swapped) unstuffed scan in global memory; replaces
// BitStream (file.hpp:122-165).  128-bit loads into a four-word register queue feed a 64-bit
// left-aligned window: one LDG.128 per 128 bits consumed, each lane streaming through its own
// cache lines (L1 keeps the line for the 7 following loads).
struct BitReader {
  const uint4 *base;      // word 0 of the image's unstuffed scan (16-byte aligned)
  const uint4 *ptr;       // next vector to load
  uint4 q;
  int qn;                 // words left in q
  uint64_t buf;
  int avail;              // valid bits in buf
  uint32_t wpos;          // bit index of the next word that will leave the queue
  B2_MEMBER uint32_t next_word() {
    if (qn == 0) { q = b2_ld_nc_v4(ptr); ++ptr; qn = 4; }
    const uint32_t w = q.x;
    q.x = q.y; q.y = q.z; q.z = q.w; --qn;
    wpos += 32;
    return w;
  }
  B2_MEMBER void seek(uint32_t bit) {      // bit position relative to `base`
    const uint32_t wi = bit >> 5;
    ptr = base + (wi >> 2);
    q = b2_ld_nc_v4(ptr); ++ptr; qn = 4;
    for (uint32_t k = 0; k < (wi & 3u); ++k) { q.x = q.y; q.y = q.z; q.z = q.w; --qn; }
    wpos = wi << 5;
    const uint64_t w0 = next_word(), w1 = next_word();
    const uint32_t o = bit & 31u;
    buf = ((w0 << 32) | w1) << o;
    avail = 64 - (int)o;
  }
  B2_MEMBER void refill() {
    if (avail <= 32) { const uint64_t w = next_word(); buf |= w << (32 - avail); avail += 32; }
  }
  B2_MEMBER uint32_t peek() const { return (uint32_t)(buf >> 32); }
  B2_MEMBER void skip(uint32_t n) { buf <<= n; avail -= (int)n; }
  B2_MEMBER uint32_t pos() const { return wpos - (uint32_t)avail; }
};

B2_DEVICE int extend(uint32_t w, uint32_t len, uint32_t size) {    // jpeg.cpp:335-343, 388-397
  const uint32_t v = (w << len) >> (32 - size);
  return (v >> (size - 1)) ? (int)v : (int)v - (int)((1u << size) - 1);
}

// True when a block that ended at bit `pe` was the last one of the segment ending at `B`:
// it reached/passed B, or only the 1-bit padding of the last byte is left (T.81 F.1.2.3; a
// complete Huffman code is never all ones, so padding cannot be taken for another block).
B2_DEVICE bool segment_done(BitReader &r, uint32_t pe, uint32_t B) {
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

struct SubResult { SubState start, e; int nb, d0, d1, d2, flag; };

// What the write pass needs on top of the walk itself.
struct WriteCtx {
  uint4 *stage4;          // this thread's staging block: chunk k at stage4[k * kCT]
  int16_t *obase;         // coefficient buffer of the image
  int bidx, limit;        // next output block, first block past the current segment's quota
  int seg_blocks, nseg, nblocks;
  int pred[3];
  uint32_t err;
};

// The one decode loop of stage E: a flat loop over SYMBOLS (DC and AC alike) so that the lanes of a
// warp stay in step; block ends are handled inside it.  Walks the blocks that start in
// [from.p, hi).  While `lo` has not been reached the walk is only the self-synchronising run-in;
// at the first block start >= lo the counters restart and that state is reported as R.start.
//   WRITE=false: counts blocks and sums DC differences (per component) since the last segment
//                start met on the way (R.flag = one was met).
//   WRITE=true : also extracts the values into the staging block and stores finished blocks.
template <bool WRITE>
B2_DEVICE void decode_span(const EntropyShared &S, const uint32_t *__restrict__ luts, BitReader &r,
                           const uint32_t *__restrict__ seg, int nseg, SubState from, uint32_t lo, uint32_t hi,
                           SubResult &R, WriteCtx *W) {
  uint32_t p = from.p, blk = from.blk;
  bool started = p >= lo;
  R.start = from; R.e = from; R.nb = R.d0 = R.d1 = R.d2 = 0; R.flag = 0;
  if (hi > S.us_bits) hi = S.us_bits;
  if (p >= hi) return;
  int k = seg_find(seg, 0, nseg - 1, p);
  uint32_t B = seg[k + 1];
  r.seek(p);
  int tab = S.tab_of[blk];
  uint32_t z = 0;
  uint16_t *stage = WRITE ? reinterpret_cast<uint16_t *>(W->stage4) : nullptr;
  for (;;) {
    r.refill();
    const uint32_t w = r.peek();
    const bool is_dc = z == 0;
    const int slot = is_dc ? (tab & 0xff) : ((tab >> 8) & 0xff);
    uint32_t e = luts[slot * kLutSize + (w >> (32 - kLutBits))];
    if (e == 0) e = slow_lookup(S.slow[slot], w, is_dc);
    const uint32_t adv = e & 0xff, zinc = (e >> 8) & 0xff, len = (e >> 16) & 0xff, run = e >> 24;
    if (is_dc) {
      const int diff = adv > len ? extend(w, len, adv - len) : 0;     // jpeg.cpp:335-345
      const int c = tab >> 16;
      if (WRITE) {
        if (run >= kRunBadCode) W->err |= run == kRunBadCode ? ST_BAD_CODE : ST_BAD_DC_LEN;
        const int dc = (W->pred[c] += diff);
        if (dc < -32768 || dc > 32767) W->err |= ST_DC_RANGE;
        stage[0] = (uint16_t)dc;
      } else {
        if (c == 0) R.d0 += diff; else if (c == 1) R.d1 += diff; else R.d2 += diff;
      }
    } else if (WRITE) {
      if (run >= kRunBadCode) W->err |= run == kRunBadCode ? ST_BAD_CODE : ST_BAD_AC_LEN;
      else {
        const uint32_t zr = z + run;
        if (zr > 63) W->err |= ST_BAD_RUN;                               // jpeg.cpp:372-376
        else if (adv > len) stage[(zr >> 3) * (kCT * 8) + (zr & 7)] = (uint16_t)extend(w, len, adv - len);
      }
    }
    r.skip(adv);
    z += zinc;
    if (z < 64) continue;
    // ---- block end ----
    const uint32_t pe = r.pos();
    if (WRITE) {
      if (W->bidx < W->limit) {                  // blocks past the segment's quota are trailing garbage
        uint4 *o = reinterpret_cast<uint4 *>(W->obase + (int64_t)W->bidx * 64);
#pragma unroll
        for (int q = 0; q < 8; ++q) o[q ^ (W->bidx & 7)] = W->stage4[q * kCT];   // chunk swizzle, see kernels_pixel.cuh
      }
#pragma unroll
      for (int q = 0; q < 8; ++q) W->stage4[q * kCT] = make_uint4(0, 0, 0, 0);
      ++W->bidx;
    }
    ++R.nb;
    if (segment_done(r, pe, B)) {                // last block of its segment (jpeg.cpp:419-425)
      if (WRITE) {
        if (pe > B) W->err |= ST_OVERRUN;
        if (W->bidx < W->limit) W->err |= ST_BLOCK_COUNT;
        if (W->bidx > W->limit && k + 1 < W->nseg) W->err |= ST_BLOCK_COUNT;
      }
      p = B; blk = 0;
      R.nb = R.d0 = R.d1 = R.d2 = 0; R.flag = 1;
      if (B < S.us_bits) {
        ++k; B = seg[k + 1];
        r.seek(p);
        if (WRITE) {
          W->bidx = k * W->seg_blocks;
          W->limit = (k + 1 < W->nseg) ? (k + 1) * W->seg_blocks : W->nblocks;
          W->pred[0] = W->pred[1] = W->pred[2] = 0;
        }
      }
    } else {
      p = pe;
      blk = blk + 1 == (uint32_t)S.bpm ? 0 : blk + 1;
    }
    if (!started && p >= lo) {                   // the run-in is over: this is the sub's start state
      started = true;
      R.start.p = p; R.start.blk = blk;
      R.nb = R.d0 = R.d1 = R.d2 = 0; R.flag = 0;
    }
    if (p >= hi) break;
    tab = S.tab_of[blk];
    z = 0;
  }
  R.e.p = p; R.e.blk = blk;
}

// Segmented-sum monoid over (flag, nb, d0, d1, d2): a flag restarts the sums.
struct Seg5 { int nb, d0, d1, d2, flag; };
B2_DEVICE Seg5 seg5_zero() { Seg5 z; z.nb = z.d0 = z.d1 = z.d2 = z.flag = 0; return z; }
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
B2_DEVICE SubInfo info_of(const SubResult &R) {
  Seg5 s; s.nb = R.nb; s.d0 = R.d0; s.d1 = R.d1; s.d2 = R.d2; s.flag = R.flag;
  return seg5_to(s);
}
// CTA-wide exclusive segmented scan; *total = combine of all threads' values.
B2_DEVICE Seg5 seg5_cta_excl(const Seg5 &v, Seg5 *total) {
  B2_SHARED Seg5 wsum[kCT / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const Seg5 inc = seg5_warp_incl(v);
  if (lane == 31) wsum[wid] = inc;
  __syncthreads();
  Seg5 base = seg5_zero(), tot = seg5_zero();
  for (int w = 0; w < kCT / 32; ++w) { if (w < wid) base = seg5_combine(base, wsum[w]); tot = seg5_combine(tot, wsum[w]); }
  __syncthreads();
  Seg5 prev = seg5_shfl_up(inc, 1);
  if (lane == 0) prev = seg5_zero();
  *total = tot;
  return seg5_combine(base, prev);
}

// Common CTA prologue: image lookup and table build.  Returns false (uniformly) when the CTA lies
// past the end of its image's unstuffed scan.
B2_DEVICE bool entropy_prologue(EntropyShared &S, uint32_t *luts, const ImageDev *__restrict__ imgs,
                                const int32_t *__restrict__ cta_base, int nimg, const TableSet *__restrict__ tsets,
                                const ImageOut *__restrict__ out, ImageDev &im, uint32_t &chunk_bit0) {
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
    int cdc[3] = {0, 0, 0}, cac[3] = {0, 0, 0};
    for (int c = 0; c < im.ncomp; ++c) {
      const int td = im.td[c] & 3, ta = im.ta[c] & 3;
      if (slot_of_dc[td] < 0) slot_of_dc[td] = nslots++;
      if (slot_of_ac[ta] < 0) slot_of_ac[ta] = nslots++;
      cdc[c] = slot_of_dc[td]; cac[c] = slot_of_ac[ta];
    }
    S.nslots = nslots;
    const int nluma = im.ncomp == 1 ? 1 : im.hs * im.vs;
    for (int b = 0; b < im.bpm; ++b) {
      const int c = b < nluma ? 0 : b - nluma + 1;
      S.tab_of[b] = cdc[c] | (cac[c] << 8) | (c << 16);
    }
    S.bpm = im.bpm;
  }
  __syncthreads();
  {
    bool done_dc[4] = {false, false, false, false}, done_ac[4] = {false, false, false, false};
    for (int c = 0; c < im.ncomp; ++c) {
      const int td = im.td[c] & 3, ta = im.ta[c] & 3;
      const int t = S.tab_of[c == 0 ? 0 : im.bpm - im.ncomp + c];   // a block of component c
      if (!done_dc[td]) { done_dc[td] = true; build_table(ts.dc[td], true, luts + (t & 0xff) * kLutSize, S.slow[t & 0xff]); }
      if (!done_ac[ta]) { done_ac[ta] = true; build_table(ts.ac[ta], false, luts + ((t >> 8) & 0xff) * kLutSize, S.slow[(t >> 8) & 0xff]); }
    }
  }
  __syncthreads();
  return true;
}

B2_HD size_t entropy_smem_bytes(int nslots, bool staging) {
  return sizeof(uint32_t) * (size_t)nslots * kLutSize + (staging ? (size_t)kCT * 128 : 0) + 16;
}

// Jacobi repair inside a CTA: re-decode every sub whose start state differs from its predecessor's
// end state, the needy subs compacted onto the first threads so that whole warps stay busy.
// S.e / S.used / S.info hold the current states; thread 0's predecessor is S.cta_in if have_in.
B2_DEVICE void entropy_repair(EntropyShared &S, const uint32_t *luts, BitReader &rd, const uint32_t *seg, int nseg,
                              uint32_t chunk_bit0, bool have_in) {
  const int tid = threadIdx.x;
  for (int round = 0; round <= kCT; ++round) {
    SubState prev; bool need = false;
    if (tid > 0 || have_in) {
      prev = tid == 0 ? S.cta_in : S.e[tid - 1];
      const SubState u = S.used[tid];
      const uint32_t lo = chunk_bit0 + (uint32_t)tid * kSubBits;
      need = lo < S.us_bits && (prev.p != u.p || prev.blk != u.blk);
    }
    if (tid == 0) S.count = 0;
    __syncthreads();
    if (need) S.list[atomicAdd(&S.count, 1)] = tid;
    __syncthreads();
    const int cnt = S.count;
#ifdef B200JPEG_EMU
    if (tid == 0 && getenv("B2_EMU_TRACE")) fprintf(stderr, "cta %u repair round %d: %d subs\n", blockIdx.x, round, cnt);
#endif
    if (cnt == 0) break;
    SubResult R; int j = -1; SubState from;
    if (tid < cnt) {
      j = S.list[tid];
      from = j == 0 ? S.cta_in : S.e[j - 1];
      const uint32_t lo = chunk_bit0 + (uint32_t)j * kSubBits;
      decode_span<false>(S, luts, rd, seg, nseg, from, lo, lo + kSubBits, R, nullptr);
    }
    __syncthreads();                               // every S.e read of this round precedes every write
    if (j >= 0) { S.e[j] = R.e; S.used[j] = from; S.info[j] = info_of(R); }
    __syncthreads();
  }
}

// ---- sync / fix -----------------------------------------------------------------------------
// FIRST=true : run-in + own sub from a guess (or the last segment start), then repair in the CTA.
// FIRST=false: take the previous CTA's settled end state (bnd_in) and repair, if it differs.
template <bool FIRST>
B2_GLOBAL void __launch_bounds__(kCT)
k_entropy_sync(const ImageDev *__restrict__ imgs, const int32_t *__restrict__ cta_base, int nimg,
               const TableSet *__restrict__ tsets, const ImageOut *__restrict__ out,
               const uint8_t *__restrict__ d_us, const uint32_t *__restrict__ seg_start,
               SubState *__restrict__ sub_e, SubState *__restrict__ sub_used, SubInfo *__restrict__ sub_info,
               SubInfo *__restrict__ cta_tot, const SubState *__restrict__ bnd_in, SubState *__restrict__ bnd_out,
               uint32_t *__restrict__ changed_flag, int nslots_max) {
  B2_DYN_SMEM(uint32_t, luts);
  B2_SHARED EntropyShared S;
  const int tid = threadIdx.x;
  (void)nslots_max;

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
  const bool live = entropy_prologue(S, luts, imgs, cta_base, nimg, tsets, out, im, chunk_bit0);
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
  const uint32_t lo = chunk_bit0 + (uint32_t)tid * kSubBits;
  BitReader rd; rd.base = reinterpret_cast<const uint4 *>(d_us + im.raw_off);

  if (FIRST) {
    SubResult R;
    SubState from; from.p = S.us_bits; from.blk = 0;
    if (lo < S.us_bits) {
      // start at the last segment start at or before lo if it is within reach (exact), else guess
      const int k = seg_find(seg, 0, im.nseg - 1, lo);
      const uint32_t sb = seg[k];
      from.p = (lo >= (uint32_t)kOverlapBits && lo - (uint32_t)kOverlapBits > sb) ? lo - (uint32_t)kOverlapBits : sb;
    }
    decode_span<false>(S, luts, rd, seg, im.nseg, from, lo, lo + kSubBits, R, nullptr);
    S.e[tid] = R.e; S.used[tid] = R.start; S.info[tid] = info_of(R);
  } else {
    S.e[tid] = sub_e[gsub]; S.used[tid] = sub_used[gsub]; S.info[tid] = sub_info[gsub];
    if (tid == 0) S.cta_in = bnd_in[cta - 1];
  }
  __syncthreads();
  entropy_repair(S, luts, rd, seg, im.nseg, chunk_bit0, !FIRST);
  const SubState e = S.e[tid];
  sub_e[gsub] = e; sub_used[gsub] = S.used[tid];
  const SubInfo mine = S.info[tid];
  sub_info[gsub] = mine;
  Seg5 total;
  seg5_cta_excl(seg5_from(mine), &total);
  if (tid == kCT - 1) {
    if (!FIRST) {
      const SubState old = bnd_in[cta];
      if (old.p != e.p || old.blk != e.blk) atomicOr(changed_flag, 1u);
    }
    bnd_out[cta] = e;
  }
  if (tid == 0) cta_tot[cta] = seg5_to(total);
}

// ---- scan over CTA totals: one warp per image ----------------------------------------------
B2_GLOBAL void k_entropy_scan(const ImageDev *__restrict__ imgs, int nimg, const SubInfo *__restrict__ cta_tot,
                              SubInfo *__restrict__ cta_pre) {
  const int i = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (i >= nimg) return;
  const int c0 = imgs[i].cta_base, nc = imgs[i].ncta;
  Seg5 carry = seg5_zero();
  for (int b = 0; b < nc; b += 32) {
    const int c = b + lane;
    Seg5 v = seg5_zero();
    if (c < nc) v = seg5_from(cta_tot[c0 + c]);
    const Seg5 inc = seg5_warp_incl(v);
    Seg5 prev = seg5_shfl_up(inc, 1);
    if (lane == 0) prev = seg5_zero();
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
  B2_DYN_SMEM(uint32_t, luts);
  B2_SHARED EntropyShared S;
  uint4 *stage4 = reinterpret_cast<uint4 *>(luts + (size_t)nslots_max * kLutSize);
  const int tid = threadIdx.x;
  ImageDev im; uint32_t chunk_bit0;
  if (!entropy_prologue(S, luts, imgs, cta_base, nimg, tsets, out, im, chunk_bit0)) return;
  const int cta = (int)blockIdx.x, local_cta = cta - im.cta_base;
  const int gsub = im.sub_base + local_cta * kCT + tid;
  const uint32_t *seg = seg_start + im.seg_base;
  const uint32_t lo = chunk_bit0 + (uint32_t)tid * kSubBits;

  // first output block and DC predictors of this sub
  Seg5 total;
  Seg5 pre = seg5_cta_excl(seg5_from(sub_info[gsub]), &total);
  pre = seg5_combine(seg5_from(cta_pre[cta]), pre);

  SubState start;
  if (local_cta == 0 && tid == 0) { start.p = 0; start.blk = 0; }
  else start = sub_e[gsub - 1];

  // staging: chunk k (8 coefficients) of thread t at stage4[k*kCT + t]
#pragma unroll
  for (int k = 0; k < 8; ++k) stage4[k * kCT + tid] = make_uint4(0, 0, 0, 0);

  WriteCtx W;
  W.stage4 = stage4 + tid;
  W.obase = coef + im.coef_base * 64;
  W.seg_blocks = im.restart_interval * im.bpm;                       // 0 when there is one segment
  W.nseg = im.nseg; W.nblocks = im.nblocks;
  W.err = 0;
  if (lo < S.us_bits && start.p < lo + kSubBits && start.p < S.us_bits) {
    const int k = seg_find(seg, 0, im.nseg - 1, start.p);
    W.bidx = k * W.seg_blocks + pre.nb;
    W.limit = (k + 1 < im.nseg) ? (k + 1) * W.seg_blocks : im.nblocks;
    W.pred[0] = pre.d0; W.pred[1] = pre.d1; W.pred[2] = pre.d2;
    BitReader rd; rd.base = reinterpret_cast<const uint4 *>(d_us + im.raw_off);
    SubResult R;
    decode_span<true>(S, luts, rd, seg, im.nseg, start, lo, lo + kSubBits, R, &W);
  }
  if (W.err) atomicOr(&out[S.img].status, W.err);
}

}  // namespace b2
