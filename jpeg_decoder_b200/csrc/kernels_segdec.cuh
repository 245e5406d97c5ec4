// kernels_segdec.cuh -- stage E, segment-parallel path: one lane per restart segment (or per
// whole scan when the file has no RSTn), decoding the unstuffed scan into the final coefficient
// blocks in ONE pass.  Replaces, for batches that offer thousands of independently decodable
// segments, BitStream (file.hpp:122-165) and Image::decodeHuffman / decodeMCUComponent /
// getNextSymbol (jpeg.cpp:300-446); stage U (kernels_unstuff.cuh) has already done what
// JPEGFile::readImageData (file.hpp:59-104) does and found where every segment starts.
//
// Why this shape: a segment starts byte-aligned with the DC predictors reset (jpeg.cpp:419-425)
// and its first block index is known a priori (k * RI * blocks-per-MCU), so nothing has to be
// synchronised or prefix-summed -- the self-synchronising path (kernels_entropy.cuh) pays three
// passes over the stream for that.  The price is one long dependent chain per lane and, with a few
// thousand segments per batch, only about one warp per SM sub-partition: the kernel is bound by the
// LATENCY of that chain (position -> window -> table entry -> position), not by issue slots or memory.
// Two things follow.
//
// (1) Every step is the same for every lane.  Events that are rare for one lane (bit-buffer refill,
//     end of block) happen in almost every step for SOME lane of a warp, and a branch then costs all 32:
//   * no bit buffer: the 32-bit window is re-read each step from the bit position, out of a
//     per-lane ring of the next 32 stream words in shared memory laid out [word][lane] (bank =
//     lane: conflict-free whatever the lanes' positions);
//   * the ring is topped up by cp.async (LDGSTS, no register in between) under a per-lane
//     PREDICATE once every four steps -- four steps consume at most 108 bits, one top-up brings 128;
//   * DC and AC tables are looked up together and selected afterwards.
//
// (2) Only what the chain needs stays on the chain.  Each DECODER warp is paired with a WRITER warp of
//     the same CTA.  The decoder walks the codes -- window, table entry, advance, block/table
//     bookkeeping -- and pushes (window, entry) per lane and step into a shared-memory FIFO; the writer
//     pops them and does everything that merely CONSUMES a symbol: magnitude extension (jpeg.cpp:335-343,
//     388-397), DC prediction (jpeg.cpp:344-345), zig-zag placement, the stores, zero-filling the next
//     block, the range checks.  The two warps run on the idle issue slots of their sub-partitions in
//     parallel, so the dependent chain per symbol is roughly halved.  Hand-over is a pair of counters
//     (st.release / ld.acquire in shared memory), one update per round of four steps.
#pragma once
#include "platform.h"
#include "dev_types.h"
#include "kernels_unstuff.cuh"
#include "kernels_huff.cuh"

namespace b2 {

#ifndef B2_SEG_PAIRS
#define B2_SEG_PAIRS 2
#endif
#ifndef B2_SEG_FIFO
#define B2_SEG_FIFO 8
#endif
constexpr int kSegPairs = B2_SEG_PAIRS;       // decoder/writer warp pairs per CTA
constexpr int kSegThreads = kSegPairs * 64;   // threads per CTA of k_entropy_seg: warps [0,kSegPairs) decode, the rest write
constexpr int kSegUnits = kSegPairs * 32;     // segments per CTA
constexpr uint32_t kSegNoUnit = 0xffffffffu;
constexpr int kSegWords = 32;                 // stream words per lane in the window ring
constexpr int kSegFifo = B2_SEG_FIFO;         // rounds (of four steps) the FIFO between a decoder and its writer holds
constexpr int kSegRoundWords = 4 * 2 * 32;    // one round: 4 steps x (window, entry) x 32 lanes
constexpr uint32_t kSegFin = 0x80000000u;     // set in the producer counter with the last round

B2_HD size_t segdec_smem_bytes() {
  return sizeof(uint32_t) * ((size_t)kHTabsPerSet * kHTab + (size_t)kSegPairs * (kSegWords * 32 + kSegFifo * kSegRoundWords + 32));
}

B2_DEVICE int seg_extend(uint32_t w, uint32_t len, uint32_t size) {    // jpeg.cpp:335-343, 388-397; size >= 1
  const uint32_t v = (w << len) >> (32u - size);
  return (v >> (size - 1u)) ? (int)v : (int)v - (int)((1u << size) - 1u);
}

// Coefficients go straight to global memory: at the start of a block its 128-byte line is zeroed with
// eight 128-bit stores (MCU[] is value-initialised, jpeg.cpp:407), then every non-zero coefficient is one
// 2-byte store into that line (same thread, same addresses: program order holds; L2 merges the line).
// Chunk c of block b lives at chunk c ^ (b & 7) (kernels_pixel.cuh).
B2_DEVICE void seg_zero_block(int16_t *__restrict__ coef, int64_t gb) {
  uint4 *o = reinterpret_cast<uint4 *>(coef) + gb * 8;
#pragma unroll
  for (int q = 0; q < 8; ++q) o[q] = make_uint4(0, 0, 0, 0);
}

// What both warps of a pair know about a lane's segment before the first symbol.
struct SegLane {
  bool live;                 // there is something to decode
  int img, bidx, bend;       // image, first block of the segment, first block past it
  int bpm, nluma;
  int64_t gb;                // global index of the first block in the coefficient buffer
};
B2_DEVICE SegLane seg_lane(const uint2 unit, const ImageDev *__restrict__ imgs, const ImageOut *__restrict__ out) {
  SegLane L;
  L.live = false; L.img = 0; L.bidx = L.bend = 0; L.bpm = 1; L.nluma = 1; L.gb = 0;
  if (unit.x == kSegNoUnit) return L;
  L.img = (int)unit.x;
  const ImageDev &im = imgs[L.img];
  if (out[L.img].status & (ST_BAD_MARKER | ST_NO_EOI | ST_BAD_RST_COUNT)) return L;   // nothing to decode
  const int k = (int)unit.y;
  L.bpm = im.bpm; L.nluma = im.ncomp == 1 ? 1 : im.hs * im.vs;
  const int seg_blocks = im.restart_interval * L.bpm;                                  // 0 when there is one segment
  L.bidx = k * seg_blocks;
  L.bend = (k + 1 < im.nseg) ? L.bidx + seg_blocks : im.nblocks;
  L.gb = im.coef_base + L.bidx;
  L.live = L.bidx < L.bend;
  return L;
}

// units[u] = (image, segment); the units of one CTA share a table set (the planner pads each
// table-set group to a multiple of kSegUnits with kSegNoUnit).  d_us / seg_start: stage U's output.
B2_GLOBAL void __launch_bounds__(kSegThreads)
k_entropy_seg(const uint2 *__restrict__ units, const ImageDev *__restrict__ imgs, const TableSet *__restrict__ tsets,
              const uint32_t *__restrict__ luts, const uint32_t *__restrict__ seg_start,
              const uint8_t *__restrict__ d_us, ImageOut *__restrict__ out, int16_t *__restrict__ coef) {
  B2_DYN_SMEM(uint32_t, lut);                                   // [8][kHTab]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool decoder = warp < kSegPairs;
  const int pair = decoder ? warp : warp - kSegPairs;
  uint32_t *pbase = lut + kHTabsPerSet * kHTab + pair * (kSegWords * 32 + kSegFifo * kSegRoundWords + 32);
  uint32_t *ringl = pbase + lane;                               // window ring: word j of this lane at ringl[(j & 31) * 32]
  uint32_t *fifo = pbase + kSegWords * 32;                      // [round][step][window|entry][lane]
  uint32_t *prod = fifo + kSegFifo * kSegRoundWords;            // rounds produced (| kSegFin), then rounds consumed
  uint32_t *cons = prod + 1;
  {
    // the CTA's table set: that of its first unit (never padding)
    const int set = imgs[units[(size_t)blockIdx.x * kSegUnits].x].tset;
    const uint4 *src = reinterpret_cast<const uint4 *>(luts + (size_t)set * kHTabsPerSet * kHTab);
    uint4 *dst = reinterpret_cast<uint4 *>(lut);
    for (int t = 0; t < kHTabsPerSet; ++t) {
      const bool used = t < 4 ? tsets[set].dc[t].set : tsets[set].ac[t - 4].set;
      if (!used) continue;
      for (int k = tid; k < kHTab / 4; k += kSegThreads) dst[t * (kHTab / 4) + k] = b2_ld_nc_v4(src + t * (kHTab / 4) + k);
    }
    if (decoder && lane == 0) { *prod = 0; *cons = 0; }
  }
  __syncthreads();

  const uint2 unit = units[(size_t)blockIdx.x * kSegUnits + pair * 32 + lane];
  const SegLane L = seg_lane(unit, imgs, out);
  int left = L.live ? L.bend - L.bidx : 0;       // blocks of the segment still to finish
  int blk = 0, comp = 0;                         // block inside the MCU, its component
  uint32_t z = 0, err = 0;
  // the inner loops address shared memory through 32-bit window addresses (register + immediate)
  const uint32_t fifo_a = b2_dsm_addr(fifo) + 4u * (uint32_t)lane;

  if (decoder) {
    // =========================== the chain: codes only ===========================
    const uint32_t lut_a = b2_dsm_addr(lut), ring_a = b2_dsm_addr(ringl);
    uint32_t dca0 = lut_a, dca1 = lut_a, dca2 = lut_a, aca0 = lut_a, aca1 = lut_a, aca2 = lut_a;   // table addresses per component
    uint32_t p = 0, pend = 0, rw = 0;    // bit position / end of the segment / next word to fetch (all in the unstuffed scan)
    const uint32_t *us = reinterpret_cast<const uint32_t *>(d_us);
    bool done = !L.live;
    if (L.live) {
      const ImageDev &im = imgs[L.img];
      // DC table td -> slot td, AC table ta -> slot 4+ta
      dca0 += (uint32_t)(im.td[0] & 3) * (kHTab * 4); dca1 += (uint32_t)(im.td[1] & 3) * (kHTab * 4); dca2 += (uint32_t)(im.td[2] & 3) * (kHTab * 4);
      aca0 += (uint32_t)(4 + (im.ta[0] & 3)) * (kHTab * 4); aca1 += (uint32_t)(4 + (im.ta[1] & 3)) * (kHTab * 4);
      aca2 += (uint32_t)(4 + (im.ta[2] & 3)) * (kHTab * 4);
      us = reinterpret_cast<const uint32_t *>(d_us + im.raw_off);
      p = seg_start[im.seg_base + (int)unit.y];
      pend = seg_start[im.seg_base + (int)unit.y + 1];
      rw = (p >> 5) & ~3u;
      for (int i = 0; i < kSegWords; ++i) b2_cp_async4(ringl + ((rw + i) & (kSegWords - 1)) * 32, us + rw + i);
      rw += kSegWords;
    }
    b2_cp_async_commit();
    b2_cp_async_wait<0>();
    uint32_t dca = dca0, aca = aca0;
    uint32_t rounds = 0;
    for (;;) {
      // room in the FIFO: the writer has finished round (rounds - kSegFifo)
      while ((int)(rounds - b2_ld_acquire_smem(cons)) >= kSegFifo) {}
      // ---- top-up: once per four steps, the same instructions for every lane, predicated per lane ----
      if (!done && rw - (p >> 5) <= (uint32_t)(kSegWords - 4)) {
#pragma unroll
        for (int i = 0; i < 4; ++i) b2_cp_async4(ringl + ((rw + i) & (kSegWords - 1)) * 32, us + rw + i);
        rw += 4;
        b2_cp_async_commit();        // groups are per thread: only real top-ups count
      }
      // the window (words wi, wi+1) is more than 24 words behind rw, i.e. older than the 5 newest groups:
      // a copy has ~20 words (~140 steps) to land before it is needed
      b2_cp_async_wait<5>();
      const uint32_t slot_a = fifo_a + (rounds & (uint32_t)(kSegFifo - 1)) * (kSegRoundWords * 4);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        // a finished lane keeps stepping in place: its entry is forced to 0 (no advance, nothing for the writer)
        const uint32_t wi = p >> 5;
        const uint32_t w0 = b2_lds(ring_a + ((wi & 31u) << 7)), w1 = b2_lds(ring_a + (((wi + 1u) & 31u) << 7));
        const uint32_t w = __funnelshift_l(w1, w0, p);
        const uint32_t i4 = (w >> (30 - kHLutBits)) & (uint32_t)((kHL1 - 1) * 4);
        const uint32_t e_ac = b2_lds(aca + i4), e_dc = b2_lds(dca + i4);            // both at once: the select is off the critical path
        uint32_t e = z == 0 ? e_dc : e_ac;
        e = done ? 0u : e;
        if (e >= kHOverflow) {                                                       // rare: long code, or an error entry
          const bool is_dc = z == 0;
          if (e & kHLong) e = b2_lds((is_dc ? dca : aca) + 4u * ((e & 0xffffu) + ((w >> (32 - kHLutBits - 6)) & 63u)));
          if (e & kHOverflow) {
            const ImageDev &im = imgs[L.img];
            e = huff_slow(is_dc ? tsets[im.tset].dc[im.td[comp] & 3] : tsets[im.tset].ac[im.ta[comp] & 3], w >> 16, is_dc);
          }
          if (e & kHBadCode) err |= ST_BAD_CODE;
          if (e & kHBadSize) err |= is_dc ? ST_BAD_DC_LEN : ST_BAD_AC_LEN;
        }
        b2_sts(slot_a + (uint32_t)u * 256u, w);
        b2_sts(slot_a + (uint32_t)u * 256u + 128u, e);
        p += e & 31u;
        z += (e >> 10) & 127u;
        if (z >= 64u) {
          // ---- block end: the next block's tables; the last block ends the lane ----
          z = 0;
          if (--left == 0) {
            done = true;
            if (p > pend) err |= ST_OVERRUN;                                        // the last block ran past the segment's end (file.hpp:36-40)
            else if (pend - p > 7u) err |= ST_BLOCK_COUNT;                          // more than padding is left: the segment holds more blocks
          } else {
            blk = blk + 1 == L.bpm ? 0 : blk + 1;
            comp = blk < L.nluma ? 0 : blk - L.nluma + 1;
            dca = comp == 0 ? dca0 : (comp == 1 ? dca1 : dca2);
            aca = comp == 0 ? aca0 : (comp == 1 ? aca1 : aca2);
          }
        }
      }
      const bool fin = __all_sync(0xffffffffu, done);   // also orders the lanes' FIFO stores before lane 0's release
      ++rounds;
      if (lane == 0) b2_st_release_smem(prod, rounds | (fin ? kSegFin : 0u));
      if (fin) break;
    }
  } else {
    // =========================== off the chain: values and stores ===========================
    int16_t *bp = coef + L.gb * 64;                    // the block being filled
    uint32_t key = ((uint32_t)L.gb & 7u) << 3;         // its chunk swizzle: coefficient pos lives at pos ^ key
    int pcur = 0, pred0 = 0, pred1 = 0, pred2 = 0;     // pcur: predictor of the current block's component
    uint32_t dcbad = 0;                                // != 0 once an absolute DC left the int16 range
    if (L.live) seg_zero_block(bp, 0);
    uint32_t r = 0;
    for (;;) {
      uint32_t pr;
      do pr = b2_ld_acquire_smem(prod); while ((pr & ~kSegFin) <= r && !(pr & kSegFin));
      if ((pr & ~kSegFin) <= r) break;                                              // the last round has been consumed
      const uint32_t slot_a = fifo_a + (r & (uint32_t)(kSegFifo - 1)) * (kSegRoundWords * 4);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t w = b2_lds(slot_a + (uint32_t)u * 256u), e = b2_lds(slot_a + (uint32_t)u * 256u + 128u);
        if (e != 0u) {
          const uint32_t adv = e & 31u, len = (e >> 5) & 31u, zinc = (e >> 10) & 127u;
          const uint32_t size = adv - len;
          // EXTEND (jpeg.cpp:335-343, 388-397), branch-free: v < 2^(size-1) <=> 2v <= 2^size - 1
          const uint32_t v = b2_shr(w << len, 32u - size), t = b2_shl(1u, size) - 1u;
          const int val = (int)v - (v + v > t ? 0 : (int)t);
          const bool is_dc = z == 0;
          const uint32_t zn = z + zinc;
          pcur += is_dc ? val : 0;                                                  // jpeg.cpp:344-345
          const int o = is_dc ? pcur : val;
          const uint32_t pos = is_dc ? 0u : zn - 1u;                                // z + run
          if ((is_dc || size != 0u) && pos < 64u) bp[pos ^ key] = (int16_t)o;
          z = zn;
          if (z >= 64u) {
            // ---- block end: EOB, or the index ran to 64.  jpeg.cpp:372-376: a run that reaches 64 is an error ----
            if (!(e & kHEob) && !(size != 0u && z == 64u)) err |= ST_BAD_RUN;
            dcbad |= (uint32_t)(pcur + 32768) >> 16;
            z = 0;
            if (--left != 0) {
              pred0 = comp == 0 ? pcur : pred0; pred1 = comp == 1 ? pcur : pred1; pred2 = comp == 2 ? pcur : pred2;
              blk = blk + 1 == L.bpm ? 0 : blk + 1;
              comp = blk < L.nluma ? 0 : blk - L.nluma + 1;
              pcur = comp == 0 ? pred0 : (comp == 1 ? pred1 : pred2);
              bp += 64; key = (key + 8u) & 0x38u;
              seg_zero_block(bp, 0);
            }
          }
        }
      }
      __syncwarp();                                     // every lane has read the round before it is handed back
      ++r;
      if (lane == 0) b2_st_release_smem(cons, r);
    }
    if (dcbad) err |= ST_DC_RANGE;
  }
  if (err) atomicOr(&out[L.img].status, err);
}

}  // namespace b2
