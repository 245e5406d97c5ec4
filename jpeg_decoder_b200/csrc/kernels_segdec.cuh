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
// (2) Only what the chain needs stays on the chain.  A segment is worked on by FOUR warps of the same CTA
//     (32 segments each):
//       decoder   walks the codes -- window, table entry, advance, block/table bookkeeping -- and pushes
//                 (window, entry | zig-zag index, block number) per lane and step into a shared-memory FIFO;
//       writer A/B take the even / the odd rounds of the FIFO and place the AC coefficients: magnitude
//                 extension (jpeg.cpp:388-397), zig-zag position, one 2-byte store.  They keep no state
//                 between symbols -- everything they need is in the FIFO entry -- so two of them (or more)
//                 work in parallel;
//       DC warp   sees every round but only the DC symbols: extension, prediction per component
//                 (jpeg.cpp:335-345), store, range check.
//     The warps run on the idle issue slots of their sub-partitions in parallel; the dependent chain per
//     symbol is what the decoder alone executes.  Hand-over is a producer counter and one consumer counter
//     per warp (st.release / ld.acquire in shared memory), one update per round of four steps.
//     The coefficient buffer is zero-filled beforehand (a memset on a side stream, pipeline_run.h): MCU[] is
//     value-initialised in the reference (jpeg.cpp:407) and the warps only store what is not zero.
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
#define B2_SEG_FIFO 4
#endif
constexpr int kSegPairs = B2_SEG_PAIRS;       // groups of four warps (decoder, writer A, writer B, DC) per CTA
constexpr int kSegRoles = 4;
constexpr int kSegThreads = kSegPairs * kSegRoles * 32;   // threads per CTA of k_entropy_seg
constexpr int kSegUnits = kSegPairs * 32;     // segments per CTA
constexpr uint32_t kSegNoUnit = 0xffffffffu;
constexpr int kSegWords = 32;                 // stream words per lane in the window ring
constexpr int kSegFifo = B2_SEG_FIFO;         // rounds the FIFO holds
constexpr int kSegSteps = 8;                  // symbols per lane and round: the per-round work (top-up, hand-over) is paid once per 8 steps
constexpr int kSegStepBytes = 3 * 128;        // one step: (window, entry, block number) x 32 lanes
constexpr int kSegRoundBytes = kSegSteps * kSegStepBytes;
constexpr uint32_t kSegFin = 0x80000000u;     // set in the producer counter with the last round
#ifndef B2_SEG_UNROLL
#define B2_SEG_UNROLL 2
#endif
#ifndef B2_SEG_POLL_NS
#define B2_SEG_POLL_NS 200
#endif
constexpr int kSegUnroll = B2_SEG_UNROLL;      // how far the decoder's step loop is unrolled (instruction-cache footprint)
constexpr unsigned kSegPollNs = B2_SEG_POLL_NS;   // a warp that finds nothing to do sleeps this long before it looks again
// third FIFO word of a step: zig-zag index before the symbol | two symbols taken << 6 | block number << 7
constexpr uint32_t kSegPairBit = 1u << 6;
constexpr int kSegNbShift = 7;
// per group: window ring | FIFO | per-lane table addresses by block-in-MCU | counters
constexpr int kSegGroupWords = kSegWords * 32 + kSegFifo * kSegRoundBytes / 4 + kMaxBpm * 32 + 32;

B2_HD size_t segdec_smem_bytes() {
  return sizeof(uint32_t) * ((size_t)kHTabsPerSet * kHTab + (size_t)kSegPairs * kSegGroupWords);
}

// What every warp of a group knows about a lane's segment before the first symbol.
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
  const int pair = warp % kSegPairs, role = warp / kSegPairs;    // role 0 decoder, 1/2 AC writers, 3 DC warp
  uint32_t *pbase = lut + kHTabsPerSet * kHTab + pair * kSegGroupWords;
  uint32_t *ringl = pbase + lane;                               // window ring: word j of this lane at ringl[(j & 31) * 32]
  uint32_t *fifo = pbase + kSegWords * 32;                      // [round][step][window|entry|block][lane]
  uint32_t *tb = fifo + kSegFifo * kSegRoundBytes / 4;          // [block-in-MCU][lane]: DC table offset | AC table offset << 16
  uint32_t *ctr = tb + kMaxBpm * 32;                            // [0] rounds produced (| kSegFin), [1..3] next round of writer A, B, DC warp
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
    if (role == 0 && lane == 0) { ctr[0] = 0; ctr[1] = 0; ctr[2] = 1; ctr[3] = 0; }
  }
  __syncthreads();

  const uint2 unit = units[(size_t)blockIdx.x * kSegUnits + pair * 32 + lane];
  const SegLane L = seg_lane(unit, imgs, out);
  uint32_t err = 0;
  // the inner loops address shared memory through 32-bit window addresses (register + immediate)
  const uint32_t fifo_a = b2_dsm_addr(fifo) + 4u * (uint32_t)lane, lut_a = b2_dsm_addr(lut);
  int16_t *const base = coef + L.gb * 64;          // block 0 of the segment; block n, coefficient pos: base[n*64 + (pos ^ key(n))]
  const uint32_t g0 = (uint32_t)L.gb & 7u;         // key(n) = ((g0 + n) & 7) << 3

  if (role == 0) {
    // =========================== the chain: codes only ===========================
    const uint32_t ring_a = b2_dsm_addr(ringl), tb_a = b2_dsm_addr(tb) + 4u * (uint32_t)lane;
    uint32_t p = 0, pend = 0, rw = 0;    // bit position / end of the segment / next word to fetch (all in the unstuffed scan)
    const uint32_t *us = reinterpret_cast<const uint32_t *>(d_us);
    const uint32_t total = L.live ? (uint32_t)(L.bend - L.bidx) : 0u;
    bool done = !L.live;
    if (L.live) {
      const ImageDev &im = imgs[L.img];
      for (int b = 0; b < L.bpm; ++b) {             // DC table td -> slot td, AC table ta -> slot 4+ta
        const int c = b < L.nluma ? 0 : b - L.nluma + 1;
        tb[b * 32 + lane] = (uint32_t)(im.td[c] & 3) * (kHTab * 4) | ((uint32_t)(4 + (im.ta[c] & 3)) * (kHTab * 4)) << 16;
      }
      us = reinterpret_cast<const uint32_t *>(d_us + im.raw_off);
      p = seg_start[im.seg_base + (int)unit.y];
      pend = seg_start[im.seg_base + (int)unit.y + 1];
      rw = (p >> 5) & ~3u;
      for (int i = 0; i < kSegWords; ++i) b2_cp_async4(ringl + ((rw + i) & (kSegWords - 1)) * 32, us + rw + i);
      rw += kSegWords;
    } else tb[lane] = 0;
    b2_cp_async_commit();
    b2_cp_async_wait<0>();
    const uint32_t tb_end = tb_a + (uint32_t)L.bpm * 128u;
    uint32_t tba = tb_a;                             // this lane's table entry of the current block-in-MCU
    uint32_t z = 0, nb = 0;                          // zig-zag index, blocks finished
    uint32_t rounds = 0;
    for (;;) {
      // the consumers' counters are read now and looked at after the round: room for the NEXT round
      const uint32_t ca = b2_ld_acquire_smem(ctr + 1), cb = b2_ld_acquire_smem(ctr + 2), cc = b2_ld_acquire_smem(ctr + 3);
      // ---- top-up: once per round, the same instructions for every lane, predicated per lane (no branch).
      //      A round consumes at most 8 x 27 bits < 7 words; a top-up brings 8 ----
      const bool need = !done && rw - (p >> 5) <= (uint32_t)(kSegWords - 8);
#pragma unroll
      for (int i = 0; i < 8; ++i) b2_cp_async4_if(need, ring_a + (((rw + i) & (kSegWords - 1)) << 7), us + rw + i);
      rw += need ? 8u : 0u;
      b2_cp_async_commit();
      // rw is at least 25 words ahead of the window and the round needs at most 8 more: everything but the two
      // newest groups must have landed -- copies issued two rounds (16 steps) ago
      b2_cp_async_wait<2>();
      const uint32_t slot_a = fifo_a + (rounds & (uint32_t)(kSegFifo - 1)) * kSegRoundBytes;
#pragma unroll kSegUnroll
      for (int u = 0; u < kSegSteps; ++u) {
        // a finished lane keeps stepping in place: its entry is forced to 0 (no advance, nothing for the writers)
        const uint32_t t = b2_lds(tba);                                              // table offsets of this block: DC | AC << 16
        const uint32_t cur = lut_a + (z == 0 ? (t & 0xffffu) : (t >> 16));           // chosen while the window is still on its way
        const uint32_t wi = p >> 5;
        const uint32_t w0 = b2_lds(ring_a + ((wi & 31u) << 7)), w1 = b2_lds(ring_a + (((wi + 1u) & 31u) << 7));
        const uint32_t w = __funnelshift_l(w1, w0, p);
        const uint32_t i4 = (w >> (30 - kHLutBits)) & (uint32_t)((kHL1 - 1) * 4);
        uint32_t e = b2_lds(cur + i4);
        e = done ? 0u : e;
        if (e >= kHOverflow) {                                                       // rare: long code, or an error entry
          const bool is_dc = z == 0;
          if (e & kHLong) e = b2_lds(cur + 4u * ((e & 0xffffu) + ((w >> (32 - kHLutBits - 6)) & 63u)));
          if (e & kHOverflow) {
            const ImageDev &im = imgs[L.img];
            const int blk = (int)((tba - tb_a) >> 7), comp = blk < L.nluma ? 0 : blk - L.nluma + 1;
            e = huff_slow(is_dc ? tsets[im.tset].dc[im.td[comp] & 3] : tsets[im.tset].ac[im.ta[comp] & 3], w >> 16, is_dc);
          }
          if (e & kHBadCode) err |= ST_BAD_CODE;
          if (e & kHBadSize) err |= is_dc ? ST_BAD_DC_LEN : ST_BAD_AC_LEN;
        }
        b2_sts(slot_a + (uint32_t)u * kSegStepBytes, w);
        // an AC entry may describe TWO symbols (kernels_huff.cuh); both are taken unless the first ends the block
        const uint32_t z1 = huff_zinc(e), a2 = huff_adv2(e);
        const bool pair = a2 != 0u && z + z1 < 64u;
        b2_sts(slot_a + (uint32_t)u * kSegStepBytes + 128u, e);
        b2_sts(slot_a + (uint32_t)u * kSegStepBytes + 256u, z | (pair ? kSegPairBit : 0u) | (nb << kSegNbShift));
        p += huff_adv(e) + (pair ? a2 : 0u);
        z += z1 + (pair ? huff_zinc2(e) : 0u);
        // ---- block end without a branch: some lane ends a block in four steps out of five ----
        const bool end = z >= 64u;
        const uint32_t tbn = tba + 128u == tb_end ? tb_a : tba + 128u;
        z = end ? 0u : z;
        tba = end ? tbn : tba;
        nb += end ? 1u : 0u;
        if (end && nb == total) {                                                    // once per lane: the segment is finished
          done = true;
          if (p > pend) err |= ST_OVERRUN;                                          // the last block ran past the segment's end (file.hpp:36-40)
          else if (pend - p > 7u) err |= ST_BLOCK_COUNT;                            // more than padding is left: the segment holds more blocks
        }
      }
      const bool fin = __all_sync(0xffffffffu, done);   // also orders the lanes' FIFO stores before lane 0's release
      ++rounds;
      if (lane == 0) b2_st_release_smem(ctr, rounds | (fin ? kSegFin : 0u));
      if (fin) break;
      // room in the FIFO for the next round: every consumer is past round (rounds - kSegFifo)
      uint32_t m = ca < cb ? (ca < cc ? ca : cc) : (cb < cc ? cb : cc);
      while ((int)(rounds - m) >= kSegFifo) {
        b2_backoff(kSegPollNs);
        const uint32_t a = b2_ld_acquire_smem(ctr + 1), b = b2_ld_acquire_smem(ctr + 2), c = b2_ld_acquire_smem(ctr + 3);
        m = a < b ? (a < c ? a : c) : (b < c ? b : c);
      }
    }
  } else if (role < 3) {
    // =========================== AC coefficients: stateless, every other round ===========================
    uint32_t r = (uint32_t)(role - 1);
    for (;;) {
      uint32_t pr;
      while (pr = b2_ld_acquire_smem(ctr), (pr & ~kSegFin) <= r && !(pr & kSegFin)) b2_backoff(kSegPollNs);
      if ((pr & ~kSegFin) <= r) break;                                              // no such round: the decoder has finished
      const uint32_t slot_a = fifo_a + (r & (uint32_t)(kSegFifo - 1)) * kSegRoundBytes;
      uint32_t fw[kSegSteps], fe[kSegSteps], fn[kSegSteps];
#pragma unroll
      for (int u = 0; u < kSegSteps; ++u) {
        fw[u] = b2_lds(slot_a + (uint32_t)u * kSegStepBytes); fe[u] = b2_lds(slot_a + (uint32_t)u * kSegStepBytes + 128u);
        fn[u] = b2_lds(slot_a + (uint32_t)u * kSegStepBytes + 256u);
      }
      __syncwarp();                                     // every lane has read the round before it is handed back
      if (lane == 0) b2_st_release_smem(ctr + role, r + 2u);
#pragma unroll
      for (int u = 0; u < kSegSteps; ++u) {
        const uint32_t e = fe[u], z = fn[u] & 63u, nblk = fn[u] >> kSegNbShift;
        if (z != 0u) {                                  // AC (entry 0 and DC symbols carry z = 0)
          int16_t *const bp = base + nblk * 64u;
          const uint32_t key = ((g0 + nblk) & 7u) << 3;
          const uint32_t adv = huff_adv(e), len = huff_len(e), zn = z + huff_zinc(e);
          const uint32_t size = adv - len;
          // EXTEND (jpeg.cpp:388-397), branch-free: v < 2^(size-1) <=> 2v <= 2^size - 1
          const uint32_t v = b2_shr(fw[u] << len, 32u - size), t = b2_shl(1u, size) - 1u;
          const int val = (int)v - (v + v > t ? 0 : (int)t);
          const uint32_t pos = zn - 1u;                                             // z + run
          if (size != 0u && pos < 64u) bp[pos ^ key] = (int16_t)val;
          // EOB, or the index ran to 64; a run that reaches past 63 is an error (jpeg.cpp:372-376)
          if (zn >= 64u && !(e & kHEob) && !(size != 0u && zn == 64u)) err |= ST_BAD_RUN;
          if (fn[u] & kSegPairBit) {
            // the second symbol of the entry: its bits follow the first one's; its details come from the same table
            const uint32_t w2 = fw[u] << adv;
            const uint32_t e2 = b2_lds(lut_a + (4u + huff_tslot(e)) * (uint32_t)(kHTab * 4) + ((w2 >> (30 - kHLutBits)) & (uint32_t)((kHL1 - 1) * 4)));
            const uint32_t adv2 = huff_adv(e2), len2 = huff_len(e2), zn2 = zn + huff_zinc(e2), size2 = adv2 - len2;
            const uint32_t v2 = b2_shr(w2 << len2, 32u - size2), t2 = b2_shl(1u, size2) - 1u;
            const int val2 = (int)v2 - (v2 + v2 > t2 ? 0 : (int)t2);
            const uint32_t pos2 = zn2 - 1u;
            if (size2 != 0u && pos2 < 64u) bp[pos2 ^ key] = (int16_t)val2;
            if (zn2 >= 64u && !(e2 & kHEob) && !(size2 != 0u && zn2 == 64u)) err |= ST_BAD_RUN;
          }
        }
      }
      r += 2u;
    }
  } else {
    // =========================== DC coefficients: prediction per component ===========================
    int pred0 = 0, pred1 = 0, pred2 = 0;
    int blk = 0;
    uint32_t dcbad = 0;                                // != 0 once an absolute DC left the int16 range
    uint32_t r = 0;
    for (;;) {
      uint32_t pr;
      while (pr = b2_ld_acquire_smem(ctr), (pr & ~kSegFin) <= r && !(pr & kSegFin)) b2_backoff(kSegPollNs);
      if ((pr & ~kSegFin) <= r) break;
      const uint32_t slot_a = fifo_a + (r & (uint32_t)(kSegFifo - 1)) * kSegRoundBytes;
      uint32_t fw[kSegSteps], fe[kSegSteps], fn[kSegSteps];
#pragma unroll
      for (int u = 0; u < kSegSteps; ++u) {
        fw[u] = b2_lds(slot_a + (uint32_t)u * kSegStepBytes); fe[u] = b2_lds(slot_a + (uint32_t)u * kSegStepBytes + 128u);
        fn[u] = b2_lds(slot_a + (uint32_t)u * kSegStepBytes + 256u);
      }
      __syncwarp();
      ++r;
      if (lane == 0) b2_st_release_smem(ctr + 3, r);
#pragma unroll
      for (int u = 0; u < kSegSteps; ++u) {
        const uint32_t e = fe[u];
        if (e != 0u && (fn[u] & 63u) == 0u) {                                       // a DC symbol opens a block
          const uint32_t adv = huff_adv(e), len = huff_len(e), size = adv - len, nblk = fn[u] >> kSegNbShift;
          const uint32_t v = b2_shr(fw[u] << len, 32u - size), t = b2_shl(1u, size) - 1u;
          const int diff = (int)v - (v + v > t ? 0 : (int)t);                       // jpeg.cpp:335-343
          const int comp = blk < L.nluma ? 0 : blk - L.nluma + 1;
          const int dc = (comp == 0 ? pred0 : (comp == 1 ? pred1 : pred2)) + diff;  // jpeg.cpp:344-345
          pred0 = comp == 0 ? dc : pred0; pred1 = comp == 1 ? dc : pred1; pred2 = comp == 2 ? dc : pred2;
          dcbad |= (uint32_t)(dc + 32768) >> 16;
          base[nblk * 64u + (((g0 + nblk) & 7u) << 3)] = (int16_t)dc;               // position 0 of chunk 0 ^ key
          blk = blk + 1 == L.bpm ? 0 : blk + 1;
        }
      }
    }
    if (dcbad) err |= ST_DC_RANGE;
  }
  if (err) atomicOr(&out[L.img].status, err);
}

}  // namespace b2
