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
#ifndef B2_SEG_ACW
#define B2_SEG_ACW 2
#endif
constexpr int kSegAcW = B2_SEG_ACW;           // AC writer warps per group: writer w takes the rounds r with r % kSegAcW == w
constexpr int kSegRoles = kSegAcW + 2;        // decoder, AC writers, DC warp
constexpr int kSegThreads = kSegPairs * kSegRoles * 32;   // threads per CTA of k_entropy_seg
constexpr int kSegUnits = kSegPairs * 32;     // segments per CTA
constexpr uint32_t kSegNoUnit = 0xffffffffu;
constexpr int kSegWords = 32;                 // stream words per lane in the window ring
constexpr int kSegFifo = B2_SEG_FIFO;         // rounds the FIFO holds
constexpr int kSegSteps = 8;                  // symbols per lane and round: the per-round work (top-up, hand-over) is paid once per 8 steps
constexpr int kSegStepBytes = 3 * 128;        // one step: (window, entry, block number) x 32 lanes
constexpr int kSegSlots = kSegSteps + 1;      // FIFO slots per round: the steps + one for symbols the first-level table does not resolve
constexpr int kSegRoundBytes = kSegSlots * kSegStepBytes;
constexpr uint32_t kSegFin = 0x80000000u;     // set in the producer counter with the last round
#ifndef B2_SEG_UNROLL
#define B2_SEG_UNROLL 2
#endif
#ifndef B2_SEG_POLL_NS
#define B2_SEG_POLL_NS 200
#endif
#ifndef B2_SEG_DIAG
#define B2_SEG_DIAG 0     // measurements only: 1 = consumers acknowledge rounds without working on them, 2 = decoder alone (no FIFO)
#endif
#ifndef B2_SEG_ROLEMAP
#define B2_SEG_ROLEMAP 0  // 0: warp = role * pairs + pair; 1: roles rotated per pair so that the decoders sit on different sub-partitions
#endif
#if B2_SEG_DIAG == 3       // 3 = everything but the AC stores (the condition keeps the arithmetic alive)
#define B2_SEG_NOSTORE(w) && (w) == 0x12345678u
#else
#define B2_SEG_NOSTORE(w)
#endif
constexpr int kSegUnroll = B2_SEG_UNROLL;      // how far the decoder's step loop is unrolled (instruction-cache footprint)
constexpr unsigned kSegPollNs = B2_SEG_POLL_NS;   // a warp that finds nothing to do sleeps this long before it looks again
// third FIFO word of a step: zig-zag index before the symbol | two symbols taken << 6 | block number << 7
constexpr uint32_t kSegPairBit = 1u << 6;
constexpr int kSegNbShift = 7;
// per group: window ring | FIFO | per-lane table addresses by block-in-MCU | counters
constexpr int kSegGroupWords = (kSegWords + 1) * 32 + kSegFifo * kSegRoundBytes / 4 + kMaxBpm * 32 + 32;

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
#if B2_SEG_ROLEMAP == 1
  const int pair = warp / kSegRoles, role = (warp + pair) % kSegRoles;   // sub-partition = warp % 4: pair 0's decoder on 0, pair 1's on 3
#elif B2_SEG_ROLEMAP == 2
  const int pair = warp / kSegRoles, role = (warp + pair + (blockIdx.x >= 148u ? 2 : 0)) % kSegRoles;   // second CTA of an SM: decoders on 2 and 1
#else
  const int pair = warp % kSegPairs, role = warp / kSegPairs;    // role 0 decoder, 1/2 AC writers, 3 DC warp
#endif
  uint32_t *pbase = lut + kHTabsPerSet * kHTab + pair * kSegGroupWords;
  uint32_t *ringl = pbase + lane;                               // window ring: word j of this lane at ringl[(j & 31) * 32]
  uint32_t *fifo = pbase + (kSegWords + 1) * 32;                      // [round][step][window|entry|block][lane]
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
    if (role == 0 && lane <= kSegAcW + 1) ctr[lane] = lane >= 1 && lane <= kSegAcW ? (uint32_t)(lane - 1) : 0u;
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
    // The eight steps of a round are ONE basic block: no branch, table and window loads that the compiler may
    // schedule freely (b2_lds_ro), so the off-chain work of step i overlaps the chain of step i+1.  What needs a
    // branch is taken out of the steps: a lane that meets an entry the first-level table cannot resolve (code longer
    // than kHLutBits -- one symbol in 200 --, an error entry) STAYS where it is for the rest of the round (its
    // entries read 0: no advance, nothing for the consumers) and the symbol is resolved after the last step, in the
    // round's extra slot.
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
      for (int i = 0; i < kSegWords; ++i) {
        b2_cp_async4(ringl + ((rw + i) & (kSegWords - 1)) * 32, us + rw + i);
        if (((rw + i) & (kSegWords - 1)) == 0u) b2_cp_async4(ringl + kSegWords * 32, us + rw + i);   // slot 32 repeats slot 0
      }
      rw += kSegWords;
    } else tb[lane] = 0;
    b2_cp_async_commit();
    b2_cp_async_wait<0>();
    __syncwarp();
    // window addresses, taken AFTER the tables and the ring were filled (b2_lds_ro), and kept in registers
    const uint32_t ring_a = b2_launder(b2_dsm_addr(ringl)), tb_a = b2_launder(b2_dsm_addr(tb) + 4u * (uint32_t)lane);
    const uint32_t lut_r = b2_launder(lut_a);
    const uint32_t tb_end = tb_a + (uint32_t)L.bpm * 128u;
    uint32_t tba = tb_a;                             // this lane's table entry of the current block-in-MCU
    uint32_t z = 0, nb = 0;                          // zig-zag index, blocks finished
    uint32_t rounds = 0;
    // what one symbol does to the lane's state (entry e; two symbols when the entry pairs them)
    auto advance = [&](uint32_t e, uint32_t &fn) {
      const uint32_t z1 = huff_zinc(e), a2 = huff_adv2(e);
      const bool pair = a2 != 0u && z + z1 < 64u;    // an AC entry may describe TWO symbols; both are taken unless the first ends the block
      fn = z | (pair ? kSegPairBit : 0u) | (nb << kSegNbShift);
      p += huff_adv(e) + (pair ? a2 : 0u);
      z += z1 + (pair ? huff_zinc2(e) : 0u);
      // ---- block end without a branch: some lane ends a block in four steps out of five ----
      const bool end = z >= 64u;
      const uint32_t tbn = tba + 128u == tb_end ? tb_a : tba + 128u;
      z = end ? 0u : z;
      tba = end ? tbn : tba;
      nb += end ? 1u : 0u;
      // ---- once per lane: the segment is finished.  The last block ran past the segment's end (file.hpp:36-40),
      //      or more than padding is left: the segment holds more blocks ----
      const bool fin_now = end && nb == total;
      const uint32_t fe = p > pend ? (uint32_t)ST_OVERRUN : (pend - p > 7u ? (uint32_t)ST_BLOCK_COUNT : 0u);
      err |= fin_now ? fe : 0u;
      done = done || fin_now;
    };
    for (;;) {
      // the consumers' counters are read now and looked at after the round: room for the NEXT round
      uint32_t m = b2_ld_acquire_smem(ctr + 1);
#pragma unroll
      for (int c = 2; c <= kSegAcW + 1; ++c) { const uint32_t v = b2_ld_acquire_smem(ctr + c); m = v < m ? v : m; }
      // ---- top-up: once per round, the same instructions for every lane, predicated per lane (no branch).
      //      A round consumes at most 9 x 27 bits < 8 words; a top-up brings 8 ----
      const bool need = !done && rw - (p >> 5) <= (uint32_t)(kSegWords - 8);
#pragma unroll
      for (int i = 0; i < 8; ++i) b2_cp_async4_if(need, ring_a + (((rw + i) & (kSegWords - 1)) << 7), us + rw + i);
      // slot 32 repeats slot 0 (the window's second word is always at +128); rw is a multiple of 4
      b2_cp_async4_if(need && (rw & (kSegWords - 1)) == 0u, ring_a + (kSegWords << 7), us + rw);
      b2_cp_async4_if(need && ((rw + 4u) & (kSegWords - 1)) == 0u, ring_a + (kSegWords << 7), us + rw + 4);
      rw += need ? 8u : 0u;
      b2_cp_async_commit();
      // rw is at least 25 words ahead of the window and the round needs at most 8 more: everything but the two
      // newest groups must have landed -- copies issued two rounds (16 steps) ago
      b2_cp_async_wait<2>();
      const uint32_t slot_a = fifo_a + (rounds & (uint32_t)(kSegFifo - 1)) * kSegRoundBytes;
      uint32_t w = 0, cur = 0, eraw = 0;
#pragma unroll kSegUnroll
      for (int u = 0; u < kSegSteps; ++u) {
        const uint32_t t = b2_lds_ro(tba);                                           // table offsets of this block: DC | AC << 16
        cur = lut_r + (z == 0 ? (t & 0xffffu) : (t >> 16));                          // chosen while the window is still on its way
        const uint32_t wa = ring_a + ((p << 2) & (uint32_t)((kSegWords - 1) << 7));
        const uint32_t w0 = b2_lds_ro(wa), w1 = b2_lds_ro(wa + 128u);
        w = __funnelshift_l(w1, w0, p);
        eraw = b2_lds_ro(cur + ((w >> (30 - kHLutBits)) & (uint32_t)((kHL1 - 1) * 4)));
        // a finished lane, and one that waits for the extra slot, step in place: entry 0
        const uint32_t e = (done || eraw >= kHOverflow) ? 0u : eraw;
        uint32_t fn;
        advance(e, fn);
#if B2_SEG_DIAG != 2
        b2_sts(slot_a + (uint32_t)u * kSegStepBytes, w);
        b2_sts(slot_a + (uint32_t)u * kSegStepBytes + 128u, e);
        b2_sts(slot_a + (uint32_t)u * kSegStepBytes + 256u, fn);
#endif
      }
      {
        // ---- the extra slot: whoever is waiting in front of a symbol the first-level table did not resolve ----
        const bool stuck = !done && eraw >= kHOverflow;                              // then w, cur and eraw are still those of p
        uint32_t e = 0, fn = 0;
        if (__any_sync(0xffffffffu, stuck)) {
          if (stuck) {
            const bool is_dc = z == 0;
            e = eraw;
            if (e & kHLong) e = b2_lds(cur + 4u * ((e & 0xffffu) + ((w >> (32 - kHLutBits - 6)) & 63u)));
            if (e & kHOverflow) {
              const ImageDev &im = imgs[L.img];
              const int blk = (int)((tba - tb_a) >> 7), comp = blk < L.nluma ? 0 : blk - L.nluma + 1;
              e = huff_slow(is_dc ? tsets[im.tset].dc[im.td[comp] & 3] : tsets[im.tset].ac[im.ta[comp] & 3], w >> 16, is_dc);
            }
            if (e & kHBadCode) err |= ST_BAD_CODE;
            if (e & kHBadSize) err |= is_dc ? ST_BAD_DC_LEN : ST_BAD_AC_LEN;
            advance(e, fn);
          }
        }
#if B2_SEG_DIAG != 2
        b2_sts(slot_a + (uint32_t)kSegSteps * kSegStepBytes, w);
        b2_sts(slot_a + (uint32_t)kSegSteps * kSegStepBytes + 128u, e);
        b2_sts(slot_a + (uint32_t)kSegSteps * kSegStepBytes + 256u, fn);
#endif
      }
      const bool fin = __all_sync(0xffffffffu, done);   // also orders the lanes' FIFO stores before lane 0's release
      ++rounds;
      if (lane == 0) b2_st_release_smem(ctr, rounds | (fin ? kSegFin : 0u));
      if (fin) break;
      // room in the FIFO for the next round: every consumer is past round (rounds - kSegFifo)
#if B2_SEG_DIAG == 2
      m = rounds;
#endif
      while ((int)(rounds - m) >= kSegFifo) {
        b2_backoff(kSegPollNs);
        m = b2_ld_acquire_smem(ctr + 1);
#pragma unroll
        for (int c = 2; c <= kSegAcW + 1; ++c) { const uint32_t v = b2_ld_acquire_smem(ctr + c); m = v < m ? v : m; }
      }
    }
  }
#if B2_SEG_DIAG == 2
  else if (role > 0) return;
#endif
  else if (role <= kSegAcW) {
    // =========================== AC coefficients: stateless, every other round ===========================
    uint32_t r = (uint32_t)(role - 1);
    for (;;) {
      uint32_t pr;
      while (pr = b2_ld_acquire_smem(ctr), (pr & ~kSegFin) <= r && !(pr & kSegFin)) b2_backoff(kSegPollNs);
      if ((pr & ~kSegFin) <= r) break;                                              // no such round: the decoder has finished
      const uint32_t slot_a = fifo_a + (r & (uint32_t)(kSegFifo - 1)) * kSegRoundBytes;
      uint32_t fw[kSegSlots], fe[kSegSlots], fn[kSegSlots];
#pragma unroll
      for (int u = 0; u < kSegSlots; ++u) {
        fw[u] = b2_lds(slot_a + (uint32_t)u * kSegStepBytes); fe[u] = b2_lds(slot_a + (uint32_t)u * kSegStepBytes + 128u);
        fn[u] = b2_lds(slot_a + (uint32_t)u * kSegStepBytes + 256u);
      }
      __syncwarp();                                     // every lane has read the round before it is handed back
      if (lane == 0) b2_st_release_smem(ctr + role, r + (uint32_t)kSegAcW);
#pragma unroll
      for (int u = 0; u < (B2_SEG_DIAG == 1 || B2_SEG_DIAG == 5 ? 0 : kSegSlots); ++u) {
        const uint32_t e = fe[u], z = fn[u] & 63u, nblk = fn[u] >> kSegNbShift;
        if (z != 0u && e != 0u) {                       // AC (DC symbols carry z = 0, a lane that steps in place entry 0)
          int16_t *const bp = base + nblk * 64u;
          const uint32_t key = ((g0 + nblk) & 7u) << 3;
          const uint32_t adv = huff_adv(e), len = huff_len(e), zn = z + huff_zinc(e);
          const uint32_t size = adv - len;
          // EXTEND (jpeg.cpp:388-397), branch-free: v < 2^(size-1) <=> 2v <= 2^size - 1
          const uint32_t v = b2_shr(fw[u] << len, 32u - size), t = b2_shl(1u, size) - 1u;
          const int val = (int)v - (v + v > t ? 0 : (int)t);
          const uint32_t pos = zn - 1u;                                             // z + run
          if (size != 0u && pos < 64u B2_SEG_NOSTORE(fw[u])) bp[pos ^ key] = (int16_t)val;
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
            if (size2 != 0u && pos2 < 64u B2_SEG_NOSTORE(w2)) bp[pos2 ^ key] = (int16_t)val2;
            if (zn2 >= 64u && !(e2 & kHEob) && !(size2 != 0u && zn2 == 64u)) err |= ST_BAD_RUN;
          }
        }
      }
      r += (uint32_t)kSegAcW;
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
      uint32_t fw[kSegSlots], fe[kSegSlots], fn[kSegSlots];
#pragma unroll
      for (int u = 0; u < kSegSlots; ++u) {
        fw[u] = b2_lds(slot_a + (uint32_t)u * kSegStepBytes); fe[u] = b2_lds(slot_a + (uint32_t)u * kSegStepBytes + 128u);
        fn[u] = b2_lds(slot_a + (uint32_t)u * kSegStepBytes + 256u);
      }
      __syncwarp();
      ++r;
      if (lane == 0) b2_st_release_smem(ctr + kSegAcW + 1, r);
#pragma unroll
      for (int u = 0; u < (B2_SEG_DIAG == 1 || B2_SEG_DIAG == 4 ? 0 : kSegSlots); ++u) {
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
