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
// passes over the stream for that.  The price is one long dependent chain per lane (position -> window -> table
// entry -> position), and all the chains of a batch are resident at once: the kernel takes as long as one chain,
// i.e. steps per segment x cycles per step, plus whatever else competes for the decoder's sub-partition.
//
// (1) Every step is the same for every lane, and the eight steps of a round are one basic block.  Events that are
//     rare for one lane (bit-buffer refill, end of block, a long code) happen in almost every step for SOME lane of a
//     warp, and a branch then costs all 32:
//   * no bit buffer: the 32-bit window is re-read each step from the bit position, out of a
//     per-lane ring of the next 32 stream words in shared memory laid out [word][lane] (bank =
//     lane: conflict-free whatever the lanes' positions; slot 32 repeats slot 0, the second word is always at +128);
//   * the ring is topped up by cp.async (LDGSTS, no register in between) under a per-lane
//     PREDICATE once per round -- nine symbols consume at most 243 bits, one top-up brings 256;
//   * DC or AC table by a select, block end and segment end by selects;
//   * an entry the first-level table cannot resolve (a code longer than kHLutBits: 0.5 % of the symbols) makes the
//     lane wait in place for the round's NINTH step, which is a step like the others plus a predicated look into the
//     table's second level: the waiting lane gets its symbol, every other lane one more symbol (a branch taken by
//     "whoever is stuck" ran in 72 % of the rounds and cost as much as one and a half steps while decoding nothing for
//     31 lanes: 5.8 % of the kernel);
//   * AC entries that hold two complete symbols inside the index bits are taken in one step (kernels_huff.cuh).
//
// (2) Only what the chain needs stays on the chain.  A group of 32 segments is worked on by four warps of the
//     same CTA:
//       decoder   walks the codes -- window, table entry, advance, block/table bookkeeping -- and pushes
//                 (window, entry, zig-zag index | block number) per lane and step into a shared-memory FIFO;
//       writers   take every kSegAcW-th round of the FIFO: magnitude extension (jpeg.cpp:388-397), zig-zag
//                 position.  They keep no state between symbols -- everything they need is in the FIFO entry; a
//                 DC symbol is a value for position 0 like any other: its DIFFERENCE goes into the block;
//       flusher   sends finished blocks to global memory (3).  A lane's blocks pass through it in order, so this is
//                 where the DC prediction per component (jpeg.cpp:335-345) and its range check happen.  It never
//                 touches the FIFO: the writer of a round passes on what the flusher needs of it (blocks finished per
//                 lane, last round or not), so that the decoder waits for the writers only.
//     Every hand-over is an mbarrier in shared memory: a waiting warp is suspended by the hardware.
//     What bounds the kernel (ncu, profiles/): the four warps of a group share one SM sub-partition, whose integer
//     pipe takes one warp instruction every two cycles and is busy 66-74 % of the time -- the decoder's chain and the
//     consumers' arithmetic compete for it, and instruction COUNT, not latency hiding, is what is left to gain.
//
// (3) Coefficients leave the SM as whole 128-byte blocks.  A 2-byte store per coefficient from 32 lanes that work on
//     32 different segments is 32 sectors per instruction: one L1 tag cycle each (233 M sectors per bench step = 0.8 ms
//     of every SM's load/store path, in front of which the decoder's shared-memory loads queue), a read-modify-write
//     in L2 for every sector, and a 1.6 GB memset beforehand because only non-zero values were stored.  Instead the
//     writers put the values into a per-lane ring of kSegAsm block images in SHARED memory (st.shared.u16), and the
//     flusher, once a round's writer is through, copies the blocks that round finished to global
//     memory with 128-bit stores -- eight threads per block, full lines, zeros included, no memset, no
//     read-modify-write -- and clears the images.  A writer checks, per lane, that the blocks of its round are less than kSegAsm ahead of the
//     flushed ones (they almost always are: a lane finishes 0.65 blocks per round) and waits otherwise.
//
//     Tried and dropped: decoders and consumers on different SMs (thread-block clusters, FIFO through DSMEM with
//     st.async + mbarrier transaction counts).  It works, but a consumer SM then carries the scattered stores of
//     eight groups (1.74 ms of tag cycles alone, measured 4.1 ms), and with block assembly its shared memory does not
//     hold eight groups' images (profiles/r01e_segdec_experiments.md).
#pragma once
#include "platform.h"
#include "dev_types.h"
#include "kernels_unstuff.cuh"
#include "kernels_huff.cuh"

namespace b2 {

#ifndef B2_SEG_GROUPS
#ifdef B200JPEG_EMU
#define B2_SEG_GROUPS 2
#else
#define B2_SEG_GROUPS 4
#endif
#endif
#ifndef B2_SEG_FIFO
#define B2_SEG_FIFO 2
#endif
#ifndef B2_SEG_ACW
#define B2_SEG_ACW 2
#endif
#ifndef B2_SEG_UNROLL
#define B2_SEG_UNROLL 8
#endif
#ifndef B2_SEG_POLL_NS
#define B2_SEG_POLL_NS 200
#endif
#ifndef B2_SEG_DIAG
#define B2_SEG_DIAG 0     // measurements only (scripts/build_variant.sh): 7 = cycle counters per role and part of a round
                          // (scripts/dev/seg_diag.py); wrong results on purpose: 5 = the writers do not compute, 8 = the flusher
                          // does not copy, 9 = no DC prediction
#endif
constexpr int kSegGroups = B2_SEG_GROUPS;     // groups (32 segments: one decoder warp + its consumers) per CTA
constexpr int kSegAcW = B2_SEG_ACW;           // AC writer warps per group: writer w takes the rounds r with r % kSegAcW == w
constexpr int kSegRoles = kSegAcW + 2;        // decoder, writers, flusher
constexpr int kSegThreads = kSegGroups * kSegRoles * 32;   // threads per CTA of k_entropy_seg
constexpr int kSegUnits = kSegGroups * 32;    // segments per CTA
constexpr uint32_t kSegNoUnit = 0xffffffffu;
constexpr int kSegWords = 32;                 // stream words per lane in the window ring (+ one slot that repeats slot 0)
constexpr int kSegFifo = B2_SEG_FIFO;         // rounds the FIFO holds
constexpr int kSegSteps = 8;                  // symbols per lane and round: the per-round work (top-up, hand-over) is paid once per 8 steps
constexpr int kSegSlots = kSegSteps + 1;      // FIFO slots per round: the steps + the ninth, which also looks into the tables' second level
constexpr int kSegSlotBytes = 3 * 128;        // one slot: (window, entry, zig-zag index | block number) x 32 lanes
constexpr int kSegRoundBytes = kSegSlots * kSegSlotBytes + 128;   // + the round's last word per lane (kSegTailNbShift)
constexpr int kSegMaxFin = (kSegSlots + 1) / 2;   // blocks a lane can FINISH in one round: a block is a DC symbol and at least one more
constexpr int kSegCopyDepth = 3;              // blocks per thread-octet and pass of the flusher's copy loop (2, 3, 4: the same time; 6: 2-4 % slower)
constexpr int kSegAsm = 8;                    // block images per lane in the assembly ring; a round touches at most 6 blocks of a lane
constexpr int kSegUnroll = B2_SEG_UNROLL;     // how far the decoder's step loop is unrolled (instruction-cache footprint)
constexpr unsigned kSegPollNs = B2_SEG_POLL_NS;   // a warp that finds nothing to do sleeps this long before it looks again
// third FIFO word of a step: zig-zag index before the symbol | two symbols taken << 6 | block number << 7
constexpr uint32_t kSegPairBit = 1u << 6;
constexpr int kSegNbShift = 7;
// last word of a round, per lane: blocks finished after the round << 4 | blocks finished IN the round << 1 | the decoder has finished
constexpr uint32_t kSegFinFlag = 1u;
constexpr int kSegTailNbShift = 4;
// what the flusher publishes per lane (fl): blocks of the lane that have left, low kSegFlBits bits | rounds it has finished
// above them.  Readers only ever need how far THEY are ahead of it -- a few blocks, a few rounds -- so the low bits of the
// differences are the differences, however long the unit is.  One word, so that one release (a fence) per round does
constexpr int kSegFlBits = 24;
// a writer waits on full[s] by phase parity, so it must see EVERY phase of the slots it reads: a slot belongs to one writer
static_assert(kSegFifo % kSegAcW == 0, "a FIFO slot always belongs to the same writer (a reader must not skip a barrier phase)");
static_assert((kSegFifo & (kSegFifo - 1)) == 0 && (kSegAsm & (kSegAsm - 1)) == 0, "ring sizes are powers of two");
static_assert(kSegAsm >= (kSegSlots + 1) / 2 + 1, "the blocks one round touches must fit the assembly ring");

constexpr int kSegWorkedDepth = 8;            // rounds between a writer and the flusher (worked[] barriers, side ring)
// shared memory of a group (32-bit words), after the CTA's copy of the lookup tables:
//   window ring | per-lane table offsets by block-in-MCU | FIFO | block images [image][lane][32 words] |
//   flushed-block count per lane | mbarriers | flush list
constexpr int kSegOffTb = (kSegWords + 1) * 32;
constexpr int kSegOffFifo = kSegOffTb + kMaxBpm * 32;
constexpr int kSegOffAsm = kSegOffFifo + kSegFifo * kSegRoundBytes / 4;
constexpr int kSegOffFl = kSegOffAsm + kSegAsm * 32 * 32;
constexpr int kSegOffBar = kSegOffFl + 32;    // mbarriers (2 words each)
constexpr int kSegOffList = kSegOffBar + 32;  // the flusher's list of finished blocks, a byte each
constexpr int kSegOffSide = kSegOffList + 32 * ((kSegSlots + 1) / 2 + 1) / 4;   // what the flusher needs of a round, a byte per lane (below)
constexpr int kSegGroupWords = kSegOffSide + kSegWorkedDepth * 32 / 4;
static_assert(kSegOffFifo % 4 == 0 && kSegOffAsm % 4 == 0 && kSegGroupWords % 4 == 0, "16-byte alignment of the FIFO and the block images");

B2_HD size_t segdec_smem_bytes() {
  return sizeof(uint32_t) * ((size_t)kHTabsPerSet * kHTab + (size_t)kSegGroups * kSegGroupWords);
}

// What every warp of a group knows about a lane's unit before the first symbol.  A unit is a whole restart segment
// (units[] = (image, segment), built by the planner: start byte-aligned, predictors zero, first block k * RI * blocks
// per MCU -- jpeg.cpp:419-425) or, on the self-synchronising path, a run of subs whose start state
// kernels_entropy.cuh has settled (XUnit: any bit position, any block-in-MCU phase, predictors = the running sums).
struct SegLane {
  bool live;                 // there is something to decode
  int img, bidx, bend;       // image, first block of the unit, first block past it
  int bpm, nluma;
  int64_t gb;                // global index of the first block in the coefficient buffer
  uint32_t p0, pend;         // bit range in the image's unstuffed scan
  int blk0;                  // block-in-MCU index of the first block
  int pred[3];               // DC predictors before the first block
};
B2_DEVICE SegLane seg_lane(const uint2 unit, const ImageDev *__restrict__ imgs, const ImageOut *__restrict__ out,
                           const uint32_t *__restrict__ seg_start) {
  SegLane L;
  L.live = false; L.img = 0; L.bidx = L.bend = 0; L.bpm = 1; L.nluma = 1; L.gb = 0;
  L.p0 = L.pend = 0; L.blk0 = 0; L.pred[0] = L.pred[1] = L.pred[2] = 0;
  if (unit.x == kSegNoUnit) return L;
  L.img = (int)unit.x;
  const ImageDev &im = imgs[L.img];
  const int k = (int)unit.y;
  L.bpm = im.bpm; L.nluma = im.ncomp == 1 ? 1 : im.hs * im.vs;
  const int seg_blocks = im.restart_interval * L.bpm;                                  // 0 when there is one segment
  L.bidx = k * seg_blocks;
  L.bend = (k + 1 < im.nseg) ? L.bidx + seg_blocks : im.nblocks;
  L.gb = im.coef_base + L.bidx;
  // a scan stage U gave up on is not decoded: its blocks read zero
  L.live = L.bidx < L.bend && !(out[L.img].status & (ST_BAD_MARKER | ST_NO_EOI | ST_BAD_RST_COUNT));
  if (L.live) { L.p0 = seg_start[im.seg_base + k]; L.pend = seg_start[im.seg_base + k + 1]; }
  return L;
}
// Unit slot `slot` of the self-synchronising path: slots are numbered batch-wide, image after image in the planner's
// table-set order (x_unit_base: n+1 prefixes, x_img: the image at each position); segunit[seg_base + nseg] is the
// number of units the image really has.
B2_DEVICE SegLane seg_lane_x(int slot, const ImageDev *__restrict__ imgs, const int32_t *__restrict__ x_unit_base,
                             const int32_t *__restrict__ x_img, int nimg, const XUnit *__restrict__ xunits,
                             const uint32_t *__restrict__ segunit) {
  SegLane L;
  L.live = false; L.img = 0; L.bidx = L.bend = 0; L.bpm = 1; L.nluma = 1; L.gb = 0;
  L.p0 = L.pend = 0; L.blk0 = 0; L.pred[0] = L.pred[1] = L.pred[2] = 0;
  const int s = find_owner(x_unit_base, nimg, slot);
  L.img = x_img[s];
  const ImageDev &im = imgs[L.img];
  L.bpm = im.bpm; L.nluma = im.ncomp == 1 ? 1 : im.hs * im.vs;
  if ((uint32_t)(slot - x_unit_base[s]) >= segunit[im.seg_base + im.nseg]) return L;   // a reserved slot the image does not use
  const XUnit u = xunits[slot];
  L.bidx = (int)u.bidx; L.bend = (int)u.bend;
  L.gb = im.coef_base + L.bidx;
  L.live = L.bidx < L.bend;
  if (!L.live) return L;     // a unit without a block start (a segment's short tail inside its last block): nothing to decode,
                             // and the idle lane must step in place over row 0 of its table column -- the only row it writes
  L.p0 = u.p0; L.pend = u.pend; L.blk0 = (int)u.blk0;
  L.pred[0] = u.pred[0]; L.pred[1] = u.pred[1]; L.pred[2] = u.pred[2];
  return L;
}

// ------------------------------------------------------------------ the FIFO between a decoder and its consumers
// Round r lives at (r % kSegFifo) * kSegRoundBytes: kSegSlots slots of three words x 32 lanes, then one word per lane
// (kSegTailNbShift).  Nobody polls: every hand-over is an mbarrier in shared memory,
// a waiting warp is suspended by the hardware (spinning consumers took two fifths of the kernel's issue slots).
//   full[s]    1 arrival: the decoder has written the round in slot s
//   empty[s]   1 arrival: the round's writer has the round in registers.  The flusher does not read the FIFO: with two
//              rounds of depth, a decoder that had to wait for the flusher to START a round before it could write the
//              next but one found no room in five rounds out of six.  What the flusher needs of a round -- how many
//              blocks each lane finished in it, and whether it is the last -- the writer passes on, a byte per lane,
//              in a ring as deep as worked[] (side)
//   worked[i]  1 arrival: the writer of round r (i = r % kSegWorkedDepth) has put its values into the block images and
//              its byte into the side ring -- what the flusher waits for.  A writer does not begin round r before the
//              flusher has finished round r - kSegWorkedDepth (the flusher's round counter, in the high bits of its per-lane word: kSegFlBits), so
//              a phase cannot complete twice before the flusher has looked at it
constexpr int kSegBars = 2 * kSegFifo + kSegWorkedDepth;
static_assert(2 * kSegBars <= 32, "32 words are set aside for the mbarriers");
struct SegFifo {
  uint32_t fifo_a;           // window address of slot 0 of round 0, this lane
  uint32_t bar_a;            // window address of full[0]
  B2_MEMBER uint32_t round_a(uint32_t r) const { return fifo_a + (r & (uint32_t)(kSegFifo - 1)) * kSegRoundBytes; }
  B2_MEMBER uint32_t full(uint32_t r) const { return bar_a + 8u * (r & (uint32_t)(kSegFifo - 1)); }
  B2_MEMBER uint32_t empty(uint32_t r) const { return bar_a + 8u * ((uint32_t)kSegFifo + (r & (uint32_t)(kSegFifo - 1))); }
  B2_MEMBER uint32_t worked(uint32_t r) const { return bar_a + 8u * ((uint32_t)(2 * kSegFifo) + (r & (uint32_t)(kSegWorkedDepth - 1))); }
  B2_MEMBER void init() const {                                  // one thread
    for (uint32_t s = 0; s < (uint32_t)kSegFifo; ++s) { b2_mbar_init_a(full(s), 1); b2_mbar_init_a(empty(s), 1); }
    for (uint32_t i = 0; i < (uint32_t)kSegWorkedDepth; ++i) b2_mbar_init_a(worked(i), 1);
  }
  // ---- producer ----
  B2_MEMBER void wait_room(uint32_t r) const {                   // round r may be written once round r - kSegFifo has been read
    if (r >= (uint32_t)kSegFifo) b2_mbar_wait_a(empty(r), ((r / (uint32_t)kSegFifo) - 1u) & 1u);
  }
  B2_MEMBER bool has_room(uint32_t r) const {                    // the same question without waiting
    return r < (uint32_t)kSegFifo || b2_mbar_test_a(empty(r), ((r / (uint32_t)kSegFifo) - 1u) & 1u);
  }
  B2_MEMBER void publish(uint32_t r, int lane) const {           // every lane's stores of the round are ordered before lane 0's arrive
    __syncwarp();
    if (lane == 0) b2_mbar_arrive_a(full(r));
  }
  // ---- consumers ----
  B2_MEMBER void wait(uint32_t r) const {
    b2_mbar_wait_a(full(r), (r / (uint32_t)kSegFifo) & 1u);
  }
  B2_MEMBER void release(uint32_t r, int lane) const {           // the warp has the round in registers
    __syncwarp();
    if (lane == 0) b2_mbar_arrive_a(empty(r));
  }
  B2_MEMBER void worked_arrive(uint32_t r, int lane) const {     // the warp's values of round r are in the block images
    __syncwarp();
    if (lane == 0) b2_mbar_arrive_a(worked(r));
  }
  B2_MEMBER void worked_wait(uint32_t r) const { b2_mbar_wait_a(worked(r), (r / (uint32_t)kSegWorkedDepth) & 1u); }
};

#if B2_SEG_DIAG == 7
// measurements only: cycles the decoder warps spend per part of a round (summed over warps), rounds, warps
__device__ unsigned long long g_seg_diag[24];   // decoder 0..8, writers 10..13, flusher 16..19
#define B2_SEG_TICK(k) do { const long long t_ = clock64(); diag[k] += (unsigned long long)(t_ - tprev); tprev = t_; } while (0)
#else
#define B2_SEG_TICK(k) do { } while (0)
#endif
// =========================== the chain: codes only ===========================
// The eight steps of a round are ONE basic block: no branch, table and window loads that the compiler may
// schedule freely (b2_lds_ro), so the off-chain work of step i overlaps the chain of step i+1.  What needs a
// branch is taken out of the steps: a lane that meets an entry the first-level table cannot resolve (code longer
// than kHLutBits -- one symbol in 200 --, an error entry) STAYS where it is for the rest of the round (its
// entries read 0: no advance, nothing for the consumers) and the symbol is resolved in the round's ninth step.
B2_DEVICE void seg_decode(const SegFifo &ff, int lane, const SegLane &L, const ImageDev *__restrict__ imgs,
                          const TableSet *__restrict__ tsets,
                          const uint8_t *__restrict__ d_us, uint32_t *ringl, uint32_t *tb, uint32_t lut_a, bool ovf_possible,
                          uint32_t &err) {
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
    p = L.p0;
    pend = L.pend;
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
  uint32_t tba = tb_a + (uint32_t)L.blk0 * 128u;   // this lane's table entry of the current block-in-MCU
  uint32_t z = 0, nb = 0;                          // zig-zag index, blocks finished
  // what one symbol does to the lane's state (entry e; two symbols when the entry pairs them)
  auto advance = [&](uint32_t e, uint32_t &fn) {
    // an AC entry may describe TWO symbols; both are taken unless the first could end the block: it moves the index by 16
    // at most, so below 48 it cannot -- a test on z alone, made while the entry is still on its way (the test on
    // z + z1 sat between the entry and the next window address: three more links in the chain)
    const uint32_t z1 = huff_zinc(e), a2 = huff_adv2(e) & (z < 48u ? 15u : 0u);
    const bool pair = a2 != 0u;
    fn = z | (pair ? kSegPairBit : 0u) | (nb << kSegNbShift);
    p += huff_adv(e) + a2;
    z += z1 + (pair ? huff_zinc2(e) : 0u);
    // ---- block end without a branch: some lane ends a block in four steps out of five ----
    const bool end = z >= 64u;
    const uint32_t tbn = tba + 128u == tb_end ? tb_a : tba + 128u;
    z = end ? 0u : z;
    tba = end ? tbn : tba;
    nb += end ? 1u : 0u;
    // ---- once per lane: the segment is finished (a finished lane no longer moves: p stays where the last block ended) ----
    done = done || (end && nb == total);
  };
  uint32_t r = 0;
#if B2_SEG_DIAG == 7
  unsigned long long diag[6] = {0, 0, 0, 0, 0, 0};
  long long tprev = clock64();
  uint32_t nroom = 0, lane_rounds = 0;     // lane_rounds: rounds in which THIS lane still had something to decode
#endif
  for (;; ++r) {
    const bool room = ff.has_room(r);     // asked before the top-up, needed after it: the barrier's answer takes ~90 cycles
    // ---- top-up: once per round, the same instructions for every lane, predicated per lane (no branch).
    //      A round consumes at most 9 x 27 bits < 8 words; a top-up brings 8.
    //      Nothing is fetched from beyond the segment's end (+ the window's look-ahead): a lane whose data ran out
    //      (truncated or corrupt scan) goes on over what the ring holds until its blocks are counted off, and ends
    //      with ST_OVERRUN -- it never reads outside the scan buffer, however many blocks the header promised ----
    const bool need = !done && rw - (p >> 5) <= (uint32_t)(kSegWords - 8) && rw <= (pend >> 5) + 8u;
    // (rw is a multiple of 4 and so is the ring size: four consecutive words never wrap)
    b2_cp_async4x4_if(need, ring_a + ((rw & (kSegWords - 1)) << 7), us + rw);
    b2_cp_async4x4_if(need, ring_a + (((rw + 4u) & (kSegWords - 1)) << 7), us + rw + 4);
    // slot 32 repeats slot 0 (the window's second word is always at +128); rw is a multiple of 4
    b2_cp_async4_if(need && (rw & (kSegWords - 1)) == 0u, ring_a + (kSegWords << 7), us + rw);
    b2_cp_async4_if(need && ((rw + 4u) & (kSegWords - 1)) == 0u, ring_a + (kSegWords << 7), us + rw + 4);
    rw += need ? 8u : 0u;
    b2_cp_async_commit();
    B2_SEG_TICK(0);
    if (!room) ff.wait_room(r);
#if B2_SEG_DIAG == 7
    nroom += room ? 0u : 1u;
    lane_rounds += done ? 0u : 1u;
#endif
    B2_SEG_TICK(1);
    const uint32_t slot_a = ff.round_a(r);
    // rw is at least 25 words ahead of the window and the round needs at most 8 more: everything but the two
    // newest groups must have landed -- copies issued two rounds (16 steps) ago
    b2_cp_async_wait<2>();
    B2_SEG_TICK(2);
    uint32_t w = 0, cur = 0, eraw = 0;
    const uint32_t nb0 = nb;                       // blocks finished before this round
    // The FIFO stores of a step are issued in the NEXT step, behind its loads: shared-memory stores and loads stay in
    // program order (the assembler cannot know that FIFO and window ring never overlap), so two stores in front of the
    // window loads are eight cycles of issue time in the chain; behind them they fall into the loads' shadow.
    uint32_t e_prev = 0, fn_prev = 0;
#pragma unroll kSegUnroll
    for (int u = 0; u < kSegSteps; ++u) {
      const uint32_t wa = b2_and_x4_add<(uint32_t)((kSegWords - 1) << 5)>(p, ring_a);      // ring_a + 128 * ((p >> 5) & 31)
      const uint32_t w0 = b2_lds_ro(wa), w1 = b2_lds_ro(wa + 128u);
      const uint32_t t = b2_lds_ro(tba);                                           // table offsets of this block: DC | AC << 16
      if (u > 0) {
        b2_sts(slot_a + (uint32_t)(u - 1) * kSegSlotBytes + 128u, e_prev);
        b2_sts(slot_a + (uint32_t)(u - 1) * kSegSlotBytes + 256u, fn_prev);
      }
      cur = lut_r + (z == 0 ? (t & 0xffffu) : (t >> 16));                          // chosen while the window is still on its way
      w = __funnelshift_l(w1, w0, p);
      eraw = b2_lds_ro(b2_shr_x4_add<32 - kHLutBits>(w, cur));
      b2_sts(slot_a + (uint32_t)u * kSegSlotBytes, w);
      // a finished lane steps in place: entry 0.  So does one that waits for the ninth step, by itself: the low bits of
      // a first-level entry with a flag are zero (kernels_huff.cuh), and the consumers find nothing in them either
      const uint32_t e = done ? 0u : eraw;
      uint32_t fn;
      advance(e, fn);
      e_prev = e; fn_prev = fn;
    }
    B2_SEG_TICK(3);
    // ---- the ninth step: a step like the others for every lane, plus the table's second level.  A lane that has been
    //      waiting in front of a long code since some earlier step reads the same first-level entry again and gets its
    //      symbol here; every other lane simply decodes one more symbol (and, if that one is long, does not wait at all).
    //      What two levels cannot settle -- kHOverflow, only in tables with more long-code prefixes than kHL2Blocks --
    //      takes a branch that is uniform for the CTA (ovf_possible); error entries are ordinary entries with flags ----
    uint32_t e, fn;
    {
      const uint32_t wa = ring_a + ((p << 2) & (uint32_t)((kSegWords - 1) << 7));
      const uint32_t w0 = b2_lds_ro(wa), w1 = b2_lds_ro(wa + 128u);
      const uint32_t t = b2_lds_ro(tba);
      b2_sts(slot_a + (uint32_t)(kSegSteps - 1) * kSegSlotBytes + 128u, e_prev);
      b2_sts(slot_a + (uint32_t)(kSegSteps - 1) * kSegSlotBytes + 256u, fn_prev);
      cur = lut_r + (z == 0 ? (t & 0xffffu) : (t >> 16));
      w = __funnelshift_l(w1, w0, p);
      eraw = b2_lds_ro(cur + ((w >> (30 - kHLutBits)) & (uint32_t)((kHL1 - 1) * 4)));
      const bool is_dc = z == 0;
      // (not for a finished or idle lane: what it reads need not be a table entry)
      const uint32_t e1 = eraw;
      eraw = b2_lds_ro_if(!done && (e1 & kHLong) != 0u, cur + 4u * (huff_l2_base(e1) + ((w >> (32 - kHLutBits - 6)) & 63u)), e1);
      eraw = huff_is_l1_error(e1) ? huff_l1_error(e1) : eraw;
      if (ovf_possible) {
        if (!done && (eraw & kHOverflow)) {
          const ImageDev &im = imgs[L.img];
          const int blk = (int)((tba - tb_a) >> 7), comp = blk < L.nluma ? 0 : blk - L.nluma + 1;
          eraw = huff_slow(is_dc ? tsets[im.tset].dc[im.td[comp] & 3] : tsets[im.tset].ac[im.ta[comp] & 3], w >> 16, is_dc);
        }
      }
      e = done ? 0u : eraw;
      err |= (e & kHBadCode) ? (uint32_t)ST_BAD_CODE : 0u;
      err |= (e & kHBadSize) ? (uint32_t)(is_dc ? ST_BAD_DC_LEN : ST_BAD_AC_LEN) : 0u;
      advance(e, fn);
    }
    B2_SEG_TICK(4);
    const bool fin = __all_sync(0xffffffffu, done);
    b2_sts(slot_a + (uint32_t)kSegSteps * kSegSlotBytes, w);
    b2_sts(slot_a + (uint32_t)kSegSteps * kSegSlotBytes + 128u, e);
    b2_sts(slot_a + (uint32_t)kSegSteps * kSegSlotBytes + 256u, fn);
    b2_sts(slot_a + (uint32_t)kSegSlots * kSegSlotBytes, (nb << kSegTailNbShift) | ((nb - nb0) << 1) | (fin ? kSegFinFlag : 0u));
    ff.publish(r, lane);
    B2_SEG_TICK(5);
    if (fin) break;
  }
#if B2_SEG_DIAG == 7
  if (lane == 0) {
    for (int k = 0; k < 6; ++k) atomicAdd(&g_seg_diag[k], diag[k]);
    atomicAdd(&g_seg_diag[6], (unsigned long long)(r + 1));
    atomicAdd(&g_seg_diag[7], 1ull);
    atomicAdd(&g_seg_diag[8], (unsigned long long)nroom);
  }
  {
    unsigned long long lr = lane_rounds;       // sum over the warp's lanes
    for (int d = 16; d > 0; d >>= 1) lr += __shfl_xor_sync(0xffffffffu, lr, d);
    if (lane == 0) atomicAdd(&g_seg_diag[9], lr);
  }
#endif
  // the last block ran past the segment's end (file.hpp:36-40), or more than padding is left: the segment holds more blocks
  if (L.live) err |= p > pend ? (uint32_t)ST_OVERRUN : (pend - p > 7u ? (uint32_t)ST_BLOCK_COUNT : 0u);
  // the AC writers whose turn did not come with the last round each get an empty one that tells them to stop
  for (int k = 1; k < kSegAcW; ++k) {
    ++r;
    ff.wait_room(r);
    const uint32_t slot_a = ff.round_a(r);
    for (int u = 0; u < kSegSlots; ++u) { b2_sts(slot_a + (uint32_t)u * kSegSlotBytes + 128u, 0u); b2_sts(slot_a + (uint32_t)u * kSegSlotBytes + 256u, 0u); }
    b2_sts(slot_a + (uint32_t)kSegSlots * kSegSlotBytes, (nb << kSegTailNbShift) | kSegFinFlag);
    ff.publish(r, lane);
  }
}

// A coefficient goes into the lane's image of its block: [image][lane][64 x int16].  The image is the block as it will
// sit in global memory, i.e. position pos at pos ^ key, key = (global block index & 7) << 3 (the pixel stage's swizzle),
// and the image number is global block index & 7 as well, so both come out of one product: with t = gn & 7 the byte
// address is asm_a + (2 pos ^ t * 0x1010) -- t << 12 selects the image, t << 4 is the key in bytes.
B2_DEVICE uint32_t seg_asm_addr(uint32_t asm_a, uint32_t gn, uint32_t pos) {
  return asm_a + ((pos + pos) ^ ((gn & (uint32_t)(kSegAsm - 1)) * 0x1010u));
}
static_assert(kSegAsm == 8, "image number and swizzle key share the low three bits of the global block index");
// EXTEND (jpeg.cpp:335-343, 388-397) of the `size` bits that follow the first `len` bits of w, branch-free: the value
// v counts from -(2^size - 1) when its first bit is 0
B2_DEVICE int seg_extend(uint32_t w, uint32_t len, uint32_t size) {
  const uint32_t x = w << len, cnt = 32u - size;
  const uint32_t v = b2_shr(x, cnt), t = b2_shr(0xffffffffu, cnt);
  return (int)v - (int)(t & ~(uint32_t)((int)x >> 31));
}

// =========================== AC coefficients: stateless, every kSegAcW-th round ===========================
// magnitude extension (jpeg.cpp:388-397), zig-zag position, one 2-byte store into the block image; everything a symbol
// needs is in its FIFO entry, so the writers keep no state between symbols and several of them work in parallel.
B2_DEVICE void seg_ac_writer(const SegFifo &ff, int w, int lane, uint32_t asm_a, const uint32_t *fl,
                             uint32_t side_a, uint32_t g0, uint32_t &err) {
#if B2_SEG_DIAG == 7
  unsigned long long diag[4] = {0, 0, 0, 0};
  long long tprev = clock64();
#endif
  for (uint32_t r = (uint32_t)w;; r += (uint32_t)kSegAcW) {
    ff.wait(r);
    B2_SEG_TICK(0);
    const uint32_t ra = ff.round_a(r);
    uint32_t fw[kSegSlots], fe[kSegSlots], fn[kSegSlots];
#pragma unroll
    for (int u = 0; u < kSegSlots; ++u) {
      fw[u] = b2_lds(ra + (uint32_t)u * kSegSlotBytes); fe[u] = b2_lds(ra + (uint32_t)u * kSegSlotBytes + 128u);
      fn[u] = b2_lds(ra + (uint32_t)u * kSegSlotBytes + 256u);
    }
    const uint32_t tail = b2_lds(ra + (uint32_t)kSegSlots * kSegSlotBytes);
    ff.release(r, lane);                                          // every lane has read the round before it is handed back
    // the blocks of this round (the highest is the one in progress after it) must fit the ring of images in front of
    // the flushed ones, and the round itself the ring of rounds between the writers and the flusher
    B2_SEG_TICK(2);
    // (one word says both, kSegFlBits: both differences are small, so their low bits are the differences)
    for (;;) {
      const uint32_t f = b2_ld_acquire_smem(fl);
      const bool wait = (((tail >> kSegTailNbShift) - f) & ((1u << kSegFlBits) - 1u)) >= (uint32_t)kSegAsm ||
                        ((r - (f >> kSegFlBits)) & ((1u << (32 - kSegFlBits)) - 1u)) >= (uint32_t)kSegWorkedDepth;
      if (!__any_sync(0xffffffffu, wait)) break;
      b2_backoff(kSegPollNs);
    }
    B2_SEG_TICK(1);
    uint32_t bad = 0;
#pragma unroll
    for (int u = 0; u < (B2_SEG_DIAG == 5 ? 0 : kSegSlots); ++u) {
      // no branch: every lane does the arithmetic of both symbols, the stores are predicated.  A DC symbol is just a
      // value for position 0 (z = 0, its entry moves the index by 1): its DIFFERENCE goes into the image, the flusher
      // turns it into the coefficient.  (A lane that steps in place has entry 0, i.e. size 0: no store, no run error.)
      const uint32_t e = fe[u], z = fn[u] & 63u;
      const bool two = (fn[u] & kSegPairBit) != 0u;
      const uint32_t gn = g0 + (fn[u] >> kSegNbShift);
      const uint32_t adv = huff_adv(e), size = huff_size(e), pos = z + huff_zinc(e) - 1u;   // z + run
      if (size != 0u && pos < 64u) b2_sts_u16(seg_asm_addr(asm_a, gn, pos), (uint32_t)seg_extend(fw[u], adv - size, size));
      // a run that reaches past 63 without being EOB is an error (jpeg.cpp:372-376); a coefficient AT 63 is not
      bad |= (e & kHEob) ? 0u : pos + (size == 0u ? 1u : 0u);
      // the second symbol of the entry (kernels_huff.cuh): its bits follow the first one's
      const uint32_t adv2 = huff_adv2(e), size2 = huff_size2(e), z2 = huff_zinc2_raw(e), pos2 = pos + z2;
      if (two && size2 != 0u && pos2 < 64u) b2_sts_u16(seg_asm_addr(asm_a, gn, pos2), (uint32_t)seg_extend(fw[u] << adv, adv2 - size2, size2));
      bad |= (!two || z2 == 0u) ? 0u : pos2 + (size2 == 0u ? 1u : 0u);    // z2 = 0: EOB
    }
    if (bad > 63u) err |= ST_BAD_RUN;                 // (positions are below 128: OR-ing them keeps "some position > 63")
    b2_sts_u8(side_a + (r & (uint32_t)(kSegWorkedDepth - 1)) * 32u + (uint32_t)lane, tail & ((1u << kSegTailNbShift) - 1u));
    ff.worked_arrive(r, lane);
    B2_SEG_TICK(2);
#if B2_SEG_DIAG == 7
    ++diag[3];
#endif
    if (tail & kSegFinFlag) break;
  }
#if B2_SEG_DIAG == 7
  if (lane == 0) for (int k = 0; k < 4; ++k) atomicAdd(&g_seg_diag[10 + k], diag[k]);
#endif
}

// =========================== the way out ===========================
// Once a round's writer is through with it, the blocks the round finished leave (DC prediction first): eight threads
// per block, 16 bytes each -- whole 128-byte lines -- and the image is cleared.
B2_DEVICE void seg_flusher(const SegFifo &ff, int lane, const SegLane &L, uint32_t asm_g /*image 0, lane 0*/, uint32_t list_a,
                           uint32_t side_a, uint32_t *fl, uint4 *__restrict__ coef4, uint32_t &err) {
  const uint32_t gblk = (uint32_t)L.gb;              // the coefficient buffer holds fewer than 2^32 blocks (512 GB)
  if (__any_sync(0xffffffffu, !L.live && L.bidx < L.bend)) {
    // a scan that is not decoded (stage U found it broken) still gets its blocks: zeros
    if (!L.live)
      for (uint32_t b = 0; b < (uint32_t)(L.bend - L.bidx); ++b)
        for (int c = 0; c < 8; ++c) coef4[(size_t)(gblk + b) * 8 + c] = make_uint4(0u, 0u, 0u, 0u);
  }
  char *cbase = reinterpret_cast<char *>(coef4) + ((lane & 7) << 4);   // this thread's sixteen bytes of block 0
  b2_keep(cbase);                                    // (kept in registers: the compiler would rather rebuild it in every pass)
  uint32_t flushed = 0;                              // blocks of this lane that have left
  const uint32_t asm_a = asm_g + ((uint32_t)lane << 7);
  int pred0 = L.pred[0], pred1 = L.pred[1], pred2 = L.pred[2];   // DC prediction per component (jpeg.cpp:335-345), restarted with the segment
  int blk = L.blk0;                                  // block-in-MCU of the lane's next block to leave
  uint32_t dcbad = 0;                                // != 0 once an absolute DC left the int16 range
#if B2_SEG_DIAG == 7
  unsigned long long diag[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  long long tprev = clock64();
#endif
  for (uint32_t r = 0;; ++r) {
    ff.worked_wait(r);                               // the round's writer is through
    B2_SEG_TICK(1);
    const uint32_t tail = b2_lds_u8(side_a + (r & (uint32_t)(kSegWorkedDepth - 1)) * 32u + (uint32_t)lane);
    const uint32_t cnt = tail >> 1, upto = flushed + cnt;                  // at most kSegMaxFin: a block takes two steps or more
    // who has a k-th finished block: everything below is counted off these masks (no vote inside a loop, no shuffle chain)
    uint32_t has[kSegMaxFin], excl = 0, total = 0;
#pragma unroll
    for (int k = 0; k < kSegMaxFin; ++k) has[k] = __ballot_sync(0xffffffffu, (uint32_t)k < cnt);
#pragma unroll
    for (int k = 0; k < kSegMaxFin; ++k) { excl += (uint32_t)__popc(has[k] & ((1u << lane) - 1u)); total += (uint32_t)__popc(has[k]); }
    // DC: the writers left each block's DC DIFFERENCE at position 0 of its image; a lane's blocks leave in order, so
    // this is where prediction happens -- once per block; the differences are fetched first, the sums follow
    int diff[kSegMaxFin];
#pragma unroll
    for (int k = 0; k < kSegMaxFin; ++k) {
      diff[k] = 0;
      if (B2_SEG_DIAG != 9 && has[k] != 0u) { if ((uint32_t)k < cnt) diff[k] = (int)(int16_t)b2_lds_u16(seg_asm_addr(asm_a, gblk + flushed + (uint32_t)k, 0u)); }
    }
#pragma unroll
    for (int k = 0; k < kSegMaxFin; ++k) {
      if (B2_SEG_DIAG != 9 && has[k] != 0u) {
        if ((uint32_t)k < cnt) {
          const int comp = blk < L.nluma ? 0 : blk - L.nluma + 1;
          const int dc = (comp == 0 ? pred0 : (comp == 1 ? pred1 : pred2)) + diff[k];
          pred0 = comp == 0 ? dc : pred0; pred1 = comp == 1 ? dc : pred1; pred2 = comp == 2 ? dc : pred2;
          dcbad |= (uint32_t)(dc + 32768);           // a bit above the low sixteen: the value is outside int16
          b2_sts_u16(seg_asm_addr(asm_a, gblk + flushed + (uint32_t)k, 0u), (uint32_t)dc);
          blk = blk + 1 == L.bpm ? 0 : blk + 1;
        }
      }
    }
    B2_SEG_TICK(4);
    // the round's finished blocks of all lanes as one list (a byte each: lane << 3 | number among the lane's blocks), so
    // that every pass of the copy loop moves whole blocks whatever lanes they belong to
#pragma unroll
    for (uint32_t k = 0; k < (uint32_t)kSegMaxFin; ++k)
      if (k < cnt) b2_sts_u8(list_a + excl + k, ((uint32_t)lane << 3) | k);
    __syncwarp();
    B2_SEG_TICK(5);
    // eight threads per block and kSegCopyDepth blocks per thread and pass: the loads of a pass are issued before its stores
    for (uint32_t q0 = 0; q0 < (B2_SEG_DIAG == 8 ? 0u : total); q0 += 4u * (uint32_t)kSegCopyDepth) {
      uint32_t gbv[kSegCopyDepth], sav[kSegCopyDepth];
      uint4 v[kSegCopyDepth];
#pragma unroll
      for (int j = 0; j < kSegCopyDepth; ++j) {
        const uint32_t q = q0 + 4u * (uint32_t)j + ((uint32_t)lane >> 3);
        const uint32_t ent = q < total ? b2_lds_u8(list_a + q) : 0u;
        const uint32_t ls = ent >> 3;
        gbv[j] = __shfl_sync(0xffffffffu, gblk + flushed, (int)ls) + (ent & 7u);   // global block index: its low bits name the image
        sav[j] = asm_g + ((gbv[j] & (uint32_t)(kSegAsm - 1)) << 12) + (ls << 7) + ((uint32_t)(lane & 7) << 4);
      }
#pragma unroll
      for (int j = 0; j < kSegCopyDepth; ++j)
        if (q0 + 4u * (uint32_t)j + ((uint32_t)lane >> 3) < total) v[j] = b2_lds_v4(sav[j]);
#pragma unroll
      for (int j = 0; j < kSegCopyDepth; ++j)
        if (q0 + 4u * (uint32_t)j + ((uint32_t)lane >> 3) < total) {
          b2_stg_v4(cbase + ((size_t)gbv[j] << 7), v[j]);                        // one multiply-add wide per address
          b2_sts_v4(sav[j], make_uint4(0u, 0u, 0u, 0u));
        }
    }
    B2_SEG_TICK(6);
#if B2_SEG_DIAG == 7
    diag[7] += total;
#endif
    flushed = upto;
    __syncwarp();                                    // the images are cleared before anybody learns that they are free
    b2_st_release_smem(fl, (flushed & ((1u << kSegFlBits) - 1u)) | ((r + 1u) << kSegFlBits));
    B2_SEG_TICK(2);
#if B2_SEG_DIAG == 7
    ++diag[3];
#endif
    if (tail & kSegFinFlag) break;
  }
#if B2_SEG_DIAG == 7
  if (lane == 0) for (int k = 0; k < 8; ++k) atomicAdd(&g_seg_diag[16 + k], diag[k]);
#endif
  if (dcbad >> 16) err |= ST_DC_RANGE;
}

// units[u] = (image, segment); the units of one CTA share a table set (the planner pads each table-set group to a
// multiple of kSegUnits with kSegNoUnit).  units == nullptr: the self-synchronising path's unit slots (seg_lane_x).
// d_us / seg_start: stage U's output.  Every block of every unit is written (whole, zeros included): the coefficient
// buffer needs no clearing.
B2_GLOBAL void __launch_bounds__(kSegThreads)
k_entropy_seg(const uint2 *__restrict__ units, const ImageDev *__restrict__ imgs, const TableSet *__restrict__ tsets,
              const uint32_t *__restrict__ luts, const uint32_t *__restrict__ seg_start,
              const uint8_t *__restrict__ d_us, ImageOut *__restrict__ out, int16_t *__restrict__ coef,
              const int32_t *__restrict__ x_unit_base, const int32_t *__restrict__ x_img, int nimg,
              const XUnit *__restrict__ xunits, const uint32_t *__restrict__ segunit) {
  B2_DYN_SMEM(uint32_t, lut);                                   // [8][kHTab], then the groups
  uint32_t seen = 0;                                            // OR of the table entries this thread copies
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = warp % kSegGroups, role = warp / kSegGroups;    // role 0 decoder, 1 .. kSegAcW writers, kSegAcW + 1 flusher
  uint32_t *const grp = lut + kHTabsPerSet * kHTab + g * kSegGroupWords;
  uint32_t *const ringl = grp + lane;                           // window ring: word j of this lane at ringl[(j & 31) * 32]
  uint32_t *const tb = grp + kSegOffTb;                         // [block-in-MCU][lane]: DC table offset | AC table offset << 16
  uint32_t *const fl = grp + kSegOffFl + lane;                  // blocks of this lane that have left the assembly ring
  SegFifo ff;
  ff.fifo_a = b2_dsm_addr(grp + kSegOffFifo) + 4u * (uint32_t)lane;
  ff.bar_a = b2_dsm_addr(grp + kSegOffBar);
  {
    // the CTA's table set: that of its first unit (never padding)
    const int set = units ? imgs[units[(size_t)blockIdx.x * kSegUnits].x].tset
                          : imgs[x_img[find_owner(x_unit_base, nimg, (int)blockIdx.x * kSegUnits)]].tset;
    const uint4 *src = reinterpret_cast<const uint4 *>(luts + (size_t)set * kHTabsPerSet * kHTab);
    uint4 *dst = reinterpret_cast<uint4 *>(lut);
    for (int t = 0; t < kHTabsPerSet; ++t) {
      const bool used = t < 4 ? tsets[set].dc[t].set : tsets[set].ac[t - 4].set;
      if (!used) continue;
      for (int k = tid; k < kHTab / 4; k += kSegThreads) {
        const uint4 v = b2_ld_nc_v4(src + t * (kHTab / 4) + k);
        dst[t * (kHTab / 4) + k] = v;
        seen |= v.x | v.y | v.z | v.w;
      }
    }
    // block images and flushed counts start at zero
    for (int gg = 0; gg < kSegGroups; ++gg) {
      uint4 *z = reinterpret_cast<uint4 *>(lut + kHTabsPerSet * kHTab + gg * kSegGroupWords + kSegOffAsm);
      for (int k = tid; k < (kSegAsm * 32 * 32 + 32) / 4; k += kSegThreads) z[k] = make_uint4(0u, 0u, 0u, 0u);
    }
    if (role == 0 && lane == 0) { ff.init(); b2_mbar_init_fence(); }
  }
  // does any table of the set hold codes that neither level resolves (kernels_huff.cuh: kHOverflow)?  uniform for the CTA
  const bool ovf_possible = __syncthreads_or((int)(seen & kHOverflow)) != 0;

  const int slot = (int)blockIdx.x * kSegUnits + g * 32 + lane;
  const SegLane L = units ? seg_lane(units[slot], imgs, out, seg_start)
                          : seg_lane_x(slot, imgs, x_unit_base, x_img, nimg, xunits, segunit);
  uint32_t err = 0;
  // the inner loops address shared memory through 32-bit window addresses (register + immediate)
  const uint32_t lut_a = b2_dsm_addr(lut), asm_g = b2_dsm_addr(grp + kSegOffAsm);
  const uint32_t g0 = (uint32_t)L.gb & 7u;           // block n of the segment: coefficient pos at pos ^ key(n), key(n) = ((g0 + n) & 7) << 3
  if (role == 0) seg_decode(ff, lane, L, imgs, tsets, d_us, ringl, tb, lut_a, ovf_possible, err);
  else if (role <= kSegAcW) seg_ac_writer(ff, role - 1, lane, asm_g + ((uint32_t)lane << 7), fl, b2_dsm_addr(grp + kSegOffSide), g0, err);
  else seg_flusher(ff, lane, L, asm_g, b2_dsm_addr(grp + kSegOffList), b2_dsm_addr(grp + kSegOffSide), fl, reinterpret_cast<uint4 *>(coef), err);
  if (err) atomicOr(&out[L.img].status, err);
}

}  // namespace b2
