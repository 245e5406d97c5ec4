// kernels_entropy.cuh -- stage E for scans WITHOUT (enough) restart markers: the self-synchronising pass.
// Together with the segment decoder (kernels_segdec.cuh) it replaces Image::decodeHuffman / decodeMCUComponent /
// getNextSymbol (jpeg.cpp:300-446) and BitStream::getBit/getBitN/align (file.hpp:130-164).
//
// The reference walks the scan bit by bit, one block after another; where a block starts is only known once every
// block before it has been decoded.  Here every restart segment (the whole scan when the file has no RSTn) is cut into
// fixed-size subsequences ("subs", 2^sub_shift bits, numbered batch-wide), one THREAD each:
//
//   plan   (k_xplan)   subs and units per segment, from the segment table stage U left behind.
//   sync   (k_xsync<1>) each thread finds the blocks that START inside its sub: how many, the sum of their DC
//          differences per component, and the first block start at/after the sub's end (its END STATE: bit position +
//          block-in-MCU index).  Its own start state is the predecessor's end state, which is not known yet, so the
//          thread starts `overlap` bits EARLIER from a guess (or from the segment's start, which is exact) and lets the
//          Huffman code self-synchronise on the way.  Thread t then checks its start against thread t-1's end; the few
//          subs that disagree are walked again (Jacobi iteration, compacted so that whole warps stay busy).
//   fix    (k_xsync<0>) the same check across CTA boundaries (one more launch per hop; two are issued, the host
//          launches more while a flag says state still moved).
//   scan   (k_xscan)   segmented prefix sums over the CTA totals;
//   units  (k_xunits)  ... and inside each CTA give every sub its first block and its DC predictors
//          (jpeg.cpp:344-345: a predictor is the running sum of the differences since the segment's start).  Runs of
//          `unit_subs` consecutive subs of a segment become UNITS (XUnit): bit range, block range, block-in-MCU phase,
//          predictors -- exactly what a lane of the segment decoder needs.
//   write  k_entropy_seg (kernels_segdec.cuh) decodes the units into coefficient blocks: there is ONE writer for both
//          entropy paths.  A restart segment is simply a unit whose start state needs no synchronisation.
//
// The walk itself (x_walk) is the segment decoder's chain: a round of eight identical, branch-free steps per lane (window
// from a per-lane shared-memory ring topped up by predicated cp.async, two-level table in shared memory, paired AC
// symbols, block end and sub end by selects), rare events (codes longer than the first-level index, the last bits of a
// segment) resolved in the round's extra slot behind one warp-uniform branch.  It extracts no coefficients: per DC
// symbol one magnitude extension for the sums, nothing for AC symbols.
// Algorithmic bytes: (1 + overlap / sub size) x scan bytes in (sync) + 36 B per sub of state.
#pragma once
#include <type_traits>
#include "platform.h"
#include "dev_types.h"
#include "kernels_unstuff.cuh"   // find_owner
#include "kernels_huff.cuh"
#include "kernels_segdec.cuh"    // seg_extend, the ring conventions

namespace b2 {

constexpr int kXT = 128;                     // subs (= threads) per CTA of the sync kernels
constexpr int kXWarps = kXT / 32;
constexpr int kXSteps = 8;                   // symbols per lane and round
constexpr int kChunkWords = 64 + 8;          // read-ahead slack past the end of a scan buffer (words)
constexpr int kXMinShift = 9, kXMaxShift = 16;

// dynamic shared memory (32-bit words): the table set's used tables (compacted), then per warp the window ring
// [word][lane] (+ the slot that repeats slot 0) and a [block-in-MCU][lane] column: DC table offset | component in its low two
// bits | AC table offset << 16
constexpr int kXRingWords = (kSegWords + 1) * 32;
constexpr int kXWarpWords = kXRingWords + kMaxBpm * 32;
static_assert((kHTab * 4) % 4 == 0 && kHTab * 4 >= 4, "the low two bits of a table's byte offset are free: they carry the component");
B2_HD size_t xsync_smem_bytes(int nslots) {
  return sizeof(uint32_t) * ((size_t)nslots * kHTab + (size_t)kXWarps * kXWarpWords);
}

// ---------------------------------------------------------------- where a sub lies
struct XLoc {
  bool valid, first, last;       // the slot holds a sub / it is the first / last sub of its segment
  int img, k;                    // image, segment
  uint32_t segp, B, lo, hi;      // segment [segp, B) and sub [lo, hi), bit positions in the image's unstuffed scan
  uint32_t jl;                   // number of the sub inside its segment
};
B2_DEVICE XLoc x_locate(int gsub, const XParams &xp, const int32_t *__restrict__ x_sub_base, const int32_t *__restrict__ x_img,
                        const ImageDev *__restrict__ imgs, const uint32_t *__restrict__ seg_start,
                        const uint32_t *__restrict__ segsub) {
  XLoc L;
  L.valid = L.first = L.last = false; L.img = 0; L.k = 0; L.segp = L.B = L.lo = L.hi = 0; L.jl = 0;
  if (gsub >= xp.nsub_slots) return L;
  const int s = find_owner(x_sub_base, xp.nimg, gsub);
  L.img = x_img[s];
  const ImageDev &im = imgs[L.img];
  const uint32_t j = (uint32_t)(gsub - x_sub_base[s]);
  const uint32_t *ss = segsub + im.seg_base;
  if (j >= ss[im.nseg]) return L;                       // a reserved slot the image does not use
  int a = 0, b = im.nseg - 1;                           // largest k with ss[k] <= j
  while (a < b) {
    const int mid = (a + b + 1) >> 1;
    if (ss[mid] <= j) a = mid; else b = mid - 1;
  }
  L.k = a;
  L.jl = j - ss[a];
  L.segp = seg_start[im.seg_base + a]; L.B = seg_start[im.seg_base + a + 1];
  L.lo = L.segp + (L.jl << xp.sub_shift);
  const uint32_t h = L.lo + (1u << xp.sub_shift);
  L.hi = h < L.B ? h : L.B;
  L.first = L.jl == 0u; L.last = j + 1u == ss[a + 1];
  L.valid = true;
  return L;
}

// ---------------------------------------------------------------- plan: subs and units per segment (one warp per image)
B2_GLOBAL void k_xplan(XParams xp, const ImageDev *__restrict__ imgs, ImageOut *__restrict__ out,
                       const uint32_t *__restrict__ seg_start, uint32_t *__restrict__ segsub, uint32_t *__restrict__ segunit) {
  const int i = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
  if (i >= xp.nimg) return;
  const ImageDev &im = imgs[i];
  const bool dead = out[i].status != 0;                 // stage U gave up on the scan: no subs, no units
  const uint32_t *seg = seg_start + im.seg_base;
  uint32_t cs = 0, cu = 0;
  for (int b = 0; b <= im.nseg; b += 32) {
    const int k = b + lane;
    uint32_t ns = 0, nu = 0;
    if (k < im.nseg && !dead) {
      const uint32_t len = seg[k + 1] - seg[k];
      // a segment without a single bit (a file cut right behind an RSTn) has no sub, and so nobody who would count its
      // blocks: the header promises at least one MCU here
      if (len == 0u) atomicOr(&out[i].status, (uint32_t)ST_BLOCK_COUNT);
      ns = (len + (1u << xp.sub_shift) - 1u) >> xp.sub_shift;
      nu = (ns + (uint32_t)xp.unit_subs - 1u) / (uint32_t)xp.unit_subs;
    }
    uint32_t is = ns, iu = nu;
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t os = __shfl_up_sync(0xffffffffu, is, d), ou = __shfl_up_sync(0xffffffffu, iu, d);
      if (lane >= d) { is += os; iu += ou; }
    }
    if (k <= im.nseg) { segsub[im.seg_base + k] = cs + is - ns; segunit[im.seg_base + k] = cu + iu - nu; }
    cs += __shfl_sync(0xffffffffu, is, 31);
    cu += __shfl_sync(0xffffffffu, iu, 31);
  }
}

// ---------------------------------------------------------------- the walk
struct XWalkIn {
  bool active;               // there is something to walk
  bool started;              // the start is a block start inside the sub (exact or the predecessor's end state), not a run-in
  const uint32_t *us;        // the image's unstuffed scan (words)
  uint32_t usbits;           // its length
  uint32_t p, blk;           // where to start: bit position, block-in-MCU index
  uint32_t lo, hi, B;        // the sub, the end of its segment
  int bpm;
  const ImageDev *im;        // for the per-length search of pathological tables only
};
struct XWalkOut { uint32_t sp, sblk, ep, eblk; int nb, d0, d1, d2; };
constexpr int kXOverNb = 1 << 28;            // "blocks" of a sub whose walk ran past its segment's end (see x_walk)

// All 32 lanes of a warp together.  ringl: this lane's column of the warp's window ring; tbl: this lane's column of the
// warp's table-offset entries, already filled for the lane's image ([blk] = DC table byte offset | component | AC << 16).  Counts the blocks that start in [sp, ep), sp = first block start >= lo, ep = first block
// start >= hi (or the segment's end), and sums their DC differences (jpeg.cpp:335-345) per component.
B2_DEVICE void x_walk(int lane, uint32_t *ringl, uint32_t *tbl, uint32_t lut_a, const XWalkIn &I,
                      const TableSet *__restrict__ tsets, XWalkOut &O) {
  (void)lane;
  uint32_t p = I.p, z = 0, nb = 0, rw = 0;
  int dsum = 0, d1 = 0, d2 = 0;
  bool started = I.started, done = !I.active;
  const uint32_t lo = I.lo, hi = I.hi, B = I.B;
  const uint32_t *us = I.us;
  if (I.active) {
    rw = (p >> 5) & ~3u;
    for (int i = 0; i < kSegWords; ++i) {
      b2_cp_async4(ringl + ((rw + i) & (kSegWords - 1)) * 32, us + rw + i);
      if (((rw + i) & (kSegWords - 1)) == 0u) b2_cp_async4(ringl + kSegWords * 32, us + rw + i);   // slot 32 repeats slot 0
    }
    rw += kSegWords;
  } else {                                              // an idle lane steps in place over entry 0 of table 0: keep its reads inside the CTA's shared memory
    tbl[0] = 0u;
  }
  b2_cp_async_commit();
  b2_cp_async_wait<0>();
  __syncwarp();
  const uint32_t ring_a = b2_launder(b2_dsm_addr(ringl)), tb_a = b2_launder(b2_dsm_addr(tbl));
  const uint32_t lut_r = b2_launder(lut_a);
  const uint32_t tb_end = tb_a + (uint32_t)(I.active ? I.bpm : 1) * 128u;
  uint32_t tba = tb_a + (I.active ? I.blk : 0u) * 128u;
  uint32_t sp = p, sba = tba;
  // what one symbol does to the lane's state (entry e; two symbols when the entry pairs them)
  // (run_in: a std::true_type while some lane of the warp is still before its sub's first block -- the bookkeeping of
  //  that moment is a sixth of a step's instructions, and three rounds in four no lane needs it)
  auto advance = [&](uint32_t e, uint32_t w, bool is_dc, uint32_t comp, auto run_in) {
    const uint32_t size = huff_size(e);
    int diff = seg_extend(w, huff_adv(e) - size, size);            // jpeg.cpp:335-343; 0 for size 0
    diff = is_dc ? diff : 0;
    // per component by multiply-add (comp is 0, 1 or 2: its two bits are the factors) -- on the FMA pipe, which this
    // kernel leaves idle, instead of a compare and a select on the integer pipe, which it fills
    dsum += diff; d1 += diff * (int)(comp & 1u); d2 += diff * (int)(comp >> 1);
    // both symbols of a paired entry unless the first could end the block (kernels_segdec.cuh: a test on z alone)
    const uint32_t z1 = huff_zinc(e), a2 = huff_adv2(e) & (z < 48u ? 15u : 0u);
    const bool pair = a2 != 0u;
    p += huff_adv(e) + a2;
    z += z1 + (pair ? huff_zinc2(e) : 0u);
    const bool end = z >= 64u;
    const uint32_t tbn = tba + 128u == tb_end ? tb_a : tba + 128u;
    z = end ? 0u : z;
    tba = end ? tbn : tba;
    nb += (end && p <= B) ? 1u : 0u;                               // a "block" that runs past the segment's end is padding
    if (decltype(run_in)::value) {
      const bool ns = end && !started && p >= lo;                  // the run-in is over: the block that starts here is the sub's first
      started = started || ns;
      sp = ns ? p : sp; sba = ns ? tba : sba;
      nb = ns ? 0u : nb; dsum = ns ? 0 : dsum; d1 = ns ? 0 : d1; d2 = ns ? 0 : d2;
    }
    done = done || (end && p >= hi);
  };
  bool all_started = false;                            // warp-uniform: every lane is past its run-in (or has nothing to walk)
  bool padblk = false;                                 // the segment's last bits (at most seven) were taken for a block start
  for (;;) {
    // top-up, once per round: the same instructions for every lane, predicated per lane (kernels_segdec.cuh)
    const bool need = !done && rw - (p >> 5) <= (uint32_t)(kSegWords - 8) && rw <= (I.usbits >> 5) + 8u;
    b2_cp_async4x4_if(need, ring_a + ((rw & (kSegWords - 1)) << 7), us + rw);
    b2_cp_async4x4_if(need, ring_a + (((rw + 4u) & (kSegWords - 1)) << 7), us + rw + 4);
    b2_cp_async4_if(need && (rw & (kSegWords - 1)) == 0u, ring_a + (kSegWords << 7), us + rw);
    b2_cp_async4_if(need && ((rw + 4u) & (kSegWords - 1)) == 0u, ring_a + (kSegWords << 7), us + rw + 4);
    rw += need ? 8u : 0u;
    b2_cp_async_commit();
    b2_cp_async_wait<2>();
    uint32_t w = 0, cur = 0, eraw = 0, comp = 0;
    bool is_dc = false, nearf = false;
    auto steps = [&](auto run_in) {
#pragma unroll
    for (int u = 0; u < kXSteps; ++u) {
      const uint32_t t = b2_lds_ro(tba);
      comp = t & 3u;
      is_dc = z == 0u;
      cur = lut_r + (is_dc ? (t & 0xfffcu) : (t >> 16));
      const uint32_t wa = b2_and_x4_add<(uint32_t)((kSegWords - 1) << 5)>(p, ring_a);      // ring_a + 128 * ((p >> 5) & 31)
      const uint32_t w0 = b2_lds_ro(wa), w1 = b2_lds_ro(wa + 128u);
      w = __funnelshift_l(w1, w0, p);
      eraw = b2_lds_ro(b2_shr_x4_add<32 - kHLutBits>(w, cur));
      // a block start inside the segment's last seven bits: padding or a very short block -- the extra slot decides
      nearf = is_dc && p + 7u >= B;
      // (an entry the first table level cannot resolve stands still by itself: kernels_huff.cuh)
      const uint32_t e = (done || nearf) ? 0u : eraw;
      advance(e, w, is_dc, comp, run_in);
    }
    };
    if (!all_started) all_started = __all_sync(0xffffffffu, started || done);
    if (all_started) steps(std::false_type()); else steps(std::true_type());
    // ---- the extra slot: lanes waiting in front of something the steps do not handle ----
    const bool stuck = !done && (nearf || eraw >= kHOverflow);
    if (__any_sync(0xffffffffu, stuck)) {
      if (stuck) {
        bool fin = false;
        if (nearf) {
          // T.81 F.1.2.3: the last byte of a segment is padded with 1-bits, and no code is all ones.  Anything else in
          // those bits is taken for a block (the reference counts blocks and ignores the padding, jpeg.cpp:419-425);
          // if it then runs past the segment's end it was padding after all (advance: p > B is not counted)
          const uint32_t n = p < B ? B - p : 0u;
          fin = n == 0u || (w >> (32u - n)) == ((1u << n) - 1u);
        }
        if (fin) {
          if (!started) { started = true; nb = 0u; dsum = d1 = d2 = 0; sp = B; sba = tb_a; }
          p = B; tba = tb_a; z = 0u; done = true;
        } else {
          padblk = padblk || nearf;
          uint32_t e = eraw;
          if (e & kHLong) e = b2_lds(cur + 4u * (huff_l2_base(e) + ((w >> (32 - kHLutBits - 6)) & 63u)));
          else if (huff_is_l1_error(e)) e = huff_l1_error(e);
          if (e & kHOverflow) {
            const int ci = (int)comp;
            e = huff_slow(is_dc ? tsets[I.im->tset].dc[I.im->td[ci] & 3] : tsets[I.im->tset].ac[I.im->ta[ci] & 3], w >> 16, is_dc);
          }
          advance(e, w, is_dc, comp, std::true_type());
        }
      }
    }
    if (__all_sync(0xffffffffu, done)) break;
  }
  b2_cp_async_wait<0>();                             // nothing of this walk lands in the ring once the next one fills it
  O.sp = sp; O.sblk = (sba - tb_a) >> 7;
  O.ep = p; O.eblk = (tba - tb_a) >> 7;
  if (O.sp >= B) { O.sp = B; O.sblk = 0; }           // the segment is over: the next block is the next segment's first
  if (O.ep >= B) { O.ep = B; O.eblk = 0; }
  O.nb = (int)nb; O.d1 = d1; O.d2 = d2; O.d0 = dsum - d1 - d2;
  // A block that began more than padding before the segment's end and ran past it is not one of the blocks counted above
  // -- if the header's count is already complete without it, nobody would notice that more than padding is left over
  // (the segment decoder says ST_BLOCK_COUNT to this; found by scripts/dev/soak.py).  A count no segment can reach makes
  // k_xunits say the same; its clamps keep every unit inside the segment's blocks.  Only a walk that met a block start
  // inside its own sub (sp < hi) knows what follows that start: a run-in that never synchronised says nothing.
  if (I.active && p > B && !padblk && sp < hi) O.nb = kXOverNb;
  if (!I.active) { O.sp = O.ep = I.p; O.sblk = O.eblk = I.blk; O.nb = O.d0 = O.d1 = O.d2 = 0; }
}

// ---------------------------------------------------------------- segmented sums over (flag, nb, d0, d1, d2): a flag restarts them
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
// CTA-wide exclusive segmented scan; *total = combine of all threads' values.
B2_DEVICE Seg5 seg5_cta_excl(const Seg5 &v, Seg5 *total) {
  B2_SHARED Seg5 wsum[kXWarps];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const Seg5 inc = seg5_warp_incl(v);
  if (lane == 31) wsum[wid] = inc;
  __syncthreads();
  Seg5 base = seg5_zero(), tot = seg5_zero();
  for (int w = 0; w < kXWarps; ++w) { if (w < wid) base = seg5_combine(base, wsum[w]); tot = seg5_combine(tot, wsum[w]); }
  __syncthreads();
  Seg5 prev = seg5_shfl_up(inc, 1);
  if (lane == 0) prev = seg5_zero();
  *total = tot;
  return seg5_combine(base, prev);
}

// ---------------------------------------------------------------- sync / fix
struct XShared {
  SubState e[kXT];                 // end state per sub
  SubState used[kXT];              // start state the end state was computed from
  SubInfo info[kXT];               // blocks / DC sums of the sub; bit 31 = first sub of its segment
  int32_t list[kXT];               // compacted indices of subs to walk again
  int32_t img[kXT];                // where each sub lies (x_locate), for whichever lane walks it
  uint32_t lo[kXT], hi[kXT], B[kXT];
  uint32_t flags[kXT];             // 1 valid, 2 first of its segment
  int32_t tabslot[kHTabsPerSet];   // table -> slot in the CTA's compacted copy (-1: not in the set)
  SubState cta_in;
  int32_t count, set, skip;
};

// this lane's columns of the warp's table entries, for image `im` (DC table td -> slot, AC table ta -> slot)
B2_DEVICE void x_set_tb(uint32_t *tbl, const ImageDev &im, const int32_t *tabslot) {
  const int nluma = im.ncomp == 1 ? 1 : im.hs * im.vs;
  for (int b = 0; b < im.bpm; ++b) {
    const int c = b < nluma ? 0 : b - nluma + 1;
    const int sd = tabslot[im.td[c] & 3], sa = tabslot[4 + (im.ta[c] & 3)];
    tbl[b * 32] = (uint32_t)(sd < 0 ? 0 : sd) * (kHTab * 4) | (uint32_t)c | ((uint32_t)(sa < 0 ? 0 : sa) * (kHTab * 4)) << 16;
  }
}

// FIRST=true : run-in + own sub from a guess (or an exact start), then repair inside the CTA.
// FIRST=false: take the previous CTA's settled end state (bnd_in) and repair, if it differs from what was used.
template <bool FIRST>
B2_GLOBAL void __launch_bounds__(kXT)
k_xsync(XParams xp, const ImageDev *__restrict__ imgs, const int32_t *__restrict__ x_sub_base, const int32_t *__restrict__ x_img,
        const TableSet *__restrict__ tsets, const uint32_t *__restrict__ luts, const ImageOut *__restrict__ out,
        const uint8_t *__restrict__ d_us, const uint32_t *__restrict__ seg_start, const uint32_t *__restrict__ segsub,
        SubState *__restrict__ sub_e, SubState *__restrict__ sub_used, SubInfo *__restrict__ sub_info,
        SubInfo *__restrict__ cta_tot, const SubState *__restrict__ bnd_in, SubState *__restrict__ bnd_out,
        uint32_t *__restrict__ changed_flag) {
  B2_DYN_SMEM(uint32_t, lut);
  B2_SHARED XShared S;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int cta = (int)blockIdx.x, gsub = cta * kXT + tid;
  const XLoc loc = x_locate(gsub, xp, x_sub_base, x_img, imgs, seg_start, segsub);

  if (tid == 0) {
    S.set = imgs[x_img[find_owner(x_sub_base, xp.nimg, cta * kXT)]].tset;
    S.skip = 0;
    if (!FIRST) {
      // cheap uniform exit: this CTA's first sub does not depend on the previous CTA, or nothing upstream moved
      int skip = 1;
      if (cta > 0 && loc.valid && !loc.first) {
        const SubState in = bnd_in[cta - 1], used = sub_used[gsub];
        skip = in.p == used.p && in.blk == used.blk;
        S.cta_in = in;
      }
      if (skip) bnd_out[cta] = bnd_in[cta];
      S.skip = skip;
    }
    int n = 0;
    for (int t = 0; t < kHTabsPerSet; ++t) {
      const bool used = t < 4 ? tsets[S.set].dc[t].set : tsets[S.set].ac[t - 4].set;
      S.tabslot[t] = used ? n++ : -1;
    }
  }
  __syncthreads();
  if (!FIRST && S.skip) return;

  {  // the table set's tables, compacted
    const uint4 *src = reinterpret_cast<const uint4 *>(luts + (size_t)S.set * kHTabsPerSet * kHTab);
    uint4 *dst = reinterpret_cast<uint4 *>(lut);
    for (int t = 0; t < kHTabsPerSet; ++t) {
      const int slot = S.tabslot[t];
      if (slot < 0) continue;
      for (int k = tid; k < kHTab / 4; k += kXT) dst[slot * (kHTab / 4) + k] = b2_ld_nc_v4(src + t * (kHTab / 4) + k);
    }
  }
  int nslots = 0;
  for (int t = 0; t < kHTabsPerSet; ++t) nslots += S.tabslot[t] >= 0;
  uint32_t *wbase = lut + (size_t)nslots * kHTab + (size_t)warp * kXWarpWords;
  uint32_t *ringl = wbase + lane, *tbl = wbase + kXRingWords + lane;
  const uint32_t lut_a = b2_dsm_addr(lut);

  S.img[tid] = loc.img; S.lo[tid] = loc.lo; S.hi[tid] = loc.hi; S.B[tid] = loc.B;
  S.flags[tid] = (loc.valid ? 1u : 0u) | (loc.first ? 2u : 0u);
  __syncthreads();                                   // tables and locations are in place

  auto walk_in = [&](int j, SubState from, bool started) {
    XWalkIn I;
    const ImageDev &im = imgs[S.img[j]];
    I.active = (S.flags[j] & 1u) != 0u && from.p < S.hi[j];
    I.started = started;
    I.us = reinterpret_cast<const uint32_t *>(d_us + im.raw_off);
    I.usbits = out[S.img[j]].us_len * 8u;
    I.p = from.p; I.blk = from.blk;
    I.lo = S.lo[j]; I.hi = S.hi[j]; I.B = S.B[j];
    I.bpm = im.bpm; I.im = &im;
    return I;
  };

  if (FIRST) {
    SubState from; from.p = loc.lo; from.blk = 0;
    bool started = true;
    if (loc.valid && !loc.first) {
      // start `overlap` bits early from a guess; from the segment's start if that is within reach (then the run-in is exact)
      started = false;
      from.p = (loc.lo - loc.segp > (uint32_t)xp.overlap) ? loc.lo - (uint32_t)xp.overlap : loc.segp;
    }
    if (loc.valid) x_set_tb(tbl, imgs[loc.img], S.tabslot);
    const XWalkIn I = walk_in(tid, from, started);
    XWalkOut R;
    x_walk(lane, ringl, tbl, lut_a, I, tsets, R);
    SubState e, u; e.p = R.ep; e.blk = R.eblk; u.p = R.sp; u.blk = R.sblk;
    Seg5 s; s.nb = R.nb; s.d0 = R.d0; s.d1 = R.d1; s.d2 = R.d2; s.flag = loc.first ? 1 : 0;
    if (!loc.valid) { e.p = e.blk = u.p = u.blk = 0; s = seg5_zero(); }
    S.e[tid] = e; S.used[tid] = u; S.info[tid] = seg5_to(s);
  } else {
    S.e[tid] = sub_e[gsub]; S.used[tid] = sub_used[gsub]; S.info[tid] = sub_info[gsub];
  }
  __syncthreads();

  // ---- Jacobi repair inside the CTA: walk again every sub whose start state differs from its predecessor's end state,
  //      the needy subs compacted onto the first threads so that whole warps stay busy ----
  const bool have_in = !FIRST;
  for (int round = 0; round <= kXT; ++round) {
    bool need = false;
    if ((S.flags[tid] & 3u) == 1u && (tid > 0 || have_in)) {     // valid and not the first sub of its segment
      const SubState prev = tid == 0 ? S.cta_in : S.e[tid - 1], u = S.used[tid];
      need = prev.p != u.p || prev.blk != u.blk;
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
    XWalkOut R; int j = -1; SubState from; from.p = from.blk = 0;
    if (warp * 32 < cnt) {                              // warp-uniform: the warp holds at least one sub to walk
      XWalkIn I; I.active = false; I.started = true; I.us = nullptr; I.usbits = 0; I.p = I.blk = 0; I.lo = I.hi = I.B = 0; I.bpm = 1; I.im = &imgs[0];
      if (tid < cnt) {
        j = S.list[tid];
        from = j == 0 ? S.cta_in : S.e[j - 1];
        x_set_tb(tbl, imgs[S.img[j]], S.tabslot);
        I = walk_in(j, from, true);
      }
      x_walk(lane, ringl, tbl, lut_a, I, tsets, R);
    }
    __syncthreads();                                    // every S.e read of this round precedes every write
    if (j >= 0) {
      SubState e; e.p = R.ep; e.blk = R.eblk;
      Seg5 s; s.nb = R.nb; s.d0 = R.d0; s.d1 = R.d1; s.d2 = R.d2; s.flag = 0;
      S.e[j] = e; S.used[j] = from; S.info[j] = seg5_to(s);
    }
    __syncthreads();
  }

  const SubState e = S.e[tid];
  sub_e[gsub] = e; sub_used[gsub] = S.used[tid];
  const SubInfo mine = S.info[tid];
  sub_info[gsub] = mine;
  Seg5 total;
  seg5_cta_excl(seg5_from(mine), &total);
  if (tid == kXT - 1) {
    if (!FIRST) {
      const SubState old = bnd_in[cta];
      if (old.p != e.p || old.blk != e.blk) atomicOr(changed_flag, 1u);
    }
    bnd_out[cta] = e;
  }
  if (tid == 0) cta_tot[cta] = seg5_to(total);
}

// ---------------------------------------------------------------- scan over the CTA totals of the whole batch (one warp)
B2_GLOBAL void k_xscan(int ncta, const SubInfo *__restrict__ cta_tot, SubInfo *__restrict__ cta_pre) {
  const int lane = threadIdx.x & 31;
  if (threadIdx.x >= 32) return;
  Seg5 carry = seg5_zero();
  for (int b = 0; b < ncta; b += 32) {
    const int c = b + lane;
    Seg5 v = seg5_zero();
    if (c < ncta) v = seg5_from(cta_tot[c]);
    const Seg5 inc = seg5_warp_incl(v);
    Seg5 prev = seg5_shfl_up(inc, 1);
    if (lane == 0) prev = seg5_zero();
    if (c < ncta) cta_pre[c] = seg5_to(seg5_combine(carry, prev));
    carry = seg5_combine(carry, seg5_shfl(inc, 31));
  }
}

// ---------------------------------------------------------------- units: what the segment decoder's lanes will decode
// Unit ul of a segment = its subs [ul * unit_subs, (ul + 1) * unit_subs).  The thread of a unit's first sub writes where
// and in which state the unit starts, the thread of its last sub where it ends; the thread of a segment's last sub checks
// that the segment held the blocks the header promises (RI MCUs; the rest of the image for the last segment).
B2_GLOBAL void __launch_bounds__(kXT)
k_xunits(XParams xp, const ImageDev *__restrict__ imgs, const int32_t *__restrict__ x_sub_base, const int32_t *__restrict__ x_img,
         const int32_t *__restrict__ x_unit_base, ImageOut *__restrict__ out, const uint32_t *__restrict__ seg_start,
         const uint32_t *__restrict__ segsub, const uint32_t *__restrict__ segunit, const SubState *__restrict__ sub_e,
         const SubState *__restrict__ sub_used, const SubInfo *__restrict__ sub_info, const SubInfo *__restrict__ cta_pre,
         XUnit *__restrict__ xunits) {
  const int tid = threadIdx.x, cta = (int)blockIdx.x, gsub = cta * kXT + tid;
  const XLoc loc = x_locate(gsub, xp, x_sub_base, x_img, imgs, seg_start, segsub);
  const SubInfo mine = sub_info[gsub];
  Seg5 total;
  Seg5 pre = seg5_cta_excl(seg5_from(mine), &total);
  pre = seg5_combine(seg5_from(cta_pre[cta]), pre);
  if (!loc.valid) return;
  if (loc.first) pre = seg5_zero();                     // the sums before a segment's first sub belong to the segment before
  const ImageDev &im = imgs[loc.img];
  const int seg_blocks = im.restart_interval * im.bpm;  // 0 when there is one segment
  const int base = loc.k * seg_blocks;
  const int limit = (loc.k + 1 < im.nseg) ? base + seg_blocks : im.nblocks;
  const int nb = mine.nb_flag & 0x7fffffff;
  int first_blk = base + pre.nb, past_blk = first_blk + nb;
  if (loc.last && past_blk != limit) atomicOr(&out[loc.img].status, (uint32_t)ST_BLOCK_COUNT);
  first_blk = first_blk < limit ? first_blk : limit;    // whatever a corrupt scan yields, no lane writes outside its segment's blocks
  past_blk = past_blk < limit ? past_blk : limit;
  const uint32_t us = (uint32_t)xp.unit_subs, ul = loc.jl / us, r = loc.jl - ul * us;
  const int s = find_owner(x_sub_base, xp.nimg, gsub);
  XUnit &u = xunits[x_unit_base[s] + (int)segunit[im.seg_base + loc.k] + (int)ul];
  if (r == 0u) {
    const SubState st = sub_used[gsub];                 // == the predecessor's end state once the repair has settled
    u.p0 = loc.first ? loc.segp : st.p;
    u.blk0 = loc.first ? 0u : st.blk;
    u.bidx = (uint32_t)first_blk;
    u.pred[0] = pre.d0; u.pred[1] = pre.d1; u.pred[2] = pre.d2;
  }
  if (r == us - 1u || loc.last) {
    u.bend = (uint32_t)past_blk;
    u.pend = loc.last ? loc.B : sub_e[gsub].p;
  }
}

}  // namespace b2
