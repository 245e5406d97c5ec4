// pipeline.h -- batch planner and kernel launch sequence of the scan-to-pixels path.
//
// This is what stands where Image::process_image_data() (jpeg.cpp:755-789) stands in the
// reference: it takes the parsed state of n images (b200jpeg_image_desc, the mirror of Image's
// private members) plus their file bytes and runs
//     U  unstuff + segment discovery      (file.hpp:59-104)
//     E  entropy decode                   (jpeg.cpp:300-446)
//     P  dequant + IDCT + upsample + RGB  (jpeg.cpp:511-753, display.hpp:19-34)
// Compiled twice from the same source: by nvcc into libb200jpeg.so (device memory, CUDA
// launches) and, with -DB200JPEG_EMU, by g++ into the test-only emulation library.
#pragma once
#include <stdint.h>
#include <string.h>
#include <string>
#include <vector>
#include <algorithm>

#include "../../include/b200jpeg.h"
#include "platform.h"
#include "dev_types.h"
#include "kernels_unstuff.cuh"
#include "kernels_huff.cuh"
#include "kernels_segdec.cuh"
#include "kernels_entropy.cuh"
#include "kernels_pixel.cuh"
#include "kernels_prog.cuh"

namespace b2 {

// ---------------------------------------------------------------- memory backend
#ifdef B200JPEG_EMU
struct Backend {
  std::string err;
  bool ok() const { return err.empty(); }
  void *alloc(size_t n) { return calloc(n ? n : 1, 1); }
  void release(void *p) { free(p); }
  void *host_alloc(size_t n) { return calloc(n ? n : 1, 1); }
  void host_release(void *p) { free(p); }
  void h2d(void *d, const void *s, size_t n) { memcpy(d, s, n); }
  void d2h(void *d, const void *s, size_t n) { memcpy(d, s, n); }
  void d2h_2d(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h) {
    for (size_t y = 0; y < h; ++y) memcpy((char *)d + y * dp, (const char *)s + y * sp, w);
  }
  void fill(void *d, int v, size_t n) { memset(d, v, n); }
  void sync() {}
  void check(const char *) {}
  cudaStream_t stream = nullptr;
};
#else
struct Backend {
  std::string err;
  cudaStream_t stream = nullptr;
  bool ok() const { return err.empty(); }
  void fail(const char *what, cudaError_t e) {
    if (err.empty()) err = std::string(what) + ": " + cudaGetErrorString(e);
  }
  void *alloc(size_t n) {
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, n ? n : 1);
    if (e != cudaSuccess) { fail("cudaMalloc", e); return nullptr; }
    return p;
  }
  void release(void *p) { if (p) cudaFree(p); }
  void *host_alloc(size_t n) {   // pinned: copies into it are truly asynchronous
    void *p = nullptr;
    cudaError_t e = cudaHostAlloc(&p, n ? n : 1, cudaHostAllocDefault);
    if (e != cudaSuccess) { fail("cudaHostAlloc", e); return nullptr; }
    return p;
  }
  void host_release(void *p) { if (p) cudaFreeHost(p); }
  void h2d(void *d, const void *s, size_t n) {
    if (!n) return;
    cudaError_t e = cudaMemcpyAsync(d, s, n, cudaMemcpyHostToDevice, stream);
    if (e != cudaSuccess) fail("cudaMemcpyAsync H2D", e);
  }
  void d2h(void *d, const void *s, size_t n) {
    if (!n) return;
    cudaError_t e = cudaMemcpyAsync(d, s, n, cudaMemcpyDeviceToHost, stream);
    if (e != cudaSuccess) fail("cudaMemcpyAsync D2H", e);
  }
  void d2h_2d(void *d, size_t dp, const void *s, size_t sp, size_t w, size_t h) {
    cudaError_t e = cudaMemcpy2DAsync(d, dp, s, sp, w, h, cudaMemcpyDeviceToHost, stream);
    if (e != cudaSuccess) fail("cudaMemcpy2DAsync D2H", e);
  }
  void fill(void *d, int v, size_t n) {
    if (!n) return;
    cudaError_t e = cudaMemsetAsync(d, v, n, stream);
    if (e != cudaSuccess) fail("cudaMemsetAsync", e);
  }
  void sync() {
    cudaError_t e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) fail("cudaStreamSynchronize", e);
  }
  void check(const char *what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) fail(what, e);
  }
};
#endif

// A device buffer that only ever grows (arena reused across batches).
struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
  bool reserve(Backend &be, size_t n) {
    if (n <= cap) return true;
    be.release(p);
    cap = n + n / 8 + 4096;
    p = be.alloc(cap);
    if (!p) { cap = 0; return false; }
    return true;
  }
  void destroy(Backend &be) { be.release(p); p = nullptr; cap = 0; }
  template <class T> T *as() const { return reinterpret_cast<T *>(p); }
};

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Sampling formats the pixel kernel is instantiated for.
enum PixFormat { PF_444 = 0, PF_422 = 1, PF_440 = 2, PF_420 = 3, PF_GRAY = 4, PF_COUNT = 5 };
inline int pix_tm(int f) {   // MCUs per pixel tile, PixCfg::kTM = 32 / blocks per MCU
  switch (f) { case PF_444: return 10; case PF_422: case PF_440: return 8; case PF_420: return 5; default: return 32; }
}

struct BatchPlan {
  int n = 0;
  std::vector<ImageDev> imgs;
  std::vector<TableSet> tsets;
  std::vector<int32_t> tile_base;                     // n+1 exclusive prefixes
  std::vector<int32_t> grp_img[PF_COUNT], grp_pre[PF_COUNT];
  PixUni uni[PF_COUNT];                               // arithmetic tile lookup for uniform groups
  std::vector<int> bwr, bhr;                          // reference block grid (coef export)
  size_t raw_bytes = 0, rgb_bytes = 0;
  int64_t nblocks = 0, pixels = 0, scan_bytes = 0;
  int ntiles = 0, nsegs = 0, nslots_max = 2;
  // entropy path of the batch (kPathSegment: one lane per restart segment; kPathSync: self-synchronising subs first)
  int path = kPathSync;
  std::vector<uint2> units;                           // segment path: (image, segment), grouped by table set, padded per group
  int nunit_ctas = 0;                                 // CTAs of k_entropy_seg (either path)
  // self-synchronising path (kernels_entropy.cuh): sub and unit slots, numbered batch-wide in table-set order
  XParams x;
  std::vector<int32_t> x_img, x_sub_base, x_unit_base;   // image at each position; n+1 prefixes of the reserved slots
};

// Sub and unit sizes of the self-synchronising path, from the batch's scan bits (measured: profiles/r02_xsweep.md).
// Both passes are bound by instruction issue along one dependent chain per lane, so what matters is (a) how many bits a
// lane walks -- sub + run-in for the sync pass, unit for the write pass: the shorter, the sooner the last wave ends -- and
// (b) that there are enough lanes to fill the device once (148 SMs x 4 CTAs x 128 threads = 75,776 sync threads).  So:
// sub = unit = the smallest power of two that keeps the batch within one wave of sync threads, between 2^9 (a single
// small picture: latency is one chain) and 2^13 (beyond that the write pass's tail grows faster than the sync pass's
// run-in overhead shrinks); run-in 2048 bits (4096 converges no better: the same one repair round, rarely two).
inline void choose_x_params(int64_t total_bits, XParams &x) {
#ifdef B2_X_FORCE_SHIFT                                // test builds: tiny subs, no run-in (exercises repair and fix)
  x.sub_shift = B2_X_FORCE_SHIFT; x.unit_subs = 2; x.overlap = 0;
  (void)total_bits;
#else
  int ss = kXMinShift;
  while (ss < 13 && (total_bits >> ss) > (int64_t)148 * 4 * kXT * 11 / 10) ++ss;   // up to 10 % over one wave: 78 k subs of 2^12 beat 39 k of 2^13 (C4)
  x.sub_shift = ss; x.unit_subs = 1; x.overlap = 2048;
#ifdef B200JPEG_TUNING
  if (const char *ev = getenv("B200JPEG_XSUB")) { const int v = atoi(ev); if (v >= kXMinShift && v <= kXMaxShift) x.sub_shift = v; }
  if (const char *ev = getenv("B200JPEG_XUNIT")) { const int v = atoi(ev); if (v >= 1 && v <= 64) x.unit_subs = v; }
  if (const char *ev = getenv("B200JPEG_XOVERLAP")) { const int v = atoi(ev); if (v >= 0 && v <= 65536) x.overlap = v; }
#endif
#endif
}

// Which entropy decoder takes the batch.  The segment path runs one long dependent chain per
// segment: its time is ~0.55 us per byte of the LONGEST segment (measured: 4.8 ms for 8.8 KB segments)
// whatever the batch holds, until the SMs fill up; the self-synchronising path costs ~0.034 us per KiB
// of the WHOLE batch (5.2 ms for 153 MB).  So the segment path wins once the batch holds >= ~16000
// segments' worth of its longest segment (profiles/README.md).
constexpr int64_t kSegPathRatio = 16000;
enum { kForceAuto = 0, kForceSegment = 1, kForceSync = 2 };
inline int choose_path(const b200jpeg_image_desc *descs, int n, int force) {
  if (force == kForceSegment) return kPathSegment;
  if (force == kForceSync) return kPathSync;
  int64_t total = 0, longest = 1;
  for (int i = 0; i < n; ++i) {
    const b200jpeg_image_desc &d = descs[i];
    const int hs = d.numComponents == 1 ? 1 : d.horizontalSamplingFactor, vs = d.numComponents == 1 ? 1 : d.verticalSamplingFactor;
    if (hs < 1 || vs < 1) return kPathSync;
    const int64_t mcus = (int64_t)((d.mcuWidth + hs - 1) / hs) * ((d.mcuHeight + vs - 1) / vs);
    const int64_t nseg = d.restartInterval > 0 ? (mcus + d.restartInterval - 1) / d.restartInterval : 1;
    total += d.scan_length;
    const int64_t unit = d.scan_length / (nseg > 0 ? nseg : 1) + 1;
    if (unit > longest) longest = unit;
  }
  return total >= kSegPathRatio * longest ? kPathSegment : kPathSync;
}

// Build the plan.  Returns 0 or a B200JPEG_E_* code.
// progressive: the descriptors are the FRAMES of progressive files (b200jpeg_progressive_desc::frame): Huffman tables travel
// with the scans, not with the frame, and the entropy stage is kernels_prog.cuh
inline int make_plan(const b200jpeg_image_desc *descs, int n, const size_t *file_sizes, BatchPlan &P, int force_path = kForceAuto,
                     bool progressive = false) {
  P = BatchPlan();
  P.n = n;
  P.path = choose_path(descs, n, force_path);
  P.imgs.resize(n);
  P.tile_base.assign(n + 1, 0);
  P.bwr.resize(n); P.bhr.resize(n);
  int64_t coef_base = 0;
  size_t raw_off = 0, rgb_off = 0;
  int seg_base = 0;
  memset(&P.x, 0, sizeof P.x);
  P.x.nimg = n;
  if (P.path == kPathSync) {
    int64_t bits = 0;
    for (int i = 0; i < n; ++i) bits += (int64_t)(descs[i].scan_length > 0 ? descs[i].scan_length : 0) * 8;
    choose_x_params(bits, P.x);
  }
  for (int i = 0; i < n; ++i) {
    const b200jpeg_image_desc &d = descs[i];
    ImageDev &im = P.imgs[i];
    memset(&im, 0, sizeof im);
    const int nc = d.numComponents, hs = d.horizontalSamplingFactor, vs = d.verticalSamplingFactor;
    if ((nc != 1 && nc != 3) || hs < 1 || hs > 2 || vs < 1 || vs > 2) return B200JPEG_E_ARG;
    if (d.image_width <= 0 || d.image_height <= 0) return B200JPEG_E_ARG;
    if (nc == 3) {   // the kernels take the subsampled component first (jpeg.cpp:110 allows id==1 anywhere)
      if (d.color_components[0].horizontal_sampling_factor != hs || d.color_components[0].vertical_sampling_factor != vs)
        return B200JPEG_E_UNSUPPORTED;
      for (int c = 1; c < 3; ++c)
        if (d.color_components[c].horizontal_sampling_factor != 1 || d.color_components[c].vertical_sampling_factor != 1)
          return B200JPEG_E_UNSUPPORTED;
    }
    if (d.scan_offset < 0 || d.scan_length < 0 || (size_t)(d.scan_offset + d.scan_length) > file_sizes[i]) return B200JPEG_E_ARG;
    if (d.scan_length >= (int64_t)1 << 28) return B200JPEG_E_UNSUPPORTED;   // bit positions are 32-bit
    im.width = d.image_width; im.height = d.image_height;
    im.hs = nc == 1 ? 1 : hs; im.vs = nc == 1 ? 1 : vs; im.ncomp = nc;
    im.bpm = nc == 1 ? 1 : hs * vs + 2;
    im.mcux = (d.mcuWidth + im.hs - 1) / im.hs;      // the x/y loops of jpeg.cpp:415-417
    im.mcuy = (d.mcuHeight + im.vs - 1) / im.vs;
    im.restart_interval = d.restartInterval;
    for (int c = 0; c < nc; ++c) {
      const b200jpeg_component &cc = d.color_components[c];
      if (cc.quantizationTableID < 0 || cc.quantizationTableID > 3 || cc.huffmanDCTTableID < 0 || cc.huffmanDCTTableID > 3 ||
          cc.huffmanACTableID < 0 || cc.huffmanACTableID > 3)
        return B200JPEG_E_TABLE_ID;                        // jpeg.cpp:156-160, 205-209
      if (!d.quantization_tables[cc.quantizationTableID].set ||
          (!progressive && (!d.huffmanDCTTable[cc.huffmanDCTTableID].set || !d.huffmanACTable[cc.huffmanACTableID].set)))
        return B200JPEG_E_TABLE_MISSING;                   // jpeg.cpp:757-774
      im.tq[c] = (int8_t)d.color_components[c].quantizationTableID;
      im.td[c] = (int8_t)d.color_components[c].huffmanDCTTableID;
      im.ta[c] = (int8_t)d.color_components[c].huffmanACTableID;
    }
    // table set, shared with earlier images when byte-identical
    TableSet ts; memset(&ts, 0, sizeof ts);
    for (int t = 0; t < 4; ++t) {
      if (d.quantization_tables[t].set) for (int k = 0; k < 64; ++k) {
        const int32_t q = d.quantization_tables[t].table[k];
        if (q < 0 || q > 65535) return B200JPEG_E_ARG;     // 8-bit in the reference (types.hpp:86-92), 16-bit at most (T.81 B.2.4.1)
        ts.q[t][k] = (uint16_t)q;
      }
      const b200jpeg_huffman_table *src[2] = {&d.huffmanDCTTable[t], &d.huffmanACTable[t]};
      HuffSpec *dst[2] = {&ts.dc[t], &ts.ac[t]};
      for (int k = 0; k < 2; ++k) if (src[k]->set) {
        // descriptors may be hand-built (C ABI, b200jpeg.hpp) without b200jpeg_parse_header's checks: the device
        // code indexes symbols[] by these counts (jpeg.cpp:170-181 bounds them the same way)
        if (src[k]->offset[0] != 0 || src[k]->offset[16] > 176) return B200JPEG_E_DHT;
        for (int l = 1; l <= 16; ++l) if (src[k]->offset[l] < src[k]->offset[l - 1]) return B200JPEG_E_DHT;
        memcpy(dst[k]->offset, src[k]->offset, 17); memcpy(dst[k]->symbols, src[k]->symbols, 176); dst[k]->set = 1;
      }
    }
    int found = -1;
    for (int t = (int)P.tsets.size() - 1; t >= 0 && t >= (int)P.tsets.size() - 4; --t)
      if (!memcmp(&P.tsets[t], &ts, sizeof ts)) { found = t; break; }
    if (found < 0) { P.tsets.push_back(ts); found = (int)P.tsets.size() - 1; }
    im.tset = found;
    {
      int nslots = 0;                                  // tables the set holds: what an entropy CTA keeps in shared memory
      for (int t = 0; t < 4; ++t) nslots += (P.tsets[found].dc[t].set != 0) + (P.tsets[found].ac[t].set != 0);
      if (nslots > P.nslots_max) P.nslots_max = nslots;
    }
    im.raw_len = (uint32_t)d.scan_length;
    im.raw_off = raw_off;
    raw_off += align_up((size_t)d.scan_length + 64, 256);
    const int64_t mcus = (int64_t)im.mcux * im.mcuy;
    im.nseg = d.restartInterval > 0 ? (int)((mcus + d.restartInterval - 1) / d.restartInterval) : 1;
    im.seg_base = seg_base; seg_base += im.nseg + 1;
    im.ntiles = (int)((d.scan_length + kUTile - 1) / kUTile);
    im.tile_base = P.tile_base[i]; P.tile_base[i + 1] = im.tile_base + im.ntiles;
    im.path = P.path;
    // slots reserved on the self-synchronising path: the unstuffed scan is at most as long as the stuffed one, and
    // every segment ends with at most one partial sub / unit (bases are assigned below, in table-set order)
    im.nsub_max = P.path == kPathSync ? (int)((((int64_t)d.scan_length * 8) >> P.x.sub_shift) + im.nseg + 1) : 0;
    im.nunit_max = P.path == kPathSync ? im.nsub_max / P.x.unit_subs + im.nseg + 1 : 0;
    im.nblocks = (int32_t)(mcus * im.bpm);
    im.coef_base = coef_base; coef_base += (im.nblocks + 7) & ~7;   // multiple of 8: the chunk swizzle key is (block & 7)
    im.rgb_pitch = (uint32_t)align_up((size_t)3 * im.width, 16);
    im.rgb_off = rgb_off; rgb_off += align_up((size_t)im.rgb_pitch * im.height, 256);
    P.bwr[i] = d.mcuWidthReal; P.bhr[i] = d.mcuHeightReal;
    P.pixels += (int64_t)im.width * im.height;
    P.scan_bytes += d.scan_length;
    const int f = nc == 1 ? PF_GRAY : (hs == 2 ? (vs == 2 ? PF_420 : PF_422) : (vs == 2 ? PF_440 : PF_444));
    if (P.grp_pre[f].empty()) P.grp_pre[f].push_back(0);
    const int tiles = im.mcuy * ((im.mcux + pix_tm(f) - 1) / pix_tm(f));
    P.grp_img[f].push_back(i);
    P.grp_pre[f].push_back(P.grp_pre[f].back() + tiles);
  }
  for (int f = 0; f < PF_COUNT; ++f) {                // uniform groups: see PixUni
    PixUni &u = P.uni[f];
    memset(&u, 0, sizeof u);
    const std::vector<int32_t> &gi = P.grp_img[f];
    if (gi.empty()) continue;
    const ImageDev &a = P.imgs[gi[0]];
    const int tpr = (a.mcux + pix_tm(f) - 1) / pix_tm(f), tpi = tpr * a.mcuy;
    bool ok = tpr < 65536 && tpi < 65536 && (int64_t)tpi * (int64_t)gi.size() < (1 << 24);
    const int64_t cstride = gi.size() > 1 ? P.imgs[gi[1]].coef_base - a.coef_base : 0;
    const uint64_t rstride = gi.size() > 1 ? P.imgs[gi[1]].rgb_off - a.rgb_off : 0;
    for (size_t k = 0; ok && k < gi.size(); ++k) {
      const ImageDev &b = P.imgs[gi[k]];
      ok = gi[k] == gi[0] + (int)k && b.mcux == a.mcux && b.mcuy == a.mcuy && b.width == a.width && b.height == a.height &&
           b.rgb_pitch == a.rgb_pitch && b.tset == a.tset && !memcmp(b.tq, a.tq, sizeof a.tq) &&
           b.coef_base == a.coef_base + (int64_t)k * cstride && b.rgb_off == a.rgb_off + (uint64_t)k * rstride;
    }
    if (!ok) continue;
    u.enabled = 1; u.img0 = gi[0]; u.mcux = a.mcux; u.tpr = tpr; u.tpi = tpi;
    u.width = a.width; u.height = a.height; u.pitch = a.rgb_pitch;
    u.qkey = (a.tset << 6) | (a.tq[0] & 3) | ((a.tq[1] & 3) << 2) | ((a.tq[2] & 3) << 4);
    u.magic_tpi = (((uint64_t)1 << 40) + (uint64_t)tpi - 1) / (uint64_t)tpi;
    u.magic_tpr = (((uint64_t)1 << 40) + (uint64_t)tpr - 1) / (uint64_t)tpr;
    u.coef_base0 = a.coef_base; u.coef_stride = cstride; u.rgb_off0 = a.rgb_off; u.rgb_stride = rstride;
  }
  std::vector<int> order(n);                          // images by table set: every entropy CTA sees one set
  for (int i = 0; i < n; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return P.imgs[a].tset < P.imgs[b].tset; });
  if (P.path == kPathSync) {
    // sub slots and unit slots batch-wide, image after image; where the table set changes the numbering skips to the
    // next multiple of the CTA size (the skipped slots belong to the image before them and stay unused)
    P.x_img.assign(order.begin(), order.end());
    P.x_sub_base.assign(n + 1, 0); P.x_unit_base.assign(n + 1, 0);
    int64_t sb = 0, ub = 0;
    for (int oi = 0; oi < n; ++oi) {
      ImageDev &im = P.imgs[order[oi]];
      if (oi > 0 && im.tset != P.imgs[order[oi - 1]].tset) { sb = (int64_t)align_up((size_t)sb, kXT); ub = (int64_t)align_up((size_t)ub, kSegUnits); }
      im.sub_base = (int32_t)sb; im.unit_base = (int32_t)ub;
      P.x_sub_base[oi] = (int32_t)sb; P.x_unit_base[oi] = (int32_t)ub;
      sb += im.nsub_max; ub += im.nunit_max;
      if (sb > (int64_t)1 << 30 || ub > (int64_t)1 << 30) return B200JPEG_E_UNSUPPORTED;
    }
    sb = (int64_t)align_up((size_t)sb, kXT); ub = (int64_t)align_up((size_t)ub, kSegUnits);
    P.x_sub_base[n] = (int32_t)sb; P.x_unit_base[n] = (int32_t)ub;
    P.x.nsub_slots = (int32_t)sb; P.x.nunit_slots = (int32_t)ub;
    P.x.ncta = (int32_t)(sb / kXT);
    P.nunit_ctas = (int)(ub / kSegUnits);
  }
  if (P.path == kPathSegment) {                       // units grouped by table set
    int cur_set = -1;
    for (int oi = 0; oi < n; ++oi) {
      const ImageDev &im = P.imgs[order[oi]];
      if (im.tset != cur_set) {
        while (P.units.size() % kSegUnits) P.units.push_back(make_uint2(kSegNoUnit, 0));
        cur_set = im.tset;
      }
      for (int k = 0; k < im.nseg; ++k) P.units.push_back(make_uint2((uint32_t)order[oi], (uint32_t)k));
    }
    while (P.units.size() % kSegUnits) P.units.push_back(make_uint2(kSegNoUnit, 0));
    P.nunit_ctas = (int)(P.units.size() / kSegUnits);
  }
  P.raw_bytes = raw_off; P.rgb_bytes = rgb_off; P.nblocks = coef_base;
  P.ntiles = P.tile_base[n]; P.nsegs = seg_base;
  return 0;
}

}  // namespace b2
