// pipeline_run.h -- device arenas, upload, the launch sequence, download.  See pipeline.h.
#pragma once
#include <stdlib.h>
#include "pipeline.h"

namespace b2 {

constexpr int kFixLaunches = 2;       // cross-CTA repair launches issued per run
constexpr int kMaxExtraFix = 64;      // further repair rounds if the last launch still moved state

struct Pipeline {
  Backend be;
  BatchPlan plan;
  bool prepared = false, ran = false;
  // arenas
  DevBuf d_raw, d_us, d_imgs, d_tsets, d_out, d_tile_base, d_tile_cnt, d_tile_pre, d_seg;
  DevBuf d_sub_e, d_sub_used, d_sub_info, d_cta_tot, d_cta_pre, d_bnd[2], d_flags, d_coef, d_rgb, d_export;
  DevBuf d_segsub, d_segunit, d_x_img, d_x_sub_base, d_x_unit_base, d_xunits;   // self-synchronising path
  DevBuf d_pimgs, d_ptables;                                                      // progressive files (kernels_prog.cuh)
  int prog_max_scans = 0, prog_levels = 0;
  DevBuf d_grp_img[PF_COUNT], d_grp_pre[PF_COUNT];
  DevBuf d_units, d_luts;
  int force_path = kForceAuto;
  bool copies_only = false;             // b200jpeg_copy_floor: run() launches nothing
  // small results (per-image ImageOut, convergence flags) land in PINNED host memory, copied right behind
  // the kernels and BEFORE the bulk RGB copies: a pageable read-back would queue behind the other
  // pipeline's RGB traffic in the D2H engine and stall the host for milliseconds
  ImageOut *h_out = nullptr;
  uint32_t *h_flags = nullptr;
  size_t h_out_cap = 0;
  bool results_enqueued = false;
  int launches = 0, sync_rounds = 0;
  int pix_grid_cap[PF_COUNT] = {2, 2, 2, 2, 2};   // resident CTAs of k_pixel per format (emulation: 2, so the tile loop is exercised)
  int64_t h2d_bytes = 0, d2h_bytes = 0;
  bool timing = false;
  float last_ms[B200JPEG_T_COUNT] = {0, 0, 0, 0, 0};
#ifndef B200JPEG_EMU
  cudaEvent_t ev[B200JPEG_T_COUNT + 1] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
#endif

  void destroy() {
    DevBuf *all[] = {&d_raw, &d_us, &d_imgs, &d_tsets, &d_out, &d_tile_base, &d_tile_cnt, &d_tile_pre,
                     &d_seg, &d_sub_e, &d_sub_used, &d_sub_info, &d_cta_tot, &d_cta_pre, &d_bnd[0], &d_bnd[1],
                     &d_flags, &d_coef, &d_rgb, &d_export, &d_units, &d_luts,
                     &d_segsub, &d_segunit, &d_x_img, &d_x_sub_base, &d_x_unit_base, &d_xunits, &d_pimgs, &d_ptables};
    for (DevBuf *b : all) b->destroy(be);
    for (int f = 0; f < PF_COUNT; ++f) { d_grp_img[f].destroy(be); d_grp_pre[f].destroy(be); }
    be.host_release(h_out); be.host_release(h_flags);
    h_out = nullptr; h_flags = nullptr; h_out_cap = 0;
  }

  static constexpr size_t kFlagWords = kFixLaunches + kMaxExtraFix + 2;
  bool reserve_host(int n) {
    if (!h_flags) h_flags = static_cast<uint32_t *>(be.host_alloc(4 * kFlagWords));
    if ((size_t)n > h_out_cap) {
      be.host_release(h_out);
      h_out_cap = (size_t)n + (size_t)n / 4 + 16;
      h_out = static_cast<ImageOut *>(be.host_alloc(sizeof(ImageOut) * h_out_cap));
    }
    return h_flags && h_out;
  }
  void enqueue_small_results() {      // flags + per-image ImageOut, asynchronous
    be.d2h(h_flags, d_flags.p, 4 * kFlagWords);
    be.d2h(h_out, d_out.p, sizeof(ImageOut) * plan.n);
    results_enqueued = true;
  }

  // plan + H2D of the scans and descriptors
  int prepare(const b200jpeg_image_desc *descs, int n, const uint8_t *const *files, const size_t *sizes, bool progressive = false) {
    prepared = ran = false;
    prog_max_scans = 0;
    if (n <= 0 || !descs || !files || !sizes) return B200JPEG_E_ARG;
    const int rc = make_plan(descs, n, sizes, plan, progressive ? (int)kForceSegment : force_path, progressive);
    if (rc) return rc;
    const BatchPlan &P = plan;
    bool ok = true;
    ok &= d_raw.reserve(be, P.raw_bytes + 1024);
    ok &= d_us.reserve(be, P.raw_bytes + (size_t)kChunkWords * 4 + 1024);
    ok &= d_luts.reserve(be, sizeof(uint32_t) * (size_t)kHTabsPerSet * kHTab * P.tsets.size());
    if (!P.units.empty()) ok &= d_units.reserve(be, sizeof(uint2) * P.units.size());
    ok &= d_imgs.reserve(be, sizeof(ImageDev) * n);
    ok &= d_tsets.reserve(be, sizeof(TableSet) * P.tsets.size());
    ok &= d_out.reserve(be, sizeof(ImageOut) * n);
    ok &= d_tile_base.reserve(be, 4 * (size_t)(n + 1));
    ok &= d_tile_cnt.reserve(be, 4 * (size_t)P.ntiles);
    ok &= d_tile_pre.reserve(be, 8 * (size_t)P.ntiles);
    ok &= d_seg.reserve(be, 4 * (size_t)P.nsegs);
    if (P.path == kPathSync) {
      ok &= d_sub_e.reserve(be, sizeof(SubState) * (size_t)P.x.nsub_slots);
      ok &= d_sub_used.reserve(be, sizeof(SubState) * (size_t)P.x.nsub_slots);
      ok &= d_sub_info.reserve(be, sizeof(SubInfo) * (size_t)P.x.nsub_slots);
      ok &= d_cta_tot.reserve(be, sizeof(SubInfo) * (size_t)P.x.ncta);
      ok &= d_cta_pre.reserve(be, sizeof(SubInfo) * (size_t)P.x.ncta);
      ok &= d_bnd[0].reserve(be, sizeof(SubState) * (size_t)P.x.ncta);
      ok &= d_bnd[1].reserve(be, sizeof(SubState) * (size_t)P.x.ncta);
      ok &= d_segsub.reserve(be, 4 * (size_t)P.nsegs);
      ok &= d_segunit.reserve(be, 4 * (size_t)P.nsegs);
      ok &= d_x_img.reserve(be, 4 * (size_t)n);
      ok &= d_x_sub_base.reserve(be, 4 * (size_t)(n + 1));
      ok &= d_x_unit_base.reserve(be, 4 * (size_t)(n + 1));
      ok &= d_xunits.reserve(be, sizeof(XUnit) * (size_t)P.x.nunit_slots);
    }
    ok &= d_flags.reserve(be, 4 * (size_t)(kFixLaunches + kMaxExtraFix + 2));
    ok &= d_coef.reserve(be, 128 * (size_t)P.nblocks);
    ok &= d_rgb.reserve(be, P.rgb_bytes);
    for (int f = 0; f < PF_COUNT; ++f) if (!P.grp_img[f].empty()) {
      ok &= d_grp_img[f].reserve(be, 4 * P.grp_img[f].size());
      ok &= d_grp_pre[f].reserve(be, 4 * P.grp_pre[f].size());
    }
    ok &= reserve_host(n);
    if (!ok) return be.ok() ? B200JPEG_E_NOMEM : B200JPEG_E_CUDA;
    h2d_bytes = 0;
    // scans: one copy per image, adjacent source ranges merged into one
    for (int i = 0; i < n;) {
      const uint8_t *src = files[i] + descs[i].scan_offset;
      size_t len = (size_t)descs[i].scan_length;
      int j = i + 1;
      // merge when the next image's scan sits exactly where the device layout expects it
      while (j < n && files[j] + descs[j].scan_offset == src + (P.imgs[j].raw_off - P.imgs[i].raw_off)) {
        len = (size_t)(P.imgs[j].raw_off - P.imgs[i].raw_off) + (size_t)descs[j].scan_length;
        ++j;
      }
      be.h2d(d_raw.as<uint8_t>() + P.imgs[i].raw_off, src, len);
      h2d_bytes += (int64_t)len;
      i = j;
    }
    be.h2d(d_imgs.p, P.imgs.data(), sizeof(ImageDev) * n);
    be.h2d(d_tsets.p, P.tsets.data(), sizeof(TableSet) * P.tsets.size());
    be.h2d(d_tile_base.p, P.tile_base.data(), 4 * (size_t)(n + 1));
    if (P.path == kPathSync) {
      be.h2d(d_x_img.p, P.x_img.data(), 4 * (size_t)n);
      be.h2d(d_x_sub_base.p, P.x_sub_base.data(), 4 * (size_t)(n + 1));
      be.h2d(d_x_unit_base.p, P.x_unit_base.data(), 4 * (size_t)(n + 1));
      h2d_bytes += (int64_t)(4 * (size_t)(3 * n + 2));
    }
    if (!P.units.empty()) { be.h2d(d_units.p, P.units.data(), sizeof(uint2) * P.units.size()); h2d_bytes += (int64_t)(sizeof(uint2) * P.units.size()); }
    h2d_bytes += (int64_t)(sizeof(ImageDev) * n + sizeof(TableSet) * P.tsets.size() + 4 * (size_t)(n + 1));
    for (int f = 0; f < PF_COUNT; ++f) if (!P.grp_img[f].empty()) {
      be.h2d(d_grp_img[f].p, P.grp_img[f].data(), 4 * P.grp_img[f].size());
      be.h2d(d_grp_pre[f].p, P.grp_pre[f].data(), 4 * P.grp_pre[f].size());
      h2d_bytes += (int64_t)(4 * (P.grp_img[f].size() + P.grp_pre[f].size()));
    }
    if (!be.ok()) return B200JPEG_E_CUDA;
    prepared = true;
    return 0;
  }

  // Progressive files: the plan comes from the frames; the scans and their tables go to the device beside it.
  int prepare_progressive(const b200jpeg_progressive_desc *descs, int n, const uint8_t *const *files, const size_t *sizes) {
    if (n <= 0 || !descs || !files || !sizes) return B200JPEG_E_ARG;
    std::vector<b200jpeg_image_desc> frames((size_t)n);
    std::vector<ProgImageDev> pimgs((size_t)n);
    std::vector<HuffSpec> tables;
    int max_scans = 0, max_levels = 0;
    for (int i = 0; i < n; ++i) {
      const b200jpeg_progressive_desc &d = descs[i];
      if (d.nscans < 1 || d.nscans > kProgMaxScans || d.ntables < 0 || d.ntables > B200JPEG_PROG_MAX_TABLES) return B200JPEG_E_ARG;
      frames[(size_t)i] = d.frame;
      ProgImageDev &pi = pimgs[(size_t)i];
      memset(&pi, 0, sizeof pi);
      pi.nscans = d.nscans; pi.width = d.frame.image_width; pi.height = d.frame.image_height;
      const int tbase = (int)tables.size();
      for (int t = 0; t < d.ntables; ++t) {
        const b200jpeg_huffman_table &src = d.tables[t];
        if (src.offset[0] != 0 || src.offset[16] > 176) return B200JPEG_E_DHT;
        for (int l = 1; l <= 16; ++l) if (src.offset[l] < src.offset[l - 1]) return B200JPEG_E_DHT;
        HuffSpec h; memset(&h, 0, sizeof h);
        memcpy(h.offset, src.offset, 17); memcpy(h.symbols, src.symbols, 176); h.set = 1;
        tables.push_back(h);
      }
      for (int k = 0; k < d.nscans; ++k) {
        const b200jpeg_scan &S = d.scans[k];
        ProgScanDev &o = pi.scans[k];
        if (S.ns < 1 || S.ns > d.frame.numComponents || S.ss < 0 || S.ss > S.se || S.se > 63 || (S.ss == 0 && S.se != 0) ||
            (S.ss != 0 && S.ns != 1) || S.al < 0 || S.al > 13 || S.ah < 0 || S.ah > 13 || S.restart_interval < 0)
          return B200JPEG_E_SOS;
        if (S.offset < d.frame.scan_offset || S.length < 0 || (size_t)(S.offset + S.length) > sizes[i]) return B200JPEG_E_ARG;
        o.ns = S.ns; o.ss = S.ss; o.se = S.se; o.ah = S.ah; o.al = S.al; o.restart_interval = S.restart_interval;
        o.offset = (uint32_t)(S.offset - d.frame.scan_offset); o.length = (uint32_t)S.length;
        for (int j = 0; j < 3; ++j) {
          o.comp[j] = 0; o.dc_table[j] = o.ac_table[j] = -1;
          if (j >= S.ns) continue;
          if (S.comp[j] < 0 || S.comp[j] >= d.frame.numComponents) return B200JPEG_E_SOS;
          o.comp[j] = S.comp[j];
          const bool need_dc = S.ss == 0 && S.ah == 0, need_ac = S.ss != 0;
          if ((need_dc && (S.dc_table[j] < 0 || S.dc_table[j] >= d.ntables)) || (need_ac && (S.ac_table[j] < 0 || S.ac_table[j] >= d.ntables)))
            return B200JPEG_E_TABLE_MISSING;
          if (need_dc) o.dc_table[j] = tbase + S.dc_table[j];
          if (need_ac) o.ac_table[j] = tbase + S.ac_table[j];
        }
        // dependency level: after every earlier scan that shares a component and whose band overlaps this one
        o.level = 0;
        for (int t = 0; t < k; ++t) {
          const ProgScanDev &e = pi.scans[t];
          bool shares = false;
          for (int a = 0; a < o.ns; ++a) for (int b = 0; b < e.ns; ++b) shares = shares || o.comp[a] == e.comp[b];
          if (shares && o.ss <= e.se && e.ss <= o.se && e.level + 1 > o.level) o.level = e.level + 1;
        }
        if (o.level + 1 > max_levels) max_levels = o.level + 1;
      }
      if (d.nscans > max_scans) max_scans = d.nscans;
    }
    const int rc = prepare(frames.data(), n, files, sizes, true);
    if (rc) return rc;
    prepared = false;
    bool ok = d_pimgs.reserve(be, sizeof(ProgImageDev) * (size_t)n) && d_ptables.reserve(be, sizeof(HuffSpec) * (tables.size() + 1));
    if (!ok) return be.ok() ? B200JPEG_E_NOMEM : B200JPEG_E_CUDA;
    be.h2d(d_pimgs.p, pimgs.data(), sizeof(ProgImageDev) * (size_t)n);
    if (!tables.empty()) be.h2d(d_ptables.p, tables.data(), sizeof(HuffSpec) * tables.size());
    be.sync();                                     // pimgs / tables are host temporaries
    h2d_bytes += (int64_t)(sizeof(ProgImageDev) * (size_t)n + sizeof(HuffSpec) * tables.size());
    if (!be.ok()) return B200JPEG_E_CUDA;
    prog_max_scans = max_scans; prog_levels = max_levels;
    prepared = true;
    return 0;
  }

  int run_progressive() {
    if (!prepared || prog_max_scans <= 0) return B200JPEG_E_STATE;
    const BatchPlan &P = plan;
    launches = 0; sync_rounds = 0;
    results_enqueued = false;
    mark(0);
    be.fill(d_out.p, 0xff, sizeof(ImageOut) * P.n);
    be.fill(d_flags.p, 0, 4 * (size_t)(kFixLaunches + kMaxExtraFix + 2));
    B2_LAUNCH(k_reset_out, (P.n + 255) / 256, 256, 0, be.stream, d_out.as<ImageOut>(), P.n);
    ++launches;
    be.fill(d_coef.p, 0, 128 * (size_t)P.nblocks);   // scans only ever write the coefficients they carry (MCU[] starts at zero, jpeg.cpp:407)
    mark(1); mark(2);
    for (int lv = 0; lv < prog_levels; ++lv) {        // one warp per (image, scan); a level's scans do not depend on each other
      B2_LAUNCH(k_prog_scan, P.n * prog_max_scans, 32, 0, be.stream, d_imgs.as<ImageDev>(), d_pimgs.as<ProgImageDev>(), P.n,
                prog_max_scans, lv, d_ptables.as<HuffSpec>(), d_raw.as<uint8_t>(), d_coef.as<int16_t>(), d_out.as<ImageOut>());
      ++launches;
    }
    mark(3);
    launch_pixels();
    mark(5);
    be.check("kernel launch");
    ran = true;
    return be.ok() ? 0 : B200JPEG_E_CUDA;
  }

  void mark(int k) {
#ifndef B200JPEG_EMU
    if (timing) cudaEventRecord(ev[k], be.stream);
#else
    (void)k;
#endif
  }

  // Tensor map of a uniform group's RGB output: u32 elements, dims (pitch/4, height, images), strides (pitch,
  // image stride); box = one pixel tile.  The pixel kernel then stores a tile with one TMA instruction.
  B2TensorMap tmap[PF_COUNT];
  template <int HS, int VS, int NC>
  void make_tmap(int f) {
    plan.uni[f].tmap_ok = 0;
#ifndef B200JPEG_EMU
    typedef PixCfg<HS, VS, NC> C;
    const PixUni &u = plan.uni[f];
    if (!u.enabled || plan.grp_img[f].empty()) return;
    typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                 const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn encode = nullptr;
    if (!encode) {
      void *fn = nullptr;
      cudaDriverEntryPointQueryResult q;
      if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) { cudaGetLastError(); return; }
      encode = reinterpret_cast<EncodeFn>(fn);
    }
    const uint64_t nimg = plan.grp_img[f].size();
    const uint64_t istride = nimg > 1 ? u.rgb_stride : align_up((size_t)u.pitch * u.height, 256);
    if ((u.pitch & 15u) || (istride & 15u) || C::kRowBytes / 4 > 256) return;
    const cuuint64_t dims[3] = {u.pitch / 4, (cuuint64_t)u.height, nimg};
    const cuuint64_t strides[2] = {u.pitch, istride};
    const cuuint32_t box[3] = {(cuuint32_t)(C::kRowBytes / 4), (cuuint32_t)C::kH, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    const CUresult r = encode(&tmap[f].map, CU_TENSOR_MAP_DATA_TYPE_UINT32, 3, d_rgb.as<uint8_t>() + u.rgb_off0, dims, strides, box, estr,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r == CUDA_SUCCESS) plan.uni[f].tmap_ok = 1;
#endif
  }

  template <int HS, int VS, int NC>
  void launch_pixel(int f) {
    if (plan.grp_img[f].empty()) return;
    make_tmap<HS, VS, NC>(f);
    const BatchPlan &P = plan;
    const int ngrp = (int)P.grp_img[f].size();
    const int tiles = P.grp_pre[f].back();
    const size_t smem = PixCfg<HS, VS, NC>::kSmemBytes;
    // persistent: one CTA slot per resident CTA of the device (pix_grid_cap[f], set at create), never more
    // CTAs than there are groups of kPixWarps tiles
    int grid = (tiles + kPixWarps - 1) / kPixWarps;
    if (pix_grid_cap[f] > 0 && grid > pix_grid_cap[f]) grid = pix_grid_cap[f];
    if (grid <= 0) return;
    B2_LAUNCH((k_pixel<HS, VS, NC>), grid, kPixThreads, smem, be.stream, d_imgs.as<ImageDev>(),
              d_grp_img[f].as<int32_t>(), d_grp_pre[f].as<int32_t>(), ngrp, P.uni[f], d_tsets.as<TableSet>(),
              d_out.as<ImageOut>(), d_coef.as<int16_t>(), d_rgb.as<uint8_t>(), tmap[f]);
    ++launches;
  }

  void launch_fix(int k) {   // k-th repair launch: reads boundary states k-1, writes k
    const BatchPlan &P = plan;
    B2_LAUNCH((k_xsync<false>), P.x.ncta, kXT, xsync_smem_bytes(P.nslots_max), be.stream, P.x, d_imgs.as<ImageDev>(),
              d_x_sub_base.as<int32_t>(), d_x_img.as<int32_t>(), d_tsets.as<TableSet>(), d_luts.as<uint32_t>(),
              d_out.as<ImageOut>(), d_us.as<uint8_t>(), d_seg.as<uint32_t>(), d_segsub.as<uint32_t>(),
              d_sub_e.as<SubState>(), d_sub_used.as<SubState>(), d_sub_info.as<SubInfo>(), d_cta_tot.as<SubInfo>(),
              d_bnd[(k + 1) & 1].as<SubState>(), d_bnd[k & 1].as<SubState>(), d_flags.as<uint32_t>() + k);
    ++launches; ++sync_rounds;
  }

  void launch_seg(bool xunits) {   // the one writer of coefficient blocks: restart segments, or the units of the self-synchronising path
    const BatchPlan &P = plan;
    if (P.nunit_ctas <= 0) return;
    B2_LAUNCH(k_entropy_seg, P.nunit_ctas, kSegThreads, segdec_smem_bytes(), be.stream,
              xunits ? (const uint2 *)nullptr : d_units.as<uint2>(), d_imgs.as<ImageDev>(), d_tsets.as<TableSet>(),
              d_luts.as<uint32_t>(), d_seg.as<uint32_t>(), d_us.as<uint8_t>(), d_out.as<ImageOut>(), d_coef.as<int16_t>(),
              d_x_unit_base.as<int32_t>(), d_x_img.as<int32_t>(), P.n, d_xunits.as<XUnit>(), d_segunit.as<uint32_t>());
    ++launches;
  }

  void launch_tail(int last_fix) {   // scan + units + write + pixel
    const BatchPlan &P = plan;
    (void)last_fix;
    B2_LAUNCH(k_xscan, 1, 32, 0, be.stream, P.x.ncta, d_cta_tot.as<SubInfo>(), d_cta_pre.as<SubInfo>());
    ++launches;
    mark(2);
    B2_LAUNCH(k_xunits, P.x.ncta, kXT, 0, be.stream, P.x, d_imgs.as<ImageDev>(), d_x_sub_base.as<int32_t>(),
              d_x_img.as<int32_t>(), d_x_unit_base.as<int32_t>(), d_out.as<ImageOut>(), d_seg.as<uint32_t>(),
              d_segsub.as<uint32_t>(), d_segunit.as<uint32_t>(), d_sub_e.as<SubState>(), d_sub_used.as<SubState>(),
              d_sub_info.as<SubInfo>(), d_cta_pre.as<SubInfo>(), d_xunits.as<XUnit>());
    ++launches;
    launch_seg(true);
    mark(3);
    launch_pixels();
  }

  void launch_pixels() {
    launch_pixel<1, 1, 3>(PF_444);
    launch_pixel<2, 1, 3>(PF_422);
    launch_pixel<1, 2, 3>(PF_440);
    launch_pixel<2, 2, 3>(PF_420);
    launch_pixel<1, 1, 1>(PF_GRAY);
    mark(4);
  }

  // every kernel of the path, asynchronously on be.stream
  int run() {
    if (!prepared) return B200JPEG_E_STATE;
    const BatchPlan &P = plan;
    launches = 0; sync_rounds = 0;
    results_enqueued = false;
    if (copies_only) {                   // the copies of a batch without its kernels: flags read "converged"
      be.fill(d_flags.p, 0, 4 * (size_t)(kFixLaunches + kMaxExtraFix + 2));
      ran = true;
      return be.ok() ? 0 : B200JPEG_E_CUDA;
    }
    mark(0);
    // ImageOut: end_pos/err_pos = kNone, the rest 0
    be.fill(d_out.p, 0xff, sizeof(ImageOut) * P.n);
    be.fill(d_flags.p, 0, 4 * (size_t)(kFixLaunches + kMaxExtraFix + 2));
    B2_LAUNCH(k_reset_out, (P.n + 255) / 256, 256, 0, be.stream, d_out.as<ImageOut>(), P.n);
    ++launches;
    B2_LAUNCH(k_build_luts, (int)P.tsets.size() * kHTabsPerSet, 256, 0, be.stream, d_tsets.as<TableSet>(), d_luts.as<uint32_t>());
    ++launches;
    if (P.ntiles > 0) {
      B2_LAUNCH(k_unstuff_count, P.ntiles, kUThreads, 0, be.stream, d_imgs.as<ImageDev>(), d_tile_base.as<int32_t>(),
                P.n, d_raw.as<uint8_t>(), d_tile_cnt.as<uint32_t>(), d_out.as<ImageOut>());
      B2_LAUNCH(k_unstuff_scan, (P.n + 7) / 8, 256, 0, be.stream, d_imgs.as<ImageDev>(), P.n,
                d_tile_cnt.as<uint32_t>(), d_tile_pre.as<uint2>());
      B2_LAUNCH(k_unstuff_scatter, P.ntiles, kUThreads, 0, be.stream, d_imgs.as<ImageDev>(),
                d_tile_base.as<int32_t>(), P.n, d_raw.as<uint8_t>(), d_tile_pre.as<uint2>(), d_us.as<uint8_t>(),
                d_seg.as<uint32_t>(), d_out.as<ImageOut>());
      launches += 3;
    }
    B2_LAUNCH(k_unstuff_finish, (P.n + 255) / 256, 256, 0, be.stream, d_imgs.as<ImageDev>(), P.n, d_seg.as<uint32_t>(),
              d_out.as<ImageOut>());
    ++launches;
    mark(1);
    if (P.path == kPathSegment) {
      mark(2);
      launch_seg(false);
      mark(3);
      launch_pixels();
      mark(5);
      be.check("kernel launch");
      ran = true;
      return be.ok() ? 0 : B200JPEG_E_CUDA;
    }
    if (P.x.ncta > 0) {
      B2_LAUNCH(k_xplan, (P.n + 7) / 8, 256, 0, be.stream, P.x, d_imgs.as<ImageDev>(), d_out.as<ImageOut>(),
                d_seg.as<uint32_t>(), d_segsub.as<uint32_t>(), d_segunit.as<uint32_t>());
      ++launches;
      B2_LAUNCH((k_xsync<true>), P.x.ncta, kXT, xsync_smem_bytes(P.nslots_max), be.stream, P.x, d_imgs.as<ImageDev>(),
                d_x_sub_base.as<int32_t>(), d_x_img.as<int32_t>(), d_tsets.as<TableSet>(), d_luts.as<uint32_t>(),
                d_out.as<ImageOut>(), d_us.as<uint8_t>(), d_seg.as<uint32_t>(), d_segsub.as<uint32_t>(),
                d_sub_e.as<SubState>(), d_sub_used.as<SubState>(), d_sub_info.as<SubInfo>(), d_cta_tot.as<SubInfo>(),
                (const SubState *)nullptr, d_bnd[1].as<SubState>(), d_flags.as<uint32_t>());
      ++launches;
      for (int k = 0; k < kFixLaunches; ++k) launch_fix(k);   // launch k reads d_bnd[(k+1)&1], writes d_bnd[k&1]
      launch_tail(kFixLaunches - 1);
    } else { mark(2); mark(3); mark(4); }
    mark(5);
    be.check("kernel launch");
    ran = true;
    return be.ok() ? 0 : B200JPEG_E_CUDA;
  }

  // wait; if self-synchronisation had not settled, keep repairing and redo the tail
  int finish() {
    if (!ran) return B200JPEG_E_STATE;
    const BatchPlan &P = plan;
    int k = kFixLaunches - 1;
    for (;;) {
      if (!results_enqueued) enqueue_small_results();
      be.sync();
      results_enqueued = false;
      if (!be.ok()) return B200JPEG_E_CUDA;
      if (P.path != kPathSync || P.x.ncta == 0 || h_flags[k] == 0) break;
      if (k + 1 >= kFixLaunches + kMaxExtraFix) {
        B2_LAUNCH(k_raise_status, (P.n + 255) / 256, 256, 0, be.stream, d_out.as<ImageOut>(), P.n, (uint32_t)ST_NOT_CONVERGED);
        launch_pixels();                 // the images are dead now (kStDead): their pixels leave as zeros
        enqueue_small_results();
        be.sync();
        results_enqueued = false;
        break;
      }
      ++k;
      launch_fix(k);
      launch_tail(k);
    }
#ifndef B200JPEG_EMU
    if (timing) {
      for (int s = 0; s < 4; ++s) cudaEventElapsedTime(&last_ms[s], ev[s], ev[s + 1]);
      cudaEventElapsedTime(&last_ms[B200JPEG_T_TOTAL], ev[0], ev[5]);
    }
#endif
    return be.ok() ? 0 : B200JPEG_E_CUDA;
  }

  // asynchronous D2H of every RGB image on the pipeline's stream (no wait)
  void enqueue_rgb(uint8_t *const *rgb_out) {
    const BatchPlan &P = plan;
    if (!results_enqueued) enqueue_small_results();
    if (rgb_out) for (int i = 0; i < P.n;) {
      if (!rgb_out[i]) { ++i; continue; }
      const ImageDev &im = P.imgs[i];
      const uint8_t *src = d_rgb.as<uint8_t>() + im.rgb_off;
      const size_t row = (size_t)3 * im.width;
      if (im.rgb_pitch != row) {
        be.d2h_2d(rgb_out[i], row, src, im.rgb_pitch, row, im.height);
        d2h_bytes += (int64_t)(row * im.height);
        ++i;
        continue;
      }
      // dense rows: images that follow each other both on the device and in the caller's memory leave as ONE copy
      size_t len = row * im.height;
      int j = i + 1;
      while (j < P.n && rgb_out[j] && P.imgs[j].rgb_pitch == (size_t)3 * P.imgs[j].width &&
             P.imgs[j].rgb_off == im.rgb_off + len && rgb_out[j] == rgb_out[i] + len && len < ((size_t)64 << 20)) {
        len += (size_t)3 * P.imgs[j].width * P.imgs[j].height;
        ++j;
      }
      be.d2h(rgb_out[i], src, len);
      d2h_bytes += (int64_t)len;
      i = j;
    }
  }

  // second half of a pipelined decode: the RGB copies were enqueued right behind the kernels
  // (enqueue_rgb); wait, and only if the self-synchronising decoder needed more repair rounds than
  // were launched (rare) copy again after the repair
  int complete(uint8_t *const *rgb_out, int32_t *status) {
    const int before = launches;
    const int rc = finish();
    if (rc) return rc;
    if (launches != before) { enqueue_rgb(rgb_out); be.sync(); }
    if (status) for (int i = 0; i < plan.n; ++i) status[i] = (int32_t)h_out[i].status;
    return be.ok() ? 0 : B200JPEG_E_CUDA;
  }

  int fetch(uint8_t *const *rgb_out, int16_t *const *coef_out, int32_t *status) {
    const int rc = finish();
    if (rc) return rc;
    const BatchPlan &P = plan;
    d2h_bytes = 0;
    if (status) for (int i = 0; i < P.n; ++i) status[i] = (int32_t)h_out[i].status;
    enqueue_rgb(rgb_out);
    if (coef_out) for (int i = 0; i < P.n; ++i) if (coef_out[i]) {
      const ImageDev &im = P.imgs[i];
      const size_t bytes = (size_t)P.bwr[i] * P.bhr[i] * 3 * 64 * sizeof(int16_t);
      if (!d_export.reserve(be, bytes)) return B200JPEG_E_NOMEM;
      be.fill(d_export.p, 0, bytes);   // MCU[] is value-initialised, jpeg.cpp:407
      if (h_out[i].status & kStDead) {  // blocks of this image were not all written this batch: zeros, never an earlier batch's
        be.d2h(coef_out[i], d_export.p, bytes);
        be.sync();
        d2h_bytes += (int64_t)bytes;
        continue;
      }
      #ifdef B200JPEG_EMU
      const int egrid = 4;
#else
      const int egrid = im.nblocks / 4 + 1 < 2368 ? im.nblocks / 4 + 1 : 2368;
#endif
      B2_LAUNCH(k_export_coef, egrid, 256, 0, be.stream, im, P.bwr[i], d_coef.as<int16_t>(),
                d_export.as<int16_t>());
      be.d2h(coef_out[i], d_export.p, bytes);
      be.sync();
      d2h_bytes += (int64_t)bytes;
    }
    be.sync();
    return be.ok() ? 0 : B200JPEG_E_CUDA;
  }
};

}  // namespace b2
