// b200jpeg.cu -- the C ABI of include/b200jpeg.h over the CUDA pipeline (sm_100a only).
// No CPU path: without a usable GPU every decode entry point returns B200JPEG_E_CUDA.
#include "pipeline_run.h"

#include <chrono>
#include <mutex>
#include <new>
#include <thread>

struct b200jpeg_ctx {
  int device = 0;
  b2::Pipeline pipe;         // the batch of prepare / run / fetch; even chunks of decode_batch
  static constexpr int kExtra = 3;
  b2::Pipeline extra[kExtra]; // further chunks of decode_batch in flight (own streams and arenas)
  bool agg_valid = false;    // counters of the last pipelined decode_batch
  b200jpeg_batch_stats agg;
  std::string last_error;
};

namespace {

template <class K>
bool allow_smem(K kernel, size_t bytes, std::string &err) {
  cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) { err = std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e); return false; }
  return true;
}

// k_pixel is persistent: allow its shared memory and size its grid to the CTAs the device keeps resident
template <int HS, int VS, int NC>
bool setup_pixel(b200jpeg_ctx *c, int f, int sms) {
  const size_t smem = b2::PixCfg<HS, VS, NC>::kSmemBytes;
  if (!allow_smem(b2::k_pixel<HS, VS, NC>, smem, c->last_error)) return false;
  int per_sm = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, b2::k_pixel<HS, VS, NC>, b2::kPixThreads, smem);
  if (e != cudaSuccess || per_sm < 1) { c->last_error = "k_pixel does not fit on an SM"; return false; }
  c->pipe.pix_grid_cap[f] = per_sm * sms;
  for (b2::Pipeline &x : c->extra) x.pix_grid_cap[f] = per_sm * sms;
  return true;
}

int fail2(b200jpeg_ctx *c, b2::Pipeline &p, int rc) {
  if (rc == B200JPEG_E_CUDA && !p.be.err.empty()) c->last_error = p.be.err;
  return rc;
}

int fail(b200jpeg_ctx *c, int rc) {
  if (rc == B200JPEG_E_CUDA && !c->pipe.be.err.empty()) c->last_error = c->pipe.be.err;
  return rc;
}

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace

extern "C" {

int b200jpeg_create(int device, b200jpeg_ctx **out) {
  if (!out) return B200JPEG_E_ARG;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) { cudaGetLastError(); return B200JPEG_E_CUDA; }
  b200jpeg_ctx *c = new (std::nothrow) b200jpeg_ctx();
  if (!c) return B200JPEG_E_NOMEM;
  c->device = device;
  DeviceGuard g(device);
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || prop.major < 10) {
    delete c;   // the binary holds sm_100a code only
    return B200JPEG_E_CUDA;
  }
  if (cudaStreamCreateWithFlags(&c->pipe.be.stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return B200JPEG_E_CUDA; }
  for (b2::Pipeline &x : c->extra)
    if (cudaStreamCreateWithFlags(&x.be.stream, cudaStreamNonBlocking) != cudaSuccess) { b200jpeg_destroy(c); return B200JPEG_E_CUDA; }
  for (int i = 0; i <= B200JPEG_T_COUNT; ++i) {
    cudaEventCreate(&c->pipe.ev[i]);
    for (b2::Pipeline &x : c->extra) cudaEventCreate(&x.ev[i]);
  }
  const size_t es = b2::xsync_smem_bytes(b2::kHTabsPerSet);
  bool ok = allow_smem(b2::k_xsync<true>, es, c->last_error) && allow_smem(b2::k_xsync<false>, es, c->last_error) &&
            allow_smem(b2::k_entropy_seg, b2::segdec_smem_bytes(), c->last_error);
  ok = ok && setup_pixel<1, 1, 3>(c, b2::PF_444, prop.multiProcessorCount) && setup_pixel<2, 1, 3>(c, b2::PF_422, prop.multiProcessorCount) &&
       setup_pixel<1, 2, 3>(c, b2::PF_440, prop.multiProcessorCount) && setup_pixel<2, 2, 3>(c, b2::PF_420, prop.multiProcessorCount) &&
       setup_pixel<1, 1, 1>(c, b2::PF_GRAY, prop.multiProcessorCount);
  if (!ok) { b200jpeg_destroy(c); return B200JPEG_E_CUDA; }
  *out = c;
  return B200JPEG_OK;
}

void b200jpeg_destroy(b200jpeg_ctx *c) {
  if (!c) return;
  DeviceGuard g(c->device);
  for (b2::Pipeline *p : {&c->pipe, &c->extra[0], &c->extra[1], &c->extra[2]}) {
    if (p->be.stream) cudaStreamSynchronize(p->be.stream);
    p->destroy();
    for (int i = 0; i <= B200JPEG_T_COUNT; ++i) if (p->ev[i]) cudaEventDestroy(p->ev[i]);
    if (p->be.stream) cudaStreamDestroy(p->be.stream);
  }
  delete c;
}

const char *b200jpeg_last_error(const b200jpeg_ctx *c) { return c ? c->last_error.c_str() : "null context"; }

int b200jpeg_batch_prepare(b200jpeg_ctx *c, const b200jpeg_image_desc *descs, int n, const uint8_t *const *files,
                           const size_t *sizes) {
  if (!c) return B200JPEG_E_ARG;
  DeviceGuard g(c->device);
  c->pipe.be.err.clear();
  c->agg_valid = false;
  return fail(c, c->pipe.prepare(descs, n, files, sizes));
}

int b200jpeg_batch_run(b200jpeg_ctx *c) {
  if (!c) return B200JPEG_E_ARG;
  DeviceGuard g(c->device);
  return fail(c, c->pipe.run());
}

int b200jpeg_batch_fetch(b200jpeg_ctx *c, uint8_t *const *rgb_out, int16_t *const *coef_out, int32_t *status) {
  if (!c) return B200JPEG_E_ARG;
  DeviceGuard g(c->device);
  return fail(c, c->pipe.fetch(rgb_out, coef_out, status));
}

int b200jpeg_sync(b200jpeg_ctx *c) {
  if (!c) return B200JPEG_E_ARG;
  DeviceGuard g(c->device);
  c->pipe.be.sync();
  return fail(c, c->pipe.be.ok() ? 0 : B200JPEG_E_CUDA);
}

// Host buffers in, host buffers out.  Large batches are cut into chunks that rotate over kSlots pipelines (own
// stream, own arenas): while the RGB of earlier chunks travels device -> host, the files of later ones travel
// host -> device and their kernels run, so the call is bounded by the slower PCIe direction (the RGB) instead of the
// sum of copy + decode + copy.  Why four pipelines and small chunks: with the device -> host link saturated, host ->
// device copies crawl (their read requests queue behind the posted writes: 19 MB took 2.1 ms in the CUPTI trace), so
// a chunk needs about as long as a chunk's RGB copy to get ready -- with only two pipelines the RGB engine ran dry
// for 0.6 ms per pair of chunks; and the first RGB byte leaves once the first (small) chunk is decoded.
// The entropy decoder is chosen ONCE per call, for what is resident on the device at a time: nslots chunks (not one chunk,
// which would ignore its neighbours in flight; not the whole call, whose chunks never meet: a 16-image chunk handed to the
// segment decoder takes a full chain's time, 1.7 ms, for 1088 segments -- measured 31.39 ms per C2 call against 31.12).
// copies_only: the same copies with no kernel in between (b200jpeg_copy_floor).
static int decode_batch_impl(b200jpeg_ctx *c, const b200jpeg_image_desc *descs, int n, const uint8_t *const *files,
                             const size_t *sizes, uint8_t *const *rgb_out, int16_t *const *coef_out, int32_t *status,
                             bool copies_only) {
  if (!c || n <= 0 || !descs || !files || !sizes) return B200JPEG_E_ARG;
  constexpr int kSlots = 1 + b200jpeg_ctx::kExtra;
  int chunk = (n + 15) / 16;
  if (chunk < 8) chunk = 8;
  if (chunk > 32) chunk = 32;
  {
    // ... in images; in BYTES a chunk should carry about 100 MB of pixels (16 pictures of 1080p): with thousands of small
    // pictures the per-chunk costs (stream waits, plan, launches) otherwise show -- 16,384 x 512^2 in chunks of 32: 355 ms,
    // in chunks of 128: 332 ms against a copy floor of 314.  At least eight chunks stay, for the overlap.
    uint64_t px = 0;
    for (int i = 0; i < n; ++i) px += (uint64_t)descs[i].image_width * descs[i].image_height;
    const uint64_t per_image = 3 * px / (uint64_t)n + 1;
    int by_bytes = (int)(((uint64_t)100 << 20) / per_image);
    if (by_bytes > 256) by_bytes = 256;
    if (by_bytes > n / 8) by_bytes = n / 8;
    if (by_bytes > chunk) chunk = by_bytes;
  }
  int nslots = kSlots;
#ifdef B200JPEG_TUNING                                 // measurement builds only (scripts/dev): chunking knobs
  if (const char *ev = getenv("B200JPEG_CHUNK")) { const int v = atoi(ev); if (v > 0) chunk = v; }
  if (const char *ev = getenv("B200JPEG_SLOTS")) { const int v = atoi(ev); if (v >= 2 && v <= kSlots) nslots = v; }
#endif
  if ((coef_out || n <= chunk) && !copies_only) {
    int rc = b200jpeg_batch_prepare(c, descs, n, files, sizes);
    if (rc) return rc;
    rc = b200jpeg_batch_run(c);
    if (rc) return rc;
    return b200jpeg_batch_fetch(c, rgb_out, coef_out, status);
  }
  DeviceGuard g(c->device);
  b2::Pipeline *slot[kSlots] = {&c->pipe, &c->extra[0], &c->extra[1], &c->extra[2]};
  int first[kSlots] = {0, 0, 0, 0};
  bool busy[kSlots] = {false, false, false, false};
  b200jpeg_batch_stats a;
  memset(&a, 0, sizeof a);
  const int user_force = c->pipe.force_path;
  const int window = n < chunk * nslots ? n : chunk * nslots;
  const int call_path = b2::choose_path(descs, window, user_force) == b2::kPathSegment ? b2::kForceSegment : b2::kForceSync;
  for (b2::Pipeline *p : slot) { p->force_path = call_path; p->copies_only = copies_only; }
  const bool timing0 = c->pipe.timing;     // per-stage event times are a single-batch facility
  int rc = 0;
  auto drain = [&](int k) {
    if (!busy[k]) return 0;
    busy[k] = false;
    b2::Pipeline &p = *slot[k];
    const int r = p.complete(rgb_out ? rgb_out + first[k] : nullptr, status ? status + first[k] : nullptr);
    a.d2h_bytes += p.d2h_bytes; a.kernel_launches += p.launches; a.sync_rounds += p.sync_rounds;
    return r ? fail2(c, p, r) : 0;
  };
  int k = 0;
  for (int at = 0; at < n && !rc; at += chunk, k = (k + 1) % nslots) {
    const int cnt = n - at < chunk ? n - at : chunk;
    rc = drain(k);
    if (rc) break;
    b2::Pipeline &p = *slot[k];
    p.be.err.clear();
    p.timing = false;
    rc = p.prepare(descs + at, cnt, files + at, sizes + at);
    if (!rc) rc = p.run();
    if (rc) { rc = fail2(c, p, rc); break; }
    p.d2h_bytes = 0;
    p.enqueue_rgb(rgb_out ? rgb_out + at : nullptr);
    first[k] = at; busy[k] = true;
    a.entropy_path = p.plan.path;
    a.images += cnt; a.pixels += p.plan.pixels; a.scan_bytes += p.plan.scan_bytes; a.coded_blocks += p.plan.nblocks;
    a.h2d_bytes += p.h2d_bytes;
    for (int i = 0; i < cnt; ++i) a.rgb_bytes += (int64_t)3 * p.plan.imgs[i].width * p.plan.imgs[i].height;
  }
  for (int j = 0; j < nslots; ++j) {                 // oldest first
    const int r = drain((k + j) % nslots);
    if (!rc) rc = r;
  }
  for (b2::Pipeline *p : slot) { p->force_path = user_force; p->copies_only = false; }
  c->agg = a; c->agg_valid = true;
  c->pipe.timing = timing0;
  return rc;
}

int b200jpeg_decode_batch(b200jpeg_ctx *c, const b200jpeg_image_desc *descs, int n, const uint8_t *const *files,
                          const size_t *sizes, uint8_t *const *rgb_out, int16_t *const *coef_out, int32_t *status) {
  return decode_batch_impl(c, descs, n, files, sizes, rgb_out, coef_out, status, false);
}

int b200jpeg_copy_floor(b200jpeg_ctx *c, const b200jpeg_image_desc *descs, int n, const uint8_t *const *files,
                        const size_t *sizes, uint8_t *const *rgb_out, float *ms) {
  const auto t0 = std::chrono::steady_clock::now();
  const int rc = decode_batch_impl(c, descs, n, files, sizes, rgb_out, nullptr, nullptr, true);
  if (ms) *ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
  return rc;
}

// ---- several devices: one context + one host thread per entry of `devices`, contiguous image ranges ----
struct b200jpeg_multi {
  std::vector<b200jpeg_ctx *> ctx;
  std::vector<int> devices;
};

int b200jpeg_multi_create(const int *devices, int ndev, b200jpeg_multi **out) {
  if (!out || !devices || ndev <= 0 || ndev > 64) return B200JPEG_E_ARG;
  *out = nullptr;
  b200jpeg_multi *m = new (std::nothrow) b200jpeg_multi();
  if (!m) return B200JPEG_E_NOMEM;
  for (int k = 0; k < ndev; ++k) {
    b200jpeg_ctx *c = nullptr;
    const int rc = b200jpeg_create(devices[k], &c);
    if (rc) { b200jpeg_multi_destroy(m); return rc; }
    m->ctx.push_back(c); m->devices.push_back(devices[k]);
  }
  *out = m;
  return B200JPEG_OK;
}

void b200jpeg_multi_destroy(b200jpeg_multi *m) {
  if (!m) return;
  for (b200jpeg_ctx *c : m->ctx) b200jpeg_destroy(c);
  delete m;
}

int b200jpeg_multi_count(const b200jpeg_multi *m) { return m ? (int)m->ctx.size() : 0; }
b200jpeg_ctx *b200jpeg_multi_ctx(b200jpeg_multi *m, int k) { return m && k >= 0 && k < (int)m->ctx.size() ? m->ctx[(size_t)k] : nullptr; }

int b200jpeg_multi_decode_batch(b200jpeg_multi *m, const b200jpeg_image_desc *descs, int n, const uint8_t *const *files,
                                const size_t *sizes, uint8_t *const *rgb_out, int32_t *status) {
  if (!m || m->ctx.empty() || n <= 0 || !descs || !files || !sizes) return B200JPEG_E_ARG;
  const int nd = (int)m->ctx.size();
  std::vector<int> rcs((size_t)nd, 0);
  auto part = [&](int k) {                            // images [lo, hi) of device k: sizes differ by at most one
    const int q = n / nd, r = n % nd;
    const int lo = k * q + (k < r ? k : r), hi = lo + q + (k < r ? 1 : 0);
    if (hi > lo)
      rcs[(size_t)k] = b200jpeg_decode_batch(m->ctx[(size_t)k], descs + lo, hi - lo, files + lo, sizes + lo,
                                            rgb_out ? rgb_out + lo : nullptr, nullptr, status ? status + lo : nullptr);
  };
  std::vector<std::thread> pool;
  for (int k = 1; k < nd; ++k) pool.emplace_back(part, k);
  part(0);
  for (std::thread &t : pool) t.join();
  for (int k = 0; k < nd; ++k) if (rcs[(size_t)k]) return rcs[(size_t)k];
  return B200JPEG_OK;
}

int b200jpeg_decode_batch_multi(const int *devices, int ndev, const b200jpeg_image_desc *descs, int n,
                                const uint8_t *const *files, const size_t *sizes, uint8_t *const *rgb_out, int32_t *status) {
  static std::mutex mu;                               // one call at a time through the process-wide contexts
  static b200jpeg_multi *cached = nullptr;
  if (!devices || ndev <= 0) return B200JPEG_E_ARG;
  std::lock_guard<std::mutex> lock(mu);
  if (cached && (cached->devices.size() != (size_t)ndev || !std::equal(cached->devices.begin(), cached->devices.end(), devices))) {
    b200jpeg_multi_destroy(cached);
    cached = nullptr;
  }
  if (!cached) {
    const int rc = b200jpeg_multi_create(devices, ndev, &cached);
    if (rc) return rc;
  }
  return b200jpeg_multi_decode_batch(cached, descs, n, files, sizes, rgb_out, status);
}

// Progressive files (SOF2): see b200jpeg.h and kernels_prog.cuh.
int b200jpeg_decode_progressive(b200jpeg_ctx *c, const b200jpeg_progressive_desc *descs, int n, const uint8_t *const *files,
                                const size_t *sizes, uint8_t *const *rgb_out, int16_t *const *coef_out, int32_t *status) {
  if (!c || n <= 0 || !descs || !files || !sizes) return B200JPEG_E_ARG;
  DeviceGuard g(c->device);
  c->pipe.be.err.clear();
  c->agg_valid = false;
  int rc = c->pipe.prepare_progressive(descs, n, files, sizes);
  if (rc) return fail(c, rc);
  rc = c->pipe.run_progressive();
  if (rc) return fail(c, rc);
  return fail(c, c->pipe.fetch(rgb_out, coef_out, status));
}

int b200jpeg_device_rgb(b200jpeg_ctx *c, int i, void **dev_ptr, size_t *pitch) {
  if (!c || !c->pipe.prepared || i < 0 || i >= c->pipe.plan.n) return B200JPEG_E_ARG;
  if (dev_ptr) *dev_ptr = c->pipe.d_rgb.as<uint8_t>() + c->pipe.plan.imgs[i].rgb_off;
  if (pitch) *pitch = c->pipe.plan.imgs[i].rgb_pitch;
  return B200JPEG_OK;
}

void *b200jpeg_stream(b200jpeg_ctx *c) { return c ? (void *)c->pipe.be.stream : nullptr; }

int b200jpeg_batch_get_stats(const b200jpeg_ctx *c, b200jpeg_batch_stats *s) {
  if (!c || !s) return B200JPEG_E_ARG;
  if (c->agg_valid) { *s = c->agg; return B200JPEG_OK; }
  const b2::BatchPlan &P = c->pipe.plan;
  s->images = P.n; s->pixels = P.pixels; s->scan_bytes = P.scan_bytes; s->coded_blocks = P.nblocks;
  s->rgb_bytes = 0;
  for (int i = 0; i < P.n; ++i) s->rgb_bytes += (int64_t)3 * P.imgs[i].width * P.imgs[i].height;
  s->h2d_bytes = c->pipe.h2d_bytes; s->d2h_bytes = c->pipe.d2h_bytes;
  s->kernel_launches = c->pipe.launches; s->sync_rounds = c->pipe.sync_rounds;
  s->entropy_path = P.path; s->reserved = 0;
  return B200JPEG_OK;
}

int b200jpeg_set_entropy_path(b200jpeg_ctx *c, int mode) {
  if (!c || mode < 0 || mode > 2) return B200JPEG_E_ARG;
  c->pipe.force_path = mode;
  for (b2::Pipeline &x : c->extra) x.force_path = mode;
  return B200JPEG_OK;
}

int b200jpeg_set_timing(b200jpeg_ctx *c, int enabled) {
  if (!c) return B200JPEG_E_ARG;
  c->pipe.timing = enabled != 0;
  return B200JPEG_OK;
}

int b200jpeg_last_timing(b200jpeg_ctx *c, float *ms, int count) {
  if (!c || !ms) return B200JPEG_E_ARG;
  for (int i = 0; i < count && i < B200JPEG_T_COUNT; ++i) ms[i] = c->pipe.last_ms[i];
  return B200JPEG_OK;
}

void *b200jpeg_host_alloc(size_t bytes) {
  void *p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}

void b200jpeg_host_free(void *p) { if (p) cudaFreeHost(p); }

}  // extern "C"

// ------------------------------------------------------------------------------------------------------------------
// b200jpeg_export_coefficients_noalloc -- no counterpart in the reference; added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// A variant of ONE kernel that uses scratch memory: k_export_coef, which rebuilds the reference's MCU[] layout
// (types.hpp:32-50) from the device's scan-order coefficient blocks.  b200jpeg_batch_fetch() stages that array in
// Pipeline::d_export, a buffer it cudaMalloc()s on demand; here the staging area is the file-scope array below, which
// every GPU that loads this library holds until the process exits (B200JPEG_NOALLOC_MAX_BYTES = 16 MiB: block
// positions x 384 bytes, i.e. images up to 43,690 block positions -- 1920x1080 4:2:0 needs 12.5 MB).
//
// In host code `g_noalloc_scratch` is a host-side stand-in, NOT a device address: the address a kernel may use comes
// from cudaGetSymbolAddress(), asked once per device with that device current (noalloc_scratch_of).
// The call allocates nothing: no cudaMalloc / cudaFree / cudaHostAlloc, no allocating library routine.  Every call on a
// device shares the one array and the array keeps the last call's contents, so each call clears the bytes it is about
// to use first (MCU[] is value-initialised, jpeg.cpp:407: positions the scan never writes must read zero).  Calls that
// overlap on the same GPU -- two host threads, two contexts on one device -- RACE on it: serialise them.
namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[B200JPEG_NOALLOC_MAX_BYTES];

constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noalloc_mu;
void *g_noalloc_addr[kNoAllocMaxDevices];          // device address of the array on each device, once asked
bool g_noalloc_known[kNoAllocMaxDevices];

// set-up, once per device, with that device current
void *noalloc_scratch_of(int device, std::string &err) {
  if (device < 0 || device >= kNoAllocMaxDevices) { err = "device index out of range"; return nullptr; }
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_known[device]) {
    void *ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr) {
      err = std::string("cudaGetSymbolAddress(g_noalloc_scratch): ") + cudaGetErrorString(e);
      cudaGetLastError();
      return nullptr;
    }
    g_noalloc_addr[device] = ptr;
    g_noalloc_known[device] = true;
  }
  return g_noalloc_addr[device];
}

}  // namespace

extern "C" int b200jpeg_export_coefficients_noalloc(b200jpeg_ctx *c, int i, int16_t *coef_out, size_t coef_out_bytes) {
  if (!c || !coef_out) return B200JPEG_E_ARG;
  b2::Pipeline &p = c->pipe;
  if (!p.prepared || !p.ran) return B200JPEG_E_STATE;                 // needs a batch that b200jpeg_batch_run() has launched
  if (i < 0 || i >= p.plan.n) return B200JPEG_E_ARG;
  const size_t bytes = (size_t)p.plan.bwr[(size_t)i] * p.plan.bhr[(size_t)i] * 3 * 64 * sizeof(int16_t);
  if (bytes > (size_t)B200JPEG_NOALLOC_MAX_BYTES) {
    c->last_error = "b200jpeg_export_coefficients_noalloc: the image's MCU array (" + std::to_string(bytes) +
                    " bytes) exceeds the static scratch of " + std::to_string((size_t)B200JPEG_NOALLOC_MAX_BYTES) +
                    " bytes (B200JPEG_NOALLOC_MAX_BYTES); use b200jpeg_batch_fetch()";
    return B200JPEG_E_TOO_LARGE;
  }
  if (coef_out_bytes < bytes) { c->last_error = "b200jpeg_export_coefficients_noalloc: output buffer too small"; return B200JPEG_E_ARG; }
  DeviceGuard g(c->device);
  void *scratch = noalloc_scratch_of(c->device, c->last_error);
  if (!scratch) return B200JPEG_E_CUDA;
  const int rc = p.finish();                                          // waits for the batch; reads status into pinned memory prepare() holds
  if (rc) return fail(c, rc);
  const b2::ImageDev &im = p.plan.imgs[(size_t)i];
  p.be.fill(scratch, 0, bytes);                                       // the array keeps the last call's contents
  if (!(p.h_out[i].status & b2::kStDead)) {                           // blocks not all written this batch: zeros, as batch_fetch
    const int egrid = im.nblocks / 4 + 1 < 2368 ? im.nblocks / 4 + 1 : 2368;
    B2_LAUNCH(b2::k_export_coef, egrid, 256, 0, p.be.stream, im, p.plan.bwr[(size_t)i], p.d_coef.as<int16_t>(),
              static_cast<int16_t *>(scratch));
  }
  p.be.d2h(coef_out, scratch, bytes);
  p.be.sync();
  p.be.check("b200jpeg_export_coefficients_noalloc");
  return fail(c, p.be.ok() ? 0 : B200JPEG_E_CUDA);
}

#if B2_SEG_DIAG == 7
// measurements only (build_variant.sh -DB2_SEG_DIAG=7): the decoder warps' cycle counters, read and cleared
extern "C" int b200jpeg_debug_seg_diag(unsigned long long *out24) {
  unsigned long long z[24] = {};
  if (cudaMemcpyFromSymbol(out24, b2::g_seg_diag, sizeof z) != cudaSuccess) return -1;
  return cudaMemcpyToSymbol(b2::g_seg_diag, z, sizeof z) == cudaSuccess ? 0 : -1;
}
#endif
