// b200jpeg.cu -- the C ABI of include/b200jpeg.h over the CUDA pipeline (sm_100a only).
// No CPU path: without a usable GPU every decode entry point returns B200JPEG_E_CUDA.
#include "pipeline_run.h"

#include <new>

struct b200jpeg_ctx {
  int device = 0;
  b2::Pipeline pipe;
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
  return true;
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
  for (int i = 0; i <= B200JPEG_T_COUNT; ++i) cudaEventCreate(&c->pipe.ev[i]);
  const size_t es = b2::entropy_smem_bytes(b2::kMaxSlots, true);
  bool ok = allow_smem(b2::k_entropy_sync<true>, es, c->last_error) && allow_smem(b2::k_entropy_sync<false>, es, c->last_error) &&
            allow_smem(b2::k_entropy_write, es, c->last_error) &&
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
  if (c->pipe.be.stream) cudaStreamSynchronize(c->pipe.be.stream);
  c->pipe.destroy();
  for (int i = 0; i <= B200JPEG_T_COUNT; ++i) if (c->pipe.ev[i]) cudaEventDestroy(c->pipe.ev[i]);
  if (c->pipe.be.stream) cudaStreamDestroy(c->pipe.be.stream);
  delete c;
}

const char *b200jpeg_last_error(const b200jpeg_ctx *c) { return c ? c->last_error.c_str() : "null context"; }

int b200jpeg_batch_prepare(b200jpeg_ctx *c, const b200jpeg_image_desc *descs, int n, const uint8_t *const *files,
                           const size_t *sizes) {
  if (!c) return B200JPEG_E_ARG;
  DeviceGuard g(c->device);
  c->pipe.be.err.clear();
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

int b200jpeg_decode_batch(b200jpeg_ctx *c, const b200jpeg_image_desc *descs, int n, const uint8_t *const *files,
                          const size_t *sizes, uint8_t *const *rgb_out, int16_t *const *coef_out, int32_t *status) {
  int rc = b200jpeg_batch_prepare(c, descs, n, files, sizes);
  if (rc) return rc;
  rc = b200jpeg_batch_run(c);
  if (rc) return rc;
  return b200jpeg_batch_fetch(c, rgb_out, coef_out, status);
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
  const b2::BatchPlan &P = c->pipe.plan;
  s->images = P.n; s->pixels = P.pixels; s->scan_bytes = P.scan_bytes; s->coded_blocks = P.nblocks;
  s->rgb_bytes = 0;
  for (int i = 0; i < P.n; ++i) s->rgb_bytes += (int64_t)3 * P.imgs[i].width * P.imgs[i].height;
  s->h2d_bytes = c->pipe.h2d_bytes; s->d2h_bytes = c->pipe.d2h_bytes;
  s->kernel_launches = c->pipe.launches; s->sync_rounds = c->pipe.sync_rounds;
  return B200JPEG_OK;
}

int b200jpeg_set_entropy_path(b200jpeg_ctx *c, int mode) {
  if (!c || mode < 0 || mode > 2) return B200JPEG_E_ARG;
  c->pipe.force_path = mode;
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
