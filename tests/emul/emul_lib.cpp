// tests/emul/emul_lib.cpp -- TEST INFRASTRUCTURE ONLY (CPU tier).
// Compiles the product's kernel sources and launch sequence (jpeg_decoder_b200/csrc/pipeline*.h)
// against the SIMT emulation in cuda_emul.h, so that -m "not gpu" tests can check kernel LOGIC
// against the oracle without a GPU.  Built into tests/emul/_build/libb200jpeg_emul.so by
// tests/conftest.py; never part of libb200jpeg.so.
#define B200JPEG_EMU 1
#include "../../jpeg_decoder_b200/csrc/pipeline_run.h"

extern "C" int emu_decode_batch(const b200jpeg_image_desc *descs, int n, const uint8_t *const *files,
                                const size_t *sizes, uint8_t *const *rgb_out, int16_t *const *coef_out,
                                int32_t *status, int32_t *info /* out [0]=launches [1]=sync_rounds, in [2]=entropy path */) {
  b2::Pipeline P;
  if (info) P.force_path = info[2];      // in: 0 auto, 1 segment path, 2 self-synchronising path
  int rc = P.prepare(descs, n, files, sizes);
  if (rc) return rc;
  rc = P.run();
  if (rc) return rc;
  rc = P.fetch(rgb_out, coef_out, status);
  if (info) { info[0] = P.launches; info[1] = P.sync_rounds; }
  P.destroy();
  return rc;
}

extern "C" int emu_decode_progressive(const b200jpeg_progressive_desc *descs, int n, const uint8_t *const *files, const size_t *sizes,
                                      uint8_t *const *rgb_out, int16_t *const *coef_out, int32_t *status) {
  b2::Pipeline P;
  int rc = P.prepare_progressive(descs, n, files, sizes);
  if (rc) return rc;
  rc = P.run_progressive();
  if (rc) return rc;
  rc = P.fetch(rgb_out, coef_out, status);
  P.destroy();
  return rc;
}
