// tests/seam/ref_seam_post.h -- TEST INFRASTRUCTURE: the drop-in proof.
//
// One translation unit = oracle/ref_wrapper_pre.h + the reference's jpeg.cpp (streamed by tests/seam/build.sh exactly as
// oracle/Makefile does: its X11 display include swapped for types.hpp + a headless displayImage) + this file.  In ONE
// process it
//   1. lets the REFERENCE's own marker reader and read_* functions parse the file up to SOS
//      (Marker::Marker file.hpp:192-258, read_sof / read_huffman_table / read_quantization_table / read_sos / read_dri,
//      jpeg.cpp:50-298) -- the state process_image_data() finds at the seam (jpeg.cpp:21-34, 858-863);
//   2. runs fill_desc() (tests/seam/fill_desc.inc = the stub INTEGRATION.md section 2 gives a maintainer) on the
//      reference's Image members and calls b200jpeg_decode_batch() through the C ABI of libb200jpeg.so;
//   3. runs the reference's own decodeHuffman() / dequantize() / inverseDCT() / YCbCrToRGB() (jpeg.cpp:785-788) on the
//      same object and compares: coefficients after decodeHuffman() must be bit-exact, RGB within +-1 LSB
//      (max |err| and the number of differing samples are printed).
// Usage: ref_seam <file.jpg> [...]   -> one JSON line per file; exit code 0 iff every file passed.
#undef main
#undef class

#include "b200jpeg.h"
#include "fill_desc.inc"
#ifdef SEAM_EMU   // CPU-tier variant: the same seam against the SIMT emulation of the kernels (tests/emul), to check fill_desc() without a GPU
extern "C" int emu_decode_batch(const b200jpeg_image_desc *descs, int n, const uint8_t *const *files, const size_t *sizes,
                                uint8_t *const *rgb_out, int16_t *const *coef_out, int32_t *status, int32_t *info);
#endif

namespace {

bool seam_parse_until_sos(Image *im) {            // the dispatch of Image::readJPEG (jpeg.cpp:834-896) up to process_image_data()
  while (true) {
    if (im->file->eof()) return false;
    unsigned char *b = im->file->read(1);
    if (b[0] != 0xff) return false;
    Marker *marker = new Marker(im->file);
    const int t = marker->type;
    if (t == MarkerType::PAD) continue;
    if (t == MarkerType::SOF) im->read_sof(marker);
    else if (t == MarkerType::DHT) im->read_huffman_table(marker);
    else if (t == MarkerType::DQT) im->read_quantization_table(marker);
    else if (t == MarkerType::DRI) im->read_dri(marker);
    else if (t == MarkerType::APP) im->read_application_headers(marker);
    else if (t == MarkerType::SOS) { im->read_sos(marker); return true; }
    else if (t == MarkerType::SOI || t == MarkerType::DRM || t == MarkerType::META) { /* ignored */ }
    else return false;
    delete marker;
  }
}

// the two things the reference does not keep: the file's bytes, and where the scan starts in them
bool seam_load(const char *path, std::vector<uint8_t> &bytes, size_t &scan_offset) {
  std::ifstream f(path, std::ios::binary);
  if (!f) return false;
  bytes.assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
  size_t pos = 2;
  while (pos + 4 <= bytes.size()) {
    if (bytes[pos] != 0xff) return false;
    const int m = bytes[pos + 1];
    if (m == 0xff) { ++pos; continue; }
    const size_t len = ((size_t)bytes[pos + 2] << 8) | bytes[pos + 3];
    pos += 2 + len;
    if (m == 0xda) { scan_offset = pos; return true; }
  }
  return false;
}

int seam_one(b200jpeg_ctx *ctx, const char *path) {
  std::vector<uint8_t> bytes;
  size_t scan_offset = 0;
  if (!seam_load(path, bytes, scan_offset)) { printf("{\"file\":\"%s\",\"error\":\"cannot load\"}\n", path); return 1; }
  Image *im = new Image(path);                                   // jpeg.cpp:797-807
  if (!seam_parse_until_sos(im)) { printf("{\"file\":\"%s\",\"error\":\"reference parser did not reach SOS\"}\n", path); return 1; }

  b200jpeg_image_desc d;
  fill_desc(im, d);
  d.scan_offset = (int64_t)scan_offset;
  d.scan_length = (int64_t)(bytes.size() - scan_offset);
  const size_t npos = (size_t)im->mcuWidthReal * im->mcuHeightReal;
  std::vector<uint8_t> rgb((size_t)3 * im->image_width * im->image_height);
  std::vector<int16_t> coef(npos * 3 * 64);
  const uint8_t *files[1] = {bytes.data()};
  const size_t sizes[1] = {bytes.size()};
  uint8_t *out[1] = {rgb.data()};
  int16_t *co[1] = {coef.data()};
  int32_t status = -1;
#ifdef SEAM_EMU
  int32_t info[3] = {0, 0, 0};
  const int rc = emu_decode_batch(&d, 1, files, sizes, out, co, &status, info);
#else
  const int rc = b200jpeg_decode_batch(ctx, &d, 1, files, sizes, out, co, &status);
#endif
  if (rc || status) { printf("{\"file\":\"%s\",\"error\":\"b200jpeg_decode_batch rc=%d status=%d: %s\"}\n", path, rc, status, ctx ? b200jpeg_last_error(ctx) : ""); return 1; }

  // the reference's own stages on the same object (jpeg.cpp:785-788)
  im->mcus = im->decodeHuffman();
  long long coef_bad = 0;
  for (size_t p = 0; p < npos; ++p)
    for (int c = 0; c < 3; ++c)
      for (int k = 0; k < 64; ++k) coef_bad += im->mcus[p][c][k] != (int)coef[(p * 3 + c) * 64 + k];
  im->dequantize();
  im->inverseDCT();
  im->YCbCrToRGB();
  int maxerr = 0;
  long long ndiff = 0;
  const int W = im->image_width, H = im->image_height;
  for (int y = 0; y < H; ++y)
    for (int x = 0; x < W; ++x) {                                  // the gather of display.hpp:19-34
      const MCU &m = im->mcus[(y / 8) * im->mcuWidthReal + (x / 8)];
      const int q = (y % 8) * 8 + (x % 8);
      const int ref[3] = {m.r[q], m.g[q], m.b[q]};
      for (int c = 0; c < 3; ++c) {
        const int e = abs(ref[c] - (int)rgb[((size_t)y * W + x) * 3 + c]);
        if (e > maxerr) maxerr = e;
        ndiff += e != 0;
      }
    }
  const bool ok = coef_bad == 0 && maxerr <= 1;
  printf("{\"file\":\"%s\",\"w\":%d,\"h\":%d,\"hs\":%d,\"vs\":%d,\"ri\":%d,\"coef_mismatch\":%lld,\"rgb_maxerr\":%d,\"rgb_differing\":%lld,\"ok\":%s}\n",
         path, W, H, im->horizontalSamplingFactor, im->verticalSamplingFactor, im->restartInterval, coef_bad, maxerr, ndiff,
         ok ? "true" : "false");
  return ok ? 0 : 1;
}

}  // namespace

int main(int argc, char **argv) {
  std::cout.setstate(std::ios_base::failbit);      // the reference logs every table to cout
  if (argc < 2) { fprintf(stderr, "usage: ref_seam <file.jpg> [...]\n"); return 64; }
  b200jpeg_ctx *ctx = nullptr;
#ifndef SEAM_EMU
  const int rc = b200jpeg_create(0, &ctx);
  if (rc) { fprintf(stderr, "ref_seam: b200jpeg_create failed (%d): no usable sm_100 GPU, and there is no CPU fallback\n", rc); return 2; }
#endif
  int bad = 0;
  for (int i = 1; i < argc; ++i) bad += seam_one(ctx, argv[i]);
#ifndef SEAM_EMU
  b200jpeg_destroy(ctx);
#endif
  return bad ? 1 : 0;
}
