// oracle/ref_wrapper_post.h -- TEST INFRASTRUCTURE: second half of the wrapper TU (see ref_wrapper_pre.h).
#undef main
#undef class

namespace {

typedef std::chrono::steady_clock Clock;
double ms_since(Clock::time_point t0) {
  return std::chrono::duration<double, std::milli>(Clock::now() - t0).count();
}

// The marker dispatch of Image::readJPEG (jpeg.cpp:834-896) up to, but not including,
// process_image_data(), so that the stages can be run and dumped separately.
bool parse_until_sos(Image *im) {
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
    else return false;  // EOI / INVALID / anything else ends parsing (jpeg.cpp:878-896)
    delete marker;
  }
}

void gather_rgb(const Image *im, std::vector<uint8_t> &rgb) {
  const int W = im->image_width, H = im->image_height;
  rgb.resize((size_t)W * H * 3);
  for (int y = 0; y < H; ++y)
    for (int x = 0; x < W; ++x) {  // display.hpp:19-34
      const MCU &m = im->mcus[(y / 8) * im->mcuWidthReal + (x / 8)];
      const int p = (y % 8) * 8 + (x % 8);
      uint8_t *o = &rgb[((size_t)y * W + x) * 3];
      o[0] = (uint8_t)m.r[p]; o[1] = (uint8_t)m.g[p]; o[2] = (uint8_t)m.b[p];
    }
}

bool write_file(const char *path, const void *p, size_t n) {
  FILE *f = fopen(path, "wb");
  if (!f) return false;
  const bool ok = fwrite(p, 1, n, f) == n;
  fclose(f);
  return ok;
}

int mode_dump(const char *in, const char *coef_path, const char *rgb_path) {
  auto t0 = Clock::now();
  Image *im = new Image(in);
  if (!parse_until_sos(im)) { fprintf(stderr, "ref_decode: no SOS reached\n"); return 2; }
  const double t_parse = ms_since(t0);
  t0 = Clock::now(); im->mcus = im->decodeHuffman(); const double t_huff = ms_since(t0);
  const size_t npos = (size_t)im->mcuWidthReal * im->mcuHeightReal;
  if (!write_file(coef_path, im->mcus, npos * sizeof(MCU))) return 3;
  t0 = Clock::now(); im->dequantize();  const double t_deq = ms_since(t0);
  t0 = Clock::now(); im->inverseDCT();  const double t_idct = ms_since(t0);
  t0 = Clock::now(); im->YCbCrToRGB();  const double t_col = ms_since(t0);
  std::vector<uint8_t> rgb; gather_rgb(im, rgb);
  if (!write_file(rgb_path, rgb.data(), rgb.size())) return 3;
  fprintf(stdout,
          "{\"w\":%d,\"h\":%d,\"bw\":%d,\"bh\":%d,\"bwr\":%d,\"bhr\":%d,\"hs\":%d,\"vs\":%d,\"ri\":%d,"
          "\"scan_unstuffed\":%u,\"sizeof_mcu\":%zu,"
          "\"ms\":{\"parse_unstuff\":%.3f,\"huffman\":%.3f,\"dequant\":%.3f,\"idct\":%.3f,\"colour\":%.3f}}\n",
          im->image_width, im->image_height, im->mcuWidth, im->mcuHeight, im->mcuWidthReal, im->mcuHeightReal,
          im->horizontalSamplingFactor, im->verticalSamplingFactor, im->restartInterval,
          im->file->size, sizeof(MCU), t_parse, t_huff, t_deq, t_idct, t_col);
  return 0;
}

int mode_full(const char *in, const char *rgb_path) {
  Image *im = new Image(in);
  im->readJPEG();
  if (im->mcus == nullptr) { fprintf(stderr, "ref_decode: nothing decoded\n"); return 2; }
  std::vector<uint8_t> rgb; gather_rgb(im, rgb);
  if (!write_file(rgb_path, rgb.data(), rgb.size())) return 3;
  fprintf(stdout, "{\"w\":%d,\"h\":%d}\n", im->image_width, im->image_height);
  return 0;
}

int mode_bench(int reps, int nfiles, char **files) {
  double seconds = 0.0; long long pixels = 0; long long images = 0;
  for (int r = 0; r < reps; ++r)
    for (int i = 0; i < nfiles; ++i) {
      auto t0 = Clock::now();
      Image *im = new Image(files[i]);
      im->readJPEG();                       // the reference's whole path, jpeg.cpp:826-907 -> 755-789
      seconds += ms_since(t0) * 1e-3;
      if (im->mcus == nullptr) { fprintf(stderr, "ref_decode: %s not decoded\n", files[i]); return 2; }
      pixels += (long long)im->image_width * im->image_height; ++images;
      delete[] im->mcus;                    // the reference never frees it (jpeg.cpp:909-913)
      delete im;
    }
  fprintf(stdout, "{\"images\":%lld,\"pixels\":%lld,\"seconds\":%.6f}\n", images, pixels, seconds);
  return 0;
}

}  // namespace

int main(int argc, char **argv) {
  std::cout.setstate(std::ios_base::failbit);  // the reference logs every table to cout
  if (argc >= 5 && !strcmp(argv[1], "dump")) return mode_dump(argv[2], argv[3], argv[4]);
  if (argc >= 4 && !strcmp(argv[1], "full")) return mode_full(argv[2], argv[3]);
  if (argc >= 4 && !strcmp(argv[1], "bench")) return mode_bench(atoi(argv[2]), argc - 3, argv + 3);
  fprintf(stderr, "usage: ref_decode dump <in.jpg> <out.coef> <out.rgb> | full <in.jpg> <out.rgb> | bench <reps> <file>...\n");
  return 64;
}
