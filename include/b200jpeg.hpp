// b200jpeg.hpp -- C++ host surface over the C ABI (b200jpeg.h), keeping the reference's names.
//
// aswanthabam/JPEG_Decoder is one translation unit: class Image (jpeg.cpp:19-914) reads the markers
// into HuffmanTable (include/huffman.hpp:8-49), QuantizationTable / ColorComponent / MCU
// (include/types.hpp:32-110) and then runs its private process_image_data() (jpeg.cpp:755-789) on the
// CPU.  A maintainer who wants the B200 path keeps exactly that surface:
//
//   b200jpeg::Image img("picture.jpg");   // Image::Image, jpeg.cpp:797-807 (SOI check)
//   img.readJPEG();                       // jpeg.cpp:826-907: markers -> tables -> process_image_data()
//   img.image_width, img.image_height, img.mcuWidthReal, img.mcuHeightReal      // jpeg.cpp:792-795
//   img.rgb();                            // headless replacement of display() (display.hpp:15-120)
//   img.mcus();                           // the MCU array after YCbCrToRGB, gathered as display.hpp:19-34 reads it
//   img.saveToBMP("out.bmp");             // jpeg.cpp:809-816, with the channel order put right
//
// Differences, all deliberate: the file is read once into memory (the reference allocates per byte,
// file.hpp:26-52); errors throw b200jpeg::Error carrying the B200JPEG_E_* code where the reference
// prints and calls exit(1); process_image_data() runs on the GPU through b200jpeg_decode_batch() and
// FAILS (throws) when no B200-class device is present -- there is no CPU fallback.
// Batches: b200jpeg::Image::decode_batch() decodes many files in one call (images are independent, the
// library cuts them into chunks that rotate over four internal pipelines).
#ifndef B200JPEG_HPP
#define B200JPEG_HPP
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

#include "b200jpeg.h"

namespace b200jpeg {

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string &what) : std::runtime_error(what + " (b200jpeg error " + std::to_string(c) + ")"), code(c) {}
};

// include/types.hpp:23-31
static const int zigZagMap[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                  41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                  30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// include/huffman.hpp:8-49.  Same members; symbols hold up to 176 entries because read_huffman_table
// admits that many (jpeg.cpp:177).
class HuffmanTable {
 public:
  unsigned char offset[17] = {0};
  unsigned char symbols[176] = {0};
  unsigned int codes[176] = {0};
  bool set = false;
  void generateCodes() {                       // huffman.hpp:17-29
    unsigned int code = 0;
    for (int i = 0; i < 16; i++) {
      for (int j = offset[i]; j < offset[i + 1]; j++) codes[j] = code++;
      code <<= 1;
    }
  }
};

// include/types.hpp:79-110: de-zig-zagged on construction, operator[] in natural order.
class QuantizationTable {
  int table[64];
  int header;

 public:
  bool set = false;
  QuantizationTable() : header(0) { std::memset(table, 0, sizeof table); }
  QuantizationTable(int header_, const unsigned char *raw) : header(header_) {
    for (int i = 0; i < 64; i++) table[zigZagMap[i]] = (int)raw[i];
    set = true;
  }
  int operator[](int index) const { return table[index]; }
  int id() const { return header & 15; }
};

// include/types.hpp:69-77
struct ColorComponent {
  int horizontal_sampling_factor = 0;
  int vertical_sampling_factor = 0;
  int quantizationTableID = 0;
  int huffmanDCTTableID = 0;
  int huffmanACTableID = 0;
  bool set = false;
};

// include/types.hpp:32-67: one 8x8 block position; y/cb/cr before YCbCrToRGB, r/g/b after.
struct MCU {
  union { int y[64] = {0}; int r[64]; };
  union { int cb[64] = {0}; int g[64]; };
  union { int cr[64] = {0}; int b[64]; };
  int *operator[](int index) { return index == 0 ? y : index == 1 ? cb : index == 2 ? cr : nullptr; }
};

// include/file.hpp:167-181.  Same enumerators, same values.
enum MarkerType { SOI = 0xd8, APP = 0xe0, DQT = 0xdb, SOF = 0xc0, DHT = 0xc4, SOS = 0xda, EOI = 0xd9, DRI = 0xdd, DRM = 0xd0,
                  META = -2, PAD = 0x00, INVALID = -1 };

// include/file.hpp:183-280 over a file that is already in memory: the byte after an 0xFF is classified exactly as
// Marker::Marker does (PAD, META for APPn, DRM for RSTn with length = n, SOS, the table/frame markers, INVALID for the
// rest -- COM included, file.hpp:252-253), `length` counts its own two bytes, `data` is the payload.  Nothing is copied
// and nothing leaks (the reference allocates per read, file.hpp:267); at SOS the scan is NOT consumed (the reference runs
// readImageData() there, file.hpp:214-219): `scan` is where the entropy-coded bytes start, for the GPU to unstuff.
class Marker {
  int _data_pointer = 0;

 public:
  unsigned char marker = 0;
  int length = 0;
  const unsigned char *data = nullptr;
  int type = MarkerType::INVALID;
  size_t scan = 0;     // SOS only: offset of the first entropy-coded byte in the file
  size_t next = 0;     // offset of the byte after this marker segment

  // `pos`: offset of the marker byte (the one after 0xFF) in file[0, n)
  Marker(const unsigned char *file, size_t n, size_t pos) {
    if (pos >= n) throw Error(B200JPEG_E_TRUNCATED, "marker reader: end of file");
    marker = file[pos];
    next = pos + 1;
    if (marker == 0x00) { type = MarkerType::PAD; return; }
    if (marker >= 0xd0 && marker < 0xd8) { type = MarkerType::DRM; length = marker - 0xd0; return; }
    // SOI / EOI stand alone (T.81 B.1.1.3); the reference would take the two bytes behind them for a length
    // (file.hpp:255-256) but never meets SOI (the constructor consumes it, jpeg.cpp:797-807) and stops at EOI
    if (marker == 0xd8) { type = MarkerType::SOI; return; }
    if (marker == 0xd9) { type = MarkerType::EOI; return; }
    if (pos + 3 > n) throw Error(B200JPEG_E_TRUNCATED, "marker reader: segment length past the end of the file");
    length = (file[pos + 1] << 8) + file[pos + 2];
    if (length < 2 || pos + 1 + (size_t)length > n) throw Error(B200JPEG_E_TRUNCATED, "marker reader: segment past the end of the file");
    data = file + pos + 3;
    next = pos + 1 + (size_t)length;
    if (marker >= 224 && marker < 240) type = MarkerType::META;
    else if (marker == 0xda) { type = MarkerType::SOS; scan = next; }
    else switch (marker) {
      case 0xdb: type = MarkerType::DQT; break;
      case 0xc0: case 0xc2: type = MarkerType::SOF; break;
      case 0xc4: type = MarkerType::DHT; break;
      case 0xd9: type = MarkerType::EOI; break;
      case 0xd8: type = MarkerType::SOI; break;
      case 0xdd: type = MarkerType::DRI; break;
      default: type = MarkerType::INVALID;
    }
  }
  // file.hpp:260-274: the next n payload bytes (nullptr past the end, where the reference logs an error)
  const unsigned char *read(int n = 1) {
    if (_data_pointer + n > length - 2) return nullptr;
    const unsigned char *res = data + _data_pointer;
    _data_pointer += n;
    return res;
  }
};

// One GPU context per device, shared by the Images of a process (b200jpeg_create / _destroy).
class Context {
  b200jpeg_ctx *ctx_ = nullptr;

 public:
  explicit Context(int device = 0) {
    const int rc = b200jpeg_create(device, &ctx_);
    if (rc) throw Error(rc, "b200jpeg_create: no usable sm_100 GPU (the decode path has no CPU fallback)");
  }
  ~Context() { b200jpeg_destroy(ctx_); }
  Context(const Context &) = delete;
  Context &operator=(const Context &) = delete;
  b200jpeg_ctx *get() const { return ctx_; }
  static Context &shared() { static Context c(0); return c; }   // device 0; other devices: Context(k), or decode_files(.., devices)
};

// Pinned host memory (b200jpeg_host_alloc): copies to and from it are asynchronous, which is what lets the library
// overlap the files going in, the kernels and the pixels coming out.
class PinnedBuffer {
  uint8_t *p_ = nullptr;
  size_t n_ = 0;

 public:
  PinnedBuffer() = default;
  explicit PinnedBuffer(size_t n) : p_(static_cast<uint8_t *>(b200jpeg_host_alloc(n))), n_(n) {
    if (!p_) throw Error(B200JPEG_E_NOMEM, "b200jpeg_host_alloc");
  }
  ~PinnedBuffer() { b200jpeg_host_free(p_); }
  PinnedBuffer(const PinnedBuffer &) = delete;
  PinnedBuffer &operator=(const PinnedBuffer &) = delete;
  PinnedBuffer(PinnedBuffer &&o) noexcept : p_(o.p_), n_(o.n_) { o.p_ = nullptr; o.n_ = 0; }
  PinnedBuffer &operator=(PinnedBuffer &&o) noexcept { if (this != &o) { b200jpeg_host_free(p_); p_ = o.p_; n_ = o.n_; o.p_ = nullptr; o.n_ = 0; } return *this; }
  uint8_t *data() const { return p_; }
  size_t size() const { return n_; }
};

// What the reference's main() does for one path (jpeg.cpp:916-929: Image ctor, readJPEG) for MANY files in one go, the
// way the device wants to be fed: every file read straight into ONE pinned arena by a pool of
// host threads (the reference reads byte by byte through an ifstream, file.hpp:26-52), every header parsed on host
// threads (b200jpeg_parse_headers), every scan decoded by b200jpeg_decode_batch -- over several GPUs when `devices`
// names more than one (contiguous image ranges, one context + host thread each) -- into ONE pinned RGB arena.
class Batch {
 public:
  std::vector<std::string> paths;
  std::vector<b200jpeg_image_desc> descs;
  std::vector<int32_t> parse_rc, status;       // per file: header parse code (0 = ok), decode status bits (0 = decoded)
  std::vector<size_t> file_off, file_len, rgb_off;
  PinnedBuffer files, pixels;
  double ms_read = 0, ms_parse = 0, ms_alloc = 0, ms_decode = 0;   // where decode_files spent its time: file read into the pinned arena
                                                                   // (incl. its allocation), header parse, output allocation, GPU decode host -> host
  int width(size_t i) const { return descs[i].image_width; }
  int height(size_t i) const { return descs[i].image_height; }
  bool ok(size_t i) const { return parse_rc[i] == 0 && status[i] == 0; }
  const uint8_t *rgb(size_t i) const { return pixels.data() + rgb_off[i]; }   // interleaved R,G,B, pitch 3*width
  void savePPM(size_t i, const std::string &filename) const {
    std::ofstream o(filename, std::ios::binary);
    o << "P6\n" << width(i) << " " << height(i) << "\n255\n";
    o.write(reinterpret_cast<const char *>(rgb(i)), (std::streamsize)3 * width(i) * height(i));
  }
};

// decode_files_into: the same into a Batch that is REUSED from call to call -- its pinned arenas are kept and only grow
// (pinning 1.6 GB takes ~0.6 s; a server that decodes batch after batch must not pay that every time).
inline void decode_files_into(Batch &b, const std::vector<std::string> &paths, const std::vector<int> &devices = std::vector<int>(1, 0),
                              int parse_flags = B200JPEG_PARSE_REFERENCE, int host_threads = 0);
inline Batch decode_files(const std::vector<std::string> &paths, const std::vector<int> &devices = std::vector<int>(1, 0),
                          int parse_flags = B200JPEG_PARSE_REFERENCE, int host_threads = 0) {
  Batch b;
  decode_files_into(b, paths, devices, parse_flags, host_threads);
  return b;
}
inline void decode_files_into(Batch &b, const std::vector<std::string> &paths, const std::vector<int> &devices, int parse_flags,
                              int host_threads) {
  b.paths = paths;
  const size_t n = paths.size();
  b.descs.resize(n); b.parse_rc.assign(n, 0); b.status.assign(n, 0);
  b.file_off.assign(n, 0); b.file_len.assign(n, 0); b.rgb_off.assign(n, 0);
  b.ms_read = b.ms_parse = b.ms_alloc = b.ms_decode = 0;
  if (n == 0) return;
  typedef std::chrono::steady_clock Clock;
  auto since = [](Clock::time_point t0) { return std::chrono::duration<double, std::milli>(Clock::now() - t0).count(); };
  Clock::time_point t0 = Clock::now();
  if (host_threads <= 0) host_threads = (int)std::thread::hardware_concurrency();
  if (host_threads < 1) host_threads = 1;
  if ((size_t)host_threads > n) host_threads = (int)n;
  auto parallel = [&](auto &&fn) {                     // fn(i) for every file, files handed out through one counter
    std::atomic<size_t> next(0);
    auto work = [&]() { for (size_t i; (i = next.fetch_add(1)) < n;) fn(i); };
    std::vector<std::thread> pool;
    for (int t = 1; t < host_threads; ++t) pool.emplace_back(work);
    work();
    for (std::thread &t : pool) t.join();
  };
  parallel([&](size_t i) {                             // sizes first: the arena is allocated once
    std::ifstream f(paths[i], std::ios::binary | std::ios::ate);
    b.file_len[i] = f ? (size_t)f.tellg() : 0;
  });
  size_t total = 0;
  for (size_t i = 0; i < n; ++i) { b.file_off[i] = total; total += (b.file_len[i] + 255) / 256 * 256; }
  if (b.files.size() < total + 256) b.files = PinnedBuffer(total + 256 + total / 8);
  parallel([&](size_t i) {
    std::ifstream f(paths[i], std::ios::binary);
    if (f && b.file_len[i]) f.read(reinterpret_cast<char *>(b.files.data() + b.file_off[i]), (std::streamsize)b.file_len[i]);
    if (!f) b.file_len[i] = 0;
  });
  std::vector<const uint8_t *> fptr(n);
  for (size_t i = 0; i < n; ++i) fptr[i] = b.files.data() + b.file_off[i];
  b.ms_read = since(t0); t0 = Clock::now();
  b200jpeg_parse_headers(fptr.data(), b.file_len.data(), (int)n, parse_flags, host_threads, b.descs.data(), b.parse_rc.data());
  b.ms_parse = since(t0); t0 = Clock::now();
  // the images whose header parsed go to the device; the others keep their parse code
  std::vector<size_t> live;
  size_t rgb_total = 0;
  for (size_t i = 0; i < n; ++i) if (b.parse_rc[i] == 0) {
    live.push_back(i);
    b.rgb_off[i] = rgb_total;
    rgb_total += ((size_t)3 * b.descs[i].image_width * b.descs[i].image_height + 255) / 256 * 256;
  }
  // progressive files (refused by b200jpeg_parse_headers like the reference refuses them): under B200JPEG_PARSE_SPEC they
  // are parsed as such and decoded by b200jpeg_decode_progressive on the first device, after the baseline ones
  std::vector<size_t> prog;
  std::vector<b200jpeg_progressive_desc> pd;
  if (parse_flags & B200JPEG_PARSE_SPEC)
    for (size_t i = 0; i < n; ++i) if (b.parse_rc[i] == B200JPEG_E_PROGRESSIVE) {
      pd.emplace_back();
      const int rc = b200jpeg_parse_progressive(fptr[i], b.file_len[i], &pd.back());
      if (rc) { b.parse_rc[i] = rc; pd.pop_back(); continue; }
      b.parse_rc[i] = 0;
      b.descs[i] = pd.back().frame;
      prog.push_back(i);
      b.rgb_off[i] = rgb_total;
      rgb_total += ((size_t)3 * b.descs[i].image_width * b.descs[i].image_height + 255) / 256 * 256;
    }
  if (b.pixels.size() < rgb_total + 256) b.pixels = PinnedBuffer(rgb_total + 256 + rgb_total / 8);
  b.ms_alloc = since(t0); t0 = Clock::now();
  if (!prog.empty()) {
    const size_t m = prog.size();
    std::vector<const uint8_t *> f(m);
    std::vector<size_t> sz(m);
    std::vector<uint8_t *> out(m);
    std::vector<int32_t> st(m, 0);
    for (size_t k = 0; k < m; ++k) { f[k] = fptr[prog[k]]; sz[k] = b.file_len[prog[k]]; out[k] = b.pixels.data() + b.rgb_off[prog[k]]; }
    Context own(devices.empty() ? 0 : devices[0]);
    const int rc = b200jpeg_decode_progressive(own.get(), pd.data(), (int)m, f.data(), sz.data(), out.data(), nullptr, st.data());
    if (rc) throw Error(rc, std::string("decode_files (progressive): ") + b200jpeg_last_error(own.get()));
    for (size_t k = 0; k < m; ++k) b.status[prog[k]] = st[k];
  }
  if (live.empty()) return;
  const size_t m = live.size();
  std::vector<b200jpeg_image_desc> d(m);
  std::vector<const uint8_t *> f(m);
  std::vector<size_t> sz(m);
  std::vector<uint8_t *> out(m);
  std::vector<int32_t> st(m, 0);
  for (size_t k = 0; k < m; ++k) {
    const size_t i = live[k];
    d[k] = b.descs[i]; f[k] = fptr[i]; sz[k] = b.file_len[i]; out[k] = b.pixels.data() + b.rgb_off[i];
  }
  const std::vector<int> devs = devices.empty() ? std::vector<int>(1, 0) : devices;
  const int rc = b200jpeg_decode_batch_multi(devs.data(), (int)devs.size(), d.data(), (int)m, f.data(), sz.data(),
                                             out.data(), st.data());
  if (rc) throw Error(rc, "decode_files: b200jpeg_decode_batch_multi failed (no usable sm_100 GPU? there is no CPU fallback)");
  for (size_t k = 0; k < m; ++k) b.status[live[k]] = st[k];
  b.ms_decode = since(t0);
}

class Image {
  std::string filename_;
  std::vector<uint8_t> file_;          // the whole file, read once
  b200jpeg_image_desc desc_;           // Image's private state at the seam (jpeg.cpp:21-34)
  std::vector<uint8_t> rgb_;           // interleaved R,G,B, row-major, top-down
  std::vector<int16_t> coef_;          // quantised coefficients, MCU-array layout, kept on request
  int status_ = 0;
  bool parsed_ = false, decoded_ = false;

  void load_tables() {                 // struct mirror -> the reference's objects
    for (int t = 0; t < 4; ++t) {
      const b200jpeg_huffman_table *src[2] = {&desc_.huffmanDCTTable[t], &desc_.huffmanACTable[t]};
      HuffmanTable *dst[2] = {&huffmanDCTTable[t], &huffmanACTable[t]};
      for (int k = 0; k < 2; ++k) {
        std::memcpy(dst[k]->offset, src[k]->offset, 17);
        std::memcpy(dst[k]->symbols, src[k]->symbols, 176);
        dst[k]->set = src[k]->set != 0;
        if (dst[k]->set) dst[k]->generateCodes();
      }
      if (desc_.quantization_tables[t].set) {
        unsigned char raw[64];
        for (int i = 0; i < 64; ++i) raw[i] = (unsigned char)desc_.quantization_tables[t].table[zigZagMap[i]];
        quantization_tables[t] = QuantizationTable(t, raw);
      }
    }
    for (int c = 0; c < 3; ++c) {
      const b200jpeg_component &s = desc_.color_components[c];
      color_components[c].horizontal_sampling_factor = s.horizontal_sampling_factor;
      color_components[c].vertical_sampling_factor = s.vertical_sampling_factor;
      color_components[c].quantizationTableID = s.quantizationTableID;
      color_components[c].huffmanDCTTableID = s.huffmanDCTTableID;
      color_components[c].huffmanACTableID = s.huffmanACTableID;
      color_components[c].set = c < desc_.numComponents;
    }
  }

 public:
  // jpeg.cpp:792-795
  int image_width = 0, image_height = 0, mcuWidthReal = 0, mcuHeightReal = 0;
  // jpeg.cpp:22-27 (public here: they are the drop-in seam)
  QuantizationTable quantization_tables[4];
  ColorComponent color_components[3];
  HuffmanTable huffmanDCTTable[4], huffmanACTable[4];
  int restartInterval = 0;
  int parse_flags = B200JPEG_PARSE_REFERENCE;   // B200JPEG_PARSE_SPEC: grey scans, COM segments, ...
  bool keep_coefficients = false;               // also fetch the MCU array after decodeHuffman()

  // Image::Image, jpeg.cpp:797-807
  explicit Image(const std::string &filename) : filename_(filename) {
    std::memset(&desc_, 0, sizeof desc_);
    std::ifstream f(filename, std::ios::binary);
    if (!f) throw Error(B200JPEG_E_NOT_JPEG, "cannot open " + filename);
    file_.assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
    if (file_.size() < 2 || file_[0] != 0xff || file_[1] != 0xd8) throw Error(B200JPEG_E_NOT_JPEG, "Invalid JPEG file");
  }
  Image(const uint8_t *bytes, size_t n) : filename_("<memory>"), file_(bytes, bytes + n) {
    std::memset(&desc_, 0, sizeof desc_);
    if (n < 2 || bytes[0] != 0xff || bytes[1] != 0xd8) throw Error(B200JPEG_E_NOT_JPEG, "Invalid JPEG file");
  }

  // The marker loop up to SOS (jpeg.cpp:826-867) -- host only, no GPU needed.
  void readHeaders() {
    const int rc = b200jpeg_parse_header(file_.data(), file_.size(), parse_flags, &desc_);
    if (rc) throw Error(rc, "readJPEG: " + filename_);
    image_width = desc_.image_width; image_height = desc_.image_height;
    mcuWidthReal = desc_.mcuWidthReal; mcuHeightReal = desc_.mcuHeightReal;
    restartInterval = desc_.restartInterval;
    load_tables();
    parsed_ = true;
  }

  // jpeg.cpp:755-789, on the GPU: decodeHuffman + dequantize + inverseDCT + YCbCrToRGB.
  void process_image_data(Context &ctx = Context::shared()) {
    if (!parsed_) readHeaders();
    rgb_.assign((size_t)3 * image_width * image_height, 0);
    const uint8_t *files[1] = {file_.data()};
    const size_t sizes[1] = {file_.size()};
    uint8_t *out[1] = {rgb_.data()};
    int16_t *co[1] = {nullptr};
    if (keep_coefficients) { coef_.assign((size_t)mcuWidthReal * mcuHeightReal * 3 * 64, 0); co[0] = coef_.data(); }
    int32_t st = 0;
    const int rc = b200jpeg_decode_batch(ctx.get(), &desc_, 1, files, sizes, out, keep_coefficients ? co : nullptr, &st);
    if (rc) throw Error(rc, std::string("process_image_data: ") + b200jpeg_last_error(ctx.get()));
    status_ = st;
    if (st) throw Error(B200JPEG_E_STATE, "process_image_data: corrupt scan, status bits " + std::to_string(st));
    decoded_ = true;
  }

  // jpeg.cpp:826-907: headers, then the scan (the reference calls process_image_data at SOS, :858-863).
  // Progressive files (SOF2): the reference refuses them (jpeg.cpp:69-73) and so does this call under
  // B200JPEG_PARSE_REFERENCE; with parse_flags = B200JPEG_PARSE_SPEC the refusal is lifted (readProgressive).
  void readJPEG(Context &ctx = Context::shared()) {
    b200jpeg_image_desc probe;
    if ((parse_flags & B200JPEG_PARSE_SPEC) &&
        b200jpeg_parse_header(file_.data(), file_.size(), parse_flags, &probe) == B200JPEG_E_PROGRESSIVE) { readProgressive(ctx); return; }
    readHeaders();
    process_image_data(ctx);
  }

  // SOF2: frame + scans through b200jpeg_parse_progressive / b200jpeg_decode_progressive (T.81 Annex G on the GPU).
  void readProgressive(Context &ctx = Context::shared()) {
    std::vector<b200jpeg_progressive_desc> pd(1);
    int rc = b200jpeg_parse_progressive(file_.data(), file_.size(), &pd[0]);
    if (rc) throw Error(rc, "readJPEG (progressive): " + filename_);
    desc_ = pd[0].frame;
    image_width = desc_.image_width; image_height = desc_.image_height;
    mcuWidthReal = desc_.mcuWidthReal; mcuHeightReal = desc_.mcuHeightReal;
    restartInterval = desc_.restartInterval;
    load_tables();
    parsed_ = true;
    rgb_.assign((size_t)3 * image_width * image_height, 0);
    const uint8_t *files[1] = {file_.data()};
    const size_t sizes[1] = {file_.size()};
    uint8_t *out[1] = {rgb_.data()};
    int16_t *co[1] = {nullptr};
    if (keep_coefficients) { coef_.assign((size_t)mcuWidthReal * mcuHeightReal * 3 * 64, 0); co[0] = coef_.data(); }
    int32_t st = 0;
    rc = b200jpeg_decode_progressive(ctx.get(), pd.data(), 1, files, sizes, out, keep_coefficients ? co : nullptr, &st);
    if (rc) throw Error(rc, std::string("readJPEG (progressive): ") + b200jpeg_last_error(ctx.get()));
    status_ = st;
    if (st) throw Error(B200JPEG_E_STATE, "readJPEG (progressive): corrupt scan, status bits " + std::to_string(st));
    decoded_ = true;
  }

  // Many images, one call: headers on the host, every scan through ONE b200jpeg_decode_batch()
  // (images are independent; the library cuts the batch into chunks that rotate over four internal pipelines).  Returns the per-image
  // status words (0 = decoded); images[i]->rgb() holds the pixels.  Nothing throws for a corrupt scan.
  static std::vector<int> decode_batch(const std::vector<Image *> &images, Context &ctx = Context::shared()) {
    const int n = (int)images.size();
    std::vector<b200jpeg_image_desc> descs((size_t)n);
    std::vector<const uint8_t *> files((size_t)n);
    std::vector<size_t> sizes((size_t)n);
    std::vector<uint8_t *> out((size_t)n);
    std::vector<int32_t> st((size_t)n, 0);
    for (int i = 0; i < n; ++i) {
      Image &im = *images[(size_t)i];
      if (!im.parsed_) im.readHeaders();
      im.rgb_.assign((size_t)3 * im.image_width * im.image_height, 0);
      descs[(size_t)i] = im.desc_; files[(size_t)i] = im.file_.data(); sizes[(size_t)i] = im.file_.size(); out[(size_t)i] = im.rgb_.data();
    }
    if (n == 0) return std::vector<int>();
    const int rc = b200jpeg_decode_batch(ctx.get(), descs.data(), n, files.data(), sizes.data(), out.data(), nullptr, st.data());
    if (rc) throw Error(rc, std::string("decode_batch: ") + b200jpeg_last_error(ctx.get()));
    for (int i = 0; i < n; ++i) { images[(size_t)i]->status_ = st[(size_t)i]; images[(size_t)i]->decoded_ = st[(size_t)i] == 0; }
    return std::vector<int>(st.begin(), st.end());
  }

  // The marker segments of the file up to and including SOS, as the reference's marker reader classifies them
  // (Marker, include/file.hpp:183-258) -- the loop of readJPEG (jpeg.cpp:826-867) without the dispatch.
  std::vector<Marker> markers() const {
    std::vector<Marker> out;
    size_t pos = 2;                                    // behind SOI, where readJPEG starts (jpeg.cpp:797-807, 826)
    while (pos + 1 < file_.size()) {
      if (file_[pos] != 0xff) break;                   // jpeg.cpp:899-903: "Invalid Marker"
      Marker m(file_.data(), file_.size(), pos + 1);
      out.push_back(m);
      pos = m.next;
      if (m.type == MarkerType::SOS || m.type == MarkerType::EOI || m.type == MarkerType::INVALID) break;
    }
    return out;
  }

  // Results decoded elsewhere (b200jpeg::decode_files) handed to an Image, for its sinks (saveToBMP, mcus, ...).
  void adopt(const b200jpeg_image_desc &d, const uint8_t *rgb) {
    desc_ = d;
    image_width = d.image_width; image_height = d.image_height;
    mcuWidthReal = d.mcuWidthReal; mcuHeightReal = d.mcuHeightReal;
    restartInterval = d.restartInterval;
    load_tables();
    rgb_.assign(rgb, rgb + (size_t)3 * image_width * image_height);
    parsed_ = decoded_ = true; status_ = 0;
  }

  // Many images over several GPUs: contiguous ranges of `images`, one context + host thread per device
  // (b200jpeg_decode_batch_multi).  Same results as decode_batch(images).
  static std::vector<int> decode_batch(const std::vector<Image *> &images, const std::vector<int> &devices) {
    const int n = (int)images.size();
    if (n == 0) return std::vector<int>();
    std::vector<b200jpeg_image_desc> descs((size_t)n);
    std::vector<const uint8_t *> files((size_t)n);
    std::vector<size_t> sizes((size_t)n);
    std::vector<uint8_t *> out((size_t)n);
    std::vector<int32_t> st((size_t)n, 0);
    for (int i = 0; i < n; ++i) {
      Image &im = *images[(size_t)i];
      if (!im.parsed_) im.readHeaders();
      im.rgb_.assign((size_t)3 * im.image_width * im.image_height, 0);
      descs[(size_t)i] = im.desc_; files[(size_t)i] = im.file_.data(); sizes[(size_t)i] = im.file_.size(); out[(size_t)i] = im.rgb_.data();
    }
    const int rc = b200jpeg_decode_batch_multi(devices.data(), (int)devices.size(), descs.data(), n, files.data(), sizes.data(), out.data(), st.data());
    if (rc) throw Error(rc, "decode_batch: b200jpeg_decode_batch_multi failed (no usable sm_100 GPU? there is no CPU fallback)");
    for (int i = 0; i < n; ++i) { images[(size_t)i]->status_ = st[(size_t)i]; images[(size_t)i]->decoded_ = st[(size_t)i] == 0; }
    return std::vector<int>(st.begin(), st.end());
  }

  const b200jpeg_image_desc &desc() const { return desc_; }
  const std::vector<uint8_t> &file() const { return file_; }
  int status() const { return status_; }
  // Headless display(): interleaved RGB8, pitch 3*image_width (the display.hpp:27-34 gather flattened).
  const std::vector<uint8_t> &rgb() const { return rgb_; }
  const std::vector<int16_t> &coefficients() const { return coef_; }

  // The reference's MCU array after YCbCrToRGB (r/g/b planes per block position), rebuilt from rgb().
  std::vector<MCU> mcus() const {
    std::vector<MCU> m((size_t)mcuWidthReal * mcuHeightReal);
    for (int yy = 0; yy < image_height; ++yy)
      for (int xx = 0; xx < image_width; ++xx) {
        MCU &b = m[(size_t)(yy / 8) * mcuWidthReal + xx / 8];
        const uint8_t *p = &rgb_[((size_t)yy * image_width + xx) * 3];
        const int k = (yy % 8) * 8 + xx % 8;
        b.r[k] = p[0]; b.g[k] = p[1]; b.b[k] = p[2];
      }
    return m;
  }

  // jpeg.cpp:448-509 / :809-816 with what it gets wrong put right: BITMAPCOREHEADER, bottom-up rows,
  // B,G,R byte order, rows padded to 4 bytes.
  void saveToBMP(const std::string &filename) const {
    if (!decoded_) throw Error(B200JPEG_E_STATE, "saveToBMP before readJPEG");
    const int pad = (4 - (image_width * 3) % 4) % 4;
    const uint32_t size = 14 + 12 + (uint32_t)(image_width * 3 + pad) * image_height;
    std::ofstream o(filename, std::ios::binary);
    auto put16 = [&](int v) { o.put((char)(v & 255)); o.put((char)((v >> 8) & 255)); };
    auto put32 = [&](uint32_t v) { put16((int)(v & 0xffff)); put16((int)(v >> 16)); };
    o.put('B'); o.put('M'); put32(size); put32(0); put32(26);
    put32(12); put16(image_width); put16(image_height); put16(1); put16(24);
    for (int yy = image_height - 1; yy >= 0; --yy) {
      for (int xx = 0; xx < image_width; ++xx) {
        const uint8_t *p = &rgb_[((size_t)yy * image_width + xx) * 3];
        o.put((char)p[2]); o.put((char)p[1]); o.put((char)p[0]);
      }
      for (int k = 0; k < pad; ++k) o.put(0);
    }
  }
  void savePPM(const std::string &filename) const {
    if (!decoded_) throw Error(B200JPEG_E_STATE, "savePPM before readJPEG");
    std::ofstream o(filename, std::ios::binary);
    o << "P6\n" << image_width << " " << image_height << "\n255\n";
    o.write(reinterpret_cast<const char *>(rgb_.data()), (std::streamsize)rgb_.size());
  }
};

}  // namespace b200jpeg
#endif
