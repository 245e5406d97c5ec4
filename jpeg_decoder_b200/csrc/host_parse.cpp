// host_parse.cpp -- host-side marker reader / header parser behind b200jpeg_parse_header().
//
// Mirrors, on one contiguous buffer instead of a per-byte ifstream, what the reference does
// between Image::Image (jpeg.cpp:797-807) and the call of process_image_data() (jpeg.cpp:861):
//   Marker::Marker          include/file.hpp:192-258   marker classification + length + payload
//   Image::readJPEG         jpeg.cpp:826-907           dispatch loop
//   read_sof                jpeg.cpp:67-146
//   read_huffman_table      jpeg.cpp:148-196  (+ HuffmanTable::generateCodes, huffman.hpp:17-29)
//   read_quantization_table jpeg.cpp:197-231  (+ QuantizationTable ctor, types.hpp:86-92)
//   read_sos                jpeg.cpp:233-287
//   read_dri                jpeg.cpp:289-298
//   table checks            jpeg.cpp:757-774
// Every exit(1) of the reference becomes a negative return code.  The scan itself is NOT
// touched here: readImageData (file.hpp:59-104) is replaced by the unstuff kernels.
#include "../../include/b200jpeg.h"
#include <string.h>
#include <atomic>
#include <thread>
#include <vector>

namespace {

const int kZigZag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                         41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                         30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

inline int be16(const uint8_t *p) { return (p[0] << 8) | p[1]; }

int parse_sof(b200jpeg_image_desc *d, int marker, const uint8_t *p, int len, bool spec, int *comp_id) {
  if (marker == 0xc2) return B200JPEG_E_PROGRESSIVE;
  if (len < 6) return B200JPEG_E_TRUNCATED;
  d->image_height = be16(p + 1);
  d->image_width = be16(p + 3);
  d->mcuWidth = (d->image_width + 7) / 8;
  d->mcuHeight = (d->image_height + 7) / 8;
  d->mcuWidthReal = d->mcuWidth;
  d->mcuHeightReal = d->mcuHeight;
  const int nc = p[5];
  d->numComponents = nc;
  if (nc != 3 && !(spec && nc == 1)) return B200JPEG_E_COMPONENTS;
  if (len < 6 + 3 * nc) return B200JPEG_E_TRUNCATED;
  for (int i = 0; i < nc; ++i) {
    const int id = p[6 + 3 * i];
    comp_id[i] = id;
    if (id == 0) d->zeroBased = 1;
    if (!spec && !d->zeroBased && id > 3) return B200JPEG_E_COMPONENT_ID;
    if (!spec && d->zeroBased && id > 2) return B200JPEG_E_COMPONENT_ID;
    b200jpeg_component *c = &d->color_components[i];
    c->horizontal_sampling_factor = p[7 + 3 * i] >> 4;
    c->vertical_sampling_factor = p[7 + 3 * i] & 15;
    c->quantizationTableID = p[8 + 3 * i];
    if (nc == 1) {  // T.81 A.2.2: a one-component scan has one block per MCU whatever Hi,Vi say
      c->horizontal_sampling_factor = c->vertical_sampling_factor = 1;
      d->horizontalSamplingFactor = d->verticalSamplingFactor = 1;
      continue;
    }
    // the reference takes identifier 1 for luma (jpeg.cpp:110); T.81 B.2.2 allows any byte, so under
    // B200JPEG_PARSE_SPEC luma is the frame's first component
    if (spec ? i == 0 : id == 1) {
      const int h = c->horizontal_sampling_factor, v = c->vertical_sampling_factor;
      if ((h != 1 && h != 2) || (v != 1 && v != 2)) return B200JPEG_E_SAMPLING;
      if (h == 2 && d->mcuWidth % 2 == 1) d->mcuWidthReal += 1;
      if (v == 2 && d->mcuHeight % 2 == 1) d->mcuHeightReal += 1;
      d->horizontalSamplingFactor = h;
      d->verticalSamplingFactor = v;
      if (i != 0 && (h != 1 || v != 1)) return B200JPEG_E_UNSUPPORTED;
    } else if (c->horizontal_sampling_factor != 1 || c->vertical_sampling_factor != 1) {
      return B200JPEG_E_SAMPLING;
    }
  }
  return 0;
}

int parse_dht(b200jpeg_image_desc *d, const uint8_t *p, int len) {
  while (len > 0) {
    if (len < 17) return B200JPEG_E_DHT;
    const int id = p[0] & 15, is_ac = p[0] >> 4;
    if (id > 3) return B200JPEG_E_TABLE_ID;
    b200jpeg_huffman_table *t = is_ac ? &d->huffmanACTable[id] : &d->huffmanDCTTable[id];
    t->offset[0] = 0;
    int all = 0;
    for (int i = 1; i <= 16; ++i) { all += p[i]; t->offset[i] = (uint8_t)all; }
    if (all > 176) return B200JPEG_E_DHT;
    if (len < 17 + all) return B200JPEG_E_DHT;
    memset(t->symbols, 0, sizeof t->symbols);
    for (int i = 0; i < all; ++i) t->symbols[i] = p[17 + i];
    t->set = 1;
    uint32_t code = 0;  // generateCodes(), huffman.hpp:17-29
    memset(t->codes, 0, sizeof t->codes);
    for (int i = 0; i < 16; ++i) {
      for (int j = t->offset[i]; j < t->offset[i + 1]; ++j) t->codes[j] = code++;
      code <<= 1;
    }
    p += 17 + all;
    len -= 17 + all;
  }
  return len != 0 ? B200JPEG_E_DHT : 0;
}

int parse_dqt(b200jpeg_image_desc *d, const uint8_t *p, int len, bool spec) {
  while (len > 0) {
    const int id = p[0] & 15, wide = p[0] >> 4;
    ++p; --len;
    if (id > 3) return B200JPEG_E_TABLE_ID;
    uint8_t raw[64];
    if (wide) {
      if (len < 128) return B200JPEG_E_TRUNCATED;
      for (int i = 0; i < 64; ++i) raw[i] = p[2 * i + 1];  // jpeg.cpp:216 keeps the low byte only
      if (spec) {                                          // T.81 B.2.4.1, Pq = 1: the full 16-bit quantiser
        b200jpeg_quant_table *q = &d->quantization_tables[id];
        for (int i = 0; i < 64; ++i) q->table[kZigZag[i]] = be16(p + 2 * i);
        q->set = 1;
        p += 128; len -= 128;
        continue;
      }
      p += 128; len -= 128;
    } else {
      if (len < 64) return B200JPEG_E_TRUNCATED;
      memcpy(raw, p, 64);
      p += 64; len -= 64;
    }
    b200jpeg_quant_table *q = &d->quantization_tables[id];
    for (int i = 0; i < 64; ++i) q->table[kZigZag[i]] = raw[i];  // types.hpp:86-92
    q->set = 1;
  }
  return 0;
}

int parse_sos(b200jpeg_image_desc *d, const uint8_t *p, int len, bool spec, const int *comp_id) {
  if (len < 1) return B200JPEG_E_SOS;
  const int nc = p[0];
  if (nc != d->numComponents || (nc != 3 && nc != 1)) return B200JPEG_E_COMPONENTS;
  if (len < 1 + 2 * nc + 3) return B200JPEG_E_SOS;
  for (int i = 0; i < nc; ++i) {  // bound by position, the id byte is ignored (jpeg.cpp:241-247)
    // B200JPEG_PARSE_SPEC: bound by component selector (T.81 B.2.3); the selectors must name the frame's
    // components in the frame's order -- the only scan the interleaved decode loop (jpeg.cpp:415-435) takes
    if (spec && p[1 + 2 * i] != comp_id[i]) return B200JPEG_E_SOS;
    d->color_components[i].huffmanDCTTableID = p[2 + 2 * i] >> 4;
    d->color_components[i].huffmanACTableID = p[2 + 2 * i] & 15;
  }
  const uint8_t *q = p + 1 + 2 * nc;
  if (q[0] != 0 || q[1] != 63) return B200JPEG_E_SOS;
  if (q[2] != 0) return B200JPEG_E_SOS;
  if (len + 2 - 6 - 2 * nc != 0) return B200JPEG_E_SOS;
  return 0;
}

}  // namespace

extern "C" int b200jpeg_parse_header(const uint8_t *file, size_t len, int flags, b200jpeg_image_desc *d) {
  if (!file || !d) return B200JPEG_E_ARG;
  memset(d, 0, sizeof *d);
  const bool spec = (flags & B200JPEG_PARSE_SPEC) != 0;
  if (len < 2 || file[0] != 0xff || file[1] != 0xd8) return B200JPEG_E_NOT_JPEG;
  size_t pos = 2;
  bool got_sos = false;
  int comp_id[3] = {-1, -1, -1};                            // SOF component identifiers (used in SPEC mode)
  while (!got_sos) {
    if (pos >= len) return B200JPEG_E_NO_SOS;
    if (file[pos++] != 0xff) return B200JPEG_E_NO_SOS;       // jpeg.cpp:899-903
    if (pos >= len) return B200JPEG_E_TRUNCATED;
    const int m = file[pos++];
    if (m == 0x00) continue;                                 // PAD, file.hpp:196-200
    if (m >= 0xd0 && m < 0xd8) continue;                     // DRM, file.hpp:208-213
    if (spec && m == 0xff) { --pos; continue; }              // fill byte, T.81 B.1.1.2
    if (pos + 2 > len) return B200JPEG_E_TRUNCATED;
    const int seglen = be16(file + pos);
    const uint8_t *p = file + pos + 2;
    const int plen = seglen - 2;
    if (plen < 0 || pos + (size_t)seglen > len) return B200JPEG_E_TRUNCATED;
    pos += (size_t)seglen;
    if (m >= 0xe0 && m < 0xf0) continue;                     // META, file.hpp:201-207
    int rc = 0;
    switch (m) {
      case 0xc0: case 0xc2: rc = parse_sof(d, m, p, plen, spec, comp_id); break;
      case 0xc4: rc = parse_dht(d, p, plen); break;
      case 0xdb: rc = parse_dqt(d, p, plen, spec); break;
      case 0xdd:
        if (plen != 2) rc = B200JPEG_E_DRI; else d->restartInterval = be16(p);
        break;
      case 0xda: rc = parse_sos(d, p, plen, spec, comp_id); got_sos = true; break;
      case 0xd8: break;                                      // SOI is only logged, jpeg.cpp:838-841
      default:
        if (spec && m != 0xd9 && !(m >= 0xc1 && m <= 0xcf)) break;   // COM and friends
        if (spec && m != 0xd9) { rc = B200JPEG_E_PROGRESSIVE; break; }
        rc = B200JPEG_E_NO_SOS;                              // EOI / INVALID, jpeg.cpp:878-887
    }
    if (rc) return rc;
  }
  if (d->numComponents == 0 || d->horizontalSamplingFactor == 0 || d->verticalSamplingFactor == 0)
    return B200JPEG_E_SAMPLING;
  for (int i = 0; i < d->numComponents; ++i) {               // jpeg.cpp:757-774
    const b200jpeg_component &c = d->color_components[i];
    if (c.quantizationTableID > 3 || c.huffmanDCTTableID > 3 || c.huffmanACTableID > 3) return B200JPEG_E_TABLE_ID;
    if (!d->quantization_tables[c.quantizationTableID].set || !d->huffmanDCTTable[c.huffmanDCTTableID].set ||
        !d->huffmanACTable[c.huffmanACTableID].set)
      return B200JPEG_E_TABLE_MISSING;
  }
  d->scan_offset = (int64_t)pos;
  d->scan_length = (int64_t)(len - pos);
  return B200JPEG_OK;
}

// Progressive files (SOF2): frame + scan list + table pool, see b200jpeg.h.  The marker walk is the one above; SOF2 has
// the layout of SOF0 (T.81 B.2.2), every DHT appends to the pool, every SOS records which pool entries are in force.
extern "C" int b200jpeg_parse_progressive(const uint8_t *file, size_t len, b200jpeg_progressive_desc *P) {
  if (!file || !P) return B200JPEG_E_ARG;
  memset(P, 0, sizeof *P);
  if (len < 2 || file[0] != 0xff || file[1] != 0xd8) return B200JPEG_E_NOT_JPEG;
  b200jpeg_image_desc *d = &P->frame;
  int comp_id[3] = {-1, -1, -1}, cur_dc[4] = {-1, -1, -1, -1}, cur_ac[4] = {-1, -1, -1, -1};
  bool got_frame = false;
  size_t pos = 2;
  while (pos + 1 < len) {
    if (file[pos] != 0xff) return B200JPEG_E_NO_SOS;
    const int m = file[pos + 1];
    if (m == 0xff) { ++pos; continue; }
    pos += 2;
    if (m == 0xd9) break;
    if (m == 0x00 || (m >= 0xd0 && m <= 0xd8)) continue;
    if (pos + 2 > len) return B200JPEG_E_TRUNCATED;
    const int seglen = be16(file + pos);
    const uint8_t *p = file + pos + 2;
    const int plen = seglen - 2;
    if (plen < 0 || pos + (size_t)seglen > len) return B200JPEG_E_TRUNCATED;
    pos += (size_t)seglen;
    int rc = 0;
    switch (m) {
      case 0xc0: return B200JPEG_E_ARG;                      // a baseline file: b200jpeg_parse_header
      case 0xc2: rc = parse_sof(d, 0xc0, p, plen, true, comp_id); got_frame = rc == 0; break;
      case 0xc1: case 0xc3: case 0xc5: case 0xc6: case 0xc7: case 0xc9: case 0xca: case 0xcb: case 0xcd: case 0xce: case 0xcf:
        return B200JPEG_E_UNSUPPORTED;                       // extended / lossless / arithmetic frames
      case 0xdb: rc = parse_dqt(d, p, plen, true); break;
      case 0xdd: if (plen != 2) rc = B200JPEG_E_DRI; else d->restartInterval = be16(p); break;
      case 0xc4: {                                           // every table of the segment goes to the pool
        const uint8_t *q = p; int left = plen;
        while (left > 0) {
          if (left < 17) return B200JPEG_E_DHT;
          int all = 0;
          for (int i = 1; i <= 16; ++i) all += q[i];
          if (all > 176 || left < 17 + all) return B200JPEG_E_DHT;
          if (P->ntables >= B200JPEG_PROG_MAX_TABLES) return B200JPEG_E_UNSUPPORTED;
          b200jpeg_image_desc tmp_holder;                    // parse_dht writes into a descriptor: borrow its slot 0 / class
          memset(&tmp_holder, 0, sizeof tmp_holder);
          uint8_t one[17 + 176];
          memcpy(one, q, (size_t)(17 + all));
          const int id = q[0] & 15, is_ac = q[0] >> 4;
          if (id > 3) return B200JPEG_E_TABLE_ID;
          rc = parse_dht(&tmp_holder, one, 17 + all);
          if (rc) return rc;
          P->tables[P->ntables] = is_ac ? tmp_holder.huffmanACTable[id] : tmp_holder.huffmanDCTTable[id];
          (is_ac ? cur_ac : cur_dc)[id] = P->ntables++;
          q += 17 + all; left -= 17 + all;
        }
        break;
      }
      case 0xda: {
        if (!got_frame) return B200JPEG_E_NO_SOS;
        if (P->nscans >= B200JPEG_PROG_MAX_SCANS) return B200JPEG_E_UNSUPPORTED;
        if (plen < 1) return B200JPEG_E_SOS;
        b200jpeg_scan *S = &P->scans[P->nscans];
        S->ns = p[0];
        if (S->ns < 1 || S->ns > d->numComponents || plen != 1 + 2 * S->ns + 3) return B200JPEG_E_SOS;
        const uint8_t *q = p + 1 + 2 * S->ns;
        S->ss = q[0]; S->se = q[1]; S->ah = q[2] >> 4; S->al = q[2] & 15;
        if (S->ss > S->se || S->se > 63 || (S->ss == 0 && S->se != 0) || (S->ss != 0 && S->ns != 1) || S->al > 13) return B200JPEG_E_SOS;
        for (int j = 0; j < 3; ++j) { S->comp[j] = 0; S->dc_table[j] = S->ac_table[j] = -1; }
        for (int j = 0; j < S->ns; ++j) {
          int ci = -1;
          for (int i = 0; i < d->numComponents; ++i) if (comp_id[i] == p[1 + 2 * j]) ci = i;
          if (ci < 0) return B200JPEG_E_SOS;
          const int td = p[2 + 2 * j] >> 4, ta = p[2 + 2 * j] & 15;
          if (td > 3 || ta > 3) return B200JPEG_E_TABLE_ID;
          S->comp[j] = ci;
          if (S->ss == 0 && S->ah == 0) { if (cur_dc[td] < 0) return B200JPEG_E_TABLE_MISSING; S->dc_table[j] = cur_dc[td]; }
          if (S->ss != 0) { if (cur_ac[ta] < 0) return B200JPEG_E_TABLE_MISSING; S->ac_table[j] = cur_ac[ta]; }
        }
        S->restart_interval = d->restartInterval;
        S->offset = (int64_t)pos;
        size_t e = pos;                                      // the scan runs up to the next marker that is not RSTn / stuffing
        while (e + 1 < len && !(file[e] == 0xff && file[e + 1] != 0x00 && file[e + 1] != 0xff && !(file[e + 1] >= 0xd0 && file[e + 1] <= 0xd7))) ++e;
        if (e + 1 >= len) e = len;
        S->length = (int64_t)(e - pos);
        if (P->nscans == 0) d->scan_offset = (int64_t)pos;
        pos = e;
        ++P->nscans;
        break;
      }
      default: break;                                        // APPn, COM, ...
    }
    if (rc) return rc;
  }
  if (!got_frame || P->nscans == 0) return B200JPEG_E_NO_SOS;
  if (d->numComponents == 0 || d->horizontalSamplingFactor == 0 || d->verticalSamplingFactor == 0) return B200JPEG_E_SAMPLING;
  for (int i = 0; i < d->numComponents; ++i) {
    const int tq = d->color_components[i].quantizationTableID;
    if (tq < 0 || tq > 3) return B200JPEG_E_TABLE_ID;
    if (!d->quantization_tables[tq].set) return B200JPEG_E_TABLE_MISSING;
  }
  d->scan_length = (int64_t)len - d->scan_offset;
  return B200JPEG_OK;
}

// n headers on several host threads: files are handed out in blocks of 16 through one atomic counter
extern "C" int b200jpeg_parse_headers(const uint8_t *const *files, const size_t *sizes, int n, int flags, int threads,
                                      b200jpeg_image_desc *descs, int32_t *rcs) {
  if (n < 0 || (n > 0 && (!files || !sizes || !descs))) return B200JPEG_E_ARG;
  if (threads <= 0) threads = (int)std::thread::hardware_concurrency();
  if (threads < 1) threads = 1;
  if (threads > 64) threads = 64;
  if (threads > (n + 31) / 32) threads = (n + 31) / 32;      // a header takes a few microseconds: no thread for less than 32
  std::atomic<int> next(0), failed(0);
  auto work = [&]() {
    for (;;) {
      const int lo = next.fetch_add(16);
      if (lo >= n) return;
      const int hi = lo + 16 < n ? lo + 16 : n;
      for (int i = lo; i < hi; ++i) {
        const int rc = b200jpeg_parse_header(files[i], sizes[i], flags, &descs[i]);
        if (rcs) rcs[i] = rc;
        if (rc) failed.fetch_add(1);
      }
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < threads; ++t) pool.emplace_back(work);
  work();
  for (std::thread &t : pool) t.join();
  return failed.load();
}

extern "C" int b200jpeg_abi_version(void) { return B200JPEG_ABI_VERSION; }
