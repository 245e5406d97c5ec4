// dev_types.h -- plain structs shared by the host planner and the kernels.
//
// Vocabulary follows the reference: a "block" is one 8x8 unit (what jpeg.cpp calls an MCU
// position, types.hpp:32-50), an "MCU" here is the JPEG MCU (hs*vs luma blocks + Cb + Cr),
// a "segment" is the entropy-coded data between two RSTn markers (file.hpp:83-86,
// jpeg.cpp:419-425), a "sub" (subsequence) is a fixed-size slice of a segment that one thread
// walks in the self-synchronising pass, a "unit" is what one lane of the segment decoder turns into
// coefficient blocks: a whole segment, or a run of consecutive subs whose start state the
// self-synchronising pass has settled.
#pragma once
#include <stdint.h>

namespace b2 {

constexpr int kMaxBpm = 6;          // blocks per MCU: hs*vs + 2 with hs,vs in {1,2}; 1 for grey
constexpr uint32_t kNone = 0xffffffffu;
enum : int32_t { kPathSync = 0, kPathSegment = 1 };   // kernels_entropy.cuh / kernels_segdec.cuh

// per-image status bits (0 = decoded)
enum : uint32_t {
  ST_BAD_MARKER = 1u << 0,     // FF xx other than 00/D0-D7/FF/D9 inside the scan (file.hpp:91-95)
  ST_NO_EOI = 1u << 1,         // no FF D9 after the scan (file.hpp:59-104 never returns)
  ST_BAD_RST_COUNT = 1u << 2,  // number of RSTn markers != ceil(MCUs/RI)-1
  ST_BAD_CODE = 1u << 3,       // no Huffman code matches (jpeg.cpp:300-320 returns 255)
  ST_BAD_DC_LEN = 1u << 4,     // DC category > 11 (jpeg.cpp:330-334)
  ST_BAD_RUN = 1u << 5,        // zero run leaves the block (jpeg.cpp:372-376)
  ST_BAD_AC_LEN = 1u << 6,     // AC size > 10 (jpeg.cpp:381-385)
  ST_OVERRUN = 1u << 7,        // a block runs past the end of its segment (file.hpp:36-40)
  ST_BLOCK_COUNT = 1u << 8,    // a segment holds fewer blocks than RI MCUs
  ST_DC_RANGE = 1u << 9,       // absolute DC does not fit int16 (reference keeps int32)
  ST_NOT_CONVERGED = 1u << 10, // self-synchronisation needed more rounds than were launched
  ST_UNSUPPORTED = 1u << 11,   // geometry outside what the kernels implement
};

// DHT as parsed into HuffmanTable (huffman.hpp:8-15) minus the derivable `codes`.
struct HuffSpec {
  uint8_t offset[17];     // cumulative counts, offset[i+1]-offset[i] codes of length i+1
  uint8_t symbols[176];   // jpeg.cpp:177 allows up to 176
  uint8_t set;
  uint8_t pad[2];
};

// One distinct set of tables; images of a batch share it when their DQT/DHT are identical.
struct TableSet {
  uint16_t q[4][64];      // QuantizationTable::table (natural order, types.hpp:86-92); 8-bit in the reference, 16-bit under PARSE_SPEC
  HuffSpec dc[4], ac[4];
};

struct ImageDev {
  int32_t width, height;
  int32_t mcux, mcuy;           // JPEG MCUs per row / rows
  int32_t hs, vs, ncomp, bpm;
  int32_t restart_interval;     // in MCUs, 0 = none
  int32_t tset;                 // index into the TableSet array
  int8_t tq[4], td[4], ta[4];   // per component, bound by position (jpeg.cpp:241-247)
  uint32_t raw_len;             // stuffed bytes available from the first scan byte
  uint32_t pad0;
  uint64_t raw_off;             // byte offset of the scan in d_raw; unstuffed copy at the same offset in d_us
  int32_t seg_base, nseg;       // segment table entries [seg_base, seg_base+nseg]
  int32_t tile_base, ntiles;    // unstuff tiles (kUTile raw bytes each)
  int32_t sub_base, nsub_max;   // self-synchronising path: first sub slot of the image (global numbering) / slots reserved
  int32_t unit_base, nunit_max; // self-synchronising path: first unit slot / slots reserved
  int32_t nblocks;              // coded blocks = mcux*mcuy*bpm
  int32_t path;                 // kPathSync / kPathSegment: which entropy decoder takes this image
  int64_t coef_base;            // first block of this image in the coefficient buffer
  uint64_t rgb_off;             // byte offset of this image in the RGB buffer
  uint32_t rgb_pitch;           // bytes per row on the device (3*width rounded up to 16)
  uint32_t pad2;
};

// Written by the kernels, one per image.
struct ImageOut {
  uint32_t end_pos;   // raw position of the D9 of the first FF D9 (kNone if absent)
  uint32_t err_pos;   // raw position of the first illegal marker byte (kNone if none)
  uint32_t us_len;    // unstuffed bytes before EOI
  uint32_t nrst;      // RSTn markers before EOI
  uint32_t status;
  uint32_t pad[3];
};

// Per-subsequence synchronisation state.
struct SubState { uint32_t p, blk; };           // first block start at/after the sub's end
struct SubInfo { int32_t nb_flag, d0, d1, d2; };  // blocks / DC sums since the last segment start; bit31 = a segment starts here

// What a lane of the segment decoder needs to decode a run of subs of the self-synchronising path (written by
// k_xunits once the start states are settled): bit range, first / past-the-last block of the image, block-in-MCU phase
// and DC predictors at the start.
struct XUnit {
  uint32_t p0, pend;          // first bit / where the next unit starts (the segment's end for the last one)
  uint32_t bidx, bend;        // blocks [bidx, bend) of the image
  int32_t pred[3];            // DC predictors at p0 (jpeg.cpp:344-345: the running sums since the segment's start)
  uint32_t blk0;              // block-in-MCU index of the block that starts at p0
};
// Batch-wide parameters of the self-synchronising path (chosen by the planner from the batch's size).
struct XParams {
  int32_t sub_shift;          // log2 of the sub size in bits
  int32_t unit_subs;          // subs per unit (units never span segments)
  int32_t overlap;            // run-in in bits before a sub whose start is a guess
  int32_t nimg;
  int32_t nsub_slots, nunit_slots;   // totals (multiples of the CTA sizes)
  int32_t ncta;               // CTAs of the sync kernels
  int32_t pad;
};

}  // namespace b2
