/* b200jpeg.h -- C ABI of the B200-native scan-to-pixels path.
 *
 * Drop-in boundary for aswanthabam/JPEG_Decoder's hot path.  The reference has no FFI; its
 * path is four private member functions run by Image::process_image_data() (jpeg.cpp:755-789),
 * called once from readJPEG() at the SOS marker (jpeg.cpp:858-863).  Everything the reference
 * holds at that seam is mirrored here as plain structs (no C++ or torch types):
 *
 *   b200jpeg_huffman_table  <- HuffmanTable            include/huffman.hpp:8-15
 *   b200jpeg_quant_table    <- QuantizationTable       include/types.hpp:79-96
 *   b200jpeg_component      <- ColorComponent          include/types.hpp:69-77
 *   b200jpeg_image_desc     <- Image's private state   jpeg.cpp:21-34, 792-795
 *   b200jpeg_mcu            <- MCU                     include/types.hpp:32-67
 *
 * and the stages map as
 *
 *   b200jpeg_parse_header() <- Image::Image + readJPEG up to SOS   jpeg.cpp:797-807, 826-867
 *                              (read_sof :67, read_huffman_table :148, read_quantization_table :197,
 *                               read_sos :233, read_dri :289; Marker::Marker file.hpp:192-258)
 *   b200jpeg_decode_batch() <- process_image_data()                jpeg.cpp:755-789
 *                              = readImageData file.hpp:59-104 + decodeHuffman :405-446
 *                              + dequantize :572-590 + inverseDCT :735-753 + YCbCrToRGB :544-561
 *                              + the RGB gather of display.hpp:19-34
 *
 * Error convention: the reference logs and calls exit(1); this library never exits and never
 * throws across the boundary.  Calls return 0 or a negative B200JPEG_E_* code; per-image
 * problems come back in status[] (bit mask, 0 = decoded).
 *
 * There is no CPU fallback: every decode entry point fails with B200JPEG_E_CUDA if no
 * sm_100-class device is usable.
 */
#ifndef B200JPEG_H
#define B200JPEG_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define B200JPEG_ABI_VERSION 2

/* ---- return codes (calls) ---- */
enum {
  B200JPEG_OK = 0,
  B200JPEG_E_ARG = -1,            /* null/invalid argument */
  B200JPEG_E_CUDA = -2,           /* CUDA error or no device (see b200jpeg_last_error) */
  B200JPEG_E_NOMEM = -3,
  B200JPEG_E_STATE = -4,          /* run/fetch without a prepared batch */
  /* header parser: one code per exit(1) of the reference */
  B200JPEG_E_NOT_JPEG = -10,      /* jpeg.cpp:800-805 */
  B200JPEG_E_PROGRESSIVE = -11,   /* jpeg.cpp:69-73 */
  B200JPEG_E_COMPONENTS = -12,    /* jpeg.cpp:83-87, 236-240 */
  B200JPEG_E_SAMPLING = -13,      /* jpeg.cpp:110-136 */
  B200JPEG_E_TABLE_ID = -14,      /* jpeg.cpp:156-160, 205-209 */
  B200JPEG_E_DHT = -15,           /* jpeg.cpp:177-181, 190-194 */
  B200JPEG_E_SOS = -16,           /* jpeg.cpp:255-264, 282-286 */
  B200JPEG_E_DRI = -17,           /* jpeg.cpp:292-296 */
  B200JPEG_E_NO_SOS = -18,        /* readJPEG returned before SOS, jpeg.cpp:878-903 */
  B200JPEG_E_TABLE_MISSING = -19, /* jpeg.cpp:757-774 */
  B200JPEG_E_TRUNCATED = -20,     /* header runs past the end of the buffer */
  B200JPEG_E_COMPONENT_ID = -21,  /* jpeg.cpp:95-104 */
  B200JPEG_E_UNSUPPORTED = -22,   /* accepted by the reference's parser but outside the kernels:
                                     the subsampled component is not the first one */
  B200JPEG_E_TOO_LARGE = -23      /* b200jpeg_export_coefficients_noalloc: beyond B200JPEG_NOALLOC_MAX_BYTES */
};

/* ---- per-image status bits written by the kernels (status[i] == 0: decoded) ---- */
enum {
  B200JPEG_ST_BAD_MARKER = 1 << 0,     /* file.hpp:91-95 */
  B200JPEG_ST_NO_EOI = 1 << 1,         /* file.hpp:59-104 would never return */
  B200JPEG_ST_BAD_RST_COUNT = 1 << 2,  /* RSTn markers != ceil(MCUs/RI)-1 */
  B200JPEG_ST_BAD_CODE = 1 << 3,       /* jpeg.cpp:300-320 */
  B200JPEG_ST_BAD_DC_LEN = 1 << 4,     /* jpeg.cpp:330-334 */
  B200JPEG_ST_BAD_RUN = 1 << 5,        /* jpeg.cpp:372-376 */
  B200JPEG_ST_BAD_AC_LEN = 1 << 6,     /* jpeg.cpp:381-385 */
  B200JPEG_ST_OVERRUN = 1 << 7,        /* file.hpp:36-40 */
  B200JPEG_ST_BLOCK_COUNT = 1 << 8,
  B200JPEG_ST_DC_RANGE = 1 << 9,
  B200JPEG_ST_NOT_CONVERGED = 1 << 10,
  B200JPEG_ST_UNSUPPORTED = 1 << 11
};

/* parser flags */
enum {
  B200JPEG_PARSE_REFERENCE = 0, /* accept exactly what the reference accepts */
  B200JPEG_PARSE_SPEC = 1       /* superset: 1-component grey, COM/unknown
                                   segments skipped, fill bytes before markers, any component identifiers
                                   (the frame's first component is luma; tables bound by the scan's
                                   component selectors, which must follow the frame's order; reference:
                                   id 1 = luma, bound by position, jpeg.cpp:110, 241-247), 16-bit DQT
                                   values kept whole (reference: low byte, jpeg.cpp:216) */
};

/* HuffmanTable, include/huffman.hpp:8-15.  offset[] are cumulative counts (jpeg.cpp:170-176),
 * codes[] what generateCodes() assigns (huffman.hpp:17-29). */
typedef struct {
  uint8_t offset[17];
  uint8_t symbols[176];   /* `char symbols[162]` in the reference; jpeg.cpp:177 admits up to 176 */
  uint32_t codes[176];
  uint8_t set;
} b200jpeg_huffman_table;

/* QuantizationTable, include/types.hpp:79-96: de-zig-zagged at construction (natural order). */
typedef struct {
  int32_t table[64];
  uint8_t set;
} b200jpeg_quant_table;

/* ColorComponent, include/types.hpp:69-77 */
typedef struct {
  int32_t horizontal_sampling_factor;
  int32_t vertical_sampling_factor;
  int32_t quantizationTableID;
  int32_t huffmanDCTTableID;
  int32_t huffmanACTableID;
} b200jpeg_component;

/* MCU, include/types.hpp:32-67: one 8x8 block position; y/cb/cr before YCbCrToRGB, r/g/b after. */
typedef struct {
  int32_t y[64], cb[64], cr[64];
} b200jpeg_mcu;

/* The state Image holds when process_image_data() is entered (jpeg.cpp:21-34, 792-795). */
typedef struct {
  int32_t image_width, image_height;          /* jpeg.cpp:792-793 */
  int32_t mcuWidth, mcuHeight;                /* ceil(/8), jpeg.cpp:30-31, 77-78 */
  int32_t mcuWidthReal, mcuHeightReal;        /* jpeg.cpp:794-795, 118-125 */
  int32_t horizontalSamplingFactor;           /* jpeg.cpp:32, 126 */
  int32_t verticalSamplingFactor;             /* jpeg.cpp:33, 127 */
  int32_t numComponents;                      /* jpeg.cpp:29 */
  int32_t restartInterval;                    /* jpeg.cpp:27, 291 */
  int32_t zeroBased;                          /* jpeg.cpp:28 */
  b200jpeg_component color_components[3];     /* jpeg.cpp:23 */
  b200jpeg_quant_table quantization_tables[4];/* jpeg.cpp:22 */
  b200jpeg_huffman_table huffmanDCTTable[4];  /* jpeg.cpp:24 */
  b200jpeg_huffman_table huffmanACTable[4];   /* jpeg.cpp:25 */
  int64_t scan_offset;   /* offset of the first entropy-coded byte in the file; where
                            Marker::Marker starts readImageData(), file.hpp:214-219 */
  int64_t scan_length;   /* bytes from scan_offset to the end of the file (stuffed, markers and
                            EOI included; the kernels find FF D9 themselves) */
} b200jpeg_image_desc;

typedef struct b200jpeg_ctx b200jpeg_ctx;

/* Host-side header parse (no GPU needed).  Mirrors readJPEG()'s marker loop up to and including
 * read_sos(), returning a code where the reference calls exit(1). */
int b200jpeg_parse_header(const uint8_t *file, size_t len, int flags, b200jpeg_image_desc *desc);

/* The same for n files at once on `threads` host threads (0 = one per hardware thread, capped by the work):
 * what a batch caller runs where the reference runs its per-byte marker reader n times (file.hpp:26-52,
 * jpeg.cpp:826-867).  rcs[i] = the code b200jpeg_parse_header would return for file i (rcs may be NULL);
 * returns the number of files that failed, or B200JPEG_E_ARG. */
int b200jpeg_parse_headers(const uint8_t *const *files, const size_t *sizes, int n, int flags, int threads,
                           b200jpeg_image_desc *descs, int32_t *rcs);

/* One context per GPU: owns a stream, device arenas and pinned staging.  Use from one host
 * thread at a time. */
int b200jpeg_create(int device, b200jpeg_ctx **ctx);
void b200jpeg_destroy(b200jpeg_ctx *ctx);
const char *b200jpeg_last_error(const b200jpeg_ctx *ctx);

/* Whole path, HOST buffers in and out (the reference-facing call):
 *   file_bytes[i]/file_sizes[i]  the complete JPEG file i (pageable or pinned host memory)
 *   rgb_out[i]   host buffer of 3*W*H bytes: row-major, top-down, interleaved R,G,B, pitch 3*W
 *                (the display.hpp:27-34 gather flattened); may be NULL to skip the copy
 *   coef_out[i]  optional (array or entry may be NULL): host buffer of
 *                mcuWidthReal*mcuHeightReal b200jpeg_mcu-shaped int16 triples, i.e.
 *                int16 [pos][3][64] natural order -- the MCU array after decodeHuffman()
 *   status[i]    per-image status bits
 * Equivalent to prepare + run + fetch below. */
int b200jpeg_decode_batch(b200jpeg_ctx *ctx, const b200jpeg_image_desc *descs, int n,
                          const uint8_t *const *file_bytes, const size_t *file_sizes,
                          uint8_t *const *rgb_out, int16_t *const *coef_out, int32_t *status);

/* Several GPUs of one box: images share nothing, so a batch is cut into contiguous image ranges, one per device,
 * each decoded by its own context on its own host thread (own stream, pinned staging and arenas); no data moves
 * between devices.  A device may be named more than once (two contexts on it).  The reference's main() decodes
 * the files of one process one after another (jpeg.cpp:916-929); this is that loop over N GPUs. */
typedef struct b200jpeg_multi b200jpeg_multi;
int b200jpeg_multi_create(const int *devices, int ndev, b200jpeg_multi **multi);
void b200jpeg_multi_destroy(b200jpeg_multi *multi);
int b200jpeg_multi_count(const b200jpeg_multi *multi);
b200jpeg_ctx *b200jpeg_multi_ctx(b200jpeg_multi *multi, int k);   /* context k (for stats, entropy path, last_error) */
/* arguments as b200jpeg_decode_batch (no coefficient output); returns the first error of any device, else 0 */
int b200jpeg_multi_decode_batch(b200jpeg_multi *multi, const b200jpeg_image_desc *descs, int n,
                                const uint8_t *const *file_bytes, const size_t *file_sizes,
                                uint8_t *const *rgb_out, int32_t *status);
/* one call, no handle: contexts for `devices` are created on first use and kept for the process */
int b200jpeg_decode_batch_multi(const int *devices, int ndev, const b200jpeg_image_desc *descs, int n,
                                const uint8_t *const *file_bytes, const size_t *file_sizes,
                                uint8_t *const *rgb_out, int32_t *status);

/* Measurement aid: exactly the host<->device copies b200jpeg_decode_batch would make for this batch (same chunks,
 * same pipelines, same buffers), with no kernel launched in between -- the PCIe floor under the call.  rgb_out
 * receives whatever the device arena holds.  *ms = wall time of the call in milliseconds. */
int b200jpeg_copy_floor(b200jpeg_ctx *ctx, const b200jpeg_image_desc *descs, int n,
                        const uint8_t *const *file_bytes, const size_t *file_sizes, uint8_t *const *rgb_out, float *ms);

/* ---- progressive DCT files (SOF2): SURVEY.md 8(f) rank 4.  The reference REFUSES them (jpeg.cpp:69-73, "Progressive
 * JPEG not supported") although it ships two as fixtures (images/prograssive-sample*.jpg); b200jpeg_parse_header keeps
 * returning B200JPEG_E_PROGRESSIVE for them, and these two calls lift the refusal.  A progressive file is a frame
 * (same fields as SOF0, jpeg.cpp:50-138) followed by several scans (ITU-T T.81 Annex G), each a band of coefficients
 * (Ss..Se) at a bit position (Ah, Al) for one or all components, each with the Huffman tables in force when it starts
 * (DHT segments may redefine a table between scans, hence the pool).  After the last scan the coefficients are what a
 * baseline file of the same picture holds, and the rest of the path (dequantize, inverseDCT, YCbCrToRGB) is the same. */
#define B200JPEG_PROG_MAX_SCANS 48
#define B200JPEG_PROG_MAX_TABLES 48
typedef struct {
  int32_t ns;                          /* components in the scan: 1, or all of them for DC scans */
  int32_t comp[3];                     /* index into frame.color_components */
  int32_t dc_table[3], ac_table[3];    /* per scan component: index into tables[], -1 where the scan needs none */
  int32_t ss, se, ah, al;              /* spectral selection and successive approximation (T.81 B.2.3) */
  int32_t restart_interval;            /* DRI in force (jpeg.cpp:285-298), in MCUs of THIS scan */
  int64_t offset, length;              /* the scan's entropy-coded bytes in the file: stuffed, RSTn included, up to the next marker */
} b200jpeg_scan;
typedef struct {
  b200jpeg_image_desc frame;           /* geometry, components (sampling, quantisation ids), quantisation tables; its Huffman
                                          tables are unused, scan_offset/scan_length span from the first scan to the end of the file */
  int32_t nscans, ntables;
  b200jpeg_scan scans[B200JPEG_PROG_MAX_SCANS];
  b200jpeg_huffman_table tables[B200JPEG_PROG_MAX_TABLES];
} b200jpeg_progressive_desc;
/* host only; B200JPEG_E_UNSUPPORTED for frames the kernels do not take (only the first component may be subsampled, by 1 or
 * 2; more scans / tables than the limits above), B200JPEG_E_NOT_JPEG .. as b200jpeg_parse_header, B200JPEG_E_ARG for a
 * baseline (SOF0) file */
int b200jpeg_parse_progressive(const uint8_t *file, size_t len, b200jpeg_progressive_desc *desc);
/* arguments as b200jpeg_decode_batch.  One GPU thread walks one scan of one image (scans of an image run one after another,
 * images side by side): correct, not fast -- see DESIGN.md. */
int b200jpeg_decode_progressive(b200jpeg_ctx *ctx, const b200jpeg_progressive_desc *descs, int n,
                                const uint8_t *const *file_bytes, const size_t *file_sizes,
                                uint8_t *const *rgb_out, int16_t *const *coef_out, int32_t *status);

/* The same in three steps, so that a caller can keep a batch resident in HBM:
 *   prepare: plan the batch, copy the scans (and the small descriptors) host -> device
 *   run    : launch every kernel of the path on the context's stream (asynchronous)
 *   fetch  : wait, then copy RGB / coefficients / status device -> host */
int b200jpeg_batch_prepare(b200jpeg_ctx *ctx, const b200jpeg_image_desc *descs, int n,
                           const uint8_t *const *file_bytes, const size_t *file_sizes);
int b200jpeg_batch_run(b200jpeg_ctx *ctx);
int b200jpeg_batch_fetch(b200jpeg_ctx *ctx, uint8_t *const *rgb_out, int16_t *const *coef_out,
                         int32_t *status);
int b200jpeg_sync(b200jpeg_ctx *ctx);

/* Device-resident results of the current batch (valid until the next prepare). */
int b200jpeg_device_rgb(b200jpeg_ctx *ctx, int i, void **dev_ptr, size_t *pitch_bytes);
void *b200jpeg_stream(b200jpeg_ctx *ctx);   /* cudaStream_t */

/* Counters of the current batch (after b200jpeg_decode_batch: summed over its chunks). */
typedef struct {
  int64_t images, pixels, scan_bytes, coded_blocks, rgb_bytes;
  int64_t h2d_bytes, d2h_bytes;
  int32_t kernel_launches;      /* kernels launched by the last b200jpeg_batch_run */
  int32_t sync_rounds;          /* fix launches the last run needed */
  int32_t entropy_path;         /* decoder that took the batch: 0 self-synchronising, 1 segment-parallel */
  int32_t reserved;
} b200jpeg_batch_stats;
int b200jpeg_batch_get_stats(const b200jpeg_ctx *ctx, b200jpeg_batch_stats *stats);

/* Entropy decoder of the following batches.  AUTO picks per batch: SEGMENT (one thread per restart
 * segment, one pass) when the batch holds enough segments to fill the GPU, else SYNC (self-synchronising
 * sub-sequences inside each scan).  Both give identical results; the override exists for tests and
 * measurements. */
enum { B200JPEG_ENTROPY_AUTO = 0, B200JPEG_ENTROPY_SEGMENT = 1, B200JPEG_ENTROPY_SYNC = 2 };
int b200jpeg_set_entropy_path(b200jpeg_ctx *ctx, int mode);

/* Per-stage CUDA-event times of the last b200jpeg_batch_run (milliseconds): unstuff, entropy
 * sync, entropy write, pixel, total.  Only recorded when enabled (adds event records). */
enum { B200JPEG_T_UNSTUFF = 0, B200JPEG_T_SYNC = 1, B200JPEG_T_WRITE = 2, B200JPEG_T_PIXEL = 3,
       B200JPEG_T_TOTAL = 4, B200JPEG_T_COUNT = 5 };
int b200jpeg_set_timing(b200jpeg_ctx *ctx, int enabled);
int b200jpeg_last_timing(b200jpeg_ctx *ctx, float *ms, int count);

/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The coefficients of image i of the CURRENT batch (after b200jpeg_batch_run) in the layout b200jpeg_batch_fetch()
 * gives for coef_out[i] -- int16 [mcuWidthReal*mcuHeightReal][3][64], natural order, the MCU array after
 * decodeHuffman() -- but staged in a fixed array of B200JPEG_NOALLOC_MAX_BYTES that lives in the library's device
 * image (every GPU that loads the library holds it until exit) instead of a buffer allocated on demand: the call
 * performs NO allocation of any kind.  coef_out is a host buffer of coef_out_bytes >= positions * 384 bytes.
 * An image whose array is larger is refused with B200JPEG_E_TOO_LARGE (text in b200jpeg_last_error).
 * All calls on one GPU share the array: calls that overlap on the same GPU (other host threads, other contexts on
 * that device) race -- serialise them.  Returns 0 or a B200JPEG_E_* code. */
#define B200JPEG_NOALLOC_MAX_BYTES (16u << 20)   /* 43,690 block positions: e.g. 1920x1080 4:2:0 (12.5 MB) */
int b200jpeg_export_coefficients_noalloc(b200jpeg_ctx *ctx, int i, int16_t *coef_out, size_t coef_out_bytes);

/* Pinned host memory helpers (cudaHostAlloc / cudaFreeHost) for zero-staging copies. */
void *b200jpeg_host_alloc(size_t bytes);
void b200jpeg_host_free(void *p);

int b200jpeg_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif
