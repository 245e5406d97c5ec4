/* oracle/jpeg_oracle.h -- TEST INFRASTRUCTURE: plain-C restatement of the reference's
 * scan-to-pixels path (aswanthabam/JPEG_Decoder, /root/reference).  It exists to CHECK the
 * CUDA path; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it.  Nothing under jpeg_decoder_b200/ links or calls it.
 *
 * Parity status: PINNED.  tests/test_oracle.py compares this port with the untouched
 * reference (oracle/_ref/ref_decode, built by oracle/Makefile from the reference's own
 * sources) on the six baseline fixtures and on synthetic streams: coefficient dumps
 * byte-identical, RGB byte-identical; digests of those runs are committed in tests/golden/.
 */
#ifndef B200JPEG_ORACLE_H
#define B200JPEG_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum {
  ORACLE_OK = 0,
  ORACLE_E_NOT_JPEG = -1,        /* jpeg.cpp:800-805 */
  ORACLE_E_PROGRESSIVE = -2,     /* jpeg.cpp:69-73 */
  ORACLE_E_COMPONENTS = -3,      /* jpeg.cpp:83-87, 236-240 */
  ORACLE_E_SAMPLING = -4,        /* jpeg.cpp:110-136 */
  ORACLE_E_TABLE_ID = -5,        /* jpeg.cpp:156-160, 205-209 */
  ORACLE_E_DHT = -6,             /* jpeg.cpp:177-181, 190-194 */
  ORACLE_E_SOS = -7,             /* jpeg.cpp:255-264, 282-286 */
  ORACLE_E_DRI = -8,             /* jpeg.cpp:292-296 */
  ORACLE_E_NO_SOS = -9,          /* readJPEG returned before SOS (jpeg.cpp:878-903) */
  ORACLE_E_TABLE_MISSING = -10,  /* jpeg.cpp:757-774 */
  ORACLE_E_SCAN_MARKER = -11,    /* file.hpp:91-95 */
  ORACLE_E_NO_EOI = -12,         /* file.hpp:59-104 never terminates without FF D9 */
  ORACLE_E_DC_LENGTH = -13,      /* jpeg.cpp:330-334 */
  ORACLE_E_AC_RUN = -14,         /* jpeg.cpp:372-376 */
  ORACLE_E_AC_LENGTH = -15,      /* jpeg.cpp:381-385 */
  ORACLE_E_OVERRUN = -16,        /* file.hpp:36-40 "Reading of unavailable data" */
  ORACLE_E_TRUNCATED = -17,      /* header runs past the end of the file */
  ORACLE_E_COMPONENT_ID = -18    /* jpeg.cpp:95-104 */
};

/* ORACLE_F_SPEC: behave as ITU-T T.81 where the reference deviates or rejects (SURVEY.md
 * 8(f) rank 1): restart every RI JPEG-MCUs regardless of sampling (reference: jpeg.cpp:414,419
 * is wrong for 4:2:0 unless RI is whole MCU rows), 1-component grayscale accepted
 * (reference: jpeg.cpp:83-87 rejects), COM/unknown segments skipped, component identifiers may be any
 * byte -- the first frame component is the (possibly subsampled) luma, and the scan's selectors must name
 * the frame's components in order (reference: identifier 1 is luma, tables bound by position,
 * jpeg.cpp:95-110, 241-247), a 16-bit DQT (Pq = 1) keeps its full values (reference: low byte,
 * jpeg.cpp:216).  Without the flag the port reproduces the reference verbatim, quirks included. */
enum { ORACLE_F_SPEC = 1 };

typedef struct {
  int width, height;      /* jpeg.cpp:75-76 */
  int bw, bh;             /* mcuWidth/mcuHeight = ceil(/8), jpeg.cpp:77-78 */
  int bwr, bhr;           /* mcuWidthReal/mcuHeightReal, jpeg.cpp:118-125 */
  int hs, vs;             /* luma sampling factors, jpeg.cpp:126-127 */
  int ncomp;
  int restart_interval;   /* jpeg.cpp:291 */
  int scan_offset;        /* file offset of the first entropy-coded byte */
  int scan_unstuffed;     /* bytes in clean_data, file.hpp:102-103 */
  double ms_parse_unstuff, ms_huffman, ms_dequant, ms_idct, ms_colour;
} oracle_info;

/* Whole path.  On success *coef (if non-NULL) receives a malloc'd int32 dump laid out as the
 * reference's MCU array after decodeHuffman(): [bwr*bhr][3][64], natural order
 * (types.hpp:32-50, jpeg.cpp:398), and *rgb (if non-NULL) a malloc'd uint8 [H][W][3] gathered
 * as display.hpp:19-34.  Free both with oracle_free(). */
int oracle_decode(const uint8_t *file, size_t len, int flags, oracle_info *info,
                  int32_t **coef, uint8_t **rgb);
/* Progressive DCT (SOF2), T.81 Annex G.  PORT ONLY: the reference rejects SOF2 (jpeg.cpp:69-73); pinned through baseline
 * twins of the same coefficients that the reference binary does decode (tests/test_oracle.py).  Same outputs as
 * oracle_decode; info->scan_unstuffed holds the number of scans. */
int oracle_decode_progressive(const uint8_t *file, size_t len, int flags, oracle_info *info,
                              int32_t **coef, uint8_t **rgb);
void oracle_free(void *p);

/* Stage-level entry points for unit tests. */
void oracle_idct_block(int32_t *block64);                  /* jpeg.cpp:594-732 */
void oracle_dequant_block(int32_t *block64, const int32_t *q64); /* jpeg.cpp:563-569 */
/* Y,Cb,Cr -> R,G,B for one pixel, jpeg.cpp:521-538 */
void oracle_ycc_to_rgb(int32_t y, int32_t cb, int32_t cr, int32_t *r, int32_t *g, int32_t *b);
/* The 14 IDCT constants as float bit patterns: m0,m1,m2,m3,m4,m5,s0..s7 (types.hpp:5-19). */
void oracle_idct_constants(uint32_t out14[14]);
/* Byte unstuffing / RST stripping exactly as file.hpp:59-104; returns bytes written or <0. */
long oracle_unstuff(const uint8_t *scan, size_t len, uint8_t *out, size_t *consumed);
extern const int oracle_zigzag[64];                        /* types.hpp:23-31 */

#ifdef __cplusplus
}
#endif
#endif
