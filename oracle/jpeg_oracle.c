/* oracle/jpeg_oracle.c -- TEST INFRASTRUCTURE (see jpeg_oracle.h).
 *
 * Plain-C restatement of the reference's scan-to-pixels path, function by function, with the
 * reference file:line each piece follows.  It is deliberately as naive as the reference
 * (bit-serial reader, linear code search): it is a checker, not a decoder to ship.
 * Build: oracle/Makefile (gcc -O2 -ffp-contract=off; no -march=native, no -ffast-math).
 */
#include "jpeg_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* types.hpp:23-31 */
const int oracle_zigzag[64] = {
    0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
    41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
    30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

/* ---- IDCT constants, computed the way types.hpp:5-19 computes them (double -> float) ---- */
static float M0, M1, M2, M3, M4, M5, S0, S1, S2, S3, S4, S5, S6, S7;
static int consts_ready = 0;
static void init_consts(void) {
  if (consts_ready) return;
  M0 = (float)(2.0 * cos(1.0 / 16.0 * 2.0 * M_PI));
  M1 = (float)(2.0 * cos(2.0 / 16.0 * 2.0 * M_PI));
  M3 = (float)(2.0 * cos(2.0 / 16.0 * 2.0 * M_PI));
  M5 = (float)(2.0 * cos(3.0 / 16.0 * 2.0 * M_PI));
  M2 = M0 - M5; /* float - float, types.hpp:9 */
  M4 = M0 + M5; /* types.hpp:10 */
  S0 = (float)(cos(0.0 / 16.0 * M_PI) / sqrt(8));
  S1 = (float)(cos(1.0 / 16.0 * M_PI) / 2.0);
  S2 = (float)(cos(2.0 / 16.0 * M_PI) / 2.0);
  S3 = (float)(cos(3.0 / 16.0 * M_PI) / 2.0);
  S4 = (float)(cos(4.0 / 16.0 * M_PI) / 2.0);
  S5 = (float)(cos(5.0 / 16.0 * M_PI) / 2.0);
  S6 = (float)(cos(6.0 / 16.0 * M_PI) / 2.0);
  S7 = (float)(cos(7.0 / 16.0 * M_PI) / 2.0);
  consts_ready = 1;
}
void oracle_idct_constants(uint32_t out[14]) {
  init_consts();
  const float v[14] = {M0, M1, M2, M3, M4, M5, S0, S1, S2, S3, S4, S5, S6, S7};
  memcpy(out, v, sizeof v);
}

/* ---- table objects: huffman.hpp:8-29, types.hpp:69-96 ---- */
typedef struct {
  unsigned char offset[17];
  unsigned char symbols[176];
  unsigned int codes[176];
  int set;
} htab_t;
typedef struct { int table[64]; int set; } qtab_t;
typedef struct { int h, v, tq, td, ta; } comp_t;

static void generate_codes(htab_t *t) { /* huffman.hpp:17-29 */
  unsigned int code = 0;
  for (int i = 0; i < 16; i++) {
    for (int j = t->offset[i]; j < t->offset[i + 1]; j++) t->codes[j] = code++;
    code <<= 1;
  }
}

typedef struct {
  const uint8_t *d; size_t n, pos;
  int flags;
  int width, height, bw, bh, bwr, bhr, hs, vs, ncomp, ri, zero_based;
  comp_t comp[3];
  int comp_id[3];             /* SOF component identifiers (only ORACLE_F_SPEC binds tables by them) */
  qtab_t q[4];
  htab_t dc[4], ac[4];
  /* BitStream state, file.hpp:122-165 */
  const uint8_t *clean; size_t clean_n, clean_pos;
  unsigned char st; int st_c;
  int err;
} dec_t;

static int be16(const uint8_t *p) { return (p[0] << 8) | p[1]; }

/* ---- segment parsers: jpeg.cpp:67-298 ---- */
static int read_sof(dec_t *D, int marker, const uint8_t *p, int len) {
  if (marker == 0xc2) return ORACLE_E_PROGRESSIVE;           /* jpeg.cpp:69-73 */
  if (len < 6) return ORACLE_E_TRUNCATED;
  D->height = be16(p + 1); D->width = be16(p + 3);           /* :75-76 */
  D->bw = (D->width + 7) / 8; D->bh = (D->height + 7) / 8;   /* :77-78 */
  D->bwr = D->bw; D->bhr = D->bh;
  const int nc = p[5]; D->ncomp = nc;
  const int spec = D->flags & ORACLE_F_SPEC;
  if (nc != 3 && !(spec && nc == 1)) return ORACLE_E_COMPONENTS; /* :83-87 */
  if (len < 6 + 3 * nc) return ORACLE_E_TRUNCATED;
  for (int i = 0; i < nc; i++) {
    const int id = p[6 + 3 * i];
    D->comp_id[i] = id;
    if (id == 0) D->zero_based = 1;                          /* :91-94 */
    if (!spec && !D->zero_based && id > 3) return ORACLE_E_COMPONENT_ID;
    if (!spec && D->zero_based && id > 2) return ORACLE_E_COMPONENT_ID;
    comp_t *c = &D->comp[i];
    c->h = p[7 + 3 * i] >> 4; c->v = p[7 + 3 * i] & 15; c->tq = p[8 + 3 * i]; /* :105-109 */
    if (nc == 1) { c->h = c->v = 1; D->hs = D->vs = 1; continue; } /* T.81 A.2.2: one block per MCU */
    /* the reference takes "identifier 1" for the luma component (:110); T.81 B.2.2 lets identifiers be any
     * byte, so under ORACLE_F_SPEC it is the FIRST component of the frame */
    if (spec ? i == 0 : id == 1) {                           /* :110-128 */
      if ((c->h != 1 && c->h != 2) || (c->v != 1 && c->v != 2)) return ORACLE_E_SAMPLING;
      if (c->h == 2 && D->bw % 2 == 1) D->bwr += 1;
      if (c->v == 2 && D->bh % 2 == 1) D->bhr += 1;
      D->hs = c->h; D->vs = c->v;
    } else if (c->h != 1 || c->v != 1) return ORACLE_E_SAMPLING; /* :129-136 */
  }
  return 0;
}

static int read_dht(dec_t *D, const uint8_t *p, int len) {   /* jpeg.cpp:148-196 */
  while (len > 0) {
    if (len < 17) return ORACLE_E_DHT;
    const int id = p[0] & 15, is_ac = p[0] >> 4;
    if (id > 3) return ORACLE_E_TABLE_ID;
    htab_t *t = is_ac ? &D->ac[id] : &D->dc[id];
    t->offset[0] = 0;
    int all = 0;
    for (int i = 1; i <= 16; i++) { all += p[i]; t->offset[i] = (unsigned char)all; }
    if (all > 176) return ORACLE_E_DHT;                      /* :177-181 */
    if (len < 17 + all) return ORACLE_E_DHT;
    for (int i = 0; i < all; i++) t->symbols[i] = p[17 + i];
    t->set = 1; generate_codes(t);
    p += 17 + all; len -= 17 + all;
  }
  return len != 0 ? ORACLE_E_DHT : 0;                        /* :190-194 */
}

static int read_dqt(dec_t *D, const uint8_t *p, int len) {   /* jpeg.cpp:197-231 */
  while (len > 0) {
    const int id = p[0] & 15, wide = p[0] >> 4;
    p++; len--;
    if (id > 3) return ORACLE_E_TABLE_ID;
    unsigned char raw[64];
    if (wide) {
      if (len < 128) return ORACLE_E_TRUNCATED;
      for (int i = 0; i < 64; i++) raw[i] = (unsigned char)((p[2 * i] << 8) + p[2 * i + 1]); /* :216 keeps the low byte */
      if (D->flags & ORACLE_F_SPEC) {                         /* T.81 B.2.4.1, Pq = 1: the full 16-bit quantiser */
        for (int i = 0; i < 64; i++) D->q[id].table[oracle_zigzag[i]] = (p[2 * i] << 8) + p[2 * i + 1];
        D->q[id].set = 1; p += 128; len -= 128;
        continue;
      }
      p += 128; len -= 128;
    } else {
      if (len < 64) return ORACLE_E_TRUNCATED;
      memcpy(raw, p, 64); p += 64; len -= 64;
    }
    for (int i = 0; i < 64; i++) D->q[id].table[oracle_zigzag[i]] = raw[i]; /* types.hpp:86-92 */
    D->q[id].set = 1;
  }
  return 0;
}

static int read_sos(dec_t *D, const uint8_t *p, int len) {   /* jpeg.cpp:233-287 */
  if (len < 1) return ORACLE_E_SOS;
  const int nc = p[0];
  if (nc != D->ncomp || (nc != 3 && nc != 1)) return ORACLE_E_COMPONENTS; /* :236-240 */
  if (len < 1 + 2 * nc + 3) return ORACLE_E_SOS;
  for (int i = 0; i < nc; i++) {                              /* tables bound by position, :241-247 */
    /* ORACLE_F_SPEC: by component selector (T.81 B.2.3); the selectors must name the frame's components in
     * the frame's order, anything else is not a scan the interleaved decode loop (:415-435) can take */
    if ((D->flags & ORACLE_F_SPEC) && p[1 + 2 * i] != D->comp_id[i]) return ORACLE_E_SOS;
    D->comp[i].td = p[2 + 2 * i] >> 4; D->comp[i].ta = p[2 + 2 * i] & 15;
  }
  const uint8_t *q = p + 1 + 2 * nc;
  if (q[0] != 0 || q[1] != 63) return ORACLE_E_SOS;           /* :255-259 */
  if (q[2] != 0) return ORACLE_E_SOS;                         /* :260-264 */
  if (len + 2 - 6 - 2 * nc != 0) return ORACLE_E_SOS;         /* :282-286 (marker->length counts itself) */
  return 0;
}

/* ---- byte unstuffing, file.hpp:59-104 ---- */
long oracle_unstuff(const uint8_t *s, size_t n, uint8_t *out, size_t *consumed) {
  size_t i = 0, o = 0;
  if (n < 2) return ORACLE_E_NO_EOI;
  unsigned char current = s[i++], last;
  for (;;) {
    last = current;
    if (i >= n) return ORACLE_E_NO_EOI;
    current = s[i++];
    if (last == 0xff) {
      if (current == 0xd9) break;                             /* :74-77 */
      else if (current == 0x00) {                             /* :78-82 */
        out[o++] = last;
        if (i >= n) return ORACLE_E_NO_EOI;
        current = s[i++];
      } else if (current >= 0xd0 && current <= 0xd7) {        /* :83-86, index never checked */
        if (i >= n) return ORACLE_E_NO_EOI;
        current = s[i++];
      } else if (current == 0xff) continue;                   /* :87-90 */
      else return ORACLE_E_SCAN_MARKER;                       /* :91-95 */
    } else out[o++] = last;
  }
  if (consumed) *consumed = i;
  return (long)o;
}

/* ---- bit reader, file.hpp:130-164 ---- */
static int get_bit(dec_t *D) {
  if (D->st_c < 0) {
    if (D->clean_pos >= D->clean_n) { D->err = ORACLE_E_OVERRUN; return 0; } /* file.hpp:36-40 */
    D->st = D->clean[D->clean_pos++];
    D->st_c = 7;
  }
  const int bit = (D->st >> D->st_c) & 1;
  D->st_c--;
  return bit;
}
static int get_bit_n(dec_t *D, int n) {
  int v = 0;
  for (int i = 0; i < n; i++) v = (v << 1) | get_bit(D);
  return v;
}

/* jpeg.cpp:300-320; the -1 sentinel leaves as unsigned char 255 */
static int next_symbol(dec_t *D, const htab_t *t) {
  unsigned int cur = 0;
  for (int i = 0; i < 16; i++) {
    cur = (cur << 1) | (unsigned)get_bit(D);
    if (D->err) return 255;
    for (int j = t->offset[i]; j < t->offset[i + 1]; j++)
      if (t->codes[j] == cur) return t->symbols[j];
  }
  return 255;
}

/* jpeg.cpp:322-403 */
static int decode_block(dec_t *D, int32_t *c, int *prev_dc, const htab_t *dct, const htab_t *act) {
  const int length = next_symbol(D, dct);
  if (D->err) return D->err;
  if (length > 11) return ORACLE_E_DC_LENGTH;                 /* :330-334 */
  int coeff = get_bit_n(D, length);
  if (length != 0 && coeff < (1 << (length - 1))) coeff -= (1 << length) - 1; /* :340-343 */
  c[0] = coeff + *prev_dc; *prev_dc = c[0];                    /* :344-345 */
  int i = 1;
  while (i < 64) {
    const int symbol = next_symbol(D, act);
    if (D->err) return D->err;
    if (symbol == 0x00) { for (; i < 64; i++) c[oracle_zigzag[i]] = 0; return 0; } /* :356-363 */
    int zeros = symbol >> 4;
    const int clen = symbol & 15;
    if (symbol == 0xf0) zeros = 16;                           /* :368-371 */
    if (i + zeros >= 64) return ORACLE_E_AC_RUN;              /* :372-376 */
    for (int j = 0; j < zeros; j++, i++) c[oracle_zigzag[i]] = 0;
    if (clen > 10) return ORACLE_E_AC_LENGTH;                 /* :381-385 */
    if (clen != 0) {
      coeff = get_bit_n(D, clen);
      if (coeff < (1 << (clen - 1))) coeff -= (1 << clen) - 1; /* :394-397 */
      c[oracle_zigzag[i]] = coeff;                            /* :398 */
      i++;
    }
    if (D->err) return D->err;
  }
  return 0;
}

#define POS(D, y, x) ((size_t)(y) * (D)->bwr + (x))
#define PLANE(m, pos, i) ((m) + ((pos) * 3 + (i)) * 64)

/* jpeg.cpp:405-446 */
static int decode_huffman(dec_t *D, int32_t *mcus) {
  int prev[3] = {0, 0, 0};
  const int ri_ref = D->ri * D->hs * D->vs;                   /* :414 */
  long mcu_index = 0;
  for (int y = 0; y < D->bh; y += D->vs)
    for (int x = 0; x < D->bw; x += D->hs, ++mcu_index) {
      int restart;
      if (D->flags & ORACLE_F_SPEC) restart = D->ri != 0 && mcu_index % D->ri == 0; /* T.81 E.2.4 */
      else restart = ri_ref != 0 && (y * D->bwr + x) % ri_ref == 0;                /* :419 */
      if (restart) { prev[0] = prev[1] = prev[2] = 0; D->st_c = -1; }               /* :421-424 */
      for (int i = 0; i < D->ncomp; ++i)
        for (int v = 0; v < D->comp[i].v; ++v)
          for (int h = 0; h < D->comp[i].h; ++h) {
            const int e = decode_block(D, PLANE(mcus, POS(D, y + v, x + h), i), &prev[i],
                                       &D->dc[D->comp[i].td], &D->ac[D->comp[i].ta]);
            if (e) return e;
          }
    }
  return 0;
}

void oracle_dequant_block(int32_t *c, const int32_t *q) {     /* jpeg.cpp:563-569 */
  for (int i = 0; i < 64; ++i) c[i] *= q[i];
}

/* One AAN 1-D pass, jpeg.cpp:598-662 (columns) == :666-730 (rows); results truncate to int. */
static void idct_1d(int32_t *c, int i0, int stride) {
  const float g0 = c[i0 + 0 * stride] * S0;
  const float g1 = c[i0 + 4 * stride] * S4;
  const float g2 = c[i0 + 2 * stride] * S2;
  const float g3 = c[i0 + 6 * stride] * S6;
  const float g4 = c[i0 + 5 * stride] * S5;
  const float g5 = c[i0 + 1 * stride] * S1;
  const float g6 = c[i0 + 7 * stride] * S7;
  const float g7 = c[i0 + 3 * stride] * S3;
  const float f4 = g4 - g7, f5 = g5 + g6, f6 = g5 - g6, f7 = g4 + g7;
  const float e2 = g2 - g3, e3 = g2 + g3, e5 = f5 - f7, e7 = f5 + f7, e8 = f4 + f6;
  const float d2 = e2 * M1, d4 = f4 * M2, d5 = e5 * M3, d6 = f6 * M4, d8 = e8 * M5;
  const float c0 = g0 + g1, c1 = g0 - g1, c2 = d2 - e3, c4 = d4 + d8, c5 = d5 + e7, c6 = d6 - d8;
  const float c8 = c5 - c6;
  const float b0 = c0 + e3, b1 = c1 + c2, b2 = c1 - c2, b3 = c0 - e3, b4 = c4 - c8, b6 = c6 - e7;
  c[i0 + 0 * stride] = (int32_t)(b0 + e7);
  c[i0 + 1 * stride] = (int32_t)(b1 + b6);
  c[i0 + 2 * stride] = (int32_t)(b2 + c8);
  c[i0 + 3 * stride] = (int32_t)(b3 + b4);
  c[i0 + 4 * stride] = (int32_t)(b3 - b4);
  c[i0 + 5 * stride] = (int32_t)(b2 - c8);
  c[i0 + 6 * stride] = (int32_t)(b1 - b6);
  c[i0 + 7 * stride] = (int32_t)(b0 - e7);
}
void oracle_idct_block(int32_t *c) {                          /* jpeg.cpp:594-732 */
  init_consts();
  for (int i = 0; i < 8; ++i) idct_1d(c, i, 8);               /* columns, :596-663 */
  for (int i = 0; i < 8; ++i) idct_1d(c, i * 8, 1);           /* rows, :664-731 */
}

void oracle_ycc_to_rgb(int32_t y, int32_t cb, int32_t cr, int32_t *r, int32_t *g, int32_t *b) {
  int rr = y + 1.402f * cr + 128;                             /* jpeg.cpp:521 */
  int gg = y - 0.344f * cb - 0.714f * cr + 128;               /* :522 */
  int bb = y + 1.772f * cb + 128;                             /* :523 */
  if (rr < 0) rr = 0; if (rr > 255) rr = 255;                 /* :524-535 */
  if (gg < 0) gg = 0; if (gg > 255) gg = 255;
  if (bb < 0) bb = 0; if (bb > 255) bb = 255;
  *r = rr; *g = gg; *b = bb;
}

/* jpeg.cpp:511-541; in place through the unions of types.hpp:32-50 */
static void ycc_block(dec_t *D, int32_t *mcus, size_t pos, size_t cpos, unsigned v, unsigned h) {
  int32_t *Y = PLANE(mcus, pos, 0), *G = PLANE(mcus, pos, 1), *B = PLANE(mcus, pos, 2);
  const int32_t *CB = PLANE(mcus, cpos, 1), *CR = PLANE(mcus, cpos, 2);
  for (unsigned y = 7; y < 8; --y)
    for (unsigned x = 7; x < 8; --x) {
      const unsigned p = y * 8 + x;
      const unsigned cp = (y / D->vs + 4 * v) * 8 + (x / D->hs + 4 * h);  /* :518-520 */
      int32_t r, g, b;
      if (D->ncomp == 1) oracle_ycc_to_rgb(Y[p], 0, 0, &r, &g, &b);
      else oracle_ycc_to_rgb(Y[p], CB[cp], CR[cp], &r, &g, &b);
      Y[p] = r; G[p] = g; B[p] = b;
    }
}

static double now_ms(void) {
  struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

/* dequantize + inverseDCT + YCbCrToRGB + the RGB gather over a decoded MCU array (jpeg.cpp:785-788 after decodeHuffman,
 * display.hpp:19-34); shared by the baseline and the progressive decoder.  tt = times after dequant / idct / colour. */
static void finish_pixels(dec_t *D, int32_t *mcus, int32_t **coef_out, uint8_t **rgb_out, double tt[3]) {
  const size_t npos = (size_t)D->bwr * D->bhr;
  if (coef_out) {
    *coef_out = (int32_t *)malloc(npos * 3 * 64 * sizeof(int32_t));
    memcpy(*coef_out, mcus, npos * 3 * 64 * sizeof(int32_t));
  }
  /* dequantize(), jpeg.cpp:572-590 -- same loop nest as decodeHuffman */
  for (int y = 0; y < D->bh; y += D->vs)
    for (int x = 0; x < D->bw; x += D->hs)
      for (int i = 0; i < D->ncomp; ++i)
        for (int v = 0; v < D->comp[i].v; ++v)
          for (int h = 0; h < D->comp[i].h; ++h)
            oracle_dequant_block(PLANE(mcus, POS(D, y + v, x + h), i), D->q[D->comp[i].tq].table);
  tt[0] = now_ms();
  /* inverseDCT(), jpeg.cpp:735-753 */
  for (int y = 0; y < D->bh; y += D->vs)
    for (int x = 0; x < D->bw; x += D->hs)
      for (int i = 0; i < D->ncomp; ++i)
        for (int v = 0; v < D->comp[i].v; ++v)
          for (int h = 0; h < D->comp[i].h; ++h)
            oracle_idct_block(PLANE(mcus, POS(D, y + v, x + h), i));
  tt[1] = now_ms();
  /* YCbCrToRGB(), jpeg.cpp:544-561: v,h descending so the anchor block converts last */
  for (int y = 0; y < D->bh; y += D->vs)
    for (int x = 0; x < D->bw; x += D->hs) {
      const size_t cpos = POS(D, y, x);
      for (unsigned v = D->vs - 1; v < (unsigned)D->vs; --v)
        for (unsigned h = D->hs - 1; h < (unsigned)D->hs; --h)
          ycc_block(D, mcus, POS(D, y + v, x + h), cpos, v, h);
    }
  tt[2] = now_ms();
  if (rgb_out) {                                                       /* display.hpp:19-34 */
    uint8_t *rgb = (uint8_t *)malloc((size_t)D->width * D->height * 3);
    for (int y = 0; y < D->height; ++y)
      for (int x = 0; x < D->width; ++x) {
        const size_t pos = POS(D, y / 8, x / 8); const int p = (y % 8) * 8 + (x % 8);
        uint8_t *o = rgb + ((size_t)y * D->width + x) * 3;
        o[0] = (uint8_t)PLANE(mcus, pos, 0)[p]; o[1] = (uint8_t)PLANE(mcus, pos, 1)[p];
        o[2] = (uint8_t)PLANE(mcus, pos, 2)[p];
      }
    *rgb_out = rgb;
  }
}

int oracle_decode(const uint8_t *file, size_t len, int flags, oracle_info *info, int32_t **coef_out,
                  uint8_t **rgb_out) {
  init_consts();
  dec_t *D = (dec_t *)calloc(1, sizeof(dec_t));
  D->d = file; D->n = len; D->flags = flags; D->st_c = -1;
  const int spec = flags & ORACLE_F_SPEC;
  int rc = 0, got_sos = 0;
  uint8_t *clean = NULL; int32_t *mcus = NULL;
  double t0 = now_ms();
  if (coef_out) *coef_out = NULL;
  if (rgb_out) *rgb_out = NULL;
  if (len < 2 || file[0] != 0xff || file[1] != 0xd8) { rc = ORACLE_E_NOT_JPEG; goto done; } /* jpeg.cpp:800-805 */
  D->pos = 2;
  /* Image::readJPEG, jpeg.cpp:826-907, over Marker::Marker, file.hpp:192-258 */
  while (!got_sos) {
    if (D->pos >= len) { rc = ORACLE_E_NO_SOS; goto done; }
    if (file[D->pos++] != 0xff) { rc = ORACLE_E_NO_SOS; goto done; }   /* :899-903 */
    if (D->pos >= len) { rc = ORACLE_E_TRUNCATED; goto done; }
    const int m = file[D->pos++];
    if (m == 0x00) continue;                                           /* PAD, file.hpp:196-200 */
    if (m >= 0xd0 && m < 0xd8) continue;                               /* DRM, file.hpp:208-213 */
    if (spec && m == 0xff) { D->pos--; continue; }                     /* fill bytes, T.81 B.1.1.2 */
    if (D->pos + 2 > len) { rc = ORACLE_E_TRUNCATED; goto done; }
    const int seglen = be16(file + D->pos);
    const uint8_t *p = file + D->pos + 2;
    const int plen = seglen - 2;
    if (plen < 0 || D->pos + (size_t)seglen > len) { rc = ORACLE_E_TRUNCATED; goto done; }
    D->pos += seglen;
    if (m >= 0xe0 && m < 0xf0) continue;                               /* META, file.hpp:201-207 */
    switch (m) {
      case 0xc0: case 0xc2: rc = read_sof(D, m, p, plen); break;
      case 0xc4: rc = read_dht(D, p, plen); break;
      case 0xdb: rc = read_dqt(D, p, plen); break;
      case 0xdd:                                                       /* jpeg.cpp:289-298 */
        if (plen != 2) rc = ORACLE_E_DRI; else D->ri = be16(p);
        break;
      case 0xda: rc = read_sos(D, p, plen); got_sos = 1; break;
      case 0xd8: break;                                                /* SOI: logged, jpeg.cpp:838-841 */
      default:
        if (spec && m != 0xd9 && !(m >= 0xc1 && m <= 0xcf)) break;     /* COM etc. skipped */
        if (spec && m != 0xd9) { rc = ORACLE_E_PROGRESSIVE; break; }   /* other SOFn */
        rc = ORACLE_E_NO_SOS;                                          /* EOI / INVALID end parsing, :878-887 */
    }
    if (rc) goto done;
  }
  if (D->ncomp == 0 || D->hs == 0 || D->vs == 0) { rc = ORACLE_E_SAMPLING; goto done; }
  /* jpeg.cpp:757-774 */
  for (int i = 0; i < D->ncomp; i++) {
    if (D->comp[i].tq > 3 || D->comp[i].td > 3 || D->comp[i].ta > 3) { rc = ORACLE_E_TABLE_ID; goto done; }
    if (!D->q[D->comp[i].tq].set || !D->dc[D->comp[i].td].set || !D->ac[D->comp[i].ta].set) {
      rc = ORACLE_E_TABLE_MISSING; goto done;
    }
  }
  /* Marker ctor on SOS -> readImageData, file.hpp:214-219 */
  const size_t scan_off = D->pos;
  clean = (uint8_t *)malloc(len - scan_off + 1);
  {
    const long n = oracle_unstuff(file + scan_off, len - scan_off, clean, NULL);
    if (n < 0) { rc = (int)n; goto done; }
    D->clean = clean; D->clean_n = (size_t)n; D->clean_pos = 0;
  }
  const double t1 = now_ms();
  const size_t npos = (size_t)D->bwr * D->bhr;
  mcus = (int32_t *)calloc(npos * 3 * 64, sizeof(int32_t));            /* jpeg.cpp:407: value-initialised */
  rc = decode_huffman(D, mcus);
  if (rc) goto done;
  const double t2 = now_ms();
  double tt[3];
  finish_pixels(D, mcus, coef_out, rgb_out, tt);
  const double t3 = tt[0], t4 = tt[1], t5 = tt[2];
  if (info) {
    info->ms_parse_unstuff = t1 - t0; info->ms_huffman = t2 - t1; info->ms_dequant = t3 - t2;
    info->ms_idct = t4 - t3; info->ms_colour = t5 - t4;
    info->scan_offset = (int)scan_off; info->scan_unstuffed = (int)D->clean_n;
  }
done:
  if (info) {
    info->width = D->width; info->height = D->height; info->bw = D->bw; info->bh = D->bh;
    info->bwr = D->bwr; info->bhr = D->bhr; info->hs = D->hs; info->vs = D->vs;
    info->ncomp = D->ncomp; info->restart_interval = D->ri;
  }
  free(clean); free(mcus); free(D);
  return rc;
}

/* ======================================================================================================================
 * Progressive DCT (SOF2) -- ITU-T T.81 Annex G restatement.  PORT ONLY: the reference rejects SOF2 (jpeg.cpp:69-73), so
 * there is no reference run of these functions.  What pins them instead (tests/test_oracle.py): a progressive file and
 * its baseline twin written by the same encoder from the same pixels carry the SAME quantised coefficients, and the twin
 * IS decoded by the reference binary -- coefficient dump and RGB must be identical.  Everything after the coefficients
 * (dequantize, inverseDCT, YCbCrToRGB) is the reference's own path above (finish_pixels).
 * Supported frames are the ones the rest of the port supports: 1 component, or 3 with only the first one subsampled.
 * ==================================================================================================================== */
typedef struct { int ns, ci[3], ss, se, ah, al; } pscan_t;

/* entropy-coded bytes of ONE scan: FF00 -> FF, RSTn dropped (file.hpp:78-86), stops in front of any other marker */
static long prog_unstuff(const uint8_t *s, size_t n, uint8_t *out, size_t *consumed) {
  size_t i = 0, o = 0;
  while (i < n) {
    const unsigned char c = s[i];
    if (c != 0xff) { out[o++] = c; i++; continue; }
    if (i + 1 >= n) break;
    const unsigned char m = s[i + 1];
    if (m == 0x00) { out[o++] = 0xff; i += 2; }
    else if (m >= 0xd0 && m <= 0xd7) i += 2;
    else if (m == 0xff) i += 1;
    else break;
  }
  *consumed = i;
  return (long)o;
}

static int prog_receive_extend(dec_t *D, int s) {            /* T.81 F.2.2.1 EXTEND, as jpeg.cpp:335-343 */
  int v = get_bit_n(D, s);
  if (s != 0 && v < (1 << (s - 1))) v -= (1 << s) - 1;
  return v;
}

/* G.1.2.1 / G.1.2.2 / G.1.2.3: one 8x8 block of one scan.  c = natural-order coefficients of the block. */
static int prog_block(dec_t *D, const pscan_t *S, int32_t *c, int *pred, const htab_t *dct, const htab_t *act, int *eobrun) {
  if (S->ss == 0) {                                           /* DC scan */
    if (S->ah == 0) {
      const int s = next_symbol(D, dct);
      if (D->err) return D->err;
      if (s > 11) return ORACLE_E_DC_LENGTH;
      *pred += prog_receive_extend(D, s);
      c[0] = *pred * (1 << S->al);
    } else if (get_bit(D)) c[0] |= 1 << S->al;
    return D->err;
  }
  if (S->ah == 0) {                                           /* AC first scan, G.1.2.2 */
    if (*eobrun > 0) { --*eobrun; return 0; }
    for (int k = S->ss; k <= S->se; ++k) {
      const int rs = next_symbol(D, act);
      if (D->err) return D->err;
      const int r = rs >> 4, sz = rs & 15;
      if (sz) {
        k += r;
        if (k > 63) return ORACLE_E_AC_RUN;
        c[oracle_zigzag[k]] = prog_receive_extend(D, sz) * (1 << S->al);
      } else if (r == 15) k += 15;
      else { *eobrun = 1 << r; if (r) *eobrun += get_bit_n(D, r); --*eobrun; break; }
    }
    return D->err;
  }
  /* AC refinement scan, G.1.2.3 */
  const int p1 = 1 << S->al, m1 = -(1 << S->al);
  int k = S->ss;
  if (*eobrun == 0) {
    for (; k <= S->se; ++k) {
      const int rs = next_symbol(D, act);
      if (D->err) return D->err;
      int r = rs >> 4, sz = rs & 15;
      if (sz) sz = get_bit(D) ? p1 : m1;                      /* a newly non-zero coefficient: its sign */
      else if (r != 15) { *eobrun = 1 << r; if (r) *eobrun += get_bit_n(D, r); break; }
      do {                                                    /* skip r still-zero coefficients; correction bits for the others */
        int32_t *q = &c[oracle_zigzag[k]];
        if (*q != 0) { if (get_bit(D) && (*q & p1) == 0) *q += *q >= 0 ? p1 : m1; }
        else if (--r < 0) break;
        ++k;
      } while (k <= S->se);
      if (sz && k <= 63) c[oracle_zigzag[k]] = sz;
    }
  }
  if (*eobrun > 0) {                                          /* rest of the band: correction bits only */
    for (; k <= S->se; ++k) {
      int32_t *q = &c[oracle_zigzag[k]];
      if (*q != 0 && get_bit(D) && (*q & p1) == 0) *q += *q >= 0 ? p1 : m1;
    }
    --*eobrun;
  }
  return D->err;
}

static int prog_scan(dec_t *D, const pscan_t *S, int32_t *mcus) {
  int pred[3] = {0, 0, 0}, eobrun = 0;
  long unit = 0;                                              /* MCUs decoded in this scan, for the restart interval (E.2.4) */
#define PROG_RESTART() do { if (D->ri && unit && unit % D->ri == 0) { pred[0] = pred[1] = pred[2] = 0; eobrun = 0; D->st_c = -1; } } while (0)
  if (S->ns > 1) {                                            /* interleaved (DC scans): MCU order as jpeg.cpp:415-435 */
    for (int y = 0; y < D->bh; y += D->vs)
      for (int x = 0; x < D->bw; x += D->hs, ++unit) {
        PROG_RESTART();
        for (int j = 0; j < S->ns; ++j) {
          const int i = S->ci[j];
          for (int v = 0; v < D->comp[i].v; ++v)
            for (int h = 0; h < D->comp[i].h; ++h) {
              const int e = prog_block(D, S, PLANE(mcus, POS(D, y + v, x + h), i), &pred[i], &D->dc[D->comp[i].td],
                                       &D->ac[D->comp[i].ta], &eobrun);
              if (e) return e;
            }
        }
      }
    return 0;
  }
  /* one component: its blocks in raster order over ceil(component size / 8) (A.2.3); a chroma block lives at the anchor
   * position of its MCU in the reference's layout (jpeg.cpp:426-435) */
  const int i = S->ci[0];
  const int lum = D->comp[i].h == D->hs && D->comp[i].v == D->vs && (i == 0);
  const int cw = lum ? D->width : (D->width + D->hs - 1) / D->hs, ch = lum ? D->height : (D->height + D->vs - 1) / D->vs;
  const int nbx = (cw + 7) / 8, nby = (ch + 7) / 8;
  for (int by = 0; by < nby; ++by)
    for (int bx = 0; bx < nbx; ++bx, ++unit) {
      PROG_RESTART();
      const size_t pos = lum ? POS(D, by, bx) : POS(D, by * D->vs, bx * D->hs);
      const int e = prog_block(D, S, PLANE(mcus, pos, i), &pred[i], &D->dc[D->comp[i].td], &D->ac[D->comp[i].ta], &eobrun);
      if (e) return e;
    }
  return 0;
#undef PROG_RESTART
}

int oracle_decode_progressive(const uint8_t *file, size_t len, int flags, oracle_info *info, int32_t **coef_out,
                              uint8_t **rgb_out) {
  init_consts();
  dec_t *D = (dec_t *)calloc(1, sizeof(dec_t));
  D->d = file; D->n = len; D->flags = flags | ORACLE_F_SPEC; D->st_c = -1;
  int rc = 0, got_frame = 0, nscans = 0;
  uint8_t *clean = NULL; int32_t *mcus = NULL;
  if (coef_out) *coef_out = NULL;
  if (rgb_out) *rgb_out = NULL;
  if (len < 2 || file[0] != 0xff || file[1] != 0xd8) { rc = ORACLE_E_NOT_JPEG; goto done; }
  size_t pos = 2;
  clean = (uint8_t *)malloc(len + 1);
  for (;;) {
    if (pos + 1 >= len) break;
    if (file[pos] != 0xff) { rc = ORACLE_E_NO_SOS; goto done; }
    const int m = file[pos + 1];
    if (m == 0xff) { pos += 1; continue; }
    pos += 2;
    if (m == 0xd9) break;                                     /* EOI */
    if (m == 0x00 || (m >= 0xd0 && m <= 0xd8)) continue;
    if (pos + 2 > len) { rc = ORACLE_E_TRUNCATED; goto done; }
    const int seglen = be16(file + pos);
    const uint8_t *p = file + pos + 2;
    const int plen = seglen - 2;
    if (plen < 0 || pos + (size_t)seglen > len) { rc = ORACLE_E_TRUNCATED; goto done; }
    pos += seglen;
    switch (m) {
      case 0xc2:                                              /* SOF2: same layout as SOF0 (B.2.2) */
        rc = read_sof(D, 0xc0, p, plen); got_frame = 1;
        if (!rc) mcus = (int32_t *)calloc((size_t)D->bwr * D->bhr * 3 * 64, sizeof(int32_t));
        break;
      case 0xc0: rc = ORACLE_E_PROGRESSIVE; break;            /* a baseline file: use oracle_decode */
      case 0xc4: rc = read_dht(D, p, plen); break;
      case 0xdb: rc = read_dqt(D, p, plen); break;
      case 0xdd: if (plen != 2) rc = ORACLE_E_DRI; else D->ri = be16(p); break;
      case 0xda: {
        if (!got_frame || !mcus) { rc = ORACLE_E_NO_SOS; break; }
        pscan_t S;
        if (plen < 1) { rc = ORACLE_E_SOS; break; }
        S.ns = p[0];
        if (S.ns < 1 || S.ns > D->ncomp || plen != 1 + 2 * S.ns + 3) { rc = ORACLE_E_SOS; break; }
        for (int j = 0; j < S.ns && !rc; ++j) {
          int ci = -1;
          for (int i = 0; i < D->ncomp; ++i) if (D->comp_id[i] == p[1 + 2 * j]) ci = i;
          if (ci < 0) { rc = ORACLE_E_SOS; break; }
          S.ci[j] = ci;
          D->comp[ci].td = p[2 + 2 * j] >> 4; D->comp[ci].ta = p[2 + 2 * j] & 15;
          if (D->comp[ci].td > 3 || D->comp[ci].ta > 3) rc = ORACLE_E_TABLE_ID;
        }
        if (rc) break;
        const uint8_t *q = p + 1 + 2 * S.ns;
        S.ss = q[0]; S.se = q[1]; S.ah = q[2] >> 4; S.al = q[2] & 15;
        if (S.ss > S.se || S.se > 63 || (S.ss == 0 && S.se != 0) || (S.ss != 0 && S.ns != 1) || S.al > 13) { rc = ORACLE_E_SOS; break; }
        for (int j = 0; j < S.ns; ++j) {                      /* the tables this scan names must exist (jpeg.cpp:757-774) */
          const comp_t *c = &D->comp[S.ci[j]];
          if (S.ss == 0 && S.ah == 0 && !D->dc[c->td].set) rc = ORACLE_E_TABLE_MISSING;
          if (S.ss != 0 && !D->ac[c->ta].set) rc = ORACLE_E_TABLE_MISSING;
        }
        if (rc) break;
        size_t used = 0;
        const long n = prog_unstuff(file + pos, len - pos, clean, &used);
        D->clean = clean; D->clean_n = (size_t)n; D->clean_pos = 0; D->st_c = -1; D->err = 0;
        rc = prog_scan(D, &S, mcus);
        pos += used; ++nscans;
        break;
      }
      default: break;                                         /* APPn, COM, ... */
    }
    if (rc) goto done;
  }
  if (!got_frame || !nscans) { rc = ORACLE_E_NO_SOS; goto done; }
  for (int i = 0; i < D->ncomp; i++) if (D->comp[i].tq > 3 || !D->q[D->comp[i].tq].set) { rc = ORACLE_E_TABLE_MISSING; goto done; }
  {
    double tt[3];
    finish_pixels(D, mcus, coef_out, rgb_out, tt);
  }
done:
  if (info) {
    memset(info, 0, sizeof *info);
    info->width = D->width; info->height = D->height; info->bw = D->bw; info->bh = D->bh;
    info->bwr = D->bwr; info->bhr = D->bhr; info->hs = D->hs; info->vs = D->vs;
    info->ncomp = D->ncomp; info->restart_interval = D->ri; info->scan_unstuffed = nscans;
  }
  free(clean); free(mcus); free(D);
  return rc;
}

void oracle_free(void *p) { free(p); }
