/*
 * jpeg_oracle.c -- CPU definition of the JPEG half of the hot path (kernel d).
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.h).
 *
 * PARITY UNPINNED against the reference: image-stitch delegates every JPEG
 * computation to the un-vendored package jpeg-encoder-wasm 0.1.0
 * (github:jburnhams/jpeg-encoder#9267440ca1bd050e198fb05aab36beee1a32a458,
 * package.json:114); neither its source nor its .wasm is under /root/reference,
 * and no reference test pins coefficients, bytes or PSNR
 * (tests/unit/jpeg-structure.test.ts:52-237 checks SOI/EOI/decodability/dims).
 * Call sites this file stands in for: src/jpeg-encoder.ts:131-136 (constructor
 * width,height,Rgba,quality), :162 (encode_strip of <= 8 RGBA rows), :180
 * (finish -> EOI); strip feed src/image-concat-core.ts:881-914, :1362-1393.
 *
 * The definition fixed here (and pinned instead against libjpeg-turbo through
 * PIL in tests/test_jpeg_oracle.py -- decoded pixels identical to PIL's own
 * 4:4:4 baseline encoding at the same quality):
 *   - baseline sequential DCT, 8-bit, 3 components Y Cb Cr, 4:4:4, one MCU =
 *     one 8x8 block of each component; the alpha byte is dropped
 *   - JFIF full-range BT.601 colour transform in libjpeg's 16-bit fixed point
 *   - libjpeg jfdctint ("islow") forward DCT, operation for operation
 *   - Annex K quantisation tables scaled with the libjpeg quality rule,
 *     round-half-away division
 *   - Annex K Huffman tables
 *   - partial edge blocks replicate the last column / row (src/jpeg-encoder.ts:152-153)
 *   - restart interval = one MCU row (DRI): every MCU row is an independent,
 *     byte-aligned entropy segment followed by RSTm -- the unit of GPU and
 *     multi-GPU parallelism.  This changes bytes, not coefficients.
 */
#include "oracle.h"

#include <stdlib.h>
#include <string.h>

static const uint8_t k_lum_q[64] = {16, 11, 10, 16, 24,  40,  51,  61,  12, 12, 14, 19, 26,  58,  60,  55,
                                    14, 13, 16, 24, 40,  57,  69,  56,  14, 17, 22, 29, 51,  87,  80,  62,
                                    18, 22, 37, 56, 68,  109, 103, 77,  24, 35, 55, 64, 81,  104, 113, 92,
                                    49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
static const uint8_t k_chr_q[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99,
                                    24, 26, 56, 99, 99, 99, 99, 99, 47, 66, 99, 99, 99, 99, 99, 99,
                                    99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
                                    99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
static const uint8_t k_zigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,
                                     12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,  7,  14, 21, 28,
                                     35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
                                     58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

static const uint8_t k_dc_lum_bits[16] = {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
static const uint8_t k_dc_chr_bits[16] = {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
static const uint8_t k_dc_vals[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
static const uint8_t k_ac_lum_bits[16] = {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d};
static const uint8_t k_ac_lum_vals[162] = {
    0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71,
    0x14, 0x32, 0x81, 0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72,
    0x82, 0x09, 0x0a, 0x16, 0x17, 0x18, 0x19, 0x1a, 0x25, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x34, 0x35, 0x36, 0x37,
    0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59,
    0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x83,
    0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3,
    0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3,
    0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe1, 0xe2,
    0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf1, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};
static const uint8_t k_ac_chr_bits[16] = {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77};
static const uint8_t k_ac_chr_vals[162] = {
    0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22,
    0x32, 0x81, 0x08, 0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1,
    0x0a, 0x16, 0x24, 0x34, 0xe1, 0x25, 0xf1, 0x17, 0x18, 0x19, 0x1a, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x35, 0x36,
    0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58,
    0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a,
    0x82, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a,
    0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba,
    0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda,
    0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};

/* libjpeg jpeg_quality_scaling + jpeg_add_quant_table(force_baseline). Natural order. */
void orc_jpeg_quant_tables(int quality, uint16_t lum[64], uint16_t chr[64]) {
  if (quality <= 0) quality = 1;
  if (quality > 100) quality = 100;
  const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
  for (int i = 0; i < 64; i++) {
    long a = ((long)k_lum_q[i] * scale + 50L) / 100L, b = ((long)k_chr_q[i] * scale + 50L) / 100L;
    if (a <= 0) a = 1; if (a > 255) a = 255;
    if (b <= 0) b = 1; if (b > 255) b = 255;
    lum[i] = (uint16_t)a; chr[i] = (uint16_t)b;
  }
}

/* libjpeg jccolor.c rgb_ycc_convert, SCALEBITS = 16 */
void orc_jpeg_rgb_to_ycc(int r, int g, int b, int *y, int *cb, int *cr) {
  *y = (19595 * r + 38470 * g + 7471 * b + 32768) >> 16;
  *cb = (-11059 * r - 21709 * g + 32768 * b + (128 << 16) + 32767) >> 16;
  *cr = (32768 * r - 27439 * g - 5329 * b + (128 << 16) + 32767) >> 16;
}

#define DESCALE(x, n) (((x) + (1 << ((n)-1))) >> (n))
/* libjpeg jfdctint.c jpeg_fdct_islow: data = samples - 128, in place; output scaled by 8 */
void orc_jpeg_fdct_islow(int data[64]) {
  enum { CB = 13, P1 = 2 };
  int *d = data;
  for (int pass = 0; pass < 2; pass++) {
    for (int i = 0; i < 8; i++) {
      int *p = pass == 0 ? d + 8 * i : d + i;
      const int s = pass == 0 ? 1 : 8;
      int tmp0 = p[0] + p[7 * s], tmp7 = p[0] - p[7 * s], tmp1 = p[s] + p[6 * s], tmp6 = p[s] - p[6 * s];
      int tmp2 = p[2 * s] + p[5 * s], tmp5 = p[2 * s] - p[5 * s], tmp3 = p[3 * s] + p[4 * s], tmp4 = p[3 * s] - p[4 * s];
      int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
      if (pass == 0) { p[0] = (tmp10 + tmp11) << P1; p[4 * s] = (tmp10 - tmp11) << P1; }
      else { p[0] = DESCALE(tmp10 + tmp11, P1); p[4 * s] = DESCALE(tmp10 - tmp11, P1); }
      const int sh = pass == 0 ? CB - P1 : CB + P1;
      int z1 = (tmp12 + tmp13) * 4433;
      p[2 * s] = DESCALE(z1 + tmp13 * 6270, sh);
      p[6 * s] = DESCALE(z1 + tmp12 * (-15137), sh);
      z1 = tmp4 + tmp7;
      int z2 = tmp5 + tmp6, z3 = tmp4 + tmp6, z4 = tmp5 + tmp7, z5 = (z3 + z4) * 9633;
      tmp4 *= 2446; tmp5 *= 16819; tmp6 *= 25172; tmp7 *= 12299;
      z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
      z3 += z5; z4 += z5;
      p[7 * s] = DESCALE(tmp4 + z1 + z3, sh);
      p[5 * s] = DESCALE(tmp5 + z2 + z4, sh);
      p[3 * s] = DESCALE(tmp6 + z2 + z3, sh);
      p[s] = DESCALE(tmp7 + z1 + z4, sh);
    }
  }
}

/* libjpeg jcdctmgr.c forward_DCT quantisation: divisor = quantval << 3, round half away from zero */
int orc_jpeg_quantize(int coef, int quantval) {
  const int q = quantval << 3;
  int t = coef;
  if (t < 0) { t = -t; t += q >> 1; t /= q; return -t; }
  t += q >> 1;
  return t / q;
}

typedef struct { uint16_t code[256]; uint8_t size[256]; } huff_t;
static void build_huff(const uint8_t bits[16], const uint8_t *vals, huff_t *h) {
  memset(h, 0, sizeof *h);
  int code = 0, k = 0;
  for (int len = 1; len <= 16; len++) {
    for (int i = 0; i < bits[len - 1]; i++, k++) { h->code[vals[k]] = (uint16_t)code; h->size[vals[k]] = (uint8_t)len; code++; }
    code <<= 1;
  }
}

/* Quantised coefficients of MCU rows covering canvas rows [y0,y1): layout [mcu_row][mcu][comp][64], zigzag order.
 * rgba: rows y0.. of the canvas (only rows < height are read; replicated below), pitch bytes apart. */
int64_t orc_jpeg_coefficients(const uint8_t *rgba, size_t pitch, uint32_t width, uint32_t height, uint32_t y0,
                              uint32_t y1, int quality, int16_t *coefs) {
  if (y0 % 8 || y0 >= y1 || y1 > height || (y1 % 8 && y1 != height)) return ORC_E_INVALID;
  uint16_t ql[64], qc[64];
  orc_jpeg_quant_tables(quality, ql, qc);
  const uint32_t mcus = (width + 7) / 8, mrows = (y1 - y0 + 7) / 8;
  for (uint32_t mr = 0; mr < mrows; mr++)
    for (uint32_t m = 0; m < mcus; m++) {
      int blk[3][64];
      for (int yy = 0; yy < 8; yy++) {
        uint32_t y = y0 + mr * 8 + yy;
        if (y > height - 1) y = height - 1;
        for (int xx = 0; xx < 8; xx++) {
          uint32_t x = m * 8 + xx;
          if (x > width - 1) x = width - 1;
          const uint8_t *px = rgba + (size_t)(y - y0) * pitch + (size_t)x * 4;
          int Y, Cb, Cr;
          orc_jpeg_rgb_to_ycc(px[0], px[1], px[2], &Y, &Cb, &Cr);
          blk[0][yy * 8 + xx] = Y - 128; blk[1][yy * 8 + xx] = Cb - 128; blk[2][yy * 8 + xx] = Cr - 128;
        }
      }
      for (int c = 0; c < 3; c++) {
        orc_jpeg_fdct_islow(blk[c]);
        int16_t *o = coefs + (((size_t)mr * mcus + m) * 3 + c) * 64;
        const uint16_t *q = c == 0 ? ql : qc;
        for (int k = 0; k < 64; k++) o[k] = (int16_t)orc_jpeg_quantize(blk[c][k_zigzag[k]], q[k_zigzag[k]]);
      }
    }
  return (int64_t)mrows * mcus * 3 * 64;
}

typedef struct { uint8_t *out; size_t cap, n; uint64_t acc; int nbits; int overflow; } bitw_t;
static void put_byte(bitw_t *w, uint8_t b) {
  if (w->n < w->cap) w->out[w->n] = b; else w->overflow = 1;
  w->n++;
}
static void put_bits(bitw_t *w, uint32_t code, int size) {
  if (!size) return;
  w->acc = (w->acc << size) | (code & ((1u << size) - 1u));
  w->nbits += size;
  while (w->nbits >= 8) {
    const uint8_t b = (uint8_t)(w->acc >> (w->nbits - 8));
    put_byte(w, b);
    if (b == 0xFF) put_byte(w, 0);
    w->nbits -= 8;
  }
}
static int bit_size(int v) { int n = 0; if (v < 0) v = -v; while (v) { n++; v >>= 1; } return n; }

/* Entropy-coded bytes of the MCU rows covering canvas rows [y0,y1): per MCU row the DC predictors restart
 * at 0, the segment is padded with 1-bits to a byte boundary and followed by RSTm (m = global MCU row index
 * mod 8) unless it is the last MCU row of the image.  Returns the byte count (may exceed cap: nothing is
 * written past cap) or a negative error. */
int64_t orc_jpeg_encode_rows(const uint8_t *rgba, size_t pitch, uint32_t width, uint32_t height, uint32_t y0,
                             uint32_t y1, int quality, uint8_t *out, size_t cap) {
  if (y0 % 8 || y0 >= y1 || y1 > height || (y1 % 8 && y1 != height)) return ORC_E_INVALID;
  const uint32_t mcus = (width + 7) / 8, mrows = (y1 - y0 + 7) / 8, total_mrows = (height + 7) / 8;
  int16_t *coefs = (int16_t *)malloc((size_t)mrows * mcus * 3 * 64 * sizeof(int16_t));
  if (!coefs) return ORC_E_INVALID;
  orc_jpeg_coefficients(rgba, pitch, width, height, y0, y1, quality, coefs);
  huff_t dcl, dcc, acl, acc_;
  build_huff(k_dc_lum_bits, k_dc_vals, &dcl); build_huff(k_dc_chr_bits, k_dc_vals, &dcc);
  build_huff(k_ac_lum_bits, k_ac_lum_vals, &acl); build_huff(k_ac_chr_bits, k_ac_chr_vals, &acc_);
  bitw_t w; memset(&w, 0, sizeof w); w.out = out; w.cap = cap;
  for (uint32_t mr = 0; mr < mrows; mr++) {
    int pred[3] = {0, 0, 0};
    for (uint32_t m = 0; m < mcus; m++)
      for (int c = 0; c < 3; c++) {
        const int16_t *b = coefs + (((size_t)mr * mcus + m) * 3 + c) * 64;
        const huff_t *dc = c == 0 ? &dcl : &dcc, *ac = c == 0 ? &acl : &acc_;
        int diff = b[0] - pred[c]; pred[c] = b[0];
        int s = bit_size(diff);
        put_bits(&w, dc->code[s], dc->size[s]);
        if (s) put_bits(&w, (uint32_t)(diff < 0 ? diff - 1 : diff), s);
        int run = 0;
        for (int k = 1; k < 64; k++) {
          const int v = b[k];
          if (!v) { run++; continue; }
          while (run > 15) { put_bits(&w, ac->code[0xF0], ac->size[0xF0]); run -= 16; }
          s = bit_size(v);
          put_bits(&w, ac->code[(run << 4) | s], ac->size[(run << 4) | s]);
          put_bits(&w, (uint32_t)(v < 0 ? v - 1 : v), s);
          run = 0;
        }
        if (run) put_bits(&w, ac->code[0], ac->size[0]);
      }
    if (w.nbits) put_bits(&w, 0x7F, 8 - w.nbits);  /* pad with 1-bits */
    const uint32_t gr = y0 / 8 + mr;
    if (gr + 1 < total_mrows) { put_byte(&w, 0xFF); put_byte(&w, (uint8_t)(0xD0 + (gr & 7))); }
  }
  free(coefs);
  return (int64_t)w.n;
}

static size_t put_marker_seg(uint8_t *o, size_t n, size_t cap, uint8_t marker, const uint8_t *payload, size_t len) {
  const size_t need = 4 + len;
  if (n + need <= cap) {
    o[n] = 0xFF; o[n + 1] = marker; o[n + 2] = (uint8_t)((len + 2) >> 8); o[n + 3] = (uint8_t)((len + 2) & 0xFF);
    memcpy(o + n + 4, payload, len);
  }
  return n + need;
}

/* SOI, APP0 (JFIF 1.01), DQT x2, SOF0, DHT x4, DRI (one MCU row), SOS. Returns the byte count. */
int64_t orc_jpeg_header(uint32_t width, uint32_t height, int quality, uint8_t *out, size_t cap) {
  if (!width || !height || width > 65535 || height > 65535) return ORC_E_INVALID;
  uint16_t ql[64], qc[64];
  orc_jpeg_quant_tables(quality, ql, qc);
  size_t n = 0;
  uint8_t buf[512];
  if (cap >= 2) { out[0] = 0xFF; out[1] = 0xD8; }
  n = 2;
  const uint8_t app0[14] = {'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0};
  n = put_marker_seg(out, n, cap, 0xE0, app0, 14);
  for (int t = 0; t < 2; t++) {
    buf[0] = (uint8_t)t;
    for (int k = 0; k < 64; k++) buf[1 + k] = (uint8_t)(t == 0 ? ql : qc)[k_zigzag[k]];
    n = put_marker_seg(out, n, cap, 0xDB, buf, 65);
  }
  {
    const uint8_t sof[15] = {8, (uint8_t)(height >> 8), (uint8_t)height, (uint8_t)(width >> 8), (uint8_t)width, 3,
                             1, 0x11, 0, 2, 0x11, 1, 3, 0x11, 1};
    n = put_marker_seg(out, n, cap, 0xC0, sof, 15);
  }
  const uint8_t *bits[4] = {k_dc_lum_bits, k_ac_lum_bits, k_dc_chr_bits, k_ac_chr_bits};
  const uint8_t *vals[4] = {k_dc_vals, k_ac_lum_vals, k_dc_vals, k_ac_chr_vals};
  const uint8_t cls[4] = {0x00, 0x10, 0x01, 0x11};
  for (int t = 0; t < 4; t++) {
    int cnt = 0;
    for (int i = 0; i < 16; i++) cnt += bits[t][i];
    buf[0] = cls[t];
    memcpy(buf + 1, bits[t], 16);
    memcpy(buf + 17, vals[t], (size_t)cnt);
    n = put_marker_seg(out, n, cap, 0xC4, buf, 17 + (size_t)cnt);
  }
  {
    const uint32_t ri = (width + 7) / 8;
    const uint8_t dri[2] = {(uint8_t)(ri >> 8), (uint8_t)ri};
    n = put_marker_seg(out, n, cap, 0xDD, dri, 2);
  }
  {
    const uint8_t sos[10] = {3, 1, 0x00, 2, 0x11, 3, 0x11, 0, 63, 0};
    n = put_marker_seg(out, n, cap, 0xDA, sos, 10);
  }
  return (int64_t)n;
}
