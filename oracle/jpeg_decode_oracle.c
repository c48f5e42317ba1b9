/* jpeg_decode_oracle.c -- TEST INFRASTRUCTURE (see oracle.h): CPU restatement of baseline JPEG decoding as the
 * reference gets it on Node: src/decoders/jpeg-decoder.ts:196-215 hands the file to sharp (libvips -> libjpeg-turbo)
 * with default settings and takes RGBA8 (or gray8) rows.  libjpeg-turbo is a third-party dependency that is not
 * under /root/reference; this file restates its published algorithm for the default decode path:
 *
 *   entropy decoding   jdhuff.c   (Huffman, sequential DCT, restart intervals; ITU-T T.81 Annex F.2)
 *   dequantise + IDCT  jidctint.c (jpeg_idct_islow: JDCT_ISLOW is the default dct_method)
 *   chroma upsampling  jdsample.c (do_fancy_upsampling = TRUE: h2v1_fancy_upsample, h2v2_fancy_upsample; box
 *                                  replication when the downsampled width is <= 2), context rows of jdmainct.c
 *   colour             jdcolor.c  (ycc_rgb_convert, 16-bit fixed point tables)
 *
 * Parity is pinned against Pillow 12.2.0 / libjpeg-turbo 3.1.4.1 in this container: tests/golden/jpeg_decode holds
 * JPEG files and the pixels Pillow decodes from them (tests/golden/make_jpeg_decode_golden.py), and
 * tests/test_jpeg_decode_oracle.py compares this file with them byte for byte.
 *
 * Scope: SOF0/SOF1, 8-bit, one interleaved scan, 1 (gray) or 3 (YCbCr) components, luma sampling 1x1, 2x1 or 2x2
 * with 1x1 chroma.  Everything else returns an error code (see jdec_message). */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum {
  JDEC_OK = 0, JDEC_E_SOI, JDEC_E_TRUNC_SOF, JDEC_E_NO_SOF, JDEC_E_PROGRESSIVE, JDEC_E_PRECISION, JDEC_E_COMPONENTS,
  JDEC_E_SAMPLING, JDEC_E_TABLE, JDEC_E_SCAN, JDEC_E_TRUNCATED, JDEC_E_HUFF, JDEC_E_RST, JDEC_E_COLORSPACE, JDEC_E_NOMEM
};

const char *jdec_message(int code) {
  switch (code) {
    case JDEC_OK: return "ok";
    case JDEC_E_SOI: return "Invalid JPEG: missing SOI marker";          /* jpeg-decoder.ts:49 */
    case JDEC_E_TRUNC_SOF: return "Invalid JPEG: truncated SOF marker";  /* :74 */
    case JDEC_E_NO_SOF: return "Invalid JPEG: no SOF marker found";      /* :93 */
    case JDEC_E_PROGRESSIVE: return "Unsupported JPEG: only baseline and extended sequential Huffman frames (SOF0, SOF1) are decoded on the device";
    case JDEC_E_PRECISION: return "Unsupported JPEG: sample precision must be 8 bits";
    case JDEC_E_COMPONENTS: return "Unsupported JPEG: 1 (gray) or 3 (YCbCr) components in one interleaved scan";
    case JDEC_E_SAMPLING: return "Unsupported JPEG: luma sampling must be 1x1, 2x1 or 2x2 with 1x1 chroma";
    case JDEC_E_TABLE: return "Invalid JPEG: bad or missing quantisation or Huffman table";
    case JDEC_E_SCAN: return "Invalid JPEG: bad scan header";
    case JDEC_E_TRUNCATED: return "Invalid JPEG: entropy-coded data ends early";
    case JDEC_E_HUFF: return "Invalid JPEG: bad Huffman code";
    case JDEC_E_RST: return "Invalid JPEG: restart marker missing or out of order";
    case JDEC_E_COLORSPACE: return "Unsupported JPEG: 3-component files must be YCbCr";
    default: return "out of memory";
  }
}

static const uint8_t k_natural[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                      41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                      30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

typedef struct {
  int present;
  uint8_t bits[17];     /* bits[l] = codes of length l */
  uint8_t vals[256];
  /* canonical decoding (T.81 F.2.2.3): */
  int32_t maxcode[18];  /* largest code of length l, -1 if none */
  int32_t valptr[17];   /* index of the first value of length l */
  int32_t mincode[17];
} Huff;

typedef struct {
  uint32_t width, height, ncomp, hmax, vmax;
  struct { uint32_t id, h, v, tq, td, ta; uint32_t wb, hb; /* blocks per row / column incl. MCU padding */
           uint32_t dw, dh; /* downsampled size in samples */ } c[3];
  uint16_t q[4][64];    /* natural order */
  int qset[4];
  Huff dc[4], ac[4];
  uint32_t restart;     /* MCUs per restart interval, 0 = none */
  uint32_t mcus_x, mcus_y;
  const uint8_t *ecs;
  size_t ecs_len;
} Frame;

static void huff_derive(Huff *h) {
  int code = 0, k = 0;
  for (int l = 1; l <= 16; l++) {
    h->valptr[l] = k;
    h->mincode[l] = code;
    k += h->bits[l];
    code += h->bits[l];
    h->maxcode[l] = h->bits[l] ? code - 1 : -1;
    code <<= 1;
  }
  h->maxcode[17] = 0x7fffffff;
}

/* Markers up to and including SOS; the reference's own header scan (jpeg-decoder.ts:46-94) gives the first three errors. */
static int parse(const uint8_t *d, size_t n, Frame *f) {
  memset(f, 0, sizeof *f);
  if (n < 2 || d[0] != 0xff || d[1] != 0xd8) return JDEC_E_SOI;
  size_t o = 2;
  int have_sof = 0, adobe = -1, jfif = 0;
  while (o + 1 < n) {
    if (d[o] != 0xff) { o++; continue; }
    const uint8_t m = d[o + 1];
    o += 2;
    if (m == 0xff) { o--; continue; }
    if (m == 0x00 || m == 0x01 || (m >= 0xd0 && m <= 0xd8)) continue;
    if (o + 2 > n) break;
    const size_t len = ((size_t)d[o] << 8) | d[o + 1];
    if (len < 2 || o + len > n) {
      if (m >= 0xc0 && m <= 0xcf && m != 0xc4 && m != 0xc8 && m != 0xcc) return JDEC_E_TRUNC_SOF;
      return JDEC_E_TRUNCATED;
    }
    const uint8_t *p = d + o + 2;
    const size_t pl = len - 2;
    if (m == 0xc0 || m == 0xc1) {
      if (pl < 6) return JDEC_E_TRUNC_SOF;
      if (p[0] != 8) return JDEC_E_PRECISION;
      f->height = ((uint32_t)p[1] << 8) | p[2];
      f->width = ((uint32_t)p[3] << 8) | p[4];
      f->ncomp = p[5];
      if (f->ncomp != 1 && f->ncomp != 3) return JDEC_E_COMPONENTS;
      if (pl < 6 + 3 * (size_t)f->ncomp || !f->width || !f->height) return JDEC_E_TRUNC_SOF;
      for (uint32_t i = 0; i < f->ncomp; i++) {
        f->c[i].id = p[6 + 3 * i];
        f->c[i].h = p[7 + 3 * i] >> 4;
        f->c[i].v = p[7 + 3 * i] & 15;
        f->c[i].tq = p[8 + 3 * i];
        if (f->c[i].tq > 3) return JDEC_E_TABLE;
      }
      have_sof = 1;
    } else if (m >= 0xc2 && m <= 0xcf && m != 0xc4 && m != 0xc8 && m != 0xcc) {
      return JDEC_E_PROGRESSIVE;
    } else if (m == 0xc4) {
      size_t q = 0;
      while (q + 17 <= pl) {
        const int tc = p[q] >> 4, th = p[q] & 15;
        if (tc > 1 || th > 3) return JDEC_E_TABLE;
        Huff *h = tc ? &f->ac[th] : &f->dc[th];
        memset(h, 0, sizeof *h);
        int total = 0;
        for (int l = 1; l <= 16; l++) { h->bits[l] = p[q + l]; total += p[q + l]; }
        if (total > 256 || q + 17 + (size_t)total > pl) return JDEC_E_TABLE;
        memcpy(h->vals, p + q + 17, (size_t)total);
        h->present = 1;
        huff_derive(h);
        q += 17 + (size_t)total;
      }
    } else if (m == 0xdb) {
      size_t q = 0;
      while (q < pl) {
        const int pq = p[q] >> 4, tq = p[q] & 15;
        if (tq > 3 || pq > 1 || q + 1 + (pq ? 128u : 64u) > pl) return JDEC_E_TABLE;
        for (int k = 0; k < 64; k++)
          f->q[tq][k_natural[k]] = pq ? (uint16_t)((p[q + 1 + 2 * k] << 8) | p[q + 2 + 2 * k]) : p[q + 1 + k];
        f->qset[tq] = 1;
        q += 1 + (pq ? 128u : 64u);
      }
    } else if (m == 0xdd) {
      if (pl < 2) return JDEC_E_TRUNCATED;
      f->restart = ((uint32_t)p[0] << 8) | p[1];
    } else if (m == 0xe0) {
      if (pl >= 5 && !memcmp(p, "JFIF", 5)) jfif = 1;
    } else if (m == 0xee) {
      if (pl >= 12 && !memcmp(p, "Adobe", 5)) adobe = p[11];
    } else if (m == 0xda) {
      if (!have_sof) return JDEC_E_NO_SOF;
      if (pl < 1 || p[0] != f->ncomp || pl < 1 + 2 * (size_t)f->ncomp + 3) return p[0] != f->ncomp ? JDEC_E_COMPONENTS : JDEC_E_SCAN;
      for (uint32_t i = 0; i < f->ncomp; i++) {
        if (p[1 + 2 * i] != f->c[i].id) return JDEC_E_SCAN;
        f->c[i].td = p[2 + 2 * i] >> 4;
        f->c[i].ta = p[2 + 2 * i] & 15;
        if (f->c[i].td > 3 || f->c[i].ta > 3 || !f->dc[f->c[i].td].present || !f->ac[f->c[i].ta].present || !f->qset[f->c[i].tq])
          return JDEC_E_TABLE;
      }
      const uint8_t *s = p + 1 + 2 * f->ncomp;
      if (s[0] != 0 || s[1] != 63 || s[2] != 0) return JDEC_E_SCAN;
      /* colour space the way jdapimin.c default_decompress_parms guesses it */
      if (f->ncomp == 3 && !jfif) {
        if (adobe == 0) return JDEC_E_COLORSPACE;
        if (adobe < 0 && f->c[0].id == 'R' && f->c[1].id == 'G' && f->c[2].id == 'B') return JDEC_E_COLORSPACE;
      }
      if (f->ncomp == 1) { f->c[0].h = f->c[0].v = 1; }   /* a single-component scan is never interleaved */
      else {
        if (f->c[1].h != 1 || f->c[1].v != 1 || f->c[2].h != 1 || f->c[2].v != 1) return JDEC_E_SAMPLING;
        if (!((f->c[0].h == 1 && f->c[0].v == 1) || (f->c[0].h == 2 && f->c[0].v == 1) || (f->c[0].h == 2 && f->c[0].v == 2)))
          return JDEC_E_SAMPLING;
      }
      f->hmax = f->c[0].h; f->vmax = f->c[0].v;
      f->mcus_x = (f->width + 8 * f->hmax - 1) / (8 * f->hmax);
      f->mcus_y = (f->height + 8 * f->vmax - 1) / (8 * f->vmax);
      for (uint32_t i = 0; i < f->ncomp; i++) {
        f->c[i].wb = f->mcus_x * f->c[i].h;
        f->c[i].hb = f->mcus_y * f->c[i].v;
        f->c[i].dw = (f->width * f->c[i].h + f->hmax - 1) / f->hmax;
        f->c[i].dh = (f->height * f->c[i].v + f->vmax - 1) / f->vmax;
      }
      f->ecs = d + o + len;
      f->ecs_len = n - (o + len);
      return JDEC_OK;
    }
    o += len;
  }
  return have_sof ? JDEC_E_TRUNCATED : JDEC_E_NO_SOF;
}

/* ---- bit reader over the entropy-coded segment (T.81 F.2.2.5, byte stuffing B.1.1.5) ---- */
typedef struct {
  const uint8_t *p, *end;
  uint32_t acc;
  int cnt;
  int hit_marker;   /* a marker other than a stuffed zero was met: zeros are fed from here on (jdhuff.c) */
  int fed;          /* zero bits fed since then */
} Bits;

static void fill(Bits *b) {
  while (b->cnt <= 24) {
    uint32_t byte = 0;
    if (!b->hit_marker && b->p < b->end) {
      byte = *b->p;
      if (byte == 0xff) {
        if (b->p + 1 < b->end && b->p[1] == 0x00) b->p += 2;
        else { b->hit_marker = 1; byte = 0; }
      } else b->p++;
    } else b->hit_marker = 1;
    if (b->hit_marker) b->fed += 8;
    b->acc |= byte << (24 - b->cnt);
    b->cnt += 8;
  }
}
static uint32_t getbits(Bits *b, int n) {
  if (!n) return 0;
  fill(b);
  const uint32_t v = b->acc >> (32 - n);
  b->acc <<= n;
  b->cnt -= n;
  return v;
}
static int decode(Bits *b, const Huff *h) {
  fill(b);
  int32_t code = 0;
  for (int l = 1; l <= 16; l++) {
    code = (int32_t)(b->acc >> (32 - l));
    if (h->maxcode[l] >= 0 && code <= h->maxcode[l] && code >= h->mincode[l]) {
      b->acc <<= l;
      b->cnt -= l;
      return h->vals[h->valptr[l] + code - h->mincode[l]];
    }
  }
  return -1;
}
static int extend(uint32_t r, int s) { return r < (1u << (s - 1)) ? (int)r - (int)((1u << s) - 1u) : (int)r; }

/* ---- jidctint.c jpeg_idct_islow: dequantised 8x8 block -> 64 samples ---- */
#define FIX_0_298631336 2446
#define FIX_0_390180644 3196
#define FIX_0_541196100 4433
#define FIX_0_765366865 6270
#define FIX_0_899976223 7373
#define FIX_1_175875602 9633
#define FIX_1_501321110 12299
#define FIX_1_847759065 15137
#define FIX_1_961570560 16069
#define FIX_2_053119869 16819
#define FIX_2_562915447 20995
#define FIX_3_072711026 25172
#define DESCALE(x, n) (((x) + ((int32_t)1 << ((n) - 1))) >> (n))

static uint8_t idct_limit(int32_t x) {   /* range_limit[x & RANGE_MASK] of the post-IDCT table (jdmaster.c) */
  x &= 1023;
  if (x < 128) return (uint8_t)(x + 128);
  if (x < 512) return 255;
  if (x < 896) return 0;
  return (uint8_t)(x - 896);
}

static void idct_islow(const int16_t *coef, const uint16_t *q, uint8_t *out /* 64, row major */) {
  int32_t ws[64];
  for (int c = 0; c < 8; c++) {
    int32_t in[8];
    for (int r = 0; r < 8; r++) in[r] = (int32_t)coef[8 * r + c] * (int32_t)q[8 * r + c];
    int32_t z2 = in[2], z3 = in[6];
    int32_t z1 = (z2 + z3) * FIX_0_541196100;
    int32_t tmp2 = z1 + z3 * (-FIX_1_847759065);
    int32_t tmp3 = z1 + z2 * FIX_0_765366865;
    z2 = in[0]; z3 = in[4];
    int32_t tmp0 = (z2 + z3) * 8192, tmp1 = (z2 - z3) * 8192;
    const int32_t tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    tmp0 = in[7]; tmp1 = in[5]; tmp2 = in[3]; tmp3 = in[1];
    z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2;
    int32_t z4 = tmp1 + tmp3;
    const int32_t z5 = (z3 + z4) * FIX_1_175875602;
    tmp0 *= FIX_0_298631336; tmp1 *= FIX_2_053119869; tmp2 *= FIX_3_072711026; tmp3 *= FIX_1_501321110;
    z1 *= -FIX_0_899976223; z2 *= -FIX_2_562915447; z3 *= -FIX_1_961570560; z4 *= -FIX_0_390180644;
    z3 += z5; z4 += z5;
    tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
    ws[0 * 8 + c] = DESCALE(tmp10 + tmp3, 11); ws[7 * 8 + c] = DESCALE(tmp10 - tmp3, 11);
    ws[1 * 8 + c] = DESCALE(tmp11 + tmp2, 11); ws[6 * 8 + c] = DESCALE(tmp11 - tmp2, 11);
    ws[2 * 8 + c] = DESCALE(tmp12 + tmp1, 11); ws[5 * 8 + c] = DESCALE(tmp12 - tmp1, 11);
    ws[3 * 8 + c] = DESCALE(tmp13 + tmp0, 11); ws[4 * 8 + c] = DESCALE(tmp13 - tmp0, 11);
  }
  for (int r = 0; r < 8; r++) {
    const int32_t *w = ws + 8 * r;
    int32_t z2 = w[2], z3 = w[6];
    int32_t z1 = (z2 + z3) * FIX_0_541196100;
    int32_t tmp2 = z1 + z3 * (-FIX_1_847759065);
    int32_t tmp3 = z1 + z2 * FIX_0_765366865;
    int32_t tmp0 = (w[0] + w[4]) * 8192, tmp1 = (w[0] - w[4]) * 8192;
    const int32_t tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
    tmp0 = w[7]; tmp1 = w[5]; tmp2 = w[3]; tmp3 = w[1];
    z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2;
    int32_t z4 = tmp1 + tmp3;
    const int32_t z5 = (z3 + z4) * FIX_1_175875602;
    tmp0 *= FIX_0_298631336; tmp1 *= FIX_2_053119869; tmp2 *= FIX_3_072711026; tmp3 *= FIX_1_501321110;
    z1 *= -FIX_0_899976223; z2 *= -FIX_2_562915447; z3 *= -FIX_1_961570560; z4 *= -FIX_0_390180644;
    z3 += z5; z4 += z5;
    tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
    uint8_t *o = out + 8 * r;
    o[0] = idct_limit(DESCALE(tmp10 + tmp3, 18)); o[7] = idct_limit(DESCALE(tmp10 - tmp3, 18));
    o[1] = idct_limit(DESCALE(tmp11 + tmp2, 18)); o[6] = idct_limit(DESCALE(tmp11 - tmp2, 18));
    o[2] = idct_limit(DESCALE(tmp12 + tmp1, 18)); o[5] = idct_limit(DESCALE(tmp12 - tmp1, 18));
    o[3] = idct_limit(DESCALE(tmp13 + tmp0, 18)); o[4] = idct_limit(DESCALE(tmp13 - tmp0, 18));
  }
}

static uint8_t clamp8(int v) { return (uint8_t)(v < 0 ? 0 : v > 255 ? 255 : v); }

/* One upsampled chroma sample at full-resolution (x, y): jdsample.c.  plane: the component's samples with `stride`
 * bytes per row; dw x dh: its downsampled size (the real samples; the rest of the plane is block padding). */
static int chroma_at(const uint8_t *plane, size_t stride, uint32_t dw, uint32_t dh, uint32_t hs, uint32_t vs, uint32_t x, uint32_t y) {
  if (hs == 1 && vs == 1) return plane[(size_t)y * stride + x];
  const uint32_t cx = x >> 1;
  if (hs == 2 && vs == 1) {
    const uint8_t *r = plane + (size_t)y * stride;
    if (dw <= 2) return r[cx];                                  /* h2v1_upsample: replication */
    if (x == 0) return r[0];                                    /* h2v1_fancy_upsample */
    if (x == 2 * dw - 1) return r[dw - 1];
    if (x & 1) return (r[cx] * 3 + r[cx + 1] + 2) >> 2;
    return (r[cx] * 3 + r[cx - 1] + 1) >> 2;
  }
  /* h2v2 */
  const uint32_t cy = y >> 1;
  if (dw <= 2) return plane[(size_t)cy * stride + cx];          /* h2v2_upsample: replication */
  const uint32_t ny = (y & 1) ? (cy + 1 < dh ? cy + 1 : dh - 1) : (cy ? cy - 1 : 0);   /* context rows, edges replicated (jdmainct.c) */
  const uint8_t *r0 = plane + (size_t)cy * stride, *r1 = plane + (size_t)ny * stride;
  const int cur = r0[cx] * 3 + r1[cx];
  if (x == 0) return (cur * 4 + 8) >> 4;
  if (x == 2 * dw - 1) return (cur * 4 + 7) >> 4;
  if (x & 1) return (cur * 3 + (r0[cx + 1] * 3 + r1[cx + 1]) + 7) >> 4;
  return (cur * 3 + (r0[cx - 1] * 3 + r1[cx - 1]) + 8) >> 4;
}

/* Geometry without decoding: 0 or an error code. */
int jdec_info(const uint8_t *data, size_t len, uint32_t *width, uint32_t *height, uint32_t *channels) {
  Frame f;
  const int rc = parse(data, len, &f);
  if (rc) return rc;
  *width = f.width; *height = f.height; *channels = f.ncomp == 1 ? 1 : 4;
  return JDEC_OK;
}

/* Decodes into out: width * height * channels bytes, RGBA8 (alpha 255) or gray8, rows top to bottom. */
int jdec_decode(const uint8_t *data, size_t len, uint8_t *out, size_t cap) {
  Frame f;
  int rc = parse(data, len, &f);
  if (rc) return rc;
  const uint32_t ch = f.ncomp == 1 ? 1 : 4;
  if (cap < (size_t)f.width * f.height * ch) return JDEC_E_NOMEM;
  uint8_t *plane[3] = {0, 0, 0};
  size_t stride[3];
  for (uint32_t i = 0; i < f.ncomp; i++) {
    stride[i] = (size_t)f.c[i].wb * 8;
    plane[i] = (uint8_t *)malloc(stride[i] * f.c[i].hb * 8);
    if (!plane[i]) { for (uint32_t k = 0; k < i; k++) free(plane[k]); return JDEC_E_NOMEM; }
  }
  Bits b;
  memset(&b, 0, sizeof b);
  b.p = f.ecs; b.end = f.ecs + f.ecs_len;
  int pred[3] = {0, 0, 0};
  uint32_t to_go = f.restart, next_rst = 0;
  for (uint32_t my = 0; my < f.mcus_y && !rc; my++) {
    for (uint32_t mx = 0; mx < f.mcus_x && !rc; mx++) {
      if (f.restart && to_go == 0) {   /* process_restart (jdhuff.c): byte-align, expect RSTn */
        if (b.fed > b.cnt) { rc = JDEC_E_TRUNCATED; break; }   /* the interval used bits that are not there */
        b.acc = 0; b.cnt = 0; b.fed = 0;
        const uint8_t *p = b.p;
        while (p < b.end && *p != 0xff) p++;                 /* the marker follows the interval's last byte */
        while (p + 1 < b.end && p[1] == 0xff) p++;
        if (p + 1 >= b.end || p[1] != 0xd0 + next_rst) { rc = JDEC_E_RST; break; }
        b.p = p + 2; b.hit_marker = 0;
        next_rst = (next_rst + 1) & 7;
        pred[0] = pred[1] = pred[2] = 0;
        to_go = f.restart;
      }
      for (uint32_t i = 0; i < f.ncomp && !rc; i++) {
        for (uint32_t v = 0; v < f.c[i].v && !rc; v++) {
          for (uint32_t h = 0; h < f.c[i].h; h++) {
            int16_t blk[64];
            memset(blk, 0, sizeof blk);
            int s = decode(&b, &f.dc[f.c[i].td]);
            if (s < 0 || s > 15) { rc = JDEC_E_HUFF; break; }
            if (s) pred[i] += extend(getbits(&b, s), s);
            blk[0] = (int16_t)pred[i];
            for (int k = 1; k < 64; k++) {
              s = decode(&b, &f.ac[f.c[i].ta]);
              if (s < 0) { rc = JDEC_E_HUFF; break; }
              const int r = s >> 4;
              s &= 15;
              if (s) {
                k += r;
                if (k > 63) { rc = JDEC_E_HUFF; break; }
                blk[k_natural[k]] = (int16_t)extend(getbits(&b, s), s);
              } else {
                if (r != 15) break;
                k += 15;
              }
            }
            if (rc) break;
            uint8_t px[64];
            idct_islow(blk, f.q[f.c[i].tq], px);
            const uint32_t bx = mx * f.c[i].h + h, by = my * f.c[i].v + v;
            for (int r = 0; r < 8; r++) memcpy(plane[i] + ((size_t)by * 8 + r) * stride[i] + (size_t)bx * 8, px + 8 * r, 8);
          }
        }
      }
      if (f.restart) to_go--;
    }
  }
  if (!rc && b.fed > b.cnt) rc = JDEC_E_TRUNCATED;
  if (!rc) {
    for (uint32_t y = 0; y < f.height; y++) {
      uint8_t *o = out + (size_t)y * f.width * ch;
      for (uint32_t x = 0; x < f.width; x++) {
        const int Y = plane[0][(size_t)y * stride[0] + x];
        if (f.ncomp == 1) { o[x] = (uint8_t)Y; continue; }
        const int cb = chroma_at(plane[1], stride[1], f.c[1].dw, f.c[1].dh, f.hmax, f.vmax, x, y) - 128;
        const int cr = chroma_at(plane[2], stride[2], f.c[2].dw, f.c[2].dh, f.hmax, f.vmax, x, y) - 128;
        /* jdcolor.c build_ycc_rgb_table / ycc_rgb_convert, SCALEBITS 16; >> is arithmetic */
        o[4 * x + 0] = clamp8(Y + ((91881 * cr + 32768) >> 16));
        o[4 * x + 1] = clamp8(Y + ((-22554 * cb + 32768 - 46802 * cr) >> 16));
        o[4 * x + 2] = clamp8(Y + ((116130 * cb + 32768) >> 16));
        o[4 * x + 3] = 255;
      }
    }
  }
  for (uint32_t i = 0; i < f.ncomp; i++) free(plane[i]);
  return rc;
}
