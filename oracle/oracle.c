/*
 * oracle.c -- CPU restatement of image-stitch's PNG pixel pipeline.
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  Plain C99, single thread, written
 * to mirror the reference's arithmetic and control flow function by function
 * so that it can stand in for the TypeScript implementation where no JS
 * runtime exists.  Reference: jburnhams/image-stitch v1.1.53.
 */
#include "oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

/* JavaScript Math.round: nearest integer, ties towards +infinity. */
static double js_round(double x) {
  double r = floor(x);
  if (x - r >= 0.5) r += 1.0;
  return r;
}

/* ------------------------------------------------------------------ */
/* png-filter.ts                                                       */
/* ------------------------------------------------------------------ */

/* png-filter.ts:16-25 -- ties resolve a, then b, then c. */
int orc_paeth_predictor(int a, int b, int c) {
  int p = a + b - c;
  int pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  if (pa <= pb && pa <= pc) return a;
  if (pb <= pc) return b;
  return c;
}

/* utils.ts:92-101 */
int orc_samples_per_pixel(int color_type) {
  switch (color_type) {
    case 0: return 1;
    case 2: return 3;
    case 3: return 1;
    case 4: return 2;
    case 6: return 4;
    default: return -1;
  }
}

/* png-filter.ts:186-211 */
int orc_bytes_per_pixel(int bit_depth, int color_type) {
  int spp = orc_samples_per_pixel(color_type);
  if (spp < 0) return -1;
  return (spp * bit_depth + 7) / 8;
}

size_t orc_scanline_bytes(uint32_t width, int bit_depth, int color_type) {
  int spp = orc_samples_per_pixel(color_type);
  if (spp < 0) return 0;
  return (size_t)(((uint64_t)width * (uint64_t)bit_depth * (uint64_t)spp + 7) / 8);
}

/* png-filter.ts:34-83 */
int orc_unfilter_scanline(int filter_type, const uint8_t *in, const uint8_t *prev,
                          int bpp, size_t n, uint8_t *out) {
  size_t i;
  size_t b = (size_t)bpp;
  switch (filter_type) {
    case 0: /* :43-45 */
      memcpy(out, in, n);
      return ORC_OK;
    case 1: /* :47-52 */
      for (i = 0; i < n; i++) {
        int left = i >= b ? out[i - b] : 0;
        out[i] = (uint8_t)((in[i] + left) & 0xff);
      }
      return ORC_OK;
    case 2: /* :54-59 */
      for (i = 0; i < n; i++) {
        int up = prev ? prev[i] : 0;
        out[i] = (uint8_t)((in[i] + up) & 0xff);
      }
      return ORC_OK;
    case 3: /* :61-67 */
      for (i = 0; i < n; i++) {
        int left = i >= b ? out[i - b] : 0;
        int up = prev ? prev[i] : 0;
        out[i] = (uint8_t)((in[i] + ((left + up) >> 1)) & 0xff);
      }
      return ORC_OK;
    case 4: /* :69-76 */
      for (i = 0; i < n; i++) {
        int left = i >= b ? out[i - b] : 0;
        int up = prev ? prev[i] : 0;
        int ul = (prev && i >= b) ? prev[i - b] : 0;
        out[i] = (uint8_t)((in[i] + orc_paeth_predictor(left, up, ul)) & 0xff);
      }
      return ORC_OK;
    default: /* :78-79 */
      return ORC_E_FILTER;
  }
}

/* png-filter.ts:165-171: bytes are read as signed, cost is the absolute value */
static uint64_t signed_abs_sum(const uint8_t *p, size_t n) {
  uint64_t s = 0;
  size_t i;
  for (i = 0; i < n; i++) s += p[i] > 127 ? (uint64_t)(256 - p[i]) : (uint64_t)p[i];
  return s;
}

/* png-filter.ts:148-181 with the candidate builders :88-142 */
int orc_filter_scanline(const uint8_t *row, const uint8_t *prev, int bpp, size_t n,
                        uint8_t *out, uint8_t *scratch, uint64_t *sums) {
  uint8_t *f_sub = scratch, *f_up = scratch + n, *f_avg = scratch + 2 * n, *f_pae = scratch + 3 * n;
  const uint8_t *cand[5];
  uint64_t cost[5];
  size_t i, b = (size_t)bpp;
  int best = 0, k;
  for (i = 0; i < n; i++) {
    int x = row[i];
    int left = i >= b ? row[i - b] : 0;
    int up = prev ? prev[i] : 0;
    int ul = (prev && i >= b) ? prev[i - b] : 0;
    f_sub[i] = (uint8_t)((x - left) & 0xff);                            /* :88-95 */
    f_up[i] = (uint8_t)((x - up) & 0xff);                               /* :100-107 */
    f_avg[i] = (uint8_t)((x - ((left + up) >> 1)) & 0xff);              /* :112-124 */
    f_pae[i] = (uint8_t)((x - orc_paeth_predictor(left, up, ul)) & 0xff); /* :129-142 */
  }
  cand[0] = row; cand[1] = f_sub; cand[2] = f_up; cand[3] = f_avg; cand[4] = f_pae;
  for (k = 0; k < 5; k++) cost[k] = signed_abs_sum(cand[k], n);
  /* :160-178: strict '<' from Infinity, so the first minimum wins */
  for (k = 1; k < 5; k++) if (cost[k] < cost[best]) best = k;
  memcpy(out, cand[best], n);
  if (sums) for (k = 0; k < 5; k++) sums[k] = cost[k];
  return best;
}

/* ------------------------------------------------------------------ */
/* adam7.ts                                                            */
/* ------------------------------------------------------------------ */

static const int A7_XS[7] = {0, 4, 0, 2, 0, 1, 0}; /* adam7.ts:23-31 */
static const int A7_YS[7] = {0, 0, 4, 0, 2, 0, 1};
static const int A7_DX[7] = {8, 8, 4, 4, 2, 2, 1};
static const int A7_DY[7] = {8, 8, 8, 4, 4, 2, 2};

/* adam7.ts:36-43: ceil((dim-start)/step) clamped at 0 */
static uint32_t a7_extent(uint32_t dim, int start, int step) {
  if ((int64_t)dim <= start) return 0;
  return (uint32_t)(((int64_t)dim - start + step - 1) / step);
}

/* adam7.ts:52-110 */
int orc_deinterlace_adam7(const uint8_t *data, size_t len, uint32_t width, uint32_t height,
                          int bit_depth, int color_type, uint8_t *out,
                          int *err_pass, int *err_line, int *err_filter) {
  int bpp = orc_bytes_per_pixel(bit_depth, color_type);
  int spp = orc_samples_per_pixel(color_type);
  size_t final_len, src = 0;
  uint8_t *line_a, *line_b;
  int pass;
  if (bpp < 0) return ORC_E_INVALID;
  final_len = orc_scanline_bytes(width, bit_depth, color_type);
  memset(out, 0, (size_t)height * final_len);
  line_a = (uint8_t *)calloc(final_len + 8, 1);
  line_b = (uint8_t *)calloc(final_len + 8, 1);
  if (!line_a || !line_b) { free(line_a); free(line_b); return ORC_E_INVALID; }

  for (pass = 0; pass < 7; pass++) {
    uint32_t pw = a7_extent(width, A7_XS[pass], A7_DX[pass]);
    uint32_t ph = a7_extent(height, A7_YS[pass], A7_DY[pass]);
    size_t pass_len;
    uint8_t *cur = line_a, *prev = NULL;
    uint32_t py;
    if (pw == 0 || ph == 0) continue; /* :68-70 */
    pass_len = (size_t)(((uint64_t)pw * (uint64_t)bit_depth * (uint64_t)spp + 7) / 8);
    for (py = 0; py < ph; py++) {
      int ft, rc;
      size_t avail, fy, px;
      if (src >= len) { /* :79-82 */
        if (err_pass) *err_pass = pass + 1;
        if (err_line) *err_line = (int)py;
        free(line_a); free(line_b);
        return ORC_E_TRUNCATED;
      }
      ft = data[src++];
      /* :87-88 `slice` silently shortens a final partial line; missing bytes
       * then read as undefined -> 0 when scattered (:141-143, :172). */
      avail = len - src < pass_len ? len - src : pass_len;
      memset(cur, 0, pass_len);
      rc = orc_unfilter_scanline(ft, data + src, prev, bpp, avail, cur);
      src += avail;
      if (rc != ORC_OK) {
        if (err_filter) *err_filter = ft;
        free(line_a); free(line_b);
        return rc;
      }
      fy = (size_t)A7_YS[pass] + (size_t)py * (size_t)A7_DY[pass];
      if (bit_depth < 8) { /* distributeSubBytePixels :158-184, MSB first */
        int ppb = 8 / bit_depth;
        int mask = (1 << bit_depth) - 1;
        for (px = 0; px < pw; px++) {
          size_t fx = (size_t)A7_XS[pass] + px * (size_t)A7_DX[pass];
          int s_shift = (ppb - 1 - (int)(px % (size_t)ppb)) * bit_depth;
          int v = (cur[px / (size_t)ppb] >> s_shift) & mask;
          size_t di = fy * final_len + fx / (size_t)ppb;
          int d_shift = (ppb - 1 - (int)(fx % (size_t)ppb)) * bit_depth;
          out[di] = (uint8_t)((out[di] & ~(mask << d_shift)) | (v << d_shift));
        }
      } else { /* distributePassPixels :130-152 */
        for (px = 0; px < pw; px++) {
          size_t fx = (size_t)A7_XS[pass] + px * (size_t)A7_DX[pass];
          memcpy(out + fy * final_len + fx * (size_t)bpp, cur + px * (size_t)bpp, (size_t)bpp);
        }
      }
      prev = cur; /* :92: per-pass predictor chain */
      cur = (cur == line_a) ? line_b : line_a;
    }
  }
  free(line_a); free(line_b);
  return ORC_OK;
}

/* png-decoder.ts:63-229 */
int orc_decode_scanlines(const uint8_t *data, size_t len, uint32_t width, uint32_t height,
                         int bit_depth, int color_type, int interlace, uint8_t *out,
                         uint32_t *rows_decoded, int *err_filter) {
  int bpp = orc_bytes_per_pixel(bit_depth, color_type);
  size_t n = orc_scanline_bytes(width, bit_depth, color_type);
  uint32_t y;
  if (rows_decoded) *rows_decoded = 0;
  if (bpp < 0) return ORC_E_INVALID;
  if (interlace == 1) { /* :73-90 */
    int rc = orc_deinterlace_adam7(data, len, width, height, bit_depth, color_type, out,
                                   NULL, NULL, err_filter);
    if (rc == ORC_OK && rows_decoded) *rows_decoded = height;
    return rc;
  }
  /* :165-205: a line is consumed only when 1+n bytes are buffered */
  for (y = 0; y < height; y++) {
    const uint8_t *line = data + (size_t)y * (n + 1);
    int rc;
    if ((size_t)(y + 1) * (n + 1) > len) return ORC_E_TRUNCATED; /* :209-211 */
    rc = orc_unfilter_scanline(line[0], line + 1, y ? out + (size_t)(y - 1) * n : NULL, bpp, n,
                               out + (size_t)y * n);
    if (rc != ORC_OK) {
      if (err_filter) *err_filter = line[0];
      return rc;
    }
    if (rows_decoded) *rows_decoded = y + 1;
  }
  return ORC_OK;
}

/* ------------------------------------------------------------------ */
/* pixel-ops.ts                                                        */
/* ------------------------------------------------------------------ */

/* pixel-ops.ts:312-326: Math.round(value * toMax / fromMax) in doubles */
uint32_t orc_scale_sample(uint32_t value, int from_bits, int to_bits) {
  double from_max, to_max;
  if (from_bits == to_bits) return value;
  from_max = (double)((1u << from_bits) - 1u);
  to_max = (double)((1u << to_bits) - 1u);
  return (uint32_t)js_round(((double)value * to_max) / from_max);
}

static uint32_t be16(const uint8_t *p) { return ((uint32_t)p[0] << 8) | p[1]; }

/* Extension (see oracle.h): colour type 3 -> RGBA8 per the PNG specification. */
void orc_expand_palette_scanline(const uint8_t *src, uint32_t width, int bit_depth, const uint8_t *plte,
                                 size_t plte_len, const uint8_t *trns, size_t trns_len, uint8_t *dst) {
  const size_t entries = plte_len / 3;
  const uint32_t mask = (1u << bit_depth) - 1u;
  uint32_t x;
  for (x = 0; x < width; x++) {
    const size_t bit = (size_t)x * (size_t)bit_depth;
    const uint32_t idx = (src[bit >> 3] >> (8 - (int)(bit & 7) - bit_depth)) & mask;
    uint8_t *o = dst + (size_t)x * 4;
    if (idx < entries) {
      o[0] = plte[idx * 3]; o[1] = plte[idx * 3 + 1]; o[2] = plte[idx * 3 + 2];
    } else {
      o[0] = o[1] = o[2] = 0;
    }
    o[3] = idx < trns_len ? trns[idx] : 255;
  }
}

/* pixel-ops.ts:496-633 */
int orc_convert_scanline(const uint8_t *src, uint32_t width, int sbd, int sct,
                         int tbd, int tct, uint8_t *dst) {
  uint32_t x;
  size_t tbpp = tbd == 16 ? 8 : 4;
  if (sbd == tbd && sct == tct) { /* :505-507 */
    memcpy(dst, src, orc_scanline_bytes(width, sbd, sct));
    return ORC_OK;
  }
  if (tct != 6) return ORC_E_TARGET; /* :510-512 */
  if (!(sct == 0 || sct == 2 || sct == 4 || sct == 6)) return ORC_E_COLOR; /* :609-611 */
  for (x = 0; x < width; x++) {
    uint32_t r = 0, g = 0, b = 0, a = 255;
    if (sct == 0) { /* :522-540 */
      uint32_t gray;
      if (sbd == 16) {
        gray = orc_scale_sample(be16(src + (size_t)x * 2), 16, tbd);
      } else if (sbd == 8) {
        gray = orc_scale_sample(src[x], 8, tbd);
      } else {
        size_t byte_off = ((size_t)x * (size_t)sbd) / 8;
        int bit_off = (int)(((size_t)x * (size_t)sbd) % 8);
        uint32_t mask = (1u << sbd) - 1u;
        gray = orc_scale_sample((src[byte_off] >> (8 - bit_off - sbd)) & mask, sbd, tbd);
      }
      r = g = b = gray;
      a = tbd == 16 ? 0xFFFFu : 0xFFu;
    } else if (sct == 2) { /* :541-565 */
      if (sbd == 16) {
        const uint8_t *p = src + (size_t)x * 6;
        r = be16(p); g = be16(p + 2); b = be16(p + 4);
        if (tbd == 8) {
          r = orc_scale_sample(r, 16, 8); g = orc_scale_sample(g, 16, 8); b = orc_scale_sample(b, 16, 8);
        }
      } else {
        const uint8_t *p = src + (size_t)x * 3;
        r = p[0]; g = p[1]; b = p[2];
        if (tbd == 16) {
          r = orc_scale_sample(r, 8, 16); g = orc_scale_sample(g, 8, 16); b = orc_scale_sample(b, 8, 16);
        }
      }
      a = tbd == 16 ? 0xFFFFu : 0xFFu;
    } else if (sct == 4) { /* :566-582 */
      if (sbd == 16) {
        const uint8_t *p = src + (size_t)x * 4;
        uint32_t gray = be16(p);
        r = g = b = tbd == 16 ? gray : orc_scale_sample(gray, 16, 8);
        a = be16(p + 2);
        if (tbd == 8) a = orc_scale_sample(a, 16, 8);
      } else {
        const uint8_t *p = src + (size_t)x * 2;
        r = g = b = tbd == 16 ? orc_scale_sample(p[0], 8, 16) : p[0];
        a = tbd == 16 ? orc_scale_sample(p[1], 8, 16) : p[1];
      }
    } else { /* sct == 6, :583-608 */
      if (sbd == 16) {
        const uint8_t *p = src + (size_t)x * 8;
        r = be16(p); g = be16(p + 2); b = be16(p + 4); a = be16(p + 6);
        if (tbd == 8) {
          r = orc_scale_sample(r, 16, 8); g = orc_scale_sample(g, 16, 8);
          b = orc_scale_sample(b, 16, 8); a = orc_scale_sample(a, 16, 8);
        }
      } else {
        const uint8_t *p = src + (size_t)x * 4;
        r = p[0]; g = p[1]; b = p[2]; a = p[3];
        if (tbd == 16) {
          r = orc_scale_sample(r, 8, 16); g = orc_scale_sample(g, 8, 16);
          b = orc_scale_sample(b, 8, 16); a = orc_scale_sample(a, 8, 16);
        }
      }
    }
    { /* :614-629 */
      uint8_t *q = dst + (size_t)x * tbpp;
      if (tbd == 16) {
        q[0] = (uint8_t)(r >> 8); q[1] = (uint8_t)r; q[2] = (uint8_t)(g >> 8); q[3] = (uint8_t)g;
        q[4] = (uint8_t)(b >> 8); q[5] = (uint8_t)b; q[6] = (uint8_t)(a >> 8); q[7] = (uint8_t)a;
      } else {
        q[0] = (uint8_t)r; q[1] = (uint8_t)g; q[2] = (uint8_t)b; q[3] = (uint8_t)a;
      }
    }
  }
  return ORC_OK;
}

/* pixel-ops.ts:646-736.  All arithmetic in IEEE doubles in the reference's
 * association order; compile with -ffp-contract=off so no FMA is formed. */
void orc_composite_scanline(uint8_t *dest, const uint8_t *source, uint32_t start_x,
                            uint32_t source_width, int bytes_per_pixel, int use_alpha_blending) {
  size_t bpp = (size_t)bytes_per_pixel;
  size_t offset = (size_t)start_x * bpp;
  uint32_t px;
  int c;
  if (!use_alpha_blending) { /* :657-661 */
    memcpy(dest + offset, source, (size_t)source_width * bpp);
    return;
  }
  for (px = 0; px < source_width; px++) {
    uint8_t *d = dest + offset + (size_t)px * bpp;
    const uint8_t *s = source + (size_t)px * bpp;
    if (bytes_per_pixel == 8) { /* :668-703 */
      double sa = (double)be16(s + 6) / 65535.0;
      if (sa >= 0.9999) {
        memcpy(d, s, 8);
      } else if (sa > 0.0001) {
        double da = (double)be16(d + 6) / 65535.0;
        double oa = sa + da * (1.0 - sa);
        if (oa > 0.0001) {
          for (c = 0; c < 3; c++) {
            double sv = (double)be16(s + c * 2), dv = (double)be16(d + c * 2);
            double blended = (sv * sa + dv * da * (1.0 - sa)) / oa;
            double v = blended < 65535.0 ? blended : 65535.0; /* Math.min(65535, blended) */
            uint32_t val;
            v = v > 0.0 ? v : 0.0;                            /* Math.max(0, ...) */
            val = (uint32_t)js_round(v);
            d[c * 2] = (uint8_t)(val >> 8);
            d[c * 2 + 1] = (uint8_t)(val & 0xff);
          }
          {
            uint32_t av = (uint32_t)js_round(oa * 65535.0);
            d[6] = (uint8_t)(av >> 8);
            d[7] = (uint8_t)(av & 0xff);
          }
        }
      }
    } else { /* :704-733 */
      double sa = (double)s[3] / 255.0;
      if (sa >= 0.9999) {
        memcpy(d, s, 4);
      } else if (sa > 0.0001) {
        double da = (double)d[3] / 255.0;
        double oa = sa + da * (1.0 - sa);
        if (oa > 0.0001) {
          for (c = 0; c < 3; c++) {
            double sv = (double)s[c], dv = (double)d[c];
            double blended = (sv * sa + dv * da * (1.0 - sa)) / oa;
            double v = blended < 255.0 ? blended : 255.0;
            v = v > 0.0 ? v : 0.0;
            d[c] = (uint8_t)js_round(v);
          }
          d[3] = (uint8_t)js_round(oa * 255.0);
        }
      }
    }
  }
}

/* ------------------------------------------------------------------ */
/* image-concat-core.ts row assembly                                   */
/* ------------------------------------------------------------------ */

static int tile_order(const void *pa, const void *pb) {
  const orc_tile *a = *(const orc_tile *const *)pa, *b = *(const orc_tile *const *)pb;
  /* positioned-layout.ts:229-234 */
  if (a->z_index != b->z_index) return a->z_index < b->z_index ? -1 : 1;
  if (a->image_idx != b->image_idx) return a->image_idx < b->image_idx ? -1 : 1;
  return 0;
}

/* Builds canvas row y into `row` (W*bpp bytes). `conv` is scratch for the
 * widest tile row in output format. */
static int compose_row(uint32_t canvas_w, int out_bit_depth, const uint8_t background[8],
                       int alpha_blend, const orc_tile **sorted, size_t n_tiles,
                       uint32_t y, uint8_t *row, uint8_t *conv) {
  size_t bpp = out_bit_depth == 16 ? 8 : 4;
  uint32_t x;
  size_t k;
  /* createTransparentScanline image-concat-core.ts:96-102 */
  for (x = 0; x < canvas_w; x++) memcpy(row + (size_t)x * bpp, background, bpp);
  for (k = 0; k < n_tiles; k++) {
    const orc_tile *t = sorted[k];
    uint32_t local_y, src_y;
    size_t src_row_bytes;
    int rc;
    if (t->vis_w == 0 || t->vis_h == 0) continue; /* fully clipped tiles are dropped */
    if ((int64_t)y < t->dst_y || (int64_t)y >= (int64_t)t->dst_y + t->vis_h) continue;
    local_y = (uint32_t)((int64_t)y - t->dst_y);
    src_y = local_y + t->src_off_y; /* :588-589 */
    src_row_bytes = orc_scanline_bytes(t->width, (int)t->bit_depth, (int)t->color_type);
    rc = orc_convert_scanline(t->pixels + (size_t)src_y * src_row_bytes, t->width, (int)t->bit_depth,
                              (int)t->color_type, out_bit_depth, 6, conv); /* :629-636 */
    if (rc != ORC_OK) return rc;
    /* extractScanlinePortion pixel-ops.ts:747-756 + compositeScanline */
    orc_composite_scanline(row, conv + (size_t)t->src_off_x * bpp, (uint32_t)t->dst_x, t->vis_w,
                           (int)bpp, alpha_blend);
  }
  return ORC_OK;
}

int64_t orc_render_rows(uint32_t canvas_w, uint32_t canvas_h, int out_bit_depth,
                        const uint8_t background[8], int alpha_blend,
                        const orc_tile *tiles, size_t n_tiles,
                        uint32_t y0, uint32_t y1, int out_kind, uint8_t *out, size_t cap) {
  size_t bpp = out_bit_depth == 16 ? 8 : 4;
  size_t row_bytes = (size_t)canvas_w * bpp;
  size_t out_row = row_bytes + (out_kind == 0 ? 1 : 0);
  size_t max_conv = row_bytes, k;
  const orc_tile **sorted;
  uint8_t *cur, *prev, *conv, *scratch;
  uint32_t y;
  int64_t written = 0;
  int rc = ORC_OK;
  if (y1 > canvas_h || y0 > y1) return ORC_E_INVALID;
  if ((size_t)(y1 - y0) * out_row > cap) return ORC_E_INVALID;
  for (k = 0; k < n_tiles; k++) {
    size_t need = (size_t)tiles[k].width * bpp;
    size_t need_src = orc_scanline_bytes(tiles[k].width, (int)tiles[k].bit_depth, (int)tiles[k].color_type);
    if (need > max_conv) max_conv = need;
    if (need_src > max_conv) max_conv = need_src;
  }
  sorted = (const orc_tile **)malloc((n_tiles + 1) * sizeof(*sorted));
  cur = (uint8_t *)malloc(row_bytes + 8);
  prev = (uint8_t *)malloc(row_bytes + 8);
  conv = (uint8_t *)malloc(max_conv + 8);
  scratch = (uint8_t *)malloc(4 * row_bytes + 8);
  if (!sorted || !cur || !prev || !conv || !scratch) { rc = ORC_E_INVALID; goto done; }
  for (k = 0; k < n_tiles; k++) sorted[k] = &tiles[k];
  qsort(sorted, n_tiles, sizeof(*sorted), tile_order);

  if (out_kind == 0 && y0 > 0) { /* predictor row for a band that does not start at 0 */
    rc = compose_row(canvas_w, out_bit_depth, background, alpha_blend, sorted, n_tiles, y0 - 1, prev, conv);
    if (rc != ORC_OK) goto done;
  }
  for (y = y0; y < y1; y++) {
    uint8_t *dst = out + (size_t)(y - y0) * out_row;
    rc = compose_row(canvas_w, out_bit_depth, background, alpha_blend, sorted, n_tiles, y, cur, conv);
    if (rc != ORC_OK) goto done;
    if (out_kind == 0) {
      /* filterScanline + filter byte, image-concat-core.ts:527-542, :1276-1303 */
      int ft = orc_filter_scanline(cur, y > 0 ? prev : NULL, (int)bpp, row_bytes, dst + 1, scratch, NULL);
      uint8_t *tmp;
      dst[0] = (uint8_t)ft;
      tmp = prev; prev = cur; cur = tmp;
    } else {
      memcpy(dst, cur, row_bytes);
    }
    written += (int64_t)out_row;
  }
done:
  free(sorted); free(cur); free(prev); free(conv); free(scratch);
  return rc == ORC_OK ? written : rc;
}
