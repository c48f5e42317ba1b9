/*
 * oracle.h -- CPU restatement of image-stitch's pixel pipeline.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it, and only as the checker or the timed
 * CPU baseline.  The product path (image_stitch_b200/) never links, imports
 * or executes this code.
 *
 * Every function cites the reference file:line (relative to the
 * jburnhams/image-stitch tree, v1.1.53) whose behaviour it restates.
 *
 * Parity status: the PNG half is pinned against the reference's own golden
 * outputs (tests/utils/fixtures/expected-outputs/example1..5.png), its unit
 * test known answers and the PngSuite JSON goldens -- see
 * tests/test_oracle_golden.py.  The JPEG half (oracle/jpeg_oracle.c) is
 * "parity unpinned": the reference delegates to the un-vendored
 * jpeg-encoder-wasm 0.1.0 package, see that file's header.
 */
#ifndef STITCH_ORACLE_H
#define STITCH_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* error codes shared by the oracle entry points */
#define ORC_OK 0
#define ORC_E_FILTER (-4)    /* "Unknown filter type: N"            png-filter.ts:78-79 */
#define ORC_E_COLOR (-5)     /* "Unsupported source color type: N"  pixel-ops.ts:609-611 */
#define ORC_E_TRUNCATED (-6) /* adam7.ts:80-82, png-decoder.ts:209-211 */
#define ORC_E_TARGET (-9)    /* "Only conversion to RGBA ..."       pixel-ops.ts:510-512 */
#define ORC_E_INVALID (-1)

/* png-filter.ts:16-25 */
int orc_paeth_predictor(int a, int b, int c);

/* png-filter.ts:186-211; returns -1 for an unknown colour type */
int orc_bytes_per_pixel(int bit_depth, int color_type);
/* utils.ts:92-101; returns -1 for an unknown colour type */
int orc_samples_per_pixel(int color_type);
/* ceil(width*bitDepth*spp/8): png-decoder.ts:68-70, image-concat-core.ts:435-438 */
size_t orc_scanline_bytes(uint32_t width, int bit_depth, int color_type);

/* png-filter.ts:34-83.  prev may be NULL (first line).  out must not alias in.
 * Returns ORC_E_FILTER for filter_type > 4. */
int orc_unfilter_scanline(int filter_type, const uint8_t *in, const uint8_t *prev,
                          int bpp, size_t n, uint8_t *out);

/* png-filter.ts:148-181 (+ 88-142).  Writes the n filtered bytes of the chosen
 * candidate into out and returns the filter type 0..4.  scratch must hold 4*n
 * bytes (the four non-identity candidates, as the reference materialises them).
 * sums (optional, 5 entries) receives the five heuristic costs. */
int orc_filter_scanline(const uint8_t *row, const uint8_t *prev, int bpp, size_t n,
                        uint8_t *out, uint8_t *scratch, uint64_t *sums);

/* adam7.ts:52-110 (+115-184).  out must hold height*scanline_bytes and is
 * zero-filled first like `new Uint8Array`.  On truncation returns
 * ORC_E_TRUNCATED and sets err_pass (1-based) / err_line like the thrown
 * message "Unexpected end of decompressed data at pass P, line L". */
int orc_deinterlace_adam7(const uint8_t *data, size_t len, uint32_t width, uint32_t height,
                          int bit_depth, int color_type, uint8_t *out,
                          int *err_pass, int *err_line, int *err_filter);

/* png-decoder.ts:63-229: inflated IDAT stream -> dense unfiltered rows.
 * interlace=1 goes through orc_deinterlace_adam7.  rows_decoded receives the
 * number of scanlines produced before an error. */
int orc_decode_scanlines(const uint8_t *data, size_t len, uint32_t width, uint32_t height,
                         int bit_depth, int color_type, int interlace, uint8_t *out,
                         uint32_t *rows_decoded, int *err_filter);

/* pixel-ops.ts:312-326 */
uint32_t orc_scale_sample(uint32_t value, int from_bits, int to_bits);

/* pixel-ops.ts:496-633.  dst holds width*4 (8-bit) or width*8 (16-bit) bytes.
 * The identity case copies (the reference returns the same array). */
int orc_convert_scanline(const uint8_t *src, uint32_t width, int src_bit_depth, int src_color_type,
                         int tgt_bit_depth, int tgt_color_type, uint8_t *dst);

/* EXTENSION, beyond the reference: the reference throws `Unsupported source color type: 3`
 * (pixel-ops.ts:609-611); this expands one row of palette indices (bit depth 1, 2, 4 or 8, MSB first like
 * pixel-ops.ts:540-552 reads sub-byte gray) to RGBA8 the way the PNG specification defines colour type 3:
 * RGB from PLTE entry i, alpha from tRNS entry i where present, else 255; an index past the end of PLTE gives
 * opaque black.  Not a parity claim: pinned by the third-party pngsuite/json goldens of the 63 palette files. */
void orc_expand_palette_scanline(const uint8_t *src, uint32_t width, int bit_depth, const uint8_t *plte,
                                 size_t plte_len, const uint8_t *trns, size_t trns_len, uint8_t *dst);

/* pixel-ops.ts:646-736 */
void orc_composite_scanline(uint8_t *dest, const uint8_t *source, uint32_t start_x,
                            uint32_t source_width, int bytes_per_pixel, int use_alpha_blending);

/* One placed input of a stitch job; the output of the host's layout stage:
 * grid cells (image-concat-core.ts:132-261) or clipped positioned images
 * (positioned-layout.ts:107-190). */
typedef struct orc_tile {
  uint32_t width, height;        /* full source image size */
  uint32_t bit_depth, color_type;
  const uint8_t *pixels;         /* dense unfiltered rows in SOURCE format */
  int32_t dst_x, dst_y;          /* clipped top-left on the canvas */
  uint32_t src_off_x, src_off_y; /* ClippedImageInfo.sourceOffsetX/Y */
  uint32_t vis_w, vis_h;         /* visible size */
  int32_t z_index;
  uint32_t image_idx;
} orc_tile;

/* Restates the per-row loops of generateFilteredScanlines
 * (image-concat-core.ts:389-545), generatePositionedScanlines (:551-685) and
 * the filter step of streamPositionedPngData (:1276-1303) for rows [y0,y1).
 *  - every row starts as `background` (createTransparentScanline :96-102,
 *    padScanline :107-127, right pad :515-525)
 *  - tiles are applied in (z_index, image_idx) order (positioned-layout.ts:229-234)
 *    through convertScanline -> extractScanlinePortion -> compositeScanline
 *  - alpha_blend=0 is `dest.set` (grid mode never overlaps, so this is also
 *    combineScanlines :76-91)
 *  - out_kind 0: rows of [filter byte + W*bpp] with the previous RAW output
 *    row as predictor (row 0: none); out_kind 1: raw rows W*bpp.
 * The filter predictor for y0>0 is recomputed from row y0-1.
 * Returns bytes written or a negative error. */
int64_t orc_render_rows(uint32_t canvas_w, uint32_t canvas_h, int out_bit_depth,
                        const uint8_t background[8], int alpha_blend,
                        const orc_tile *tiles, size_t n_tiles,
                        uint32_t y0, uint32_t y1, int out_kind, uint8_t *out, size_t cap);

/* ---- JPEG half (jpeg_oracle.c; parity unpinned against the reference, pinned against libjpeg) ---- */
void orc_jpeg_quant_tables(int quality, uint16_t lum[64], uint16_t chr[64]);
void orc_jpeg_rgb_to_ycc(int r, int g, int b, int *y, int *cb, int *cr);
void orc_jpeg_fdct_islow(int data[64]);
int orc_jpeg_quantize(int coef, int quantval);
int64_t orc_jpeg_coefficients(const uint8_t *rgba, size_t pitch, uint32_t width, uint32_t height, uint32_t y0,
                              uint32_t y1, int quality, int16_t *coefs);
int64_t orc_jpeg_encode_rows(const uint8_t *rgba, size_t pitch, uint32_t width, uint32_t height, uint32_t y0,
                             uint32_t y1, int quality, uint8_t *out, size_t cap);
int64_t orc_jpeg_header(uint32_t width, uint32_t height, int quality, uint8_t *out, size_t cap);

#ifdef __cplusplus
}
#endif
#endif
