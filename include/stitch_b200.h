/*
 * stitch_b200.h -- C ABI of libstitch_b200.so, the B200-native replacement for
 * image-stitch's pixel pipeline (inflated scanlines in, filtered PNG scanlines
 * / raw RGBA rows / entropy-coded JPEG MCUs out).
 *
 * The reference (jburnhams/image-stitch v1.1.53, TypeScript) has no FFI seam
 * around this path: the functions below replace ordinary TS functions that the
 * row generators call inline.  Each entry point names what it replaces
 * (paths relative to the reference tree).  INTEGRATION.md shows the N-API
 * addon and the Python/ctypes binding over this header.
 *
 * Conventions: every function returns STITCH_OK (0) or a negative
 * stitch_status; stitch_last_error(ctx) returns the reference's own message
 * text for that failure ("Unknown filter type: 9", "Unsupported source color
 * type: 3", ...).  All pointers are plain host pointers unless the name says
 * "device".  No torch / C++ types cross this boundary.
 *
 * Threading: calls on one ctx are serialised by the caller.  One ctx drives
 * one CUDA device (one process per GPU; row bands of one canvas are spread
 * over ranks by the host).
 */
#ifndef STITCH_B200_H
#define STITCH_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define STITCH_ABI_VERSION 1

typedef enum stitch_status {
  STITCH_OK = 0,
  STITCH_E_INVALID = -1,   /* bad argument / descriptor */
  STITCH_E_CUDA = -2,      /* CUDA runtime failure (message carries cudaGetErrorString) */
  STITCH_E_NOMEM = -3,
  STITCH_E_FILTER = -4,    /* "Unknown filter type: N"            src/png-filter.ts:78-79 */
  STITCH_E_COLOR = -5,     /* "Unsupported source color type: N"  src/pixel-ops.ts:609-611 */
  STITCH_E_TRUNCATED = -6, /* "Expected H scanlines, decoded N" src/decoders/png-decoder.ts:209-211,
                              "Unexpected end of decompressed data at pass P, line L" src/adam7.ts:80-82 */
  STITCH_E_STATE = -7,     /* call order violated */
  STITCH_E_CAPACITY = -8,  /* caller buffer too small */
  STITCH_E_INFLATE = -9,   /* "Inflate error: <zlib message>"   src/streaming-inflate.ts:60-62, :69-71 */
  STITCH_E_NODEVICE = -10, /* no CUDA device: there is no CPU fallback */
  STITCH_E_JPEG = -11      /* "Invalid JPEG: ..." src/decoders/jpeg-decoder.ts:49,74,93; "Unsupported JPEG: ..." for what the
                              device decoder does not take (progressive, 12-bit, CMYK, unusual sampling) */
} stitch_status;

typedef enum stitch_out_kind {
  STITCH_OUT_PNG_FILTERED = 0, /* rows of [filter byte + W*bpp]: filterScanline src/png-filter.ts:148-181 */
  STITCH_OUT_RAW_RGBA = 1,     /* composited rows W*bpp: generateRawScanlines src/image-concat-core.ts:691-831 */
  STITCH_OUT_JPEG_ENTROPY = 2, /* entropy-coded MCU rows: StreamingJpegEncoder.encode_strip, src/jpeg-encoder.ts:155-166 */
  STITCH_OUT_PNG_DEFLATE = 3   /* the filtered rows of kind 0, DEFLATE-compressed on the device: what
                                  StreamingDeflator.push/flush emit (src/streaming-deflate.ts:130-185), as
                                  byte-aligned non-final blocks; inflates to the same rows, bytes differ from pako's */
} stitch_out_kind;

typedef struct stitch_ctx stitch_ctx;
typedef struct stitch_job stitch_job;

/* One stitch job = one output canvas. Replaces the per-call state of
 * CoreStreamingConcatenator.streamGridMode / streamPositionedMode
 * (src/image-concat-core.ts:1008-1217). */
typedef struct stitch_job_desc {
  uint32_t canvas_w, canvas_h;
  uint32_t out_bit_depth;  /* 8 | 16: determineCommonFormat src/pixel-ops.ts:293-307; forced 8 for JPEG */
  uint32_t out_kind;       /* stitch_out_kind */
  uint32_t alpha_blend;    /* 1: alpha-over (positioned default); 0: dest.set (grid mode, enableAlphaBlending:false) */
  uint32_t jpeg_quality;   /* 1..100, JPEG only (src/jpeg-encoder.ts:109-114) */
  uint8_t background[8];   /* one pixel in OUTPUT format: getTransparentColor src/pixel-ops.ts:255-287 */
  uint32_t band_y0, band_y1; /* rows this job will ever be asked for (multi-GPU row band); 0,0 = whole canvas */
  uint32_t reserved[4];
} stitch_job_desc;

/* One placed input. Produced by the host from calculateLayout
 * (src/image-concat-core.ts:132-204) or clipImagesToCanvas
 * (src/positioned-layout.ts:107-190). */
typedef struct stitch_tile_desc {
  uint32_t width, height;        /* source image size */
  uint32_t bit_depth;            /* 1,2,4,8,16 */
  uint32_t color_type;           /* 0,2,4,6 (3 = palette raises STITCH_E_COLOR like the reference) */
  uint32_t interlace;            /* 1: Adam7 stream, src/adam7.ts:52-110 */
  uint32_t filtered;             /* 1: data = inflated IDAT stream with filter bytes (src/decoders/png-decoder.ts:63-229)
                                    0: data = dense unfiltered rows (foreign ImageDecoder.scanlines())
                                    2: data = the zlib stream itself (the concatenated IDAT payloads,
                                       src/png-parser.ts / png-decoder.ts:100-117 before createDecompressionStream):
                                       inflated on the device, then treated like 1 */
  int32_t dst_x, dst_y;          /* clipped top-left on the canvas, >= 0 */
  uint32_t src_off_x, src_off_y; /* ClippedImageInfo.sourceOffsetX/Y */
  uint32_t vis_w, vis_h;         /* visible size after clipping */
  int32_t z_index;               /* composite order key 1 (src/positioned-layout.ts:229-234) */
  uint32_t image_idx;            /* composite order key 2, and the "#N" of error messages */
} stitch_tile_desc;

/* Per-call device timings (CUDA events on the ctx stream), milliseconds. */
typedef struct stitch_timings {
  float h2d_ms;       /* tile uploads still pending when the call started */
  float decode_ms;    /* kernel (a): unfilter + Adam7 */
  float compose_ms;   /* kernels (b)+(c) fused, or (b) then (c), or (b)+(d) */
  float d2h_ms;       /* band download */
  float total_ms;
  uint32_t launches;  /* kernels launched by this call */
  uint32_t reserved;
} stitch_timings;

/* JPEG band result header (STITCH_OUT_JPEG_ENTROPY).  The restart interval is
 * one MCU row (the DRI segment of stitch_jpeg_header), so a band's payload is a
 * byte-aligned run of complete entropy segments, each followed by its RSTm
 * marker (none after the image's last MCU row): bands are spliced by plain
 * concatenation in ascending row order, then EOI. */
/* DEFLATE band result header (STITCH_OUT_PNG_DEFLATE), followed by n_bytes of complete, byte-aligned,
 * non-final deflate blocks.  The host writes the zlib header (78 01), the bands in ascending row order, the
 * final empty block (01 00 00 FF FF) and the big-endian Adler-32 of all raw rows, combining the per-band
 * values with adler32_combine (RFC 1950).  crc32 is the CRC-32 (zlib's crc32) of the n_bytes compressed bytes,
 * computed on the device, so that the IDAT chunk's check value (src/png-writer.ts:12-30, src/utils.ts:18-24:
 * CRC over the chunk type and data) is crc32_combine(crc32("IDAT" + prefix), crc32, n_bytes) without a host
 * pass over the data. */
typedef struct stitch_deflate_band_info {
  uint64_t n_bytes;   /* compressed bytes that follow */
  uint64_t n_raw;     /* raw bytes they inflate to: (y1-y0) * (1 + W*bpp) */
  uint32_t adler;     /* Adler-32 of those raw bytes alone */
  uint32_t crc32;     /* CRC-32 of the n_bytes compressed bytes alone */
} stitch_deflate_band_info;

typedef struct stitch_jpeg_band_info {
  uint64_t n_bits;     /* 8 * payload bytes that follow this header (stuffing and RST markers included) */
  int32_t last_dc[3];  /* always 0: every restart segment starts from zero DC predictors */
  uint32_t reserved;
} stitch_jpeg_band_info;

/* ---- context ------------------------------------------------------------ */

/* device_id: CUDA ordinal.  Fails with STITCH_E_NODEVICE when no GPU is
 * visible -- by design there is no CPU path behind this ABI. */
int stitch_ctx_create(int device_id, stitch_ctx **out);
void stitch_ctx_destroy(stitch_ctx *ctx);
/* Use an existing cudaStream_t (e.g. torch's current stream) for all work. */
int stitch_ctx_set_stream(stitch_ctx *ctx, void *cuda_stream);
const char *stitch_last_error(const stitch_ctx *ctx);
int stitch_abi_version(void);

/* ---- job ---------------------------------------------------------------- */

int stitch_job_create(stitch_ctx *ctx, const stitch_job_desc *desc, stitch_job **out);
void stitch_job_destroy(stitch_job *job);

/* Registers one input and starts its upload on the context's upload stream
 * (kernels of earlier bands and downloads keep running meanwhile).  Pageable
 * `data` is borrowed until the call returns; page-locked `data` (cudaHostAlloc /
 * cudaHostRegister) is read asynchronously and must stay valid until the next
 * stitch_job_decode / stitch_job_run_band that needs the tile returns.  Validates what
 * the reference validates while reading: colour type (pixel-ops.ts:609-611),
 * stream length (png-decoder.ts:209-211, adam7.ts:79-82); filter bytes are
 * validated on the device by stitch_job_decode. */
int stitch_job_add_tile(stitch_job *job, const stitch_tile_desc *tile, const uint8_t *data, size_t len);

/* EXTENSION, beyond the reference: a colour type 3 tile together with the bodies of its PLTE
 * (3 bytes per entry, 1..256 entries) and tRNS (1 alpha byte per entry, may be NULL/0) chunks.  The reference
 * has no such path -- convertScanline throws `Unsupported source color type: 3` (pixel-ops.ts:609-611), which
 * is what stitch_job_add_tile keeps returning for colour type 3.  Here the index plane (bit depth 1, 2, 4, 8;
 * filtered 0, 1 or 2; interlaced or not) is decoded by kernel (a) and expanded to RGBA8 on the device as the PNG
 * specification defines: RGB = PLTE[i], A = tRNS[i] or 255; an index past the end of PLTE gives opaque black.
 * Checked against the third-party pngsuite/json goldens, not against the reference. */
int stitch_job_add_tile_palette(stitch_job *job, const stitch_tile_desc *tile, const uint8_t *data, size_t len,
                                const uint8_t *plte, size_t plte_len, const uint8_t *trns, size_t trns_len);

/* Same, but `device_data` already lives in this ctx's device memory (16-byte
 * aligned, stays valid and unmodified until the job is destroyed). */
int stitch_job_add_tile_device(stitch_job *job, const stitch_tile_desc *tile, const void *device_data, size_t len);

/* Kernel (a): unfilterScanline (src/png-filter.ts:34-83) over every row of
 * every pending tile + deinterlaceAdam7 (src/adam7.ts:52-110).  Idempotent
 * unless `force` is set (bench: re-run with inputs resident in HBM). */
int stitch_job_decode(stitch_job *job, int force, stitch_timings *t);

/* SURVEY 8f.2, the step before kernel (a): inflates every tile that was added as a zlib stream
 * (filtered == 2) and is not inflated yet -- all of them again with `force` -- one warp per stream, and checks
 * the Adler-32 trailers.  What createDecompressionStream('deflate') / pako.Inflate do on the host
 * (src/streaming-inflate.ts:23-76).  stitch_job_decode / stitch_job_run_band call it on their own for the
 * tiles they need; this entry point exists for timing the step alone.  Failures: STITCH_E_INFLATE with
 * zlib's message, STITCH_E_TRUNCATED when the stream holds fewer rows than the header promises. */
int stitch_job_inflate(stitch_job *job, int force, stitch_timings *t);

/* Forgets that the job's tiles were decoded (no device work): the next
 * stitch_job_run_band unfilters again what its rows need, in the same stream
 * submission as the composite -- one synchronisation per band instead of two.
 * For repeated measurements with the inputs resident in HBM. */
int stitch_job_invalidate(stitch_job *job);

/* Kernels (b)+(c) (PNG), (b) (RAW) or (b)+(d) (JPEG) for canvas rows
 * [y0,y1): convertScanline (src/pixel-ops.ts:496-633), row assembly
 * (src/image-concat-core.ts:404-545, :551-685), compositeScanline
 * (src/pixel-ops.ts:646-736), filterScanline (src/png-filter.ts:148-181).
 * Runs stitch_job_decode first when needed.
 *  out != NULL: the band is copied to host memory `out` (cap bytes):
 *      PNG  (y1-y0) rows of [filter byte + W*bpp]
 *      RAW  (y1-y0) rows of W*bpp
 *      JPEG stitch_jpeg_band_info followed by the band's entropy bytes
 *           (y0 % 8 == 0; y1 % 8 == 0 or y1 == canvas_h)
 *      DEFLATE stitch_deflate_band_info followed by the band's deflate blocks
 *           (cap: at most (y1-y0)*(1+W*bpp) * 1.001 + 1 KiB)
 *  out == NULL: the result stays on the device (stitch_job_band_device).
 * The PNG predictor row for y0 > 0 is recomputed on the device (1-row halo).
 * With a host destination a PNG / RAW band is produced in chunks of about
 * 64 MB, each downloaded on its own stream while the next one is computed and
 * later tiles are still uploading; only the tiles a chunk touches are decoded
 * before it.  The call returns when `out` is complete.  In that case
 * t->decode_ms covers the first chunk's decode only. */
int stitch_job_run_band(stitch_job *job, uint32_t y0, uint32_t y1, uint8_t *out, size_t cap,
                        size_t *out_len, stitch_timings *t);

/* Device address / pitch of the last computed band (row r of the band starts
 * at ptr + r*pitch; for PNG the filter byte is at ptr + r*pitch + lead). */
int stitch_job_band_device(stitch_job *job, void **ptr, size_t *pitch, size_t *lead, size_t *row_bytes);

/* First decode error seen on the device, in the reference's wording. */
int stitch_job_status(stitch_job *job);

/* ---- 8(f).4 extension: baseline JPEG tiles, decoded on the device ---------------------------------------------
 * The reference decodes JPEG inputs on the host through sharp / jpeg-js / the browser (src/decoders/jpeg-decoder.ts:
 * 196-281) and feeds RGBA8 (or gray8) rows to the same hot path; here the file crosses the bus as it is and
 * libjpeg-turbo's default decode (what sharp gives on Node) is redone on the GPU: Huffman decoding per restart
 * interval, jpeg_idct_islow, fancy chroma upsampling, fixed-point YCbCr -> RGB (csrc/jpeg_decode.cu).
 * Accepted: SOF0 / SOF1, 8-bit, one interleaved scan, gray or YCbCr, luma sampling 1x1, 2x1, 2x2 with 1x1 chroma;
 * anything else fails with STITCH_E_JPEG and a message that says what. */

/* Dimensions and channel count (1 gray, 4 RGBA) without decoding: parseJpegHeader + getHeader,
 * src/decoders/jpeg-decoder.ts:46-94, :298-315, same messages.  ctx may be NULL (no message text then). */
int stitch_jpeg_info(stitch_ctx *ctx, const uint8_t *data, size_t len, uint32_t *width, uint32_t *height, uint32_t *channels);

/* Adds the JPEG file `data` as a tile.  Of `tile` the placement fields are used (dst_x/y, src_off_x/y, vis_w/h,
 * z_index, image_idx); width and height must be 0 or the frame's.  The file is parsed and queued for upload and
 * decoding before the call returns; decode errors of the entropy-coded data surface at the first run_band that
 * needs the tile (STITCH_E_JPEG).  At most 4096 JPEG tiles per job. */
int stitch_job_add_tile_jpeg(stitch_job *job, const stitch_tile_desc *tile, const uint8_t *data, size_t len);

/* The decoder alone: the whole image as dense rows, width * height * channels bytes (what JpegDecoder.scanlines()
 * yields row by row, src/decoders/jpeg-decoder.ts:317-347). */
int stitch_decode_jpeg(stitch_ctx *ctx, const uint8_t *data, size_t len, uint8_t *out, size_t cap, uint32_t *width,
                       uint32_t *height, uint32_t *channels);

/* ---- stand-alone pieces (unit-level parity) ------------------------------- */

/* Kernel (a) alone: inflated stream -> dense unfiltered rows in SOURCE format
 * (what ImageDecoder.scanlines() yields, src/decoders/types.ts:55-61). */
int stitch_decode_image(stitch_ctx *ctx, const stitch_tile_desc *tile, const uint8_t *data, size_t len,
                        uint8_t *out, size_t cap, size_t *out_len);

/* Inflate alone: one zlib stream (RFC 1950) -> at most cap bytes.  STITCH_E_CAPACITY when the stream holds
 * more (out is filled, *out_len == cap). */
int stitch_inflate(stitch_ctx *ctx, const uint8_t *zdata, size_t zlen, uint8_t *out, size_t cap, size_t *out_len);

/* Kernel (c) alone: filterScanline over `rows` rows of `row_bytes` bytes.
 * prev_row (may be NULL) is the raw row above rows[0]. out: rows*(1+row_bytes). */
int stitch_filter_rows(stitch_ctx *ctx, const uint8_t *rows, const uint8_t *prev_row, uint32_t n_rows,
                       size_t row_bytes, uint32_t bpp, uint8_t *out, size_t cap);

/* Kernel (c) again, for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same result as stitch_filter_rows, but d_rows / d_prev_row / d_out are DEVICE pointers the caller
 * owns (row r of the input at d_rows + r*in_pitch; of the output at d_out + r*out_pitch: filter-type byte, then
 * row_bytes filtered bytes; d_prev_row may be NULL), and the kernels' scratch -- five 64-bit cost sums per row --
 * is a static 1 MiB __device__ array inside the library, so the call performs no cudaMalloc/cudaFree/pool
 * allocation.  Limits: n_rows <= STITCH_NOALLOC_MAX_ROWS and row_bytes <= STITCH_NOALLOC_MAX_ROW_BYTES per call
 * (STITCH_E_CAPACITY beyond).  All calls on one GPU share the array: the call clears what it uses, runs on the
 * context's stream and returns after synchronising it; calls that overlap on the same GPU race. */
#define STITCH_NOALLOC_MAX_ROWS 16384u
#define STITCH_NOALLOC_MAX_ROW_BYTES (1u << 25)
int stitch_filter_rows_noalloc(stitch_ctx *ctx, const uint8_t *d_rows, size_t in_pitch, const uint8_t *d_prev_row,
                               uint32_t n_rows, size_t row_bytes, uint32_t bpp, uint8_t *d_out, size_t out_pitch);

/* Device-side proof of the alpha-over fast path (compositeScanline, src/pixel-ops.ts:646-736): compares the
 * library's blend with the literal f64 arithmetic of the reference ON THE DEVICE.  mode 0: 8-bit, every
 * (srcAlpha, dstAlpha, src, dst) -- 2^32 tuples (amount = k > 1 checks every k-th (sA, dA, s) only); mode 1:
 * 16-bit, `amount` hashed samples (0 = 2^28) including every alpha within 8 of the thresholds 65529 and 7.
 * first[4] receives one offending tuple.  stitch_selftest_blend8_samples returns n triples {dst, src, blended}
 * of hashed pixels for a host-side check against another implementation.  Test infrastructure. */
int stitch_selftest_blend(stitch_ctx *ctx, uint32_t mode, uint64_t amount, uint64_t *checked, uint64_t *mismatches,
                          uint32_t first[4]);
int stitch_selftest_blend8_samples(stitch_ctx *ctx, uint32_t n, uint32_t seed, uint32_t *out_triples);

/* ---- JPEG container helpers (host side of src/jpeg-encoder.ts) ------------ */

/* Writes SOI + JFIF APP0 + DQT + SOF0 + DHT + SOS for the encoder definition in
 * DESIGN.md (what the external encoder emits with its first strip). */
int stitch_jpeg_header(uint32_t width, uint32_t height, uint32_t quality, uint8_t *out, size_t cap, size_t *out_len);

#ifdef __cplusplus
}
#endif
#endif /* STITCH_B200_H */
