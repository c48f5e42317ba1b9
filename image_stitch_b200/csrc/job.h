// job.h -- private state shared by the translation units of libstitch_b200.so (not part of the C ABI)
#pragma once
#include "../../include/stitch_b200.h"
#include "common.cuh"
#include "kernels.h"

#include <string>
#include <utility>
#include <vector>

// ---------------------------------------------------------------------------
struct stitch_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  int sm_count = 148;
  std::string err;
  cudaEvent_t ev[6] = {};
  // Host<->device copies run beside the kernels: uploads of tile bytes on `up`, downloads of finished row
  // chunks on `down`; ordering is by events (one per tile, one per chunk).
  cudaStream_t up = nullptr, down = nullptr;
  cudaEvent_t ev_chunk = nullptr, ev_down = nullptr;
  // Small host tables (tile lists, chain tables, bins) reach the device through this pinned arena and a copy
  // KERNEL, not through cudaMemcpyAsync: the host-to-device copy engine serves its queue in order, and a job's
  // tile uploads (hundreds of MB, queued by add_tile) would hold a 1 KB table back until they are all through.
  uint8_t *h_arena = nullptr; size_t arena_cap = 0, arena_used = 0;
  uint32_t aux_launches = 0;   // copy / fill kernels launched since the current API call began (they count as launches)
  stitch::DevStatus *d_status = nullptr;   // device error word
  stitch::DevStatus *h_status = nullptr;   // pinned mirror
  // one pinned staging chunk kept between jobs (jpeg_decode.cu: pinning 32 MB costs about 10 ms)
  uint8_t *pinned_spare = nullptr; size_t pinned_spare_cap = 0;
  // JPEG tiles are decoded on a few streams in turn, so that the small kernels of several tiles overlap
  cudaStream_t jd_streams[4] = {};
  uint32_t jd_next = 0;
};

struct TileRec {
  stitch_tile_desc d;
  uint8_t *d_src = nullptr;   // uploaded bytes (inflated stream or dense rows); owned unless external
  bool src_external = false;
  size_t src_len = 0;
  uint8_t *d_pix = nullptr;   // dense unfiltered rows (== d_src for filtered==0)
  uint32_t pitch = 0;
  uint8_t *d_pass = nullptr;  // Adam7 pass scratch
  std::vector<stitch::UnfStream> streams;
  bool has_a7 = false;
  stitch::A7Image a7;
  bool decoded = false;
  uint8_t *d_z = nullptr;     // zlib stream as uploaded (descriptor said filtered == 2); d_src is its inflate target
  bool z_external = false;
  size_t z_len = 0;
  bool inflated = true;       // false: d_src is still to be produced from d_z (inflate.cu)
  // 8f.4 extension (colour type 3): the unfilter chains write the index plane d_idx, expand_palette_kernel makes d_pix (RGBA8)
  bool has_pal = false, pal_opaque = true;
  uint8_t *d_idx = nullptr;   // == d_src for filtered == 0
  uint32_t idx_pitch = 0;
  uint32_t pal[256];          // RGBA8 in memory order; entries past PLTE are opaque black
  cudaEvent_t up_event = nullptr;  // recorded on ctx->up after the tile's bytes were queued
  bool up_waited = true;           // ctx->stream already waits for up_event
};

struct stitch_job {
  stitch_ctx *ctx = nullptr;
  stitch_job_desc desc;
  uint32_t bpp = 4;  // output bytes per pixel
  std::vector<TileRec> tiles;
  bool finalized = false;
  // device-side job state
  stitch::DevTile *d_tiles = nullptr;
  uint32_t *d_bin_start = nullptr, *d_bin_tiles = nullptr;
  stitch::UnfStream *d_streams = nullptr; size_t streams_cap = 0;
  stitch::UnfStream *d_slow = nullptr; size_t slow_cap = 0;          // chains with a truncated final row (legacy kernel)
  stitch::UnfUnit *d_units = nullptr; size_t units_cap = 0;
  uint8_t *d_uprog = nullptr; size_t uprog_cap = 0;          // wavefront kernel: ticket (256 B) | seam mailbox
  std::vector<stitch::UnfStream> h_fast;  // chains whose tables are on the device (skip the upload when unchanged)
  uint32_t n_fast_units = 0;
  uint32_t seam_epoch = 0;                                   // tag of the last decode that used the mailbox
  stitch::A7Image *d_a7 = nullptr; size_t a7_cap = 0;
  stitch::PalImage *d_pal = nullptr; size_t pal_cap = 0;
  uint32_t bins_x = 0, bins_y = 0, bin_w_log2 = 9, bin_h_log2 = 4, bin_y_origin = 0;
  // inflate (8f.2): stream table, one result slot per tile, pinned mirror, tiles whose result is not checked yet
  stitch::InfStream *d_inf = nullptr; size_t inf_cap = 0;
  stitch::InfResult *d_infres = nullptr; size_t infres_cap = 0;
  stitch::InfResult *h_infres = nullptr; size_t h_infres_cap = 0;
  std::vector<uint32_t> inf_pending;
  uint32_t *d_sync = nullptr; size_t sync_words = 0;   // fused kernel: rowsum | band_done | ticket
  bool force_two_pass = false;
  // strip.cu: TMA tensor maps (one per tile), per-box-column interval plan, background row
  void *d_maps = nullptr, *d_ivs = nullptr, *d_col_start = nullptr, *d_lists = nullptr, *d_bgrow = nullptr;
  uint32_t strip_ncol = 0;
  bool strip_ok = false;
  // band buffers
  uint8_t *d_plans = nullptr; size_t plans_bytes = 0;  // fused kernel: per-unit plan records
  uint8_t *d_unit_left = nullptr; size_t unit_left_bytes = 0;  // fused kernel, plan mode: left pixels of every unit
  uint8_t *d_defl = nullptr; size_t defl_bytes = 0;    // deflate: compressed band / chunk (two buffers, used in turn)
  uint8_t *d_defl_b = nullptr; size_t defl_b_bytes = 0;
  uint8_t *d_pack[2] = {nullptr, nullptr}; size_t pack_bytes[2] = {0, 0};   // PNG rows of a chunk without row padding: one flat download
  cudaEvent_t ev_pack[2] = {nullptr, nullptr};                                // the download of d_pack[i] has finished
  cudaEvent_t ev_defl[2] = {nullptr, nullptr};         // the download of buffer i has finished
  uint8_t *d_defl_scratch = nullptr; size_t defl_scratch_bytes = 0;
  unsigned long long *h_defl = nullptr;                // pinned: {bytes, adler}
  uint8_t *d_ctab = nullptr; size_t ctab_bytes = 0;    // canvas-as-one-tile tables: DevTile | bin_start[2] | bin_tiles[1]
  std::vector<uint32_t> busy_bins;                     // prefix sum over bin rows of bins that need the generic compositor
  uint8_t *d_canvas = nullptr; size_t canvas_bytes = 0; size_t canvas_pitch = 0;
  uint8_t *d_out = nullptr; size_t out_bytes = 0;
  size_t out_pitch = 0, out_lead = 0, out_row_bytes = 0;
  uint32_t last_y0 = 0, last_y1 = 0;
  int status = STITCH_OK;
  // JPEG scratch (jpeg.cu): coefficients, bit offsets, per-row sizes, unstuffed stream, pinned totals
  uint8_t *d_jpeg[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  size_t jpeg_cap[5] = {0, 0, 0, 0, 0};
  unsigned long long *h_jpeg_totals = nullptr;  // pinned: [0] unstuffed bytes, [1] output bytes
  int jpeg_tables_quality = -1;
  uint32_t jpeg_skip_bands = 0, jpeg_overflow_streak = 0;   // jpeg.cu: bands that skip the one-pass route after overflows
  bool jpeg_overflowed = false;                 // jpeg.cu: the try without a canvas overflowed a chunk
  bool jpeg_huff_ready = false;                 // jpeg.cu: the Annex K tables are in constant memory, the MCU kernel's shared memory is opted in
  // JPEG tiles (jpeg_decode.cu): one pinned status word per tile, written by the tile's last decode kernel;
  // jd_pending: (slot, tile) pairs not checked yet
  uint32_t *h_jdstat = nullptr; size_t jdstat_cap = 0, jdstat_used = 0;
  std::vector<std::pair<uint32_t, uint32_t>> jd_pending;
  std::vector<std::pair<uint8_t *, size_t>> pinned;   // pinned staging chunks (jpeg_decode.cu), bump-allocated
  size_t pinned_used = 0;
};


int stitch_fail(stitch_ctx *ctx, int code, const char *fmt, ...);
// a tile whose dense rows a kernel chain on stream `produced_on` is producing at d_rows (cudaMallocAsync'ed there; the job owns it)
int stitch_add_tile_produced(stitch_job *job, const stitch_tile_desc *tile, uint8_t *d_rows, size_t len, cudaStream_t produced_on);
// tile k's dense rows -> host (after its upload / decode chain), with the job's pending decode errors checked
int stitch_job_read_tile(stitch_job *job, uint32_t k, uint8_t *out, size_t cap);
const char *stitch_jpeg_decode_message(uint32_t code);
int stitch_dev_alloc(stitch_ctx *ctx, void **p, size_t bytes);
void stitch_dev_free(stitch_ctx *ctx, void *p);
int stitch_ensure_buffer(stitch_ctx *ctx, uint8_t **p, size_t *have, size_t need);
// dst <- src (host, any memory) / dst <- 32-bit value, on ctx->stream, without the copy engines.  bytes % 4 == 0.
int stitch_upload_small(stitch_ctx *ctx, void *d_dst, const void *src, size_t bytes);
int stitch_fill_words(stitch_ctx *ctx, void *d_dst, uint32_t value, size_t bytes);
// pinned host memory <- device, written by a kernel (a few words: sizes, check values, status), so that the read-back
// does not queue behind the band downloads on the device-to-host copy engine.  Valid after a stream synchronisation.
int stitch_download_small(stitch_ctx *ctx, void *h_pinned_dst, const void *d_src, size_t bytes);
void stitch_arena_reset(stitch_ctx *ctx);   // after a stream synchronisation: the staged tables have been read

#define STITCH_CK(call)                                                                        \
  do {                                                                                         \
    cudaError_t e_ = (call);                                                                   \
    if (e_ != cudaSuccess)                                                                     \
      return stitch_fail(ctx, e_ == cudaErrorMemoryAllocation ? STITCH_E_NOMEM : STITCH_E_CUDA, \
                         "CUDA error %s at %s:%d (%s)", cudaGetErrorName(e_), __FILE__, __LINE__, cudaGetErrorString(e_)); \
  } while (0)

// jpeg.cu: kernel (d) over composited RGBA8 rows [y0,y1) held at `canvas` (row y0 first).  Leaves the
// entropy-coded bytes in job->d_out and their count in *out_bytes (stream synchronised once inside).
// deflate.cu: filtered rows -> deflate blocks at d_dst; *crc_raw is the CRC-32 remainder of those bytes from a zero
// register (chunks concatenate with stitch_crc_raw_concat, stitch_crc_finish turns the total into zlib's crc32)
int stitch_deflate_band(stitch_job *job, const uint8_t *d_rows, uint32_t n_rows, size_t host_row, uint8_t *d_dst,
                        size_t *out_bytes, uint32_t *adler, uint32_t *crc_raw, uint32_t *launches);
size_t stitch_deflate_bound(size_t raw_bytes);
uint32_t stitch_crc_raw_concat(uint32_t ra, uint32_t rb, uint64_t len_b);
uint32_t stitch_crc_finish(uint32_t raw, uint64_t n);
// strip_plan.cu: interval plan + tensor maps for strip.cu; dt = the tiles back to front as uploaded to d_tiles
int stitch_build_strip_plan(stitch_job *job, std::vector<stitch::DevTile> &dt, const std::vector<uint32_t> &src_words,
                            const std::vector<uint32_t> &src_rows);
bool stitch_encode_tile_maps(const stitch::DevTile &t, uint32_t words_per_row, uint32_t rows, void *out3);
// canvas == nullptr: try the band from the tiles themselves (an exact mosaic of RGBA8 tiles, no blending); when
// that is not possible *need_canvas comes back true and the caller composites and calls again with the canvas.
int stitch_jpeg_encode_band(stitch_job *job, uint32_t y0, uint32_t y1, const uint8_t *canvas, size_t canvas_pitch,
                            size_t *out_bytes, uint32_t *launches, bool *need_canvas);
