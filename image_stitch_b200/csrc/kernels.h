// kernels.h -- launcher prototypes shared by the translation units of libstitch_b200.so
#pragma once
#include "common.cuh"

#include <mutex>

struct stitch_job;
struct stitch_timings;

namespace stitch {
// cudaFuncSetAttribute, occupancy figures and __device__/__constant__ symbols are PER DEVICE: one-time set-up is
// remembered per device (a process may drive several: ConcatOptions.devices, one stitch_ctx per device), under a
// lock because the contexts' threads launch concurrently.  value_plus_1[dev] == 0: not done yet on that device.
struct PerDevice {
  std::mutex mu;
  int value_plus_1[64] = {};
  template <class F>   // F: cudaError_t(int *value), run once per device with that device current
  cudaError_t get(int *value, F init) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!value_plus_1[dev]) {
      int v = 0;
      if ((e = init(&v)) != cudaSuccess) return e;
      value_plus_1[dev] = v + 1;
    }
    *value = value_plus_1[dev] - 1;
    return cudaSuccess;
  }
};
}  // namespace stitch

namespace stitch {

struct ComposeParams {
  const DevTile *tiles;       // sorted by (z_index, image_idx)
  const uint32_t *bin_start;  // CSR over bins, bins_x * bins_y + 1 entries
  const uint32_t *bin_tiles;  // tile indices, back to front within a bin
  uint32_t bins_x, bin_w_log2, bin_h_log2;
  uint32_t bin_y_origin;      // canvas row of bin row 0
  uint32_t canvas_w;
  uint32_t y_begin, y_end;    // canvas rows to produce
  uint32_t alpha_blend;
  uint32_t bg_lo, bg_hi;      // background pixel, memory order
  uint8_t *out;               // row (y - y_begin) at out + (y - y_begin) * out_pitch
  size_t out_pitch;
};

struct FilterParams {
  const uint8_t *rows;   // raw rows; row r at rows + r*in_pitch
  size_t in_pitch;
  const uint8_t *prev0;  // raw row above row 0, or nullptr (previousLine === null)
  uint32_t n_rows;
  uint32_t row_bytes;
  uint32_t bpp;
  uint8_t *out;          // row r: filter byte at out + r*out_pitch + lead, data right after it
  size_t out_pitch;
  uint32_t lead;
};

// Fused composite + filter-select (fused.cu).  Bins are FB_R rows x FB_SWW words; each bin's tile list
// also holds tiles touching the row above / the pixel left of it.  Units (one per warp) are
// fused_unit_rows() x fused_unit_words() and never straddle a bin.
struct StripIv;
struct FusedParams {
  const DevTile *tiles;
  const uint32_t *bin_start;
  const uint32_t *bin_tiles;
  uint32_t bins_x;
  uint32_t canvas_w;
  uint32_t y_begin, y_end;    // canvas rows to produce
  uint32_t band0;             // absolute index (y / FB_R) of the first band
  uint32_t n_bands, nseg;    // unit bands in [y_begin,y_end), units per band
  uint32_t alpha_blend;
  uint32_t bg_lo, bg_hi;
  unsigned long long *rowsum; // [(y_end - y_begin)][8], zeroed: low half = cost sum, high half = units arrived
  uint32_t *ticket;           // zeroed
  void *plans;                // fused_plan_bytes(n_bands * nseg) of scratch: per-unit plan records
  uint32_t single_bin;        // 1: bin_start/bin_tiles describe one bin that covers the whole canvas
  const void *maps;           // per tile three 128-byte CUtensorMaps (boxes of 1, 4 and 9 rows of 128 words), or nullptr
  // The host's interval plan (strip_plan.cu), or nullptr.  With it the kernel needs no bins: a unit is one 128-word
  // column of the plan, its rows are composed interval by interval (background / one tile / layers that all span
  // the column / generic), and a band's units are dealt out most expensive first.
  const StripIv *ivs;
  const uint32_t *col_start;
  const uint32_t *lists;
  uint32_t *unit_rows;        // plan mode scratch: fused_unit_rows_bytes(n_units), prepared rows of the units that need them
  uint32_t *unit_left;        // plan mode scratch: fused_unit_left_bytes(n_units), the pixels left of every unit | prep_list
  uint32_t *prep_list;        // plan mode scratch (tail of unit_left's buffer): the prepare kernel's tasks
  uint32_t *prep_count;       // zeroed: entries of prep_list
  uint8_t *out;               // row r: filter byte at out + r*out_pitch + lead
  size_t out_pitch;
  uint32_t lead;
};
cudaError_t launch_fused(const FusedParams &p, bool out16, int sm_count, cudaStream_t st);
size_t fused_plan_bytes(uint32_t n_units);
size_t fused_unit_rows_bytes(uint32_t n_units);
size_t fused_unit_left_bytes(uint32_t n_units);
cudaError_t launch_compose_canvas(const FusedParams &p, bool out16, uint8_t *canvas, size_t pitch, uint32_t ys,
                                  uint32_t ye, cudaStream_t st);
uint32_t fused_rows_per_band();
uint32_t fused_words_per_seg();
uint32_t fused_unit_rows();
uint32_t fused_unit_words();

// Row-pipelined composite + filter-select (strip.cu), kernels (b)+(c) of north_star with TMA-staged row bands.
// A canvas row is cut into BOX columns of 256 words (1 KiB).  For every box column the host lists the y-intervals
// over which the set of tiles touching the column does not change (StripIv) and what a box row of that interval
// is made of:
//   SK_DIRECT  one native-format tile covers the box and overwrites what is below: one 2-D TMA box load
//   SK_BG      no tile: a bulk copy of the background row
//   SK_SHIFT   the same, but the box starts at a source byte that is not a multiple of 16 (a tiled TMA load faults
//              on such a start coordinate -- measured on B200), or the tile's pitch / width rule TMA out: the
//              producer warp copies the box with 4-byte cp.async completing on the same mbarrier
//   SK_COMP    anything else: normalise + alpha-over by the CTA's warps over the interval's tile list
struct StripIv {
  uint32_t y_end;             // exclusive canvas row where the interval ends
  uint32_t kind;              // StripKind
  uint32_t tile;              // SK_DIRECT, SK_SHIFT: index into tiles / maps
  int32_t cw, cy;             // SK_DIRECT, SK_SHIFT: the box's first word is word cw of source row (y + cy)
  uint32_t list_off, n_list;  // SK_COMP: tiles touching the column in this interval, back to front (into lists)
                              // SK_DIRECT, SK_SHIFT: list_off = words of the box that exist in the source row (<= 128)
  uint32_t pad;               // SK_COMP: bit 1 = every tile of the list is in the output's format, bit 0 = and spans the whole column
};
enum StripKind : uint32_t { SK_BG = 0, SK_DIRECT = 1, SK_COMP = 2, SK_SHIFT = 3 };
struct StripParams {
  const DevTile *tiles;
  const void *maps;           // three 128-byte CUtensorMaps per tile (32-bit words): boxes of 128 words x 1, 4, 9 rows
  const StripIv *ivs;
  const uint32_t *col_start;  // [ncol + 1]: intervals of box column c are ivs[col_start[c] .. col_start[c + 1])
  const uint32_t *lists;
  const uint8_t *bg_row;      // 1 KiB of background pixels, 16-byte aligned
  uint32_t canvas_w;
  uint32_t y_begin, y_end;    // canvas rows to produce
  uint32_t nsx, nry, ncol;    // column strips x row ranges (one CTA each); box columns of the canvas
  uint32_t rows_per_range;
  uint32_t alpha_blend;
  uint32_t bg_lo, bg_hi;
  unsigned long long *rowsum; // [(y_end - y_begin)][8], zeroed: low half = cost sum, high half = warps arrived
  uint8_t *out;               // row r: filter byte at out + r*out_pitch + lead
  size_t out_pitch;
  uint32_t lead;
};
uint32_t strip_box_words();
// q = 0: the launcher picks the quads per thread.  cudaErrorInvalidConfiguration: the row is too wide for one
// resident grid (the caller falls back to composite-then-filter).
cudaError_t launch_strip(StripParams p, bool out16, int q, int device, int sm_count, cudaStream_t st);
// Normalise + composite rows [ys, ye) into a dense canvas in HBM, driven by the same interval plan (strip.cu).
cudaError_t launch_compose_plan(StripParams p, bool out16, uint8_t *canvas, size_t pitch, uint32_t ys, uint32_t ye, cudaStream_t st);

cudaError_t launch_unfilter(const UnfStream *d_streams, uint32_t n_streams, DevStatus *d_status, cudaStream_t st);
// d_seam: mailbox of unfilter_wave_seam_bytes() per chain, zeroed once; epoch: a value never used with it before (> 0).
size_t unfilter_wave_seam_bytes(uint32_t rows, uint32_t row_bytes);
cudaError_t launch_unfilter_wave(const UnfStream *d_streams, const UnfUnit *d_units, uint32_t n_units, void *d_seam,
                                 uint32_t epoch, uint32_t *d_ticket, DevStatus *d_status, int sm_count,
                                 uint32_t uniform_bpp /* 0 = mixed */, cudaStream_t st);
cudaError_t launch_adam7_gather(const A7Image *d_images, uint32_t n_images, uint32_t max_row_bytes,
                                uint32_t max_height, cudaStream_t st);
cudaError_t launch_expand_palette(const PalImage *d_images, uint32_t n_images, uint32_t max_width, uint32_t max_height,
                                  cudaStream_t st);
cudaError_t launch_compose(const ComposeParams &p, bool out16, cudaStream_t st);
cudaError_t launch_filter_rows(const FilterParams &p, cudaStream_t st);

// inflate.cu: one zlib stream (concatenated IDAT payloads of one PNG) per warp.
struct InfStream {
  const uint8_t *src;   // zlib stream, 4-byte aligned, readable up to src_len rounded up to 4
  uint8_t *dst;         // inflated bytes, 16-byte aligned, writable up to dst_cap rounded up to 16
  uint32_t src_len;
  uint32_t dst_cap;     // bytes the caller wants: decoding stops there
  uint32_t slot;        // index of this stream's InfResult
  uint32_t pad;
};
struct InfResult {
  uint32_t status;        // 0 or InfStatus
  uint32_t out_len;       // bytes written (<= dst_cap)
  uint32_t adler_stored;  // the stream's trailer
  uint32_t flags;         // bit 0: stopped at dst_cap before the stream ended; bit 1: trailer read
};
enum InfStatus : uint32_t {
  INF_OK = 0, INF_E_HEADER, INF_E_METHOD, INF_E_WINDOW, INF_E_DICT, INF_E_BLOCK_TYPE, INF_E_STORED, INF_E_TOO_MANY,
  INF_E_CODE_LENGTHS, INF_E_BIT_REPEAT, INF_E_NO_EOB, INF_E_LITLEN_SET, INF_E_DIST_SET, INF_E_LITLEN_CODE,
  INF_E_DIST_CODE, INF_E_FAR, INF_E_DATA_CHECK, INF_E_EOF
};
cudaError_t launch_inflate(const InfStream *d_streams, InfResult *d_results, uint32_t n_streams, cudaStream_t st);
const char *inflate_message(uint32_t status);

}  // namespace stitch
