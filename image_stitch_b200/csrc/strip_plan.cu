// strip_plan.cu -- host side of the row-pipelined composite+filter kernel (strip.cu): the per-box-column interval
// plan and the TMA tensor maps of the tiles.  The plan is the GPU-shaped form of buildScanlineIndex
// (src/positioned-layout.ts:201-242: per output row, the intersecting inputs back to front): per 1 KiB box column
// of the canvas, the y-intervals over which the set of tiles touching the column does not change, each with the
// tiles back to front and a classification (direct copy of one tile / background / needs compositing).
#include "job.h"

#include <cuda.h>   // CUtensorMap and enums only: the encoder is looked up at run time, libcuda is not linked

#include <algorithm>
#include <cstring>

using namespace stitch;

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;   // a process-wide function pointer of the driver, not per-device state
  static bool tried = false;
  if (!tried) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
    else cudaGetLastError();
    tried = true;
  }
  return fn;
}

// Rows of 32-bit words, box of one column (128 words) x box_rows rows; elements outside the tile read as zero.
static bool encode_tile_map(const DevTile &t, uint32_t words_per_row, uint32_t rows, uint32_t box_rows, CUtensorMap *out) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return false;
  if ((reinterpret_cast<uintptr_t>(t.pix) & 15u) || (t.pitch & 15u) || !words_per_row || !rows) return false;
  const cuuint64_t dims[2] = {words_per_row, rows};
  const cuuint64_t strides[1] = {t.pitch};
  const cuuint32_t box[2] = {strip_box_words(), box_rows};
  const cuuint32_t estr[2] = {1, 1};
  return fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<uint8_t *>(t.pix), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// Three maps (1, 4, 9 rows) for one tile into out[3]; false when the tile cannot be a TMA source.
bool stitch_encode_tile_maps(const DevTile &t, uint32_t words_per_row, uint32_t rows, void *out) {
  static_assert(sizeof(CUtensorMap) == 128, "one 128-byte descriptor");
  CUtensorMap *m = static_cast<CUtensorMap *>(out);
  return words_per_row >= strip_box_words() && encode_tile_map(t, words_per_row, rows, 1, m) &&
         encode_tile_map(t, words_per_row, rows, 4, m + 1) && encode_tile_map(t, words_per_row, rows, 9, m + 2);
}

template <class T>
static int upload_vec(stitch_ctx *ctx, const std::vector<T> &v, void **d_ptr) {
  stitch_dev_free(ctx, *d_ptr);
  *d_ptr = nullptr;
  int rc = stitch_dev_alloc(ctx, d_ptr, std::max<size_t>(v.size() * sizeof(T), 16));
  if (rc) return rc;
  static_assert(sizeof(T) % 4 == 0, "tables are copied word by word");
  return v.empty() ? STITCH_OK : stitch_upload_small(ctx, *d_ptr, v.data(), v.size() * sizeof(T));
}

// dt: the job's tiles back to front (uploaded to job->d_tiles AFTER this call: dt[k].pad receives "has valid tensor
// maps"); src_words/src_rows: each tile's full size.
int stitch_build_strip_plan(stitch_job *job, std::vector<DevTile> &dt, const std::vector<uint32_t> &src_words,
                            const std::vector<uint32_t> &src_rows) {
  stitch_ctx *ctx = job->ctx;
  const stitch_job_desc &jd = job->desc;
  const uint32_t bpw = job->bpp / 4, box_px = strip_box_words() / bpw;
  const uint32_t ncol = (jd.canvas_w + box_px - 1) / box_px;
  const uint32_t native = jd.out_bit_depth == 16 ? (uint32_t)FMT_RGBA16 : (uint32_t)FMT_RGBA8;
  job->strip_ok = false;

  std::vector<CUtensorMap> maps(std::max<size_t>(3 * dt.size(), 1));   // per tile: boxes of 1, 4 and 9 rows
  std::vector<uint8_t> direct_ok(dt.size(), 0);
  for (size_t k = 0; k < dt.size(); k++) {
    memset(&maps[3 * k], 0, 3 * sizeof(CUtensorMap));
    direct_ok[k] = dt[k].fmt == native && src_words[k] >= strip_box_words() &&
                   encode_tile_map(dt[k], src_words[k], src_rows[k], 1, &maps[3 * k]) &&
                   encode_tile_map(dt[k], src_words[k], src_rows[k], 4, &maps[3 * k + 1]) &&
                   encode_tile_map(dt[k], src_words[k], src_rows[k], 9, &maps[3 * k + 2]);
    dt[k].pad = direct_ok[k];   // the kernels may use the tile's maps
  }

  std::vector<std::vector<uint32_t>> col_tiles(ncol);
  for (uint32_t k = 0; k < dt.size(); k++) {
    const DevTile &t = dt[k];
    if (t.x1 <= t.x0 || t.y1 <= t.y0) continue;
    const uint32_t c0 = (uint32_t)t.x0 / box_px, c1 = std::min(((uint32_t)t.x1 - 1u) / box_px, ncol - 1);
    for (uint32_t c = c0; c <= c1; c++) col_tiles[c].push_back(k);
  }
  std::vector<StripIv> ivs;
  std::vector<uint32_t> col_start(ncol + 1, 0), lists;
  std::vector<uint32_t> ys, cur;
  for (uint32_t c = 0; c < ncol; c++) {
    col_start[c] = (uint32_t)ivs.size();
    const int32_t X0 = (int32_t)(c * box_px), X1 = (int32_t)std::min<uint64_t>((uint64_t)X0 + box_px, jd.canvas_w);
    ys.clear();
    ys.push_back(0); ys.push_back(jd.canvas_h);
    for (uint32_t k : col_tiles[c]) {
      ys.push_back((uint32_t)std::min<int64_t>(std::max<int32_t>(dt[k].y0, 0), jd.canvas_h));
      ys.push_back((uint32_t)std::min<int64_t>(std::max<int32_t>(dt[k].y1, 0), jd.canvas_h));
    }
    std::sort(ys.begin(), ys.end());
    ys.erase(std::unique(ys.begin(), ys.end()), ys.end());
    size_t first_of_col = ivs.size();
    size_t prev_list_begin = 0, prev_list_len = 0;
    for (size_t i = 0; i + 1 < ys.size(); i++) {
      const int32_t ya = (int32_t)ys[i], yb = (int32_t)ys[i + 1];
      cur.clear();
      for (uint32_t k : col_tiles[c]) if (dt[k].y0 <= ya && dt[k].y1 >= yb) cur.push_back(k);
      // everything below the top-most tile that covers the whole box and overwrites is hidden
      size_t k0 = 0;
      for (size_t q = cur.size(); q-- > 0;) {
        const DevTile &t = dt[cur[q]];
        if ((!jd.alpha_blend || t.opaque) && t.x0 <= X0 && t.x1 >= X1) { k0 = q; break; }
      }
      cur.erase(cur.begin(), cur.begin() + (ptrdiff_t)k0);
      StripIv iv{};
      iv.y_end = (uint32_t)yb;
      if (cur.empty()) {
        iv.kind = SK_BG;
      } else {
        const DevTile &t = dt[cur.back()];
        const bool top_covers = (!jd.alpha_blend || t.opaque) && t.x0 <= X0 && t.x1 >= X1;
        if (top_covers && t.fmt == native) {
          iv.tile = cur.back();
          iv.cw = (int32_t)((X0 - t.x0 + t.sx0) * (int32_t)bpw);
          iv.cy = t.sy0 - t.y0;
          // a tiled TMA load needs its start coordinate on a 16-byte boundary (it faults otherwise: measured)
          iv.kind = direct_ok[cur.back()] && (iv.cw & 3) == 0 ? SK_DIRECT : SK_SHIFT;
          iv.list_off = std::min<uint32_t>(strip_box_words(), src_words[cur.back()] - (uint32_t)iv.cw);
        } else {
          iv.kind = SK_COMP;
          iv.n_list = (uint32_t)cur.size();
          // bit 1: every layer is in the output's own format (word loads); bit 0: and spans the whole column (no per-pixel tests)
          bool words = true, spans = true;
          for (uint32_t k : cur) {
            words = words && dt[k].fmt == native && (dt[k].pitch & 3u) == 0;
            spans = spans && dt[k].x0 <= X0 && dt[k].x1 >= X1;
          }
          iv.pad = (words && spans ? 1u : 0u) | (words ? 2u : 0u);
        }
      }
      // merge with the interval above when it asks for the same thing
      bool merged = false;
      if (ivs.size() > first_of_col) {
        StripIv &pv = ivs.back();
        if (pv.kind == iv.kind) {
          if (iv.kind == SK_BG) merged = true;
          else if (iv.kind == SK_DIRECT || iv.kind == SK_SHIFT) merged = pv.tile == iv.tile;
          else merged = prev_list_len == cur.size() && std::equal(cur.begin(), cur.end(), lists.begin() + (ptrdiff_t)prev_list_begin);
        }
        if (merged) pv.y_end = iv.y_end;
      }
      if (!merged) {
        if (iv.kind == SK_COMP) {
          iv.list_off = (uint32_t)lists.size();
          prev_list_begin = lists.size(); prev_list_len = cur.size();
          lists.insert(lists.end(), cur.begin(), cur.end());
        }
        ivs.push_back(iv);
      }
    }
  }
  col_start[ncol] = (uint32_t)ivs.size();
  if (lists.empty()) lists.push_back(0);

  std::vector<uint32_t> bg(strip_box_words());
  uint32_t lo, hi;
  memcpy(&lo, jd.background, 4); memcpy(&hi, jd.background + 4, 4);
  for (uint32_t i = 0; i < bg.size(); i++) bg[i] = bpw == 2 ? ((i & 1u) ? hi : lo) : lo;

  int rc;
  static_assert(sizeof(CUtensorMap) == 128, "one 128-byte descriptor per tile");
  static_assert(sizeof(StripIv) == 32, "two 16-byte loads per interval");
  if ((rc = upload_vec(ctx, maps, &job->d_maps))) return rc;
  if ((rc = upload_vec(ctx, ivs, &job->d_ivs))) return rc;
  if ((rc = upload_vec(ctx, col_start, &job->d_col_start))) return rc;
  if ((rc = upload_vec(ctx, lists, &job->d_lists))) return rc;
  if ((rc = upload_vec(ctx, bg, &job->d_bgrow))) return rc;
  job->strip_ncol = ncol;
  job->strip_ok = true;
  return STITCH_OK;
}
