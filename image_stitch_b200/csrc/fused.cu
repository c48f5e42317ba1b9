// fused.cu -- kernels (b)+(c) in ONE pass over HBM: normalise + composite the
// tiles under a canvas patch, compute the five PNG filter costs
// (src/png-filter.ts:148-181), agree on the per-row winner across all the
// warps that share a canvas row, and write the filtered scanlines.  HBM
// traffic is "read every visible tile pixel once, write every output byte
// once" (the 1-row halo above a patch is an L2 hit: the patch above is being
// processed at the same time).
//
// Work decomposition: a UNIT is FU_R canvas rows x 128 32-bit words (one
// 16-byte quad per lane per row) and belongs to ONE WARP; warps are independent
// workers with a private shared-memory stash, so there is no CTA barrier
// anywhere.  A row's cost is a sum over all units of its band: units publish
// partial sums with red.global.add and bump a per-band counter; a warp writes
// its rows once every unit of the band has arrived (the 64-bit sums carry their
// own arrival count).  Units are dealt round-robin to a cooperatively launched
// (fully resident) grid with at least as many warps as a band has units; see
// fused_kernel for why no wait cycle can form.
//
//   phase A   raw stash <- cp.async of the one tile that covers the unit (grid
//             mosaics, opaque layers), else the generic per-quad compositor
//   sweep     rows top to bottom, previous row carried in registers already
//             unpacked to half2; per word: None/Sub/Up/Average costs on packed
//             bytes (VABSDIFF4 + accumulate), Paeth two bytes at a time in
//             half2 (exact for 9-bit integers); the Paeth residual falls out of
//             the half2 pipeline as (x - pred + 1280) whose low byte is the
//             filtered byte, and is stashed for phase B
//   phase B   winner per row (first minimum: None, Sub, Up, Average, Paeth),
//             one 128-bit store per lane per row
#include "common.cuh"
#include "kernels.h"
#include "filter_math.cuh"
#include "async_copy.cuh"

namespace stitch {

static constexpr int FB_R = 16;     // bin height (host binning granularity), rows
static constexpr int FB_SWW = 512;  // bin width in 32-bit words

static constexpr int FU_R = 8;                       // rows per unit
static constexpr int FU_SW = 128;                    // words per unit row
static constexpr int FU_WARPS = 24;                  // workers per CTA (one CTA per SM): what 227 KB of shared memory hold; 80 registers each
static constexpr int FU_RAW_WORDS = (FU_R + 1) * FU_SW;  // row 0 = predictor row above the unit
static constexpr int FU_LC_WORDS = ((FU_R + 1) * 2 + 3) / 4 * 4;  // two words left of every row
static constexpr int FU_WARP_WORDS = (FU_RAW_WORDS + FU_LC_WORDS + 31) / 32 * 32;  // stashes start on 128-byte boundaries (TMA destinations)
static constexpr int FU_SLOTS = 2;                   // stashes per warp
static constexpr size_t FU_CTA_SMEM = (size_t)FU_WARPS * FU_SLOTS * FU_WARP_WORDS * 4 + 16 + FU_WARPS * FU_SLOTS * 8;  // + TMEM base address + one mbarrier per stash
// The Paeth residuals of a unit (FU_R rows x 16 bytes per lane) wait for the row winners in TENSOR MEMORY:
// lane l of warp w owns TMEM lane 32 (w % 4) + l (a warp reaches only that quarter), warps w / 4 = 0..3 use
// disjoint column ranges, one unit takes FU_R * 4 columns.  This is plain scratch use of
// TMEM (tcgen05.st / tcgen05.ld, no MMA): it takes the residuals out of shared memory, which is what limits
// the number of resident warps.
static constexpr int FU_TMEM_SLOT_COLS = FU_R * 4;
static constexpr int FU_TMEM_COLS = 512;             // power of two >= (FU_WARPS / 4) * FU_SLOTS * FU_TMEM_SLOT_COLS
static constexpr int FU_MAX_CTAS = 512 / FU_TMEM_COLS;  // more CTAs per SM than this would spin in tcgen05.alloc
static_assert((FU_WARPS / 4) * FU_SLOTS * FU_TMEM_SLOT_COLS <= FU_TMEM_COLS && FU_WARPS % 4 == 0, "TMEM stash layout");
static_assert(FB_R % FU_R == 0 && FB_SWW % FU_SW == 0, "a unit must lie inside one bin");

__device__ __forceinline__ void fu_cp_async16(void *smem, const void *gmem) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}

__device__ __forceinline__ void fu_cp_async4(uint32_t sa, const void *gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void fu_cp_async16s(uint32_t sa, const void *gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}

// One row of the sweep.  GUARD: the unit reaches past the right edge of the canvas.
// hb/hlu: the row above, unpacked (input); hx/hlc: this row, unpacked (output, becomes the next row's hb/hlu).
template <int BPW, bool GUARD>
__device__ __forceinline__ void sweep_row(const uint4 cur, const uint32_t lc0, const uint32_t lc1, const uint4 up,
                                          const H4 (&hb)[4], const H4 (&hlu)[2], H4 (&hx)[4], H4 (&hlc)[2], int qw,
                                          int wmax, uint32_t pf_taddr, uint32_t (&sum)[5]) {
  const uint32_t x[4] = {cur.x, cur.y, cur.z, cur.w}, b[4] = {up.x, up.y, up.z, up.w};
  const uint32_t lc[2] = {lc0, lc1};
  if (BPW == 2) hlc[0] = unpack_h4(lc0);
  hlc[1] = unpack_h4(lc1);
  uint32_t s0 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0, pf[4];
#pragma unroll
  for (int j = 0; j < 4; j++) {
    hx[j] = unpack_h4(x[j]);
    const uint32_t a = j >= BPW ? x[j - BPW] : lc[2 - BPW + j];
    const H4 &ha = j >= BPW ? hx[j - BPW] : hlc[2 - BPW + j];
    const H4 &hc = j >= BPW ? hb[j - BPW] : hlu[2 - BPW + j];
    pf[j] = pack_h4(paeth_resid_h2(hx[j].lo, ha.lo, hb[j].lo, hc.lo), paeth_resid_h2(hx[j].hi, ha.hi, hb[j].hi, hc.hi));
    if (!GUARD || qw + j < wmax) {  // words past the canvas edge carry no cost
      s0 = acc_abs_s4(x[j], s0);
      s1 = acc_abs_s4(absdiff_u4(x[j], a), s1);
      s2 = acc_abs_s4(absdiff_u4(x[j], b[j]), s2);
      s3 = acc_abs_s4(absdiff_u4(x[j], avg4(a, b[j])), s3);
      s4 = acc_abs_s4(pf[j], s4);
    }
  }
  tmem_st4(pf_taddr, make_uint4(pf[0], pf[1], pf[2], pf[3]));
  sum[0] = s0; sum[1] = s1; sum[2] = s2; sum[3] = s3; sum[4] = s4;
}

// Row r of the unit: loads, left context, costs, warp reduction, one red.global per candidate.
template <int BPW>
__device__ __forceinline__ void sweep_step(const uint32_t *raw_row, const uint32_t *lcs_row, uint32_t pf_row,
                                           unsigned long long *rs_row, uint32_t lane, const uint32_t (&lmask)[5], bool guard,
                                           int qw, int wmax, uint4 &up, const H4 (&hb)[4], const H4 (&hlu)[2],
                                           H4 (&hx)[4], H4 (&hlc)[2]) {
  const uint4 cur = *reinterpret_cast<const uint4 *>(raw_row + 4 * lane);
  uint32_t c0 = BPW == 2 ? __shfl_up_sync(0xffffffffu, cur.z, 1) : 0u;
  uint32_t c1 = __shfl_up_sync(0xffffffffu, cur.w, 1);
  if (lane == 0u) { c0 = lcs_row[0]; c1 = lcs_row[1]; }
  uint32_t sum[5];
  if (guard) sweep_row<BPW, true>(cur, c0, c1, up, hb, hlu, hx, hlc, qw, wmax, pf_row, sum);
  else sweep_row<BPW, false>(cur, c0, c1, up, hb, hlu, hx, hlc, qw, wmax, pf_row, sum);
  uint32_t v = 0u;
#pragma unroll
  for (int f = 0; f < 5; f++) v |= __reduce_add_sync(0xffffffffu, sum[f]) & lmask[f];
  if (lane < 5u) atomicAdd(rs_row + lane, (1ull << 32) | (unsigned long long)v);  // sum | arrival count
  up = cur;
}

// Everything a warp needs to know about one unit; built by unit_plan() one iteration ahead so that the
// bin -> list -> tile load chain overlaps with the cp.async traffic of the unit before it.
struct UnitPlan {
  uint32_t seg;
  int ya, nrows, xw0;
  const uint32_t *list;
  uint32_t n_list;
  const uint8_t *fast_src;  // tile byte that lands on (X0, ya) when the unit is a plain window of one tile
  uint32_t fast_pitch;
  bool halo_fast, left_fast;
  bool tma_ok;              // the covering tile has tensor maps
  uint32_t tile;            // the covering tile (index into tiles / maps)
  int32_t c0, c1;           // TMA coordinates of the predictor row's first word: word in the source row, source row
  uint32_t iv, iv_left;     // plan mode: the intervals of the unit's column and of the column left of it that hold row max(ya - 1, 0)
  uint32_t u, src_kind;     // plan mode: index of the unit's record; 4 = one TMA box from the tile, 8 = background only, 0 = prepared rows
};

// Rows and columns of unit u (pure arithmetic).
__device__ __forceinline__ void unit_geom(const FusedParams &p, const uint32_t u, UnitPlan &pl, const uint32_t seg_of_rec = 0xffffffffu) {
  const uint32_t band = u / p.nseg;
  pl.seg = seg_of_rec != 0xffffffffu ? seg_of_rec : u - band * p.nseg;   // plan mode deals a band's units out of order
  const uint32_t band_abs = p.band0 + band;
  const int ya = max((int)p.y_begin, (int)(band_abs * FU_R));
  const int yb = min((int)p.y_end, (int)((band_abs + 1) * FU_R));
  pl.ya = ya; pl.nrows = yb - ya;
  pl.xw0 = (int)pl.seg * FU_SW;
}

template <int BPW>
__device__ __forceinline__ UnitPlan unit_plan(const FusedParams &p, const uint32_t u) {
  constexpr int PXQ = 4 / BPW;  // pixels per quad
  UnitPlan pl;
  unit_geom(p, u, pl);
  const int ya = pl.ya, yb = pl.ya + pl.nrows, xw0 = pl.xw0;
  const uint32_t bin = p.single_bin ? 0u : ((uint32_t)ya / FB_R) * p.bins_x + (uint32_t)(xw0 / FB_SWW);
  const uint32_t b0 = __ldg(p.bin_start + bin), n_list = __ldg(p.bin_start + bin + 1) - b0;
  const uint32_t *list = p.bin_tiles + b0;
  pl.list = list; pl.n_list = n_list;
  const int X0 = xw0 / BPW, X1 = min((int)p.canvas_w, (xw0 + FU_SW) / BPW);  // unit's pixel columns
  pl.fast_src = nullptr; pl.fast_pitch = 0; pl.halo_fast = false; pl.left_fast = false; pl.tma_ok = false; pl.tile = 0; pl.c0 = 0; pl.c1 = 0;
  for (int k = (int)n_list - 1; k >= 0; k--) {
    const DevTile &t = p.tiles[__ldg(list + k)];
    const int tx0 = t.x0, ty0 = t.y0, tx1 = t.x1, ty1 = t.y1;
    if (ty0 >= yb || ty1 <= ya || tx0 >= X1 || tx1 <= X0) continue;
    // top-most tile that touches the unit
    const bool covers = ty0 <= ya && ty1 >= yb && tx0 <= X0 && tx1 >= X1;
    const bool overwrite = p.alpha_blend == 0u || t.opaque;
    const bool native = BPW == 1 ? t.fmt == FMT_RGBA8 : t.fmt == FMT_RGBA16;
    if (covers && overwrite && native && ((X1 - X0) % PXQ) == 0) {
      const uint32_t pitch = t.pitch;
      const uint8_t *src = t.pix + (size_t)(ya - ty0 + t.sy0) * pitch + (size_t)(X0 - tx0 + t.sx0) * (4u * BPW);
      if (((reinterpret_cast<uintptr_t>(src) | pitch) & 15u) == 0) {
        pl.fast_src = src; pl.fast_pitch = pitch;
        pl.tile = __ldg(list + k);
        pl.tma_ok = t.pad != 0u && p.maps != nullptr;
        pl.c0 = (X0 - tx0 + t.sx0) * BPW; pl.c1 = ya - 1 - ty0 + t.sy0;
        pl.halo_fast = ty0 <= ya - 1;
        pl.left_fast = tx0 <= X0 - 1;
        // The predictor row ya-1 and the pixel column X0-1 lie outside the unit: a tile ABOVE this one in z order
        // that ends exactly at the unit's top or left edge does not touch the unit (it was skipped above) but is
        // what the canvas shows there, so the copies straight out of this tile would be wrong.
        for (int q = k + 1; q < (int)n_list && (pl.halo_fast || pl.left_fast); q++) {
          const DevTile &h = p.tiles[__ldg(list + q)];
          if (h.x0 >= h.x1 || h.y0 >= h.y1) continue;
          if (h.y0 <= ya - 1 && h.y1 > ya - 1 && h.x0 < X1 && h.x1 > X0) pl.halo_fast = false;
          if (h.x0 <= X0 - 1 && h.x1 > X0 - 1 && h.y0 < yb && h.y1 > ya - 1) pl.left_fast = false;
        }
      }
    }
    break;
  }
  return pl;
}

// The plans depend on the job's geometry only, not on pixel data: a small kernel writes one 32-byte record
// per unit ahead of the fused kernel, whose warps then need a single load (instead of the dependent
// bin -> list -> tile chain) between taking a ticket and issuing the unit's copies.
struct __align__(16) PlanRec {
  const uint8_t *fast_src;
  uint32_t fast_pitch;
  uint32_t list_off;   // into bin_tiles
  uint32_t n_flags;    // n_list << 3 | bit 0 halo_fast | bit 1 left_fast | bit 2 tma_ok
  uint32_t tile;
  int32_t c0, c1;      // see UnitPlan
};
static_assert(sizeof(PlanRec) == 32, "two 16-byte loads per record");

template <int BPW>
__global__ void fused_plan_kernel(const FusedParams p, PlanRec *recs, const uint32_t n_units) {
  const uint32_t u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= n_units) return;
  const UnitPlan pl = unit_plan<BPW>(p, u);
  PlanRec r;
  r.fast_src = pl.fast_src; r.fast_pitch = pl.fast_pitch;
  r.list_off = (uint32_t)(pl.list - p.bin_tiles);
  r.n_flags = (pl.n_list << 3) | (pl.halo_fast ? 1u : 0u) | (pl.left_fast ? 2u : 0u) | (pl.tma_ok ? 4u : 0u);
  r.tile = pl.tile; r.c0 = pl.c0; r.c1 = pl.c1;
  recs[u] = r;
}

enum : uint32_t { PK_TMA = 4u, PK_BG = 8u, PK_COPY = 16u, PK_LEFT = 32u };   // plan mode: flags of a record, see unit_class()
struct PlanRaw { uint4 a, b; };
__device__ __forceinline__ PlanRaw plan_fetch(const FusedParams &p, const uint32_t u) {
  const uint4 *q = reinterpret_cast<const uint4 *>(reinterpret_cast<const PlanRec *>(p.plans) + u);
  PlanRaw r;
  r.a = __ldg(q); r.b = __ldg(q + 1);
  return r;
}
template <bool PLAN>
__device__ __forceinline__ UnitPlan plan_make(const FusedParams &p, const uint32_t u, const PlanRaw &r) {
  UnitPlan pl;
  if (PLAN) {   // see the record layout above unit_class()
    unit_geom(p, u, pl, r.a.z);
    pl.iv = 0; pl.iv_left = 0; pl.u = u; pl.src_kind = r.b.x;
    pl.fast_src = reinterpret_cast<const uint8_t *>(((unsigned long long)r.a.y << 32) | r.a.x);   // PK_COPY, PK_LEFT
    pl.fast_pitch = r.a.w; pl.list = nullptr; pl.n_list = 0;
    pl.tma_ok = (r.b.x & PK_TMA) != 0u; pl.halo_fast = pl.tma_ok; pl.left_fast = (r.b.x & PK_LEFT) != 0u;
    pl.tile = r.b.y; pl.c0 = (int32_t)r.b.z; pl.c1 = (int32_t)r.b.w;
    return pl;
  }
  unit_geom(p, u, pl);
  pl.iv = 0; pl.iv_left = 0; pl.u = u; pl.src_kind = 0;
  pl.fast_src = reinterpret_cast<const uint8_t *>(((unsigned long long)r.a.y << 32) | r.a.x);
  pl.fast_pitch = r.a.z;
  pl.list = p.bin_tiles + r.a.w;
  pl.n_list = r.b.x >> 3;
  pl.tma_ok = (r.b.x & 4u) != 0u;
  pl.halo_fast = (r.b.x & 1u) != 0u;
  pl.left_fast = (r.b.x & 2u) != 0u;
  pl.tile = r.b.y; pl.c0 = (int32_t)r.b.z; pl.c1 = (int32_t)r.b.w;
  return pl;
}

// ---- plan mode: the host's interval plan (strip_plan.cu) instead of bins --------------------------------------
// A unit is one 128-word column of the plan.  Three kernels: fused_plan_iv_kernel classifies the units of a band
// (one CTA per band) and writes their records; fused_prep_kernel (a warp per unit) normalises and composites the
// rows of the units that are not a plain window of one tile, and the left pixels that cannot be copied from the
// unit's own tile, into scratch; the fused kernel then gets every unit by one bulk copy (or a few cp.async) and
// every unit costs it the same -- nobody waits for a warp that is still blending.
//
// Record of unit u (32 bytes, as two uint4 a, b); b.x = flags:
//   PK_TMA     the unit and its predictor row are a window of one tile that has tensor maps: b.y = tile,
//              b.z, b.w = box coordinates
//   PK_COPY    the same without tensor maps or with a source offset that is not a multiple of 16 bytes (a tiled TMA
//              load faults on it): the lanes copy the window with cp.async; a.x, a.y = address of the window's
//              first word, a.w = pitch
//   PK_BG      background only: filled by the lanes
//   none       prepared rows: one bulk copy of the unit's slot in unit_rows
//   PK_LEFT    the pixel left of the unit comes from the same source (the tile's row, or background); with PK_TMA,
//              a.x, a.y, a.w are then as for PK_COPY.  Otherwise it is read from unit_left.
// a.z = the unit's column.  Where a.x, a.y hold no address they are the intervals of the unit's column and of the
// column left of it that hold the predictor row (for the prepare kernel); PK_COPY keeps the latter in b.y.

static constexpr int FU_PREP_WORDS = (FU_R + 1) * FU_SW;   // a unit's slot of prepared rows
static constexpr int FU_LEFT_WORDS = (FU_R + 1) * 2;       // a unit's left pixels
static constexpr int FP_ROWS = 1;           // prepare kernel: rows per task
static constexpr uint32_t FP_LEFT = 15u;   // prepare kernel: a task is unit << 4 | part; parts 0 .. (FU_R + 1) / rows - 1 are rows, FP_LEFT the left pixels

struct IvCur {
  uint32_t idx, end_idx;
  uint32_t y_end, kind, tile, list_off, n_list, flags;
  int32_t cw, cy;
  const uint8_t *pix;   // SK_DIRECT, SK_SHIFT: the tile's rows
  uint32_t pitch;
};
__device__ __forceinline__ void iv_load(const FusedParams &p, IvCur &c) {
  const uint4 *q = reinterpret_cast<const uint4 *>(p.ivs) + 2u * c.idx;
  const uint4 a = __ldg(q), b = __ldg(q + 1);
  c.y_end = a.x; c.kind = a.y; c.tile = a.z; c.cw = (int32_t)a.w;
  c.cy = (int32_t)b.x; c.list_off = b.y; c.n_list = b.z; c.flags = b.w;
  if (c.kind == SK_DIRECT || c.kind == SK_SHIFT) {
    const DevTile *t = p.tiles + c.tile;
    c.pix = t->pix; c.pitch = t->pitch;
  }
}
__device__ __forceinline__ void iv_seek(const FusedParams &p, IvCur &c, int y) {
  while ((uint32_t)y >= c.y_end && c.idx + 1u < c.end_idx) { c.idx++; iv_load(p, c); }
}
// index of the interval of column col that holds row y
__device__ __forceinline__ uint32_t iv_find(const FusedParams &p, uint32_t col, uint32_t y) {
  uint32_t lo = __ldg(p.col_start + col), hi = __ldg(p.col_start + col + 1) - 1u;
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (__ldg(reinterpret_cast<const uint32_t *>(p.ivs) + 8u * mid) > y) hi = mid; else lo = mid + 1u;
  }
  return lo;
}

// The record of the unit (column seg, rows [ya, yb)); returns its class for the dealing order:
// 0: some row is composited, 1: prepared from copies and background, 2: needs no preparation.
__device__ __forceinline__ uint32_t unit_class(const FusedParams &p, uint32_t seg, int ya, int yb, PlanRaw &r) {
  const uint32_t y_first = (uint32_t)max(ya - 1, 0);
  IvCur c;
  c.idx = iv_find(p, seg, y_first); c.end_idx = __ldg(p.col_start + seg + 1);
  iv_load(p, c);
  const uint32_t iv_left = seg ? iv_find(p, seg - 1u, y_first) : 0u;
  r.a = make_uint4(c.idx, iv_left, seg, 0u);
  r.b = make_uint4(0u, 0u, 0u, 0u);
  const bool whole = c.y_end >= (uint32_t)yb;   // one interval holds the predictor row and the unit
  // does the column left of the unit show the same source in these rows?  (column 0: the pixel is 0, read from unit_left)
  bool left_same = false;
  if (seg && whole) {
    const uint4 *q = reinterpret_cast<const uint4 *>(p.ivs) + 2u * iv_left;
    const uint4 la = __ldg(q);
    left_same = la.x >= (uint32_t)yb && (c.kind == SK_BG ? la.y == SK_BG
                                                           : (la.y == SK_DIRECT || la.y == SK_SHIFT) && la.z == c.tile);
  }
  if (whole && c.kind == SK_BG) {
    r.b.x = PK_BG | (left_same ? PK_LEFT : 0u);
    return 2u;
  }
  if (whole && ya > 0 && (c.kind == SK_DIRECT || c.kind == SK_SHIFT) && c.list_off == (uint32_t)FU_SW) {
    const unsigned long long src = reinterpret_cast<unsigned long long>(c.pix + (size_t)(ya - 1 + c.cy) * c.pitch) + 4ull * (unsigned)c.cw;
    if (c.kind == SK_DIRECT && p.maps != nullptr && p.tiles[c.tile].pad != 0u) {
      r.b = make_uint4(PK_TMA, c.tile, (uint32_t)c.cw, (uint32_t)(ya - 1 + c.cy));
      if (left_same) { r.a.x = (uint32_t)src; r.a.y = (uint32_t)(src >> 32); r.a.w = c.pitch; r.b.x |= PK_LEFT; }
    } else {
      r.a.x = (uint32_t)src; r.a.y = (uint32_t)(src >> 32); r.a.w = c.pitch;
      r.b.x = PK_COPY | (left_same ? PK_LEFT : 0u); r.b.y = iv_left;
    }
    return 2u;
  }
  uint32_t cls = 1u;
  for (;;) {
    if (c.kind == SK_COMP) cls = 0u;
    if (c.y_end >= (uint32_t)yb || c.idx + 1u >= c.end_idx) break;
    c.idx++; iv_load(p, c);
  }
  return cls;
}

template <int BPW>
__global__ void __launch_bounds__(256) fused_plan_iv_kernel(const FusedParams p, PlanRec *recs, const uint32_t FP_PARTS) {
  __shared__ uint32_t cnt[4], pos[4], list_base;
  const uint32_t band = blockIdx.x, band_abs = p.band0 + band;
  const int ya = max((int)p.y_begin, (int)(band_abs * FU_R)), yb = min((int)p.y_end, (int)((band_abs + 1) * FU_R));
  if (threadIdx.x < 4) { cnt[threadIdx.x] = 0; pos[threadIdx.x] = 0; }
  __syncthreads();
  // A band's units are dealt out in record order: those that need preparing go first (composited rows ahead of
  // copied ones, then those that only need their left pixels), so that they form one range of the band's records.
  PlanRaw r;
  uint32_t cls0 = 0;   // this thread's first unit: classified once
  for (uint32_t seg = threadIdx.x, i = 0; seg < p.nseg; seg += blockDim.x, i++) {
    uint32_t cls = unit_class(p, seg, ya, yb, r);
    if (cls == 2u && (r.b.x & PK_LEFT)) cls = 3u;
    if (i == 0) cls0 = cls;
    atomicAdd(&cnt[cls], 1u);
  }
  const PlanRaw r0 = r;
  __syncthreads();
  const uint32_t base[4] = {0u, cnt[0], cnt[0] + cnt[1], cnt[0] + cnt[1] + cnt[2]};
  const uint32_t n_prep = base[3], first = band * p.nseg;
  if (threadIdx.x == 0) {
    const uint32_t nt = base[2] * (FP_PARTS + 1) + (n_prep - base[2]);
    list_base = nt ? atomicAdd(p.prep_count, nt) : 0u;
  }
  uint4 *out = reinterpret_cast<uint4 *>(recs + first);
  for (uint32_t seg = threadIdx.x, i = 0; seg < p.nseg; seg += blockDim.x, i++) {
    uint32_t cls = cls0;
    if (p.nseg > blockDim.x) {
      cls = unit_class(p, seg, ya, yb, r);
      if (cls == 2u && (r.b.x & PK_LEFT)) cls = 3u;
    } else r = r0;
    const uint32_t at = base[cls] + atomicAdd(&pos[cls], 1u);
    out[2u * at] = r.a; out[2u * at + 1u] = r.b;
  }
  __syncthreads();
  // tasks of the band: classes 0 and 1 (the first n_rows_units records) have FP_PARTS row parts and the left pixels,
  // class 2 the left pixels only
  const uint32_t n_rows_units = base[2], n_tasks = n_rows_units * (FP_PARTS + 1) + (n_prep - n_rows_units);
  for (uint32_t j = threadIdx.x; j < n_tasks; j += blockDim.x) {
    uint32_t task;
    if (j < n_rows_units * (FP_PARTS + 1)) {
      const uint32_t part = j % (FP_PARTS + 1);
      task = ((first + j / (FP_PARTS + 1)) << 4) | (part == FP_PARTS ? FP_LEFT : part);
    } else task = ((first + n_rows_units + (j - n_rows_units * (FP_PARTS + 1))) << 4) | FP_LEFT;
    p.prep_list[list_base + j] = task;
  }
}

struct FuCtx {   // what the generic compositor reads, by value (see compose_quad)
  const DevTile *tiles;
  uint32_t alpha_blend, bg_lo, bg_hi, canvas_w;
};
template <int BPW>
static __device__ __noinline__ uint4 fu_compose_quad(const FuCtx c, const uint32_t *list, uint32_t n_list, int xw, int y) {
  return compose_quad<BPW>(c, list, n_list, xw, y);
}

static constexpr int FP_LAYERS = 2;                        // layers of a row loaded together

// One layer's words under this lane's quad: which of them the layer covers (bit j: word j), and their values.
template <int BPW>
__device__ __forceinline__ uint32_t layer_fetch(const DevTile *t, const int x, const int y, const int n_canvas, uint4 &s) {
  constexpr int PX = 4 / BPW;
  const int tx0 = t->x0, tx1 = t->x1;
  uint32_t m = 0u;
#pragma unroll
  for (int i = 0; i < PX; i++)
    if (x + i >= tx0 && x + i < tx1 && i * BPW < n_canvas) m |= (BPW == 2 ? 3u : 1u) << (i * BPW);
  s = make_uint4(0u, 0u, 0u, 0u);
  if (m) {
    const uint32_t *src = reinterpret_cast<const uint32_t *>(t->pix + (size_t)(y - t->y0 + t->sy0) * t->pitch) + (x - tx0 + t->sx0) * BPW;
    if (m == 15u && (reinterpret_cast<uintptr_t>(src) & 15u) == 0) s = __ldg(reinterpret_cast<const uint4 *>(src));
    else {
      if (m & 1u) s.x = __ldg(src);
      if (m & 2u) s.y = __ldg(src + 1);
      if (m & 4u) s.z = __ldg(src + 2);
      if (m & 8u) s.w = __ldg(src + 3);
    }
  }
  return m;
}
// The blend of one layer into a quad: out of line, so that the kernel holds one copy of the arithmetic (with a
// copy per layer and word the prepare kernel spent half its time waiting for instructions).
template <int BPW>
static __device__ __noinline__ uint4 layer_blend(uint4 v, const uint4 s, const uint32_t m) {
  if (BPW == 1) {
    if (m & 1u) v.x = blend8(v.x, s.x);
    if (m & 2u) v.y = blend8(v.y, s.y);
    if (m & 4u) v.z = blend8(v.z, s.z);
    if (m & 8u) v.w = blend8(v.w, s.w);
  } else {
    if (m & 1u) { const uint2 a = blend16(make_uint2(v.x, v.y), make_uint2(s.x, s.y)); v.x = a.x; v.y = a.y; }
    if (m & 4u) { const uint2 a = blend16(make_uint2(v.z, v.w), make_uint2(s.z, s.w)); v.z = a.x; v.w = a.y; }
  }
  return v;
}
template <int BPW>
__device__ __forceinline__ void layer_apply(uint4 &v, const uint4 s, const uint32_t m, const bool overwrite) {
  if (m == 0u) return;
  if (overwrite) {
    if (m & 1u) v.x = s.x;
    if (m & 2u) v.y = s.y;
    if (m & 4u) v.z = s.z;
    if (m & 8u) v.w = s.w;
  } else {
    v = layer_blend<BPW>(v, s, m);
  }
}

// Row y (>= 0) of the unit's column, composed from the interval c that holds it: one quad per lane.  All loads of
// a row (up to FP_LAYERS layers at a time) are issued before the first blend.
template <int BPW>
__device__ __forceinline__ uint4 unit_compose_row(const FusedParams &p, const IvCur &c, const int xw0, const int y, const uint32_t lane) {
  const int wmax = (int)p.canvas_w * BPW, qw = xw0 + 4 * (int)lane;
  if (qw >= wmax) return make_uint4(0u, 0u, 0u, 0u);   // nothing right of the canvas
  const int n_canvas = min(4, wmax - qw);               // words of this lane's quad that exist in the canvas
  uint4 v = BPW == 2 ? make_uint4(p.bg_lo, p.bg_hi, p.bg_lo, p.bg_hi) : make_uint4(p.bg_lo, p.bg_lo, p.bg_lo, p.bg_lo);
  if (c.kind == SK_DIRECT || c.kind == SK_SHIFT) {
    const uint32_t *src = reinterpret_cast<const uint32_t *>(c.pix + (size_t)(y + c.cy) * c.pitch) + c.cw + 4 * (int)lane;
    const int n = min(n_canvas, (int)c.list_off - 4 * (int)lane);   // words that exist in the source row too
    if (n == 4 && (reinterpret_cast<uintptr_t>(src) & 15u) == 0) v = __ldg(reinterpret_cast<const uint4 *>(src));
    else { v.x = n > 0 ? __ldg(src) : 0u; v.y = n > 1 ? __ldg(src + 1) : 0u; v.z = n > 2 ? __ldg(src + 2) : 0u; v.w = n > 3 ? __ldg(src + 3) : 0u; }
  } else if (c.kind == SK_COMP && (c.flags & 1u)) {
    // every layer spans the column in the output's format: layer after layer, no per-pixel tests
    for (uint32_t k = 0; k < c.n_list; k++) {
      const DevTile *t = p.tiles + __ldg(p.lists + c.list_off + k);
      const uint32_t *src = reinterpret_cast<const uint32_t *>(t->pix + (size_t)(y - t->y0 + t->sy0) * t->pitch) +
                            (xw0 / BPW - t->x0 + t->sx0) * BPW + 4 * (int)lane;
      uint4 s;
      if (n_canvas < 4) {   // the canvas (and possibly the tile's row) ends inside this quad: no loads past it
        s.x = __ldg(src); s.y = n_canvas > 1 ? __ldg(src + 1) : 0u; s.z = n_canvas > 2 ? __ldg(src + 2) : 0u; s.w = 0u;
      } else if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
        s = __ldg(reinterpret_cast<const uint4 *>(src));
      } else {
        s.x = __ldg(src); s.y = __ldg(src + 1); s.z = __ldg(src + 2); s.w = __ldg(src + 3);
      }
      if (p.alpha_blend == 0u || t->opaque) v = s;
      else v = layer_blend<BPW>(v, s, 15u);
    }
  } else if (c.kind == SK_COMP && (c.flags & 2u)) {
    // every layer is in the output's format: word loads under a per-lane coverage mask
    const int x = qw / BPW;
    for (uint32_t k0 = 0; k0 < c.n_list; k0 += FP_LAYERS) {
      uint4 s[FP_LAYERS];
      uint32_t m[FP_LAYERS], ow = 0u;
#pragma unroll
      for (int k = 0; k < FP_LAYERS; k++) {
        m[k] = 0u;
        if (k0 + k < c.n_list) {
          const DevTile *t = p.tiles + __ldg(p.lists + c.list_off + k0 + k);
          if (p.alpha_blend == 0u || t->opaque) ow |= 1u << k;
          m[k] = layer_fetch<BPW>(t, x, y, n_canvas, s[k]);
        }
      }
#pragma unroll
      for (int k = 0; k < FP_LAYERS; k++) layer_apply<BPW>(v, s[k], m[k], (ow >> k) & 1u);
    }
  } else if (c.kind == SK_COMP) {
    FuCtx cc;
    cc.tiles = p.tiles; cc.alpha_blend = p.alpha_blend; cc.bg_lo = p.bg_lo; cc.bg_hi = p.bg_hi; cc.canvas_w = p.canvas_w;
    v = fu_compose_quad<BPW>(cc, p.lists + c.list_off, c.n_list, qw, y);
  }
  if (n_canvas < 4) {
    if (n_canvas < 2) v.y = 0u;
    if (n_canvas < 3) v.z = 0u;
    v.w = 0u;
  }
  return v;
}

// The pixel just left of the unit in rows ya-1 .. yb-1, through the plan of the column left of it (lane r: row
// ya-1+r), into the unit's 2 (FU_R + 1) words of unit_left.
template <int BPW>
__device__ __forceinline__ void unit_left_plan(const FusedParams &p, const UnitPlan &pl, uint32_t *lcs, const uint32_t lane) {
  if ((int)lane > pl.nrows) return;
  const int y = pl.ya - 1 + (int)lane;
  uint32_t c0 = 0u, c1 = 0u;
  if (y >= 0 && pl.seg > 0u) {
    IvCur h;
    h.idx = pl.iv_left; h.end_idx = __ldg(p.col_start + pl.seg);
    iv_load(p, h);
    iv_seek(p, h, y);
    if (h.kind == SK_BG) {
      c1 = BPW == 2 ? p.bg_hi : p.bg_lo; c0 = p.bg_lo;
    } else if (h.kind == SK_DIRECT || h.kind == SK_SHIFT) {
      const uint32_t *lp = reinterpret_cast<const uint32_t *>(h.pix + (size_t)(y + h.cy) * h.pitch) + h.cw + FU_SW;
      c1 = __ldg(lp - 1);
      if (BPW == 2) c0 = __ldg(lp - 2);
    } else {
      FuCtx cc;
      cc.tiles = p.tiles; cc.alpha_blend = p.alpha_blend; cc.bg_lo = p.bg_lo; cc.bg_hi = p.bg_hi; cc.canvas_w = p.canvas_w;
      const uint4 v = fu_compose_quad<BPW>(cc, p.lists + h.list_off, h.n_list, pl.xw0 - 4, y);
      c1 = v.w; c0 = v.z;
    }
  }
  lcs[lane * 2] = BPW == 2 ? c0 : 0u;
  lcs[lane * 2 + 1] = c1;
}

// The prepare kernel: a fixed grid of warps over the tasks the plan kernel listed.  A task is a unit and a part:
// part k < FP_PARTS = rows ya-1+FP_ROWS k .. of the unit into its slot of unit_rows (FU_R + 1 rows of 128 words:
// what the TMA box of a plain unit looks like), part FP_PARTS = the unit's left pixels.  Rows are split into parts
// so that the machine is full of short, independent load chains.
template <int BPW, int ROWS, int MINB>
__global__ void __launch_bounds__(256, MINB) fused_prep_kernel(const FusedParams p) {
  const uint32_t lane = threadIdx.x & 31u, n_warps = gridDim.x * (blockDim.x >> 5);
  const uint32_t n_tasks = __ldg(p.prep_count);
  for (uint32_t t = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); t < n_tasks; t += n_warps) {
    const uint32_t task = __ldg(p.prep_list + t), u = task >> 4, part = task & 15u;
    const PlanRaw r = plan_fetch(p, u);
    const uint32_t flags = r.b.x;
    UnitPlan pl;
    unit_geom(p, u, pl, r.a.z);
    if (part == FP_LEFT) {
      pl.iv_left = (flags & PK_COPY) ? r.b.y : r.a.y;
      unit_left_plan<BPW>(p, pl, p.unit_left + (size_t)u * FU_LEFT_WORDS, lane);
      continue;
    }
    IvCur c;
    c.idx = r.a.x; c.end_idx = __ldg(p.col_start + pl.seg + 1);
    iv_load(p, c);
    const int r0 = (int)part * ROWS, r1 = min(r0 + ROWS, pl.nrows + 1);
    uint32_t *dst = p.unit_rows + (size_t)u * FU_PREP_WORDS + r0 * FU_SW + 4u * lane;
    for (int rr = r0; rr < r1; rr++, dst += FU_SW) {
      const int y = pl.ya - 1 + rr;
      uint4 v = make_uint4(0u, 0u, 0u, 0u);   // previousLine === null above row 0
      if (y >= 0) {
        iv_seek(p, c, y);
        v = unit_compose_row<BPW>(p, c, pl.xw0, y, lane);
      }
      *reinterpret_cast<uint4 *>(dst) = v;
    }
  }
}

// ---- phase A: raw stash <- composited pixels of rows ya-1 .. yb-1 (cp.async left in flight) ----
template <int BPW, bool PLAN>
__device__ __forceinline__ bool unit_load(const FusedParams &p, const UnitPlan &pl, uint32_t *stash, const uint32_t bar,
                                          const uint32_t lane) {
  uint32_t *raw = stash, *lcs = stash + FU_RAW_WORDS;
  const int wmax = (int)p.canvas_w * BPW;  // words per canvas row
  const int ya = pl.ya, nrows = pl.nrows, xw0 = pl.xw0, qw = xw0 + 4 * (int)lane;
  const uint8_t *fast_src = pl.fast_src;
  const uint32_t fast_pitch = pl.fast_pitch;
  // A plain window of one tile, predictor row included: ONE 2-D TMA box (128 words x 9 rows) from the tile's tensor
  // map, issued by lane 0, completing on the stash's mbarrier.  (Rows past the unit and words past the canvas or
  // the tile are loaded or zero-filled and never used.)
  if (PLAN) {
    // one TMA box of the unit's tile, one bulk copy of its prepared rows, cp.async by the lanes, or background
    const uint32_t k = pl.src_kind;
    const bool bulk = (k & (PK_BG | PK_COPY)) == 0u;
    if (bulk) {
      if (lane == 0u) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the emit's reads of this stash, then the bulk write
        mbar_arrive_tx(bar, FU_PREP_WORDS * 4);
        if (k & PK_TMA)
          tma_box_2d((uint32_t)__cvta_generic_to_shared(raw), reinterpret_cast<const uint8_t *>(p.maps) + (size_t)pl.tile * 384u + 256u,
                     pl.c0, pl.c1, bar);
        else
          bulk_copy((uint32_t)__cvta_generic_to_shared(raw), p.unit_rows + (size_t)pl.u * FU_PREP_WORDS, FU_PREP_WORDS * 4, bar);
      }
    } else if (k & PK_COPY) {
      const uint8_t *src = fast_src + 16u * lane;
      const uint32_t sa = (uint32_t)__cvta_generic_to_shared(raw + 4 * lane);
      if (((reinterpret_cast<uintptr_t>(fast_src) | fast_pitch) & 15u) == 0) {
        for (int rr = 0; rr <= nrows; rr++, src += fast_pitch) fu_cp_async16s(sa + (uint32_t)rr * (FU_SW * 4), src);
      } else {
        for (int rr = 0; rr <= nrows; rr++, src += fast_pitch) {
#pragma unroll
          for (int j = 0; j < 4; j++) fu_cp_async4(sa + (uint32_t)rr * (FU_SW * 4) + 4u * j, src + 4 * j);
        }
      }
    } else {
      uint4 v = make_uint4(0u, 0u, 0u, 0u);
      if (qw < wmax) {
        v = BPW == 2 ? make_uint4(p.bg_lo, p.bg_hi, p.bg_lo, p.bg_hi) : make_uint4(p.bg_lo, p.bg_lo, p.bg_lo, p.bg_lo);
        if (qw + 1 >= wmax) v.y = 0u;
        if (qw + 2 >= wmax) v.z = 0u;
        if (qw + 3 >= wmax) v.w = 0u;
      }
      uint32_t *dst = raw + 4 * lane;
      *reinterpret_cast<uint4 *>(dst) = ya == 0 ? make_uint4(0u, 0u, 0u, 0u) : v;   // previousLine === null
      for (int rr = 0; rr < nrows; rr++) { dst += FU_SW; *reinterpret_cast<uint4 *>(dst) = v; }
    }
    if ((int)lane <= nrows) {   // the pixel left of the unit, rows ya-1 .. yb-1
      const uint32_t sa = (uint32_t)__cvta_generic_to_shared(lcs + 2 * lane);
      if (!(k & PK_LEFT)) {
        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(p.unit_left + (size_t)pl.u * FU_LEFT_WORDS + 2 * lane) : "memory");
      } else if (k & PK_BG) {
        const bool none = ya == 0 && lane == 0u;
        lcs[2 * lane] = none || BPW == 1 ? 0u : p.bg_lo;
        lcs[2 * lane + 1] = none ? 0u : (BPW == 2 ? p.bg_hi : p.bg_lo);
      } else {
        const uint8_t *lp = fast_src + (size_t)lane * fast_pitch;
        if (BPW == 2) asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(sa), "l"(lp - 8) : "memory");
        else fu_cp_async4(sa + 4u, lp - 4);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    return bulk;
  }
  const bool tma = pl.tma_ok && pl.halo_fast && ya > 0;
  if (tma) {
    if (lane == 0u) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the emit's reads of this stash, then the bulk write
      mbar_arrive_tx(bar, (FU_R + 1) * FU_SW * 4);
      tma_box_2d((uint32_t)__cvta_generic_to_shared(raw), reinterpret_cast<const uint8_t *>(p.maps) + (size_t)pl.tile * 384u + 256u,
                 pl.c0, pl.c1, bar);
    }
  } else {
    uint32_t *dst = raw + 4 * lane;
    const bool inside = qw < wmax;
    // predictor row
    if (ya == 0 || !inside) *reinterpret_cast<uint4 *>(dst) = make_uint4(0u, 0u, 0u, 0u);  // previousLine === null
    else if (fast_src && pl.halo_fast) fu_cp_async16(dst, fast_src - fast_pitch + 16u * lane);
    else *reinterpret_cast<uint4 *>(dst) = compose_quad<BPW>(p, pl.list, pl.n_list, qw, ya - 1);
    dst += FU_SW;
    if (!inside) {
      for (int rr = 0; rr < nrows; rr++, dst += FU_SW) *reinterpret_cast<uint4 *>(dst) = make_uint4(0u, 0u, 0u, 0u);
    } else if (fast_src) {
      const uint8_t *src = fast_src + 16u * lane;
      for (int rr = 0; rr < nrows; rr++, dst += FU_SW, src += fast_pitch) fu_cp_async16(dst, src);
    } else {
      for (int rr = 0; rr < nrows; rr++, dst += FU_SW)
        *reinterpret_cast<uint4 *>(dst) = compose_quad<BPW>(p, pl.list, pl.n_list, qw, ya + rr);
    }
  }
  if ((int)lane <= nrows) {  // the pixel just left of the unit (0 at the canvas edge)
    const int y = ya - 1 + (int)lane;
    uint32_t c0 = 0u, c1 = 0u;
    if (y >= 0 && xw0 > 0) {
      if (fast_src && pl.left_fast && (lane > 0 || pl.halo_fast)) {
        const uint32_t *lp = reinterpret_cast<const uint32_t *>(fast_src + (ptrdiff_t)((int)lane - 1) * (ptrdiff_t)fast_pitch);
        c1 = __ldg(lp - 1);
        if (BPW == 2) c0 = __ldg(lp - 2);
      } else {
        const uint4 v = compose_quad<BPW>(p, pl.list, pl.n_list, (xw0 - BPW) & ~3, y);
        c1 = v.w;
        if (BPW == 2) c0 = v.z;
      }
    }
    lcs[lane * 2] = c0;
    lcs[lane * 2 + 1] = c1;
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  return tma;
}

// ---- sweep: filter costs of every row, Paeth residuals into the stash, partial sums published ----
template <int BPW>
__device__ __forceinline__ void unit_sweep(const FusedParams &p, const UnitPlan &pl, uint32_t *stash, const uint32_t pf_taddr,
                                           const uint32_t lane, const uint32_t (&lmask)[5], const bool tma, const uint32_t bar,
                                           const uint32_t parity) {
  uint32_t *raw = stash, *lcs = stash + FU_RAW_WORDS;
  const int wmax = (int)p.canvas_w * BPW;
  const int nrows = pl.nrows, qw = pl.xw0 + 4 * (int)lane;
  const bool guard = pl.xw0 + FU_SW > wmax;
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  if (tma) mbar_wait(bar, parity);
  __syncwarp();
  uint4 up = *reinterpret_cast<const uint4 *>(raw + 4 * lane);
  uint32_t l0 = __shfl_up_sync(0xffffffffu, up.z, 1), l1 = __shfl_up_sync(0xffffffffu, up.w, 1);
  if (lane == 0u) { l0 = lcs[0]; l1 = lcs[1]; }
  H4 hA[4], hlA[2], hB[4], hlB[2];  // two unpacked rows, used alternately as "above" and "current"
  hA[0] = unpack_h4(up.x); hA[1] = unpack_h4(up.y); hA[2] = unpack_h4(up.z); hA[3] = unpack_h4(up.w);
  hlA[0] = unpack_h4(l0); hlA[1] = unpack_h4(l1);
  const uint32_t *raw_row = raw + FU_SW;
  const uint32_t *lcs_row = lcs + 2;
  uint32_t pf_row = pf_taddr;
  unsigned long long *rs_row = p.rowsum + (size_t)(pl.ya - (int)p.y_begin) * 8;
  int r = 0;
  for (; r + 2 <= nrows; r += 2) {
    sweep_step<BPW>(raw_row, lcs_row, pf_row, rs_row, lane, lmask, guard, qw, wmax, up, hA, hlA, hB, hlB);
    sweep_step<BPW>(raw_row + FU_SW, lcs_row + 2, pf_row + 4, rs_row + 8, lane, lmask, guard, qw, wmax, up, hB, hlB,
                    hA, hlA);
    raw_row += 2 * FU_SW; lcs_row += 4; pf_row += 8; rs_row += 16;
  }
  if (r < nrows) sweep_step<BPW>(raw_row, lcs_row, pf_row, rs_row, lane, lmask, guard, qw, wmax, up, hA, hlA, hB, hlB);
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// What phase B needs to know about a unit after its sweep is over.
struct UnitRef {
  uint32_t seg;
  int ya, nrows;
};

// ---- phase B, part 1: wait until every unit of the band has added its sums; lane r gets row r's winner ----
// Each 64-bit row-sum word carries its own arrival count in the high half, so the sums need no fence
// and no separate "band done" flag: lane r reads row r's five words until all five are complete.
__device__ __forceinline__ uint32_t unit_wait(const FusedParams &p, const UnitRef &ref, const uint32_t lane) {
  uint32_t my_ft = 0u;
  for (;;) {
    bool ready = true;
    if ((int)lane < ref.nrows) {
      const unsigned long long *rs = p.rowsum + (size_t)(ref.ya + (int)lane - (int)p.y_begin) * 8;
      unsigned long long w0, w1, w2, w3, w4;
      asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(rs) : "memory");
      asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w2), "=l"(w3) : "l"(rs + 2) : "memory");
      asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w4) : "l"(rs + 4) : "memory");
      const uint32_t n = p.nseg;
      ready = (uint32_t)(w0 >> 32) == n && (uint32_t)(w1 >> 32) == n && (uint32_t)(w2 >> 32) == n &&
              (uint32_t)(w3 >> 32) == n && (uint32_t)(w4 >> 32) == n;
      uint32_t best = (uint32_t)w0;
      my_ft = 0u;
      if ((uint32_t)w1 < best) { best = (uint32_t)w1; my_ft = 1u; }   // strict '<': first minimum wins (png-filter.ts:174)
      if ((uint32_t)w2 < best) { best = (uint32_t)w2; my_ft = 2u; }
      if ((uint32_t)w3 < best) { best = (uint32_t)w3; my_ft = 3u; }
      if ((uint32_t)w4 < best) { best = (uint32_t)w4; my_ft = 4u; }
    }
    if (__all_sync(0xffffffffu, ready)) break;
    __nanosleep(100);
  }
  return my_ft;
}

// Winners of rows 0..FU_R-1 (lane r holds row r's), 3 bits each, as a warp-uniform word.
__device__ __forceinline__ uint32_t pack_winners(const uint32_t my_ft, const uint32_t lane) {
  return __reduce_or_sync(0xffffffffu, lane < (uint32_t)FU_R ? my_ft << (3u * lane) : 0u);
}

// ---- phase B, part 2: rows [r0, r1) of the unit: the winner's bytes, one 128-bit store per lane per row ----
// ftpack: the winners of the unit's rows, 3 bits each, the same in every lane (so every branch is uniform).
template <int BPW>
__device__ __forceinline__ void unit_emit(const FusedParams &p, const UnitRef &ref, const uint32_t *stash,
                                          const uint32_t pf_taddr, const uint32_t lane, const uint32_t my_ft,
                                          const uint32_t ftpack, const int r0, const int r1) {
  const uint32_t *raw = stash, *lcs = stash + FU_RAW_WORDS;
  const int wmax = (int)p.canvas_w * BPW;
  const int qw = (int)ref.seg * FU_SW + 4 * (int)lane;
  uint8_t *orow = p.out + (size_t)(ref.ya + r0 - (int)p.y_begin) * p.out_pitch + p.lead;
  // the filter-type bytes of the rows: lane r holds row r's
  if (ref.seg == 0u && (int)lane >= r0 && (int)lane < r1) orow[(size_t)((int)lane - r0) * p.out_pitch] = (uint8_t)my_ft;
  uint8_t *odat = orow + 1 + 4 * (size_t)qw;
  const uint32_t *cur_row = raw + (r0 + 1) * FU_SW + 4 * lane;
  const bool full = ((int)ref.seg + 1) * FU_SW <= wmax;  // every lane stores a whole quad
  for (int r = r0; r < r1; r++, odat += p.out_pitch, cur_row += FU_SW) {
    const uint32_t ft = (ftpack >> (3 * r)) & 7u;
    uint4 o;
    if (ft == 4u) {
      o = tmem_ld4(pf_taddr + 4u * (uint32_t)r);
    } else {
      o = *reinterpret_cast<const uint4 *>(cur_row);
      if (ft != 0u) {
        uint4 pr;  // the predictor bytes
        if (ft == 2u) {
          pr = *reinterpret_cast<const uint4 *>(cur_row - FU_SW);
        } else {
          uint32_t c0 = BPW == 2 ? __shfl_up_sync(0xffffffffu, o.z, 1) : 0u;
          uint32_t c1 = __shfl_up_sync(0xffffffffu, o.w, 1);
          if (lane == 0u) { c0 = lcs[2 * (r + 1)]; c1 = lcs[2 * (r + 1) + 1]; }
          pr = BPW == 1 ? make_uint4(c1, o.x, o.y, o.z) : make_uint4(c0, c1, o.x, o.y);
          if (ft == 3u) {
            const uint4 upq = *reinterpret_cast<const uint4 *>(cur_row - FU_SW);
            pr = make_uint4(avg4(pr.x, upq.x), avg4(pr.y, upq.y), avg4(pr.z, upq.z), avg4(pr.w, upq.w));
          }
        }
        o = make_uint4(sub4(o.x, pr.x), sub4(o.y, pr.y), sub4(o.z, pr.z), sub4(o.w, pr.w));
      }
    }
    if (full) {
      *reinterpret_cast<uint4 *>(odat) = o;
    } else if (qw < wmax) {
      uint32_t *od = reinterpret_cast<uint32_t *>(odat);
      od[0] = o.x;
      if (qw + 1 < wmax) od[1] = o.y;
      if (qw + 2 < wmax) od[2] = o.z;
      if (qw + 3 < wmax) od[3] = o.w;
    }
  }
}

// Software pipeline per warp (two stashes), units handed out by an atomic ticket:
//
//   sweep(u_k) -> wait(u_k-1) -> take ticket u_k+1 -> emit(u_k-1), the ticket and the plan record of
//   u_k+1 arriving meanwhile -> load(u_k+1) into the stash just emitted -> sweep(u_k+1) ...
//
// so the band of u_k-1 has had a whole sweep's time to complete and wait() normally returns at once.
//
// No wait cycle can form.  A warp blocks only in wait(), and whenever it does every ticket it has taken is
// already swept (it takes the next one after wait() returns and sweeps it before it can block again: emit,
// load and sweep never wait).  Suppose every running warp were blocked, and let B be the lowest band any
// of them waits for.  Each blocked warp holds two tickets (u_k-1 and u_k) in bands >= B.  If some ticket of B
// had not been taken yet, all held tickets would lie in B itself, but 2 x (running warps) > units per band
// (launch_fused checks it; the cooperative launch makes the whole grid resident).  So all of B is taken;
// its units are swept (held by blocked or finished warps), B is complete, and the warps waiting for it
// are not blocked after all.
template <int BPW, bool PLAN>
__global__ void __launch_bounds__(FU_WARPS * 32, 1) fused_kernel(const FusedParams p) {
  extern __shared__ __align__(16) uint32_t fu_smem[];
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  uint32_t *cur_stash = fu_smem + warp * (FU_SLOTS * FU_WARP_WORDS), *prev_stash = cur_stash + FU_WARP_WORDS;
  // tensor memory for the residual stash: warp 0 allocates for the CTA, every warp uses its own lane quarter
  uint32_t *tmem_slot = fu_smem + FU_WARPS * FU_SLOTS * FU_WARP_WORDS;
  // one mbarrier per stash: the unit's TMA box completes on it
  uint32_t cur_bar = (uint32_t)__cvta_generic_to_shared(tmem_slot) + 16u + warp * (FU_SLOTS * 8u), prev_bar = cur_bar + 8u;
  uint32_t cur_par = 0u, prev_par = 0u;
  if (lane == 0u) {
    mbar_init(cur_bar, 1u); mbar_init(prev_bar, 1u);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 0u) {
    const uint32_t sa = (uint32_t)__cvta_generic_to_shared(tmem_slot);
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(sa), "n"(FU_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  uint32_t cur_pf = tmem_base + (((warp & 3u) * 32u) << 16) + (warp >> 2) * (FU_SLOTS * FU_TMEM_SLOT_COLS);
  uint32_t prev_pf = cur_pf + FU_TMEM_SLOT_COLS;

  const uint32_t n_units = p.n_bands * p.nseg;
  const uint32_t n_warps = gridDim.x * FU_WARPS;
  uint32_t lmask[5];  // lane f keeps candidate f's row sum
#pragma unroll
  for (int f = 0; f < 5; f++) {
    lmask[f] = lane == (uint32_t)f ? 0xffffffffu : 0u;
    asm volatile("" : "+r"(lmask[f]));  // keep them as registers: one LOP3 per sum instead of ISETP + SEL + move
  }

  uint32_t u = blockIdx.x * FU_WARPS + warp;  // first unit: the warp's own index; tickets continue from n_warps
  if (u < n_units) {
    UnitPlan plan = plan_make<PLAN>(p, u, plan_fetch(p, u));
    bool cur_tma = unit_load<BPW, PLAN>(p, plan, cur_stash, cur_bar, lane);
    UnitRef prev;
    prev.seg = 0; prev.ya = 0; prev.nrows = 0;
    bool have_prev = false;
    for (;;) {
      unit_sweep<BPW>(p, plan, cur_stash, cur_pf, lane, lmask, cur_tma, cur_bar, cur_par);
      if (cur_tma) cur_par ^= 1u;
      uint32_t ft = 0u;
      if (have_prev) ft = unit_wait(p, prev, lane);
      const uint32_t ftpack = pack_winners(ft, lane);
      uint32_t t = 0u;
      if (lane == 0u) t = atomicAdd(p.ticket, 1u);
      const int half = prev.nrows >> 1;
      if (have_prev) unit_emit<BPW>(p, prev, prev_stash, prev_pf, lane, ft, ftpack, 0, half);
      u = n_warps + __shfl_sync(0xffffffffu, t, 0);
      const bool more = u < n_units;
      PlanRaw raw_next;
      if (more) raw_next = plan_fetch(p, u);
      if (have_prev) unit_emit<BPW>(p, prev, prev_stash, prev_pf, lane, ft, ftpack, half, prev.nrows);
      __syncwarp();  // prev_stash is free again
      prev.seg = plan.seg; prev.ya = plan.ya; prev.nrows = plan.nrows;
      have_prev = true;
      { uint32_t *sw = cur_stash; cur_stash = prev_stash; prev_stash = sw; }
      { const uint32_t sw = cur_pf; cur_pf = prev_pf; prev_pf = sw; }
      { const uint32_t sw = cur_bar; cur_bar = prev_bar; prev_bar = sw; }
      { const uint32_t sw = cur_par; cur_par = prev_par; prev_par = sw; }
      if (!more) break;
      plan = plan_make<PLAN>(p, u, raw_next);
      cur_tma = unit_load<BPW, PLAN>(p, plan, cur_stash, cur_bar, lane);
    }
    const uint32_t ft = unit_wait(p, prev, lane);
    unit_emit<BPW>(p, prev, prev_stash, prev_pf, lane, ft, pack_winners(ft, lane), 0, prev.nrows);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0u) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(FU_TMEM_COLS) : "memory");
  }
}

uint32_t fused_rows_per_band() { return FB_R; }
uint32_t fused_words_per_seg() { return FB_SWW; }
uint32_t fused_unit_rows() { return FU_R; }
uint32_t fused_unit_words() { return FU_SW; }

// Normalise + composite rows [ys, ye) into a dense canvas in HBM, one 16-byte quad per thread.  Used ahead of
// the fused kernel when many units of a band are not plain windows of one tile (overlaps, blending, foreign
// formats): the composite is then perfectly parallel work with no cross-warp waits, and the fused kernel
// reads the canvas back as a single opaque tile on its cp.async path.
template <int BPW>
__global__ void __launch_bounds__(256) compose_canvas_kernel(const FusedParams p, uint8_t *canvas, const size_t pitch,
                                                             const uint32_t ys, const uint32_t ye) {
  const int xw = 4 * (int)(blockIdx.x * blockDim.x + threadIdx.x);
  if (xw >= (int)p.canvas_w * BPW) return;
  // a thread walks down FU_R rows of one bin: the bin's tile list is looked up once
  const uint32_t y0 = max(ys, (ys / FU_R + blockIdx.y) * FU_R), y1 = min(ye, (ys / FU_R + blockIdx.y + 1u) * FU_R);
  if (y0 >= y1) return;
  const uint32_t bin = (y0 / FB_R) * p.bins_x + (uint32_t)(xw / FB_SWW);
  const uint32_t b0 = __ldg(p.bin_start + bin), n_list = __ldg(p.bin_start + bin + 1) - b0;
  uint8_t *dst = canvas + (size_t)(y0 - ys) * pitch + (size_t)xw * 4u;
  for (uint32_t y = y0; y < y1; y++, dst += pitch)
    *reinterpret_cast<uint4 *>(dst) = compose_quad<BPW>(p, p.bin_tiles + b0, n_list, xw, (int)y);
}

cudaError_t launch_compose_canvas(const FusedParams &p, bool out16, uint8_t *canvas, size_t pitch, uint32_t ys,
                                  uint32_t ye, cudaStream_t st) {
  if (ye <= ys) return cudaSuccess;
  const uint32_t quads = (p.canvas_w * (out16 ? 2u : 1u) + 3u) / 4u;
  const uint32_t groups = (ye - 1u) / FU_R - ys / FU_R + 1u;  // FU_R-row groups on the absolute grid
  const dim3 grid((quads + 255u) / 256u, groups);
  if (groups > 65535u) return cudaErrorInvalidConfiguration;
  if (out16) compose_canvas_kernel<2><<<grid, 256, 0, st>>>(p, canvas, pitch, ys, ye);
  else compose_canvas_kernel<1><<<grid, 256, 0, st>>>(p, canvas, pitch, ys, ye);
  return cudaGetLastError();
}

size_t fused_plan_bytes(uint32_t n_units) { return (size_t)n_units * sizeof(PlanRec); }
size_t fused_unit_rows_bytes(uint32_t n_units) { return (size_t)n_units * FU_PREP_WORDS * 4; }
size_t fused_unit_left_bytes(uint32_t n_units) { return (size_t)n_units * (FU_LEFT_WORDS + FU_R + 2) * 4; }   // + the list of tasks

template <int BPW, bool PLAN>
static cudaError_t fused_occupancy(int *occ) {
  static PerDevice once;   // shared-memory opt-in and occupancy, per device
  return once.get(occ, [](int *v) {
    cudaError_t e = cudaFuncSetAttribute(fused_kernel<BPW, PLAN>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FU_CTA_SMEM);
    return e != cudaSuccess ? e : cudaOccupancyMaxActiveBlocksPerMultiprocessor(v, fused_kernel<BPW, PLAN>, FU_WARPS * 32, FU_CTA_SMEM);
  });
}

cudaError_t launch_fused(const FusedParams &p, bool out16, int sm_count, cudaStream_t st) {
  const bool plan = p.ivs != nullptr;
  int occ = 0;
  cudaError_t pe = out16 ? (plan ? fused_occupancy<2, true>(&occ) : fused_occupancy<2, false>(&occ))
                         : (plan ? fused_occupancy<1, true>(&occ) : fused_occupancy<1, false>(&occ));
  if (pe != cudaSuccess) return pe;
  if (occ < 1) return cudaErrorLaunchOutOfResources;
  if (occ > FU_MAX_CTAS) occ = FU_MAX_CTAS;
  const uint32_t n_units = p.n_bands * p.nseg;
  const uint32_t want = (n_units + FU_WARPS - 1) / FU_WARPS, resident = (uint32_t)(occ * sm_count);
  const uint32_t grid = want < resident ? want : resident;
  // wait-freedom needs 2 x warps > units per band (see fused_kernel); wider rows take the two-pass path
  if (2u * grid * FU_WARPS <= p.nseg) return cudaErrorInvalidConfiguration;
  PlanRec *recs = reinterpret_cast<PlanRec *>(p.plans);
  if (plan) {
    if (n_units >= (1u << 28)) return cudaErrorInvalidConfiguration;   // a task is unit << 4 | part
    if (out16) fused_plan_iv_kernel<2><<<p.n_bands, 256, 0, st>>>(p, recs, (FU_R + 1) / FP_ROWS);
    else fused_plan_iv_kernel<1><<<p.n_bands, 256, 0, st>>>(p, recs, (FU_R + 1) / FP_ROWS);
    // a fixed grid of warps over the listed tasks (measured: 1 or 3 rows per task and 3 to 6 CTAs per SM all take the
    // same time on C3 -- the kernel is bound by the blend arithmetic, not by occupancy)
    if (out16) fused_prep_kernel<2, FP_ROWS, 4><<<sm_count * 4, 256, 0, st>>>(p);
    else fused_prep_kernel<1, FP_ROWS, 4><<<sm_count * 4, 256, 0, st>>>(p);
  } else {
    if (out16) fused_plan_kernel<2><<<(n_units + 255u) / 256u, 256, 0, st>>>(p, recs, n_units);
    else fused_plan_kernel<1><<<(n_units + 255u) / 256u, 256, 0, st>>>(p, recs, n_units);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  // Cooperative launch: the runtime guarantees that every CTA of the grid is resident (or fails),
  // which the wait-freedom argument above fused_kernel relies on.
  FusedParams pp = p;
  void *args[] = {&pp};
  const void *fn = out16 ? (plan ? (const void *)fused_kernel<2, true> : (const void *)fused_kernel<2, false>)
                         : (plan ? (const void *)fused_kernel<1, true> : (const void *)fused_kernel<1, false>);
  return cudaLaunchCooperativeKernel(fn, dim3(grid), dim3(FU_WARPS * 32), args, FU_CTA_SMEM, st);
}

}  // namespace stitch
