// strip.cu -- kernels (b)+(c) in ONE pass over HBM, row-pipelined: normalise + composite the tiles under a
// canvas row (src/image-concat-core.ts:404-545, :551-685; src/pixel-ops.ts:496-736), compute the five PNG filter
// costs (src/png-filter.ts:148-181), agree on the row's winner across all the warps that share the row, write the
// filtered scanline.  HBM traffic: every visible tile pixel read once, every output byte written once.
//
// Decomposition.  A canvas row is cut into COLUMNS of 128 words (one 16-byte quad per lane), the rows
// [y_begin, y_end) into nry RANGES; warp (column, range) walks down its column one row at a time and is its own
// producer and consumer around a private ring of 16 row slots in shared memory (4 stages of 4 rows):
//
//   stage in   lane 0 issues, per row, what the host's plan says for the column: a 2-D TMA box load (128 words x 1
//              row) straight from the one tile that covers it, or a bulk copy of the background row; a column whose
//              source is not 16-byte aligned (a tiled TMA load faults on such a start coordinate) is copied by all
//              lanes with 4-byte cp.async; a column that needs compositing (tile edges, overlap, blending, foreign
//              formats) is composed by the warp, one quad per lane.  The pixel left of the column comes separately
//              (4-byte cp.async by lane 0).  Everything completes on the stage's mbarrier.
//   sweep      row y: the row above is carried unpacked in registers; costs of None/Sub/Up/Average on packed bytes
//              (VABSDIFF4), Paeth two bytes at a time in half2, its residual parked in TENSOR MEMORY; five warp sums
//              leave with one red.global.add.u64 each (the high half counts arrivals: no fence, no flag)
//   emit       row y-LAG: its sums are complete once every column of the range has swept it -- which they did while
//              this warp swept the LAG rows since --, winner = first minimum (None, Sub, Up, Average, Paeth);
//              Paeth comes from tensor memory, the others are recomputed from the raw rows still in the ring
//
// Warps wait for one another only through the row sums, LAG rows late; every warp of the grid must be resident
// (cooperative launch).  No wait cycle can form: a warp blocks only in the wait for the sums of row r, and every
// warp reaches the sweep of row r without waiting for anything but rows before r.
#include "common.cuh"
#include "kernels.h"
#include "filter_math.cuh"
#include "async_copy.cuh"

namespace stitch {

static constexpr int SK_WARPS = 16;                      // warps per CTA (one CTA per SM)
static constexpr int SK_THREADS = SK_WARPS * 32;
static constexpr int SK_BOX = 128;                       // words per column = per TMA box (512 bytes)
static constexpr int SK_RING = 16;                       // row slots per warp
static constexpr int SK_SROWS = 4;                       // rows per stage
static constexpr int SK_NST = SK_RING / SK_SROWS;        // stages
static constexpr int SK_TROWS = 2 * SK_SROWS;            // rows of Paeth residuals in tensor memory: the group swept and the one before
static constexpr int SK_W_ROWS = SK_RING * SK_BOX * 4;   // per-warp shared memory: rows | left pixels | stage barriers
static constexpr int SK_W_HALO = SK_W_ROWS;              //   left pixel of row slot r: 8 bytes at + 8 r (words c0, c1)
static constexpr int SK_W_BAR = SK_W_HALO + SK_RING * 8;
static constexpr int SK_W_BYTES = (SK_W_BAR + SK_NST * 8 + 127) / 128 * 128;
static constexpr int SK_SMEM = SK_WARPS * SK_W_BYTES + 16;          // + tensor-memory base address
static constexpr int SK_TMEM_COLS = 256;                 // >= (SK_WARPS / 4) x TROWS x 4 words
static_assert((SK_WARPS / 4) * SK_TROWS * 4 <= SK_TMEM_COLS && SK_WARPS % 4 == 0, "tensor memory layout");
static_assert(SK_NST == 4 && SK_SROWS == 4, "the group schedule below: sweep g, emit g-1, refill the stage of g-2 with g+2");

uint32_t strip_box_words() { return SK_BOX; }

// one out-of-line copy of the generic compositor per kernel (it carries the f64 blend); what it reads of the
// kernel's parameters travels by value, so that the parameter block is not copied to the stack
struct ComposeCtx {
  const DevTile *tiles;
  uint32_t alpha_blend, bg_lo, bg_hi, canvas_w;
};
__device__ __forceinline__ ComposeCtx compose_ctx(const StripParams &p) {
  ComposeCtx c;
  c.tiles = p.tiles; c.alpha_blend = p.alpha_blend; c.bg_lo = p.bg_lo; c.bg_hi = p.bg_hi; c.canvas_w = p.canvas_w;
  return c;
}
template <int BPW>
static __device__ __noinline__ uint4 strip_compose_quad(const ComposeCtx c, const uint32_t *list, uint32_t n_list, int xw, int y) {
  return compose_quad<BPW>(c, list, n_list, xw, y);
}

// One quad of one row: costs into s[], Paeth residual to tensor memory; hx/hlc = this row unpacked (the next row's
// "above").  hb/hlu: the row above, unpacked.  lc0/lc1: the two words left of the quad (lc0 only for 8-byte pixels).
template <int BPW, bool GUARD>
__device__ __forceinline__ void sweep_quad(const uint4 cur, const uint32_t lc0, const uint32_t lc1, const uint4 up,
                                           const H4 (&hb)[4], const H4 (&hlu)[2], H4 (&hx)[4], H4 (&hlc)[2],
                                           const int qw, const int wmax, const uint32_t taddr, uint32_t (&s)[5]) {
  const uint32_t x[4] = {cur.x, cur.y, cur.z, cur.w}, b[4] = {up.x, up.y, up.z, up.w};
  const uint32_t lc[2] = {lc0, lc1};
  if (BPW == 2) hlc[0] = unpack_h4(lc0);
  hlc[1] = unpack_h4(lc1);
  uint32_t pf[4];
#pragma unroll
  for (int j = 0; j < 4; j++) {
    hx[j] = unpack_h4(x[j]);
    const uint32_t a = j >= BPW ? x[j - BPW] : lc[2 - BPW + j];
    const H4 &ha = j >= BPW ? hx[j - BPW] : hlc[2 - BPW + j];
    const H4 &hc = j >= BPW ? hb[j - BPW] : hlu[2 - BPW + j];
    pf[j] = pack_h4(paeth_resid_h2(hx[j].lo, ha.lo, hb[j].lo, hc.lo), paeth_resid_h2(hx[j].hi, ha.hi, hb[j].hi, hc.hi));
    if (!GUARD || qw + j < wmax) {  // words past the canvas edge carry no cost
      s[0] = acc_abs_s4(x[j], s[0]);
      s[1] = acc_abs_s4(absdiff_u4(x[j], a), s[1]);
      s[2] = acc_abs_s4(absdiff_u4(x[j], b[j]), s[2]);
      s[3] = acc_abs_s4(absdiff_u4(x[j], avg4(a, b[j])), s[3]);
      s[4] = acc_abs_s4(pf[j], s[4]);
    }
  }
  tmem_st4(taddr, make_uint4(pf[0], pf[1], pf[2], pf[3]));
}

// ---- the host's plan, walked downwards ---------------------------------------------------------------------
struct IvCursor {
  uint32_t idx, end_idx;   // current interval, one past the column's last
  uint32_t y_end, kind, tile, list_off, n_list, flags;
  int32_t cw, cy;
  const uint8_t *pix;      // SK_DIRECT, SK_SHIFT: the tile's rows
  uint32_t pitch;
};
__device__ __forceinline__ void iv_load(const StripParams &p, IvCursor &c) {
  const uint4 *q = reinterpret_cast<const uint4 *>(p.ivs + c.idx);
  const uint4 a = __ldg(q), b = __ldg(q + 1);
  c.y_end = a.x; c.kind = a.y; c.tile = a.z; c.cw = (int32_t)a.w;
  c.cy = (int32_t)b.x; c.list_off = b.y; c.n_list = b.z; c.flags = b.w;
  if (c.kind == SK_DIRECT || c.kind == SK_SHIFT) {
    const DevTile *t = p.tiles + c.tile;
    c.pix = t->pix; c.pitch = t->pitch;
  }
}
__device__ __forceinline__ void iv_open(const StripParams &p, IvCursor &c, uint32_t col) {
  c.idx = __ldg(p.col_start + col); c.end_idx = __ldg(p.col_start + col + 1);
  iv_load(p, c);
}
__device__ __forceinline__ void iv_seek(const StripParams &p, IvCursor &c, int y) {
  while ((uint32_t)y >= c.y_end && c.idx + 1u < c.end_idx) { c.idx++; iv_load(p, c); }
}

__device__ __forceinline__ void cp_async8(uint32_t dst, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}

// The cursors live in shared memory between stage issues (12 words each), so that they hold no registers
// while the warp sweeps.
__device__ __forceinline__ void cur_save(uint32_t a, const IvCursor &c) {
  sts128(a, make_uint4(c.idx, c.end_idx, c.y_end, c.kind));
  sts128(a + 16u, make_uint4(c.tile, c.list_off, c.n_list, (uint32_t)c.cw));
  const unsigned long long px = reinterpret_cast<unsigned long long>(c.pix);
  sts128(a + 32u, make_uint4((uint32_t)c.cy, c.pitch, (uint32_t)px, (uint32_t)(px >> 32)));
}
__device__ __forceinline__ void cur_load(uint32_t a, IvCursor &c) {
  const uint4 x = lds128(a), y = lds128(a + 16u), z = lds128(a + 32u);
  c.idx = x.x; c.end_idx = x.y; c.y_end = x.z; c.kind = x.w;
  c.tile = y.x; c.list_off = y.y; c.n_list = y.z; c.cw = (int32_t)y.w;
  c.cy = (int32_t)z.x; c.pitch = z.y;
  c.pix = reinterpret_cast<const uint8_t *>(((unsigned long long)z.w << 32) | z.z);
}

// Shared-memory addresses of one warp's private area.
struct WarpSm {
  uint32_t rows, halo, bar, cur;   // cur: two saved cursors (the column's, the left neighbour's)
};

// The pixel left of the column in row y -> the slot's halo words (c0 at +0, c1 at +4).  One lane.
template <int BPW>
__device__ __forceinline__ void halo_issue(const StripParams &p, IvCursor &hc, const bool has_left, const uint32_t col,
                                           const int y, const uint32_t halo) {
  if (!has_left || y < 0) { sts32(halo, 0u); sts32(halo + 4u, 0u); return; }
  iv_seek(p, hc, y);
  if (hc.kind == SK_DIRECT || hc.kind == SK_SHIFT) {
    const uint8_t *src = hc.pix + (size_t)(y + hc.cy) * hc.pitch + (size_t)(hc.cw + SK_BOX - BPW) * 4u;
    if (BPW == 2) cp_async8(halo, src); else cp_async4(halo + 4u, src);
  } else if (hc.kind == SK_COMP) {
    const uint4 v = strip_compose_quad<BPW>(compose_ctx(p), p.lists + hc.list_off, hc.n_list, (int)(col * SK_BOX - 4u), y);
    sts32(halo, v.z); sts32(halo + 4u, v.w);
  } else {
    sts32(halo, BPW == 2 ? p.bg_lo : 0u); sts32(halo + 4u, BPW == 2 ? p.bg_hi : p.bg_lo);
  }
}

// Ring entries 4j .. 4j+3 (rows ya-1+k) of the warp's column into their slots; completes on the stage's mbarrier.
// Lanes 0..3 fetch the left pixels of the four rows, each with its own copy of the left neighbour's cursor.
template <int BPW>
__device__ __forceinline__ void issue_stage(const StripParams &p, const WarpSm ws, const uint32_t lane, const uint32_t col,
                                            const int ya, const int n_k, const int j) {
  const int k0 = j * SK_SROWS;
  if (k0 >= n_k) return;
  const int y0 = ya - 1 + k0;
  const int n_r = min(SK_SROWS, n_k - k0);   // rows of the stage that exist
  IvCursor bc;
  cur_load(ws.cur, bc);
  const uint32_t bar = ws.bar + 8u * ((uint32_t)j & (SK_NST - 1));
  const uint32_t slot0 = (uint32_t)k0 & (SK_RING - 1);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // our reads of these slots, then bulk writes into them
  if ((int)lane < n_r) {
    IvCursor hc;
    const bool has_left = col > 0u;
    if (has_left) cur_load(ws.cur + 48u, hc);
    halo_issue<BPW>(p, hc, has_left, col, y0 + (int)lane, ws.halo + 8u * (slot0 + lane));
    if ((int)lane == n_r - 1 && has_left) cur_save(ws.cur + 48u, hc);   // the furthest one
  }
  if (y0 >= 0) iv_seek(p, bc, y0);
  if (y0 >= 0 && bc.kind == SK_DIRECT && (uint32_t)(y0 + n_r - 1) < bc.y_end) {
    // the whole stage lies in one tile: one box of 128 words x 4 rows (rows past the range or the tile are never read)
    if (lane == 0u) {
      mbar_expect_tx(bar, SK_SROWS * SK_BOX * 4);
      tma_box_2d(ws.rows + slot0 * (SK_BOX * 4), reinterpret_cast<const uint8_t *>(p.maps) + (size_t)bc.tile * 384u + 128u, bc.cw, y0 + bc.cy, bar);
    }
  } else {
#pragma unroll 1
    for (int r = 0; r < n_r; r++) {
      const int y = y0 + r;
      const uint32_t row = ws.rows + (slot0 + (uint32_t)r) * (SK_BOX * 4);
      if (y < 0) {   // previousLine === null: zeros
        sts128(row + 16u * lane, make_uint4(0u, 0u, 0u, 0u));
        continue;
      }
      iv_seek(p, bc, y);
      if (bc.kind == SK_DIRECT) {
        if (lane == 0u) {
          mbar_expect_tx(bar, SK_BOX * 4);
          tma_box_2d(row, reinterpret_cast<const uint8_t *>(p.maps) + (size_t)bc.tile * 384u, bc.cw, y + bc.cy, bar);
        }
      } else if (bc.kind == SK_BG) {
        if (lane == 0u) {
          mbar_expect_tx(bar, SK_BOX * 4);
          bulk_copy(row, p.bg_row, SK_BOX * 4, bar);
        }
      } else if (bc.kind == SK_SHIFT) {
        const uint8_t *src = bc.pix + (size_t)(y + bc.cy) * bc.pitch + (size_t)bc.cw * 4u + 16u * lane;
#pragma unroll
        for (int i = 0; i < 4; i++)
          if (4u * lane + i < bc.list_off) cp_async4(row + 16u * lane + 4u * i, src + 4 * i);
      } else {
        const uint4 v = strip_compose_quad<BPW>(compose_ctx(p), p.lists + bc.list_off, bc.n_list, (int)(col * SK_BOX + 4u * lane), y);
        sts128(row + 16u * lane, v);
      }
    }
  }
  __syncwarp();
  mbar_arrive_after_cp_async(bar);   // every lane once per stage: when its own copies, if any, have landed
  cur_save(ws.cur, bc);              // (all lanes hold the same cursor: same value, same address)
}

// Winners of the SK_SROWS ring entries k0 .. k0+3 (first minimum in the order None, Sub, Up, Average, Paeth;
// png-filter.ts:174) once all `total` columns have added their sums: lane 8 r + f holds word f of row r.  Returns
// the winners packed 3 bits per row, the same in every lane.  pre: the lane's word as read earlier.
__device__ __forceinline__ uint32_t strip_winners(const StripParams &p, const int y0, const int k0, const int n_k,
                                                  const uint32_t lane, const uint32_t total, unsigned long long w) {
  const uint32_t r = lane >> 3, f = lane & 7u;
  const bool mine = f < 5u && k0 + (int)r >= 1 && k0 + (int)r < n_k;   // entry 0 is the predictor row: never emitted
  const unsigned long long *rs = p.rowsum + (size_t)(y0 + (int)r - (int)p.y_begin) * 8 + f;
  for (;;) {
    const bool ok = !mine || (uint32_t)(w >> 32) == total;
    if (__all_sync(0xffffffffu, ok)) break;
#ifdef SK_EXPERIMENT_NOWAIT
    break;
#endif
    __nanosleep(100);
    if (mine) asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(rs) : "memory");
  }
  uint32_t key = mine ? (((uint32_t)w << 3) | f) : 0xffffffffu;
  key = min(key, __shfl_xor_sync(0xffffffffu, key, 1));
  key = min(key, __shfl_xor_sync(0xffffffffu, key, 2));
  key = min(key, __shfl_xor_sync(0xffffffffu, key, 4));
  return __reduce_or_sync(0xffffffffu, f == 0u ? (key & 7u) << (3u * r) : 0u);
}

// The row above, as the sweep of the next row needs it.
struct RowCarry {
  uint4 raw;     // the lane's quad, packed
  H4 hb[4];      // ... unpacked to half2 (1024 + byte)
  H4 hl[2];      // the two words left of the quad, unpacked ([0] only for 8-byte pixels)
};

// The two words left of the lane's quad: the lane to the left has them; lane 0 reads the slot's halo words.
template <int BPW>
__device__ __forceinline__ void left_words(const uint4 q, const uint32_t halo, const uint32_t lane, uint32_t &c0, uint32_t &c1) {
  c0 = BPW == 2 ? __shfl_up_sync(0xffffffffu, q.z, 1) : 0u;
  c1 = __shfl_up_sync(0xffffffffu, q.w, 1);
  if (lane == 0u) {
    if (BPW == 2) c0 = lds32(halo);
    c1 = lds32(halo + 4u);
  }
}

// Sweep of one row: `ab` = the row above (in), `cu` = this row (out).  Adds the warp's sums to the row's words.
template <int BPW>
__device__ __forceinline__ void strip_sweep(const StripParams &p, const RowCarry &ab, RowCarry &cu, const uint32_t row,
                                            const uint32_t halo, const uint32_t lane, const int qw, const int wmax,
                                            const bool guard, const uint32_t taddr, const int y, const uint32_t (&lmask)[5]) {
  uint32_t s[5] = {0u, 0u, 0u, 0u, 0u};
  const uint4 cur = lds128(row);
  uint32_t lc0, lc1;
  left_words<BPW>(cur, halo, lane, lc0, lc1);
  if (guard) sweep_quad<BPW, true>(cur, lc0, lc1, ab.raw, ab.hb, ab.hl, cu.hb, cu.hl, qw, wmax, taddr, s);
  else sweep_quad<BPW, false>(cur, lc0, lc1, ab.raw, ab.hb, ab.hl, cu.hb, cu.hl, qw, wmax, taddr, s);
  cu.raw = cur;
  uint32_t v = 0u;
#pragma unroll
  for (int f = 0; f < 5; f++) v |= __reduce_add_sync(0xffffffffu, s[f]) & lmask[f];
  if (lane < 5u) atomicAdd(p.rowsum + (size_t)(y - (int)p.y_begin) * 8 + lane, (1ull << 32) | (unsigned long long)v);
}

// Row y with winner ft: the filtered bytes, one 128-bit store per lane.  cur/up: the row and the row above in the ring.
template <int BPW>
__device__ __forceinline__ void strip_emit(const StripParams &p, const uint32_t ft, const int y, const uint32_t cur_row,
                                           const uint32_t up_row, const uint32_t halo, const uint32_t taddr,
                                           const uint32_t lane, const uint32_t col, const int qw, const int wmax) {
  uint8_t *orow = p.out + (size_t)(y - (int)p.y_begin) * p.out_pitch + p.lead;
  if (col == 0u && lane == 0u) orow[0] = (uint8_t)ft;
  uint4 o;
  if (ft == 4u) {
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");   // the row's residuals were stored LAG rows ago
    o = tmem_ld4(taddr);
  } else {
    o = lds128(cur_row);
    if (ft != 0u) {
      uint4 pr;
      if (ft == 2u) {
        pr = lds128(up_row);
      } else {
        uint32_t c0, c1;
        left_words<BPW>(o, halo, lane, c0, c1);
        pr = BPW == 1 ? make_uint4(c1, o.x, o.y, o.z) : make_uint4(c0, c1, o.x, o.y);
        if (ft == 3u) {
          const uint4 upq = lds128(up_row);
          pr = make_uint4(avg4(pr.x, upq.x), avg4(pr.y, upq.y), avg4(pr.z, upq.z), avg4(pr.w, upq.w));
        }
      }
      o = make_uint4(sub4(o.x, pr.x), sub4(o.y, pr.y), sub4(o.z, pr.z), sub4(o.w, pr.w));
    }
  }
  uint8_t *od = orow + 1 + 4 * (size_t)qw;
  if (qw + 4 <= wmax) {
    *reinterpret_cast<uint4 *>(od) = o;
  } else if (qw < wmax) {
    uint32_t *ow = reinterpret_cast<uint32_t *>(od);
    ow[0] = o.x;
    if (qw + 1 < wmax) ow[1] = o.y;
    if (qw + 2 < wmax) ow[2] = o.z;
  }
  __syncwarp();   // the partial stores may have split the warp; what follows is .sync.aligned
}

template <int BPW>
__global__ void __launch_bounds__(SK_THREADS, 1) strip_kernel(const StripParams p) {
  extern __shared__ __align__(128) uint8_t sk_smem[];
  const uint32_t smem = (uint32_t)__cvta_generic_to_shared(sk_smem);
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  WarpSm ws;
  ws.rows = smem + warp * SK_W_BYTES; ws.halo = ws.rows + SK_W_HALO; ws.bar = ws.rows + SK_W_BAR;
  ws.cur = smem + SK_WARPS * SK_W_BYTES + 16u + warp * 96u;
  const uint32_t tmem_slot = smem + SK_WARPS * SK_W_BYTES;
  if (lane == 0u) {
    for (int s = 0; s < SK_NST; s++) mbar_init(ws.bar + 8u * s, 32u);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 0u) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(SK_TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = lds32(tmem_slot);

  // consecutive warps take consecutive columns of one range
  const uint32_t wid = blockIdx.x * SK_WARPS + warp;
  const uint32_t ry = wid / p.ncol, col = wid - ry * p.ncol;
  if (ry < p.nry) {
    const int ya = (int)(p.y_begin + ry * p.rows_per_range);
    const int yb = (int)min(p.y_end, (uint32_t)ya + p.rows_per_range);
    const int n_k = yb - ya + 1;   // ring entries: rows ya-1 .. yb-1
    const int qw = (int)(col * SK_BOX + 4u * lane), wmax = (int)p.canvas_w * BPW;
    const bool guard = (int)(col * SK_BOX) + SK_BOX > wmax;
    const uint32_t total = p.ncol;
    // tensor memory: lane quarter warp % 4, column block warp / 4, one 4-word row per ring entry mod TROWS
    const uint32_t tm = tmem_base + (((warp & 3u) * 32u) << 16) + (warp >> 2) * (SK_TROWS * 4u);
    uint32_t lmask[5];
#pragma unroll
    for (int f = 0; f < 5; f++) {
      lmask[f] = lane == (uint32_t)f ? 0xffffffffu : 0u;
      asm volatile("" : "+r"(lmask[f]));
    }
    {  // the cursors start at the columns' first intervals
      IvCursor c;
      iv_open(p, c, col);
      cur_save(ws.cur, c);
      if (lane == 0u && col > 0u) { iv_open(p, c, col - 1u); cur_save(ws.cur + 48u, c); }
      __syncwarp();
    }
    auto row_addr = [&](int k) { return ws.rows + ((uint32_t)k & (SK_RING - 1)) * (SK_BOX * 4) + 16u * lane; };
    auto halo_addr = [&](int k) { return ws.halo + ((uint32_t)k & (SK_RING - 1)) * 8u; };
    auto tm_row = [&](int k) { return tm + ((uint32_t)k & (SK_TROWS - 1)) * 4u; };
    auto wait_stage = [&](int j) { mbar_wait(ws.bar + 8u * ((uint32_t)j & (SK_NST - 1)), ((uint32_t)j / SK_NST) & 1u); };
    // lane 8 r + f reads word f of the sums of ring entry k0 + r
    auto poll = [&](int k0) {
      unsigned long long w = 0ull;
      const int k = k0 + (int)(lane >> 3);
      if ((lane & 7u) < 5u && k >= 1 && k < n_k)
        asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p.rowsum + (size_t)(ya - 1 + k - (int)p.y_begin) * 8 + (lane & 7u)) : "memory");
      return w;
    };
    auto sweep = [&](int k, const RowCarry &ab, RowCarry &cu) {
      strip_sweep<BPW>(p, ab, cu, row_addr(k), halo_addr(k), lane, qw, wmax, guard, tm_row(k), ya - 1 + k, lmask);
    };
    auto emit_group = [&](int g, unsigned long long pre) {   // ring entries 4g .. 4g+3
      const int k0 = g * SK_SROWS;
      const uint32_t ftpack = strip_winners(p, ya - 1 + k0, k0, n_k, lane, total, pre);
#pragma unroll
      for (int r = 0; r < SK_SROWS; r++) {
        const int k = k0 + r;
        if (k >= 1 && k < n_k)
          strip_emit<BPW>(p, (ftpack >> (3 * r)) & 7u, ya - 1 + k, row_addr(k), row_addr(k - 1), halo_addr(k), tm_row(k), lane, col, qw, wmax);
      }
    };
    for (int j = 0; j < SK_NST - 1; j++) issue_stage<BPW>(p, ws, lane, col, ya, n_k, j);   // stages 0, 1, 2
    RowCarry ca, cb;
    wait_stage(0);
    {
      ca.raw = lds128(row_addr(0));
      ca.hb[0] = unpack_h4(ca.raw.x); ca.hb[1] = unpack_h4(ca.raw.y); ca.hb[2] = unpack_h4(ca.raw.z); ca.hb[3] = unpack_h4(ca.raw.w);
      uint32_t c0, c1;
      left_words<BPW>(ca.raw, halo_addr(0), lane, c0, c1);
      ca.hl[0] = unpack_h4(c0); ca.hl[1] = unpack_h4(c1);
    }
    // Group g = ring entries 4g .. 4g+3.  Per iteration: sweep group g, emit group g-1 (its sums have had a group's
    // time to complete), then refill the stage of group g-2 -- nothing reads it any more -- with group g+2.
    const int n_g = (n_k + SK_SROWS - 1) / SK_SROWS;
    for (int g = 0; g < n_g; g++) {
      const int k0 = g * SK_SROWS;
      if (g > 0) wait_stage(g);
      unsigned long long pre = 0ull;
      if (k0 + 3 < n_k) {   // a whole group: straight-line
        if (g > 0) sweep(k0, ca, cb); else cb = ca;
        sweep(k0 + 1, cb, ca);
        sweep(k0 + 2, ca, cb);
        if (g > 0) pre = poll(k0 - SK_SROWS);   // read late (the others have had three rows' time), used after one more row
        sweep(k0 + 3, cb, ca);
      } else {
        for (int k = k0 > 0 ? k0 : 1; k < n_k; k++) { sweep(k, ca, cb); ca = cb; }
        if (g > 0) pre = poll(k0 - SK_SROWS);
      }
      if (g > 0) emit_group(g - 1, pre);
      if (g > 0) issue_stage<BPW>(p, ws, lane, col, ya, n_k, g + 2);
    }
    emit_group(n_g - 1, poll((n_g - 1) * SK_SROWS));
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0u) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(SK_TMEM_COLS) : "memory");
  }
}

// ---- plan-driven compositor: rows [ys, ye) of the canvas into HBM ----------------------------------------------
// Used when a band goes "via canvas" (many composited boxes) and for the JPEG feed.  A warp owns one 128-word column
// over CP_ROWS rows, so the plan's interval makes the kind of work warp-uniform: a background column is a stream of
// constant stores, a column inside one tile a copy (16-byte loads when the source is aligned, else word loads), a
// column whose layers all span it in the output's format blends layer after layer without per-pixel tests, and only
// columns with tile edges or foreign formats walk their -- short, exact -- tile list per quad.
static constexpr int CP_ROWS = 16;

template <int BPW>
__global__ void __launch_bounds__(128) compose_plan_kernel(const StripParams p, uint8_t *canvas, const size_t pitch,
                                                           const uint32_t ys, const uint32_t ye) {
  const uint32_t lane = threadIdx.x, col = blockIdx.x;
  const uint32_t y0 = ys + (blockIdx.y * blockDim.y + threadIdx.y) * CP_ROWS;
  if (y0 >= ye) return;
  const uint32_t y1 = min(ye, y0 + CP_ROWS);
  const int wmax = (int)p.canvas_w * BPW, qw = (int)(col * SK_BOX + 4u * lane);
  if (qw >= wmax) return;
  // the column's interval that holds row y0: binary search over y_end
  IvCursor c;
  c.idx = __ldg(p.col_start + col); c.end_idx = __ldg(p.col_start + col + 1);
  {
    uint32_t lo = c.idx, hi = c.end_idx - 1u;
    while (lo < hi) {
      const uint32_t mid = (lo + hi) >> 1;
      if (__ldg(&p.ivs[mid].y_end) > y0) hi = mid; else lo = mid + 1u;
    }
    c.idx = lo;
  }
  iv_load(p, c);
  const ComposeCtx cc = compose_ctx(p);
  uint8_t *dst = canvas + (size_t)(y0 - ys) * pitch + (size_t)qw * 4u;
  const bool partial = qw + 4 > wmax;   // the canvas ends inside this quad: words past it are written as zeros
  for (uint32_t y = y0; y < y1; y++, dst += pitch) {
    iv_seek(p, c, (int)y);
    uint4 v;
    if (c.kind == SK_BG) {
      v = BPW == 2 ? make_uint4(p.bg_lo, p.bg_hi, p.bg_lo, p.bg_hi) : make_uint4(p.bg_lo, p.bg_lo, p.bg_lo, p.bg_lo);
    } else if (c.kind == SK_DIRECT) {
      v = __ldg(reinterpret_cast<const uint4 *>(c.pix + (size_t)((int)y + c.cy) * c.pitch + (size_t)c.cw * 4u) + lane);
    } else if (c.kind == SK_SHIFT) {
      const uint32_t *src = reinterpret_cast<const uint32_t *>(c.pix + (size_t)((int)y + c.cy) * c.pitch) + c.cw + 4 * (int)lane;
      const uint32_t n = c.list_off > 4u * lane ? c.list_off - 4u * lane : 0u;   // words of the quad that exist in the source row
      v.x = n > 0u ? __ldg(src) : 0u; v.y = n > 1u ? __ldg(src + 1) : 0u; v.z = n > 2u ? __ldg(src + 2) : 0u; v.w = n > 3u ? __ldg(src + 3) : 0u;
    } else if (c.flags & 1u) {
      // every layer spans the column in the output's format: background, then layer after layer, no per-pixel tests
      v = BPW == 2 ? make_uint4(p.bg_lo, p.bg_hi, p.bg_lo, p.bg_hi) : make_uint4(p.bg_lo, p.bg_lo, p.bg_lo, p.bg_lo);
      for (uint32_t k = 0; k < c.n_list; k++) {
        const DevTile *t = p.tiles + __ldg(p.lists + c.list_off + k);
        const uint32_t *src = reinterpret_cast<const uint32_t *>(t->pix + (size_t)((int)y - t->y0 + t->sy0) * t->pitch) +
                              ((int)(col * (SK_BOX / BPW)) - t->x0 + t->sx0) * BPW + 4 * (int)lane;
        uint4 s;
        if (partial) {   // the canvas (and possibly the tile's row) ends inside this quad: no loads past it
          s.x = __ldg(src); s.y = qw + 1 < wmax ? __ldg(src + 1) : 0u; s.z = qw + 2 < wmax ? __ldg(src + 2) : 0u; s.w = 0u;
        } else if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
          s = __ldg(reinterpret_cast<const uint4 *>(src));
        } else {
          s.x = __ldg(src); s.y = __ldg(src + 1); s.z = __ldg(src + 2); s.w = __ldg(src + 3);
        }
        if (p.alpha_blend == 0u || t->opaque) v = s;
        else if (BPW == 1) v = make_uint4(blend8(v.x, s.x), blend8(v.y, s.y), blend8(v.z, s.z), blend8(v.w, s.w));
        else {
          const uint2 a = blend16(make_uint2(v.x, v.y), make_uint2(s.x, s.y)), b = blend16(make_uint2(v.z, v.w), make_uint2(s.z, s.w));
          v = make_uint4(a.x, a.y, b.x, b.y);
        }
      }
    } else {
      v = compose_quad<BPW>(cc, p.lists + c.list_off, c.n_list, qw, (int)y);
    }
    if (partial) {
      if (qw + 1 >= wmax) v.y = 0u;
      if (qw + 2 >= wmax) v.z = 0u;
      if (qw + 3 >= wmax) v.w = 0u;
    }
    *reinterpret_cast<uint4 *>(dst) = v;
  }
}

cudaError_t launch_compose_plan(StripParams p, bool out16, uint8_t *canvas, size_t pitch, uint32_t ys, uint32_t ye, cudaStream_t st) {
  if (ye <= ys) return cudaSuccess;
  const uint32_t row_words = p.canvas_w * (out16 ? 2u : 1u);
  const uint32_t ncol = (row_words + SK_BOX - 1) / SK_BOX;
  const uint32_t chunks = (ye - ys + CP_ROWS - 1) / CP_ROWS;
  const dim3 block(32, 4), grid(ncol, (chunks + 3u) / 4u);
  if (grid.y > 65535u) return cudaErrorInvalidConfiguration;
  if (out16) compose_plan_kernel<2><<<grid, block, 0, st>>>(p, canvas, pitch, ys, ye);
  else compose_plan_kernel<1><<<grid, block, 0, st>>>(p, canvas, pitch, ys, ye);
  return cudaGetLastError();
}

template <int BPW>
static cudaError_t launch_strip_t(StripParams &p, int sm_count, cudaStream_t st) {
  const uint32_t row_words = p.canvas_w * BPW, n_rows = p.y_end - p.y_begin;
  const int smem_bytes = SK_SMEM + SK_WARPS * 96;
  p.ncol = (row_words + SK_BOX - 1) / SK_BOX;
  p.nsx = p.ncol;
  cudaError_t e = cudaFuncSetAttribute(strip_kernel<BPW>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);   // per device: at every launch
  if (e != cudaSuccess) return e;
  int occ = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, strip_kernel<BPW>, SK_THREADS, smem_bytes);
  if (e != cudaSuccess) return e;
  if (occ < 1) return cudaErrorLaunchOutOfResources;
  const uint32_t resident = (uint32_t)sm_count * SK_WARPS;   // one CTA per SM (it owns the SM's tensor-memory columns)
  if (p.ncol > resident) return cudaErrorInvalidConfiguration;
  uint32_t nry = resident / p.ncol;
  if (nry > n_rows) nry = n_rows;
  p.rows_per_range = (n_rows + nry - 1) / nry;
  p.nry = (n_rows + p.rows_per_range - 1) / p.rows_per_range;   // no empty range
  const uint32_t ctas = (p.ncol * p.nry + SK_WARPS - 1) / SK_WARPS;
  void *args[] = {&p};
  return cudaLaunchCooperativeKernel((const void *)strip_kernel<BPW>, dim3(ctas), dim3(SK_THREADS), args, (size_t)smem_bytes, st);
}

cudaError_t launch_strip(StripParams p, bool out16, int q, int device, int sm_count, cudaStream_t st) {
  (void)device; (void)q;
  const uint32_t row_words = p.canvas_w * (out16 ? 2u : 1u);
  if (p.y_end <= p.y_begin || row_words == 0u) return cudaSuccess;
  if ((size_t)row_words * 4 >= ((size_t)1 << 22)) return cudaErrorInvalidConfiguration;   // cost sums must stay below 2^29
  return out16 ? launch_strip_t<2>(p, sm_count, st) : launch_strip_t<1>(p, sm_count, st);
}

}  // namespace stitch
