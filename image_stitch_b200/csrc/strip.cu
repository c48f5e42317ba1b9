// strip.cu -- kernels (b)+(c) in ONE pass over HBM, row-pipelined: normalise + composite the tiles under a
// canvas row (src/image-concat-core.ts:404-545, :551-685; src/pixel-ops.ts:496-736), compute the five PNG filter
// costs (src/png-filter.ts:148-181), agree on the row's winner across every CTA that shares the row, write the
// filtered scanline.  HBM traffic: every visible tile pixel read once, every output byte written once.
//
// Decomposition.  Rows [y_begin, y_end) are cut into nry row RANGES, a canvas row into nsx column STRIPS of
// SW = Q * 2048 words; CTA (sx, ry) walks down its strip of its range one row at a time.  A CTA is 16 consumer
// warps and one producer warp around a ring of NSLOT row slots in shared memory:
//
//   producer   per row: waits until the slot is free; lane b owns box b (1 KiB) of the strip row and issues what
//              the host's plan says for it -- a 2-D TMA box load straight from the one tile that covers it (any
//              pixel alignment: the tensor map does the address arithmetic), a bulk copy of the background row,
//              or nothing: boxes that need compositing are flagged for the consumers.  Lane 0 also fetches the
//              pixel left of the strip.  Loads complete on the slot's "full" mbarrier.
//   consumers  per row y: (1) composite the flagged boxes of row y+1 (normalise + alpha-over per 16-byte quad),
//              arrive on its "full" barrier; (2) wait for row y and sweep it: thread t owns quads t, t+512, ... of
//              the strip row, the row above is carried unpacked in registers, costs of None/Sub/Up/Average on
//              packed bytes (VABSDIFF4), Paeth two bytes at a time in half2, its residual parked in TENSOR MEMORY;
//              (3) publish the warp's five partial sums with one red.global.add.u64 each (the high half counts
//              arrivals, so the sums need no fence and no flag); (4) read row y-1's sums -- complete once every
//              warp of every strip has swept y-1, which they did while this warp swept y --, pick the winner
//              (first minimum: None, Sub, Up, Average, Paeth) and emit row y-1: Paeth from tensor memory, the
//              others recomputed from the raw rows still in the ring; (5) release row y-2's slot.
//
// Every CTA of the grid must be resident (the CTAs of a range wait for one another's sums): cooperative launch.
// No wait cycle can form: the sums of row r need every strip to have swept r, which needs row r's slot, which is
// free once row r-NSLOT+... was emitted, which needs the sums of an EARLIER row: induction on the row index.
#include "common.cuh"
#include "kernels.h"
#include "filter_math.cuh"

namespace stitch {

static constexpr int SK_WARPS = 16;                      // consumer warps
static constexpr int SK_CONS = SK_WARPS * 32;            // consumer threads
static constexpr int SK_THREADS = SK_CONS + 32;          // + the producer warp
static constexpr int SK_BOX = 256;                       // words per TMA box (1 KiB)
static constexpr int SK_NSLOT = 5;                       // rows y-2 .. y+2
static constexpr int SK_LEAD = 128;                      // bytes ahead of a slot's row; the pixel left of the strip ends at +128

template <int Q> struct SKGeo {
  static constexpr int SW = Q * SK_CONS * 4;             // words per strip row
  static constexpr int NB = SW / SK_BOX;                 // boxes per strip row (8 Q)
  static constexpr int SLOT_BYTES = SK_LEAD + SW * 4;
  static constexpr int OFF_BAR = SK_NSLOT * SLOT_BYTES;  // full[NSLOT] | empty[NSLOT] | planned[NSLOT]
  static constexpr int OFF_PLAN = OFF_BAR + 3 * SK_NSLOT * 8 + 8;   // per slot: 32 interval indices + mask (+ pad)
  static constexpr int PLAN_WORDS = 36;
  static constexpr int OFF_TMEM = OFF_PLAN + SK_NSLOT * PLAN_WORDS * 4;
  static constexpr int SMEM = OFF_TMEM + 16;
  static constexpr int TMEM_COLS = Q * 32 < 32 ? 32 : Q * 32;       // 4 warp groups x 2 rows x Q quads x 4 words
  static_assert(NB <= 32, "one producer lane per box");
  static_assert((TMEM_COLS & (TMEM_COLS - 1)) == 0 && TMEM_COLS <= 512, "tensor memory columns: a power of two");
  static_assert(SLOT_BYTES % 128 == 0, "TMA destinations are 128-byte aligned");
};

uint32_t strip_box_words() { return SK_BOX; }

// ---- mbarrier / TMA / shared-memory primitives (shared::cta addresses as 32-bit) ------------------------------
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {   // no arrival: only more bytes to wait for
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// arrives once all cp.async copies this thread has issued so far have landed (at once if there are none)
__device__ __forceinline__ void mbar_arrive_after_cp_async(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0u;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try(bar, parity)) {}
}
__device__ __forceinline__ void tma_box_2d(uint32_t dst, const void *map, int c0, int c1, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
__device__ __forceinline__ void bulk_copy(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ uint4 lds128(uint32_t a) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t lds32(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
  return v;
}
__device__ __forceinline__ void sts128(uint32_t a, const uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}

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
  uint32_t y_end, kind, tile, list_off, n_list;
  int32_t cw, cy;
  const uint8_t *pix;      // SK_DIRECT, SK_SHIFT: the tile's rows
  uint32_t pitch;
};
__device__ __forceinline__ void iv_load(const StripParams &p, IvCursor &c) {
  const uint4 *q = reinterpret_cast<const uint4 *>(p.ivs + c.idx);
  const uint4 a = __ldg(q), b = __ldg(q + 1);
  c.y_end = a.x; c.kind = a.y; c.tile = a.z; c.cw = (int32_t)a.w;
  c.cy = (int32_t)b.x; c.list_off = b.y; c.n_list = b.z;
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

// The pixel left of the strip in row y (the last pixel of box column `ch`): c0 = word -2 (8-byte pixels), c1 = word -1.
template <int BPW>
__device__ __forceinline__ uint2 halo_pixel(const StripParams &p, IvCursor &hc, const uint32_t ch, const int y) {
  if (y < 0) return make_uint2(0u, 0u);
  iv_seek(p, hc, y);
  if (hc.kind == SK_DIRECT || hc.kind == SK_SHIFT) {
    const uint32_t *src = reinterpret_cast<const uint32_t *>(hc.pix + (size_t)(y + hc.cy) * hc.pitch) + (hc.cw + SK_BOX - BPW);
    return BPW == 2 ? make_uint2(__ldg(src), __ldg(src + 1)) : make_uint2(0u, __ldg(src));
  }
  if (hc.kind == SK_COMP) {
    const uint4 v = strip_compose_quad<BPW>(compose_ctx(p), p.lists + hc.list_off, hc.n_list, (int)(ch * SK_BOX + SK_BOX - 4), y);
    return make_uint2(v.z, v.w);
  }
  return BPW == 2 ? make_uint2(p.bg_lo, p.bg_hi) : make_uint2(0u, p.bg_lo);
}

// ---- producer warp: rows ya-1 .. yb-1 of the strip into the ring ------------------------------------------------
template <int BPW, int Q>
__device__ __forceinline__ void strip_producer(const StripParams &p, const uint32_t smem, const uint32_t lane,
                                               const uint32_t sx, const int ya, const int n_k) {
  using G = SKGeo<Q>;
  const uint32_t bar_full = smem + G::OFF_BAR, bar_empty = bar_full + 8 * SK_NSLOT, bar_plan = bar_empty + 8 * SK_NSLOT;
  const uint32_t col = sx * G::NB + lane;
  const bool has_col = lane < (uint32_t)G::NB && col < p.ncol;
  IvCursor bc{}, hc{};
  if (has_col) iv_open(p, bc, col);
  const bool has_halo = lane == 0u && sx > 0u;
  const uint32_t ch = sx * G::NB - 1u;
  if (has_halo) iv_open(p, hc, ch);
  uint2 halo = make_uint2(0u, 0u);
  if (has_halo) halo = halo_pixel<BPW>(p, hc, ch, ya - 1);
  for (int k = 0; k < n_k; k++) {
    const int y = ya - 1 + k;
    const uint32_t slot = (uint32_t)k % SK_NSLOT, it = (uint32_t)k / SK_NSLOT;
    if (k >= SK_NSLOT) mbar_wait(bar_empty + 8u * slot, (it & 1u) ^ 1u);   // the row that used the slot was emitted
    const uint32_t row = smem + slot * G::SLOT_BYTES + SK_LEAD;
    const uint32_t full = bar_full + 8u * slot;
    const uint32_t plan = smem + G::OFF_PLAN + slot * (G::PLAN_WORDS * 4);
    if (lane == 0u) {   // the pixel left of the strip (fetched one row ahead)
      if (BPW == 2) sts32(row - 8u, halo.x);
      sts32(row - 4u, halo.y);
    }
    bool comp = false, shift = false;
    const uint8_t *sp = nullptr;   // SK_SHIFT: first byte of the box in the source row
    if (has_col && y >= 0) {
      iv_seek(p, bc, y);
      const uint32_t dst = row + lane * (SK_BOX * 4);
      if (bc.kind == SK_DIRECT) {
        mbar_expect_tx(full, SK_BOX * 4);
        tma_box_2d(dst, reinterpret_cast<const uint8_t *>(p.maps) + (size_t)bc.tile * 128u, bc.cw, y + bc.cy, full);
      } else if (bc.kind == SK_BG) {
        mbar_expect_tx(full, SK_BOX * 4);
        bulk_copy(dst, p.bg_row, SK_BOX * 4, full);
      } else if (bc.kind == SK_SHIFT) {
        shift = true;
        sp = bc.pix + (size_t)(y + bc.cy) * bc.pitch + (size_t)bc.cw * 4u;
      } else {
        comp = true;
        sts32(plan + lane * 4u, bc.idx);
      }
    }
    // boxes whose source is not 16-byte aligned: the whole warp copies them word by word, asynchronously
    for (uint32_t m = __ballot_sync(0xffffffffu, shift); m; m &= m - 1u) {
      const int b = __ffs((int)m) - 1;
      const unsigned long long spb = __shfl_sync(0xffffffffu, (unsigned long long)sp, b);
      const uint32_t nw = __shfl_sync(0xffffffffu, bc.list_off, b);
      const uint32_t dstb = row + (uint32_t)b * (SK_BOX * 4) + lane * 4u;
      const uint8_t *srcb = reinterpret_cast<const uint8_t *>(spb) + lane * 4u;
#pragma unroll
      for (int i = 0; i < SK_BOX / 32; i++)
        if (lane + 32u * i < nw) cp_async4(dstb + 128u * i, srcb + 128 * i);
    }
    mbar_arrive_after_cp_async(full);   // every lane arrives once per row: when its copies, if any, have landed
    const uint32_t mask = __ballot_sync(0xffffffffu, comp);
    if (lane == 0u) sts32(plan + 32u * 4u, mask);
    mbar_arrive(bar_plan + 8u * slot);
    if (has_halo && k + 1 < n_k) halo = halo_pixel<BPW>(p, hc, ch, y + 1);
  }
}

// ---- consumers: their share of the compositing of ring entry k (row ya-1+k) --------------------------------------
template <int BPW, int Q>
__device__ __forceinline__ void strip_compose_share(const StripParams &p, const uint32_t smem, const uint32_t tid,
                                                    const uint32_t sx, const int ya, const int k) {
  using G = SKGeo<Q>;
  const uint32_t bar_full = smem + G::OFF_BAR, bar_plan = bar_full + 16 * SK_NSLOT;
  const int y = ya - 1 + k;
  const uint32_t slot = (uint32_t)k % SK_NSLOT, it = (uint32_t)k / SK_NSLOT;
  const uint32_t row = smem + slot * G::SLOT_BYTES + SK_LEAD;
  const uint32_t plan = smem + G::OFF_PLAN + slot * (G::PLAN_WORDS * 4);
  mbar_wait(bar_plan + 8u * slot, it & 1u);   // the slot is free and the producer has said what it does not load
  if (y < 0) {   // previousLine === null: zeros
#pragma unroll
    for (int i = 0; i < Q; i++) sts128(row + 16u * (tid + SK_CONS * i), make_uint4(0u, 0u, 0u, 0u));
  } else {
    const uint32_t mask = lds32(plan + 32u * 4u);
    if (mask) {
#pragma unroll
      for (int i = 0; i < Q; i++) {
        const uint32_t b = (tid >> 6) + 8u * i;
        if ((mask >> b) & 1u) {   // uniform for the warp: a box is two warps wide
          const StripIv &iv = p.ivs[lds32(plan + b * 4u)];
          const uint32_t j = tid + SK_CONS * i;
          const uint4 v = strip_compose_quad<BPW>(compose_ctx(p), p.lists + iv.list_off, iv.n_list, (int)(sx * G::SW + 4u * j), y);
          sts128(row + 16u * j, v);
        }
      }
    }
  }
  __syncwarp();
  if ((tid & 31u) == 0u) mbar_arrive(bar_full + 8u * slot);
}

// Winner of row r (first minimum in the order None, Sub, Up, Average, Paeth; png-filter.ts:174) once all `total`
// warps have added their sums.  pre: lane f's word as read earlier (count in the high half).
__device__ __forceinline__ uint32_t strip_winner(const StripParams &p, const int r, const uint32_t lane,
                                                 const uint32_t total, unsigned long long w) {
  const unsigned long long *rs = p.rowsum + (size_t)(r - (int)p.y_begin) * 8 + (lane < 5u ? lane : 0u);
  for (;;) {
    const bool ok = lane >= 5u || (uint32_t)(w >> 32) == total;
    if (__all_sync(0xffffffffu, ok)) break;
    __nanosleep(40);
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(rs) : "memory");
  }
  const uint32_t key = lane < 5u ? (((uint32_t)w << 3) | lane) : 0xffffffffu;
  return __reduce_min_sync(0xffffffffu, key) & 7u;
}

// The row above, as the sweep of the next row needs it.
template <int Q> struct RowCarry {
  uint4 raw[Q];     // the thread's quads, packed
  H4 hb[Q][4];      // ... unpacked to half2 (1024 + byte)
  H4 hl[Q][2];      // the two words left of each quad, unpacked ([0] only for 8-byte pixels)
};

template <int BPW, int Q>
__device__ __forceinline__ void carry_init(RowCarry<Q> &c, const uint32_t row, const uint32_t tid) {
#pragma unroll
  for (int i = 0; i < Q; i++) {
    const uint32_t a = row + 16u * (tid + SK_CONS * i);
    c.raw[i] = lds128(a);
    c.hb[i][0] = unpack_h4(c.raw[i].x); c.hb[i][1] = unpack_h4(c.raw[i].y);
    c.hb[i][2] = unpack_h4(c.raw[i].z); c.hb[i][3] = unpack_h4(c.raw[i].w);
    if (BPW == 2) c.hl[i][0] = unpack_h4(lds32(a - 8u));
    c.hl[i][1] = unpack_h4(lds32(a - 4u));
  }
}

// Sweep of one row: `ab` = the row above (in), `cu` = this row (out).  Adds the warp's sums to the row's words.
template <int BPW, int Q>
__device__ __forceinline__ void strip_sweep(const StripParams &p, const RowCarry<Q> &ab, RowCarry<Q> &cu, const uint32_t row,
                                            const uint32_t tid, const int xw0, const int wmax, const bool guard,
                                            const uint32_t taddr, const int y, const uint32_t (&lmask)[5]) {
  uint32_t s[5] = {0u, 0u, 0u, 0u, 0u};
#pragma unroll
  for (int i = 0; i < Q; i++) {
    const uint32_t j = tid + SK_CONS * i, a = row + 16u * j;
    const uint4 cur = lds128(a);
    const uint32_t lc1 = lds32(a - 4u);
    const uint32_t lc0 = BPW == 2 ? lds32(a - 8u) : 0u;
    const int qw = xw0 + 4 * (int)j;
    if (guard) sweep_quad<BPW, true>(cur, lc0, lc1, ab.raw[i], ab.hb[i], ab.hl[i], cu.hb[i], cu.hl[i], qw, wmax, taddr + 4u * i, s);
    else sweep_quad<BPW, false>(cur, lc0, lc1, ab.raw[i], ab.hb[i], ab.hl[i], cu.hb[i], cu.hl[i], qw, wmax, taddr + 4u * i, s);
    cu.raw[i] = cur;
  }
  uint32_t v = 0u;
#pragma unroll
  for (int f = 0; f < 5; f++) v |= __reduce_add_sync(0xffffffffu, s[f]) & lmask[f];
  const uint32_t lane = tid & 31u;
  if (lane < 5u) atomicAdd(p.rowsum + (size_t)(y - (int)p.y_begin) * 8 + lane, (1ull << 32) | (unsigned long long)v);
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
}

// Row y with winner ft: the filtered bytes, one 128-bit store per quad.  cur/up: the row and the row above in the ring.
template <int BPW, int Q>
__device__ __forceinline__ void strip_emit(const StripParams &p, const uint32_t ft, const int y, const uint32_t cur_row,
                                           const uint32_t up_row, const uint32_t taddr, const uint32_t tid, const uint32_t sx,
                                           const int xw0, const int wmax) {
  uint8_t *orow = p.out + (size_t)(y - (int)p.y_begin) * p.out_pitch + p.lead;
  if (sx == 0u && tid == 0u) orow[0] = (uint8_t)ft;
#pragma unroll
  for (int i = 0; i < Q; i++) {
    const uint32_t j = tid + SK_CONS * i, a = cur_row + 16u * j, b = up_row + 16u * j;
    const int qw = xw0 + 4 * (int)j;
    uint4 o;
    if (ft == 4u) {
      __syncwarp();   // tcgen05.ld is .sync.aligned: the partial stores of the quad before may have split the warp
      o = tmem_ld4(taddr + 4u * i);
    } else {
      o = lds128(a);
      if (ft != 0u) {
        uint4 pr;
        if (ft == 2u) {
          pr = lds128(b);
        } else {
          const uint32_t c1 = lds32(a - 4u);
          pr = BPW == 1 ? make_uint4(c1, o.x, o.y, o.z) : make_uint4(lds32(a - 8u), c1, o.x, o.y);
          if (ft == 3u) {
            const uint4 upq = lds128(b);
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
  }
}

template <int BPW, int Q>
__global__ void __launch_bounds__(SK_THREADS, 1) strip_kernel(const StripParams p) {
  using G = SKGeo<Q>;
  extern __shared__ __align__(128) uint8_t sk_smem[];
  const uint32_t smem = (uint32_t)__cvta_generic_to_shared(sk_smem);
  const uint32_t tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
  const uint32_t ry = blockIdx.x / p.nsx, sx = blockIdx.x - ry * p.nsx;
  const int ya = (int)(p.y_begin + ry * p.rows_per_range);
  const int yb = (int)min(p.y_end, (uint32_t)ya + p.rows_per_range);
  const int n_k = yb - ya + 1;   // ring entries: rows ya-1 .. yb-1
  const uint32_t bar_full = smem + G::OFF_BAR, bar_empty = bar_full + 8 * SK_NSLOT, bar_plan = bar_empty + 8 * SK_NSLOT;
  if (tid == 0u) {
    for (int s = 0; s < SK_NSLOT; s++) {
      mbar_init(bar_full + 8u * s, 32u + SK_WARPS);   // every producer lane + every consumer warp
      mbar_init(bar_empty + 8u * s, SK_WARPS);
      mbar_init(bar_plan + 8u * s, 32u);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  const uint32_t tmem_slot = smem + G::OFF_TMEM;
  if (warp == 0u) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "n"(G::TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = lds32(tmem_slot);

  if (warp == (uint32_t)SK_WARPS) {
    strip_producer<BPW, Q>(p, smem, lane, sx, ya, n_k);
  } else {
    const int xw0 = (int)(sx * G::SW), wmax = (int)p.canvas_w * BPW;
    const bool guard = xw0 + G::SW > wmax;
    const uint32_t total = p.nsx * SK_WARPS;
    // tensor memory: lane quarter warp % 4; column block (warp / 4, row parity)
    const uint32_t tm = tmem_base + (((warp & 3u) * 32u) << 16) + (warp >> 2) * (2u * Q * 4u);
    uint32_t lmask[5];
#pragma unroll
    for (int f = 0; f < 5; f++) {
      lmask[f] = lane == (uint32_t)f ? 0xffffffffu : 0u;
      asm volatile("" : "+r"(lmask[f]));
    }
    auto slot_row = [&](int k) { return smem + ((uint32_t)k % SK_NSLOT) * G::SLOT_BYTES + SK_LEAD; };
    auto wait_full = [&](int k) { mbar_wait(bar_full + 8u * ((uint32_t)k % SK_NSLOT), ((uint32_t)k / SK_NSLOT) & 1u); };
    auto poll = [&](int r) {
      unsigned long long w = 0ull;
      if (lane < 5u) asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(p.rowsum + (size_t)(r - (int)p.y_begin) * 8 + lane) : "memory");
      return w;
    };
    // one step of the pipeline for ring entry k (row y = ya-1+k), k >= 1
    auto step = [&](int k, const RowCarry<Q> &ab, RowCarry<Q> &cu) {
      const int y = ya - 1 + k;
      if (k + 1 < n_k) strip_compose_share<BPW, Q>(p, smem, tid, sx, ya, k + 1);
      unsigned long long pre = 0ull;
      if (k >= 2) pre = poll(y - 1);
      wait_full(k);
      strip_sweep<BPW, Q>(p, ab, cu, slot_row(k), tid, xw0, wmax, guard, tm + (uint32_t)(k & 1) * (Q * 4u), y, lmask);
      if (k >= 2) {
        const uint32_t ft = strip_winner(p, y - 1, lane, total, pre);
        strip_emit<BPW, Q>(p, ft, y - 1, slot_row(k - 1), slot_row(k - 2), tm + (uint32_t)((k - 1) & 1) * (Q * 4u), tid, sx, xw0, wmax);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // our reads of the slot, then the producer's bulk writes
        __syncwarp();
        if (lane == 0u) mbar_arrive(bar_empty + 8u * ((uint32_t)(k - 2) % SK_NSLOT));
      }
    };
    strip_compose_share<BPW, Q>(p, smem, tid, sx, ya, 0);
    strip_compose_share<BPW, Q>(p, smem, tid, sx, ya, 1);
    RowCarry<Q> ca, cb;
    wait_full(0);
    carry_init<BPW, Q>(ca, slot_row(0), tid);
    int k = 1;
    for (; k + 1 < n_k; k += 2) {
      step(k, ca, cb);
      step(k + 1, cb, ca);
    }
    if (k < n_k) { step(k, ca, cb); k++; }
    {  // the last row
      const int y = ya - 1 + (n_k - 1);
      const uint32_t ft = strip_winner(p, y, lane, total, poll(y));
      strip_emit<BPW, Q>(p, ft, y, slot_row(n_k - 1), slot_row(n_k - 2), tm + (uint32_t)((n_k - 1) & 1) * (Q * 4u), tid, sx, xw0, wmax);
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0u) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(G::TMEM_COLS) : "memory");
  }
}

template <int BPW, int Q>
static cudaError_t launch_strip_t(StripParams &p, int sm_count, cudaStream_t st) {
  using G = SKGeo<Q>;
  const uint32_t row_words = p.canvas_w * BPW, n_rows = p.y_end - p.y_begin;
  p.nsx = (row_words + G::SW - 1) / G::SW;
  cudaError_t e = cudaFuncSetAttribute(strip_kernel<BPW, Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM);   // per device: set at every launch
  if (e != cudaSuccess) return e;
  int occ = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, strip_kernel<BPW, Q>, SK_THREADS, G::SMEM);
  if (e != cudaSuccess) return e;
  if (occ < 1) return cudaErrorLaunchOutOfResources;
  const uint32_t resident = (uint32_t)sm_count;   // one CTA per SM: a CTA owns its SM's tensor-memory columns
  if (p.nsx > resident) return cudaErrorInvalidConfiguration;
  uint32_t nry = resident / p.nsx;
  if (nry > n_rows) nry = n_rows;
  p.rows_per_range = (n_rows + nry - 1) / nry;
  p.nry = (n_rows + p.rows_per_range - 1) / p.rows_per_range;   // no empty range
  void *args[] = {&p};
  return cudaLaunchCooperativeKernel((const void *)strip_kernel<BPW, Q>, dim3(p.nsx * p.nry), dim3(SK_THREADS), args, G::SMEM, st);
}

cudaError_t launch_strip(StripParams p, bool out16, int q, int device, int sm_count, cudaStream_t st) {
  (void)device;
  const uint32_t row_words = p.canvas_w * (out16 ? 2u : 1u);
  if (p.y_end <= p.y_begin || row_words == 0u) return cudaSuccess;
  if ((size_t)row_words * 4 >= ((size_t)1 << 22)) return cudaErrorInvalidConfiguration;   // cost sums must stay below 2^29
  if (q == 0) {   // wide rows: two quads per thread (half the strips, half the per-row overhead per quad)
    const uint32_t s1 = (row_words + 2047u) / 2048u, s2 = (row_words + 4095u) / 4096u;
    q = (s2 >= 2u && s2 * 2u <= s1 + (s1 + 9u) / 10u) ? 2 : 1;
  }
  if (out16) return q == 2 ? launch_strip_t<2, 2>(p, sm_count, st) : launch_strip_t<2, 1>(p, sm_count, st);
  return q == 2 ? launch_strip_t<1, 2>(p, sm_count, st) : launch_strip_t<1, 1>(p, sm_count, st);
}

}  // namespace stitch
