// fused.cu -- kernels (b)+(c) in ONE pass over HBM: normalise + composite the
// tiles of a canvas segment into shared memory, compute the five PNG filter
// costs (src/png-filter.ts:148-181), agree on the per-row winner across the
// CTAs that share a row, and write the filtered scanlines.  HBM traffic is
// "read every visible tile pixel once, write every output byte once".
//
// Work decomposition: an ITEM is FB_R canvas rows x FB_SWW 32-bit words
// (2 KB of every row).  A row's cost is a sum over all items of its band, so
// items publish partial sums with red.global.add and bump a per-band counter;
// an item writes its rows once the counter says the band is complete.  Items
// are handed out in band-major order by an atomic ticket to a persistent grid
// that is fully resident (grid <= SMs x occupancy, and >= items per band), so
// every item a CTA waits for is already running: no deadlock, no grid sync.
//
// Arithmetic: bytes stay packed four to a register.  cost(candidate) =
// sum |signed byte| = VABSDIFF4.S8.ACC(|x - pred|, 0) with |x - pred| from
// VABSDIFF4.U8 (both native, 64 lanes/clk/SM on the ALU pipe); the Paeth
// predictor is evaluated two bytes at a time in half2 (exact for 9-bit
// integers) on the FMA pipe, which runs concurrently with the ALU pipe.
#include <cuda_fp16.h>

#include "common.cuh"
#include "kernels.h"

namespace stitch {

static constexpr int FB_R = 16;            // rows per item
static constexpr int FB_SWW = 512;         // 32-bit words per item row (2 KB): 512 px RGBA8 / 256 px RGBA16
static constexpr int FB_QPR = FB_SWW / 4;  // 16-byte quads per item row
static constexpr int FB_PITCH = FB_SWW + 4;  // raw row pitch in words: [.. left context][FB_SWW words]
static constexpr int FB_THREADS = 256;
static constexpr int FB_RAW_WORDS = (FB_R + 1) * FB_PITCH;
static constexpr int FB_PF_WORDS = FB_R * FB_SWW;
static constexpr size_t FB_SMEM = (size_t)(FB_RAW_WORDS + FB_PF_WORDS) * 4;

// ---- packed-byte helpers ------------------------------------------------------
__device__ __forceinline__ uint32_t absdiff_u4(uint32_t a, uint32_t b) { return __vabsdiffu4(a, b); }
// acc + sum over bytes of |sext8(v)|
__device__ __forceinline__ uint32_t acc_abs_s4(uint32_t v, uint32_t acc) {
  uint32_t r;
  asm("vabsdiff4.u32.s32.s32.add %0, %1, %2, %3;" : "=r"(r) : "r"(v), "r"(0u), "r"(acc));
  return r;
}
__device__ __forceinline__ uint32_t favg4(uint32_t a, uint32_t b) { return avg4(a, b); }

// ---- phase A: composite one 16-byte quad of canvas row y ------------------------
template <int BPW>
__device__ __forceinline__ uint4 compose_quad(const FusedParams &p, const uint32_t *list, uint32_t n_list, int xw, int y) {
  constexpr int PX = 4 / BPW;  // pixels per quad
  const int x = xw / BPW;
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < PX; i++) {
    if (BPW == 1) w[i] = p.bg_lo; else { w[2 * i] = p.bg_lo; w[2 * i + 1] = p.bg_hi; }
  }
  // start at the top-most tile that covers the whole quad and hides everything below it
  int k0 = 0;
  for (int k = (int)n_list - 1; k > 0; k--) {
    const DevTile &t = p.tiles[list[k]];
    if ((p.alpha_blend == 0u || t.opaque) && y >= t.y0 && y < t.y1 && x >= t.x0 && x + PX <= t.x1) { k0 = k; break; }
  }
  for (int k = k0; k < (int)n_list; k++) {
    const DevTile &t = p.tiles[list[k]];
    if (y < t.y0 || y >= t.y1 || x + PX <= t.x0 || x >= t.x1) continue;
    const uint32_t sy = (uint32_t)(y - t.y0 + t.sy0);
    const int sx0 = x - t.x0 + t.sx0;
    const bool inside = x >= t.x0 && x + PX <= t.x1;
    if (BPW == 1) {
      uint32_t s[4];
      if (inside && t.fmt == FMT_RGBA8) {
        const uint8_t *src = t.pix + (size_t)sy * t.pitch + 4u * (uint32_t)sx0;
        if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
          const uint4 v = __ldg(reinterpret_cast<const uint4 *>(src));
          s[0] = v.x; s[1] = v.y; s[2] = v.z; s[3] = v.w;
        } else {
#pragma unroll
          for (int i = 0; i < 4; i++) s[i] = __ldg(reinterpret_cast<const uint32_t *>(src) + i);
        }
#pragma unroll
        for (int i = 0; i < 4; i++) w[i] = p.alpha_blend ? blend8(w[i], s[i]) : s[i];
      } else {
#pragma unroll
        for (int i = 0; i < 4; i++) {
          if (x + i >= t.x0 && x + i < t.x1) {
            const uint32_t sp = fetch_rgba8(t, (uint32_t)(sx0 + i), sy);
            w[i] = p.alpha_blend ? blend8(w[i], sp) : sp;
          }
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 2; i++) {
        if (x + i >= t.x0 && x + i < t.x1) {
          const uint2 sp = fetch_rgba16(t, (uint32_t)(sx0 + i), sy);
          const uint2 d = p.alpha_blend ? blend16(make_uint2(w[2 * i], w[2 * i + 1]), sp) : sp;
          w[2 * i] = d.x; w[2 * i + 1] = d.y;
        }
      }
    }
  }
  // words past the right edge of the canvas do not exist
  const int wmax = (int)p.canvas_w * BPW;
#pragma unroll
  for (int j = 0; j < 4; j++) if (xw + j >= wmax) w[j] = 0u;
  return make_uint4(w[0], w[1], w[2], w[3]);
}

// ---- phase B helper: filtered quad for a known winner -----------------------------
template <int BPW>
__device__ __forceinline__ uint4 filter_quad(uint32_t ft, const uint4 &cur, const uint4 &up, const uint32_t *lc,
                                             const uint4 &pf) {
  const uint32_t x[4] = {cur.x, cur.y, cur.z, cur.w}, b[4] = {up.x, up.y, up.z, up.w};
  uint32_t o[4];
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const uint32_t a = j >= BPW ? x[j - BPW] : lc[j];
    switch (ft) {
      case 1: o[j] = sub4(x[j], a); break;
      case 2: o[j] = sub4(x[j], b[j]); break;
      case 3: o[j] = sub4(x[j], favg4(a, b[j])); break;
      default: o[j] = x[j]; break;
    }
  }
  if (ft == 4u) return pf;
  return make_uint4(o[0], o[1], o[2], o[3]);
}

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

template <int BPW>
__global__ void __launch_bounds__(FB_THREADS, 3) fused_kernel(const FusedParams p) {
  extern __shared__ __align__(16) uint32_t smem[];
  uint32_t *raw = smem;                  // (FB_R+1) rows: row 0 = predictor row, words [4-BPW,4) = left context
  uint32_t *pfw = smem + FB_RAW_WORDS;   // Paeth-filtered words of the item
  __shared__ uint32_t s_item;
  __shared__ uint32_t s_sum[FB_R][5];
  const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const uint32_t n_items = p.n_bands * p.nseg;
  const int wmax = (int)p.canvas_w * BPW;  // words per canvas row

  for (;;) {
    __syncthreads();  // previous item's smem is no longer in use
    if (tid == 0) s_item = atomicAdd(p.ticket, 1u);
    if (tid < FB_R * 5) (&s_sum[0][0])[tid] = 0u;
    __syncthreads();
    const uint32_t item = s_item;
    if (item >= n_items) break;
    const uint32_t band = item / p.nseg, seg = item - band * p.nseg;
    const uint32_t band_abs = p.band0 + band;
    const int ya = max((int)p.y_begin, (int)(band_abs * FB_R));
    const int yb = min((int)p.y_end, (int)((band_abs + 1) * FB_R));
    const int nrows = yb - ya;
    const int xw0 = (int)seg * FB_SWW;
    const uint32_t bin = band_abs * p.bins_x + seg;
    const uint32_t b0 = p.bin_start[bin], n_list = p.bin_start[bin + 1] - b0;
    const uint32_t *list = p.bin_tiles + b0;

    // ---- phase A: composite (nrows + 1) x FB_SWW words into shared memory ----
    for (int idx = (int)tid; idx < (nrows + 1) * FB_QPR; idx += FB_THREADS) {
      const int rr = idx / FB_QPR, q = idx - rr * FB_QPR;
      const int y = ya - 1 + rr;
      uint4 v = make_uint4(0u, 0u, 0u, 0u);  // previousLine === null above row 0: predictor bytes are 0
      if (y >= 0 && xw0 + 4 * q < wmax) v = compose_quad<BPW>(p, list, n_list, xw0 + 4 * q, y);
      *reinterpret_cast<uint4 *>(raw + rr * FB_PITCH + 4 + 4 * q) = v;
    }
    if ((int)tid <= nrows) {  // left context: the pixel just left of the segment (0 at the canvas edge)
      const int y = ya - 1 + (int)tid;
      uint32_t c0 = 0u, c1 = 0u;
      if (y >= 0 && xw0 > 0) {
        const int xw = xw0 - BPW, qa = xw & ~3;
        const uint4 v = compose_quad<BPW>(p, list, n_list, qa, y);
        const uint32_t vv[4] = {v.x, v.y, v.z, v.w};
        if (BPW == 1) c1 = vv[3]; else { c0 = vv[2]; c1 = vv[3]; }
      }
      raw[tid * FB_PITCH + 2] = c0;
      raw[tid * FB_PITCH + 3] = c1;
    }
    __syncthreads();

    // ---- filter costs: warp = (128-word strip, half of the rows), one quad per lane, sweeping down ----
    {
      const uint32_t strip = warp & 3u, grp = warp >> 2;
      const int r_begin = (int)grp * (FB_R / 2), r_end = min(nrows, r_begin + FB_R / 2);
      const int qw = (int)(strip * 128u + lane * 4u);  // word offset of my quad inside the item row
      uint4 up = make_uint4(0, 0, 0, 0);
      uint32_t lu[2] = {0u, 0u};  // words preceding `up` (context of the row above)
      H4 hb[4], hlu[2];           // unpacked up / up-left
      for (int r = r_begin; r < r_end; r++) {
        const uint32_t *row_up = raw + r * FB_PITCH + 4, *row_cur = row_up + FB_PITCH;
        if (r == r_begin) {
          up = *reinterpret_cast<const uint4 *>(row_up + qw);
          lu[0] = row_up[qw - 2]; lu[1] = row_up[qw - 1];
          hb[0] = unpack_h4(up.x); hb[1] = unpack_h4(up.y); hb[2] = unpack_h4(up.z); hb[3] = unpack_h4(up.w);
          hlu[0] = unpack_h4(lu[0]); hlu[1] = unpack_h4(lu[1]);
        }
        const uint4 cur = *reinterpret_cast<const uint4 *>(row_cur + qw);
        const uint32_t lc[2] = {row_cur[qw - 2], row_cur[qw - 1]};
        const uint32_t x[4] = {cur.x, cur.y, cur.z, cur.w}, b[4] = {up.x, up.y, up.z, up.w};
        H4 hx[4], hlc[2];
        hlc[0] = unpack_h4(lc[0]); hlc[1] = unpack_h4(lc[1]);
        uint32_t s0 = 0, s1 = 0, s2 = 0, s3 = 0, s4 = 0, pf[4];
#pragma unroll
        for (int j = 0; j < 4; j++) {
          hx[j] = unpack_h4(x[j]);
          const uint32_t a = j >= BPW ? x[j - BPW] : lc[2 - BPW + j];
          const H4 &ha = j >= BPW ? hx[j - BPW] : hlc[2 - BPW + j];
          const H4 &hc = j >= BPW ? hb[j - BPW] : hlu[2 - BPW + j];
          const uint32_t pk = paeth4_h(ha, hb[j], hc);
          pf[j] = sub4(x[j], pk);
          if (xw0 + qw + j < wmax) {  // words past the canvas edge carry no cost
            s0 = acc_abs_s4(x[j], s0);
            s1 = acc_abs_s4(absdiff_u4(x[j], a), s1);
            s2 = acc_abs_s4(absdiff_u4(x[j], b[j]), s2);
            s3 = acc_abs_s4(absdiff_u4(x[j], favg4(a, b[j])), s3);
            s4 = acc_abs_s4(pf[j], s4);
          }
        }
        *reinterpret_cast<uint4 *>(pfw + r * FB_SWW + qw) = make_uint4(pf[0], pf[1], pf[2], pf[3]);
        s0 = __reduce_add_sync(0xffffffffu, s0); s1 = __reduce_add_sync(0xffffffffu, s1);
        s2 = __reduce_add_sync(0xffffffffu, s2); s3 = __reduce_add_sync(0xffffffffu, s3);
        s4 = __reduce_add_sync(0xffffffffu, s4);
        if (lane < 5u) {
          const uint32_t v = lane == 0 ? s0 : lane == 1 ? s1 : lane == 2 ? s2 : lane == 3 ? s3 : s4;
          atomicAdd(&s_sum[r][lane], v);
        }
        // this row becomes the row above
        up = cur; lu[0] = lc[0]; lu[1] = lc[1];
#pragma unroll
        for (int j = 0; j < 4; j++) hb[j] = hx[j];
        hlu[0] = hlc[0]; hlu[1] = hlc[1];
      }
    }
    __syncthreads();
    // publish the item's partial sums, then count the item as done for its band
    if ((int)tid < nrows * 5) {
      const int r = (int)tid / 5, f = (int)tid - r * 5;
      atomicAdd(p.rowsum + (size_t)(ya + r - (int)p.y_begin) * 8 + f, s_sum[r][f]);
      __threadfence();
    }
    __syncthreads();
    if (tid == 0) {
      atomicAdd(p.band_done + band, 1u);
      while (ld_volatile_u32(p.band_done + band) < p.nseg) __nanosleep(100);
      __threadfence();
    }
    __syncthreads();

    // ---- phase B: one warp per row writes the winner's bytes ----
    for (int r = (int)warp; r < nrows; r += FB_THREADS / 32) {
      const uint32_t *rs = p.rowsum + (size_t)(ya + r - (int)p.y_begin) * 8;
      uint32_t best = __ldcg(rs), ft = 0;
#pragma unroll
      for (uint32_t f = 1; f < 5; f++) {
        const uint32_t v = __ldcg(rs + f);
        if (v < best) { best = v; ft = f; }   // strict '<': first minimum wins (png-filter.ts:174)
      }
      uint8_t *orow = p.out + (size_t)(ya + r - (int)p.y_begin) * p.out_pitch + p.lead;
      if (seg == 0 && lane == 0) orow[0] = (uint8_t)ft;
      uint32_t *odat = reinterpret_cast<uint32_t *>(orow + 1) + xw0;
      const uint32_t *row_up = raw + r * FB_PITCH + 4, *row_cur = row_up + FB_PITCH;
      for (int q = (int)lane; q < FB_QPR; q += 32) {
        const int qw = 4 * q;
        if (xw0 + qw >= wmax) break;
        const uint4 cur = *reinterpret_cast<const uint4 *>(row_cur + qw);
        const uint4 up = *reinterpret_cast<const uint4 *>(row_up + qw);
        const uint4 pf = *reinterpret_cast<const uint4 *>(pfw + r * FB_SWW + qw);
        const uint32_t lc[2] = {row_cur[qw - 2], row_cur[qw - 1]};
        const uint32_t lcc[2] = {lc[2 - BPW], lc[1]};
        const uint4 o = filter_quad<BPW>(ft, cur, up, lcc, pf);
        if (xw0 + qw + 4 <= wmax) {
          *reinterpret_cast<uint4 *>(odat + qw) = o;
        } else {
          const uint32_t ov[4] = {o.x, o.y, o.z, o.w};
          for (int j = 0; xw0 + qw + j < wmax; j++) odat[qw + j] = ov[j];
        }
      }
    }
  }
}

cudaError_t launch_fused(const FusedParams &p, bool out16, int sm_count, cudaStream_t st) {
  static int occ8 = -1, occ16 = -1;
  if (occ8 < 0) {
    cudaFuncSetAttribute(fused_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FB_SMEM);
    cudaFuncSetAttribute(fused_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FB_SMEM);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ8, fused_kernel<1>, FB_THREADS, FB_SMEM);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ16, fused_kernel<2>, FB_THREADS, FB_SMEM);
  }
  const int occ = out16 ? occ16 : occ8;
  if (occ < 1) return cudaErrorLaunchOutOfResources;
  const uint32_t resident = (uint32_t)(occ * sm_count);
  if (resident < p.nseg) return cudaErrorInvalidConfiguration;  // caller falls back to the two-pass path
  const uint32_t n_items = p.n_bands * p.nseg;
  const uint32_t grid = n_items < resident ? n_items : resident;
  if (out16) fused_kernel<2><<<grid, FB_THREADS, FB_SMEM, st>>>(p);
  else fused_kernel<1><<<grid, FB_THREADS, FB_SMEM, st>>>(p);
  return cudaGetLastError();
}

uint32_t fused_rows_per_band() { return FB_R; }
uint32_t fused_words_per_seg() { return FB_SWW; }

}  // namespace stitch
