// compose.cu -- kernel (b): normalise + place/composite tiles onto canvas rows
// (src/pixel-ops.ts:496-736, src/image-concat-core.ts:404-545, :551-685) and
// kernel (c): PNG filter selection (src/png-filter.ts:148-181).
#include "common.cuh"
#include "kernels.h"

namespace stitch {


// One canvas pixel: start from the background (createTransparentScanline,
// image-concat-core.ts:96-102) and apply every covering tile back to front.
template <bool OUT16>
__device__ __forceinline__ void compose_pixel(const ComposeParams &p, const uint32_t *list, uint32_t n_list, int x,
                                              int y, uint32_t &lo, uint32_t &hi) {
  lo = p.bg_lo; hi = p.bg_hi;
  for (uint32_t k = 0; k < n_list; k++) {
    const DevTile &t = p.tiles[list[k]];
    if (x < t.x0 || x >= t.x1 || y < t.y0 || y >= t.y1) continue;
    const uint32_t sx = (uint32_t)(x - t.x0 + t.sx0), sy = (uint32_t)(y - t.y0 + t.sy0);
    if (OUT16) {
      uint2 s = fetch_rgba16(t, sx, sy);
      uint2 d = p.alpha_blend ? blend16(make_uint2(lo, hi), s) : s;
      lo = d.x; hi = d.y;
    } else {
      uint32_t s = fetch_rgba8(t, sx, sy);
      lo = p.alpha_blend ? blend8(lo, s) : s;
    }
  }
}

template <bool OUT16>
__global__ void __launch_bounds__(128) compose_kernel(const ComposeParams p) {
  constexpr int PX = OUT16 ? 2 : 4;  // pixels per thread = 16 bytes of output
  const uint32_t chunk = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t x0 = chunk * PX;
  if (x0 >= p.canvas_w) return;
  for (uint32_t y = p.y_begin + blockIdx.y; y < p.y_end; y += gridDim.y) {
    const uint32_t bin = ((y - p.bin_y_origin) >> p.bin_h_log2) * p.bins_x + (x0 >> p.bin_w_log2);
    const uint32_t b0 = p.bin_start[bin], b1 = p.bin_start[bin + 1];
    const uint32_t *list = p.bin_tiles + b0;
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < PX; i++) {
      uint32_t lo = p.bg_lo, hi = p.bg_hi;
      if (x0 + i < p.canvas_w) compose_pixel<OUT16>(p, list, b1 - b0, (int)(x0 + i), (int)y, lo, hi);
      if (OUT16) { w[2 * i] = lo; w[2 * i + 1] = hi; } else { w[i] = lo; }
    }
    uint8_t *dst = p.out + (size_t)(y - p.y_begin) * p.out_pitch + (size_t)x0 * (OUT16 ? 8 : 4);
    if (x0 + PX <= p.canvas_w) {
      *reinterpret_cast<uint4 *>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
    } else {
      const uint32_t n_words = (p.canvas_w - x0) * (OUT16 ? 2u : 1u);
      for (uint32_t i = 0; i < n_words; i++) reinterpret_cast<uint32_t *>(dst)[i] = w[i];
    }
  }
}

cudaError_t launch_compose(const ComposeParams &p, bool out16, cudaStream_t st) {
  if (p.y_end <= p.y_begin || p.canvas_w == 0) return cudaSuccess;
  const uint32_t px = out16 ? 2u : 4u;
  const uint32_t chunks = (p.canvas_w + px - 1u) / px;
  const uint32_t rows = p.y_end - p.y_begin;
  dim3 grid((chunks + 127u) / 128u, rows < 32768u ? rows : 32768u);
  if (out16) compose_kernel<true><<<grid, 128, 0, st>>>(p);
  else compose_kernel<false><<<grid, 128, 0, st>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// kernel (c), stand-alone form: one CTA per output row.
// cost(candidate) = sum over bytes of |signed byte| (png-filter.ts:165-171);
// first minimum in the order None, Sub, Up, Average, Paeth wins (:174).
// ---------------------------------------------------------------------------

__device__ __forceinline__ uint32_t cost1(uint32_t v) { return v > 127u ? 256u - v : v; }

static constexpr int FILT_THREADS = 256;

__device__ __forceinline__ void block_reduce5(uint32_t s[5], uint32_t *smem /* 5*8 */) {
#pragma unroll
  for (int k = 0; k < 5; k++) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s[k] += __shfl_xor_sync(0xffffffffu, s[k], o);
  }
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < 5; k++) smem[k * 8 + warp] = s[k];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 5; k++) {
    uint32_t t = 0;
#pragma unroll
    for (int w = 0; w < FILT_THREADS / 32; w++) t += smem[k * 8 + w];
    s[k] = t;
  }
}

__device__ __forceinline__ uint32_t pick_filter(const uint32_t s[5]) {
  uint32_t best = 0;
#pragma unroll
  for (int k = 1; k < 5; k++) if (s[k] < s[best]) best = k;
  return best;
}

__global__ void __launch_bounds__(FILT_THREADS) filter_rows_kernel(const FilterParams p) {
  __shared__ uint32_t red[40];
  for (uint32_t r = blockIdx.x; r < p.n_rows; r += gridDim.x) {
    const uint8_t *cur = p.rows + (size_t)r * p.in_pitch;
    const uint8_t *up = r ? cur - p.in_pitch : p.prev0;
    const uint32_t n = p.row_bytes, bpp = p.bpp;
    uint32_t s[5] = {0, 0, 0, 0, 0};
    for (uint32_t i = threadIdx.x; i < n; i += FILT_THREADS) {
      const int x = cur[i];
      const int a = i >= bpp ? cur[i - bpp] : 0;
      const int b = up ? up[i] : 0;
      const int c = (up && i >= bpp) ? up[i - bpp] : 0;
      s[0] += cost1((uint32_t)x);
      s[1] += cost1((uint32_t)(x - a) & 0xffu);
      s[2] += cost1((uint32_t)(x - b) & 0xffu);
      s[3] += cost1((uint32_t)(x - ((a + b) >> 1)) & 0xffu);
      s[4] += cost1((uint32_t)(x - paeth1(a, b, c)) & 0xffu);
    }
    block_reduce5(s, red);
    const uint32_t ft = pick_filter(s);
    uint8_t *dst = p.out + (size_t)r * p.out_pitch + p.lead;
    if (threadIdx.x == 0) dst[0] = (uint8_t)ft;
    dst += 1;
    for (uint32_t i = threadIdx.x; i < n; i += FILT_THREADS) {
      const int x = cur[i];
      const int a = i >= bpp ? cur[i - bpp] : 0;
      const int b = up ? up[i] : 0;
      const int c = (up && i >= bpp) ? up[i - bpp] : 0;
      int pr;
      switch (ft) {
        case 1: pr = a; break;
        case 2: pr = b; break;
        case 3: pr = (a + b) >> 1; break;
        case 4: pr = paeth1(a, b, c); break;
        default: pr = 0; break;
      }
      dst[i] = (uint8_t)((x - pr) & 0xff);
    }
    __syncthreads();
  }
}

cudaError_t launch_filter_rows(const FilterParams &p, cudaStream_t st) {
  if (p.n_rows == 0) return cudaSuccess;
  const uint32_t grid = p.n_rows < 148u * 8u * 64u ? p.n_rows : 148u * 8u * 64u;
  filter_rows_kernel<<<grid, FILT_THREADS, 0, st>>>(p);
  return cudaGetLastError();
}

}  // namespace stitch
