// decode.cu -- kernel (a): PNG unfilter (src/png-filter.ts:34-83) as a skewed
// wavefront, plus the Adam7 gather (src/adam7.ts:52-184).
//
// Dependency structure of unfiltering: byte i of row y needs byte i-bpp of the
// same row (Sub/Average/Paeth) and bytes i, i-bpp of row y-1 (Up/Average/Paeth).
// So one image (or one Adam7 pass) is a single chain in y, but pixel x of row
// y only needs pixel <= x of row y-1: rows can run concurrently when row y
// stays at least one pixel behind row y-1.  Lane k of a warp owns row 32g+k and
// at step t handles pixel t-k; a warp is a 32-row diagonal wavefront, the
// "up" pixel arrives from lane k-1 by one shuffle per step.  Consecutive
// 32-row groups of one chain run on different warps of the same CTA and are
// chained through the already written output row in global memory, guarded by
// a per-group progress word in shared memory, so all groups of an image are
// in flight at once (32 rows x 32 groups deep for a 1024-row tile).
#include "common.cuh"
#include "kernels.h"

namespace stitch {

static constexpr int UNF_WARPS = 16;  // warps per CTA = row groups in flight per chain

__device__ __forceinline__ uint32_t paeth4(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t r = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    int aa = (a >> (8 * k)) & 0xff, bb = (b >> (8 * k)) & 0xff, cc = (c >> (8 * k)) & 0xff;
    r |= (uint32_t)paeth1(aa, bb, cc) << (8 * k);
  }
  return r;
}

// Predictor for one pixel held as up to 8 bytes in (lo,hi).
__device__ __forceinline__ void predict(uint32_t ft, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                        uint32_t c_lo, uint32_t c_hi, uint32_t &p_lo, uint32_t &p_hi, bool wide) {
  switch (ft) {
    case 1: p_lo = a_lo; p_hi = a_hi; break;
    case 2: p_lo = b_lo; p_hi = b_hi; break;
    case 3: p_lo = avg4(a_lo, b_lo); p_hi = wide ? avg4(a_hi, b_hi) : 0u; break;
    case 4: p_lo = paeth4(a_lo, b_lo, c_lo); p_hi = wide ? paeth4(a_hi, b_hi, c_hi) : 0u; break;
    default: p_lo = 0u; p_hi = 0u; break;
  }
}

template <int BPP>
__device__ __forceinline__ void load_px(const uint8_t *p, uint32_t nbytes, uint32_t &lo, uint32_t &hi) {
  lo = 0u; hi = 0u;
#pragma unroll
  for (int k = 0; k < BPP; k++) {
    if ((uint32_t)k < nbytes) {
      uint32_t v = p[k];
      if (k < 4) lo |= v << (8 * k); else hi |= v << (8 * (k - 4));
    }
  }
}
template <int BPP>
__device__ __forceinline__ void store_px(uint8_t *p, uint32_t nbytes, uint32_t lo, uint32_t hi) {
#pragma unroll
  for (int k = 0; k < BPP; k++) {
    if ((uint32_t)k < nbytes) p[k] = (uint8_t)((k < 4 ? lo >> (8 * k) : hi >> (8 * (k - 4))) & 0xffu);
  }
}

// progress word: (group << 32) | pixels_done_of_the_group's_last_row
__device__ __forceinline__ unsigned long long ld_progress(const volatile unsigned long long *p) { return *p; }

template <int BPP>
__device__ void unfilter_chain(const UnfStream &s, volatile unsigned long long *progress, DevStatus *status) {
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  const uint32_t n_groups = (s.rows + 31u) / 32u;
  const uint32_t npx = (s.row_bytes + BPP - 1) / BPP;
  constexpr bool wide = BPP > 4;

  for (uint32_t g = warp; g < n_groups; g += UNF_WARPS) {
    const uint32_t r = g * 32u + lane;
    const bool valid = r < s.rows;
    const uint8_t *src_row = s.src + (size_t)(valid ? r : 0u) * s.src_stride;
    uint8_t *dst_row = s.dst + (size_t)(valid ? r : 0u) * s.dst_pitch;
    const uint8_t *up_row = s.dst + (size_t)(g * 32u - (g ? 1u : 0u)) * s.dst_pitch;  // row above lane 0
    uint32_t ft = valid ? src_row[0] : 0u;
    if (ft > 4u) {  // png-filter.ts:78-79
      // smallest key wins: first tile, then first row; low byte carries the offending value
      unsigned long long e = ((unsigned long long)(s.tile & 0x7fffffu) << 40) | ((unsigned long long)r << 8) | (ft & 0xffu);
      atomicMin(&status->first_error, e);
      ft = 0u;
    }
    // bytes of this row that exist in the stream (only an Adam7 final row can be short)
    const uint32_t row_avail = (valid && r == s.rows - 1u) ? s.last_row_bytes : s.row_bytes;
    const bool has_up = r > 0u;  // previousLine === null on the first row of the chain

    uint32_t left_lo = 0, left_hi = 0;  // my pixel x-1
    uint32_t ul_lo = 0, ul_hi = 0;      // row above, pixel x-1
    uint32_t out_lo = 0, out_hi = 0;    // what I produced in the previous step
    const volatile unsigned long long *dep = &progress[(g + UNF_WARPS - 1u) % UNF_WARPS];
    uint32_t dep_ready = 0;             // pixels of the row above lane 0 known to be complete

    const uint32_t n_steps = npx + 31u;
    for (uint32_t t = 0; t < n_steps; t++) {
      // lane 0 needs pixel t of the previous group's last row
      if (g > 0u && t < npx && t >= dep_ready) {
        const unsigned long long want = ((unsigned long long)(g - 1u) << 32) | (unsigned long long)(t + 1u);
        unsigned long long have;
        while ((have = ld_progress(dep)) < want) __nanosleep(40);
        dep_ready = (uint32_t)(have >> 32) > g - 1u ? npx : (uint32_t)(have & 0xffffffffu);
        __threadfence_block();
      }
      // "up" pixel: produced by lane-1 in the previous step
      uint32_t up_lo = __shfl_up_sync(0xffffffffu, out_lo, 1);
      uint32_t up_hi = wide ? __shfl_up_sync(0xffffffffu, out_hi, 1) : 0u;
      const int x = (int)t - (int)lane;
      const bool active = valid && x >= 0 && (uint32_t)x < npx;
      if (lane == 0u) {
        up_lo = 0u; up_hi = 0u;
        if (g > 0u && active) load_px<BPP>(up_row + (size_t)x * BPP, s.row_bytes - (uint32_t)x * BPP, up_lo, up_hi);
      }
      if (!has_up) { up_lo = 0u; up_hi = 0u; }
      if (active) {
        const uint32_t off = (uint32_t)x * BPP;
        const uint32_t nb = off < row_avail ? min((uint32_t)BPP, row_avail - off) : 0u;
        uint32_t f_lo, f_hi, p_lo, p_hi;
        load_px<BPP>(src_row + 1 + off, nb, f_lo, f_hi);
        predict(ft, left_lo, left_hi, up_lo, up_hi, ul_lo, ul_hi, p_lo, p_hi, wide);
        out_lo = add4(f_lo, p_lo);
        out_hi = wide ? add4(f_hi, p_hi) : 0u;
        if (nb < (uint32_t)BPP) {  // bytes past a truncated final row read as 0 (adam7.ts:87-88,141-143)
          const uint32_t keep_lo = nb >= 4u ? 0xffffffffu : ((1u << (8u * nb)) - 1u);
          const uint32_t keep_hi = nb <= 4u ? 0u : (nb >= 8u ? 0xffffffffu : ((1u << (8u * (nb - 4u))) - 1u));
          out_lo &= keep_lo; out_hi &= keep_hi;
        }
        store_px<BPP>(dst_row + off, min((uint32_t)BPP, s.row_bytes - off), out_lo, out_hi);
        left_lo = out_lo; left_hi = out_hi;
        ul_lo = up_lo; ul_hi = up_hi;
      }
      // publish progress of the group's last row (lane 31, or the last valid lane of a short group)
      if ((t & 31u) == 31u || t == n_steps - 1u) {
        __threadfence_block();
        __syncwarp();
        if (lane == 31u) {
          const uint32_t last_lane = min(31u, s.rows - 1u - g * 32u);
          const int done = (int)t + 1 - (int)last_lane;  // pixels finished on that row
          const uint32_t d = done < 0 ? 0u : min((uint32_t)done, npx);
          progress[g % UNF_WARPS] = ((unsigned long long)g << 32) | d;
        }
      }
    }
  }
}

__global__ void __launch_bounds__(UNF_WARPS * 32) unfilter_kernel(const UnfStream *streams, DevStatus *status) {
  __shared__ unsigned long long progress[UNF_WARPS];
  if (threadIdx.x < UNF_WARPS) progress[threadIdx.x] = 0ull;
  __syncthreads();
  const UnfStream s = streams[blockIdx.x];
  if (s.rows == 0u || s.row_bytes == 0u) return;
  switch (s.bpp) {
    case 1: unfilter_chain<1>(s, progress, status); break;
    case 2: unfilter_chain<2>(s, progress, status); break;
    case 3: unfilter_chain<3>(s, progress, status); break;
    case 4: unfilter_chain<4>(s, progress, status); break;
    case 6: unfilter_chain<6>(s, progress, status); break;
    case 8: unfilter_chain<8>(s, progress, status); break;
    default: break;
  }
}

// ---------------------------------------------------------------------------
// Adam7 gather: one thread per output byte.  Restates distributePassPixels /
// distributeSubBytePixels (adam7.ts:115-184) in gather form: the pass a pixel
// belongs to is a function of (x & 7, y & 7).
// ---------------------------------------------------------------------------
__constant__ uint8_t c_a7_pass[64] = {
    0, 5, 3, 5, 1, 5, 3, 5, 6, 6, 6, 6, 6, 6, 6, 6, 4, 5, 4, 5, 4, 5, 4, 5, 6, 6, 6, 6, 6, 6, 6, 6,
    2, 5, 3, 5, 2, 5, 3, 5, 6, 6, 6, 6, 6, 6, 6, 6, 4, 5, 4, 5, 4, 5, 4, 5, 6, 6, 6, 6, 6, 6, 6, 6};
__constant__ uint8_t c_a7_xs[7] = {0, 4, 0, 2, 0, 1, 0};
__constant__ uint8_t c_a7_ys[7] = {0, 0, 4, 0, 2, 0, 1};
__constant__ uint8_t c_a7_dx_log[7] = {3, 3, 2, 2, 1, 1, 0};
__constant__ uint8_t c_a7_dy_log[7] = {3, 3, 3, 2, 2, 1, 1};

__global__ void __launch_bounds__(256) adam7_gather_kernel(const A7Image *images) {
  const A7Image &im = images[blockIdx.z];
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;  // output byte in the row
  if (im.bit_depth >= 8u && (im.bpp & 3u) == 0u) {
    // pixels of 4 or 8 bytes: a thread moves one aligned 32-bit word (pass rows and output rows are 16-byte
    // aligned), so the stores of a warp are 128 contiguous bytes
    const uint32_t wpp = im.bpp >> 2;  // words per pixel
    if (i * 4u >= im.row_bytes) return;
    const uint32_t x = i / wpp, sub = i - x * wpp;
    for (uint32_t y = blockIdx.y; y < im.height; y += gridDim.y) {
      const uint32_t p = c_a7_pass[(y & 7u) * 8u + (x & 7u)];
      const uint32_t px = (x - c_a7_xs[p]) >> c_a7_dx_log[p], py = (y - c_a7_ys[p]) >> c_a7_dy_log[p];
      const uint32_t v = *reinterpret_cast<const uint32_t *>(im.pass_pix[p] + (size_t)py * im.pass_pitch[p] + (size_t)(px * wpp + sub) * 4u);
      *reinterpret_cast<uint32_t *>(im.dst + (size_t)y * im.dst_pitch + (size_t)i * 4u) = v;
    }
    return;
  }
  if (im.bit_depth >= 8u && im.bpp > 1u) {
    // pixels of 2, 3 or 6 bytes: a thread moves one pixel, the pass lookup is done once per pixel
    if (i >= im.width) return;
    const uint32_t bpp = im.bpp;
    for (uint32_t y = blockIdx.y; y < im.height; y += gridDim.y) {
      const uint32_t p = c_a7_pass[(y & 7u) * 8u + (i & 7u)];
      const uint32_t px = (i - c_a7_xs[p]) >> c_a7_dx_log[p], py = (y - c_a7_ys[p]) >> c_a7_dy_log[p];
      const uint8_t *src = im.pass_pix[p] + (size_t)py * im.pass_pitch[p] + (size_t)px * bpp;
      uint8_t *dst = im.dst + (size_t)y * im.dst_pitch + (size_t)i * bpp;
      if ((bpp & 1u) == 0u) {
        for (uint32_t b = 0; b < bpp; b += 2u) *reinterpret_cast<uint16_t *>(dst + b) = *reinterpret_cast<const uint16_t *>(src + b);
      } else {
        for (uint32_t b = 0; b < bpp; b++) dst[b] = src[b];
      }
    }
    return;
  }
  if (i >= im.row_bytes) return;
  for (uint32_t y = blockIdx.y; y < im.height; y += gridDim.y) {
    uint32_t v = 0;
    if (im.bit_depth >= 8u) {
      const uint32_t x = i / im.bpp, b = i - x * im.bpp;
      const uint32_t p = c_a7_pass[(y & 7u) * 8u + (x & 7u)];
      const uint32_t px = (x - c_a7_xs[p]) >> c_a7_dx_log[p], py = (y - c_a7_ys[p]) >> c_a7_dy_log[p];
      v = im.pass_pix[p][(size_t)py * im.pass_pitch[p] + px * im.bpp + b];
    } else {
      const uint32_t bd = im.bit_depth, ppb = 8u / bd, mask = (1u << bd) - 1u;
      for (uint32_t k = 0; k < ppb; k++) {
        const uint32_t x = i * ppb + k;
        if (x >= im.width) break;
        const uint32_t p = c_a7_pass[(y & 7u) * 8u + (x & 7u)];
        const uint32_t px = (x - c_a7_xs[p]) >> c_a7_dx_log[p], py = (y - c_a7_ys[p]) >> c_a7_dy_log[p];
        const uint32_t sb = im.pass_pix[p][(size_t)py * im.pass_pitch[p] + px / ppb];
        const uint32_t pv = (sb >> ((ppb - 1u - (px % ppb)) * bd)) & mask;
        v |= pv << ((ppb - 1u - k) * bd);
      }
    }
    im.dst[(size_t)y * im.dst_pitch + i] = (uint8_t)v;
  }
}

// ---- palette expansion (an extension: pixel-ops.ts:609-611 throws on colour type 3) ---------
// PNG specification, colour type 3: RGB from PLTE[index], alpha from tRNS[index] or 255.  The host has merged
// both into one 256-entry RGBA table; a block keeps it in shared memory, a thread expands four pixels into one
// 16-byte store.  Reads 1 B (or less) and writes 4 B per pixel.
__global__ void __launch_bounds__(256) expand_palette_kernel(const PalImage *__restrict__ images) {
  const PalImage &im = images[blockIdx.z];
  __shared__ uint32_t pal[256];
  pal[threadIdx.x] = im.pal[threadIdx.x];
  __syncthreads();
  const uint32_t x0 = (blockIdx.x * 256u + threadIdx.x) * 4u;
  if (x0 >= im.width) return;
  const uint32_t bd = im.bit_depth, mask = (1u << bd) - 1u;
  for (uint32_t y = blockIdx.y; y < im.height; y += gridDim.y) {
    const uint8_t *row = im.idx + (size_t)y * im.idx_pitch;
    uint32_t px[4];
#pragma unroll
    for (uint32_t k = 0; k < 4u; k++) {
      const uint32_t x = x0 + k < im.width ? x0 + k : im.width - 1u;
      const uint32_t bit = x * bd;
      px[k] = pal[(row[bit >> 3] >> (8u - (bit & 7u) - bd)) & mask];
    }
    uint8_t *o = im.dst + (size_t)y * im.dst_pitch + (size_t)x0 * 4u;
    if (x0 + 4u <= im.width) *reinterpret_cast<uint4 *>(o) = make_uint4(px[0], px[1], px[2], px[3]);
    else
      for (uint32_t k = 0; x0 + k < im.width; k++) reinterpret_cast<uint32_t *>(o)[k] = px[k];
  }
}

// ---- host launchers ---------------------------------------------------------

cudaError_t launch_unfilter(const UnfStream *d_streams, uint32_t n_streams, DevStatus *d_status, cudaStream_t st) {
  if (n_streams == 0) return cudaSuccess;
  unfilter_kernel<<<n_streams, UNF_WARPS * 32, 0, st>>>(d_streams, d_status);
  return cudaGetLastError();
}

cudaError_t launch_adam7_gather(const A7Image *d_images, uint32_t n_images, uint32_t max_row_bytes,
                                uint32_t max_height, cudaStream_t st) {
  if (n_images == 0) return cudaSuccess;
  dim3 grid((max_row_bytes + 255u) / 256u, max_height < 1024u ? (max_height ? max_height : 1u) : 1024u, n_images);
  adam7_gather_kernel<<<grid, 256, 0, st>>>(d_images);
  return cudaGetLastError();
}

cudaError_t launch_expand_palette(const PalImage *d_images, uint32_t n_images, uint32_t max_width, uint32_t max_height,
                                  cudaStream_t st) {
  if (n_images == 0) return cudaSuccess;
  dim3 grid((max_width + 1023u) / 1024u, max_height < 1024u ? (max_height ? max_height : 1u) : 1024u, n_images);
  expand_palette_kernel<<<grid, 256, 0, st>>>(d_images);
  return cudaGetLastError();
}

}  // namespace stitch
