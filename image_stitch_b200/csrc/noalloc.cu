// noalloc.cu -- stitch_filter_rows_noalloc: kernel (c) (filterScanline, src/png-filter.ts:148-181) for callers
// that cannot allocate device memory (CUDA graph capture, tight serving loops).  No counterpart in the
// reference.
//
// The split-row form of filter selection needs scratch: every row is cut into segments, each CTA adds the five
// candidate costs of its segment to the row's sums, and a second kernel picks the winner once all segments are
// in.  Those sums live in a file-scope static __device__ array of a fixed, modest size (1 MiB) which every GPU
// that loads the library holds until exit.  The entry point calls no cudaMalloc / cudaFree / pool routine; the
// caller passes device pointers for the rows and for the output.  Every call shares the array and it keeps the
// last call's contents, so each call clears the part it uses first; calls that overlap on one GPU race.
#include "job.h"

#include <mutex>

namespace stitch {

static constexpr uint32_t NA_MAX_ROWS = STITCH_NOALLOC_MAX_ROWS;   // rows per call
static constexpr uint32_t NA_SEG = 8192;                            // bytes of a row per CTA
static constexpr int NA_THREADS = 256;

// [row][8]: five 64-bit cost sums per row (None, Sub, Up, Average, Paeth), padded to 64 bytes
static __device__ __align__(256) unsigned long long g_na_rowsum[(size_t)NA_MAX_ROWS * 8];

__device__ __forceinline__ uint32_t na_cost1(uint32_t v) { return v > 127u ? 256u - v : v; }

__global__ void na_clear_kernel(unsigned long long *sums, const uint32_t n_words) {
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += gridDim.x * blockDim.x) sums[i] = 0ull;
}

// grid (segments, rows): costs of one segment of one row, added to the row's sums
__global__ void __launch_bounds__(NA_THREADS) na_cost_kernel(const uint8_t *rows, const size_t in_pitch, const uint8_t *prev0,
                                                             const uint32_t row_bytes, const uint32_t bpp,
                                                             unsigned long long *sums) {
  __shared__ uint32_t red[5][NA_THREADS / 32];
  const uint32_t r = blockIdx.y;
  const uint8_t *cur = rows + (size_t)r * in_pitch;
  const uint8_t *up = r ? cur - in_pitch : prev0;   // previousLine === null for the first row without prev0
  const uint32_t i0 = blockIdx.x * NA_SEG, i1 = min(row_bytes, i0 + NA_SEG);
  uint32_t s[5] = {0, 0, 0, 0, 0};
  for (uint32_t i = i0 + threadIdx.x; i < i1; i += NA_THREADS) {
    const int x = cur[i];
    const int a = i >= bpp ? cur[i - bpp] : 0;
    const int b = up ? up[i] : 0;
    const int c = (up && i >= bpp) ? up[i - bpp] : 0;
    s[0] += na_cost1((uint32_t)x);
    s[1] += na_cost1((uint32_t)(x - a) & 0xffu);
    s[2] += na_cost1((uint32_t)(x - b) & 0xffu);
    s[3] += na_cost1((uint32_t)(x - ((a + b) >> 1)) & 0xffu);
    s[4] += na_cost1((uint32_t)(x - paeth1(a, b, c)) & 0xffu);
  }
#pragma unroll
  for (int k = 0; k < 5; k++) {
    s[k] = __reduce_add_sync(0xffffffffu, s[k]);
    if ((threadIdx.x & 31u) == 0u) red[k][threadIdx.x >> 5] = s[k];
  }
  __syncthreads();
  if (threadIdx.x < 5u) {
    unsigned long long t = 0;
    for (int w = 0; w < NA_THREADS / 32; w++) t += red[threadIdx.x][w];
    atomicAdd(sums + (size_t)r * 8 + threadIdx.x, t);
  }
}

// grid (segments, rows): winner of the row (first minimum: None, Sub, Up, Average, Paeth; png-filter.ts:174),
// then the segment's filtered bytes; segment 0 also writes the filter-type byte
__global__ void __launch_bounds__(NA_THREADS) na_emit_kernel(const uint8_t *rows, const size_t in_pitch, const uint8_t *prev0,
                                                             const uint32_t row_bytes, const uint32_t bpp,
                                                             const unsigned long long *sums, uint8_t *out,
                                                             const size_t out_pitch) {
  const uint32_t r = blockIdx.y;
  const unsigned long long *rs = sums + (size_t)r * 8;
  uint32_t ft = 0;
  unsigned long long best = rs[0];
#pragma unroll
  for (uint32_t k = 1; k < 5; k++) if (rs[k] < best) { best = rs[k]; ft = k; }
  const uint8_t *cur = rows + (size_t)r * in_pitch;
  const uint8_t *up = r ? cur - in_pitch : prev0;
  uint8_t *dst = out + (size_t)r * out_pitch;
  if (blockIdx.x == 0 && threadIdx.x == 0) dst[0] = (uint8_t)ft;
  dst += 1;
  const uint32_t i0 = blockIdx.x * NA_SEG, i1 = min(row_bytes, i0 + NA_SEG);
  for (uint32_t i = i0 + threadIdx.x; i < i1; i += NA_THREADS) {
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
}

// The array's name is a host-side stand-in in host code: the device address is asked for once per device, with
// that device current, and cached.
static std::mutex g_na_mutex;
static unsigned long long *g_na_ptr[64] = {};

static cudaError_t na_scratch(int device, unsigned long long **ptr) {
  if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_na_mutex);
  if (!g_na_ptr[device]) {
    void *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_na_rowsum);   // the caller made `device` current
    if (e != cudaSuccess) return e;
    g_na_ptr[device] = static_cast<unsigned long long *>(p);
  }
  *ptr = g_na_ptr[device];
  return cudaSuccess;
}

}  // namespace stitch

using namespace stitch;

extern "C" int stitch_filter_rows_noalloc(stitch_ctx *ctx, const uint8_t *d_rows, size_t in_pitch, const uint8_t *d_prev_row,
                                          uint32_t n_rows, size_t row_bytes, uint32_t bpp, uint8_t *d_out, size_t out_pitch) {
  if (!ctx) return STITCH_E_INVALID;
  if (!d_rows || !d_out || !n_rows || !row_bytes || !bpp)
    return stitch_fail(ctx, STITCH_E_INVALID, "stitch_filter_rows_noalloc: null buffer or empty shape");
  if (n_rows > NA_MAX_ROWS)
    return stitch_fail(ctx, STITCH_E_CAPACITY, "stitch_filter_rows_noalloc: %u rows exceed the static scratch (at most %u rows per call)",
                       n_rows, NA_MAX_ROWS);
  if (row_bytes > STITCH_NOALLOC_MAX_ROW_BYTES)
    return stitch_fail(ctx, STITCH_E_CAPACITY, "stitch_filter_rows_noalloc: rows of %zu bytes exceed the limit of %u bytes",
                       row_bytes, (unsigned)STITCH_NOALLOC_MAX_ROW_BYTES);
  if (in_pitch < row_bytes || out_pitch < row_bytes + 1)
    return stitch_fail(ctx, STITCH_E_INVALID, "stitch_filter_rows_noalloc: pitch smaller than a row");
  STITCH_CK(cudaSetDevice(ctx->device));
  unsigned long long *sums = nullptr;
  STITCH_CK(na_scratch(ctx->device, &sums));
  const uint32_t n_words = n_rows * 8u;
  const dim3 grid((unsigned)((row_bytes + NA_SEG - 1) / NA_SEG), n_rows);
  na_clear_kernel<<<(n_words + 255u) / 256u, 256, 0, ctx->stream>>>(sums, n_words);   // the last call's sums are still there
  na_cost_kernel<<<grid, NA_THREADS, 0, ctx->stream>>>(d_rows, in_pitch, d_prev_row, (uint32_t)row_bytes, bpp, sums);
  na_emit_kernel<<<grid, NA_THREADS, 0, ctx->stream>>>(d_rows, in_pitch, d_prev_row, (uint32_t)row_bytes, bpp, sums, d_out, out_pitch);
  STITCH_CK(cudaGetLastError());
  STITCH_CK(cudaStreamSynchronize(ctx->stream));
  return STITCH_OK;
}
