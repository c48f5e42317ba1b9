// api.cu -- C ABI of libstitch_b200.so (include/stitch_b200.h): contexts, jobs,
// tile upload, binning, kernel launches, band download.  Host-side C++ only
// orchestrates; every pixel is touched by the CUDA kernels in decode.cu,
// compose.cu, fused.cu and jpeg.cu.  There is deliberately no CPU fallback.
#include "job.h"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace stitch;

static int fail(stitch_ctx *ctx, int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (ctx) ctx->err = buf;
  // a failed runtime call leaves its code as the thread's "last error" until someone reads it: read it here, or the
  // next call's cudaGetLastError() after its launches reports this failure as its own (found by tools/jpeg_corrupt_fuzz.py)
  if (code == STITCH_E_CUDA || code == STITCH_E_NOMEM) cudaGetLastError();
  return code;
}
int stitch_fail(stitch_ctx *ctx, int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (ctx) ctx->err = buf;
  if (code == STITCH_E_CUDA || code == STITCH_E_NOMEM) cudaGetLastError();   // see fail()
  return code;
}

#define CK(call)                                                                         \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess)                                                               \
      return fail(ctx, e_ == cudaErrorMemoryAllocation ? STITCH_E_NOMEM : STITCH_E_CUDA, \
                  "CUDA error %s at %s:%d (%s)", cudaGetErrorName(e_), __FILE__, __LINE__, cudaGetErrorString(e_)); \
  } while (0)

static inline size_t round_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static int samples_per_pixel(uint32_t ct) {
  switch (ct) { case 0: return 1; case 2: return 3; case 3: return 1; case 4: return 2; case 6: return 4; default: return -1; }
}
static size_t scanline_bytes(uint32_t w, uint32_t bd, uint32_t ct) {
  return (size_t)(((uint64_t)w * bd * (uint64_t)samples_per_pixel(ct) + 7) / 8);
}

static int dev_alloc(stitch_ctx *ctx, void **p, size_t bytes) {
  CK(cudaMallocAsync(p, bytes ? bytes : 16, ctx->stream));
  return STITCH_OK;
}
static void dev_free(stitch_ctx *ctx, void *p) {
  if (p) cudaFreeAsync(p, ctx->stream);
}

// ---- small tables and fills without the copy engines (see stitch_ctx::h_arena) ------------------------------------
__global__ void copy_words_kernel(uint32_t *dst, const uint32_t *src, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}
__global__ void fill_words_kernel(uint32_t *dst, uint32_t v, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = v;
}
// Rows of `row_bytes` bytes at src + r * pitch  ->  the same rows back to back (the host's layout, whose rows of
// 1 + W*bpp bytes have no alignment), four bytes per thread.
__global__ void pack_rows_kernel(uint32_t *dst, const uint8_t *src, size_t pitch, uint32_t row_bytes, size_t n_words, size_t total) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += (size_t)gridDim.x * blockDim.x) {
    const size_t o = i * 4;
    size_t r = o / row_bytes;
    uint32_t c = (uint32_t)(o - r * row_bytes), w = 0;
    const uint8_t *p = src + r * pitch;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      if (o + k < total) w |= (uint32_t)p[c] << (8 * k);
      if (++c == row_bytes) { c = 0; p += pitch; }
    }
    dst[i] = w;
  }
}
static constexpr size_t ARENA_BYTES = (size_t)32 << 20;

int stitch_upload_small(stitch_ctx *ctx, void *d_dst, const void *src, size_t bytes) {
  if (!bytes) return STITCH_OK;
  const size_t need = round_up(bytes, 16);
  if ((bytes & 3u) || !ctx->h_arena || ctx->arena_used + need > ctx->arena_cap) {   // too large for the arena: the copy engine after all
    CK(cudaMemcpyAsync(d_dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return STITCH_OK;
  }
  uint8_t *stage = ctx->h_arena + ctx->arena_used;   // pinned host memory is addressable from kernels (unified addressing)
  ctx->arena_used += need;
  memcpy(stage, src, bytes);
  const size_t n = bytes / 4;
  const unsigned blocks = (unsigned)std::min<size_t>((n + 255) / 256, 1024);
  copy_words_kernel<<<blocks, 256, 0, ctx->stream>>>((uint32_t *)d_dst, (const uint32_t *)stage, n);
  CK(cudaGetLastError());
  ctx->aux_launches++;
  return STITCH_OK;
}

int stitch_fill_words(stitch_ctx *ctx, void *d_dst, uint32_t value, size_t bytes) {
  const size_t n = bytes / 4;
  if (!n) return STITCH_OK;
  const unsigned blocks = (unsigned)std::min<size_t>((n + 255) / 256, 2048);
  fill_words_kernel<<<blocks, 256, 0, ctx->stream>>>((uint32_t *)d_dst, value, n);
  CK(cudaGetLastError());
  ctx->aux_launches++;
  return STITCH_OK;
}

int stitch_download_small(stitch_ctx *ctx, void *h_pinned_dst, const void *d_src, size_t bytes) {
  const size_t n = bytes / 4;
  if (!n) return STITCH_OK;
  copy_words_kernel<<<(unsigned)std::min<size_t>((n + 255) / 256, 64), 256, 0, ctx->stream>>>((uint32_t *)h_pinned_dst, (const uint32_t *)d_src, n);
  CK(cudaGetLastError());
  ctx->aux_launches++;
  return STITCH_OK;
}

void stitch_arena_reset(stitch_ctx *ctx) { ctx->arena_used = 0; }

// ---------------------------------------------------------------------------
extern "C" int stitch_abi_version(void) { return STITCH_ABI_VERSION; }

extern "C" int stitch_ctx_create(int device_id, stitch_ctx **out) {
  if (!out) return STITCH_E_INVALID;
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0 || device_id < 0 || device_id >= n) {
    cudaGetLastError();
    return STITCH_E_NODEVICE;
  }
  stitch_ctx *ctx = new stitch_ctx();
  ctx->device = device_id;
  if (cudaSetDevice(device_id) != cudaSuccess) { delete ctx; return STITCH_E_NODEVICE; }
  cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device_id);
  if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return STITCH_E_CUDA; }
  ctx->own_stream = true;
  for (auto &ev : ctx->ev) cudaEventCreate(&ev);
  cudaStreamCreateWithFlags(&ctx->up, cudaStreamNonBlocking);
  cudaStreamCreateWithFlags(&ctx->down, cudaStreamNonBlocking);
  cudaEventCreateWithFlags(&ctx->ev_chunk, cudaEventDisableTiming);
  cudaEventCreateWithFlags(&ctx->ev_down, cudaEventDisableTiming);
  if (!ctx->up || !ctx->down || !ctx->ev_chunk || !ctx->ev_down) { delete ctx; return STITCH_E_CUDA; }
  // keep freed device memory cached in the pool: jobs reuse it without new cudaMalloc calls
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, device_id) == cudaSuccess) {
    uint64_t thr = UINT64_MAX;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
  }
  cudaMalloc((void **)&ctx->d_status, sizeof(DevStatus));
  cudaHostAlloc((void **)&ctx->h_status, sizeof(DevStatus), cudaHostAllocDefault);
  if (!ctx->d_status || !ctx->h_status) { delete ctx; return STITCH_E_NOMEM; }
  if (cudaHostAlloc((void **)&ctx->h_arena, ARENA_BYTES, cudaHostAllocDefault) == cudaSuccess) ctx->arena_cap = ARENA_BYTES;
  else { ctx->h_arena = nullptr; cudaGetLastError(); }
  *out = ctx;
  return STITCH_OK;
}

extern "C" void stitch_ctx_destroy(stitch_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  cudaStreamSynchronize(ctx->up);
  cudaStreamSynchronize(ctx->down);
  for (auto &ev : ctx->ev) if (ev) cudaEventDestroy(ev);
  cudaEventDestroy(ctx->ev_chunk); cudaEventDestroy(ctx->ev_down);
  cudaStreamDestroy(ctx->up); cudaStreamDestroy(ctx->down);
  for (cudaStream_t js : ctx->jd_streams) if (js) cudaStreamDestroy(js);
  if (ctx->pinned_spare) cudaFreeHost(ctx->pinned_spare);
  if (ctx->d_status) cudaFree(ctx->d_status);
  if (ctx->h_status) cudaFreeHost(ctx->h_status);
  if (ctx->h_arena) cudaFreeHost(ctx->h_arena);
  if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
}

extern "C" int stitch_ctx_set_stream(stitch_ctx *ctx, void *cuda_stream) {
  if (!ctx) return STITCH_E_INVALID;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
  ctx->stream = (cudaStream_t)cuda_stream;
  ctx->own_stream = false;
  return STITCH_OK;
}

extern "C" const char *stitch_last_error(const stitch_ctx *ctx) { return ctx ? ctx->err.c_str() : "no context"; }

// ---------------------------------------------------------------------------
extern "C" int stitch_job_create(stitch_ctx *ctx, const stitch_job_desc *desc, stitch_job **out) {
  if (!ctx || !desc || !out) return STITCH_E_INVALID;
  *out = nullptr;
  if (desc->canvas_w == 0 || desc->canvas_h == 0) return fail(ctx, STITCH_E_INVALID, "Invalid dimensions: %ux%u", desc->canvas_w, desc->canvas_h);
  if (desc->out_bit_depth != 8 && desc->out_bit_depth != 16) return fail(ctx, STITCH_E_INVALID, "out_bit_depth must be 8 or 16");
  if (desc->out_kind > STITCH_OUT_PNG_DEFLATE) return fail(ctx, STITCH_E_INVALID, "unknown out_kind %u", desc->out_kind);
  if (desc->out_kind == STITCH_OUT_JPEG_ENTROPY) {
    if (desc->out_bit_depth != 8) return fail(ctx, STITCH_E_INVALID, "JPEG output requires 8-bit RGBA");
    if (desc->jpeg_quality < 1 || desc->jpeg_quality > 100)
      return fail(ctx, STITCH_E_INVALID, "Invalid quality: %u (must be 1-100)", desc->jpeg_quality);
  }
  if ((uint64_t)desc->canvas_w * (desc->out_bit_depth == 16 ? 8 : 4) >= (1ull << 25))
    return fail(ctx, STITCH_E_INVALID, "canvas rows above 32 MiB are not supported");
  stitch_job *job = new stitch_job();
  job->ctx = ctx;
  job->desc = *desc;
  if (job->desc.band_y1 == 0 || job->desc.band_y1 > desc->canvas_h) { job->desc.band_y0 = 0; job->desc.band_y1 = desc->canvas_h; }
  if (job->desc.band_y0 >= job->desc.band_y1) { delete job; return fail(ctx, STITCH_E_INVALID, "empty band"); }
  job->bpp = desc->out_bit_depth == 16 ? 8 : 4;
  { const char *e = getenv("STITCH_B200_TWO_PASS"); job->force_two_pass = e && e[0] == '1'; }
  cudaSetDevice(ctx->device);
  *out = job;
  return STITCH_OK;
}

static void free_tile(stitch_ctx *ctx, TileRec &t) {
  if (t.up_event) {  // the upload may still be in flight: frees below are ordered on ctx->stream
    if (!t.up_waited) cudaStreamWaitEvent(ctx->stream, t.up_event, 0);
    cudaEventDestroy(t.up_event);
    t.up_event = nullptr; t.up_waited = true;
  }
  if (t.d_pix && t.d_pix != t.d_src) dev_free(ctx, t.d_pix);
  if (t.d_src && !t.src_external) dev_free(ctx, t.d_src);
  if (t.d_z && !t.z_external) dev_free(ctx, t.d_z);
  dev_free(ctx, t.d_pass);
  if (t.d_idx && t.d_idx != t.d_src) dev_free(ctx, t.d_idx);
  t.d_pix = t.d_src = t.d_pass = t.d_z = t.d_idx = nullptr;
}

extern "C" void stitch_job_destroy(stitch_job *job) {
  if (!job) return;
  stitch_ctx *ctx = job->ctx;
  cudaSetDevice(ctx->device);
  for (auto &t : job->tiles) free_tile(ctx, t);
  dev_free(ctx, job->d_tiles); dev_free(ctx, job->d_bin_start); dev_free(ctx, job->d_bin_tiles);
  dev_free(ctx, job->d_streams); dev_free(ctx, job->d_a7); dev_free(ctx, job->d_pal);
  dev_free(ctx, job->d_slow); dev_free(ctx, job->d_units); dev_free(ctx, job->d_uprog);
  dev_free(ctx, job->d_canvas); dev_free(ctx, job->d_out); dev_free(ctx, job->d_sync); dev_free(ctx, job->d_plans); dev_free(ctx, job->d_ctab);
  dev_free(ctx, job->d_defl); dev_free(ctx, job->d_defl_b); dev_free(ctx, job->d_defl_scratch);
  dev_free(ctx, job->d_inf); dev_free(ctx, job->d_infres);
  dev_free(ctx, job->d_unit_left); dev_free(ctx, job->d_maps); dev_free(ctx, job->d_ivs); dev_free(ctx, job->d_col_start); dev_free(ctx, job->d_lists); dev_free(ctx, job->d_bgrow);
  dev_free(ctx, job->d_pack[0]); dev_free(ctx, job->d_pack[1]);
  for (auto &e : job->ev_pack) if (e) cudaEventDestroy(e);
  if (job->h_infres) cudaFreeHost(job->h_infres);
  if (job->h_jdstat) cudaFreeHost(job->h_jdstat);
  if (!job->pinned.empty())   // copies out of the chunks may still be queued
    for (cudaStream_t js : ctx->jd_streams) if (js) cudaStreamSynchronize(js);
  for (auto &pc : job->pinned) {
    if (!ctx->pinned_spare) { ctx->pinned_spare = pc.first; ctx->pinned_spare_cap = pc.second; }
    else cudaFreeHost(pc.first);
  }
  for (auto &e : job->ev_defl) if (e) cudaEventDestroy(e);
  if (job->h_defl) cudaFreeHost(job->h_defl);
  for (auto &q : job->d_jpeg) dev_free(ctx, q);
  if (job->h_jpeg_totals) cudaFreeHost(job->h_jpeg_totals);
  cudaStreamSynchronize(ctx->stream);
  delete job;
}

// Adam7 geometry (src/adam7.ts:23-43)
static const int A7_XS[7] = {0, 4, 0, 2, 0, 1, 0}, A7_YS[7] = {0, 0, 4, 0, 2, 0, 1};
static const int A7_DX[7] = {8, 8, 4, 4, 2, 2, 1}, A7_DY[7] = {8, 8, 8, 4, 4, 2, 2};
static uint32_t a7_extent(uint32_t dim, int start, int step) {
  return (int64_t)dim <= start ? 0u : (uint32_t)(((int64_t)dim - start + step - 1) / step);
}

// Colour type / bit depth pairs the path takes: the reference's (common.cuh src_fmt), and colour type 3 only
// for a tile that came with its palette (stitch_job_add_tile_palette, the 8f.4 extension).
static bool format_ok(const stitch_tile_desc &d, bool has_pal) {
  if (d.color_type == 3) return has_pal && (d.bit_depth == 1 || d.bit_depth == 2 || d.bit_depth == 4 || d.bit_depth == 8);
  return samples_per_pixel(d.color_type) > 0 && src_fmt(d.color_type, d.bit_depth) != FMT_INVALID;
}

// Validates the descriptor and stream length the way the reference does while
// reading, and prepares the unfilter chains.  d_src must already be set.
static int prepare_tile(stitch_ctx *ctx, stitch_job *job, TileRec &t, uint32_t tile_no) {
  const stitch_tile_desc &d = t.d;
  const int spp = samples_per_pixel(d.color_type);
  if (!format_ok(d, t.has_pal)) {
    if ((d.color_type == 3 && !t.has_pal) || spp < 0)  // pixel-ops.ts:609-611 (palette reaches convertScanline and throws)
      return fail(ctx, STITCH_E_COLOR, "Unsupported source color type: %u", d.color_type);
    return fail(ctx, STITCH_E_INVALID, "Unsupported bit depth %u for color type %u", d.bit_depth, d.color_type);
  }
  if (d.width == 0 || d.height == 0) return fail(ctx, STITCH_E_INVALID, "tile #%u has zero size", d.image_idx + 1);
  if (d.vis_w && d.vis_h) {
    if (d.dst_x < 0 || d.dst_y < 0 || (uint64_t)d.dst_x + d.vis_w > job->desc.canvas_w ||
        (uint64_t)d.dst_y + d.vis_h > job->desc.canvas_h || (uint64_t)d.src_off_x + d.vis_w > d.width ||
        (uint64_t)d.src_off_y + d.vis_h > d.height)
      return fail(ctx, STITCH_E_INVALID, "tile #%u rectangle is outside the canvas or its source", d.image_idx + 1);
  }
  const size_t rb = scanline_bytes(d.width, d.bit_depth, d.color_type);
  const uint32_t fbpp = (uint32_t)((spp * d.bit_depth + 7) / 8);
  if (rb >= (1ull << 31) || (t.has_pal && d.width >= (1u << 29))) return fail(ctx, STITCH_E_INVALID, "tile row too large");
  t.streams.clear();
  t.has_a7 = false;
  // the plane the unfilter chains (and the Adam7 gather) write: the pixels themselves, or palette indices
  uint8_t *plane = nullptr;
  uint32_t plane_pitch = (uint32_t)round_up(rb, 16);
  auto alloc_planes = [&]() -> int {
    int rc = STITCH_OK;
    if (t.has_pal) {
      t.pitch = (uint32_t)round_up((size_t)d.width * 4u, 16);
      t.idx_pitch = plane_pitch;
      if (d.filtered && (rc = dev_alloc(ctx, (void **)&t.d_idx, (size_t)plane_pitch * d.height))) return rc;
      plane = t.d_idx;
    } else {
      t.pitch = plane_pitch;
    }
    if ((rc = dev_alloc(ctx, (void **)&t.d_pix, (size_t)t.pitch * d.height))) return rc;
    if (!t.has_pal) plane = t.d_pix;
    return rc;
  };
  if (!d.filtered) {
    if (t.src_len < rb * d.height)  // image-concat-core.ts:427-448 (decoder ended early / wrong row length)
      return fail(ctx, STITCH_E_TRUNCATED, "Expected %u scanlines, decoded %zu", d.height, rb ? t.src_len / rb : 0);
    if (t.has_pal) {  // raw index rows: only the expansion is left to do
      plane_pitch = (uint32_t)rb;
      int rc = alloc_planes();
      if (rc) return rc;
      t.d_idx = t.d_src;
      t.decoded = false;
      return STITCH_OK;
    }
    t.d_pix = t.d_src;
    t.pitch = (uint32_t)rb;
    t.decoded = true;
    return STITCH_OK;
  }
  if (!d.interlace) {
    if (t.src_len < (rb + 1) * (size_t)d.height)  // png-decoder.ts:209-211
      return fail(ctx, STITCH_E_TRUNCATED, "Expected %u scanlines, decoded %zu", d.height, t.src_len / (rb + 1));
    int rc = alloc_planes();
    if (rc) return rc;
    UnfStream s{};
    s.src = t.d_src; s.dst = plane; s.src_stride = (uint32_t)rb + 1; s.row_bytes = (uint32_t)rb;
    s.dst_pitch = plane_pitch; s.rows = d.height; s.bpp = fbpp; s.last_row_bytes = (uint32_t)rb;
    s.tile = tile_no; s.pass = 0;
    s.src_end = t.d_src + round_up(t.src_len, 16);
    t.streams.push_back(s);
    return STITCH_OK;
  }
  // Adam7: walk the passes exactly like deinterlaceAdam7 (adam7.ts:61-107)
  size_t off = 0, scratch = 0;
  size_t pass_off[7], pass_scr[7];
  uint32_t pw[7], ph[7], prb[7], ppitch[7];
  uint32_t last_avail[7];
  for (int p = 0; p < 7; p++) {
    pw[p] = a7_extent(d.width, A7_XS[p], A7_DX[p]);
    ph[p] = a7_extent(d.height, A7_YS[p], A7_DY[p]);
    prb[p] = (uint32_t)(((uint64_t)pw[p] * d.bit_depth * spp + 7) / 8);
    ppitch[p] = (uint32_t)round_up(prb[p], 16);
    pass_off[p] = off; pass_scr[p] = scratch; last_avail[p] = prb[p];
    if (pw[p] == 0 || ph[p] == 0) { pw[p] = ph[p] = 0; continue; }  // adam7.ts:68-70
    for (uint32_t y = 0; y < ph[p]; y++) {
      if (off >= t.src_len)  // adam7.ts:79-82
        return fail(ctx, STITCH_E_TRUNCATED, "Unexpected end of decompressed data at pass %d, line %u", p + 1, y);
      off += 1;
      const size_t avail = std::min((size_t)prb[p], t.src_len - off);  // :87-88 slice shortens the last line
      if (y == ph[p] - 1) last_avail[p] = (uint32_t)avail;
      else if (avail < prb[p]) {  // a short line that is not the stream's last: the next line read throws
        return fail(ctx, STITCH_E_TRUNCATED, "Unexpected end of decompressed data at pass %d, line %u", p + 1, y + 1);
      }
      off += avail;
    }
    scratch += (size_t)ppitch[p] * ph[p];
  }
  int rc = dev_alloc(ctx, (void **)&t.d_pass, scratch + 16);
  if (rc) return rc;
  if ((rc = alloc_planes())) return rc;
  memset(&t.a7, 0, sizeof t.a7);
  for (int p = 0; p < 7; p++) {
    t.a7.pass_pix[p] = t.d_pass + pass_scr[p];
    t.a7.pass_pitch[p] = ppitch[p]; t.a7.pass_w[p] = pw[p]; t.a7.pass_h[p] = ph[p];
    if (!pw[p] || !ph[p]) continue;
    UnfStream s{};
    s.src = t.d_src + pass_off[p]; s.dst = t.d_pass + pass_scr[p]; s.src_stride = prb[p] + 1; s.row_bytes = prb[p];
    s.dst_pitch = ppitch[p]; s.rows = ph[p]; s.bpp = fbpp; s.last_row_bytes = last_avail[p];
    s.tile = tile_no; s.pass = (uint32_t)p + 1;
    s.src_end = t.d_src + round_up(t.src_len, 16);
    t.streams.push_back(s);
  }
  t.a7.dst = plane; t.a7.dst_pitch = plane_pitch; t.a7.width = d.width; t.a7.height = d.height;
  t.a7.bit_depth = d.bit_depth; t.a7.bpp = fbpp; t.a7.row_bytes = (uint32_t)rb;
  t.has_a7 = true;
  return STITCH_OK;
}

// Bytes of the inflated IDAT stream a descriptor promises: filter byte + row for every row of the image
// (png-decoder.ts:92) or of every non-empty Adam7 pass (adam7.ts:61-107).  0 for descriptors prepare_tile rejects.
static size_t inflated_len(const stitch_tile_desc &d, bool has_pal) {
  const int spp = samples_per_pixel(d.color_type);
  if (!format_ok(d, has_pal) || !d.width || !d.height) return 0;
  if (!d.interlace) return (scanline_bytes(d.width, d.bit_depth, d.color_type) + 1) * (size_t)d.height;
  size_t total = 0;
  for (int p = 0; p < 7; p++) {
    const uint32_t pw = a7_extent(d.width, A7_XS[p], A7_DX[p]), ph = a7_extent(d.height, A7_YS[p], A7_DY[p]);
    if (!pw || !ph) continue;
    total += (size_t)ph * (1 + (size_t)(((uint64_t)pw * d.bit_depth * spp + 7) / 8));
  }
  return total;
}

static int add_tile_common(stitch_job *job, const stitch_tile_desc *tile, const uint8_t *host, const void *dev, size_t len,
                           const uint32_t *pal = nullptr, bool pal_opaque = true, bool own_dev = false, cudaStream_t produced_on = nullptr) {
  if (!job || !tile || (!host && !dev && len)) return STITCH_E_INVALID;
  stitch_ctx *ctx = job->ctx;
  cudaSetDevice(ctx->device);
  if (tile->filtered > 2u) return fail(ctx, STITCH_E_INVALID, "tile.filtered must be 0, 1 or 2");
  const bool z = tile->filtered == 2u;  // 8f.2: `data` is the zlib stream; the inflated stream is made on the device
  TileRec t;
  t.d = *tile;
  if (pal) {
    t.has_pal = true;
    t.pal_opaque = pal_opaque;
    memcpy(t.pal, pal, sizeof t.pal);
  }
  if (z) {
    t.d.filtered = 1;
    t.src_len = inflated_len(*tile, t.has_pal);
    t.z_len = len;
    t.inflated = false;
    if (len >= (1ull << 32) || t.src_len >= (1ull << 32)) return fail(ctx, STITCH_E_INVALID, "zlib streams of 4 GiB and more are not supported");
    if (dev) { t.d_z = (uint8_t *)dev; t.z_external = true; }
    else CK(cudaMallocAsync((void **)&t.d_z, round_up(len, 16) + 32, ctx->up));
    CK(cudaMallocAsync((void **)&t.d_src, round_up(t.src_len, 16) + 32, ctx->up));
  } else {
    t.src_len = len;
    if (dev) {
      t.d_src = (uint8_t *)dev;
      t.src_external = !own_dev;   // own_dev: rows a kernel chain on ctx->up is producing (a decoded JPEG tile)
    } else {
      // +32: the unfilter kernel may read a few bytes past a row with vector loads.  Allocated in the order of
      // the upload stream, where the bytes are written; ctx->stream sees it through the tile's upload event.
      CK(cudaMallocAsync((void **)&t.d_src, round_up(len, 16) + 32, ctx->up));
    }
  }
  int rc = prepare_tile(ctx, job, t, (uint32_t)job->tiles.size());
  if (rc) {
    // rows a kernel chain is still producing stay the caller's to free, on the chain's stream (it frees them when this
    // call fails; freeing them here as well was a double free whose error surfaced in the NEXT call: tools/jpeg_corrupt_fuzz.py)
    if (own_dev) t.src_external = true;
    free_tile(ctx, t);
    return rc;
  }
  if (host && len) {
    cudaError_t e = cudaMemcpyAsync(z ? t.d_z : t.d_src, host, len, cudaMemcpyHostToDevice, ctx->up);
    if (e != cudaSuccess) { free_tile(ctx, t); return fail(ctx, STITCH_E_CUDA, "H2D failed: %s", cudaGetErrorString(e)); }
  }
  if (!dev || z || own_dev) {  // also orders the allocation of an inflate target made on the upload stream
    cudaError_t e = cudaEventCreateWithFlags(&t.up_event, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventRecord(t.up_event, own_dev && produced_on ? produced_on : ctx->up);
    if (e != cudaSuccess) { if (own_dev) t.src_external = true; free_tile(ctx, t); return fail(ctx, STITCH_E_CUDA, "H2D failed: %s", cudaGetErrorString(e)); }
    t.up_waited = false;
  }
  job->tiles.push_back(std::move(t));
  job->finalized = false;
  return STITCH_OK;
}

extern "C" int stitch_job_add_tile(stitch_job *job, const stitch_tile_desc *tile, const uint8_t *data, size_t len) {
  return add_tile_common(job, tile, data, nullptr, len);
}
// 8f.4 extension: a colour type 3 tile with its PLTE (3 bytes per entry) and tRNS (1 byte per entry) chunk bodies.
extern "C" int stitch_job_add_tile_palette(stitch_job *job, const stitch_tile_desc *tile, const uint8_t *data, size_t len,
                                           const uint8_t *plte, size_t plte_len, const uint8_t *trns, size_t trns_len) {
  if (!job || !tile) return STITCH_E_INVALID;
  stitch_ctx *ctx = job->ctx;
  if (tile->color_type != 3) return fail(ctx, STITCH_E_INVALID, "a palette was given for color type %u", tile->color_type);
  if (!plte || plte_len == 0 || plte_len % 3 != 0 || plte_len > 768)  // PNG specification 11.2.3
    return fail(ctx, STITCH_E_INVALID, "PLTE must hold 1 to 256 entries of 3 bytes, got %zu bytes", plte_len);
  if ((!trns && trns_len) || trns_len > plte_len / 3)              // 11.3.2.1
    return fail(ctx, STITCH_E_INVALID, "tRNS holds %zu entries, PLTE %zu", trns_len, plte_len / 3);
  uint32_t pal[256];
  bool opaque = true;
  for (size_t i = 0; i < 256; i++) {
    uint32_t v = 0xFF000000u;
    if (i < plte_len / 3) v |= (uint32_t)plte[3 * i] | ((uint32_t)plte[3 * i + 1] << 8) | ((uint32_t)plte[3 * i + 2] << 16);
    if (i < trns_len) { v = (v & 0x00FFFFFFu) | ((uint32_t)trns[i] << 24); opaque = opaque && trns[i] == 255; }
    pal[i] = v;
  }
  return add_tile_common(job, tile, data, nullptr, len, pal, opaque);
}
int stitch_add_tile_produced(stitch_job *job, const stitch_tile_desc *tile, uint8_t *d_rows, size_t len, cudaStream_t produced_on) {
  return add_tile_common(job, tile, nullptr, d_rows, len, nullptr, true, true, produced_on);
}
extern "C" int stitch_job_add_tile_device(stitch_job *job, const stitch_tile_desc *tile, const void *device_data, size_t len) {
  if (((uintptr_t)device_data & 15u) != 0) return fail(job ? job->ctx : nullptr, STITCH_E_INVALID, "device_data must be 16-byte aligned");
  return add_tile_common(job, tile, nullptr, device_data, len);
}

// ---------------------------------------------------------------------------
template <class T>
static int upload_array(stitch_ctx *ctx, const std::vector<T> &v, T **d_ptr, size_t *cap) {
  if (v.size() > *cap) {
    dev_free(ctx, *d_ptr);
    *d_ptr = nullptr;
    *cap = 0;
    int rc = dev_alloc(ctx, (void **)d_ptr, v.size() * sizeof(T));
    if (rc) return rc;
    *cap = v.size();
  }
  static_assert(sizeof(T) % 4 == 0, "tables are copied word by word");
  if (!v.empty()) return stitch_upload_small(ctx, *d_ptr, v.data(), v.size() * sizeof(T));
  return STITCH_OK;
}

static int reset_status(stitch_ctx *ctx) {
  return stitch_fill_words(ctx, ctx->d_status, 0xffffffffu, sizeof(DevStatus));
}

// Results of the inflate launches since the last check (stream synchronised by the caller): zlib's message
// the way the reference reports it (`Inflate error: ${msg}`, streaming-inflate.ts:60-62), or the row-count
// errors of png-decoder.ts:209-211 / adam7.ts:79-82 when the stream ends early.
static int check_inflate(stitch_ctx *ctx, stitch_job *job) {
  int rc = STITCH_OK;
  for (uint32_t idx : job->inf_pending) {
    if (rc) break;
    const InfResult &r = job->h_infres[idx];
    const TileRec &t = job->tiles[idx];
    if (r.status) {   // (even with every byte of the image out: the window in which the error was met may have made some of them)
      job->status = STITCH_E_INFLATE;
      rc = fail(ctx, STITCH_E_INFLATE, "Inflate error: %s", inflate_message(r.status));
    } else if (r.out_len < t.src_len) {
      job->status = STITCH_E_TRUNCATED;
      const size_t rb = scanline_bytes(t.d.width, t.d.bit_depth, t.d.color_type);
      if (!t.d.interlace) rc = fail(ctx, STITCH_E_TRUNCATED, "Expected %u scanlines, decoded %zu", t.d.height, (size_t)r.out_len / (rb + 1));
      else {
        const int spp = samples_per_pixel(t.d.color_type);
        size_t off = 0;
        for (int p = 0; p < 7 && !rc; p++) {
          const uint32_t pw = a7_extent(t.d.width, A7_XS[p], A7_DX[p]), ph = a7_extent(t.d.height, A7_YS[p], A7_DY[p]);
          if (!pw || !ph) continue;
          const size_t line = 1 + (size_t)(((uint64_t)pw * t.d.bit_depth * spp + 7) / 8);
          for (uint32_t y = 0; y < ph; y++, off += line)
            if (off + line > r.out_len) { rc = fail(ctx, STITCH_E_TRUNCATED, "Unexpected end of decompressed data at pass %d, line %u", p + 1, y); break; }
        }
      }
    }
  }
  job->inf_pending.clear();
  return rc;
}

// Reads the device error word (stream must be synchronised by the caller).
static int check_status(stitch_ctx *ctx, stitch_job *job) {
  if (job && !job->inf_pending.empty()) {
    const int rc = check_inflate(ctx, job);
    if (rc) return rc;
  }
  if (job && !job->jd_pending.empty()) {   // JPEG tiles whose decode chain ctx->stream has waited for
    size_t keep = 0;
    int rc = STITCH_OK;
    for (const auto &pr : job->jd_pending) {
      if (pr.second < job->tiles.size() && !job->tiles[pr.second].up_waited) { job->jd_pending[keep++] = pr; continue; }
      const uint32_t code = job->h_jdstat[pr.first];
      if (code && !rc) {
        job->status = STITCH_E_JPEG;
        rc = fail(ctx, STITCH_E_JPEG, "%s", stitch_jpeg_decode_message(code));
      }
    }
    job->jd_pending.resize(keep);
    if (rc) return rc;
  }
  const unsigned long long e = ctx->h_status->first_error;
  if (e == ~0ull) return STITCH_OK;
  const uint32_t ft = (uint32_t)(e & 0xffu);
  if (job) job->status = STITCH_E_FILTER;
  return fail(ctx, STITCH_E_FILTER, "Unknown filter type: %u", ft);  // png-filter.ts:78-79
}

// ctx->stream waits for the uploads of the tiles that touch canvas rows [ya, yb).
static int wait_uploads(stitch_job *job, uint32_t ya, uint32_t yb) {
  stitch_ctx *ctx = job->ctx;
  for (auto &t : job->tiles) {
    if (t.up_waited) continue;
    const int64_t ty0 = t.d.dst_y, ty1 = (int64_t)t.d.dst_y + t.d.vis_h;
    if (t.d.vis_h && (ty1 <= (int64_t)ya || ty0 >= (int64_t)yb)) continue;
    CK(cudaStreamWaitEvent(ctx->stream, t.up_event, 0));
    t.up_waited = true;
  }
  return STITCH_OK;
}

// 8f.2: inflates, one warp per stream, the zlib tiles that touch canvas rows [ya, yb) and are not inflated yet
// (all of them again with `force`).  Results are read by check_status after the caller's next synchronisation.
static int inflate_tiles(stitch_job *job, int force, uint32_t *launches, uint32_t ya = 0, uint32_t yb = 0xffffffffu) {
  stitch_ctx *ctx = job->ctx;
  std::vector<InfStream> v;
  for (uint32_t i = 0; i < job->tiles.size(); i++) {
    TileRec &t = job->tiles[i];
    if (!t.d_z || (t.inflated && !force)) continue;
    if (t.d.vis_h && ((int64_t)t.d.dst_y + t.d.vis_h <= (int64_t)ya || (int64_t)t.d.dst_y >= (int64_t)yb)) continue;
    InfStream s{};
    s.src = t.d_z; s.dst = t.d_src; s.src_len = (uint32_t)t.z_len; s.dst_cap = (uint32_t)t.src_len; s.slot = i;
    v.push_back(s);
    t.inflated = true;
    t.decoded = false;
  }
  if (v.empty()) return STITCH_OK;
  int rc = upload_array(ctx, v, &job->d_inf, &job->inf_cap);
  if (rc) return rc;
  const size_t n = job->tiles.size();
  if (n > job->infres_cap) {
    dev_free(ctx, job->d_infres); job->d_infres = nullptr; job->infres_cap = 0;
    if ((rc = dev_alloc(ctx, (void **)&job->d_infres, n * sizeof(InfResult)))) return rc;
    job->infres_cap = n;
  }
  if (n > job->h_infres_cap) {
    InfResult *h = nullptr;
    CK(cudaHostAlloc((void **)&h, n * sizeof(InfResult), cudaHostAllocDefault));
    if (job->h_infres) {  // results that are still to be checked live in the old mirror until the next sync copies them again
      CK(cudaStreamSynchronize(ctx->stream));
      memcpy(h, job->h_infres, job->h_infres_cap * sizeof(InfResult));
      cudaFreeHost(job->h_infres);
    }
    job->h_infres = h; job->h_infres_cap = n;
  }
  CK(launch_inflate(job->d_inf, job->d_infres, (uint32_t)v.size(), ctx->stream));
  if (launches) *launches += 2;
  // the results of this launch lie between its first and last slot (tiles are visited in order)
  if ((rc = stitch_download_small(ctx, job->h_infres + v.front().slot, job->d_infres + v.front().slot,
                                  (size_t)(v.back().slot - v.front().slot + 1) * sizeof(InfResult)))) return rc;
  for (auto &s : v) job->inf_pending.push_back(s.slot);
  return STITCH_OK;
}

// Unfilters (and de-interlaces) the tiles that touch canvas rows [ya, yb) and are not decoded yet.
static int decode_tiles(stitch_job *job, int force, uint32_t *launches, uint32_t ya = 0, uint32_t yb = 0xffffffffu) {
  stitch_ctx *ctx = job->ctx;
  int wrc = wait_uploads(job, ya, yb);
  if (wrc) return wrc;
  if ((wrc = inflate_tiles(job, 0, launches, ya, yb))) return wrc;
  std::vector<UnfStream> streams;
  std::vector<A7Image> a7;
  std::vector<PalImage> pal;
  uint32_t a7_max_rb = 0, a7_max_h = 0, pal_max_w = 0, pal_max_h = 0;
  for (auto &t : job->tiles) {
    if (!t.d.filtered && !t.has_pal) continue;
    if (t.decoded && !force) continue;
    if (t.d.vis_h && ((int64_t)t.d.dst_y + t.d.vis_h <= (int64_t)ya || (int64_t)t.d.dst_y >= (int64_t)yb)) continue;
    for (auto &s : t.streams) streams.push_back(s);
    if (t.has_a7) {
      a7.push_back(t.a7);
      a7_max_rb = std::max(a7_max_rb, t.a7.row_bytes);
      a7_max_h = std::max(a7_max_h, t.a7.height);
    }
    if (t.has_pal) {
      PalImage pi{};
      pi.idx = t.d_idx; pi.dst = t.d_pix; pi.idx_pitch = t.idx_pitch; pi.dst_pitch = t.pitch;
      pi.width = t.d.width; pi.height = t.d.height; pi.bit_depth = t.d.bit_depth;
      memcpy(pi.pal, t.pal, sizeof pi.pal);
      pal.push_back(pi);
      pal_max_w = std::max(pal_max_w, t.d.width);
      pal_max_h = std::max(pal_max_h, t.d.height);
    }
    t.decoded = true;
  }
  auto expand_palettes = [&]() -> int {  // after the index planes are whole
    if (pal.empty()) return STITCH_OK;
    int rc = upload_array(ctx, pal, &job->d_pal, &job->pal_cap);
    if (rc) return rc;
    CK(launch_expand_palette(job->d_pal, (uint32_t)pal.size(), pal_max_w, pal_max_h, ctx->stream));
    if (launches) (*launches)++;
    return STITCH_OK;
  };
  if (streams.empty()) return expand_palettes();
  // Chains whose final row is cut short (Adam7 slice semantics, adam7.ts:87-88) keep the legacy kernel;
  // everything else runs on the wavefront kernel, one unit per 32-row group, handed out group-major so
  // that all chains advance together and a unit's predecessor is always already running.
  std::vector<UnfStream> fast, slow;
  for (auto &s : streams) (s.last_row_bytes == s.row_bytes ? fast : slow).push_back(s);
  if (!fast.empty()) {
    std::stable_sort(fast.begin(), fast.end(), [](const UnfStream &a, const UnfStream &b) { return a.rows > b.rows; });
    uint32_t max_groups = 0;
    size_t seam_blocks = 0;  // mailbox offsets in units of one 32-word block (256 B)
    for (auto &s : fast) {
      s.prog_base = (uint32_t)seam_blocks;
      seam_blocks += unfilter_wave_seam_bytes(s.rows, s.row_bytes) / 256u;
      max_groups = std::max(max_groups, (s.rows + 31u) / 32u);
    }
    if (seam_blocks > 0xffffffffull) return fail(ctx, STITCH_E_INVALID, "too many unfilter blocks in one decode");
    // decode(force) of an unchanged tile set (benchmarks, retries) finds its tables already on the device
    const bool same = job->d_streams && job->d_units && job->h_fast.size() == fast.size() &&
                      memcmp(job->h_fast.data(), fast.data(), fast.size() * sizeof(UnfStream)) == 0;
    int rc = STITCH_OK;
    if (!same) {
      std::vector<UnfUnit> units;
      for (uint32_t g = 0; g < max_groups; g++)
        for (uint32_t i = 0; i < fast.size(); i++) {
          if (g >= (fast[i].rows + 31u) / 32u) break;  // sorted by rows: no later chain has this group either
          units.push_back(UnfUnit{i, g});
        }
      job->h_fast.clear();
      if ((rc = upload_array(ctx, fast, &job->d_streams, &job->streams_cap))) return rc;
      if ((rc = upload_array(ctx, units, &job->d_units, &job->units_cap))) return rc;
      job->h_fast = fast;
      job->n_fast_units = (uint32_t)units.size();
    }
    const size_t bytes = 256u + seam_blocks * 256u;
    if (bytes > job->uprog_cap) {
      dev_free(ctx, job->d_uprog); job->d_uprog = nullptr; job->uprog_cap = 0;
      if ((rc = dev_alloc(ctx, (void **)&job->d_uprog, bytes))) return rc;
      job->uprog_cap = bytes;
      if ((rc = stitch_fill_words(ctx, job->d_uprog, 0u, bytes))) return rc;  // tags of a fresh mailbox: epoch 0 = never written
    }
    if (++job->seam_epoch == 0u) {  // 2^32 decodes later the tags would repeat: start over
      if ((rc = stitch_fill_words(ctx, job->d_uprog, 0u, job->uprog_cap))) return rc;
      job->seam_epoch = 1u;
    }
    if ((rc = stitch_fill_words(ctx, job->d_uprog, 0u, 4))) return rc;  // ticket
    uint32_t uniform_bpp = fast[0].bpp;
    for (auto &s : fast) if (s.bpp != uniform_bpp) uniform_bpp = 0;
    CK(launch_unfilter_wave(job->d_streams, job->d_units, job->n_fast_units, job->d_uprog + 256, job->seam_epoch,
                            reinterpret_cast<uint32_t *>(job->d_uprog), ctx->d_status, ctx->sm_count, uniform_bpp,
                            ctx->stream));
    if (launches) (*launches)++;
  }
  if (!slow.empty()) {
    int rc = upload_array(ctx, slow, &job->d_slow, &job->slow_cap);
    if (rc) return rc;
    CK(launch_unfilter(job->d_slow, (uint32_t)slow.size(), ctx->d_status, ctx->stream));
    if (launches) (*launches)++;
  }
  if (!a7.empty()) {
    int rc = upload_array(ctx, a7, &job->d_a7, &job->a7_cap);
    if (rc) return rc;
    CK(launch_adam7_gather(job->d_a7, (uint32_t)a7.size(), a7_max_rb, a7_max_h, ctx->stream));
    if (launches) (*launches)++;
  }
  return expand_palettes();
}

extern "C" int stitch_job_inflate(stitch_job *job, int force, stitch_timings *t) {
  if (!job) return STITCH_E_INVALID;
  stitch_ctx *ctx = job->ctx;
  cudaSetDevice(ctx->device);
  if (ctx->arena_used) { cudaStreamSynchronize(ctx->stream); stitch_arena_reset(ctx); }  // an earlier call left early
  ctx->aux_launches = 0;
  uint32_t launches = 0;
  int rc = wait_uploads(job, 0, 0xffffffffu);
  if (rc) return rc;
  if ((rc = reset_status(ctx))) return rc;
  CK(cudaEventRecord(ctx->ev[0], ctx->stream));
  if ((rc = inflate_tiles(job, force, &launches))) return rc;
  CK(cudaEventRecord(ctx->ev[1], ctx->stream));
  CK(cudaMemcpyAsync(ctx->h_status, ctx->d_status, sizeof(DevStatus), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  stitch_arena_reset(ctx);
  if (t) {
    memset(t, 0, sizeof *t);
    cudaEventElapsedTime(&t->decode_ms, ctx->ev[0], ctx->ev[1]);
    t->total_ms = t->decode_ms;
    t->launches = launches + ctx->aux_launches;
  }
  return check_status(ctx, job);
}

extern "C" int stitch_inflate(stitch_ctx *ctx, const uint8_t *zdata, size_t zlen, uint8_t *out, size_t cap, size_t *out_len) {
  if (!ctx || (!zdata && zlen) || (!out && cap)) return STITCH_E_INVALID;
  if (out_len) *out_len = 0;
  if (zlen >= (1ull << 32) || cap >= (1ull << 32)) return fail(ctx, STITCH_E_INVALID, "zlib streams of 4 GiB and more are not supported");
  cudaSetDevice(ctx->device);
  uint8_t *d_z = nullptr, *d_o = nullptr;
  InfStream *d_s = nullptr;
  InfResult *d_r = nullptr;
  int rc = dev_alloc(ctx, (void **)&d_z, round_up(zlen, 16) + 32);
  if (!rc) rc = dev_alloc(ctx, (void **)&d_o, round_up(cap, 16) + 32);
  if (!rc) rc = dev_alloc(ctx, (void **)&d_s, sizeof(InfStream));
  if (!rc) rc = dev_alloc(ctx, (void **)&d_r, sizeof(InfResult));
  InfResult r{};
  if (!rc) {
    InfStream s{};
    s.src = d_z; s.dst = d_o; s.src_len = (uint32_t)zlen; s.dst_cap = (uint32_t)cap; s.slot = 0;
    cudaError_t e = zlen ? cudaMemcpyAsync(d_z, zdata, zlen, cudaMemcpyHostToDevice, ctx->stream) : cudaSuccess;
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_s, &s, sizeof s, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = launch_inflate(d_s, d_r, 1, ctx->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(&r, d_r, sizeof r, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    if (e == cudaSuccess && r.status == 0 && r.out_len) {
      e = cudaMemcpyAsync(out, d_o, r.out_len, cudaMemcpyDeviceToHost, ctx->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    }
    if (e != cudaSuccess) rc = fail(ctx, STITCH_E_CUDA, "inflate failed: %s", cudaGetErrorString(e));
  }
  dev_free(ctx, d_z); dev_free(ctx, d_o); dev_free(ctx, d_s); dev_free(ctx, d_r);
  if (rc) return rc;
  if (r.status) return fail(ctx, STITCH_E_INFLATE, "Inflate error: %s", inflate_message(r.status));
  if (out_len) *out_len = r.out_len;
  return (r.flags & 1u) ? fail(ctx, STITCH_E_CAPACITY, "output buffer too small") : STITCH_OK;
}

extern "C" int stitch_job_invalidate(stitch_job *job) {
  if (!job) return STITCH_E_INVALID;
  for (auto &t : job->tiles) t.decoded = false;
  return STITCH_OK;
}

extern "C" int stitch_job_decode(stitch_job *job, int force, stitch_timings *t) {
  if (!job) return STITCH_E_INVALID;
  stitch_ctx *ctx = job->ctx;
  cudaSetDevice(ctx->device);
  if (ctx->arena_used) { cudaStreamSynchronize(ctx->stream); stitch_arena_reset(ctx); }  // an earlier call left early
  ctx->aux_launches = 0;
  uint32_t launches = 0;
  int rc = reset_status(ctx);
  if (rc) return rc;
  CK(cudaEventRecord(ctx->ev[0], ctx->stream));
  rc = decode_tiles(job, force, &launches);
  if (rc) return rc;
  CK(cudaEventRecord(ctx->ev[1], ctx->stream));
  CK(cudaMemcpyAsync(ctx->h_status, ctx->d_status, sizeof(DevStatus), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  stitch_arena_reset(ctx);
  if (t) {
    memset(t, 0, sizeof *t);
    cudaEventElapsedTime(&t->decode_ms, ctx->ev[0], ctx->ev[1]);
    t->total_ms = t->decode_ms;
    t->launches = launches + ctx->aux_launches;
  }
  return check_status(ctx, job);
}

// Sort tiles back-to-front and bin their rectangles (positioned-layout.ts:201-242
// builds a per-row list; a 2-D bin grid serves the same purpose for a GPU).
static int finalize_job(stitch_job *job) {
  stitch_ctx *ctx = job->ctx;
  const stitch_job_desc &jd = job->desc;
  std::vector<uint32_t> order;
  for (uint32_t i = 0; i < job->tiles.size(); i++)
    if (job->tiles[i].d.vis_w && job->tiles[i].d.vis_h) order.push_back(i);
  std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) {
    const stitch_tile_desc &x = job->tiles[a].d, &y = job->tiles[b].d;
    if (x.z_index != y.z_index) return x.z_index < y.z_index;
    return x.image_idx < y.image_idx;
  });
  std::vector<DevTile> dt(order.size());
  for (size_t k = 0; k < order.size(); k++) {
    const TileRec &t = job->tiles[order[k]];
    DevTile &o = dt[k];
    o.pix = t.d_pix; o.pitch = t.pitch; o.fmt = t.has_pal ? (uint32_t)FMT_RGBA8 : src_fmt(t.d.color_type, t.d.bit_depth);
    o.x0 = t.d.dst_x; o.y0 = t.d.dst_y; o.x1 = t.d.dst_x + (int32_t)t.d.vis_w; o.y1 = t.d.dst_y + (int32_t)t.d.vis_h;
    o.sx0 = (int32_t)t.d.src_off_x; o.sy0 = (int32_t)t.d.src_off_y;
    o.opaque = t.has_pal ? (t.pal_opaque ? 1u : 0u) : (t.d.color_type == 0 || t.d.color_type == 2) ? 1u : 0u;
    o.pad = 0;
  }
  // Bins are the fused kernel's items (absolute grid: FB_R rows x FB_SWW words).  A bin also lists
  // tiles that touch the row above it or the pixel left of it (filter predictors are recomputed).
  const uint32_t bh = fused_rows_per_band(), bw = fused_words_per_seg() / (job->bpp / 4);
  job->bin_h_log2 = 0; while ((1u << job->bin_h_log2) < bh) job->bin_h_log2++;
  job->bin_w_log2 = 0; while ((1u << job->bin_w_log2) < bw) job->bin_w_log2++;
  job->bin_y_origin = 0;
  job->bins_x = (jd.canvas_w + bw - 1) / bw;
  job->bins_y = (jd.canvas_h + bh - 1) / bh;
  const size_t n_bins = (size_t)job->bins_x * job->bins_y;
  std::vector<uint32_t> start(n_bins + 1, 0);
  auto bin_range = [&](const DevTile &o, uint32_t &bx0, uint32_t &bx1, uint32_t &by0, uint32_t &by1) -> bool {
    if (o.x1 <= o.x0 || o.y1 <= o.y0) return false;
    bx0 = (uint32_t)o.x0 >> job->bin_w_log2; bx1 = std::min((uint32_t)o.x1 >> job->bin_w_log2, job->bins_x - 1);
    by0 = (uint32_t)o.y0 >> job->bin_h_log2; by1 = std::min((uint32_t)o.y1 >> job->bin_h_log2, job->bins_y - 1);
    return true;
  };
  for (const DevTile &o : dt) {
    uint32_t bx0, bx1, by0, by1;
    if (!bin_range(o, bx0, bx1, by0, by1)) continue;
    for (uint32_t by = by0; by <= by1; by++)
      for (uint32_t bx = bx0; bx <= bx1; bx++) start[(size_t)by * job->bins_x + bx + 1]++;
  }
  for (size_t i = 0; i < n_bins; i++) start[i + 1] += start[i];
  std::vector<uint32_t> fill(start.begin(), start.end() - 1), list(start[n_bins] ? start[n_bins] : 1);
  for (uint32_t k = 0; k < dt.size(); k++) {
    uint32_t bx0, bx1, by0, by1;
    if (!bin_range(dt[k], bx0, bx1, by0, by1)) continue;
    for (uint32_t by = by0; by <= by1; by++)
      for (uint32_t bx = bx0; bx <= bx1; bx++) list[fill[(size_t)by * job->bins_x + bx]++] = k;
  }
  // A bin is "plain" when it is empty or when its top-most tile covers it, overwrites what is below and is in
  // the output's own format: its units then take the fused kernel's copy path.  render_rows() switches to
  // composite-then-filter when too many bins of a band are not.
  job->busy_bins.assign(job->bins_y + 1, 0);
  const uint32_t native = jd.out_bit_depth == 16 ? (uint32_t)FMT_RGBA16 : (uint32_t)FMT_RGBA8;
  for (uint32_t by = 0; by < job->bins_y; by++) {
    uint32_t busy = 0;
    for (uint32_t bx = 0; bx < job->bins_x; bx++) {
      const size_t bin = (size_t)by * job->bins_x + bx;
      if (start[bin + 1] == start[bin]) continue;
      const DevTile &o = dt[list[start[bin + 1] - 1]];
      const int32_t X0 = (int32_t)(bx << job->bin_w_log2), X1 = std::min<int64_t>((int64_t)X0 + bw, jd.canvas_w);
      const int32_t Y0 = (int32_t)(by << job->bin_h_log2), Y1 = std::min<int64_t>((int64_t)Y0 + bh, jd.canvas_h);
      const bool covers = o.x0 <= X0 && o.x1 >= X1 && o.y0 <= Y0 && o.y1 >= Y1;
      const bool overwrite = !jd.alpha_blend || o.opaque;
      const bool aligned = covers && o.pitch % 16u == 0 &&
                           (reinterpret_cast<uintptr_t>(o.pix) + (size_t)(X0 - o.x0 + o.sx0) * job->bpp) % 16u == 0;
      if (!(covers && overwrite && aligned && o.fmt == native)) busy++;
    }
    job->busy_bins[by + 1] = job->busy_bins[by] + busy;
  }
  int rc;
  {
    // tensor maps of the tiles (dt[k].pad <- valid) and the interval plan of strip.cu, ahead of the tile table's upload
    std::vector<uint32_t> src_words(order.size()), src_rows(order.size());
    for (size_t k = 0; k < order.size(); k++) {
      src_words[k] = job->tiles[order[k]].d.width * (job->bpp / 4);
      src_rows[k] = job->tiles[order[k]].d.height;
    }
    if ((rc = stitch_build_strip_plan(job, dt, src_words, src_rows))) return rc;
  }
  dev_free(ctx, job->d_tiles); dev_free(ctx, job->d_bin_start); dev_free(ctx, job->d_bin_tiles);
  job->d_tiles = nullptr; job->d_bin_start = nullptr; job->d_bin_tiles = nullptr;
  if ((rc = dev_alloc(ctx, (void **)&job->d_tiles, dt.size() * sizeof(DevTile)))) return rc;
  if ((rc = dev_alloc(ctx, (void **)&job->d_bin_start, start.size() * sizeof(uint32_t)))) return rc;
  if ((rc = dev_alloc(ctx, (void **)&job->d_bin_tiles, list.size() * sizeof(uint32_t)))) return rc;
  static_assert(sizeof(DevTile) % 4 == 0, "tables are copied word by word");
  if ((rc = stitch_upload_small(ctx, job->d_tiles, dt.data(), dt.size() * sizeof(DevTile)))) return rc;
  if ((rc = stitch_upload_small(ctx, job->d_bin_start, start.data(), start.size() * sizeof(uint32_t)))) return rc;
  if ((rc = stitch_upload_small(ctx, job->d_bin_tiles, list.data(), list.size() * sizeof(uint32_t)))) return rc;
  // the host vectors are pageable: the async copies above have already staged them
  CK(cudaStreamSynchronize(ctx->stream));
  job->finalized = true;
  return STITCH_OK;
}

int stitch_dev_alloc(stitch_ctx *ctx, void **p, size_t bytes) { return dev_alloc(ctx, p, bytes); }
void stitch_dev_free(stitch_ctx *ctx, void *p) { dev_free(ctx, p); }

static int ensure_buffer(stitch_ctx *ctx, uint8_t **p, size_t *have, size_t need) {
  if (need <= *have) return STITCH_OK;
  dev_free(ctx, *p);
  *p = nullptr; *have = 0;
  int rc = dev_alloc(ctx, (void **)p, need);
  if (rc) return rc;
  *have = need;
  return STITCH_OK;
}

int stitch_ensure_buffer(stitch_ctx *ctx, uint8_t **p, size_t *have, size_t need) { return ensure_buffer(ctx, p, have, need); }

// Adler-32 of A+B from adler32(A), adler32(B), len(B) (zlib's adler32_combine)
static uint32_t adler32_combine(uint32_t adler1, uint32_t adler2, uint64_t len2) {
  const uint64_t M = 65521;
  const uint64_t a1 = adler1 & 0xffffu, b1 = adler1 >> 16, a2 = adler2 & 0xffffu, b2 = adler2 >> 16;
  const uint64_t a = (a1 + a2 + M - 1) % M;
  const uint64_t b = (b1 + b2 + (len2 % M) * ((a1 + M - 1) % M)) % M;
  return (uint32_t)((b << 16) | a);
}

static uint32_t native_fmt(uint32_t out_bit_depth) { return out_bit_depth == 16 ? (uint32_t)FMT_RGBA16 : (uint32_t)FMT_RGBA8; }

static void bg_words(const stitch_job_desc &d, uint32_t &lo, uint32_t &hi) {
  memcpy(&lo, d.background, 4);
  memcpy(&hi, d.background + 4, 4);
  if (d.out_bit_depth != 16) hi = 0;
}

static StripParams strip_params(const stitch_job *job) {
  const stitch_job_desc &jd = job->desc;
  StripParams sp{};
  sp.tiles = job->d_tiles; sp.maps = job->d_maps; sp.ivs = reinterpret_cast<const StripIv *>(job->d_ivs);
  sp.col_start = reinterpret_cast<const uint32_t *>(job->d_col_start); sp.lists = reinterpret_cast<const uint32_t *>(job->d_lists);
  sp.bg_row = reinterpret_cast<const uint8_t *>(job->d_bgrow);
  sp.canvas_w = jd.canvas_w; sp.ncol = job->strip_ncol; sp.alpha_blend = jd.alpha_blend;
  bg_words(jd, sp.bg_lo, sp.bg_hi);
  return sp;
}

// Rows [a, b) of a PNG-filtered or raw band into d_rows (row pitch job->out_pitch): decode the tiles the rows
// need, composite, filter.  Everything is queued on ctx->stream; nothing is waited for.
static int render_rows(stitch_job *job, uint32_t a, uint32_t b, uint8_t *d_rows, uint32_t *launches, cudaEvent_t ev_decoded) {
  stitch_ctx *ctx = job->ctx;
  const stitch_job_desc &jd = job->desc;
  const bool png = jd.out_kind == STITCH_OUT_PNG_FILTERED || jd.out_kind == STITCH_OUT_PNG_DEFLATE;
  const size_t row_bytes = (size_t)jd.canvas_w * job->bpp;
  const size_t pitch16 = round_up(row_bytes, 16);
  const uint32_t n_rows = b - a;
  int rc;
  if ((rc = decode_tiles(job, 0, launches, a ? a - 1 : 0, b))) return rc;  // +1 row: the PNG predictor row above
  if (ev_decoded) CK(cudaEventRecord(ev_decoded, ctx->stream));
  if (!job->finalized && (rc = finalize_job(job))) return rc;

  ComposeParams cp{};
  cp.tiles = job->d_tiles; cp.bin_start = job->d_bin_start; cp.bin_tiles = job->d_bin_tiles;
  cp.bins_x = job->bins_x; cp.bin_w_log2 = job->bin_w_log2; cp.bin_h_log2 = job->bin_h_log2;
  cp.bin_y_origin = job->bin_y_origin; cp.canvas_w = jd.canvas_w; cp.alpha_blend = jd.alpha_blend;
  bg_words(jd, cp.bg_lo, cp.bg_hi);
  if (!png) {
    cp.y_begin = a; cp.y_end = b; cp.out = d_rows; cp.out_pitch = pitch16;
    CK(launch_compose(cp, jd.out_bit_depth == 16, ctx->stream));
    *launches += 1;
    return STITCH_OK;
  }
  if (!job->force_two_pass) {
    // fused single pass: rowsum[n_rows][8] (64-bit) | ticket
    const uint32_t bh = fused_unit_rows();
    const uint32_t band0 = a / bh, n_bands = (b - 1) / bh - band0 + 1;
    const size_t words = (size_t)n_rows * 16 + 4;
    if (words > job->sync_words) {
      dev_free(ctx, job->d_sync); job->d_sync = nullptr; job->sync_words = 0;
      if ((rc = dev_alloc(ctx, (void **)&job->d_sync, words * 4))) return rc;
      job->sync_words = words;
    }
    if ((rc = stitch_fill_words(ctx, job->d_sync, 0u, words * 4))) return rc;
    FusedParams fp{};
    fp.tiles = job->d_tiles; fp.bin_start = job->d_bin_start; fp.bin_tiles = job->d_bin_tiles;
    fp.bins_x = job->bins_x; fp.canvas_w = jd.canvas_w; fp.y_begin = a; fp.y_end = b; fp.band0 = band0;
    fp.n_bands = n_bands;
    fp.nseg = (uint32_t)((row_bytes / 4 + fused_unit_words() - 1) / fused_unit_words()); fp.alpha_blend = jd.alpha_blend;
    bg_words(jd, fp.bg_lo, fp.bg_hi);
    fp.rowsum = reinterpret_cast<unsigned long long *>(job->d_sync); fp.ticket = job->d_sync + (size_t)n_rows * 16;
    fp.out = d_rows; fp.out_pitch = job->out_pitch; fp.lead = (uint32_t)job->out_lead;
    fp.maps = job->strip_ok ? job->d_maps : nullptr;
    // STITCH_STRIP=1 selects the row-pipelined kernel (strip.cu: parity-green, slower than the unit kernel on
    // B200 -- DESIGN.md section 3d); read at every call so that tests can run both in one process
    const char *strip_env = getenv("STITCH_STRIP");
    if (job->strip_ok && strip_env && strip_env[0] == '1') {
      StripParams sp = strip_params(job);
      sp.y_begin = a; sp.y_end = b;
      sp.rowsum = reinterpret_cast<unsigned long long *>(job->d_sync);
      sp.out = d_rows; sp.out_pitch = job->out_pitch; sp.lead = (uint32_t)job->out_lead;
      static const int q_env = getenv("STITCH_STRIP_Q") ? atoi(getenv("STITCH_STRIP_Q")) : 0;   // tuning
      const cudaError_t se = launch_strip(sp, jd.out_bit_depth == 16, q_env, ctx->device, ctx->sm_count, ctx->stream);
      if (se == cudaSuccess) { *launches += 1; return STITCH_OK; }
      if (se != cudaErrorInvalidConfiguration) CK(se);
      cudaGetLastError();
    }
    if ((rc = ensure_buffer(ctx, &job->d_plans, &job->plans_bytes, fused_plan_bytes(fp.n_bands * fp.nseg)))) return rc;
    fp.plans = job->d_plans;
    // How many of the band's bins need the generic compositor?  Few: one fused pass straight from the tiles.
    // Many: composite the band (plus the predictor row above it) into HBM first and give the fused kernel that
    // canvas as its only tile -- 8 more bytes of HBM traffic per pixel, but the expensive, uneven part of the
    // work no longer sits inside the kernel whose warps wait for one another.
    const uint32_t by0 = a >> job->bin_h_log2, by1 = (b - 1) >> job->bin_h_log2;
    const uint64_t busy = job->busy_bins.empty() ? 0 : job->busy_bins[by1 + 1] - job->busy_bins[by0];
    bool via_canvas = busy * 8 > (uint64_t)(by1 - by0 + 1) * job->bins_x;
    uint32_t extra = 0;
    // With an interval plan the band needs no canvas: the unit kernel in plan mode takes plain windows of one tile
    // straight from the tiles (TMA box or cp.async), fills background itself, and reads only the units that need
    // normalising or compositing from scratch a prepare kernel wrote (fused.cu).  Measured on C3: 0.43 ms
    // against 0.62 ms per 4096-row band without blending, 0.96 against 0.84 ms with zIndex blending (the prepare
    // kernel repeats each unit's predictor row and is bound by the blend arithmetic) -- so blended bands still go
    // through the canvas.  STITCH_SINGLE_PASS=0 / 1 forces the canvas / the plan mode for busy bands, 2 the plan
    // mode for every band (tests, measurements); read at every call.
    const char *sp_env = getenv("STITCH_SINGLE_PASS");
    const char sp = sp_env ? sp_env[0] : '\0';
    const bool plan_mode = job->strip_ok && (sp == '2' || (via_canvas && (sp == '1' || (sp != '0' && !jd.alpha_blend))));
    if (plan_mode) {
      fp.ivs = reinterpret_cast<const StripIv *>(job->d_ivs); fp.col_start = reinterpret_cast<const uint32_t *>(job->d_col_start);
      fp.lists = reinterpret_cast<const uint32_t *>(job->d_lists);
      // the units' prepared rows share the band canvas's buffer (the two routes never run together)
      const uint32_t n_units = fp.n_bands * fp.nseg;
      if ((rc = ensure_buffer(ctx, &job->d_canvas, &job->canvas_bytes, fused_unit_rows_bytes(n_units)))) return rc;
      if ((rc = ensure_buffer(ctx, &job->d_unit_left, &job->unit_left_bytes, fused_unit_left_bytes(n_units)))) return rc;
      fp.unit_rows = reinterpret_cast<uint32_t *>(job->d_canvas); fp.unit_left = reinterpret_cast<uint32_t *>(job->d_unit_left);
      fp.prep_list = fp.unit_left + (size_t)n_units * 2u * (fused_unit_rows() + 1u);   // behind the left pixels
      fp.prep_count = fp.ticket + 1;   // zeroed with the row sums
      via_canvas = false; extra = 1;
    }
    if (via_canvas) {
      const uint32_t ys = a ? a - 1 : 0;
      job->canvas_pitch = pitch16;
      if ((rc = ensure_buffer(ctx, &job->d_canvas, &job->canvas_bytes, pitch16 * (size_t)(b - ys)))) return rc;
      if ((rc = ensure_buffer(ctx, &job->d_ctab, &job->ctab_bytes, 1024))) return rc;
      if (job->strip_ok) CK(launch_compose_plan(strip_params(job), jd.out_bit_depth == 16, job->d_canvas, pitch16, ys, b, ctx->stream));
      else CK(launch_compose_canvas(fp, jd.out_bit_depth == 16, job->d_canvas, pitch16, ys, b, ctx->stream));
      struct { uint8_t maps[384]; DevTile t; uint32_t pad[4]; uint32_t bin_start[2]; uint32_t bin_tiles[2]; } tab;   // maps first: 128-byte aligned
      memset(&tab, 0, sizeof tab);
      tab.t.pix = job->d_canvas; tab.t.pitch = (uint32_t)pitch16; tab.t.fmt = native_fmt(jd.out_bit_depth);
      tab.t.x0 = 0; tab.t.y0 = (int32_t)ys; tab.t.x1 = (int32_t)jd.canvas_w; tab.t.y1 = (int32_t)b;
      tab.t.opaque = 1;
      tab.t.pad = stitch_encode_tile_maps(tab.t, (uint32_t)(pitch16 / 4), b - ys, tab.maps) ? 1u : 0u;   // the canvas as a TMA source
      tab.bin_start[1] = 1;
      static_assert(sizeof tab % 4 == 0, "tables are copied word by word");
      if ((rc = stitch_upload_small(ctx, job->d_ctab, &tab, sizeof tab))) return rc;  // staged on return
      fp.tiles = reinterpret_cast<const DevTile *>(job->d_ctab + offsetof(decltype(tab), t));
      fp.maps = job->d_ctab;
      fp.bin_start = reinterpret_cast<const uint32_t *>(job->d_ctab + offsetof(decltype(tab), bin_start));
      fp.bin_tiles = reinterpret_cast<const uint32_t *>(job->d_ctab + offsetof(decltype(tab), bin_tiles));
      fp.single_bin = 1; fp.alpha_blend = 0;
      extra = 1;
    }
    cudaError_t fe = launch_fused(fp, jd.out_bit_depth == 16, ctx->sm_count, ctx->stream);
    if (fe == cudaSuccess) { *launches += 2 + extra; return STITCH_OK; }
    if (fe != cudaErrorInvalidConfiguration) CK(fe);
    cudaGetLastError();
  }
  // rows too wide for the fused kernel: composite to HBM, then filter-select
  const uint32_t ys = a ? a - 1 : 0;  // 1-row halo: the predictor row is recomputed, not exchanged
  job->canvas_pitch = pitch16;
  if ((rc = ensure_buffer(ctx, &job->d_canvas, &job->canvas_bytes, pitch16 * (size_t)(b - ys)))) return rc;
  cp.y_begin = ys; cp.y_end = b; cp.out = job->d_canvas; cp.out_pitch = pitch16;
  CK(launch_compose(cp, jd.out_bit_depth == 16, ctx->stream));
  FilterParams fp{};
  fp.rows = job->d_canvas + (size_t)(a - ys) * pitch16; fp.in_pitch = pitch16;
  fp.prev0 = a ? job->d_canvas : nullptr;
  fp.n_rows = n_rows; fp.row_bytes = (uint32_t)row_bytes; fp.bpp = job->bpp;
  fp.out = d_rows; fp.out_pitch = job->out_pitch; fp.lead = (uint32_t)job->out_lead;
  CK(launch_filter_rows(fp, ctx->stream));
  *launches += 2;
  return STITCH_OK;
}

// Device -> host on the download stream, in pieces the size of the job's tile uploads.  The link is shared copy
// by copy: while uploads of one tile each are still in flight, downloads of the same size get their half of it,
// larger ones crowd the uploads out and smaller ones are crowded out themselves (tools/bus_bidir.py).
static int download_pieces(stitch_job *job, uint8_t *host, const uint8_t *dev, size_t bytes) {
  stitch_ctx *ctx = job->ctx;
  size_t piece = bytes, up_bytes = 0, up_n = 0;
  for (const auto &t : job->tiles) if (t.up_event) { up_bytes += t.d_z ? t.z_len : t.src_len; up_n++; }
  if (up_n) piece = std::max<size_t>(up_bytes / up_n, (size_t)1 << 20);
  if (const char *e = getenv("STITCH_DOWN_PIECE")) piece = std::max<size_t>(4096, strtoull(e, nullptr, 10));  // tuning
  for (size_t o = 0; o < bytes; o += piece)
    CK(cudaMemcpyAsync(host + o, dev + o, std::min(piece, bytes - o), cudaMemcpyDeviceToHost, ctx->down));
  return STITCH_OK;
}

static int run_band_impl(stitch_job *job, uint32_t y0, uint32_t y1, uint8_t *out, size_t cap, size_t *out_len, stitch_timings *tm);

extern "C" int stitch_job_run_band(stitch_job *job, uint32_t y0, uint32_t y1, uint8_t *out, size_t cap,
                                   size_t *out_len, stitch_timings *tm) {
  if (!job) return STITCH_E_INVALID;
  const int rc = run_band_impl(job, y0, y1, out, cap, out_len, tm);
  if (rc != STITCH_OK) {
    // An error exit may leave earlier chunks' downloads in flight towards the caller's buffer, kernels queued and
    // staged tables unread: drain all three streams before the caller may free anything.
    stitch_ctx *ctx = job->ctx;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->down);
    cudaStreamSynchronize(ctx->stream);
    cudaGetLastError();
    stitch_arena_reset(ctx);
  }
  return rc;
}

static int run_band_impl(stitch_job *job, uint32_t y0, uint32_t y1, uint8_t *out, size_t cap, size_t *out_len, stitch_timings *tm) {
  stitch_ctx *ctx = job->ctx;
  const stitch_job_desc &jd = job->desc;
  cudaSetDevice(ctx->device);
  if (ctx->arena_used) { cudaStreamSynchronize(ctx->stream); stitch_arena_reset(ctx); }  // an earlier call left early
  ctx->aux_launches = 0;
  if (out_len) *out_len = 0;
  if (y0 >= y1 || y0 < jd.band_y0 || y1 > jd.band_y1)
    return fail(ctx, STITCH_E_INVALID, "rows [%u,%u) are outside the job's band [%u,%u)", y0, y1, jd.band_y0, jd.band_y1);
  const bool jpeg = jd.out_kind == STITCH_OUT_JPEG_ENTROPY;
  if (jpeg && ((y0 & 7u) || ((y1 & 7u) && y1 != jd.canvas_h)))
    return fail(ctx, STITCH_E_INVALID, "JPEG bands must start on a multiple of 8 rows and end on one or at the canvas bottom");
  const size_t row_bytes = (size_t)jd.canvas_w * job->bpp;
  const uint32_t n_rows = y1 - y0;
  const bool deflate = jd.out_kind == STITCH_OUT_PNG_DEFLATE;
  const bool png = jd.out_kind == STITCH_OUT_PNG_FILTERED || deflate;
  const size_t host_row = row_bytes + (png ? 1 : 0);
  if (out && !jpeg && !deflate && cap < host_row * n_rows) return fail(ctx, STITCH_E_CAPACITY, "output buffer too small: need %zu bytes", host_row * n_rows);
  size_t jpeg_bytes = 0;
  const size_t pitch16 = round_up(row_bytes, 16);

  uint32_t launches = 0;
  int rc = reset_status(ctx);
  if (rc) return rc;
  CK(cudaEventRecord(ctx->ev[0], ctx->stream));
  CK(cudaEventRecord(ctx->ev[1], ctx->stream));
  // zlib tiles: a stream takes as long to inflate alone as among hundreds (a warp pair each), so every tile of
  // the JOB that is still compressed is inflated in ONE launch, ahead of the first band that needs any of them
  for (auto &t : job->tiles)
    if (t.d_z && !t.inflated) {
      if ((rc = wait_uploads(job, 0, 0xffffffffu))) return rc;
      if ((rc = inflate_tiles(job, 0, &launches))) return rc;
      break;
    }
  if (jpeg) {
    // With a host destination the band is coded in sub-bands of about 512 MB of pixels: the bytes of sub-band k go
    // down the bus (two device buffers take turns) while sub-band k+1 is coded and while later tiles still come up.
    // (tools/jpeg_d2h.py: 32768^2 with its 285 MB downloaded 11.8 ms in one piece, 7.6 ms at 512 MB, 7.9 at 256 MB or
    // 1 GB, 11.0 at 64 MB -- a sub-band costs about 70 us of host work and one synchronisation)
    uint32_t chunk = n_rows;
    if (out) {
      size_t chunk_bytes = (size_t)512 << 20;
      if (const char *e = getenv("STITCH_JPEG_CHUNK_BYTES")) chunk_bytes = std::max<size_t>(1, strtoull(e, nullptr, 10));  // tuning / tests
      const size_t want = chunk_bytes / std::max<size_t>(row_bytes, 1);
      chunk = (uint32_t)std::min<size_t>(n_rows, std::max<size_t>(8, want / 8 * 8));
    }
    bool first = true;
    uint32_t n_sub = 0;
    size_t total = 0;
    for (uint32_t a = y0; a < y1; a += chunk) {
      const uint32_t b = (uint32_t)std::min<uint64_t>((uint64_t)a + chunk, y1);
      if ((rc = decode_tiles(job, 0, &launches, a, b))) return rc;
      if (first) CK(cudaEventRecord(ctx->ev[2], ctx->stream));
      first = false;
      if (!job->finalized && (rc = finalize_job(job))) return rc;
      const uint32_t k = n_sub++ & 1u;
      if (out) {   // the coder writes into job->d_out: lend it the buffer whose turn it is
        if (job->ev_pack[k]) CK(cudaStreamWaitEvent(ctx->stream, job->ev_pack[k], 0));  // its last download is over
        std::swap(job->d_out, job->d_pack[k]);
        std::swap(job->out_bytes, job->pack_bytes[k]);
      }
      // an exact mosaic of RGBA8 tiles is coded straight from the tiles (jpeg.cu); anything else is composited
      // into HBM first, then kernel (d) reads the canvas; MCU rows are independent entropy segments
      size_t sub_bytes = 0;
      bool need_canvas = false;
      rc = stitch_jpeg_encode_band(job, a, b, nullptr, 0, &sub_bytes, &launches, &need_canvas);
      if (!rc && need_canvas) {
        job->canvas_pitch = pitch16;
        rc = ensure_buffer(ctx, &job->d_canvas, &job->canvas_bytes, pitch16 * (size_t)(b - a));
        if (!rc) {  // the quad compositor of fused.cu (16-byte loads inside a tile) feeding the canvas
          FusedParams fq{};
          fq.tiles = job->d_tiles; fq.bin_start = job->d_bin_start; fq.bin_tiles = job->d_bin_tiles;
          fq.bins_x = job->bins_x; fq.canvas_w = jd.canvas_w; fq.alpha_blend = jd.alpha_blend;
          bg_words(jd, fq.bg_lo, fq.bg_hi);
          cudaError_t ce = job->strip_ok ? launch_compose_plan(strip_params(job), false, job->d_canvas, pitch16, a, b, ctx->stream)
                                         : launch_compose_canvas(fq, false, job->d_canvas, pitch16, a, b, ctx->stream);
          if (ce != cudaSuccess) rc = fail(ctx, STITCH_E_CUDA, "compositor launch failed: %s", cudaGetErrorString(ce));
          launches += 1;
        }
        if (!rc) rc = stitch_jpeg_encode_band(job, a, b, job->d_canvas, pitch16, &sub_bytes, &launches, nullptr);
      }
      if (out) {
        std::swap(job->d_out, job->d_pack[k]);
        std::swap(job->out_bytes, job->pack_bytes[k]);
      }
      if (rc) return rc;
      if (out) {
        const size_t need = sizeof(stitch_jpeg_band_info) + total + sub_bytes;
        if (cap < need) return fail(ctx, STITCH_E_CAPACITY, "output buffer too small: need %zu bytes", need);
        CK(cudaEventRecord(ctx->ev_chunk, ctx->stream));
        CK(cudaStreamWaitEvent(ctx->down, ctx->ev_chunk, 0));
        if (sub_bytes && (rc = download_pieces(job, out + sizeof(stitch_jpeg_band_info) + total, job->d_pack[k], sub_bytes))) return rc;
        if (!job->ev_pack[k]) CK(cudaEventCreateWithFlags(&job->ev_pack[k], cudaEventDisableTiming));
        CK(cudaEventRecord(job->ev_pack[k], ctx->down));
      }
      total += sub_bytes;
    }
    jpeg_bytes = total;
    job->out_pitch = jpeg_bytes; job->out_lead = 0; job->out_row_bytes = jpeg_bytes;   // without a host destination: the one sub-band, in d_out
    CK(cudaEventRecord(ctx->ev[3], ctx->stream));
    if (out) {
      CK(cudaEventRecord(ctx->ev_down, ctx->down));
      CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_down, 0));
      stitch_jpeg_band_info info;
      memset(&info, 0, sizeof info);
      info.n_bits = (uint64_t)jpeg_bytes * 8;  // every MCU row restarts the DC predictors: last_dc stays 0
      memcpy(out, &info, sizeof info);
      if (out_len) *out_len = sizeof info + jpeg_bytes;
    }
  } else {
    if (png) { job->out_pitch = 16 + pitch16; job->out_lead = 15; }
    else { job->out_pitch = pitch16; job->out_lead = 0; }
    job->out_row_bytes = host_row;
    if ((rc = ensure_buffer(ctx, &job->d_out, &job->out_bytes, job->out_pitch * (size_t)n_rows))) return rc;
    // With a host destination the band is produced in chunks of about 64 MB: chunk k goes down the bus on its
    // own stream while chunk k+1 is computed and while later tiles are still coming up (add_tile queued
    // them on the upload stream), so the bus is busy in both directions and the kernels hide behind it.
    uint32_t chunk = n_rows;
    if (out) {
      size_t chunk_bytes = (size_t)64 << 20;
      if (const char *e = getenv("STITCH_CHUNK_BYTES")) chunk_bytes = std::max<size_t>(1, strtoull(e, nullptr, 10));  // tuning / tests
      const size_t want = chunk_bytes / std::max<size_t>(host_row, 1);
      chunk = (uint32_t)std::min<size_t>(n_rows, std::max<size_t>(16, want / 16 * 16));
    }
    bool first = true;
    uint32_t n_chunks = 0, z_adler = 1, z_crc_raw = 0;
    size_t z_total = 0;
    // STITCH_TRACE=1: when every chunk was rendered and when it had left the device, on stderr (debugging aid)
    const bool trace = getenv("STITCH_TRACE") != nullptr;
    std::vector<cudaEvent_t> tr_render, tr_down;
    cudaEvent_t tr_up = nullptr;
    if (trace) { cudaEventCreate(&tr_up); cudaEventRecord(tr_up, ctx->up); }
    for (uint32_t a = y0; a < y1; a += chunk) {
      const uint32_t b = (uint32_t)std::min<uint64_t>((uint64_t)a + chunk, y1);
      uint8_t *d_rows = job->d_out + (size_t)(a - y0) * job->out_pitch;
      if ((rc = render_rows(job, a, b, d_rows, &launches, first ? ctx->ev[2] : nullptr))) return rc;
      first = false;
      if (trace) { cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, ctx->stream); tr_render.push_back(e); }
      if (deflate) {
        // 8(f).1: the chunk leaves the device compressed.  Two device buffers take turns, so chunk k goes
        // down the bus while chunk k+1 is rendered and compressed.
        const uint32_t k = n_chunks++ & 1u;
        uint8_t **zbuf = k ? &job->d_defl_b : &job->d_defl;
        size_t *zcap = k ? &job->defl_b_bytes : &job->defl_bytes;
        if (job->ev_defl[k]) CK(cudaStreamWaitEvent(ctx->stream, job->ev_defl[k], 0));  // its last download is over
        if ((rc = ensure_buffer(ctx, zbuf, zcap, stitch_deflate_bound(host_row * (size_t)(b - a))))) return rc;
        size_t zbytes = 0;
        uint32_t adler = 1, crc_raw = 0;
        if ((rc = stitch_deflate_band(job, d_rows, b - a, host_row, *zbuf, &zbytes, &adler, &crc_raw, &launches))) return rc;
        z_adler = adler32_combine(z_adler, adler, (uint64_t)host_row * (b - a));
        z_crc_raw = stitch_crc_raw_concat(z_crc_raw, crc_raw, zbytes);
        if (out) {
          if (cap < sizeof(stitch_deflate_band_info) + z_total + zbytes)
            return fail(ctx, STITCH_E_CAPACITY, "output buffer too small: need more than %zu bytes", sizeof(stitch_deflate_band_info) + z_total + zbytes);
          if ((rc = download_pieces(job, out + sizeof(stitch_deflate_band_info) + z_total, *zbuf, zbytes))) return rc;
          if (!job->ev_defl[k]) CK(cudaEventCreateWithFlags(&job->ev_defl[k], cudaEventDisableTiming));
          CK(cudaEventRecord(job->ev_defl[k], ctx->down));
        }
        z_total += zbytes;
      } else if (out && png) {
        // the filtered rows are packed on the device (the host rows have no padding) and leave as flat copies,
        // cut to the size of the tile uploads they share the link with (download_pieces)
        const uint32_t k = n_chunks++ & 1u;
        const size_t bytes = host_row * (size_t)(b - a);
        if (job->ev_pack[k]) CK(cudaStreamWaitEvent(ctx->stream, job->ev_pack[k], 0));  // its last download is over
        if ((rc = ensure_buffer(ctx, &job->d_pack[k], &job->pack_bytes[k], round_up(bytes, 16)))) return rc;
        const size_t n_words = (bytes + 3) / 4;
        pack_rows_kernel<<<(unsigned)std::min<size_t>((n_words + 255) / 256, 148 * 16), 256, 0, ctx->stream>>>(
            (uint32_t *)job->d_pack[k], d_rows + job->out_lead, job->out_pitch, (uint32_t)host_row, n_words, bytes);
        CK(cudaGetLastError());
        launches += 1;
        CK(cudaEventRecord(ctx->ev_chunk, ctx->stream));
        CK(cudaStreamWaitEvent(ctx->down, ctx->ev_chunk, 0));
        // The link is shared copy by copy: with uploads of one tile each still in flight, downloads of the same
        // size get their half of it, larger ones crowd the uploads out and smaller ones are crowded out themselves.
        if ((rc = download_pieces(job, out + (size_t)(a - y0) * host_row, job->d_pack[k], bytes))) return rc;
        if (!job->ev_pack[k]) CK(cudaEventCreateWithFlags(&job->ev_pack[k], cudaEventDisableTiming));
        CK(cudaEventRecord(job->ev_pack[k], ctx->down));
        if (trace) { cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, ctx->down); tr_down.push_back(e); }
      } else if (out) {
        CK(cudaEventRecord(ctx->ev_chunk, ctx->stream));
        CK(cudaStreamWaitEvent(ctx->down, ctx->ev_chunk, 0));
        CK(cudaMemcpy2DAsync(out + (size_t)(a - y0) * host_row, host_row, d_rows + job->out_lead, job->out_pitch, host_row,
                             b - a, cudaMemcpyDeviceToHost, ctx->down));
        if (trace) { cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, ctx->down); tr_down.push_back(e); }
      }
    }
    if (trace) {
      cudaStreamSynchronize(ctx->stream); cudaStreamSynchronize(ctx->down); cudaStreamSynchronize(ctx->up);
      float ms = 0;
      cudaEventElapsedTime(&ms, ctx->ev[0], tr_up);
      fprintf(stderr, "[stitch trace] uploads queued before run_band finished at %.2f ms\n", ms);
      for (size_t i = 0; i < tr_render.size(); i++) {
        float r = 0, d = 0;
        cudaEventElapsedTime(&r, ctx->ev[0], tr_render[i]);
        if (i < tr_down.size()) cudaEventElapsedTime(&d, ctx->ev[0], tr_down[i]);
        fprintf(stderr, "[stitch trace] chunk %zu rendered %.2f ms, downloaded %.2f ms\n", i, r, d);
      }
      for (auto e : tr_render) cudaEventDestroy(e);
      for (auto e : tr_down) cudaEventDestroy(e);
      cudaEventDestroy(tr_up);
    }
    if (deflate) {
      CK(cudaEventRecord(ctx->ev[3], ctx->stream));
      if (out) {
        CK(cudaEventRecord(ctx->ev_down, ctx->down));
        CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_down, 0));
        stitch_deflate_band_info info;
        memset(&info, 0, sizeof info);
        info.n_bytes = z_total; info.n_raw = (uint64_t)host_row * n_rows; info.adler = z_adler;
        info.crc32 = stitch_crc_finish(z_crc_raw, z_total);
        memcpy(out, &info, sizeof info);
        if (out_len) *out_len = sizeof info + z_total;
      }
    } else {
    CK(cudaEventRecord(ctx->ev[3], ctx->stream));
    if (out) {
      CK(cudaEventRecord(ctx->ev_down, ctx->down));
      CK(cudaStreamWaitEvent(ctx->stream, ctx->ev_down, 0));
      if (out_len) *out_len = host_row * n_rows;
    }
    }
  }
  CK(cudaEventRecord(ctx->ev[4], ctx->stream));
  CK(cudaMemcpyAsync(ctx->h_status, ctx->d_status, sizeof(DevStatus), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  stitch_arena_reset(ctx);
  job->last_y0 = y0; job->last_y1 = y1;
  if (tm) {
    // decode_ms covers the first chunk's decode only when the band is produced in several chunks
    memset(tm, 0, sizeof *tm);
    cudaEventElapsedTime(&tm->h2d_ms, ctx->ev[0], ctx->ev[1]);
    cudaEventElapsedTime(&tm->decode_ms, ctx->ev[1], ctx->ev[2]);
    cudaEventElapsedTime(&tm->compose_ms, ctx->ev[2], ctx->ev[3]);
    cudaEventElapsedTime(&tm->d2h_ms, ctx->ev[3], ctx->ev[4]);
    cudaEventElapsedTime(&tm->total_ms, ctx->ev[0], ctx->ev[4]);
    tm->launches = launches + ctx->aux_launches;
  }
  return check_status(ctx, job);
}

extern "C" int stitch_job_band_device(stitch_job *job, void **ptr, size_t *pitch, size_t *lead, size_t *row_bytes) {
  if (!job || !job->d_out) return STITCH_E_STATE;
  if (ptr) *ptr = job->d_out;
  if (pitch) *pitch = job->out_pitch;
  if (lead) *lead = job->out_lead;
  if (row_bytes) *row_bytes = job->out_row_bytes;
  return STITCH_OK;
}

int stitch_job_read_tile(stitch_job *job, uint32_t k, uint8_t *out, size_t cap) {
  if (!job || !out || k >= job->tiles.size()) return STITCH_E_INVALID;
  stitch_ctx *ctx = job->ctx;
  cudaSetDevice(ctx->device);
  TileRec &t = job->tiles[k];
  if (!t.decoded || !t.d_pix) return fail(ctx, STITCH_E_STATE, "tile #%u is not decoded", k + 1);
  const size_t rb = scanline_bytes(t.d.width, t.d.bit_depth, t.d.color_type);
  if (cap < rb * t.d.height) return fail(ctx, STITCH_E_CAPACITY, "output buffer too small: need %zu bytes", rb * t.d.height);
  if (!t.up_waited) { CK(cudaStreamWaitEvent(ctx->stream, t.up_event, 0)); t.up_waited = true; }
  CK(cudaMemcpy2DAsync(out, rb, t.d_pix, t.pitch, rb, t.d.height, cudaMemcpyDeviceToHost, ctx->stream));
  int rc = reset_status(ctx);
  if (rc) return rc;
  CK(cudaMemcpyAsync(ctx->h_status, ctx->d_status, sizeof(DevStatus), cudaMemcpyDeviceToHost, ctx->stream));
  CK(cudaStreamSynchronize(ctx->stream));
  stitch_arena_reset(ctx);
  return check_status(ctx, job);
}

extern "C" int stitch_job_status(stitch_job *job) { return job ? job->status : STITCH_E_INVALID; }

// ---------------------------------------------------------------------------
extern "C" int stitch_decode_image(stitch_ctx *ctx, const stitch_tile_desc *tile, const uint8_t *data, size_t len,
                                   uint8_t *out, size_t cap, size_t *out_len) {
  if (!ctx || !tile || !out) return STITCH_E_INVALID;
  if (out_len) *out_len = 0;
  stitch_job_desc jd{};
  jd.canvas_w = std::max(1u, tile->width); jd.canvas_h = std::max(1u, tile->height);
  jd.out_bit_depth = 8; jd.out_kind = STITCH_OUT_RAW_RGBA;
  stitch_job *job = nullptr;
  int rc = stitch_job_create(ctx, &jd, &job);
  if (rc) return rc;
  stitch_tile_desc td = *tile;
  td.dst_x = td.dst_y = 0; td.src_off_x = td.src_off_y = 0; td.vis_w = td.width; td.vis_h = td.height;
  rc = stitch_job_add_tile(job, &td, data, len);
  if (!rc) rc = stitch_job_decode(job, 0, nullptr);
  if (!rc) {
    const TileRec &t = job->tiles[0];
    const size_t rb = scanline_bytes(td.width, td.bit_depth, td.color_type);
    if (cap < rb * td.height) rc = fail(ctx, STITCH_E_CAPACITY, "output buffer too small: need %zu bytes", rb * td.height);
    else {
      cudaError_t e = cudaMemcpy2DAsync(out, rb, t.d_pix, t.pitch, rb, td.height, cudaMemcpyDeviceToHost, ctx->stream);
      if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
      if (e != cudaSuccess) rc = fail(ctx, STITCH_E_CUDA, "D2H failed: %s", cudaGetErrorString(e));
      else if (out_len) *out_len = rb * td.height;
    }
  }
  stitch_job_destroy(job);
  return rc;
}

extern "C" int stitch_filter_rows(stitch_ctx *ctx, const uint8_t *rows, const uint8_t *prev_row, uint32_t n_rows,
                                  size_t row_bytes, uint32_t bpp, uint8_t *out, size_t cap) {
  if (!ctx || !rows || !out || !n_rows || !row_bytes || !bpp) return STITCH_E_INVALID;
  if (cap < (row_bytes + 1) * n_rows) return fail(ctx, STITCH_E_CAPACITY, "output buffer too small");
  if (row_bytes >= (1ull << 25)) return fail(ctx, STITCH_E_INVALID, "rows above 32 MiB are not supported");
  cudaSetDevice(ctx->device);
  const size_t pitch = round_up(row_bytes, 16), opitch = 16 + pitch;
  uint8_t *d_in = nullptr, *d_out = nullptr;
  int rc = dev_alloc(ctx, (void **)&d_in, pitch * ((size_t)n_rows + 1));
  if (rc) return rc;
  rc = dev_alloc(ctx, (void **)&d_out, opitch * n_rows);
  if (rc) { dev_free(ctx, d_in); return rc; }
  cudaError_t e = cudaSuccess;
  if (prev_row) e = cudaMemcpyAsync(d_in, prev_row, row_bytes, cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess)
    e = cudaMemcpy2DAsync(d_in + pitch, pitch, rows, row_bytes, row_bytes, n_rows, cudaMemcpyHostToDevice, ctx->stream);
  FilterParams fp{};
  fp.rows = d_in + pitch; fp.in_pitch = pitch; fp.prev0 = prev_row ? d_in : nullptr; fp.n_rows = n_rows;
  fp.row_bytes = (uint32_t)row_bytes; fp.bpp = bpp; fp.out = d_out; fp.out_pitch = opitch; fp.lead = 15;
  if (e == cudaSuccess) e = launch_filter_rows(fp, ctx->stream);
  if (e == cudaSuccess)
    e = cudaMemcpy2DAsync(out, row_bytes + 1, d_out + 15, opitch, row_bytes + 1, n_rows, cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  dev_free(ctx, d_in); dev_free(ctx, d_out);
  if (e != cudaSuccess) return fail(ctx, STITCH_E_CUDA, "filter_rows failed: %s", cudaGetErrorString(e));
  return STITCH_OK;
}
