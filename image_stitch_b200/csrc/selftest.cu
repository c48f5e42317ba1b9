// selftest.cu -- device-side proof obligations of the alpha-over fast path (src/pixel-ops.ts:646-736).
// blend8 / blend16 (common.cuh) replace the reference's divisions by cheaper forms and fall back to the literal
// arithmetic near rounding boundaries; this kernel compares them with the literal forms (blend8_exact,
// blend16_exact) on the device: every (srcAlpha, dstAlpha, src, dst) of the 8-bit case -- 2^32 tuples -- and a
// strided sample of the 16-bit case with every alpha near the reference's thresholds.  Test infrastructure that
// has to live in the library because the functions under test are device code.
#include "job.h"

namespace stitch {

struct SelfTestResult {
  unsigned long long checked, mismatches;
  uint32_t first[4];   // first mismatch seen: dst, src (8-bit) or dst.x, dst.y, src.x, src.y (16-bit)
};

__device__ __forceinline__ void st_report(SelfTestResult *r, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  if (atomicAdd(&r->mismatches, 1ull) == 0ull) { r->first[0] = a; r->first[1] = b; r->first[2] = c; r->first[3] = d; }
}

// One thread per (sa, da, s): d runs over 0..255 on channel 0; channels 1 and 2 carry other (s, d) pairs of the
// same alphas, so each call of blend8 checks three tuples (channel 0 alone is exhaustive).
__global__ void __launch_bounds__(256) selftest_blend8_kernel(SelfTestResult *res, const uint32_t stride) {
  unsigned long long checked = 0;
  for (unsigned long long idx = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; idx * stride < (1ull << 24);
       idx += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t t = (uint32_t)(idx * stride);
    const uint32_t sa = t >> 16, da = (t >> 8) & 0xFFu, s = t & 0xFFu;
    const uint32_t src = s | ((s ^ 0x5Au) << 8) | ((255u - s) << 16) | (sa << 24);
    for (uint32_t d = 0; d < 256u; d++) {
      const uint32_t dst = d | (((d + 85u) & 0xFFu) << 8) | (((d * 7u + 3u) & 0xFFu) << 16) | (da << 24);
      const uint32_t fast = blend8(dst, src);
      uint32_t want;
      if (sa == 255u) want = src;          // srcAlpha >= 0.9999 (:707-711)
      else if (sa == 0u) want = dst;       // srcAlpha > 0.0001 fails (:712, :732)
      else want = blend8_exact(dst, src);
      if (fast != want) st_report(res, dst, src, fast, want);
      checked++;
    }
  }
  checked = __reduce_add_sync(0xffffffffu, (uint32_t)checked);
  if ((threadIdx.x & 31u) == 0u) atomicAdd(&res->checked, checked);
}

__device__ __forceinline__ uint32_t st_hash(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
__device__ __forceinline__ uint32_t st_be16(uint32_t v) { return ((v >> 8) & 0xFFu) | ((v & 0xFFu) << 8); }

// Sample n of the 16-bit case: alphas and samples from a hash of n; every 4th sample pins the source alpha to a
// value within +-8 of the reference's thresholds (65529: copy, 7: skip), every 16th also the destination alpha.
__global__ void __launch_bounds__(256) selftest_blend16_kernel(SelfTestResult *res, const unsigned long long n_samples) {
  unsigned long long checked = 0;
  for (unsigned long long n = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; n < n_samples; n += (unsigned long long)gridDim.x * blockDim.x) {
    const uint32_t h0 = st_hash((uint32_t)n), h1 = st_hash(h0 ^ (uint32_t)(n >> 32) ^ 0x9e3779b9u), h2 = st_hash(h1 + 0x85ebca6bu);
    uint32_t sa = h0 & 0xFFFFu, da = h0 >> 16;
    if ((n & 3ull) == 1ull) sa = ((n >> 2) & 1ull) ? 65521u + (uint32_t)((n >> 3) % 15ull) : (uint32_t)((n >> 3) % 16ull);
    if ((n & 15ull) == 5ull) da = ((n >> 4) & 1ull) ? 65535u - (uint32_t)((n >> 5) % 9ull) : (uint32_t)((n >> 5) % 9ull);
    const uint32_t s0 = h1 & 0xFFFFu, s1 = h1 >> 16, s2 = h2 & 0xFFFFu;
    const uint32_t d0 = h2 >> 16, d1 = st_hash(h2) & 0xFFFFu, d2 = st_hash(h2) >> 16;
    const uint2 src = make_uint2(st_be16(s0) | (st_be16(s1) << 16), st_be16(s2) | (st_be16(sa) << 16));
    const uint2 dst = make_uint2(st_be16(d0) | (st_be16(d1) << 16), st_be16(d2) | (st_be16(da) << 16));
    const uint2 fast = blend16(dst, src);
    uint2 want;
    if (sa >= 65529u) want = src;        // sa / 65535 >= 0.9999 (:670-677)
    else if (sa < 7u) want = dst;        // sa / 65535 > 0.0001 fails
    else want = blend16_exact(dst, src);
    if (fast.x != want.x || fast.y != want.y) st_report(res, dst.x, dst.y, src.x, src.y);
    checked++;
  }
  checked = __reduce_add_sync(0xffffffffu, (uint32_t)checked);
  if ((threadIdx.x & 31u) == 0u) atomicAdd(&res->checked, checked);
}

// Random tuples for the host to check against the ORACLE (tests only): out[i] = {dst, src, blend8(dst, src)}.
__global__ void selftest_blend8_samples_kernel(uint32_t *out, const uint32_t n, const uint32_t seed) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t dst = st_hash(i * 2u + seed), src = st_hash(i * 2u + 1u + seed * 0x9e3779b9u);
  out[3 * i] = dst; out[3 * i + 1] = src; out[3 * i + 2] = blend8(dst, src);
}

}  // namespace stitch

using namespace stitch;

extern "C" int stitch_selftest_blend(stitch_ctx *ctx, uint32_t mode, uint64_t amount, uint64_t *checked, uint64_t *mismatches,
                                     uint32_t first[4]) {
  if (!ctx || !checked || !mismatches) return STITCH_E_INVALID;
  if (mode > 1u) return stitch_fail(ctx, STITCH_E_INVALID, "stitch_selftest_blend: mode must be 0 (8-bit) or 1 (16-bit)");
  STITCH_CK(cudaSetDevice(ctx->device));
  SelfTestResult *d_res = nullptr, h{};
  int rc = stitch_dev_alloc(ctx, (void **)&d_res, sizeof(SelfTestResult));
  if (rc) return rc;
  STITCH_CK(cudaMemsetAsync(d_res, 0, sizeof(SelfTestResult), ctx->stream));
  const unsigned blocks = (unsigned)ctx->sm_count * 8u;
  if (mode == 0u) {
    const uint32_t stride = (uint32_t)(amount ? amount : 1u);   // 1 = all 2^32 tuples; k = every k-th (sa, da, s)
    selftest_blend8_kernel<<<blocks, 256, 0, ctx->stream>>>(d_res, stride);
  } else {
    selftest_blend16_kernel<<<blocks, 256, 0, ctx->stream>>>(d_res, amount ? amount : (1ull << 28));
  }
  STITCH_CK(cudaGetLastError());
  STITCH_CK(cudaMemcpyAsync(&h, d_res, sizeof h, cudaMemcpyDeviceToHost, ctx->stream));
  STITCH_CK(cudaStreamSynchronize(ctx->stream));
  stitch_dev_free(ctx, d_res);
  *checked = h.checked; *mismatches = h.mismatches;
  if (first) for (int i = 0; i < 4; i++) first[i] = h.first[i];
  return STITCH_OK;
}

extern "C" int stitch_selftest_blend8_samples(stitch_ctx *ctx, uint32_t n, uint32_t seed, uint32_t *out_triples) {
  if (!ctx || !out_triples || !n) return STITCH_E_INVALID;
  STITCH_CK(cudaSetDevice(ctx->device));
  uint32_t *d = nullptr;
  int rc = stitch_dev_alloc(ctx, (void **)&d, (size_t)n * 12);
  if (rc) return rc;
  selftest_blend8_samples_kernel<<<(n + 255u) / 256u, 256, 0, ctx->stream>>>(d, n, seed);
  STITCH_CK(cudaGetLastError());
  STITCH_CK(cudaMemcpyAsync(out_triples, d, (size_t)n * 12, cudaMemcpyDeviceToHost, ctx->stream));
  STITCH_CK(cudaStreamSynchronize(ctx->stream));
  stitch_dev_free(ctx, d);
  return STITCH_OK;
}
