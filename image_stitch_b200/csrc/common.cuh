// common.cuh -- shared device structs and per-pixel device functions.
// sm_100a only.  Reference citations are relative to jburnhams/image-stitch v1.1.53.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace stitch {

// ---- source pixel formats (colour type x bit depth the reference accepts) ----
enum SrcFmt : uint32_t {
  FMT_RGBA8 = 0, FMT_RGB8, FMT_GA8, FMT_G8, FMT_G4, FMT_G2, FMT_G1,
  FMT_RGBA16, FMT_RGB16, FMT_GA16, FMT_G16, FMT_INVALID
};

__host__ __device__ inline uint32_t src_fmt(uint32_t color_type, uint32_t bit_depth) {
  switch (color_type) {
    case 6: return bit_depth == 8 ? FMT_RGBA8 : bit_depth == 16 ? FMT_RGBA16 : FMT_INVALID;
    case 2: return bit_depth == 8 ? FMT_RGB8 : bit_depth == 16 ? FMT_RGB16 : FMT_INVALID;
    case 4: return bit_depth == 8 ? FMT_GA8 : bit_depth == 16 ? FMT_GA16 : FMT_INVALID;
    case 0:
      switch (bit_depth) {
        case 8: return FMT_G8; case 16: return FMT_G16; case 4: return FMT_G4;
        case 2: return FMT_G2; case 1: return FMT_G1; default: return FMT_INVALID;
      }
    default: return FMT_INVALID;
  }
}

// One placed input as the compose kernels see it (sorted by (z_index, image_idx)).
struct DevTile {
  const uint8_t *pix;  // dense unfiltered rows in SOURCE format
  uint32_t pitch;      // bytes between rows
  uint32_t fmt;        // SrcFmt
  int32_t x0, y0;      // clipped top-left on the canvas
  int32_t x1, y1;      // exclusive bottom-right on the canvas
  int32_t sx0, sy0;    // source pixel that lands on (x0,y0): sourceOffsetX/Y
  uint32_t opaque;     // 1: every pixel has alpha == max (no alpha channel, or verified by decode)
  uint32_t pad;        // 1: the tile has valid TMA tensor maps (strip_plan.cu), 0: none
};

// One dependency chain for kernel (a): a non-interlaced image or one Adam7 pass.
struct UnfStream {
  const uint8_t *src;      // first filter byte of the chain
  uint8_t *dst;            // dense output rows
  uint32_t src_stride;     // 1 + row_bytes
  uint32_t row_bytes;
  uint32_t dst_pitch;
  uint32_t rows;
  uint32_t bpp;            // filter distance: ceil(spp*bitDepth/8), png-filter.ts:186-211
  uint32_t last_row_bytes; // bytes available for the final row (adam7.ts:87-88 slice semantics)
  uint32_t tile;           // for error reports
  uint32_t pass;           // 0 = not interlaced, 1..7 Adam7 pass
  const uint8_t *src_end;  // 16-byte granules starting at or after this address are never read
  uint32_t prog_base;      // index of this chain's first progress word (one per 32-row group)
  uint32_t pad;
};

// One work unit of the wavefront unfilter kernel: a 32-row group of one chain.
struct UnfUnit {
  uint32_t stream;
  uint32_t group;
};

// Adam7 gather parameters for one interlaced image.
struct A7Image {
  const uint8_t *pass_pix[7];
  uint32_t pass_pitch[7];
  uint32_t pass_w[7], pass_h[7];
  uint8_t *dst;
  uint32_t dst_pitch;
  uint32_t width, height;
  uint32_t bit_depth, bpp;  // bpp in bytes (>=1)
  uint32_t row_bytes;
};

// One colour type 3 image for expand_palette_kernel (8f.4 extension; the reference throws on palette input).
struct PalImage {
  const uint8_t *idx;   // dense rows of palette indices, MSB first below 8 bits
  uint8_t *dst;         // RGBA8 rows
  uint32_t idx_pitch, dst_pitch;
  uint32_t width, height;
  uint32_t bit_depth, pad;
  uint32_t pal[256];    // RGBA8 in memory order
};

// device error word: 0 = none, else (kind<<56 | tile<<24 | row/aux)
struct DevStatus {
  unsigned long long first_error;
};

// ---------------------------------------------------------------------------
// convertScanline (src/pixel-ops.ts:496-633) for one pixel.
// 16->8 uses Math.round(v*255/65535) == (v+128)/257 for all v (exhaustively
// checked in tests/test_oracle_unit.py); n-bit -> 8 is v*255/(2^n-1) exactly.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t scale16to8(uint32_t v) { return (v + 128u) / 257u; }

__device__ __forceinline__ uint32_t ld_be16(const uint8_t *p) { return ((uint32_t)p[0] << 8) | p[1]; }

// Returns the pixel as RGBA8 in memory order (byte0=R ... byte3=A).
__device__ __forceinline__ uint32_t fetch_rgba8(const DevTile &t, uint32_t sx, uint32_t sy) {
  const uint8_t *row = t.pix + (size_t)sy * t.pitch;
  switch (t.fmt) {
    case FMT_RGBA8: return *reinterpret_cast<const uint32_t *>(row + 4u * sx);
    case FMT_RGB8: {
      const uint8_t *p = row + 3u * sx;
      return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | 0xFF000000u;
    }
    case FMT_GA8: {
      uint32_t ga = *reinterpret_cast<const uint16_t *>(row + 2u * sx);
      uint32_t g = ga & 0xFFu, a = ga >> 8;
      return g * 0x010101u | (a << 24);
    }
    case FMT_G8: return (uint32_t)row[sx] * 0x010101u | 0xFF000000u;
    case FMT_G4: {
      uint32_t v = (row[sx >> 1] >> ((~sx & 1u) * 4u)) & 0xFu;
      return (v * 17u) * 0x010101u | 0xFF000000u;
    }
    case FMT_G2: {
      uint32_t v = (row[sx >> 2] >> ((~sx & 3u) * 2u)) & 0x3u;
      return (v * 85u) * 0x010101u | 0xFF000000u;
    }
    case FMT_G1: {
      uint32_t v = (row[sx >> 3] >> (~sx & 7u)) & 0x1u;
      return (v * 255u) * 0x010101u | 0xFF000000u;
    }
    case FMT_RGBA16: {
      const uint8_t *p = row + 8u * sx;
      uint2 w = *reinterpret_cast<const uint2 *>(p);  // pitch and 8*sx keep this 8-byte aligned
      uint32_t r = __byte_perm(w.x, 0, 0x4401), g = __byte_perm(w.x, 0, 0x4423);
      uint32_t b = __byte_perm(w.y, 0, 0x4401), a = __byte_perm(w.y, 0, 0x4423);
      return scale16to8(r) | (scale16to8(g) << 8) | (scale16to8(b) << 16) | (scale16to8(a) << 24);
    }
    case FMT_RGB16: {
      const uint8_t *p = row + 6u * sx;
      return scale16to8(ld_be16(p)) | (scale16to8(ld_be16(p + 2)) << 8) | (scale16to8(ld_be16(p + 4)) << 16) |
             0xFF000000u;
    }
    case FMT_GA16: {
      const uint8_t *p = row + 4u * sx;
      uint32_t g = scale16to8(ld_be16(p)), a = scale16to8(ld_be16(p + 2));
      return g * 0x010101u | (a << 24);
    }
    case FMT_G16: {
      uint32_t g = scale16to8(ld_be16(row + 2u * sx));
      return g * 0x010101u | 0xFF000000u;
    }
    default: return 0u;
  }
}

// Returns the pixel as RGBA16 in memory order: x = bytes R_hi R_lo G_hi G_lo, y = B_hi B_lo A_hi A_lo.
// 8->16 is Math.round(v*65535/255) == v*257, i.e. the byte repeated (pixel-ops.ts:312-326).
__device__ __forceinline__ uint2 fetch_rgba16(const DevTile &t, uint32_t sx, uint32_t sy) {
  const uint8_t *row = t.pix + (size_t)sy * t.pitch;
  switch (t.fmt) {
    case FMT_RGBA16: return *reinterpret_cast<const uint2 *>(row + 8u * sx);
    case FMT_RGB16: {
      const uint8_t *p = row + 6u * sx;
      uint32_t x = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
      uint32_t y = (uint32_t)p[4] | ((uint32_t)p[5] << 8) | 0xFFFF0000u;
      return make_uint2(x, y);
    }
    case FMT_GA16: {
      uint32_t w = *reinterpret_cast<const uint32_t *>(row + 4u * sx);  // g_hi g_lo a_hi a_lo
      uint32_t gg = __byte_perm(w, 0, 0x1010);
      return make_uint2(gg, __byte_perm(w, 0, 0x3210) & 0xFFFF0000u | (gg & 0xFFFFu));
    }
    case FMT_G16: {
      uint32_t g = *reinterpret_cast<const uint16_t *>(row + 2u * sx);
      uint32_t gg = g | (g << 16);
      return make_uint2(gg, (gg & 0xFFFFu) | 0xFFFF0000u);
    }
    default: {
      uint32_t p8 = fetch_rgba8(t, sx, sy);  // every <=8-bit format: sample s -> bytes [s,s]
      uint32_t x = __byte_perm(p8, 0, 0x1100);
      uint32_t y = __byte_perm(p8, 0, 0x3322);
      return make_uint2(x, y);
    }
  }
}

// ---------------------------------------------------------------------------
// compositeScanline (src/pixel-ops.ts:646-736) for one pixel, bit-exact:
// IEEE doubles, the reference's association order, explicit round-to-nearest
// intrinsics so that no FMA contraction can change a result.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t js_round_pos(double v) {  // Math.round for v >= 0
  double r = floor(v);
  return (uint32_t)r + ((__dsub_rn(v, r) >= 0.5) ? 1u : 0u);
}

// The reference's arithmetic, operation by operation (the slow, always-right path).
static __device__ __noinline__ uint32_t blend8_exact(uint32_t dst, uint32_t src) {
  uint32_t sa = src >> 24;
  double sA = __ddiv_rn((double)sa, 255.0);
  double dA = __ddiv_rn((double)(dst >> 24), 255.0);
  double om = __dsub_rn(1.0, sA);
  double oA = __dadd_rn(sA, __dmul_rn(dA, om));  // :715
  // oA >= sA >= 1/255 > 0.0001, so the :717 guard always passes here
  uint32_t out = 0;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    double s = (double)((src >> (8 * c)) & 0xFFu), d = (double)((dst >> (8 * c)) & 0xFFu);
    double v = __ddiv_rn(__dadd_rn(__dmul_rn(s, sA), __dmul_rn(__dmul_rn(d, dA), om)), oA);  // :722-723
    v = v < 255.0 ? v : 255.0;
    v = v > 0.0 ? v : 0.0;
    out |= js_round_pos(v) << (8 * c);
  }
  out |= js_round_pos(__dmul_rn(oA, 255.0)) << 24;  // :728
  return out;
}

// ---- the same results with far fewer instructions ---------------------------------------------
// (1) x / 255 and x / 65535 for integer x: q0 = x*y, r = x - b*q0 (exact in one FMA), q = q0 + r*y with
//     y = RN(1/b) is the correctly rounded quotient (Markstein); checked against IEEE division for every
//     x in tests/test_oracle_unit.py::test_division_by_constant_is_exact.
// (2) (num / oA) is only ever rounded to an integer: num * (1/oA) with a reciprocal good to ~2^-52 lands
//     within 1e-10 of the reference's quotient, so both round to the same integer unless the value sits
//     within 1e-6 of a .5 boundary -- in which case the pixel is redone by the exact path above.
// (3) Math.round(v) for 0 <= v < 2^31: n = RN-to-integer(v) by adding 2^52, then +1 on an exact .5 tie that
//     went down.
__device__ __forceinline__ double u32_to_f64(uint32_t x) {  // exact
  return __dsub_rn(__hiloint2double(0x43300000, (int)x), 4503599627370496.0);
}
__device__ __forceinline__ double div_const(double a, double b, double y) {
  const double q0 = __dmul_rn(a, y);
  return __fma_rn(__fma_rn(-b, q0, a), y, q0);
}
__device__ __forceinline__ double rcp_near(double x) {  // x in [2^-20, 2]: relative error of a few 2^-53
  double r = (double)__frcp_rn((float)x);
  double e = __fma_rn(-x, r, 1.0);
  r = __fma_rn(e, r, r);
  e = __fma_rn(-x, r, 1.0);
  return __fma_rn(e, r, r);
}
__device__ __forceinline__ uint32_t round_int(double v, double &diff) {  // nearest integer, diff = v - it
  const double t = __dadd_rn(v, 4503599627370496.0);
  diff = __dsub_rn(v, __dsub_rn(t, 4503599627370496.0));
  return (uint32_t)__double2loint(t);
}
__device__ __forceinline__ uint32_t js_round_magic(double v) {  // Math.round, exact
  double diff;
  const uint32_t n = round_int(v, diff);
  return n + (diff == 0.5 ? 1u : 0u);
}
// One pixel of compositeScanline after the trivial-alpha exits.  Returns false when a channel is too close to
// a rounding boundary to be trusted.
template <int MAXV>
__device__ __forceinline__ bool blend_fast(uint32_t sa, uint32_t da, const uint32_t (&s)[3], const uint32_t (&d)[3],
                                           uint32_t (&o)[3], uint32_t &oa) {
  constexpr double B = (double)MAXV;
  const double Y = MAXV == 255 ? 0x1.0101010101010p-8 : 0x1.0001000100010p-16;  // RN(1/B)
  const double sA = div_const(u32_to_f64(sa), B, Y), dA = div_const(u32_to_f64(da), B, Y);
  const double om = __dsub_rn(1.0, sA);
  const double dAom = __dmul_rn(dA, om);
  const double oA = __dadd_rn(sA, dAom);
  const double r = rcp_near(oA);
  bool sure = true;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const double num = __dadd_rn(__dmul_rn(u32_to_f64(s[c]), sA), __dmul_rn(__dmul_rn(u32_to_f64(d[c]), dA), om));
    double diff;
    o[c] = round_int(__dmul_rn(num, r), diff);
    sure = sure && fabs(diff) < 0.499999;
  }
  oa = js_round_magic(__dmul_rn(oA, B));
  return sure;
}

// 8-bit alpha-over in exact integer arithmetic.  With sa, da the alphas and s, d one channel, the reference's value is
// the rational v = N / D,  N = 255 s sa + d da (255 - sa),  D = 255 sa + da (255 - sa)  (both below 2^26), rounded half
// up; its f64 evaluation is within 1e-10 of v, and v is at least 1 / (2 D) > 3e-6 away from every .5 boundary unless it
// sits exactly on one -- so floor((2 N + D) / (2 D)) is the reference's byte except on exact ties, which are redone by
// the literal arithmetic (only there can the rounding errors of the f64 sequence decide).  The output alpha is
// round(D / 255), which can never tie (2 D = 255 (2 k + 1) has no solution).  Checked against blend8_exact for all
// 2^32 (sa, da, s, d) on the device: tests/test_gpu_blend_proof.py.
__device__ __forceinline__ uint32_t blend8(uint32_t dst, uint32_t src) {
  const uint32_t sa = src >> 24;
  if (sa == 255u) return src;  // srcAlpha >= 0.9999 <=> 255   (:707-711)
  if (sa == 0u) return dst;    // srcAlpha > 0.0001 fails only for 0 (:712, :732)
  const uint32_t da = dst >> 24;
  if (da == 255u) {
    // over an opaque pixel N / D = X / 255 with X = s sa + d (255 - sa): round half up = floor((2 X + 255) / 510), and
    // 2 X + 255 is odd, so there is no tie; the result is opaque (D = 255^2)
    // Red and blue travel in the 16-bit halves of one word (X <= 255^2 fits), and
    // floor((2 X + 255) / 510) = floor((X + 127) / 255) = (Y + 1 + (Y >> 8)) >> 8 with Y = X + 127 <= 65152.
    const uint32_t na = 255u - sa;
    const uint32_t rb = (src & 0x00FF00FFu) * sa + (dst & 0x00FF00FFu) * na + 0x007F007Fu;
    const uint32_t g = ((src >> 8) & 0xFFu) * sa + ((dst >> 8) & 0xFFu) * na + 0x7Fu;
    const uint32_t out = 0xFF000000u | (((rb + 0x00010001u + ((rb >> 8) & 0x00FF00FFu)) >> 8) & 0x00FF00FFu) |
                         (((g + 1u + (g >> 8)) >> 8) << 8);
    return out;
  }
  const uint32_t k1 = 255u * sa, k2 = da * (255u - sa), D = k1 + k2, D2 = 2u * D;   // 255 <= D <= 65025
  // M + 1/2 < 2^32 / D2 < M + 3: floor(T M / 2^32) is floor(T / D2) or one less for every T < 2^27 (the reciprocal
  // of the exactly converted D2 is within 2^-24 of the true one, the truncation takes less than one more)
  const uint32_t M = (uint32_t)(__frcp_rn((float)D2) * 4294967296.f) - 1u;
  uint32_t out = ((D2 + 255u) / 510u) << 24;
  bool tie = false;
#pragma unroll
  for (int c = 0; c < 3; c++) {
    const uint32_t s = (src >> (8 * c)) & 0xFFu, d = (dst >> (8 * c)) & 0xFFu;
    const uint32_t T = 2u * (s * k1 + d * k2) + D;           // 2 N + D < 2^27
    uint32_t q = __umulhi(T, M);
    uint32_t rem = T - q * D2;                                // in [0, 2 D2)
    if (rem >= D2) { q++; rem -= D2; }
    tie = tie || rem == 0u;
    out |= q << (8 * c);
  }
  if (tie) return blend8_exact(dst, src);
  return out;
}

__device__ __forceinline__ uint32_t be16_of(uint32_t w, int hi_half) {  // sample from the byte image
  return hi_half ? __byte_perm(w, 0, 0x4423) : __byte_perm(w, 0, 0x4401);
}
__device__ __forceinline__ uint32_t to_be16(uint32_t v) { return ((v >> 8) & 0xFFu) | ((v & 0xFFu) << 8); }

static __device__ __noinline__ uint2 blend16_exact(uint2 dst, uint2 src) {
  uint32_t sa = be16_of(src.y, 1);
  double sA = __ddiv_rn((double)sa, 65535.0);
  double dA = __ddiv_rn((double)be16_of(dst.y, 1), 65535.0);
  double om = __dsub_rn(1.0, sA);
  double oA = __dadd_rn(sA, __dmul_rn(dA, om));
  uint32_t ch[3];
#pragma unroll
  for (int c = 0; c < 3; c++) {
    uint32_t sw = c < 2 ? src.x : src.y, dw = c < 2 ? dst.x : dst.y;
    double s = (double)be16_of(sw, c & 1), d = (double)be16_of(dw, c & 1);
    double v = __ddiv_rn(__dadd_rn(__dmul_rn(s, sA), __dmul_rn(__dmul_rn(d, dA), om)), oA);
    v = v < 65535.0 ? v : 65535.0;
    v = v > 0.0 ? v : 0.0;
    ch[c] = js_round_pos(v);
  }
  uint32_t av = js_round_pos(__dmul_rn(oA, 65535.0));
  return make_uint2(to_be16(ch[0]) | (to_be16(ch[1]) << 16), to_be16(ch[2]) | (to_be16(av) << 16));
}

__device__ __forceinline__ uint2 blend16(uint2 dst, uint2 src) {
  const uint32_t sa = be16_of(src.y, 1);
  // (sa/65535 >= 0.9999) <=> sa >= 65529 ; (sa/65535 > 0.0001) <=> sa >= 7  (:670-677)
  if (sa >= 65529u) return src;
  if (sa < 7u) return dst;
  const uint32_t s[3] = {be16_of(src.x, 0), be16_of(src.x, 1), be16_of(src.y, 0)};
  const uint32_t d[3] = {be16_of(dst.x, 0), be16_of(dst.x, 1), be16_of(dst.y, 0)};
  uint32_t o[3], oa;
  if (!blend_fast<65535>(sa, be16_of(dst.y, 1), s, d, o, oa)) return blend16_exact(dst, src);
  return make_uint2(to_be16(o[0]) | (to_be16(o[1]) << 16), to_be16(o[2]) | (to_be16(oa) << 16));
}

// ---------------------------------------------------------------------------
// PNG filter arithmetic on 4 packed bytes (src/png-filter.ts:16-25, :88-142)
// ---------------------------------------------------------------------------
__device__ __forceinline__ int paeth1(int a, int b, int c) {
  int p = a + b - c;
  int pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

// ---------------------------------------------------------------------------
// Packed-byte helpers shared by the unfilter and filter-selection kernels.
// Four bytes live in one register; the Paeth predictor is evaluated two bytes
// at a time in half2 (1024+v is exact in fp16, so are all differences).
// ---------------------------------------------------------------------------
}  // namespace stitch
#include <cuda_fp16.h>
namespace stitch {

struct H4 { __half2 lo, hi; };  // bytes 0,1 and bytes 2,3 as 1024+v
__device__ __forceinline__ H4 unpack_h4(uint32_t w) {
  H4 r;
  uint32_t a = __byte_perm(w, 0x64646464u, 0x4140), b = __byte_perm(w, 0x64646464u, 0x4342);
  r.lo = *reinterpret_cast<__half2 *>(&a);
  r.hi = *reinterpret_cast<__half2 *>(&b);
  return r;
}
// paethPredictor (png-filter.ts:16-25) on two lanes; ties a, then b, then c.
__device__ __forceinline__ __half2 paeth_h2(__half2 a, __half2 b, __half2 c) {
  const __half2 d1 = __hsub2(b, c);  // p - a
  const __half2 d2 = __hsub2(a, c);  // p - b
  const __half2 s = __hadd2(d1, d2);  // p - c
  const __half2 pb = __habs2(d2), pc = __habs2(s);
  const __half2 t1 = __hle2(pb, pc);                        // pb <= pc
  const __half2 t2 = __hle2(__habs2(d1), __hmin2(pb, pc));  // pa <= pb && pa <= pc
  const __half2 pk1 = __hfma2(t1, d1, c);                   // t1 ? b : c
  return __hfma2(t2, __hsub2(a, pk1), pk1);                 // t2 ? a : pk1
}
__device__ __forceinline__ uint32_t pack_h4(__half2 lo, __half2 hi) {  // low byte of each 1024+v
  return __byte_perm(*reinterpret_cast<uint32_t *>(&lo), *reinterpret_cast<uint32_t *>(&hi), 0x6420);
}
__device__ __forceinline__ uint32_t paeth4_h(const H4 &a, const H4 &b, const H4 &c) {
  return pack_h4(paeth_h2(a.lo, b.lo, c.lo), paeth_h2(a.hi, b.hi, c.hi));
}
// per-byte (a + b) mod 256
__device__ __forceinline__ uint32_t add4(uint32_t a, uint32_t b) {
  return ((a & 0x7f7f7f7fu) + (b & 0x7f7f7f7fu)) ^ ((a ^ b) & 0x80808080u);
}
// per-byte floor((a + b) / 2)
// (a & b) + (((a ^ b) & 0xfe..fe) >> 1): the masked xor is one LOP3, the shift-and-add one IMAD.HI (x >> 1 is the
// high word of x * 2^31), which also moves an instruction from the ALU pipe to the FMA pipe.
__device__ __forceinline__ uint32_t avg4(uint32_t a, uint32_t b) {
  uint32_t t, r;
  asm("lop3.b32 %0, %1, %2, %3, 0x28;" : "=r"(t) : "r"(a), "r"(b), "r"(0xfefefefeu));
  asm("mad.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(t), "r"(0x80000000u), "r"(a & b));
  return r;
}
// per-byte (x - p) mod 256
__device__ __forceinline__ uint32_t sub4(uint32_t x, uint32_t p) {
  return ((x | 0x80808080u) - (p & 0x7f7f7f7fu)) ^ ((x ^ ~p) & 0x80808080u);
}

}  // namespace stitch
