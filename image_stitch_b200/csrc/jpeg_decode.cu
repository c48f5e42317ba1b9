// jpeg_decode.cu -- SURVEY.md 8(f).4: baseline JPEG tiles are decoded on the device, so that they cross the bus
// compressed and enter the compositor as the RGBA8 (or gray8) rows the reference's JPEG decoder yields
// (src/decoders/jpeg-decoder.ts:284-347: whole-image decode, then rows; on Node the pixels are sharp's, i.e.
// libjpeg-turbo's default path: Huffman decoding, jpeg_idct_islow, fancy chroma upsampling, fixed-point YCbCr -> RGB).
// The arithmetic follows that published algorithm bit for bit; oracle/jpeg_decode_oracle.c is its CPU restatement,
// pinned against libjpeg-turbo 3.1.4.1 (tests/golden/jpeg_decode), and tests/test_gpu_jpeg_decode.py compares the two.
//
//   host      marker segments (SOF0/SOF1, DQT, DHT, DRI, SOS) are parsed and the restart markers of the entropy-coded
//             segment located while the file is still in host memory -- container work, like cutting a PNG into chunks
//   entropy   one thread per restart interval (DC predictors restart there, T.81 F.2.1.3): 9-bit look-ahead tables in
//             shared memory, canonical walk for longer codes; coefficients go to HBM de-zigzagged.  A file without
//             restart markers is one interval, i.e. one thread: tiles then decode side by side, not inside a tile
//   IDCT      one thread per 8x8 block: dequantise + jpeg_idct_islow, samples into the component's plane
//   colour    one thread per 4 output pixels: h2v1 / h2v2 fancy upsampling (replication when the downsampled width is
//             <= 2), context rows replicated at the image edges, YCbCr -> RGBA8; gray is copied
//
// Everything runs on the job's upload stream right behind the file's H2D copy; the tile's upload event covers it.
#include "job.h"

#include <cstring>
#include <vector>

namespace stitch {
namespace {

enum JdErr : uint32_t {
  JD_OK = 0, JD_E_SOI, JD_E_TRUNC_SOF, JD_E_NO_SOF, JD_E_PROGRESSIVE, JD_E_PRECISION, JD_E_COMPONENTS, JD_E_SAMPLING,
  JD_E_TABLE, JD_E_SCAN, JD_E_TRUNCATED, JD_E_HUFF, JD_E_RST, JD_E_COLORSPACE
};

const char *jd_message(uint32_t code) {
  switch (code) {
    case JD_E_SOI: return "Invalid JPEG: missing SOI marker";          // jpeg-decoder.ts:49
    case JD_E_TRUNC_SOF: return "Invalid JPEG: truncated SOF marker";  // :74
    case JD_E_NO_SOF: return "Invalid JPEG: no SOF marker found";      // :93
    case JD_E_PROGRESSIVE: return "Unsupported JPEG: only baseline and extended sequential Huffman frames (SOF0, SOF1) are decoded on the device";
    case JD_E_PRECISION: return "Unsupported JPEG: sample precision must be 8 bits";
    case JD_E_COMPONENTS: return "Unsupported JPEG: 1 (gray) or 3 (YCbCr) components in one interleaved scan";
    case JD_E_SAMPLING: return "Unsupported JPEG: luma sampling must be 1x1, 2x1 or 2x2 with 1x1 chroma";
    case JD_E_TABLE: return "Invalid JPEG: bad or missing quantisation or Huffman table";
    case JD_E_SCAN: return "Invalid JPEG: bad scan header";
    case JD_E_TRUNCATED: return "Invalid JPEG: entropy-coded data ends early";
    case JD_E_HUFF: return "Invalid JPEG: bad Huffman code";
    case JD_E_RST: return "Invalid JPEG: restart marker missing or out of order";
    case JD_E_COLORSPACE: return "Unsupported JPEG: 3-component files must be YCbCr";
    default: return "ok";
  }
}

constexpr int LOOK = 9;   // look-ahead bits of the Huffman tables

struct JdHuff {
  uint16_t look[1 << LOOK];   // code bits << 8 | symbol for codes of up to LOOK bits, 0 for longer ones
  int32_t maxcode[18];        // canonical decoding (T.81 F.2.2.3): largest code of a length, -1 if none
  int32_t valptr[17];
  int32_t mincode[17];
  uint8_t vals[256];
};
struct JdComp {
  uint32_t h, v;              // sampling factors
  uint32_t wb, hb;            // blocks per row / column, MCU padding included
  uint32_t dw, dh;            // downsampled size in samples
  uint32_t block0;            // the component's first block in `coefs`
  uint32_t plane_off, stride; // its sample plane in `planes`
};
struct JdImage {
  const uint8_t *ecs;         // entropy-coded segment
  uint32_t ecs_len;
  uint32_t width, height, ncomp, hmax, vmax, mcus_x, mcus_y;
  uint32_t restart;           // MCUs per restart interval (the whole scan when the file has none)
  uint32_t n_intervals;
  const uint32_t *rst_off;    // byte offset in ecs of every interval's first byte
  JdComp c[3];
  int16_t *coefs;             // 64 per block, natural order
  uint8_t *planes;
  uint8_t *out;               // dense rows: RGBA8 (3 components) or gray8
  uint32_t *status;           // first error (JdErr), 0 = none
  uint16_t q[3][64];          // per component, natural order
  JdHuff huff[6];             // component c: DC at 2 c, AC at 2 c + 1
};

__constant__ uint8_t c_natural[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                      41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                      30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
const uint8_t h_natural[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                               41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                               30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// ---- entropy decoding -----------------------------------------------------------------------------------------
// MSB-first bit reader over one restart interval (T.81 F.2.2.5; a stuffed zero follows every 0xFF data byte, B.1.1.5).
// At a marker or at the end of the data zeros are fed, as jdhuff.c does; `fed` counts them so that an interval that
// used bits which are not there is reported.
struct JdBits {
  const uint8_t *p, *end;
  unsigned long long acc;   // next bit at bit 63
  int cnt, fed;
  bool stop;
  __device__ __forceinline__ void fill() {
    while (cnt <= 48) {
      unsigned long long byte = 0;
      if (!stop && p < end) {
        byte = *p;
        if (byte == 0xffu) {
          if (p + 1 < end && p[1] == 0u) p += 2;
          else { stop = true; byte = 0; }
        } else p++;
      } else stop = true;
      if (stop) fed += 8;
      acc |= byte << (56 - cnt);
      cnt += 8;
    }
  }
  __device__ __forceinline__ uint32_t peek(int n) const { return (uint32_t)(acc >> (64 - n)); }
  __device__ __forceinline__ void skip(int n) { acc <<= n; cnt -= n; }
};

__device__ __forceinline__ int jd_decode(JdBits &b, const JdHuff &h) {
  if (b.cnt < 16) b.fill();
  const uint32_t e = h.look[b.peek(LOOK)];
  if (e) { b.skip((int)(e >> 8)); return (int)(e & 0xffu); }
  for (int l = LOOK + 1; l <= 16; l++) {
    const int32_t code = (int32_t)b.peek(l);
    if (h.maxcode[l] >= 0 && code <= h.maxcode[l] && code >= h.mincode[l]) {
      b.skip(l);
      return h.vals[h.valptr[l] + code - h.mincode[l]];
    }
  }
  return -1;
}
__device__ __forceinline__ int jd_receive_extend(JdBits &b, int s) {   // F.2.2.1 RECEIVE + EXTEND
  if (b.cnt < s) b.fill();
  const uint32_t r = b.peek(s);
  b.skip(s);
  return r < (1u << (s - 1)) ? (int)r - (int)((1u << s) - 1u) : (int)r;
}

constexpr int JH_THREADS = 64;

__global__ void __launch_bounds__(JH_THREADS) jd_huff_kernel(const JdImage *img_p) {
  __shared__ JdHuff sh[6];
  const JdImage &img = *img_p;
  {
    const uint32_t n = img.ncomp * 2u * (uint32_t)(sizeof(JdHuff) / 4);
    const uint32_t *src = reinterpret_cast<const uint32_t *>(img.huff);
    uint32_t *dst = reinterpret_cast<uint32_t *>(sh);
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= img.n_intervals) return;
  const uint32_t total = img.mcus_x * img.mcus_y;
  const uint32_t m0 = k * img.restart, m1 = min(total, m0 + img.restart);
  JdBits b;
  b.p = img.ecs + img.rst_off[k];
  b.end = img.ecs + (k + 1u < img.n_intervals ? img.rst_off[k + 1u] - 2u : img.ecs_len);   // up to the next interval's marker
  b.acc = 0; b.cnt = 0; b.fed = 0; b.stop = false;
  int pred[3] = {0, 0, 0};
  uint32_t err = 0;
  uint32_t mx = m0 % img.mcus_x, my = m0 / img.mcus_x;
  for (uint32_t m = m0; m < m1 && !err; m++) {
    for (uint32_t i = 0; i < img.ncomp && !err; i++) {
      const JdComp &c = img.c[i];
      for (uint32_t v = 0; v < c.v && !err; v++) {
        for (uint32_t hh = 0; hh < c.h; hh++) {
          int16_t *blk = img.coefs + ((size_t)c.block0 + (size_t)(my * c.v + v) * c.wb + (mx * c.h + hh)) * 64;
          int s = jd_decode(b, sh[2 * i]);
          if (s < 0 || s > 15) { err = JD_E_HUFF; break; }
          if (s) pred[i] += jd_receive_extend(b, s);
          blk[0] = (int16_t)pred[i];
          for (int kk = 1; kk < 64; kk++) {
            s = jd_decode(b, sh[2 * i + 1]);
            if (s < 0) { err = JD_E_HUFF; break; }
            const int r = s >> 4;
            s &= 15;
            if (s) {
              kk += r;
              if (kk > 63) { err = JD_E_HUFF; break; }
              blk[c_natural[kk]] = (int16_t)jd_receive_extend(b, s);
            } else {
              if (r != 15) break;
              kk += 15;
            }
          }
          if (err) break;
        }
      }
    }
    if (++mx == img.mcus_x) { mx = 0; my++; }
  }
  if (!err && b.fed > b.cnt) err = JD_E_TRUNCATED;   // the interval used bits that are not there
  if (err) atomicCAS(img.status, 0u, err);
}

// ---- jidctint.c jpeg_idct_islow -----------------------------------------------------------------------------------
#define JD_DESCALE(x, n) (((x) + (1 << ((n) - 1))) >> (n))
__device__ __forceinline__ void jd_idct_1d(int &d0, int &d1, int &d2, int &d3, int &d4, int &d5, int &d6, int &d7, const int shift) {
  int z2 = d2, z3 = d6;
  int z1 = (z2 + z3) * 4433;
  int tmp2 = z1 + z3 * (-15137);
  int tmp3 = z1 + z2 * 6270;
  int tmp0 = (d0 + d4) * 8192, tmp1 = (d0 - d4) * 8192;
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  tmp0 = d7; tmp1 = d5; tmp2 = d3; tmp3 = d1;
  z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2;
  int z4 = tmp1 + tmp3;
  const int z5 = (z3 + z4) * 9633;
  tmp0 *= 2446; tmp1 *= 16819; tmp2 *= 25172; tmp3 *= 12299;
  z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
  z3 += z5; z4 += z5;
  tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
  d0 = (tmp10 + tmp3 + (1 << (shift - 1))) >> shift; d7 = (tmp10 - tmp3 + (1 << (shift - 1))) >> shift;
  d1 = (tmp11 + tmp2 + (1 << (shift - 1))) >> shift; d6 = (tmp11 - tmp2 + (1 << (shift - 1))) >> shift;
  d2 = (tmp12 + tmp1 + (1 << (shift - 1))) >> shift; d5 = (tmp12 - tmp1 + (1 << (shift - 1))) >> shift;
  d3 = (tmp13 + tmp0 + (1 << (shift - 1))) >> shift; d4 = (tmp13 - tmp0 + (1 << (shift - 1))) >> shift;
}
__device__ __forceinline__ uint32_t jd_limit(int x) {   // range_limit[x & RANGE_MASK] of the post-IDCT table (jdmaster.c)
  x &= 1023;
  return x < 128 ? (uint32_t)(x + 128) : x < 512 ? 255u : x < 896 ? 0u : (uint32_t)(x - 896);
}

__global__ void __launch_bounds__(128) jd_idct_kernel(const JdImage *img_p, const uint32_t n_blocks) {
  const JdImage &img = *img_p;
  const uint32_t blk = blockIdx.x * blockDim.x + threadIdx.x;
  if (blk >= n_blocks) return;
  uint32_t ci = 0;
  if (img.ncomp == 3) ci = blk >= img.c[2].block0 ? 2u : blk >= img.c[1].block0 ? 1u : 0u;
  const JdComp &c = img.c[ci];
  const uint32_t b = blk - c.block0, bx = b % c.wb, by = b / c.wb;
  int d[64];
  const uint4 *src = reinterpret_cast<const uint4 *>(img.coefs + (size_t)blk * 64);
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const uint4 v = src[i];
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int j = 0; j < 4; j++) {
      d[8 * i + 2 * j] = (int)(int16_t)(w[j] & 0xffffu) * (int)img.q[ci][8 * i + 2 * j];
      d[8 * i + 2 * j + 1] = (int)(int16_t)(w[j] >> 16) * (int)img.q[ci][8 * i + 2 * j + 1];
    }
  }
#pragma unroll
  for (int col = 0; col < 8; col++)   // pass 1: columns, CONST_BITS - PASS1_BITS = 11
    jd_idct_1d(d[col], d[8 + col], d[16 + col], d[24 + col], d[32 + col], d[40 + col], d[48 + col], d[56 + col], 11);
  uint8_t *o = img.planes + c.plane_off + (size_t)by * 8 * c.stride + (size_t)bx * 8;
#pragma unroll
  for (int r = 0; r < 8; r++) {       // pass 2: rows, CONST_BITS + PASS1_BITS + 3 = 18
    jd_idct_1d(d[8 * r], d[8 * r + 1], d[8 * r + 2], d[8 * r + 3], d[8 * r + 4], d[8 * r + 5], d[8 * r + 6], d[8 * r + 7], 18);
    const uint32_t lo = jd_limit(d[8 * r]) | (jd_limit(d[8 * r + 1]) << 8) | (jd_limit(d[8 * r + 2]) << 16) | (jd_limit(d[8 * r + 3]) << 24);
    const uint32_t hi = jd_limit(d[8 * r + 4]) | (jd_limit(d[8 * r + 5]) << 8) | (jd_limit(d[8 * r + 6]) << 16) | (jd_limit(d[8 * r + 7]) << 24);
    *reinterpret_cast<uint2 *>(o + (size_t)r * c.stride) = make_uint2(lo, hi);
  }
}

// ---- jdsample.c + jdcolor.c ---------------------------------------------------------------------------------------
__device__ __forceinline__ int jd_chroma(const uint8_t *plane, const uint32_t stride, const uint32_t dw, const uint32_t dh,
                                         const uint32_t hs, const uint32_t vs, const uint32_t x, const uint32_t y) {
  if (hs == 1u) return plane[(size_t)y * stride + x];
  const uint32_t cx = x >> 1;
  if (vs == 1u) {
    const uint8_t *r = plane + (size_t)y * stride;
    if (dw <= 2u) return r[cx];                       // h2v1_upsample: replication
    if (x == 0u) return r[0];                         // h2v1_fancy_upsample
    if (x == 2u * dw - 1u) return r[dw - 1u];
    if (x & 1u) return (r[cx] * 3 + r[cx + 1u] + 2) >> 2;
    return (r[cx] * 3 + r[cx - 1u] + 1) >> 2;
  }
  const uint32_t cy = y >> 1;
  if (dw <= 2u) return plane[(size_t)cy * stride + cx];   // h2v2_upsample: replication
  // h2v2_fancy_upsample; the context rows above the first and below the last real row repeat them (jdmainct.c)
  const uint32_t ny = (y & 1u) ? (cy + 1u < dh ? cy + 1u : dh - 1u) : (cy ? cy - 1u : 0u);
  const uint8_t *r0 = plane + (size_t)cy * stride, *r1 = plane + (size_t)ny * stride;
  const int cur = r0[cx] * 3 + r1[cx];
  if (x == 0u) return (cur * 4 + 8) >> 4;
  if (x == 2u * dw - 1u) return (cur * 4 + 7) >> 4;
  if (x & 1u) return (cur * 3 + (r0[cx + 1u] * 3 + r1[cx + 1u]) + 7) >> 4;
  return (cur * 3 + (r0[cx - 1u] * 3 + r1[cx - 1u]) + 8) >> 4;
}
__device__ __forceinline__ uint32_t jd_clamp(int v) { return (uint32_t)min(max(v, 0), 255); }

__global__ void __launch_bounds__(256) jd_color_kernel(const JdImage *img_p) {
  const JdImage &img = *img_p;
  const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= img.width) return;
  const uint32_t Y = img.planes[img.c[0].plane_off + (size_t)y * img.c[0].stride + x];
  if (img.ncomp == 1u) { img.out[(size_t)y * img.width + x] = (uint8_t)Y; return; }
  const int cb = jd_chroma(img.planes + img.c[1].plane_off, img.c[1].stride, img.c[1].dw, img.c[1].dh, img.hmax, img.vmax, x, y) - 128;
  const int cr = jd_chroma(img.planes + img.c[2].plane_off, img.c[2].stride, img.c[2].dw, img.c[2].dh, img.hmax, img.vmax, x, y) - 128;
  // jdcolor.c build_ycc_rgb_table / ycc_rgb_convert (SCALEBITS 16, arithmetic shifts)
  const uint32_t r = jd_clamp((int)Y + ((91881 * cr + 32768) >> 16));
  const uint32_t g = jd_clamp((int)Y + ((-22554 * cb + 32768 - 46802 * cr) >> 16));
  const uint32_t bl = jd_clamp((int)Y + ((116130 * cb + 32768) >> 16));
  reinterpret_cast<uint32_t *>(img.out)[(size_t)y * img.width + x] = r | (g << 8) | (bl << 16) | 0xFF000000u;
}

__global__ void jd_status_out_kernel(uint32_t *h_dst, const uint32_t *d_src) { *h_dst = *d_src; }

// ---- host: the file's marker segments ---------------------------------------------------------------------------
struct HostHuff {
  bool present = false;
  uint8_t bits[17] = {};
  uint8_t vals[256] = {};
};
struct HostFrame {
  uint32_t width = 0, height = 0, ncomp = 0, hmax = 1, vmax = 1, restart = 0, mcus_x = 0, mcus_y = 0;
  struct { uint32_t id, h, v, tq, td, ta; } c[3] = {};
  uint16_t q[4][64] = {};
  bool qset[4] = {};
  HostHuff dc[4], ac[4];
  size_t ecs_off = 0, ecs_len = 0;
  std::vector<uint32_t> rst_off;   // interval starts within the segment
};

void derive(const HostHuff &h, JdHuff &o) {
  memset(&o, 0, sizeof o);
  int code = 0, k = 0;
  for (int l = 1; l <= 16; l++) {
    o.valptr[l] = k;
    o.mincode[l] = code;
    for (int i = 0; i < h.bits[l]; i++, k++, code++) {
      if (l <= LOOK && code < (1 << l)) {
        const int first = code << (LOOK - l);
        for (int j = 0; j < (1 << (LOOK - l)); j++) o.look[first + j] = (uint16_t)((l << 8) | h.vals[k]);
      }
    }
    o.maxcode[l] = h.bits[l] ? code - 1 : -1;
    code <<= 1;
  }
  o.maxcode[17] = 0x7fffffff;
  memcpy(o.vals, h.vals, 256);
}

// Everything up to the end of the entropy-coded segment.  `header_only`: stop at the frame header (the reference's
// parseJpegHeader, jpeg-decoder.ts:46-94: dimensions and channel count, with its three messages).
uint32_t parse(const uint8_t *d, size_t n, HostFrame &f, bool header_only) {
  if (n < 2 || d[0] != 0xff || d[1] != 0xd8) return JD_E_SOI;
  size_t o = 2;
  bool have_sof = false, jfif = false;
  int adobe = -1;
  while (o + 1 < n) {
    if (d[o] != 0xff) { o++; continue; }
    const uint8_t m = d[o + 1];
    o += 2;
    if (m == 0xff) { o--; continue; }
    if (m == 0x00 || m == 0x01 || (m >= 0xd0 && m <= 0xd8)) continue;
    if (o + 2 > n) break;
    const size_t len = ((size_t)d[o] << 8) | d[o + 1];
    const bool sof = m >= 0xc0 && m <= 0xcf && m != 0xc4 && m != 0xc8 && m != 0xcc;
    if (len < 2 || o + len > n) return sof ? JD_E_TRUNC_SOF : JD_E_TRUNCATED;
    const uint8_t *p = d + o + 2;
    const size_t pl = len - 2;
    if (sof && header_only) {   // the reference's scan takes every frame type and checks nothing else (:70-82)
      if (pl < 6) return JD_E_TRUNC_SOF;
      f.height = ((uint32_t)p[1] << 8) | p[2];
      f.width = ((uint32_t)p[3] << 8) | p[4];
      f.ncomp = p[5];
      return JD_OK;
    }
    if (m == 0xc0 || m == 0xc1) {
      if (pl < 6) return JD_E_TRUNC_SOF;
      if (p[0] != 8) return JD_E_PRECISION;
      f.height = ((uint32_t)p[1] << 8) | p[2];
      f.width = ((uint32_t)p[3] << 8) | p[4];
      f.ncomp = p[5];
      if (f.ncomp != 1 && f.ncomp != 3) return JD_E_COMPONENTS;
      if (pl < 6 + 3 * (size_t)f.ncomp || !f.width || !f.height) return JD_E_TRUNC_SOF;
      for (uint32_t i = 0; i < f.ncomp; i++) {
        f.c[i].id = p[6 + 3 * i]; f.c[i].h = p[7 + 3 * i] >> 4; f.c[i].v = p[7 + 3 * i] & 15; f.c[i].tq = p[8 + 3 * i];
        if (f.c[i].tq > 3) return JD_E_TABLE;
      }
      have_sof = true;
    } else if (sof) {
      return JD_E_PROGRESSIVE;
    } else if (m == 0xc4) {
      size_t q = 0;
      while (q + 17 <= pl) {
        const int tc = p[q] >> 4, th = p[q] & 15;
        if (tc > 1 || th > 3) return JD_E_TABLE;
        HostHuff &h = tc ? f.ac[th] : f.dc[th];
        h = HostHuff();
        int total = 0;
        for (int l = 1; l <= 16; l++) { h.bits[l] = p[q + l]; total += p[q + l]; }
        if (total > 256 || q + 17 + (size_t)total > pl) return JD_E_TABLE;
        memcpy(h.vals, p + q + 17, (size_t)total);
        h.present = true;
        q += 17 + (size_t)total;
      }
    } else if (m == 0xdb) {
      size_t q = 0;
      while (q < pl) {
        const int pq = p[q] >> 4, tq = p[q] & 15;
        if (tq > 3 || pq > 1 || q + 1 + (pq ? 128u : 64u) > pl) return JD_E_TABLE;
        for (int k = 0; k < 64; k++)
          f.q[tq][h_natural[k]] = pq ? (uint16_t)((p[q + 1 + 2 * k] << 8) | p[q + 2 + 2 * k]) : p[q + 1 + k];
        f.qset[tq] = true;
        q += 1 + (pq ? 128u : 64u);
      }
    } else if (m == 0xdd) {
      if (pl < 2) return JD_E_TRUNCATED;
      f.restart = ((uint32_t)p[0] << 8) | p[1];
    } else if (m == 0xe0) {
      if (pl >= 5 && !memcmp(p, "JFIF", 5)) jfif = true;
    } else if (m == 0xee) {
      if (pl >= 12 && !memcmp(p, "Adobe", 5)) adobe = p[11];
    } else if (m == 0xda) {
      if (!have_sof) return JD_E_NO_SOF;
      if (pl < 1 || p[0] != f.ncomp) return JD_E_COMPONENTS;
      if (pl < 1 + 2 * (size_t)f.ncomp + 3) return JD_E_SCAN;
      for (uint32_t i = 0; i < f.ncomp; i++) {
        if (p[1 + 2 * i] != f.c[i].id) return JD_E_SCAN;
        f.c[i].td = p[2 + 2 * i] >> 4; f.c[i].ta = p[2 + 2 * i] & 15;
        if (f.c[i].td > 3 || f.c[i].ta > 3 || !f.dc[f.c[i].td].present || !f.ac[f.c[i].ta].present || !f.qset[f.c[i].tq]) return JD_E_TABLE;
      }
      const uint8_t *s = p + 1 + 2 * f.ncomp;
      if (s[0] != 0 || s[1] != 63 || s[2] != 0) return JD_E_SCAN;
      if (f.ncomp == 3 && !jfif) {   // the colour space the way jdapimin.c guesses it
        if (adobe == 0) return JD_E_COLORSPACE;
        if (adobe < 0 && f.c[0].id == 'R' && f.c[1].id == 'G' && f.c[2].id == 'B') return JD_E_COLORSPACE;
      }
      if (f.ncomp == 1) f.c[0].h = f.c[0].v = 1;   // a single-component scan is never interleaved
      else {
        if (f.c[1].h != 1 || f.c[1].v != 1 || f.c[2].h != 1 || f.c[2].v != 1) return JD_E_SAMPLING;
        if (!((f.c[0].h == 1 && f.c[0].v == 1) || (f.c[0].h == 2 && f.c[0].v == 1) || (f.c[0].h == 2 && f.c[0].v == 2))) return JD_E_SAMPLING;
      }
      f.hmax = f.c[0].h; f.vmax = f.c[0].v;
      f.mcus_x = (f.width + 8 * f.hmax - 1) / (8 * f.hmax);
      f.mcus_y = (f.height + 8 * f.vmax - 1) / (8 * f.vmax);
      f.ecs_off = o + len;
      // the segment ends at the first marker that is neither a stuffed zero nor RSTn; RSTn start the intervals
      const uint32_t total = f.mcus_x * f.mcus_y;
      const uint32_t per = f.restart ? f.restart : total;
      const uint32_t n_int = (total + per - 1) / per;
      f.rst_off.assign(1, 0u);
      size_t e = f.ecs_off;
      while (e < n) {
        const uint8_t *ff = static_cast<const uint8_t *>(memchr(d + e, 0xff, n - e));
        if (!ff) { e = n; break; }
        e = (size_t)(ff - d);
        if (e + 1 >= n) { e = n; break; }
        const uint8_t mk = d[e + 1];
        if (mk == 0x00) { e += 2; continue; }
        if (mk == 0xff) { e += 1; continue; }
        if (mk >= 0xd0 && mk <= 0xd7) {
          if (f.rst_off.size() >= n_int || mk != 0xd0 + ((f.rst_off.size() - 1) & 7)) return JD_E_RST;
          f.rst_off.push_back((uint32_t)(e + 2 - f.ecs_off));
          e += 2;
          continue;
        }
        break;   // EOI or another segment
      }
      f.ecs_len = e - f.ecs_off;
      if (f.rst_off.size() != n_int) return f.restart && e < n ? JD_E_RST : JD_E_TRUNCATED;
      if (f.ecs_len >= (1ull << 32)) return JD_E_TRUNCATED;
      f.restart = per;
      return JD_OK;
    }
    o += len;
  }
  return have_sof ? JD_E_TRUNCATED : JD_E_NO_SOF;
}

}  // namespace
}  // namespace stitch

using namespace stitch;

const char *stitch_jpeg_decode_message(uint32_t code) { return jd_message(code); }

// jpeg-decoder.ts:46-94 (parseJpegHeader) and :298-315 (getHeader: 1 channel for gray, else 4).
extern "C" int stitch_jpeg_info(stitch_ctx *ctx, const uint8_t *data, size_t len, uint32_t *width, uint32_t *height, uint32_t *channels) {
  if (!data || !width || !height || !channels) return STITCH_E_INVALID;
  HostFrame f;
  const uint32_t e = parse(data, len, f, true);
  if (e) return ctx ? stitch_fail(ctx, STITCH_E_JPEG, "%s", jd_message(e)) : STITCH_E_JPEG;
  *width = f.width; *height = f.height; *channels = f.ncomp == 1 ? 1u : 4u;
  return STITCH_OK;
}

extern "C" int stitch_job_add_tile_jpeg(stitch_job *job, const stitch_tile_desc *tile, const uint8_t *data, size_t len) {
  if (!job || !tile || !data) return STITCH_E_INVALID;
  stitch_ctx *ctx = job->ctx;
  cudaSetDevice(ctx->device);
  HostFrame f;
  const uint32_t pe = parse(data, len, f, false);
  if (pe) return stitch_fail(ctx, STITCH_E_JPEG, "%s", jd_message(pe));
  if ((tile->width && tile->width != f.width) || (tile->height && tile->height != f.height))
    return stitch_fail(ctx, STITCH_E_INVALID, "tile #%u: the descriptor says %ux%u, the JPEG frame %ux%u", tile->image_idx + 1,
                       tile->width, tile->height, f.width, f.height);
  // the image's tables
  std::vector<uint8_t> host(sizeof(JdImage));
  JdImage &img = *reinterpret_cast<JdImage *>(host.data());
  memset(&img, 0, sizeof img);
  img.width = f.width; img.height = f.height; img.ncomp = f.ncomp; img.hmax = f.hmax; img.vmax = f.vmax;
  img.mcus_x = f.mcus_x; img.mcus_y = f.mcus_y; img.restart = f.restart; img.n_intervals = (uint32_t)f.rst_off.size();
  img.ecs_len = (uint32_t)f.ecs_len;
  size_t n_blocks = 0, plane_bytes = 0;
  for (uint32_t i = 0; i < f.ncomp; i++) {
    JdComp &c = img.c[i];
    c.h = f.c[i].h; c.v = f.c[i].v;
    c.wb = f.mcus_x * c.h; c.hb = f.mcus_y * c.v;
    c.dw = (f.width * c.h + f.hmax - 1) / f.hmax; c.dh = (f.height * c.v + f.vmax - 1) / f.vmax;
    c.block0 = (uint32_t)n_blocks; c.plane_off = (uint32_t)plane_bytes; c.stride = c.wb * 8;
    n_blocks += (size_t)c.wb * c.hb;
    plane_bytes += (size_t)c.stride * c.hb * 8;
    memcpy(img.q[i], f.q[f.c[i].tq], sizeof img.q[i]);
    derive(f.dc[f.c[i].td], img.huff[2 * i]);
    derive(f.ac[f.c[i].ta], img.huff[2 * i + 1]);
  }
  if (n_blocks >= (1u << 26) || plane_bytes >= (1ull << 32)) return stitch_fail(ctx, STITCH_E_INVALID, "JPEG tile too large");
  const uint32_t ch = f.ncomp == 1 ? 1u : 4u;
  const size_t out_bytes = (size_t)f.width * f.height * ch;
  // status slot of this tile (pinned, written by the last kernel below)
  if (!job->h_jdstat) {
    STITCH_CK(cudaHostAlloc((void **)&job->h_jdstat, 4096 * sizeof(uint32_t), cudaHostAllocDefault));
    job->jdstat_cap = 4096;
  }
  if (job->jdstat_used >= job->jdstat_cap) return stitch_fail(ctx, STITCH_E_INVALID, "more than %zu JPEG tiles in one job", job->jdstat_cap);
  const uint32_t slot = (uint32_t)job->jdstat_used++;
  job->h_jdstat[slot] = 0;

  cudaStream_t st = ctx->up;
  uint8_t *d_out = nullptr, *d_ecs = nullptr, *d_planes = nullptr;
  int16_t *d_coefs = nullptr;
  uint32_t *d_rst = nullptr, *d_stat = nullptr;
  JdImage *d_img = nullptr;
  auto release = [&](bool all) {
    if (d_ecs) cudaFreeAsync(d_ecs, st);
    if (d_planes) cudaFreeAsync(d_planes, st);
    if (d_coefs) cudaFreeAsync(d_coefs, st);
    if (d_rst) cudaFreeAsync(d_rst, st);
    if (d_stat) cudaFreeAsync(d_stat, st);
    if (d_img) cudaFreeAsync(d_img, st);
    if (all && d_out) cudaFreeAsync(d_out, st);
  };
  cudaError_t e = cudaMallocAsync((void **)&d_out, out_bytes + 32, st);
  if (e == cudaSuccess) e = cudaMallocAsync((void **)&d_ecs, f.ecs_len + 16, st);
  if (e == cudaSuccess) e = cudaMallocAsync((void **)&d_planes, plane_bytes, st);
  if (e == cudaSuccess) e = cudaMallocAsync((void **)&d_coefs, n_blocks * 128, st);
  if (e == cudaSuccess) e = cudaMallocAsync((void **)&d_rst, f.rst_off.size() * 4, st);
  if (e == cudaSuccess) e = cudaMallocAsync((void **)&d_stat, 4, st);
  if (e == cudaSuccess) e = cudaMallocAsync((void **)&d_img, sizeof(JdImage), st);
  if (e == cudaSuccess) {
    img.ecs = d_ecs; img.rst_off = d_rst; img.coefs = d_coefs; img.planes = d_planes; img.out = d_out; img.status = d_stat;
    e = cudaMemcpyAsync(d_ecs, data + f.ecs_off, f.ecs_len, cudaMemcpyHostToDevice, st);
  }
  // pageable sources: the runtime stages them before the call returns
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_rst, f.rst_off.data(), f.rst_off.size() * 4, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_img, &img, sizeof img, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_stat, 0, 4, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_coefs, 0, n_blocks * 128, st);
  if (e == cudaSuccess) {
    jd_huff_kernel<<<(img.n_intervals + JH_THREADS - 1) / JH_THREADS, JH_THREADS, 0, st>>>(d_img);
    jd_idct_kernel<<<(unsigned)((n_blocks + 127) / 128), 128, 0, st>>>(d_img, (uint32_t)n_blocks);
    jd_color_kernel<<<dim3((f.width + 255) / 256, f.height), 256, 0, st>>>(d_img);
    jd_status_out_kernel<<<1, 1, 0, st>>>(job->h_jdstat + slot, d_stat);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) {
    release(true);
    return stitch_fail(ctx, e == cudaErrorMemoryAllocation ? STITCH_E_NOMEM : STITCH_E_CUDA, "JPEG tile: %s", cudaGetErrorString(e));
  }
  release(false);
  stitch_tile_desc d = *tile;
  d.width = f.width; d.height = f.height; d.bit_depth = 8; d.color_type = f.ncomp == 1 ? 0u : 6u; d.interlace = 0; d.filtered = 0;
  const int rc = stitch_add_tile_produced(job, &d, d_out, out_bytes);
  if (rc) { cudaFreeAsync(d_out, st); return rc; }
  job->jd_pending.push_back({slot, (uint32_t)job->tiles.size() - 1u});
  return STITCH_OK;
}

// The decoder as an operator (the reference's JpegDecoder: getHeader() + scanlines(), jpeg-decoder.ts:284-347): the
// whole image, RGBA8 or gray8 rows.
extern "C" int stitch_decode_jpeg(stitch_ctx *ctx, const uint8_t *data, size_t len, uint8_t *out, size_t cap, uint32_t *width,
                                  uint32_t *height, uint32_t *channels) {
  if (!ctx || !data || !out || !width || !height || !channels) return STITCH_E_INVALID;
  int rc = stitch_jpeg_info(ctx, data, len, width, height, channels);
  if (rc) return rc;
  if (cap < (size_t)*width * *height * *channels)
    return stitch_fail(ctx, STITCH_E_CAPACITY, "output buffer too small: need %zu bytes", (size_t)*width * *height * *channels);
  stitch_job_desc jd;
  memset(&jd, 0, sizeof jd);
  jd.canvas_w = *width; jd.canvas_h = *height; jd.out_bit_depth = 8; jd.out_kind = STITCH_OUT_RAW_RGBA; jd.band_y0 = 0; jd.band_y1 = *height;
  stitch_job *job = nullptr;
  if ((rc = stitch_job_create(ctx, &jd, &job))) return rc;
  stitch_tile_desc td;
  memset(&td, 0, sizeof td);
  td.vis_w = *width; td.vis_h = *height;
  rc = stitch_job_add_tile_jpeg(job, &td, data, len);
  if (!rc) rc = stitch_job_read_tile(job, 0, out, cap);
  stitch_job_destroy(job);
  return rc;
}
