// jpeg.cu -- kernel (d): RGB->YCbCr, 8x8 forward DCT, quantisation and Huffman
// coding of composited RGBA8 rows (what StreamingJpegEncoder.encode_strip does
// in the reference, src/jpeg-encoder.ts:155-166; strip feed
// src/image-concat-core.ts:881-914, :1362-1393).
//
// The reference's JPEG arithmetic lives in the un-vendored jpeg-encoder-wasm
// package, so the definition implemented here is the one DESIGN.md section 6
// fixes (and tests pin against libjpeg-turbo): baseline 4:4:4, JFIF BT.601
// fixed-point colour transform, libjpeg "islow" integer FDCT, Annex K tables
// with libjpeg quality scaling, edge replication, restart interval = one MCU
// row.  Every MCU row is therefore an independent, byte-aligned entropy
// segment: the natural unit of parallelism for CTAs, bands and GPUs.
//
//   J1 dct      thread = one 8x8 block of one component: colour transform, FDCT,
//               quantise (exact reciprocal multiply), zig-zag int16 to HBM
//   J2 lengths  CTA = one MCU row: code length of every block, block-wide scan
//               -> bit offset of every block inside its row segment
//   J3 rowscan  sizes of the row segments -> byte offsets (unstuffed stream)
//   J4 emit     thread = one block: Huffman codes written at their bit offset
//               (64-bit accumulator, atomicOr on the shared boundary words)
//   J5 ffcount  CTA = one MCU row: 0xFF bytes of its segment
//   J6 rowscan  final offsets (stuffing + RSTm)
//   J7 stuff    CTA = one MCU row: byte stuffing, RSTm marker, final bytes
#include "job.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cstring>

using namespace stitch;

namespace {

const uint8_t k_lum_q[64] = {16, 11, 10, 16, 24,  40,  51,  61,  12, 12, 14, 19, 26,  58,  60,  55,
                             14, 13, 16, 24, 40,  57,  69,  56,  14, 17, 22, 29, 51,  87,  80,  62,
                             18, 22, 37, 56, 68,  109, 103, 77,  24, 35, 55, 64, 81,  104, 113, 92,
                             49, 64, 78, 87, 103, 121, 120, 101, 72, 92, 95, 98, 112, 100, 103, 99};
const uint8_t k_chr_q[64] = {17, 18, 24, 47, 99, 99, 99, 99, 18, 21, 26, 66, 99, 99, 99, 99, 24, 26, 56, 99, 99, 99,
                             99, 99, 47, 66, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99,
                             99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99, 99};
const uint8_t k_zigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                              41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                              30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
const uint8_t k_dc_lum_bits[16] = {0, 1, 5, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 0, 0};
const uint8_t k_dc_chr_bits[16] = {0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0};
const uint8_t k_dc_vals[12] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11};
const uint8_t k_ac_lum_bits[16] = {0, 2, 1, 3, 3, 2, 4, 3, 5, 5, 4, 4, 0, 0, 1, 0x7d};
const uint8_t k_ac_lum_vals[162] = {
    0x01, 0x02, 0x03, 0x00, 0x04, 0x11, 0x05, 0x12, 0x21, 0x31, 0x41, 0x06, 0x13, 0x51, 0x61, 0x07, 0x22, 0x71,
    0x14, 0x32, 0x81, 0x91, 0xa1, 0x08, 0x23, 0x42, 0xb1, 0xc1, 0x15, 0x52, 0xd1, 0xf0, 0x24, 0x33, 0x62, 0x72,
    0x82, 0x09, 0x0a, 0x16, 0x17, 0x18, 0x19, 0x1a, 0x25, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x34, 0x35, 0x36, 0x37,
    0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58, 0x59,
    0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a, 0x83,
    0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a, 0xa2, 0xa3,
    0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba, 0xc2, 0xc3,
    0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda, 0xe1, 0xe2,
    0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf1, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};
const uint8_t k_ac_chr_bits[16] = {0, 2, 1, 2, 4, 4, 3, 4, 7, 5, 4, 4, 0, 1, 2, 0x77};
const uint8_t k_ac_chr_vals[162] = {
    0x00, 0x01, 0x02, 0x03, 0x11, 0x04, 0x05, 0x21, 0x31, 0x06, 0x12, 0x41, 0x51, 0x07, 0x61, 0x71, 0x13, 0x22,
    0x32, 0x81, 0x08, 0x14, 0x42, 0x91, 0xa1, 0xb1, 0xc1, 0x09, 0x23, 0x33, 0x52, 0xf0, 0x15, 0x62, 0x72, 0xd1,
    0x0a, 0x16, 0x24, 0x34, 0xe1, 0x25, 0xf1, 0x17, 0x18, 0x19, 0x1a, 0x26, 0x27, 0x28, 0x29, 0x2a, 0x35, 0x36,
    0x37, 0x38, 0x39, 0x3a, 0x43, 0x44, 0x45, 0x46, 0x47, 0x48, 0x49, 0x4a, 0x53, 0x54, 0x55, 0x56, 0x57, 0x58,
    0x59, 0x5a, 0x63, 0x64, 0x65, 0x66, 0x67, 0x68, 0x69, 0x6a, 0x73, 0x74, 0x75, 0x76, 0x77, 0x78, 0x79, 0x7a,
    0x82, 0x83, 0x84, 0x85, 0x86, 0x87, 0x88, 0x89, 0x8a, 0x92, 0x93, 0x94, 0x95, 0x96, 0x97, 0x98, 0x99, 0x9a,
    0xa2, 0xa3, 0xa4, 0xa5, 0xa6, 0xa7, 0xa8, 0xa9, 0xaa, 0xb2, 0xb3, 0xb4, 0xb5, 0xb6, 0xb7, 0xb8, 0xb9, 0xba,
    0xc2, 0xc3, 0xc4, 0xc5, 0xc6, 0xc7, 0xc8, 0xc9, 0xca, 0xd2, 0xd3, 0xd4, 0xd5, 0xd6, 0xd7, 0xd8, 0xd9, 0xda,
    0xe2, 0xe3, 0xe4, 0xe5, 0xe6, 0xe7, 0xe8, 0xe9, 0xea, 0xf2, 0xf3, 0xf4, 0xf5, 0xf6, 0xf7, 0xf8, 0xf9, 0xfa};

// libjpeg quality scaling (jpeg_quality_scaling + force_baseline), natural order
void quant_tables(int quality, uint16_t lum[64], uint16_t chr[64]) {
  if (quality <= 0) quality = 1;
  if (quality > 100) quality = 100;
  const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
  for (int i = 0; i < 64; i++) {
    long a = ((long)k_lum_q[i] * scale + 50L) / 100L, b = ((long)k_chr_q[i] * scale + 50L) / 100L;
    lum[i] = (uint16_t)(a <= 0 ? 1 : a > 255 ? 255 : a);
    chr[i] = (uint16_t)(b <= 0 ? 1 : b > 255 ? 255 : b);
  }
}

// Device tables.  c_q8/c_recip: [component class][zig-zag index] divisor (quantval << 3) and its
// reciprocal ceil(2^32 / divisor); floor(n * recip / 2^32) == n / divisor for n < 2^32 / divisor, and
// |coefficient| + divisor/2 < 2^17 here.
__constant__ uint32_t c_q8[2][64];
__constant__ uint32_t c_recip[2][64];
// Huffman: entry = (code << 8) | size ; [0] luminance, [1] chrominance
__constant__ uint32_t c_dc_tab[2][16];
__constant__ uint32_t c_ac_tab[2][256];

void build_huff(const uint8_t bits[16], const uint8_t *vals, uint32_t *tab, int n_tab) {
  for (int i = 0; i < n_tab; i++) tab[i] = 0;
  uint32_t code = 0;
  int k = 0;
  for (int len = 1; len <= 16; len++) {
    for (int i = 0; i < bits[len - 1]; i++, k++) { tab[vals[k]] = (code << 8) | (uint32_t)len; code++; }
    code <<= 1;
  }
}

// zig-zag position -> natural index, foldable at compile time inside unrolled device loops
__host__ __device__ constexpr int zz_nat(int k) {
  constexpr int t[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                         41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                         30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
  return t[k];
}

// ---- J1: colour transform + FDCT + quantisation ----------------------------------------------------------
#define JD(x, n) (((x) + (1 << ((n)-1))) >> (n))
__device__ __forceinline__ void fdct_1d(int &d0, int &d1, int &d2, int &d3, int &d4, int &d5, int &d6, int &d7,
                                        const bool first) {
  constexpr int CB = 13, P1 = 2;
  int tmp0 = d0 + d7, tmp7 = d0 - d7, tmp1 = d1 + d6, tmp6 = d1 - d6;
  int tmp2 = d2 + d5, tmp5 = d2 - d5, tmp3 = d3 + d4, tmp4 = d3 - d4;
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  const int sh = first ? CB - P1 : CB + P1;
  d0 = first ? (tmp10 + tmp11) << P1 : JD(tmp10 + tmp11, P1);
  d4 = first ? (tmp10 - tmp11) << P1 : JD(tmp10 - tmp11, P1);
  int z1 = (tmp12 + tmp13) * 4433;
  d2 = JD(z1 + tmp13 * 6270, sh);
  d6 = JD(z1 + tmp12 * (-15137), sh);
  z1 = tmp4 + tmp7;
  int z2 = tmp5 + tmp6, z3 = tmp4 + tmp6, z4 = tmp5 + tmp7;
  const int z5 = (z3 + z4) * 9633;
  tmp4 *= 2446; tmp5 *= 16819; tmp6 *= 25172; tmp7 *= 12299;
  z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;
  z3 += z5; z4 += z5;
  d7 = JD(tmp4 + z1 + z3, sh);
  d5 = JD(tmp5 + z2 + z4, sh);
  d3 = JD(tmp6 + z2 + z3, sh);
  d1 = JD(tmp7 + z1 + z4, sh);
}

struct JpegGeom {
  uint32_t width, height;   // canvas
  uint32_t y0;              // first canvas row of the band (multiple of 8)
  uint32_t mcus, mrows;     // MCUs per row, MCU rows in the band
  uint32_t rows_in_band;    // canvas rows held at `canvas`
};

// One CTA = 128 MCUs of one MCU row (8 canvas rows x 1024 pixels), one thread per MCU.  The 32 KB of pixels are staged
// in shared memory by coalesced 16-byte cp.async (the threads' own reads are 32 bytes at a stride of 32: half-used
// sectors, and the second and third component pass used to depend on L1 keeping them), then the three components of an
// MCU are done by the same thread one after the other from there.
static constexpr int JD_MCUS = 128;
__global__ void __launch_bounds__(JD_MCUS, 6) jpeg_dct_kernel(const uint8_t *canvas, size_t pitch, JpegGeom g, int16_t *coefs,
                                                              unsigned long long *nzmask) {
  __shared__ __align__(16) uint8_t px_sm[8][JD_MCUS * 32];
  const uint32_t m = blockIdx.x * JD_MCUS + threadIdx.x;  // MCU in the row
  const uint32_t x0 = blockIdx.x * JD_MCUS * 8u;          // first pixel column of the CTA
  const uint32_t px_here = min(g.width - x0, (uint32_t)JD_MCUS * 8u);   // pixels of the CTA's span that exist
  for (uint32_t mr = blockIdx.y; mr < g.mrows; mr += gridDim.y) {
    __syncthreads();   // the previous MCU row's reads are over
    for (uint32_t c = threadIdx.x; c < 8u * (JD_MCUS * 32u / 16u); c += JD_MCUS) {
      const uint32_t yy = c / (JD_MCUS * 2u), q = c % (JD_MCUS * 2u);   // row, 16-byte chunk (4 pixels) in the row
      if (q * 4u >= px_here) continue;
      uint32_t y = g.y0 + mr * 8u + yy;
      y = min(y, g.height - 1u) - g.y0;  // replicate the last image row below the bottom edge
      const uint32_t sa = (uint32_t)__cvta_generic_to_shared(&px_sm[yy][q * 16u]);
      const uint8_t *src = canvas + (size_t)y * pitch + (size_t)x0 * 4u + q * 16u;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(src) : "memory");
    }
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (m >= g.mcus) continue;
    for (uint32_t comp = 0; comp < 3u; comp++) {  // 0 Y, 1 Cb, 2 Cr
      int d[64];
#pragma unroll
      for (int yy = 0; yy < 8; yy++) {
        uint32_t px[8];
        if (m * 8u + 8u <= g.width) {
          const uint4 a = *reinterpret_cast<const uint4 *>(&px_sm[yy][threadIdx.x * 32u]), b = *reinterpret_cast<const uint4 *>(&px_sm[yy][threadIdx.x * 32u + 16u]);
          px[0] = a.x; px[1] = a.y; px[2] = a.z; px[3] = a.w; px[4] = b.x; px[5] = b.y; px[6] = b.z; px[7] = b.w;
        } else {
#pragma unroll
          for (int xx = 0; xx < 8; xx++) {
            const uint32_t x = min(m * 8u + xx, g.width - 1u);  // replicate the last column
            px[xx] = *reinterpret_cast<const uint32_t *>(&px_sm[yy][(x - x0) * 4u]);
          }
        }
#pragma unroll
        for (int xx = 0; xx < 8; xx++) {
          const int r = px[xx] & 0xff, gg = (px[xx] >> 8) & 0xff, b = (px[xx] >> 16) & 0xff;  // alpha dropped
          int v;
          if (comp == 0) v = (19595 * r + 38470 * gg + 7471 * b + 32768) >> 16;
          else if (comp == 1) v = (-11059 * r - 21709 * gg + 32768 * b + (128 << 16) + 32767) >> 16;
          else v = (32768 * r - 27439 * gg - 5329 * b + (128 << 16) + 32767) >> 16;
          d[yy * 8 + xx] = v - 128;
        }
      }
#pragma unroll
      for (int i = 0; i < 8; i++)
        fdct_1d(d[8 * i], d[8 * i + 1], d[8 * i + 2], d[8 * i + 3], d[8 * i + 4], d[8 * i + 5], d[8 * i + 6], d[8 * i + 7], true);
#pragma unroll
      for (int i = 0; i < 8; i++)
        fdct_1d(d[i], d[8 + i], d[16 + i], d[24 + i], d[32 + i], d[40 + i], d[48 + i], d[56 + i], false);
      const int cls = comp ? 1 : 0;
      uint32_t packed[32];
      uint32_t nz_lo = 0u, nz_hi = 0u;  // bit k: AC coefficient k (zigzag order) is not zero
#pragma unroll
      for (int k = 0; k < 64; k++) {
        const int c = d[zz_nat(k)];
        const uint32_t q8 = c_q8[cls][k], n = (uint32_t)abs(c) + (q8 >> 1);
        const int qv = (int)__umulhi(n, c_recip[cls][k]);
        const uint32_t v16 = (uint32_t)(c < 0 ? -qv : qv) & 0xffffu;
        if (k & 1) packed[k >> 1] |= v16 << 16; else packed[k >> 1] = v16;
        if (k > 0 && qv != 0) { if (k < 32) nz_lo |= 1u << (k & 31); else nz_hi |= 1u << (k & 31); }
      }
      const size_t blk_idx = ((size_t)mr * g.mcus + m) * 3 + comp;
      nzmask[blk_idx] = ((unsigned long long)nz_hi << 32) | nz_lo;
      uint4 *o = reinterpret_cast<uint4 *>(coefs + blk_idx * 64);
#pragma unroll
      for (int k = 0; k < 8; k++) o[k] = make_uint4(packed[4 * k], packed[4 * k + 1], packed[4 * k + 2], packed[4 * k + 3]);
    }
  }
}

// ---- Huffman helpers ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t mag_bits(int v) { return 32u - (uint32_t)__clz(abs(v)); }  // 0 for v == 0

struct HuffSmem {
  uint32_t dc[2][16];
  uint32_t ac[2][256];
};
// The MCU kernel's form of the tables: entry = ((code << s) << 6) | (code length + s), s = the symbol's low 4 bits
// (its magnitude bits go into the s low bits of the code field).
__device__ __forceinline__ void load_huff_sized(HuffSmem &h) {
  for (uint32_t i = threadIdx.x; i < 32; i += blockDim.x) {
    const uint32_t e = (&c_dc_tab[0][0])[i], s = i & 15u;
    (&h.dc[0][0])[i] = (((e >> 8) << s) << 6) | ((e & 0xffu) + s);
  }
  for (uint32_t i = threadIdx.x; i < 512; i += blockDim.x) {
    const uint32_t e = (&c_ac_tab[0][0])[i], s = i & 15u;
    (&h.ac[0][0])[i] = (((e >> 8) << s) << 6) | ((e & 0xffu) + s);
  }
}
__device__ __forceinline__ void load_huff(HuffSmem &h) {
  for (uint32_t i = threadIdx.x; i < 32; i += blockDim.x) (&h.dc[0][0])[i] = (&c_dc_tab[0][0])[i];
  for (uint32_t i = threadIdx.x; i < 512; i += blockDim.x) (&h.ac[0][0])[i] = (&c_ac_tab[0][0])[i];
}

// Walks one block's symbols.  EMIT: void(uint32_t code, uint32_t size) with size <= 27.
// w: the block's 32 coefficient words (shared memory), mask: its non-zero AC positions.  Each thread steps
// through ITS non-zero coefficients, so a warp iterates max-over-lanes(non-zeros) times, not 63 times with
// a handful of lanes alive in each.
template <class EMIT>
__device__ __forceinline__ void code_block(const uint32_t *w, unsigned long long mask, int pred, const uint32_t *dc,
                                           const uint32_t *ac, EMIT emit) {
  const int dcv = (int)(int16_t)(w[0] & 0xffffu);
  const int diff = dcv - pred;
  uint32_t s = mag_bits(diff);
  uint32_t e = dc[s];
  // code then magnitude bits (negative values: low bits of diff - 1)
  emit(((e >> 8) << s) | ((uint32_t)(diff < 0 ? diff - 1 : diff) & ((1u << s) - 1u)), (e & 0xffu) + s);
  int prev = 0;
  while (mask) {
    const int k = __ffsll((long long)mask) - 1;
    mask &= mask - 1ull;
    uint32_t run = (uint32_t)(k - prev - 1);
    prev = k;
    const uint32_t ww = w[k >> 1];
    const int v = (int)(int16_t)((k & 1) ? (ww >> 16) : (ww & 0xffffu));
    while (run > 15u) { e = ac[0xF0]; emit(e >> 8, e & 0xffu); run -= 16u; }
    s = mag_bits(v);
    e = ac[(run << 4) | s];
    emit(((e >> 8) << s) | ((uint32_t)(v < 0 ? v - 1 : v) & ((1u << s) - 1u)), (e & 0xffu) + s);
  }
  if (prev != 63) { e = ac[0]; emit(e >> 8, e & 0xffu); }
}

// 256 consecutive blocks (32 KB) -> shared memory, coalesced; block t at words [33 t, 33 t + 32) (odd stride:
// the threads' accesses to "their" block spread over the banks).
static constexpr int JB_STRIDE = 33;
__device__ __forceinline__ void stage_blocks(uint32_t *sm, const int16_t *coefs, size_t first_blk, size_t n_blk_total) {
  const uint4 *src = reinterpret_cast<const uint4 *>(coefs + first_blk * 64);
  const size_t avail = (n_blk_total - first_blk) * 8;  // 16-byte pieces left
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const uint32_t i = (uint32_t)j * 256u + threadIdx.x;
    if (i < avail) {
      const uint4 v = __ldg(src + i);
      uint32_t *d = sm + (i >> 3) * JB_STRIDE + (i & 7u) * 4u;
      d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
    }
  }
}

// DC predictor of block t of a staged chunk: previous MCU, same component; a restart resets it to 0.
// blk_in_row: the block's index inside its MCU row; the first 3 blocks of a chunk look at global memory.
__device__ __forceinline__ int block_pred(const uint32_t *sm, uint32_t t, const int16_t *coefs, size_t blk, uint32_t blk_in_row) {
  if (blk_in_row < 3u) return 0;
  if (t >= 3u) return (int)(int16_t)(sm[(t - 3u) * JB_STRIDE] & 0xffffu);
  return (int)__ldg(coefs + (blk - 3) * 64);
}

// ---- J2: code lengths + scan, one CTA per MCU row -------------------------------------------------------
__global__ void __launch_bounds__(256) jpeg_len_kernel(const int16_t *coefs, const unsigned long long *nzmask, JpegGeom g,
                                                       uint32_t *bit_off, uint32_t *row_bits) {
  __shared__ HuffSmem hs;
  __shared__ uint32_t warp_tot[8];
  __shared__ uint32_t s_base;
  __shared__ uint32_t sblk[256 * JB_STRIDE];
  load_huff(hs);
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  for (uint32_t mr = blockIdx.x; mr < g.mrows; mr += gridDim.x) {
    const uint32_t n_blk = g.mcus * 3u;
    const size_t row0 = (size_t)mr * n_blk;
    for (uint32_t b0 = 0; b0 < n_blk; b0 += 256u) {
      stage_blocks(sblk, coefs, row0 + b0, row0 + n_blk);
      __syncthreads();
      const uint32_t b = b0 + threadIdx.x;
      uint32_t len = 0;
      if (b < n_blk) {
        const uint32_t cls = (b % 3u) ? 1u : 0u;
        code_block(sblk + threadIdx.x * JB_STRIDE, nzmask[row0 + b], block_pred(sblk, threadIdx.x, coefs, row0 + b, b),
                   hs.dc[cls], hs.ac[cls], [&](uint32_t, uint32_t size) { len += size; });
      }
      // block-wide exclusive scan
      uint32_t incl = len;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += t; }
      if (lane == 31u) warp_tot[warp] = incl;
      __syncthreads();
      uint32_t wbase = 0;
#pragma unroll
      for (int k = 0; k < 8; k++) if ((uint32_t)k < warp) wbase += warp_tot[k];
      const uint32_t base = s_base;
      if (b < n_blk) bit_off[row0 + b] = base + wbase + incl - len;
      __syncthreads();
      if (threadIdx.x == 255u) s_base = base + wbase + incl;
      __syncthreads();
    }
    if (threadIdx.x == 0) { row_bits[mr] = s_base; s_base = 0; }
    __syncthreads();
  }
}

// ---- J3/J6: exclusive scan over the MCU rows of the band (single CTA) ------------------------------------
// mode 0: size = ceil(bits / 8)          -> base[], totals[0]
// mode 1: size = bytes + ff + (rst ? 2)   -> base[], totals[1]
__global__ void __launch_bounds__(1024) jpeg_rowscan_kernel(const uint32_t *row_bits, const uint32_t *row_ff, uint32_t n_rows,
                                                            uint32_t first_mrow, uint32_t total_mrows, int mode,
                                                            unsigned long long *base, unsigned long long *totals) {
  __shared__ unsigned long long warp_tot[32];
  __shared__ unsigned long long s_base;
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  for (uint32_t r0 = 0; r0 < n_rows; r0 += 1024u) {
    const uint32_t r = r0 + threadIdx.x;
    unsigned long long sz = 0;
    if (r < n_rows) {
      sz = (row_bits[r] + 7u) >> 3;
      if (mode == 1) sz += row_ff[r] + ((first_mrow + r + 1u < total_mrows) ? 2u : 0u);
    }
    unsigned long long incl = sz;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += t; }
    if (lane == 31u) warp_tot[warp] = incl;
    __syncthreads();
    unsigned long long wbase = 0;
    for (uint32_t k = 0; k < warp; k++) wbase += warp_tot[k];
    const unsigned long long b = s_base;
    if (r < n_rows) base[r] = b + wbase + incl - sz;
    __syncthreads();
    if (threadIdx.x == 1023u) s_base = b + wbase + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) totals[mode] = s_base;
}

// ---- J4: emit, thread = one block ------------------------------------------------------------------------
struct BitSink {
  uint32_t *words;          // unstuffed stream as 32-bit words (stream byte j = byte j of memory)
  unsigned long long pos;   // next bit position in the stream
  unsigned long long acc;   // pending bits, right-aligned
  uint32_t nacc;
  __device__ __forceinline__ void flush_word(uint32_t be, uint32_t first_bit_in_word) {
    // `be` holds stream bits MSB-first starting at first_bit_in_word of the word at pos/32
    (void)first_bit_in_word;
    atomicOr(words + (pos >> 5), __byte_perm(be, 0, 0x0123));
  }
  __device__ __forceinline__ void put(uint32_t code, uint32_t size) {
    acc = (acc << size) | code;
    nacc += size;
    while (nacc >= 32u || (nacc && ((pos & 31u) + nacc >= 32u))) {
      // fill the current word up to its end
      const uint32_t room = 32u - (uint32_t)(pos & 31u);
      const uint32_t take = room;  // nacc >= room here
      const uint32_t bits = (uint32_t)(acc >> (nacc - take)) & (take == 32u ? 0xffffffffu : ((1u << take) - 1u));
      flush_word(bits, 0);  // right-aligned within the word's remaining `room` bits == correct position
      pos += take;
      nacc -= take;
    }
  }
  __device__ __forceinline__ void finish() {
    if (nacc) {
      const uint32_t room = 32u - (uint32_t)(pos & 31u);  // nacc < room
      const uint32_t bits = ((uint32_t)acc & ((1u << nacc) - 1u)) << (room - nacc);
      flush_word(bits, 0);
      pos += nacc;
      nacc = 0;
    }
  }
};

__global__ void __launch_bounds__(256) jpeg_emit_kernel(const int16_t *coefs, const unsigned long long *nzmask, JpegGeom g,
                                                        const uint32_t *bit_off, const uint32_t *row_bits,
                                                        const unsigned long long *row_base, uint32_t *ubuf) {
  __shared__ HuffSmem hs;
  __shared__ uint32_t sblk[256 * JB_STRIDE];
  load_huff(hs);
  const uint32_t n_blk = g.mcus * 3u;
  const size_t total = (size_t)g.mrows * n_blk;
  for (size_t c0 = (size_t)blockIdx.x * 256u; c0 < total; c0 += (size_t)gridDim.x * 256u) {
    __syncthreads();  // the previous chunk is done with sblk (and the tables are loaded)
    stage_blocks(sblk, coefs, c0, total);
    __syncthreads();
    const size_t i = c0 + threadIdx.x;
    if (i >= total) continue;
    const uint32_t mr = (uint32_t)(i / n_blk), b = (uint32_t)(i - (size_t)mr * n_blk);
    const uint32_t cls = (b % 3u) ? 1u : 0u;
    BitSink sink;
    sink.words = ubuf; sink.pos = row_base[mr] * 8ull + bit_off[i]; sink.acc = 0; sink.nacc = 0;
    code_block(sblk + threadIdx.x * JB_STRIDE, nzmask[i], block_pred(sblk, threadIdx.x, coefs, i, b), hs.dc[cls], hs.ac[cls],
               [&](uint32_t code, uint32_t size) { sink.put(code, size); });
    if (b == n_blk - 1u) {  // pad the row segment with 1-bits to a byte boundary
      const uint32_t pad = (8u - (row_bits[mr] & 7u)) & 7u;
      if (pad) sink.put((1u << pad) - 1u, pad);
    }
    sink.finish();
  }
}

// ---- J5: count 0xFF bytes per MCU row ---------------------------------------------------------------------
__global__ void __launch_bounds__(256) jpeg_ffcount_kernel(const uint8_t *ubuf, const uint32_t *row_bits,
                                                           const unsigned long long *row_base, uint32_t n_rows, uint32_t *row_ff) {
  __shared__ uint32_t s_cnt;
  for (uint32_t r = blockIdx.x; r < n_rows; r += gridDim.x) {
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    const uint8_t *p = ubuf + row_base[r];
    const uint32_t n = (row_bits[r] + 7u) >> 3;
    uint32_t c = 0;
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) c += p[i] == 0xFFu;
    c = __reduce_add_sync(0xffffffffu, c);
    if ((threadIdx.x & 31u) == 0 && c) atomicAdd(&s_cnt, c);
    __syncthreads();
    if (threadIdx.x == 0) row_ff[r] = s_cnt;
    __syncthreads();
  }
}

// ---- J7: byte stuffing + RSTm ------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) jpeg_stuff_kernel(const uint8_t *ubuf, const uint32_t *row_bits,
                                                         const unsigned long long *row_base, const unsigned long long *out_base,
                                                         uint32_t n_rows, uint32_t first_mrow, uint32_t total_mrows, uint8_t *out) {
  __shared__ uint32_t warp_tot[8];
  __shared__ uint32_t s_ff;
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  for (uint32_t r = blockIdx.x; r < n_rows; r += gridDim.x) {
    if (threadIdx.x == 0) s_ff = 0;
    __syncthreads();
    const uint8_t *p = ubuf + row_base[r];
    uint8_t *o = out + out_base[r];
    const uint32_t n = (row_bits[r] + 7u) >> 3;
    for (uint32_t i0 = 0; i0 < n; i0 += 256u) {
      const uint32_t i = i0 + threadIdx.x;
      const uint32_t byte = i < n ? p[i] : 0u;
      const uint32_t is_ff = byte == 0xFFu ? 1u : 0u;
      uint32_t incl = is_ff;
#pragma unroll
      for (int k = 1; k < 32; k <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, k); if (lane >= (uint32_t)k) incl += t; }
      if (lane == 31u) warp_tot[warp] = incl;
      __syncthreads();
      uint32_t wbase = 0;
#pragma unroll
      for (int k = 0; k < 8; k++) if ((uint32_t)k < warp) wbase += warp_tot[k];
      const uint32_t before = s_ff + wbase + incl - is_ff;
      if (i < n) {
        o[i + before] = (uint8_t)byte;
        if (is_ff) o[i + before + 1] = 0;
      }
      __syncthreads();
      if (threadIdx.x == 255u) s_ff += wbase + incl;
      __syncthreads();
    }
    if (threadIdx.x == 0 && first_mrow + r + 1u < total_mrows) {
      o[n + s_ff] = 0xFF;
      o[n + s_ff + 1] = (uint8_t)(0xD0u + ((first_mrow + r) & 7u));
    }
    __syncthreads();
  }
}

// =========================================================================================================
// The one-pass route: colour, FDCT, quantisation AND Huffman coding in one kernel, with no coefficients in HBM.
//
//   M1 mcu      CTA = 128 MCUs of one MCU row ("chunk"), thread = one MCU.  The quantised coefficients never
//               leave the SM: the unrolled zig-zag loop keeps the non-zero ones, 16 bits each, in the thread's
//               private column of shared memory, a short loop codes them into the thread's bit accumulator,
//               whose full words come up the same column behind them.  The MCU lengths are scanned across the
//               CTA, the private words are shifted into place in a chunk buffer (the pixel staging area, dead by
//               then) and the chunk goes to its fixed slot in HBM.  An exact mosaic of RGBA8 tiles is read in
//               place (jpeg_direct_plan), anything else from the composited canvas.
//   M2 rowinfo  CTA = one MCU row: bit offsets of its chunks, size of the row segment, its 0xFF bytes
//   J6 rowscan  final offsets (stuffing + RSTm), as on the other route
//   M3 stuff    CTA = one MCU row: chunks funnelled together 32 bits at a time, stuffed, RSTm
//
// A thread holds JM_WORDS words of code for its MCU; an MCU that needs more is coded again into a slot in HBM
// (jm_mcu_slow).  A chunk holds JM_CHUNK_BYTES; a band with a chunk beyond that, or with more MCUs in slots than
// there are slots (noise at quality 95 and up), raises a flag and is done again by the J1..J7 route above.
static constexpr int JM_MCUS = 128;
static constexpr int JM_WORDS = 40;                       // 1280 bits per MCU (20 per pixel); what four CTAs per SM leave room for
static constexpr int JM_GAP = 3;                          // words between a block's first entry and the code before it (the pending word, the DC code, slack)
static constexpr int JM_PX = JM_MCUS * 8 + 8;             // staged pixels per row: the MCU on the left, then the chunk's
static constexpr uint32_t JM_CHUNK_BYTES = 8192;          // 8 bits per pixel on average over a chunk
static constexpr uint32_t JM_CHUNK_WORDS = JM_CHUNK_BYTES / 4;

__constant__ uint8_t c_zigzag_dev[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                         41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                         30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct JpegQuant {   // zig-zag order; [0] luminance, [1] chrominance; divisor = quantval << 3 as in c_q8
  uint2 hr[2][64];         // x: divisor / 2, y: ceil(2^32 / divisor) -- one 64-bit constant load for the pair
};

// Direct staging: where the band is an exact mosaic of RGBA8 tiles (nothing overlaps, nothing shows the background,
// no blending) the MCU kernel reads the tiles themselves and no canvas is composited.  The band's rows are cut
// at every tile edge into groups; inside a group the same tiles cover every row, left to right.
struct JmSeg { uint32_t x0, pitch; const uint8_t *p; };          // p: the pixel at column x0 of the group's first row
struct JmGroup { uint32_t row0, row1, seg0, nseg; };             // canvas rows [row0, row1)
struct JmDirect {
  const JmGroup *groups; const JmSeg *segs;
  const uint32_t *mrow_group;   // group of every MCU row of the band
  const uint32_t *chunk_seg;    // [group][chunk]: index in segs of the tile at the chunk's first staged column
};

struct JmSmem {
  uint8_t px[8][JM_PX * 4];              // RGBA rows; reused as the chunk buffer (words) once the pixels are consumed
  uint32_t stage[JM_WORDS][JM_MCUS];     // word j of thread t's code at [j][t]
  HuffSmem hs;
  uint32_t warp_tot[JM_MCUS / 32];
};

// v - 128 for one pixel word (R | G << 8 | B << 16 | A << 24); the level shift rides in the rounding constant
__device__ __forceinline__ int jm_y(uint32_t px) {
  uint32_t v;
  asm("dp2a.lo.u32.u32 %0, %1, %2, %3;" : "=r"(v) : "r"(19595u | (38470u << 16)), "r"(px), "r"(32768u - (128u << 16)));
  asm("dp2a.hi.u32.u32 %0, %1, %2, %3;" : "=r"(v) : "r"(7471u), "r"(px), "r"(v));
  return (int)v >> 16;
}
// chroma: (32767 - N) >> 16 with N = a1.lo R + a1.hi G + a2.lo B  (signed 16-bit weights: the + 32768 of the JFIF
// formulas does not fit, its negative does, so the weights are the negated ones)
__device__ __forceinline__ int jm_c(uint32_t px, uint32_t a1, uint32_t a2) {
  int n;
  asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(n) : "r"(a1), "r"(px), "r"(0));
  asm("dp2a.hi.s32.u32 %0, %1, %2, %3;" : "=r"(n) : "r"(a2), "r"(px), "r"(n));
  return (32767 - n) >> 16;
}
static constexpr uint32_t JM_CB1 = 11059u | (21709u << 16), JM_CB2 = 0x8000u;    // N = 11059 R + 21709 G - 32768 B
static constexpr uint32_t JM_CR1 = 0x8000u | (27439u << 16), JM_CR2 = 5329u;     // N = -32768 R + 27439 G + 5329 B

struct JmBits {
  unsigned long long acc;   // pending bits in the low nacc bits (above them: bits already stored)
  uint32_t nacc, wr;        // wr: words produced so far
  uint32_t ovf;             // a word had no room: the MCU goes to the other route
  uint32_t *col;            // &stage[0][t]
  // size <= 32; lim: first word of the column that is not free (JM_WORDS, or where the block's unread coefficients start)
  __device__ __forceinline__ void put(uint32_t code, uint32_t size, uint32_t lim) {
    acc = (acc << size) | code;
    nacc += size;
    if (nacc >= 32u) {
      nacc -= 32u;
      if (wr < lim) col[wr * JM_MCUS] = (uint32_t)(acc >> nacc); else ovf = 1u;
      wr++;
    }
  }
};

// An MCU whose code does not fit its column (the corner where four tiles meet inside a block; a patch of noise): the
// thread codes it again, from its staged pixels, into a slot of worst-case size in HBM -- rolled loops, the block in
// local memory, a step per coefficient: slow, and rare.  Returns the length in bits.
static constexpr uint32_t JM_ESC_WORDS = 160;     // 3 x (20 bits of DC + 63 x 26 bits of AC) = 4974 bits
static constexpr uint32_t JM_ESC_SLOTS = 16384;   // more MCUs than that in a band: the band goes to the J1..J7 route
__device__ __noinline__ uint32_t jm_mcu_slow(const JmSmem &sm, const JpegQuant &q, const uint32_t t, const bool has_left, uint32_t *esc) {
  unsigned long long acc = 0;
  uint32_t nacc = 0, wr = 0;
  auto put = [&](uint32_t code, uint32_t size) {
    acc = (acc << size) | code;
    nacc += size;
    if (nacc >= 32u) { nacc -= 32u; esc[wr++] = (uint32_t)(acc >> nacc); }
  };
  for (uint32_t comp = 0; comp < 3u; comp++) {
    const uint32_t cls = comp ? 1u : 0u;
    const uint32_t a1 = comp == 1u ? JM_CB1 : JM_CR1, a2 = comp == 1u ? JM_CB2 : JM_CR2;
    int d[64];
    int pred = 0;
    if (has_left) {   // the quantised DC of the MCU on the left: the sum of its samples (staged 8 pixels before this thread's)
      int s2 = 0;
#pragma unroll 1
      for (uint32_t i = 0; i < 64u; i++) {
        const uint32_t px = *reinterpret_cast<const uint32_t *>(&sm.px[i >> 3][t * 32u + (i & 7u) * 4u]);
        s2 += comp == 0u ? jm_y(px) : jm_c(px, a1, a2);
      }
      const int qa = (int)__umulhi((uint32_t)abs(s2) + q.hr[cls][0].x, q.hr[cls][0].y);
      pred = s2 < 0 ? -qa : qa;
    }
#pragma unroll 1
    for (uint32_t i = 0; i < 64u; i++) {
      const uint32_t px = *reinterpret_cast<const uint32_t *>(&sm.px[i >> 3][32u + t * 32u + (i & 7u) * 4u]);
      d[i] = comp == 0u ? jm_y(px) : jm_c(px, a1, a2);
    }
#pragma unroll 1
    for (int i = 0; i < 8; i++)
      fdct_1d(d[8 * i], d[8 * i + 1], d[8 * i + 2], d[8 * i + 3], d[8 * i + 4], d[8 * i + 5], d[8 * i + 6], d[8 * i + 7], true);
#pragma unroll 1
    for (int i = 0; i < 8; i++)
      fdct_1d(d[i], d[8 + i], d[16 + i], d[24 + i], d[32 + i], d[40 + i], d[48 + i], d[56 + i], false);
    const uint32_t *dct = sm.hs.dc[cls], *act = sm.hs.ac[cls];
    {
      const int qa = (int)__umulhi((uint32_t)abs(d[0]) + q.hr[cls][0].x, q.hr[cls][0].y);
      const int diff = (d[0] < 0 ? -qa : qa) - pred;
      const uint32_t a = (uint32_t)abs(diff), sz = 32u - (uint32_t)__clz((int)a), e = dct[sz];
      put((e >> 6) | ((a ^ (uint32_t)(diff >> 31)) & ((1u << sz) - 1u)), e & 63u);
    }
    int prev = 0;
#pragma unroll 1
    for (int k = 1; k < 64; k++) {
      const int c = d[c_zigzag_dev[k]];
      const uint32_t a = __umulhi((uint32_t)abs(c) + q.hr[cls][k].x, q.hr[cls][k].y);
      if (a == 0u) continue;
      uint32_t run = (uint32_t)(k - 1 - prev);
      prev = k;
      const uint32_t z = act[0xF0];
      while (run > 15u) { put(z >> 6, z & 63u); run -= 16u; }
      const uint32_t sz = 32u - (uint32_t)__clz((int)a), e = act[(run << 4) | sz];
      put((e >> 6) | ((a ^ (uint32_t)(c >> 31)) & ((1u << sz) - 1u)), e & 63u);
    }
    if (prev != 63) { const uint32_t e = act[0]; put(e >> 6, e & 63u); }
  }
  if (nacc) esc[wr] = (uint32_t)acc << (32u - nacc);
  return wr * 32u + nacc;
}

__global__ void __launch_bounds__(JM_MCUS, 4) jpeg_mcu_kernel(const uint8_t *canvas, size_t pitch, JpegGeom g,
                                                              const __grid_constant__ JpegQuant q, const JmDirect dir, uint32_t *chunk_words,
                                                              uint32_t *chunk_bits, uint32_t *esc_pool, unsigned long long *overflow) {
  extern __shared__ __align__(16) uint8_t jm_raw[];
  JmSmem &sm = *reinterpret_cast<JmSmem *>(jm_raw);
  load_huff_sized(sm.hs);
  const uint32_t t = threadIdx.x, lane = t & 31u, warp = t >> 5;
  const uint32_t m = blockIdx.x * JM_MCUS + t;          // MCU in the row
  const uint32_t x0 = blockIdx.x * JM_MCUS * 8u;        // first pixel column of the chunk
  const uint32_t px_here = min(g.width - x0, (uint32_t)JM_MCUS * 8u) + 8u;   // staged pixels that exist (with the 8 on the left)
  const bool active = m < g.mcus;
  const bool warp_active = blockIdx.x * JM_MCUS + warp * 32u < g.mcus;
  uint32_t *cw = reinterpret_cast<uint32_t *>(&sm.px[0][0]);
  for (uint32_t mr = blockIdx.y; mr < g.mrows; mr += gridDim.y) {
    // the previous chunk has left the buffer (and the tables are loaded); and once the band is lost to the other route
    // (a chunk beyond its 8 KB, no slot left) the chunks not yet begun are not worth coding
    if (__syncthreads_or(t == 0u && *reinterpret_cast<const volatile unsigned long long *>(overflow) != 0ull)) {
      if (t == 0u) for (; mr < g.mrows; mr += gridDim.y) chunk_bits[(size_t)mr * gridDim.x + blockIdx.x] = 0u;   // the row kernels still run over the band
      return;
    }
    // Staging: a lane takes a 16-byte piece (4 pixels) of the chunk's span and copies its 8 rows; the 256 pieces of the
    // chunk are spread over the 4 warps (2 x 32 consecutive pieces each), the 2 pieces of the MCU on the left go to
    // two lanes of warp 0.  Rows past the bottom edge repeat the last image row.
    const uint32_t yf = g.y0 + mr * 8u, rows_here = min(8u, g.height - yf);
    JmGroup grp;
    uint32_t si0 = 0, gi = 0;
    if (dir.groups) {
      gi = __ldg(dir.mrow_group + mr);
      grp = dir.groups[gi];
      si0 = __ldg(dir.chunk_seg + (size_t)gi * gridDim.x + blockIdx.x);
    }
    auto stage_piece = [&](const uint32_t p4) {
      if (p4 * 4u >= px_here) return;
      const uint32_t x = x0 + p4 * 4u - 8u;
      const uint32_t sa = (uint32_t)__cvta_generic_to_shared(&sm.px[0][p4 * 16u]);
      const uint8_t *src;
      size_t pit;
      if (dir.groups) {
        const uint32_t seg_end = grp.seg0 + grp.nseg;
        uint32_t si = si0;
        while (si + 1u < seg_end && __ldg(&dir.segs[si + 1u].x0) <= x) si++;
        const JmSeg sg = dir.segs[si];
        const uint32_t x_end = si + 1u < seg_end ? __ldg(&dir.segs[si + 1u].x0) : g.width;   // where the tile's columns end
        pit = sg.pitch;
        src = sg.p + (size_t)(yf - grp.row0) * pit + (size_t)(x - sg.x0) * 4u;
        if (yf + rows_here <= grp.row1 && x + 4u <= x_end && (((reinterpret_cast<uintptr_t>(src) | pit) & 15u) != 0u)) {
          // one tile, but its rows are not 16-byte aligned here (an odd width or position): four 4-byte copies a row
#pragma unroll
          for (uint32_t k = 0; k < 8u; k++) {
#pragma unroll
            for (uint32_t j = 0; j < 4u; j++)
              asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa + k * (uint32_t)(JM_PX * 4) + j * 4u), "l"(src + j * 4u) : "memory");
            if (k + 1u < rows_here) src += pit;
          }
          return;
        }
        if (yf + rows_here > grp.row1 || x + 4u > x_end) {
          // The general case, pixel by pixel: the MCU row crosses a tile edge (its rows belong to two groups), or the piece
          // does (two tiles, or the canvas ends inside it).  Rare: tile edges.
          uint32_t gk = gi;
          JmGroup gr = grp;
          for (uint32_t k = 0; k < 8u; k++) {
            const uint32_t y = yf + min(k, rows_here - 1u);
            while (y >= gr.row1) gr = dir.groups[++gk];
            for (uint32_t j = 0; j < 4u && x + j < g.width; j++) {
              uint32_t lo = 0, hi = gr.nseg;   // the group's tile at column x + j
              while (hi - lo > 1u) { const uint32_t mid = (lo + hi) >> 1; if (__ldg(&dir.segs[gr.seg0 + mid].x0) <= x + j) lo = mid; else hi = mid; }
              const JmSeg s1 = dir.segs[gr.seg0 + lo];
              const uint8_t *s = s1.p + (size_t)(y - gr.row0) * s1.pitch + (size_t)(x + j - s1.x0) * 4u;
              asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(sa + k * (uint32_t)(JM_PX * 4) + j * 4u), "l"(s) : "memory");
            }
          }
          return;
        }
      } else {
        pit = pitch;
        src = canvas + (size_t)(yf - g.y0) * pit + (size_t)x * 4u;
      }
#pragma unroll
      for (uint32_t k = 0; k < 8u; k++) {
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa + k * (uint32_t)(JM_PX * 4)), "l"(src) : "memory");
        if (k + 1u < rows_here) src += pit;
      }
    };
    stage_piece(2u + warp * 64u + lane);
    stage_piece(2u + warp * 64u + 32u + lane);
    if (warp == 0u && lane < 2u && x0 != 0u) stage_piece(lane);
    asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if ((g.width & 7u) && x0 + (uint32_t)JM_MCUS * 8u >= g.width) {   // the chunk holds the image's partial MCU: replicate the last column
      if (t < 56u) {
        const uint32_t yy = t / 7u, p = g.width - x0 + 8u + t % 7u;   // staged pixel index of a column past the edge
        if (p < ((g.width + 7u) & ~7u) - x0 + 8u)
          *reinterpret_cast<uint32_t *>(&sm.px[yy][p * 4u]) = *reinterpret_cast<const uint32_t *>(&sm.px[yy][(g.width - 1u - x0 + 8u) * 4u]);
      }
      __syncthreads();
    }
    JmBits b;
    b.acc = 0; b.nacc = 0; b.wr = 0; b.ovf = 0; b.col = &sm.stage[0][t];
    if (warp_active) {
      uint8_t *colb = reinterpret_cast<uint8_t *>(b.col);
      for (uint32_t comp = 0; comp < 3u; comp++) {  // 0 Y, 1 Cb, 2 Cr
        const uint32_t cls = comp ? 1u : 0u;
        const uint32_t a1 = comp == 1u ? JM_CB1 : JM_CR1, a2 = comp == 1u ? JM_CB2 : JM_CR2;
        // DC of the MCU on the warp's left (lane 0's predictor): the DC term of the islow FDCT is exactly the
        // sum of the 64 level-shifted samples, so the warp adds them up, two pixels a lane
        int left_dc = 0;
        if (blockIdx.x != 0u || warp != 0u) {
          const uint2 two = *reinterpret_cast<const uint2 *>(&sm.px[lane >> 2][(warp * 256u + (lane & 3u) * 2u) * 4u]);
          int s2 = comp == 0u ? jm_y(two.x) + jm_y(two.y) : jm_c(two.x, a1, a2) + jm_c(two.y, a1, a2);
          s2 = __reduce_add_sync(0xffffffffu, s2);
          const int qa = (int)__umulhi((uint32_t)abs(s2) + q.hr[cls][0].x, q.hr[cls][0].y);
          left_dc = s2 < 0 ? -qa : qa;
        }
        int d[64];
#pragma unroll
        for (int yy = 0; yy < 8; yy++) {
          const uint4 a = *reinterpret_cast<const uint4 *>(&sm.px[yy][32u + t * 32u]), c4 = *reinterpret_cast<const uint4 *>(&sm.px[yy][48u + t * 32u]);
          const uint32_t px[8] = {a.x, a.y, a.z, a.w, c4.x, c4.y, c4.z, c4.w};
          if (comp == 0u) {
#pragma unroll
            for (int xx = 0; xx < 8; xx++) d[yy * 8 + xx] = jm_y(px[xx]);
          } else {
#pragma unroll
            for (int xx = 0; xx < 8; xx++) d[yy * 8 + xx] = jm_c(px[xx], a1, a2);
          }
        }
#pragma unroll
        for (int i = 0; i < 8; i++)
          fdct_1d(d[8 * i], d[8 * i + 1], d[8 * i + 2], d[8 * i + 3], d[8 * i + 4], d[8 * i + 5], d[8 * i + 6], d[8 * i + 7], true);
#pragma unroll
        for (int i = 0; i < 8; i++)
          fdct_1d(d[i], d[8 + i], d[16 + i], d[24 + i], d[32 + i], d[40 + i], d[48 + i], d[56 + i], false);
        const uint32_t *dct = sm.hs.dc[cls], *act = sm.hs.ac[cls];
        const int qa0 = (int)__umulhi((uint32_t)abs(d[0]) + q.hr[cls][0].x, q.hr[cls][0].y);
        const int dcq = d[0] < 0 ? -qa0 : qa0;
        // Quantise the 63 AC terms in straight-line code; the non-zero ones go, 16 bits each, into the thread's column
        // of code words a little above the code written so far, in zig-zag order upwards: entry e in half e & 1 of word
        // first + e / 2.  The walk below reads them from the bottom while the code comes up behind it (a word of code
        // may only be stored below the word of the next unread entry), so the entries' room is used twice.  Their
        // zig-zag positions stay in a 64-bit mask.  Entries that find no room below the column's end are dropped and
        // counted: the MCU is then coded again into a slot (jm_mcu_slow), like one whose code catches up with its entries.
        uint32_t nz_lo = 0, nz_hi = 0;
        const uint32_t first = min(b.wr + (uint32_t)JM_GAP, (uint32_t)JM_WORDS);
        const uint32_t col32 = (uint32_t)__cvta_generic_to_shared(colb);
        const uint32_t wa_last = col32 + (JM_WORDS - 1) * JM_MCUS * 4 + 2;   // the column's last place: entries beyond it pile up there
        uint32_t wa = min(col32 + first * (JM_MCUS * 4), wa_last), wstep = 2;
#pragma unroll
        for (int k = 1; k < 64; k++) {
          const int c = d[zz_nat(k)];
          const uint2 hr = q.hr[cls][k];
          const uint32_t a = __umulhi((uint32_t)abs(c) + hr.x, hr.y);
          if (a != 0u) {
            if (k < 32) nz_lo |= 1u << k; else nz_hi |= 1u << (k - 32);
            asm volatile("st.shared.u16 [%0], %1;" ::"r"(wa), "h"((uint16_t)(a ^ (uint32_t)(c >> 31))) : "memory");   // a, or ~a for a negative term: its magnitude bits as they are coded
            wa = min(wa + wstep, wa_last);
            wstep = JM_MCUS * 4 - wstep;   // 2, 510, 2, ...
          }
        }
        const uint32_t n_ent = (uint32_t)(__popc(nz_lo) + __popc(nz_hi));
        if (n_ent > 2u * ((uint32_t)JM_WORDS - first)) b.ovf = 1u;   // entries without a place of their own
        int pred = __shfl_up_sync(0xffffffffu, dcq, 1);
        if (lane == 0u) pred = left_dc;
        if (b.ovf) continue;   // the MCU will be coded again (and its entries may be incomplete): no walk
        {  // DC difference against the MCU on the left (0 at the start of the row: a restart interval)
          const int diff = dcq - pred;
          const uint32_t a = (uint32_t)abs(diff), s = 32u - (uint32_t)__clz((int)a), e = dct[s];
          b.put((e >> 6) | ((a ^ (uint32_t)(diff >> 31)) & ((1u << s) - 1u)), e & 63u, n_ent ? first : (uint32_t)JM_WORDS);
        }
        int prev = 0;
        const uint8_t *rp = colb + first * (JM_MCUS * 4);
        int rstep = 2;
        uint32_t left = n_ent;
        // each lane walks ITS non-zero terms: zig-zag positions 1..31, then 32..63
#pragma unroll 1
        for (int half = 0; half < 2; half++) {
          uint32_t mk = half ? nz_hi : nz_lo;
          const int kb = half * 32;
          while (mk) {
            const int k = kb + __ffs((int)mk) - 1;
            mk &= mk - 1u;
            const uint32_t x = *reinterpret_cast<const uint16_t *>(rp);
            rp += rstep;
            rstep = JM_MCUS * 4 - rstep;
            left--;
            const uint32_t lim = left ? (uint32_t)(rp - colb) / (JM_MCUS * 4u) : (uint32_t)JM_WORDS;   // the word of the next unread entry
            uint32_t run = (uint32_t)(k - 1 - prev);
            prev = k;
            if (run > 15u) {
              const uint32_t z = act[0xF0];
              do { b.put(z >> 6, z & 63u, lim); run -= 16u; } while (run > 15u);
            }
            const uint32_t a = (x ^ (0u - (x >> 15))) & 0x7fffu;
            const uint32_t s = 32u - (uint32_t)__clz((int)a), e = act[(run << 4) | s];
            b.put((e >> 6) | (x & ((1u << s) - 1u)), e & 63u, lim);
          }
        }
        if (prev != 63) { const uint32_t e = act[0]; b.put(e >> 6, e & 63u, (uint32_t)JM_WORDS); }
      }
      if (b.nacc) {   // left-aligned tail, zeros below
        if (b.wr < (uint32_t)JM_WORDS) b.col[b.wr * JM_MCUS] = (uint32_t)b.acc << (32u - b.nacc); else b.ovf = 1u;
      }
    }
    // lengths -> bit offsets inside the chunk
    uint32_t len = active ? b.wr * 32u + b.nacc : 0u;
    const bool no_room = active && (b.ovf || len > (uint32_t)JM_WORDS * 32u);
    const uint32_t *esc = nullptr;   // where this MCU's code is when it is not in its column
    int hard = 0;
    if (__syncthreads_or(no_room)) {   // rare.  The pixels are still staged: the threads concerned code their MCUs again, into HBM
      if (no_room) {
        const unsigned long long slot = atomicAdd(overflow + 1, 1ull);
        if (slot < JM_ESC_SLOTS) {
          uint32_t *e = esc_pool + slot * JM_ESC_WORDS;
          len = jm_mcu_slow(sm, q, t, blockIdx.x != 0u || t != 0u, e);
          esc = e;
        } else {
          hard = 1;
        }
      }
      hard = __syncthreads_or(hard);
    }
    uint32_t incl = len;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += v; }
    if (lane == 31u) sm.warp_tot[warp] = incl;
    __syncthreads();
    uint32_t off = incl - len, total = 0;
#pragma unroll
    for (uint32_t k = 0; k < (uint32_t)JM_MCUS / 32u; k++) { const uint32_t w = sm.warp_tot[k]; if (k < warp) off += w; total += w; }
    const size_t ci = (size_t)mr * gridDim.x + blockIdx.x;
    if (hard || total > JM_CHUNK_BYTES * 8u) {   // uniform over the CTA: the band goes to the other route
      if (t == 0u) { atomicOr(overflow, 1ull); chunk_bits[ci] = 0u; }
      continue;
    }
    const uint32_t nw = (total + 31u) >> 5;
    for (uint32_t i = t; i <= nw; i += JM_MCUS) cw[i] = 0u;
    __syncthreads();
    for (uint32_t j = 0; j * 32u < len; j++) {
      const uint32_t w = esc ? esc[j] : sm.stage[j][t], p = off + j * 32u, sh = p & 31u;
      atomicOr(&cw[p >> 5], w >> sh);
      if (sh) atomicOr(&cw[(p >> 5) + 1u], w << (32u - sh));
    }
    __syncthreads();
    uint32_t *dst = chunk_words + ci * JM_CHUNK_WORDS;
    for (uint32_t i = t; i < nw; i += JM_MCUS) dst[i] = __byte_perm(cw[i], 0, 0x0123);   // stream byte j = byte j of memory
    if (t == 0u) chunk_bits[ci] = total;
  }
}

// ---- M2 / M3: the chunks of an MCU row, read as one bit stream -------------------------------------------
static constexpr uint32_t JM_MAX_CHUNKS = 64;   // 8191 MCUs at most (16-bit SOF0 width)

// Bit offsets of the row's chunks (exclusive scan of their lengths) into off[0 .. n]; warp 0 does it.
__device__ __forceinline__ void jm_row_offsets(const uint32_t *bits, uint32_t n, uint32_t *off) {
  if (threadIdx.x < 32u) {
    const uint32_t lane = threadIdx.x;
    const uint32_t a = lane < n ? bits[lane] : 0u, c = lane + 32u < n ? bits[lane + 32u] : 0u;
    uint32_t ia = a, ic = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t va = __shfl_up_sync(0xffffffffu, ia, o), vc = __shfl_up_sync(0xffffffffu, ic, o);
      if (lane >= (uint32_t)o) { ia += va; ic += vc; }
    }
    const uint32_t first = __shfl_sync(0xffffffffu, ia, 31);
    off[lane] = ia - a;
    off[lane + 32u] = first + ic - c;
    if (lane == 31u) off[64] = first + ic;
  }
}

// The row's stream read 32 bits at a time from a bit position that is a multiple of 32, MSB first; zero past
// `total`, then the 1-bits that pad the segment to a byte boundary.
struct JmRowReader {
  const uint32_t *row_words, *off;
  uint32_t n, total, c, pos;
  __device__ __forceinline__ void seek(uint32_t b) {
    uint32_t lo = 0, hi = n;   // largest c with off[c] <= b
    while (hi - lo > 1u) { const uint32_t mid = (lo + hi) >> 1; if (off[mid] <= b) lo = mid; else hi = mid; }
    c = lo; pos = b;
  }
  __device__ __forceinline__ uint32_t next() {
    const uint32_t b = pos;
    uint32_t out = 0, got = 0;
    while (got < 32u && b + got < total) {
      const uint32_t lb = b + got - off[c], avail = off[c + 1u] - off[c] - lb;
      if (avail == 0u) { c++; continue; }
      const uint32_t take = min(32u - got, avail), sh = lb & 31u;
      const uint32_t *w = row_words + (size_t)c * JM_CHUNK_WORDS + (lb >> 5);
      const uint32_t w0 = __byte_perm(__ldg(w), 0, 0x0123), w1 = sh + take > 32u ? __byte_perm(__ldg(w + 1), 0, 0x0123) : 0u;
      uint32_t v = __funnelshift_l(w1, w0, sh);
      if (take < 32u) v &= ~(0xffffffffu >> take);
      out |= v >> got;
      got += take;
      if (take == avail) c++;
    }
    if (b < total && total - b < 32u && (total & 7u)) {
      const uint32_t from = total - b, to = ((total + 7u) & ~7u) - b;   // 1 <= to <= 32
      out |= (0xffffffffu >> from) & (to == 32u ? 0xffffffffu : ~(0xffffffffu >> to));
    }
    pos = b + 32u;
    return out;
  }
  // the next 128 bits: five loads and four funnel shifts where they lie inside one chunk (nearly always)
  __device__ __forceinline__ void next4(uint32_t w[4]) {
    const uint32_t lb = pos - off[c];
    if (pos + 128u <= total && lb + 128u <= off[c + 1u] - off[c]) {
      const uint32_t *p = row_words + (size_t)c * JM_CHUNK_WORDS + (lb >> 5);
      const uint32_t sh = lb & 31u;
      uint32_t v[5];
#pragma unroll
      for (int j = 0; j < 5; j++) v[j] = __byte_perm(__ldg(p + j), 0, 0x0123);   // p[4] is inside the slot (or the 16 spare bytes behind the last one); unused when sh == 0
#pragma unroll
      for (int j = 0; j < 4; j++) w[j] = __funnelshift_l(v[j + 1], v[j], sh);
      pos += 128u;
    } else {
#pragma unroll
      for (int j = 0; j < 4; j++) w[j] = next();
    }
  }
};

static constexpr uint32_t JM_MAX_ROUNDS = JM_MAX_CHUNKS * JM_CHUNK_BYTES / 4096u;   // rounds of 4096 stream bytes in a row

// M2: per MCU row the size of its segment, and the 0xFF bytes of every round of 4096 bytes of it (so that M3 can do
// the rounds independently of each other).
__global__ void __launch_bounds__(256) jpeg_rowinfo_kernel(const uint32_t *chunk_words, const uint32_t *chunk_bits, uint32_t n_chunks,
                                                           uint32_t n_rows, uint32_t *row_bits, uint32_t *row_ff, uint32_t *round_ff) {
  __shared__ uint32_t off[JM_MAX_CHUNKS + 1];
  __shared__ uint32_t s_round[JM_MAX_ROUNDS];
  __shared__ uint32_t s_tot[4];
  for (uint32_t r = blockIdx.x; r < n_rows; r += gridDim.x) {
    __syncthreads();
    if (threadIdx.x < JM_MAX_ROUNDS) s_round[threadIdx.x] = 0;
    jm_row_offsets(chunk_bits + (size_t)r * n_chunks, n_chunks, off);
    __syncthreads();
    const uint32_t total = off[n_chunks], n_rounds = (total + 32767u) >> 15;
    const uint32_t *rw = chunk_words + (size_t)r * n_chunks * JM_CHUNK_WORDS;
    JmRowReader rd;
    rd.row_words = rw; rd.off = off; rd.n = n_chunks; rd.total = total;
    for (uint32_t k = 0; k < n_rounds; k++) {   // 16 bytes of the stream a thread
      const uint32_t i = k * 1024u + threadIdx.x * 4u;
      uint32_t c = 0;
      if (i * 32u < total) {
        rd.seek(i * 32u);
        uint32_t w[4];
        rd.next4(w);
#pragma unroll
        for (int j = 0; j < 4; j++) c += (uint32_t)__popc(__vcmpeq4(w[j], 0xffffffffu)) >> 3;
      }
      c = __reduce_add_sync(0xffffffffu, c);
      if ((threadIdx.x & 31u) == 0 && c) atomicAdd(&s_round[k], c);
    }
    __syncthreads();
    uint32_t v = 0;
    if (threadIdx.x < JM_MAX_ROUNDS) { v = s_round[threadIdx.x]; if (threadIdx.x < n_rounds) round_ff[(size_t)r * JM_MAX_ROUNDS + threadIdx.x] = v; }
    if (threadIdx.x < 128u) {   // JM_MAX_ROUNDS == 128: four warps hold the counts
      v = __reduce_add_sync(0xffffffffu, v);
      if ((threadIdx.x & 31u) == 0) s_tot[threadIdx.x >> 5] = v;
    }
    __syncthreads();
    if (threadIdx.x == 0) { row_bits[r] = total; row_ff[r] = s_tot[0] + s_tot[1] + s_tot[2] + s_tot[3]; }
  }
}
static_assert(JM_MAX_ROUNDS == 128u, "jpeg_rowinfo_kernel adds the round counts up with four warps");

// M3: a CTA takes rounds of 4096 stream bytes of an MCU row (blockIdx.y, then every gridDim.y-th): funnels the chunks
// together, stuffs, writes.  The 0xFF bytes of the rounds before come from M2, so rounds do not wait for each other.
__global__ void __launch_bounds__(256) jpeg_stuff_chunks_kernel(const uint32_t *chunk_words, const uint32_t *chunk_bits, uint32_t n_chunks,
                                                                const unsigned long long *out_base, const uint32_t *round_ff, uint32_t n_rows,
                                                                uint32_t first_mrow, uint32_t total_mrows, uint8_t *out) {
  __shared__ uint32_t off[JM_MAX_CHUNKS + 1];
  __shared__ uint32_t warp_tot[8];
  __shared__ uint32_t s_before[4];
  __shared__ __align__(16) uint8_t sbuf[16 + 8192 + 16];   // one round's stuffed bytes, placed so that shared and global addresses agree mod 16
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  for (uint32_t r = blockIdx.x; r < n_rows; r += gridDim.x) {
    __syncthreads();
    jm_row_offsets(chunk_bits + (size_t)r * n_chunks, n_chunks, off);
    __syncthreads();
    const uint32_t total = off[n_chunks], n = (total + 7u) >> 3, n_rounds = (total + 32767u) >> 15;
    const uint32_t *rw = chunk_words + (size_t)r * n_chunks * JM_CHUNK_WORDS;
    uint8_t *o = out + out_base[r];
    JmRowReader rd;
    rd.row_words = rw; rd.off = off; rd.n = n_chunks; rd.total = total;
    for (uint32_t k = blockIdx.y; k < n_rounds; k += gridDim.y) {
      const uint32_t i0 = k * 1024u, i = i0 + threadIdx.x * 4u;   // the thread's 4 words: 16 bytes of the stream
      uint32_t w[4] = {0u, 0u, 0u, 0u}, cnt = 0;
      if (i * 32u < total) {
        rd.seek(i * 32u);
        rd.next4(w);
#pragma unroll
        for (int j = 0; j < 4; j++) cnt += (uint32_t)__popc(__vcmpeq4(w[j], 0xffffffffu)) >> 3;   // bytes past the end are 0
      }
      uint32_t incl = cnt;
#pragma unroll
      for (int q = 1; q < 32; q <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, incl, q); if (lane >= (uint32_t)q) incl += v; }
      if (lane == 31u) warp_tot[warp] = incl;
      if (threadIdx.x < 128u) {   // 0xFF bytes of the rounds before this one
        uint32_t v = threadIdx.x < k ? round_ff[(size_t)r * JM_MAX_ROUNDS + threadIdx.x] : 0u;
        v = __reduce_add_sync(0xffffffffu, v);
        if (lane == 0u) s_before[warp] = v;
      }
      __syncthreads();
      uint32_t wbase = 0, round_ff_here = 0;
#pragma unroll
      for (int q = 0; q < 8; q++) { const uint32_t v = warp_tot[q]; if ((uint32_t)q < warp) wbase += v; round_ff_here += v; }
      const uint32_t before = s_before[0] + s_before[1] + s_before[2] + s_before[3];
      uint8_t *dst = o + (size_t)i0 * 4u + before;                      // where the round's bytes go
      const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 15u);
      const uint32_t round_bytes = min(4096u, n - i0 * 4u) + round_ff_here;
      uint32_t pos = mis + (i - i0) * 4u + wbase + incl - cnt;
      if (i * 4u + 16u <= n) {
        // The thread's 16 bytes as they are, at its place: words where the buffer allows, bytes at the two ends.  A
        // thread with 0xFF bytes owns cnt more bytes behind them and puts things right below.
        uint32_t v[4];
#pragma unroll
        for (int j = 0; j < 4; j++) v[j] = __byte_perm(w[j], 0, 0x0123);   // memory order
        const uint32_t a = pos & 3u;
        if (a == 0u) {
#pragma unroll
          for (int j = 0; j < 4; j++) *reinterpret_cast<uint32_t *>(&sbuf[pos + 4u * j]) = v[j];
        } else {
          const uint32_t head = 4u - a, sh = head * 8u;   // head bytes of v[0], then three whole words, then a bytes of v[3]
#pragma unroll
          for (uint32_t bb = 0; bb < 3u; bb++) if (bb < head) sbuf[pos + bb] = (uint8_t)(v[0] >> (8u * bb));
#pragma unroll
          for (int j = 0; j < 3; j++) *reinterpret_cast<uint32_t *>(&sbuf[pos + head + 4u * j]) = __funnelshift_r(v[j], v[j + 1], sh);
#pragma unroll
          for (uint32_t bb = 0; bb < 3u; bb++) if (bb < a) sbuf[pos + 16u - a + bb] = (uint8_t)(v[3] >> (sh + 8u * bb));
        }
        if (cnt) {   // from the first 0xFF on: a zero behind each, everything after it moves up
          uint32_t zc = 0;
#pragma unroll 1
          for (uint32_t j = 0; j < 4u; j++) {
            const uint32_t wj = j == 0u ? w[0] : j == 1u ? w[1] : j == 2u ? w[2] : w[3];
            if (__vcmpeq4(wj, 0xffffffffu) == 0u) {
              if (zc) {
#pragma unroll
                for (uint32_t bb = 0; bb < 4u; bb++) sbuf[pos + 4u * j + bb + zc] = (uint8_t)(wj >> (24u - 8u * bb));
              }
            } else {
#pragma unroll
              for (uint32_t bb = 0; bb < 4u; bb++) {
                const uint32_t byte = (wj >> (24u - 8u * bb)) & 0xffu;
                sbuf[pos + 4u * j + bb + zc] = (uint8_t)byte;
                if (byte == 0xffu) { zc++; sbuf[pos + 4u * j + bb + zc] = 0; }
              }
            }
          }
        }
      } else if (i * 4u < n) {   // the row's last bytes
#pragma unroll 1
        for (uint32_t jb = 0; jb < 16u && i * 4u + jb < n; jb++) {
          const uint32_t wj = (jb >> 2) == 0u ? w[0] : (jb >> 2) == 1u ? w[1] : (jb >> 2) == 2u ? w[2] : w[3];
          const uint32_t byte = (wj >> (24u - 8u * (jb & 3u))) & 0xffu;
          sbuf[pos++] = (uint8_t)byte;
          if (byte == 0xffu) sbuf[pos++] = 0;
        }
      }
      __syncthreads();
      // copy out: 16 bytes a thread where the piece lies wholly inside the round, single bytes at its two ends
      for (uint32_t q = threadIdx.x * 16u; q < mis + round_bytes; q += 4096u) {
        if (q >= mis && q + 16u <= mis + round_bytes) {
          *reinterpret_cast<uint4 *>(dst - mis + q) = *reinterpret_cast<const uint4 *>(&sbuf[q]);
        } else {
          for (uint32_t e = max(q, mis); e < min(q + 16u, mis + round_bytes); e++) dst[e - mis] = sbuf[e];
        }
      }
      if (k + 1u == n_rounds && threadIdx.x == 0 && first_mrow + r + 1u < total_mrows) {   // the last round ends the segment: RSTm
        dst[round_bytes] = 0xFF;
        dst[round_bytes + 1u] = (uint8_t)(0xD0u + ((first_mrow + r) & 7u));
      }
      __syncthreads();
    }
  }
}

int upload_tables(stitch_job *job) {
  stitch_ctx *ctx = job->ctx;
  const int quality = (int)job->desc.jpeg_quality;
  uint16_t ql[64], qc[64];
  quant_tables(quality, ql, qc);
  uint32_t q8[2][64], rc[2][64];
  for (int k = 0; k < 64; k++) {
    q8[0][k] = (uint32_t)ql[k_zigzag[k]] << 3;
    q8[1][k] = (uint32_t)qc[k_zigzag[k]] << 3;
    for (int c = 0; c < 2; c++) rc[c][k] = (uint32_t)(((1ull << 32) + q8[c][k] - 1) / q8[c][k]);
  }
  uint32_t dc[2][16], ac[2][256];
  build_huff(k_dc_lum_bits, k_dc_vals, dc[0], 16); build_huff(k_dc_chr_bits, k_dc_vals, dc[1], 16);
  build_huff(k_ac_lum_bits, k_ac_lum_vals, ac[0], 256); build_huff(k_ac_chr_bits, k_ac_chr_vals, ac[1], 256);
  // constant memory is per device and shared by every job of this process: uploads are stream-ordered
  STITCH_CK(cudaMemcpyToSymbolAsync(c_q8, q8, sizeof q8, 0, cudaMemcpyHostToDevice, ctx->stream));
  STITCH_CK(cudaMemcpyToSymbolAsync(c_recip, rc, sizeof rc, 0, cudaMemcpyHostToDevice, ctx->stream));
  STITCH_CK(cudaMemcpyToSymbolAsync(c_dc_tab, dc, sizeof dc, 0, cudaMemcpyHostToDevice, ctx->stream));
  STITCH_CK(cudaMemcpyToSymbolAsync(c_ac_tab, ac, sizeof ac, 0, cudaMemcpyHostToDevice, ctx->stream));
  STITCH_CK(cudaStreamSynchronize(ctx->stream));  // the host arrays above are on the stack
  return STITCH_OK;
}

}  // namespace

// Is the band [y0, y1) an exact mosaic of RGBA8 tiles the MCU kernel can read in place?  If so the groups and
// their tiles go to the device (scratch [3], which only the J1..J7 route uses otherwise) and *dir describes them.
static int jpeg_direct_plan(stitch_job *job, uint32_t y0, uint32_t y1, JmDirect *dir, bool *ok) {
  stitch_ctx *ctx = job->ctx;
  const stitch_job_desc &jd = job->desc;
  *ok = false;
  if (jd.alpha_blend || jd.out_bit_depth != 8) return STITCH_OK;
  if (const char *e = getenv("STITCH_JPEG_DIRECT")) if (atoi(e) == 0) return STITCH_OK;
  struct Rect { uint32_t xa, xb, ya, yb; const TileRec *t; };
  std::vector<Rect> rs;
  std::vector<uint32_t> cuts{y0, y1};
  for (const TileRec &t : job->tiles) {
    const stitch_tile_desc &d = t.d;
    if (!d.vis_w || !d.vis_h) continue;
    const uint32_t ya = (uint32_t)d.dst_y, yb = ya + d.vis_h, xa = (uint32_t)d.dst_x, xb = xa + d.vis_w;
    if (yb <= y0 || ya >= y1) continue;
    if (d.color_type != 6 || d.bit_depth != 8 || t.has_pal || !t.d_pix) return STITCH_OK;
    if ((reinterpret_cast<uintptr_t>(t.d_pix) & 3u) || (t.pitch & 3u)) return STITCH_OK;   // pixels are read as words (16 bytes at a time where the rows allow)
    rs.push_back({xa, xb, ya, yb, &t});
    if (ya > y0) cuts.push_back(ya);
    if (yb < y1) cuts.push_back(yb);
    if (rs.size() > 4096) return STITCH_OK;
  }
  // (tiles whose rows are not 16-byte aligned come in by 4-byte copies: tools/jpeg_bench.py 1023 takes 2.45 ms from the
  // tiles and 2.38 ms through the compositor's canvas, against 1.7 ms from aligned tiles -- no reason to prefer the canvas)
  std::sort(cuts.begin(), cuts.end());
  cuts.erase(std::unique(cuts.begin(), cuts.end()), cuts.end());
  if (cuts.size() * rs.size() > ((size_t)4 << 20)) return STITCH_OK;   // the loop below would cost more host time than the compositor saves
  std::vector<JmGroup> groups;
  std::vector<JmSeg> segs;
  std::vector<const Rect *> row;
  for (size_t k = 0; k + 1 < cuts.size(); k++) {
    const uint32_t ra = cuts[k], rb = cuts[k + 1];
    row.clear();
    for (const Rect &r : rs) if (r.ya <= ra && r.yb >= rb) row.push_back(&r);
    std::sort(row.begin(), row.end(), [](const Rect *a, const Rect *b) { return a->xa < b->xa; });
    uint32_t x = 0;
    JmGroup g{ra, rb, (uint32_t)segs.size(), (uint32_t)row.size()};
    for (const Rect *r : row) {
      if (r->xa != x) return STITCH_OK;   // a gap (the background shows) or an overlap
      const stitch_tile_desc &d = r->t->d;
      segs.push_back({r->xa, r->t->pitch, r->t->d_pix + (size_t)(d.src_off_y + (ra - r->ya)) * r->t->pitch + (size_t)d.src_off_x * 4u});
      x = r->xb;
    }
    if (x != jd.canvas_w || segs.size() > 65536) return STITCH_OK;
    groups.push_back(g);
  }
  // a tile that only touches a group partly was dropped from it above and would leave a gap, so reaching here
  // means every row of the band is tiled exactly
  const uint32_t mrows = (y1 - y0 + 7) / 8, n_chunks = ((jd.canvas_w + 7) / 8 + JM_MCUS - 1) / JM_MCUS;
  if (groups.size() * n_chunks > (1u << 20)) return STITCH_OK;
  std::vector<uint32_t> idx(mrows + groups.size() * n_chunks);   // mrow_group | chunk_seg
  for (uint32_t gi = 0, mr = 0; gi < groups.size(); gi++) {
    for (; mr < mrows && y0 + mr * 8u < groups[gi].row1; mr++) idx[mr] = gi;
    uint32_t si = groups[gi].seg0;
    for (uint32_t c = 0; c < n_chunks; c++) {
      const uint32_t xf = c ? c * JM_MCUS * 8u - 8u : 0u;
      while (si + 1 < groups[gi].seg0 + groups[gi].nseg && segs[si + 1].x0 <= xf) si++;
      idx[mrows + (size_t)gi * n_chunks + c] = si;
    }
  }
  const size_t gb = (groups.size() * sizeof(JmGroup) + 15) / 16 * 16, sb = segs.size() * sizeof(JmSeg), need = gb + sb + idx.size() * 4;
  int rc;
  if ((rc = stitch_ensure_buffer(ctx, &job->d_jpeg[3], &job->jpeg_cap[3], need))) return rc;
  STITCH_CK(cudaMemcpyAsync(job->d_jpeg[3], groups.data(), groups.size() * sizeof(JmGroup), cudaMemcpyHostToDevice, ctx->stream));   // pageable: staged before the call returns
  STITCH_CK(cudaMemcpyAsync(job->d_jpeg[3] + gb, segs.data(), sb, cudaMemcpyHostToDevice, ctx->stream));
  STITCH_CK(cudaMemcpyAsync(job->d_jpeg[3] + gb + sb, idx.data(), idx.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
  dir->groups = reinterpret_cast<const JmGroup *>(job->d_jpeg[3]);
  dir->segs = reinterpret_cast<const JmSeg *>(job->d_jpeg[3] + gb);
  dir->mrow_group = reinterpret_cast<const uint32_t *>(job->d_jpeg[3] + gb + sb);
  dir->chunk_seg = dir->mrow_group + mrows;
  *ok = true;
  return STITCH_OK;
}

// The one-pass route over a band.  Scratch: [0] chunk slots, [1] bits per chunk, [2] per-row sizes and totals.
static int encode_band_one_pass(stitch_job *job, const JpegGeom &g, uint32_t first_mrow, uint32_t total_mrows, const uint8_t *canvas,
                                size_t canvas_pitch, const JmDirect &dir, size_t *out_bytes, uint32_t *launches, bool *overflow) {
  stitch_ctx *ctx = job->ctx;
  int rc;
  static_assert(sizeof(JmSmem) <= 56 * 1024, "four CTAs of the MCU kernel per SM");
  const uint32_t n_chunks = (g.mcus + JM_MCUS - 1) / JM_MCUS;
  if (n_chunks > JM_MAX_CHUNKS) { *overflow = true; return STITCH_OK; }
  if (!job->jpeg_huff_ready) {  // Annex K tables: the same for every job and quality, so no job can disturb another's
    uint32_t dc[2][16], ac[2][256];
    build_huff(k_dc_lum_bits, k_dc_vals, dc[0], 16); build_huff(k_dc_chr_bits, k_dc_vals, dc[1], 16);
    build_huff(k_ac_lum_bits, k_ac_lum_vals, ac[0], 256); build_huff(k_ac_chr_bits, k_ac_chr_vals, ac[1], 256);
    STITCH_CK(cudaMemcpyToSymbolAsync(c_dc_tab, dc, sizeof dc, 0, cudaMemcpyHostToDevice, ctx->stream));   // pageable source: staged before the call returns
    STITCH_CK(cudaMemcpyToSymbolAsync(c_ac_tab, ac, sizeof ac, 0, cudaMemcpyHostToDevice, ctx->stream));
    STITCH_CK(cudaFuncSetAttribute(jpeg_mcu_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(JmSmem)));
    job->jpeg_huff_ready = true;
  }
  JpegQuant q;
  {
    uint16_t ql[64], qc[64];
    quant_tables((int)job->desc.jpeg_quality, ql, qc);
    for (int k = 0; k < 64; k++) {
      const uint32_t d0 = (uint32_t)ql[k_zigzag[k]] << 3, d1 = (uint32_t)qc[k_zigzag[k]] << 3;
      q.hr[0][k] = make_uint2(d0 >> 1, (uint32_t)(((1ull << 32) + d0 - 1) / d0));
      q.hr[1][k] = make_uint2(d1 >> 1, (uint32_t)(((1ull << 32) + d1 - 1) / d1));
    }
  }
  const size_t n_slots = (size_t)g.mrows * n_chunks;
  if ((rc = stitch_ensure_buffer(ctx, &job->d_jpeg[0], &job->jpeg_cap[0], n_slots * JM_CHUNK_BYTES + 16))) return rc;
  const size_t bits_bytes = (n_slots * 4 + 15) / 16 * 16;   // bits per chunk, then the slots of MCUs that did not fit their columns
  if ((rc = stitch_ensure_buffer(ctx, &job->d_jpeg[1], &job->jpeg_cap[1], bits_bytes + (size_t)JM_ESC_SLOTS * JM_ESC_WORDS * 4))) return rc;
  const size_t rows_pad = ((size_t)g.mrows + 3) / 4 * 4;
  if ((rc = stitch_ensure_buffer(ctx, &job->d_jpeg[2], &job->jpeg_cap[2], rows_pad * 24 + 64))) return rc;
  if (!job->h_jpeg_totals) STITCH_CK(cudaHostAlloc((void **)&job->h_jpeg_totals, 32, cudaHostAllocDefault));
  uint32_t *chunk_words = reinterpret_cast<uint32_t *>(job->d_jpeg[0]);
  uint32_t *chunk_bits = reinterpret_cast<uint32_t *>(job->d_jpeg[1]);
  if ((rc = stitch_ensure_buffer(ctx, &job->d_jpeg[4], &job->jpeg_cap[4], (size_t)g.mrows * JM_MAX_ROUNDS * 4))) return rc;
  uint32_t *round_ff = reinterpret_cast<uint32_t *>(job->d_jpeg[4]);
  uint32_t *row_bits = reinterpret_cast<uint32_t *>(job->d_jpeg[2]);
  uint32_t *row_ff = row_bits + rows_pad;
  unsigned long long *row_base = reinterpret_cast<unsigned long long *>(row_ff + rows_pad);
  unsigned long long *out_base = row_base + rows_pad;
  unsigned long long *totals = out_base + rows_pad;   // [1] output bytes, [2] overflow flag, [3] MCUs coded into the slots
  STITCH_CK(cudaMemsetAsync(totals, 0, 32, ctx->stream));
  {
    dim3 grid(n_chunks, g.mrows < 16384u ? g.mrows : 16384u);
    jpeg_mcu_kernel<<<grid, JM_MCUS, sizeof(JmSmem), ctx->stream>>>(canvas, canvas_pitch, g, q, dir, chunk_words, chunk_bits,
                                                                    reinterpret_cast<uint32_t *>(job->d_jpeg[1] + bits_bytes), totals + 2);
    STITCH_CK(cudaGetLastError());
  }
  const uint32_t row_grid = g.mrows < 148u * 8u ? g.mrows : 148u * 8u;
  jpeg_rowinfo_kernel<<<row_grid, 256, 0, ctx->stream>>>(chunk_words, chunk_bits, n_chunks, g.mrows, row_bits, row_ff, round_ff);
  STITCH_CK(cudaGetLastError());
  jpeg_rowscan_kernel<<<1, 1024, 0, ctx->stream>>>(row_bits, row_ff, g.mrows, first_mrow, total_mrows, 1, out_base, totals);
  STITCH_CK(cudaGetLastError());
  STITCH_CK(cudaMemcpyAsync(job->h_jpeg_totals, totals, 32, cudaMemcpyDeviceToHost, ctx->stream));
  STITCH_CK(cudaStreamSynchronize(ctx->stream));  // the size of the band's bytes decides the size of the output buffer
  if (launches) *launches += 3;
  if (getenv("STITCH_JPEG_DEBUG"))
    fprintf(stderr, "[stitch jpeg] band of %u MCU rows x %u MCUs: %llu MCUs coded into slots, overflow %llu, %s\n", g.mrows, g.mcus,
            job->h_jpeg_totals[3], job->h_jpeg_totals[2], dir.groups ? "tiles read in place" : "from a canvas");
  if (job->h_jpeg_totals[2]) { *overflow = true; return STITCH_OK; }
  const size_t n_out = (size_t)job->h_jpeg_totals[1];
  if ((rc = stitch_ensure_buffer(ctx, &job->d_out, &job->out_bytes, n_out + 64))) return rc;
  {  // rounds of 4096 bytes are independent: enough CTAs per row that the device is full whatever the band's height
    const uint32_t max_rounds = (n_chunks * JM_CHUNK_BYTES + 4095u) / 4096u;
    uint32_t per_row = (4u * 148u * 8u + g.mrows - 1u) / g.mrows;
    per_row = per_row < 1u ? 1u : per_row > max_rounds ? max_rounds : per_row > 8u ? 8u : per_row;
    dim3 grid(g.mrows < 16384u ? g.mrows : 16384u, per_row);
    jpeg_stuff_chunks_kernel<<<grid, 256, 0, ctx->stream>>>(chunk_words, chunk_bits, n_chunks, out_base, round_ff, g.mrows, first_mrow, total_mrows, job->d_out);
  }
  STITCH_CK(cudaGetLastError());
  if (launches) *launches += 1;
  *out_bytes = n_out;
  return STITCH_OK;
}

int stitch_jpeg_encode_band(stitch_job *job, uint32_t y0, uint32_t y1, const uint8_t *canvas, size_t canvas_pitch,
                            size_t *out_bytes, uint32_t *launches, bool *need_canvas) {
  stitch_ctx *ctx = job->ctx;
  const stitch_job_desc &jd = job->desc;
  int rc;
  JpegGeom g;
  g.width = jd.canvas_w; g.height = jd.canvas_h; g.y0 = y0; g.mcus = (jd.canvas_w + 7) / 8; g.mrows = (y1 - y0 + 7) / 8;
  g.rows_in_band = y1 - y0;
  const uint32_t total_mrows = (jd.canvas_h + 7) / 8, first_mrow = y0 / 8;
  // STITCH_JPEG_ONE_PASS: 0 = the J1..J7 route only, 2 = the one-pass route only (content it cannot hold is an
  // error: tests), otherwise one pass first and J1..J7 for a band that overflowed
  int one_pass = 1;
  if (const char *e = getenv("STITCH_JPEG_ONE_PASS")) one_pass = atoi(e);
  if (need_canvas) *need_canvas = false;
  // A band that overflowed the one-pass route has cost its MCU kernel for nothing, and busy content comes in runs:
  // after an overflow the next 1, 2, 4, 8 bands (doubling while it keeps happening) go straight to J1..J7.
  const bool backing_off = one_pass == 1 && job->jpeg_skip_bands > 0;
  if (!canvas) {   // first try: the tiles themselves, if the band is an exact mosaic of them
    JmDirect dir{};
    bool ok = false;
    if (one_pass && !backing_off && (rc = jpeg_direct_plan(job, y0, y1, &dir, &ok))) return rc;
    bool overflow = false;
    if (ok && (rc = encode_band_one_pass(job, g, first_mrow, total_mrows, nullptr, 0, dir, out_bytes, launches, &overflow))) return rc;
    if (!ok || overflow) *need_canvas = true;
    job->jpeg_overflowed = overflow;   // the canvas would overflow the same chunks: the second call goes straight to J1..J7
    if (ok && !overflow) job->jpeg_overflow_streak = 0;
    return STITCH_OK;
  }
  const bool known_overflow = job->jpeg_overflowed || backing_off;
  if (backing_off) job->jpeg_skip_bands--;
  if (job->jpeg_overflowed) {
    job->jpeg_overflow_streak = job->jpeg_overflow_streak < 3u ? job->jpeg_overflow_streak + 1u : 3u;
    job->jpeg_skip_bands = 1u << job->jpeg_overflow_streak >> 1;   // 1, 2, 4
  }
  job->jpeg_overflowed = false;
  if (known_overflow && one_pass == 2)
    return stitch_fail(ctx, STITCH_E_CAPACITY, "JPEG band [%u,%u): more than %d bits in an MCU or %u bytes in a chunk of %d MCUs (STITCH_JPEG_ONE_PASS=2)",
                       y0, y1, JM_WORDS * 32, JM_CHUNK_BYTES, JM_MCUS);
  if (one_pass && !known_overflow) {
    bool overflow = false;
    if ((rc = encode_band_one_pass(job, g, first_mrow, total_mrows, canvas, canvas_pitch, JmDirect{}, out_bytes, launches, &overflow))) return rc;
    if (!overflow) { job->jpeg_overflow_streak = 0; return STITCH_OK; }
    job->jpeg_overflow_streak = job->jpeg_overflow_streak < 3u ? job->jpeg_overflow_streak + 1u : 3u;
    job->jpeg_skip_bands = 1u << job->jpeg_overflow_streak >> 1;
    if (one_pass == 2) {
      return stitch_fail(ctx, STITCH_E_CAPACITY, "JPEG band [%u,%u): more than %d bits in an MCU or %u bytes in a chunk of %d MCUs (STITCH_JPEG_ONE_PASS=2)",
                         y0, y1, JM_WORDS * 32, JM_CHUNK_BYTES, JM_MCUS);
    }
  }
  // tables are re-uploaded per band: another job (another quality) may have run on this device in between
  if ((rc = upload_tables(job))) return rc;
  const size_t n_blk = (size_t)g.mrows * g.mcus * 3;
  // scratch: [0] coefficients, [1] bit offsets, [2] per-row: bits | ff | base(u64) | out_base(u64) | totals(u64 x2), [3] unstuffed
  if ((rc = stitch_ensure_buffer(ctx, &job->d_jpeg[0], &job->jpeg_cap[0], n_blk * 128))) return rc;
  if ((rc = stitch_ensure_buffer(ctx, &job->d_jpeg[1], &job->jpeg_cap[1], n_blk * 4))) return rc;
  const size_t rows_pad = ((size_t)g.mrows + 3) / 4 * 4;
  if ((rc = stitch_ensure_buffer(ctx, &job->d_jpeg[2], &job->jpeg_cap[2], rows_pad * 24 + 64))) return rc;
  if (!job->h_jpeg_totals) STITCH_CK(cudaHostAlloc((void **)&job->h_jpeg_totals, 32, cudaHostAllocDefault));
  if ((rc = stitch_ensure_buffer(ctx, &job->d_jpeg[4], &job->jpeg_cap[4], n_blk * 8))) return rc;
  unsigned long long *nzmask = reinterpret_cast<unsigned long long *>(job->d_jpeg[4]);
  int16_t *coefs = reinterpret_cast<int16_t *>(job->d_jpeg[0]);
  uint32_t *bit_off = reinterpret_cast<uint32_t *>(job->d_jpeg[1]);
  uint32_t *row_bits = reinterpret_cast<uint32_t *>(job->d_jpeg[2]);
  uint32_t *row_ff = row_bits + rows_pad;
  unsigned long long *row_base = reinterpret_cast<unsigned long long *>(row_ff + rows_pad);
  unsigned long long *out_base = row_base + rows_pad;
  unsigned long long *totals = out_base + rows_pad;

  {
    dim3 grid((g.mcus + 127u) / 128u, g.mrows < 4096u ? g.mrows : 4096u);
    jpeg_dct_kernel<<<grid, JD_MCUS, 0, ctx->stream>>>(canvas, canvas_pitch, g, coefs, nzmask);
    STITCH_CK(cudaGetLastError());
  }
  const uint32_t row_grid = g.mrows < 148u * 8u ? g.mrows : 148u * 8u;
  jpeg_len_kernel<<<row_grid, 256, 0, ctx->stream>>>(coefs, nzmask, g, bit_off, row_bits);
  STITCH_CK(cudaGetLastError());
  jpeg_rowscan_kernel<<<1, 1024, 0, ctx->stream>>>(row_bits, row_ff, g.mrows, first_mrow, total_mrows, 0, row_base, totals);
  STITCH_CK(cudaGetLastError());
  STITCH_CK(cudaMemcpyAsync(job->h_jpeg_totals, totals, 8, cudaMemcpyDeviceToHost, ctx->stream));
  STITCH_CK(cudaStreamSynchronize(ctx->stream));  // the unstuffed size decides the scratch and output sizes
  const size_t ubytes = (size_t)job->h_jpeg_totals[0];
  const size_t uwords = (ubytes + 3) / 4 + 2;
  if ((rc = stitch_ensure_buffer(ctx, &job->d_jpeg[3], &job->jpeg_cap[3], uwords * 4))) return rc;
  if ((rc = stitch_ensure_buffer(ctx, &job->d_out, &job->out_bytes, 2 * ubytes + 2 * (size_t)g.mrows + 64))) return rc;
  { int frc = stitch_fill_words(ctx, job->d_jpeg[3], 0u, (size_t)uwords * 4); if (frc) return frc; }
  {
    const size_t want = (n_blk + 255) / 256;
    const uint32_t grid = (uint32_t)(want < 148u * 16u ? want : 148u * 16u);
    jpeg_emit_kernel<<<grid, 256, 0, ctx->stream>>>(coefs, nzmask, g, bit_off, row_bits, row_base, reinterpret_cast<uint32_t *>(job->d_jpeg[3]));
    STITCH_CK(cudaGetLastError());
  }
  jpeg_ffcount_kernel<<<row_grid, 256, 0, ctx->stream>>>(job->d_jpeg[3], row_bits, row_base, g.mrows, row_ff);
  STITCH_CK(cudaGetLastError());
  jpeg_rowscan_kernel<<<1, 1024, 0, ctx->stream>>>(row_bits, row_ff, g.mrows, first_mrow, total_mrows, 1, out_base, totals);
  STITCH_CK(cudaGetLastError());
  jpeg_stuff_kernel<<<row_grid, 256, 0, ctx->stream>>>(job->d_jpeg[3], row_bits, row_base, out_base, g.mrows, first_mrow, total_mrows, job->d_out);
  STITCH_CK(cudaGetLastError());
  STITCH_CK(cudaMemcpyAsync(job->h_jpeg_totals + 1, totals + 1, 8, cudaMemcpyDeviceToHost, ctx->stream));
  STITCH_CK(cudaStreamSynchronize(ctx->stream));
  *out_bytes = (size_t)job->h_jpeg_totals[1];
  if (launches) *launches += 7;
  return STITCH_OK;
}

// ---- container helper: what the external encoder emits in front of the first strip -----------------------
static size_t put_seg(uint8_t *o, size_t n, size_t cap, uint8_t marker, const uint8_t *payload, size_t len) {
  if (n + 4 + len <= cap) {
    o[n] = 0xFF; o[n + 1] = marker; o[n + 2] = (uint8_t)((len + 2) >> 8); o[n + 3] = (uint8_t)((len + 2) & 0xFF);
    memcpy(o + n + 4, payload, len);
  }
  return n + 4 + len;
}

extern "C" int stitch_jpeg_header(uint32_t width, uint32_t height, uint32_t quality, uint8_t *out, size_t cap, size_t *out_len) {
  if (!out || !out_len) return STITCH_E_INVALID;
  *out_len = 0;
  // JPEG stores dimensions in 16 bits; the reference only checks > 0 (src/jpeg-encoder.ts:109-111)
  if (!width || !height || width > 65535u || height > 65535u || quality < 1 || quality > 100) return STITCH_E_INVALID;
  uint16_t ql[64], qc[64];
  quant_tables((int)quality, ql, qc);
  uint8_t buf[512];
  size_t n = 2;
  if (cap >= 2) { out[0] = 0xFF; out[1] = 0xD8; }
  const uint8_t app0[14] = {'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0};
  n = put_seg(out, n, cap, 0xE0, app0, 14);
  for (int t = 0; t < 2; t++) {
    buf[0] = (uint8_t)t;
    for (int k = 0; k < 64; k++) buf[1 + k] = (uint8_t)(t == 0 ? ql : qc)[k_zigzag[k]];
    n = put_seg(out, n, cap, 0xDB, buf, 65);
  }
  const uint8_t sof[15] = {8, (uint8_t)(height >> 8), (uint8_t)height, (uint8_t)(width >> 8), (uint8_t)width, 3,
                           1, 0x11, 0, 2, 0x11, 1, 3, 0x11, 1};
  n = put_seg(out, n, cap, 0xC0, sof, 15);
  const uint8_t *bits[4] = {k_dc_lum_bits, k_ac_lum_bits, k_dc_chr_bits, k_ac_chr_bits};
  const uint8_t *vals[4] = {k_dc_vals, k_ac_lum_vals, k_dc_vals, k_ac_chr_vals};
  const uint8_t cls[4] = {0x00, 0x10, 0x01, 0x11};
  for (int t = 0; t < 4; t++) {
    int cnt = 0;
    for (int i = 0; i < 16; i++) cnt += bits[t][i];
    buf[0] = cls[t];
    memcpy(buf + 1, bits[t], 16);
    memcpy(buf + 17, vals[t], (size_t)cnt);
    n = put_seg(out, n, cap, 0xC4, buf, 17 + (size_t)cnt);
  }
  const uint32_t ri = (width + 7) / 8;  // restart interval: one MCU row
  const uint8_t dri[2] = {(uint8_t)(ri >> 8), (uint8_t)ri};
  n = put_seg(out, n, cap, 0xDD, dri, 2);
  const uint8_t sos[10] = {3, 1, 0x00, 2, 0x11, 3, 0x11, 0, 63, 0};
  n = put_seg(out, n, cap, 0xDA, sos, 10);
  if (n > cap) return STITCH_E_CAPACITY;
  *out_len = n;
  return STITCH_OK;
}
