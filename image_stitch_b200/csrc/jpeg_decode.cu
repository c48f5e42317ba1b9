// jpeg_decode.cu -- SURVEY.md 8(f).4: baseline JPEG tiles are decoded on the device, so that they cross the bus
// compressed and enter the compositor as the RGBA8 (or gray8) rows the reference's JPEG decoder yields
// (src/decoders/jpeg-decoder.ts:284-347: whole-image decode, then rows; on Node the pixels are sharp's, i.e.
// libjpeg-turbo's default path: Huffman decoding, jpeg_idct_islow, fancy chroma upsampling, fixed-point YCbCr -> RGB).
// The arithmetic follows that published algorithm bit for bit; oracle/jpeg_decode_oracle.c is its CPU restatement,
// pinned against libjpeg-turbo 3.1.4.1 (tests/golden/jpeg_decode), and tests/test_gpu_jpeg_decode.py compares the two.
//
//   host      marker segments (SOF0/SOF1, DQT, DHT, DRI, SOS) are parsed and the entropy-coded segment is copied into
//             pinned staging memory without its framing (the zero stuffed after every 0xFF data byte, the RSTn markers
//             between restart intervals) -- container work, like cutting a PNG into chunks; no code is decoded
//   entropy   speculative parallel Huffman walk inside every restart interval (see "entropy decoding" below):
//             subsequences of 1024 bits, one thread each, relaxation rounds in one cooperative kernel until the
//             hand-over states settle, a prefix sum for block positions, a last walk that writes the coefficients
//             de-zigzagged, and a prefix sum for the DC predictions; 11-bit look-ahead tables in shared memory.  One
//             thread per restart interval decodes the file only if the walk has not settled after 32 rounds
//   IDCT      one thread per 8x8 block: dequantise + jpeg_idct_islow, samples into the component's plane
//   colour    one thread per 4 output pixels: h2v1 / h2v2 fancy upsampling (replication when the downsampled width is
//             <= 2), context rows replicated at the image edges, YCbCr -> RGBA8; gray is copied
//
// Everything runs on one of four decode streams of the context (taken in turn, so that the small kernels of several
// tiles overlap) right behind the file's H2D copy; the tile's upload event is recorded there and covers it.
#include "job.h"
#include "kernels.h"

#include <algorithm>
#include <cstdlib>
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

constexpr int LOOK = 11;           // look-ahead bits of the Huffman tables (longer codes are rare: out-of-line canonical walk)
constexpr uint32_t SUB_BITS = 1024;  // bits per subsequence of the speculative walk
constexpr int SYNC_ROUNDS = 32;      // relaxation rounds before the serial decoder takes over (a settled walk skips the rest)
constexpr uint32_t BAD = 0xffffffffu;
constexpr uint32_t DC_CHUNK = 8;     // MCUs per chunk of the DC prefix sum

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
// One subsequence of the entropy-coded data: the symbols that START in [start_bit, end_bit) of its restart interval.
struct JdSub {
  uint32_t start_bit, end_bit;
  uint32_t interval;
  uint32_t first;             // 1: the interval's first subsequence -- its start state is exact
};
struct JdState {
  uint32_t bp;                // bit position of the next symbol (BAD: the walk met an impossible code)
  uint32_t pos;               // block within the MCU << 8 | zig-zag index of the next coefficient (0: the DC code is next)
};
struct JdImage {
  const uint32_t *ecs;        // entropy-coded data with byte stuffing and restart markers removed (host), big-endian bit order
  uint32_t width, height, ncomp, hmax, vmax, mcus_x, mcus_y;
  uint32_t restart;           // MCUs per restart interval (the whole scan when the file has none)
  uint32_t n_intervals, n_subs, bpm;   // bpm: blocks per MCU
  uint32_t mcu_comp[6], mcu_bx[6], mcu_by[6];   // block j of an MCU: component, block column / row inside the MCU
  const uint32_t *int_bit;    // [n_intervals + 1] first bit of every interval (byte aligned), end of the data
  const uint32_t *int_sub;    // [n_intervals + 1] first subsequence of every interval
  const JdSub *subs;
  JdState *st_in[2];          // start state of every subsequence, double-buffered by round parity
  JdState *st_out;            // state after its last symbol
  JdState *last_in;           // the start state st_out was walked from
  uint32_t *nblk, *base;      // blocks completed inside the subsequence; sequence number of its first block
  uint32_t *flags;            // [r] round r changed something; [SYNC_ROUNDS] 1 = the walk settled; [SYNC_ROUNDS + 1] grid barrier
  int *dc_sum;                // [3][chunks] DC prefix sum scratch
  JdComp c[3];
  int16_t *coefs;             // 64 per block, natural order
  uint8_t *planes;
  uint8_t *out;               // dense rows: RGBA8 (3 components) or gray8
  uint32_t *status;           // first error (JdErr), 0 = none
  uint16_t q[3][64];          // per component, natural order
  JdHuff huff[6];             // component c: DC at 2 c, AC at 2 c + 1
};

__constant__ uint8_t c_natural[80] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                      41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                      30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63,
                                      63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};
const uint8_t h_natural[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                               41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                               30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

// ---- entropy decoding -----------------------------------------------------------------------------------------
// Huffman decoding is serial (a code's position depends on every code before it), and most files carry no restart
// markers, so the parallelism inside an interval comes from speculation (the approach of Weissenberger & Schmidt,
// "Accelerating JPEG decompression on GPUs"): the interval's bits are cut into subsequences of SUB_BITS; a thread
// walks its subsequence from a guessed state (a block starting exactly there) and hands the state it ends in to the
// next one.  Huffman-coded data resynchronises after a few symbols, so after a few rounds of "walk again from what
// my predecessor ended in" nothing changes any more: that fixed point is the true decode, because the first
// subsequence of every interval starts exact and every other start state is its predecessor's end state.  Then the
// blocks a subsequence completes are counted, a prefix sum gives every subsequence its first block, and one more
// walk writes the coefficients (DC differences; the predictions are a second prefix sum).  A file whose walk has not
// settled after SYNC_ROUNDS rounds is decoded by one thread per restart interval instead.

// The 32 bits that start at bit bp.
__device__ __forceinline__ uint32_t jd_window(const uint32_t *ecs, const uint32_t bp) {
  const uint32_t i = bp >> 5;
  const uint32_t w0 = __byte_perm(__ldg(ecs + i), 0u, 0x0123), w1 = __byte_perm(__ldg(ecs + i + 1u), 0u, 0x0123);
  return __funnelshift_l(w1, w0, bp & 31u);
}
__device__ __forceinline__ int jd_extend(const uint32_t r, const uint32_t s) {   // F.2.2.1 EXTEND
  return r < (1u << (s - 1u)) ? (int)r - (int)((1u << s) - 1u) : (int)r;
}
__device__ __forceinline__ int16_t *jd_block(const JdImage &img, const uint32_t g, const uint32_t j) {
  const uint32_t mcu = g / img.bpm, mx = mcu % img.mcus_x, my = mcu / img.mcus_x;
  const JdComp &c = img.c[img.mcu_comp[j]];
  return img.coefs + ((size_t)c.block0 + (size_t)(my * c.v + img.mcu_by[j]) * c.wb + (mx * c.h + img.mcu_bx[j])) * 64;
}

// Canonical walk (T.81 F.2.2.3) for codes longer than the look-ahead: code bits << 8 | symbol, 0 = no such code.
static __device__ __noinline__ uint32_t jd_long_code(const JdHuff &h, const uint32_t w) {
  for (int l = LOOK + 1; l <= 16; l++) {
    const int32_t code = (int32_t)(w >> (32 - l));
    if (h.maxcode[l] >= 0 && code <= h.maxcode[l] && code >= h.mincode[l]) return ((uint32_t)l << 8) | h.vals[h.valptr[l] + code - h.mincode[l]];
  }
  return 0u;
}

// Walks the symbols that start in [s.bp, limit): EMIT writes coefficients (DC as the difference, F.2.1.3), starting
// with block sequence number g and never past g_end (the interval's padding follows its last block).  end_bit: the
// interval's end.  Returns the state after the last symbol and the number of blocks completed; err != 0 (EMIT only):
// the data is corrupt.  DC_ABS: the caller keeps the predictors (serial decoder).
template <bool EMIT, bool DC_ABS>
__device__ __forceinline__ JdState jd_walk(const JdImage &img, const JdHuff *sh, JdState s, const uint32_t limit, const uint32_t end_bit,
                                           uint32_t g, const uint32_t g_end, uint32_t &nblk, uint32_t &err, int *pred) {
  // what every symbol needs, in registers: blocks per MCU, the component of every block of an MCU (2 bits each), and
  // a 64-bit window of the data that is topped up a word at a time (the next bit is bit 63)
  const uint32_t *__restrict__ ecs = img.ecs;
  const uint32_t bpm = img.bpm;
  uint32_t comp_map = 0;
  for (uint32_t j = 0; j < bpm; j++) comp_map |= img.mcu_comp[j] << (2u * j);
  uint32_t bp = s.bp, bim = s.pos >> 8, k = s.pos & 0xffu;
  uint32_t comp = (comp_map >> (2u * bim)) & 3u;
  uint32_t wi = bp >> 5;
  unsigned long long buf = (unsigned long long)__byte_perm(__ldg(ecs + wi), 0u, 0x0123) << (32u + (bp & 31u));
  uint32_t have = 32u - (bp & 31u);
  wi++;
  nblk = 0;
  int16_t *blk = nullptr;
  if (EMIT && g < g_end) blk = jd_block(img, g, bim);
  while (bp < limit) {
    if (EMIT && g >= g_end) break;
    if (have < 32u) {
      buf |= (unsigned long long)__byte_perm(__ldg(ecs + wi), 0u, 0x0123) << (32u - have);
      have += 32u; wi++;
    }
    const uint32_t w = (uint32_t)(buf >> 32);
    const JdHuff &h = sh[2u * comp + (k ? 1u : 0u)];
    const uint32_t e = h.look[w >> (32 - LOOK)];
    uint32_t len = e >> 8, sym = e & 0xffu;
    if (!e) {   // a code longer than the look-ahead
      const uint32_t e2 = jd_long_code(h, w);
      len = e2 >> 8; sym = e2 & 0xffu;
      if (!len) { if (EMIT) err = JD_E_HUFF; s.bp = BAD; return s; }
    }
    // One symbol, without branches on its kind (the 32 lanes of a warp sit at different kinds of symbol: with a branch
    // per kind the warp pays for every kind at every step).
    const bool isdc = k == 0u;
    const uint32_t sbits = isdc ? sym : (sym & 15u), run = sym >> 4, total = len + sbits;
    const uint32_t k_coef = isdc ? 0u : k + run;                       // the coefficient this symbol sets (when sbits != 0)
    const uint32_t k_next = isdc ? 1u : (sbits ? k_coef + 1u : (run == 15u ? k + 16u : 64u));
    if (sbits > 15u || (sbits && k_coef > 63u)) { if (EMIT) err = JD_E_HUFF; s.bp = BAD; return s; }
    if (bp + total > end_bit) {   // the interval's padding (all ones: never a whole code), or data that ends early
      if (EMIT) err = JD_E_TRUNCATED;
      break;
    }
    if (EMIT) {
      const uint32_t vbits = sbits ? (w << len) >> (32u - sbits) : 0u;
      int v = sbits ? jd_extend(vbits, sbits) : 0;
      if (DC_ABS && isdc) { pred[comp] += v; v = pred[comp]; }
      if (isdc || sbits) blk[c_natural[k_coef]] = (int16_t)v;
    }
    bp += total; buf <<= total; have -= total;
    const bool done = k_next >= 64u;
    k = done ? 0u : k_next;
    nblk += done ? 1u : 0u;
    g += done ? 1u : 0u;
    bim = done ? (bim + 1u == bpm ? 0u : bim + 1u) : bim;
    comp = (comp_map >> (2u * bim)) & 3u;
    if (EMIT && done && g < g_end) blk = jd_block(img, g, bim);
  }
  s.bp = bp; s.pos = (bim << 8) | k;
  return s;
}

__device__ __forceinline__ void jd_load_tables(const JdImage &img, JdHuff *sh) {
  const uint32_t n = img.ncomp * 2u * (uint32_t)(sizeof(JdHuff) / 4);
  const uint32_t *src = reinterpret_cast<const uint32_t *>(img.huff);
  uint32_t *dst = reinterpret_cast<uint32_t *>(sh);
  for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
  __syncthreads();
}

constexpr int JH_THREADS = 128;

// The relaxation rounds, in one resident grid (cooperative launch) with a grid barrier between rounds: in every
// round each subsequence whose start state changed is walked again from the state its predecessor ended in.
// Ends when a round changes nothing (flags[SYNC_ROUNDS] = 1) or after SYNC_ROUNDS rounds.
__device__ __forceinline__ void jd_grid_barrier(uint32_t *counter, const uint32_t target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(counter, 1u);
    while (*reinterpret_cast<volatile uint32_t *>(counter) < target) __nanosleep(64);
    __threadfence();
  }
  __syncthreads();
}
__global__ void __launch_bounds__(JH_THREADS) jd_sync_kernel(const JdImage *img_p) {
  __shared__ JdHuff sh[6];
  const JdImage &img = *img_p;
  jd_load_tables(img, sh);
  uint32_t *barrier = img.flags + SYNC_ROUNDS + 1;
  for (uint32_t round = 0; round < (uint32_t)SYNC_ROUNDS; round++) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < img.n_subs; i += gridDim.x * blockDim.x) {
      const JdSub sub = img.subs[i];
      JdState in;
      if (sub.first || round == 0u) { in.bp = sub.start_bit; in.pos = 0u; }   // exact / the guess: a block starts here
      else {
        in = img.st_in[round & 1u][i];
        if (in.bp == BAD) { in.bp = sub.start_bit; in.pos = 0u; }
      }
      const JdState before = img.st_out[i];
      JdState out = before;
      const JdState last = img.last_in[i];
      if (round == 0u || last.bp != in.bp || last.pos != in.pos) {   // the same start gives the same walk: only changed starts are walked again
        uint32_t nblk, err = 0;
        out = jd_walk<false, false>(img, sh, in, sub.end_bit, img.int_bit[sub.interval + 1u], 0u, 0u, nblk, err, nullptr);
        if (before.bp != out.bp || before.pos != out.pos || img.nblk[i] != nblk) img.flags[round] = 1u;
        img.st_out[i] = out;
        img.nblk[i] = nblk;
        img.last_in[i] = in;
      }
      if (i + 1u < img.n_subs) img.st_in[(round + 1u) & 1u][i + 1u] = out;   // (the next interval's first subsequence ignores it)
    }
    jd_grid_barrier(barrier, (round + 1u) * gridDim.x);
    if (*reinterpret_cast<volatile uint32_t *>(img.flags + round) == 0u) {   // nothing changed: the walk has settled
      if (blockIdx.x == 0 && threadIdx.x == 0) img.flags[SYNC_ROUNDS] = 1u;
      return;
    }
  }
}

// Per restart interval (one CTA): has the walk settled; first block of every subsequence; does the count add up.
__global__ void __launch_bounds__(256) jd_scan_kernel(const JdImage *img_p) {
  __shared__ uint32_t warp_sum[8], carry;
  const JdImage &img = *img_p;
  if (img.flags[SYNC_ROUNDS] == 0u) return;   // not settled: the serial decoder takes over
  const uint32_t iv = blockIdx.x, s0 = img.int_sub[iv], s1 = img.int_sub[iv + 1u];
  const uint32_t total = img.mcus_x * img.mcus_y, m0 = iv * img.restart, m1 = min(total, m0 + img.restart);
  if (threadIdx.x == 0) carry = m0 * img.bpm;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  for (uint32_t at = s0; at < s1; at += blockDim.x) {
    const uint32_t i = at + threadIdx.x;
    const uint32_t v = i < s1 ? img.nblk[i] : 0u;
    uint32_t x = v;
    for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if ((int)lane >= d) x += y; }
    if (lane == 31u) warp_sum[warp] = x;
    __syncthreads();
    uint32_t before = carry;
    for (uint32_t w2 = 0; w2 < warp; w2++) before += warp_sum[w2];
    if (i < s1) img.base[i] = before + x - v;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1u) carry = before + x;
    __syncthreads();
  }
  if (threadIdx.x == 0 && carry != m1 * img.bpm) atomicCAS(img.status, 0u, carry < m1 * img.bpm ? (uint32_t)JD_E_TRUNCATED : (uint32_t)JD_E_HUFF);
}

// The last walk: coefficients into HBM (DC differences).
__global__ void __launch_bounds__(JH_THREADS) jd_emit_kernel(const JdImage *img_p) {
  __shared__ JdHuff sh[6];
  const JdImage &img = *img_p;
  if (img.flags[SYNC_ROUNDS] == 0u) return;
  jd_load_tables(img, sh);
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= img.n_subs) return;
  const JdSub sub = img.subs[i];
  JdState in;
  if (sub.first) { in.bp = sub.start_bit; in.pos = 0u; }
  else in = img.st_in[0][i];
  if (in.bp == BAD) { atomicCAS(img.status, 0u, (uint32_t)JD_E_HUFF); return; }
  const uint32_t total = img.mcus_x * img.mcus_y, m1 = min(total, (sub.interval + 1u) * img.restart);
  uint32_t nblk, err = 0;
  jd_walk<true, false>(img, sh, in, sub.end_bit, img.int_bit[sub.interval + 1u], img.base[i], m1 * img.bpm, nblk, err, nullptr);
  if (err) atomicCAS(img.status, 0u, err);
}

// DC predictions (F.2.1.3.1: a running sum per component, restarting with every interval) as a prefix sum over chunks
// of DC_CHUNK MCUs: chunk sums, their exclusive scan per interval, then the running sum inside every chunk.
__device__ __forceinline__ void jd_chunk_range(const JdImage &img, const uint32_t chunk, uint32_t &m0, uint32_t &m1) {
  const uint32_t cpi = (img.restart + DC_CHUNK - 1u) / DC_CHUNK, iv = chunk / cpi, piece = chunk - iv * cpi;
  const uint32_t total = img.mcus_x * img.mcus_y, i0 = iv * img.restart, i1 = min(total, i0 + img.restart);
  m0 = min(i1, i0 + piece * DC_CHUNK); m1 = min(i1, m0 + DC_CHUNK);
}
template <int PHASE>
__global__ void __launch_bounds__(128) jd_dc_kernel(const JdImage *img_p, const uint32_t n_chunks) {
  const JdImage &img = *img_p;
  if (img.flags[SYNC_ROUNDS] == 0u) return;   // the serial decoder wrote the predictions itself
  const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
  const uint32_t cpi = (img.restart + DC_CHUNK - 1u) / DC_CHUNK;
  if (PHASE == 1) {   // one warp per interval and component: exclusive scan of its chunk sums
    const uint32_t wid = t >> 5, lane = threadIdx.x & 31u;
    if (wid >= img.n_intervals * img.ncomp) return;
    const uint32_t iv = wid / img.ncomp, comp = wid - iv * img.ncomp;
    int *sums = img.dc_sum + (size_t)comp * n_chunks + (size_t)iv * cpi;
    int carry = 0;
    for (uint32_t at = 0; at < cpi; at += 32u) {
      const uint32_t p = at + lane;
      const int v = p < cpi ? sums[p] : 0;
      int x = v;
      for (int d = 1; d < 32; d <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, d); if ((int)lane >= d) x += y; }
      if (p < cpi) sums[p] = carry + x - v;
      carry += __shfl_sync(0xffffffffu, x, 31);
    }
    return;
  }
  if (t >= n_chunks * img.ncomp) return;
  const uint32_t chunk = t / img.ncomp, comp = t - chunk * img.ncomp;
  uint32_t m0, m1;
  jd_chunk_range(img, chunk, m0, m1);
  int run = PHASE == 2 ? img.dc_sum[(size_t)comp * n_chunks + chunk] : 0;
  for (uint32_t m = m0; m < m1; m++) {
    for (uint32_t j = 0; j < img.bpm; j++) {
      if (img.mcu_comp[j] != comp) continue;
      int16_t *blk = jd_block(img, m * img.bpm, j);
      run += blk[0];
      if (PHASE == 2) blk[0] = (int16_t)run;
    }
  }
  if (PHASE == 0) img.dc_sum[(size_t)comp * n_chunks + chunk] = run;
}

// The serial decoder: one thread per restart interval.  Runs only when the speculative walk did not settle.
__global__ void __launch_bounds__(JH_THREADS) jd_serial_kernel(const JdImage *img_p) {
  __shared__ JdHuff sh[6];
  const JdImage &img = *img_p;
  if (img.flags[SYNC_ROUNDS] != 0u) return;
  jd_load_tables(img, sh);
  const uint32_t iv = blockIdx.x * blockDim.x + threadIdx.x;
  if (iv >= img.n_intervals) return;
  const uint32_t total = img.mcus_x * img.mcus_y, m0 = iv * img.restart, m1 = min(total, m0 + img.restart);
  JdState in;
  in.bp = img.int_bit[iv]; in.pos = 0u;
  int pred[3] = {0, 0, 0};
  uint32_t nblk, err = 0;
  const uint32_t end_bit = img.int_bit[iv + 1u];
  jd_walk<true, true>(img, sh, in, end_bit, end_bit, m0 * img.bpm, m1 * img.bpm, nblk, err, pred);
  if (!err && nblk != (m1 - m0) * img.bpm) err = JD_E_TRUNCATED;
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
  size_t ecs_off = 0;              // the entropy-coded segment starts here
  uint32_t n_intervals = 0;
  std::vector<uint32_t> int_byte;  // [n_intervals + 1] interval starts in the UNSTUFFED data, its length (unstuff())
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
      const uint32_t total = f.mcus_x * f.mcus_y;
      if (!f.restart) f.restart = total;
      f.n_intervals = (total + f.restart - 1) / f.restart;
      return JD_OK;
    }
    o += len;
  }
  return have_sof ? JD_E_TRUNCATED : JD_E_NO_SOF;
}

// The entropy-coded segment without its framing (T.81 B.1.1.5: the zero after every 0xFF data byte; B.2.1: the RSTn
// markers between restart intervals), copied to dst (capacity: the rest of the file).  The segment ends at the first
// marker that is neither.  Fills f.int_byte.
uint32_t unstuff(const uint8_t *d, size_t n, HostFrame &f, uint8_t *dst) {
  f.int_byte.assign(1, 0u);
  size_t e = f.ecs_off, out = 0;
  while (e < n) {
    const uint8_t *ff = static_cast<const uint8_t *>(memchr(d + e, 0xff, n - e));
    const size_t stop = ff ? (size_t)(ff - d) : n;
    memcpy(dst + out, d + e, stop - e);
    out += stop - e;
    e = stop;
    if (e + 1 >= n) { e = n; break; }
    const uint8_t mk = d[e + 1];
    if (mk == 0x00) { dst[out++] = 0xff; e += 2; continue; }
    if (mk == 0xff) { e += 1; continue; }   // fill byte
    if (mk >= 0xd0 && mk <= 0xd7) {
      if (f.int_byte.size() >= f.n_intervals || mk != 0xd0 + ((f.int_byte.size() - 1) & 7)) return JD_E_RST;
      f.int_byte.push_back((uint32_t)out);
      e += 2;
      continue;
    }
    break;   // EOI or another segment
  }
  if (f.int_byte.size() != f.n_intervals) return f.n_intervals > 1 && e < n ? JD_E_RST : JD_E_TRUNCATED;
  if (out >= (1ull << 28)) return JD_E_TRUNCATED;   // bit positions are 32-bit
  f.int_byte.push_back((uint32_t)out);
  memset(dst + out, 0, 8);   // the readers look one word ahead
  return JD_OK;
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

// Pinned staging memory of the job (files are unstuffed into it and copied from it asynchronously; pageable sources
// would make every copy wait for the decode kernels queued before it).  Lives until the job is destroyed.
static uint8_t *job_pinned(stitch_job *job, size_t bytes) {
  bytes = (bytes + 255) & ~(size_t)255;
  if (job->pinned.empty() || job->pinned_used + bytes > job->pinned.back().second) {
    size_t cap = std::max<size_t>(bytes, (size_t)32 << 20);
    void *p = nullptr;
    stitch_ctx *ctx = job->ctx;
    if (ctx->pinned_spare && ctx->pinned_spare_cap >= bytes) {   // the chunk the last job left behind
      p = ctx->pinned_spare; cap = ctx->pinned_spare_cap;
      ctx->pinned_spare = nullptr; ctx->pinned_spare_cap = 0;
    } else if (cudaHostAlloc(&p, cap, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    job->pinned.push_back({static_cast<uint8_t *>(p), cap});
    job->pinned_used = 0;
  }
  uint8_t *r = job->pinned.back().first + job->pinned_used;
  job->pinned_used += bytes;
  return r;
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
  // host block (pinned): image tables | interval bits | interval subsequences | subsequences | unstuffed data
  const size_t raw_cap = len - f.ecs_off + 16;
  uint8_t *h_ecs = job_pinned(job, raw_cap);
  if (!h_ecs) return stitch_fail(ctx, STITCH_E_NOMEM, "JPEG tile: no pinned staging memory");
  const uint32_t ue = unstuff(data, len, f, h_ecs);
  if (ue) return stitch_fail(ctx, STITCH_E_JPEG, "%s", jd_message(ue));
  const uint32_t n_int = f.n_intervals;
  const size_t ecs_bytes = ((size_t)f.int_byte[n_int] + 8 + 3) & ~(size_t)3;
  std::vector<JdSub> subs;
  std::vector<uint32_t> int_bit(n_int + 1), int_sub(n_int + 1);
  // STITCH_JPEG_SUB_BITS: tuning and tests (tiny subsequences do not settle and exercise the serial decoder); read at every call
  const char *sb_env = getenv("STITCH_JPEG_SUB_BITS");
  const uint32_t sub_bits = sb_env ? (uint32_t)std::min(1 << 30, std::max(64, atoi(sb_env))) : SUB_BITS;
  for (uint32_t iv = 0; iv < n_int; iv++) {
    const uint32_t b0 = f.int_byte[iv] * 8u, b1 = f.int_byte[iv + 1] * 8u;
    int_bit[iv] = b0;
    int_sub[iv] = (uint32_t)subs.size();
    uint32_t at = b0;
    do {
      JdSub sb;
      sb.start_bit = at; sb.end_bit = std::min(b1, at + sub_bits); sb.interval = iv; sb.first = at == b0 ? 1u : 0u;
      subs.push_back(sb);
      at += sub_bits;
    } while (at < b1);
  }
  int_bit[n_int] = f.int_byte[n_int] * 8u;
  int_sub[n_int] = (uint32_t)subs.size();
  const uint32_t n_subs = (uint32_t)subs.size();

  size_t n_blocks = 0, plane_bytes = 0;
  JdImage img;
  memset(&img, 0, sizeof img);
  img.width = f.width; img.height = f.height; img.ncomp = f.ncomp; img.hmax = f.hmax; img.vmax = f.vmax;
  img.mcus_x = f.mcus_x; img.mcus_y = f.mcus_y; img.restart = f.restart; img.n_intervals = n_int; img.n_subs = n_subs;
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
    for (uint32_t v = 0; v < c.v; v++)
      for (uint32_t hh = 0; hh < c.h; hh++) { img.mcu_comp[img.bpm] = i; img.mcu_bx[img.bpm] = hh; img.mcu_by[img.bpm] = v; img.bpm++; }
  }
  if (n_blocks >= (1u << 26) || plane_bytes >= (1ull << 32)) return stitch_fail(ctx, STITCH_E_INVALID, "JPEG tile too large");
  const uint32_t ch = f.ncomp == 1 ? 1u : 4u;
  const size_t out_bytes = (size_t)f.width * f.height * ch;
  const uint32_t cpi = (f.restart + DC_CHUNK - 1) / DC_CHUNK, n_chunks = n_int * cpi;
  // status slot of this tile (pinned, written by the last kernel below)
  if (!job->h_jdstat) {
    STITCH_CK(cudaHostAlloc((void **)&job->h_jdstat, 4096 * sizeof(uint32_t), cudaHostAllocDefault));
    job->jdstat_cap = 4096;
  }
  if (job->jdstat_used >= job->jdstat_cap) return stitch_fail(ctx, STITCH_E_INVALID, "more than %zu JPEG tiles in one job", job->jdstat_cap);
  const uint32_t slot = (uint32_t)job->jdstat_used++;
  job->h_jdstat[slot] = 0;

  // device block: [uploaded: image | int_bit | int_sub | subs | data] [scratch: states, counts, flags, status, DC sums]
  auto up16 = [](size_t x) { return (x + 15) & ~(size_t)15; };
  const size_t o_img = 0, o_ibit = up16(sizeof(JdImage)), o_isub = o_ibit + up16((n_int + 1) * 4), o_subs = o_isub + up16((n_int + 1) * 4);
  const size_t o_ecs = o_subs + up16((size_t)n_subs * sizeof(JdSub)), up_bytes = o_ecs + up16(ecs_bytes);
  const size_t o_in0 = up_bytes, o_in1 = o_in0 + up16((size_t)n_subs * 8), o_out = o_in1 + up16((size_t)n_subs * 8);
  const size_t o_last = o_out + up16((size_t)n_subs * 8);
  const size_t o_nblk = o_last + up16((size_t)n_subs * 8), o_base = o_nblk + up16((size_t)n_subs * 4), o_flags = o_base + up16((size_t)n_subs * 4);
  const size_t o_stat = o_flags + up16((SYNC_ROUNDS + 2) * 4), o_dc = o_stat + 16, dev_bytes = o_dc + up16((size_t)n_chunks * 3 * 4);
  uint8_t *h_up = job_pinned(job, o_ecs);   // tables (the data is already pinned at h_ecs)
  if (!h_up) return stitch_fail(ctx, STITCH_E_NOMEM, "JPEG tile: no pinned staging memory");

  // the decode chains of successive tiles take turns on a few streams (each behind the uploads queued so far)
  cudaStream_t &slot_stream = ctx->jd_streams[ctx->jd_next++ & 3u];
  if (!slot_stream) STITCH_CK(cudaStreamCreateWithFlags(&slot_stream, cudaStreamNonBlocking));
  cudaStream_t st = slot_stream;
  uint8_t *d_out = nullptr, *d_blk = nullptr, *d_planes = nullptr;
  int16_t *d_coefs = nullptr;
  cudaError_t e = cudaMallocAsync((void **)&d_out, out_bytes + 32, st);
  if (e == cudaSuccess) e = cudaMallocAsync((void **)&d_blk, dev_bytes, st);
  if (e == cudaSuccess) e = cudaMallocAsync((void **)&d_planes, plane_bytes, st);
  if (e == cudaSuccess) e = cudaMallocAsync((void **)&d_coefs, n_blocks * 128, st);
  if (e == cudaSuccess) {
    img.ecs = reinterpret_cast<const uint32_t *>(d_blk + o_ecs);
    img.int_bit = reinterpret_cast<const uint32_t *>(d_blk + o_ibit); img.int_sub = reinterpret_cast<const uint32_t *>(d_blk + o_isub);
    img.subs = reinterpret_cast<const JdSub *>(d_blk + o_subs);
    img.st_in[0] = reinterpret_cast<JdState *>(d_blk + o_in0); img.st_in[1] = reinterpret_cast<JdState *>(d_blk + o_in1);
    img.st_out = reinterpret_cast<JdState *>(d_blk + o_out); img.last_in = reinterpret_cast<JdState *>(d_blk + o_last);
    img.nblk = reinterpret_cast<uint32_t *>(d_blk + o_nblk); img.base = reinterpret_cast<uint32_t *>(d_blk + o_base);
    img.flags = reinterpret_cast<uint32_t *>(d_blk + o_flags); img.status = reinterpret_cast<uint32_t *>(d_blk + o_stat);
    img.dc_sum = reinterpret_cast<int *>(d_blk + o_dc);
    img.coefs = d_coefs; img.planes = d_planes; img.out = d_out;
    memcpy(h_up + o_img, &img, sizeof img);
    memcpy(h_up + o_ibit, int_bit.data(), (n_int + 1) * 4);
    memcpy(h_up + o_isub, int_sub.data(), (n_int + 1) * 4);
    memcpy(h_up + o_subs, subs.data(), (size_t)n_subs * sizeof(JdSub));
    e = cudaMemcpyAsync(d_blk, h_up, o_ecs, cudaMemcpyHostToDevice, st);
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_blk + o_ecs, h_ecs, ecs_bytes, cudaMemcpyHostToDevice, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_blk + o_out, 0xff, up16((size_t)n_subs * 8), st);                 // st_out: BAD
  if (e == cudaSuccess) e = cudaMemsetAsync(d_blk + o_nblk, 0, o_dc - o_nblk, st);                              // counts, flags, status
  if (e == cudaSuccess) e = cudaMemsetAsync(d_coefs, 0, n_blocks * 128, st);
  if (e == cudaSuccess) {
    const JdImage *d_img = reinterpret_cast<const JdImage *>(d_blk);
    const unsigned sub_grid = (n_subs + JH_THREADS - 1) / JH_THREADS;
    {
      static PerDevice occ_once;
      int occ = 0;
      e = occ_once.get(&occ, [](int *v) { return cudaOccupancyMaxActiveBlocksPerMultiprocessor(v, jd_sync_kernel, JH_THREADS, 0); });
      const unsigned resident = (unsigned)std::max(1, occ) * (unsigned)ctx->sm_count;
      const JdImage *arg = d_img;
      void *args[] = {(void *)&arg};
      if (e == cudaSuccess)
        e = cudaLaunchCooperativeKernel((const void *)jd_sync_kernel, dim3(std::min(sub_grid, resident)), dim3(JH_THREADS), args, 0, st);
    }
    jd_scan_kernel<<<n_int, 256, 0, st>>>(d_img);
    jd_emit_kernel<<<sub_grid, JH_THREADS, 0, st>>>(d_img);
    jd_dc_kernel<0><<<(n_chunks * f.ncomp + 127) / 128, 128, 0, st>>>(d_img, n_chunks);
    jd_dc_kernel<1><<<(n_int * f.ncomp * 32 + 127) / 128, 128, 0, st>>>(d_img, n_chunks);
    jd_dc_kernel<2><<<(n_chunks * f.ncomp + 127) / 128, 128, 0, st>>>(d_img, n_chunks);
    jd_serial_kernel<<<(n_int + JH_THREADS - 1) / JH_THREADS, JH_THREADS, 0, st>>>(d_img);
    jd_idct_kernel<<<(unsigned)((n_blocks + 127) / 128), 128, 0, st>>>(d_img, (uint32_t)n_blocks);
    jd_color_kernel<<<dim3((f.width + 255) / 256, f.height), 256, 0, st>>>(d_img);
    jd_status_out_kernel<<<1, 1, 0, st>>>(job->h_jdstat + slot, img.status);
    e = cudaGetLastError();
  }
  if (d_blk) cudaFreeAsync(d_blk, st);
  if (d_planes) cudaFreeAsync(d_planes, st);
  if (d_coefs) cudaFreeAsync(d_coefs, st);
  if (e != cudaSuccess) {
    if (d_out) cudaFreeAsync(d_out, st);
    return stitch_fail(ctx, e == cudaErrorMemoryAllocation ? STITCH_E_NOMEM : STITCH_E_CUDA, "JPEG tile: %s", cudaGetErrorString(e));
  }
  stitch_tile_desc d = *tile;
  d.width = f.width; d.height = f.height; d.bit_depth = 8; d.color_type = f.ncomp == 1 ? 0u : 6u; d.interlace = 0; d.filtered = 0;
  const int rc = stitch_add_tile_produced(job, &d, d_out, out_bytes, st);
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
