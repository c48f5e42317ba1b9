// deflate.cu -- SURVEY.md 8(f).1: the filtered scanlines of a band become a zlib-compatible DEFLATE stream
// on the device, so that compressed bytes cross the bus and the host's deflate (src/streaming-deflate.ts,
// by far the slowest host stage) disappears.  The bytes need not equal pako's: the requirement is that
// they inflate to the same filtered rows.
//
// Filtered PNG rows are residuals: entropy coding gets almost everything, string matching almost nothing
// (zlib level 6 reaches 1.77 on the benchmark tiles, a per-block Huffman code 1.95).  So the stream is a
// sequence of independent 32 KiB blocks, each a dynamic-Huffman block of literals only:
//
//   D1 plan     per block: byte histogram (per-warp shared-memory bins), optimal code lengths (rank sort +
//               two-queue Huffman by one thread, 257 symbols), zlib-style limiting to 15 bits, canonical
//               codes bit-reversed for the LSB-first stream; coded size, or "stored" when coding does not pay;
//               partial Adler-32 sums
//   D2 scan     exclusive scan of the block sizes, Adler-32 of the band
//   D3 encode   per block: raw bytes staged in shared memory, per-thread bit offsets by a block scan, codes
//               OR-ed into a shared-memory image of the block, header in front, Z_SYNC_FLUSH-style empty
//               stored block behind (so every block ends on a byte boundary and blocks concatenate), copy out
//
// Format references: RFC 1950 (zlib wrapper, Adler-32), RFC 1951 sections 3.2.4, 3.2.7.
#include "job.h"

#include <cstring>

namespace {
using namespace stitch;

constexpr int DB = 32768;            // raw bytes per block (a stored block holds at most 65535)
constexpr int DT = 256;              // threads per CTA
constexpr int PER_T = DB / DT;       // 128 bytes per thread in D3
constexpr int NSYM = 257;            // literals + end-of-block
constexpr int TAB = 260;             // table stride in words
constexpr uint32_t ADLER_M = 65521u;
constexpr uint32_t HDR_BITS = 3 + 5 + 5 + 4 + 19 * 3 + 258 * 4;  // 1106

struct BandView {
  const uint8_t *base;   // first byte of row 0 (filter byte)
  size_t pitch;          // bytes between rows on the device
  uint32_t row_bytes;    // bytes per row in the stream
  unsigned long long total;
};

__device__ __forceinline__ const uint8_t *byte_addr(const BandView &v, unsigned long long o) {
  const unsigned long long r = o / v.row_bytes;
  return v.base + r * v.pitch + (o - r * v.row_bytes);
}

// Calls fn(address of stream byte o, o - o0) for every o in [o0, o1), consecutive threads on consecutive bytes.
template <class F>
__device__ __forceinline__ void for_bytes(const BandView &v, unsigned long long o0, unsigned long long o1, F fn) {
  unsigned long long o = o0;
  while (o < o1) {
    const unsigned long long r = o / v.row_bytes, c = o - r * v.row_bytes;
    const unsigned long long n = min((unsigned long long)v.row_bytes - c, o1 - o);
    const uint8_t *p = v.base + r * v.pitch + c;
    for (unsigned long long j = threadIdx.x; j < n; j += DT) fn(p[j], (uint32_t)(o - o0 + j));
    o += n;
  }
}

// ---- D1 ----------------------------------------------------------------------------------------------------
struct PlanSmem {
  uint32_t hist[DT / 32][NSYM + 3];
  uint32_t freq[NSYM + 3];
  uint16_t sorted[NSYM + 3];      // symbols by (freq, symbol) ascending
  uint32_t weight[2 * NSYM];      // tree nodes: leaves in sorted order, then internal nodes
  uint16_t parent[2 * NSYM];
  uint8_t len[NSYM + 3];
  uint32_t red[DT / 32][3];
};

__global__ void __launch_bounds__(DT) deflate_plan_kernel(const BandView v, uint32_t *tabs, uint32_t *sizes, uint32_t *adler_parts) {
  __shared__ PlanSmem s;
  const uint32_t t = threadIdx.x, warp = t >> 5, lane = t & 31u;
  const unsigned long long o0 = (unsigned long long)blockIdx.x * DB, o1 = min(o0 + DB, v.total);
  const uint32_t L = (uint32_t)(o1 - o0);
  for (uint32_t i = t; i < (DT / 32) * (NSYM + 3); i += DT) (&s.hist[0][0])[i] = 0u;
  __syncthreads();
  unsigned long long s1 = 0, s2 = 0;  // Adler partial sums "from zero": sum b, sum (L - j) b
  for_bytes(v, o0, o1, [&](uint8_t b, uint32_t j) {
    atomicAdd(&s.hist[warp][b], 1u);
    s1 += b;
    s2 += (unsigned long long)(L - j) * b;
  });
  __syncthreads();
  for (uint32_t i = t; i < NSYM; i += DT) {
    uint32_t f = 0;
#pragma unroll
    for (int w = 0; w < DT / 32; w++) f += s.hist[w][i];
    s.freq[i] = i == 256u ? 1u : f;
  }
  // Adler partials -> one pair per block
  s1 %= ADLER_M; s2 %= ADLER_M;
  for (int o = 16; o > 0; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
  if (lane == 0u) { s.red[warp][0] = (uint32_t)(s1 % ADLER_M); s.red[warp][1] = (uint32_t)(s2 % ADLER_M); }
  __syncthreads();
  // rank sort: position of symbol i among all symbols ordered by (freq, symbol)
  for (uint32_t i = t; i < NSYM; i += DT) {
    const uint32_t fi = s.freq[i];
    uint32_t rank = 0;
    for (uint32_t k = 0; k < NSYM; k++) {
      const uint32_t fk = s.freq[k];
      rank += (fk < fi || (fk == fi && k < i)) ? 1u : 0u;
    }
    s.sorted[rank] = (uint16_t)i;
  }
  __syncthreads();
  if (t == 0u) {
    uint32_t a = 0, b = 0;
    for (int w = 0; w < DT / 32; w++) { a += s.red[w][0]; b += s.red[w][1]; }
    adler_parts[2 * blockIdx.x] = a % ADLER_M;
    adler_parts[2 * blockIdx.x + 1] = b % ADLER_M;
    // ---- Huffman code lengths (one thread; 257 symbols) ----
    uint32_t z = 0;
    while (z < NSYM && s.freq[s.sorted[z]] == 0u) z++;  // unused symbols sort first
    const uint32_t n = NSYM - z;                          // >= 2: at least one literal and end-of-block
    for (uint32_t i = 0; i < n; i++) s.weight[i] = s.freq[s.sorted[z + i]];
    uint32_t leaf = 0, inode = n, made = n;               // two queues: leaves [leaf, n), internal [inode, made)
    for (uint32_t k = 0; k + 1 < n; k++) {
      uint32_t pick[2];
      for (int q = 0; q < 2; q++) {
        if (leaf < n && (inode >= made || s.weight[leaf] <= s.weight[inode])) pick[q] = leaf++;
        else pick[q] = inode++;
      }
      s.weight[made] = s.weight[pick[0]] + s.weight[pick[1]];
      s.parent[pick[0]] = (uint16_t)made; s.parent[pick[1]] = (uint16_t)made;
      made++;
    }
    // depths: a parent is always created after its children, so walk down from the root (node made-1)
    uint32_t bl_count[32];
    for (int i = 0; i < 32; i++) bl_count[i] = 0;
    // reuse weight[] for depths
    s.weight[made - 1] = 0;
    for (int i = (int)made - 2; i >= 0; i--) s.weight[i] = s.weight[s.parent[i]] + 1u;
    uint32_t overflow = 0;
    for (uint32_t i = 0; i < n; i++) {
      uint32_t d = s.weight[i];
      if (d > 15u) { d = 15u; overflow++; }
      bl_count[d]++;
    }
    // zlib's gen_bitlen repair: move one leaf down from a shorter length for every two overflowing ones
    while (overflow > 0) {
      uint32_t bits = 14;
      while (bl_count[bits] == 0u) bits--;
      bl_count[bits]--; bl_count[bits + 1] += 2u; bl_count[15]--;
      overflow -= overflow >= 2u ? 2u : overflow;
    }
    // hand the lengths out again: the least frequent symbols get the longest codes
    for (uint32_t i = 0; i < NSYM; i++) s.len[i] = 0;
    uint32_t i = 0;
    for (int bits = 15; bits >= 1; bits--)
      for (uint32_t c = 0; c < bl_count[bits]; c++) s.len[s.sorted[z + i++]] = (uint8_t)bits;
    // canonical codes (RFC 1951 3.2.2) into weight[] by symbol
    uint32_t next_code[16], code = 0;
    uint32_t cnt[16];
    for (int b2 = 0; b2 < 16; b2++) cnt[b2] = 0;
    for (uint32_t sy = 0; sy < NSYM; sy++) cnt[s.len[sy]]++;
    cnt[0] = 0;
    for (int b2 = 1; b2 < 16; b2++) { code = (code + cnt[b2 - 1]) << 1; next_code[b2] = code; }
    for (uint32_t sy = 0; sy < NSYM; sy++) {
      const uint32_t l = s.len[sy];
      s.weight[sy] = l ? next_code[l]++ : 0u;
    }
  }
  __syncthreads();
  // table entry: code bit-reversed (Huffman codes go MSB first into an LSB-first stream) | length << 16
  unsigned long long bits = 0;
  for (uint32_t i = t; i < NSYM; i += DT) {
    const uint32_t l = s.len[i];
    const uint32_t rev = l ? (__brev(s.weight[i]) >> (32u - l)) : 0u;
    tabs[(size_t)blockIdx.x * TAB + i] = rev | (l << 16);
    bits += (unsigned long long)s.freq[i] * l;
  }
  for (int o = 16; o > 0; o >>= 1) bits += __shfl_xor_sync(0xffffffffu, bits, o);
  if (lane == 0u) s.red[warp][2] = (uint32_t)bits;
  __syncthreads();
  if (t == 0u) {
    uint32_t tot = 0;
    for (int w = 0; w < DT / 32; w++) tot += s.red[w][2];
    const uint32_t coded = (HDR_BITS + tot + 3u + 7u) / 8u + 4u;  // block + empty stored block after it
    const uint32_t stored = 5u + L + 5u;                           // stored block + the same empty stored block
    // bit 31: the block is written as a stored block
    sizes[blockIdx.x] = coded < stored ? coded : (stored | 0x80000000u);
  }
}

// ---- D2: exclusive scan of the sizes, Adler-32 of the band ---------------------------------------------------
__global__ void __launch_bounds__(1024) deflate_scan_kernel(const uint32_t *sizes, const uint32_t *adler_parts, uint32_t n_blocks,
                                                            unsigned long long total, unsigned long long *offsets,
                                                            unsigned long long *result /* [0] bytes, [1] adler */) {
  __shared__ unsigned long long warp_tot[32];
  __shared__ unsigned long long s_base;
  __shared__ unsigned long long red[2][32];
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  unsigned long long a_sum = 0, b_sum = 0;
  for (uint32_t b0 = 0; b0 < n_blocks; b0 += 1024u) {
    const uint32_t b = b0 + threadIdx.x;
    unsigned long long sz = 0;
    if (b < n_blocks) {
      sz = sizes[b] & 0x7fffffffu;
      // bytes of the band after this block weigh its byte sum once more each (Adler-32's second sum)
      const unsigned long long end = min((unsigned long long)(b + 1u) * DB, total);
      a_sum += adler_parts[2 * b];
      b_sum += (adler_parts[2 * b + 1] + ((total - end) % ADLER_M) * adler_parts[2 * b]) % ADLER_M;
    }
    unsigned long long incl = sz;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const unsigned long long x = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += x; }
    if (lane == 31u) warp_tot[warp] = incl;
    __syncthreads();
    unsigned long long wbase = 0;
    for (uint32_t k = 0; k < warp; k++) wbase += warp_tot[k];
    const unsigned long long base = s_base;
    if (b < n_blocks) offsets[b] = base + wbase + incl - sz;
    __syncthreads();
    if (threadIdx.x == 1023u) s_base = base + wbase + incl;
    __syncthreads();
  }
  a_sum %= ADLER_M; b_sum %= ADLER_M;
  for (int o = 16; o > 0; o >>= 1) { a_sum += __shfl_xor_sync(0xffffffffu, a_sum, o); b_sum += __shfl_xor_sync(0xffffffffu, b_sum, o); }
  if (lane == 0u) { red[0][warp] = a_sum % ADLER_M; red[1][warp] = b_sum % ADLER_M; }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long a = 0, b = 0;
    for (int w = 0; w < 32; w++) { a += red[0][w]; b += red[1][w]; }
    a = (1ull + a) % ADLER_M;                   // Adler-32 starts at a = 1, which also adds `total` to b
    b = (total % ADLER_M + b) % ADLER_M;
    result[0] = s_base;
    result[1] = (b << 16) | a;
    result[2] = 0;  // CRC-32 remainder, accumulated by crc32_kernel
    result[3] = 0;
  }
}

// ---- D3 ----------------------------------------------------------------------------------------------------
constexpr int RAW_STRIDE = PER_T / 4 + 1;         // 33 words per thread: conflict-free private runs
constexpr int OUT_WORDS = (DB + 512) / 4;         // a coded block is smaller than a stored one (D1), header included

struct EncSmem {
  uint32_t raw[DT * RAW_STRIDE];
  uint32_t out[OUT_WORDS];
  uint32_t tab[TAB];
  uint32_t warp_tot[DT / 32];
};

__device__ __forceinline__ void or_bits(uint32_t *out, uint32_t pos, uint32_t value, uint32_t nbits) {  // nbits <= 32
  const unsigned long long v = (unsigned long long)value << (pos & 31u);
  atomicOr(out + (pos >> 5), (uint32_t)v);
  if ((pos & 31u) + nbits > 32u) atomicOr(out + (pos >> 5) + 1, (uint32_t)(v >> 32));
}

__global__ void __launch_bounds__(DT) deflate_encode_kernel(const BandView v, const uint32_t *tabs, const uint32_t *sizes,
                                                            const unsigned long long *offsets, uint8_t *dst) {
  extern __shared__ __align__(16) uint8_t enc_smem[];
  EncSmem &s = *reinterpret_cast<EncSmem *>(enc_smem);
  const uint32_t t = threadIdx.x, warp = t >> 5, lane = t & 31u;
  const unsigned long long o0 = (unsigned long long)blockIdx.x * DB, o1 = min(o0 + DB, v.total);
  const uint32_t L = (uint32_t)(o1 - o0);
  const uint32_t size = sizes[blockIdx.x] & 0x7fffffffu;
  const bool stored = (sizes[blockIdx.x] >> 31) != 0u;
  uint8_t *out = dst + offsets[blockIdx.x];
  if (stored) {  // 00 LEN NLEN raw..., then the empty stored block
    if (t == 0u) {
      out[0] = 0; out[1] = (uint8_t)(L & 0xffu); out[2] = (uint8_t)(L >> 8); out[3] = (uint8_t)(~L & 0xffu); out[4] = (uint8_t)((~L >> 8) & 0xffu);
      uint8_t *e = out + 5 + L;
      e[0] = 0; e[1] = 0; e[2] = 0; e[3] = 0xff; e[4] = 0xff;
    }
    for_bytes(v, o0, o1, [&](uint8_t b, uint32_t j) { out[5 + j] = b; });
    return;
  }
  for (uint32_t i = t; i < TAB; i += DT) s.tab[i] = i < NSYM ? tabs[(size_t)blockIdx.x * TAB + i] : 0u;
  for (uint32_t i = t; i < OUT_WORDS; i += DT) s.out[i] = 0u;
  uint8_t *rawb = reinterpret_cast<uint8_t *>(s.raw);
  for_bytes(v, o0, o1, [&](uint8_t b, uint32_t j) { rawb[(j / PER_T) * (RAW_STRIDE * 4) + (j % PER_T)] = b; });
  __syncthreads();
  // my run of the block: bytes [t * PER_T, ...)
  const uint32_t first = t * PER_T, mine = first < L ? min((uint32_t)PER_T, L - first) : 0u;
  const uint8_t *my = rawb + t * (RAW_STRIDE * 4);
  uint32_t bits = 0;
  for (uint32_t j = 0; j < mine; j++) bits += s.tab[my[j]] >> 16;
  uint32_t incl = bits;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const uint32_t x = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= (uint32_t)o) incl += x; }
  if (lane == 31u) s.warp_tot[warp] = incl;
  __syncthreads();
  uint32_t pos = HDR_BITS + incl - bits;
  for (uint32_t k = 0; k < warp; k++) pos += s.warp_tot[k];
  // header: BFINAL 0, BTYPE 10, HLIT 0 (257 codes), HDIST 0 (1 code), HCLEN 15 (19 lengths), the code-length
  // code gives 4 bits to each of the symbols 0..15 and none to 16, 17, 18 (so a length l is sent as the 4-bit
  // code l, MSB first); then 257 literal/length lengths and one distance length of 0
  if (t == 0u) {
    or_bits(s.out, 0, 0u | (2u << 1) | (0u << 3) | (0u << 8) | (15u << 13), 17);
    // order 16 17 18 0 8 7 9 6 10 5 11 4 12 3 13 2 14 1 15: three zeros, then sixteen fours
    uint32_t p = 17 + 9;
    for (int k = 0; k < 16; k++, p += 3) or_bits(s.out, p, 4u, 3);
  }
  for (uint32_t i = t; i < 258u; i += DT) {
    const uint32_t l = i < NSYM ? (s.tab[i] >> 16) : 0u;
    or_bits(s.out, 17u + 57u + 4u * i, __brev(l) >> 28, 4);
  }
  // my codes
  {
    unsigned long long acc = 0;
    uint32_t nb = pos & 31u, word = pos >> 5;
    for (uint32_t j = 0; j < mine; j++) {
      const uint32_t e = s.tab[my[j]];
      acc |= (unsigned long long)(e & 0xffffu) << nb;
      nb += e >> 16;
      if (nb >= 32u) { atomicOr(s.out + word, (uint32_t)acc); word++; acc >>= 32; nb -= 32u; }
    }
    if (nb) atomicOr(s.out + word, (uint32_t)acc);
  }
  if (t == DT - 1u) {  // end of block, then the empty stored block: 000, pad to a byte, 00 00 FF FF
    uint32_t end = pos + bits;  // thread DT-1 is last in stream order (runs past L are empty)
    const uint32_t e = s.tab[256];
    or_bits(s.out, end, e & 0xffffu, e >> 16);
    end += (e >> 16) + 3u;
    const uint32_t byte = (end + 7u) >> 3;
    or_bits(s.out, (byte + 2u) * 8u, 0xffffu, 16);
  }
  __syncthreads();
  const uint8_t *ob = reinterpret_cast<const uint8_t *>(s.out);
  for (uint32_t i = t; i < size; i += DT) out[i] = ob[i];
}


// ---- CRC-32 of the compressed bytes (the IDAT chunk's check value, src/png-writer.ts:12-30, src/utils.ts:18-24) ----
// The register update is linear over GF(2): state' = state * x^(8 len) + R(M), with R(M) the remainder computed
// from a zero state.  So every 128-byte slice gets its own R by table lookups, a CTA folds its 256 slices in a
// tree (multiply the left half by x^(8 * its distance to the end), add), one thread moves the CTA's value to
// the end of the buffer the same way and XORs it into the result.  Bits are reflected as in zlib: bit 31 is
// x^0, multiplying by x is a right shift with the polynomial folded in.
constexpr uint32_t CRC_POLY = 0xedb88320u;
constexpr int CRC_SLICE = 128;               // bytes per thread
constexpr int CRC_CTA = DT * CRC_SLICE;      // 32 KiB per CTA
constexpr int CRC_POWS = 48;                 // x^(8 * 2^j), j < 48

struct CrcTables {
  uint32_t t[4][256];       // slicing-by-4
  uint32_t pow8[CRC_POWS];  // x^(8 * 2^j) mod P
};
__device__ CrcTables g_crc;

__host__ __device__ inline uint32_t crc_mulmod(uint32_t a, uint32_t b) {
  uint32_t p = 0;
  for (int j = 0; j < 32; j++) {  // a's coefficient of x^j sits in bit 31 - j
    if ((a >> (31 - j)) & 1u) p ^= b;
    b = (b >> 1) ^ ((b & 1u) ? CRC_POLY : 0u);  // b * x
  }
  return p;
}
// x^(8 n) mod P from the table of x^(8 * 2^j)
__host__ __device__ inline uint32_t crc_xpow8(const uint32_t *pow8, unsigned long long n) {
  uint32_t p = 0x80000000u;  // 1
  for (int j = 0; n != 0 && j < CRC_POWS; j++, n >>= 1)
    if (n & 1ull) p = crc_mulmod(p, pow8[j]);
  return p;
}

// result[0] = number of bytes (written by the scan kernel), result[2] ^= R(dst[0 .. n))
__global__ void __launch_bounds__(DT) crc32_kernel(const uint8_t *dst, unsigned long long *result) {
  __shared__ uint32_t tab[4][256];
  __shared__ uint32_t vals[DT];
  const unsigned long long n = result[0];
  const unsigned long long c0 = (unsigned long long)blockIdx.x * CRC_CTA;
  if (c0 >= n) return;
  const uint32_t t = threadIdx.x;
  for (uint32_t i = t; i < 1024u; i += DT) (&tab[0][0])[i] = (&g_crc.t[0][0])[i];
  __syncthreads();
  const unsigned long long avail = n - c0;
  const bool full = avail >= (unsigned long long)CRC_CTA;
  uint32_t c = 0u;
  if (full) {
    const uint4 *q = reinterpret_cast<const uint4 *>(dst + c0 + (size_t)t * CRC_SLICE);
#pragma unroll
    for (int k = 0; k < CRC_SLICE / 16; k++) {
      const uint4 w = q[k];
      const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
      for (int i = 0; i < 4; i++) {
        c ^= ww[i];
        c = tab[3][c & 0xffu] ^ tab[2][(c >> 8) & 0xffu] ^ tab[1][(c >> 16) & 0xffu] ^ tab[0][c >> 24];
      }
    }
  } else {
    // the tail: right-aligned in the CTA's window (leading zeros leave a zero state unchanged), byte by byte
    const long long first = (long long)t * CRC_SLICE - ((long long)CRC_CTA - (long long)avail);
    for (int k = 0; k < CRC_SLICE; k++) {
      const long long o = first + k;
      if (o >= 0) c = tab[0][(c ^ dst[c0 + (unsigned long long)o]) & 0xffu] ^ (c >> 8);
    }
  }
  vals[t] = c;
  __syncthreads();
#pragma unroll 1
  for (int k = 0; k < 8; k++) {  // slices 2^k apart: the left one is 128 * 2^k bytes further from the end
    const uint32_t stride = 1u << k;
    if ((t & (2u * stride - 1u)) == 0u) vals[t] = crc_mulmod(vals[t], g_crc.pow8[7 + k]) ^ vals[t + stride];
    __syncthreads();
  }
  if (t == 0u) {
    const unsigned long long after = full ? n - c0 - (unsigned long long)CRC_CTA : 0ull;
    const uint32_t v = crc_mulmod(vals[0], crc_xpow8(g_crc.pow8, after));
    atomicXor(reinterpret_cast<unsigned int *>(result + 2), v);
  }
}

CrcTables g_crc_host;

static void crc_tables_fill() {
  CrcTables &h = g_crc_host;
  for (uint32_t b = 0; b < 256; b++) {
    uint32_t c = b;
    for (int k = 0; k < 8; k++) c = (c >> 1) ^ ((c & 1u) ? CRC_POLY : 0u);
    h.t[0][b] = c;
  }
  for (uint32_t b = 0; b < 256; b++)
    for (int k = 1; k < 4; k++) h.t[k][b] = h.t[0][h.t[k - 1][b] & 0xffu] ^ (h.t[k - 1][b] >> 8);
  uint32_t p = 0x80000000u >> 8;  // x^8
  for (int j = 0; j < CRC_POWS; j++) { h.pow8[j] = p; p = crc_mulmod(p, p); }
}
void crc_tables_init() {   // host tables: once per process, whichever thread comes first
  static std::once_flag flag;
  std::call_once(flag, crc_tables_fill);
}

}  // namespace

// Compresses n_rows filtered rows that start at d_rows (row pitch job->out_pitch, filter byte at out_lead)
// into d_dst.  *out_bytes and *adler describe the result.  Ends with one small D2H read and a stream sync:
// the caller needs the size on the host to move the bytes on.
int stitch_deflate_band(stitch_job *job, const uint8_t *d_rows, uint32_t n_rows, size_t host_row, uint8_t *d_dst,
                        size_t *out_bytes, uint32_t *adler, uint32_t *crc_raw, uint32_t *launches) {
  stitch_ctx *ctx = job->ctx;
  int rc;
  BandView v;
  v.base = d_rows + job->out_lead; v.pitch = job->out_pitch; v.row_bytes = (uint32_t)host_row;
  v.total = (unsigned long long)host_row * n_rows;
  const uint32_t n_blocks = (uint32_t)((v.total + DB - 1) / DB);
  const size_t scratch = (size_t)n_blocks * (TAB * 4 + 4 + 8 + 8) + 96;
  {  // the CRC tables are a __device__ symbol and the encoder's shared memory an attribute: both per device
    static PerDevice once;
    int unused = 0;
    STITCH_CK(once.get(&unused, [](int *) {
      crc_tables_init();
      cudaError_t e = cudaMemcpyToSymbol(g_crc, &g_crc_host, sizeof(CrcTables));
      return e != cudaSuccess ? e : cudaFuncSetAttribute(deflate_encode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(EncSmem));
    }));
  }
  if ((rc = stitch_ensure_buffer(ctx, &job->d_defl_scratch, &job->defl_scratch_bytes, scratch))) return rc;
  if (!job->h_defl) STITCH_CK(cudaHostAlloc((void **)&job->h_defl, 32, cudaHostAllocDefault));
  unsigned long long *offsets = reinterpret_cast<unsigned long long *>(job->d_defl_scratch);
  unsigned long long *result = offsets + n_blocks;  // [0] bytes, [1] adler, [2] crc remainder, [3] spare
  uint32_t *adler_parts = reinterpret_cast<uint32_t *>(result + 4);
  uint32_t *sizes = adler_parts + 2 * (size_t)n_blocks;
  uint32_t *tabs = sizes + n_blocks;
  deflate_plan_kernel<<<n_blocks, DT, 0, ctx->stream>>>(v, tabs, sizes, adler_parts);
  STITCH_CK(cudaGetLastError());
  deflate_scan_kernel<<<1, 1024, 0, ctx->stream>>>(sizes, adler_parts, n_blocks, v.total, offsets, result);
  STITCH_CK(cudaGetLastError());
  deflate_encode_kernel<<<n_blocks, DT, sizeof(EncSmem), ctx->stream>>>(v, tabs, sizes, offsets, d_dst);
  STITCH_CK(cudaGetLastError());
  // CRC-32 remainder of the compressed bytes (they are still in L2)
  const size_t bound = stitch_deflate_bound((size_t)v.total);
  crc32_kernel<<<(unsigned)((bound + CRC_CTA - 1) / CRC_CTA), DT, 0, ctx->stream>>>(d_dst, result);
  STITCH_CK(cudaGetLastError());
  if ((rc = stitch_download_small(ctx, job->h_defl, result, 24))) return rc;   // by a kernel: not behind the downloads in flight
  STITCH_CK(cudaStreamSynchronize(ctx->stream));
  *out_bytes = (size_t)job->h_defl[0];
  *adler = (uint32_t)job->h_defl[1];
  *crc_raw = (uint32_t)job->h_defl[2];
  if (launches) *launches += 4;
  return STITCH_OK;
}

// R(A || B) from R(A), R(B) and len(B); and the standard CRC-32 (zlib.crc32) of n bytes from their remainder R.
uint32_t stitch_crc_raw_concat(uint32_t ra, uint32_t rb, uint64_t len_b) {
  crc_tables_init();
  return crc_mulmod(ra, crc_xpow8(g_crc_host.pow8, len_b)) ^ rb;
}
uint32_t stitch_crc_finish(uint32_t raw, uint64_t n) {
  crc_tables_init();
  return raw ^ crc_mulmod(0xffffffffu, crc_xpow8(g_crc_host.pow8, n)) ^ 0xffffffffu;
}

// worst case of stitch_deflate_band's output: every block stored
size_t stitch_deflate_bound(size_t raw_bytes) { return raw_bytes + (raw_bytes / DB + 1) * 10 + 64; }
