// inflate.cu -- SURVEY.md 8(f).2: the zlib streams of the input PNGs (concatenated IDAT payloads) are inflated
// on the device, so that tiles cross the bus compressed and the host's inflate (src/streaming-inflate.ts:23-76,
// src/decoders/png-decoder.ts:63-229 up to the point where rows are unfiltered) disappears.  The result is the
// inflated, still filtered stream that kernel (a) starts from, written straight into the tile's device buffer.
//
// DEFLATE decoding is serial inside a stream (a code's position depends on every code before it), so the
// parallelism across tiles is one stream per warp PAIR, and the work inside a stream is made parallel by
// speculation:
//
//   walk     a window of 32 x 768 code bits is cut into 32 subsequences; lane i walks subsequence i from a guessed
//            start and keeps its symbols as records.  Huffman codes resynchronise, so the walk usually ends on a true
//            code boundary; lanes whose start differs from their left neighbour's end walk again until nothing
//            changes (exact by induction from lane 0), else the consistent prefix of lanes is taken
//   apply    every lane writes its own records' bytes into a shared-memory ring holding the last 32 KiB of output;
//            copies from bytes another lane produces in the same window stay 16-bit pointers, resolved for all bytes
//            in parallel by pointer jumping; the ring goes to HBM in 16-byte granules per lane
//   pipeline the first warp of the pair walks window w+1 while the second applies window w (named barriers)
//
//   tables   per dynamic block, built by the warp in shared memory: canonical code from the code lengths
//            (count by length, stable rank by __match_any), an 11-bit single-symbol table filled symbol-parallel,
//            then an 11-bit MULTI-LITERAL table made from it by lookups: an entry carries up to three
//            literals when their codes fit in the index (filtered PNG rows are nearly all literals with 4-9 bit
//            codes, so one shared-memory round trip of the serial chain yields 2+ bytes); codes longer than the
//            index take a canonical bit-by-bit walk (count/first-code, RFC 1951 3.2.2)
//   checks   everything zlib's inflate checks, with zlib's messages (what pako reports through
//            `Inflate error: ${msg}`, streaming-inflate.ts:60-62): header, block type, stored lengths, code-length
//            sets (over-subscribed / incomplete), missing end-of-block, invalid codes, distance too far back, and
//            the Adler-32 trailer (a second, data-parallel kernel over the inflated bytes)
//
// Format references: RFC 1950, RFC 1951.
#include "job.h"

#include <cstring>

namespace stitch {
namespace {

constexpr int TB = 11;                 // literal/length table index bits
constexpr int TD = 9;                  // distance table index bits
constexpr int INF_STREAMS = 2;         // streams per CTA, two warps each: one walks the codes, one produces the bytes
// Speculative lane-parallel Huffman walk: a window of 32 x SUBBITS code bits is cut into 32 subsequences, one
// per lane; a lane keeps at most SUBCAP symbol records of its subsequence.
constexpr uint32_t SUBBITS = 768u;
constexpr uint32_t SUBCAP = 96u;
constexpr uint32_t QHEAD = 16u;          // record rows in front of a lane's list: a re-walked prefix may be longer than the one it replaces
constexpr uint32_t QROWS = SUBCAP + QHEAD;
constexpr uint32_t LANE_OUT = 256u;      // a lane stops once its records produce this many bytes (a match may add 257 more)
constexpr uint32_t RES = 0xffffu;        // src[]: this byte of the window is final
constexpr uint32_t ZWORDS = 32u * SUBBITS / 32u + 16u;   // staged code words: the window, its 127-bit lead-in, a symbol of overrun, one word of look-ahead
constexpr uint32_t MAX_ROUNDS = 8u;
constexpr uint32_t F_EOB = 1u << 8, F_ERR = 1u << 9;
constexpr uint32_t F_WIDE = 1u << 10;   // a merged record list that makes more than 512 bytes: its lane is left to the next window
// The last 32 KiB of output (the farthest a match may reach) live in a per-warp shared-memory ring, so a match
// copy never waits for L2; the ring is longer than that by one window's output and a flush unit, so nothing that is
// still to be read or written out is overwritten.  Global memory is written from the ring, 512 bytes (one uint4 per
// lane) at a time.
constexpr uint32_t OUTCAP = 8192u;   // most output bytes one window of subsequences may produce
constexpr uint32_t WIN = 32768u + OUTCAP + 512u;
static_assert(WIN % 16u == 0u, "ring: granule-aligned");
constexpr uint32_t FULL = 0xffffffffu;
constexpr uint32_t E_LONG = 0x30u;     // entry kind 3: code longer than the index, or no code at all

// entry (literal/length table): bits 0-3 code bits consumed, 4-5 kind (0 literals, 1 length, 2 end of block,
// 3 long/invalid); literals: 6-7 count, 8-31 up to three bytes; length: 8-11 extra bits, 16-24 base
// entry (distance table): bits 0-3 code bits, 4-5 kind (0 ok, 3 long/invalid), 8-11 extra bits, 16-31 base
__constant__ uint16_t c_lbase[31] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99,
                                     115, 131, 163, 195, 227, 258, 0, 0};
__constant__ uint8_t c_lext[31] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0, 0, 0};
__constant__ uint16_t c_dbase[32] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025,
                                     1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577, 0, 0};
__constant__ uint8_t c_dext[32] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13, 0, 0};
__constant__ uint8_t c_clorder[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};

struct Canon {            // one canonical Huffman code (RFC 1951 3.2.2)
  uint32_t cnt[16];       // codes per length
  uint32_t offs[17];      // index of the first symbol of a length in `sym`
  uint32_t next[16];      // first code of a length
};

// One message from the warp that walks the codes to the warp that produces the bytes.
enum MsgKind : uint32_t { MSG_WINDOW = 0, MSG_STORED = 1, MSG_END = 2 };
struct Msg {
  uint32_t kind;
  uint32_t n_lanes, total;   // WINDOW: lanes of the chain, bytes they produce
  uint32_t a, b, c;          // STORED: byte offset in the stream, length.  END: status, Adler-32 trailer, flags
  uint32_t n[32], off[32];   // WINDOW: records and first output byte of every lane
  uint32_t base[32];         // WINDOW: row of the lane's first record
};

struct WarpSmem {
  uint32_t lt[1 << TB];
  uint32_t dt[1 << TD];
  uint16_t l1[1 << TB];
  uint16_t lsym[288], dsym[32];
  Canon lc, dc;
  uint8_t lens[320];      // literal/length code lengths, then the distance code lengths
  uint8_t clt[128];       // code-length code: symbol | bits << 5
  uint32_t pad[3];
  __align__(16) uint32_t zw[ZWORDS];        // code words of the current window
  uint32_t q[2][QROWS * 32u];               // symbol records (two windows in flight), record k of lane i at q[.][(base_i + k) * 32 + i]
  uint16_t bpos[SUBCAP * 32u];              // first walk of the window: the bit each record starts at
  Msg msg[2];                               // what the walking warp hands to the applying warp with them
  __align__(16) uint16_t src[OUTCAP];       // per output byte of the window: RES, or the window byte it copies
  __align__(16) uint8_t win[WIN];
};
static_assert(sizeof(WarpSmem) % 16 == 0, "rings must stay 16-byte aligned");

// ring -> global, whole 512-byte units below `upto` (all of it with `all`); granules may run past `upto` into
// the 16 bytes of slack every destination buffer has
__device__ __forceinline__ void ring_flush(const uint8_t *win, uint8_t *dst, uint32_t &flushed, uint32_t upto, bool all, uint32_t lane) {
  while (flushed + 512u <= upto) {
    const uint32_t g = flushed + 16u * lane;
    *reinterpret_cast<uint4 *>(dst + g) = *reinterpret_cast<const uint4 *>(win + g % WIN);
    flushed += 512u;
  }
  if (all && flushed < upto) {
    const uint32_t g = flushed + 16u * lane;
    if (g < upto) *reinterpret_cast<uint4 *>(dst + g) = *reinterpret_cast<const uint4 *>(win + g % WIN);
    flushed = upto;
  }
}

// ---- bit reader: identical in every lane ------------------------------------------------------------------
struct Bits {
  const uint32_t *w;
  uint32_t last;   // index of the last readable word
  uint32_t wi;     // index of the next word to fetch
  uint32_t nx;     // fetched, not yet in bb
  unsigned long long bb;
  uint32_t bc;

  __device__ __forceinline__ void init(const uint8_t *src, uint32_t src_len, uint32_t byte_off) {
    w = reinterpret_cast<const uint32_t *>(src);
    last = src_len ? (src_len - 1u) >> 2 : 0u;
    const uint32_t k = byte_off >> 2, sh = 8u * (byte_off & 3u);
    bb = (unsigned long long)(__ldg(w + min(k, last)) >> sh);
    bc = 32u - sh;
    nx = __ldg(w + min(k + 1u, last));
    wi = k + 2u;
  }
  // at least 33 valid bits afterwards
  __device__ __forceinline__ void refill() {
    if (bc <= 32u) {
      bb |= (unsigned long long)nx << bc;
      bc += 32u;
      nx = __ldg(w + min(wi, last));
      wi++;
    }
  }
  __device__ __forceinline__ uint32_t peek(uint32_t n) const { return (uint32_t)bb & ((1u << n) - 1u); }
  __device__ __forceinline__ void drop(uint32_t n) { bb >>= n; bc -= n; }
  __device__ __forceinline__ uint32_t take(uint32_t n) { const uint32_t v = peek(n); drop(n); return v; }
  __device__ __forceinline__ unsigned long long consumed_bits() const { return (unsigned long long)(wi - 1u) * 32ull - bc; }
};

// Bit-by-bit canonical walk (codes are packed most significant bit first, RFC 1951 3.1.1).  Returns the symbol
// and its length, or -1 when no code of up to 15 bits matches (incomplete code).
__device__ __forceinline__ int canon_walk(uint32_t bits, const Canon &c, const uint16_t *sym, uint32_t &len_out) {
  int code = 0, first = 0, index = 0;
  for (uint32_t len = 1; len <= 15u; len++) {
    code |= (int)(bits & 1u);
    bits >>= 1;
    const int count = (int)c.cnt[len];
    if (code - count < first) { len_out = len; return sym[index + (code - first)]; }
    index += count; first += count;
    first <<= 1; code <<= 1;
  }
  return -1;
}

// Warp-wide: counts, validity (zlib inftrees.c: over-subscribed or incomplete sets are rejected, except an
// incomplete distance/length set that consists of a single code... zlib allows that for distances only),
// offsets, first codes, symbols sorted by (length, symbol).  Returns 0 ok, 1 over-subscribed, 2 incomplete.
__device__ int canon_build(const uint8_t *lens, uint32_t n, Canon &c, uint16_t *sym, uint32_t lane) {
  if (lane < 16u) c.cnt[lane] = 0u;
  __syncwarp();
  for (uint32_t i = lane; i < n; i += 32u) atomicAdd(&c.cnt[lens[i]], 1u);
  __syncwarp();
  int left = 1, bad = 0;
  uint32_t off = 0, code = 0;
  for (uint32_t len = 1; len <= 15u; len++) {   // redundantly in every lane (broadcast reads)
    const uint32_t k = c.cnt[len];
    left = (left << 1) - (int)k;
    if (left < 0) bad = 1;
    code = (code + (len > 1u ? c.cnt[len - 1u] : 0u)) << 1;
    if (lane == 0u) { c.offs[len] = off; c.next[len] = code; }
    off += k;
  }
  if (lane == 0u) { c.offs[16] = off; c.offs[0] = 0u; }
  __syncwarp();
  if (bad) return 1;
  // stable ranks inside chunks of 32 symbols; running offsets in registers of the lanes 1..15 (lane L owns length L)
  uint32_t run = lane >= 1u && lane <= 15u ? c.offs[lane] : 0u;
  for (uint32_t base = 0; base < n; base += 32u) {
    const uint32_t i = base + lane;
    const uint32_t l = i < n ? lens[i] : 0u;
    const uint32_t peers = __match_any_sync(FULL, l);
    const uint32_t rank = __popc(peers & ((1u << lane) - 1u));
    const uint32_t start = __shfl_sync(FULL, run, l);   // lane l holds the running offset of length l
    if (l) sym[start + rank] = (uint16_t)i;
    // lane L adds the number of symbols of length L seen in this chunk
    uint32_t add = 0;
    for (uint32_t L = 1; L <= 15u; L++) {
      const uint32_t m = __ballot_sync(FULL, l == L);
      if (lane == L) add = __popc(m);
    }
    run += add;
  }
  __syncwarp();
  return left > 0 ? 2 : 0;
}

// Symbol-parallel fill of a single-symbol table with `bits` index bits: `put(index, symbol, length)`.
template <class Put>
__device__ __forceinline__ void canon_fill(const Canon &c, const uint16_t *sym, const uint8_t *lens, int bits, uint32_t lane, Put put) {
  // symbols with at least 32 replicas: the warp shares each symbol's replicas
  const int wide = bits - 5;
  const uint32_t n_wide = wide >= 1 ? c.offs[wide + 1] : 0u;
  for (uint32_t i = 0; i < n_wide; i++) {
    const uint32_t s = sym[i], L = lens[s];
    const uint32_t code = c.next[L] + (i - c.offs[L]);
    const uint32_t r = __brev(code) >> (32u - L);
    for (uint32_t k = lane; k < (1u << (bits - L)); k += 32u) put(r | (k << L), s, L);
  }
  const uint32_t n_all = c.offs[bits + 1];
  for (uint32_t i = n_wide + lane; i < n_all; i += 32u) {
    const uint32_t s = sym[i], L = lens[s];
    const uint32_t code = c.next[L] + (i - c.offs[L]);
    const uint32_t r = __brev(code) >> (32u - L);
    for (uint32_t k = 0; k < (1u << (bits - L)); k++) put(r | (k << L), s, L);
  }
}

// Tables of one block from sm.lens[0..nlen) and sm.lens[nlen..nlen+ndist).  Returns 0 or an InfStatus.
__device__ int build_tables(WarpSmem &sm, uint32_t nlen, uint32_t ndist, uint32_t lane) {
  int rc = canon_build(sm.lens, nlen, sm.lc, sm.lsym, lane);
  // zlib (inftrees.c): an incomplete set is accepted only when it is a single one-bit code, or no code at all
  // (a block without matches has no distance code; the end-of-block check has already excluded the literal side)
  if (rc == 1 || (rc == 2 && !(sm.lc.offs[16] == 1u && sm.lc.cnt[1] == 1u))) return INF_E_LITLEN_SET;
  rc = canon_build(sm.lens + nlen, ndist, sm.dc, sm.dsym, lane);
  if (rc == 1 || (rc == 2 && !(sm.dc.offs[16] == 0u || (sm.dc.offs[16] == 1u && sm.dc.cnt[1] == 1u)))) return INF_E_DIST_SET;
  for (uint32_t i = lane; i < (1u << TB); i += 32u) sm.l1[i] = 0;
  for (uint32_t i = lane; i < (1u << TD); i += 32u) sm.dt[i] = E_LONG;
  __syncwarp();
  uint16_t *l1 = sm.l1;
  canon_fill(sm.lc, sm.lsym, sm.lens, TB, lane, [l1](uint32_t idx, uint32_t s, uint32_t L) { l1[idx] = (uint16_t)(L | (s << 4)); });
  uint32_t *dt = sm.dt;
  canon_fill(sm.dc, sm.dsym, sm.lens + nlen, TD, lane, [dt](uint32_t idx, uint32_t s, uint32_t L) {
    dt[idx] = s < 30u ? (L | ((uint32_t)c_dext[s] << 8) | ((uint32_t)c_dbase[s] << 16)) : E_LONG;
  });
  __syncwarp();
  // multi-literal table from the single-symbol one
  for (uint32_t idx = lane; idx < (1u << TB); idx += 32u) {
    const uint32_t e1 = l1[idx];
    uint32_t e = E_LONG;
    if (e1) {
      const uint32_t L1 = e1 & 15u, s1 = e1 >> 4;
      if (s1 < 256u) {
        uint32_t nb = L1, cnt = 1u, lits = s1;
        const uint32_t e2 = l1[idx >> nb];
        if (e2 && (e2 >> 4) < 256u && nb + (e2 & 15u) <= (uint32_t)TB) {
          lits |= (e2 >> 4) << 8; nb += e2 & 15u; cnt = 2u;
          const uint32_t e3 = l1[idx >> nb];
          if (e3 && (e3 >> 4) < 256u && nb + (e3 & 15u) <= (uint32_t)TB) { lits |= (e3 >> 4) << 16; nb += e3 & 15u; cnt = 3u; }
        }
        e = nb | (cnt << 6) | (lits << 8);
      } else if (s1 == 256u) {
        e = L1 | 0x20u;
      } else if (s1 < 286u) {
        e = L1 | 0x10u | ((uint32_t)c_lext[s1 - 257u] << 8) | ((uint32_t)c_lbase[s1 - 257u] << 16);
      }
    }
    sm.lt[idx] = e;
  }
  __syncwarp();
  return 0;
}

// Dynamic block header (RFC 1951 3.2.7): code lengths into sm.lens.  Uniform in all lanes.
__device__ int read_dynamic_header(WarpSmem &sm, Bits &br, uint32_t lane, uint32_t &nlen, uint32_t &ndist) {
  br.refill();
  nlen = br.take(5) + 257u;
  ndist = br.take(5) + 1u;
  const uint32_t ncode = br.take(4) + 4u;
  if (nlen > 286u || ndist > 30u) return INF_E_TOO_MANY;
  uint8_t *cl = sm.lens + 300;  // 19 code-length code lengths live in the tail of lens for a moment
  if (lane < 19u) cl[lane] = 0;
  __syncwarp();
  for (uint32_t i = 0; i < ncode; i++) {
    br.refill();
    const uint32_t v = br.take(3);
    if (lane == 0u) cl[c_clorder[i]] = (uint8_t)v;
  }
  __syncwarp();
  // the code-length code: counts and first codes redundantly per lane, 7-bit table filled by the symbols' lanes
  uint32_t cnt[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (uint32_t i = 0; i < 19u; i++) cnt[cl[i]]++;
  int left = 1;
  uint32_t next[8], code = 0;
  next[0] = 0;
  for (uint32_t len = 1; len <= 7u; len++) {
    left = (left << 1) - (int)cnt[len];
    if (left < 0) return INF_E_CODE_LENGTHS;
    code = (code + (len > 1u ? cnt[len - 1u] : 0u)) << 1;
    next[len] = code;
  }
  if (left > 0) return INF_E_CODE_LENGTHS;  // zlib: incomplete code-length code is an error (type CODES)
  for (uint32_t i = lane; i < 128u; i += 32u) sm.clt[i] = 0;
  __syncwarp();
  if (lane < 19u) {
    const uint32_t L = cl[lane];
    if (L) {
      uint32_t c = next[L];
      for (uint32_t j = 0; j < lane; j++) c += cl[j] == L ? 1u : 0u;
      const uint32_t r = __brev(c) >> (32u - L);
      for (uint32_t k = 0; k < (1u << (7u - L)); k++) sm.clt[r | (k << L)] = (uint8_t)(lane | (L << 5));
    }
  }
  __syncwarp();
  // the nlen + ndist code lengths, run-length coded
  const uint32_t total = nlen + ndist;
  uint32_t i = 0, prev = 0;
  while (i < total) {
    br.refill();
    const uint32_t e = sm.clt[br.peek(7)];
    if (e == 0u) return INF_E_CODE_LENGTHS;
    br.drop(e >> 5);
    const uint32_t s = e & 31u;
    if (s < 16u) {
      if (lane == 0u) sm.lens[i] = (uint8_t)s;
      prev = s; i++;
    } else {
      uint32_t rep, val = 0;
      if (s == 16u) {
        if (i == 0u) return INF_E_BIT_REPEAT;
        val = prev; rep = 3u + br.take(2);
      } else if (s == 17u) rep = 3u + br.take(3);
      else rep = 11u + br.take(7);
      if (i + rep > total) return INF_E_BIT_REPEAT;
      if (lane < rep) sm.lens[i + lane] = (uint8_t)val;
      if (lane + 32u < rep) sm.lens[i + lane + 32u] = (uint8_t)val;
      if (lane + 64u < rep) sm.lens[i + lane + 64u] = (uint8_t)val;
      if (lane + 96u < rep) sm.lens[i + lane + 96u] = (uint8_t)val;
      if (lane + 128u < rep) sm.lens[i + lane + 128u] = (uint8_t)val;
      prev = val; i += rep;
    }
  }
  __syncwarp();
  if (sm.lens[256] == 0) return INF_E_NO_EOB;
  return 0;
}


// ---- lane-private bit reader over the staged window ---------------------------------------------------------
// A lane's place in the staged code words: just the bit position.  Every symbol fetches the 64 bits from there
// (three shared-memory words, two funnel shifts) -- fewer dependent instructions per symbol than keeping a shifting
// bit buffer up to date, and the dependent chain is what a single walking warp is bound by.
struct LaneBits {
  const uint32_t *zw;
  uint32_t p;
  __device__ __forceinline__ void init(const uint32_t *z, uint32_t rel_bit) { zw = z; p = rel_bit; }
  __device__ __forceinline__ uint32_t pos() const { return p; }
};

// One lane walks the codes from window bit `start` until it reaches `limit`, the end of the block, an invalid code,
// or SUBCAP records.  Records: literals = count << 24 | up to three bytes; match = 1 << 31 | (length - 3) << 15 |
// (distance - 1).  Returns the bit after the last symbol taken; `n_flag` = records | F_EOB | F_ERR | status << 16.
// One symbol at the lane's position: 0 = a record (literals or a match) and the bytes it makes, 1 = end of block,
// 2 = invalid code (`flag` says which).
__device__ __forceinline__ uint32_t walk_one(WarpSmem &sm, LaneBits &lb, const bool live, uint32_t &rec, uint32_t &add, uint32_t &flag) {
  // Written without branches on the symbol's kind: the lanes of a warp step together, and with 32 of them some lane
  // has a match in nearly every step, so the distance lookup is paid anyway -- selects are cheaper than the two
  // diverging paths.  Only codes longer than the table index branch off.  A lane that is not `live` goes through
  // the same instructions and takes no bits.
  const uint32_t wi = lb.p >> 5, sh = lb.p & 31u;
  const uint32_t w0 = lb.zw[wi], w1 = lb.zw[wi + 1u], w2 = lb.zw[wi + 2u];
  const uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);   // the 64 bits from p on
  uint32_t e = sm.lt[lo & ((1u << TB) - 1u)];
  if (live && (e & E_LONG) == E_LONG) {
    uint32_t L = 0;
    const int sy = canon_walk(lo, sm.lc, sm.lsym, L);
    if (sy < 0 || sy >= 286) { flag = F_ERR | ((uint32_t)INF_E_LITLEN_CODE << 16); return 2u; }
    if (sy < 256) e = L | (1u << 6) | ((uint32_t)sy << 8);
    else if (sy == 256) e = L | 0x20u;
    else e = L | 0x10u | ((uint32_t)c_lext[sy - 257] << 8) | ((uint32_t)c_lbase[sy - 257] << 16);
  }
  const uint32_t nb = e & 15u, kind = (e >> 4) & 3u;
  const bool is_len = live && kind == 1u;
  const uint32_t lext = is_len ? (e >> 8) & 15u : 0u;
  const uint32_t len = (e >> 16) + ((lo >> nb) & ((1u << lext) - 1u));    // code <= 15 bits, extra <= 5
  const uint32_t k = nb + lext;
  const uint32_t dbits = __funnelshift_r(lo, hi, k);                      // the 32 bits after the length
  uint32_t d = sm.dt[dbits & ((1u << TD) - 1u)];
  if (is_len && (d & E_LONG) == E_LONG) {
    uint32_t L = 0;
    const int sy = canon_walk(dbits, sm.dc, sm.dsym, L);
    if (sy < 0 || sy >= 30) { flag = F_ERR | ((uint32_t)INF_E_DIST_CODE << 16); return 2u; }
    d = L | ((uint32_t)c_dext[sy] << 8) | ((uint32_t)c_dbase[sy] << 16);
  }
  const uint32_t dn = d & 15u, dext = (d >> 8) & 15u;
  const uint32_t dist = (d >> 16) + ((dbits >> dn) & ((1u << dext) - 1u));  // code <= 15 bits, extra <= 13
  const uint32_t cnt = (e >> 6) & 3u;
  rec = is_len ? (1u << 31) | ((len - 3u) << 15) | (dist - 1u) : (cnt << 24) | (e >> 8);
  add = is_len ? len : cnt;
  lb.p += live ? (is_len ? k + dn + dext : nb) : 0u;
  if (live && kind == 2u) { flag = F_EOB; return 1u; }
  return 0u;
}

__device__ __forceinline__ uint32_t rec_bytes(uint32_t rec) { return (int)rec >= 0 ? rec >> 24 : ((rec >> 15) & 0xffu) + 3u; }

// `mask`: the lanes that walk in this round.  They step together: the loop condition is a vote, so the lanes meet
// again at every symbol instead of each running its own copy of the loop.  `first`: the window's first walk, which
// also notes where every record starts.
// Every lane of the warp calls (the votes are full-warp votes); only the `active` ones walk.  (A step that does
// one table lookup -- literal/length OR distance, as a two-state machine -- was measured: cheaper per step, but the
// steps of a round are then max over lanes of symbols + matches, and the round got 25 % slower.)
__device__ __forceinline__ uint32_t walk_sub(WarpSmem &sm, uint32_t *qbuf, uint32_t start, uint32_t limit, uint32_t lane, bool active,
                                             bool first, uint32_t old_end, uint32_t &n_flag, uint32_t &n_out) {
  LaneBits lb;
  lb.init(sm.zw, start);
  uint32_t n = 0, flag = 0, out = 0;
  uint32_t *q = qbuf + QHEAD * 32u + lane;
  uint16_t *bp = sm.bpos + lane;
  bool live = active;
  for (;;) {
    const uint32_t at = lb.pos();
    live = live && at < limit && n < SUBCAP && out < LANE_OUT;
    if (!__any_sync(FULL, live)) break;
    uint32_t rec, add;
    live = walk_one(sm, lb, live, rec, add, flag) == 0u && live;
    if (live) {
      q[n * 32u] = rec;
      if (first) bp[n * 32u] = (uint16_t)at;
    }
    n += live ? 1u : 0u;
    out += live ? add : 0u;
  }
  if (!active) return old_end;
  n_flag = n | flag;
  n_out = out;
  return lb.pos();
}

// What a lane keeps of its FIRST walk of the window, for the walks that follow: the records in rows QHEAD.. of its
// column with the bit each starts at (bpos), of which those from `valid` on are still in place.
struct FirstWalk {
  uint32_t n, out, flag, end;   // records, bytes they make, F_EOB / F_ERR | status, bit after the last symbol
  uint32_t valid;               // records below this index have been overwritten by an earlier merge
};

// A later walk of the window: from the lane's corrected start, and only until it steps on a bit where a (still
// intact) record of the first walk starts -- from there on the two walks are the same walk.  The new records are
// then moved in front of the old tail; `base` is the row the merged list starts at.  A walk that never
// meets the first one replaces it (and the lane has no first walk to meet any more).
__device__ __forceinline__ uint32_t walk_again(WarpSmem &sm, uint32_t *qbuf, uint32_t start, uint32_t limit, uint32_t lane, bool active,
                                               FirstWalk &fw, uint32_t old_end, uint32_t &n_flag, uint32_t &n_out, uint32_t &base) {
  LaneBits lb;
  lb.init(sm.zw, start);
  const uint32_t n_old = fw.n;
  uint32_t m = 0, flag = 0, out = 0, j = fw.valid;
  uint32_t *q = qbuf + lane;
  const uint16_t *bp = sm.bpos + lane;
  // The new records go to rows 0.. of the lane's column, the first walk's live in rows QHEAD..: the first walk's
  // record j is still needed as long as the new walk has not passed it, so row m may be written while m < QHEAD + j.
  // A new walk that makes so many more records than the old one gives up meeting it (`alone`).
  bool live = active, met = false, alone = false;
  for (;;) {
    {
      const uint32_t at = lb.pos();
      live = live && at < limit && m < SUBCAP && out < LANE_OUT;
      if (live && !alone) {
        // both walks move forward: j follows by at most two records a step (once the walks have met they advance a
        // record a step each, so a j that lags catches up and the meeting is seen a little later at worst)
        j += j < n_old && bp[j * 32u] < at ? 1u : 0u;
        j += j < n_old && bp[j * 32u] < at ? 1u : 0u;
        if (j < n_old && bp[j * 32u] == at) { met = true; live = false; }
        alone = m >= QHEAD + j;
      }
    }
    if (!__any_sync(FULL, live)) break;
    uint32_t rec, add;
    live = walk_one(sm, lb, live, rec, add, flag) == 0u && live;
    if (live) q[m * 32u] = rec;
    m += live ? 1u : 0u;
    out += live ? add : 0u;
  }
  if (!active) return old_end;
  if (met) {
    uint32_t tail = 0;   // bytes of the first walk's records from j on (those rows are untouched)
    for (uint32_t k = j; k < n_old; k++) tail += rec_bytes(q[(QHEAD + k) * 32u]);
    base = QHEAD + j - m;   // m <= QHEAD + j: every row was written under that condition
    if (base) for (uint32_t t = m; t-- > 0u;) q[(base + t) * 32u] = q[t * 32u];
    fw.valid = j;
    n_flag = (m + n_old - j) | fw.flag;
    n_out = out + tail;
    if (n_out > 512u) n_flag |= F_WIDE;   // the applier packs a match's offset inside its lane into 9 bits
    return fw.end;
  }
  fw.n = 0u;   // the first walk's records are gone
  fw.valid = 0u;
  base = 0u;
  n_flag = m | flag;
  n_out = out;
  return lb.pos();
}

// ---- two warps per stream ---------------------------------------------------------------------------------------
// The WALKER reads headers, builds tables and walks the codes of a window into records; the APPLIER turns the
// records of the window before into bytes (ring, pointer jumping, HBM) meanwhile.  They meet at four named
// barriers per stream (slot s of 2: FULL_s "records are there", EMPTY_s "slot may be refilled"); a barrier counts 64
// threads, 32 that arrive and go on and 32 that wait.
__device__ __forceinline__ void bar_wait(uint32_t id) { asm volatile("bar.sync %0, 64;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void bar_signal(uint32_t id) { asm volatile("bar.arrive %0, 64;" ::"r"(id) : "memory"); }

struct Link {
  uint32_t full0, empty0;   // barrier ids of slot 0; slot 1 is +1
  uint32_t slot;
};

__device__ void walker(WarpSmem &sm, const InfStream &st, Link lk, const uint32_t lane) {
  const uint32_t cap = st.dst_cap;
  unsigned long long opos = 0;   // bytes produced so far (the applier's `pos` after it has caught up)
  int err = 0;
  uint32_t adler_stored = 0, flags = 0;
  Bits br;
  // every message: wait until the slot is free, fill it, hand it over
  auto acquire = [&]() -> Msg & { bar_wait(lk.empty0 + lk.slot); return sm.msg[lk.slot]; };
  auto publish = [&]() { __syncwarp(); bar_signal(lk.full0 + lk.slot); lk.slot ^= 1u; };

  if (st.src_len < 2u) { err = INF_E_EOF; goto done; }
  br.init(st.src, st.src_len, 0u);
  {
    const uint32_t cmf = br.take(8), flg = br.take(8);
    if (((cmf << 8) | flg) % 31u) { err = INF_E_HEADER; goto done; }
    if ((cmf & 15u) != 8u) { err = INF_E_METHOD; goto done; }
    if ((cmf >> 4) > 7u) { err = INF_E_WINDOW; goto done; }
    if (flg & 0x20u) { err = INF_E_DICT; goto done; }
  }
  for (;;) {  // blocks
    br.refill();
    const uint32_t bfinal = br.take(1), btype = br.take(2);
    if (btype == 3u) { err = INF_E_BLOCK_TYPE; goto done; }
    if (btype == 0u) {
      br.drop(br.bc & 7u);
      br.refill();
      const uint32_t len = br.take(16);
      br.refill();
      const uint32_t nlen = br.take(16);
      if ((len ^ 0xffffu) != nlen) { err = INF_E_STORED; goto done; }
      const unsigned long long at = br.consumed_bits() >> 3;
      if (at + len > st.src_len) { err = INF_E_EOF; goto done; }
      Msg &m = acquire();
      if (lane == 0u) { m.kind = MSG_STORED; m.a = (uint32_t)at; m.b = len; }
      publish();
      opos += len;
      br.init(st.src, st.src_len, (uint32_t)at + len);
    } else {
      if (btype == 1u) {
        for (uint32_t i = lane; i < 288u; i += 32u) sm.lens[i] = i < 144u ? 8 : i < 256u ? 9 : i < 280u ? 7 : 8;
        sm.lens[288 + lane] = 5;  // 32 distance codes of 5 bits (30 and 31 never valid)
        __syncwarp();
        err = build_tables(sm, 288u, 32u, lane);
      } else {
        uint32_t nlen, ndist;
        err = read_dynamic_header(sm, br, lane, nlen, ndist);
        if (!err) err = build_tables(sm, nlen, ndist, lane);
      }
      if (err) goto done;
      // ---- symbols of the block, a window of 32 subsequences at a time ----
      // Lane i walks the codes of subsequence i from a guessed start (a multiple of SUBBITS; lane 0's start is
      // exact).  Huffman codes resynchronise: a walk that starts off a code boundary usually falls onto real
      // boundaries before its subsequence ends, so its END is right although its start was not.  Each round,
      // the lanes whose start differs from their left neighbour's end walk again from that end; when no start
      // changes any more every lane's records are the true symbols (by induction from lane 0).  After MAX_ROUNDS
      // the consistent prefix of lanes is used and the next window starts where it ended.
      unsigned long long bitpos = br.consumed_bits();
      bool eob = false;
      const uint32_t last_gran = (st.src_len - 1u) >> 4;
      while (!eob) {
        const uint32_t a0 = (uint32_t)(bitpos >> 7);         // 16-byte granule the window starts in
        const uint32_t rel0 = (uint32_t)(bitpos & 127ull);
        for (uint32_t t = lane; t < ZWORDS / 4u; t += 32u)
          reinterpret_cast<uint4 *>(sm.zw)[t] = __ldg(reinterpret_cast<const uint4 *>(st.src) + min(a0 + t, last_gran));
        {  // the window after this one most likely starts 32 subsequences on: have its words in L1 by then
          const uint32_t g = min(a0 + 32u * SUBBITS / 128u + 8u * lane, last_gran);
          if (lane < ZWORDS / 32u + 1u) asm volatile("prefetch.global.L1 [%0];" ::"l"(reinterpret_cast<const uint4 *>(st.src) + g));
        }
        Msg &m = acquire();    // also orders the staging stores before the walks
        uint32_t *qbuf = sm.q[lk.slot];
        __syncwarp();
        uint32_t start = rel0 + lane * SUBBITS;
        const uint32_t limit = rel0 + (lane + 1u) * SUBBITS;
        uint32_t end = 0, n_flag = 0, n_out = 0, base = QHEAD;
        bool walk = true;
        FirstWalk fw;
        uint32_t bad = 0;   // lanes whose start is not their left neighbour's end
        for (uint32_t round = 0;; round++) {
          if (round == 0u) {
            end = walk_sub(sm, qbuf, start, limit, lane, true, true, end, n_flag, n_out);
            fw.n = n_flag & 0xffu; fw.out = n_out; fw.flag = n_flag & ~0xffu; fw.end = end; fw.valid = 0u;
          } else {
            end = walk_again(sm, qbuf, start, limit, lane, walk, fw, end, n_flag, n_out, base);
          }
          __syncwarp();
          const uint32_t left_end = __shfl_up_sync(FULL, end, 1);
          const uint32_t left_flag = __shfl_up_sync(FULL, n_flag, 1);
          // nothing follows a subsequence that ended the block or failed: lanes to its right are not part of the chain
          const bool chained = lane > 0u && (left_flag & (F_EOB | F_ERR)) == 0u;
          walk = chained && start != left_end;
          bad = __ballot_sync(FULL, walk);
          if (bad == 0u || round + 1u >= MAX_ROUNDS) break;
          if (walk) start = left_end;
        }
        // lanes of the chain: up to the first stop (inclusive) and before the first inconsistent start
        const uint32_t stops = __ballot_sync(FULL, (n_flag & (F_EOB | F_ERR)) != 0u);
        uint32_t n_lanes = bad ? (uint32_t)__ffs(bad) - 1u : 32u;
        if (stops) n_lanes = min(n_lanes, (uint32_t)__ffs(stops));
        const uint32_t wide = __ballot_sync(FULL, (n_flag & F_WIDE) != 0u);   // never lane 0: it does not walk twice
        if (wide) n_lanes = min(n_lanes, (uint32_t)__ffs(wide) - 1u);
        uint32_t off = n_out;   // inclusive scan of the lanes' output sizes
#pragma unroll
        for (uint32_t d = 1; d < 32u; d <<= 1) {
          const uint32_t v = __shfl_up_sync(FULL, off, d);
          if (lane >= d) off += v;
        }
        n_lanes = min(n_lanes, (uint32_t)__popc(__ballot_sync(FULL, off <= OUTCAP)));  // sizes are positive sums: a prefix of lanes
        const uint32_t total = __shfl_sync(FULL, off, n_lanes - 1u);
        m.n[lane] = n_flag & 0xffu;
        m.off[lane] = off - n_out;
        m.base[lane] = base;
        if (lane == 0u) { m.kind = MSG_WINDOW; m.n_lanes = n_lanes; m.total = total; }
        publish();
        opos += total;
        const uint32_t last_flag = __shfl_sync(FULL, n_flag, n_lanes - 1u);
        bitpos = ((unsigned long long)a0 << 7) + __shfl_sync(FULL, end, n_lanes - 1u);
        if (last_flag & F_ERR) { err = (int)(last_flag >> 16); break; }
        eob = (last_flag & F_EOB) != 0u;
        if (bitpos > 8ull * st.src_len) { err = INF_E_EOF; break; }
        if (opos > cap) { flags |= 1u; break; }  // more than the caller asked for: the rest is not needed
      }
      if (err || (flags & 1u)) goto done;
      br.init(st.src, st.src_len, (uint32_t)(bitpos >> 3));
      br.drop((uint32_t)(bitpos & 7ull));
    }
    if (br.consumed_bits() > 8ull * st.src_len) { err = INF_E_EOF; goto done; }
    if (bfinal) break;
    if (opos > cap) { flags |= 1u; goto done; }
  }
  {  // Adler-32 trailer (RFC 1950), big-endian, on the next byte boundary
    const unsigned long long at = (br.consumed_bits() + 7ull) >> 3;
    if (at + 4ull > st.src_len) { err = INF_E_EOF; goto done; }
    const uint8_t *t = st.src + at;
    adler_stored = ((uint32_t)__ldg(t) << 24) | ((uint32_t)__ldg(t + 1) << 16) | ((uint32_t)__ldg(t + 2) << 8) | __ldg(t + 3);
    flags |= 2u;  // trailer read: the Adler kernel compares
  }
done:
  Msg &m = acquire();
  if (lane == 0u) { m.kind = MSG_END; m.a = (uint32_t)err; m.b = adler_stored; m.c = flags; }
  publish();
}

__device__ void applier(WarpSmem &sm, const InfStream &st, InfResult *results, Link lk, const uint32_t lane) {
  uint8_t *dst = st.dst;
  uint8_t *win = sm.win;
  uint16_t *src = sm.src;
  const uint32_t cap = st.dst_cap;
  uint32_t pos = 0, flushed = 0, wpos = 0;  // wpos: pos modulo the ring
  uint32_t err = 0, end_status = 0, adler_stored = 0, flags = 0;
  bar_signal(lk.empty0);       // both slots start empty
  bar_signal(lk.empty0 + 1u);
  for (;;) {
    bar_wait(lk.full0 + lk.slot);
    const Msg &m = sm.msg[lk.slot];
    const uint32_t kind = m.kind;
    if (kind == MSG_END) { end_status = m.a; adler_stored = m.b; flags = m.c; break; }
    if (kind == MSG_STORED) {
      const uint8_t *s = st.src + m.a;
      const uint32_t len = m.b;
      bar_signal(lk.empty0 + lk.slot);
      lk.slot ^= 1u;
      if (err) continue;
      for (uint32_t done = 0; done < len;) {   // through the ring, a piece at a time (later matches may reach back into it)
        const uint32_t n = min(len - done, 4096u);
        for (uint32_t j = lane; j < n; j += 32u) win[(pos + j) % WIN] = __ldg(s + done + j);
        pos += n; done += n;
        __syncwarp();
        ring_flush(win, dst, flushed, min(pos, cap), false, lane);
        __syncwarp();
      }
      wpos = pos % WIN;
      continue;
    }
    // ---- a window ----
    // Every byte of the window starts out final.  Pass L: every lane of the chain writes the literals of its own
    // records and moves its matches, with the offset they write to, to the head of its own record column.  Pass M:
    // the matches, in order.  (One pass over the records would run the match code in every step in which any lane
    // has a match -- nearly all of them -- although a lane has only a handful.)  A match that copies from bytes
    // another lane is producing in this same window leaves 16-bit pointers, resolved afterwards by pointer jumping.
    const uint32_t n_lanes = m.n_lanes, total = m.total;
    const uint32_t off = m.off[lane], ni = m.n[lane], base = m.base[lane];
    uint32_t *col = sm.q[lk.slot] + base * 32u + lane;
    bool far = false;
    uint32_t plo = OUTCAP, phi = 0;   // the pointers this lane leaves lie in [plo, phi)
    for (uint32_t t = lane; t * 8u < total; t += 32u) reinterpret_cast<uint4 *>(src)[t] = make_uint4(~0u, ~0u, ~0u, ~0u);
    __syncwarp();
    {
      uint32_t o = off, nm = 0;
      const uint32_t my_n = lane < n_lanes && !err ? ni : 0u;
      for (uint32_t k = 0; __any_sync(FULL, k < my_n); k++) {   // a vote per record: the lanes stay in step
        const bool on = k < my_n;            // selects and predicated stores, no diverging paths
        const uint32_t r = on ? col[k * 32u] : 0u;
        const bool lit = on && (int)r >= 0, mat = on && (int)r < 0;
        const uint32_t cnt = r >> 24;
        uint32_t s0 = wpos + o;
        s0 = s0 >= WIN ? s0 - WIN : s0;
        const uint32_t s1 = s0 + 1u == WIN ? 0u : s0 + 1u, s2 = s1 + 1u == WIN ? 0u : s1 + 1u;
        if (lit) win[s0] = (uint8_t)r;
        if (lit && cnt > 1u) win[s1] = (uint8_t)(r >> 8);
        if (lit && cnt > 2u) win[s2] = (uint8_t)(r >> 16);
        const uint32_t orel = o - off;   // < 512: records start below LANE_OUT, merged lists are kept below by the walker
        far = far || (mat && orel >= 512u);
        if (mat) col[nm * 32u] = (r << 9) | (orel & 511u);   // (length - 3) << 24 | (distance - 1) << 9 | offset; slot nm <= k is consumed
        nm += mat ? 1u : 0u;
        o += lit ? cnt : mat ? ((r >> 15) & 0xffu) + 3u : 0u;
      }
      for (uint32_t i = 0; __any_sync(FULL, i < nm); i++) {
        if (i >= nm) continue;
        const uint32_t c = col[i * 32u];
        const uint32_t len = (c >> 24) + 3u, dist = ((c >> 9) & 0x7fffu) + 1u;
        o = off + (c & 511u);
        far = far || dist > pos + o;
        // The byte copied lies `dist` back in the ring whether it is history or was made in this window.  History is
        // final; a byte this lane made itself is final or a pointer to follow; another lane's byte is pointed to.
        const uint32_t o0 = o;
        int rel = (int)o - (int)dist;
        uint32_t sl = wpos + o;
        sl = sl >= WIN ? sl - WIN : sl;
        int sp = (int)wpos + rel;
        sp = sp < 0 ? sp + (int)WIN : sp >= (int)WIN ? sp - (int)WIN : sp;
        bool left = false;   // pointers left behind
        for (uint32_t t = 0; t < len; t++) {
          uint32_t p = rel < 0 ? RES : (uint32_t)rel;
          if (rel >= (int)off) p = src[rel];
          if (p == RES) win[sl] = win[sp];
          else { src[o] = (uint16_t)p; left = true; }
          o++; rel++;
          sl = sl + 1u == WIN ? 0u : sl + 1u;
          sp = sp + 1 == (int)WIN ? 0 : sp + 1;
        }
        if (left) { plo = min(plo, o0); phi = max(phi, o); }
      }
    }
    __syncwarp();
    bar_signal(lk.empty0 + lk.slot);   // the records are consumed: the walker may refill the slot
    lk.slot ^= 1u;
    if (err) continue;
    if (__any_sync(FULL, far)) { err = INF_E_FAR; continue; }
    plo = __reduce_min_sync(FULL, plo) & ~31u;
    phi = __reduce_max_sync(FULL, phi);
    while (plo < phi) {   // pointer jumping: a pointer always leads to a lower lane's region, so at most 5 rounds
      uint32_t olo = OUTCAP, ohi = 0;   // the pointers still open after this round
      for (uint32_t x0 = plo; x0 < phi; x0 += 64u) {   // two groups of 32 a step: their loads overlap
        uint32_t x[2], p[2], qn[2];
        uint8_t v[2];
#pragma unroll
        for (int g = 0; g < 2; g++) {
          x[g] = x0 + 32u * g + lane;
          p[g] = x[g] < phi ? src[x[g]] : RES;
        }
        if (!__any_sync(FULL, p[0] != RES || p[1] != RES)) continue;   // no pointer here
#pragma unroll
        for (int g = 0; g < 2; g++) {
          const uint32_t pp = p[g] == RES ? 0u : p[g];
          qn[g] = src[pp];
          uint32_t ps = wpos + pp;
          ps = ps >= WIN ? ps - WIN : ps;
          v[g] = win[ps];
        }
        __syncwarp();
#pragma unroll
        for (int g = 0; g < 2; g++) {
          if (p[g] != RES) {
            if (qn[g] == RES) {
              uint32_t sl = wpos + x[g];
              sl = sl >= WIN ? sl - WIN : sl;
              win[sl] = v[g];
              src[x[g]] = (uint16_t)RES;
            } else {
              src[x[g]] = (uint16_t)qn[g];
              olo = min(olo, x[g]); ohi = max(ohi, x[g] + 1u);
            }
          }
        }
        __syncwarp();
      }
      plo = __reduce_min_sync(FULL, olo) & ~31u;
      phi = __reduce_max_sync(FULL, ohi);
    }
    pos += total; wpos += total;
    if (wpos >= WIN) wpos -= WIN;
    ring_flush(win, dst, flushed, min(pos, cap), false, lane);
  }
  __syncwarp();
  ring_flush(win, dst, flushed, min(pos, cap), true, lane);
  if (lane == 0u) {
    InfResult r;
    r.status = err ? err : end_status; r.out_len = pos < cap ? pos : cap; r.adler_stored = adler_stored;
    r.flags = err ? 0u : flags;
    results[st.slot] = r;
  }
}

__global__ void __launch_bounds__(INF_STREAMS * 64) inflate_kernel(const InfStream *streams, InfResult *results, uint32_t n_streams) {
  extern __shared__ __align__(16) unsigned char inf_smem_raw[];
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  const uint32_t pair = warp >> 1;
  const uint32_t sid = blockIdx.x * INF_STREAMS + pair;
  if (sid >= n_streams) return;
  WarpSmem &sm = reinterpret_cast<WarpSmem *>(inf_smem_raw)[pair];
  const InfStream st = streams[sid];
  Link lk;
  lk.full0 = 1u + 4u * pair; lk.empty0 = 3u + 4u * pair; lk.slot = 0u;   // barrier 0 is __syncthreads'
  if (warp & 1u) applier(sm, st, results, lk, lane);
  else walker(sm, st, lk, lane);
}

// Adler-32 of every inflated stream against its trailer: one CTA per stream.  For bytes d[0..N):
// s1 = 1 + sum d[i], s2 = N + sum (N - i) d[i]  (mod 65521).
constexpr uint32_t ADLER_MOD = 65521u;
__global__ void __launch_bounds__(256) inflate_adler_kernel(const InfStream *streams, InfResult *results) {
  const uint32_t sid = blockIdx.x, slot = streams[sid].slot;
  const InfResult r = results[slot];
  if (r.status != 0u || (r.flags & 2u) == 0u || (r.flags & 1u) != 0u) return;
  const uint8_t *d = streams[sid].dst;
  const uint32_t N = r.out_len;
  unsigned long long s1 = 0, s2 = 0;
  const uint32_t n16 = N / 16u;
  const uint4 *q = reinterpret_cast<const uint4 *>(d);   // tile buffers are 16-byte aligned
  for (uint32_t k = threadIdx.x; k < n16; k += blockDim.x) {
    const uint4 v = q[k];
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t a = 0, b = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const uint32_t b0 = w[j] & 255u, b1 = (w[j] >> 8) & 255u, b2 = (w[j] >> 16) & 255u, b3 = w[j] >> 24;
      a += b0 + b1 + b2 + b3;
      b += (16u - 4u * j) * b0 + (15u - 4u * j) * b1 + (14u - 4u * j) * b2 + (13u - 4u * j) * b3;
    }
    const uint32_t m = (N - 16u * k - 16u) % ADLER_MOD;
    s1 += a;
    s2 += (unsigned long long)m * a + b;
    if ((k / blockDim.x & 1023u) == 1023u) s2 %= ADLER_MOD;
  }
  if (threadIdx.x == 0) {
    for (uint32_t i = n16 * 16u; i < N; i++) { s1 += d[i]; s2 += (unsigned long long)(N - i) * d[i]; }
  }
  s1 %= ADLER_MOD; s2 %= ADLER_MOD;
  __shared__ unsigned long long sh1[256], sh2[256];
  sh1[threadIdx.x] = s1; sh2[threadIdx.x] = s2;
  __syncthreads();
  for (uint32_t s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) { sh1[threadIdx.x] += sh1[threadIdx.x + s]; sh2[threadIdx.x] += sh2[threadIdx.x + s]; }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const uint32_t a = (uint32_t)((1ull + sh1[0]) % ADLER_MOD), b = (uint32_t)((N % ADLER_MOD + sh2[0]) % ADLER_MOD);
    if (((b << 16) | a) != r.adler_stored) results[slot].status = INF_E_DATA_CHECK;
  }
}

}  // namespace

cudaError_t launch_inflate(const InfStream *d_streams, InfResult *d_results, uint32_t n_streams, cudaStream_t st) {
  if (!n_streams) return cudaSuccess;
  static PerDevice once;   // shared-memory opt-in, per device
  const size_t smem = sizeof(WarpSmem) * INF_STREAMS;
  int unused = 0;
  const cudaError_t pe = once.get(&unused, [smem](int *) {
    return cudaFuncSetAttribute(inflate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  });
  if (pe != cudaSuccess) return pe;
  inflate_kernel<<<(n_streams + INF_STREAMS - 1) / INF_STREAMS, INF_STREAMS * 64, smem, st>>>(d_streams, d_results, n_streams);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  inflate_adler_kernel<<<n_streams, 256, 0, st>>>(d_streams, d_results);
  return cudaGetLastError();
}

const char *inflate_message(uint32_t status) {
  switch (status) {   // zlib's own texts (inflate.c, inftrees.c), which pako passes through
    case INF_E_HEADER: return "incorrect header check";
    case INF_E_METHOD: return "unknown compression method";
    case INF_E_WINDOW: return "invalid window size";
    case INF_E_DICT: return "need dictionary";
    case INF_E_BLOCK_TYPE: return "invalid block type";
    case INF_E_STORED: return "invalid stored block lengths";
    case INF_E_TOO_MANY: return "too many length or distance symbols";
    case INF_E_CODE_LENGTHS: return "invalid code lengths set";
    case INF_E_BIT_REPEAT: return "invalid bit length repeat";
    case INF_E_NO_EOB: return "invalid code -- missing end-of-block";
    case INF_E_LITLEN_SET: return "invalid literal/lengths set";
    case INF_E_DIST_SET: return "invalid distances set";
    case INF_E_LITLEN_CODE: return "invalid literal/length code";
    case INF_E_DIST_CODE: return "invalid distance code";
    case INF_E_FAR: return "invalid distance too far back";
    case INF_E_DATA_CHECK: return "incorrect data check";
    case INF_E_EOF: return "unexpected end of file";
    default: return "unknown inflate error";
  }
}

}  // namespace stitch
