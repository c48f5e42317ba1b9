// unfilter_wave.cu -- kernel (a), fast path: PNG unfilter (src/png-filter.ts:34-83)
// as a warp-level diagonal wavefront with shared-memory staging.
//
// Dependencies: byte i of row y needs byte i-bpp of row y (Sub/Average/Paeth)
// and bytes i, i-bpp of row y-1 (Up/Average/Paeth).  One image (or one Adam7
// pass) is therefore a single chain in y, but rows can overlap in time when
// row y stays one word behind row y-1.
//
//   unit      = one 32-row group of one chain; one WARP runs one unit at a time
//   lane l    = row 32g+l; at step T it produces 32-bit word T-l of its row
//   up word   = lane l-1's previous result, one SHFL per step; lane 0 reads the
//               last row of group g-1 (the "seam" row, produced by another warp)
//   staging   = the rows of the inflated stream are unaligned (stride 1+rowBytes),
//               so every 32-word column block of the 32 rows is fetched with
//               16-byte cp.async granules (coalesced, 144 B per row incl. slack)
//               into a 3-slot ring in shared memory; results overwrite the slot
//               in place and leave as full 128-bit coalesced stores.
//   seam      = a finished 32-word block of a group's last row is also written to a
//               mailbox as 32 x (data, tag) pairs, one aligned 8-byte store per
//               lane.  The consumer polls the pairs until every tag equals the
//               epoch of this decode: data and tag travel in one single-copy
//               atomic access, so no fence, no flag word and no acquire are
//               needed (the LL protocol of collective libraries).
//   schedule  = units are handed out group-major by an atomic ticket, so a unit
//               only ever waits for a unit with an earlier ticket, which is
//               already running (a waiting warp holds no unstarted ticket).
//   step      = filters None/Sub/Up/Paeth are all "Paeth with some inputs forced
//               to zero" (the PRMT that unpacks bytes to half2 does the forcing),
//               Average is computed next to it and selected per lane: one
//               branch-free instruction stream for all 32 rows whatever their
//               filter types, fully unrolled over the 32 steps of a block.
//               Filter distances of 1, 2, 3 and 6 bytes use the same predictor in
//               rounds over the bytes of a word (byte_word below).
//
// HBM traffic is the algorithmic minimum plus the seam rows (2/32 of the pixel
// bytes written, 2/32 read): every stream byte is read once, every pixel byte
// written once.
#include "common.cuh"
#include "kernels.h"
#include <type_traits>

namespace stitch {

static constexpr int UW_WARPS = 4;                          // independent workers per CTA
static constexpr int UW_ROW_BYTES = 144;                    // 32 words + 16 B alignment slack
static constexpr int UW_SLOT_BYTES = 33 * UW_ROW_BYTES;     // row 0 = seam row above the group, rows 1..32 = lanes
static constexpr int UW_WARP_SMEM = 3 * UW_SLOT_BYTES;      // 14256 B
static constexpr size_t UW_CTA_SMEM = (size_t)UW_WARPS * UW_WARP_SMEM;
static constexpr uint32_t FULL = 0xffffffffu;

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
// One (data, tag) pair of the seam mailbox: a single 8-byte access at gpu scope.
__device__ __forceinline__ uint2 ld_pair(const uint2 *p) {
  uint2 v;
  asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_pair(uint2 *p, uint32_t data, uint32_t tag) {
  asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(data), "r"(tag) : "memory");
}

// byte `idx` of the 12-byte window [w2 | w1 | w0]
template <int IDX>
__device__ __forceinline__ uint32_t win_byte(uint32_t w2, uint32_t w1, uint32_t w0) {
  static_assert(IDX >= 0 && IDX < 12, "window index");
  const uint32_t w = IDX < 4 ? w2 : IDX < 8 ? w1 : w0;
  return (w >> (8 * (IDX & 3))) & 0xffu;
}

template <int BPP, int K>
__device__ __forceinline__ void unf_byte(uint32_t ft, uint32_t f, uint32_t up, uint32_t o1, uint32_t o2, uint32_t u1,
                                         uint32_t u2, uint32_t &cur) {
  constexpr int IDX = 8 + K - BPP;
  const uint32_t a = win_byte<IDX>(o2, o1, cur), c = win_byte<IDX>(u2, u1, up);
  const uint32_t b = (up >> (8 * K)) & 0xffu, x = (f >> (8 * K)) & 0xffu;
  uint32_t pr;
  switch (ft) {
    case 1: pr = a; break;
    case 2: pr = b; break;
    case 3: pr = (a + b) >> 1; break;
    case 4: pr = (uint32_t)paeth1((int)a, (int)b, (int)c); break;
    default: pr = 0u; break;
  }
  cur |= ((x + pr) & 0xffu) << (8 * K);
}

// One 32-bit word of unfilterScanline for filter distances that are not a multiple of 4 (byte-serial).
// o1/o2: my previous two result words, u1/u2: the previous two words of the row above, up: the word above.
template <int BPP>
__device__ __forceinline__ uint32_t unf_word(uint32_t ft, uint32_t f, uint32_t up, uint32_t o1, uint32_t o2, uint32_t u1,
                                             uint32_t u2) {
  uint32_t cur = 0u;
  unf_byte<BPP, 0>(ft, f, up, o1, o2, u1, u2, cur);
  unf_byte<BPP, 1>(ft, f, up, o1, o2, u1, u2, cur);
  unf_byte<BPP, 2>(ft, f, up, o1, o2, u1, u2, cur);
  unf_byte<BPP, 3>(ft, f, up, o1, o2, u1, u2, cur);
  return cur;
}

// ---------------------------------------------------------------------------
// Branch-free step for filter distances 4 and 8 (whole words).
// ---------------------------------------------------------------------------
// Bytes become halves 1024+v through PRMT with the constant word 0x00000064:
// selector nibble 4 picks 0x64, nibble 5 picks 0x00, so 0x4140 / 0x4342 unpack
// bytes (0,1) / (2,3) and 0x4545 yields 1024+0 twice, i.e. "this input is zero".
__device__ __forceinline__ __half2 unpack2(uint32_t w, uint32_t sel) {
  const uint32_t r = __byte_perm(w, 0x64u, sel);
  return *reinterpret_cast<const __half2 *>(&r);
}
struct LaneFilter {
  uint32_t a_lo, a_hi, b_lo, b_hi, c_lo, c_hi;  // PRMT selectors for the left, up and up-left inputs
  bool avg;
};
__device__ __forceinline__ LaneFilter lane_filter(uint32_t ft) {
  LaneFilter k;
  const bool ka = ft == 1u || ft == 3u || ft == 4u;  // Sub, Average, Paeth read the left pixel
  const bool kb = ft >= 2u;                          // Up, Average, Paeth read the pixel above
  const bool kc = ft == 4u;                          // Paeth reads the pixel above-left
  k.a_lo = ka ? 0x4140u : 0x4545u; k.a_hi = ka ? 0x4342u : 0x4545u;
  k.b_lo = kb ? 0x4140u : 0x4545u; k.b_hi = kb ? 0x4342u : 0x4545u;
  k.c_lo = kc ? 0x4140u : 0x4545u; k.c_hi = kc ? 0x4342u : 0x4545u;
  k.avg = ft == 3u;
  return k;
}
// 1024 + floor((a + b) / 2): a + (b - 1024) is exact, and t/2 + 511.75 = 1023.75 + (a+b)/2 rounds (one
// rounding, ulp 1, ties to even only at a+b = 0) to 1024 + floor((a+b)/2).
__device__ __forceinline__ __half2 avg_h2(__half2 a, __half2 b) {
  const __half2 t = __hadd2(a, __hsub2(b, __float2half2_rn(1024.f)));
  return __hfma2(t, __float2half2_rn(0.5f), __float2half2_rn(511.75f));
}
template <int D>
struct WaveState {       // D = filter distance in words
  __half2 a_lo[D], a_hi[D];  // my last D results (left pixel), masked by the filter type
  __half2 c_lo[D], c_hi[D];  // the last D words of the row above (above-left pixel), masked
  uint32_t last;             // my last result, packed (what the lane below reads)
};
// f: filtered word; up: word above.  Returns the reconstructed word and advances the state.
// PAETH = false: no row of the warp's group uses the Paeth filter (a warp-uniform fact, checked once per unit).
// None, Sub, Up and Average then share one form: t = 1024 + a' + b' with the unused input forced to zero is the
// predictor itself for None/Sub/Up and halves to the Average predictor -- a third fewer instructions per step.
template <int D, bool PAETH>
__device__ __forceinline__ uint32_t wave_word(const LaneFilter &k, WaveState<D> &s, uint32_t f, uint32_t up) {
  const __half2 m1024 = __float2half2_rn(-1024.f);
  const __half2 b_lo = unpack2(up, k.b_lo), b_hi = unpack2(up, k.b_hi);
  const __half2 f_lo = __hadd2(unpack2(f, 0x4140u), m1024), f_hi = __hadd2(unpack2(f, 0x4342u), m1024);
  const __half2 t_lo = __hadd2(s.a_lo[0], __hadd2(b_lo, m1024)), t_hi = __hadd2(s.a_hi[0], __hadd2(b_hi, m1024));
  const __half2 v_lo = __hfma2(t_lo, __float2half2_rn(0.5f), __float2half2_rn(511.75f));  // see avg_h2
  const __half2 v_hi = __hfma2(t_hi, __float2half2_rn(0.5f), __float2half2_rn(511.75f));
  __half2 p_lo, p_hi;
  if (PAETH) {
    p_lo = paeth_h2(s.a_lo[0], b_lo, s.c_lo[0]);
    p_hi = paeth_h2(s.a_hi[0], b_hi, s.c_hi[0]);
  } else {
    p_lo = t_lo; p_hi = t_hi;
  }
  const __half2 r_lo = __hadd2(k.avg ? v_lo : p_lo, f_lo), r_hi = __hadd2(k.avg ? v_hi : p_hi, f_hi);
  const uint32_t o = pack_h4(r_lo, r_hi);  // low byte of each 1024 + (pred + x): the sum mod 256
#pragma unroll
  for (int d = 0; d + 1 < D; d++) {
    s.a_lo[d] = s.a_lo[d + 1]; s.a_hi[d] = s.a_hi[d + 1];
    if (PAETH) { s.c_lo[d] = s.c_lo[d + 1]; s.c_hi[d] = s.c_hi[d + 1]; }
  }
  s.a_lo[D - 1] = unpack2(o, k.a_lo); s.a_hi[D - 1] = unpack2(o, k.a_hi);
  if (PAETH) { s.c_lo[D - 1] = unpack2(up, k.c_lo); s.c_hi[D - 1] = unpack2(up, k.c_hi); }
  s.last = o;
  return o;
}

// ---------------------------------------------------------------------------
// Branch-free step for filter distances that are not a whole number of words.
// ---------------------------------------------------------------------------
// State is kept packed (my previous two results, the previous two words above); the lane's filter type is
// applied as three all-or-nothing byte masks on the packed operands.  A word is produced in rounds: a round
// holds the bytes whose left neighbour (distance BPP) is already known -- 1 round for BPP 6 (all four bytes
// look into earlier words), 2 for BPP 2 and 3, 4 for BPP 1 -- and each round is the same half2 predictor as
// above on one or two bytes.
struct ByteState {
  uint32_t o1, o2, u1, u2, last;
};
struct ByteFilter {
  uint32_t mA, mB, mC;  // 0 or ~0: does this row's filter read the left / up / up-left byte
  bool avg;
};
__device__ __forceinline__ ByteFilter byte_filter(uint32_t ft) {
  ByteFilter k;
  k.mA = (ft == 1u || ft == 3u || ft == 4u) ? ~0u : 0u;
  k.mB = ft >= 2u ? ~0u : 0u;
  k.mC = ft == 4u ? ~0u : 0u;
  k.avg = ft == 3u;
  return k;
}
// 1024 + (pred + x) for two bytes; a, b, c as 1024+v (masked operands are 1024), x as plain v.
template <bool PAETH>
__device__ __forceinline__ __half2 pred_add(__half2 a, __half2 b, __half2 c, __half2 x, bool avg) {
  const __half2 t = __hadd2(a, __hadd2(b, __float2half2_rn(-1024.f)));
  const __half2 v = __hfma2(t, __float2half2_rn(0.5f), __float2half2_rn(511.75f));
  const __half2 p = PAETH ? paeth_h2(a, b, c) : t;
  return __hadd2(avg ? v : p, x);
}
__device__ __forceinline__ __half2 plain2(uint32_t w, uint32_t sel) {  // bytes -> plain values (not 1024-biased)
  return __hadd2(unpack2(w, sel), __float2half2_rn(-1024.f));
}
template <int BPP, bool PAETH>
__device__ __forceinline__ uint32_t byte_word(const ByteFilter &k, ByteState &s, uint32_t f, uint32_t up) {
  const uint32_t B = up & k.mB, Cu = up & k.mC, C1 = s.u1 & k.mC, C2 = s.u2 & k.mC, A1 = s.o1 & k.mA, A2 = s.o2 & k.mA;
  const __half2 zero = __float2half2_rn(1024.f);
  uint32_t o;
  if (BPP == 6) {
    const uint32_t A = __byte_perm(A2, A1, 0x5432), C = __byte_perm(C2, C1, 0x5432);
    const __half2 r0 = pred_add<PAETH>(unpack2(A, 0x4140u), unpack2(B, 0x4140u), unpack2(C, 0x4140u), plain2(f, 0x4140u), k.avg);
    const __half2 r1 = pred_add<PAETH>(unpack2(A, 0x4342u), unpack2(B, 0x4342u), unpack2(C, 0x4342u), plain2(f, 0x4342u), k.avg);
    o = pack_h4(r0, r1);
  } else if (BPP == 3) {
    const __half2 r0 = pred_add<PAETH>(unpack2(A1, 0x4241u), unpack2(B, 0x4140u), unpack2(C1, 0x4241u), plain2(f, 0x4140u), k.avg);
    const __half2 r1 = pred_add<PAETH>(unpack2(A1, 0x4543u), unpack2(B, 0x4542u), unpack2(C1, 0x4543u), plain2(f, 0x4542u), k.avg);
    // byte 3: left = byte 0 of this word (low half of r0, brought back to 1024 + v), up-left = byte 0 above
    const uint32_t r0w = *reinterpret_cast<const uint32_t *>(&r0);
    const __half2 a3 = k.mA ? unpack2(r0w, 0x4540u) : zero;
    const __half2 r2 = pred_add<PAETH>(a3, unpack2(B, 0x4543u), unpack2(Cu, 0x4540u), plain2(f, 0x4543u), k.avg);
    const uint32_t t = __byte_perm(r0w, *reinterpret_cast<const uint32_t *>(&r1), 0x4420);
    o = __byte_perm(t, *reinterpret_cast<const uint32_t *>(&r2), 0x4210);
  } else if (BPP == 2) {
    const __half2 r0 = pred_add<PAETH>(unpack2(A1, 0x4342u), unpack2(B, 0x4140u), unpack2(C1, 0x4342u), plain2(f, 0x4140u), k.avg);
    const uint32_t r0w = *reinterpret_cast<const uint32_t *>(&r0);
    const uint32_t p0 = __byte_perm(r0w, 0u, 0x4420) & 0xffffu;  // bytes 0, 1 of the result
    const __half2 a1 = k.mA ? unpack2(p0, 0x4140u) : zero;
    const __half2 r1 = pred_add<PAETH>(a1, unpack2(B, 0x4342u), unpack2(Cu, 0x4140u), plain2(f, 0x4342u), k.avg);
    o = pack_h4(r0, r1);
  } else {  // BPP == 1
    __half2 r[4];
    r[0] = pred_add<PAETH>(unpack2(A1, 0x4543u), unpack2(B, 0x4540u), unpack2(C1, 0x4543u), plain2(f, 0x4540u), k.avg);
#pragma unroll
    for (int j = 1; j < 4; j++) {
      const uint32_t prev = *reinterpret_cast<const uint32_t *>(&r[j - 1]);
      const __half2 a = k.mA ? unpack2(prev, 0x4540u) : zero;
      r[j] = pred_add<PAETH>(a, unpack2(B, 0x4540u + (uint32_t)j), unpack2(Cu, 0x4540u + (uint32_t)(j - 1)),
                             plain2(f, 0x4540u + (uint32_t)j), k.avg);
    }
    const uint32_t t0 = __byte_perm(*reinterpret_cast<const uint32_t *>(&r[0]), *reinterpret_cast<const uint32_t *>(&r[1]), 0x4440);
    const uint32_t t1 = __byte_perm(*reinterpret_cast<const uint32_t *>(&r[2]), *reinterpret_cast<const uint32_t *>(&r[3]), 0x4440);
    o = __byte_perm(t0, t1, 0x5410);
  }
  s.o2 = s.o1; s.o1 = o; s.u2 = s.u1; s.u1 = up; s.last = o;
  return o;
}

struct UwParams {
  const UnfStream *streams;
  const UnfUnit *units;
  uint32_t n_units;
  uint2 *seam;     // mailbox: 32 pairs per (chain, group, column block)
  uint32_t epoch;  // tag of this decode (the mailbox is zeroed once, epochs start at 1)
  uint32_t *ticket;
  DevStatus *status;
};

template <int BPP>
__device__ void run_unit(const UnfStream &s, const uint32_t g, uint8_t *ws, uint2 *seam, const uint32_t epoch,
                         DevStatus *status) {
  constexpr bool WORDS = BPP == 4 || BPP == 8;  // filter distance is a whole number of words
  constexpr int D = WORDS ? BPP / 4 : 1;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t r = g * 32u + lane;
  const bool valid = r < s.rows;
  const uint32_t rows_g = min(32u, s.rows - g * 32u);
  const uint32_t nwords = (s.row_bytes + 3u) >> 2;
  const uint32_t nblocks = (nwords + 31u) >> 5;
  const uint32_t n_groups = (s.rows + 31u) >> 5;
  const bool publish = g + 1u < n_groups;  // implies rows_g == 32
  const bool has_above = g > 0u;
  uint2 *my_seam = seam + ((size_t)s.prog_base + (size_t)g * nblocks) * 32u + lane;  // block b at +32*b
  const uint2 *ab_seam = my_seam - (size_t)nblocks * 32u;                            // same, group above

  // my row inside the stream: filter byte, then row_bytes filtered bytes at an arbitrary alignment
  const uint8_t *src_row = s.src + (size_t)(valid ? r : 0u) * s.src_stride + 1;
  uint32_t ft = valid ? src_row[-1] : 0u;
  if (ft > 4u) {  // png-filter.ts:78-79; smallest (tile,row) wins, low byte carries the offending value
    const unsigned long long e =
        ((unsigned long long)(s.tile & 0x7fffffu) << 40) | ((unsigned long long)r << 8) | (ft & 0xffu);
    atomicMin(&status->first_error, e);
    ft = 0u;
  }
  const uint32_t sh = (uint32_t)(reinterpret_cast<uintptr_t>(src_row) & 15u);
  const uint32_t sel = 0x3210u + 0x1111u * (sh & 3u);  // PRMT selector: 4 bytes starting at byte (sh & 3)
  const uint32_t my_row_off = (lane + 1u) * UW_ROW_BYTES;
  const uint32_t in_off = my_row_off + (sh & ~3u);

  // Granule q = 32*it + lane of a block (q = 9*row + col) lands at slot + 144 + 16*q; its source offset from
  // s.src for block 0 is computed once per unit (rows past the end get an offset that never passes the bound).
  // Offsets are relative to the unit's first row so that they fit 32 bits for streams of any length.
  const uint8_t *const unit_src = s.src + (size_t)(g * 32u) * s.src_stride;
  const ptrdiff_t end_far = s.src_end - unit_src;
  const int32_t end_off = (int32_t)(end_far < 0x3fffffff ? end_far : 0x3fffffff);
  int32_t goff[9];
#pragma unroll
  for (int it = 0; it < 9; it++) {
    const uint32_t q = (uint32_t)it * 32u + lane;
    const uint32_t row = q / 9u, col = q - row * 9u;
    const uint32_t ro = row * s.src_stride + 1u;
    const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(unit_src) + ro) & 15u;
    goff[it] = row < rows_g ? (int32_t)(ro - mis + col * 16u) : 0x40000000;
  }
  auto issue_block = [&](uint32_t b) {
    uint8_t *slot = ws + (b % 3u) * UW_SLOT_BYTES + UW_ROW_BYTES + 16u * lane;
#pragma unroll
    for (int it = 0; it < 9; it++) {
      const int32_t o = goff[it] + (int32_t)(b * 128u);
      if (o < end_off) cp_async16(slot + it * 512, unit_src + o);
    }
    cp_async_commit();
  };
  // Block b is complete (every lane has left it): pixels to HBM, last row also to the seam mailbox.
  uint8_t *const st_dst = s.dst + (size_t)(g * 32u + (lane >> 3)) * s.dst_pitch + (lane & 7u) * 16u;
  const uint32_t st_src = ((lane >> 3) + 1u) * UW_ROW_BYTES + (lane & 7u) * 16u;
  const size_t pitch4 = (size_t)s.dst_pitch * 4u;
  auto store_block = [&](uint32_t b) {
    const uint8_t *slot = ws + (b % 3u) * UW_SLOT_BYTES;
    if (publish) st_pair(my_seam + b * 32u, *reinterpret_cast<const uint32_t *>(slot + 32u * UW_ROW_BYTES + lane * 4u), epoch);
    if (b * 128u + (lane & 7u) * 16u < s.dst_pitch) {
      uint8_t *d = st_dst + b * 128u;
      if (rows_g == 32u) {
#pragma unroll
        for (int it = 0; it < 8; it++)
          *reinterpret_cast<uint4 *>(d + it * pitch4) = *reinterpret_cast<const uint4 *>(slot + st_src + it * 4 * UW_ROW_BYTES);
      } else {
#pragma unroll
        for (int it = 0; it < 8; it++)
          if ((uint32_t)it * 4u + (lane >> 3) < rows_g)
            *reinterpret_cast<uint4 *>(d + it * pitch4) = *reinterpret_cast<const uint4 *>(slot + st_src + it * 4 * UW_ROW_BYTES);
      }
    }
  };

  const uint32_t total_steps = nwords + rows_g - 1u;
  const uint32_t nphases = (total_steps + 31u) >> 5;
  uint32_t stored = 0u;
  // state of the byte-distance path
  ByteState bs;
  bs.o1 = bs.o2 = bs.u1 = bs.u2 = bs.last = 0u;
  const ByteFilter bf = byte_filter(ft);
  // state of the word path
  const LaneFilter lf = lane_filter(ft);
  const bool any_paeth = __any_sync(FULL, ft == 4u);
  WaveState<D> st;
#pragma unroll
  for (int d = 0; d < D; d++) st.a_lo[d] = st.a_hi[d] = st.c_lo[d] = st.c_hi[d] = __float2half2_rn(1024.f);
  st.last = 0u;

  if (!has_above) {  // first group of the chain: the row above is all zero (png-filter.ts:41)
#pragma unroll
    for (int k = 0; k < 3; k++) *reinterpret_cast<uint32_t *>(ws + k * UW_SLOT_BYTES + lane * 4u) = 0u;
  }
  {  // slot 2 is what lanes read before their first word (see the step loop)
    uint4 *z = reinterpret_cast<uint4 *>(ws + 2 * UW_SLOT_BYTES + UW_ROW_BYTES);
#pragma unroll
    for (int k = 0; k < 9; k++) z[k * 32 + lane] = make_uint4(0u, 0u, 0u, 0u);
  }
  issue_block(0u);
  uint2 ab = make_uint2(0u, epoch);
  if (has_above) ab = ld_pair(ab_seam);
  for (uint32_t p = 0; p < nphases; p++) {
    cp_async_wait_all();
    __syncwarp();
    if (p >= 2u && stored < nblocks) {
      store_block(stored);
      stored++;
      __syncwarp();  // the slot is refilled below
    }
    if (p + 1u < nblocks) issue_block(p + 1u);
    const uint32_t cur_slot = (p % 3u) * UW_SLOT_BYTES, prev_slot = ((p + 2u) % 3u) * UW_SLOT_BYTES;
    if (has_above && p < nblocks) {  // seam block p -> row 0 of the current slot; block p+1 is requested early
      while (!__all_sync(FULL, ab.y == epoch)) {
        __nanosleep(100);
        if (ab.y != epoch) ab = ld_pair(ab_seam + p * 32u);
      }
      *reinterpret_cast<uint32_t *>(ws + cur_slot + lane * 4u) = ab.x;
      if (p + 1u < nblocks) ab = ld_pair(ab_seam + (p + 1u) * 32u);
      __syncwarp();
    }
    // Whole-word formats run every block through the branch-free steps, including the blocks where some lanes
    // have not started or have already finished.  A lane that has not started reads zeros (slot 2 was cleared,
    // the lane above still produces zeros), so it produces zeros and its state stays "left = above-left = 0"
    // until its first real word; a lane past its last word, or without a row, computes garbage that nobody
    // reads: store_block never writes rows >= rows_g or columns >= dst_pitch, blocks >= nblocks are never
    // stored, and the lane below read this lane's last real word one step before.
    if (WORDS) {
      // lane l works in the current slot from step l on, before that in the tail of the previous one
      const uint32_t in_cur = cur_slot + in_off - 4u * lane, in_prev = prev_slot + in_off - 4u * lane + 128u;
      const uint32_t out_cur = cur_slot + my_row_off - 4u * lane, out_prev = prev_slot + my_row_off - 4u * lane + 128u;
      auto block_steps = [&](auto paeth_tag) {
        constexpr bool PAETH = decltype(paeth_tag)::value;
#pragma unroll
        for (int c8 = 0; c8 < 4; c8++) {
          uint32_t fw[8], sw[8];
          *reinterpret_cast<uint4 *>(&sw[0]) = *reinterpret_cast<const uint4 *>(ws + cur_slot + c8 * 32);
          *reinterpret_cast<uint4 *>(&sw[4]) = *reinterpret_cast<const uint4 *>(ws + cur_slot + c8 * 32 + 16);
#pragma unroll
          for (int k = 0; k < 8; k++) {
            const int i = c8 * 8 + k;
            const uint8_t *ip = ws + ((int)lane <= i ? in_cur : in_prev) + 4 * i;
            fw[k] = __byte_perm(*reinterpret_cast<const uint32_t *>(ip), *reinterpret_cast<const uint32_t *>(ip + 4), sel);
          }
#pragma unroll
          for (int k = 0; k < 8; k++) {
            const int i = c8 * 8 + k;
            uint32_t up = __shfl_up_sync(FULL, st.last, 1);
            if (lane == 0u) up = sw[k];
            const uint32_t o = wave_word<D, PAETH>(lf, st, fw[k], up);
            *reinterpret_cast<uint32_t *>(ws + ((int)lane <= i ? out_cur : out_prev) + 4 * i) = o;
          }
        }
      };
      if (any_paeth) block_steps(std::true_type{}); else block_steps(std::false_type{});
    } else {
      // same walk with the byte-distance step; 8 steps unrolled, the loop over the four chunks is kept rolled
      const uint32_t in_cur = cur_slot + in_off - 4u * lane, in_prev = prev_slot + in_off - 4u * lane + 128u;
      const uint32_t out_cur = cur_slot + my_row_off - 4u * lane, out_prev = prev_slot + my_row_off - 4u * lane + 128u;
      auto block_steps = [&](auto paeth_tag) {
        constexpr bool PAETH = decltype(paeth_tag)::value;
#pragma unroll 1
        for (uint32_t c8 = 0; c8 < 4u; c8++) {
          uint32_t fw[8], sw[8];
          *reinterpret_cast<uint4 *>(&sw[0]) = *reinterpret_cast<const uint4 *>(ws + cur_slot + c8 * 32u);
          *reinterpret_cast<uint4 *>(&sw[4]) = *reinterpret_cast<const uint4 *>(ws + cur_slot + c8 * 32u + 16u);
#pragma unroll
          for (int k = 0; k < 8; k++) {
            const uint32_t i = c8 * 8u + (uint32_t)k;
            const uint8_t *ip = ws + (lane <= i ? in_cur : in_prev) + 4u * i;
            fw[k] = __byte_perm(*reinterpret_cast<const uint32_t *>(ip), *reinterpret_cast<const uint32_t *>(ip + 4), sel);
          }
#pragma unroll
          for (int k = 0; k < 8; k++) {
            const uint32_t i = c8 * 8u + (uint32_t)k;
            uint32_t up = __shfl_up_sync(FULL, bs.last, 1);
            if (lane == 0u) up = sw[k];
            const uint32_t o = byte_word<WORDS ? 1 : BPP, PAETH>(bf, bs, fw[k], up);
            *reinterpret_cast<uint32_t *>(ws + (lane <= i ? out_cur : out_prev) + 4u * i) = o;
          }
        }
      };
      if (any_paeth) block_steps(std::true_type{}); else block_steps(std::false_type{});
    }
  }
  __syncwarp();
  while (stored < nblocks) {
    store_block(stored);
    stored++;
  }
  __syncwarp();  // the ring is reused by the next unit
}

// KBPP > 0: every chain of the launch has this filter distance (the common case: one pixel format per job),
// which keeps the specialised kernels small.  KBPP == 0: mixed formats, dispatch per unit.
template <int KBPP>
__global__ void __launch_bounds__(UW_WARPS * 32, 4)
    unfilter_wave_kernel(const UwParams p) {
  extern __shared__ __align__(16) uint8_t uw_smem[];
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  uint8_t *ws = uw_smem + warp * UW_WARP_SMEM;
  for (;;) {
    uint32_t u = 0u;
    if (lane == 0u) u = atomicAdd(p.ticket, 1u);
    u = __shfl_sync(FULL, u, 0);
    if (u >= p.n_units) break;
    const UnfUnit un = p.units[u];
    const UnfStream s = p.streams[un.stream];
    if (KBPP > 0) {
      run_unit<KBPP == 0 ? 4 : KBPP>(s, un.group, ws, p.seam, p.epoch, p.status);
    } else {
      switch (s.bpp) {
        case 1: run_unit<1>(s, un.group, ws, p.seam, p.epoch, p.status); break;
        case 2: run_unit<2>(s, un.group, ws, p.seam, p.epoch, p.status); break;
        case 3: run_unit<3>(s, un.group, ws, p.seam, p.epoch, p.status); break;
        case 4: run_unit<4>(s, un.group, ws, p.seam, p.epoch, p.status); break;
        case 6: run_unit<6>(s, un.group, ws, p.seam, p.epoch, p.status); break;
        case 8: run_unit<8>(s, un.group, ws, p.seam, p.epoch, p.status); break;
        default: break;
      }
    }
  }
}

template <int KBPP>
static cudaError_t launch_uw(const UwParams &p, int sm_count, cudaStream_t st) {
  static PerDevice once;   // one per instantiation: shared-memory opt-in and occupancy, per device
  int occ = 0;
  const cudaError_t pe = once.get(&occ, [](int *v) {
    cudaError_t e = cudaFuncSetAttribute(unfilter_wave_kernel<KBPP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)UW_CTA_SMEM);
    return e != cudaSuccess ? e : cudaOccupancyMaxActiveBlocksPerMultiprocessor(v, unfilter_wave_kernel<KBPP>, UW_WARPS * 32, UW_CTA_SMEM);
  });
  if (pe != cudaSuccess) return pe;
  if (occ < 1) return cudaErrorLaunchOutOfResources;
  // Residency is not needed for correctness (see "schedule" above); the grid is sized to it for balance.
  const uint32_t resident = (uint32_t)(occ * sm_count);
  const uint32_t want = (p.n_units + UW_WARPS - 1) / UW_WARPS;
  unfilter_wave_kernel<KBPP><<<want < resident ? want : resident, UW_WARPS * 32, UW_CTA_SMEM, st>>>(p);
  return cudaGetLastError();
}

size_t unfilter_wave_seam_bytes(uint32_t rows, uint32_t row_bytes) {
  const size_t groups = (rows + 31u) / 32u, nblocks = ((row_bytes + 3u) / 4u + 31u) / 32u;
  return groups * nblocks * 32u * sizeof(uint2);
}

cudaError_t launch_unfilter_wave(const UnfStream *d_streams, const UnfUnit *d_units, uint32_t n_units, void *d_seam,
                                 uint32_t epoch, uint32_t *d_ticket, DevStatus *d_status, int sm_count,
                                 uint32_t uniform_bpp, cudaStream_t st) {
  if (n_units == 0) return cudaSuccess;
  const UwParams p{d_streams, d_units, n_units, reinterpret_cast<uint2 *>(d_seam), epoch, d_ticket, d_status};
  switch (uniform_bpp) {
    case 1: return launch_uw<1>(p, sm_count, st);
    case 2: return launch_uw<2>(p, sm_count, st);
    case 3: return launch_uw<3>(p, sm_count, st);
    case 4: return launch_uw<4>(p, sm_count, st);
    case 6: return launch_uw<6>(p, sm_count, st);
    case 8: return launch_uw<8>(p, sm_count, st);
    default: return launch_uw<0>(p, sm_count, st);
  }
}

}  // namespace stitch
