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
//               last row of group g-1 (written to HBM by another warp, guarded
//               by a per-group progress word)
//   staging   = the rows of the inflated stream are unaligned (stride 1+rowBytes),
//               so every 32-word column block of the 32 rows is fetched with
//               16-byte cp.async granules (coalesced, 144 B per row incl. slack)
//               into a 3-slot ring in shared memory; results overwrite the slot
//               in place and leave as full 128-bit coalesced stores.
//   schedule  = units are handed out group-major by an atomic ticket to a grid
//               that is fully resident, so a unit's predecessor is always
//               already running (no deadlock) and all chains advance together.
//
// HBM traffic is the algorithmic minimum (+1/32 for the group seam rows): every
// stream byte is read once, every pixel byte written once.
#include "common.cuh"
#include "kernels.h"

namespace stitch {

static constexpr int UW_WARPS = 4;                          // independent workers per CTA
static constexpr int UW_ROW_BYTES = 144;                    // 32 words + 16 B alignment slack
static constexpr int UW_SLOT_BYTES = 33 * UW_ROW_BYTES;     // row 0 = row above the group, rows 1..32 = lanes
static constexpr int UW_WARP_SMEM = 3 * UW_SLOT_BYTES;      // 14256 B
static constexpr size_t UW_CTA_SMEM = (size_t)UW_WARPS * UW_WARP_SMEM;

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
  const uint32_t sa = (uint32_t)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(uint32_t *p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
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

// One 32-bit word of unfilterScanline.  o1/o2: my previous two result words, u1/u2: the previous two words
// of the row above, up: the word above.
template <int BPP>
__device__ __forceinline__ uint32_t unf_word(uint32_t ft, uint32_t f, uint32_t up, uint32_t o1, uint32_t o2, uint32_t u1,
                                             uint32_t u2) {
  if constexpr (BPP == 4 || BPP == 8) {
    const uint32_t a = BPP == 4 ? o1 : o2, c = BPP == 4 ? u1 : u2;
    uint32_t pred;
    switch (ft) {
      case 1: pred = a; break;
      case 2: pred = up; break;
      case 3: pred = avg4(a, up); break;
      case 4: pred = paeth4_h(unpack_h4(a), unpack_h4(up), unpack_h4(c)); break;
      default: pred = 0u; break;
    }
    return add4(f, pred);
  } else {
    uint32_t cur = 0u;
    unf_byte<BPP, 0>(ft, f, up, o1, o2, u1, u2, cur);
    unf_byte<BPP, 1>(ft, f, up, o1, o2, u1, u2, cur);
    unf_byte<BPP, 2>(ft, f, up, o1, o2, u1, u2, cur);
    unf_byte<BPP, 3>(ft, f, up, o1, o2, u1, u2, cur);
    return cur;
  }
}

struct UwParams {
  const UnfStream *streams;
  const UnfUnit *units;
  uint32_t n_units;
  uint32_t *progress;  // one word per (chain, group): column blocks of the group already in HBM
  uint32_t *ticket;
  DevStatus *status;
};

template <int BPP>
__device__ void run_unit(const UnfStream &s, const uint32_t g, uint8_t *ws, uint32_t *progress, DevStatus *status) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t r = g * 32u + lane;
  const bool valid = r < s.rows;
  const uint32_t rows_g = min(32u, s.rows - g * 32u);
  const uint32_t nwords = (s.row_bytes + 3u) >> 2;
  const uint32_t nblocks = (nwords + 31u) >> 5;
  const uint32_t n_groups = (s.rows + 31u) >> 5;
  const bool publish = g + 1u < n_groups;
  uint32_t *my_prog = progress + s.prog_base + g;
  const uint32_t *prev_prog = my_prog - 1;
  const bool has_above = g > 0u;
  uint32_t known = 0u;  // blocks of the group above known to be in HBM

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
  const uint8_t *above_row = has_above ? s.dst + (size_t)(g * 32u - 1u) * s.dst_pitch : nullptr;

  auto wait_above = [&](uint32_t need) {
    if (has_above && known < need) {
      uint32_t v = 0u;
      if (lane == 0u) {
        while ((v = ld_acquire_u32(prev_prog)) < need) __nanosleep(100);
      }
      known = __shfl_sync(0xffffffffu, v, 0);
    }
  };
  auto issue_block = [&](uint32_t b) {
    uint8_t *slot = ws + (b % 3u) * UW_SLOT_BYTES;
#pragma unroll
    for (int it = 0; it < 9; it++) {
      const uint32_t q = (uint32_t)it * 32u + lane;  // 288 granules: 32 rows x 9
      const uint32_t row = q / 9u, col = q - row * 9u;
      const uint32_t rr = g * 32u + row;
      if (rr < s.rows) {
        const uint8_t *rp = s.src + (size_t)rr * s.src_stride + 1;
        const uint8_t *gp = rp - (reinterpret_cast<uintptr_t>(rp) & 15u) + (size_t)b * 128u + col * 16u;
        if (gp < s.src_end) cp_async16(slot + (row + 1u) * UW_ROW_BYTES + col * 16u, gp);
      }
    }
    if (has_above && lane < 8u) {
      const uint32_t off = b * 128u + lane * 16u;
      if (off < s.dst_pitch) cp_async16(slot + lane * 16u, above_row + off);
    }
    cp_async_commit();
  };
  auto store_block = [&](uint32_t b) {
    const uint8_t *slot = ws + (b % 3u) * UW_SLOT_BYTES;
#pragma unroll
    for (int it = 0; it < 8; it++) {
      const uint32_t row = (uint32_t)it * 4u + (lane >> 3), col = lane & 7u;
      const uint32_t rr = g * 32u + row, off = b * 128u + col * 16u;
      if (rr < s.rows && off < s.dst_pitch) {
        const uint4 v = *reinterpret_cast<const uint4 *>(slot + (row + 1u) * UW_ROW_BYTES + col * 16u);
        *reinterpret_cast<uint4 *>(s.dst + (size_t)rr * s.dst_pitch + off) = v;
      }
    }
  };

  const uint32_t total_steps = nwords + rows_g - 1u;
  const uint32_t nphases = (total_steps + 31u) >> 5;
  uint32_t stored = 0u;
  uint32_t o1 = 0u, o2 = 0u, u1 = 0u, u2 = 0u, last = 0u;

  wait_above(1u);
  issue_block(0u);
  for (uint32_t p = 0; p < nphases; p++) {
    cp_async_wait_all();
    __syncwarp();
    if (p >= 2u && stored < nblocks) {  // block p-2 is complete: every lane has left it
      store_block(stored);
      stored++;
      __syncwarp();
      if (publish && lane == 0u) st_release_u32(my_prog, stored);
    }
    if (p + 1u < nblocks) {
      wait_above(p + 2u);
      issue_block(p + 1u);
    }
    const uint32_t cur_slot = (p % 3u) * UW_SLOT_BYTES, prev_slot = ((p + 2u) % 3u) * UW_SLOT_BYTES;
#pragma unroll 4
    for (uint32_t i = 0; i < 32u; i++) {
      uint32_t up = __shfl_up_sync(0xffffffffu, last, 1);
      const uint32_t j = p * 32u + i - lane;  // wraps to a huge value while the lane has not started
      if (valid && j < nwords) {
        const uint32_t wi = j & 31u;
        const uint32_t slot_off = lane <= i ? cur_slot : prev_slot;
        const uint8_t *ip = ws + slot_off + in_off + wi * 4u;
        const uint32_t f = __byte_perm(*reinterpret_cast<const uint32_t *>(ip),
                                       *reinterpret_cast<const uint32_t *>(ip + 4), sel);
        if (lane == 0u) up = has_above ? *reinterpret_cast<const uint32_t *>(ws + slot_off + wi * 4u) : 0u;
        const uint32_t o = unf_word<BPP>(ft, f, up, o1, o2, u1, u2);
        *reinterpret_cast<uint32_t *>(ws + slot_off + my_row_off + wi * 4u) = o;
        o2 = o1; o1 = o; u2 = u1; u1 = up; last = o;
      }
    }
  }
  __syncwarp();
  while (stored < nblocks) {
    store_block(stored);
    stored++;
  }
  __syncwarp();
  if (publish && lane == 0u) st_release_u32(my_prog, nblocks);
  __syncwarp();  // the ring is reused by the next unit
}

__global__ void __launch_bounds__(UW_WARPS * 32) unfilter_wave_kernel(const UwParams p) {
  extern __shared__ __align__(16) uint8_t uw_smem[];
  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
  uint8_t *ws = uw_smem + warp * UW_WARP_SMEM;
  for (;;) {
    uint32_t u = 0u;
    if (lane == 0u) u = atomicAdd(p.ticket, 1u);
    u = __shfl_sync(0xffffffffu, u, 0);
    if (u >= p.n_units) break;
    const UnfUnit un = p.units[u];
    const UnfStream s = p.streams[un.stream];
    switch (s.bpp) {
      case 1: run_unit<1>(s, un.group, ws, p.progress, p.status); break;
      case 2: run_unit<2>(s, un.group, ws, p.progress, p.status); break;
      case 3: run_unit<3>(s, un.group, ws, p.progress, p.status); break;
      case 4: run_unit<4>(s, un.group, ws, p.progress, p.status); break;
      case 6: run_unit<6>(s, un.group, ws, p.progress, p.status); break;
      case 8: run_unit<8>(s, un.group, ws, p.progress, p.status); break;
      default: break;
    }
  }
}

cudaError_t launch_unfilter_wave(const UnfStream *d_streams, const UnfUnit *d_units, uint32_t n_units,
                                 uint32_t *d_progress, uint32_t *d_ticket, DevStatus *d_status, int sm_count,
                                 cudaStream_t st) {
  if (n_units == 0) return cudaSuccess;
  static int occ = -1;
  if (occ < 0) {
    cudaFuncSetAttribute(unfilter_wave_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)UW_CTA_SMEM);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, unfilter_wave_kernel, UW_WARPS * 32, UW_CTA_SMEM);
  }
  if (occ < 1) return cudaErrorLaunchOutOfResources;
  UwParams p{d_streams, d_units, n_units, d_progress, d_ticket, d_status};
  const uint32_t resident = (uint32_t)(occ * sm_count);  // every CTA of the grid is resident: see the header
  const uint32_t want = (n_units + UW_WARPS - 1) / UW_WARPS;
  unfilter_wave_kernel<<<want < resident ? want : resident, UW_WARPS * 32, UW_CTA_SMEM, st>>>(p);
  return cudaGetLastError();
}

}  // namespace stitch
