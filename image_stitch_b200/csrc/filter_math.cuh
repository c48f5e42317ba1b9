// filter_math.cuh -- device helpers shared by the composite+filter kernels (fused.cu, strip.cu):
// packed-byte cost sums, the half2 Paeth residual, tensor-memory scratch accessors and the generic
// per-quad compositor (normalise + alpha-over, src/pixel-ops.ts:496-736).
#pragma once
#include "common.cuh"

namespace stitch {

// ---- packed-byte helpers ------------------------------------------------------
__device__ __forceinline__ uint32_t absdiff_u4(uint32_t a, uint32_t b) { return __vabsdiffu4(a, b); }
// acc + sum over bytes of |sext8(v)|  ==  acc + sum of min(b, 256 - b)   (png-filter.ts:165-171)
__device__ __forceinline__ uint32_t acc_abs_s4(uint32_t v, uint32_t acc) {
  uint32_t r;
  asm("vabsdiff4.u32.s32.s32.add %0, %1, %2, %3;" : "=r"(r) : "r"(v), "r"(0u), "r"(acc));
  return r;
}

__device__ __forceinline__ void tmem_st4(uint32_t taddr, const uint4 v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};"
               ::"r"(taddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 tmem_ld4(uint32_t taddr) {
  uint4 v;
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  return v;
}

// sat(|a| - |b|) on two lanes: 1 where |a| > |b|, else 0 (the operands are integers).  A comparison that
// runs on the FMA pipe (HADD2.SAT) instead of the half-rate ALU pipe (HSET2), which is the one this kernel fills.
__device__ __forceinline__ __half2 abs_gt_h2(__half2 a, __half2 b) {
  uint32_t r;
  asm("{ .reg .b32 ta, tb;\n\t"
      "abs.f16x2 ta, %1;\n\t"
      "abs.f16x2 tb, %2;\n\t"
      "sub.sat.f16x2 %0, ta, tb; }" : "=r"(r) : "r"(*reinterpret_cast<uint32_t *>(&a)), "r"(*reinterpret_cast<uint32_t *>(&b)));
  return *reinterpret_cast<__half2 *>(&r);
}

// Paeth-filtered bytes of two lanes: low byte of (x - paeth(a,b,c) + 1280) in each half.
// png-filter.ts:16-25: a if pa <= pb && pa <= pc, else b if pb <= pc, else c.
__device__ __forceinline__ __half2 paeth_resid_h2(__half2 x, __half2 a, __half2 b, __half2 c) {
  const __half2 d1 = __hsub2(b, c);   // p - a
  const __half2 d2 = __hsub2(a, c);   // p - b
  const __half2 s = __hadd2(d1, d2);  // p - c
  const __half2 u = abs_gt_h2(d2, s);                       // pb > pc: c beats b
  const __half2 m = __hmin2(__habs2(d2), __habs2(s));
  const __half2 v = abs_gt_h2(d1, m);                       // pa > min(pb, pc): not a
  const __half2 pk1 = __hfma2(__hneg2(u), d1, b);           // u ? c : b
  const __half2 sel = __hfma2(v, __hsub2(pk1, a), a);       // v ? pk1 : a
  return __hadd2(__hsub2(x, sel), __float2half2_rn(1280.f));  // exact, in [1025, 1535]
}

// ---- generic compositor: one 16-byte quad of canvas row y ------------------------
template <int BPW, class P>
__device__ __forceinline__ uint4 compose_quad(const P &p, const uint32_t *list, uint32_t n_list, int xw, int y) {
  constexpr int PX = 4 / BPW;  // pixels per quad
  const int x = xw / BPW;
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < PX; i++) {
    if (BPW == 1) w[i] = p.bg_lo; else { w[2 * i] = p.bg_lo; w[2 * i + 1] = p.bg_hi; }
  }
  // start at the top-most tile that covers the whole quad and hides everything below it
  int k0 = 0;
  for (int k = (int)n_list - 1; k > 0; k--) {
    const DevTile &t = p.tiles[list[k]];
    if ((p.alpha_blend == 0u || t.opaque) && y >= t.y0 && y < t.y1 && x >= t.x0 && x + PX <= t.x1) { k0 = k; break; }
  }
  for (int k = k0; k < (int)n_list; k++) {
    const DevTile &t = p.tiles[list[k]];
    if (y < t.y0 || y >= t.y1 || x + PX <= t.x0 || x >= t.x1) continue;
    const uint32_t sy = (uint32_t)(y - t.y0 + t.sy0);
    const int sx0 = x - t.x0 + t.sx0;
    const bool inside = x >= t.x0 && x + PX <= t.x1;
    if (BPW == 1) {
      uint32_t s[4];
      if (inside && t.fmt == FMT_RGBA8) {
        const uint8_t *src = t.pix + (size_t)sy * t.pitch + 4u * (uint32_t)sx0;
        if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
          const uint4 v = __ldg(reinterpret_cast<const uint4 *>(src));
          s[0] = v.x; s[1] = v.y; s[2] = v.z; s[3] = v.w;
        } else {
#pragma unroll
          for (int i = 0; i < 4; i++) s[i] = __ldg(reinterpret_cast<const uint32_t *>(src) + i);
        }
#pragma unroll
        for (int i = 0; i < 4; i++) w[i] = p.alpha_blend ? blend8(w[i], s[i]) : s[i];
      } else {
#pragma unroll
        for (int i = 0; i < 4; i++) {
          if (x + i >= t.x0 && x + i < t.x1) {
            const uint32_t sp = fetch_rgba8(t, (uint32_t)(sx0 + i), sy);
            w[i] = p.alpha_blend ? blend8(w[i], sp) : sp;
          }
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 2; i++) {
        if (x + i >= t.x0 && x + i < t.x1) {
          const uint2 sp = fetch_rgba16(t, (uint32_t)(sx0 + i), sy);
          const uint2 d = p.alpha_blend ? blend16(make_uint2(w[2 * i], w[2 * i + 1]), sp) : sp;
          w[2 * i] = d.x; w[2 * i + 1] = d.y;
        }
      }
    }
  }
  // words past the right edge of the canvas do not exist
  const int wmax = (int)p.canvas_w * BPW;
#pragma unroll
  for (int j = 0; j < 4; j++) if (xw + j >= wmax) w[j] = 0u;
  return make_uint4(w[0], w[1], w[2], w[3]);
}

}  // namespace stitch
