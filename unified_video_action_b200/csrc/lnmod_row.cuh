// One row of a = LayerNorm_{eps=1e-6}(h) [* ln_w + ln_b] * (1 + scale) + shift (modulate DL:93-94, ResBlock
// DL:163-166, FinalLayer DL:185-188) by one warp: two-pass statistics in registers, NV = W / 128 float4 per lane.
// Shared by k_lnmod (elementwise.cuh) and by the LayerNorm-modulate pre-phase of the latency-path GEMM
// (umma_gemm.cuh), so both produce the same bits.  shift / scale: TA, row stride ldmod; stats (optional): mean, rstd.
#pragma once
#include "common.cuh"

namespace uva {

template <typename TA, int NV>
__device__ __forceinline__ void lnmod_row(const float* __restrict__ h, const float* __restrict__ ln_w,
                                          const float* __restrict__ ln_b, const TA* __restrict__ shift,
                                          const TA* __restrict__ scale, int64_t ldmod, TA* __restrict__ a,
                                          float* __restrict__ stats, int64_t r, int lane) {
  constexpr int W = NV * 128;
  float4 v[NV];
  uint2 ush[NV], usc[NV];  // bf16 shift / scale: requested together with h (one memory round trip per row, not two)
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) v[i] = *reinterpret_cast<const float4*>(h + r * W + (lane + 32 * i) * 4);
  if constexpr (sizeof(TA) == 2) {
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      ush[i] = *reinterpret_cast<const uint2*>(shift + r * ldmod + (lane + 32 * i) * 4);
      usc[i] = *reinterpret_cast<const uint2*>(scale + r * ldmod + (lane + 32 * i) * 4);
    }
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  const float mean = warp_sum(s) * (1.0f / W);
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const float dx = v[i].x - mean, dy = v[i].y - mean, dz = v[i].z - mean, dw = v[i].w - mean;
    ss += (dx * dx + dy * dy) + (dz * dz + dw * dw);
  }
  const float rstd = rsqrtf(warp_sum(ss) * (1.0f / W) + 1e-6f);
  if (stats && lane == 0) {
    stats[2 * r] = mean;
    stats[2 * r + 1] = rstd;
  }
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int j = (lane + 32 * i) * 4;
    float n[4] = {(v[i].x - mean) * rstd, (v[i].y - mean) * rstd, (v[i].z - mean) * rstd,
                  (v[i].w - mean) * rstd};
    if (ln_w) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(ln_w + j));
      const float4 b = __ldg(reinterpret_cast<const float4*>(ln_b + j));
      n[0] = n[0] * g.x + b.x; n[1] = n[1] * g.y + b.y; n[2] = n[2] * g.z + b.z; n[3] = n[3] * g.w + b.w;
    }
    float sh[4], sc[4];
    if constexpr (sizeof(TA) == 2) {
      const uint2 us = ush[i], uc = usc[i];
      const __nv_bfloat162 s0 = *reinterpret_cast<const __nv_bfloat162*>(&us.x), s1 = *reinterpret_cast<const __nv_bfloat162*>(&us.y);
      const __nv_bfloat162 c0 = *reinterpret_cast<const __nv_bfloat162*>(&uc.x), c1 = *reinterpret_cast<const __nv_bfloat162*>(&uc.y);
      sh[0] = __low2float(s0); sh[1] = __high2float(s0); sh[2] = __low2float(s1); sh[3] = __high2float(s1);
      sc[0] = __low2float(c0); sc[1] = __high2float(c0); sc[2] = __low2float(c1); sc[3] = __high2float(c1);
    } else {
      const float4 fs = *reinterpret_cast<const float4*>(shift + r * ldmod + j);
      const float4 fc = *reinterpret_cast<const float4*>(scale + r * ldmod + j);
      sh[0] = fs.x; sh[1] = fs.y; sh[2] = fs.z; sh[3] = fs.w;
      sc[0] = fc.x; sc[1] = fc.y; sc[2] = fc.z; sc[3] = fc.w;
    }
    TA o[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) o[q] = from_f32<TA>(n[q] * (1.0f + sc[q]) + sh[q]);
    if constexpr (sizeof(TA) == 2) {
      *reinterpret_cast<uint2*>(a + r * W + j) = *reinterpret_cast<uint2*>(o);
    } else {
      *reinterpret_cast<float4*>(a + r * W + j) = *reinterpret_cast<float4*>(o);
    }
  }
}

}  // namespace uva
