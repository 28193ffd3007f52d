// Row kernels of the fused DiffActLoss pre-net ("conv_fc", DAL:41-61 / 112-124 == 169-181).  The convolution and
// the Linears run on the tcgen05 GEMMs (umma_gemm2.cuh with the conv gather, umma_gemm.cuh); these kernels are the
// glue between them: operand packing, pooling layout, ReLU / bias, the 4 -> 16 interpolation over time, and the
// corresponding backward pieces.
#pragma once
#include "common.cuh"

namespace uva {

// conv.weight fp32 [co][ci][3][3] -> forward operand Wf[co][tap * C + ci] and data-gradient operand
// Wd[ci][(8 - tap) * C + co] (taps flipped: the data gradient of a 3x3 / padding-1 convolution is the same gather
// with the kernel rotated by 180 degrees), both bf16, K-major.
__global__ void k_pack_conv_w(const float* __restrict__ w, bf16* __restrict__ wf, bf16* __restrict__ wd, int C) {
  const int64_t n = (int64_t)C * C * 9;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int tap = (int)(i % 9);
    const int64_t r = i / 9;
    const int ci = (int)(r % C), co = (int)(r / C);
    const bf16 v = __float2bfloat16_rn(w[i]);
    wf[(int64_t)co * 9 * C + (int64_t)tap * C + ci] = v;
    wd[(int64_t)ci * 9 * C + (int64_t)(8 - tap) * C + co] = v;
  }
}

// pooled fp32 [img][16 regions][C] -> bf16 [img][c * 16 + region]  ("b c w h -> b (c w h)", DAL:116)
__global__ void k_pool_to_bf16(const float* __restrict__ pooled, bf16* __restrict__ out, int64_t imgs, int C) {
  const int64_t n = imgs * 16 * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int reg = (int)(i % 16);
    const int64_t r = i / 16;
    const int c = (int)(r % C);
    const int64_t img = r / C;
    out[i] = __float2bfloat16_rn(pooled[(img * 16 + reg) * C + c]);
  }
}

// out(bf16) = max(bias + sum_s part[s], 0): the split-K partial planes of fc.0 reduced in a FIXED order (bit-reproducible,
// unlike an atomic reduction) -> operand of fc.2.  part: [S][plane_rows][C], rows <= plane_rows.
__global__ void k_splitk_bias_relu_bf16(const float* __restrict__ part, const float* __restrict__ bias, bf16* __restrict__ out,
                                        int S, int64_t plane_rows, int64_t rows, int C) {
  const int64_t n = rows * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    float s = bias[i % C];
    for (int k = 0; k < S; ++k) s += part[(int64_t)k * plane_rows * C + i];
    out[i] = __float2bfloat16_rn(fmaxf(s, 0.f));
  }
}

// interpolate = Linear(n_frames -> n_actions) over the time axis (DAL:118-121):
// out[b, j, c] = bi[j] + sum_t Wi[j, t] f2[b, t, c]   -> bf16 operand of refine.0
__global__ void k_interp(const float* __restrict__ f2, const float* __restrict__ wi, const float* __restrict__ bi,
                         bf16* __restrict__ out, int64_t B, int T, int J, int C) {
  const int64_t n = B * J * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    const int64_t r = i / C;
    const int j = (int)(r % J);
    const int64_t b = r / J;
    float s = bi[j];
    for (int t = 0; t < T; ++t) s = fmaf(wi[j * T + t], f2[(b * T + t) * C + c], s);
    out[i] = __float2bfloat16_rn(s);
  }
}

// ---- backward -------------------------------------------------------------------------------------------------
__global__ void k_f32_to_bf16(const float* __restrict__ x, bf16* __restrict__ y, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    y[i] = __float2bfloat16_rn(x[i]);
}
// dx(bf16) = dy * (act > 0)   (ReLU backward; act is the bf16 ReLU OUTPUT kept by the forward)
__global__ void k_relu_bwd_bf16(const float* __restrict__ dy, const bf16* __restrict__ act, bf16* __restrict__ dx, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    dx[i] = __float2bfloat16_rn(__bfloat162float(act[i]) > 0.f ? dy[i] : 0.f);
}

// interpolate backward: df2[b, t, c] = sum_j Wi[j, t] d[b, j, c] (bf16 + the column sums of df2 come later);
// dWi[j, t] += sum_{b, c} d[b, j, c] f2[b, t, c];  dbi[j] += sum_{b, c} d[b, j, c].   One block per (b, 256 channels).
constexpr int IB_MAXT = 8, IB_MAXJ = 16;
__global__ void __launch_bounds__(256)
k_interp_bwd(const float* __restrict__ d, const float* __restrict__ f2, const float* __restrict__ wi, bf16* __restrict__ df2,
             float* __restrict__ dwi, float* __restrict__ dbi, int64_t B, int T, int J, int C) {
  const int64_t b = blockIdx.y;
  const int c = blockIdx.x * 256 + threadIdx.x;
  float dj[IB_MAXJ], ft[IB_MAXT];
  const bool ok = c < C;
  for (int j = 0; j < J; ++j) dj[j] = ok ? d[(b * J + j) * C + c] : 0.f;
  for (int t = 0; t < T; ++t) ft[t] = ok ? f2[(b * T + t) * C + c] : 0.f;
  if (ok)
    for (int t = 0; t < T; ++t) {
      float s = 0.f;
      for (int j = 0; j < J; ++j) s = fmaf(wi[j * T + t], dj[j], s);
      df2[(b * T + t) * C + c] = __float2bfloat16_rn(s);
    }
  __shared__ float red[8][IB_MAXJ * (IB_MAXT + 1)];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int j = 0; j < J; ++j) {
    for (int t = 0; t <= T; ++t) {
      float v = (t < T) ? dj[j] * ft[t] : dj[j];
      v = warp_sum(v);
      if (lane == 0) red[w][j * (T + 1) + t] = v;
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < J * (T + 1); i += 256) {
    float s = 0.f;
    for (int k = 0; k < 8; ++k) s += red[k][i];
    const int j = i / (T + 1), t = i % (T + 1);
    if (t < T) { if (dwi) atomicAdd(dwi + j * T + t, s); }
    else if (dbi) atomicAdd(dbi + j, s);
  }
}

// AdaptiveAvgPool(4,4) + ReLU backward on 16 x 16 maps: dconv[img, px, c] = y[img, px, c] > 0 ?
// dpooled[img, c * 16 + region(px)] / 16 : 0   (px = w * 16 + h, region = (w / 4) * 4 + h / 4)
// 8 channels (16 bytes of y / dconv) per thread; C % 8 == 0.
__global__ void k_pool_relu_bwd(const float* __restrict__ dpooled, const bf16* __restrict__ y, bf16* __restrict__ dconv,
                                int64_t imgs, int C) {
  const int c8 = C / 8;
  const int64_t n = imgs * 256 * c8;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % c8) * 8;
    const int64_t r = i / c8;
    const int px = (int)(r % 256);
    const int64_t img = r / 256;
    const int reg = ((px >> 4) >> 2) * 4 + ((px & 15) >> 2);
    const uint4 yv = *reinterpret_cast<const uint4*>(y + r * C + c);
    const __nv_bfloat162* y2 = reinterpret_cast<const __nv_bfloat162*>(&yv);
    const float* gp = dpooled + img * 16 * C + (int64_t)c * 16 + reg;
    uint4 ov;
    uint32_t* o = reinterpret_cast<uint32_t*>(&ov);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 f = __bfloat1622float2(y2[k]);
      const float g0 = __ldg(gp + (2 * k) * 16) * 0.0625f, g1 = __ldg(gp + (2 * k + 1) * 16) * 0.0625f;
      o[k] = pack_bf16x2(f.x > 0.f ? g0 : 0.f, f.y > 0.f ? g1 : 0.f);
    }
    *reinterpret_cast<uint4*>(dconv + r * C + c) = ov;
  }
}

// dW[co][ci][tap] += dWp[co][tap * C + ci]   (packed weight-gradient accumulator -> reference layout)
__global__ void k_unpack_conv_wgrad(const float* __restrict__ dwp, float* __restrict__ dw, int C) {
  const int64_t n = (int64_t)C * C * 9;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int tap = (int)(i % 9);
    const int64_t r = i / 9;
    const int ci = (int)(r % C), co = (int)(r / C);
    dw[i] += dwp[(int64_t)co * 9 * C + (int64_t)tap * C + ci];
  }
}

}  // namespace uva
