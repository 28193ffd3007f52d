// Row-wise / elementwise kernels of the diffusion head (HBM-bound; vectorised, one warp per row
// where a row reduction is needed).  Reference arithmetic cited per kernel (DL = diffusion_loss.py,
// GD = diffusion/gaussian_diffusion.py).
#pragma once
#include "common.cuh"
#include "lnmod_row.cuh"

namespace uva {

// ------------------------------------------------------------------------------------------
// weight packing
// ------------------------------------------------------------------------------------------
// dst[r, k] (bf16, rows_pad x cols) = src[r, k] (fp32, rows x cols), zero rows for r >= rows.
__global__ void k_pack_bf16(const float* __restrict__ src, bf16* __restrict__ dst, int rows,
                            int rows_pad, int cols) {
  const int64_t n = (int64_t)rows_pad * cols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / cols);
    dst[i] = (r < rows) ? __float2bfloat16_rn(src[i]) : __float2bfloat16_rn(0.f);
  }
}
// AdaLN modulation weight of one block ([nrow = 3W or 2W, W] fp32: shift | scale | gate rows) -> bf16 in
// tile order for the fused LayerNorm-modulate epilogue: 256-row tiles of (128 shift rows | the matching
// 128 scale rows), then the gate rows unchanged.  The bias gets the same permutation (fp32).
__global__ void k_pack_ada_perm(const float* __restrict__ src, const float* __restrict__ bsrc,
                                bf16* __restrict__ dst, float* __restrict__ bdst, int nrow, int W) {
  const int64_t n = (int64_t)nrow * W;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / W);
    const int k = (int)(i - (int64_t)r * W);
    int sr = r;
    if (r < 2 * W) {
      const int j = r / 256, w = r % 256;
      sr = (w < 128) ? j * 128 + w : W + j * 128 + (w - 128);
    }
    dst[i] = __float2bfloat16_rn(src[(int64_t)sr * W + k]);
    if (k == 0) bdst[r] = bsrc[sr];
  }
}

// One launch re-packs every parameter of a head (uva_head_load_weights): tensor t = blockIdx.y is read once and
// written as (a) the fp32 copy in the handle's own layout, (b) its bf16 GEMM operand (Linear weights, bf16 mode),
// (c) for the AdaLN modulation Linears the bf16 copy in (shift | scale)-tile row order of k_pack_ada_perm, resp.
// the fp32 bias in the same order.  Replaces ~50 device-to-device copies + ~25 pack launches per call.
constexpr int LOAD_MAX_JOBS = 40;
struct LoadJob {
  const float* src;
  float* dst32;
  bf16* dst16;    // may be null
  bf16* perm16;   // may be null: weight rows permuted (element i = row * cols + k)
  float* perm32;  // may be null: bias permuted (element i = row)
  int64_t n;
  int cols;       // row length for perm16 (ignored otherwise)
  int W;          // model width: rows [0, W) shift, [W, 2W) scale, rest unchanged
};
struct LoadBatch {
  LoadJob job[LOAD_MAX_JOBS];
  int count;
};
__device__ __forceinline__ int ada_perm_row(int sr, int W) {
  if (sr < W) return (sr >> 7) * 256 + (sr & 127);
  if (sr < 2 * W) return ((sr - W) >> 7) * 256 + 128 + ((sr - W) & 127);
  return sr;
}
__global__ void __launch_bounds__(256) k_load_params(const __grid_constant__ LoadBatch b) {
  if ((int)blockIdx.y >= b.count) return;
  const LoadJob& j = b.job[blockIdx.y];
  const bool vec = ((reinterpret_cast<uintptr_t>(j.src) | reinterpret_cast<uintptr_t>(j.dst32) | reinterpret_cast<uintptr_t>(j.dst16) |
                     reinterpret_cast<uintptr_t>(j.perm16)) & 15) == 0 && (j.cols % 4 == 0 || j.perm16 == nullptr);
  const int64_t n4 = vec ? j.n / 4 : 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(j.src)[i];
    reinterpret_cast<float4*>(j.dst32)[i] = v;
    const uint2 pk = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
    if (j.dst16) reinterpret_cast<uint2*>(j.dst16)[i] = pk;
    if (j.perm16) {
      const int64_t e = 4 * i;
      const int sr = (int)(e / j.cols);
      const int k = (int)(e - (int64_t)sr * j.cols);
      *reinterpret_cast<uint2*>(j.perm16 + (int64_t)ada_perm_row(sr, j.W) * j.cols + k) = pk;
    }
    if (j.perm32) {
      const int e = (int)(4 * i);
      j.perm32[ada_perm_row(e, j.W)] = v.x; j.perm32[ada_perm_row(e + 1, j.W)] = v.y;
      j.perm32[ada_perm_row(e + 2, j.W)] = v.z; j.perm32[ada_perm_row(e + 3, j.W)] = v.w;
    }
  }
  for (int64_t i = 4 * n4 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < j.n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = j.src[i];
    j.dst32[i] = v;
    if (j.dst16) j.dst16[i] = __float2bfloat16_rn(v);
    if (j.perm16) {
      const int sr = (int)(i / j.cols);
      j.perm16[(int64_t)ada_perm_row(sr, j.W) * j.cols + (i - (int64_t)sr * j.cols)] = __float2bfloat16_rn(v);
    }
    if (j.perm32) j.perm32[ada_perm_row((int)i, j.W)] = v;
  }
}

template <typename TO>
__global__ void k_convert(const float* __restrict__ src, TO* __restrict__ dst, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = from_f32<TO>(src[i]);
}

// ------------------------------------------------------------------------------------------
// ys = SiLU(cond_embed(c) + time_embed(t))   — first op of every adaLN_modulation (DL:158-160,
// 181-183, 271-273).  temb is the [T_train, W] table; the timestep of row r is t_rows[r] when
// t_rows != null, else t_all.   y_out (optional, TA) keeps y for the backward.
// ------------------------------------------------------------------------------------------
// SiLU of the sampling-path row kernels: the bf16 operand takes the one-MUFU form the GEMM epilogues use (the kernel was
// bound by the ~20 instructions per element of __expf + IEEE division, not by HBM); fp32 mode keeps the accurate form.
template <typename TA>
__device__ __forceinline__ float silu_out(float x) {
  if constexpr (sizeof(TA) == 2) return silu_tanh(x);
  else return silu_f(x);
}

template <typename TA>
__global__ void k_ysilu(const float* __restrict__ cemb, const float* __restrict__ temb,
                        const int64_t* __restrict__ t_rows, int t_all, TA* __restrict__ ys,
                        TA* __restrict__ y_out, int64_t rows, int W) {
  const int wq = W / 4;
  const int64_t n = rows * wq;
  if (wq == (int)blockDim.x && y_out == nullptr) {
    // sampling form (one block = one row, two rows per iteration): no 64-bit division per element, both 16-byte loads
    // of the pair requested together
    const int j = (int)threadIdx.x * 4;
    for (int64_t r = blockIdx.x; r < rows; r += 2 * (int64_t)gridDim.x) {
      const int64_t r2 = r + gridDim.x;
      const bool two = r2 < rows;
      const float4 c0 = *reinterpret_cast<const float4*>(cemb + r * W + j);
      const float4 c1 = two ? *reinterpret_cast<const float4*>(cemb + r2 * W + j) : make_float4(0.f, 0.f, 0.f, 0.f);
      const int64_t t0 = t_rows ? t_rows[r] : (int64_t)t_all;
      const int64_t t1 = (t_rows && two) ? t_rows[r2] : t0;
      const float4 e0 = __ldg(reinterpret_cast<const float4*>(temb + t0 * W + j));
      const float4 e1 = __ldg(reinterpret_cast<const float4*>(temb + t1 * W + j));
      TA o[4];
      o[0] = from_f32<TA>(silu_out<TA>(c0.x + e0.x)); o[1] = from_f32<TA>(silu_out<TA>(c0.y + e0.y));
      o[2] = from_f32<TA>(silu_out<TA>(c0.z + e0.z)); o[3] = from_f32<TA>(silu_out<TA>(c0.w + e0.w));
      if constexpr (sizeof(TA) == 2) *reinterpret_cast<uint2*>(ys + r * W + j) = *reinterpret_cast<uint2*>(o);
      else *reinterpret_cast<float4*>(ys + r * W + j) = *reinterpret_cast<float4*>(o);
      if (two) {
        o[0] = from_f32<TA>(silu_out<TA>(c1.x + e1.x)); o[1] = from_f32<TA>(silu_out<TA>(c1.y + e1.y));
        o[2] = from_f32<TA>(silu_out<TA>(c1.z + e1.z)); o[3] = from_f32<TA>(silu_out<TA>(c1.w + e1.w));
        if constexpr (sizeof(TA) == 2) *reinterpret_cast<uint2*>(ys + r2 * W + j) = *reinterpret_cast<uint2*>(o);
        else *reinterpret_cast<float4*>(ys + r2 * W + j) = *reinterpret_cast<float4*>(o);
      }
    }
    return;
  }
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / wq;
    const int j = (int)(i - r * wq) * 4;
    const int64_t t = t_rows ? t_rows[r] : (int64_t)t_all;
    const float4 c = *reinterpret_cast<const float4*>(cemb + r * W + j);
    const float4 e = __ldg(reinterpret_cast<const float4*>(temb + t * W + j));
    float y[4] = {c.x + e.x, c.y + e.y, c.z + e.z, c.w + e.w};
    TA o[4];
    if (y_out) {
#pragma unroll
      for (int q = 0; q < 4; ++q) o[q] = from_f32<TA>(y[q]);
      if constexpr (sizeof(TA) == 2) *reinterpret_cast<uint2*>(y_out + r * W + j) = *reinterpret_cast<uint2*>(o);
      else *reinterpret_cast<float4*>(y_out + r * W + j) = *reinterpret_cast<float4*>(o);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) o[q] = from_f32<TA>(y_out ? silu_f(y[q]) : silu_out<TA>(y[q]));  // (training keeps the accurate form)
    if constexpr (sizeof(TA) == 2) {
      *reinterpret_cast<uint2*>(ys + r * W + j) = *reinterpret_cast<uint2*>(o);
    } else {
      *reinterpret_cast<float4*>(ys + r * W + j) = *reinterpret_cast<float4*>(o);
    }
  }
}

// The same for a group of denoise steps at once (latency path: the condition embedding does not change
// over the chain, so SiLU(c + temb[t_k]) of every step k is known up front): ys[(g, r), :] for step
// k0 + g, g < G; tsteps[k] = original-process timestep of loop iteration k.
__global__ void k_ysilu_steps(const float* __restrict__ cemb, const float* __restrict__ temb,
                              const int* __restrict__ tsteps, int k0, int G, bf16* __restrict__ ys,
                              int64_t rows, int W) {
  const int wq = W / 4;
  const int64_t per = rows * wq;
  const int64_t n = per * G;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int g = (int)(i / per);
    const int64_t rem = i - (int64_t)g * per;
    const int64_t r = rem / wq;
    const int j = (int)(rem - r * wq) * 4;
    const int64_t t = tsteps[k0 + g];
    const float4 c = *reinterpret_cast<const float4*>(cemb + r * W + j);
    const float4 e = __ldg(reinterpret_cast<const float4*>(temb + t * W + j));
    bf16 o[4] = {__float2bfloat16_rn(silu_tanh(c.x + e.x)), __float2bfloat16_rn(silu_tanh(c.y + e.y)),
                 __float2bfloat16_rn(silu_tanh(c.z + e.z)), __float2bfloat16_rn(silu_tanh(c.w + e.w))};
    *reinterpret_cast<uint2*>(ys + ((int64_t)g * rows + r) * W + j) = *reinterpret_cast<uint2*>(o);
  }
}

// ------------------------------------------------------------------------------------------
// input_proj (DL:233, 268): h[r, :] = x[src(r), :] . Wi^T + bi,  K = C <= 16 (CUDA cores, fp32:
// the diffusion state is never rounded to bf16).  With CFG (half > 0) the net input of row r is
// x[r % half] (forward_with_cfg, DL:286-287).  One block = 32 rows x W columns; each thread keeps
// the weights of 4 output columns in registers.
// ------------------------------------------------------------------------------------------
constexpr int IP_ROWS = 64;
__global__ void __launch_bounds__(256)
k_input_proj(const float* __restrict__ x, const float* __restrict__ Wi, const float* __restrict__ bi,
             float* __restrict__ h, int64_t rows, int C, int W, int64_t half, float* __restrict__ stats,
             int dup = 0) {
  // dup (CFG, rows = 2 * half): the uncond row r + half has the input of the cond row r (DL:286-287), so a block
  // covers rows of the first half only and writes every result (and its row sums) to both rows
  __shared__ __align__(16) float xs[IP_ROWS][16];  // the block's input rows, zero padded to 16 channels
  const int64_t r0 = (int64_t)blockIdx.x * IP_ROWS;
  const int64_t left = (dup ? half : rows) - r0;
  const int64_t twin = dup ? half : 0;
  const int nr = left < IP_ROWS ? (int)left : IP_ROWS;
  for (int i = threadIdx.x; i < IP_ROWS * 16; i += blockDim.x) {
    const int rr = i >> 4, c = i & 15;
    float v = 0.f;
    if (rr < nr && c < C) {
      const int64_t r = r0 + rr;
      const int64_t rs = (half > 0 && r >= half) ? (r - half) : r;  // rows < 2 * half
      v = __ldg(x + rs * C + c);
    }
    xs[rr][c] = v;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  for (int j0 = threadIdx.x * 4; j0 < W; j0 += blockDim.x * 4) {
    float w[4][16];
    float b[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      b[q] = bi[j0 + q];
      if (C == 16) {  // 4 x 16-byte loads per output column
        const float4* wp = reinterpret_cast<const float4*>(Wi + (int64_t)(j0 + q) * 16);
#pragma unroll
        for (int c4 = 0; c4 < 4; ++c4) {
          const float4 t = __ldg(wp + c4);
          w[q][4 * c4] = t.x; w[q][4 * c4 + 1] = t.y; w[q][4 * c4 + 2] = t.z; w[q][4 * c4 + 3] = t.w;
        }
      } else {
#pragma unroll
        for (int c = 0; c < 16; ++c) w[q][c] = (c < C) ? Wi[(int64_t)(j0 + q) * C + c] : 0.f;
      }
    }
    float* hp = h + r0 * W + j0;
    float* sp = stats != nullptr ? stats + (r0 * (W / 128) + j0 / 128) * 2 : nullptr;
    // two rows per iteration (independent FMA chains); padded rows compute zeros and are not stored
#pragma unroll 1
    for (int rr = 0; rr < nr; rr += 2) {
      float acc[2][4];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const float4* xp = reinterpret_cast<const float4*>(xs[rr + u]);
        float xv[16];
#pragma unroll
        for (int c4 = 0; c4 < 4; ++c4) {
          const float4 t = xp[c4];
          xv[4 * c4] = t.x; xv[4 * c4 + 1] = t.y; xv[4 * c4 + 2] = t.z; xv[4 * c4 + 3] = t.w;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[u][q] = b[q];
#pragma unroll
        for (int c = 0; c < 16; ++c) {
#pragma unroll
          for (int q = 0; q < 4; ++q) acc[u][q] = fmaf(xv[c], w[q][c], acc[u][q]);
        }
      }
      const bool two = rr + 1 < nr;  // block-uniform
      *reinterpret_cast<float4*>(hp + (int64_t)rr * W) = make_float4(acc[0][0], acc[0][1], acc[0][2], acc[0][3]);
      if (two) *reinterpret_cast<float4*>(hp + (int64_t)(rr + 1) * W) = make_float4(acc[1][0], acc[1][1], acc[1][2], acc[1][3]);
      if (twin) {
        *reinterpret_cast<float4*>(hp + (twin + rr) * W) = make_float4(acc[0][0], acc[0][1], acc[0][2], acc[0][3]);
        if (two) *reinterpret_cast<float4*>(hp + (twin + rr + 1) * W) = make_float4(acc[1][0], acc[1][1], acc[1][2], acc[1][3]);
      }
      if (sp != nullptr) {  // row sums for the LayerNorm of the first block: one (sum, sum of squares) per 128 columns
        float s0 = (acc[0][0] + acc[0][1]) + (acc[0][2] + acc[0][3]);
        float q0 = (acc[0][0] * acc[0][0] + acc[0][1] * acc[0][1]) + (acc[0][2] * acc[0][2] + acc[0][3] * acc[0][3]);
        float s1 = (acc[1][0] + acc[1][1]) + (acc[1][2] + acc[1][3]);
        float q1 = (acc[1][0] * acc[1][0] + acc[1][1] * acc[1][1]) + (acc[1][2] * acc[1][2] + acc[1][3] * acc[1][3]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          s0 += __shfl_xor_sync(0xffffffffu, s0, o);
          q0 += __shfl_xor_sync(0xffffffffu, q0, o);
          s1 += __shfl_xor_sync(0xffffffffu, s1, o);
          q1 += __shfl_xor_sync(0xffffffffu, q1, o);
        }
        if (lane == 0) {
          *reinterpret_cast<float2*>(sp + (int64_t)rr * (W / 128) * 2) = make_float2(s0, q0);
          if (two) *reinterpret_cast<float2*>(sp + (int64_t)(rr + 1) * (W / 128) * 2) = make_float2(s1, q1);
          if (twin) {
            *reinterpret_cast<float2*>(sp + (twin + rr) * (W / 128) * 2) = make_float2(s0, q0);
            if (two) *reinterpret_cast<float2*>(sp + (twin + rr + 1) * (W / 128) * 2) = make_float2(s1, q1);
          }
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// a = LayerNorm_{eps=1e-6}(h) [* ln_w + ln_b] * (1 + scale) + shift      (modulate DL:93-94,
// ResBlock DL:163-166, FinalLayer DL:185-188).  One warp per row, two-pass statistics in
// registers.  NV = W / 128 float4 per lane.  shift/scale: TA, row stride ldmod.
// Optionally stores mean / rstd (training backward).
// ------------------------------------------------------------------------------------------
template <typename TA, int NV>
__global__ void __launch_bounds__(256)
k_lnmod(const float* __restrict__ h, const float* __restrict__ ln_w, const float* __restrict__ ln_b,
        const TA* __restrict__ shift, const TA* __restrict__ scale, int64_t ldmod,
        TA* __restrict__ a, float* __restrict__ stats, int64_t rows) {
  griddep_wait();
  griddep_launch();
  const int lane = threadIdx.x & 31;
  const int64_t warp_global = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t warps_total = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp_global; r < rows; r += warps_total)
    lnmod_row<TA, NV>(h, ln_w, ln_b, shift, scale, ldmod, a, stats, r, lane);
}

// Generic-width variant (W multiple of 32, any size; three passes through L1) for test geometries.
template <typename TA>
__global__ void __launch_bounds__(256)
k_lnmod_generic(const float* __restrict__ h, const float* __restrict__ ln_w,
                const float* __restrict__ ln_b, const TA* __restrict__ shift,
                const TA* __restrict__ scale, int64_t ldmod, TA* __restrict__ a,
                float* __restrict__ stats, int64_t rows, int W) {
  const int lane = threadIdx.x & 31;
  const int64_t warp_global = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t warps_total = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp_global; r < rows; r += warps_total) {
    float s = 0.f;
    for (int j = lane; j < W; j += 32) s += h[r * W + j];
    const float mean = warp_sum(s) / (float)W;
    float ss = 0.f;
    for (int j = lane; j < W; j += 32) {
      const float d = h[r * W + j] - mean;
      ss += d * d;
    }
    const float rstd = rsqrtf(warp_sum(ss) / (float)W + 1e-6f);
    if (stats && lane == 0) {
      stats[2 * r] = mean;
      stats[2 * r + 1] = rstd;
    }
    for (int j = lane; j < W; j += 32) {
      float n = (h[r * W + j] - mean) * rstd;
      if (ln_w) n = n * ln_w[j] + ln_b[j];
      a[r * W + j] = from_f32<TA>(n * (1.0f + to_f32(scale[r * ldmod + j])) + to_f32(shift[r * ldmod + j]));
    }
  }
}

// ------------------------------------------------------------------------------------------
// Latency path (rows <= 1024, W = 1024): input_proj (DL:268) and the LayerNorm-modulate of the first ResBlock
// (DL:163-165) in one kernel.  One block = IPL_ROWS rows x all W columns (thread t owns columns 4t .. 4t+3 and
// keeps their input_proj weights in registers), so the row statistics are a block reduction: h = x . Wi^T + bi
// is written as the fp32 residual stream and a = LN(h) * ln_w + ln_b) * (1 + scale) + shift as the bf16 operand of
// the first GEMM.  Same two-pass statistics as k_lnmod.  Saves one dependent launch per denoise step and spreads
// the projection over rows / 8 blocks instead of rows / 64.
// ------------------------------------------------------------------------------------------
constexpr int IPL_ROWS = 8;
__global__ void __launch_bounds__(256)
k_input_proj_lnmod(const float* __restrict__ x, const float* __restrict__ Wi, const float* __restrict__ bi,
                   float* __restrict__ h, bf16* __restrict__ a, const float* __restrict__ ln_w,
                   const float* __restrict__ ln_b, const bf16* __restrict__ shift, const bf16* __restrict__ scale,
                   int64_t ldmod, int64_t rows, int C, int64_t half) {
  constexpr int W = 1024;
  __shared__ __align__(16) float xs[IPL_ROWS][16];
  __shared__ float red[2][IPL_ROWS][8];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t r0 = (int64_t)blockIdx.x * IPL_ROWS;
  const int nr = (rows - r0 < IPL_ROWS) ? (int)(rows - r0) : IPL_ROWS;
  griddep_wait();
  griddep_launch();
  if (tid < IPL_ROWS * 16) {
    const int rr = tid >> 4, c = tid & 15;
    float v = 0.f;
    if (rr < nr && c < C) {
      const int64_t r = r0 + rr;
      const int64_t rs = (half > 0 && r >= half) ? (r - half) : r;
      v = __ldg(x + rs * C + c);
    }
    xs[rr][c] = v;
  }
  const int j0 = tid * 4;
  float w[4][16], b[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {  // weights do not depend on the previous kernel: loaded before griddep_wait
    b[q] = __ldg(bi + j0 + q);
#pragma unroll
    for (int c = 0; c < 16; ++c) w[q][c] = (c < C) ? __ldg(Wi + (int64_t)(j0 + q) * C + c) : 0.f;
  }
  // issued early: the modulation and LayerNorm vectors of this thread's columns
  const float4 g4 = __ldg(reinterpret_cast<const float4*>(ln_w + j0));
  const float4 b4 = __ldg(reinterpret_cast<const float4*>(ln_b + j0));
  __syncthreads();
  float acc[IPL_ROWS][4];
#pragma unroll
  for (int rr = 0; rr < IPL_ROWS; ++rr) {
#pragma unroll
    for (int q = 0; q < 4; ++q) acc[rr][q] = b[q];
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      const float xv = xs[rr][c];
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[rr][q] = fmaf(xv, w[q][c], acc[rr][q]);
    }
    const float s = warp_sum((acc[rr][0] + acc[rr][1]) + (acc[rr][2] + acc[rr][3]));
    if (lane == 0) red[0][rr][warp] = s;
  }
  __syncthreads();
  float mean[IPL_ROWS];
#pragma unroll
  for (int rr = 0; rr < IPL_ROWS; ++rr) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += red[0][rr][i];
    mean[rr] = s * (1.0f / W);
    const float d0 = acc[rr][0] - mean[rr], d1 = acc[rr][1] - mean[rr], d2 = acc[rr][2] - mean[rr], d3 = acc[rr][3] - mean[rr];
    const float qv = warp_sum((d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3));
    if (lane == 0) red[1][rr][warp] = qv;
  }
  __syncthreads();
#pragma unroll
  for (int rr = 0; rr < IPL_ROWS; ++rr) {
    if (rr < nr) {
      float qs = 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) qs += red[1][rr][i];
      const float rstd = rsqrtf(qs * (1.0f / W) + 1e-6f);
      const int64_t r = r0 + rr;
      *reinterpret_cast<float4*>(h + r * W + j0) = make_float4(acc[rr][0], acc[rr][1], acc[rr][2], acc[rr][3]);
      const uint2 us = *reinterpret_cast<const uint2*>(shift + r * ldmod + j0);
      const uint2 uc = *reinterpret_cast<const uint2*>(scale + r * ldmod + j0);
      const float2 s01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&us.x));
      const float2 s23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&us.y));
      const float2 c01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&uc.x));
      const float2 c23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&uc.y));
      const float n0 = (acc[rr][0] - mean[rr]) * rstd * g4.x + b4.x, n1 = (acc[rr][1] - mean[rr]) * rstd * g4.y + b4.y;
      const float n2 = (acc[rr][2] - mean[rr]) * rstd * g4.z + b4.z, n3 = (acc[rr][3] - mean[rr]) * rstd * g4.w + b4.w;
      *reinterpret_cast<uint2*>(a + r * W + j0) =
          make_uint2(pack_bf16x2(n0 * (1.0f + c01.x) + s01.x, n1 * (1.0f + c01.y) + s01.y),
                     pack_bf16x2(n2 * (1.0f + c23.x) + s23.x, n3 * (1.0f + c23.y) + s23.y));
    }
  }
}

// ------------------------------------------------------------------------------------------
// One DDPM ancestral step on the net output (p_mean_variance GD:260-346 for LEARNED_RANGE /
// EPSILON, q_posterior_mean_variance GD:238-258, p_sample GD:395-444) with the CFG mix of
// forward_with_cfg (DL:288-293) and the temperature scaling folded in.
//   net [rows, ldn] (eps | v) and x_cur / x_next [rows, C] are chunk-local; noise is the caller's
//   [N, C] step-noise slab.  half > 0: chunk rows are cond || uncond (half each); eps = u + s (c - u)
//   from rows (r % half, r % half + half); chunk row r is global row g0 + r (r < half) or
//   g0 + ghalf + (r - half).
// ------------------------------------------------------------------------------------------
struct StepCoef {
  float sqrt_recip, sqrt_recipm1, coef1, coef2, min_log, max_log;
  int nonzero;
};

// Counter-based noise for the sampler when the caller passes no step_noise ("NULL ->
// in-kernel Philox(seed, offset)"): Philox4x32-10 (Salmon et al., SC'11; the generator behind
// torch.randn on CUDA) keyed by the 64-bit seed, counter = (element index / 4, loop iteration k, call
// offset), two Box-Muller pairs per block.  The value of element e of step k depends on nothing else, so
// row chunking, the two-lane overlap and multi-GPU row sharding (with global row indices) do not change
// the samples.  oracle.philox_normal restates it in numpy.
struct PhiloxSeed {
  uint32_t k0, k1;  // seed low / high
  uint32_t offset;  // per-call counter word
  uint32_t on;      // 0: noise comes from memory
};

__device__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// standard normal number `e` of loop iteration `step`
__device__ __forceinline__ float philox_normal(const PhiloxSeed& s, uint32_t step, uint64_t e) {
  uint32_t r[4];
  const uint64_t blk = e >> 2;
  philox4x32_10((uint32_t)blk, (uint32_t)(blk >> 32), step, s.offset, s.k0, s.k1, r);
  const int lane = (int)(e & 3);
  const uint32_t a = r[lane & 2], b = r[(lane & 2) + 1];
  const float u1 = ((float)(a >> 8) + 1.0f) * (1.0f / 16777216.0f);  // (0, 1]
  const float u2 = (float)(b >> 8) * (1.0f / 16777216.0f);           // [0, 1)
  const float rad = sqrtf(-2.0f * logf(u1));
  float sn, cs;
  sincospif(2.0f * u2, &sn, &cs);
  return rad * ((lane & 1) ? sn : cs);
}

__global__ void k_philox_fill(PhiloxSeed s, uint32_t step0, int64_t per_step, int64_t n, float* __restrict__ out) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t j = i / per_step;
    out[i] = philox_normal(s, step0 + (uint32_t)j, (uint64_t)(i - j * per_step));
  }
}

// x_{i-1} of one element from the (CFG-mixed) eps, the variance channel v, x_i and the noise draw (from memory at
// `noise[e]`, or Philox element e): learned-range variance GD:283-291, x0-from-eps GD:348-353 (+ clip GD:293-297),
// posterior mean GD:238-258, ancestral step GD:431-444 with the temperature scaling.  Shared by k_ddpm_step and
// the fused final-layer epilogue (final_ddpm.cuh) so that both give the same bits.
// (the arithmetic, with the noise draw nz already in hand: ddpm_update and the vectorised epilogue of final_ddpm.cuh share it)
__device__ __forceinline__ float ddpm_update_nz(float eps, float v, float x, float nz, float temperature, int clip,
                                                const StepCoef& k) {
  const float frac = (v + 1.0f) / 2.0f;
  const float logvar = frac * k.max_log + (1.0f - frac) * k.min_log;
  float x0 = k.sqrt_recip * x - k.sqrt_recipm1 * eps;
  if (clip) x0 = fminf(fmaxf(x0, -1.0f), 1.0f);
  const float mean = k.coef1 * x0 + k.coef2 * x;
  float out = mean;
  if (k.nonzero) out = mean + expf(0.5f * logvar) * nz * temperature;
  return out;
}
__device__ __forceinline__ float ddpm_update(float eps, float v, float x, const float* __restrict__ noise,
                                             const PhiloxSeed& ps, uint32_t step, int64_t e, float temperature,
                                             int clip, const StepCoef& k) {
  float nz = 0.f;
  if (k.nonzero) nz = ps.on ? philox_normal(ps, step, (uint64_t)e) : noise[e];
  return ddpm_update_nz(eps, v, x, nz, temperature, clip, k);
}

__global__ void k_ddpm_step(const float* __restrict__ net, int ldn, const float* __restrict__ x_cur,
                            const float* __restrict__ noise, float* __restrict__ x_next, int64_t rows,
                            int C, int64_t half, int64_t g0, int64_t ghalf, float cfg,
                            float temperature, int clip, StepCoef k, PhiloxSeed ps = PhiloxSeed{0, 0, 0, 0},
                            uint32_t step = 0) {
  const int64_t n = rows * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / C;
    const int c = (int)(i - r * C);
    float eps;
    if (half > 0) {
      const int64_t rc = r % half;
      const float ce = net[rc * ldn + c];
      const float ue = net[(rc + half) * ldn + c];
      eps = ue + cfg * (ce - ue);
    } else {
      eps = net[r * ldn + c];
    }
    const float v = net[r * ldn + C + c];
    // noise lives in the caller's [N, C] layout: chunk-local row r -> global row
    const int64_t gr = g0 + ((half > 0 && r >= half) ? ghalf + (r - half) : r);
    x_next[i] = ddpm_update(eps, v, x_cur[i], noise, ps, step, gr * C + c, temperature, clip, k);
  }
}

// ------------------------------------------------------------------------------------------
// Decoder tail -> condition rows (MAR.forward_mae_decoder, mar_con_unified.py:702-724):
//   z[r, :] = LayerNorm_eps(x_dec[b, skip + tok, :]) * ln_w + ln_b + (temporal[tok / S, :] + spatial[tok % S, :])
// with r = b * T*S + tok (row_index[r] replaces r when given: the gather z[mask_to_pred.nonzero()] of MAR:1090).
// One warp per row, the row stays in registers (two-pass statistics like ATen's LayerNorm), and the result is
// written ONCE as the operand type of cond_embed (bf16 in bf16 mode): the fp32 z of the reference never exists.
// ------------------------------------------------------------------------------------------
struct DecoderTail {
  const float* x;          // [B, L, Z] decoder output BEFORE decoder_norm
  const float* ln_w;       // [Z] decoder_norm.weight
  const float* ln_b;       // [Z] decoder_norm.bias
  const float* temporal;   // [T, Z] diffusion_temporal_embed
  const float* spatial;    // [S, Z] diffusion_spatial_embed
  const int64_t* row_index;  // optional [N]: token-grid index (b * T*S + tok) of output row r
  float eps;
  int T, S;
  int64_t L, skip;         // tokens per sample in x, leading tokens dropped (buffer_size_text, MAR:705-707)
};
static const int DT_MAX_V4 = 8;  // Z <= 1024

template <typename TO>
__global__ void k_decoder_tail(DecoderTail p, int64_t row0, int64_t rows, int Z, TO* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int nv = Z / 128;  // float4 per lane
  const int64_t warps = (int64_t)gridDim.x * (blockDim.x >> 5);
  for (int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); r < rows; r += warps) {
    const int64_t g = p.row_index ? p.row_index[row0 + r] : row0 + r;
    const int64_t TS = (int64_t)p.T * p.S;
    const int64_t b = g / TS;
    const int tok = (int)(g - b * TS);
    const float* xr = p.x + (b * p.L + p.skip + tok) * Z;
    const float* te = p.temporal + (size_t)(tok / p.S) * Z;
    const float* se = p.spatial + (size_t)(tok % p.S) * Z;
    float4 v[DT_MAX_V4];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < DT_MAX_V4; ++i)
      if (i < nv) {
        v[i] = *reinterpret_cast<const float4*>(xr + 128 * i + 4 * lane);
        s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
      }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s / (float)Z;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < DT_MAX_V4; ++i)
      if (i < nv) {
        const float a = v[i].x - mean, bq = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
        q += (a * a + bq * bq) + (c * c + d * d);
      }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q / (float)Z + p.eps);
#pragma unroll
    for (int i = 0; i < DT_MAX_V4; ++i)
      if (i < nv) {
        const int j = 128 * i + 4 * lane;
        const float4 w = __ldg(reinterpret_cast<const float4*>(p.ln_w + j));
        const float4 bb = __ldg(reinterpret_cast<const float4*>(p.ln_b + j));
        const float4 t4 = __ldg(reinterpret_cast<const float4*>(te + j));
        const float4 s4 = __ldg(reinterpret_cast<const float4*>(se + j));
        float o4[4];
        o4[0] = ((v[i].x - mean) * rstd * w.x + bb.x) + (t4.x + s4.x);
        o4[1] = ((v[i].y - mean) * rstd * w.y + bb.y) + (t4.y + s4.y);
        o4[2] = ((v[i].z - mean) * rstd * w.z + bb.z) + (t4.z + s4.z);
        o4[3] = ((v[i].w - mean) * rstd * w.w + bb.w) + (t4.w + s4.w);
        TO* dst = out + r * Z + j;
        if constexpr (sizeof(TO) == 2) {
          TO o[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) o[k] = from_f32<TO>(o4[k]);
          *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<uint2*>(o);
        } else {
          *reinterpret_cast<float4*>(dst) = make_float4(o4[0], o4[1], o4[2], o4[3]);
        }
      }
  }
}

}  // namespace uva
