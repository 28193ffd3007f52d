// Training kernels of the diffusion head: q_sample, the MSE + VLB row loss with its gradient, the
// masked mean, and the row-wise / reduction parts of the backward pass.  The GEMM-shaped parts of the
// backward (data and weight gradients) run on the GEMM cores (umma_gemm.cuh / sgemm.cuh).
// Reference arithmetic: GD = diffusion/gaussian_diffusion.py, DU = diffusion/diffusion_utils.py,
// DL = diffusion_loss.py.
#pragma once
#include "common.cuh"

namespace uva {

// ------------------------------------------------------------------------------------------
// q_sample (GD:220-236): x_t = sqrt(acp[t]) x0 + sqrt(1 - acp[t]) noise, t per row.
// ------------------------------------------------------------------------------------------
__global__ void k_qsample(const float* __restrict__ x0, const float* __restrict__ noise,
                          const int64_t* __restrict__ t, const float* __restrict__ sqrt_ac,
                          const float* __restrict__ sqrt_1mac, float* __restrict__ xt, int64_t rows, int C) {
  const int64_t n = rows * C;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / C;
    const int64_t tt = t[r];
    xt[i] = sqrt_ac[tt] * x0[i] + sqrt_1mac[tt] * noise[i];
  }
}

// ------------------------------------------------------------------------------------------
// Row loss of training_losses (GD:746-818, LossType.MSE + LEARNED_RANGE + EPSILON):
//   mse = mean_C (noise - eps)^2                                              GD:810
//   vb  = _vb_terms_bpd with eps detached (GD:789-796, 713-744):
//         t > 0 : mean_C normal_kl(q posterior, p(mean(eps), logvar(v))) / ln 2   DU:10-36
//         t = 0 : mean_C -discretized_gaussian_log_likelihood(x0; mean, 0.5 logvar) / ln 2   DU:47-73
//   row_loss = mse + vb                                                       GD:812
// and d row_loss / d net output (eps through the MSE only, v through the VLB only).
// One thread per row (C <= 16).  dnet: [rows, ldn] (columns >= 2C are written as 0).
// ------------------------------------------------------------------------------------------
struct TrainTables {
  const float *sqrt_ac, *sqrt_1mac, *sqrt_recip, *sqrt_recipm1, *coef1, *coef2, *post_logvar, *log_beta;
};

__device__ __forceinline__ void approx_cdf(float x, float& cdf, float& dcdf) {
  // DU:39-44: 0.5 (1 + tanh(sqrt(2/pi) (x + 0.044715 x^3)))
  const float k = 0.7978845608028654f;
  const float th = tanhf(k * (x + 0.044715f * x * x * x));
  cdf = 0.5f * (1.0f + th);
  dcdf = 0.5f * (1.0f - th * th) * k * (1.0f + 3.0f * 0.044715f * x * x);
}

__global__ void k_loss_rows(const float* __restrict__ net, int ldn, const float* __restrict__ x0,
                            const float* __restrict__ xt, const float* __restrict__ noise,
                            const int64_t* __restrict__ t, TrainTables tb, int64_t rows, int C,
                            float* __restrict__ row_loss, float* __restrict__ dnet) {
  const float inv_ln2 = 1.4426950408889634f;
  for (int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; r < rows; r += (int64_t)gridDim.x * blockDim.x) {
    const int64_t tt = t[r];
    const float min_log = tb.post_logvar[tt], max_log = tb.log_beta[tt];
    const float srec = tb.sqrt_recip[tt], srecm1 = tb.sqrt_recipm1[tt];
    const float c1 = tb.coef1[tt], c2 = tb.coef2[tt];
    const float dlv_dv = 0.5f * (max_log - min_log);
    float mse = 0.f, vb = 0.f;
    for (int c = 0; c < C; ++c) {
      const float eps = net[r * ldn + c];
      const float v = net[r * ldn + C + c];
      const float nz = noise[r * C + c];
      const float x = xt[r * C + c];
      const float xs = x0[r * C + c];
      const float d = nz - eps;
      mse += d * d;
      const float frac = (v + 1.0f) / 2.0f;
      const float lv = frac * max_log + (1.0f - frac) * min_log;             // GD:304-309
      const float px0 = srec * x - srecm1 * eps;                              // GD:348-353 (no clip, GD:793)
      const float mean = c1 * px0 + c2 * x;                                   // GD:238-258
      const float tmean = c1 * xs + c2 * x;
      float term, dterm_dlv;
      if (tt != 0) {
        const float e1 = expf(min_log - lv);
        const float dm = tmean - mean;
        const float e2 = expf(-lv);
        term = 0.5f * (-1.0f + lv - min_log + e1 + dm * dm * e2);             // DU:10-36
        dterm_dlv = 0.5f * (1.0f - e1 - dm * dm * e2);
      } else {
        // DU:47-73 with log_scales = 0.5 lv; d/d log_scales of the *_in arguments is -(*_in)
        const float cx = xs - mean;
        const float inv_std = expf(-0.5f * lv);
        const float pin = inv_std * (cx + 1.0f / 255.0f);
        const float min_ = inv_std * (cx - 1.0f / 255.0f);
        float cp, dcp, cm, dcm;
        approx_cdf(pin, cp, dcp);
        approx_cdf(min_, cm, dcm);
        float ll, dll_dls;
        if (xs < -0.999f) {
          ll = logf(fmaxf(cp, 1e-12f));
          dll_dls = (cp >= 1e-12f) ? (dcp * (-pin)) / cp : 0.f;
        } else if (xs > 0.999f) {
          const float om = 1.0f - cm;
          ll = logf(fmaxf(om, 1e-12f));
          dll_dls = (om >= 1e-12f) ? (-dcm * (-min_)) / om : 0.f;
        } else {
          const float dl = cp - cm;
          ll = logf(fmaxf(dl, 1e-12f));
          dll_dls = (dl >= 1e-12f) ? (dcp * (-pin) - dcm * (-min_)) / dl : 0.f;
        }
        term = -ll;
        dterm_dlv = -0.5f * dll_dls;
      }
      vb += term;
      dnet[r * ldn + c] = 2.0f * (eps - nz) / (float)C;
      dnet[r * ldn + C + c] = dterm_dlv * dlv_dv * inv_ln2 / (float)C;
    }
    for (int c = 2 * C; c < ldn; ++c) dnet[r * ldn + c] = 0.f;
    row_loss[r] = mse / (float)C + (vb / (float)C) * inv_ln2;
  }
}

// loss = sum(row_loss * mask) / sum(mask)  (DL:64-66), or mean(row_loss) when mask == null
// (DAL:162-164).  One block; out[0] = loss, out[1] = denominator (kept for the backward).
__global__ void __launch_bounds__(1024)
k_loss_reduce(const float* __restrict__ row_loss, const float* __restrict__ mask, int64_t rows,
              float* __restrict__ loss_out, float* __restrict__ denom_out) {
  __shared__ double sh_s[32], sh_m[32];
  double s = 0.0, m = 0.0;
  for (int64_t r = threadIdx.x; r < rows; r += blockDim.x) {
    const float w = mask ? mask[r] : 1.0f;
    s += (double)(row_loss[r] * w);
    m += (double)w;
  }
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    m += __shfl_xor_sync(0xffffffffu, m, o);
  }
  if ((threadIdx.x & 31) == 0) { sh_s[threadIdx.x >> 5] = s; sh_m[threadIdx.x >> 5] = m; }
  __syncthreads();
  if (threadIdx.x < 32) {
    s = (threadIdx.x < (blockDim.x >> 5)) ? sh_s[threadIdx.x] : 0.0;
    m = (threadIdx.x < (blockDim.x >> 5)) ? sh_m[threadIdx.x] : 0.0;
    for (int o = 16; o > 0; o >>= 1) {
      s += __shfl_xor_sync(0xffffffffu, s, o);
      m += __shfl_xor_sync(0xffffffffu, m, o);
    }
    if (threadIdx.x == 0) {
      loss_out[0] = (float)s / (float)m;
      denom_out[0] = (float)m;
    }
  }
}

// d loss / d row_loss[r] = dloss * mask[r] / sum(mask) (or dloss / rows); scales the rows of dnet.
// out_b (may be null): the same rows as the bf16 operand of the final layer's tensor-core backward; columns
// >= c_valid (padding of the 32-wide net-output buffer) are written as zero.
__global__ void k_dnet_scale(const float* __restrict__ dnet, const float* __restrict__ mask,
                             const float* __restrict__ denom, const float* __restrict__ dloss,
                             float* __restrict__ out, int64_t rows, int ldn, bf16* __restrict__ out_b = nullptr,
                             int c_valid = 0) {
  const int64_t n = rows * ldn;
  const float g = dloss[0] / denom[0];
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / ldn;
    const float v = dnet[i] * (mask ? mask[r] * g : g);
    out[i] = v;
    if (out_b) out_b[i] = __float2bfloat16_rn(((int)(i - r * ldn) < c_valid) ? v : 0.f);
  }
}

}  // namespace uva
