// Fused AdamW + EMA step over the head's parameter tensors.
//
// The reference trains with torch.optim.AdamW (policy/unified_video_action_policy.py:343-360: decoupled
// weight decay, none on 1-D tensors / biases, :325-341) and keeps an exponential moving average of the
// weights (model/autoregressive/ema_model.py:58-89: ema = ema * decay + p * (1 - decay)).  In eager
// PyTorch that is ~10 passes over the 35.7 M head parameters; here it is ONE pass: read p, g, m, v, ema,
// write p, m, v, ema (36 bytes per parameter, HBM-bound).
#pragma once
#include "common.cuh"

namespace uva {

constexpr int OPT_MAX_TENSORS = 48;

struct AdamWBatch {
  float* p[OPT_MAX_TENSORS];
  const float* g[OPT_MAX_TENSORS];
  float* m[OPT_MAX_TENSORS];
  float* v[OPT_MAX_TENSORS];
  float* ema[OPT_MAX_TENSORS];  // null: no EMA copy for this tensor
  int64_t n[OPT_MAX_TENSORS];
  float wd[OPT_MAX_TENSORS];    // weight decay of this tensor
  int count;
  float lr, beta1, beta2, eps, ema_decay, grad_scale;
  float bc1, bc2_sqrt;          // 1 - beta1^step, sqrt(1 - beta2^step)
};

// torch.optim.AdamW (single-tensor form, amsgrad = False, maximize = False):
//   p *= 1 - lr * wd;  m = b1 m + (1 - b1) g;  v = b2 v + (1 - b2) g^2;
//   denom = sqrt(v) / sqrt(1 - b2^t) + eps;  p -= (lr / (1 - b1^t)) * m / denom
__device__ __forceinline__ void adamw_one(float& p, float g, float& m, float& v, float* ema, const AdamWBatch& b, float wd) {
  g *= b.grad_scale;
  p *= 1.0f - b.lr * wd;
  m = m + (1.0f - b.beta1) * (g - m);       // exp_avg.lerp_(grad, 1 - beta1)
  v = b.beta2 * v + (1.0f - b.beta2) * g * g;
  const float denom = sqrtf(v) / b.bc2_sqrt + b.eps;
  p -= (b.lr / b.bc1) * (m / denom);
  if (ema != nullptr) *ema = *ema * b.ema_decay + p * (1.0f - b.ema_decay);
}

__global__ void __launch_bounds__(256) k_adamw_ema(const __grid_constant__ AdamWBatch b) {
  const int t = blockIdx.y;
  if (t >= b.count) return;
  float* p = b.p[t];
  const float* g = b.g[t];
  float* m = b.m[t];
  float* v = b.v[t];
  float* e = b.ema[t];
  const int64_t n = b.n[t];
  const float wd = b.wd[t];
  const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                     reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(e)) & 15) == 0;
  const int64_t n4 = vec ? n / 4 : 0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 P = reinterpret_cast<float4*>(p)[i];
    const float4 G = reinterpret_cast<const float4*>(g)[i];
    float4 M = reinterpret_cast<float4*>(m)[i];
    float4 V = reinterpret_cast<float4*>(v)[i];
    float4 E = e ? reinterpret_cast<float4*>(e)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    adamw_one(P.x, G.x, M.x, V.x, e ? &E.x : nullptr, b, wd);
    adamw_one(P.y, G.y, M.y, V.y, e ? &E.y : nullptr, b, wd);
    adamw_one(P.z, G.z, M.z, V.z, e ? &E.z : nullptr, b, wd);
    adamw_one(P.w, G.w, M.w, V.w, e ? &E.w : nullptr, b, wd);
    reinterpret_cast<float4*>(p)[i] = P;
    reinterpret_cast<float4*>(m)[i] = M;
    reinterpret_cast<float4*>(v)[i] = V;
    if (e) reinterpret_cast<float4*>(e)[i] = E;
  }
  for (int64_t i = 4 * n4 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    adamw_one(p[i], g[i], m[i], v[i], e ? e + i : nullptr, b, wd);
}

}  // namespace uva
