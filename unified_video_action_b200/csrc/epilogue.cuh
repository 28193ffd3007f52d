// GEMM epilogues shared by the fp32 FFMA GEMM (sgemm.cuh) and the tcgen05 GEMM (umma_gemm.cuh).
// An epilogue receives accumulator values acc[row, col] (fp32, bias NOT yet added) and performs
// the fused tail of the layer.  `apply` handles one element, `apply2` two adjacent columns
// (col even) so 16-bit outputs can be written as packed pairs.
#pragma once
#include "common.cuh"

namespace uva {

struct EpiArgs {
  const float* bias = nullptr;  // [N] fp32 or null
  void* out = nullptr;          // primary output
  int64_t ldo = 0;              // row stride of out (elements)
  const void* aux = nullptr;    // second input (gate / saved activation), element type = TA
  int64_t ldaux = 0;
  void* out2 = nullptr;         // optional second output (pre-activation / un-gated MLP output)
  int64_t ldo2 = 0;
  const float* res = nullptr;   // residual input of the gated-residual kinds (row stride ldo); null = out
  int32_t n_valid = 0;          // valid output columns (columns >= n_valid are padding)
  float alpha = 1.0f;
  // LayerNorm coupling (bf16 sampling path): the gated-residual kinds add the row sums (sum, sum of
  // squares) of their output to stats[row][2]; EPI_LNMOD reads them back (see umma_gemm2.cuh)
  float* stats = nullptr;
  const float* ln_w = nullptr;  // [W] LayerNorm affine (null: none)
  const float* ln_b = nullptr;
  int32_t ln_tiles = 0;         // EPI_LNMOD: number of leading 256-column tiles that are (shift | scale) pairs
  int32_t ln_width = 0;         // EPI_LNMOD: W
  // EPI_DSILU_ACT on the CTA-pair kernel: column sums of the fp32 result (the bias gradient of the Linear in front of
  // the SiLU) are reduced over each warp's 32 rows with shuffles and added here with one atomic per column; null: off
  float* colsum = nullptr;
  // deterministic split-K (single-CTA kernel, EPI_STORE_F32): split s of a tile stores its partial sums to row
  // (row + s * split_rows) of `out`; the caller reduces the split_k partial planes in a fixed order afterwards
  int32_t split_k = 0;
  int64_t split_rows = 0;
  int32_t conv_cpb = 0;  // implicit-GEMM convolution operands: 64-channel blocks per tap (C_in / 64)
  // LayerNorm-modulate pre-phase of the latency-path GEMM1 (umma_gemm_tn_kernel<.., EPI_SILU_ACT>, one wave, K = 1024):
  // before the main loop the CTAs of a 128-row panel produce the A operand a = LN(h) (1 + scale) + shift themselves
  // (a few rows each), meet at a panel barrier (pre_bar: {count, generation} per panel) and only then start their TMA
  // loads of A.  Replaces one k_lnmod launch per block and denoise step.  pre_h == null: off.
  const float* pre_h = nullptr;      // [M, K] fp32 residual stream
  const void* pre_shift = nullptr;   // TA, row stride pre_ldm
  const void* pre_scale = nullptr;
  int64_t pre_ldm = 0;
  const float* pre_lnw = nullptr;    // LayerNorm affine or null
  const float* pre_lnb = nullptr;
  void* pre_a = nullptr;             // [M, K] TA: the tensor the A map points at
  unsigned int* pre_bar = nullptr;   // [m_tiles][2], zero-initialised once, self-resetting
};

enum EpiKind {
  EPI_STORE_F32 = 0,  // out(f32)[r,c]  = acc + bias
  EPI_STORE_ACT = 1,  // out(TA)[r,c]   = acc + bias
  EPI_SILU_ACT = 2,   // out(TA)[r,c]   = SiLU(acc + bias)                       (DL:152-156)
  EPI_GATED_RES = 3,  // out(f32)[r,c] += aux(TA)[r,c] * (acc + bias)            (DL:167)
  EPI_SILU_SAVE = 4,  // out(TA) = SiLU(p), out2(TA) = p = acc + bias            (training fwd)
  EPI_ACCUM_F32 = 5,  // out(f32)[r,c] += alpha * acc                            (wgrad accumulate)
  EPI_DSILU_ACT = 6,  // out(TA)[r,c]   = (acc) * SiLU'(aux(TA)[r,c])            (dgrad through SiLU)
  EPI_ATOMIC_F32 = 7, // atomicAdd(out(f32)[r,c], alpha * acc)                   (split-K wgrad)
  EPI_GATED_SAVE = 8, // q = acc + bias; out2(TA) = q; out(f32) = res + aux * q  (training fwd, DL:167)
  EPI_RELU_ACT = 10,  // out(TA)[r,c]   = max(acc + bias, 0)                      (pre-net Linear + ReLU, DAL:47-55)
  EPI_CONVPOOL = 11,  // CTA-pair kernel, implicit 3x3 convolution on 16 x 16 maps: y = max(acc + bias, 0); out2(TA) = y
                      // (optional, for the backward), out(f32)[image][4x4 region][c] += y / 16   (Conv2d + ReLU +
                      // AdaptiveAvgPool2d(4,4) of the DiffActLoss pre-net, DAL:41-46)
  EPI_LNMOD = 9,      // AdaLN GEMM with the LayerNorm-modulate fused (CTA-pair kernel only): tiles < ln_tiles
                      // hold 128 shift | 128 scale columns -> out(TA) = LN(res)[*ln_w+ln_b]*(1+scale)+shift;
                      // the other tiles (gate) are stored to out2(TA)                         (DL:93-94, 163-166)
};

template <int KIND, typename TA>
struct Epi {
  __device__ static __forceinline__ void apply(const EpiArgs& e, int64_t r, int c, float acc) {
    if (c >= e.n_valid) return;
    const float b = (e.bias != nullptr) ? __ldg(e.bias + c) : 0.0f;
    if constexpr (KIND == EPI_STORE_F32) {
      reinterpret_cast<float*>(e.out)[r * e.ldo + c] = acc + b;
    } else if constexpr (KIND == EPI_STORE_ACT) {
      reinterpret_cast<TA*>(e.out)[r * e.ldo + c] = from_f32<TA>(acc + b);
    } else if constexpr (KIND == EPI_SILU_ACT) {
      reinterpret_cast<TA*>(e.out)[r * e.ldo + c] = from_f32<TA>(silu_f(acc + b));
    } else if constexpr (KIND == EPI_RELU_ACT) {
      reinterpret_cast<TA*>(e.out)[r * e.ldo + c] = from_f32<TA>(fmaxf(acc + b, 0.f));
    } else if constexpr (KIND == EPI_GATED_RES || KIND == EPI_GATED_SAVE) {
      const float g = to_f32(reinterpret_cast<const TA*>(e.aux)[r * e.ldaux + c]);
      float* o = reinterpret_cast<float*>(e.out) + r * e.ldo + c;
      const float* rp = (e.res != nullptr ? e.res : reinterpret_cast<const float*>(e.out)) + r * e.ldo + c;
      if constexpr (KIND == EPI_GATED_SAVE) reinterpret_cast<TA*>(e.out2)[r * e.ldo2 + c] = from_f32<TA>(acc + b);
      *o = *rp + g * (acc + b);
    } else if constexpr (KIND == EPI_SILU_SAVE) {
      const float p = acc + b;
      reinterpret_cast<TA*>(e.out2)[r * e.ldo2 + c] = from_f32<TA>(p);
      reinterpret_cast<TA*>(e.out)[r * e.ldo + c] = from_f32<TA>(silu_f(p));
    } else if constexpr (KIND == EPI_ACCUM_F32) {
      float* o = reinterpret_cast<float*>(e.out) + r * e.ldo + c;
      *o = *o + e.alpha * acc;
    } else if constexpr (KIND == EPI_DSILU_ACT) {
      const float p = to_f32(reinterpret_cast<const TA*>(e.aux)[r * e.ldaux + c]);
      reinterpret_cast<TA*>(e.out)[r * e.ldo + c] = from_f32<TA>(acc * dsilu_f(p));
    } else if constexpr (KIND == EPI_ATOMIC_F32) {
      atomicAdd(reinterpret_cast<float*>(e.out) + r * e.ldo + c, e.alpha * acc);
    }
  }

  // Two adjacent columns (c even, row strides even): packed stores for 16-bit outputs.
  __device__ static __forceinline__ void apply2(const EpiArgs& e, int64_t r, int c, float a0,
                                                float a1) {
    if constexpr (sizeof(TA) == 2 &&
                  (KIND == EPI_STORE_ACT || KIND == EPI_SILU_ACT || KIND == EPI_SILU_SAVE ||
                   KIND == EPI_DSILU_ACT)) {
      if (c + 1 >= e.n_valid) {
        apply(e, r, c, a0);
        apply(e, r, c + 1, a1);
        return;
      }
      float b0 = 0.f, b1 = 0.f;
      if (e.bias != nullptr) {
        const float2 bb = __ldg(reinterpret_cast<const float2*>(e.bias + c));
        b0 = bb.x;
        b1 = bb.y;
      }
      float v0 = a0 + b0, v1 = a1 + b1;
      if constexpr (KIND == EPI_SILU_SAVE) {
        *reinterpret_cast<uint32_t*>(reinterpret_cast<bf16*>(e.out2) + r * e.ldo2 + c) =
            pack_bf16x2(v0, v1);
      }
      if constexpr (KIND == EPI_SILU_ACT || KIND == EPI_SILU_SAVE) {
        v0 = silu_f(v0);
        v1 = silu_f(v1);
      }
      if constexpr (KIND == EPI_DSILU_ACT) {
        const __nv_bfloat162 p =
            *reinterpret_cast<const __nv_bfloat162*>(reinterpret_cast<const bf16*>(e.aux) + r * e.ldaux + c);
        v0 = a0 * dsilu_f(__low2float(p));
        v1 = a1 * dsilu_f(__high2float(p));
      }
      *reinterpret_cast<uint32_t*>(reinterpret_cast<bf16*>(e.out) + r * e.ldo + c) =
          pack_bf16x2(v0, v1);
    } else {
      apply(e, r, c, a0);
      apply(e, r, c + 1, a1);
    }
  }
};

}  // namespace uva
