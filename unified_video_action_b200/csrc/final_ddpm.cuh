// FinalLayer linear (DL:188) with the DDPM ancestral step folded into its epilogue (north_star: "fold the DDPM
// posterior update, the CFG mix and the temperature scaling into the final-layer epilogue").
//
//   net[r, 0:2C] = a[r, :] . Wlin^T + blin           (a = LayerNorm-modulated h, bf16; Wlin padded to 32 rows)
//   eps = CFG mix of the cond / uncond rows (DL:288-292), v = net[r, C:2C]
//   x_next[r, :] = ddpm_update(eps, v, x_cur[r, :], noise)      (GD:260-346, 395-444; elementwise.cuh)
//
// The net output never reaches memory.  With CFG a chunk's rows are cond (first half) || uncond (second half) and
// row r needs the eps of its partner r +- half, which lives in another 128-row tile.  One work item therefore
// computes BOTH tiles of a pair: the TMA producer loads the cond box [128 rows, 64 k] and the uncond box at row
// offset +half (any offset: TMA coordinates need no alignment), the MMA thread issues both into adjacent TMEM
// column ranges (cond: 32 columns, uncond: 32 columns), and the epilogue thread of lane row i holds the four
// vectors (cond eps, cond v, uncond eps, uncond v) of the pair, mixes, and writes x_next for both rows.
//
// Same warp roles and pipeline as umma_gemm_tn_kernel (umma_gemm.cuh): warp 0 TMA, warp 1 tcgen05.mma issue,
// warp 2 TMEM allocation, warps 4-7 epilogue (one per TMEM lane quarter); two accumulator stages.
#pragma once
#include "common.cuh"
#include "elementwise.cuh"
#include "umma_gemm.cuh"

namespace uva {

constexpr int FD_THREADS = 256;
constexpr int FD_BN = 32;  // padded output columns (2C <= 32)

struct FinalDdpmArgs {
  const float* bias;   // [2C]
  const float* x_cur;  // [rows, C] chunk-local state
  float* x_next;       // [rows, C]
  const float* noise;  // caller's [N, C] slab of this step, or null (Philox)
  PhiloxSeed ps;
  uint32_t step;
  int64_t rows, half, g0, ghalf;
  int C, K, clip;
  float cfg, temperature;
  StepCoef kc;
};

template <bool CFG>
struct FinalDdpmCfg {
  static constexpr int kStageBytesA = UG_BM * UG_BK * 2;
  static constexpr int kStageBytesB = FD_BN * UG_BK * 2;
  static constexpr int kStageBytes = (CFG ? 2 : 1) * kStageBytesA + kStageBytesB;
  static constexpr int kStages = 5;
  static constexpr int kAccCols = (CFG ? 2 : 1) * FD_BN;       // TMEM columns of one accumulator stage
  static constexpr uint32_t kTmemCols = 2 * kAccCols < 32 ? 32 : 2 * kAccCols;  // 64 or 128
  static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + 256;
};

template <bool CFG>
__global__ void __launch_bounds__(FD_THREADS, 1)
umma_final_ddpm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                       const __grid_constant__ FinalDdpmArgs p) {
  using Cfg = FinalDdpmCfg<CFG>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kStages;
  uint64_t* tfull_bar = bars + 2 * kStages;
  uint64_t* tempty_bar = bars + 2 * kStages + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t span = CFG ? p.half : p.rows;  // rows that own a work item
  const int64_t items = (span + UG_BM - 1) / UG_BM;
  const int num_kb = (p.K + UG_BK - 1) / UG_BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 4);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_wait();
  griddep_launch();

  if (warp == 0) {
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int64_t item = blockIdx.x; item < items; item += gridDim.x) {
        const int32_t m0 = (int32_t)(item * UG_BM);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint8_t* sa = smem + stage * Cfg::kStageBytes;
          mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)Cfg::kStageBytes);
          tma_load_2d(sa, &tmA, &full_bar[stage], kb * UG_BK, m0);
          if constexpr (CFG) tma_load_2d(sa + Cfg::kStageBytesA, &tmA, &full_bar[stage], kb * UG_BK, m0 + (int32_t)p.half);
          tma_load_2d(sa + (CFG ? 2 : 1) * Cfg::kStageBytesA, &tmB, &full_bar[stage], kb * UG_BK, 0);
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = umma_idesc_bf16(UG_BM, FD_BN, false, false);
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    for (int64_t item = blockIdx.x; item < items; item += gridDim.x) {
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * Cfg::kAccCols);
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
          const uint64_t da0 = umma_desc_k_sw128(sa);
          const uint64_t da1 = umma_desc_k_sw128(sa + Cfg::kStageBytesA);
          const uint64_t db = umma_desc_k_sw128(sa + (CFG ? 2 : 1) * Cfg::kStageBytesA);
#pragma unroll
          for (int k = 0; k < UG_BK / 16; ++k) {
            const uint32_t accum = (kb > 0 || k > 0) ? 1u : 0u;
            umma_bf16(d_tmem, da0 + 2ull * (uint64_t)k, db + 2ull * (uint64_t)k, idesc, accum);
            if constexpr (CFG) umma_bf16(d_tmem + FD_BN, da1 + 2ull * (uint64_t)k, db + 2ull * (uint64_t)k, idesc, accum);
          }
          umma_commit(&empty_bar[stage]);
          if (kb == num_kb - 1) umma_commit(&tfull_bar[acc]);
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    const int C = p.C;
    int acc = 0;
    uint32_t acc_phase = 0;
    float b[FD_BN];
#pragma unroll
    for (int i = 0; i < FD_BN; ++i) b[i] = (i < 2 * C) ? __ldg(p.bias + i) : 0.f;
    const bool vec_ok = (((uintptr_t)p.x_cur | (uintptr_t)p.x_next | (uintptr_t)(p.ps.on ? nullptr : p.noise)) & 15u) == 0;
    for (int64_t item = blockIdx.x; item < items; item += gridDim.x) {
      const int64_t r = item * UG_BM + q * 32 + lane;  // cond row (CFG) / row
      const bool ok = r < span;
      // The row's state (and its noise draws, when they come from memory) are requested BEFORE the accumulator is waited
      // for - as 16-byte loads when C = 16 and the pointers allow it, as independent 4-byte loads otherwise - and the
      // result leaves as 16-byte (or independent 4-byte) stores.  The first form of this epilogue walked the channels
      // one dependent global access after the other (load x, load noise, store): ~30 us per 128-row item at C = 16 for an
      // item whose operands arrive in ~6 us (ncu: dram 29 %, sm 8 %)
      const bool vec = (C == 16) && vec_ok;
      const bool mem_noise = p.kc.nonzero && !p.ps.on;
      float x0[16], x1[16], n0[16], n1[16];
      if (ok) {
        if (vec) {
          const float4* px0 = reinterpret_cast<const float4*>(p.x_cur + r * 16);
#pragma unroll
          for (int i = 0; i < 4; ++i) { const float4 t = px0[i]; x0[4 * i] = t.x; x0[4 * i + 1] = t.y; x0[4 * i + 2] = t.z; x0[4 * i + 3] = t.w; }
          if constexpr (CFG) {
            const float4* px1 = reinterpret_cast<const float4*>(p.x_cur + (r + p.half) * 16);
#pragma unroll
            for (int i = 0; i < 4; ++i) { const float4 t = px1[i]; x1[4 * i] = t.x; x1[4 * i + 1] = t.y; x1[4 * i + 2] = t.z; x1[4 * i + 3] = t.w; }
          }
          if (mem_noise) {
            const float4* pn0 = reinterpret_cast<const float4*>(p.noise + (p.g0 + r) * 16);
#pragma unroll
            for (int i = 0; i < 4; ++i) { const float4 t = pn0[i]; n0[4 * i] = t.x; n0[4 * i + 1] = t.y; n0[4 * i + 2] = t.z; n0[4 * i + 3] = t.w; }
            if constexpr (CFG) {
              const float4* pn1 = reinterpret_cast<const float4*>(p.noise + (p.g0 + p.ghalf + r) * 16);
#pragma unroll
              for (int i = 0; i < 4; ++i) { const float4 t = pn1[i]; n1[4 * i] = t.x; n1[4 * i + 1] = t.y; n1[4 * i + 2] = t.z; n1[4 * i + 3] = t.w; }
            }
          }
        } else {
#pragma unroll
          for (int c = 0; c < 16; ++c) {
            if (c < C) {
              x0[c] = p.x_cur[r * C + c];
              if constexpr (CFG) x1[c] = p.x_cur[(r + p.half) * C + c];
              if (mem_noise) {
                n0[c] = p.noise[(p.g0 + r) * C + c];
                if constexpr (CFG) n1[c] = p.noise[(p.g0 + p.ghalf + r) * C + c];
              }
            }
          }
        }
      }
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * Cfg::kAccCols);
      uint32_t v0[32], v1[32];
      tmem_ld_32x32(t_addr, v0);
      if constexpr (CFG) tmem_ld_32x32(t_addr + FD_BN, v1);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);  // the accumulator is in registers now
      if (ok) {
        float o0[16], o1[16];
#pragma unroll
        for (int c = 0; c < FD_BN / 2; ++c) {
          if (c < C) {
            // static register indexing: c is a compile-time constant after unrolling, C only gates
            float ce = 0.f, cv = 0.f, ue = 0.f, uv = 0.f;
#pragma unroll
            for (int j = 0; j < FD_BN; ++j) {
              if (j == c) { ce = __uint_as_float(v0[j]) + b[j]; if constexpr (CFG) ue = __uint_as_float(v1[j]) + b[j]; }
              if (j == C + c) { cv = __uint_as_float(v0[j]) + b[j]; if constexpr (CFG) uv = __uint_as_float(v1[j]) + b[j]; }
            }
            float z0 = 0.f, z1 = 0.f;
            if (p.kc.nonzero) {
              z0 = p.ps.on ? philox_normal(p.ps, p.step, (uint64_t)((p.g0 + r) * C + c)) : n0[c];
              if constexpr (CFG) z1 = p.ps.on ? philox_normal(p.ps, p.step, (uint64_t)((p.g0 + p.ghalf + r) * C + c)) : n1[c];
            }
            if constexpr (CFG) {
              const float eps = ue + p.cfg * (ce - ue);
              o0[c] = ddpm_update_nz(eps, cv, x0[c], z0, p.temperature, p.clip, p.kc);
              o1[c] = ddpm_update_nz(eps, uv, x1[c], z1, p.temperature, p.clip, p.kc);
            } else {
              o0[c] = ddpm_update_nz(ce, cv, x0[c], z0, p.temperature, p.clip, p.kc);
            }
          }
        }
        if (vec) {
          float4* po0 = reinterpret_cast<float4*>(p.x_next + r * 16);
#pragma unroll
          for (int i = 0; i < 4; ++i) po0[i] = make_float4(o0[4 * i], o0[4 * i + 1], o0[4 * i + 2], o0[4 * i + 3]);
          if constexpr (CFG) {
            float4* po1 = reinterpret_cast<float4*>(p.x_next + (r + p.half) * 16);
#pragma unroll
            for (int i = 0; i < 4; ++i) po1[i] = make_float4(o1[4 * i], o1[4 * i + 1], o1[4 * i + 2], o1[4 * i + 3]);
          }
        } else {
#pragma unroll
          for (int c = 0; c < 16; ++c) {
            if (c < C) {
              p.x_next[r * C + c] = o0[c];
              if constexpr (CFG) p.x_next[(r + p.half) * C + c] = o1[c];
            }
          }
        }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<Cfg::kTmemCols>(tmem_base);
  }
}

}  // namespace uva
