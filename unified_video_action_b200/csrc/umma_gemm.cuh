// tcgen05 / TMEM GEMM core (sm_100a): D[M,N] = A[M,K] . B[N,K]^T with bf16 operands (both
// K-contiguous), fp32 accumulation in tensor memory and a fused epilogue.
//
//   * persistent: grid = min(#tiles, #SMs); tile = 128 rows x BN columns; static round-robin
//     schedule, rasterised so that concurrently running CTAs share B (weight) tiles and a small
//     window of A row-panels (L2 reuse);
//   * warp 0  : TMA producer — 128x64 (A) and BNx64 (B) bf16 boxes, 128-byte swizzle, into a
//                kStages-deep shared-memory ring (full/empty mbarriers);
//   * warp 1  : MMA issuer — one elected thread issues 4 x tcgen05.mma (M=128, N=BN, K=16) per
//                64-wide k-block, tcgen05.commit releases the smem slot / publishes the accumulator;
//   * warp 2  : TMEM allocator (2 accumulator stages of BN columns: epilogue of tile i overlaps
//                the MMAs of tile i+1);
//   * warps 4-7: epilogue — tcgen05.ld (lane = row), transposed through a per-warp padded smem
//                patch so that global accesses are row-contiguous (coalesced), fused epilogue math.
#pragma once
#include "common.cuh"
#include "epilogue.cuh"

namespace uva {

constexpr int UG_BM = 128;
constexpr int UG_BK = 64;  // 64 bf16 = 128 bytes = one swizzle row
constexpr int UG_THREADS = 256;

template <int BN>
struct UmmaCfg {
  static constexpr int kStageBytesA = UG_BM * UG_BK * 2;
  static constexpr int kStageBytesB = BN * UG_BK * 2;
  static constexpr int kStageBytes = kStageBytesA + kStageBytesB;
  static constexpr int kStages = (BN >= 256) ? 4 : (BN >= 128 ? 6 : 8);
  static constexpr int kPatchBytes = 4 * 32 * 33 * 4;  // 4 epilogue warps x [32][33] words
  static constexpr int kBarBytes = 256;
  static constexpr int kSmemBytes = 1024 /*align slack*/ + kStages * kStageBytes + kPatchBytes + kBarBytes;
  static constexpr uint32_t kTmemCols = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128 : (2 * BN <= 256) ? 256 : 512;
};

// tile id -> (m_blk, n_blk): groups of `gm` row panels; inside a group the row panel varies
// fastest, so ~gm CTAs share one weight tile and the group's A panels stay L2-resident.
__device__ __forceinline__ void ug_tile_coords(int64_t tile, int64_t m_tiles, int n_tiles, int gm,
                                               int64_t& m_blk, int& n_blk) {
  const int64_t per_group = (int64_t)gm * n_tiles;
  const int64_t g = tile / per_group;
  const int64_t rem = tile - g * per_group;
  const int64_t m_first = g * gm;
  int64_t rows_in_group = m_tiles - m_first;
  if (rows_in_group > gm) rows_in_group = gm;
  n_blk = (int)(rem / rows_in_group);
  m_blk = m_first + (rem - (int64_t)n_blk * rows_in_group);
}

template <int BN, int KIND, typename TA>
__global__ void __launch_bounds__(UG_THREADS, 1)
umma_gemm_tn_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    int64_t M, int N, int K, int gm, EpiArgs ep) {
  using Cfg = UmmaCfg<BN>;
  constexpr int kStages = Cfg::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stage_base = smem;
  uint32_t* patch = reinterpret_cast<uint32_t*>(smem + kStages * Cfg::kStageBytes);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes + Cfg::kPatchBytes);
  uint64_t* full_bar = bars;                  // [kStages]
  uint64_t* empty_bar = bars + kStages;       // [kStages]
  uint64_t* tfull_bar = bars + 2 * kStages;   // [2]
  uint64_t* tempty_bar = bars + 2 * kStages + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int64_t m_tiles = (M + UG_BM - 1) / UG_BM;
  const int n_tiles = N / BN;
  const int64_t num_tiles = m_tiles * n_tiles;
  const int num_kb = K / UG_BK;

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

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int64_t m_blk; int n_blk;
        ug_tile_coords(tile, m_tiles, n_tiles, gm, m_blk, n_blk);
        const int32_t m0 = (int32_t)(m_blk * UG_BM);
        const int32_t n0 = n_blk * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint8_t* sa = stage_base + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kStageBytesA;
          mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)Cfg::kStageBytes);
          tma_load_2d(sa, &tmA, &full_bar[stage], kb * UG_BK, m0);
          tma_load_2d(sb, &tmB, &full_bar[stage], kb * UG_BK, n0);
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = umma_idesc_bf16(UG_BM, BN);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(stage_base + stage * Cfg::kStageBytes);
          const uint32_t sb = sa + Cfg::kStageBytesA;
          const uint64_t da = umma_desc_k_sw128(sa);
          const uint64_t db = umma_desc_k_sw128(sb);
#pragma unroll
          for (int k = 0; k < UG_BK / 16; ++k) {
            // advance 16 bf16 = 32 bytes inside the 128-byte swizzle row: +2 in 16-byte units
            umma_bf16(d_tmem, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc,
                      (kb > 0 || k > 0) ? 1u : 0u);
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
    // ===================== epilogue =====================
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    uint32_t* my_patch = patch + q * (32 * 33);
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int64_t tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int64_t m_blk; int n_blk;
      ug_tile_coords(tile, m_tiles, n_tiles, gm, m_blk, n_blk);
      const int64_t row0 = m_blk * UG_BM + q * 32;
      const int n0 = n_blk * BN;
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
      constexpr bool kPairs = (sizeof(TA) == 2) && (KIND == EPI_STORE_ACT || KIND == EPI_SILU_ACT ||
                                                    KIND == EPI_SILU_SAVE || KIND == EPI_DSILU_ACT);
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32x32(t_addr + (uint32_t)c0, v);
        tmem_ld_wait();
        // thread = row `lane`; write my 32 columns into patch[lane][0..31] (row stride 33 words)
#pragma unroll
        for (int j = 0; j < 32; ++j) my_patch[lane * 33 + j] = v[j];
        __syncwarp();
        if constexpr (kPairs) {
          // lane handles columns (2*(lane&15), +1) of rows rr + (lane>>4): 2 rows per pass
#pragma unroll 4
          for (int rr = 0; rr < 32; rr += 2) {
            const int r = rr + (lane >> 4);
            const int cc = 2 * (lane & 15);
            const float a0 = __uint_as_float(my_patch[r * 33 + cc]);
            const float a1 = __uint_as_float(my_patch[r * 33 + cc + 1]);
            const int64_t grow = row0 + r;
            if (grow < M) Epi<KIND, TA>::apply2(ep, grow, n0 + c0 + cc, a0, a1);
          }
        } else {
#pragma unroll 4
          for (int r = 0; r < 32; ++r) {
            const float a0 = __uint_as_float(my_patch[r * 33 + lane]);
            const int64_t grow = row0 + r;
            if (grow < M) Epi<KIND, TA>::apply(ep, grow, n0 + c0 + lane, a0);
          }
        }
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
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
