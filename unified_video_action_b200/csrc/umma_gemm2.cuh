// CTA-pair tcgen05 GEMM (sm_100a, cta_group::2): D[M,N] = sum_k A(m,k) B(n,k), bf16 operands, fp32
// accumulation in tensor memory, fused row-segment epilogue (RowEpi of umma_gemm.cuh).
//
// A cluster of two CTAs (one TPC) owns a 256 x 256 output tile: CTA r loads the A rows
// [256 m_blk + 128 r, +128) and the B rows [256 n_blk + 128 r, +128) of every k-block, the leader CTA
// issues ONE tcgen05.mma.cta_group::2 (M = 256, N = 256, K = 16) per 16 k that reads A from each
// CTA's own shared memory and the two B halves from both, and each CTA's tensor memory receives
// the accumulators of its own 128 rows.  Per SM and per 128 tensor-pipe cycles that is 8 KB of shared
// memory operand reads instead of the 12 KB of the single-CTA 128 x 256 tile, and half the L2 -> SMEM
// traffic for B.
//
//   warp 0 : TMA producer (both CTAs; completion bytes of both land on the LEADER's full barrier)
//   warp 1 : MMA issuer (leader CTA only); tcgen05.commit multicasts to the barriers of both CTAs
//   warp 2 : tensor-memory allocator (cta_group::2 allocation, both CTAs); afterwards the store thread of the TMA-fed epilogues
//   warps 4-11: epilogue of this CTA's 128 rows; accumulator stage release = remote arrive on the
//           leader's barrier (16 arrivals: 8 warps x 2 CTAs)
#pragma once
#include "pair_epilogue.cuh"
#include "umma_gemm.cuh"

namespace uva {

// TMA-fed epilogue (EPI_GATED_RES): the residual tile h (fp32) and the gate tile (bf16) are prefetched by
// warp 3 into a ring of [128 rows x 32 columns] boxes while the tile's MMAs run; the epilogue warps
// update the h box in place and one thread per warp group writes it back with a TMA store.
// EPI_LNMOD uses the same ring: h boxes are prefetched for the (shift | scale) tiles, and every output
// box (a, gate: bf16) is staged in the 8 KB part of a slot and written with a TMA store.
struct EpiMaps {
  CUtensorMap h_in;   // fp32 [M, W], box 32 x 128, 128-byte swizzle (residual stream input)
  CUtensorMap h_out;  // GATED_RES: fp32 output, same geometry as h_in.  LNMOD: bf16 gate output [M, W], 64-byte swizzle
  CUtensorMap g;      // GATED_RES: bf16 gate input [M, N], box 32 x 128, 64-byte swizzle.  LNMOD: bf16 `a` output [M, W]
  CUtensorMap b2;     // GATED_SAVE: bf16 output of the un-gated MLP result q
  // training kinds: DSILU_ACT g = aux input, h_out = bf16 output; SILU_SAVE h_out = out, g = out2 (both bf16);
  // GATED_SAVE h_in = res, h_out = out (fp32), g = gate input, b2 = out2
};

template <int BN, int KIND = 0>
struct Umma2Cfg {
  static constexpr int kBNH = BN / 2;  // B rows per CTA
  static constexpr int kStageBytesA = UG_BM * UG_BK * 2;
  static constexpr int kStageBytesB = kBNH * UG_BK * 2;
  static constexpr int kStageBytes = kStageBytesA + kStageBytesB;
  static constexpr bool kTmaEpi = (KIND == EPI_GATED_RES || KIND == EPI_LNMOD || KIND == EPI_SILU_ACT || KIND == EPI_STORE_ACT ||
                                   KIND == EPI_DSILU_ACT || KIND == EPI_SILU_SAVE || KIND == EPI_GATED_SAVE);
  static constexpr bool kTmaOutOnly = (KIND == EPI_SILU_ACT || KIND == EPI_STORE_ACT || KIND == EPI_SILU_SAVE);  // ring = output staging only
  static constexpr int kStages = kTmaEpi ? 4 : 5;
  static constexpr int kRing = 4;                         // epilogue prefetch ring (kTmaEpi)
  static constexpr int kSlotH = UG_BM * 32 * 4;           // 16 KB
  static constexpr int kSlotG = UG_BM * 32 * 2;           //  8 KB
  static constexpr int kSlotBytes = kSlotH + kSlotG;
  static constexpr int kEpiBytes = kTmaEpi ? kRing * kSlotBytes : 8 * PE_WARP_BYTES;
  static constexpr int kBarBytes = 256;
  static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kEpiBytes + kBarBytes;
  static constexpr uint32_t kTmemCols = 512;  // 2 accumulator stages x 256 columns
};

// CL = 2: one pair per cluster.  CL = 4: two pairs per cluster work on the SAME 256 rows and on two
// adjacent 256-column tiles; every A box is loaded once and multicast to both pairs (each CTA fetches
// 64 of the 128 rows its half needs), so a CTA pulls 24 KB instead of 32 KB per k-block through the
// L2 -> SMEM path, which is what bounds the CL = 2 kernel (profiles/r01_ncu_full_gemm2_v3_summary.txt).
// The cluster size is a launch attribute (cudaLaunchAttributeClusterDimension).
// CONV (the DiffActLoss pre-net's Conv2d(C, C, 3, padding 1) on 16 x 16 maps as an implicit GEMM, NHWC bf16 input,
// CL = 2 only):  1: the A operand is gathered - row m = pixel (image m / 256, w = (m % 256) / 16, h = m % 16),
// k-block kb = (tap kb / cpb, channels 64 (kb % cpb) ..): one 4D TMA box {64 c, 16 h, 8 w, 1 image} at
// (c0, kw - 1, w0 + kh - 1, image) per CTA, the padding is TMA's out-of-range zero fill (forward and, with the taps
// flipped in the packed weight, data gradient).  2: the B operand (MN-major, weight-gradient form) is gathered the
// same way: k = pixel, n = (tap, c_in), boxes {64 c, 16 h, 4 w, 1 image}.
template <int BN, int KIND, typename TA, bool A_MN = false, bool B_MN = false, int CL = 2, int CONV = 0>
__global__ void __launch_bounds__(UG_THREADS, 1)
umma_gemm2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ EpiMaps em, int64_t M, int N, int K, int gm, int splits, EpiArgs ep) {
  static_assert(BN == 256, "CTA-pair kernel: 256 x 256 tiles");
  using Cfg = Umma2Cfg<BN, KIND>;
  constexpr int kStages = Cfg::kStages;
  constexpr int BM2 = 2 * UG_BM;
  constexpr uint32_t kAtomBytes = 64 * UG_BK * 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stage_base = smem;
  uint8_t* epi_smem = smem + kStages * Cfg::kStageBytes;  // per-warp staging tiles or the TMA ring (1024-aligned)
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_smem + Cfg::kEpiBytes);
  uint64_t* full_bar = bars;                      // [kStages] (used in the leader CTA)
  uint64_t* empty_bar = bars + kStages;           // [kStages] (each CTA its own, multicast commit)
  uint64_t* tfull_bar = bars + 2 * kStages;       // [2]       (each CTA its own, multicast commit)
  uint64_t* tempty_bar = bars + 2 * kStages + 2;  // [2]       (used in the leader CTA)
  uint64_t* efull_bar = bars + 2 * kStages + 4;                  // [kRing] epilogue prefetch ring (this CTA)
  uint64_t* eempty_bar = bars + 2 * kStages + 4 + Cfg::kRing;    // [kRing]
  uint64_t* sready_bar = bars + 2 * kStages + 4 + 2 * Cfg::kRing;  // [kRing] box staged by the epilogue warps, ready for its TMA store
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4 + 3 * Cfg::kRing);
  static_assert((2 * kStages + 4 + 3 * Cfg::kRing) * 8 + 4 <= Cfg::kBarBytes, "barrier area");

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  static_assert(CL == 2 || CL == 4, "cluster of one or two CTA pairs");
  const uint32_t crank = cluster_ctarank();
  const uint32_t rank = crank & 1u;   // position inside the pair
  const uint32_t pair = crank >> 1;   // pair inside the cluster (0 when CL == 2)
  const bool leader = rank == 0;
  const int64_t cluster_id = blockIdx.x / CL;
  const int64_t num_clusters = gridDim.x / CL;
  const int64_t m_tiles = (M + BM2 - 1) / BM2;
  constexpr int kPairs = CL / 2;
  const int n_tiles = N / BN / kPairs;  // column tiles (CL == 2) or pairs of adjacent column tiles (CL == 4)
  const int64_t num_items = m_tiles * n_tiles * splits;
  const uint16_t pair_mask = (uint16_t)(3u << (2 * pair));
  constexpr uint16_t kAllMask = (uint16_t)((1u << CL) - 1u);
  const int num_kb_total = (K + UG_BK - 1) / UG_BK;
  const int kb_per = (num_kb_total + splits - 1) / splits;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if constexpr (Cfg::kTmaEpi) {
      tma_prefetch_desc(&em.h_in);
      tma_prefetch_desc(&em.h_out);
      tma_prefetch_desc(&em.g);
      if constexpr (KIND == EPI_GATED_SAVE) tma_prefetch_desc(&em.b2);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], kPairs);  // every pair that receives this CTA's boxes releases the slot
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 16);
    }
    for (int s = 0; s < Cfg::kRing; ++s) {
      mbar_init(&efull_bar[s], 1);
      mbar_init(&eempty_bar[s], 1);
      mbar_init(&sready_bar[s], 1);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_2sm<Cfg::kTmemCols>(tmem_slot);
  __syncwarp();
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_wait();    // programmatic dependent launch (when the launch asks for it): the prologue above overlapped the
  griddep_launch();  // previous kernel's tail; no global memory is touched before this point

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int64_t item = cluster_id; item < num_items; item += num_clusters) {
        int64_t m_blk; int n_blk;
        const int64_t tile = item / splits;
        const int split = (int)(item - tile * splits);
        ug_tile_coords(tile, m_tiles, n_tiles, gm, m_blk, n_blk);
        const int32_t m0 = (int32_t)(m_blk * BM2) + (int32_t)rank * UG_BM;
        const int32_t n0 = (n_blk * kPairs + (int32_t)pair) * BN + (int32_t)rank * Cfg::kBNH;
        const int kb0 = split * kb_per;
        const int kb1 = (kb0 + kb_per < num_kb_total) ? kb0 + kb_per : num_kb_total;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint8_t* sa = stage_base + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kStageBytesA;
          if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2u * (uint32_t)Cfg::kStageBytes);
          if constexpr (CL == 4) {
            // this CTA fetches rows [64 pair, +64) of its half's 128 A rows for both pairs (box: 64 x 64)
            const uint16_t mc = (uint16_t)(5u << rank);
            if constexpr (A_MN) tma_load_2d_2sm_mc(sa + pair * kAtomBytes, &tmA, &full_bar[stage], m0 + 64 * (int32_t)pair, kb * UG_BK, mc);
            else tma_load_2d_2sm_mc(sa + pair * kAtomBytes, &tmA, &full_bar[stage], kb * UG_BK, m0 + 64 * (int32_t)pair, mc);
          } else if constexpr (CONV == 1) {
            static_assert(CONV != 1 || (!A_MN && CL == 2), "conv gather of A: K-major, one pair per cluster");
            const int tap = kb / ep.conv_cpb;
            const int32_t c0 = (kb - tap * ep.conv_cpb) * UG_BK;
            tma_load_4d_2sm(sa, &tmA, &full_bar[stage], c0, tap % 3 - 1, (int32_t)rank * 8 + tap / 3 - 1, (int32_t)m_blk);
          } else if constexpr (A_MN) {
#pragma unroll
            for (int a = 0; a < UG_BM / 64; ++a) tma_load_2d_2sm(sa + a * kAtomBytes, &tmA, &full_bar[stage], m0 + 64 * a, kb * UG_BK);
          } else {
            tma_load_2d_2sm(sa, &tmA, &full_bar[stage], kb * UG_BK, m0);
          }
          if constexpr (CONV == 2) {
            static_assert(CONV != 2 || (B_MN && CL == 2), "conv gather of B: MN-major, one pair per cluster");
            const int cin = ep.conv_cpb * UG_BK;
#pragma unroll
            for (int a = 0; a < Cfg::kBNH / 64; ++a) {
              const int nidx = n0 + 64 * a;  // = tap * C_in + c_in
              const int tap = nidx / cin;
              tma_load_4d_2sm(sb + a * kAtomBytes, &tmB, &full_bar[stage], nidx - tap * cin, tap % 3 - 1,
                              (kb & 3) * 4 + tap / 3 - 1, kb >> 2);
            }
          } else if constexpr (B_MN) {
#pragma unroll
            for (int a = 0; a < Cfg::kBNH / 64; ++a) tma_load_2d_2sm(sb + a * kAtomBytes, &tmB, &full_bar[stage], n0 + 64 * a, kb * UG_BK);
          } else {
            tma_load_2d_2sm(sb, &tmB, &full_bar[stage], kb * UG_BK, n0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1 && leader) {
    // ===================== MMA issuer (leader CTA) =====================
    constexpr uint32_t idesc = umma_idesc_bf16(BM2, BN, A_MN, B_MN);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int64_t item = cluster_id; item < num_items; item += num_clusters) {
      const int split = (int)(item % splits);
      const int kb0 = split * kb_per;
      const int kb1 = (kb0 + kb_per < num_kb_total) ? kb0 + kb_per : num_kb_total;
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_u32(stage_base + stage * Cfg::kStageBytes);
          const uint32_t sb = sa + Cfg::kStageBytesA;
          const uint64_t da = A_MN ? umma_desc_mn_sw128(sa, kAtomBytes) : umma_desc_k_sw128(sa);
          const uint64_t db = B_MN ? umma_desc_mn_sw128(sb, kAtomBytes) : umma_desc_k_sw128(sb);
          constexpr uint64_t ka = A_MN ? 128 : 2;
          constexpr uint64_t kbs = B_MN ? 128 : 2;
#pragma unroll
          for (int k = 0; k < UG_BK / 16; ++k)
            umma_bf16_2sm(d_tmem, da + ka * (uint64_t)k, db + kbs * (uint64_t)k, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          umma_commit_2sm(&empty_bar[stage], kAllMask);
          if (kb == kb1 - 1) umma_commit_2sm(&tfull_bar[acc], pair_mask);
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  } else if (warp == 2) {
    // ===================== store thread (TMA-fed epilogue kinds) =====================
    // The epilogue warps only stage 32-column boxes in the ring; this thread issues their TMA stores and frees a
    // slot when its store has read it, so no epilogue warp ever waits for a store (they used to: ~0.3 us per box).
    if constexpr (Cfg::kTmaEpi) {
      if (elect_one()) {
        uint32_t seq = 0;
        int prev_slot = -1;
        for (int64_t item = cluster_id; item < num_items; item += num_clusters) {
          int64_t m_blk; int n_blk;
          ug_tile_coords(item / splits, m_tiles, n_tiles, gm, m_blk, n_blk);
          n_blk = n_blk * kPairs + (int)pair;
          const int32_t r0cta = (int32_t)(m_blk * BM2) + (int32_t)rank * UG_BM;
          const bool ln_tile = (KIND == EPI_LNMOD) && n_blk < ep.ln_tiles;
          const int nblk = ln_tile ? 4 : BN / 32;
          for (int blk = 0; blk < nblk; ++blk, ++seq) {
            const int slot = (int)(seq % Cfg::kRing);
            const uint32_t ph = (seq / Cfg::kRing) & 1u;
            mbar_wait(&sready_bar[slot], ph);
            uint8_t* sp = epi_smem + slot * Cfg::kSlotBytes;
            const int32_t cc = n_blk * BN + 32 * blk;
            if constexpr (KIND == EPI_LNMOD) {
              if (ln_tile) tma_store_2d(&em.g, sp + Cfg::kSlotH, n_blk * 128 + 32 * blk, r0cta);
              else tma_store_2d(&em.h_out, sp + Cfg::kSlotH, (n_blk - ep.ln_tiles) * BN + 32 * blk, r0cta);
            } else if constexpr (KIND == EPI_GATED_RES) {
              tma_store_2d(&em.h_out, sp, cc, r0cta);
            } else if constexpr (KIND == EPI_SILU_SAVE) {
              tma_store_2d(&em.h_out, sp + Cfg::kSlotH, cc, r0cta);
              tma_store_2d(&em.g, sp, cc, r0cta);
            } else if constexpr (KIND == EPI_GATED_SAVE) {
              tma_store_2d(&em.h_out, sp, cc, r0cta);
              tma_store_2d(&em.b2, sp + Cfg::kSlotH, cc, r0cta);
            } else {  // SILU_ACT, STORE_ACT, DSILU_ACT: one bf16 box
              tma_store_2d(&em.h_out, sp + Cfg::kSlotH, cc, r0cta);
            }
            bulk_commit_group();
            if (prev_slot >= 0) {
              bulk_wait_read<1>();  // every store but the one just issued has read its box
              mbar_arrive(&eempty_bar[prev_slot]);
            }
            prev_slot = slot;
          }
        }
        if (prev_slot >= 0) {
          bulk_wait_read<0>();
          mbar_arrive(&eempty_bar[prev_slot]);
        }
        bulk_wait_all();  // this thread's TMA stores have landed
      }
    }
  } else if (warp == 3) {
    // ===================== epilogue operand prefetch (TMA-fed epilogue kinds) =====================
    if constexpr (Cfg::kTmaEpi) {
      if (elect_one()) {
        uint32_t seq = 0;
        for (int64_t item = cluster_id; item < num_items; item += num_clusters) {
          int64_t m_blk; int n_blk;
          ug_tile_coords(item / splits, m_tiles, n_tiles, gm, m_blk, n_blk);
          n_blk = n_blk * kPairs + (int)pair;
          const int32_t r0 = (int32_t)(m_blk * BM2) + (int32_t)rank * UG_BM;
          const int32_t c0 = n_blk * BN;
          const bool ln_tile = (KIND == EPI_LNMOD) && n_blk < ep.ln_tiles;
          const int nblk = ln_tile ? 4 : BN / 32;
          for (int blk = 0; blk < nblk; ++blk, ++seq) {
            const int slot = (int)(seq % Cfg::kRing);
            const uint32_t ph = (seq / Cfg::kRing) & 1u;
            mbar_wait(&eempty_bar[slot], ph ^ 1u);
            uint8_t* sp = epi_smem + slot * Cfg::kSlotBytes;
            if constexpr (Cfg::kTmaOutOnly) {
              mbar_arrive(&efull_bar[slot]);  // relays "slot free" to the epilogue warps
            } else if constexpr (KIND == EPI_LNMOD) {
              if (ln_tile) {  // residual columns of this tile's 128 features
                mbar_arrive_expect_tx(&efull_bar[slot], (uint32_t)Cfg::kSlotH);
                tma_load_2d(sp, &em.h_in, &efull_bar[slot], n_blk * 128 + 32 * blk, r0);
              } else {
                mbar_arrive(&efull_bar[slot]);  // gate tile: the slot is only an output staging buffer
              }
            } else if constexpr (KIND == EPI_DSILU_ACT) {
              mbar_arrive_expect_tx(&efull_bar[slot], (uint32_t)Cfg::kSlotG);
              tma_load_2d(sp + Cfg::kSlotH, &em.g, &efull_bar[slot], c0 + 32 * blk, r0);
            } else {
              mbar_arrive_expect_tx(&efull_bar[slot], (uint32_t)Cfg::kSlotBytes);
              tma_load_2d(sp, &em.h_in, &efull_bar[slot], c0 + 32 * blk, r0);
              tma_load_2d(sp + Cfg::kSlotH, &em.g, &efull_bar[slot], c0 + 32 * blk, r0);
            }
          }
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue (this CTA's 128 rows) =====================
    const int q = warp & 3;          // TMEM lane quarter
    const int cg = (warp - 4) >> 2;  // column half of the tile
    constexpr int kColsPerGroup = BN / 2;
    uint8_t* wsm = epi_smem + (Cfg::kTmaEpi ? 0 : (warp - 4) * PE_WARP_BYTES);
    const WarpStage ws{reinterpret_cast<float*>(wsm), reinterpret_cast<uint32_t*>(wsm + 32 * PE_F32_LD * 4)};
    constexpr bool kStaged = (KIND == EPI_LNMOD || KIND == EPI_GATED_RES || KIND == EPI_SILU_ACT || KIND == EPI_STORE_ACT);
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t tile_seq = 0;  // tiles this CTA has finished (ring position of the TMA-fed epilogue)
    uint32_t seq_base = 0;  // EPI_LNMOD: ring boxes consumed by the finished tiles (4 or 8 per tile)
    for (int64_t item = cluster_id; item < num_items; item += num_clusters, ++tile_seq) {
      int64_t m_blk; int n_blk;
      ug_tile_coords(item / splits, m_tiles, n_tiles, gm, m_blk, n_blk);
      n_blk = n_blk * kPairs + (int)pair;
      const int64_t row0 = m_blk * BM2 + (int64_t)rank * UG_BM + q * 32;  // first row of this warp
      const int64_t left = M - row0;
      const int rows_valid = left >= 32 ? 32 : (left > 0 ? (int)left : 0);
      const int64_t row = row0 + lane;
      const bool row_ok = lane < rows_valid;
      const int n0 = n_blk * BN + cg * kColsPerGroup;
      const uint32_t t_tile = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
      if constexpr (KIND == EPI_LNMOD) {
        const int W = ep.ln_width;
        const int r = q * 32 + lane;  // row inside this CTA's 128-row box
        const uint32_t rsw = (uint32_t)((r >> 1) & 3);
        if (n_blk < ep.ln_tiles) {
          // (shift | scale) tile: warp group cg takes the 32-feature boxes j = cg, cg + 2 of features f0 .. f0+127
          const int f0 = n_blk * 128;
          float mean = 0.f, rstd = 0.f;
          if (row_ok) {  // row sums written by the producer of h, one (sum, sum of squares) per 128 columns
            const float4* sp4 = reinterpret_cast<const float4*>(ep.stats + row * (W / 64));
            float s = 0.f, ss = 0.f;
            for (int i = 0; i < W / 256; ++i) {
              const float4 t = sp4[i];
              s += t.x + t.z;
              ss += t.y + t.w;
            }
            mean = s / (float)W;
            rstd = rsqrtf(fmaxf(ss / (float)W - mean * mean, 0.f) + 1e-6f);
          }
          mbar_wait(&tfull_bar[acc], acc_phase);
          tc_fence_after();
#pragma unroll 1
          for (int bi = 0; bi < 2; ++bi) {
            const int j = 2 * bi + cg;
            const uint32_t seq = seq_base + (uint32_t)j;
            const int slot = (int)(seq % Cfg::kRing);
            const uint32_t ph = (seq / Cfg::kRing) & 1u;
            uint8_t* sp = epi_smem + slot * Cfg::kSlotBytes;
            uint32_t sh[32], sc[32];
            tmem_ld_32x32(t_tile + (uint32_t)(32 * j), sh);
            tmem_ld_32x32(t_tile + (uint32_t)(128 + 32 * j), sc);
            mbar_wait(&efull_bar[slot], ph);
            float4 hv[8];
            const uint8_t* hrow = sp + r * 128;
#pragma unroll
            for (int i = 0; i < 8; ++i) hv[i] = *reinterpret_cast<const float4*>(hrow + ((i ^ (r & 7)) << 4));
            tmem_ld_wait();
            const float* bsh = ep.bias + n_blk * 256 + 32 * j;
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 b1 = __ldg(reinterpret_cast<const float4*>(bsh) + i);
              const float4 b2 = __ldg(reinterpret_cast<const float4*>(bsh + 128) + i);
              float n0_ = (hv[i].x - mean) * rstd, n1_ = (hv[i].y - mean) * rstd;
              float n2_ = (hv[i].z - mean) * rstd, n3_ = (hv[i].w - mean) * rstd;
              if (ep.ln_w != nullptr) {
                const float4 g = __ldg(reinterpret_cast<const float4*>(ep.ln_w + f0 + 32 * j) + i);
                const float4 bb = __ldg(reinterpret_cast<const float4*>(ep.ln_b + f0 + 32 * j) + i);
                n0_ = fmaf(n0_, g.x, bb.x); n1_ = fmaf(n1_, g.y, bb.y); n2_ = fmaf(n2_, g.z, bb.z); n3_ = fmaf(n3_, g.w, bb.w);
              }
              const float o0 = fmaf(n0_, 1.0f + (__uint_as_float(sc[4 * i]) + b2.x), __uint_as_float(sh[4 * i]) + b1.x);
              const float o1 = fmaf(n1_, 1.0f + (__uint_as_float(sc[4 * i + 1]) + b2.y), __uint_as_float(sh[4 * i + 1]) + b1.y);
              const float o2 = fmaf(n2_, 1.0f + (__uint_as_float(sc[4 * i + 2]) + b2.z), __uint_as_float(sh[4 * i + 2]) + b1.z);
              const float o3 = fmaf(n3_, 1.0f + (__uint_as_float(sc[4 * i + 3]) + b2.w), __uint_as_float(sh[4 * i + 3]) + b1.w);
              pk[2 * i] = pack_bf16x2(o0, o1);
              pk[2 * i + 1] = pack_bf16x2(o2, o3);
            }
            uint8_t* orow = sp + Cfg::kSlotH + r * 64;
#pragma unroll
            for (int i = 0; i < 4; ++i)
              *reinterpret_cast<uint4*>(orow + ((i ^ rsw) << 4)) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
            fence_proxy_async();
            named_bar_sync(1 + cg, 128);
            if (q == 0 && lane == 0) mbar_arrive(&sready_bar[slot]);  // hand the box to the store thread (warp 2)
          }
          seq_base += 4;
        } else {
          // gate tile: bf16 boxes into out2 (h_out map)
          mbar_wait(&tfull_bar[acc], acc_phase);
          tc_fence_after();
#pragma unroll 1
          for (int bi = 0; bi < BN / 64; ++bi) {
            const int blk = 2 * bi + cg;
            const uint32_t seq = seq_base + (uint32_t)blk;
            const int slot = (int)(seq % Cfg::kRing);
            const uint32_t ph = (seq / Cfg::kRing) & 1u;
            uint8_t* sp = epi_smem + slot * Cfg::kSlotBytes;
            uint32_t v[32];
            tmem_ld_32x32(t_tile + (uint32_t)(32 * blk), v);
            float b[32];
            bias32_ldg(ep.bias + n_blk * 256 + 32 * blk, b);
            mbar_wait(&efull_bar[slot], ph);  // slot free (relayed by the prefetch warp)
            tmem_ld_wait();
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) pk[i] = pack_bf16x2(__uint_as_float(v[2 * i]) + b[2 * i], __uint_as_float(v[2 * i + 1]) + b[2 * i + 1]);
            uint8_t* orow = sp + Cfg::kSlotH + r * 64;
#pragma unroll
            for (int i = 0; i < 4; ++i)
              *reinterpret_cast<uint4*>(orow + ((i ^ rsw) << 4)) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
            fence_proxy_async();
            named_bar_sync(1 + cg, 128);
            if (q == 0 && lane == 0) mbar_arrive(&sready_bar[slot]);  // hand the box to the store thread (warp 2)
          }
          seq_base += BN / 32;
        }
      } else if constexpr (KIND == EPI_GATED_RES) {
        // h[rows, cols] = h + gate * (acc + bias) on the fp32 residual stream + row sums for the next LayerNorm.
        // Warp group cg takes the 32-column boxes blk = cg, cg + 2, ... of the prefetch ring.
        const int r = q * 32 + lane;  // row inside this CTA's 128-row box
        float s = 0.f, ss = 0.f;
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
#pragma unroll 1
        for (int bi = 0; bi < BN / 64; ++bi) {
          const int blk = 2 * bi + cg;
          const uint32_t seq = tile_seq * (BN / 32) + (uint32_t)blk;
          const int slot = (int)(seq % Cfg::kRing);
          const uint32_t ph = (seq / Cfg::kRing) & 1u;
          uint8_t* sp = epi_smem + slot * Cfg::kSlotBytes;
          uint32_t v[32];
          tmem_ld_32x32(t_tile + (uint32_t)(blk * 32), v);
          float b[32];
          bias32_ldg(ep.bias != nullptr ? ep.bias + n_blk * BN + blk * 32 : nullptr, b);
          mbar_wait(&efull_bar[slot], ph);
          float4 hv[8];
          uint4 gv[4];
          uint8_t* hrow = sp + r * 128;
          const uint8_t* grow = sp + Cfg::kSlotH + r * 64;
#pragma unroll
          for (int i = 0; i < 8; ++i) hv[i] = *reinterpret_cast<const float4*>(hrow + ((i ^ (r & 7)) << 4));
#pragma unroll
          for (int i = 0; i < 4; ++i) gv[i] = *reinterpret_cast<const uint4*>(grow + ((i ^ ((r >> 1) & 3)) << 4));
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint4 gq = gv[i >> 1];
            const uint32_t w0 = (i & 1) ? gq.z : gq.x, w1 = (i & 1) ? gq.w : gq.y;
            const float2 ga = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w0));
            const float2 gb = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w1));
            float4 o = hv[i];
            o.x = fmaf(ga.x, __uint_as_float(v[4 * i]) + b[4 * i], o.x);
            o.y = fmaf(ga.y, __uint_as_float(v[4 * i + 1]) + b[4 * i + 1], o.y);
            o.z = fmaf(gb.x, __uint_as_float(v[4 * i + 2]) + b[4 * i + 2], o.z);
            o.w = fmaf(gb.y, __uint_as_float(v[4 * i + 3]) + b[4 * i + 3], o.w);
            s += (o.x + o.y) + (o.z + o.w);
            ss += (o.x * o.x + o.y * o.y) + (o.z * o.z + o.w * o.w);
            *reinterpret_cast<float4*>(hrow + ((i ^ (r & 7)) << 4)) = o;
          }
          fence_proxy_async();
          named_bar_sync(1 + cg, 128);
          if (q == 0 && lane == 0) mbar_arrive(&sready_bar[slot]);  // hand the box to the store thread (warp 2)
        }
        if (ep.stats != nullptr && row_ok)
          *reinterpret_cast<float2*>(ep.stats + (row * (ep.ln_width / 128) + n_blk * 2 + cg) * 2) = make_float2(s, ss);
      } else if constexpr (KIND == EPI_SILU_ACT || KIND == EPI_STORE_ACT) {
        // out(bf16) = f(acc + bias): 32-column boxes staged in the ring and written with TMA stores (h_out map)
        const int r = q * 32 + lane;
        const uint32_t rsw = (uint32_t)((r >> 1) & 3);
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
#pragma unroll 1
        for (int bi = 0; bi < BN / 64; ++bi) {
          const int blk = 2 * bi + cg;
          const uint32_t seq = tile_seq * (BN / 32) + (uint32_t)blk;
          const int slot = (int)(seq % Cfg::kRing);
          const uint32_t ph = (seq / Cfg::kRing) & 1u;
          uint8_t* sp = epi_smem + slot * Cfg::kSlotBytes;
          uint32_t v[32];
          tmem_ld_32x32(t_tile + (uint32_t)(32 * blk), v);
          float b[32];
          bias32_ldg(ep.bias != nullptr ? ep.bias + n_blk * BN + 32 * blk : nullptr, b);
          tmem_ld_wait();
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float x0 = __uint_as_float(v[2 * i]) + b[2 * i], x1 = __uint_as_float(v[2 * i + 1]) + b[2 * i + 1];
            if constexpr (KIND == EPI_SILU_ACT) {
              x0 = silu_tanh(x0);
              x1 = silu_tanh(x1);
            }
            pk[i] = pack_bf16x2(x0, x1);
          }
          mbar_wait(&efull_bar[slot], ph);  // slot free (relayed by warp 3)
          uint8_t* orow = sp + Cfg::kSlotH + r * 64;
#pragma unroll
          for (int i = 0; i < 4; ++i)
            *reinterpret_cast<uint4*>(orow + ((i ^ rsw) << 4)) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
          fence_proxy_async();
          named_bar_sync(1 + cg, 128);
          if (q == 0 && lane == 0) mbar_arrive(&sready_bar[slot]);  // hand the box to the store thread (warp 2)
        }
      } else if constexpr (KIND == EPI_DSILU_ACT || KIND == EPI_SILU_SAVE || KIND == EPI_GATED_SAVE) {
        // training kinds on the same ring (32-column boxes, warp group cg takes blk = cg, cg + 2, ...):
        //   DSILU_ACT : out(bf16) = acc * SiLU'(aux)            aux box prefetched (g map), result in place
        //   SILU_SAVE : out2 = p = acc + bias, out = SiLU(p)      two bf16 output boxes (g part: out, h part: out2)
        //   GATED_SAVE: q = acc + bias -> out2(bf16); out(f32) = res + gate * q   (res / gate prefetched)
        const int r = q * 32 + lane;
        const uint32_t rsw = (uint32_t)((r >> 1) & 3);
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
#pragma unroll 1
        for (int bi = 0; bi < BN / 64; ++bi) {
          const int blk = 2 * bi + cg;
          const uint32_t seq = tile_seq * (BN / 32) + (uint32_t)blk;
          const int slot = (int)(seq % Cfg::kRing);
          const uint32_t ph = (seq / Cfg::kRing) & 1u;
          uint8_t* sp = epi_smem + slot * Cfg::kSlotBytes;
          uint8_t* grow = sp + Cfg::kSlotH + r * 64;
          uint32_t v[32];
          tmem_ld_32x32(t_tile + (uint32_t)(32 * blk), v);
          float b[32];
          bias32_ldg((KIND != EPI_DSILU_ACT && ep.bias != nullptr) ? ep.bias + n_blk * BN + 32 * blk : nullptr, b);
          mbar_wait(&efull_bar[slot], ph);
          uint32_t pk[16];
          if constexpr (KIND == EPI_DSILU_ACT) {
            uint4 av[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) av[i] = *reinterpret_cast<const uint4*>(grow + ((i ^ rsw) << 4));
            tmem_ld_wait();
            const uint32_t* aw = reinterpret_cast<const uint32_t*>(av);
            if (ep.colsum == nullptr) {
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const float2 pq = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&aw[i]));
                pk[i] = pack_bf16x2(__uint_as_float(v[2 * i]) * dsilu_tanh(pq.x), __uint_as_float(v[2 * i + 1]) * dsilu_tanh(pq.y));
              }
            } else {
              // the same products, plus their column sums over the warp's 32 rows (rows past M are zero: TMA
              // zero-fills both operands): a transposing butterfly, step k exchanges half of the live values with
              // lane ^ (16 >> k) so that after 5 steps lane l holds the sum of column l -> 31 shuffles
              // and adds per thread instead of 160, then one atomic per column and warp
              float f[32];
#pragma unroll
              for (int i = 0; i < 16; ++i) {
                const float2 pq = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&aw[i]));
                f[2 * i] = __uint_as_float(v[2 * i]) * dsilu_tanh(pq.x);
                f[2 * i + 1] = __uint_as_float(v[2 * i + 1]) * dsilu_tanh(pq.y);
                pk[i] = pack_bf16x2(f[2 * i], f[2 * i + 1]);
              }
#pragma unroll
              for (int half = 16; half >= 1; half >>= 1) {
                const bool up = (lane & half) != 0;
#pragma unroll
                for (int i = 0; i < half; ++i) {
                  const float give = up ? f[i] : f[i + half];
                  const float keep = up ? f[i + half] : f[i];
                  f[i] = keep + __shfl_xor_sync(0xffffffffu, give, half);
                }
              }
              // lane bit `half` selected the upper half of the live columns at every step: lane l ends with column l
              atomicAdd(ep.colsum + n_blk * BN + 32 * blk + lane, f[0]);
            }
          } else if constexpr (KIND == EPI_SILU_SAVE) {
            tmem_ld_wait();
            uint32_t pp[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float x0 = __uint_as_float(v[2 * i]) + b[2 * i], x1 = __uint_as_float(v[2 * i + 1]) + b[2 * i + 1];
              pp[i] = pack_bf16x2(x0, x1);
              pk[i] = pack_bf16x2(silu_tanh(x0), silu_tanh(x1));  // 3 instructions per element: the 8 epilogue warps are latency bound
            }
            uint8_t* prow = sp + r * 64;  // first 8 KB of the h part: the pre-activation box
#pragma unroll
            for (int i = 0; i < 4; ++i)
              *reinterpret_cast<uint4*>(prow + ((i ^ rsw) << 4)) = make_uint4(pp[4 * i], pp[4 * i + 1], pp[4 * i + 2], pp[4 * i + 3]);
          } else {
            float4 hv[8];
            uint4 gv[4];
            uint8_t* hrow = sp + r * 128;
#pragma unroll
            for (int i = 0; i < 8; ++i) hv[i] = *reinterpret_cast<const float4*>(hrow + ((i ^ (r & 7)) << 4));
#pragma unroll
            for (int i = 0; i < 4; ++i) gv[i] = *reinterpret_cast<const uint4*>(grow + ((i ^ rsw) << 4));
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const uint4 gq = gv[i >> 1];
              const uint32_t w0 = (i & 1) ? gq.z : gq.x, w1 = (i & 1) ? gq.w : gq.y;
              const float2 ga = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w0));
              const float2 gb = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&w1));
              const float q0 = __uint_as_float(v[4 * i]) + b[4 * i], q1 = __uint_as_float(v[4 * i + 1]) + b[4 * i + 1];
              const float q2 = __uint_as_float(v[4 * i + 2]) + b[4 * i + 2], q3 = __uint_as_float(v[4 * i + 3]) + b[4 * i + 3];
              pk[2 * i] = pack_bf16x2(q0, q1);
              pk[2 * i + 1] = pack_bf16x2(q2, q3);
              float4 o = hv[i];
              o.x = fmaf(ga.x, q0, o.x); o.y = fmaf(ga.y, q1, o.y); o.z = fmaf(gb.x, q2, o.z); o.w = fmaf(gb.y, q3, o.w);
              *reinterpret_cast<float4*>(hrow + ((i ^ (r & 7)) << 4)) = o;
            }
          }
#pragma unroll
          for (int i = 0; i < 4; ++i)
            *reinterpret_cast<uint4*>(grow + ((i ^ rsw) << 4)) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
          fence_proxy_async();
          named_bar_sync(1 + cg, 128);
          if (q == 0 && lane == 0) mbar_arrive(&sready_bar[slot]);  // hand the box to the store thread (warp 2)
        }
      } else if constexpr (KIND == EPI_CONVPOOL) {
        // y = max(acc + bias, 0) per pixel (row) and channel; 4 x 4 average pool: this CTA's 128 rows are 8 w-rows x 16 h
        // of one image, warp q holds w = 2q, 2q + 1 (lane >> 4) and h = lane & 15, so a pool region is the 8 lanes with
        // the same (lane >> 2) & 3 of TWO warps: three shuffle steps (lane bits 0, 1, 4), then one atomic per channel
        // from lanes 0, 4, 8, 12 (the two warps of a region meet in the atomics)
        const int64_t img = m_blk;
        const int region0 = ((int)rank * 2 + (q >> 1)) * 4;  // pool row (w / 4) * 4
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
#pragma unroll 1
        for (int c0 = 0; c0 < kColsPerGroup; c0 += 32) {
          uint32_t v[32];
          tmem_ld_32x32(t_tile + (uint32_t)(cg * kColsPerGroup + c0), v);
          float b[32];
          bias32_ldg(ep.bias + n0 + c0, b);
          tmem_ld_wait();
          float y[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) y[i] = fmaxf(__uint_as_float(v[i]) + b[i], 0.f);
          if (ep.out2 != nullptr) {  // kept for the backward (its sign is the ReLU mask)
            uint4* o2 = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(ep.out2) + row * ep.ldo2 + n0 + c0);
#pragma unroll
            for (int i = 0; i < 4; ++i)
              o2[i] = make_uint4(pack_bf16x2(y[8 * i], y[8 * i + 1]), pack_bf16x2(y[8 * i + 2], y[8 * i + 3]),
                                 pack_bf16x2(y[8 * i + 4], y[8 * i + 5]), pack_bf16x2(y[8 * i + 6], y[8 * i + 7]));
          }
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            float t = y[i];
            t += __shfl_xor_sync(0xffffffffu, t, 1);
            t += __shfl_xor_sync(0xffffffffu, t, 2);
            t += __shfl_xor_sync(0xffffffffu, t, 16);
            y[i] = t;
          }
          if ((lane & 0x13) == 0) {
            float* o = reinterpret_cast<float*>(ep.out) + (img * 16 + region0 + (lane >> 2)) * (int64_t)ep.ldo + n0 + c0;
#pragma unroll
            for (int i = 0; i < 32; ++i) atomicAdd(o + i, y[i] * 0.0625f);
          }
        }
      } else {
        typename RowEpi<KIND, TA>::Pre pre;
        if (RowEpi<KIND, TA>::kNeedsPrefetch && row_ok) RowEpi<KIND, TA>::prefetch(ep, row, n0, pre);
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();
#pragma unroll 1
        for (int c0 = 0; c0 < kColsPerGroup; c0 += 32) {
          uint32_t v[32];
          tmem_ld_32x32(t_tile + (uint32_t)(cg * kColsPerGroup + c0), v);
          tmem_ld_wait();
          if (row_ok) RowEpi<KIND, TA>::store(ep, row, n0 + c0, v, pre);
          if (RowEpi<KIND, TA>::kNeedsPrefetch && row_ok && c0 + 32 < kColsPerGroup)
            RowEpi<KIND, TA>::prefetch(ep, row, n0 + c0 + 32, pre);
        }
        if (row_ok) RowEpi<KIND, TA>::finish(ep, row, pre);
      }
      (void)kStaged;
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&tempty_bar[acc], crank & ~1u);
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_2sm<Cfg::kTmemCols>(tmem_base);
  }
}

}  // namespace uva
