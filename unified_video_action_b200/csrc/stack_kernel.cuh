// Persistent layer-chain kernel of the bf16 throughput path (sm_100a): ALL tcgen05 GEMMs of one denoise step of
// SimpleMLPAdaLN (DL:261-283: per ResBlock the AdaLN modulation GEMM with the LayerNorm-modulate epilogue, GEMM1 + SiLU,
// GEMM2 + gated residual; then the final layer's AdaLN GEMM) in ONE launch, scheduled panel-major so that the
// activations of a row panel stay in the 126 MB L2 between the layers instead of round-tripping HBM.
//
// Why: as separate launches over a 65 536-row chunk every layer streams its activations from / to HBM (11.3 GB per
// denoise step; the gated-residual GEMM alone 771 MB per launch at 79 % of the measured HBM peak with the tensor pipe
// 45 % active, profiles/r01_ncu_full_v12_summary.txt).  Here the rows are cut into GROUPS of `group` 256-row tiles
// (default 16 = 4096 rows = 48 MB of h / a / u / gate / ys); the item list is
//     for group: for layer (19 of them): for m-tile in group: for n-tile of the layer
// and the persistent clusters (one CTA pair per TPC, 256 x 256 tiles, cta_group::2 — the pipeline of
// umma_gemm2.cuh) take items round-robin in that order.  A layer's tile may only start when the tiles it reads are
// complete; completion is a counter per (layer, m-tile) in global memory:
//   * the epilogue warps only compute and stage 32-column boxes in shared memory; ONE store thread per CTA (warp 2)
//     issues the TMA stores, frees the staging slots when the stores have read them, waits for an item's stores to
//     COMPLETE (cp.async.bulk.wait_group 0) and then signals  red.release.gpu  (2 signals per item: one per CTA) -
//     no epilogue warp ever waits for a store;
//   * the TMA producer polls  ld.acquire.gpu  on the counter of the layer that writes its A operand
//     (u for GEMM2, a for GEMM1; the AdaLN GEMM reads ys, which exists before the launch, so 61 % of the MMA work
//     never waits), issues  fence.proxy.async  and only then loads A — the weight box of the same pipeline stage is
//     requested BEFORE the poll, weights depend on nothing;
//   * the epilogue-operand prefetch warp polls the counter of the layer that wrote h / gate (and whose reads of the
//     buffers this item overwrites must have finished: write-after-read on a, gate, u) before it loads the h / gate
//     boxes or relays a free staging slot; the epilogue warps poll it once more before they read the row statistics.
// Items are ordered so that every dependency has a smaller item index, every cluster works through its items in
// order and all clusters are co-resident (persistent grid <= cudaOccupancyMaxActiveClusters), so the item with the
// smallest index that is not finished can always make progress: no deadlock.  Every wait is bounded (trap after 4 s).
#pragma once
#include "umma_gemm2.cuh"

namespace uva {

constexpr int SK_MAX_LAYERS = 25;  // 3 * depth + 1, depth <= 8
enum { SK_LNMOD = 0, SK_SILU = 1, SK_GATED = 2 };

struct StackLayer {
  int kind;           // SK_*
  int n_tiles;        // 256-column output tiles
  int ln_tiles;       // SK_LNMOD: leading tiles that are (shift | scale) pairs producing `a`; the rest write the gate
  int a_src;          // A operand: 0 = ys, 1 = a, 2 = u
  int w_row0;         // first row of this layer's weight in the [rows, W] weight map
  int dep_layer;      // layer whose completion (per m-tile) the A operand needs; -1: none
  int dep_count;      // = 2 * n_tiles of that layer (one signal per CTA of the pair and tile)
  int epi_dep_layer;  // layer whose completion the epilogue operands / overwritten buffers need; -1: none
  int epi_dep_count;
  int pad_;
  const float* bias;      // [256 * n_tiles]
  const float* ln_w;      // SK_LNMOD: LayerNorm affine [W] or null
  const float* ln_b;
  const float* stats_in;  // SK_LNMOD: row sums of h, [M][W/64] (sum, sum of squares per 128 columns)
  float* stats_out;       // SK_GATED: row sums of the new h
};

struct StackArgs {
  StackLayer L[SK_MAX_LAYERS];
  int n_layers;
  int items_per_mtile;  // sum of n_tiles over the layers
  int group;            // > 0: m-tiles per group (group-major order); 0: diagonal wavefront order
  int W;
  // diagonal order: P[l] = items of the layers < l of one m-tile, S[k] = sum_{i<k} P[i+1], Q[k] = sum_{i=1..k} P[i]
  int P[SK_MAX_LAYERS + 1], S[SK_MAX_LAYERS + 1], Q[SK_MAX_LAYERS + 1];
  int64_t M;
  unsigned int* done;   // [n_layers][m_tiles], zero at launch
};

struct StackMaps {
  CUtensorMap ys, a_ld, u_ld;  // A operands: bf16 [M, W], box 64 x 128, 128-byte swizzle
  CUtensorMap w;               // all weights: bf16 [rows, W], box 64 x 128, 128-byte swizzle
  CUtensorMap h;               // fp32 residual stream [M, W], box 32 x 128, 128-byte swizzle (load and store)
  CUtensorMap a_st, gate, u_st;  // bf16 [M, W], box 32 x 128, 64-byte swizzle (epilogue stores; gate also loaded)
};

struct SkItem {
  int layer, n;
  int64_t m;
};

// Item order.  group > 0: for group: for layer: for m-tile: for n-tile (a layer's tiles wait for the previous layer
// of the same group: with 64 tiles per small layer and 74 clusters most of a layer boundary is a stall).
// group == 0 (default): DIAGONAL WAVEFRONT.  Wave t holds, for every m-tile j with 0 <= t - j < n_layers, the tiles
// of layer l = t - j: each m-tile advances one layer per wave and the m-tiles in flight are all at different layers,
// so a tile and the tiles it depends on (same m-tile, previous layer) are a whole wave (all layers once = 128 tiles
// for depth 6) apart in the item order, while only n_layers m-tiles (19 x 256 rows = 58 MB of activations, plus the
// 71.5 MB of weights) are live at any time: they stay in L2.
//   items before wave t:  C(t) = A(t) - B(t),  A(t) = t <= NL ? S[t] : S[NL] + (t - NL) ipm,  B(t) = t <= MT ? 0 : Q[t - MT]
__device__ __forceinline__ int64_t sk_wave_start(const StackArgs& sa, int64_t t, int64_t MT) {
  const int NL = sa.n_layers;
  const int64_t a = t <= NL ? (int64_t)sa.S[t] : (int64_t)sa.S[NL] + (t - NL) * sa.items_per_mtile;
  const int64_t b = t <= MT ? 0 : (int64_t)sa.Q[t - MT];
  return a - b;
}

__device__ __forceinline__ SkItem sk_decode(const StackArgs& sa, int64_t item, int64_t m_tiles) {
  SkItem it;
  if (sa.group <= 0) {
    const int NL = sa.n_layers;
    int64_t lo = 0, hi = m_tiles + NL - 1;  // waves [lo, hi): largest t with C(t) <= item
    while (hi - lo > 1) {
      const int64_t mid = (lo + hi) >> 1;
      if (sk_wave_start(sa, mid, m_tiles) <= item) lo = mid;
      else hi = mid;
    }
    const int64_t t = lo;
    const int l_lo = (int)(t - m_tiles + 1 > 0 ? t - m_tiles + 1 : 0);
    const int l_hi = (int)(t < NL - 1 ? t : NL - 1);
    const int r = (int)(item - sk_wave_start(sa, t, m_tiles)) + sa.P[l_lo];
    int l = l_lo;
    while (l < l_hi && r >= sa.P[l + 1]) ++l;
    it.layer = l;
    it.m = t - l;
    it.n = r - sa.P[l];
    return it;
  }
  const int64_t per_group = (int64_t)sa.group * sa.items_per_mtile;
  const int64_t g = item / per_group;
  int64_t rem = item - g * per_group;
  const int64_t m_first = g * sa.group;
  int64_t gm = m_tiles - m_first;
  if (gm > sa.group) gm = sa.group;
  int l = 0;
  for (; l < sa.n_layers - 1; ++l) {
    const int64_t cnt = gm * sa.L[l].n_tiles;
    if (rem < cnt) break;
    rem -= cnt;
  }
  const int nt = sa.L[l].n_tiles;
  it.layer = l;
  it.m = m_first + rem / nt;
  it.n = (int)(rem % nt);
  return it;
}

__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// spin until *p >= target (acquire, gpu scope); bounded like mbar_wait
__device__ __forceinline__ void sk_wait_done(const unsigned int* p, unsigned int target) {
  uint32_t spins = 0;
  uint64_t t0 = 0;
  for (;;) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    if (v >= target) return;
#if !(defined(UVA_DEBUG) && defined(UVA_EXP_SK_NO_SLEEP))
    __nanosleep(40);
#endif
    if ((++spins & 0xFFu) == 0) {
      const uint64_t now = global_timer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > 4000000000ull) __trap();
    }
  }
}
// publish: everything this thread wrote or observed before (its completed TMA stores, the row statistics of the
// threads it met at the named barrier) happens-before a reader that acquires the counter
__device__ __forceinline__ void sk_signal(unsigned int* p) {
  fence_proxy_async_all();
  asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p) : "memory");
}

using SkCfg = Umma2Cfg<256, EPI_GATED_RES>;  // 4 operand stages + the 4-slot epilogue ring (h box 16 KB + bf16 box 8 KB)

__global__ void __launch_bounds__(UG_THREADS, 1)
umma_stack_kernel(const __grid_constant__ StackMaps tm, const __grid_constant__ StackArgs sa) {
  using Cfg = SkCfg;
  constexpr int BN = 256;
  constexpr int kStages = Cfg::kStages;
  constexpr int BM2 = 2 * UG_BM;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stage_base = smem;
  uint8_t* epi_smem = smem + kStages * Cfg::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_smem + Cfg::kEpiBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kStages;
  uint64_t* tfull_bar = bars + 2 * kStages;
  uint64_t* tempty_bar = bars + 2 * kStages + 2;
  uint64_t* efull_bar = bars + 2 * kStages + 4;
  uint64_t* eempty_bar = bars + 2 * kStages + 4 + Cfg::kRing;
  uint64_t* sready_bar = bars + 2 * kStages + 4 + 2 * Cfg::kRing;  // [kRing] box staged, ready for its TMA store
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4 + 3 * Cfg::kRing);
  static_assert((2 * kStages + 4 + 3 * Cfg::kRing) * 8 + 4 <= Cfg::kBarBytes, "barrier area");

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = cluster_ctarank();
  const uint32_t rank = crank & 1u;
  const bool leader = rank == 0;
  const int64_t cluster_id = blockIdx.x / 2;
  const int64_t num_clusters = gridDim.x / 2;
  const int64_t m_tiles = (sa.M + BM2 - 1) / BM2;
  const int64_t num_items = m_tiles * sa.items_per_mtile;
  const int num_kb = sa.W / UG_BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tm.ys); tma_prefetch_desc(&tm.a_ld); tma_prefetch_desc(&tm.u_ld); tma_prefetch_desc(&tm.w);
    tma_prefetch_desc(&tm.h); tma_prefetch_desc(&tm.a_st); tma_prefetch_desc(&tm.gate); tma_prefetch_desc(&tm.u_st);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
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
  griddep_wait();
  griddep_launch();

  if (warp == 0) {
    // ===================== TMA producer (both CTAs) =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int64_t item = cluster_id; item < num_items; item += num_clusters) {
        const SkItem it = sk_decode(sa, item, m_tiles);
        const StackLayer& L = sa.L[it.layer];
        const CUtensorMap* mA = L.a_src == 0 ? &tm.ys : (L.a_src == 1 ? &tm.a_ld : &tm.u_ld);
        const int32_t m0 = (int32_t)(it.m * BM2) + (int32_t)rank * UG_BM;
        const int32_t n0 = L.w_row0 + it.n * BN + (int32_t)rank * Cfg::kBNH;
        bool ready = L.dep_layer < 0;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint8_t* sA = stage_base + stage * Cfg::kStageBytes;
          uint8_t* sB = sA + Cfg::kStageBytesA;
          if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2u * (uint32_t)Cfg::kStageBytes);
          tma_load_2d_2sm(sB, &tm.w, &full_bar[stage], kb * UG_BK, n0);  // weights depend on nothing: request first
          if (!ready) {
            sk_wait_done(sa.done + (int64_t)L.dep_layer * m_tiles + it.m, (unsigned int)L.dep_count);
            fence_proxy_async_all();
            ready = true;
          }
          tma_load_2d_2sm(sA, mA, &full_bar[stage], kb * UG_BK, m0);
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1 && leader) {
    // ===================== MMA issuer (leader CTA) =====================
    constexpr uint32_t idesc = umma_idesc_bf16(BM2, BN, false, false);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int64_t item = cluster_id; item < num_items; item += num_clusters) {
      mbar_wait(&tempty_bar[acc], acc_phase ^ 1u);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa_ = smem_u32(stage_base + stage * Cfg::kStageBytes);
          const uint32_t sb_ = sa_ + Cfg::kStageBytesA;
          const uint64_t da = umma_desc_k_sw128(sa_);
          const uint64_t db = umma_desc_k_sw128(sb_);
#pragma unroll
          for (int k = 0; k < UG_BK / 16; ++k)
            umma_bf16_2sm(d_tmem, da + 2ull * (uint64_t)k, db + 2ull * (uint64_t)k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
          umma_commit_2sm(&empty_bar[stage], 3);
          if (kb == num_kb - 1) umma_commit_2sm(&tfull_bar[acc], 3);
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  } else if (warp == 2) {
    // ===================== store thread: TMA stores of the staged boxes, slot release, completion signal =====================
    if (elect_one()) {
      uint32_t seq = 0;
      for (int64_t item = cluster_id; item < num_items; item += num_clusters) {
        const SkItem it = sk_decode(sa, item, m_tiles);
        const StackLayer& L = sa.L[it.layer];
        const int32_t r0cta = (int32_t)(it.m * BM2) + (int32_t)rank * UG_BM;
        const bool ln_tile = L.kind == SK_LNMOD && it.n < L.ln_tiles;
        const int nblk = ln_tile ? 4 : BN / 32;
        const CUtensorMap* mo = ln_tile ? &tm.a_st : (L.kind == SK_GATED ? &tm.h : (L.kind == SK_SILU ? &tm.u_st : &tm.gate));
        const int32_t c0 = ln_tile ? it.n * 128 : (L.kind == SK_LNMOD ? (it.n - L.ln_tiles) * BN : it.n * BN);
        const uint32_t src_off = L.kind == SK_GATED ? 0u : (uint32_t)Cfg::kSlotH;
        int prev_slot = -1;
        for (int blk = 0; blk < nblk; ++blk, ++seq) {
          const int slot = (int)(seq % Cfg::kRing);
          const uint32_t ph = (seq / Cfg::kRing) & 1u;
          mbar_wait(&sready_bar[slot], ph);
          tma_store_2d(mo, epi_smem + slot * Cfg::kSlotBytes + src_off, c0 + 32 * blk, r0cta);
          bulk_commit_group();
          if (prev_slot >= 0) {
            bulk_wait_read<1>();  // every store but the one just issued has read its box
            mbar_arrive(&eempty_bar[prev_slot]);
          }
          prev_slot = slot;
        }
        bulk_wait_read<0>();
        mbar_arrive(&eempty_bar[prev_slot]);
#if defined(UVA_DEBUG) && defined(UVA_EXP_SK_NO_COMPLETION)  // timing experiment only (results may be wrong)
        (void)0;
#else
        bulk_wait_all();  // this CTA's half of the item is in memory
#endif
        sk_signal(sa.done + (int64_t)it.layer * m_tiles + it.m);
      }
    }
  } else if (warp == 3) {
    // ===================== epilogue operand prefetch / staging-slot relay =====================
    if (elect_one()) {
      uint32_t seq = 0;
      for (int64_t item = cluster_id; item < num_items; item += num_clusters) {
        const SkItem it = sk_decode(sa, item, m_tiles);
        const StackLayer& L = sa.L[it.layer];
        const int32_t r0 = (int32_t)(it.m * BM2) + (int32_t)rank * UG_BM;
        const bool ln_tile = L.kind == SK_LNMOD && it.n < L.ln_tiles;
        const int nblk = ln_tile ? 4 : BN / 32;
        bool ready = L.epi_dep_layer < 0;
        for (int blk = 0; blk < nblk; ++blk, ++seq) {
          const int slot = (int)(seq % Cfg::kRing);
          const uint32_t ph = (seq / Cfg::kRing) & 1u;
          mbar_wait(&eempty_bar[slot], ph ^ 1u);
          if (!ready) {
            // h / gate of this m-tile are complete, and the layer that READ the buffer this item overwrites is done
            sk_wait_done(sa.done + (int64_t)L.epi_dep_layer * m_tiles + it.m, (unsigned int)L.epi_dep_count);
            fence_proxy_async_all();
            ready = true;
          }
          uint8_t* sp = epi_smem + slot * Cfg::kSlotBytes;
          if (ln_tile) {
            mbar_arrive_expect_tx(&efull_bar[slot], (uint32_t)Cfg::kSlotH);
            tma_load_2d(sp, &tm.h, &efull_bar[slot], it.n * 128 + 32 * blk, r0);
          } else if (L.kind == SK_GATED) {
            mbar_arrive_expect_tx(&efull_bar[slot], (uint32_t)Cfg::kSlotBytes);
            tma_load_2d(sp, &tm.h, &efull_bar[slot], it.n * BN + 32 * blk, r0);
            tma_load_2d(sp + Cfg::kSlotH, &tm.gate, &efull_bar[slot], it.n * BN + 32 * blk, r0);
          } else {
            mbar_arrive(&efull_bar[slot]);  // output-only staging: relays "slot free (and overwrite allowed)"
          }
        }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue (this CTA's 128 rows) =====================
    const int q = warp & 3;
    const int cg = (warp - 4) >> 2;
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t seq_base = 0;
    const int W = sa.W;
    const int r = q * 32 + lane;  // row inside this CTA's 128-row box
    const uint32_t rsw = (uint32_t)((r >> 1) & 3);
    for (int64_t item = cluster_id; item < num_items; item += num_clusters) {
      const SkItem it = sk_decode(sa, item, m_tiles);
      const StackLayer& L = sa.L[it.layer];
      const int n_blk = it.n;
      const int64_t row0 = it.m * BM2 + (int64_t)rank * UG_BM + q * 32;
      const int64_t left = sa.M - row0;
      const int rows_valid = left >= 32 ? 32 : (left > 0 ? (int)left : 0);
      const int64_t row = row0 + lane;
      const bool row_ok = lane < rows_valid;
      const uint32_t t_tile = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
      if (L.kind == SK_LNMOD && n_blk < L.ln_tiles) {
        // (shift | scale) tile -> a = LN(h) (1 + scale) + shift for features f0 .. f0 + 127
        const int f0 = n_blk * 128;
        float mean = 0.f, rstd = 0.f;
        if (L.epi_dep_layer >= 0) {
          if (lane == 0) sk_wait_done(sa.done + (int64_t)L.epi_dep_layer * m_tiles + it.m, (unsigned int)L.epi_dep_count);
          __syncwarp();
        }
        if (row_ok) {
          const float4* sp4 = reinterpret_cast<const float4*>(L.stats_in + row * (W / 64));
          float s = 0.f, ss = 0.f;
          for (int i = 0; i < W / 256; ++i) {
            const float4 t = __ldcg(sp4 + i);
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
          const float* bsh = L.bias + n_blk * 256 + 32 * j;
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 b1 = __ldg(reinterpret_cast<const float4*>(bsh) + i);
            const float4 b2 = __ldg(reinterpret_cast<const float4*>(bsh + 128) + i);
            float n0_ = (hv[i].x - mean) * rstd, n1_ = (hv[i].y - mean) * rstd;
            float n2_ = (hv[i].z - mean) * rstd, n3_ = (hv[i].w - mean) * rstd;
            if (L.ln_w != nullptr) {
              const float4 g = __ldg(reinterpret_cast<const float4*>(L.ln_w + f0 + 32 * j) + i);
              const float4 bb = __ldg(reinterpret_cast<const float4*>(L.ln_b + f0 + 32 * j) + i);
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
          if (q == 0 && lane == 0) mbar_arrive(&sready_bar[slot]);  // hand the box to the store thread
        }
        seq_base += 4;
      } else if (L.kind == SK_LNMOD || L.kind == SK_SILU) {
        // bf16 output tile: gate = acc + bias (AdaLN gate third) or u = SiLU(acc + bias) (GEMM1)
        const bool silu = L.kind == SK_SILU;
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
          bias32_ldg(L.bias + n_blk * BN + 32 * blk, b);
          tmem_ld_wait();
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float x0 = __uint_as_float(v[2 * i]) + b[2 * i], x1 = __uint_as_float(v[2 * i + 1]) + b[2 * i + 1];
            if (silu) {
              x0 = silu_tanh(x0);
              x1 = silu_tanh(x1);
            }
            pk[i] = pack_bf16x2(x0, x1);
          }
          mbar_wait(&efull_bar[slot], ph);  // slot free, and the previous readers of the output buffer are done
          uint8_t* orow = sp + Cfg::kSlotH + r * 64;
#pragma unroll
          for (int i = 0; i < 4; ++i)
            *reinterpret_cast<uint4*>(orow + ((i ^ rsw) << 4)) = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
          fence_proxy_async();
          named_bar_sync(1 + cg, 128);
          if (q == 0 && lane == 0) mbar_arrive(&sready_bar[slot]);  // hand the box to the store thread
        }
        seq_base += BN / 32;
      } else {
        // SK_GATED: h += gate * (acc + bias) on the fp32 residual stream, in place, + row sums for the next LayerNorm
        float s = 0.f, ss = 0.f;
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
          tmem_ld_32x32(t_tile + (uint32_t)(blk * 32), v);
          float b[32];
          bias32_ldg(L.bias + n_blk * BN + blk * 32, b);
          mbar_wait(&efull_bar[slot], ph);
          float4 hv[8];
          uint4 gv[4];
          uint8_t* hrow = sp + r * 128;
          const uint8_t* grow = sp + Cfg::kSlotH + r * 64;
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
            float4 o = hv[i];
            o.x = fmaf(ga.x, __uint_as_float(v[4 * i]) + b[4 * i], o.x);
            o.y = fmaf(ga.y, __uint_as_float(v[4 * i + 1]) + b[4 * i + 1], o.y);
            o.z = fmaf(gb.x, __uint_as_float(v[4 * i + 2]) + b[4 * i + 2], o.z);
            o.w = fmaf(gb.y, __uint_as_float(v[4 * i + 3]) + b[4 * i + 3], o.w);
            s += (o.x + o.y) + (o.z + o.w);
            ss += (o.x * o.x + o.y * o.y) + (o.z * o.z + o.w * o.w);
            *reinterpret_cast<float4*>(hrow + ((i ^ (r & 7)) << 4)) = o;
          }
          // row sums of the new h: written before the last box is handed over, so that the store thread's
          // completion signal (release) covers them
          if (bi == BN / 64 - 1 && L.stats_out != nullptr && row_ok)
            *reinterpret_cast<float2*>(L.stats_out + (row * (W / 128) + n_blk * 2 + cg) * 2) = make_float2(s, ss);
          fence_proxy_async();
          named_bar_sync(1 + cg, 128);
          if (q == 0 && lane == 0) mbar_arrive(&sready_bar[slot]);  // hand the box to the store thread
        }
        seq_base += BN / 32;
      }
      // the accumulator stage is drained: hand it back before waiting for the stores
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
