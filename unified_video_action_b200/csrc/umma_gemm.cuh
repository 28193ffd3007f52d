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
//   * warps 4-11: epilogue — 8 warps = 4 TMEM lane quarters x 2 column halves; tcgen05.ld gives
//                each thread 32 consecutive columns of its own row; the fused epilogue works on that
//                row segment with 16-byte vector loads/stores (operand loads are issued before the
//                TMEM load so their latency overlaps it).
#pragma once
#include "common.cuh"
#include "epilogue.cuh"
#include "lnmod_row.cuh"

namespace uva {

constexpr int UG_BM = 128;
constexpr int UG_BK = 64;  // 64 bf16 = 128 bytes = one swizzle row
constexpr int UG_THREADS = 384;

template <int BN>
struct UmmaCfg {
  static constexpr int kStageBytesA = UG_BM * UG_BK * 2;
  static constexpr int kStageBytesB = BN * UG_BK * 2;
  static constexpr int kStageBytes = kStageBytesA + kStageBytesB;
  static constexpr int kStages = (BN >= 256) ? 4 : (BN >= 128 ? 6 : 8);
  static constexpr int kBarBytes = 256;
  static constexpr int kSmemBytes = 1024 /*align slack*/ + kStages * kStageBytes + kBarBytes;
  static constexpr uint32_t kTmemCols = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128 : (2 * BN <= 256) ? 256 : 512;
  static constexpr int kColGroups = (BN >= 64) ? 2 : 1;  // epilogue warps per lane quarter
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

// ------------------------------------------------------------------------------------------
// Row-segment epilogue: thread owns acc[0..31] = D[row, col0 .. col0+31].
// ------------------------------------------------------------------------------------------
template <int KIND, typename TA>
struct RowEpi {
  static constexpr bool kNeedsPrefetch = (KIND == EPI_GATED_RES || KIND == EPI_GATED_SAVE);
  struct Pre {
    float4 h[8];
    uint4 g[4];
    float s = 0.f, ss = 0.f;  // row sums of the outputs written so far (LayerNorm statistics)
  };
  // after the last chunk of a tile: publish the row sums
  __device__ static __forceinline__ void finish(const EpiArgs& e, int64_t row, Pre& p) {
    if constexpr (KIND == EPI_GATED_RES || KIND == EPI_GATED_SAVE) {
      if (e.stats != nullptr) {
        atomicAdd(e.stats + 2 * row, p.s);
        atomicAdd(e.stats + 2 * row + 1, p.ss);
        p.s = p.ss = 0.f;
      }
    }
  }

  __device__ static __forceinline__ void prefetch(const EpiArgs& e, int64_t row, int col0, Pre& p) {
    if constexpr (KIND == EPI_GATED_RES || KIND == EPI_GATED_SAVE) {
      const float4* hp = reinterpret_cast<const float4*>((e.res != nullptr ? e.res : reinterpret_cast<const float*>(e.out)) + row * e.ldo + col0);
      const uint4* gp = reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(e.aux) + row * e.ldaux + col0);
#pragma unroll
      for (int i = 0; i < 8; ++i) p.h[i] = hp[i];
#pragma unroll
      for (int i = 0; i < 4; ++i) p.g[i] = gp[i];
    }
  }

  __device__ static __forceinline__ void bias32(const EpiArgs& e, int col0, float (&b)[32]) {
    if (e.bias != nullptr) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float4 t = __ldg(reinterpret_cast<const float4*>(e.bias + col0) + i);
        b[4 * i] = t.x; b[4 * i + 1] = t.y; b[4 * i + 2] = t.z; b[4 * i + 3] = t.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) b[i] = 0.f;
    }
  }

  __device__ static __forceinline__ void store(const EpiArgs& e, int64_t row, int col0, const uint32_t (&v)[32], Pre& p) {
    constexpr bool kVec = (sizeof(TA) == 2) || KIND == EPI_STORE_F32;  // (gated kinds: TA is bf16 here)
    if constexpr (KIND == EPI_ATOMIC_F32) {
      if (col0 + 32 <= e.n_valid) {
        float* o = reinterpret_cast<float*>(e.out) + row * e.ldo + col0;
#pragma unroll
        for (int i = 0; i < 8; ++i)
          red_add_v4(o + 4 * i, e.alpha * __uint_as_float(v[4 * i]), e.alpha * __uint_as_float(v[4 * i + 1]),
                     e.alpha * __uint_as_float(v[4 * i + 2]), e.alpha * __uint_as_float(v[4 * i + 3]));
        return;
      }
    }
    if (!kVec || col0 + 32 > e.n_valid || KIND == EPI_ACCUM_F32 || KIND == EPI_ATOMIC_F32) {
      // generic tail / rarely used kinds: scalar per-element epilogue
#pragma unroll
      for (int j = 0; j < 32; ++j) Epi<KIND, TA>::apply(e, row, col0 + j, __uint_as_float(v[j]));
      return;
    }
    float b[32];
    bias32(e, col0, b);
    if constexpr (KIND == EPI_STORE_F32) {
      float4* o = reinterpret_cast<float4*>(reinterpret_cast<float*>(e.out) + row * e.ldo + col0);
#pragma unroll
      for (int i = 0; i < 8; ++i)
        o[i] = make_float4(__uint_as_float(v[4 * i]) + b[4 * i], __uint_as_float(v[4 * i + 1]) + b[4 * i + 1],
                           __uint_as_float(v[4 * i + 2]) + b[4 * i + 2], __uint_as_float(v[4 * i + 3]) + b[4 * i + 3]);
    } else if constexpr (KIND == EPI_GATED_RES || KIND == EPI_GATED_SAVE) {
      float4* o = reinterpret_cast<float4*>(reinterpret_cast<float*>(e.out) + row * e.ldo + col0);
      const __nv_bfloat162* g2 = reinterpret_cast<const __nv_bfloat162*>(p.g);
      if constexpr (KIND == EPI_GATED_SAVE) {
        uint4* o2 = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(e.out2) + row * e.ldo2 + col0);
#pragma unroll
        for (int i = 0; i < 4; ++i)
          o2[i] = make_uint4(pack_bf16x2(__uint_as_float(v[8 * i]) + b[8 * i], __uint_as_float(v[8 * i + 1]) + b[8 * i + 1]),
                             pack_bf16x2(__uint_as_float(v[8 * i + 2]) + b[8 * i + 2], __uint_as_float(v[8 * i + 3]) + b[8 * i + 3]),
                             pack_bf16x2(__uint_as_float(v[8 * i + 4]) + b[8 * i + 4], __uint_as_float(v[8 * i + 5]) + b[8 * i + 5]),
                             pack_bf16x2(__uint_as_float(v[8 * i + 6]) + b[8 * i + 6], __uint_as_float(v[8 * i + 7]) + b[8 * i + 7]));
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float2 ga = __bfloat1622float2(g2[2 * i]);
        const float2 gb = __bfloat1622float2(g2[2 * i + 1]);
        float4 r = p.h[i];
        r.x = fmaf(ga.x, __uint_as_float(v[4 * i]) + b[4 * i], r.x);
        r.y = fmaf(ga.y, __uint_as_float(v[4 * i + 1]) + b[4 * i + 1], r.y);
        r.z = fmaf(gb.x, __uint_as_float(v[4 * i + 2]) + b[4 * i + 2], r.z);
        r.w = fmaf(gb.y, __uint_as_float(v[4 * i + 3]) + b[4 * i + 3], r.w);
        o[i] = r;
        p.s += (r.x + r.y) + (r.z + r.w);
        p.ss += (r.x * r.x + r.y * r.y) + (r.z * r.z + r.w * r.w);
      }
    } else {
      // 16-bit outputs: out = f(acc + bias) packed to bf16, 4 x 16-byte stores
      float x[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(v[j]) + b[j];
      if constexpr (KIND == EPI_SILU_SAVE) {
        uint4* o2 = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(e.out2) + row * e.ldo2 + col0);
#pragma unroll
        for (int i = 0; i < 4; ++i)
          o2[i] = make_uint4(pack_bf16x2(x[8 * i], x[8 * i + 1]), pack_bf16x2(x[8 * i + 2], x[8 * i + 3]),
                             pack_bf16x2(x[8 * i + 4], x[8 * i + 5]), pack_bf16x2(x[8 * i + 6], x[8 * i + 7]));
      }
      if constexpr (KIND == EPI_SILU_ACT || KIND == EPI_SILU_SAVE) {
#pragma unroll
        for (int j = 0; j < 32; ++j) x[j] = __fdividef(x[j], 1.0f + __expf(-x[j]));
      }
      if constexpr (KIND == EPI_RELU_ACT) {
#pragma unroll
        for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.f);
      }
      if constexpr (KIND == EPI_DSILU_ACT) {
        const uint4* ap = reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(e.aux) + row * e.ldaux + col0);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const uint4 t = ap[i];
          const __nv_bfloat162* t2 = reinterpret_cast<const __nv_bfloat162*>(&t);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float2 pq = __bfloat1622float2(t2[q]);
            x[8 * i + 2 * q] = __uint_as_float(v[8 * i + 2 * q]) * dsilu_f(pq.x);
            x[8 * i + 2 * q + 1] = __uint_as_float(v[8 * i + 2 * q + 1]) * dsilu_f(pq.y);
          }
        }
      }
      uint4* o = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(e.out) + row * e.ldo + col0);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        o[i] = make_uint4(pack_bf16x2(x[8 * i], x[8 * i + 1]), pack_bf16x2(x[8 * i + 2], x[8 * i + 3]),
                          pack_bf16x2(x[8 * i + 4], x[8 * i + 5]), pack_bf16x2(x[8 * i + 6], x[8 * i + 7]));
    }
  }
};

// A_MN / B_MN: the operand is stored [K, MN] row-major ("MN-major": activations / output gradients
// in the weight-gradient GEMM, the untransposed weight in the data-gradient GEMM); its tensor map
// has 64x64 boxes and one k-block is loaded as BM/64 (BN/64) atoms.  `splits` > 1: split-K, work
// item = (tile, split); only the reducing epilogue kinds may be used with it.
template <int BN, int KIND, typename TA, bool A_MN = false, bool B_MN = false>
__global__ void __launch_bounds__(UG_THREADS, 1)
umma_gemm_tn_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    int64_t M, int N, int K, int gm, int splits, EpiArgs ep) {
  using Cfg = UmmaCfg<BN>;
  constexpr int kStages = Cfg::kStages;
  constexpr uint32_t kAtomBytes = 64 * UG_BK * 2;  // one [64 k][64 mn] atom of an MN-major operand
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* stage_base = smem;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes);
  uint64_t* full_bar = bars;                  // [kStages]
  uint64_t* empty_bar = bars + kStages;       // [kStages]
  uint64_t* tfull_bar = bars + 2 * kStages;   // [2]
  uint64_t* tempty_bar = bars + 2 * kStages + 2;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int64_t m_tiles = (M + UG_BM - 1) / UG_BM;
  const int n_tiles = N / BN;
  const int64_t num_tiles = m_tiles * n_tiles * splits;  // work items
  const int num_kb_total = (K + UG_BK - 1) / UG_BK;       // K tail: TMA zero-fills out of bounds
  const int kb_per = (num_kb_total + splits - 1) / splits;

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
      mbar_init(&tempty_bar[s], 4 * Cfg::kColGroups);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<Cfg::kTmemCols>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_wait();    // (programmatic dependent launch: everything above overlapped the previous kernel's tail)
  griddep_launch();  // after the wait: triggering before it (next kernel two ahead) measured slower

  if constexpr (KIND == EPI_SILU_ACT && !A_MN && sizeof(TA) == 2) {
    if (ep.pre_h != nullptr) {
      // LayerNorm-modulate pre-phase (see EpiArgs): the host guarantees gridDim.x == num_tiles (one tile per CTA, all
      // CTAs co-resident), splits == 1 and K == 1024.  CTA n_blk of a panel takes rows n_blk * rp .. + rp of it.
      int64_t m_blk; int n_blk;
      ug_tile_coords((int64_t)blockIdx.x, m_tiles, n_tiles, gm, m_blk, n_blk);
      const int rp = (UG_BM + n_tiles - 1) / n_tiles;
      int64_t r1 = m_blk * UG_BM + (int64_t)(n_blk + 1) * rp;
      if (r1 > (m_blk + 1) * UG_BM) r1 = (m_blk + 1) * UG_BM;
      if (r1 > M) r1 = M;
      for (int64_t r = m_blk * UG_BM + (int64_t)n_blk * rp + warp; r < r1; r += UG_THREADS / 32)
        lnmod_row<TA, 8>(ep.pre_h, ep.pre_lnw, ep.pre_lnb, reinterpret_cast<const TA*>(ep.pre_shift),
                         reinterpret_cast<const TA*>(ep.pre_scale), ep.pre_ldm, reinterpret_cast<TA*>(ep.pre_a), nullptr, r, lane);
      __syncthreads();
      if (threadIdx.x == 0) {
        // sense-reversing barrier of the n_tiles CTAs of this panel: the generation is read before arriving (it cannot
        // change until this CTA has arrived), the count wraps to zero by itself -> nothing to reset between launches
        unsigned int* cnt = ep.pre_bar + 2 * m_blk;
        unsigned int* gen = cnt + 1;
        unsigned int g0, g;
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(g0) : "l"(gen) : "memory");
        // release: this CTA's rows of a (ordered before this thread by the bar.sync above) are visible before the arrival
        unsigned int old;
        asm volatile("atom.release.gpu.global.inc.u32 %0, [%1], %2;" : "=r"(old) : "l"(cnt), "r"((unsigned int)n_tiles - 1u) : "memory");
        if (old == (unsigned int)n_tiles - 1u) {
          asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(gen) : "memory");
        } else {
          uint32_t spins = 0;
          uint64_t t0 = 0;
          do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(g) : "l"(gen) : "memory");
            if ((++spins & 0xFFFu) == 0) {  // a lost CTA becomes a trap instead of a hung GPU
              const uint64_t now = global_timer_ns();
              if (t0 == 0) t0 = now;
              else if (now - t0 > 4000000000ull) __trap();
            }
          } while (g == g0);
        }
      }
      __syncthreads();
      // the other CTAs wrote a through the generic proxy; the TMA loads below read it through the async proxy
      asm volatile("fence.proxy.async.global;" ::: "memory");
    }
  }

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int64_t item = blockIdx.x; item < num_tiles; item += gridDim.x) {
        int64_t m_blk; int n_blk;
        const int64_t tile = item / splits;
        const int split = (int)(item - tile * splits);
        ug_tile_coords(tile, m_tiles, n_tiles, gm, m_blk, n_blk);
        const int32_t m0 = (int32_t)(m_blk * UG_BM);
        const int32_t n0 = n_blk * BN;
        const int kb0 = split * kb_per;
        const int kb1 = (kb0 + kb_per < num_kb_total) ? kb0 + kb_per : num_kb_total;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint8_t* sa = stage_base + stage * Cfg::kStageBytes;
          uint8_t* sb = sa + Cfg::kStageBytesA;
          mbar_arrive_expect_tx(&full_bar[stage], (uint32_t)Cfg::kStageBytes);
          if constexpr (A_MN) {
#pragma unroll
            for (int a = 0; a < UG_BM / 64; ++a) tma_load_2d(sa + a * kAtomBytes, &tmA, &full_bar[stage], m0 + 64 * a, kb * UG_BK);
          } else {
            tma_load_2d(sa, &tmA, &full_bar[stage], kb * UG_BK, m0);
          }
          if constexpr (B_MN) {
#pragma unroll
            for (int a = 0; a < BN / 64; ++a) tma_load_2d(sb + a * kAtomBytes, &tmB, &full_bar[stage], n0 + 64 * a, kb * UG_BK);
          } else {
            tma_load_2d(sb, &tmB, &full_bar[stage], kb * UG_BK, n0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = umma_idesc_bf16(UG_BM, BN, A_MN, B_MN);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int64_t item = blockIdx.x; item < num_tiles; item += gridDim.x) {
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
          // one UMMA consumes 16 k: K-major +32 bytes inside the 128-byte swizzle row (+2 in 16-byte
          // units); MN-major +16 k-rows of 128 bytes = 2048 bytes (+128 units)
          constexpr uint64_t ka = A_MN ? 128 : 2;
          constexpr uint64_t kbs = B_MN ? 128 : 2;
#pragma unroll
          for (int k = 0; k < UG_BK / 16; ++k) {
            umma_bf16(d_tmem, da + ka * (uint64_t)k, db + kbs * (uint64_t)k, idesc,
                      (kb > kb0 || k > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[stage]);
          if (kb == kb1 - 1) umma_commit(&tfull_bar[acc]);
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1u; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1u; }
    }
  } else if (warp >= 4 && ((warp - 4) >> 2) < Cfg::kColGroups) {
    // ===================== epilogue =====================
    const int q = warp & 3;          // TMEM lane quarter this warp may access
    const int cg = (warp - 4) >> 2;  // column group
    constexpr int kColsPerGroup = BN / Cfg::kColGroups;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int64_t item = blockIdx.x; item < num_tiles; item += gridDim.x) {
      int64_t m_blk; int n_blk;
      ug_tile_coords(item / splits, m_tiles, n_tiles, gm, m_blk, n_blk);
      const int64_t row_in = m_blk * UG_BM + q * 32 + lane;
      const bool row_ok = row_in < M;
      const int64_t row = row_in + (int64_t)(item % splits) * ep.split_rows;  // (split_rows = 0 unless partial planes are asked for)
      const int n0 = n_blk * BN + cg * kColsPerGroup;
      typename RowEpi<KIND, TA>::Pre pre;
      if (RowEpi<KIND, TA>::kNeedsPrefetch && row_ok) RowEpi<KIND, TA>::prefetch(ep, row, n0, pre);
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN + cg * kColsPerGroup);
#pragma unroll 1
      for (int c0 = 0; c0 < kColsPerGroup; c0 += 32) {
        uint32_t v[32];
        tmem_ld_32x32(t_addr + (uint32_t)c0, v);
        tmem_ld_wait();
        if (row_ok) RowEpi<KIND, TA>::store(ep, row, n0 + c0, v, pre);
        if (RowEpi<KIND, TA>::kNeedsPrefetch && row_ok && c0 + 32 < kColsPerGroup)
          RowEpi<KIND, TA>::prefetch(ep, row, n0 + c0 + 32, pre);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty_bar[acc]);
      if (row_ok) RowEpi<KIND, TA>::finish(ep, row, pre);
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
