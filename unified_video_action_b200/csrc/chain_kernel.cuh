// Persistent denoise-chain kernel for the latency regime (predict_action, SURVEY.md 8d C1: N = 16 rows,
// up to 128): ONE cooperative launch runs every denoise step of the DDPM loop (GD:446-541).
//
// The chain is 100 steps x 14 dependent layers deep and each layer is tiny (16 x 1024 x 1024 MACs), so
// the time is launch / synchronisation latency, not math.  Design:
//   * 128 CTAs (one per SM), CTA c owns output features [8c, 8c+8) of EVERY W x W layer;
//   * weight-stationary: the CTA's 8 rows of all 2*depth MLP weight matrices (12 x 8 x 1024 bf16 = 192 KB)
//     are loaded into shared memory once and stay there for all steps -> no weight traffic in the loop;
//   * activations are exchanged through L2 (h fp32 [N,W], u bf16 [N,W]) with a grid-wide barrier after
//     every layer (release / acquire on a global counter; cooperative launch guarantees co-residency);
//   * per layer a CTA stages the full 16-row activation tile in shared memory (LayerNorm-modulate fused
//     into the staging of GEMM1's operand), 8 warps split K and run mma.sync.m16n8k16 (bf16, fp32
//     accumulate), partial sums are reduced through shared memory and the epilogue (SiLU / gated
//     residual) writes the CTA's 16 x 8 output patch;
//   * the AdaLN modulation (shift / scale / gate of every block and step) comes from the table that the
//     tcgen05 GEMM computed up front for all steps (head_small.inc);
//   * final layer: CTAs 0 .. 2C-1 compute one output column each; the DDPM update (CFG mix, learned-range
//     variance, x0-from-eps, clip, posterior mean, noise; GD:260-346, 395-444) is fused with the next
//     step's input_proj.
// tcgen05 needs M >= 64 rows per instruction and a TMEM round trip per layer; at 16 rows the legacy
// warp-level MMA has the lower latency, which is what bounds this path.
#pragma once
#include "common.cuh"
#include "elementwise.cuh"

namespace uva {

constexpr int CH_CTAS = 128;     // CTAs = W / CH_F
constexpr int CH_F = 8;          // output features per CTA and layer
constexpr int CH_W = 1024;       // model width this kernel is built for
constexpr int CH_THREADS = 256;  // 8 warps: K split 8 ways
constexpr int CH_MAX_DEPTH = 6;
constexpr int CH_MT = 16;        // rows per MMA tile
constexpr int CH_SMEM_W = 2 * CH_MAX_DEPTH * CH_F * CH_W * 2;  // 196608: stationary weights
constexpr int CH_SMEM_A = CH_MT * CH_W * 2;                    //  32768: activation tile (also scratch)
constexpr int CH_SMEM_BYTES = CH_SMEM_W + CH_SMEM_A;

struct ChainArgs {
  // weights (bf16 K-major [W, W]) and fp32 vectors per block
  const bf16* w1[CH_MAX_DEPTH];
  const bf16* w2[CH_MAX_DEPTH];
  const float* b1[CH_MAX_DEPTH];
  const float* b2[CH_MAX_DEPTH];
  const float* ln_w[CH_MAX_DEPTH];
  const float* ln_b[CH_MAX_DEPTH];
  const bf16* lin_w;   // [32, W] final linear (rows >= 2C are zero)
  const float* lin_b;  // [2C]
  const float* inp_w;  // [W, C] input_proj
  const float* inp_b;  // [W]
  const bf16* mod;     // [steps, N, nmod] modulation table of all steps (bias included)
  int64_t nmod;
  const float* x0;     // [N, C] start state
  const float* noise;  // [steps, N, C]
  float* xbuf[2];      // [N, C] chain state ping-pong (global, written by CTA 0)
  float* out;          // [N, C]
  float* h;            // [N, W] fp32 residual stream
  bf16* u;             // [N, W] MLP hidden
  float* netout;       // [N, 32]
  const StepCoef* coef;  // [T_gen] indexed by timestep index i
  unsigned int* counter; // grid barrier (zeroed before the launch)
  int N, C, depth, steps, k_begin, T_gen;
  int half;            // N / 2 with CFG, else 0
  float cfg, temperature;
  int clip;
  int hold;  // cycles between a CTA's barrier arrival and its first poll (UVA_CHAIN_HOLD, default 480)
  int dbg;  // UVA_CHAIN_DBG (timing diagnosis only, results are wrong): 1 = no grid barriers, 2 = no activation staging
};

__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                               uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// grid-wide barrier: every CTA arrives once per epoch on a monotonically increasing counter
__device__ __forceinline__ void chain_grid_sync(unsigned int* counter, unsigned int& epoch, int dbg = 0, int hold = 0) {
  if (dbg & 1) { __syncthreads(); return; }
  __syncthreads();
  if (threadIdx.x == 0) {
    epoch += 1;
    const unsigned int target = epoch * gridDim.x;
    // no separate fence: bar.sync orders the CTA's writes before this thread and the release is cumulative
    // (the pattern of CUTLASS' semaphore arrive); an extra MEMBAR.GPU cost 0.15 us per barrier
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
    if (hold > 0) {
      // the first poll is held back: polls queue on the counter's L2 line in front of the arrivals of slower CTAs, and
      // a poll issued before the barrier can possibly be complete only delays it (measured: flat optimum at 384 - 576 cycles,
      // predict_action p50 5.72 -> 5.58 ms; 768 cycles is already later than the typical completion)
      const long long t1 = clock64() + hold;
      while (clock64() < t1) {}
    }
    unsigned int v;
    uint32_t spins = 0;
    uint64_t t0 = 0;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
      if ((++spins & 0xFFFu) == 0) {  // a lost CTA becomes a trap instead of a hung GPU
        const uint64_t now = global_timer_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > 4000000000ull) __trap();
      }
    } while (v < target);
  }
  __syncthreads();
}

// shared-memory tiles: rows of 1024 bf16 = 512 words; word w of row r lives at w ^ ((r & 7) << 2), so the
// MMA fragment loads (8 rows x 4 consecutive words) and the 16-byte staging stores are conflict-free
__device__ __forceinline__ uint32_t ch_word(const uint32_t* tile, int row, int w) {
  return tile[row * (CH_W / 2) + (w ^ ((row & 7) << 2))];
}

// one 16 x 8 output patch: acc[row, n] = sum_k A[row, k] * Wt[n, k]; 8 warps split K, result reduced into
// red[16][8] (fp32, aliased on the A tile AFTER every warp has finished reading it)
__device__ __forceinline__ void ch_gemm_patch(const uint32_t* A, const uint32_t* Wt, float* red, int warp, int lane) {
  const int g = lane >> 2, t = lane & 3;
  float c[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int s = 0; s < CH_W / 16 / 8; ++s) {  // 8 k-steps of 16 per warp
    const int wbase = (warp * (CH_W / 16 / 8) + s) * 8;  // word offset of this k-step
    const uint32_t a0 = ch_word(A, g, wbase + t);
    const uint32_t a1 = ch_word(A, g + 8, wbase + t);
    const uint32_t a2 = ch_word(A, g, wbase + 4 + t);
    const uint32_t a3 = ch_word(A, g + 8, wbase + 4 + t);
    const uint32_t b0 = ch_word(Wt, g, wbase + t);
    const uint32_t b1 = ch_word(Wt, g, wbase + 4 + t);
    mma_bf16_16816(c, a0, a1, a2, a3, b0, b1);
  }
  __syncthreads();  // all warps are done with A
  float* part = red + warp * (CH_MT * CH_F);
  part[g * CH_F + 2 * t] = c[0];
  part[g * CH_F + 2 * t + 1] = c[1];
  part[(g + 8) * CH_F + 2 * t] = c[2];
  part[(g + 8) * CH_F + 2 * t + 1] = c[3];
  __syncthreads();
}
__device__ __forceinline__ float ch_reduced(const float* red, int idx) {
  float s = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) s += red[w * (CH_MT * CH_F) + idx];
  return s;
}

// stage a 16-row tile of a = LayerNorm(h) [* ln_w + ln_b] * (1 + scale) + shift as bf16 (DL:93-94, 163-166,
// 185-188); warp w takes rows 2w, 2w+1; rows >= N are zero
__device__ __forceinline__ void ch_stage_lnmod(uint32_t* A, const float* h, int m0, int N, const float* ln_w,
                                               const float* ln_b, const bf16* shift, const bf16* scale,
                                               int64_t ldmod, int warp, int lane) {
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const int rl = 2 * warp + rr;
    const int row = m0 + rl;
    uint32_t* dst = A + rl * (CH_W / 2);
    const int sw = (rl & 7) << 2;
    if (row >= N) {
#pragma unroll
      for (int i = 0; i < 8; ++i) *reinterpret_cast<uint2*>(dst + ((2 * (i * 32 + lane)) ^ sw)) = make_uint2(0u, 0u);
      continue;
    }
    const float4* hp = reinterpret_cast<const float4*>(h + (int64_t)row * CH_W);
    const uint2* shp = reinterpret_cast<const uint2*>(shift + (int64_t)row * ldmod);
    const uint2* scp = reinterpret_cast<const uint2*>(scale + (int64_t)row * ldmod);
    float4 v[8];
    uint2 shv[8], scv[8];  // issued together with the h loads: one L2 round trip for the whole row
    float4 gwv[8], gbv[8];  // ... and so is the LayerNorm affine (L1 is too small here to keep it across layers)
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = __ldcg(hp + i * 32 + lane);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      shv[i] = __ldg(shp + i * 32 + lane);
      scv[i] = __ldg(scp + i * 32 + lane);
    }
    if (ln_w != nullptr) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        gwv[i] = __ldg(reinterpret_cast<const float4*>(ln_w) + i * 32 + lane);
        gbv[i] = __ldg(reinterpret_cast<const float4*>(ln_b) + i * 32 + lane);
      }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
    const float mean = warp_sum(s) * (1.0f / CH_W);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
      q += (a * a + b * b) + (c * c + d * d);
    }
    const float rstd = rsqrtf(warp_sum(q) * (1.0f / CH_W) + 1e-6f);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int e = i * 32 + lane;  // float4 index inside the row
      float n0 = (v[i].x - mean) * rstd, n1 = (v[i].y - mean) * rstd, n2 = (v[i].z - mean) * rstd, n3 = (v[i].w - mean) * rstd;
      if (ln_w != nullptr) {
        const float4 gw = gwv[i], gb = gbv[i];
        n0 = fmaf(n0, gw.x, gb.x); n1 = fmaf(n1, gw.y, gb.y); n2 = fmaf(n2, gw.z, gb.z); n3 = fmaf(n3, gw.w, gb.w);
      }
      const uint2 sh = shv[i], sc = scv[i];
      const float2 sh01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sh.x));
      const float2 sh23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sh.y));
      const float2 sc01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sc.x));
      const float2 sc23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sc.y));
      const uint32_t p0 = pack_bf16x2(fmaf(n0, 1.0f + sc01.x, sh01.x), fmaf(n1, 1.0f + sc01.y, sh01.y));
      const uint32_t p1 = pack_bf16x2(fmaf(n2, 1.0f + sc23.x, sh23.x), fmaf(n3, 1.0f + sc23.y, sh23.y));
      *reinterpret_cast<uint2*>(dst + ((2 * e) ^ sw)) = make_uint2(p0, p1);
    }
  }
}

// stage a 16-row tile of a bf16 [N, W] activation written by other CTAs (L2 loads)
__device__ __forceinline__ void ch_stage_b16(uint32_t* A, const bf16* src, int m0, int N, int tid) {
#pragma unroll
  for (int i = 0; i < CH_MT * (CH_W / 8) / CH_THREADS; ++i) {
    const int qd = i * CH_THREADS + tid;
    const int rl = qd >> 7, c = qd & 127;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (m0 + rl < N) v = __ldcg(reinterpret_cast<const uint4*>(src + (int64_t)(m0 + rl) * CH_W) + c);
    *reinterpret_cast<uint4*>(A + rl * (CH_W / 2) + ((c ^ (rl & 7)) << 2)) = v;
  }
}

// ---- cluster-cooperative staging (CL CTAs of a cluster share the work of staging one activation tile) ----
// Every CTA needs the whole 16-row tile, so without clusters each of the 128 CTAs pulls all of it from L2
// (8 MB of L2 reads per layer) and normalises all 16 rows.  In a cluster of CL CTAs, CTA `rank` loads and
// normalises only rows [rank * 16/CL, (rank+1) * 16/CL) and writes the bf16 result into the tile of every CTA of
// the cluster through distributed shared memory; a cluster barrier then publishes the tile.
__device__ __forceinline__ uint32_t ch_mapa(uint32_t local_addr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_addr), "r"(cta));
  return r;
}
__device__ __forceinline__ void ch_st_cluster_v2(uint32_t addr, uint32_t x, uint32_t y) {
  asm volatile("st.shared::cluster.v2.u32 [%0], {%1, %2};" ::"r"(addr), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ void ch_st_cluster_v4(uint32_t addr, uint4 v) {
  asm volatile("st.shared::cluster.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// same arithmetic as ch_stage_lnmod (bit-identical tile), one warp per row, rows of this CTA only
template <int CL>
__device__ __forceinline__ void ch_stage_lnmod_cl(uint32_t* A, const float* h, int m0, int N, const float* ln_w,
                                                  const float* ln_b, const bf16* shift, const bf16* scale,
                                                  int64_t ldmod, int warp, int lane, uint32_t rank) {
  constexpr int RPC = CH_MT / CL;  // rows per CTA
  if (warp >= RPC) return;
  const int rl = (int)rank * RPC + warp;
  const int row = m0 + rl;
  const int sw = (rl & 7) << 2;
  uint32_t dst[CL];
  const uint32_t local = smem_u32(A + rl * (CH_W / 2));
#pragma unroll
  for (int c = 0; c < CL; ++c) dst[c] = ch_mapa(local, (uint32_t)c);
  if (row >= N) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
#pragma unroll
      for (int c = 0; c < CL; ++c) ch_st_cluster_v2(dst[c] + 4u * (uint32_t)((2 * (i * 32 + lane)) ^ sw), 0u, 0u);
    }
    return;
  }
  const float4* hp = reinterpret_cast<const float4*>(h + (int64_t)row * CH_W);
  const uint2* shp = reinterpret_cast<const uint2*>(shift + (int64_t)row * ldmod);
  const uint2* scp = reinterpret_cast<const uint2*>(scale + (int64_t)row * ldmod);
  float4 v[8];
  uint2 shv[8], scv[8];
  float4 gwv[8], gbv[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __ldcg(hp + i * 32 + lane);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    shv[i] = __ldg(shp + i * 32 + lane);
    scv[i] = __ldg(scp + i * 32 + lane);
  }
  if (ln_w != nullptr) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      gwv[i] = __ldg(reinterpret_cast<const float4*>(ln_w) + i * 32 + lane);
      gbv[i] = __ldg(reinterpret_cast<const float4*>(ln_b) + i * 32 + lane);
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  const float mean = warp_sum(s) * (1.0f / CH_W);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
    q += (a * a + b * b) + (c * c + d * d);
  }
  const float rstd = rsqrtf(warp_sum(q) * (1.0f / CH_W) + 1e-6f);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int e = i * 32 + lane;
    float n0 = (v[i].x - mean) * rstd, n1 = (v[i].y - mean) * rstd, n2 = (v[i].z - mean) * rstd, n3 = (v[i].w - mean) * rstd;
    if (ln_w != nullptr) {
      const float4 gw = gwv[i], gb = gbv[i];
      n0 = fmaf(n0, gw.x, gb.x); n1 = fmaf(n1, gw.y, gb.y); n2 = fmaf(n2, gw.z, gb.z); n3 = fmaf(n3, gw.w, gb.w);
    }
    const uint2 sh = shv[i], sc = scv[i];
    const float2 sh01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sh.x));
    const float2 sh23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sh.y));
    const float2 sc01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sc.x));
    const float2 sc23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&sc.y));
    const uint32_t p0 = pack_bf16x2(fmaf(n0, 1.0f + sc01.x, sh01.x), fmaf(n1, 1.0f + sc01.y, sh01.y));
    const uint32_t p1 = pack_bf16x2(fmaf(n2, 1.0f + sc23.x, sh23.x), fmaf(n3, 1.0f + sc23.y, sh23.y));
    const uint32_t off = 4u * (uint32_t)((2 * e) ^ sw);
#pragma unroll
    for (int c = 0; c < CL; ++c) ch_st_cluster_v2(dst[c] + off, p0, p1);
  }
}

// rows of this CTA of a bf16 [N, W] activation -> the tile of every CTA of the cluster
template <int CL>
__device__ __forceinline__ void ch_stage_b16_cl(uint32_t* A, const bf16* src, int m0, int N, int tid, uint32_t rank) {
  constexpr int RPC = CH_MT / CL;
  const uint32_t base = smem_u32(A);
#pragma unroll
  for (int i = 0; i < (RPC * (CH_W / 8) + CH_THREADS - 1) / CH_THREADS; ++i) {
    const int qd = i * CH_THREADS + tid;
    if (qd >= RPC * (CH_W / 8)) break;
    const int rl = (int)rank * RPC + (qd >> 7), c = qd & 127;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (m0 + rl < N) v = __ldcg(reinterpret_cast<const uint4*>(src + (int64_t)(m0 + rl) * CH_W) + c);
    const uint32_t off = 4u * (uint32_t)(rl * (CH_W / 2) + ((c ^ (rl & 7)) << 2));
#pragma unroll
    for (int d = 0; d < CL; ++d) ch_st_cluster_v4(ch_mapa(base + off, (uint32_t)d), v);
  }
}

template <int CL>
__global__ void __launch_bounds__(CH_THREADS, 1) k_denoise_chain(const __grid_constant__ ChainArgs p) {
  extern __shared__ __align__(16) uint8_t ch_smem[];
  uint32_t* Wsm = reinterpret_cast<uint32_t*>(ch_smem);              // [2*depth][8][512 words]
  uint32_t* A = reinterpret_cast<uint32_t*>(ch_smem + CH_SMEM_W);    // [16][512 words]
  float* red = reinterpret_cast<float*>(A);                          // [8 warps][16][8] (after the MMAs)
  float* xs = reinterpret_cast<float*>(A);                           // [N][C] (phase 0)
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int f0 = blockIdx.x * CH_F;
  const int N = p.N, C = p.C, W = CH_W;
  const int mtiles = (N + CH_MT - 1) / CH_MT;
  unsigned int epoch = 0;
  const uint32_t crank = (CL > 1) ? cluster_ctarank() : 0u;

  // stationary small vectors: this CTA's 8 entries of every MLP bias and its 8 rows of input_proj (no global round
  // trip in the epilogues: what bounds a layer is the number of dependent memory round trips, not bytes)
  __shared__ float s_b[2 * CH_MAX_DEPTH][CH_F];
  __shared__ float s_wi[CH_F][16];
  __shared__ float s_bi[CH_F];
  if (tid < 2 * p.depth * CH_F) {
    const int l = tid >> 3, n = tid & 7;
    s_b[l][n] = __ldg(((l & 1) ? p.b2[l >> 1] : p.b1[l >> 1]) + f0 + n);
  }
  if (tid < CH_F * 16) {
    const int n = tid >> 4, c = tid & 15;
    s_wi[n][c] = (c < C) ? __ldg(p.inp_w + (int64_t)(f0 + n) * C + c) : 0.f;
    if (c == 0) s_bi[n] = __ldg(p.inp_b + f0 + n);
  }
  // stationary weights: this CTA's 8 rows of every MLP matrix
  for (int l = 0; l < 2 * p.depth; ++l) {
    const bf16* src = ((l & 1) ? p.w2[l >> 1] : p.w1[l >> 1]) + (int64_t)f0 * W;
    uint32_t* dst = Wsm + l * (CH_F * (W / 2));
#pragma unroll
    for (int i = 0; i < CH_F * (CH_W / 8) / CH_THREADS; ++i) {
      const int qd = i * CH_THREADS + tid;
      const int n = qd >> 7, c = qd & 127;
      *reinterpret_cast<uint4*>(dst + n * (W / 2) + ((c ^ (n & 7)) << 2)) =
          __ldg(reinterpret_cast<const uint4*>(src + (int64_t)n * W) + c);
    }
  }
  __syncthreads();

  for (int j = 0; j <= p.steps; ++j) {
    // ---- phase 0: state x_j (DDPM update of step j-1, GD:395-444) + input_proj of step j (DL:268) ----
    {
      const float* xprev = (j <= 1) ? p.x0 : p.xbuf[(j - 1) & 1];
      StepCoef kc = {};
      if (j > 0) kc = p.coef[p.T_gen - 1 - (p.k_begin + j - 1)];
      for (int i = tid; i < N * C; i += CH_THREADS) {
        float xn;
        if (j == 0) {
          xn = p.x0[i];
        } else {
          const int r = i / C, c = i - r * C;
          float eps;
          if (p.half > 0) {
            const int rc = r % p.half;
            const float ce = __ldcg(p.netout + rc * 32 + c);
            const float ue = __ldcg(p.netout + (rc + p.half) * 32 + c);
            eps = ue + p.cfg * (ce - ue);
          } else {
            eps = __ldcg(p.netout + r * 32 + c);
          }
          const float v = __ldcg(p.netout + r * 32 + C + c);
          const float x = (j == 1) ? p.x0[i] : __ldcg(xprev + i);
          const float frac = (v + 1.0f) / 2.0f;
          const float logvar = frac * kc.max_log + (1.0f - frac) * kc.min_log;
          float xs0 = kc.sqrt_recip * x - kc.sqrt_recipm1 * eps;
          if (p.clip) xs0 = fminf(fmaxf(xs0, -1.0f), 1.0f);
          const float mean = kc.coef1 * xs0 + kc.coef2 * x;
          xn = mean;
          if (kc.nonzero) xn = mean + expf(0.5f * logvar) * p.noise[(int64_t)(j - 1) * N * C + i] * p.temperature;
        }
        xs[i] = xn;
        if (blockIdx.x == 0) {
          if (j == p.steps) p.out[i] = xn;
          else if (j > 0) p.xbuf[j & 1][i] = xn;
        }
      }
      if (j == p.steps) break;
      __syncthreads();
      for (int it = tid; it < N * CH_F; it += CH_THREADS) {
        const int r = it >> 3, n = it & 7;
        const int rs = (p.half > 0 && r >= p.half) ? r - p.half : r;
        float acc = s_bi[n];
        for (int c = 0; c < C; ++c) acc = fmaf(xs[rs * C + c], s_wi[n][c], acc);
        p.h[(int64_t)r * W + f0 + n] = acc;
      }
      chain_grid_sync(p.counter, epoch, p.dbg, p.hold);
    }
    const bf16* modj = p.mod + (int64_t)j * N * p.nmod;
    for (int b = 0; b < p.depth; ++b) {
      const bf16* shift = modj + (int64_t)3 * W * b;
      // ---- GEMM1: u = SiLU(LNmod(h) . W1^T + b1) (DL:163-166) ----
      for (int mt = 0; mt < mtiles; ++mt) {
        const int m0 = mt * CH_MT;
        if constexpr (CL > 1) {
          if (!(p.dbg & 2)) ch_stage_lnmod_cl<CL>(A, p.h, m0, N, p.ln_w[b], p.ln_b[b], shift, shift + W, p.nmod, warp, lane, crank);
          cluster_sync_all();  // every CTA of the cluster has written its rows into this CTA's tile
        } else {
          if (!(p.dbg & 2)) ch_stage_lnmod(A, p.h, m0, N, p.ln_w[b], p.ln_b[b], shift, shift + W, p.nmod, warp, lane);
          __syncthreads();
        }
        ch_gemm_patch(A, Wsm + (2 * b) * (CH_F * (W / 2)), red, warp, lane);
        if (tid < CH_MT * CH_F) {
          const int rl = tid >> 3, n = tid & 7;
          if (m0 + rl < N) {
            const float v = ch_reduced(red, tid) + s_b[2 * b][n];
            p.u[(int64_t)(m0 + rl) * W + f0 + n] = __float2bfloat16_rn(silu_f(v));
          }
        }
        // the tile doubles as the reduction scratch: peers may only overwrite it once this CTA is done with it
        if (CL > 1 && mt + 1 < mtiles) cluster_sync_all(); else __syncthreads();
      }
      chain_grid_sync(p.counter, epoch, p.dbg, p.hold);
      // ---- GEMM2: h += gate * (u . W2^T + b2) (DL:167) ----
      for (int mt = 0; mt < mtiles; ++mt) {
        const int m0 = mt * CH_MT;
        // the epilogue's own operands (gate from the modulation table, the residual element this thread wrote itself
        // one block earlier) are requested before the staging, so they arrive under it
        float gate = 0.f, hres = 0.f;
        if (tid < CH_MT * CH_F && m0 + (tid >> 3) < N) {
          gate = __bfloat162float(shift[(int64_t)(m0 + (tid >> 3)) * p.nmod + 2 * W + f0 + (tid & 7)]);
          hres = __ldcg(p.h + (int64_t)(m0 + (tid >> 3)) * W + f0 + (tid & 7));
        }
        if constexpr (CL > 1) {
          if (!(p.dbg & 2)) ch_stage_b16_cl<CL>(A, p.u, m0, N, tid, crank);
          cluster_sync_all();
        } else {
          if (!(p.dbg & 2)) ch_stage_b16(A, p.u, m0, N, tid);
          __syncthreads();
        }
        ch_gemm_patch(A, Wsm + (2 * b + 1) * (CH_F * (W / 2)), red, warp, lane);
        if (tid < CH_MT * CH_F) {
          const int rl = tid >> 3, n = tid & 7;
          if (m0 + rl < N) {
            const float v = ch_reduced(red, tid) + s_b[2 * b + 1][n];
            p.h[(int64_t)(m0 + rl) * W + f0 + n] = fmaf(gate, v, hres);
          }
        }
        if (CL > 1 && mt + 1 < mtiles) cluster_sync_all(); else __syncthreads();
      }
      chain_grid_sync(p.counter, epoch, p.dbg, p.hold);
    }
    // ---- final layer (DL:185-189): CTA o < 2C computes output column o ----
    if ((int)blockIdx.x < 2 * C) {
      const bf16* fshift = modj + (int64_t)3 * W * p.depth;
      const uint32_t* lw = reinterpret_cast<const uint32_t*>(p.lin_w + (int64_t)blockIdx.x * W);
      uint32_t wreg[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) wreg[i] = __ldg(lw + lane + 32 * i);
      const float lb = __ldg(p.lin_b + blockIdx.x);
      for (int mt = 0; mt < mtiles; ++mt) {
        const int m0 = mt * CH_MT;
        ch_stage_lnmod(A, p.h, m0, N, nullptr, nullptr, fshift, fshift + W, p.nmod, warp, lane);
        __syncwarp();  // each warp reads back only the two rows it staged
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
          const int rl = 2 * warp + rr;
          float s = 0.f;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const uint32_t aw = ch_word(A, rl, lane + 32 * i);
            const float2 af = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&aw));
            const float2 wf = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&wreg[i]));
            s = fmaf(af.x, wf.x, s);
            s = fmaf(af.y, wf.y, s);
          }
          s = warp_sum(s);
          if (lane == 0 && m0 + rl < N) p.netout[(m0 + rl) * 32 + blockIdx.x] = s + lb;
        }
        __syncwarp();
      }
    }
    chain_grid_sync(p.counter, epoch, p.dbg, p.hold);
  }
}

}  // namespace uva
