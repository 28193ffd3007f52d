// Backward-pass row kernels and reductions of the diffusion head (autograd of DL:261-283 written out
// by hand).  TA = activation type of the arithmetic mode (float / bf16).
#pragma once
#include "common.cuh"

namespace uva {

// accurate SiLU derivative (fp32 time-embedding table path)
__device__ __forceinline__ float dsilu_acc(float x) {
  const float s = 1.0f / (1.0f + expf(-x));
  return s * (1.0f + x * (1.0f - s));
}

template <typename TA>
__device__ __forceinline__ void load4(const TA* p, float (&v)[4]);
template <>
__device__ __forceinline__ void load4<float>(const float* p, float (&v)[4]) {
  const float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
template <>
__device__ __forceinline__ void load4<bf16>(const bf16* p, float (&v)[4]) {
  const uint2 t = *reinterpret_cast<const uint2*>(p);
  const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&t.x);
  const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&t.y);
  v[0] = __low2float(a); v[1] = __high2float(a); v[2] = __low2float(b); v[3] = __high2float(b);
}
__device__ __forceinline__ void store4(float* p, const float (&v)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void store4(bf16* p, const float (&v)[4]) {
  *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]));
}

// ------------------------------------------------------------------------------------------
// FinalLayer.linear backward (DL:189): out[N,2C] = a_f[N,W] . Wl[2C,W]^T + bl, 2C <= 32: CUDA cores.
//   da[r, j]  = sum_c dnet[r, c] Wl[c, j]                      (k_final_dgrad)
//   dWl[c, j] += sum_r dnet[r, c] a_f[r, j],  dbl[c] += sum_r dnet[r, c]   (k_final_wgrad)
// ------------------------------------------------------------------------------------------
template <typename TA>
__global__ void k_final_dgrad(const float* __restrict__ dnet, int ldn, const float* __restrict__ Wl,
                              TA* __restrict__ da, int64_t rows, int W, int C2) {
  const int wq = W / 4;
  const int64_t n = rows * wq;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / wq;
    const int j = (int)(i - r * wq) * 4;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c = 0; c < C2; ++c) {
      const float d = dnet[r * ldn + c];
      const float4 w = __ldg(reinterpret_cast<const float4*>(Wl + (int64_t)c * W + j));
      acc[0] = fmaf(d, w.x, acc[0]); acc[1] = fmaf(d, w.y, acc[1]);
      acc[2] = fmaf(d, w.z, acc[2]); acc[3] = fmaf(d, w.w, acc[3]);
    }
    store4(da + r * W + j, acc);
  }
}

constexpr int FW_ROWS = 128;  // rows per block of the skinny weight-gradient kernels
constexpr int IW_ROWS = 256;  // most rows a block of k_inproj_wgrad takes (its x_t rows are staged in shared memory)
constexpr int IW_STAGES = 3;
constexpr int kIwRingBytes = IW_STAGES * 4 * 256 * 16;
template <typename TA>
__global__ void __launch_bounds__(256)
k_final_wgrad(const float* __restrict__ dnet, int ldn, const TA* __restrict__ a, float* __restrict__ dWl,
              float* __restrict__ dbl, int64_t rows, int W, int C2) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t r0 = (int64_t)blockIdx.y * FW_ROWS;
  const int64_t r1 = (r0 + FW_ROWS < rows) ? r0 + FW_ROWS : rows;
  float acc[32];
#pragma unroll
  for (int c = 0; c < 32; ++c) acc[c] = 0.f;
  if (j < W) {
    if ((ldn & 3) == 0) {  // vector loads of the dnet row (warp-uniform address: one 16-byte broadcast per 4 FMAs)
#pragma unroll 2
      for (int64_t r = r0; r < r1; ++r) {
        const float av = to_f32(a[r * W + j]);
        const float4* dp4 = reinterpret_cast<const float4*>(dnet + r * ldn);
#pragma unroll
        for (int c4 = 0; c4 < 8; ++c4) {
          if (4 * c4 < C2) {
            const float4 d = __ldg(dp4 + c4);
            acc[4 * c4] = fmaf(d.x, av, acc[4 * c4]);
            acc[4 * c4 + 1] = fmaf(d.y, av, acc[4 * c4 + 1]);
            acc[4 * c4 + 2] = fmaf(d.z, av, acc[4 * c4 + 2]);
            acc[4 * c4 + 3] = fmaf(d.w, av, acc[4 * c4 + 3]);
          }
        }
      }
    } else {
      for (int64_t r = r0; r < r1; ++r) {
        const float av = to_f32(a[r * W + j]);
#pragma unroll
        for (int c = 0; c < 32; ++c)
          if (c < C2) acc[c] = fmaf(dnet[r * ldn + c], av, acc[c]);
      }
    }
#pragma unroll
    for (int c = 0; c < 32; ++c)
      if (c < C2 && dWl) atomicAdd(dWl + (int64_t)c * W + j, acc[c]);
  }
  if (blockIdx.x == 0 && threadIdx.x < C2 && dbl) {
    float s = 0.f;
    for (int64_t r = r0; r < r1; ++r) s += dnet[r * ldn + threadIdx.x];
    atomicAdd(dbl + threadIdx.x, s);
  }
}

// rows per block of k_inproj_wgrad: one wave of ~2 blocks per SM over the 4 x gridDim.y grid, a multiple of 16, <= IW_ROWS
inline int iw_rows_per_block(int64_t rows, int W, int num_sms) {
  const int gx = (W + 255) / 256;
  int64_t gy = (2 * num_sms + gx - 1) / gx;
  int64_t rpb = (rows + gy - 1) / gy;
  rpb = (rpb + 15) / 16 * 16;
  if (rpb < 16) rpb = 16;
  if (rpb > IW_ROWS) rpb = IW_ROWS;
  return (int)rpb;
}

// ------------------------------------------------------------------------------------------
// input_proj backward (DL:268): h0 = x_t . Wi^T + bi with K = C <= 16:
//   dWi[j, c] += sum_r dh[r, j] x_t[r, c],  dbi[j] += sum_r dh[r, j].
// ------------------------------------------------------------------------------------------
// A block takes 256 columns (64 threads x 4 adjacent columns, 16-byte loads) x rows_per_block rows (4 row lanes of 64
// threads; iw_rows_per_block() sizes the grid for one wave of two blocks per SM); the 4 lanes are combined in shared memory and the block adds its [256, C] patch with coalesced
// reductions (the patch is contiguous in dWi).  W % 4 == 0.
__global__ void __launch_bounds__(256)
k_inproj_wgrad(const float* __restrict__ dh, const float* __restrict__ xt, float* __restrict__ dWi,
               float* __restrict__ dbi, int64_t rows, int W, int C, int rows_per_block) {
  __shared__ float patch[256 * 17];  // [column][16 c + bias], combined over the row lanes
  __shared__ __align__(16) float xs[IW_ROWS * 16];  // the block's x_t rows, zero padded to 16 channels
  extern __shared__ float4 ring[];  // [IW_STAGES][4][256]: kIwRingBytes of dynamic shared memory
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  const int j = blockIdx.x * 256 + tx * 4;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_block;  // rows_per_block <= IW_ROWS
  const int64_t r1 = (r0 + rows_per_block < rows) ? r0 + rows_per_block : rows;
  for (int e = threadIdx.x; e < rows_per_block * 16; e += 256) {
    const int rr = e >> 4, c = e & 15;
    xs[e] = (r0 + rr < r1 && c < C) ? xt[(r0 + rr) * C + c] : 0.f;
  }
  __syncthreads();
  float acc[4][16];
  float sb[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int c = 0; c < 16; ++c) acc[k][c] = 0.f;
  if (j < W) {
    // d h streams through a ring of IW_STAGES thread-private stages (4 rows x 16 bytes per thread and stage, cp.async):
    // the loads of the next two stages are in flight while one is consumed, without holding registers for them
    const int n_it = (int)((r1 - r0 + 15) / 16);
    auto issue = [&](int it) {
      if (it < n_it) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int64_t ru = r0 + ty + 16 * it + 4 * u;
          const bool on = ru < r1;
          const float* src = dh + (on ? ru : r0) * W + j;
          const uint32_t dst = smem_u32(&ring[((it % IW_STAGES) * 4 + u) * 256 + threadIdx.x]);
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(on ? 16 : 0) : "memory");
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (int it = 0; it < IW_STAGES - 1; ++it) issue(it);
    for (int it = 0; it < n_it; ++it) {
      issue(it + IW_STAGES - 1);
      asm volatile("cp.async.wait_group %0;" ::"n"(IW_STAGES - 1) : "memory");
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float4 d = ring[((it % IW_STAGES) * 4 + u) * 256 + threadIdx.x];
        const int rr = ty + 16 * it + 4 * u;  // (rows past the end were zero-filled)
        const float4* xp = reinterpret_cast<const float4*>(xs + (rr < IW_ROWS ? rr : 0) * 16);
        const float dv[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
        for (int c4 = 0; c4 < 4; ++c4) {
          const float4 x = xp[c4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            acc[k][4 * c4] = fmaf(dv[k], x.x, acc[k][4 * c4]);
            acc[k][4 * c4 + 1] = fmaf(dv[k], x.y, acc[k][4 * c4 + 1]);
            acc[k][4 * c4 + 2] = fmaf(dv[k], x.z, acc[k][4 * c4 + 2]);
            acc[k][4 * c4 + 3] = fmaf(dv[k], x.w, acc[k][4 * c4 + 3]);
          }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) sb[k] += dv[k];
      }
    }
  }
  // combine the 4 row lanes: lane 0 stores, the others add in turn
  for (int turn = 0; turn < 4; ++turn) {
    if (ty == turn) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float* pr = patch + (tx * 4 + k) * 17;
#pragma unroll
        for (int c = 0; c < 16; ++c) pr[c] = (turn ? pr[c] : 0.f) + acc[k][c];
        pr[16] = (turn ? pr[16] : 0.f) + sb[k];
      }
    }
    __syncthreads();
  }
  const int j0 = blockIdx.x * 256;
  if (dWi) {
    for (int e = threadIdx.x; e < 256 * C; e += 256) {
      const int col = e / C, c = e - col * C;
      if (j0 + col < W) atomicAdd(dWi + (int64_t)j0 * C + e, patch[col * 17 + c]);
    }
  }
  if (dbi && j0 + (int)threadIdx.x < W) atomicAdd(dbi + j0 + threadIdx.x, patch[threadIdx.x * 17 + 16]);
}

// ------------------------------------------------------------------------------------------
// Gated residual backward (DL:167: h' = h + gate * q):  dq = dh * gate, dgate = dh * q.
// dh itself passes through unchanged (residual).  gate / dgate live in the modulation buffers.
// ------------------------------------------------------------------------------------------
// sum_dq / sum_dgate (may be null; only when the grid stride is a multiple of W / 4, i.e. every thread keeps its
// four columns): the bias gradients  db2 += sum_r dq,  d(adaLN bias, gate part) += sum_r dgate  accumulated in
// registers over the thread's rows and added with one atomic per column and thread — the two separate
// column-sum passes over dq and dgate disappear.
// The block is (blockDim.x / (W / 4)) row lanes of W / 4 threads: the lanes are combined in shared memory before the
// atomics, so that an address sees one reduction per BLOCK - the end-of-kernel atomics of 592 blocks x 2 W columns on
// 64 L2 sectors cost 20 us of a 57 us kernel (measured, scripts/rowkernel_bench.cu); 148 blocks of 4 lanes: 5 us.
template <typename TA>
__global__ void __launch_bounds__(1024)
k_gate_bwd(const float* __restrict__ dh, const TA* __restrict__ gate, const TA* __restrict__ q,
           int64_t ldmod, TA* __restrict__ dq, TA* __restrict__ dgate, int64_t rows, int W,
           float* __restrict__ sum_dq = nullptr, float* __restrict__ sum_dgate = nullptr) {
  const int wq = W / 4;
  const int64_t n = rows * wq;
  float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
  int jcol = -1;
  if ((int)blockDim.x % wq == 0) {
    // row-lane form: every thread keeps its 4 columns; 2 rows per iteration with all 6 loads requested before the first
    // use (the kernel is bound by bytes in flight per SM, not by arithmetic)
    __shared__ float comb[2 * 1024];  // (W <= 1024 in this form: blockDim.x <= 1024)
    constexpr int U = 2;
    const int lanes = (int)blockDim.x / wq;
    const int tx = (int)threadIdx.x % wq, ty = (int)threadIdx.x / wq;
    const int j = tx * 4;
    const int64_t step = (int64_t)gridDim.x * lanes;
    for (int64_t r = (int64_t)blockIdx.x * lanes + ty; r < rows; r += U * step) {
      float4 d[U];
      float g[U][4], qq[U][4];
      bool on[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int64_t ru = r + u * step;
        on[u] = ru < rows;
        if (on[u]) {
          d[u] = *reinterpret_cast<const float4*>(dh + ru * W + j);
          load4<TA>(gate + ru * ldmod + j, g[u]);
          load4<TA>(q + ru * W + j, qq[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (on[u]) {
          const int64_t ru = r + u * step;
          const float o1[4] = {d[u].x * g[u][0], d[u].y * g[u][1], d[u].z * g[u][2], d[u].w * g[u][3]};
          const float o2[4] = {d[u].x * qq[u][0], d[u].y * qq[u][1], d[u].z * qq[u][2], d[u].w * qq[u][3]};
          store4(dq + ru * W + j, o1);
          store4(dgate + ru * ldmod + j, o2);
#pragma unroll
          for (int k = 0; k < 4; ++k) { s1[k] += o1[k]; s2[k] += o2[k]; }
        }
      }
    }
    if (sum_dq == nullptr && sum_dgate == nullptr) return;  // uniform
    for (int turn = 0; turn < lanes; ++turn) {
      if (ty == turn) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          comb[j + k] = (turn ? comb[j + k] : 0.f) + s1[k];
          comb[W + j + k] = (turn ? comb[W + j + k] : 0.f) + s2[k];
        }
      }
      __syncthreads();
    }
    for (int c = threadIdx.x; c < 2 * W; c += blockDim.x) {
      float* dst = (c < W) ? sum_dq : sum_dgate;
      if (dst) atomicAdd(dst + (c < W ? c : c - W), comb[c]);
    }
    return;
  }
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / wq;
    const int j = (int)(i - r * wq) * 4;
    jcol = j;
    const float4 d = *reinterpret_cast<const float4*>(dh + r * W + j);
    float g[4], qq[4];
    load4<TA>(gate + r * ldmod + j, g);
    load4<TA>(q + r * W + j, qq);
    const float o1[4] = {d.x * g[0], d.y * g[1], d.z * g[2], d.w * g[3]};
    const float o2[4] = {d.x * qq[0], d.y * qq[1], d.z * qq[2], d.w * qq[3]};
    store4(dq + r * W + j, o1);
    store4(dgate + r * ldmod + j, o2);
  }
  (void)jcol; (void)s1; (void)s2;  // (the generic form has no fixed columns per thread: the caller sums separately)
}

// ------------------------------------------------------------------------------------------
// Backward of a = LN(h)[* ln_w + ln_b] * (1 + scale) + shift  (modulate DL:93-94 on in_ln / norm_final):
//   dshift = da, dscale = da * nhat, dnhat = da * (1 + scale), dln_w += sum_r dnhat * xhat,
//   dln_b += sum_r dnhat, dxhat = dnhat * ln_w, dh_ln = rstd (dxhat - mean(dxhat) - xhat mean(dxhat xhat)).
// dh = (first ? 0 : dh) + dh_ln.  One warp per row, W = NV * 128; the column sums for ln_w / ln_b are
// kept in registers over the rows of a warp, combined through shared memory and added atomically.
// ------------------------------------------------------------------------------------------
// A row is split over P warps (a lane holds 32 / P columns).  The first form of this kernel kept a row per warp: ncu
// showed 255 registers per thread, 8 resident warps per SM (12 % of the maximum), dram at 50 %: every warp walked a
// chain of ~1500 dependent-issue instructions per row (82 us + a 19 us column-sum pass at 16 384 rows; an L2 prefetch of
// the next row only made it slower).  With P = 4 a thread holds 8 columns (114 registers, 2 blocks = 16 warps per SM),
// the per-row chain is a quarter as long, the row sums are combined through shared memory and a named barrier of the
// row's P warps: 59 us with all four column sums (scripts/rowkernel_bench.cu; P = 8 at 64 registers: 62 us).
// dsum ([2 W], may be null): bias gradient of the shift | scale parts of this modulation = column sums of dshift |
// dscale.  All four column sums (ln_w, ln_b, shift, scale) are register accumulators, combined per block at the end.
// GATE: the gated-residual backward of the NEXT block down the stack (k_gate_bwd: dq = dh * gate, dgate = dh * q, and their
// column sums) rides along: the d h row this kernel has just produced is still in registers, so the separate kernel's
// re-read of it (4 of its 12 bytes per element) and its launch disappear.
template <typename TA>
struct GateNext {
  const TA* gate;    // [rows, ldmod] gate third of the next block's modulation
  const TA* q;       // [rows, W] that block's mlp output before the gate
  TA* dq;            // [rows, W]
  TA* dgate;         // [rows, ldmod]
  float* sum_dq;     // [W] += column sums of dq (mlp.2 bias gradient), may be null
  float* sum_dgate;  // [W] += column sums of dgate (gate third of the adaLN bias gradient), may be null
};

template <typename TA, int NV, int P, bool GATE = false>
__global__ void __launch_bounds__(256, (P >= 8 ? (GATE ? 3 : 4) : 2))
k_lnmod_bwd_split(const TA* __restrict__ da, const float* __restrict__ h, const float* __restrict__ stats,
                  const float* __restrict__ ln_w, const float* __restrict__ ln_b, const TA* __restrict__ scale,
                  int64_t ldmod, TA* __restrict__ dshift, TA* __restrict__ dscale, float* __restrict__ dh, int first,
                  float* __restrict__ dln_w, float* __restrict__ dln_b, int64_t rows, float* __restrict__ dsum,
                  GateNext<TA> gn = GateNext<TA>{}) {
  constexpr int W = NV * 128;
  constexpr int NC = NV / P;        // 128-column chunks per warp
  constexpr int RPB = 8 / P;        // rows per block iteration
  static_assert(NV % P == 0 && 8 % P == 0, "row split");
  __shared__ float2 part[2][RPB][P];
  __shared__ float red[8][NC * 128];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int slot = wib / P, p = wib % P;
  float aw[NC][4], ab[NC][4], as[NC][4], ac[NC][4];
  float aq[GATE ? NC : 1][4], ag[GATE ? NC : 1][4];
#pragma unroll
  for (int i = 0; i < NC; ++i)
#pragma unroll
    for (int q = 0; q < 4; ++q) aw[i][q] = ab[i][q] = as[i][q] = ac[i][q] = 0.f;
#pragma unroll
  for (int i = 0; i < (GATE ? NC : 1); ++i)
#pragma unroll
    for (int q = 0; q < 4; ++q) aq[i][q] = ag[i][q] = 0.f;
  const int64_t nblk = (rows + RPB - 1) / RPB;  // block iterations (uniform over the block: the barrier needs all P warps)
  int par = 0;
  for (int64_t it = blockIdx.x; it < nblk; it += gridDim.x, par ^= 1) {
    const int64_t r = it * RPB + slot;
    const bool on = r < rows;
    float xh[NC][4], dx[NC][4];
    float4 prev[NC];
    float gt[GATE ? NC : 1][4], qn[GATE ? NC : 1][4];
    float s1 = 0.f, s2 = 0.f, rstd = 0.f;
    if (on) {
      const float mean = stats[2 * r];
      rstd = stats[2 * r + 1];
      if constexpr (GATE) {  // requested with the other operands of the row
#pragma unroll
        for (int i = 0; i < NC; ++i) {
          const int j = (lane + 32 * (i * P + p)) * 4;
          load4<TA>(gn.gate + r * ldmod + j, gt[i]);
          load4<TA>(gn.q + r * W + j, qn[i]);
        }
      }
      if (!first) {
#pragma unroll
        for (int i = 0; i < NC; ++i) prev[i] = *reinterpret_cast<const float4*>(dh + r * W + (lane + 32 * (i * P + p)) * 4);
      }
#pragma unroll
      for (int i = 0; i < NC; ++i) {
        const int j = (lane + 32 * (i * P + p)) * 4;
        const float4 hv = *reinterpret_cast<const float4*>(h + r * W + j);
        float d[4], sc[4];
        load4<TA>(da + r * W + j, d);
        load4<TA>(scale + r * ldmod + j, sc);
        xh[i][0] = (hv.x - mean) * rstd; xh[i][1] = (hv.y - mean) * rstd;
        xh[i][2] = (hv.z - mean) * rstd; xh[i][3] = (hv.w - mean) * rstd;
        float lw[4] = {1.f, 1.f, 1.f, 1.f}, lb[4] = {0.f, 0.f, 0.f, 0.f};
        if (ln_w) {
          const float4 g = __ldg(reinterpret_cast<const float4*>(ln_w + j));
          const float4 b = __ldg(reinterpret_cast<const float4*>(ln_b + j));
          lw[0] = g.x; lw[1] = g.y; lw[2] = g.z; lw[3] = g.w;
          lb[0] = b.x; lb[1] = b.y; lb[2] = b.z; lb[3] = b.w;
        }
        float dsc[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float nh = xh[i][q] * lw[q] + lb[q];
          dsc[q] = d[q] * nh;
          const float dn = d[q] * (1.0f + sc[q]);
          aw[i][q] += dn * xh[i][q];
          ab[i][q] += dn;
          as[i][q] += d[q];
          ac[i][q] += dsc[q];
          dx[i][q] = dn * lw[q];
          s1 += dx[i][q];
          s2 += dx[i][q] * xh[i][q];
        }
        store4(dshift + r * ldmod + j, d);
        store4(dscale + r * ldmod + j, dsc);
      }
    }
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    if (lane == 0) part[par][slot][p] = make_float2(s1, s2);
    asm volatile("bar.sync %0, %1;" ::"r"(1 + slot), "r"(32 * P) : "memory");
    s1 = 0.f; s2 = 0.f;
#pragma unroll
    for (int k = 0; k < P; ++k) { const float2 v = part[par][slot][k]; s1 += v.x; s2 += v.y; }
    s1 *= (1.0f / W);
    s2 *= (1.0f / W);
    if (on) {
#pragma unroll
      for (int i = 0; i < NC; ++i) {
        const int j = (lane + 32 * (i * P + p)) * 4;
        float o[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) o[q] = rstd * (dx[i][q] - s1 - xh[i][q] * s2);
        if (!first) {
          const float4 pv = prev[i];
          o[0] += pv.x; o[1] += pv.y; o[2] += pv.z; o[3] += pv.w;
        }
        store4(dh + r * W + j, o);
        if constexpr (GATE) {
          const float o1[4] = {o[0] * gt[i][0], o[1] * gt[i][1], o[2] * gt[i][2], o[3] * gt[i][3]};
          const float o2[4] = {o[0] * qn[i][0], o[1] * qn[i][1], o[2] * qn[i][2], o[3] * qn[i][3]};
          store4(gn.dq + r * W + j, o1);
          store4(gn.dgate + r * ldmod + j, o2);
#pragma unroll
          for (int q = 0; q < 4; ++q) { aq[i][q] += o1[q]; ag[i][q] += o2[q]; }
        }
      }
    }
  }
  // column sums: the RPB row slots that hold the same columns are combined in shared memory, one reduction per column
  // and block.  which: 0 ln_w, 1 ln_b, 2 shift bias, 3 scale bias
  for (int which = 0; which < (GATE ? 6 : 4); ++which) {
    float* dst = which == 0 ? dln_w : which == 1 ? dln_b : which == 2 ? dsum : which == 3 ? (dsum ? dsum + W : nullptr)
                 : which == 4 ? gn.sum_dq : gn.sum_dgate;
    if (dst == nullptr) continue;  // uniform over the grid
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NC; ++i)
#pragma unroll
      for (int q = 0; q < 4; ++q)
        red[wib][i * 128 + lane * 4 + q] = which == 0 ? aw[i][q] : which == 1 ? ab[i][q] : which == 2 ? as[i][q] : which == 3 ? ac[i][q]
                                           : which == 4 ? aq[GATE ? i : 0][q] : ag[GATE ? i : 0][q];
    __syncthreads();
    for (int c = threadIdx.x; c < W; c += blockDim.x) {
      // column c = (l + 32 (i P + pp)) 4 + q  ->  chunk = c / 128 = i P + pp, offset in chunk = c % 128
      const int chunk = c >> 7, off = c & 127;
      const int pp = chunk % P, i = chunk / P;
      float t = 0.f;
#pragma unroll
      for (int sl = 0; sl < RPB; ++sl) t += red[sl * P + pp][i * 128 + off];
      atomicAdd(dst + c, t);
    }
  }
}

// generic width (test geometries): one warp per row, no register caching, column sums by atomics
template <typename TA>
__global__ void __launch_bounds__(256)
k_lnmod_bwd_generic(const TA* __restrict__ da, const float* __restrict__ h, const float* __restrict__ stats,
                    const float* __restrict__ ln_w, const float* __restrict__ ln_b, const TA* __restrict__ scale,
                    int64_t ldmod, TA* __restrict__ dshift, TA* __restrict__ dscale, float* __restrict__ dh,
                    int first, float* __restrict__ dln_w, float* __restrict__ dln_b, int64_t rows, int W) {
  const int lane = threadIdx.x & 31;
  const int64_t warp_global = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int64_t warps_total = ((int64_t)gridDim.x * blockDim.x) >> 5;
  for (int64_t r = warp_global; r < rows; r += warps_total) {
    const float mean = stats[2 * r], rstd = stats[2 * r + 1];
    float s1 = 0.f, s2 = 0.f;
    for (int j = lane; j < W; j += 32) {
      const float xh = (h[r * W + j] - mean) * rstd;
      const float d = to_f32(da[r * W + j]);
      const float lw = ln_w ? ln_w[j] : 1.f, lb = ln_w ? ln_b[j] : 0.f;
      const float dn = d * (1.0f + to_f32(scale[r * ldmod + j]));
      dshift[r * ldmod + j] = from_f32<TA>(d);
      dscale[r * ldmod + j] = from_f32<TA>(d * (xh * lw + lb));
      if (dln_w) {
        atomicAdd(dln_w + j, dn * xh);
        atomicAdd(dln_b + j, dn);
      }
      s1 += dn * lw;
      s2 += dn * lw * xh;
    }
    s1 = warp_sum(s1) / (float)W;
    s2 = warp_sum(s2) / (float)W;
    for (int j = lane; j < W; j += 32) {
      const float xh = (h[r * W + j] - mean) * rstd;
      const float lw = ln_w ? ln_w[j] : 1.f;
      const float dn = to_f32(da[r * W + j]) * (1.0f + to_f32(scale[r * ldmod + j]));
      const float o = rstd * (dn * lw - s1 - xh * s2);
      dh[r * W + j] = (first ? 0.f : dh[r * W + j]) + o;
    }
  }
}

// ------------------------------------------------------------------------------------------
// Column sums (bias gradients): out[c] += sum_r X[r, c].  A block covers CS_ROWS rows x (32 lanes x
// 16 bytes) columns; warp w takes rows w, w + 8, ... with 8 independent 16-byte loads in flight per lane,
// the 8 warps are combined through shared memory and added with one atomic per column.
// ------------------------------------------------------------------------------------------
constexpr int CS_ROWS = 128;
template <typename TI>
__global__ void __launch_bounds__(256)
k_colsum(const TI* __restrict__ X, int64_t ld, int64_t rows, int cols, float* __restrict__ out) {
  constexpr int EPL = 16 / (int)sizeof(TI);  // elements per lane
  __shared__ float red[8][32 * EPL + 4];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int c0 = (blockIdx.x * 32 + lane) * EPL;
  const int64_t r0 = (int64_t)blockIdx.y * CS_ROWS;
  const int64_t r1 = (r0 + CS_ROWS < rows) ? r0 + CS_ROWS : rows;
  float s[EPL];
#pragma unroll
  for (int q = 0; q < EPL; ++q) s[q] = 0.f;
  if (c0 + EPL <= cols) {
    const TI* base = X + c0;
    int64_t r = r0 + w;
    for (; r + 56 < r1; r += 64) {
      uint4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = *reinterpret_cast<const uint4*>(base + (r + 8 * u) * ld);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if constexpr (sizeof(TI) == 2) {
          const __nv_bfloat162* p2 = reinterpret_cast<const __nv_bfloat162*>(&v[u]);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float2 f = __bfloat1622float2(p2[q]);
            s[2 * q] += f.x;
            s[2 * q + 1] += f.y;
          }
        } else {
          const float* pf = reinterpret_cast<const float*>(&v[u]);
#pragma unroll
          for (int q = 0; q < EPL; ++q) s[q] += pf[q];
        }
      }
    }
    for (; r < r1; r += 8) {
#pragma unroll
      for (int q = 0; q < EPL; ++q) s[q] += to_f32(base[r * ld + q]);
    }
  } else if (c0 < cols) {  // ragged column tail
    for (int64_t r = r0 + w; r < r1; r += 8)
      for (int q = 0; q < EPL && c0 + q < cols; ++q) s[q] += to_f32(X[r * ld + c0 + q]);
  }
#pragma unroll
  for (int q = 0; q < EPL; ++q) red[w][lane * EPL + q] = s[q];
  __syncthreads();
  for (int i = threadIdx.x; i < 32 * EPL; i += 256) {
    const int c = blockIdx.x * 32 * EPL + i;
    if (c < cols) {
      float t = 0.f;
#pragma unroll
      for (int w8 = 0; w8 < 8; ++w8) t += red[w8][i];
      atomicAdd(out + c, t);
    }
  }
}

// dtemb[t[r], :] += dy[r, :]   (gradient of the time-embedding table rows gathered in y = c + temb[t])
template <typename TA>
__global__ void k_temb_scatter(const TA* __restrict__ dy, const int64_t* __restrict__ t,
                               float* __restrict__ dtemb, int64_t rows, int W) {
  const int wq = W / 4;
  const int64_t n = rows * wq;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / wq;
    const int j = (int)(i - r * wq) * 4;
    float v[4];
    load4<TA>(dy + r * W + j, v);
    red_add_v4(dtemb + t[r] * W + j, v[0], v[1], v[2], v[3]);
  }
}

// out = SiLU(in) (fp32, accurate exp) and dpre = dhid * SiLU'(pre): time-embedding MLP table
__global__ void k_silu_f32(const float* __restrict__ in, float* __restrict__ out, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float x = in[i];
    out[i] = x / (1.0f + expf(-x));
  }
}
__global__ void k_dsilu_mul(float* __restrict__ d, const float* __restrict__ pre, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    d[i] *= dsilu_acc(pre[i]);
}

}  // namespace uva
