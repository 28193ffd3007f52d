// fp32 FFMA GEMM (CUDA cores): D[M,N] = A[M,K] . B[N,K]^T, both operands K-contiguous
// ("TN"), fp32 accumulate, fused epilogue.  This is the arithmetic of the fp32 mode
// (<= 1e-4 relative vs the reference, which runs fp32 with TF32 off) and of the small
// one-off GEMMs (timestep-embedding table).  TIn = float or bf16.
#pragma once
#include "common.cuh"
#include "epilogue.cuh"

namespace uva {

constexpr int SG_BM = 128, SG_BN = 128, SG_BK = 16, SG_THREADS = 256;

template <typename TIn>
__device__ __forceinline__ void sg_load4(const TIn* p, float (&v)[4]);
template <>
__device__ __forceinline__ void sg_load4<float>(const float* p, float (&v)[4]) {
  const float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
template <>
__device__ __forceinline__ void sg_load4<bf16>(const bf16* p, float (&v)[4]) {
  const uint2 t = *reinterpret_cast<const uint2*>(p);
  const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&t.x);
  const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&t.y);
  v[0] = __low2float(a); v[1] = __high2float(a); v[2] = __low2float(b); v[3] = __high2float(b);
}

// K must be a multiple of 4 (tails in K are zero-filled per 4-element group); lda/ldb multiples of 4.
// A_MN / B_MN: the operand is stored [K, MN] row-major (element (m,k) at A[k*lda + m]); then M (N)
// must be a multiple of 4 and K is free.
template <int KIND, typename TA, typename TIn, bool A_MN = false, bool B_MN = false>
__global__ void __launch_bounds__(SG_THREADS)
sgemm_tn_kernel(const TIn* __restrict__ A, int64_t lda, const TIn* __restrict__ B, int64_t ldb,
                int64_t M, int N, int K, int kper, EpiArgs ep) {
  __shared__ float As[SG_BK][SG_BM + 4];
  __shared__ float Bs[SG_BK][SG_BN + 4];
  const int tid = threadIdx.x;
  const int64_t m0 = (int64_t)blockIdx.y * SG_BM;
  const int n0 = blockIdx.x * SG_BN;
  const int ty = tid / 16, tx = tid % 16;
  const int lrow = tid / 4;        // 0..63
  const int lk = (tid % 4) * 4;    // 0,4,8,12

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  const int kbeg = blockIdx.z * kper;
  if (kbeg < K && kbeg + kper < K) K = kbeg + kper;  // this split's k range is [kbeg, K)
  for (int k0 = kbeg; k0 < K; k0 += SG_BK) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int r = lrow + 64 * h;
      const int idx = tid + SG_THREADS * h;  // MN-major: 16 k x 32 float4 along mn
      const int tk = idx / 32, tm = (idx % 32) * 4;
      float va[4] = {0.f, 0.f, 0.f, 0.f}, vb[4] = {0.f, 0.f, 0.f, 0.f};
      if constexpr (A_MN) {
        if (m0 + tm < M && k0 + tk < K) sg_load4<TIn>(A + (int64_t)(k0 + tk) * lda + m0 + tm, va);
#pragma unroll
        for (int q = 0; q < 4; ++q) As[tk][tm + q] = va[q];
      } else {
        if (m0 + r < M && k0 + lk < K) sg_load4<TIn>(A + (m0 + r) * lda + k0 + lk, va);
#pragma unroll
        for (int q = 0; q < 4; ++q) As[lk + q][r] = va[q];
      }
      if constexpr (B_MN) {
        if (n0 + tm < N && k0 + tk < K) sg_load4<TIn>(B + (int64_t)(k0 + tk) * ldb + n0 + tm, vb);
#pragma unroll
        for (int q = 0; q < 4; ++q) Bs[tk][tm + q] = vb[q];
      } else {
        if (n0 + r < N && k0 + lk < K) sg_load4<TIn>(B + (int64_t)(n0 + r) * ldb + k0 + lk, vb);
#pragma unroll
        for (int q = 0; q < 4; ++q) Bs[lk + q][r] = vb[q];
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < SG_BK; ++k) {
      float a[8], b[8];
      const float4 a0 = *reinterpret_cast<const float4*>(&As[k][ty * 8]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[k][ty * 8 + 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[k][tx * 8]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[k][tx * 8 + 4]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t r = m0 + ty * 8 + i;
    if (r >= M) continue;
#pragma unroll
    for (int j = 0; j < 8; j += 2) {
      const int c = n0 + tx * 8 + j;
      if (c < N) Epi<KIND, TA>::apply2(ep, r, c, acc[i][j], acc[i][j + 1]);
    }
  }
}

template <int KIND, typename TA, typename TIn, bool A_MN = false, bool B_MN = false>
inline cudaError_t sgemm_tn(const TIn* A, int64_t lda, const TIn* B, int64_t ldb, int64_t M, int N,
                            int64_t K, const EpiArgs& ep, cudaStream_t st) {
  if (M <= 0 || N <= 0) return cudaSuccess;
  // split-K for the reducing epilogue (weight gradients: M x N small, K = rows long)
  int splits = 1;
  if (KIND == EPI_ATOMIC_F32) {
    const int64_t tiles = ((N + SG_BN - 1) / SG_BN) * ((M + SG_BM - 1) / SG_BM);
    while (splits < 64 && tiles * splits < 296 && K / (splits * 2) >= 256) splits *= 2;
  }
  const int64_t kper = ((K + splits - 1) / splits + SG_BK - 1) / SG_BK * SG_BK;
  dim3 grid((N + SG_BN - 1) / SG_BN, (unsigned)((M + SG_BM - 1) / SG_BM), splits);
  sgemm_tn_kernel<KIND, TA, TIn, A_MN, B_MN><<<grid, SG_THREADS, 0, st>>>(A, lda, B, ldb, M, N, (int)K, (int)kper, ep);
  return cudaGetLastError();
}

}  // namespace uva
