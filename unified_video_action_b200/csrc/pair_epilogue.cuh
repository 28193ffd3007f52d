// Epilogue building blocks of the CTA-pair GEMM: warp-transposed global IO.
//
// tcgen05.ld hands every thread 32 consecutive columns of ITS OWN row, so a direct global access
// touches 32 different 128-byte lines per warp instruction.  Here every global access of an epilogue
// is made in the transposed mapping (8 lanes cover one 128-byte row segment, a warp instruction covers
// 4 full rows) and moved to / from the row-per-thread mapping through a small per-warp shared-memory
// tile (padded rows: conflict-free in both mappings).  Only __syncwarp() is needed.
#pragma once
#include "common.cuh"
#include "epilogue.cuh"

namespace uva {

constexpr int PE_F32_LD = 36;                      // floats per staged fp32 row (32 + 4 pad)
constexpr int PE_B16_LD = 20;                      // 32-bit words per staged bf16 row (16 + 4 pad)
constexpr int PE_WARP_BYTES = 32 * PE_F32_LD * 4 + 32 * PE_B16_LD * 4;  // 4608 + 2560 = 7168

struct WarpStage {
  float* f;      // [32][PE_F32_LD]
  uint32_t* b;   // [32][PE_B16_LD]
};

// ---- fp32 [32 rows x 32 cols] tiles ---------------------------------------------------------
__device__ __forceinline__ void tile_ldg_f32(const float* g, int64_t ld, int rows_valid, int lane, float4 (&r)[8]) {
  const int c = (lane & 7) * 4;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int row = (lane >> 3) + 4 * i;
    r[i] = (row < rows_valid) ? *reinterpret_cast<const float4*>(g + row * ld + c) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
}
// staged tile -> this thread's row (call after tile_ldg_f32 data has arrived)
__device__ __forceinline__ void tile_to_row_f32(float* sm, int lane, const float4 (&r)[8], float (&row)[32]) {
  const int c = (lane & 7) * 4;
  __syncwarp();  // earlier transposed-phase reads of this buffer are done
#pragma unroll
  for (int i = 0; i < 8; ++i) *reinterpret_cast<float4*>(sm + ((lane >> 3) + 4 * i) * PE_F32_LD + c) = r[i];
  __syncwarp();
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float4 t = *reinterpret_cast<const float4*>(sm + lane * PE_F32_LD + 4 * i);
    row[4 * i] = t.x; row[4 * i + 1] = t.y; row[4 * i + 2] = t.z; row[4 * i + 3] = t.w;
  }
}
// this thread's row -> coalesced global store
__device__ __forceinline__ void row_stg_f32(float* sm, float* g, int64_t ld, int rows_valid, int lane, const float (&row)[32]) {
  const int c = (lane & 7) * 4;
  __syncwarp();  // earlier transposed-phase reads of this buffer are done
#pragma unroll
  for (int i = 0; i < 8; ++i)
    *reinterpret_cast<float4*>(sm + lane * PE_F32_LD + 4 * i) = make_float4(row[4 * i], row[4 * i + 1], row[4 * i + 2], row[4 * i + 3]);
  __syncwarp();
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = (lane >> 3) + 4 * i;
    if (r < rows_valid) *reinterpret_cast<float4*>(g + r * ld + c) = *reinterpret_cast<const float4*>(sm + r * PE_F32_LD + c);
  }
}

// ---- bf16 [32 rows x 32 cols] tiles (64 bytes per row) ----------------------------------------
__device__ __forceinline__ void tile_ldg_b16(const bf16* g, int64_t ld, int rows_valid, int lane, uint4 (&r)[4]) {
  const int c = (lane & 3) * 8;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int row = (lane >> 2) + 8 * i;
    r[i] = (row < rows_valid) ? *reinterpret_cast<const uint4*>(g + row * ld + c) : make_uint4(0u, 0u, 0u, 0u);
  }
}
__device__ __forceinline__ void tile_to_row_b16(uint32_t* sm, int lane, const uint4 (&r)[4], uint32_t (&row)[16]) {
  const int c = (lane & 3) * 4;
  __syncwarp();
#pragma unroll
  for (int i = 0; i < 4; ++i) *reinterpret_cast<uint4*>(sm + ((lane >> 2) + 8 * i) * PE_B16_LD + c) = r[i];
  __syncwarp();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint4 t = *reinterpret_cast<const uint4*>(sm + lane * PE_B16_LD + 4 * i);
    row[4 * i] = t.x; row[4 * i + 1] = t.y; row[4 * i + 2] = t.z; row[4 * i + 3] = t.w;
  }
}
__device__ __forceinline__ void row_stg_b16(uint32_t* sm, bf16* g, int64_t ld, int rows_valid, int lane, const uint32_t (&row)[16]) {
  const int c = (lane & 3) * 4;
  __syncwarp();  // earlier transposed-phase reads of this buffer are done
#pragma unroll
  for (int i = 0; i < 4; ++i)
    *reinterpret_cast<uint4*>(sm + lane * PE_B16_LD + 4 * i) = make_uint4(row[4 * i], row[4 * i + 1], row[4 * i + 2], row[4 * i + 3]);
  __syncwarp();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = (lane >> 2) + 8 * i;
    if (r < rows_valid) *reinterpret_cast<uint4*>(g + r * ld + c * 2) = *reinterpret_cast<const uint4*>(sm + r * PE_B16_LD + c);
  }
}

__device__ __forceinline__ void bias32_ldg(const float* bias, float (&b)[32]) {
  if (bias != nullptr) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(bias) + i);
      b[4 * i] = t.x; b[4 * i + 1] = t.y; b[4 * i + 2] = t.z; b[4 * i + 3] = t.w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i) b[i] = 0.f;
  }
}

__device__ __forceinline__ void pack32(const float (&x)[32], uint32_t (&p)[16]) {
#pragma unroll
  for (int i = 0; i < 16; ++i) p[i] = pack_bf16x2(x[2 * i], x[2 * i + 1]);
}

}  // namespace uva
