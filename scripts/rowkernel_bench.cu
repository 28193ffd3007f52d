// Stand-alone timing of the backward row kernels (train_bwd_kernels.cuh) at the C2 shapes: each kernel alone on the
// device, CUDA events, buffers rotated so that the inputs come from HBM (3 sets of a > 126 MB working set).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o scripts/rowkernel_bench.bin scripts/rowkernel_bench.cu
//   gpurun -- ./scripts/rowkernel_bench.bin
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../unified_video_action_b200/csrc/train_bwd_kernels.cuh"
#include "../unified_video_action_b200/csrc/elementwise.cuh"

using namespace uva;

#define CK(x)                                                                       \
  do {                                                                              \
    cudaError_t e_ = (x);                                                           \
    if (e_ != cudaSuccess) {                                                        \
      fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));    \
      exit(1);                                                                      \
    }                                                                               \
  } while (0)

template <typename F>
static float time_it(const char* name, double bytes, int iters, F&& launch) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  for (int i = 0; i < 3; ++i) launch(i);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(a));
  for (int i = 0; i < iters; ++i) launch(i);
  CK(cudaEventRecord(b));
  CK(cudaDeviceSynchronize());
  CK(cudaGetLastError());
  float ms = 0.f;
  CK(cudaEventElapsedTime(&ms, a, b));
  const float us = ms * 1000.f / iters;
  printf("%-44s %8.1f us  %7.1f GB/s\n", name, us, bytes / us * 1e-3);
  return us;
}

int main(int argc, char** argv) {
  const int64_t N = argc > 1 ? atol(argv[1]) : 16384;
  const int W = 1024, SETS = 3;
  const int64_t ldm = (argc > 2 ? atol(argv[2]) : 20) * W;  // modulation row stride in units of W (20: [N, nmod], 3: block-major)
  int sms = 148;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  float *dh[SETS], *h[SETS], *stats, *xt, *dWi, *dbi, *lnw, *lnb, *dlnw, *dlnb, *dsum, *sdq, *sdg;
  bf16 *da[SETS], *q[SETS], *dq[SETS], *mod, *dmod;
  for (int s = 0; s < SETS; ++s) {
    CK(cudaMalloc(&dh[s], N * W * 4)); CK(cudaMemset(dh[s], 0, N * W * 4));
    CK(cudaMalloc(&h[s], N * W * 4)); CK(cudaMemset(h[s], 0, N * W * 4));
    CK(cudaMalloc(&da[s], N * W * 2)); CK(cudaMemset(da[s], 0, N * W * 2));
    CK(cudaMalloc(&q[s], N * W * 2)); CK(cudaMemset(q[s], 0, N * W * 2));
    CK(cudaMalloc(&dq[s], N * W * 2));
  }
  CK(cudaMalloc(&mod, N * ldm * 2)); CK(cudaMemset(mod, 0, N * ldm * 2));
  CK(cudaMalloc(&dmod, N * ldm * 2));
  CK(cudaMalloc(&stats, N * 2 * 4)); CK(cudaMemset(stats, 0, N * 2 * 4));
  CK(cudaMalloc(&xt, N * 16 * 4)); CK(cudaMemset(xt, 0, N * 16 * 4));
  CK(cudaMalloc(&dWi, W * 16 * 4)); CK(cudaMalloc(&dbi, W * 4));
  CK(cudaMalloc(&lnw, W * 4)); CK(cudaMalloc(&lnb, W * 4)); CK(cudaMalloc(&dlnw, W * 4)); CK(cudaMalloc(&dlnb, W * 4));
  CK(cudaMemset(lnw, 0, W * 4)); CK(cudaMemset(lnb, 0, W * 4));
  CK(cudaMalloc(&dsum, 2 * W * 4)); CK(cudaMalloc(&sdq, W * 4)); CK(cudaMalloc(&sdg, W * 4));
  const double RW = (double)N * W;
  printf("rows %lld, %d SMs, modulation row stride %lld\n", (long long)N, sms, (long long)ldm);

  for (int i = 0; i < 300; ++i) CK(cudaMemcpyAsync(dh[(i + 1) % SETS], dh[i % SETS], N * W * 4, cudaMemcpyDeviceToDevice, 0));  // clock ramp
  CK(cudaDeviceSynchronize());
  // ---- sampling row kernels: ys = SiLU(cemb + temb[t]) (reads fp32 h[] as cemb, writes bf16 da[])
  {
    float* temb; CK(cudaMalloc(&temb, 1000 * W * 4)); CK(cudaMemset(temb, 0, 1000 * W * 4));
    for (int mult : {8, 16}) {
      char nm[64];
      snprintf(nm, sizeof nm, "k_ysilu grid %d", mult * sms);
      time_it(nm, RW * 6, 20, [&](int i) { k_ysilu<bf16><<<mult * sms, 256>>>(h[i % SETS], temb, nullptr, 500, da[i % SETS], nullptr, N, W); });
      snprintf(nm, sizeof nm, "k_ysilu generic form, grid %d x 128", 2 * mult * sms);
      time_it(nm, RW * 6, 20, [&](int i) { k_ysilu<bf16><<<2 * mult * sms, 128>>>(h[i % SETS], temb, nullptr, 500, da[i % SETS], nullptr, N, W); });
    }
  }
  // ---- input_proj weight gradient
  for (int C : {16, 2}) {
    char nm[64];
    snprintf(nm, sizeof nm, "k_inproj_wgrad C=%d", C);
    const int rpb = iw_rows_per_block(N, W, sms);
    dim3 g((W + 255) / 256, (unsigned)((N + rpb - 1) / rpb));
    CK(cudaFuncSetAttribute(k_inproj_wgrad, cudaFuncAttributeMaxDynamicSharedMemorySize, kIwRingBytes));
    time_it(nm, RW * 4, 20, [&](int i) { k_inproj_wgrad<<<g, 256, kIwRingBytes>>>(dh[i % SETS], xt, dWi, dbi, N, W, C, rpb); });
  }
  // ---- gate backward (block b: shift | scale | gate thirds at columns 3 W b)
  for (int thr : {256, 512, 1024}) {
    for (int mult : {1, 2}) {
      const int gb = mult * sms * (thr == 256 ? 4 : thr == 512 ? 2 : 1);
      char nm[64];
      snprintf(nm, sizeof nm, "k_gate_bwd + sums, %d x %d", gb, thr);
      time_it(nm, RW * 12, 20, [&](int i) {
        k_gate_bwd<bf16><<<gb, thr>>>(dh[i % SETS], mod + 2 * W, q[i % SETS], ldm, dq[i % SETS], dmod + 2 * W, N, W, sdq, sdg);
      });
    }
  }
  time_it("k_gate_bwd no sums, 148 x 1024", RW * 12, 20, [&](int i) {
    k_gate_bwd<bf16><<<sms, 1024>>>(dh[i % SETS], mod + 2 * W, q[i % SETS], ldm, dq[i % SETS], dmod + 2 * W, N, W, nullptr, nullptr);
  });
  // plain copy of the same byte count as a yardstick (read 8 B + write 4 B per element like k_gate_bwd)
  time_it("copy yardstick (float4 in, float2 out x 2 streams)", RW * 12, 20, [&](int i) {
    cudaMemcpyAsync(dh[(i + 1) % SETS], dh[i % SETS], N * W * 4, cudaMemcpyDeviceToDevice, 0);
    cudaMemcpyAsync(dq[(i + 1) % SETS], dq[i % SETS], N * W * 2, cudaMemcpyDeviceToDevice, 0);
  });
  // ---- LayerNorm-modulate backward
  for (int blocks : {3 * sms, 4 * sms, 5 * sms, 6 * sms, 8 * sms}) {
    char nm[64];
    snprintf(nm, sizeof nm, "k_lnmod_bwd_split P=8 grid %d + all sums", blocks);
    time_it(nm, RW * 20, 20, [&](int i) {
      k_lnmod_bwd_split<bf16, 8, 8><<<blocks, 256>>>(da[i % SETS], h[i % SETS], stats, lnw, lnb, mod + W, ldm, dmod, dmod + W, dh[i % SETS], 0, dlnw, dlnb, N, dsum);
    });
  }
  for (int blocks : {2 * sms, 4 * sms}) {
    char nm[64];
    snprintf(nm, sizeof nm, "k_lnmod_bwd_split P=4 grid %d + all sums", blocks);
    time_it(nm, RW * 20, 20, [&](int i) {
      k_lnmod_bwd_split<bf16, 8, 4><<<blocks, 256>>>(da[i % SETS], h[i % SETS], stats, lnw, lnb, mod + W, ldm, dmod, dmod + W, dh[i % SETS], 0, dlnw, dlnb, N, dsum);
    });
  }
  for (int blocks : {2 * sms, 3 * sms, 4 * sms}) {
    char nm[80];
    GateNext<bf16> gn{mod + 2 * W, q[0], dq[0], dmod + 2 * W, sdq, sdg};
    snprintf(nm, sizeof nm, "k_lnmod_bwd_split P=8 + gate backward, grid %d", blocks);
    time_it(nm, RW * 28, 20, [&](int i) {
      gn.q = q[i % SETS]; gn.dq = dq[i % SETS];
      k_lnmod_bwd_split<bf16, 8, 8, true><<<blocks, 256>>>(da[i % SETS], h[i % SETS], stats, lnw, lnb, mod + W, ldm, dmod + 3 * W, dmod + 4 * W, dh[i % SETS], 0, dlnw, dlnb, N, dsum, gn);
    });
    snprintf(nm, sizeof nm, "k_lnmod_bwd_split P=4 + gate backward, grid %d", blocks);
    time_it(nm, RW * 28, 20, [&](int i) {
      gn.q = q[i % SETS]; gn.dq = dq[i % SETS];
      k_lnmod_bwd_split<bf16, 8, 4, true><<<blocks, 256>>>(da[i % SETS], h[i % SETS], stats, lnw, lnb, mod + W, ldm, dmod + 3 * W, dmod + 4 * W, dh[i % SETS], 0, dlnw, dlnb, N, dsum, gn);
    });
  }
  time_it("k_lnmod_bwd_split P=8 grid 592, no sums", RW * 20, 20, [&](int i) {
    k_lnmod_bwd_split<bf16, 8, 8><<<4 * sms, 256>>>(da[i % SETS], h[i % SETS], stats, lnw, lnb, mod + W, ldm, dmod, dmod + W, dh[i % SETS], 0, nullptr, nullptr, N, nullptr);
  });
  time_it("k_lnmod_bwd_split P=8 grid 592, first, no sums", RW * 16, 20, [&](int i) {
    k_lnmod_bwd_split<bf16, 8, 8><<<4 * sms, 256>>>(da[i % SETS], h[i % SETS], stats, nullptr, nullptr, mod + W, ldm, dmod, dmod + W, dh[i % SETS], 1, nullptr, nullptr, N, nullptr);
  });
  time_it("k_lnmod_bwd_split P=4 grid 592, first, no sums", RW * 16, 20, [&](int i) {
    k_lnmod_bwd_split<bf16, 8, 4><<<4 * sms, 256>>>(da[i % SETS], h[i % SETS], stats, nullptr, nullptr, mod + W, ldm, dmod, dmod + W, dh[i % SETS], 1, nullptr, nullptr, N, nullptr);
  });
  // ---- column sums of [N, 2 W] (what the fused sums replace)
  {
    dim3 g((2 * W + 255) / 256, (unsigned)((N + CS_ROWS - 1) / CS_ROWS));
    time_it("k_colsum [N, 2W] bf16", RW * 2 * 2, 20, [&](int i) { k_colsum<bf16><<<g, 256>>>(dmod + (i % 3) * 3 * W, ldm, N, 2 * W, dsum); });
  }
  return 0;
}
