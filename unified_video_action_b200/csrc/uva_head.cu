// libuva_head.so — host orchestration + C ABI (include/uva_head.h) of the B200-native UVA
// diffusion head.  All device work is in the kernels of this directory; this file sequences
// them on the caller's stream.  No CPU fallback exists: every compute entry point launches
// CUDA kernels or fails.
#include "../../include/uva_head.h"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"
#include "elementwise.cuh"
#include "epilogue.cuh"
#include "schedule.hpp"
#include "sgemm.cuh"
#include "train_bwd_kernels.cuh"
#include "train_kernels.cuh"
#include "umma_gemm.cuh"
#include "umma_gemm2.cuh"
#include "stack_kernel.cuh"
#include "chain_kernel.cuh"
#include "final_ddpm.cuh"
#include "optim.cuh"
#include "prenet_kernels.cuh"

#include <nvtx3/nvToolsExt.h>

using namespace uva;

// NVTX ranges around every compute entry point of the C ABI (tracing).  UVA_NVTX=1 turns them on (read
// once); a range covers the host-side enqueue, and ncu / nsys can filter on it (--nvtx --nvtx-include "uva_head_sample/").
struct NvtxRange {
  bool on;
  explicit NvtxRange(const char* name) {
    static const bool enabled = [] { const char* e = getenv("UVA_NVTX"); return e && atoi(e) != 0; }();
    on = enabled;
    if (on) nvtxRangePushA(name);
  }
  ~NvtxRange() { if (on) nvtxRangePop(); }
};

// ----------------------------------------------------------------------------------------------
// errors
// ----------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";
static int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
#define CUDA_TRY(expr)                                                                          \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess)                                                                      \
      return fail(UVA_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                  __LINE__);                                                                    \
  } while (0)
#define UVA_TRY(expr)          \
  do {                         \
    int _r = (expr);           \
    if (_r != UVA_OK) return _r; \
  } while (0)

extern "C" const char* uva_last_error(void) { return g_err; }
extern "C" const char* uva_build_info(void) {
  return "libuva_head sm_100a (tcgen05/TMEM/TMA bf16 GEMM core + fp32 FFMA mode), built " __DATE__;
}

// ----------------------------------------------------------------------------------------------
// parameter table (reference state_dict order, DL:222-235)
// ----------------------------------------------------------------------------------------------
enum { P_TE0_W = 0, P_TE0_B, P_TE2_W, P_TE2_B, P_COND_W, P_COND_B, P_INP_W, P_INP_B, P_BLOCK0 };
enum { B_LN_W = 0, B_LN_B, B_W1, B_B1, B_W2, B_B2, B_ADA_W, B_ADA_B, B_COUNT };
enum { F_LIN_W = 0, F_LIN_B, F_ADA_W, F_ADA_B, F_COUNT };
static const int kFreq = 256;  // frequency_embedding_size (DL:103)

extern "C" int32_t uva_head_num_params(int32_t depth) { return P_BLOCK0 + B_COUNT * depth + F_COUNT; }

static void param_shape(const UvaHeadDesc& d, int idx, int64_t& rows, int64_t& cols) {
  const int64_t W = d.W, C = d.C, Z = d.Z;
  rows = cols = 0;
  if (idx < P_BLOCK0) {
    switch (idx) {
      case P_TE0_W: rows = W; cols = kFreq; break;
      case P_TE0_B: rows = 1; cols = W; break;
      case P_TE2_W: rows = W; cols = W; break;
      case P_TE2_B: rows = 1; cols = W; break;
      case P_COND_W: rows = W; cols = Z; break;
      case P_COND_B: rows = 1; cols = W; break;
      case P_INP_W: rows = W; cols = C; break;
      case P_INP_B: rows = 1; cols = W; break;
    }
    return;
  }
  const int fbase = P_BLOCK0 + B_COUNT * d.depth;
  if (idx < fbase) {
    switch ((idx - P_BLOCK0) % B_COUNT) {
      case B_LN_W: case B_LN_B: case B_B1: case B_B2: rows = 1; cols = W; break;
      case B_W1: case B_W2: rows = W; cols = W; break;
      case B_ADA_W: rows = 3 * W; cols = W; break;
      case B_ADA_B: rows = 1; cols = 3 * W; break;
    }
    return;
  }
  switch (idx - fbase) {
    case F_LIN_W: rows = 2 * C; cols = W; break;
    case F_LIN_B: rows = 1; cols = 2 * C; break;
    case F_ADA_W: rows = 2 * W; cols = W; break;
    case F_ADA_B: rows = 1; cols = 2 * W; break;
  }
}

extern "C" int64_t uva_head_param_numel(const UvaHeadDesc* d, int32_t idx) {
  if (!d || idx < 0 || idx >= uva_head_num_params(d->depth)) return -1;
  int64_t r, c;
  param_shape(*d, idx, r, c);
  return r * c;
}

extern "C" int uva_schedule_table(int32_t T_train, int32_t T_gen, int32_t which, const char* name,
                                  double* out, int32_t cap) {
  if (T_train <= 1 || T_gen <= 0 || T_gen > T_train || !name || !out)
    return fail(UVA_ERR_INVALID, "uva_schedule_table: bad arguments");
  const ScheduleTables S = create_schedule(T_train, which == 0 ? T_gen : T_train);
  const std::string n(name);
  const int T = S.T();
  if (n == "timestep_map") {
    for (int i = 0; i < T && i < cap; ++i) out[i] = (double)S.timestep_map[i];
    return T;
  }
  const std::vector<double>* v = S.by_name(n);
  if (!v) return fail(UVA_ERR_INVALID, "uva_schedule_table: unknown table '%s'", name);
  for (int i = 0; i < T && i < cap; ++i) out[i] = (*v)[i];
  return T;
}

// ----------------------------------------------------------------------------------------------
// handle
// ----------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);

struct DevTables {  // fp32 device copies of one process' tables, [T] each
  float *sqrt_ac = nullptr, *sqrt_1mac = nullptr, *sqrt_recip = nullptr, *sqrt_recipm1 = nullptr,
        *coef1 = nullptr, *coef2 = nullptr, *post_logvar = nullptr, *log_beta = nullptr;
};

struct UvaPrenet;
static void prenet_free(UvaPrenet* p);
static const int kLnPrePanels = 64;  // 128-row panels the LayerNorm pre-phase barriers cover (latency path: <= small_rows / 128)
struct UvaHead {
  UvaHeadDesc d;
  int num_sms = 148;
  int gemm_sms = 148;       // SMs the persistent GEMM grids may take right now (num_sms, or less inside a training call)
  int train_reserve = 0;    // uva_head_set_train_reserve: SMs left free during loss_fwd / loss_bwd for a head on another stream
  int64_t chunk_rows = 65536;
  int nmod = 0;  // (3*depth + 2) * W : width of the concatenated AdaLN modulation output
  ScheduleTables gen, train;
  DevTables dtrain;
  float* d_tables = nullptr;  // backing store of dtrain
  float* feat = nullptr;      // [T_train, 256] sinusoidal features (DL:110-134)
  float* temb = nullptr;      // [T_train, W] TimestepEmbedder(t) table
  float* temb_hidden = nullptr;  // [T_train, W] SiLU(Linear0(feat)) (kept for the backward)
  float* temb_pre = nullptr;     // [T_train, W] Linear0(feat) before the SiLU (kept for the backward)
  bf16 *temb_hidden_b = nullptr, *temb_pre_b = nullptr, *feat_b = nullptr;  // bf16 copies: operands of the backward GEMMs (bf16 mode)
  size_t ob_te2w = 0;            // bf16 copy of time_embed.mlp.2.weight [W, W]
  // fp32 parameter copy, own layout (offsets in floats)
  float* wf = nullptr;
  size_t wf_floats = 0;
  size_t o_te0w, o_te0b, o_te2w, o_te2b, o_condw, o_condb, o_inpw, o_inpb, o_adaw, o_adab, o_linw, o_linb;
  std::vector<size_t> o_lnw, o_lnb, o_w1, o_b1, o_w2, o_b2;
  // bf16 operand copies (bf16 mode)
  bf16* wb = nullptr;
  size_t wb_elems = 0;
  size_t ob_condw, ob_adaw, ob_linw;      // [W,Z], [nmod,W], [32,W] (zero padded)
  size_t ob_adaw_p = 0;                   // [nmod,W] in (shift|scale)-tile order (k_pack_ada_perm)
  float* adab_p = nullptr;                // [nmod] bias in the same order
  std::vector<size_t> ob_w1, ob_w2;       // [W,W]
  bool weights_loaded = false;
  // Data-parallel training: while set (uva_head_set_grad_sync), loss_bwd all-reduces every contiguous range of the
  // caller's flat gradient buffer on the communicator's stream as soon as the kernels that write it are enqueued
  // (final layer, each block's MLP part, each AdaLN Linear, the embedders) - the collective hides behind the rest
  // of the backward instead of following it (collective.inc)
  struct GradSync { void* comm = nullptr; cudaStream_t cs = nullptr; int average = 1; int launched = 0; std::vector<cudaEvent_t> ev; } gsync;
  UvaPrenet* prenet = nullptr;  // fused DiffActLoss pre-net (head_prenet.inc), allocated by uva_head_prenet_load_weights
  int64_t launches = 0;
  PFN_encodeTiled encode = nullptr;
  // optional per-category GEMM timing (bench.py roofline): CUDA event pairs on the launch stream
  // second stream: two row chunks are processed concurrently so that the HBM-bound row kernels
  // of one chunk overlap the tensor-bound GEMMs of the other
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  bool overlap = true;
  // UVA_LANE_SPLIT: while two chunk lanes are in flight each lane's persistent GEMM grids take HALF of the SMs, so that
  // two GEMMs of different kinds really run side by side (the HBM-bound gated-residual GEMM of one lane next to a
  // tensor-bound one of the other) instead of alternating whole-chip kernels
  // (measured, C3 at 131072 rows: 14.14 / 14.49 -> 14.63 / 14.69 M row-steps/s under the same power cap)
  bool lane_split = true;
  bool pair = true;  // CTA-pair (cta_group::2) GEMM for M > 128
  bool quad = false;  // UVA_QUAD: clusters of two CTA pairs with multicast A boxes (forward-form GEMMs)
  bool fuse_ln = true;  // UVA_FUSE_LN: LayerNorm-modulate inside the AdaLN GEMM epilogue (bf16 sampling path)
  int grid_mult = 1;  // UVA_GRID_MULT: GEMM grids of grid_mult x #SMs CTAs (block scheduler balances the tail)
  // set by the sampler around run_stack: the final linear then runs with the DDPM update in its epilogue
  // (final_ddpm.cuh) instead of writing the net output (bf16 mode; UVA_FUSE_DDPM=0 keeps the two kernels)
  struct FinalFuse {
    bool on = false;
    const float* x_cur = nullptr; float* x_next = nullptr; const float* noise = nullptr;
    PhiloxSeed ps{0, 0, 0, 0}; uint32_t step = 0;
    int64_t half = 0, g0 = 0, ghalf = 0;
    float cfg = 1.f, temperature = 1.f; int clip = 0;
    StepCoef kc{};
  } fin;
  bool fuse_ddpm = true;
  // UVA_STACK: all tcgen05 GEMMs of a denoise step in ONE persistent launch, panel-major so that a row panel's
  // activations stay in L2 between the layers (stack_kernel.cuh).  UVA_STACK_GROUP: 256-row tiles per panel group.
  // Measured (profiles/r02_stack_*): one launch per step wins where the per-layer launches are short - 4096 rows 39.2 ->
  // 33.4 ms per 100-step call, 8192 rows 68.5 -> 61.0, 16384 rows 128.1 -> 118.0, even at 32768 - and loses from 65536
  // rows on (the rows in flight that hide the layer-to-layer dependency latency no longer fit L2), so it takes calls
  // of up to stack_rows rows (UVA_STACK_ROWS), as ONE chunk.
  bool stack = true;
  int64_t stack_rows = 32768;
  int stack_group = 32;  // 256-row tiles per group of the group-major item order; 0: diagonal wavefront order
  // set by uva_head_sample_from_decoder around the sampler: the condition rows come from the MAR decoder's
  // pre-norm output through k_decoder_tail (LayerNorm + diffusion pos-embed) instead of a z tensor
  struct TailSrc { bool on = false; DecoderTail t{}; } tail;
  bool pdl = true;       // UVA_PDL: programmatic dependent launch between the per-step kernels of the latency path
  bool pdl_big = true;   // UVA_PDL_BIG: programmatic dependent launch between the GEMMs of the chunked (throughput) path as well
  bool pdl_train = false;  // UVA_PDL_TRAIN (measured on the C2 step: 5.76 -> 5.82 ms, the side-stream head already fills the gaps; off)
  bool pdl_now = false;  // set while the latency path enqueues its chain
  // UVA_FUSE_LNPRE=1: latency path, LayerNorm-modulate as the pre-phase of GEMM1 (umma_gemm.cuh).  Bit-identical and 5
  // launches per step fewer, but measured slower on B200 (896 rows: 17.23 -> 17.52 ms per 100-step call): with
  // programmatic dependent launch the separate k_lnmod costs only its own critical path, the pre-phase adds the
  // panel barrier's two L2 round trips in front of the first TMA load -> off by default
  bool fuse_lnpre = false;
  unsigned int* lnpre_bar = nullptr;  // [kLnPrePanels][2] panel barriers of that pre-phase
  bool epi_colsum = true;   // UVA_EPI_COLSUM: bias gradients of mlp.0 / cond_embed reduced in the SiLU' GEMM epilogue (warp shuffles + atomics)
  bool fuse_gate_ln = true;  // UVA_FUSE_GATE_LN: k_gate_bwd of block b - 1 inside the LayerNorm-modulate backward of block b (78 us instead of 59 + 39)
  bool fuse_colsum = true;  // UVA_FUSE_COLSUM: bias gradients of mlp.2 / gate modulation summed inside k_gate_bwd
  bool final_tc = true;  // UVA_FINAL_TC: final-layer backward (dgrad K = 32, wgrad M = 2C) on the tcgen05 GEMMs
  bool fuse_inproj = true;  // UVA_FUSE_INPROJ: input_proj + first LayerNorm-modulate in one kernel on the latency path
  PhiloxSeed noise{0, 0, 0, 0};  // uva_head_set_noise_seed: generator of the sampler's noise when step_noise == NULL
  int dbg_skip = 0;  // -DUVA_DEBUG builds only: UVA_DEBUG_SKIP (timing diagnosis, results are wrong): 1 = no GEMM launches, 2 = no row kernels
  // latency path (head_small.inc): small batches run with the AdaLN modulation of all denoise steps
  // computed up front and the whole chain replayed from a CUDA graph
  int* d_tsteps = nullptr;       // [T_gen] original-process timestep of loop iteration k
  // UVA_SMALL_ROWS: largest batch that takes the latency path.  Measured (100-step call, cfg 1): 1536 rows 35.8 -> 27.4 ms
  // and 2048 rows 37.1 -> 29.6 ms against the chunked path (whose 256 x 256 pair tiles fill only 48 - 64 SMs there);
  // even at 3072 rows, 4 % slower at 4096.  The modulation table of all steps costs 4.1 MB per row (8.4 GB at 2048).
  int64_t small_rows = 1024;  // (2048 before the layer-chain kernel existed: 1536 rows 28.6 -> 25.9 ms, 2048 rows 32.9 -> 26.4 ms on it)
  StepCoef* d_coef = nullptr;    // [T_gen] DDPM update coefficients of the sampling process, by timestep index
  bool use_chain = true;         // UVA_CHAIN: persistent chain kernel (chain_kernel.cuh)
  int64_t chain_rows = 32;       // UVA_CHAIN_ROWS: largest batch it takes (its time grows with rows / 16)
  bool use_graph = true;         // UVA_GRAPH
  // UVA_HOIST_OVERLAP: graph path (33 .. small_rows rows): the modulation table is hoisted in half groups and the GEMM of
  // the NEXT half group runs on its own stream, on hoist_sms SMs (UVA_HOIST_SMS; 0 = sized from the row count), beside the
  // launch-bound chain of the current one, whose tiles are then picked for the remaining chain_sms SMs
  bool hoist_overlap = true;
  int hoist_sms = 0;
  int chain_sms = 0;             // SMs the chain's one-wave tile choice may fill right now (0: all)
  cudaStream_t hoist_st = nullptr;
  cudaEvent_t ev_hfork = nullptr, ev_hdone = nullptr;
  // UVA_WGRAD_STREAM: the W x W weight-gradient GEMMs of loss_bwd (64 of the 74 cluster slots each, off the dependency
  // chain) run on hoist_st beside the data-gradient GEMMs and row kernels of the chain, which wait for them only
  // before their operand buffer is rewritten one block later (measured, C2 step, 4 A/B pairs: 5.23 -> 5.06 ms)
  bool wgrad_stream = true;
  int wgrad_side_mult = 1;  // UVA_WGRAD_SIDE_MULT: grid multiplier of the side-stream weight-gradient GEMMs (4 = one item per cluster: measured 5.07 vs 5.03 ms, no gain)
  cudaEvent_t ev_wfork = nullptr, ev_w1 = nullptr, ev_w2 = nullptr;
  bool bn32 = true;              // UVA_BN32: 128 x 32 tiles for the W x W GEMMs when rows <= 128 (32 CTAs stream the weight)
  struct SmallGraph {
    const void* ws; int64_t N; int k_begin, k_end, clip; float temperature, cfg;
    cudaGraphExec_t exec; int64_t launches; int seen;
  };
  std::vector<SmallGraph> graphs;
  // training calls on small batches (the action head: 16 rows per sample) are launch bound on the HOST (~140
  // launches with their tensor maps per fwd + bwd): the second call with the same pointers captures a CUDA graph,
  // later calls replay it.  Key = every pointer and size the enqueued work depends on.  UVA_TRAIN_GRAPH=0: off.
  struct TrainGraph {
    std::vector<uintptr_t> key;
    cudaGraphExec_t exec; int64_t launches; int seen;
  };
  std::vector<TrainGraph> tgraphs;
  bool train_graph = true;
  int64_t train_graph_rows = 1024;  // UVA_TRAIN_GRAPH_ROWS: largest batch whose loss_fwd / loss_bwd replay from a CUDA graph
  bool timing = false;
  int cur_cat = 0;
  struct TimedLaunch { int cat; cudaEvent_t a, b; };
  std::vector<TimedLaunch> timed;
};
enum { CAT_ADA = 0, CAT_MLP1, CAT_MLP2, CAT_FINAL, CAT_COND, CAT_OTHER, CAT_STACK, CAT_COUNT };

#define LAUNCH_CHECK(h)                                                                       \
  do {                                                                                        \
    (h)->launches++;                                                                          \
    cudaError_t _e = cudaGetLastError();                                                      \
    if (_e != cudaSuccess)                                                                    \
      return fail(UVA_ERR_CUDA, "kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e),   \
                  __FILE__, __LINE__);                                                        \
  } while (0)

static inline int grid_for(int64_t n, int threads, int sms) {
  int64_t b = (n + threads - 1) / threads;
  const int64_t cap = (int64_t)sms * 16;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

extern "C" int64_t uva_head_launch_count(const UvaHead* h) { return h ? h->launches : 0; }

extern "C" int uva_head_set_noise_seed(UvaHead* h, uint64_t seed, uint32_t offset, int32_t enable) {
  if (!h) return fail(UVA_ERR_INVALID, "set_noise_seed: null handle");
  h->noise = PhiloxSeed{(uint32_t)seed, (uint32_t)(seed >> 32), offset, enable ? 1u : 0u};
  return UVA_OK;
}

extern "C" int uva_philox_normal(uint64_t seed, uint32_t offset, int32_t step0, int64_t per_step, int64_t n, float* out,
                                 void* stream) {
  if (!out || n < 0 || per_step <= 0 || step0 < 0) return fail(UVA_ERR_INVALID, "philox_normal: bad argument");
  if (n == 0) return UVA_OK;
  int dev = 0, sms = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const PhiloxSeed s{(uint32_t)seed, (uint32_t)(seed >> 32), offset, 1u};
  k_philox_fill<<<grid_for(n, 256, sms), 256, 0, (cudaStream_t)stream>>>(s, (uint32_t)step0, per_step, n, out);
  const cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) return fail(UVA_ERR_CUDA, "philox_normal: %s", cudaGetErrorString(e));
  return UVA_OK;
}

extern "C" int uva_head_set_train_reserve(UvaHead* h, int32_t sms) {
  if (!h || sms < 0 || sms >= h->num_sms) return fail(UVA_ERR_INVALID, "set_train_reserve: bad argument");
  h->train_reserve = sms / 2 * 2;
  return UVA_OK;
}

extern "C" int uva_head_timing_enable(UvaHead* h, int32_t enable) {
  if (!h) return fail(UVA_ERR_INVALID, "timing_enable: null handle");
  h->timing = enable != 0;
  return UVA_OK;
}

extern "C" int uva_head_timing_read(UvaHead* h, int32_t category, double* total_ms, int64_t* count) {
  if (!h || !total_ms || !count) return fail(UVA_ERR_INVALID, "timing_read: null argument");
  double ms = 0.0;
  int64_t n = 0;
  std::vector<UvaHead::TimedLaunch> keep;
  for (auto& t : h->timed) {
    if (t.cat != category) { keep.push_back(t); continue; }
    CUDA_TRY(cudaEventSynchronize(t.b));
    float e = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&e, t.a, t.b));
    ms += e;
    n++;
    cudaEventDestroy(t.a);
    cudaEventDestroy(t.b);
  }
  h->timed.swap(keep);
  *total_ms = ms;
  *count = n;
  return UVA_OK;
}

static int upload_tables(const ScheduleTables& S, DevTables& dt, float** backing) {
  const int T = S.T();
  std::vector<float> host((size_t)8 * T);
  const std::vector<double>* src[8] = {&S.sqrt_alphas_cumprod, &S.sqrt_one_minus_alphas_cumprod,
                                       &S.sqrt_recip_alphas_cumprod, &S.sqrt_recipm1_alphas_cumprod,
                                       &S.posterior_mean_coef1, &S.posterior_mean_coef2,
                                       &S.posterior_log_variance_clipped, &S.log_betas};
  for (int k = 0; k < 8; ++k)
    for (int i = 0; i < T; ++i) host[(size_t)k * T + i] = (float)(*src[k])[i];  // .float() of GD:901
  CUDA_TRY(cudaMalloc(backing, host.size() * sizeof(float)));
  CUDA_TRY(cudaMemcpy(*backing, host.data(), host.size() * sizeof(float), cudaMemcpyHostToDevice));
  float* b = *backing;
  dt.sqrt_ac = b; dt.sqrt_1mac = b + T; dt.sqrt_recip = b + 2 * T; dt.sqrt_recipm1 = b + 3 * T;
  dt.coef1 = b + 4 * T; dt.coef2 = b + 5 * T; dt.post_logvar = b + 6 * T; dt.log_beta = b + 7 * T;
  return UVA_OK;
}

extern "C" int uva_head_create(const UvaHeadDesc* desc, UvaHead** out) {
  if (!desc || !out) return fail(UVA_ERR_INVALID, "uva_head_create: null argument");
  const UvaHeadDesc& d = *desc;
  if (d.C < 1 || d.C > 16) return fail(UVA_ERR_INVALID, "target_channels C=%d unsupported (1..16)", d.C);
  if (d.depth < 1 || d.depth > 64) return fail(UVA_ERR_INVALID, "depth=%d unsupported", d.depth);
  if (d.W < 32 || d.W % 32 != 0 || d.Z < 4 || d.Z % 4 != 0)
    return fail(UVA_ERR_INVALID, "width W=%d must be a multiple of 32, Z=%d a multiple of 4", d.W, d.Z);
  if (d.mode == UVA_MODE_BF16 && (d.W % 256 != 0 || d.Z % 64 != 0))
    return fail(UVA_ERR_INVALID, "bf16 mode needs W %% 256 == 0 and Z %% 64 == 0 (W=%d, Z=%d)", d.W, d.Z);
  if (d.mode != UVA_MODE_BF16 && d.mode != UVA_MODE_FP32) return fail(UVA_ERR_INVALID, "bad mode %d", d.mode);
  if (d.T_train < 2 || d.T_gen < 1 || d.T_gen > d.T_train)
    return fail(UVA_ERR_INVALID, "bad diffusion lengths T_train=%d T_gen=%d", d.T_train, d.T_gen);
  int dev = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, dev));
  if (prop.major != 10)
    return fail(UVA_ERR_CUDA, "libuva_head is built for sm_100a only; device is sm_%d%d", prop.major, prop.minor);

  UvaHead* h = new UvaHead();
  h->d = d;
  h->num_sms = prop.multiProcessorCount;
  h->gemm_sms = h->num_sms;
  h->nmod = (3 * d.depth + 2) * d.W;
  if (const char* e = getenv("UVA_CHUNK_ROWS")) {
    long v = atol(e);
    if (v >= 256) h->chunk_rows = (v / 256) * 256;
  }
  if (const char* e = getenv("UVA_OVERLAP")) h->overlap = atoi(e) != 0;
  if (const char* e = getenv("UVA_LANE_SPLIT")) h->lane_split = atoi(e) != 0;
  if (const char* e = getenv("UVA_PAIR")) h->pair = atoi(e) != 0;
  if (const char* e = getenv("UVA_QUAD")) h->quad = atoi(e) != 0;
  if (const char* e = getenv("UVA_FUSE_LN")) h->fuse_ln = atoi(e) != 0;
  if (const char* e = getenv("UVA_STACK")) h->stack = atoi(e) != 0;
  if (const char* e = getenv("UVA_STACK_ROWS")) { long v = atol(e); if (v >= 0) h->stack_rows = v; }
  if (const char* e = getenv("UVA_STACK_GROUP")) { int v = atoi(e); if (v >= 0 && v <= 4096) h->stack_group = v; }
#ifdef UVA_DEBUG  // result-corrupting timing switches exist only in a -DUVA_DEBUG build, never in the shipped library
  if (const char* e = getenv("UVA_DEBUG_SKIP")) h->dbg_skip = atoi(e);
#endif
  if (const char* e = getenv("UVA_SMALL_ROWS")) h->small_rows = atol(e);
  if (const char* e = getenv("UVA_GRAPH")) h->use_graph = atoi(e) != 0;
  if (const char* e = getenv("UVA_CHAIN")) h->use_chain = atoi(e) != 0;
  if (const char* e = getenv("UVA_CHAIN_ROWS")) { long v = atol(e); if (v >= 1 && v <= 128) h->chain_rows = v; }
  if (const char* e = getenv("UVA_BN32")) h->bn32 = atoi(e) != 0;
  if (const char* e = getenv("UVA_FUSE_DDPM")) h->fuse_ddpm = atoi(e) != 0;
  if (const char* e = getenv("UVA_FUSE_INPROJ")) h->fuse_inproj = atoi(e) != 0;
  if (const char* e = getenv("UVA_PDL")) h->pdl = atoi(e) != 0;
  if (const char* e = getenv("UVA_HOIST_OVERLAP")) h->hoist_overlap = atoi(e) != 0;
  if (const char* e = getenv("UVA_HOIST_SMS")) h->hoist_sms = atoi(e);
  if (const char* e = getenv("UVA_WGRAD_STREAM")) h->wgrad_stream = atoi(e) != 0;
  if (const char* e = getenv("UVA_WGRAD_SIDE_MULT")) { int v = atoi(e); if (v >= 1 && v <= 64) h->wgrad_side_mult = v; }
  if (const char* e = getenv("UVA_FINAL_TC")) h->final_tc = atoi(e) != 0;
  if (const char* e = getenv("UVA_TRAIN_GRAPH")) h->train_graph = atoi(e) != 0;
  if (const char* e = getenv("UVA_TRAIN_GRAPH_ROWS")) h->train_graph_rows = atol(e);
  if (const char* e = getenv("UVA_FUSE_COLSUM")) h->fuse_colsum = atoi(e) != 0;
  if (const char* e = getenv("UVA_FUSE_GATE_LN")) h->fuse_gate_ln = atoi(e) != 0;
  if (const char* e = getenv("UVA_EPI_COLSUM")) h->epi_colsum = atoi(e) != 0;
  if (const char* e = getenv("UVA_PDL_BIG")) h->pdl_big = atoi(e) != 0;
  if (const char* e = getenv("UVA_PDL_TRAIN")) h->pdl_train = atoi(e) != 0;
  if (const char* e = getenv("UVA_FUSE_LNPRE")) h->fuse_lnpre = atoi(e) != 0;
  if (const char* e = getenv("UVA_GRID_MULT")) { int v = atoi(e); if (v >= 1 && v <= 64) h->grid_mult = v; }
  CUDA_TRY(cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking));
  CUDA_TRY(cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
  CUDA_TRY(cudaStreamCreateWithFlags(&h->hoist_st, cudaStreamNonBlocking));
  CUDA_TRY(cudaEventCreateWithFlags(&h->ev_hfork, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&h->ev_hdone, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&h->ev_wfork, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&h->ev_w1, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&h->ev_w2, cudaEventDisableTiming));
  h->gen = create_schedule(d.T_train, d.T_gen);
  h->train = create_schedule(d.T_train, d.T_train);
  int rc = upload_tables(h->train, h->dtrain, &h->d_tables);
  if (rc != UVA_OK) { delete h; return rc; }
  {
    std::vector<int> ts(d.T_gen);
    for (int k = 0; k < d.T_gen; ++k) ts[k] = h->gen.timestep_map[d.T_gen - 1 - k];  // GD:519 order, RS:122-127 map
    CUDA_TRY(cudaMalloc(&h->d_tsteps, ts.size() * sizeof(int)));
    CUDA_TRY(cudaMemcpy(h->d_tsteps, ts.data(), ts.size() * sizeof(int), cudaMemcpyHostToDevice));
    std::vector<StepCoef> kc(d.T_gen);
    for (int i = 0; i < d.T_gen; ++i) {  // .float() of the float64 tables at use (GD:901)
      kc[i].sqrt_recip = (float)h->gen.sqrt_recip_alphas_cumprod[i];
      kc[i].sqrt_recipm1 = (float)h->gen.sqrt_recipm1_alphas_cumprod[i];
      kc[i].coef1 = (float)h->gen.posterior_mean_coef1[i];
      kc[i].coef2 = (float)h->gen.posterior_mean_coef2[i];
      kc[i].min_log = (float)h->gen.posterior_log_variance_clipped[i];
      kc[i].max_log = (float)h->gen.log_betas[i];
      kc[i].nonzero = (i != 0) ? 1 : 0;
    }
    CUDA_TRY(cudaMalloc(&h->d_coef, kc.size() * sizeof(StepCoef)));
    CUDA_TRY(cudaMemcpy(h->d_coef, kc.data(), kc.size() * sizeof(StepCoef), cudaMemcpyHostToDevice));
  }

  // sinusoidal features for t = 0..T_train-1 (DL:110-134): cos || sin of t * freqs (fp32 args)
  {
    const int half = kFreq / 2;
    std::vector<float> f((size_t)d.T_train * kFreq);
    const float c = (float)(-std::log(10000.0));
    for (int t = 0; t < d.T_train; ++t)
      for (int i = 0; i < half; ++i) {
        const float fr = expf((c * (float)i) / (float)half);
        const float arg = (float)t * fr;
        f[(size_t)t * kFreq + i] = (float)std::cos((double)arg);
        f[(size_t)t * kFreq + half + i] = (float)std::sin((double)arg);
      }
    CUDA_TRY(cudaMalloc(&h->feat, f.size() * sizeof(float)));
    CUDA_TRY(cudaMemcpy(h->feat, f.data(), f.size() * sizeof(float), cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMalloc(&h->temb, (size_t)d.T_train * d.W * sizeof(float)));
    CUDA_TRY(cudaMalloc(&h->temb_hidden, (size_t)d.T_train * d.W * sizeof(float)));
    CUDA_TRY(cudaMalloc(&h->temb_pre, (size_t)d.T_train * d.W * sizeof(float)));
    if (d.mode == UVA_MODE_BF16) {
      CUDA_TRY(cudaMalloc(&h->temb_hidden_b, (size_t)d.T_train * d.W * sizeof(bf16)));
      CUDA_TRY(cudaMalloc(&h->temb_pre_b, (size_t)d.T_train * d.W * sizeof(bf16)));
      CUDA_TRY(cudaMalloc(&h->feat_b, f.size() * sizeof(bf16)));
      k_convert<bf16><<<grid_for((int64_t)f.size(), 256, h->num_sms), 256>>>(h->feat, h->feat_b, (int64_t)f.size());
      CUDA_TRY(cudaDeviceSynchronize());
    }
  }

  // fp32 parameter layout
  {
    const size_t W = d.W, C = d.C, Z = d.Z;
    size_t o = 0;
    auto take = [&](size_t n) { size_t r = o; o += (n + 63) / 64 * 64; return r; };
    h->o_te0w = take(W * kFreq); h->o_te0b = take(W); h->o_te2w = take(W * W); h->o_te2b = take(W);
    h->o_condw = take(W * Z); h->o_condb = take(W); h->o_inpw = take(W * C); h->o_inpb = take(W);
    for (int b = 0; b < d.depth; ++b) {
      h->o_lnw.push_back(take(W)); h->o_lnb.push_back(take(W));
      h->o_w1.push_back(take(W * W)); h->o_b1.push_back(take(W));
      h->o_w2.push_back(take(W * W)); h->o_b2.push_back(take(W));
    }
    h->o_adaw = take((size_t)h->nmod * W); h->o_adab = take(h->nmod);
    h->o_linw = take(2 * C * W); h->o_linb = take(64);
    h->wf_floats = o;
    CUDA_TRY(cudaMalloc(&h->wf, o * sizeof(float)));
    CUDA_TRY(cudaMemset(h->wf, 0, o * sizeof(float)));
  }
  if (d.mode == UVA_MODE_BF16) {
    const size_t W = d.W, Z = d.Z;
    size_t o = 0;
    auto take = [&](size_t n) { size_t r = o; o += (n + 511) / 512 * 512; return r; };
    h->ob_condw = take(W * Z); h->ob_adaw = take((size_t)h->nmod * W); h->ob_linw = take(32 * W);
    h->ob_adaw_p = take((size_t)h->nmod * W);
    h->ob_te2w = take(W * W);
    CUDA_TRY(cudaMalloc(&h->adab_p, (size_t)h->nmod * sizeof(float)));
    CUDA_TRY(cudaMalloc(&h->lnpre_bar, (size_t)kLnPrePanels * 2 * sizeof(unsigned int)));
    CUDA_TRY(cudaMemset(h->lnpre_bar, 0, (size_t)kLnPrePanels * 2 * sizeof(unsigned int)));
    for (int b = 0; b < d.depth; ++b) {
      h->ob_w1.push_back(take(W * W)); h->ob_w2.push_back(take(W * W));
    }
    h->wb_elems = o;
    CUDA_TRY(cudaMalloc(&h->wb, o * sizeof(bf16)));
    CUDA_TRY(cudaMemset(h->wb, 0, o * sizeof(bf16)));
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CUDA_TRY(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (!fn || qres != cudaDriverEntryPointSuccess) { delete h; return fail(UVA_ERR_CUDA, "cuTensorMapEncodeTiled not available"); }
    h->encode = (PFN_encodeTiled)fn;
  }
  *out = h;
  return UVA_OK;
}

extern "C" void uva_head_destroy(UvaHead* h) {
  if (!h) return;
  cudaFree(h->d_tables); cudaFree(h->feat); cudaFree(h->temb); cudaFree(h->temb_hidden); cudaFree(h->temb_pre);
  cudaFree(h->temb_hidden_b); cudaFree(h->temb_pre_b); cudaFree(h->feat_b);
  cudaFree(h->wf); cudaFree(h->wb); cudaFree(h->adab_p); cudaFree(h->d_tsteps); cudaFree(h->d_coef); cudaFree(h->lnpre_bar);
  for (auto& g : h->graphs)
    if (g.exec) cudaGraphExecDestroy(g.exec);
  for (auto& g : h->tgraphs)
    if (g.exec) cudaGraphExecDestroy(g.exec);
  for (auto e : h->gsync.ev) cudaEventDestroy(e);
  prenet_free(h->prenet);
  if (h->side) cudaStreamDestroy(h->side);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  if (h->hoist_st) cudaStreamDestroy(h->hoist_st);
  if (h->ev_hfork) cudaEventDestroy(h->ev_hfork);
  if (h->ev_wfork) cudaEventDestroy(h->ev_wfork);
  if (h->ev_w1) cudaEventDestroy(h->ev_w1);
  if (h->ev_w2) cudaEventDestroy(h->ev_w2);
  if (h->ev_hdone) cudaEventDestroy(h->ev_hdone);
  delete h;
}

// ----------------------------------------------------------------------------------------------
// fused AdamW + EMA step (optim.cuh)
// ----------------------------------------------------------------------------------------------
extern "C" int uva_adamw_ema_step(int32_t n_tensors, float* const* params, const float* const* grads,
                                  float* const* exp_avg, float* const* exp_avg_sq, float* const* ema,
                                  const int64_t* numel, const float* weight_decay, const UvaAdamWHyper* hp,
                                  void* stream) {
  if (n_tensors < 0 || !params || !grads || !exp_avg || !exp_avg_sq || !numel || !weight_decay || !hp)
    return fail(UVA_ERR_INVALID, "adamw_ema_step: null argument");
  if (hp->step < 1) return fail(UVA_ERR_INVALID, "adamw_ema_step: step must be >= 1 (it is the 1-based count of this update)");
  if (!(hp->beta1 >= 0.f && hp->beta1 < 1.f && hp->beta2 >= 0.f && hp->beta2 < 1.f && hp->eps >= 0.f))
    return fail(UVA_ERR_INVALID, "adamw_ema_step: bad hyper-parameters");
  cudaStream_t st = (cudaStream_t)stream;
  int num_sms = 148;
  {
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    CUDA_TRY(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
  }
  for (int t0 = 0; t0 < n_tensors; t0 += OPT_MAX_TENSORS) {
    AdamWBatch b = {};
    b.count = (n_tensors - t0 < OPT_MAX_TENSORS) ? n_tensors - t0 : OPT_MAX_TENSORS;
    int64_t biggest = 0;
    for (int i = 0; i < b.count; ++i) {
      const int t = t0 + i;
      if (!params[t] || !grads[t] || !exp_avg[t] || !exp_avg_sq[t] || numel[t] < 0)
        return fail(UVA_ERR_INVALID, "adamw_ema_step: tensor %d has a null pointer or a negative size", t);
      b.p[i] = params[t]; b.g[i] = grads[t]; b.m[i] = exp_avg[t]; b.v[i] = exp_avg_sq[t];
      b.ema[i] = ema ? ema[t] : nullptr;
      b.n[i] = numel[t]; b.wd[i] = weight_decay[t];
      if (numel[t] > biggest) biggest = numel[t];
    }
    b.lr = hp->lr; b.beta1 = hp->beta1; b.beta2 = hp->beta2; b.eps = hp->eps; b.ema_decay = hp->ema_decay;
    b.grad_scale = hp->grad_scale;
    b.bc1 = (float)(1.0 - std::pow((double)hp->beta1, (double)hp->step));
    b.bc2_sqrt = (float)std::sqrt(1.0 - std::pow((double)hp->beta2, (double)hp->step));
    int bx = (int)((biggest / 4 + 255) / 256);
    if (bx > 2 * num_sms) bx = 2 * num_sms;
    if (bx < 1) bx = 1;
    k_adamw_ema<<<dim3((unsigned)bx, (unsigned)b.count), 256, 0, st>>>(b);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(UVA_ERR_CUDA, "adamw_ema_step: launch failed: %s", cudaGetErrorString(e));
  }
  return UVA_OK;
}

#include "head_infer.inc"
#include "head_small.inc"
#include "head_sample.inc"
#include "collective.inc"
#include "head_train.inc"
#include "head_prenet.inc"

// ----------------------------------------------------------------------------------------------
// uva_head_net_forward_static: SimpleMLPAdaLN.forward for a small batch WITHOUT a caller workspace and without any
// allocation.  No counterpart in the reference: added for callers that cannot
// call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The scratch (the InferWs buffers of head_infer.inc: cemb, ys, mod, h, a, u, net output, bf16 operand rows, LayerNorm
// statistics) lives in ONE file-scope device array of kStaticScratchBytes = 16 MiB, which every GPU that loads the
// library holds until the process exits.  That bounds the batch: the buffers of round_up(N, 128) rows must fit, i.e.
// N <= 256 rows in bf16 mode and N <= 128 rows in fp32 mode at the reference geometry (W = 1024, depth = 6, Z = 768).
//
// In host code `g_static_scratch` is a host-side stand-in, not a device address: the device address is taken once per
// device, with that device current, through cudaGetSymbolAddress, and only that pointer is handed to the kernels.
// Every call (of every handle, on every stream) shares the array and it keeps the last call's contents: nothing in it
// has to start at zero for the kernels of this path (each buffer is written before it is read), the LayerNorm
// statistics and the completion counters are cleared anyway; two calls that overlap on the GPU (two streams, two
// threads, two handles) RACE on it - serialise them on one stream.
// ----------------------------------------------------------------------------------------------
static constexpr size_t kStaticScratchBytes = (size_t)16 << 20;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

static constexpr int kStaticMaxDevices = 64;
static void* g_static_ptr[kStaticMaxDevices] = {};  // device address of g_static_scratch, per device ordinal
static std::mutex g_static_mu;

// device address of the scratch array on the current device (looked up once per device)
static int static_scratch_ptr(void** out) {
  int dev = -1;
  CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kStaticMaxDevices) return fail(UVA_ERR_INVALID, "net_forward_static: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_static_scratch);
    if (e != cudaSuccess || !p)
      return fail(UVA_ERR_CUDA, "net_forward_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
    g_static_ptr[dev] = p;
  }
  *out = g_static_ptr[dev];
  return UVA_OK;
}

extern "C" int64_t uva_head_static_max_rows(const UvaHead* h) {
  if (!h) return 0;
  int64_t best = 0;
  for (int64_t R = UG_BM; R <= 4096; R += UG_BM) {
    if (carve_infer(h, nullptr, R).bytes > kStaticScratchBytes) break;
    best = R;
  }
  return best;
}

extern "C" int uva_head_net_forward_static(UvaHead* h, const float* x, const int64_t* t, const float* c, int64_t N,
                                           float* out, void* stream) {
  NvtxRange nvtx_("uva_head_net_forward_static");
  if (!h || !x || !t || !c || !out) return fail(UVA_ERR_INVALID, "net_forward_static: null argument");
  if (!h->weights_loaded) return fail(UVA_ERR_STATE, "net_forward_static: weights not loaded");
  if (N <= 0) return UVA_OK;
  const UvaHeadDesc& d = h->d;
  const int64_t R = (N + UG_BM - 1) / UG_BM * UG_BM;
  const size_t need = carve_infer(h, nullptr, R).bytes;
  if (need > kStaticScratchBytes)
    return fail(UVA_ERR_INVALID, "net_forward_static: N = %lld rows need %zu bytes of scratch, the static array holds %zu (at most %lld rows for this head; use uva_head_net_forward with a workspace)",
                (long long)N, need, kStaticScratchBytes, (long long)uva_head_static_max_rows(h));
  void* base = nullptr;
  UVA_TRY(static_scratch_ptr(&base));
  InferWs w = carve_infer(h, base, R);
  cudaStream_t st = (cudaStream_t)stream;
  // shared with every earlier call: the small accumulator-like regions start from zero
  CUDA_TRY(cudaMemsetAsync(w.stats, 0, (size_t)((uint8_t*)base + w.bytes - (uint8_t*)w.stats), st));
  UVA_TRY(compute_cemb(h, w, c, 0, 0, N, st));
  UVA_TRY(run_stack_mode(h, w, N, x, 0, t, 0, st));
  CUDA_TRY(cudaMemcpy2DAsync(out, (size_t)2 * d.C * 4, w.netout, (size_t)kNetLd * 4, (size_t)2 * d.C * 4, (size_t)N,
                             cudaMemcpyDeviceToDevice, st));
  return UVA_OK;
}
