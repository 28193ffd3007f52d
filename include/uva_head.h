/*
 * uva_head.h — C ABI of libuva_head.so: the B200-native (sm_100a) per-token diffusion head of
 * Unified Video Action (UVA).
 *
 * The reference (yuz1wan/unified_video_action) has no FFI layer for this path: its "operator
 * API" is four Python nn.Module signatures.  Every entry point below names the reference
 * code it replaces (paths relative to
 * /root/reference/unified_video_action/model/autoregressive/):
 *
 *   DL  = diffusion_loss.py          GD = diffusion/gaussian_diffusion.py
 *   DAL = diffusion_action_loss.py   RS = diffusion/respace.py
 *   DI  = diffusion/__init__.py      DU = diffusion/diffusion_utils.py
 *
 * Conventions
 *   - plain C types only; every `const float*` / `float*` / `void*` argument is a DEVICE pointer
 *     unless the comment says "host"; tensors are dense row-major;
 *   - every call enqueues work on the given cudaStream_t (passed as void*) and returns without
 *     synchronising; the caller owns every buffer, the library owns the opaque handle (packed
 *     weights, schedule tables, timestep-embedding table);
 *   - return value 0 = ok, negative = error; uva_last_error() gives the message (thread local);
 *   - there is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef UVA_HEAD_H_
#define UVA_HEAD_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct UvaHead UvaHead;

enum {
  UVA_OK = 0,
  UVA_ERR_INVALID = -1,   /* bad argument / shape / alignment (checked before any launch) */
  UVA_ERR_CUDA = -2,      /* CUDA runtime or driver error                                  */
  UVA_ERR_STATE = -3,     /* e.g. weights not loaded                                       */
  UVA_ERR_WORKSPACE = -4  /* workspace too small                                           */
};

enum {
  UVA_MODE_FP32 = 0, /* fp32 FFMA GEMMs, fp32 activations (<= 1e-4 rel. vs reference fp32)   */
  UVA_MODE_BF16 = 1  /* tcgen05/TMEM bf16 GEMMs, fp32 accumulate, fp32 residual stream + LN  */
};

/* Geometry of one head = ctor arguments of DiffLoss (DL:11-42) / DiffActLoss (DAL:12-107). */
typedef struct {
  int32_t C;       /* target_channels: net in = C, out = 2C (learn_sigma, DI:21-46)           */
  int32_t Z;       /* z_channels (condition width), multiple of 64                            */
  int32_t W;       /* width, multiple of 256                                                  */
  int32_t depth;   /* number of ResBlocks (DL:222-229)                                        */
  int32_t T_train; /* diffusion_steps of the training process (1000)                          */
  int32_t T_gen;   /* num_sampling_steps: respaced length of the sampling process (100)       */
  int32_t mode;    /* UVA_MODE_*                                                              */
} UvaHeadDesc;

/* Number of parameter tensors of SimpleMLPAdaLN in reference state_dict order (DL:222-235):
 *   time_embed.mlp.{0,2}.{weight,bias}, cond_embed.{weight,bias}, input_proj.{weight,bias},
 *   depth x res_blocks.i.{in_ln.weight, in_ln.bias, mlp.0.weight, mlp.0.bias, mlp.2.weight,
 *   mlp.2.bias, adaLN_modulation.1.weight, adaLN_modulation.1.bias},
 *   final_layer.{linear.weight, linear.bias, adaLN_modulation.1.weight, adaLN_modulation.1.bias}
 * = 8 + 8*depth + 4. */
int32_t uva_head_num_params(int32_t depth);
/* Element count of parameter `index` (same order); -1 when out of range. */
int64_t uva_head_param_numel(const UvaHeadDesc* desc, int32_t index);

/* Replaces DiffLoss.__init__ / create_diffusion (DL:11-42, DI:11-47, GD:102-202, RS:12-112):
 * builds the float64 cosine schedule tables of both processes on the host, uploads fp32 copies,
 * allocates the packed-weight storage.  Needs a CUDA device (current device is used). */
int uva_head_create(const UvaHeadDesc* desc, UvaHead** out);
void uva_head_destroy(UvaHead* h);

/* Host-only schedule tables (GD:157-202 after RS:65-91 respacing), float64, no device needed.
 * which: 0 = sampling process (T_gen entries), 1 = training process (T_train entries).
 * name : "betas", "alphas_cumprod", "alphas_cumprod_prev", "sqrt_alphas_cumprod",
 *        "sqrt_one_minus_alphas_cumprod", "sqrt_recip_alphas_cumprod",
 *        "sqrt_recipm1_alphas_cumprod", "posterior_variance", "posterior_log_variance_clipped",
 *        "posterior_mean_coef1", "posterior_mean_coef2", "timestep_map" (as doubles).
 * Writes min(cap, T) doubles to out (host); returns T or a negative error. */
int uva_schedule_table(int32_t T_train, int32_t T_gen, int32_t which, const char* name,
                       double* out /*host*/, int32_t cap);

/* Replaces nn.Module parameter access for the head: (re)packs the fp32 parameters (device
 * pointers, reference state_dict order, see uva_head_num_params) into the kernel operand copies
 * (bf16 K-major + transposed copies in bf16 mode, fp32 copy in fp32 mode) and rebuilds the
 * timestep-embedding table TimestepEmbedder(t) for t = 0..T_train-1 (DL:97-139).  Call after
 * load_state_dict and after every optimizer step. */
int uva_head_load_weights(UvaHead* h, const float* const* params /*host array of device ptrs*/,
                          int32_t n_params, void* stream);

/* Workspace size for `rows` rows.  what: 0 = net_forward / sample, 1 = loss fwd + bwd. */
size_t uva_head_workspace_bytes(const UvaHead* h, int64_t rows, int32_t what);

/* SimpleMLPAdaLN.forward (DL:261-283): x[N,C], t[N] (original-process timesteps 0..T_train-1),
 * c[N,Z] -> out[N,2C] fp32. */
int uva_head_net_forward(UvaHead* h, const float* x, const int64_t* t, const float* c, int64_t N,
                         float* out, void* ws, size_t ws_bytes, void* stream);

/* The same computation for a small batch with NO workspace argument and NO allocation of any kind (no cudaMalloc,
 * cudaFree or allocating library call).  This entry point has no counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Its scratch is one file-scope `static __device__` array of 16 MiB that
 * every GPU loading the library holds until exit; the buffers of round_up(N, 128) rows must fit in it:
 * N <= uva_head_static_max_rows(h) = 256 rows in bf16 mode / 128 rows in fp32 mode at the reference geometry
 * (W = 1024, depth = 6, Z = 768).  A larger N returns UVA_ERR_INVALID with a message that names the limit.
 * Every call shares the array, which keeps the previous call's contents (the entry point clears what it needs):
 * calls that overlap on the GPU - two streams, two threads, two handles - RACE on it; serialise them on one stream.
 * Same inputs, outputs and results as uva_head_net_forward. */
int uva_head_net_forward_static(UvaHead* h, const float* x, const int64_t* t, const float* c, int64_t N,
                                float* out, void* stream);
int64_t uva_head_static_max_rows(const UvaHead* h);

/* DiffLoss.sample / DiffActLoss.sample after the pre-net (DL:68-90, DAL:207-227) =
 * p_sample_loop (GD:446-541) over p_sample (GD:395-444) / p_mean_variance (GD:260-346) with the
 * wrapped model (RS:115-127) and optionally forward_with_cfg (DL:285-293).
 *   z[N,Z]; x_init[N,C] start noise (for cfg != 1 rows are cond || uncond and the caller passes
 *   cat([noise, noise]) like DL:71-72); step_noise[(k_end-k_begin),N,C]: entry j is the
 *   randn_like drawn at loop iteration k_begin + j (timestep index T_gen-1-k);
 *   teacher (optional, may be NULL) [(k_end-k_begin),N,C]: carried state is replaced by
 *   teacher[j] before each step (per-step parity tests);
 *   trace (optional, may be NULL) [(k_end-k_begin),N,C]: receives the state after each step;
 *   runs loop iterations k_begin..k_end-1 (0..T_gen for the whole chain); out[N,C]. */
/* step_noise may be NULL once uva_head_set_noise_seed(h, seed, offset, 1) has been called: the
 * randn_like of GD:431 is then drawn inside the DDPM-update kernel from Philox4x32-10 keyed by `seed`,
 * counter = (element index / 4, loop iteration k, offset), Box-Muller on 24-bit uniforms; the caller
 * advances `offset` per sample() call.  Element index = row * C + c in the caller's [N, C] layout, so the
 * samples do not depend on chunking.  This stream is NOT torch's randn stream: it is the mode for
 * production sampling (no [T_gen, N, C] noise tensor in HBM); parity runs pass step_noise.
 * uva_philox_normal writes the same numbers to memory: out[j * per_step + e] = draw e of step step0 + j. */
int uva_head_set_noise_seed(UvaHead* h, uint64_t seed, uint32_t offset, int32_t enable);
int uva_philox_normal(uint64_t seed, uint32_t offset, int32_t step0, int64_t per_step, int64_t n,
                      float* out, void* stream);

int uva_head_sample(UvaHead* h, const float* z, const float* x_init, const float* step_noise,
                    const float* teacher, float* trace, float temperature, float cfg_scale,
                    int32_t clip_denoised, int32_t k_begin, int32_t k_end, int64_t N, float* out,
                    void* ws, size_t ws_bytes, void* stream);

/* DiffLoss.forward / DiffActLoss.forward after the pre-net (DL:44-66, DAL:145-166) =
 * training_losses (GD:746-818: q_sample GD:220-236, MSE + VLB GD:713-744, DU:10-73) + masked
 * mean.  target[N,C], z[N,Z], mask[N] or NULL (plain mean, DAL:162-164), t[N] int64 (the draw
 * of DL:51), noise[N,C] (the draw of GD:761).  Writes loss_out[1] and, if row_loss != NULL,
 * the per-row loss [N].  Activations needed by the backward stay in ws. */
int uva_head_loss_fwd(UvaHead* h, const float* target, const float* z, const float* mask,
                      const int64_t* t, const float* noise, int64_t N, float* loss_out,
                      float* row_loss, void* ws, size_t ws_bytes, void* stream);

/* Backward of uva_head_loss_fwd (autograd of DL:44-66): dloss[1] is the incoming gradient of
 * the scalar loss.  Writes dz[N,Z]; ACCUMULATES into dparams (host array of device fp32
 * pointers, reference order; an entry may be NULL to skip).  Must follow a loss_fwd on the same
 * ws with the same inputs. */
int uva_head_loss_bwd(UvaHead* h, const float* dloss, const float* target, const float* z,
                      const float* mask, const int64_t* t, const float* noise, int64_t N,
                      float* dz, float* const* dparams /*host array of device ptrs*/,
                      void* ws, size_t ws_bytes, void* stream);

/* MAR decoder tail fused into the head.  The reference computes the head's condition as
 *   z = decoder_norm(x)[:, buffer_size_text:] + (diffusion_temporal_embed[t] + diffusion_spatial_embed[s])
 * (mar_con_unified.py:702-724) and hands the fp32 z to DiffLoss.sample.  uva_head_sample_from_decoder takes the
 * decoder output BEFORE decoder_norm instead: row r of the sampler is token tok = g % (T*S) of sample b = g / (T*S)
 * with g = row_index[r] (or r when row_index == NULL; the gather z[mask_to_pred.nonzero()] of MAR:1090), read from
 * x_dec[b, skip + tok, :]; LayerNorm + pos-embed run as the prologue of the cond_embed GEMM and write its bf16
 * operand directly, so the fp32 z (403 MB per call at BASELINE config 3) never exists.  All other arguments as
 * uva_head_sample.  uva_decoder_tail_rows writes the same rows as fp32 (for consumers that need z itself). */
typedef struct {
  const float* ln_w;       /* [Z] decoder_norm.weight */
  const float* ln_b;       /* [Z] decoder_norm.bias */
  const float* temporal;   /* [T, Z] diffusion_temporal_embed */
  const float* spatial;    /* [S, Z] diffusion_spatial_embed */
  float eps;               /* decoder_norm.eps (1e-6 for mar_base, MAR:1201) */
  int32_t T, S;            /* frames, tokens per frame */
  int64_t tokens_per_sample;  /* L: tokens per sample in x_dec (buffer_size_text + T*S with CLIP text tokens) */
  int64_t skip;            /* leading tokens dropped (buffer_size_text, MAR:705-707) */
} UvaDecoderTail;
int uva_head_sample_from_decoder(UvaHead* h, const float* x_dec, const UvaDecoderTail* tail, const int64_t* row_index,
                                 const float* x_init, const float* step_noise, const float* teacher, float* trace,
                                 float temperature, float cfg_scale, int32_t clip_denoised, int32_t k_begin,
                                 int32_t k_end, int64_t N, float* out, void* ws, size_t ws_bytes, void* stream);
int uva_decoder_tail_rows(const float* x_dec, const UvaDecoderTail* tail, const int64_t* row_index, int64_t N,
                          int32_t Z, float* z_out, void* stream);

/* DiffActLoss pre-net fused into the library.  The reference runs, in front of the action /
 * proprioception head (diffusion_action_loss.py:41-61, 112-124 == 169-181, act_model_type "conv_fc"):
 *   z [B, n_frames * 256, Z] -> n_frames 16 x 16 token maps -> Conv2d(Z, Z, 3, padding 1) + ReLU + AdaptiveAvgPool2d(4, 4)
 *   -> Linear(16 Z, Z) + ReLU + Linear(Z, Z) -> Linear(n_frames -> 16) over time -> Linear + ReLU + Linear -> [B, 16, Z]
 * as ~25 cuDNN / cuBLAS / elementwise launches.  Here the convolution is ONE implicit-GEMM launch of the CTA-pair tcgen05
 * kernel (operands gathered by 4D TMA boxes from the token tensor, padding = TMA zero fill, bias + ReLU + average pool in
 * the epilogue) and the Linears are tcgen05 GEMMs; the backward uses the same cores (data gradient = the gather with
 * flipped taps, weight gradient = gathered B operand).  bf16 operands, fp32 accumulation; handles created in bf16 mode only.
 *   params / dparams: 12 device pointers in the reference's state_dict order - conv.0.{weight [Z,Z,3,3], bias},
 *   fc.0.{weight [Z,16Z], bias}, fc.2.{weight, bias}, interpolate.{weight [16,n_frames], bias}, refine.0.{weight, bias},
 *   refine.2.{weight, bias}; gradients are ACCUMULATED into dparams (null entries are skipped), dz is written.
 *   uva_head_prenet_backward must follow a uva_head_prenet_forward(train = 1) on the same workspace. */
int32_t uva_head_prenet_num_params(void);
int uva_head_prenet_load_weights(UvaHead* h, const float* const* params, int32_t n_params, int32_t n_frames,
                                 int32_t n_actions, void* stream);
size_t uva_head_prenet_workspace_bytes(const UvaHead* h, int64_t B, int32_t train);
int uva_head_prenet_forward(UvaHead* h, const float* z, int64_t B, float* out, void* ws, size_t ws_bytes, int32_t train,
                            void* stream);
int uva_head_prenet_backward(UvaHead* h, const float* dout, int64_t B, float* dz, float* const* dparams, void* ws,
                             size_t ws_bytes, void* stream);

/* Fused AdamW + EMA update of a list of fp32 parameter tensors in ONE pass.  Replaces
 * torch.optim.AdamW.step() as configured by the reference (policy/unified_video_action_policy.py:325-360:
 * decoupled weight decay, 0 on 1-D tensors and biases) followed by EMAModel.step()
 * (model/autoregressive/ema_model.py:58-89).  Per tensor t (host arrays of device pointers):
 *   g = grads[t] * grad_scale;  p *= 1 - lr * weight_decay[t];  m = beta1 m + (1 - beta1) g;
 *   v = beta2 v + (1 - beta2) g^2;  p -= lr / (1 - beta1^step) * m / (sqrt(v) / sqrt(1 - beta2^step) + eps);
 *   ema[t] = ema[t] * ema_decay + p * (1 - ema_decay)      (skipped when ema == NULL or ema[t] == NULL)
 * step is the 1-based count of this update.  Call uva_head_load_weights afterwards to refresh the packed
 * operands. */
typedef struct {
  float lr, beta1, beta2, eps, ema_decay, grad_scale;
  int32_t step;
} UvaAdamWHyper;
int uva_adamw_ema_step(int32_t n_tensors, float* const* params, const float* const* grads,
                       float* const* exp_avg, float* const* exp_avg_sq, float* const* ema,
                       const int64_t* numel, const float* weight_decay, const UvaAdamWHyper* hp, void* stream);

/* The one collective of the path: data-parallel training all-reduces the head gradients.
 * The reference leaves this to DDP's bucketed reducer under Accelerate
 * (workspace/train_unified_video_action_workspace.py:290-292); here the fused backward writes every
 * gradient of a head into ONE flat fp32 buffer, which is reduced in place with one ncclAllReduce on
 * NVLink / NVSwitch (ncclAvg when average != 0, else ncclSum), enqueued on `stream` behind the backward.
 * NCCL is bound at run time (the libnccl.so.2 already mapped into the process, else the system one); the
 * calls fail with UVA_ERR_STATE when none can be loaded.  Sampling shards rows and needs no collective.
 *   uva_nccl_unique_id  : rank 0 fills 128 bytes (host) that the caller broadcasts to the other ranks
 *                         (any out-of-band channel: torch.distributed store, MPI, a file);
 *   uva_nccl_comm_create: collective over all ranks; the current CUDA device is the rank's GPU;
 *   uva_head_allreduce_grads: flat = device pointer to `count` fp32 values, reduced in place;
 *   uva_head_allreduce_grads_bf16: the same buffer crosses the link as bf16 (half the bytes): rounded into
 *                         `scratch` (device, count * 2 bytes), reduced there by NCCL, widened back into flat. */
int32_t uva_nccl_version(void);
int uva_nccl_unique_id(void* out128 /*host*/);
int uva_nccl_comm_create(int32_t world_size, int32_t rank, const void* id128 /*host*/, void** comm /*host*/);
int uva_nccl_comm_destroy(void* comm);
int uva_head_allreduce_grads(void* comm, float* flat, int64_t count, int32_t average, void* stream);
int uva_head_allreduce_grads_bf16(void* comm, float* flat, void* scratch, int64_t count, int32_t average, void* stream);
/* The same collective from INSIDE the backward: while a communicator is set, uva_head_loss_bwd (calls with more than
 * 1024 rows) all-reduces every contiguous range of the caller's flat gradient buffer on `comm_stream` as soon as the
 * kernels that write it are enqueued (final linear, each block's in_ln + MLP, each adaLN Linear, the embedders:
 * 3 * depth + 3 collectives), so that the all-reduce runs behind the rest of the backward instead of after it.  The
 * gradient tensors of such a range must be one contiguous span (gaps <= 3 floats, zero-filled).  comm == NULL: off.
 * uva_head_grad_sync_launched: collectives the last loss_bwd enqueued (0: small batch, reduce afterwards). */
int uva_head_set_grad_sync(UvaHead* h, void* comm, void* comm_stream, int32_t average);
int32_t uva_head_grad_sync_launched(const UvaHead* h);

/* Joint training step (MAR:770-784): the video head (16 384 rows) and the action head (256 rows) are independent
 * and run on two streams.  The video head's persistent GEMM grids would take every SM, so the small head could
 * only run between them; with a reserve of `sms` SMs (e.g. 16) the large head's loss_fwd / loss_bwd size their
 * grids for num_sms - sms and the small head's kernels run beside them.  0 = use every SM (default).  Sampling
 * and calls with <= 1024 rows are not affected. */
int uva_head_set_train_reserve(UvaHead* h, int32_t sms);

/* Number of kernel launches this handle has enqueued so far (bench.py's gpu_launches). */
int64_t uva_head_launch_count(const UvaHead* h);

/* Optional per-kernel timing for the roofline report: while enabled, every tcgen05 GEMM launch is
 * bracketed by CUDA events on the launch stream.  timing_read synchronises on the recorded events of
 * one category (0 AdaLN-modulation GEMM, 1 MLP GEMM 1, 2 MLP GEMM 2, 3 final linear, 4 cond_embed),
 * returns their summed duration and count, and forgets them.  While timing is enabled the sampler
 * runs its row chunks one after the other (no two-lane overlap), so each timed kernel has the GPU
 * to itself. */
int uva_head_timing_enable(UvaHead* h, int32_t enable);
int uva_head_timing_read(UvaHead* h, int32_t category, double* total_ms, int64_t* count);

/* Unit check of the tcgen05 GEMM core (bf16 operands, fp32 result), D[M,N]:
 *   layout 0: A[M,K], B[N,K] (forward form, D = A.B^T + bias; block_n 32 or 256, K multiple of 8);
 *   layout 1: A[M,K], B[K,N] (data-gradient form, B MN-major; block_n 256);
 *   layout 2: A[K,M], B[K,N] (weight-gradient form, both MN-major, split-K with vector reductions
 *             ADDED into D, which the caller zero-fills; block_n 256, bias ignored);
 *   layout 8, 9, 10: forms 0, 1, 2 on the CTA-pair (cta_group::2) kernel, 256 x 256 tiles;
 *   layout 16, 17, 18: the same on clusters of two pairs with multicast A boxes (N multiple of 512).
 * N multiple of block_n. */
int uva_gemm_bf16_selftest(const void* A, const void* B, const float* bias, float* D, int64_t M,
                           int32_t N, int32_t K, int32_t block_n, int32_t layout, void* stream);

const char* uva_last_error(void);
const char* uva_build_info(void);

#ifdef __cplusplus
}
#endif
#endif /* UVA_HEAD_H_ */
