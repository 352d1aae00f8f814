/*
 * rmb_b200.h -- C ABI of the B200-native batched policy-rollout engine for RoboManipBaselines' ACT hot path.
 *
 * The reference (RoboManipBaselines @ 219a5496) is pure Python and has no FFI; each entry point below replaces the
 * Python function named next to it (paths relative to the reference root). A maintainer binds this library with
 * ctypes. Conventions:
 *   - every function returns 0 on success, <0 on error; rmb_last_error() gives the message
 *   - all `d_*` pointers are DEVICE pointers owned by the caller (e.g. torch tensors); `stream` is a cudaStream_t
 *   - all work is enqueued on `stream`; no hidden synchronisation except in rmb_act_engine_create/finalize
 *   - one engine per GPU, not thread-safe per handle; sm_100a only, there is no CPU fallback
 */
#ifndef RMB_B200_H_
#define RMB_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rmb_act_engine rmb_act_engine_t;

/* model_meta_info["policy"]["args"] of policy/act/TrainAct.py:45-57 + dims that DETRVAE keeps as class attributes
 * (third_party/act/detr/models/detr_vae.py:37-46) -- carried per engine instance here. */
typedef struct rmb_act_config {
  int32_t max_batch;        /* episodes per forward (workspace is sized for this) */
  int32_t num_cams;         /* len(camera_names) */
  int32_t img_h, img_w;     /* camera resolution, multiples of 32 (480 x 640 in MujocoUR5eCable) */
  int32_t state_dim;        /* DETRVAE.STATE_DIM */
  int32_t action_dim;       /* DETRVAE.ACTION_DIM */
  int32_t hidden_dim;       /* 512 (only value supported) */
  int32_t dim_feedforward;  /* multiple of 64 (3200) */
  int32_t nheads;           /* 8 (head dim must be 64) */
  int32_t enc_layers;       /* 4 */
  int32_t dec_layers;       /* 7 -- only layer 0 is live at inference (detr_vae.py:142 takes hs[0]) */
  int32_t chunk_size;       /* num_queries (100) */
  int32_t device;           /* CUDA device ordinal */
  /* MT-ACT (policy/mt_act; the task-conditioned DETRVAE lives in the un-vendored third_party/roboagent: ARCHITECTURE
   * HYPOTHESIS, parity unpinned): */
  int32_t multi_task;       /* 1: ResNet blocks of layers 2..4 are FiLM-modulated by the projected task embedding */
  int32_t task_emb_dim;     /* 384 (all-MiniLM-L6-v2, policy/mt_act/RolloutMtAct.py:54-56); multiple of 64 */
  int32_t task_token;       /* 1: the projected task embedding is also a third prefix token (L = 3 + ...). Default 0:
                             * RolloutMtAct.draw_plot reshapes correlation_mat[2:, 1] to (15, 20 Ncam)
                             * (policy/mt_act/RolloutMtAct.py:147-155), which only fits L = 2 + 300 Ncam */
} rmb_act_config_t;

/* Last error message of the library (per process). */
const char* rmb_last_error(void);
/* Library version string and a cheap "is this a B200" probe (returns 0 if device `dev` is sm_100). */
const char* rmb_version(void);
/* 16-bit storage / tensor-core operand type the library was built with: "fp16" (default) or "bf16". Every `act`
 * buffer below holds this type; accumulation is always fp32. */
const char* rmb_act_dtype(void);
int rmb_check_device(int dev);
/* Synchronises the current device and returns the kernels' fault flag: 0 = none, 1 = an mbarrier wait inside a
 * TMA/tcgen05 pipeline timed out (the kernel drained with garbage results instead of hanging), <0 = CUDA error.
 * clear != 0 resets the flag. Results produced while the flag is set are invalid. */
int rmb_device_fault(int clear);

/* Saturation diagnostics (fp16 build: conversions clamp to +-65504 instead of overflowing to inf). With check mode on,
 * every 16-bit tensor a forward of this engine stores is scanned right after the launch that wrote it and elements at
 * the clamp value (bf16 build: inf / nan) are added to a per-process device counter. Off by default (the timed path
 * launches nothing extra). rmb_saturation_count synchronises the device and returns the counter (clear != 0 resets). */
int rmb_act_set_check_saturation(rmb_act_engine_t* e, int on);
int64_t rmb_saturation_count(int clear);

/* ---- engine lifetime: replaces ACTPolicy.__init__ (third_party/act/policy.py:10-16) + RolloutBase.load_ckpt
 *      (robo_manip_baselines/common/base/RolloutBase.py:376-385) */
int rmb_act_engine_create(const rmb_act_config_t* cfg, rmb_act_engine_t** out);
void rmb_act_engine_destroy(rmb_act_engine_t* e);
/* Hand over one tensor of the reference checkpoint (`policy_*.ckpt` state dict, key e.g.
 * "model.transformer.encoder.layers.0.linear1.weight"), fp32, host memory, `numel` elements. Unused keys are accepted
 * and ignored (the CVAE encoder, is_pad_head, decoder layers 1..6). */
int rmb_act_load_tensor(rmb_act_engine_t* e, const char* name, const float* host_data, int64_t numel);
/* Fold FrozenBatchNorm2d into the conv weights, precompute the input-independent tensors (sine position table,
 * latent token, decoder query projection), upload 16-bit weights. Must be called once after all tensors are loaded. */
int rmb_act_finalize(rmb_act_engine_t* e);

/* ---- MT-ACT: `task_emb=` of ACTPolicy.__call__ (policy/mt_act/RolloutMtAct.py:84-85). d_task_emb float32
 * [batch, task_emb_dim], one sentence embedding per episode. Computes proj_text_emb and the FiLM gamma / beta of every
 * conditioned ResNet block ONCE; they stay resident in the engine until the next call (the reference caches the
 * embedding per task text the same way, RolloutMtAct.py:115-120), so a policy step launches nothing extra. */
int rmb_act_set_task_embedding(rmb_act_engine_t* e, const float* d_task_emb, int batch, void* stream);

/* ---- ACTPolicy.__call__ inference branch (third_party/act/policy.py:18-38) -> DETRVAE.forward
 *      (third_party/act/detr/models/detr_vae.py:89-150), for `batch` episodes at once.
 * image_format 0: d_images = uint8  [batch, num_cams, H, W, 3]  (env layout; RolloutBase.get_images :479-490 fused)
 * image_format 1: d_images = float32 [batch, num_cams, 3, H, W] in [0,1] (the nn.Module's input)
 * d_qpos  float32 [batch, state_dim] (already normalised); d_a_hat float32 [batch, chunk_size, action_dim] (normalised). */
int rmb_act_forward(rmb_act_engine_t* e, const void* d_images, int image_format, const float* d_qpos, float* d_a_hat,
                    int batch, void* stream);
/* Same forward with a CUDA-event pair recorded around every kernel launch on `stream` (synchronises the stream before
 * returning). Fills out[0..n) in launch order and returns n (<= cap), <0 on error. flops / bytes are the ALGORITHMIC
 * figures of that launch: 2*rows*N*K useful FLOPs; operands read once + result written once. */
typedef struct rmb_launch_prof {
  char name[48];    /* which layer, e.g. "layer2.0.conv1 3x3" */
  char kernel[40];  /* __global__ function, e.g. "gemm_tc_kernel" */
  float ms;         /* device time between the two events */
  double flops;
  double bytes;
} rmb_launch_prof_t;
int rmb_act_forward_profile(rmb_act_engine_t* e, const void* d_images, int image_format, const float* d_qpos,
                            float* d_a_hat, int batch, void* stream, rmb_launch_prof_t* out, int cap);
/* Number of kernels rmb_act_forward launches for one call (for bench.py's gpu_launches). */
int rmb_act_forward_num_launches(const rmb_act_engine_t* e);
/* Algorithmic FLOPs of one forward for `batch` episodes (live graph only). */
double rmb_act_forward_flops(const rmb_act_engine_t* e, int batch);
/* Debug taps: copy an internal activation (fp32 view) of the last forward into d_out. which: 0 = backbone+input_proj
 * tokens [batch, L, 512], 1 = encoder memory [batch, L, 512]. Returns number of floats, <0 on error. */
int64_t rmb_act_debug_read(rmb_act_engine_t* e, int which, float* d_out, int batch, void* stream);

/* Opt-in debug output for RolloutAct.draw_plot (policy/act/RolloutAct.py:36-43,115-125): the head-averaged
 * self-attention map (L, L) of every encoder layer for ONE episode of the batch -- what the reference's forward hook
 * keeps as `correlation_mat` (nn.MultiheadAttention's averaged attn_output_weights[0]). on != 0: every following
 * rmb_act_forward also launches one small kernel per encoder layer; off (default): nothing extra is launched.
 * rmb_act_read_attention_maps copies float32 [enc_layers, L, L] of the last forward to d_out; returns floats, <0 error. */
int rmb_act_set_attention_debug(rmb_act_engine_t* e, int on, int episode);
int64_t rmb_act_read_attention_maps(rmb_act_engine_t* e, float* d_out, void* stream);

/* ---- RolloutAct.infer_policy temporal ensembling + denormalize_data
 *      (robo_manip_baselines/policy/act/RolloutAct.py:68-101; common/utils/DataUtils.py:26-40)
 * d_ring  float32 [batch, C, C, Da]   slot (n mod C) holds the chunk of policy call n
 * d_chunks float32 [batch, C, Da]     chunk predicted at this call (output of rmb_act_forward)
 * d_step  int32 [batch]               per-episode 0-based policy-call counter, advanced by one
 * d_reset uint8 [batch] or NULL       1 = episode restarts at this call (RolloutAct.reset_variables :62-66)
 * d_wtab  float64 [C, C]              wtab[(m-1)*C + r] = exp(-0.01 r)/sum_{s<m} exp(-0.01 s) (numpy-made, see host code)
 * norm_type 0 gaussian: out = d_p0*a + d_p1 (std, mean); 1 limits: out = d_p0*(a - out_min) + d_p1 (scale, min)
 * outputs (each may be NULL): d_action float64 [batch, Da] denormalised, d_action_f32 same in fp32,
 *                             d_action_norm float64 [batch, Da] ensembled but still normalised */
int rmb_ensemble_step(float* d_ring, const float* d_chunks, int32_t* d_step, const uint8_t* d_reset,
                      const double* d_wtab, int norm_type, const double* d_p0, const double* d_p1, double out_min,
                      double* d_action, float* d_action_f32, double* d_action_norm, int batch, int C, int Da,
                      void* stream);

/* ---- normalize_data / denormalize_data (common/utils/DataUtils.py:9-40), fp64 like numpy.
 * normalize: gaussian (type 0): y = (x - p0)/p1 (mean, std); limits (1): y = p1*(x - p0) + out_min (min, scale); y is fp32
 *            (RolloutBase.get_state :474-475 casts to float32).
 * denormalize: gaussian: y = p0*x + p1 (std, mean); limits: y = p0*(x - out_min) + p1 (scale, min); fp64. */
int rmb_normalize(const double* d_x, int norm_type, const double* d_p0, const double* d_p1, double out_min, float* d_y,
                  int64_t rows, int D, void* stream);
int rmb_denormalize(const double* d_x, int norm_type, const double* d_p0, const double* d_p1, double out_min,
                    double* d_y, int64_t rows, int D, void* stream);

/* ---- RolloutBase.get_images (:479-490) + ImageNet normalise (third_party/act/policy.py:20-22) alone:
 * uint8 [n_img, H, W, 3] -> act [n_img, H, W + 8, 4] (4 zero pixels either side, 4th channel zero). */
int rmb_preprocess_images_u8(const uint8_t* d_images, void* d_out_act, int n_img, int H, int W, void* stream);
int rmb_preprocess_images_f32(const float* d_images_chw, void* d_out_act, int n_img, int H, int W, void* stream);

/* ---- kernel-level entry points (used by the parity tests; same kernels the engine launches) */
/* out[M,N] = A[M,K](act) * Wt[N,K]^T(act) + bias (+ residual) (ReLU); flags: bit0 relu, bit1 out fp32, bit2 residual fp32 */
int rmb_gemm(const void* d_A, const void* d_Wt, const float* d_bias, const void* d_residual, void* d_out,
                  int64_t M, int N, int K, int flags, void* stream);
/* LayerNorm(A[M,K] Wt[512,K]^T + bias + residual) in one launch (the encoder's out-projection / FFN down-projection
 * with norm1 / norm2 fused, third_party/act/detr/models/transformer.py:172-178): residual, out fp32 [M,512]; optional
 * 16-bit copies y_act = y and ypos_act = y + pos[row % L] (pos fp32 [L,512]); eps 1e-5 */
int rmb_gemm_residual_layernorm(const void* d_A, const void* d_Wt, const float* d_bias, const float* d_residual,
                                const float* d_gamma, const float* d_beta, const float* d_pos, int L, float* d_out,
                                void* d_y_act, void* d_ypos_act, int64_t M, int K, void* stream);
/* NHWC act conv k in {1,3}, stride in {1,2}, pad k/2; Wt [Cout, k*k*Cin] tap-major; residual/out [B,OH,OW,Cout] act */
int rmb_conv2d_nhwc(const void* d_in, const void* d_Wt, const float* d_bias, const void* d_residual, void* d_out,
                         int B, int H, int W, int Cin, int Cout, int ksize, int stride, int relu, void* stream);
/* ResNet stem fused with its max-pool (the kernel the engine launches): same input / weights as rmb_conv1_stem, out
 * [B, H/4, W/4, 64] act = maxpool3x3s2(relu(conv1 + bias)), bit-identical to rmb_conv1_stem + rmb_maxpool3x3s2_nhwc */
int rmb_stem_pool(const void* d_in_padded, const void* d_Wt, const float* d_bias, void* d_out, int B, int H, int W,
                  void* stream);
/* ResNet stem on the padded NHWC4 input: Wt [64, 7*32] (see gemm_plan.h), out [B, H/2, W/2, 64] act, ReLU */
int rmb_conv1_stem(const void* d_in_padded, const void* d_Wt, const float* d_bias, void* d_out, int B, int H, int W,
                        void* stream);
int rmb_maxpool3x3s2_nhwc(const void* d_in, void* d_out, int B, int H, int W, int C, void* stream);
/* softmax(Q K^T / 8) V for 64-wide heads. Q [B*Lq (or Lq if q_shared), ldq], K [B*Lk, ldk], V [B*Lk, ldv], act;
 * head h reads columns col0 + 64h; out [B*Lq, ld_out] act */
int rmb_attention(const void* d_Q, int ldq, int q_col0, int q_shared, const void* d_K, int ldk, int k_col0,
                       const void* d_V, int ldv, int v_col0, void* d_out, int ld_out, int B, int n_heads, int Lq, int Lk,
                       void* stream);
/* Which attention kernel rmb_attention and the ACT engine launch (csrc/act_engine.cu: 1 two-sweep, 2 one-sweep with P in
 * shared memory, 3 one-sweep with P in TMEM and two query streams per CTA, 4 = 3 with three streams and 64-key blocks,
 * the default). variant 0 = query. Returns the variant now in force, or -1.
 * Captured CUDA graphs keep the kernel they were captured with. */
int rmb_set_attention_variant(int variant);
/* LayerNorm over 512 columns: x fp32 [rows,512]; outputs optional (NULL): y fp32, y act, act(y + pos[row % L]) */
int rmb_layernorm512(const float* d_x, const float* d_gamma, const float* d_beta, const float* d_pos, int L,
                     float* d_y_f32, void* d_y_act, void* d_ypos_act, int64_t rows, int do_norm, void* stream);
/* The same LayerNorm over 16-bit rows: x act [rows,512] (the pre-norm sums x + sublayer(x) that the residual GEMMs of
 * the fp16 build write in 16 bits; transformer.py:166-179 norm1 / norm2); statistics and arithmetic in fp32. */
int rmb_layernorm512_act(const void* d_x_act, const float* d_gamma, const float* d_beta, const float* d_pos, int L,
                         float* d_y_f32, void* d_y_act, void* d_ypos_act, int64_t rows, int do_norm, void* stream);

/* =====================================================================================================================
 * Diffusion Policy / DP3 denoising path.
 * The arithmetic of this path lives in sources that are NOT part of the reference mount (un-vendored submodules
 * third_party/diffusion_policy @ 7e8fea7, third_party/3D-Diffusion-Policy @ 95ad86a, and diffusers == 0.11.1); the
 * entry points below replace what the reference calls at
 *   policy/diffusion_policy/RolloutDiffusionPolicy.py:37-43,74      DDPMScheduler + policy.predict_action
 *   policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:50-56,90 DDIMScheduler + DP3.predict_action
 * with the hyper-parameters of TrainDiffusionPolicy.py:114-138 / TrainDiffusionPolicy3d.py:185-218. */
typedef struct rmb_unet_engine rmb_unet_engine_t;

/* ConditionalUnet1D(input_dim, global_cond_dim, diffusion_step_embed_dim, down_dims, kernel_size, n_groups,
 * cond_predict_scale=True) as built from model_meta_info["policy"]["args"]. */
typedef struct rmb_unet_config {
  int32_t max_batch;           /* trajectories per call (workspace is sized for this) */
  int32_t horizon;             /* T (16): power of two, divisible by 2^(n_levels-1) */
  int32_t input_dim;           /* action_dim; kernel_size * input_dim <= 64 */
  int32_t global_cond_dim;     /* obs feature dim * n_obs_steps */
  int32_t dsed;                /* diffusion_step_embed_dim (128) */
  int32_t n_levels;            /* len(down_dims), 2..4 */
  int32_t down_dims[4];        /* multiples of 64 */
  int32_t kernel_size;         /* 5 (odd) */
  int32_t n_groups;            /* 8 (only value supported) */
  int32_t num_train_timesteps; /* 100: size of the per-timestep FiLM table */
  int32_t device;
} rmb_unet_config_t;

/* One scheduler step, made on the host with the scheduler's own fp32 arithmetic (diffusers 0.11.1):
 *   x0 = pred_sample ? out : (x - s_beta*out) / s_alpha;  if clip: x0 = clamp(x0, -1, 1)
 *   kind 0 (DDPMScheduler.step, fixed_small): x' = c_x0*x0 + c_xt*x (+ sigma*noise when sigma > 0)
 *   kind 1 (DDIMScheduler.step, eta 0):       x' = c_x0*x0 + c_xt*(derive_eps ? (x - s_alpha*x0)/s_beta : out) */
typedef struct rmb_sched_step {
  int32_t t;            /* timestep fed to the UNet (0 <= t < num_train_timesteps) */
  int32_t kind;         /* 0 DDPM, 1 DDIM */
  int32_t pred_sample;  /* prediction_type: 0 "epsilon", 1 "sample" */
  int32_t clip;         /* clip_sample */
  int32_t derive_eps;
  float s_beta;         /* sqrt(1 - alphas_cumprod[t]) */
  float s_alpha;        /* sqrt(alphas_cumprod[t]) */
  float c_x0;           /* DDPM pred_original_sample_coeff | DDIM sqrt(alpha_prod_t_prev) */
  float c_xt;           /* DDPM current_sample_coeff       | DDIM sqrt(1 - alpha_prod_t_prev) */
  float sigma;          /* DDPM sqrt(variance) (0 at t = 0) */
} rmb_sched_step_t;

/* One DDPMScheduler.step / DDIMScheduler.step over n float32 elements, stand-alone (the fused form runs inside
 * rmb_unet1d_denoise): d_out[i] = step(d_x[i], d_model_out[i], d_noise[i]); d_noise may be NULL when sigma == 0. */
int rmb_scheduler_step(const float* d_x, const float* d_model_out, const float* d_noise, const rmb_sched_step_t* step,
                       float* d_out, int64_t n, void* stream);

int rmb_unet_engine_create(const rmb_unet_config_t* cfg, rmb_unet_engine_t** out);
void rmb_unet_engine_destroy(rmb_unet_engine_t* e);
/* ConditionalUnet1D state-dict tensor (key relative to the module, e.g. "down_modules.0.0.blocks.0.block.0.weight"),
 * fp32 host memory. */
int rmb_unet_load_tensor(rmb_unet_engine_t* e, const char* name, const float* host_data, int64_t numel);
int rmb_unet_finalize(rmb_unet_engine_t* e);
/* The whole conditional_sample loop (DiffusionUnetHybridImagePolicy / DP3 .conditional_sample, obs_as_global_cond):
 * d_global_cond float32 [batch, global_cond_dim]; d_x_init float32 [batch, T, input_dim] (the torch.randn trajectory);
 * d_noise float32 [n_steps, batch, T, input_dim] (DDPM variance noise of loop iteration i; may be NULL when every
 * sigma is 0); steps: HOST array of n_steps entries; d_x_out float32 [batch, T, input_dim] (may alias d_x_init).
 * n_steps * launches_per_step + 3 kernels are enqueued on `stream`; no synchronisation. */
int rmb_unet1d_denoise(rmb_unet_engine_t* e, const float* d_global_cond, const float* d_x_init, const float* d_noise,
                       const rmb_sched_step_t* steps, int n_steps, float* d_x_out, int batch, void* stream);
/* One ConditionalUnet1D.forward(sample, timestep, global_cond): d_sample / d_out float32 [batch, T, input_dim]. */
int rmb_unet1d_forward(rmb_unet_engine_t* e, const float* d_sample, int timestep, const float* d_global_cond,
                       float* d_out, int batch, void* stream);
/* Kernel launches per UNet evaluation, and algorithmic FLOPs of one evaluation for `batch` trajectories. */
int rmb_unet_launches_per_step(const rmb_unet_engine_t* e);
double rmb_unet_step_flops(const rmb_unet_engine_t* e, int batch);
/* rmb_unet1d_forward with CUDA events around every launch (see rmb_act_forward_profile). */
int rmb_unet1d_forward_profile(rmb_unet_engine_t* e, const float* d_sample, int timestep, const float* d_global_cond,
                               float* d_out, int batch, void* stream, rmb_launch_prof_t* out, int cap);

/* ---- DP3Encoder (PointNetEncoderXYZ with LayerNorm + state MLP): obs -> per-obs-step feature */
typedef struct rmb_pointnet_engine rmb_pointnet_engine_t;
typedef struct rmb_pointnet_config {
  int32_t max_clouds;    /* batch * n_obs_steps */
  int32_t num_points;    /* 1024 */
  int32_t in_channels;   /* 3 (xyz) or 6 (use_pc_color) */
  int32_t out_channels;  /* encoder_output_dim (64), <= 64 */
  int32_t state_dim;     /* 0 = no state MLP */
  int32_t device;
} rmb_pointnet_config_t;
int rmb_pointnet_engine_create(const rmb_pointnet_config_t* cfg, rmb_pointnet_engine_t** out);
void rmb_pointnet_engine_destroy(rmb_pointnet_engine_t* e);
/* DP3Encoder state-dict tensor ("extractor.mlp.0.weight", "extractor.final_projection.0.weight", "state_mlp.0.weight", ...). */
int rmb_pointnet_load_tensor(rmb_pointnet_engine_t* e, const char* name, const float* host_data, int64_t numel);
int rmb_pointnet_finalize(rmb_pointnet_engine_t* e);
/* d_points float32 [n_clouds, num_points, point_stride] (first in_channels of each point are used);
 * d_state float32 [n_clouds, state_dim] or NULL; d_out float32 [n_clouds, out_channels (+ 64 with state)]. */
int rmb_pointnet_encode(rmb_pointnet_engine_t* e, const float* d_points, int point_stride, const float* d_state,
                        float* d_out, int n_clouds, void* stream);

/* ---- Diffusion Policy image observation encoder of one camera key (robomimic VisualCore as diffusion_policy builds it:
 * CropRandomizer (centre crop at eval) -> ResNet18Conv with BatchNorm replaced by GroupNorm(C/16, C) ->
 * SpatialSoftmax(num_kp) -> Flatten -> Linear(2*num_kp, feat)); source not in the reference mount (robomimic 0.2.0). */
typedef struct rmb_imgenc_engine rmb_imgenc_engine_t;
typedef struct rmb_imgenc_config {
  int32_t max_images;      /* batch * n_obs_steps */
  int32_t in_h, in_w;      /* image the rollout feeds: model_meta_info["data"]["image_size"] (240 x 320) */
  int32_t crop_h, crop_w;  /* crop_shape (216 x 288) */
  int32_t num_kp;          /* 32 */
  int32_t feat_dim;        /* 64 */
  int32_t device;
} rmb_imgenc_config_t;
int rmb_imgenc_engine_create(const rmb_imgenc_config_t* cfg, rmb_imgenc_engine_t** out);
void rmb_imgenc_engine_destroy(rmb_imgenc_engine_t* e);
/* keys relative to obs_encoder.obs_nets.<rgb key>: "backbone.nets.0.weight", "backbone.nets.4.0.bn1.weight",
 * "pool.nets.weight", "nets.3.weight", ... */
int rmb_imgenc_load_tensor(rmb_imgenc_engine_t* e, const char* name, const float* host_data, int64_t numel);
int rmb_imgenc_finalize(rmb_imgenc_engine_t* e);
/* d_images float32 [n_images, 3, in_h, in_w] (already scaled by the caller, RolloutDiffusionPolicy.py:117-123);
 * feature j of image i is written to d_out[i * ld_out + j], j < feat_dim. */
int rmb_imgenc_encode(rmb_imgenc_engine_t* e, const float* d_images, float* d_out, int64_t ld_out, int n_images,
                      void* stream);
/* debug tap: layer4 output after the last GroupNorm + ReLU as float32 [n_images, Hf, Wf, 512]; returns floats written */
int64_t rmb_imgenc_debug_read(rmb_imgenc_engine_t* e, float* d_out, int n_images, void* stream);
double rmb_imgenc_flops(const rmb_imgenc_engine_t* e, int n_images);

/* ---- DP3 observation pipeline on the device (the step before DP3.predict_action):
 * convert_depth_image_to_pointcloud (common/utils/VisionUtils.py:55-87) -> crop_pointcloud_bb + downsample_pointcloud_fps
 * (common/utils/Vision3dUtils.py:6-25) -> normalize_data + float32 cast (RolloutDiffusionPolicy3d.py:152-153), for
 * `batch` already-resized camera frames at once. fp64 arithmetic in the reference's op order (bit-exact). */
typedef struct rmb_pointcloud_params {
  int32_t H, W;              /* data["image_size"] after cv2.resize (RolloutDiffusionPolicy3d.py:134-137) */
  int32_t num_points;        /* data["num_points"] */
  double focal_scaling;      /* (1 / tan(deg2rad(fovy) / 2)) * H / 2, made on the host with numpy (VisionUtils.py:59) */
  double near_clip, far_clip;          /* 0, +inf */
  int32_t has_min_bound, has_max_bound;
  double min_bound[3], max_bound[3];
  int32_t norm_type;         /* -1 none, 0 gaussian: (x - p0)/p1 (mean, std), 1 limits: p1*(x - p0) + out_min (min, scale) */
  double norm_p0[6], norm_p1[6], norm_out_min;
} rmb_pointcloud_params_t;
int64_t rmb_pointcloud_scratch_bytes(int batch, int H, int W, int num_points);
/* d_depth float32 [batch, H, W]; d_rgb uint8 [batch, H, W, 3]; d_out float32 [batch, num_points, 6] (xyz, rgb/255);
 * d_count int32 [batch] (points that survived clip + crop) or NULL; d_indices int32 [batch, num_points] (FPS indices
 * into the cropped cloud, -1 padded) or NULL. */
int rmb_pointcloud_from_depth(const float* d_depth, const uint8_t* d_rgb, const rmb_pointcloud_params_t* p, void* d_scratch,
                              float* d_out, int32_t* d_count, int32_t* d_indices, int batch, void* stream);

/* rmb_pointcloud_from_depth without caller-provided scratch and WITHOUT ANY ALLOCATION (no cudaMalloc / cudaFree inside):
 * for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid it). This entry point
 * has no counterpart in the reference. The scratch is a file-scope
 * static __device__ array of RMB_POINTCLOUD_STATIC_SCRATCH_BYTES (32 MiB) that every GPU loading the library holds until
 * exit; batch * H * W * 56 + batch * num_points * 4 bytes must fit (e.g. 84 clouds of 84 x 84, one of 480 x 640), larger
 * requests return -1 with a message in rmb_last_error(). Same arguments, kernel and bit-exact results as
 * rmb_pointcloud_from_depth. All calls share the one array, which keeps the last call's contents (the kernel
 * re-initialises everything it reads, nothing needs clearing): calls that OVERLAP on the GPU (two streams, parallel
 * graph branches) race -- serialise them on one stream. */
#define RMB_POINTCLOUD_STATIC_SCRATCH_BYTES (32ll << 20)
int rmb_pointcloud_from_depth_static(const float* d_depth, const uint8_t* d_rgb, const rmb_pointcloud_params_t* p,
                                     float* d_out, int32_t* d_count, int32_t* d_indices, int batch, void* stream);

/* ---- cv2.resize(img, (W2, H2)) with INTER_LINEAR on the device (RolloutDiffusionPolicy.py:117,
 * RolloutDiffusionPolicy3d.py:134-137). uint8: OpenCV's fixed-point path, bit-exact for the reference's down-scaling;
 * out_mode 0 = uint8 HWC [n, H2, W2, 3]; out_mode 1 = float32 CHW [n, 3, H2, W2] = (v * (1/255)) * scale + offset
 * (scale 2, offset -1 is the Diffusion Policy input, RolloutDiffusionPolicy.py:119-123). float32: single-channel depth. */
int rmb_resize_u8_bilinear(const uint8_t* d_src, int n, int H, int W, int H2, int W2, int out_mode, float scale,
                           float offset, void* d_dst, void* stream);
int rmb_resize_f32_bilinear(const float* d_src, int n, int H, int W, int H2, int W2, float* d_dst, void* stream);

/* ---- Per-episode rollout state of the diffusion policies on the device (B episodes at once).
 * rmb_obs_ring_push: the n_obs_steps observation buffers of RolloutDiffusionPolicy.get_state / get_images
 * (policy/diffusion_policy/RolloutDiffusionPolicy.py:87-143) and RolloutDiffusionPolicy3d.get_state / get_pointcloud
 * (policy/diffusion_policy_3d/RolloutDiffusionPolicy3d.py:103-171). d_buf [batch, n_obs_steps, row_bytes],
 * d_new [batch, row_bytes], d_mode uint8 [batch]: 0 = untouched (episode does not call the policy at this step),
 * 1 = first call since reset (every slot <- new), 2 = later call (pop(0) + append). row_bytes % 4 == 0. */
int rmb_obs_ring_push(void* d_buf, const void* d_new, const uint8_t* d_mode, int batch, int n_obs_steps,
                      int64_t row_bytes, void* stream);
/* d_dst row i <- d_src row d_idx[i] (compacts the episodes whose action FIFO is empty into a dense sub-batch). */
int rmb_gather_rows(const void* d_src, const int32_t* d_idx, void* d_dst, int n_rows, int64_t row_bytes, void* stream);
/* The float64 action FIFO of infer_policy (RolloutDiffusionPolicy.py:66-85, RolloutDiffusionPolicy3d.py:83-101):
 * d_fifo float64 [batch, n_action_steps, Da], d_pos int32 [batch] read position (n_action_steps = empty).
 * d_sub_row int32 [batch] or NULL: >= 0 -> the episode called the policy and its FIFO is refilled from
 * d_act_new + d_sub_row[b] * act_row_stride (float32 [n_action_steps, Da], cast to float64) with position 0.
 * Then one pop(0) per episode + denormalize_data (norm parameters as rmb_denormalize) -> d_action float64 [batch, Da];
 * d_action_norm (or NULL) gets the popped, still normalised action. */
int rmb_action_fifo_step(double* d_fifo, int32_t* d_pos, const float* d_act_new, int64_t act_row_stride,
                         const int32_t* d_sub_row, int norm_type, const double* d_p0, const double* d_p1, double out_min,
                         double* d_action, double* d_action_norm, int batch, int n_action_steps, int Da, void* stream);

/* ---- Batched action consumer for arm bodies (the step right after infer_policy): ArmManager.set_command_data /
 * get_command_data / inverse_kinematics (robo_manip_baselines/common/body/ArmManager.py:88-254), one damped least-squares
 * IK step per end-effector command, fp64, one thread per episode. The kinematic model is a serial chain of revolute
 * joints (what pin.buildModelFromUrdf yields for the reference's arm URDFs; Pinocchio is restated, not linked). */
#define RMB_ARM_MAX_JOINTS 8
#define RMB_ARM_MAX_GRIPPER 4
typedef struct rmb_arm_model {
  int32_t n_joints;                                  /* revolute joints of the chain (<= RMB_ARM_MAX_JOINTS) */
  int32_t ik_joint_id;                               /* ArmConfig.ik_eef_joint_id: 1-based joint whose frame is the end effector */
  double placement_R[RMB_ARM_MAX_JOINTS][9];         /* joint placement in its parent joint's frame, rotation row-major */
  double placement_p[RMB_ARM_MAX_JOINTS][3];
  double axis[RMB_ARM_MAX_JOINTS][3];                /* rotation axis in the joint's own frame (unit) */
} rmb_arm_model_t;
typedef struct rmb_arm_layout {
  int32_t n_gripper;                                 /* gripper joints of this body (<= RMB_ARM_MAX_GRIPPER) */
  int32_t eef_idx;                                   /* ArmConfig.eef_idx */
  int32_t arm_idx[RMB_ARM_MAX_JOINTS];               /* ArmConfig.arm_joint_idxes: columns of a joint-position command */
  int32_t gripper_idx[RMB_ARM_MAX_GRIPPER];          /* ArmConfig.gripper_joint_idxes */
  int32_t gripper_idx_in_gripper[RMB_ARM_MAX_GRIPPER]; /* ArmConfig.gripper_joint_idxes_in_gripper_joint_pos */
  double init_arm[RMB_ARM_MAX_JOINTS];               /* ArmConfig.init_arm_joint_pos */
  double init_gripper[RMB_ARM_MAX_GRIPPER];
  double gripper_low[RMB_ARM_MAX_GRIPPER];           /* env.action_space.low / high at gripper_joint_idxes_for_limit */
  double gripper_high[RMB_ARM_MAX_GRIPPER];
} rmb_arm_layout_t;
/* key: 0 command_joint_pos, 1 command_joint_pos_rel, 2 command_gripper_joint_pos, 3 command_eef_pose (tx ty tz qw qx qy qz),
 * 4 command_eef_pose_rel (tx ty tz roll pitch yaw), 5 reset. d_command float64 [batch, >= columns the key reads] with row
 * stride cmd_stride (elements); d_mask uint8 [batch] or NULL (0 = leave the episode alone). State, all float64 on the
 * device: d_arm_q [batch, n_joints], d_gripper_q [batch, max(n_gripper, 1)], d_target_se3 [batch, 12] (R row-major, p). */
int rmb_arm_command(const rmb_arm_model_t* model, const rmb_arm_layout_t* layout, int key, const double* d_command,
                    int cmd_stride, int is_skip, const uint8_t* d_mask, double* d_arm_q, double* d_gripper_q,
                    double* d_target_se3, int batch, void* stream);
/* get_command_data: d_joint_pos [batch, joint_pos_stride] (arm and gripper joints scattered to their command columns) and /
 * or d_eef_pose [batch, 7] = pose of the target SE3; either may be NULL. */
int rmb_arm_get_command(const rmb_arm_model_t* model, const rmb_arm_layout_t* layout, const double* d_arm_q,
                        const double* d_gripper_q, const double* d_target_se3, double* d_joint_pos, int joint_pos_stride,
                        double* d_eef_pose, int batch, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* RMB_B200_H_ */
