/* medvae_b200.h — C ABI of libmedvae_b200.so (sm_100a CUDA kernels for the conditional-VAE
 * training step of parsakzr/medvae-disentangled-multimodal).
 *
 * The reference has no FFI / plugin interface for this path: its boundary is the Python class
 * surface of src.models / src.losses.  This header is therefore the NEW native
 * boundary that sits under that surface; each entry point cites the reference code whose leaf
 * arithmetic it replaces (paths relative to the reference root).  Conventions:
 *   - every function returns 0 on success, non-zero on failure; mvae_last_error() gives the
 *     thread-local message.  Nothing falls back to the CPU: without a CUDA device every compute
 *     entry point fails.
 *   - all pointers are DEVICE pointers owned by the caller (PyTorch allocates them); no
 *     ownership is transferred; work is enqueued on `stream` (a cudaStream_t) and not synced.
 *   - activations are NHWC (channels innermost).  dtype: MVAE_F32 (verification mode) or
 *     MVAE_BF16 (bf16 storage, fp32 accumulate).  Parameters and their gradients are fp32 in the
 *     reference's own layouts (OIHW conv weights) so state_dict stays compatible.
 */
#ifndef MEDVAE_B200_H
#define MEDVAE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MVAE_ABI_VERSION 5

enum { MVAE_F32 = 0, MVAE_BF16 = 1 };

/* conv geometry (encoder_decoder.py):
 *   MVAE_CONV_S1   stride 1, padding (R-1)/2            nn.Conv2d 3x3 p1 / 1x1   (:128-146, :76-81)
 *   MVAE_CONV_DOWN F.pad(0,1,0,1) + 3x3 stride 2 pad 0  Downsample.forward        (:184-188)
 *   MVAE_CONV_UP   nearest 2x + 3x3 stride 1 pad 1      Upsample.forward          (:205-209)
 */
enum { MVAE_CONV_S1 = 0, MVAE_CONV_DOWN = 1, MVAE_CONV_UP = 2 };

/* engine selection for the conv entry points */
enum { MVAE_ENGINE_AUTO = 0, MVAE_ENGINE_SIMT = 1, MVAE_ENGINE_UMMA = 2 };

typedef struct mvae_conv_desc {
    int32_t batch;     /* B                                                        */
    int32_t in_h;      /* input H (before the implicit pad / upsample)             */
    int32_t in_w;
    int32_t in_c;      /* Cin                                                      */
    int32_t out_c;     /* Cout                                                     */
    int32_t ksize;     /* 1 or 3 (R == S)                                          */
    int32_t mode;      /* MVAE_CONV_*                                              */
    int32_t dtype;     /* MVAE_F32 | MVAE_BF16 : activation + packed-weight dtype  */
    int32_t engine;    /* MVAE_ENGINE_*                                            */
} mvae_conv_desc;

const char* mvae_last_error(void);
int mvae_abi_version(void);
/* 1 if a CUDA device with compute capability 10.x is present */
int mvae_device_ok(void);
/* number of CUDA kernels this library has launched in this process (all threads) */
uint64_t mvae_launch_count(void);
/* tuning / bring-up knobs (key, value ints; every default is 0).  2: 1 = attention without tcgen05;  3: GroupNorm chunk
 * budget in MB;  5: 1 = no two-pixel-tile stages on 128-channel tiles;  6: 1 = one filter tap per wgrad CTA;
 * 7: 1 = 256 x 256 single-buffer tiles;  8: 1 = no CTA pairs;  9: GroupNorm-backward epilogue on 8 (1) / 16 (2) warps
 * instead of the per-layer choice;  10: 1 = never fuse the GroupNorm backward into dgrad, 2 / 3 = only behind a
 * reduction longer than 1152 / 2304. */
int mvae_set_debug(int key, int value);

/* output spatial size for a conv desc */
int mvae_conv2d_out_hw(const mvae_conv_desc* d, int32_t* out_h, int32_t* out_w);
/* number of packed weight slabs ([slab][N][K]): ksize^2, or 16 for MVAE_CONV_UP */
int mvae_conv2d_num_slabs(const mvae_conv_desc* d);
/* 1 if the tcgen05 path will be used for this desc (bf16, Cin%64==0, Cout%64==0) */
int mvae_conv2d_uses_umma(const mvae_conv_desc* d);

/* nn.Conv2d weights (OIHW fp32) -> packed operands.  w_fprop: [slab][Cout][Cin], w_dgrad:
 * [slab][Cin][Cout] (either may be NULL).  `out_c_total`/`out_c_offset` let several convs
 * (AttnBlock q,k,v — encoder_decoder.py:76-79) share one packed operand of out_c_total rows. */
int mvae_conv2d_pack_weights(const mvae_conv_desc* d, const float* w_oihw, void* w_fprop,
                             void* w_dgrad, int32_t out_c_total, int32_t out_c_offset,
                             void* stream);

/* The same for ALL convs of a model in one launch (once per optimizer step).  jobs: DEVICE array of n_jobs entries
 * ordered by tile_begin; job i owns tiles [tile_begin, tile_begin + ceil(out_c/32) * ceil(in_c/32)) of the
 * launch's total_tiles.  A job writes rows [out_c_offset, out_c_offset + out_c) x columns [0, in_c) of packed
 * operands whose dimensions are out_c_total x in_c_pitch (fused q,k,v; conv_in / conv_out zero-padded to one
 * tensor-core tile: the padding is left untouched) and copies the conv's bias (NULL: zeros) to
 * bias_dst[out_c_offset ...] when bias_dst != NULL. */
typedef struct mvae_pack_job {
    const float* w_oihw;      /* [out_c][in_c][ksize][ksize] fp32 */
    void* w_fprop;            /* [n_slabs][out_c_total][in_c_pitch] or NULL */
    void* w_dgrad;            /* [n_slabs][in_c_pitch][out_c_total] or NULL */
    const float* bias_src;    /* [out_c] or NULL */
    float* bias_dst;          /* [out_c_total] or NULL */
    int32_t out_c, in_c, ksize, mode, n_slabs, out_c_total, out_c_offset, in_c_pitch, tile_begin, reserved;
} mvae_pack_job;
int mvae_conv2d_pack_weights_batched(const mvae_pack_job* jobs_dev, int32_t n_jobs, int32_t total_tiles,
                                     int32_t dtype, void* stream);

/* y = conv(x) + bias (+ residual).  Replaces nn.Conv2d.forward at encoder_decoder.py:152,162,
 * 168 (ResnetBlock), :86-88,:105 (AttnBlock 1x1), :187, :208, :308,:327,:426,:450.
 * `residual` (same shape as y, may be NULL) fuses the `x + h` of :170 / :107. */
int mvae_conv2d_fprop(const mvae_conv_desc* d, const void* x, const void* w_fprop,
                      const float* bias, const void* residual, void* y, void* stream);

/* dx = conv_transpose(dy).  The backward of the above w.r.t. its input (autograd of F.conv2d,
 * F.pad and F.interpolate in the reference). */
int mvae_conv2d_dgrad(const mvae_conv_desc* d, const void* dy, const void* w_dgrad, void* dx,
                      void* stream);

/* dW (OIHW fp32 [out_c][in_c][k][k], overwritten) and dbias (fp32 [out_c], may be NULL) from the
 * layer input x and the output gradient dy.  workspace: mvae_conv2d_wgrad_workspace_bytes(). */
size_t mvae_conv2d_wgrad_workspace_bytes(const mvae_conv_desc* d);
int mvae_conv2d_wgrad(const mvae_conv_desc* d, const void* x, const void* dy, float* dw_oihw,
                      float* dbias, void* workspace, size_t workspace_bytes, void* stream);

/* same, accumulating: dw = beta*dw + dW, dbias = beta*dbias + db (beta = 1: add into a gradient
 * buffer that already holds earlier micro-batches; beta = 0: overwrite).  dw must be contiguous OIHW. */
int mvae_conv2d_wgrad_acc(const mvae_conv_desc* d, const void* x, const void* dy, float* dw_oihw,
                          float* dbias, void* workspace, size_t workspace_bytes, float beta, void* stream);

/* GroupNorm(min(32,C), eps) [+ x*sigmoid(x)]  — Normalize + nonlinearity,
 * encoder_decoder.py:13-15,28-33 as used at :150-151,:158-159,:85,:325-326,:446-447.
 * x,y: [B, HW, C].  stats (out): [B][G][2] fp32 (mean, rstd).  scratch: fp32 workspace of
 * mvae_groupnorm_scratch_floats(batch, hw, c, groups, backward = 0 | 1) elements.
 * All reductions run in a fixed order: results are bitwise reproducible run to run. */
int mvae_groupnorm_splits(int32_t batch, int32_t hw, int32_t c);
int64_t mvae_groupnorm_scratch_floats(int32_t batch, int32_t hw, int32_t c, int32_t groups,
                                      int32_t backward);
int mvae_groupnorm_silu_fwd(const void* x, const float* gamma, const float* beta, void* y,
                            float* stats, float* scratch, int32_t batch, int32_t hw, int32_t c,
                            int32_t groups, float eps, int32_t apply_silu, int32_t dtype,
                            void* stream);
/* dx = d(GN+SiLU)(dy) + dx_add (dx_add may be NULL: fuses the residual-branch gradient);
 * dgamma/dbeta (fp32 [C]) are overwritten. */
int mvae_groupnorm_silu_bwd(const void* dy, const void* x, const float* gamma, const float* beta,
                            const float* stats, const void* dx_add, void* dx, float* dgamma,
                            float* dbeta, float* scratch, int32_t batch, int32_t hw, int32_t c,
                            int32_t groups, int32_t apply_silu, int32_t dtype, void* stream);

/* ---- GroupNorm fused into the convs either side of it (bf16 tcgen05 path only) -----------------
 * In ResnetBlock / AttnBlock / the heads every Normalize (encoder_decoder.py:150,158,85,325,446)
 * reads a conv output and (in the backward) receives its gradient from a conv dgrad.  These entry
 * points move the two REDUCTION passes of GroupNorm into those convs' epilogues, leaving one
 * streaming pass per direction:
 *   mvae_conv2d_fprop_gn   = mvae_conv2d_fprop + gn_sums[n][g] = (sum y, sum y^2) over the group
 *                            (fp32 [B][groups][2], zeroed by the call; fp32 atomics)
 *   mvae_groupnorm_apply_fwd : y = [silu](GN(x)) from those sums; writes stats (mean, rstd)
 *   mvae_conv2d_dgrad_gn   = mvae_conv2d_dgrad whose output da = dL/d[silu](GN(gn_x)) is turned
 *                            into du = dL/d(xhat*gamma+beta) in the epilogue (stored instead of da),
 *                            + gn_s12[n][g] = (sum du*gamma, sum du*gamma*xhat)
 *   mvae_groupnorm_apply_bwd : dx = rstd*(du*gamma - s1/cnt - xhat*s2/cnt) + dx_add, and in the
 *                            same pass out3[0] = dgamma, out3[1] = dbeta, out3[2] = column sums of dx
 *                            (= bias gradient of the conv that produced x).  out3: fp32 [3][C].
 * mvae_conv2d_fuses_gn(d, groups, direction): 1 if the conv (direction 0: fprop, output channels
 * normalised; 1: dgrad, input channels) can carry the hook: tcgen05 path, channels/groups a power
 * of two >= 4.  mvae_groupnorm_fused_ok(c, groups): the one-pass kernels support this shape. */
int mvae_conv2d_fuses_gn(const mvae_conv_desc* d, int32_t groups, int32_t direction);
int mvae_conv2d_fprop_gn(const mvae_conv_desc* d, const void* x, const void* w_fprop, const float* bias,
                         const void* residual, void* y, float* gn_sums, int32_t groups, void* stream);
int mvae_conv2d_dgrad_gn(const mvae_conv_desc* d, const void* dy, const void* w_dgrad, void* du,
                         const void* gn_x, const float* gn_stats, const float* gamma, const float* beta,
                         int32_t groups, int32_t apply_silu, float* gn_s12, void* stream);
int mvae_groupnorm_fused_ok(int32_t c, int32_t groups);
int64_t mvae_groupnorm_apply_bwd_scratch_floats(int32_t batch, int32_t hw, int32_t c);
int mvae_groupnorm_apply_fwd(const void* x, const float* gamma, const float* beta, void* y,
                             const float* sums, float* stats, int32_t batch, int32_t hw, int32_t c,
                             int32_t groups, float eps, int32_t apply_silu, int32_t dtype, void* stream);
int mvae_groupnorm_apply_bwd(const void* du, const void* x, const float* gamma, const float* stats,
                             const float* s12, const void* dx_add, void* dx, float* out3, float* scratch,
                             int32_t batch, int32_t hw, int32_t c, int32_t groups, int32_t dtype,
                             void* stream);

/* AttnBlock core, encoder_decoder.py:91-103: per image, single head of width C.
 * qkv: [B, T, 3C] (q|k|v along channels), out: [B, T, C].
 * probs: [B, T, T] fp32 scratch kept for the backward. */
int mvae_attn_core_fwd(const void* qkv, void* out, float* probs, int32_t batch, int32_t tokens,
                       int32_t c, int32_t dtype, void* stream);
/* dqkv from dout; dscratch: [B, T, T] fp32 */
int mvae_attn_core_bwd(const void* qkv, const float* probs, const void* dout, void* dqkv,
                       float* dscratch, int32_t batch, int32_t tokens, int32_t c, int32_t dtype,
                       void* stream);

/* The same AttnBlock core as ONE persistent tcgen05 kernel per direction (bf16 only; replaces the q/k/v -> softmax ->
 * weighted-sum part of src/models/encoder_decoder.py:91-103): the scores of an image stay in TMEM / shared memory.
 * mvae_attn_fused_ok: 1 when (tokens, c, dtype) is served by it (bf16, c % 64 == 0, tokens <= 256), else use
 * mvae_attn_core_*.  lse2: [B, T] fp32, log2-domain log-sum-exp of the scaled scores per query, written by the forward
 * and read by the backward, which recomputes the probabilities instead of storing [B, T, T].
 * Backward: out = the forward's output, dscratch: B*T floats (row dots of dout and out). */
int mvae_attn_fused_ok(int32_t tokens, int32_t c, int32_t dtype);
int mvae_attn_fused_fwd(const void* qkv, void* out, float* lse2, int32_t batch, int32_t tokens, int32_t c,
                        int32_t dtype, void* stream);
int mvae_attn_fused_bwd(const void* qkv, const void* out, const float* lse2, const void* dout, void* dqkv,
                        float* dscratch, int32_t batch, int32_t tokens, int32_t c, int32_t dtype,
                        void* stream);
/* mvae_attn_fused_bwd without a caller-provided scratch buffer.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses
 * that forbid it).  The B*T row dots live in a file-scope static __device__ array of MVAE_ATTN_STATIC_MAX_ROWS floats
 * (4 MiB, held by every GPU that loads the library until exit); the call allocates nothing (no cudaMalloc / cudaFree, no
 * allocating library routine) and returns an error when batch * tokens exceeds the maximum.  Same results as
 * mvae_attn_fused_bwd.  All calls share the array (it keeps the last call's contents; nothing needs clearing: each call
 * overwrites the range it reads): calls that overlap on the GPU — different streams without an ordering between
 * them — race. */
#define MVAE_ATTN_STATIC_MAX_ROWS 1048576
int mvae_attn_fused_bwd_static(const void* qkv, const void* out, const float* lse2, const void* dout, void* dqkv,
                               int32_t batch, int32_t tokens, int32_t c, int32_t dtype, void* stream);

/* ConditionalVAE.create_condition_map + torch.cat, conditional_vae.py:107-127 (concat path),
 * and the NCHW->NHWC conversion of x.  cond: [B, cond_dim] fp32 (one-hot or dense);
 * w: [C*64, cond_dim], b: [C*64] (condition_proj.0).  out: NHWC [B,H,W,out_c] with channels
 * [0,C) = x, [C,2C) = bilinear(ReLU(Linear(cond))) and [2C,out_c) zero.  With w == NULL only
 * the layout conversion is done (BaseVAE / BetaVAE input). */
int mvae_input_pack_fwd(const float* x_nchw, const float* cond, const float* w, const float* b,
                        void* out, int32_t batch, int32_t c, int32_t h, int32_t w_, int32_t cond_dim,
                        int32_t out_c, int32_t dtype, void* stream);
/* gradient of the condition branch: d_in NHWC [B,H,W,out_c] (channels [C,2C) are read) ->
 * dw [C*64, cond_dim], db [C*64] (fp32, overwritten). */
int mvae_input_pack_bwd(const void* d_in, const float* cond, const float* w, const float* b,
                        float* dw, float* db, int32_t batch, int32_t c, int32_t h, int32_t w_,
                        int32_t cond_dim, int32_t out_c, int32_t dtype, void* stream);

/* layout helpers: NHWC (dtype) <-> NCHW fp32, optional channel window */
int mvae_nhwc_to_nchw_f32(const void* src, float* dst, int32_t batch, int32_t h, int32_t w,
                          int32_t c_src, int32_t c_off, int32_t c_dst, int32_t dtype, void* stream);
int mvae_nchw_f32_to_nhwc(const float* src, void* dst, int32_t batch, int32_t h, int32_t w,
                          int32_t c, int32_t dtype, void* stream);

/* torch.chunk + BaseVAE.reparameterize (base_vae.py:76,83-87) + the KL integrand of
 * VAELoss (vae_losses.py:55), one pass.  moments: NHWC [B,R,R,2*zc].  eps: NCHW fp32 or NULL
 * (then Philox4x32-10 keyed by (seed, offset, elem_base + element) generates it and, if eps_out !=
 * NULL, stores it; elem_base = this call's first sample in the GLOBAL batch x zc*r*r, so the noise of a
 * sample does not depend on how the batch is split over ranks and micro-batches — SURVEY 8e).  Outputs NCHW fp32 mean/logvar/z, optionally std = exp(0.5 logvar) (std_out) and
 * NHWC z (z_nhwc, dtype), both may be NULL.  kl_sum:
 * one double, ACCUMULATED: sum 0.5(e^lv + mu^2 - 1 - lv).  clamp > 0 applies
 * clamp(mu,±clamp), clamp(logvar,±clamp) first (disentangled_conditional_vae.py:409-412). */
int mvae_reparam_kl_fwd(const void* moments, const float* eps, float* eps_out, float* mean,
                        float* logvar, float* std_out, float* z_nchw, void* z_nhwc, double* kl_sum,
                        int32_t batch, int32_t zc, int32_t r, float clamp, uint64_t seed,
                        uint64_t offset, uint64_t elem_base, const double* step_state, int32_t dtype,
                        void* stream);
/* d_moments (NHWC dtype) from d_mean, d_logvar, d_z (NCHW fp32, each may be NULL) and the
 * saved mean/logvar/eps; clamp masks the gradient like torch.clamp. */
int mvae_reparam_kl_bwd(const float* d_mean, const float* d_logvar, const float* d_z,
                        const float* mean_raw_unused, const float* logvar, const float* eps,
                        const void* moments, void* d_moments, int32_t batch, int32_t zc, int32_t r,
                        float clamp, int32_t dtype, void* stream);

/* VAELoss reconstruction term (vae_losses.py:41-48): kind 0 mse, 1 l1, 2 bce-with-logits.
 * recon, x: same-shape fp32 arrays of n elements.  loss_sum: one double, ACCUMULATED (un-normalised
 * sum).  d_unit (may be NULL): d(mean loss)/d(recon) = integrand'/n, written in the same pass. */
int mvae_recon_loss_fwd_bwd(const float* recon, const float* x, double* loss_sum, float* d_unit,
                            int64_t n, int32_t kind, void* stream);
/* KL term alone (for posteriors that did not come from mvae_reparam_kl_fwd) and its gradient:
 * d_mean = coef*mu, d_logvar = coef*0.5(e^lv-1), coef read from device (*coef_dev) */
int mvae_kl_fwd(const float* mean, const float* logvar, double* kl_sum, int64_t n, void* stream);
int mvae_kl_bwd(const float* mean, const float* logvar, const float* coef_dev, float* d_mean,
                float* d_logvar, int64_t n, void* stream);
/* out[i] = in[i] * (*scale_dev) */
int mvae_scale_by_device_scalar(const float* in, const float* scale_dev, float* out, int64_t n,
                                void* stream);

/* fused multi-tensor AdamW/Adam + global-norm clip + non-finite guard
 * (lightning_module.py:390-477: optimizer.step, clip_gradients, NaN-grad zeroing).
 * Pointer tables live on the DEVICE: params/grads/exp_avg/exp_avg_sq[i] -> fp32 arrays of
 * sizes[i] elements.  Pass 1 (mvae_grad_sqnorm) accumulates sum(g^2) of finite entries into
 * *sqnorm_dev; pass 2 applies the step with grads scaled by min(1, max_norm/(norm+1e-6)). */
int mvae_grad_sqnorm(const float* const* grads, const int64_t* sizes, int32_t n_tensors,
                     double* sqnorm_dev, void* stream);
int mvae_adamw_clip_step(float* const* params, const float* const* grads, float* const* exp_avg,
                         float* const* exp_avg_sq, const int64_t* sizes, int32_t n_tensors,
                         const double* sqnorm_dev, float max_norm, float lr, float beta1,
                         float beta2, float eps, float weight_decay, int32_t decoupled,
                         int32_t step, void* stream);

/* the same two passes over a CHUNK TABLE (device arrays): chunk i covers elements
 * [chunk_offset[i], chunk_offset[i] + chunk_len[i]) of tensor chunk_tensor[i]; equal-sized chunks
 * balance the CTAs whatever the tensor sizes, and aligned chunks use 16-byte accesses.
 * Per-tensor semantics of the reference (src/lightning_module.py:468-477, torch.optim):
 *   - pass 1 accumulates tensor_sq[t] (sum of squares) and sets tensor_bad[t] when tensor t holds ANY NaN / Inf
 *     (both arrays have n_tensors entries and must be zero on entry);
 *   - pass 2 treats the WHOLE gradient of a bad tensor as zero (as `param.grad.zero_()` does), leaves a tensor
 *     with tensor_skip[t] != 0 untouched (a parameter whose .grad is None is skipped by torch.optim: no weight
 *     decay, no moment update; tensor_skip may be NULL), takes the global norm over the remaining tensors and
 *     writes its square to *sqnorm_out (may be NULL). */
int mvae_grad_sqnorm_chunked(const float* const* grads, const int32_t* chunk_tensor,
                             const int64_t* chunk_offset, const int32_t* chunk_len, int32_t n_chunks,
                             double* tensor_sq, int32_t* tensor_bad, void* stream);
int mvae_adamw_clip_step_chunked(float* const* params, const float* const* grads, float* const* exp_avg,
                                 float* const* exp_avg_sq, const int32_t* chunk_tensor,
                                 const int64_t* chunk_offset, const int32_t* chunk_len, int32_t n_chunks,
                                 const double* tensor_sq, const int32_t* tensor_bad,
                                 const int32_t* tensor_skip, int32_t n_tensors, double* sqnorm_out,
                                 float max_norm, float lr, float beta1, float beta2, float eps,
                                 float weight_decay, int32_t decoupled, int32_t step,
                                 const double* step_state, void* stream);

/* Device-resident per-step scalars, so that a CUDA-graph capture of the training step does not bake
 * them into its kernel arguments: step_state = double[3] {optimizer step t, learning rate, Philox
 * stream position}.  Entry points that take `step_state` (may be NULL = use the by-value arguments)
 * read t / lr from it (mvae_adamw_clip_step_chunked) or add its position to `offset`
 * (mvae_reparam_kl_fwd, mvae_dropout_apply).  mvae_step_state_advance: t += 1, position += rng_increment. */
int mvae_step_state_advance(double* step_state, double rng_increment, void* stream);

/* DisentangledConditionalVAE per-modality input projector
 * (disentangled_conditional_vae.py:124-171): out NHWC [B,H,W,out_c]; for modalities whose
 * channel count is 1, out[:, :3] = w_m * x[:,0] + b_m (Conv2d(1,3,1)), else x passes through.
 * x: NCHW fp32 with xc channels; idx: int64 [B]; proj_w: [n_mod][3], proj_b: [n_mod][3];
 * is_gray: int32 [n_mod]. */
int mvae_dis_input_fwd(const float* x_nchw, const int64_t* idx, const float* proj_w,
                       const float* proj_b, const int32_t* is_gray, void* out, int32_t batch,
                       int32_t xc, int32_t h, int32_t w, int32_t n_mod, int32_t out_c,
                       int32_t dtype, void* stream);
int mvae_dis_input_bwd(const void* d_in, const float* x_nchw, const int64_t* idx,
                       const int32_t* is_gray, float* d_proj_w, float* d_proj_b, int32_t batch,
                       int32_t xc, int32_t h, int32_t w, int32_t n_mod, int32_t out_c,
                       int32_t dtype, void* stream);

/* DisentangledConditionalVAE.decode per-modality heads (disentangled_conditional_vae.py:241-303):
 * modality_decoders[m] = Conv2d(3,3,3,1,1) -> ReLU -> Conv2d(3,3,3,1,1), then for 1-channel modalities
 * modality_output_projectors[m] = Conv2d(3,1,1), zero-padded to out_c channels (1 if the whole batch
 * is gray, else 3).  r_nhwc: decoder output [B,H,W,r_channels] (dtype; first 3 channels used).
 * Stacked parameters: w1,w2 [n_mod][3][3][3][3]; b1,b2 [n_mod][3]; proj_w [n_mod][3]; proj_b [n_mod].
 * hid, y: fp32 [B,H,W,3] kept for the backward.  out: NCHW fp32 [B,out_c,H,W]. */
int mvae_dis_head_fwd(const void* r_nhwc, int32_t r_channels, const int64_t* idx, const float* w1,
                      const float* b1, const float* w2, const float* b2, const float* proj_w,
                      const float* proj_b, const int32_t* is_gray, float* hid, float* y, float* out_nchw,
                      int32_t batch, int32_t h, int32_t w, int32_t n_mod, int32_t out_c, int32_t dtype,
                      void* stream);
/* grads: fp32 [n_mod][172] = (dw1 81, db1 3, dw2 81, db2 3, dproj_w 3, dproj_b 1), overwritten.
 * d_r_nhwc: channels 0..2 are written (caller zero-fills the tile padding).  scratch: 6*B*H*W floats. */
int mvae_dis_head_bwd(const float* d_out_nchw, const void* r_nhwc, int32_t r_channels, const float* hid,
                      const float* y, const int64_t* idx, const float* w1, const float* w2,
                      const float* proj_w, const int32_t* is_gray, void* d_r_nhwc, float* grads,
                      float* scratch, int32_t batch, int32_t h, int32_t w, int32_t n_mod, int32_t out_c,
                      int32_t dtype, void* stream);

/* modality_separation_loss + contrastive_loss (:305-386) on z_flat fp32 [B][flat_dim] (the NCHW latent
 * flattened as partition_latent does, :195-206): z_modality = z_flat[:, shared : shared+modality_dim].
 * aux: fp32 workspace of mvae_dis_aux_floats() elements; afterwards aux[0] = separation loss,
 * aux[1] = contrastive loss (the rest is kept for the backward).  Backward: g_*_dev are device
 * scalars (NULL = 0); dz_flat columns [shared, shared+modality_dim) are written. */
int64_t mvae_dis_aux_floats(int32_t batch, int32_t modality_dim, int32_t n_mod);
int mvae_dis_aux_loss_fwd(const float* z_flat, const int64_t* idx, float* aux, int32_t batch,
                          int64_t flat_dim, int32_t shared_dim, int32_t modality_dim, int32_t n_mod,
                          float temperature, void* stream);
int mvae_dis_aux_loss_bwd(const float* z_flat, const int64_t* idx, const float* aux, const float* g_sep_dev,
                          const float* g_con_dev, float* dz_flat, int32_t batch, int64_t flat_dim,
                          int32_t shared_dim, int32_t modality_dim, int32_t n_mod, float temperature,
                          void* stream);
/* DisentangledVAELoss total with its NaN/Inf guards (:527-565) from four device scalars: out5 =
 * (total, recon, kl, separation, contrastive) after the guards; mask4[i] = 1 where gradient flows. */
int mvae_dis_loss_total(const float* recon, const float* kl, const float* sep, const float* con,
                        float w_recon, float w_kl, float w_sep, float w_con, float* out5, float* mask4,
                        void* stream);

/* ---- callers either side of the training step ------------------------------------ */

/* LinearAttention / LinAttnBlock core (src/models/encoder_decoder.py:36-65; heads=1, dim_head=C) on the
 * fused NHWC qkv tensor [B][T][3C] (to_qkv output): ks = softmax over tokens of k, ctx = ks^T v,
 * out = q ctx.  ks: fp32 [B][T][C], ctx: fp32 [B][C][C], both kept for the backward.
 * Backward scratch: dctx fp32 [B][C][C], dks fp32 [B][T][C]. */
int mvae_linattn_core_fwd(const void* qkv, void* out, float* ks, float* ctx, int32_t batch, int32_t tokens,
                          int32_t c, int32_t dtype, void* stream);
int mvae_linattn_core_bwd(const void* qkv, const float* ks, const float* ctx, const void* dout, void* dqkv,
                          float* dctx, float* dks, int32_t batch, int32_t tokens, int32_t c, int32_t dtype,
                          void* stream);

/* nn.Dropout inside ResnetBlock (src/models/encoder_decoder.py:164): y = x * keep / (1-p) over n elements,
 * keep from Philox4x32-10 (seed, offset, elem_base + element) or, if mask_in != NULL, from that byte mask
 * (verification hook).  elem_base = index of the first element in the GLOBAL batch (first sample x elements per
 * sample): a sample's mask then does not depend on the rank count or the micro-batch split.  mask_out (may be
 * NULL) receives the mask used.  The backward is the same call on the incoming gradient with the same
 * (seed, offset, elem_base) or mask. */
int mvae_dropout_apply(const void* x, void* y, int64_t n, float p, uint64_t seed, uint64_t offset,
                       uint64_t elem_base, const double* step_state, const uint8_t* mask_in,
                       uint8_t* mask_out, int32_t dtype, void* stream);

/* FiLMLayer (src/models/conditional_vae.py:191-203): out = features * (Ws c + bs) + (Wt c + bt).
 * features/out: NCHW fp32 [B][C][HW]; cond: [B][K]; Ws, Wt: [C][K]; st: fp32 [2][B][C] receives the
 * per-sample scale and shift (kept for the backward).  Backward writes d_features, the four parameter
 * gradients and (if d_cond != NULL) the condition gradient; dst: fp32 scratch [2][B][C]. */
int mvae_film_fwd(const float* features, const float* cond, const float* w_scale, const float* b_scale,
                  const float* w_shift, const float* b_shift, float* out, float* st, int32_t batch,
                  int32_t c, int32_t hw, int32_t k, void* stream);
int mvae_film_bwd(const float* d_out, const float* features, const float* cond, const float* w_scale,
                  const float* w_shift, const float* st, float* d_features, float* dst, float* d_w_scale,
                  float* d_b_scale, float* d_w_shift, float* d_b_shift, float* d_cond, int32_t batch,
                  int32_t c, int32_t hw, int32_t k, void* stream);

/* Batch feed (src/data/medmnist_data.py:186-251 __getitem__, :372-380 Normalize(0.5, 0.5), :16-72
 * mixed_modality_collate_fn) over a staged uint8 batch src [B][H][W][src_stride_c] (1 or 3 wide):
 * per sample, src_c[b] in {1,3} valid source channels and tgt_c[b] in {1,3} the modality's channel count;
 * u8/255, RGB->gray 0.299/0.587/0.114 or gray->RGB repeat, (v-0.5)/0.5, zero-pad to out_c channels.
 * x: NCHW fp32 [B][out_c][H][W]; onehot (may be NULL): fp32 [B][n_mod] from idx.  Bit-identical to the
 * reference's fp32 host arithmetic. */
int mvae_batch_feed_u8(const uint8_t* src, const int32_t* src_c, const int32_t* tgt_c, const int64_t* idx,
                       float* x, float* onehot, int32_t batch, int32_t h, int32_t w, int32_t src_stride_c,
                       int32_t out_c, int32_t n_mod, void* stream);

/* Validation metrics (src/utils/metrics.py:14-102; torchmetrics PSNR / SSIM with data_range as given,
 * gaussian 11x11 sigma 1.5, k1 0.01, k2 0.03).  All results stay on the device.
 * recon: original/recon NCHW fp32 with planes = B*C images of H x W; out4 = (mse, mae, psnr, ssim);
 *        ws: mvae_recon_metrics_ws_doubles() doubles.
 * kl:    mean/logvar fp32 [B][C][S] (S = product of trailing dims; a "sample" of kl_mean / kl_std is one
 *        (b, s) pair, as `sum(dim=1)` makes it); out4 = (kl_total, kl_mean, kl_std, kl_per_dim_mean).
 * latent: z fp32 [B][D]; out3 = (latent_mean_avg, latent_std_avg, latent_sparsity(|z| < thr)).
 * kl / latent ws: mvae_stat_metrics_ws_doubles() doubles. */
int64_t mvae_recon_metrics_ws_doubles(int32_t planes, int32_t h, int32_t w);
int64_t mvae_stat_metrics_ws_doubles(void);
int mvae_recon_metrics(const float* original, const float* recon, int32_t planes, int32_t h, int32_t w,
                       float data_range, double* ws, float* out4, void* stream);
int mvae_kl_metrics(const float* mean, const float* logvar, int32_t batch, int32_t c, int32_t s, double* ws,
                    float* out4, void* stream);
int mvae_latent_metrics(const float* z, int32_t batch, int64_t d, float thr, double* ws, float* out3,
                        void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MEDVAE_B200_H */
