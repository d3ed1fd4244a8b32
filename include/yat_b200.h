/*
 * yat_b200 — C-ABI of the B200-native diffusion-training hot path.
 *
 * The reference (frutiemax92/YAT) is 100 % Python; its hot path is reached through a Python
 * plugin surface, not an FFI.  This header is the boundary a binding for that surface sits
 * on: plain pointers and sizes, no torch types.  Every entry point
 *   - takes device pointers owned by the caller (inputs are never written, outputs are
 *     caller-allocated), an explicit CUDA stream (a cudaStream_t passed as void*),
 *   - never allocates, never synchronises, and is CUDA-graph capturable,
 *   - returns 0 on success or a negative yat_status; yat_last_error() gives the text.
 * There is no CPU fallback behind any of them.
 *
 * Reference interfaces replaced (paths relative to the upstream tree):
 *   yat_noise_flowmatch   train_sana.py:183-207,217  train_sd35.py:180-185,192
 *   yat_noise_ddpm        train_pixart_sigma.py:171-176,183
 *   yat_loss_fused        train_sana.py:182,218  train_sd35.py:179,193
 *                         train_pixart_sigma.py:170,182-184  README.md:81-82 (preservation)
 *   yat_repa_cos          common/repa.py:137-154 (re-specified)
 *   yat_linattn_fwd/bwd   diffusers SanaLinearAttnProcessor2_0, reached from
 *                         utils/patch_sana_attention_layers.py:94
 *   yat_ln_modulate_*     utils/patch_sana_attention_layers.py:84-90,104-105 (+ norm_out)
 *   yat_gate_residual_*   utils/patch_sana_attention_layers.py:95,113
 *   yat_glumb_mid_*       GLUMBConv middle, utils/patch_sana_attention_layers.py:106-111
 *   yat_pad_embeddings    train_sana.py:167-179  train_pixart_sigma.py:158-169
 *   yat_grad_clip_coef / yat_adamw_step   common/trainer.py:245-248,347-348 (clip_grad_norm_ + AdamW.step)
 *   yat_bucket_assign     common/trainer.py:109-120  common/bucket_sampler.py:314-322
 */
#ifndef YAT_B200_H_
#define YAT_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  YAT_OK = 0,
  YAT_ERR_BAD_ARG = -1,      /* null pointer, non-positive size, misaligned pointer */
  YAT_ERR_UNSUPPORTED = -2,  /* shape / dtype outside what the kernels implement     */
  YAT_ERR_CUDA = -3          /* a CUDA call failed; see yat_last_error()             */
} yat_status;

/* element types of latent-shaped tensors */
#define YAT_DTYPE_BF16 0
#define YAT_DTYPE_F32 1

/* loss target modes */
#define YAT_TARGET_VELOCITY 0 /* target = noise - x0  (flow matching: SANA, SD3.5) */
#define YAT_TARGET_EPS 1      /* target = noise       (DDPM eps-prediction: PixArt)  */

/* loss output semantics */
#define YAT_LOSS_F32 0  /* MSELoss(pred.float(), target.float())      train_sana.py:218 */
#define YAT_LOSS_LOWP 1 /* MSELoss(pred.to(bf16), target) in bf16      train_sd35.py:193 */

int yat_version(void);
const char* yat_last_error(void);

/* ---- (1) fused flow-matching noising ------------------------------------------------
 * Per sample b: u -> idx=(long)(u*T) (clamped to T-1) -> t=timesteps_tab[idx] ->
 * sigma=cast(sigmas_tab[idx]); x_t=(1-sigma)*x0+sigma*eps with the reference's per-op
 * rounding in the latent dtype; target=eps-x0.
 *   x0            [B, chw] latents (dtype), read-only
 *   noise_in      [B, chw] or NULL.  NULL => eps generated in-kernel (Philox4x32-10 keyed
 *                 by seed, counter (quad, sample, offset); see csrc/philox.cuh)
 *   u_in          [B] fp32 in (0,1] or NULL.  NULL => u=sigmoid(N(0,1)) from the same stream
 *   sigmas_tab, timesteps_tab   fp32 [T] device tables (built once by the host side)
 *   x_t           [B, chw] out
 *   target        [B, chw] out or NULL
 *   noise_out     [B, chw] out or NULL (generated eps, if the caller wants it materialised)
 *   idx_out int64 [B], t_out fp32 [B], sigma_out fp32 [B] (value after the dtype cast)
 */
int yat_noise_flowmatch(const void* x0, const void* noise_in, const float* u_in,
                        const float* sigmas_tab, const float* timesteps_tab, int T,
                        uint64_t seed, uint64_t offset, int B, int64_t chw, int dtype,
                        void* x_t, void* target, void* noise_out,
                        int64_t* idx_out, float* t_out, float* sigma_out, void* stream);

/* ---- (1b) DDPM add_noise ------------------------------------------------------------
 * t = T-1-idx (int64), x_t = sqrt(acp[t])*x0 + sqrt(1-acp[t])*eps with acp cast to the
 * latent dtype first and per-op rounding (diffusers DDPMScheduler.add_noise).
 */
int yat_noise_ddpm(const void* x0, const void* noise_in, const float* u_in,
                   const float* alphas_cumprod_tab, int T,
                   uint64_t seed, uint64_t offset, int B, int64_t chw, int dtype,
                   void* x_t, void* noise_out, int64_t* idx_out, int64_t* t_out, void* stream);

/* ---- (2) fused denoising loss, forward + backward in one pass ----------------------------
 *   pred        [B, pred_channels, hw]  model output (bf16); only channels [0, C) are used
 *               (PixArt learned-sigma head: pred_channels = 2C), dpred of the rest is 0
 *   x0          [B, C, hw] latents (bf16); may be NULL when target_mode == EPS
 *   noise       [B, C, hw] eps (bf16) or NULL => regenerated from (seed, offset)
 *   pred_frozen [B, pred_channels, hw] or NULL (use_preservation reconstruction term)
 *   weights     fp32 [B] or NULL (=> all ones, the reference's plain mean)
 *   rho         preservation_ratio;  gscale  upstream gradient of the scalar loss
 *               (accelerate's 1/gradient_accumulation_steps)
 *   loss_out    fp32 [3]: total, noise term, preservation term
 *   dpred       [B, pred_channels, hw] bf16 out or NULL (forward only)
 *   workspace   device scratch of yat_loss_workspace_bytes(), zero-initialised once by the
 *               caller; the kernel leaves it zeroed for the next call
 */
int64_t yat_loss_workspace_bytes(void);
int yat_loss_fused(const void* pred, const void* x0, const void* noise, const void* pred_frozen,
                   const float* weights, float rho, float gscale, int target_mode, int loss_mode,
                   uint64_t seed, uint64_t offset,
                   int B, int C, int pred_channels, int64_t hw,
                   float* loss_out, void* dpred, void* workspace, void* stream);

/* ---- (2b) REPA cosine alignment, forward + backward ----------------------------------
 *   z   [B, gh*gw, E] bf16 projected student tokens;  y [B, th*tw, E] teacher (y_dtype)
 *   loss_out fp32 [1] = -(1/(B*th*tw)) sum cos(pool(z), y);  dz [B, gh*gw, E] bf16 or NULL,
 *   already multiplied by gscale.  gbuf: fp32 scratch [B, th*tw, E] (pooled-gradient staging).
 */
int64_t yat_repa_workspace_bytes(int B, int th, int tw, int E);
int yat_repa_cos(const void* z, const void* y, int y_dtype, int B, int gh, int gw, int th, int tw,
                 int E, float gscale, float* loss_out, void* dz, void* workspace, void* stream);

/* ---- (3) SANA ReLU linear attention ---------------------------------------------------
 *   q,k,v,o,(do,dq,dk,dv)  [B, N, H*D] bf16 row-major (outputs of to_q/to_k/to_v), D == 32
 *   O_n = relu(Q_n)(relu(K)^T V) / (relu(Q_n) . sum_m relu(K_m) + 1e-15)
 *   tcgen05/TMEM kernels (csrc/linattn_sm100.cu); bwd recomputes K^T V, no saved state.
 */
int yat_linattn_fwd(const void* q, const void* k, const void* v, int B, int N, int H, int D,
                    void* o, void* stream);
int yat_linattn_bwd(const void* q, const void* k, const void* v, const void* dout,
                    int B, int N, int H, int D, void* dq, void* dk, void* dv, void* stream);
/* Same kernels on column slices of wider row-major buffers: q, k, v are [B, N, H*D] views with a row pitch of `ld`
 * elements (the three thirds of ONE fused q|k|v projection output [B, N, 3*H*D]), dq, dk, dv likewise with `ld_grad`
 * (so the backward writes straight into the gradient of that fused projection); o and dout stay dense. */
int yat_linattn_fwd_ld(const void* q, const void* k, const void* v, int64_t ld, int B, int N, int H, int D,
                       void* o, void* stream);
int yat_linattn_bwd_ld(const void* q, const void* k, const void* v, int64_t ld, const void* dout, int B, int N, int H,
                       int D, void* dq, void* dk, void* dv, int64_t ld_grad, void* stream);
/* CUDA-core fp32 restatement of the same math; exists so the GPU tests can cross-check the
 * tensor-core kernels at full size, where the CPU oracle is too slow.  Not a product path. */
int yat_linattn_fwd_simt(const void* q, const void* k, const void* v, int B, int N, int H, int D,
                         void* o, void* stream);
int yat_linattn_bwd_simt(const void* q, const void* k, const void* v, const void* dout,
                         int B, int N, int H, int D, void* dq, void* dk, void* dv, void* stream);

/* ---- (next, N1) fused neighbours of the attention / feed-forward sub-layers -----------------
 * SANA block body, utils/patch_sana_attention_layers.py:84-95,104-113 (adaLN-single modulate, gate,
 * residual) and SanaModulatedNorm (norm_out).  x,y,res,out,dx,dy,dout: [B, N, C] bf16 token-major;
 * scale/shift/gate: per-sample rows, element (b, c) at ptr[b * mod_batch_stride + c] (bf16; they are
 * chunks of one [B, 6, C] tensor); C % 8 == 0, C <= 4096.
 *   ln_modulate     y   = LayerNorm(x; eps, no affine) * (1 + scale_b) + shift_b
 *                   fwd optionally saves mean / rstd fp32 [B*N] for the backward
 *                   bwd: dx, dscale [B, C] bf16 = sum_n dy * xhat, dshift [B, C] bf16 = sum_n dy
 *   gate_residual   out = res + gate_b * y
 *                   bwd: dy = gate_b * dout, dgate [B, C] bf16 = sum_n dout * y  (dres == dout)
 * workspace: yat_block_workspace_bytes(B, N, C) bytes of device scratch (split partial sums; the
 * reduction order is fixed, results are run-to-run identical).
 */
int64_t yat_block_workspace_bytes(int B, int N, int C);
int yat_ln_modulate_fwd(const void* x, const void* scale, const void* shift, int64_t mod_batch_stride,
                        int B, int N, int C, float eps, void* y, float* mean, float* rstd, void* stream);
int yat_ln_modulate_bwd(const void* x, const void* dy, const void* scale, int64_t mod_batch_stride,
                        const float* mean, const float* rstd, int B, int N, int C,
                        void* dx, void* dscale, void* dshift, void* workspace, void* stream);
/* the same two ops with the neighbouring residual arithmetic folded in:
 *   yat_add_ln_modulate_fwd  xsum = bf16(x + x2) (the attn2 residual add, :103), y = ln_modulate(xsum): one pass
 *   yat_ln_modulate_bwd_res  dx = dres + d ln_modulate / dx: the gradient that reaches the same tensor through the
 *                            residual branch is added in the same pass (dres may be NULL)                        */
int yat_add_ln_modulate_fwd(const void* x, const void* x2, const void* scale, const void* shift,
                            int64_t mod_batch_stride, int B, int N, int C, float eps, void* xsum, void* y,
                            float* mean, float* rstd, void* stream);
int yat_ln_modulate_bwd_res(const void* x, const void* dy, const void* dres, const void* scale,
                            int64_t mod_batch_stride, const float* mean, const float* rstd, int B, int N, int C,
                            void* dx, void* dscale, void* dshift, void* workspace, void* stream);
int yat_gate_residual_fwd(const void* res, const void* y, const void* gate, int64_t mod_batch_stride,
                          int B, int N, int C, void* out, void* stream);
int yat_gate_residual_bwd(const void* dout, const void* y, const void* gate, int64_t mod_batch_stride,
                          int B, int N, int C, void* dy, void* dgate, void* workspace, void* stream);

/* ---- (next, N1) bias gradient of a Linear layer: out[c] = sum_r a[r][c] ----------------------
 * torch autograd's `grad_output.sum(0)` of every biased projection of the block (attention.to_out / attn2.to_q,k,v,
 * diffusers Attention as used by utils/patch_sana_attention_layers.py:84-115).  a [rows, C] bf16 row-major,
 * out [C] bf16, fp32 accumulation in a fixed order; workspace = yat_colsum_workspace_bytes(rows, C) bytes.
 */
int64_t yat_colsum_workspace_bytes(int64_t rows, int C);
int yat_colsum_bf16(const void* a, int64_t rows, int C, void* out, void* workspace, void* stream);

/* yat_colsum_bf16_static: the same two kernels with the split partials in a file-scope device array inside the library
 * instead of a caller workspace, for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  NO COUNTERPART IN THE REFERENCE.
 * It allocates nothing (no cudaMalloc / cudaFree / allocating library call); the array is yat_colsum_static_max_bytes()
 * = 2 MiB per GPU that loads the library, held until exit.  Shapes whose partials do not fit (splits * C * 4 B, the
 * figure yat_colsum_workspace_bytes reports) are rejected with YAT_ERR_UNSUPPORTED and a message naming both sizes;
 * every SANA / PixArt / SD3.5 width of this build fits (C <= 11200 at any row count).  All calls share the array:
 * nothing has to start at zero (each call overwrites what it reads), but two calls that OVERLAP on the GPU (different
 * streams, or graphs replayed concurrently) race - serialise them.  The address is resolved once per device on the
 * first call with that device current (cudaGetSymbolAddress): make that first call outside stream capture.
 * Result: bit-identical to yat_colsum_bf16.
 */
int64_t yat_colsum_static_max_bytes(void);
int yat_colsum_bf16_static(const void* a, int64_t rows, int C, void* out, void* stream);

/* ---- (next, N1) fused middle of GLUMBConv ---------------------------------------------------
 * utils/patch_sana_attention_layers.py:106-111 -> GLUMBConv: SiLU, depthwise 3x3 (+bias, zero pad 1),
 * chunk, x * SiLU(gate).  Token-major (NHWC):
 *   u [B, H, W, 2*Hd] bf16 (conv_inverted output)   w_dw [2*Hd, 9] bf16   b_dw [2*Hd] bf16
 *   mid [B, H, W, Hd] bf16 (conv_point input);  Hd % 8 == 0 (any H, W: the image is cut into 4-row strips)
 * fwd optionally writes c [B, H, W, 2*Hd] bf16 = the depthwise-conv output (what autograd would keep);
 * bwd: (u, c, dmid) -> du [B,H,W,2*Hd], dw [2*Hd, 9] bf16, db [2*Hd] bf16, and (du_colsum, may be NULL)
 * sum over tokens of du [2*Hd] bf16 = the bias gradient of conv_inverted, formed from the fp32 values in the same pass;
 * workspace = yat_glumb_workspace_bytes(B, H, Hd) bytes (fixed-order partial sums).
 */
int64_t yat_glumb_workspace_bytes(int B, int H, int Hd);
int yat_glumb_mid_fwd(const void* u, const void* w_dw, const void* b_dw, int B, int H, int W, int Hd,
                      void* mid, void* c_out, void* stream);
int yat_glumb_mid_bwd(const void* u, const void* c, const void* w_dw, const void* dmid,
                      int B, int H, int W, int Hd, void* du, void* dw, void* db, void* du_colsum,
                      void* workspace, void* stream);

/* ---- (next) ragged text-embedding pad + mask -------------------------------------------
 *   packed [sum L_i, E] bf16, offsets int32 [B+1] -> padded [B, Lmax, E] bf16 (zero fill),
 *   mask int64 [B, Lmax] (1 for real tokens).  train_sana.py:167-179.
 */
int yat_pad_embeddings(const void* packed, const int32_t* offsets, int B, int Lmax, int E,
                       void* padded, int64_t* mask, void* stream);

/* ---- head padding for the cross-attention SDPA --------------------------------------------------
 * The -10000 key-padding bias of the text cross-attention (utils/patched_sana_transformer.py:275-277 ->
 * AttnProcessor2_0) is folded into one extra head column so the library SDPA runs without a mask tensor:
 *   head_pad    x [rows, H*d] bf16 -> out [rows, H, da]: out[..,:d] = x, out[..,d] = tail_row[row] (if given) else
 *               tail_const, rest 0;    head_unpad  xa [B, T, H, >=d] with element strides -> out [B*T, H*d].
 * d, da and the strides are multiples of 8 elements.  Each is the other's backward.
 */
int yat_head_pad(const void* x, const void* tail_row, float tail_const, int64_t rows, int H, int d, int da,
                 void* out, void* stream);
int yat_head_unpad(const void* xa, int64_t stride_b, int64_t stride_t, int64_t stride_h, int B, int T, int H, int d,
                   void* out, void* stream);

/* ---- optimiser step: global-norm clip + AdamW (common/trainer.py:347-348) ---------------------
 * The parameters of one group are described by device tables built once by the caller:
 *   tensor_tab [T][4] int64 = {param ptr, exp_avg ptr, exp_avg_sq ptr, numel}
 *   grad_ptrs  [T]    int64 = gradient pointer of each tensor (refreshed every step)
 *   chunks     [n_chunks][2] int32 = {tensor index, chunk index}; chunk c of a tensor covers elements
 *              [c * yat_optim_chunk_elems(), ...) of it
 * yat_grad_clip_coef: norm_coef_out[0] = ||g||_2 over all tensors (fp32 accumulation, fixed order),
 *   norm_coef_out[1] = min(1, max_norm / (norm + 1e-6))  (torch.nn.utils.clip_grad_norm_).
 *   workspace = yat_optim_workspace_bytes(n_chunks) bytes, zero before the first call.
 * yat_adamw_step: torch.optim.AdamW arithmetic (decoupled weight decay, bias correction for `step`
 *   >= 1) in fp32 on bf16 / fp32 storage; the gradient is first scaled by *clip_coef (device scalar,
 *   may be NULL = 1) and rounded to its dtype, as clip_grad_norm_ would have left it.  p, m, v are
 *   updated in place; gradients are not written.
 */
int yat_optim_chunk_elems(void);
int64_t yat_optim_workspace_bytes(int n_chunks);
int yat_grad_clip_coef(const int64_t* tensor_tab, const int64_t* grad_ptrs, const int32_t* chunks, int n_chunks,
                       int dtype, float max_norm, float* norm_coef_out, void* workspace, void* stream);
int yat_adamw_step(const int64_t* tensor_tab, const int64_t* grad_ptrs, const int32_t* chunks, int n_chunks,
                   int dtype, const float* clip_coef, double lr, double beta1, double beta2, double eps,
                   double weight_decay, int64_t step, void* stream);

/* ---- EMA of the weights (common/trainer.py:264-268 EMAModel(decay=0.999), :350-351 ema_model.step every
 * optimiser step).  diffusers' EMAModel.step is, per tensor, s.sub_(one_minus_decay * (s - p)) - three tensor ops,
 * each rounded to the storage dtype - with p the value the optimiser just stored; the arithmetic here keeps those
 * roundings.  one_minus_decay = 1 - get_decay(optimization_step) is formed by the caller (yat_b200/ema.py).
 * yat_adamw_ema_step: yat_adamw_step + the shadow update in the SAME pass (+4 B/param instead of a separate
 *   6 B/param pass); ema_ptrs [T] int64 = shadow pointer of each tensor of tensor_tab (NULL = plain AdamW).
 * yat_ema_step: the update alone; tensor_tab rows = {param ptr, shadow ptr, unused, numel}.
 */
int yat_adamw_ema_step(const int64_t* tensor_tab, const int64_t* grad_ptrs, const int32_t* chunks, int n_chunks,
                       int dtype, const float* clip_coef, double lr, double beta1, double beta2, double eps,
                       double weight_decay, int64_t step, const int64_t* ema_ptrs, double ema_one_minus_decay,
                       void* stream);
int yat_ema_step(const int64_t* tensor_tab, const int32_t* chunks, int n_chunks, int dtype, double one_minus_decay,
                 void* stream);

/* ---- (next, N1) projection GEMM on tcgen05:  c[M, N] = a[M, K] . w[N, K]^T (+ bias[N]) -----------------------
 * The `x @ W^T` form of torch F.linear, i.e. every projection of the SANA block (diffusers Attention to_q / to_k /
 * to_v / to_out and the token-major 1x1 convolutions of GLUMBConv, utils/patch_sana_attention_layers.py:84-113).
 * a [M, K] and w [N, K] bf16 row-major with row pitches lda / ldw (elements, multiples of 8), bias bf16 [N] or NULL
 * (added in fp32 before the single rounding), c [M, N] bf16 with pitch ldc (multiple of 16, c 32-byte aligned).
 * fp32 accumulation in TMEM, cta_group::2 (two CTAs per 256 x tile_n tile), TMA operand staging, persistent grid.
 * tile_n: 256, 224 or 128; 0 = pick (224 when N is a multiple of 224 but not of 256: all SANA-1.6B widths).
 */
int yat_gemm_bf16_tn(const void* a, const void* w, const void* bias, void* c, int64_t M, int N, int K,
                     int64_t lda, int64_t ldw, int64_t ldc, int tile_n, void* stream);

/* ---- (next, N1) to_q | to_k | to_v projection with the linear attention's phase 1 as its epilogue ----------------
 * Replaces `attn.to_q / to_k / to_v` + the first half of diffusers' SanaLinearAttnProcessor2_0 (reached from
 * utils/patch_sana_attention_layers.py:94) by ONE tcgen05 kernel:
 *   qkv   [B, N, 3*H*32] bf16 = x [B, N, Cin] . w_qkv[3*H*32, Cin]^T      (q | k | v thirds; what the backward reads)
 *   state [B, 2*ceil(N/256), H, 33, 32] fp32: per 128-token tile and head, rows 0..31 = relu(k)^T v, row 32 = sum relu(k)
 * yat_qkv_linattn_state_bytes gives the size of `state`.  yat_linattn_fwd_state is the attention forward that consumes
 * it: it streams only q (row pitch ld elements) and sums the tile partials in fixed order -> o [B, N, H*32] bf16.
 * No bias (SANA's attention_bias is False); head_dim 32 only.
 */
int64_t yat_qkv_linattn_state_bytes(int B, int N, int H);
int yat_qkv_linattn_fwd(const void* x, const void* w_qkv, int B, int N, int H, int D, int Cin, void* qkv, void* state,
                        void* stream);
int yat_linattn_fwd_state(const void* q, int64_t ld, const void* state, int B, int N, int H, int D, void* o,
                          void* stream);

/* ---- cross-attention of the block (attn2): softmax(scale q k^T + bias) v, few keys ------------------------------------
 * Replaces F.scaled_dot_product_attention inside diffusers' AttnProcessor2_0 for `self.attn2(...)` of the SANA / PixArt
 * block (utils/patch_sana_attention_layers.py:100-106; the mask argument is diffusers' additive bias, 0 on caption tokens
 * and -10000 on padding).  q, o [B, Nq, H*d] bf16 (q row pitch ldq elements, o dense), k, v [B, L, H*d] bf16 (row pitch
 * ldkv), bias fp32 [B, L] or NULL, lse fp32 [B, H, Nq] or NULL (ln sum_j exp(scale q.k_j + bias_j), what the backward
 * reads).  d a multiple of 8, <= 128; L <= yat_xattn_max_keys() (384: a whole row of scores lives in tensor memory).
 */
int yat_xattn_max_keys(void);
int yat_xattn_fwd(const void* q, const void* k, const void* v, const void* bias, int B, int H, int Nq, int L, int d,
                  int64_t ldq, int64_t ldkv, float scale, void* o, void* lse, void* stream);

/* yat_xattn_bwd: the backward of the same attention (the autograd of F.scaled_dot_product_attention for `self.attn2(...)`,
 * utils/patch_sana_attention_layers.py:100-106).  o, lse: what yat_xattn_fwd returned; d_o [B, Nq, H*d] dense bf16;
 * dq [B, Nq, H*d], dk, dv [B, L, H*d] dense bf16 (dk and dv may both be NULL: only dq is formed);
 * workspace = yat_xattn_bwd_workspace_bytes(B, H, Nq) bytes (D = rowsum(dO * O), fp32).
 */
int64_t yat_xattn_bwd_workspace_bytes(int B, int H, int Nq);
int yat_xattn_bwd(const void* q, const void* k, const void* v, const void* bias, const void* o, const void* lse,
                  const void* d_o, int B, int H, int Nq, int L, int d, int64_t ldq, int64_t ldkv, float scale,
                  void* dq, void* dk, void* dv, void* workspace, void* stream);

/* ---- bucket assignment (host function; float64 compare, first-minimum wins) ---------------
 *   out[i] = argmin_j |keys[j] - ratios[i]| with strict '<' in table order, or -1 if nkeys==0.
 */
int yat_bucket_assign(const double* ratios, int n, const double* keys, int nkeys, int32_t* out);

/* ---- test hook: raw Philox words for (seed, offset, sample, stream), quads [0, n_quads) ----- */
int yat_philox_raw(uint64_t seed, uint64_t offset, uint32_t sample, uint32_t stream_id,
                   int n_quads, uint32_t* out /* device [n_quads,4] */, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* YAT_B200_H_ */
