/* odvae_b200 — C ABI of the B200-native OD-VAE AutoencoderKL hot path.
 *
 * Every entry point is `extern "C"`, takes plain device pointers + sizes + a CUDA stream handle
 * (`void* stream` == cudaStream_t), enqueues work on that stream and returns immediately with a
 * status code (0 = ok, negative = error).  Nothing here allocates, synchronises or throws.
 * There is NO CPU fallback: on a device that is not sm_100 the calls return ODVAE_ERR_ARCH.
 *
 * Activation layout inside the library is NHWC ("channels-last"); `act_dtype` selects the element
 * type of activation tensors: ODVAE_F32 (validation mode, SIMT kernels) or ODVAE_BF16 (the timed
 * mode; convolutions / GEMMs run on tcgen05 tensor cores with fp32 accumulation in TMEM).
 * Parameters and parameter gradients are fp32 in the reference's own layouts (OIHW conv weights,
 * [out,in] linear weights) so reference checkpoints load unchanged.
 *
 * Each function cites the reference interface it replaces; paths are relative to the reference
 * root, LD/ = submodules/latent-diffusion/.
 */
#ifndef ODVAE_B200_H
#define ODVAE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default)
#endif

#define ODVAE_OK 0
#define ODVAE_ERR_SHAPE (-1)     /* unsupported / inconsistent shape or argument */
#define ODVAE_ERR_ARCH (-2)      /* device is not sm_100 */
#define ODVAE_ERR_WORKSPACE (-3) /* workspace too small */
#define ODVAE_ERR_CUDA (-4)      /* a CUDA runtime/driver call failed (see odvae_last_cuda_error) */

#define ODVAE_F32 0
#define ODVAE_BF16 1
#define ODVAE_F16 2 /* IEEE half: forward activations / packed forward weights of the mixed 16-bit mode */

/* conv geometry selectors (LD/ldm/modules/diffusionmodules/model.py) */
#define ODVAE_CONV_3x3_S1 0  /* Conv2d(k=3,s=1,p=1): model.py:92-106,383-387,487-491,529-533 */
#define ODVAE_CONV_1x1 1     /* Conv2d(k=1): nin_shortcut model.py:115-119, q/k/v/proj :156-175 */
#define ODVAE_CONV_3x3_S2 2  /* F.pad(0,1,0,1) + Conv2d(k=3,s=2,p=0): Downsample model.py:72-76 */

/* conv implementation selectors */
#define ODVAE_IMPL_SIMT 0 /* CUDA-core implicit GEMM, any dtype (validation / fp32 mode) */
#define ODVAE_IMPL_TC 1   /* tcgen05 + TMA implicit GEMM, bf16 only */

const char* odvae_status_string(int status);
const char* odvae_last_cuda_error(void);
/* 1 when the current device is compute capability 10.x */
int odvae_device_supported(void);
int odvae_version(void);
/* number of kernels this library has launched in this process (all streams, all threads) */
int64_t odvae_launch_count(void);
/* test / profiling switch: 0 makes the 3x3 stride-1 convs at W = 256 use the per-tap TMA gather instead of the
 * shared-memory halo slabs (tests compare the two paths). Returns ODVAE_OK. */
int odvae_set_conv_halo(int enabled);

/* ---------------------------------------------------------------------------------------------
 * Layout / dtype conversion at the module boundary (AE.py:223, LD/ldm/models/autoencoder.py:344-349)
 * ------------------------------------------------------------------------------------------- */
/* src: NCHW fp32 [B,C,H,W] -> dst: NHWC act_dtype [B,H,W,Cpad], channels >= C zero-filled.
 * Optional affine `dst = src*scale + shift` fuses the reference's _rescale (AE.py:434-436). */
int odvae_nchw_to_nhwc(const float* src, void* dst, int B, int C, int H, int W, int Cpad, int act_dtype,
                       float scale, float shift, void* stream);
/* src: NHWC act_dtype [B,H,W,Cpad] -> dst NCHW fp32 [B,C,H,W] (first C channels) */
int odvae_nhwc_to_nchw(const void* src, float* dst, int B, int C, int H, int W, int Cpad, int act_dtype,
                       void* stream);
/* global min and max of an fp32 tensor (AE.py:436 `_rescale`): out[0]=min, out[1]=max */
int odvae_minmax(const float* src, int64_t n, float* out2, void* workspace, size_t ws_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Convolution (torch.nn.Conv2d call sites).  x: NHWC [B,H,W,CinP]; y: NHWC
 * [B,Ho,Wo,CoutP].  Weights are pre-packed by odvae_pack_conv_weight from the fp32 OIHW master.
 * ------------------------------------------------------------------------------------------- */
/* Packs fp32 OIHW [Cout,Cin,k,k] (k from geom) into the layouts the kernels consume.
 *   which = 0: forward   [CoutP][taps][CinP]   (K-major rows per output channel)
 *   which = 1: dgrad     [CinP][taps'][CoutP]  (taps rotated 180 degrees); for ODVAE_CONV_3x3_S2 four
 *              input-parity-phase matrices [CinP][nt*CoutP] (nt = 4,2,2,1) are concatenated
 * act_dtype selects bf16 / fp16 (TC path) or f32 (SIMT path) packing. Padded rows/cols are zero. */
int odvae_pack_conv_weight(const float* w_oihw, void* packed, int Cout, int Cin, int geom, int CoutP, int CinP,
                           int which, int act_dtype, void* stream);
/* Multi-tensor form of odvae_pack_conv_weight: ONE launch re-packs every entry of a device-resident table (the packed
 * 16-bit copies of all conv weights after an optimizer step). Row layout, 10 x int64 per entry:
 *   { w_oihw pointer, packed pointer, Cout, Cin, k (1|3), CoutP, CinP, which | (stride-2 dgrad ? 2 : 0), act_dtype,
 *     first block }, first block = running sum of odvae_pack_conv_weights_mt_blocks over the previous entries;
 * n_blocks = the total. Same element mapping as odvae_pack_conv_weight (bit-identical output). */
int64_t odvae_pack_conv_weights_mt_blocks(int CoutP, int CinP, int geom);
int odvae_pack_conv_weights_mt(const int64_t* table_dev, int n_entries, int64_t n_blocks, void* stream);
/* 1 when fwd, dgrad and wgrad of this conv can all run on the tcgen05 engine (TMA tiling constraints:
 * CinP % 64 == 0, CoutP % 64 == 0 for dgrad, power-of-two output width, >= 128 output pixels per image). */
int odvae_conv2d_tc_supported(int B, int H, int W, int CinP, int CoutP, int geom);
/* y = conv(x, w) + bias (+ residual).  `residual` (same layout and format as y) may be NULL.
 * act_dtype: format of x and w_packed (f32 | bf16 | fp16); out_dtype: format of the NHWC y / residual (f32 iff
 * act_dtype is f32; bf16 or fp16 otherwise, independent of the operand format).
 * out_nchw_f32 != 0 writes y as NCHW fp32 [B,Cout,Ho,Wo] instead (decoder.conv_out -> dec_obj). */
int odvae_conv2d_fwd(const void* x, const void* w_packed, const float* bias, const void* residual, void* y,
                     int B, int H, int W, int CinP, int CoutP, int Cout, int geom, int act_dtype, int out_dtype,
                     int impl, int out_nchw_f32, void* stream);
/* dx = conv_transpose(dy, w) (+ dx_add).  dy: NHWC [B,Ho,Wo,CoutP]; dx: NHWC [B,H,W,CinP].
 * `w_packed` is the which=1 packing. */
int odvae_conv2d_dgrad(const void* dy, const void* w_packed_dgrad, const void* dx_add, void* dx, int B, int H,
                       int W, int CinP, int CoutP, int geom, int act_dtype, int impl, void* stream);
/* dw (fp32 OIHW [Cout,Cin,kh,kw]) = sum over batch/pixels of dy (x) x ; db[Cout] = sum dy.
 * Deterministic: split-K partials go to `workspace` and are reduced in a fixed order.
 * accumulate != 0 adds into dw/db instead of overwriting. db may be NULL.
 * out_scale multiplies dw and db: the inverse of the static power-of-two scale that 16-bit activation gradients carry in
 * the all-fp16 mode (1.0 otherwise). */
size_t odvae_conv2d_wgrad_workspace_bytes(int B, int H, int W, int CinP, int CoutP, int geom, int impl);
int odvae_conv2d_wgrad(const void* x, const void* dy, float* dw_oihw, float* db, int B, int H, int W, int CinP,
                       int CoutP, int Cin, int Cout, int geom, int act_dtype, int impl, int accumulate, float out_scale,
                       void* workspace, size_t ws_bytes, void* stream);

/* nearest-neighbour 2x upsample (F.interpolate, model.py:54) and its adjoint (2x2 sum-pool) */
int odvae_upsample2x(const void* x, void* y, int B, int H, int W, int C, int act_dtype, void* stream);
int odvae_upsample2x_bwd(const void* dy, void* dx, int B, int H, int W, int C, int act_dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Batched GEMM on tcgen05 (attention QK^T / PV and their gradients; torch.bmm at model.py:186-198).
 * D[b] = alpha * op(A[b]) * op(B[b])^T (+bias[n]) (+residual), bf16 in, fp32 accumulate.
 *   a_mn = 0: A stored [M][K] (K contiguous, ld = lda);  a_mn = 1: A stored [K][M] (M contiguous)
 *   b_mn = 0: B stored [N][K] (K contiguous, ld = ldb);  b_mn = 1: B stored [K][N] (N contiguous)
 * out_dtype: ODVAE_F32 or ODVAE_BF16.  strides are in elements.
 * ------------------------------------------------------------------------------------------- */
int odvae_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K, int batch, int64_t lda,
                    int64_t ldb, int64_t ldd, int64_t strideA, int64_t strideB, int64_t strideD, int a_mn,
                    int b_mn, int out_dtype, float alpha, const float* bias, const void* residual, void* stream);

/* Same contraction with the operand format chosen by the caller: a_dtype == b_dtype in {ODVAE_BF16, ODVAE_F16}
 * (measured on B200: tcgen05.mma kind::f16 with DIFFERENT A and B formats raises an illegal-instruction trap, so a
 * mixed pair returns ODVAE_ERR_SHAPE); out_dtype in {ODVAE_F32, ODVAE_BF16, ODVAE_F16} is independent of the
 * operand format; residual (if any) has the output dtype. */
int odvae_gemm_16(const void* A, const void* B, void* D, int M, int N, int K, int batch, int64_t lda, int64_t ldb,
                  int64_t ldd, int64_t strideA, int64_t strideB, int64_t strideD, int a_mn, int b_mn, int a_dtype,
                  int b_dtype, int out_dtype, float alpha, const float* bias, const void* residual, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Fused attention core of AttnBlock (model.py:187-197: w_ = softmax(q^T k * C^-0.5), single head), tcgen05 only.
 * odvae_attn_scores_softmax: P[b] = softmax_rows(scale * q[b] k[b]^T) written ONCE in `dtype` (bf16 / fp16); the
 *   fp32 score matrix never reaches HBM. Two passes of the q k^T GEMM: pass 1's epilogue leaves per-row
 *   (max, sum exp) partials per column tile in the workspace, pass 2's epilogue combines them in a fixed order and
 *   writes the normalised probabilities. q, k: [B][N][C] views with row strides ldq / ldk and batch strides.
 * odvae_attn_dsoftmax: dS[b] = P[b] * (dO[b] v[b]^T - delta[b][:,None]) * scale, the softmax backward fused into the
 *   dP GEMM's epilogue; delta[b][i] = sum_c dO[b][i][c] O[b][i][c] (= sum_j P_ij dP_ij), from odvae_rowdot.
 * ------------------------------------------------------------------------------------------- */
size_t odvae_attn_softmax_workspace_bytes(int B, int N);
int odvae_attn_scores_softmax(const void* q, const void* k, void* P, int B, int N, int C, int64_t ldq, int64_t ldk,
                              int64_t strideq, int64_t stridek, float scale, int dtype, void* workspace,
                              size_t ws_bytes, void* stream);
int odvae_attn_dsoftmax(const void* dO, const void* v, const void* P, const float* delta, void* dS, int B, int N,
                        int C, int64_t ldo, int64_t ldv, int64_t strideo, int64_t stridev, float scale, int dtype,
                        void* stream);
/* out[r] = sum_c a[r*lda + c] * b[r*ldb + c], c < C (C % 8 == 0), 16-bit inputs of `dtype`, fp32 accumulate */
int odvae_rowdot(const void* a, const void* b, float* out, int64_t rows, int C, int64_t lda, int64_t ldb, int dtype,
                 void* stream);

/* ---------------------------------------------------------------------------------------------
 * GroupNorm(32, C, eps=1e-6, affine) [+ SiLU]  (Normalize model.py:38-39, nonlinearity :33-35)
 * x, y: NHWC [B,HW,C]; stats: fp32 [B,G,2] = (mean, rstd).
 * ------------------------------------------------------------------------------------------- */
size_t odvae_groupnorm_workspace_bytes(int B, int HW, int C, int G);
int odvae_groupnorm_stats(const void* x, float* stats, int B, int HW, int C, int G, float eps, int act_dtype,
                          void* workspace, size_t ws_bytes, void* stream);
/* Convolution forward (same arguments as odvae_conv2d_fwd, NHWC output only) that ALSO produces the GroupNorm
 * statistics of its output y -- Normalize of the next ResnetBlock / AttnBlock (model.py:38-39,121-141) -- from the fp32
 * accumulators in the epilogue, so the separate statistics pass over y disappears. Served by the tcgen05 swapped-role
 * tiles (Cout = 128 layers); odvae_conv2d_fwd_gnstats_workspace_bytes returns 0 for shapes it does not serve and the
 * caller then uses odvae_conv2d_fwd + odvae_groupnorm_stats. stats: fp32 [B,G,2] = (mean, rstd). */
size_t odvae_conv2d_fwd_gnstats_workspace_bytes(int B, int H, int W, int CinP, int CoutP, int geom, int G);
int odvae_conv2d_fwd_gnstats(const void* x, const void* w_packed, const float* bias, const void* residual, void* y,
                             float* stats, int B, int H, int W, int CinP, int CoutP, int geom, int G, float eps,
                             int act_dtype, int out_dtype, void* workspace, size_t ws_bytes, void* stream);
/* y (and, if y2 != NULL, a second copy y2 in out2_dtype) = [silu](GN(x)). Format triples (in, out, out2):
 * f32->f32; bf16->bf16(+bf16); fp16->fp16(+bf16: the copy a weight-gradient GEMM contracts with bf16 dy);
 * fp16->bf16 (attention blocks of the mixed mode).
 * overflow_flags (may be NULL): int32[2] on the device. fp16 inputs: [0] is set to 1 when x holds a value at +-65504, i.e. its
 * producer's store saturated (forward range overflow); odvae_groupnorm_bwd sets [1] likewise for an fp16 dy (gradient scale too
 * large). The flags are sticky; the host reads and clears them (PoseAutoencoder.check_overflow). */
int odvae_groupnorm_apply(const void* x, const float* stats, const float* gamma, const float* beta, void* y, void* y2,
                          int B, int HW, int C, int G, int silu, int in_dtype, int out_dtype, int out2_dtype,
                          int32_t* overflow_flags, void* stream);
/* GroupNorm(32, eps)+[SiLU] statistics AND apply in one launch for small fp16 feature maps (HW <= 1024, C % 64 == 0: the
 * 32x32 / 16x16 levels of model.py:38-39,33-35 whose three-launch form is launch-latency bound): y = [silu](GN(x)), stats =
 * fp32 [B,G,2] (mean, rstd) for the backward. Served: act_dtype fp16, HW <= 1024, C % 64 == 0, channels per group in
 * {2,4,...,64}; anything else returns ODVAE_ERR_SHAPE (use odvae_groupnorm_stats + odvae_groupnorm_apply).
 * odvae_groupnorm_bwd picks its one-launch form for the same shapes by itself (fp16 x and fp16 gradients). */
int odvae_groupnorm_fwd_small(const void* x, const float* gamma, const float* beta, void* y, float* stats, int B, int HW,
                              int C, int G, float eps, int silu, int act_dtype, int32_t* overflow_flags, void* stream);
/* dx = d/dx [silu](GN(x)) . dy (+ dx_add);  dgamma/dbeta fp32 [C] (accumulate != 0 adds).
 * act_dtype = format of x, grad_dtype = format of dy / dx_add / dx: (f32,f32), (bf16,bf16), (fp16,bf16), (fp16,fp16).
 * pgrad_scale multiplies dgamma / dbeta (inverse of the static activation-gradient scale of the all-fp16 mode). */
int odvae_groupnorm_bwd(const void* x, const float* stats, const float* gamma, const float* beta, const void* dy,
                        const void* dx_add, void* dx, float* dgamma, float* dbeta, int B, int HW, int C, int G,
                        int silu, int act_dtype, int grad_dtype, int accumulate, float pgrad_scale, int32_t* overflow_flags, void* workspace, size_t ws_bytes,
                        void* stream);
/* GroupNorm+SiLU backward with the reduce pass fused into the producing dgrad (model.py:121-141 backward: conv -> GroupNorm ->
 * SiLU chains of the 128-channel ResnetBlocks): odvae_conv2d_dgrad_gnbwd is odvae_conv2d_dgrad (stride-1 geometries, fp16) that
 * ALSO writes, per 128-pixel box and channel, (sum dy', sum dy' xhat) for the GroupNorm whose OUTPUT gradient it produces --
 * gn_x = that GroupNorm's input (same NHWC layout as dx), gn_stats / gamma / beta as in odvae_groupnorm_bwd, part = fp32
 * [B * boxes * CinP * 2] with boxes = H * W / 128. ..._workspace_bytes returns part's size, or 0 when the shape is not served
 * (tcgen05 swapped-role tiles with a pixel stride of 128 elements only). odvae_groupnorm_bwd_from_partials then finishes the
 * backward (finalize + one pass over x and dy); same outputs as odvae_groupnorm_bwd to fp16 rounding of the elementwise chain. */
size_t odvae_conv2d_dgrad_gnbwd_workspace_bytes(int B, int H, int W, int CinP, int CoutP, int geom, int G);
int odvae_conv2d_dgrad_gnbwd(const void* dy, const void* w_packed_dgrad, void* dx, const void* gn_x, const float* gn_stats,
                             const float* gamma, const float* beta, float* part, size_t part_bytes, int B, int H, int W,
                             int CinP, int CoutP, int geom, int G, int act_dtype, void* stream);
int odvae_groupnorm_bwd_from_partials(const void* x, const float* stats, const float* gamma, const float* beta,
                                      const void* dy, const void* dx_add, void* dx, float* dgamma, float* dbeta,
                                      const float* part, int nbox, int B, int HW, int C, int G, int accumulate,
                                      float pgrad_scale, int32_t* overflow_flags, void* workspace, size_t ws_bytes,
                                      void* stream);

/* ---------------------------------------------------------------------------------------------
 * Attention core of AttnBlock (model.py:183-198): single head, scale = C^-0.5.
 * qkv: NHWC-token layout [B,N,3C] (q | k | v per token).  P: bf16/f32 [B,N,N] probabilities (saved
 * for backward).  o: [B,N,C].
 * ------------------------------------------------------------------------------------------- */
/* row softmax of fp32 scores S[rows][n] * scale -> P (act_dtype), in place allowed when f32 */
int odvae_softmax_rows(const float* S, void* P, int64_t rows, int n, float scale, int act_dtype, void* stream);
/* dS = P * (dP - rowsum(dP*P)) * scale ; dP fp32 [rows][n], P act_dtype, dS act_dtype */
int odvae_softmax_rows_bwd(const float* dP, const void* P, void* dS, int64_t rows, int n, float scale,
                           int act_dtype, void* stream);
/* SIMT batched GEMM for the f32 validation mode (same operand conventions as odvae_gemm_bf16) */
int odvae_gemm_simt(const void* A, const void* B, void* D, int M, int N, int K, int batch, int64_t lda,
                    int64_t ldb, int64_t ldd, int64_t strideA, int64_t strideB, int64_t strideD, int a_mn,
                    int b_mn, int in_dtype, int out_dtype, float alpha, const float* bias, const void* residual,
                    void* stream);

/* ---------------------------------------------------------------------------------------------
 * Latent sampling + KL (DiagonalGaussianDistribution DIST.py:24-46, dropout/noise AE.py:226-241)
 * moments: NHWC act_dtype [B,HW,ld_moments], channels [0,Cz) = mean, [Cz,2Cz) = logvar, rest padding.
 * z / eps / keep / noise may be NULL (mode() / no dropout / no noise); dz / dkl may be NULL in bwd.
 * outputs (all fp32, NCHW [B,Cz,h,w] = the module-boundary layout): mean, logvar (clamped).
 * z: NHWC act_dtype [B,HW,CzPad] = (mean + std*eps)*keep*keep_scale + noise + enc_pose (zero padded).
 * kl[B] = 0.5*sum(mean^2 + var - 1 - logvar).  eps/noise fp32 NCHW, keep uint8 NCHW, enc_pose fp32 [B,Cz*HW]
 * (NCHW flattened; may be NULL). keep_scale_dev (may be NULL): device float that overrides `keep_scale` -- the dropout
 * ramp changes 1/(1-p) every step, and a CUDA-graph replay of the train step reads it from memory.
 * ------------------------------------------------------------------------------------------- */
int odvae_sample_kl_fwd(const void* moments, const float* eps, const uint8_t* keep, const float* noise,
                        const float* enc_pose, float keep_scale, float* mean, float* logvar, void* z,
                        float* kl, int B, int HW, int Cz, int CzPad, int ld_moments, int act_dtype,
                        const float* keep_scale_dev, void* stream);
/* dmoments = d(z)/d(moments)^T dz + dkl[b] * dKL/dmoments (+ direct grads on mean/logvar outputs).
 * dz: NHWC act_dtype [B,HW,CzPad]; denc_pose fp32 [B,Cz*HW] (NCHW flattened) = dz * dz_scale.
 * dz_scale / out_scale: the decoder-side gradient arrives multiplied by a static scale (dz_scale = its inverse) and the
 * encoder-side gradient leaves multiplied by out_scale (all-fp16 mode; 1.0 otherwise). act_dtype f32 / bf16 / fp16. */
int odvae_sample_kl_bwd(const void* dz, const float* dkl, const float* dmean_direct, const float* dlogvar_direct,
                        const float* mean, const float* logvar, const float* eps, const uint8_t* keep,
                        float keep_scale, void* dmoments, float* denc_pose, int B, int HW, int Cz, int CzPad,
                        int ld_moments, int act_dtype, float dz_scale, float out_scale, const float* keep_scale_dev,
                        void* stream);

/* ---------------------------------------------------------------------------------------------
 * Masked L1 reconstruction term (PoseLoss._get_rec_loss/_get_nll_loss LOSS.py:134-158, masks :250-257)
 * x, xhat: NCHW fp32 [B,3,H,W]; mask2d: fp32 [B,1,H,W]; sums[b] = sum |x*m - xhat*m| (0 if !use_pixel).
 * sgn (int8 NCHW) = sign(x*m - xhat*m), saved for backward (the backward multiplies by the float mask, so a soft,
 * e.g. bilinearly resized, mask differentiates exactly; the reference dataset's masks are {0,1}).
 * ------------------------------------------------------------------------------------------- */
int odvae_l1_masked_fwd(const float* x, const float* xhat, const float* mask2d, float* sums, int8_t* sgn, int B,
                        int C, int HW, void* workspace, size_t ws_bytes, void* stream);
size_t odvae_l1_masked_workspace_bytes(int B, int C, int HW);
/* dxhat[b,c,p] = -sgn * mask2d[b,p] * coef[b]  (coef folds mask_bg, 1/(exp(logvar)+eps), 1/sum(mask_bg), upstream grad) */
int odvae_l1_masked_bwd(const int8_t* sgn, const float* mask2d, const float* coef, float* dxhat, int B, int C, int HW, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Fused Adam (AE.py:365-377: lr, betas=(0.5,0.9), eps=1e-8) + global-norm clip (CFG:140) on flat
 * fp32 buffers.
 * ------------------------------------------------------------------------------------------- */
size_t odvae_sumsq_workspace_bytes(int64_t n);
/* out1[0] (+)= sum g^2, fixed-order two-stage reduction */
int odvae_sumsq(const float* g, int64_t n, float* out1, int accumulate, void* workspace, size_t ws_bytes,
                void* stream);
/* p -= lr/bc1 * m / (sqrt(v)/sqrt(bc2) + eps) after m,v updates with g * min(1, max_norm/(sqrt(gnorm_sq)+1e-6));
 * gnorm_sq may be NULL (no clipping). bias_corr{1,2} = 1 - beta^t computed by the caller. */
int odvae_adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                    float eps, float bias_corr1, float bias_corr2, const float* gnorm_sq, float max_norm,
                    void* stream);

/* Multi-tensor variants over a device-resident chunk table (no flattening copies): `table` is
 * nrows x {uint64 p, uint64 g, uint64 m, uint64 v, int64 count} (40 bytes per row, count <= 65536);
 * one CTA per row.  odvae_mt_sumsq: out1[0] = sum over all rows of g^2 (fixed order; workspace holds
 * nrows floats; p/m/v entries ignored).  odvae_mt_adam_step: Adam on every row with the gradient scaled by
 * grad_scale (1/world_size after a sum-allreduce) and clipped by max_norm against gnorm_sq (taken over
 * the UNSCALED gradients; NULL = no clipping). */
int odvae_mt_sumsq(const void* table, int nrows, float* out1, void* workspace, size_t ws_bytes, void* stream);
int odvae_mt_adam_step(const void* table, int nrows, float lr, float beta1, float beta2, float eps,
                       float bias_corr1, float bias_corr2, const float* gnorm_sq, float max_norm,
                       float grad_scale, void* stream);
/* Variant with DEVICE-resident step counters, one per parameter (torch.optim.Adam's state["step"] as a device float): table7
 * rows = {p, g, m, v, count, step_ptr, active_ptr}; the kernel computes 1 - beta^t from *step_ptr, and skips rows whose
 * *active_ptr (uint8, may be 0 = always active) is 0 -- a parameter without a gradient this step is neither updated nor aged
 * (torch skips .grad is None). odvae_mt_step_incr advances steps[i] (+1 where active[i]) and runs first. Lets a CUDA-graph
 * replay of the whole train step advance the optimizer without the host. */
int odvae_mt_step_incr(float* steps, const uint8_t* active, int n, void* stream);
int odvae_mt_adam_step_dev(const void* table7, int nrows, float lr, float beta1, float beta2, float eps,
                           const float* gnorm_sq, float max_norm, float grad_scale, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Small heads (csrc/heads.cu): pose-MLP glue and the whole PoseLoss reduction tail, one launch each way.
 * ------------------------------------------------------------------------------------------- */
/* PoseEncoderSpatialVAE tile+add (pose_encoder.py:121-125): h[b][p*F+f] = hx[p*F+f] + hz[b][f];
 * bwd: dhx[P*F] = sum_b dh, dhz[B][F] = sum_p dh. */
int odvae_tile_add_fwd(const float* hx, const float* hz, float* h, int B, int P, int F, void* stream);
int odvae_tile_add_bwd(const float* dh, float* dhx, float* dhz, int B, int P, int F, void* stream);
/* PoseAutoencoder._decode_pose (AE.py:126-160): zz [B][16+ncls] = (mu[8] | logvar[8] | logits[ncls]) ->
 * mean/logvar [B][8] (logvar clamped to [-30,20]) and dec_pose [B][8+ncls] = (mu + exp(logvar/2)*eps | logits).
 * eps == NULL gives mode().  bwd: any of d_dec_pose / d_mean / d_logvar may be NULL. */
int odvae_pose_head_fwd(const float* zz, const float* eps, float* dec_pose, float* mean, float* logvar, int B,
                        int ncls, void* stream);
int odvae_pose_head_bwd(const float* zz, const float* eps, const float* d_dec_pose, const float* d_mean,
                        const float* d_logvar, float* dzz, int B, int ncls, void* stream);
/* PoseLoss.forward tail, optimizer_idx 0 (LOSS.py:221-228,260-278,307-321): mask_bg = class_gt != background_idx,
 * focal class loss (mmdet FocalLoss() defaults), masked MSE bbox / fill-factor, masked L1 + SmoothL1(sin) pose,
 * NLL from the per-sample masked-L1 sums (l1_sums[B], NULL = pixel loss off; npix = C*H*W per sample) with the
 * learned scalar `logvar`, masked KL(obj) (kl_obj[B], may be NULL), per-class prior KL of the bbox posterior
 * (prior [L][8][2] = (mean, clamped logvar); prior_row[b] = table row or -1 for the "background" label; the
 * [8,8] cross term of src/util/distributions.py:29-41 is reproduced).  Terms divided by sum(mask_bg) are 0
 * when it is 0.  weights7 (HOST pointer) = {nll, kl_obj, pose, class, bbox, fill, kl_bbox} weights of `total`.
 * out16: 0 total 1 nll 2 kl_obj 3 rec_mean 4 pose 5 class 6 bbox 7 fill 8 kl_bbox 9-12 mean t1,t2,t3,v3
 * 13 sum(mask_bg).  bwd takes dout16 (gradient of every out entry) and writes the input gradients
 * (d_l1_sums / d_kl_obj may be NULL). */
int odvae_loss_heads_fwd(const float* dec_pose, const float* bbox_mean, const float* bbox_logvar,
                         const float* l1_sums, const float* kl_obj, const float* logvar, const float* pose_gt,
                         const float* bbox_gt, const float* fill_gt, const int64_t* class_gt, const int* prior_row,
                         const float* prior, int B, int ncls, int background_idx, float npix, const float* weights7,
                         float* out16, void* stream);
int odvae_loss_heads_bwd(const float* dec_pose, const float* bbox_mean, const float* bbox_logvar,
                         const float* l1_sums, const float* kl_obj, const float* logvar, const float* pose_gt,
                         const float* bbox_gt, const float* fill_gt, const int64_t* class_gt, const int* prior_row,
                         const float* prior, int B, int ncls, int background_idx, float npix, const float* weights7,
                         const float* out16, const float* dout16, float* d_dec_pose, float* d_bbox_mean,
                         float* d_bbox_logvar, float* d_l1_sums, float* d_kl_obj, float* d_logvar, void* stream);

/* generic elementwise helpers on activations (residual adds that cannot be fused) */
int odvae_add(const void* a, const void* b, void* out, int64_t n, int act_dtype, void* stream);

/* dst = 2*(src - min)/(max - min) - 1 with minmax = {min, max} on the device (AE.py:434-436 `_rescale`) */
int odvae_rescale(const float* src, float* dst, int64_t n, const float* minmax, void* stream);
/* Input path in ONE pass (AE.py:296-304,434-436 + LD autoencoder.py:344-349): src NCHW fp32 [B,C,H,W] ->
 * dst_nchw = 2*(src-min)/(max-min)-1 (fp32 NCHW: the loss's rgb_gt) AND dst_nhwc = the same values as NHWC act_dtype
 * [B,H,W,Cpad] (channels >= C zero): encoder.conv_in's operand. Same arithmetic as odvae_rescale (bit-identical values). */
int odvae_rescale_nhwc(const float* src, const float* minmax, float* dst_nchw, void* dst_nhwc, int B, int C, int H, int W,
                       int Cpad, int act_dtype, void* stream);

/* Thin-channel 3x3 / stride 1 / pad 1 im2col for the 3-channel edge convs (encoder.conv_in, LD model.py:383-387; the
 * backward of decoder.conv_out, LD model.py:529-533): src NCHW fp32 [B,C,H,W], C <= 4 -> col NHWC act_dtype [B,H,W,64] with
 * col[.., t*4 + c] = scale * src[b, c, y + s*(t/3-1), x + s*(t%3-1)], s = -1 when flip (data-gradient operand: dy as seen
 * from the input pixel), zero outside the image, in channels >= C and in columns 36..63. The conv then is a 1x1 conv /
 * GEMM over K = 64: y = col W1 with W1[co][t*4+c] = w[co][c][t]; dx = col(dy, flip) Wd with Wd[ci][t*4+co] = w[co][ci][t];
 * dW from the 1x1 weight gradient of (col, dy) resp. (x, col(dy, flip)). 16-bit formats only. */
int odvae_im2col3x3_thin(const float* src, void* col, int B, int C, int H, int W, int flip, float scale, int act_dtype,
                         void* stream);

/* Pose MLP activations on fp32 [n] (pose_decoder.py:82-94 Tanh, pose_encoder.py:82-98 SiLU):
 * act = 1 tanh, 2 silu.  bwd: dx = dy * act'(x) with x the pre-activation. */
int odvae_act_fwd(const float* x, float* y, int64_t n, int act, void* stream);
int odvae_act_bwd(const float* x, const float* dy, float* dx, int64_t n, int act, void* stream);

/* 16-bit format conversion, n % 8 == 0: (ODVAE_F16 -> ODVAE_BF16) or (ODVAE_BF16 -> ODVAE_F16, saturating).
 * Mixed mode: the bf16 copy of an fp16 forward activation that a weight-gradient GEMM contracts with bf16 dy. */
int odvae_cast16(const void* src, void* dst, int64_t n, int src_dtype, int dst_dtype, void* stream);

/* ---------------------------------------------------------------------------------------------
 * LPIPS perceptual distance. Call site: PoseLoss._get_rec_loss (src/modules/losses/contperceptual.py:141-145)
 * -> taming.modules.losses.lpips.LPIPS (un-vendored; public definition in SURVEY appendix B). The VGG16 3x3 convs run
 * through odvae_conv2d_fwd / odvae_conv2d_dgrad; these are the kernels around them. dtype: ODVAE_F32 or ODVAE_BF16,
 * tensors NHWC with C % 8 == 0 (bf16) / C % 4 == 0 (f32).
 * ------------------------------------------------------------------------------------------- */
/* nn.ReLU(inplace=True) of torchvision's vgg16.features; bwd: dx = dy where the ReLU OUTPUT y > 0 (dx may alias dy) */
int odvae_relu(void* x, int64_t n, int dtype, void* stream);
int odvae_relu_bwd(const void* y, const void* dy, void* dx, int64_t n, int dtype, void* stream);
/* nn.MaxPool2d(2, 2): x [B,H,W,C] -> y [B,H/2,W/2,C]; bwd routes dy to the first maximum of each window (torch order) */
int odvae_maxpool2x2(const void* x, void* y, int B, int H, int W, int C, int dtype, void* stream);
int odvae_maxpool2x2_bwd(const void* x, const void* dy, void* dx, int B, int H, int W, int C, int dtype, void* stream);
/* One LPIPS level: out[b] (+)= mean_p sum_c w[c] * (f0/(|f0|_c + 1e-10) - f1/(|f1|_c + 1e-10))^2 over features
 * f0, f1 [B][HW][C] (normalize_tensor + squared difference + NetLinLayer 1x1 conv + spatial_average);
 * bwd: df1 (+)= dout[b] * d out[b] / d f1 (the reconstruction's features; the input's tower needs no gradient). */
size_t odvae_lpips_level_workspace_bytes(int B);
int odvae_lpips_level_fwd(const void* f0, const void* f1, const float* w, float* out, int B, int HW, int C, int dtype,
                          int accumulate, void* workspace, size_t ws_bytes, void* stream);
int odvae_lpips_level_bwd(const void* f0, const void* f1, const float* w, const float* dout, void* df1, int B, int HW,
                          int C, int dtype, int accumulate, void* stream);

/* out[c] (+)= sum_r m[r*ld + c]  (nn.Linear bias gradients); dtype = ODVAE_F32 / ODVAE_BF16 */
size_t odvae_colsum_workspace_bytes(int cols);
int odvae_colsum(const void* m, int64_t rows, int cols, int ld, int dtype, float* out, int accumulate,
                 void* workspace, size_t ws_bytes, void* stream);
/* odvae_colsum without a workspace and without any allocation. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, serving loops, harnesses that
 * forbid it). The stage-1 partial sums live in a file-scope `static __device__` array inside the library (592 x
 * ODVAE_COLSUM_STATIC_MAX_COLS floats = 1.16 MiB, held by every GPU that loads the library until exit); the call performs
 * no cudaMalloc / cudaFree and calls no allocating library routine. Same kernels, same summation order, same result bits
 * as odvae_colsum. cols > ODVAE_COLSUM_STATIC_MAX_COLS returns ODVAE_ERR_WORKSPACE (odvae_last_cuda_error() says why).
 * All calls on one device SHARE the array: calls that may overlap on the GPU (different streams without an ordering
 * between them, concurrent graph branches) race on it -- order them, or use odvae_colsum with separate workspaces.
 * Nothing in the array needs clearing between calls (every partial that is read was written by the same call). */
#define ODVAE_COLSUM_STATIC_MAX_COLS 512
int odvae_colsum_static(const void* m, int64_t rows, int cols, int ld, int dtype, float* out, int accumulate,
                        void* stream);

/* ---------------------------------------------------------------------------------------------
 * Upsample: F.interpolate(scale 2, nearest) + Conv2d(k=3, s=1, p=1) (LD model.py:53-57) WITHOUT the upsampled tensor: output
 * pixel (2i+py, 2j+px) only ever reads 2x2 low-resolution pixels, so the layer is four parity-phase 2x2 convolutions with
 * pre-summed weights over x: NHWC [B,H,W,CinP] -> y: NHWC [B,2H,2W,CoutP] (16/36 of the FLOPs of the literal form).
 *   odvae_pack_upconv_weight: fp32 OIHW [Cout,Cin,3,3] -> which = 0: [4 phases][CoutP][4 taps][CinP] (forward);
 *                             which = 1: [CinP][16 (phase, tap)][CoutP] (dgrad: ONE launch over the stride-2 phase view of dy)
 *   odvae_upconv_wgrad: four 4-tap weight-gradient GEMMs (dy sub-sampled per phase) folded back onto the 3x3 filter;
 *                       deterministic; out_scale / accumulate as in odvae_conv2d_wgrad.
 * ------------------------------------------------------------------------------------------- */
int odvae_upconv_tc_supported(int B, int H, int W, int CinP, int CoutP);
int odvae_pack_upconv_weight(const float* w_oihw, void* packed, int Cout, int Cin, int CoutP, int CinP, int which,
                             int act_dtype, void* stream);
int odvae_upconv_fwd(const void* x, const void* w_phase_packed, const float* bias, void* y, int B, int H, int W, int CinP,
                     int CoutP, int act_dtype, int impl, void* stream);
int odvae_upconv_dgrad(const void* dy, const void* w_dgrad_packed, void* dx, int B, int H, int W, int CinP, int CoutP,
                       int act_dtype, int impl, void* stream);
size_t odvae_upconv_wgrad_workspace_bytes(int B, int H, int W, int CinP, int CoutP, int Cin, int Cout, int impl);
int odvae_upconv_wgrad(const void* x, const void* dy, float* dw_oihw, float* db, int B, int H, int W, int CinP, int CoutP,
                       int Cin, int Cout, int act_dtype, int impl, int accumulate, float out_scale, void* workspace,
                       size_t ws_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * PatchGAN discriminator. Call sites: src/modules/losses/contperceptual.py:280-301 (generator side:
 * logits_fake = discriminator(reconstructions)) and :352-375 (discriminator side) ->
 * taming.modules.discriminator.model.NLayerDiscriminator (un-vendored; public definition in SURVEY appendix B:
 * Conv2d(k=4, s=2|1, p=1) [+ BatchNorm2d, training mode] + LeakyReLU(0.2)). Each conv = im2col + GEMM
 * (odvae_gemm_16 / odvae_gemm_simt); these are the kernels around the GEMMs. dtypes: ODVAE_F32 / BF16 / F16.
 * ------------------------------------------------------------------------------------------- */
/* patches[m][(r*4+s)*C + c] = act(src[b, oy*stride-1+r, ox*stride-1+s, c]) (zero outside the image), m = (b, oy, ox),
 * Ho = (H-2)/stride + 1.  src_nchw = 0: src NHWC [B,H,W,C] (C % 8 == 0, Kp == 16 C), act(v) = lrelu(v*scale[c]+shift[c])
 * when scale != NULL (the previous layer's train-mode BatchNorm folded in), lrelu(v) when slope != 1, else identity.
 * src_nchw = 1: src NCHW fp32 [B,C,H,W] with small C (the image), times mask[B,H,W] when given (`x * mask_2d_bbox`,
 * contperceptual.py:250-257), Kp >= 16 C (columns beyond are zero). */
int odvae_patchgan_im2col(const void* src, const float* mask, const float* scale, const float* shift, float slope,
                          void* dst, int B, int H, int W, int C, int stride, int Kp, int src_nchw, int src_dtype,
                          int dst_dtype, void* stream);
/* gather-form inverse: dst[b,iy,ix,c] = lrelu'(u) * sum_{(r,s) reaching (iy,ix)} dP[m][(r*4+s)*C + c], u = ypre*scale+shift
 * (scale NULL: u = ypre).  dst_nchw = 1: gradient w.r.t. the NCHW fp32 image (times mask), no activation derivative. */
int odvae_patchgan_col2im(const void* dP, const void* ypre, const float* scale, const float* shift, float slope,
                          const float* mask, void* dst, int B, int H, int W, int C, int stride, int Kp, int dst_nchw,
                          int dp_dtype, int y_dtype, int dst_dtype, void* stream);
/* torch.nn.BatchNorm2d in training mode over y[M][C] (M = B*H*W rows): batch mean / biased variance -> scale = gamma*rstd,
 * shift = beta - mean*scale (consumed by the next odvae_patchgan_im2col), mean and rstd saved for backward; running_mean,
 * running_var (unbiased, momentum) and num_batches_tracked updated in place when given. Deterministic two-stage sums. */
size_t odvae_bn_workspace_bytes(int64_t M, int C);
int odvae_bn_train_stats(const void* y, int64_t M, int C, int dtype, const float* gamma, const float* beta, float eps,
                         float momentum, float* running_mean, float* running_var, int64_t* num_batches_tracked,
                         float* scale, float* shift, float* mean, float* rstd, void* workspace, size_t ws_bytes,
                         void* stream);
/* dyp (in: gradient w.r.t. the BatchNorm output, already times lrelu'; out: gradient w.r.t. y), dgamma / dbeta (+)= */
int odvae_bn_train_bwd(const void* y, void* dyp, int64_t M, int C, int y_dtype, int g_dtype, const float* gamma,
                       const float* mean, const float* rstd, float* dgamma, float* dbeta, int accumulate,
                       void* workspace, size_t ws_bytes, void* stream);
/* fp32 OIHW [Cout][Cin][4][4] -> [CoutP][Kp], k = (r*4+s)*Cin + c, zero padded; and the inverse for weight gradients:
 * dw (+)= sum over nsplit slabs part[split][CoutP][Kp] (fixed order) */
int odvae_pack_k4_weight(const float* w, void* out, int Cout, int Cin, int CoutP, int Kp, int dtype, void* stream);
int odvae_unpack_k4_wgrad(const float* part, int nsplit, float* dw, int Cout, int Cin, int CoutP, int Kp, int accumulate,
                          void* stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* ODVAE_B200_H */
