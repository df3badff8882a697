// C ABI for convolution and batched GEMM (see include/odvae_b200.h).
#include "common.cuh"
#include "engine_host.h"
#include "skinny_gemm.cuh"

using namespace odvae;

extern "C" {

int odvae_set_conv_halo(int enabled) {
  set_conv_halo_enabled(enabled);
  return ODVAE_OK;
}

int odvae_conv2d_tc_supported(int B, int H, int W, int CinP, int CoutP, int geom) {
  EngineDesc d;
  if (describe_conv_fwd(d, nullptr, nullptr, nullptr, nullptr, nullptr, B, H, W, CinP, CoutP, CoutP, geom, 0) != ODVAE_OK) return 0;
  if (!d.tc_ok) return 0;
  EngineDesc ds[4];
  int n = 0;
  if (describe_conv_dgrad(ds, &n, nullptr, nullptr, 2, nullptr, nullptr, B, H, W, CinP, CoutP, geom) != ODVAE_OK) return 0;
  for (int i = 0; i < n; ++i)
    if (!ds[i].tc_ok) return 0;
  if (describe_conv_wgrad(d, nullptr, nullptr, nullptr, B, H, W, CinP, CoutP, geom, ODVAE_IMPL_TC) != ODVAE_OK) return 0;
  return d.tc_ok ? 1 : 0;
}

int odvae_conv2d_fwd(const void* x, const void* w_packed, const float* bias, const void* residual, void* y, int B,
                     int H, int W, int CinP, int CoutP, int Cout, int geom, int act_dtype, int out_dtype, int impl,
                     int out_nchw_f32, void* stream) {
  EngineDesc d;
  int st = describe_conv_fwd(d, x, w_packed, bias, residual, y, B, H, W, CinP, CoutP, Cout, geom, out_nchw_f32);
  if (st != ODVAE_OK) return st;
  if (!out_nchw_f32) {
    // NHWC output (and residual) format: fp32 mode keeps fp32; 16-bit modes choose bf16 / fp16 independently of the
    // operand format (mixed mode: bf16 attention projections add into the fp16 residual stream)
    if ((act_dtype == ODVAE_F32) != (out_dtype == ODVAE_F32)) return ODVAE_ERR_SHAPE;
    if (out_dtype == ODVAE_F16) d.p.out_mode = OUT_F16;
  }
  return run_engine(d, impl, act_dtype, static_cast<cudaStream_t>(stream));
}

static bool gnstats_served(const EngineDesc& d, int CoutP, int G) {
  return d.tc_ok && d.p.swap && CoutP % G == 0;
}

size_t odvae_conv2d_fwd_gnstats_workspace_bytes(int B, int H, int W, int CinP, int CoutP, int geom, int G) {
  EngineDesc d;
  if (G <= 0 || describe_conv_fwd(d, nullptr, nullptr, nullptr, nullptr, nullptr, B, H, W, CinP, CoutP, CoutP, geom, 0) != ODVAE_OK)
    return 0;
  if (!gnstats_served(d, CoutP, G)) return 0;
  return static_cast<size_t>(B) * d.p.tiles_x * d.p.tiles_y * CoutP * 2 * sizeof(float);
}

int odvae_conv2d_fwd_gnstats(const void* x, const void* w_packed, const float* bias, const void* residual, void* y,
                             float* stats, int B, int H, int W, int CinP, int CoutP, int geom, int G, float eps,
                             int act_dtype, int out_dtype, void* workspace, size_t ws_bytes, void* stream) {
  EngineDesc d;
  int st = describe_conv_fwd(d, x, w_packed, bias, residual, y, B, H, W, CinP, CoutP, CoutP, geom, 0);
  if (st != ODVAE_OK) return st;
  if (G <= 0 || !gnstats_served(d, CoutP, G)) return ODVAE_ERR_SHAPE;
  if (act_dtype == ODVAE_F32 || out_dtype == ODVAE_F32) return ODVAE_ERR_SHAPE;
  const int boxes = d.p.tiles_x * d.p.tiles_y;
  if (ws_bytes < static_cast<size_t>(B) * boxes * CoutP * 2 * sizeof(float)) return ODVAE_ERR_WORKSPACE;
  if (out_dtype == ODVAE_F16) d.p.out_mode = OUT_F16;
  d.p.gn_part = reinterpret_cast<float*>(workspace);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  st = run_engine(d, ODVAE_IMPL_TC, act_dtype, s);
  if (st != ODVAE_OK) return st;
  const int Ho = geom == ODVAE_CONV_3x3_S2 ? H / 2 : H, Wo = geom == ODVAE_CONV_3x3_S2 ? W / 2 : W;
  return launch_gn_stats_finalize(d.p.gn_part, stats, B, boxes, Ho * Wo, CoutP, G, eps, s);
}

int odvae_conv2d_dgrad(const void* dy, const void* w_packed_dgrad, const void* dx_add, void* dx, int B, int H, int W,
                       int CinP, int CoutP, int geom, int act_dtype, int impl, void* stream) {
  EngineDesc ds[4];
  int n = 0;
  int st = describe_conv_dgrad(ds, &n, dy, w_packed_dgrad, act_dtype == ODVAE_F32 ? 4 : 2, dx_add, dx, B, H, W, CinP,
                               CoutP, geom);
  if (st != ODVAE_OK) return st;
  for (int i = 0; i < n; ++i) {
    if (act_dtype == ODVAE_F16) ds[i].p.out_mode = OUT_F16;  // dx (and dx_add) in the operand format
    st = run_engine(ds[i], impl, act_dtype, static_cast<cudaStream_t>(stream));
    if (st != ODVAE_OK) return st;
  }
  return ODVAE_OK;
}

// dgrad of a 128-channel 3x3 / 1x1 stride-1 conv whose result is the gradient of a GroupNorm+SiLU output: the same launch also
// writes the two GroupNorm-backward sums per (pixel box, channel) (engine_desc.h gnb_*), so odvae_groupnorm_bwd_from_partials
// needs no reduce pass over x and dy. Served by the tcgen05 swapped-role tiles with the fp16 epilogue only.
static bool gnbwd_served(const EngineDesc* ds, int n, int CinP, int G) {
  return n == 1 && ds[0].tc_ok && ds[0].p.swap && G > 0 && CinP % G == 0 && ds[0].p.oX == 128 && TC_FAST_EPI;
}

size_t odvae_conv2d_dgrad_gnbwd_workspace_bytes(int B, int H, int W, int CinP, int CoutP, int geom, int G) {
  EngineDesc ds[4];
  int n = 0;
  if (describe_conv_dgrad(ds, &n, nullptr, nullptr, 2, nullptr, nullptr, B, H, W, CinP, CoutP, geom) != ODVAE_OK) return 0;
  if (!gnbwd_served(ds, n, CinP, G)) return 0;
  return static_cast<size_t>(B) * ds[0].p.tiles_x * ds[0].p.tiles_y * CinP * 2 * sizeof(float);
}

int odvae_conv2d_dgrad_gnbwd(const void* dy, const void* w_packed_dgrad, void* dx, const void* gn_x, const float* gn_stats,
                             const float* gamma, const float* beta, float* part, size_t part_bytes, int B, int H, int W,
                             int CinP, int CoutP, int geom, int G, int act_dtype, void* stream) {
  if (act_dtype != ODVAE_F16) return ODVAE_ERR_SHAPE;
  EngineDesc ds[4];
  int n = 0;
  int st = describe_conv_dgrad(ds, &n, dy, w_packed_dgrad, 2, nullptr, dx, B, H, W, CinP, CoutP, geom);
  if (st != ODVAE_OK) return st;
  if (!gnbwd_served(ds, n, CinP, G)) return ODVAE_ERR_SHAPE;
  TcParams& p = ds[0].p;
  if (part_bytes < static_cast<size_t>(B) * p.tiles_x * p.tiles_y * CinP * 2 * sizeof(float)) return ODVAE_ERR_WORKSPACE;
  p.out_mode = OUT_F16;
  p.gnb_x = gn_x;
  p.gnb_stats = gn_stats;
  p.gnb_gamma = gamma;
  p.gnb_beta = beta;
  p.gnb_part = part;
  p.gnb_cpg = CinP / G;
  return run_engine(ds[0], ODVAE_IMPL_TC, act_dtype, static_cast<cudaStream_t>(stream));
}

size_t odvae_conv2d_wgrad_workspace_bytes(int B, int H, int W, int CinP, int CoutP, int geom, int impl) {
  EngineDesc d;
  if (describe_conv_wgrad(d, nullptr, nullptr, nullptr, B, H, W, CinP, CoutP, geom, impl) != ODVAE_OK) return 0;
  const size_t part = static_cast<size_t>(d.p.ntaps) * d.p.split_k * CoutP * CinP * sizeof(float);
  return part + colsum_workspace_bytes(CoutP) + 256;
}

int odvae_conv2d_wgrad(const void* x, const void* dy, float* dw_oihw, float* db, int B, int H, int W, int CinP,
                       int CoutP, int Cin, int Cout, int geom, int act_dtype, int impl, int accumulate, float out_scale,
                       void* workspace, size_t ws_bytes, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  EngineDesc d;
  float* ws = reinterpret_cast<float*>(workspace);
  int st = describe_conv_wgrad(d, x, dy, ws, B, H, W, CinP, CoutP, geom, impl);
  if (st != ODVAE_OK) return st;
  const size_t part = static_cast<size_t>(d.p.ntaps) * d.p.split_k * CoutP * CinP * sizeof(float);
  if (ws_bytes < part + colsum_workspace_bytes(CoutP)) return ODVAE_ERR_WORKSPACE;
  float* cpart = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + ((part + 255) / 256) * 256);
  // tcgen05 path: the bias gradient (column sums of dy) is accumulated by the weight-gradient kernel's idle epilogue
  // warps from the dy tiles already in shared memory -- no second pass over dy
  const int cs_rows = (d.p.ntaps / (d.p.wg_T > 0 ? d.p.wg_T : 1)) * d.p.split_k * 2;
  const bool fused_db = db != nullptr && impl == ODVAE_IMPL_TC && d.tc_ok &&
                        static_cast<size_t>(cs_rows) * Cout * sizeof(float) <= colsum_workspace_bytes(CoutP);
  if (fused_db) {
    d.p.colsum = cpart;
    d.p.colsum_cols = Cout;
  }
  st = run_engine(d, impl, act_dtype, s);
  if (st != ODVAE_OK) return st;
  // split-K partials -> dw, and (fused_db) the partial column sums of dy -> db, in ONE launch
  st = launch_wgrad_reduce(ws, dw_oihw, Cout, Cin, d.p.ntaps, CoutP, CinP, d.p.split_k, accumulate, s,
                           fused_db ? cpart : nullptr, fused_db ? cs_rows : 0, db, out_scale);
  if (st != ODVAE_OK || fused_db) return st;
  if (db != nullptr) {
    const int Ho = geom == ODVAE_CONV_3x3_S2 ? H / 2 : H, Wo = geom == ODVAE_CONV_3x3_S2 ? W / 2 : W;
    st = launch_colsum(dy, static_cast<long long>(B) * Ho * Wo, Cout, CoutP, act_dtype, db, accumulate, cpart, s, out_scale);
  }
  return st;
}

// ---- nearest-2x upsample + 3x3 conv as parity-phase 2x2 convs over the low-resolution input (LD model.py:53-57) ----
int odvae_upconv_tc_supported(int B, int H, int W, int CinP, int CoutP) {
  EngineDesc ds[4];
  if (describe_upconv_fwd(ds, nullptr, nullptr, 2, nullptr, nullptr, B, H, W, CinP, CoutP) != ODVAE_OK) return 0;
  for (int i = 0; i < 4; ++i)
    if (!ds[i].tc_ok) return 0;
  EngineDesc d;
  if (describe_upconv_dgrad(d, nullptr, nullptr, nullptr, B, H, W, CinP, CoutP) != ODVAE_OK || !d.tc_ok) return 0;
  if (describe_upconv_wgrad(d, 0, nullptr, nullptr, 2, nullptr, B, H, W, CinP, CoutP, ODVAE_IMPL_TC) != ODVAE_OK) return 0;
  return d.tc_ok ? 1 : 0;
}

int odvae_upconv_fwd(const void* x, const void* w_phase_packed, const float* bias, void* y, int B, int H, int W, int CinP,
                     int CoutP, int act_dtype, int impl, void* stream) {
  EngineDesc ds[4];
  int st = describe_upconv_fwd(ds, x, w_phase_packed, act_dtype == ODVAE_F32 ? 4 : 2, bias, y, B, H, W, CinP, CoutP);
  if (st != ODVAE_OK) return st;
  for (int i = 0; i < 4; ++i) {
    if (act_dtype == ODVAE_F16) ds[i].p.out_mode = OUT_F16;
    st = run_engine(ds[i], impl, act_dtype, static_cast<cudaStream_t>(stream));
    if (st != ODVAE_OK) return st;
  }
  return ODVAE_OK;
}

int odvae_upconv_dgrad(const void* dy, const void* w_dgrad_packed, void* dx, int B, int H, int W, int CinP, int CoutP,
                       int act_dtype, int impl, void* stream) {
  EngineDesc d;
  int st = describe_upconv_dgrad(d, dy, w_dgrad_packed, dx, B, H, W, CinP, CoutP);
  if (st != ODVAE_OK) return st;
  if (act_dtype == ODVAE_F16) d.p.out_mode = OUT_F16;
  return run_engine(d, impl, act_dtype, static_cast<cudaStream_t>(stream));
}

static size_t upconv_phase_ws(const EngineDesc& d, int CoutP, int CinP) {
  return ((static_cast<size_t>(d.p.ntaps) * d.p.split_k * CoutP * CinP * sizeof(float) + 255) / 256) * 256;
}

size_t odvae_upconv_wgrad_workspace_bytes(int B, int H, int W, int CinP, int CoutP, int Cin, int Cout, int impl) {
  EngineDesc d;
  if (describe_upconv_wgrad(d, 0, nullptr, nullptr, 2, nullptr, B, H, W, CinP, CoutP, impl) != ODVAE_OK) return 0;
  // split-K partials of one phase (reused) + colsum partials + the 4 folded phase results [4][Cout][Cin][4] + [4][Cout]
  return upconv_phase_ws(d, CoutP, CinP) + colsum_workspace_bytes(CoutP) + 256 +
         (static_cast<size_t>(16) * Cout * Cin + 4 * static_cast<size_t>(Cout)) * sizeof(float) + 256;
}

int odvae_upconv_wgrad(const void* x, const void* dy, float* dw_oihw, float* db, int B, int H, int W, int CinP, int CoutP,
                       int Cin, int Cout, int act_dtype, int impl, int accumulate, float out_scale, void* workspace,
                       size_t ws_bytes, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (ws_bytes < odvae_upconv_wgrad_workspace_bytes(B, H, W, CinP, CoutP, Cin, Cout, impl) - 256) return ODVAE_ERR_WORKSPACE;
  uint8_t* base = reinterpret_cast<uint8_t*>(workspace);
  float* ws = reinterpret_cast<float*>(base);
  EngineDesc d0;
  int st = describe_upconv_wgrad(d0, 0, x, dy, act_dtype == ODVAE_F32 ? 4 : 2, ws, B, H, W, CinP, CoutP, impl);
  if (st != ODVAE_OK) return st;
  const size_t part = upconv_phase_ws(d0, CoutP, CinP);
  float* cpart = reinterpret_cast<float*>(base + part);
  float* dwp = reinterpret_cast<float*>(base + part + ((colsum_workspace_bytes(CoutP) + 255) / 256) * 256);
  float* dbp = dwp + static_cast<size_t>(16) * Cout * Cin;
  for (int ph = 0; ph < 4; ++ph) {
    EngineDesc d;
    st = describe_upconv_wgrad(d, ph, x, dy, act_dtype == ODVAE_F32 ? 4 : 2, ws, B, H, W, CinP, CoutP, impl);
    if (st != ODVAE_OK) return st;
    const int cs_rows = d.p.ntaps * d.p.split_k * 2;
    const bool fused_db = db != nullptr && impl == ODVAE_IMPL_TC && d.tc_ok &&
                          static_cast<size_t>(cs_rows) * Cout * sizeof(float) <= colsum_workspace_bytes(CoutP);
    if (fused_db) {
      d.p.colsum = cpart;
      d.p.colsum_cols = Cout;
    }
    st = run_engine(d, impl, act_dtype, s);
    if (st != ODVAE_OK) return st;
    float* dbp_ph = dbp + static_cast<size_t>(ph) * Cout;
    st = launch_wgrad_reduce(ws, dwp + static_cast<size_t>(ph) * 4 * Cout * Cin, Cout, Cin, 4, CoutP, CinP, d.p.split_k, 0, s,
                             fused_db ? cpart : nullptr, fused_db ? cs_rows : 0, db != nullptr ? dbp_ph : nullptr, 1.0f);
    if (st != ODVAE_OK) return st;
    if (db != nullptr && !fused_db) {
      // column sums of this phase's quarter of dy: a strided gather -> go through the generic colsum on the phase rows
      // (SIMT / fp32 validation path only): rows of one phase are not contiguous, so sum the whole dy once instead
      if (ph == 0) {
        st = launch_colsum(dy, static_cast<long long>(B) * 4 * H * W, Cout, CoutP, act_dtype, dbp_ph, 0, cpart, s, 1.0f);
        if (st != ODVAE_OK) return st;
      } else {
        ODVAE_CUDA_TRY(cudaMemsetAsync(dbp_ph, 0, static_cast<size_t>(Cout) * sizeof(float), s));
      }
    }
  }
  return launch_upconv_wgrad_fold(dwp, db != nullptr ? dbp : nullptr, dw_oihw, db, Cout, Cin, accumulate, out_scale, s);
}

int odvae_gemm_bf16(const void* A, const void* B, void* D, int M, int N, int K, int batch, int64_t lda, int64_t ldb,
                    int64_t ldd, int64_t strideA, int64_t strideB, int64_t strideD, int a_mn, int b_mn, int out_dtype,
                    float alpha, const float* bias, const void* residual, void* stream) {
  EngineDesc d;
  int st = describe_gemm(d, A, B, D, M, N, K, batch, lda, ldb, ldd, strideA, strideB, strideD, a_mn, b_mn,
                         out_dtype == ODVAE_F32 ? OUT_F32 : OUT_BF16, alpha, bias, residual);
  if (st != ODVAE_OK) return st;
  return run_engine(d, ODVAE_IMPL_TC, ODVAE_BF16, static_cast<cudaStream_t>(stream));
}

int odvae_gemm_16(const void* A, const void* B, void* D, int M, int N, int K, int batch, int64_t lda, int64_t ldb,
                  int64_t ldd, int64_t strideA, int64_t strideB, int64_t strideD, int a_mn, int b_mn, int a_dtype,
                  int b_dtype, int out_dtype, float alpha, const float* bias, const void* residual, void* stream) {
  if ((a_dtype != ODVAE_BF16 && a_dtype != ODVAE_F16) || a_dtype != b_dtype) return ODVAE_ERR_SHAPE;
  EngineDesc d;
  const int om = out_dtype == ODVAE_F32 ? OUT_F32 : (out_dtype == ODVAE_F16 ? OUT_F16 : OUT_BF16);
  int st = describe_gemm(d, A, B, D, M, N, K, batch, lda, ldb, ldd, strideA, strideB, strideD, a_mn, b_mn, om, alpha,
                         bias, residual);
  if (st != ODVAE_OK) return st;
  return run_engine(d, ODVAE_IMPL_TC, a_dtype, static_cast<cudaStream_t>(stream));
}

// ---- fused attention core (tcgen05 only): scores + softmax in two GEMM passes, softmax backward in the dP GEMM ----
static int attn_nt(int N) {
  EngineDesc d;
  if (describe_gemm(d, nullptr, nullptr, nullptr, N, N, 64, 1, 64, 64, N, 0, 0, 0, 0, 0, OUT_BF16, 1.0f, nullptr,
                    nullptr) != ODVAE_OK)
    return 0;
  return d.p.num_n_tiles;
}

static constexpr int ATTN_PARTS = tc_epi_warps(TC_SPEC_GEMM_SOFTMAX) / 4;  // column parts of a row-statistics tile

size_t odvae_attn_softmax_workspace_bytes(int B, int N) {
  // row partials (ATTN_PARTS column parts per tile, one per epilogue warp of a lane quadrant) + the combined log-sum-exp per row
  return static_cast<size_t>(B) * attn_nt(N) * ATTN_PARTS * N * 2 * sizeof(float) + static_cast<size_t>(B) * N * sizeof(float) + 256;
}

int odvae_attn_scores_softmax(const void* q, const void* k, void* P, int B, int N, int C, int64_t ldq, int64_t ldk,
                              int64_t strideq, int64_t stridek, float scale, int dtype, void* workspace,
                              size_t ws_bytes, void* stream) {
  if ((dtype != ODVAE_BF16 && dtype != ODVAE_F16) || N % 256 != 0 || C % 64 != 0) return ODVAE_ERR_SHAPE;
  if (ws_bytes < odvae_attn_softmax_workspace_bytes(B, N) - 256) return ODVAE_ERR_WORKSPACE;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int om = dtype == ODVAE_F16 ? OUT_F16 : OUT_BF16;
  const int slots = attn_nt(N) * ATTN_PARTS;
  float* part = reinterpret_cast<float*>(workspace);
  float* lse2 = part + static_cast<size_t>(B) * slots * N * 2;
  for (int pass = 0; pass < 2; ++pass) {
    EngineDesc d;
    int st = describe_gemm(d, q, k, P, N, N, C, B, ldq, ldk, N, strideq, stridek, static_cast<int64_t>(N) * N, 0, 0, om,
                           scale, nullptr, nullptr);
    if (st != ODVAE_OK) return st;
    d.p.epi = pass == 0 ? EPI_ROWSTATS : EPI_SOFTMAX;
    d.p.aux_out = part;
    d.p.aux_f32 = lse2;
    st = run_engine(d, ODVAE_IMPL_TC, dtype, s);
    if (st != ODVAE_OK) return st;
    if (pass == 0) {
      st = launch_attn_stats_combine(part, lse2, B, slots, N, s);
      if (st != ODVAE_OK) return st;
    }
  }
  return ODVAE_OK;
}

int odvae_attn_dsoftmax(const void* dO, const void* v, const void* P, const float* delta, void* dS, int B, int N,
                        int C, int64_t ldo, int64_t ldv, int64_t strideo, int64_t stridev, float scale, int dtype,
                        void* stream) {
  if ((dtype != ODVAE_BF16 && dtype != ODVAE_F16) || N % 256 != 0 || C % 64 != 0) return ODVAE_ERR_SHAPE;
  EngineDesc d;
  const int om = dtype == ODVAE_F16 ? OUT_F16 : OUT_BF16;
  int st = describe_gemm(d, dO, v, dS, N, N, C, B, ldo, ldv, N, strideo, stridev, static_cast<int64_t>(N) * N, 0, 0, om,
                         scale, nullptr, nullptr);
  if (st != ODVAE_OK) return st;
  d.p.epi = EPI_DSOFTMAX;
  d.p.aux_f32 = delta;
  d.p.aux16 = P;
  return run_engine(d, ODVAE_IMPL_TC, dtype, static_cast<cudaStream_t>(stream));
}

int odvae_gemm_simt(const void* A, const void* B, void* D, int M, int N, int K, int batch, int64_t lda, int64_t ldb,
                    int64_t ldd, int64_t strideA, int64_t strideB, int64_t strideD, int a_mn, int b_mn, int in_dtype,
                    int out_dtype, float alpha, const float* bias, const void* residual, void* stream) {
  if (in_dtype == ODVAE_F32 && out_dtype == ODVAE_F32 && batch == 1 && !a_mn && residual == nullptr && M <= 64 &&
      K >= 256 && N >= 1) {
    // pose-MLP Linears (M = batch): K split over the warps of a CTA, one CTA per narrow column strip
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const float* a = reinterpret_cast<const float*>(A);
    const float* w = reinterpret_cast<const float*>(B);
    float* dst = reinterpret_cast<float*>(D);
    if (!b_mn) {
      skinny_gemm_nt_kernel<<<(N + 7) / 8, 256, 0, s>>>(a, w, bias, dst, M, N, K, lda, ldb, ldd, alpha);
      return check_launch("skinny_gemm_nt_kernel");
    }
    skinny_gemm_nn_kernel<<<dim3((N + 31) / 32, (M + 31) / 32), 256, 0, s>>>(a, w, bias, dst, M, N, K, lda, ldb, ldd, alpha);
    return check_launch("skinny_gemm_nn_kernel");
  }
  EngineDesc d;
  // out_dtype: ODVAE_F32 -> fp32 output; otherwise the activation dtype (= in_dtype)
  int st = describe_gemm(d, A, B, D, M, N, K, batch, lda, ldb, ldd, strideA, strideB, strideD, a_mn, b_mn,
                         (out_dtype == ODVAE_F32 && in_dtype != ODVAE_F32) ? OUT_F32 : OUT_BF16, alpha, bias, residual);
  if (st != ODVAE_OK) return st;
  return run_engine(d, ODVAE_IMPL_SIMT, in_dtype, static_cast<cudaStream_t>(stream));
}

}  // extern "C"
