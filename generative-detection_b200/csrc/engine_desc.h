// Problem description shared by the tcgen05 engine and the SIMT validation engine.
#pragma once
#include <stdint.h>

// A/B switch: the compile-time-specialised swapped-role conv epilogue (fp16, pixel stride 128; tc_engine.cuh `emit`), which the
// W = 128 halo tiles and the fused GroupNorm-backward sums build on
#ifndef TC_FAST_EPI
#define TC_FAST_EPI 1
#endif

namespace odvae {

enum { KIND_GEMM = 0, KIND_CONV = 1, KIND_WGRAD = 2 };
// instances of tc_engine_kernel<SPEC>: (kind, operand roles, epilogue) fixed at compile time
enum {
  TC_SPEC_CONV = 0,          // KIND_CONV, pixels on M
  TC_SPEC_CONV_SWAP = 1,     // KIND_CONV, channels on M / pixels on N (halo slabs, GroupNorm statistics, staged residual)
  TC_SPEC_WGRAD = 2,         // KIND_WGRAD
  TC_SPEC_GEMM = 3,          // KIND_GEMM, EPI_NONE
  TC_SPEC_GEMM_SOFTMAX = 4,  // KIND_GEMM, EPI_ROWSTATS / EPI_SOFTMAX
  TC_SPEC_GEMM_DSOFTMAX = 5, // KIND_GEMM, EPI_DSOFTMAX
  TC_SPEC_CONV_SWAP_GNB = 6, // TC_SPEC_CONV_SWAP whose fp16 epilogue also produces the GroupNorm-backward sums (gnb_*); no residual,
                             // no forward statistics, no staged / generic epilogue in this instance (register budget)
  TC_NUM_SPECS = 7
};
// Epilogue warps per CTA (after warp 0 = TMA producer, warp 1 = MMA issuer): two per TMEM lane quadrant, half the accumulator
// columns each. The softmax-backward instance (its epilogue streams a P tile from HBM and a dS tile back) runs 16: four per
// quadrant, 64 columns each -- 1.20 -> 1.09 ms per B = 64 block together with the stationary B panel below; the forward
// softmax instances measured 3 % slower with 16 (their epilogue is issue-bound: 10.7 k warp instructions per tile).
constexpr int tc_epi_warps(int spec) { return spec == TC_SPEC_GEMM_DSOFTMAX ? 16 : 8; }
constexpr int tc_threads(int spec) { return 64 + 32 * tc_epi_warps(spec); }
enum { OUT_BF16 = 0, OUT_F32 = 1, OUT_NCHW_F32 = 2, OUT_F16 = 3 };
enum { FMT_F16 = 0, FMT_BF16 = 1 };
// fused attention epilogues of KIND_GEMM (see tc_engine.cuh): row statistics / softmax / softmax backward
enum { EPI_NONE = 0, EPI_ROWSTATS = 1, EPI_SOFTMAX = 2, EPI_DSOFTMAX = 3 };  // tcgen05 kind::f16 operand format codes

struct TcParams {
  int kind, a_mn, b_mn, BN, stages;
  int a_fmt, b_fmt;  // FMT_* of the A / B operands
  int num_m_tiles, num_n_tiles, num_z;  // z: GEMM batch | CONV 1 | WGRAD ntaps*split_k
  int num_k_iters;
  // spatial operand (CONV: A, WGRAD: B): coords (c, x, [ph], y, b)
  int sp_rank, cchunks, ntaps;
  int tap_dx[16], tap_dy[16], tap_c[16], tap_ph[16];  // up to 16 taps (the 2x-upsample conv's single-launch dgrad)
  // CONV m-tile geometry (box bw x bh = 128 output pixels)
  int bw_log2, tiles_x, tiles_y;
  int swap;  // CONV with swapped roles: M = 128 output channels (weights), N = 256 output pixels (2 pixel boxes)
  // CONV halo mode (swapped 3x3 stride-1 at W = 256): the N tile is ONE image row; each input row (+1 pixel of halo on
  // both sides) is fetched once per 64-channel chunk into a shared-memory slab and the three horizontally shifted taps
  // read it through operand descriptors that start 0 / 1 / 2 pixel rows into the slab. `stages` then counts the
  // 16 KiB weight stages and `halo_slabs` the slab ring.
  int halo, halo_slabs;
  // halo == 2: the W = 128 form. The N tile is two image rows (y0, y0 + 1); a slab holds input rows (y0 + ky - 1, y0 + ky) of
  // one 64-channel chunk pixel-interleaved -- slab row 2 (x + 1) + r -- fetched by ONE TMA box (64 ch, 2 rows, 130 px) over a
  // (c, y, x, b) view of the NHWC tensor, so accumulator column n = 2 x + row reads slab row n + 2 kx: a horizontal tap is a
  // 256-byte shift of the operand descriptor. The epilogue maps column n back to (x = n / 2, row = n % 2). Operand traffic
  // per 128 x 256 tile: 288 KB of weights + 200 KB of slabs instead of + 576 KB of nine shifted gathers (these layers were
  // bound by the ~72 GB/s an SM gets from L2: 860 TF/s). halo2_ok: conv_common's candidate mark, resolved in launch_tc.
  int halo2_ok;
  // WGRAD k-chunk geometry (box kbw x kbh = 64 output pixels)
  int kbw_log2, kch_x, kch_y, split_k, kch_per_split;
  int wg_T, kch_total;  // taps accumulated per tile (1 or 3 = one filter row); pixel chunks over all splits
  // epilogue
  int out_mode, n_valid, m_valid, out_h, out_w;
  long long ldd, strideD, oB, oY, oX, o0;
  float alpha;
  int epi;               // EPI_*
  const float* aux_f32;  // SOFTMAX: row partials [z][n_tile][M][2] (max, sum) in the log2 domain; DSOFTMAX: delta [z][M]
  float* aux_out;        // ROWSTATS: row partials (same layout)
  const void* aux16;     // DSOFTMAX: P, same layout / format as D
  const float* bias;
  const void* residual;
  void* D;
  // WGRAD: bias gradient fused into the weight-gradient GEMM. The epilogue warps (idle during the K loop) sum the dy
  // tiles from shared memory; tile (tap group g, split s) covers the K chunks kit % ngroups == g of its split and
  // writes colsum[((g * split_k + s) * 2 + pixel half) * colsum_cols + channel] (fixed order, no atomics).
  float* colsum;
  int colsum_cols;
  // CONV (swapped roles): GroupNorm statistics of the OUTPUT fused into the epilogue. Every epilogue thread owns one
  // channel and 128 pixels of a pixel box: it writes (sum y, sum y^2) of its fp32 results to
  // gn_part[((image * boxes_per_image + box) * n_valid + channel) * 2] -- the layout gn_stats_finalize reduces.
  float* gn_part;
  // CONV (swapped roles): byte offset (from the 1024-aligned shared-memory base) of the 8 x 2 KiB per-warp staging
  // buffers the epilogue transposes 32 pixels x 32 channels through, so that the 16-bit results leave as 16-byte vector
  // stores (and the residual arrives as 16-byte vector loads); 0 = scalar 2-byte path.
  // CONV (swapped roles, fp16 fast epilogue; the dgrad that produces a GroupNorm+SiLU output gradient): the two sums of the
  // GroupNorm BACKWARD -- per (pixel box, channel): sum dy', sum dy' xhat with dy' = dy * silu'(gamma xhat + beta), dy = the value
  // this epilogue STORES, xhat from the GroupNorm input gnb_x (same NHWC layout as D) and gnb_stats [image][group](mean, rstd)
  // -- in packed half2 over pixel pairs like csrc/groupnorm.cu's gn_reduce_h2_kernel, written to
  // gnb_part[((image * boxes + box) * n_valid + channel) * 2]. The standalone reduce pass over x and dy disappears.
  const void* gnb_x;
  const float* gnb_stats;
  const float* gnb_gamma;
  const float* gnb_beta;
  float* gnb_part;
  int gnb_cpg;  // channels per group
  int epi_off;
  int res_prefetch;  // staged residual epilogue: L2-prefetch the tile's residual lines at tile start (ODVAE_RES_PREFETCH)
  // GEMM with a stationary B panel (K <= 256, both operands K-major): a CTA walks RUNS of bstat_mt consecutive m-tiles
  // that share (n_tile, z); the BN x K panel of B is loaded once per run into shared memory (num_k_iters chunks of
  // BN x 64) and only the 16 KiB A tiles stream through the stage ring, which cuts the L2 -> SM operand traffic of a
  // K = 256 score GEMM from 192 KiB to ~68 KiB per 128 x 256 tile (lts throughput 42 % -> 10 %). Used for dO v^T + softmax
  // backward, whose epilogue needs the L2 bandwidth for the P / dS tiles; the q k^T passes measured 4 % slower with it (the
  // panel reload at every run boundary stalls the MMAs) and keep the plain schedule. 0 = every tile loads both operands.
  int bstat_mt, bstat_runs;
  // GEMM with TWO m-tiles per CTA tile (256 x 256 outputs, K streamed): both 128-row A tiles of a stage are multiplied with
  // the SAME B tile (two MMAs per K step into the two 256-column accumulators), so B crosses the L2 -> SM fabric once per
  // 256 rows instead of once per 128. The attention GEMMs P v, P^T dO, dS k, dS^T q (M = K = 4096, N = 256) stream a 2 GB
  // P / dS matrix from HBM but moved 6.3 GB through L2 -> SM at the fabric's ~11 TB/s ceiling (B = 2/3 of it, re-fetched by
  // each of the 32 m-tiles of an image): 4.4 TB/s of DRAM. num_m_tiles counts PAIRS when set.
  int mpair;
};


// Strided view of an operand (elements, innermost dimension first, strides[0] == 1).
//   spatial operands: rank 4 (C, W, H, B) or rank 5 (2C, W/2, 2, H/2, B) for the stride-2 phase view
//   GEMM operands   : rank 3 (inner, outer, batch)
struct OperandView {
  const void* ptr;
  int rank;
  long long dims[5];
  long long strides[5];
};

struct EngineDesc {
  TcParams p;
  OperandView A, B;
  uint32_t boxA[5], boxB[5];  // TMA boxes (tcgen05 engine only)
  int tc_ok;                  // shapes satisfy the tcgen05 engine's tiling constraints
  // SIMT-only geometry (does not depend on TMA boxes)
  int conv_out_h, conv_out_w, conv_batch;  // output pixel grid of the gather (CONV: rows, WGRAD: K)
  int chan_per_tap;
  long long gemm_lda, gemm_ldb, gemm_k, gemm_n, gemm_strideA, gemm_strideB;
};

}  // namespace odvae
