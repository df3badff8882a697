// fp32 GEMMs with a tiny M (the pose MLPs: M = batch <= 64 rows against 500..4096-wide weight matrices,
// pose_decoder.py:82-97 / pose_encoder.py:82-131). The 64x64-tile SIMT engine launches only N/64 CTAs for these
// (8 CTAs for 4096 -> 500: 0.93 ms); here the K dimension is split over the warps of a CTA and the CTA owns a narrow
// strip of output columns, so the weight matrix streams through ~64 SMs. Deterministic: fixed-order reductions.
#pragma once
#include "common.cuh"

namespace odvae {

// D[m][n] = alpha * sum_k A[m][k] * W[n][k] (+ bias[n]);  A: [M][lda], W: [N][ldb] (K-major).  CTA = 8 columns.
__global__ void __launch_bounds__(256)
skinny_gemm_nt_kernel(const float* __restrict__ A, const float* __restrict__ W, const float* __restrict__ bias,
                      float* __restrict__ D, int M, int N, int K, long long lda, long long ldb, long long ldd,
                      float alpha) {
  __shared__ float red[8][8][8];  // [warp][m in group][n]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * 8;
  const int kper = (((K + 7) / 8) + 31) / 32 * 32;  // k range of a warp, a multiple of 32
  const int kbeg = warp * kper, kend = min(K, kbeg + kper);
  for (int m0 = 0; m0 < M; m0 += 8) {
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.0f;
    for (int k = kbeg + lane; k < kend; k += 32) {
      float w[8], a[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) w[j] = (n0 + j < N) ? __ldg(W + (n0 + j) * ldb + k) : 0.0f;
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = (m0 + i < M) ? __ldg(A + (m0 + i) * lda + k) : 0.0f;
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float v = warp_sum(acc[i][j]);
        if (lane == 0) red[warp][i][j] = v;
      }
    __syncthreads();
    if (threadIdx.x < 64) {
      const int i = threadIdx.x >> 3, j = threadIdx.x & 7;
      float v = 0.0f;
#pragma unroll
      for (int w8 = 0; w8 < 8; ++w8) v += red[w8][i][j];
      if (m0 + i < M && n0 + j < N) D[(m0 + i) * ldd + n0 + j] = v * alpha + (bias ? __ldg(bias + n0 + j) : 0.0f);
    }
    __syncthreads();
  }
}

// D[m][n] = alpha * sum_k A[m][k] * W[k][n] (+ bias[n]);  W: [K][ldb] (N contiguous).  CTA = 32 columns x 32 rows.
__global__ void __launch_bounds__(256)
skinny_gemm_nn_kernel(const float* __restrict__ A, const float* __restrict__ W, const float* __restrict__ bias,
                      float* __restrict__ D, int M, int N, int K, long long lda, long long ldb, long long ldd,
                      float alpha) {
  __shared__ __align__(16) float As[8][32][16];  // per warp: 32 rows x 16 k
  __shared__ float red[32][33];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x * 32 + lane;
  const int m0 = blockIdx.y * 32;
  const int kper = (((K + 7) / 8) + 15) / 16 * 16;
  const int kbeg = warp * kper, kend = min(K, kbeg + kper);
  float acc[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) acc[i] = 0.0f;
  for (int k0 = kbeg; k0 < kend; k0 += 16) {
    // stage A[m0 .. m0+31][k0 .. k0+15]: lane -> row `lane`, 16 consecutive k
#pragma unroll
    for (int kk = 0; kk < 16; ++kk)
      As[warp][lane][kk] = (m0 + lane < M && k0 + kk < kend) ? __ldg(A + (m0 + lane) * lda + k0 + kk) : 0.0f;
    __syncwarp();
#pragma unroll
    for (int k4 = 0; k4 < 16; k4 += 4) {
      float w[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) w[t] = (n < N && k0 + k4 + t < kend) ? __ldg(W + (k0 + k4 + t) * ldb + n) : 0.0f;
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const float4 a = *reinterpret_cast<const float4*>(&As[warp][i][k4]);
        acc[i] = fmaf(a.x, w[0], acc[i]);
        acc[i] = fmaf(a.y, w[1], acc[i]);
        acc[i] = fmaf(a.z, w[2], acc[i]);
        acc[i] = fmaf(a.w, w[3], acc[i]);
      }
    }
    __syncwarp();
  }
  // fixed-order reduction over the 8 warps through one 32x32 buffer
  for (int w8 = 0; w8 < 8; ++w8) {
    if (warp == w8) {
#pragma unroll
      for (int i = 0; i < 32; ++i) red[i][lane] = (w8 == 0 ? 0.0f : red[i][lane]) + acc[i];
    }
    __syncthreads();
  }
  if (warp == 0 && n < N) {
    const float b = bias ? __ldg(bias + n) : 0.0f;
    for (int i = 0; i < 32 && m0 + i < M; ++i) D[(m0 + i) * ldd + n] = red[i][lane] * alpha + b;
  }
}

}  // namespace odvae
