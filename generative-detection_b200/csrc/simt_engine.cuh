// CUDA-core (SIMT) engine with the same problem description as the tcgen05 engine.
// Purpose: (1) fp32 validation mode (<=1e-4 parity), (2) shapes the TMA tiling cannot express,
// (3) an independent check of the tensor-core path. 64x64x16 tiles, 4x4 outputs per thread, fp32 accumulate.
#pragma once
#include "common.cuh"
#include "engine_desc.h"

namespace odvae {

struct SimtParams {
  TcParams p;
  OperandView A, B;
  int out_h, out_w;       // pixel grid of the gather rows (CONV) / K (WGRAD)
  int chan_per_tap;       // channels of the gathered operand per tap (CONV K = ntaps*chan_per_tap)
  long long M, N, K;      // GEMM view of the problem (per z)
  long long lda, ldb, strideA, strideB;
  long long k_per_split;  // WGRAD
};

template <typename T>
__device__ __forceinline__ float spatial_fetch(const OperandView& v, int c, int x, int ph, int y, int b) {
  const T* ptr = reinterpret_cast<const T*>(v.ptr);
  if (v.rank == 5) {
    if (x < 0 || x >= v.dims[1] || y < 0 || y >= v.dims[3]) return 0.0f;
    return to_f32<T>(ptr[c + x * v.strides[1] + ph * v.strides[2] + y * v.strides[3] + b * v.strides[4]]);
  }
  if (x < 0 || x >= v.dims[1] || y < 0 || y >= v.dims[2]) return 0.0f;
  return to_f32<T>(ptr[c + x * v.strides[1] + y * v.strides[2] + b * v.strides[3]]);
}

template <typename T>
__device__ __forceinline__ float simt_load_a(const SimtParams& s, long long m, long long k, int z) {
  const TcParams& p = s.p;
  if (m >= s.M || k >= s.K) return 0.0f;
  if (p.kind == KIND_GEMM) {
    const T* a = reinterpret_cast<const T*>(s.A.ptr) + z * s.strideA;
    return to_f32<T>(p.a_mn ? a[k * s.lda + m] : a[m * s.lda + k]);
  } else if (p.kind == KIND_CONV) {
    const int hw = s.out_h * s.out_w;
    const int b = static_cast<int>(m / hw);
    const int r = static_cast<int>(m - static_cast<long long>(b) * hw);
    const int y = r / s.out_w, x = r - y * s.out_w;
    const int tap = static_cast<int>(k / s.chan_per_tap);
    const int c = static_cast<int>(k - static_cast<long long>(tap) * s.chan_per_tap);
    return spatial_fetch<T>(s.A, c + p.tap_c[tap], x + p.tap_dx[tap], p.tap_ph[tap], y + p.tap_dy[tap], b);
  } else {  // WGRAD: A(m = co, k = output pixel) = dy[pixel][co], through the operand view (strided for the upsample phases)
    const int hw = s.out_h * s.out_w;
    const int b = static_cast<int>(k / hw);
    const int r = static_cast<int>(k - static_cast<long long>(b) * hw);
    const int y = r / s.out_w, x = r - y * s.out_w;
    return spatial_fetch<T>(s.A, static_cast<int>(m), x, 0, y, b);
  }
}

template <typename T>
__device__ __forceinline__ float simt_load_b(const SimtParams& s, long long n, long long k, int z) {
  const TcParams& p = s.p;
  if (n >= s.N || k >= s.K) return 0.0f;
  if (p.kind == KIND_GEMM) {
    const T* bp = reinterpret_cast<const T*>(s.B.ptr) + z * s.strideB;
    return to_f32<T>(p.b_mn ? bp[k * s.ldb + n] : bp[n * s.ldb + k]);
  } else if (p.kind == KIND_CONV) {
    const T* bp = reinterpret_cast<const T*>(s.B.ptr);
    return to_f32<T>(bp[n * s.ldb + k]);
  } else {
    const int tap = z / p.split_k;
    const int hw = s.out_h * s.out_w;
    const int b = static_cast<int>(k / hw);
    const int r = static_cast<int>(k - static_cast<long long>(b) * hw);
    const int y = r / s.out_w, x = r - y * s.out_w;
    return spatial_fetch<T>(s.B, static_cast<int>(n) + p.tap_c[tap], x + p.tap_dx[tap], p.tap_ph[tap],
                            y + p.tap_dy[tap], b);
  }
}

// T: operand format; TO: format of a 16-bit / "activation dtype" output and of its residual
template <typename T, typename TO>
__global__ void __launch_bounds__(256) simt_engine_kernel(const __grid_constant__ SimtParams s) {
  const TcParams& p = s.p;
  __shared__ float As[16][64 + 4];
  __shared__ float Bs[16][64 + 4];
  const int z = blockIdx.z;
  const long long m0 = static_cast<long long>(blockIdx.y) * 64;
  const long long n0 = static_cast<long long>(blockIdx.x) * 64;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.0f;

  long long kbeg = 0, kend = s.K;
  if (p.kind == KIND_WGRAD) {
    const int split = z % p.split_k;
    kbeg = split * s.k_per_split;
    kend = kbeg + s.k_per_split < s.K ? kbeg + s.k_per_split : s.K;
  }
  for (long long k0 = kbeg; k0 < kend; k0 += 16) {
    // 64x16 elements per operand, 256 threads -> 4 each
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = threadIdx.x + i * 256;
      const int kk = e & 15, mm = e >> 4;
      const long long k = k0 + kk;
      As[kk][mm] = (k < kend) ? simt_load_a<T>(s, m0 + mm, k, z) : 0.0f;
      Bs[kk][mm] = (k < kend) ? simt_load_b<T>(s, n0 + mm, k, z) : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  // epilogue (same semantics as the tensor-core engine)
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + ty * 4 + i;
    if (m >= s.M) continue;
    long long row_off;
    int b = 0, yy = 0, xx = 0;
    if (p.kind == KIND_GEMM) {
      row_off = static_cast<long long>(z) * p.strideD + m * p.ldd;
    } else if (p.kind == KIND_CONV) {
      const int hw = s.out_h * s.out_w;
      b = static_cast<int>(m / hw);
      const int r = static_cast<int>(m - static_cast<long long>(b) * hw);
      yy = r / s.out_w;
      xx = r - yy * s.out_w;
      row_off = static_cast<long long>(b) * p.oB + static_cast<long long>(yy) * p.oY + static_cast<long long>(xx) * p.oX + p.o0;
    } else {
      row_off = (static_cast<long long>(z) * p.m_valid + m) * p.ldd;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long n = n0 + tx * 4 + j;
      if (n >= p.n_valid) continue;
      float v = acc[i][j] * p.alpha;
      if (p.bias != nullptr) v += __ldg(p.bias + n);
      if (p.out_mode == OUT_BF16 || p.out_mode == OUT_F16) {  // "activation dtype" output
        if (p.residual != nullptr) v += to_f32<TO>(reinterpret_cast<const TO*>(p.residual)[row_off + n]);
        reinterpret_cast<TO*>(p.D)[row_off + n] = from_f32<TO>(v);
      } else if (p.out_mode == OUT_F32) {
        if (p.residual != nullptr) v += reinterpret_cast<const float*>(p.residual)[row_off + n];
        reinterpret_cast<float*>(p.D)[row_off + n] = v;
      } else {
        const long long off = ((static_cast<long long>(b) * p.n_valid + n) * p.out_h + yy) * p.out_w + xx;
        reinterpret_cast<float*>(p.D)[off] = v;
      }
    }
  }
}

}  // namespace odvae
