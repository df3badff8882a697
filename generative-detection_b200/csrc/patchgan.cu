// PatchGAN discriminator kernels: taming.modules.discriminator.model.NLayerDiscriminator as called from
// src/modules/losses/contperceptual.py:280-301 (generator side) and :352-375 (discriminator side).
//
// Every 4x4 convolution (stride 2 or 1, padding 1) runs as  patches = im2col(activation)  ->  GEMM on the tcgen05
// engine (odvae_gemm_16; odvae_gemm_simt in the fp32 validation mode).  The kernels here are the memory-bound pieces
// around those GEMMs:
//   * im2col with the PREVIOUS layer's BatchNorm affine + LeakyReLU(0.2) applied on the fly (the normalised activation
//     is never written), and the first layer's `x * mask_2d_bbox` fused into the gather from the NCHW fp32 image;
//   * train-mode BatchNorm statistics (deterministic two-stage column sums; running stats + num_batches_tracked updated
//     like torch.nn.BatchNorm2d) and the BatchNorm backward (reduce + apply);
//   * col2im (gather form, fixed summation order) fused with the LeakyReLU derivative;
//   * weight packing OIHW -> [CoutP][(r, s, c)] and the inverse for weight gradients (fixed-order sum over GEMM splits).
// Everything is 16-byte vectorised over channels; no atomics.
#include "common.cuh"

namespace odvae {

template <typename T> struct PV;  // 8-element channel vector of T
template <> struct PV<float> {
  static __device__ __forceinline__ void load(const float* p, float (&v)[8]) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  static __device__ __forceinline__ void store(float* p, const float (&v)[8]) {
    reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
  }
};
template <> struct PV<bf16> {
  static __device__ __forceinline__ void load(const bf16* p, float (&v)[8]) {
    const uint4 r = __ldg(reinterpret_cast<const uint4*>(p));
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = unpack_bf16x2(w[i]);
      v[2 * i] = f.x;
      v[2 * i + 1] = f.y;
    }
  }
  static __device__ __forceinline__ void store(bf16* p, const float (&v)[8]) {
    *reinterpret_cast<uint4*>(p) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                              pack_bf16x2(v[6], v[7]));
  }
};
template <> struct PV<f16> {
  static __device__ __forceinline__ void load(const f16* p, float (&v)[8]) {
    const uint4 r = __ldg(reinterpret_cast<const uint4*>(p));
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = unpack_f16x2(w[i]);
      v[2 * i] = f.x;
      v[2 * i + 1] = f.y;
    }
  }
  static __device__ __forceinline__ void store(f16* p, const float (&v)[8]) {
    *reinterpret_cast<uint4*>(p) = make_uint4(pack_f16x2(v[0], v[1]), pack_f16x2(v[2], v[3]), pack_f16x2(v[4], v[5]),
                                              pack_f16x2(v[6], v[7]));
  }
};

__device__ __forceinline__ float lrelu(float u, float slope) { return u > 0.0f ? u : u * slope; }

// ---------------------------------------------------------------------------------------------
// im2col, NHWC source: dst[m][(r*4+s)*C + c] = act(src[b, oy*st-1+r, ox*st-1+s, c]), zero outside the image
// (Conv2d zero-pads the ACTIVATED tensor).  act(v) = lrelu(v*scale[c] + shift[c]) when `scale` is given, lrelu(v) when
// only `slope != 1`, identity otherwise.
// ---------------------------------------------------------------------------------------------
template <typename TI, typename TO>
__global__ void __launch_bounds__(256)
im2col_k4_nhwc(const TI* __restrict__ src, const float* __restrict__ scale, const float* __restrict__ shift, float slope,
               TO* __restrict__ dst, int B, int H, int W, int C, int st, int Ho, int Wo) {
  const int cv = C >> 3;
  const long long total = static_cast<long long>(B) * Ho * Wo * 16 * cv;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += gridDim.x * 256ll) {
    const int c8 = static_cast<int>(i % cv);
    const long long q = i / cv;
    const int t = static_cast<int>(q & 15);
    const long long m = q >> 4;
    const int ox = static_cast<int>(m % Wo);
    const long long mm = m / Wo;
    const int oy = static_cast<int>(mm % Ho);
    const int b = static_cast<int>(mm / Ho);
    const int iy = oy * st - 1 + (t >> 2), ix = ox * st - 1 + (t & 3);
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (iy >= 0 && iy < H && ix >= 0 && ix < W) {
      PV<TI>::load(src + ((static_cast<long long>(b) * H + iy) * W + ix) * C + c8 * 8, v);
      if (scale != nullptr) {
        float sc[8], sh[8];
        PV<float>::load(scale + c8 * 8, sc);
        PV<float>::load(shift + c8 * 8, sh);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = lrelu(fmaf(v[j], sc[j], sh[j]), slope);
      } else if (slope != 1.0f) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = lrelu(v[j], slope);
      }
    }
    PV<TO>::store(dst + (m * 16 + t) * C + c8 * 8, v);
  }
}

// first layer: NCHW fp32 image with C <= 4 channels (x optional mask[b, iy, ix]); K = 16*C padded to Kp (multiple of 8)
template <typename TO>
__global__ void __launch_bounds__(256)
im2col_k4_nchw(const float* __restrict__ src, const float* __restrict__ mask, TO* __restrict__ dst, int B, int H, int W,
               int C, int st, int Ho, int Wo, int Kp) {
  const int kv = Kp >> 3;
  const long long total = static_cast<long long>(B) * Ho * Wo * kv;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += gridDim.x * 256ll) {
    const int k8 = static_cast<int>(i % kv);
    const long long m = i / kv;
    const int ox = static_cast<int>(m % Wo);
    const long long mm = m / Wo;
    const int oy = static_cast<int>(mm % Ho);
    const int b = static_cast<int>(mm / Ho);
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int k = k8 * 8 + j;
      const int t = k / C, c = k - t * C;
      const int iy = oy * st - 1 + (t >> 2), ix = ox * st - 1 + (t & 3);
      float x = 0.0f;
      if (t < 16 && iy >= 0 && iy < H && ix >= 0 && ix < W) {
        x = __ldg(src + ((static_cast<long long>(b) * C + c) * H + iy) * W + ix);
        if (mask != nullptr) x *= __ldg(mask + (static_cast<long long>(b) * H + iy) * W + ix);
      }
      v[j] = x;
    }
    PV<TO>::store(dst + m * Kp + k8 * 8, v);
  }
}

// ---------------------------------------------------------------------------------------------
// col2im (gather form): dact[b, iy, ix, c] = sum over the taps (r, s) that reach (iy, ix) of dP[m(oy, ox)][(r*4+s)*C + c]
// in fixed (r, s) order, times the LeakyReLU derivative of this layer's pre-activation u = ypre*scale + shift.
// ---------------------------------------------------------------------------------------------
template <typename TP, typename TY, typename TO>
__global__ void __launch_bounds__(256)
col2im_k4_nhwc(const TP* __restrict__ dP, const TY* __restrict__ ypre, const float* __restrict__ scale,
               const float* __restrict__ shift, float slope, TO* __restrict__ dst, int B, int H, int W, int C, int st,
               int Ho, int Wo) {
  const int cv = C >> 3;
  const long long total = static_cast<long long>(B) * H * W * cv;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += gridDim.x * 256ll) {
    const int c8 = static_cast<int>(i % cv);
    const long long pix = i / cv;
    const int ix = static_cast<int>(pix % W);
    const long long pp = pix / W;
    const int iy = static_cast<int>(pp % H);
    const int b = static_cast<int>(pp / H);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int r = 0; r < 4; ++r) {
      const int ny = iy + 1 - r;
      if (ny < 0 || (ny % st) != 0) continue;
      const int oy = ny / st;
      if (oy >= Ho) continue;
      for (int s = 0; s < 4; ++s) {
        const int nx = ix + 1 - s;
        if (nx < 0 || (nx % st) != 0) continue;
        const int ox = nx / st;
        if (ox >= Wo) continue;
        float v[8];
        PV<TP>::load(dP + (((static_cast<long long>(b) * Ho + oy) * Wo + ox) * 16 + (r * 4 + s)) * C + c8 * 8, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += v[j];
      }
    }
    float u[8];
    PV<TY>::load(ypre + pix * C + c8 * 8, u);
    if (scale != nullptr) {
      float sc[8], sh[8];
      PV<float>::load(scale + c8 * 8, sc);
      PV<float>::load(shift + c8 * 8, sh);
#pragma unroll
      for (int j = 0; j < 8; ++j) u[j] = fmaf(u[j], sc[j], sh[j]);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = u[j] > 0.0f ? acc[j] : acc[j] * slope;
    PV<TO>::store(dst + pix * C + c8 * 8, acc);
  }
}

// first layer: gradient w.r.t. the NCHW fp32 image (x optional mask)
template <typename TP>
__global__ void __launch_bounds__(256)
col2im_k4_nchw(const TP* __restrict__ dP, const float* __restrict__ mask, float* __restrict__ dst, int B, int H, int W,
               int C, int st, int Ho, int Wo, int Kp) {
  const long long total = static_cast<long long>(B) * C * H * W;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += gridDim.x * 256ll) {
    const int ix = static_cast<int>(i % W);
    long long q = i / W;
    const int iy = static_cast<int>(q % H);
    q /= H;
    const int c = static_cast<int>(q % C);
    const int b = static_cast<int>(q / C);
    float acc = 0.0f;
    for (int r = 0; r < 4; ++r) {
      const int ny = iy + 1 - r;
      if (ny < 0 || (ny % st) != 0) continue;
      const int oy = ny / st;
      if (oy >= Ho) continue;
      for (int s = 0; s < 4; ++s) {
        const int nx = ix + 1 - s;
        if (nx < 0 || (nx % st) != 0) continue;
        const int ox = nx / st;
        if (ox >= Wo) continue;
        acc += to_f32<TP>(dP[((static_cast<long long>(b) * Ho + oy) * Wo + ox) * Kp + (r * 4 + s) * C + c]);
      }
    }
    if (mask != nullptr) acc *= __ldg(mask + (static_cast<long long>(b) * H + iy) * W + ix);
    dst[i] = acc;
  }
}

// ---------------------------------------------------------------------------------------------
// Column reductions over y[M][C] (row-major, C % 8 == 0): partial sums per block of rows, fixed order.
// MODE 0: (sum y, sum y^2)                                   -- BatchNorm statistics
// MODE 1: (sum dy', sum dy' * xhat), xhat = (y - mean) * rstd -- BatchNorm backward
// part layout: [block][2][C]
// ---------------------------------------------------------------------------------------------
static constexpr int BN_ROWS_PER_BLOCK = 512;

template <typename TY, typename TD, int MODE>
__global__ void __launch_bounds__(256)
bn_colreduce(const TY* __restrict__ y, const TD* __restrict__ dyp, const float* __restrict__ mean,
             const float* __restrict__ rstd, long long M, int C, float* __restrict__ part) {
  const int cv = C >> 3;                // 8-channel vectors per row (8 .. 64)
  const int lanes = 256 / cv;           // rows in flight per pass
  const int c8 = threadIdx.x % cv, lane = threadIdx.x / cv;
  const long long r0 = static_cast<long long>(blockIdx.x) * BN_ROWS_PER_BLOCK;
  const long long r1 = r0 + BN_ROWS_PER_BLOCK < M ? r0 + BN_ROWS_PER_BLOCK : M;
  float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, q[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float mu[8], rs[8];
  if (MODE == 1 && lane < lanes) {
    PV<float>::load(mean + c8 * 8, mu);
    PV<float>::load(rstd + c8 * 8, rs);
  }
  if (lane < lanes) {
    for (long long r = r0 + lane; r < r1; r += lanes) {
      float v[8];
      PV<TY>::load(y + r * C + c8 * 8, v);
      if (MODE == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          a[j] += v[j];
          q[j] = fmaf(v[j], v[j], q[j]);
        }
      } else {
        float d[8];
        PV<TD>::load(dyp + r * C + c8 * 8, d);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          a[j] += d[j];
          q[j] = fmaf(d[j], (v[j] - mu[j]) * rs[j], q[j]);
        }
      }
    }
  }
  // combine the row lanes in fixed order through shared memory
  __shared__ float sm[2][256][8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sm[0][threadIdx.x][j] = a[j];
    sm[1][threadIdx.x][j] = q[j];
  }
  __syncthreads();
  if (threadIdx.x < cv) {
    for (int w = 0; w < 2; ++w) {
      float tot[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int l = 0; l < lanes; ++l) {
#pragma unroll
        for (int j = 0; j < 8; ++j) tot[j] += sm[w][l * cv + threadIdx.x][j];
      }
      PV<float>::store(part + (static_cast<long long>(blockIdx.x) * 2 + w) * C + threadIdx.x * 8, tot);
    }
  }
}

// BatchNorm2d (training mode) finalize: mean / biased variance -> (scale, shift) of the affine the next im2col applies;
// running_mean / running_var (unbiased, momentum) and num_batches_tracked updated like torch.nn.BatchNorm2d.
__global__ void bn_stats_final(const float* __restrict__ part, int nblk, long long M, int C, const float* __restrict__ gamma,
                               const float* __restrict__ beta, float eps, float momentum, float* __restrict__ rmean,
                               float* __restrict__ rvar, long long* __restrict__ nbt, float* __restrict__ scale,
                               float* __restrict__ shift, float* __restrict__ mean_out, float* __restrict__ rstd_out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c == 0 && nbt != nullptr) nbt[0] += 1;
  if (c >= C) return;
  double s = 0.0, q = 0.0;
  for (int b = 0; b < nblk; ++b) {
    s += static_cast<double>(part[(static_cast<long long>(b) * 2) * C + c]);
    q += static_cast<double>(part[(static_cast<long long>(b) * 2 + 1) * C + c]);
  }
  const double mean = s / static_cast<double>(M);
  double var = q / static_cast<double>(M) - mean * mean;
  if (var < 0.0) var = 0.0;
  const float rstd = static_cast<float>(1.0 / sqrt(var + static_cast<double>(eps)));
  const float sc = gamma[c] * rstd;
  scale[c] = sc;
  shift[c] = beta[c] - static_cast<float>(mean) * sc;
  mean_out[c] = static_cast<float>(mean);
  rstd_out[c] = rstd;
  if (rmean != nullptr) {
    const double unb = M > 1 ? var * static_cast<double>(M) / static_cast<double>(M - 1) : var;
    rmean[c] = (1.0f - momentum) * rmean[c] + momentum * static_cast<float>(mean);
    rvar[c] = (1.0f - momentum) * rvar[c] + momentum * static_cast<float>(unb);
  }
}

// BatchNorm backward finalize: s1 = sum dy', s2 = sum dy' xhat  ->  dbeta, dgamma, and the per-channel coefficients
// coef[0][c] = gamma rstd, coef[1][c] = s1 / M, coef[2][c] = s2 / M of  dy = gamma rstd (dy' - s1/M - xhat s2/M)
__global__ void bn_bwd_final(const float* __restrict__ part, int nblk, long long M, int C, const float* __restrict__ gamma,
                             const float* __restrict__ rstd, float* __restrict__ dgamma, float* __restrict__ dbeta,
                             int accumulate, float* __restrict__ coef) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  double s1 = 0.0, s2 = 0.0;
  for (int b = 0; b < nblk; ++b) {
    s1 += static_cast<double>(part[(static_cast<long long>(b) * 2) * C + c]);
    s2 += static_cast<double>(part[(static_cast<long long>(b) * 2 + 1) * C + c]);
  }
  if (dgamma != nullptr) {
    dgamma[c] = (accumulate ? dgamma[c] : 0.0f) + static_cast<float>(s2);
    dbeta[c] = (accumulate ? dbeta[c] : 0.0f) + static_cast<float>(s1);
  }
  coef[c] = gamma[c] * rstd[c];
  coef[C + c] = static_cast<float>(s1 / static_cast<double>(M));
  coef[2 * C + c] = static_cast<float>(s2 / static_cast<double>(M));
}

template <typename TY, typename TD>
__global__ void __launch_bounds__(256)
bn_bwd_apply(const TY* __restrict__ y, TD* __restrict__ dyp, const float* __restrict__ mean, const float* __restrict__ rstd,
             const float* __restrict__ coef, long long M, int C) {
  const int cv = C >> 3;
  const long long total = M * cv;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < total; i += gridDim.x * 256ll) {
    const int c8 = static_cast<int>(i % cv);
    float v[8], d[8], mu[8], rs[8], k0[8], k1[8], k2[8];
    PV<TY>::load(y + i * 8, v);
    PV<TD>::load(dyp + i * 8, d);
    PV<float>::load(mean + c8 * 8, mu);
    PV<float>::load(rstd + c8 * 8, rs);
    PV<float>::load(coef + c8 * 8, k0);
    PV<float>::load(coef + C + c8 * 8, k1);
    PV<float>::load(coef + 2 * C + c8 * 8, k2);
#pragma unroll
    for (int j = 0; j < 8; ++j) d[j] = k0[j] * (d[j] - k1[j] - (v[j] - mu[j]) * rs[j] * k2[j]);
    PV<TD>::store(dyp + i * 8, d);
  }
}

// ---------------------------------------------------------------------------------------------
// Weights: OIHW fp32 [Cout][Cin][4][4] <-> packed [CoutP][Kp], k = (r*4+s)*Cin + c (rows / columns beyond are zero)
// ---------------------------------------------------------------------------------------------
template <typename TO>
__global__ void pack_k4_weight(const float* __restrict__ w, TO* __restrict__ out, int Cout, int Cin, int CoutP, int Kp) {
  const long long total = static_cast<long long>(CoutP) * Kp;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int k = static_cast<int>(i % Kp), o = static_cast<int>(i / Kp);
    const int t = k / Cin, c = k - t * Cin;
    float v = 0.0f;
    if (o < Cout && t < 16) v = __ldg(w + (static_cast<long long>(o) * Cin + c) * 16 + t);
    out[i] = from_f32<TO>(v);
  }
}

// dw[o][c][r][s] (+)= sum_split part[split][o][(r*4+s)*Cin + c]
__global__ void unpack_k4_wgrad(const float* __restrict__ part, int S, float* __restrict__ dw, int Cout, int Cin, int CoutP,
                                int Kp, int accumulate) {
  const long long total = static_cast<long long>(Cout) * Cin * 16;
  const long long slab = static_cast<long long>(CoutP) * Kp;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int t = static_cast<int>(i & 15);
    const long long oc = i >> 4;
    const int c = static_cast<int>(oc % Cin), o = static_cast<int>(oc / Cin);
    const long long src = static_cast<long long>(o) * Kp + t * Cin + c;
    float acc = 0.0f;
    for (int s = 0; s < S; ++s) acc += part[s * slab + src];
    dw[i] = accumulate ? dw[i] + acc : acc;
  }
}

static int grid_for(long long work_items, int per_block = 256) {
  long long blocks = (work_items + per_block - 1) / per_block;
  const long long cap = 16ll * num_sms();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

}  // namespace odvae

using namespace odvae;

#define PG_DISPATCH1(code, T, ...)                    \
  switch (code) {                                     \
    case ODVAE_F32: { typedef float T; __VA_ARGS__; } break; \
    case ODVAE_BF16: { typedef bf16 T; __VA_ARGS__; } break; \
    case ODVAE_F16: { typedef f16 T; __VA_ARGS__; } break;   \
    default: return ODVAE_ERR_SHAPE;                  \
  }

extern "C" {

int odvae_patchgan_im2col(const void* src, const float* mask, const float* scale, const float* shift, float slope,
                          void* dst, int B, int H, int W, int C, int stride, int Kp, int src_nchw, int src_dtype,
                          int dst_dtype, void* stream) {
  if (B <= 0 || H <= 0 || W <= 0 || C <= 0 || (stride != 1 && stride != 2) || Kp % 8 != 0) return ODVAE_ERR_SHAPE;
  const int Ho = (H + 2 - 4) / stride + 1, Wo = (W + 2 - 4) / stride + 1;
  if (Ho <= 0 || Wo <= 0) return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long M = static_cast<long long>(B) * Ho * Wo;
  if (src_nchw) {
    if (src_dtype != ODVAE_F32 || Kp < 16 * C || scale != nullptr) return ODVAE_ERR_SHAPE;
    PG_DISPATCH1(dst_dtype, TO, (im2col_k4_nchw<TO><<<grid_for(M * (Kp / 8)), 256, 0, st>>>(
                                    static_cast<const float*>(src), mask, static_cast<TO*>(dst), B, H, W, C, stride, Ho,
                                    Wo, Kp)))
    return check_launch("im2col_k4_nchw");
  }
  if (C % 8 != 0 || Kp != 16 * C || mask != nullptr) return ODVAE_ERR_SHAPE;
  const int g = grid_for(M * 16 * (C / 8));
  PG_DISPATCH1(src_dtype, TI,
               PG_DISPATCH1(dst_dtype, TO, (im2col_k4_nhwc<TI, TO><<<g, 256, 0, st>>>(
                                               static_cast<const TI*>(src), scale, shift, slope, static_cast<TO*>(dst), B,
                                               H, W, C, stride, Ho, Wo))))
  return check_launch("im2col_k4_nhwc");
}

int odvae_patchgan_col2im(const void* dP, const void* ypre, const float* scale, const float* shift, float slope,
                          const float* mask, void* dst, int B, int H, int W, int C, int stride, int Kp, int dst_nchw,
                          int dp_dtype, int y_dtype, int dst_dtype, void* stream) {
  if (B <= 0 || H <= 0 || W <= 0 || C <= 0 || (stride != 1 && stride != 2) || Kp % 8 != 0) return ODVAE_ERR_SHAPE;
  const int Ho = (H + 2 - 4) / stride + 1, Wo = (W + 2 - 4) / stride + 1;
  if (Ho <= 0 || Wo <= 0) return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dst_nchw) {
    if (dst_dtype != ODVAE_F32 || Kp < 16 * C) return ODVAE_ERR_SHAPE;
    const long long n = static_cast<long long>(B) * C * H * W;
    PG_DISPATCH1(dp_dtype, TP, (col2im_k4_nchw<TP><<<grid_for(n), 256, 0, st>>>(static_cast<const TP*>(dP), mask,
                                                                              static_cast<float*>(dst), B, H, W, C,
                                                                              stride, Ho, Wo, Kp)))
    return check_launch("col2im_k4_nchw");
  }
  if (C % 8 != 0 || Kp != 16 * C || ypre == nullptr || mask != nullptr) return ODVAE_ERR_SHAPE;
  const int g = grid_for(static_cast<long long>(B) * H * W * (C / 8));
  // the gradient tensors share one dtype (dP in, dact out); ypre is the forward format
  if (dp_dtype != dst_dtype) return ODVAE_ERR_SHAPE;
  PG_DISPATCH1(dp_dtype, TP,
               PG_DISPATCH1(y_dtype, TY, (col2im_k4_nhwc<TP, TY, TP><<<g, 256, 0, st>>>(
                                             static_cast<const TP*>(dP), static_cast<const TY*>(ypre), scale, shift, slope,
                                             static_cast<TP*>(dst), B, H, W, C, stride, Ho, Wo))))
  return check_launch("col2im_k4_nhwc");
}

size_t odvae_bn_workspace_bytes(int64_t M, int C) {
  const long long nblk = (M + BN_ROWS_PER_BLOCK - 1) / BN_ROWS_PER_BLOCK;
  return static_cast<size_t>(nblk) * 2 * C * sizeof(float) + static_cast<size_t>(3) * C * sizeof(float) + 256;
}

int odvae_bn_train_stats(const void* y, int64_t M, int C, int dtype, const float* gamma, const float* beta, float eps,
                         float momentum, float* running_mean, float* running_var, int64_t* num_batches_tracked,
                         float* scale, float* shift, float* mean, float* rstd, void* workspace, size_t ws_bytes,
                         void* stream) {
  if (M <= 0 || C <= 0 || C % 8 != 0 || C > 2048) return ODVAE_ERR_SHAPE;
  if (ws_bytes < odvae_bn_workspace_bytes(M, C) - 256) return ODVAE_ERR_WORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nblk = static_cast<int>((M + BN_ROWS_PER_BLOCK - 1) / BN_ROWS_PER_BLOCK);
  float* part = static_cast<float*>(workspace);
  PG_DISPATCH1(dtype, TY, (bn_colreduce<TY, TY, 0><<<nblk, 256, 0, st>>>(static_cast<const TY*>(y), nullptr, nullptr,
                                                                        nullptr, M, C, part)))
  int s = check_launch("bn_colreduce<0>");
  if (s != ODVAE_OK) return s;
  bn_stats_final<<<(C + 127) / 128, 128, 0, st>>>(part, nblk, M, C, gamma, beta, eps, momentum, running_mean, running_var,
                                                 reinterpret_cast<long long*>(num_batches_tracked), scale, shift, mean,
                                                 rstd);
  return check_launch("bn_stats_final");
}

int odvae_bn_train_bwd(const void* y, void* dyp, int64_t M, int C, int y_dtype, int g_dtype, const float* gamma,
                       const float* mean, const float* rstd, float* dgamma, float* dbeta, int accumulate,
                       void* workspace, size_t ws_bytes, void* stream) {
  if (M <= 0 || C <= 0 || C % 8 != 0 || C > 2048) return ODVAE_ERR_SHAPE;
  if (ws_bytes < odvae_bn_workspace_bytes(M, C) - 256) return ODVAE_ERR_WORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int nblk = static_cast<int>((M + BN_ROWS_PER_BLOCK - 1) / BN_ROWS_PER_BLOCK);
  float* part = static_cast<float*>(workspace);
  float* coef = part + static_cast<size_t>(nblk) * 2 * C;
  PG_DISPATCH1(y_dtype, TY,
               PG_DISPATCH1(g_dtype, TD, (bn_colreduce<TY, TD, 1><<<nblk, 256, 0, st>>>(
                                             static_cast<const TY*>(y), static_cast<const TD*>(dyp), mean, rstd, M, C,
                                             part))))
  int s = check_launch("bn_colreduce<1>");
  if (s != ODVAE_OK) return s;
  bn_bwd_final<<<(C + 127) / 128, 128, 0, st>>>(part, nblk, M, C, gamma, rstd, dgamma, dbeta, accumulate, coef);
  s = check_launch("bn_bwd_final");
  if (s != ODVAE_OK) return s;
  const int g = grid_for(M * (C / 8));
  PG_DISPATCH1(y_dtype, TY,
               PG_DISPATCH1(g_dtype, TD, (bn_bwd_apply<TY, TD><<<g, 256, 0, st>>>(static_cast<const TY*>(y),
                                                                                static_cast<TD*>(dyp), mean, rstd, coef,
                                                                                M, C))))
  return check_launch("bn_bwd_apply");
}

int odvae_pack_k4_weight(const float* w, void* out, int Cout, int Cin, int CoutP, int Kp, int dtype, void* stream) {
  if (Cout <= 0 || Cin <= 0 || CoutP < Cout || Kp < 16 * Cin) return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int g = grid_for(static_cast<long long>(CoutP) * Kp);
  PG_DISPATCH1(dtype, TO, (pack_k4_weight<TO><<<g, 256, 0, st>>>(w, static_cast<TO*>(out), Cout, Cin, CoutP, Kp)))
  return check_launch("pack_k4_weight");
}

int odvae_unpack_k4_wgrad(const float* part, int nsplit, float* dw, int Cout, int Cin, int CoutP, int Kp, int accumulate,
                          void* stream) {
  if (Cout <= 0 || Cin <= 0 || CoutP < Cout || Kp < 16 * Cin || nsplit <= 0) return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unpack_k4_wgrad<<<grid_for(static_cast<long long>(Cout) * Cin * 16), 256, 0, st>>>(part, nsplit, dw, Cout, Cin, CoutP, Kp,
                                                                                  accumulate);
  return check_launch("unpack_k4_wgrad");
}

}  // extern "C"
