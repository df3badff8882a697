// GroupNorm(32 groups, eps, affine) [+ SiLU] forward / backward on NHWC activations.
// Reference: Normalize (LD model.py:38-39) + nonlinearity x*sigmoid(x) (model.py:33-35).
// Memory-bound: every pass reads whole pixels (all channels, 16-byte vectors, fully coalesced);
// per-(image,channel) partial sums go through a small fp32 workspace and are combined in a fixed
// order (deterministic, no atomics).
#include "common.cuh"
#include "engine_host.h"
#include "tc_primitives.cuh"

#include <stdlib.h>
#include <type_traits>

namespace odvae {

template <typename T> struct VecN;
template <> struct VecN<float> { static constexpr int V = 4; };
template <> struct VecN<bf16> { static constexpr int V = 8; };
template <> struct VecN<f16> { static constexpr int V = 8; };

static constexpr int GN_THREADS = 256;
// tuning knobs: minimum resident CTAs per SM (caps registers at 85) and loads in flight. Measured alone on B200, 128 ch x
// 256^2 x 64, fp16 (tools/gn_bench.py + tools/gpu_gn_variants.sh, round 2): apply 0.576 -> 0.354 ms (6.06 TB/s = 92 % of the HBM
// copy peak) with 3 CTAs/SM; backward 1.283 -> 1.056 ms with 3 CTAs/SM on both passes and 2 loads in flight per tensor in
// the reduce pass; 4 CTAs/SM spills and is slower everywhere.
#ifndef GN_APPLY_MINB
#define GN_APPLY_MINB 3
#endif
#ifndef GN_BWD_MINB
#define GN_BWD_MINB 3
#endif
#ifndef GN_RED_MINB
#define GN_RED_MINB 3
#endif
#ifndef GN_BWD_U
#define GN_BWD_U 2
#endif
#ifndef GN_RED_U
#define GN_RED_U 2
#endif

// MODE 0: per-(b,c) sums of (x, x^2)          -> forward statistics
// MODE 1: per-(b,c) sums of (dy', dy' * xhat)  -> backward reductions, dy' = dy * silu'(u) when silu
// part layout: [B][nsplit][C][2]
template <typename T, typename TG, int MODE>
__global__ void __launch_bounds__(GN_THREADS, GN_RED_MINB)
gn_reduce_kernel(const T* __restrict__ x, const TG* __restrict__ dy, const float* __restrict__ stats,
                 const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ part, int HW,
                 int C, int G, int silu, int nsplit, int nitems) {
  constexpr int V = VecN<T>::V;
  const int cv = C / V;
  const int rows = GN_THREADS / cv;
  const int col = threadIdx.x % cv, row = threadIdx.x / cv;
  const int c0 = col * V;
  extern __shared__ float red[];
  // persistent grid (one resident wave): work items (image b, pixel range `split`) are taken grid-stride, so the
  // kernel has no partial last wave whatever the batch size
  for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
  const int b = item / nsplit, split = item - b * nsplit;
  const int per = (HW + nsplit - 1) / nsplit;
  const int p0 = split * per;
  const int p1 = p0 + per < HW ? p0 + per : HW;

  float s0[V], s1[V];
#pragma unroll
  for (int j = 0; j < V; ++j) s0[j] = s1[j] = 0.0f;
  float mu[V], rs[V], sc[V], sh[V];
  if (MODE == 1) {
    const int cpg = C / G;
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const int c = c0 + j;
      const int g = c / cpg;
      mu[j] = __ldg(stats + (static_cast<long long>(b) * G + g) * 2);
      rs[j] = __ldg(stats + (static_cast<long long>(b) * G + g) * 2 + 1);
      const float gm = __ldg(gamma + c);
      sc[j] = rs[j] * gm;
      sh[j] = __ldg(beta + c) - mu[j] * sc[j];
    }
  }
  if (row < rows) {
    const T* xb = x + static_cast<long long>(b) * HW * C + c0;
    const TG* dyb = MODE == 1 ? dy + static_cast<long long>(b) * HW * C + c0 : nullptr;
    constexpr int U = MODE == 0 ? 4 : GN_RED_U;  // forward statistics: one tensor, 4 loads in flight
    int p = p0 + row;
    for (; p + (U - 1) * rows < p1; p += U * rows) {
      T xv[U][V];
      TG dv[U][V];
#pragma unroll
      for (int u = 0; u < U; ++u)
        *reinterpret_cast<uint4*>(xv[u]) = __ldg(reinterpret_cast<const uint4*>(xb + static_cast<long long>(p + u * rows) * C));
      if (MODE == 1) {
#pragma unroll
        for (int u = 0; u < U; ++u)
          *reinterpret_cast<uint4*>(dv[u]) = __ldg(reinterpret_cast<const uint4*>(dyb + static_cast<long long>(p + u * rows) * C));
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const float v = to_f32<T>(xv[u][j]);
          if (MODE == 0) {
            s0[j] += v;
            s1[j] = fmaf(v, v, s1[j]);
          } else {
            float d = to_f32<TG>(dv[u][j]);
            if (silu) d *= silu_grad_t<sizeof(T) == 2>(fmaf(v, sc[j], sh[j]));
            s0[j] += d;
            s1[j] = fmaf(d, (v - mu[j]) * rs[j], s1[j]);
          }
        }
      }
    }
    for (; p < p1; p += rows) {
      T xv[V];
      *reinterpret_cast<uint4*>(xv) = __ldg(reinterpret_cast<const uint4*>(xb + static_cast<long long>(p) * C));
      if (MODE == 0) {
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const float v = to_f32<T>(xv[j]);
          s0[j] += v;
          s1[j] = fmaf(v, v, s1[j]);
        }
      } else {
        TG dv[V];
        *reinterpret_cast<uint4*>(dv) = __ldg(reinterpret_cast<const uint4*>(dyb + static_cast<long long>(p) * C));
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const float v = to_f32<T>(xv[j]);
          float d = to_f32<TG>(dv[j]);
          if (silu) d *= silu_grad_t<sizeof(T) == 2>(fmaf(v, sc[j], sh[j]));
          s0[j] += d;
          s1[j] = fmaf(d, (v - mu[j]) * rs[j], s1[j]);
        }
      }
    }
  }
  // block reduction over rows: red[2][rows][C]
  float* r0 = red;
  float* r1 = red + rows * C;
  if (row < rows) {
#pragma unroll
    for (int j = 0; j < V; ++j) {
      r0[row * C + c0 + j] = s0[j];
      r1[row * C + c0 + j] = s1[j];
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += GN_THREADS) {
    float a = 0.0f, bb = 0.0f;
    for (int r = 0; r < rows; ++r) {
      a += r0[r * C + c];
      bb += r1[r * C + c];
    }
    float* dst = part + ((static_cast<long long>(b) * nsplit + split) * C + c) * 2;
    dst[0] = a;
    dst[1] = bb;
  }
  __syncthreads();  // red[] is reused by the next item
  }
}

// forward finalize: grid (image, group block), one warp per group; stats[b][g] = (mean, rstd).
// The conv-epilogue statistics arrive as 512 partials per image and channel (one per 128-pixel box at 256x256): with one
// CTA per image this kernel took 28 us per launch (ncu launch list s10); spreading the groups over gridDim.y and reading
// float2 with 4 independent accumulator pairs per lane makes it latency- instead of serial-add-bound.
__global__ void gn_stats_finalize_kernel(const float* __restrict__ part, float* __restrict__ stats, int nsplit, int HW,
                                         int C, int G, float eps) {
  const int b = blockIdx.x;
  const int cpg = C / G;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const float2* part2 = reinterpret_cast<const float2*>(part);
  for (int g = blockIdx.y * nwarps + warp; g < G; g += gridDim.y * nwarps) {
    // lanes over (split, channel) pairs in a fixed order; 4 independent chains per lane, combined in a fixed order
    float s[4] = {0.0f, 0.0f, 0.0f, 0.0f}, ss[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    const int n = nsplit * cpg;
    int i = lane;
    for (; i + 96 < n; i += 128) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int k = i + 32 * u;
        const int sp = k / cpg, j = k - sp * cpg;
        const float2 v = __ldg(part2 + (static_cast<long long>(b) * nsplit + sp) * C + g * cpg + j);
        s[u] += v.x;
        ss[u] += v.y;
      }
    }
    for (; i < n; i += 32) {
      const int sp = i / cpg, j = i - sp * cpg;
      const float2 v = __ldg(part2 + (static_cast<long long>(b) * nsplit + sp) * C + g * cpg + j);
      s[0] += v.x;
      ss[0] += v.y;
    }
    float st = warp_sum((s[0] + s[1]) + (s[2] + s[3]));
    float sst = warp_sum((ss[0] + ss[1]) + (ss[2] + ss[3]));
    if (lane == 0) {
      const float nn = static_cast<float>(cpg) * HW;
      const float mean = st / nn;
      float var = sst / nn - mean * mean;
      var = var < 0.0f ? 0.0f : var;
      stats[(static_cast<long long>(b) * G + g) * 2] = mean;
      stats[(static_cast<long long>(b) * G + g) * 2 + 1] = rsqrtf(var + eps);
    }
  }
}

// y = [silu]((x - mean) * rstd * gamma + beta)
// The grid-stride (gridDim.x * GN_THREADS vectors) is a multiple of C/V, so every thread keeps ONE channel vector for
// the whole loop: scale/shift live in registers and the loop body is load -> fma (+SiLU) -> store.
// TO / TO2: formats of y and of the optional second copy y2 (mixed mode: fp16 for the forward conv, bf16 for wgrad);
// sizeof(T) == sizeof(TO) == sizeof(TO2) so all three share one vector index.
template <typename T, typename TO, typename TO2>
__global__ void __launch_bounds__(GN_THREADS, GN_APPLY_MINB)
gn_apply_kernel(const T* __restrict__ x, const float* __restrict__ stats, const float* __restrict__ gamma,
                const float* __restrict__ beta, TO* __restrict__ y, TO2* __restrict__ y2, int HW, int C, int G,
                int silu, int* __restrict__ overflow) {
  // fp16 inputs only: a value stored at +-65504 is a SATURATED store of its producer (common.cuh sat_f16) -- every tensor
  // of the residual stream passes through here, so this is where the forward's range overflow is detected
  constexpr bool kCheck = std::is_same<T, f16>::value;
  bool sat = false;
  constexpr int V = VecN<T>::V;
  const int b = blockIdx.y;
  const int cpg = C / G;
  const int cv = C / V;
  long long i = blockIdx.x * static_cast<long long>(GN_THREADS) + threadIdx.x;
  const int c0 = static_cast<int>(i % cv) * V;
  float scale[V], shift[V];
#pragma unroll
  for (int j = 0; j < V; ++j) {
    const int c = c0 + j;
    const int g = c / cpg;
    const float mean = __ldg(stats + (static_cast<long long>(b) * G + g) * 2);
    const float rstd = __ldg(stats + (static_cast<long long>(b) * G + g) * 2 + 1);
    scale[j] = rstd * __ldg(gamma + c);
    shift[j] = __ldg(beta + c) - mean * scale[j];
  }
  const long long nvec = static_cast<long long>(HW) * cv;
  const uint4* xb = reinterpret_cast<const uint4*>(x + static_cast<long long>(b) * HW * C);
  uint4* yb = reinterpret_cast<uint4*>(y + static_cast<long long>(b) * HW * C);
  uint4* y2b = y2 ? reinterpret_cast<uint4*>(y2 + static_cast<long long>(b) * HW * C) : nullptr;
  const long long stride = static_cast<long long>(gridDim.x) * GN_THREADS;
  constexpr int U = 4;
  for (; i + (U - 1) * stride < nvec; i += U * stride) {
    T xv[U][V];
#pragma unroll
    for (int u = 0; u < U; ++u) *reinterpret_cast<uint4*>(xv[u]) = __ldg(xb + i + u * stride);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      TO yv[V];
      TO2 y2v[V];
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const float xf = to_f32<T>(xv[u][j]);
        if (kCheck) sat |= fabsf(xf) >= 65504.0f;
        float v = fmaf(xf, scale[j], shift[j]);
        if (silu) v = silu_t<sizeof(T) == 2>(v);
        yv[j] = from_f32<TO>(v);
        y2v[j] = from_f32<TO2>(v);
      }
      yb[i + u * stride] = *reinterpret_cast<const uint4*>(yv);
      if (y2b) y2b[i + u * stride] = *reinterpret_cast<const uint4*>(y2v);
    }
  }
  for (; i < nvec; i += stride) {
    T xv[V];
    TO yv[V];
    TO2 y2v[V];
    *reinterpret_cast<uint4*>(xv) = __ldg(xb + i);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const float xf = to_f32<T>(xv[j]);
      if (kCheck) sat |= fabsf(xf) >= 65504.0f;
      float v = fmaf(xf, scale[j], shift[j]);
      if (silu) v = silu_t<sizeof(T) == 2>(v);
      yv[j] = from_f32<TO>(v);
      y2v[j] = from_f32<TO2>(v);
    }
    yb[i] = *reinterpret_cast<const uint4*>(yv);
    if (y2b) y2b[i] = *reinterpret_cast<const uint4*>(y2v);
  }
  if (kCheck && sat && overflow != nullptr) overflow[0] = 1;  // benign race: every writer stores the same value
}

// backward finalize A: one CTA per image. AB[b][c] = (sum dy', sum dy' xhat); coef[b][g] = (c2, c3)
//   dx = c1 * dy' + c2 * x + c3,  c1 = rstd*gamma_c, c2 = -rstd^2 s2/n, c3 = rstd (mean rstd s2 - s1)/n
__global__ void gn_bwd_finalize_a_kernel(const float* __restrict__ part, const float* __restrict__ stats,
                                         const float* __restrict__ gamma, float* __restrict__ AB,
                                         float* __restrict__ coef, int nsplit, int HW, int C, int G) {
  extern __shared__ float sm[];  // [C][2] gamma-weighted sums
  const int b = blockIdx.x;
  const int cpg = C / G;
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    float a = 0.0f, bb = 0.0f;
    for (int sp = 0; sp < nsplit; ++sp) {
      const float* src = part + ((static_cast<long long>(b) * nsplit + sp) * C + c) * 2;
      a += src[0];
      bb += src[1];
    }
    AB[(static_cast<long long>(b) * C + c) * 2] = a;
    AB[(static_cast<long long>(b) * C + c) * 2 + 1] = bb;
    const float gm = __ldg(gamma + c);
    sm[2 * c] = gm * a;
    sm[2 * c + 1] = gm * bb;
  }
  __syncthreads();
  for (int g = threadIdx.x; g < G; g += blockDim.x) {
    float s1 = 0.0f, s2 = 0.0f;
    for (int j = 0; j < cpg; ++j) {
      s1 += sm[2 * (g * cpg + j)];
      s2 += sm[2 * (g * cpg + j) + 1];
    }
    const float mean = stats[(static_cast<long long>(b) * G + g) * 2];
    const float rstd = stats[(static_cast<long long>(b) * G + g) * 2 + 1];
    const float inv_n = 1.0f / (static_cast<float>(cpg) * HW);
    coef[(static_cast<long long>(b) * G + g) * 2] = -rstd * rstd * s2 * inv_n;
    coef[(static_cast<long long>(b) * G + g) * 2 + 1] = rstd * (mean * rstd * s2 - s1) * inv_n;
  }
}

// backward finalize B: dgamma[c] = sum_b AB[b][c][1], dbeta[c] = sum_b AB[b][c][0]
__global__ void gn_bwd_finalize_b_kernel(const float* __restrict__ AB, float* __restrict__ dgamma,
                                         float* __restrict__ dbeta, int B, int C, int accumulate, float scale) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float a = 0.0f, bb = 0.0f;
  for (int b = 0; b < B; ++b) {
    a += AB[(static_cast<long long>(b) * C + c) * 2];
    bb += AB[(static_cast<long long>(b) * C + c) * 2 + 1];
  }
  a *= scale;
  bb *= scale;
  dbeta[c] = accumulate ? dbeta[c] + a : a;
  dgamma[c] = accumulate ? dgamma[c] + bb : bb;
}

// backward finalize from PER-BOX partials (the sums a dgrad epilogue produced, tc_engine.cuh gnb_*: part[(b * nbox + box) * C + c] =
// (sum dy', sum dy' xhat), 512 boxes per 256 x 256 image): grid (image, group blocks), one warp per group; lane = (channel of the
// group, box phase); four independent chains per lane, fixed-order xor-shuffle trees. Writes AB[b][c] and coef[b][g] like
// gn_bwd_finalize_a_kernel.
__global__ void gn_bwd_finalize_boxes_kernel(const float* __restrict__ part, const float* __restrict__ stats,
                                             const float* __restrict__ gamma, float* __restrict__ AB,
                                             float* __restrict__ coef, int nbox, int HW, int C, int G) {
  const int b = blockIdx.x;
  const int cpg = C / G;  // power of two <= 32 (host-checked)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int j = lane % cpg, s0 = lane / cpg, step = 32 / cpg;
  for (int g = blockIdx.y * nwarps + warp; g < G; g += gridDim.y * nwarps) {
    const float2* p2 = reinterpret_cast<const float2*>(part) + static_cast<long long>(b) * nbox * C + g * cpg + j;
    float a[4] = {0.0f, 0.0f, 0.0f, 0.0f}, bb[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    int sp = s0;
    for (; sp + 3 * step < nbox; sp += 4 * step) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float2 v = __ldg(p2 + static_cast<long long>(sp + u * step) * C);
        a[u] += v.x;
        bb[u] += v.y;
      }
    }
    for (; sp < nbox; sp += step) {
      const float2 v = __ldg(p2 + static_cast<long long>(sp) * C);
      a[0] += v.x;
      bb[0] += v.y;
    }
    float sa = (a[0] + a[1]) + (a[2] + a[3]), sb = (bb[0] + bb[1]) + (bb[2] + bb[3]);
    for (int o = cpg; o < 32; o <<= 1) {  // box phases of the same channel
      sa += __shfl_xor_sync(0xffffffffu, sa, o);
      sb += __shfl_xor_sync(0xffffffffu, sb, o);
    }
    if (lane < cpg) {
      AB[(static_cast<long long>(b) * C + g * cpg + j) * 2] = sa;
      AB[(static_cast<long long>(b) * C + g * cpg + j) * 2 + 1] = sb;
    }
    const float gm = __ldg(gamma + g * cpg + j);
    float ga = gm * sa, gb = gm * sb;
    for (int o = 1; o < cpg; o <<= 1) {  // channels of the group
      ga += __shfl_xor_sync(0xffffffffu, ga, o);
      gb += __shfl_xor_sync(0xffffffffu, gb, o);
    }
    if (lane == 0) {
      const float mean = stats[(static_cast<long long>(b) * G + g) * 2];
      const float rstd = stats[(static_cast<long long>(b) * G + g) * 2 + 1];
      const float inv_n = 1.0f / (static_cast<float>(cpg) * HW);
      coef[(static_cast<long long>(b) * G + g) * 2] = -rstd * rstd * gb * inv_n;
      coef[(static_cast<long long>(b) * G + g) * 2 + 1] = rstd * (mean * rstd * gb - ga) * inv_n;
    }
  }
}

// The same sums inside the apply kernels (one launch less per GroupNorm backward: 67 per train step): the first
// ceil(C / 32) CTAs of the grid (in linear order) each own 32 channels; warp w adds images w, w + 8, ... and warp 0 combines
// the eight partials in a fixed order before the CTA starts on its share of dx. AB == nullptr: nothing to do.
__device__ __forceinline__ void gn_bwd_param_grads(const float* __restrict__ AB, float* __restrict__ dgamma,
                                                   float* __restrict__ dbeta, int B, int C, int accumulate,
                                                   float scale) {
  if (AB == nullptr) return;
  const int nblk = gridDim.x * gridDim.y, blk = blockIdx.y * gridDim.x + blockIdx.x;
  const int ncb = (C + 31) / 32;
  if (blk >= ncb) return;  // block-uniform
  __shared__ float pg[8][32][2];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;  // GN_THREADS = 256 -> 8 warps
  for (int cb = blk; cb < ncb; cb += nblk) {
    const int c = cb * 32 + lane;
    float a = 0.0f, bb = 0.0f;
    if (c < C) {
      for (int b = w; b < B; b += 8) {
        const float2 v = __ldg(reinterpret_cast<const float2*>(AB) + static_cast<long long>(b) * C + c);
        a += v.x;
        bb += v.y;
      }
    }
    pg[w][lane][0] = a;
    pg[w][lane][1] = bb;
    __syncthreads();
    if (w == 0 && c < C) {
      a = 0.0f;
      bb = 0.0f;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        a += pg[k][lane][0];
        bb += pg[k][lane][1];
      }
      a *= scale;
      bb *= scale;
      dbeta[c] = accumulate ? dbeta[c] + a : a;
      dgamma[c] = accumulate ? dgamma[c] + bb : bb;
    }
    __syncthreads();
  }
}

// dx = c1 * dy' + c2 * x + c3 (+ dx_add); one channel vector per thread for the whole loop (see gn_apply_kernel)
template <typename T, typename TG>
__global__ void __launch_bounds__(GN_THREADS, GN_BWD_MINB)
gn_bwd_apply_kernel(const T* __restrict__ x, const TG* __restrict__ dy, const TG* __restrict__ dx_add,
                    const float* __restrict__ stats, const float* __restrict__ gamma, const float* __restrict__ beta,
                    const float* __restrict__ coef, TG* __restrict__ dx, int HW, int C, int G, int silu,
                    int* __restrict__ overflow, const float* __restrict__ AB, float* __restrict__ dgamma,
                    float* __restrict__ dbeta, int B, int accumulate, float pscale) {
  constexpr int V = VecN<T>::V;
  constexpr bool kCheck = std::is_same<TG, f16>::value;  // fp16 gradients: a saturated dy means the gradient scale is too large
  bool sat = false;
  gn_bwd_param_grads(AB, dgamma, dbeta, B, C, accumulate, pscale);
  const int b = blockIdx.y;
  const int cpg = C / G;
  const int cv = C / V;
  long long i = blockIdx.x * static_cast<long long>(GN_THREADS) + threadIdx.x;
  const int c0 = static_cast<int>(i % cv) * V;
  float scale[V], shift[V], c2[V], c3[V];
#pragma unroll
  for (int j = 0; j < V; ++j) {
    const int c = c0 + j;
    const int g = c / cpg;
    const float mean = __ldg(stats + (static_cast<long long>(b) * G + g) * 2);
    const float rstd = __ldg(stats + (static_cast<long long>(b) * G + g) * 2 + 1);
    scale[j] = rstd * __ldg(gamma + c);
    shift[j] = __ldg(beta + c) - mean * scale[j];
    c2[j] = __ldg(coef + (static_cast<long long>(b) * G + g) * 2);
    c3[j] = __ldg(coef + (static_cast<long long>(b) * G + g) * 2 + 1);
  }
  const long long nvec = static_cast<long long>(HW) * cv;
  const long long base = static_cast<long long>(b) * HW * C;
  const uint4* xb = reinterpret_cast<const uint4*>(x + base);
  const uint4* dyb = reinterpret_cast<const uint4*>(dy + base);
  const uint4* ab = dx_add ? reinterpret_cast<const uint4*>(dx_add + base) : nullptr;
  uint4* dxb = reinterpret_cast<uint4*>(dx + base);
  const long long stride = static_cast<long long>(gridDim.x) * GN_THREADS;
  constexpr int U = GN_BWD_U;
  for (; i < nvec; i += U * stride) {
    T xv[U][V];
    TG dv[U][V], av[U][V];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long k = i + u * stride;
      if (k < nvec) {
        *reinterpret_cast<uint4*>(xv[u]) = __ldg(xb + k);
        *reinterpret_cast<uint4*>(dv[u]) = __ldg(dyb + k);
        if (ab) *reinterpret_cast<uint4*>(av[u]) = __ldg(ab + k);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long k = i + u * stride;
      if (k < nvec) {
        TG ov[V];
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const float xf = to_f32<T>(xv[u][j]);
          float d = to_f32<TG>(dv[u][j]);
          if (kCheck) sat |= fabsf(d) >= 65504.0f;
          if (silu) d *= silu_grad_t<sizeof(T) == 2>(fmaf(xf, scale[j], shift[j]));
          float o = fmaf(scale[j], d, fmaf(c2[j], xf, c3[j]));
          if (ab) o += to_f32<TG>(av[u][j]);
          ov[j] = from_f32<TG>(o);
        }
        dxb[k] = *reinterpret_cast<const uint4*>(ov);
      }
    }
  }
  if (kCheck && sat && overflow != nullptr) overflow[1] = 1;
}

// ------------------------------------------------------------------------------------------------------------------
// Packed-half2 backward for fp16 activations + fp16 gradients (the timed mode). The fp32 kernels above spend ~14 (reduce)
// and ~18 (apply) instructions per element -- conversions, a scalar MUFU and scalar FMAs -- and run at 0.67-0.69 of the
// HBM copy peak: issue-bound, not memory-bound (round-2 launch list). Here every elementwise step works on TWO channels per
// instruction (HFMA2 / HMNMX2, tanh.approx.f16x2 = one MUFU per pair) and only the sums are fp32:
//   xc = x - fp16(mean)              exact where it matters (Sterbenz), so no cancellation is left for the constants
//   xh = xc * rstd + (fp16(mean) - mean) * rstd
//   uh = xh * gamma/2 + beta/2 = u / 2,  t = tanh(uh),  w = 1 - t^2,  q = t + uh w,  silu'(u) = 0.5 + 0.5 q
// Reduce: d' / 64 and d' xh / 64 are summed in half2 over 8 pixels (the 1/64 keeps eight near-saturated terms finite),
// then flushed into fp32 accumulators; the partial sums leave the kernel multiplied by 64 again.
// Apply: dx = c1 d' + (k2 xh + k3) (+ dx_add), k2 = -rstd S2/n, k3 = -rstd S1/n (the c2 x + c3 of the fp32 kernel
// rewritten on xh so that no large terms cancel in fp16), clamped to +-65504 like every fp16 store of this library.
// ------------------------------------------------------------------------------------------------------------------
#ifndef GN_H2
#define GN_H2 1
#endif
#ifndef GN_FUSE_PGRAD
#define GN_FUSE_PGRAD 1
#endif
#ifndef GN_H2_RED_U
#define GN_H2_RED_U 4
#endif
#ifndef GN_H2_RED_MINB
#define GN_H2_RED_MINB 2
#endif
#ifndef GN_H2_BWD_MINB
#define GN_H2_BWD_MINB 3
#endif
#ifndef GN_H2_BWD_U
#define GN_H2_BWD_U 2
#endif

__device__ __forceinline__ __half2 h2_tanh(__half2 v) {
  uint32_t r, a = *reinterpret_cast<uint32_t*>(&v);
  asm("tanh.approx.f16x2 %0, %1;" : "=r"(r) : "r"(a));
  return *reinterpret_cast<__half2*>(&r);
}
__device__ __forceinline__ __half2 h2_bits(uint32_t v) { return *reinterpret_cast<__half2*>(&v); }
__device__ __forceinline__ uint32_t h2_u32(__half2 v) { return *reinterpret_cast<uint32_t*>(&v); }

// per-thread constants of one channel pair
struct GnH2Const {
  __half2 muh, rs, k0, hg, hb;
};
__device__ __forceinline__ void gn_h2_consts(GnH2Const (&k)[4], const float* __restrict__ stats,
                                             const float* __restrict__ gamma, const float* __restrict__ beta, int b,
                                             int c0, int cpg, int G) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float muh[2], rs[2], k0[2], hg[2], hb[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int c = c0 + 2 * j + e;
      const int g = c / cpg;
      const float mean = __ldg(stats + (static_cast<long long>(b) * G + g) * 2);
      const float rstd = __ldg(stats + (static_cast<long long>(b) * G + g) * 2 + 1);
      muh[e] = __half2float(__float2half_rn(sat_f16(mean)));
      rs[e] = rstd;
      k0[e] = (muh[e] - mean) * rstd;
      hg[e] = 0.5f * __ldg(gamma + c);
      hb[e] = 0.5f * __ldg(beta + c);
    }
    k[j].muh = __floats2half2_rn(muh[0], muh[1]);
    k[j].rs = __floats2half2_rn(rs[0], rs[1]);
    k[j].k0 = __floats2half2_rn(k0[0], k0[1]);
    k[j].hg = __floats2half2_rn(hg[0], hg[1]);
    k[j].hb = __floats2half2_rn(hb[0], hb[1]);
  }
}
// xh and silu'(u) * (0.5 / gs) ... returned as q so that the caller picks the final scale: silu' = 0.5 + 0.5 q
template <bool SILU>
__device__ __forceinline__ void gn_h2_core(const GnH2Const& k, __half2 x, __half2& xh, __half2& q) {
  xh = __hfma2(__hsub2(x, k.muh), k.rs, k.k0);
  if (SILU) {
    const __half2 uh = __hfma2(xh, k.hg, k.hb);
    const __half2 t = h2_tanh(uh);
    const __half2 w = __hfma2(__hneg2(t), t, __float2half2_rn(1.0f));
    q = __hfma2(uh, w, t);
  }
}

// eight pixels of one channel vector: half2 partial sums, one fp32 flush. GUARD: only the first n (< 8) pixels exist
// (a missing pixel is loaded as dy = 0 and contributes nothing).
template <bool SILU, bool GUARD>
__device__ __forceinline__ void gn_h2_reduce8(const GnH2Const (&k)[4], const uint4* __restrict__ xp,
                                              const uint4* __restrict__ dp, int step, int n, __half2 inv,
                                              float (&s0)[8], float (&s1)[8]) {
  constexpr int U = GN_H2_RED_U, NH = 8 / U;
  __half2 a0[4], a1[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) a0[j] = a1[j] = __float2half2_rn(0.0f);
#pragma unroll
  for (int h = 0; h < NH; ++h) {
    uint4 xv[U], dv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int e = h * U + u;
      if (!GUARD || e < n) {
        xv[u] = __ldg(xp + e * step);
        dv[u] = __ldg(dp + e * step);
      } else {
        xv[u] = dv[u] = make_uint4(0u, 0u, 0u, 0u);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint32_t xw[4] = {xv[u].x, xv[u].y, xv[u].z, xv[u].w};
      const uint32_t dw[4] = {dv[u].x, dv[u].y, dv[u].z, dv[u].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        __half2 xh, q;
        gn_h2_core<SILU>(k[j], h2_bits(xw[j]), xh, q);
        const __half2 ds = SILU ? __hmul2(h2_bits(dw[j]), __hfma2(q, inv, inv)) : __hmul2(h2_bits(dw[j]), inv);
        a0[j] = __hadd2(a0[j], ds);
        a1[j] = __hfma2(ds, xh, a1[j]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float2 f0 = __half22float2(a0[j]), f1 = __half22float2(a1[j]);
    s0[2 * j] += f0.x;
    s0[2 * j + 1] += f0.y;
    s1[2 * j] += f1.x;
    s1[2 * j + 1] += f1.y;
  }
}

// part layout as gn_reduce_kernel MODE 1: [B][nsplit][C][2] = (sum dy', sum dy' xhat)
template <bool SILU>
__global__ void __launch_bounds__(GN_THREADS, GN_H2_RED_MINB)
gn_reduce_h2_kernel(const __half* __restrict__ x, const __half* __restrict__ dy, const float* __restrict__ stats,
                    const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ part, int HW,
                    int C, int G, int nsplit, int nitems) {
  constexpr int V = 8, U = GN_H2_RED_U, NH = 8 / U;  // U loads in flight per tensor, fp32 flush every U * NH = 8 pixels
  const int cv = C / V;
  const int rows = GN_THREADS / cv;
  const int col = threadIdx.x % cv, row = threadIdx.x / cv;
  const int c0 = col * V;
  extern __shared__ float red[];
  const __half2 inv = __float2half2_rn(SILU ? 1.0f / 128.0f : 1.0f / 64.0f);
  for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
    const int b = item / nsplit, split = item - b * nsplit;
    const int per = (HW + nsplit - 1) / nsplit;
    const int p0 = split * per;
    const int p1 = p0 + per < HW ? p0 + per : HW;
    GnH2Const k[4];
    gn_h2_consts(k, stats, gamma, beta, b, c0, C / G, G);
    float s0[V], s1[V];
#pragma unroll
    for (int j = 0; j < V; ++j) s0[j] = s1[j] = 0.0f;
    if (row < rows) {
      const uint4* xb = reinterpret_cast<const uint4*>(x + static_cast<long long>(b) * HW * C + c0);
      const uint4* dyb = reinterpret_cast<const uint4*>(dy + static_cast<long long>(b) * HW * C + c0);
      const int cvs = C / V;  // uint4 stride of one pixel
      const int step = rows * cvs;                        // uint4 distance between this thread's pixels
      int n = p1 > p0 + row ? (p1 - p0 - row + rows - 1) / rows : 0;  // pixels of this thread
      const uint4* xp = xb + (p0 + row) * cvs;
      const uint4* dp = dyb + (p0 + row) * cvs;
      for (; n >= NH * U; n -= NH * U) {  // whole groups of 8 pixels: no bounds checks
        gn_h2_reduce8<SILU, false>(k, xp, dp, step, NH * U, inv, s0, s1);
        xp += NH * U * step;
        dp += NH * U * step;
      }
      if (n > 0) gn_h2_reduce8<SILU, true>(k, xp, dp, step, n, inv, s0, s1);
    }
    float* r0 = red;
    float* r1 = red + rows * C;
    if (row < rows) {
#pragma unroll
      for (int j = 0; j < V; ++j) {
        r0[row * C + c0 + j] = s0[j];
        r1[row * C + c0 + j] = s1[j];
      }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += GN_THREADS) {
      float a = 0.0f, bb = 0.0f;
      for (int r = 0; r < rows; ++r) {
        a += r0[r * C + c];
        bb += r1[r * C + c];
      }
      float* dst = part + ((static_cast<long long>(b) * nsplit + split) * C + c) * 2;
      dst[0] = 64.0f * a;
      dst[1] = 64.0f * bb;
    }
    __syncthreads();  // red[] is reused by the next item
  }
}

// dx = c1 dy' + k2 xhat + k3 (+ dx_add), one channel vector per thread for the whole loop (see gn_apply_kernel)
template <bool SILU, bool ADD>
__global__ void __launch_bounds__(GN_THREADS, GN_H2_BWD_MINB)
gn_bwd_apply_h2_kernel(const __half* __restrict__ x, const __half* __restrict__ dy, const __half* __restrict__ dx_add,
                       const float* __restrict__ stats, const float* __restrict__ gamma, const float* __restrict__ beta,
                       const float* __restrict__ coef, __half* __restrict__ dx, int HW, int C, int G,
                       int* __restrict__ overflow, const float* __restrict__ AB, float* __restrict__ dgamma,
                       float* __restrict__ dbeta, int B, int accumulate, float pscale) {
  constexpr int V = 8, U = GN_H2_BWD_U;
  gn_bwd_param_grads(AB, dgamma, dbeta, B, C, accumulate, pscale);
  const int b = blockIdx.y;
  const int cpg = C / G;
  const int cv = C / V;
  int i = blockIdx.x * GN_THREADS + threadIdx.x;  // vector index inside the image: HW * C / 8 < 2^31
  const int c0 = (i % cv) * V;
  GnH2Const k[4];
  gn_h2_consts(k, stats, gamma, beta, b, c0, cpg, G);
  __half2 c1[4], k2[4], k3[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float a[2], b2[2], b3[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int c = c0 + 2 * j + e;
      const int g = c / cpg;
      const float mean = __ldg(stats + (static_cast<long long>(b) * G + g) * 2);
      const float rstd = __ldg(stats + (static_cast<long long>(b) * G + g) * 2 + 1);
      const float cc2 = __ldg(coef + (static_cast<long long>(b) * G + g) * 2);      // -rstd^2 S2 / n
      const float cc3 = __ldg(coef + (static_cast<long long>(b) * G + g) * 2 + 1);  // rstd (mean rstd S2 - S1) / n
      a[e] = rstd * __ldg(gamma + c);
      b2[e] = __fdividef(cc2, rstd);     // c2 x + c3 = (c2 / rstd) xhat + (c3 + c2 mean)
      b3[e] = fmaf(cc2, mean, cc3);
    }
    c1[j] = __floats2half2_rn(sat_f16(a[0]), sat_f16(a[1]));
    k2[j] = __floats2half2_rn(sat_f16(b2[0]), sat_f16(b2[1]));
    k3[j] = __floats2half2_rn(sat_f16(b3[0]), sat_f16(b3[1]));
  }
  const int nvec = HW * cv;
  const long long base = static_cast<long long>(b) * HW * C;
  const uint4* xb = reinterpret_cast<const uint4*>(x + base);
  const uint4* dyb = reinterpret_cast<const uint4*>(dy + base);
  const uint4* ab = ADD ? reinterpret_cast<const uint4*>(dx_add + base) : nullptr;
  uint4* dxb = reinterpret_cast<uint4*>(dx + base);
  const int stride = gridDim.x * GN_THREADS;
  const __half2 half = __float2half2_rn(0.5f), hmax = __float2half2_rn(65504.0f), hmin = __float2half2_rn(-65504.0f);
  __half2 amax = __float2half2_rn(0.0f);  // running max |dy|: a saturated dy means the gradient scale is too large
  for (; i < nvec; i += U * stride) {
    uint4 xv[U], dv[U], av[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int kk = i + u * stride;
      if (kk < nvec) {
        xv[u] = __ldg(xb + kk);
        dv[u] = __ldg(dyb + kk);
        if (ADD) av[u] = __ldg(ab + kk);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int kk = i + u * stride;
      if (kk < nvec) {
        const uint32_t xw[4] = {xv[u].x, xv[u].y, xv[u].z, xv[u].w};
        const uint32_t dw[4] = {dv[u].x, dv[u].y, dv[u].z, dv[u].w};
        const uint32_t aw[4] = {ADD ? av[u].x : 0u, ADD ? av[u].y : 0u, ADD ? av[u].z : 0u, ADD ? av[u].w : 0u};
        uint32_t ow[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          __half2 xh, q;
          gn_h2_core<SILU>(k[j], h2_bits(xw[j]), xh, q);
          __half2 d = h2_bits(dw[j]);
          amax = __hmax2(amax, __habs2(d));
          if (SILU) d = __hmul2(d, __hfma2(q, half, half));
          __half2 o = __hfma2(c1[j], d, __hfma2(k2[j], xh, k3[j]));
          if (ADD) o = __hadd2(o, h2_bits(aw[j]));
          ow[j] = h2_u32(__hmin2(__hmax2(o, hmin), hmax));
        }
        dxb[kk] = make_uint4(ow[0], ow[1], ow[2], ow[3]);
      }
    }
  }
  if (overflow != nullptr && (__hge(__low2half(amax), __low2half(hmax)) || __hge(__high2half(amax), __low2half(hmax))))
    overflow[1] = 1;  // benign race: every writer stores the same value
}

// ------------------------------------------------------------------------------------------------------------------
// Small feature maps (HW <= 1024: the 32x32 and 16x16 levels, 31 of the 67 GroupNorm sites): ONE launch per direction
// instead of three (statistics / finalize / apply, each 5-15 us of mostly launch latency for 8-30 us of memory time).
// CTA = (image, 64-channel slab = 128 bytes of every pixel; 64 channels are whole groups). Phase 1 reads the slab and
// reduces per channel in registers -> shared memory (fixed order), phase 2 re-reads it (<= 256 KiB per CTA: L1 / L2 hits)
// and writes the result. fp16 activations and fp16 gradients only (the timed mode), fp32 arithmetic.
// Backward: dgamma / dbeta need the sums over ALL images; the B CTAs of a channel slab take a ticket from a file-scope
// device counter after publishing their (sum dy', sum dy' xhat) rows, and the LAST one adds the rows in image order
// (deterministic: the order does not depend on who is last) and resets the counter. Backward launches of this kernel
// that overlap on the GPU (two streams) would share the counters: the callers issue GroupNorm backward on one stream.
// ------------------------------------------------------------------------------------------------------------------
#ifndef GN_SMALL
#define GN_SMALL 1
#endif
static constexpr int GN_SMALL_MAX_HW = 1024;
__device__ unsigned int gn_small_ticket[64];  // one per 64-channel slab (C <= 4096); zero at load, zero after every launch

struct GnSmallShared {
  float r0[32][65], r1[32][65];  // per pixel-phase partial sums of the slab's 64 channels (+1: no bank conflicts)
  float c0[64], c1[64];          // per channel
  float ga[32], gb[32];          // per group of the slab: (mean, rstd) forward / (c2, c3) backward
  int last;
};

// per-thread slab coordinates: 8 lanes cover the 128 bytes of a pixel, 32 pixel phases
#define GN_SMALL_COORDS                                                              \
  const int b = blockIdx.y, slab = blockIdx.x;                                       \
  const int sub = threadIdx.x & 7, prow = threadIdx.x >> 3;                          \
  const int c0 = slab * 64 + sub * 8;                                                \
  const int cpg = C / G, cvs = C / 8;                                                \
  const int gps = 64 / cpg; /* groups per slab */                                    \
  const long long base = (static_cast<long long>(b) * HW * C + c0) / 8;

__device__ __forceinline__ void gn_small_block_sums(GnSmallShared& sh, const float (&s0)[8], const float (&s1)[8], int sub,
                                                    int prow) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sh.r0[prow][sub * 8 + j] = s0[j];
    sh.r1[prow][sub * 8 + j] = s1[j];
  }
  __syncthreads();
  if (threadIdx.x < 64) {
    float a = 0.0f, bb = 0.0f;
#pragma unroll 8
    for (int r = 0; r < 32; ++r) {
      a += sh.r0[r][threadIdx.x];
      bb += sh.r1[r][threadIdx.x];
    }
    sh.c0[threadIdx.x] = a;
    sh.c1[threadIdx.x] = bb;
  }
  __syncthreads();
}

// KEEP: HW <= 256 -> a thread owns at most 8 pixels of its channel vector, loaded ONCE (all loads in flight together) and kept
// in registers across the reduction; otherwise phase 2 re-reads the slab (L1 / L2 hits), 4 loads in flight per tensor.
template <bool SILU, bool KEEP>
__global__ void __launch_bounds__(256, 2)
gn_small_fwd_kernel(const __half* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                    __half* __restrict__ y, float* __restrict__ stats, int HW, int C, int G, float eps,
                    int* __restrict__ overflow) {
  __shared__ GnSmallShared sh;
  GN_SMALL_COORDS
  const uint4* xb = reinterpret_cast<const uint4*>(x) + base;
  uint4* yb = reinterpret_cast<uint4*>(y) + base;
  constexpr int NP = KEEP ? 8 : 4;  // pixels per thread and round
  float s0[8], s1[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s0[j] = s1[j] = 0.0f;
  bool sat = false;
  uint4 v[NP];
  for (int p = prow; p < HW; p += 32 * NP) {
#pragma unroll
    for (int u = 0; u < NP; ++u) {
      v[u] = make_uint4(0u, 0u, 0u, 0u);
      if (p + 32 * u < HW) v[u] = __ldg(xb + static_cast<long long>(p + 32 * u) * cvs);
    }
#pragma unroll
    for (int u = 0; u < NP; ++u) {
      const __half* hv = reinterpret_cast<const __half*>(&v[u]);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float f = __half2float(hv[j]);
        sat |= fabsf(f) >= 65504.0f;
        s0[j] += f;
        s1[j] = fmaf(f, f, s1[j]);
      }
    }
    if (KEEP) break;  // HW <= 32 * 8: one round, v[] stays live
  }
  gn_small_block_sums(sh, s0, s1, sub, prow);
  if (threadIdx.x < gps) {
    float a = 0.0f, bb = 0.0f;
    for (int j = 0; j < cpg; ++j) {
      a += sh.c0[threadIdx.x * cpg + j];
      bb += sh.c1[threadIdx.x * cpg + j];
    }
    const float nn = static_cast<float>(cpg) * HW;
    const float mean = a / nn;
    float var = bb / nn - mean * mean;
    var = var < 0.0f ? 0.0f : var;
    const float rstd = rsqrtf(var + eps);
    sh.ga[threadIdx.x] = mean;
    sh.gb[threadIdx.x] = rstd;
    const int g = slab * gps + threadIdx.x;
    stats[(static_cast<long long>(b) * G + g) * 2] = mean;
    stats[(static_cast<long long>(b) * G + g) * 2 + 1] = rstd;
  }
  __syncthreads();
  float scale[8], shift[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int cl = sub * 8 + j;
    const float mean = sh.ga[cl / cpg], rstd = sh.gb[cl / cpg];
    scale[j] = rstd * __ldg(gamma + c0 + j);
    shift[j] = __ldg(beta + c0 + j) - mean * scale[j];
  }
  for (int p = prow; p < HW; p += 32 * NP) {
    if (!KEEP) {
#pragma unroll
      for (int u = 0; u < NP; ++u)
        if (p + 32 * u < HW) v[u] = __ldg(xb + static_cast<long long>(p + 32 * u) * cvs);
    }
#pragma unroll
    for (int u = 0; u < NP; ++u) {
      if (p + 32 * u < HW) {
        const __half* hv = reinterpret_cast<const __half*>(&v[u]);
        __half ov[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float r = fmaf(__half2float(hv[j]), scale[j], shift[j]);
          if (SILU) r = silu_t<true>(r);
          ov[j] = from_f32<f16>(r);
        }
        yb[static_cast<long long>(p + 32 * u) * cvs] = *reinterpret_cast<const uint4*>(ov);
      }
    }
    if (KEEP) break;
  }
  if (sat && overflow != nullptr) overflow[0] = 1;  // benign race: every writer stores the same value
}

template <bool SILU, bool ADD, bool KEEP>
__global__ void __launch_bounds__(256, 2)
gn_small_bwd_kernel(const __half* __restrict__ x, const __half* __restrict__ dy, const __half* __restrict__ dx_add,
                    const float* __restrict__ stats, const float* __restrict__ gamma, const float* __restrict__ beta,
                    __half* __restrict__ dx, float* __restrict__ AB, float* __restrict__ dgamma,
                    float* __restrict__ dbeta, int B, int HW, int C, int G, int accumulate, float pscale,
                    int* __restrict__ overflow) {
  __shared__ GnSmallShared sh;
  GN_SMALL_COORDS
  const uint4* xb = reinterpret_cast<const uint4*>(x) + base;
  const uint4* dyb = reinterpret_cast<const uint4*>(dy) + base;
  const uint4* ab = ADD ? reinterpret_cast<const uint4*>(dx_add) + base : nullptr;
  uint4* dxb = reinterpret_cast<uint4*>(dx) + base;
  constexpr int NP = KEEP ? 8 : 4;
  float mu[8], rs[8], sc[8], shf[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int g = (c0 + j) / cpg;
    mu[j] = __ldg(stats + (static_cast<long long>(b) * G + g) * 2);
    rs[j] = __ldg(stats + (static_cast<long long>(b) * G + g) * 2 + 1);
    sc[j] = rs[j] * __ldg(gamma + c0 + j);
    shf[j] = __ldg(beta + c0 + j) - mu[j] * sc[j];
  }
  float s0[8], s1[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s0[j] = s1[j] = 0.0f;
  bool sat = false;
  uint4 xv[NP], dv[NP];
  for (int p = prow; p < HW; p += 32 * NP) {
#pragma unroll
    for (int u = 0; u < NP; ++u) {
      xv[u] = dv[u] = make_uint4(0u, 0u, 0u, 0u);  // a missing pixel has dy = 0
      if (p + 32 * u < HW) {
        xv[u] = __ldg(xb + static_cast<long long>(p + 32 * u) * cvs);
        dv[u] = __ldg(dyb + static_cast<long long>(p + 32 * u) * cvs);
      }
    }
#pragma unroll
    for (int u = 0; u < NP; ++u) {
      const __half* hx = reinterpret_cast<const __half*>(&xv[u]);
      __half* hd = reinterpret_cast<__half*>(&dv[u]);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xf = __half2float(hx[j]);
        float d = __half2float(hd[j]);
        sat |= fabsf(d) >= 65504.0f;
        if (SILU) d *= silu_grad_t<true>(fmaf(xf, sc[j], shf[j]));
        s0[j] += d;
        s1[j] = fmaf(d, (xf - mu[j]) * rs[j], s1[j]);
      }
    }
    if (KEEP) break;
  }
  gn_small_block_sums(sh, s0, s1, sub, prow);
  if (threadIdx.x < 64) {  // publish this image's row of the slab for the dgamma / dbeta sums
    float2* dst = reinterpret_cast<float2*>(AB) + static_cast<long long>(b) * C + slab * 64 + threadIdx.x;
    *dst = make_float2(sh.c0[threadIdx.x], sh.c1[threadIdx.x]);
  }
  if (threadIdx.x < gps) {
    float a = 0.0f, bb = 0.0f;
    for (int j = 0; j < cpg; ++j) {
      const float gm = __ldg(gamma + slab * 64 + threadIdx.x * cpg + j);
      a += gm * sh.c0[threadIdx.x * cpg + j];
      bb += gm * sh.c1[threadIdx.x * cpg + j];
    }
    const int g = slab * gps + threadIdx.x;
    const float mean = __ldg(stats + (static_cast<long long>(b) * G + g) * 2);
    const float rstd = __ldg(stats + (static_cast<long long>(b) * G + g) * 2 + 1);
    const float inv_n = 1.0f / (static_cast<float>(cpg) * HW);
    sh.ga[threadIdx.x] = -rstd * rstd * bb * inv_n;                  // c2
    sh.gb[threadIdx.x] = rstd * (mean * rstd * bb - a) * inv_n;      // c3
  }
  __syncthreads();
  float c2[8], c3[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    c2[j] = sh.ga[(sub * 8 + j) / cpg];
    c3[j] = sh.gb[(sub * 8 + j) / cpg];
  }
  for (int p = prow; p < HW; p += 32 * NP) {
    uint4 av[NP];
#pragma unroll
    for (int u = 0; u < NP; ++u) {
      if (p + 32 * u < HW) {
        if (!KEEP) {
          xv[u] = __ldg(xb + static_cast<long long>(p + 32 * u) * cvs);
          dv[u] = __ldg(dyb + static_cast<long long>(p + 32 * u) * cvs);
        }
        if (ADD) av[u] = __ldg(ab + static_cast<long long>(p + 32 * u) * cvs);
      }
    }
#pragma unroll
    for (int u = 0; u < NP; ++u) {
      if (p + 32 * u < HW) {
        const __half* hx = reinterpret_cast<const __half*>(&xv[u]);
        const __half* hd = reinterpret_cast<const __half*>(&dv[u]);
        const __half* ha = reinterpret_cast<const __half*>(&av[u]);
        __half ov[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xf = __half2float(hx[j]);
          float d = __half2float(hd[j]);
          if (SILU) d *= silu_grad_t<true>(fmaf(xf, sc[j], shf[j]));
          float o = fmaf(sc[j], d, fmaf(c2[j], xf, c3[j]));
          if (ADD) o += __half2float(ha[j]);
          ov[j] = from_f32<f16>(o);
        }
        dxb[static_cast<long long>(p + 32 * u) * cvs] = *reinterpret_cast<const uint4*>(ov);
      }
    }
    if (KEEP) break;
  }
  if (sat && overflow != nullptr) overflow[1] = 1;
  // ---- dgamma / dbeta: the last CTA of this slab (over the B images) adds the published rows in image order ----
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int t = atomicAdd(&gn_small_ticket[slab], 1u);
    sh.last = (t == static_cast<unsigned int>(B) - 1u) ? 1 : 0;
  }
  __syncthreads();
  if (!sh.last) return;
  __threadfence();
  {
    const int ch = threadIdx.x & 63, part = threadIdx.x >> 6;  // 4 image slices per channel
    float a = 0.0f, bb = 0.0f;
    for (int i = part; i < B; i += 4) {
      const float2 v = __ldcg(reinterpret_cast<const float2*>(AB) + static_cast<long long>(i) * C + slab * 64 + ch);
      a += v.x;
      bb += v.y;
    }
    sh.r0[part][ch] = a;
    sh.r1[part][ch] = bb;
  }
  __syncthreads();
  if (threadIdx.x < 64) {
    const float a = ((sh.r0[0][threadIdx.x] + sh.r0[1][threadIdx.x]) + (sh.r0[2][threadIdx.x] + sh.r0[3][threadIdx.x])) * pscale;
    const float bb = ((sh.r1[0][threadIdx.x] + sh.r1[1][threadIdx.x]) + (sh.r1[2][threadIdx.x] + sh.r1[3][threadIdx.x])) * pscale;
    const int c = slab * 64 + threadIdx.x;
    dbeta[c] = accumulate ? dbeta[c] + a : a;
    dgamma[c] = accumulate ? dgamma[c] + bb : bb;
  }
  if (threadIdx.x == 0) gn_small_ticket[slab] = 0u;  // ready for the next launch (stream order)
}
#undef GN_SMALL_COORDS

static bool gn_small_ok(int HW, int C, int G) {
  return GN_SMALL && HW <= GN_SMALL_MAX_HW && C % 64 == 0 && C <= 4096 && C % G == 0 && 64 % (C / G) == 0 && (C / G) >= 2;
}

// Pixel-range splits per image for the reduce kernels. The reduce grid is ONE resident wave of `slots` CTAs taking
// (image, split) items grid-stride; B * nsplit is chosen near 4 items per CTA so that items / slots falls just below
// an integer (no CTA does a whole extra item while the others idle). slots <= 0: conservative 2 CTAs per SM.
static int gn_nsplit(int B, int HW, int slots) {
  if (slots <= 0) slots = 2 * num_sms();
  int maxs = HW / 32;
  if (maxs < 1) maxs = 1;
  if (maxs > 128) maxs = 128;
  int lo = (2 * slots + B - 1) / B, hi = (5 * slots) / B;
  if (lo < 1) lo = 1;
  if (hi < lo) hi = lo;
  if (lo > maxs) lo = maxs;
  if (hi > maxs) hi = maxs;
  int best = lo;
  double best_eff = -1.0;
  for (int ns = lo; ns <= hi; ++ns) {
    const double w = static_cast<double>(B) * ns / slots;
    const double waves = w <= 1.0 ? 1.0 : static_cast<double>(static_cast<long long>(w + 0.999999));
    const double eff = w <= 1.0 ? w : w / waves;
    if (eff >= best_eff) {
      best_eff = eff;
      best = ns;
    }
  }
  return best;
}
// resident CTAs of a reduce-kernel instantiation (cached per kernel pointer / smem size)
template <typename K>
static int gn_slots(K kernel, size_t smem) {
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, GN_THREADS, smem) != cudaSuccess || per_sm < 1) per_sm = 1;
  return per_sm * num_sms();
}
static int gn_ws_nsplit(int B, int HW) {  // upper bound of nsplit over the reduce modes / team path (workspace sizing)
  int maxs = HW / 32;
  if (maxs < 1) maxs = 1;
  return maxs > 320 ? 320 : maxs;
}
static bool gn_shape_ok(int C, int G, int V) {
  if (C % G != 0 || C % V != 0) return false;
  const int cv = C / V;
  return cv <= GN_THREADS && (GN_THREADS % cv) == 0 && C <= 2048;
}
static int gn_apply_blocks(int B, int HW, int C, int V) {
  long long nvec = static_cast<long long>(HW) * (C / V);
  long long blocks = (nvec + GN_THREADS * 4 - 1) / (GN_THREADS * 4);
  long long cap = (8ll * num_sms() + B - 1) / B;
  if (cap < 1) cap = 1;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

int launch_gn_stats_finalize(const float* part, float* stats, int B, int nsplit, int HW, int C, int G, float eps,
                             cudaStream_t st) {
  gn_stats_finalize_kernel<<<dim3(B, (G + 7) / 8), 256, 0, st>>>(part, stats, nsplit, HW, C, G, eps);
  return check_launch("gn_stats_finalize_kernel");
}

}  // namespace odvae

using namespace odvae;

extern "C" {


size_t odvae_groupnorm_workspace_bytes(int B, int HW, int C, int G) {
  const int ns = gn_ws_nsplit(B, HW);
  return (static_cast<size_t>(B) * ns * C * 2 + static_cast<size_t>(B) * C * 2 + static_cast<size_t>(B) * G * 2) * sizeof(float) + 256;
}

int odvae_groupnorm_stats(const void* x, float* stats, int B, int HW, int C, int G, float eps, int act_dtype,
                          void* workspace, size_t ws_bytes, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int V = act_dtype == ODVAE_F32 ? 4 : 8;
  if (!gn_shape_ok(C, G, V)) return ODVAE_ERR_SHAPE;
  if (ws_bytes < static_cast<size_t>(B) * gn_ws_nsplit(B, HW) * C * 2 * sizeof(float)) return ODVAE_ERR_WORKSPACE;
  float* part = reinterpret_cast<float*>(workspace);
  const int rows = GN_THREADS / (C / V);
  const size_t smem = static_cast<size_t>(2) * rows * C * sizeof(float);
  int ns;
#define GN_LAUNCH_REDUCE0(T)                                                                                          \
  {                                                                                                                   \
    const int slots = gn_slots(gn_reduce_kernel<T, bf16, 0>, smem);                                                   \
    ns = gn_nsplit(B, HW, slots);                                                                                     \
    const int items = B * ns;                                                                                         \
    gn_reduce_kernel<T, bf16, 0><<<items < slots ? items : slots, GN_THREADS, smem, st>>>(                             \
        reinterpret_cast<const T*>(x), nullptr, nullptr, nullptr, nullptr, part, HW, C, G, 0, ns, items);             \
  }
  if (act_dtype == ODVAE_BF16) GN_LAUNCH_REDUCE0(bf16)
  else if (act_dtype == ODVAE_F16) GN_LAUNCH_REDUCE0(f16)
  else {
    const int slots = gn_slots(gn_reduce_kernel<float, float, 0>, smem);
    ns = gn_nsplit(B, HW, slots);
    const int items = B * ns;
    gn_reduce_kernel<float, float, 0><<<items < slots ? items : slots, GN_THREADS, smem, st>>>(
        reinterpret_cast<const float*>(x), nullptr, nullptr, nullptr, nullptr, part, HW, C, G, 0, ns, items);
  }
#undef GN_LAUNCH_REDUCE0
  int s = check_launch("gn_reduce_kernel<0>");
  if (s != ODVAE_OK) return s;
  gn_stats_finalize_kernel<<<dim3(B, (G + 7) / 8), 256, 0, st>>>(part, stats, ns, HW, C, G, eps);
  return check_launch("gn_stats_finalize_kernel");
}

// (in, out, out2) format combinations the modes use: f32 -> f32; bf16 -> bf16; fp16 -> fp16 (+ bf16 copy for the
// weight-gradient GEMM); fp16 -> bf16 (attention blocks run in bf16 inside the mixed mode)
int odvae_groupnorm_apply(const void* x, const float* stats, const float* gamma, const float* beta, void* y, void* y2,
                          int B, int HW, int C, int G, int silu, int in_dtype, int out_dtype, int out2_dtype,
                          int32_t* overflow_flags, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int V = in_dtype == ODVAE_F32 ? 4 : 8;
  if (!gn_shape_ok(C, G, V)) return ODVAE_ERR_SHAPE;
  dim3 grid(gn_apply_blocks(B, HW, C, V), B);
  if (in_dtype == ODVAE_F32 && out_dtype == ODVAE_F32 && y2 == nullptr)
    gn_apply_kernel<float, float, float><<<grid, GN_THREADS, 0, st>>>(reinterpret_cast<const float*>(x), stats, gamma, beta, reinterpret_cast<float*>(y), nullptr, HW, C, G, silu, overflow_flags);
  else if (in_dtype == ODVAE_BF16 && out_dtype == ODVAE_BF16 && (y2 == nullptr || out2_dtype == ODVAE_BF16))
    gn_apply_kernel<bf16, bf16, bf16><<<grid, GN_THREADS, 0, st>>>(reinterpret_cast<const bf16*>(x), stats, gamma, beta, reinterpret_cast<bf16*>(y), reinterpret_cast<bf16*>(y2), HW, C, G, silu, overflow_flags);
  else if (in_dtype == ODVAE_F16 && out_dtype == ODVAE_F16 && (y2 == nullptr || out2_dtype == ODVAE_BF16))
    gn_apply_kernel<f16, f16, bf16><<<grid, GN_THREADS, 0, st>>>(reinterpret_cast<const f16*>(x), stats, gamma, beta, reinterpret_cast<f16*>(y), reinterpret_cast<bf16*>(y2), HW, C, G, silu, overflow_flags);
  else if (in_dtype == ODVAE_F16 && out_dtype == ODVAE_BF16 && y2 == nullptr)
    gn_apply_kernel<f16, bf16, bf16><<<grid, GN_THREADS, 0, st>>>(reinterpret_cast<const f16*>(x), stats, gamma, beta, reinterpret_cast<bf16*>(y), nullptr, HW, C, G, silu, overflow_flags);
  else
    return ODVAE_ERR_SHAPE;
  return check_launch("gn_apply_kernel");
}

// statistics + apply in ONE launch for small fp16 feature maps (gn_small_fwd_kernel); ODVAE_ERR_SHAPE when the shape is not
// served (the caller then uses odvae_groupnorm_stats + odvae_groupnorm_apply).
int odvae_groupnorm_fwd_small(const void* x, const float* gamma, const float* beta, void* y, float* stats, int B, int HW,
                              int C, int G, float eps, int silu, int act_dtype, int32_t* overflow_flags, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (act_dtype != ODVAE_F16 || !gn_small_ok(HW, C, G)) return ODVAE_ERR_SHAPE;
  dim3 grid(C / 64, B);
#define GN_LAUNCH_SMALL_FWD(S, K)                                                                                        \
  gn_small_fwd_kernel<S, K><<<grid, 256, 0, st>>>(reinterpret_cast<const __half*>(x), gamma, beta,                        \
                                                  reinterpret_cast<__half*>(y), stats, HW, C, G, eps, overflow_flags);
  const bool keep = HW <= 256;
  if (silu && keep) GN_LAUNCH_SMALL_FWD(true, true)
  else if (silu) GN_LAUNCH_SMALL_FWD(true, false)
  else if (keep) GN_LAUNCH_SMALL_FWD(false, true)
  else GN_LAUNCH_SMALL_FWD(false, false)
#undef GN_LAUNCH_SMALL_FWD
  return check_launch("gn_small_fwd_kernel");
}
// act_dtype: format of x (the saved forward input); grad_dtype: format of dy / dx_add / dx.
// Supported pairs: (f32, f32), (bf16, bf16), (fp16, bf16), (fp16, fp16). pgrad_scale multiplies dgamma / dbeta (the inverse
// of the static scale 16-bit activation gradients travel with in the all-fp16 mode; dx keeps the scale).
int odvae_groupnorm_bwd(const void* x, const float* stats, const float* gamma, const float* beta, const void* dy,
                        const void* dx_add, void* dx, float* dgamma, float* dbeta, int B, int HW, int C, int G,
                        int silu, int act_dtype, int grad_dtype, int accumulate, float pgrad_scale,
                        int32_t* overflow_flags, void* workspace, size_t ws_bytes, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int V = act_dtype == ODVAE_F32 ? 4 : 8;
  if (!gn_shape_ok(C, G, V)) return ODVAE_ERR_SHAPE;
  const int combo = (act_dtype == ODVAE_F32 && grad_dtype == ODVAE_F32)     ? 0
                    : (act_dtype == ODVAE_BF16 && grad_dtype == ODVAE_BF16) ? 1
                    : (act_dtype == ODVAE_F16 && grad_dtype == ODVAE_BF16)  ? 2
                    : (act_dtype == ODVAE_F16 && grad_dtype == ODVAE_F16)   ? 3
                                                                            : -1;
  if (combo < 0) return ODVAE_ERR_SHAPE;
  if (ws_bytes < odvae_groupnorm_workspace_bytes(B, HW, C, G) - 256) return ODVAE_ERR_WORKSPACE;
  float* part = reinterpret_cast<float*>(workspace);
  float* AB = part + static_cast<size_t>(B) * gn_ws_nsplit(B, HW) * C * 2;
  float* coef = AB + static_cast<size_t>(B) * C * 2;
  if (combo == 3 && gn_small_ok(HW, C, G)) {  // small feature maps: one launch (see gn_small_bwd_kernel)
    dim3 grid(C / 64, B);
#define GN_LAUNCH_SMALL_BWD(S, A)                                                                                        \
    if (HW <= 256) GN_LAUNCH_SMALL_BWD2(S, A, true) else GN_LAUNCH_SMALL_BWD2(S, A, false)
#define GN_LAUNCH_SMALL_BWD2(S, A, K)                                                                                    \
    gn_small_bwd_kernel<S, A, K><<<grid, 256, 0, st>>>(                                                                      \
        reinterpret_cast<const __half*>(x), reinterpret_cast<const __half*>(dy), reinterpret_cast<const __half*>(dx_add), \
        stats, gamma, beta, reinterpret_cast<__half*>(dx), AB, dgamma, dbeta, B, HW, C, G, accumulate, pgrad_scale,        \
        overflow_flags);
    if (silu && dx_add) GN_LAUNCH_SMALL_BWD(true, true)
    else if (silu) GN_LAUNCH_SMALL_BWD(true, false)
    else if (dx_add) GN_LAUNCH_SMALL_BWD(false, true)
    else GN_LAUNCH_SMALL_BWD(false, false)
#undef GN_LAUNCH_SMALL_BWD
#undef GN_LAUNCH_SMALL_BWD2
    return check_launch("gn_small_bwd_kernel");
  }
  const int rows = GN_THREADS / (C / V);
  const size_t smem = static_cast<size_t>(2) * rows * C * sizeof(float);
  int ns;
#define GN_LAUNCH_REDUCE1(T, TG)                                                                                      \
  {                                                                                                                   \
    const int slots = gn_slots(gn_reduce_kernel<T, TG, 1>, smem);                                                     \
    ns = gn_nsplit(B, HW, slots);                                                                                     \
    const int items = B * ns;                                                                                         \
    gn_reduce_kernel<T, TG, 1><<<items < slots ? items : slots, GN_THREADS, smem, st>>>(                               \
        reinterpret_cast<const T*>(x), reinterpret_cast<const TG*>(dy), stats, gamma, beta, part, HW, C, G, silu, ns, \
        items);                                                                                                       \
  }
#define GN_LAUNCH_REDUCE_H2(S)                                                                                        \
  {                                                                                                                   \
    const int slots = gn_slots(gn_reduce_h2_kernel<S>, smem);                                                         \
    ns = gn_nsplit(B, HW, slots);                                                                                     \
    const int items = B * ns;                                                                                         \
    gn_reduce_h2_kernel<S><<<items < slots ? items : slots, GN_THREADS, smem, st>>>(                                   \
        reinterpret_cast<const __half*>(x), reinterpret_cast<const __half*>(dy), stats, gamma, beta, part, HW, C, G,  \
        ns, items);                                                                                                   \
  }
  if (combo == 1) GN_LAUNCH_REDUCE1(bf16, bf16)
  else if (combo == 2) GN_LAUNCH_REDUCE1(f16, bf16)
  else if (combo == 3 && GN_H2) {
    if (silu) GN_LAUNCH_REDUCE_H2(true)
    else GN_LAUNCH_REDUCE_H2(false)
  }
  else if (combo == 3) GN_LAUNCH_REDUCE1(f16, f16)
  else GN_LAUNCH_REDUCE1(float, float)
#undef GN_LAUNCH_REDUCE_H2
#undef GN_LAUNCH_REDUCE1
  int s = check_launch("gn_reduce_kernel<1>");
  if (s != ODVAE_OK) return s;
  gn_bwd_finalize_a_kernel<<<B, 256, static_cast<size_t>(2) * C * sizeof(float), st>>>(part, stats, gamma, AB, coef, ns, HW, C, G);
  s = check_launch("gn_bwd_finalize_a_kernel");
  if (s != ODVAE_OK) return s;
#if GN_FUSE_PGRAD
  const float* ABp = AB;  // dgamma / dbeta are summed by the first CTAs of the apply kernel
#else
  const float* ABp = nullptr;
  gn_bwd_finalize_b_kernel<<<(C + 127) / 128, 128, 0, st>>>(AB, dgamma, dbeta, B, C, accumulate, pgrad_scale);
  s = check_launch("gn_bwd_finalize_b_kernel");
  if (s != ODVAE_OK) return s;
#endif
  dim3 grid2(gn_apply_blocks(B, HW, C, V), B);
#define GN_LAUNCH_APPLY(T, TG)                                                                                         \
  gn_bwd_apply_kernel<T, TG><<<grid2, GN_THREADS, 0, st>>>(reinterpret_cast<const T*>(x), reinterpret_cast<const TG*>(dy), \
                                                           reinterpret_cast<const TG*>(dx_add), stats, gamma, beta, coef,  \
                                                           reinterpret_cast<TG*>(dx), HW, C, G, silu, overflow_flags,  \
                                                           ABp, dgamma, dbeta, B, accumulate, pgrad_scale);
#define GN_LAUNCH_APPLY_H2(S, A)                                                                                       \
  gn_bwd_apply_h2_kernel<S, A><<<grid2, GN_THREADS, 0, st>>>(                                                          \
      reinterpret_cast<const __half*>(x), reinterpret_cast<const __half*>(dy), reinterpret_cast<const __half*>(dx_add), \
      stats, gamma, beta, coef, reinterpret_cast<__half*>(dx), HW, C, G, overflow_flags, ABp, dgamma, dbeta, B,       \
      accumulate, pgrad_scale);
  if (combo == 1) GN_LAUNCH_APPLY(bf16, bf16)
  else if (combo == 2) GN_LAUNCH_APPLY(f16, bf16)
  else if (combo == 3 && GN_H2) {
    if (silu && dx_add) GN_LAUNCH_APPLY_H2(true, true)
    else if (silu) GN_LAUNCH_APPLY_H2(true, false)
    else if (dx_add) GN_LAUNCH_APPLY_H2(false, true)
    else GN_LAUNCH_APPLY_H2(false, false)
  }
  else if (combo == 3) GN_LAUNCH_APPLY(f16, f16)
  else GN_LAUNCH_APPLY(float, float)
#undef GN_LAUNCH_APPLY_H2
#undef GN_LAUNCH_APPLY
  return check_launch("gn_bwd_apply_kernel");
}

// GroupNorm(+SiLU) backward whose two sums arrive as per-box partials from the dgrad that produced dy (odvae_conv2d_dgrad_gnbwd):
// finalize (coefficients, per-image channel sums) + ONE pass over x and dy (apply, dgamma / dbeta summed by its first CTAs).
// fp16 x and fp16 gradients, SiLU on, channels per group a power of two <= 32.
int odvae_groupnorm_bwd_from_partials(const void* x, const float* stats, const float* gamma, const float* beta,
                                      const void* dy, const void* dx_add, void* dx, float* dgamma, float* dbeta,
                                      const float* part, int nbox, int B, int HW, int C, int G, int accumulate,
                                      float pgrad_scale, int32_t* overflow_flags, void* workspace, size_t ws_bytes,
                                      void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (!gn_shape_ok(C, G, 8) || !GN_H2) return ODVAE_ERR_SHAPE;
  const int cpg = C / G;
  if (cpg > 32 || (cpg & (cpg - 1)) != 0 || nbox < 1) return ODVAE_ERR_SHAPE;
  if (ws_bytes < odvae_groupnorm_workspace_bytes(B, HW, C, G) - 256) return ODVAE_ERR_WORKSPACE;
  float* wsf = reinterpret_cast<float*>(workspace);
  float* AB = wsf + static_cast<size_t>(B) * gn_ws_nsplit(B, HW) * C * 2;
  float* coef = AB + static_cast<size_t>(B) * C * 2;
  gn_bwd_finalize_boxes_kernel<<<dim3(B, (G + 7) / 8), 256, 0, st>>>(part, stats, gamma, AB, coef, nbox, HW, C, G);
  int s = check_launch("gn_bwd_finalize_boxes_kernel");
  if (s != ODVAE_OK) return s;
  dim3 grid2(gn_apply_blocks(B, HW, C, 8), B);
  if (dx_add)
    gn_bwd_apply_h2_kernel<true, true><<<grid2, GN_THREADS, 0, st>>>(
        reinterpret_cast<const __half*>(x), reinterpret_cast<const __half*>(dy), reinterpret_cast<const __half*>(dx_add), stats,
        gamma, beta, coef, reinterpret_cast<__half*>(dx), HW, C, G, overflow_flags, AB, dgamma, dbeta, B, accumulate, pgrad_scale);
  else
    gn_bwd_apply_h2_kernel<true, false><<<grid2, GN_THREADS, 0, st>>>(
        reinterpret_cast<const __half*>(x), reinterpret_cast<const __half*>(dy), nullptr, stats, gamma, beta, coef,
        reinterpret_cast<__half*>(dx), HW, C, G, overflow_flags, AB, dgamma, dbeta, B, accumulate, pgrad_scale);
  return check_launch("gn_bwd_apply_h2_kernel");
}

}  // extern "C"
