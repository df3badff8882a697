// LPIPS perceptual distance: the pointwise / reduction kernels around the VGG16 conv stack.
// Reference call site: PoseLoss._get_rec_loss (src/modules/losses/contperceptual.py:141-145) -> taming.modules.losses.
// lpips.LPIPS (un-vendored; public definition restated in SURVEY appendix B): VGG16 features (3x3 convs + ReLU, 2x2 max
// pools) of input and reconstruction, channel-wise unit normalisation, squared difference, non-negative 1x1 "lin"
// weights, spatial mean, sum over the 5 levels. The convs run on the conv engine (api_conv.cu); here:
//   relu / relu_bwd, maxpool2x2 / _bwd on NHWC, and the per-level distance reduction + its gradient w.r.t. the
//   reconstruction's features. All reductions are fixed-order (deterministic). dtype: ODVAE_F32 or ODVAE_BF16.
#include "common.cuh"
#include "engine_host.h"

namespace odvae {

template <typename T> struct LpVec;
template <> struct LpVec<float> { static constexpr int V = 4; };
template <> struct LpVec<bf16> { static constexpr int V = 8; };

template <typename T>
__global__ void relu_kernel(T* __restrict__ x, long long nvec) {
  constexpr int V = LpVec<T>::V;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    T v[V];
    *reinterpret_cast<uint4*>(v) = reinterpret_cast<const uint4*>(x)[i];
#pragma unroll
    for (int j = 0; j < V; ++j) v[j] = from_f32<T>(fmaxf(to_f32<T>(v[j]), 0.0f));
    reinterpret_cast<uint4*>(x)[i] = *reinterpret_cast<const uint4*>(v);
  }
}

// dx = dy where y > 0 (y = the ReLU output), in place on dy when dx == dy
template <typename T>
__global__ void relu_bwd_kernel(const T* __restrict__ y, const T* dy, T* dx, long long nvec) {
  constexpr int V = LpVec<T>::V;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    T yv[V], dv[V];
    *reinterpret_cast<uint4*>(yv) = __ldg(reinterpret_cast<const uint4*>(y) + i);
    *reinterpret_cast<uint4*>(dv) = reinterpret_cast<const uint4*>(dy)[i];
#pragma unroll
    for (int j = 0; j < V; ++j)
      if (!(to_f32<T>(yv[j]) > 0.0f)) dv[j] = from_f32<T>(0.0f);
    reinterpret_cast<uint4*>(dx)[i] = *reinterpret_cast<const uint4*>(dv);
  }
}

// NHWC 2x2 / stride 2 max pool; one thread per (output pixel, channel vector)
template <typename T>
__global__ void maxpool2_kernel(const T* __restrict__ x, T* __restrict__ y, int B, int Ho, int Wo, int C) {
  constexpr int V = LpVec<T>::V;
  const int cv = C / V;
  const long long total = static_cast<long long>(B) * Ho * Wo * cv;
  const long long rowv = 2LL * Wo * cv;  // vectors per input row
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % cv);
    long long q = i / cv;
    const int xo = static_cast<int>(q % Wo);
    const long long r = q / Wo;  // b * Ho + yo
    const uint4* src = reinterpret_cast<const uint4*>(x) + (2 * r) * rowv + (2LL * xo) * cv + c;
    T a[4][V];
    *reinterpret_cast<uint4*>(a[0]) = __ldg(src);
    *reinterpret_cast<uint4*>(a[1]) = __ldg(src + cv);
    *reinterpret_cast<uint4*>(a[2]) = __ldg(src + rowv);
    *reinterpret_cast<uint4*>(a[3]) = __ldg(src + rowv + cv);
    T o[V];
#pragma unroll
    for (int j = 0; j < V; ++j)
      o[j] = from_f32<T>(fmaxf(fmaxf(to_f32<T>(a[0][j]), to_f32<T>(a[1][j])), fmaxf(to_f32<T>(a[2][j]), to_f32<T>(a[3][j]))));
    reinterpret_cast<uint4*>(y)[i] = *reinterpret_cast<const uint4*>(o);
  }
}

// dx[window] = dy at the first maximum of the window (scan order (0,0),(0,1),(1,0),(1,1), as torch), 0 elsewhere
template <typename T>
__global__ void maxpool2_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dy, T* __restrict__ dx, int B, int Ho,
                                    int Wo, int C) {
  constexpr int V = LpVec<T>::V;
  const int cv = C / V;
  const long long total = static_cast<long long>(B) * Ho * Wo * cv;
  const long long rowv = 2LL * Wo * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % cv);
    long long q = i / cv;
    const int xo = static_cast<int>(q % Wo);
    const long long r = q / Wo;
    const long long off = (2 * r) * rowv + (2LL * xo) * cv + c;
    const uint4* src = reinterpret_cast<const uint4*>(x) + off;
    T a[4][V], g[V], o[4][V];
    *reinterpret_cast<uint4*>(a[0]) = __ldg(src);
    *reinterpret_cast<uint4*>(a[1]) = __ldg(src + cv);
    *reinterpret_cast<uint4*>(a[2]) = __ldg(src + rowv);
    *reinterpret_cast<uint4*>(a[3]) = __ldg(src + rowv + cv);
    *reinterpret_cast<uint4*>(g) = __ldg(reinterpret_cast<const uint4*>(dy) + i);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      int best = 0;
      float bv = to_f32<T>(a[0][j]);
#pragma unroll
      for (int k = 1; k < 4; ++k) {
        const float v = to_f32<T>(a[k][j]);
        if (v > bv) {
          bv = v;
          best = k;
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) o[k][j] = (k == best) ? g[j] : from_f32<T>(0.0f);
    }
    uint4* dst = reinterpret_cast<uint4*>(dx) + off;
    dst[0] = *reinterpret_cast<const uint4*>(o[0]);
    dst[cv] = *reinterpret_cast<const uint4*>(o[1]);
    dst[rowv] = *reinterpret_cast<const uint4*>(o[2]);
    dst[rowv + cv] = *reinterpret_cast<const uint4*>(o[3]);
  }
}

// One LPIPS level. f0 / f1: [B][HW][C] features of input / reconstruction; w: [C] lin weights.
// d(p) = sum_c w_c (f0_c / (|f0| + eps) - f1_c / (|f1| + eps))^2 ; part[b][split] = sum over the split's pixels.
// One warp per pixel, lanes stride the channel vectors; CTA = 8 warps over a pixel range.
static constexpr float LPIPS_EPS = 1e-10f;
static constexpr int LP_SPLITS = 64;

template <typename T>
__global__ void __launch_bounds__(256)
lpips_level_fwd_kernel(const T* __restrict__ f0, const T* __restrict__ f1, const float* __restrict__ w,
                       float* __restrict__ part, int HW, int C) {
  constexpr int V = LpVec<T>::V;
  const int b = blockIdx.y, split = blockIdx.x, nsplit = gridDim.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int per = (HW + nsplit - 1) / nsplit;
  const int p0 = split * per, p1 = p0 + per < HW ? p0 + per : HW;
  const int cv = C / V;
  float acc = 0.0f;
  for (int p = p0 + warp; p < p1; p += 8) {
    const uint4* a = reinterpret_cast<const uint4*>(f0 + (static_cast<long long>(b) * HW + p) * C);
    const uint4* c = reinterpret_cast<const uint4*>(f1 + (static_cast<long long>(b) * HW + p) * C);
    float s0 = 0.0f, s1 = 0.0f;
    for (int i = lane; i < cv; i += 32) {
      T av[V], bv[V];
      *reinterpret_cast<uint4*>(av) = __ldg(a + i);
      *reinterpret_cast<uint4*>(bv) = __ldg(c + i);
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const float x0 = to_f32<T>(av[j]), x1 = to_f32<T>(bv[j]);
        s0 = fmaf(x0, x0, s0);
        s1 = fmaf(x1, x1, s1);
      }
    }
    s0 = warp_sum(s0);
    s1 = warp_sum(s1);
    const float r0 = 1.0f / (sqrtf(s0) + LPIPS_EPS), r1 = 1.0f / (sqrtf(s1) + LPIPS_EPS);
    float d = 0.0f;
    for (int i = lane; i < cv; i += 32) {  // second read hits L1/L2 (one pixel row of <= 2 KB)
      T av[V], bv[V];
      *reinterpret_cast<uint4*>(av) = __ldg(a + i);
      *reinterpret_cast<uint4*>(bv) = __ldg(c + i);
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const float e = to_f32<T>(av[j]) * r0 - to_f32<T>(bv[j]) * r1;
        d = fmaf(__ldg(w + i * V + j) * e, e, d);
      }
    }
    acc += warp_sum(d);
  }
  __shared__ float sm[8];
  if (lane == 0) sm[warp] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += sm[k];
    part[static_cast<long long>(b) * nsplit + split] = t;
  }
}

// out[b] (+)= (sum_split part[b][split]) / HW
__global__ void lpips_level_final_kernel(const float* __restrict__ part, float* __restrict__ out, int B, int nsplit,
                                         float inv_hw, int accumulate) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  float t = 0.0f;
  for (int s = 0; s < nsplit; ++s) t += part[static_cast<long long>(b) * nsplit + s];
  t *= inv_hw;
  out[b] = accumulate ? out[b] + t : t;
}

// d f1 = dout[b] / HW * d(d(p)) / d f1 ; with u = f1 / (n + eps), n = |f1|, e_c = -2 w_c (f0_c r0 - u_c):
//   df1_j = e_j / (n + eps) - (sum_c e_c f1_c) f1_j / (n (n + eps)^2);  written (accumulate: added) in T
template <typename T>
__global__ void __launch_bounds__(256)
lpips_level_bwd_kernel(const T* __restrict__ f0, const T* __restrict__ f1, const float* __restrict__ w,
                       const float* __restrict__ dout, T* __restrict__ df1, int HW, int C, int accumulate) {
  constexpr int V = LpVec<T>::V;
  const int b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cv = C / V;
  const float g = __ldg(dout + b) / static_cast<float>(HW);
  for (int p = blockIdx.x * 8 + warp; p < HW; p += gridDim.x * 8) {
    const long long base = (static_cast<long long>(b) * HW + p) * C;
    const uint4* a = reinterpret_cast<const uint4*>(f0 + base);
    const uint4* c = reinterpret_cast<const uint4*>(f1 + base);
    float s0 = 0.0f, s1 = 0.0f;
    for (int i = lane; i < cv; i += 32) {
      T av[V], bv[V];
      *reinterpret_cast<uint4*>(av) = __ldg(a + i);
      *reinterpret_cast<uint4*>(bv) = __ldg(c + i);
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const float x0 = to_f32<T>(av[j]), x1 = to_f32<T>(bv[j]);
        s0 = fmaf(x0, x0, s0);
        s1 = fmaf(x1, x1, s1);
      }
    }
    s0 = warp_sum(s0);
    s1 = warp_sum(s1);
    const float n1 = sqrtf(s1);
    const float r0 = 1.0f / (sqrtf(s0) + LPIPS_EPS), r1 = 1.0f / (n1 + LPIPS_EPS);
    float dot = 0.0f;  // sum_c e_c f1_c
    for (int i = lane; i < cv; i += 32) {
      T av[V], bv[V];
      *reinterpret_cast<uint4*>(av) = __ldg(a + i);
      *reinterpret_cast<uint4*>(bv) = __ldg(c + i);
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const float x1 = to_f32<T>(bv[j]);
        const float e = -2.0f * __ldg(w + i * V + j) * (to_f32<T>(av[j]) * r0 - x1 * r1);
        dot = fmaf(e, x1, dot);
      }
    }
    dot = warp_sum(dot);
    const float k2 = n1 > 0.0f ? dot * r1 * r1 / n1 : 0.0f;
    uint4* dst = reinterpret_cast<uint4*>(df1 + base);
    for (int i = lane; i < cv; i += 32) {
      T av[V], bv[V], ov[V];
      *reinterpret_cast<uint4*>(av) = __ldg(a + i);
      *reinterpret_cast<uint4*>(bv) = __ldg(c + i);
      if (accumulate) *reinterpret_cast<uint4*>(ov) = dst[i];
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const float x1 = to_f32<T>(bv[j]);
        const float e = -2.0f * __ldg(w + i * V + j) * (to_f32<T>(av[j]) * r0 - x1 * r1);
        float o = g * (e * r1 - k2 * x1);
        if (accumulate) o += to_f32<T>(ov[j]);
        ov[j] = from_f32<T>(o);
      }
      dst[i] = *reinterpret_cast<const uint4*>(ov);
    }
  }
}

static int lp_grid(long long work, int threads) {
  long long blocks = (work + threads - 1) / threads;
  const long long cap = 16LL * num_sms();
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : static_cast<int>(blocks);
}

}  // namespace odvae

using namespace odvae;

extern "C" {

int odvae_relu(void* x, int64_t n, int dtype, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == ODVAE_BF16) {
    if (n % 8) return ODVAE_ERR_SHAPE;
    relu_kernel<bf16><<<lp_grid(n / 8, 256), 256, 0, st>>>(reinterpret_cast<bf16*>(x), n / 8);
  } else if (dtype == ODVAE_F32) {
    if (n % 4) return ODVAE_ERR_SHAPE;
    relu_kernel<float><<<lp_grid(n / 4, 256), 256, 0, st>>>(reinterpret_cast<float*>(x), n / 4);
  } else {
    return ODVAE_ERR_SHAPE;
  }
  return check_launch("relu_kernel");
}

int odvae_relu_bwd(const void* y, const void* dy, void* dx, int64_t n, int dtype, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == ODVAE_BF16) {
    if (n % 8) return ODVAE_ERR_SHAPE;
    relu_bwd_kernel<bf16><<<lp_grid(n / 8, 256), 256, 0, st>>>(reinterpret_cast<const bf16*>(y), reinterpret_cast<const bf16*>(dy), reinterpret_cast<bf16*>(dx), n / 8);
  } else if (dtype == ODVAE_F32) {
    if (n % 4) return ODVAE_ERR_SHAPE;
    relu_bwd_kernel<float><<<lp_grid(n / 4, 256), 256, 0, st>>>(reinterpret_cast<const float*>(y), reinterpret_cast<const float*>(dy), reinterpret_cast<float*>(dx), n / 4);
  } else {
    return ODVAE_ERR_SHAPE;
  }
  return check_launch("relu_bwd_kernel");
}

int odvae_maxpool2x2(const void* x, void* y, int B, int H, int W, int C, int dtype, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if ((H | W) & 1) return ODVAE_ERR_SHAPE;
  const int V = dtype == ODVAE_BF16 ? 8 : 4;
  if (C % V || (dtype != ODVAE_BF16 && dtype != ODVAE_F32)) return ODVAE_ERR_SHAPE;
  const long long total = static_cast<long long>(B) * (H / 2) * (W / 2) * (C / V);
  if (dtype == ODVAE_BF16)
    maxpool2_kernel<bf16><<<lp_grid(total, 256), 256, 0, st>>>(reinterpret_cast<const bf16*>(x), reinterpret_cast<bf16*>(y), B, H / 2, W / 2, C);
  else
    maxpool2_kernel<float><<<lp_grid(total, 256), 256, 0, st>>>(reinterpret_cast<const float*>(x), reinterpret_cast<float*>(y), B, H / 2, W / 2, C);
  return check_launch("maxpool2_kernel");
}

int odvae_maxpool2x2_bwd(const void* x, const void* dy, void* dx, int B, int H, int W, int C, int dtype, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if ((H | W) & 1) return ODVAE_ERR_SHAPE;
  const int V = dtype == ODVAE_BF16 ? 8 : 4;
  if (C % V || (dtype != ODVAE_BF16 && dtype != ODVAE_F32)) return ODVAE_ERR_SHAPE;
  const long long total = static_cast<long long>(B) * (H / 2) * (W / 2) * (C / V);
  if (dtype == ODVAE_BF16)
    maxpool2_bwd_kernel<bf16><<<lp_grid(total, 256), 256, 0, st>>>(reinterpret_cast<const bf16*>(x), reinterpret_cast<const bf16*>(dy), reinterpret_cast<bf16*>(dx), B, H / 2, W / 2, C);
  else
    maxpool2_bwd_kernel<float><<<lp_grid(total, 256), 256, 0, st>>>(reinterpret_cast<const float*>(x), reinterpret_cast<const float*>(dy), reinterpret_cast<float*>(dx), B, H / 2, W / 2, C);
  return check_launch("maxpool2_bwd_kernel");
}

size_t odvae_lpips_level_workspace_bytes(int B) { return static_cast<size_t>(B) * LP_SPLITS * sizeof(float); }

int odvae_lpips_level_fwd(const void* f0, const void* f1, const float* w, float* out, int B, int HW, int C, int dtype,
                          int accumulate, void* workspace, size_t ws_bytes, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int V = dtype == ODVAE_BF16 ? 8 : 4;
  if (C % V || (dtype != ODVAE_BF16 && dtype != ODVAE_F32) || B <= 0 || HW <= 0) return ODVAE_ERR_SHAPE;
  if (ws_bytes < odvae_lpips_level_workspace_bytes(B)) return ODVAE_ERR_WORKSPACE;
  int nsplit = LP_SPLITS;
  if (nsplit > (HW + 7) / 8) nsplit = (HW + 7) / 8;
  float* part = reinterpret_cast<float*>(workspace);
  const dim3 grid(nsplit, B);
  if (dtype == ODVAE_BF16)
    lpips_level_fwd_kernel<bf16><<<grid, 256, 0, st>>>(reinterpret_cast<const bf16*>(f0), reinterpret_cast<const bf16*>(f1), w, part, HW, C);
  else
    lpips_level_fwd_kernel<float><<<grid, 256, 0, st>>>(reinterpret_cast<const float*>(f0), reinterpret_cast<const float*>(f1), w, part, HW, C);
  int s = check_launch("lpips_level_fwd_kernel");
  if (s != ODVAE_OK) return s;
  lpips_level_final_kernel<<<(B + 127) / 128, 128, 0, st>>>(part, out, B, nsplit, 1.0f / static_cast<float>(HW), accumulate);
  return check_launch("lpips_level_final_kernel");
}

int odvae_lpips_level_bwd(const void* f0, const void* f1, const float* w, const float* dout, void* df1, int B, int HW,
                          int C, int dtype, int accumulate, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int V = dtype == ODVAE_BF16 ? 8 : 4;
  if (C % V || (dtype != ODVAE_BF16 && dtype != ODVAE_F32) || B <= 0 || HW <= 0) return ODVAE_ERR_SHAPE;
  int gx = (HW + 7) / 8;
  const int cap = (16 * num_sms() + B - 1) / B;
  if (gx > cap) gx = cap;
  const dim3 grid(gx < 1 ? 1 : gx, B);
  if (dtype == ODVAE_BF16)
    lpips_level_bwd_kernel<bf16><<<grid, 256, 0, st>>>(reinterpret_cast<const bf16*>(f0), reinterpret_cast<const bf16*>(f1), w, dout, reinterpret_cast<bf16*>(df1), HW, C, accumulate);
  else
    lpips_level_bwd_kernel<float><<<grid, 256, 0, st>>>(reinterpret_cast<const float*>(f0), reinterpret_cast<const float*>(f1), w, dout, reinterpret_cast<float*>(df1), HW, C, accumulate);
  return check_launch("lpips_level_bwd_kernel");
}

}  // extern "C"
