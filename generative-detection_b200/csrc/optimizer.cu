// Fused Adam + global-norm gradient clipping over flat fp32 buffers.
// Reference: torch.optim.Adam(lr, betas=(0.5, 0.9)) at src/models/autoencoder.py:365-377 and
// gradient_clip_val: 1.0 at configs/autoencoder/pose/autoencoder_kl_16x16x16.yaml:140 (Lightning's
// clip_grad_norm_: g *= min(1, max_norm / (||g||_2 + 1e-6))).
// Memory-bound: 16-byte vectors, grid = multiple of the SM count, fixed-order reductions.
#include "common.cuh"

namespace odvae {

static constexpr int SUMSQ_BLOCKS_PER_SM = 4;

__global__ void __launch_bounds__(256) sumsq_stage1(const float* __restrict__ g, long long n, float* __restrict__ part) {
  float acc = 0.0f;
  const long long nv = n >> 2;
  const float4* g4 = reinterpret_cast<const float4*>(g);
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < nv; i += gridDim.x * 256ll) {
    const float4 v = __ldg(g4 + i);
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  for (long long i = (nv << 2) + blockIdx.x * 256ll + threadIdx.x; i < n; i += gridDim.x * 256ll) acc += g[i] * g[i];
  __shared__ float sm[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float r = threadIdx.x < 8 ? sm[threadIdx.x] : 0.0f;
    r = warp_sum(r);
    if (threadIdx.x == 0) part[blockIdx.x] = r;
  }
}
__global__ void sumsq_stage2(const float* __restrict__ part, int nblk, float* __restrict__ out, int accumulate) {
  // single warp, fixed order
  float acc = 0.0f;
  for (int i = threadIdx.x; i < nblk; i += 32) acc += part[i];
  acc = warp_sum(acc);
  if (threadIdx.x == 0) out[0] = accumulate ? out[0] + acc : acc;
}

__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
            long long n, float lr, float b1, float b2, float eps, float bc1, float bc2,
            const float* __restrict__ gnorm_sq, float max_norm) {
  float clip = 1.0f;
  if (gnorm_sq != nullptr && max_norm > 0.0f) {
    const float c = max_norm / (sqrtf(gnorm_sq[0]) + 1e-6f);
    clip = c < 1.0f ? c : 1.0f;
  }
  const float step = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  const long long nv = n >> 2;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < nv; i += gridDim.x * 256ll) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = __ldg(reinterpret_cast<const float4*>(g) + i);
    float4 mv = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float* pp = reinterpret_cast<float*>(&pv);
    const float* gg = reinterpret_cast<const float*>(&gv);
    float* mm = reinterpret_cast<float*>(&mv);
    float* vvp = reinterpret_cast<float*>(&vv);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gr = gg[j] * clip;
      mm[j] = b1 * mm[j] + (1.0f - b1) * gr;
      vvp[j] = b2 * vvp[j] + (1.0f - b2) * gr * gr;
      const float denom = sqrtf(vvp[j]) * inv_sqrt_bc2 + eps;
      pp[j] -= step * mm[j] / denom;
    }
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(m)[i] = mv;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  for (long long i = (nv << 2) + blockIdx.x * 256ll + threadIdx.x; i < n; i += gridDim.x * 256ll) {
    const float gr = g[i] * clip;
    const float mi = b1 * m[i] + (1.0f - b1) * gr;
    const float vi = b2 * v[i] + (1.0f - b2) * gr * gr;
    m[i] = mi;
    v[i] = vi;
    p[i] -= step * mi / (sqrtf(vi) * inv_sqrt_bc2 + eps);
  }
}


// ---------------------------------------------------------------------------------------------
// Multi-tensor variants: one launch over a device-resident chunk table (no flattening copies).
// table row = {p, g, m, v (device addresses), count}; one CTA per row; rows are <= 64 Ki elements.
// ---------------------------------------------------------------------------------------------
struct MtRow {
  unsigned long long p, g, m, v;
  long long n;
};

__global__ void __launch_bounds__(256) mt_sumsq_stage1(const MtRow* __restrict__ rows, float* __restrict__ part) {
  const MtRow r = rows[blockIdx.x];
  const float* g = reinterpret_cast<const float*>(r.g);
  float acc = 0.0f;
  if ((r.g & 15ull) == 0) {
    const long long nv = r.n >> 2;
    const float4* g4 = reinterpret_cast<const float4*>(g);
    for (long long i = threadIdx.x; i < nv; i += 256) {
      const float4 v = __ldg(g4 + i);
      acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    for (long long i = (nv << 2) + threadIdx.x; i < r.n; i += 256) acc += g[i] * g[i];
  } else {
    for (long long i = threadIdx.x; i < r.n; i += 256) acc += g[i] * g[i];
  }
  __shared__ float sm[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < 8 ? sm[threadIdx.x] : 0.0f;
    t = warp_sum(t);
    if (threadIdx.x == 0) part[blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(256)
mt_adam_kernel(const MtRow* __restrict__ rows, float lr, float b1, float b2, float eps, float bc1, float bc2,
               const float* __restrict__ gnorm_sq, float max_norm, float grad_scale) {
  const MtRow r = rows[blockIdx.x];
  float* p = reinterpret_cast<float*>(r.p);
  const float* g = reinterpret_cast<const float*>(r.g);
  float* m = reinterpret_cast<float*>(r.m);
  float* v = reinterpret_cast<float*>(r.v);
  float clip = grad_scale;
  if (gnorm_sq != nullptr && max_norm > 0.0f) {
    // the norm was taken over the unscaled gradients: ||s*g|| = s*||g||
    const float c = max_norm / (sqrtf(gnorm_sq[0]) * grad_scale + 1e-6f);
    clip = c < 1.0f ? c * grad_scale : grad_scale;
  }
  const float step = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  long long done = 0;
  if (((r.p | r.g | r.m | r.v) & 15ull) == 0) {
    const long long nv = r.n >> 2;
    for (long long i = threadIdx.x; i < nv; i += 256) {
      float4 pv = reinterpret_cast<float4*>(p)[i];
      const float4 gv = __ldg(reinterpret_cast<const float4*>(g) + i);
      float4 mv = reinterpret_cast<float4*>(m)[i];
      float4 vv = reinterpret_cast<float4*>(v)[i];
      float* pp = reinterpret_cast<float*>(&pv);
      const float* gg = reinterpret_cast<const float*>(&gv);
      float* mm = reinterpret_cast<float*>(&mv);
      float* vvp = reinterpret_cast<float*>(&vv);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float gr = gg[j] * clip;
        mm[j] = b1 * mm[j] + (1.0f - b1) * gr;
        vvp[j] = b2 * vvp[j] + (1.0f - b2) * gr * gr;
        pp[j] -= step * mm[j] / (sqrtf(vvp[j]) * inv_sqrt_bc2 + eps);
      }
      reinterpret_cast<float4*>(p)[i] = pv;
      reinterpret_cast<float4*>(m)[i] = mv;
      reinterpret_cast<float4*>(v)[i] = vv;
    }
    done = nv << 2;
  }
  for (long long i = done + threadIdx.x; i < r.n; i += 256) {
    const float gr = g[i] * clip;
    const float mi = b1 * m[i] + (1.0f - b1) * gr;
    const float vi = b2 * v[i] + (1.0f - b2) * gr * gr;
    m[i] = mi;
    v[i] = vi;
    p[i] -= step * mi / (sqrtf(vi) * inv_sqrt_bc2 + eps);
  }
}

// ---------------------------------------------------------------------------------------------
// Device-resident step counters (one per PARAMETER): the bias corrections 1 - beta^t are computed in the kernel from
// steps[i], so a CUDA-graph replay of the optimizer advances them without the host, and a parameter whose gradient first
// appears at step k (the decoder after the encoder-pretrain phase) starts its own count there -- torch.optim.Adam skips
// parameters whose .grad is None, it does not age them. row = {p, g, m, v, n, step (float*), active (uint8*)}.
// ---------------------------------------------------------------------------------------------
struct MtRow7 {
  unsigned long long p, g, m, v;
  long long n;
  unsigned long long step, active;
};

__global__ void mt_step_incr_kernel(float* __restrict__ steps, const uint8_t* __restrict__ active, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && (active == nullptr || active[i])) steps[i] += 1.0f;
}

__global__ void __launch_bounds__(256)
mt_adam_dev_kernel(const MtRow7* __restrict__ rows, float lr, float b1, float b2, float eps,
                   const float* __restrict__ gnorm_sq, float max_norm, float grad_scale) {
  const MtRow7 r = rows[blockIdx.x];
  if (r.active != 0ull && *reinterpret_cast<const uint8_t*>(r.active) == 0) return;  // no gradient this step: skipped
  const float t = *reinterpret_cast<const float*>(r.step);  // already incremented for this step
  const float bc1 = 1.0f - powf(b1, t), bc2 = 1.0f - powf(b2, t);
  float* p = reinterpret_cast<float*>(r.p);
  const float* g = reinterpret_cast<const float*>(r.g);
  float* m = reinterpret_cast<float*>(r.m);
  float* v = reinterpret_cast<float*>(r.v);
  float clip = grad_scale;
  if (gnorm_sq != nullptr && max_norm > 0.0f) {
    const float c = max_norm / (sqrtf(gnorm_sq[0]) * grad_scale + 1e-6f);
    clip = c < 1.0f ? c * grad_scale : grad_scale;
  }
  const float step = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  long long done = 0;
  if (((r.p | r.g | r.m | r.v) & 15ull) == 0) {
    const long long nv = r.n >> 2;
    for (long long i = threadIdx.x; i < nv; i += 256) {
      float4 pv = reinterpret_cast<float4*>(p)[i];
      const float4 gv = __ldg(reinterpret_cast<const float4*>(g) + i);
      float4 mv = reinterpret_cast<float4*>(m)[i];
      float4 vv = reinterpret_cast<float4*>(v)[i];
      float* pp = reinterpret_cast<float*>(&pv);
      const float* gg = reinterpret_cast<const float*>(&gv);
      float* mm = reinterpret_cast<float*>(&mv);
      float* vvp = reinterpret_cast<float*>(&vv);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float gr = gg[j] * clip;
        mm[j] = b1 * mm[j] + (1.0f - b1) * gr;
        vvp[j] = b2 * vvp[j] + (1.0f - b2) * gr * gr;
        pp[j] -= step * mm[j] / (sqrtf(vvp[j]) * inv_sqrt_bc2 + eps);
      }
      reinterpret_cast<float4*>(p)[i] = pv;
      reinterpret_cast<float4*>(m)[i] = mv;
      reinterpret_cast<float4*>(v)[i] = vv;
    }
    done = nv << 2;
  }
  for (long long i = done + threadIdx.x; i < r.n; i += 256) {
    const float gr = g[i] * clip;
    const float mi = b1 * m[i] + (1.0f - b1) * gr;
    const float vi = b2 * v[i] + (1.0f - b2) * gr * gr;
    m[i] = mi;
    v[i] = vi;
    p[i] -= step * mi / (sqrtf(vi) * inv_sqrt_bc2 + eps);
  }
}

}  // namespace odvae

using namespace odvae;

extern "C" {

size_t odvae_sumsq_workspace_bytes(int64_t n) {
  (void)n;
  return static_cast<size_t>(num_sms()) * SUMSQ_BLOCKS_PER_SM * sizeof(float) + 256;
}

int odvae_sumsq(const float* g, int64_t n, float* out1, int accumulate, void* workspace, size_t ws_bytes,
                void* stream) {
  const int nblk = num_sms() * SUMSQ_BLOCKS_PER_SM;
  if (ws_bytes < static_cast<size_t>(nblk) * sizeof(float)) return ODVAE_ERR_WORKSPACE;
  if ((reinterpret_cast<uintptr_t>(g) & 15) != 0) return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* part = reinterpret_cast<float*>(workspace);
  sumsq_stage1<<<nblk, 256, 0, st>>>(g, n, part);
  int s = check_launch("sumsq_stage1");
  if (s != ODVAE_OK) return s;
  sumsq_stage2<<<1, 32, 0, st>>>(part, nblk, out1, accumulate);
  return check_launch("sumsq_stage2");
}

int odvae_adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float beta1, float beta2,
                    float eps, float bias_corr1, float bias_corr2, const float* gnorm_sq, float max_norm,
                    void* stream) {
  if (n <= 0) return ODVAE_ERR_SHAPE;
  if (((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
        reinterpret_cast<uintptr_t>(v)) & 15) != 0)
    return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  long long blocks = ((n >> 2) + 255) / 256;
  const long long cap = 8ll * num_sms();
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  adam_kernel<<<static_cast<int>(blocks), 256, 0, st>>>(p, g, m, v, n, lr, beta1, beta2, eps, bias_corr1, bias_corr2,
                                                       gnorm_sq, max_norm);
  return check_launch("adam_kernel");
}

int odvae_mt_sumsq(const void* table, int nrows, float* out1, void* workspace, size_t ws_bytes, void* stream) {
  if (nrows <= 0) return ODVAE_ERR_SHAPE;
  if (ws_bytes < static_cast<size_t>(nrows) * sizeof(float)) return ODVAE_ERR_WORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* part = reinterpret_cast<float*>(workspace);
  mt_sumsq_stage1<<<nrows, 256, 0, st>>>(reinterpret_cast<const MtRow*>(table), part);
  int s = check_launch("mt_sumsq_stage1");
  if (s != ODVAE_OK) return s;
  sumsq_stage2<<<1, 32, 0, st>>>(part, nrows, out1, 0);
  return check_launch("sumsq_stage2");
}

int odvae_mt_adam_step(const void* table, int nrows, float lr, float beta1, float beta2, float eps, float bias_corr1,
                       float bias_corr2, const float* gnorm_sq, float max_norm, float grad_scale, void* stream) {
  if (nrows <= 0) return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  mt_adam_kernel<<<nrows, 256, 0, st>>>(reinterpret_cast<const MtRow*>(table), lr, beta1, beta2, eps, bias_corr1,
                                        bias_corr2, gnorm_sq, max_norm, grad_scale);
  return check_launch("mt_adam_kernel");
}

int odvae_mt_step_incr(float* steps, const uint8_t* active, int n, void* stream) {
  if (n <= 0) return ODVAE_ERR_SHAPE;
  mt_step_incr_kernel<<<(n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(steps, active, n);
  return check_launch("mt_step_incr_kernel");
}

int odvae_mt_adam_step_dev(const void* table7, int nrows, float lr, float beta1, float beta2, float eps,
                           const float* gnorm_sq, float max_norm, float grad_scale, void* stream) {
  if (nrows <= 0) return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  mt_adam_dev_kernel<<<nrows, 256, 0, st>>>(reinterpret_cast<const MtRow7*>(table7), lr, beta1, beta2, eps, gnorm_sq,
                                            max_norm, grad_scale);
  return check_launch("mt_adam_dev_kernel");
}

}  // extern "C"
