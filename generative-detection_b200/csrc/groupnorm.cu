// GroupNorm(32 groups, eps, affine) [+ SiLU] forward / backward on NHWC activations.
// Reference: Normalize (LD model.py:38-39) + nonlinearity x*sigmoid(x) (model.py:33-35).
// Memory-bound: every pass reads whole pixels (all channels, 16-byte vectors, fully coalesced);
// per-(image,channel) partial sums go through a small fp32 workspace and are combined in a fixed
// order (deterministic, no atomics).
#include "common.cuh"
#include "engine_host.h"
#include "tc_primitives.cuh"

#include <stdlib.h>

namespace odvae {

template <typename T> struct VecN;
template <> struct VecN<float> { static constexpr int V = 4; };
template <> struct VecN<bf16> { static constexpr int V = 8; };
template <> struct VecN<f16> { static constexpr int V = 8; };

static constexpr int GN_THREADS = 256;

// MODE 0: per-(b,c) sums of (x, x^2)          -> forward statistics
// MODE 1: per-(b,c) sums of (dy', dy' * xhat)  -> backward reductions, dy' = dy * silu'(u) when silu
// part layout: [B][nsplit][C][2]
template <typename T, typename TG, int MODE>
__global__ void __launch_bounds__(GN_THREADS)
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
    constexpr int U = 4;
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
__global__ void __launch_bounds__(GN_THREADS)
gn_apply_kernel(const T* __restrict__ x, const float* __restrict__ stats, const float* __restrict__ gamma,
                const float* __restrict__ beta, TO* __restrict__ y, TO2* __restrict__ y2, int HW, int C, int G,
                int silu) {
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
        float v = fmaf(to_f32<T>(xv[u][j]), scale[j], shift[j]);
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
      float v = fmaf(to_f32<T>(xv[j]), scale[j], shift[j]);
      if (silu) v = silu_t<sizeof(T) == 2>(v);
      yv[j] = from_f32<TO>(v);
      y2v[j] = from_f32<TO2>(v);
    }
    yb[i] = *reinterpret_cast<const uint4*>(yv);
    if (y2b) y2b[i] = *reinterpret_cast<const uint4*>(y2v);
  }
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
                                         float* __restrict__ dbeta, int B, int C, int accumulate) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float a = 0.0f, bb = 0.0f;
  for (int b = 0; b < B; ++b) {
    a += AB[(static_cast<long long>(b) * C + c) * 2];
    bb += AB[(static_cast<long long>(b) * C + c) * 2 + 1];
  }
  dbeta[c] = accumulate ? dbeta[c] + a : a;
  dgamma[c] = accumulate ? dgamma[c] + bb : bb;
}

// dx = c1 * dy' + c2 * x + c3 (+ dx_add); one channel vector per thread for the whole loop (see gn_apply_kernel)
template <typename T, typename TG>
__global__ void __launch_bounds__(GN_THREADS)
gn_bwd_apply_kernel(const T* __restrict__ x, const TG* __restrict__ dy, const TG* __restrict__ dx_add,
                    const float* __restrict__ stats, const float* __restrict__ gamma, const float* __restrict__ beta,
                    const float* __restrict__ coef, TG* __restrict__ dx, int HW, int C, int G, int silu) {
  constexpr int V = VecN<T>::V;
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
  constexpr int U = 2;
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
          if (silu) d *= silu_grad_t<sizeof(T) == 2>(fmaf(xf, scale[j], shift[j]));
          float o = fmaf(scale[j], d, fmaf(c2[j], xf, c3[j]));
          if (ab) o += to_f32<TG>(av[u][j]);
          ov[j] = from_f32<TG>(o);
        }
        dxb[k] = *reinterpret_cast<const uint4*>(ov);
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Backward in ONE cooperative launch with L2 reuse ("team" path).
// The two-kernel path reads x and dy twice from HBM (reduce, then apply): 5 E s bytes against the algorithmic 3 E s.
// Here the resident grid is split into `nteams` teams (one CTA of every team per SM). A team walks over items of `ipg`
// consecutive images sized to a fraction of the L2 (~32 MB of x + dy): phase 1 reduces the item's pixel ranges
// (copies tagged L2::evict_last), a team barrier, a few duty CTAs combine the partials per (image, group) in a fixed
// order, a second barrier, then phase 2 re-reads the SAME pixel range newest-first out of L2 (copies/stores tagged
// evict_first) and writes dx. ncu: 2.23 GB read + 1.04 GB written per launch at 128 ch x 256^2 x 64 images, against
// 4.30 + 1.04 GB for the two-kernel path.
// Every CTA's pixel range is one contiguous byte range of the NHWC tensor, so the data moves with 1-D bulk async copies
// (cp.async.bulk, 8 KB per tensor and chunk) through a 4-stage shared-memory ring with mbarrier completion. The ring
// runs ahead across the phase boundary: the first chunks of phase 2 are already in flight while the team waits for the
// coefficients, and the next item's phase 1 starts while phase 2 drains - the barriers hide behind the copies.
// Barriers are monotonically increasing arrival counters in the workspace (zeroed by a memset before the launch);
// all sums are fixed-order (deterministic).
// ------------------------------------------------------------------------------------------------------------------
static constexpr int GT_CHUNK_BYTES = 8192;  // per tensor and chunk: 512 16-byte vectors = 2 per thread
static constexpr int GT_MAX_STAGES = 4;  // the host picks the deepest ring that still fits two CTAs per SM

__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
// 1-D bulk async copy global -> shared, completion counted in bytes on an mbarrier
__device__ __forceinline__ void bulk_load(void* dst_smem, const void* src, uint32_t bytes, uint64_t* bar, uint64_t pol,
                                          bool hint) {
  if (hint)
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
            smem_u32(dst_smem)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
        : "memory");
  else
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void st16_hint(void* ptr, const uint4& v, uint64_t pol, bool hint) {
  if (hint)
    asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(ptr), "r"(v.x), "r"(v.y), "r"(v.z),
                 "r"(v.w), "l"(pol)
                 : "memory");
  else
    *reinterpret_cast<uint4*>(ptr) = v;
}
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// all threads of the CTA call; returns after the counter reached `target` (bounded: a protocol bug traps, never hangs)
__device__ __forceinline__ void team_wait(const unsigned* ctr, unsigned target) {
  if (threadIdx.x == 0) {
    if (ld_acquire_u32(ctr) < target) {
      const uint64_t t0 = globaltimer_ns();
      uint32_t spins = 0;
      while (ld_acquire_u32(ctr) < target) {
        if ((++spins & 0xffu) == 0 && globaltimer_ns() - t0 > ODVAE_WAIT_TIMEOUT_NS) __trap();
      }
    }
  }
  __syncthreads();
}
__device__ __forceinline__ void team_arrive(unsigned* ctr) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(ctr, 1u);
  }
}

#ifdef ODVAE_GN_TEAM_TIMING
__device__ unsigned long long gt_dbg[2 * 16 * 8];  // [cta slot][item][stamp]
#define GT_STAMP(k)                                                                                   \
  if (threadIdx.x == 0 && team == 0 && (j == 0 || j == 100) && (it / nteams) < 16)                     \
    gt_dbg[((j == 0 ? 0 : 1) * 16 + it / nteams) * 8 + (k)] = globaltimer_ns();
#else
#define GT_STAMP(k)
#endif
// ctr layout: [team][2] = (A: CTAs that finished phase 1, B: (image, group) pairs finalized)
template <typename T, typename TG>
__global__ void __launch_bounds__(GN_THREADS, 2)
gn_bwd_team_kernel(const T* __restrict__ x, const TG* __restrict__ dy, const TG* __restrict__ dx_add,
                   const float* __restrict__ stats, const float* __restrict__ gamma, const float* __restrict__ beta,
                   float* part, float* AB, float* coef, TG* __restrict__ dx, unsigned* ctr, int B, int HW, int C, int G,
                   int silu, int ipg, int spi, int nteams, int hint, int GT_STAGES) {
  static_assert(sizeof(T) == sizeof(TG), "x and dy chunks share one vector index");
  constexpr int V = VecN<T>::V;
  const int tsize = gridDim.x / nteams;
  const int team = blockIdx.x / tsize;
  if (team >= nteams) return;
  const int j = blockIdx.x - team * tsize;
  unsigned* ctrA = ctr + 2 * team;
  unsigned* ctrB = ctrA + 1;
  const int cpg = C / G;
  const int cv = C / V;
  const int rows = GN_THREADS / cv;
  const int col = threadIdx.x % cv, row = threadIdx.x / cv;
  const int c0 = col * V;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  extern __shared__ __align__(128) unsigned char gt_smem[];
  const int GT_STAGE_BYTES = (dx_add ? 3 : 2) * GT_CHUNK_BYTES;                      // x | dy [| dx_add]
  unsigned char* ring = gt_smem;                                                     // [GT_STAGES][x | dy | add]
  float* red = reinterpret_cast<float*>(gt_smem + GT_STAGES * GT_STAGE_BYTES);       // [2][rows][C]
  uint64_t* full = reinterpret_cast<uint64_t*>(red + 2 * rows * C);                  // [GT_STAGES]
  const unsigned NS = static_cast<unsigned>(GT_STAGES);
  const bool use_hint = hint != 0;
  const uint64_t pol_keep = l2_policy_evict_last();
  const uint64_t pol_drop = l2_policy_evict_first();
  const int nitems = (B + ipg - 1) / ipg;
  const int img = j / spi, split = j - img * spi;
  const int per = (HW + spi - 1) / spi;
  const int p0 = split * per;
  const int p1 = p0 + per < HW ? p0 + per : HW;
  const int len = (img < ipg && p1 > p0) ? p1 - p0 : 0;
  const int CH = GT_CHUNK_BYTES / (C * static_cast<int>(sizeof(T)));  // pixels per chunk
  const int nch = (len + CH - 1) / CH;
  const float inv_n = 1.0f / (static_cast<float>(cpg) * HW);

  if (threadIdx.x == 0) {
    for (int s = 0; s < GT_STAGES; ++s) mbar_init(full + s, 1);
    fence_barrier_init();
  }
  __syncthreads();

  // producer iterator (thread 0 only): item, phase, chunk, running sequence number
  int p_it = team, p_phase = 0, p_i = 0;
  unsigned p_n = 0;
  auto produce_next = [&]() {
    // skip items in which this CTA has no pixels (inactive image slot of a partial last item)
    while (p_it < nitems) {
      const int nimg_p = B - p_it * ipg < ipg ? B - p_it * ipg : ipg;
      if (nch > 0 && img < nimg_p) break;
      p_it += nteams;
    }
    if (p_it >= nitems) return;
    const int c = p_phase == 0 ? p_i : nch - 1 - p_i;
    const int px0 = p0 + c * CH;
    const int npx = px0 + CH <= p1 ? CH : p1 - px0;
    const uint32_t bytes = static_cast<uint32_t>(npx) * C * sizeof(T);
    const long long off = (static_cast<long long>(p_it * ipg + img) * HW + px0) * C;
    const int s = p_n % NS;
    unsigned char* st = ring + s * GT_STAGE_BYTES;
    const bool add = p_phase == 1 && dx_add != nullptr;
    mbar_arrive_expect_tx(full + s, bytes * (add ? 3u : 2u));
    const uint64_t pol = p_phase == 0 ? pol_keep : pol_drop;
    bulk_load(st, x + off, bytes, full + s, pol, use_hint);
    bulk_load(st + GT_CHUNK_BYTES, dy + off, bytes, full + s, pol, use_hint);
    if (add) bulk_load(st + 2 * GT_CHUNK_BYTES, dx_add + off, bytes, full + s, pol_drop, use_hint);
    ++p_n;
    if (++p_i == nch) {
      p_i = 0;
      if (++p_phase == 2) {
        p_phase = 0;
        p_it += nteams;
      }
    }
  };
  // start the teams half a beat apart: team t begins once team t-1 has finished its first phase 1
  if (team > 0) team_wait(ctr + 2 * (team - 1), static_cast<unsigned>(tsize));
  if (threadIdx.x == 0) {
    for (int s = 0; s < GT_STAGES; ++s) produce_next();
  }

  unsigned c_n = 0;  // consumer sequence number
  unsigned targetA = 0, targetB = 0;
  for (int it = team; it < nitems; it += nteams) {
    const int b0 = it * ipg;
    const int nimg = B - b0 < ipg ? B - b0 : ipg;
    const int b = b0 + img;
    const int my_nch = img < nimg ? nch : 0;
    float mu[V], rs[V], sc[V], sh[V];
    GT_STAMP(0)
    // ---------------- phase 1: per-(image, split, channel) sums of (dy', dy' xhat) ----------------
    if (img < nimg) {
      float s0[V], s1[V];
#pragma unroll
      for (int q = 0; q < V; ++q) {
        s0[q] = s1[q] = 0.0f;
        const int c = c0 + q;
        const int g = c / cpg;
        mu[q] = __ldg(stats + (static_cast<long long>(b) * G + g) * 2);
        rs[q] = __ldg(stats + (static_cast<long long>(b) * G + g) * 2 + 1);
        sc[q] = rs[q] * __ldg(gamma + c);
        sh[q] = __ldg(beta + c) - mu[q] * sc[q];
      }
      for (int i = 0; i < my_nch; ++i, ++c_n) {
        const int s = c_n % NS;
        mbar_wait(full + s, (c_n / NS) & 1);
        const int px0 = p0 + i * CH;
        const int nvec = (px0 + CH <= p1 ? CH : p1 - px0) * cv;
        const uint4* sx = reinterpret_cast<const uint4*>(ring + s * GT_STAGE_BYTES);
        const uint4* sd = sx + GT_CHUNK_BYTES / 16;
#pragma unroll
        for (int u = 0; u < GT_CHUNK_BYTES / 16 / GN_THREADS; ++u) {
          const int v = threadIdx.x + u * GN_THREADS;
          if (v < nvec) {
            T xv[V];
            TG dv[V];
            *reinterpret_cast<uint4*>(xv) = sx[v];
            *reinterpret_cast<uint4*>(dv) = sd[v];
#pragma unroll
            for (int q = 0; q < V; ++q) {
              const float xf = to_f32<T>(xv[q]);
              float d = to_f32<TG>(dv[q]);
              if (silu) d *= silu_grad_t<sizeof(T) == 2>(fmaf(xf, sc[q], sh[q]));
              s0[q] += d;
              s1[q] = fmaf(d, (xf - mu[q]) * rs[q], s1[q]);
            }
          }
        }
        __syncthreads();  // stage s is free again
        if (threadIdx.x == 0) produce_next();
      }
      GT_STAMP(1)
      float* r0 = red;
      float* r1 = red + rows * C;
#pragma unroll
      for (int q = 0; q < V; ++q) {
        r0[row * C + c0 + q] = s0[q];
        r1[row * C + c0 + q] = s1[q];
      }
      __syncthreads();
      for (int c = threadIdx.x; c < C; c += GN_THREADS) {
        float a = 0.0f, bb = 0.0f;
        for (int r = 0; r < rows; ++r) {
          a += r0[r * C + c];
          bb += r1[r * C + c];
        }
        __stcg(reinterpret_cast<float2*>(part + ((static_cast<long long>(b) * spi + split) * C + c) * 2),
               make_float2(a, bb));
      }
    }
    GT_STAMP(2)
    team_arrive(ctrA);
    GT_STAMP(3)
    targetA += static_cast<unsigned>(tsize);
    // ---------------- duty: combine the partials of (image, group) pairs, fixed order ----------------
    const int npairs = nimg * G;
    if (j < npairs) {
      team_wait(ctrA, targetA);
      for (int pr = j; pr < npairs; pr += tsize) {
        const int bi = b0 + pr / G, g = pr % G;
        // (channel, value) combos of the group over the warps; lanes stride over the splits
        for (int cb = warp; cb < cpg * 2; cb += GN_THREADS / 32) {
          const int ci = cb >> 1, vi = cb & 1;
          const int c = g * cpg + ci;
          float sacc = 0.0f;
          for (int sp = lane; sp < spi; sp += 32)
            sacc += __ldcg(part + ((static_cast<long long>(bi) * spi + sp) * C + c) * 2 + vi);
          sacc = warp_sum(sacc);
          if (lane == 0) {
            AB[(static_cast<long long>(bi) * C + c) * 2 + vi] = sacc;
            red[cb] = __ldg(gamma + c) * sacc;
          }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
          float s1 = 0.0f, s2 = 0.0f;
          for (int ci = 0; ci < cpg; ++ci) {
            s1 += red[2 * ci];
            s2 += red[2 * ci + 1];
          }
          const float mean = __ldg(stats + (static_cast<long long>(bi) * G + g) * 2);
          const float rstd = __ldg(stats + (static_cast<long long>(bi) * G + g) * 2 + 1);
          float2 cf;
          cf.x = -rstd * rstd * s2 * inv_n;
          cf.y = rstd * (mean * rstd * s2 - s1) * inv_n;
          __stcg(reinterpret_cast<float2*>(coef + (static_cast<long long>(bi) * G + g) * 2), cf);
        }
        team_arrive(ctrB);
      }
    }
    targetB += static_cast<unsigned>(npairs);
    GT_STAMP(4)
    team_wait(ctrB, targetB);
    GT_STAMP(5)
    // ---------------- phase 2: dx = c1 dy' + c2 x + c3 (+ dx_add), newest pixels first ----------------
    if (my_nch > 0) {
      float c2[V], c3[V];
#pragma unroll
      for (int q = 0; q < V; ++q) {
        const int g = (c0 + q) / cpg;
        const float2 cf = __ldcg(reinterpret_cast<const float2*>(coef + (static_cast<long long>(b) * G + g) * 2));
        c2[q] = cf.x;
        c3[q] = cf.y;
      }
      for (int i = 0; i < my_nch; ++i, ++c_n) {
        const int s = c_n % NS;
        mbar_wait(full + s, (c_n / NS) & 1);
        const int px0 = p0 + (my_nch - 1 - i) * CH;
        const int nvec = (px0 + CH <= p1 ? CH : p1 - px0) * cv;
        const uint4* sx = reinterpret_cast<const uint4*>(ring + s * GT_STAGE_BYTES);
        const uint4* sd = sx + GT_CHUNK_BYTES / 16;
        const uint4* sa = sd + GT_CHUNK_BYTES / 16;
        uint4* dxv = reinterpret_cast<uint4*>(dx + (static_cast<long long>(b) * HW + px0) * C);
#pragma unroll
        for (int u = 0; u < GT_CHUNK_BYTES / 16 / GN_THREADS; ++u) {
          const int v = threadIdx.x + u * GN_THREADS;
          if (v < nvec) {
            T xv[V];
            TG dv[V], av[V], ov[V];
            *reinterpret_cast<uint4*>(xv) = sx[v];
            *reinterpret_cast<uint4*>(dv) = sd[v];
            if (dx_add) *reinterpret_cast<uint4*>(av) = sa[v];
#pragma unroll
            for (int q = 0; q < V; ++q) {
              const float xf = to_f32<T>(xv[q]);
              float d = to_f32<TG>(dv[q]);
              if (silu) d *= silu_grad_t<sizeof(T) == 2>(fmaf(xf, sc[q], sh[q]));
              float o = fmaf(sc[q], d, fmaf(c2[q], xf, c3[q]));
              if (dx_add) o += to_f32<TG>(av[q]);
              ov[q] = from_f32<TG>(o);
            }
            st16_hint(dxv + v, *reinterpret_cast<const uint4*>(ov), pol_drop, use_hint);
          }
        }
        __syncthreads();
        if (threadIdx.x == 0) produce_next();
      }
    }
    GT_STAMP(6)
  }
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
static long long env_ll(const char* name, long long dflt) {
  const char* v = getenv(name);
  return v && *v ? atoll(v) : dflt;
}
// Team-path plan for the backward: images per item, splits per image, team size. Returns false -> two-kernel path.
struct GnTeamPlan {
  int nteams, tsize, ipg, spi, hint;
};
template <typename K>
static bool gn_team_plan(K kernel, size_t smem, int B, int HW, int C, size_t bytes_per_image, GnTeamPlan* plan) {
  // opt-in (default off): measured on B200 the team path moves 3.27 GB instead of 5.34 GB per launch (128 ch x 256^2 x
  // 64 images) but runs 1.37 ms against 1.03 ms for the two-kernel path - see profiles/README.md "GroupNorm backward"
  const long long enable = env_ll("ODVAE_GN_BWD_TEAM", 0);
  const long long item_mb = env_ll("ODVAE_GN_TEAM_ITEM_MB", 34);
  const long long min_total_mb = env_ll("ODVAE_GN_TEAM_MIN_MB", 96);
  const long long hint = env_ll("ODVAE_GN_TEAM_HINTS", 1);
  if (!enable) return false;
  const size_t total = bytes_per_image * B;
  if (total < static_cast<size_t>(min_total_mb) << 20) return false;  // already L2-resident between the two kernels
  int per_sm = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, GN_THREADS, smem) != cudaSuccess || per_sm < 2) return false;
  const long long teams = env_ll("ODVAE_GN_TEAMS", 2);
  plan->nteams = teams < 1 ? 1 : (teams > 2 ? 2 : static_cast<int>(teams));
  plan->tsize = 2 * num_sms() / plan->nteams;
  plan->hint = hint ? 1 : 0;
  int ipg = static_cast<int>((static_cast<size_t>(item_mb) << 20) / bytes_per_image);
  if (ipg < 1) ipg = 1;
  if (ipg > B) ipg = B;
  int max_spi = gn_ws_nsplit(B, HW);
  // at least 8 splits per image so a team's CTAs all have work
  while (ipg > 1 && plan->tsize / ipg < 8) --ipg;
  int spi = plan->tsize / ipg;
  if (spi > max_spi) spi = max_spi;
  if (spi < 1) return false;
  plan->ipg = ipg;
  plan->spi = spi;
  return true;
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

#ifdef ODVAE_GN_TEAM_TIMING
__attribute__((visibility("default"))) int odvae_debug_gn_team_timing(unsigned long long* out) {
  return cudaMemcpyFromSymbol(out, gt_dbg, sizeof(gt_dbg)) == cudaSuccess ? 0 : -1;
}
#endif

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
                          void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int V = in_dtype == ODVAE_F32 ? 4 : 8;
  if (!gn_shape_ok(C, G, V)) return ODVAE_ERR_SHAPE;
  dim3 grid(gn_apply_blocks(B, HW, C, V), B);
  if (in_dtype == ODVAE_F32 && out_dtype == ODVAE_F32 && y2 == nullptr)
    gn_apply_kernel<float, float, float><<<grid, GN_THREADS, 0, st>>>(reinterpret_cast<const float*>(x), stats, gamma, beta, reinterpret_cast<float*>(y), nullptr, HW, C, G, silu);
  else if (in_dtype == ODVAE_BF16 && out_dtype == ODVAE_BF16 && (y2 == nullptr || out2_dtype == ODVAE_BF16))
    gn_apply_kernel<bf16, bf16, bf16><<<grid, GN_THREADS, 0, st>>>(reinterpret_cast<const bf16*>(x), stats, gamma, beta, reinterpret_cast<bf16*>(y), reinterpret_cast<bf16*>(y2), HW, C, G, silu);
  else if (in_dtype == ODVAE_F16 && out_dtype == ODVAE_F16 && (y2 == nullptr || out2_dtype == ODVAE_BF16))
    gn_apply_kernel<f16, f16, bf16><<<grid, GN_THREADS, 0, st>>>(reinterpret_cast<const f16*>(x), stats, gamma, beta, reinterpret_cast<f16*>(y), reinterpret_cast<bf16*>(y2), HW, C, G, silu);
  else if (in_dtype == ODVAE_F16 && out_dtype == ODVAE_BF16 && y2 == nullptr)
    gn_apply_kernel<f16, bf16, bf16><<<grid, GN_THREADS, 0, st>>>(reinterpret_cast<const f16*>(x), stats, gamma, beta, reinterpret_cast<bf16*>(y), nullptr, HW, C, G, silu);
  else
    return ODVAE_ERR_SHAPE;
  return check_launch("gn_apply_kernel");
}

// act_dtype: format of x (the saved forward input); grad_dtype: format of dy / dx_add / dx.
// Supported pairs: (f32, f32), (bf16, bf16), (fp16, bf16).
int odvae_groupnorm_bwd(const void* x, const float* stats, const float* gamma, const float* beta, const void* dy,
                        const void* dx_add, void* dx, float* dgamma, float* dbeta, int B, int HW, int C, int G,
                        int silu, int act_dtype, int grad_dtype, int accumulate, void* workspace, size_t ws_bytes,
                        void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int V = act_dtype == ODVAE_F32 ? 4 : 8;
  if (!gn_shape_ok(C, G, V)) return ODVAE_ERR_SHAPE;
  const int combo = (act_dtype == ODVAE_F32 && grad_dtype == ODVAE_F32)     ? 0
                    : (act_dtype == ODVAE_BF16 && grad_dtype == ODVAE_BF16) ? 1
                    : (act_dtype == ODVAE_F16 && grad_dtype == ODVAE_BF16)  ? 2
                                                                            : -1;
  if (combo < 0) return ODVAE_ERR_SHAPE;
  if (ws_bytes < odvae_groupnorm_workspace_bytes(B, HW, C, G) - 256) return ODVAE_ERR_WORKSPACE;
  float* part = reinterpret_cast<float*>(workspace);
  float* AB = part + static_cast<size_t>(B) * gn_ws_nsplit(B, HW) * C * 2;
  float* coef = AB + static_cast<size_t>(B) * C * 2;
  const int rows = GN_THREADS / (C / V);
  const size_t smem = static_cast<size_t>(2) * rows * C * sizeof(float);
  {
    // one cooperative launch with L2 reuse when the tensors are far larger than the L2
    const size_t esz_x = act_dtype == ODVAE_F32 ? 4 : 2, esz_g = grad_dtype == ODVAE_F32 ? 4 : 2;
    const size_t bytes_per_image = static_cast<size_t>(HW) * C * (esz_x + esz_g);
    unsigned* ctr = reinterpret_cast<unsigned*>(coef + static_cast<size_t>(B) * G * 2);
    GnTeamPlan plan;
    int launched = 0;
    int nstages = GT_MAX_STAGES;
    size_t smem_team = 0;
#define GN_TRY_TEAM(T, TG)                                                                                            \
  {                                                                                                                   \
    auto kfn = gn_bwd_team_kernel<T, TG>;                                                                             \
    static bool attr_set = false; /* benign race: idempotent */                                                      \
    if (!attr_set) {                                                                                                  \
      ODVAE_CUDA_TRY(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024));             \
      attr_set = true;                                                                                                \
    }                                                                                                                 \
    bool ok = false;                                                                                                  \
    if (ws_bytes >= odvae_groupnorm_workspace_bytes(B, HW, C, G))                                                     \
      for (nstages = GT_MAX_STAGES; nstages >= 2 && !ok; --nstages) {                                                 \
        smem_team = static_cast<size_t>(nstages) * (dx_add ? 3 : 2) * GT_CHUNK_BYTES + smem + 64;                                       \
        ok = smem_team <= 120 * 1024 && gn_team_plan(kfn, smem_team, B, HW, C, bytes_per_image, &plan);               \
        if (ok) break;                                                                                                \
      }                                                                                                               \
    if (ok) {                                                                                                         \
      ODVAE_CUDA_TRY(cudaMemsetAsync(ctr, 0, 64, st));                                                                \
      const T* xa = reinterpret_cast<const T*>(x);                                                                    \
      const TG* dya = reinterpret_cast<const TG*>(dy);                                                                \
      const TG* adda = reinterpret_cast<const TG*>(dx_add);                                                           \
      TG* dxa = reinterpret_cast<TG*>(dx);                                                                            \
      void* args[] = {&xa, &dya, &adda, &stats, &gamma, &beta, &part, &AB, &coef, &dxa, &ctr, &B, &HW, &C, &G, &silu,  \
                      &plan.ipg, &plan.spi, &plan.nteams, &plan.hint, &nstages};                                                \
      ODVAE_CUDA_TRY(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kfn), dim3(plan.nteams * plan.tsize),        \
                                                 dim3(GN_THREADS), args, smem_team, st));                             \
      launched = 1;                                                                                                   \
    }                                                                                                                 \
  }
    if (combo == 1) GN_TRY_TEAM(bf16, bf16)
    else if (combo == 2) GN_TRY_TEAM(f16, bf16)
    else GN_TRY_TEAM(float, float)
#undef GN_TRY_TEAM
    if (launched) {
      int s = check_launch("gn_bwd_team_kernel");
      if (s != ODVAE_OK) return s;
      gn_bwd_finalize_b_kernel<<<(C + 127) / 128, 128, 0, st>>>(AB, dgamma, dbeta, B, C, accumulate);
      return check_launch("gn_bwd_finalize_b_kernel");
    }
  }
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
  if (combo == 1) GN_LAUNCH_REDUCE1(bf16, bf16)
  else if (combo == 2) GN_LAUNCH_REDUCE1(f16, bf16)
  else GN_LAUNCH_REDUCE1(float, float)
#undef GN_LAUNCH_REDUCE1
  int s = check_launch("gn_reduce_kernel<1>");
  if (s != ODVAE_OK) return s;
  gn_bwd_finalize_a_kernel<<<B, 256, static_cast<size_t>(2) * C * sizeof(float), st>>>(part, stats, gamma, AB, coef, ns, HW, C, G);
  s = check_launch("gn_bwd_finalize_a_kernel");
  if (s != ODVAE_OK) return s;
  gn_bwd_finalize_b_kernel<<<(C + 127) / 128, 128, 0, st>>>(AB, dgamma, dbeta, B, C, accumulate);
  s = check_launch("gn_bwd_finalize_b_kernel");
  if (s != ODVAE_OK) return s;
  dim3 grid2(gn_apply_blocks(B, HW, C, V), B);
  if (combo == 1)
    gn_bwd_apply_kernel<bf16, bf16><<<grid2, GN_THREADS, 0, st>>>(reinterpret_cast<const bf16*>(x), reinterpret_cast<const bf16*>(dy), reinterpret_cast<const bf16*>(dx_add), stats, gamma, beta, coef, reinterpret_cast<bf16*>(dx), HW, C, G, silu);
  else if (combo == 2)
    gn_bwd_apply_kernel<f16, bf16><<<grid2, GN_THREADS, 0, st>>>(reinterpret_cast<const f16*>(x), reinterpret_cast<const bf16*>(dy), reinterpret_cast<const bf16*>(dx_add), stats, gamma, beta, coef, reinterpret_cast<bf16*>(dx), HW, C, G, silu);
  else
    gn_bwd_apply_kernel<float, float><<<grid2, GN_THREADS, 0, st>>>(reinterpret_cast<const float*>(x), reinterpret_cast<const float*>(dy), reinterpret_cast<const float*>(dx_add), stats, gamma, beta, coef, reinterpret_cast<float*>(dx), HW, C, G, silu);
  return check_launch("gn_bwd_apply_kernel");
}

}  // extern "C"
