// "Small heads": the per-sample scalar work around the pose MLPs and the whole PoseLoss reduction tail.
// Reference: PoseAutoencoder._decode_pose / _decode_pose_to_distribution (src/models/autoencoder.py:126-160),
// PoseEncoderSpatialVAE.forward tile+add (src/modules/autoencodermodules/pose_encoder.py:121-125),
// PoseLoss.forward and helpers (src/modules/losses/contperceptual.py:111-132,147-164,176-212,214-321),
// DiagonalGaussianDistribution.kl(other) (src/util/distributions.py:29-41), mmdet FocalLoss() defaults.
// The reference runs these as dozens of tiny launches plus a Python loop; here each direction is ONE launch of
// one CTA with fixed-order reductions (bit-stable, no atomics, no host sync for the `sum(mask_bg) > 0` branches).
#include "common.cuh"

namespace odvae {

static constexpr int HEADS_THREADS = 256;
static constexpr int NBBOX = 8;  // POSE_6D_DIM(4) + LHW_DIM(3) + FILL_FACTOR_DIM(1)

__device__ __forceinline__ float block_sum(float v, float* sm) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sm[w] = v;
  __syncthreads();
  float r = l < (HEADS_THREADS / 32) ? sm[l] : 0.0f;
  return warp_sum(r);
}

// ---------------------------------------------------------------------------------------------
// tile + add of the pose encoder: h[b][p*F+f] = hx[p*F+f] + hz[b][f]
// ---------------------------------------------------------------------------------------------
__global__ void tile_add_fwd_kernel(const float* __restrict__ hx, const float* __restrict__ hz, float* __restrict__ h,
                                    int B, int P, int F) {
  const long long n = static_cast<long long>(B) * P * F;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int b = static_cast<int>(i / (P * F));
    const int r = static_cast<int>(i - static_cast<long long>(b) * P * F);
    h[i] = hx[r] + hz[b * F + (r % F)];
  }
}
// dhx[r] = sum_b dh[b][r] ; dhz[b][f] = sum_p dh[b][p*F+f]   (one thread per output, fixed order)
__global__ void tile_add_bwd_kernel(const float* __restrict__ dh, float* __restrict__ dhx, float* __restrict__ dhz,
                                    int B, int P, int F) {
  const int PF = P * F;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < PF) {
    float a = 0.0f;
    for (int b = 0; b < B; ++b) a += dh[static_cast<long long>(b) * PF + i];
    dhx[i] = a;
  } else if (i < PF + B * F) {
    const int j = i - PF;
    const int b = j / F, f = j - b * F;
    float a = 0.0f;
    for (int p = 0; p < P; ++p) a += dh[static_cast<long long>(b) * PF + p * F + f];
    dhz[j] = a;
  }
}

// ---------------------------------------------------------------------------------------------
// pose head: zz[B][2*8+ncls] -> bbox posterior (mean, clamped logvar) and dec_pose = [mean + std*eps | logits]
// ---------------------------------------------------------------------------------------------
__global__ void pose_head_fwd_kernel(const float* __restrict__ zz, const float* __restrict__ eps,
                                     float* __restrict__ dec_pose, float* __restrict__ mean, float* __restrict__ logvar,
                                     int B, int ncls) {
  const int ld = 2 * NBBOX + ncls, ldo = NBBOX + ncls;
  const int n = B * ldo;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int b = i / ldo, j = i - b * ldo;
    if (j < NBBOX) {
      const float mu = zz[b * ld + j];
      const float lv = fminf(fmaxf(zz[b * ld + NBBOX + j], -30.0f), 20.0f);
      mean[b * NBBOX + j] = mu;
      logvar[b * NBBOX + j] = lv;
      const float e = eps != nullptr ? eps[b * NBBOX + j] : 0.0f;  // eps == NULL: mode()
      dec_pose[i] = mu + expf(0.5f * lv) * e;
    } else {
      dec_pose[i] = zz[b * ld + 2 * NBBOX + (j - NBBOX)];
    }
  }
}
// dzz from d(dec_pose), d(mean), d(logvar) (any may be NULL)
__global__ void pose_head_bwd_kernel(const float* __restrict__ zz, const float* __restrict__ eps,
                                     const float* __restrict__ ddec, const float* __restrict__ dmean,
                                     const float* __restrict__ dlogvar, float* __restrict__ dzz, int B, int ncls) {
  const int ld = 2 * NBBOX + ncls, ldo = NBBOX + ncls;
  const int n = B * ld;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int b = i / ld, j = i - b * ld;
    float g = 0.0f;
    if (j < NBBOX) {
      if (ddec) g += ddec[b * ldo + j];
      if (dmean) g += dmean[b * NBBOX + j];
    } else if (j < 2 * NBBOX) {
      const int a = j - NBBOX;
      const float raw = zz[i];
      if (raw >= -30.0f && raw <= 20.0f) {  // torch.clamp passes the gradient on the closed interval
        const float lv = raw;
        if (ddec && eps) g += ddec[b * ldo + a] * eps[b * NBBOX + a] * 0.5f * expf(0.5f * lv);
        if (dlogvar) g += dlogvar[b * NBBOX + a];
      }
    } else {
      if (ddec) g += ddec[b * ldo + NBBOX + (j - 2 * NBBOX)];
    }
    dzz[i] = g;
  }
}

// ---------------------------------------------------------------------------------------------
// loss tail. out[] indices (fp32):
//   0 total  1 nll  2 kl_obj  3 rec_mean  4 pose  5 class  6 bbox  7 fill  8 kl_bbox
//   9 t1_mean 10 t2_mean 11 t3_mean 12 v3_mean 13 n_fg (= sum mask_bg)
// w[] (host-side weights folded into `total`): 0 nll 1 kl_obj 2 pose 3 class 4 bbox 5 fill 6 kl_bbox
// ---------------------------------------------------------------------------------------------
struct LossHeadsArgs {
  const float* dec_pose;    // [B][8+ncls]
  const float* bbox_mean;   // [B][8]
  const float* bbox_logvar; // [B][8]
  const float* l1_sums;     // [B] or NULL (pixel loss off)
  const float* kl_obj;      // [B] or NULL
  const float* logvar;      // scalar parameter
  const float* pose_gt;     // [B][4]
  const float* bbox_gt;     // [B][3]
  const float* fill_gt;     // [B]
  const long long* class_gt;  // [B]
  const int* prior_row;     // [B] row of the per-class prior table, -1 = "background" label (skipped)
  const float* prior;       // [L][8][2] (mean, clamped logvar)
  int B, ncls, background_idx;
  float npix;               // C*H*W elements per sample of the reconstruction
  float w[7];
  float focal_gamma, focal_alpha;
};

__device__ __forceinline__ float softplus_neg_abs(float x) { return log1pf(expf(-fabsf(x))); }

__global__ void __launch_bounds__(HEADS_THREADS) loss_heads_fwd_kernel(const LossHeadsArgs a, float* __restrict__ out) {
  __shared__ float sm[HEADS_THREADS / 32];
  const int B = a.B, ncls = a.ncls, ldo = NBBOX + ncls;
  const float lv = a.logvar[0];
  const float inv_var = 1.0f / (expf(lv) + 1e-8f);
  float s_nfg = 0, s_nll = 0, s_kl = 0, s_rec = 0, s_pose = 0, s_cls = 0, s_bbox = 0, s_fill = 0, s_klb = 0;
  float s_t1 = 0, s_t2 = 0, s_t3 = 0, s_v3 = 0;
  for (int b = threadIdx.x; b < B; b += HEADS_THREADS) {
    const float m = a.class_gt[b] != a.background_idx ? 1.0f : 0.0f;
    const float* dp = a.dec_pose + b * ldo;
    s_nfg += m;
    const float l1 = a.l1_sums ? a.l1_sums[b] : 0.0f;
    s_rec += l1;
    s_nll += m * (l1 * inv_var + lv * a.npix);
    if (a.kl_obj) s_kl += m * a.kl_obj[b];
    const float t1 = fabsf(a.pose_gt[b * 4 + 0] - dp[0]);
    const float t2 = fabsf(a.pose_gt[b * 4 + 1] - dp[1]);
    const float t3 = fabsf(a.pose_gt[b * 4 + 2] - dp[2]);
    const float d = sinf(a.pose_gt[b * 4 + 3]) - sinf(dp[3]);
    const float v3 = fabsf(d) < 1.0f ? 0.5f * d * d : fabsf(d) - 0.5f;
    s_t1 += t1; s_t2 += t2; s_t3 += t3; s_v3 += v3;
    s_pose += m * (t1 + t2 + t3 + v3);
    float bb = 0.0f;
    for (int j = 0; j < 3; ++j) {
      const float e = a.bbox_gt[b * 3 + j] - dp[4 + j];
      bb += e * e;
    }
    s_bbox += m * bb;
    const float ef = a.fill_gt[b] - dp[7];
    s_fill += m * ef * ef;
    const long long t = a.class_gt[b];
    for (int c = 0; c < ncls; ++c) {
      const float x = dp[NBBOX + c];
      const float y = (c == t) ? 1.0f : 0.0f;
      const float p = 1.0f / (1.0f + expf(-x));
      const float pt = (1.0f - p) * y + p * (1.0f - y);
      const float wf = (a.focal_alpha * y + (1.0f - a.focal_alpha) * (1.0f - y)) * powf(pt, a.focal_gamma);
      const float bce = fmaxf(x, 0.0f) - x * y + softplus_neg_abs(x);
      s_cls += bce * wf;
    }
    const int row = a.prior_row[b];
    if (row >= 0) {
      const float* pr = a.prior + row * NBBOX * 2;
      float acc = 0.0f;
      for (int i = 0; i < NBBOX; ++i) {
        const float mu = a.bbox_mean[b * NBBOX + i], l = a.bbox_logvar[b * NBBOX + i], var = expf(l);
        for (int j = 0; j < NBBOX; ++j) {
          const float pm = pr[2 * j], pl = pr[2 * j + 1];
          const float ipv = 1.0f / (expf(pl) + 1e-5f);
          acc += (mu - pm) * (mu - pm) * ipv + var * ipv - 1.0f - l + pl;
        }
      }
      s_klb += 0.5f * acc;
    }
  }
  const float nfg = block_sum(s_nfg, sm);
  const float nll = block_sum(s_nll, sm), kl = block_sum(s_kl, sm), rec = block_sum(s_rec, sm);
  const float pose = block_sum(s_pose, sm), cls = block_sum(s_cls, sm), bbox = block_sum(s_bbox, sm);
  const float fill = block_sum(s_fill, sm), klb = block_sum(s_klb, sm);
  const float t1 = block_sum(s_t1, sm), t2 = block_sum(s_t2, sm), t3 = block_sum(s_t3, sm), v3 = block_sum(s_v3, sm);
  if (threadIdx.x == 0) {
    const float inv = nfg > 0.0f ? 1.0f / nfg : 0.0f;  // reference: torch.tensor(0.0) when sum(mask_bg) == 0
    const float o_nll = nll * inv, o_kl = kl * inv, o_pose = pose * inv, o_bbox = bbox * inv, o_fill = fill * inv;
    const float o_klb = klb * inv, o_cls = cls / (static_cast<float>(B) * ncls);
    out[1] = o_nll; out[2] = o_kl; out[3] = rec / (static_cast<float>(B) * a.npix);
    out[4] = o_pose; out[5] = o_cls; out[6] = o_bbox; out[7] = o_fill; out[8] = o_klb;
    out[9] = t1 / B; out[10] = t2 / B; out[11] = t3 / B; out[12] = v3 / B; out[13] = nfg;
    out[0] = a.w[2] * o_pose + a.w[0] * o_nll + a.w[3] * o_cls + a.w[4] * o_bbox + a.w[5] * o_fill +
             a.w[1] * o_kl + a.w[6] * o_klb;
    out[14] = out[15] = 0.0f;
  }
}

// dout[16] = upstream gradient of every out[] entry; effective coefficient of a term = dout[term] + dout[0]*w[term]
__global__ void __launch_bounds__(HEADS_THREADS)
loss_heads_bwd_kernel(const LossHeadsArgs a, const float* __restrict__ out, const float* __restrict__ dout,
                      float* __restrict__ d_dec_pose, float* __restrict__ d_mean, float* __restrict__ d_blogvar,
                      float* __restrict__ d_l1, float* __restrict__ d_kl, float* __restrict__ d_logvar) {
  __shared__ float sm[HEADS_THREADS / 32];
  const int B = a.B, ncls = a.ncls, ldo = NBBOX + ncls;
  const float nfg = out[13];
  const float inv = nfg > 0.0f ? 1.0f / nfg : 0.0f;
  const float g0 = dout[0];
  const float c_nll = (dout[1] + g0 * a.w[0]) * inv, c_kl = (dout[2] + g0 * a.w[1]) * inv;
  const float c_pose = (dout[4] + g0 * a.w[2]) * inv, c_cls = (dout[5] + g0 * a.w[3]) / (static_cast<float>(B) * ncls);
  const float c_bbox = (dout[6] + g0 * a.w[4]) * inv, c_fill = (dout[7] + g0 * a.w[5]) * inv;
  const float c_klb = (dout[8] + g0 * a.w[6]) * inv;
  const float c_rec = dout[3] / (static_cast<float>(B) * a.npix);
  const float c_t1 = dout[9] / B, c_t2 = dout[10] / B, c_t3 = dout[11] / B, c_v3 = dout[12] / B;
  const float lv = a.logvar[0];
  const float ev = expf(lv);
  const float inv_var = 1.0f / (ev + 1e-8f);
  float s_dlv = 0.0f;
  for (int b = threadIdx.x; b < B; b += HEADS_THREADS) {
    const float m = a.class_gt[b] != a.background_idx ? 1.0f : 0.0f;
    const float* dp = a.dec_pose + b * ldo;
    float* g = d_dec_pose + b * ldo;
    const float l1 = a.l1_sums ? a.l1_sums[b] : 0.0f;
    if (d_l1) d_l1[b] = c_nll * m * inv_var + c_rec;
    if (d_kl) d_kl[b] = c_kl * m;
    s_dlv += c_nll * m * (a.npix - l1 * ev * inv_var * inv_var);
    const float cp = c_pose * m;
    for (int k = 0; k < 3; ++k) {
      const float e = a.pose_gt[b * 4 + k] - dp[k];
      const float sg = e > 0.0f ? 1.0f : (e < 0.0f ? -1.0f : 0.0f);
      const float ct = k == 0 ? c_t1 : (k == 1 ? c_t2 : c_t3);
      g[k] = -(cp + ct) * sg;
    }
    const float d = sinf(a.pose_gt[b * 4 + 3]) - sinf(dp[3]);
    const float dv = fabsf(d) < 1.0f ? d : (d > 0.0f ? 1.0f : -1.0f);
    g[3] = -(cp + c_v3) * dv * cosf(dp[3]);
    for (int j = 0; j < 3; ++j) g[4 + j] = -2.0f * c_bbox * m * (a.bbox_gt[b * 3 + j] - dp[4 + j]);
    g[7] = -2.0f * c_fill * m * (a.fill_gt[b] - dp[7]);
    const long long t = a.class_gt[b];
    for (int c = 0; c < ncls; ++c) {
      const float x = dp[NBBOX + c];
      const float y = (c == t) ? 1.0f : 0.0f;
      const float p = 1.0f / (1.0f + expf(-x));
      const float pt = (1.0f - p) * y + p * (1.0f - y);
      const float af = a.focal_alpha * y + (1.0f - a.focal_alpha) * (1.0f - y);
      const float wf = af * powf(pt, a.focal_gamma);
      const float bce = fmaxf(x, 0.0f) - x * y + softplus_neg_abs(x);
      const float dpt = (1.0f - 2.0f * y) * p * (1.0f - p);
      const float dwf = pt > 0.0f ? af * a.focal_gamma * powf(pt, a.focal_gamma - 1.0f) * dpt : 0.0f;
      g[NBBOX + c] = c_cls * (wf * (p - y) + bce * dwf);
    }
    const int row = a.prior_row[b];
    for (int i = 0; i < NBBOX; ++i) {
      float gm = 0.0f, gl = 0.0f;
      if (row >= 0) {
        const float* pr = a.prior + row * NBBOX * 2;
        const float mu = a.bbox_mean[b * NBBOX + i], var = expf(a.bbox_logvar[b * NBBOX + i]);
        for (int j = 0; j < NBBOX; ++j) {
          const float ipv = 1.0f / (expf(pr[2 * j + 1]) + 1e-5f);
          gm += (mu - pr[2 * j]) * ipv;
          gl += 0.5f * (var * ipv - 1.0f);
        }
      }
      d_mean[b * NBBOX + i] = c_klb * gm;
      d_blogvar[b * NBBOX + i] = c_klb * gl;
    }
  }
  const float dlv = block_sum(s_dlv, sm);
  if (threadIdx.x == 0) d_logvar[0] = dlv;
}

}  // namespace odvae

using namespace odvae;

static LossHeadsArgs make_args(const float* dec_pose, const float* bbox_mean, const float* bbox_logvar,
                               const float* l1_sums, const float* kl_obj, const float* logvar, const float* pose_gt,
                               const float* bbox_gt, const float* fill_gt, const int64_t* class_gt,
                               const int* prior_row, const float* prior, int B, int ncls, int background_idx,
                               float npix, const float* weights7) {
  LossHeadsArgs a;
  a.dec_pose = dec_pose; a.bbox_mean = bbox_mean; a.bbox_logvar = bbox_logvar; a.l1_sums = l1_sums; a.kl_obj = kl_obj;
  a.logvar = logvar; a.pose_gt = pose_gt; a.bbox_gt = bbox_gt; a.fill_gt = fill_gt;
  a.class_gt = reinterpret_cast<const long long*>(class_gt);
  a.prior_row = prior_row; a.prior = prior; a.B = B; a.ncls = ncls; a.background_idx = background_idx; a.npix = npix;
  for (int i = 0; i < 7; ++i) a.w[i] = weights7[i];
  a.focal_gamma = 2.0f;
  a.focal_alpha = 0.25f;
  return a;
}

extern "C" {

int odvae_tile_add_fwd(const float* hx, const float* hz, float* h, int B, int P, int F, void* stream) {
  if (B <= 0 || P <= 0 || F <= 0) return ODVAE_ERR_SHAPE;
  const long long n = static_cast<long long>(B) * P * F;
  int blocks = static_cast<int>((n + 255) / 256);
  if (blocks > 4 * num_sms()) blocks = 4 * num_sms();
  tile_add_fwd_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(hx, hz, h, B, P, F);
  return check_launch("tile_add_fwd_kernel");
}

int odvae_tile_add_bwd(const float* dh, float* dhx, float* dhz, int B, int P, int F, void* stream) {
  if (B <= 0 || P <= 0 || F <= 0) return ODVAE_ERR_SHAPE;
  const int n = P * F + B * F;
  tile_add_bwd_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(dh, dhx, dhz, B, P, F);
  return check_launch("tile_add_bwd_kernel");
}

int odvae_pose_head_fwd(const float* zz, const float* eps, float* dec_pose, float* mean, float* logvar, int B,
                        int ncls, void* stream) {
  if (B <= 0 || ncls <= 0) return ODVAE_ERR_SHAPE;
  const int n = B * (NBBOX + ncls);
  pose_head_fwd_kernel<<<(n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(zz, eps, dec_pose, mean, logvar,
                                                                                        B, ncls);
  return check_launch("pose_head_fwd_kernel");
}

int odvae_pose_head_bwd(const float* zz, const float* eps, const float* d_dec_pose, const float* d_mean,
                        const float* d_logvar, float* dzz, int B, int ncls, void* stream) {
  if (B <= 0 || ncls <= 0) return ODVAE_ERR_SHAPE;
  const int n = B * (2 * NBBOX + ncls);
  pose_head_bwd_kernel<<<(n + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(zz, eps, d_dec_pose, d_mean,
                                                                                        d_logvar, dzz, B, ncls);
  return check_launch("pose_head_bwd_kernel");
}

int odvae_loss_heads_fwd(const float* dec_pose, const float* bbox_mean, const float* bbox_logvar,
                         const float* l1_sums, const float* kl_obj, const float* logvar, const float* pose_gt,
                         const float* bbox_gt, const float* fill_gt, const int64_t* class_gt, const int* prior_row,
                         const float* prior, int B, int ncls, int background_idx, float npix, const float* weights7,
                         float* out16, void* stream) {
  if (B <= 0 || ncls <= 0 || weights7 == nullptr) return ODVAE_ERR_SHAPE;
  const LossHeadsArgs a = make_args(dec_pose, bbox_mean, bbox_logvar, l1_sums, kl_obj, logvar, pose_gt, bbox_gt, fill_gt,
                                    class_gt, prior_row, prior, B, ncls, background_idx, npix, weights7);
  loss_heads_fwd_kernel<<<1, HEADS_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(a, out16);
  return check_launch("loss_heads_fwd_kernel");
}

int odvae_loss_heads_bwd(const float* dec_pose, const float* bbox_mean, const float* bbox_logvar,
                         const float* l1_sums, const float* kl_obj, const float* logvar, const float* pose_gt,
                         const float* bbox_gt, const float* fill_gt, const int64_t* class_gt, const int* prior_row,
                         const float* prior, int B, int ncls, int background_idx, float npix, const float* weights7,
                         const float* out16, const float* dout16, float* d_dec_pose, float* d_bbox_mean,
                         float* d_bbox_logvar, float* d_l1_sums, float* d_kl_obj, float* d_logvar, void* stream) {
  if (B <= 0 || ncls <= 0 || weights7 == nullptr) return ODVAE_ERR_SHAPE;
  const LossHeadsArgs a = make_args(dec_pose, bbox_mean, bbox_logvar, l1_sums, kl_obj, logvar, pose_gt, bbox_gt, fill_gt,
                                    class_gt, prior_row, prior, B, ncls, background_idx, npix, weights7);
  loss_heads_bwd_kernel<<<1, HEADS_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
      a, out16, dout16, d_dec_pose, d_bbox_mean, d_bbox_logvar, d_l1_sums, d_kl_obj, d_logvar);
  return check_launch("loss_heads_bwd_kernel");
}

}  // extern "C"
