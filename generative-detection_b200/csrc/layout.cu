// Layout / dtype conversion, weight packing, nearest-2x upsample and small reductions.
// All memory-bound: 16-byte vector accesses, grids sized in multiples of the SM count.
#include <mutex>

#include "common.cuh"

namespace odvae {

template <typename T> struct Vec16;  // 16-byte vector of T
template <> struct Vec16<float> { static constexpr int N = 4; };
template <> struct Vec16<bf16> { static constexpr int N = 8; };
template <> struct Vec16<f16> { static constexpr int N = 8; };

// ---------------------------------------------------------------------------------------------
// NCHW fp32 <-> NHWC T
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ src, T* __restrict__ dst, int B, int C, int HW,
                                    int Cpad, float scale, float shift) {
  const long long total = static_cast<long long>(B) * HW;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int b = static_cast<int>(i / HW);
    const int p = static_cast<int>(i - static_cast<long long>(b) * HW);
    const float* s = src + (static_cast<long long>(b) * C) * HW + p;
    T* d = dst + i * Cpad;
    // 32 bytes (one full sector) per store: a thread owns a whole pixel, so 16-byte stores would leave every
    // instruction with half-filled sectors
    constexpr int V2 = 2 * Vec16<T>::N;
    for (int c0 = 0; c0 < Cpad; c0 += V2) {
      __align__(16) T tmp[V2];
#pragma unroll
      for (int j = 0; j < V2; ++j) {
        const int c = c0 + j;
        float v = (c < C) ? __ldg(s + static_cast<long long>(c) * HW) * scale + shift : 0.0f;
        tmp[j] = from_f32<T>(v);
      }
      if (c0 + V2 <= Cpad) {
        st_global_256(d + c0, *reinterpret_cast<const uint32_t(*)[8]>(tmp));
      } else {
        *reinterpret_cast<uint4*>(d + c0) = *reinterpret_cast<const uint4*>(tmp);
      }
    }
  }
}

// _rescale + NCHW fp32 copy + NHWC conversion in one pass (f4): thread = pixel
template <typename T>
__global__ void rescale_nhwc_kernel(const float* __restrict__ src, const float* __restrict__ mm, float* __restrict__ dst_nchw,
                                    T* __restrict__ dst_nhwc, int B, int C, int HW, int Cpad) {
  const float mn = mm[0], range = mm[1] - mm[0];
  const long long total = static_cast<long long>(B) * HW;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int b = static_cast<int>(i / HW);
    const int p = static_cast<int>(i - static_cast<long long>(b) * HW);
    const long long plane0 = (static_cast<long long>(b) * C) * HW + p;
    T* d = dst_nhwc + i * Cpad;
    constexpr int V2 = 2 * Vec16<T>::N;
    for (int c0 = 0; c0 < Cpad; c0 += V2) {
      __align__(16) T tmp[V2];
#pragma unroll
      for (int j = 0; j < V2; ++j) {
        const int c = c0 + j;
        float v = 0.0f;
        if (c < C) {
          v = 2.0f * (__ldg(src + plane0 + static_cast<long long>(c) * HW) - mn) / range - 1.0f;  // == rescale_kernel
          dst_nchw[plane0 + static_cast<long long>(c) * HW] = v;
        }
        tmp[j] = from_f32<T>(v);
      }
      if (c0 + V2 <= Cpad) {
        st_global_256(d + c0, *reinterpret_cast<const uint32_t(*)[8]>(tmp));
      } else {
        *reinterpret_cast<uint4*>(d + c0) = *reinterpret_cast<const uint4*>(tmp);
      }
    }
  }
}

// Thin-channel 3x3 (stride 1, pad 1) im2col: src NCHW fp32 [B,C,H,W] with C <= 4 -> col NHWC [B,H,W,64] 16-bit,
// col[b,y,x, t*4 + c] = scale * src[b,c, y + s*(t/3 - 1), x + s*(t%3 - 1)] (zero outside the image; columns 36.. and
// channels >= C zero), s = +1 for a forward / weight-gradient operand, -1 for the data-gradient operand (dy seen from the
// input pixel). The 3-channel edge convs (encoder.conv_in, decoder.conv_out) then run as ONE K = 64 GEMM step per tile instead
// of nine 64-channel-padded taps. Eight lanes share a pixel: lane j writes the 16-byte chunk of taps 2j, 2j+1, so a warp
// stores 512 contiguous bytes.
template <typename T>
__global__ void im2col3_thin_kernel(const float* __restrict__ src, T* __restrict__ col, int C, int H, int W, int sgn,
                                    float scale) {
  // grid (x chunks, y / 4, b): no index divisions (a flat 64-bit index cost more instructions than the loads and stores), and
  // four rows per thread so that 24 loads are in flight per thread
  constexpr int ROWS = 4;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  const int x = q >> 3, j = q & 7;
  if (x >= W) return;
  const int y0 = blockIdx.y * ROWS, b = blockIdx.z;
  const long long HW = static_cast<long long>(H) * W;
  const float* img = src + static_cast<long long>(b) * C * HW;
  float f[ROWS][8];
#pragma unroll
  for (int r = 0; r < ROWS; ++r) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int t = 2 * j + h;
      const int ty = t / 3, tx = t - 3 * ty;
      const int yy = y0 + r + sgn * (ty - 1), xx = x + sgn * (tx - 1);
      const bool ok = t < 9 && yy >= 0 && yy < H && xx >= 0 && xx < W;
      const float* px = img + static_cast<long long>(yy) * W + xx;
#pragma unroll
      for (int c = 0; c < 4; ++c) f[r][4 * h + c] = (ok && c < C) ? __ldg(px + c * HW) : 0.0f;
    }
  }
#pragma unroll
  for (int r = 0; r < ROWS; ++r) {
    if (y0 + r >= H) break;
    __align__(16) T v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = from_f32<T>(scale * f[r][k]);
    *reinterpret_cast<uint4*>(col + ((static_cast<long long>(b) * H + y0 + r) * W + x) * 64 + 8 * j) =
        *reinterpret_cast<const uint4*>(v);
  }
}

template <typename T>
__global__ void nhwc_to_nchw_kernel(const T* __restrict__ src, float* __restrict__ dst, int B, int C, int HW,
                                    int Cpad) {
  const long long total = static_cast<long long>(B) * HW;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int b = static_cast<int>(i / HW);
    const int p = static_cast<int>(i - static_cast<long long>(b) * HW);
    const T* s = src + i * Cpad;
    float* d = dst + (static_cast<long long>(b) * C) * HW + p;
    constexpr int V = Vec16<T>::N;
    for (int c0 = 0; c0 < C; c0 += V) {
      T tmp[V];
      *reinterpret_cast<uint4*>(tmp) = *reinterpret_cast<const uint4*>(s + c0);
#pragma unroll
      for (int j = 0; j < V; ++j)
        if (c0 + j < C) d[static_cast<long long>(c0 + j) * HW] = to_f32<T>(tmp[j]);
    }
  }
}

// ---------------------------------------------------------------------------------------------
// min / max of an fp32 tensor (two-stage, deterministic)
// ---------------------------------------------------------------------------------------------
__global__ void minmax_stage1(const float* __restrict__ src, long long n, float* __restrict__ part) {
  float mn = INFINITY, mx = -INFINITY;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float v = __ldg(src + i);
    mn = fminf(mn, v);
    mx = fmaxf(mx, v);
  }
  __shared__ float smn[32], smx[32];
  mn = warp_min(mn);
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) {
    smn[threadIdx.x >> 5] = mn;
    smx[threadIdx.x >> 5] = mx;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    const int nw = blockDim.x >> 5;
    mn = threadIdx.x < nw ? smn[threadIdx.x] : INFINITY;
    mx = threadIdx.x < nw ? smx[threadIdx.x] : -INFINITY;
    mn = warp_min(mn);
    mx = warp_max(mx);
    if (threadIdx.x == 0) {
      part[2 * blockIdx.x] = mn;
      part[2 * blockIdx.x + 1] = mx;
    }
  }
}
__global__ void minmax_stage2(const float* __restrict__ part, int nblk, float* __restrict__ out2) {
  float mn = INFINITY, mx = -INFINITY;
  for (int i = threadIdx.x; i < nblk; i += blockDim.x) {
    mn = fminf(mn, part[2 * i]);
    mx = fmaxf(mx, part[2 * i + 1]);
  }
  __shared__ float smn[32], smx[32];
  mn = warp_min(mn);
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) {
    smn[threadIdx.x >> 5] = mn;
    smx[threadIdx.x >> 5] = mx;
  }
  __syncthreads();
  if (threadIdx.x < 32) {
    const int nw = blockDim.x >> 5;
    mn = threadIdx.x < nw ? smn[threadIdx.x] : INFINITY;
    mx = threadIdx.x < nw ? smx[threadIdx.x] : -INFINITY;
    mn = warp_min(mn);
    mx = warp_max(mx);
    if (threadIdx.x == 0) {
      out2[0] = mn;
      out2[1] = mx;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Conv weight packing from the fp32 OIHW master
// ---------------------------------------------------------------------------------------------
// which=0: P[co][tap][ci] = w[co][ci][r][s]
// which=1 (s1 / 1x1): P[ci][tap'][co] = w[co][ci][kh-1-r'][kw-1-s']
// which=1 (s2): 4 parity-phase matrices [CinP][nt*CoutP], phases (py,px) in order 00,01,10,11,
//               taps of a phase in (r,s) ascending order: entry = w[co][ci][r][s]
// packed element e of one conv -> its value from the OIHW master (0 in the padding)
__device__ __forceinline__ float pack_conv_weight_value(const float* __restrict__ w, long long e, int Cout, int Cin,
                                                        int k, int CoutP, int CinP, int which, int s2) {
  const int taps = k * k;
  {
    int co, ci, r, s;
    if (which == 0) {
      ci = static_cast<int>(e % CinP);
      long long q = e / CinP;
      int tap = static_cast<int>(q % taps);
      co = static_cast<int>(q / taps);
      r = tap / k;
      s = tap % k;
    } else if (!s2) {
      co = static_cast<int>(e % CoutP);
      long long q = e / CoutP;
      int tap = static_cast<int>(q % taps);
      ci = static_cast<int>(q / taps);
      r = k - 1 - tap / k;
      s = k - 1 - tap % k;
    } else {
      // phase sizes: nt = 4,2,2,1
      const long long unit = static_cast<long long>(CinP) * CoutP;
      long long rem = e;
      int py = 0, px = 0, nt = 4;
      const int nts[4] = {4, 2, 2, 1};
      int ph = 0;
      for (; ph < 4; ++ph) {
        if (rem < nts[ph] * unit) break;
        rem -= nts[ph] * unit;
      }
      py = ph >> 1;
      px = ph & 1;
      nt = nts[ph];
      co = static_cast<int>(rem % CoutP);
      long long q = rem / CoutP;
      int t = static_cast<int>(q % nt);
      ci = static_cast<int>(q / nt);
      // t-th tap with (r&1)==py, (s&1)==px in ascending (r,s)
      const int ns = px == 0 ? 2 : 1;
      const int ri = t / ns, si = t % ns;
      r = py == 0 ? 2 * ri : 1;
      s = px == 0 ? 2 * si : 1;
    }
    float v = 0.0f;
    if (co < Cout && ci < Cin) v = __ldg(w + ((static_cast<long long>(co) * Cin + ci) * k + r) * k + s);
    return v;
  }
}

template <typename T>
__global__ void pack_conv_weight_kernel(const float* __restrict__ w, T* __restrict__ P, int Cout, int Cin, int k,
                                        int CoutP, int CinP, int which, int s2) {
  const long long total = static_cast<long long>(CoutP) * CinP * k * k;
  for (long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<long long>(gridDim.x) * blockDim.x)
    P[e] = from_f32<T>(pack_conv_weight_value(w, e, Cout, Cin, k, CoutP, CinP, which, s2));
}

// Multi-tensor variant: ONE launch re-packs every conv weight of the model after the optimizer step (188 packed copies
// in the train step; as separate launches they cost 1.7 ms of 5-10 us kernels, ncu launch list s10).
// table row (int64 x PACK_MT_COLS): w ptr, P ptr, Cout, Cin, k, CoutP, CinP, which | s2 << 1, dtype code, first block.
// Block b serves PACK_MT_ELEMS consecutive packed elements of the entry whose block range contains b.
static constexpr int PACK_MT_COLS = 10;
static constexpr int PACK_MT_ELEMS = 4096;
__global__ void __launch_bounds__(256)
pack_conv_weight_mt_kernel(const long long* __restrict__ table, int n_entries) {
  int lo = 0, hi = n_entries - 1;
  const long long blk = blockIdx.x;
  while (lo < hi) {  // last entry with first_block <= blk
    const int mid = (lo + hi + 1) >> 1;
    if (__ldg(table + static_cast<long long>(mid) * PACK_MT_COLS + 9) <= blk) lo = mid; else hi = mid - 1;
  }
  const long long* row = table + static_cast<long long>(lo) * PACK_MT_COLS;
  const float* w = reinterpret_cast<const float*>(__ldg(row + 0));
  void* P = reinterpret_cast<void*>(__ldg(row + 1));
  const int Cout = static_cast<int>(__ldg(row + 2)), Cin = static_cast<int>(__ldg(row + 3));
  const int k = static_cast<int>(__ldg(row + 4)), CoutP = static_cast<int>(__ldg(row + 5));
  const int CinP = static_cast<int>(__ldg(row + 6)), flags = static_cast<int>(__ldg(row + 7));
  const int dtype = static_cast<int>(__ldg(row + 8));
  const long long total = static_cast<long long>(CoutP) * CinP * k * k;
  const long long e0 = (blk - __ldg(row + 9)) * PACK_MT_ELEMS;
#pragma unroll 4
  for (int i = threadIdx.x; i < PACK_MT_ELEMS; i += 256) {
    const long long e = e0 + i;
    if (e >= total) break;
    const float v = pack_conv_weight_value(w, e, Cout, Cin, k, CoutP, CinP, flags & 1, (flags >> 1) & 1);
    if (dtype == ODVAE_BF16) reinterpret_cast<bf16*>(P)[e] = from_f32<bf16>(v);
    else if (dtype == ODVAE_F16) reinterpret_cast<f16*>(P)[e] = from_f32<f16>(v);
    else reinterpret_cast<float*>(P)[e] = v;
  }
}

// ---------------------------------------------------------------------------------------------
// nearest-2x upsample + 3x3 conv as parity-phase 2x2 convs (engine_host.cu describe_upconv_*): weight pre-sums and the fold of
// the 16 (phase, tap) weight-gradient matrices back onto the 3x3 filter. r -> a(py, r): py = 0: {0 -> 0; 1, 2 -> 1},
// py = 1: {0, 1 -> 0; 2 -> 1} (which low-resolution row the upsampled row 2i + py + r - 1 falls on).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ int up_tap_of(int parity, int r) { return parity == 0 ? (r == 0 ? 0 : 1) : (r == 2 ? 1 : 0); }

__device__ __forceinline__ float upconv_presum(const float* __restrict__ w, int co, int ci, int Cin, int py, int px, int a,
                                               int b) {
  float v = 0.0f;
  for (int r = 0; r < 3; ++r) {
    if (up_tap_of(py, r) != a) continue;
    for (int s = 0; s < 3; ++s)
      if (up_tap_of(px, s) == b) v += __ldg(w + (static_cast<long long>(co) * Cin + ci) * 9 + r * 3 + s);
  }
  return v;
}

// which = 0: fwd [phase][CoutP][tap = a*2+b][CinP];  which = 1: dgrad [CinP][t16 = ((py*2+a)*2+px)*2+b][CoutP]
template <typename T>
__global__ void pack_upconv_weight_kernel(const float* __restrict__ w, T* __restrict__ P, int Cout, int Cin, int CoutP,
                                          int CinP, int which) {
  const long long total = 16ll * CoutP * CinP;
  for (long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<long long>(gridDim.x) * blockDim.x) {
    int co, ci, py, px, a, b;
    if (which == 0) {
      ci = static_cast<int>(e % CinP);
      long long q = e / CinP;
      const int t = static_cast<int>(q & 3);
      q >>= 2;
      co = static_cast<int>(q % CoutP);
      const int ph = static_cast<int>(q / CoutP);
      py = ph >> 1; px = ph & 1; a = t >> 1; b = t & 1;
    } else {
      co = static_cast<int>(e % CoutP);
      long long q = e / CoutP;
      const int t = static_cast<int>(q & 15);
      ci = static_cast<int>(q >> 4);
      py = t >> 3; a = (t >> 2) & 1; px = (t >> 1) & 1; b = t & 1;
    }
    const float v = (co < Cout && ci < Cin) ? upconv_presum(w, co, ci, Cin, py, px, a, b) : 0.0f;
    P[e] = from_f32<T>(v);
  }
}

// dwp: [4 phases][Cout][Cin][4 taps] (wgrad_reduce layout with taps = 4); dbp: [4][Cout] or null
__global__ void upconv_wgrad_fold_kernel(const float* __restrict__ dwp, const float* __restrict__ dbp, float* __restrict__ dw,
                                         float* __restrict__ db, int Cout, int Cin, int accumulate, float scale) {
  const long long total = static_cast<long long>(Cout) * Cin * 9;
  const long long slab = static_cast<long long>(Cout) * Cin * 4;
  for (long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int rs = static_cast<int>(e % 9);
    const long long oc = e / 9;
    const int r = rs / 3, s = rs - r * 3;
    float acc = 0.0f;
#pragma unroll
    for (int ph = 0; ph < 4; ++ph)
      acc += dwp[ph * slab + oc * 4 + up_tap_of(ph >> 1, r) * 2 + up_tap_of(ph & 1, s)];
    acc *= scale;
    dw[e] = accumulate ? dw[e] + acc : acc;
  }
  if (db != nullptr && dbp != nullptr) {
    const long long c = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    if (c < Cout) {
      const float t = ((dbp[c] + dbp[Cout + c]) + (dbp[2 * Cout + c] + dbp[3 * Cout + c])) * scale;
      db[c] = accumulate ? db[c] + t : t;
    }
  }
}

// dw[co][ci][tap] (+)= sum_split ws[(tap*split_k + split)][co][ci]   (ws rows: CoutP x CinP)
// One thread sums VEC consecutive ci over all splits, 8 independent (vector) loads in flight, added in split order
// (the order is part of the bit-stability contract). The workspace was written by the preceding weight-gradient kernel
// and is mostly L2-resident, so the kernel is latency-bound: wide loads and deep unrolling are what shorten it.
// The bias gradient rides along: the weight-gradient kernel leaves `cs_rows` partial column sums per output channel
// (`cpart`, see TcParams::colsum); the first Cout threads of the grid add them up in row order (8 loads in flight).
template <int VEC>
__global__ void wgrad_reduce_kernel(const float* __restrict__ ws, float* __restrict__ dw, int Cout, int Cin,
                                    int taps, int CoutP, int CinP, int split_k, int accumulate,
                                    const float* __restrict__ cpart, int cs_rows, float* __restrict__ db, float scale) {
  if (cpart != nullptr) {
    const long long c = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    if (c < Cout) {
      float t = 0.0f;
      int r = 0;
      for (; r + 7 < cs_rows; r += 8) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = __ldg(cpart + static_cast<long long>(r + u) * Cout + c);
#pragma unroll
        for (int u = 0; u < 8; ++u) t += v[u];
      }
      for (; r < cs_rows; ++r) t += __ldg(cpart + static_cast<long long>(r) * Cout + c);
      t *= scale;
      db[c] = accumulate ? db[c] + t : t;
    }
  }
  const int cin_v = Cin / VEC;
  const long long total = static_cast<long long>(Cout) * cin_v * taps;
  const long long sstride = static_cast<long long>(CoutP) * CinP;
  for (long long e = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; e < total;
       e += static_cast<long long>(gridDim.x) * blockDim.x) {
    // iterate with ci fastest for coalesced workspace reads
    const int ci = static_cast<int>(e % cin_v) * VEC;
    long long q = e / cin_v;
    const int tap = static_cast<int>(q % taps);
    const int co = static_cast<int>(q / taps);
    float acc[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) acc[j] = 0.0f;
    const float* src = ws + (static_cast<long long>(tap) * split_k * CoutP + co) * CinP + ci;
    constexpr int U = 8;
    int s = 0;
    for (; s + U - 1 < split_k; s += U) {
      float v[U][VEC];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (VEC == 4) *reinterpret_cast<float4*>(v[u]) = __ldg(reinterpret_cast<const float4*>(src + (s + u) * sstride));
        else v[u][0] = __ldg(src + (s + u) * sstride);
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int j = 0; j < VEC; ++j) acc[j] += v[u][j];
    }
    for (; s < split_k; ++s) {
#pragma unroll
      for (int j = 0; j < VEC; ++j) acc[j] += __ldg(src + s * sstride + j);
    }
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      float* d = dw + (static_cast<long long>(co) * Cin + ci + j) * taps + tap;
      *d = accumulate ? (*d + acc[j] * scale) : acc[j] * scale;
    }
  }
}

// column sums of a [rows][ld] matrix (bias gradients): stage 1 partials [nblk][cols], stage 2 final
template <typename T>
__global__ void colsum_stage1(const T* __restrict__ m, long long rows, int cols, int ld, float* __restrict__ part) {
  // blockDim = (32, 8): x over columns (strided), y over rows
  const long long rows_per_blk = (rows + gridDim.x - 1) / gridDim.x;
  const long long r0 = blockIdx.x * rows_per_blk;
  const long long r1 = r0 + rows_per_blk < rows ? r0 + rows_per_blk : rows;
  __shared__ float sm[8][33];
  for (int c0 = 0; c0 < cols; c0 += 32) {
    const int c = c0 + threadIdx.x;
    float acc = 0.0f;
    if (c < cols)
      for (long long r = r0 + threadIdx.y; r < r1; r += 8) acc += to_f32<T>(m[r * ld + c]);
    sm[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && c < cols) {
      float t = 0.0f;
#pragma unroll
      for (int j = 0; j < 8; ++j) t += sm[j][threadIdx.x];
      part[static_cast<long long>(blockIdx.x) * cols + c] = t;
    }
    __syncthreads();
  }
}
// vectorised variant: every thread streams 16-byte vectors of whole rows (ld % V == 0, ld / V <= 256); the block's
// row lanes are combined through shared memory in a fixed order. Columns >= cols (channel padding) are dropped.
template <typename T, int V>
__global__ void __launch_bounds__(256) colsum_stage1_vec(const T* __restrict__ m, long long rows, int cols, int ld,
                                                          float* __restrict__ part) {
  const int nvec = ld / V;
  const int lanes = 256 / nvec;  // row lanes per block
  const int cv = threadIdx.x % nvec, lane = threadIdx.x / nvec;
  const long long rows_per_blk = (rows + gridDim.x - 1) / gridDim.x;
  const long long r0 = blockIdx.x * rows_per_blk;
  const long long r1 = r0 + rows_per_blk < rows ? r0 + rows_per_blk : rows;
  float acc[V];
#pragma unroll
  for (int j = 0; j < V; ++j) acc[j] = 0.0f;
  if (lane < lanes) {
    const T* base = m + static_cast<long long>(cv) * V;
    long long r = r0 + lane;
    for (; r + 3ll * lanes < r1; r += 4ll * lanes) {
      T v0[V], v1[V], v2[V], v3[V];
      *reinterpret_cast<uint4*>(v0) = __ldg(reinterpret_cast<const uint4*>(base + r * ld));
      *reinterpret_cast<uint4*>(v1) = __ldg(reinterpret_cast<const uint4*>(base + (r + lanes) * ld));
      *reinterpret_cast<uint4*>(v2) = __ldg(reinterpret_cast<const uint4*>(base + (r + 2ll * lanes) * ld));
      *reinterpret_cast<uint4*>(v3) = __ldg(reinterpret_cast<const uint4*>(base + (r + 3ll * lanes) * ld));
#pragma unroll
      for (int j = 0; j < V; ++j)
        acc[j] += (to_f32<T>(v0[j]) + to_f32<T>(v1[j])) + (to_f32<T>(v2[j]) + to_f32<T>(v3[j]));
    }
    for (; r < r1; r += lanes) {
      T v0[V];
      *reinterpret_cast<uint4*>(v0) = __ldg(reinterpret_cast<const uint4*>(base + r * ld));
#pragma unroll
      for (int j = 0; j < V; ++j) acc[j] += to_f32<T>(v0[j]);
    }
  }
  extern __shared__ float csm[];  // [lanes][ld]
  if (lane < lanes) {
#pragma unroll
    for (int j = 0; j < V; ++j) csm[lane * ld + cv * V + j] = acc[j];
  }
  __syncthreads();
  for (int c = threadIdx.x; c < cols; c += 256) {
    float t = 0.0f;
    for (int l = 0; l < lanes; ++l) t += csm[l * ld + c];
    part[static_cast<long long>(blockIdx.x) * cols + c] = t;
  }
}
// blockDim = (32, 8): x over columns, y over 8 row lanes of the partial list; lanes are combined in a fixed order
__global__ void colsum_stage2(const float* __restrict__ part, int nblk, int cols, float* __restrict__ out,
                              int accumulate, float scale = 1.0f) {
  const int c = blockIdx.x * 32 + threadIdx.x;
  __shared__ float sm[8][33];
  float t = 0.0f;
  if (c < cols) {
    int i = threadIdx.y;
    for (; i + 24 < nblk; i += 32) {  // four independent loads in flight, added in index order
      const float a0 = __ldg(part + static_cast<long long>(i) * cols + c);
      const float a1 = __ldg(part + static_cast<long long>(i + 8) * cols + c);
      const float a2 = __ldg(part + static_cast<long long>(i + 16) * cols + c);
      const float a3 = __ldg(part + static_cast<long long>(i + 24) * cols + c);
      t += a0; t += a1; t += a2; t += a3;
    }
    for (; i < nblk; i += 8) t += __ldg(part + static_cast<long long>(i) * cols + c);
  }
  sm[threadIdx.y][threadIdx.x] = t;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
    float r = 0.0f;
#pragma unroll
    for (int j = 0; j < 8; ++j) r += sm[j][threadIdx.x];
    r *= scale;
    out[c] = accumulate ? out[c] + r : r;
  }
}

// ---------------------------------------------------------------------------------------------
// nearest 2x upsample and adjoint, NHWC
// ---------------------------------------------------------------------------------------------
// one thread per INPUT vector: 1 load, 4 stores (the 2x2 block), 32-bit index math - the per-output-vector form spent
// its time in 64-bit div/mod chains (2.7 TB/s in the s10 launch list)
template <typename T>
__global__ void upsample2x_kernel(const T* __restrict__ x, T* __restrict__ y, int B, int H, int W, int C) {
  constexpr int V = Vec16<T>::N;
  const int cv = C / V;
  const long long rows = static_cast<long long>(B) * H;  // input rows
  const int row_vecs = W * cv;                           // vectors per input row
  const uint4* xv = reinterpret_cast<const uint4*>(x);
  uint4* yv = reinterpret_cast<uint4*>(y);
  for (long long r = blockIdx.y; r < rows; r += gridDim.y) {
    const uint4* src = xv + r * row_vecs;
    // input row r = (b, yi) -> output rows 2*(b*H + yi) and +1 of the [B*2H][2W][cv] output
    uint4* d0 = yv + (2 * r) * (2LL * row_vecs);
    uint4* d1 = d0 + 2LL * row_vecs;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < row_vecs; i += gridDim.x * blockDim.x) {
      const int xi = i / cv, c = i - xi * cv;
      const uint4 v = __ldg(src + i);
      const int o = (2 * xi) * cv + c;
      d0[o] = v;
      d0[o + cv] = v;
      d1[o] = v;
      d1[o + cv] = v;
    }
  }
}
template <typename T>
__global__ void upsample2x_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, int B, int H, int W, int C) {
  constexpr int V = Vec16<T>::N;
  const int cv = C / V;
  const long long total = static_cast<long long>(B) * H * W * cv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % cv);
    long long q = i / cv;
    const int xi = static_cast<int>(q % W);
    q /= W;
    const int yi = static_cast<int>(q % H);
    const int b = static_cast<int>(q / H);
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.0f;
#pragma unroll
    for (int dyy = 0; dyy < 2; ++dyy)
#pragma unroll
      for (int dxx = 0; dxx < 2; ++dxx) {
        T tmp[V];
        *reinterpret_cast<uint4*>(tmp) = __ldg(reinterpret_cast<const uint4*>(
            dy + ((static_cast<long long>(b) * 2 * H + 2 * yi + dyy) * 2 * W + 2 * xi + dxx) * C) + c);
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] += to_f32<T>(tmp[j]);
      }
    T out[V];
#pragma unroll
    for (int j = 0; j < V; ++j) out[j] = from_f32<T>(acc[j]);
    reinterpret_cast<uint4*>(dx)[i] = *reinterpret_cast<const uint4*>(out);
  }
}

template <typename T>
__global__ void add_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ o, long long nvec) {
  constexpr int V = Vec16<T>::N;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    T ta[V], tb[V], to[V];
    *reinterpret_cast<uint4*>(ta) = __ldg(reinterpret_cast<const uint4*>(a) + i);
    *reinterpret_cast<uint4*>(tb) = __ldg(reinterpret_cast<const uint4*>(b) + i);
#pragma unroll
    for (int j = 0; j < V; ++j) to[j] = from_f32<T>(to_f32<T>(ta[j]) + to_f32<T>(tb[j]));
    reinterpret_cast<uint4*>(o)[i] = *reinterpret_cast<const uint4*>(to);
  }
}

// fp16 <-> bf16 over 16-byte vectors (SRC_HALF: fp16 -> bf16, else bf16 -> fp16 with saturation)
template <bool SRC_HALF>
__global__ void __launch_bounds__(256) cast16_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, long long nvec) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nvec;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const uint4 v = __ldg(src + i);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_16x2(w[j], SRC_HALF);
      o[j] = pack_16x2(f.x, f.y, !SRC_HALF);
    }
    dst[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

static int grid_for(long long work_items, int threads, int per_sm = 8) {
  long long blocks = (work_items + threads - 1) / threads;
  long long cap = static_cast<long long>(num_sms()) * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

// internal entry points used by api.cu ----------------------------------------------------------
int launch_wgrad_reduce(const float* ws, float* dw, int Cout, int Cin, int taps, int CoutP, int CinP, int split_k,
                        int accumulate, cudaStream_t st, const float* cpart, int cs_rows, float* db, float scale) {
  const long long total = static_cast<long long>(Cout) * Cin * taps;
  if (Cin % 4 == 0 && CinP % 4 == 0 && (reinterpret_cast<uintptr_t>(ws) & 15) == 0) {
    int grid = grid_for(total / 4, 128);
    if (cpart != nullptr && static_cast<long long>(grid) * 128 < Cout) grid = (Cout + 127) / 128;
    wgrad_reduce_kernel<4><<<grid, 128, 0, st>>>(ws, dw, Cout, Cin, taps, CoutP, CinP, split_k, accumulate, cpart, cs_rows, db, scale);
  } else {
    int grid = grid_for(total, 256);
    if (cpart != nullptr && static_cast<long long>(grid) * 256 < Cout) grid = (Cout + 255) / 256;
    wgrad_reduce_kernel<1><<<grid, 256, 0, st>>>(ws, dw, Cout, Cin, taps, CoutP, CinP, split_k, accumulate, cpart, cs_rows, db, scale);
  }
  return check_launch("wgrad_reduce_kernel");
}

int launch_upconv_wgrad_fold(const float* dwp, const float* dbp, float* dw, float* db, int Cout, int Cin, int accumulate,
                             float scale, cudaStream_t st) {
  long long total = static_cast<long long>(Cout) * Cin * 9;
  int grid = grid_for(total, 256);
  if (static_cast<long long>(grid) * 256 < Cout) grid = (Cout + 255) / 256;
  upconv_wgrad_fold_kernel<<<grid, 256, 0, st>>>(dwp, dbp, dw, db, Cout, Cin, accumulate, scale);
  return check_launch("upconv_wgrad_fold_kernel");
}

// final stage alone: out[c] (+)= sum_r part[r][c] -- the tcgen05 weight-gradient kernel writes the partials itself
int launch_colsum_final(const float* part, int nrows, int cols, float* out, int accumulate, cudaStream_t st) {
  colsum_stage2<<<(cols + 31) / 32, dim3(32, 8), 0, st>>>(part, nrows, cols, out, accumulate);
  return check_launch("colsum_stage2");
}

static constexpr int COLSUM_BLOCKS = 592;
size_t colsum_workspace_bytes(int cols) { return static_cast<size_t>(COLSUM_BLOCKS) * cols * sizeof(float); }
int launch_colsum(const void* m, long long rows, int cols, int ld, int act_dtype, float* out, int accumulate,
                  float* part, cudaStream_t st, float scale) {
  if (act_dtype != ODVAE_BF16 && act_dtype != ODVAE_F32 && act_dtype != ODVAE_F16) return ODVAE_ERR_SHAPE;
  int nblk = COLSUM_BLOCKS;
  if (rows < nblk * 8) nblk = static_cast<int>((rows + 7) / 8);
  if (nblk < 1) nblk = 1;
  const int V = act_dtype == ODVAE_F32 ? 4 : 8;
  const bool vec = ld % V == 0 && ld / V <= 256 && 256 % (ld / V) == 0 && (reinterpret_cast<uintptr_t>(m) & 15) == 0 &&
                   static_cast<size_t>(256 / (ld / V)) * ld * sizeof(float) <= 48 * 1024;
  if (vec) {
    const size_t smem = static_cast<size_t>(256 / (ld / V)) * ld * sizeof(float);
    if (act_dtype == ODVAE_BF16)
      colsum_stage1_vec<bf16, 8><<<nblk, 256, smem, st>>>(reinterpret_cast<const bf16*>(m), rows, cols, ld, part);
    else if (act_dtype == ODVAE_F16)
      colsum_stage1_vec<f16, 8><<<nblk, 256, smem, st>>>(reinterpret_cast<const f16*>(m), rows, cols, ld, part);
    else
      colsum_stage1_vec<float, 4><<<nblk, 256, smem, st>>>(reinterpret_cast<const float*>(m), rows, cols, ld, part);
  } else {
    dim3 blk(32, 8);
    if (act_dtype == ODVAE_BF16)
      colsum_stage1<bf16><<<nblk, blk, 0, st>>>(reinterpret_cast<const bf16*>(m), rows, cols, ld, part);
    else if (act_dtype == ODVAE_F16)
      colsum_stage1<f16><<<nblk, blk, 0, st>>>(reinterpret_cast<const f16*>(m), rows, cols, ld, part);
    else
      colsum_stage1<float><<<nblk, blk, 0, st>>>(reinterpret_cast<const float*>(m), rows, cols, ld, part);
  }
  int s = check_launch("colsum_stage1");
  if (s != ODVAE_OK) return s;
  colsum_stage2<<<(cols + 31) / 32, dim3(32, 8), 0, st>>>(part, nblk, cols, out, accumulate, scale);
  return check_launch("colsum_stage2");
}

// ---------------------------------------------------------------------------------------------
// Column sums without any allocation (odvae_colsum_static): the stage-1 partials live in a file-scope device array instead
// of a caller-provided workspace, for callers that cannot allocate (CUDA-graph capture, serving loops). Every GPU that loads
// the library holds the array (COLSUM_BLOCKS x ODVAE_COLSUM_STATIC_MAX_COLS floats = 1.16 MiB) until exit. In host code
// `g_colsum_static_scratch` is only the symbol's host-side stand-in: the device address comes from cudaGetSymbolAddress,
// looked up once per device with that device current.
// ---------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_colsum_static_scratch[COLSUM_BLOCKS * ODVAE_COLSUM_STATIC_MAX_COLS];

static constexpr int STATIC_MAX_DEVICES = 64;
static float* g_colsum_static_addr[STATIC_MAX_DEVICES] = {};
static std::mutex g_colsum_static_mutex;

// device address of the scratch array on the CURRENT device (nullptr + last error set on failure)
static float* colsum_static_scratch() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= STATIC_MAX_DEVICES) {
    set_last_cuda_error(e != cudaSuccess ? e : cudaErrorInvalidDevice, "odvae_colsum_static: cudaGetDevice");
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(g_colsum_static_mutex);
  if (g_colsum_static_addr[dev] == nullptr) {
    float* ptr = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_colsum_static_scratch);
    if (e != cudaSuccess || ptr == nullptr) {
      set_last_cuda_error(e != cudaSuccess ? e : cudaErrorInvalidSymbol,
                          "odvae_colsum_static: cudaGetSymbolAddress(g_colsum_static_scratch)");
      return nullptr;
    }
    g_colsum_static_addr[dev] = ptr;
  }
  return g_colsum_static_addr[dev];
}
}  // namespace odvae

using namespace odvae;

extern "C" {

int odvae_nchw_to_nhwc(const float* src, void* dst, int B, int C, int H, int W, int Cpad, int act_dtype, float scale,
                       float shift, void* stream) {
  if (B <= 0 || C <= 0 || Cpad < C) return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long total = static_cast<long long>(B) * H * W;
  if (act_dtype == ODVAE_BF16) {
    if (Cpad % 8) return ODVAE_ERR_SHAPE;
    nchw_to_nhwc_kernel<bf16><<<grid_for(total, 256), 256, 0, st>>>(src, reinterpret_cast<bf16*>(dst), B, C, H * W, Cpad, scale, shift);
  } else if (act_dtype == ODVAE_F16) {
    if (Cpad % 8) return ODVAE_ERR_SHAPE;
    nchw_to_nhwc_kernel<f16><<<grid_for(total, 256), 256, 0, st>>>(src, reinterpret_cast<f16*>(dst), B, C, H * W, Cpad, scale, shift);
  } else {
    if (Cpad % 4) return ODVAE_ERR_SHAPE;
    nchw_to_nhwc_kernel<float><<<grid_for(total, 256), 256, 0, st>>>(src, reinterpret_cast<float*>(dst), B, C, H * W, Cpad, scale, shift);
  }
  return check_launch("nchw_to_nhwc_kernel");
}

int odvae_rescale_nhwc(const float* src, const float* minmax, float* dst_nchw, void* dst_nhwc, int B, int C, int H, int W,
                       int Cpad, int act_dtype, void* stream) {
  if (B <= 0 || C <= 0 || Cpad < C) return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long total = static_cast<long long>(B) * H * W;
  if (act_dtype == ODVAE_BF16) {
    if (Cpad % 8) return ODVAE_ERR_SHAPE;
    rescale_nhwc_kernel<bf16><<<grid_for(total, 256), 256, 0, st>>>(src, minmax, dst_nchw, reinterpret_cast<bf16*>(dst_nhwc), B, C, H * W, Cpad);
  } else if (act_dtype == ODVAE_F16) {
    if (Cpad % 8) return ODVAE_ERR_SHAPE;
    rescale_nhwc_kernel<f16><<<grid_for(total, 256), 256, 0, st>>>(src, minmax, dst_nchw, reinterpret_cast<f16*>(dst_nhwc), B, C, H * W, Cpad);
  } else {
    if (Cpad % 4) return ODVAE_ERR_SHAPE;
    rescale_nhwc_kernel<float><<<grid_for(total, 256), 256, 0, st>>>(src, minmax, dst_nchw, reinterpret_cast<float*>(dst_nhwc), B, C, H * W, Cpad);
  }
  return check_launch("rescale_nhwc_kernel");
}

int odvae_im2col3x3_thin(const float* src, void* col, int B, int C, int H, int W, int flip, float scale, int act_dtype,
                         void* stream) {
  if (B <= 0 || C <= 0 || C > 4 || H <= 0 || W <= 0 || H > 4 * 65535 || B > 65535) return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int sgn = flip ? -1 : 1;
  const dim3 grid(static_cast<unsigned>((static_cast<long long>(W) * 8 + 255) / 256), static_cast<unsigned>((H + 3) / 4),
                  static_cast<unsigned>(B));
  if (act_dtype == ODVAE_BF16)
    im2col3_thin_kernel<bf16><<<grid, 256, 0, st>>>(src, reinterpret_cast<bf16*>(col), C, H, W, sgn, scale);
  else if (act_dtype == ODVAE_F16)
    im2col3_thin_kernel<f16><<<grid, 256, 0, st>>>(src, reinterpret_cast<f16*>(col), C, H, W, sgn, scale);
  else
    return ODVAE_ERR_SHAPE;
  return check_launch("im2col3_thin_kernel");
}

int odvae_nhwc_to_nchw(const void* src, float* dst, int B, int C, int H, int W, int Cpad, int act_dtype, void* stream) {
  if (B <= 0 || C <= 0 || Cpad < C) return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long total = static_cast<long long>(B) * H * W;
  if (act_dtype == ODVAE_BF16) {
    if (Cpad % 8) return ODVAE_ERR_SHAPE;
    nhwc_to_nchw_kernel<bf16><<<grid_for(total, 256), 256, 0, st>>>(reinterpret_cast<const bf16*>(src), dst, B, C, H * W, Cpad);
  } else if (act_dtype == ODVAE_F16) {
    if (Cpad % 8) return ODVAE_ERR_SHAPE;
    nhwc_to_nchw_kernel<f16><<<grid_for(total, 256), 256, 0, st>>>(reinterpret_cast<const f16*>(src), dst, B, C, H * W, Cpad);
  } else {
    if (Cpad % 4) return ODVAE_ERR_SHAPE;
    nhwc_to_nchw_kernel<float><<<grid_for(total, 256), 256, 0, st>>>(reinterpret_cast<const float*>(src), dst, B, C, H * W, Cpad);
  }
  return check_launch("nhwc_to_nchw_kernel");
}

int odvae_minmax(const float* src, int64_t n, float* out2, void* workspace, size_t ws_bytes, void* stream) {
  const int nblk = 2 * num_sms();
  if (ws_bytes < static_cast<size_t>(nblk) * 2 * sizeof(float)) return ODVAE_ERR_WORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  minmax_stage1<<<nblk, 512, 0, st>>>(src, n, reinterpret_cast<float*>(workspace));
  int s = check_launch("minmax_stage1");
  if (s != ODVAE_OK) return s;
  minmax_stage2<<<1, 256, 0, st>>>(reinterpret_cast<const float*>(workspace), nblk, out2);
  return check_launch("minmax_stage2");
}

int odvae_pack_conv_weight(const float* w_oihw, void* packed, int Cout, int Cin, int geom, int CoutP, int CinP,
                           int which, int act_dtype, void* stream) {
  if (CoutP < Cout || CinP < Cin) return ODVAE_ERR_SHAPE;
  const int k = geom == ODVAE_CONV_1x1 ? 1 : 3;
  const int s2 = (geom == ODVAE_CONV_3x3_S2 && which == 1) ? 1 : 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long total = static_cast<long long>(CoutP) * CinP * k * k;
  if (act_dtype == ODVAE_BF16)
    pack_conv_weight_kernel<bf16><<<grid_for(total, 256), 256, 0, st>>>(w_oihw, reinterpret_cast<bf16*>(packed), Cout, Cin, k, CoutP, CinP, which, s2);
  else if (act_dtype == ODVAE_F16)
    pack_conv_weight_kernel<f16><<<grid_for(total, 256), 256, 0, st>>>(w_oihw, reinterpret_cast<f16*>(packed), Cout, Cin, k, CoutP, CinP, which, s2);
  else
    pack_conv_weight_kernel<float><<<grid_for(total, 256), 256, 0, st>>>(w_oihw, reinterpret_cast<float*>(packed), Cout, Cin, k, CoutP, CinP, which, s2);
  return check_launch("pack_conv_weight_kernel");
}

int odvae_pack_upconv_weight(const float* w_oihw, void* packed, int Cout, int Cin, int CoutP, int CinP, int which,
                             int act_dtype, void* stream) {
  if (Cout <= 0 || Cin <= 0 || CoutP < Cout || CinP < Cin || (which != 0 && which != 1)) return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const long long total = 16ll * CoutP * CinP;
  if (act_dtype == ODVAE_BF16)
    pack_upconv_weight_kernel<bf16><<<grid_for(total, 256), 256, 0, st>>>(w_oihw, reinterpret_cast<bf16*>(packed), Cout, Cin, CoutP, CinP, which);
  else if (act_dtype == ODVAE_F16)
    pack_upconv_weight_kernel<f16><<<grid_for(total, 256), 256, 0, st>>>(w_oihw, reinterpret_cast<f16*>(packed), Cout, Cin, CoutP, CinP, which);
  else
    pack_upconv_weight_kernel<float><<<grid_for(total, 256), 256, 0, st>>>(w_oihw, reinterpret_cast<float*>(packed), Cout, Cin, CoutP, CinP, which);
  return check_launch("pack_upconv_weight_kernel");
}

int64_t odvae_pack_conv_weights_mt_blocks(int CoutP, int CinP, int geom) {
  const int k = geom == ODVAE_CONV_1x1 ? 1 : 3;
  const long long total = static_cast<long long>(CoutP) * CinP * k * k;
  return (total + PACK_MT_ELEMS - 1) / PACK_MT_ELEMS;
}

int odvae_pack_conv_weights_mt(const int64_t* table_dev, int n_entries, int64_t n_blocks, void* stream) {
  if (n_entries <= 0 || n_blocks <= 0) return ODVAE_OK;
  if (n_blocks > 0x7fffffffLL) return ODVAE_ERR_SHAPE;
  pack_conv_weight_mt_kernel<<<static_cast<unsigned>(n_blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const long long*>(table_dev), n_entries);
  return check_launch("pack_conv_weight_mt_kernel");
}

int odvae_upsample2x(const void* x, void* y, int B, int H, int W, int C, int act_dtype, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int V = (act_dtype == ODVAE_BF16 || act_dtype == ODVAE_F16) ? 8 : 4;
  if (C % V) return ODVAE_ERR_SHAPE;
  // grid: x over the vectors of one input row, y over (image, input row) grid-stride; ~8 CTAs per SM in total
  const int row_vecs = W * (C / V);
  const int gx = (row_vecs + 255) / 256;
  long long gy = (8LL * num_sms() + gx - 1) / gx;
  const long long rows = static_cast<long long>(B) * H;
  if (gy > rows) gy = rows;
  if (gy > 65535) gy = 65535;
  if (gy < 1) gy = 1;
  const dim3 grid(gx, static_cast<unsigned>(gy));
  if (V == 8)  // a copy: any 16-bit format
    upsample2x_kernel<bf16><<<grid, 256, 0, st>>>(reinterpret_cast<const bf16*>(x), reinterpret_cast<bf16*>(y), B, H, W, C);
  else
    upsample2x_kernel<float><<<grid, 256, 0, st>>>(reinterpret_cast<const float*>(x), reinterpret_cast<float*>(y), B, H, W, C);
  return check_launch("upsample2x_kernel");
}

int odvae_upsample2x_bwd(const void* dy, void* dx, int B, int H, int W, int C, int act_dtype, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (act_dtype == ODVAE_BF16) {
    if (C % 8) return ODVAE_ERR_SHAPE;
    const long long total = static_cast<long long>(B) * H * W * (C / 8);
    upsample2x_bwd_kernel<bf16><<<grid_for(total, 256), 256, 0, st>>>(reinterpret_cast<const bf16*>(dy), reinterpret_cast<bf16*>(dx), B, H, W, C);
  } else if (act_dtype == ODVAE_F16) {
    if (C % 8) return ODVAE_ERR_SHAPE;
    const long long total = static_cast<long long>(B) * H * W * (C / 8);
    upsample2x_bwd_kernel<f16><<<grid_for(total, 256), 256, 0, st>>>(reinterpret_cast<const f16*>(dy), reinterpret_cast<f16*>(dx), B, H, W, C);
  } else {
    if (C % 4) return ODVAE_ERR_SHAPE;
    const long long total = static_cast<long long>(B) * H * W * (C / 4);
    upsample2x_bwd_kernel<float><<<grid_for(total, 256), 256, 0, st>>>(reinterpret_cast<const float*>(dy), reinterpret_cast<float*>(dx), B, H, W, C);
  }
  return check_launch("upsample2x_bwd_kernel");
}

int odvae_cast16(const void* src, void* dst, int64_t n, int src_dtype, int dst_dtype, void* stream) {
  if (n % 8) return ODVAE_ERR_SHAPE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (src_dtype == ODVAE_F16 && dst_dtype == ODVAE_BF16)
    cast16_kernel<true><<<grid_for(n / 8, 256), 256, 0, st>>>(reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4*>(dst), n / 8);
  else if (src_dtype == ODVAE_BF16 && dst_dtype == ODVAE_F16)
    cast16_kernel<false><<<grid_for(n / 8, 256), 256, 0, st>>>(reinterpret_cast<const uint4*>(src), reinterpret_cast<uint4*>(dst), n / 8);
  else
    return ODVAE_ERR_SHAPE;
  return check_launch("cast16_kernel");
}

int odvae_add(const void* a, const void* b, void* out, int64_t n, int act_dtype, void* stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (act_dtype == ODVAE_BF16) {
    if (n % 8) return ODVAE_ERR_SHAPE;
    add_kernel<bf16><<<grid_for(n / 8, 256), 256, 0, st>>>(reinterpret_cast<const bf16*>(a), reinterpret_cast<const bf16*>(b), reinterpret_cast<bf16*>(out), n / 8);
  } else {
    if (n % 4) return ODVAE_ERR_SHAPE;
    add_kernel<float><<<grid_for(n / 4, 256), 256, 0, st>>>(reinterpret_cast<const float*>(a), reinterpret_cast<const float*>(b), reinterpret_cast<float*>(out), n / 4);
  }
  return check_launch("add_kernel");
}

int odvae_colsum_static(const void* m, int64_t rows, int cols, int ld, int dtype, float* out, int accumulate, void* stream) {
  if (rows <= 0 || cols <= 0 || ld < cols) return ODVAE_ERR_SHAPE;
  if (cols > ODVAE_COLSUM_STATIC_MAX_COLS) {
    set_last_cuda_error(cudaErrorInvalidValue,
                        "odvae_colsum_static: cols exceeds ODVAE_COLSUM_STATIC_MAX_COLS (512), the static scratch array "
                        "cannot hold the partial sums; use odvae_colsum with a workspace");
    return ODVAE_ERR_WORKSPACE;
  }
  float* part = colsum_static_scratch();  // device address of the file-scope array on the current device
  if (part == nullptr) return ODVAE_ERR_CUDA;
  // stage 1 writes every partial stage 2 reads ([nblk][cols], both derived from this call's arguments), so nothing in the
  // array has to start at zero; what an earlier call left behind is simply overwritten or never read
  return launch_colsum(m, rows, cols, ld, dtype, out, accumulate, part, static_cast<cudaStream_t>(stream), 1.0f);
}

}  // extern "C"
