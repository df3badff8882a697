// Shared helpers for the odvae_b200 CUDA library.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/odvae_b200.h"

namespace odvae {

void set_last_cuda_error(cudaError_t e, const char* where);
int check_launch(const char* where);  // returns ODVAE_OK or ODVAE_ERR_CUDA after a kernel launch
int num_sms();
unsigned long long launch_count();

#define ODVAE_CUDA_TRY(expr)                       \
  do {                                             \
    cudaError_t _e = (expr);                       \
    if (_e != cudaSuccess) {                       \
      ::odvae::set_last_cuda_error(_e, #expr);     \
      return ODVAE_ERR_CUDA;                       \
    }                                              \
  } while (0)

typedef __nv_bfloat16 bf16;
typedef __half f16;

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<bf16>(bf16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f32<f16>(f16 v) { return __half2float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
// fp16 stores saturate (|v| > 65504 -> +-65504) instead of producing inf
__device__ __forceinline__ float sat_f16(float v) { return fminf(fmaxf(v, -65504.0f), 65504.0f); }
// (one F2FP.SATFINITE instruction: |v| > 65504 and inf -> +-65504; the min/max pair of sat_f16 cost 2 more per value)
template <> __device__ __forceinline__ f16 from_f32<f16>(float v) {
  uint16_t r;
  asm("cvt.rn.satfinite.f16.f32 %0, %1;" : "=h"(r) : "f"(v));
  return __ushort_as_half(r);
}
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f32<bf16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t v) {
  __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162*>(&v);
  return __bfloat1622float2(h);
}
__device__ __forceinline__ uint32_t pack_f16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ float2 unpack_f16x2(uint32_t v) {
  __half2 h = *reinterpret_cast<__half2*>(&v);
  return __half22float2(h);
}
// runtime-selected 16-bit packing (engine epilogues): half = true -> IEEE fp16, else bf16
__device__ __forceinline__ uint32_t pack_16x2(float lo, float hi, bool half) {
  return half ? pack_f16x2(lo, hi) : pack_bf16x2(lo, hi);
}
__device__ __forceinline__ float2 unpack_16x2(uint32_t v, bool half) {
  return half ? unpack_f16x2(v) : unpack_bf16x2(v);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

__device__ __forceinline__ float silu_f(float u) { return __fdividef(u, 1.0f + __expf(-u)); }
// d/du [u * sigmoid(u)]
__device__ __forceinline__ float silu_grad_f(float u) {
  float s = __fdividef(1.0f, 1.0f + __expf(-u));
  return s * (1.0f + u * (1.0f - s));
}
// 16-bit modes: sigmoid(u) = 0.5 + 0.5 tanh(u / 2) with ONE MUFU op (tanh.approx, |rel err| < 2^-10.9, below the
// fp16/bf16 rounding of the stored result) instead of ex2 + rcp; the fp32 validation mode keeps the exact form.
__device__ __forceinline__ float fast_sigmoid(float u) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * u));
  return fmaf(t, 0.5f, 0.5f);
}
template <bool FAST> __device__ __forceinline__ float silu_t(float u) { return FAST ? u * fast_sigmoid(u) : silu_f(u); }
template <bool FAST> __device__ __forceinline__ float silu_grad_t(float u) {
  if (!FAST) return silu_grad_f(u);
  const float s = fast_sigmoid(u);
  return s * fmaf(u, 1.0f - s, 1.0f);
}

// ----------------------------------------------------------------------------------------------
// 256-bit global accesses (sm_100: LDG/STG.E.ENL2.256). A thread that owns one matrix row touches 32 contiguous bytes
// = one full L2 sector per instruction; as two 16-byte accesses every instruction moved half-filled sectors
// (ncu, attention softmax pass: 4.43 GB of L1->L2 write traffic for 2.10 GB stored). Falls back to 2 x 128 bit when the
// address is not 32-byte aligned.
// ----------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_global_256(void* ptr, const uint32_t (&o)[8]) {
  if ((reinterpret_cast<uintptr_t>(ptr) & 31u) == 0) {
    asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(ptr), "r"(o[0]), "r"(o[1]), "r"(o[2]),
                 "r"(o[3]), "r"(o[4]), "r"(o[5]), "r"(o[6]), "r"(o[7])
                 : "memory");
  } else {
    reinterpret_cast<uint4*>(ptr)[0] = make_uint4(o[0], o[1], o[2], o[3]);
    reinterpret_cast<uint4*>(ptr)[1] = make_uint4(o[4], o[5], o[6], o[7]);
  }
}
__device__ __forceinline__ void ld_global_nc_256(const void* ptr, uint32_t (&r)[8]) {
  if ((reinterpret_cast<uintptr_t>(ptr) & 31u) == 0) {
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "l"(ptr));
  } else {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(ptr)), b = __ldg(reinterpret_cast<const uint4*>(ptr) + 1);
    r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w;
    r[4] = b.x; r[5] = b.y; r[6] = b.z; r[7] = b.w;
  }
}

static inline int ceil_div(int64_t a, int64_t b) { return static_cast<int>((a + b - 1) / b); }

}  // namespace odvae
