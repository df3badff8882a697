// Hardware probe (not product code): can a tcgen05 SWIZZLE_128B K-major operand descriptor start at an arbitrary
// 128-byte row of a TMA-written tile (rows shifted by s), and can TMA write a swizzled box at a destination that is
// only 128-byte aligned? Both are needed to serve the 3 horizontally shifted taps of a 3x3 conv from ONE halo tile.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I../../generative-detection_b200/csrc halo_probe.cu -o halo_probe
#define ODVAE_WAIT_TIMEOUT_NS 300000000ull
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_bf16.h>
#include "tc_primitives.cuh"

using namespace odvae;
typedef __nv_bfloat16 bf16;

static const int ROWS = 264, NV = 12;  // variants per test

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA128,
             const __grid_constant__ CUtensorMap tmB, float* out, int mode) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                 // up to 264 rows x 128 B = 33792 B
  uint8_t* sB = smem + 36 * 1024;     // 64 rows x 128 B
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 48 * 1024);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_slot, 64);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  if (threadIdx.x == 0) {
    if (mode == 0) {  // one 136-row box at the aligned base
      mbar_arrive_expect_tx(&bars[0], 136 * 128 + 64 * 128);
      tma_load_2d(sA, &tmA, &bars[0], 0, 0);
    } else {          // box 1: rows 0..129 at base; box 2: rows 130..257 at base + 130 * 128 (128-byte aligned only)
      mbar_arrive_expect_tx(&bars[0], 130 * 128 + 128 * 128 + 64 * 128);
      tma_load_2d(sA, &tmA, &bars[0], 0, 0);  // tmA box is 136 rows in mode 0, 130 rows in mode 1 (host picks)
      tma_load_2d(sA + 130 * 128, &tmA128, &bars[0], 0, 130);
    }
    tma_load_2d(sB, &tmB, &bars[0], 0, 0);
  }
  uint32_t phase = 0;
  for (int v = 0; v < NV; ++v) {
    // mode 0: s = v / 2 + (v >= 8 ? 4 : 0) rows of shift, base_offset = (v & 1) ? (s & 7) : 0
    int s = mode == 0 ? (v >> 1) + (v >= 8 ? 4 : 0) : (v >> 1) * 23 + 1;
    int bo = (v & 1) ? (s & 7) : 0;
    if (threadIdx.x == 0) {
      if (v == 0) mbar_wait(&bars[0], 0);
      tc_fence_after();
      const uint32_t idesc = make_idesc_bf16(128, 64, 0, 0);
      for (int k = 0; k < 4; ++k) {
        const uint64_t ad = make_smem_desc_sw128(smem_u32(sA) + s * 128 + k * 32, 0, 1024, bo);
        const uint64_t bd = make_smem_desc_sw128(smem_u32(sB) + k * 32, 0, 1024);
        umma_bf16(tmem, ad, bd, idesc, k != 0);
      }
      umma_commit(&bars[1]);
    }
    mbar_wait(&bars[1], phase);
    phase ^= 1;
    tc_fence_after();
    for (int c = 0; c < 64; c += 16) {
      uint32_t r[16];
      tmem_ld_x16(tmem + (static_cast<uint32_t>(warp * 32) << 16) + c, r);
      tmem_ld_wait();
      for (int i = 0; i < 16; ++i) out[(static_cast<size_t>(v) * 128 + warp * 32 + lane) * 64 + c + i] = __uint_as_float(r[i]);
    }
    tc_fence_before();
    __syncthreads();
  }
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 64);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static CUtensorMap make2d(EncodeTiledFn fn, void* ptr, int rows, int boxrows) {
  CUtensorMap tm;
  cuuint64_t gdim[2] = {64, static_cast<cuuint64_t>(rows)};
  cuuint64_t gstr[1] = {128};
  cuuint32_t box[2] = {64, static_cast<cuuint32_t>(boxrows)};
  cuuint32_t es[2] = {1, 1};
  CUresult r = fn(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, ptr, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("encode failed %d (box rows %d)\n", (int)r, boxrows); exit(2); }
  return tm;
}

int main() {
  void* ptr = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q);
  EncodeTiledFn fn = reinterpret_cast<EncodeTiledFn>(ptr);
  std::vector<bf16> hA(ROWS * 64), hB(64 * 64);
  std::vector<float> fA(ROWS * 64);
  for (int r = 0; r < ROWS; ++r)
    for (int k = 0; k < 64; ++k) {
      float v = static_cast<float>((r * 64 + k) % 251 - 125);
      fA[r * 64 + k] = v;
      hA[r * 64 + k] = __float2bfloat16(v);
    }
  for (int n = 0; n < 64; ++n)
    for (int k = 0; k < 64; ++k) hB[n * 64 + k] = __float2bfloat16(n == k ? 1.0f : 0.0f);
  bf16 *dA, *dB;
  float* dOut;
  cudaMalloc(&dA, hA.size() * 2);
  cudaMalloc(&dB, hB.size() * 2);
  cudaMalloc(&dOut, sizeof(float) * NV * 128 * 64);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  std::vector<float> hOut(NV * 128 * 64);
  for (int mode = 0; mode < 2; ++mode) {
    CUtensorMap tmA = make2d(fn, dA, ROWS, mode == 0 ? 136 : 130);
    CUtensorMap tmA128 = make2d(fn, dA, ROWS, 128);
    CUtensorMap tmB = make2d(fn, dB, 64, 64);
    cudaMemset(dOut, 0xff, sizeof(float) * NV * 128 * 64);
    probe_kernel<<<1, 128, 64 * 1024>>>(tmA, tmA128, tmB, dOut, mode);
    cudaError_t e = cudaDeviceSynchronize();
    printf("mode %d: launch %s\n", mode, cudaGetErrorString(e));
    if (e != cudaSuccess) return 1;
    cudaMemcpy(hOut.data(), dOut, hOut.size() * 4, cudaMemcpyDeviceToHost);
    for (int v = 0; v < NV; ++v) {
      int s = mode == 0 ? (v >> 1) + (v >= 8 ? 4 : 0) : (v >> 1) * 23 + 1;
      int bo = (v & 1) ? (s & 7) : 0;
      int bad = 0, first = -1;
      for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 64; ++n)
          if (hOut[(v * 128 + m) * 64 + n] != fA[(s + m) * 64 + n]) { if (first < 0) first = m * 64 + n; ++bad; }
      printf("  mode %d shift %3d base_offset %d: %s (%d mismatches, first at row %d col %d)\n", mode, s, bo,
             bad ? "WRONG" : "ok", bad, first < 0 ? -1 : first / 64, first < 0 ? -1 : first % 64);
    }
  }
  return 0;
}
