// Hardware probe (not product code): MN-major SWIZZLE_128B B operand read from a TMA-written pixel slab with
// LBO = 128 B, i.e. consecutive 64-wide N chunks are the SAME 64 channels shifted by one pixel row (K index), and a
// start address that is any 128-byte row of the slab. Needed to serve the 3 horizontal taps of the weight-gradient
// GEMM from one halo slab with a single N = 192 MMA.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -I../../generative-detection_b200/csrc mn_probe.cu -o mn_probe
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
static const int ROWS = 96, NV = 6;

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmI, float* out) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sX = smem;               // 96 rows x 128 B
  uint8_t* sI = smem + 16 * 1024;   // A: 128 x 64 K-major selector
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 40 * 1024);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(tmem_slot, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&bars[0], ROWS * 128 + 128 * 128);
    tma_load_2d(sX, &tmX, &bars[0], 0, 0);
    tma_load_2d(sI, &tmI, &bars[0], 0, 0);
  }
  uint32_t phase = 0;
  for (int v = 0; v < NV; ++v) {
    const int s = v * 3;  // start row
    if (threadIdx.x == 0) {
      if (v == 0) mbar_wait(&bars[0], 0);
      tc_fence_after();
      const uint32_t idesc = make_idesc_bf16(128, 192, 0, 1);
      for (int k = 0; k < 4; ++k) {
        const uint64_t ad = make_smem_desc_sw128(smem_u32(sI) + k * 32, 0, 1024);
        const uint64_t bd = make_smem_desc_sw128(smem_u32(sX) + s * 128 + k * 2048, 128, 1024);
        umma_bf16(tmem, ad, bd, idesc, k != 0);
      }
      umma_commit(&bars[1]);
    }
    mbar_wait(&bars[1], phase);
    phase ^= 1;
    tc_fence_after();
    for (int c = 0; c < 192; c += 16) {
      uint32_t r[16];
      tmem_ld_x16(tmem + (static_cast<uint32_t>(warp * 32) << 16) + c, r);
      tmem_ld_wait();
      for (int i = 0; i < 16; ++i) out[(static_cast<size_t>(v) * 128 + warp * 32 + lane) * 192 + c + i] = __uint_as_float(r[i]);
    }
    tc_fence_before();
    __syncthreads();
  }
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem, 256);
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
  if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); exit(2); }
  return tm;
}

int main() {
  void* ptr = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q);
  EncodeTiledFn fn = reinterpret_cast<EncodeTiledFn>(ptr);
  std::vector<bf16> hX(ROWS * 64), hI(128 * 64);
  std::vector<float> fX(ROWS * 64);
  for (int r = 0; r < ROWS; ++r)
    for (int c = 0; c < 64; ++c) {
      float v = static_cast<float>((r * 64 + c) % 251 - 125);
      fX[r * 64 + c] = v;
      hX[r * 64 + c] = __float2bfloat16(v);
    }
  for (int m = 0; m < 128; ++m)
    for (int k = 0; k < 64; ++k) hI[m * 64 + k] = __float2bfloat16(m == k ? 1.0f : 0.0f);
  bf16 *dX, *dI;
  float* dOut;
  cudaMalloc(&dX, hX.size() * 2);
  cudaMalloc(&dI, hI.size() * 2);
  cudaMalloc(&dOut, sizeof(float) * NV * 128 * 192);
  cudaMemcpy(dX, hX.data(), hX.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dI, hI.data(), hI.size() * 2, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  CUtensorMap tmX = make2d(fn, dX, ROWS, ROWS);
  CUtensorMap tmI = make2d(fn, dI, 128, 128);
  cudaMemset(dOut, 0xff, sizeof(float) * NV * 128 * 192);
  probe_kernel<<<1, 128, 64 * 1024>>>(tmX, tmI, dOut);
  cudaError_t e = cudaDeviceSynchronize();
  printf("launch %s\n", cudaGetErrorString(e));
  if (e != cudaSuccess) return 1;
  std::vector<float> hOut(NV * 128 * 192);
  cudaMemcpy(hOut.data(), dOut, hOut.size() * 4, cudaMemcpyDeviceToHost);
  for (int v = 0; v < NV; ++v) {
    const int s = v * 3;
    int bad = 0, first = -1;
    for (int m = 0; m < 64; ++m)        // D[m][j*64 + c] = X[s + m + j][c]
      for (int n = 0; n < 192; ++n) {
        const int j = n / 64, c = n % 64;
        if (hOut[(v * 128 + m) * 192 + n] != fX[(s + m + j) * 64 + c]) { if (first < 0) first = m * 192 + n; ++bad; }
      }
    printf("  start row %2d: %s (%d mismatches, first at m %d n %d)\n", s, bad ? "WRONG" : "ok", bad,
           first < 0 ? -1 : first / 192, first < 0 ? -1 : first % 192);
  }
  return 0;
}
