// Probe: does cuTensorMapEncodeTiled accept a (c, y, x) dimension order over an NHWC tensor (stride of dim 2 smaller than the
// stride of dim 1), and does a box (64, 2, 130) land in shared memory as [x][y][c] rows? (halo slab for W = 128 conv tiles)
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__global__ void k(const __grid_constant__ CUtensorMap tm, unsigned short* out, int x0, int y0) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ __align__(8) unsigned long long bar;
  unsigned bar_a = (unsigned)__cvta_generic_to_shared(&bar);
  unsigned dst = (unsigned)__cvta_generic_to_shared(smem);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(64 * 2 * 130 * 2));
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(dst), "l"(&tm), "r"(0), "r"(y0), "r"(x0), "r"(bar_a) : "memory");
  }
  unsigned ok = 0;
  while (!ok) {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(bar_a));
  }
  for (int i = threadIdx.x; i < 64 * 2 * 130; i += blockDim.x) out[i] = reinterpret_cast<unsigned short*>(smem)[i];
}
int main() {
  const int H = 16, W = 128, C = 64;
  std::vector<__half> h(H * W * C);
  for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) for (int c = 0; c < C; ++c) h[(y * W + x) * C + c] = __float2half((float)(y * 1000 + x) + c * 0.0f);
  __half* d; cudaMalloc(&d, h.size() * 2); cudaMemcpy(d, h.data(), h.size() * 2, cudaMemcpyHostToDevice);
  void* fp = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q);
  EncodeFn fn = (EncodeFn)fp;
  CUtensorMap tm;
  cuuint64_t gdim[3] = {(cuuint64_t)C, (cuuint64_t)H, (cuuint64_t)W};
  cuuint64_t gstr[2] = {(cuuint64_t)W * C * 2, (cuuint64_t)C * 2};
  cuuint32_t box[3] = {64, 2, 130}, es[3] = {1, 1, 1};
  CUresult r = fn(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, d, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode result %d\n", (int)r);
  if (r != CUDA_SUCCESS) return 0;
  unsigned short* out; cudaMalloc(&out, 64 * 2 * 130 * 2);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 40000);
  k<<<1, 128, 40000>>>(tm, out, -1, 3);
  cudaError_t e = cudaDeviceSynchronize();
  printf("kernel: %s\n", cudaGetErrorString(e));
  std::vector<unsigned short> o(64 * 2 * 130);
  cudaMemcpy(o.data(), out, o.size() * 2, cudaMemcpyDeviceToHost);
  // smem row r (128 B) -> expect (x = r / 2 - 1, y = 3 + r % 2); channel 0 of a SW128 row sits at 16-byte chunk (0 ^ (r % 8))
  for (int r = 0; r < 8; ++r) {
    __half v = *reinterpret_cast<__half*>(&o[r * 64 + ((0 ^ (r & 7)) * 8)]);
    printf("row %d: %.0f (expect x=%d y=%d -> %d)\n", r, __half2float(v), r / 2 - 1, 3 + (r & 1), (r / 2 - 1) < 0 ? 0 : (3 + (r & 1)) * 1000 + r / 2 - 1);
  }
  for (int r = 256; r < 260; ++r) {
    __half v = *reinterpret_cast<__half*>(&o[r * 64 + ((0 ^ (r & 7)) * 8)]);
    printf("row %d: %.0f (x=%d y=%d)\n", r, __half2float(v), r / 2 - 1, 3 + (r & 1));
  }
  return 0;
}
