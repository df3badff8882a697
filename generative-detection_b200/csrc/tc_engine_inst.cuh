// One instance of the tcgen05 engine kernel per translation unit (tc_engine_inst<N>.cu) so that `make -j` compiles them
// in parallel; engine_host.cu picks the instance from (kind, swap, epi).
#pragma once
#include <cuda_runtime.h>
#include <mutex>
#include "tc_engine.cuh"

#define ODVAE_TC_ENGINE_INSTANCE(SPEC)                                                                             \
  namespace odvae {                                                                                                 \
  cudaError_t tc_engine_launch_##SPEC(int grid, size_t smem, cudaStream_t st, const CUtensorMap& tmA,               \
                                      const CUtensorMap& tmB, const TcParams& p) {                                  \
    static std::once_flag once;                                                                                     \
    static cudaError_t attr_err = cudaSuccess;                                                                      \
    std::call_once(once, [] {                                                                                       \
      attr_err = cudaFuncSetAttribute(tc_engine_kernel<SPEC>, cudaFuncAttributeMaxDynamicSharedMemorySize,          \
                                      227 * 1024);                                                                  \
    });                                                                                                             \
    if (attr_err != cudaSuccess) return attr_err;                                                                   \
    tc_engine_kernel<SPEC><<<grid, tc_threads(SPEC), smem, st>>>(tmA, tmB, p);                                            \
    return cudaSuccess;                                                                                             \
  }                                                                                                                 \
  }
