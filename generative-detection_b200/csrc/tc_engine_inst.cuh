// One instance of the tcgen05 engine kernel per translation unit (tc_engine_inst<N>.cu) so that `make -j` compiles them
// in parallel; engine_host.cu picks the instance from (kind, swap, epi).
#pragma once
#include <cuda_runtime.h>
#include <cstdlib>
#include <mutex>
#include "tc_engine.cuh"

namespace odvae {

// ODVAE_PDL [0]: launch the engine kernels with programmatic stream serialization -- the kernel's prologue (barrier init, TMEM
// allocation, descriptor prefetch) may run while the previous kernel of the stream drains; the kernel itself waits
// (griddepcontrol.wait, tc_engine.cuh) before it touches global memory, so results do not depend on the switch. Measured on
// one box (round 2): 140.4 / 142.3 ms per train step without, 142.3 / 141.8 ms with -- no gain, so it stays off.
inline bool tc_pdl_enabled() {
  static const bool on = [] {
    const char* v = getenv("ODVAE_PDL");
    return v && v[0] == '1';
  }();
  return on;
}

template <int SPEC>
cudaError_t tc_engine_launch(int grid, size_t smem, cudaStream_t st, const CUtensorMap& tmA, const CUtensorMap& tmB,
                             const TcParams& p) {
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    attr_err = cudaFuncSetAttribute(tc_engine_kernel<SPEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
  });
  if (attr_err != cudaSuccess) return attr_err;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned>(grid));
  cfg.blockDim = dim3(static_cast<unsigned>(tc_threads(SPEC)));
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = tc_pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, tc_engine_kernel<SPEC>, tmA, tmB, p);
}

}  // namespace odvae

#define ODVAE_TC_ENGINE_INSTANCE(SPEC)                                                                \
  namespace odvae {                                                                                    \
  cudaError_t tc_engine_launch_##SPEC(int grid, size_t smem, cudaStream_t st, const CUtensorMap& tmA,  \
                                      const CUtensorMap& tmB, const TcParams& p) {                     \
    return tc_engine_launch<SPEC>(grid, smem, st, tmA, tmB, p);                                        \
  }                                                                                                    \
  }
