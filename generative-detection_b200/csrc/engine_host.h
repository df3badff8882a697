// Internal (C++) interface of the GEMM/conv engines; the C ABI in api_conv.cu wraps these.
#pragma once
#include <cuda_runtime.h>

#include "engine_desc.h"

namespace odvae {
int run_engine(EngineDesc& d, int impl, int act_dtype, cudaStream_t stream);
int describe_gemm(EngineDesc& d, const void* A, const void* B, void* D, int M, int N, int K, int batch, long long lda,
                  long long ldb, long long ldd, long long strideA, long long strideB, long long strideD, int a_mn,
                  int b_mn, int out_mode, float alpha, const float* bias, const void* residual);
int describe_conv_fwd(EngineDesc& d, const void* x, const void* w_packed, const float* bias, const void* residual,
                      void* y, int B, int H, int W, int CinP, int CoutP, int Cout, int geom, int out_nchw_f32);
int describe_conv_dgrad(EngineDesc* ds, int* count, const void* dy, const void* w_packed_dgrad, int w_elem_bytes,
                        const void* dx_add, void* dx, int B, int H, int W, int CinP, int CoutP, int geom);
int describe_conv_wgrad(EngineDesc& d, const void* x, const void* dy, float* ws, int B, int H, int W, int CinP,
                        int CoutP, int geom, int impl);
// nearest-2x upsample + 3x3 conv (LD model.py:53-57) as parity-phase 2x2 convs over the LOW-resolution input:
//   fwd: 4 descriptors (one per output phase (py, px)), dgrad: 1 descriptor with 16 taps, wgrad: 4 descriptors (4 taps each)
int describe_upconv_fwd(EngineDesc* ds, const void* x, const void* w_phase_packed, int w_elem_bytes, const float* bias,
                        void* y, int B, int H, int W, int CinP, int CoutP);
int describe_upconv_dgrad(EngineDesc& d, const void* dy, const void* w_dgrad_packed, void* dx, int B, int H, int W,
                          int CinP, int CoutP);
int describe_upconv_wgrad(EngineDesc& d, int phase, const void* x, const void* dy, int dy_elem_bytes, float* ws, int B,
                          int H, int W, int CinP, int CoutP, int impl);
int set_conv_halo_enabled(int on);  // returns the previous setting
// groupnorm.cu: stats[b][g] = (mean, rstd) from per-(image, split, channel) partial sums
int launch_gn_stats_finalize(const float* part, float* stats, int B, int nsplit, int HW, int C, int G, float eps,
                             cudaStream_t st);
// pointwise.cu
int launch_attn_stats_combine(const float* part, float* lse2, int B, int slots, int N, cudaStream_t st);
// layout.cu
int launch_wgrad_reduce(const float* ws, float* dw, int Cout, int Cin, int taps, int CoutP, int CinP, int split_k,
                        int accumulate, cudaStream_t st, const float* cpart = nullptr, int cs_rows = 0,
                        float* db = nullptr, float scale = 1.0f);
int launch_upconv_wgrad_fold(const float* dwp, const float* dbp, float* dw, float* db, int Cout, int Cin, int accumulate,
                             float scale, cudaStream_t st);
size_t colsum_workspace_bytes(int cols);
int launch_colsum_final(const float* part, int nrows, int cols, float* out, int accumulate, cudaStream_t st);
int launch_colsum(const void* m, long long rows, int cols, int ld, int act_dtype, float* out, int accumulate,
                  float* part, cudaStream_t st, float scale = 1.0f);
}  // namespace odvae
