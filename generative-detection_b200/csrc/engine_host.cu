// Host side of the GEMM/conv engines: problem description, TMA tensor maps, tile/BN/split choice,
// launches of the tcgen05 engine (bf16) and of the SIMT validation engine (f32 or bf16).
#include <cstdio>
#include <cuda.h>
#include <cudaTypedefs.h>

#include <mutex>

#include "engine_host.h"

#include <stdlib.h>
#include "simt_engine.cuh"
#include "tc_engine.cuh"

namespace odvae {  // tc_engine_inst<N>.cu
#define ODVAE_DECL_TC_LAUNCH(N)                                                                    \
  cudaError_t tc_engine_launch_##N(int grid, size_t smem, cudaStream_t st, const CUtensorMap& tmA, \
                                   const CUtensorMap& tmB, const TcParams& p);
ODVAE_DECL_TC_LAUNCH(0) ODVAE_DECL_TC_LAUNCH(1) ODVAE_DECL_TC_LAUNCH(2)
ODVAE_DECL_TC_LAUNCH(3) ODVAE_DECL_TC_LAUNCH(4) ODVAE_DECL_TC_LAUNCH(5) ODVAE_DECL_TC_LAUNCH(6)
#undef ODVAE_DECL_TC_LAUNCH
}  // namespace odvae

namespace odvae {

// ---------------------------------------------------------------------------------------------
// cuTensorMapEncodeTiled via the runtime's driver entry point (no link-time libcuda dependency)
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) fn = reinterpret_cast<EncodeTiledFn>(ptr);
  });
  return fn;
}

static int make_tmap_bf16(CUtensorMap* tm, const OperandView& v, const uint32_t* box) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) {
    set_last_cuda_error(cudaErrorUnknown, "cuTensorMapEncodeTiled entry point unavailable");
    return ODVAE_ERR_CUDA;
  }
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5];
  cuuint32_t es[5];
  for (int i = 0; i < v.rank; ++i) {
    gdim[i] = static_cast<cuuint64_t>(v.dims[i]);
    bx[i] = box[i];
    es[i] = 1;
  }
  for (int i = 1; i < v.rank; ++i) {
    gstr[i - 1] = static_cast<cuuint64_t>(v.strides[i]) * 2;
    if (gstr[i - 1] % 16 != 0) return ODVAE_ERR_SHAPE;
  }
  if ((reinterpret_cast<uintptr_t>(v.ptr) & 15) != 0) return ODVAE_ERR_SHAPE;
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, static_cast<cuuint32_t>(v.rank), const_cast<void*>(v.ptr),
                  gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_ERROR_INVALID_CONTEXT) {
    // a thread that has not touched the CUDA runtime yet (e.g. the first node an autograd worker thread runs): the driver
    // entry point above needs a current context, which any runtime call binds
    cudaFree(nullptr);
    r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, static_cast<cuuint32_t>(v.rank), const_cast<void*>(v.ptr), gdim, gstr, bx,
           es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS) {
    static thread_local char msg[320];
    int n = snprintf(msg, sizeof(msg), "cuTensorMapEncodeTiled failed (CUresult %d): rank %d ptr %p dims", static_cast<int>(r),
                     v.rank, v.ptr);
    for (int i = 0; i < v.rank && n < static_cast<int>(sizeof(msg)); ++i)
      n += snprintf(msg + n, sizeof(msg) - n, " %lld", static_cast<long long>(v.dims[i]));
    if (n < static_cast<int>(sizeof(msg))) n += snprintf(msg + n, sizeof(msg) - n, " strides");
    for (int i = 0; i < v.rank && n < static_cast<int>(sizeof(msg)); ++i)
      n += snprintf(msg + n, sizeof(msg) - n, " %lld", static_cast<long long>(v.strides[i]));
    if (n < static_cast<int>(sizeof(msg))) n += snprintf(msg + n, sizeof(msg) - n, " box");
    for (int i = 0; i < v.rank && n < static_cast<int>(sizeof(msg)); ++i) n += snprintf(msg + n, sizeof(msg) - n, " %u", box[i]);
    set_last_cuda_error(cudaErrorInvalidValue, msg);
    return ODVAE_ERR_CUDA;
  }
  return ODVAE_OK;
}

// test / profiling switch (odvae_set_conv_halo): 1 = never pick the halo conv path
static int g_disable_halo = 0;
int set_conv_halo_enabled(int on) {
  const int prev = g_disable_halo ? 0 : 1;
  g_disable_halo = on ? 0 : 1;
  return prev;
}

static int pick_bn(long long n) {
  const int cands[] = {256, 128, 64, 32, 16};
  for (int c : cands)
    if (n % c == 0) return c;
  return 0;
}
static bool is_pow2(int v) { return v > 0 && (v & (v - 1)) == 0; }
static int ilog2(int v) {
  int l = 0;
  while ((1 << l) < v) ++l;
  return l;
}

// BN here = B-operand rows per stage (WGRAD tap groups: wg_T * BN)
static size_t tc_smem_bytes(int stages, int BN) {
  return static_cast<size_t>(stages) * (TC_A_BYTES + static_cast<size_t>(BN) * 128) + (2 * stages + 4) * 8 + 16 + 1024;
}
static int tc_pick_stages(int BN, int num_k_iters) {
  const size_t budget = 224 * 1024;
  int s = static_cast<int>((budget - 2048) / (TC_A_BYTES + static_cast<size_t>(BN) * 128));
  if (s > 8) s = 8;
  if (s > num_k_iters) s = num_k_iters;
  if (s < 1) s = 1;
  return s;
}

static size_t tc_halo_smem_bytes(int stages, int slabs) {
  return static_cast<size_t>(stages) * TC_A_BYTES + static_cast<size_t>(slabs) * TC_SLAB_BYTES +
         (2 * stages + 6 + 2 * slabs) * 8 + 16 + 1024;
}

static int launch_tc(EngineDesc& d, cudaStream_t stream) {
  if (!d.tc_ok) return ODVAE_ERR_SHAPE;
  TcParams& p = d.p;
  OperandView viewA = d.A;
  uint32_t boxA[5] = {d.boxA[0], d.boxA[1], d.boxA[2], d.boxA[3], d.boxA[4]};
  {
    // W = 128 halo tiles (conv_common marked the candidate): needs the interleaved epilogue = fp16 output, pixel stride 128
    static const bool halo2_on = [] {
      const char* v = getenv("ODVAE_CONV_HALO2");
      return !(v && v[0] == '0');
    }();
    if (halo2_on && TC_FAST_EPI && p.kind == KIND_CONV && p.halo2_ok && p.halo == 0 && p.out_mode == OUT_F16 && p.oX == 128 &&
        p.a_fmt == FMT_F16) {
      p.halo = 2;
      p.halo_slabs = 4;
      // (c, x, y, b) -> (c, y, x, b): a box of 64 channels x 2 rows x 130 pixels lands in shared memory as [x][row][c]
      const long long d1 = viewA.dims[1], s1 = viewA.strides[1];
      viewA.dims[1] = viewA.dims[2];
      viewA.strides[1] = viewA.strides[2];
      viewA.dims[2] = d1;
      viewA.strides[2] = s1;
      boxA[0] = 64; boxA[1] = 2; boxA[2] = TC_HALO_BOX; boxA[3] = 1;
    }
  }
  CUtensorMap tmA, tmB;
  int st = make_tmap_bf16(&tmA, viewA, boxA);
  if (st != ODVAE_OK) return st;
  st = make_tmap_bf16(&tmB, d.B, d.boxB);
  if (st != ODVAE_OK) return st;
  const int b_rows = p.kind == KIND_WGRAD ? p.BN * p.wg_T : p.BN;
  size_t smem;
  if (p.halo && p.kind == KIND_CONV) {
    p.stages = 5;
    smem = tc_halo_smem_bytes(p.stages, p.halo_slabs);
  } else if (p.halo) {  // WGRAD: stage = dy tile + one 9 KiB slab per 64-channel chunk
    const size_t stage_bytes = TC_A_BYTES + static_cast<size_t>(p.BN / 64) * TC_WSLAB_BYTES;
    int s = static_cast<int>((224 * 1024 - 2048) / stage_bytes);
    if (s > 8) s = 8;
    if (s > p.num_k_iters) s = p.num_k_iters;
    p.stages = s < 1 ? 1 : s;
    smem = p.stages * stage_bytes + (2 * p.stages + 4) * 8 + 16 + 1024;
  } else {
    p.stages = tc_pick_stages(b_rows, p.num_k_iters);
    smem = tc_smem_bytes(p.stages, b_rows);
  }
  // dO v^T + softmax backward (K = C = 256, 32 x 16 tiles per image): keep the v panel in shared memory for a run of
  // m-tiles, so that the L2 -> SM path carries the P / dS tiles of the epilogue instead of re-fetched operands (engine_desc.h)
  p.bstat_mt = p.bstat_runs = 0;
  {
    static const bool bstat_on = [] {
      const char* v = getenv("ODVAE_GEMM_BSTAT");
      return !(v && v[0] == '0');
    }();
    const size_t panel = static_cast<size_t>(p.num_k_iters) * p.BN * 128;
    if (bstat_on && p.kind == KIND_GEMM && p.epi == EPI_DSOFTMAX && !p.a_mn && !p.b_mn && panel <= 128 * 1024 && p.num_m_tiles >= 8) {
      const int sms = num_sms();
      // longest run length (a divisor of the m-tile count) whose last wave of runs still fills >= 90 % of the SMs
      int best = 0;
      for (int mt = p.num_m_tiles; mt >= 4; --mt) {
        if (p.num_m_tiles % mt) continue;
        const long long runs = static_cast<long long>(p.num_n_tiles) * p.num_z * (p.num_m_tiles / mt);
        const long long waves = (runs + sms - 1) / sms;
        if (runs >= sms && static_cast<double>(runs) / static_cast<double>(waves * sms) >= 0.9) {
          best = mt;
          break;
        }
      }
      if (best > 0) {
        p.bstat_mt = best;
        p.bstat_runs = static_cast<int>(static_cast<long long>(p.num_n_tiles) * p.num_z * (p.num_m_tiles / best));
        int s = static_cast<int>((224 * 1024 - 2048 - panel) / TC_A_BYTES);
        if (s > 8) s = 8;
        p.stages = s;
        smem = static_cast<size_t>(p.stages) * TC_A_BYTES + panel + (2 * p.stages + 8) * 8 + 16 + 1024;
      }
    }
  }
  // K-streaming GEMMs with N = 256 (the attention P / dS GEMMs): two m-tiles per CTA tile share every B tile (engine_desc.h)
  p.mpair = 0;
  {
    // ODVAE_GEMM_MPAIR: 0 = off, 2 = whenever the shape allows (tests), default = only where the pair tiles still fill
    // the GPU four times over (read per call: the tests switch it)
    const char* mv = getenv("ODVAE_GEMM_MPAIR");
    const int mpair_mode = mv ? atoi(mv) : 1;
    const long long pairs = static_cast<long long>(p.num_m_tiles / 2) * p.num_n_tiles * p.num_z;
    if (mpair_mode != 0 && p.kind == KIND_GEMM && p.epi == EPI_NONE && p.bstat_mt == 0 && p.BN == 256 &&
        (p.num_m_tiles % 2) == 0 && ((p.num_k_iters >= 16 && pairs >= 4ll * num_sms()) || mpair_mode == 2)) {
      p.mpair = 1;
      p.num_m_tiles /= 2;
      const size_t stage_bytes = 2 * static_cast<size_t>(TC_A_BYTES) + static_cast<size_t>(p.BN) * 128;
      int s = static_cast<int>((224 * 1024 - 2048) / stage_bytes);
      if (s > p.num_k_iters) s = p.num_k_iters;
      p.stages = s;
      smem = static_cast<size_t>(p.stages) * stage_bytes + (2 * p.stages + 4) * 8 + 16 + 1024;
    }
  }
  // swapped-role conv epilogue: 8 x 2 KiB staging buffers for 16-byte vector stores / residual loads
  p.epi_off = 0;
  {
    // default OFF since round 2: with the tile's residual lines prefetched into L2 at tile start (res_prefetch below) the
    // direct epilogue (2-byte loads, 64 contiguous bytes per warp and pixel) measured 1.13-1.15 ms for 128->128 3x3 @256^2 x 64
    // with a residual against 1.25-1.26 ms through the staging buffers (which also cost the halo tiles one weight stage)
    static const bool epi_stage_on = [] {
      const char* v = getenv("ODVAE_EPI_STAGE");
      return v && v[0] == '1';
    }();
    const bool aligned = (p.n_valid % 8) == 0 && (p.oX % 8) == 0 && (p.oY % 8) == 0 && (p.oB % 8) == 0 && (p.o0 % 8) == 0;
    // (only where a residual is added: plain stores measured no faster through the staging buffer)
    if (epi_stage_on && p.halo != 2 && p.kind == KIND_CONV && p.swap && aligned && p.residual != nullptr &&
        (p.out_mode == OUT_BF16 || p.out_mode == OUT_F16)) {
      const size_t limit = 227 * 1024;
      if (p.halo && smem + 16384 + 128 > limit) {  // give one weight stage to the epilogue
        p.stages -= 1;
        smem = tc_halo_smem_bytes(p.stages, p.halo_slabs);
      }
      const size_t off = (smem - 1024 + 127) & ~static_cast<size_t>(127);
      if (off + 16384 + 1024 <= limit && p.stages >= 2) {
        p.epi_off = static_cast<int>(off);
        smem = off + 16384 + 1024;
      }
    }
    static const bool res_prefetch_on = [] {
      const char* v = getenv("ODVAE_RES_PREFETCH");
      return !(v && v[0] == '0');
    }();
    p.res_prefetch = (res_prefetch_on && p.kind == KIND_CONV && p.swap && p.residual != nullptr) ? 1 : 0;
  }
  const long long total = static_cast<long long>(p.num_m_tiles) * p.num_n_tiles * p.num_z;
  if (total <= 0 || p.num_k_iters <= 0) return ODVAE_ERR_SHAPE;
  int grid = num_sms();
  if (total < grid) grid = static_cast<int>(total);
  if (p.bstat_mt > 0 && p.bstat_runs < grid) grid = p.bstat_runs;
  // the instance of tc_engine_kernel<SPEC> compiled for this (kind, operand roles, epilogue)
  int spec;
  if (p.kind == KIND_CONV) {
    spec = p.swap ? (p.gnb_part != nullptr ? TC_SPEC_CONV_SWAP_GNB : TC_SPEC_CONV_SWAP) : TC_SPEC_CONV;
    if (p.gnb_part != nullptr && !(p.swap && TC_FAST_EPI && p.out_mode == OUT_F16 && p.oX == 128 && p.residual == nullptr &&
                                   p.gn_part == nullptr))
      return ODVAE_ERR_SHAPE;  // the fused GroupNorm-backward sums exist in the fp16 fast epilogue only
  } else if (p.kind == KIND_WGRAD) {
    spec = TC_SPEC_WGRAD;
  } else {
    spec = p.epi == EPI_NONE ? TC_SPEC_GEMM : (p.epi == EPI_DSOFTMAX ? TC_SPEC_GEMM_DSOFTMAX : TC_SPEC_GEMM_SOFTMAX);
  }
  if (p.kind != KIND_GEMM && (p.epi != EPI_NONE || (p.swap && p.kind != KIND_CONV))) return ODVAE_ERR_SHAPE;
  typedef cudaError_t (*LaunchFn)(int, size_t, cudaStream_t, const CUtensorMap&, const CUtensorMap&, const TcParams&);
  static const LaunchFn launchers[TC_NUM_SPECS] = {tc_engine_launch_0, tc_engine_launch_1, tc_engine_launch_2,
                                                   tc_engine_launch_3, tc_engine_launch_4, tc_engine_launch_5,
                                                   tc_engine_launch_6};
  const cudaError_t attr_err = launchers[spec](grid, smem, stream, tmA, tmB, p);
  if (attr_err != cudaSuccess) {
    set_last_cuda_error(attr_err, "cudaFuncSetAttribute(tc_engine_kernel)");
    return ODVAE_ERR_CUDA;
  }
  return check_launch("tc_engine_kernel");
}

static int launch_simt(const EngineDesc& d, int act_dtype, cudaStream_t stream) {
  SimtParams s = {};
  s.p = d.p;
  s.A = d.A;
  s.B = d.B;
  const TcParams& p = d.p;
  s.out_h = d.conv_out_h;
  s.out_w = d.conv_out_w;
  s.lda = d.gemm_lda;
  s.ldb = d.gemm_ldb;
  if (p.kind == KIND_GEMM) {
    s.M = p.m_valid;
    s.N = p.n_valid;
    s.K = d.gemm_k;
    s.strideA = d.gemm_strideA;
    s.strideB = d.gemm_strideB;
  } else if (p.kind == KIND_CONV) {
    s.M = static_cast<long long>(d.conv_batch) * d.conv_out_h * d.conv_out_w;
    s.N = d.gemm_n;
    s.chan_per_tap = d.chan_per_tap;
    s.K = static_cast<long long>(p.ntaps) * d.chan_per_tap;
  } else {
    s.M = p.m_valid;
    s.N = p.n_valid;
    s.K = static_cast<long long>(d.conv_batch) * d.conv_out_h * d.conv_out_w;
    s.k_per_split = (s.K + p.split_k - 1) / p.split_k;
  }
  dim3 grid(static_cast<unsigned>((s.N + 63) / 64), static_cast<unsigned>((s.M + 63) / 64), static_cast<unsigned>(p.num_z));
  if (grid.y > 65535 || grid.z > 65535) return ODVAE_ERR_SHAPE;
  const bool out_half = p.out_mode == OUT_F16;
  if (act_dtype == ODVAE_BF16) {
    if (out_half) simt_engine_kernel<bf16, f16><<<grid, 256, 0, stream>>>(s);
    else simt_engine_kernel<bf16, bf16><<<grid, 256, 0, stream>>>(s);
  } else if (act_dtype == ODVAE_F16) {
    if (out_half) simt_engine_kernel<f16, f16><<<grid, 256, 0, stream>>>(s);
    else simt_engine_kernel<f16, bf16><<<grid, 256, 0, stream>>>(s);
  } else {
    simt_engine_kernel<float, float><<<grid, 256, 0, stream>>>(s);
  }
  return check_launch("simt_engine_kernel");
}

int run_engine(EngineDesc& d, int impl, int act_dtype, cudaStream_t stream) {
  if (impl == ODVAE_IMPL_TC) {
    if (act_dtype != ODVAE_BF16 && act_dtype != ODVAE_F16) return ODVAE_ERR_SHAPE;
    d.p.a_fmt = d.p.b_fmt = (act_dtype == ODVAE_F16 ? FMT_F16 : FMT_BF16);  // one operand format per MMA (B200 traps on mixes)
    return launch_tc(d, stream);
  }
  return launch_simt(d, act_dtype, stream);
}

// ---------------------------------------------------------------------------------------------
// Batched GEMM description
// ---------------------------------------------------------------------------------------------
int describe_gemm(EngineDesc& d, const void* A, const void* B, void* D, int M, int N, int K, int batch, long long lda,
                  long long ldb, long long ldd, long long strideA, long long strideB, long long strideD, int a_mn,
                  int b_mn, int out_mode, float alpha, const float* bias, const void* residual) {
  if (M <= 0 || N <= 0 || K <= 0 || batch <= 0) return ODVAE_ERR_SHAPE;
  d = EngineDesc{};
  TcParams& p = d.p;
  int BN = pick_bn(N);
  if (BN == 0) BN = N < 64 ? ((N + 15) / 16) * 16 : 64;
  if (b_mn && (BN % 64 != 0)) BN = 64;
  d.tc_ok = 1;
  if (lda % 8 || ldb % 8) d.tc_ok = 0;
  if (batch > 1 && (strideA % 8 || strideB % 8)) d.tc_ok = 0;
  if ((out_mode == OUT_BF16 || out_mode == OUT_F16) && (ldd % 8 != 0 || strideD % 8 != 0)) d.tc_ok = 0;
  if (out_mode == OUT_F32 && (ldd % 4 != 0 || strideD % 4 != 0)) d.tc_ok = 0;
  const long long bsA = batch > 1 ? strideA : lda * (a_mn ? K : M);
  const long long bsB = batch > 1 ? strideB : ldb * (b_mn ? K : N);
  d.A.ptr = A;
  d.A.rank = 3;
  d.A.strides[0] = 1;
  d.A.strides[1] = lda;
  d.A.strides[2] = bsA;
  d.A.dims[2] = batch;
  if (!a_mn) {
    d.A.dims[0] = K;
    d.A.dims[1] = M;
    d.boxA[0] = 64; d.boxA[1] = 128; d.boxA[2] = 1;
  } else {
    d.A.dims[0] = M;
    d.A.dims[1] = K;
    d.boxA[0] = 64; d.boxA[1] = 64; d.boxA[2] = 1;
  }
  d.B.ptr = B;
  d.B.rank = 3;
  d.B.strides[0] = 1;
  d.B.strides[1] = ldb;
  d.B.strides[2] = bsB;
  d.B.dims[2] = batch;
  if (!b_mn) {
    d.B.dims[0] = K;
    d.B.dims[1] = N;
    d.boxB[0] = 64; d.boxB[1] = static_cast<uint32_t>(BN); d.boxB[2] = 1;
  } else {
    d.B.dims[0] = N;
    d.B.dims[1] = K;
    d.boxB[0] = 64; d.boxB[1] = 64; d.boxB[2] = 1;
  }
  p.kind = KIND_GEMM;
  p.a_fmt = p.b_fmt = FMT_BF16;
  p.a_mn = a_mn;
  p.b_mn = b_mn;
  p.BN = BN;
  p.num_m_tiles = (M + 127) / 128;
  p.num_n_tiles = (N + BN - 1) / BN;
  p.num_z = batch;
  p.num_k_iters = (K + 63) / 64;
  p.out_mode = out_mode;
  p.n_valid = N;
  p.m_valid = M;
  p.ldd = ldd;
  p.strideD = strideD;
  p.alpha = alpha;
  p.bias = bias;
  p.residual = residual;
  p.D = D;
  d.gemm_lda = lda;
  d.gemm_ldb = ldb;
  d.gemm_k = K;
  d.gemm_strideA = strideA;
  d.gemm_strideB = strideB;
  return ODVAE_OK;
}

// ---------------------------------------------------------------------------------------------
// Spatial views over NHWC activations
// ---------------------------------------------------------------------------------------------
static void view4(OperandView& v, const void* ptr, int B, int H, int W, int C) {
  v.ptr = ptr;
  v.rank = 4;
  v.dims[0] = C; v.dims[1] = W; v.dims[2] = H; v.dims[3] = B;
  v.strides[0] = 1; v.strides[1] = C; v.strides[2] = static_cast<long long>(W) * C;
  v.strides[3] = static_cast<long long>(H) * W * C;
}
// stride-2 phase view of [B,H,W,C]: element (c', j, ph, i, b) = x[b, 2i+ph, j, c'] with c' in [0,2C) covering
// pixels (2j, 2j+1)
static void view5_s2(OperandView& v, const void* ptr, int B, int H, int W, int C) {
  v.ptr = ptr;
  v.rank = 5;
  v.dims[0] = 2 * C; v.dims[1] = W / 2; v.dims[2] = 2; v.dims[3] = H / 2; v.dims[4] = B;
  v.strides[0] = 1; v.strides[1] = 2 * C; v.strides[2] = static_cast<long long>(W) * C;
  v.strides[3] = 2ll * W * C; v.strides[4] = static_cast<long long>(H) * W * C;
}
static void spatial_box(uint32_t* box, int rank, int bw, int bh) {
  if (rank == 5) {
    box[0] = 64; box[1] = bw; box[2] = 1; box[3] = bh; box[4] = 1;
  } else {
    box[0] = 64; box[1] = bw; box[2] = bh; box[3] = 1;
  }
}
static bool tile_geometry(int Ho, int Wo, int pixels, int* bw, int* bh) {
  if (!is_pow2(Wo)) return false;
  int w = Wo < pixels ? Wo : pixels;
  int h = pixels / w;
  if (h > Ho || Ho % h != 0) return false;
  *bw = w;
  *bh = h;
  return true;
}

static void fill_taps_fwd(TcParams& p, int geom, int CinP) {
  if (geom == ODVAE_CONV_1x1) {
    p.ntaps = 1;
    p.tap_dx[0] = p.tap_dy[0] = p.tap_c[0] = p.tap_ph[0] = 0;
    p.sp_rank = 4;
  } else if (geom == ODVAE_CONV_3x3_S1) {
    p.ntaps = 9;
    for (int r = 0; r < 3; ++r)
      for (int s = 0; s < 3; ++s) {
        const int t = r * 3 + s;
        p.tap_dx[t] = s - 1;
        p.tap_dy[t] = r - 1;
        p.tap_c[t] = 0;
        p.tap_ph[t] = 0;
      }
    p.sp_rank = 4;
  } else {  // 3x3 stride 2, zero pad bottom/right (model.py:72-76), gathered through the phase view
    p.ntaps = 9;
    for (int r = 0; r < 3; ++r)
      for (int s = 0; s < 3; ++s) {
        const int t = r * 3 + s;
        p.tap_dx[t] = s >> 1;
        p.tap_dy[t] = r >> 1;
        p.tap_c[t] = (s & 1) * CinP;
        p.tap_ph[t] = r & 1;
      }
    p.sp_rank = 5;
  }
}

static void conv_common(EngineDesc& d, int Bn, int Ho, int Wo, int gatherC, int N, int out_mode) {
  TcParams& p = d.p;
  int bw = 0, bh = 0;
  const bool geo = tile_geometry(Ho, Wo, 128, &bw, &bh);
  const int BN = pick_bn(N);
  d.tc_ok = geo && gatherC % 64 == 0 && BN != 0;
  p.kind = KIND_CONV;
  p.a_fmt = p.b_fmt = FMT_BF16;
  p.BN = BN ? BN : 16;
  p.cchunks = gatherC / 64;
  p.num_k_iters = p.ntaps * p.cchunks;
  if (geo) {
    p.bw_log2 = ilog2(bw);
    p.tiles_x = Wo / bw;
    p.tiles_y = Ho / bh;
    p.num_m_tiles = Bn * p.tiles_x * p.tiles_y;
    spatial_box(d.boxA, p.sp_rank, bw, bh);
  }
  p.num_n_tiles = BN ? N / BN : 1;
  p.num_z = 1;
  p.m_valid = p.num_m_tiles * 128;
  p.out_mode = out_mode;
  // Swapped roles for the N = 128 layers: a 128 x 128 tile reads A and B from shared memory at the full 128 B/clk
  // of the SM (measured ~750 TF/s), a 128 x 256 tile at 96 B/clk (~1250 TF/s). With M = 128 output channels
  // (weights) and N = 256 pixels the same layer runs the wide tile; the epilogue stores channel-contiguous.
  p.swap = 0;
  if (d.tc_ok && geo && BN == 128 && N == 128 && (out_mode == OUT_BF16 || out_mode == OUT_NCHW_F32) && bw >= 32 &&
      (p.num_m_tiles % 2) == 0) {
    p.swap = 1;
    p.BN = 256;
    p.num_n_tiles = p.num_m_tiles / 2;  // pairs of pixel boxes
    p.num_m_tiles = N / 128;            // tiles of 128 output channels
  }
  // Halo mode: 256 consecutive pixels of an image row (W = 256: the whole row, W = 512: half a row) are one N tile; the 9 taps re-read 3 input rows from shared memory instead of 9 shifted
  // boxes from L2 (the 128-channel layers at 256x256 were L2->SM bound: 14.5 GB per launch at ~10.7 TB/s)
  p.halo = 0;
  if (p.swap && !g_disable_halo && p.ntaps == 9 && p.sp_rank == 4 && Wo % 256 == 0 && bw == 128 && p.tap_dx[0] == -1 &&
      p.tap_dy[0] == -1 && p.tap_dx[8] == 1 && p.tap_dy[8] == 1) {
    p.halo = 1;
    p.halo_slabs = 4;
    d.boxA[0] = 64; d.boxA[1] = TC_HALO_BOX; d.boxA[2] = 1; d.boxA[3] = 1;
  }
  // Halo mode for W = 128 (engine_desc.h `halo == 2`): the N tile is TWO image rows; one slab = input rows (y, y + 1) of a
  // 64-channel chunk, pixel-interleaved ([x][row][c], written by ONE TMA box over a (c, y, x, b) view of the tensor), so the
  // tile's 256 pixels (x, row) are 256 consecutive slab rows and a horizontal tap is a shift of two rows. Candidate only:
  // launch_tc turns it on when the output format allows the interleaved epilogue (fp16, pixel stride 128).
  p.halo2_ok = 0;
  if (p.swap && !g_disable_halo && p.ntaps == 9 && p.sp_rank == 4 && Wo == 128 && bw == 128 && (Ho % 2) == 0 &&
      out_mode == OUT_BF16 && p.tap_dx[0] == -1 && p.tap_dy[0] == -1 && p.tap_dx[8] == 1 && p.tap_dy[8] == 1)
    p.halo2_ok = 1;
  p.alpha = 1.0f;
  d.conv_batch = Bn;
  d.conv_out_h = Ho;
  d.conv_out_w = Wo;
  d.chan_per_tap = gatherC;
  d.gemm_n = N;
  d.gemm_ldb = static_cast<long long>(p.ntaps) * gatherC;
  // weight matrix [N][ntaps*gatherC], K-major
  d.B.rank = 2;
  d.B.dims[0] = d.gemm_ldb;
  d.B.dims[1] = N;
  d.B.strides[0] = 1;
  d.B.strides[1] = d.gemm_ldb;
  d.boxB[0] = 64;
  d.boxB[1] = p.swap ? 128u : static_cast<uint32_t>(p.BN);  // swapped: 128 weight rows fill the A slot
}

int describe_conv_fwd(EngineDesc& d, const void* x, const void* w_packed, const float* bias, const void* residual,
                      void* y, int B, int H, int W, int CinP, int CoutP, int Cout, int geom, int out_nchw_f32) {
  if (geom == ODVAE_CONV_3x3_S2 && ((H | W) & 1)) return ODVAE_ERR_SHAPE;
  if (out_nchw_f32 && residual) return ODVAE_ERR_SHAPE;
  const int Ho = geom == ODVAE_CONV_3x3_S2 ? H / 2 : H, Wo = geom == ODVAE_CONV_3x3_S2 ? W / 2 : W;
  d = EngineDesc{};
  TcParams& p = d.p;
  fill_taps_fwd(p, geom, CinP);
  if (geom == ODVAE_CONV_3x3_S2)
    view5_s2(d.A, x, B, H, W, CinP);
  else
    view4(d.A, x, B, H, W, CinP);
  conv_common(d, B, Ho, Wo, CinP, CoutP, out_nchw_f32 ? OUT_NCHW_F32 : OUT_BF16);
  d.B.ptr = w_packed;
  p.n_valid = out_nchw_f32 ? Cout : CoutP;
  p.out_h = Ho;
  p.out_w = Wo;
  p.ldd = CoutP;
  p.oB = static_cast<long long>(Ho) * Wo * CoutP;
  p.oY = static_cast<long long>(Wo) * CoutP;
  p.oX = CoutP;
  p.o0 = 0;
  p.bias = bias;
  p.residual = residual;
  p.D = y;
  return ODVAE_OK;
}

// dgrad: up to 4 engine launches (stride-2 runs one per input parity phase)
int describe_conv_dgrad(EngineDesc* ds, int* count, const void* dy, const void* w_packed_dgrad, int w_elem_bytes,
                        const void* dx_add, void* dx, int B, int H, int W, int CinP, int CoutP, int geom) {
  if (geom != ODVAE_CONV_3x3_S2) {
    EngineDesc& d = ds[0];
    d = EngineDesc{};
    TcParams& p = d.p;
    fill_taps_fwd(p, geom, CoutP);  // the 180-degree rotation is baked into the packed weights
    view4(d.A, dy, B, H, W, CoutP);
    conv_common(d, B, H, W, CoutP, CinP, OUT_BF16);
    d.B.ptr = w_packed_dgrad;
    p.n_valid = CinP;
    p.out_h = H;
    p.out_w = W;
    p.ldd = CinP;
    p.oB = static_cast<long long>(H) * W * CinP;
    p.oY = static_cast<long long>(W) * CinP;
    p.oX = CinP;
    p.residual = dx_add;
    p.D = dx;
    *count = 1;
    return ODVAE_OK;
  }
  if ((H | W) & 1) return ODVAE_ERR_SHAPE;
  const int Ho = H / 2, Wo = W / 2;
  long long w_off = 0;
  int n = 0;
  for (int py = 0; py < 2; ++py)
    for (int px = 0; px < 2; ++px) {
      EngineDesc& d = ds[n++];
      d = EngineDesc{};
      TcParams& p = d.p;
      p.sp_rank = 4;
      int nt = 0;
      for (int r = 0; r < 3; ++r)
        for (int s = 0; s < 3; ++s)
          if ((r & 1) == py && (s & 1) == px) {
            p.tap_dx[nt] = -(s >> 1);
            p.tap_dy[nt] = -(r >> 1);
            p.tap_c[nt] = 0;
            p.tap_ph[nt] = 0;
            ++nt;
          }
      p.ntaps = nt;
      view4(d.A, dy, B, Ho, Wo, CoutP);
      conv_common(d, B, Ho, Wo, CoutP, CinP, OUT_BF16);
      d.B.ptr = reinterpret_cast<const uint8_t*>(w_packed_dgrad) + w_off * w_elem_bytes;
      w_off += static_cast<long long>(nt) * CoutP * CinP;
      p.n_valid = CinP;
      p.out_h = Ho;
      p.out_w = Wo;
      p.ldd = CinP;
      p.oB = static_cast<long long>(H) * W * CinP;
      p.oY = 2ll * W * CinP;
      p.oX = 2ll * CinP;
      p.o0 = (static_cast<long long>(py) * W + px) * CinP;
      p.residual = dx_add;
      p.D = dx;
    }
  *count = 4;
  return ODVAE_OK;
}

// wgrad: ws[tap][split][CoutP][CinP] (fp32) = sum over the split's output pixels of dy^T x_shifted
int describe_conv_wgrad(EngineDesc& d, const void* x, const void* dy, float* ws, int B, int H, int W, int CinP,
                        int CoutP, int geom, int impl) {
  if (geom == ODVAE_CONV_3x3_S2 && ((H | W) & 1)) return ODVAE_ERR_SHAPE;
  const int Ho = geom == ODVAE_CONV_3x3_S2 ? H / 2 : H, Wo = geom == ODVAE_CONV_3x3_S2 ? W / 2 : W;
  d = EngineDesc{};
  TcParams& p = d.p;
  p.kind = KIND_WGRAD;
  p.a_fmt = p.b_fmt = FMT_BF16;
  p.a_mn = 1;
  p.b_mn = 1;
  fill_taps_fwd(p, geom, CinP);
  view4(d.A, dy, B, Ho, Wo, CoutP);
  if (geom == ODVAE_CONV_3x3_S2)
    view5_s2(d.B, x, B, H, W, CinP);
  else
    view4(d.B, x, B, H, W, CinP);
  int kbw = 0, kbh = 0;
  const bool geo = tile_geometry(Ho, Wo, 64, &kbw, &kbh);
  const long long kpix = static_cast<long long>(B) * Ho * Wo;
  d.tc_ok = geo && CinP % 64 == 0 && CoutP % 8 == 0 && kpix % 64 == 0;
  const bool tc = impl == ODVAE_IMPL_TC && d.tc_ok;
  // tcgen05: a tile accumulates one whole filter row (3 taps) so that the dy tile is fetched once per 3 taps and
  // the MMAs run N = 256 + 128 wide; 1x1 convs keep one tap with the widest BN that divides CinP.
  const int T = (tc && p.ntaps == 9) ? 3 : 1;
  int BN = CinP % 256 == 0 ? 256 : (CinP % 128 == 0 ? 128 : 64);
  if (T == 3 && BN > 128) BN = 128;
  p.BN = BN;
  p.wg_T = T;
  const long long base_tiles = static_cast<long long>(p.ntaps / T) * ((CoutP + 127) / 128) * ((CinP + BN - 1) / BN);
  int split = 1;
  if (tc) {
    const long long kch = kpix / 64;
    // one wave of tiles over the SMs; at least 8 pixel chunks per tile; uneven split (last split = remainder)
    long long want = num_sms() / base_tiles;
    if (want > kch / 8) want = kch / 8;
    if (want < 1) want = 1;
    const long long kps = (kch + want - 1) / want;
    split = static_cast<int>((kch + kps - 1) / kps);
    p.kbw_log2 = ilog2(kbw);
    p.kch_x = Wo / kbw;
    p.kch_y = Ho / kbh;
    p.kch_per_split = static_cast<int>(kps);
    p.kch_total = static_cast<int>(kch);
    p.num_k_iters = p.kch_per_split;
    spatial_box(d.boxA, 4, kbw, kbh);
    spatial_box(d.boxB, p.sp_rank, kbw, kbh);
    // Halo mode (3x3 stride 1, K chunk = 64 pixels of one image row): the x tile is fetched once per filter row with one
    // halo pixel on each side instead of three shifted copies (stage 64 KB -> 34 KB at Cin block 128)
    if (T == 3 && geom == ODVAE_CONV_3x3_S1 && kbw >= 16 && !g_disable_halo) {
      p.halo = 1;
      d.boxB[0] = 64; d.boxB[1] = static_cast<uint32_t>(kbw + 2); d.boxB[2] = static_cast<uint32_t>(kbh); d.boxB[3] = 1;
    }
  } else {
    // SIMT: 64x64 output tiles; split K so the grid fills the machine
    const long long tiles = static_cast<long long>(p.ntaps) * ((CoutP + 63) / 64) * ((CinP + 63) / 64);
    while (tiles * split < 4 * num_sms() && kpix / (2 * split) >= 256) split *= 2;
  }
  p.split_k = split;
  p.num_m_tiles = (CoutP + 127) / 128;
  p.num_n_tiles = (CinP + BN - 1) / BN;
  p.num_z = (p.ntaps / T) * split;
  p.out_mode = OUT_F32;
  p.n_valid = CinP;
  p.m_valid = CoutP;
  p.ldd = CinP;
  p.alpha = 1.0f;
  p.D = ws;
  d.conv_batch = B;
  d.conv_out_h = Ho;
  d.conv_out_w = Wo;
  d.gemm_lda = CoutP;
  return ODVAE_OK;
}

// ---------------------------------------------------------------------------------------------
// nearest-2x upsample + 3x3 conv as parity-phase 2x2 convs. Output pixel (2i+py, 2j+px) of the conv over the
// upsampled tensor reads low-resolution rows {i-1, i} (py = 0) or {i, i+1} (py = 1), likewise columns: 4 taps with
// pre-summed weights instead of 9 taps over a 4x larger tensor -- 16/36 of the FLOPs and no upsampled tensor in HBM.
// tap t = a*2 + b: row offset a - (1 - py), column offset b - (1 - px).
// ---------------------------------------------------------------------------------------------
static void fill_taps_up(TcParams& p, int py, int px) {
  p.ntaps = 4;
  p.sp_rank = 4;
  for (int a = 0; a < 2; ++a)
    for (int b = 0; b < 2; ++b) {
      const int t = a * 2 + b;
      p.tap_dy[t] = a - (1 - py);
      p.tap_dx[t] = b - (1 - px);
      p.tap_c[t] = 0;
      p.tap_ph[t] = 0;
    }
}

int describe_upconv_fwd(EngineDesc* ds, const void* x, const void* w_phase_packed, int w_elem_bytes, const float* bias,
                        void* y, int B, int H, int W, int CinP, int CoutP) {
  for (int py = 0; py < 2; ++py)
    for (int px = 0; px < 2; ++px) {
      EngineDesc& d = ds[py * 2 + px];
      d = EngineDesc{};
      TcParams& p = d.p;
      fill_taps_up(p, py, px);
      view4(d.A, x, B, H, W, CinP);
      conv_common(d, B, H, W, CinP, CoutP, OUT_BF16);
      d.B.ptr = reinterpret_cast<const uint8_t*>(w_phase_packed) +
                static_cast<long long>(py * 2 + px) * CoutP * 4 * CinP * w_elem_bytes;
      p.n_valid = CoutP;
      p.out_h = 2 * H;
      p.out_w = 2 * W;
      p.ldd = CoutP;
      p.oB = 4ll * H * W * CoutP;
      p.oY = 2ll * (2 * W) * CoutP;
      p.oX = 2ll * CoutP;
      p.o0 = (static_cast<long long>(py) * 2 * W + px) * CoutP;
      p.bias = bias;
      p.D = y;
    }
  return ODVAE_OK;
}

// dx[i, j] = sum over the 16 (phase, tap) pairs of Wp^T dy[2i' + py, 2j' + px]: ONE launch over the stride-2 phase view of dy
// (c', j, ph, i, b). (py, a) -> (row offset, ph): (0,0) -> (+1, 0), (0,1) -> (0, 0), (1,0) -> (0, 1), (1,1) -> (-1, 1); columns
// alike with the pixel-pair channel offset px * CoutP. tap t = ((py*2 + a)*2 + px)*2 + b  (the dgrad weight packing's order).
int describe_upconv_dgrad(EngineDesc& d, const void* dy, const void* w_dgrad_packed, void* dx, int B, int H, int W,
                          int CinP, int CoutP) {
  d = EngineDesc{};
  TcParams& p = d.p;
  p.ntaps = 16;
  p.sp_rank = 5;
  static const int off[2][2] = {{1, 0}, {0, -1}};  // [parity][a] -> low-resolution offset
  for (int py = 0; py < 2; ++py)
    for (int a = 0; a < 2; ++a)
      for (int px = 0; px < 2; ++px)
        for (int b = 0; b < 2; ++b) {
          const int t = ((py * 2 + a) * 2 + px) * 2 + b;
          p.tap_dy[t] = off[py][a];
          p.tap_ph[t] = py;
          p.tap_dx[t] = off[px][b];
          p.tap_c[t] = px * CoutP;
        }
  view5_s2(d.A, dy, B, 2 * H, 2 * W, CoutP);
  conv_common(d, B, H, W, CoutP, CinP, OUT_BF16);
  d.B.ptr = w_dgrad_packed;
  p.n_valid = CinP;
  p.out_h = H;
  p.out_w = W;
  p.ldd = CinP;
  p.oB = static_cast<long long>(H) * W * CinP;
  p.oY = static_cast<long long>(W) * CinP;
  p.oX = CinP;
  p.D = dx;
  return ODVAE_OK;
}

// ws[tap][split][CoutP][CinP] for ONE phase: dy sub-sampled at (2i + py, 2j + px) (a strided 4-D view), x shifted by the tap
int describe_upconv_wgrad(EngineDesc& d, int phase, const void* x, const void* dy, int dy_elem_bytes, float* ws, int B,
                          int H, int W, int CinP, int CoutP, int impl) {
  const int py = phase >> 1, px = phase & 1;
  d = EngineDesc{};
  TcParams& p = d.p;
  p.kind = KIND_WGRAD;
  p.a_fmt = p.b_fmt = FMT_BF16;
  p.a_mn = 1;
  p.b_mn = 1;
  fill_taps_up(p, py, px);
  // dy phase view: (c, j, i, b) -> dy[b, 2i + py, 2j + px, c]
  d.A.ptr = reinterpret_cast<const uint8_t*>(dy) + (static_cast<long long>(py) * 2 * W + px) * CoutP * dy_elem_bytes;
  d.A.rank = 4;
  d.A.dims[0] = CoutP; d.A.dims[1] = W; d.A.dims[2] = H; d.A.dims[3] = B;
  d.A.strides[0] = 1; d.A.strides[1] = 2ll * CoutP; d.A.strides[2] = 2ll * (2 * W) * CoutP;
  d.A.strides[3] = 4ll * H * W * CoutP;
  view4(d.B, x, B, H, W, CinP);
  int kbw = 0, kbh = 0;
  const bool geo = tile_geometry(H, W, 64, &kbw, &kbh);
  const long long kpix = static_cast<long long>(B) * H * W;
  d.tc_ok = geo && CinP % 64 == 0 && CoutP % 8 == 0 && kpix % 64 == 0;
  const bool tc = impl == ODVAE_IMPL_TC && d.tc_ok;
  int BN = CinP % 256 == 0 ? 256 : (CinP % 128 == 0 ? 128 : 64);
  p.BN = BN;
  p.wg_T = 1;
  const long long base_tiles = static_cast<long long>(p.ntaps) * ((CoutP + 127) / 128) * ((CinP + BN - 1) / BN);
  int split = 1;
  if (tc) {
    const long long kch = kpix / 64;
    long long want = num_sms() / base_tiles;
    if (want > kch / 8) want = kch / 8;
    if (want < 1) want = 1;
    const long long kps = (kch + want - 1) / want;
    split = static_cast<int>((kch + kps - 1) / kps);
    p.kbw_log2 = ilog2(kbw);
    p.kch_x = W / kbw;
    p.kch_y = H / kbh;
    p.kch_per_split = static_cast<int>(kps);
    p.kch_total = static_cast<int>(kch);
    p.num_k_iters = p.kch_per_split;
    spatial_box(d.boxA, 4, kbw, kbh);
    spatial_box(d.boxB, 4, kbw, kbh);
  } else {
    const long long tiles = static_cast<long long>(p.ntaps) * ((CoutP + 63) / 64) * ((CinP + 63) / 64);
    while (tiles * split < 4 * num_sms() && kpix / (2 * split) >= 256) split *= 2;
  }
  p.split_k = split;
  p.num_m_tiles = (CoutP + 127) / 128;
  p.num_n_tiles = (CinP + BN - 1) / BN;
  p.num_z = p.ntaps * split;
  p.out_mode = OUT_F32;
  p.n_valid = CinP;
  p.m_valid = CoutP;
  p.ldd = CinP;
  p.alpha = 1.0f;
  p.D = ws;
  d.conv_batch = B;
  d.conv_out_h = H;
  d.conv_out_w = W;
  d.gemm_lda = CoutP;
  return ODVAE_OK;
}

}  // namespace odvae
