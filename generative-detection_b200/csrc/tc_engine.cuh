// The tcgen05 GEMM engine: one persistent, warp-specialised kernel that serves
//   KIND_GEMM  : batched GEMM with K-major or MN-major operands (attention QK^T / PV / gradients)
//   KIND_CONV  : implicit-GEMM convolution forward / dgrad (A gathered by TMA from NHWC with OOB zero fill)
//   KIND_WGRAD : convolution weight gradient (both operands MN-major, K = pixels, deterministic split-K)
// Tile = 128 (M, TMEM lanes) x BN (<=256) with BLOCK_K = 64 bf16 (one 128-byte swizzle atom per row).
#pragma once
#include <type_traits>

#include "common.cuh"
#include "tc_primitives.cuh"
#include "engine_desc.h"

// A/B switch: the compile-time-specialised swapped-role conv epilogue (fp16, pixel stride 128) -- see `emit` below
#ifndef TC_FAST_EPI
#define TC_FAST_EPI 1
#endif

namespace odvae {

static constexpr int TC_BLOCK_M = 128;
static constexpr int TC_BLOCK_K = 64;
static constexpr int TC_A_BYTES = TC_BLOCK_M * TC_BLOCK_K * 2;  // 16 KiB
static constexpr int TC_THREADS = 320;  // warp 0: TMEM alloc + TMA producer; warp 1: MMA issuer; warps 2-9: epilogue
// (the softmax-backward instance runs 16 epilogue warps: tc_epi_warps / tc_threads in engine_desc.h)
static constexpr int TC_TMEM_COLS = 512;
// halo slab: one input row of 256 pixels + 1 halo pixel on each side, 128 B (64 channels) per pixel, written by two
// 130-pixel TMA boxes (x = -1..128 and 129..258; out-of-image pixels are zero-filled); padded to a 1 KiB multiple
static constexpr int TC_HALO_BOX = 130;
static constexpr int TC_SLAB_BYTES = 33 * 1024;
// weight-gradient halo slab: the K chunk's 64 output pixels (kbw x kbh, kbw = 64 / 32 / 16) with 1 halo pixel on each
// side of every row, one TMA box of (kbw + 2) x kbh pixels: 66 / 68 / 72 rows of 128 B
static constexpr int TC_WSLAB_BYTES = 9 * 1024;

struct TileCoord {
  int m_tile, n_tile, z;
  int b, y0, x0;  // CONV
  int tap, split; // WGRAD
};

// pixel box `pb` (128 output pixels) of a CONV problem -> (image, first row, first column)
__device__ __forceinline__ void decode_pixel_box(const TcParams& p, int pb, int& b, int& y0, int& x0) {
  const int per_img = p.tiles_x * p.tiles_y;
  b = pb / per_img;
  const int q = pb - b * per_img;
  const int ty = q / p.tiles_x, tx = q - ty * p.tiles_x;
  x0 = tx << p.bw_log2;
  y0 = ty * (TC_BLOCK_M >> p.bw_log2);
}

__device__ __forceinline__ TileCoord decode_tile(const TcParams& p, const int kind, const bool swap, int t) {
  TileCoord c;
  c.n_tile = t % p.num_n_tiles;
  int r = t / p.num_n_tiles;
  c.m_tile = r % p.num_m_tiles;
  c.z = r / p.num_m_tiles;
  c.b = c.y0 = c.x0 = c.tap = c.split = 0;
  if (kind == KIND_CONV) {
    // swapped roles: the N tile is a pair of pixel boxes (2*n_tile, 2*n_tile + 1); (b, y0, x0) = the first one
    decode_pixel_box(p, swap ? 2 * c.n_tile : c.m_tile, c.b, c.y0, c.x0);
  } else if (kind == KIND_WGRAD) {
    const int grp = c.z / p.split_k;
    c.split = c.z - grp * p.split_k;
    c.tap = grp * p.wg_T;  // first tap of this tile's tap group
  }
  return c;
}

// Tile s of this CTA's walk. Default: tiles blockIdx.x, + gridDim.x, ... B-stationary GEMM (bstat_mt > 0): runs blockIdx.x,
// + gridDim.x, ... of bstat_mt consecutive m-tiles with (n_tile, z) fixed; `mi` = position inside the run. -1 = done.
__device__ __forceinline__ int sched_tile(const TcParams& p, const int total_tiles, const int s, int& mi) {
  if (p.bstat_mt <= 0) {
    mi = 0;
    const long long t = static_cast<long long>(blockIdx.x) + static_cast<long long>(s) * gridDim.x;
    return t < total_tiles ? static_cast<int>(t) : -1;
  }
  const int rs = s / p.bstat_mt;
  mi = s - rs * p.bstat_mt;
  const long long run = static_cast<long long>(blockIdx.x) + static_cast<long long>(rs) * gridDim.x;
  if (run >= p.bstat_runs) return -1;
  const int nz = p.num_n_tiles * p.num_z;
  const int part = static_cast<int>(run / nz);
  const int rem = static_cast<int>(run - static_cast<long long>(part) * nz);
  const int z = rem / p.num_n_tiles, n = rem - z * p.num_n_tiles;
  return (z * p.num_m_tiles + part * p.bstat_mt + mi) * p.num_n_tiles + n;  // decode_tile's order
}

// k-iterations of one tile: WGRAD splits its pixel chunks unevenly (the last split takes the remainder)
__device__ __forceinline__ int tile_k_iters(const TcParams& p, const int kind, const TileCoord& c) {
  if (kind != KIND_WGRAD) return p.num_k_iters;
  const int rem = p.kch_total - c.split * p.kch_per_split;
  return rem < p.kch_per_split ? rem : p.kch_per_split;
}

__device__ __forceinline__ void spatial_load(void* dst, const CUtensorMap* tm, uint64_t* bar, int rank, int c,
                                             int x, int ph, int y, int b) {
  if (rank == 5)
    tma_load_5d(dst, tm, bar, c, x, ph, y, b);
  else
    tma_load_4d(dst, tm, bar, c, x, y, b);
}

// SPEC: TC_SPEC_* (engine_desc.h). One instance per .cu file (tc_engine_inst*.cu) so they compile in parallel.
template <int SPEC>
__global__ void __launch_bounds__(tc_threads(SPEC), 1)
tc_engine_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ TcParams p) {
  // The operation kind, the operand roles and the epilogue are compile-time constants of each instance: every mode gets
  // its own register allocation (the monolithic kernel spilled at the 168-register cap of a 320-thread CTA).
  constexpr bool kGnb = SPEC == TC_SPEC_CONV_SWAP_GNB;
  constexpr int kKind = (SPEC <= TC_SPEC_CONV_SWAP || kGnb) ? KIND_CONV : (SPEC == TC_SPEC_WGRAD ? KIND_WGRAD : KIND_GEMM);
  constexpr bool kSwap = SPEC == TC_SPEC_CONV_SWAP || kGnb;
  constexpr int kEpiThreads = 32 * tc_epi_warps(SPEC);
  const int epi = (SPEC <= TC_SPEC_GEMM || kGnb) ? static_cast<int>(EPI_NONE)
                  : SPEC == TC_SPEC_GEMM_DSOFTMAX ? static_cast<int>(EPI_DSOFTMAX)
                  : (p.epi == EPI_ROWSTATS ? static_cast<int>(EPI_ROWSTATS) : static_cast<int>(EPI_SOFTMAX));
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // manual 1024-byte alignment (SWIZZLE_128B atoms)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int stages = p.stages;
  const int BN = p.BN;
  const int wgT = kKind == KIND_WGRAD ? p.wg_T : 1;
  // WGRAD halo mode: per 64-channel chunk ONE slab of 64 + 2 pixels (TC_WSLAB_BYTES) serves the 3 taps of a filter row
  const bool wg_halo = kKind == KIND_WGRAD && p.halo;
  const uint32_t b_bytes = wg_halo ? static_cast<uint32_t>(BN >> 6) * TC_WSLAB_BYTES
                                   : static_cast<uint32_t>(BN) * 128u * static_cast<uint32_t>(wgT);  // per stage
  const int acc_cols = BN * wgT;                      // TMEM columns of one accumulator
  const int num_acc = 2 * acc_cols <= TC_TMEM_COLS ? 2 : 1;
  // m-tile pairs (engine_desc.h `mpair`): a stage holds two 16 KiB A tiles
  const bool mpair = kKind == KIND_GEMM && p.mpair != 0;
  const uint32_t a_stage = mpair ? 2u * TC_A_BYTES : static_cast<uint32_t>(TC_A_BYTES);
  uint8_t* sA = smem;
  uint8_t* sB = smem + static_cast<size_t>(stages) * a_stage;
  // halo mode: sA = weight stages, sB = ring of input-row slabs
  const uint32_t wslab_tx = static_cast<uint32_t>(((1 << p.kbw_log2) + 2) * (TC_BLOCK_K >> p.kbw_log2)) * 128u;
  const bool cv_halo = kKind == KIND_CONV && p.halo;
  // stationary B panel: sB = num_k_iters chunks of BN x 64, guarded by sfull_bar[0] / sempty_bar[0]
  const bool bstat = kKind == KIND_GEMM && p.bstat_mt > 0;
  const uint32_t bchunk = static_cast<uint32_t>(BN) * 128u;
  const int nslabs = cv_halo ? p.halo_slabs : (bstat ? 1 : 0);
  uint64_t* bars = reinterpret_cast<uint64_t*>(
      sB + (cv_halo ? static_cast<size_t>(nslabs) * TC_SLAB_BYTES
                    : (bstat ? static_cast<size_t>(p.num_k_iters) * bchunk : static_cast<size_t>(stages) * b_bytes)));
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + stages;
  uint64_t* tfull_bar = bars + 2 * stages;
  uint64_t* tempty_bar = bars + 2 * stages + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * stages + 4);
  uint64_t* sfull_bar = bars + 2 * stages + 6;  // halo slabs: filled by TMA / released by the MMA issuer
  uint64_t* sempty_bar = sfull_bar + nslabs;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_tiles = p.num_m_tiles * p.num_n_tiles * p.num_z;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < stages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], p.colsum != nullptr ? 9 : 1);  // + the 8 epilogue warps that sum the dy tiles
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tfull_bar[i], 1);
      mbar_init(&tempty_bar[i], kEpiThreads);
    }
    for (int i = 0; i < nslabs; ++i) {
      mbar_init(&sfull_bar[i], 1);
      mbar_init(&sempty_bar[i], 1);
    }
    fence_barrier_init();
  }
  __syncwarp();
  if (warp == 0) {
    tmem_alloc(tmem_slot, TC_TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Programmatic dependent launch (tc_engine_inst.cuh): everything above overlapped the previous kernel's tail. Let the next
  // kernel of the stream be scheduled as SMs free up, then wait until the previous grid has completed and its writes are visible
  // (both are no-ops when the launch did not carry the attribute).
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");

  if (warp == 0) {
    // ============================ TMA producer ============================
    if (lane == 0 && cv_halo) {
      // halo conv: per tile (256 consecutive pixels of one image row, starting at tc.x0) and 64-channel chunk, 3 input-row slabs, each followed by the weight
      // tiles of its 3 horizontally shifted taps -- the same order the MMA issuer consumes them in
      int ws = 0, ss = 0;
      uint32_t wph = 0, sph = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const TileCoord tc = decode_tile(p, kKind, kSwap, t);
        const int m0 = tc.m_tile * TC_BLOCK_M;
        for (int cc = 0; cc < p.cchunks; ++cc) {
          for (int ry = 0; ry < 3; ++ry) {
            mbar_wait(&sempty_bar[ss], sph ^ 1u);
            uint64_t* sb = &sfull_bar[ss];
            mbar_arrive_expect_tx(sb, 2u * TC_HALO_BOX * 128u);
            uint8_t* dst = sB + static_cast<size_t>(ss) * TC_SLAB_BYTES;
            if (p.halo == 2) {
              // W = 128: rows (y0 + ry - 1, y0 + ry), pixels -1 .. 128, pixel-interleaved; the view's coordinates are (c, y, x, b)
              tma_load_4d(dst, &tmA, sb, cc * 64, tc.y0 + ry - 1, -1, tc.b);
            } else {
              tma_load_4d(dst, &tmA, sb, cc * 64, tc.x0 - 1, tc.y0 + ry - 1, tc.b);
              tma_load_4d(dst + TC_HALO_BOX * 128, &tmA, sb, cc * 64, tc.x0 + TC_HALO_BOX - 1, tc.y0 + ry - 1, tc.b);
            }
            if (++ss == nslabs) {
              ss = 0;
              sph ^= 1u;
            }
            for (int rx = 0; rx < 3; ++rx) {
              mbar_wait(&empty_bar[ws], wph ^ 1u);
              uint64_t* fb = &full_bar[ws];
              mbar_arrive_expect_tx(fb, TC_A_BYTES);
              tma_load_2d(sA + static_cast<size_t>(ws) * TC_A_BYTES, &tmB, fb, ((ry * 3 + rx) * p.cchunks + cc) * TC_BLOCK_K, m0);
              if (++ws == stages) {
                ws = 0;
                wph ^= 1u;
              }
            }
          }
        }
      }
    } else if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      uint32_t bph = 0;
      for (int s = 0;; ++s) {
        int mi;
        const int t = sched_tile(p, total_tiles, s, mi);
        if (t < 0) break;
        const TileCoord tc = decode_tile(p, kKind, kSwap, t);
        const int m0 = tc.m_tile * (mpair ? 2 * TC_BLOCK_M : TC_BLOCK_M);
        const int n0 = tc.n_tile * BN;
        const int nk = tile_k_iters(p, kKind, tc);
        if (bstat && mi == 0) {
          // new run: the panel may be overwritten once every MMA of the previous run has retired
          mbar_wait(&sempty_bar[0], bph ^ 1u);
          mbar_arrive_expect_tx(&sfull_bar[0], static_cast<uint32_t>(nk) * bchunk);
          for (int kit = 0; kit < nk; ++kit)
            tma_load_3d(sB + static_cast<size_t>(kit) * bchunk, &tmB, &sfull_bar[0], kit * TC_BLOCK_K, n0, tc.z);
          bph ^= 1u;
        }
        for (int kit = 0; kit < nk; ++kit) {
          mbar_wait(&empty_bar[stage], phase ^ 1u);
          uint64_t* fb = &full_bar[stage];
          mbar_arrive_expect_tx(fb, wg_halo ? TC_A_BYTES + static_cast<uint32_t>(BN >> 6) * wslab_tx
                                            : (bstat ? a_stage : a_stage + b_bytes));
          uint8_t* a_dst = sA + static_cast<size_t>(stage) * a_stage;
          uint8_t* b_dst = sB + static_cast<size_t>(stage) * b_bytes;
          if (kKind == KIND_GEMM) {
            const int k0 = kit * TC_BLOCK_K;
            for (int h = 0; h < (mpair ? 2 : 1); ++h) {
              uint8_t* ah = a_dst + h * TC_A_BYTES;
              const int mh = m0 + h * TC_BLOCK_M;
              if (!p.a_mn) {
                tma_load_3d(ah, &tmA, fb, k0, mh, tc.z);
              } else {
                tma_load_3d(ah, &tmA, fb, mh, k0, tc.z);
                tma_load_3d(ah + 8192, &tmA, fb, mh + 64, k0, tc.z);
              }
            }
            if (bstat) {
              // B is the run's stationary panel
            } else if (!p.b_mn) {
              tma_load_3d(b_dst, &tmB, fb, k0, n0, tc.z);
            } else {
              for (int j = 0; j < (BN >> 6); ++j) tma_load_3d(b_dst + j * 8192, &tmB, fb, n0 + 64 * j, k0, tc.z);
            }
          } else if (kKind == KIND_CONV) {
            const int tap = kit / p.cchunks;
            const int cc = kit - tap * p.cchunks;
            if (!kSwap) {
              spatial_load(a_dst, &tmA, fb, p.sp_rank, cc * 64 + p.tap_c[tap], tc.x0 + p.tap_dx[tap], p.tap_ph[tap],
                           tc.y0 + p.tap_dy[tap], tc.b);
              tma_load_2d(b_dst, &tmB, fb, kit * TC_BLOCK_K, n0);
            } else {
              // A slot <- 128 output channels of the weight matrix; B slot <- two 128-pixel boxes (N = 256 pixels)
              tma_load_2d(a_dst, &tmB, fb, kit * TC_BLOCK_K, m0);
              int b1, y1, x1;
              decode_pixel_box(p, 2 * tc.n_tile + 1, b1, y1, x1);
              spatial_load(b_dst, &tmA, fb, p.sp_rank, cc * 64 + p.tap_c[tap], tc.x0 + p.tap_dx[tap], p.tap_ph[tap],
                           tc.y0 + p.tap_dy[tap], tc.b);
              spatial_load(b_dst + TC_A_BYTES, &tmA, fb, p.sp_rank, cc * 64 + p.tap_c[tap], x1 + p.tap_dx[tap],
                           p.tap_ph[tap], y1 + p.tap_dy[tap], b1);
            }
          } else {  // KIND_WGRAD
            const int q = tc.split * p.kch_per_split + kit;
            const int per_img = p.kch_x * p.kch_y;
            const int b = q / per_img;
            const int r = q - b * per_img;
            const int cy = r / p.kch_x, cx = r - cy * p.kch_x;
            const int x0 = cx << p.kbw_log2;
            const int y0 = cy * (TC_BLOCK_K >> p.kbw_log2);
            tma_load_4d(a_dst, &tmA, fb, m0, x0, y0, b);
            tma_load_4d(a_dst + 8192, &tmA, fb, m0 + 64, x0, y0, b);
            if (wg_halo) {
              // one slab per 64-channel chunk: pixels x0 - 1 .. x0 + 64 of input row y0 + dy (zero-filled outside)
              for (int j = 0; j < (BN >> 6); ++j)
                tma_load_4d(b_dst + j * TC_WSLAB_BYTES, &tmB, fb, n0 + 64 * j, x0 - 1, y0 + p.tap_dy[tc.tap], b);
            }
            // wg_T tap tiles back to back: [tap][BN/64 chunks of 64 channels x 64 pixels]
            for (int tt = 0; !wg_halo && tt < wgT; ++tt) {
              const int tap = tc.tap + tt;
              for (int j = 0; j < (BN >> 6); ++j)
                spatial_load(b_dst + (tt * (BN >> 6) + j) * 8192, &tmB, fb, p.sp_rank, n0 + 64 * j + p.tap_c[tap],
                             x0 + p.tap_dx[tap], p.tap_ph[tap], y0 + p.tap_dy[tap], b);
            }
          }
          if (++stage == stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ============================ MMA issuer ============================
    if (lane == 0) {
      // one accumulator = acc_cols columns, written by MMAs of N <= 256 over consecutive 64-wide B chunks
      const int n_first = acc_cols > 256 ? 256 : acc_cols;
      const int n_rest = acc_cols - n_first;
      const uint32_t idesc = make_idesc_16(TC_BLOCK_M, static_cast<uint32_t>(n_first), p.a_mn, p.b_mn, p.a_fmt, p.b_fmt);
      const uint32_t idesc2 = make_idesc_16(TC_BLOCK_M, static_cast<uint32_t>(n_rest > 0 ? n_rest : 16), p.a_mn, p.b_mn, p.a_fmt, p.b_fmt);
      const uint32_t idesc_h = make_idesc_16(TC_BLOCK_M, 192u, p.a_mn, p.b_mn, p.a_fmt, p.b_fmt);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      int ss = 0;
      uint32_t sph = 0;
      for (int t = blockIdx.x; cv_halo && t < total_tiles; t += gridDim.x) {
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1u);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * acc_cols);
        uint32_t accum = 0;
        for (int it = 0; it < 3 * p.cchunks; ++it) {  // (channel chunk, input row) slabs
          mbar_wait(&sfull_bar[ss], sph);
          tc_fence_after();
          const uint32_t slab = smem_u32(sB + static_cast<size_t>(ss) * TC_SLAB_BYTES);
          for (int rx = 0; rx < 3; ++rx) {
            mbar_wait(&full_bar[stage], phase);
            tc_fence_after();
            const uint32_t a_base = smem_u32(sA + static_cast<size_t>(stage) * TC_A_BYTES);
            // pixel x + (rx - 1) <-> slab row x + rx (halo == 2: two interleaved image rows, slab row 2 (x + rx) + r)
            const uint32_t b_base = slab + static_cast<uint32_t>(rx) * (p.halo == 2 ? 256u : 128u);
#pragma unroll
            for (int k = 0; k < TC_BLOCK_K / 16; ++k) {
              umma_bf16(d_tmem, make_smem_desc_sw128(a_base + k * 32, 0, 1024), make_smem_desc_sw128(b_base + k * 32, 0, 1024),
                        idesc, accum);
              accum = 1u;
            }
            umma_commit(&empty_bar[stage]);
            if (++stage == stages) {
              stage = 0;
              phase ^= 1u;
            }
          }
          umma_commit(&sempty_bar[ss]);
          if (++ss == nslabs) {
            ss = 0;
            sph ^= 1u;
          }
        }
        umma_commit(&tfull_bar[acc]);
        if (++acc == num_acc) {
          acc = 0;
          acc_phase ^= 1u;
        }
      }
      uint32_t bph = 0;
      for (int s = 0; !cv_halo; ++s) {
        int mi;
        const int t = sched_tile(p, total_tiles, s, mi);
        if (t < 0) break;
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1u);
        if (mpair) mbar_wait(&tempty_bar[1], acc_phase ^ 1u);  // a pair tile owns both accumulators (acc stays 0)
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + static_cast<uint32_t>(acc * acc_cols);
        const TileCoord tcm = decode_tile(p, kKind, kSwap, t);
        const int nk = tile_k_iters(p, kKind, tcm);
        if (bstat && mi == 0) {
          mbar_wait(&sfull_bar[0], bph);
          tc_fence_after();
          bph ^= 1u;
        }
        for (int kit = 0; kit < nk; ++kit) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_base = smem_u32(sA + static_cast<size_t>(stage) * a_stage);
          const uint32_t b_base = bstat ? smem_u32(sB + static_cast<size_t>(kit) * bchunk)
                                        : smem_u32(sB + static_cast<size_t>(stage) * b_bytes);
#pragma unroll
          for (int k = 0; k < TC_BLOCK_K / 16; ++k) {
            const uint64_t adesc = p.a_mn ? make_smem_desc_sw128(a_base + k * 2048, 8192, 1024)
                                          : make_smem_desc_sw128(a_base + k * 32, 0, 1024);
            const uint64_t bdesc = p.b_mn ? make_smem_desc_sw128(b_base + k * 2048, 8192, 1024)
                                          : make_smem_desc_sw128(b_base + k * 32, 0, 1024);
            if (wg_halo) {
              // per 64-channel chunk one N = 192 MMA: the three 64-wide N chunks are the same channels read 0 / 1 / 2
              // pixel rows into the slab (LBO = 128 B), i.e. the taps dx = -1, 0, +1
              // K step k = output pixels 16k .. 16k + 15 of the chunk = slab rows (row r, x offset xo) + tap shift
              const int r16 = (16 * k) >> p.kbw_log2, xo = (16 * k) & ((1 << p.kbw_log2) - 1);
              const uint32_t koff = static_cast<uint32_t>(r16 * ((1 << p.kbw_log2) + 2) + xo) * 128u;
              for (int j = 0; j < (BN >> 6); ++j) {
                const uint64_t bd = make_smem_desc_sw128(b_base + j * TC_WSLAB_BYTES + koff, 128, 1024);
                umma_bf16(d_tmem + static_cast<uint32_t>(j * 192), adesc, bd, idesc_h, (kit | k) != 0 ? 1u : 0u);
              }
              continue;
            }
            umma_bf16(d_tmem, adesc, bdesc, idesc, (kit | k) != 0 ? 1u : 0u);
            if (mpair) {  // rows 128..255 of the pair: second A tile of the stage, same B, second accumulator
              const uint32_t a1 = a_base + TC_A_BYTES;
              const uint64_t adesc1 = p.a_mn ? make_smem_desc_sw128(a1 + k * 2048, 8192, 1024)
                                             : make_smem_desc_sw128(a1 + k * 32, 0, 1024);
              umma_bf16(d_tmem + static_cast<uint32_t>(acc_cols), adesc1, bdesc, idesc, (kit | k) != 0 ? 1u : 0u);
            }
            if (n_rest > 0) {  // WGRAD tap groups only (MN-major B): chunks 4.. of the stage
              const uint64_t bdesc2 = make_smem_desc_sw128(b_base + 4 * 8192 + k * 2048, 8192, 1024);
              umma_bf16(d_tmem + 256u, adesc, bdesc2, idesc2, (kit | k) != 0 ? 1u : 0u);
            }
          }
          umma_commit(&empty_bar[stage]);  // frees this smem stage once the MMAs above retire
          if (++stage == stages) {
            stage = 0;
            phase ^= 1u;
          }
        }
        umma_commit(&tfull_bar[acc]);  // accumulator complete -> epilogue
        if (mpair) {
          umma_commit(&tfull_bar[1]);
          acc_phase ^= 1u;
          continue;
        }
        if (bstat && mi == p.bstat_mt - 1) umma_commit(&sempty_bar[0]);  // run complete -> the panel may be replaced
        if (++acc == num_acc) {
          acc = 0;
          acc_phase ^= 1u;
        }
      }
    }
    __syncwarp();
  } else if (warp >= 2) {
    // ============================ epilogue (8 warps) ============================
    // A warp may only touch the TMEM lane quadrant (warp id % 4); two warps share a quadrant and split the
    // accumulator's columns in halves, so all four SM sub-partitions carry two epilogue warps each.
    const int q = warp & 3;         // TMEM lane quadrant this warp may access
    const int r = q * 32 + lane;    // row within the tile
    // (`half` = column part: 0..1 with 8 epilogue warps, 0..3 with the 16 of the attention-score instances)
    constexpr int kParts = tc_epi_warps(SPEC) / 4;
    const int half = (warp - 2) >> 2;
    const bool split_cols = (acc_cols % (16 * kParts)) == 0;
    const int nhalf = split_cols ? kParts : 1;
    const int c_begin = split_cols ? half * (acc_cols / kParts) : 0;
    const int c_end = split_cols ? c_begin + (acc_cols / kParts) : (half == 0 ? acc_cols : 0);
    int acc = 0;
    uint32_t acc_phase = 0;
    int cs_stage = 0;
    uint32_t cs_phase = 0;
    for (int s2 = 0;; ++s2) {
      // m-tile pairs: two passes per scheduled tile, rows 0..127 from accumulator 0, rows 128..255 from accumulator 1
      const int s = mpair ? (s2 >> 1) : s2;
      const int mh = mpair ? (s2 & 1) : 0;
      int mi;
      const int t = sched_tile(p, total_tiles, s, mi);
      if (t < 0) break;
      const TileCoord tc = decode_tile(p, kKind, kSwap, t);
      const int m0 = (mpair ? 2 * tc.m_tile + mh : tc.m_tile) * TC_BLOCK_M;
      const int n0 = tc.n_tile * BN;

      if (kKind == KIND_WGRAD && p.colsum != nullptr) {
        // ---- fused bias gradient: column sums of the dy tiles (A operand, MN-major SW128: pixel row k at k * 128 B,
        // 16-byte chunk (channel / 8) XOR (k % 8)); thread = (channel, pixel half), fixed summation order ----
        const int nk = tile_k_iters(p, kKind, tc);
        const int ngroups = p.ntaps / wgT;
        const int grp = tc.tap / wgT;
        const int ew = warp - 2;
        const int ch = (ew & 3) * 32 + lane;
        const int ph = ew >> 2;
        const int cc = ch & 63;
        float csum = 0.0f;
        for (int kit = 0; kit < nk; ++kit) {
          mbar_wait(&full_bar[cs_stage], cs_phase);
          if (tc.n_tile == 0 && (kit % ngroups) == grp) {
            const uint8_t* a = sA + static_cast<size_t>(cs_stage) * TC_A_BYTES + (ch >> 6) * 8192;
#pragma unroll 8
            for (int i = 0; i < 32; ++i) {
              const int k = ph * 32 + i;
              const uint16_t raw = *reinterpret_cast<const uint16_t*>(a + k * 128 + ((((cc >> 3) ^ (k & 7)) << 4) | ((cc & 7) << 1)));
              csum += unpack_16x2(static_cast<uint32_t>(raw), p.a_fmt == FMT_F16).x;
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&empty_bar[cs_stage]);
          if (++cs_stage == stages) {
            cs_stage = 0;
            cs_phase ^= 1u;
          }
        }
        if (tc.n_tile == 0 && m0 + ch < p.colsum_cols)
          p.colsum[(static_cast<long long>(grp * p.split_k + tc.split) * 2 + ph) * p.colsum_cols + m0 + ch] = csum;
      }

      bool row_ok = true;
      long long row_off = 0;  // element offset of (row, column n0) in D / residual
      int yy = 0, xx = 0;
      if (kKind == KIND_GEMM) {
        row_ok = (m0 + r) < p.m_valid;
        row_off = static_cast<long long>(tc.z) * p.strideD + static_cast<long long>(m0 + r) * p.ldd + n0;
      } else if (kKind == KIND_CONV) {
        yy = tc.y0 + (r >> p.bw_log2);
        xx = tc.x0 + (r & ((1 << p.bw_log2) - 1));
        row_off = static_cast<long long>(tc.b) * p.oB + static_cast<long long>(yy) * p.oY +
                  static_cast<long long>(xx) * p.oX + p.o0 + n0;
      } else {
        row_ok = (m0 + r) < p.m_valid;
        row_off = ((static_cast<long long>(tc.tap) * p.split_k + tc.split) * p.m_valid + (m0 + r)) * p.ldd + n0;
      }
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + static_cast<uint32_t>(acc * acc_cols);
      const long long row_off0 = row_off;
      const long long tap_stride = static_cast<long long>(p.split_k) * p.m_valid * p.ldd;  // WGRAD: next tap's slab

      if (kKind == KIND_CONV && kSwap) {
        // ---- swapped-role conv: lane = output channel, column = output pixel. This warp owns the 128 pixels of
        // pixel box (2 * n_tile + half); for every pixel the 32 lanes write 32 consecutive channels (one 64-byte
        // segment), and bias / per-channel work is a per-thread scalar.
        // halo == 2 (W = 128, two interleaved rows): accumulator column n = 2 x + row; this warp's 128 columns are pixels
        // x = 64 half .. 64 half + 63 of rows pb_y and pb_y + 1
        const bool il_rt = p.halo == 2;
        int pb_b, pb_y, pb_x;
        decode_pixel_box(p, il_rt ? 2 * tc.n_tile : 2 * tc.n_tile + half, pb_b, pb_y, pb_x);
        if (il_rt) pb_x = 64 * half;
        const int n = m0 + r;
        const bool ch_ok = n < p.n_valid;
        const float bias = (p.bias != nullptr && ch_ok) ? __ldg(p.bias + n) : 0.0f;
        const bool halfp_rt = p.out_mode == OUT_F16;
        uint16_t* D16 = reinterpret_cast<uint16_t*>(p.D);
        const uint16_t* R16 = kGnb ? nullptr : reinterpret_cast<const uint16_t*>(p.residual);
        const long long base0 = static_cast<long long>(pb_b) * p.oB + p.o0 + n;
        const long long sx_rt = p.oX;
        const long long sx = sx_rt;
        float gs0 = 0.0f, gs1 = 0.0f, gq0 = 0.0f, gq1 = 0.0f;  // fused GroupNorm statistics (two chains each)
        // fused GroupNorm-backward sums (engine_desc.h gnb_*): this thread's channel constants for the tile's image
        const bool gnb = kGnb && ch_ok;
        const uint16_t* X16 = reinterpret_cast<const uint16_t*>(p.gnb_x);
        __half2 b_muh, b_rs, b_k0, b_hg, b_hb;
        float bs0 = 0.0f, bs1 = 0.0f;
        if (gnb) {
          const int G = p.n_valid / p.gnb_cpg;
          const float* st2 = p.gnb_stats + (static_cast<long long>(pb_b) * G + n / p.gnb_cpg) * 2;
          const float mean = __ldg(st2), rstd = __ldg(st2 + 1);
          const float muh = __half2float(__float2half_rn(mean));
          b_muh = __float2half2_rn(muh);
          b_rs = __float2half2_rn(rstd);
          b_k0 = __float2half2_rn((muh - mean) * rstd);
          b_hg = __float2half2_rn(0.5f * __ldg(p.gnb_gamma + n));
          b_hb = __float2half2_rn(0.5f * __ldg(p.gnb_beta + n));
        }
        // staged path: the warp's 32 pixels x 32 channels go through a private 2 KiB buffer ([pixel][32 channels], 64-byte
        // rows) and leave as 16-byte vectors: lane l, vector k covers pixel (l >> 2) + 8 k, channels 8 (l & 3) .. + 7
        uint8_t* stg = (!kGnb && p.epi_off) ? smem + p.epi_off + (warp - 2) * 2048 : nullptr;
        const int vch = m0 + q * 32 + (lane & 3) * 8;  // first channel of this lane's vectors
        uint4 rv[4] = {};  // residual vectors of the batch about to be emitted
        auto load_rv = [&](int pi0) {
          const int yy = pb_y + (pi0 >> p.bw_log2), xx = pb_x + (pi0 & ((1 << p.bw_log2) - 1));
          const long long vb = static_cast<long long>(pb_b) * p.oB + p.o0 + vch + static_cast<long long>(yy) * p.oY +
                               static_cast<long long>(xx + (lane >> 2)) * sx;
          if (vch < p.n_valid) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              rv[k] = __ldg(reinterpret_cast<const uint4*>(R16 + vb + static_cast<long long>(8 * k) * sx));
          }
        };
        // `fast` (std::true_type): IEEE fp16 output (and residual) with a pixel stride of exactly 128 elements, direct path.
        // Both become compile-time constants, so the 2-byte accesses of a batch use ONE base register + immediate offsets
        // and the format select disappears: 3 issue slots per element instead of 7.5 (6 instead of 13.5 with a residual;
        // SASS of the generic path: predicated bf16 AND fp16 packs with four FMNMX clamps per pair, IADD3 + IADD3.X per access).
        auto emit = [&](uint32_t (&v)[32], int pi0, auto fast, auto ilv) {
          constexpr bool kFast = decltype(fast)::value;
          constexpr bool kIl = decltype(ilv)::value;  // interleaved rows (only together with kFast)
          const bool halfp = kFast ? true : halfp_rt;
          const long long sx = kFast ? 128ll : sx_rt;
          const int yy = kIl ? pb_y : pb_y + (pi0 >> p.bw_log2);
          const int xx = kIl ? pb_x + (pi0 >> 1) : pb_x + (pi0 & ((1 << p.bw_log2) - 1));
          const long long base = base0 + static_cast<long long>(yy) * p.oY + static_cast<long long>(xx) * sx;
          const long long base1 = base + p.oY;  // kIl: odd columns are the pixels of the second row
          // element i of the batch: next pixel in x, or (x + i / 2, row i % 2) when the rows are interleaved
          auto at = [&](int i) -> long long { return kIl ? ((i & 1) ? base1 : base) + (i >> 1) * sx : base + i * sx; };
          if (p.out_mode == OUT_NCHW_F32) {
            // thin output conv (decoder.conv_out, Cout = 3 padded to the 128-row weight tile): lane = channel, the 32
            // pixels of the batch are consecutive in x -> 128 contiguous bytes of the NCHW fp32 result per valid lane
            if (ch_ok) {
              float* dstf = reinterpret_cast<float*>(p.D) +
                            ((static_cast<long long>(pb_b) * p.n_valid + n) * p.out_h + yy) * p.out_w + xx;
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                uint32_t ow[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) ow[j] = __float_as_uint(fmaf(__uint_as_float(v[8 * k + j]), p.alpha, bias));
                st_global_256(dstf + 8 * k, ow);
              }
            }
          } else if (stg != nullptr) {
            const long long vbase = static_cast<long long>(pb_b) * p.oB + p.o0 + vch + static_cast<long long>(yy) * p.oY +
                                    static_cast<long long>(xx + (lane >> 2)) * sx;
            float f[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] = fmaf(__uint_as_float(v[i]), p.alpha, bias);
            if (R16 != nullptr) {
              // residual: 4 coalesced 16-byte vectors per lane (requested one batch ahead, see load_rv) -> staging
              // buffer -> this lane's channel column
#pragma unroll
              for (int k = 0; k < 4; ++k)
                *reinterpret_cast<uint4*>(stg + ((lane >> 2) + 8 * k) * 64 + (lane & 3) * 16) = rv[k];
              __syncwarp();
              if (pi0 + 32 < 128) load_rv(pi0 + 32);
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                const uint16_t rr = *reinterpret_cast<const uint16_t*>(stg + i * 64 + lane * 2);
                f[i] += unpack_16x2(static_cast<uint32_t>(rr), halfp).x;
              }
              __syncwarp();
            }
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              const uint32_t pk = pack_16x2(f[i], f[i + 1], halfp);
              *reinterpret_cast<uint16_t*>(stg + i * 64 + lane * 2) = static_cast<uint16_t>(pk & 0xffffu);
              *reinterpret_cast<uint16_t*>(stg + (i + 1) * 64 + lane * 2) = static_cast<uint16_t>(pk >> 16);
              if (p.gn_part != nullptr && ch_ok) {
                const float2 r = unpack_16x2(pk, halfp);
                gs0 += r.x;
                gq0 = fmaf(r.x, r.x, gq0);
                gs1 += r.y;
                gq1 = fmaf(r.y, r.y, gq1);
              }
            }
            __syncwarp();
            if (vch < p.n_valid) {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const uint4 ov = *reinterpret_cast<const uint4*>(stg + ((lane >> 2) + 8 * k) * 64 + (lane & 3) * 16);
                *reinterpret_cast<uint4*>(D16 + vbase + static_cast<long long>(8 * k) * sx) = ov;
              }
            }
            __syncwarp();
          } else if (ch_ok) {
            float f[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) f[i] = fmaf(__uint_as_float(v[i]), p.alpha, bias);
            uint16_t rr[32];  // residual values, or (gnb) the GroupNorm input at the same positions
            if (R16 != nullptr) {
#pragma unroll
              for (int i = 0; i < 32; ++i) rr[i] = __ldg(R16 + at(i));
#pragma unroll
              for (int i = 0; i < 32; ++i) f[i] += unpack_16x2(static_cast<uint32_t>(rr[i]), halfp).x;
            } else if (kFast && gnb) {
#pragma unroll
              for (int i = 0; i < 32; ++i) rr[i] = __ldg(X16 + at(i));
            }
            __half2 ba0 = __float2half2_rn(0.0f), ba1 = __float2half2_rn(0.0f);
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              const uint32_t pk = pack_16x2(f[i], f[i + 1], halfp);
              D16[at(i)] = static_cast<uint16_t>(pk & 0xffffu);
              D16[at(i + 1)] = static_cast<uint16_t>(pk >> 16);
              if (!kGnb && p.gn_part != nullptr) {  // statistics of the STORED (rounded) values, as the separate pass sees them
                const float2 r = unpack_16x2(pk, halfp);
                gs0 += r.x;
                gq0 = fmaf(r.x, r.x, gq0);
                gs1 += r.y;
                gq1 = fmaf(r.y, r.y, gq1);
              }
              if (kFast && gnb) {
                // GroupNorm backward sums on the pair (two pixels of this thread's channel), dy = the stored pair
                const uint32_t xw = static_cast<uint32_t>(rr[i]) | (static_cast<uint32_t>(rr[i + 1]) << 16);
                const __half2 xh = __hfma2(__hsub2(*reinterpret_cast<const __half2*>(&xw), b_muh), b_rs, b_k0);
                const __half2 uh = __hfma2(xh, b_hg, b_hb);
                uint32_t tw, uw = *reinterpret_cast<const uint32_t*>(&uh);
                asm("tanh.approx.f16x2 %0, %1;" : "=r"(tw) : "r"(uw));
                const __half2 t = *reinterpret_cast<const __half2*>(&tw);
                const __half2 wq = __hfma2(uh, __hfma2(__hneg2(t), t, __float2half2_rn(1.0f)), t);
                const __half2 inv = __float2half2_rn(1.0f / 128.0f);
                const __half2 ds = __hmul2(*reinterpret_cast<const __half2*>(&pk), __hfma2(wq, inv, inv));  // dy' / 64
                ba0 = __hadd2(ba0, ds);
                ba1 = __hfma2(ds, xh, ba1);
                if ((i & 6) == 6) {  // every 4 pairs: flush the half2 partial sums into fp32
                  const float2 f0 = __half22float2(ba0), f1 = __half22float2(ba1);
                  bs0 += f0.x + f0.y;
                  bs1 += f1.x + f1.y;
                  ba0 = ba1 = __float2half2_rn(0.0f);
                }
              }
            }
          }
        };
        const uint16_t* P16 = kGnb ? X16 : R16;  // what the epilogue will read per element: residual, or the GroupNorm input
        if (P16 != nullptr) {
          if (stg != nullptr) load_rv(0);  // in flight while the accumulator completes
          if ((kGnb || p.res_prefetch) && ch_ok) {
            // this warp's residual columns for the whole tile (staged path: the three later 32-pixel batches): pull their lines
            // into L2 now -- the tile's MMAs take ~6 us -- so that the epilogue's loads see L2 latency, not DRAM latency
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (j == 0 && stg != nullptr) continue;
              const int pi = 32 * j + lane;
              const int yy = il_rt ? pb_y + (pi & 1) : pb_y + (pi >> p.bw_log2);
              const int xx = il_rt ? pb_x + (pi >> 1) : pb_x + (pi & ((1 << p.bw_log2) - 1));
              const uint16_t* a = P16 + static_cast<long long>(pb_b) * p.oB + p.o0 + (m0 + q * 32) +
                                  static_cast<long long>(yy) * p.oY + static_cast<long long>(xx) * sx;
              asm volatile("prefetch.global.L2 [%0];" ::"l"(a));
            }
          }
        }
        auto drain = [&](auto fast, auto ilv) {
          mbar_wait(&tfull_bar[acc], acc_phase);
          tc_fence_after();
          uint32_t va[32], vb[32];
          tmem_ld_x32(t_row + static_cast<uint32_t>(c_begin), va);
          tmem_ld_wait();
          tmem_ld_x32(t_row + static_cast<uint32_t>(c_begin + 32), vb);
          emit(va, 0, fast, ilv);
          tmem_ld_wait();
          tmem_ld_x32(t_row + static_cast<uint32_t>(c_begin + 64), va);
          emit(vb, 32, fast, ilv);
          tmem_ld_wait();
          tmem_ld_x32(t_row + static_cast<uint32_t>(c_begin + 96), vb);
          emit(va, 64, fast, ilv);
          tmem_ld_wait();
          emit(vb, 96, fast, ilv);
        };
        if (TC_FAST_EPI && il_rt)  // (the host turns halo == 2 on only where the fast path's conditions hold)
          drain(std::true_type{}, std::true_type{});
        else if (kGnb || (TC_FAST_EPI && halfp_rt && sx_rt == 128 && stg == nullptr && p.out_mode != OUT_NCHW_F32))
          drain(std::true_type{}, std::false_type{});  // (the host picks the GNB instance only under these conditions)
        else
          drain(std::false_type{}, std::false_type{});
        if (!kGnb && p.gn_part != nullptr && ch_ok) {
          const int boxes = p.tiles_x * p.tiles_y;  // pixel boxes per image
          const int box = 2 * tc.n_tile + half - pb_b * boxes;
          float2* dst = reinterpret_cast<float2*>(p.gn_part) + (static_cast<long long>(pb_b) * boxes + box) * p.n_valid + n;
          *dst = make_float2(gs0 + gs1, gq0 + gq1);
        }
        if (gnb) {
          const int boxes = p.tiles_x * p.tiles_y;
          const int box = 2 * tc.n_tile + half - pb_b * boxes;
          float2* dst = reinterpret_cast<float2*>(p.gnb_part) + (static_cast<long long>(pb_b) * boxes + box) * p.n_valid + n;
          *dst = make_float2(64.0f * bs0, 64.0f * bs1);
        }
      } else if (epi != EPI_NONE) {
        // ---- fused attention epilogues (KIND_GEMM, BN = 256; one accumulator row and kCols columns per thread) ----
        // Global reads this tile needs (softmax partials / delta / the P tile) are issued BEFORE waiting for the
        // accumulator, so their latency hides behind the tile's MMAs; TMEM is drained 32 (16) columns at a time with the
        // next tcgen05.ld in flight while the current chunk is processed.
        constexpr int kCols = 256 / kParts;  // columns per warp (the host checks N % 256 == 0)
        constexpr int kPV = kCols / 16;      // 32-byte vectors of P per thread
        const float s2 = p.alpha * 1.4426950408889634f;  // scores live in the log2 domain: exp(a*s) = 2^(s2*s)
        const bool halfp = p.out_mode == OUT_F16;
        float mx = -INFINITY, l = 0.0f, lse2 = 0.0f, delta = 0.0f;  // mx: running max of the RAW accumulators
        uint32_t pp[kPV][8];  // DSOFTMAX: this thread's kCols P values
        if (epi == EPI_SOFTMAX && row_ok) {
          // lse2[row] = max + log2(sum 2^(s2*s - max)), combined from the row partials by attn_stats_combine_kernel
          lse2 = __ldg(p.aux_f32 + static_cast<long long>(tc.z) * p.m_valid + (m0 + r));
        } else if (epi == EPI_DSOFTMAX && row_ok) {
          delta = __ldg(p.aux_f32 + static_cast<long long>(tc.z) * p.m_valid + (m0 + r));
          const uint16_t* psrc = reinterpret_cast<const uint16_t*>(p.aux16) + row_off + c_begin;
#pragma unroll
          for (int i = 0; i < kPV; ++i) ld_global_nc_256(psrc + 16 * i, pp[i]);
          // P streams from HBM (2 GB per block): pull the NEXT tile's segment of this thread's row into L2 now, a whole
          // tile time ahead of the loads above
          int mi_n;
          const int tn = sched_tile(p, total_tiles, s + 1, mi_n);
          if (tn >= 0) {
            const TileCoord tcn = decode_tile(p, kKind, kSwap, tn);
            if (tcn.m_tile * TC_BLOCK_M + r < p.m_valid)
              prefetch_l2(reinterpret_cast<const uint16_t*>(p.aux16) + static_cast<long long>(tcn.z) * p.strideD +
                          static_cast<long long>(tcn.m_tile * TC_BLOCK_M + r) * p.ldd + tcn.n_tile * BN + c_begin);
          }
        }
        mbar_wait(&tfull_bar[acc], acc_phase);
        tc_fence_after();

        auto process = [&](uint32_t (&v)[32], int cc) {
          if (epi == EPI_ROWSTATS) {
            // online (max, sum 2^(s2*(v - max))) on raw accumulators: s2 > 0, so the max commutes with the scale
            // (four independent chains each: a 31-deep fmax chain left the MUFU pipe idle for its whole latency)
            float c0 = __uint_as_float(v[0]), c1 = __uint_as_float(v[1]), c2 = __uint_as_float(v[2]), c3 = __uint_as_float(v[3]);
#pragma unroll
            for (int i = 4; i < 32; i += 4) {
              c0 = fmaxf(c0, __uint_as_float(v[i]));
              c1 = fmaxf(c1, __uint_as_float(v[i + 1]));
              c2 = fmaxf(c2, __uint_as_float(v[i + 2]));
              c3 = fmaxf(c3, __uint_as_float(v[i + 3]));
            }
            const float mn = fmaxf(mx, fmaxf(fmaxf(c0, c1), fmaxf(c2, c3)));
            const float off = -mn * s2;
            float a0 = 0.0f, a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              a0 += fast_exp2(fmaf(__uint_as_float(v[i]), s2, off));
              a1 += fast_exp2(fmaf(__uint_as_float(v[i + 1]), s2, off));
              a2 += fast_exp2(fmaf(__uint_as_float(v[i + 2]), s2, off));
              a3 += fast_exp2(fmaf(__uint_as_float(v[i + 3]), s2, off));
            }
            l = l * fast_exp2((mx - mn) * s2) + ((a0 + a1) + (a2 + a3));
            mx = mn;
          } else if (row_ok) {
            uint32_t o[16];
            {
#pragma unroll
              for (int i = 0; i < 16; ++i)
                o[i] = pack_16x2(fast_exp2(fmaf(__uint_as_float(v[2 * i]), s2, -lse2)),
                                 fast_exp2(fmaf(__uint_as_float(v[2 * i + 1]), s2, -lse2)), halfp);
            }
            uint16_t* dst = reinterpret_cast<uint16_t*>(p.D) + row_off + cc;
            const uint32_t o_lo[8] = {o[0], o[1], o[2], o[3], o[4], o[5], o[6], o[7]};
            const uint32_t o_hi[8] = {o[8], o[9], o[10], o[11], o[12], o[13], o[14], o[15]};
            st_global_256(dst, o_lo);
            st_global_256(dst + 16, o_hi);
          }
        };
        if (epi == EPI_DSOFTMAX) {
          // dS = P * (dP - delta) * alpha: chunks of 16 columns, the next tcgen05.ld in flight during each chunk
          uint32_t da[16], db[16];
          tmem_ld_x16(t_row + static_cast<uint32_t>(c_begin), da);
#pragma unroll
          for (int j = 0; j < kPV; ++j) {
            uint32_t (&cur)[16] = (j & 1) ? db : da;
            uint32_t (&nxt)[16] = (j & 1) ? da : db;
            tmem_ld_wait();
            if (j < kPV - 1) tmem_ld_x16(t_row + static_cast<uint32_t>(c_begin + 16 * (j + 1)), nxt);
            if (row_ok) {
              const uint32_t (&pw)[8] = pp[j];
              uint32_t o[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const float2 pv = unpack_16x2(pw[i], halfp);
                o[i] = pack_16x2(pv.x * p.alpha * (__uint_as_float(cur[2 * i]) - delta),
                                 pv.y * p.alpha * (__uint_as_float(cur[2 * i + 1]) - delta), halfp);
              }
              st_global_256(reinterpret_cast<uint16_t*>(p.D) + row_off + c_begin + 16 * j, o);
            }
          }
        } else {
          uint32_t va[32], vb[32];
          tmem_ld_x32(t_row + static_cast<uint32_t>(c_begin), va);
#pragma unroll
          for (int j = 0; j < kCols / 32; ++j) {
            uint32_t (&cur)[32] = (j & 1) ? vb : va;
            uint32_t (&nxt)[32] = (j & 1) ? va : vb;
            tmem_ld_wait();
            if (j < kCols / 32 - 1) tmem_ld_x32(t_row + static_cast<uint32_t>(c_begin + 32 * (j + 1)), nxt);
            process(cur, c_begin + 32 * j);
          }
        }
        if (epi == EPI_ROWSTATS && row_ok) {
          float2* dst = reinterpret_cast<float2*>(p.aux_out) +
                        (static_cast<long long>(tc.z) * p.num_n_tiles * nhalf + tc.n_tile * nhalf + half) * p.m_valid +
                        (m0 + r);
          *dst = make_float2(mx * s2, l);  // (max, sum) in the scaled log2 domain
        }
      } else {
      mbar_wait(&tfull_bar[acc], acc_phase);
      tc_fence_after();
      for (int cc = c_begin; cc < c_end; cc += 16) {
        uint32_t v[16];
        tmem_ld_x16(t_row + static_cast<uint32_t>(cc), v);
        tmem_ld_wait();
        int c = cc;  // column within this tap's BN-wide block
        if (wg_halo) {  // column = [channel chunk j][tap][64 channels]
          const int j = cc / 192, rem = cc - j * 192;
          const int tt = rem >> 6;
          c = j * 64 + (rem & 63);
          row_off = row_off0 + tt * tap_stride;
        } else if (wgT > 1) {
          const int tt = cc / BN;
          c = cc - tt * BN;
          row_off = row_off0 + tt * tap_stride;
        }
        const int ncol = n0 + c;  // global column of v[0]
        if (row_ok && ncol < p.n_valid) {
          float f[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) f[i] = __uint_as_float(v[i]) * p.alpha;
          const bool full = (ncol + 16 <= p.n_valid);
          if (p.bias != nullptr) {
            if (full) {
#pragma unroll
              for (int i = 0; i < 16; i += 4) {
                float4 bv = __ldg(reinterpret_cast<const float4*>(p.bias + ncol + i));
                f[i] += bv.x; f[i + 1] += bv.y; f[i + 2] += bv.z; f[i + 3] += bv.w;
              }
            } else {
              for (int i = 0; i < 16; ++i)
                if (ncol + i < p.n_valid) f[i] += __ldg(p.bias + ncol + i);
            }
          }
          if (p.out_mode == OUT_BF16 || p.out_mode == OUT_F16) {
            const bool half = p.out_mode == OUT_F16;  // 16-bit output (and residual) format
            uint16_t* dst = reinterpret_cast<uint16_t*>(p.D) + row_off + c;
            if (full) {
              if (p.residual != nullptr) {
                uint32_t rw[8];
                ld_global_nc_256(reinterpret_cast<const uint16_t*>(p.residual) + row_off + c, rw);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                  float2 rv = unpack_16x2(rw[i], half);
                  f[2 * i] += rv.x;
                  f[2 * i + 1] += rv.y;
                }
              }
              uint32_t ow[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) ow[i] = pack_16x2(f[2 * i], f[2 * i + 1], half);
              st_global_256(dst, ow);
            } else {
              for (int i = 0; i < 16; ++i)
                if (ncol + i < p.n_valid) {
                  float val = f[i];
                  if (p.residual != nullptr)
                    val += unpack_16x2(reinterpret_cast<const uint16_t*>(p.residual)[row_off + c + i], half).x;
                  dst[i] = static_cast<uint16_t>(pack_16x2(val, 0.0f, half) & 0xffffu);
                }
            }
          } else if (p.out_mode == OUT_F32) {
            float* dst = reinterpret_cast<float*>(p.D) + row_off + c;
            if (full) {
              if (p.residual != nullptr) {
                const float4* rp = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.residual) + row_off + c);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                  float4 rv = __ldg(rp + i);
                  f[4 * i] += rv.x; f[4 * i + 1] += rv.y; f[4 * i + 2] += rv.z; f[4 * i + 3] += rv.w;
                }
              }
#pragma unroll
              for (int i = 0; i < 2; ++i) {
                uint32_t ow[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) ow[j] = __float_as_uint(f[8 * i + j]);
                st_global_256(dst + 8 * i, ow);
              }
            } else {
              for (int i = 0; i < 16; ++i)
                if (ncol + i < p.n_valid) {
                  float val = f[i];
                  if (p.residual != nullptr) val += reinterpret_cast<const float*>(p.residual)[row_off + c + i];
                  dst[i] = val;
                }
            }
          } else {  // OUT_NCHW_F32: D[b][n][yy][xx], consecutive lanes -> consecutive xx (coalesced)
            float* base = reinterpret_cast<float*>(p.D);
            for (int i = 0; i < 16; ++i)
              if (ncol + i < p.n_valid) {
                long long off = ((static_cast<long long>(tc.b) * p.n_valid + (ncol + i)) * p.out_h + yy) * p.out_w + xx;
                base[off] = f[i];
              }
          }
        }
      }
      }
      tc_fence_before();
      mbar_arrive(&tempty_bar[acc]);
      if (++acc == num_acc) {
        acc = 0;
        acc_phase ^= 1u;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, TC_TMEM_COLS);
  }
}

}  // namespace odvae
