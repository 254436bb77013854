// Block-sparse GEMM mainloop for sm_100a: one persistent, warp-specialised kernel template
// that serves SDD, DSD and DDS with all transpose flags.
//
// Replaces the reference's BlockGemm (sputnik/block/cutlass/block_gemm.h:749-878), its
// BlockTileAccessIterator / DependentTileAccessIterator (block_tile_access_iterator.h:178-235,
// dependent_tile_access_iterator.h:120-133) and BlockTileOutputIterator
// (block_tile_output_iterator.h:159-199).  Semantics kept: alpha = 1, beta = 0, fp32
// accumulate, round-to-nearest 16-bit output, every output tile fully overwritten (zero
// tiles for empty block rows / columns).
//
// Design (not a translation):
//  * every nonzero 128x128 block is contracted by tcgen05.mma (UMMA 128 x BN x 16) with the
//    accumulator in TMEM, double-buffered so the epilogue of tile i overlaps tile i+1;
//  * operand tiles are staged by TMA (128 B swizzle) whose coordinates come from the block
//    index metadata, which the producer warp prefetches into shared memory 32 blocks at a
//    time (the reference pays a dependent __ldg per block on the critical path);
//  * transposed access (trans_a on a sparse A, non-transposed sparse B) is a TMA gather
//    through block_offsets into an MN-major UMMA descriptor: no transpose is materialised;
//  * CTAs are persistent (grid = #SMs) and walk a tile list.
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "ptx.cuh"

namespace bsmm {

enum : int { kOpSDD = 0, kOpDSD = 1, kOpDDS = 2 };

constexpr int kBlock = 128;       // sparse block edge == UMMA M
constexpr int kBK = 64;           // K elements per pipeline stage (128 B of 16-bit data)
constexpr int kThreads = 256;     // warp 0 TMA, 1 MMA, 2 TMEM alloc, 3 idle, 4-7 epilogue
constexpr int kSmemBudget = 200 * 1024;

struct GemmParams {
  // DSD / DDS: the walk over the sparse operand, already resolved by the host to the row view
  // (offsets, indices, sequential blocks) or the column view (offsets_t, indices_t,
  // block_offsets).
  const int* walk_offsets;
  const short* walk_indices;
  const int* walk_blocks;  // nullptr: block id == position in the walk
  // SDD: coordinates of output block t.
  const short* sdd_rows;
  const short* sdd_cols;
  void* c;
  long long ldc;    // elements
  int m, n;         // dense output extent (clipping)
  int tiles_major;  // SDD: nnz blocks; DSD: block rows of op(A); DDS: block cols of op(B)
  int tiles_minor;  // SDD: 1; DSD: ceil(n / BN); DDS: ceil(m / 128)
  int k_steps;      // SDD: ceil(k / 64)
  uint32_t idesc;
};

template <int BN>
struct Tiling {
  static constexpr int kABytes = kBlock * kBK * 2;
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = kSmemBudget / kStageBytes;
  static constexpr int kTmemCols = 2 * BN;  // two accumulator stages (power of two >= 32)
  // stages + barriers (full, empty, tmem_full[2], tmem_empty[2]) + tmem ptr + index staging
  static constexpr int kAuxBytes = (2 * kStages + 4) * 8 + 16 + 2 * 2 * 32 * 4;
  static constexpr int kSmemBytes = kStages * kStageBytes + kAuxBytes + 1024;  // + align slack
};

template <bool BF16>
__device__ __forceinline__ uint32_t pack2(uint32_t lo_bits, uint32_t hi_bits) {
  float lo = __uint_as_float(lo_bits), hi = __uint_as_float(hi_bits);
  if constexpr (BF16) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  } else {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
}

template <int OP, bool A_MN, bool B_MN, int BN, bool BF16>
__global__ void __launch_bounds__(kThreads, 1)
bsmm_gemm_kernel(const __grid_constant__ CUtensorMap map_a,
                 const __grid_constant__ CUtensorMap map_b, const GemmParams p) {
  using T = Tiling<BN>;
  constexpr int S = T::kStages;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* aux = smem + S * T::kStageBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(aux);
  uint64_t* empty_bar = full_bar + S;
  uint64_t* tfull_bar = empty_bar + S;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  int* s_idx = reinterpret_cast<int*>(tmem_ptr + 4);  // [2][32] dense-side block coordinate
  int* s_blk = s_idx + 64;                            // [2][32] block id in the value array

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_major * p.tiles_minor;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&map_a);
    ptx::prefetch_tensormap(&map_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      ptx::mbar_init(&full_bar[i], 1);
      ptx::mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&tfull_bar[i], 1);
      ptx::mbar_init(&tempty_bar[i], 4);  // one arrive per epilogue warp
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc(tmem_ptr, T::kTmemCols);
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===================================================================== TMA producer
    int stage = 0;
    uint32_t phase = 0;
    int buf = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      const int major = t / p.tiles_minor;
      const int minor = t - major * p.tiles_minor;
      if constexpr (OP == kOpSDD) {
        if (lane == 0) {
          const int r0 = int(__ldg(p.sdd_rows + t)) * kBlock;
          const int c0 = int(__ldg(p.sdd_cols + t)) * kBlock;
          for (int s = 0; s < p.k_steps; ++s) {
            const int k0 = s * kBK;
            ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* sa = smem + stage * T::kStageBytes;
            uint8_t* sb = sa + T::kABytes;
            ptx::mbar_arrive_expect_tx(&full_bar[stage], T::kStageBytes);
            if constexpr (A_MN) {
              ptx::tma_load_2d(&map_a, &full_bar[stage], sa, r0, k0);
              ptx::tma_load_2d(&map_a, &full_bar[stage], sa + 8192, r0 + 64, k0);
            } else {
              ptx::tma_load_2d(&map_a, &full_bar[stage], sa, k0, r0);
            }
            if constexpr (B_MN) {
#pragma unroll
              for (int h = 0; h < BN / 64; ++h)
                ptx::tma_load_2d(&map_b, &full_bar[stage], sb + h * 8192, c0 + h * 64, k0);
            } else {
              ptx::tma_load_2d(&map_b, &full_bar[stage], sb, k0, c0);
            }
            if (++stage == S) { stage = 0; phase ^= 1; }
          }
        }
      } else {
        const int j0 = __ldg(p.walk_offsets + major);
        const int j1 = __ldg(p.walk_offsets + major + 1);
        const int d0 = minor * (OP == kOpDSD ? BN : kBlock);  // dense-side tile origin
        for (int jc = j0; jc < j1; jc += 32) {
          // Prefetch the next 32 block coordinates of this walk into shared memory.
          const int j = jc + lane;
          if (j < j1) {
            s_idx[buf * 32 + lane] = int(__ldg(p.walk_indices + j));
            s_blk[buf * 32 + lane] = p.walk_blocks ? __ldg(p.walk_blocks + j) : j;
          }
          __syncwarp();
          if (lane == 0) {
            const int cnt = min(32, j1 - jc);
            for (int i = 0; i < cnt; ++i) {
              const int blk_row0 = s_blk[buf * 32 + i] * kBlock;
              const int dense_k0 = s_idx[buf * 32 + i] * kBlock;
#pragma unroll
              for (int kh = 0; kh < 2; ++kh) {
                ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                uint8_t* sa = smem + stage * T::kStageBytes;
                uint8_t* sb = sa + T::kABytes;
                ptx::mbar_arrive_expect_tx(&full_bar[stage], T::kStageBytes);
                const int krow = dense_k0 + kh * kBK;
                // Sparse operand: A for DSD, B for DDS. K-major reads the block as stored;
                // MN-major reads it transposed (64-row slabs of the stored block).
                {
                  constexpr bool SP_MN = (OP == kOpDSD) ? A_MN : B_MN;
                  const CUtensorMap* mp = (OP == kOpDSD) ? &map_a : &map_b;
                  uint8_t* dst = (OP == kOpDSD) ? sa : sb;
                  if constexpr (SP_MN) {
                    ptx::tma_load_2d(mp, &full_bar[stage], dst, 0, blk_row0 + kh * kBK);
                    ptx::tma_load_2d(mp, &full_bar[stage], dst + 8192, 64, blk_row0 + kh * kBK);
                  } else {
                    ptx::tma_load_2d(mp, &full_bar[stage], dst, kh * kBK, blk_row0);
                  }
                }
                // Dense operand, gathered at the block's coordinate.
                if constexpr (OP == kOpDSD) {
                  if constexpr (B_MN) {
#pragma unroll
                    for (int h = 0; h < BN / 64; ++h)
                      ptx::tma_load_2d(&map_b, &full_bar[stage], sb + h * 8192, d0 + h * 64,
                                       krow);
                  } else {
                    ptx::tma_load_2d(&map_b, &full_bar[stage], sb, krow, d0);
                  }
                } else {
                  if constexpr (A_MN) {
                    ptx::tma_load_2d(&map_a, &full_bar[stage], sa, d0, krow);
                    ptx::tma_load_2d(&map_a, &full_bar[stage], sa + 8192, d0 + 64, krow);
                  } else {
                    ptx::tma_load_2d(&map_a, &full_bar[stage], sa, krow, d0);
                  }
                }
                if (++stage == S) { stage = 0; phase ^= 1; }
              }
            }
          }
          __syncwarp();
          buf ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================================== MMA issuer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      constexpr uint32_t kAdvA = A_MN ? (2048 >> 4) : (32 >> 4);
      constexpr uint32_t kAdvB = B_MN ? (2048 >> 4) : (32 >> 4);
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        int nsteps;
        if constexpr (OP == kOpSDD) {
          nsteps = p.k_steps;
        } else {
          const int major = t / p.tiles_minor;
          nsteps = 2 * (__ldg(p.walk_offsets + major + 1) - __ldg(p.walk_offsets + major));
        }
        ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);
        ptx::tc_fence_after_sync();
        const uint32_t tmem_d = tmem_base + uint32_t(as * BN);
        if (nsteps == 0) {
          ptx::mbar_arrive(&tfull_bar[as]);  // nothing accumulated: epilogue writes zeros
        }
        for (int s = 0; s < nsteps; ++s) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after_sync();
          const uint32_t sa = ptx::smem_u32(smem + stage * T::kStageBytes);
          const uint32_t sb = sa + T::kABytes;
          const uint64_t da = ptx::make_smem_desc(sa, A_MN ? 8192 : 16, 1024);
          const uint64_t db = ptx::make_smem_desc(sb, B_MN ? 8192 : 16, 1024);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            ptx::umma_f16(tmem_d, da + uint64_t(k * kAdvA), db + uint64_t(k * kAdvB), p.idesc,
                          uint32_t((s | k) != 0));
          }
          ptx::umma_commit(&empty_bar[stage]);
          if (s == nsteps - 1) ptx::umma_commit(&tfull_bar[as]);
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ===================================================================== epilogue
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    int as = 0;
    uint32_t aphase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      const int major = t / p.tiles_minor;
      const int minor = t - major * p.tiles_minor;
      int nsteps, row0, col0;
      long long ld;
      uint16_t* out;
      int row_lim, col_lim;
      if constexpr (OP == kOpSDD) {
        nsteps = p.k_steps;
        out = reinterpret_cast<uint16_t*>(p.c) + size_t(t) * (kBlock * kBlock);
        ld = kBlock; row0 = 0; col0 = 0; row_lim = kBlock; col_lim = kBlock;
      } else {
        nsteps = 2 * (__ldg(p.walk_offsets + major + 1) - __ldg(p.walk_offsets + major));
        out = reinterpret_cast<uint16_t*>(p.c);
        ld = p.ldc; row_lim = p.m; col_lim = p.n;
        if constexpr (OP == kOpDSD) { row0 = major * kBlock; col0 = minor * BN; }
        else                        { row0 = minor * kBlock; col0 = major * kBlock; }
      }
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after_sync();
      const int row = row0 + q * 32 + lane;
      uint16_t* orow = out + size_t(row) * size_t(ld);
      const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * BN);
#pragma unroll 1
      for (int cc = 0; cc < BN / 32; ++cc) {
        uint32_t v[32];
        if (nsteps > 0) {
          ptx::tmem_ld_32x32b_x32(taddr + uint32_t(cc * 32), v);
          ptx::tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) v[i] = 0u;
        }
        if (row < row_lim) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const int col = col0 + cc * 32 + g * 8;
            if (col < col_lim) {  // n % 8 == 0: an 8-column group is all in or all out
              uint4 w;
              w.x = pack2<BF16>(v[g * 8 + 0], v[g * 8 + 1]);
              w.y = pack2<BF16>(v[g * 8 + 2], v[g * 8 + 3]);
              w.z = pack2<BF16>(v[g * 8 + 4], v[g * 8 + 5]);
              w.w = pack2<BF16>(v[g * 8 + 6], v[g * 8 + 7]);
              *reinterpret_cast<uint4*>(orow + col) = w;
            }
          }
        }
      }
      ptx::tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty_bar[as]);
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
  }

  __syncwarp();
  ptx::tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) {
    ptx::tc_fence_after_sync();
    ptx::tmem_dealloc(tmem_base, T::kTmemCols);
  }
}

}  // namespace bsmm
