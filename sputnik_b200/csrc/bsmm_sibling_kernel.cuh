// The sparse-output / sparse x sparse siblings of the path: SSD, SDS and DSS on one persistent,
// warp-specialised tcgen05 kernel.
//
// Replaces the reference's ssd / sds / dss launchers
// (sputnik/block/{ssd,sds,dss}/cutlass/*_mixed_b128_128x128x32x5_{nn,nt,tn,tt}_align8.cu), its
// index intersection (sputnik/block/cutlass/index_merge.h:91-227, block_tile_union_iterator.h)
// and, for the sparse outputs, BlockTileOutputIterator (block_tile_output_iterator.h:159-199).
// Semantics kept: alpha = 1, beta = 0, fp32 accumulate, every output tile fully overwritten.
//
//   SSD  C(sparse) = op(A sparse) op(B)        tile = block t of C; K walk = op(A)'s row of t
//   SDS  C(sparse) = op(A) op(B sparse)        tile = block t of C; K walk = op(B)'s column of t
//   DSS  C(dense)  = op(A sparse) op(B sparse) tile = (block row i, block column j) of C;
//                                              K walk = row i of op(A)  INTERSECTED  with
//                                              column j of op(B)
//
// Design (not a translation):
//  * the reference intersects 256-bit K-contiguous bitmasks per CTA (built by a host round trip,
//    bitmask.cu:7-45) and recovers list positions with a popcount prefix per index; here the
//    producer warp scatters the N-side walk of the tile into a shared-memory position table
//    (block id + 1 per K block), then streams the M-side walk through it 32 entries at a time
//    (ballot + prefix popcount compaction).  No bitmask, no K limit of 32768, unordered index
//    lists are fine;
//  * the surviving (M-side block, N-side block) pairs drive TMA directly: a sparse operand is a
//    2-D [nnz_blocks * 128][128] array, read as stored (K-major or MN-major UMMA descriptor —
//    the index-indirected gather, nothing is transposed in memory);
//  * SSD pairs blocks t, t + 1 of the same block row into one 128 x 256 UMMA tile (they share
//    the sparse walk), exactly like the SDD kernel; SDS / DSS tiles are 128 x 128;
//  * producer -> MMA / epilogue hand-off carries the tile id, its K-step count and the pair
//    flag through a small shared-memory queue (the count of a DSS tile is only known after
//    the intersection).
#pragma once

#include "bsmm_kernel.cuh"

namespace bsmm {

enum : int { kSibSSD = 0, kSibSDS = 1, kSibDSS = 2 };

constexpr int kSibMaxKBlocks = 2048;  // DSS position table: K <= 262144

struct SiblingParams {
  int op;
  // Walk of op(A)'s block row (SSD, DSS) and of op(B)'s block column (SDS, DSS), resolved by the
  // host to the row view (offsets, indices, blocks == nullptr) or the column view (offsets_t,
  // indices_t, block_offsets) of the stored matrix.
  const int* a_offsets;
  const short* a_indices;
  const int* a_blocks;
  const int* b_offsets;
  const short* b_indices;
  const int* b_blocks;
  // Sparse output (SSD, SDS): coordinates of block t.
  const short* c_rows;
  const short* c_cols;
  void* c;
  long long ldc;   // DSS: elements per row of C
  int num_tiles;   // SSD / SDS: nnz blocks of C; DSS: block rows x block columns of C
  int tiles_n;     // DSS: block columns of C
  int single;      // SSD: never pair blocks (few tiles: run them all concurrently)
  uint32_t idesc;       // UMMA N = 256 (paired SSD tiles)
  uint32_t idesc_half;  // UMMA N = 128
  int bf16;
  unsigned int* tile_counter;  // dynamic scheduler (nullptr: static round-robin)
};

struct SiblingTiling {
  using T = Tiling<256>;
  static constexpr int kTableBytes = kSibMaxKBlocks * 4;
  static constexpr int kEpiBytes = 4 * 4096;  // one staging tile per epilogue warp (no TMA store)
  static constexpr int kAuxBytes = (2 * T::kStages + 4) * 8 + 16 + 2 * 2 * 32 * 4 +
                                   kTileQ * (16 + 3 * 4) + 64;
  static constexpr int kSmemBytes =
      T::kStages * T::kStageBytes + kEpiBytes + kTableBytes + kAuxBytes + 1024;
  static_assert(kSmemBytes <= 227 * 1024, "shared memory of one CTA");
};

template <bool A_MN, bool B_MN>
__global__ void __launch_bounds__(kThreads, 1)
bsmm_sibling_kernel(const __grid_constant__ CUtensorMap map_a,
                    const __grid_constant__ CUtensorMap map_b, const SiblingParams p) {
  using T = Tiling<256>;
  constexpr int S = T::kStages;
  constexpr int BN = 256;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* epi_smem = smem + S * T::kStageBytes;
  int* s_tab = reinterpret_cast<int*>(epi_smem + SiblingTiling::kEpiBytes);  // [kSibMaxKBlocks]
  uint8_t* aux = reinterpret_cast<uint8_t*>(s_tab) + SiblingTiling::kTableBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(aux);
  uint64_t* empty_bar = full_bar + S;
  uint64_t* tfull_bar = empty_bar + S;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  int* s_p = reinterpret_cast<int*>(tmem_ptr + 4);  // [2][32] M-side coordinate of a K block
  int* s_q = s_p + 64;                              // [2][32] N-side coordinate
  uint64_t* tq_full = reinterpret_cast<uint64_t*>(s_q + 64);
  uint64_t* tq_empty = tq_full + kTileQ;
  int* s_tile = reinterpret_cast<int*>(tq_empty + kTileQ);
  int* s_steps = s_tile + kTileQ;
  int* s_pair = s_steps + kTileQ;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.num_tiles;

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
    for (int i = 0; i < kTileQ; ++i) {
      ptx::mbar_init(&tq_full[i], 1);
      ptx::mbar_init(&tq_empty[i], 5);  // MMA thread + 4 epilogue warps
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc(tmem_ptr, T::kTmemCols);
  if (warp == 3 && p.op == kSibDSS) {
    for (int i = lane; i < kSibMaxKBlocks; i += 32) s_tab[i] = 0;
  }
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;
  ptx::grid_dep_wait();
  ptx::grid_dep_launch();

  if (warp == 0) {
    // ===================================================================== TMA producer
    int stage = 0;
    uint32_t phase = 0;
    int buf = 0;
    int tq = 0, t_static = int(blockIdx.x);
    uint32_t tqphase = 0;
    const unsigned draw_limit = unsigned(num_tiles) + gridDim.x - 1u;
    for (;;) {
      int t = 0;
      if (lane == 0) {
        ptx::mbar_wait(&tq_empty[tq], tqphase ^ 1);
        for (;;) {
          if (p.tile_counter) {
            t = int(atomicInc(p.tile_counter, draw_limit));
          } else {
            t = t_static < num_tiles ? t_static : num_tiles;
            t_static += int(gridDim.x);
          }
          // SSD tile ids are block ids: an odd block in the same block row as its even neighbour
          // belongs to that neighbour's pair tile.
          if (p.op == kSibSSD && t < num_tiles && !p.single && (t & 1) &&
              __ldg(p.c_rows + t - 1) == __ldg(p.c_rows + t))
            continue;
          break;
        }
      }
      t = __shfl_sync(0xffffffffu, t, 0);
      if (t >= num_tiles) {
        if (lane == 0) {
          s_tile[tq] = t;
          s_steps[tq] = 0;
          s_pair[tq] = 0;
          ptx::mbar_arrive(&tq_full[tq]);
        }
        break;
      }
      // ---- tile coordinates and walks (whole warp, converged)
      int i, j, j1 = 0;
      bool paired = false;
      if (p.op == kSibDSS) {
        i = t / p.tiles_n;
        j = t - i * p.tiles_n;
      } else {
        i = int(__ldg(p.c_rows + t));
        j = int(__ldg(p.c_cols + t));
        if (p.op == kSibSSD && !p.single && !(t & 1) && t + 1 < num_tiles &&
            int(__ldg(p.c_rows + t + 1)) == i) {
          paired = true;
          j1 = int(__ldg(p.c_cols + t + 1));
        }
      }
      int a0 = 0, a1 = 0, b0 = 0, b1 = 0;
      if (p.op != kSibSDS) { a0 = __ldg(p.a_offsets + i); a1 = __ldg(p.a_offsets + i + 1); }
      if (p.op != kSibSSD) { b0 = __ldg(p.b_offsets + j); b1 = __ldg(p.b_offsets + j + 1); }
      int nblocks;
      if (p.op == kSibSSD) {
        nblocks = a1 - a0;
      } else if (p.op == kSibSDS) {
        nblocks = b1 - b0;
      } else {
        // Column j of op(B) into the position table, then count the hits of row i of op(A).
        for (int w = b0 + lane; w < b1; w += 32)
          s_tab[int(__ldg(p.b_indices + w))] = (p.b_blocks ? __ldg(p.b_blocks + w) : w) + 1;
        __syncwarp();
        nblocks = 0;
        for (int wc = a0; wc < a1; wc += 32) {
          const int w = wc + lane;
          const bool hit = w < a1 && s_tab[int(__ldg(p.a_indices + w))] != 0;
          nblocks += __popc(__ballot_sync(0xffffffffu, hit));
        }
      }
      if (lane == 0) {
        s_tile[tq] = t;
        s_steps[tq] = 2 * nblocks;
        s_pair[tq] = paired ? 1 : 0;
        ptx::mbar_arrive(&tq_full[tq]);
      }
      if (++tq == kTileQ) { tq = 0; tqphase ^= 1; }

      // ---- K blocks: the driving walk, 32 entries at a time
      const int w0 = p.op == kSibSDS ? b0 : a0;
      const int w1 = p.op == kSibSDS ? b1 : a1;
      const uint32_t bytes = T::kABytes + (paired ? T::kBBytes : T::kBBytes / 2);
      for (int wc = w0; wc < w1; wc += 32) {
        const int w = wc + lane;
        bool valid = w < w1;
        int pv = 0, qv = 0;
        if (valid) {
          if (p.op == kSibSSD) {
            pv = p.a_blocks ? __ldg(p.a_blocks + w) : w;  // block of A
            qv = int(__ldg(p.a_indices + w));             // K block -> rows of op(B)
          } else if (p.op == kSibSDS) {
            pv = int(__ldg(p.b_indices + w));             // K block -> columns of op(A)
            qv = p.b_blocks ? __ldg(p.b_blocks + w) : w;  // block of B
          } else {
            const int hit = s_tab[int(__ldg(p.a_indices + w))];
            valid = hit != 0;
            pv = p.a_blocks ? __ldg(p.a_blocks + w) : w;
            qv = hit - 1;
          }
        }
        const unsigned m = __ballot_sync(0xffffffffu, valid);
        if (valid) {
          const int pos = __popc(m & ((1u << lane) - 1u));
          s_p[buf * 32 + pos] = pv;
          s_q[buf * 32 + pos] = qv;
        }
        __syncwarp();
        if (lane == 0) {
          const int cnt = __popc(m);
          for (int e = 0; e < cnt; ++e) {
            const int pe = s_p[buf * 32 + e], qe = s_q[buf * 32 + e];
            int m0, mk, n0, nk;
            if (p.op == kSibSDS) { m0 = i * kBlock; mk = pe * kBlock; }       // dense op(A)
            else if (A_MN)       { m0 = 0;          mk = pe * kBlock; }       // block stored [k][m]
            else                 { m0 = pe * kBlock; mk = 0; }                // block stored [m][k]
            if (p.op == kSibSSD) { n0 = j * kBlock; nk = qe * kBlock; }       // dense op(B)
            else if (B_MN)       { n0 = 0;          nk = qe * kBlock; }       // block stored [k][n]
            else                 { n0 = qe * kBlock; nk = 0; }                // block stored [n][k]
#pragma unroll
            for (int kh = 0; kh < 2; ++kh) {
              ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
              uint8_t* sa = smem + stage * T::kStageBytes;
              uint8_t* sb = sa + T::kABytes;
              ptx::mbar_arrive_expect_tx(&full_bar[stage], bytes);
              load_mside<A_MN>(&map_a, &full_bar[stage], sa, m0, mk + kh * kBK);
              load_nside<B_MN, 128>(&map_b, &full_bar[stage], sb, n0, nk + kh * kBK);
              if (paired)
                load_nside<B_MN, 128>(&map_b, &full_bar[stage], sb + 16384, j1 * kBlock,
                                      nk + kh * kBK);
              if (++stage == S) { stage = 0; phase ^= 1; }
            }
          }
        }
        __syncwarp();
        buf ^= 1;
      }
      if (p.op == kSibDSS) {  // leave the table clean for the next tile
        for (int w = b0 + lane; w < b1; w += 32) s_tab[int(__ldg(p.b_indices + w))] = 0;
        __syncwarp();
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
      int tq = 0;
      uint32_t tqphase = 0;
      for (;;) {
        ptx::mbar_wait(&tq_full[tq], tqphase);
        const int t = s_tile[tq];
        const int nsteps = s_steps[tq];
        const uint32_t idesc = s_pair[tq] ? p.idesc : p.idesc_half;
        ptx::mbar_arrive(&tq_empty[tq]);
        if (++tq == kTileQ) { tq = 0; tqphase ^= 1; }
        if (t >= num_tiles) break;
        ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);
        ptx::tc_fence_after_sync();
        if (nsteps == 0) ptx::mbar_arrive(&tfull_bar[as]);  // epilogue writes zeros
        const uint32_t tmem_d = tmem_base + uint32_t(as * BN);
        for (int s = 0; s < nsteps; ++s) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after_sync();
          const uint32_t sa = ptx::smem_u32(smem + stage * T::kStageBytes);
          const uint32_t sb = sa + T::kABytes;
          const uint64_t da = ptx::make_smem_desc(sa, A_MN ? 8192 : 16, 1024);
          const uint64_t db = ptx::make_smem_desc(sb, B_MN ? 8192 : 16, 1024);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            ptx::umma_f16(tmem_d, da + uint64_t(k * kAdvA), db + uint64_t(k * kAdvB), idesc,
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
    uint8_t* tw = epi_smem + q * 4096;
    uint16_t* const cbase = reinterpret_cast<uint16_t*>(p.c);
    int as = 0;
    uint32_t aphase = 0;
    int tq = 0;
    uint32_t tqphase = 0;
    for (;;) {
      ptx::mbar_wait(&tq_full[tq], tqphase);
      const int t = s_tile[tq];
      const int nsteps = s_steps[tq];
      const int paired = s_pair[tq];
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tq_empty[tq]);
      if (++tq == kTileQ) { tq = 0; tqphase ^= 1; }
      if (t >= num_tiles) break;
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after_sync();
      // 128 x 128 row-major tile(s): block t (and t + 1) of a sparse C, or tile (i, j) of a
      // dense C.  64 columns (128 B) at a time through a per-warp XOR-swizzled staging tile so
      // that global stores are whole 128 B lines.
      uint16_t* tile_base;
      long long ld;
      if (p.op == kSibDSS) {
        const int i = t / p.tiles_n, j = t - i * p.tiles_n;
        tile_base = cbase + size_t(i) * kBlock * size_t(p.ldc) + size_t(j) * kBlock;
        ld = p.ldc;
      } else {
        tile_base = cbase + size_t(t) * (kBlock * kBlock);
        ld = kBlock;
      }
      const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * BN);
      const int nch = paired ? 4 : 2;
#pragma unroll 1
      for (int ch = 0; ch < nch; ++ch) {
        const int seg = lane & 7;
        uint16_t* o = tile_base + size_t(ch >> 1) * (kBlock * kBlock) + size_t(q * 32) * size_t(ld) +
                      (ch & 1) * 64 + seg * 8;
        const bool accum = (p.bf16 & kFmtAccum) != 0;
        uint4 old[8];
        if (accum) {  // BSMM_FLAG_ACCUMULATE: old lines requested before the accumulator is read
#pragma unroll
          for (int e = 0; e < 8; ++e)
            old[e] = *reinterpret_cast<const uint4*>(o + size_t((lane >> 3) + 4 * e) * size_t(ld));
        }
        uint32_t v[64];
        if (nsteps > 0) {
          ptx::tmem_ld_32x32b_x32(taddr + uint32_t(ch * 64), *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
          ptx::tmem_ld_32x32b_x32(taddr + uint32_t(ch * 64 + 32), *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
          ptx::tmem_ld_wait();
        } else {
#pragma unroll
          for (int e = 0; e < 64; ++e) v[e] = 0u;
        }
        stage_chunk64(*reinterpret_cast<const uint32_t(*)[64]>(&v[0]), tw, lane, p.bf16);
        __syncwarp();
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const int r = (lane >> 3) + 4 * e;
          uint4 w = *reinterpret_cast<const uint4*>(tw + r * 128 + ((seg ^ (r & 7)) << 4));
          if (accum) w = add4(w, old[e], p.bf16);
          *reinterpret_cast<uint4*>(o + size_t(r) * size_t(ld)) = w;
        }
        __syncwarp();
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

// Bitmask of a topology, K-contiguous, bit-identical to the reference's BitMatrix
// (sputnik/block/bitmask/bit_matrix.h:12-62) as filled by Bitmask() (bitmask.cu:7-45, HOST code
// with two blocking copies): word [r * words_per_row + c / 64] has bit (c % 64) set for every
// block (r, c) of the walked view.  One warp per walk row.
__global__ void __launch_bounds__(256)
bitmask_kernel(const int* __restrict__ offsets, const short* __restrict__ indices, int walk_rows,
               int words_per_row, unsigned long long* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int r = warp; r < walk_rows; r += nwarps) {
    unsigned long long* row = out + size_t(r) * words_per_row;
    for (int i = lane; i < words_per_row; i += 32) row[i] = 0ull;
    __syncwarp();
    const int w1 = __ldg(offsets + r + 1);
    for (int w = __ldg(offsets + r) + lane; w < w1; w += 32) {
      const int c = int(__ldg(indices + w));
      atomicOr(row + (c >> 6), 1ull << (c & 63));
    }
    __syncwarp();
  }
}

}  // namespace bsmm
