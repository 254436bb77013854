// Split-K for the remainder wave of the 2-SM DSD / DDS kernel (bsmm_pair_kernel.cuh).
//
// Why: the tile-per-cluster schedule quantises into waves.  C2 (the headline) is 128 tiles of
// 256 x 256 on P = 74 CTA pairs: one full wave and a remainder of R = 54 tiles, during which 20
// pairs idle (13.5 % of the machine, however fast the mainloop is: profiles/r01_trace_*.txt).
//
// Schedule: the W = floor(tiles / P) full waves run exactly as in the classic kernel (tile t on
// cluster t mod P, neighbouring clusters streaming the same operands in lockstep — which is what
// keeps the L2 -> SM feed merged).  Each of the R remainder tiles is cut ONCE along its walk:
//   clusters 0 .. R-1   take the head, blocks [0, h), of remainder tile c      ("finisher");
//   clusters R .. P-1   share the tails, blocks [h, n), round-robin: g = ceil(R / (P - R))
//                       tails each                                              ("writers");
// with h = n g / (1 + g), so that a finisher's h blocks take as long as a writer's g tails.  A
// writer parks its fp32 accumulators in the tile's workspace slot and release-stores the launch
// epoch to the slot's flags; the finisher, after its own mainloop, acquires the flags, adds the
// partial into TMEM (tcgen05.ld / add / tcgen05.st) and runs the normal epilogue.  C2: h = 24 of
// n = 32, the remainder costs 24 block-times instead of 32 (56 instead of 64 in total) for
// 54 x 256 KiB of partials.
//
// Writers never wait and run before (or while) the finishers do, so there is no deadlock as long
// as all clusters are co-resident (grid <= cudaOccupancyMaxActiveClusters, like the classic
// kernel).  The summation order is fixed by (topology, P): results are deterministic, and differ
// from the classic schedule only in fp32 association.
//
// (A full stream-K schedule — equal contiguous unit ranges per cluster — was built first and
// measured: C2 56.3 -> 81.0 us.  Ranges that start at unrelated walk positions break the lockstep
// between clusters, profiles/r01_streamk_ab.txt.)
//
// Tiles of empty walk rows are zero-filled by cluster t mod P before its epilogue loop and never
// appear as pieces.
#pragma once

#include "bsmm_pair_kernel.cuh"

namespace bsmm {

struct SkPiece {
  int pair, minor;   // tile
  int kb, ke, n;     // blocks [kb, ke) of the n in the pair's walk
  int slot;          // workspace slot of the tile (remainder tiles only)
};

// Iterator over the pieces of cluster `cid`.  Every role of a CTA runs its own copy; all of them
// (and both CTAs of the pair) see the same sequence.
struct SkWalk {
  const int* offs;
  int tm, P, cid;
  int t;        // next classic tile
  int T0, R;    // first remainder tile, number of remainder tiles
  int g;        // tails per writer cluster
  int j;        // writers: next remainder tile; finishers: R once done

  __device__ __forceinline__ void init(const GemmParams& p, int cid_, int P_) {
    offs = p.walk_offsets;
    tm = p.tiles_minor;
    P = P_;
    cid = cid_;
    const int tiles = (p.tiles_major >> 1) * tm;
    T0 = tiles / P * P;
    R = tiles - T0;
    g = R > 0 && R < P ? (R + (P - R) - 1) / (P - R) : 0;
    t = cid;
    j = cid < R ? cid : cid - R;
  }

  __device__ __forceinline__ int head(int n) const {  // blocks the finisher keeps
    const int h = (n * g + g) / (1 + g);              // ~ n g / (1 + g), rounded up
    return h < 1 ? 1 : (h > n ? n : h);
  }

  __device__ __forceinline__ bool next(SkPiece& pc) {
    while (t < T0) {
      const int pair = t / tm;
      const int n = __ldg(offs + 2 * pair + 1) - __ldg(offs + 2 * pair);
      const int minor = t - pair * tm;
      t += P;
      if (n == 0) continue;
      pc.pair = pair; pc.minor = minor; pc.kb = 0; pc.ke = pc.n = n; pc.slot = 0;
      return true;
    }
    while (j < R) {
      const int tile = T0 + j;
      const int pair = tile / tm;
      const int n = __ldg(offs + 2 * pair + 1) - __ldg(offs + 2 * pair);
      pc.pair = pair; pc.minor = tile - pair * tm; pc.n = n; pc.slot = j;
      if (cid < R) {
        j = R;  // a finisher has exactly one remainder piece
        if (n == 0) return false;
        pc.kb = 0; pc.ke = g > 0 ? head(n) : n;
        return true;
      }
      j += P - R;
      if (n == 0 || head(n) >= n) continue;  // nothing left for a tail
      pc.kb = head(n); pc.ke = n;
      return true;
    }
    return false;
  }
};

// Workspace slot of (remainder tile, rank): fp32 [256 columns][128 rows] — column-major so that the 32
// lanes of an epilogue warp (one accumulator row each) touch 128 contiguous bytes per column.
constexpr size_t kSkSlotFloats = size_t(kBlock) * PairTiling::kBN;

// p.sk_ws / p.sk_flags / p.sk_epoch: workspace, flags [remainder tile][rank][epilogue warp] and
// the value that marks a slot as written by THIS launch.
template <bool A_MN, bool B_MN, bool TRANS_OUT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
bsmm_pair_sk_kernel(const __grid_constant__ CUtensorMap map_a,
                    const __grid_constant__ CUtensorMap map_b, const GemmParams p) {
  using T = PairTiling;
  constexpr int S = T::kStages;
  constexpr int BN = T::kBN;

  if (gated_off(p)) return;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* epi_smem = smem + S * T::kStageBytes;
  uint8_t* aux = epi_smem + T::kEpiBytes;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(aux);
  uint64_t* empty_bar = full_bar + S;
  uint64_t* tfull_bar = empty_bar + S;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  int* s_idx = reinterpret_cast<int*>(tmem_ptr + 4);  // [2][32] dense-side block coordinate
  int* s_blk = s_idx + 64;                            // [2][32] this CTA's block id

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const bool leader = rank == 0;
  const int cid = int(ptx::cluster_id_x());
  const int P = int(ptx::num_clusters_x());

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
      ptx::mbar_init(&tempty_bar[i], 8);  // 4 epilogue warps of each CTA of the pair
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc_pair(tmem_ptr, T::kTmemCols);
  ptx::tc_fence_before_sync();
  ptx::cluster_sync();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;
  ptx::grid_dep_wait();
  ptx::grid_dep_launch();


  if (warp == 0) {
    // ===================================================================== TMA producer
    int stage = 0;
    uint32_t phase = 0;
    int buf = 0;
    SkWalk w;
    w.init(p, cid, P);
    SkPiece pc;
    while (w.next(pc)) {
      const int l0 = __ldg(p.walk_offsets + 2 * pc.pair) + pc.kb;
      const int own0 = __ldg(p.walk_offsets + 2 * pc.pair + int(rank)) + pc.kb;
      const int nblk = pc.ke - pc.kb;
      const int d0 = pc.minor * BN + int(rank) * (BN / 2);  // this CTA's half of the dense tile
      for (int ic = 0; ic < nblk; ic += 32) {
        if (ic + lane < nblk) {
          s_idx[buf * 32 + lane] = int(__ldg(p.walk_indices + l0 + ic + lane));
          s_blk[buf * 32 + lane] =
              p.walk_blocks ? __ldg(p.walk_blocks + own0 + ic + lane) : own0 + ic + lane;
        }
        __syncwarp();
        if (lane == 0) {
          const int cnt = min(32, nblk - ic);
          for (int i = 0; i < cnt; ++i) {
            const int blk_row0 = s_blk[buf * 32 + i] * kBlock;
            const int dense_k0 = s_idx[buf * 32 + i] * kBlock;
#pragma unroll
            for (int kh = 0; kh < 2; ++kh) {
              ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
              uint8_t* sa = smem + stage * T::kStageBytes;
              uint8_t* sb = sa + T::kABytes;
              if (leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * T::kStageBytes);
              const uint32_t bar = ptx::map_to_rank(&full_bar[stage], 0);
              if constexpr (A_MN) {
                ptx::tma_load_2d_pair(&map_a, bar, sa, 0, blk_row0 + kh * kBK);
                ptx::tma_load_2d_pair(&map_a, bar, sa + 8192, 64, blk_row0 + kh * kBK);
              } else {
                ptx::tma_load_2d_pair(&map_a, bar, sa, kh * kBK, blk_row0);
              }
              const int k0 = dense_k0 + kh * kBK;
              if constexpr (B_MN) {
                ptx::tma_load_2d_pair(&map_b, bar, sb, d0, k0);
                ptx::tma_load_2d_pair(&map_b, bar, sb + 8192, d0 + 64, k0);
              } else {
                ptx::tma_load_2d_pair(&map_b, bar, sb, k0, d0);
              }
              if (++stage == S) { stage = 0; phase ^= 1; }
            }
          }
        }
        __syncwarp();
        buf ^= 1;
      }
    }
  } else if (warp == 1) {
    // ===================================================================== MMA issuer (leader)
    if (leader && lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      constexpr uint32_t kAdvA = A_MN ? (2048 >> 4) : (32 >> 4);
      constexpr uint32_t kAdvB = B_MN ? (2048 >> 4) : (32 >> 4);
      SkWalk w;
      w.init(p, cid, P);
      SkPiece pc;
      while (w.next(pc)) {
        const int nsteps = 2 * (pc.ke - pc.kb);  // > 0: a piece holds at least one block
        ptx::mbar_wait_cluster(&tempty_bar[as], aphase ^ 1);
        ptx::tc_fence_after_sync();
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
            ptx::umma_f16_pair(tmem_d, da + uint64_t(k * kAdvA), db + uint64_t(k * kAdvB), p.idesc,
                               uint32_t((s | k) != 0));
          }
          ptx::umma_commit_pair(&empty_bar[stage]);
          if (s == nsteps - 1) ptx::umma_commit_pair(&tfull_bar[as]);
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ===================================================================== epilogue
    const int q = warp & 3;
    // tiles of empty walk rows (no unit belongs to them): zeros, dealt round-robin
    {
      const int num_tiles = (p.tiles_major >> 1) * p.tiles_minor;
      for (int t = cid; t < num_tiles; t += P) {
        const int pair = t / p.tiles_minor;
        if (__ldg(p.walk_offsets + 2 * pair + 1) == __ldg(p.walk_offsets + 2 * pair))
          store_dsd_tile<TRANS_OUT>(p, 0u, epi_smem, q, lane, 2 * pair + int(rank),
                                    t - pair * p.tiles_minor, false);
      }
    }
    int as = 0;
    uint32_t aphase = 0;
    SkWalk w;
    w.init(p, cid, P);
    SkPiece pc;
    while (w.next(pc)) {
      const int major = 2 * pc.pair + int(rank);
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after_sync();
      const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * BN);
      if (pc.kb > 0) {
        // the tail of a remainder tile: park the accumulators for the tile's finisher
        float* slot = p.sk_ws + (size_t(pc.slot) * 2 + rank) * kSkSlotFloats + q * 32 + lane;
#pragma unroll 1
        for (int cc = 0; cc < BN / 32; ++cc) {
          uint32_t v[32];
          ptx::tmem_ld_32x32b_x32(taddr + uint32_t(cc * 32), v);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j)
            __stcg(slot + size_t(cc * 32 + j) * kBlock, __uint_as_float(v[j]));
        }
        __threadfence();
        __syncwarp();
        if (lane == 0)
          ptx::st_release_gpu(p.sk_flags + (size_t(pc.slot) * 2 + rank) * 4 + q, p.sk_epoch);
      } else {
        if (pc.ke < pc.n) {
          // the head of a remainder tile: fold in the tail's partial accumulators
          const unsigned int* flag = p.sk_flags + (size_t(pc.slot) * 2 + rank) * 4 + q;
          if (lane == 0) {
            while (ptx::ld_acquire_gpu(flag) != p.sk_epoch) __nanosleep(40);
          }
          __syncwarp();
          __threadfence();
          const float* slot = p.sk_ws + (size_t(pc.slot) * 2 + rank) * kSkSlotFloats + q * 32 + lane;
#pragma unroll 1
          for (int cc = 0; cc < BN / 32; ++cc) {
            uint32_t v[32];
            ptx::tmem_ld_32x32b_x32(taddr + uint32_t(cc * 32), v);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j)
              v[j] = __float_as_uint(__uint_as_float(v[j]) +
                                     __ldcg(slot + size_t(cc * 32 + j) * kBlock));
            ptx::tmem_st_32x32b_x32(taddr + uint32_t(cc * 32), v);
          }
          ptx::tmem_st_wait();
        }
        store_dsd_tile<TRANS_OUT>(p, taddr, epi_smem, q, lane, major, pc.minor, true);
      }
      ptx::tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(ptx::map_to_rank(&tempty_bar[as], 0));
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
  }

  __syncwarp();
  ptx::tc_fence_before_sync();
  ptx::cluster_sync();  // no CTA may exit (or free TMEM) while its peer still signals it
  if (warp == 2) {
    ptx::tc_fence_after_sync();
    ptx::tmem_dealloc_pair(tmem_base, T::kTmemCols);
  }
}

}  // namespace bsmm
