// 2-SM variant of the DSD / DDS mainloop: tcgen05.mma.cta_group::2 (UMMA 256 x 256 x 16) over a
// CTA pair, for "twin" block rows — two block rows (or, on the transposed walk, block columns)
// of the sparse operand whose index lists are identical, so that at every step of the walk the
// two sparse blocks multiply the SAME panel of the dense operand.  dMoE topologies consist of
// such rows (all block rows of an expert share the expert's column range).
//
// Why: the 1-SM kernel (bsmm_kernel.cuh) is bound by the L2 -> SM feed, not by the tensor pipe
// (profiles/r01_ncu_c2_dsd_full.txt: 48 KiB per 128x256x64 stage = 96 B/clk/SM needed, ~69
// delivered).  Here each CTA of the pair loads its own 128 x 64 sparse tile and only HALF of the
// 256 x 64 dense tile: 32 KiB per stage (64 B/clk/SM at full tensor rate), and the ring is 6
// stages deep instead of 4.
//
// Roles per CTA: warp 0 TMA producer (both CTAs), warp 1 MMA issuer (leader CTA only), warp 2
// TMEM allocator, warps 4-7 epilogue (each CTA drains its own 128 accumulator rows).
// Barriers live at the same shared-memory offsets in both CTAs:
//   full[s]   leader's only: 1 arrival (leader's expect_tx of 64 KiB) + both CTAs' TMA bytes;
//   empty[s]  both: multicast tcgen05.commit from the leader;
//   tfull[a]  both: multicast tcgen05.commit from the leader (accumulator stage a complete);
//   tempty[a] leader's only: 8 arrivals (4 epilogue warps of each CTA).
// Replaces, like bsmm_kernel.cuh, the reference's BlockGemm + indexed iterators
// (sputnik/block/cutlass/block_gemm.h:749-878, dependent_tile_access_iterator.h:120-133).
#pragma once
#include <type_traits>

#include "bsmm_kernel.cuh"

namespace bsmm {

struct PairTiling {
  static constexpr int kBN = 256;
  static constexpr int kABytes = kBlock * kBK * 2;        // 16 KiB: this CTA's sparse tile
  static constexpr int kBBytes = (kBN / 2) * kBK * 2;     // 16 KiB: this CTA's half of the dense tile
  static constexpr int kStageBytes = kABytes + kBBytes;   // 32 KiB
  static constexpr int kStages = 6;
  static constexpr int kTmemCols = 2 * kBN;
  static constexpr int kEpiWarpBytes = 8192;   // double staging buffer per epilogue warp (tma_chunk64)
  static constexpr int kEpiBytes = 4 * kEpiWarpBytes;
  // barriers, TMEM pointer, index chunks [2][32] x 2, and the DSD kernel's tile-decode ring
  // (2 full + 2 empty barriers, 2 x 4 ints of tile header, 2 x 32 x 2 ints of first chunk)
  static constexpr int kAuxBytes = (2 * kStages + 5) * 8 + 16 + 2 * 2 * 32 * 4 + 4 * 8 + 2 * 4 * 4 +
                                   2 * 2 * 32 * 4;
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + kAuxBytes + 1024;
  static_assert(kSmemBytes <= 227 * 1024, "shared memory of one CTA");
};

// One epilogue warp stores 32 accumulator rows x 64 columns, row-major: TMEM -> registers ->
// cvt.rn -> XOR-swizzled staging tile -> whole 128 B lines (8 lanes x 16 B per row, 4 rows per
// instruction).  `o` = this lane's segment of row 0 (already offset by (lane & 7) * 8 columns).
__device__ __forceinline__ void store_chunk64(uint32_t taddr, uint8_t* tw, int lane, uint16_t* o,
                                              size_t ld, bool col_ok, bool nonempty, int bf16,
                                              int rows_ok = 32, void* const* rows = nullptr,
                                              int col = 0, int cols_left = 64,
                                              const CUtensorMap* map = nullptr, int tc0 = 0,
                                              int tc1 = 0, int* flip = nullptr) {
  if (map && !rows && (bf16 & kFmtTma)) {  // TMA-store epilogue (see tma_chunk64)
    tma_chunk64(taddr, tw, lane, map, tc0, tc1, nonempty, bf16, flip);
    return;
  }
  const int seg = lane & 7;
  if ((bf16 & (kFmtDirect | kFmtAccum)) == kFmtDirect) {
    // direct epilogue (see store_row64_direct): lane = row, four 32 B stores from registers.
    // `cols_left` = valid columns from the chunk's first column.
    int cols = lane < rows_ok ? min(64, cols_left) : 0;
    uint16_t* rp;
    if (rows) {
      uint16_t* base = static_cast<uint16_t*>(ptx::ldg_ptr(rows + lane));
      if (!base) cols = 0;
      rp = base + (col - seg * 8);
    } else {
      rp = o - seg * 8 + size_t(lane) * ld;
    }
    uint32_t v[64];
    if (nonempty) {
      ptx::tmem_ld_32x32b_x32(taddr, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
      ptx::tmem_ld_32x32b_x32(taddr + 32u, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
      ptx::tmem_ld_wait();
    } else {
#pragma unroll
      for (int i = 0; i < 64; ++i) v[i] = 0u;
    }
    store_row64_direct(v, rp, cols, bf16);
    return;
  }
  // BSMM_FLAG_ACCUMULATE: the old lines of the chunk are requested before the accumulator is read
  const bool accum = (bf16 & kFmtAccum) && !rows && col_ok;
  uint4 old[8];
  if (accum) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = (lane >> 3) + 4 * i;
      old[i] = r < rows_ok ? *reinterpret_cast<const uint4*>(o + size_t(r) * ld)
                           : make_uint4(0u, 0u, 0u, 0u);
    }
  }
  // scattered rows: the 8 row addresses this lane stores to are requested up front, so their
  // latency hides under the TMEM read + conversion (they used to sit right before each store)
  uint16_t* rp[8];
  if (rows) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      rp[i] = static_cast<uint16_t*>(ptx::ldg_ptr(rows + (lane >> 3) + 4 * i));
  }
  uint32_t v[64];
  if (nonempty && !BSMM_DBG(2)) {
    ptx::tmem_ld_32x32b_x32(taddr, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
    ptx::tmem_ld_32x32b_x32(taddr + 32u, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
    ptx::tmem_ld_wait();
  } else {
#pragma unroll
    for (int i = 0; i < 64; ++i) v[i] = 0u;
  }
  stage_chunk64(v, tw, lane, bf16);
  __syncwarp();
  if (accum) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = (lane >> 3) + 4 * i;
      const uint4 w = *reinterpret_cast<const uint4*>(tw + r * 128 + ((seg ^ (r & 7)) << 4));
      if (r < rows_ok) *reinterpret_cast<uint4*>(o + size_t(r) * ld) = add4(w, old[i], bf16);
    }
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int r = (lane >> 3) + 4 * i;
      const uint4 w = *reinterpret_cast<const uint4*>(tw + r * 128 + ((seg ^ (r & 7)) << 4));
      if (rows) {  // scattered rows (bsmm_dsd_scatter): per-row base pointers, null = drop
        if (col_ok && rp[i]) *reinterpret_cast<uint4*>(rp[i] + col) = w;
      } else if (col_ok && r < rows_ok && !BSMM_DBG(1)) {
        *reinterpret_cast<uint4*>(o + size_t(r) * ld) = w;
      }
    }
  }
  __syncwarp();
}

// Stores one 128 x (<= 256) accumulator (TMEM lanes = rows of the sparse dimension) of a DSD-like
// tile.  TRANS_OUT = false: C[major*128 + lane][minor*256 + col]; true: the transpose (DDS).
// Same data path as the epilogue of bsmm_gemm_kernel.
// `part`: 0 = the whole tile, 1 = its first 128 columns, 2 = the rest (the last tile of a CTA is
// drained by two warps per lane quarter: the epilogue warp and the by then idle warp q).
template <bool TRANS_OUT>
__device__ __forceinline__ void store_dsd_tile(const GemmParams& p, uint32_t taddr,
                                               uint8_t* epi_smem, int q, int lane, int major,
                                               int minor, bool nonempty, int part = 0,
                                               const CUtensorMap* map = nullptr,
                                               int* flip = nullptr) {
  // `epi_smem`: the staging memory of lane quarter q starts at epi_smem + q * kEpiWarpBytes
  constexpr int BN = PairTiling::kBN;
  uint16_t* const cbase = reinterpret_cast<uint16_t*>(p.c);
  const int ncc = min(BN / 32, (p.dense_lim - minor * BN + 31) / 32);
  const int cc0 = part == 2 ? min(ncc, BN / 64) : 0;           // in 32-column units
  const int cc1 = part == 1 ? min(ncc, BN / 64) : ncc;
  if constexpr (!TRANS_OUT) {
    const int ch0 = (cc0 + 1) >> 1, nch = (cc1 + 1) >> 1;  // odd cc0: that chunk is part 1's
#pragma unroll 1
    for (int ch = ch0; ch < nch; ++ch) {
      const int col = minor * BN + ch * 64 + (lane & 7) * 8;
      uint16_t* o = cbase + size_t(major * kBlock + q * 32) * size_t(p.ldc) + col;
      store_chunk64(taddr + uint32_t(ch * 64), epi_smem + q * PairTiling::kEpiWarpBytes, lane, o,
                    size_t(p.ldc), col < p.dense_lim /* extents are multiples of 8 */, nonempty,
                    p.bf16, 32, p.c_rows ? p.c_rows + major * kBlock + q * 32 : nullptr, col,
                    p.dense_lim - (minor * BN + ch * 64), map, minor * BN + ch * 64,
                    major * kBlock + q * 32, flip);
    }
  } else {
#pragma unroll 1
    for (int cc = cc0; cc < cc1; ++cc) {
      const int drow0 = minor * BN + cc * 32;
      if (map && (p.bf16 & kFmtTma)) {  // [dense extent][sparse extent] map, box 32 x 32
        tma_chunk32_tr(taddr + uint32_t(cc * 32), epi_smem + q * PairTiling::kEpiWarpBytes, lane,
                       map, major * kBlock + q * 32, drow0, nonempty, p.bf16, flip);
        continue;
      }
      const int seg = lane & 3;
      uint16_t* o = cbase + size_t(major * kBlock + q * 32 + seg * 8);
      const bool accum = (p.bf16 & kFmtAccum) != 0;
      uint4 old[4];
      if (accum) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int r = drow0 + (lane >> 2) + 8 * i;
          old[i] = r < p.dense_lim ? *reinterpret_cast<const uint4*>(o + size_t(r) * size_t(p.ldc))
                                   : make_uint4(0u, 0u, 0u, 0u);
        }
      }
      uint32_t v[32];
      if (nonempty) {
        ptx::tmem_ld_32x32b_x32(taddr + uint32_t(cc * 32), v);
        ptx::tmem_ld_wait();
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = 0u;
      }
      uint16_t* tw = reinterpret_cast<uint16_t*>(epi_smem + q * PairTiling::kEpiWarpBytes);
      stage_chunk32_tr(v, tw, lane, p.bf16);
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int r = (lane >> 2) + 8 * i;
        uint4 w = *reinterpret_cast<const uint4*>(tw + r * 32 + seg * 8);
        if (accum) w = add4(w, old[i], p.bf16);
        if (drow0 + r < p.dense_lim)
          *reinterpret_cast<uint4*>(o + size_t(drow0 + r) * size_t(p.ldc)) = w;
      }
      __syncwarp();
    }
  }
}

// Tile t -> (pair of twin walk rows, 256-wide dense tile), rasterised so that concurrently running
// clusters share operands: consecutive pairs with the same walk length and first index (the block
// rows of one dMoE expert; at most 8 pairs, inside an aligned window so that every tile agrees on
// the group) form a group whose tiles are walked PAIR-fastest.  The 74 resident clusters then hold
// all g pairs of an expert for 74 / g dense tiles, instead of all dense tiles of 74 / tiles_minor
// pairs: a panel of the dense operand (w2 rows of an expert x 256 columns, 7 MiB at C5) is
// streamed from HBM once per expert instead of once per wave that touches the expert (C5 DSD:
// 1.84 GB of DRAM traffic for 1.24 GB of operands before, profiles/r02_ncu_c5_dsd_e8.txt).
// Any grouping gives a bijection on the tile range, so correctness never depends on it.
// Every thread computes it for itself (a few L1-resident loads per tile).  PEEK: the pre-wait
// prefetch variant (L2-only loads, indices clamped).
template <bool PEEK = false>
__device__ __forceinline__ void pair_dsd_tile(const GemmParams& p, int t, int* pair, int* minor) {
  auto ld32 = [&](const int* q) { return PEEK ? ptx::ld_cg_s32(q) : __ldg(q); };
  auto ld16 = [&](const short* q) { return PEEK ? ptx::ld_cg_s16(q) : int(__ldg(q)); };
  auto first = [&](int o, int n) {
    if (n <= 0) return -1;
    return ld16(p.walk_indices + (PEEK ? min(max(o, 0), (1 << 17) - 1) : o));
  };
  const int tm = p.tiles_minor, npairs = p.tiles_major >> 1;
  const int pr = t / tm;
  if (!p.raster) {   // BSMM_RASTER=0: plain pair-major order (A/B runs)
    *pair = pr;
    *minor = t - pr * tm;
    return;
  }
  const int w0 = pr & ~7, w1 = min(w0 + 8, npairs);
  // the window's 8 (offset, length, first index) triples: independent loads, two round trips
  int o[8], n[8], f[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const bool in = w0 + i < w1;
    o[i] = in ? ld32(p.walk_offsets + 2 * (w0 + i)) : 0;
    n[i] = in ? ld32(p.walk_offsets + 2 * (w0 + i) + 1) - o[i] : -2;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) f[i] = first(o[i], n[i]);
  const int me = pr - w0;
  int n_me = 0, f_me = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i)
    if (i == me) { n_me = n[i]; f_me = f[i]; }
  unsigned same = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) same |= unsigned(n[i] == n_me && f[i] == f_me) << i;
  int gs = me, ge = me + 1;
  while (gs > 0 && ((same >> (gs - 1)) & 1u)) --gs;
  while (ge < 8 && ((same >> ge) & 1u)) ++ge;
  gs += w0;
  ge += w0;
  const int g = ge - gs, u = t - gs * tm;
  *pair = gs + u % g;
  *minor = u / g;
}

// A_MN / B_MN / TRANS_OUT as in bsmm_gemm_kernel<kOpDSD, ...>.  Grid = 2 x (number of resident
// clusters); cluster c walks tiles c, c + #clusters, ...; tile = (pair of walk rows, 256-wide
// dense tile).  p.tiles_major = number of walk rows (even), p.idesc = UMMA 256 x 256.
template <bool A_MN, bool B_MN, bool TRANS_OUT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
bsmm_pair_kernel(const __grid_constant__ CUtensorMap map_a,
                 const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ CUtensorMap map_c, const GemmParams p) {
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
  uint64_t* last_bar = tempty_bar + 2;  // both CTAs: the cluster's LAST accumulator is complete
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(last_bar + 1);
  int* s_idx = reinterpret_cast<int*>(tmem_ptr + 4);  // [2][32] dense-side block coordinate
  int* s_blk = s_idx + 64;                            // [2][32] this CTA's block id
  // Tile-decode ring: warp 3 resolves the NEXT tile (walk offsets -> first 32 block coordinates:
  // two dependent round trips) while the producer is still issuing the current one, so the
  // producer never waits on metadata at a tile boundary (short transposed walks: 16 k-steps).
  uint64_t* meta_full = reinterpret_cast<uint64_t*>(s_blk + 64);
  uint64_t* meta_empty = meta_full + 2;
  int* s_meta = reinterpret_cast<int*>(meta_empty + 2);  // [2][4]: minor, l0, nblk, own0
  int* m_idx = s_meta + 8;                               // [2][32]
  int* m_blk = m_idx + 64;                               // [2][32]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (threadIdx.x == 0) BSMM_STAMP(0);
  const uint32_t rank = ptx::cluster_ctarank();
  const bool leader = rank == 0;
  const int num_tiles = (p.tiles_major >> 1) * p.tiles_minor;
  const int t_first = int(ptx::cluster_id_x());
  const int t_step = int(ptx::num_clusters_x());

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&map_a);
    ptx::prefetch_tensormap(&map_b);
    ptx::prefetch_tensormap(&map_c);
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
    ptx::mbar_init(last_bar, 1);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&meta_full[i], 1);
      ptx::mbar_init(&meta_empty[i], 1);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc_pair(tmem_ptr, T::kTmemCols);
  ptx::tc_fence_before_sync();
  ptx::cluster_sync();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;
  // Everything above touched only parameters and shared memory: with programmatic dependent
  // launch it overlaps the tail of the previous kernel in the stream.  So does the L2 prefetch of
  // the first tile's first operand boxes (see bsmm_gemm_kernel: hint-only metadata peek).
  if (warp == 3 && p.prefetch && p.bk == kBK && t_first < num_tiles) {
    int pair, minor_pf;
    pair_dsd_tile<true>(p, t_first, &pair, &minor_pf);
    const int d0 = minor_pf * BN + int(rank) * (BN / 2);
    const int l0 = ptx::ld_cg_s32(p.walk_offsets + 2 * pair);
    const int l1 = ptx::ld_cg_s32(p.walk_offsets + 2 * pair + 1);
    const int own0 = leader ? l0 : l1;
    const int half = lane & 1, i = lane >> 1;
    if (l0 >= 0 && i < kPrefetchBlocks && l0 + i < l1 && l1 <= (1 << 17)) {
      const int k0 = (ptx::ld_cg_s16(p.walk_indices + l0 + i) & 0x7fff) * kBlock + half * kBK;
      const int blk =
          (p.walk_blocks ? (ptx::ld_cg_s32(p.walk_blocks + own0 + i) & 0x1ffff) : own0 + i) * kBlock;
      if constexpr (A_MN) {
        ptx::tma_prefetch_2d(&map_a, 0, blk + half * kBK);
        ptx::tma_prefetch_2d(&map_a, 64, blk + half * kBK);
      } else {
        ptx::tma_prefetch_2d(&map_a, half * kBK, blk);
      }
      if constexpr (B_MN) {
        ptx::tma_prefetch_2d(&map_b, d0, k0);
        ptx::tma_prefetch_2d(&map_b, d0 + 64, k0);
      } else {
        ptx::tma_prefetch_2d(&map_b, k0, d0);
      }
    }
  }
  ptx::grid_dep_wait();
  ptx::grid_dep_launch();
  if (threadIdx.x == 0) BSMM_STAMP(1);

  if (warp == 0) {
    // ===================================================================== TMA producer
    int stage = 0;
    uint32_t phase = 0;
    int buf = 0;
    int mslot = 0;
    uint32_t mphase = 0;
    for (int t = t_first; t < num_tiles; t += t_step) {
      // tile header and first 32 block coordinates: decoded ahead by warp 3 (below)
      ptx::mbar_wait(&meta_full[mslot], mphase);
      const int minor = s_meta[mslot * 4 + 0];
      const int l0 = s_meta[mslot * 4 + 1];
      const int nblk = s_meta[mslot * 4 + 2];
      const int own0 = s_meta[mslot * 4 + 3];
      const int d0 = minor * BN + int(rank) * (BN / 2);  // this CTA's half of the dense tile
      if (lane == 0 && t == t_first) BSMM_STAMP(2);
      for (int ic = 0; ic < nblk; ic += 32) {
        const int* c_idx = m_idx + mslot * 32;
        const int* c_blk = m_blk + mslot * 32;
        if (ic > 0) {   // walks longer than 32 blocks: further chunks are loaded here
          if (ic + lane < nblk) {
            s_idx[buf * 32 + lane] = int(__ldg(p.walk_indices + l0 + ic + lane));
            s_blk[buf * 32 + lane] =
                p.walk_blocks ? __ldg(p.walk_blocks + own0 + ic + lane) : own0 + ic + lane;
          }
          c_idx = s_idx + buf * 32;
          c_blk = s_blk + buf * 32;
          __syncwarp();
        }
        if (lane == 0 && t == t_first && ic == 0) BSMM_STAMP(3);
        if (lane == 0) {
          const int cnt = min(32, nblk - ic);
          // specialised on the operand width (see bsmm_gemm_kernel's producer)
          auto blk_loop = [&](auto bkc) {
            constexpr int bk = decltype(bkc)::value;
            [[maybe_unused]] constexpr bool two_slabs = bk == kBK;   // MN-major: slabs of bk elements x bk k-rows
            for (int i = 0; i < cnt; ++i) {
              const int blk_row0 = c_blk[i] * kBlock;
              const int dense_k0 = c_idx[i] * kBlock;
#pragma unroll
              for (int kh = 0; kh < kBlock / bk; ++kh) {   // 2 stages per block (1: e4m3)
                ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                uint8_t* sa = smem + stage * T::kStageBytes;
                uint8_t* sb = sa + T::kABytes;
                if (leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * T::kStageBytes);
                const uint32_t bar = ptx::map_to_rank(&full_bar[stage], 0);
                if constexpr (A_MN) {
                  ptx::tma_load_2d_pair(&map_a, bar, sa, 0, blk_row0 + kh * bk);
                  if constexpr (two_slabs)
                    ptx::tma_load_2d_pair(&map_a, bar, sa + 8192, 64, blk_row0 + kh * bk);
                } else {
                  ptx::tma_load_2d_pair(&map_a, bar, sa, kh * bk, blk_row0);
                }
                const int k0 = dense_k0 + kh * bk;
                if constexpr (B_MN) {
                  ptx::tma_load_2d_pair(&map_b, bar, sb, d0, k0);
                  if constexpr (two_slabs) ptx::tma_load_2d_pair(&map_b, bar, sb + 8192, d0 + 64, k0);
                } else {
                  ptx::tma_load_2d_pair(&map_b, bar, sb, k0, d0);
                }
                if (++stage == S) { stage = 0; phase ^= 1; }
              }
            }
          };
          if (p.f8) blk_loop(std::integral_constant<int, 2 * kBK>{});
          else      blk_loop(std::integral_constant<int, kBK>{});
        }
        __syncwarp();
        buf ^= 1;
      }
      if (lane == 0) ptx::mbar_arrive(&meta_empty[mslot]);  // after the __syncwarp: slot consumed
      if (++mslot == 2) { mslot = 0; mphase ^= 1; }
    }
  } else if (warp == 1) {
    // ===================================================================== MMA issuer (leader)
    if (leader && lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      // the walk length of tile t + t_step is requested while tile t is being issued
      auto walk_len = [&](int t) {
        int pair, minor_unused;
        pair_dsd_tile(p, t, &pair, &minor_unused);
        return __ldg(p.walk_offsets + 2 * pair + 1) - __ldg(p.walk_offsets + 2 * pair);
      };
      int len_next = t_first < num_tiles ? walk_len(t_first) : 0;
      for (int t = t_first; t < num_tiles; t += t_step) {
        const int nsteps = (p.f8 ? 1 : 2) * len_next;
        if (t + t_step < num_tiles) len_next = walk_len(t + t_step);
        ptx::mbar_wait_cluster(&tempty_bar[as], aphase ^ 1);
        ptx::tc_fence_after_sync();
        const bool last = t + t_step >= num_tiles;
        if (nsteps == 0) {  // nothing accumulated: both epilogues write zeros
          ptx::mbar_arrive(&tfull_bar[as]);
          ptx::mbar_arrive_cluster(ptx::map_to_rank(&tfull_bar[as], 1));
          if (last) {
            ptx::mbar_arrive(last_bar);
            ptx::mbar_arrive_cluster(ptx::map_to_rank(last_bar, 1));
          }
        }
        const uint32_t tmem_d = tmem_base + uint32_t(as * BN);
        // specialised on the operand width: constant descriptor advances, no per-step branch
        auto mma_loop = [&](auto f8c) {
          constexpr bool F8 = decltype(f8c)::value;
          constexpr int bk = F8 ? 2 * kBK : kBK;
          constexpr uint32_t kAdvA = A_MN ? uint32_t(bk * 2) : (32 >> 4);   // see bsmm_gemm_kernel
          constexpr uint32_t kAdvB = B_MN ? uint32_t(bk * 2) : (32 >> 4);
          constexpr uint32_t lbo_mn = uint32_t(bk) * 128u;
          for (int s = 0; s < nsteps; ++s) {
            ptx::mbar_wait(&full_bar[stage], phase);
            ptx::tc_fence_after_sync();
            if (t == t_first && s == 0) BSMM_STAMP(4);
            const uint32_t sa = ptx::smem_u32(smem + stage * T::kStageBytes);
            const uint32_t sb = sa + T::kABytes;
            const uint64_t da = ptx::make_smem_desc(sa, A_MN ? lbo_mn : 16, 1024);
            const uint64_t db = ptx::make_smem_desc(sb, B_MN ? lbo_mn : 16, 1024);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              if constexpr (F8)
                ptx::umma_f8_pair(tmem_d, da + uint64_t(k * kAdvA), db + uint64_t(k * kAdvB),
                                  p.idesc, uint32_t((s | k) != 0));
              else
                ptx::umma_f16_pair(tmem_d, da + uint64_t(k * kAdvA), db + uint64_t(k * kAdvB),
                                   p.idesc, uint32_t((s | k) != 0));
            }
            ptx::umma_commit_pair(&empty_bar[stage]);
            if (s == nsteps - 1) {
              ptx::umma_commit_pair(&tfull_bar[as]);
              if (last) ptx::umma_commit_pair(last_bar);  // wakes the helper warps (tail drain)
            }
            if (++stage == S) { stage = 0; phase ^= 1; }
          }
        };
        if (p.f8) mma_loop(std::true_type{}); else mma_loop(std::false_type{});
        if (++as == 2) { as = 0; aphase ^= 1; }
        BSMM_STAMP(5);
      }
    }
  } else if (warp >= 4) {
    // ===================================================================== epilogue
    const int q = warp & 3;
    int as = 0;
    uint32_t aphase = 0;
    int epi_flip = 0;
    for (int t = t_first; t < num_tiles; t += t_step) {
      int pair, minor;
      pair_dsd_tile(p, t, &pair, &minor);
      const int major = 2 * pair + int(rank);
      const bool nonempty =
          __ldg(p.walk_offsets + 2 * pair + 1) != __ldg(p.walk_offsets + 2 * pair);
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after_sync();
      if (warp == 4 && lane == 0 && t == t_first) BSMM_STAMP(6);
      const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * BN);
      // the CTA's last tile: warps 0-3 (idle by now) drain columns 128.. (see below)
      const bool last = t + t_step >= num_tiles;
      store_dsd_tile<TRANS_OUT>(p, taddr, epi_smem, q, lane, major, minor, nonempty, last ? 1 : 0,
                                &map_c, &epi_flip);
      ptx::tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(ptx::map_to_rank(&tempty_bar[as], 0));
      if (++as == 2) { as = 0; aphase ^= 1; }
      if (warp == 4 && lane == 0) BSMM_STAMP(t == t_first ? 7 : 8);
    }
    tma_epilogue_done(lane);
  }
  if (warp == 3) {
    // ===================================================================== tile decoder
    int mslot = 0;
    uint32_t mphase = 0;
    for (int t = t_first; t < num_tiles; t += t_step) {
      int pair, minor;
      pair_dsd_tile(p, t, &pair, &minor);
      // Walk length and dense coordinates come from the pair's first row (twins: identical);
      // the blocks are this CTA's own row's.
      const int l0 = __ldg(p.walk_offsets + 2 * pair);
      const int nblk = __ldg(p.walk_offsets + 2 * pair + 1) - l0;
      const int own0 = __ldg(p.walk_offsets + 2 * pair + int(rank));
      int idx = 0, blk = 0;
      if (lane < nblk) {
        idx = int(__ldg(p.walk_indices + l0 + lane));
        blk = p.walk_blocks ? __ldg(p.walk_blocks + own0 + lane) : own0 + lane;
      }
      ptx::mbar_wait(&meta_empty[mslot], mphase ^ 1);  // the loads above are in flight meanwhile
      m_idx[mslot * 32 + lane] = idx;
      m_blk[mslot * 32 + lane] = blk;
      if (lane == 0) {
        s_meta[mslot * 4 + 0] = minor;
        s_meta[mslot * 4 + 1] = l0;
        s_meta[mslot * 4 + 2] = nblk;
        s_meta[mslot * 4 + 3] = own0;
      }
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&meta_full[mslot]);
      if (++mslot == 2) { mslot = 0; mphase ^= 1; }
    }
  }
  if (warp < 4 && t_first < num_tiles) {
    // Tail: nothing overlaps the last tile's epilogue, so the producer / MMA / allocator warps
    // help with it.  Warp q reads the same TMEM lane quarter as epilogue warp 4 + q, the other
    // half of the columns; the pipeline stages are free by now (the accumulator-complete commit
    // follows the last MMA's reads) and serve as staging memory.
    __syncwarp();
    const int count = (num_tiles - 1 - t_first) / t_step + 1;  // tiles of this cluster
    const int t = t_first + (count - 1) * t_step;
    const int as = (count - 1) & 1;
    int pair, minor;
    pair_dsd_tile(p, t, &pair, &minor);
    const bool nonempty =
        __ldg(p.walk_offsets + 2 * pair + 1) != __ldg(p.walk_offsets + 2 * pair);
    ptx::mbar_wait(last_bar, 0);  // single-use barrier: no phase to track across tiles
    ptx::tc_fence_after_sync();
    const uint32_t taddr = tmem_base + (uint32_t(warp * 32) << 16) + uint32_t(as * BN);
    int epi_flip = 0;
    store_dsd_tile<TRANS_OUT>(p, taddr, smem, warp, lane, 2 * pair + int(rank), minor, nonempty, 2,
                              &map_c, &epi_flip);
    ptx::tc_fence_before_sync();
    tma_epilogue_done(lane);
  }

  __syncwarp();
  ptx::tc_fence_before_sync();
  ptx::cluster_sync();  // no CTA may exit (or free TMEM) while its peer still signals it
  if (threadIdx.x == 0) BSMM_STAMP(9);
  if (warp == 2) {
    ptx::tc_fence_after_sync();
    ptx::tmem_dealloc_pair(tmem_base, T::kTmemCols);
  }
}

// ------------------------------------------------------------------------------------ SDD
// 2-SM SDD for twin block rows with an even number of blocks each (BSMM_FLAG_TWIN_ROWS |
// BSMM_FLAG_EVEN_ROWS): tile t = the 2 x 2 clique {rows 2p, 2p+1} x {two consecutive blocks c0, c1
// of those rows}.  CTA r supplies the 128 rows of op(A) of block row 2p + r and the column panel
// of op(B) for block c_r; it receives the accumulator rows of its block row for BOTH column
// panels, i.e. two blocks that are adjacent in BCSR storage.  Blocks of a twin pair occupy
// [offsets[2p], offsets[2p] + 2n), so with all n even tile t owns leader blocks 4t' .. and the
// pair is found from row_indices[4t] (no prefix sum needed).
struct PairSddTile {
  int row0;      // element row of this CTA's block row
  int col_own;   // element column of the panel this CTA loads
  int block0;    // first of this CTA's two output blocks
};

// Warp-collective.  Tile t -> clique, rasterised so that concurrently running clusters share
// operand panels: consecutive twin pairs with the same length and first column (the block rows of
// one dMoE expert; at most 8 pairs, inside an aligned window so every tile agrees on the group)
// form a group, and the group's tiles are walked column-pair-major — 74 clusters then cover all g
// row pairs x 74/g column pairs instead of 74 column pairs of ONE row pair, so each panel of
// op(B) is fetched from HBM once per group instead of once per row pair.  The remap is a bijection
// on the group's tile range whatever the hints say (all its pairs have the same length).
// PEEK = true: the pre-wait prefetch variant — L2-only loads, and every index is clamped so that
// whatever the (possibly not yet final) metadata holds, no load leaves the arrays.
template <bool PEEK = false>
__device__ __forceinline__ PairSddTile pair_sdd_tile(const GemmParams& p, int t, int rank,
                                                     int lane) {
  auto ld16 = [&](const short* q) { return PEEK ? ptx::ld_cg_s16(q) : int(__ldg(q)); };
  auto ld32 = [&](const int* q) { return PEEK ? ptx::ld_cg_s32(q) : __ldg(q); };
  // blocks per tile: 4 (a 2 x 2 clique), or 2 with BSMM_FLAG_SWIGLU (rows 2p, 2p + 1 of ONE output
  // block column j; CTA r then loads panel 2j + r of op(B): gate | up)
  const int sh = p.swiglu ? 1 : 2;
  auto blk = [&](int b) { return PEEK ? min(max(b, 0), (p.tiles_major << sh) - 1) : b; };
  int pr = ld16(p.sdd_rows + (t << sh)) >> 1;
  if (PEEK) pr = min(max(pr, 0), (p.sparse_lim >> 8) - 1);
  const int w0 = pr & ~7;
  const int q = w0 + (lane & 7);
  const bool valid = q < (p.sparse_lim >> 8);
  const int bq = valid ? ld32(p.walk_offsets + 2 * q) : -1;
  const int nq = valid ? ld32(p.walk_offsets + 2 * q + 1) - bq : -1;
  const int fq = nq > 0 ? ld16(p.sdd_cols + blk(bq)) : -1;
  const int me = pr - w0;
  const int n = __shfl_sync(0xffffffffu, nq, me);
  const int f = __shfl_sync(0xffffffffu, fq, me);
  const unsigned same = __ballot_sync(0xffffffffu, valid && nq == n && fq == f) & 0xffu;
  int gs = me, ge = me + 1;
  while (gs > 0 && ((same >> (gs - 1)) & 1u)) --gs;
  while (ge < 8 && ((same >> ge) & 1u)) ++ge;
  const int g = ge - gs;
  const int u = t - (__shfl_sync(0xffffffffu, bq, gs) >> sh);
  const int pl = PEEK ? min(max(gs + u % g, 0), 7) : gs + u % g;
  const int j = (u / g) << (sh - 1);
  const int base = __shfl_sync(0xffffffffu, bq, pl);
  PairSddTile r;
  r.row0 = (2 * (w0 + pl) + rank) * kBlock;
  if (p.swiglu)
    r.col_own = (2 * (ld16(p.sdd_cols + blk(base + j)) & 0x7fff) + rank) * kBlock;
  else
    r.col_own = (ld16(p.sdd_cols + blk(base + j + rank)) & 0x7fff) * kBlock;
  r.block0 = base + rank * n + j;
  return r;
}

// p.walk_offsets = c.offsets, p.sdd_rows = c.row_indices, p.sdd_cols = c.indices,
// p.tiles_major = nnz blocks / 4, p.k_steps = ceil(k / 64), p.idesc = UMMA 256 x 256.
template <bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
bsmm_pair_sdd_kernel(const __grid_constant__ CUtensorMap map_a,
                     const __grid_constant__ CUtensorMap map_b,
                     const __grid_constant__ CUtensorMap map_c, const GemmParams p) {
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
  uint64_t* last_bar = tempty_bar + 2;  // both CTAs: the cluster's LAST accumulator is complete
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(last_bar + 1);
  // Tile-decode ring (see bsmm_pair_kernel): warp 3 resolves the next tile's clique — three
  // dependent metadata round trips — while the producer is still issuing the current tile.
  uint64_t* meta_full = reinterpret_cast<uint64_t*>(tmem_ptr + 4);
  uint64_t* meta_empty = meta_full + 2;
  int* s_meta = reinterpret_cast<int*>(meta_empty + 2);  // [2][4]: row0, col_own, block0

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = int(ptx::cluster_ctarank());
  const bool leader = rank == 0;
  int num_tiles = p.tiles_major;
  const int t_first = int(ptx::cluster_id_x());
  const int t_step = int(ptx::num_clusters_x());

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&map_a);
    ptx::prefetch_tensormap(&map_b);
    ptx::prefetch_tensormap(&map_c);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      ptx::mbar_init(&full_bar[i], 1);
      ptx::mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&tfull_bar[i], 1);
      ptx::mbar_init(&tempty_bar[i], 8);
    }
    ptx::mbar_init(last_bar, 1);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&meta_full[i], 1);
      ptx::mbar_init(&meta_empty[i], 1);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc_pair(tmem_ptr, T::kTmemCols);
  ptx::tc_fence_before_sync();
  ptx::cluster_sync();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;
  // Everything above touched only parameters and shared memory: with programmatic dependent
  // launch it overlaps the tail of the previous kernel in the stream.
  if (warp == 3 && p.prefetch && p.bk == kBK && t_first < num_tiles && !p.nnz_dev) {
    // L2 prefetch of the first tile's first k-steps before the wait (hint-only metadata peek)
    const PairSddTile tile = pair_sdd_tile<true>(p, t_first, rank, lane);
    if (lane < min(p.k_steps, kPrefetchSteps)) {
      const int k0 = lane * kBK;
      if constexpr (A_MN) {
        ptx::tma_prefetch_2d(&map_a, tile.row0, k0);
        ptx::tma_prefetch_2d(&map_a, tile.row0 + 64, k0);
      } else {
        ptx::tma_prefetch_2d(&map_a, k0, tile.row0);
      }
      if constexpr (B_MN) {
        ptx::tma_prefetch_2d(&map_b, tile.col_own, k0);
        ptx::tma_prefetch_2d(&map_b, tile.col_own + 64, k0);
      } else {
        ptx::tma_prefetch_2d(&map_b, k0, tile.col_own);
      }
    }
  }
  ptx::grid_dep_wait();
  ptx::grid_dep_launch();
  if (p.nnz_dev) num_tiles = min(num_tiles, __ldg(p.nnz_dev) >> (p.swiglu ? 1 : 2));  // DEVICE_NNZ

  if (warp == 0) {
    // ===================================================================== TMA producer
    int stage = 0;
    uint32_t phase = 0;
    int mslot = 0;
    uint32_t mphase = 0;
    for (int t = t_first; t < num_tiles; t += t_step) {
      ptx::mbar_wait(&meta_full[mslot], mphase);   // decoded ahead by warp 3 (below)
      PairSddTile tile;
      tile.row0 = s_meta[mslot * 4 + 0];
      tile.col_own = s_meta[mslot * 4 + 1];
      tile.block0 = s_meta[mslot * 4 + 2];
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&meta_empty[mslot]);
      if (++mslot == 2) { mslot = 0; mphase ^= 1; }
      if (lane == 0) {
        for (int s = 0; s < p.k_steps; ++s) {
          const int k0 = s * p.bk;
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * T::kStageBytes;
          uint8_t* sb = sa + T::kABytes;
          if (leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * T::kStageBytes);
          const uint32_t bar = ptx::map_to_rank(&full_bar[stage], 0);
          const bool two_slabs = p.bk == kBK;
          if constexpr (A_MN) {  // op(A) stored [k][m]: two [64 k][64 m] slabs (one of 128: e4m3)
            ptx::tma_load_2d_pair(&map_a, bar, sa, tile.row0, k0);
            if (two_slabs) ptx::tma_load_2d_pair(&map_a, bar, sa + 8192, tile.row0 + 64, k0);
          } else {
            ptx::tma_load_2d_pair(&map_a, bar, sa, k0, tile.row0);
          }
          if constexpr (B_MN) {  // op(B) stored [k][n]
            ptx::tma_load_2d_pair(&map_b, bar, sb, tile.col_own, k0);
            if (two_slabs) ptx::tma_load_2d_pair(&map_b, bar, sb + 8192, tile.col_own + 64, k0);
          } else {
            ptx::tma_load_2d_pair(&map_b, bar, sb, k0, tile.col_own);
          }
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================================== MMA issuer (leader)
    if (leader && lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      const uint32_t kAdvA = A_MN ? uint32_t(p.bk * 2) : (32 >> 4);   // see bsmm_gemm_kernel
      const uint32_t kAdvB = B_MN ? uint32_t(p.bk * 2) : (32 >> 4);
      const uint32_t lbo_mn = uint32_t(p.bk) * 128u;
      for (int t = t_first; t < num_tiles; t += t_step) {
        ptx::mbar_wait_cluster(&tempty_bar[as], aphase ^ 1);
        ptx::tc_fence_after_sync();
        const uint32_t tmem_d = tmem_base + uint32_t(as * BN);
        for (int s = 0; s < p.k_steps; ++s) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after_sync();
          const uint32_t sa = ptx::smem_u32(smem + stage * T::kStageBytes);
          const uint32_t sb = sa + T::kABytes;
          const uint64_t da = ptx::make_smem_desc(sa, A_MN ? lbo_mn : 16, 1024);
          const uint64_t db = ptx::make_smem_desc(sb, B_MN ? lbo_mn : 16, 1024);
          if (p.f8) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma_f8_pair(tmem_d, da + uint64_t(k * kAdvA), db + uint64_t(k * kAdvB), p.idesc,
                                uint32_t((s | k) != 0));
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma_f16_pair(tmem_d, da + uint64_t(k * kAdvA), db + uint64_t(k * kAdvB), p.idesc,
                                 uint32_t((s | k) != 0));
          }
          ptx::umma_commit_pair(&empty_bar[stage]);
          if (s == p.k_steps - 1) {
            ptx::umma_commit_pair(&tfull_bar[as]);
            if (t + t_step >= num_tiles) ptx::umma_commit_pair(last_bar);  // tail drain
          }
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ===================================================================== epilogue
    const int q = warp & 3;
    int as = 0;
    uint32_t aphase = 0;
    uint16_t* const cbase = reinterpret_cast<uint16_t*>(p.c);
    int epi_flip = 0;
    for (int t = t_first; t < num_tiles; t += t_step) {
      const PairSddTile tile = pair_sdd_tile(p, t, rank, lane);
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after_sync();
      const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * BN);
      if (p.swiglu) {   // one block per CTA: silu(columns 0..127) * columns 128..255
#pragma unroll 1
        for (int ch = 0; ch < 2; ++ch)
          tma_chunk64_swiglu(taddr + uint32_t(ch * 64), epi_smem + q * T::kEpiWarpBytes, lane, &map_c,
                             ch * 64, tile.block0 * kBlock + q * 32, true, p.bf16, &epi_flip);
        ptx::tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive_cluster(ptx::map_to_rank(&tempty_bar[as], 0));
        if (++as == 2) { as = 0; aphase ^= 1; }
        continue;
      }
      // the CTA's last tile: warps 0-3 store the second block (columns 128..255, below)
      const int nch = t + t_step >= num_tiles ? 2 : 4;
#pragma unroll 1
      for (int ch = 0; ch < nch; ++ch) {
        // columns [0,128) -> block0, [128,256) -> block0 + 1 (adjacent in BCSR storage)
        uint16_t* o = cbase + (size_t(tile.block0) + size_t(ch >> 1)) * (kBlock * kBlock) +
                      size_t(q * 32) * kBlock + (ch & 1) * 64 + (lane & 7) * 8;
        store_chunk64(taddr + uint32_t(ch * 64), epi_smem + q * T::kEpiWarpBytes, lane, o, kBlock,
                      true, true, p.bf16, 32, nullptr, 0, 64, &map_c, (ch & 1) * 64,
                      (tile.block0 + (ch >> 1)) * kBlock + q * 32, &epi_flip);
      }
      ptx::tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(ptx::map_to_rank(&tempty_bar[as], 0));
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    tma_epilogue_done(lane);
  }
  if (warp == 3) {
    // ===================================================================== tile decoder
    int mslot = 0;
    uint32_t mphase = 0;
    for (int t = t_first; t < num_tiles; t += t_step) {
      const PairSddTile tile = pair_sdd_tile(p, t, rank, lane);
      ptx::mbar_wait(&meta_empty[mslot], mphase ^ 1);
      if (lane == 0) {
        s_meta[mslot * 4 + 0] = tile.row0;
        s_meta[mslot * 4 + 1] = tile.col_own;
        s_meta[mslot * 4 + 2] = tile.block0;
        ptx::mbar_arrive(&meta_full[mslot]);
      }
      if (++mslot == 2) { mslot = 0; mphase ^= 1; }
    }
  }
  if (warp < 4 && t_first < num_tiles && !p.swiglu) {
    // Tail drain (as in bsmm_pair_kernel): the idle warps store the last tile's second block.
    __syncwarp();
    const int count = (num_tiles - 1 - t_first) / t_step + 1;
    const int t = t_first + (count - 1) * t_step;
    const int as = (count - 1) & 1;
    int epi_flip = 0;
    const PairSddTile tile = pair_sdd_tile(p, t, rank, lane);
    ptx::mbar_wait(last_bar, 0);
    ptx::tc_fence_after_sync();
    const uint32_t taddr = tmem_base + (uint32_t(warp * 32) << 16) + uint32_t(as * BN);
    uint16_t* const cbase = reinterpret_cast<uint16_t*>(p.c);
#pragma unroll 1
    for (int ch = 2; ch < 4; ++ch) {
      uint16_t* o = cbase + (size_t(tile.block0) + size_t(ch >> 1)) * (kBlock * kBlock) +
                    size_t(warp * 32) * kBlock + (ch & 1) * 64 + (lane & 7) * 8;
      store_chunk64(taddr + uint32_t(ch * 64), smem + warp * T::kEpiWarpBytes, lane, o, kBlock, true,
                    true, p.bf16, 32, nullptr, 0, 64, &map_c, (ch & 1) * 64,
                    (tile.block0 + (ch >> 1)) * kBlock + warp * 32, &epi_flip);
    }
    ptx::tc_fence_before_sync();
    tma_epilogue_done(lane);
  }

  __syncwarp();
  ptx::tc_fence_before_sync();
  ptx::cluster_sync();
  if (warp == 2) {
    ptx::tc_fence_after_sync();
    ptx::tmem_dealloc_pair(tmem_base, T::kTmemCols);
  }
}

// ------------------------------------------------------------------------------------ DDS
// 2-SM DDS computed directly (not as a transposed DSD): M side = 256 rows of op(A) (128 per CTA),
// N side = the blocks of two TWIN walk rows of the sparse operand (block columns 2c, 2c+1 of B,
// or block rows when B is used transposed), one block per CTA.  The accumulator of CTA r is
// C[m0 + 128 r .. +128][256 c .. +256]: row-major, 512 B runs, no transposing epilogue.
//   A_MN: op(A) stored [k][m];  B_MN: the sparse blocks are read as stored [k][n] (B not
//   transposed: walk by block column through offsets_t / indices_t / block_offsets).
// p.tiles_major = walk rows (even), p.tiles_minor = ceil(m / 256), p.dense_lim = m,
// p.idesc = UMMA 256 x 256 with (A_MN, B_MN).
template <bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
bsmm_pair_dds_kernel(const __grid_constant__ CUtensorMap map_a,
                     const __grid_constant__ CUtensorMap map_b,
                     const __grid_constant__ CUtensorMap map_c, const GemmParams p) {
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
  uint64_t* last_bar = tempty_bar + 2;  // both CTAs: the cluster's LAST accumulator is complete
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(last_bar + 1);
  int* s_idx = reinterpret_cast<int*>(tmem_ptr + 4);  // [2][32] k block of each step
  int* s_blk = s_idx + 64;                            // [2][32] this CTA's block id

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = int(ptx::cluster_ctarank());
  const bool leader = rank == 0;
  const int m_tiles = p.tiles_minor;
  const int num_tiles = (p.tiles_major >> 1) * m_tiles;
  const int t_first = int(ptx::cluster_id_x());
  const int t_step = int(ptx::num_clusters_x());

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&map_a);
    ptx::prefetch_tensormap(&map_b);
    ptx::prefetch_tensormap(&map_c);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      ptx::mbar_init(&full_bar[i], 1);
      ptx::mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&tfull_bar[i], 1);
      ptx::mbar_init(&tempty_bar[i], 8);
    }
    ptx::mbar_init(last_bar, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc_pair(tmem_ptr, T::kTmemCols);
  ptx::tc_fence_before_sync();
  ptx::cluster_sync();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;
  // Everything above touched only parameters and shared memory: with programmatic dependent
  // launch it overlaps the tail of the previous kernel in the stream.
  if (warp == 3 && p.prefetch && p.bk == kBK && t_first < num_tiles) {
    // L2 prefetch of the first tile's first operand boxes before the wait (hint-only peek)
    const int cp = t_first / m_tiles;
    const int m0 = (t_first - cp * m_tiles) * 256 + rank * kBlock;
    const int l0 = ptx::ld_cg_s32(p.walk_offsets + 2 * cp);
    const int l1 = ptx::ld_cg_s32(p.walk_offsets + 2 * cp + 1);
    const int own0 = leader ? l0 : l1;
    const int half = lane & 1, i = lane >> 1;
    if (l0 >= 0 && i < kPrefetchBlocks && l0 + i < l1 && l1 <= (1 << 17)) {
      const int k0 = (ptx::ld_cg_s16(p.walk_indices + l0 + i) & 0x7fff) * kBlock + half * kBK;
      const int blk =
          (p.walk_blocks ? (ptx::ld_cg_s32(p.walk_blocks + own0 + i) & 0x1ffff) : own0 + i) * kBlock;
      if constexpr (A_MN) {
        ptx::tma_prefetch_2d(&map_a, m0, k0);
        ptx::tma_prefetch_2d(&map_a, m0 + 64, k0);
      } else {
        ptx::tma_prefetch_2d(&map_a, k0, m0);
      }
      if constexpr (B_MN) {
        ptx::tma_prefetch_2d(&map_b, 0, blk + half * kBK);
        ptx::tma_prefetch_2d(&map_b, 64, blk + half * kBK);
      } else {
        ptx::tma_prefetch_2d(&map_b, half * kBK, blk);
      }
    }
  }
  ptx::grid_dep_wait();
  ptx::grid_dep_launch();

  if (warp == 0) {
    // ===================================================================== TMA producer
    int stage = 0;
    uint32_t phase = 0;
    int buf = 0;
    for (int t = t_first; t < num_tiles; t += t_step) {
      const int cp = t / m_tiles;                                   // twin pair of walk rows
      const int m0 = (t - cp * m_tiles) * 256 + rank * kBlock;      // this CTA's rows of op(A)
      const int l0 = __ldg(p.walk_offsets + 2 * cp);
      const int nblk = __ldg(p.walk_offsets + 2 * cp + 1) - l0;
      const int own0 = __ldg(p.walk_offsets + 2 * cp + rank);
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
#pragma unroll 2
            for (int kh = 0; kh * p.bk < kBlock; ++kh) {   // 2 stages per block (1: e4m3)
              ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
              uint8_t* sa = smem + stage * T::kStageBytes;
              uint8_t* sb = sa + T::kABytes;
              if (leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * T::kStageBytes);
              const uint32_t bar = ptx::map_to_rank(&full_bar[stage], 0);
              const int k0 = dense_k0 + kh * p.bk;
              const bool two_slabs = p.bk == kBK;
              if constexpr (A_MN) {  // op(A) stored [k][m]
                ptx::tma_load_2d_pair(&map_a, bar, sa, m0, k0);
                if (two_slabs) ptx::tma_load_2d_pair(&map_a, bar, sa + 8192, m0 + 64, k0);
              } else {               // stored [m][k]
                ptx::tma_load_2d_pair(&map_a, bar, sa, k0, m0);
              }
              if constexpr (B_MN) {  // block stored [k][n]: two [64 k][64 n] slabs
                ptx::tma_load_2d_pair(&map_b, bar, sb, 0, blk_row0 + kh * p.bk);
                if (two_slabs) ptx::tma_load_2d_pair(&map_b, bar, sb + 8192, 64, blk_row0 + kh * p.bk);
              } else {               // block stored [n][k]
                ptx::tma_load_2d_pair(&map_b, bar, sb, kh * p.bk, blk_row0);
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
      const uint32_t kAdvA = A_MN ? uint32_t(p.bk * 2) : (32 >> 4);   // see bsmm_gemm_kernel
      const uint32_t kAdvB = B_MN ? uint32_t(p.bk * 2) : (32 >> 4);
      const uint32_t lbo_mn = uint32_t(p.bk) * 128u;
      for (int t = t_first; t < num_tiles; t += t_step) {
        const int cp = t / m_tiles;
        const int nsteps = (p.f8 ? 1 : 2) *
            (__ldg(p.walk_offsets + 2 * cp + 1) - __ldg(p.walk_offsets + 2 * cp));
        ptx::mbar_wait_cluster(&tempty_bar[as], aphase ^ 1);
        ptx::tc_fence_after_sync();
        const bool last = t + t_step >= num_tiles;
        if (nsteps == 0) {
          ptx::mbar_arrive(&tfull_bar[as]);
          ptx::mbar_arrive_cluster(ptx::map_to_rank(&tfull_bar[as], 1));
          if (last) {
            ptx::mbar_arrive(last_bar);
            ptx::mbar_arrive_cluster(ptx::map_to_rank(last_bar, 1));
          }
        }
        const uint32_t tmem_d = tmem_base + uint32_t(as * BN);
        for (int s = 0; s < nsteps; ++s) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after_sync();
          const uint32_t sa = ptx::smem_u32(smem + stage * T::kStageBytes);
          const uint32_t sb = sa + T::kABytes;
          const uint64_t da = ptx::make_smem_desc(sa, A_MN ? lbo_mn : 16, 1024);
          const uint64_t db = ptx::make_smem_desc(sb, B_MN ? lbo_mn : 16, 1024);
          if (p.f8) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma_f8_pair(tmem_d, da + uint64_t(k * kAdvA), db + uint64_t(k * kAdvB), p.idesc,
                                uint32_t((s | k) != 0));
          } else {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              ptx::umma_f16_pair(tmem_d, da + uint64_t(k * kAdvA), db + uint64_t(k * kAdvB), p.idesc,
                                 uint32_t((s | k) != 0));
          }
          ptx::umma_commit_pair(&empty_bar[stage]);
          if (s == nsteps - 1) {
            ptx::umma_commit_pair(&tfull_bar[as]);
            if (last) ptx::umma_commit_pair(last_bar);  // wakes the helper warps (tail drain)
          }
          if (++stage == S) { stage = 0; phase ^= 1; }
        }
        if (++as == 2) { as = 0; aphase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ===================================================================== epilogue
    const int q = warp & 3;
    int as = 0;
    uint32_t aphase = 0;
    uint16_t* const cbase = reinterpret_cast<uint16_t*>(p.c);
    int epi_flip = 0;
    for (int t = t_first; t < num_tiles; t += t_step) {
      const int cp = t / m_tiles;
      const int row0 = (t - cp * m_tiles) * 256 + rank * kBlock + q * 32;
      const bool nonempty =
          __ldg(p.walk_offsets + 2 * cp + 1) != __ldg(p.walk_offsets + 2 * cp);
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after_sync();
      const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16) + uint32_t(as * BN);
      const int rows_ok = p.dense_lim - row0;  // <= 0: this warp's rows are all out of range
      // the CTA's last tile: warps 0-3 drain columns 128.. (below)
      const int nch = t + t_step >= num_tiles ? 2 : 4;
#pragma unroll 1
      for (int ch = 0; ch < nch; ++ch) {
        uint16_t* o = cbase + size_t(row0) * size_t(p.ldc) + size_t(cp) * 256 + ch * 64 +
                      (lane & 7) * 8;
        store_chunk64(taddr + uint32_t(ch * 64), epi_smem + q * T::kEpiWarpBytes, lane, o,
                      size_t(p.ldc), true, nonempty, p.bf16, rows_ok, nullptr, 0, 64, &map_c,
                      cp * 256 + ch * 64, row0, &epi_flip);
      }
      ptx::tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(ptx::map_to_rank(&tempty_bar[as], 0));
      if (++as == 2) { as = 0; aphase ^= 1; }
    }
    tma_epilogue_done(lane);
  }
  if (warp < 4 && t_first < num_tiles) {
    // Tail drain (as in bsmm_pair_kernel): the idle warps store columns 128..255 of the last
    // tile, staged through the (by now free) first pipeline stage.
    __syncwarp();
    const int count = (num_tiles - 1 - t_first) / t_step + 1;
    const int t = t_first + (count - 1) * t_step;
    const int as = (count - 1) & 1;
    int epi_flip = 0;
    const int cp = t / m_tiles;
    const int row0 = (t - cp * m_tiles) * 256 + rank * kBlock + warp * 32;
    const bool nonempty =
        __ldg(p.walk_offsets + 2 * cp + 1) != __ldg(p.walk_offsets + 2 * cp);
    ptx::mbar_wait(last_bar, 0);
    ptx::tc_fence_after_sync();
    const uint32_t taddr = tmem_base + (uint32_t(warp * 32) << 16) + uint32_t(as * BN);
    uint16_t* const cbase = reinterpret_cast<uint16_t*>(p.c);
#pragma unroll 1
    for (int ch = 2; ch < 4; ++ch) {
      uint16_t* o = cbase + size_t(row0) * size_t(p.ldc) + size_t(cp) * 256 + ch * 64 +
                    (lane & 7) * 8;
      store_chunk64(taddr + uint32_t(ch * 64), smem + warp * T::kEpiWarpBytes, lane, o,
                    size_t(p.ldc), true, nonempty, p.bf16, p.dense_lim - row0, nullptr, 0, 64, &map_c,
                    cp * 256 + ch * 64, row0, &epi_flip);
    }
    ptx::tc_fence_before_sync();
    tma_epilogue_done(lane);
  }

  __syncwarp();
  ptx::tc_fence_before_sync();
  ptx::cluster_sync();
  if (warp == 2) {
    ptx::tc_fence_after_sync();
    ptx::tmem_dealloc_pair(tmem_base, T::kTmemCols);
  }
}

}  // namespace bsmm
