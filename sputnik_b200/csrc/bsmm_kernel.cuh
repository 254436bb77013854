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
//  * every nonzero 128x128 block is contracted by tcgen05.mma (UMMA 128 x 256 x 16) with the
//    accumulator in TMEM, double-buffered (2 x 256 columns) so the epilogue of tile i overlaps
//    the mainloop of tile i+1.  A 128 x 128 UMMA tile would read shared memory at the SM's full
//    128 B/clk; every op is therefore shaped so that the N side is 256 wide:
//      - DSD: M side = the sparse block, N side = 256 columns of the dense operand;
//      - DDS: computed as the transposed DSD problem C^T = op(B)^T op(A)^T (M side = the sparse
//        block, N side = 256 ROWS of op(A)) with a transposing epilogue;
//      - SDD: M side = 128 rows of op(A), N side = the column panels of TWO consecutive nonzero
//        blocks (when they share a block row; otherwise two 128-wide passes);
//  * operand tiles are staged by TMA (128 B swizzle) whose coordinates come from the block
//    index metadata, which the producer warp prefetches into shared memory 32 blocks at a
//    time (the reference pays a dependent __ldg per block on the critical path);
//  * transposed access (trans_a on a sparse A, non-transposed sparse B) is a TMA gather
//    through block_offsets into an MN-major UMMA descriptor: no transpose is materialised;
//  * CTAs are persistent (grid = #SMs) and walk a tile list.
#pragma once

#include <type_traits>

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "ptx.cuh"

namespace bsmm {

enum : int { kOpSDD = 0, kOpDSD = 1 };

constexpr int kBlock = 128;       // sparse block edge == UMMA M
constexpr int kBK = 64;           // K elements per pipeline stage of 16-bit data (128 B per row);
                                  // 8-bit operands (e4m3) take 128: GemmParams::bk.  A stage is
                                  // always 128 B wide, so every byte count below is type-free.
constexpr int kThreads = 256;     // warp 0 TMA, 1 MMA, 2 TMEM alloc, 3 idle, 4-7 epilogue
constexpr int kSmemBudget = 200 * 1024;
constexpr int kTileQ = 4;         // depth of the in-CTA tile-id queue (producer -> MMA, epilogue)

struct GemmParams {
  // DSD (and DDS as transposed DSD): the walk over the sparse operand, already resolved by the
  // host to the row view (offsets, indices, sequential blocks) or the column view (offsets_t,
  // indices_t, block_offsets).
  const int* walk_offsets;
  const short* walk_indices;
  const int* walk_blocks;  // nullptr: block id == position in the walk
  // SDD: coordinates of output block t.
  const short* sdd_rows;
  const short* sdd_cols;
  void* c;
  // DSD with scattered output rows (bsmm_dsd_scatter): row i of C starts at c_rows[i] (local or
  // peer memory), a null entry drops the row; `c` / `ldc` are then unused.
  void* const* c_rows;
  long long ldc;    // elements
  int sparse_lim;   // extent of C along the sparse (M-side) dimension (always a multiple of 128)
  int dense_lim;    // extent of C along the dense (N-side) dimension (clipping)
  int tiles_major;  // SDD: nnz blocks (tile ids are block ids); DSD: block rows/cols of the walk
  int tiles_minor;  // SDD: 1; DSD: ceil(dense extent / BN)
  int k_steps;      // SDD: ceil(k / 64)
  int nb;           // SDD: nnz blocks
  uint32_t idesc;       // UMMA instruction descriptor with N = BN
  uint32_t idesc_half;  // SDD: N = 128 (unpaired blocks)
  int bf16;             // output format word: kFmtBf16 | activation << 1 | kFmtAccum
  int sdd_single;       // SDD: one block per tile (few blocks: all of them run concurrently)
  // SDD with BSMM_FLAG_SWIGLU: op(B) holds, for output block column j, the gate panel at block
  // column 2j and the up panel at 2j + 1; every tile is ONE output block computed as a 128 x 256
  // accumulator (gate | up) and stored as silu(gate) * up.
  int swiglu;
  // Dynamic tile scheduler: CTAs draw tile ids from this counter (nullptr: static round-robin).
  // It wraps back to 0 with the last draw of the launch (atomicInc limit), so it needs no reset.
  unsigned int* tile_counter;
  // SDD with BSMM_FLAG_DEVICE_NNZ: the live block count is read here (c.offsets + block rows);
  // `nb` is then only the capacity of the index / value arrays.
  const int* nnz_dev;
  // Stream-K schedule of the 2-SM DSD kernel (bsmm_pair_sk_kernel.cuh): fp32 partial-tile
  // workspace, its flags and the value that marks a slot as written by this launch.
  float* sk_ws;
  unsigned int* sk_flags;
  unsigned int sk_epoch;
  // Device-side kernel selection for callers that give no topology hints: the launch runs iff
  // ((*gate & gate_mask) == gate_mask) == (gate_want != 0); otherwise every CTA returns at once.
  // The word is written by topology_gate_kernel, launched just before (bsmm_api.cu, auto_pair).
  const unsigned int* gate;
  unsigned int gate_mask;
  int gate_want;
  // Prefetch the first tile's operands into L2 before griddepcontrol.wait (see the kernels).
  int prefetch;
  // Operand element type: K elements per 128 B stage row (64: fp16 / bf16, 128: e4m3) and whether
  // the MMAs are kind::f8f6f4.  Outputs are 16-bit either way.
  int bk;
  int f8;
  // 2-SM DSD: group-rasterised tile order (pair_dsd_tile); 0 = plain pair-major order
  int raster;
  // Whole-operand L2 warm-up (1-SM kernel, operands that fit L2 together): before
  // griddepcontrol.wait every CTA bulk-prefetches its 1/grid slice of both input operands, so the
  // compulsory first touch of each line streams in at DRAM bandwidth from the start of the launch
  // instead of stalling a pipeline stage for a DRAM round trip the first time a tile needs it.
  const void* warm_ptr[2];
  unsigned long long warm_bytes[2];  // 0: off
  // 1-SM DSD, BN = 256, dynamic schedule: tile ids >= split_from are HALF tiles (128 dense columns,
  // two per full tile) — the tail wave is drawn at half granularity.  0 = no split.
  int split_from;
#ifdef BSMM_TRACE
  unsigned long long* trace;  // [grid][16] globaltimer stamps (tools/trace build only)
#endif
};

// True when a gated launch must not run.  Called first thing by every thread of every GEMM kernel
// (before any barrier, TMEM or cluster operation, so returning is safe); waits for the previous
// grid — the one that wrote the gate word — when a gate is present.
__device__ __forceinline__ bool gated_off(const GemmParams& p) {
  if (!p.gate) return false;
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const bool on = (*reinterpret_cast<const volatile unsigned int*>(p.gate) & p.gate_mask) ==
                  p.gate_mask;
  return on != (p.gate_want != 0);
}

#ifdef BSMM_TRACE
// tools/trace.py experiments: bit 0 = epilogues skip their global stores, bit 1 = they skip the
// TMEM reads (zeros instead).  Trace build only.
__device__ int g_bsmm_dbg = 0;
#define BSMM_DBG(bit) (g_bsmm_dbg & (bit))
#else
#define BSMM_DBG(bit) 0
#endif

#ifdef BSMM_TRACE
__device__ __forceinline__ void trace_stamp(const GemmParams& p, int slot) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  if (p.trace) p.trace[size_t(blockIdx.x) * 16 + slot] = t;
}
#define BSMM_STAMP(slot) trace_stamp(p, slot)
#else
#define BSMM_STAMP(slot)
#endif

template <int BN>
struct Tiling {
  static constexpr int kABytes = kBlock * kBK * 2;
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = kSmemBudget / kStageBytes;
  static constexpr int kTmemCols = 2 * BN;  // two accumulator stages (power of two >= 32)
  // stages + barriers (full, empty, tmem_full[2], tmem_empty[2]) + tmem ptr + index staging
  static constexpr int kAuxBytes =
      (2 * kStages + 4) * 8 + 16 + 2 * 2 * 32 * 4 + kTileQ * (4 + 16) + 8 + 16;
  // epilogue staging: per epilogue warp TWO tiles of 32 rows x 128 B (row-major store; the TMA
  // store of one is still reading while the next is being filled) or of 32 x 32 16-bit (DDS)
  static constexpr int kEpiWarpBytes = 8192;
  static constexpr int kEpiBytes = 4 * kEpiWarpBytes;
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + kAuxBytes + 1024;
  static_assert(kSmemBytes <= 227 * 1024, "shared memory of one CTA");
};

// Output format word (GemmParams::bf16, SiblingParams::bf16): bit 0 = bf16 (else fp16), bits 1-2 =
// fused activation on the fp32 accumulator, bit 3 = accumulate into C.
enum : int { kFmtBf16 = 1, kFmtActShift = 1, kFmtActMask = 3, kFmtAccum = 8, kFmtDirect = 16, kFmtTma = 32 };
enum : int { kActNone = 0, kActRelu = 1, kActGelu = 2 };

// GELU, tanh form (what MegaBlocks' dMoE MLP applies to the SDD output): 0.5 x (1 + tanh(u)),
// u = sqrt(2/pi) (x + 0.044715 x^3); tanh.approx.f32 (rel. error ~2^-11, below 16-bit rounding).
__device__ __forceinline__ float gelu_tanh(float x) {
  const float u = 0.7978845608028654f * x * fmaf(0.044715f * x, x, 1.0f);
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(u));
  return 0.5f * x * (1.0f + t);
}

// ACT = false is the reference's epilogue (one cvt.rn per pair); the branch on the activation is
// taken once per 64-column chunk (stage_chunk64 / stage_chunk32_t below), never per element:
// ptxas if-converts a per-element branch and the plain path then pays for the predicated-off
// GELU (measured: C2 52.8 -> 57.5 us).
template <bool ACT, int BF = -1>
__device__ __forceinline__ uint32_t pack2(uint32_t lo_bits, uint32_t hi_bits, int fmt) {
  float lo = __uint_as_float(lo_bits), hi = __uint_as_float(hi_bits);
  if constexpr (ACT) {
    if (((fmt >> kFmtActShift) & kFmtActMask) == kActRelu) {
      lo = fmaxf(lo, 0.0f);
      hi = fmaxf(hi, 0.0f);
    } else {
      lo = gelu_tanh(lo);
      hi = gelu_tanh(hi);
    }
  }
  // BF = 1 / 0: the output type is a compile-time fact of the caller's chunk-level branch (ptxas
  // if-converts a per-element test into BOTH conversions plus a select); -1: decided here.
  if (BF == 1 || (BF < 0 && (fmt & kFmtBf16))) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
  __half2 v = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}

// 8 packed 16-bit values + the 8 already at `dst` (kFmtAccum: C += op(A) op(B)), rounded once
// more to the 16-bit type — bit for bit what the unfused sequence (product to a temporary, then an
// elementwise add) stores, without the temporary and its two extra passes over C.
__device__ __forceinline__ uint32_t add2(uint32_t a, uint32_t b, int fmt) {
  if (fmt & kFmtBf16) {
    const float2 x = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&a));
    const float2 y = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&b));
    __nv_bfloat162 v = __floats2bfloat162_rn(x.x + y.x, x.y + y.y);
    return *reinterpret_cast<uint32_t*>(&v);
  }
  const float2 x = __half22float2(*reinterpret_cast<const __half2*>(&a));
  const float2 y = __half22float2(*reinterpret_cast<const __half2*>(&b));
  __half2 v = __floats2half2_rn(x.x + y.x, x.y + y.y);
  return *reinterpret_cast<uint32_t*>(&v);
}
// The old values of a chunk are loaded BEFORE its accumulator is read from TMEM (all loads of the
// chunk in flight together, their latency under the tcgen05.ld / convert / staging work); loading
// each line right before its store made the accumulating epilogue 3x slower than the plain one.
__device__ __forceinline__ uint4 add4(uint4 w, const uint4& old, int fmt) {
  w.x = add2(w.x, old.x, fmt);
  w.y = add2(w.y, old.y, fmt);
  w.z = add2(w.z, old.z, fmt);
  w.w = add2(w.w, old.w, fmt);
  return w;
}

// 64 fp32 accumulator columns of this thread's row -> 16-bit, into the warp's XOR-swizzled
// staging tile (row = lane, 128 B per row).
template <bool ACT, int BF>
__device__ __forceinline__ void stage_chunk64_t(const uint32_t (&v)[64], uint8_t* tw, int lane,
                                                int fmt) {
#pragma unroll
  for (int g = 0; g < 8; ++g) {
    uint4 w;
    w.x = pack2<ACT, BF>(v[g * 8 + 0], v[g * 8 + 1], fmt);
    w.y = pack2<ACT, BF>(v[g * 8 + 2], v[g * 8 + 3], fmt);
    w.z = pack2<ACT, BF>(v[g * 8 + 4], v[g * 8 + 5], fmt);
    w.w = pack2<ACT, BF>(v[g * 8 + 6], v[g * 8 + 7], fmt);
    *reinterpret_cast<uint4*>(tw + lane * 128 + ((g ^ (lane & 7)) << 4)) = w;
  }
}
__device__ __forceinline__ void stage_chunk64(const uint32_t (&v)[64], uint8_t* tw, int lane,
                                              int fmt) {
  if ((fmt >> kFmtActShift) & kFmtActMask) {   // activations: the type test stays per element
    stage_chunk64_t<true, -1>(v, tw, lane, fmt);
  } else if (fmt & kFmtBf16) {
    stage_chunk64_t<false, 1>(v, tw, lane, fmt);
  } else {
    stage_chunk64_t<false, 0>(v, tw, lane, fmt);
  }
}
// Direct epilogue (kFmtDirect): the thread owns one accumulator row; its 64 fp32 columns become
// 128 contiguous bytes of C, stored from registers as four 32 B stores (whole L2 sectors) — no
// shared-memory staging.  Why: the mainloop saturates the shared-memory pipe (UMMA operand reads
// + TMA writes), and an epilogue that round-trips its tile through shared memory crawls behind
// it (tools/trace.py: 4.9 us per 128 x 256 tile under a running mainloop, 1.8 us alone; neither
// the TMEM reads nor the global stores account for it).  `row` = address of the row's first
// column of this chunk, `cols` = valid columns from there (multiple of 8, may be <= 0).
template <bool ACT>
__device__ __forceinline__ void store_row64_direct_t(const uint32_t (&v)[64], uint16_t* row,
                                                     int cols, int fmt) {
  const bool wide = (reinterpret_cast<uintptr_t>(row) & 31) == 0;
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    if (g * 16 >= cols) break;
    uint32_t w[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) w[i] = pack2<ACT>(v[g * 16 + 2 * i], v[g * 16 + 2 * i + 1], fmt);
    if (wide && g * 16 + 16 <= cols) {
      ptx::st_global_v8(row + g * 16, w);
    } else {
      ptx::st_global_v4(row + g * 16, w[0], w[1], w[2], w[3]);
      if (g * 16 + 8 < cols) ptx::st_global_v4(row + g * 16 + 8, w[4], w[5], w[6], w[7]);
    }
  }
}
__device__ __forceinline__ void store_row64_direct(const uint32_t (&v)[64], uint16_t* row, int cols,
                                                   int fmt) {
  if ((fmt >> kFmtActShift) & kFmtActMask) store_row64_direct_t<true>(v, row, cols, fmt);
  else store_row64_direct_t<false>(v, row, cols, fmt);
}

// TMA-store epilogue (kFmtTma): the warp converts its 32 x 64 sub-tile into the XOR-swizzled
// staging tile (exactly the CU_TENSOR_MAP_SWIZZLE_128B layout when the tile is 1024 B aligned)
// and ONE lane hands it to the TMA unit: a bulk tensor store — or, for BSMM_FLAG_ACCUMULATE, a bulk
// tensor reduce-add performed at the L2, so the SM never reads the old C.  Compared with the
// ld.shared -> st.global loop this removes 16 dependent memory instructions per chunk from a warp
// that runs alone on its scheduler (the epilogue is instruction-latency bound: ~230 instructions
// per chunk at an IPC of 0.2, profiles/r02_ncu_c4_dsd_10_source_top.txt), clips ragged edges in
// hardware and runs asynchronously.  The staging tile is reused once the previous store has
// READ it (wait_group.read), which the next chunk's TMEM load + conversion normally covers.
// (c0, c1) = (column, row) of the sub-tile in the output's tensor map (box 64 x 32).
// `flip` (per-thread state, or null): `tw` is a double buffer of 2 x 4 KiB used alternately, so
// the chunk being staged only waits for the store issued TWO chunks ago to have read its source.
// (Measured neutral, profiles/r02_epi_double_buffer_ab.txt: the ~2000 clk a chunk takes under a
// running mainloop are TMEM reads competing with the MMAs, profiles/r02_epilogue_bound.txt.)
__device__ __forceinline__ void tma_chunk64(uint32_t taddr, uint8_t* tw, int lane,
                                            const CUtensorMap* map, int c0, int c1, bool nonempty,
                                            int fmt, int* flip = nullptr) {
  if (flip) {
    tw += *flip * 4096;
    *flip ^= 1;
    if (lane == 0) ptx::tma_store_wait_read<1>();
  } else if (lane == 0) {
    ptx::tma_store_wait_read<0>();
  }
  __syncwarp();
  uint32_t v[64];
  if (nonempty && !BSMM_DBG(2)) {
    ptx::tmem_ld_32x32b_x32(taddr, *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
    ptx::tmem_ld_32x32b_x32(taddr + 32u, *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
    ptx::tmem_ld_wait();
  } else {
#pragma unroll
    for (int i = 0; i < 64; ++i) v[i] = 0u;
  }
  if (!BSMM_DBG(4)) stage_chunk64(v, tw, lane, fmt);
  ptx::fence_proxy_async_smem();
  __syncwarp();
  if (lane == 0 && !BSMM_DBG(1)) {
    if (fmt & kFmtAccum) ptx::tma_reduce_add_2d(map, tw, c0, c1);
    else ptx::tma_store_2d(map, tw, c0, c1);
    ptx::tma_store_commit();
  }
}

__device__ __forceinline__ float silu_mul(uint32_t gate_bits, uint32_t up_bits) {
  const float g = __uint_as_float(gate_bits);
  return __fdividef(g, 1.0f + __expf(-g)) * __uint_as_float(up_bits);
}
// SwiGLU chunk: 64 output columns = silu(gate columns) * (up columns), gate at `taddr`, up at
// `taddr + 128` (the other half of the 128 x 256 accumulator); same staging + bulk store as
// tma_chunk64.  Two 32-column pieces keep the live registers at 64 + 64.
__device__ __forceinline__ void tma_chunk64_swiglu(uint32_t taddr, uint8_t* tw, int lane,
                                                   const CUtensorMap* map, int c0, int c1,
                                                   bool nonempty, int fmt, int* flip) {
  tw += *flip * 4096;
  *flip ^= 1;
  if (lane == 0) ptx::tma_store_wait_read<1>();
  __syncwarp();
  uint32_t v[64];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    uint32_t g[32], u[32];
    if (nonempty) {
      ptx::tmem_ld_32x32b_x32(taddr + uint32_t(h * 32), g);
      ptx::tmem_ld_32x32b_x32(taddr + uint32_t(128 + h * 32), u);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) v[h * 32 + i] = __float_as_uint(silu_mul(g[i], u[i]));
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[h * 32 + i] = 0u;
    }
  }
  stage_chunk64(v, tw, lane, fmt & ~(kFmtActMask << kFmtActShift));
  ptx::fence_proxy_async_smem();
  __syncwarp();
  if (lane == 0) {
    ptx::tma_store_2d(map, tw, c0, c1);
    ptx::tma_store_commit();
  }
}

// 32 fp32 columns of this thread's row -> 16-bit, transposed into a [32][32] staging tile.
template <bool ACT>
__device__ __forceinline__ void stage_chunk32_tr_t(const uint32_t (&v)[32], uint16_t* tw, int lane,
                                                   int fmt) {
#pragma unroll
  for (int jj = 0; jj < 16; ++jj) {
    const uint32_t w = pack2<ACT>(v[2 * jj], v[2 * jj + 1], fmt);
    tw[(2 * jj) * 32 + lane] = uint16_t(w & 0xffffu);
    tw[(2 * jj + 1) * 32 + lane] = uint16_t(w >> 16);
  }
}
__device__ __forceinline__ void stage_chunk32_tr(const uint32_t (&v)[32], uint16_t* tw, int lane,
                                                 int fmt) {
  if ((fmt >> kFmtActShift) & kFmtActMask) stage_chunk32_tr_t<true>(v, tw, lane, fmt);
  else stage_chunk32_tr_t<false>(v, tw, lane, fmt);
}

// Transposed output (DDS as transposed DSD): the warp's 32 x 32 transposed tile [dense row][sparse
// column], 64 B per row, unswizzled, stored / reduce-added by TMA (box 32 x 32).
__device__ __forceinline__ void tma_chunk32_tr(uint32_t taddr, uint8_t* tw, int lane,
                                               const CUtensorMap* map, int c0, int c1,
                                               bool nonempty, int fmt, int* flip = nullptr) {
  if (flip) {   // two 2 KiB tiles used alternately (see tma_chunk64)
    tw += *flip * 2048;
    *flip ^= 1;
    if (lane == 0) ptx::tma_store_wait_read<1>();
  } else if (lane == 0) {
    ptx::tma_store_wait_read<0>();
  }
  __syncwarp();
  uint32_t v[32];
  if (nonempty) {
    ptx::tmem_ld_32x32b_x32(taddr, v);
    ptx::tmem_ld_wait();
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = 0u;
  }
  stage_chunk32_tr(v, reinterpret_cast<uint16_t*>(tw), lane, fmt);
  ptx::fence_proxy_async_smem();
  __syncwarp();
  if (lane == 0) {
    if (fmt & kFmtAccum) ptx::tma_reduce_add_2d(map, tw, c0, c1);
    else ptx::tma_store_2d(map, tw, c0, c1);
    ptx::tma_store_commit();
  }
}
// Every lane that may have issued bulk stores waits for them before its CTA exits.
__device__ __forceinline__ void tma_epilogue_done(int lane) {
  if (lane == 0) ptx::tma_store_wait_all<0>();
}

// N-side (dense / second operand) tile load: `rows` N-rows starting at n0, 64 k starting at k0.
// K-major: boxes [128 rows][64 k] (the map's box), 16 KiB each; MN-major: 64-column slabs of
// [64 k][64 n], 8 KiB each.
// `bk` = elements per 128 B (64 / 128): an MN-major slab is [bk k][bk n] = bk * 128 bytes.
// BK is a template parameter: the single producer thread is the critical path of short tiles and
// a run-time slab count (a division and a rolled loop) cost the 16-bit kernels 5-25 %
// (profiles/r02_regression_ab.txt).
template <bool MN, int ROWS, int BK = kBK>
__device__ __forceinline__ void load_nside(const CUtensorMap* map, uint64_t* bar, uint8_t* dst,
                                           int n0, int k0) {
  if constexpr (MN) {
#pragma unroll
    for (int h = 0; h < ROWS / BK; ++h)
      ptx::tma_load_2d(map, bar, dst + h * BK * 128, n0 + h * BK, k0);
  } else {
#pragma unroll
    for (int h = 0; h < (ROWS + 127) / 128; ++h)   // boxes of (at most) 128 rows, 16 KiB apart
      ptx::tma_load_2d(map, bar, dst + h * 16384, k0, n0 + h * 128);
  }
}

// M-side tile load: 128 rows starting at m0, 64 k starting at k0.
template <bool MN, int BK = kBK>
__device__ __forceinline__ void load_mside(const CUtensorMap* map, uint64_t* bar, uint8_t* dst,
                                           int m0, int k0) {
  if constexpr (MN) {
    ptx::tma_load_2d(map, bar, dst, m0, k0);
    if constexpr (BK == kBK) ptx::tma_load_2d(map, bar, dst + 8192, m0 + 64, k0);
  } else {
    ptx::tma_load_2d(map, bar, dst, k0, m0);
  }
}

// L2 prefetch twins of load_nside / load_mside (same boxes, no destination).
template <bool MN, int ROWS>
__device__ __forceinline__ void prefetch_nside(const CUtensorMap* map, int n0, int k0,
                                               int bk = kBK) {
  if constexpr (MN) {
#pragma unroll 4
    for (int h = 0; h < ROWS / bk; ++h) ptx::tma_prefetch_2d(map, n0 + h * bk, k0);
  } else {
#pragma unroll
    for (int h = 0; h < (ROWS + 127) / 128; ++h) ptx::tma_prefetch_2d(map, k0, n0 + h * 128);
  }
}
template <bool MN>
__device__ __forceinline__ void prefetch_mside(const CUtensorMap* map, int m0, int k0,
                                               int bk = kBK) {
  if constexpr (MN) {
    ptx::tma_prefetch_2d(map, m0, k0);
    if (bk == kBK) ptx::tma_prefetch_2d(map, m0 + 64, k0);
  } else {
    ptx::tma_prefetch_2d(map, k0, m0);
  }
}
// Blocks (DSD walks) / k-steps (SDD) of a CTA's FIRST tile whose operand boxes are prefetched into
// L2 before griddepcontrol.wait, i.e. while the previous grid in the stream is still draining.
constexpr int kPrefetchBlocks = 3;
constexpr int kPrefetchSteps = 6;

// OP = kOpSDD: A_MN / B_MN describe op(A) (M side) and op(B) (N side, two 128-wide panels).
// OP = kOpDSD: A_MN describes the sparse block (M side), B_MN the dense operand (N side).
// TRANS_OUT (DSD only): the accumulator holds C^T (lanes = index along the sparse dimension,
// columns = index along the dense dimension) and is stored transposed — this is DDS.
// CL (DSD only): thread-block cluster size along the dense dimension.  The CL CTAs of a cluster
// work on CL adjacent dense-side tiles of the SAME block row / column, so they need the same
// sparse block at every step: each CTA fetches 1/CL of it and TMA-multicasts the slice to the
// whole cluster (tools/tma_bw.cu: the L2 -> SM feed, not the tensor pipe, is what bounds a
// 128x256x64 stage when every CTA pulls its own 48 KiB).  Stage release is multicast too, which
// keeps the cluster in lockstep.
template <int OP, bool A_MN, bool B_MN, int BN, bool TRANS_OUT, int CL>
__global__ void __launch_bounds__(kThreads, 1)
bsmm_gemm_kernel(const __grid_constant__ CUtensorMap map_a,
                 const __grid_constant__ CUtensorMap map_b,
                 const __grid_constant__ CUtensorMap map_c, const GemmParams p) {
  using T = Tiling<BN>;
  constexpr int S = T::kStages;
  static_assert(OP != kOpSDD || BN == 256, "SDD tiles are block pairs");
  static_assert(CL == 1 || OP == kOpDSD, "clusters share the sparse block of a DSD walk");
  static_assert(CL == 1 || CL == 2 || CL == 4, "cluster size");

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
  int* s_blk = s_idx + 64;                            // [2][32] block id in the value array
  uint64_t* tq_full = reinterpret_cast<uint64_t*>(s_blk + 64);  // tile-id queue
  uint64_t* tq_empty = tq_full + kTileQ;
  int* s_tile = reinterpret_cast<int*>(tq_empty + kTileQ);
  // tail drain: epilogue warp 4 posts {last tile, accumulator stage, its phase} for warps 0-3
  uint64_t* help_bar = reinterpret_cast<uint64_t*>(s_tile + kTileQ);
  int* s_last = reinterpret_cast<int*>(help_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (threadIdx.x == 0) BSMM_STAMP(0);
  // Tile walk: CL == 1: tile t -> CTA t mod grid.  CL > 1: tile GROUP (major, CL adjacent minors)
  // -> cluster; this CTA takes minor = group * CL + rank.  (tiles_minor is a multiple of CL.)
  const int crank = CL > 1 ? int(ptx::cluster_ctarank()) : 0;
  int num_tiles = p.tiles_major * (p.tiles_minor / CL);
  int nb = p.nb;
  const int t_first = CL > 1 ? int(ptx::cluster_id_x()) : int(blockIdx.x);
  const int t_step = CL > 1 ? int(ptx::num_clusters_x()) : int(gridDim.x);
  const int groups_minor = p.tiles_minor / CL;
  constexpr uint16_t kMask = uint16_t((1u << CL) - 1u);
  // DSD tile id -> (walk row, first dense column, width in columns).  With a tail-wave split
  // (bsmm_api.cu) ids >= p.split_from are half tiles: full tile split_from + u / 2, half u & 1.
  constexpr bool kSplit = OP == kOpDSD && CL == 1 && BN == 256;
  if constexpr (kSplit) {
    if (p.split_from > 0) num_tiles += num_tiles - p.split_from;
  }
  auto dsd_tile = [&](int t, int& major, int& col0, int& width) {
    int f = t, half = 0;
    width = BN;
    if constexpr (kSplit) {
      if (p.split_from > 0 && t >= p.split_from) {
        const int u = t - p.split_from;
        f = p.split_from + (u >> 1);
        half = u & 1;
        width = BN / 2;
      }
    }
    major = f / groups_minor;
    col0 = ((f - major * groups_minor) * CL + crank) * BN + half * (BN / 2);
  };
  (void)dsd_tile;  // SDD instantiations do not use it

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&map_a);
    ptx::prefetch_tensormap(&map_b);
    ptx::prefetch_tensormap(&map_c);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < S; ++i) {
      ptx::mbar_init(&full_bar[i], 1);
      ptx::mbar_init(&empty_bar[i], CL);  // released by the MMA warp of every CTA in the cluster
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&tfull_bar[i], 1);
      ptx::mbar_init(&tempty_bar[i], 4);  // one arrive per epilogue warp
    }
    for (int i = 0; i < kTileQ; ++i) {
      ptx::mbar_init(&tq_full[i], 1);
      ptx::mbar_init(&tq_empty[i], 5);  // MMA thread + 4 epilogue warps
    }
    ptx::mbar_init(help_bar, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc(tmem_ptr, T::kTmemCols);
  ptx::tc_fence_before_sync();
  if constexpr (CL > 1) ptx::cluster_sync();  // peers' barriers must be initialised too
  else __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_ptr;
  // Everything above touched only parameters and shared memory: with programmatic dependent
  // launch it overlaps the tail of the previous kernel in the stream.  So does this: the
  // otherwise idle warp 3 (NOT the producer, whose own path to the first load must not grow
  // when there is nothing to overlap with) peeks at the metadata of the CTA's first tile (L2-only loads; the values are
  // used for nothing but prefetch coordinates and everything is read again after the wait) and
  // pulls the first operand boxes into L2, so that the first pipeline stages fill at L2 latency
  // instead of behind three dependent HBM round trips (offsets -> indices -> data: 3.3 us).
  if (CL == 1 && warp == 2 && p.prefetch) {
    // whole-operand warm-up: this CTA's slice of each operand, 8 KiB per request, 32 lanes
    constexpr unsigned long long kChunk = 8192;
#pragma unroll
    for (int o = 0; o < 2; ++o) {
      const unsigned long long total = p.warm_bytes[o];
      if (total == 0) continue;
      const unsigned long long per = ((total + gridDim.x - 1) / gridDim.x + kChunk - 1) / kChunk * kChunk;
      const unsigned long long lo = per * blockIdx.x;
      const unsigned long long hi = lo + per < total ? lo + per : total;
      for (unsigned long long off = lo + kChunk * lane; off < hi; off += kChunk * 32) {
        const unsigned long long len = hi - off < kChunk ? hi - off : kChunk;
        ptx::bulk_prefetch_l2(static_cast<const uint8_t*>(p.warm_ptr[o]) + off, uint32_t(len & ~15ull));
      }
    }
  }
  if (CL == 1 && warp == 3 && p.prefetch && t_first < num_tiles && !p.nnz_dev) {
    if constexpr (OP == kOpSDD) {
      const int r0 = (ptx::ld_cg_s16(p.sdd_rows + t_first) & 0x7fff) * kBlock;
      const int c0 = (ptx::ld_cg_s16(p.sdd_cols + t_first) & 0x7fff) * kBlock;
      const int c1 = t_first + 1 < nb ? (ptx::ld_cg_s16(p.sdd_cols + t_first + 1) & 0x7fff) * kBlock : c0;
      if (lane < min(p.k_steps, kPrefetchSteps)) {
        prefetch_mside<A_MN>(&map_a, r0, lane * p.bk, p.bk);
        prefetch_nside<B_MN, 128>(&map_b, c0, lane * p.bk, p.bk);
        if (!p.sdd_single) prefetch_nside<B_MN, 128>(&map_b, c1, lane * p.bk, p.bk);
      }
    } else {
      const int major = t_first / groups_minor;
      const int d0 = (t_first - major * groups_minor) * BN;
      const int j0 = ptx::ld_cg_s32(p.walk_offsets + major);
      const int j1 = ptx::ld_cg_s32(p.walk_offsets + major + 1);
      const int half = lane & 1, i = lane >> 1;   // lane = (block of the walk, k half)
      if (j0 >= 0 && i < kPrefetchBlocks && j0 + i < j1 && j1 <= (1 << 17) &&
          half * p.bk < kBlock) {
        const int idx = ptx::ld_cg_s16(p.walk_indices + j0 + i) & 0x7fff;
        const int blk = p.walk_blocks ? (ptx::ld_cg_s32(p.walk_blocks + j0 + i) & 0x1ffff) : j0 + i;
        if constexpr (A_MN) prefetch_mside<true>(&map_a, 0, blk * kBlock + half * p.bk, p.bk);
        else                prefetch_mside<false>(&map_a, blk * kBlock, half * p.bk, p.bk);
        prefetch_nside<B_MN, BN>(&map_b, d0, idx * kBlock + half * p.bk, p.bk);
      }
    }
  }
  ptx::grid_dep_wait();
  ptx::grid_dep_launch();
  if (threadIdx.x == 0) BSMM_STAMP(1);
  if constexpr (OP == kOpSDD) {
    if (p.nnz_dev) {  // device-sized topology: blocks past offsets[block_rows] do not exist
      nb = min(nb, __ldg(p.nnz_dev));
      num_tiles = nb;
    }
  }

  // Stores (part 0) a whole accumulator tile, or its first 128 columns (1) / the rest (2), from
  // accumulator stage `das`, for the TMEM lane quarter `dq`, staging through `tw_base` (4 KiB).
  uint16_t* const cbase = reinterpret_cast<uint16_t*>(p.c);
  int epi_flip = 0;  // which half of this warp's double staging buffer the next TMA chunk takes
  auto drain_tile = [&](int t, int das, int dq, uint8_t* tw_base, int part) {
      int nsteps, ncc = BN / 32;
      int major = 0, col0 = 0;   // DSD: walk row and first dense column of the tile
      if constexpr (OP == kOpSDD) {
        nsteps = p.k_steps;
        if (p.swiglu) {
          // one output block per tile: columns [0, 128) of the accumulator gate columns
          // [128, 256); stored whole by the epilogue warps (the tail-drain helpers, part 2, have
          // nothing to add)
          if (part == 2) return;
          const uint32_t ta_ = tmem_base + (uint32_t(dq * 32) << 16) + uint32_t(das * BN);
#pragma unroll 1
          for (int ch = 0; ch < 2; ++ch)
            tma_chunk64_swiglu(ta_ + uint32_t(ch * 64), tw_base, lane, &map_c, ch * 64,
                               t * kBlock + dq * 32, nsteps > 0, p.bf16, &epi_flip);
          return;
        }
        const bool paired = !p.sdd_single && !(t & 1) && t + 1 < nb &&
                            __ldg(p.sdd_rows + t + 1) == __ldg(p.sdd_rows + t);
        if (!paired) ncc = 4;  // a single block
      } else {
        int width;
        dsd_tile(t, major, col0, width);
        nsteps = 2 * (__ldg(p.walk_offsets + major + 1) - __ldg(p.walk_offsets + major));
        ncc = max(0, min(width / 32, (p.dense_lim - col0 + 31) / 32));
      }
      const int cc0 = part == 2 ? min(ncc, 4) : 0;   // in 32-column units
      const int cc1 = part == 1 ? min(ncc, 4) : ncc;
      const uint32_t taddr = tmem_base + (uint32_t(dq * 32) << 16) + uint32_t(das * BN);
      const bool use_tma = (p.bf16 & kFmtTma) && (OP == kOpSDD || p.c_rows == nullptr);
      if constexpr (OP == kOpSDD || !TRANS_OUT) {
        // Row-major store.  The thread owns accumulator row `lrow`; 64 columns (128 B) at a time
        // go through a per-warp XOR-swizzled staging tile so that global stores are whole
        // 128 B lines (8 lanes x 16 B per row, 4 rows per instruction).
        uint8_t* tw = tw_base;
        const int nch = (cc1 + 1) >> 1;
#pragma unroll 1
        for (int ch = (cc0 + 1) >> 1; ch < nch; ++ch) {  // (odd cc0: that chunk belongs to part 1)
          if (use_tma) {
            if constexpr (OP == kOpSDD)  // output map: [nb * 128][128]
              tma_chunk64(taddr + uint32_t(ch * 64), tw, lane, &map_c, (ch & 1) * 64,
                          (t + (ch >> 1)) * kBlock + dq * 32, nsteps > 0, p.bf16, &epi_flip);
            else                         // output map: [sparse extent][dense extent]
              tma_chunk64(taddr + uint32_t(ch * 64), tw, lane, &map_c, col0 + ch * 64,
                          major * kBlock + dq * 32, nsteps > 0, p.bf16, &epi_flip);
            continue;
          }
          if ((p.bf16 & (kFmtDirect | kFmtAccum)) == kFmtDirect) {
            // direct epilogue: this thread stores its own row from registers
            uint16_t* rp;
            int cols = 64;
            if constexpr (OP == kOpSDD) {
              rp = cbase + (size_t(t) + size_t(ch >> 1)) * (kBlock * kBlock) +
                   size_t(dq * 32 + lane) * kBlock + (ch & 1) * 64;
            } else {
              const int c64 = col0 + ch * 64;
              cols = min(64, p.dense_lim - c64);
              const int row = major * kBlock + dq * 32 + lane;
              if (p.c_rows) {
                uint16_t* base = static_cast<uint16_t*>(ptx::ldg_ptr(p.c_rows + row));
                if (!base) cols = 0;
                rp = base + c64;
              } else {
                rp = cbase + size_t(row) * size_t(p.ldc) + c64;
              }
            }
            uint32_t v[64];
            if (nsteps > 0) {
              ptx::tmem_ld_32x32b_x32(taddr + uint32_t(ch * 64), *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
              ptx::tmem_ld_32x32b_x32(taddr + uint32_t(ch * 64 + 32), *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
              ptx::tmem_ld_wait();
            } else {
#pragma unroll
              for (int i = 0; i < 64; ++i) v[i] = 0u;
            }
            store_row64_direct(*reinterpret_cast<const uint32_t(*)[64]>(&v[0]), rp, cols, p.bf16);
            continue;
          }
          const int seg = lane & 7;
          uint16_t* o = nullptr;
          long long ld = 0;
          bool col_ok = true;
          int col = 0;
          if constexpr (OP == kOpSDD) {
            // columns [0,128) -> block t, [128,256) -> block t+1; blocks are contiguous in C.
            o = cbase + (size_t(t) + size_t(ch >> 1)) * (kBlock * kBlock) +
                size_t(dq * 32) * kBlock + (ch & 1) * 64 + seg * 8;
            ld = kBlock;
          } else {
            col = col0 + ch * 64 + seg * 8;
            col_ok = col < p.dense_lim;  // extents are multiples of 8
            o = cbase + size_t(major * kBlock + dq * 32) * size_t(p.ldc) + col;
            ld = p.ldc;
          }
          const bool scatter = OP != kOpSDD && p.c_rows != nullptr;
          const bool accum = (p.bf16 & kFmtAccum) && !scatter && col_ok;
          uint16_t* rp8[8];   // scattered rows: addresses requested before the TMEM read
          if (scatter) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              rp8[i] = static_cast<uint16_t*>(
                  ptx::ldg_ptr(p.c_rows + major * kBlock + dq * 32 + (lane >> 3) + 4 * i));
          }
          uint4 old[8];
          if (accum) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              old[i] = *reinterpret_cast<const uint4*>(o + size_t((lane >> 3) + 4 * i) * size_t(ld));
          }
          uint32_t v[64];
          if (nsteps > 0 && !BSMM_DBG(2)) {
            ptx::tmem_ld_32x32b_x32(taddr + uint32_t(ch * 64), *reinterpret_cast<uint32_t(*)[32]>(&v[0]));
            ptx::tmem_ld_32x32b_x32(taddr + uint32_t(ch * 64 + 32), *reinterpret_cast<uint32_t(*)[32]>(&v[32]));
            ptx::tmem_ld_wait();
          } else {
#pragma unroll
            for (int i = 0; i < 64; ++i) v[i] = 0u;
          }
          stage_chunk64(*reinterpret_cast<const uint32_t(*)[64]>(&v[0]), tw, lane, p.bf16);
          __syncwarp();
          if (accum) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int r = (lane >> 3) + 4 * i;
              const uint4 w = *reinterpret_cast<const uint4*>(tw + r * 128 + ((seg ^ (r & 7)) << 4));
              *reinterpret_cast<uint4*>(o + size_t(r) * size_t(ld)) = add4(w, old[i], p.bf16);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int r = (lane >> 3) + 4 * i;
              const uint4 w = *reinterpret_cast<const uint4*>(tw + r * 128 + ((seg ^ (r & 7)) << 4));
              if (scatter) {
                if (col_ok && rp8[i]) *reinterpret_cast<uint4*>(rp8[i] + col) = w;
              } else if (col_ok && !BSMM_DBG(1)) {
                *reinterpret_cast<uint4*>(o + size_t(r) * size_t(ld)) = w;
              }
            }
          }
          __syncwarp();
        }
      } else {
#pragma unroll 1
        for (int cc = cc0; cc < cc1; ++cc) {
          // C[dense index][sparse index].  The thread holds 32 dense-side values of ONE
          // sparse-side column; a per-warp 32 x 32 transpose through shared memory turns that
          // into 64 B runs along the sparse dimension, stored 16 B per lane (8 rows / instr).
          const int drow0 = col0 + cc * 32;
          if (use_tma) {  // output map: [dense extent][sparse extent], box 32 x 32
            tma_chunk32_tr(taddr + uint32_t(cc * 32), tw_base, lane, &map_c,
                           major * kBlock + dq * 32, drow0, nsteps > 0, p.bf16, &epi_flip);
            continue;
          }
          const int seg = lane & 3;
          uint16_t* o = cbase + size_t(major * kBlock + dq * 32 + seg * 8);
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
          if (nsteps > 0) {
            ptx::tmem_ld_32x32b_x32(taddr + uint32_t(cc * 32), v);
            ptx::tmem_ld_wait();
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = 0u;
          }
          uint16_t* tw = reinterpret_cast<uint16_t*>(tw_base);
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
  };

  if (warp == 0) {
    // ===================================================================== TMA producer
    int stage = 0;
    uint32_t phase = 0;
    int buf = 0;
    // Tile ids go to the other roles through a small shared-memory queue.  Dynamic mode: drawn
    // from a global counter in launch order, so CTAs that finish early take more tiles (rows of
    // very different nnz) while concurrently running CTAs still work on neighbouring tiles.
    int tq = 0, t_static = t_first;
    uint32_t tqphase = 0;
    // The first tile of a CTA is its block index (no global round trip before the first load);
    // the counter then hands out tiles gridDim.x .. num_tiles - 1 and one sentinel per CTA:
    // num_tiles draws in all, so it wraps to 0 with the last one (atomicInc limit).  Not with a
    // device-sized topology, where num_tiles may turn out smaller than the grid.
    const bool static_first = p.tile_counter != nullptr && p.nnz_dev == nullptr;
    const unsigned draw_limit =
        static_first ? unsigned(num_tiles) - 1u : unsigned(num_tiles) + gridDim.x - 1u;
    bool first_draw = static_first;
    for (;;) {
      int t = 0;
      if (lane == 0) {
        ptx::mbar_wait(&tq_empty[tq], tqphase ^ 1);
        for (;;) {
          if (first_draw) {
            t = int(blockIdx.x);
            first_draw = false;
          } else if (p.tile_counter) {
            t = int(atomicInc(p.tile_counter, draw_limit)) + (static_first ? int(gridDim.x) : 0);
          } else {
            t = t_static < num_tiles ? t_static : num_tiles;
            t_static += t_step;
          }
          if constexpr (OP == kOpSDD) {
            // SDD tile ids are block ids.  An odd block in the same block row as its even
            // neighbour belongs to that neighbour's pair tile: draw again.
            if (t < num_tiles && !p.sdd_single && !p.swiglu && (t & 1) &&
                __ldg(p.sdd_rows + t - 1) == __ldg(p.sdd_rows + t))
              continue;
          }
          break;
        }
        s_tile[tq] = t;
        ptx::mbar_arrive(&tq_full[tq]);
      }
      t = __shfl_sync(0xffffffffu, t, 0);
      if (++tq == kTileQ) { tq = 0; tqphase ^= 1; }
      if (t >= num_tiles) break;
      if constexpr (OP == kOpSDD) {
        // Tile coordinates are computed by the whole (converged) warp; only the barrier / TMA
        // issue is predicated on lane 0.
        const int r0 = int(__ldg(p.sdd_rows + t)) * kBlock;
        int c0 = int(__ldg(p.sdd_cols + t)) * kBlock;
        bool paired = false;
        int c1 = 0;
        if (p.swiglu) {   // gate panel at block column 2j, up panel at 2j + 1 of op(B)
          c0 *= 2;
          c1 = c0 + kBlock;
          paired = true;
        } else if (!p.sdd_single && !(t & 1) && t + 1 < nb &&
                   __ldg(p.sdd_rows + t + 1) == __ldg(p.sdd_rows + t)) {
          paired = true;
          c1 = int(__ldg(p.sdd_cols + t + 1)) * kBlock;
        }
        const uint32_t bytes = T::kABytes + (paired ? T::kBBytes : T::kBBytes / 2);
        __syncwarp();
        if (lane == 0) {
          if (t == t_first) BSMM_STAMP(2);
          // specialised on the operand width (compile-time slab count, see load_nside)
          auto k_loop = [&](auto bkc) {
            constexpr int bk = decltype(bkc)::value;
            for (int s = 0; s < p.k_steps; ++s) {
              const int k0 = s * bk;
              ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
              uint8_t* sa = smem + stage * T::kStageBytes;
              uint8_t* sb = sa + T::kABytes;
              ptx::mbar_arrive_expect_tx(&full_bar[stage], bytes);
              load_mside<A_MN, bk>(&map_a, &full_bar[stage], sa, r0, k0);
              load_nside<B_MN, 128, bk>(&map_b, &full_bar[stage], sb, c0, k0);
              if (paired) load_nside<B_MN, 128, bk>(&map_b, &full_bar[stage], sb + 16384, c1, k0);
              if (++stage == S) { stage = 0; phase ^= 1; }
            }
          };
          if (p.f8) k_loop(std::integral_constant<int, 2 * kBK>{});
          else      k_loop(std::integral_constant<int, kBK>{});
        }
        __syncwarp();
      } else {
        int major, d0, width;   // d0: dense-side tile origin
        dsd_tile(t, major, d0, width);
        const int j0 = __ldg(p.walk_offsets + major);
        const int j1 = __ldg(p.walk_offsets + major + 1);
        const uint32_t stage_tx = width == BN ? T::kStageBytes : T::kStageBytes - T::kBBytes / 2;
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
            auto blk_loop = [&](auto bkc) {
            constexpr int bk = decltype(bkc)::value;
            for (int i = 0; i < cnt; ++i) {
              const int blk_row0 = s_blk[buf * 32 + i] * kBlock;
              const int dense_k0 = s_idx[buf * 32 + i] * kBlock;
#pragma unroll
              for (int kh = 0; kh < kBlock / bk; ++kh) {   // 2 stages per block (1: e4m3)
                ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                uint8_t* sa = smem + stage * T::kStageBytes;
                uint8_t* sb = sa + T::kABytes;
                ptx::mbar_arrive_expect_tx(&full_bar[stage], stage_tx);
                // Sparse block: K-major reads it as stored; MN-major reads it transposed
                // (64-row slabs of the stored block) — the index-indirected gather.
                if constexpr (CL == 1) {
                  if constexpr (A_MN) load_mside<true, bk>(&map_a, &full_bar[stage], sa, 0, blk_row0 + kh * bk);
                  else                load_mside<false, bk>(&map_a, &full_bar[stage], sa, blk_row0, kh * bk);
                } else if constexpr (!A_MN) {
                  // slice = 128 / CL rows of the [128 rows][64 k] tile
                  constexpr int R = kBlock / CL;
                  ptx::tma_load_2d_mcast(&map_a, &full_bar[stage], sa + crank * R * 128, kh * kBK,
                                         blk_row0 + crank * R, kMask);
                } else {
                  // tile = two slabs [64 k][64 m]; slice = slab (CL = 2) or half a slab (CL = 4)
                  constexpr int R = 2 * kBK / CL;  // k rows per slice
                  const int slab = (CL == 2) ? crank : (crank >> 1);
                  const int sub = (CL == 2) ? 0 : (crank & 1);
                  ptx::tma_load_2d_mcast(&map_a, &full_bar[stage], sa + slab * 8192 + sub * R * 128,
                                         slab * 64, blk_row0 + kh * kBK + sub * R, kMask);
                }
                // Dense operand, gathered at the block's coordinate.
                if (kSplit && width != BN)
                  load_nside<B_MN, 128, bk>(&map_b, &full_bar[stage], sb, d0, dense_k0 + kh * bk);
                else
                  load_nside<B_MN, BN, bk>(&map_b, &full_bar[stage], sb, d0, dense_k0 + kh * bk);
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
      }
    }
  } else if (warp == 1) {
    // ===================================================================== MMA issuer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      // per-MMA advance of the descriptors' start address (>> 4): 32 B along K for K-major
      // operands, 16 (32: e4m3) k-rows of 128 B for MN-major ones; LBO of MN-major = one slab
      int tq = 0;
      uint32_t tqphase = 0;
      for (;;) {
        ptx::mbar_wait(&tq_full[tq], tqphase);
        const int t = s_tile[tq];
        ptx::mbar_arrive(&tq_empty[tq]);
        if (++tq == kTileQ) { tq = 0; tqphase ^= 1; }
        if (t >= num_tiles) break;
        int nsteps;
        uint32_t idesc = p.idesc;
        if constexpr (OP == kOpSDD) {
          nsteps = p.k_steps;
          const bool paired = p.swiglu || (!p.sdd_single && !(t & 1) && t + 1 < nb &&
                                           __ldg(p.sdd_rows + t + 1) == __ldg(p.sdd_rows + t));
          if (!paired) idesc = p.idesc_half;  // a single block: N = 128
        } else {
          int major, col0_unused, width;
          dsd_tile(t, major, col0_unused, width);
          if (width != BN) idesc = p.idesc_half;  // tail-wave half tile: N = 128
          nsteps = (p.f8 ? 1 : 2) * (__ldg(p.walk_offsets + major + 1) - __ldg(p.walk_offsets + major));
        }
        ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);
        ptx::tc_fence_after_sync();
        if (nsteps == 0) {
          ptx::mbar_arrive(&tfull_bar[as]);  // nothing accumulated: epilogue writes zeros
        }
        const uint32_t tmem_d = tmem_base + uint32_t(as * BN);
        // specialised on the operand width: constant descriptor advances, no per-step branch
        auto mma_loop = [&](auto f8c) {
          constexpr bool F8 = decltype(f8c)::value;
          constexpr int bk = F8 ? 2 * kBK : kBK;
          constexpr uint32_t kAdvA = A_MN ? uint32_t(bk * 2) : (32 >> 4);
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
                ptx::umma_f8(tmem_d, da + uint64_t(k * kAdvA), db + uint64_t(k * kAdvB), idesc,
                             uint32_t((s | k) != 0));
              else
                ptx::umma_f16(tmem_d, da + uint64_t(k * kAdvA), db + uint64_t(k * kAdvB), idesc,
                              uint32_t((s | k) != 0));
            }
            if constexpr (CL > 1) ptx::umma_commit_mcast(&empty_bar[stage], kMask);
            else ptx::umma_commit(&empty_bar[stage]);
            if (s == nsteps - 1) ptx::umma_commit(&tfull_bar[as]);
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
    const int q = warp & 3;  // TMEM lane quarter this warp may read
    int as = 0;
    uint32_t aphase = 0;
    int tq = 0;
    uint32_t tqphase = 0;
    auto pop = [&]() {
      ptx::mbar_wait(&tq_full[tq], tqphase);
      const int id = s_tile[tq];
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tq_empty[tq]);
      if (++tq == kTileQ) { tq = 0; tqphase ^= 1; }
      return id;
    };
    // The next id is popped before the current tile is drained: a sentinel there marks the
    // current tile as the CTA's last one, which warps 0-3 help to store (tail drain, below).
    int cur = pop();
    if (CL == 1 && cur >= num_tiles && warp == 4 && lane == 0) {
      s_last[0] = -1;  // this CTA drew no tile at all
      ptx::mbar_arrive(help_bar);
    }
    while (cur < num_tiles) {
      const int t = cur;
      const int nxt = pop();
      const bool last = CL == 1 && nxt >= num_tiles;
      if (last && warp == 4 && lane == 0) {
        s_last[0] = t; s_last[1] = as; s_last[2] = int(aphase);
        ptx::mbar_arrive(help_bar);
      }
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after_sync();
      if (warp == 4 && lane == 0 && t == t_first) BSMM_STAMP(6);
      drain_tile(t, as, q, epi_smem + q * T::kEpiWarpBytes, last ? 1 : 0);
      ptx::tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty_bar[as]);
      if (++as == 2) { as = 0; aphase ^= 1; }
      if (warp == 4 && lane == 0) BSMM_STAMP(t == t_first ? 7 : 8);
      cur = nxt;
    }
    tma_epilogue_done(lane);
  }
  if constexpr (CL == 1) {
    if (warp < 4) {
      // Tail drain: nothing overlaps the epilogue of the CTA's last tile, so the producer / MMA /
      // allocator / spare warps store its columns 128.. (warp w reads the TMEM lane quarter of
      // epilogue warp 4 + w; the first pipeline stage is free by now and serves as staging).
      __syncwarp();
      ptx::mbar_wait(help_bar, 0);
      const int t = s_last[0];
      if (t >= 0) {
        const int las = s_last[1];
        ptx::mbar_wait(&tfull_bar[las], uint32_t(s_last[2]));
        ptx::tc_fence_after_sync();
        drain_tile(t, las, warp, smem + warp * T::kEpiWarpBytes, 2);
        ptx::tc_fence_before_sync();
        tma_epilogue_done(lane);
      }
    }
  }

  __syncwarp();
  ptx::tc_fence_before_sync();
  if constexpr (CL > 1) ptx::cluster_sync();  // no CTA may exit while peers still signal it
  else __syncthreads();
  if (threadIdx.x == 0) BSMM_STAMP(9);
  if (warp == 2) {
    ptx::tc_fence_after_sync();
    ptx::tmem_dealloc(tmem_base, T::kTmemCols);
  }
}

}  // namespace bsmm
