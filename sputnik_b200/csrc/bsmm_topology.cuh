// Device-side topology kernels: row_indices and the transposed-iteration metadata.
//
// Replaces sputnik/block/row_indices/row_indices.cu:7-36 (device kernel, one CTA per block
// row) and sputnik/block/transpose/transpose.cu:69-125 (HOST code: blocking D2H copies,
// std::stable_sort argsort, gathers, histogram + cumsum, H2D copies).  Results are
// bit-identical to the reference: blocks ordered by (block column, position in `data`),
// which is exactly what a stable sort of the column indices produces.
#pragma once

#include <climits>
#include <cstdint>
#include <cuda_runtime.h>

namespace bsmm {

// Largest r with offsets[r] <= j, i.e. the block row that owns block j (empty rows skipped).
__device__ __forceinline__ int row_of_block(const int* __restrict__ offsets, int block_rows,
                                            int j) {
  int lo = 0, hi = block_rows;  // invariant: offsets[lo] <= j < offsets[hi]
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(offsets + mid) <= j) lo = mid; else hi = mid;
  }
  return lo;
}

// One thread per nonzero block (coalesced int16 stores), instead of one CTA per block row.
__global__ void row_indices_kernel(const int* __restrict__ offsets, int block_rows, int nb,
                                   short* __restrict__ row_indices) {
  for (int j = blockIdx.x * blockDim.x + threadIdx.x; j < nb; j += gridDim.x * blockDim.x)
    row_indices[j] = short(row_of_block(offsets, block_rows, j));
}

// bsmm_dsd_scatter on a matrix without blocks: zeros to every non-null row (the GEMM kernels
// need at least one block for their tensor maps).
__global__ void scatter_zero_rows_kernel(void* const* rows, int nrows, int cols8) {
  const long long total = 1ll * nrows * cols8;
  for (long long i = blockIdx.x * 1ll * blockDim.x + threadIdx.x; i < total;
       i += 1ll * gridDim.x * blockDim.x) {
    const int r = int(i / cols8), c = int(i - 1ll * r * cols8);
    uint4* rp = reinterpret_cast<uint4*>(__ldg(reinterpret_cast<const unsigned long long*>(rows + r)));
    if (rp) rp[c] = make_uint4(0u, 0u, 0u, 0u);
  }
}

// bsmm_gather_rows: dst[r][:] = *src_rows[r] (row_bytes each; a null source gives a zero row).
// One warp per row, 4 x 16 B per lane in flight before the first store: the sources may be
// peer-GPU memory (NVLink round trip ~2-3 us), so bytes in flight are what buys bandwidth.
__global__ void __launch_bounds__(256)
gather_rows_kernel(void* const* __restrict__ src_rows, int rows, int vec_per_row,
                   uint4* __restrict__ dst) {
  const int lane = threadIdx.x & 31;
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int r = warp; r < rows; r += nwarps) {
    const uint4* src = reinterpret_cast<const uint4*>(
        __ldg(reinterpret_cast<const unsigned long long*>(src_rows + r)));
    uint4* out = dst + size_t(r) * vec_per_row;
    if (!src) {
      for (int i = lane; i < vec_per_row; i += 32) out[i] = make_uint4(0u, 0u, 0u, 0u);
      continue;
    }
    int i = lane;
    for (; i + 96 < vec_per_row; i += 128) {
      const uint4 a = __ldg(src + i), b = __ldg(src + i + 32), c = __ldg(src + i + 64),
                  d = __ldg(src + i + 96);
      out[i] = a; out[i + 32] = b; out[i + 64] = c; out[i + 96] = d;
    }
    for (; i < vec_per_row; i += 32) out[i] = __ldg(src + i);
  }
}

// Topology analysis for the hint flags of include/bsmm.h.  `result` starts as the candidate set
// (twin | even); every violation clears its bit.  One thread per pair of walk rows 2p, 2p+1.
__global__ void topology_hints_kernel(const int* __restrict__ offsets,
                                      const short* __restrict__ indices, int walk_rows,
                                      unsigned twin_bit, unsigned even_bit,
                                      unsigned* __restrict__ result) {
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; 2 * p < walk_rows;
       p += gridDim.x * blockDim.x) {
    const int o0 = __ldg(offsets + 2 * p), o1 = __ldg(offsets + 2 * p + 1);
    const int n0 = o1 - o0;
    if (n0 & 1) atomicAnd(result, ~even_bit);
    if (2 * p + 1 >= walk_rows) {  // odd row count: the last row has no twin
      atomicAnd(result, ~twin_bit);
      continue;
    }
    const int n1 = __ldg(offsets + 2 * p + 2) - o1;
    if (n1 & 1) atomicAnd(result, ~even_bit);
    bool same = n0 == n1;
    for (int i = 0; same && i < n0; ++i) same = __ldg(indices + o0 + i) == __ldg(indices + o1 + i);
    if (!same) atomicAnd(result, ~twin_bit);
  }
}

// One-CTA variant of the analysis for device-side kernel selection (GemmParams::gate): writes
// the final word itself, so the slot needs no initialisation and no host round trip.
__global__ void __launch_bounds__(1024)
topology_gate_kernel(const int* __restrict__ offsets, const short* __restrict__ indices,
                     int walk_rows, unsigned twin_bit, unsigned even_bit,
                     unsigned* __restrict__ out) {
  __shared__ unsigned bad;
  // Launched with programmatic stream serialisation: wait for the previous grid (it may have
  // written the topology), then let the gated GEMM launches behind this one become resident and
  // run their prologues; they block in griddepcontrol.wait until this grid has completed.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (threadIdx.x == 0) bad = (walk_rows & 1) ? twin_bit : 0u;
  __syncthreads();
  for (int p = threadIdx.x; 2 * p + 1 < walk_rows; p += blockDim.x) {
    const int o0 = __ldg(offsets + 2 * p), o1 = __ldg(offsets + 2 * p + 1);
    const int n0 = o1 - o0, n1 = __ldg(offsets + 2 * p + 2) - o1;
    unsigned b = ((n0 | n1) & 1) ? even_bit : 0u;
    bool same = n0 == n1;
    for (int i = 0; same && i < n0; ++i) same = __ldg(indices + o0 + i) == __ldg(indices + o1 + i);
    if (!same) b |= twin_bit;
    if (b) atomicOr(&bad, b);
  }
  __syncthreads();
  if (threadIdx.x == 0) *out = (twin_bit | even_bit) & ~bad;
}

// dMoE topology from the router histogram, entirely on the device (bsmm_moe_topology, bsmm.h).
// Expert e owns n_e = ceil(tokens_e / (128 * align)) * align block rows starting at block row
// start_e, and the F block columns [e F, (e+1) F): a block-diagonal matrix with E dense
// n_e x F rectangles.  Every CTA scans the E counts itself (E <= 4096; a few hundred cycles) and
// then fills a grid-stride share of the arrays, so the whole build is one launch.  The arrays
// are sized for `cap_rows` block rows: rows past the live total are empty (offsets stay at the
// live block count) and the tail of the index arrays is zero-filled.
constexpr int kMoeThreads = 256;
constexpr int kMoeMaxExperts = 4096;

__global__ void __launch_bounds__(kMoeThreads)
moe_topology_kernel(const int* __restrict__ tokens_per_expert, int E, int F, int align,
                    int cap_rows, int* __restrict__ offsets, short* __restrict__ indices,
                    short* __restrict__ row_indices, int* __restrict__ offsets_t,
                    short* __restrict__ indices_t, int* __restrict__ block_offsets,
                    int* __restrict__ padded_bins) {
  __shared__ int start[kMoeMaxExperts + 1];  // first block row of expert e (clamped to cap_rows)
  if (threadIdx.x < 32) {
    const int lane = threadIdx.x;
    int carry = 0;
    long long carry_tokens = 0;  // unclamped padded token rows, reported through padded_bins
    for (int base = 0; base < E; base += 32) {
      const int e = base + lane;
      int n = 0;
      if (e < E) {
        const int tok = max(__ldg(tokens_per_expert + e), 0);
        n = (tok + 128 * align - 1) / (128 * align) * align;
      }
      int incl = n;
      for (int d = 1; d < 32; d <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += up;
      }
      if (e < E) {
        start[e] = min(carry + incl - n, cap_rows);
        if (padded_bins && blockIdx.x == 0)
          padded_bins[e] = int(min(carry_tokens + 128ll * (incl - n), 2147483647ll));
      }
      const int total = __shfl_sync(0xffffffffu, incl, 31);
      carry_tokens += 128ll * total;
      carry = min(carry + total, cap_rows);  // experts past the capacity get no rows
    }
    if (lane == 0) {
      start[E] = carry;
      if (padded_bins && blockIdx.x == 0)
        padded_bins[E] = int(min(carry_tokens, 2147483647ll));
    }
  }
  __syncthreads();
  const int R = start[E];  // live block rows
  const int nnz = R * F, cap_nnz = cap_rows * F;
  const int tid = blockIdx.x * blockDim.x + threadIdx.x, nthreads = gridDim.x * blockDim.x;
  // last e with start[e] <= r (experts with no rows share a start; the last one owns the row)
  auto expert_of_row = [&](int r) {
    int lo = 0, hi = E;  // invariant: start[lo] <= r < start[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (start[mid] <= r) lo = mid; else hi = mid;
    }
    return lo;
  };
  for (int r = tid; r <= cap_rows; r += nthreads) offsets[r] = min(r, R) * F;
  for (int i = tid; i < cap_nnz; i += nthreads) {
    short col = 0, row = 0;
    if (i < nnz) {
      const int r = i / F, j = i - r * F;
      col = short(expert_of_row(r) * F + j);
      row = short(r);
    }
    indices[i] = col;
    row_indices[i] = row;
  }
  if (offsets_t) {
    // column view: expert e's columns hold rows start_e .. start_e + n_e, ascending (what the
    // stable sort of transpose.cu:87-104 gives); its segment starts at F * start_e
    for (int c = tid; c <= E * F; c += nthreads) {
      if (c == E * F) { offsets_t[c] = nnz; continue; }
      const int e = c / F, j = c - e * F;
      offsets_t[c] = F * start[e] + j * (start[e + 1] - start[e]);
    }
    for (int i = tid; i < cap_nnz; i += nthreads) {
      short row = 0;
      int blk = 0;
      if (i < nnz) {
        const int e = expert_of_row(i / F);  // segment e covers [F start_e, F start_{e+1})
        const int n = start[e + 1] - start[e];
        const int u = i - F * start[e], j = u / n, k = u - j * n;
        row = short(start[e] + k);
        blk = (start[e] + k) * F + j;
      }
      indices_t[i] = row;
      block_offsets[i] = blk;
    }
  }
}

// Exchange tables of an expert-sharded dMoE layer, on the device (bsmm_moe_exchange, bsmm.h).
// counts[s * E + e] = token rows rank s holds for global expert e (every rank's rows are sorted by
// expert).  This rank owns experts [e_lo, e_lo + EL).  Its local layout is expert-major,
// source-rank-minor, each expert padded to a multiple of 128 * align rows (exactly the row layout
// bsmm_moe_topology builds from tokens_per_expert = the per-expert sums written here).  For every
// local row i < cap_rows * 128 the kernel finds (source rank s, row r in s's input) and writes
// bases[t][s] + r * pitch into table t (t < ntables: the gather tables of x / dy and the scatter
// tables of y / dx differ only in their base addresses), or 0 for a padding row / a row beyond
// the capacity.  One launch, no host round trip.
constexpr int kMoeExchangeMaxRanks = 64;
constexpr int kMoeExchangeMaxTables = 4;
struct MoeExchangeBases {
  unsigned long long base[kMoeExchangeMaxTables][kMoeExchangeMaxRanks];
  unsigned long long* table[kMoeExchangeMaxTables];
};

__global__ void __launch_bounds__(kMoeThreads)
moe_exchange_kernel(const int* __restrict__ counts, int G, int E, int e_lo, int EL, int align,
                    int cap_rows, unsigned long long pitch, int ntables, MoeExchangeBases b,
                    int* __restrict__ tokens_per_expert) {
  extern __shared__ int ex_sm[];
  int* start = ex_sm;                 // [EL + 1] first padded token row of local expert e
  int* cnt = ex_sm + EL + 1;          // [EL * G] counts[s][e_lo + e], expert-major
  int* src_row0 = cnt + EL * G;       // [EL * G] first row of (s, e) in rank s's input
  const int tid = threadIdx.x;
  // rows of rank s before expert e_lo + e: prefix over ALL global experts (E <= 4096, G <= 64)
  for (int i = tid; i < EL * G; i += blockDim.x) {
    const int e = i / G, s = i - e * G;
    int before = 0;
    for (int g = 0; g < e_lo + e; ++g) before += max(__ldg(counts + s * E + g), 0);
    cnt[i] = max(__ldg(counts + s * E + e_lo + e), 0);
    src_row0[i] = before;
  }
  __syncthreads();
  if (tid == 0) {
    int run = 0;
    for (int e = 0; e < EL; ++e) {
      int tok = 0;
      for (int s = 0; s < G; ++s) tok += cnt[e * G + s];
      start[e] = run;
      if (blockIdx.x == 0 && tokens_per_expert) tokens_per_expert[e] = tok;
      const int q = 128 * align;
      run += (tok + q - 1) / q * q;
    }
    start[EL] = run;
  }
  __syncthreads();
  const int total_rows = cap_rows * 128;
  for (int i = blockIdx.x * blockDim.x + tid; i < total_rows; i += gridDim.x * blockDim.x) {
    unsigned long long off = 0;
    int src = -1;
    if (i < start[EL]) {
      int lo = 0, hi = EL;  // last e with start[e] <= i
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (start[mid] <= i) lo = mid; else hi = mid;
      }
      int j = i - start[lo];
      for (int s = 0; s < G; ++s) {
        const int c = cnt[lo * G + s];
        if (j < c) {
          src = s;
          off = (unsigned long long)(src_row0[lo * G + s] + j) * pitch;
          break;
        }
        j -= c;
      }
    }
    for (int t = 0; t < ntables; ++t) b.table[t][i] = src >= 0 ? b.base[t][src] + off : 0ull;
  }
}

// ---------------------------------------------------------------------------------------------
// Multi-CTA transposed metadata (the default path; the single-CTA kernel below remains the
// fallback for shapes this one does not cover).  One launch, no atomics on global memory, no
// workspace, deterministic:
//   * CTA b OWNS the block columns [b * cpc, (b + 1) * cpc).  It scans ALL nb column indices
//     (int16, L2-resident: 2 nb bytes per CTA) and keeps only its own — counting on the way how
//     many indices lie below its range, which is the start of its slice of every output array;
//   * pass 0 counts its columns (shared-memory histogram, touched only by matches) -> offsets_t;
//   * pass 1 compacts the matching block ids IN STORAGE ORDER into a shared-memory list (warp
//     scan of per-thread match counts + a scan over the 32 warp totals per 8192 indices), then
//     one warp per column pulls its blocks out of the list in order (ballot + prefix popcount):
//     ascending storage position inside every column == the reference's std::stable_sort
//     (transpose.cu:11-19, 87-104).  indices_t = owning row of each block by binary search.
// A column range that holds more blocks than the list is split into rounds (another scan per
// round); a SINGLE column longer than the list is streamed straight to global memory (inside one
// column, list order is output order).
constexpr int kTrThreads = 1024;
constexpr int kTrPerThread = 8;                         // indices per thread per iteration
constexpr int kTrChunk = kTrThreads * kTrPerThread;     // 8192 indices per iteration
constexpr int kTrListCap = 20 * 1024;                   // entries of the compaction list (80 KiB)
constexpr int kTrMaxOwnCols = 256;

__global__ void __launch_bounds__(kTrThreads, 1)
transpose_metadata_mc_kernel(const int* __restrict__ offsets, const short* __restrict__ indices,
                             int block_rows, int block_cols, int nb, int cpc,
                             int* __restrict__ offsets_t, short* __restrict__ indices_t,
                             int* __restrict__ block_offsets) {
  extern __shared__ unsigned tr_list[];   // kTrListCap entries: (column - round start) << 17 | block id
  __shared__ int hist[kTrMaxOwnCols + 1];
  __shared__ int wtot[2][32];
  __shared__ int s_lt, s_cb;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // Programmatic stream serialisation (see topology_gate_kernel): the product that follows may
  // start its prologue while this grid runs; it waits for this grid's completion itself.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int c_lo = blockIdx.x * cpc;
  if (c_lo >= block_cols) return;
  const int ncols = min(cpc, block_cols - c_lo);
  const bool vec_ok = (reinterpret_cast<uintptr_t>(indices) & 15) == 0;

  // 8 consecutive indices starting at `base` (vector load when possible); entries past nb read
  // as 32768 (never a valid column, never below any range).
  auto load8 = [&](int base, int (&c)[kTrPerThread]) {
    if (vec_ok && base + kTrPerThread <= nb) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(indices + base));
      const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        c[2 * e] = int(short(w[e] & 0xffffu));
        c[2 * e + 1] = int(short(w[e] >> 16));
      }
    } else {
#pragma unroll
      for (int e = 0; e < kTrPerThread; ++e)
        c[e] = base + e < nb ? int(__ldg(indices + base + e)) : 32768;
    }
  };

  for (int i = tid; i <= ncols; i += kTrThreads) hist[i] = 0;
  if (tid == 0) s_lt = 0;
  __syncthreads();
  // ---- pass 0: blocks below the range, blocks per own column
  int lt = 0;
  for (int base = tid * kTrPerThread; base < nb; base += kTrChunk) {
    int c[kTrPerThread];
    load8(base, c);
#pragma unroll
    for (int e = 0; e < kTrPerThread; ++e) {
      lt += int(c[e] < c_lo);
      const unsigned d = unsigned(c[e] - c_lo);
      if (d < unsigned(ncols)) atomicAdd(&hist[d], 1);
    }
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) lt += __shfl_xor_sync(0xffffffffu, lt, d);
  if (lane == 0 && lt) atomicAdd(&s_lt, lt);
  __syncthreads();
  // exclusive scan of the (<= 256) own counts by warp 0: 8 per lane
  if (warp == 0) {
    int v[8], sum = 0;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int i = lane * 8 + e;
      v[e] = i < ncols ? hist[i] : 0;
      sum += v[e];
    }
    int incl = sum;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int up = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += up;
    }
    int run = incl - sum;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int i = lane * 8 + e;
      if (i < ncols) hist[i] = run;
      run += v[e];
    }
    if (lane == 31) hist[ncols] = incl;   // blocks in the whole own range
  }
  __syncthreads();
  const int out0 = s_lt;                  // first output position of the own range
  for (int i = tid; i < ncols; i += kTrThreads) offsets_t[c_lo + i] = out0 + hist[i];
  if (tid == 0 && c_lo + ncols == block_cols) offsets_t[block_cols] = out0 + hist[ncols];

  // ---- rounds over sub-ranges [ca, cb) of the own columns that fit the list
  int ca = 0;
  while (ca < ncols) {
    if (tid == 0) {
      int cb = ca + 1;
      while (cb < ncols && hist[cb + 1] - hist[ca] <= kTrListCap) ++cb;
      s_cb = cb;
    }
    __syncthreads();
    const int cb = s_cb;
    const int n_round = hist[cb] - hist[ca];
    if (n_round == 0) { __syncthreads(); ca = cb; continue; }
    const bool streaming = n_round > kTrListCap;   // a single long column
    const int round_out = out0 + hist[ca];
    int running = 0;
    for (int it = 0; it * kTrChunk < nb; ++it) {
      const int base = it * kTrChunk + tid * kTrPerThread;
      int c[kTrPerThread];
      unsigned mask = 0;
      if (base < nb) {
        load8(base, c);
#pragma unroll
        for (int e = 0; e < kTrPerThread; ++e)
          if (unsigned(c[e] - c_lo - ca) < unsigned(cb - ca)) mask |= 1u << e;
      }
      const int cnt = __popc(mask);
      int incl = cnt;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += up;
      }
      if (lane == 31) wtot[it & 1][warp] = incl;
      __syncthreads();
      const int w = wtot[it & 1][lane];
      int wincl = w;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, wincl, d);
        if (lane >= d) wincl += up;
      }
      const int wprefix = __shfl_sync(0xffffffffu, wincl - w, warp);
      const int total = __shfl_sync(0xffffffffu, wincl, 31);
      int pos = running + wprefix + incl - cnt;
#pragma unroll
      for (int e = 0; e < kTrPerThread; ++e) {
        if (mask & (1u << e)) {
          const int j = base + e;
          if (!streaming) {
            tr_list[pos] = (unsigned(c[e] - c_lo - ca) << 17) | unsigned(j);
          } else {
            block_offsets[round_out + pos] = j;
            indices_t[round_out + pos] = short(row_of_block(offsets, block_rows, j));
          }
          ++pos;
        }
      }
      running += total;
    }
    __syncthreads();
    if (!streaming) {
      for (int col = ca + warp; col < cb; col += kTrThreads / 32) {
        const int n_col = hist[col + 1] - hist[col];
        const int col_out = out0 + hist[col];
        int found = 0;
        for (int i0 = 0; i0 < n_round && found < n_col; i0 += 32) {
          const int i = i0 + lane;
          const unsigned ent = i < n_round ? tr_list[i] : 0xffffffffu;
          const bool m = i < n_round && int(ent >> 17) == col - ca;
          const unsigned bal = __ballot_sync(0xffffffffu, m);
          if (m) {
            const int p = col_out + found + __popc(bal & ((1u << lane) - 1u));
            const int j = int(ent & 0x1ffffu);
            block_offsets[p] = j;
            indices_t[p] = short(row_of_block(offsets, block_rows, j));
          }
          found += __popc(bal);
        }
      }
      __syncthreads();
    }
    ca = cb;
  }
}

constexpr int kTransposeThreads = 1024;

// Single-CTA counting sort by block column.  Shared memory: hist[block_cols + 1] then
// seg[seg_cap] (scratch for columns holding more than 32 blocks).
__global__ void __launch_bounds__(kTransposeThreads, 1)
transpose_metadata_kernel(const int* __restrict__ offsets, const short* __restrict__ indices,
                          int block_rows, int block_cols, int nb, int seg_cap,
                          int* __restrict__ offsets_t, short* __restrict__ indices_t,
                          int* __restrict__ block_offsets) {
  extern __shared__ int tsm[];
  int* hist = tsm;
  int* seg = tsm + block_cols + 1;
  __shared__ int warp_sums[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  for (int c = tid; c <= block_cols; c += kTransposeThreads) hist[c] = 0;
  __syncthreads();
  for (int j = tid; j < nb; j += kTransposeThreads) atomicAdd(&hist[int(__ldg(indices + j))], 1);
  __syncthreads();

  // Exclusive scan of hist[0, block_cols): contiguous chunk per thread + block scan of sums.
  const int per = (block_cols + kTransposeThreads - 1) / kTransposeThreads;
  const int c_lo = min(tid * per, block_cols), c_hi = min(c_lo + per, block_cols);
  int local = 0;
  for (int c = c_lo; c < c_hi; ++c) local += hist[c];
  int incl = local;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int up = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += up;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = warp_sums[lane];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int up = __shfl_up_sync(0xffffffffu, w, d);
      if (lane >= d) w += up;
    }
    warp_sums[lane] = w;  // inclusive over warps
  }
  __syncthreads();
  int run = incl - local + (warp ? warp_sums[warp - 1] : 0);
  for (int c = c_lo; c < c_hi; ++c) {
    const int cnt = hist[c];
    hist[c] = run;  // becomes the scatter cursor
    offsets_t[c] = run;
    run += cnt;
  }
  if (tid == 0) offsets_t[block_cols] = nb;
  __syncthreads();

  // Scatter (order inside a column is fixed up below).
  for (int j = tid; j < nb; j += kTransposeThreads) {
    const int pos = atomicAdd(&hist[int(__ldg(indices + j))], 1);
    block_offsets[pos] = j;
  }
  __syncthreads();

  // Stable order inside a column == ascending position in `data`.  Columns with <= 32
  // blocks: one warp ranks them with shuffles.
  for (int c = warp; c < block_cols; c += kTransposeThreads / 32) {
    const int s = offsets_t[c], e = offsets_t[c + 1], n = e - s;
    if (n > 1 && n <= 32) {
      const int v = lane < n ? block_offsets[s + lane] : INT_MAX;
      int rank = 0;
#pragma unroll
      for (int i = 0; i < 32; ++i) rank += int(__shfl_sync(0xffffffffu, v, i) < v);
      if (lane < n) block_offsets[s + rank] = v;
    }
  }
  __syncthreads();
  // Longer columns: the whole CTA ranks one column at a time through shared memory.
  for (int c = 0; c < block_cols; ++c) {
    const int s = offsets_t[c], n = offsets_t[c + 1] - s;
    if (n <= 32) continue;  // uniform across the CTA
    if (n <= seg_cap) {
      for (int i = tid; i < n; i += kTransposeThreads) seg[i] = block_offsets[s + i];
      __syncthreads();
      for (int i = tid; i < n; i += kTransposeThreads) {
        const int v = seg[i];
        int rank = 0;
        for (int x = 0; x < n; ++x) rank += int(seg[x] < v);
        block_offsets[s + rank] = v;
      }
      __syncthreads();
    } else {
      if (tid == 0) {  // insertion sort: only for columns that do not fit shared memory
        for (int i = 1; i < n; ++i) {
          const int v = block_offsets[s + i];
          int x = i - 1;
          while (x >= 0 && block_offsets[s + x] > v) {
            block_offsets[s + x + 1] = block_offsets[s + x];
            --x;
          }
          block_offsets[s + x + 1] = v;
        }
      }
      __syncthreads();
    }
  }
  __syncthreads();
  for (int i = tid; i < nb; i += kTransposeThreads)
    indices_t[i] = short(row_of_block(offsets, block_rows, block_offsets[i]));
}

}  // namespace bsmm
