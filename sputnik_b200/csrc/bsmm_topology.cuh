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
