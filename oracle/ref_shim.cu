// ref_shim.cu — extern "C" doorway into the UNMODIFIED reference, compiled from the sources where
// they lie under /root/reference (see oracle/Makefile; output oracle/_ref/libsputnik_ref.so).
// TEST INFRASTRUCTURE ONLY.  Contains no reference code: it only calls the reference's public
// API (sputnik/sputnik.h:18-25, block/transpose/transpose.h:10, block/row_indices/row_indices.h:10).
//
// Used by: tests (GPU parity of our kernels and of the device-side transpose against the
// reference's own kernels / host Transpose()), and bench.py's legacy-GPU baseline line.
#include <cuda_runtime.h>

#include "sputnik/block/arguments.h"
#include "sputnik/block/dds/dds.h"
#include "sputnik/block/dsd/dsd.h"
#include "sputnik/block/row_indices/row_indices.h"
#include "sputnik/block/sdd/sdd.h"
#include "sputnik/block/transpose/transpose.h"

namespace sb = sputnik::block;

namespace {
sb::BlockMatrix make_block(int rows, int cols, int nonzeros, void* data, void* offsets,
                           void* indices, void* offsets_t, void* indices_t, void* block_offsets,
                           void* row_indices) {
  sb::BlockMatrix m(rows, cols, sb::BlockSize::k128, nonzeros, data, offsets, indices);
  m.offsets_t = offsets_t;
  m.indices_t = indices_t;
  m.block_offsets = block_offsets;
  m.row_indices = row_indices;
  return m;
}
}  // namespace

extern "C" {

// fp16 only: the reference instantiates cutlass::half_t kernels and nothing else.
int ref_sdd(int a_rows, int a_cols, void* a, int ta, int b_rows, int b_cols, void* b, int tb,
            int c_rows, int c_cols, int c_nonzeros, void* c_data, void* c_offsets,
            void* c_indices, void* c_row_indices, void* stream) {
  sb::Matrix ma(a_rows, a_cols, a), mb(b_rows, b_cols, b);
  sb::BlockMatrix mc = make_block(c_rows, c_cols, c_nonzeros, c_data, c_offsets, c_indices,
                                  nullptr, nullptr, nullptr, c_row_indices);
  return int(sb::Matmul(ma, ta != 0, mb, tb != 0, mc, static_cast<cudaStream_t>(stream)));
}

int ref_dsd(int a_rows, int a_cols, int a_nonzeros, void* a_data, void* a_offsets,
            void* a_indices, void* a_offsets_t, void* a_indices_t, void* a_block_offsets, int ta,
            int b_rows, int b_cols, void* b, int tb, int c_rows, int c_cols, void* c,
            int create_metadata, void* stream) {
  sb::BlockMatrix ma = make_block(a_rows, a_cols, a_nonzeros, a_data, a_offsets, a_indices,
                                  a_offsets_t, a_indices_t, a_block_offsets, nullptr);
  sb::Matrix mb(b_rows, b_cols, b), mc(c_rows, c_cols, c);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  return int(create_metadata ? sb::Matmul(ma, ta != 0, mb, tb != 0, mc, s)
                             : sb::MatmulEx(ma, ta != 0, mb, tb != 0, mc, s));
}

int ref_dds(int a_rows, int a_cols, void* a, int ta, int b_rows, int b_cols, int b_nonzeros,
            void* b_data, void* b_offsets, void* b_indices, void* b_offsets_t, void* b_indices_t,
            void* b_block_offsets, int tb, int c_rows, int c_cols, void* c, int create_metadata,
            void* stream) {
  sb::Matrix ma(a_rows, a_cols, a), mc(c_rows, c_cols, c);
  sb::BlockMatrix mb = make_block(b_rows, b_cols, b_nonzeros, b_data, b_offsets, b_indices,
                                  b_offsets_t, b_indices_t, b_block_offsets, nullptr);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  return int(create_metadata ? sb::Matmul(ma, ta != 0, mb, tb != 0, mc, s)
                             : sb::MatmulEx(ma, ta != 0, mb, tb != 0, mc, s));
}

int ref_transpose(int rows, int cols, int nonzeros, void* offsets, void* indices, void* offsets_t,
                  void* indices_t, void* block_offsets, void* stream) {
  sb::BlockMatrix m = make_block(rows, cols, nonzeros, nullptr, offsets, indices, offsets_t,
                                 indices_t, block_offsets, nullptr);
  return int(sb::Transpose(m, static_cast<cudaStream_t>(stream)));
}

int ref_row_indices(int rows, int cols, int nonzeros, void* offsets, void* indices,
                    short* row_indices, void* stream) {
  sb::BlockMatrix m = make_block(rows, cols, nonzeros, nullptr, offsets, indices, nullptr,
                                 nullptr, nullptr, nullptr);
  return int(sb::RowIndices(m, row_indices, static_cast<cudaStream_t>(stream)));
}

}  // extern "C"
