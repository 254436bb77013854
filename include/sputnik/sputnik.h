// sputnik/sputnik.h — header-only C++ shim over libbsmm.so (include/bsmm.h).
//
// Source-level drop-in for the block-sparse SDD / DSD / DDS path of tgale96/sputnik (and its
// siblings SSD / SDS / DSS, block/{ssd,sds,dss}/*.h, block/bitmask/bitmask.h): code written
// against `#include "sputnik/sputnik.h"` and `namespace sputnik::block` (reference
// sputnik/sputnik.h:18-25, sputnik/block/arguments.h:13-274, block/{sdd,dsd,dds}/*.h,
// block/row_indices/row_indices.h:10, block/transpose/transpose.h:10) recompiles unchanged and
// links with -lbsmm instead of -lsputnik.  Everything here forwards to the C-ABI; there is no
// computation in this header.
//
// Fidelity notes:
//  * the reference has no dtype field (fp16 is implied by its instantiations): the reference
//    signatures below compute in fp16; bf16 is reached through the extra overloads that take a
//    trailing `DType` (an extension);
//  * error behaviour is mapped back onto the reference's: cudaErrorNotSupported for block sizes
//    other than 128 (sdd.cu:11-14) and for misaligned problems (the launchers' can_implement
//    failure, sdd_mixed_b128_128x128x32x5_nn_align8.cu:96-99); abort() with a message where the
//    reference hits SPUTNIK_LOG(FATAL) / SPUTNIK_CHECK (missing metadata pointer, shape
//    mismatch: dsd_..._tn_align8.cu:73-75, block/sdd/cutlass/sdd.cu:66-70, logging.cc:14-18);
//  * Transpose() runs on the device and does not stall the stream (the reference's host version
//    blocks on two D2H copies, transpose.cu:78-89); results are bit-identical.
#ifndef BSMM_SPUTNIK_SHIM_H_
#define BSMM_SPUTNIK_SHIM_H_

#include <cstdio>
#include <cstdlib>

#include <cuda_runtime.h>

#include "../bsmm.h"

namespace sputnik {
namespace block {

enum class BlockSize { kNone = 0, k16 = 16, k32 = 32, k64 = 64, k128 = 128 };

enum class DType { kHalf = BSMM_F16, kBfloat16 = BSMM_BF16 };

inline int AsInt(BlockSize b) {
  switch (b) {
    case BlockSize::k16: case BlockSize::k32: case BlockSize::k64: case BlockSize::k128:
      return static_cast<int>(b);
    default:
      return 0;
  }
}

inline BlockSize AsBlockSize(int b) {
  if (b == 16 || b == 32 || b == 64 || b == 128) return static_cast<BlockSize>(b);
  std::fprintf(stderr, "[FATAL] Invalid block size.\n");
  std::abort();
}

// BCSR handle: same public fields, same five constructors (arguments.h:48-153).
struct BlockMatrix {
  int rows, cols, nonzeros;
  BlockSize block_size;
  void* data;
  void* offsets;
  void* indices;
  void* offsets_t;
  void* indices_t;
  void* block_offsets;
  void* row_indices;
  void* bitmask;  // filled by Bitmask(); the library's own DSS does not read it
  bool create_metadata = true;
  // Extensions (include/bsmm.h BSMM_FLAG_TWIN_ROWS / TWIN_COLS / EVEN_ROWS): promises about the
  // topology that select the 2-SM kernels.  Off by default: unmodified callers never set them.
  bool twin_rows = false;
  bool twin_cols = false;
  bool even_rows = false;

  BlockMatrix(int r, int c, BlockSize bs, int nnz, void const* d, void const* off, void const* idx)
      : BlockMatrix(r, c, bs, nnz, d, off, idx, nullptr, nullptr, nullptr, nullptr, nullptr) {}

  BlockMatrix(int r, int c, BlockSize bs, int nnz, void const* d, void const* off, void const* idx,
              void const* row_idx)
      : BlockMatrix(r, c, bs, nnz, d, off, idx, nullptr, nullptr, nullptr, row_idx, nullptr) {}

  BlockMatrix(int r, int c, BlockSize bs, int nnz, void const* d, void const* off, void const* idx,
              void const* off_t, void const* idx_t, void const* blk_off)
      : BlockMatrix(r, c, bs, nnz, d, off, idx, off_t, idx_t, blk_off, nullptr, nullptr) {}

  BlockMatrix(int r, int c, BlockSize bs, int nnz, void const* d, void const* off, void const* idx,
              void const* off_t, void const* idx_t, void const* blk_off, void const* mask)
      : BlockMatrix(r, c, bs, nnz, d, off, idx, off_t, idx_t, blk_off, nullptr, mask) {}

 private:
  BlockMatrix(int r, int c, BlockSize bs, int nnz, void const* d, void const* off, void const* idx,
              void const* off_t, void const* idx_t, void const* blk_off, void const* row_idx,
              void const* mask)
      : rows(r), cols(c), nonzeros(nnz), block_size(bs),
        data(const_cast<void*>(d)), offsets(const_cast<void*>(off)),
        indices(const_cast<void*>(idx)), offsets_t(const_cast<void*>(off_t)),
        indices_t(const_cast<void*>(idx_t)), block_offsets(const_cast<void*>(blk_off)),
        row_indices(const_cast<void*>(row_idx)), bitmask(const_cast<void*>(mask)) {}
};

struct Matrix {
  int rows, cols;
  void* data;
  Matrix(int r, int c, void const* d) : rows(r), cols(c), data(const_cast<void*>(d)) {}
};

// arguments.h:164-231
struct MatmulShape {
  int m, n, k, lda, ldb, ldc;
  MatmulShape(int a_rows, int a_cols, bool ta, int b_rows, int b_cols, bool tb)
      : m(ta ? a_cols : a_rows), n(tb ? b_rows : b_cols), k(ta ? a_rows : a_cols),
        lda(a_cols), ldb(b_cols), ldc(tb ? b_rows : b_cols) {}
  template <class A, class B>
  MatmulShape(const A& a, bool ta, const B& b, bool tb)
      : MatmulShape(a.rows, a.cols, ta, b.rows, b.cols, tb) {}
};

template <class A, class B, class C>
inline bool ValidMatmul(const A& a, bool ta, const B& b, bool tb, const C& c) {
  MatmulShape s(a, ta, b, tb);
  return (ta ? a.rows : a.cols) == s.k && (tb ? b.cols : b.rows) == s.k &&
         (tb ? b.rows : b.cols) == s.n && c.rows == s.m && c.cols == s.n;
}

namespace detail {

inline bsmm_block_matrix Pod(const BlockMatrix& a) {
  return bsmm_block_matrix{a.rows, a.cols, a.nonzeros, AsInt(a.block_size), a.data, a.offsets,
                           a.indices, a.offsets_t, a.indices_t, a.block_offsets, a.row_indices};
}
inline bsmm_matrix Pod(const Matrix& a) { return bsmm_matrix{a.rows, a.cols, a.data}; }

inline unsigned Hints(const BlockMatrix& a) {
  return (a.twin_rows ? BSMM_FLAG_TWIN_ROWS : 0u) | (a.twin_cols ? BSMM_FLAG_TWIN_COLS : 0u) |
         (a.even_rows ? BSMM_FLAG_EVEN_ROWS : 0u);
}

[[noreturn]] inline void Fatal(const char* what, int status) {
  std::fprintf(stderr, "[FATAL] sputnik::block: %s (%s)\n", what, bsmm_status_string(status));
  std::abort();
}

// Known limit: DSS needs K <= 262144 (2048 blocks; the reference's index merge allows 32768 blocks,
// index_merge.h:103) — beyond it bsmm_dss returns BSMM_ERR_RANGE, mapped to cudaErrorNotSupported.
// bsmm_status -> the reference's observable behaviour.
inline cudaError_t Map(int status, const char* what) {
  switch (status) {
    case BSMM_OK: return cudaSuccess;
    case BSMM_ERR_BLOCK_SIZE:
    case BSMM_ERR_MISALIGNED:
    case BSMM_ERR_DTYPE:
    case BSMM_ERR_RANGE: return cudaErrorNotSupported;
    case BSMM_ERR_CUDA: return static_cast<cudaError_t>(bsmm_last_cuda_error());
    case BSMM_ERR_HINT: return cudaErrorInvalidValue;  // BSMM_VALIDATE_HINTS=1 (debug) only
    default: Fatal(what, status);  // missing metadata / shape mismatch: the reference aborts
  }
}

inline void CudaCheck(cudaError_t e) {
  if (e != cudaSuccess) {
    std::fprintf(stderr, "[FATAL] CUDA Error: %s\n", cudaGetErrorString(e));
    std::abort();
  }
}

}  // namespace detail

// ---- SDD: sdd/sdd.h:10-15
// `epilogue` (extension, on every DType overload): BSMM_FLAG_ACCUMULATE / RELU / GELU, or-ed.
inline cudaError_t Matmul(const Matrix a, bool transpose_a, const Matrix b, bool transpose_b,
                          BlockMatrix c, cudaStream_t stream, DType dtype,
                          unsigned epilogue = BSMM_FLAG_NONE) {
  bsmm_matrix pa = detail::Pod(a), pb = detail::Pod(b);
  bsmm_block_matrix pc = detail::Pod(c);
  return detail::Map(bsmm_sdd(&pa, transpose_a, &pb, transpose_b, &pc, static_cast<int>(dtype),
                              detail::Hints(c) | epilogue, stream), "SDD");
}
inline cudaError_t Matmul(const Matrix a, bool transpose_a, const Matrix b, bool transpose_b,
                          BlockMatrix c, cudaStream_t stream) {
  return Matmul(a, transpose_a, b, transpose_b, c, stream, DType::kHalf);
}

// ---- DSD: dsd/dsd.h:10-22.  Matmul builds the transposed metadata when it needs it and
// a.create_metadata is set (dsd.cu:9-18); MatmulEx takes it as valid (dsd.cu:20-27).
inline cudaError_t MatmulEx(const BlockMatrix a, bool transpose_a, const Matrix b,
                            bool transpose_b, Matrix c, cudaStream_t stream, DType dtype,
                            unsigned epilogue = BSMM_FLAG_NONE) {
  bsmm_block_matrix pa = detail::Pod(a);
  bsmm_matrix pb = detail::Pod(b), pc = detail::Pod(c);
  return detail::Map(bsmm_dsd(&pa, transpose_a, &pb, transpose_b, &pc, static_cast<int>(dtype),
                              detail::Hints(a) | epilogue, stream), "DSD");
}
inline cudaError_t MatmulEx(const BlockMatrix a, bool transpose_a, const Matrix b,
                            bool transpose_b, Matrix c, cudaStream_t stream) {
  return MatmulEx(a, transpose_a, b, transpose_b, c, stream, DType::kHalf);
}
inline cudaError_t Matmul(const BlockMatrix a, bool transpose_a, const Matrix b, bool transpose_b,
                          Matrix c, cudaStream_t stream, DType dtype,
                          unsigned epilogue = BSMM_FLAG_NONE) {
  bsmm_block_matrix pa = detail::Pod(a);
  bsmm_matrix pb = detail::Pod(b), pc = detail::Pod(c);
  const unsigned flags = (a.create_metadata ? BSMM_FLAG_CREATE_METADATA : BSMM_FLAG_NONE) |
                         detail::Hints(a) | epilogue;
  return detail::Map(bsmm_dsd(&pa, transpose_a, &pb, transpose_b, &pc, static_cast<int>(dtype),
                              flags, stream), "DSD");
}
inline cudaError_t Matmul(const BlockMatrix a, bool transpose_a, const Matrix b, bool transpose_b,
                          Matrix c, cudaStream_t stream) {
  return Matmul(a, transpose_a, b, transpose_b, c, stream, DType::kHalf);
}

// ---- DDS: dds/dds.h:10-22 (dds.cu:7-24)
inline cudaError_t MatmulEx(const Matrix a, bool transpose_a, const BlockMatrix b,
                            bool transpose_b, Matrix c, cudaStream_t stream, DType dtype,
                            unsigned epilogue = BSMM_FLAG_NONE) {
  bsmm_matrix pa = detail::Pod(a), pc = detail::Pod(c);
  bsmm_block_matrix pb = detail::Pod(b);
  return detail::Map(bsmm_dds(&pa, transpose_a, &pb, transpose_b, &pc, static_cast<int>(dtype),
                              detail::Hints(b) | epilogue, stream), "DDS");
}
inline cudaError_t MatmulEx(const Matrix a, bool transpose_a, const BlockMatrix b,
                            bool transpose_b, Matrix c, cudaStream_t stream) {
  return MatmulEx(a, transpose_a, b, transpose_b, c, stream, DType::kHalf);
}
inline cudaError_t Matmul(const Matrix a, bool transpose_a, const BlockMatrix b, bool transpose_b,
                          Matrix c, cudaStream_t stream, DType dtype,
                          unsigned epilogue = BSMM_FLAG_NONE) {
  bsmm_matrix pa = detail::Pod(a), pc = detail::Pod(c);
  bsmm_block_matrix pb = detail::Pod(b);
  const unsigned flags = (b.create_metadata ? BSMM_FLAG_CREATE_METADATA : BSMM_FLAG_NONE) |
                         detail::Hints(b) | epilogue;
  return detail::Map(bsmm_dds(&pa, transpose_a, &pb, transpose_b, &pc, static_cast<int>(dtype),
                              flags, stream), "DDS");
}
inline cudaError_t Matmul(const Matrix a, bool transpose_a, const BlockMatrix b, bool transpose_b,
                          Matrix c, cudaStream_t stream) {
  return Matmul(a, transpose_a, b, transpose_b, c, stream, DType::kHalf);
}

// ---- SSD: ssd/ssd.h (ssd.cu:7-24): sparse x dense -> sparse (sampled at c's topology)
inline cudaError_t Matmul(const BlockMatrix a, bool transpose_a, const Matrix b, bool transpose_b,
                          BlockMatrix c, cudaStream_t stream, DType dtype) {
  bsmm_block_matrix pa = detail::Pod(a), pc = detail::Pod(c);
  bsmm_matrix pb = detail::Pod(b);
  const unsigned flags = a.create_metadata ? BSMM_FLAG_CREATE_METADATA : BSMM_FLAG_NONE;
  return detail::Map(bsmm_ssd(&pa, transpose_a, &pb, transpose_b, &pc, static_cast<int>(dtype),
                              flags, stream), "SSD");
}
inline cudaError_t Matmul(const BlockMatrix a, bool transpose_a, const Matrix b, bool transpose_b,
                          BlockMatrix c, cudaStream_t stream) {
  return Matmul(a, transpose_a, b, transpose_b, c, stream, DType::kHalf);
}
inline cudaError_t MatmulEx(const BlockMatrix a, bool transpose_a, const Matrix b,
                            bool transpose_b, BlockMatrix c, cudaStream_t stream,
                            DType dtype = DType::kHalf) {
  BlockMatrix acp = a;
  acp.create_metadata = false;
  return Matmul(acp, transpose_a, b, transpose_b, c, stream, dtype);
}

// ---- SDS: sds/sds.h (sds.cu:7-24): dense x sparse -> sparse
inline cudaError_t Matmul(const Matrix a, bool transpose_a, const BlockMatrix b, bool transpose_b,
                          BlockMatrix c, cudaStream_t stream, DType dtype) {
  bsmm_matrix pa = detail::Pod(a);
  bsmm_block_matrix pb = detail::Pod(b), pc = detail::Pod(c);
  const unsigned flags = b.create_metadata ? BSMM_FLAG_CREATE_METADATA : BSMM_FLAG_NONE;
  return detail::Map(bsmm_sds(&pa, transpose_a, &pb, transpose_b, &pc, static_cast<int>(dtype),
                              flags, stream), "SDS");
}
inline cudaError_t Matmul(const Matrix a, bool transpose_a, const BlockMatrix b, bool transpose_b,
                          BlockMatrix c, cudaStream_t stream) {
  return Matmul(a, transpose_a, b, transpose_b, c, stream, DType::kHalf);
}
inline cudaError_t MatmulEx(const Matrix a, bool transpose_a, const BlockMatrix b,
                            bool transpose_b, BlockMatrix c, cudaStream_t stream,
                            DType dtype = DType::kHalf) {
  BlockMatrix bcp = b;
  bcp.create_metadata = false;
  return Matmul(a, transpose_a, bcp, transpose_b, c, stream, dtype);
}

// ---- DSS: dss/dss.h (dss.cu:9-27): sparse x sparse -> dense.  create_metadata is per matrix
// (dss_mixed_b128_128x128x32x5_tn_align8.cu:85-97): the column views asked for are built here,
// then the product runs on valid metadata.  The reference's bitmask buffers are not needed
// (the intersection runs inside the kernel); a null `bitmask` is therefore not an error.
inline cudaError_t Matmul(const BlockMatrix a, bool transpose_a, const BlockMatrix b,
                          bool transpose_b, Matrix c, cudaStream_t stream, DType dtype) {
  bsmm_block_matrix pa = detail::Pod(a), pb = detail::Pod(b);
  bsmm_matrix pc = detail::Pod(c);
  int st = bsmm_dss_check(&pa, transpose_a, &pb, transpose_b, &pc, static_cast<int>(dtype), 0u);
  if (st == BSMM_OK && transpose_a && a.create_metadata) st = bsmm_transpose_metadata(&pa, stream);
  if (st == BSMM_OK && !transpose_b && b.create_metadata) st = bsmm_transpose_metadata(&pb, stream);
  if (st != BSMM_OK) return detail::Map(st, "DSS");
  return detail::Map(bsmm_dss(&pa, transpose_a, &pb, transpose_b, &pc, static_cast<int>(dtype),
                              BSMM_FLAG_NONE, stream), "DSS");
}
inline cudaError_t Matmul(const BlockMatrix a, bool transpose_a, const BlockMatrix b,
                          bool transpose_b, Matrix c, cudaStream_t stream) {
  return Matmul(a, transpose_a, b, transpose_b, c, stream, DType::kHalf);
}
inline cudaError_t MatmulEx(const BlockMatrix a, bool transpose_a, const BlockMatrix b,
                            bool transpose_b, Matrix c, cudaStream_t stream,
                            DType dtype = DType::kHalf) {
  BlockMatrix acp = a, bcp = b;
  acp.create_metadata = false;
  bcp.create_metadata = false;
  return Matmul(acp, transpose_a, bcp, transpose_b, c, stream, dtype);
}

// ---- bitmask/bitmask.h: Bitmask (bitmask.cu:7-45) on the device, Allocate / FreeBitmaskBuffers
inline cudaError_t Bitmask(BlockMatrix m, cudaStream_t stream) {
  bsmm_block_matrix pm = detail::Pod(m);
  return detail::Map(bsmm_bitmask(&pm, m.bitmask, stream), "Bitmask");
}
inline void AllocateBitmaskBuffers(BlockMatrix& m) {
  bsmm_block_matrix pm = detail::Pod(m);
  size_t bytes = 0;
  if (bsmm_bitmask_bytes(&pm, &bytes) != BSMM_OK)
    detail::Fatal("AllocateBitmaskBuffers", BSMM_ERR_BLOCK_SIZE);
  detail::CudaCheck(cudaMalloc(&m.bitmask, bytes));
}
inline void FreeBitmaskBuffers(BlockMatrix& m) {
  if (m.bitmask) detail::CudaCheck(cudaFree(m.bitmask));
}

// ---- row_indices/row_indices.h:10
inline cudaError_t RowIndices(BlockMatrix a, short* row_indices, cudaStream_t stream) {
  bsmm_block_matrix pa = detail::Pod(a);
  return detail::Map(bsmm_row_indices(&pa, row_indices, stream), "RowIndices");
}

// ---- transpose/transpose.h:10
inline cudaError_t Transpose(BlockMatrix a, cudaStream_t stream) {
  bsmm_block_matrix pa = detail::Pod(a);
  return detail::Map(bsmm_transpose_metadata(&pa, stream), "Transpose");
}

// ---- extension (no reference counterpart): fill a.twin_rows / twin_cols / even_rows by a device
// analysis of the topology (bsmm_analyze_topology; synchronous, once per topology).  twin_cols
// needs valid transposed metadata: call after Transpose().
inline cudaError_t AnalyzeTopology(BlockMatrix& a, cudaStream_t stream) {
  bsmm_block_matrix pa = detail::Pod(a);
  unsigned flags = 0;
  cudaError_t e = detail::Map(bsmm_analyze_topology(&pa, &flags, stream), "AnalyzeTopology");
  if (e != cudaSuccess) return e;
  a.twin_rows = flags & BSMM_FLAG_TWIN_ROWS;
  a.twin_cols = flags & BSMM_FLAG_TWIN_COLS;
  a.even_rows = flags & BSMM_FLAG_EVEN_ROWS;
  return cudaSuccess;
}

// ---- caller-side buffer helpers (arguments.h:233-274): explicit cudaMalloc / cudaFree.
inline void AllocateTransposeBuffers(BlockMatrix& a) {
  bsmm_block_matrix pa = detail::Pod(a);
  size_t off = 0, idx = 0, blk = 0;
  if (bsmm_transpose_buffer_bytes(&pa, &off, &idx, &blk) != BSMM_OK)
    detail::Fatal("AllocateTransposeBuffers", BSMM_ERR_BLOCK_SIZE);
  detail::CudaCheck(cudaMalloc(&a.offsets_t, off));
  detail::CudaCheck(cudaMalloc(&a.indices_t, idx));
  detail::CudaCheck(cudaMalloc(&a.block_offsets, blk));
}
inline void FreeTransposeBuffers(BlockMatrix& a) {
  if (a.offsets_t) detail::CudaCheck(cudaFree(a.offsets_t));
  if (a.indices_t) detail::CudaCheck(cudaFree(a.indices_t));
  if (a.block_offsets) detail::CudaCheck(cudaFree(a.block_offsets));
}
inline void AllocateRowIndicesBuffer(BlockMatrix& a) {
  const int bs = AsInt(a.block_size);
  detail::CudaCheck(cudaMalloc(&a.row_indices, size_t(a.nonzeros) / (bs * bs) * sizeof(short)));
}
inline void FreeRowIndicesBuffer(BlockMatrix& a) {
  if (a.row_indices) detail::CudaCheck(cudaFree(a.row_indices));
}

}  // namespace block
}  // namespace sputnik

#endif  // BSMM_SPUTNIK_SHIM_H_
