// The reference's dMoE-shaped DSD benchmark (sputnik/block/dsd/dsd_benchmark.cu:18-118: random
// operands, warm-up, timed back-to-back launches, nonzero-block flops) written against the
// reference's C++ interface and run through include/sputnik/sputnik.h + libbsmm.so.  fp16 (the
// reference's type), prebuilt transposed metadata (MatmulEx), NO topology hints: what an
// unmodified caller of the reference gets after relinking.  `shim_bench [experts rows_per_expert
// ffn hidden iters]`; defaults = BASELINE C2's shape (8 x 1024 rows, ffn 4096, hidden 1024).
#include <cstdio>
#include <cstdlib>
#include <vector>

#include <cuda_fp16.h>

#include "sputnik/sputnik.h"

using namespace sputnik::block;

#define CK(x)                                                                   \
  do {                                                                          \
    cudaError_t e_ = (x);                                                       \
    if (e_ != cudaSuccess) {                                                    \
      std::fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); \
      return 1;                                                                 \
    }                                                                           \
  } while (0)

int main(int argc, char** argv) {
  const int experts = argc > 1 ? std::atoi(argv[1]) : 8;
  const int rows_pe = argc > 2 ? std::atoi(argv[2]) : 1024;
  const int ffn = argc > 3 ? std::atoi(argv[3]) : 4096;
  const int hidden = argc > 4 ? std::atoi(argv[4]) : 1024;
  const int iters = argc > 5 ? std::atoi(argv[5]) : 200;
  const int B = 128, M = experts * rows_pe, K = experts * ffn, N = hidden;
  const int br = M / B, fc = ffn / B, nb = br * fc;
  std::vector<int> offsets(br + 1);
  std::vector<short> indices(nb);
  for (int r = 0; r <= br; ++r) offsets[r] = r * fc;
  for (int r = 0; r < br; ++r)
    for (int j = 0; j < fc; ++j) indices[r * fc + j] = short((r / (rows_pe / B)) * fc + j);
  std::vector<__half> hv(size_t(nb) * B * B), hb(size_t(K) * N);
  unsigned s = 12345u;
  auto rnd = [&] { s = s * 1664525u + 1013904223u; return __float2half((float(s >> 8) / 8388608.0f) - 1.0f); };
  for (auto& x : hv) x = rnd();
  for (auto& x : hb) x = rnd();
  __half *dv, *db, *dc;
  int* doff;
  short* didx;
  CK(cudaMalloc(&dv, hv.size() * 2)); CK(cudaMalloc(&db, hb.size() * 2));
  CK(cudaMalloc(&dc, size_t(M) * N * 2));
  CK(cudaMalloc(&doff, offsets.size() * 4)); CK(cudaMalloc(&didx, indices.size() * 2));
  CK(cudaMemcpy(dv, hv.data(), hv.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(doff, offsets.data(), offsets.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(didx, indices.data(), indices.size() * 2, cudaMemcpyHostToDevice));

  BlockMatrix a(M, K, BlockSize::k128, nb * B * B, dv, doff, didx);
  AllocateTransposeBuffers(a);
  CK(Transpose(a, nullptr));
  Matrix b(K, N, db), c(M, N, dc);
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const double flops = 2.0 * nb * B * B * double(N);
  for (int mode = 0; mode < 2; ++mode) {
    if (mode == 1) CK(AnalyzeTopology(a, nullptr));  // one call per topology: hints filled in
    for (int i = 0; i < 10; ++i) CK(MatmulEx(a, false, b, false, c, nullptr));
    CK(cudaEventRecord(e0));
    for (int i = 0; i < iters; ++i) CK(MatmulEx(a, false, b, false, c, nullptr));
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    std::printf("%-34s M=%d K=%d N=%d nnz_blocks=%d: %8.2f us  %8.1f TFLOP/s\n",
                mode ? "MatmulEx after AnalyzeTopology()" : "MatmulEx, unmodified caller", M, K, N,
                nb, ms * 1e3 / iters, flops / (ms * 1e-3 / iters) / 1e12);
  }
  FreeTransposeBuffers(a);
  cudaFree(dv); cudaFree(db); cudaFree(dc); cudaFree(doff); cudaFree(didx);
  return 0;
}
