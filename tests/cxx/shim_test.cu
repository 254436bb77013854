// A caller written against the reference's C++ interface (`#include "sputnik/sputnik.h"`,
// namespace sputnik::block), in the shape of the reference's own tests
// (sputnik/block/dsd/dsd_test.cu:155-200, sdd_test.cu:148-200, dds_test.cu:155-200), compiled
// against include/sputnik/sputnik.h and linked with libbsmm.so.
//
//   shim_test abi   no GPU needed: overload resolution, argument records, error mapping
//   shim_test gpu   runs SDD / DSD (n and t, Matmul and MatmulEx) / DDS on the device through the
//                   shim and checks them against a plain host loop (fp16, the reference's type)
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include <cuda_fp16.h>

#include "sputnik/sputnik.h"

using namespace sputnik::block;

#define EXPECT(cond)                                                          \
  do {                                                                        \
    if (!(cond)) {                                                            \
      std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);  \
      return 1;                                                               \
    }                                                                         \
  } while (0)

static int AbiOnly() {
  EXPECT(AsInt(BlockSize::k128) == 128 && AsInt(BlockSize::kNone) == 0);
  EXPECT(AsBlockSize(64) == BlockSize::k64);
  void* fake = reinterpret_cast<void*>(0x10000);
  BlockMatrix a(256, 384, BlockSize::k128, 3 * 128 * 128, fake, fake, fake);
  EXPECT(a.create_metadata && a.offsets_t == nullptr && a.row_indices == nullptr);
  BlockMatrix c(256, 384, BlockSize::k128, 3 * 128 * 128, fake, fake, fake, fake);
  EXPECT(c.row_indices == fake);
  BlockMatrix t(256, 384, BlockSize::k128, 3 * 128 * 128, fake, fake, fake, fake, fake, fake);
  EXPECT(t.block_offsets == fake && t.bitmask == nullptr);
  Matrix b(384, 72, fake), out(256, 72, fake);
  MatmulShape s(a, false, b, false);
  EXPECT(s.m == 256 && s.n == 72 && s.k == 384 && s.lda == 384 && s.ldb == 72 && s.ldc == 72);
  EXPECT(ValidMatmul(a, false, b, false, out));
  EXPECT(!ValidMatmul(a, true, b, false, out));
  // Block sizes other than 128: cudaErrorNotSupported (sdd.cu:11-14, dsd.cu:14, dds.cu:12).
  BlockMatrix a64(256, 384, BlockSize::k64, 3 * 128 * 128, fake, fake, fake);
  EXPECT(Matmul(a64, false, b, false, out, nullptr) == cudaErrorNotSupported);
  EXPECT(MatmulEx(a64, false, b, false, out, nullptr) == cudaErrorNotSupported);
  Matrix lhs(72, 256, fake), out2(72, 384, fake);
  EXPECT(Matmul(lhs, false, a64, false, out2, nullptr) == cudaErrorNotSupported);
  BlockMatrix c64(256, 384, BlockSize::k64, 3 * 128 * 128, fake, fake, fake, fake);
  EXPECT(Matmul(Matrix(256, 64, fake), false, Matrix(64, 384, fake), false, c64, nullptr) ==
         cudaErrorNotSupported);
  // Misaligned problem (k % 8 != 0): the launcher's can_implement failure.
  EXPECT(Matmul(a, false, Matrix(384, 70, fake), false, Matrix(256, 70, fake), nullptr) ==
         cudaErrorNotSupported);
  // bf16 extension overloads resolve.
  EXPECT(Matmul(a64, false, b, false, out, nullptr, DType::kBfloat16) == cudaErrorNotSupported);
  std::printf("shim abi ok\n");
  return 0;
}

// ---------------------------------------------------------------------------- device part
struct Lcg {
  uint64_t s;
  float next() {  // U(-1, 1)
    s = s * 6364136223846793005ull + 1442695040888963407ull;
    return float((s >> 40) & 0xffffff) / float(1 << 23) - 1.0f;
  }
};

template <class T>
static T* ToDevice(const std::vector<T>& h) {
  T* d = nullptr;
  cudaMalloc(&d, h.size() * sizeof(T) + 16);
  cudaMemcpy(d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice);
  return d;
}

static std::vector<float> Round(std::vector<float> v) {
  for (auto& x : v) x = __half2float(__float2half(x));
  return v;
}
static std::vector<__half> Half(const std::vector<float>& v) {
  std::vector<__half> h(v.size());
  for (size_t i = 0; i < v.size(); ++i) h[i] = __float2half(v[i]);
  return h;
}

static int CheckNear(const std::vector<__half>& got, const std::vector<double>& want, const char* what) {
  double worst = 0;
  for (size_t i = 0; i < want.size(); ++i) {
    const double g = __half2float(got[i]);
    if (std::isnan(g)) { std::fprintf(stderr, "%s: NaN at %zu\n", what, i); return 1; }
    worst = std::fmax(worst, std::fabs(g - want[i]));
  }
  std::printf("%-28s max abs err %.4f\n", what, worst);
  return worst <= 5e-2 ? 0 : 1;  // the reference's tolerance (dsd_test.cu:192)
}

static int Gpu() {
  // 3 x 4 block grid, unordered indices in row 0, empty row 1 (cases the reference's tests
  // exercise through random topologies).
  const int br = 3, bc = 4, B = 128, n = 72;
  std::vector<int> offsets = {0, 3, 3, 5};
  std::vector<short> indices = {2, 0, 3, 1, 2};
  const int nb = 5, M = br * B, K = bc * B;
  Lcg rng{42};
  std::vector<float> vals(size_t(nb) * B * B), dense(size_t(K) * n), denseT(size_t(M) * n);
  for (auto& x : vals) x = rng.next();
  for (auto& x : dense) x = rng.next();
  for (auto& x : denseT) x = rng.next();
  vals = Round(vals); dense = Round(dense); denseT = Round(denseT);
  std::vector<float> full(size_t(M) * K, 0.f);  // densified A
  for (int r = 0; r < br; ++r)
    for (int j = offsets[r]; j < offsets[r + 1]; ++j)
      for (int y = 0; y < B; ++y)
        for (int x = 0; x < B; ++x)
          full[size_t(r * B + y) * K + indices[j] * B + x] = vals[(size_t(j) * B + y) * B + x];

  __half* d_vals = ToDevice(Half(vals));
  int* d_off = ToDevice(offsets);
  short* d_idx = ToDevice(indices);
  __half* d_dense = ToDevice(Half(dense));
  __half* d_denseT = ToDevice(Half(denseT));
  std::vector<__half> nanfill(size_t(std::max(M, K)) * std::max(n, K), __float2half(NAN));
  int rc = 0;

  BlockMatrix a(M, K, BlockSize::k128, nb * B * B, d_vals, d_off, d_idx);
  AllocateTransposeBuffers(a);
  AllocateRowIndicesBuffer(a);
  EXPECT(RowIndices(a, static_cast<short*>(a.row_indices), nullptr) == cudaSuccess);
  std::vector<short> rows(nb);
  cudaMemcpy(rows.data(), a.row_indices, nb * sizeof(short), cudaMemcpyDeviceToHost);
  const short want_rows[5] = {0, 0, 0, 2, 2};
  EXPECT(std::memcmp(rows.data(), want_rows, sizeof(want_rows)) == 0);

  {  // DSD nn
    __half* d_c = ToDevice(nanfill);
    EXPECT(Matmul(a, false, Matrix(K, n, d_dense), false, Matrix(M, n, d_c), nullptr) == cudaSuccess);
    std::vector<__half> got(size_t(M) * n);
    cudaMemcpy(got.data(), d_c, got.size() * 2, cudaMemcpyDeviceToHost);
    std::vector<double> want(size_t(M) * n, 0.0);
    for (int i = 0; i < M; ++i)
      for (int l = 0; l < K; ++l) {
        const float av = full[size_t(i) * K + l];
        if (av != 0.f)
          for (int j = 0; j < n; ++j) want[size_t(i) * n + j] += double(av * dense[size_t(l) * n + j]);
      }
    rc |= CheckNear(got, want, "DSD nn (Matmul)");
    // extension: fused accumulate through the shim's epilogue argument (C += A B, so C = 2 A B)
    EXPECT(Matmul(a, false, Matrix(K, n, d_dense), false, Matrix(M, n, d_c), nullptr, DType::kHalf,
                  BSMM_FLAG_ACCUMULATE) == cudaSuccess);
    cudaMemcpy(got.data(), d_c, got.size() * 2, cudaMemcpyDeviceToHost);
    std::vector<double> want2(want);
    for (auto& x : want2) x *= 2.0;
    rc |= CheckNear(got, want2, "DSD nn (+= via epilogue)");
    cudaFree(d_c);
  }
  {  // DSD tn: Matmul builds the transposed metadata (create_metadata), MatmulEx then reuses it
    std::vector<double> want(size_t(K) * n, 0.0);
    for (int l = 0; l < M; ++l)
      for (int i = 0; i < K; ++i) {
        const float av = full[size_t(l) * K + i];
        if (av != 0.f)
          for (int j = 0; j < n; ++j) want[size_t(i) * n + j] += double(av * denseT[size_t(l) * n + j]);
      }
    for (int ex = 0; ex < 2; ++ex) {
      __half* d_c = ToDevice(nanfill);
      cudaError_t e = ex ? MatmulEx(a, true, Matrix(M, n, d_denseT), false, Matrix(K, n, d_c), nullptr)
                         : Matmul(a, true, Matrix(M, n, d_denseT), false, Matrix(K, n, d_c), nullptr);
      EXPECT(e == cudaSuccess);
      std::vector<__half> got(size_t(K) * n);
      cudaMemcpy(got.data(), d_c, got.size() * 2, cudaMemcpyDeviceToHost);
      rc |= CheckNear(got, want, ex ? "DSD tn (MatmulEx)" : "DSD tn (Matmul)");
      cudaFree(d_c);
    }
  }
  {  // DDS nn: C (n x K) = lhs (n x M) * A — walks A by block column (transposed metadata)
    std::vector<float> lhs(size_t(n) * M);
    for (auto& x : lhs) x = rng.next();
    lhs = Round(lhs);
    __half* d_lhs = ToDevice(Half(lhs));
    __half* d_c = ToDevice(nanfill);
    EXPECT(Matmul(Matrix(n, M, d_lhs), false, a, false, Matrix(n, K, d_c), nullptr) == cudaSuccess);
    std::vector<__half> got(size_t(n) * K);
    cudaMemcpy(got.data(), d_c, got.size() * 2, cudaMemcpyDeviceToHost);
    std::vector<double> want(size_t(n) * K, 0.0);
    for (int i = 0; i < n; ++i)
      for (int l = 0; l < M; ++l)
        for (int j = 0; j < K; ++j)
          want[size_t(i) * K + j] += double(lhs[size_t(i) * M + l] * full[size_t(l) * K + j]);
    rc |= CheckNear(got, want, "DDS nn (Matmul)");
    cudaFree(d_c); cudaFree(d_lhs);
  }
  {  // SDD nt: C(topology of a) = X (M x 64) * Y^T, Y stored K x 64
    const int kk = 64;
    std::vector<float> x(size_t(M) * kk), y(size_t(K) * kk);
    for (auto& v : x) v = rng.next();
    for (auto& v : y) v = rng.next();
    x = Round(x); y = Round(y);
    __half* d_x = ToDevice(Half(x));
    __half* d_y = ToDevice(Half(y));
    __half* d_c = ToDevice(std::vector<__half>(size_t(nb) * B * B, __float2half(NAN)));
    BlockMatrix c(M, K, BlockSize::k128, nb * B * B, d_c, d_off, d_idx, a.row_indices);
    EXPECT(Matmul(Matrix(M, kk, d_x), false, Matrix(K, kk, d_y), true, c, nullptr) == cudaSuccess);
    std::vector<__half> got(size_t(nb) * B * B);
    cudaMemcpy(got.data(), d_c, got.size() * 2, cudaMemcpyDeviceToHost);
    std::vector<double> want(got.size(), 0.0);
    for (int r = 0; r < br; ++r)
      for (int j = offsets[r]; j < offsets[r + 1]; ++j)
        for (int yy = 0; yy < B; ++yy)
          for (int xx = 0; xx < B; ++xx) {
            double acc = 0;
            for (int l = 0; l < kk; ++l)
              acc += double(x[size_t(r * B + yy) * kk + l] * y[size_t(indices[j] * B + xx) * kk + l]);
            want[(size_t(j) * B + yy) * B + xx] = acc;
          }
    rc |= CheckNear(got, want, "SDD nt (Matmul)");
    cudaFree(d_x); cudaFree(d_y); cudaFree(d_c);
  }
  {  // DSS tn (sibling): C (K x K, dense) = A^T * A, both operands walked by block column;
     // Bitmask of the transposed view (the reference's DSS needs it; ours only offers it)
    __half* d_c = ToDevice(std::vector<__half>(size_t(K) * K, __float2half(NAN)));
    EXPECT(Matmul(a, true, a, false, Matrix(K, K, d_c), nullptr) == cudaSuccess);
    std::vector<__half> got(size_t(K) * K);
    cudaMemcpy(got.data(), d_c, got.size() * 2, cudaMemcpyDeviceToHost);
    std::vector<double> want(got.size(), 0.0);
    for (int l = 0; l < M; ++l)
      for (int i = 0; i < K; ++i) {
        const float ai = full[size_t(l) * K + i];
        if (ai == 0.f) continue;
        for (int j = 0; j < K; ++j) want[size_t(i) * K + j] += double(ai * full[size_t(l) * K + j]);
      }
    rc |= CheckNear(got, want, "DSS tn (Matmul)");
    EXPECT(MatmulEx(a, true, a, false, Matrix(K, K, d_c), nullptr) == cudaSuccess);
    cudaMemcpy(got.data(), d_c, got.size() * 2, cudaMemcpyDeviceToHost);
    rc |= CheckNear(got, want, "DSS tn (MatmulEx)");
    AllocateBitmaskBuffers(a);
    EXPECT(Bitmask(a, nullptr) == cudaSuccess);
    unsigned long long mask[4] = {0, 0, 0, 0};  // transposed view: 4 walk rows x 1 word
    cudaMemcpy(mask, a.bitmask, sizeof(mask), cudaMemcpyDeviceToHost);
    const unsigned long long want_mask[4] = {1ull, 4ull, 5ull, 1ull};  // cols 0..3 -> block rows
    EXPECT(std::memcmp(mask, want_mask, sizeof(mask)) == 0);
    FreeBitmaskBuffers(a);
    cudaFree(d_c);
  }
  {  // AnalyzeTopology: the 3x4 test topology has no twins; a 2x2 dense one has all three hints
    EXPECT(AnalyzeTopology(a, nullptr) == cudaSuccess);
    EXPECT(!a.twin_rows && !a.twin_cols && !a.even_rows);
    std::vector<int> o2 = {0, 2, 4};
    std::vector<short> i2 = {0, 1, 0, 1};
    int* d_o2 = ToDevice(o2);
    short* d_i2 = ToDevice(i2);
    BlockMatrix t(2 * B, 2 * B, BlockSize::k128, 4 * B * B, nullptr, d_o2, d_i2);
    AllocateTransposeBuffers(t);
    EXPECT(Transpose(t, nullptr) == cudaSuccess);
    EXPECT(AnalyzeTopology(t, nullptr) == cudaSuccess);
    EXPECT(t.twin_rows && t.twin_cols && t.even_rows);
    FreeTransposeBuffers(t);
    cudaFree(d_o2); cudaFree(d_i2);
  }
  EXPECT(cudaDeviceSynchronize() == cudaSuccess);
  FreeTransposeBuffers(a);
  FreeRowIndicesBuffer(a);
  if (rc == 0) std::printf("shim gpu ok\n");
  return rc;
}

int main(int argc, char** argv) {
  if (argc > 1 && std::strcmp(argv[1], "gpu") == 0) return Gpu();
  return AbiOnly();
}
