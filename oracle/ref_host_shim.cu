// ref_host_shim.cu — extern "C" doorway into the reference's HOST test utilities
// (sputnik::Matrix::operator*, T(), Mul(): sputnik/matrix_utils.h:364-391, matrix_utils.cu:595-643;
// sputnik::block::ToMatrix: sputnik/block/matrix_utils.h:81-112).  TEST INFRASTRUCTURE ONLY;
// contains no reference code, only calls into it.  Built by `make -C oracle ref_host`.
#include <cstring>

#include "sputnik/matrix_utils.h"

extern "C" {

// out[m x n] = a[m x k] * b[k x n] through sputnik::Matrix::operator*.
void ref_host_matmul(const float* a, const float* b, float* out, int m, int k, int n) {
  sputnik::Matrix ma(m, k), mb(k, n);
  std::memcpy(ma.Values(), a, sizeof(float) * size_t(m) * k);
  std::memcpy(mb.Values(), b, sizeof(float) * size_t(k) * n);
  sputnik::Matrix mc = ma * mb;
  std::memcpy(out, mc.Values(), sizeof(float) * size_t(m) * n);
}

// out[cols x rows] = in[rows x cols].T() through sputnik::Matrix::T().
void ref_host_transpose(const float* in, float* out, int rows, int cols) {
  sputnik::Matrix mi(rows, cols);
  std::memcpy(mi.Values(), in, sizeof(float) * size_t(rows) * cols);
  sputnik::Matrix mo = mi.T();
  std::memcpy(out, mo.Values(), sizeof(float) * size_t(rows) * cols);
}

// inout *= other through sputnik::Matrix::Mul().
void ref_host_mul(float* inout, const float* other, int rows, int cols) {
  sputnik::Matrix a(rows, cols), b(rows, cols);
  std::memcpy(a.Values(), inout, sizeof(float) * size_t(rows) * cols);
  std::memcpy(b.Values(), other, sizeof(float) * size_t(rows) * cols);
  a.Mul(b);
  std::memcpy(inout, a.Values(), sizeof(float) * size_t(rows) * cols);
}

}  // extern "C"
