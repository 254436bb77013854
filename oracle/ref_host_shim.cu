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

// ---- handle-based doorway (bench.py --impl reference): operands are converted to
// sputnik::Matrix ONCE, outside the timed loop; the timed calls are the reference's own
// operator* / T() / Mul() and nothing else.
void* ref_host_new(const float* values, int rows, int cols) {
  auto* m = new sputnik::Matrix(rows, cols);
  if (values) std::memcpy(m->Values(), values, sizeof(float) * size_t(rows) * cols);
  return m;
}
void ref_host_free(void* m) { delete static_cast<sputnik::Matrix*>(m); }
int ref_host_rows(void* m) { return static_cast<sputnik::Matrix*>(m)->Rows(); }
int ref_host_cols(void* m) { return static_cast<sputnik::Matrix*>(m)->Columns(); }
void ref_host_get(void* m, float* out) {
  auto* p = static_cast<sputnik::Matrix*>(m);
  std::memcpy(out, p->Values(), sizeof(float) * size_t(p->Rows()) * p->Columns());
}
// new handle = a * b (Matrix::operator*, matrix_utils.h:376-391)
void* ref_host_mm(void* a, void* b) {
  return new sputnik::Matrix(*static_cast<sputnik::Matrix*>(a) * *static_cast<sputnik::Matrix*>(b));
}
// new handle = a.T() (matrix_utils.cu:635-643)
void* ref_host_t(void* a) { return new sputnik::Matrix(static_cast<sputnik::Matrix*>(a)->T()); }
// a *= b elementwise (Matrix::Mul, the SDD mask of sdd_test.cu:170-171,193)
void ref_host_mul_inplace(void* a, void* b) {
  static_cast<sputnik::Matrix*>(a)->Mul(*static_cast<sputnik::Matrix*>(b));
}

}  // extern "C"
