// ccicp2d/matrix.h -- header-compatible stand-in for the reference's Matrix
// (ccicp2d/include/ccicp2d/matrix.h:48-133) as far as the ICP call site needs it
// (ccicp2d/src/icpTools.cpp:168-176, 195-197): a dense row-major double matrix
// with public val / m / n, construction, copy, assignment, eye and zero.  The
// numerical methods of the reference class (svd, inv, lu ...) are not part of
// the drop-in surface: the B200 path evaluates its 2x2 / 3x3 algebra on the
// device (slam_b200/csrc/ccicp_pair_main.cuh).
#ifndef CCICP2D_B200_MATRIX_H
#define CCICP2D_B200_MATRIX_H
#include <stdint.h>
#include <cstdlib>
#include <cstring>

typedef double FLOAT;

class Matrix {
 public:
  Matrix() : val(0), m(0), n(0) {}
  Matrix(const int32_t m_, const int32_t n_) : val(0), m(0), n(0) { allocate(m_, n_); }
  Matrix(const int32_t m_, const int32_t n_, const FLOAT* v) : val(0), m(0), n(0) {
    allocate(m_, n_);
    if (m * n > 0) std::memcpy(val[0], v, sizeof(FLOAT) * (size_t)m * n);
  }
  Matrix(const Matrix& M) : val(0), m(0), n(0) { copy_from(M); }
  ~Matrix() { release(); }
  Matrix& operator=(const Matrix& M) {
    if (this != &M) { release(); copy_from(M); }
    return *this;
  }
  void zero() { if (m * n > 0) std::memset(val[0], 0, sizeof(FLOAT) * (size_t)m * n); }
  void eye() { zero(); for (int32_t i = 0; i < (m < n ? m : n); i++) val[i][i] = 1.0; }
  static Matrix eye(const int32_t k) { Matrix M(k, k); M.eye(); return M; }

  FLOAT** val;      // row pointers into one contiguous block, like the reference (matrix.cpp:829-839)
  int32_t m, n;

 private:
  void allocate(const int32_t m_, const int32_t n_) {
    m = m_ < 0 ? -m_ : m_; n = n_ < 0 ? -n_ : n_;
    if (m == 0 || n == 0) { val = 0; return; }
    val = (FLOAT**)std::malloc(sizeof(FLOAT*) * (size_t)m);
    val[0] = (FLOAT*)std::calloc((size_t)m * n, sizeof(FLOAT));
    for (int32_t i = 1; i < m; i++) val[i] = val[i - 1] + n;
  }
  void release() {
    if (val) { std::free(val[0]); std::free(val); val = 0; }
    m = n = 0;
  }
  void copy_from(const Matrix& M) {
    allocate(M.m, M.n);
    if (m * n > 0) std::memcpy(val[0], M.val[0], sizeof(FLOAT) * (size_t)m * n);
  }
};
#endif
