// exact_solve.cuh -- CCICP_F_EXACT_SUMS: the reference's own evaluation order for the statistics and the solve of
// one fitStep, so that R,t come out BIT-IDENTICAL to the reference (not just within 1e-5 m / 1e-6 rad):
//   centroids  : running sums over the accepted pairs in scene order, GA block first (icpPointToPoint.cpp:80-86,
//                115-121), then mu /= inIDX (:149-150)
//   H          : q = p - ones*mu, H = ~q_t * q_m, k innermost and sequential, each product rounded before it is
//                added (Matrix::operator*, matrix.cpp:268-282; no FMA in the reference's x86-64 build)
//   R_, t_     : Numerical-Recipes svdcmp as in Matrix::svd (matrix.cpp:582-810, pythag :848-856) specialised to
//                2x2, R_ = V * ~U, t_ = mu_m - R_ * mu_t (icpPointToPoint.cpp:160-163)
//   compose    : R = R_*R, t = R_*t + t_, delta = max(l2norm(R_ - I), l2norm(t_)) (:166-171, matrix.cpp:354-360)
// Built with -fmad=false; double sqrt and division are IEEE-exact on the device, so every intermediate equals the
// host's.  The sequential sums cannot be parallelised without changing their rounding; the four independent
// accumulators of each pass run on four lanes of one warp.
#pragma once
#include <cuda_runtime.h>

namespace ccicp {
namespace exact {

__device__ __forceinline__ double nr_sign(double a, double b) { return b >= 0.0 ? fabs(a) : -fabs(a); }
__device__ __forceinline__ double nr_sqr(double a) { return a == 0.0 ? 0.0 : a * a; }
__device__ __forceinline__ double nr_pythag(double a, double b) {     // matrix.cpp:848-856
  const double absa = fabs(a), absb = fabs(b);
  if (absa > absb) return absa * sqrt(1.0 + nr_sqr(absb / absa));
  return absb == 0.0 ? 0.0 : absb * sqrt(1.0 + nr_sqr(absa / absb));
}

// Matrix::svd for a 2x2 (matrix.cpp:582-810).  H row-major in, U and V row-major out.
__device__ __noinline__ void svd2(const double* H, double* Uo, double* Vo) {
  double u00 = H[0], u01 = H[1], u10 = H[2], u11 = H[3];
  double v00 = 0.0, v01 = 0.0, v10 = 0.0, v11 = 0.0;
  double w0, w1, rv10, rv11;
  double anorm, c, f, g, h, s, scale, x, y, z;
  // ---- Householder reduction to bidiagonal form (:597-646), i = 0
  g = scale = anorm = 0.0;
  rv10 = scale * g;
  g = s = scale = 0.0;
  scale += fabs(u00); scale += fabs(u10);
  if (scale) {
    u00 /= scale; s += u00 * u00;
    u10 /= scale; s += u10 * u10;
    f = u00;
    g = -nr_sign(sqrt(s), f);
    h = f * g - s;
    u00 = f - g;
    {   // j = 1
      s = 0.0; s += u00 * u01; s += u10 * u11;
      f = s / h;
      u01 += f * u00; u11 += f * u10;
    }
    u00 *= scale; u10 *= scale;
  }
  w0 = scale * g;
  g = s = scale = 0.0;
  {   // i != n-1: row part, l = 1
    scale += fabs(u01);
    if (scale) {
      u01 /= scale; s += u01 * u01;
      f = u01;
      g = -nr_sign(sqrt(s), f);
      h = f * g - s;
      u01 = f - g;
      const double r1 = u01 / h;           // rv1[1]
      rv11 = r1;
      {   // j = 1
        s = 0.0; s += u11 * u01;
        u11 += s * r1;
      }
      u01 *= scale;
    } else {
      rv11 = 0.0;                          // placeholder, overwritten below (rv1[1] = scale*g at i = 1)
    }
  }
  { const double a2 = fabs(w0) + fabs(rv10); anorm = anorm > a2 ? anorm : a2; }
  // i = 1
  rv11 = scale * g;
  g = s = scale = 0.0;
  scale += fabs(u11);
  if (scale) {
    u11 /= scale; s += u11 * u11;
    f = u11;
    g = -nr_sign(sqrt(s), f);
    h = f * g - s;
    u11 = f - g;
    u11 *= scale;
  }
  w1 = scale * g;
  g = s = scale = 0.0;
  { const double a2 = fabs(w1) + fabs(rv11); anorm = anorm > a2 ? anorm : a2; }
  // ---- accumulation of right-hand transformations (:647-662); entering with g = 0 (set at the end of i = 1), l = 2
  v11 = 1.0;                               // i = 1
  g = rv11;
  {   // i = 0, l = 1
    if (g) {
      v10 = (u01 / u01) / g;
      s = 0.0; s += u01 * v11;
      v11 += s * v10;
    }
    v01 = v10 = 0.0;
    v00 = 1.0;
    g = rv10;
  }
  // ---- accumulation of left-hand transformations (:663-678)
  {   // i = 1, l = 2
    g = w1;
    if (g) {
      g = 1.0 / g;
      u11 *= g;
    } else {
      u11 = 0.0;
    }
    ++u11;
  }
  {   // i = 0, l = 1
    g = w0;
    u01 = 0.0;
    if (g) {
      g = 1.0 / g;
      {   // j = 1
        s = 0.0; s += u10 * u11;
        f = (s / u00) * g;
        u01 += f * u00; u11 += f * u10;
      }
      u00 *= g; u10 *= g;
    } else {
      u00 = 0.0; u10 = 0.0;
    }
    ++u00;
  }
  // ---- diagonalisation of the bidiagonal form (:679-765); arrays indexed through small helpers
  double w[2] = {w0, w1}, rv1[2] = {rv10, rv11};
  double U[2][2] = {{u00, u01}, {u10, u11}}, V[2][2] = {{v00, v01}, {v10, v11}};
  for (int k = 1; k >= 0; k--) {
    for (int its = 0; its < 30; its++) {
      int flag = 1, l, nm = 0;
      for (l = k; l >= 0; l--) {
        nm = l - 1;
        if ((double)(fabs(rv1[l]) + anorm) == anorm) { flag = 0; break; }
        if ((double)(fabs(w[nm < 0 ? 0 : nm]) + anorm) == anorm) break;    // rv1[0] == 0, so l = 0 always breaks above
      }
      if (flag) {
        c = 0.0; s = 1.0;
        for (int i = l; i <= k; i++) {
          f = s * rv1[i];
          rv1[i] = c * rv1[i];
          if ((double)(fabs(f) + anorm) == anorm) break;
          g = w[i];
          h = nr_pythag(f, g);
          w[i] = h;
          h = 1.0 / h;
          c = g * h;
          s = -f * h;
          for (int j = 0; j < 2; j++) {
            y = U[j][nm]; z = U[j][i];
            U[j][nm] = y * c + z * s;
            U[j][i] = z * c - y * s;
          }
        }
      }
      z = w[k];
      if (l == k) {
        if (z < 0.0) { w[k] = -z; V[0][k] = -V[0][k]; V[1][k] = -V[1][k]; }
        break;
      }
      // only reachable with k = 1, l = 0
      x = w[l];
      nm = k - 1;
      y = w[nm];
      g = rv1[nm];
      h = rv1[k];
      f = ((y - z) * (y + z) + (g - h) * (g + h)) / (2.0 * h * y);
      g = nr_pythag(f, 1.0);
      f = ((x - z) * (x + z) + h * ((y / (f + nr_sign(g, f))) - h)) / x;
      c = s = 1.0;
      for (int j = l; j <= nm; j++) {
        const int i = j + 1;
        g = rv1[i];
        y = w[i];
        h = s * g;
        g = c * g;
        z = nr_pythag(f, h);
        rv1[j] = z;
        c = f / z;
        s = h / z;
        f = x * c + g * s;
        g = g * c - x * s;
        h = y * s;
        y *= c;
        for (int jj = 0; jj < 2; jj++) {
          x = V[jj][j]; z = V[jj][i];
          V[jj][j] = x * c + z * s;
          V[jj][i] = z * c - x * s;
        }
        z = nr_pythag(f, h);
        w[j] = z;
        if (z) { z = 1.0 / z; c = f * z; s = h * z; }
        f = c * g + s * y;
        x = c * y - s * g;
        for (int jj = 0; jj < 2; jj++) {
          y = U[jj][j]; z = U[jj][i];
          U[jj][j] = y * c + z * s;
          U[jj][i] = z * c - y * s;
        }
      }
      rv1[l] = 0.0;
      rv1[k] = f;
      w[k] = x;
    }
  }
  // ---- sort by decreasing singular value (:767-790) and sign flips (:791-799)
  if (w[0] < w[1]) {
    for (int k = 0; k < 2; k++) {
      double a = U[k][1]; U[k][1] = U[k][0]; U[k][0] = a;
      a = V[k][1]; V[k][1] = V[k][0]; V[k][0] = a;
    }
  }
  for (int k = 0; k < 2; k++) {
    int neg = 0;
    for (int i = 0; i < 2; i++) { if (U[i][k] < 0.0) neg++; if (V[i][k] < 0.0) neg++; }
    if (neg > 2) {
      for (int i = 0; i < 2; i++) { U[i][k] = -U[i][k]; V[i][k] = -V[i][k]; }
    }
  }
  Uo[0] = U[0][0]; Uo[1] = U[0][1]; Uo[2] = U[1][0]; Uo[3] = U[1][1];
  Vo[0] = V[0][0]; Vo[1] = V[0][1]; Vo[2] = V[1][0]; Vo[3] = V[1][1];
}

}  // namespace exact
}  // namespace ccicp
