// features.cuh -- per-point normal and curvature of a 2-D scan (north_star extra; NO counterpart on the reference's
// built path).  The reference's only normal code is the dead ccicp2d/src/icpPointToPlane.cpp:276-348 (computeNormal:
// PCA over the 10 nearest kd-tree neighbours, normal = eigenvector of the smallest eigenvalue; not in its CMake build
// and no longer compiling against icp.h) -- it is the shape this follows, in 2-D:
//
//   neighbours of point i = the points of the same class block within `k` positions of i (a scan arrives in beam order,
//   so index neighbours are the spatial neighbours along the surface) and within `radius` of it; fewer than 3 (i
//   included) -> no feature (curvature NaN, normal 0).  C = scatter matrix of those points about their mean,
//   l1 >= l2 its eigenvalues (closed form),  curvature = l2 / (l1 + l2)  (0 on a straight wall, 0.5 isotropic),
//   normal = unit eigenvector of l2, turned towards the sensor origin.
//
// All arithmetic is FP64 in a fixed order (neighbours ascending), so oracle/curv_port.py reproduces it to the last bit
// of the float outputs.  Loads are 16-byte double2 vector loads, consecutive threads read consecutive points.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace ccicp {
namespace feat {

__device__ __forceinline__ void point_feature(const double2* __restrict__ pts, int i, int b0, int b1, int k, double r2,
                                              float& curv, float& nx, float& ny) {
  const double2 pi = __ldg(pts + i);
  const int lo = max(i - k, b0), hi = min(i + k, b1 - 1);
  double sx = 0.0, sy = 0.0;
  int n = 0;
  for (int j = lo; j <= hi; j++) {
    const double2 p = __ldg(pts + j);
    const double dx = p.x - pi.x, dy = p.y - pi.y;
    if (dx * dx + dy * dy <= r2) { sx += dx; sy += dy; n++; }
  }
  curv = CUDART_NAN_F; nx = 0.0f; ny = 0.0f;
  if (n < 3) return;
  const double mx = sx / (double)n, my = sy / (double)n;
  double cxx = 0.0, cxy = 0.0, cyy = 0.0;
  for (int j = lo; j <= hi; j++) {
    const double2 p = __ldg(pts + j);
    const double dx = p.x - pi.x, dy = p.y - pi.y;
    if (dx * dx + dy * dy <= r2) {
      const double ux = dx - mx, uy = dy - my;
      cxx += ux * ux; cxy += ux * uy; cyy += uy * uy;
    }
  }
  const double tr = cxx + cyy, hd = (cxx - cyy) * 0.5;
  const double disc = sqrt(hd * hd + cxy * cxy);
  double l2 = tr * 0.5 - disc;
  if (l2 < 0.0) l2 = 0.0;
  curv = tr > 0.0 ? (float)(l2 / tr) : 0.0f;
  // eigenvector of l2: (cxy, l2 - cxx) and (l2 - cyy, cxy) are parallel; take the longer one
  const double ax = cxy, ay = l2 - cxx, bx = l2 - cyy, by = cxy;
  const double na = ax * ax + ay * ay, nb = bx * bx + by * by;
  double vx = na >= nb ? ax : bx, vy = na >= nb ? ay : by;
  const double nv = sqrt(vx * vx + vy * vy);
  if (!(nv > 0.0)) return;                                   // isotropic scatter: no direction
  vx /= nv; vy /= nv;
  if (vx * pi.x + vy * pi.y > 0.0) { vx = -vx; vy = -vy; }   // towards the sensor (the origin of the scan frame)
  nx = (float)vx; ny = (float)vy;
}

// one thread per point; off[2*n_clouds+1]: GA block then NGA block per cloud (the CSR the registration calls take)
__global__ void scan_features_kernel(const double2* __restrict__ pts, const int32_t* __restrict__ off, int n_blocks, int n_points,
                                     int k, double r2, float2* __restrict__ normal, float* __restrict__ curvature) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_points) return;
  int lo = 0, hi = n_blocks;                  // class block of point i: largest b with off[b] <= i
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(off + mid) <= i) lo = mid; else hi = mid;
  }
  float c, nx, ny;
  point_feature(pts, i, __ldg(off + lo), __ldg(off + lo + 1), k, r2, c, nx, ny);
  curvature[i] = c;
  if (normal) normal[i] = make_float2(nx, ny);
}

}  // namespace feat
}  // namespace ccicp
