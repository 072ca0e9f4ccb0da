// lum.cuh -- 3-D Lu-Milios edge information, batched: graph_slam's computeEdgeInformationLUM
// (graph_slam/src/graphSlamTools.cpp:108-214), the step after registration in graph_slam.
//
// One CTA per edge.  For every source point (already transformed by the registration result) the nearest target
// point is found by an exact brute-force sweep over shared-memory tiles of the target cloud (float squared distance
// ((dx*dx)+dy*dy)+dz*dz like FLANN's L2_Simple, ties -> lowest target index); pairs closer than max_corr_distance
// give aver = 0.5f*(s+t), diff = s-t in float (Eigen::Vector3f, :130-137).  The 15 sums behind M'M and M'Z are
// FP64 block reductions of the reference's float products (:150-171, formulas kept literally), thread 0 solves
// D = MM^-1 * MZ (Gauss-Jordan with partial pivoting; the reference uses Eigen's 6x6 inverse), a second pass
// accumulates s^2 (:187-193) and the result is MM * (1.0f / ss), or the identity when ss < 1e-13 / not finite
// (:196-201).  The voxel-filtered clouds this runs on (0.5 m leaves, graphSlamTools.cpp:262-281) have a few thousand
// points, so the n*m sweep (~25 M distance evaluations per edge) is a few microseconds of FP32 work per SM.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace lum {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kTile = 2048;            // target points per shared-memory tile (24 KB as SoA)
constexpr int kSums = 15;

struct Args {
  const float* source_xyz; const int32_t* source_off;
  const float* target_xyz; const int32_t* target_off;
  float* nn_dist; int32_t* nn_idx;     // scratch [total source points]
  double* info; int32_t* num_corr; uint32_t* status;
  double max_corr_distance;
  int32_t n_edges;
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <int N>
__device__ __forceinline__ void block_sum(double (&v)[N], double* red, int tid) {
#pragma unroll
  for (int k = 0; k < N; k++) v[k] = warp_sum(v[k]);
  __syncthreads();
  if ((tid & 31) == 0) {
#pragma unroll
    for (int k = 0; k < N; k++) red[(tid >> 5) * N + k] = v[k];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < N; k++) {
    double t = 0.0;
    for (int w = 0; w < kWarps; w++) t += red[w * N + k];
    v[k] = t;
  }
}

// 6x6 inverse times vector: solves MM * D = MZ by Gauss-Jordan with partial pivoting.  Returns false if singular.
__device__ inline bool solve6(const double* MM, const double* MZ, double* D) {
  double A[6][7];
  for (int i = 0; i < 6; i++) { for (int j = 0; j < 6; j++) A[i][j] = MM[i * 6 + j]; A[i][6] = MZ[i]; }
  for (int c = 0; c < 6; c++) {
    int piv = c;
    double best = fabs(A[c][c]);
    for (int r = c + 1; r < 6; r++) if (fabs(A[r][c]) > best) { best = fabs(A[r][c]); piv = r; }
    if (!(best > 0.0)) return false;
    if (piv != c) for (int j = 0; j < 7; j++) { const double t = A[c][j]; A[c][j] = A[piv][j]; A[piv][j] = t; }
    const double inv = 1.0 / A[c][c];
    for (int j = 0; j < 7; j++) A[c][j] *= inv;
    for (int r = 0; r < 6; r++) {
      if (r == c) continue;
      const double f = A[r][c];
      if (f != 0.0) for (int j = 0; j < 7; j++) A[r][j] -= f * A[c][j];
    }
  }
  for (int i = 0; i < 6; i++) D[i] = A[i][6];
  return true;
}

// un-fused float ((dx*dx + dy*dy) + dz*dz), the order the sums of squares are taken in throughout this file
__device__ __forceinline__ float dist3(const float3& p, float x, float y, float z) {
  const float dx = __fsub_rn(p.x, x), dy = __fsub_rn(p.y, y), dz = __fsub_rn(p.z, z);
  return __fadd_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)), __fmul_rn(dz, dz));
}
// lowest index in the group of four starting at g whose distance equals the group minimum
__device__ __forceinline__ int first_equal(const float3& p, const float* tx, const float* ty, const float* tz, int g, float best) {
#pragma unroll
  for (int k = 0; k < 3; k++)
    if (dist3(p, tx[g + k], ty[g + k], tz[g + k]) == best) return g + k;
  return g + 3;
}

// kHaveNN: nn_dist / nn_idx were filled by the hash-grid search (lum_grid.cuh); the brute-force sweep is skipped.
template <bool kHaveNN>
__global__ void __launch_bounds__(kThreads) lum_edge_kernel(const Args a) {
  __shared__ __align__(16) float tx[kTile], ty[kTile], tz[kTile];
  __shared__ double red[kWarps * kSums];
  __shared__ double Dsh[6];
  __shared__ int flag_sh;
  const int tid = threadIdx.x;
  for (int e = blockIdx.x; e < a.n_edges; e += gridDim.x) {
    const int s0 = a.source_off[e], ns = a.source_off[e + 1] - s0;
    const int t0 = a.target_off[e], nt = a.target_off[e + 1] - t0;
    const float3* src = (const float3*)a.source_xyz + s0;
    // ---- exact nearest target point of every source point
    if (!kHaveNN) for (int i = tid; i < ns; i += kThreads) { a.nn_dist[s0 + i] = CUDART_INF_F; a.nn_idx[s0 + i] = -1; }
    for (int tb = 0; tb < (kHaveNN ? 0 : nt); tb += kTile) {
      const int m = min(kTile, nt - tb);
      __syncthreads();
      for (int j = tid; j < m; j += kThreads) {
        tx[j] = a.target_xyz[3 * (size_t)(t0 + tb + j)];
        ty[j] = a.target_xyz[3 * (size_t)(t0 + tb + j) + 1];
        tz[j] = a.target_xyz[3 * (size_t)(t0 + tb + j) + 2];
      }
      const int m4 = (m + 3) & ~3;               // pad the last group of four: distance +inf, never selected
      if (tid < m4 - m) { tx[m + tid] = CUDART_INF_F; ty[m + tid] = CUDART_INF_F; tz[m + tid] = CUDART_INF_F; }
      __syncthreads();
      // Two source points per thread share every target load; per group of four targets only the group minimum is
      // compared with the incumbent (3 FMNMX + 1 compare instead of 4 compare + 8 select), the index inside the winning
      // group is resolved once per tile by recomputing its four distances (bit-identical: same un-fused operations).
      for (int i = tid; i < ns; i += 2 * kThreads) {
        const int i1 = i + kThreads;
        const bool has1 = i1 < ns;
        const float3 p0 = src[i], p1 = has1 ? src[i1] : make_float3(CUDART_NAN_F, CUDART_NAN_F, CUDART_NAN_F);
        float best0 = a.nn_dist[s0 + i], best1 = has1 ? a.nn_dist[s0 + i1] : 0.0f;
        int g0 = -1, g1 = -1;
#pragma unroll 2
        for (int j = 0; j < m4; j += 4) {
          const float4 X = *reinterpret_cast<const float4*>(&tx[j]), Y = *reinterpret_cast<const float4*>(&ty[j]),
                       Z = *reinterpret_cast<const float4*>(&tz[j]);
          const float a0 = dist3(p0, X.x, Y.x, Z.x), a1 = dist3(p0, X.y, Y.y, Z.y), a2 = dist3(p0, X.z, Y.z, Z.z),
                      a3 = dist3(p0, X.w, Y.w, Z.w);
          const float b0 = dist3(p1, X.x, Y.x, Z.x), b1 = dist3(p1, X.y, Y.y, Z.y), b2 = dist3(p1, X.z, Y.z, Z.z),
                      b3 = dist3(p1, X.w, Y.w, Z.w);
          const float ga = fminf(fminf(a0, a1), fminf(a2, a3)), gb = fminf(fminf(b0, b1), fminf(b2, b3));
          if (ga < best0) { best0 = ga; g0 = j; }
          if (gb < best1) { best1 = gb; g1 = j; }
        }
        if (g0 >= 0) { a.nn_dist[s0 + i] = best0; a.nn_idx[s0 + i] = tb + first_equal(p0, tx, ty, tz, g0, best0); }
        if (g1 >= 0) { a.nn_dist[s0 + i1] = best1; a.nn_idx[s0 + i1] = tb + first_equal(p1, tx, ty, tz, g1, best1); }
      }
    }
    // ---- sums of M'M and M'Z (graphSlamTools.cpp:150-171), float products widened to double
    const double gate = a.max_corr_distance * a.max_corr_distance;
    double v[kSums];
#pragma unroll
    for (int k = 0; k < kSums; k++) v[k] = 0.0;
    int cnt = 0;
    for (int i = tid; i < ns; i += kThreads) {
      const int bi = a.nn_idx[s0 + i];
      if (bi < 0 || !((double)a.nn_dist[s0 + i] < gate)) continue;
      const float3 s = src[i];
      const float3 t = ((const float3*)a.target_xyz)[t0 + bi];
      const float ax = __fmul_rn(0.5f, __fadd_rn(s.x, t.x)), ay = __fmul_rn(0.5f, __fadd_rn(s.y, t.y)), az = __fmul_rn(0.5f, __fadd_rn(s.z, t.z));
      const float dx = __fsub_rn(s.x, t.x), dy = __fsub_rn(s.y, t.y), dz = __fsub_rn(s.z, t.z);
      cnt++;
      v[0] += (double)ax; v[1] += (double)ay; v[2] += (double)az;
      v[3] += (double)__fmul_rn(ax, az);                                    // MM(3,4) -=
      v[4] += (double)__fmul_rn(ax, ay);                                    // MM(3,5) -=
      v[5] += (double)__fmul_rn(ay, az);                                    // MM(4,5) -=
      v[6] += (double)__fadd_rn(__fmul_rn(ay, ay), __fmul_rn(az, az));      // MM(3,3) +=
      v[7] += (double)__fadd_rn(__fmul_rn(ax, ax), __fmul_rn(ay, ay));      // MM(4,4) +=
      v[8] += (double)__fadd_rn(__fmul_rn(ax, ax), __fmul_rn(az, az));      // MM(5,5) +=
      v[9] += (double)dx; v[10] += (double)dy; v[11] += (double)dz;         // MZ(0..2)
      v[12] += (double)__fsub_rn(__fmul_rn(ay, dz), __fmul_rn(az, dy));     // MZ(3)
      v[13] += (double)__fsub_rn(__fmul_rn(ax, dy), __fmul_rn(ay, dx));     // MZ(4)
      v[14] += (double)__fsub_rn(__fmul_rn(az, dx), __fmul_rn(ax, dz));     // MZ(5)
    }
    double cntd[1] = {(double)cnt};
    block_sum<kSums>(v, red, tid);
    block_sum<1>(cntd, red, tid);
    const int num = (int)cntd[0];
    double MM[36];
#pragma unroll
    for (int k = 0; k < 36; k++) MM[k] = 0.0;
    MM[0 * 6 + 4] = -v[1]; MM[0 * 6 + 5] = v[2]; MM[1 * 6 + 3] = -v[2]; MM[1 * 6 + 4] = v[0]; MM[2 * 6 + 3] = v[1]; MM[2 * 6 + 5] = -v[0];
    MM[3 * 6 + 4] = -v[3]; MM[3 * 6 + 5] = -v[4]; MM[4 * 6 + 5] = -v[5];
    MM[3 * 6 + 3] = v[6]; MM[4 * 6 + 4] = v[7]; MM[5 * 6 + 5] = v[8];
    MM[0] = MM[7] = MM[14] = (double)(float)num;                            // :174
    MM[4 * 6 + 0] = MM[0 * 6 + 4]; MM[5 * 6 + 0] = MM[0 * 6 + 5]; MM[3 * 6 + 1] = MM[1 * 6 + 3]; MM[4 * 6 + 1] = MM[1 * 6 + 4];
    MM[3 * 6 + 2] = MM[2 * 6 + 3]; MM[5 * 6 + 2] = MM[2 * 6 + 5]; MM[4 * 6 + 3] = MM[3 * 6 + 4]; MM[5 * 6 + 3] = MM[3 * 6 + 5];
    MM[5 * 6 + 4] = MM[4 * 6 + 5];
    if (tid == 0) {
      const double MZ[6] = {v[9], v[10], v[11], v[12], v[13], v[14]};
      double D[6] = {0, 0, 0, 0, 0, 0};
      flag_sh = solve6(MM, MZ, D) ? 1 : 0;
      for (int k = 0; k < 6; k++) Dsh[k] = D[k];
    }
    __syncthreads();
    const bool solved = flag_sh != 0;
    double D[6];
#pragma unroll
    for (int k = 0; k < 6; k++) D[k] = Dsh[k];
    // ---- s^2 (:187-193): each pair contributes float(sum of three double squares)
    double ss[1] = {0.0};
    for (int i = tid; i < ns; i += kThreads) {
      const int bi = a.nn_idx[s0 + i];
      if (bi < 0 || !((double)a.nn_dist[s0 + i] < gate)) continue;
      const float3 s = src[i];
      const float3 t = ((const float3*)a.target_xyz)[t0 + bi];
      const float ax = __fmul_rn(0.5f, __fadd_rn(s.x, t.x)), ay = __fmul_rn(0.5f, __fadd_rn(s.y, t.y)), az = __fmul_rn(0.5f, __fadd_rn(s.z, t.z));
      const float dx = __fsub_rn(s.x, t.x), dy = __fsub_rn(s.y, t.y), dz = __fsub_rn(s.z, t.z);
      const double e0 = (double)dx - (D[0] + (double)az * D[5] - (double)ay * D[4]);
      const double e1 = (double)dy - (D[1] + (double)ax * D[4] - (double)az * D[3]);
      const double e2 = (double)dz - (D[2] + (double)ay * D[3] - (double)ax * D[5]);
      ss[0] += (double)(float)(e0 * e0 + e1 * e1 + e2 * e2);
    }
    block_sum<1>(ss, red, tid);
    const float ssf = (float)ss[0];
    const bool bad = !solved || ssf < 0.0000000000001f || !isfinite(ssf);          // :196
    if (tid < 36) {
      double o;
      if (bad) o = (tid / 6 == tid % 6) ? 1.0 : 0.0;
      else o = MM[tid] * (double)(1.0f / ssf);                                       // :204
      a.info[(size_t)e * 36 + tid] = o;
    }
    if (tid == 0) { a.num_corr[e] = num; a.status[e] = bad ? 1u : 0u; }
    __syncthreads();
  }
}

}  // namespace lum
