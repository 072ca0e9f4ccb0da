// lum_grid.cuh -- nearest target point of every source point through a 3-D uniform hash grid, for the clouds
// computeEdgeInformationLUM (graph_slam/src/graphSlamTools.cpp:108-214) really runs on: voxel-filtered Velodyne
// clouds of tens of thousands of points, where the reference uses pcl::KdTreeFLANN (:120, :130) and the brute-force
// sweep of lum.cuh is O(n * m).  The 3-D analogue of the pair kernel's cell table:
//
//   * a pair counts only if its squared distance is below max_corr_distance^2 (:131-133), so the grid cell is
//     h = max_corr_distance * (1 + 1e-3) and the nearest neighbour, IF it counts, lies in the 3 x 3 x 3 cells around the
//     query; a source point whose block holds nothing below the gate gets no neighbour (the brute-force path reports a
//     farther one that the gate then rejects: same information matrix);
//   * per edge a chained hash table in global memory: cell (ix, iy, iz) -> bucket  hash & (H - 1),  H a power of two
//     >= 2 x points; points are pushed with atomicExch on the bucket head (one kernel over all target points of the
//     batch), so the build needs no sort; a bucket may hold points of other cells (they are candidates like any other
//     point: every candidate's true float distance is evaluated, so extra candidates never change the arg-min);
//   * arg-min of the same un-fused float expression as lum.cuh, ties -> lowest target index (list order is irrelevant).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace lum {

struct GridArgs {
  const float* source_xyz; const int32_t* source_off;
  const float* target_xyz; const int32_t* target_off;
  const int64_t* table_off;        // [n_edges + 1] first bucket of every edge's table (sizes are powers of two)
  int32_t* head;                   // [table_off[n_edges]] bucket -> last pushed target point (edge-local index), -1 = empty
  int32_t* next;                   // [total target points] chain
  float* nn_dist; int32_t* nn_idx; // [total source points]
  float inv_h; float gate;         // 1 / cell size; max_corr_distance^2 as float (rounded up)
  int32_t n_edges;
};

__device__ __forceinline__ uint32_t cell_hash(int ix, int iy, int iz) {
  return (uint32_t)ix * 73856093u ^ (uint32_t)iy * 19349663u ^ (uint32_t)iz * 83492791u;
}
__device__ __forceinline__ int edge_of(const int32_t* off, int n_edges, int i) {      // largest e with off[e] <= i
  int lo = 0, hi = n_edges;
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (__ldg(off + mid) <= i) lo = mid; else hi = mid; }
  return lo;
}

__global__ void lum_grid_build_kernel(const GridArgs a, int n_target) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_target) return;
  const int e = edge_of(a.target_off, a.n_edges, i);
  const float3 p = ((const float3*)a.target_xyz)[i];
  if (!(isfinite(p.x) && isfinite(p.y) && isfinite(p.z))) { a.next[i] = -2; return; }
  const int ix = (int)floorf(p.x * a.inv_h), iy = (int)floorf(p.y * a.inv_h), iz = (int)floorf(p.z * a.inv_h);
  const int64_t t0 = a.table_off[e];
  const uint32_t mask = (uint32_t)(a.table_off[e + 1] - t0) - 1u;
  a.next[i] = atomicExch(&a.head[t0 + (cell_hash(ix, iy, iz) & mask)], i - __ldg(a.target_off + e));
}

__global__ void lum_grid_nn_kernel(const GridArgs a, int n_source) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_source) return;
  const int e = edge_of(a.source_off, a.n_edges, i);
  const float3 p = ((const float3*)a.source_xyz)[i];
  float best = CUDART_INF_F;
  int bidx = -1;
  const int tbase = __ldg(a.target_off + e), nt = __ldg(a.target_off + e + 1) - tbase;
  if (nt > 0 && isfinite(p.x) && isfinite(p.y) && isfinite(p.z)) {
    const int cx = (int)floorf(p.x * a.inv_h), cy = (int)floorf(p.y * a.inv_h), cz = (int)floorf(p.z * a.inv_h);
    const int64_t t0 = a.table_off[e];
    const uint32_t mask = (uint32_t)(a.table_off[e + 1] - t0) - 1u;
    const float3* tgt = (const float3*)a.target_xyz + tbase;
    const int32_t* nxt = a.next + tbase;
    for (int dz = -1; dz <= 1; dz++)
      for (int dy = -1; dy <= 1; dy++)
        for (int dx = -1; dx <= 1; dx++) {
          int j = a.head[t0 + (cell_hash(cx + dx, cy + dy, cz + dz) & mask)];
          while (j >= 0) {
            const float3 t = tgt[j];
            const float ex = __fsub_rn(p.x, t.x), ey = __fsub_rn(p.y, t.y), ez = __fsub_rn(p.z, t.z);
            const float d = __fadd_rn(__fadd_rn(__fmul_rn(ex, ex), __fmul_rn(ey, ey)), __fmul_rn(ez, ez));
            if (d < best || (d == best && j < bidx)) { best = d; bidx = j; }
            j = nxt[j];
          }
        }
    if (!(best < a.gate)) { best = CUDART_INF_F; bidx = -1; }       // beyond the gate the block is not exhaustive: report nothing
  }
  a.nn_dist[i] = best; a.nn_idx[i] = bidx;
}

}  // namespace lum
