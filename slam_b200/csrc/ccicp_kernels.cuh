// ccicp_kernels.cuh -- sm_100a device code of the batched two-class 2-D ICP.
//
// One CTA registers one scan pair (all of its seeds) completely on chip: the
// model cloud is rounded to float (icp.cpp:51-60), binned into a per-class
// uniform grid by a single-digit radix (counting) sort on cell id in shared
// memory, and then every iteration of IcpPointToPoint::fitStep
// (icpPointToPoint.cpp:33-172) runs without leaving the SM: FP64 transform ->
// float query (a4), exact float nearest neighbour over the grid (a5), squared
// distance gate (a6), FP64 warp->block reductions of centroids and of the
// centred 2x2 cross-covariance (a7), closed-form equivalent of R_ = V*U^T (a8)
// and the compose + delta step (a9).  (aN = SURVEY.md section 8(a) rows.)
//
// Built with -fmad=false: every float/double expression below is evaluated
// un-fused, in the order written, exactly like the reference's x86-64 SSE2
// build; the __f*_rn / __d*_rn intrinsics on the parity-critical expressions
// make that explicit.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

#include "../../include/ccicp2d_b200.h"
#include "kdreplay.h"

namespace ccicp {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr unsigned kNoSlot = 0xFFFFu;

struct GridP {          // per class
  float minx, miny, inv_h, h;
  int32_t G, base, n, first;   // cells per side, first cell id, model points, first model point (class offset)
};

struct IcpArgs {
  const double* model_xy; const int32_t* model_off;
  const double* scene_xy; const int32_t* scene_off;
  const double* init_Rt;
  ccicp_result* out;
  int32_t* corr_idx;             // may be NULL
  int32_t* work_counter;
  char* scratch; size_t scratch_per_cta;   // tie-replay trees
  int32_t n_pairs, n_seeds, seeds_per_item, items_per_pair;
  int32_t max_iter; uint32_t flags;
  double min_delta, indist;
  int64_t total_scene;           // corr_idx seed stride (scene points in the whole batch)
  int32_t mcap, scap, ccap;      // shared-memory capacities: model points, scene points, cells
  int32_t max_model;             // scratch sizing
};

// cells per side for a class of n model points (0 = class never searched,
// icpPointToPoint.cpp:59,93 skip a class whose model has <= 3 points)
__host__ __device__ inline int grid_side(int n) {
  if (n <= 3) return 0;
  if (n <= 32) return 8;
  if (n <= 128) return 16;
  if (n <= 512) return 32;
  return 64;
}

__device__ __forceinline__ int f2ord(float f) {   // monotone float -> int
  int b = __float_as_int(f);
  return b ^ ((b >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float ord2f(int o) {
  return __int_as_float(o ^ ((o >> 31) & 0x7fffffff));
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- shared memory carve-up ------------------------------------------------
struct Smem {
  double2* scene;      // [scap]   (aliased during the grid build by tmp_pts / tmp_cell / tmp_rank)
  float2* mpts;        // [mcap]   model points, float, sorted by cell
  uint16_t* slots;     // [scap]   matched sorted-slot per scene point, kNoSlot = none
  uint16_t* morig;     // [mcap]   sorted slot -> per-class original model index
  uint16_t* cstart;    // [ccap+2] cell -> first sorted slot (exclusive prefix), packed counters while building
  double* red;         // [kWarps*10]
  double* rtprev;      // [8] pre-update R,t of the last successful iteration
  GridP* grid;         // [2]
  int* misc;           // [16]
};

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

__host__ __device__ inline size_t smem_bytes(int mcap, int scap, int ccap) {
  size_t region0 = (size_t)scap * 16;
  size_t need_tmp = (size_t)mcap * 12;
  if (need_tmp > region0) region0 = need_tmp;
  size_t b = align_up(region0, 16);
  b += align_up((size_t)mcap * 8, 16);
  b += align_up((size_t)scap * 2, 16);
  b += align_up((size_t)mcap * 2, 16);
  b += align_up((size_t)(ccap + 4) * 2, 16);
  b += (kWarps * 10 + 8) * 8;
  b += 2 * sizeof(GridP);
  b += 16 * 4;
  return b;
}

__device__ inline Smem carve(char* base, int mcap, int scap, int ccap) {
  Smem s;
  size_t region0 = (size_t)scap * 16;
  size_t need_tmp = (size_t)mcap * 12;
  if (need_tmp > region0) region0 = need_tmp;
  size_t o = 0;
  s.scene = (double2*)(base + o); o += align_up(region0, 16);
  s.mpts = (float2*)(base + o); o += align_up((size_t)mcap * 8, 16);
  s.slots = (uint16_t*)(base + o); o += align_up((size_t)scap * 2, 16);
  s.morig = (uint16_t*)(base + o); o += align_up((size_t)mcap * 2, 16);
  s.cstart = (uint16_t*)(base + o); o += align_up((size_t)(ccap + 4) * 2, 16);
  s.red = (double*)(base + o); o += kWarps * 10 * 8;
  s.rtprev = (double*)(base + o); o += 8 * 8;
  s.grid = (GridP*)(base + o); o += 2 * sizeof(GridP);
  s.misc = (int*)(base + o);
  return s;
}

// misc[] slots
enum { M_WORK = 0, M_BBOX = 1 /* 8 ints: per class minx,miny,maxx,maxy (ordered) */, M_TREE = 9 /* 2 */,
       M_NNODE = 11 /* 2 */, M_REPLAYS = 13, M_TREEFAIL = 14 };

// ---- exact float nearest neighbour over one class grid ---------------------
// Returns the float-argmin (identical arithmetic to kdtree.cpp:610-614:
// dis = fl(fl(dx*dx) + fl(dy*dy)), dx = fl(m - q)), its sorted slot, and
// whether >= 2 model points share that minimum.  Exact for every point whose
// dis could pass the gate: the ring walk stops only when a conservative lower
// bound on the distance of every unvisited cell exceeds min(best, gate).
__device__ __forceinline__ void scan_slots(const float2* __restrict__ mpts, int s, int e, float qx, float qy,
                                           float& best, int& bslot, int& tie) {
  for (int j = s; j < e; j++) {
    float2 p = mpts[j];
    float dx = __fsub_rn(p.x, qx);
    float dy = __fsub_rn(p.y, qy);
    float d = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
    if (d < best) { best = d; bslot = j; tie = 0; }
    else if (d == best) tie = 1;
  }
}

__device__ __forceinline__ void nn_search(const float2* __restrict__ mpts, const uint16_t* __restrict__ cstart,
                                          const GridP& g, float qx, float qy, float gate_f,
                                          float& best, int& bslot, int& tie) {
  best = CUDART_INF_F; bslot = -1; tie = 0;
  const int G = g.G;
  float ux = __fmul_rn(__fsub_rn(qx, g.minx), g.inv_h);
  float uy = __fmul_rn(__fsub_rn(qy, g.miny), g.inv_h);
  ux = fminf(fmaxf(ux, -1.0e6f), 1.0e6f);
  uy = fminf(fmaxf(uy, -1.0e6f), 1.0e6f);
  const float fxl = floorf(ux), fyl = floorf(uy);
  const int cx = (int)fxl, cy = (int)fyl;
  const float fx = ux - fxl, fy = uy - fyl;
  const float m = fminf(fminf(fx, 1.0f - fx), fminf(fy, 1.0f - fy));
  const float slop = 0.01f + 4e-7f * fmaxf(fabsf(ux), fabsf(uy));
  int r0 = max(max(-cx, cx - (G - 1)), max(-cy, cy - (G - 1)));
  r0 = max(r0, 0);
  const int rend = max(max(cx, G - 1 - cx), max(cy, G - 1 - cy));
  const uint16_t* cs = cstart + g.base;
  for (int r = r0; r <= rend; r++) {
    if (r > 0) {
      float lbc = (float)(r - 1) + m - slop;
      if (lbc > 0.0f) {
        float lb = lbc * g.h * 0.99999f;
        if (lb * lb > fminf(best, gate_f)) break;
      }
    }
    const int xa = max(cx - r, 0), xb = min(cx + r, G - 1);
    const int ya = cy - r, yb = cy + r;
    if (xa <= xb) {
      if (ya >= 0 && ya < G) scan_slots(mpts, cs[ya * G + xa], cs[ya * G + xb + 1], qx, qy, best, bslot, tie);
      if (r > 0 && yb >= 0 && yb < G) scan_slots(mpts, cs[yb * G + xa], cs[yb * G + xb + 1], qx, qy, best, bslot, tie);
    }
    if (r > 0) {
      const int y0 = max(ya + 1, 0), y1 = min(yb - 1, G - 1);
      const int xl = cx - r, xr = cx + r;
      const bool hl = xl >= 0 && xl < G, hr = xr >= 0 && xr < G;
      for (int y = y0; y <= y1; y++) {
        if (hl) scan_slots(mpts, cs[y * G + xl], cs[y * G + xl + 1], qx, qy, best, bslot, tie);
        if (hr) scan_slots(mpts, cs[y * G + xr], cs[y * G + xr + 1], qx, qy, best, bslot, tie);
      }
    }
  }
}

// a4: q = float((r00*x + r01*y) + t0), FP64 un-fused (icpPointToPoint.cpp:69-70)
__device__ __forceinline__ void transform(const double* Rt, double x, double y, float& qx, float& qy) {
  qx = __double2float_rn(__dadd_rn(__dadd_rn(__dmul_rn(Rt[0], x), __dmul_rn(Rt[1], y)), Rt[4]));
  qy = __double2float_rn(__dadd_rn(__dadd_rn(__dmul_rn(Rt[2], x), __dmul_rn(Rt[3], y)), Rt[5]));
}

}  // namespace ccicp
