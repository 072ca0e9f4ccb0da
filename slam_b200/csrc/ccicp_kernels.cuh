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

#include "nn_grid.h"

namespace ccicp {

// Two instantiations of the same pair kernel (template <kThreads, kGlobal>):
//   on-chip : 256 threads, the pair's working set in dynamic shared memory (scans up to ~6000 points per cloud)
//   L2      : 1024 threads (64 registers), one CTA per SM, the working set in a per-CTA global workspace that stays L2/L1-resident
//             (scans up to CCICP_MAX_POINTS; config C4 = 32768 points per cloud)
#ifndef CCICP_THREADS_ONCHIP
#define CCICP_THREADS_ONCHIP 256
#endif
constexpr int kThreadsOnChip = CCICP_THREADS_ONCHIP;
#ifndef CCICP_L2_THREADS
#define CCICP_L2_THREADS 1024
#endif
constexpr int kThreadsL2 = CCICP_L2_THREADS;
constexpr int kRedStride = 10;     // doubles per warp in the reduction scratch (8 sums + 2 ints)

struct IcpArgs {
  const double* model_xy; const int32_t* model_off;
  const double* scene_xy; const int32_t* scene_off;
  const double* init_Rt;
  ccicp_result* out;
  int32_t* corr_idx;             // may be NULL
  int32_t* work_counter;
  char* scratch; size_t scratch_per_cta;   // tie-replay trees
  int32_t stage_cap;             // L2 variant: elements of the tie replay's shared-memory staging arrays
  char* workspace; size_t workspace_per_cta;   // L2 variant: the carved working set of every resident CTA
  int32_t n_pairs, n_seeds, seeds_per_item, items_per_pair;
  int32_t pair_base;             // this launch registers pairs [pair_base, pair_base + n_pairs)
  int32_t max_iter; uint32_t flags;
  double min_delta, indist;
  double trim_frac;              // CCICP_F_TRIM: fraction of the gated correspondences kept per iteration
  double curv_tol, feat_radius; int32_t feat_k, pad_feat;   // CCICP_F_CURV_GATE: |curvature difference| allowed, feature neighbourhood
  int64_t total_scene;           // corr_idx seed stride (scene points in the whole batch)
  int32_t mcap, scap, ccap;      // shared-memory capacities: model points, scene points, cells
  int32_t max_model;             // scratch sizing
  const int32_t* model_id; const int32_t* scene_id;      // scan-table form: pair p registers scan scene_id[p] against scan model_id[p]
                                                         // (both index the same offsets / xy arrays); NULL = pair p's own blocks
  const int32_t* item_list; const int32_t* item_count;   // exact-order re-run: listed problem ids (pair * n_seeds + seed) and their count
  unsigned long long* diag;      // diagnostics instantiation: {distance evaluations, searched queries, accepted pairs}
};

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

// ---- working-set carve-up (dynamic shared memory, or the CTA's global workspace) ---
struct Smem {
  const double2* scene;// [scap]   scene points (on-chip: a copy in shared memory; L2 variant: the input array itself)
  float2* mpts;        // [mcap]   model points, float, sorted by cell
  uint16_t* slots;     // [scap]   matched sorted-slot per scene point, kNoSlot = none
  uint16_t* morig;     // [mcap]   sorted slot -> per-class original model index
  uint16_t* kinv;      // [mcap]   original model index (GA block then NGA block) -> sorted slot
  uint16_t* kind;      // [mcap]   tie replay: the reference kd-tree's index permutation (built lazily)
  uint16_t* cstart;    // [ccap+2] cell -> first sorted slot (exclusive prefix), packed counters while building
  uint32_t* tiebits;   // [scap/32] scene points whose float minimum is shared (tie replay work list)
  char* region0;       // scene copy (on-chip) aliased with the sort temporaries tmp_pts / tmp_cell / tmp_rank
  double* red;         // [2][kWarps*kRedStride] double-buffered by iteration parity   (always shared memory)
  double* rtprev;      // [8] pre-update R,t of the last successful iteration          (always shared memory)
  GridP* grid;         // [2]                                                          (always shared memory)
  int* misc;           // [M_COUNT]                                                         (always shared memory)
};

__host__ __device__ inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// bytes of the carved working set; global_ws = L2 variant (no scene copy)
__host__ __device__ inline size_t workset_bytes(int mcap, int scap, int ccap, bool global_ws) {
  size_t region0 = global_ws ? 0 : (size_t)scap * 16;
  size_t need_tmp = (size_t)mcap * 14;
  if (need_tmp > region0) region0 = need_tmp;
  size_t b = align_up(region0, 16);
  b += align_up((size_t)mcap * 8, 16);
  b += align_up((size_t)scap * 2, 16);
  b += 3 * align_up((size_t)mcap * 2, 16);
  b += align_up((size_t)(ccap + 4) * 2, 16);
  b += align_up((size_t)(scap / 32 + 1) * 4, 16);
  return align_up(b, 128);
}

template <bool kGlobal>
__device__ __forceinline__ Smem carve(char* base, int mcap, int scap, int ccap, double* red, double* rtprev, GridP* grid,
                                      int* misc) {
  Smem s;
  size_t region0 = kGlobal ? 0 : (size_t)scap * 16;
  size_t need_tmp = (size_t)mcap * 14;
  if (need_tmp > region0) region0 = need_tmp;
  size_t o = 0;
  s.region0 = base + o; s.scene = (const double2*)(base + o); o += align_up(region0, 16);
  s.mpts = (float2*)(base + o); o += align_up((size_t)mcap * 8, 16);
  s.slots = (uint16_t*)(base + o); o += align_up((size_t)scap * 2, 16);
  s.morig = (uint16_t*)(base + o); o += align_up((size_t)mcap * 2, 16);
  s.kinv = (uint16_t*)(base + o); o += align_up((size_t)mcap * 2, 16);
  s.kind = (uint16_t*)(base + o); o += align_up((size_t)mcap * 2, 16);
  s.cstart = (uint16_t*)(base + o); o += align_up((size_t)(ccap + 4) * 2, 16);
  s.tiebits = (uint32_t*)(base + o);
  s.red = red; s.rtprev = rtprev; s.grid = grid; s.misc = misc;
  if (!kGlobal) {
    // tell the compiler these are shared-memory addresses so it emits LDS/STS, not generic LD/ST
    __builtin_assume(__isShared(s.scene)); __builtin_assume(__isShared(s.region0)); __builtin_assume(__isShared(s.mpts));
    __builtin_assume(__isShared(s.slots)); __builtin_assume(__isShared(s.morig)); __builtin_assume(__isShared(s.kinv));
    __builtin_assume(__isShared(s.kind)); __builtin_assume(__isShared(s.cstart)); __builtin_assume(__isShared(s.tiebits));
  }
  __builtin_assume(__isShared(s.red)); __builtin_assume(__isShared(s.rtprev)); __builtin_assume(__isShared(s.grid));
  __builtin_assume(__isShared(s.misc));
  return s;
}

// misc[] slots
enum { M_WORK = 0, M_BBOX = 1 /* 8 ints: per class minx,miny,maxx,maxy (ordered) */, M_TREE = 9 /* 2 */,
       M_NNODE = 11 /* 2 */, M_REPLAYS = 13, M_TREEFAIL = 14, M_BCAST = 15 /* tie replay */,
       M_TIEFLAG = 16 /* 2: by iteration parity */, M_COUNT = 24 };

// a4: q = float((r00*x + r01*y) + t0), FP64 un-fused (icpPointToPoint.cpp:69-70)
__device__ __forceinline__ void transform(const double* Rt, double x, double y, float& qx, float& qy) {
  qx = __double2float_rn(__dadd_rn(__dadd_rn(__dmul_rn(Rt[0], x), __dmul_rn(Rt[1], y)), Rt[4]));
  qy = __double2float_rn(__dadd_rn(__dadd_rn(__dmul_rn(Rt[2], x), __dmul_rn(Rt[3], y)), Rt[5]));
}

}  // namespace ccicp
