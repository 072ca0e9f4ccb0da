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

#ifndef CCICP_GBIG
#define CCICP_GBIG 64      // cells per side for classes with > 512 model points
#endif

namespace ccicp {

// Two instantiations of the same pair kernel (template <kThreads, kGlobal>):
//   on-chip : 256 threads, the pair's working set in dynamic shared memory (scans up to ~6000 points per cloud)
//   L2      : 1024 threads (64 registers), one CTA per SM, the working set in a per-CTA global workspace that stays L2/L1-resident
//             (scans up to CCICP_MAX_POINTS; config C4 = 32768 points per cloud)
constexpr int kThreadsOnChip = 256;
#ifndef CCICP_L2_THREADS
#define CCICP_L2_THREADS 1024
#endif
constexpr int kThreadsL2 = CCICP_L2_THREADS;
#ifndef CCICP_CAND_UNROLL
#define CCICP_CAND_UNROLL 2
#endif
constexpr int kCandUnroll = CCICP_CAND_UNROLL;   // candidate loop of nn_search
#ifndef CCICP_ROW_LOOPS
#define CCICP_ROW_LOOPS 1
#endif
#ifndef CCICP_ROW_UNROLL
#define CCICP_ROW_UNROLL 4
#endif
constexpr int kRowUnroll = CCICP_ROW_UNROLL;
#ifndef CCICP_CAND_UNROLL_DENSE
#define CCICP_CAND_UNROLL_DENSE 4
#endif
constexpr int kCandUnrollDense = CCICP_CAND_UNROLL_DENSE;   // L2 variant: the candidates come from global memory, more loads in flight
constexpr int kRedStride = 10;     // doubles per warp in the reduction scratch (8 sums + 2 ints)
constexpr unsigned kNoSlot = 0xFFFFu;

struct GridP {          // per class
  float minx, miny, inv_h, h;
  float cx, cy; int32_t pad0, pad1;   // bbox centre: pivot of the shifted sums
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
  int32_t stage_cap;             // L2 variant: elements of the tie replay's shared-memory staging arrays
  char* workspace; size_t workspace_per_cta;   // L2 variant: the carved working set of every resident CTA
  int32_t n_pairs, n_seeds, seeds_per_item, items_per_pair;
  int32_t pair_base;             // this launch registers pairs [pair_base, pair_base + n_pairs)
  int32_t max_iter; uint32_t flags;
  double min_delta, indist;
  double trim_frac;              // CCICP_F_TRIM: fraction of the gated correspondences kept per iteration
  int64_t total_scene;           // corr_idx seed stride (scene points in the whole batch)
  int32_t mcap, scap, ccap;      // shared-memory capacities: model points, scene points, cells
  int32_t max_model;             // scratch sizing
};

// cells per side for a class of n model points (0 = class never searched,
// icpPointToPoint.cpp:59,93 skip a class whose model has <= 3 points)
#ifndef CCICP_GHUGE
#define CCICP_GHUGE 1024   // cells per side for classes with > 8192 model points (L2 variant only; C4 wave: 256 -> 30.2, 512 -> 25.4, 1024 -> 23.7 ms)
#endif
__host__ __device__ inline int grid_side(int n) {
  if (n <= 3) return 0;
  if (n <= 32) return 8;
  if (n <= 128) return 16;
  if (n <= 512) return 32;
  if (n <= 2048) return CCICP_GBIG;
  if (n <= 8192) return 128;
  return CCICP_GHUGE;
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
  int* misc;           // [16]                                                         (always shared memory)
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
       M_NNODE = 11 /* 2 */, M_REPLAYS = 13, M_TREEFAIL = 14 };

// ---- exact float nearest neighbour over one class grid ---------------------
// Returns the float-argmin (identical arithmetic to kdtree.cpp:610-614:
// dis = fl(fl(dx*dx) + fl(dy*dy)), dx = fl(m - q)), its sorted slot, and
// whether >= 2 distinct model points share that minimum.
//
// Row sweep in chunks of three grid rows.  An upper bound `best` comes from the
// previous iteration's match (warm start) or, failing that, from probing the
// home cell and then the 3x3 block.  A row is cut to the cell span that can
// still hold a point with dis <= lim = min(best, gate); cells are sorted
// row-major, so the span is ONE contiguous slot range (2 table reads).  The three
// ranges of a chunk are walked by one branch-free loop, so the lanes of a warp
// stay converged.  All bounds are conservative w.r.t. the float arithmetic
// (slop terms): a point is skipped only if it provably cannot win, tie or pass
// the gate.
struct NNQuery {
  float qx, qy, ux, uy, slop, h, inv_h;
  int G;
  const uint16_t* cs;
};

// (no __restrict__/__ldg on mpts/cstart: in the L2 variant they are global memory written earlier by this kernel)
// kDense (L2 variant, densely sampled scans): doubling probe blocks + chord pruning; the on-chip variant keeps the
// cheaper home cell / 3x3 probes and a bounding-box sweep (measured: the extra arithmetic costs 4 % on 1081-beam pairs).
template <bool kDense>
__device__ __forceinline__ void nn_search(const float2* mpts, const uint16_t* cstart,
                                          const GridP& g, float qx, float qy, float gate_f, unsigned warm,
                                          float& best, int& bslot, int& tie) {
  best = CUDART_INF_F; bslot = -1;
  float second = CUDART_INF_F;      // second smallest distance seen (with multiplicity): tie <=> second == best
  const int G = g.G;
  const float inv_h = g.inv_h;
  const uint16_t* cs = cstart + g.base;
  const float ux = fminf(fmaxf(__fmul_rn(__fsub_rn(qx, g.minx), inv_h), -1.0e6f), 1.0e6f);
  const float uy = fminf(fmaxf(__fmul_rn(__fsub_rn(qy, g.miny), inv_h), -1.0e6f), 1.0e6f);
  const float slop = 0.01f + 4e-7f * fmaxf(fabsf(ux), fabsf(uy));
  // Without a warm start an upper bound B is probed first: the home cell, then blocks of radius 1, 2, 4, ... cells
  // (r < 0 below), until a block holds a point or the gate radius is covered.  The final pass (r = -1) is the exact
  // sweep of every cell that can hold a point with dis <= min(B+, gate).  The column span of a chunk of rows is cut
  // to the chord of that disc at the chunk's distance from the query (not just to the disc's bounding box): for a
  // query far from a densely sampled wall this is the difference between hundreds of candidates and a handful.
  const float gate_cells = __fsqrt_ru(gate_f) * inv_h;
  int r = 0;
  if (warm != kNoSlot) {
    const float2 p = mpts[warm];
    const float dx = __fsub_rn(p.x, qx), dy = __fsub_rn(p.y, qy);
    best = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
    r = -1;
  }
  for (;;) {
    int xa, xb, ya, yb;
    float su = 0.0f;
    if (r >= 0) {
      if (best < CUDART_INF_F || (kDense ? (float)r > gate_cells : r > 1)) { r = -1; continue; }
      const int cx = (int)floorf(ux), cy = (int)floorf(uy);
      xa = cx - r; xb = cx + r; ya = cy - r; yb = cy + r;
    } else {
      // keep only the bound: restart from the next float above B so every point is tested exactly once
      best = __int_as_float(__float_as_int(best) + (best < CUDART_INF_F ? 1 : 0));
      second = CUDART_INF_F; bslot = -1;
      su = __fsqrt_ru(fminf(best, gate_f)) * inv_h * 1.00001f + slop;   // search radius in cells, rounded up
      xa = (int)floorf(ux - su); xb = (int)floorf(ux + su);
      ya = (int)floorf(uy - su); yb = (int)floorf(uy + su);
    }
    xa = max(xa, 0); xb = min(xb, G - 1); ya = max(ya, 0); yb = min(yb, G - 1);
    if (xa <= xb) {
      const bool wide = kDense && r < 0 && yb - ya >= 3;
      for (int y = ya; y <= yb; y += 3) {
        int xs = xa, xe = xb;
        if (!kDense && y > ya) {            // later chunks: shrink the column span to the current bound
          const float s2 = __fsqrt_ru(fminf(best, gate_f)) * inv_h * 1.00001f + slop;
          xa = max(xa, (int)floorf(ux - s2)); xb = min(xb, (int)floorf(ux + s2));
          if (xa > xb) break;
          xs = xa; xe = xb;
        }
        if (wide) {                         // chord of the current disc over rows [y, y+3)
          su = __fsqrt_ru(fminf(best, gate_f)) * inv_h * 1.00001f + slop;
          const float dyc = fmaxf(fmaxf((float)y - uy, uy - (float)(y + 3)), 0.0f) - slop;
          const float dyc0 = fmaxf(dyc, 0.0f);
          const float rem = __fmaf_ru(-dyc0, dyc0 * 0.99999f, su * su);
          if (rem < 0.0f) continue;
          const float sx = __fsqrt_ru(rem) * 1.00001f + slop;
          xs = max(xa, (int)floorf(ux - sx)); xe = min(xb, (int)floorf(ux + sx));
          if (xs > xe) continue;
        }
        const uint16_t* r0 = cs + y * G;
        const int j0 = r0[xs], e0 = r0[xe + 1];
        int j1 = 0, e1 = 0, j2 = 0, e2 = 0;
        if (y + 1 <= yb) { j1 = r0[G + xs]; e1 = r0[G + xe + 1]; }
        if (y + 2 <= yb) { j2 = r0[2 * G + xs]; e2 = r0[2 * G + xe + 1]; }
        const int n0 = e0 - j0, n01 = n0 + (e1 - j1), tot = n01 + (e2 - j2);
        const int o1 = j1 - n0, o2 = j2 - n01;
        // On chip: one plain loop per row -- no index mapping (13 instead of 18.5 instructions per candidate) but three
        // trip counts per warp (+22 % trips): +3.5 % on C2.  The L2 variant keeps the flat loop over the three ranges (all
        // lanes of a warp run the same number of trips, more loads in flight): row loops cost it 13 %.
        if constexpr (CCICP_ROW_LOOPS && !kDense) {
          const int jb[3] = {j0, j1, j2}, je[3] = {e0, e1, e2};
#pragma unroll
          for (int k = 0; k < 3; k++) {
#pragma unroll kRowUnroll
            for (int j = jb[k]; j < je[k]; j++) {
              const float2 p = mpts[j];
              const float dx = __fsub_rn(p.x, qx), dy = __fsub_rn(p.y, qy);
              const float d = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
              bslot = d < best ? j : bslot;
              second = fminf(second, fmaxf(d, best));
              best = fminf(best, d);
            }
          }
        } else {
#if CCICP_TWO_RANGE
        if (n01 == tot) {                     // at most two rows hold candidates: one select less per candidate
#pragma unroll (kDense ? kCandUnrollDense : kCandUnroll)
          for (int t = 0; t < tot; t++) {
            const int j = t + (t < n0 ? j0 : o1);
            const float2 p = mpts[j];
            const float dx = __fsub_rn(p.x, qx), dy = __fsub_rn(p.y, qy);
            const float d = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
            bslot = d < best ? j : bslot;
            second = fminf(second, fmaxf(d, best));
            best = fminf(best, d);
          }
          continue;
        }
#endif
#pragma unroll (kDense ? kCandUnrollDense : kCandUnroll)
        for (int t = 0; t < tot; t++) {
          int off = o2;
          off = t < n01 ? o1 : off;
          off = t < n0 ? j0 : off;
          const int j = t + off;
          const float2 p = mpts[j];
          const float dx = __fsub_rn(p.x, qx), dy = __fsub_rn(p.y, qy);
          const float d = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
          bslot = d < best ? j : bslot;
          second = fminf(second, fmaxf(d, best));
          best = fminf(best, d);
        }
        }
      }
    }
    if (r < 0) break;
    r = r == 0 ? 1 : 2 * r;
  }
  tie = (bslot >= 0 && second == best) ? 1 : 0;
}

// a4: q = float((r00*x + r01*y) + t0), FP64 un-fused (icpPointToPoint.cpp:69-70)
__device__ __forceinline__ void transform(const double* Rt, double x, double y, float& qx, float& qy) {
  qx = __double2float_rn(__dadd_rn(__dadd_rn(__dmul_rn(Rt[0], x), __dmul_rn(Rt[1], y)), Rt[4]));
  qy = __double2float_rn(__dadd_rn(__dadd_rn(__dmul_rn(Rt[2], x), __dmul_rn(Rt[3], y)), Rt[5]));
}

}  // namespace ccicp
