// nn_grid.h -- exact float nearest neighbour over one class grid: the search routine of the pair kernel.
//
// nn_search sweeps every cell that can hold a point with dis <= min(bound, gate) and returns the float arg-min of the
// reference's distance expression (kdtree.cpp:610-614: dis = fl(fl(dx*dx) + fl(dy*dy)), dx = fl(m - q)), its sorted
// slot, and whether >= 2 distinct model points share that minimum (those queries are re-answered by the kd-tree
// replay).
//
// The file compiles as CUDA device code (ccicp_kernels.cuh; the library is built with -fmad=false) and as plain host
// C++ (tests/host/nn_grid_host.cpp, -ffp-contract=off) so that exactness is checked on the CPU against a brute-force
// arg-min over millions of queries without a GPU.  Product code, not part of oracle/.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#include <math_constants.h>
#define NN_HD __host__ __device__ __forceinline__
#define NN_D __device__ __forceinline__
#else
#include <cmath>
#include <cstring>
#define NN_HD inline
#define NN_D inline
struct float2 { float x, y; };
struct double2 { double x, y; };
#ifndef CUDART_INF_F
#define CUDART_INF_F (__builtin_inff())
#endif
// host stand-ins of the rounding intrinsics (this translation unit is compiled with -ffp-contract=off)
inline float __fmul_rn(float a, float b) { return a * b; }
inline float __fadd_rn(float a, float b) { return a + b; }
inline float __fsub_rn(float a, float b) { return a - b; }
inline int __float_as_int(float f) { int i; memcpy(&i, &f, 4); return i; }
inline float __int_as_float(int i) { float f; memcpy(&f, &i, 4); return f; }
inline float __fsqrt_ru(float x) {
  float r = sqrtf(x);
  if ((double)r * (double)r < (double)x) r = nextafterf(r, __builtin_inff());
  return r;
}
inline float __fmaf_ru(float a, float b, float c) {
  const double d = (double)a * (double)b + (double)c;
  float r = (float)d;
  if ((double)r < d) r = nextafterf(r, __builtin_inff());
  return r;
}
template <class T> inline T min(T a, T b) { return a < b ? a : b; }
template <class T> inline T max(T a, T b) { return a > b ? a : b; }
#endif

#ifndef CCICP_GBIG
#define CCICP_GBIG 64      // cells per side for classes with > 512 model points
#endif
#ifndef CCICP_GHUGE
#define CCICP_GHUGE 512    // cells per side for classes with > 8192 model points (L2 variant only; one C4 wave with the 8-slot chunk loop: 256 -> 12.7, 384 -> 12.2, 512 -> 11.9, 768 -> 12.7, 1024 -> 13.9, 1536 -> 17.0 ms; the flat loop before it preferred 1024)
#endif
#ifndef CCICP_CAND_UNROLL
#define CCICP_CAND_UNROLL 2
#endif
#ifndef CCICP_ROW_LOOPS
#define CCICP_ROW_LOOPS 4   // on-chip candidate loop: 4 = one loop over unmasked 4-slot chunks (shipped), 1 = one plain loop per row,
#endif                     // 0 = flat loop over the three ranges (the L2 variant always uses that one)
#ifndef CCICP_ROW_UNROLL
#define CCICP_ROW_UNROLL 4
#endif
#ifndef CCICP_CAND_UNROLL_DENSE
#define CCICP_CAND_UNROLL_DENSE 4
#endif

namespace ccicp {

constexpr int kCandUnroll = CCICP_CAND_UNROLL;              // candidate loop of nn_search
constexpr int kRowUnroll = CCICP_ROW_UNROLL;
#ifndef CCICP_CHUNK_MIN2
#define CCICP_CHUNK_MIN2 0         // on chip (4-slot chunks): the deferred slot look-up costs more than the tournament saves (3.03 against 2.95 ms)
#endif
#ifndef CCICP_CHUNK_MIN2_DENSE
#define CCICP_CHUNK_MIN2_DENSE 1   // large-scan variants (8-slot chunks): 13.9 against 14.2 ms per C4 wave
#endif
#ifndef CCICP_DENSE_CHUNKS
#define CCICP_DENSE_CHUNKS 1
#endif
#ifndef CCICP_ROW_CHUNK_DENSE
#define CCICP_ROW_CHUNK_DENSE 8
#endif
#ifndef CCICP_ROW_CHUNK
#define CCICP_ROW_CHUNK 4
#endif
constexpr int kRowChunkDense = CCICP_ROW_CHUNK_DENSE;     // L2 variant (one C4 wave: flat loop 16.6 ms, chunks of 4: 15.0, of 8: 14.2)
constexpr int kRowChunk = CCICP_ROW_CHUNK;                 // CCICP_ROW_LOOPS == 4: slots per chunk (2: 3.33 ms, 3: 3.13 ms, 4: 3.00 ms on C2)
constexpr int kCandUnrollDense = CCICP_CAND_UNROLL_DENSE;   // L2 variant: the candidates come from global memory, more loads in flight
constexpr unsigned kNoSlot = 0xFFFFu;

struct GridP {          // per class
  float minx, miny, inv_h, h;
  float cx, cy; int32_t pad0, pad1;   // bbox centre: pivot of the shifted sums
  int32_t G, base, n, first;   // cells per side, first cell id, model points, first model point (class offset)
};

// cells per side for a class of n model points (0 = class never searched,
// icpPointToPoint.cpp:59,93 skip a class whose model has <= 3 points)
NN_HD int grid_side(int n) {
  if (n <= 3) return 0;
  if (n <= 32) return 8;
  if (n <= 128) return 16;
  if (n <= 512) return 32;
  if (n <= 2048) return CCICP_GBIG;
  if (n <= 8192) return 128;
  return CCICP_GHUGE;
}

// cell coordinates of a model point (the grid build's binning rule)
NN_D void cell_of(const GridP& g, float x, float y, int& ix, int& iy) {
  ix = (int)__fmul_rn(__fsub_rn(x, g.minx), g.inv_h);
  iy = (int)__fmul_rn(__fsub_rn(y, g.miny), g.inv_h);
  ix = min(max(ix, 0), g.G - 1); iy = min(max(iy, 0), g.G - 1);
}

// The reference's distance (kdtree.cpp:610-614): dis = fl(fl(dx*dx) + fl(dy*dy)), dx = fl(m.x - q.x), dy = fl(m.y - q.y).
// On sm_100 the two subtractions and the two multiplications are one packed instruction each (sub/mul.rn.f32x2: IEEE
// round-to-nearest per element, subnormals kept -- the same bits as the scalar form, 3 instructions instead of 5).
struct QPack {
#if defined(__CUDA_ARCH__)
  unsigned long long v;
#else
  float x, y;
#endif
};
NN_D QPack pack_query(float qx, float qy) {
  QPack q;
#if defined(__CUDA_ARCH__)
  asm("mov.b64 %0, {%1, %2};" : "=l"(q.v) : "f"(qx), "f"(qy));
#else
  q.x = qx; q.y = qy;
#endif
  return q;
}
NN_D float dist2(const float2 p, const QPack& q) {
#if defined(__CUDA_ARCH__)
  unsigned long long pv, d;
  float sx, sy;
  asm("mov.b64 %0, {%1, %2};" : "=l"(pv) : "f"(p.x), "f"(p.y));
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(pv), "l"(q.v));
  asm("mul.rn.f32x2 %0, %1, %1;" : "=l"(d) : "l"(d));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(sx), "=f"(sy) : "l"(d));
  return __fadd_rn(sx, sy);
#else
  const float dx = __fsub_rn(p.x, q.x), dy = __fsub_rn(p.y, q.y);
  return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
#endif
}

// Upper bound of sqrt(x) for the search radius: sqrt.approx (1 ulp) -- the callers multiply by 1.00001, far above that
// error, so the IEEE round-up sequence of __fsqrt_ru (10 instructions) is not needed.  Host build: the exact form.
NN_D float sqrt_up(float x) {
#if defined(__CUDA_ARCH__)
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));   // ftz: a subnormal bound reads as 0, the slop term (>= 0.01 cell) covers it
  return r;
#else
  return __fsqrt_ru(x);
#endif
}
NN_D int floor_int(float x) {      // (int)floorf(x) as one conversion
#if defined(__CUDA_ARCH__)
  return __float2int_rd(x);
#else
  return (int)floorf(x);
#endif
}

// ---- pieces of the chunked candidate loop ----------------------------------------------------------------------------
// distances of the kCh slots from j on; slots past the class's last one (cend) read the last point and count as +inf
template <int kCh, bool kCurv>
NN_D void chunk_dists(const float2* mpts, const float* mcurv, int j, int cend, const QPack& q2, float qc, float ctol,
                      float (&d)[kCh]) {
  if (j + kCh <= cend) {
#pragma unroll
    for (int c = 0; c < kCh; c++) {
      d[c] = dist2(mpts[j + c], q2);
      if (kCurv) d[c] = fabsf(mcurv[j + c] - qc) > ctol ? CUDART_INF_F : d[c];
    }
  } else {
#pragma unroll
    for (int c = 0; c < kCh; c++) {
      const int jc = min(j + c, cend - 1);
      d[c] = dist2(mpts[jc], q2);
      if (kCurv) d[c] = fabsf(mcurv[jc] - qc) > ctol ? CUDART_INF_F : d[c];
      d[c] = j + c < cend ? d[c] : CUDART_INF_F;
    }
  }
}
// the two smallest of d[B .. B+N) (with multiplicity: lo == s iff the minimum is attained twice), by a min/max tournament:
// 7 instructions for four values and 17 for eight, where tracking (best, second, slot) per candidate costs 5 each
template <int B, int N, int kCh>
NN_D void two_smallest(const float (&d)[kCh], float& lo, float& s) {
  if constexpr (N == 1) { lo = d[B]; s = CUDART_INF_F; }
  else if constexpr (N == 2) { lo = fminf(d[B], d[B + 1]); s = fmaxf(d[B], d[B + 1]); }
  else {
    float la, sa, lb, sb;
    two_smallest<B, N / 2, kCh>(d, la, sa);
    two_smallest<B + N / 2, N - N / 2, kCh>(d, lb, sb);
    lo = fminf(la, lb);
    s = fminf(fmaxf(la, lb), fminf(sa, sb));
  }
}

// ---- exact float nearest neighbour over one class grid: row sweep -------------------------------------------------
// Row sweep in chunks of three grid rows.  An upper bound `best` comes from the previous iteration's match (warm
// start) or, failing that, from probing the home cell and then the 3x3 block.  A row is cut to the cell span that can
// still hold a point with dis <= lim = min(best, gate); cells are sorted row-major, so the span is ONE contiguous slot
// range (2 table reads).  All bounds are conservative w.r.t. the float arithmetic (slop terms): a point is skipped only
// if it provably cannot win, tie or pass the gate.
// (no __restrict__/__ldg on mpts/cstart: in the L2 variant they are global memory written earlier by this kernel)
// kDense (L2 variant, densely sampled scans): doubling probe blocks + chord pruning; the on-chip variant keeps the
// cheaper home cell / 3x3 probes and a bounding-box sweep (measured: the extra arithmetic costs 4 % on 1081-beam pairs).
// kDiag: count the distance evaluations into `evals` (diagnostics instantiation only).
// kCurv (CCICP_F_CURV_GATE, an extra without reference counterpart): a candidate whose curvature mcurv[j] differs from
// the query's `qc` by more than `ctol` is not a candidate at all (NaN = no feature: compatible with everything).
template <bool kDense, bool kDiag = false, bool kCurv = false, int kLoops = CCICP_ROW_LOOPS>
NN_D void nn_search(const float2* mpts, const uint16_t* cstart, const GridP& g, float qx, float qy, float gate_f,
                    unsigned warm, float& best, int& bslot, int& tie, int& evals, const float* mcurv = nullptr,
                    float qc = 0.0f, float ctol = 0.0f) {
  best = CUDART_INF_F; bslot = -1;
  int jbest = -1;                   // two-smallest mode: first slot of the chunk that holds the minimum
  float second = CUDART_INF_F;      // second smallest distance seen (with multiplicity): tie <=> second == best
  const int G = g.G;
  const float inv_h = g.inv_h;
  const uint16_t* cs = cstart + g.base;
  const float ux = fminf(fmaxf(__fmul_rn(__fsub_rn(qx, g.minx), inv_h), -1.0e6f), 1.0e6f);
  const float uy = fminf(fmaxf(__fmul_rn(__fsub_rn(qy, g.miny), inv_h), -1.0e6f), 1.0e6f);
  const float slop = 0.01f + 4e-7f * fmaxf(fabsf(ux), fabsf(uy));
  const QPack q2 = pack_query(qx, qy);
  const int cend = kLoops == 4 ? (int)cs[G * G] : 0;      // one past the class's last slot
  // Without a warm start an upper bound B is probed first: the home cell, then blocks of radius 1, 2, 4, ... cells
  // (r < 0 below), until a block holds a point or the gate radius is covered.  The final pass (r = -1) is the exact
  // sweep of every cell that can hold a point with dis <= min(B+, gate).  The column span of a chunk of rows is cut
  // to the chord of that disc at the chunk's distance from the query (not just to the disc's bounding box): for a
  // query far from a densely sampled wall this is the difference between hundreds of candidates and a handful.
  const float gate_cells = __fsqrt_ru(gate_f) * inv_h;
  int r = 0;
  if (warm != kNoSlot) {
    best = dist2(mpts[warm], q2);
    r = -1;
  }
  for (;;) {
    int xa, xb, ya, yb;
    float su = 0.0f;
    if (r >= 0) {
      if (best < CUDART_INF_F || (kDense ? (float)r > gate_cells : r > 1)) { r = -1; continue; }
      const int cx = floor_int(ux), cy = floor_int(uy);
      xa = cx - r; xb = cx + r; ya = cy - r; yb = cy + r;
    } else {
      // keep only the bound: restart from the next float above B so every point is tested exactly once
      best = __int_as_float((int)min((unsigned)__float_as_int(best) + 1u, 0x7f800000u));   // next float above B, +inf stays
      second = CUDART_INF_F; bslot = -1; jbest = -1;
      su = sqrt_up(fminf(best, gate_f)) * inv_h * 1.00001f + slop;   // search radius in cells, rounded up
      xa = floor_int(ux - su); xb = floor_int(ux + su);
      ya = floor_int(uy - su); yb = floor_int(uy + su);
    }
    xa = max(xa, 0); xb = min(xb, G - 1); ya = max(ya, 0); yb = min(yb, G - 1);
    int cov = 0;      // kLoops == 4: one past the last slot this pass has evaluated
    if (xa <= xb) {
      const bool wide = kDense && r < 0 && yb - ya >= 3;
      for (int y = ya; y <= yb; y += 3) {
        int xs = xa, xe = xb;
        if (!kDense && y > ya) {            // later chunks: shrink the column span to the current bound
          const float s2 = sqrt_up(fminf(best, gate_f)) * inv_h * 1.00001f + slop;
          xa = max(xa, floor_int(ux - s2)); xb = min(xb, floor_int(ux + s2));
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
        if (kDiag) evals += tot;
        // History of this loop: a flat loop over the three ranges (18.5 instructions per candidate: the range of a candidate
        // is picked per candidate), then on chip one plain unrolled loop per row (13 per candidate, but three trip counts per
        // warp and a remainder path that a 3.6-candidate row rarely leaves), now chunks (below) in both variants.
        if (kLoops == 4 && (!kDense || CCICP_DENSE_CHUNKS)) {
          constexpr int kCh = kDense ? kRowChunkDense : kRowChunk;
          // ONE loop over chunks of kCh slots covering the three rows.  A chunk that runs past the end of its row
          // evaluates the slots that follow: points of the SAME class outside the box, which provably cannot win or tie
          // (every point with dis <= bound lies inside the box), so nothing is masked -- except in the last slots of the
          // class, where the chunk would reach the other class / the end of the array (cold path, masked).
          // Rows ascend in slot order, so the slots a chunk evaluates past its row's end may be the first slots of the
          // next row (a vertical wall: nothing between them): `cov` = end of the slots already evaluated in this pass;
          // a row starts there at the earliest, so no point is evaluated twice (a duplicate of the best would read as a tie).
          const int s0 = max(j0, cov), c0 = (max(e0 - s0, 0) + kCh - 1) / kCh;
          cov = c0 ? s0 + c0 * kCh : cov;
          const int s1 = max(j1, cov), c1 = (max(e1 - s1, 0) + kCh - 1) / kCh;
          cov = c1 ? s1 + c1 * kCh : cov;
          const int s2 = max(j2, cov), c2 = (max(e2 - s2, 0) + kCh - 1) / kCh;
          cov = c2 ? s2 + c2 * kCh : cov;
          const int c01 = c0 + c1, nchunk = c01 + c2;
          const int vb1 = s1 - c0 * kCh, vb2 = s2 - c01 * kCh;
#pragma unroll 1
          for (int t = 0; t < nchunk; t++) {
            int vb = t >= c0 ? vb1 : s0;
            vb = t >= c01 ? vb2 : vb;
            const int j = vb + t * kCh;
            float d[kCh];
            chunk_dists<kCh, kCurv>(mpts, mcurv, j, cend, q2, qc, ctol, d);
            if (kDense ? CCICP_CHUNK_MIN2_DENSE : CCICP_CHUNK_MIN2) {          // per chunk: its two smallest distances; the slot is resolved after the sweep
              float lo, s2;
              two_smallest<0, kCh, kCh>(d, lo, s2);
              jbest = lo < best ? j : jbest;
              second = fminf(second, fminf(fmaxf(best, lo), s2));
              best = fminf(best, lo);
            } else {
#pragma unroll
              for (int c = 0; c < kCh; c++) {
                bslot = d[c] < best ? j + c : bslot;
                second = fminf(second, fmaxf(d[c], best));
                best = fminf(best, d[c]);
              }
            }
          }
        } else if (kLoops && !kDense) {
          // (kLoops == 1: the instantiation of multi-seed launches -- cold queries and far seeds sweep long rows, where a
          // plain unrolled loop per row is cheaper than the chunk bookkeeping: C5 867 k against 837 k problems/s)
          const int jb[3] = {j0, j1, j2}, je[3] = {e0, e1, e2};
#pragma unroll
          for (int k = 0; k < 3; k++) {
#pragma unroll kRowUnroll
            for (int j = jb[k]; j < je[k]; j++) {
              float d = dist2(mpts[j], q2);
              if (kCurv) d = fabsf(mcurv[j] - qc) > ctol ? CUDART_INF_F : d;
              bslot = d < best ? j : bslot;
              second = fminf(second, fmaxf(d, best));
              best = fminf(best, d);
            }
          }
        } else {
#pragma unroll (kDense ? kCandUnrollDense : kCandUnroll)
          for (int t = 0; t < tot; t++) {
            int off = o2;
            off = t < n01 ? o1 : off;
            off = t < n0 ? j0 : off;
            const int j = t + off;
            float d = dist2(mpts[j], q2);
            if (kCurv) d = fabsf(mcurv[j] - qc) > ctol ? CUDART_INF_F : d;
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
  if ((kDense ? CCICP_CHUNK_MIN2_DENSE : CCICP_CHUNK_MIN2) && kLoops == 4 && (!kDense || CCICP_DENSE_CHUNKS) && jbest >= 0) {
    // the chunk that lowered the bound last holds the minimum: its first slot at that distance is the answer
    constexpr int kCh = kDense ? kRowChunkDense : kRowChunk;
    float d[kCh];
    chunk_dists<kCh, kCurv>(mpts, mcurv, jbest, cend, q2, qc, ctol, d);
#pragma unroll
    for (int c = kCh - 1; c >= 0; c--) bslot = d[c] == best ? jbest + c : bslot;
  }
  tie = (bslot >= 0 && second == best) ? 1 : 0;
}

}  // namespace ccicp
