// ccicp_pair_kernel.cuh -- the persistent one-CTA-per-pair ICP kernel.
#pragma once
#include "ccicp_kernels.cuh"

namespace ccicp {

// ---- grid build: counting (single-digit radix) sort of the model on cell id -
// a1 (icp.cpp:51-60): model doubles are rounded to float once, here.
// In-place exclusive scan of n_ent uint16 counters in GLOBAL memory (L2 variant: up to 2 x 1024^2 cells).  Every warp
// owns a contiguous region and walks it in 256-entry steps, each lane loading 8 consecutive counters as one 16-byte
// vector, so both passes are fully coalesced (a thread-contiguous chunk per thread touches one 32-byte sector per
// 2-byte counter and made the 512^2 table cost 2.3 ms per pair).
template <int kThreads>
__device__ __forceinline__ void scan_cells_global(uint16_t* a, int n_ent, uint32_t* wtot, int tid) {
  constexpr int kWarps = kThreads / 32;
  const int w = tid >> 5, lane = tid & 31;
  const int per_warp = (((n_ent + kWarps - 1) / kWarps) + 255) & ~255;
  const int beg = min(w * per_warp, n_ent), end = min(beg + per_warp, n_ent);
  uint32_t total = 0;
  for (int base = beg; base < end; base += 256) {
    const int i0 = base + lane * 8;
    uint32_t sum = 0;
    if (i0 + 8 <= end) {
      const uint4 v = *(const uint4*)(a + i0);
      sum = (v.x & 0xffffu) + (v.x >> 16) + (v.y & 0xffffu) + (v.y >> 16) + (v.z & 0xffffu) + (v.z >> 16) + (v.w & 0xffffu) + (v.w >> 16);
    } else {
      for (int k = 0; k < 8; k++) if (i0 + k < end) sum += a[i0 + k];
    }
    total += sum;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
  __syncthreads();
  if (lane == 0) wtot[w] = total;
  __syncthreads();
  uint32_t carry = 0;
  for (int k = 0; k < w; k++) carry += wtot[k];
  for (int base = beg; base < end; base += 256) {
    const int i0 = base + lane * 8;
    uint32_t c[8];
    const bool vec = i0 + 8 <= end;
    if (vec) {
      const uint4 v = *(const uint4*)(a + i0);
      c[0] = v.x & 0xffffu; c[1] = v.x >> 16; c[2] = v.y & 0xffffu; c[3] = v.y >> 16;
      c[4] = v.z & 0xffffu; c[5] = v.z >> 16; c[6] = v.w & 0xffffu; c[7] = v.w >> 16;
    } else {
#pragma unroll
      for (int k = 0; k < 8; k++) c[k] = (i0 + k < end) ? a[i0 + k] : 0u;
    }
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) sum += c[k];
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    uint32_t run = carry + incl - sum;
    carry += __shfl_sync(0xffffffffu, incl, 31);
    uint32_t o8[8];
#pragma unroll
    for (int k = 0; k < 8; k++) { o8[k] = run & 0xffffu; run += c[k]; }
    if (vec) {
      uint4 v;
      v.x = o8[0] | (o8[1] << 16); v.y = o8[2] | (o8[3] << 16); v.z = o8[4] | (o8[5] << 16); v.w = o8[6] | (o8[7] << 16);
      *(uint4*)(a + i0) = v;
    } else {
#pragma unroll
      for (int k = 0; k < 8; k++) if (i0 + k < end) a[i0 + k] = (uint16_t)o8[k];
    }
  }
}

template <int kThreads, bool kGlobal>
__device__ __forceinline__ void build_grid(const IcpArgs& a, const Smem& s, int m0, int nGA, int nNGA, int tid) {
  constexpr int kWarps = kThreads / 32;
  const int nm = nGA + nNGA;
  const int G0 = grid_side(nGA), G1 = grid_side(nNGA);
  const int ncell = G0 * G0 + G1 * G1;
  float2* tmp_pts = (float2*)s.region0;
  uint32_t* tmp_cell = (uint32_t*)(tmp_pts + a.mcap);
  uint16_t* tmp_rank = (uint16_t*)(tmp_cell + a.mcap);
  uint32_t* cnt32 = (uint32_t*)s.cstart;
  for (int j = tid; j < (ncell + 4) / 2; j += kThreads) cnt32[j] = 0u;
  if (tid < 8) s.misc[M_BBOX + tid] = (tid & 2) ? INT_MIN : INT_MAX;
  if (tid < 2) { s.misc[M_TREE + tid] = 0; s.misc[M_NNODE + tid] = 0; }
  __syncthreads();

  // load + round + per-class bounding box
  int bb[8] = {INT_MAX, INT_MAX, INT_MIN, INT_MIN, INT_MAX, INT_MAX, INT_MIN, INT_MIN};
  const double2* src = (const double2*)a.model_xy + m0;
  for (int i = tid; i < nm; i += kThreads) {
    double2 p = __ldg(src + i);
    float2 f = make_float2(__double2float_rn(p.x), __double2float_rn(p.y));
    tmp_pts[i] = f;
    const int c4 = (i >= nGA) ? 4 : 0;
    const int ox = f2ord(f.x), oy = f2ord(f.y);
    bb[c4 + 0] = min(bb[c4 + 0], ox); bb[c4 + 1] = min(bb[c4 + 1], oy);
    bb[c4 + 2] = max(bb[c4 + 2], ox); bb[c4 + 3] = max(bb[c4 + 3], oy);
  }
#pragma unroll
  for (int k = 0; k < 8; k++) {
    int v = (k & 2) ? __reduce_max_sync(0xffffffffu, bb[k]) : __reduce_min_sync(0xffffffffu, bb[k]);
    if ((tid & 31) == 0) {
      if (k & 2) atomicMax(&s.misc[M_BBOX + k], v); else atomicMin(&s.misc[M_BBOX + k], v);
    }
  }
  __syncthreads();

  if (tid < 2) {
    GridP g;
    const int n = tid ? nNGA : nGA;
    g.G = tid ? G1 : G0;
    g.base = tid ? G0 * G0 : 0;
    g.n = n;
    g.first = tid ? nGA : 0;
    g.minx = g.miny = 0.0f; g.h = 1.0f; g.inv_h = 1.0f; g.cx = g.cy = 0.0f; g.pad0 = g.pad1 = 0;
    if (g.G > 0) {
      const float minx = ord2f(s.misc[M_BBOX + 4 * tid + 0]), miny = ord2f(s.misc[M_BBOX + 4 * tid + 1]);
      const float maxx = ord2f(s.misc[M_BBOX + 4 * tid + 2]), maxy = ord2f(s.misc[M_BBOX + 4 * tid + 3]);
      float ext = fmaxf(maxx - minx, maxy - miny);
      if (!(ext > 1.0e-6f)) ext = 1.0e-6f;
      g.minx = minx; g.miny = miny;
      g.cx = 0.5f * minx + 0.5f * maxx; g.cy = 0.5f * miny + 0.5f * maxy;
      g.h = ext / (float)g.G;
      g.inv_h = (float)g.G / ext;
    }
    s.grid[tid] = g;
  }
  __syncthreads();

  // histogram: two 16-bit counters per 32-bit word; the returned old value is the rank in the cell
  for (int i = tid; i < nm; i += kThreads) {
    const GridP& g = s.grid[i >= nGA];
    if (g.G == 0) continue;
    float2 f = tmp_pts[i];
    int ix, iy;
    cell_of(g, f.x, f.y, ix, iy);
    const int cell = g.base + iy * g.G + ix;
    const int sh = (cell & 1) * 16;
    uint32_t old = atomicAdd(&cnt32[cell >> 1], 1u << sh);
    tmp_cell[i] = (uint32_t)cell;
    tmp_rank[i] = (uint16_t)((old >> sh) & 0xffffu);
  }
  __syncthreads();

  // exclusive scan of the ncell+1 counters, in place
  if (kGlobal) {
    scan_cells_global<kThreads>(s.cstart, ncell + 1, (uint32_t*)s.red, tid);
  } else {
    const int n_ent = ncell + 1;
    const int ch = (n_ent + kThreads - 1) / kThreads;
    const int beg = min(tid * ch, n_ent), end = min(beg + ch, n_ent);
    uint32_t sum = 0;
    for (int j = beg; j < end; j++) sum += s.cstart[j];
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
      if ((tid & 31) >= o) incl += v;
    }
    uint32_t* wtot = (uint32_t*)s.red;
    if ((tid & 31) == 31) wtot[tid >> 5] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (int w = 0; w < kWarps; w++) woff += (w < (tid >> 5)) ? wtot[w] : 0u;
    uint32_t run = woff + incl - sum;
    for (int j = beg; j < end; j++) { uint32_t c = s.cstart[j]; s.cstart[j] = (uint16_t)run; run += c; }
  }
  __syncthreads();

  for (int i = tid; i < nm; i += kThreads) {
    const int cls = i >= nGA;
    if (s.grid[cls].G == 0) continue;
    const int slot = s.cstart[tmp_cell[i]] + tmp_rank[i];
    s.mpts[slot] = tmp_pts[i];
    s.morig[slot] = (uint16_t)(i - (cls ? nGA : 0));
    s.kinv[i] = (uint16_t)slot;
  }
  __syncthreads();
}

// ---- tie replay (rare): rebuild the reference kd-tree along the visited paths, re-answer tied queries
// The build reproduces the reference's algorithm (sequential float mean, in-order partition in its data-parallel
// form); the points are reached via kinv -> mpts, the permutation is kind, the node records go to a per-CTA global
// scratch.
struct SmemPts {
  const float2* mpts; const uint16_t* inv;
  __device__ __forceinline__ kdreplay::Pt operator()(int i) const {
    const float2 v = mpts[inv[i]];
    kdreplay::Pt p; p.x = v.x; p.y = v.y;
    return p;
  }
};

__host__ __device__ inline size_t replay_nodes_bytes(int max_model) {
  return align_up((size_t)(2 * max_model + 16) * sizeof(kdreplay::LNode), 16);
}
// per CTA: the lazily built tree nodes of both classes + the rank tables of the cooperative partition
__host__ __device__ inline size_t replay_rank_bytes(int max_model) { return align_up((size_t)(max_model + 16) * sizeof(uint16_t), 16); }
// + the global staging pair (float + u16 per model point) of splits that exceed the shared-memory staging arrays
__host__ __device__ inline size_t replay_scratch_bytes(int max_model) {
  return replay_nodes_bytes(max_model) + replay_rank_bytes(max_model) + align_up((size_t)(max_model + 16) * 6, 16);
}

// ---- cooperative tie replay: the whole CTA walks the tied queries (listed in s.tiebits) one at a time ----------
// Both kernel variants use it.  (A one-thread-per-class replay used to serve the on-chip variant: 0.1-0.3 ms for the
// first tie of a pair, which made a tie in one of the last pairs of a launch the kernel's tail.)
template <int kThreads>
struct CtaCoop {
  int tid, nt;
  float* sv; uint16_t* si; int cap; int* bcast;
  float* gsv; uint16_t* gsi; int gcap;  // global staging for ranges beyond the shared-memory arrays (large classes)
  uint16_t* rank;                     // [>= class size] global scratch: rank tables of the data-parallel partition
  float* fred;                        // [kThreads/32][8] shared
  __device__ __forceinline__ void sync() { __syncthreads(); }
  // block-wide exclusive prefix sum over tid; the total goes to every thread
  __device__ __forceinline__ int exscan(int v, int* total) {
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if ((tid & 31) >= o) incl += t; }
    int* wsum = (int*)fred;
    __syncthreads();
    if ((tid & 31) == 31) wsum[tid >> 5] = incl;
    __syncthreads();
    int woff = 0, tot = 0;
    for (int w = 0; w < kThreads / 32; w++) { const int c = wsum[w]; woff += w < (tid >> 5) ? c : 0; tot += c; }
    *total = tot;
    return woff + incl - v;
  }
  __device__ __forceinline__ void minmax4(float* lo, float* hi) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
        hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
      }
    }
    __syncthreads();
    if ((tid & 31) == 0) {
#pragma unroll
      for (int k = 0; k < 4; k++) { fred[(tid >> 5) * 8 + k] = lo[k]; fred[(tid >> 5) * 8 + 4 + k] = hi[k]; }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; k++) {
      float l = fred[k], h = fred[4 + k];
      for (int w = 1; w < kThreads / 32; w++) { l = fminf(l, fred[w * 8 + k]); h = fmaxf(h, fred[w * 8 + 4 + k]); }
      lo[k] = l; hi[k] = h;
    }
  }
};

// Inlined on purpose: as an out-of-line call its register footprint decided which of the caller's loop-carried values
// ptxas kept in local memory for their WHOLE live range -- with the data-parallel partition inside, the no-tie path of
// the C4 wave went from 17.7 to 22.4 ms; inlined it is 17.4 ms (the cold block is spilled around, not through).
template <int kThreads>
__device__ __forceinline__ void replay_ties_coop(const IcpArgs& a, const Smem& s, const double* Rt, int ns, int nsGA, int tid,
                                              char* stage, int stage_cap, float* fred, const double2* scene,
                                              uint16_t* rank_base) {
  __syncthreads();
  for (int cls = 0; cls < 2; cls++) {
    if (s.grid[cls].G == 0) continue;
    const GridP g = s.grid[cls];
    kdreplay::LNode* nodes = (kdreplay::LNode*)(a.scratch + (size_t)blockIdx.x * a.scratch_per_cta) + 2 * g.first + 8 * cls;
    const int max_nodes = 2 * g.n + 8;
    const SmemPts pts{s.mpts, s.kinv + g.first};
    uint16_t* ind = s.kind + g.first;
    CtaCoop<kThreads> co;
    co.tid = tid; co.nt = kThreads; co.cap = stage_cap;
    co.sv = (float*)stage; co.si = (uint16_t*)(stage + (size_t)stage_cap * 4);
    co.bcast = &s.misc[M_BCAST]; co.fred = fred;
    {
      char* gs = a.scratch + (size_t)blockIdx.x * a.scratch_per_cta + replay_nodes_bytes(a.max_model) + replay_rank_bytes(a.max_model);
      co.gsv = (float*)gs; co.gsi = (uint16_t*)(gs + (size_t)(a.max_model + 16) * 4); co.gcap = a.max_model + 16;
    }
    co.rank = (rank_base ? rank_base
                         : (uint16_t*)(a.scratch + (size_t)blockIdx.x * a.scratch_per_cta + replay_nodes_bytes(a.max_model))) + g.first;
    int* nn = &s.misc[M_NNODE + cls];
    int state = s.misc[M_TREE + cls];
    const int i0 = cls ? nsGA : 0, i1 = cls ? ns : nsGA;
    for (int w = i0 >> 5; w <= (i1 - 1) >> 5 && i1 > i0; w++) {
      uint32_t bits = s.tiebits[w];
      while (bits) {
        const int b = __ffs(bits) - 1;
        bits &= bits - 1;
        const int i = (w << 5) + b;
        if (i < i0 || i >= i1) continue;
        if (state == 0) state = (kdreplay::lazy_init_coop(pts, g.n, ind, nodes, nn, co) == kdreplay::OK) ? 1 : 2;
        if (state != 1) continue;                  // tree unavailable: keep the grid's answer
        const double2 sp = scene[i];
        float qx, qy;
        transform(Rt, sp.x, sp.y, qx, qy);
        int32_t o = -1; float d = 0.0f;
        const int rc = kdreplay::lazy_nearest_coop(pts, ind, nodes, max_nodes, nn, qx, qy, &o, &d, co);
        if (rc != kdreplay::OK || o < 0) { state = 2; continue; }
        if (tid == 0) { s.slots[i] = s.kinv[g.first + o]; s.misc[M_REPLAYS] += 1; }
      }
    }
    __syncthreads();
    if (tid == 0) {
      s.misc[M_TREE + cls] = state;
      if (state == 2) s.misc[M_TREEFAIL] |= 1;
    }
  }
  __syncthreads();
  for (int w = tid; w <= (ns - 1) >> 5; w += kThreads) s.tiebits[w] = 0u;
  __syncthreads();
}

// Shifted one-pass statistics (a7).  With pivot c (centre of the model's bounding box) every thread
// accumulates  S_m = sum(m - c), S_t = sum(q - c), P = sum (q - c)(m - c)^T ; then
//   mu_m = c + S_m/n, mu_t = c + S_t/n, H = P - n (S_t/n)(S_m/n)^T
// which equals the reference's centred H = sum (q - mu_t)(m - mu_m)^T (icpPointToPoint.cpp:149-159) in exact
// arithmetic and differs from its sequential FP64 evaluation only by rounding (~1e-16 relative).
__device__ __forceinline__ void accum_corr(double (&v)[8], const float2 m, float qx, float qy, double pcx, double pcy) {
  const double m0 = (double)m.x - pcx, m1 = (double)m.y - pcy;
  const double t0 = (double)qx - pcx, t1 = (double)qy - pcy;
  v[0] += m0; v[1] += m1; v[2] += t0; v[3] += t1;
  v[4] += t0 * m0; v[5] += t0 * m1; v[6] += t1 * m0; v[7] += t1 * m1;
}

// partial sums recomputed from slots[] (used after a tie replay)
template <int kThreads>
__device__ __forceinline__ void sums_from_slots(const Smem& s, const double* Rt, int ns, int tid, double pcx, double pcy,
                                                double (&v)[8], int& cnt) {
#pragma unroll
  for (int k = 0; k < 8; k++) v[k] = 0.0;
  cnt = 0;
  for (int i = tid; i < ns; i += kThreads) {
    const unsigned slot = s.slots[i];
    if (slot == kNoSlot) continue;
    double2 sp = s.scene[i];
    float qx, qy;
    transform(Rt, sp.x, sp.y, qx, qy);
    accum_corr(v, s.mpts[slot], qx, qy, pcx, pcy);
    cnt++;
  }
}

}  // namespace ccicp
