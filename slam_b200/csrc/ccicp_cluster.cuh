// ccicp_cluster.cuh -- one thread-block CLUSTER per scan pair, for pairs whose working set does not fit one SM's shared
// memory (config C4: 32768 + 32768 points; the reference's own size class, up to 20000 points per class,
// ccicp2d/include/ccicp2d/icpTools.h:21).
//
// Same algorithm and the same arithmetic as icp_pair_body (ccicp_pair_main.cuh): grid build by a counting sort on cell
// id, per iteration transform -> exact float NN (nn_search) -> gate -> shifted FP64 sums -> closed-form solve, kd-tree
// replay for float ties.  What changes is who does it:
//   * kCluster CTAs (1024 threads each) share ONE pair: every loop over model / scene points is strided over the whole
//     cluster, so a pair gets kCluster SMs instead of one and only ~148 / kCluster pairs are in flight;
//   * the pair's working set (float model sorted by cell, cell table, slots, index maps) lives in a per-cluster global
//     workspace: ~35 working sets of 1.9 MB stay L2-resident (126 MB), where the one-CTA-per-pair variant's 148 did not
//     (4.7 GB of DRAM reads per 148-pair wave for 155 MB of input);
//   * the FP64 reduction spans the cluster: every CTA reduces its warps' partials in shared memory, publishes ONE
//     partial, barrier.cluster, then one warp of every CTA reads the kCluster partials through distributed shared memory
//     in rank order and the CTA solves redundantly (identical bits in every CTA, no broadcast).
// The tie replay runs on rank 0 (it is CTA-cooperative) while the other ranks wait at the cluster barrier.
#pragma once
#include <cooperative_groups.h>

#include "ccicp_pair_main.cuh"

namespace ccicp {

#ifndef CCICP_CLUSTER
#define CCICP_CLUSTER 4
#endif
constexpr int kCluster = CCICP_CLUSTER;
constexpr int kThreadsCl = 1024;
namespace cg = cooperative_groups;

// ints at the head of the per-cluster workspace (global memory, shared by the CTAs of the cluster)
enum { CH_BBOX = 0 /* 8 */, CH_TREE = 8 /* 2 */, CH_NNODE = 10 /* 2 */, CH_REPLAYS = 12, CH_TREEFAIL = 13, CH_COUNT = 32 };

struct ClPartial {            // what one CTA publishes per reduction (read by every CTA through DSMEM)
  double v[8];
  int n, tie, pad0, pad1;
};

// cluster-wide sum of NV doubles per thread: warp butterfly -> per-warp partials -> CTA partial -> DSMEM gather in rank order
template <int NV>
__device__ __forceinline__ void cluster_sum(cg::cluster_group& cl, double* red, ClPartial* part, ClPartial* tot, double (&v)[NV],
                                            int& n, int& flag, int tid) {
  constexpr int kWarps = kThreadsCl / 32;
  const int lane = tid & 31, w = tid >> 5;
#pragma unroll
  for (int k = 0; k < NV; k++) v[k] = warp_sum(v[k]);
  const int wn = __reduce_add_sync(0xffffffffu, n);
  const int wf = __any_sync(0xffffffffu, flag != 0);
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) red[w * kRedStride + k] = v[k];
    ((int*)&red[w * kRedStride + 8])[0] = wn; ((int*)&red[w * kRedStride + 8])[1] = wf;
  }
  __syncthreads();
  if (tid < 32) {
    double t = 0.0;
    if (lane < NV) {
#pragma unroll 4
      for (int q = 0; q < kWarps; q++) t += red[q * kRedStride + lane];
      part->v[lane] = t;
    }
    const int cn = __reduce_add_sync(0xffffffffu, ((const int*)&red[lane * kRedStride + 8])[0]);     // kWarps == 32
    const int cf = __any_sync(0xffffffffu, ((const int*)&red[lane * kRedStride + 8])[1] != 0);
    if (lane == 0) { part->n = cn; part->tie = cf; }
  }
  cl.sync();
  if (tid < 32) {                  // one warp reads the kCluster partials through DSMEM (rank order), the CTA reads the totals locally
    double t = 0.0;
    int cn = 0, cf = 0;
#pragma unroll
    for (int r = 0; r < kCluster; r++) {
      const ClPartial* p = cl.map_shared_rank(part, r);
      if (lane < NV) t += p->v[lane];
      cn += p->n; cf |= p->tie;
    }
    if (lane < NV) tot->v[lane] = t;
    if (lane == 0) { tot->n = cn; tot->tie = cf; }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < NV; k++) v[k] = tot->v[k];
  n = tot->n; flag = tot->tie;
}

// in-place exclusive scan of n_ent uint16 counters in the cluster's workspace, rank r scanning the r-th contiguous part
__device__ __forceinline__ void cluster_scan_cells(cg::cluster_group& cl, uint16_t* a, int n_ent, uint32_t* wtot, uint32_t* ctot,
                                                   int tid, int rank) {
  constexpr int kWarps = kThreadsCl / 32;
  const int per = (((n_ent + kCluster - 1) / kCluster) + 255) & ~255;          // 256-entry aligned parts: vector loads stay aligned
  const int rb = min(rank * per, n_ent), re = min(rb + per, n_ent);
  const int w = tid >> 5, lane = tid & 31;
  const int per_warp = ((((re - rb) + kWarps - 1) / kWarps) + 255) & ~255;
  const int beg = min(rb + w * per_warp, re), end = min(beg + per_warp, re);
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
  if (tid == 0) { uint32_t t = 0; for (int k = 0; k < kWarps; k++) t += wtot[k]; *ctot = t; }
  cl.sync();
  uint32_t carry = 0;
  for (int r = 0; r < rank; r++) carry += *cl.map_shared_rank(ctot, r);
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

// ---- grid build over the cluster (a1 + the counting sort; same binning rule and layout as build_grid) ---------------
__device__ __forceinline__ void cluster_build_grid(cg::cluster_group& cl, const IcpArgs& a, const Smem& s, int* hdr, uint32_t* ctot,
                                                   int m0, int nGA, int nNGA, int tid, int rank) {
  const int ctid = rank * kThreadsCl + tid, cstride = kCluster * kThreadsCl;
  const int nm = nGA + nNGA;
  const int G0 = grid_side(nGA), G1 = grid_side(nNGA);
  const int ncell = G0 * G0 + G1 * G1;
  float2* tmp_pts = (float2*)s.region0;
  uint32_t* tmp_cell = (uint32_t*)(tmp_pts + a.mcap);
  uint16_t* tmp_rank = (uint16_t*)(tmp_cell + a.mcap);
  uint32_t* cnt32 = (uint32_t*)s.cstart;
  for (int j = ctid; j < (ncell + 4) / 2; j += cstride) cnt32[j] = 0u;
  if (rank == 0 && tid < CH_COUNT) hdr[tid] = (tid < 8) ? ((tid & 2) ? INT_MIN : INT_MAX) : 0;
  cl.sync();

  int bb[8] = {INT_MAX, INT_MAX, INT_MIN, INT_MIN, INT_MAX, INT_MAX, INT_MIN, INT_MIN};
  const double2* src = (const double2*)a.model_xy + m0;
  for (int i = ctid; i < nm; i += cstride) {
    const double2 p = __ldg(src + i);
    const float2 f = make_float2(__double2float_rn(p.x), __double2float_rn(p.y));
    tmp_pts[i] = f;
    const int c4 = (i >= nGA) ? 4 : 0;
    const int ox = f2ord(f.x), oy = f2ord(f.y);
    bb[c4 + 0] = min(bb[c4 + 0], ox); bb[c4 + 1] = min(bb[c4 + 1], oy);
    bb[c4 + 2] = max(bb[c4 + 2], ox); bb[c4 + 3] = max(bb[c4 + 3], oy);
  }
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const int v = (k & 2) ? __reduce_max_sync(0xffffffffu, bb[k]) : __reduce_min_sync(0xffffffffu, bb[k]);
    if ((tid & 31) == 0) { if (k & 2) atomicMax(&hdr[CH_BBOX + k], v); else atomicMin(&hdr[CH_BBOX + k], v); }
  }
  cl.sync();

  if (tid < 2) {                    // every CTA derives the same grid parameters into its own shared memory
    GridP g;
    const int n = tid ? nNGA : nGA;
    g.G = tid ? G1 : G0;
    g.base = tid ? G0 * G0 : 0;
    g.n = n;
    g.first = tid ? nGA : 0;
    g.minx = g.miny = 0.0f; g.h = 1.0f; g.inv_h = 1.0f; g.cx = g.cy = 0.0f; g.pad0 = g.pad1 = 0;
    if (g.G > 0) {
      const volatile int* hb = hdr + CH_BBOX + 4 * tid;
      const float minx = ord2f(hb[0]), miny = ord2f(hb[1]), maxx = ord2f(hb[2]), maxy = ord2f(hb[3]);
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

  for (int i = ctid; i < nm; i += cstride) {
    const GridP& g = s.grid[i >= nGA];
    if (g.G == 0) continue;
    const float2 f = tmp_pts[i];
    int ix, iy;
    cell_of(g, f.x, f.y, ix, iy);
    const int cell = g.base + iy * g.G + ix;
    const int sh = (cell & 1) * 16;
    const uint32_t old = atomicAdd(&cnt32[cell >> 1], 1u << sh);
    tmp_cell[i] = (uint32_t)cell;
    tmp_rank[i] = (uint16_t)((old >> sh) & 0xffffu);
  }
  cl.sync();
  cluster_scan_cells(cl, s.cstart, ncell + 1, (uint32_t*)s.red, ctot, tid, rank);
  cl.sync();
  for (int i = ctid; i < nm; i += cstride) {
    const int cls = i >= nGA;
    if (s.grid[cls].G == 0) continue;
    const int slot = s.cstart[tmp_cell[i]] + tmp_rank[i];
    s.mpts[slot] = tmp_pts[i];
    s.morig[slot] = (uint16_t)(i - (cls ? nGA : 0));
    s.kinv[i] = (uint16_t)slot;
  }
  cl.sync();
}

// ---- the kernel ------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreadsCl, 1) icp_pair_cluster_kernel(const IcpArgs a) {
  constexpr int kWarps = kThreadsCl / 32;
  cg::cluster_group cl = cg::this_cluster();
  extern __shared__ __align__(16) char smem_raw[];        // staging of the cooperative tie replay (rank 0 uses it)
  __shared__ double red_s[2 * kWarps * kRedStride];
  __shared__ double rtprev_s[8 + kWarps * 4];
  __shared__ GridP grid_s[2];
  __shared__ int misc_s[M_COUNT];
  __shared__ ClPartial part_s[2];
  __shared__ ClPartial tot_s;
  __shared__ uint32_t ctot_s;
  const int tid = threadIdx.x;
  const int rank = (int)cl.block_rank();
  const int cluster_id = blockIdx.x / kCluster;
  const int ctid = rank * kThreadsCl + tid, cstride = kCluster * kThreadsCl;
  char* wbase = a.workspace + (size_t)cluster_id * a.workspace_per_cta;
  int* hdr = (int*)wbase;
  Smem s = carve<true>(wbase + CH_COUNT * sizeof(int), a.mcap, a.scap, a.ccap, red_s, rtprev_s, grid_s, misc_s);
  const float gate_f = __double2float_ru(a.indist);
  const bool replay_on = !(a.flags & CCICP_F_NO_TIE_REPLAY);
  const int n_items = a.n_pairs * a.items_per_pair;
  int par = 0;                       // parity of the published partial (toggles with every cluster_sum)
  int dummy_evals = 0;

  for (;;) {
    cl.sync();
    if (rank == 0 && tid == 0) {
      const int it = atomicAdd(a.work_counter, 1);
      for (int r = 0; r < kCluster; r++) cl.map_shared_rank(misc_s, r)[M_WORK] = it;
    }
    cl.sync();
    const int item = misc_s[M_WORK];
    if (item >= n_items) break;
    const int pair = a.pair_base + item / a.items_per_pair;
    const int seed0 = (item % a.items_per_pair) * a.seeds_per_item;
    const int seed1 = min(seed0 + a.seeds_per_item, a.n_seeds);
    const int mi = a.model_id ? a.model_id[pair] : pair, si = a.scene_id ? a.scene_id[pair] : pair;
    const int m0 = a.model_off[2 * mi], nmGA = a.model_off[2 * mi + 1] - m0, nmNGA = a.model_off[2 * mi + 2] - a.model_off[2 * mi + 1];
    const int s0 = a.scene_off[2 * si], nsGA = a.scene_off[2 * si + 1] - s0, nsNGA = a.scene_off[2 * si + 2] - a.scene_off[2 * si + 1];
    const int ns = nsGA + nsNGA;
    const bool model_ok = nmGA + nmNGA >= 5, scene_ok = ns >= 5;
    const bool fits = grid_side(nmGA) * grid_side(nmGA) + grid_side(nmNGA) * grid_side(nmNGA) <= a.ccap &&
                      nmGA + nmNGA <= a.mcap && ns <= a.scap;
    const bool run = model_ok && scene_ok && fits;
    if (run) {
      cluster_build_grid(cl, a, s, hdr, &ctot_s, m0, nmGA, nmNGA, tid, rank);
      s.scene = (const double2*)a.scene_xy + s0;           // read in place (L2) every iteration
    }

    for (int seed = seed0; seed < seed1; seed++) {
      const size_t prob = (size_t)pair * a.n_seeds + seed;
      double Rt[6];
#pragma unroll
      for (int k = 0; k < 6; k++) Rt[k] = __ldg(a.init_Rt + prob * 6 + k);
      uint32_t status = 0;
      int iters = 0, num_corr = 0;
      double last_delta = 0.0;
      bool have_ok = false, last_failed = false;
      cl.sync();
      if (run) {
        if (rank == 0 && tid == 0) { hdr[CH_REPLAYS] = 0; hdr[CH_TREEFAIL] = 0; }
        for (int i = ctid; i < ns; i += cstride) s.slots[i] = (uint16_t)kNoSlot;
        for (int w = ctid; w <= (ns - 1) >> 5 && ns > 0; w += cstride) s.tiebits[w] = 0u;
      }
      cl.sync();

      if (!model_ok) status |= CCICP_ST_TOO_FEW_MODEL;
      else if (!scene_ok) status |= CCICP_ST_TOO_FEW_SCENE;
      else if (!fits) status |= CCICP_ST_HINT_TOO_SMALL;
      else {
        const GridP& gp = s.grid[nmNGA >= nmGA];
        const double pcx = (double)gp.cx, pcy = (double)gp.cy;
        for (int iter = 0; iter < a.max_iter; iter++) {
          int pend = 0;
          for (int i = ctid; i < ns; i += cstride) {          // loop A: search
            const double2 sp = s.scene[i];
            float qx, qy;
            transform(Rt, sp.x, sp.y, qx, qy);
            const GridP& g = s.grid[i >= nsGA];
            unsigned slot = kNoSlot;
            if (g.G > 0) {
              float best; int bslot, tie;
              nn_search<true, false>(s.mpts, s.cstart, g, qx, qy, gate_f, s.slots[i], best, bslot, tie, dummy_evals);
              if (bslot >= 0 && (double)best < a.indist) {
                slot = (unsigned)bslot;
                if (tie && replay_on) { pend = 1; atomicOr(&s.tiebits[i >> 5], 1u << (i & 31)); }
              }
            }
            s.slots[i] = (uint16_t)slot;
          }
          double v[8];
          int cnt = 0;
#pragma unroll
          for (int k = 0; k < 8; k++) v[k] = 0.0;
          for (int i = ctid; i < ns; i += cstride) {          // loop B: shifted sums of the accepted pairs
            const unsigned slot = s.slots[i];
            if (slot == kNoSlot) continue;
            const double2 sp = s.scene[i];
            float qx, qy;
            transform(Rt, sp.x, sp.y, qx, qy);
            accum_corr(v, s.mpts[slot], qx, qy, pcx, pcy);
            cnt++;
          }
          int n = cnt, any_tie = pend;
          cluster_sum<8>(cl, red_s, &part_s[par], &tot_s, v, n, any_tie, tid); par ^= 1;
          if (any_tie) {                         // rare: rank 0 replays the reference kd-tree for every tied query
            if (rank == 0) {
              misc_s[M_TREE] = hdr[CH_TREE]; misc_s[M_TREE + 1] = hdr[CH_TREE + 1];      // lazily built tree state of this pair
              misc_s[M_NNODE] = hdr[CH_NNODE]; misc_s[M_NNODE + 1] = hdr[CH_NNODE + 1];
              misc_s[M_REPLAYS] = 0; misc_s[M_TREEFAIL] = 0;
              __syncthreads();
              replay_ties_coop<kThreadsCl>(a, s, Rt, ns, nsGA, tid, smem_raw, a.stage_cap, (float*)rtprev_s + 16, s.scene, nullptr);
              if (tid == 0) {
                hdr[CH_TREE] = misc_s[M_TREE]; hdr[CH_TREE + 1] = misc_s[M_TREE + 1];
                hdr[CH_NNODE] = misc_s[M_NNODE]; hdr[CH_NNODE + 1] = misc_s[M_NNODE + 1];
                hdr[CH_REPLAYS] += misc_s[M_REPLAYS]; hdr[CH_TREEFAIL] |= misc_s[M_TREEFAIL];
              }
            }
            status |= CCICP_ST_TIE_REPLAY;
            cl.sync();
            cnt = 0;
#pragma unroll
            for (int k = 0; k < 8; k++) v[k] = 0.0;
            for (int i = ctid; i < ns; i += cstride) {
              const unsigned slot = s.slots[i];
              if (slot == kNoSlot) continue;
              const double2 sp = s.scene[i];
              float qx, qy;
              transform(Rt, sp.x, sp.y, qx, qy);
              accum_corr(v, s.mpts[slot], qx, qy, pcx, pcy);
              cnt++;
            }
            n = cnt; int none = 0;
            cluster_sum<8>(cl, red_s, &part_s[par], &tot_s, v, n, none, tid); par ^= 1;
          }
          iters++;
          if (n == 0) {
            last_delta = -1.0; last_failed = true;
            if (!(-1.0 < a.min_delta)) iters = a.max_iter;
            break;
          }
          last_failed = false; have_ok = true; num_corr = n;
          const double dm0 = v[0] / (double)n, dm1 = v[1] / (double)n, dt0 = v[2] / (double)n, dt1 = v[3] / (double)n;
          const double mu_m[2] = {pcx + dm0, pcy + dm1};
          const double mu_t[2] = {pcx + dt0, pcy + dt1};
          const double H[4] = {v[4] - v[2] * dm0, v[5] - v[2] * dm1, v[6] - v[3] * dm0, v[7] - v[3] * dm1};
          if (tid == 0) {
#pragma unroll
            for (int q = 0; q < 6; q++) s.rtprev[q] = Rt[q];
          }
          last_delta = solve_and_compose(H, mu_m, mu_t, Rt, status);
          if (last_delta < a.min_delta) { status |= CCICP_ST_CONVERGED; break; }
        }
        if (!have_ok && iters > 0) status |= CCICP_ST_NO_CORR;
      }

      // ---- outputs
      cl.sync();
      if (a.corr_idx) {
        int32_t* dst = a.corr_idx + (size_t)seed * (size_t)a.total_scene + s0;
        for (int i = ctid; i < ns; i += cstride) {
          int32_t o = -1;
          if (run && iters > 0) {
            const unsigned slot = s.slots[i];
            if (slot != kNoSlot) o = s.morig[slot];
          }
          dst[i] = o;
        }
      }
      double info[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      double score = CUDART_INF;
      if (have_ok) {
        double P[6];
        __syncthreads();
#pragma unroll
        for (int q = 0; q < 6; q++) P[q] = s.rtprev[q];
        if (last_failed) {       // regenerate the slots of the last successful iteration (ties: grid order)
          for (int i = ctid; i < ns; i += cstride) {
            const double2 sp = s.scene[i];
            float qx, qy;
            transform(P, sp.x, sp.y, qx, qy);
            const GridP& g = s.grid[i >= nsGA];
            unsigned slot = kNoSlot;
            if (g.G > 0) {
              float best; int bslot, tie;
              nn_search<true, false>(s.mpts, s.cstart, g, qx, qy, gate_f, kNoSlot, best, bslot, tie, dummy_evals);
              if (bslot >= 0 && (double)best < a.indist) slot = (unsigned)bslot;
            }
            s.slots[i] = (uint16_t)slot;
          }
        }
        const bool want_info = (a.flags & CCICP_F_INFO) != 0;
        const bool fixed_dy = (a.flags & CCICP_F_INFO_FIXED_DY) != 0;
        double v[7] = {0, 0, 0, 0, 0, 0, 0};   // sdis, sx, sy, xpy, mz0, mz1, mz2
        for (int i = ctid; i < ns; i += cstride) {
          const unsigned slot = s.slots[i];
          if (slot == kNoSlot) continue;
          const double2 sp = s.scene[i];
          float qx, qy;
          transform(P, sp.x, sp.y, qx, qy);
          const float2 m = s.mpts[slot];
          const float fdx = __fsub_rn(m.x, qx), fdy = __fsub_rn(m.y, qy);
          v[0] += (double)__fadd_rn(__fmul_rn(fdx, fdx), __fmul_rn(fdy, fdy));
          if (want_info) {      // icpPointToPoint.cpp:250-275
            const double ax = (double)m.x, ay = (double)m.y, bx = (double)qx, by = (double)qy;
            const double x = (ax + bx) / 2.0, y = (ay + by) / 2.0;
            const double dx = ax - bx, dy = fixed_dy ? ay - by : ax - bx;
            v[1] += x; v[2] += y; v[3] += x * x + y * y;
            v[4] += dx; v[5] += dy; v[6] += -y * dx + x * dy;
          }
        }
        int z0 = 0, z1 = 0;
        cluster_sum<7>(cl, red_s, &part_s[par], &tot_s, v, z0, z1, tid); par ^= 1;
        const double nn = (double)num_corr;
        score = v[0] / nn;
        if (want_info) {
          const double sx = v[1], sy = v[2], xpy = v[3];
          const double sch = xpy - (sx * sx + sy * sy) / nn;
          const double D2 = (v[6] + (sy * v[4] - sx * v[5]) / nn) / sch;
          const double D0 = (v[4] + sy * D2) / nn, D1 = (v[5] - sx * D2) / nn;
          double ss[1] = {0.0};
          for (int i = ctid; i < ns; i += cstride) {
            const unsigned slot = s.slots[i];
            if (slot == kNoSlot) continue;
            const double2 sp = s.scene[i];
            float qx, qy;
            transform(P, sp.x, sp.y, qx, qy);
            const float2 m = s.mpts[slot];
            const double ax = (double)m.x, ay = (double)m.y, bx = (double)qx, by = (double)qy;
            const double x = (ax + bx) / 2.0, y = (ay + by) / 2.0;
            const double tx = (ax - bx - D0 + y * D2), ty = (ay - by - D1 - x * D2);
            ss[0] += tx * tx + ty * ty;
          }
          cluster_sum<1>(cl, red_s, &part_s[par], &tot_s, ss, z0, z1, tid); par ^= 1;
          const double sconst = 1.0 / (ss[0] / (double)(2 * num_corr - 3));
          if (!(sch != 0.0) || !(ss[0] > 0.0)) status |= CCICP_ST_INFO_SINGULAR;
          info[0] = nn * sconst; info[1] = 0.0 * sconst; info[2] = -sy * sconst;
          info[3] = 0.0 * sconst; info[4] = nn * sconst; info[5] = sx * sconst;
          info[6] = -sy * sconst; info[7] = sx * sconst; info[8] = xpy * sconst;
        }
      }
      if (rank == 0 && tid == 0) {
        if (run && hdr[CH_TREEFAIL]) status |= CCICP_ST_TREE_OVERFLOW;
        ccicp_result r;
#pragma unroll
        for (int q = 0; q < 4; q++) r.R[q] = Rt[q];
        r.t[0] = Rt[4]; r.t[1] = Rt[5];
#pragma unroll
        for (int q = 0; q < 9; q++) r.info[q] = info[q];
        r.last_delta = last_delta; r.score = score;
        r.num_corr = num_corr; r.iters = iters; r.status = status; r.best_seed = seed;
        r.n_tie_replays = run ? hdr[CH_REPLAYS] : 0; r.reserved = 0;
        a.out[prob] = r;
      }
    }
  }
}

}  // namespace ccicp
