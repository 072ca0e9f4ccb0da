// ccicp_pair_main.cuh -- kernel entry: persistent CTAs pull (pair, seed-chunk)
// work items from an atomic counter; each item = grid build once + all
// iterations of every seed in the chunk, entirely in shared memory/registers.
#pragma once
#include "ccicp_pair_kernel.cuh"
#ifndef CCICP_SOLVE_ALL
#define CCICP_SOLVE_ALL 0
#endif
#include "exact_solve.cuh"
#include "features.cuh"

#ifndef CCICP_SPARSE_LOG2
#define CCICP_SPARSE_LOG2 5
#endif
#ifndef CCICP_SPARSE_FIRST
#define CCICP_SPARSE_FIRST 1      // first iteration of a seed: sparse cold pass, then every point starts from its block's sample
#endif

namespace ccicp {

// block-wide sum of up to NV doubles per thread (+ two ints); result broadcast
// to every thread in a fixed order (deterministic run to run).
template <int NV, int kWarps>
__device__ __forceinline__ void block_reduce(double* red, int off, double (&v)[NV], int tid) {
#pragma unroll
  for (int k = 0; k < NV; k++) v[k] = warp_sum(v[k]);
  if ((tid & 31) == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) red[(tid >> 5) * 10 + off + k] = v[k];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < NV; k++) {
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < kWarps; w++) t += red[w * 10 + off + k];
    v[k] = t;
  }
}

// Sum of 8 doubles per lane over the warp by a halving butterfly: after the xor-16 / 8 / 4 rounds every lane carries
// ONE of the 8 values (4 + 2 + 1 exchanges instead of 8 x 5), two plain rounds finish it.  Returns the warp total of
// value index lane >> 2 (valid in every lane; lanes with (lane & 3) == 0 publish it).  Fixed order: deterministic.
__device__ __forceinline__ double warp_sum8(const double (&v)[8], int lane) {
  const bool h16 = (lane & 16) != 0, h8 = (lane & 8) != 0, h4 = (lane & 4) != 0;
  double a[4], b[2];
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const double keep = h16 ? v[k + 4] : v[k], send = h16 ? v[k] : v[k + 4];
    a[k] = keep + __shfl_xor_sync(0xffffffffu, send, 16);
  }
#pragma unroll
  for (int k = 0; k < 2; k++) {
    const double keep = h8 ? a[k + 2] : a[k], send = h8 ? a[k] : a[k + 2];
    b[k] = keep + __shfl_xor_sync(0xffffffffu, send, 8);
  }
  const double keep = h4 ? b[1] : b[0], send = h4 ? b[0] : b[1];
  double c = keep + __shfl_xor_sync(0xffffffffu, send, 4);
  c += __shfl_xor_sync(0xffffffffu, c, 2);
  c += __shfl_xor_sync(0xffffffffu, c, 1);
  return c;
}

// Every warp publishes its partial sums (8 doubles) and its count of accepted pairs; ONE warp then adds the partials in
// warp order and solves (solve_warp0), so the FP64 tail of an iteration is issued once per CTA, not once per warp.
__device__ __forceinline__ void publish_partials(double* red, const double (&v)[8], int cnt, int tid) {
  const int lane = tid & 31;
  const double c = warp_sum8(v, lane);
  const int wc = __reduce_add_sync(0xffffffffu, cnt);
  double* r = red + (tid >> 5) * kRedStride;
  if ((lane & 3) == 0) r[lane >> 2] = c;
  if (lane == 0) ((int*)&r[8])[0] = wc;
}

// warp 0, after the barrier: total count n and the 8 totals (in every lane)
template <int kWarps>
__device__ __forceinline__ int gather_partials(const double* red, double (&v)[8], int lane) {
  const int n = __reduce_add_sync(0xffffffffu, lane < kWarps ? ((const int*)&red[lane * kRedStride + 8])[0] : 0);
  double t = 0.0;
#pragma unroll
  for (int w = 0; w < kWarps; w++) t += red[w * kRedStride + (lane & 7)];
#pragma unroll
  for (int k = 0; k < 8; k++) v[k] = __shfl_sync(0xffffffffu, t, k);
  return n;
}

// a8 + a9.  Closed form of R_ = V*U^T for the 2x2 H = sum q_t^T q_m
// (icpPointToPoint.cpp:159-163): a rotation when det H > 0, the improper
// (reflection) polar factor when det H < 0 -- the reference applies no det fix.
__device__ __forceinline__ double solve_and_compose(const double* H, const double* mu_m, const double* mu_t,
                                                    double* Rt, uint32_t& status) {
  const double det = H[0] * H[3] - H[1] * H[2];
  // Rank-deficient or nearly so (one or two correspondences, collinear ones): the sign of det H is rounding noise and
  // the reference's SVD picks its factor by its own internals, so the closed form is not comparable here.  Flagged;
  // the host side re-runs flagged problems with the exact-order instantiation (see launch_icp).
  if (fabs(det) <= 1e-9 * (H[0] * H[0] + H[1] * H[1] + H[2] * H[2] + H[3] * H[3])) status |= CCICP_ST_DEGENERATE;
  double r00, r01, r10, r11;
  if (det >= 0.0) {
    const double A = H[0] + H[3], B = H[1] - H[2];
    const double r = sqrt(A * A + B * B);
    if (!(r > 0.0)) status |= CCICP_ST_DEGENERATE;
    const double c = r > 0.0 ? A / r : 1.0, s = r > 0.0 ? B / r : 0.0;
    r00 = c; r01 = -s; r10 = s; r11 = c;
  } else {
    status |= CCICP_ST_REFLECTION;
    const double A = H[0] - H[3], B = H[1] + H[2];
    const double r = sqrt(A * A + B * B);
    const double c = r > 0.0 ? A / r : 1.0, s = r > 0.0 ? B / r : 0.0;
    r00 = c; r01 = s; r10 = s; r11 = -c;
  }
  const double t0 = mu_m[0] - (r00 * mu_t[0] + r01 * mu_t[1]);
  const double t1 = mu_m[1] - (r10 * mu_t[0] + r11 * mu_t[1]);
  // compose: R = R_*R ; t = R_*t + t_   (icpPointToPoint.cpp:166-167)
  const double n00 = r00 * Rt[0] + r01 * Rt[2], n01 = r00 * Rt[1] + r01 * Rt[3];
  const double n10 = r10 * Rt[0] + r11 * Rt[2], n11 = r10 * Rt[1] + r11 * Rt[3];
  const double nt0 = (r00 * Rt[4] + r01 * Rt[5]) + t0, nt1 = (r10 * Rt[4] + r11 * Rt[5]) + t1;
  Rt[0] = n00; Rt[1] = n01; Rt[2] = n10; Rt[3] = n11; Rt[4] = nt0; Rt[5] = nt1;
  // delta = max(||R_-I||_F, ||t_||_2)   (icpPointToPoint.cpp:170, matrix.cpp:354-360)
  const double a = r00 - 1.0, d = r11 - 1.0;
  const double nr = sqrt(a * a + r01 * r01 + r10 * r10 + d * d);
  const double nt = sqrt(t0 * t0 + t1 * t1);
  return nr > nt ? nr : nt;
}

// CCICP_F_EXACT_SUMS: statistics and solve of one fitStep in the reference's evaluation order (exact_solve.cuh).
// Warp 0 does it: per chunk of 32 scene points every lane prepares one pair (transform, model point) in parallel,
// then the accepted pairs are fed IN ORDER to the accumulators by warp shuffles, so the only serial chain is the
// FP64 add itself.  Lane L carries accumulator L & 3 (the other lanes hold redundant copies); lane 0 runs the SVD.
// ex[0..3] = mu_m, mu_t; ex[4..7] = H; ex[8..13] = new R,t; ex[14] = delta; ex[15] = status bits.
// Stages the accepted pairs of scene points [i0, i0+32) in order into buf[k][4] = {m.x, m.y, q.x, q.y} as doubles
// (the float -> double widening is exact, :83-84); returns how many there are.
__device__ __forceinline__ int exact_stage(const Smem& s, const double* Rt, int ns, int i, int lane, double* buf) {
  bool valid = false;
  float qx = 0.0f, qy = 0.0f;
  float2 m = make_float2(0.0f, 0.0f);
  if (i < ns) {
    const unsigned slot = s.slots[i];
    if (slot != kNoSlot) {
      valid = true;
      const double2 sp = s.scene[i];
      transform(Rt, sp.x, sp.y, qx, qy);
      m = s.mpts[slot];
    }
  }
  const unsigned mask = __ballot_sync(0xffffffffu, valid);
  if (valid) {
    double* d = buf + 4 * __popc(mask & ((1u << lane) - 1u));
    d[0] = (double)m.x; d[1] = (double)m.y; d[2] = (double)qx; d[3] = (double)qy;
  }
  __syncwarp();
  return __popc(mask);
}

__device__ __noinline__ void exact_step(const Smem& s, const double* Rt, int ns, int n, int tid, double* ex) {
  if (tid < 32) {
    const int lane = tid, q4 = lane & 3;
    double* buf = s.red;                                   // both parity buffers are idle here: 160 >= 128 doubles
    double acc = 0.0;
    for (int i0 = 0; i0 < ns; i0 += 32) {                  // icpPointToPoint.cpp:80-86, 115-121 (scene order, GA first)
      const int cnt = exact_stage(s, Rt, ns, i0 + lane, lane, buf);
#pragma unroll 4
      for (int k = 0; k < cnt; k++) acc += buf[4 * k + q4];
      __syncwarp();
    }
    const double mu = acc / (double)n;                     // :149-150
    const double mu_m0 = __shfl_sync(0xffffffffu, mu, 0), mu_m1 = __shfl_sync(0xffffffffu, mu, 1);
    const double mu_t0 = __shfl_sync(0xffffffffu, mu, 2), mu_t1 = __shfl_sync(0xffffffffu, mu, 3);
    if (lane < 4) ex[lane] = mu;
    // H[ct][cm] = sum_k (p_t[k][ct] - mu_t[ct]) * (p_m[k][cm] - mu_m[cm]), lane = 2*ct + cm
    const int cm = lane & 1, ct = (lane >> 1) & 1;
    const double my_mu_m = cm ? mu_m1 : mu_m0, my_mu_t = ct ? mu_t1 : mu_t0;
    double h = 0.0;
    for (int i0 = 0; i0 < ns; i0 += 32) {
      const int cnt = exact_stage(s, Rt, ns, i0 + lane, lane, buf);
#pragma unroll 4
      for (int k = 0; k < cnt; k++) {
        const double pm = buf[4 * k + cm], pt = buf[4 * k + 2 + ct];
        h += __dmul_rn(pt - my_mu_t, pm - my_mu_m);
      }
      __syncwarp();
    }
    if (lane < 4) ex[4 + lane] = h;
    __syncwarp();
    if (tid == 0) {
      const double H[4] = {ex[4], ex[5], ex[6], ex[7]};
      const double mu_m2[2] = {ex[0], ex[1]}, mu_t2[2] = {ex[2], ex[3]};
      double U[4], V[4];
      exact::svd2(H, U, V);
      double R_[4], t_[2], Rn[4], tn[2];
      for (int i = 0; i < 2; i++)                          // R_ = V * ~U (:162)
        for (int j = 0; j < 2; j++) { double c = 0.0; for (int k = 0; k < 2; k++) c += V[2 * i + k] * U[2 * j + k]; R_[2 * i + j] = c; }
      for (int i = 0; i < 2; i++) { double c = 0.0; for (int k = 0; k < 2; k++) c += R_[2 * i + k] * mu_t2[k]; t_[i] = mu_m2[i] - c; }
      for (int i = 0; i < 2; i++)                          // R = R_*R, t = R_*t + t_ (:166-167)
        for (int j = 0; j < 2; j++) { double c = 0.0; for (int k = 0; k < 2; k++) c += R_[2 * i + k] * Rt[2 * k + j]; Rn[2 * i + j] = c; }
      for (int i = 0; i < 2; i++) { double c = 0.0; for (int k = 0; k < 2; k++) c += R_[2 * i + k] * Rt[4 + k]; tn[i] = c + t_[i]; }
      double nr = 0.0, nt = 0.0;                           // :170, matrix.cpp:354-360
      for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++) { const double v = R_[2 * i + j] - (i == j ? 1.0 : 0.0); nr += v * v; }
      for (int i = 0; i < 2; i++) nt += t_[i] * t_[i];
      nr = sqrt(nr); nt = sqrt(nt);
      ex[8] = Rn[0]; ex[9] = Rn[1]; ex[10] = Rn[2]; ex[11] = Rn[3]; ex[12] = tn[0]; ex[13] = tn[1];
      ex[14] = nr < nt ? nt : nr;
      const double det = R_[0] * R_[3] - R_[1] * R_[2];
      const double detH = H[0] * H[3] - H[1] * H[2];
      uint32_t st = 0;
      if (det < 0.0) st |= CCICP_ST_REFLECTION;
      if (detH == 0.0) st |= CCICP_ST_DEGENERATE;
      ex[15] = (double)st;
    }
  }
}

// CCICP_F_TRIM (a north_star extra; the reference has only the fixed gate): keep the `trim_frac` fraction of this
// iteration's gated correspondences with the smallest squared distances.  The threshold is the k-th smallest distance,
// k = ceil(trim_frac * n_gated), found by a 4 x 8-bit radix select over the (non-negative, hence order-preserving) float
// bit patterns in shared memory; correspondences at the threshold distance are all kept.  A trimmed scene point is
// treated like an ungated one from here on (slot = none: no part in the sums, the indices, the score, the information
// matrix); its match only serves as the starting bound of the next iteration's search.  The distances are computed once
// per call with the search's own float expression and kept as the select's keys.
template <int kThreads>
__device__ __noinline__ void trim_select(const IcpArgs& a, Smem& s, int ns, int tid, uint32_t* th, uint32_t* keys) {
  // keys[i] = bit pattern of the gated squared distance of scene point i, left there by the search loop that found it
  // (0xFFFFFFFF = not gated: never matches a prefix because a finite float's top byte is < 0xFF); every thread reads
  // back only what it wrote
  uint32_t prefix = 0, mask = 0;
  for (int pass = 3; pass >= 0; pass--) {
    __syncthreads();
    if (tid < 256) th[tid] = 0;
    __syncthreads();
    for (int i = tid; i < ns; i += kThreads) {
      const uint32_t key = keys[i];
      if (key != 0xFFFFFFFFu && (key & mask) == prefix) atomicAdd(&th[(key >> (8 * pass)) & 255u], 1u);
    }
    __syncthreads();
    if (tid < 32) {
      uint32_t c[8], sum = 0;
#pragma unroll
      for (int q = 0; q < 8; q++) { c[q] = th[tid * 8 + q]; sum += c[q]; }
      uint32_t incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (tid >= o) incl += t; }
      const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
      uint32_t k = th[257];
      __syncwarp();                                        // every lane has read the carried rank before one lane replaces it
      if (pass == 3) {                                     // first pass: all gated correspondences are counted
        const double want = ceil(a.trim_frac * (double)total);
        k = want < 1.0 ? 1u : (want > (double)total ? total : (uint32_t)want);
      }
      const uint32_t excl = incl - sum;
      if (total > 0 && excl < k && k <= incl) {            // exactly one lane
        uint32_t run = excl;
        int dgt = 0;
#pragma unroll
        for (int q = 0; q < 8; q++) { if (run + c[q] < k) { run += c[q]; dgt = q + 1; } else break; }
        th[256] = (uint32_t)(tid * 8 + dgt);
        th[257] = k - run;                                 // rank inside the chosen digit
      }
      if (tid == 0) { th[258] = total; if (pass == 3) th[259] = (k >= total) ? 1u : 0u; }
    }
    __syncthreads();
    if (th[258] == 0 || th[259]) {                         // nothing gated, or every gated pair is kept (trim_frac = 1):
      for (int i = tid; i < ns; i += kThreads) keys[i] = kNoSlot;      // nothing to trim, no further passes (uniform)
      return;
    }
    prefix |= th[256] << (8 * pass);
    mask |= 0xFFu << (8 * pass);
  }
  // From here on keys[i] is the search's starting slot for the next iteration: the match of a trimmed point (still
  // the best guess; the search is exact whatever the start), none otherwise.  Without it the trimmed points paid for a
  // cold search every iteration (C2, trim_frac 0.7: 9.6 ms against 6.3 ms with nothing trimmed).
  for (int i = tid; i < ns; i += kThreads) {
    const uint32_t key = keys[i];
    uint32_t start = kNoSlot;
    if (key != 0xFFFFFFFFu && key > prefix) {
      start = s.slots[i];
      s.slots[i] = (uint16_t)kNoSlot;
      atomicAnd(&s.tiebits[i >> 5], ~(1u << (i & 31)));
    }
    keys[i] = start;
  }
}

template <int kThreads, bool kGlobal, bool kExact, bool kTrim = false, bool kDiag = false, bool kCurv = false, int kLoops = CCICP_ROW_LOOPS>
__device__ __forceinline__ void icp_pair_body(const IcpArgs& a) {
  constexpr int kWarps = kThreads / 32;
  extern __shared__ __align__(16) char smem_raw[];
  __shared__ double red_s[2 * kWarps * kRedStride];
  __shared__ double rtprev_s[8 + kWarps * 4];     // + float[kWarps][8] min/max / scan scratch of the tie replay
  __shared__ double exact_s[kExact ? 16 : 1];
  __shared__ GridP grid_s[2];
  __shared__ int misc_s[M_COUNT];
  __shared__ double sol_s[8];          // warp 0 -> CTA: new R,t, delta, {n, status bits} of the iteration
  __shared__ uint32_t trim_s[kTrim ? 264 : 1];
  char* base = kGlobal ? a.workspace + (size_t)blockIdx.x * a.workspace_per_cta : smem_raw;
  Smem s = carve<kGlobal>(base, a.mcap, a.scap, a.ccap, red_s, rtprev_s, grid_s, misc_s);
  uint32_t* trim_keys = nullptr;       // kTrim: [scap] behind the carved working set
  if constexpr (kTrim) {
    trim_keys = (uint32_t*)(base + workset_bytes(a.mcap, a.scap, a.ccap, kGlobal));
    if (!kGlobal) __builtin_assume(__isShared(trim_keys));
  }
  // kCurv (CCICP_F_CURV_GATE): curvature of every model point (by sorted slot) and of every scene point, behind the
  // carved working set; computed once per pair from the FP64 inputs (features.cuh)
  float* mcurv = nullptr; float* scurv = nullptr;
  if constexpr (kCurv) {
    mcurv = (float*)(base + workset_bytes(a.mcap, a.scap, a.ccap, kGlobal));
    scurv = mcurv + a.mcap;
    if (!kGlobal) { __builtin_assume(__isShared(mcurv)); __builtin_assume(__isShared(scurv)); }
  }
  const float ctol = kCurv ? (float)a.curv_tol : 0.0f;
  const int tid = threadIdx.x;
  const float gate_f = __double2float_ru(a.indist);
  const bool replay_on = !(a.flags & CCICP_F_NO_TIE_REPLAY) && !kCurv;      // curvature-gated ties: lowest model index (below)
  // a.item_list (exact-order re-run of flagged problems): one listed (pair, seed) problem per work item, the count
  // was left on the device by collect_retry_kernel
  const bool listed = a.item_list != nullptr;
  const int n_items = listed ? *a.item_count : a.n_pairs * a.items_per_pair;
  int diag_evals = 0, diag_queries = 0, diag_inliers = 0;      // kDiag only (dead otherwise)

  for (;;) {
    __syncthreads();
    if (tid == 0) s.misc[M_WORK] = atomicAdd(a.work_counter, 1);
    __syncthreads();
    const int item = s.misc[M_WORK];
    if (item >= n_items) break;
    int pair, seed0, seed1;
    if (listed) {
      const int prob = a.item_list[item];
      pair = prob / a.n_seeds; seed0 = prob - pair * a.n_seeds; seed1 = seed0 + 1;
    } else {
      pair = a.pair_base + item / a.items_per_pair;
      seed0 = (item % a.items_per_pair) * a.seeds_per_item;
      seed1 = min(seed0 + a.seeds_per_item, a.n_seeds);
    }

    const int mi = a.model_id ? a.model_id[pair] : pair, si = a.scene_id ? a.scene_id[pair] : pair;
    const int m0 = a.model_off[2 * mi], nmGA = a.model_off[2 * mi + 1] - m0, nmNGA = a.model_off[2 * mi + 2] - a.model_off[2 * mi + 1];
    const int s0 = a.scene_off[2 * si], nsGA = a.scene_off[2 * si + 1] - s0, nsNGA = a.scene_off[2 * si + 2] - a.scene_off[2 * si + 1];
    const int ns = nsGA + nsNGA;
    const bool model_ok = nmGA + nmNGA >= 5, scene_ok = ns >= 5;
    // capacity of this launch's working set (a caller-supplied cell hint may be too small for this pair)
    const bool fits = grid_side(nmGA) * grid_side(nmGA) + grid_side(nmNGA) * grid_side(nmNGA) <= a.ccap &&
                      nmGA + nmNGA <= a.mcap && ns <= a.scap;

    if (model_ok && scene_ok && fits) {
      build_grid<kThreads, kGlobal>(a, s, m0, nmGA, nmNGA, tid);
      const double2* src = (const double2*)a.scene_xy + s0;
      if (kGlobal) {
        s.scene = src;             // read the (L2-resident) input in place every iteration
      } else {
        double2* dst = (double2*)s.region0;
        for (int i = tid; i < ns; i += kThreads) dst[i] = __ldg(src + i);
        // each thread only ever reads back the scene points it loaded itself
      }
      if constexpr (kCurv) {
        const double r2 = a.feat_radius * a.feat_radius;
        const double2* msrc = (const double2*)a.model_xy + m0;
        for (int i = tid; i < nmGA + nmNGA; i += kThreads) {
          const bool nga = i >= nmGA;
          if (s.grid[nga].G == 0) continue;                  // class never searched: its points have no slot
          float c, nx, ny;
          feat::point_feature(msrc, i, nga ? nmGA : 0, nga ? nmGA + nmNGA : nmGA, a.feat_k, r2, c, nx, ny);
          mcurv[s.kinv[i]] = c;
        }
        for (int i = tid; i < ns; i += kThreads) {
          const bool nga = i >= nsGA;
          float c, nx, ny;
          feat::point_feature(src, i, nga ? nsGA : 0, nga ? ns : nsGA, a.feat_k, r2, c, nx, ny);
          scurv[i] = c;
        }
        __syncthreads();
      }
    }

    for (int seed = seed0; seed < seed1; seed++) {
      const size_t prob = (size_t)pair * a.n_seeds + seed;
      double Rt[6];
#pragma unroll
      for (int k = 0; k < 6; k++) Rt[k] = __ldg(a.init_Rt + prob * 6 + k);
      uint32_t status = listed ? (uint32_t)CCICP_ST_EXACT_RETRY : 0u;
      int iters = 0, num_corr = 0;
      double last_delta = 0.0;
      bool have_ok = false, last_failed = false;
      __syncthreads();
      if (tid == 0) { s.misc[M_REPLAYS] = 0; s.misc[M_TREEFAIL] = 0; s.misc[M_TIEFLAG] = 0; s.misc[M_TIEFLAG + 1] = 0; }
      if (model_ok && scene_ok && fits) {       // (a pair that does not fit must not touch the carved arrays: ns may exceed scap)
        for (int i = tid; i < ns; i += kThreads) s.slots[i] = (uint16_t)kNoSlot;   // no warm start for a new seed
        if constexpr (kTrim) { for (int i = tid; i < ns; i += kThreads) trim_keys[i] = kNoSlot; }
        for (int w = tid; w <= (ns - 1) >> 5 && ns > 0; w += kThreads) s.tiebits[w] = 0u;
      }
      __syncthreads();

      if (!model_ok) status |= CCICP_ST_TOO_FEW_MODEL;
      else if (!scene_ok) status |= CCICP_ST_TOO_FEW_SCENE;
      else if (!fits) status |= CCICP_ST_HINT_TOO_SMALL;
      else {
        const GridP& gp = s.grid[nmNGA >= nmGA];
        const double pcx = (double)gp.cx, pcy = (double)gp.cy;     // pivot of the shifted sums
        for (int iter = 0; iter < a.max_iter; iter++) {
          const int par = iter & 1;
          // ---- a4..a6: transform, nearest neighbour, gate
          int pend = 0;
          // (large-scan variants only: on chip the sparse pass and its barrier cost 1 % more than they save; not with the
          // curvature gate, where a borrowed point need not be a candidate of this query at all)
          constexpr bool kSparseFirst = CCICP_SPARSE_FIRST && kGlobal && !kCurv;
          constexpr int kSparseMask = (1 << CCICP_SPARSE_LOG2) - 1;
          // First iteration of a seed: nobody has a warm start.  A sparse pass (phase 0) searches every 2^CCICP_SPARSE_LOG2-th scene point
          // cold; the full pass (phase 1) then starts a point without a match of its own from the match of the sample
          // of its block -- a model point a few beams away bounds the search almost as well as a warm start.
          // Later iterations run phase 1 only (a point that lost its match borrows the sample's the same way).
          for (int phase = (kSparseFirst && iter == 0) ? 0 : 1; phase < 2; phase++) {
          const int nq = phase ? ns : (ns + kSparseMask) >> CCICP_SPARSE_LOG2;
          for (int q = tid; q < nq; q += kThreads) {          // loop A: search only (few live registers)
            const int i = phase ? q : q << CCICP_SPARSE_LOG2;
            const double2 sp = s.scene[i];
            float qx, qy;
            transform(Rt, sp.x, sp.y, qx, qy);
            const GridP& g = s.grid[i >= nsGA];
            unsigned slot = kNoSlot;
            if (g.G > 0) {
              float best; int bslot, tie;
              unsigned start = s.slots[i];
              if constexpr (kTrim) { if (start == kNoSlot) start = trim_keys[i] & 0xFFFFu; }
              if (kSparseFirst && start == kNoSlot && phase && ((i & ~kSparseMask) >= nsGA) == (i >= nsGA)) start = s.slots[i & ~kSparseMask];
              if constexpr (kCurv) {
                const float qc = scurv[i];
                nn_search<kGlobal, false, true, kLoops>(s.mpts, s.cstart, g, qx, qy, gate_f, start, best, bslot, tie, diag_evals, mcurv, qc, ctol);
                if (tie) {     // several compatible model points at the float minimum (rare): the lowest model index wins
                  const QPack q2 = pack_query(qx, qy);
                  int bo = 0x7fffffff;
                  for (int j = s.cstart[g.base]; j < s.cstart[g.base + g.G * g.G]; j++)
                    if (!(fabsf(mcurv[j] - qc) > ctol) && dist2(s.mpts[j], q2) == best && (int)s.morig[j] < bo) { bo = s.morig[j]; bslot = j; }
                  tie = 0;
                }
              } else {
                nn_search<kGlobal, kDiag, false, kLoops>(s.mpts, s.cstart, g, qx, qy, gate_f, start, best, bslot, tie, diag_evals);
              }
              if (kDiag) diag_queries++;
              if (bslot >= 0 && (double)best < a.indist) {
                slot = (unsigned)bslot;
                if (tie && replay_on) { pend = 1; atomicOr(&s.tiebits[i >> 5], 1u << (i & 31)); }
              }
              if constexpr (kTrim) trim_keys[i] = slot != kNoSlot ? __float_as_uint(best) : 0xFFFFFFFFu;   // the select's key
            } else if constexpr (kTrim) trim_keys[i] = 0xFFFFFFFFu;
            s.slots[i] = (uint16_t)slot;
          }
          if (!phase) __syncthreads();
          }
          if constexpr (kTrim) trim_select<kThreads>(a, s, ns, tid, trim_s, trim_keys);
          // ---- a7: shifted sums of the accepted pairs; every warp publishes its partials, warp 0 adds them and solves
          double v[8];
          int cnt;
          sums_from_slots<kThreads>(s, Rt, ns, tid, pcx, pcy, v, cnt);        // loop B: converged accumulation
          publish_partials(s.red, v, cnt, tid);
          if (__any_sync(0xffffffffu, pend != 0) && (tid & 31) == 0) s.misc[M_TIEFLAG + par] = 1;
          __syncthreads();
          const int any_tie = s.misc[M_TIEFLAG + par];
          if (tid == 0) s.misc[M_TIEFLAG + (par ^ 1)] = 0;   // the next iteration's flag (last read one barrier pair ago)
          if (any_tie) {                         // rare: replay the reference kd-tree for the tied queries
            if (kGlobal) {
              replay_ties_coop<kThreads>(a, s, Rt, ns, nsGA, tid, smem_raw, a.stage_cap, (float*)rtprev_s + 16, s.scene, nullptr);
            } else {
              // On chip the whole CTA replays too (one thread per class used to: 0.1-0.3 ms for the first tie of a pair,
              // and a tie in one of the last pairs of a launch is the kernel's tail -- the 14 replays of the C2 batch
              // cost 0.17 ms of 3.85).  The staging arrays and the rank table take the place of the scene copy, which is
              // read from the input for the tied queries and reloaded afterwards (same thread -> same points).
              const double2* src = (const double2*)a.scene_xy + s0;
              __syncthreads();
              replay_ties_coop<kThreads>(a, s, Rt, ns, nsGA, tid, s.region0, a.mcap, (float*)rtprev_s + 16, src,
                                         (uint16_t*)(s.region0 + (size_t)a.mcap * 6));
              double2* dst = (double2*)s.region0;
              for (int i = tid; i < ns; i += kThreads) dst[i] = __ldg(src + i);
            }
            status |= CCICP_ST_TIE_REPLAY;
            sums_from_slots<kThreads>(s, Rt, ns, tid, pcx, pcy, v, cnt);
            publish_partials(s.red, v, cnt, tid);
            __syncthreads();
          }
#if CCICP_SOLVE_ALL
          // (experiment) every warp adds the partials and solves redundantly: no second barrier, 8x the FP64 tail
          int n;
          {
            n = gather_partials<kWarps>(s.red, v, tid & 31);
            if (n > 0) {
              const double dm0 = v[0] / (double)n, dm1 = v[1] / (double)n, dt0 = v[2] / (double)n, dt1 = v[3] / (double)n;
              const double mu_m[2] = {pcx + dm0, pcy + dm1};
              const double mu_t[2] = {pcx + dt0, pcy + dt1};
              const double H[4] = {v[4] - v[2] * dm0, v[5] - v[2] * dm1, v[6] - v[3] * dm0, v[7] - v[3] * dm1};
              if (tid == 0) {
#pragma unroll
                for (int q = 0; q < 6; q++) s.rtprev[q] = Rt[q];
              }
              last_delta = solve_and_compose(H, mu_m, mu_t, Rt, status);
            }
          }
          iters++;
          if (n == 0) {
            last_delta = -1.0; last_failed = true;
            if (!(-1.0 < a.min_delta)) iters = a.max_iter;
            break;
          }
          last_failed = false; have_ok = true; num_corr = n;
          if (kDiag) diag_inliers += (tid == 0) ? n : 0;
          if (last_delta < a.min_delta) { status |= CCICP_ST_CONVERGED; break; }
#else
          // ---- a7..a9 on ONE warp: totals, centroids, centred cross-covariance, solve, compose
          if (tid < 32) {
            const int n = gather_partials<kWarps>(s.red, v, tid);
            uint32_t st = 0;
            double delta = -1.0;
            double Rn[6];
#pragma unroll
            for (int q = 0; q < 6; q++) Rn[q] = Rt[q];
            if (n > 0) {
              if constexpr (kExact) {
                exact_step(s, Rt, ns, n, tid, exact_s);
                __syncwarp();
#pragma unroll
                for (int q = 0; q < 6; q++) Rn[q] = exact_s[8 + q];
                delta = exact_s[14];
                st = (uint32_t)exact_s[15];
              } else {
                const double dm0 = v[0] / (double)n, dm1 = v[1] / (double)n, dt0 = v[2] / (double)n, dt1 = v[3] / (double)n;
                const double mu_m[2] = {pcx + dm0, pcy + dm1};
                const double mu_t[2] = {pcx + dt0, pcy + dt1};
                const double H[4] = {v[4] - v[2] * dm0, v[5] - v[2] * dm1, v[6] - v[3] * dm0, v[7] - v[3] * dm1};
                delta = solve_and_compose(H, mu_m, mu_t, Rn, st);
              }
            }
            if (tid == 0) {
              if (n > 0) {
#pragma unroll
                for (int q = 0; q < 6; q++) s.rtprev[q] = Rt[q];     // pre-update R,t of the last successful iteration
              }
#pragma unroll
              for (int q = 0; q < 6; q++) sol_s[q] = Rn[q];
              sol_s[6] = delta;
              ((int*)&sol_s[7])[0] = n; ((uint32_t*)&sol_s[7])[1] = st;
            }
          }
          __syncthreads();
          const int n = ((const int*)&sol_s[7])[0];
          iters++;
          if (n == 0) {                          // fitStep returns -1 before touching R,t (icpPointToPoint.cpp:128-131)
            last_delta = -1.0; last_failed = true;
            if (!(-1.0 < a.min_delta)) iters = a.max_iter;   // every remaining fitStep repeats this one
            break;
          }
          last_failed = false; have_ok = true; num_corr = n;
#pragma unroll
          for (int q = 0; q < 6; q++) Rt[q] = sol_s[q];
          last_delta = sol_s[6];
          status |= ((const uint32_t*)&sol_s[7])[1];
          if (kDiag) diag_inliers += (tid == 0) ? n : 0;
          if (last_delta < a.min_delta) { status |= CCICP_ST_CONVERGED; break; }
#endif
        }
        if (!have_ok && iters > 0) status |= CCICP_ST_NO_CORR;   // max_iter <= 0: no fitStep ran, the reference reports nothing
      }

      // ---- outputs: correspondences of the last executed iteration
      __syncthreads();
      if (a.corr_idx) {
        int32_t* dst = a.corr_idx + (size_t)seed * (size_t)a.total_scene + s0;
        for (int i = tid; i < ns; i += kThreads) {
          int32_t o = -1;
          if (model_ok && scene_ok && fits && iters > 0) {
            const unsigned slot = s.slots[i];
            if (slot != kNoSlot) o = s.morig[slot];
          }
          dst[i] = o;
        }
      }

      // ---- score + a10 getEdgeWeight from the last successful iteration's correspondences
      double info[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
      double score = CUDART_INF;
      if (have_ok) {
        double P[6];
#pragma unroll
        for (int q = 0; q < 6; q++) P[q] = s.rtprev[q];
        if (last_failed) {       // regenerate the slots of the last successful iteration (ties: grid order)
          __syncthreads();
          for (int i = tid; i < ns; i += kThreads) {
            const double2 sp = s.scene[i];
            float qx, qy;
            transform(P, sp.x, sp.y, qx, qy);
            const GridP& g = s.grid[i >= nsGA];
            unsigned slot = kNoSlot;
            if (g.G > 0) {
              float best; int bslot, tie;
              if constexpr (kCurv) {
                const float qc = scurv[i];
                nn_search<kGlobal, false, true, kLoops>(s.mpts, s.cstart, g, qx, qy, gate_f, kNoSlot, best, bslot, tie, diag_evals, mcurv, qc, ctol);
                if (tie) {
                  const QPack q2 = pack_query(qx, qy);
                  int bo = 0x7fffffff;
                  for (int j = s.cstart[g.base]; j < s.cstart[g.base + g.G * g.G]; j++)
                    if (!(fabsf(mcurv[j] - qc) > ctol) && dist2(s.mpts[j], q2) == best && (int)s.morig[j] < bo) { bo = s.morig[j]; bslot = j; }
                }
              } else {
                nn_search<kGlobal, false, false, kLoops>(s.mpts, s.cstart, g, qx, qy, gate_f, kNoSlot, best, bslot, tie, diag_evals);
              }
              if (bslot >= 0 && (double)best < a.indist) slot = (unsigned)bslot;
              if constexpr (kTrim) trim_keys[i] = slot != kNoSlot ? __float_as_uint(best) : 0xFFFFFFFFu;
            } else if constexpr (kTrim) trim_keys[i] = 0xFFFFFFFFu;
            s.slots[i] = (uint16_t)slot;
          }
          if constexpr (kTrim) trim_select<kThreads>(a, s, ns, tid, trim_s, trim_keys);
        }
        const bool want_info = (a.flags & CCICP_F_INFO) != 0;
        const bool fixed_dy = (a.flags & CCICP_F_INFO_FIXED_DY) != 0;
        double v[7] = {0, 0, 0, 0, 0, 0, 0};   // sdis, sx, sy, xpy, mz0, mz1, mz2
        for (int i = tid; i < ns; i += kThreads) {
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
        block_reduce<7, kWarps>(s.red, 0, v, tid);
        const double nn = (double)num_corr;
        score = v[0] / nn;
        if (want_info) {
          // MM = [[n,0,-sy],[0,n,sx],[-sy,sx,xpy]]; D = MM^-1 * MZ (icpPointToPoint.cpp:277-285)
          const double sx = v[1], sy = v[2], xpy = v[3];
          const double sch = xpy - (sx * sx + sy * sy) / nn;
          const double D2 = (v[6] + (sy * v[4] - sx * v[5]) / nn) / sch;
          const double D0 = (v[4] + sy * D2) / nn, D1 = (v[5] - sx * D2) / nn;
          double ss[1] = {0.0};
          __syncthreads();
          for (int i = tid; i < ns; i += kThreads) {
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
          block_reduce<1, kWarps>(s.red, 8, ss, tid);
          const double sconst = 1.0 / (ss[0] / (double)(2 * num_corr - 3));
          if (!(sch != 0.0) || !(ss[0] > 0.0)) status |= CCICP_ST_INFO_SINGULAR;
          info[0] = nn * sconst; info[1] = 0.0 * sconst; info[2] = -sy * sconst;
          info[3] = 0.0 * sconst; info[4] = nn * sconst; info[5] = sx * sconst;
          info[6] = -sy * sconst; info[7] = sx * sconst; info[8] = xpy * sconst;
        }
      }
      if (tid == 0) {
        if (s.misc[M_TREEFAIL]) status |= CCICP_ST_TREE_OVERFLOW;
        ccicp_result r;
#pragma unroll
        for (int q = 0; q < 4; q++) r.R[q] = Rt[q];
        r.t[0] = Rt[4]; r.t[1] = Rt[5];
#pragma unroll
        for (int q = 0; q < 9; q++) r.info[q] = info[q];
        r.last_delta = last_delta; r.score = score;
        r.num_corr = num_corr; r.iters = iters; r.status = status; r.best_seed = seed;
        r.n_tie_replays = s.misc[M_REPLAYS]; r.reserved = 0;
        a.out[prob] = r;
      }
    }
  }
  if constexpr (kDiag) {      // diagnostics instantiation: totals of this launch (bench.py's roofline_useful)
    const unsigned long long ev = (unsigned long long)__reduce_add_sync(0xffffffffu, (unsigned)diag_evals);
    const unsigned long long qu = (unsigned long long)__reduce_add_sync(0xffffffffu, (unsigned)diag_queries);
    const unsigned long long in = (unsigned long long)__reduce_add_sync(0xffffffffu, (unsigned)diag_inliers);
    if ((tid & 31) == 0 && a.diag) { atomicAdd(a.diag + 0, ev); atomicAdd(a.diag + 1, qu); atomicAdd(a.diag + 2, in); }
  }
}

__global__ void __launch_bounds__(kThreadsOnChip, CCICP_MIN_BLOCKS) icp_pair_kernel(const IcpArgs a) {
  icp_pair_body<kThreadsOnChip, false, false>(a);
}

#ifndef CCICP_SEEDS_LOOPS
#define CCICP_SEEDS_LOOPS 1
#endif
// Multi-seed launches (config C5: far hypotheses, cold starts): the candidate loop of long rows (nn_grid.h, kLoops = 1).
__global__ void __launch_bounds__(kThreadsOnChip, CCICP_MIN_BLOCKS) icp_pair_seeds_kernel(const IcpArgs a) {
  icp_pair_body<kThreadsOnChip, false, false, false, false, false, CCICP_SEEDS_LOOPS>(a);
}

// Large scans (config C4): same algorithm, working set in the CTA's global workspace (L2/L1 resident).
__global__ void __launch_bounds__(kThreadsL2, 1) icp_pair_l2_kernel(const IcpArgs a) {
  icp_pair_body<kThreadsL2, true, false>(a);
}

// Diagnostics instantiation (ccicp_params.flags & CCICP_F_DIAG_COUNTS): same algorithm, counts distance evaluations,
// searched queries and accepted pairs of the launch; never the timed kernel.
__global__ void __launch_bounds__(kThreadsOnChip, CCICP_MIN_BLOCKS) icp_pair_diag_kernel(const IcpArgs a) {
  icp_pair_body<kThreadsOnChip, false, false, false, true>(a);
}
__global__ void __launch_bounds__(kThreadsL2, 1) icp_pair_l2_diag_kernel(const IcpArgs a) {
  icp_pair_body<kThreadsL2, true, false, false, true>(a);
}

// CCICP_F_CURV_GATE instantiations (north_star extra: curvature-gated correspondence search)
__global__ void __launch_bounds__(kThreadsOnChip, CCICP_MIN_BLOCKS) icp_pair_curv_kernel(const IcpArgs a) {
  icp_pair_body<kThreadsOnChip, false, false, false, false, true>(a);
}
__global__ void __launch_bounds__(kThreadsL2, 1) icp_pair_l2_curv_kernel(const IcpArgs a) {
  icp_pair_body<kThreadsL2, true, false, false, false, true>(a);
}

// CCICP_F_EXACT_SUMS instantiations (separate kernels so the default ones carry none of their register pressure).
__global__ void __launch_bounds__(kThreadsOnChip, CCICP_MIN_BLOCKS) icp_pair_exact_kernel(const IcpArgs a) {
  icp_pair_body<kThreadsOnChip, false, true>(a);
}
__global__ void __launch_bounds__(kThreadsL2, 1) icp_pair_l2_exact_kernel(const IcpArgs a) {
  icp_pair_body<kThreadsL2, true, true>(a);
}
// CCICP_F_TRIM instantiations (north_star extra: trimmed outlier rejection)
__global__ void __launch_bounds__(kThreadsOnChip, CCICP_MIN_BLOCKS) icp_pair_trim_kernel(const IcpArgs a) {
  icp_pair_body<kThreadsOnChip, false, false, true>(a);
}
__global__ void __launch_bounds__(kThreadsL2, 1) icp_pair_l2_trim_kernel(const IcpArgs a) {
  icp_pair_body<kThreadsL2, true, false, true>(a);
}

// Problems of this launch whose default-mode solve met a (nearly) rank-deficient H: listed for the exact-order re-run.
__global__ void collect_retry_kernel(const ccicp_result* __restrict__ res, int first_prob, int n_prob, int32_t* __restrict__ list,
                                     int32_t* __restrict__ count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_prob) return;
  if (res[first_prob + i].status & CCICP_ST_DEGENERATE) list[atomicAdd(count, 1)] = first_prob + i;
}

// best-of-seeds: one warp per pair.  Default: arg-min of score over seeds with num_corr >= min_corr_frac * n_scene,
// ties -> lowest seed index.  first_below >= 0 (CCICP_F_FIRST_BELOW): the lowest seed index whose normalised score
// score / n_scene is below the threshold (global_match.cpp:131-151); if there is none, the arg-min with
// CCICP_ST_NO_SEED_BELOW.
__global__ void select_best_seed_kernel(const ccicp_result* __restrict__ all, ccicp_result* __restrict__ out,
                                        const int32_t* __restrict__ scene_off, const int32_t* __restrict__ scene_id,
                                        int n_pairs, int n_seeds, double min_corr_frac, double first_below) {
  const int pair = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (pair >= n_pairs) return;
  const int sid = scene_id ? scene_id[pair] : pair;
  const double n_scene = (double)(scene_off[2 * sid + 2] - scene_off[2 * sid]);
  const double need = min_corr_frac * n_scene;
  double best = CUDART_INF;
  int bseed = 0x7fffffff, fseed = 0x7fffffff;
  for (int sd = lane; sd < n_seeds; sd += 32) {
    const ccicp_result& r = all[(size_t)pair * n_seeds + sd];
    double sc = ((double)r.num_corr >= need && r.num_corr > 0) ? r.score : CUDART_INF;
    if (sc < best) { best = sc; bseed = sd; }      // ascending sd per lane: first minimum kept
    if (first_below >= 0.0 && sc / n_scene < first_below && sd < fseed) fseed = sd;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    double ob = __shfl_xor_sync(0xffffffffu, best, o);
    int os = __shfl_xor_sync(0xffffffffu, bseed, o);
    if (ob < best || (ob == best && os < bseed)) { best = ob; bseed = os; }
    fseed = min(fseed, __shfl_xor_sync(0xffffffffu, fseed, o));
  }
  if (bseed == 0x7fffffff) bseed = 0;               // nothing qualified: report seed 0
  const bool want_first = first_below >= 0.0;
  const int pick = (want_first && fseed != 0x7fffffff) ? fseed : bseed;
  const ccicp_result* src = &all[(size_t)pair * n_seeds + pick];
  ccicp_result* dst = &out[pair];
  const int4* s4 = (const int4*)src;
  int4* d4 = (int4*)dst;
  for (int q = lane; q < (int)(sizeof(ccicp_result) / 16); q += 32) d4[q] = s4[q];
  __syncwarp();
  if (lane == 0) {
    if (!(best < CUDART_INF) && pick == bseed) dst->score = CUDART_INF;
    if (want_first && fseed == 0x7fffffff) dst->status |= CCICP_ST_NO_SEED_BELOW;
  }
}

}  // namespace ccicp
