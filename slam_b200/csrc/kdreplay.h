// kdreplay.h -- tie-resolution slow path: a from-scratch restatement of the
// reference's deterministic kd-tree (M. Kennel's kdtree2 as vendored in
// ccicp2d), used ONLY for queries whose float-minimum distance is shared by two
// or more model points.  For those queries the reference's answer depends on
// its traversal order (a later-visited candidate with dis <= best replaces the
// incumbent, ccicp2d/src/kdtree.cpp:613-619, 666-675), so the exact tree
// (build: kdtree.cpp:114-233, 235-269, 294-317) and the exact traversal
// (search: kdtree.cpp:515-581, leaf scan: :583-683 with nn == 1) are replayed.
//
// The file compiles both as CUDA device code (included by ccicp_kernels.cu,
// which is built with -fmad=false so no float expression is contracted) and as
// plain host C++ (tests/host/kdreplay_host.cpp) so the logic is checked on the
// CPU against the compiled reference without a GPU.  It is product code, not
// part of oracle/.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define KD_HD __host__ __device__ __forceinline__
#define KD_FN __host__ __device__
#define KD_COLD __host__ __device__ __noinline__
#else
#define KD_HD inline
#define KD_FN inline
#define KD_COLD inline
#endif

namespace kdreplay {

struct Pt { float x, y; };

enum { BUCKET = 12, STACK = 64 };
enum { OK = 0, ERR_DEPTH = 1, ERR_NODES = 2, ERR_DEGENERATE = 3 };

// Point access goes through an accessor `P` with `Pt operator()(int original_index) const`, and the index
// permutation through any integer array type `I` (the device build keeps both in shared memory).
struct PtrPts {
  const Pt* p;
  KD_HD Pt operator()(int i) const { return p[i]; }
};
template <class P>
KD_HD float coord(const P& p, int i, int c) { Pt v = p(i); return c == 0 ? v.x : v.y; }

// min/max of coordinate c over ind[l..u] (kdtree.cpp:235-269; min/max are
// order independent so the pairwise walk is not reproduced).
template <class P, class I>
KD_HD void spread(const P& pts, const I* ind, int c, int l, int u, float* lo, float* hi) {
  float smin = coord(pts, ind[l], c), smax = smin;
  for (int i = l + 1; i <= u; i++) {
    float v = coord(pts, ind[i], c);
    if (v < smin) smin = v;
    if (v > smax) smax = v;
  }
  *lo = smin; *hi = smax;
}

KD_HD float dis_from_bnd(float x, float amin, float amax) {
  if (x > amax) return x - amax;
  if (x < amin) return amin - x;
  return 0.0f;
}

// ---------------------------------------------------------------------------------------------------------
// Lazy variant used by the kernel: the tree is materialised only along the paths the tied queries actually
// traverse.  A node is split the first time a search reaches it; the split reproduces exactly what the eager
// build would have produced for that node because (a) the partition of a node depends only on the order its
// ancestors' partitions left in ind[l..u], (b) the "approximate box" a node inherits (kdtree.cpp:153-168) is
// carried down explicitly in abox, and (c) every bounding box the search consults (cut_val_left/right,
// box_in_search_range) is the TRUE box of a range, which is computed directly from the points of that range.
struct LNode {
  float tbox[4];             // true bounding box of ind[l..u]: {x.lo, x.hi, y.lo, y.hi}
  float abox[4];             // the box the eager build would hold for this node while choosing its cut dimension
  float cut_val, cut_val_left, cut_val_right;
  int32_t cut_dim;           // -1 terminal, -2 internal but not split yet, 0/1 split on that coordinate
  int32_t l, u, left, right;
};
enum { LEAF = -1, UNSPLIT = -2 };

template <class P, class I>
KD_FN int lazy_split(const P& pts, I* ind, LNode* nodes, int max_nodes, int* n_nodes, int id) {
  LNode nd = nodes[id];
  const int l = nd.l, u = nd.u;
  int c = -1;
  float maxspread = 0.0f;
  for (int i = 0; i < 2; i++) {
    float sprd = nd.abox[2 * i + 1] - nd.abox[2 * i];
    if (sprd > maxspread) { maxspread = sprd; c = i; }
  }
  if (c < 0) return ERR_DEGENERATE;
  float sum = 0.0f;
  for (int k = l; k <= u; k++) sum += coord(pts, ind[k], c);
  const float average = sum / (float)(u - l + 1);
  int lb = l, ub = u;
  while (lb < ub) {
    if (coord(pts, ind[lb], c) <= average) {
      lb++;
    } else {
      I tmp = ind[lb]; ind[lb] = ind[ub]; ind[ub] = tmp;
      ub--;
    }
  }
  const int m = (coord(pts, ind[lb], c) <= average) ? lb : lb - 1;
  int child[2] = {-1, -1};
  for (int sde = 0; sde < 2; sde++) {
    const int cl = sde ? m + 1 : l, cu = sde ? u : m;
    if (cu < cl) continue;
    if (cl == l && cu == u) return ERR_DEGENERATE;       // the reference would recurse forever here
    if (*n_nodes >= max_nodes) return ERR_NODES;
    LNode ch;
    ch.l = cl; ch.u = cu; ch.left = ch.right = -1;
    ch.cut_val = ch.cut_val_left = ch.cut_val_right = 0.0f;
    spread(pts, ind, 0, cl, cu, &ch.tbox[0], &ch.tbox[1]);
    spread(pts, ind, 1, cl, cu, &ch.tbox[2], &ch.tbox[3]);
    for (int i = 0; i < 4; i++) ch.abox[i] = nd.abox[i];
    ch.abox[2 * c] = ch.tbox[2 * c]; ch.abox[2 * c + 1] = ch.tbox[2 * c + 1];
    ch.cut_dim = (cu - cl <= BUCKET) ? LEAF : UNSPLIT;
    child[sde] = *n_nodes;
    nodes[(*n_nodes)++] = ch;
  }
  nd.left = child[0]; nd.right = child[1];
  if (nd.right < 0) {
    nd.cut_val = nodes[nd.left].tbox[2 * c + 1];
    nd.cut_val_left = nd.cut_val_right = nd.cut_val;
  } else if (nd.left < 0) {
    nd.cut_val = nodes[nd.right].tbox[2 * c + 1];
    nd.cut_val_left = nd.cut_val_right = nd.cut_val;
  } else {
    nd.cut_val_right = nodes[nd.right].tbox[2 * c];
    nd.cut_val_left = nodes[nd.left].tbox[2 * c + 1];
    nd.cut_val = (float)((double)(nd.cut_val_left + nd.cut_val_right) / 2.0);
  }
  nd.cut_dim = c;
  nodes[id] = nd;
  return OK;
}

KD_HD bool lbox_in_range(const LNode& nd, float qx, float qy, float ballsize) {
  float d0 = dis_from_bnd(qx, nd.tbox[0], nd.tbox[1]);
  float dis2 = 0.0f;
  dis2 += d0 * d0;
  if (dis2 > ballsize) return false;
  float d1 = dis_from_bnd(qy, nd.tbox[2], nd.tbox[3]);
  dis2 += d1 * d1;
  if (dis2 > ballsize) return false;
  return true;
}

// ---------------------------------------------------------------------------------------------------------
// Cooperative variant (large scans, L2 kernel): every thread of the CTA runs the SAME traversal on the same
// query (all control flow is uniform); a node split is shared out: the gather of the cut coordinate and the
// bounding boxes of the children are data-parallel, and so is the in-order partition (kdtree.cpp:294-317) in its
// rank-table form (part_* below); only the float mean (a sequential sum, kdtree.cpp:181-186) runs on one thread.
// Everything works on a staged copy (`sv`, `si`).
//
// `C` (the cooperation policy) provides:
//   int tid, nt;                      this thread, threads in the group
//   float* sv; uint16_t* si; int cap; staging arrays for `cap` elements (shared memory on the device)
//   float* gsv; uint16_t* gsi; int gcap; a second, larger pair for ranges beyond `cap` (global memory; gcap = 0: none)
//   int* bcast;                       one int visible to the whole group
//   void sync();                      group barrier (also orders memory)
//   void minmax4(float* lo, float* hi) group-wide min of lo[0..3] / max of hi[0..3], result in every thread
//   int exscan(int v, int* total)     group-wide exclusive prefix sum over tid, the total in every thread (has barriers)
//   uint16_t* rank;                   `cap` entries for the partition's rank tables (NULL: sequential partition on thread 0)
// With nt == 1 (host build, tests/host/kdreplay_host.cpp) this is a plain sequential program.
template <class P, class I, class C>
KD_FN void coop_boxes(const P& pts, const uint16_t* si, const float* sv, int c, int k0, int k1, C& co, float* tbox) {
  float lo[4] = {3.0e38f, 3.0e38f, 3.0e38f, 3.0e38f}, hi[4] = {-3.0e38f, -3.0e38f, -3.0e38f, -3.0e38f};
  for (int k = k0 + co.tid; k <= k1; k += co.nt) {
    const float vc = sv[k];
    const float vo = coord(pts, (int)si[k], 1 - c);
    const float x = c == 0 ? vc : vo, y = c == 0 ? vo : vc;
    lo[0] = x < lo[0] ? x : lo[0]; hi[0] = x > hi[0] ? x : hi[0];
    lo[1] = y < lo[1] ? y : lo[1]; hi[1] = y > hi[1] ? y : hi[1];
  }
  co.minmax4(lo, hi);
  tbox[0] = lo[0]; tbox[1] = hi[0]; tbox[2] = lo[1]; tbox[3] = hi[1];
}

template <class P, class I, class C>
KD_FN int lazy_init_coop(const P& pts, int n, I* ind, LNode* nodes, int* n_nodes, C& co) {
  for (int i = co.tid; i < n; i += co.nt) ind[i] = (I)i;
  if (n <= 0) { if (co.tid == 0) *n_nodes = 0; co.sync(); return OK; }
  float lo[4] = {3.0e38f, 3.0e38f, 3.0e38f, 3.0e38f}, hi[4] = {-3.0e38f, -3.0e38f, -3.0e38f, -3.0e38f};
  for (int i = co.tid; i < n; i += co.nt) {
    const Pt v = pts(i);
    lo[0] = v.x < lo[0] ? v.x : lo[0]; hi[0] = v.x > hi[0] ? v.x : hi[0];
    lo[1] = v.y < lo[1] ? v.y : lo[1]; hi[1] = v.y > hi[1] ? v.y : hi[1];
  }
  co.minmax4(lo, hi);
  if (co.tid == 0) {
    LNode r;
    r.l = 0; r.u = n - 1; r.left = r.right = -1;
    r.cut_val = r.cut_val_left = r.cut_val_right = 0.0f;
    r.tbox[0] = lo[0]; r.tbox[1] = hi[0]; r.tbox[2] = lo[1]; r.tbox[3] = hi[1];
    for (int i = 0; i < 4; i++) r.abox[i] = r.tbox[i];
    r.cut_dim = (n - 1 <= BUCKET) ? LEAF : UNSPLIT;
    nodes[0] = r;
    *n_nodes = 1;
  }
  co.sync();
  return OK;
}

// The reference's in-place partition (select_on_coordinate_value, kdtree.cpp:294-317)
//     lb = l, ub = u;  while (lb < ub) { if (v[lb] <= alpha) lb++; else { swap(lb, ub); ub--; } }
// examines every element exactly once, in an order that only depends on the outcomes: after an element that stays
// (<= alpha) the next one examined is the next untouched element from the FRONT, after one that is moved away the
// next is the next untouched element from the BACK (the one just swapped in).  Kept elements end up at the front in
// examination order, moved ones at the back in examination order from the right end, and the split index is
// (#kept - 1).  That is what this branch-free formulation writes (both streams are pre-loaded, so the serial chain is
// compare -> select instead of load -> compare -> branch -> swap).  In place: a front slot is written only after its
// element was consumed, a back slot in the same step its element is consumed (already held in a register).
KD_HD int partition_stream(float* v, uint16_t* ix, int n, float alpha) {
  if (n <= 0) return -1;
  int f = 1, bk = n - 1;                 // next untouched element of the front / back stream
  int gl = 0, br = n - 1;                // next write slot for kept / moved elements
  float cv = v[0]; uint16_t ci = ix[0];
  float fv = n > 1 ? v[1] : 0.0f; uint16_t fi = n > 1 ? ix[1] : (uint16_t)0;
  float bv = v[n - 1]; uint16_t bi = ix[n - 1];
  for (int t = 0; t < n; t++) {
    const bool keep = cv <= alpha;
    const int dst = keep ? gl : br;
    v[dst] = cv; ix[dst] = ci;
    gl += keep ? 1 : 0; br -= keep ? 0 : 1;
    cv = keep ? fv : bv; ci = keep ? fi : bi;
    if (keep) { f++; if (f < n) { fv = v[f]; fi = ix[f]; } }
    else { bk--; if (bk >= 0) { bv = v[bk]; bi = ix[bk]; } }
  }
  return gl - 1;
}

// ---- data-parallel form of the same partition ---------------------------------------------------------------
// The examination order above is an interleaving of two fixed streams -- the front stream a[0], a[1], ... and the
// back stream a[n-1], a[n-2], ... -- so where every element ends up follows from prefix counts alone.  Call an
// element "big" if !(v <= alpha) (it is moved away) and "kept" otherwise.  The front stream advances after every
// kept element, so front element i is examined right after the i-th keep: a kept front element stays at slot i.
// The k-th big front element (index P[k]) starts round k: it goes to the slot just left of the (k-1)-th kept
// element from the right (index Q[k-1], Q[0] = n), then the back stream is consumed down to the k-th kept element
// from the right (Q[k]); the big elements passed on the way each move one slot to the left and a[Q[k]] lands in
// slot P[k].  Round k starts iff P[k] < Q[k-1] (monotone in k: K = number of rounds that start).  The last round
// either completes (Q[K] > P[K]: the front stream then runs over kept elements up to F = Q[K]) or the back stream
// meets the front inside it (F = P[K] + 1).  Elements below F belong to the front stream, the others to the back
// stream.  T[0..NB) = P[1..NB], T[NB..n) = Q[1..NL] are the two rank tables (NB big, NL = n - NB kept elements).
// Checked against partition_stream() element for element in tests/test_kdreplay_cpu.py.
KD_HD bool part_big(float v, float alpha) { return !(v <= alpha); }

KD_HD int part_count(const float* v, int k0, int k1, float alpha) {
  int c = 0;
  for (int k = k0; k < k1; k++) c += part_big(v[k], alpha) ? 1 : 0;
  return c;
}

// bb0 = big elements in [0, k0)
KD_HD void part_fill(const float* v, int k0, int k1, float alpha, int bb0, int NB, int n, uint16_t* T) {
  const int NL = n - NB;
  int bb = bb0;
  for (int x = k0; x < k1; x++) {
    if (part_big(v[x], alpha)) T[bb++] = (uint16_t)x;
    else T[NB + (NL - (x - bb) - 1)] = (uint16_t)x;
  }
}

KD_HD bool part_round_starts(const uint16_t* T, int NB, int n, int k) {      // 1 <= k <= NB
  const int NL = n - NB;
  const int Pk = (int)T[k - 1];
  const int Qk1 = k == 1 ? n : (k - 1 <= NL ? (int)T[NB + k - 2] : -1);
  return Pk < Qk1;
}

KD_HD int part_front(const uint16_t* T, int NB, int n, int K) {
  if (K == 0) return n;
  const int NL = n - NB;
  const int PK = (int)T[K - 1];
  if (K <= NL) { const int QK = (int)T[NB + K - 1]; if (QK > PK) return QK; }
  return PK + 1;
}

// slot of element x after the partition; bb = big elements in [0, x)
KD_HD int part_dest(const uint16_t* T, int NB, int n, int F, int x, bool big, int bb) {
  if (x < F) {
    if (!big) return x;
    return (bb == 0 ? n : (int)T[NB + bb - 1]) - 1;          // Q[k-1] - 1 with k = bb + 1
  }
  if (big) return x - 1;
  const int NL = n - NB;
  return (int)T[NL - (x - bb) - 1];                           // P[k] with k = (kept elements right of x) + 1
}

KD_HD int kd_f2i(float f) { union { float f; int i; } u; u.f = f; return u.i; }
KD_HD float kd_i2f(int i) { union { float f; int i; } u; u.i = i; return u.f; }

// child boxes by membership (kept -> left child, big -> right child); the order inside a child does not matter
template <class P, class I, class C>
KD_FN void coop_boxes_split(const P& pts, const uint16_t* si, const float* sv, int c, int n, float alpha, C& co,
                            float* lbox, float* rbox) {
  float lo[4] = {3.0e38f, 3.0e38f, 3.0e38f, 3.0e38f}, hi[4] = {-3.0e38f, -3.0e38f, -3.0e38f, -3.0e38f};
  for (int k = co.tid; k < n; k += co.nt) {
    const float vc = sv[k];
    const float vo = coord(pts, (int)si[k], 1 - c);
    const float x = c == 0 ? vc : vo, y = c == 0 ? vo : vc;
    const bool big = part_big(vc, alpha);                   // (no dynamic indexing: the arrays stay in registers)
    const float xl = big ? 3.0e38f : x, yl = big ? 3.0e38f : y, xr = big ? x : 3.0e38f, yr = big ? y : 3.0e38f;
    const float xL = big ? -3.0e38f : x, yL = big ? -3.0e38f : y, xR = big ? x : -3.0e38f, yR = big ? y : -3.0e38f;
    lo[0] = xl < lo[0] ? xl : lo[0]; lo[1] = yl < lo[1] ? yl : lo[1]; lo[2] = xr < lo[2] ? xr : lo[2]; lo[3] = yr < lo[3] ? yr : lo[3];
    hi[0] = xL > hi[0] ? xL : hi[0]; hi[1] = yL > hi[1] ? yL : hi[1]; hi[2] = xR > hi[2] ? xR : hi[2]; hi[3] = yR > hi[3] ? yR : hi[3];
  }
  co.minmax4(lo, hi);
  lbox[0] = lo[0]; lbox[1] = hi[0]; lbox[2] = lo[1]; lbox[3] = hi[1];
  rbox[0] = lo[2]; rbox[1] = hi[2]; rbox[2] = lo[3]; rbox[3] = hi[3];
}

// The group-wide partition of the staged range: writes the moved elements' indices to out[] (out[x] already holds
// si[x]) and returns the split index.  Out of line: it is the rare path of a register-tight kernel.
template <class I, class C>
KD_COLD int partition_ranked(const float* sv, const uint16_t* si, int n, float alpha, I* out, C& co) {
  // every thread owns a contiguous run of elements (odd length: conflict-free shared memory)
  const int chunk = ((n + co.nt - 1) / co.nt) | 1;
  const int k0 = co.tid * chunk < n ? co.tid * chunk : n, k1 = k0 + chunk < n ? k0 + chunk : n;
  uint16_t* T = co.rank;
  int NB = 0;
  const int bb0 = co.exscan(part_count(sv, k0, k1, alpha), &NB);
  part_fill(sv, k0, k1, alpha, bb0, NB, n, T);
  co.sync();
  int kc = 0;
  for (int k = 1 + co.tid; k <= NB; k += co.nt) kc += part_round_starts(T, NB, n, k) ? 1 : 0;
  int K = 0;
  co.exscan(kc, &K);
  const int F = part_front(T, NB, n, K);
  int bb = bb0;
  for (int x = k0; x < k1; x++) {
    const bool big = part_big(sv[x], alpha);
    const int d = part_dest(T, NB, n, F, x, big, bb);
    if (d != x) out[d] = (I)si[x];
    bb += big ? 1 : 0;
  }
  return n - NB - 1;
}

// Splits node `id` (uniform call).  Ranges larger than the staging capacity fall back to the sequential split
// on thread 0.
template <class P, class I, class C>
KD_FN int lazy_split_coop(const P& pts, I* ind, LNode* nodes, int max_nodes, int* n_nodes, int id, C& co) {
  const LNode nd0 = nodes[id];
  const int l = nd0.l, u = nd0.u, n = u - l + 1;
  // staging arrays: the fast ones (shared memory on the device) when the range fits, else the group's slower, larger
  // pair (global memory: the top one or two splits of a large class); only a range beyond both is split sequentially
  float* sv = co.sv;
  uint16_t* si = co.si;
  if (n > co.cap) {
    if (n <= co.gcap) { sv = co.gsv; si = co.gsi; }
    else {
      if (co.tid == 0) *co.bcast = lazy_split(pts, ind, nodes, max_nodes, n_nodes, id);
      co.sync();
      const int rc = *co.bcast;
      co.sync();
      return rc;
    }
  }
  int c = -1;
  float maxspread = 0.0f;
  for (int i = 0; i < 2; i++) {
    float sprd = nd0.abox[2 * i + 1] - nd0.abox[2 * i];
    if (sprd > maxspread) { maxspread = sprd; c = i; }
  }
  if (c < 0) return ERR_DEGENERATE;
  for (int k = co.tid; k < n; k += co.nt) {                 // gather the cut coordinate (data-parallel)
    const int o = (int)ind[l + k];
    si[k] = (uint16_t)o;
    sv[k] = coord(pts, o, c);
  }
  co.sync();
  // the float mean is a sequential sum by definition (:181-186): one thread adds, in order.  A range staged in the second
  // (global-memory) pair is fed to it through the fast arrays tile by tile -- a dependent add per global load would make
  // the top split of a 30000-point class cost as much as the rest of the replay.
  {
    float sum = 0.0f;                                       // meaningful on thread 0 only
    const bool tiled = sv != co.sv;
    for (int base = 0; base < n; base += tiled ? co.cap : n) {
      const int m = tiled ? (n - base < co.cap ? n - base : co.cap) : n;
      const float* tv = sv;
      if (tiled) {
        for (int k = co.tid; k < m; k += co.nt) co.sv[k] = sv[base + k];
        co.sync();
        tv = co.sv;
      }
      if (co.tid == 0) {
        int k = 0;
        for (; k + 8 <= m; k += 8) {                        // loads first, then the add chain in order
          const float a0 = tv[k], a1 = tv[k + 1], a2 = tv[k + 2], a3 = tv[k + 3];
          const float a4 = tv[k + 4], a5 = tv[k + 5], a6 = tv[k + 6], a7 = tv[k + 7];
          sum += a0; sum += a1; sum += a2; sum += a3; sum += a4; sum += a5; sum += a6; sum += a7;
        }
        for (; k < m; k++) sum += tv[k];
      }
      if (tiled) co.sync();
    }
    if (co.tid == 0) *co.bcast = kd_f2i(sum / (float)n);
  }
  co.sync();
  const float alpha = kd_i2f(*co.bcast);
  int m;
  bool by_flag = false;
  if (co.rank != nullptr) {
    m = partition_ranked<I, C>(sv, si, n, alpha, ind + l, co);
    by_flag = true;
  } else {
    if (co.tid == 0) *co.bcast = partition_stream(sv, si, n, alpha);
    co.sync();
    m = *co.bcast;                                          // last index (staged numbering) of the left child
    for (int k = co.tid; k < n; k += co.nt) ind[l + k] = (I)si[k];
  }
  if (m >= n - 1 || m < 0) {
    // one side empty: the other child would be the node itself -> the reference recurses forever
    co.sync();
    return ERR_DEGENERATE;
  }
  if (*n_nodes + 2 > max_nodes) { co.sync(); return ERR_NODES; }
  float lbox[4], rbox[4];
  if (by_flag) {
    coop_boxes_split<P, I, C>(pts, si, sv, c, n, alpha, co, lbox, rbox);
  } else {
    coop_boxes<P, I, C>(pts, si, sv, c, 0, m, co, lbox);
    coop_boxes<P, I, C>(pts, si, sv, c, m + 1, n - 1, co, rbox);
  }
  co.sync();                                                // everyone is done with the staging arrays and *n_nodes
  if (co.tid == 0) {
    LNode nd = nd0;
    const int base = *n_nodes;
    for (int sde = 0; sde < 2; sde++) {
      LNode ch;
      ch.l = sde ? l + m + 1 : l; ch.u = sde ? u : l + m; ch.left = ch.right = -1;
      ch.cut_val = ch.cut_val_left = ch.cut_val_right = 0.0f;
      for (int i = 0; i < 4; i++) { ch.tbox[i] = sde ? rbox[i] : lbox[i]; ch.abox[i] = nd.abox[i]; }
      ch.abox[2 * c] = ch.tbox[2 * c]; ch.abox[2 * c + 1] = ch.tbox[2 * c + 1];
      ch.cut_dim = (ch.u - ch.l <= BUCKET) ? LEAF : UNSPLIT;
      nodes[base + sde] = ch;
    }
    nd.left = base; nd.right = base + 1;
    nd.cut_val_right = rbox[2 * c];
    nd.cut_val_left = lbox[2 * c + 1];
    nd.cut_val = (float)((double)(nd.cut_val_left + nd.cut_val_right) / 2.0);
    nd.cut_dim = c;
    nodes[id] = nd;
    *n_nodes = base + 2;
  }
  co.sync();
  return OK;
}

// Exact replay of n_nearest(q, 1), executed redundantly (and uniformly) by every thread of the group; only the
// node splits are shared out.  Same traversal as lazy_nearest().
template <class P, class I, class C>
KD_FN int lazy_nearest_coop(const P& pts, I* ind, LNode* nodes, int max_nodes, int* n_nodes,
                            float qx, float qy, int32_t* idx_out, float* dis_out, C& co) {
  struct Pending { int node; float extra; };
  Pending st[STACK];
  int sp = 0;
  float ballsize = 1.0e38f;
  bool have = false;
  int32_t best = -1;
  float bestd = 0.0f;
  *idx_out = -1; *dis_out = 0.0f;
  if (*n_nodes <= 0) return OK;
  int cur = 0;
  for (;;) {
    while (cur >= 0 && nodes[cur].cut_dim != LEAF) {
      if (nodes[cur].cut_dim == UNSPLIT) {
        int rc = lazy_split_coop(pts, ind, nodes, max_nodes, n_nodes, cur, co);
        if (rc != OK) return rc;
      }
      const LNode nd = nodes[cur];
      float qval = nd.cut_dim == 0 ? qx : qy;
      int closer, farther;
      float extra;
      if (qval < nd.cut_val) { closer = nd.left; farther = nd.right; extra = nd.cut_val_right - qval; }
      else { closer = nd.right; farther = nd.left; extra = qval - nd.cut_val_left; }
      if (sp >= STACK) return ERR_DEPTH;
      st[sp].node = farther; st[sp].extra = extra; sp++;
      cur = closer;
    }
    if (cur >= 0) {
      const int l = nodes[cur].l, u = nodes[cur].u;
      for (int i = l; i <= u; i++) {
        int o = (int)ind[i];
        const Pt po = pts(o);
        float dx = po.x - qx;
        float dis = 0.0f;
        dis += dx * dx;
        if (dis > ballsize) continue;
        float dy = po.y - qy;
        dis += dy * dy;
        if (dis > ballsize) continue;
        have = true; best = o; bestd = dis; ballsize = dis;
      }
    }
    bool found = false;
    while (sp > 0) {
      sp--;
      int far_ = st[sp].node;
      float extra = st[sp].extra;
      if (far_ >= 0 && extra * extra < ballsize && lbox_in_range(nodes[far_], qx, qy, ballsize)) {
        cur = far_; found = true; break;
      }
    }
    if (!found) break;
  }
  if (have) { *idx_out = best; *dis_out = bestd; }
  return OK;
}

}  // namespace kdreplay
