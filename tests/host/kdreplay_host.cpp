#include <algorithm>
// Host build of the product's tie-replay kd-tree (slam_b200/csrc/kdreplay.h) so
// its build + traversal can be checked on the CPU against the compiled
// reference (oracle/_ref) without a GPU.  Test scaffolding only.
#include <cstdint>
#include <cstring>
#include <vector>
#include "kdreplay_ref.h"

extern "C" {

// Same record layout as oracle/ref_probe.cpp:dump_node (10 floats per node, preorder).
int kdr_tree(const double* pts, int n, int32_t* ind_out, float* nodes_out, int max_nodes) {
  std::vector<kdreplay::Pt> p(n > 0 ? n : 1);
  for (int i = 0; i < n; i++) { p[i].x = (float)pts[2 * i]; p[i].y = (float)pts[2 * i + 1]; }
  std::vector<kdreplay::Node> nodes(2 * (n > 0 ? n : 1) + 2);
  int nn = 0;
  int rc = kdreplay::build(kdreplay::PtrPts{p.data()}, n, ind_out, nodes.data(), (int)nodes.size(), &nn);
  if (rc != kdreplay::OK) return -1000 - rc;
  if (nn > max_nodes) return -nn;
  for (int k = 0; k < nn; k++) {
    const kdreplay::Node& nd = nodes[k];
    bool leaf = nd.cut_dim < 0;
    float* r = nodes_out + 10 * k;
    r[0] = leaf ? -1.0f : (float)nd.cut_dim;
    r[1] = nd.cut_val;
    r[2] = leaf ? 0.0f : nd.cut_val_left;
    r[3] = leaf ? 0.0f : nd.cut_val_right;
    r[4] = (float)nd.l; r[5] = (float)nd.u;
    r[6] = nd.box[0]; r[7] = nd.box[1]; r[8] = nd.box[2]; r[9] = nd.box[3];
  }
  return nn;
}

int kdr_nn(const double* pts, int n, const float* q, int nq, int32_t* idx, float* dis) {
  std::vector<kdreplay::Pt> p(n > 0 ? n : 1);
  for (int i = 0; i < n; i++) { p[i].x = (float)pts[2 * i]; p[i].y = (float)pts[2 * i + 1]; }
  std::vector<kdreplay::Node> nodes(2 * (n > 0 ? n : 1) + 2);
  std::vector<int32_t> ind(n > 0 ? n : 1);
  int nn = 0;
  int rc = kdreplay::build(kdreplay::PtrPts{p.data()}, n, ind.data(), nodes.data(), (int)nodes.size(), &nn);
  if (rc != kdreplay::OK) return -1000 - rc;
  for (int i = 0; i < nq; i++) {
    rc = kdreplay::nearest(kdreplay::PtrPts{p.data()}, ind.data(), nodes.data(), nn, q[2 * i], q[2 * i + 1], &idx[i], &dis[i]);
    if (rc != kdreplay::OK) return -2000 - rc;
  }
  return 0;
}


// Lazy tree (what the kernel uses): answers must equal the reference's for any query order.
int kdr_nn_lazy(const double* pts, int n, const float* q, int nq, int32_t* idx, float* dis, int* nodes_built) {
  std::vector<kdreplay::Pt> p(n > 0 ? n : 1);
  for (int i = 0; i < n; i++) { p[i].x = (float)pts[2 * i]; p[i].y = (float)pts[2 * i + 1]; }
  std::vector<kdreplay::LNode> nodes(2 * (n > 0 ? n : 1) + 8);
  std::vector<uint16_t> ind(n > 0 ? n : 1);
  int nn = 0;
  kdreplay::PtrPts acc{p.data()};
  int rc = kdreplay::lazy_init(acc, n, ind.data(), nodes.data(), &nn);
  if (rc != kdreplay::OK) return -1000 - rc;
  for (int i = 0; i < nq; i++) {
    rc = kdreplay::lazy_nearest(acc, ind.data(), nodes.data(), (int)nodes.size(), &nn, q[2 * i], q[2 * i + 1], &idx[i], &dis[i]);
    if (rc != kdreplay::OK) return -2000 - rc;
  }
  if (nodes_built) *nodes_built = nn;
  return 0;
}


// Cooperative lazy tree (the L2 kernel's path) run as a one-thread group; `cap` is the staging capacity, so a
// small value also exercises the sequential fall-back for oversized ranges.
struct HostCoop {
  int tid = 0, nt = 1;
  float* sv; uint16_t* si; int cap; int* bcast;
  float* gsv = nullptr; uint16_t* gsi = nullptr; int gcap = 0;
  uint16_t* rank = nullptr;
  void sync() {}
  void minmax4(float*, float*) {}
  int exscan(int v, int* total) { *total = v; return 0; }
};

// The rank-table partition run by `nt` emulated threads (phase by phase, like the barriers of the device code),
// and the reference's sequential partition on the same input.  ix_* receive the permuted element indices; returns
// the split index of the parallel form, *m_seq that of the sequential one.
int kdr_partition_both(const float* v, int n, float alpha, int nt, uint16_t* ix_par, uint16_t* ix_seq, int* m_seq) {
  std::vector<float> sv(v, v + n);
  for (int k = 0; k < n; k++) ix_seq[k] = (uint16_t)k;
  *m_seq = kdreplay::partition_stream(sv.data(), ix_seq, n, alpha);
  std::vector<uint16_t> T(n > 0 ? n : 1);
  const int chunk = ((n + nt - 1) / nt) | 1;
  std::vector<int> cnt(nt), bb0(nt), k0(nt), k1(nt);
  int NB = 0;
  for (int t = 0; t < nt; t++) {
    k0[t] = t * chunk < n ? t * chunk : n; k1[t] = k0[t] + chunk < n ? k0[t] + chunk : n;
    cnt[t] = kdreplay::part_count(v, k0[t], k1[t], alpha);
  }
  for (int t = 0; t < nt; t++) { bb0[t] = NB; NB += cnt[t]; }
  for (int t = 0; t < nt; t++) kdreplay::part_fill(v, k0[t], k1[t], alpha, bb0[t], NB, n, T.data());
  int K = 0;
  for (int t = 0; t < nt; t++)
    for (int k = 1 + t; k <= NB; k += nt) K += kdreplay::part_round_starts(T.data(), NB, n, k) ? 1 : 0;
  const int F = kdreplay::part_front(T.data(), NB, n, K);
  for (int k = 0; k < n; k++) ix_par[k] = (uint16_t)k;          // like ind[l + x] == si[x] after the gather
  for (int t = 0; t < nt; t++) {
    int bb = bb0[t];
    for (int x = k0[t]; x < k1[t]; x++) {
      const bool big = kdreplay::part_big(v[x], alpha);
      const int d = kdreplay::part_dest(T.data(), NB, n, F, x, big, bb);
      if (d != x) ix_par[d] = (uint16_t)x;
      bb += big ? 1 : 0;
    }
  }
  return n - NB - 1;
}

int kdr_nn_lazy_coop_g(const double* pts, int n, const float* q, int nq, int32_t* idx, float* dis, int cap, int gcap, int* nodes_built);
int kdr_nn_lazy_coop(const double* pts, int n, const float* q, int nq, int32_t* idx, float* dis, int cap, int* nodes_built) {
  return kdr_nn_lazy_coop_g(pts, n, q, nq, idx, dis, cap, 0, nodes_built);
}
// gcap > 0: ranges beyond `cap` are staged in a second pair of arrays of gcap elements (the device's global staging)
int kdr_nn_lazy_coop_g(const double* pts, int n, const float* q, int nq, int32_t* idx, float* dis, int cap, int gcap, int* nodes_built) {
  std::vector<kdreplay::Pt> p(n > 0 ? n : 1);
  for (int i = 0; i < n; i++) { p[i].x = (float)pts[2 * i]; p[i].y = (float)pts[2 * i + 1]; }
  std::vector<kdreplay::LNode> nodes(2 * (n > 0 ? n : 1) + 8);
  std::vector<uint16_t> ind(n > 0 ? n : 1);
  std::vector<float> sv(cap > 0 ? cap : 1);
  std::vector<uint16_t> si(cap > 0 ? cap : 1);
  int nn = 0, bc = 0;
  std::vector<uint16_t> rank(std::max(std::max(cap, gcap), 1));
  std::vector<float> gsv(gcap > 0 ? gcap : 1);
  std::vector<uint16_t> gsi(gcap > 0 ? gcap : 1);
  HostCoop co; co.sv = sv.data(); co.si = si.data(); co.cap = cap; co.bcast = &bc;
  co.gsv = gsv.data(); co.gsi = gsi.data(); co.gcap = gcap;
  co.rank = cap & 1 ? nullptr : rank.data();             // an odd capacity selects the sequential partition
  kdreplay::PtrPts acc{p.data()};
  int rc = kdreplay::lazy_init_coop(acc, n, ind.data(), nodes.data(), &nn, co);
  if (rc != kdreplay::OK) return -1000 - rc;
  for (int i = 0; i < nq; i++) {
    rc = kdreplay::lazy_nearest_coop(acc, ind.data(), nodes.data(), (int)nodes.size(), &nn, q[2 * i], q[2 * i + 1], &idx[i], &dis[i], co);
    if (rc != kdreplay::OK) return -2000 - rc;
  }
  if (nodes_built) *nodes_built = nn;
  return 0;
}

}
