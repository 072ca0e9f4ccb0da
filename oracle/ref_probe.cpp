// oracle/ref_probe.cpp -- TEST INFRASTRUCTURE ONLY.
//
// C-ABI harness around the UNMODIFIED reference sources
//   /root/reference/ccicp2d/src/{icp,icpPointToPoint,kdtree,matrix}.cpp
// which oracle/Makefile compiles where they lie (nothing is copied into this
// repo) into oracle/_ref/libccicp_ref.so.  The product never links, imports
// or executes this file; only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs do.
//
// The reference is driven exactly the way its only call site does
// (ccicp2d/src/icpTools.cpp:187-188): construct IcpPointToPoint per call
// (kd-trees rebuilt), then fit(T_GA, T_NGA, nGA, nNGA, R, t, indist, 0).
//
// This translation unit is compiled with -fno-access-control so it can call
// the private virtual fitStep (icpPointToPoint.h:37) and read the protected
// members (icp.h:79-99) and KDTree internals (kdtree.h:132-139) without
// editing any reference file.
#include <cstdint>
#include <cstring>
#include <vector>

#include "ccicp2d/icpPointToPoint.h"

namespace {

struct Pose2 {
  Matrix R, t;
  Pose2() : R(2, 2), t(2, 1) {}
  void load(const double* Rt) {
    R.val[0][0] = Rt[0]; R.val[0][1] = Rt[1];
    R.val[1][0] = Rt[2]; R.val[1][1] = Rt[3];
    t.val[0][0] = Rt[4]; t.val[1][0] = Rt[5];
  }
  void store(double* Rt) const {
    Rt[0] = R.val[0][0]; Rt[1] = R.val[0][1];
    Rt[2] = R.val[1][0]; Rt[3] = R.val[1][1];
    Rt[4] = t.val[0][0]; Rt[5] = t.val[1][0];
  }
};

// The reference constructor leaves the trees NULL and M_*_SIZE uninitialised
// when nGA+nNGA < 5 (icp.cpp:37-43) and fit() would then dereference them;
// the facade never gets there with real data, so the probe refuses instead.
inline bool model_ok(int nmGA, int nmNGA) { return nmGA + nmNGA >= 5; }

double dummy_pt[2] = {0.0, 0.0};
inline double* ptr_or_dummy(const double* p) {
  return p ? const_cast<double*>(p) : dummy_pt;
}

// Replays one class's queries for the running R,t exactly as fitStep does
// (icpPointToPoint.cpp:69-76 / 104-111) and records the model index or -1.
void replay_class(kdtree::KDTree* tree, int model_n, const double* T, int n,
                  const Matrix& R, const Matrix& t, double inDist,
                  int32_t* corr, float* dis_out, uint8_t* tie_out) {
  std::vector<float> query(2);
  kdtree::KDTreeResultVector result;
  double r00 = R.val[0][0]; double r01 = R.val[0][1];
  double r10 = R.val[1][0]; double r11 = R.val[1][1];
  double t0 = t.val[0][0]; double t1 = t.val[1][0];
  for (int i = 0; i < n; i++) {
    if (model_n <= 3) {  // class skipped entirely (icpPointToPoint.cpp:59,93)
      corr[i] = -1;
      if (dis_out) dis_out[i] = -1.0f;
      if (tie_out) tie_out[i] = 0;
      continue;
    }
    query[0] = r00 * T[i * 2 + 0] + r01 * T[i * 2 + 1] + t0;
    query[1] = r10 * T[i * 2 + 0] + r11 * T[i * 2 + 1] + t1;
    tree->n_nearest(query, 1, result);
    corr[i] = (result[0].dis < inDist) ? result[0].idx : -1;
    if (dis_out) dis_out[i] = result[0].dis;
    if (tie_out) {
      // brute-force count of model points sharing the float minimum
      int shared = 0;
      for (int m = 0; m < model_n; m++) {
        float dx = tree->the_data[m][0] - query[0];
        float dy = tree->the_data[m][1] - query[1];
        float d = dx * dx;
        d += dy * dy;
        if (d == result[0].dis) shared++;
      }
      tie_out[i] = (shared > 1) ? 1 : 0;
    }
  }
}

void dump_node(kdtree::KDTreeNode* nd, std::vector<float>& rec) {
  // preorder; 10 floats per node:
  // cut_dim, cut_val, cut_val_left, cut_val_right, l, u, box0.lo, box0.hi, box1.lo, box1.hi
  // followed by a leaf flag folded into cut_dim = -1 for terminals
  bool leaf = (nd->left == NULL && nd->right == NULL);
  rec.push_back(leaf ? -1.0f : (float)nd->cut_dim);
  rec.push_back(nd->cut_val);
  rec.push_back(leaf ? 0.0f : nd->cut_val_left);
  rec.push_back(leaf ? 0.0f : nd->cut_val_right);
  rec.push_back((float)nd->l);
  rec.push_back((float)nd->u);
  rec.push_back(nd->box[0].lower);
  rec.push_back(nd->box[0].upper);
  rec.push_back(nd->box[1].lower);
  rec.push_back(nd->box[1].upper);
  if (nd->left) dump_node(nd->left, rec);
  if (nd->right) dump_node(nd->right, rec);
}

}  // namespace

extern "C" {

// Status codes shared with the port (oracle/ccicp_port.h).
enum { CCREF_OK = 0, CCREF_TOO_FEW_MODEL = 1, CCREF_TOO_FEW_SCENE = 2, CCREF_NO_CORR = 4 };

// One registration through the public API only.  Rt = {r00,r01,r10,r11,t0,t1}
// in/out.  eW may be NULL.  Returns a CCREF_* status.
int ccref_fit(const double* mGA, int nmGA, const double* mNGA, int nmNGA,
              const double* sGA, int nsGA, const double* sNGA, int nsNGA,
              int max_iter, double min_delta, double indist,
              double* Rt, int* num_corr, double* eW) {
  if (num_corr) *num_corr = 0;
  if (!model_ok(nmGA, nmNGA)) return CCREF_TOO_FEW_MODEL;
  if (nsGA + nsNGA < 5) return CCREF_TOO_FEW_SCENE;  // icp.cpp:100-103: R,t untouched
  IcpPointToPoint icp(ptr_or_dummy(mGA), ptr_or_dummy(mNGA), nmGA, nmNGA, 2);
  icp.setMaxIterations(max_iter);
  icp.setMinDeltaParam(min_delta);
  icp.numCorr = 0;  // uninitialised in the reference until the first successful fitStep
  Pose2 p;
  p.load(Rt);
  icp.fit(ptr_or_dummy(sGA), ptr_or_dummy(sNGA), nsGA, nsNGA, p.R, p.t, indist, 0.0);
  p.store(Rt);
  if (num_corr) *num_corr = icp.getNumberCorrespondences();
  if (icp.cModelPoints.m == 0) return CCREF_NO_CORR;
  if (eW) icp.getEdgeWeight(eW);
  return CCREF_OK;
}

// Lock-step trace: runs fitStep one iteration at a time (equivalent to
// fitIterate, icp.cpp:116-122) and records, per executed iteration k:
//   Rt_trace[k]   running R,t BEFORE the step ([n_iter+1][6]; slot iters_run = final)
//   corr[k]       per scene point (GA block then NGA block) model index or -1
//   dis[k]        float squared NN distance (optional, NULL to skip)
//   tie[k]        1 where >=2 model points share the float minimum (optional)
//   ncorr[k], delta[k]
// Returns a CCREF_* status; *iters_run = number of fitStep calls made.
int ccref_trace(const double* mGA, int nmGA, const double* mNGA, int nmNGA,
                const double* sGA, int nsGA, const double* sNGA, int nsNGA,
                int max_iter, double min_delta, double indist, const double* Rt0,
                double* Rt_trace, int32_t* corr, float* dis, uint8_t* tie,
                int32_t* ncorr, double* delta, int* iters_run, double* eW) {
  *iters_run = 0;
  memcpy(Rt_trace, Rt0, 6 * sizeof(double));
  if (!model_ok(nmGA, nmNGA)) return CCREF_TOO_FEW_MODEL;
  if (nsGA + nsNGA < 5) return CCREF_TOO_FEW_SCENE;
  IcpPointToPoint icp(ptr_or_dummy(mGA), ptr_or_dummy(mNGA), nmGA, nmNGA, 2);
  icp.numCorr = 0;
  double* tGA = ptr_or_dummy(sGA);
  double* tNGA = ptr_or_dummy(sNGA);
  Pose2 p;
  p.load(Rt0);
  const int ns = nsGA + nsNGA;
  int status = CCREF_OK;
  int k = 0;
  for (; k < max_iter; k++) {
    p.store(Rt_trace + 6 * k);
    replay_class(icp.M_tree_GA, nmGA, tGA, nsGA, p.R, p.t, indist,
                 corr + (size_t)k * ns, dis ? dis + (size_t)k * ns : NULL,
                 tie ? tie + (size_t)k * ns : NULL);
    replay_class(icp.M_tree_NGA, nmNGA, tNGA, nsNGA, p.R, p.t, indist,
                 corr + (size_t)k * ns + nsGA, dis ? dis + (size_t)k * ns + nsGA : NULL,
                 tie ? tie + (size_t)k * ns + nsGA : NULL);
    double d = icp.fitStep(tGA, tNGA, nsGA, nsNGA, p.R, p.t, indist);
    delta[k] = d;
    int nc = 0;
    for (int i = 0; i < ns; i++) nc += (corr[(size_t)k * ns + i] >= 0);
    ncorr[k] = nc;
    if (d < min_delta) { k++; break; }
  }
  *iters_run = k;
  p.store(Rt_trace + 6 * k);
  if (icp.cModelPoints.m == 0) status |= CCREF_NO_CORR;
  else if (eW) icp.getEdgeWeight(eW);
  return status;
}

// kd-tree structure dump for validating restated / device tree builders.
// ind_out[n]; nodes_out[10*max_nodes] preorder records (see dump_node).
int ccref_tree(const double* pts, int n, int32_t* ind_out, float* nodes_out, int max_nodes) {
  if (n <= 0) return 0;
  kdtree::KDTreeArray data;
  data.resize(boost::extents[n][2]);
  for (int i = 0; i < n; i++) {
    data[i][0] = (float)pts[2 * i];
    data[i][1] = (float)pts[2 * i + 1];
  }
  kdtree::KDTree tree(data);
  for (int i = 0; i < n; i++) ind_out[i] = tree.ind[i];
  std::vector<float> rec;
  dump_node(tree.root, rec);
  int nn = (int)(rec.size() / 10);
  if (nn > max_nodes) return -nn;
  memcpy(nodes_out, rec.data(), rec.size() * sizeof(float));
  return nn;
}

// Raw 1-NN queries (float coords) against a tree built from pts.
void ccref_nn(const double* pts, int n, const float* q, int nq, int32_t* idx, float* dis) {
  kdtree::KDTreeArray data;
  data.resize(boost::extents[n][2]);
  for (int i = 0; i < n; i++) {
    data[i][0] = (float)pts[2 * i];
    data[i][1] = (float)pts[2 * i + 1];
  }
  kdtree::KDTree tree(data);
  std::vector<float> qv(2);
  kdtree::KDTreeResultVector res;
  for (int i = 0; i < nq; i++) {
    qv[0] = q[2 * i]; qv[1] = q[2 * i + 1];
    tree.n_nearest(qv, 1, res);
    idx[i] = res[0].idx; dis[i] = res[0].dis;
  }
}

// Timed throughput loop for the CPU baseline: `reps` full registrations
// (construct + fit, like icpTools.cpp:187-188) of the same pair.
int ccref_bench(const double* mGA, int nmGA, const double* mNGA, int nmNGA,
                const double* sGA, int nsGA, const double* sNGA, int nsNGA,
                int max_iter, double min_delta, double indist, const double* Rt0, double* Rt_out) {
  double Rt[6];
  memcpy(Rt, Rt0, sizeof(Rt));
  int nc;
  int st = ccref_fit(mGA, nmGA, mNGA, nmNGA, sGA, nsGA, sNGA, nsNGA, max_iter, min_delta,
                     indist, Rt, &nc, NULL);
  if (Rt_out) memcpy(Rt_out, Rt, sizeof(Rt));
  return st;
}

}  // extern "C"
