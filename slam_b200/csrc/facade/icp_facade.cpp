// icp_facade.cpp -- libccicp2d.so: the reference's C++ class surface
// (include/ccicp2d/*.h) implemented on the C ABI of libccicp2d_b200.so.
// Error behaviour mirrors the reference: nothing throws or returns a code;
// problems are logged (the reference uses ROS_ERROR/WARN_STREAM, icp.cpp:31,39,
// 90-101) and the outputs are left untouched.
#include <cstdio>
#include <cstring>
#include <atomic>

#include "ccicp2d/icpPointToPoint.h"
#include "ccicp2d_b200.h"

namespace {
// One ccicp_ctx per HOST THREAD (thread_local, destroyed when the thread exits): two threads calling fit() at the same
// time run on their own streams and buffers.  The reference is not re-entrant at all -- matrix.cpp:26-31 keeps
// file-static scratch variables, so even distinct Matrix objects race across threads (SURVEY.md 8(b) "Threading").
std::atomic<int> g_device{0};
struct ThreadCtx {
  ccicp_ctx* ctx = nullptr;
  int device = -1;
  ~ThreadCtx() { if (ctx) ccicp_destroy(ctx); }
};
thread_local ThreadCtx t_ctx;

ccicp_ctx* context() {
  const int want = g_device.load();
  if (t_ctx.ctx && t_ctx.device != want) { ccicp_destroy(t_ctx.ctx); t_ctx.ctx = nullptr; }
  if (!t_ctx.ctx) {
    int rc = ccicp_create(&t_ctx.ctx, want);
    if (rc != CCICP_OK) {
      std::fprintf(stderr, "[ccicp2d] ERROR: cannot open CUDA device %d: %s\n", want, ccicp_strerror(rc));
      t_ctx.ctx = nullptr;
    }
    t_ctx.device = want;
  }
  return t_ctx.ctx;
}
}  // namespace

void Icp::setDevice(int device) { g_device.store(device); }   // takes effect at each thread's next fit()

Icp::Icp(double* M_GA, double* M_NGA, const int32_t M_GA_num, const int32_t M_NGA_num, const int32_t dim_)
    : M_GA_SIZE(0), M_NGA_SIZE(0), dim(dim_), sub_step(10), max_iter(20), min_delta(1e-6), numCorr(0),
      valid_(false), last_status(0), last_iters(0) {
  std::memset(info_, 0, sizeof(info_));
  if (dim != 2) {   // the reference accepts 3 here but its fitStep only handles 2-D data (icpPointToPoint.cpp:54-56)
    std::fprintf(stderr, "[ccicp2d] ERROR: this build registers data of dimensionality 2 only\n");
    return;
  }
  if (M_GA_num < 0 || M_NGA_num < 0 || M_GA_num + M_NGA_num < 5) {
    std::fprintf(stderr, "[ccicp2d] ERROR: LIBICP works only with at least 5 model points\n");
    return;
  }
  M_GA_SIZE = M_GA_num; M_NGA_SIZE = M_NGA_num;
  model_.resize(2 * (size_t)(M_GA_num + M_NGA_num));
  if (M_GA_num) std::memcpy(model_.data(), M_GA, sizeof(double) * 2 * (size_t)M_GA_num);
  if (M_NGA_num) std::memcpy(model_.data() + 2 * (size_t)M_GA_num, M_NGA, sizeof(double) * 2 * (size_t)M_NGA_num);
  valid_ = true;
}

Icp::~Icp() {}

void Icp::fit(double* T_GA, double* T_NGA, const int32_t T_GA_num, const int32_t T_NGA_num, Matrix& R, Matrix& t,
              double indist, double /*h_dist*/) {
  if (!valid_) { std::fprintf(stderr, "[ccicp2d] ERROR: Model not initialized\n"); return; }
  if (T_NGA_num < 5) std::fprintf(stderr, "[ccicp2d] WARN: Template has %d NGA points\n", T_NGA_num);
  if (T_GA_num < 5) std::fprintf(stderr, "[ccicp2d] WARN: Template has %d GA points\n", T_GA_num);
  if (T_GA_num < 0 || T_NGA_num < 0 || T_GA_num + T_NGA_num < 5) {
    std::fprintf(stderr, "[ccicp2d] ERROR: Total template has %d points\n", T_GA_num + T_NGA_num);
    return;                                           // R,t untouched (icp.cpp:100-103)
  }
  if (R.m != 2 || R.n != 2 || t.m != 2 || t.n != 1) {
    std::fprintf(stderr, "[ccicp2d] ERROR: R must be 2x2 and t 2x1\n");
    return;
  }
  std::vector<double> scene(2 * (size_t)(T_GA_num + T_NGA_num));
  if (T_GA_num) std::memcpy(scene.data(), T_GA, sizeof(double) * 2 * (size_t)T_GA_num);
  if (T_NGA_num) std::memcpy(scene.data() + 2 * (size_t)T_GA_num, T_NGA, sizeof(double) * 2 * (size_t)T_NGA_num);
  const int32_t moff[3] = {0, M_GA_SIZE, M_GA_SIZE + M_NGA_SIZE};
  const int32_t soff[3] = {0, T_GA_num, T_GA_num + T_NGA_num};
  double Rt[6] = {R.val[0][0], R.val[0][1], R.val[1][0], R.val[1][1], t.val[0][0], t.val[1][0]};
  ccicp_params prm;
  ccicp_default_params(&prm);
  prm.max_iter = max_iter; prm.min_delta = min_delta; prm.indist = indist; prm.flags = CCICP_F_INFO;
  ccicp_result res;
  std::memset(&res, 0, sizeof(res));
  ccicp_ctx* ctx = context();
  if (!ctx) return;
  const int rc = ccicp_register_batch(ctx, 1, 1, model_.data(), moff, scene.data(), soff, Rt, &prm, &res, nullptr);
  if (rc != CCICP_OK) {
    std::fprintf(stderr, "[ccicp2d] ERROR: registration failed: %s %s\n", ccicp_strerror(rc), ccicp_last_cuda_error(ctx));
    return;
  }
  last_status = res.status; last_iters = res.iters;
  if (res.status & CCICP_ST_NO_CORR) std::fprintf(stderr, "[ccicp2d] WARN: NO CORRESPONDENCES!!\n");
  R.val[0][0] = res.R[0]; R.val[0][1] = res.R[1]; R.val[1][0] = res.R[2]; R.val[1][1] = res.R[3];
  t.val[0][0] = res.t[0]; t.val[1][0] = res.t[1];
  if (!(res.status & CCICP_ST_NO_CORR)) {             // numCorr / correspondences change only on a successful fitStep
    numCorr = res.num_corr;
    std::memcpy(info_, res.info, sizeof(info_));
  }
}

void Icp::getEdgeWeight(double* eW) { std::memcpy(eW, info_, sizeof(info_)); }
int Icp::getNumberCorrespondences(void) { return numCorr; }
