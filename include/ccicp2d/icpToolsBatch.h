// ccicp2d/icpToolsBatch.h -- the reference's immediate caller, batched.
//
// CCICPBatch mirrors CCICP::doICPMatch(model_GA, model_NGA, scene_GA, scene_NGA, initPose)
// (ccicp2d/include/ccicp2d/icpTools.h:44-47, ccicp2d/src/icpTools.cpp:123-218) without pulling PCL / ROS / Eigen
// headers into this library: the methods are templates, so the reference's own argument types work unchanged --
//   point containers : anything indexable whose elements give x = e(0), y = e(1)     (std::vector<Eigen::Vector2d>)
//   poses            : anything with .pose.position.{x,y,z} and .pose.orientation.{x,y,z,w}
//                      (geometry_msgs::PoseStamped)
// Same semantics: R = rot(yaw(initPose)), t = (x, y); fewer than 5 scene points -> orientation.w = 9999 and nothing
// else set (icpTools.cpp:179-184); otherwise IcpPointToPoint + fit(..., indist 5, 0) with the Icp defaults
// (max_iter 20, min_delta 1e-6), result x,y from t, z / roll / pitch carried over from initPose, yaw = atan2(R10,R00).
//
//   one pair  : PoseStamped out = batch.doICPMatch(mGA, mNGA, sGA, sNGA, initPose);       // drop-in, synchronous
//   many pairs: int id = batch.add(mGA, mNGA, sGA, sNGA, initPose); ... batch.run(); batch.result(id, out);
// run() registers every queued pair in ONE ccicp_match_batch call (one CTA per pair on the GPU).
#ifndef CCICP2D_B200_ICP_TOOLS_BATCH_H
#define CCICP2D_B200_ICP_TOOLS_BATCH_H
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "ccicp2d_b200.h"

#ifndef ICP_MAX_PTS
#define ICP_MAX_PTS 20000      // ccicp2d/include/ccicp2d/icpTools.h:21
#endif

// Growable array of doubles in page-locked host memory (ccicp_host_alloc): the packed scan coordinates are what the
// batch call uploads chunk by chunk under the running kernels, which only overlaps from pinned memory.  Falls back to
// malloc when pinned memory is not available.
class CCICPPinnedDoubles {
 public:
  CCICPPinnedDoubles() : p_(nullptr), n_(0), cap_(0), pinned_(false) {}
  ~CCICPPinnedDoubles() { release(); }
  CCICPPinnedDoubles(const CCICPPinnedDoubles&) = delete;
  CCICPPinnedDoubles& operator=(const CCICPPinnedDoubles&) = delete;
  void clear() { n_ = 0; }
  size_t size() const { return n_; }
  const double* data() const { return p_; }
  void push2(double a, double b) {
    if (n_ + 2 > cap_) grow(n_ + 2);
    p_[n_++] = a; p_[n_++] = b;
  }

 private:
  void grow(size_t need) {
    size_t cap = cap_ ? 2 * cap_ : 4096;
    while (cap < need) cap *= 2;
    void* q = nullptr;
    bool pinned = ccicp_host_alloc(&q, cap * sizeof(double)) == CCICP_OK && q;
    if (!pinned) q = std::malloc(cap * sizeof(double));
    if (!q) { std::fprintf(stderr, "CCICPBatch: out of host memory\n"); std::abort(); }
    if (n_) std::memcpy(q, p_, n_ * sizeof(double));
    release();
    p_ = (double*)q; cap_ = cap; pinned_ = pinned;
  }
  void release() {
    if (p_) { if (pinned_) ccicp_host_free(p_); else std::free(p_); }
    p_ = nullptr; cap_ = 0;
  }
  double* p_;
  size_t n_, cap_;
  bool pinned_;
};

class CCICPBatch {
 public:
  explicit CCICPBatch(int device = 0) : ctx_(nullptr), status_(ccicp_create(&ctx_, device)) {
    ccicp_default_params(&prm_);        // max_iter 20, min_delta 1e-6 (icp.cpp:27), indist 5 (icpTools.cpp:188)
    clear();
    if (status_ != CCICP_OK) std::fprintf(stderr, "CCICPBatch: %s\n", ccicp_strerror(status_));
  }
  ~CCICPBatch() { ccicp_destroy(ctx_); }
  CCICPBatch(const CCICPBatch&) = delete;
  CCICPBatch& operator=(const CCICPBatch&) = delete;

  int status() const { return status_; }              // CCICP_E_* of the last call (0 = OK); never throws / exits
  ccicp_params& params() { return prm_; }
  size_t size() const { return init_.size(); }

  void clear() {
    model_xy_.clear(); scene_xy_.clear(); init_.clear(); out_.clear(); rec_.clear();
    model_off_.assign(1, 0); scene_off_.assign(1, 0);
  }

  // Queue one pair; returns its index in the batch.
  template <class PtsA, class PtsB, class PtsC, class PtsD, class PoseStampedT>
  int add(const PtsA& model_GA, const PtsB& model_NGA, const PtsC& scene_GA, const PtsD& scene_NGA,
          const PoseStampedT& initPose) {
    append(model_GA, model_xy_, model_off_);
    append(model_NGA, model_xy_, model_off_);
    append(scene_GA, scene_xy_, scene_off_);
    append(scene_NGA, scene_xy_, scene_off_);
    ccicp_pose p;
    p.px = initPose.pose.position.x; p.py = initPose.pose.position.y; p.pz = initPose.pose.position.z;
    p.qx = initPose.pose.orientation.x; p.qy = initPose.pose.orientation.y;
    p.qz = initPose.pose.orientation.z; p.qw = initPose.pose.orientation.w;
    init_.push_back(p);
    return (int)init_.size() - 1;
  }

  // Register everything queued.  Returns a CCICP_E_* code.
  int run() {
    if (status_ != CCICP_OK && !ctx_) return status_;
    const int32_t n = (int32_t)init_.size();
    out_.resize(init_.size());
    rec_.resize(init_.size());
    status_ = ccicp_match_batch(ctx_, n, model_xy_.data(), model_off_.data(), scene_xy_.data(), scene_off_.data(),
                                init_.data(), &prm_, out_.data(), rec_.data());
    if (status_ != CCICP_OK)
      std::fprintf(stderr, "CCICPBatch::run: %s %s\n", ccicp_strerror(status_), ccicp_last_cuda_error(ctx_));
    return status_;
  }

  // Result of pair `id` after run(), written into a PoseStamped-like object the way doICPMatch fills its return value.
  template <class PoseStampedT>
  void result(int id, PoseStampedT& out) const {
    const ccicp_pose& p = out_[(size_t)id];
    out.pose.position.x = p.px; out.pose.position.y = p.py; out.pose.position.z = p.pz;
    out.pose.orientation.x = p.qx; out.pose.orientation.y = p.qy; out.pose.orientation.z = p.qz;
    out.pose.orientation.w = p.qw;
  }
  const ccicp_result& record(int id) const { return rec_[(size_t)id]; }

  // The reference's signature: one pair, synchronous.
  template <class PtsA, class PtsB, class PtsC, class PtsD, class PoseStampedT>
  PoseStampedT doICPMatch(const PtsA& model_GA, const PtsB& model_NGA, const PtsC& scene_GA, const PtsD& scene_NGA,
                          PoseStampedT& initPose) {
    clear();
    const int id = add(model_GA, model_NGA, scene_GA, scene_NGA, initPose);
    PoseStampedT res = PoseStampedT();
    if ((int64_t)(scene_off_[2] - scene_off_[0]) < 5)
      std::fprintf(stderr, "ERROR: Total Scene has %d points\n", (int)(scene_off_[2] - scene_off_[0]));
    if (run() == CCICP_OK) result(id, res);
    else res.pose.orientation.w = 9999;      // the caller's failure test (scan_registration.cpp:161)
    return res;
  }

 private:
  template <class Pts>
  static void append(const Pts& pts, CCICPPinnedDoubles& xy, std::vector<int32_t>& off) {
    const size_t n = pts.size();
    for (size_t i = 0; i < n; i++) xy.push2(pts[i](0), pts[i](1));
    off.push_back(off.back() + (int32_t)n);
  }

  ccicp_ctx* ctx_;
  int status_;
  ccicp_params prm_;
  CCICPPinnedDoubles model_xy_, scene_xy_;
  std::vector<int32_t> model_off_, scene_off_;
  std::vector<ccicp_pose> init_, out_;
  std::vector<ccicp_result> rec_;
};
#endif
