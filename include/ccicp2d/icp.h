// ccicp2d/icp.h -- header-compatible re-declaration of the reference's Icp base
// class (ccicp2d/include/ccicp2d/icp.h:33-101) on top of the C ABI in
// ccicp2d_b200.h, so that ccicp2d/src/icpTools.cpp:187-188 compiles and runs
// unchanged against libccicp2d.so.
//
//   reference                              here
//   Icp::Icp           icp.cpp:26-70       copies the model (float rounding + grid build happen on the device)
//   setMaxIterations / setMinDeltaParam / setSubsamplingStep   icp.h:48-54   same setters, same defaults
//   Icp::fit           icp.cpp:80-114      one ccicp_register_batch() call with n_pairs = 1
//   getEdgeWeight / getNumberCorrespondences  icpPointToPoint.cpp:233-320    served from the record fit() got back
#ifndef CCICP2D_B200_ICP_H
#define CCICP2D_B200_ICP_H
#include <stdint.h>
#include <vector>

#include "ccicp2d/matrix.h"

class Icp {
 public:
  Icp(double* M_GA, double* M_NGA, const int32_t M_GA_num, const int32_t M_NGA_num, const int32_t dim);
  virtual ~Icp();

  void setSubsamplingStep(int32_t val) { sub_step = val; }   // accepted and unused, like the reference
  void setMaxIterations(int32_t val) { max_iter = val; }
  void setMinDeltaParam(double val) { min_delta = val; }

  // M = R*T + t.  R (2x2) and t (2x1) are in/out; indist gates the SQUARED nearest-neighbour distance
  // (icpPointToPoint.cpp:76); h_dist is accepted and ignored (icp.cpp:80).
  void fit(double* T_GA, double* T_NGA, const int32_t T_GA_num, const int32_t T_NGA_num, Matrix& R, Matrix& t,
           double indist, double h_dist);

  void getEdgeWeight(double* eW);          // 9 doubles, row-major 3x3
  int getNumberCorrespondences(void);

  // extras (not in the reference): outcome of the last fit
  uint32_t lastStatus() const { return last_status; }     // CCICP_ST_* bits
  int32_t lastIterations() const { return last_iters; }
  // device used by every Icp object of the process (default 0); call before the first fit
  static void setDevice(int device);

 protected:
  int M_GA_SIZE;
  int M_NGA_SIZE;
  int32_t dim;
  int32_t sub_step;
  int32_t max_iter;
  double min_delta;
  int numCorr;

 private:
  std::vector<double> model_;      // GA block then NGA block, interleaved xy (copied like icp.cpp:51-60)
  bool valid_;
  double info_[9];
  uint32_t last_status;
  int32_t last_iters;
};
#endif
