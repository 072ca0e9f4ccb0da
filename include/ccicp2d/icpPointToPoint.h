// ccicp2d/icpPointToPoint.h -- header-compatible IcpPointToPoint
// (ccicp2d/include/ccicp2d/icpPointToPoint.h:26-41).  The per-iteration work
// (fitStep, icpPointToPoint.cpp:33-172) runs inside the CUDA kernel; this class
// only keeps the reference's constructor / method signatures.
#ifndef CCICP2D_B200_ICP_POINT_TO_POINT_H
#define CCICP2D_B200_ICP_POINT_TO_POINT_H
#include "ccicp2d/icp.h"

class IcpPointToPoint : public Icp {
 public:
  IcpPointToPoint(double* M_GA, double* M_NGA, const int32_t M_GA_num, const int32_t M_NGA_num, const int32_t dim)
      : Icp(M_GA, M_NGA, M_GA_num, M_NGA_num, dim) {}
  virtual ~IcpPointToPoint() {}
};
#endif
