// pose_math.h -- host-side pose <-> (R,t) conversions of the batched doICPMatch entry point.
//
// The reference (ccicp2d/src/icpTools.cpp:168-176, 195-212) does these with ROS tf, a third-party dependency
// that is NOT vendored and not version-pinned (ccicp2d/package.xml:10-28 lists names only).  What follows restates
// the published tf / LinearMath algorithms the call sites use:
//   tf::quaternionMsgToTF     (tf/transform_datatypes.h): normalise if |len2 - 1| > 0.1
//   tf::Matrix3x3::setRotation(LinearMath/Matrix3x3.h):   rotation matrix of a quaternion, s = 2 / len2
//   tf::Matrix3x3::getEulerYPR(LinearMath/Matrix3x3.h):   solution 1, with the gimbal-lock branch
//   tf::getYaw                (tf/transform_datatypes.h): = getRPY(...).yaw
//   tf::Quaternion::setRPY    (LinearMath/Quaternion.h):  half-angle products
// Plain double arithmetic with libm, like the reference's host code.
#pragma once
#include <cmath>

#include "../../include/ccicp2d_b200.h"

namespace ccicp_pose_math {

struct Ypr { double yaw, pitch, roll; };

inline Ypr euler_ypr(double x, double y, double z, double w) {
  double len2 = x * x + y * y + z * z + w * w;
  if (std::fabs(len2 - 1.0) > 0.1f) {          // QUATERNION_TOLERANCE is the float literal 0.1f
    const double inv = 1.0 / std::sqrt(len2);
    x *= inv; y *= inv; z *= inv; w *= inv;
    len2 = x * x + y * y + z * z + w * w;
  }
  const double s = 2.0 / len2;
  const double xs = x * s, ys = y * s, zs = z * s;
  const double wx = w * xs, wy = w * ys, wz = w * zs;
  const double xx = x * xs, xy = x * ys, xz = x * zs;
  const double yy = y * ys, yz = y * zs, zz = z * zs;
  // rows of the rotation matrix: only the entries getEulerYPR reads
  const double m00 = 1.0 - (yy + zz), m10 = xy + wz;
  const double m20 = xz - wy, m21 = yz + wx, m22 = 1.0 - (xx + yy);
  Ypr e;
  if (std::fabs(m20) >= 1.0) {
    e.yaw = 0.0;
    const double delta = std::atan2(m21, m22);
    e.pitch = m20 < 0.0 ? M_PI / 2.0 : -M_PI / 2.0;
    e.roll = delta;
  } else {
    e.pitch = -std::asin(m20);
    const double cp = std::cos(e.pitch);
    e.roll = std::atan2(m21 / cp, m22 / cp);
    e.yaw = std::atan2(m10 / cp, m00 / cp);
  }
  return e;
}

// icpTools.cpp:168-176
inline void pose_to_Rt(const ccicp_pose& p, double* Rt) {
  const double theta = euler_ypr(p.qx, p.qy, p.qz, p.qw).yaw;
  Rt[0] = std::cos(theta); Rt[1] = -std::sin(theta);
  Rt[2] = std::sin(theta); Rt[3] = std::cos(theta);
  Rt[4] = p.px; Rt[5] = p.py;
}

// icpTools.cpp:195-212
inline void Rt_to_pose(const double* Rt, const ccicp_pose& init, ccicp_pose* out) {
  const double corr_yaw = std::atan2(Rt[2], Rt[0]);
  const Ypr e = euler_ypr(init.qx, init.qy, init.qz, init.qw);
  const double hy = corr_yaw * 0.5, hp = e.pitch * 0.5, hr = e.roll * 0.5;
  const double cy = std::cos(hy), sy = std::sin(hy);
  const double cp = std::cos(hp), sp = std::sin(hp);
  const double cr = std::cos(hr), sr = std::sin(hr);
  out->px = Rt[4]; out->py = Rt[5]; out->pz = init.pz;
  out->qx = sr * cp * cy - cr * sp * sy;
  out->qy = cr * sp * cy + sr * cp * sy;
  out->qz = cr * cp * sy - sr * sp * cy;
  out->qw = cr * cp * cy + sr * sp * sy;
}

inline void sentinel(ccicp_pose* out) {        // icpTools.cpp:179-184: default-constructed pose, w = 9999
  out->px = out->py = out->pz = 0.0;
  out->qx = out->qy = out->qz = 0.0;
  out->qw = 9999.0;
}

}  // namespace ccicp_pose_math
