// callsite_demo.cpp -- the reference's call site (ccicp2d/src/icpTools.cpp:168-197), stand-alone.
// This ONE source is compiled twice without change:
//   tests/host/callsite_b200   against include/ccicp2d/*.h + libccicp2d.so      (the B200 drop-in)
//   oracle/_ref/callsite_ref   against /root/reference/ccicp2d/include + the unmodified reference sources
// and tests/test_facade_gpu.py checks that both print the same registration.  Test scaffolding.
//
// usage: callsite_demo <pair.bin> [max_iter min_delta]
// pair.bin: int32 nmGA,nmNGA,nsGA,nsNGA; double x,y,yaw (initial pose); then the four xy arrays (doubles).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "ccicp2d/icpPointToPoint.h"

int main(int argc, char** argv) {
  if (argc < 2) return 2;
  FILE* f = std::fopen(argv[1], "rb");
  if (!f) return 3;
  int32_t n[4];
  double pose[3];
  if (std::fread(n, 4, 4, f) != 4 || std::fread(pose, 8, 3, f) != 3) return 4;
  std::vector<double> a[4];
  for (int k = 0; k < 4; k++) {
    a[k].resize(2 * (size_t)n[k] + 2);
    if (n[k] && std::fread(a[k].data(), 16, n[k], f) != (size_t)n[k]) return 5;
  }
  std::fclose(f);

  // ccicp2d/src/icpTools.cpp:168-176
  Matrix rot(2, 2);
  Matrix trans(2, 1);
  trans.val[0][0] = pose[0];
  trans.val[1][0] = pose[1];
  rot.val[0][0] = std::cos(pose[2]); rot.val[0][1] = -std::sin(pose[2]);
  rot.val[1][0] = std::sin(pose[2]); rot.val[1][1] = std::cos(pose[2]);

  // ccicp2d/src/icpTools.cpp:187-188
  IcpPointToPoint icp(a[0].data(), a[1].data(), n[0], n[1], (int32_t)(2));
  if (argc >= 4) { icp.setMaxIterations(std::atoi(argv[2])); icp.setMinDeltaParam(std::atof(argv[3])); }
  icp.fit(a[2].data(), a[3].data(), n[2], n[3], rot, trans, 5, 0);

  double eW[9];
  icp.getEdgeWeight(eW);
  // ccicp2d/src/icpTools.cpp:195-197
  std::printf("%.17g %.17g %.17g\n", trans.val[0][0], trans.val[1][0], std::atan2(rot.val[1][0], rot.val[0][0]));
  std::printf("%d\n", icp.getNumberCorrespondences());
  for (int k = 0; k < 9; k++) std::printf("%.17g%c", eW[k], k == 8 ? '\n' : ' ');
  return 0;
}
