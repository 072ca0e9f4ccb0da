// oracle/classify_probe.cpp -- TEST INFRASTRUCTURE ONLY.
// C entry point around the reference's own CCICP::classifyPoints: the function body is #included from
// oracle/_ref/classify_ref_body.inc, which oracle/Makefile cuts out of /root/reference/ccicp2d/src/icpTools.cpp
// (lines 36-103) at build time.  Pins slam_b200's ccicp_classify_points to the reference.
#include "shim_pcl/classify_shim.h"
#include "_ref/classify_ref_body.inc"

extern "C" {
// xyz: float [n][3].  Outputs (capacity n): out_xyz [k][3], out_ga [k]; returns k = points kept.
// The reference converts floor(NaN) to int (undefined): callers pass finite coordinates only.
int ccref_classify(const float* xyz, int n, float* out_xyz, uint8_t* out_ga) {
  pcl::PointCloud<PointXYZGD>::Ptr cloud(new pcl::PointCloud<PointXYZGD>);
  cloud->points.resize(n > 0 ? n : 0);
  for (int i = 0; i < n; i++) {
    PointXYZGD p;
    p.x = xyz[3 * i]; p.y = xyz[3 * i + 1]; p.z = xyz[3 * i + 2]; p.pad_ = 1.0f; p.ground_adj = 0; p.drivable = 0;
    cloud->points[i] = p;
  }
  CCICP c;
  c.classifyPoints(cloud);
  const int k = (int)cloud->points.size();
  for (int i = 0; i < k; i++) {
    out_xyz[3 * i] = cloud->points[i].x; out_xyz[3 * i + 1] = cloud->points[i].y; out_xyz[3 * i + 2] = cloud->points[i].z;
    out_ga[i] = (uint8_t)(cloud->points[i].ground_adj != 0);
  }
  return k;
}
}
